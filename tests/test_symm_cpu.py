"""Host-side pieces of the symmetric heap that need no GPU: the SCM_RIGHTS file-descriptor exchange that carries the
VMM allocation / multicast handles between ranks, and the segment-pinning rule of the allocator."""
import os

import torch.multiprocessing as mp

import _workers as W
from conftest import free_port


def _fd_worker(rank, world, port, out_dir):
    import torch.distributed as dist

    from distributed_training_b200.parallel.symm import exchange_fds

    W.init(rank, world, port)
    try:
        fd = os.memfd_create(f"rank{rank}")
        os.write(fd, f"hello from {rank}".encode())
        fds = exchange_fds(fd, rank, world, 1)
        for q in range(world):
            assert os.pread(fds[q], 64, 0) == f"hello from {q}".encode()
        only_root = exchange_fds(fd if rank == 0 else None, rank, world, 2)     # multicast handle: rank 0 only
        assert only_root[0] is not None and all(only_root[q] is None for q in range(1, world))
        assert os.pread(only_root[0], 64, 0) == b"hello from 0"
        open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_fd_exchange_between_ranks(tmp_path):
    world = 3
    mp.spawn(_fd_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
