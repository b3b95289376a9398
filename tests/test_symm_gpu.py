"""Single-GPU checks of the symmetric heap's allocator rules (the multi-rank heap paths are in test_comm_gpu.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_reserved_segment_is_preferred_by_following_allocations():
    """``reserve(n)`` must pin the segment it prepared.  First-fit over all segments put a
    weight's shard in an older segment with a hole and its full/grad buffers in the reserved one; the fused kernels
    then combined one segment's descriptor with another segment's offsets."""
    from distributed_training_b200.parallel.symm import SymmComm

    comm = SymmComm(torch.device("cuda", 0), segment_bytes=8 << 20)
    a = comm.alloc(1 << 20, torch.bfloat16)                 # 2 MiB in segment 0, 6 MiB left
    seg0 = comm.locate(a)[0]
    seg = comm.reserve(7 << 20)                             # does not fit in segment 0 -> new segment
    assert seg is not seg0 and len(comm.segments) == 2
    b = comm.alloc(1 << 20, torch.bfloat16)                 # would still fit in segment 0's hole
    c = comm.alloc(1 << 20, torch.bfloat16)
    assert comm.locate(b)[0] is seg and comm.locate(c)[0] is seg
    assert comm.locate(b)[1] + (2 << 20) == comm.locate(c)[1]
    # a fused call that needs both operands in one segment now succeeds across the old hole
    b.fill_(1.0)
    c.fill_(2.0)
    comm.all_reduce_sgd_(b, c, 0.5, 0.0).wait()             # world 1: params = c - 0.5 * b
    torch.cuda.synchronize()
    assert float(c.float().min()) == 1.5 == float(c.float().max())
    comm.close()


def test_native_sharded_module_spanning_small_segments():
    """A model whose sharded state does not fit the default segment gets ONE segment sized for it, and the module
    refuses to run with offsets from mixed segments."""
    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.parallel.sharded_native import NativeShardedModule
    from distributed_training_b200.parallel.symm import SymmComm

    comm = SymmComm(torch.device("cuda", 0), segment_bytes=1 << 20)     # far smaller than the model
    comm.alloc(1 << 18, torch.bfloat16)                                  # leaves a hole in segment 0
    torch.manual_seed(0)
    model = MLP(256, 512, 256, num_layers=3, dtype=torch.bfloat16).cuda()
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    wrapped = NativeShardedModule(model, comm, OptimizerSpec("adamw", lr=1e-2))
    segs = {id(sw.segment) for sw in wrapped.weights.values()}
    assert len(segs) == 1 and wrapped.segment is next(iter(wrapped.weights.values())).segment
    for pool in (wrapped.full_pool, wrapped.grad_pool):
        for slots in pool.values():
            for s in slots:
                assert comm.locate(s.buf)[0] is wrapped.segment
    x = torch.randn(128, 256, device="cuda").to(torch.bfloat16)
    y = torch.randint(0, 256, (128,), device="cuda")
    l0 = float(wrapped.forward_backward(x, y))
    for _ in range(10):
        loss = wrapped.forward_backward(x, y)
    torch.cuda.synchronize()
    assert float(loss) < l0
    comm.close()


def test_loader_device_gather_matches_host_gather():
    """The kernel that assembles a batch straight out of pinned host memory (dtb_gather_rows on the copy stream)
    yields exactly the batches of the host index_select + staged-copy path, ragged last batch included."""
    from distributed_training_b200.data import PooledWideDataset, ShardedLoader, ShardedSampler, TokenDataset

    dev = torch.device("cuda", 0)
    for ds in (PooledWideDataset(1000, 264, 10, pool=128), TokenDataset(100, 33, 1000)):
        host = ShardedLoader(ds, 32, sampler=ShardedSampler(ds, 2, 1, seed=3), device=dev, device_gather=False)
        gpu = ShardedLoader(ds, 32, sampler=ShardedSampler(ds, 2, 1, seed=3), device=dev)
        assert gpu.device_gather and not host.device_gather
        a, b = list(host), list(gpu)
        torch.cuda.synchronize()
        assert len(a) == len(b) and a[-1][0].shape[0] < 32
        for (xa, ya), (xb, yb) in zip(a, b):
            assert xb.is_cuda and torch.equal(xa, xb) and torch.equal(ya, yb)
        again = list(gpu)                                   # staging slots and events are reused across epochs
        torch.cuda.synchronize()
        assert all(torch.equal(p[0], q[0]) for p, q in zip(b, again))
