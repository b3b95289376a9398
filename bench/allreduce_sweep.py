"""BASELINE.json config 5: gradient all-reduce bus bandwidth, 1 KB - 1 GB, ours vs NCCL.

    python -m torch.distributed.run --nproc-per-node N bench/allreduce_sweep.py [--dtype bf16]

Device-timed (CUDA events), max over ranks, busbw = 2 (W-1)/W * bytes / time."""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def timed(fn, iters):
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--ctas", type=int, default=0)
    ns = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("cpu:gloo,cuda:nccl")
    from distributed_training_b200.ops import native
    from distributed_training_b200.parallel.symm import SymmComm

    dtype = torch.bfloat16 if ns.dtype == "bf16" else torch.float32
    esize = 2 if dtype == torch.bfloat16 else 4
    comm = SymmComm(torch.device("cuda", torch.cuda.current_device()), segment_bytes=(ns.max_mb << 20) + (64 << 20),
                    comm_ctas=ns.ctas, overlap_stream=False)
    sizes = [1 << p for p in range(10, 31) if (1 << p) <= (ns.max_mb << 20)]
    big = comm.alloc((ns.max_mb << 20) // esize, dtype)
    nccl_buf = torch.empty((ns.max_mb << 20) // esize, dtype=dtype, device="cuda")
    rows = []
    for nbytes in sizes:
        n = nbytes // esize
        buf = big[:n]
        seg, off = comm.locate(buf)
        iters = 50 if nbytes <= (1 << 22) else (20 if nbytes <= (1 << 26) else 5)
        ours = lambda: native.comm_all_reduce_avg(seg.desc, off, n, dtype == torch.float32, 0, ns.ctas)
        ref_t = nccl_buf[:n]

        def nccl():
            dist.all_reduce(ref_t)
            ref_t.div_(world)

        for fn in (ours, nccl):
            for _ in range(3):
                fn()
        t_ours, t_nccl = timed(ours, iters), timed(nccl, iters)
        factor = 2 * (world - 1) / world * nbytes / 1e6
        row = {"bytes": nbytes, "world": world, "dtype": ns.dtype, "ours_us": t_ours * 1e3, "nccl_us": t_nccl * 1e3,
               "ours_busbw_gbs": factor / t_ours, "nccl_busbw_gbs": factor / t_nccl}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    comm.check_errors()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
