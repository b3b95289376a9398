"""BASELINE.json config 5: gradient all-reduce bus bandwidth, 1 KB - 1 GB, ours vs NCCL.

    python -m torch.distributed.run --nproc-per-node N bench/allreduce_sweep.py [--dtype bf16] [--out FILE]

Per size: our default dispatch ("auto": one-shot <= 64 KiB, NVLS multimem two-shot when the heap has a multicast
mapping and world >= 4, else peer-load two-shot), the peer-load two-shot and the NVLS two-shot forced, and NCCL
(``all_reduce`` + ``div_``, what the reference's DDP does: /root/reference/src/playground/ddp_script.py:150-154).
Device-timed (CUDA events), max over ranks, busbw = 2 (W-1)/W * bytes / time.

``--p2p`` additionally measures the NVLink denominators used in BASELINE.md: a peer -> local copy-engine copy and
an SM copy kernel (one direction, and both directions at once), plus broadcast and all-gather rates."""
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
    ap.add_argument("--min-kb", type=int, default=1)
    ap.add_argument("--ctas", type=int, default=0)
    ap.add_argument("--out", default="")
    ap.add_argument("--p2p", action="store_true")
    ap.add_argument("--nccl", type=int, default=1)
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
    sizes = [1 << p for p in range(10, 31) if (ns.min_kb << 10) <= (1 << p) <= (ns.max_mb << 20)]
    big = comm.alloc((ns.max_mb << 20) // esize, dtype)
    nccl_buf = torch.empty((ns.max_mb << 20) // esize, dtype=dtype, device="cuda")
    fp32 = dtype == torch.float32
    out = open(ns.out, "a") if (ns.out and rank == 0) else None

    def emit(row):
        if rank == 0:
            line = json.dumps(row)
            print(line, flush=True)
            if out:
                out.write(line + "\n")
                out.flush()

    seg, base_off = comm.locate(big)
    emit({"heap": comm.backend, "multicast": comm.multicast, "nvls_default": comm.nvls_active, "world": world})
    for nbytes in sizes:
        n = nbytes // esize
        iters = 50 if nbytes <= (1 << 22) else (20 if nbytes <= (1 << 26) else 5)
        factor = 2 * (world - 1) / world * nbytes / 1e6
        row = {"bytes": nbytes, "world": world, "dtype": ns.dtype}
        variants = {"auto": (0, ns.ctas), "p2p": (2 if nbytes > (64 << 10) else 1, ns.ctas)}
        if comm.multicast:
            variants["nvls"] = (3, ns.ctas)
            if nbytes >= (8 << 20):
                variants["nvls_x16"] = (3, 16 | native.EXCLUSIVE)
                variants["nvls_x8"] = (3, 8 | native.EXCLUSIVE)
                variants["p2p_x16"] = (2, 16 | native.EXCLUSIVE)
        for name, (algo, ctas) in variants.items():
            fn = lambda a=algo, c=ctas: native.comm_all_reduce_avg(seg.desc, base_off, n, fp32, a, c)
            for _ in range(3):
                fn()
            t = timed(fn, iters)
            row[f"{name}_us"] = round(t * 1e3, 2)
            row[f"{name}_busbw_gbs"] = round(factor / t, 1)
        if ns.nccl:
            ref_t = nccl_buf[:n]

            def nccl():
                dist.all_reduce(ref_t)
                ref_t.div_(world)

            for _ in range(3):
                nccl()
            t = timed(nccl, iters)
            row["nccl_us"] = round(t * 1e3, 2)
            row["nccl_busbw_gbs"] = round(factor / t, 1)
        emit(row)

    if ns.p2p and world > 1:
        nbytes = min(256 << 20, (ns.max_mb << 20) // 2)
        peer = (rank + 1) % world
        local = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        remote_ptr = seg.peer_ptrs[peer] + base_off

        def ce_pull():
            native.debug_memcpy_async(local.data_ptr(), remote_ptr, nbytes)

        def ce_push():
            native.debug_memcpy_async(remote_ptr, local.data_ptr(), nbytes)

        for name, fn in (("copy_engine_pull", ce_pull), ("copy_engine_push", ce_push)):
            for _ in range(2):
                fn()
            t = timed(fn, 10)
            emit({"p2p": name, "bytes": nbytes, "gbs_per_direction": round(nbytes / 1e6 / t, 1), "world": world,
                  "note": "every rank copies from/to rank+1 at the same time"})
        # SM-issued peer loads: the broadcast / all-gather kernels give the one-direction inbound rates
        n16 = nbytes // world // 16 * 16
        fn = lambda: native.comm_broadcast(seg.desc, base_off, nbytes, 0, ns.ctas)
        for _ in range(2):
            fn()
        t = timed(fn, 10)
        emit({"p2p": "broadcast", "bytes": nbytes, "gbs_inbound_per_gpu": round(nbytes / 1e6 / t, 1), "world": world,
              "nvls": comm.nvls_active})
        so, fo = base_off, base_off + nbytes          # shard = head of the buffer, gathered copy = its second half
        fn = lambda: native.comm_all_gather(seg.desc, fo, so, n16, 1, ns.ctas)
        for _ in range(2):
            fn()
        t = timed(fn, 10)
        emit({"p2p": "all_gather", "bytes_out": n16 * world,
              "gbs_inbound_per_gpu": round(n16 * (world - 1) / 1e6 / t, 1), "world": world, "nvls": comm.nvls_active})
    comm.check_errors()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
