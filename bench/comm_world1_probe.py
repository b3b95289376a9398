"""One-GPU launches of the communication kernels and the all-gather GEMM, for ncu.

ncu replays every kernel many times, so it cannot profile a kernel that waits for other ranks; with world size 1 the
same kernels run their complete local path (flag barrier on their own pad, loads, reduction arithmetic, stores) and
can be captured:

    ncu --set full --clock-control none --import-source on -k regex:'allreduce|reduce_scatter|gemm_kernel' \
        -o gpurun_out/comm_world1 python bench/comm_world1_probe.py

What the capture shows is the SM side of the kernels (instruction mix, registers, issue rate, local HBM traffic);
NVLink counters stay at zero, the multi-GPU rates are in profiles/allreduce_sweep_*_r2.jsonl.
"""
from __future__ import annotations

import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402
from distributed_training_b200.parallel.symm import SymmComm  # noqa: E402


def main() -> None:
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    nbytes = 128 << 20
    comm = SymmComm(dev, segment_bytes=4 * nbytes, overlap_stream=False)
    n = nbytes // 2
    grads = comm.alloc(n, torch.bfloat16)
    params = comm.alloc(n, torch.bfloat16)
    grads.normal_()
    params.normal_()
    seg, g_off = comm.locate(grads)
    _, p_off = comm.locate(params)
    master = params.float()
    m, v = torch.zeros_like(master), torch.zeros_like(master)
    for _ in range(2):
        native.comm_all_reduce_avg(seg.desc, g_off, n, False, 2, 0)                          # peer-load two-shot
        native.comm_all_reduce_sgd(seg.desc, g_off, p_off, n, 1e-3, 0.0, 0)                  # + fused SGD
        native.comm_all_reduce_sgd(seg.desc, g_off, p_off, n, 1e-3, 0.0, 4 | native.EXCLUSIVE)   # whole-SM CTAs
        native.comm_rs_optim(seg.desc, g_off, p_off, master, m, v, n, 1, 1e-3, 0.9, 0.999, 1e-8, 0.1, 0.1, 0.001,
                             0.0, 0)                                                         # reduce-scatter + AdamW
        native.comm_broadcast(seg.desc, g_off, nbytes, 0, 0)
    torch.cuda.synchronize()
    comm.check_errors()
    # all-gather fused into the GEMM that consumes the gathered weight (GPT-2 XL fc: 4096 x 6400 x 1600)
    mm, nn, kk = 4096, 6400, 1600
    x = torch.randn(mm, kk, device=dev).to(torch.bfloat16)
    shard = (torch.randn(nn, kk, device=dev) * 0.02).to(torch.bfloat16)
    full = torch.empty_like(shard)
    out = torch.empty(mm, nn, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(native.gemm_allgather_flag_ints(nn), dtype=torch.int32, device=dev)
    for _ in range(2):
        native.gemm_allgather(x, full, out, [shard.data_ptr()], flags, 0, nn, m=mm, n=nn, k=kk)
    torch.cuda.synchronize()
    assert torch.equal(full, shard)
    comm.close()
    print("ok")


if __name__ == "__main__":
    main()
