"""How much does a memory-streaming kernel (the all-reduce's traffic) slow a co-resident tcgen05 GEMM?  One GPU.

For each GEMM layout (forward/dgrad-shaped and wgrad-shaped) run a train of GEMMs on the main stream while a side
stream keeps (a) nothing, (b) the stream-probe kernel with C CTAs of 256 threads, (c) copy-engine device-to-device
copies busy.  Reports ms per GEMM and the side stream's achieved GB/s.
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402

BF = torch.bfloat16


def main() -> None:
    native.require()
    dev = "cuda"
    M, D = 4096, 8192
    x, w, dy = (torch.randn(M, D, device=dev).to(BF), torch.randn(D, D, device=dev).to(BF) * 0.01,
                torch.randn(M, D, device=dev).to(BF))
    out, dw = torch.empty(M, D, device=dev, dtype=BF), torch.empty(D, D, device=dev, dtype=BF)
    gemms = {
        "fwd 4096x8192x8192": lambda: native.gemm(x, w, out, m=M, n=D, k=D),
        "dgrad 4096x8192x8192": lambda: native.gemm(dy, w, out, m=M, n=D, k=D, b_mn=True),
        "wgrad 8192x8192x4096": lambda: native.gemm(dy, x, dw, m=D, n=D, k=M, a_mn=True, b_mn=True),
    }
    nbytes = 128 << 20
    bufs = [torch.zeros(nbytes // 2, device=dev, dtype=BF) for _ in range(4)]
    side = torch.cuda.Stream()
    n_gemm = 16

    def side_probe(ctas):
        def run(reps):
            for _ in range(reps):
                native.debug_stream_probe(bufs[0], bufs[1], bufs[2], bufs[3], nbytes, ctas)
        return run, 5 * nbytes

    def side_ce(reps):
        for _ in range(reps):
            bufs[3].copy_(bufs[2], non_blocking=True)      # same dtype, contiguous -> cudaMemcpyAsync D2D

    sides = {"none": (None, 0)}
    for c in (32, 64, 128):
        sides[f"probe {c} CTAs"] = side_probe(c)
    sides["copy engine D2D"] = (side_ce, 2 * nbytes)

    # side-stream throughput alone
    for sname, (fn, bytes_per) in sides.items():
        if fn is None:
            continue
        with torch.cuda.stream(side):
            fn(2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(10)
            e1.record()
        torch.cuda.synchronize()
        print(json.dumps({"side_alone": sname, "ms_per_rep": round(e0.elapsed_time(e1) / 10, 4),
                          "gb_per_s": round(bytes_per * 10 / e0.elapsed_time(e1) / 1e6, 1)}), flush=True)

    for gname, g in gemms.items():
        for _ in range(3):
            g()
        torch.cuda.synchronize()
        for sname, (fn, bytes_per) in sides.items():
            reps = 80
            torch.cuda.synchronize()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if fn is not None:
                with torch.cuda.stream(side):
                    s0.record()
                    fn(reps)
                    s1.record()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n_gemm):
                g()
            e1.record()
            torch.cuda.synchronize()
            rec = {"gemm": gname, "side": sname, "gemm_ms": round(e0.elapsed_time(e1) / n_gemm, 4)}
            if fn is not None:
                # the side train may outlast the GEMMs; its own average rate is still what ran beside them
                rec["side_gb_per_s"] = round(bytes_per * reps / s0.elapsed_time(s1) / 1e6, 1)
                rec["side_total_ms"] = round(s0.elapsed_time(s1), 2)
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
