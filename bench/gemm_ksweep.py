"""Fixed M x N, growing K: time(K) = overhead + epilogue tail + K * slope.  Separates per-launch / per-tile
costs from mainloop throughput for the small-K GEMMs of the GPT-2 configs."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402

BF = torch.bfloat16
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)


def timed(fn, iters=12):
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


M, N = 8192, 2304
for K in (64, 256, 768, 1536, 3072, 6144):
    a = torch.randn(M, K, device="cuda").to(BF)
    b = torch.randn(N, K, device="cuda").to(BF)
    bias = torch.randn(N, device="cuda").to(BF)
    out = torch.empty(M, N, device="cuda", dtype=BF)
    pre = torch.empty(M, N, device="cuda", dtype=BF)
    ref = torch.empty(M, N, device="cuda", dtype=BF)
    rec = {"M": M, "N": N, "K": K}
    for name, kw in (("cg2_bn256", dict(cta_pair=2, tile_n=256)), ("cg1_bn256", dict(cta_pair=1, tile_n=256)),
                     ("cg1_bn128", dict(cta_pair=1, tile_n=128))):
        rec[name] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, **kw)), 1)
    rec["cg2_bias"] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, bias=bias, cta_pair=2, tile_n=256)), 1)
    rec["cg2_bias_gelu_d2"] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, bias=bias, act=1, d2=pre,
                                                              cta_pair=2, tile_n=256)), 1)
    rec["cublas"] = round(timed(lambda: torch.mm(a, b.t(), out=ref)), 1)
    rec["cublas_bias"] = round(timed(lambda: torch.addmm(bias, a, b.t(), out=ref)), 1)
    best = min(rec["cg2_bn256"], rec["cg1_bn256"], rec["cg1_bn128"])
    rec["tflops_best_ours"] = round(2.0 * M * N * K / best / 1e6, 1)
    rec["tflops_cublas"] = round(2.0 * M * N * K / rec["cublas"] / 1e6, 1)
    print(json.dumps(rec), flush=True)
