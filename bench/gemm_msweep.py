"""K = 64 (epilogue only), N fixed, growing M: slope = per-tile epilogue cost, intercept = launch + pipeline fill."""
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


N, K = 2304, 64
for M in (256, 2048, 4096, 8192, 16384, 32768):
    a = torch.randn(M, K, device="cuda").to(BF)
    b = torch.randn(N, K, device="cuda").to(BF)
    out = torch.empty(M, N, device="cuda", dtype=BF)
    ref = torch.empty(M, N, device="cuda", dtype=BF)
    o32 = torch.empty(M, N, device="cuda", dtype=torch.float32) if M <= 16384 else None
    rec = {"M": M, "tiles_per_cluster_cg2": round(((M + 255) // 256) * 9 / 74, 2)}
    rec["cg2"] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, cta_pair=2, tile_n=256)), 1)
    rec["cg1_256"] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, cta_pair=1, tile_n=256)), 1)
    rec["cg1_128"] = round(timed(lambda: native.gemm(a, b, out, m=M, n=N, k=K, cta_pair=1, tile_n=128)), 1)
    rec["cublas"] = round(timed(lambda: torch.mm(a, b.t(), out=ref)), 1)
    rec["memset_same_bytes"] = round(timed(lambda: out.zero_()), 1)
    rec["copy_same_bytes"] = round(timed(lambda: out.copy_(ref)), 1)
    print(json.dumps(rec), flush=True)
