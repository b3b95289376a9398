"""Every single-GPU hot kernel once, at the shape it has in the benchmark configs (wide MLP 8x8192 batch 4096;
GPT-2 small batch 8x1024).  Two uses:

    python bench/kernel_zoo.py                 # CUDA-event time per kernel (L2 flushed before each), JSON lines
    ncu --set full --clock-control none --import-source on --profile-from-start off -o zoo \
        python bench/kernel_zoo.py --ncu

With ``--ncu`` each kernel runs once inside a cudaProfilerStart/Stop window (after an unprofiled warm-up), so the
report contains exactly one instance of every kernel and nothing from torch.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402

BF, F32 = torch.bfloat16, torch.float32


def build_cases():
    dev = "cuda"
    rn = lambda *s: torch.randn(*s, device=dev).to(BF)  # noqa: E731
    cases = []

    # ---- wide MLP (batch 4096, width 8192)
    M = 4096
    D = 8192
    x, w, b = rn(M, D), rn(D, D) * 0.01, rn(D)
    pre, h, dy, dx, dw = torch.empty(M, D, device=dev, dtype=BF), torch.empty(M, D, device=dev, dtype=BF), rn(M, D), \
        torch.empty(M, D, device=dev, dtype=BF), torch.empty(D, D, device=dev, dtype=BF)
    fl = 2.0 * M * D * D
    cases.append(("gemm fwd 4096x8192x8192 +bias+gelu (pre-act saved)", fl, 0,
                  lambda: native.gemm(x, w, h, m=M, n=D, k=D, bias=b, act=1, d2=pre)))
    cases.append(("gemm dgrad 4096x8192x8192 *gelu'(aux)", fl, 0,
                  lambda: native.gemm(dy, w, dx, m=M, n=D, k=D, b_mn=True, aux=pre, act=2)))
    cases.append(("gemm wgrad 8192x8192x4096", fl, 0,
                  lambda: native.gemm(dy, x, dw, m=D, n=D, k=M, a_mn=True, b_mn=True)))
    cases.append(("gemm wgrad 8192x8192x4096 fused SGD (W -= lr dW)", fl, 0,
                  lambda: native.gemm(dy, x, w, m=D, n=D, k=M, a_mn=True, b_mn=True, aux=w, alpha=-1e-6, beta=1.0)))
    db = torch.empty(D, device=dev, dtype=BF)
    bsum = torch.zeros(D, device=dev, dtype=F32)
    cases.append(("gemm dgrad 4096x8192x8192 *gelu'(aux) + column sums (bias grad) in the epilogue", fl, 0,
                  lambda: native.gemm(dy, w, dx, m=M, n=D, k=D, b_mn=True, aux=pre, act=2, colsum=bsum)))
    cases.append(("take fp32 colsum -> bf16 bias grad 8192", 0, D * 10, lambda: native.take_f32_to_bf16(bsum, db)))
    cases.append(("colsum 4096x8192 (bias grad)", 0, M * D * 2, lambda: native.colsum_split(dy, db)))
    logits, tgt, rl = rn(M, D), torch.randint(0, D, (M,), device=dev), torch.empty(M, device=dev, dtype=F32)
    cases.append(("cross-entropy fwd+bwd 4096x8192", 0, 2 * M * D * 2,
                  lambda: native.cross_entropy_fwd_bwd(logits, tgt, rl, D, 1.0 / M)))

    # ---- GPT-2 small (B=8, T=1024, d=768, H=12, V=50304)
    Bt, T, d, H, V = 8, 1024, 768, 12, 50304
    R = Bt * T
    xa, wqkv, bqkv = rn(R, d), rn(3 * d, d) * 0.02, rn(3 * d)
    qkv = torch.empty(Bt, T, 3 * d, device=dev, dtype=BF)
    cases.append(("gemm qkv 8192x2304x768 +bias", 2.0 * R * 3 * d * d, 0,
                  lambda: native.gemm(xa, wqkv, qkv, m=R, n=3 * d, k=d, bias=bqkv)))
    wfc, hfc, prefc = rn(4 * d, d) * 0.02, torch.empty(R, 4 * d, device=dev, dtype=BF), \
        torch.empty(R, 4 * d, device=dev, dtype=BF)
    cases.append(("gemm mlp-fc 8192x3072x768 +bias+gelu", 2.0 * R * 4 * d * d, 0,
                  lambda: native.gemm(xa, wfc, hfc, m=R, n=4 * d, k=d, bias=rn(4 * d), act=1, d2=prefc)))
    wte, lg = rn(V, d) * 0.02, torch.empty(R, V, device=dev, dtype=BF)
    cases.append(("gemm lm-head 8192x50304x768", 2.0 * R * V * d, 0,
                  lambda: native.gemm(xa, wte, lg, m=R, n=V, k=d)))
    dfc, dwfc = rn(R, 4 * d), torch.empty(4 * d, d, device=dev, dtype=BF)
    cases.append(("gemm wgrad 3072x768x8192 (split-K by cost model)", 2.0 * R * 4 * d * d, 0,
                  lambda: native.gemm(dfc, xa, dwfc, m=4 * d, n=d, k=R, a_mn=True, b_mn=True)))
    att, lse = torch.empty(Bt, T, d, device=dev, dtype=BF), torch.empty(Bt, H, T, device=dev, dtype=F32)
    qkv_in = rn(Bt, T, 3 * d)
    att_fl = 4.0 * Bt * H * T * T * 64 / 2          # causal: half of the square
    cases.append(("attention fwd B8 H12 T1024 hd64 causal", att_fl, 0,
                  lambda: native.attention_fwd(qkv_in, att, lse, H)))
    datt, dqkv = rn(Bt, T, d), torch.empty(Bt, T, 3 * d, device=dev, dtype=BF)
    scratch = torch.empty(native.attention_bwd_scratch_floats(Bt, T, H), device=dev, dtype=F32)
    native.attention_fwd(qkv_in, att, lse, H)
    cases.append(("attention bwd B8 H12 T1024 hd64 causal", att_fl * 2.5, 0,
                  lambda: native.attention_bwd(qkv_in, att, datt, lse, dqkv, scratch, H)))
    g, be = rn(d), rn(d)
    y, xres = torch.empty(R, d, device=dev, dtype=BF), torch.empty(R, d, device=dev, dtype=BF)
    mean, rstd = torch.empty(R, device=dev, dtype=F32), torch.empty(R, device=dev, dtype=F32)
    res = rn(R, d)
    cases.append(("layernorm fwd +residual 8192x768", 0, 4 * R * d * 2,
                  lambda: native.layernorm_fwd(xa, res, g, be, y, xres, mean, rstd)))
    native.layernorm_fwd(xa, res, g, be, y, xres, mean, rstd)
    dxl, dg, dbe = torch.empty(R, d, device=dev, dtype=BF), torch.empty(d, device=dev, dtype=F32), \
        torch.empty(d, device=dev, dtype=F32)
    lsc = torch.empty(native.layernorm_bwd_scratch_floats(R, d), device=dev, dtype=F32)
    dyl = rn(R, d)
    cases.append(("layernorm bwd 8192x768", 0, 4 * R * d * 2,
                  lambda: native.layernorm_bwd(dyl, xres, g, mean, rstd, res, dxl, dg, dbe, lsc)))
    tg2, rl2 = torch.randint(0, 50257, (R,), device=dev), torch.empty(R, device=dev, dtype=F32)
    cases.append(("cross-entropy fwd+bwd 8192x50304 (in place)", 0, 2 * R * V * 2,
                  lambda: native.cross_entropy_fwd_bwd(lg, tg2, rl2, 50257, 1.0 / R)))
    tok, wpe, emb = torch.randint(0, 50257, (Bt, T), device=dev), rn(T, d), torch.empty(Bt, T, d, device=dev, dtype=BF)
    cases.append(("embedding fwd 8x1024x768", 0, 2 * R * d * 2, lambda: native.embedding_fwd(tok, wte, wpe, emb)))

    # ---- optimizers over GPT-2-small-sized state (124M params, bf16 params + fp32 master/m/v)
    n = 124_000_000
    p, gr = torch.zeros(n, device=dev, dtype=BF), torch.zeros(n, device=dev, dtype=BF)
    ma, m1, v1 = torch.zeros(n, device=dev, dtype=F32), torch.zeros(n, device=dev, dtype=F32), \
        torch.zeros(n, device=dev, dtype=F32)
    cases.append(("AdamW multi-tensor 124M (bf16 p,g; fp32 master,m,v)", 0, n * (2 + 2 + 3 * 8),
                  lambda: native.multi_tensor_adamw([p], [gr], [m1], [v1], [ma], 1e-3, 0.9, 0.999, 1e-8, 0.01,
                                                    0.1, 0.001)))
    cases.append(("SGD multi-tensor 124M (bf16 p,g)", 0, n * 6,
                  lambda: native.multi_tensor_sgd([p], [gr], None, None, 0.01, 0.0, 0.0)))
    return cases


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ncu", action="store_true")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="",
                    help="comma-separated substrings: run only the kernels whose name matches one")
    ns = ap.parse_args()
    native.require()
    peaks = json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text()) \
        if (Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").exists() else {}
    cases = build_cases()
    if ns.only:
        keys = [k.strip().lower() for k in ns.only.split(",") if k.strip()]
        cases = [c for c in cases if any(k in c[0].lower() for k in keys)]
    for _, _, _, fn in cases:          # warm-up (descriptor caches, first-touch)
        fn()
    torch.cuda.synchronize()
    if ns.ncu:
        torch.cuda.profiler.start()
        for name, _, _, fn in cases:
            fn()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        print("profiled", len(cases), "cases")
        return
    flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
    for name, flops, nbytes, fn in cases:
        ts = []
        for _ in range(ns.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        us = ts[len(ts) // 2]
        rec = {"kernel": name, "us_median": round(us, 1), "us_min": round(ts[0], 1)}
        if flops:
            rec["tflops"] = round(flops / us / 1e6, 1)
        if nbytes:
            rec["gb_per_s"] = round(nbytes / us / 1e3, 1)
        print(json.dumps(rec), flush=True)
    print(json.dumps({"peaks": peaks}))


if __name__ == "__main__":
    main()
