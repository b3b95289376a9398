"""GEMM timings on the GPT-2 small layer shapes (R = 8192 tokens), with the epilogues the model uses."""
import json, os, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native

def t(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3

R, C = 8192, 768
bf = dict(dtype=torch.bfloat16, device="cuda")
rows = []
for name, (m, n, k) in {"qkv": (R, 3*C, C), "proj": (R, C, C), "fc": (R, 4*C, C), "proj2": (R, C, 4*C)}.items():
    x, w, b = torch.randn(m, k, **bf), torch.randn(n, k, **bf) * 0.02, torch.randn(n, **bf)
    out, z, res = torch.empty(m, n, **bf), torch.empty(m, n, **bf), torch.randn(m, n, **bf)
    dy, dx = torch.randn(m, n, **bf), torch.empty(m, k, **bf)
    dw, pre = torch.empty(n, k, **bf), torch.randn(m, k, **bf)
    r = {"layer": name, "mode": os.environ.get("DTB_GEMM_DIRECT_EPILOGUE", "0"),
         "fwd_plain": t(lambda: native.gemm(x, w, out, m=m, n=n, k=k)),
         "fwd_bias": t(lambda: native.gemm(x, w, out, m=m, n=n, k=k, bias=b)),
         "fwd_bias_res": t(lambda: native.gemm(x, w, out, m=m, n=n, k=k, bias=b, aux=res)),
         "fwd_bias_gelu_z": t(lambda: native.gemm(x, w, out, m=m, n=n, k=k, bias=b, act=1, d2=z)),
         "dgrad": t(lambda: native.gemm(dy, w, dx, m=m, n=k, k=n, b_mn=True)),
         "dgrad_dgelu": t(lambda: native.gemm(dy, w, dx, m=m, n=k, k=n, b_mn=True, aux=pre, act=2)),
         "wgrad": t(lambda: native.gemm(dy, x, dw, m=n, n=k, k=m, a_mn=True, b_mn=True)),
         "cublas_fwd": t(lambda: torch.addmm(b, x, w.t(), out=out)),
         "cublas_dgrad": t(lambda: torch.matmul(dy, w, out=dx)),
         "cublas_wgrad": t(lambda: torch.matmul(dy.t(), x, out=dw))}
    print(json.dumps({k_: (round(v, 1) if isinstance(v, float) else v) for k_, v in r.items()}), flush=True)
