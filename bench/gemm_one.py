"""Run one GEMM configuration a few times (target for `ncu --set full -k regex:gemm_kernel`)."""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_training_b200.ops import native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--m", type=int, default=8192)
ap.add_argument("--n", type=int, default=8192)
ap.add_argument("--k", type=int, default=8192)
ap.add_argument("--cg", type=int, default=2)
ap.add_argument("--bn", type=int, default=256)
ap.add_argument("--amn", type=int, default=0)
ap.add_argument("--bmn", type=int, default=0)
ap.add_argument("--iters", type=int, default=6)
ns = ap.parse_args()
a = torch.randn((ns.k, ns.m) if ns.amn else (ns.m, ns.k), device="cuda").to(torch.bfloat16)
b = torch.randn((ns.k, ns.n) if ns.bmn else (ns.n, ns.k), device="cuda").to(torch.bfloat16)
out = torch.empty(ns.m, ns.n, dtype=torch.bfloat16, device="cuda")
for _ in range(ns.iters):
    native.gemm(a, b, out, m=ns.m, n=ns.n, k=ns.k, a_mn=bool(ns.amn), b_mn=bool(ns.bmn), cta_pair=ns.cg, tile_n=ns.bn)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
