"""Tiny-shape pass through every single-GPU kernel family, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck --error-exitcode 9 python bench/sanitize_smoke.py

Shapes are as small as the kernels allow (the sanitizer slows kernels 10-100x) but still hit: cta_group::1 and ::2
GEMMs, K-/MN-major operands, ragged M/N/K tails, split-K, every epilogue, attention fwd/bwd, LayerNorm, both CE
kernels, MSE, embedding, colsum, optimizers.  Results are compared with PyTorch so a silent out-of-bounds read that
happens to be harmless still has to produce the right numbers.
"""
from __future__ import annotations

import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from distributed_training_b200.models.gpt2 import GPT2, GPT2Config  # noqa: E402
from distributed_training_b200.models.mlp import MLP  # noqa: E402
from distributed_training_b200.ops import native  # noqa: E402
from distributed_training_b200.optim import FusedAdamW, FusedSGD  # noqa: E402


def check(name: str, got: torch.Tensor, want: torch.Tensor, tol: float) -> None:
    err = (got.float() - want.float()).abs().max().item()
    scale = want.float().abs().max().item() + 1e-6
    assert err <= tol * scale, f"{name}: max err {err:.4g} vs scale {scale:.4g}"
    print(f"  {name}: rel err {err / scale:.2e}")


def gemms() -> None:
    torch.manual_seed(0)
    dev = "cuda"
    for (m, n, k) in [(128, 128, 64), (300, 200, 136), (512, 512, 256), (256, 384, 1024)]:
        a = torch.randn(m, k, device=dev).bfloat16()
        b = torch.randn(n, k, device=dev).bfloat16()
        bias = torch.randn(n, device=dev).bfloat16()
        out = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        native.gemm(a, b, out, m=m, n=n, k=k, bias=bias)
        check(f"gemm {m}x{n}x{k} +bias", out, a.float() @ b.float().t() + bias.float(), 2e-2)
        # wgrad layout: both operands MN-major (dW[n,k] = dY[m,n]^T X[m,k]); split-K forced on the long one
        dw = torch.empty(n, k, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(m, n, device=dev).bfloat16()
        native.gemm(dy, a, dw, m=n, n=k, k=m, a_mn=True, b_mn=True, split_k=2 if m >= 512 else 0)
        check(f"wgrad {n}x{k}x{m}", dw, dy.float().t() @ a.float(), 2e-2)
        # dgrad layout: B MN-major (dX[m,k] = dY[m,n] W[n,k])
        dx = torch.empty(m, k, device=dev, dtype=torch.bfloat16)
        native.gemm(dy, b, dx, m=m, n=k, k=n, b_mn=True)
        check(f"dgrad {m}x{k}x{n}", dx, dy.float() @ b.float(), 2e-2)


def models() -> None:
    torch.manual_seed(1)
    mlp = MLP(256, 512, 128, num_layers=3, dtype=torch.bfloat16).cuda()
    opt = FusedSGD(list(mlp.parameters()), lr=0.05)
    x = torch.randn(192, 256, device="cuda").bfloat16()
    y = torch.randint(0, 128, (192,), device="cuda")
    losses = []
    for _ in range(3):
        losses.append(float(mlp.forward_backward(x, y)))
        opt.step()
    assert losses[-1] < losses[0], losses
    print(f"  mlp native step: loss {losses[0]:.4f} -> {losses[-1]:.4f}")

    cfg = GPT2Config(vocab_size=1000, n_positions=128, n_embd=128, n_layer=2, n_head=2)
    gpt = GPT2(cfg, dtype=torch.bfloat16).cuda()
    assert gpt.native_ready()
    opt2 = FusedAdamW(list(gpt.parameters()), lr=1e-3)
    tok = torch.randint(0, 1000, (2, 128), device="cuda")
    tgt = torch.randint(0, 1000, (2, 128), device="cuda")
    losses = []
    for _ in range(3):
        losses.append(float(gpt.forward_backward(tok, tgt)))
        opt2.step()
    assert losses[-1] < losses[0], losses
    print(f"  gpt2 native step: loss {losses[0]:.4f} -> {losses[-1]:.4f}")


def main() -> None:
    native.require()
    torch.cuda.set_device(0)
    gemms()
    models()
    torch.cuda.synchronize()
    print(f"sanitize smoke ok: {native.launch_count()} native launches")


if __name__ == "__main__":
    main()
