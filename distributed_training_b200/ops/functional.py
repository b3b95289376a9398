"""Autograd-visible wrappers over the sm_100a kernels.

These make the kernels usable from ordinary ``nn.Module`` code (the generic tier of the framework:
arbitrary models under ``ReplicatedModule`` / ``ShardedModule`` with autograd), in contrast to the
hand-scheduled ``forward_backward`` steps of the built-in models.  Every function dispatches on the
input: CUDA tensors of a supported dtype go to the library (never to a PyTorch fallback — a missing
library on a GPU machine raises), CPU tensors use the PyTorch reference implementation so the same
model code runs in the gloo plumbing tests.

Reference call sites replaced: ``self.model(source)`` → ``addmm`` (distributed_trainer.py:162),
``F.cross_entropy`` (:163), ``nn.MSELoss`` (playground/ddp_script.py:135).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from . import native


def _use_native(*tensors: torch.Tensor) -> bool:
    if not all(t is None or t.is_cuda for t in tensors):
        return False
    native.require()
    return True


class _LinearBF16(torch.autograd.Function):
    """y = x W^T + b on tcgen05; backward = dgrad + wgrad GEMMs and a column-sum."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]).contiguous()
        m, k = x2.shape
        n = weight.shape[0]
        out = torch.empty(m, n, dtype=torch.bfloat16, device=x.device)
        native.gemm(x2, weight, out, m=m, n=n, k=k, bias=bias)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.in_shape = shape
        return out.view(*shape[:-1], n)

    @staticmethod
    def backward(ctx, gy):
        x2, weight = ctx.saved_tensors
        m, k = x2.shape
        n = weight.shape[0]
        gy2 = gy.reshape(m, n).contiguous()
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = torch.empty(m, k, dtype=torch.bfloat16, device=gy.device)
            native.gemm(gy2, weight, gx, m=m, n=k, k=n, b_mn=True)
            gx = gx.view(ctx.in_shape)
        if ctx.needs_input_grad[1]:
            gw = torch.empty(n, k, dtype=torch.bfloat16, device=gy.device)
            native.gemm(gy2, x2, gw, m=n, n=k, k=m, a_mn=True, b_mn=True)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = torch.empty(n, dtype=torch.bfloat16, device=gy.device)
            native.colsum_split(gy2, gb)
        return gx, gw, gb


class _LinearSmallF32(torch.autograd.Function):
    """fp32 SIMT Linear for shapes far below a tensor-core tile (the reference's Linear(20, 1))."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        y = torch.empty(x2.shape[0], weight.shape[0], dtype=torch.float32, device=x.device)
        native.linear_small_fwd(x2, weight.contiguous(), bias, y)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.in_shape = x.shape
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, gy):
        x2, weight = ctx.saved_tensors
        gy2 = gy.reshape(x2.shape[0], weight.shape[0]).contiguous()
        gx = torch.empty_like(x2) if ctx.needs_input_grad[0] else None
        gw = torch.empty_like(weight)
        gb = torch.empty(weight.shape[0], dtype=torch.float32, device=gy.device) if ctx.has_bias else None
        native.linear_small_bwd(x2, weight.contiguous(), gy2, gx, gw, gb)
        return (gx.view(ctx.in_shape) if gx is not None else None), gw, gb


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not _use_native(x, weight, bias):
        return F.linear(x, weight, bias)
    n, k = weight.shape
    if x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and k % 8 == 0 and n % 8 == 0:
        return _LinearBF16.apply(x, weight, bias)
    if x.dtype == torch.float32 and weight.dtype == torch.float32 and n * k <= (1 << 16):
        return _LinearSmallF32.apply(x, weight, bias)
    raise NotImplementedError(
        f"no native Linear for dtype {x.dtype}/{weight.dtype} with weight {tuple(weight.shape)}: "
        "use bf16 with multiples of 8, or fp32 with at most 65536 weights")


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xc = x.contiguous()
        y = torch.empty_like(xc)
        native.gelu_fwd(xc, y)
        ctx.save_for_backward(xc)
        return y

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gx = torch.empty_like(x)
        native.gelu_bwd(x, gy.contiguous(), gx)
        return gx


def gelu(x: torch.Tensor) -> torch.Tensor:
    """tanh-approximated GELU."""
    if _use_native(x) and x.dtype == torch.bfloat16:
        return _Gelu.apply(x)
    return F.gelu(x, approximate="tanh")


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]).contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        native.layernorm_fwd(x2, None, weight, bias, y, None, mean, rstd, eps)
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.in_shape = shape
        return y.view(shape)

    @staticmethod
    def backward(ctx, gy):
        x2, weight, mean, rstd = ctx.saved_tensors
        rows, cols = x2.shape
        gy2 = gy.reshape(rows, cols).contiguous()
        gx = torch.empty_like(x2)
        dg = torch.empty(cols, dtype=torch.float32, device=gy.device)
        db = torch.empty(cols, dtype=torch.float32, device=gy.device)
        scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, cols), dtype=torch.float32, device=gy.device)
        native.layernorm_bwd(gy2, x2, weight, mean, rstd, None, gx, dg, db, scratch)
        return gx.view(ctx.in_shape), dg.to(weight.dtype), db.to(weight.dtype), None


def layer_norm(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    if _use_native(x, weight, bias) and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and x.shape[-1] <= 2048:
        return _LayerNorm.apply(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


class _CrossEntropy(torch.autograd.Function):
    """Mean softmax cross-entropy with class-index targets; forward already computes dlogits."""

    @staticmethod
    def forward(ctx, logits, targets):
        rows, vocab = logits.shape
        work = logits.detach().clone()
        row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        native.cross_entropy_fwd_bwd(work, targets.contiguous(), row_loss, vocab, 1.0 / rows)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        native.mean_f32(row_loss, loss)
        ctx.save_for_backward(work)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return dlogits * g.to(dlogits.dtype), None


class _SoftCrossEntropy(torch.autograd.Function):
    """``F.cross_entropy`` with float targets of the logits' shape (what the reference's trainer calls)."""

    @staticmethod
    def forward(ctx, logits, targets):
        rows, cols = logits.shape
        row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        d = torch.empty_like(logits)
        native.soft_cross_entropy_fwd_bwd(logits.contiguous(), targets.contiguous(), row_loss, d, 1.0 / rows)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        native.mean_f32(row_loss, loss)
        ctx.save_for_backward(d)
        return loss

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return d * g, None


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    if _use_native(logits, targets):
        if targets.dtype == torch.int64 and logits.dtype == torch.bfloat16 and logits.dim() == 2 \
                and logits.shape[1] % 8 == 0:
            return _CrossEntropy.apply(logits, targets)
        if targets.dtype == torch.float32 and logits.dtype == torch.float32 and logits.dim() == 2:
            return _SoftCrossEntropy.apply(logits, targets)
    if targets.dtype in (torch.int64, torch.int32):
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), targets.view(-1))
    return F.cross_entropy(logits.float(), targets.float())


class _Mse(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        n = pred.numel()
        s = torch.empty((), dtype=torch.float32, device=pred.device)
        d = torch.empty_like(pred)
        native.mse_fwd_bwd(pred.contiguous(), target.contiguous(), s, d, 1.0 / n)
        ctx.save_for_backward(d)
        return s / n

    @staticmethod
    def backward(ctx, g):
        (d,) = ctx.saved_tensors
        return d * g, None


def mse_loss(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    if _use_native(pred, target) and pred.dtype == torch.float32 and target.dtype == torch.float32 \
            and pred.shape == target.shape:
        return _Mse.apply(pred, target)
    return F.mse_loss(pred.float(), target.float())


class NativeLinear(nn.Linear):
    """``nn.Linear`` whose forward/backward run on the library's kernels when the tensors are on a GPU.
    Same parameters and ``state_dict`` keys as ``nn.Linear``, so snapshots stay interchangeable."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D102
        return linear(x, self.weight, self.bias)
