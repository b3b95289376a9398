"""Fused multi-tensor optimizers.

The reference steps ``torch.optim.SGD(lr)`` (/root/reference/src/distributed_trainer.py:200,
/root/reference/src/playground/ddp_script.py:134), which on GPU is a chain of
``_foreach_*`` kernels.  ``FusedSGD`` / ``FusedAdamW`` update every parameter of
a group with one launch of the sm_100a multi-tensor kernel
(``csrc/ops/optim.cu``) when the tensors are CUDA, and fall back to the same
math with ``torch._foreach`` on CPU so the gloo plumbing path and the unit
tests exercise identical semantics.

bf16 parameters can carry fp32 master weights (``master_weights=True``): the
kernel reads the gradient (bf16 or fp32), updates the fp32 master and moments
and writes the rounded bf16 parameter in the same pass.
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional

import torch

from .parallel.strategy import OptimizerSpec


def _native():
    from .ops import native

    return native if native.available() else None


class _FusedBase(torch.optim.Optimizer):
    def __init__(self, params: Iterable[torch.Tensor], defaults: Dict[str, Any],
                 master_weights: bool = False) -> None:
        super().__init__(params, defaults)
        self.master_weights = master_weights
        self._step_count = 0
        self._already_updated: set = set()

    def mark_updated(self, params) -> None:
        """Parameters whose update for the current step was applied by a fused kernel (gradient
        all-reduce + SGD, or the wgrad GEMM epilogue); the next ``step()`` leaves them alone."""
        self._already_updated.update(id(p) for p in params)

    def _pending(self, group) -> List[torch.Tensor]:
        return [p for p in group["params"] if p.grad is not None and id(p) not in self._already_updated]

    @torch.no_grad()
    def step_params(self, params) -> None:
        """Apply THIS step's update to ``params`` now, on the current stream; the following ``step()`` skips them.

        The data-parallel wrapper calls this per gradient bucket right behind the bucket's all-reduce, so the
        optimizer runs under the rest of backward instead of after it (the reference steps the whole model after
        ``loss.backward()``, /root/reference/src/distributed_trainer.py:164-165)."""
        wanted = {id(p) for p in params if p.grad is not None} - self._already_updated
        if not wanted:
            return
        for group in self.param_groups:
            chosen = [p for p in group["params"] if id(p) in wanted]
            if chosen:
                self._apply(group, chosen, self._step_count + 1)
        self._already_updated.update(wanted)

    def _apply(self, group, params: List[torch.Tensor], t: int) -> None:  # pragma: no cover - interface
        raise NotImplementedError

    def state_dict(self) -> Dict[str, Any]:
        sd = super().state_dict()
        sd["fused_step_count"] = self._step_count      # AdamW bias correction must continue, not restart at t = 1
        return sd

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        state_dict = dict(state_dict)
        self._step_count = int(state_dict.pop("fused_step_count", self._step_count))
        super().load_state_dict(state_dict)

    def zero_grad(self, set_to_none: bool = False) -> None:  # keep bucket views attached
        if set_to_none:
            return super().zero_grad(set_to_none=True)
        grads = [p.grad for g in self.param_groups for p in g["params"] if p.grad is not None]
        if grads:
            torch._foreach_zero_(grads)

    def _master(self, p: torch.Tensor) -> Optional[torch.Tensor]:
        if not self.master_weights or p.dtype == torch.float32:
            return None
        st = self.state[p]
        if "master" not in st:
            st["master"] = p.detach().float().clone()
        return st["master"]


class FusedSGD(_FusedBase):
    """``p -= lr * (g + wd * p)`` with optional momentum, one launch per group."""

    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, weight_decay: float = 0.0,
                 master_weights: bool = False) -> None:
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), master_weights)

    @property
    def fusable_sgd(self) -> bool:
        """Stateless update (no momentum, no master copy): can be folded into a reduction / GEMM epilogue."""
        return not self.master_weights and all(g["momentum"] == 0.0 for g in self.param_groups)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self._step_count += 1
        for group in self.param_groups:
            params = self._pending(group)
            if params:
                self._apply(group, params, self._step_count)
        self._already_updated.clear()
        return loss

    def _apply(self, group, params: List[torch.Tensor], t: int) -> None:
        native = _native()
        lr, mom, wd = group["lr"], group["momentum"], group["weight_decay"]
        grads = [p.grad for p in params]
        if native is not None and params[0].is_cuda:
            bufs = None
            if mom != 0.0:
                bufs = []
                for p in params:
                    st = self.state[p]
                    if "momentum_buffer" not in st:
                        st["momentum_buffer"] = torch.zeros_like(p, dtype=torch.float32)
                    bufs.append(st["momentum_buffer"])
            masters = [self._master(p) for p in params] if self.master_weights else None
            native.multi_tensor_sgd(params, grads, masters, bufs, lr, mom, wd)
            return
        # reference math on CPU
        targets = [self._master(p) if self._master(p) is not None else p for p in params]
        g32 = [g.to(t_.dtype) for g, t_ in zip(grads, targets)]
        if wd != 0.0:
            g32 = torch._foreach_add(g32, targets, alpha=wd)
        if mom != 0.0:
            bufs = []
            for p, g in zip(params, g32):
                st = self.state[p]
                if "momentum_buffer" not in st:
                    st["momentum_buffer"] = torch.zeros_like(g)
                bufs.append(st["momentum_buffer"])
            torch._foreach_mul_(bufs, mom)
            torch._foreach_add_(bufs, g32)
            g32 = bufs
        torch._foreach_add_(targets, g32, alpha=-lr)
        for p, t_ in zip(params, targets):
            if t_ is not p:
                p.copy_(t_)


class FusedAdamW(_FusedBase):
    """Decoupled-weight-decay Adam; fp32 moments, optional fp32 master weights."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, master_weights: bool = True) -> None:
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay),
                         master_weights)

    def _moments(self, p: torch.Tensor):
        st = self.state[p]
        if "exp_avg" not in st:
            st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
            st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
        return st["exp_avg"], st["exp_avg_sq"]

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self._step_count += 1
        for group in self.param_groups:
            params = self._pending(group)
            if params:
                self._apply(group, params, self._step_count)
        self._already_updated.clear()
        return loss

    def _apply(self, group, params: List[torch.Tensor], t: int) -> None:
        native = _native()
        lr, (b1, b2), eps, wd = group["lr"], group["betas"], group["eps"], group["weight_decay"]
        grads = [p.grad for p in params]
        ms, vs = zip(*[self._moments(p) for p in params])
        masters = [self._master(p) for p in params]
        bc1, bc2 = 1.0 - b1 ** t, 1.0 - b2 ** t
        if native is not None and params[0].is_cuda:
            native.multi_tensor_adamw(params, grads, list(ms), list(vs),
                                      masters if self.master_weights else None,
                                      lr, b1, b2, eps, wd, bc1, bc2)
            return
        for p, g, m, v, w in zip(params, grads, ms, vs, masters):
            target = w if w is not None else p
            g32 = g.float()
            m.mul_(b1).add_(g32, alpha=1.0 - b1)
            v.mul_(b2).addcmul_(g32, g32, value=1.0 - b2)
            x = target.float() if target.dtype != torch.float32 else target
            x = x * (1.0 - lr * wd) - (lr / bc1) * m / ((v / bc2).sqrt() + eps)
            target.copy_(x)
            if w is not None:
                p.copy_(w)


def build_fused_optimizer(params: List[torch.Tensor], spec: OptimizerSpec,
                          master_weights: Optional[bool] = None):
    if spec.name == "sgd":
        return FusedSGD(params, lr=spec.lr, momentum=spec.momentum, weight_decay=spec.weight_decay,
                        master_weights=bool(master_weights))
    if spec.name == "adamw":
        return FusedAdamW(params, lr=spec.lr, betas=spec.betas, eps=spec.eps,
                          weight_decay=spec.weight_decay,
                          master_weights=True if master_weights is None else master_weights)
    raise ValueError(f"Unknown optimizer {spec.name!r}; expected 'sgd' or 'adamw'")
