"""FSDP-style fully sharded data parallelism (ZeRO-3), built from collectives.

The reference's ``FSDPStrategy`` wraps the whole model in one FSDP1 flat
parameter (/root/reference/src/dist_strategy/fsdp_strategy.py:20-26), which
gives one all-gather before forward, one before backward and one
reduce-scatter after it — no overlap — and is marked work-in-progress in its
roadmap.  This implementation follows the roadmap's intent:

* the model is cut into *units* (one per transformer block / linear layer);
  every unit's parameters are flattened, padded and split evenly, each rank
  keeping 1/W as its **shard** (bf16/fp32 as the model's dtype);
* before a unit runs forward (and again before its backward when
  ``reshard_after_forward``) the shards are all-gathered into the unit's full
  buffer; module parameters are *views* into that buffer;
* when the last gradient of a unit lands, the unit's flat gradient is
  reduce-scattered (mean) to the owning ranks, overlapping the remaining
  backward; with the fused optimizer the shard update runs right there
  (``dtb_comm_rs_optim`` in ``csrc/comm/collectives.cu`` does reduce-scatter +
  AdamW in one kernel on B200: the native path, ``sharded_native.py``);
* ``state_dict()`` is a collective that returns the full, *unprefixed*
  parameter names (what FSDP's FULL_STATE_DICT produces), so snapshots load
  into a bare model, into the DDP strategy, or into the reference.

Fixes relative to the reference: parameters are broadcast from rank 0 before
sharding (Q8), every rank enters the state-dict gather (Q6), the optimizer is
built on the shards by the strategy (Q7).
"""
from __future__ import annotations

import logging
import math
from collections import OrderedDict
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import nn

from .comm import Communicator, make_communicator
from .strategy import DistributedStrategy, OptimizerSpec, atomic_torch_save

logger = logging.getLogger(__name__)

SHARD_ALIGN_ELEMS = 64  # every rank's shard starts on a 128-byte boundary (bf16)


def default_units(module: nn.Module) -> List[nn.Module]:
    """Sharding units: a model may define ``fsdp_units()``; otherwise every
    parameter-owning child, descending through pure containers."""
    if hasattr(module, "fsdp_units"):
        return list(module.fsdp_units())
    units: List[nn.Module] = []

    def visit(m: nn.Module) -> None:
        own = list(m.parameters(recurse=False))
        children = list(m.children())
        if isinstance(m, (nn.ModuleList, nn.Sequential)) or (not own and children and m is module):
            for c in children:
                visit(c)
        elif own or any(True for _ in m.parameters()):
            units.append(m)

    visit(module)
    seen = {id(p) for u in units for p in u.parameters()}
    if any(id(p) not in seen for p in module.parameters()):
        return [module]  # root owns parameters directly: single unit, like the reference
    return units or [module]


class ShardUnit:
    """One flat-parameter group: shard, gathered full buffer, flat gradient."""

    def __init__(self, name: str, module: nn.Module, params: List[Tuple[str, nn.Parameter]],
                 comm: Communicator) -> None:
        self.name = name
        self.module = module
        self.comm = comm
        self.param_names = [n for n, _ in params]
        self.params = [p for _, p in params]
        self.dtype = self.params[0].dtype
        if any(p.dtype != self.dtype for p in self.params):
            raise ValueError(f"unit {name}: mixed parameter dtypes are not supported")
        self.offsets: List[int] = []
        total = 0
        for p in self.params:
            self.offsets.append(total)
            total += (p.numel() + 7) // 8 * 8          # 16-byte aligned slots
        per_rank = math.ceil(total / comm.world_size)
        per_rank = (per_rank + SHARD_ALIGN_ELEMS - 1) // SHARD_ALIGN_ELEMS * SHARD_ALIGN_ELEMS
        self.shard_numel = per_rank
        self.full_numel = per_rank * comm.world_size
        self.full = comm.alloc(self.full_numel, self.dtype)
        self.grad_full = comm.alloc(self.full_numel, self.dtype)
        self.shard = comm.alloc(per_rank, self.dtype)
        self.grad_shard = torch.zeros(per_rank, dtype=torch.float32, device=self.full.device)
        with torch.no_grad():
            for p, off in zip(self.params, self.offsets):
                self.full[off: off + p.numel()].copy_(p.detach().reshape(-1))
        lo = comm.rank * per_rank
        self.shard.copy_(self.full[lo: lo + per_rank])
        for p, off in zip(self.params, self.offsets):
            p.data = self.full[off: off + p.numel()].view(p.shape)
            p.grad = self.grad_full[off: off + p.numel()].view(p.shape)
        self.gathered = True
        self.pending = len(self.params)
        self.ready: set = set()
        self.reduced = False

    # ---- collectives ----------------------------------------------------
    def gather(self) -> None:
        if not self.gathered:
            self.comm.all_gather_into(self.full, self.shard)
            self.gathered = True

    def release(self) -> None:
        """Mark the full buffer stale (its storage is kept and reused)."""
        self.gathered = False

    def reduce_gradients(self) -> None:
        tmp = self.grad_shard if self.grad_shard.dtype == self.grad_full.dtype else \
            torch.empty(self.shard_numel, dtype=self.grad_full.dtype, device=self.grad_full.device)
        self.comm.reduce_scatter_avg_into(tmp, self.grad_full)
        if tmp is not self.grad_shard:
            self.grad_shard.copy_(tmp)
        self.reduced = True

    def reset(self, zero_grads: bool = True) -> None:
        self.pending = len(self.params)
        self.ready.clear()
        self.reduced = False
        if zero_grads:      # autograd accumulates in place into the attached views
            self.grad_full.zero_()


class ShardedModule(nn.Module):
    """Module whose parameters are sharded across ranks between uses."""

    def __init__(self, module: nn.Module, comm: Communicator, units: Optional[List[nn.Module]] = None,
                 reshard_after_forward: bool = True, init_sync: bool = True) -> None:
        super().__init__()
        self.inner = module          # not named "module": FULL_STATE_DICT keys are unprefixed
        self.comm = comm
        self.reshard_after_forward = reshard_after_forward
        if init_sync:
            comm.broadcast_([p.data for p in module.parameters()] + [b.data for b in module.buffers()], src=0)
        names = {id(p): n for n, p in module.named_parameters()}
        self.units: List[ShardUnit] = []
        claimed: set = set()
        for k, unit_module in enumerate(units if units is not None else default_units(module)):
            params = [(names[id(p)], p) for p in unit_module.parameters()
                      if id(p) not in claimed and p.requires_grad]
            if not params:
                continue
            claimed.update(id(p) for _, p in params)
            unit = ShardUnit(f"unit{k}", unit_module, params, comm)
            self.units.append(unit)
            unit_module.register_forward_pre_hook(self._pre_forward(unit))
            unit_module.register_forward_hook(self._post_forward(unit))
            unit_module.register_full_backward_pre_hook(self._pre_backward(unit))
            for i, p in enumerate(unit.params):
                p.register_post_accumulate_grad_hook(self._grad_hook(unit, i))
        stray = [n for n, p in module.named_parameters() if id(p) not in claimed and p.requires_grad]
        if stray:
            raise ValueError(f"parameters not covered by any sharding unit: {stray[:4]}...")
        self._callback_queued = False
        self.fused_optimizer = None   # set by ShardedOptimizer when it updates inside backward
        self.require_backward_grad_sync = True

    # ---- hooks ------------------------------------------------------------
    def _pre_forward(self, unit: ShardUnit):
        def hook(_m, _args):
            unit.gather()
        return hook

    def _post_forward(self, unit: ShardUnit):
        def hook(_m, _args, _out):
            if self.reshard_after_forward and torch.is_grad_enabled() and len(self.units) > 1:
                unit.release()
        return hook

    def _pre_backward(self, unit: ShardUnit):
        def hook(_m, _grad_out):
            unit.gather()
        return hook

    def _grad_hook(self, unit: ShardUnit, index: int):
        def hook(param: torch.Tensor) -> None:
            off = unit.offsets[index]
            view = unit.grad_full[off: off + param.numel()].view(param.shape)
            if param.grad is not None and param.grad.data_ptr() != view.data_ptr():
                view.copy_(param.grad)
                param.grad = view
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
            if index not in unit.ready:
                unit.ready.add(index)
                unit.pending -= 1
                if unit.pending == 0 and self.require_backward_grad_sync:
                    self._unit_done(unit)
        return hook

    def _unit_done(self, unit: ShardUnit) -> None:
        unit.reduce_gradients()
        unit.release()
        if self.fused_optimizer is not None:
            self.fused_optimizer.update_unit(unit)

    def _finalize(self) -> None:
        self._callback_queued = False
        if not self.require_backward_grad_sync:
            return
        for unit in self.units:
            if not unit.reduced:
                for i, p in enumerate(unit.params):
                    if i not in unit.ready:
                        off = unit.offsets[i]
                        unit.grad_full[off: off + p.numel()].zero_()
                self._unit_done(unit)

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        return self.inner(*args, **kwargs)

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor, **kwargs: Any) -> torch.Tensor:
        """Native models: fused all-gather→GEMM / GEMM→reduce-scatter path."""
        if not hasattr(self.inner, "forward_backward"):
            raise AttributeError(f"{type(self.inner).__name__} has no native forward_backward")
        return self.inner.forward_backward(inputs, targets, shard_ctx=self, **kwargs)

    @property
    def native_step(self) -> bool:
        return hasattr(self.inner, "forward_backward") and getattr(self.inner, "native_ready", lambda: False)() \
            and getattr(self.inner, "supports_sharded_native", False)

    def begin_step(self) -> None:
        for unit in self.units:
            unit.reset()

    # ---- state dict ---------------------------------------------------------
    def gather_all(self) -> None:
        for unit in self.units:
            unit.gathered = False
            unit.gather()

    def state_dict(self, *args: Any, destination=None, prefix: str = "", keep_vars: bool = False):
        """Collective: full unsharded state with the wrapped model's own key names (CPU tensors)."""
        self.gather_all()
        out = OrderedDict() if destination is None else destination
        inner = self.inner.state_dict(prefix=prefix, keep_vars=False)
        for k, v in inner.items():
            out[k] = v.detach().to("cpu", copy=True)
        return out

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        self.gather_all()
        result = self.inner.load_state_dict(state, strict=strict)
        # parameters are views into unit.full, so the load landed there; refresh the shards
        for unit in self.units:
            lo = self.comm.rank * unit.shard_numel
            unit.shard.copy_(unit.full[lo: lo + unit.shard_numel])
        if self.fused_optimizer is not None:
            self.fused_optimizer.refresh_master()
        return result


class ShardedOptimizer:
    """Optimizer over the local shards (fp32 master + moments per shard).

    ``step()`` consumes each unit's reduce-scattered fp32 gradient shard.  With
    ``update_in_backward`` the update of a unit runs as soon as its gradients
    have been reduced, overlapping the rest of backward; ``step()`` then only
    advances the step counter.
    """

    def __init__(self, model: ShardedModule, spec: OptimizerSpec, update_in_backward: bool = False) -> None:
        self.model = model
        self.spec = spec
        self.update_in_backward = update_in_backward
        self.step_count = 0
        self.state: Dict[str, Dict[str, torch.Tensor]] = {}
        for unit in model.units:
            st = {"master": unit.shard.detach().float().clone()}
            if spec.name == "adamw":
                st["exp_avg"] = torch.zeros_like(st["master"])
                st["exp_avg_sq"] = torch.zeros_like(st["master"])
            elif spec.momentum:
                st["momentum_buffer"] = torch.zeros_like(st["master"])
            self.state[unit.name] = st
        self.param_groups = [{"lr": spec.lr, "weight_decay": spec.weight_decay, "params": []}]
        if update_in_backward:
            model.fused_optimizer = self
        self._updated: set = set()

    def refresh_master(self) -> None:
        for unit in self.model.units:
            self.state[unit.name]["master"].copy_(unit.shard.float())

    def zero_grad(self, set_to_none: bool = False) -> None:
        """Gradients are cleared by ``step()``; kept for the torch optimizer call sequence."""
        return None

    @torch.no_grad()
    def update_unit(self, unit: ShardUnit) -> None:
        if unit.name in self._updated:
            return
        self._updated.add(unit.name)
        spec, st = self.spec, self.state[unit.name]
        lr = self.param_groups[0]["lr"]
        wd = self.param_groups[0]["weight_decay"]
        t = self.step_count + 1
        g, w = unit.grad_shard, st["master"]
        from ..ops import native

        if native.available() and w.is_cuda:
            native.shard_update(unit.shard, w, g, st, spec.name, lr, wd, spec.momentum, spec.betas, spec.eps, t)
        elif spec.name == "adamw":
            b1, b2 = spec.betas
            st["exp_avg"].mul_(b1).add_(g, alpha=1 - b1)
            st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
            denom = (st["exp_avg_sq"] / (1 - b2 ** t)).sqrt_().add_(spec.eps)
            w.mul_(1 - lr * wd).addcdiv_(st["exp_avg"], denom, value=-lr / (1 - b1 ** t))
            unit.shard.copy_(w)
        else:
            d = g if wd == 0 else g.add(w, alpha=wd)
            if spec.momentum:
                d = st["momentum_buffer"].mul_(spec.momentum).add_(d)
            w.add_(d, alpha=-lr)
            unit.shard.copy_(w)
        unit.release()

    def step(self, closure=None):
        for unit in self.model.units:
            if unit.reduced:
                self.update_unit(unit)
        self.step_count += 1
        self.model.begin_step()
        self._updated.clear()
        return None

    def state_dict(self) -> Dict[str, Any]:
        """COLLECTIVE (every rank must call it): the unsharded optimizer state -- each unit's fp32 master and
        moments concatenated over ranks in shard order -- on rank 0; other ranks get an empty ``state``.  The file
        can therefore be loaded at any world size (``load_state_dict`` re-slices it)."""
        comm = self.model.comm
        full: Dict[str, Dict[str, torch.Tensor]] = {}
        for unit in self.model.units:
            for n, t in self.state[unit.name].items():
                parts = comm.gather_to_root(t)
                if parts is not None:
                    full.setdefault(unit.name, {})[n] = torch.cat([p.view(-1) for p in parts])
        return {"step": self.step_count, "format": "full", "world_size": comm.world_size, "state": full}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        """Every rank loads the same full-state dict and keeps its own slice."""
        if sd.get("format") != "full":
            raise ValueError("sharded optimizer state must be a full-state dict written by ShardedOptimizer.state_dict "
                             "(per-rank shard dumps of earlier versions cannot be re-sharded)")
        self.step_count = int(sd["step"])
        rank = self.model.comm.rank
        for unit in self.model.units:
            n = unit.shard_numel
            for name, t in sd["state"].get(unit.name, {}).items():
                dst = self.state[unit.name][name]
                flat = t.view(-1)
                piece = flat[rank * n: (rank + 1) * n]
                dst.zero_()
                dst[: piece.numel()].copy_(piece)


class FSDPStrategy(DistributedStrategy):
    """Sharded strategy; constructor signature as the reference's ``FSDPStrategy(is_gpu)``
    (/root/reference/src/dist_strategy/fsdp_strategy.py:13-15)."""

    collective_state_dict = True

    def __init__(self, is_gpu: bool = False, comm: Optional[Communicator] = None,
                 reshard_after_forward: bool = True) -> None:
        self.is_gpu = is_gpu
        self.comm = comm
        self.reshard_after_forward = reshard_after_forward

    def prepare_model(self, model: nn.Module, device: Any) -> nn.Module:
        if self.comm is None:
            dev = torch.device("cuda", torch.cuda.current_device()) if self.is_gpu else torch.device("cpu")
            self.comm = make_communicator(dev)
        if (getattr(self.comm, "supports_fused_step", False) and getattr(model, "supports_sharded_native", False)
                and model.native_ready()):
            # B200 path: all-gather fused into the forward GEMMs, reduce-scatter fused with the optimizer
            from .sharded_native import NativeShardedModule

            return NativeShardedModule(model, self.comm)
        return ShardedModule(model, self.comm, reshard_after_forward=self.reshard_after_forward)

    def save_checkpoint(self, model: nn.Module, state_dict: dict, path: str) -> None:
        """``state_dict['MODEL_STATE']`` must come from the collective ``model.state_dict()``
        (``ModelCheckpoint.save`` calls it on every rank); only rank 0 writes."""
        if self.comm is None or self.comm.rank == 0:
            atomic_torch_save(state_dict, path)

    def load_checkpoint(self, model: nn.Module, checkpoint: dict, device: Any) -> Tuple[nn.Module, int]:
        model.load_state_dict(checkpoint["MODEL_STATE"])
        return model, int(checkpoint["EPOCHS_RUN"])

    def build_optimizer(self, model: nn.Module, spec: OptimizerSpec) -> Any:
        from .sharded_native import NativeShardedModule, NativeShardedOptimizer

        if isinstance(model, NativeShardedModule):
            return NativeShardedOptimizer(model, spec)
        if not isinstance(model, ShardedModule):
            raise TypeError("FSDPStrategy.build_optimizer expects the module returned by prepare_model")
        return ShardedOptimizer(model, spec, update_in_backward=True)

    def adopt_optimizer(self, model: nn.Module, optimizer: Any) -> Any:
        if isinstance(optimizer, torch.optim.Optimizer):
            return self.build_optimizer(model, OptimizerSpec.from_torch(optimizer))
        return optimizer

    def barrier(self) -> None:
        if self.comm is not None:
            self.comm.barrier()
