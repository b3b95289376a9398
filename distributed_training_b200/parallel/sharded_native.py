"""Sharded (ZeRO-3 / FSDP-style) training on the fused B200 kernels.

This is the native counterpart of :class:`ShardedModule` for models that run their own
``forward_backward`` (the MLP and GPT-2).  Sharding is per parameter, along dim 0:

* every large 2-D weight ``[N, K]`` keeps only ``N / W`` rows per rank (bf16 shard in the symmetric heap,
  fp32 master + optimizer moments for those rows in ordinary memory);
* **forward**: the first GEMM that consumes a weight gathers it *inside the GEMM kernel*
  (``dtb_gemm_bf16_allgather``: gather warps pull the peers' shards over NVLink piece by piece while the
  tensor cores already work on tiles whose rows have landed); the gathered matrix stays resident for the
  backward dgrad;
* **backward**: the wgrad GEMM writes the full ``dW`` into a symmetric gradient buffer and
  ``dtb_comm_rs_optim`` reduce-scatters it (mean, fixed rank order), applies SGD/AdamW to the fp32 master rows
  and writes the new bf16 shard in one kernel on the communication stream, overlapping the rest of backward;
* small parameters (biases, LayerNorm) stay replicated: their gradients share one flat heap buffer that is
  all-reduced once at the end of backward and stepped by the fused multi-tensor optimizer.

Reference: ``FSDPStrategy`` wraps torch FSDP1 (/root/reference/src/dist_strategy/fsdp_strategy.py:20-26) with
one flat parameter, i.e. all-gather → compute → reduce-scatter with no overlap and (Q7) an optimizer that
points at dead parameters; here the strategy owns shards and optimizer state.
"""
from __future__ import annotations

import logging
from collections import OrderedDict
from typing import Any, Dict, List, Optional

import torch
from torch import nn

from ..ops import native
from .strategy import OptimizerSpec

logger = logging.getLogger(__name__)

SHARD_ROW_MULTIPLE = 16          # gather pieces are 16 rows
MIN_SHARDED_NUMEL = 1 << 16      # smaller tensors stay replicated


class _ShardedWeight:
    """Book-keeping for one row-sharded matrix."""

    def __init__(self, name: str, param: nn.Parameter, comm: Any, spec_name: str) -> None:
        n, k = param.shape
        w, r = comm.world_size, comm.rank
        self.name, self.param = name, param
        self.rows, self.cols, self.rows_per_rank = n, k, n // w
        numel_shard = self.rows_per_rank * k
        self.shard = comm.alloc(numel_shard, torch.bfloat16).view(self.rows_per_rank, k)
        self.full = comm.alloc(n * k, torch.bfloat16).view(n, k)
        self.grad = comm.alloc(n * k, torch.bfloat16).view(n, k)
        with torch.no_grad():
            self.full.copy_(param.data)
            self.shard.copy_(self.full[r * self.rows_per_rank:(r + 1) * self.rows_per_rank])
        param.data = self.full                      # the module sees the gathered matrix
        self.master = self.shard.float().clone()
        self.m = torch.zeros_like(self.master) if spec_name == "adamw" else None
        self.v = torch.zeros_like(self.master) if spec_name == "adamw" else None
        seg, off = comm.locate(self.shard)
        self.segment = seg
        self.shard_offset = off
        self.grad_offset = comm.locate(self.grad)[1]
        self.full_offset = comm.locate(self.full)[1]
        self.peer_shard_ptrs = [seg.peer_ptrs[q] + off for q in range(w)]
        self.flags = torch.zeros(native.gemm_allgather_flag_ints(n), dtype=torch.int32, device=param.device)
        self.gathered = True                        # `full` currently holds every rank's rows
        self.updated = False
        self.gather_args = None                     # prebuilt GatherArgs (first fused gather builds it)


class NativeShardedModule(nn.Module):
    """Native-step model with row-sharded weights; also the model's ``grad_sink``."""

    native_step = True

    def __init__(self, module: nn.Module, comm: Any, spec: Optional[OptimizerSpec] = None) -> None:
        super().__init__()
        if not getattr(module, "supports_sharded_native", False) or not module.native_ready():
            raise TypeError(f"{type(module).__name__} has no native sharded path")
        self.inner = module
        self.comm = comm
        self.spec = spec or OptimizerSpec("adamw", lr=1e-4)
        self.step_count = 0
        w = comm.world_size
        comm.broadcast_([p.data for p in module.parameters()] + [b.data for b in module.buffers()], src=0)

        named = list(module.named_parameters())
        big = [(n, p) for n, p in named if self._shardable(p, w)]
        small = [(n, p) for n, p in named if not self._shardable(p, w)]
        # one segment for everything the fused kernels address by offset
        need = sum(p.numel() * 2 * 2 + p.numel() * 2 // w + 1024 for _, p in big)
        need += sum(p.numel() * 2 + 512 for _, p in small) + (1 << 20)
        comm.reserve(need)
        self.weights: "OrderedDict[int, _ShardedWeight]" = OrderedDict()
        for n, p in big:
            self.weights[id(p)] = _ShardedWeight(n, p, comm, self.spec.name)
            if self.spec.name == "sgd" and self.spec.momentum:
                sw = self.weights[id(p)]
                sw.m = torch.zeros_like(sw.master)
        # replicated small parameters: gradients live back to back in one heap buffer
        self.small_params: List[nn.Parameter] = [p for _, p in small]
        offsets, total = [], 0
        for p in self.small_params:
            offsets.append(total)
            total += (p.numel() + 63) // 64 * 64
        self.small_flat = comm.alloc(max(total, 64), torch.bfloat16)
        self.small_views: Dict[int, torch.Tensor] = {}
        for p, off in zip(self.small_params, offsets):
            view = self.small_flat[off: off + p.numel()].view(p.shape)
            self.small_views[id(p)] = view
            p.grad = view
        from ..optim import build_fused_optimizer

        self.small_optimizer = build_fused_optimizer(self.small_params, self.spec, master_weights=False) \
            if self.small_params else None
        self._pending_events: List[torch.cuda.Event] = []
        self.last_exposed_comm_ms = 0.0
        self._stall_events = None
        logger.info(f"[rank {comm.rank}] native sharded module: {len(self.weights)} sharded matrices "
                    f"({sum(s.rows * s.cols for s in self.weights.values()) / 1e6:.1f} M params), "
                    f"{len(self.small_params)} replicated small tensors")

    def reset_optimizer(self, spec: OptimizerSpec) -> None:
        """(Re)create optimizer state for ``spec`` (called when the optimizer is built after wrapping)."""
        from ..optim import build_fused_optimizer

        self.spec = spec
        adam = spec.name == "adamw"
        for sw in self.weights.values():
            sw.m = torch.zeros_like(sw.master) if (adam or spec.momentum) else None
            sw.v = torch.zeros_like(sw.master) if adam else None
        self.small_optimizer = build_fused_optimizer(self.small_params, spec, master_weights=False) \
            if self.small_params else None

    @staticmethod
    def _shardable(p: torch.Tensor, world: int) -> bool:
        return (p.dim() == 2 and p.dtype == torch.bfloat16 and p.numel() >= MIN_SHARDED_NUMEL
                and p.shape[0] % (SHARD_ROW_MULTIPLE * world) == 0 and p.shape[1] % 8 == 0)

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        self.gather_all()
        return self.inner(*args, **kwargs)

    # ------------------------------------------------------------ gather side
    def gather_all(self) -> None:
        for sw in self.weights.values():
            self.ensure_gathered(sw.param)

    def ensure_gathered(self, param: nn.Parameter) -> None:
        """Stand-alone all-gather (embedding tables, checkpoints): peer-store kernel into ``full``."""
        sw = self.weights.get(id(param))
        if sw is None or sw.gathered:
            return
        # channel 1: this runs on the compute stream and must not share barrier slots with the
        # reduce-scatter kernels of the communication stream (channel 0)
        native.comm_all_gather(sw.segment.desc, sw.full_offset, sw.shard_offset, sw.shard.numel(), 2,
                               self.comm.comm_ctas, channel=1)
        sw.gathered = True

    def gemm_forward(self, x: torch.Tensor, weight: nn.Parameter, out: torch.Tensor, *, m: int, n: int, k: int,
                     **kw: Any) -> torch.Tensor:
        """``out = epilogue(x W^T)``; the first use of a sharded ``W`` in a step gathers it inside the GEMM."""
        sw = self.weights.get(id(weight))
        if sw is None or sw.gathered:
            return native.gemm(x, weight, out, m=m, n=n, k=k, **kw)
        if sw.gather_args is None:      # constant per weight: peer shard pointers, flag scratch, geometry
            sw.gather_args = native.make_gather_args(sw.peer_shard_ptrs, sw.flags, self.comm.rank, sw.rows_per_rank)
        native.gemm_allgather(x, sw.full, out, sw.peer_shard_ptrs, sw.flags, self.comm.rank, sw.rows_per_rank,
                              m=m, n=n, k=k, bias=kw.get("bias"), aux=kw.get("aux"), act=kw.get("act", 0),
                              d2=kw.get("d2"), beta=kw.get("beta", 1.0), max_ctas=kw.get("max_ctas", 0),
                              gather_args=sw.gather_args)
        sw.gathered = True
        return out

    # -------------------------------------------------------------- grad side
    def grad_view(self, param: nn.Parameter) -> torch.Tensor:
        sw = self.weights.get(id(param))
        return sw.grad if sw is not None else self.small_views[id(param)]

    def mark_ready(self, param: nn.Parameter) -> None:
        """A sharded weight's full gradient is queued: reduce-scatter + optimizer step on the comm stream."""
        sw = self.weights.get(id(param))
        if sw is None or sw.updated:
            return
        spec = self.spec
        t = self.step_count + 1
        adam = spec.name == "adamw"
        b1, b2 = spec.betas
        comm = self.comm
        stream = comm.stream if comm.stream is not None else torch.cuda.current_stream()
        stream.wait_stream(torch.cuda.current_stream())
        # launched straight onto the communication stream (no torch.cuda.stream context: ~10 us of host time each,
        # 200 of these per GPT-2 XL step)
        native.comm_rs_optim(sw.segment.desc, sw.grad_offset, sw.shard_offset, sw.master, sw.m, sw.v,
                             sw.shard.numel(), 1 if adam else 0, spec.lr, b1, b2, spec.eps, spec.weight_decay,
                             1.0 - b1 ** t, 1.0 - b2 ** t, spec.momentum, comm.comm_ctas, stream=stream.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(stream)
        self._pending_events.append(ev)
        sw.updated = True
        sw.gathered = False            # the shards changed: `full` is stale until the next gather

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor, **kwargs: Any) -> torch.Tensor:
        loss = kwargs.pop("loss", None)
        if loss is not None and "loss" in getattr(self.inner.forward_backward, "__code__", type("c", (), {"co_varnames": ()})).co_varnames:
            kwargs["loss"] = loss          # the model validates it (train.loss reaches the native step)
        elif loss not in (None, "cross_entropy"):
            raise ValueError(f"{type(self.inner).__name__}'s native step trains with cross_entropy; got loss={loss!r}")
        for sw in self.weights.values():
            sw.updated = False
        loss = self.inner.forward_backward(inputs, targets, grad_sink=self, **kwargs)
        self.finish_step()
        return loss

    def finish_step(self) -> None:
        """Replicated small tensors: one all-reduce + fused step; then join the communication stream."""
        for sw in self.weights.values():
            if not sw.updated:          # weight that received no gradient this step
                sw.grad.zero_()
                self.mark_ready(sw.param)
        if self.small_params:
            self.comm.all_reduce_avg_(self.small_flat).wait()
            self.small_optimizer.step()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        for ev in self._pending_events:
            torch.cuda.current_stream().wait_event(ev)
        stop.record()
        self._stall_events = (start, stop)
        self._pending_events.clear()
        self.step_count += 1

    def exposed_comm_ms(self) -> float:
        if self._stall_events is not None:
            start, stop = self._stall_events
            stop.synchronize()
            self.last_exposed_comm_ms = start.elapsed_time(stop)
            self._stall_events = None
        return self.last_exposed_comm_ms

    # ------------------------------------------------------------- state dict
    def state_dict(self, *args: Any, destination=None, prefix: str = "", keep_vars: bool = False):
        """Collective: full, unprefixed state (FULL_STATE_DICT layout) as CPU tensors."""
        self.gather_all()
        torch.cuda.current_stream().synchronize()
        out = OrderedDict() if destination is None else destination
        for k, v in self.inner.state_dict(prefix=prefix).items():
            out[k] = v.detach().to("cpu", copy=True)
        return out

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        result = self.inner.load_state_dict(state, strict=strict)
        r = self.comm.rank
        with torch.no_grad():
            for sw in self.weights.values():
                sw.shard.copy_(sw.full[r * sw.rows_per_rank:(r + 1) * sw.rows_per_rank])
                sw.master.copy_(sw.shard.float())
                sw.gathered = True
        return result


class NativeShardedOptimizer:
    """Optimizer facade for :class:`NativeShardedModule`: the updates themselves run inside backward
    (reduce-scatter + SGD/AdamW kernel per weight, fused multi-tensor step for the small tensors)."""

    def __init__(self, model: NativeShardedModule, spec: OptimizerSpec) -> None:
        self.model = model
        model.reset_optimizer(spec)
        self.param_groups = [{"lr": spec.lr, "weight_decay": spec.weight_decay, "params": []}]

    def zero_grad(self, set_to_none: bool = False) -> None:
        return None

    def step(self, closure=None):
        spec = self.model.spec
        spec.lr = self.param_groups[0]["lr"]          # schedulers edit param_groups
        spec.weight_decay = self.param_groups[0]["weight_decay"]
        if self.model.small_optimizer is not None:
            for g in self.model.small_optimizer.param_groups:
                g["lr"], g["weight_decay"] = spec.lr, spec.weight_decay
        return None

    def state_dict(self) -> Dict[str, Any]:
        """COLLECTIVE (every rank must call it): full optimizer state on rank 0 -- per sharded weight the ``[N, K]``
        fp32 master and moments (row shards concatenated in rank order), plus the replicated small-tensor optimizer's
        state.  Other ranks return an empty ``state``.  ``load_state_dict`` re-slices by rows, so the file resumes at
        any world size that divides the row counts."""
        m = self.model
        full: Dict[str, Dict[str, torch.Tensor]] = {}
        for sw in m.weights.values():
            for key, t in (("master", sw.master), ("exp_avg", sw.m), ("exp_avg_sq", sw.v)):
                if t is None:
                    continue
                parts = m.comm.gather_to_root(t)
                if parts is not None:
                    full.setdefault(sw.name, {})[key] = torch.cat(parts, dim=0)
        small = m.small_optimizer.state_dict() if (m.small_optimizer is not None and m.comm.rank == 0) else None
        return {"step": m.step_count, "format": "full", "world_size": m.comm.world_size, "state": full, "small": small}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        """Every rank loads the same full-state dict and keeps its own rows."""
        if sd.get("format") != "full":
            raise ValueError("sharded optimizer state must be a full-state dict written by "
                             "NativeShardedOptimizer.state_dict (per-rank shard dumps cannot be re-sharded)")
        m = self.model
        m.step_count = int(sd["step"])
        r = m.comm.rank
        by_name = {sw.name: sw for sw in m.weights.values()}
        for name, st in sd["state"].items():
            sw = by_name[name]
            rows = slice(r * sw.rows_per_rank, (r + 1) * sw.rows_per_rank)
            sw.master.copy_(st["master"][rows])
            if sw.m is not None and "exp_avg" in st:
                sw.m.copy_(st["exp_avg"][rows])
            if sw.v is not None and "exp_avg_sq" in st:
                sw.v.copy_(st["exp_avg_sq"][rows])
        if sd.get("small") is not None and m.small_optimizer is not None:
            m.small_optimizer.load_state_dict(sd["small"])
