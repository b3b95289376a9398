"""Sharded (ZeRO-3 / FSDP-style) training on the fused B200 kernels.

This is the native counterpart of :class:`ShardedModule` for models that run their own
``forward_backward`` (the MLP and GPT-2).  Sharding is per parameter, along dim 0:

* every large 2-D weight ``[N, K]`` keeps only ``N / W`` rows per rank (bf16 shard in the symmetric heap,
  fp32 master + optimizer moments for those rows in ordinary memory);
* **FULL_SHARD memory** (what ``FSDP(model)`` defaults to in the reference,
  /root/reference/src/dist_strategy/fsdp_strategy.py:20-26): the unsharded weight and the unsharded gradient exist
  only transiently, in small pools of heap buffers shared by all weights of the same size (``pool_depth`` slots per
  size class), so per-rank memory is shards + optimizer state + a few layers' worth of full matrices;
* **forward**: the first GEMM that consumes a weight gathers it *inside the GEMM kernel*
  (``dtb_gemm_bf16_allgather``: gather warps pull the peers' shards over NVLink piece by piece while the
  tensor cores already work on tiles whose rows have landed) into a pool slot, which the next weight of that size
  overwrites (resharding after forward is implicit: nothing is kept);
* **backward**: the weight is re-gathered for its dgrad (FSDP's BACKWARD_PRE) by the stand-alone all-gather kernel
  (one ``multimem.st`` per vector on NVLS) on a side stream, *prefetched* while the previous layer's backward GEMMs
  run (the backward order is learnt in the first step); the wgrad GEMM writes the full ``dW`` into a pooled
  symmetric gradient buffer and ``dtb_comm_rs_optim`` reduce-scatters it (mean), applies SGD/AdamW to the fp32
  master rows and writes the new bf16 shard in one kernel on the communication stream, overlapping the rest of
  backward; the slot is reused once that kernel has finished (it ends with a rank barrier, so every owner has read it);
* small parameters (biases, LayerNorm) stay replicated: their gradients share one flat heap buffer that is
  all-reduced once at the end of backward and stepped by the fused multi-tensor optimizer.

Reference: ``FSDPStrategy`` wraps torch FSDP1 (/root/reference/src/dist_strategy/fsdp_strategy.py:20-26) with
one flat parameter, i.e. all-gather → compute → reduce-scatter with no overlap and (Q7) an optimizer that
points at dead parameters; here the strategy owns shards and optimizer state.
"""
from __future__ import annotations

import logging
import os
from collections import OrderedDict
from typing import Any, Dict, List, Optional

import torch
from torch import nn

from ..ops import native
from .strategy import OptimizerSpec

logger = logging.getLogger(__name__)

SHARD_ROW_MULTIPLE = 16          # gather pieces are 16 rows
MIN_SHARDED_NUMEL = 1 << 16      # smaller tensors stay replicated


def padded_rows_per_rank(rows: int, world: int) -> int:
    per = (rows + world - 1) // world
    return (per + SHARD_ROW_MULTIPLE - 1) // SHARD_ROW_MULTIPLE * SHARD_ROW_MULTIPLE


class _Slot:
    """One pooled heap buffer (a full weight or a full gradient of one size class)."""

    def __init__(self, buf: torch.Tensor, offset: int) -> None:
        self.buf = buf
        self.offset = offset                 # byte offset in the symmetric segment (identical on every rank)
        self.owner: Optional["_ShardedWeight"] = None
        self.free_event: Optional[torch.cuda.Event] = None    # recorded when the last asynchronous user finished
        self.write_event: Optional[torch.cuda.Event] = None   # last side-stream gather that wrote this slot
        self.stamp = 0                       # last-use counter (LRU choice inside a class)


class _ShardedWeight:
    """Book-keeping for one row-sharded matrix: the persistent state is the bf16 row shard (symmetric heap) and
    the fp32 master / moments of those rows; ``full`` / ``grad`` point at pool slots only while they are live."""

    def __init__(self, name: str, param: nn.Parameter, comm: Any, spec_name: str, segment: Any) -> None:
        n, k = param.shape
        w, r = comm.world_size, comm.rank
        self.name, self.param = name, param
        # every rank holds the same number of rows, a multiple of the 16-row gather piece; when N does not divide
        # (GPT-2 XL at 8 ranks: 4800 -> 608 rows/rank, 1600 -> 208) the last rank's shard ends in zero padding rows
        self.rows, self.cols, self.rows_per_rank = n, k, padded_rows_per_rank(n, w)
        self.padded_rows = self.rows_per_rank * w
        numel_shard = self.rows_per_rank * k
        self.shard = comm.alloc(numel_shard, torch.bfloat16).view(self.rows_per_rank, k)
        with torch.no_grad():
            mine = param.data[r * self.rows_per_rank:(r + 1) * self.rows_per_rank]
            self.shard.zero_()
            self.shard[: mine.shape[0]].copy_(mine)
        self.master = self.shard.float().clone()
        self.m = torch.zeros_like(self.master) if spec_name == "adamw" else None
        self.v = torch.zeros_like(self.master) if spec_name == "adamw" else None
        seg, off = comm.locate(self.shard)
        if seg is not segment:
            raise RuntimeError(f"{name}: shard landed in symmetric segment {seg.index}, expected {segment.index} "
                               "(the fused kernels address shard, full weight and gradient by offset in ONE segment)")
        self.segment = seg
        self.shard_offset = off
        self.peer_shard_ptrs = [seg.peer_ptrs[q] + off for q in range(w)]
        self.flags = torch.zeros(native.gemm_allgather_flag_ints(n), dtype=torch.int32, device=param.device)
        self.full_slot: Optional[_Slot] = None      # pool slot that currently holds every rank's rows, if any
        self.grad_slot: Optional[_Slot] = None
        self.updated = False
        self.gather_args = None                     # prebuilt GatherArgs per pool slot

    @property
    def gathered(self) -> bool:
        return self.full_slot is not None and self.full_slot.owner is self

    @property
    def full(self) -> torch.Tensor:
        assert self.gathered, f"{self.name}: full weight is not resident"
        return self.full_slot.buf[: self.rows * self.cols].view(self.rows, self.cols)

    @property
    def grad(self) -> torch.Tensor:
        assert self.grad_slot is not None and self.grad_slot.owner is self, f"{self.name}: no gradient buffer"
        return self.grad_slot.buf[: self.rows * self.cols].view(self.rows, self.cols)


class NativeShardedModule(nn.Module):
    """Native-step model with row-sharded weights; also the model's ``grad_sink``.

    Protocol the native models follow: ``gemm_forward(x, W, out, ...)`` for a Linear forward (gathers ``W`` inside
    the GEMM), ``weight_view(W)`` wherever else the full matrix is read (dgrad, embedding lookup, LM head),
    ``grad_view(W)`` for the buffer a wgrad kernel writes, ``mark_ready(W)`` once that gradient is complete.
    """

    native_step = True

    def __init__(self, module: nn.Module, comm: Any, spec: Optional[OptimizerSpec] = None, pool_depth: int = 3,
                 reshard_after_forward: bool = True, prefetch_ahead: int = 2) -> None:
        super().__init__()
        if not getattr(module, "supports_sharded_native", False) or not module.native_ready():
            raise TypeError(f"{type(module).__name__} has no native sharded path")
        self.inner = module
        self.comm = comm
        if hasattr(module, "recompute_gelu") and os.environ.get("DTB_RECOMPUTE_GELU", "1") != "0":
            module.recompute_gelu = True      # the sharded strategy is chosen for memory: trade ~2 % time for 25 % of
            module._ws = {}                   # the saved activations (DTB_RECOMPUTE_GELU=0 keeps gelu(z) per block)
        self.spec = spec or OptimizerSpec("adamw", lr=1e-4)
        self.step_count = 0
        self.reshard_after_forward = reshard_after_forward
        self.prefetch_ahead = max(int(prefetch_ahead), 0)
        w = comm.world_size
        named = list(module.named_parameters())
        big = [(n, p) for n, p in named if self._shardable(p, w)]
        small = [(n, p) for n, p in named if not self._shardable(p, w)]
        classes: Dict[int, int] = {}        # pool size class = padded element count of the full matrix
        for _, p in big:
            key = padded_rows_per_rank(p.shape[0], w) * w * p.shape[1]
            classes[key] = classes.get(key, 0) + 1
        depth = {numel: (min(max(pool_depth, 1), count) if reshard_after_forward else count)
                 for numel, count in classes.items()}
        # one segment for everything the fused kernels address by offset: shards, full-weight and gradient pools,
        # the flat buffer of the small replicated gradients
        need = sum(padded_rows_per_rank(p.shape[0], w) * p.shape[1] * 2 + 512 for _, p in big)
        need += sum(2 * d * (numel * 2 + 512) for numel, d in depth.items())
        need += sum(p.numel() * 2 + 512 for _, p in small) + (1 << 20)
        # the init broadcast's staging slab comes out of the same segment (reserved first, so nothing else is mapped)
        staging = min(256 << 20, sum(p.numel() * p.element_size() for p in module.parameters()) + (1 << 20))
        segment = comm.reserve(need + (staging if w > 1 else 0) + (1 << 20))
        comm.broadcast_([p.data for p in module.parameters()] + [b.data for b in module.buffers()], src=0)
        if segment.nbytes - segment.used < need:
            segment = comm.reserve(need)
        self.segment = segment
        self.weights: "OrderedDict[int, _ShardedWeight]" = OrderedDict()
        for n, p in big:
            sw = _ShardedWeight(n, p, comm, self.spec.name, segment)
            if self.spec.name == "sgd" and self.spec.momentum:
                sw.m = torch.zeros_like(sw.master)
            self.weights[id(p)] = sw
        self.full_pool: Dict[int, List[_Slot]] = {}
        self.grad_pool: Dict[int, List[_Slot]] = {}
        for pool in (self.full_pool, self.grad_pool):
            for numel, d in depth.items():
                for _ in range(d):
                    buf = comm.alloc(numel, torch.bfloat16)
                    seg, off = comm.locate(buf)
                    if seg is not segment:
                        raise RuntimeError("sharded pools must share the shards' symmetric segment")
                    buf.zero_()       # padding rows past N stay zero for good: no kernel writes them with data
                    pool.setdefault(numel, []).append(_Slot(buf, off))
        # the full-size original storage goes away: the parameter keeps its shape on a one-element placeholder
        for sw in self.weights.values():
            sw.param.data = torch.empty(1, dtype=torch.bfloat16, device=sw.param.device).expand(sw.rows, sw.cols)
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
        # stand-alone gathers (backward re-gather, checkpoints) all go through ONE side stream on barrier channel 1,
        # so every rank issues them in the same order and they never share flags with the reduce-scatters (channel 0)
        self.gather_stream = torch.cuda.Stream(comm.device, priority=-1)
        self._access: List[tuple] = []          # (weight id, kind) in the order of this step's accesses
        self._plan: List[tuple] = []            # the first step's access order: what to prefetch next
        self._cursor = 0
        self._gather_events: Dict[int, torch.cuda.Event] = {}   # weight id -> gather in flight on the side stream
        pooled = sum(len(v) * k * 2 for k, v in self.full_pool.items()) * 2
        logger.info(f"[rank {comm.rank}] native sharded module: {len(self.weights)} sharded matrices "
                    f"({sum(s.rows * s.cols for s in self.weights.values()) / 1e6:.1f} M params), "
                    f"{len(self.small_params)} replicated small tensors; full/grad pools {pooled / 2**20:.0f} MiB "
                    f"({'FULL_SHARD' if reshard_after_forward else 'weights kept resident'})")

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
                and p.shape[0] >= SHARD_ROW_MULTIPLE * world and p.shape[1] % 8 == 0)

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        """Autograd / inference path of the wrapped model: materialises every full weight in ordinary memory for
        the call (evaluation convenience -- training goes through ``forward_backward``)."""
        saved = []
        for sw in self.weights.values():
            saved.append((sw, sw.param.data))
            sw.param.data = self.weight_view(sw.param).clone()
        try:
            return self.inner(*args, **kwargs)
        finally:
            for sw, placeholder in saved:
                sw.param.data = placeholder

    # ------------------------------------------------------------ pools
    def _note(self, sw: _ShardedWeight, kind: str) -> int:
        """Record this access (first step) / advance the cursor over the learnt plan; returns the access index."""
        i = self._cursor
        self._cursor += 1
        self._access.append((id(sw.param), kind))
        return i

    def _take_slot(self, pool: Dict[int, List[_Slot]], sw: _ShardedWeight, index: int,
                   for_index: int) -> Optional[_Slot]:
        """A slot of ``sw``'s size class whose content is no longer needed: unowned first, else the one whose
        (planned) use lies furthest in the past.  Slots stamped at or after ``index`` are in use / prefetched for
        later and are not taken (returns None then, unless this is the access that needs it now)."""
        slots = pool[sw.padded_rows * sw.cols]
        free = [s for s in slots if s.owner is None]
        if free:
            slot = free[0]
        else:
            old = [s for s in slots if s.stamp < index]
            if not old:
                if for_index > index:
                    return None                    # a prefetch never steals from nearer uses
                old = slots                        # needed now: steal the slot planned furthest ahead
                slot = max(old, key=lambda s: s.stamp)
            else:
                slot = min(old, key=lambda s: s.stamp)
        if slot.owner is not None and slot.owner is not sw:
            victim = slot.owner
            if pool is self.full_pool:
                victim.full_slot = None
                self._gather_events.pop(id(victim.param), None)
            else:
                victim.grad_slot = None
        slot.owner = sw
        slot.stamp = for_index
        return slot

    def _launch_gather(self, sw: _ShardedWeight, slot: _Slot) -> None:
        """Stand-alone all-gather of ``sw`` into ``slot`` on the side stream, ordered after everything queued on the
        compute stream so far (the slot's previous readers) -- locally; the kernel's entry barrier extends that to
        the peers, which store into this rank's slot."""
        gs = self.gather_stream
        gs.wait_stream(torch.cuda.current_stream())
        native.comm_all_gather(sw.segment.desc, slot.offset, sw.shard_offset, sw.shard.numel(), 2,
                               self.comm.comm_ctas, channel=1, stream=gs.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(gs)
        self._gather_events[id(sw.param)] = ev
        slot.write_event = ev
        sw.full_slot = slot

    def _prefetch(self, index: int) -> None:
        """Start the gathers of the next ``prefetch_ahead`` planned ``view`` accesses that are not resident."""
        if self.prefetch_ahead == 0 or index >= len(self._plan) or self._plan[index] != self._access[index]:
            return                  # no plan yet, or this step deviates from it: gather on demand
        for j in range(index + 1, min(index + 1 + self.prefetch_ahead, len(self._plan))):
            pid, kind = self._plan[j]
            nxt = self.weights[pid]
            if kind != "view" or nxt.gathered:
                if nxt.gathered:
                    nxt.full_slot.stamp = max(nxt.full_slot.stamp, j)
                continue
            slot = self._take_slot(self.full_pool, nxt, index, j)
            if slot is None:
                break
            self._launch_gather(nxt, slot)

    # ------------------------------------------------------------ gather side
    def weight_view(self, param: nn.Parameter) -> torch.Tensor:
        """The full ``[N, K]`` matrix of ``param``, resident in a pool slot (gathered on demand, or already
        prefetched); plain parameters are returned unchanged.  Valid until a later access of the same size class
        needs the slot: consume it with kernels queued right away."""
        sw = self.weights.get(id(param))
        if sw is None:
            return param
        i = self._note(sw, "view")
        if not sw.gathered:
            slot = self._take_slot(self.full_pool, sw, i, i)
            self._launch_gather(sw, slot)
        else:
            sw.full_slot.stamp = i
        ev = self._gather_events.pop(id(param), None)
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)
        self._prefetch(i)
        return sw.full

    def ensure_gathered(self, param: nn.Parameter) -> None:
        self.weight_view(param)

    def gather_all(self) -> None:
        """Only meaningful when weights are kept resident (``reshard_after_forward=False``)."""
        for sw in self.weights.values():
            self.weight_view(sw.param)

    def gemm_forward(self, x: torch.Tensor, weight: nn.Parameter, out: torch.Tensor, *, m: int, n: int, k: int,
                     **kw: Any) -> torch.Tensor:
        """``out = epilogue(x W^T)``; a sharded ``W`` that is not resident is gathered inside the GEMM."""
        sw = self.weights.get(id(weight))
        if sw is None:
            return native.gemm(x, weight, out, m=m, n=n, k=k, **kw)
        i = self._note(sw, "fused")
        if sw.gathered:
            sw.full_slot.stamp = i
            ev = self._gather_events.pop(id(weight), None)
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)
            self._prefetch(i)
            return native.gemm(x, sw.full, out, m=m, n=n, k=k, **kw)
        slot = self._take_slot(self.full_pool, sw, i, i)
        sw.full_slot = slot
        if slot.write_event is not None:      # a side-stream gather into this slot must not still be writing it
            torch.cuda.current_stream().wait_event(slot.write_event)
            slot.write_event = None
        if sw.gather_args is None:
            sw.gather_args = {}
        ga = sw.gather_args.get(id(slot))
        if ga is None:      # constant per (weight, slot): peer shard pointers, flag scratch, geometry
            ga = sw.gather_args[id(slot)] = native.make_gather_args(sw.peer_shard_ptrs, sw.flags, self.comm.rank,
                                                                    sw.rows_per_rank)
        native.gemm_allgather(x, sw.full, out, sw.peer_shard_ptrs, sw.flags, self.comm.rank, sw.rows_per_rank,
                              m=m, n=n, k=k, bias=kw.get("bias"), aux=kw.get("aux"), act=kw.get("act", 0),
                              d2=kw.get("d2"), beta=kw.get("beta", 1.0), max_ctas=kw.get("max_ctas", 0),
                              gather_args=ga)
        self._prefetch(i)
        return out

    # -------------------------------------------------------------- grad side
    def grad_view(self, param: nn.Parameter) -> torch.Tensor:
        sw = self.weights.get(id(param))
        if sw is None:
            return self.small_views[id(param)]
        if sw.grad_slot is None or sw.grad_slot.owner is not sw:
            slot = self._take_slot(self.grad_pool, sw, self._cursor, self._cursor)
            if slot.free_event is not None:       # the reduce-scatter that last read this slot must have finished
                torch.cuda.current_stream().wait_event(slot.free_event)
                slot.free_event = None
            sw.grad_slot = slot
        sw.grad_slot.stamp = self._cursor
        return sw.grad

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
        slot = sw.grad_slot
        assert slot is not None and slot.owner is sw, f"{sw.name}: mark_ready without a gradient buffer"
        stream = comm.stream if comm.stream is not None else torch.cuda.current_stream()
        stream.wait_stream(torch.cuda.current_stream())
        # launched straight onto the communication stream (no torch.cuda.stream context: ~10 us of host time each,
        # 200 of these per GPT-2 XL step)
        native.comm_rs_optim(sw.segment.desc, slot.offset, sw.shard_offset, sw.master, sw.m, sw.v,
                             sw.shard.numel(), 1 if adam else 0, spec.lr, b1, b2, spec.eps, spec.weight_decay,
                             1.0 - b1 ** t, 1.0 - b2 ** t, spec.momentum, comm.comm_ctas, stream=stream.cuda_stream)
        ev = torch.cuda.Event()
        ev.record(stream)
        self._pending_events.append(ev)
        slot.free_event = ev            # the kernel ends with a rank barrier: every owner has read this buffer by then
        slot.owner = None
        sw.grad_slot = None
        sw.updated = True
        if sw.full_slot is not None and sw.full_slot.owner is sw:     # the shards changed: any resident copy is stale
            sw.full_slot.owner = None
        sw.full_slot = None
        self._gather_events.pop(id(param), None)

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor, **kwargs: Any) -> torch.Tensor:
        loss = kwargs.pop("loss", None)
        if loss not in (None, "cross_entropy"):
            raise ValueError(f"{type(self.inner).__name__}'s native step trains with cross_entropy; got loss={loss!r}")
        for sw in self.weights.values():
            sw.updated = False
        self._cursor = 0
        for pool in (self.full_pool, self.grad_pool):       # stamps are access indices of ONE step
            for slots in pool.values():
                for s in slots:
                    s.stamp = -1
        self._access = []
        loss_t = self.inner.forward_backward(inputs, targets, grad_sink=self, **kwargs)
        self.finish_step()
        return loss_t

    def finish_step(self) -> None:
        """Replicated small tensors: one all-reduce + fused step; then join the communication stream."""
        for sw in self.weights.values():
            if not sw.updated:          # weight that received no gradient this step
                self.grad_view(sw.param).zero_()
                self.mark_ready(sw.param)
        if self.small_params:
            self.comm.all_reduce_avg_(self.small_flat).wait()
            self.small_optimizer.step()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        for ev in self._pending_events:
            torch.cuda.current_stream().wait_event(ev)
        torch.cuda.current_stream().wait_stream(self.gather_stream)
        stop.record()
        self._stall_events = (start, stop)
        self._pending_events.clear()
        self._gather_events.clear()
        if self._access != self._plan:
            self._plan = list(self._access)      # steps walk the weights in the same order: next step prefetches by it
        self._access = []
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
        """Collective: full, unprefixed state (FULL_STATE_DICT layout) as CPU tensors.  Sharded weights are
        gathered one at a time through the pool, so the call needs no more device memory than a training step."""
        out = OrderedDict() if destination is None else destination
        plan, self._plan = self._plan, []                  # no prefetching outside the step schedule
        try:
            params = dict(self.inner.named_parameters())
            for k, v in self.inner.state_dict(prefix=prefix, keep_vars=True).items():
                p = params.get(k[len(prefix):])
                if p is not None and id(p) in self.weights:
                    full = self.weight_view(p)
                    out[k] = full.detach().to("cpu", copy=True)        # synchronises the current stream
                else:
                    out[k] = v.detach().to("cpu", copy=True)
        finally:
            self._plan = plan
            self._access = []
        return out

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        state = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        own = dict(self.inner.named_parameters())
        own.update(dict(self.inner.named_buffers()))
        missing = [k for k in own if k not in state]
        unexpected = [k for k in state if k not in own]
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing keys {missing[:4]}, unexpected keys {unexpected[:4]}")
        r = self.comm.rank
        with torch.no_grad():
            for k, v in state.items():
                t = own.get(k)
                if t is None:
                    continue
                sw = self.weights.get(id(t))
                if sw is None:
                    t.data.copy_(v)
                    continue
                rows = v[r * sw.rows_per_rank:(r + 1) * sw.rows_per_rank]
                sw.shard.zero_()
                sw.shard[: rows.shape[0]].copy_(rows)
                sw.master.copy_(sw.shard.float())
                if sw.full_slot is not None and sw.full_slot.owner is sw:
                    sw.full_slot.owner = None
                sw.full_slot = None
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)


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
                    full.setdefault(sw.name, {})[key] = torch.cat(parts, dim=0)[: sw.rows]     # drop padding rows
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
            for dst, key in ((sw.master, "master"), (sw.m, "exp_avg"), (sw.v, "exp_avg_sq")):
                if dst is not None and key in st:
                    piece = st[key][rows]
                    dst.zero_()
                    dst[: piece.shape[0]].copy_(piece)
        if sd.get("small") is not None and m.small_optimizer is not None:
            m.small_optimizer.load_state_dict(sd["small"])
