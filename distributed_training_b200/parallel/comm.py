"""Communicator abstraction used by the strategies.

Two implementations:

* :class:`TorchDistComm` – ``torch.distributed`` collectives.  This is the CPU
  (gloo) path of BASELINE.json config 1 and mirrors what the reference's
  playground does by hand (``dist.broadcast`` per parameter,
  ``dist.all_reduce(SUM)`` + ``/= world_size``;
  /root/reference/src/playground/ddp_script.py:120-121,149-154).
* :class:`SymmComm` (``parallel/symm.py``) – the B200 path: buffers live in a
  symmetric heap mapped into every rank, collectives are single sm_100a
  kernels doing peer loads/stores over NVLink with flag barriers.

Both hand out *communication buffers* (``alloc``); gradient buckets and
sharded parameters are carved from those so a collective never stages through
a copy.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


class Work:
    """Handle for an in-flight collective."""

    def wait(self) -> None:  # pragma: no cover - interface
        raise NotImplementedError


class _DoneWork(Work):
    def wait(self) -> None:
        return None


class _TorchWork(Work):
    def __init__(self, handle, post=None) -> None:
        self._handle = handle
        self._post = post

    def wait(self) -> None:
        if self._handle is not None:
            self._handle.wait()
            self._handle = None
        if self._post is not None:
            self._post()
            self._post = None


class Communicator(ABC):
    rank: int
    world_size: int
    device: torch.device

    @abstractmethod
    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """Zero-initialised 1-D buffer usable as a collective operand."""

    @abstractmethod
    def broadcast_(self, tensors: Sequence[torch.Tensor], src: int = 0) -> None:
        """Overwrite ``tensors`` on every rank with ``src``'s values."""

    @abstractmethod
    def all_reduce_avg_(self, buf: torch.Tensor) -> Work:
        """In-place mean over ranks; asynchronous, ordered after work already
        queued on the caller's current stream."""

    @abstractmethod
    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor) -> None:
        """``out[r*n:(r+1)*n] = shard@rank r`` for every rank ``r``."""

    @abstractmethod
    def reduce_scatter_avg_into(self, out: torch.Tensor, full: torch.Tensor) -> None:
        """``out = mean_r(full@r[rank*n:(rank+1)*n])``."""

    @abstractmethod
    def barrier(self) -> None:
        ...

    # -- conveniences ------------------------------------------------------
    def reserve(self, nbytes: int) -> None:
        """Hint: the next allocations total ``nbytes`` and should be co-located."""
        return None

    def then(self, work: Optional["Work"], fn) -> Optional["Work"]:
        """Run ``fn()`` after ``work`` (a collective handle or None) and return a handle for both.  Stream-ordered
        communicators run it on their communication stream so it overlaps the caller's compute; this default just
        waits and calls."""
        if work is not None:
            work.wait()
        fn()
        return None

    def all_reduce_max_scalar(self, value: float) -> float:
        """Max of a host scalar over ranks (timing: max over ranks)."""
        if self.world_size == 1 or not dist.is_initialized():
            return float(value)
        t = torch.tensor([value], dtype=torch.float64)
        if dist.get_backend() == "nccl":
            t = t.to(self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_to_root(self, t: torch.Tensor, root: int = 0) -> Optional[List[torch.Tensor]]:
        """Host-side gather of equally shaped tensors over the bootstrap process group: rank ``root`` gets the list
        of every rank's tensor (CPU copies), the others ``None``.  Checkpointing only -- never on the step path."""
        if self.world_size == 1 or not dist.is_initialized():
            return [t.detach().to("cpu", copy=True)]
        on_cpu = "gloo" in str(dist.get_backend())
        src = t.detach().contiguous().to("cpu" if on_cpu else self.device)
        out = [torch.empty_like(src) for _ in range(self.world_size)] if self.rank == root else None
        dist.gather(src, out, dst=root)
        return [o.cpu() for o in out] if out is not None else None

    def close(self) -> None:
        return None


class SingleProcessComm(Communicator):
    """World size 1: every collective is the identity."""

    def __init__(self, device: torch.device) -> None:
        self.rank, self.world_size, self.device = 0, 1, torch.device(device)

    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def broadcast_(self, tensors, src: int = 0) -> None:
        return None

    def all_reduce_avg_(self, buf: torch.Tensor) -> Work:
        return _DoneWork()

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor) -> None:
        out.view(-1)[: shard.numel()].copy_(shard.view(-1))

    def reduce_scatter_avg_into(self, out: torch.Tensor, full: torch.Tensor) -> None:
        out.view(-1).copy_(full.view(-1)[: out.numel()])

    def barrier(self) -> None:
        return None


class TorchDistComm(Communicator):
    """Collectives through the default ``torch.distributed`` process group."""

    #: parameters are coalesced into slabs of at most this many bytes for the
    #: init broadcast (torch DDP uses a 250 MiB buffer for the same purpose)
    BROADCAST_SLAB_BYTES = 250 << 20

    def __init__(self, device: torch.device, group: Optional[dist.ProcessGroup] = None) -> None:
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world_size = dist.get_world_size(group)
        self.device = torch.device(device)

    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def broadcast_(self, tensors: Sequence[torch.Tensor], src: int = 0) -> None:
        if self.world_size == 1:
            return
        by_kind = {}
        for t in tensors:
            by_kind.setdefault((t.dtype, t.device), []).append(t)
        for group in by_kind.values():
            slab: List[torch.Tensor] = []
            size = 0
            for t in group + [None]:  # type: ignore[list-item]
                if t is not None and (not slab or size + t.numel() * t.element_size() <= self.BROADCAST_SLAB_BYTES):
                    slab.append(t)
                    size += t.numel() * t.element_size()
                    continue
                if slab:
                    flat = torch.cat([s.detach().reshape(-1) for s in slab])
                    dist.broadcast(flat, src=src, group=self.group)
                    offset = 0
                    with torch.no_grad():
                        for s in slab:
                            s.copy_(flat[offset: offset + s.numel()].view_as(s))
                            offset += s.numel()
                slab, size = ([t], t.numel() * t.element_size()) if t is not None else ([], 0)

    def all_reduce_avg_(self, buf: torch.Tensor) -> Work:
        if self.world_size == 1:
            return _DoneWork()
        world = self.world_size
        # bf16 SUM is not supported by every gloo build: reduce in fp32 and cast back
        if buf.device.type == "cpu" and buf.dtype in (torch.bfloat16, torch.float16):
            tmp = buf.float()
            handle = dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            return _TorchWork(handle, post=lambda: buf.copy_(tmp.div_(world)))
        handle = dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return _TorchWork(handle, post=lambda: buf.div_(world))

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor) -> None:
        if self.world_size == 1:
            out.view(-1)[: shard.numel()].copy_(shard.view(-1))
            return
        n = shard.numel()
        chunks = [out.view(-1)[r * n:(r + 1) * n] for r in range(self.world_size)]
        if out.device.type == "cpu":
            tmp = [torch.empty_like(shard.view(-1)) for _ in range(self.world_size)]
            dist.all_gather(tmp, shard.contiguous().view(-1), group=self.group)
            for c, t in zip(chunks, tmp):
                c.copy_(t)
        else:
            dist.all_gather_into_tensor(out.view(-1)[: n * self.world_size], shard.contiguous().view(-1),
                                        group=self.group)

    def reduce_scatter_avg_into(self, out: torch.Tensor, full: torch.Tensor) -> None:
        n = out.numel()
        if self.world_size == 1:
            out.view(-1).copy_(full.view(-1)[:n])
            return
        if full.device.type == "cpu":
            # gloo has no reduce_scatter: all-reduce then slice (plumbing path only)
            tmp = full.detach().float().clone() if full.dtype != torch.float32 else full.detach().clone()
            dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=self.group)
            out.view(-1).copy_(tmp.view(-1)[self.rank * n:(self.rank + 1) * n].div_(self.world_size))
        else:
            dist.reduce_scatter_tensor(out.view(-1), full.view(-1)[: n * self.world_size].contiguous(),
                                       op=dist.ReduceOp.SUM, group=self.group)
            out.div_(self.world_size)

    def barrier(self) -> None:
        if self.world_size > 1:
            dist.barrier(group=self.group)


def make_communicator(device: torch.device, prefer_native: bool = True) -> Communicator:
    """Pick the communicator for ``device``.

    CUDA -> symmetric-memory kernels (``SymmComm``); CPU -> gloo.  World size
    1 without a process group gets the identity communicator.
    """
    device = torch.device(device)
    if device.type == "cuda" and prefer_native:
        from .symm import SymmComm

        return SymmComm(device)
    if not (dist.is_available() and dist.is_initialized()):
        return SingleProcessComm(device)
    return TorchDistComm(device)
