"""Symmetric-memory communicator for B200 (NVLink 5 / NVSwitch peer access).

Every rank allocates identically sized *segments* of device memory and maps
all peers' segments, so rank ``r``'s buffer at ``(segment, offset)`` has a
known address in every process.  Collectives are single kernels from
``csrc/comm/collectives.cu`` that read and write those peer addresses
directly; ``torch.distributed`` (gloo) only bootstraps (handle exchange,
host-side scalars).

Two heap back ends (``DTB200_HEAP=auto|vmm|ipc``):

* **vmm** (default when the driver supports it and world > 1): segments are
  ``cuMemCreate`` allocations exported as POSIX file descriptors; the fds are
  passed between the ranks over Unix sockets (``SCM_RIGHTS``), imported and
  mapped.  All ranks' pages are additionally bound to one **multicast
  object** whose mapping (``dtb_comm.mc_base``) is the NVLS address: the
  ``multimem.ld_reduce`` / ``multimem.st`` kernels reduce inside the NVSwitch
  and write every replica with one store.
* **ipc**: ``cudaMalloc`` + CUDA IPC handles (single rank, or drivers without
  VMM/fd support).  No multicast mapping; the peer-load/store kernels run.

Replaces ``ProcessGroupNCCL`` as used by the reference
(/root/reference/src/distributed_trainer.py:61-62).

Buffers handed out by :meth:`SymmComm.alloc` are ordinary torch tensors backed
by segment memory (via ``__cuda_array_interface__``), so gradient buckets and
parameter shards can live in them with no staging copies.  Allocation is a
deterministic bump allocator: every rank must call ``alloc`` with the same
sizes in the same order (true for replicas of one model).
"""
from __future__ import annotations

import logging
import os
import socket
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import native
from .comm import Communicator, Work

logger = logging.getLogger(__name__)

ALIGN = 256
DEFAULT_SEGMENT_BYTES = int(os.environ.get("DTB200_SEGMENT_MB", "1024")) << 20


class _RawBuffer:
    """Exposes a raw device range to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int, owner: object) -> None:
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None,
        }


@dataclass
class Segment:
    index: int
    nbytes: int
    local_ptr: int
    peer_ptrs: List[int]
    signal_ptrs: List[int]
    desc: native.CommDesc
    used: int = 0
    base_tensor: Optional[torch.Tensor] = None
    backend: str = "ipc"
    mc_ptr: int = 0                       # NVLS multicast mapping of the segment (0 = none)
    total_bytes: int = 0
    handles: List[int] = field(default_factory=list)      # VMM allocation handles (own + imported)
    mc_handle: int = 0


def _fd_address(rank: int, seq: int) -> str:
    port = os.environ.get("MASTER_PORT", "0")
    return f"\0dtb200-{os.getuid()}-{port}-{seq}-{rank}"       # abstract namespace: nothing to unlink


def exchange_fds(fd: Optional[int], rank: int, world: int, seq: int) -> List[Optional[int]]:
    """All-to-all of one file descriptor per rank over Unix sockets (``SCM_RIGHTS`` -- VMM allocation handles are
    POSIX fds, and this is the one fd transport every container allows).  ``None`` = this rank contributes nothing.
    Returns, for every rank, that rank's descriptor as received in this process (own slot = ``fd``).  Collective over
    the default (bootstrap) process group; ``seq`` must be the same on all ranks and fresh per call."""
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(_fd_address(rank, seq))
    server.listen(world)
    dist.barrier()                                   # every server is listening
    clients = {}
    for q in range(world):
        if q != rank:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(_fd_address(q, seq))           # queues in q's backlog; q accepts below
            clients[q] = c
    for _ in range(world - 1):
        conn, _addr = server.accept()
        with conn:
            if fd is None:
                conn.sendall(struct.pack("i", 0))
            else:
                socket.send_fds(conn, [struct.pack("i", 1)], [fd])
    out: List[Optional[int]] = [None] * world
    out[rank] = fd
    for q, c in clients.items():
        with c:
            msg, fds, _flags, _addr = socket.recv_fds(c, 4, 1)
            has = struct.unpack("i", msg)[0] if len(msg) == 4 else 0
            out[q] = fds[0] if (has and fds) else None
    server.close()
    dist.barrier()
    return out


class _StreamWork(Work):
    """Completion = an event on the communication stream; ``wait`` orders the caller's stream after it."""

    def __init__(self, event: "torch.cuda.Event") -> None:
        self.event = event

    def wait(self) -> None:
        torch.cuda.current_stream().wait_event(self.event)


class SymmComm(Communicator):
    """Peer-memory communicator; one per process (one process per GPU)."""

    #: parameters may be re-homed into the heap so reduction + optimizer step fuse (``all_reduce_sgd_``)
    supports_fused_step = True

    def __init__(self, device: torch.device, segment_bytes: int = DEFAULT_SEGMENT_BYTES,
                 comm_ctas: int = 0, overlap_stream: bool = True) -> None:
        native.require()
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        if dist.is_available() and dist.is_initialized():
            self.rank, self.world_size = dist.get_rank(), dist.get_world_size()
        else:
            self.rank, self.world_size = 0, 1
        if self.world_size not in (1, 2, 4, 8):
            raise ValueError(f"SymmComm supports 1, 2, 4 or 8 ranks per node, got {self.world_size}")
        self.segment_bytes = int(segment_bytes)
        # CTA budget of a collective (0 = library default of 64): with 2 ranks each rank moves half of every
        # buffer, so more CTAs are needed to keep enough bytes in flight
        self.comm_ctas = comm_ctas if comm_ctas else (128 if self.world_size == 2 else 0)
        #: > 0: while a native step whose backward GEMMs leave that many SMs free is running (ReplicatedModule sets
        #: ``exclusive_now``), large bf16 all-reduces run there on whole-SM CTAs (clusters of two) at their stand-alone
        #: speed instead of sharing SMs with GEMM CTAs.  DTB_COMM_EXCLUSIVE_SMS overrides (0 = always co-resident);
        #: the default is resolved with the first segment (``_default_exclusive_sms``): the NVLS kernels issue one
        #: multimem.ld_reduce + one multimem.st per 16 bytes, so 4 SMs carry the reduction at link rate and the GEMMs
        #: keep 144 (8 GPUs, wide MLP: 10.22 ms/step with 4 SMs, 10.64 with 8, 10.79 with 16, 10.95 co-resident); the
        #: peer-load kernels need 16 (4 GPUs, round 1: 10.33 ms vs 11.27 co-resident).
        env_excl = os.environ.get("DTB_COMM_EXCLUSIVE_SMS")
        self.exclusive_sms = int(env_excl) if env_excl is not None else (16 if self.world_size > 1 else 0)
        self._exclusive_from_env = env_excl is not None
        self.exclusive_min_bytes = 8 << 20
        self.exclusive_now = False     # set by the data-parallel wrapper while a step that left SMs free is running
        #: > 0: SM count that replaces ``exclusive_sms`` for the reductions at the end of backward (set by the wrapper)
        self.exclusive_boost = 0
        #: opt-in (DTB_COMM_CE=1): large bf16 all-reduces move their NVLink legs with the copy engines
        #: (dtb_comm_all_reduce_ce).  Local DMA traffic barely slows a GEMM, but measured at 2 GPUs the peer copies ran
        #: at ~1/3 of their solo rate beside the GEMMs and the step was 0.4 ms slower than with co-resident comm CTAs
        self.copy_engine = os.environ.get("DTB_COMM_CE", "0") == "1"
        self.copy_engine_min_bytes = 16 << 20
        self._ce_staging: List[torch.Tensor] = []
        self.segments: List[Segment] = []
        self._owner_of: Dict[int, Tuple[Segment, int]] = {}   # storage data_ptr -> (segment, offset)
        self.stream = torch.cuda.Stream(self.device, priority=-1) if overlap_stream else None
        self._closed = False
        self._scratch: Optional[torch.Tensor] = None
        self._preferred: Optional[Segment] = None
        self._fd_seq = 0
        self.backend: Optional[str] = None          # "vmm" | "ipc", decided collectively with the first segment
        self.multicast = False
        #: spin-watchdog budget of every collective, in SM clocks (0 = library default, ~20 s); DTB_COMM_TIMEOUT_MS
        self.timeout_cycles = int(float(os.environ.get("DTB_COMM_TIMEOUT_MS", "0")) * 1.9e6)

    # ------------------------------------------------------------------ heap
    def _exchange(self, payload: bytes) -> List[bytes]:
        if self.world_size == 1:
            return [payload]
        gathered: List[Optional[bytes]] = [None] * self.world_size
        dist.all_gather_object(gathered, payload)
        return [g for g in gathered]  # type: ignore[misc]

    def _exchange_fds(self, fd: Optional[int]) -> List[Optional[int]]:
        self._fd_seq += 1
        return exchange_fds(fd, self.rank, self.world_size, self._fd_seq)

    def _all_ok(self, ok: bool) -> bool:
        if self.world_size == 1:
            return ok
        flags: List[Optional[bool]] = [None] * self.world_size
        dist.all_gather_object(flags, bool(ok))
        return all(bool(f) for f in flags)

    def _pick_backend(self) -> str:
        want = os.environ.get("DTB200_HEAP", "auto").lower()
        if want not in ("auto", "vmm", "ipc"):
            raise ValueError(f"DTB200_HEAP must be auto, vmm or ipc, got {want!r}")
        if want == "ipc" or (want == "auto" and self.world_size == 1):
            self.multicast = False
            return "ipc"
        vmm_ok, mc_ok = native.symm_vmm_probe()
        vmm_all = self._all_ok(vmm_ok)
        self.multicast = self._all_ok(mc_ok) and self.world_size > 1 and os.environ.get("DTB200_MULTICAST", "1") != "0"
        if not vmm_all:
            if want == "vmm":
                raise RuntimeError("DTB200_HEAP=vmm but the driver reports no VMM / fd-handle support")
            logger.warning("driver reports no VMM / POSIX-fd handle support: symmetric heap falls back to CUDA IPC")
            self.multicast = False
            return "ipc"
        return "vmm"

    def _new_segment(self, min_bytes: int) -> Segment:
        if self.backend is None:
            self.backend = self._pick_backend()
        nbytes = max(self.segment_bytes, (min_bytes + ALIGN - 1) // ALIGN * ALIGN)
        if self.backend == "vmm":
            seg = self._new_segment_vmm(nbytes)
        else:
            seg = self._new_segment_ipc(nbytes)
        self.segments.append(seg)
        if len(self.segments) == 1 and not self._exclusive_from_env and self.nvls_active:
            self.exclusive_sms = 4
        if self.world_size > 1:
            dist.barrier()     # every rank has mapped every peer before the first kernel touches it
        logger.info(f"[rank {self.rank}] symmetric segment {seg.index}: {seg.nbytes / 2**20:.0f} MiB mapped on "
                    f"{self.world_size} ranks ({seg.backend}{', NVLS multicast' if seg.mc_ptr else ''})")
        return seg

    def _describe(self, index: int, nbytes: int, ptr: int, peer_ptrs: List[int], mc_ptr: int, backend: str) -> Segment:
        desc = native.CommDesc()
        for r in range(self.world_size):
            desc.peers[r] = peer_ptrs[r]
            desc.signals[r] = peer_ptrs[r] + nbytes
        desc.mc_base = mc_ptr if mc_ptr else None
        desc.rank, desc.world, desc.timeout_cycles = self.rank, self.world_size, self.timeout_cycles
        seg = Segment(index, nbytes, ptr, peer_ptrs, [p + nbytes for p in peer_ptrs], desc, backend=backend,
                      mc_ptr=mc_ptr, total_bytes=nbytes + native.SIGNAL_PAD_BYTES)
        seg.base_tensor = torch.as_tensor(_RawBuffer(ptr, nbytes, seg), device=self.device)
        return seg

    def _new_segment_ipc(self, nbytes: int) -> Segment:
        total = nbytes + native.SIGNAL_PAD_BYTES
        ptr, handle = native.symm_alloc(total)       # zero-filled, synchronised
        handles = self._exchange(handle)
        peer_ptrs = []
        for r, h in enumerate(handles):
            peer_ptrs.append(ptr if r == self.rank else native.symm_open(h))
        return self._describe(len(self.segments), nbytes, ptr, peer_ptrs, 0, "ipc")

    def _new_segment_vmm(self, nbytes: int) -> Segment:
        """cuMemCreate + fd exchange + (when supported) one multicast object bound to every rank's pages."""
        w, r = self.world_size, self.rank
        gran = native.symm_vmm_granularity(w, self.multicast)
        total = (nbytes + native.SIGNAL_PAD_BYTES + gran - 1) // gran * gran
        nbytes = total - native.SIGNAL_PAD_BYTES          # the signal pad sits at the end of the rounded-up range
        handle, fd = native.symm_vmm_create(total)
        fds = self._exchange_fds(fd)
        handles, peer_ptrs = [], []
        for q in range(w):
            h = handle if q == r else native.symm_vmm_import(fds[q])
            handles.append(h)
            peer_ptrs.append(native.symm_vmm_map(h, total, gran))
            if q != r and fds[q] is not None:
                os.close(fds[q])
        os.close(fd)
        ptr = peer_ptrs[r]
        whole = torch.as_tensor(_RawBuffer(ptr, total, None), device=self.device)
        whole.zero_()
        torch.cuda.synchronize(self.device)
        dist.barrier()
        mc_ptr, mc_handle = 0, 0
        if self.multicast:
            mc_ptr, mc_handle = self._bind_multicast(handle, total, gran)
        seg = self._describe(len(self.segments), nbytes, ptr, peer_ptrs, mc_ptr, "vmm")
        seg.total_bytes, seg.handles, seg.mc_handle = total, handles, mc_handle
        return seg

    def _bind_multicast(self, mem_handle: int, total: int, gran: int) -> Tuple[int, int]:
        """(multicast address, handle) of this segment, or (0, 0) when any rank fails a step: every rank then runs
        the peer-load/store kernels on the unicast mappings (decided collectively, so nobody waits for a peer that
        gave up)."""
        w, r = self.world_size, self.rank
        mc_handle, fd, ok = 0, None, True
        if r == 0:
            try:
                mc_handle, fd = native.symm_mc_create(total, w)
            except RuntimeError as exc:
                logger.warning(f"cuMulticastCreate failed ({exc}); NVLS disabled")
                ok = False
        if not self._all_ok(ok):
            self.multicast = False
            return 0, 0
        fds = self._exchange_fds(fd)
        stage = "import"
        try:
            if r != 0:
                mc_handle = native.symm_vmm_import(fds[0])
                os.close(fds[0])
            else:
                os.close(fd)
            stage = "add_device"
            native.symm_mc_add_device(mc_handle)
        except RuntimeError as exc:
            logger.warning(f"[rank {r}] multicast {stage} failed ({exc}); NVLS disabled")
            ok = False
        if not self._all_ok(ok):            # also the barrier "every device was added" that must precede binding
            self.multicast = False
            return 0, 0
        mc_ptr = 0
        try:
            native.symm_mc_bind(mc_handle, mem_handle, total)
            mc_ptr = native.symm_vmm_map(mc_handle, total, gran)
        except RuntimeError as exc:
            logger.warning(f"[rank {r}] multicast bind/map failed ({exc}); NVLS disabled")
            ok = False
        if not self._all_ok(ok):
            self.multicast = False
            return 0, 0
        return mc_ptr, mc_handle

    @property
    def nvls_active(self) -> bool:
        """True when the bandwidth-bound collectives run on the multimem (NVLS) kernels: a multicast mapping exists
        and the library's policy (``use_nvls`` in collectives.cu: world >= 4, or ``DTB_COMM_NVLS=1``) selects them."""
        mode = os.environ.get("DTB_COMM_NVLS", "")
        if not self.multicast or self.world_size < 2 or mode == "0":
            return False
        return mode == "1" or self.world_size >= 4

    def reserve(self, nbytes: int) -> Segment:
        """Make sure ONE segment has ``nbytes`` free and make it the preferred segment of the allocations that
        follow, so buffers a fused kernel addresses by offset (gradient bucket + parameters, shard + full weight +
        gradient) end up in the same segment even when older segments still have holes."""
        nbytes = (nbytes + ALIGN - 1) // ALIGN * ALIGN
        seg = next((s for s in self.segments if s.nbytes - s.used >= nbytes), None)
        if seg is None:
            seg = self._new_segment(nbytes)
        self._preferred = seg
        return seg

    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        esize = torch.empty((), dtype=dtype).element_size()
        nbytes = (numel * esize + ALIGN - 1) // ALIGN * ALIGN
        order = ([self._preferred] if self._preferred is not None else []) + \
            [s for s in self.segments if s is not self._preferred]
        seg = next((s for s in order if s.nbytes - s.used >= nbytes), None)
        if seg is None:
            seg = self._new_segment(nbytes)
        offset = seg.used
        seg.used += nbytes
        flat = seg.base_tensor[offset: offset + nbytes].view(dtype)[:numel]  # type: ignore[index]
        self._owner_of[flat.data_ptr()] = (seg, offset)
        return flat

    def locate(self, t: torch.Tensor) -> Tuple[Segment, int]:
        """(segment, byte offset) of a tensor that lives in the symmetric heap."""
        ptr = t.data_ptr()
        hit = self._owner_of.get(ptr)
        if hit is not None:
            return hit
        for seg in self.segments:
            if seg.local_ptr <= ptr < seg.local_ptr + seg.nbytes:
                return seg, ptr - seg.local_ptr
        raise ValueError("tensor is not in the symmetric heap (allocate it with SymmComm.alloc)")

    def in_heap(self, t: torch.Tensor) -> bool:
        ptr = t.data_ptr()
        return any(seg.local_ptr <= ptr < seg.local_ptr + seg.nbytes for seg in self.segments)

    def scratch(self, nbytes: int) -> torch.Tensor:
        """A reusable heap buffer for operands that were not allocated in the heap."""
        if self._scratch is None or self._scratch.numel() < nbytes:
            self._scratch = self.alloc(max(nbytes, 1 << 20), torch.uint8)
        return self._scratch[:nbytes]

    # ------------------------------------------------------------ collectives
    def _on_comm_stream(self):
        """Order the comm stream after the caller's current stream and switch to it."""
        if self.stream is None:
            return torch.cuda.stream(torch.cuda.current_stream())
        self.stream.wait_stream(torch.cuda.current_stream())
        return torch.cuda.stream(self.stream)

    def _comm_stream_after_current(self) -> "torch.cuda.Stream":
        """The stream collectives launch on, ordered after everything queued on the caller's stream.  Kernels are
        launched onto it by raw handle: no ``torch.cuda.stream`` context (~10 us of host time per collective)."""
        current = torch.cuda.current_stream()
        if self.stream is None:
            return current
        self.stream.wait_stream(current)
        return self.stream

    def broadcast_(self, tensors: Sequence[torch.Tensor], src: int = 0) -> None:
        """Tensors that already live in the heap (parameters re-homed next to their gradient buckets) are broadcast
        IN PLACE, one kernel per tensor and no staging; anything else is coalesced into heap slabs, one peer-store
        (or multimem.st) kernel per slab, and copied back."""
        if self.world_size == 1:
            return
        staged: List[torch.Tensor] = []
        for t in tensors:
            if t.is_contiguous() and self.in_heap(t) and t.data_ptr() % 16 == 0:
                seg, off = self.locate(t)
                nbytes = (t.numel() * t.element_size() + 15) // 16 * 16     # heap slots are padded to 256 B
                native.comm_broadcast(seg.desc, off, nbytes, src, self.comm_ctas)
            else:
                staged.append(t)
        # coalesce the rest into heap slabs, one kernel per slab
        slab_cap = 256 << 20
        if staged:      # size the staging buffer once (scratch() never shrinks or frees)
            self.scratch(min(slab_cap, sum((t.numel() * t.element_size() + 15) // 16 * 16 for t in staged)))
        group: List[torch.Tensor] = []
        size = 0

        def flush() -> None:
            nonlocal group, size
            if not group:
                return
            stage = self.scratch((size + 15) // 16 * 16)
            seg, off = self.locate(stage)
            cursor = 0
            views = []
            for t in group:
                nb = t.numel() * t.element_size()
                view = stage[cursor: cursor + nb].view(t.dtype).view(t.shape)
                if self.rank == src:
                    view.copy_(t)
                views.append((t, view))
                cursor += (nb + 15) // 16 * 16
            native.comm_broadcast(seg.desc, off, (cursor + 15) // 16 * 16, src, self.comm_ctas)
            if self.rank != src:
                with torch.no_grad():
                    for t, view in views:
                        t.copy_(view)
            group, size = [], 0

        for t in staged:
            nb = (t.numel() * t.element_size() + 15) // 16 * 16
            if size + nb > slab_cap:
                flush()
            group.append(t)
            size += nb
        flush()
        torch.cuda.current_stream().synchronize()

    def _ctas_for(self, nbytes: int) -> int:
        if self.exclusive_sms > 0 and self.exclusive_now and nbytes >= self.exclusive_min_bytes:
            return (self.exclusive_boost or self.exclusive_sms) | native.EXCLUSIVE
        return self.comm_ctas

    def _ce_scratch(self, nbytes: int) -> Optional[torch.Tensor]:
        """Local staging for the copy-engine path, or None when this buffer should take the SM path."""
        w = self.world_size
        if not self.copy_engine or w < 2 or nbytes < self.copy_engine_min_bytes or nbytes % (w * 32):
            return None
        need = nbytes // w * (w - 1)
        if not self._ce_staging or self._ce_staging[-1].numel() < need:
            # never freed or shrunk while the communicator lives: kernels on the comm stream may still read it
            self._ce_staging.append(torch.empty(need, dtype=torch.uint8, device=self.device))
        return self._ce_staging[-1]

    def all_reduce_avg_(self, buf: torch.Tensor) -> Work:
        if buf.dtype not in (torch.bfloat16, torch.float32):
            raise TypeError(f"all_reduce_avg_ supports bf16/fp32, got {buf.dtype}")
        seg, off = self.locate(buf)
        esize = buf.element_size()
        numel = (buf.numel() * esize + 15) // 16 * 16 // esize      # heap slots are padded to 256 B
        staging = self._ce_scratch(numel * esize) if buf.dtype == torch.bfloat16 else None
        stream = self._comm_stream_after_current()
        if staging is not None:
            native.comm_all_reduce_ce(seg.desc, off, off, False, numel, staging, max_ctas=self.comm_ctas,
                                      stream=stream.cuda_stream)
        else:
            native.comm_all_reduce_avg(seg.desc, off, numel, buf.dtype == torch.float32, 0,
                                       self._ctas_for(numel * esize), stream=stream.cuda_stream)
        event = torch.cuda.Event()
        event.record(stream)
        return _StreamWork(event)

    def then(self, work: Optional[Work], fn) -> Optional[Work]:
        """``fn()`` on the communication stream, behind ``work`` (same stream) and everything already queued on the
        caller's stream; e.g. the optimizer update of a bucket right after its all-reduce."""
        if self.stream is None:
            return super().then(work, fn)
        with self._on_comm_stream():
            fn()
            event = torch.cuda.Event()
            event.record()
        return _StreamWork(event)

    def all_reduce_sgd_(self, grads: torch.Tensor, params: torch.Tensor, lr: float, weight_decay: float) -> Work:
        """Average ``grads`` over ranks and apply ``p = p*(1-lr*wd) - lr*g`` to ``params`` (same layout, both in
        the same heap segment) in one kernel; every rank ends with identical new parameters."""
        if grads.dtype != torch.bfloat16 or params.dtype != torch.bfloat16:
            raise TypeError("all_reduce_sgd_ is a bf16 path")
        seg, g_off = self.locate(grads)
        seg_p, p_off = self.locate(params)
        if seg is not seg_p:
            raise ValueError("all_reduce_sgd_: grads and params must live in the same symmetric segment")
        numel = (grads.numel() * 2 + 15) // 16 * 8
        staging = self._ce_scratch(numel * 2)
        stream = self._comm_stream_after_current()
        if staging is not None:
            native.comm_all_reduce_ce(seg.desc, g_off, p_off, True, numel, staging, lr, weight_decay,
                                      max_ctas=self.comm_ctas, stream=stream.cuda_stream)
        else:
            native.comm_all_reduce_sgd(seg.desc, g_off, p_off, numel, lr, weight_decay,
                                       self._ctas_for(numel * 2), stream=stream.cuda_stream)
        event = torch.cuda.Event()
        event.record(stream)
        return _StreamWork(event)

    def all_gather_into(self, out: torch.Tensor, shard: torch.Tensor) -> None:
        seg_o, off_o = self.locate(out)
        seg_s, off_s = self.locate(shard)
        if seg_o is not seg_s:
            raise ValueError("all_gather_into: out and shard must live in the same symmetric segment")
        native.comm_all_gather(seg_o.desc, off_o, off_s, shard.numel(), shard.element_size(), self.comm_ctas)

    def reduce_scatter_avg_into(self, out: torch.Tensor, full: torch.Tensor) -> None:
        seg, off = self.locate(full)
        native.comm_reduce_scatter_avg(seg.desc, out, off, out.numel(), full.dtype == torch.float32, self.comm_ctas)

    def barrier(self) -> None:
        if self.world_size == 1 or not self.segments:
            if self.world_size > 1:
                dist.barrier()
            return
        native.comm_barrier(self.segments[0].desc, 7)
        torch.cuda.current_stream().synchronize()

    def check_errors(self) -> None:
        for seg in self.segments:
            flag = native.comm_error_flag(seg.desc)
            if flag:
                raise RuntimeError(f"collective watchdog fired on rank {self.rank}: peer {flag - 1} never arrived")

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
            if self.world_size > 1 and dist.is_initialized():
                dist.barrier()
        except Exception:  # pragma: no cover
            pass
        for seg in self.segments:
            for r, p in enumerate(seg.peer_ptrs):
                if r != self.rank:
                    try:
                        if seg.backend == "vmm":
                            native.symm_vmm_unmap(p, seg.total_bytes)
                            native.symm_vmm_release(seg.handles[r])
                        else:
                            native.symm_close(p)
                    except Exception:  # pragma: no cover
                        pass
        # the local allocations (and the multicast mapping of them) are released at process exit: live tensors may
        # still point into them
