"""Hand-rolled replicated data parallelism.

What the reference gets from ``torch.nn.parallel.DistributedDataParallel``
(/root/reference/src/dist_strategy/ddp_strategy.py:18) and what its playground
spells out by hand (/root/reference/src/playground/ddp_script.py:120-121,149-154)
is implemented here directly:

* at wrap time rank 0's parameters and buffers are broadcast to every rank;
* gradients live in flat *buckets* (reverse registration order, 1 MiB first
  bucket, ``bucket_cap_mb`` afterwards); each ``param.grad`` is a view into
  its bucket, so reduced values are visible to the optimizer without a copy;
* a post-accumulate-grad hook per parameter counts the bucket down and
  launches its all-reduce the moment the last gradient lands, so
  communication overlaps the rest of backward; averaging (and, on B200, the
  bf16/fp32 cast) happens inside the collective;
* an autograd-engine callback joins all in-flight buckets at the end of
  backward, so ``loss.backward(); optimizer.step()`` works unchanged.

Models that implement ``forward_backward(inputs, targets, grad_sink=...)``
(the native MLP / GPT-2) bypass autograd: their wgrad kernels write straight
into the bucket views and call :meth:`ReplicatedModule.mark_ready`.

The wrapper is communicator-agnostic: gloo on CPU (``TorchDistComm``), the
peer-memory kernels on B200 (``SymmComm``).
"""
from __future__ import annotations

import contextlib
import os
import zlib
import logging
import time
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
from torch import nn

from .buckets import BucketPlan, plan_buckets
from .comm import Communicator, Work, make_communicator
from .strategy import DistributedStrategy, OptimizerSpec, atomic_torch_save

logger = logging.getLogger(__name__)


class GradBucket:
    """Flat gradient buffer plus the bookkeeping for one reduction per step."""

    def __init__(self, plan: BucketPlan, buffer: torch.Tensor, params: List[nn.Parameter]) -> None:
        self.plan = plan
        self.buffer = buffer
        self.param_buffer: Optional[torch.Tensor] = None     # parameters with the same slot layout (fused step)
        self.views: Dict[int, torch.Tensor] = {}
        for slot in plan.slots:
            self.views[slot.param_index] = buffer[slot.offset: slot.offset + slot.numel].view(slot.shape)
        self.expected = len(plan.slots)
        self.pending = self.expected
        self.ready: set = set()
        self.work: Optional[Work] = None
        self.launched = False

    def reset(self) -> None:
        self.pending = self.expected
        self.ready.clear()
        self.work = None
        self.launched = False


class ReplicatedModule(nn.Module):
    """``module`` replicated on every rank with bucketed, overlapped gradient averaging.

    ``state_dict()`` keys carry the ``module.`` prefix, exactly like torch's
    DDP wrapper, so snapshots are interchangeable with the reference's
    (SURVEY §5.4).
    """

    def async_wgrad_mode(self) -> Optional[str]:
        """How a native model may overlap its weight-gradient GEMMs with the dgrad chain (``grad_view`` / ``mark_ready``
        may be called with a side stream current; the reduction is ordered behind whatever stream is current).
        One replica: "full" (the chains never join).  Several replicas: one stream.  Both overlapped forms were
        measured slower at 4 GPUs because the waiting CTAs of the second GEMM take the SMs that the backward GEMMs
        leave to the whole-SM reduction CTAs: "full" 10.38 ms/step against 10.26-10.35 (128 MiB reductions 0.9-1.15 ms
        instead of 0.48, profiles/timeline_wide_mlp_n4_async_x8.json), "forkjoin" (W_i beside D_i, join per layer,
        reduction queued at the join) 10.92-10.95 against 10.22-10.28.  DTB_ASYNC_WGRAD_MULTI = full | forkjoin turns
        them on."""
        if self.comm.world_size == 1:
            return "full"
        env = os.environ.get("DTB_ASYNC_WGRAD_MULTI", "0")
        return {"full": "full", "1": "full", "forkjoin": "forkjoin"}.get(env)

    def __init__(self, module: nn.Module, comm: Communicator, bucket_cap_mb: float = 25.0,
                 first_bucket_mb: float = 1.0, init_sync: bool = True,
                 grad_dtype: Optional[torch.dtype] = None) -> None:
        super().__init__()
        self.module = module
        self.comm = comm
        self.require_backward_grad_sync = True
        self._params: List[nn.Parameter] = [p for p in module.parameters() if p.requires_grad]
        self._index_of: Dict[int, int] = {id(p): i for i, p in enumerate(self._params)}
        if init_sync:
            self.verify_replicas()

        plans = plan_buckets(self._params, int(bucket_cap_mb * (1 << 20)),
                             int(first_bucket_mb * (1 << 20)), reverse=True, grad_dtype=grad_dtype)
        # On the peer-memory communicator the parameters themselves are re-homed into the heap, laid out
        # like their gradient bucket, so a bucket's reduction can be fused with the optimizer step.
        self.params_in_heap = bool(getattr(comm, "supports_fused_step", False)) and grad_dtype is None
        comm.reserve(sum(pl.nbytes + 256 for pl in plans) * (2 if self.params_in_heap else 1))
        self.buckets: List[GradBucket] = []
        self._bucket_of: Dict[int, GradBucket] = {}
        for plan in plans:
            bucket = GradBucket(plan, comm.alloc(plan.numel, plan.dtype), self._params)
            if self.params_in_heap:
                bucket.param_buffer = comm.alloc(plan.numel, plan.dtype)
                with torch.no_grad():
                    for slot in plan.slots:
                        p = self._params[slot.param_index]
                        home = bucket.param_buffer[slot.offset: slot.offset + slot.numel].view(slot.shape)
                        home.copy_(p.data)
                        p.data = home
            self.buckets.append(bucket)
            for idx in plan.param_indices:
                self._bucket_of[idx] = bucket
        if init_sync:
            # after re-homing: heap-resident parameter buffers are broadcast in place (one kernel per bucket, no
            # staging copies); buffers and parameters outside the heap go through a staging slab
            self.sync_parameters()
        self._optimizer: Any = None
        self._hook_handles = []
        for i, p in enumerate(self._params):
            self._hook_handles.append(p.register_post_accumulate_grad_hook(self._make_hook(i)))
        self.attach_grad_views()
        self._callback_queued = False
        self._in_manual_step = False
        self._chunk_works: List[Work] = []
        #: seconds (CPU) or milliseconds of stream stall (CUDA) the last step spent
        #: waiting for gradient communication after backward had finished
        self.last_exposed_comm_ms = 0.0
        self._stall_events: Optional[Tuple[Any, Any]] = None

    # ------------------------------------------------------------------ setup
    def verify_replicas(self) -> None:
        """Fail fast if ranks wrapped different architectures (torch's ``_verify_param_shape_across_processes``,
        SURVEY N1): a CRC of every parameter's name, shape and dtype must agree on all ranks."""
        if self.comm.world_size == 1:
            return
        desc = ";".join(f"{n}:{tuple(p.shape)}:{p.dtype}" for n, p in self.module.named_parameters())
        crc = float(zlib.crc32(desc.encode()))           # < 2**32, exact in float64
        hi = self.comm.all_reduce_max_scalar(crc)
        lo = -self.comm.all_reduce_max_scalar(-crc)
        if hi != lo:
            raise RuntimeError(
                f"rank {self.comm.rank}: replicas disagree on parameter names/shapes/dtypes "
                f"(fingerprint {int(crc):#x}, range {int(lo):#x}..{int(hi):#x})")

    def sync_parameters(self, src: int = 0) -> None:
        """Broadcast rank ``src``'s parameters and buffers (DDP init sync)."""
        homed = {id(p) for b in getattr(self, "buckets", []) if b.param_buffer is not None
                 for p in (self._params[i] for i in b.plan.param_indices)}
        tensors = [b.param_buffer for b in getattr(self, "buckets", []) if b.param_buffer is not None]
        tensors += [p.data for p in self.module.parameters() if id(p) not in homed]
        tensors += [b.data for b in self.module.buffers()]
        self.comm.broadcast_(tensors, src=src)

    def attach_grad_views(self) -> None:
        """Point every ``param.grad`` at its bucket view (zero-filled)."""
        for i, p in enumerate(self._params):
            view = self._bucket_of[i].views[i]
            if p.grad is not None and p.grad.data_ptr() != view.data_ptr():
                view.copy_(p.grad)
            p.grad = view if view.dtype == p.dtype else None

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self) -> Iterator[None]:
        """Skip gradient averaging inside the context (gradient accumulation)."""
        previous = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = previous

    # ------------------------------------------------------------ autograd path
    def _make_hook(self, index: int):
        def hook(param: torch.Tensor) -> None:
            if not self.require_backward_grad_sync or self._in_manual_step:
                return
            bucket = self._bucket_of[index]
            view = bucket.views[index]
            grad = param.grad
            if grad is not None and grad.data_ptr() != view.data_ptr():
                # optimizer.zero_grad(set_to_none=True) detached the view: land the fresh
                # gradient in the bucket and re-attach so the optimizer reads the reduced value
                view.copy_(grad)
                if view.dtype == param.dtype:
                    param.grad = view
            if not self._callback_queued:
                self._callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            self._count_down(bucket, index)

        return hook

    def _count_down(self, bucket: GradBucket, index: int) -> None:
        if index in bucket.ready:
            return
        bucket.ready.add(index)
        bucket.pending -= 1
        if bucket.pending == 0:
            self._launch(bucket)

    # ------------------------------------------------------------- fused step
    def attach_optimizer(self, optimizer: Any) -> None:
        """Let gradient reduction and the optimizer step run as one kernel where that is exact:
        plain SGD (no momentum, no master weights) on bf16 parameters of a native-step model."""
        self._optimizer = optimizer

    def _fused_sgd_hyper(self) -> Optional[Tuple[float, float]]:
        opt = self._optimizer
        if opt is None or not getattr(opt, "fusable_sgd", False) or not self.params_in_heap or not self.native_step:
            return None
        group = opt.param_groups[0]
        return float(group["lr"]), float(group["weight_decay"])

    def _launch(self, bucket: GradBucket) -> None:
        if self.comm.world_size > 1:      # a single replica already holds the mean
            hyper = self._fused_sgd_hyper() if self._in_manual_step else None
            if hyper is not None and bucket.param_buffer is not None and bucket.plan.dtype == torch.bfloat16:
                bucket.work = self.comm.all_reduce_sgd_(bucket.buffer, bucket.param_buffer, *hyper)
                self._optimizer.mark_updated(self._params[i] for i in bucket.plan.param_indices)
            else:
                bucket.work = self.comm.all_reduce_avg_(bucket.buffer)
                self._step_behind(bucket, [self._params[i] for i in bucket.plan.param_indices])
        elif self._in_manual_step:
            self._step_behind(bucket, [self._params[i] for i in bucket.plan.param_indices])
        bucket.launched = True

    def _step_behind(self, bucket: GradBucket, params: List[nn.Parameter]) -> None:
        """Stateful optimizers (AdamW, SGD with momentum / master weights) cannot be folded into the reduction kernel,
        but their update of a bucket can run right behind that bucket's all-reduce on the communication stream, under
        the rest of backward, instead of as one pass over the whole model afterwards.  Native models only: they mark a
        parameter ready after its last use in backward, so nothing reads the old value any more."""
        opt = self._optimizer
        if getattr(opt, "fusable_sgd", False):
            return          # stateless SGD: fused into the reduction / wgrad epilogue; leftovers are one tiny launch
        if (opt is None or not self._in_manual_step or not hasattr(opt, "step_params") or not self.params_in_heap
                or any(self._bucket_of[self._index_of[id(p)]].views[self._index_of[id(p)]].dtype != p.dtype
                       for p in params)):
            return
        for p in params:       # gradients live in the bucket views
            if p.grad is None:
                p.grad = self._bucket_of[self._index_of[id(p)]].views[self._index_of[id(p)]]
        bucket.work = self.comm.then(bucket.work, lambda: opt.step_params(params))

    def _finalize_backward(self) -> None:
        self._callback_queued = False
        self.finish_gradient_sync()

    # --------------------------------------------------------------- manual path
    def grad_view(self, param: nn.Parameter) -> torch.Tensor:
        """Bucket slice a native wgrad kernel should write ``param``'s gradient into."""
        i = self._index_of[id(param)]
        return self._bucket_of[i].views[i]

    def mark_ready(self, param: nn.Parameter) -> None:
        """Declare ``param``'s gradient final for this step (native models call this)."""
        if not self.require_backward_grad_sync:
            return
        i = self._index_of[id(param)]
        self._count_down(self._bucket_of[i], i)

    #: weights smaller than this are reduced in one piece even when they end backward
    TAIL_CHUNK_MIN_BYTES = 32 << 20
    #: pieces a weight gradient at the end of backward is produced (and reduced) in; DTB_DDP_TAIL_CHUNKS overrides.
    #: Device timeline at 4 GPUs (profiles/timeline_wide_mlp_n4_sync_x4.json): with only the LAST wgrad cut in 8, the
    #: bucket before it was still reduced in one piece after its GEMM (0.47 ms) and the eight 16 MiB pieces queued up
    #: behind it at 78 us each: 0.52 ms of reduction after the last GEMM.  Eight 1024-row GEMMs also took 0.57 ms
    #: where the whole wgrad takes 0.36 (128 tiles on 72 CTA pairs = two waves each).  Four 2048-row pieces for each
    #: of the last TAIL_LAYERS layers: the pieces are 4 full waves on 66 pairs and every reduction but the last one
    #: runs under a GEMM.
    TAIL_CHUNKS = int(os.environ.get("DTB_DDP_TAIL_CHUNKS", "4"))
    #: how many layers at the end of backward get that treatment (DTB_DDP_TAIL_LAYERS)
    TAIL_LAYERS = int(os.environ.get("DTB_DDP_TAIL_LAYERS", "2"))
    #: whole-SM CTAs of the reductions in the tail (DTB_DDP_TAIL_SMS): the GEMM pieces there leave that many SMs free
    #: (``tail_gemm_ctas``) at no cost in waves, and nothing hides a slow reduction any more
    TAIL_SMS = int(os.environ.get("DTB_DDP_TAIL_SMS", "16"))

    def begin_tail(self) -> None:
        """Called by a native model before the first chunked weight gradient: from here on the reductions are on the
        critical path, so they get ``TAIL_SMS`` whole SMs instead of ``exclusive_sms``."""
        if getattr(self.comm, "exclusive_now", False) and self.TAIL_SMS > int(getattr(self.comm, "exclusive_sms", 0)):
            self.comm.exclusive_boost = self.TAIL_SMS

    def tail_gemm_ctas(self, cap: int) -> int:
        """CTA cap for the GEMM pieces of the tail (leaves ``TAIL_SMS`` SMs to the reductions)."""
        if not getattr(self.comm, "exclusive_boost", 0):
            return cap
        sms = torch.cuda.get_device_properties(self.comm.device).multi_processor_count
        want = max(2, (sms - self.comm.exclusive_boost) & ~1)
        return min(cap, want) if cap else want

    def tail_chunks(self, param: nn.Parameter) -> int:
        """How many row chunks the *last* weight gradient of backward should be produced in.  Its reduction
        cannot overlap anything, so cutting it in pieces lets all but the last piece hide behind the
        remaining wgrad GEMMs."""
        i = self._index_of[id(param)]
        bucket = self._bucket_of[i]
        if self.comm.world_size == 1 or bucket.expected != 1 or param.dim() != 2 or param.shape[0] % self.TAIL_CHUNKS:
            return 1
        return self.TAIL_CHUNKS if bucket.plan.nbytes >= self.TAIL_CHUNK_MIN_BYTES else 1

    def mark_ready_chunk(self, param: nn.Parameter, chunk: int, n_chunks: int) -> None:
        """Rows ``[chunk, chunk+1) * rows/n_chunks`` of ``param``'s gradient are final: reduce just that slice
        (the parameter must own its bucket, see :meth:`tail_chunks`)."""
        if not self.require_backward_grad_sync:
            return
        i = self._index_of[id(param)]
        bucket = self._bucket_of[i]
        assert bucket.expected == 1 and bucket.plan.slots[0].offset == 0
        per = param.numel() // n_chunks
        lo, hi = chunk * per, (chunk + 1) * per
        hyper = self._fused_sgd_hyper() if self._in_manual_step else None
        if hyper is not None and bucket.param_buffer is not None and bucket.plan.dtype == torch.bfloat16:
            work = self.comm.all_reduce_sgd_(bucket.buffer[lo:hi], bucket.param_buffer[lo:hi], *hyper)
            if chunk == n_chunks - 1:
                self._optimizer.mark_updated([param])
        else:
            work = self.comm.all_reduce_avg_(bucket.buffer[lo:hi])
        self._chunk_works.append(work)
        if chunk == n_chunks - 1:
            bucket.ready.add(i)
            bucket.pending = 0
            bucket.launched = True
            if hyper is None:
                self._step_behind(bucket, [param])

    def forward_backward(self, inputs: torch.Tensor, targets: torch.Tensor, **kwargs: Any) -> torch.Tensor:
        """One fused forward+backward of a native model with overlapped reduction."""
        if not hasattr(self.module, "forward_backward"):
            raise AttributeError(f"{type(self.module).__name__} has no native forward_backward")
        self._in_manual_step = True
        try:
            hyper = self._fused_sgd_hyper()
            if hyper is not None and self.comm.world_size == 1:
                # single replica: the wgrad GEMM epilogue applies W -= lr*(dW + wd*W) itself
                kwargs["fused_sgd"] = hyper
                kwargs["on_updated"] = self._optimizer.mark_updated
            # exclusive-mode reduction: backward GEMMs leave `exclusive_sms` SMs free and the large all-reduces run
            # there on whole-SM CTAs at their stand-alone speed instead of fighting a GEMM CTA for its SM
            excl = int(getattr(self.comm, "exclusive_sms", 0) or 0)
            use_excl = (excl > 0 and self.comm.world_size > 1 and self.comm.device.type == "cuda"
                        and getattr(self.module, "supports_bwd_max_ctas", False))
            if use_excl:
                sms = torch.cuda.get_device_properties(self.comm.device).multi_processor_count
                kwargs.setdefault("bwd_max_ctas", max(2, (sms - excl) & ~1))
                self.comm.exclusive_now = True
            loss = self.module.forward_backward(inputs, targets, grad_sink=self, **kwargs)
            self.finish_gradient_sync()
        finally:
            self._in_manual_step = False
            if hasattr(self.comm, "exclusive_now"):
                self.comm.exclusive_now = False
                self.comm.exclusive_boost = 0
        return loss

    def finish_gradient_sync(self) -> None:
        """Join every bucket; unused parameters contribute zeros."""
        if not self.require_backward_grad_sync:
            return
        for bucket in self.buckets:
            if not bucket.launched:
                for idx, view in bucket.views.items():
                    if idx not in bucket.ready:
                        view.zero_()
                self._launch(bucket)
        cuda = self.comm.device.type == "cuda"
        if cuda:
            start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            start.record()
        t0 = time.perf_counter()
        for bucket in self.buckets:
            if bucket.work is not None:
                bucket.work.wait()
        for work in self._chunk_works:
            work.wait()
        self._chunk_works.clear()
        if cuda:
            stop.record()
            self._stall_events = (start, stop)
        else:
            self.last_exposed_comm_ms = (time.perf_counter() - t0) * 1e3
        for i, p in enumerate(self._params):
            view = self._bucket_of[i].views[i]
            if view.dtype != p.dtype:            # reduced in a narrower wire dtype
                if p.grad is None:
                    p.grad = view.to(p.dtype)
                else:
                    p.grad.copy_(view)
            elif p.grad is None:
                p.grad = view
        for bucket in self.buckets:
            bucket.reset()

    def exposed_comm_ms(self) -> float:
        """Time the compute stream stalled on gradient communication last step."""
        if self._stall_events is not None:
            start, stop = self._stall_events
            stop.synchronize()
            self.last_exposed_comm_ms = start.elapsed_time(stop)
            self._stall_events = None
        return self.last_exposed_comm_ms

    @property
    def native_step(self) -> bool:
        """True when the wrapped model runs its own fused forward+backward (no autograd)."""
        return hasattr(self.module, "forward_backward") and getattr(self.module, "native_ready", lambda: False)()

    def extra_repr(self) -> str:
        sizes = ", ".join(f"{b.plan.nbytes / 2**20:.1f}MiB" for b in self.buckets)
        return f"world={self.comm.world_size}, buckets=[{sizes}]"


class DDPStrategy(DistributedStrategy):
    """Replicated strategy; constructor signature as the reference's
    ``DDPStrategy(local_rank, is_gpu)`` (/root/reference/src/dist_strategy/ddp_strategy.py:10-13)."""

    def __init__(self, local_rank: int = 0, is_gpu: bool = False, bucket_cap_mb: float = 25.0,
                 comm: Optional[Communicator] = None, first_bucket_mb: float = 1.0) -> None:
        self.local_rank = local_rank
        self.is_gpu = is_gpu
        self.bucket_cap_mb = bucket_cap_mb
        self.first_bucket_mb = first_bucket_mb
        self.comm = comm

    def _device(self, device: Any) -> torch.device:
        if self.is_gpu:
            return torch.device("cuda", self.local_rank if not isinstance(device, torch.device) or device.index is None
                                else device.index)
        return torch.device("cpu")

    def prepare_model(self, model: nn.Module, device: Any) -> nn.Module:
        if self.comm is None:
            self.comm = make_communicator(self._device(device))
        return ReplicatedModule(model, self.comm, bucket_cap_mb=self.bucket_cap_mb,
                                first_bucket_mb=self.first_bucket_mb)

    def save_checkpoint(self, model: nn.Module, state_dict: dict, path: str) -> None:
        atomic_torch_save(state_dict, path)

    def load_checkpoint(self, model: nn.Module, checkpoint: dict, device: Any) -> Tuple[nn.Module, int]:
        state = checkpoint["MODEL_STATE"]
        wrapped = isinstance(model, ReplicatedModule)
        has_prefix = all(k.startswith("module.") for k in state) and len(state) > 0
        if wrapped and not has_prefix:      # snapshot from the sharded strategy / a bare model
            state = {f"module.{k}": v for k, v in state.items()}
        elif not wrapped and has_prefix:
            state = {k[len("module."):]: v for k, v in state.items()}
        model.load_state_dict(state)
        return model, int(checkpoint["EPOCHS_RUN"])

    def build_optimizer(self, model: nn.Module, spec: OptimizerSpec) -> Any:
        from ..optim import build_fused_optimizer

        return build_fused_optimizer([p for p in model.parameters() if p.requires_grad], spec)

    def barrier(self) -> None:
        if self.comm is not None:
            self.comm.barrier()
