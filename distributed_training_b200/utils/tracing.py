"""NVTX ranges and step timers (the reference has no tracing of any kind).

``DTB200_NVTX=1`` turns the ranges on; they cost nothing otherwise.  Ranges are placed around the phases a
timeline needs to explain a step: ``h2d``, ``forward_backward`` (native) or ``forward`` / ``backward`` (autograd),
``grad_sync`` (waiting for bucket reductions), ``optimizer``.  Capture with ``nsys`` where available, or read the
ranges in ``ncu --nvtx``.
"""
from __future__ import annotations

import contextlib
import os
from typing import Iterator

import torch

ENABLED = os.environ.get("DTB200_NVTX", "0") == "1"


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    if ENABLED and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class KernelTimeline:
    """Device-side timeline of the native launches of a few steps, without a profiler.

    While active, selected wrappers of ``ops.native`` are bracketed by CUDA events on the stream they launch on;
    ``rows()`` returns ``{"name", "detail", "stream", "start_ms", "ms"}`` relative to the first launch.  Events
    cost ~2 µs of host time each and do not serialise streams, so overlap between the compute and comm streams is
    visible as overlapping intervals.  Used by ``bench/step_timeline.py`` to see WHICH kernels stretch when the
    gradient reduction runs beside them.
    """

    DEFAULT = ("gemm", "gemm_allgather", "colsum", "colsum_split", "cross_entropy_fwd_bwd", "mse_fwd_bwd",
               "attention_fwd", "attention_bwd", "layernorm_fwd", "layernorm_bwd", "embedding_fwd", "embedding_bwd",
               "multi_tensor_sgd", "multi_tensor_adamw", "cast_f32_to_bf16", "take_f32_to_bf16", "mean_f32",
               "shard_update", "comm_all_reduce_avg", "comm_all_reduce_sgd", "comm_all_reduce_ce",
               "comm_all_gather", "comm_reduce_scatter_avg", "comm_rs_optim", "comm_broadcast")

    def __init__(self, names=DEFAULT) -> None:
        self.names = tuple(names)
        self._saved = {}
        self._records = []

    def __enter__(self) -> "KernelTimeline":
        from ..ops import native

        for n in self.names:
            orig = getattr(native, n, None)
            if orig is None:
                continue
            self._saved[n] = orig
            setattr(native, n, self._wrap(n, orig))
        return self

    def __exit__(self, *exc) -> None:
        from ..ops import native

        for n, orig in self._saved.items():
            setattr(native, n, orig)
        self._saved.clear()

    def _wrap(self, name, fn):
        def traced(*args, **kwargs):
            # collectives are launched onto the communication stream by raw handle (``stream=``)
            raw = kwargs.get("stream")
            stream = torch.cuda.ExternalStream(raw) if isinstance(raw, int) and raw else torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            out = fn(*args, **kwargs)
            e1.record(stream)
            detail = ""
            if name.startswith("gemm"):
                detail = f"{kwargs.get('m')}x{kwargs.get('n')}x{kwargs.get('k')}"
                if kwargs.get("batch", (1, 1)) != (1, 1):
                    detail += f" b{kwargs['batch']}"
            elif "numel" in kwargs:
                detail = str(kwargs["numel"])
            elif name.startswith("comm_all_reduce") and len(args) >= 3:
                detail = str(args[3] if name.endswith("sgd") else args[2])
            self._records.append((name, detail, stream.cuda_stream, e0, e1))
            return out

        return traced

    def rows(self):
        torch.cuda.synchronize()
        if not self._records:
            return []
        t0 = self._records[0][3]
        streams = {}
        out = []
        for name, detail, sid, e0, e1 in self._records:
            out.append({"name": name, "detail": detail, "stream": streams.setdefault(sid, len(streams)),
                        "start_ms": round(t0.elapsed_time(e0), 4), "ms": round(e0.elapsed_time(e1), 4)})
        return out
