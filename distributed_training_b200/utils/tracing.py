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
