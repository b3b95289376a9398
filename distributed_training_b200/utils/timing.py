"""Device-side step timing: CUDA events on the launching stream, barrier + synchronize on both
sides, max over ranks (BASELINE.json metric definition)."""
from __future__ import annotations

from typing import Callable

import torch
import torch.distributed as dist


def barrier_sync() -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def max_over_ranks(value: float) -> float:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def time_steps_ms(step: Callable[[int], None], steps: int) -> float:
    """Total device milliseconds for ``steps`` calls of ``step(i)``, max over ranks."""
    barrier_sync()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(steps):
        step(i)
    stop.record()
    torch.cuda.synchronize()
    total = start.elapsed_time(stop)
    barrier_sync()
    return max_over_ranks(total)
