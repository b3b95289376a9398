"""Gradient bucket planning.

torch's DDP ``Reducer`` (what the reference runs implicitly through
``DDP(model)`` at /root/reference/src/dist_strategy/ddp_strategy.py:18) walks
the parameters in *reverse* registration order, closes a first small bucket at
1 MiB so communication starts early in backward, and caps later buckets at
``bucket_cap_mb`` (25 MiB).  This planner keeps those rules and adds two that
matter for the peer-memory kernels:

* a parameter that alone reaches the cap gets a bucket of its own (the
  128 MiB weights of the wide MLP are reduced the moment their wgrad lands);
* every slot starts on a 128-byte boundary, so the all-reduce kernels can use
  16-byte vector loads on every rank without a ragged head.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Sequence, Tuple

import torch

FIRST_BUCKET_BYTES = 1 << 20
DEFAULT_BUCKET_CAP_BYTES = 25 << 20
SLOT_ALIGN_BYTES = 128


@dataclass
class BucketSlot:
    """Where one parameter's gradient lives inside its bucket."""

    param_index: int
    offset: int          # elements from bucket start
    numel: int
    shape: Tuple[int, ...]


@dataclass
class BucketPlan:
    index: int
    dtype: torch.dtype
    device: torch.device
    slots: List[BucketSlot] = field(default_factory=list)
    numel: int = 0       # padded length in elements

    @property
    def nbytes(self) -> int:
        return self.numel * torch.empty((), dtype=self.dtype).element_size()

    @property
    def param_indices(self) -> List[int]:
        return [s.param_index for s in self.slots]


def _align(n_elems: int, elem_size: int) -> int:
    per = max(SLOT_ALIGN_BYTES // elem_size, 1)
    return (n_elems + per - 1) // per * per


def plan_buckets(params: Sequence[torch.Tensor], bucket_cap_bytes: int = DEFAULT_BUCKET_CAP_BYTES,
                 first_bucket_bytes: int = FIRST_BUCKET_BYTES, reverse: bool = True,
                 grad_dtype: "torch.dtype | None" = None) -> List[BucketPlan]:
    """Assign every tensor in ``params`` to a bucket.

    ``params`` is given in registration order; with ``reverse=True`` (default)
    they are visited last-to-first, the order backward produces gradients.
    Buckets are keyed by (dtype, device): tensors of different kinds never
    share one.  ``grad_dtype`` overrides the bucket element type (e.g. reduce
    fp32 master gradients as bf16 on the wire).
    """
    order: Iterable[int] = range(len(params) - 1, -1, -1) if reverse else range(len(params))
    open_buckets: Dict[Tuple[torch.dtype, torch.device], BucketPlan] = {}
    open_payload: Dict[Tuple[torch.dtype, torch.device], int] = {}
    closed_count: Dict[Tuple[torch.dtype, torch.device], int] = {}
    result: List[BucketPlan] = []

    def close(key) -> None:
        bucket = open_buckets.pop(key)
        open_payload.pop(key)
        closed_count[key] = closed_count.get(key, 0) + 1
        bucket.index = len(result)
        result.append(bucket)

    for i in order:
        p = params[i]
        dtype = grad_dtype or p.dtype
        key = (dtype, p.device)
        esize = torch.empty((), dtype=dtype).element_size()
        nbytes = p.numel() * esize
        limit = first_bucket_bytes if closed_count.get(key, 0) == 0 else bucket_cap_bytes
        if key in open_buckets and open_payload[key] + nbytes > limit:
            close(key)
            limit = bucket_cap_bytes
        if key not in open_buckets:
            open_buckets[key] = BucketPlan(index=-1, dtype=dtype, device=p.device)
            open_payload[key] = 0
        bucket = open_buckets[key]
        bucket.slots.append(BucketSlot(i, bucket.numel, p.numel(), tuple(p.shape)))
        bucket.numel += _align(p.numel(), esize)
        open_payload[key] += nbytes
        if open_payload[key] >= limit:
            close(key)
    for key in list(open_buckets):
        close(key)
    return result
