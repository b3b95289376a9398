"""Bucket planning rules: reverse order, 1 MiB first bucket, cap, oversize -> own bucket."""
import torch

from distributed_training_b200.parallel.buckets import plan_buckets


def _params(sizes, dtype=torch.float32):
    return [torch.empty(s, dtype=dtype) for s in sizes]


def test_reverse_order_and_first_bucket_cap():
    mib = (1 << 20) // 4
    ps = _params([mib // 2] * 6)                     # six 0.5 MiB tensors
    plans = plan_buckets(ps, bucket_cap_bytes=2 << 20, first_bucket_bytes=1 << 20)
    assert plans[0].param_indices == [5, 4]          # last parameters first, closed at 1 MiB
    assert plans[1].param_indices == [3, 2, 1, 0]    # then the 2 MiB cap
    assert sorted(i for p in plans for i in p.param_indices) == list(range(6))


def test_oversize_param_gets_own_bucket():
    mib = (1 << 20) // 4
    ps = _params([mib // 4, 40 * mib, mib // 4, mib // 4])
    plans = plan_buckets(ps)                          # 25 MiB cap
    big = [p for p in plans if 1 in p.param_indices]
    assert len(big) == 1 and big[0].param_indices == [1]


def test_dtypes_never_share_a_bucket_and_slots_are_aligned():
    ps = [torch.empty(100, dtype=torch.float32), torch.empty(100, dtype=torch.bfloat16),
          torch.empty(33, dtype=torch.float32)]
    plans = plan_buckets(ps)
    for plan in plans:
        assert len({ps[i].dtype for i in plan.param_indices}) == 1
        esize = torch.empty((), dtype=plan.dtype).element_size()
        for slot in plan.slots:
            assert (slot.offset * esize) % 128 == 0
        assert plan.numel * esize % 128 == 0


def test_wide_mlp_layout():
    # 8 x (8192x8192 weight + 8192 bias) in bf16: every weight is 128 MiB -> own bucket
    ps = []
    for _ in range(8):
        ps += [torch.empty(0, dtype=torch.bfloat16).new_empty((8192, 8192), device="meta"),
               torch.empty(0, dtype=torch.bfloat16).new_empty((8192,), device="meta")]
    plans = plan_buckets(ps)
    weight_buckets = [p for p in plans if any(ps[i].dim() == 2 for i in p.param_indices)]
    assert len(weight_buckets) == 8
    assert all(len([i for i in p.param_indices if ps[i].dim() == 2]) == 1 for p in weight_buckets)
