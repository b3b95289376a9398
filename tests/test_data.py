"""Dataset + sampler: parity with torch's DistributedSampler (SURVEY §4.3 unit tier, C9/C10)."""
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from distributed_training_b200.data import (MyTrainDataset, ShardedLoader, ShardedSampler, TokenDataset,
                                            WideDataset)


def test_dataset_shapes_and_seed():
    a, b = MyTrainDataset(64, seed=3), MyTrainDataset(64, seed=3)
    x, y = a[5]
    assert x.shape == (20,) and y.shape == (1,) and len(a) == 64
    assert torch.equal(a.inputs, b.inputs)                        # every rank builds the same data (Q12)
    assert 0.0 <= float(a.inputs.min()) and float(a.inputs.max()) <= 1.0
    assert len(a.data) == 64 and torch.equal(a.data[5][0], x)     # reference's list-of-tuples view


@pytest.mark.parametrize("n,world,shuffle,drop_last", [(2048, 2, True, False), (1000, 4, True, False),
                                                         (1001, 4, False, False), (1003, 4, True, True),
                                                         (7, 8, True, False)])
def test_sampler_matches_torch(n, world, shuffle, drop_last):
    data = list(range(n))
    for epoch in (0, 3):
        for rank in range(world):
            ref = DistributedSampler(data, num_replicas=world, rank=rank, shuffle=shuffle, seed=11,
                                     drop_last=drop_last)
            ours = ShardedSampler(data, num_replicas=world, rank=rank, shuffle=shuffle, seed=11,
                                  drop_last=drop_last)
            ref.set_epoch(epoch)
            ours.set_epoch(epoch)
            assert list(ref) == list(ours) and len(ref) == len(ours)


def test_reference_default_step_count():
    ds = MyTrainDataset(2048)
    loader = ShardedLoader(ds, 32, sampler=ShardedSampler(ds, num_replicas=2, rank=0))
    assert len(loader.sampler) == 1024 and len(loader) == 32      # verified value in SURVEY C10


def test_loader_covers_shard_exactly_once():
    ds = MyTrainDataset(100, seed=1)
    seen = []
    for rank in range(4):
        sampler = ShardedSampler(ds, num_replicas=4, rank=rank, shuffle=True, seed=5)
        sampler.set_epoch(2)
        for xb, yb in ShardedLoader(ds, 8, sampler=sampler):
            assert xb.shape[1:] == (20,) and yb.shape[1:] == (1,)
            seen.append(xb)
    allx = torch.cat(seen)
    assert allx.shape[0] == 100
    assert torch.equal(allx.sum(0), ds.inputs.sum(0)) or torch.allclose(allx.sum(0), ds.inputs.sum(0))


def test_token_and_wide_datasets():
    t = TokenDataset(4, seq_len=16, vocab_size=50)
    assert t.inputs.shape == (4, 16) and torch.equal(t.inputs[:, 1:], t.targets[:, :-1])
    w = WideDataset(4, features=64, num_classes=10)
    assert w.inputs.dtype == torch.bfloat16 and w.targets.dtype == torch.int64 and int(w.targets.max()) < 10
