"""Synthetic datasets, rank sharding and the batch loader.

Parity targets
--------------
* ``MyTrainDataset(size)`` yields ``(rand(20), rand(1))`` pairs like
  /root/reference/src/data_utils.py:7-16, but from a seeded generator so every
  rank sees the same dataset and keeps the samples in two dense
  tensors instead of a Python list of tuples.
* ``ShardedSampler`` produces exactly the index stream of
  ``torch.utils.data.DistributedSampler`` (shuffle by ``seed + epoch``,
  wrap-around padding to ``ceil(N / world)``, stride-``world`` subsampling),
  which is what the reference builds in ``prepare_dataloader``
  (/root/reference/src/distributed_trainer.py:204-211).
* ``ShardedLoader`` replaces ``DataLoader``: whole batches are gathered with one
  ``index_select`` into pinned staging buffers and copied with
  ``non_blocking=True`` one batch ahead of the consumer.
"""
from __future__ import annotations

import math
from typing import Iterator, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch.utils.data import Dataset, Sampler


class TensorPairDataset(Dataset):
    """Map-style dataset backed by two dense tensors (inputs, targets)."""

    inputs: torch.Tensor
    targets: torch.Tensor

    def __len__(self) -> int:
        return int(self.inputs.shape[0])

    def __getitem__(self, index: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.inputs[index], self.targets[index]

    def tensors(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.inputs, self.targets


class MyTrainDataset(TensorPairDataset):
    """Uniform-random regression pairs: 20 features -> 1 target."""

    def __init__(self, size: int, seed: Optional[int] = 0, input_size: int = 20,
                 output_size: int = 1) -> None:
        self.size = int(size)
        gen = None
        if seed is not None:
            gen = torch.Generator().manual_seed(int(seed))
        self.inputs = torch.rand(self.size, input_size, generator=gen)
        self.targets = torch.rand(self.size, output_size, generator=gen)

    @property
    def data(self) -> List[Tuple[torch.Tensor, torch.Tensor]]:
        """List-of-tuples view, the storage format the reference exposes."""
        return [(self.inputs[i], self.targets[i]) for i in range(self.size)]


class LinearTeacherDataset(TensorPairDataset):
    """Regression targets from a fixed random linear teacher plus noise.

    Used by tests and the plumbing config so that "loss goes down" is a
    meaningful assertion (pure-noise targets only converge to the mean).
    """

    def __init__(self, size: int, input_size: int = 20, output_size: int = 1,
                 noise: float = 0.01, seed: int = 0) -> None:
        gen = torch.Generator().manual_seed(int(seed))
        self.inputs = torch.randn(size, input_size, generator=gen)
        teacher = torch.randn(input_size, output_size, generator=gen) / math.sqrt(input_size)
        self.targets = self.inputs @ teacher + noise * torch.randn(size, output_size, generator=gen)


class WideDataset(TensorPairDataset):
    """``[size, features]`` activations with integer class targets (wide-MLP config)."""

    def __init__(self, size: int, features: int = 8192, num_classes: int = 8192,
                 dtype: torch.dtype = torch.bfloat16, seed: int = 0) -> None:
        gen = torch.Generator().manual_seed(int(seed))
        self.inputs = torch.randn(size, features, generator=gen).to(dtype)
        self.targets = torch.randint(0, num_classes, (size,), generator=gen, dtype=torch.int64)


class TokenDataset(TensorPairDataset):
    """Synthetic token sequences; targets are the inputs shifted by one."""

    def __init__(self, size: int, seq_len: int = 1024, vocab_size: int = 50257,
                 seed: int = 0) -> None:
        gen = torch.Generator().manual_seed(int(seed))
        stream = torch.randint(0, vocab_size, (size, seq_len + 1), generator=gen, dtype=torch.int64)
        self.inputs = stream[:, :-1].contiguous()
        self.targets = stream[:, 1:].contiguous()


class ShardedSampler(Sampler[int]):
    """Rank-strided shard of a (optionally shuffled) permutation.

    Index-for-index compatible with ``DistributedSampler`` for the same
    ``(num_replicas, rank, shuffle, seed, drop_last, epoch)``.
    """

    def __init__(self, dataset: Sequence, num_replicas: Optional[int] = None,
                 rank: Optional[int] = None, shuffle: bool = True, seed: int = 0,
                 drop_last: bool = False) -> None:
        if num_replicas is None:
            num_replicas = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        if not 0 <= rank < num_replicas:
            raise ValueError(f"rank {rank} outside [0, {num_replicas})")
        self.dataset = dataset
        self.num_replicas = int(num_replicas)
        self.rank = int(rank)
        self.shuffle = shuffle
        self.seed = int(seed)
        self.drop_last = drop_last
        self.epoch = 0
        n = len(dataset)
        if drop_last and n % self.num_replicas:
            self.num_samples = n // self.num_replicas
        else:
            self.num_samples = (n + self.num_replicas - 1) // self.num_replicas
        self.total_size = self.num_samples * self.num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def indices(self) -> torch.Tensor:
        n = len(self.dataset)
        if self.shuffle:
            gen = torch.Generator().manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=gen)
        else:
            order = torch.arange(n)
        if self.drop_last:
            order = order[: self.total_size]
        elif self.total_size > n:
            reps = math.ceil((self.total_size - n) / max(n, 1))
            order = torch.cat([order, order.repeat(reps)[: self.total_size - n]])
        return order[self.rank: self.total_size: self.num_replicas]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return self.num_samples


class ShardedLoader:
    """Batch iterator over a :class:`TensorPairDataset` shard.

    On CUDA the next batch is gathered into pinned host memory and its
    host-to-device copy is issued on a side stream while the current batch is
    being consumed; the consumer stream waits on the copy's event, never on
    the host.  ``device=None`` yields CPU batches (the ``Trainer`` moves them).
    """

    def __init__(self, dataset: Dataset, batch_size: int, sampler: Optional[ShardedSampler] = None,
                 device: Optional[torch.device] = None, drop_last: bool = False,
                 pin_memory: Optional[bool] = None) -> None:
        self.dataset = dataset
        self.batch_size = int(batch_size)
        self.sampler = sampler if sampler is not None else ShardedSampler(dataset, 1, 0, shuffle=False)
        self.device = torch.device(device) if device is not None else None
        self.drop_last = drop_last
        cuda = self.device is not None and self.device.type == "cuda"
        self.pin_memory = cuda if pin_memory is None else (pin_memory and torch.cuda.is_available())
        self._copy_stream = torch.cuda.Stream(self.device) if cuda else None
        self._staging: List[Optional[Tuple[torch.Tensor, torch.Tensor]]] = [None, None]

    def __len__(self) -> int:
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _gather(self, idx: torch.Tensor, slot: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if hasattr(self.dataset, "tensors"):
            xs, ys = self.dataset.tensors()  # type: ignore[attr-defined]
            if self.pin_memory and idx.numel() == self.batch_size:
                if self._staging[slot] is None:
                    self._staging[slot] = (
                        torch.empty((self.batch_size, *xs.shape[1:]), dtype=xs.dtype).pin_memory(),
                        torch.empty((self.batch_size, *ys.shape[1:]), dtype=ys.dtype).pin_memory(),
                    )
                sx, sy = self._staging[slot]  # type: ignore[misc]
                torch.index_select(xs, 0, idx, out=sx)
                torch.index_select(ys, 0, idx, out=sy)
                return sx, sy
            return xs.index_select(0, idx), ys.index_select(0, idx)
        items = [self.dataset[int(i)] for i in idx]
        return torch.stack([it[0] for it in items]), torch.stack([it[1] for it in items])

    def _to_device(self, batch: Tuple[torch.Tensor, torch.Tensor]):
        if self.device is None:
            return batch, None
        if self._copy_stream is None:
            return (batch[0].to(self.device), batch[1].to(self.device)), None
        with torch.cuda.stream(self._copy_stream):
            out = (batch[0].to(self.device, non_blocking=True),
                   batch[1].to(self.device, non_blocking=True))
            event = torch.cuda.Event()
            event.record(self._copy_stream)
        return out, event

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        order = self.sampler.indices() if hasattr(self.sampler, "indices") else torch.tensor(list(self.sampler))
        chunks = list(order.split(self.batch_size))
        if self.drop_last and chunks and chunks[-1].numel() < self.batch_size:
            chunks.pop()
        pending = None
        for i, idx in enumerate(chunks):
            if self._copy_stream is not None and self._staging[i % 2] is not None:
                # the staging slot is reused every second batch: its previous copy must be done
                self._copy_stream.synchronize()
            nxt = self._to_device(self._gather(idx, i % 2))
            if pending is not None:
                yield self._release(pending)
            pending = nxt
        if pending is not None:
            yield self._release(pending)

    def _release(self, pending):
        batch, event = pending
        if event is not None:
            torch.cuda.current_stream(self.device).wait_event(event)
            for t in batch:
                t.record_stream(torch.cuda.current_stream(self.device))
        return batch
