"""Plain-PyTorch model/dataset definitions for the reference arm (no code from the product path).

The reference ships only ``nn.Linear(20, 1)``; BASELINE.json's larger configs need a torch model of the
named architecture to hand to the reference's ``Trainer``."""
from __future__ import annotations

import torch
from torch import nn
from torch.utils.data import Dataset


def wide_mlp_torch(width: int = 8192, depth: int = 8, dtype: torch.dtype = torch.bfloat16) -> nn.Module:
    layers = []
    for i in range(depth):
        layers.append(nn.Linear(width, width, dtype=dtype))
        if i != depth - 1:
            layers.append(nn.GELU(approximate="tanh"))
    return nn.Sequential(*layers)


class PooledClassificationDataset(Dataset):
    """``size`` samples drawn (modulo) from a small pool, so any step count fits in host memory."""

    def __init__(self, size: int, features: int, classes: int, dtype: torch.dtype, pool: int = 8192, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.size = size
        self.x = torch.randn(pool, features, generator=g).to(dtype)
        self.y = torch.randint(0, classes, (pool,), generator=g, dtype=torch.int64)

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, i: int):
        j = i % self.x.shape[0]
        return self.x[j], self.y[j]
