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


class TokenPairDataset(Dataset):
    """Synthetic next-token pairs: ``size`` sequences of ``seq_len`` tokens, targets = inputs shifted by one."""

    def __init__(self, size: int, seq_len: int = 1024, vocab: int = 50257, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        stream = torch.randint(0, vocab, (size, seq_len + 1), generator=g, dtype=torch.int64)
        self.x = stream[:, :-1].contiguous()
        self.y = stream[:, 1:].contiguous()

    def __len__(self) -> int:
        return int(self.x.shape[0])

    def __getitem__(self, i: int):
        return self.x[i], self.y[i]


class _TorchBlock(nn.Module):
    def __init__(self, c: int, heads: int, dtype: torch.dtype) -> None:
        super().__init__()
        self.heads = heads
        self.ln_1 = nn.LayerNorm(c, dtype=dtype)
        self.qkv = nn.Linear(c, 3 * c, dtype=dtype)
        self.proj = nn.Linear(c, c, dtype=dtype)
        self.ln_2 = nn.LayerNorm(c, dtype=dtype)
        self.fc = nn.Linear(c, 4 * c, dtype=dtype)
        self.fc_proj = nn.Linear(4 * c, c, dtype=dtype)

    def forward(self, x):
        b, t, c = x.shape
        q, k, v = self.qkv(self.ln_1(x)).view(b, t, 3, self.heads, c // self.heads).permute(2, 0, 3, 1, 4)
        y = nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + self.proj(y.transpose(1, 2).reshape(b, t, c))
        return x + self.fc_proj(nn.functional.gelu(self.fc(self.ln_2(x)), approximate="tanh"))


class TorchGPT2(nn.Module):
    """Plain eager GPT-2 (tied LM head, vocab padded to 50304 like the product model).  ``forward`` returns
    logits as [batch, vocab, seq] so that ``F.cross_entropy(output, targets)`` — the loss hard-wired in the
    reference's ``Trainer._run_batch`` — is the language-model loss for [batch, seq] integer targets."""

    def __init__(self, n_embd=768, n_layer=12, n_head=12, vocab=50304, n_positions=1024, dtype=torch.bfloat16):
        super().__init__()
        self.wte = nn.Embedding(vocab, n_embd, dtype=dtype)
        self.wpe = nn.Embedding(n_positions, n_embd, dtype=dtype)
        self.h = nn.ModuleList([_TorchBlock(n_embd, n_head, dtype) for _ in range(n_layer)])
        self.ln_f = nn.LayerNorm(n_embd, dtype=dtype)

    def forward(self, tokens):
        t = tokens.shape[1]
        x = self.wte(tokens) + self.wpe(torch.arange(t, device=tokens.device))[None]
        for blk in self.h:
            x = blk(x)
        return (self.ln_f(x) @ self.wte.weight.t()).transpose(1, 2)
