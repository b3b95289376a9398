"""GPT-2 (small 124M / XL 1.5B) for BASELINE.json configs 3 and 4.

Not part of the reference (its only model is ``nn.Linear(20, 1)``,
/root/reference/src/distributed_trainer.py:199); it is the transformer the
north star names.  Architecture: learned token + position embeddings, pre-LN
blocks (LN → causal MHA → residual, LN → MLP(4x, tanh-GELU) → residual), final
LN, LM head tied to the token embedding, no dropout.

``forward`` is plain PyTorch (autograd, CPU-capable).  ``forward_backward`` is
the B200 path: every matmul is the tcgen05 GEMM with fused bias / GELU /
residual epilogues, LayerNorm(+residual), attention, embedding and
cross-entropy are the library's own kernels, and backward is written out by
hand so that weight gradients land in the DDP buckets (``grad_sink``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List

import torch
import torch.nn.functional as F
from torch import nn

from ..ops import native


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    layer_norm_eps: float = 1e-5

    @property
    def padded_vocab(self) -> int:
        return (self.vocab_size + 127) // 128 * 128

    @classmethod
    def small(cls) -> "GPT2Config":
        return cls()

    @classmethod
    def xl(cls) -> "GPT2Config":
        return cls(n_embd=1600, n_layer=48, n_head=25)


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config, dtype: torch.dtype) -> None:
        super().__init__()
        c = cfg.n_embd
        self.n_head = cfg.n_head
        self.ln_1 = nn.LayerNorm(c, eps=cfg.layer_norm_eps, dtype=dtype)
        self.attn_qkv = nn.Linear(c, 3 * c, dtype=dtype)
        self.attn_proj = nn.Linear(c, c, dtype=dtype)
        self.ln_2 = nn.LayerNorm(c, eps=cfg.layer_norm_eps, dtype=dtype)
        self.mlp_fc = nn.Linear(c, 4 * c, dtype=dtype)
        self.mlp_proj = nn.Linear(4 * c, c, dtype=dtype)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, t, c = x.shape
        qkv = self.attn_qkv(self.ln_1(x))
        q, k, v = qkv.view(b, t, 3, self.n_head, c // self.n_head).permute(2, 0, 3, 1, 4)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + self.attn_proj(y.transpose(1, 2).reshape(b, t, c))
        return x + self.mlp_proj(F.gelu(self.mlp_fc(self.ln_2(x)), approximate="tanh"))


class GPT2(nn.Module):
    supports_sharded_native = True

    def __init__(self, cfg: GPT2Config, dtype: torch.dtype = torch.float32) -> None:
        super().__init__()
        self.cfg = cfg
        # the embedding table is padded to a multiple of 128 rows so the LM-head GEMM needs no edge tiles
        self.wte = nn.Embedding(cfg.padded_vocab, cfg.n_embd, dtype=dtype)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd, dtype=dtype)
        self.h = nn.ModuleList([Block(cfg, dtype) for _ in range(cfg.n_layer)])
        self.ln_f = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps, dtype=dtype)
        self.apply(self._init)
        for blk in self.h:  # GPT-2 residual-projection scaling
            for lin in (blk.attn_proj, blk.mlp_proj):
                nn.init.normal_(lin.weight, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))
        self._ws: Dict[Any, Any] = {}
        #: native step: do not keep gelu(z) per block, rebuild it in backward (-25 % activation memory for one extra
        #: elementwise kernel per block, ~2 % of a GPT-2 XL step).  The sharded strategy switches it on.
        self.recompute_gelu = False

    @staticmethod
    def _init(m: nn.Module) -> None:
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def fsdp_units(self) -> List[nn.Module]:
        return [self.wte, self.wpe, *self.h, self.ln_f]

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        """Logits over the *padded* vocabulary, ``[batch, seq, padded_vocab]``; padding logits are -inf-free
        garbage-free (they come from zero-gradient rows) and are masked out by the loss helpers."""
        b, t = tokens.shape
        pos = torch.arange(t, device=tokens.device)
        x = self.wte(tokens) + self.wpe(pos)[None]
        for blk in self.h:
            x = blk(x)
        x = self.ln_f(x)
        logits = x @ self.wte.weight.t()
        return logits[..., : self.cfg.vocab_size]

    def loss(self, tokens: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        logits = self.forward(tokens)
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), targets.view(-1))

    # ------------------------------------------------------------------ native
    def native_ready(self) -> bool:
        p = self.wte.weight
        return (p.is_cuda and p.dtype == torch.bfloat16 and native.available()
                and self.cfg.n_embd % 8 == 0 and (self.cfg.n_embd // self.cfg.n_head) == 64
                and hasattr(native, "attention_fwd"))

    def forward_backward(self, tokens: torch.Tensor, targets: torch.Tensor, loss: str = "cross_entropy",
                         grad_sink: Any = None, max_ctas: int = 0) -> torch.Tensor:
        from .gpt2_native import gpt2_forward_backward

        return gpt2_forward_backward(self, tokens, targets, grad_sink=grad_sink, max_ctas=max_ctas)
