"""Strategy interface shared by the replicated (DDP) and sharded (FSDP-style) paths.

The three abstract methods and their signatures are the reference's
(/root/reference/src/dist_strategy/dist_strategy.py:8-26).  Two optional hooks
are added because the strategy — not the caller — knows where parameters and
optimizer state physically live (the reference builds the optimizer
on parameters FSDP later replaces):

* ``build_optimizer(model, spec)`` returns the optimizer that matches the
  wrapped model (fused multi-tensor step for DDP, shard-local fused
  reduce-scatter+update for the sharded strategy);
* ``adopt_optimizer(model, optimizer)`` re-targets a user-supplied
  ``torch.optim.Optimizer`` at the wrapped model's parameters.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from dataclasses import dataclass, field
from typing import Any, Dict, Tuple

import torch
from torch.nn import Module


@dataclass
class OptimizerSpec:
    """Optimizer description that a strategy can realise on its own storage."""

    name: str = "sgd"              # sgd | adamw
    lr: float = 1e-3
    weight_decay: float = 0.0
    momentum: float = 0.0
    betas: Tuple[float, float] = (0.9, 0.999)
    eps: float = 1e-8
    extra: Dict[str, Any] = field(default_factory=dict)

    @classmethod
    def from_torch(cls, optimizer: torch.optim.Optimizer) -> "OptimizerSpec":
        group = optimizer.param_groups[0]
        if isinstance(optimizer, torch.optim.AdamW) or isinstance(optimizer, torch.optim.Adam):
            return cls("adamw", lr=group["lr"], weight_decay=group.get("weight_decay", 0.0),
                       betas=tuple(group.get("betas", (0.9, 0.999))), eps=group.get("eps", 1e-8))
        if isinstance(optimizer, torch.optim.SGD):
            return cls("sgd", lr=group["lr"], weight_decay=group.get("weight_decay", 0.0),
                       momentum=group.get("momentum", 0.0))
        raise TypeError(f"Cannot derive an OptimizerSpec from {type(optimizer).__name__}")


class DistributedStrategy(ABC):
    @abstractmethod
    def prepare_model(self, model: Module, device: Any) -> Module:
        """Wrap ``model`` for this strategy (broadcast / shard / hook up)."""

    @abstractmethod
    def save_checkpoint(self, model: Module, state_dict: dict, path: str) -> None:
        """Write ``state_dict`` (``MODEL_STATE`` + ``EPOCHS_RUN``) to ``path``."""

    @abstractmethod
    def load_checkpoint(self, model: Module, checkpoint: dict, device: Any) -> Tuple[Module, int]:
        """Load ``checkpoint['MODEL_STATE']`` into ``model``; return it and the epoch."""

    # ---- additions -------------------------------------------------------
    #: True when ``model.state_dict()`` is a collective every rank must enter
    collective_state_dict: bool = False

    def build_optimizer(self, model: Module, spec: OptimizerSpec) -> Any:
        params = [p for p in model.parameters() if p.requires_grad]
        if spec.name == "adamw":
            return torch.optim.AdamW(params, lr=spec.lr, weight_decay=spec.weight_decay,
                                     betas=spec.betas, eps=spec.eps)
        if spec.name == "sgd":
            return torch.optim.SGD(params, lr=spec.lr, weight_decay=spec.weight_decay,
                                   momentum=spec.momentum)
        raise ValueError(f"Unknown optimizer {spec.name!r}")

    def adopt_optimizer(self, model: Module, optimizer: Any) -> Any:
        return optimizer

    def barrier(self) -> None:
        return None


def atomic_torch_save(obj: Any, path: str) -> None:
    """``torch.save`` to ``path`` via a temp file + rename (no torn snapshots)."""
    import os

    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)
