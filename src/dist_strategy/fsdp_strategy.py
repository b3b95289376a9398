"""Sharded (FSDP-style) strategy under the reference's module path
(/root/reference/src/dist_strategy/fsdp_strategy.py:1-46)."""
from distributed_training_b200.parallel.sharded import FSDPStrategy, ShardedModule, ShardedOptimizer  # noqa: F401
