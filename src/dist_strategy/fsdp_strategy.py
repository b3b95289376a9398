"""Sharded (FSDP-style) strategy under the reference's module path."""
from distributed_training_b200.parallel.sharded import FSDPStrategy, ShardedModule, ShardedOptimizer  # noqa: F401
