"""Strategy ABC under the reference's module path (/root/reference/src/dist_strategy/dist_strategy.py:1-26)."""
from distributed_training_b200.parallel.strategy import DistributedStrategy, OptimizerSpec  # noqa: F401
