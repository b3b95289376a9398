"""Strategy ABC under the reference's module path."""
from distributed_training_b200.parallel.strategy import DistributedStrategy, OptimizerSpec  # noqa: F401
