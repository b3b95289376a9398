"""Hand-rolled DDP strategy under the reference's module path
(/root/reference/src/dist_strategy/ddp_strategy.py:1-32)."""
from distributed_training_b200.parallel.ddp import DDPStrategy, ReplicatedModule  # noqa: F401
