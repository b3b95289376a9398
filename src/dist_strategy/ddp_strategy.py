"""Hand-rolled DDP strategy under the reference's module path."""
from distributed_training_b200.parallel.ddp import DDPStrategy, ReplicatedModule  # noqa: F401
