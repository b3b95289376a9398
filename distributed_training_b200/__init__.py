"""distributed_training_b200 — B200-native data-parallel training framework.

Capabilities and API surface follow erfanMhi/distributed_training (Hydra
entry point, ``Trainer`` with epochs / ``save_every`` / snapshot resume, a
``DistributedStrategy`` abstraction with a hand-rolled DDP and an FSDP-style
sharded strategy, ``DistributedSampler``-style sharding, the
``{"MODEL_STATE", "EPOCHS_RUN"}`` snapshot layout); the compute and
communication hot paths are hand-written sm_100a CUDA kernels
(``csrc/`` → ``_native/libdtb200.so``).
"""
__version__ = "0.1.0"

from .parallel.strategy import DistributedStrategy, OptimizerSpec  # noqa: F401
from .parallel.ddp import DDPStrategy, ReplicatedModule  # noqa: F401
from .parallel.sharded import FSDPStrategy, ShardedModule  # noqa: F401
