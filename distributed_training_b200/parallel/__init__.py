"""Parallel strategies, gradient bucketing and communicators."""
from .buckets import plan_buckets  # noqa: F401
from .comm import Communicator, SingleProcessComm, TorchDistComm, make_communicator  # noqa: F401
from .ddp import DDPStrategy, ReplicatedModule  # noqa: F401
from .sharded import FSDPStrategy, ShardedModule, ShardedOptimizer  # noqa: F401
from .strategy import DistributedStrategy, OptimizerSpec  # noqa: F401
