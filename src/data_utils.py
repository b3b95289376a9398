"""Datasets and sharding under the reference's module name (``src.data_utils``,
/root/reference/src/data_utils.py:1-16)."""
from distributed_training_b200.data import (  # noqa: F401
    LinearTeacherDataset,
    MyTrainDataset,
    ShardedLoader,
    ShardedSampler,
    TokenDataset,
    WideDataset,
)
