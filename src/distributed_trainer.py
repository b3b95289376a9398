"""``python -m src.distributed_trainer`` / ``torchrun -m src.distributed_trainer``.

Same module path and public names as the reference entry point
(/root/reference/src/distributed_trainer.py:29-280); the implementation lives in
:mod:`distributed_training_b200.trainer`.
"""
from distributed_training_b200.trainer import (  # noqa: F401
    DistributedEnvironment,
    ModelCheckpoint,
    Trainer,
    TrainingConfig,
    build_dataset,
    build_model,
    load_train_objs,
    main,
    prepare_dataloader,
    run,
    setup_logging,
    training_config_from,
)

if __name__ == "__main__":
    main()
