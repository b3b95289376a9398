"""``python src/playground/ddp_script.py`` — the from-scratch DDP demo
(/root/reference/src/playground/ddp_script.py:1-260)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from distributed_training_b200.playground.ddp_script import (  # noqa: E402,F401
    DummyDataset,
    SimpleModel,
    cleanup,
    get_argument_parser,
    main,
    setup,
    setup_logging,
    train,
)

if __name__ == "__main__":
    main()
