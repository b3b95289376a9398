"""Import the UNMODIFIED reference package from ``baseline/_ref``.

The reference is imported as ``src.*`` (its pyproject ships ``packages=[{include="src"}]``) and
needs ``hydra`` / ``omegaconf`` at import time
(/root/reference/src/distributed_trainer.py:9,13).  Neither is installed on the image, so the
config shim registers stand-in modules first.  ``baseline/_ref`` is put at the front of
``sys.path`` and any already-imported ``src`` (this repository's compatibility package of the same
name) is evicted, so ``import src.distributed_trainer`` resolves to the reference's file.
"""
from __future__ import annotations

import importlib
import os
import sys
from pathlib import Path
from types import ModuleType
from typing import Optional

REF_ROOT = Path(__file__).resolve().parent / "_ref"


def reference_available() -> Optional[str]:
    """None when importable, else a one-line reason."""
    if not (REF_ROOT / "src" / "distributed_trainer.py").exists():
        return f"{REF_ROOT} missing: run baseline/install_reference.sh"
    return None


def load_reference() -> ModuleType:
    """Return the reference's ``src.distributed_trainer`` module."""
    why = reference_available()
    if why:
        raise ImportError(why)
    repo_root = str(Path(__file__).resolve().parent.parent)
    if repo_root not in sys.path:
        sys.path.append(repo_root)
    from distributed_training_b200 import hydra_compat

    hydra_compat.install_stub_modules()
    for name in [m for m in sys.modules if m == "src" or m.startswith("src.")]:
        del sys.modules[name]
    sys.path.insert(0, str(REF_ROOT))
    try:
        mod = importlib.import_module("src.distributed_trainer")
    finally:
        pass
    assert str(REF_ROOT) in str(Path(mod.__file__).resolve()), f"imported {mod.__file__}, not the reference"
    return mod
