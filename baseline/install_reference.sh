#!/usr/bin/env bash
# One-time offline install of the UNMODIFIED reference into baseline/_ref (git-ignored).
#
# The task's stock command
#   python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
#          --target baseline/_ref /root/reference
# fails on this image because the reference's build backend (poetry-core) is not in the offline
# wheelhouse ("ModuleNotFoundError: No module named 'poetry'").  /root/reference is read-only, so
# the install runs from a copy under /tmp whose *packaging metadata only* is swapped for a
# setuptools stanza; every file under src/ is installed byte-for-byte (verified with diff below).
# --no-deps: hydra-core is not installable offline; the baseline harness provides the Hydra shim.
set -euo pipefail
REPO="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
REF="${1:-/root/reference}"
TMP="$(mktemp -d /tmp/refcopy.XXXXXX)"
cp -r "$REF"/. "$TMP"/
cat > "$TMP/pyproject.toml" <<'TOML'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
TOML
cat > "$TMP/setup.py" <<'PY'
from setuptools import find_namespace_packages, setup

setup(name="distributed-training", version="0.1.0",
      description="Distributed training with PyTorch DDP",
      packages=find_namespace_packages(include=["src", "src.*"]))
PY
rm -rf "$REPO/baseline/_ref"
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$REPO/baseline/_ref" "$TMP"
mkdir -p "$REPO/baseline/_ref/conf"
cp -r "$REF/conf/." "$REPO/baseline/_ref/conf/"
diff -r -x __pycache__ "$REF/src" "$REPO/baseline/_ref/src" && echo "reference sources installed unmodified"
rm -rf "$TMP"
