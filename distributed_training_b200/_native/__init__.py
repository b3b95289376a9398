"""In-tree build and loading of ``libdtb200.so`` (the sm_100a kernel library).

``build()`` compiles every ``csrc/**/*.cu`` with
``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (object files under
``build/``) and links them into this directory, next to this file, so the
shared object travels with the source tree (it is git-ignored).  nvcc
cross-compiles without a GPU; the library itself links only the static CUDA
runtime, so it also *loads* on a GPU-less box.
"""
from __future__ import annotations

import concurrent.futures as _cf
import hashlib
import os
import shutil
import subprocess
from pathlib import Path
from typing import List, Optional

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent.parent
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
BUILD_DIR = REPO_ROOT / "build" / "dtb200"
LIB_PATH = PKG_DIR / "libdtb200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr", "--extended-lambda",
]


def nvcc_path() -> Optional[str]:
    cand = shutil.which("nvcc")
    if cand:
        return cand
    for p in ("/usr/local/cuda/bin/nvcc",):
        if os.path.exists(p):
            return p
    return None


def sources() -> List[Path]:
    return sorted(CSRC.rglob("*.cu"))


def _stamp(src: Path) -> str:
    h = hashlib.sha1()
    h.update(src.read_bytes())
    for hdr in sorted(list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list(INCLUDE.glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, verbose: bool) -> Path:
    rel = src.relative_to(CSRC)
    obj = BUILD_DIR / rel.with_suffix(".o")
    stamp_file = obj.with_suffix(".stamp")
    stamp = _stamp(src)
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return obj
    obj.parent.mkdir(parents=True, exist_ok=True)
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
    if verbose:
        print(" ".join(cmd), flush=True)
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{proc.stdout}\n{proc.stderr}")
    stamp_file.write_text(stamp)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile and link the library; returns its path.  Incremental unless ``force``."""
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build libdtb200.so")
    if force and BUILD_DIR.exists():
        shutil.rmtree(BUILD_DIR)
    srcs = sources()
    with _cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as pool:
        objs = list(pool.map(lambda s: _compile_one(nvcc, s, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        tmp = LIB_PATH.with_suffix(".so.tmp")
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(tmp),
               *[str(o) for o in objs], "-lcudart_static", "-ldl", "-lrt", "-lpthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        proc = subprocess.run(cmd, capture_output=True, text=True)
        if proc.returncode != 0:
            raise RuntimeError(f"link failed:\n{proc.stdout}\n{proc.stderr}")
        os.replace(tmp, LIB_PATH)
    return LIB_PATH


if __name__ == "__main__":
    import sys

    print(build(force="--force" in sys.argv, verbose=True))
