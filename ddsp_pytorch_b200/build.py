"""Builds libddsp_b200.so (hand-written sm_100a kernels + C ABI of include/ddsp_b200.h) in-tree with nvcc.

    python -m ddsp_pytorch_b200.build            # or: from ddsp_pytorch_b200.build import build_library

nvcc cross-compiles for sm_100a without a GPU; the resulting .so is git-ignored but travels with the tree.
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB_DIR = os.path.join(PKG, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libddsp_b200.so")
STAMP = LIB_PATH + ".srchash"

# no --use_fast_math: IEEE division / sqrt everywhere
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "-cudart", "static"]


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _source_hash() -> str:
    h = hashlib.sha256()
    for p in _sources() + sorted(glob.glob(os.path.join(CSRC, "*.cuh"))) + [os.path.join(PKG, "..", "include", "ddsp_b200.h")]:
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def nvcc_path() -> str:
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every csrc/*.cu into lib/libddsp_b200.so (skipped when sources are unchanged)."""
    want = _source_hash()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(STAMP) and open(STAMP).read().strip() == want:
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    cmd = [nvcc_path(), *NVCC_FLAGS, "-o", LIB_PATH, *_sources()]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    with open(STAMP, "w") as f:
        f.write(want)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
