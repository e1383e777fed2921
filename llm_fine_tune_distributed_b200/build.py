"""Builds libb200sft.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

The library is plain CUDA + the CUDA runtime: no torch headers, no pybind.  nvcc cross-compiles
without a GPU, so this runs in the CPU-only build container as well as on the B200 box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = PKG_DIR / "libb200sft.so"
SOURCES = ["flce.cu", "elementwise.cu", "rmsnorm.cu", "nvl_allreduce.cu", "optim.cu", "embed_grad.cu"]
HEADERS = ["sm100_ptx.cuh", "gemm_engine.cuh", "capi_common.h", "../../include/b200sft.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-t", "4",
    "-Xcompiler", "-fPIC", "-shared",
    "--cudart", "static",  # self-contained: shares the primary context (streams, pointers) with torch
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: set NVCC or put /usr/local/cuda/bin on PATH")


def _sources() -> list[Path]:
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    built = LIB_PATH.stat().st_mtime
    deps = _sources() + [CSRC / h for h in HEADERS]
    return any(d.exists() and d.stat().st_mtime > built for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA translation unit into one shared library; returns its path."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", str(LIB_PATH), *map(str, _sources())]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed ({res.returncode}):\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr, file=sys.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
