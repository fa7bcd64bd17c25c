"""In-tree build of libnnd_b200.so (the CUDA kernels + C ABI) with nvcc for sm_100a.

Replaces the reference's extension build dip/torch-nndistance/build.py:48-60
(setuptools CUDAExtension with no arch flags).  The library has no torch or
Python dependency: it is a plain C-ABI shared object (include/nnd_b200.h).
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
SOURCES = [PKG_DIR / "csrc" / "nnd_kernels.cu"]
HEADERS = [REPO_ROOT / "include" / "nnd_b200.h", PKG_DIR / "csrc" / "nnd_sym.cuh", PKG_DIR / "csrc" / "nnd_symc.cuh", PKG_DIR / "csrc" / "nnd_grad.cuh",
           PKG_DIR / "csrc" / "nnd_grad_exact.cuh", PKG_DIR / "csrc" / "nnd_small.cuh", PKG_DIR / "csrc" / "nnd_hint.cuh",
           PKG_DIR / "csrc" / "nnd_pruned.cuh"]
LIB_PATH = PKG_DIR / "libnnd_b200.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libnnd_b200.so cannot be built (there is no CPU fallback)")


def is_stale() -> bool:
    if not LIB_PATH.exists():
        return True
    built = LIB_PATH.stat().st_mtime
    return any(p.stat().st_mtime > built for p in SOURCES + HEADERS)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile the CUDA library if it is missing or older than its sources."""
    if not force and not is_stale():
        return LIB_PATH
    cmd = [find_nvcc(), *NVCC_FLAGS, f"-I{REPO_ROOT / 'include'}", "-o", str(LIB_PATH)]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [str(s) for s in SOURCES]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if verbose:
        print(proc.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
