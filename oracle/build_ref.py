"""oracle/build_ref.py -- TEST INFRASTRUCTURE ONLY.

Compiles the reference's own CPU implementation of nnd, UNMODIFIED and where it lies
(/root/reference/dip/torch-nndistance/src/my_lib.cpp), into oracle/_ref/nnd_ref_cpu.so
(a pybind11 torch extension exporting nnd_forward / nnd_backward, my_lib.cpp:135-138).
Flags: -O2 -ffp-contract=off, no -march (FMA contraction changes dist bits).
No reference source is copied into this repository; oracle/_ref/ is git-ignored but
travels to the GPU box with the snapshot.
"""
from __future__ import annotations

import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_SRC = Path("/root/reference/dip/torch-nndistance/src/my_lib.cpp")
OUT_DIR = HERE / "_ref"
MODULE = "nnd_ref_cpu"


def build(verbose: bool = False) -> Path | None:
    """Returns the built .so, or None when /root/reference is absent (e.g. on the GPU box)."""
    out = OUT_DIR / f"{MODULE}.so"
    if not REF_SRC.exists():
        return out if out.exists() else None
    if out.exists() and out.stat().st_mtime >= REF_SRC.stat().st_mtime:
        return out
    from torch.utils.cpp_extension import load

    OUT_DIR.mkdir(exist_ok=True)
    build_dir = OUT_DIR / "build"
    build_dir.mkdir(exist_ok=True)
    load(name=MODULE, sources=[str(REF_SRC)], extra_cflags=["-O2", "-ffp-contract=off", "-w"],
         build_directory=str(build_dir), verbose=verbose)
    shutil.copy2(build_dir / f"{MODULE}.so", out)
    shutil.rmtree(build_dir, ignore_errors=True)
    return out


CUDA_MODULE = "nnd_ref_cuda"
REF_CUDA_SRCS = [REF_SRC.parent / "my_lib_cuda.cpp", REF_SRC.parent / "nnd_cuda.cu"]


def build_cuda(verbose: bool = False) -> Path | None:
    """The reference's own CUDA half (my_lib_cuda.cpp + nnd_cuda.cu), UNMODIFIED, recompiled for sm_100a with the
    flags its build.py:48-60 uses (torch CUDAExtension defaults, i.e. nvcc's default -fmad=true).  It is the
    "reference GPU kernel on the same box" comparison row, NOT a parity oracle: nvcc contracts
    its distance sum to FMA, so its dist bits differ from the reference CPU path.  Only
    bench.py --impl reference_cuda and tests/ load it."""
    import os

    out = OUT_DIR / f"{CUDA_MODULE}.so"
    if not all(p.exists() for p in REF_CUDA_SRCS):
        return out if out.exists() else None
    if out.exists() and out.stat().st_mtime >= max(p.stat().st_mtime for p in REF_CUDA_SRCS):
        return out
    from torch.utils.cpp_extension import load

    OUT_DIR.mkdir(exist_ok=True)
    build_dir = OUT_DIR / "build_cuda"
    build_dir.mkdir(exist_ok=True)
    old = os.environ.get("TORCH_CUDA_ARCH_LIST")
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    try:
        load(name=CUDA_MODULE, sources=[str(p) for p in REF_CUDA_SRCS], extra_cflags=["-O2", "-w"],
             extra_cuda_cflags=["-w"], build_directory=str(build_dir), verbose=verbose, is_python_module=False)
    finally:
        if old is None:
            os.environ.pop("TORCH_CUDA_ARCH_LIST", None)
        else:
            os.environ["TORCH_CUDA_ARCH_LIST"] = old
    shutil.copy2(build_dir / f"{CUDA_MODULE}.so", out)
    shutil.rmtree(build_dir, ignore_errors=True)
    return out


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
    print(build_cuda(verbose="-v" in sys.argv))
