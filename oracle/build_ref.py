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


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
