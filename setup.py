"""Build step of the package (see pyproject.toml): compiles libnnd_b200.so in-tree with nvcc for sm_100a
(pointcloudregistration_b200/_build.py -- the same function __graft_entry__.build() and the tests call) before the
Python files are collected, and ships the C header next to the library.  Replaces the reference's
dip/torch-nndistance/build.py:48-60 (CUDAExtension without arch flags)."""
import shutil
import sys
from pathlib import Path

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent


class build_py_with_cuda(build_py):
    def run(self):
        sys.path.insert(0, str(ROOT))
        from pointcloudregistration_b200 import _build

        lib = _build.build_library(force=False)
        print(f"[setup] {lib}")
        inc = ROOT / "pointcloudregistration_b200" / "include"
        inc.mkdir(exist_ok=True)
        shutil.copy2(ROOT / "include" / "nnd_b200.h", inc / "nnd_b200.h")  # installed copy of the C ABI
        super().run()


setup(cmdclass={"build_py": build_py_with_cuda})
