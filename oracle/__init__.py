"""oracle -- TEST INFRASTRUCTURE ONLY (CPU checker for the nnd hot path).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
may import this package.  The product (pointcloudregistration_b200, torch_nndistance)
never does; it has no CPU path at all.

Two checkers, same results bit for bit (tests/test_oracle.py proves it):
  * `port`      : oracle/nnd_oracle.c, a plain-C restatement of my_lib.cpp:3-25,91-130
  * `reference` : oracle/_ref/nnd_ref_cpu.so, the reference's my_lib.cpp compiled unmodified
                  (present when built in the dev container; travels to the GPU box)
All functions take and return numpy arrays / CPU tensors; nothing here touches CUDA.
"""
from __future__ import annotations

import ctypes
import importlib.util
import subprocess
from ctypes import c_int, c_void_p
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
PORT_LIB = HERE / "libnnd_oracle.so"
REF_LIB = HERE / "_ref" / "nnd_ref_cpu.so"

_port = None
_ref = None


def build_port(force: bool = False) -> Path:
    src = HERE / "nnd_oracle.c"
    if force or not PORT_LIB.exists() or PORT_LIB.stat().st_mtime < src.stat().st_mtime:
        cmd = ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-o", str(PORT_LIB), str(src)]
        subprocess.run(cmd, check=True)
    return PORT_LIB


def port():
    global _port
    if _port is None:
        build_port()
        lib = ctypes.CDLL(str(PORT_LIB))
        lib.nnd_oracle_query_f32.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]
        lib.nnd_oracle_query_mt_f32.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int]
        lib.nnd_oracle_forward_f32.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int] + [c_void_p] * 4
        lib.nnd_oracle_forward_mt_f32.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int] + [c_void_p] * 4 + [c_int]
        lib.nnd_oracle_backward_f32.argtypes = [c_void_p] * 6 + [c_int, c_int, c_int, c_void_p, c_void_p]
        for f in ("nnd_oracle_query_f32", "nnd_oracle_query_mt_f32", "nnd_oracle_forward_f32",
                  "nnd_oracle_forward_mt_f32", "nnd_oracle_backward_f32"):
            getattr(lib, f).restype = c_int
        _port = lib
    return _port


def have_reference() -> bool:
    return REF_LIB.exists()


def reference():
    """The reference's own compiled my_lib.cpp (pybind module with nnd_forward / nnd_backward)."""
    global _ref
    if _ref is None:
        if not REF_LIB.exists():
            raise FileNotFoundError(f"{REF_LIB} not built (python oracle/build_ref.py, needs /root/reference)")
        import torch  # noqa: F401  (the extension links against libtorch)

        spec = importlib.util.spec_from_file_location("nnd_ref_cpu", str(REF_LIB))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _ref = mod
    return _ref


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def _p(a):
    return a.ctypes.data_as(c_void_p)


def forward(xyz1, xyz2, threads: int = 1):
    """port: dist1[b,n], dist2[b,m] float32, idx1, idx2 int32 (numpy)."""
    xyz1, xyz2 = _f32(xyz1), _f32(xyz2)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    assert xyz2.shape[0] == b and xyz1.shape[2] == 3 and xyz2.shape[2] == 3
    d1 = np.empty((b, n), np.float32)
    d2 = np.empty((b, m), np.float32)
    i1 = np.empty((b, n), np.int32)
    i2 = np.empty((b, m), np.int32)
    if threads > 1:
        port().nnd_oracle_forward_mt_f32(_p(xyz1), _p(xyz2), b, n, m, _p(d1), _p(d2), _p(i1), _p(i2), threads)
    else:
        port().nnd_oracle_forward_f32(_p(xyz1), _p(xyz2), b, n, m, _p(d1), _p(d2), _p(i1), _p(i2))
    return d1, d2, i1, i2


def query(q, t, threads: int = 1):
    """port, one direction: dist[b,nq], idx[b,nq]."""
    q, t = _f32(q), _f32(t)
    b, nq, _ = q.shape
    nt = t.shape[1]
    d = np.empty((b, nq), np.float32)
    i = np.empty((b, nq), np.int32)
    port().nnd_oracle_query_mt_f32(_p(q), _p(t), b, nq, nt, _p(d), _p(i), max(1, threads))
    return d, i


def backward(xyz1, xyz2, graddist1, graddist2, idx1, idx2):
    """port: gradxyz1[b,n,3], gradxyz2[b,m,3]."""
    xyz1, xyz2 = _f32(xyz1), _f32(xyz2)
    gd1, gd2 = _f32(graddist1), _f32(graddist2)
    idx1 = np.ascontiguousarray(idx1, dtype=np.int32)
    idx2 = np.ascontiguousarray(idx2, dtype=np.int32)
    b, n, _ = xyz1.shape
    m = xyz2.shape[1]
    g1 = np.empty((b, n, 3), np.float32)
    g2 = np.empty((b, m, 3), np.float32)
    port().nnd_oracle_backward_f32(_p(xyz1), _p(xyz2), _p(gd1), _p(gd2), _p(idx1), _p(idx2), b, n, m, _p(g1), _p(g2))
    return g1, g2


def reference_forward(xyz1, xyz2):
    """reference my_lib.cpp:28-60 through its own pybind entry point. numpy in / numpy out."""
    import torch

    ref = reference()
    x1 = torch.from_numpy(_f32(xyz1))
    x2 = torch.from_numpy(_f32(xyz2))
    b, n, _ = x1.shape
    m = x2.shape[1]
    d1 = torch.zeros(b, n)
    d2 = torch.zeros(b, m)
    i1 = torch.zeros(b, n, dtype=torch.int32)
    i2 = torch.zeros(b, m, dtype=torch.int32)
    ref.nnd_forward(x1, x2, d1, d2, i1, i2)
    return d1.numpy(), d2.numpy(), i1.numpy(), i2.numpy()


def reference_backward(xyz1, xyz2, graddist1, graddist2, idx1, idx2):
    """reference my_lib.cpp:64-133 through its own pybind entry point."""
    import torch

    ref = reference()
    x1 = torch.from_numpy(_f32(xyz1))
    x2 = torch.from_numpy(_f32(xyz2))
    g1 = torch.zeros_like(x1)
    g2 = torch.zeros_like(x2)
    ref.nnd_backward(x1, x2, g1, g2, torch.from_numpy(_f32(graddist1)), torch.from_numpy(_f32(graddist2)),
                     torch.from_numpy(np.ascontiguousarray(idx1, dtype=np.int32)),
                     torch.from_numpy(np.ascontiguousarray(idx2, dtype=np.int32)))
    return g1.numpy(), g2.numpy()
