"""`torch_nndistance_aten`-shaped native-op layer on top of the C ABI.

Mirrors the pybind module the reference builds (my_lib_cuda.cpp:74-77 and the CPU
twin my_lib.cpp:135-138): caller-allocated outputs written in place, return 1 on
success (nnd_cuda.cu:161).  CHECK_INPUT semantics (is_cuda && is_contiguous,
my_lib_cuda.cpp:4-6) raise RuntimeError like AT_ASSERTM does.
"""
from __future__ import annotations

import torch

from . import _cabi
from .functional import _stream_ptr, _workspace


def _check_input(**tensors):
    for name, t in tensors.items():
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor")  # my_lib_cuda.cpp:4
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")  # my_lib_cuda.cpp:5


def _check_dtypes(f32=(), i32=()):
    for t in f32:
        if t.dtype != torch.float32:
            raise RuntimeError(f"expected float32 tensor, got {t.dtype}")  # .data<float>() throws, nnd_cuda.cu:144-147
    for t in i32:
        if t.dtype != torch.int32:
            raise RuntimeError(f"expected int32 tensor, got {t.dtype}")  # .data<int>() throws, nnd_cuda.cu:148-149


def nnd_forward_cuda(xyz1, xyz2, dist1, dist2, idx1, idx2) -> int:
    """my_lib_cuda.cpp:25-41.  Writes dist1, dist2, idx1, idx2 in place; returns 1."""
    _check_input(xyz1=xyz1, xyz2=xyz2, dist1=dist1, dist2=dist2, idx1=idx1, idx2=idx2)
    _check_dtypes(f32=(xyz1, xyz2, dist1, dist2), i32=(idx1, idx2))
    lib = _cabi.load()
    b, n, m = xyz1.size(0), xyz1.size(1), xyz2.size(1)  # nnd_cuda.cu:140-142
    dev = xyz1.device
    with torch.cuda.device(dev):
        ws = _workspace(lib.nnd_workspace_bytes(b, n, m), dev)
        rc = lib.nnd_forward_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), b, n, m, dist1.data_ptr(), dist2.data_ptr(),
                                      idx1.data_ptr(), idx2.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_forward_cuda")
    return 1


def nnd_backward_cuda(xyz1, xyz2, gradxyz1, gradxyz2, graddist1, graddist2, idx1, idx2) -> int:
    """my_lib_cuda.cpp:44-72.  Writes gradxyz1, gradxyz2 in place; returns 1."""
    _check_input(xyz1=xyz1, xyz2=xyz2, gradxyz1=gradxyz1, gradxyz2=gradxyz2, graddist1=graddist1,
                 graddist2=graddist2, idx1=idx1, idx2=idx2)
    _check_dtypes(f32=(xyz1, xyz2, gradxyz1, gradxyz2, graddist1, graddist2), i32=(idx1, idx2))
    lib = _cabi.load()
    b, n, m = xyz1.size(0), xyz1.size(1), xyz2.size(1)
    dev = xyz1.device
    with torch.cuda.device(dev):
        rc = lib.nnd_backward_cuda_f32(xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(), graddist2.data_ptr(),
                                       idx1.data_ptr(), idx2.data_ptr(), b, n, m, gradxyz1.data_ptr(),
                                       gradxyz2.data_ptr(), _stream_ptr(dev))
    _cabi.check(rc, "nnd_backward_cuda")
    return 1


def _no_cpu(name):
    def fn(*_args, **_kwargs):
        raise RuntimeError(
            f"{name}: this is the B200 build of torch_nndistance_aten; the CPU twin (my_lib.cpp:135-138) is not "
            "provided -- move the tensors to a CUDA device")
    fn.__name__ = name
    return fn


nnd_forward = _no_cpu("nnd_forward")
nnd_backward = _no_cpu("nnd_backward")
