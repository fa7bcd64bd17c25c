"""CPU tests of the host-side mirror of the reference interface (names, validation, error behaviour)."""
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def test_drop_in_names_exist():
    import torch_nndistance as NND
    import torch_nndistance_aten as my_lib

    assert NND.__version__ == "1.0.0"  # reference torch_nndistance/__init__.py:1
    assert callable(NND.nnd) and issubclass(NND.NNDFunction, torch.autograd.Function)
    assert hasattr(NND.NNDFunction, "forward") and hasattr(NND.NNDFunction, "backward")
    assert isinstance(NND.NNDModule(), torch.nn.Module)
    for name in ("nnd_forward_cuda", "nnd_backward_cuda", "nnd_forward", "nnd_backward"):
        assert callable(getattr(my_lib, name))  # my_lib_cuda.cpp:74-77, my_lib.cpp:135-138


def test_cpu_tensors_are_rejected_not_silently_computed():
    import torch_nndistance as NND
    import torch_nndistance_aten as my_lib

    a, b = torch.rand(1, 8, 3), torch.rand(1, 9, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        NND.nnd(a, b)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        my_lib.nnd_forward_cuda(a, b, torch.zeros(1, 8), torch.zeros(1, 9), torch.zeros(1, 8, dtype=torch.int32),
                                torch.zeros(1, 9, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="B200 build"):
        my_lib.nnd_forward(a, b, None, None, None, None)


@pytest.mark.parametrize("bad, exc", [
    (lambda: (torch.rand(8, 3), torch.rand(1, 8, 3)), ValueError),
    (lambda: (torch.rand(1, 8, 4), torch.rand(1, 8, 3)), ValueError),
    (lambda: (torch.rand(1, 8, 3).double(), torch.rand(1, 8, 3)), TypeError),
    (lambda: ([[0.0, 0.0, 0.0]], torch.rand(1, 8, 3)), TypeError),
])
def test_shape_and_dtype_validation(bad, exc):
    from pointcloudregistration_b200.functional import _check_clouds

    with pytest.raises(exc):
        _check_clouds(*bad())


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from pointcloudregistration_b200 import _cabi

    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", tmp_path / "libnnd_b200.so")
    with pytest.raises(_cabi.NNDLibraryError, match="no CPU fallback"):
        _cabi.load()


def test_product_never_imports_the_oracle():
    import pathlib
    import re

    root = pathlib.Path(__file__).resolve().parent.parent
    files = list((root / "pointcloudregistration_b200").rglob("*.py")) + list((root / "torch_nndistance").rglob("*.py"))
    files.append(root / "torch_nndistance_aten.py")
    for f in files:
        src = f.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"


def test_exact_backward_switch_follows_torch_deterministic_mode():
    """functional.set_exact_backward: None follows torch.use_deterministic_algorithms, True/False override it."""
    import torch

    from pointcloudregistration_b200 import functional as F

    before = torch.are_deterministic_algorithms_enabled()
    try:
        F.set_exact_backward(None)
        torch.use_deterministic_algorithms(False)
        assert not F.exact_backward_enabled()
        torch.use_deterministic_algorithms(True)
        assert F.exact_backward_enabled()
        F.set_exact_backward(False)
        assert not F.exact_backward_enabled()
        torch.use_deterministic_algorithms(False)
        F.set_exact_backward(True)
        assert F.exact_backward_enabled()
    finally:
        F.set_exact_backward(None)
        torch.use_deterministic_algorithms(before)


def test_library_override_needs_the_debug_flag(tmp_path):
    """NND_B200_LIB alone must not redirect the product to another library; with
    NND_B200_DEBUG=1 it does, for timing experimental builds of the same ABI."""
    import os
    import subprocess
    import sys

    code = "from pointcloudregistration_b200 import _cabi; print(_cabi.LIB_PATH)"
    env = dict(os.environ, NND_B200_LIB=str(tmp_path / "other.so"), PYTHONPATH=str(ROOT))
    env.pop("NND_B200_DEBUG", None)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout
    assert out.strip().endswith("pointcloudregistration_b200/libnnd_b200.so")
    env["NND_B200_DEBUG"] = "1"
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout
    assert out.strip() == str(tmp_path / "other.so")


def test_packaging_metadata_names_the_drop_in_modules():
    """pyproject.toml installs the three import names the reference's `python build.py install` provided."""
    text = (ROOT / "pyproject.toml").read_text()
    for name in ("pointcloudregistration_b200", "torch_nndistance", "torch_nndistance_aten", "libnnd_b200.so"):
        assert name in text
    assert "build_library" in (ROOT / "setup.py").read_text()


def test_peer_exchange_is_cuda_only_and_validates_without_a_gpu(cuda_lib):
    """The exchange over peer memory needs CUDA tensors (the CPU `gloo` tests keep the collectives); its C entry points
    reject bad rank / world / pointer arguments before touching a device."""
    import torch

    from pointcloudregistration_b200 import sharded

    x = torch.zeros(1, 16, 3)
    assert sharded.PeerExchange.get(x, x, None, 2, 0, sharded.CudaCompute) is None
    assert cuda_lib.nnd_peer_max_ranks() >= 8
    assert cuda_lib.nnd_sum_p2p_cuda_f32(None, 2, 0, 16, None) == 10001
    assert cuda_lib.nnd_sum_p2p_cuda_f32(None, 0, 0, 16, None) == 10001
    assert cuda_lib.nnd_exchange_cols_p2p_cuda_f32(None, None, None, None, None, 2, 5, None, None, 1, 8, 8, 0, 8, None) == 10001
    assert cuda_lib.nnd_exchange_cols_p2p_cuda_f32(None, None, None, None, None, 99, 0, None, None, 1, 8, 8, 0, 8, None) == 10001
