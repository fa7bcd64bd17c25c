import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def cuda_lib():
    """Build (if stale) and load libnnd_b200.so; GPU tests fail loudly if it is not there."""
    from pointcloudregistration_b200 import _build, _cabi

    if _build.is_stale():
        _build.build_library()
    return _cabi.load()


@pytest.fixture(scope="session")
def device():
    import torch

    assert torch.cuda.is_available(), "GPU tests need CUDA; run them with `-m gpu` on the B200 box"
    return torch.device("cuda:0")
