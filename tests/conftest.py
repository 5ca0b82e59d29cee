import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def lib():
    """libsmi.so, built on demand (nvcc cross-compiles without a GPU)."""
    from smi_experiments_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        from smi_experiments_b200.csrc import build
        build.build()
    return _lib.load()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device (there is no CPU fallback)")
    from smi_experiments_b200 import ops
    ops.device_check()
    return torch.device("cuda:0")
