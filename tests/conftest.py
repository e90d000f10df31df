import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    from oracle import awq_oracle

    awq_oracle.build()
    return awq_oracle


@pytest.fixture(scope="session")
def cabi():
    """The C-ABI library through ctypes; GPU tests fail (not skip) if it cannot run."""
    import torch

    from sglang_awq_b200 import _C

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    L = _C.lib()
    assert L.awq_b200_check_device() == 0, L.awq_b200_last_error().decode()
    return _C
