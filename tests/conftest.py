import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def built_lib():
    """Path of the in-tree C-ABI library; builds it when missing (nvcc cross-compiles without a GPU)."""
    path = os.path.join(ROOT, "gpu-fhe_b200", "libgpufhe_b200.so")
    if not os.path.exists(path):
        import __graft_entry__ as g
        g.build()
    assert os.path.exists(path)
    return path


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    return torch.device("cuda", 0)
