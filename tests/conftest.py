import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def vgpu():
    import vgpu_b200  # noqa: F401  registers the package
    import sys as _sys
    return _sys.modules["vgpu_b200"]


@pytest.fixture(scope="session")
def cuda_dev():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test running without CUDA: there is no CPU fallback")
    return torch.device("cuda", 0)
