import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (sm_100a); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def hpe():
    import hpe_b200
    return hpe_b200


@pytest.fixture(scope="session")
def cuda_dev():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but CUDA is not available (no silent CPU fallback)")
    import hpe_b200
    hpe_b200._lib.check(hpe_b200._lib.lib().hpe_check_device(), "hpe_check_device")
    return torch.device("cuda:0")
