import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("test is marked gpu but no CUDA device is visible")
    return torch.device("cuda", 0)


@pytest.fixture(scope="session", autouse=True)
def _built_oracle():
    # the C restatement is test infrastructure; build it once per session if missing
    from oracle import c_oracle

    c_oracle.build()
    yield
