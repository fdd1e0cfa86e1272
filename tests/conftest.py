import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    """GPU tests must never pass silently on a box without a GPU: they fail (the driver selects with -m)."""
    return


@pytest.fixture(scope="session")
def cuda():
    import torch
    assert torch.cuda.is_available(), "this test needs a CUDA device"
    return torch.device("cuda:0")
