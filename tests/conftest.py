import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def hal():
    from iree_b200 import hal as h
    h.lib()
    return h


@pytest.fixture(scope="session")
def device(hal):
    return hal.Device(0)


@pytest.fixture(scope="session")
def ops(hal, device):
    return hal.Ops(device)
