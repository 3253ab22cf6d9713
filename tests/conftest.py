import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.load_c()
    return O


@pytest.fixture(scope="session")
def gpu_available():
    from gp_dla_detection_b200 import _lib
    return _lib.load().dla_device_count() > 0
