import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure): builds oracle/liborbref.so on first use."""
    from oracle import orbref
    orbref.lib()
    return orbref


@pytest.fixture(scope="session")
def gpu_lib():
    """liborbgpu through its C ABI.  Fails (not skips) if the library is missing on a GPU box."""
    from ar_orbslam2_b200 import _lib
    return _lib.lib()
