"""pytest configuration: the `gpu` marker and shared fixtures.

`-m "not gpu"` (this container, no GPU): oracle vs the reference's golden vectors, host logic,
C-ABI symbol export.  `-m gpu` (B200 box): parity of the CUDA path through the C ABI.
"""
import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run on the GPU box with -m gpu)")


@pytest.fixture(scope="session")
def q3t():
    """The product package (ctypes view of libq3t_b200.so); builds the library if it is missing."""
    from util import pkg
    m = pkg()
    if not os.path.exists(m.LIB_PATH):
        m.build()
    return m


@pytest.fixture(scope="session")
def tiny_models():
    from util import models
    return models("tiny", "f16")


@pytest.fixture(scope="session")
def full_models():
    from util import models
    return models("full", "f16")
