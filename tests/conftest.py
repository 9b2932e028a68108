import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle as _oracle

    _oracle.build()
    return _oracle


@pytest.fixture(scope="session")
def cuda():
    """The loaded C-ABI library; fails loudly (no fallback) if missing."""
    from pyopal_b200.platform import cuda as _cuda

    _cuda.load_library()
    if _cuda.device_count() <= 0:
        pytest.fail("gpu test selected but no CUDA device is visible")
    return _cuda
