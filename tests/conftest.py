import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def increasing(start_or_len, length=None):
    """BufferUtils.getIncreasingByteArray (core/common/.../util/io/BufferUtils.java:203-209)."""
    if length is None:
        start, length = 0, start_or_len
    else:
        start = start_or_len
    return bytes((k + start) & 0xFF for k in range(length))


@pytest.fixture
def inc():
    return increasing
