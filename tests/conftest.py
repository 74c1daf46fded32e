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


@pytest.fixture(autouse=True)
def _gpu_stream(request):
    """GPU tests run torch ops and libaps kernels on ONE non-default stream, so they are ordered
    (libaps treats a NULL stream as "the store's own stream", which torch's default stream does
    not synchronise with)."""
    if request.node.get_closest_marker("gpu") is None:
        yield
        return
    import torch
    s = torch.cuda.Stream()
    prev = torch.cuda.current_stream()
    torch.cuda.set_stream(s)
    yield
    torch.cuda.synchronize()
    torch.cuda.set_stream(prev)
