"""CPU: LocalCacheFileSystem.openFile with no usable cache (CACHE/LocalCacheFileSystem.java:51-70) hands out the
delegated file system's stream; position arithmetic of the stream object (read / skip / seek / remaining / getPos,
CACHE/LocalCacheFileInStream.java:108-243) does not need a store, so it is checked here without a GPU."""
import pytest

from conftest import increasing


def _stream(data, chunk=None):
    from alluxio_b200.cache import LocalCacheFileInStream
    calls = []

    def external(pos, n):
        calls.append((pos, n))
        return data[pos:pos + (min(n, chunk) if chunk else n)]

    return LocalCacheFileInStream(None, "7", len(data), external), calls


def test_bypass_reads_come_from_the_external_stream():
    data = increasing(5000)
    s, calls = _stream(data)
    assert s.bypass() and s.remaining() == 5000 and s.get_pos() == 0
    b = bytearray(1234)
    assert s.read(b) == 1234 and bytes(b) == data[:1234] and s.get_pos() == 1234
    assert s.read(b, 10, 100) == 100 and bytes(b[10:110]) == data[1234:1334]
    assert s.skip(1000) == 1000 and s.get_pos() == 2334 and s.skip(0) == 0 and s.skip(-3) == 0
    big = bytearray(10000)
    assert s.read(big) == 5000 - 2334 and bytes(big[:5000 - 2334]) == data[2334:]   # short read at the end of the file
    assert s.read(big) == -1 and s.remaining() == 0
    assert s.read(big, 0, 0) == 0
    s.seek(4999)
    assert s.read(big) == 1 and big[0] == data[4999]
    one = bytearray(3)
    assert s.positioned_read(100, one) == 3 and bytes(one) == data[100:103] and s.get_pos() == 5000
    assert s.positioned_read(5000, one) == -1
    assert s.counters() == {"BytesReadCache": 0, "BytesRequestedExternal": 0, "BytesReadExternal": 0}
    assert all(n > 0 for _, n in calls)


def test_bypass_argument_checks_and_closed_stream():
    data = increasing(300)
    s, _ = _stream(data)
    b = bytearray(100)
    for bad in ((b, -1, 1), (b, 0, -1), (b, 0, 101), (b, 101, 0)[:3]):
        if bad[2] == 0:
            continue
        with pytest.raises(ValueError):
            s.read(*bad)
    with pytest.raises(ValueError):
        s.seek(-1)
    with pytest.raises(ValueError):
        s.seek(301)
    s.seek(300)                                             # the end itself is a legal position
    with pytest.raises(ValueError):
        s.positioned_read(-1, b)
    s.close()
    with pytest.raises(RuntimeError):
        s.seek(0)
    with pytest.raises(RuntimeError):
        s.skip(1)


def test_bypass_short_external_reads_are_completed_or_fail():
    data = increasing(2000)
    s, calls = _stream(data, chunk=300)                     # the source hands out at most 300 bytes per call
    b = bytearray(1000)
    assert s.read(b) == 1000 and bytes(b) == data[:1000] and len(calls) == 4
    from alluxio_b200.cache import LocalCacheFileInStream
    dead = LocalCacheFileInStream(None, "8", 100, lambda pos, n: b"")
    with pytest.raises(IOError):
        dead.read(bytearray(10))
