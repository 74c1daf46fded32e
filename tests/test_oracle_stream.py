"""The oracle's LocalCacheFileInStream pinned by the reference's own unit tests.

Port of CTEST/LocalCacheFileInStreamTest.java (CTEST = /root/reference/core/client/fs/src/test/java/alluxio/client/file/
cache/), case by case.  The reference drives the stream over a ByteArrayCacheManager that counts pages served and pages
cached (:395-441); here the stream runs over the oracle's real LocalCacheManager + in-memory PageStore, and the same two
counts are read off it: pages cached = pages in the meta store, pages served = cache gets that returned bytes.
CPU only: the oracle is the checker of the GPU tests (tests/test_gpu_stream_factory.py), this file is what pins it."""
import random

import numpy as np
import pytest

from oracle import binding as orc

PAGE_SIZE = 1 << 20  # alluxio.user.client.cache.page.size default (LocalCacheFileInStreamTest.java:81-82)


def increasing(n, start=0):
    """BufferUtils.getIncreasingByteArray(start, len)."""
    return ((np.arange(n, dtype=np.int64) + start) & 0xFF).astype(np.uint8).tobytes()


class External:
    """ByteArrayFileSystem's stream (:446-540): positional reads of a byte array; counts calls."""

    def __init__(self, data, multi_read=False, seed=0):
        self.data, self.calls, self.multi, self.rng = data, 0, multi_read, random.Random(seed)

    def __call__(self, pos, n):
        self.calls += 1
        if self.multi and n > 1:  # MultiReadFileInStream.read (:641-644): a random part of what was asked for
            n = self.rng.randrange(1, n)
        return self.data[pos:pos + n]


class Cache:
    """One oracle manager with room for `pages` pages, plus the two counts of ByteArrayCacheManager."""

    def __init__(self, pages=64, page_size=PAGE_SIZE):
        self.page_size = page_size
        self.store = orc.PageStore.open_mem(page_size, pages * page_size)
        self.manager = orc.LocalCacheManager(self.store, page_size)
        self.streams = []

    def open(self, file_id, data, **kw):
        ext = External(data, **kw)
        s = orc.LocalCacheFileInStream(self.manager, file_id, len(data), ext, self.page_size)
        s.external = ext
        self.streams.append(s)
        return s

    @property
    def pages_cached(self):
        return self.manager.pages()


def verify_read_full_file(file_size, buffer_size, pages):
    """verifyReadFullFile (:373-391)."""
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    miss = bytearray(buffer_size)
    assert stream.read(miss) == file_size                      # cache miss
    assert bytes(miss[:file_size]) == data
    assert stream.counters()["BytesReadCache"] == 0            # mPagesServed == 0
    assert cache.pages_cached == pages
    stream.seek(0)
    hit = bytearray(buffer_size)
    calls = stream.external.calls
    assert stream.read(hit) == file_size                       # cache hit
    assert bytes(hit[:file_size]) == data
    assert stream.counters()["BytesReadCache"] == file_size    # every page served by the cache ...
    assert stream.external.calls == calls                      # ... and nothing read outside


def test_read_full_page():           # readFullPage :88-94
    verify_read_full_file(PAGE_SIZE, PAGE_SIZE, 1)


def test_read_small_page():          # readSmallPage :96-102
    verify_read_full_file(PAGE_SIZE // 5, PAGE_SIZE // 5, 1)


def test_read_partial_page():        # readPartialPage :104-131
    file_size = PAGE_SIZE
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    part, offset = file_size // 5, file_size // 5
    miss = bytearray(part)
    stream.seek(offset)
    assert stream.read(miss) == part and bytes(miss) == data[offset:offset + part]
    assert stream.counters()["BytesReadCache"] == 0 and cache.pages_cached == 1
    hit = bytearray(part)
    stream.seek(offset)
    assert stream.read(hit) == part and bytes(hit) == data[offset:offset + part]
    assert stream.counters()["BytesReadCache"] == part


def test_read_multi_page():          # readMultiPage :133-139
    verify_read_full_file(PAGE_SIZE + 10, PAGE_SIZE + 10, 2)


def test_read_multi_page_mixed():    # readMultiPageMixed :141-168
    pages = 10
    file_size = PAGE_SIZE * pages
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    rng = random.Random(7)
    cached = 0
    for i in range(pages):
        stream.seek(PAGE_SIZE * i)
        if rng.random() < 0.5:
            assert stream.read_byte() == data[i * PAGE_SIZE]
            cached += 1
    assert stream.counters()["BytesReadCache"] == 0 and cache.pages_cached == cached
    stream.seek(0)
    full = bytearray(file_size)
    assert stream.read(full) == file_size and bytes(full) == data
    assert stream.counters()["BytesReadCache"] == cached * PAGE_SIZE   # exactly the cached pages were served
    assert cache.pages_cached == pages


def test_read_oversized_buffer():                # readOversizedBuffer :170-176
    verify_read_full_file(PAGE_SIZE, PAGE_SIZE * 2, 1)


def test_read_small_page_oversized_buffer():     # readSmallPageOversizedBuffer :178-184
    verify_read_full_file(PAGE_SIZE // 3, PAGE_SIZE // 3 * 2, 1)


def test_positioned_read_partial_page():         # positionedReadPartialPage :186-213
    file_size = PAGE_SIZE
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    part, offset = file_size // 5, file_size // 5
    miss = bytearray(part)
    assert stream.positioned_read(offset, miss, 0, len(miss)) == part and bytes(miss) == data[offset:offset + part]
    assert stream.counters()["BytesReadCache"] == 0 and cache.pages_cached == 1
    hit = bytearray(part)
    assert stream.positioned_read(offset, hit, 0, len(hit)) == part and bytes(hit) == data[offset:offset + part]
    assert stream.counters()["BytesReadCache"] == part
    assert stream.get_pos() == 0                 # a positioned read leaves the stream where it was


def test_position_read_oversized_buffer():       # positionReadOversizedBuffer :215-238
    """len (2 * fileSize) exceeds b.length - off (2 * fileSize - 2): accepted, because what is actually read fits."""
    file_size = PAGE_SIZE
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    miss = bytearray(file_size * 2)
    assert stream.positioned_read(1, miss, 2, file_size * 2) == file_size - 1
    assert bytes(miss[2:file_size]) == data[1:file_size - 1]
    assert bytes(miss[2:file_size + 1]) == data[1:file_size]
    assert stream.counters()["BytesReadCache"] == 0 and cache.pages_cached == 1
    hit = bytearray(file_size * 2)
    assert stream.positioned_read(1, hit, 2, file_size * 2) == file_size - 1
    assert bytes(hit[2:file_size + 1]) == data[1:file_size]
    assert stream.counters()["BytesReadCache"] == file_size - 1


def test_read_pages_metrics():                   # readPagesMetrics :240-267
    file_size = PAGE_SIZE * 5
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data)
    read_size = file_size - 1
    miss = bytearray(read_size)
    stream.read(miss)
    c = stream.counters()
    assert (c["BytesReadCache"], c["BytesRequestedExternal"], c["BytesReadExternal"]) == (0, read_size, file_size)
    assert stream.read_byte() == data[-1]
    c = stream.counters()
    assert (c["BytesReadCache"], c["BytesRequestedExternal"], c["BytesReadExternal"]) == (1, read_size, file_size)


def test_external_store_multi_read():            # externalStoreMultiRead :269-296
    file_size = PAGE_SIZE
    data = increasing(file_size)
    cache = Cache()
    stream = cache.open("/test", data, multi_read=True, seed=3)
    miss = bytearray(file_size)
    assert stream.read(miss) == file_size and bytes(miss) == data
    assert stream.external.calls > 1             # the page came in several pieces
    assert stream.counters()["BytesReadCache"] == 0 and cache.pages_cached == 1
    stream.seek(0)
    hit = bytearray(file_size)
    assert stream.read(hit) == file_size and bytes(hit) == data
    assert stream.counters()["BytesReadCache"] == file_size


def test_read_multiple_files():                  # readMultipleFiles :298-311
    rng = random.Random(11)
    cache = Cache()
    files = {"/file%d" % i: increasing(rng.randrange(100), start=i) for i in range(10)}
    for path, data in files.items():
        stream = cache.open(path, data)
        out = bytearray()
        buf = bytearray(8192)                    # ByteStreams.toByteArray: read until -1
        while True:
            r = stream.read(buf)
            if r == -1:
                break
            out += buf[:r]
        assert bytes(out) == data
        stream.close()


# ---- the rest of the class, which the reference's tests leave to its callers ------------------------------------------

def test_argument_checks_and_end_of_file():
    """read / positionedRead :108-117, :175-184: the order of the checks."""
    data = increasing(1000)
    cache = Cache(page_size=256)
    s = cache.open("7", data)
    b = bytearray(16)
    with pytest.raises(ValueError):
        s.read(b, 0, -1)
    with pytest.raises(ValueError):
        s.read(b, -1, 4)
    with pytest.raises(ValueError):
        s.positioned_read(-1, b, 0, 4)
    assert s.read(b, 0, 0) == 0                  # len == 0 before the end-of-file test
    s.seek(1000)
    assert s.read(b, 0, 0) == 0 and s.read(b) == -1 and s.remaining() == 0
    assert s.positioned_read(1000, b) == -1 and s.positioned_read(5000, b) == -1 and s.positioned_read(999, b) == 1
    assert b[0] == data[999]


def test_seek_and_skip():
    """skip :154-162, seek :230-243."""
    data = increasing(1000)
    cache = Cache(page_size=256)
    s = cache.open("7", data)
    assert s.skip(-5) == 0 and s.skip(0) == 0 and s.get_pos() == 0
    assert s.skip(300) == 300 and s.get_pos() == 300 and s.remaining() == 700
    assert s.skip(5000) == 700 and s.get_pos() == 1000 and s.skip(1) == 0
    s.seek(1000)                                 # pos == length is allowed
    with pytest.raises(ValueError):
        s.seek(1001)
    with pytest.raises(ValueError):
        s.seek(-1)
    s.seek(10)
    b = bytearray(4)
    assert s.read(b) == 4 and bytes(b) == data[10:14] and s.get_pos() == 14
    s.close()
    with pytest.raises(RuntimeError):
        s.seek(0)
    with pytest.raises(RuntimeError):
        s.skip(1)


def test_short_external_read_is_an_io_exception_and_caches_nothing():
    """readExternalPage :297-308: a page that does not arrive whole is an IOException; nothing reaches the cache."""
    data = increasing(1000)
    cache = Cache(page_size=256)
    s = orc.LocalCacheFileInStream(cache.manager, "9", 1000, lambda pos, n: data[pos:min(pos + n, 300)], 256)
    b = bytearray(600)
    with pytest.raises(IOError):
        s.read(b)
    assert bytes(b[:256]) == data[:256] and s.get_pos() == 256   # the first page had been delivered (mPosition moved)
    assert cache.pages_cached == 1
    c = s.counters()
    assert (c["BytesRequestedExternal"], c["BytesReadExternal"]) == (256, 256 + 44)


def test_a_buffer_too_small_for_what_is_read_is_an_illegal_argument():
    """The bound that exists is LocalCacheManager.get's bytesToRead <= buffer.length - offsetInBuffer (:364-366), page
    by page: a read whose first page fits and second does not delivers the first and then throws."""
    data = increasing(1000)
    cache = Cache(page_size=256)
    s = cache.open("7", data)
    b = bytearray(300)
    with pytest.raises(ValueError):
        s.read(b, 0, 600)
    assert bytes(b[:256]) == data[:256] and s.get_pos() == 256


def test_random_walk_returns_the_file():
    """Any in-contract sequence of reads, positioned reads, seeks and skips returns the file's bytes; the cache never
    holds more than its capacity and every byte is accounted for by exactly one meter."""
    rng = random.Random(5)
    ps = 4096
    data = np.random.default_rng(5).integers(0, 256, 37 * ps + 123, dtype=np.uint8).tobytes()
    cache = Cache(pages=8, page_size=ps)
    s = cache.open("12", data)
    delivered = 0
    for _ in range(400):
        op = rng.randrange(4)
        if op == 0:
            s.seek(rng.randrange(len(data) + 1))
        elif op == 1:
            before = s.get_pos()
            assert s.skip(rng.randrange(3 * ps)) == s.get_pos() - before
        elif op == 2:
            n, off = rng.randrange(3 * ps), rng.randrange(5)
            b, pos = bytearray(n + off), s.get_pos()
            r = s.read(b, off, n)
            want = data[pos:pos + n]
            assert r == (0 if n == 0 else -1 if pos >= len(data) else len(want))
            if r > 0:
                assert bytes(b[off:off + r]) == want and s.get_pos() == pos + r
                delivered += r
        else:
            n, pos = rng.randrange(3 * ps), rng.randrange(len(data) + ps)
            b, here = bytearray(n), s.get_pos()
            r = s.positioned_read(pos, b)
            want = data[pos:pos + n]
            assert r == (0 if n == 0 else -1 if pos >= len(data) else len(want))
            assert s.get_pos() == here
            if r > 0:
                assert bytes(b[:r]) == want
                delivered += r
        assert cache.pages_cached <= 8
    c = s.counters()
    assert c["BytesReadCache"] + c["BytesRequestedExternal"] == delivered
    assert c["BytesReadCache"] > 0 and c["BytesRequestedExternal"] > 0
