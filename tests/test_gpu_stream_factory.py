"""-m gpu: the callers of the cache manager, mirrored natively:
LocalCacheFileInStream.read / seek / skip (CACHE/LocalCacheFileInStream.java:108-243; cases after
CTEST/LocalCacheFileInStreamTest.java), the bypass of LocalCacheFileSystem.openFile (:51-70), CacheManager.Factory
(CACHE/CacheManager.java:69-127), asynchronous write (CACHE/LocalCacheManager.java:214-245) and asynchronous restore
(:126-133)."""
import threading

import numpy as np
import pytest

from conftest import increasing

pytestmark = pytest.mark.gpu

PS = 1024


def PID(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def _manager(**kw):
    from alluxio_b200.cache import LocalCacheManager
    kw.setdefault("evictor", "LRU")
    return LocalCacheManager.create(PS, kw.pop("cache_size", 64 * PS), **kw)


class External:
    """The delegated file system's stream, counting what is asked of it."""

    def __init__(self, data, short=False):
        self.data, self.calls, self.bytes, self.short = data, 0, 0, short

    def __call__(self, pos, n):
        self.calls += 1
        out = self.data[pos:pos + (max(1, n // 2) if self.short else n)]
        self.bytes += len(out)
        return out


def test_sequential_read_seek_skip_and_eof():
    file_len = 5 * PS + 100
    data = np.random.default_rng(1).integers(0, 256, file_len, dtype=np.uint8).tobytes()
    m = _manager()
    ext = External(data)
    s = m.open_file("42", file_len, ext)
    assert not s.bypass() and s.remaining() == file_len and s.get_pos() == 0
    b = bytearray(700)
    got = bytearray()
    while True:                                            # readFullPage / readSmallPage / readPartialPage
        r = s.read(b)
        if r == -1:
            break
        assert 0 < r <= 700
        got += b[:r]
    assert bytes(got) == data and s.remaining() == 0 and s.get_pos() == file_len
    assert ext.bytes == file_len                           # every page fetched exactly once, whole
    c = s.counters()
    assert c["BytesReadCache"] + c["BytesRequestedExternal"] == file_len and c["BytesReadExternal"] == file_len
    assert c["BytesReadCache"] > 0
    assert m.pages() == 6 and m.bytes() == file_len        # the last page is short (:294)
    s.seek(PS + 10)                                        # second pass: all from the cache
    calls = ext.calls
    b2 = bytearray(3 * PS)
    assert s.read(b2, 5, 2 * PS) == 2 * PS and bytes(b2[5:5 + 2 * PS]) == data[PS + 10:3 * PS + 10] and ext.calls == calls
    assert s.skip(100) == 100 and s.get_pos() == 3 * PS + 110
    assert s.skip(10 ** 9) == file_len - (3 * PS + 110) and s.remaining() == 0
    assert s.read(b2) == -1
    assert s.skip(-5) == 0 and s.read(b2, 0, 0) == 0
    with pytest.raises(ValueError):
        s.seek(file_len + 1)
    with pytest.raises(ValueError):
        s.seek(-1)
    with pytest.raises(ValueError):
        s.read(b2, -1, 1)
    s.seek(0)
    with pytest.raises(ValueError):                        # (at end of file the answer would be -1 first, as in :115-117)
        s.read(b2, 0, len(b2) + 1)
    s.seek(file_len)                                       # seeking to the end is legal
    one = bytearray(1)
    assert s.positioned_read(file_len - 1, one) == 1 and one[0] == data[-1]   # positioned reads leave the position alone
    assert s.get_pos() == file_len and s.positioned_read(file_len, one) == -1
    s.close()
    with pytest.raises(RuntimeError):
        s.seek(0)
    with pytest.raises(RuntimeError):
        s.skip(1)
    m.close()


def test_short_external_read_is_an_io_error_and_caches_nothing():
    data = bytes(range(256)) * 8
    m = _manager()
    s = m.open_file("7", len(data), External(data, short=True))
    ok = m.open_file("7", len(data), External(data))
    b = bytearray(100)
    # a stream whose source returns short but non-empty reads still completes the page (the loop of :297-303)
    assert s.read(b) == 100 and bytes(b) == data[:100]
    empty = m.open_file("8", len(data), lambda pos, n: b"")
    with pytest.raises(IOError):
        empty.read(b)
    assert m.pages() == 1
    assert ok.read(b) == 100
    m.close()


def test_bypass_when_the_cache_is_not_in_use():
    from alluxio_b200.cache import LocalCacheFileInStream
    data = increasing(3000)
    ext = External(data)
    s = LocalCacheFileInStream(None, "1", len(data), ext)  # mCacheManager == null
    assert s.bypass()
    b = bytearray(3000)
    assert s.read(b, 0, 1234) == 1234 and s.read(b, 1234, 3000 - 1234) == 3000 - 1234 and bytes(b) == data
    assert s.read(b) == -1 and ext.bytes == 3000 and s.counters()["BytesReadCache"] == 0


def test_async_write_queues_rejects_and_lands():
    m = _manager(cache_size=4096 * PS, slab_slots=4096, async_write_threads=1)
    accepted = []
    for k in range(300):                                   # one worker: most puts find it busy and are refused at once
        if m.put(PID("9", k), increasing(k % 256, PS)):
            accepted.append(k)
    m.wait_async()
    assert accepted and accepted[0] == 0
    assert m.put_async_rejections() == 300 - len(accepted)
    assert m.pages() == len(accepted)
    buf = bytearray(PS)
    for k in accepted[:50]:
        assert m.get(PID("9", k), PS, buf) == PS and bytes(buf) == increasing(k % 256, PS)
    assert m.metrics()["PutErrors"] == 300 - len(accepted)
    m.close()
    m = _manager(cache_size=4096 * PS, slab_slots=4096, async_write_threads=16)
    n_ok = sum(bool(m.put(PID("5", k), increasing(PS))) for k in range(64))
    m.wait_async()
    assert m.pages() == n_ok >= 16                          # sixteen workers take at least the first sixteen
    m.close()


def test_async_restore_serves_gets_refuses_puts_until_ready():
    from alluxio_b200.cache import LocalCacheManager
    from alluxio_b200.page_store import GpuPageStore
    store = GpuPageStore(page_size=PS, cache_size=4096 * PS, slab_slots=4096)
    for k in range(2000):
        store.put(PID("3", k), increasing(k % 256, PS))
    m = LocalCacheManager.create(PS, 4096 * PS, over=store, async_restore=True)
    states = {m.state()}
    early_put = m.put(PID("4", 0), increasing(PS))          # very likely still READ_ONLY: then refused and counted
    not_ready = m.put_not_ready_errors()
    m.wait_async()
    states.add(m.state())
    assert m.state() == m.READ_WRITE and states <= {m.READ_ONLY, m.READ_WRITE}
    assert (not early_put) == (not_ready == 1)
    assert m.pages() == 2000 + (1 if early_put else 0)
    buf = bytearray(PS)
    assert m.get(PID("3", 1999), PS, buf) == PS and bytes(buf) == increasing(1999 % 256, PS)
    assert m.put(PID("4", 1), increasing(PS))
    m.close()
    store.close()


def test_factory_is_a_singleton_and_null_while_creating():
    from alluxio_b200.cache import CacheManagerFactory
    CacheManagerFactory.clear()
    got = []
    barrier = threading.Barrier(6)

    def worker():
        barrier.wait()
        got.append(CacheManagerFactory.get(PS, 256 * PS))

    ts = [threading.Thread(target=worker) for _ in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    handles = {m._h.value for m in got if m is not None}
    assert len(handles) == 1                                # one manager; late arrivals during creation saw None
    again = CacheManagerFactory.get(PS, 999 * PS)           # "conf only respected for the first call"
    assert again._h.value in handles
    assert again.put(PID("1", 1), increasing(PS)) and again.pages() == 1
    errs = CacheManagerFactory.create_errors()
    CacheManagerFactory.clear()
    with pytest.raises(Exception):                          # a failed creation is reported and counted
        CacheManagerFactory.get(0, 256 * PS)                # a page size of zero is not a configuration
    assert CacheManagerFactory.create_errors() == errs + 1
    fresh = CacheManagerFactory.get(PS, 256 * PS)
    assert fresh.pages() == 0
    CacheManagerFactory.clear()


def test_time_bound_page_store():
    """CTEST/TimeBoundPageStoreTest.java: put / get / delete pass through; a call that outlives the bound fails like
    an IOException and is counted; more concurrent calls than store threads are refused."""
    m = _manager(timeout_ms=150, timeout_threads=2)
    p = PID("0L", 0)
    page, buf = increasing(PS), bytearray(PS)
    assert m.put(p, page)                                   # put / get / delete
    assert m.get(p, PS, buf) == PS and bytes(buf) == page
    assert m.delete(p) and m.get(p, PS, buf) == 0
    assert m.store_timeouts() == {"put": 0, "get": 0, "delete": 0, "rejected": 0}
    m.set_store_delays(put_ms=600)                          # putTimeout: the manager sees a failed store write
    assert not m.put(p, page)
    assert m.store_timeouts()["put"] == 1 and m.metrics()["PutStoreWriteErrors"] == 1 and m.pages() == 0
    m.set_store_delays()
    import time
    time.sleep(0.7)                                         # let the abandoned put finish inside the store
    q = PID("0L", 1)
    assert m.put(q, page)
    m.set_store_delays(get_ms=600)                          # getTimeout: -1, and the page is dropped from the meta store
    assert m.get(q, PS, buf) == -1
    assert m.store_timeouts()["get"] == 1 and m.pages() == 0
    m.set_store_delays()
    time.sleep(0.7)
    r = PID("0L", 2)
    assert m.put(r, page)
    m.set_store_delays(delete_ms=600)                       # deleteTimeout
    assert not m.delete(r)
    assert m.store_timeouts()["delete"] == 1
    m.set_store_delays()
    time.sleep(0.7)
    # concurrentPutWithSmallThreadPool: two store threads, six callers at once, each call slower than the bound
    m.set_store_delays(put_ms=400)
    results = []
    barrier = threading.Barrier(6)

    def worker(k):
        barrier.wait()
        results.append(m.put(PID("7", k), page))

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    t = m.store_timeouts()
    assert not any(results) and t["rejected"] == 4 and t["put"] == 1 + 2
    m.set_store_delays()
    time.sleep(0.6)
    m.close()
    plain = _manager()                                      # no timeout configured: no wrapper, as with the default -1
    with pytest.raises(Exception):
        plain.store_timeouts()
    plain.close()


def test_async_cache_same_page_is_refused_while_pending():
    """LocalCacheManagerTest.asyncCacheSamePage (:563-580): a put of an id that is still queued returns false and is
    not an async rejection; the first put lands."""
    m = _manager(cache_size=256 * PS, slab_slots=256, async_write_threads=2, timeout_ms=5000)
    m.set_store_delays(put_ms=300)                          # HangingPageStore, shortened: the worker sits in the store
    p = PID("0L", 0)
    assert m.put(p, increasing(PS))
    assert not m.put(p, increasing(255, PS))                # same id, still pending
    assert m.put(PID("0L", 1), increasing(PS))              # another id takes the second worker
    assert not m.put(PID("0L", 2), increasing(PS))          # both workers busy: rejected
    assert m.put_async_rejections() == 1
    m.wait_async()
    m.set_store_delays()
    buf = bytearray(PS)
    assert m.get(p, PS, buf) == PS and bytes(buf) == increasing(PS)          # the first put's bytes
    assert m.pages() == 2
    m.close()


def test_oversized_buffers_and_several_files():
    """LocalCacheFileInStreamTest.readOversizedBuffer / readSmallPageOversizedBuffer / positionReadOversizedBuffer /
    readMultipleFiles: a buffer larger than the file returns the file; files do not see each other's pages."""
    m = _manager()
    files = {str(100 + k): np.random.default_rng(k).integers(0, 256, n, dtype=np.uint8).tobytes()
             for k, n in enumerate((PS, PS // 3, 2 * PS + 17, 5))}
    streams = {f: m.open_file(f, len(d), External(d)) for f, d in files.items()}
    for f, d in files.items():
        b = bytearray(len(d) * 2 + 7)
        assert streams[f].read(b) == len(d) and bytes(b[:len(d)]) == d and streams[f].read(b) == -1
        assert streams[f].positioned_read(0, b) == len(d) and bytes(b[:len(d)]) == d
        assert streams[f].positioned_read(len(d) - 1, b, 3) == 1 and b[3] == d[-1]
    for f, d in files.items():                              # second pass: from the cache, still each file's own bytes
        s = m.open_file(f, len(d), lambda pos, n: (_ for _ in ()).throw(AssertionError("external read on a cached file")))
        b = bytearray(len(d))
        assert s.read(b) == len(d) and bytes(b) == d
    assert m.pages() == sum((len(d) + PS - 1) // PS for d in files.values())
    m.close()


def test_batched_stream_reads_equal_per_page_reads():
    """SURVEY.md 8f-1 on the stream object: with set_batched() every read is one batched store call; bytes, position,
    counters and cache contents must be those of the per-page loop."""
    file_len = 9 * PS + 321
    data = np.random.default_rng(21).integers(0, 256, file_len, dtype=np.uint8).tobytes()
    ma, mb = _manager(), _manager()
    ea, eb = External(data), External(data)
    sa, sb = ma.open_file("55", file_len, ea), mb.open_file("55", file_len, eb)
    sb.set_batched()
    rng = np.random.default_rng(22)
    for _ in range(60):
        if rng.integers(4) == 0:
            pos = int(rng.integers(0, file_len + 1))
            sa.seek(pos); sb.seek(pos)
        n = int(rng.integers(0, 4 * PS))
        ba, bb = bytearray(n + 3), bytearray(n + 3)
        ra, rb = sa.read(ba, 3, n), sb.read(bb, 3, n)
        assert ra == rb and ba == bb and sa.get_pos() == sb.get_pos()
        if ra > 0:
            assert bytes(ba[3:3 + ra]) == data[sa.get_pos() - ra:sa.get_pos()]
    assert sa.counters() == sb.counters()
    assert (ma.pages(), ma.bytes()) == (mb.pages(), mb.bytes())
    assert ea.bytes == eb.bytes
    one = bytearray(2 * PS)
    assert sb.positioned_read(PS - 5, one) == 2 * PS and bytes(one) == data[PS - 5:3 * PS - 5]
    ma.close(); mb.close()
