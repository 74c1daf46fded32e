"""-m gpu: the host layer over the GPU store against the oracle, step by step -- the manager under random operations
(with stretches of a faulty store), the stream op by op against the oracle's restatement of LocalCacheFileInStream, the
reference's positionReadOversizedBuffer case literally, and TimeBoundPageStoreTest.concurrentPutWithLargeThreadPool.
These cases were written in round 2 after the last GPU session; they are validated on CPU over the store double
(tests/test_host_logic_cpu.py runs these same bodies) and are collected late so that they cannot mask the older suite."""
import threading

import numpy as np
import pytest

from conftest import increasing
from oracle import binding as orc
from test_gpu_cache_manager import PID, manager
from test_gpu_stream_factory import PS, External, _manager

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("evictor", ["LRU", "LFU", "FIFO"])
def test_random_operations_equal_the_oracle_manager(evictor):
    """put / get / delete in random order on a cache under pressure -- pages of every length from empty to full, partial
    and over-long gets (a short read drops the page), gets of zero bytes, offsets up to and beyond the page size, buffers
    too small, re-puts of resident pages, deletes of absent ones, and a store that fails its puts or its deletes for a
    while (FaultyPageStore, CTEST/LocalCacheManagerTest.java:660-691).  Every return value or exception, every byte
    returned and the cache's page and byte counts must equal the oracle manager's after every step."""
    rng = np.random.default_rng({"LRU": 1, "LFU": 2, "FIFO": 3}[evictor])
    ps, cap_pages = 512, 24
    ostore = orc.PageStore.open_mem(ps, cap_pages * ps)
    om = orc.LocalCacheManager(ostore, ps, evictor)
    m = manager(cache_size=cap_pages * ps, evictor=evictor, page_size=ps, slab_slots=4096)
    lengths = [0, 1, 17, ps // 2, ps - 1, ps, ps, ps, ps, ps]
    for step in range(3000):
        f, i = str(int(rng.integers(5))), int(rng.integers(20))
        op = int(rng.integers(10))
        if step % 500 == 250:                                   # a stretch with a faulty store, then back to normal
            fp, fd = bool(rng.integers(2)), bool(rng.integers(2))
            m.set_store_faulty(put=fp, delete=fd)
            ostore.set_faulty(put=fp, delete=fd)
        elif step % 500 == 330:
            m.set_store_faulty()
            ostore.set_faulty()
        if op < 4:
            n = lengths[int(rng.integers(len(lengths)))]
            data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
            r, ro = m.put(PID(f, i), data), om.put(orc.PageId(f, i), data)
            assert r == ro, (step, "put", f, i, n, r, ro)
        elif op < 9:
            want = int(rng.choice([0, 1, 17, ps // 2, ps - 1, ps, ps]))
            off = int(rng.choice([0, 0, 0, 1, ps // 2, ps, ps + 1]))
            room = want if rng.integers(5) else max(0, want - 1)
            at = int(rng.integers(3))
            b, bo = bytearray(at + room), bytearray(at + room)
            out = []
            for mgr, pid, buf in ((m, PID(f, i), b), (om, orc.PageId(f, i), bo)):
                try:
                    out.append(mgr.get(pid, want, buf, offset_in_buffer=at, page_offset=off))
                except ValueError:
                    out.append("IllegalArgumentException")
            assert out[0] == out[1], (step, "get", f, i, want, off, room, at, out)
            assert b == bo, (step, "get bytes", f, i, want, off)
        else:
            r, ro = m.delete(PID(f, i)), om.delete(orc.PageId(f, i))
            assert r == ro, (step, "delete", f, i, r, ro)
        assert (m.pages(), m.bytes()) == (om.pages(), om.bytes()), (step, op)
        if step % 50 == 49:                                     # the Client.Cache* counters both sides keep (LocalCacheManager.Metrics)
            mt, omt = m.metrics(), om.metrics()
            common = [k for k in omt if k in mt]
            assert len(common) == len(omt) == 18 and {k: mt[k] for k in common} == omt, step
    mt = m.metrics()
    assert mt["PagesEvicted"] > 100 and mt["PutStoreWriteErrors"] + mt["PutStoreDeleteErrors"] > 0 and mt["GetErrors"] > 0
    m.set_store_faulty()
    ostore.set_faulty()
    buf, obuf = bytearray(ps), bytearray(ps)
    for f in range(5):                                          # the same pages are left, with the same bytes
        for i in range(20):
            for n in (ps, ps - 1, ps // 2, 17, 1):              # find the stored length: the first n that reads whole
                r, ro = m.get(PID(str(f), i), n, buf), om.get(orc.PageId(str(f), i), n, obuf)
                assert r == ro and (r <= 0 or buf[:r] == obuf[:r])
                if r != -1:
                    break
    m.close()


def test_position_read_oversized_buffer():
    """CTEST/LocalCacheFileInStreamTest.java:215-238, literally: positionedRead(1, b, 2, 2 * fileSize) with b.length =
    2 * fileSize asks for more than b has room for behind off -- and is served, because what the file still holds fits
    (the only bound is LocalCacheManager.get's, page by page)."""
    file_size = PS
    data = increasing(file_size)
    m = _manager()
    s = m.open_file("/test", file_size, External(data))
    miss = bytearray(file_size * 2)
    assert s.positioned_read(1, miss, 2, file_size * 2) == file_size - 1
    assert bytes(miss[2:file_size + 1]) == data[1:file_size]
    assert s.counters()["BytesReadCache"] == 0 and m.pages() == 1
    hit = bytearray(file_size * 2)
    assert s.positioned_read(1, hit, 2, file_size * 2) == file_size - 1
    assert bytes(hit[2:file_size + 1]) == data[1:file_size]
    assert s.counters()["BytesReadCache"] == file_size - 1
    s.set_batched()                                            # the one-call form: same answer
    again = bytearray(file_size * 2)
    assert s.positioned_read(1, again, 2, file_size * 2) == file_size - 1 and again == hit
    m.close()


def test_stream_equals_the_oracle_stream_op_by_op():
    """Three files read through ONE small cache by the product's stream and by the oracle's restatement of
    LocalCacheFileInStream (pinned by the reference's own tests, tests/test_oracle_stream.py): the same random sequence
    of read / positionedRead / seek / skip -- buffers larger and smaller than the read included -- must give the same
    return values and exceptions, bytes, positions, meters and cache contents after every step."""
    from oracle import binding as orc
    cache_pages = 12
    m = _manager(cache_size=cache_pages * PS)
    om = orc.LocalCacheManager(orc.PageStore.open_mem(PS, cache_pages * PS), PS, "LRU")
    rng = np.random.default_rng(77)
    files = {str(300 + k): rng.integers(0, 256, n, dtype=np.uint8).tobytes() for k, n in enumerate((7 * PS + 13, 3 * PS, 9 * PS - 1))}
    mine = {f: m.open_file(f, len(d), External(d)) for f, d in files.items()}
    theirs = {f: orc.LocalCacheFileInStream(om, f, len(d), External(d), PS) for f, d in files.items()}

    def both(f, call):
        out = []
        for s in (mine[f], theirs[f]):
            try:
                out.append(call(s))
            except (ValueError, IOError, RuntimeError) as e:
                out.append(type(e).__name__)
        assert out[0] == out[1], (f, out)
        return out[0]

    names = list(files)
    for step in range(600):
        f = names[int(rng.integers(len(names)))]
        n_file = len(files[f])
        op = int(rng.integers(5))
        if op == 0:
            both(f, lambda s, p=int(rng.integers(-2, n_file + 3)): s.seek(p))
        elif op == 1:
            both(f, lambda s, n=int(rng.integers(-3, 3 * PS)): s.skip(n))
        elif op == 2 or op == 3:                                # read; one in four with a buffer that may be too small
            n, off = int(rng.integers(0, 4 * PS)), int(rng.integers(0, 4))
            room = n if rng.integers(4) else int(rng.integers(0, n + 1))
            a, b = bytearray(off + room), bytearray(off + room)
            bufs = iter((a, b))
            both(f, lambda s: s.read(next(bufs), off, n))
            assert a == b, (step, f, n, off, room)
        else:
            n, pos = int(rng.integers(0, 4 * PS)), int(rng.integers(-1, n_file + PS))
            a, b = bytearray(n + 5), bytearray(n + 5)
            bufs = iter((a, b))
            both(f, lambda s: s.positioned_read(pos, next(bufs), 5, n))
            assert a == b, (step, f, n, pos)
        assert mine[f].get_pos() == theirs[f].get_pos() and mine[f].remaining() == theirs[f].remaining()
        c, oc = mine[f].counters(), theirs[f].counters()
        assert all(c[k] == oc[k] for k in c), (step, c, oc)
        assert (m.pages(), m.bytes()) == (om.pages(), om.bytes()), step
    assert m.pages() == cache_pages or m.bytes() > (cache_pages - 1) * PS      # the cache was full: evictions took part
    buf, obuf = bytearray(PS), bytearray(PS)
    for f, d in files.items():                                  # the same pages are resident, with the same bytes
        for i in range((len(d) + PS - 1) // PS):
            want = min(PS, len(d) - i * PS)
            r, ro = m.get(PID(f, i), want, buf), om.get(orc.PageId(f, i), want, obuf)
            assert r == ro and (r <= 0 or buf[:r] == obuf[:r]), (f, i, r, ro)
    m.close()


def test_time_bound_store_concurrent_puts_with_a_large_pool():
    """TimeBoundPageStoreTest.concurrentPutWithLargeThreadPool (:135-144): as many store threads as callers and a store
    that does not hang -- every put goes through, nothing times out, nothing is rejected."""
    n = 10
    m = _manager(cache_size=256 * PS, slab_slots=256, timeout_ms=5000, timeout_threads=n)
    page = increasing(PS)
    results = [None] * n
    barrier = threading.Barrier(n)

    def worker(k):
        barrier.wait()
        results[k] = m.put(PID("0L", k), page)

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(n)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(results) and m.pages() == n
    assert m.store_timeouts() == {"put": 0, "get": 0, "delete": 0, "rejected": 0}
    buf = bytearray(PS)
    for k in range(n):
        assert m.get(PID("0L", k), PS, buf) == PS and bytes(buf) == page
    m.close()


def test_async_restore_with_more_pages_than_capacity():
    """LocalCacheManagerTest.asyncRestoreWithMorePagesThanCapacity (:513-529): the restore thread adopts what fits and
    deletes the rest from the store; the manager is READ_WRITE afterwards."""
    from alluxio_b200.cache import LocalCacheManager
    from alluxio_b200.page_store import GpuPageStore
    store = GpuPageStore(page_size=PS, cache_size=2 * PS, slab_slots=16)
    for i in range(3):
        store.put(PID("5", i), increasing(i, PS))
    m = LocalCacheManager.create(PS, 2 * PS, over=store, async_restore=True)
    m.wait_async()
    assert m.state() == m.READ_WRITE and (m.pages(), m.bytes()) == (2, 2 * PS)
    assert len(store.get_pages()) == 2
    buf = bytearray(PS)
    hits = [i for i in range(3) if m.get(PID("5", i), PS, buf) == PS and bytes(buf) == increasing(i, PS)]
    assert len(hits) == 2
    assert m.put(PID("6", 0), increasing(PS))                  # room is made the usual way: one victim
    assert (m.pages(), m.bytes()) == (2, 2 * PS)
    m.close()
    store.close()


def test_positioned_reads_through_the_manager_equal_the_oracle_stream():
    """apc_positioned_read -- LocalCacheFileInStream.positionedRead as a call on the manager (CACHE/LocalCacheFileInStream
    .java:175-223) -- against the oracle's stream over the oracle's manager, two files sharing a cache under pressure:
    bytes, end-of-file answers, how many bytes came from the cache and from outside, and the cache's size after every call."""
    cap = 10
    m = _manager(cache_size=cap * PS)
    om = orc.LocalCacheManager(orc.PageStore.open_mem(PS, cap * PS), PS, "LRU")
    rng = np.random.default_rng(31)
    files = {"811": rng.integers(0, 256, 13 * PS + 5, dtype=np.uint8).tobytes(), "812": rng.integers(0, 256, 4 * PS, dtype=np.uint8).tobytes()}
    ostreams = {f: orc.LocalCacheFileInStream(om, f, len(d), (lambda d: lambda pos, n: d[pos:pos + n])(d), PS) for f, d in files.items()}
    for step in range(300):
        f = "811" if rng.integers(3) else "812"
        d = files[f]
        pos, n = int(rng.integers(0, len(d) + PS)), int(rng.integers(0, 4 * PS))
        before = ostreams[f].counters()
        got, from_cache, from_external = m.positioned_read(f, len(d), pos, n, lambda p, k: d[p:p + k])
        b = bytearray(n)
        r = ostreams[f].positioned_read(pos, b)
        after = ostreams[f].counters()
        if r < 0:
            assert got is None, (step, f, pos, n)
        else:
            assert got == bytes(b[:r]) == d[pos:pos + n], (step, f, pos, n)
            assert from_cache == after["BytesReadCache"] - before["BytesReadCache"], (step, f, pos, n)
            assert from_external == after["BytesRequestedExternal"] - before["BytesRequestedExternal"], (step, f, pos, n)
        assert (m.pages(), m.bytes()) == (om.pages(), om.bytes()), step
    assert m.pages() == cap
    m.close()


def test_a_failed_restore_wipes_the_store_and_the_cache_goes_on_empty():
    """restoreOrInit (CACHE/LocalCacheManager.java:441-464): when restore fails, the meta store is reset, the page store
    is created anew (emptied) and the cache is READ_WRITE and empty -- LocalCacheManagerTest.syncRestoreUnknownFile
    (:402-420) ends that way.  A GPU store holds no unknown files, so the failure is injected where the reference's
    FaultyPageStore injects it: the delete of the page that does not fit (:493-497)."""
    from alluxio_b200.cache import LocalCacheManager
    from alluxio_b200.page_store import GpuPageStore
    store = GpuPageStore(page_size=PS, cache_size=2 * PS, slab_slots=16)
    ref_store = orc.PageStore.open_mem(PS, 2 * PS)
    for i in range(3):                                         # three pages, room for two
        store.put(PID("5", i), increasing(i, PS))
        ref_store.put(orc.PageId("5", i), increasing(i, PS))
    store.set_faulty(delete=True)
    ref_store.set_faulty(delete=True)
    m = LocalCacheManager.over_store(store, PS, "LRU")
    om = orc.LocalCacheManager(ref_store, PS, "LRU")
    assert m.state() == om.state() == m.READ_WRITE
    assert (m.pages(), m.bytes()) == (om.pages(), om.bytes()) == (0, 0)
    assert store.get_pages() == [] and ref_store.get_pages() == []
    store.set_faulty()
    ref_store.set_faulty()
    buf = bytearray(PS)
    assert m.get(PID("5", 0), PS, buf) == om.get(orc.PageId("5", 0), PS, bytearray(PS)) == 0
    assert m.put(PID("5", 0), increasing(9, PS)) and m.get(PID("5", 0), PS, buf) == PS and bytes(buf) == increasing(9, PS)
    m.close()
    store.close()


def test_load_cache_and_evict():
    """LocalCacheManagerIntegrationTest.loadCacheAndEvict (:117-144): a cache is filled, its manager closed and a new
    one created over the same store; the restored pages then take part in eviction like any others -- putting half a
    cache of new pages evicts exactly half of the restored ones."""
    from alluxio_b200.cache import LocalCacheManager
    from alluxio_b200.page_store import GpuPageStore
    count = 32
    page = increasing(PS)
    store = GpuPageStore(page_size=PS, cache_size=count * PS, slab_slots=128)
    m = LocalCacheManager.over_store(store, PS, "LRU")
    buf = bytearray(PS)
    for i in range(count):                                     # loadFullCache
        assert m.put(PID("0", i), page)
    for i in range(count):
        assert m.get(PID("0", i), PS, buf) == PS and bytes(buf) == page
    m.close()                                                  # the store outlives its manager, as the cache directory does
    m = LocalCacheManager.over_store(store, PS, "LRU")
    assert m.state() == m.READ_WRITE and m.pages() == count
    for i in range(count // 2):
        assert m.put(PID("1", i), page)
    evicted = 0
    for i in range(count):
        ret = m.get(PID("0", i), PS, buf)
        if ret <= 0:
            evicted += 1
            continue
        assert ret == PS and bytes(buf) == page
    assert evicted == count // 2
    for i in range(count // 2):
        assert m.get(PID("1", i), PS, buf) == PS and bytes(buf) == page
    assert len(store.get_pages()) == count
    m.close()
    store.close()
