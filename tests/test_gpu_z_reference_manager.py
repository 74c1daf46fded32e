"""-m gpu: the drop-in as it is deployed -- the REFERENCE's cache manager on top of the GPU store.

In an Alluxio client nothing above the PageStore boundary changes: the Java LocalCacheManager, its meta store and its
evictors keep running and call GpuPageStore.put / get / delete / getPages.  The JVM is not available
here, so the manager is the oracle's restatement of that Java (pinned by the reference's own LocalCacheManagerTest,
tests/test_oracle_cache_manager.py), and the store behind it is the product: `orc.PageStore.over(GpuPageStore)` hands the
five PageStore calls through to the C ABI.  Whatever the reference's manager does over its own store it must do over the
GPU store: same return values, same bytes, same residency, same pages left in the store.
(tests/test_host_logic_cpu.py also runs these cases on CPU over the store double, which proves the harness.)"""
import uuid

import numpy as np
import pytest

from conftest import increasing
from oracle import binding as orc

pytestmark = pytest.mark.gpu

PS = 512


def gpu_store(cache_pages, **kw):
    from alluxio_b200.page_store import GpuPageStore
    return GpuPageStore(page_size=PS, cache_size=cache_pages * PS, slab_slots=kw.pop("slab_slots", 4096), **kw)


def over(store, cache_pages):
    from alluxio_b200.page_store import PageId
    return orc.PageStore.over(store, PS, cache_pages * PS, make_id=PageId)


@pytest.mark.parametrize("evictor", ["LRU", "LFU"])
def test_reference_manager_over_the_gpu_store_equals_it_over_its_own_store(evictor):
    rng = np.random.default_rng({"LRU": 41, "LFU": 42}[evictor])
    cap = 24
    store = gpu_store(cap)
    a = orc.LocalCacheManager(over(store, cap), PS, evictor)                       # reference manager, GPU store
    ref_store = orc.PageStore.open_mem(PS, cap * PS)
    b = orc.LocalCacheManager(ref_store, PS, evictor)                              # reference manager, its own kind of store
    assert a.state() == b.state() == a.READ_WRITE
    lengths = [0, 1, 17, PS // 2, PS - 1, PS, PS, PS, PS]
    for step in range(2500):
        pid = orc.PageId(str(int(rng.integers(5))), int(rng.integers(20)))
        op = int(rng.integers(10))
        if step % 500 == 300:                                                       # FaultyPageStore for a while: on the GPU
            fp, fd = bool(rng.integers(2)), bool(rng.integers(2))                   # store it is the library's own injection
            store.set_faulty(put=fp, delete=fd)
            ref_store.set_faulty(put=fp, delete=fd)
        elif step % 500 == 360:
            store.set_faulty()
            ref_store.set_faulty()
        if op < 4:
            data = rng.integers(0, 256, lengths[int(rng.integers(len(lengths)))], dtype=np.uint8).tobytes()
            assert a.put(pid, data) == b.put(pid, data), (step, "put", pid)
        elif op < 9:
            want = int(rng.choice([0, 1, 17, PS // 2, PS - 1, PS, PS]))
            off = int(rng.choice([0, 0, 0, 1, PS // 2, PS]))
            at = int(rng.integers(3))
            ba, bb = bytearray(at + want), bytearray(at + want)
            out = []
            for mgr, buf in ((a, ba), (b, bb)):
                try:
                    out.append(mgr.get(pid, want, buf, offset_in_buffer=at, page_offset=off))
                except ValueError:                                                  # offset beyond a short page
                    out.append("IllegalArgumentException")
            assert out[0] == out[1] and ba == bb, (step, "get", pid, want, off, out)
        else:
            assert a.delete(pid) == b.delete(pid), (step, "delete", pid)
        assert (a.pages(), a.bytes()) == (b.pages(), b.bytes()), step
    store.set_faulty()
    ref_store.set_faulty()
    left = sorted((p.file_id, p.page_index, n) for p, n in store.get_pages())
    ref_left = sorted((p.file_id, p.page_index, n) for p, n in ref_store.get_pages())
    # The stores hold the same pages.  (More than the manager counts: a failed get only forgets the page in the meta
    # store, CACHE/LocalCacheManager.java:384-395 -- the store keeps it until the id is put again, in the reference's
    # LocalPageStore as a file, here as a slab slot.)
    assert left == ref_left and len(left) >= a.pages() > 0
    st = store.stats()
    assert st["pages"] == len(left) and st["bytes"] == sum(n for _, _, n in left) and st["crc_mismatches"] == 0
    store.close()


def test_reference_manager_restores_from_a_populated_gpu_store():
    """CTEST/LocalCacheManagerTest.java:357-369 (syncRestore) and :494-511 (more pages than capacity) with the reference's
    restore loop (CACHE/LocalCacheManager.java:466-509) walking the GPU store's getPages()."""
    from alluxio_b200.page_store import PageId
    store = gpu_store(64)
    uid = str(uuid.uuid4())
    store.put(PageId("0", 0), increasing(PS))
    store.put(PageId(uid, 3), increasing(7, 100))
    m = orc.LocalCacheManager(over(store, 64), PS, "LRU")
    assert m.state() == m.READ_WRITE and (m.pages(), m.bytes()) == (2, PS + 100)
    buf = bytearray(PS)
    assert m.get(orc.PageId("0", 0), PS, buf) == PS and bytes(buf) == increasing(PS)
    assert m.get(orc.PageId(uid, 3), 100, buf) == 100 and bytes(buf[:100]) == increasing(7, 100)
    store.close()
    store = gpu_store(2)                                                            # room for two pages, three in the store
    for i in range(3):
        store.put(PageId("5", i), increasing(i, PS))
    m = orc.LocalCacheManager(over(store, 2), PS, "LRU")
    assert m.state() == m.READ_WRITE and (m.pages(), m.bytes()) == (2, 2 * PS)
    assert len(store.get_pages()) == 2                                              # the surplus page was deleted from the store
    assert sum(m.get(orc.PageId("5", i), PS, buf) == PS for i in range(3)) == 2
    store.close()


def test_reference_read_flow_over_the_gpu_store():
    """The reference's stream (oracle restatement) over the reference's manager over the GPU store: the whole client
    read path above the boundary is the checker's, only the store is the product."""
    cap = 12
    store = gpu_store(cap)
    m = orc.LocalCacheManager(over(store, cap), PS, "LRU")
    data = np.random.default_rng(9).integers(0, 256, 30 * PS + 77, dtype=np.uint8).tobytes()
    s = orc.LocalCacheFileInStream(m, "77", len(data), lambda pos, n: data[pos:pos + n], PS)
    rng = np.random.default_rng(10)
    for _ in range(200):
        if rng.integers(3) == 0:
            s.seek(int(rng.integers(len(data) + 1)))
        pos, n = s.get_pos(), int(rng.integers(0, 4 * PS))
        b = bytearray(n)
        r = s.read(b)
        assert r == (0 if n == 0 else -1 if pos >= len(data) else min(n, len(data) - pos))
        assert r <= 0 or bytes(b[:r]) == data[pos:pos + r]
        assert m.pages() <= cap
    c = s.counters()
    assert c["BytesReadCache"] > 0 and c["BytesRequestedExternal"] > 0
    assert store.stats()["pages"] == m.pages() and store.stats()["crc_mismatches"] == 0   # no failed gets here: no orphans
    store.close()
