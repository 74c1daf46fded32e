"""-m gpu: warm restart.  A LocalPageStore survives a restart because its pages are files and
LocalCacheManager.restore re-reads them through getPages() (CACHE/LocalCacheManager.java:466-509); the GPU store
gets the same behaviour from aps_snapshot_save / aps_snapshot_load.  The restore cases follow
CTEST/LocalCacheManagerTest.java:357-369,420-432 (sync restore, restore with more pages than fit)."""
import os

import numpy as np
import pytest

from oracle import binding as orc

pytestmark = pytest.mark.gpu

PAGE = 4096


def _store(cache_pages=512, crc=True, page_size=PAGE, **kw):
    from alluxio_b200.page_store import GpuPageStore
    return GpuPageStore(page_size=page_size, cache_size=cache_pages * page_size, crc=crc, **kw)


def _pid(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def _fill(store, n, seed=11):
    """n pages of assorted lengths (0, 1, odd, full) under decimal and 32-hex file ids; returns {PageId: bytes}."""
    rng = np.random.default_rng(seed)
    want = {}
    for k in range(n):
        fid = str(k % 7) if k % 3 else "%032x" % (0xabc0000000000000000000000000000 + k % 5)
        length = [0, 1, 17, PAGE - 1, PAGE][k % 5] if k < 10 else int(rng.integers(0, PAGE + 1))
        data = orc.fill_page(seed, k % 7, k, PAGE)[:length].tobytes()
        pid = _pid(fid, k)
        store.put(pid, data)
        want[pid] = data
    return want


def test_save_load_round_trip_and_manager_restore(tmp_path):
    from alluxio_b200.cache import LocalCacheManager
    path = str(tmp_path / "cache.snap")
    a = _store()
    want = _fill(a, 300)
    assert a.save_snapshot(path) == 300
    assert not os.path.exists(path + ".tmp")
    a.close()

    b = _store()
    assert b.get_pages() == []
    assert b.load_snapshot(path) == (300, 0)
    got = dict(b.get_pages())
    assert got == {p: len(d) for p, d in want.items()}
    for pid, data in want.items():
        buf = bytearray(PAGE)
        assert b.get(pid, buf) == len(data) and bytes(buf[:len(data)]) == data
        if data:
            assert b.last_crc == orc.crc32c(np.frombuffer(data, np.uint8))
    # the manager's restore adopts what the store lists
    m = LocalCacheManager.over_store(b, PAGE, evictor="LRU")
    assert m.state() == m.READ_WRITE
    assert m.pages() == 300 and m.bytes() == sum(len(d) for d in want.values())
    pid, data = next((p, d) for p, d in want.items() if len(d) == PAGE)
    buf = bytearray(PAGE)
    assert m.get(pid, PAGE, buf) == PAGE and bytes(buf) == data
    m.close()
    b.close()


def test_load_into_smaller_store_rejects_the_overflow(tmp_path):
    path = str(tmp_path / "big.snap")
    a = _store()
    _fill(a, 200)
    assert a.save_snapshot(path) == 200
    a.close()
    b = _store(cache_pages=64, slab_slots=64)
    loaded, rejected = b.load_snapshot(path)
    assert loaded == 64 and rejected == 136 and len(b.get_pages()) == 64
    b.close()


def test_bad_snapshots_are_refused(tmp_path):
    from alluxio_b200.page_store import IOExceptionJ
    path = str(tmp_path / "s.snap")
    a = _store()
    _fill(a, 40)
    a.save_snapshot(path)
    a.close()
    raw = open(path, "rb").read()
    b = _store()
    with pytest.raises(IOExceptionJ):                       # missing file
        b.load_snapshot(str(tmp_path / "absent.snap"))
    open(path, "wb").write(raw[:len(raw) - 100])            # truncated
    with pytest.raises(IOExceptionJ):
        b.load_snapshot(path)
    open(path, "wb").write(b"not a snapshot" + raw[14:])    # wrong magic
    with pytest.raises(IOExceptionJ):
        b.load_snapshot(path)
    assert b.get_pages() == []
    # a payload byte flipped on disk: that page fails its CRC32C and is not kept, the others load
    recs = np.frombuffer(raw[len(raw) - 40 * 32:], dtype=np.dtype([("key", "<u8", 3), ("len", "<u4"), ("crc", "<u4")]))
    starts = 40 + np.concatenate([[0], np.cumsum((recs["len"].astype(np.int64) + 15) // 16 * 16)])[:-1]
    victim = int(np.argmax(recs["len"] >= 8))
    bad = bytearray(raw)
    bad[int(starts[victim]) + 3] ^= 0x40
    open(path, "wb").write(bytes(bad))
    loaded, rejected = b.load_snapshot(path)
    assert (loaded, rejected) == (39, 1)
    b.close()
    c = _store(page_size=8192)                              # another page size: like a LocalPageStore dir of another size
    open(path, "wb").write(raw)
    with pytest.raises(ValueError):
        c.load_snapshot(path)
    c.close()


def test_snapshot_without_crc_and_empty_store(tmp_path):
    path = str(tmp_path / "e.snap")
    a = _store(crc=False)
    assert a.save_snapshot(path) == 0
    b = _store(crc=False)
    assert b.load_snapshot(path) == (0, 0)
    want = _fill(a, 25)
    assert a.save_snapshot(path) == 25
    assert b.load_snapshot(path) == (25, 0)
    for pid, data in want.items():
        buf = bytearray(PAGE)
        assert b.get(pid, buf) == len(data) and bytes(buf[:len(data)]) == data
    a.close()
    b.close()


def test_interned_file_ids_survive_a_restart(tmp_path):
    """File ids that are neither decimal longs nor 32 hex digits (UUIDs, raw paths) are interned by the calling process.
    LocalPageStore keeps the id in the page's path, so getPages() returns it after a restart
    (CACHE/store/LocalPageStore.java:177-199); the snapshot carries the intern table to give the same answer."""
    from alluxio_b200.cache import LocalCacheManager
    ids = ["/data/warehouse/part-0001.parquet", "123e4567-e89b-12d3-a456-426614174000", "7", "%032x" % 0xabcdef0123456789abcdef]
    # (a) through the Python mirror of GpuPageStore
    path = str(tmp_path / "a.snap")
    a = _store()
    want = {}
    for k, f in enumerate(ids * 3):
        want[_pid(f, k)] = orc.fill_page(5, k, k, 100 + k).tobytes()
        a.put(_pid(f, k), want[_pid(f, k)])
    assert a.save_snapshot(path) == len(want)
    a.close()
    b = _store()                       # a fresh process would start with an empty intern table, as this store does
    b.put(_pid("some/other/name", 0), b"x")   # ... unless it interned something else first: then the tables must agree
    with pytest.raises(ValueError):
        b.load_snapshot(path)
    b.close()
    b = _store()
    assert b.load_snapshot(path) == (len(want), 0)
    assert dict(b.get_pages()) == {p: len(d) for p, d in want.items()}
    for pid, data in want.items():
        buf = bytearray(PAGE)
        assert b.get(pid, buf) == len(data) and bytes(buf[:len(data)]) == data
    b.close()
    # (b) through the native manager: save with its table, restore into a new store + manager
    path2 = str(tmp_path / "b.snap")
    m = LocalCacheManager(PAGE, 256 * PAGE, evictor="LRU")
    for k, f in enumerate(ids * 2):
        assert m.put(_pid(f, k), want[_pid(f, k)])
    assert m.save_snapshot(path2) == 2 * len(ids)
    m.close()
    c = _store(cache_pages=256)
    m2 = LocalCacheManager.over_snapshot(c, PAGE, path2, evictor="LRU")
    assert m2.state() == m2.READ_WRITE and m2.pages() == 2 * len(ids)
    for k, f in enumerate(ids * 2):
        data, buf = want[_pid(f, k)], bytearray(PAGE)
        assert m2.get(_pid(f, k), len(data), buf) == len(data) and bytes(buf[:len(data)]) == data
    m2.close()
    c.close()
    # (c) a snapshot written WITHOUT a table leaves interned ids out instead of restoring them under a wrong id
    from alluxio_b200 import capi
    import ctypes as C
    d = _store()
    for k, f in enumerate(ids):
        d.put(_pid(f, k), b"abc")
    n = C.c_uint64()
    capi.check(capi.lib().aps_snapshot_save(d.handle, os.fsencode(str(tmp_path / "c.snap")), C.byref(n)))
    assert n.value == 2                # the decimal and the 32-hex id
    d.close()
