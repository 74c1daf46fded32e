"""Port of CTEST/LocalCacheManagerTest.java (sync-write, sync-restore cases) and of the restore
cases in tests/src/test/java/alluxio/client/fs/LocalCacheManagerIntegrationTest.java against the
oracle.  Thread-only behaviours (async queue rejection :527-577, timeouts :618-655, the slow-scan
READ_ONLY window :400-418) need concurrency and are outside a serialised replay."""
import os
import uuid

import pytest

from conftest import increasing
from oracle import binding as orc
from oracle.binding import PageId

PAGE_SIZE = 1024
CACHE_SIZE = 512 * 1024
PAGE_ID1, PAGE_ID2 = PageId("0L", 0), PageId("1L", 1)
PAGE1, PAGE2 = increasing(PAGE_SIZE), increasing(255, PAGE_SIZE)


def page(i, n):
    return increasing(i, n)


def page_id(i, idx):
    return PageId(str(i), idx)


class Env:
    def __init__(self, tmp_path, cache_size=CACHE_SIZE, evictor="FIFO", wipe=True):
        self.root = tmp_path / "LOCAL"  # PageStore.getStorePath(type, dir), PageStore.java:85-87
        mk = orc.PageStore.create_local if wipe else orc.PageStore.open_local
        self.store = mk(self.root, page_size=PAGE_SIZE, cache_size=cache_size)
        self.mgr = orc.LocalCacheManager(self.store, PAGE_SIZE, evictor)
        self.buf = bytearray(PAGE_SIZE)


@pytest.fixture
def env(tmp_path):
    return Env(tmp_path)


def test_create_nonexistent_root_dir(tmp_path):
    e = Env(tmp_path / str(uuid.uuid4()))
    assert e.mgr.state() == orc.LocalCacheManager.READ_WRITE


def test_put_new(env):
    assert env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == len(PAGE1)
    assert bytes(env.buf) == PAGE1


def test_put_exist_keeps_old_bytes(env):
    assert env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.put(PAGE_ID1, PAGE2)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == len(PAGE1)
    assert bytes(env.buf) == PAGE1


def test_put_evict(tmp_path):
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    assert e.mgr.put(PAGE_ID1, PAGE1)
    assert e.mgr.put(PAGE_ID2, PAGE2)
    assert e.mgr.get(PAGE_ID1, len(PAGE1), e.buf) == 0
    assert e.mgr.get(PAGE_ID2, len(PAGE1), e.buf) == len(PAGE2)
    assert bytes(e.buf) == PAGE2


def test_put_small_pages(tmp_path):
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    small = 8
    n = PAGE_SIZE // small
    for i in range(n):
        assert e.mgr.put(page_id(i, 0), page(i, small))
    for i in range(n):
        buf = bytearray(small)
        assert e.mgr.get(page_id(i, 0), small, buf) == small
        assert bytes(buf) == page(i, small)


def test_evict_small_page_by_put_small_page(tmp_path):
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    small = 8
    n = PAGE_SIZE // small
    for i in range(n):
        assert e.mgr.put(page_id(i, 0), page(i, small))
    assert e.mgr.put(page_id(n, 0), page(-1, small))
    for i in range(n):
        buf = bytearray(small)
        if i == 0:
            assert e.mgr.get(page_id(i, 0), small, buf) == 0
        else:
            assert e.mgr.get(page_id(i, 0), small, buf) == small
            assert bytes(buf) == page(i, small)


def test_evict_small_pages_by_put_big_page(tmp_path):
    # one victim per put: numPages-1 failures, then success
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    small = 8
    n = PAGE_SIZE // small
    for i in range(n):
        assert e.mgr.put(page_id(i, 0), page(i, small))
    big, big_id = page(0, PAGE_SIZE), page_id(-1, 0)
    for _ in range(n - 1):
        assert not e.mgr.put(big_id, big)
    assert e.mgr.put(big_id, big)
    buf = bytearray(PAGE_SIZE)
    assert e.mgr.get(big_id, PAGE_SIZE, buf) == PAGE_SIZE
    assert bytes(buf) == big


def test_evict_big_pages_by_put_small_page(tmp_path):
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    assert e.mgr.put(page_id(-1, 0), page(0, PAGE_SIZE))
    small = increasing(8)
    for i in range(PAGE_SIZE // 8):
        assert e.mgr.put(page_id(i, 0), small)


def test_put_get_partial_page(env):
    for size in (1, PAGE_SIZE // 2, PAGE_SIZE - 1):
        pid = PageId("3", size)
        data = page(0, size)
        assert env.mgr.put(pid, data)
        buf = bytearray(size)
        assert env.mgr.get(pid, size, buf) == size
        assert bytes(buf) == data


def test_put_more_than_cache_capacity(env):
    cache_pages = CACHE_SIZE // PAGE_SIZE
    for i in range(2 * cache_pages):
        env.mgr.put(PageId("3", i), page(i, PAGE_SIZE))
        if i >= cache_pages:
            assert env.mgr.get(PageId("3", i - cache_pages), PAGE_SIZE, env.buf) == 0
            assert env.mgr.get(PageId("3", i - cache_pages + 1), PAGE_SIZE, env.buf) == PAGE_SIZE
            assert bytes(env.buf) == page(i - cache_pages + 1, PAGE_SIZE)


def test_put_large_id(env):
    for idx in (2**63 - 2, 2**63 - 1):
        pid = PageId("0", idx)
        env.mgr.put(pid, PAGE1)
        assert env.mgr.get(pid, len(PAGE1), env.buf) == PAGE_SIZE
        assert bytes(env.buf) == PAGE1


def test_get_exist_and_not_exist(env):
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == 0
    env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == PAGE_SIZE
    assert bytes(env.buf) == PAGE1


def test_get_offset(env):
    env.mgr.put(PAGE_ID1, PAGE1)
    for off in (0, 1, PAGE_SIZE // 2, PAGE_SIZE - 1):
        assert env.mgr.get(PAGE_ID1, len(PAGE1) - off, env.buf, page_offset=off) == len(PAGE1) - off
        assert bytes(env.buf[:len(PAGE1) - off]) == PAGE1[off:]


def test_get_not_enough_space_exception(env):
    with pytest.raises(ValueError):
        env.mgr.get(PAGE_ID1, len(PAGE1), bytearray(len(PAGE1) - 1))


def test_get_offset_beyond_page_size_exception(env):
    # LocalCacheManager.java:362-363
    with pytest.raises(ValueError):
        env.mgr.get(PAGE_ID1, 1, env.buf, page_offset=PAGE_SIZE + 1)


def test_get_more_than_stored_returns_minus_one_and_drops_page(env):
    # LocalCacheManager.java:384-395,544-549: ret != bytesToRead => -1 and page removed from metastore
    pid = PageId("3", 0)
    env.mgr.put(pid, page(0, 100))
    assert env.mgr.get(pid, 200, bytearray(200)) == -1
    assert env.mgr.get(pid, 100, bytearray(100)) == 0


def test_get_of_zero_bytes_returns_minus_one_and_drops_page(env):
    """LocalCacheManager.get: getPage returns bytesToRead (0), get treats bytesRead <= 0 as a failed read, removes the page
    from the meta store and returns -1 (CACHE/LocalCacheManager.java:383-395, 540-555).  The page store keeps the file."""
    assert env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, 0, env.buf) == -1
    assert env.mgr.pages() == 0 and env.mgr.bytes() == 0
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == 0      # a miss now
    assert env.mgr.put(PAGE_ID1, PAGE2)                          # re-admitted; the store's file is overwritten
    assert env.mgr.get(PAGE_ID1, len(PAGE2), env.buf) == len(PAGE2) and bytes(env.buf) == PAGE2


def test_delete(env):
    assert not env.mgr.delete(PAGE_ID1)
    env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.delete(PAGE_ID1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == 0


def test_sync_restore(tmp_path):
    e = Env(tmp_path)
    uid = PageId(str(uuid.uuid4()), 0)
    e.store.put(PAGE_ID1, PAGE1)
    e.store.put(uid, PAGE2)
    mgr = orc.LocalCacheManager(e.store, PAGE_SIZE, "FIFO")
    assert mgr.state() == orc.LocalCacheManager.READ_WRITE
    assert mgr.get(PAGE_ID1, len(PAGE1), e.buf) == len(PAGE1) and bytes(e.buf) == PAGE1
    assert mgr.get(uid, len(PAGE2), e.buf) == len(PAGE2) and bytes(e.buf) == PAGE2


def test_sync_restore_unknown_file_wipes_cache(tmp_path):
    e = Env(tmp_path)
    uid = PageId(str(uuid.uuid4()), 0)
    e.store.put(PAGE_ID1, PAGE1)
    e.store.put(uid, PAGE2)
    (e.root / "invalidPageFile").write_bytes(b"")
    mgr = orc.LocalCacheManager(e.store, PAGE_SIZE, "FIFO")
    assert mgr.state() == orc.LocalCacheManager.READ_WRITE
    assert mgr.get(PAGE_ID1, len(PAGE1), e.buf) == 0
    assert mgr.get(uid, len(PAGE2), e.buf) == 0
    assert os.listdir(e.root) == []


def test_sync_restore_with_more_pages_than_capacity(tmp_path):
    e = Env(tmp_path)
    uid = PageId(str(uuid.uuid4()), 0)
    e.store.put(PAGE_ID1, PAGE1)
    e.store.put(PAGE_ID2, PAGE2)
    e.store.put(uid, increasing(len(PAGE1) + len(PAGE2) + 1))
    e2 = Env(tmp_path, cache_size=len(PAGE1) + len(PAGE2), wipe=False)
    assert e2.mgr.get(PAGE_ID1, len(PAGE1), e2.buf) == len(PAGE1) and bytes(e2.buf) == PAGE1
    assert e2.mgr.get(PAGE_ID2, len(PAGE2), e2.buf) == len(PAGE2) and bytes(e2.buf) == PAGE2
    assert e2.mgr.get(uid, len(PAGE2), e2.buf) == 0


def test_restore_with_different_page_size_wipes(tmp_path):
    # LocalCacheManagerIntegrationTest.java:161-175: paths carry the page size, so a new page size
    # does not recognise the old files
    e = Env(tmp_path)
    e.mgr.put(PAGE_ID1, PAGE1)
    store2 = orc.PageStore.open_local(e.root, page_size=2 * PAGE_SIZE, cache_size=CACHE_SIZE)
    mgr2 = orc.LocalCacheManager(store2, 2 * PAGE_SIZE, "LRU")
    assert mgr2.state() == orc.LocalCacheManager.READ_WRITE
    assert mgr2.get(PAGE_ID1, len(PAGE1), bytearray(PAGE_SIZE)) == 0


def test_recover_cache_from_failed_put(env):
    env.store.set_faulty(put=True)
    assert not env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == 0
    assert not env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == 0
    env.store.set_faulty(put=False)
    assert env.mgr.put(PAGE_ID1, PAGE1)
    assert env.mgr.get(PAGE_ID1, len(PAGE1), env.buf) == PAGE_SIZE and bytes(env.buf) == PAGE1


def test_failed_page_store_delete_on_eviction(tmp_path):
    e = Env(tmp_path, cache_size=PAGE_SIZE)
    e.store.set_faulty(delete=True)
    assert e.mgr.put(PAGE_ID1, PAGE1)
    assert not e.mgr.put(PAGE_ID2, PAGE2)
    e.store.set_faulty(delete=False)
    assert e.mgr.put(PAGE_ID1, PAGE1)
    assert e.mgr.get(PAGE_ID1, len(PAGE1), e.buf) == PAGE_SIZE and bytes(e.buf) == PAGE1


def test_lru_in_reference_is_put_order(tmp_path):
    # LocalCacheManager.get uses hasPage, which never touches the evictor, so a hit
    # does not refresh recency and the "LRU" victim is the oldest PUT.
    e = Env(tmp_path, cache_size=2 * PAGE_SIZE, evictor="LRU")
    a, b, c = PageId("7", 0), PageId("7", 1), PageId("7", 2)
    e.mgr.put(a, PAGE1); e.mgr.put(b, PAGE1)
    assert e.mgr.get(a, PAGE_SIZE, e.buf) == PAGE_SIZE   # hit on a
    e.mgr.put(c, PAGE1)                                   # evicts a, not b
    assert e.mgr.get(a, PAGE_SIZE, e.buf) == 0
    assert e.mgr.get(b, PAGE_SIZE, e.buf) == PAGE_SIZE


def test_page_lock_id(env):
    # LocalCacheManager.java:172-174  floorMod((int)(hash + index), 1024)
    assert env.mgr.page_lock_id(PageId("0", 0)) == 48
    assert env.mgr.page_lock_id(PageId("0", 2**63 - 1)) == (48 + 2**63 - 1) % 2**32 % 1024
    h = orc.java_string_hashcode("polygenelubricants")
    assert env.mgr.page_lock_id(PageId("polygenelubricants", 5)) == ((h + 5) % 2**32) % 1024


def test_replay_matches_step_by_step_manager(tmp_path):
    import numpy as np
    rng = np.random.default_rng(3)
    n = 4000
    files = rng.integers(0, 5, n)
    pages = rng.integers(0, 40, n)
    lens = np.where(pages == 39, 100, PAGE_SIZE).astype(np.int32)
    for ev in ("LRU", "LFU", "FIFO"):
        r = orc.replay(ev, PAGE_SIZE, 64 * PAGE_SIZE, files, pages, lens)
        store = orc.PageStore.open_mem(PAGE_SIZE, 64 * PAGE_SIZE, keep_bytes=False)
        mgr = orc.LocalCacheManager(store, PAGE_SIZE, ev)
        for k in range(n):
            pid = PageId(int(files[k]), int(pages[k]))
            got = mgr.get(pid, int(lens[k]), bytearray(int(lens[k])))
            assert (got > 0) == bool(r["hit"][k])
            if got <= 0:
                assert mgr.put(pid, bytes(int(lens[k]))) == bool(r["put_ok"][k])
        assert mgr.pages() == r["final_pages"] and mgr.bytes() == r["final_bytes"]
        assert r["n_victims"] > 0


# ---- tests/src/test/java/alluxio/client/fs/LocalCacheManagerIntegrationTest.java: a cache that is closed and created again

IT_PAGE_COUNT = 32


def _load_full_cache(tmp_path, evictor="LRU"):
    """loadFullCache (:249-257): 32 pages of 1 KB into a 32 KB cache, all readable."""
    e = Env(tmp_path, cache_size=IT_PAGE_COUNT * PAGE_SIZE, evictor=evictor)
    for i in range(IT_PAGE_COUNT):
        assert e.mgr.put(page_id(0, i), PAGE1)
    for i in range(IT_PAGE_COUNT):
        assert e.mgr.get(page_id(0, i), PAGE_SIZE, e.buf) == PAGE_SIZE and bytes(e.buf) == PAGE1
    return e


def test_load_cache_and_evict(tmp_path):
    """loadCacheAndEvict (:117-144): the restored pages take part in eviction like any others."""
    _load_full_cache(tmp_path)
    e = Env(tmp_path, cache_size=IT_PAGE_COUNT * PAGE_SIZE, evictor="LRU", wipe=False)   # close + create again
    for i in range(IT_PAGE_COUNT // 2):
        assert e.mgr.put(page_id(1, i), PAGE1)
    evicted = 0
    for i in range(IT_PAGE_COUNT):
        ret = e.mgr.get(page_id(0, i), PAGE_SIZE, e.buf)
        if ret <= 0:
            evicted += 1
            continue
        assert ret == PAGE_SIZE and bytes(e.buf) == PAGE1
    assert evicted == IT_PAGE_COUNT // 2
    for i in range(IT_PAGE_COUNT // 2):
        assert e.mgr.get(page_id(1, i), PAGE_SIZE, e.buf) == PAGE_SIZE and bytes(e.buf) == PAGE1


def test_load_cache_smaller_new_cache_size(tmp_path):
    """loadCacheSmallerNewCacheSizeLocal (:183-205): half the budget keeps (at least) half of the pages."""
    _load_full_cache(tmp_path)
    e = Env(tmp_path, cache_size=IT_PAGE_COUNT * PAGE_SIZE // 2, evictor="LRU", wipe=False)
    assert e.mgr.state() == orc.LocalCacheManager.READ_WRITE
    hits = sum(e.mgr.get(page_id(0, i), PAGE_SIZE, e.buf) == PAGE_SIZE for i in range(IT_PAGE_COUNT))
    assert hits == IT_PAGE_COUNT // 2 and e.mgr.pages() == hits
    assert len(e.store.get_pages()) == hits          # the pages that did not fit were deleted from the store (:493-497)


def test_load_cache_larger_new_cache_size(tmp_path):
    """loadCacheLargerNewCacheSize (:218-222)."""
    e = Env(tmp_path, cache_size=IT_PAGE_COUNT * PAGE_SIZE)
    assert e.mgr.put(PAGE_ID1, PAGE1)
    e = Env(tmp_path, cache_size=2 * IT_PAGE_COUNT * PAGE_SIZE, wipe=False)
    assert e.mgr.get(PAGE_ID1, PAGE_SIZE, e.buf) == PAGE_SIZE and bytes(e.buf) == PAGE1
