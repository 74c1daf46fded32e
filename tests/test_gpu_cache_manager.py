"""-m gpu: the native LocalCacheManager driving the GPU page store -- the reference's
LocalCacheManagerTest cases (CTEST/LocalCacheManagerTest.java), LocalCacheFileInStreamTest-style
reads, and whole-trace replays compared element by element with the oracle (hit/miss vector, put
results, victim sequence, final residency)."""
import numpy as np
import pytest

from conftest import increasing
from oracle import binding as orc

pytestmark = pytest.mark.gpu

PAGE_SIZE = 1024
CACHE_SIZE = 512 * 1024
PAGE1, PAGE2 = increasing(PAGE_SIZE), increasing(255, PAGE_SIZE)


def PID(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def manager(cache_size=CACHE_SIZE, evictor="FIFO", page_size=PAGE_SIZE, slab_slots=0, **kw):
    from alluxio_b200.cache import LocalCacheManager
    return LocalCacheManager(page_size, cache_size, evictor=evictor, slab_slots=slab_slots, **kw)


PAGE_ID1, PAGE_ID2 = ("0L", 0), ("1L", 1)


def test_put_new_put_exist_get_delete():
    m = manager()
    p1, p2 = PID(*PAGE_ID1), PID(*PAGE_ID2)
    buf = bytearray(PAGE_SIZE)
    assert m.state() == m.READ_WRITE
    assert m.get(p1, PAGE_SIZE, buf) == 0                       # getNotExist
    assert not m.delete(p1)                                     # deleteNotExist
    assert m.put(p1, PAGE1)
    assert m.put(p1, PAGE2)                                     # putExist: old bytes stay
    assert m.get(p1, PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE1
    for off in (0, 1, PAGE_SIZE // 2, PAGE_SIZE - 1):           # getOffset
        assert m.get(p1, PAGE_SIZE - off, buf, page_offset=off) == PAGE_SIZE - off
        assert bytes(buf[:PAGE_SIZE - off]) == PAGE1[off:]
    with pytest.raises(ValueError):                             # getNotEnoughSpaceException
        m.get(p1, PAGE_SIZE, bytearray(PAGE_SIZE - 1))
    with pytest.raises(ValueError):
        m.get(p1, 1, buf, page_offset=PAGE_SIZE + 1)
    assert m.delete(p1)
    assert m.get(p1, PAGE_SIZE, buf) == 0
    assert (m.bytes(), m.pages()) == (0, 0)
    m.close()


def test_put_evict_and_more_than_capacity():
    m = manager(cache_size=PAGE_SIZE)
    p1, p2 = PID(*PAGE_ID1), PID(*PAGE_ID2)
    buf = bytearray(PAGE_SIZE)
    assert m.put(p1, PAGE1) and m.put(p2, PAGE2)
    assert m.get(p1, PAGE_SIZE, buf) == 0
    assert m.get(p2, PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE2
    m.close()
    m = manager()
    cache_pages = CACHE_SIZE // PAGE_SIZE
    for i in range(2 * cache_pages):                            # putMoreThanCacheCapacity (FIFO order)
        m.put(PID("3", i), increasing(i, PAGE_SIZE))
        if i >= cache_pages:
            assert m.get(PID("3", i - cache_pages), PAGE_SIZE, buf) == 0
            assert m.get(PID("3", i - cache_pages + 1), PAGE_SIZE, buf) == PAGE_SIZE
            assert bytes(buf) == increasing(i - cache_pages + 1, PAGE_SIZE)
    m.close()


def test_byte_based_capacity_one_victim_per_put():
    small, n = 8, PAGE_SIZE // 8
    m = manager(cache_size=PAGE_SIZE, slab_slots=n + 8)         # short pages need their own slots
    for i in range(n):                                          # putSmallPages
        assert m.put(PID(i, 0), increasing(i, small))
    for i in range(n):
        b = bytearray(small)
        assert m.get(PID(i, 0), small, b) == small and bytes(b) == increasing(i, small)
    assert m.put(PID(n, 0), increasing(-1 & 0xFF, small))       # evictSmallPageByPutSmallPage
    assert m.get(PID(0, 0), small, bytearray(small)) == 0
    big, big_id = increasing(0, PAGE_SIZE), PID(-1, 0)
    for _ in range(n - 1):                                      # evictSmallPagesByPutPigPage
        assert not m.put(big_id, big)
    assert m.put(big_id, big)
    b = bytearray(PAGE_SIZE)
    assert m.get(big_id, PAGE_SIZE, b) == PAGE_SIZE and bytes(b) == big
    for i in range(n):                                          # evictBigPagesByPutSmallPage
        assert m.put(PID(1000 + i, 0), increasing(small))
    m.close()


def test_partial_pages_large_ids_and_short_read_drops_page():
    m = manager()
    for size in (1, PAGE_SIZE // 2, PAGE_SIZE - 1):             # putGetPartialPage
        pid, data = PID("3", size), increasing(0, size)
        assert m.put(pid, data)
        b = bytearray(size)
        assert m.get(pid, size, b) == size and bytes(b) == data
    for idx in (2**63 - 2, 2**63 - 1):                          # putLargeId
        pid = PID("0", idx)
        m.put(pid, PAGE1)
        b = bytearray(PAGE_SIZE)
        assert m.get(pid, PAGE_SIZE, b) == PAGE_SIZE and bytes(b) == PAGE1
    pid = PID("4", 0)
    m.put(pid, increasing(0, 100))
    assert m.get(pid, 200, bytearray(200)) == -1                # ret != bytesToRead -> -1, page dropped
    assert m.get(pid, 100, bytearray(100)) == 0
    m.close()


def test_get_of_zero_bytes_is_an_error_and_drops_the_page():
    """LocalCacheManager.get treats bytesRead <= 0 as a failed read (LocalCacheManager.java:383-395): asking a resident
    page for 0 bytes returns -1 and removes the page from the meta store.  Result and residency afterwards must equal
    the oracle's, for a full page, for an empty page and through the batched read path."""
    ostore = orc.PageStore.open_mem(PAGE_SIZE, CACHE_SIZE)
    om = orc.LocalCacheManager(ostore, PAGE_SIZE, "LRU")
    m = manager(evictor="LRU")
    script = [("put", "7", 0, PAGE1), ("get", "7", 0, 0), ("get", "7", 0, PAGE_SIZE),      # dropped by the 0-byte get
              ("put", "7", 0, PAGE2), ("get", "7", 0, PAGE_SIZE),                           # re-admitted: store overwritten
              ("put", "8", 3, b""), ("get", "8", 3, 0), ("get", "8", 3, 0),                 # empty page: -1, then a miss
              ("put", "9", 1, PAGE1), ("get", "9", 1, 10), ("get", "9", 1, 0), ("delete", "9", 1, None)]
    for op, f, i, arg in script:
        if op == "put":
            assert m.put(PID(f, i), arg) == om.put(orc.PageId(f, i), arg)
        elif op == "delete":
            assert m.delete(PID(f, i)) == om.delete(orc.PageId(f, i))
        else:
            b1, b2 = bytearray(max(arg, 1)), bytearray(max(arg, 1))
            r1, r2 = m.get(PID(f, i), arg, b1), om.get(orc.PageId(f, i), arg, b2)
            assert r1 == r2, (op, f, i, arg, r1, r2)
            if r1 > 0:
                assert bytes(b1[:r1]) == bytes(b2[:r2])
        assert (m.pages(), m.bytes()) == (om.pages(), om.bytes())
    m.close()


def test_store_faults_leave_no_state():
    m = manager()
    p1, p2 = PID(*PAGE_ID1), PID(*PAGE_ID2)
    buf = bytearray(PAGE_SIZE)
    m.set_store_faulty(put=True)                                # recoverCacheFromFailedPut
    assert not m.put(p1, PAGE1)
    assert m.get(p1, PAGE_SIZE, buf) == 0
    assert not m.put(p1, PAGE1)
    assert m.metrics()["PutStoreWriteErrors"] == 2 and m.metrics()["PutErrors"] == 2
    m.set_store_faulty()
    assert m.put(p1, PAGE1)
    assert m.get(p1, PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE1
    m.close()
    m = manager(cache_size=PAGE_SIZE)                           # failedPageStoreDeleteOnEviction
    m.set_store_faulty(delete=True)
    assert m.put(p1, PAGE1)
    assert not m.put(p2, PAGE2)
    assert m.metrics()["PutStoreDeleteErrors"] == 1 and m.metrics()["PagesEvicted"] == 0
    m.set_store_faulty()
    assert m.put(p1, PAGE1)
    assert m.get(p1, PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE1
    m.close()


def test_lock_ids_match_oracle():
    m = manager()
    store = orc.PageStore.open_mem(PAGE_SIZE, CACHE_SIZE)
    om = orc.LocalCacheManager(store, PAGE_SIZE)
    for f, i in [("0", 0), ("0", 2**63 - 1), ("polygenelubricants", 5), ("d41d8cd98f00b204e9800998ecf8427e", 77),
                 ("/some/path/é", 12345678901234)]:
        assert m.page_lock_id(PID(f, i)) == om.page_lock_id(orc.PageId(f, i))
    m.close()


def test_positioned_read_through_the_cache():
    """CTEST/LocalCacheFileInStreamTest.java: reads across page boundaries; a miss fetches the whole
    page from outside and caches it; the last page is short."""
    ps = 4096
    m = manager(cache_size=64 * ps, evictor="LRU", page_size=ps)
    file_len = 5 * ps + 123
    data = np.random.default_rng(5).integers(0, 256, file_len, dtype=np.uint8).tobytes()
    calls = []

    def external(pos, n):
        calls.append((pos, n))
        return data[pos:pos + n]

    got, fc, fe = m.positioned_read("77", file_len, 100, 3 * ps, external)
    assert got == data[100:100 + 3 * ps] and fc == 0 and fe == 3 * ps
    assert [c[0] for c in calls] == [0, ps, 2 * ps, 3 * ps] and all(c[1] == ps for c in calls)
    calls.clear()
    got, fc, fe = m.positioned_read("77", file_len, 0, 4 * ps, external)     # all four pages now hit
    assert got == data[:4 * ps] and (fc, fe) == (4 * ps, 0) and not calls
    got, fc, fe = m.positioned_read("77", file_len, 5 * ps - 10, 1000, external)   # crosses into the short last page
    assert got == data[5 * ps - 10:] and len(got) == 133
    assert calls == [(4 * ps, ps), (5 * ps, 123)]
    got, _, _ = m.positioned_read("77", file_len, file_len, 10, external)
    assert got is None                                                        # EOF
    assert m.pages() == 6 and m.bytes() == file_len
    m.close()


@pytest.mark.parametrize("evictor", ["LRU", "LFU", "FIFO"])
def test_replay_matches_oracle_element_by_element(evictor):
    """BASELINE config 3 in miniature: 64 KiB pages, Zipf(0.99) accesses, capacity << universe; the
    evictor on the host drives GPU table inserts and erases.  Everything observable must equal the
    oracle's replay of the same trace; returned bytes are checked against the generator."""
    ps = 65536
    n_ids, cap_pages, n_acc = 6000, 400, 12000
    rng = np.random.default_rng(3)
    ranks = np.arange(1, n_ids + 1, dtype=np.float64)
    pz = ranks ** -0.99
    pz /= pz.sum()
    ids = rng.permutation(n_ids)[rng.choice(n_ids, size=n_acc, p=pz)]
    files, pages = ids // 100, ids % 100
    lens = np.where(pages == 99, 1000 + files, ps).astype(np.int32)   # last page of every file is short
    lens[(pages == 98) & (files % 3 == 0)] = 0   # empty pages: a get of 0 bytes counts as a failed read and drops the page
    cache_bytes = cap_pages * ps
    want = orc.replay(evictor, ps, cache_bytes, files, pages, lens)
    m = manager(cache_size=cache_bytes, evictor=evictor, page_size=ps, slab_slots=cap_pages + 128)
    got = m.replay(files, pages, lens, seed=0xA11C10, verify=True)
    assert got["verify_failures"] == 0
    assert np.array_equal(got["hit"], want["hit"])
    assert np.array_equal(got["put_ok"], want["put_ok"])
    assert got["n_victims"] == want["n_victims"] > 0
    assert np.array_equal(got["victims"], want["victims"])
    assert (got["final_pages"], got["final_bytes"]) == (want["final_pages"], want["final_bytes"])
    st = m.store_stats()
    assert st["pages"] == got["final_pages"] and st["bytes"] == got["final_bytes"] and st["crc_mismatches"] == 0
    m.close()


@pytest.mark.parametrize("evictor", ["LRU", "LFU"])
def test_batched_replay_equals_oracle_and_single_page_replay(evictor):
    """The batched driver (victim deletes, miss puts and hit gets issued as GPU batches) takes the same
    decisions as the reference flow: identical hit vector, put results and victim order as the oracle."""
    ps = 65536
    n_ids, cap_pages, n_acc = 20000, 1500, 60000
    rng = np.random.default_rng(11)
    pz = np.arange(1, n_ids + 1, dtype=np.float64) ** -0.99
    pz /= pz.sum()
    ids = rng.permutation(n_ids)[rng.choice(n_ids, size=n_acc, p=pz)]
    files, pages = ids // 100, ids % 100
    lens = np.where(pages == 99, 2000 + files, ps).astype(np.int32)
    lens[(pages == 98) & (files % 3 == 0)] = 0   # empty pages (LocalCacheManager.java:383-395: bytesRead <= 0 is an error)
    cache_bytes = cap_pages * ps
    want = orc.replay(evictor, ps, cache_bytes, files, pages, lens)
    m = manager(cache_size=cache_bytes, evictor=evictor, page_size=ps, slab_slots=cap_pages + 256, max_batch=1024)
    got = m.replay_batched(files, pages, lens, seed=0xA11C10, window=1024)
    assert got["bad_gets"] == 0
    assert np.array_equal(got["hit"], want["hit"])
    assert np.array_equal(got["put_ok"], want["put_ok"])
    assert got["n_victims"] == want["n_victims"] > 0
    assert np.array_equal(got["victims"], want["victims"])
    assert (got["final_pages"], got["final_bytes"]) == (want["final_pages"], want["final_bytes"])
    assert got["gets"] == int(want["hit"].sum()) and got["deletes"] == want["n_victims"]
    st = m.store_stats()
    assert st["pages"] == got["final_pages"] and st["bytes"] == got["final_bytes"] and st["crc_mismatches"] == 0
    # residency on the device equals the host's view: every id the oracle says is resident reads back
    buf = bytearray(ps)
    last_seen = {}
    for k in range(n_acc):
        last_seen[(int(files[k]), int(pages[k]))] = int(lens[k])
    resident = 0
    for (f, p), ln in list(last_seen.items())[:400]:
        r = m.get(PID(f, p), ln, buf)
        if r > 0:
            resident += 1
            assert bytes(buf[:ln]) == orc.fill_page(0xA11C10, f, p, ln).tobytes()
    assert resident > 0
    m.close()


def test_restore_over_a_populated_store():
    """CTEST/LocalCacheManagerTest.java:357-369 (syncRestore) and :494-511 (more pages than capacity): a manager
    created over a store that already holds pages adopts what fits and deletes the rest from the store."""
    import uuid
    from alluxio_b200.cache import LocalCacheManager
    from alluxio_b200.page_store import GpuPageStore
    store = GpuPageStore(page_size=PAGE_SIZE, cache_size=CACHE_SIZE)
    uid = PID(str(uuid.uuid4()), 0)
    store.put(PID(*PAGE_ID1), PAGE1)
    store.put(uid, PAGE2)
    m = LocalCacheManager.over_store(store, PAGE_SIZE, "FIFO")
    buf = bytearray(PAGE_SIZE)
    assert m.state() == m.READ_WRITE
    assert m.get(PID(*PAGE_ID1), PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE1
    assert m.get(uid, PAGE_SIZE, buf) == PAGE_SIZE and bytes(buf) == PAGE2
    assert (m.pages(), m.bytes()) == (2, 2 * PAGE_SIZE)
    m.close()
    store.close()
    # capacity for two pages, three in the store: one is discarded from the store during restore
    store = GpuPageStore(page_size=PAGE_SIZE, cache_size=2 * PAGE_SIZE, slab_slots=16)
    for i in range(3):
        store.put(PID("5", i), increasing(i, PAGE_SIZE))
    m = LocalCacheManager.over_store(store, PAGE_SIZE, "LRU")
    assert (m.pages(), m.bytes()) == (2, 2 * PAGE_SIZE)
    assert len(store.get_pages()) == 2
    hits = sum(m.get(PID("5", i), PAGE_SIZE, buf) == PAGE_SIZE for i in range(3))
    assert hits == 2
    m.close()
    store.close()


def test_no_exception_manager_and_metrics():
    m = manager(cache_size=2 * PAGE_SIZE)
    p1, p2, p3 = PID("a", 0), PID("a", 1), PID("a", 2)
    buf = bytearray(PAGE_SIZE)
    assert m.get_noexcept(p1, PAGE_SIZE, bytearray(PAGE_SIZE - 1)) == -1      # IllegalArgument swallowed
    assert m.put(p1, PAGE1) and m.put(p2, PAGE2) and m.put(p3, PAGE1)          # third put evicts one page
    assert not m.delete(PID("a", 99))
    assert m.get_noexcept(p3, PAGE_SIZE, buf) == PAGE_SIZE
    mt = m.metrics()
    assert mt["GetErrors"] == 1 and mt["BytesWrittenCache"] == 3 * PAGE_SIZE
    assert mt["PagesEvicted"] == 1 and mt["BytesEvicted"] == PAGE_SIZE
    assert mt["DeleteErrors"] == 1 and mt["DeleteNonExistingPageErrors"] == 1
    m.close()


def test_batched_positioned_read_equals_per_page_reads():
    """a multi-page read becomes one batched store call; same bytes, same cache effects."""
    ps = 4096
    file_len = 9 * ps + 777
    data = np.random.default_rng(8).integers(0, 256, file_len, dtype=np.uint8).tobytes()
    a = manager(cache_size=64 * ps, evictor="LRU", page_size=ps)
    b = manager(cache_size=64 * ps, evictor="LRU", page_size=ps)
    ext = lambda pos, n: data[pos:pos + n]  # noqa: E731
    rng = np.random.default_rng(9)
    for _ in range(40):
        pos = int(rng.integers(0, file_len))
        n = int(rng.integers(1, 4 * ps))
        ra = a.positioned_read("12", file_len, pos, n, ext)
        rb = b.positioned_read("12", file_len, pos, n, ext, batched=True)
        assert ra == rb and ra[0] == data[pos:pos + n]
    assert (a.pages(), a.bytes()) == (b.pages(), b.bytes())
    got, fc, fe = b.positioned_read("12", file_len, 0, file_len, ext, batched=True)
    got, fc, fe = b.positioned_read("12", file_len, 0, file_len, ext, batched=True)
    assert got == data and fc == file_len and fe == 0            # second full read: every page from the cache
    assert b.metrics()["BytesReadCache"] >= file_len
    a.close()
    b.close()
