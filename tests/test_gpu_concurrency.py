"""-m gpu: concurrent callers.  The reference orders operations on the SAME PageId with striped page
locks and lets different ids proceed concurrently (CACHE/LocalCacheManager.java:64-72); the store must
be safe for that.  Bit-exact replay is only defined for a serialised sequence, so these tests assert
invariants: no torn page (every byte read back belongs to one put of that id), no lost insert, byte and
page accounting consistent, CRC checks never fire."""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def PID(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def test_store_threads_on_disjoint_ids():
    from alluxio_b200.page_store import GpuPageStore, PageNotFoundException
    ps = 4096
    store = GpuPageStore(page_size=ps, cache_size=4096 * ps, slab_slots=4096)
    n_threads, per = 8, 120
    errors = []

    def worker(t):
        try:
            rng = np.random.default_rng(t)
            mine = {}
            for step in range(per * 3):
                i = int(rng.integers(per))
                pid = PID("t%d" % t, i)
                op = rng.integers(3)
                if op == 0:
                    n = int(rng.integers(1, ps + 1))
                    data = bytes([(t * 31 + i + step) & 0xFF]) * n
                    store.put(pid, data)
                    mine[i] = data
                elif op == 1:
                    buf = bytearray(ps)
                    try:
                        got = store.get(pid, buf)
                        assert i in mine and bytes(buf[:got]) == mine[i], ("torn or stale page", t, i)
                    except PageNotFoundException:
                        assert i not in mine, ("lost insert", t, i)
                else:
                    try:
                        store.delete(pid)
                        assert i in mine
                        del mine[i]
                    except PageNotFoundException:
                        assert i not in mine
            results[t] = mine
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    results = {}
    threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]
    st = store.stats()
    assert st["pages"] == sum(len(m) for m in results.values())
    assert st["bytes"] == sum(len(d) for m in results.values() for d in m.values())
    assert st["crc_mismatches"] == 0
    resident = {(p.file_id, p.page_index): n for p, n in store.get_pages()}
    for t, m in results.items():
        for i, d in m.items():
            assert resident[("t%d" % t, i)] == len(d)
    store.close()


def test_manager_threads_with_eviction():
    """Several threads read through one LocalCacheManager (get; on a miss put) over a small cache, so
    evictions race with reads of other ids.  Invariants only."""
    from alluxio_b200.cache import LocalCacheManager
    ps = 4096
    m = LocalCacheManager(ps, 64 * ps, evictor="LRU", slab_slots=128)
    errors = []

    def payload(f, i):
        return bytes([(f * 7 + i) & 0xFF]) * ps

    def worker(t):
        try:
            rng = np.random.default_rng(100 + t)
            buf = bytearray(ps)
            for _ in range(600):
                f, i = int(rng.integers(4)), int(rng.integers(60))
                pid = PID(f, i)
                got = m.get(pid, ps, buf)
                if got > 0:
                    assert got == ps and bytes(buf) == payload(f, i), ("torn page", f, i)
                else:
                    m.put(pid, payload(f, i))  # may lose a race for the slot: false is allowed
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(6)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]
    assert m.bytes() <= 64 * ps and m.pages() == m.bytes() // ps
    st = m.store_stats()
    assert st["crc_mismatches"] == 0
    assert st["pages"] == m.pages(), "meta store and GPU table disagree on residency"
    m.close()
