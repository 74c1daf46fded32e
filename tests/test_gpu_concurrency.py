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


def test_two_stores_on_one_device_put_concurrently():
    """Two stores on the same GPU, each fed 256 x 1 MiB puts from its own thread on its own stream, repeatedly.
    The put kernel's page tasks hand the slab slot over to each other; that hand-over must never depend on a CTA
    that is not running yet, or two such kernels sharing the SMs stall each other for good.  Checked: every call
    returns (the test runs under a watchdog), every page reads back bit-exact with the CRC the oracle computes."""
    import ctypes as C
    import torch
    from alluxio_b200 import capi
    from alluxio_b200.page_store import GpuPageStore
    from gpu_util import GET_REQ, PUT_REQ, RESULT, fill_pages, from_dev, to_dev
    from oracle import binding as orc
    ps, n, rounds = 1 << 20, 256, 6
    L = capi.lib()
    stores = [GpuPageStore(page_size=ps, cache_size=(n + 8) * ps, slab_slots=n + 8, max_batch=n) for _ in range(2)]
    srcs = [torch.empty(n * ps, dtype=torch.uint8, device="cuda:0") for _ in range(2)]
    for t in range(2):
        fill_pages(srcs[t], ps, ps, [t] * n, list(range(n)), 77)
    streams = [torch.cuda.Stream() for _ in range(2)]
    res = [torch.zeros(n * 8, dtype=torch.uint8, device="cuda:0") for _ in range(2)]
    reqs = []
    for t in range(2):
        r = np.zeros(n, PUT_REQ)
        r["file_lo"], r["page_index"], r["len"] = t, np.arange(n), ps
        r["src"] = srcs[t].data_ptr() + np.arange(n, dtype=np.uint64) * ps
        reqs.append(to_dev(r))
    torch.cuda.synchronize()
    errors, start = [], threading.Barrier(2)

    def worker(t):
        try:
            sp = C.c_void_p(streams[t].cuda_stream)
            start.wait()
            for _ in range(rounds):
                capi.check(L.aps_put_batch_dev(stores[t].handle, n, reqs[t].data_ptr(), res[t].data_ptr(), sp), "put")
            streams[t].synchronize()
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(t,), daemon=True) for t in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=60)
    assert not any(th.is_alive() for th in threads), "concurrent fused puts of two stores did not finish"
    assert not errors, errors
    dst = torch.empty(n * ps, dtype=torch.uint8, device="cuda:0")
    for t in range(2):
        r = from_dev(res[t], RESULT)
        assert (r["len"] == ps).all()
        g = np.zeros(n, GET_REQ)
        g["file_lo"], g["page_index"], g["nbytes"] = t, np.arange(n), ps
        g["dst"] = dst.data_ptr() + np.arange(n, dtype=np.uint64) * ps
        d_g, d_r = to_dev(g), torch.zeros(n * 8, dtype=torch.uint8, device="cuda:0")
        capi.check(L.aps_get_batch_dev(stores[t].handle, n, d_g.data_ptr(), d_r.data_ptr(), None), "get")
        torch.cuda.synchronize()
        gr = from_dev(d_r, RESULT)
        assert (gr["len"] == ps).all() and torch.equal(dst, srcs[t])
        for i in (0, 100, 255):
            assert int(gr[i]["crc"]) == int(r[i]["crc"]) == orc.crc32c(orc.fill_page(77, t, i, ps))
        st = stores[t].stats()
        assert st["pages"] == n and st["crc_mismatches"] == 0
    for s in stores:
        s.close()


def test_timeline_diagnostic_on_and_off_leaves_the_store_intact():
    """aps_debug_timeline(enable = 0) once freed the copy kernel's rank counters and the shard table as well.  Switch
    the diagnostic on, read it, switch it off, and the store -- also one with shards attached -- must still serve
    batched gets and puts and close cleanly."""
    import ctypes as C
    import torch
    from alluxio_b200 import capi
    from alluxio_b200.page_store import GpuPageStore
    from gpu_util import GET_REQ, RESULT, from_dev, to_dev
    ps, n = 65536, 64
    L = capi.lib()
    a = GpuPageStore(page_size=ps, cache_size=256 * ps, slab_slots=256, max_batch=n)
    b = GpuPageStore(page_size=ps, cache_size=256 * ps, slab_slots=256, max_batch=n)
    hs = (C.c_void_p * 2)(a.handle, b.handle)
    capi.check(L.aps_shard_attach_local(a.handle, 2, 0, hs), "attach")
    rng = np.random.default_rng(9)
    pages = {i: rng.integers(0, 256, ps, dtype=np.uint8).tobytes() for i in range(n)}
    owner = {}
    for i, d in pages.items():
        k = a.key(PID("5", i))
        owner[i] = L.aps_shard_of(C.byref(k), 2)
        (a if owner[i] == 0 else b).put(PID("5", i), d)
    dst = torch.zeros(n * ps, dtype=torch.uint8, device="cuda:0")
    g = np.zeros(n, GET_REQ)
    g["file_lo"], g["page_index"], g["nbytes"] = 5, np.arange(n), ps
    g["dst"] = dst.data_ptr() + np.arange(n, dtype=np.uint64) * ps
    d_g, d_r = to_dev(g), torch.zeros(n * 8, dtype=torch.uint8, device="cuda:0")

    def check_sharded_get():
        dst.zero_()
        capi.check(L.aps_get_batch_sharded_dev(a.handle, n, d_g.data_ptr(), d_r.data_ptr(), None), "sharded get")
        torch.cuda.synchronize()
        assert (from_dev(d_r, RESULT)["len"] == ps).all()
        host = dst.cpu().numpy()
        for i in (0, 31, 63):
            assert host[i * ps:(i + 1) * ps].tobytes() == pages[i]

    check_sharded_get()
    raw = np.zeros(1024 + 256 * 32 + 256, np.uint64)
    capi.check(L.aps_debug_timeline(a.handle, 1, None, 0))
    check_sharded_get()
    capi.check(L.aps_debug_timeline(a.handle, 1, raw.ctypes.data, raw.size))
    assert (raw[:1024].reshape(256, 4)[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF)).any()
    capi.check(L.aps_debug_timeline(a.handle, 0, None, 0))
    for _ in range(3):
        check_sharded_get()
    a.put(PID("6", 1), pages[0])
    buf = bytearray(ps)
    assert a.get(PID("6", 1), buf) == ps and bytes(buf) == pages[0]
    h = a.handle
    a.close()
    b.close()
    # AutoCloseable.close is idempotent; any other call on a closed store answers APS_ECLOSED
    assert L.aps_close(h) == capi.APS_OK
    st = capi.aps_stats_t()
    assert L.aps_stats(h, C.byref(st)) == capi.APS_ECLOSED
