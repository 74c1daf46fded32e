"""-m gpu: the HBM hash table (K1) under churn, against a Python dict."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_insert_erase_lookup_churn_matches_dict():
    import torch
    from alluxio_b200 import capi
    from alluxio_b200.page_store import GpuPageStore, PageId
    from gpu_util import KEY, PUT_REQ, RESULT, cur_stream, from_dev, to_dev
    ps = 16
    slots = 6000
    # a deliberately tight table (load factor up to ~0.73) so probe chains and tombstones happen
    store = GpuPageStore(page_size=ps, cache_size=slots * ps, slab_slots=slots, table_capacity=8192, max_batch=4096,
                         crc=False)
    rng = np.random.default_rng(11)
    src = torch.zeros(4096 * ps, dtype=torch.uint8, device="cuda:0")
    model = {}
    universe = [(int(f), int(i)) for f in range(40) for i in range(400)]

    def keys_arr(ids, dtype):
        a = np.zeros(len(ids), dtype)
        for k, (f, i) in enumerate(ids):
            kk = store.key(PageId(f, i))
            a[k]["file_hi"], a[k]["file_lo"], a[k]["page_index"] = kk.file_hi, kk.file_lo, kk.page_index
        return a

    for rnd in range(12):
        # insert a batch of distinct ids (some already present -> replace)
        pick = rng.choice(len(universe), size=2500, replace=False)
        ids = [universe[p] for p in pick]
        room = slots - len(model)
        fresh = [x for x in ids if x not in model][:max(0, room)]
        ids = fresh + [x for x in ids if x in model]
        reqs = keys_arr(ids, PUT_REQ)
        lens = rng.integers(0, ps + 1, len(ids))
        reqs["len"] = lens
        reqs["src"] = src.data_ptr()
        d_res = torch.zeros(len(ids) * 8, dtype=torch.uint8, device="cuda:0")
        d_reqs = to_dev(reqs)
        capi.check(capi.lib().aps_put_batch_dev(store.handle, len(ids), d_reqs.data_ptr(), d_res.data_ptr(), cur_stream()))
        torch.cuda.synchronize()
        res = from_dev(d_res, RESULT)
        assert (res["len"] == lens).all()
        for x, l in zip(ids, lens):
            model[x] = int(l)
        # erase a random subset + some absent ids
        present = list(model)
        kill = [present[p] for p in rng.choice(len(present), size=len(present) // 2, replace=False)]
        absent = [x for x in universe[:200] if x not in model][:50]
        d_keys = to_dev(keys_arr(kill + absent, KEY))
        d_res = torch.zeros((len(kill) + len(absent)) * 8, dtype=torch.uint8, device="cuda:0")
        capi.check(capi.lib().aps_delete_batch_dev(store.handle, len(kill) + len(absent), d_keys.data_ptr(),
                                                   d_res.data_ptr(), cur_stream()))
        torch.cuda.synchronize()
        res = from_dev(d_res, RESULT)
        assert (res["len"][:len(kill)] == 0).all() and (res["len"][len(kill):] == -capi.APS_ENOTFOUND).all()
        for x in kill:
            del model[x]
        # look everything up
        probe = universe[::3]
        d_keys = to_dev(keys_arr(probe, KEY))
        d_len = torch.zeros(len(probe), dtype=torch.int32, device="cuda:0")
        for lo in range(0, len(probe), 4096):
            m = min(4096, len(probe) - lo)
            capi.check(capi.lib().aps_lookup_batch_dev(store.handle, m, d_keys.data_ptr() + lo * 24,
                                                       d_len.data_ptr() + lo * 4, cur_stream()))
        torch.cuda.synchronize()
        assert d_len.cpu().tolist() == [model.get(x, -1) for x in probe]
        st = store.stats()
        assert st["pages"] == len(model) and st["bytes"] == sum(model.values())
        assert st["slots_free"] == slots - len(model)
    got = sorted((int(p.file_id), p.page_index, n) for p, n in store.get_pages())
    assert got == sorted((f, i, n) for (f, i), n in model.items())
    st = store.stats()
    assert st["probe_buckets"] >= st["lookups"] > 0
    store.close()


def test_table_full_reports_no_space():
    from alluxio_b200.page_store import GpuPageStore, IOExceptionJ, PageId
    store = GpuPageStore(page_size=16, cache_size=1 << 20, slab_slots=4096, table_capacity=8, crc=False)
    ok = 0
    with pytest.raises(IOExceptionJ):
        for i in range(64):
            store.put(PageId("3", i), b"x")
            ok += 1
    assert ok == 8
    for i in range(8):
        b = bytearray(1)
        assert store.get(PageId("3", i), b) == 1
    store.close()


def test_compact_removes_tombstones_and_keeps_every_page():
    import ctypes as C
    from alluxio_b200 import capi
    from alluxio_b200.page_store import GpuPageStore, PageId
    # a table of 4 buckets for 32 slots: buckets fill up completely, so deletes leave tombstones
    store = GpuPageStore(page_size=16, cache_size=1 << 20, slab_slots=32, table_capacity=32, crc=True)
    for i in range(32):
        store.put(PageId("8", i), bytes([i]) * 16)
    for i in range(0, 32, 2):
        store.delete(PageId("8", i))
    st = store.stats()
    assert st["pages"] == 16 and st["tombstones"] > 0
    capi.check(capi.lib().aps_compact(store.handle))
    st = store.stats()
    assert st["pages"] == 16 and st["tombstones"] == 0
    for i in range(32):
        buf = bytearray(16)
        if i % 2:
            assert store.get(PageId("8", i), buf) == 16 and bytes(buf) == bytes([i]) * 16
        else:
            with pytest.raises(Exception):
                store.get(PageId("8", i), buf)
    for i in range(0, 32, 2):           # the freed slots and entries are usable again
        store.put(PageId("9", i), bytes([i + 100]) * 16)
    assert store.stats()["pages"] == 32
    # host-array batches larger than max_batch are served in pieces
    import torch
    from gpu_util import GET_REQ, RESULT
    s2 = GpuPageStore(page_size=16, cache_size=1 << 20, slab_slots=64, max_batch=8, crc=True)
    for i in range(40):
        s2.put(PageId("1", i), bytes([i]) * 16)
    pinned = torch.zeros(40 * 16, dtype=torch.uint8).pin_memory()
    res = s2.get_batch([(PageId("1", i), 0, 16, pinned.data_ptr() + i * 16) for i in range(40)])
    assert [r[0] for r in res] == [16] * 40
    assert pinned.numpy().reshape(40, 16)[:, 0].tolist() == list(range(40))
    s2.close()
    store.close()
