"""-m gpu: the batched get/put/delete path (device-resident and host-array forms) against the
oracle: returned bytes, lengths, statuses and CRC32C values, bit for bit."""
import ctypes as C

import numpy as np
import pytest

from oracle import binding as orc

pytestmark = pytest.mark.gpu

SEED = 0xA11C10
MB = 1 << 20


def torch_mod():
    import torch
    return torch


def make_store(**kw):
    from alluxio_b200.page_store import GpuPageStore
    return GpuPageStore(**kw)


def key_fields(store, pid):
    k = store.key(pid)
    return k.file_hi, k.file_lo, k.page_index


def put_dev(store, pids, lens, src, stride):
    """aps_put_batch_dev with sources at src + i*stride."""
    from gpu_util import PUT_REQ, RESULT, cur_stream, from_dev, to_dev
    from alluxio_b200 import capi
    torch = torch_mod()
    reqs = np.zeros(len(pids), PUT_REQ)
    for i, pid in enumerate(pids):
        reqs[i]["file_hi"], reqs[i]["file_lo"], reqs[i]["page_index"] = key_fields(store, pid)
        reqs[i]["len"] = lens[i]
        reqs[i]["src"] = src.data_ptr() + i * stride
    d_reqs = to_dev(reqs)
    d_res = torch.zeros(len(pids) * 8, dtype=torch.uint8, device="cuda:0")
    capi.check(capi.lib().aps_put_batch_dev(store.handle, len(pids), d_reqs.data_ptr(), d_res.data_ptr(), cur_stream()))
    torch.cuda.synchronize()
    return from_dev(d_res, RESULT)


def get_dev(store, pids, offs, nbytes, dst_ptrs):
    from gpu_util import GET_REQ, RESULT, cur_stream, from_dev, to_dev
    from alluxio_b200 import capi
    torch = torch_mod()
    reqs = np.zeros(len(pids), GET_REQ)
    for i, pid in enumerate(pids):
        reqs[i]["file_hi"], reqs[i]["file_lo"], reqs[i]["page_index"] = key_fields(store, pid)
        reqs[i]["page_off"], reqs[i]["nbytes"], reqs[i]["dst"] = offs[i], nbytes[i], dst_ptrs[i]
    d_reqs = to_dev(reqs)
    d_res = torch.zeros(len(pids) * 8, dtype=torch.uint8, device="cuda:0")
    capi.check(capi.lib().aps_get_batch_dev(store.handle, len(pids), d_reqs.data_ptr(), d_res.data_ptr(), cur_stream()))
    torch.cuda.synchronize()
    return from_dev(d_res, RESULT)


def test_fill_pages_matches_oracle_generator():
    from gpu_util import fill_pages
    torch = torch_mod()
    for page_len in (8, 13, 4096, 65536 + 3):
        stride = (page_len + 15) // 16 * 16
        files, idx = [0, 3, 63, 2**31 + 5], [0, 5, 127, 2**40]
        buf = torch.zeros(len(files) * stride, dtype=torch.uint8, device="cuda:0")
        fill_pages(buf, stride, page_len, files, idx, SEED)
        host = buf.cpu().numpy()
        for p in range(len(files)):
            want = orc.fill_page(SEED, files[p], idx[p], page_len)
            assert np.array_equal(host[p * stride:p * stride + page_len], want), (page_len, p)


def test_full_pages_1mib_put_get_bytes_and_crc():
    """BASELINE config 2 in miniature: 1 MiB pages, batched put from HBM, batched get into HBM."""
    from alluxio_b200.page_store import PageId
    from gpu_util import fill_pages
    torch = torch_mod()
    n = 24
    store = make_store(page_size=MB, cache_size=32 * MB)
    files = [i % 3 for i in range(n)]
    idx = [i // 3 for i in range(n)]
    pids = [PageId(files[i], idx[i]) for i in range(n)]
    src = torch.empty(n * MB, dtype=torch.uint8, device="cuda:0")
    fill_pages(src, MB, MB, files, idx, SEED)
    res = put_dev(store, pids, [MB] * n, src, MB)
    want_crc = [orc.crc32c(orc.fill_page(SEED, files[i], idx[i], MB)) for i in range(n)]
    assert list(res["len"]) == [MB] * n
    assert list(res["crc"]) == want_crc
    order = np.random.default_rng(1).permutation(n)
    dst = torch.zeros(n * MB, dtype=torch.uint8, device="cuda:0")
    res = get_dev(store, [pids[j] for j in order], [0] * n, [MB] * n, [dst.data_ptr() + k * MB for k in range(n)])
    assert list(res["len"]) == [MB] * n
    assert list(res["crc"]) == [want_crc[j] for j in order]
    host = dst.cpu().numpy()
    for k, j in enumerate(order):
        assert np.array_equal(host[k * MB:(k + 1) * MB], orc.fill_page(SEED, files[j], idx[j], MB))
    st = store.stats()
    assert st["pages"] == n and st["bytes"] == n * MB and st["crc_mismatches"] == 0
    store.close()


@pytest.mark.parametrize("ps", [65536, 4112, 16000, 48])
def test_ragged_lengths_offsets_misses_and_unaligned_destinations(ps):
    from alluxio_b200 import capi
    from alluxio_b200.page_store import PageId
    torch = torch_mod()
    store = make_store(page_size=ps, cache_size=64 * ps, slab_slots=128)
    rng = np.random.default_rng(2)
    lens = [n for n in (0, 1, 15, 16, 17, 511, 512, 513, 16383, 16384, 16385, 32768, ps - 1, ps, 40000, 4097) if n <= ps]
    n = len(lens)
    pids = [PageId("77", i) for i in range(n)]
    payload = [rng.integers(0, 256, l, dtype=np.uint8) for l in lens]
    src = torch.zeros(n * ps, dtype=torch.uint8, device="cuda:0")
    for i, p in enumerate(payload):
        src[i * ps:i * ps + len(p)] = torch.from_numpy(p).cuda()
    res = put_dev(store, pids, lens, src, ps)
    assert list(res["len"]) == lens
    assert list(res["crc"]) == [orc.crc32c(p) for p in payload]

    # requests: whole pages, offsets, short reads, offset == len, offset > len, missing ids, odd dst alignment
    reqs = []
    for i in range(n):
        reqs.append((i, 0, ps))                      # whole page (clamped to its length)
        reqs.append((i, lens[i], 10))                # offset == length -> 0 bytes
        reqs.append((i, lens[i] + 1, 10))            # offset beyond -> EINVAL
        if lens[i] > 20:
            reqs.append((i, 7, lens[i] - 7))         # unaligned source
            reqs.append((i, 16, 5))                  # tiny
            reqs.append((i, int(rng.integers(0, lens[i])), int(rng.integers(0, ps))))
    reqs.append((None, 0, 100))                      # miss
    dst = torch.zeros(len(reqs) * (ps + 16), dtype=torch.uint8, device="cuda:0")
    pids_r, offs, nb, ptrs = [], [], [], []
    for k, (i, off, want) in enumerate(reqs):
        pids_r.append(PageId("77", 10**6) if i is None else pids[i])
        offs.append(off)
        nb.append(want)
        ptrs.append(dst.data_ptr() + k * (ps + 16) + (k % 5))  # k%5: misaligned destinations too
    res = get_dev(store, pids_r, offs, nb, ptrs)
    host = dst.cpu().numpy()
    for k, (i, off, want) in enumerate(reqs):
        if i is None:
            assert res[k]["len"] == -capi.APS_ENOTFOUND
            continue
        if off > lens[i]:
            assert res[k]["len"] == -capi.APS_EINVAL
            continue
        exp = payload[i][off:off + min(want, lens[i] - off)]
        assert res[k]["len"] == len(exp), (k, i, off, want)
        base = k * (ps + 16) + (k % 5)
        assert np.array_equal(host[base:base + len(exp)], exp), (k, i, off, want)
        assert res[k]["crc"] == orc.crc32c(exp), (k, i, off, want)
        assert not host[base + len(exp):base + len(exp) + 8].any()  # nothing written past the end
    store.close()


def test_host_array_batch_forms_with_pinned_memory():
    from alluxio_b200 import capi
    from alluxio_b200.page_store import PageId
    torch = torch_mod()
    ps = 1 << 16
    store = make_store(page_size=ps, cache_size=64 * ps)
    n = 20
    rng = np.random.default_rng(3)
    pinned_src = torch.from_numpy(rng.integers(0, 256, n * ps, dtype=np.uint8)).pin_memory()
    pids = [PageId("abc/def", i) for i in range(n)]
    lens = [ps if i % 4 else ps - 100 - i for i in range(n)]
    from gpu_util import cur_stream
    torch.cuda.synchronize()
    res = store.put_batch([(pids[i], lens[i], pinned_src.data_ptr() + i * ps) for i in range(n)], cur_stream())
    src_np = pinned_src.numpy()
    assert [r[0] for r in res] == lens
    assert [r[1] for r in res] == [orc.crc32c(src_np[i * ps:i * ps + lens[i]]) for i in range(n)]
    # duplicate ids in one put batch are refused
    with pytest.raises(capi.ApsError):
        store.put_batch([(pids[0], 1, pinned_src.data_ptr()), (pids[0], 1, pinned_src.data_ptr())])
    pinned_dst = torch.zeros(n * ps, dtype=torch.uint8).pin_memory()
    res = store.get_batch([(pids[i], 0, ps, pinned_dst.data_ptr() + i * ps) for i in range(n)], cur_stream())
    dst_np = pinned_dst.numpy()
    for i in range(n):
        assert res[i][0] == lens[i]
        assert np.array_equal(dst_np[i * ps:i * ps + lens[i]], src_np[i * ps:i * ps + lens[i]])
        assert res[i][1] == orc.crc32c(src_np[i * ps:i * ps + lens[i]])
    # gets may repeat an id
    res = store.get_batch([(pids[1], 0, 16, pinned_dst.data_ptr()), (pids[1], 16, 16, pinned_dst.data_ptr() + 16)])
    assert [r[0] for r in res] == [16, 16]
    assert np.array_equal(dst_np[:32], src_np[ps:ps + 32])
    # delete half, then look all of them up
    gone = store.delete_batch(pids[::2] + [PageId("abc/def", 999)])
    assert gone == [0] * len(pids[::2]) + [-capi.APS_ENOTFOUND]
    res = store.get_batch([(pids[i], 0, ps, pinned_dst.data_ptr() + i * ps) for i in range(n)])
    assert [r[0] for r in res] == [(-capi.APS_ENOTFOUND if i % 2 == 0 else lens[i]) for i in range(n)]
    store.close()


def test_lookup_and_standalone_crc_device_forms():
    from alluxio_b200 import capi
    from alluxio_b200.page_store import PageId
    from gpu_util import KEY, cur_stream, to_dev
    torch = torch_mod()
    ps = 4096
    store = make_store(page_size=ps, cache_size=256 * ps)
    for i in range(50):
        store.put(PageId("5", i), bytes([i]) * (i * 10))
    keys = np.zeros(80, KEY)
    for i in range(80):
        keys[i]["file_hi"], keys[i]["file_lo"], keys[i]["page_index"] = key_fields(store, PageId("5", i))
    d_keys = to_dev(keys)
    d_len = torch.zeros(80, dtype=torch.int32, device="cuda:0")
    capi.check(capi.lib().aps_lookup_batch_dev(store.handle, 80, d_keys.data_ptr(), d_len.data_ptr(), cur_stream()))
    torch.cuda.synchronize()
    assert d_len.cpu().tolist() == [i * 10 for i in range(50)] + [-1] * 30
    # stand-alone CRC over arbitrary device ranges
    rng = np.random.default_rng(4)
    blob = rng.integers(0, 256, 20 * ps, dtype=np.uint8)
    d_blob = torch.from_numpy(blob).cuda()
    lens = [0, 1, 100, 512, 4095, 4096, 1000, 16, 17, 4096]
    starts = [0, 1, 16, 4096, 8192, 12288, 20001, 32, 48, 40960]
    ptrs = torch.tensor([d_blob.data_ptr() + s for s in starts], dtype=torch.int64, device="cuda:0")
    d_lens = torch.tensor(lens, dtype=torch.int32, device="cuda:0")
    d_crc = torch.zeros(len(lens), dtype=torch.int32, device="cuda:0")
    capi.check(capi.lib().aps_crc32c_batch_dev(store.handle, len(lens), ptrs.data_ptr(), d_lens.data_ptr(),
                                               d_crc.data_ptr(), cur_stream()))
    torch.cuda.synchronize()
    got = d_crc.cpu().numpy().view(np.uint32).tolist()
    assert got == [orc.crc32c(blob[s:s + l]) for s, l in zip(starts, lens)]
    store.close()


def test_batch_of_256_full_pages_roundtrip_properties():
    """BASELINE config 2 batch shape (256 x 1 MiB) checked through size-independent properties:
    put -> get round trip is the identity, per-page CRCs equal the host-cores CRC of the same bytes,
    and the CRC of the concatenation folds out of the per-page CRCs (checksum of checksums)."""
    from alluxio_b200.page_store import PageId
    from gpu_util import fill_pages
    torch = torch_mod()
    n = 256
    store = make_store(page_size=MB, cache_size=n * MB, max_batch=256)
    files = [i % 8 for i in range(n)]
    idx = [i // 8 for i in range(n)]
    pids = [PageId(files[i], idx[i]) for i in range(n)]
    src = torch.empty(n * MB, dtype=torch.uint8, device="cuda:0")
    fill_pages(src, MB, MB, files, idx, SEED)
    res = put_dev(store, pids, [MB] * n, src, MB)
    assert (res["len"] == MB).all()
    dst = torch.zeros(n * MB, dtype=torch.uint8, device="cuda:0")
    res = get_dev(store, pids, [0] * n, [MB] * n, [dst.data_ptr() + k * MB for k in range(n)])
    assert (res["len"] == MB).all()
    assert torch.equal(src, dst)
    host = src.cpu().numpy()
    whole = 0
    for i in range(n):
        c = orc.crc32c(host[i * MB:(i + 1) * MB], "hw") or orc.crc32c(host[i * MB:(i + 1) * MB])
        assert res[i]["crc"] == c
        whole = orc.crc32c_combine(whole, int(res[i]["crc"]), MB)
    assert whole == (orc.crc32c(host, "hw") or orc.crc32c(host))
    store.close()


def test_crc_disabled_store_copies_identically():
    from alluxio_b200.page_store import PageId
    torch = torch_mod()
    ps = 1 << 16
    store = make_store(page_size=ps, cache_size=16 * ps, crc=False)
    rng = np.random.default_rng(6)
    data = rng.integers(0, 256, 8 * ps, dtype=np.uint8)
    src = torch.from_numpy(data).cuda()
    pids = [PageId("1", i) for i in range(8)]
    lens = [ps, ps - 1, 513, 0, 16384, 1, ps, 100]
    res = put_dev(store, pids, lens, src, ps)
    assert list(res["len"]) == lens
    dst = torch.zeros(8 * ps, dtype=torch.uint8, device="cuda:0")
    res = get_dev(store, pids, [0] * 8, [ps] * 8, [dst.data_ptr() + k * ps for k in range(8)])
    assert list(res["len"]) == lens and not res["crc"].any()
    host = dst.cpu().numpy()
    for i in range(8):
        assert np.array_equal(host[i * ps:i * ps + lens[i]], data[i * ps:i * ps + lens[i]])
    store.close()


def test_batched_get_into_pinned_host_memory_staged_and_direct(monkeypatch):
    """Large host-destination batches go through an HBM staging area and the copy engine; small ones, or with
    APS_NO_D2H_STAGE=1, are stored by the kernel directly.  Same bytes, same results, nothing written past len."""
    import ctypes as C
    import os
    import torch
    import gpu_util as gu
    from alluxio_b200 import capi
    from oracle import binding as orc
    L = capi.lib()
    page, n = 256 * 1024, 48                                    # 12 MiB per batch: above the staging threshold
    outs = []
    for direct in (False, True):
        if direct:
            monkeypatch.setenv("APS_NO_D2H_STAGE", "1")
        else:
            monkeypatch.delenv("APS_NO_D2H_STAGE", raising=False)
        cfg = capi.aps_config(device=0, page_size=page, cache_bytes=n * page, slab_slots=n + 8, table_capacity=0,
                              max_batch=n, flags=capi.APS_FLAG_CRC)
        h = C.c_void_p()
        capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "open")
        ids = np.arange(n, dtype=np.uint64)
        lens = np.where(ids % 5 == 2, 1000 + 7 * ids, page).astype(np.uint32)   # some short pages
        stage = torch.empty(n * page, dtype=torch.uint8, device="cuda:0")
        gu.fill_pages(stage, page, page, ids // 8, ids % 8, 55)
        pr = np.zeros(n, gu.PUT_REQ)
        pr["file_lo"], pr["page_index"], pr["len"] = ids // 8, ids % 8, lens
        pr["src"] = stage.data_ptr() + ids * page
        res = np.zeros(n, gu.RESULT)
        capi.check(L.aps_put_batch(h, n, pr.ctypes.data, res.ctypes.data, None), "put")
        host = torch.full((n * page + 4096,), 0xEE, dtype=torch.uint8).pin_memory()
        rq = np.zeros(n, gu.GET_REQ)
        rq["file_lo"], rq["page_index"], rq["nbytes"] = ids // 8, ids % 8, page
        rq["file_lo"][7] = 9999                                                  # one miss in the middle of a run
        order = np.arange(n)
        order[10], order[11] = 11, 10                                            # two destinations swapped: not adjacent
        rq["dst"] = host.data_ptr() + order.astype(np.uint64) * page
        capi.check(L.aps_get_batch(h, n, rq.ctypes.data, res.ctypes.data, None), "get")
        outs.append((res.copy(), host.numpy().copy()))
        for i in range(n):
            want_len = -capi.APS_ENOTFOUND if i == 7 else int(lens[i])
            assert res["len"][i] == want_len
            at = int(order[i]) * page
            got = host.numpy()[at:at + page]
            if i == 7:
                assert (got == 0xEE).all()
                continue
            want = orc.fill_page(55, int(ids[i]) // 8, int(ids[i]) % 8, page)[:lens[i]]
            assert np.array_equal(got[:lens[i]], want) and (got[lens[i]:] == 0xEE).all()
            assert res["crc"][i] == orc.crc32c(want)
        assert (host.numpy()[n * page:] == 0xEE).all()
        capi.check(L.aps_close(h))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
