"""-m gpu: APS_FLAG_STABLE_REQS.  With the flag a get that follows another get on the same stream reads its
requests, the table and the slab before it waits for its predecessor, and stores only afterwards.  Results must
be those of a store without the flag in every interleaving: get after get into the SAME destination (the second
one's bytes must win), get after put, after delete, after an overwrite, ragged and missing pages."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as gu
from alluxio_b200 import capi
from oracle import binding as orc

pytestmark = pytest.mark.gpu

PAGE = 64 * 1024
N = 96


def _open(flags):
    cfg = capi.aps_config(device=0, page_size=PAGE, cache_bytes=4 * N * PAGE, slab_slots=4 * N, table_capacity=0,
                          max_batch=4 * N, flags=flags)
    h = C.c_void_p()
    capi.check(capi.lib().aps_open(C.byref(cfg), C.byref(h)), "open")
    return h


def _put(h, files, idx, lens, seed, stage):
    L = capi.lib()
    gu.fill_pages(stage, PAGE, PAGE, files, idx, seed)
    r = np.zeros(len(files), gu.PUT_REQ)
    r["file_lo"], r["page_index"], r["len"] = files, idx, lens
    r["src"] = stage.data_ptr() + np.arange(len(files), dtype=np.uint64) * PAGE
    d = gu.to_dev(r)
    res = torch.zeros(len(files) * 8, dtype=torch.uint8, device="cuda:0")
    capi.check(L.aps_put_batch_dev(h, len(files), d.data_ptr(), res.data_ptr(), gu.cur_stream()), "put")
    return d, res


def _get_reqs(files, idx, dst, nbytes=PAGE):
    r = np.zeros(len(files), gu.GET_REQ)
    r["file_lo"], r["page_index"], r["nbytes"] = files, idx, nbytes
    r["dst"] = dst.data_ptr() + np.arange(len(files), dtype=np.uint64) * PAGE
    return gu.to_dev(r)


@pytest.mark.parametrize("crc", [True, False])
def test_interleavings_match_a_store_without_the_flag(crc, monkeypatch):
    L = capi.lib()
    base = capi.APS_FLAG_CRC if crc else 0
    outs = []
    # three stores: every kernel waits first (APS_NO_EARLY=1); early with re-validated requests (the default); early
    # with trusted requests (the flag)
    for flags, no_early in ((base, True), (base, False), (base | capi.APS_FLAG_STABLE_REQS, False)):
        if no_early:
            monkeypatch.setenv("APS_NO_EARLY", "1")
        else:
            monkeypatch.delenv("APS_NO_EARLY", raising=False)
        h = _open(flags)
        stage = torch.empty(N * PAGE, dtype=torch.uint8, device="cuda:0")
        dst = torch.zeros(N * PAGE, dtype=torch.uint8, device="cuda:0")
        files = np.arange(N, dtype=np.uint64) % 5
        idx = np.arange(N, dtype=np.uint64)
        lens = np.where(idx % 7 == 3, 1000 + idx, PAGE).astype(np.uint32)       # some ragged pages
        keep = [_put(h, files, idx, lens, 31, stage)]
        keep.append(_put(h, files + 100, idx, lens, 32, stage))                 # a second set, other bytes
        ra = _get_reqs(files, idx, dst)
        rb = _get_reqs(files + 100, idx, dst)
        rmiss = _get_reqs(files + 999, idx, dst)
        res = [torch.zeros(N * 8, dtype=torch.uint8, device="cuda:0") for _ in range(6)]
        sp = gu.cur_stream()
        snap = []
        torch.cuda.synchronize()
        # get A, get B into the same destination, get A again: no synchronisation in between
        capi.check(L.aps_get_batch_dev(h, N, ra.data_ptr(), res[0].data_ptr(), sp))
        capi.check(L.aps_get_batch_dev(h, N, rb.data_ptr(), res[1].data_ptr(), sp))
        snap.append(dst.clone())                                                # stream-ordered: B's bytes
        capi.check(L.aps_get_batch_dev(h, N, ra.data_ptr(), res[2].data_ptr(), sp))
        capi.check(L.aps_get_batch_dev(h, N, rmiss.data_ptr(), res[3].data_ptr(), sp))
        snap.append(dst.clone())                                                # A's bytes; the misses wrote nothing
        # overwrite set A with new bytes, then get, get
        keep.append(_put(h, files, idx, lens, 33, stage))
        capi.check(L.aps_get_batch_dev(h, N, ra.data_ptr(), res[4].data_ptr(), sp))
        snap.append(dst.clone())
        # delete half of A, then get twice
        dk = np.zeros(N // 2, gu.KEY)
        dk["file_lo"], dk["page_index"] = files[::2], idx[::2]
        d_dk = gu.to_dev(dk)
        dres = torch.zeros(N * 8, dtype=torch.uint8, device="cuda:0")
        capi.check(L.aps_delete_batch_dev(h, N // 2, d_dk.data_ptr(), dres.data_ptr(), sp))
        capi.check(L.aps_get_batch_dev(h, N, rb.data_ptr(), res[1].data_ptr(), sp))
        capi.check(L.aps_get_batch_dev(h, N, ra.data_ptr(), res[5].data_ptr(), sp))
        snap.append(dst.clone())
        torch.cuda.synchronize()
        outs.append(([gu.from_dev(r, gu.RESULT).copy() for r in res], [s.cpu().numpy() for s in snap]))
        st = capi.aps_stats_t()
        capi.check(L.aps_stats(h, C.byref(st)))
        assert st.crc_mismatches == 0
        capi.check(L.aps_close(h))
    (res0, snap0), (res1, snap1) = outs[0], outs[2]
    for other_res, other_snap in outs[1:]:
        for a, b in zip(res0, other_res):
            assert np.array_equal(a["len"], b["len"]) and np.array_equal(a["crc"], b["crc"])
        for a, b in zip(snap0, other_snap):
            assert np.array_equal(a, b)
    # and against the oracle: after "get A, get B" the destination holds B's pages
    lens = np.where(np.arange(N) % 7 == 3, 1000 + np.arange(N), PAGE)
    for i in (0, 3, 10, N - 1):
        want = orc.fill_page(32, 100 + i % 5, i, PAGE)[:lens[i]]
        assert np.array_equal(snap1[0][i * PAGE:i * PAGE + lens[i]], want)
        assert res1[1]["len"][i] == lens[i]
        if crc:
            assert res1[1]["crc"][i] == orc.crc32c(want)
    assert (res1[3]["len"] == -capi.APS_ENOTFOUND).all()
    assert (res1[5]["len"][::2] == -capi.APS_ENOTFOUND).all() and (res1[5]["len"][1::2] > 0).all()


def test_many_back_to_back_gets_large_pages():
    """The bench's shape in small: 1 MiB pages, 40 gets in a row from a pool of request batches, one destination."""
    L = capi.lib()
    page, n_res, batch = 1 << 20, 512, 64
    cfg = capi.aps_config(device=0, page_size=page, cache_bytes=n_res * page, slab_slots=n_res + 8, table_capacity=0,
                          max_batch=n_res, flags=capi.APS_FLAG_CRC | capi.APS_FLAG_STABLE_REQS)
    h = C.c_void_p()
    capi.check(L.aps_open(C.byref(cfg), C.byref(h)), "open")
    stage = torch.empty(n_res * page, dtype=torch.uint8, device="cuda:0")
    ids = np.arange(n_res, dtype=np.uint64)
    gu.fill_pages(stage, page, page, ids // 64, ids % 64, 77)
    r = np.zeros(n_res, gu.PUT_REQ)
    r["file_lo"], r["page_index"], r["len"] = ids // 64, ids % 64, page
    r["src"] = stage.data_ptr() + ids * page
    d = gu.to_dev(r)
    pres = torch.zeros(n_res * 8, dtype=torch.uint8, device="cuda:0")
    capi.check(L.aps_put_batch_dev(h, n_res, d.data_ptr(), pres.data_ptr(), gu.cur_stream()))
    rng = np.random.default_rng(4)
    dst = torch.zeros(batch * page, dtype=torch.uint8, device="cuda:0")
    picks = rng.integers(0, n_res, size=(40, batch)).astype(np.uint64)
    rq = np.zeros((40, batch), gu.GET_REQ)
    rq["file_lo"], rq["page_index"], rq["nbytes"] = picks // 64, picks % 64, page
    rq["dst"] = dst.data_ptr() + np.arange(batch, dtype=np.uint64)[None, :] * page
    d_rq = gu.to_dev(rq)
    res = torch.zeros(40 * batch * 8, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    for k in range(40):
        capi.check(L.aps_get_batch_dev(h, batch, d_rq.data_ptr() + k * batch * 40, res.data_ptr() + k * batch * 8,
                                       gu.cur_stream()))
    torch.cuda.synchronize()
    out = gu.from_dev(res, gu.RESULT).reshape(40, batch)
    want_crc = gu.from_dev(pres, gu.RESULT)["crc"]
    assert (out["len"] == page).all()
    assert np.array_equal(out["crc"], want_crc[picks.astype(np.int64)])
    got = dst.cpu().numpy().reshape(batch, page)
    src = stage.cpu().numpy().reshape(n_res, page)
    assert np.array_equal(got, src[picks[39].astype(np.int64)])                 # the last get's bytes
    st = capi.aps_stats_t()
    capi.check(L.aps_stats(h, C.byref(st)))
    assert st.crc_mismatches == 0
    capi.check(L.aps_close(h))


def test_requests_rewritten_on_the_stream_between_gets_are_honoured():
    """Without APS_FLAG_STABLE_REQS a get may read its requests early, but must not trust them: here the request
    array is overwritten by a kernel on the same stream right before every get (the early read can only see the
    previous batch), and every get must serve the batch that is current when its turn comes."""
    L = capi.lib()
    h = _open(capi.APS_FLAG_CRC)
    stage = torch.empty(N * PAGE, dtype=torch.uint8, device="cuda:0")
    files = np.arange(N, dtype=np.uint64) % 5
    idx = np.arange(N, dtype=np.uint64)
    lens = np.full(N, PAGE, np.uint32)
    keep = [_put(h, files + 10 * s_, idx, lens, 40 + s_, stage) for s_ in range(4)]
    dsts = [torch.zeros(N * PAGE, dtype=torch.uint8, device="cuda:0") for _ in range(4)]
    batches = [_get_reqs(files + 10 * s_, idx, dsts[s_]) for s_ in range(4)]
    live = torch.zeros_like(batches[0])
    res = [torch.zeros(N * 8, dtype=torch.uint8, device="cuda:0") for _ in range(12)]
    sp = gu.cur_stream()
    torch.cuda.synchronize()
    for k in range(12):
        live.copy_(batches[k % 4])                          # a kernel on the same stream rewrites the requests ...
        capi.check(L.aps_get_batch_dev(h, N, live.data_ptr(), res[k].data_ptr(), sp))   # ... that this get must use
    torch.cuda.synchronize()
    put_crc = [gu.from_dev(keep[s_][1], gu.RESULT)["crc"].copy() for s_ in range(4)]
    for k in range(12):
        r = gu.from_dev(res[k], gu.RESULT)
        assert (r["len"] == PAGE).all() and np.array_equal(r["crc"], put_crc[k % 4]), k
    for s_ in range(4):
        got = dsts[s_].cpu().numpy()
        for i in (0, N - 1):
            assert np.array_equal(got[i * PAGE:(i + 1) * PAGE], orc.fill_page(40 + s_, int(files[i]) + 10 * s_, i, PAGE))
    st = capi.aps_stats_t()
    capi.check(L.aps_stats(h, C.byref(st)))
    assert st.crc_mismatches == 0
    capi.check(L.aps_close(h))
