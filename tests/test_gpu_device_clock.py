"""-m gpu: opt-in eviction on the device (APS_FLAG_DEVICE_CLOCK).  CLOCK is an approximation of
the reference's LRUCacheEvictor (CACHE/evictor/LRUCacheEvictor.java) and is never used for the parity runs; what
is checked here is what CLOCK promises -- referenced pages survive a sweep that can be satisfied from unreferenced
ones, counts and residency stay consistent, the table stays probe-able -- and that its hit ratio on the Zipfian
trace of BASELINE configs[2] stays close to the oracle's exact LRU."""
import ctypes as C

import numpy as np
import pytest
import torch

import gpu_util as gu
from alluxio_b200 import capi
from oracle import binding as orc

pytestmark = pytest.mark.gpu

PAGE = 4096


def _open(slots, flags=capi.APS_FLAG_CRC | capi.APS_FLAG_DEVICE_CLOCK, max_batch=4096):
    cfg = capi.aps_config(device=0, page_size=PAGE, cache_bytes=slots * PAGE, slab_slots=slots, table_capacity=0,
                          max_batch=max_batch, flags=flags)
    h = C.c_void_p()
    capi.check(capi.lib().aps_open(C.byref(cfg), C.byref(h)), "open")
    return h


class Dev:
    def __init__(self, h, max_batch=4096):
        self.h, self.L = h, capi.lib()
        self.stage = torch.empty(max_batch * PAGE, dtype=torch.uint8, device="cuda:0")
        self.sink = torch.empty(max_batch * PAGE, dtype=torch.uint8, device="cuda:0")
        self.res = torch.zeros(max_batch * 8, dtype=torch.uint8, device="cuda:0")
        self.lens = torch.zeros(max_batch, dtype=torch.int32, device="cuda:0")

    def put(self, ids, seed=9):
        ids = np.asarray(ids, dtype=np.uint64)
        gu.fill_pages(self.stage, PAGE, PAGE, ids // 1000, ids % 1000, seed)
        r = np.zeros(len(ids), gu.PUT_REQ)
        r["file_lo"], r["page_index"], r["len"] = ids // 1000, ids % 1000, PAGE
        r["src"] = self.stage.data_ptr() + np.arange(len(ids), dtype=np.uint64) * PAGE
        d = gu.to_dev(r)
        capi.check(self.L.aps_put_batch_dev(self.h, len(ids), d.data_ptr(), self.res.data_ptr(), gu.cur_stream()), "put")
        torch.cuda.synchronize()
        return gu.from_dev(self.res, gu.RESULT)[:len(ids)]["len"].copy()

    def get(self, ids):
        ids = np.asarray(ids, dtype=np.uint64)
        r = np.zeros(len(ids), gu.GET_REQ)
        r["file_lo"], r["page_index"], r["nbytes"] = ids // 1000, ids % 1000, PAGE
        r["dst"] = self.sink.data_ptr() + np.arange(len(ids), dtype=np.uint64) * PAGE
        d = gu.to_dev(r)
        capi.check(self.L.aps_get_batch_dev(self.h, len(ids), d.data_ptr(), self.res.data_ptr(), gu.cur_stream()), "get")
        torch.cuda.synchronize()
        return gu.from_dev(self.res, gu.RESULT)[:len(ids)].copy()

    def present(self, ids):
        ids = np.asarray(ids, dtype=np.uint64)
        k = np.zeros(len(ids), gu.KEY)
        k["file_lo"], k["page_index"] = ids // 1000, ids % 1000
        d = gu.to_dev(k)
        capi.check(self.L.aps_lookup_batch_dev(self.h, len(ids), d.data_ptr(), self.lens.data_ptr(), gu.cur_stream()))
        torch.cuda.synchronize()
        return self.lens[:len(ids)].cpu().numpy() >= 0

    def stats(self):
        st = capi.aps_stats_t()
        capi.check(self.L.aps_stats(self.h, C.byref(st)))
        return st


def test_referenced_pages_survive_and_counts_stay_consistent():
    L = capi.lib()
    d = Dev(_open(256))
    assert (d.put(np.arange(256)) == PAGE).all()
    victims = torch.zeros(16 * 24, dtype=torch.uint8, device="cuda:0")
    count = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    capi.check(L.aps_clock_evict_dev(d.h, 16, victims.data_ptr(), count.data_ptr(), gu.cur_stream()))
    torch.cuda.synchronize()
    assert int(count.item()) == 16
    v = gu.from_dev(victims, gu.KEY)
    vid = (v["file_lo"] * 1000 + v["page_index"]).astype(np.int64)
    assert len(set(vid.tolist())) == 16 and not d.present(vid).any()            # the reported victims are gone
    st = d.stats()
    assert (st.pages, st.slots_free, st.clock_evictions) == (240, 16, 16)
    survivors = np.setdiff1d(np.arange(256), vid)
    hot = survivors[:100]
    got = d.get(hot)                                                            # referenced again
    assert (got["len"] == PAGE).all()
    assert (d.put(np.arange(5000, 5100)) == PAGE).all()                         # 100 new pages, 16 free: 84 must go
    st = d.stats()
    assert (st.pages, st.slots_free, st.clock_evictions) == (256, 0, 100)
    assert d.present(hot).all()                                                 # CLOCK's promise
    assert d.present(np.arange(5000, 5100)).all()
    assert int(d.present(survivors[100:]).sum()) == 140 - 84
    # bytes and CRCs of what is resident are intact
    got = d.get(hot[:8])
    for i, pid in enumerate(hot[:8]):
        want = orc.fill_page(9, int(pid) // 1000, int(pid) % 1000, PAGE)
        assert got["crc"][i] == orc.crc32c(want)
        assert np.array_equal(d.sink[i * PAGE:(i + 1) * PAGE].cpu().numpy(), want)
    assert st.crc_mismatches == 0
    capi.check(L.aps_close(d.h))


def test_everything_referenced_still_makes_room_and_compact_keeps_ownership():
    L = capi.lib()
    d = Dev(_open(512))
    d.put(np.arange(512))
    d.get(np.arange(512))                                                       # every bit set: the hand needs a second turn
    assert (d.put(np.arange(9000, 9300)) == PAGE).all()
    st = d.stats()
    assert (st.pages, st.clock_evictions) == (512, 300)
    assert d.present(np.arange(9000, 9300)).all()
    capi.check(L.aps_compact(d.h))                                              # entries move; slot ownership must follow
    assert d.present(np.arange(9000, 9300)).all()
    assert (d.put(np.arange(20000, 20400)) == PAGE).all()
    st = d.stats()
    assert (st.pages, st.clock_evictions, st.tombstones) == (512, 700, st.tombstones)
    assert d.present(np.arange(20000, 20400)).all()
    assert int(d.present(np.concatenate([np.arange(512), np.arange(9000, 9300)])).sum()) == 112
    capi.check(L.aps_close(d.h))


def test_needs_the_flag():
    L = capi.lib()
    h = _open(64, flags=capi.APS_FLAG_CRC)
    assert L.aps_clock_evict_dev(h, 4, None, None, None) == capi.APS_EINVAL
    capi.check(L.aps_close(h))


def test_hit_ratio_close_to_exact_lru_on_a_zipf_trace():
    """configs[2] in miniature, everything on the device: per window one batched get, the distinct misses are put
    (the store evicts by itself).  The oracle replays the same trace access by access with exact LRU."""
    cap, n_ids, n_acc, window = 2000, 20000, 120000, 500
    rng = np.random.default_rng(3)
    pz = np.arange(1, n_ids + 1, dtype=np.float64) ** -0.99
    pz /= pz.sum()
    ids = rng.permutation(n_ids)[rng.choice(n_ids, size=n_acc, p=pz)]
    want = orc.replay("LRU", PAGE, cap * PAGE, ids // 1000, ids % 1000, np.full(n_acc, PAGE, np.int32))
    lru_hits = int(want["hit"].sum())
    d = Dev(_open(cap))
    hits = 0
    for w0 in range(0, n_acc, window):
        w = ids[w0:w0 + window]
        r = d.get(w)
        miss = r["len"] < 0
        hits += int((~miss).sum())
        new = np.unique(w[miss])
        if len(new):
            assert (d.put(new) == PAGE).all()
    st = d.stats()
    assert st.pages == cap and st.crc_mismatches == 0
    # a batch cannot hit on a page it is itself about to insert, so it starts a little behind; CLOCK costs a little more
    assert hits >= 0.90 * lru_hits, (hits, lru_hits)
    capi.check(capi.lib().aps_close(d.h))
