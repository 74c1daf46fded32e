"""-m gpu: the sharded cache.  Two (or three) stores act as shards; requests for any id are served by
the owning shard's table and slab in place, in one kernel.  With a single GPU the shards share the
device (same code path, local "peer" memory); with two or more GPUs they sit on different devices and
the reads cross NVLink."""
import ctypes as C

import numpy as np
import pytest

from oracle import binding as orc

pytestmark = pytest.mark.gpu
SEED = 0xA11C10


def _run(n_shards, devices):
    import torch
    from alluxio_b200 import capi, sharding
    from alluxio_b200.page_store import GpuPageStore, PageId
    from gpu_util import GET_REQ, PUT_REQ, RESULT, from_dev
    L = capi.lib()
    ps = 1 << 16
    n_pages = 300
    stores = [GpuPageStore(page_size=ps, cache_size=n_pages * ps, device=devices[i % len(devices)]) for i in range(n_shards)]
    ids = np.arange(n_pages, dtype=np.uint64)
    files, idx = ids // np.uint64(16), ids % np.uint64(16)
    owners = sharding.shard_of_keys(np.zeros_like(ids), files, idx, n_shards)
    lens = np.where(idx == 15, 1000 + ids, ps).astype(np.uint32)     # ragged last pages
    assert len(set(owners.tolist())) == n_shards
    # puts go to the owner (single-page API, host bytes from the generator)
    for i in range(n_pages):
        page = orc.fill_page(SEED, int(files[i]), int(idx[i]), int(lens[i]))
        stores[owners[i]].put(PageId(int(files[i]), int(idx[i])), page.tobytes())
    handles = (C.c_void_p * n_shards)(*[s.handle for s in stores])
    for r in range(n_shards):
        capi.check(L.aps_shard_attach_local(stores[r].handle, n_shards, r, handles), "attach")
    # every shard asks for a mix of everyone's pages (+ one absent id)
    rng = np.random.default_rng(9)
    for r in range(n_shards):
        dev = torch.device("cuda", devices[r % len(devices)])
        with torch.cuda.device(dev):
            stream = torch.cuda.Stream(dev)
            with torch.cuda.stream(stream):
                pick = rng.permutation(n_pages)[:200]
                reqs = np.zeros(len(pick) + 1, GET_REQ)
                dst = torch.zeros((len(pick) + 1) * ps, dtype=torch.uint8, device=dev)
                reqs["file_lo"][:-1], reqs["page_index"][:-1] = files[pick], idx[pick]
                reqs["file_lo"][-1], reqs["page_index"][-1] = 999999, 3
                reqs["nbytes"] = ps
                reqs["dst"] = dst.data_ptr() + np.arange(len(reqs), dtype=np.uint64) * ps
                d_reqs = torch.from_numpy(reqs.view(np.uint8).reshape(-1).copy()).to(dev)
                d_res = torch.zeros(len(reqs) * 8, dtype=torch.uint8, device=dev)
                capi.check(L.aps_get_batch_sharded_dev(stores[r].handle, len(reqs), d_reqs.data_ptr(), d_res.data_ptr(),
                                                       C.c_void_p(stream.cuda_stream)), "sharded get")
                stream.synchronize()
                res = from_dev(d_res, RESULT)
                host = dst.cpu().numpy()
        assert res[-1]["len"] == -capi.APS_ENOTFOUND
        remote = 0
        for k, i in enumerate(pick):
            want = orc.fill_page(SEED, int(files[i]), int(idx[i]), int(lens[i]))
            assert res[k]["len"] == lens[i], (r, k)
            assert np.array_equal(host[k * ps:k * ps + lens[i]], want), (r, k)
            assert res[k]["crc"] == orc.crc32c(want), (r, k)
            remote += owners[i] != r
        assert remote > 0
        # the host-array form gives the same answers
        pinned = torch.zeros(4 * ps, dtype=torch.uint8).pin_memory()
        h = np.zeros(4, GET_REQ)
        h["file_lo"], h["page_index"], h["nbytes"] = files[pick[:4]], idx[pick[:4]], ps
        h["dst"] = pinned.data_ptr() + np.arange(4, dtype=np.uint64) * ps
        hres = np.zeros(4, RESULT)
        capi.check(L.aps_get_batch_sharded(stores[r].handle, 4, h.ctypes.data, hres.ctypes.data, None))
        assert list(hres["len"]) == [int(lens[i]) for i in pick[:4]]
        assert list(hres["crc"]) == [int(res[k]["crc"]) for k in range(4)]
    for s in stores:
        s.close()


def test_two_shards_on_one_device():
    _run(2, [0])


def test_three_shards_on_one_device():
    _run(3, [0])


def test_shards_on_different_gpus_over_nvlink():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    _run(2, [0, 1])
    if torch.cuda.device_count() >= 4:
        _run(4, [0, 1, 2, 3])


def _copy_from_device(ptr, nbytes, dev):
    """Bytes at a raw device address (memory the library allocated) as a numpy array."""
    from gpu_util import raw_tensor
    return raw_tensor(ptr, nbytes, dev).cpu().numpy()


def _run_push(n_shards, devices, depth, rounds=4):
    """The push transport: every shard posts its multi-get, the OWNERS stream the pages into the requester's landing
    buffer (CRC verified by the owner), the requester waits for their done words.  Checked against the oracle's
    generator and CRC32C, and against the pull form (aps_get_batch_sharded_dev) on the same requests."""
    import torch
    from alluxio_b200 import capi, sharding
    from alluxio_b200.page_store import GpuPageStore, PageId
    from gpu_util import GET_REQ, RESULT, from_dev
    L = capi.lib()
    ps, n_pages, nreq = 1 << 16, 300, 201
    stores = [GpuPageStore(page_size=ps, cache_size=n_pages * ps, device=devices[i % len(devices)], max_batch=256)
              for i in range(n_shards)]
    ids = np.arange(n_pages, dtype=np.uint64)
    files, idx = ids // np.uint64(16), ids % np.uint64(16)
    owners = sharding.shard_of_keys(np.zeros_like(ids), files, idx, n_shards)
    lens = np.where(idx == 15, 1000 + ids, ps).astype(np.uint32)
    for i in range(n_pages):
        page = orc.fill_page(SEED, int(files[i]), int(idx[i]), int(lens[i]))
        stores[owners[i]].put(PageId(int(files[i]), int(idx[i])), page.tobytes())
    handles = (C.c_void_p * n_shards)(*[s.handle for s in stores])
    for r in range(n_shards):
        capi.check(L.aps_push_setup(stores[r].handle, depth * nreq * ps, depth), "push setup")
    for r in range(n_shards):
        capi.check(L.aps_shard_attach_local(stores[r].handle, n_shards, r, handles), "attach")
        capi.check(L.aps_push_attach_local(stores[r].handle, n_shards, r, handles), "push attach")
    rng = np.random.default_rng(19)
    devs = [torch.device("cuda", devices[r % len(devices)]) for r in range(n_shards)]
    streams = []
    for r in range(n_shards):
        with torch.cuda.device(devs[r]):
            streams.append(torch.cuda.Stream(devs[r]))
    land = []
    for r in range(n_shards):
        nb = C.c_uint64()
        land.append((L.aps_push_landing(stores[r].handle, C.byref(nb)), nb.value))
        assert land[r][0] and land[r][1] == depth * nreq * ps
    picks = [[rng.permutation(n_pages)[:nreq - 1] for _ in range(rounds)] for _ in range(n_shards)]
    d_reqs, d_res = {}, {}
    for r in range(n_shards):
        for k in range(rounds):
            reqs = np.zeros(nreq, GET_REQ)
            pick = picks[r][k]
            reqs["file_lo"][:-1], reqs["page_index"][:-1] = files[pick], idx[pick]
            reqs["file_lo"][-1], reqs["page_index"][-1] = 999999, 3          # an id nobody holds
            reqs["nbytes"] = ps
            reqs["page_off"][5] = 100                                         # one partial read
            reqs["dst"] = land[r][0] + (k % depth) * nreq * ps + np.arange(nreq, dtype=np.uint64) * ps
            if k == 1:
                reqs["dst"][7] = 4096                                        # not in the landing buffer: refused, not posted
            with torch.cuda.device(devs[r]):
                d_reqs[r, k] = torch.from_numpy(reqs.view(np.uint8).reshape(-1).copy()).to(devs[r])
                d_res[r, k] = torch.zeros(nreq * 8, dtype=torch.uint8, device=devs[r])
    for d in set(devs):
        torch.cuda.synchronize(d)

    def sp(r):
        return C.c_void_p(streams[r].cuda_stream)

    def check(r, k):
        streams[r].synchronize()
        res = from_dev(d_res[r, k], RESULT)
        host = _copy_from_device(land[r][0] + (k % depth) * nreq * ps, nreq * ps, devs[r])
        assert res[-1]["len"] == -capi.APS_ENOTFOUND, (r, k, res["len"][-8:].tolist(), int((res["len"] == 0).sum()),
                                                       int((res["len"] == -capi.APS_ETIMEOUT).sum()), time.time() - t_round)
        remote = 0
        for q, i in enumerate(picks[r][k]):
            if k == 1 and q == 7:
                assert res[q]["len"] == -capi.APS_EINVAL
                continue
            off = 100 if q == 5 else 0
            n = max(0, int(lens[i]) - off)
            want = orc.fill_page(SEED, int(files[i]), int(idx[i]), int(lens[i]))[off:]
            assert res[q]["len"] == n, (r, k, q, res[q])
            assert np.array_equal(host[q * ps:q * ps + n], want), (r, k, q)
            assert res[q]["crc"] == orc.crc32c(want), (r, k, q)
            remote += owners[i] != r
        assert remote > 0

    import time
    t_round = time.time()
    if depth == 1:
        for k in range(rounds):
            t_round = time.time()
            for r in range(n_shards):
                capi.check(L.aps_get_batch_push_dev(stores[r].handle, nreq, d_reqs[r, k].data_ptr(), d_res[r, k].data_ptr(), sp(r)), "push get")
            for r in range(n_shards):
                check(r, k)
    else:  # two multi-gets in flight per shard: begin(k + 1) is issued before end(k)
        for r in range(n_shards):
            capi.check(L.aps_get_batch_push_begin_dev(stores[r].handle, nreq, d_reqs[r, 0].data_ptr(), sp(r)), "begin")
        for k in range(rounds):
            if k + 1 < rounds:
                for r in range(n_shards):
                    capi.check(L.aps_get_batch_push_begin_dev(stores[r].handle, nreq, d_reqs[r, k + 1].data_ptr(), sp(r)), "begin")
            for r in range(n_shards):
                capi.check(L.aps_get_batch_push_end_dev(stores[r].handle, d_res[r, k].data_ptr(), sp(r)), "end")
            for r in range(n_shards):
                check(r, k)
        assert L.aps_get_batch_push_end_dev(stores[0].handle, d_res[0, 0].data_ptr(), sp(0)) == capi.APS_EINVAL  # nothing in flight
    for s in stores:
        assert s.stats()["crc_mismatches"] == 0
        s.close()


@pytest.mark.parametrize("depth", [1, 2])
def test_push_transport_two_shards_on_one_device(depth):
    _run_push(2, [0], depth)


def test_push_transport_three_shards_on_one_device():
    _run_push(3, [0], 1)


def test_push_transport_across_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    _run_push(2, [0, 1], 1)
    _run_push(2, [0, 1], 2)
    if torch.cuda.device_count() >= 4:
        _run_push(4, [0, 1, 2, 3], 2)


def test_numpy_router_equals_library():
    from alluxio_b200 import capi, sharding
    L = capi.lib()
    rng = np.random.default_rng(1)
    hi = rng.integers(0, 2**63, 300).astype(np.uint64)
    lo = rng.integers(0, 2**63, 300).astype(np.uint64)
    ix = rng.integers(0, 2**63, 300).astype(np.uint64)
    for n in (2, 3, 8):
        o = sharding.shard_of_keys(hi, lo, ix, n)
        for i in range(300):
            k = capi.aps_key(int(hi[i]), int(lo[i]), int(ix[i]))
            assert L.aps_shard_of(C.byref(k), n) == o[i]
