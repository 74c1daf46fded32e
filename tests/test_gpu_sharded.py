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
