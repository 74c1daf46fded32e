"""-m gpu, collected last: the no-allocation CRC32C entry point (aps_crc32c_noalloc).

It serves callers that cannot take device memory from the driver at the moment they need page
checksums -- the cache is sized to fill HBM, and an allocation would also stall every concurrent
read -- so all of its working memory is a fixed array inside libaps.  The test calls it through the
C ABI twice in a row (the second call meets the first call's leftovers in that array) and requires
the very same CRCs as the existing aps_crc32c_batch_dev entry point: bit-exact, like every integer
result in this repository.  It makes no assumption about what ran before it in the process."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ranges(rng, blob_len, n, max_len):
    lens = rng.integers(0, max_len + 1, n).astype(np.uint32)
    lens[:8] = [0, 1, 15, 16, 511, 512, 513, max_len]
    starts = rng.integers(0, blob_len - max_len, n).astype(np.int64)
    starts[::3] &= ~15  # a third of the ranges start 16-byte aligned (vector path), the rest anywhere
    return starts, lens


def test_matches_allocating_entry_point_on_two_calls_in_a_row():
    import torch
    from alluxio_b200 import capi
    from alluxio_b200.page_store import GpuPageStore
    from gpu_util import cur_stream
    L = capi.lib()
    max_len = 1 << 20
    store = GpuPageStore(page_size=max_len, cache_size=8 * max_len, max_batch=1024)  # reference path needs a store
    rng = np.random.default_rng(2024)
    blob = torch.from_numpy(rng.integers(0, 256, 64 << 20, dtype=np.uint8)).cuda()
    for call, n in enumerate((700, 1024)):
        starts, lens = _ranges(rng, blob.numel(), n, max_len)
        ptrs = torch.from_numpy(starts + blob.data_ptr()).cuda()
        d_len = torch.from_numpy(lens.view(np.int32)).cuda()
        want = torch.full((n,), -1, dtype=torch.int32, device="cuda")
        got = torch.full((n,), -1, dtype=torch.int32, device="cuda")
        capi.check(L.aps_crc32c_batch_dev(store.handle, n, ptrs.data_ptr(), d_len.data_ptr(), want.data_ptr(), cur_stream()))
        capi.check(L.aps_crc32c_noalloc(0, n, ptrs.data_ptr(), d_len.data_ptr(), max_len, got.data_ptr(), cur_stream()),
                   "aps_crc32c_noalloc call %d" % call)
        torch.cuda.synchronize()
        assert torch.equal(got, want), "call %d differs from aps_crc32c_batch_dev" % call
    # and against the oracle for a handful, so both entry points are not wrong together
    from oracle import binding as orc
    host = blob.cpu().numpy()
    g = got.cpu().numpy().view(np.uint32)
    for i in (0, 1, 7, 100, 1023):
        assert int(g[i]) == orc.crc32c(host[starts[i]:starts[i] + lens[i]])
    store.close()


def test_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from alluxio_b200 import capi
    L = capi.lib()
    buf = torch.zeros(64, dtype=torch.int64, device="cuda")
    p = buf.data_ptr()
    assert L.aps_crc32c_noalloc(0, capi.APS_NOALLOC_MAX_RANGES + 1, p, p, 16, p, None) == capi.APS_ELIMIT
    assert L.aps_crc32c_noalloc(0, 1, p, p, capi.APS_NOALLOC_MAX_LEN + 1, p, None) == capi.APS_ELIMIT
    assert b"fixed scratch" in L.aps_strerror(capi.APS_ELIMIT)
    assert L.aps_crc32c_noalloc(0, 0, None, None, 0, None, None) == capi.APS_OK  # nothing to do
