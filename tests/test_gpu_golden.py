"""-m gpu: the committed checksum vectors (tests/golden/checksum_kats.json: RFC 3720 B.4, RFC 1321 A.5, the
reference's MD5 test vector) through the GPU kernels."""
import ctypes as C
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
KATS = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "checksum_kats.json")))


def test_crc32c_vectors_through_put_and_get():
    from alluxio_b200.page_store import GpuPageStore, PageId
    store = GpuPageStore(page_size=4096, cache_size=1 << 20)
    for i, v in enumerate(KATS["crc32c"]):
        data = bytes.fromhex(v["hex"])
        store.put(PageId("kat", i), data)
        buf = bytearray(max(1, len(data)))
        assert store.get(PageId("kat", i), buf, bytes_to_read=len(data)) == len(data)
        assert "%08x" % store.last_crc == v["crc32c"], v["name"]
    # the same bytes at the head of a long page still give the known answer for a partial read
    store.put(PageId("kat", 100), bytes(range(32)) + bytes(4000))
    buf = bytearray(32)
    assert store.get(PageId("kat", 100), buf, bytes_to_read=32) == 32 and "%08x" % store.last_crc == "46dd794e"
    store.close()


def test_md5_vectors():
    import torch
    from alluxio_b200 import capi
    blobs = [bytes.fromhex(v["hex"]) for v in KATS["md5"]]
    n = len(blobs)
    offs, total = [], 0
    for b in blobs:
        offs.append(total)
        total += (len(b) + 15) // 16 * 16 + 16
    host = np.zeros(total, np.uint8)
    for o, b in zip(offs, blobs):
        host[o:o + len(b)] = np.frombuffer(b, np.uint8)
    dev = torch.from_numpy(host).cuda()
    torch.cuda.synchronize()
    ptrs = (C.c_void_p * n)(*[dev.data_ptr() + o for o in offs])
    lens = (C.c_uint64 * n)(*[len(b) for b in blobs])
    out = (C.c_uint8 * (16 * n))()
    capi.check(capi.lib().aps_md5_batch(0, n, ptrs, lens, out))
    for i, v in enumerate(KATS["md5"]):
        assert bytes(out[16 * i:16 * i + 16]).hex() == v["md5"], v["name"]
