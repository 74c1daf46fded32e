"""-m gpu: the many-files MD5 kernel (K4) against the oracle's RFC 1321 restatement and hashlib.
Replaces ChecksumCommand.calculateChecksum (shell/.../ChecksumCommand.java:83-90) per file."""
import hashlib

import numpy as np
import pytest

from oracle import binding as orc

pytestmark = pytest.mark.gpu


def md5_many(blobs, misalign=0):
    import torch
    from alluxio_b200 import capi
    n = len(blobs)
    offs, total = [], 0
    for b in blobs:
        total = (total + 15) // 16 * 16 + misalign
        offs.append(total)
        total += len(b)
    host = np.zeros(total + 16, np.uint8)
    for o, b in zip(offs, blobs):
        host[o:o + len(b)] = np.frombuffer(b, np.uint8)
    dev = torch.from_numpy(host).cuda()
    torch.cuda.synchronize()
    ptrs = (__import__("ctypes").c_void_p * n)(*[dev.data_ptr() + o for o in offs])
    lens = (__import__("ctypes").c_uint64 * n)(*[len(b) for b in blobs])
    out = (__import__("ctypes").c_uint8 * (16 * n))()
    capi.check(capi.lib().aps_md5_batch(0, n, ptrs, lens, out), "aps_md5_batch")
    raw = bytes(out)
    return [raw[16 * i:16 * i + 16].hex() for i in range(n)]


def test_rfc1321_suite_and_reference_vector():
    kats = [(b"", "d41d8cd98f00b204e9800998ecf8427e"), (b"a", "0cc175b9c0f1b6a831c399e269772661"),
            (b"abc", "900150983cd24fb0d6963f7d28e17f72"), (b"message digest", "f96b697d7cb7938d525a2f31aaf161d0"),
            (b"abcdefghijklmnopqrstuvwxyz", "c3fcd3d76192e4007dfb496cca67e13b"),
            (b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789", "d174ab98d277d9f5a5611c2c9f419d9f"),
            (b"1234567890" * 8, "57edf4a22be3c955ac49da2e2107b67a"),
            (bytes(range(10)), "c56bd5480f6e5413cb62a0ad9666613a")]  # ChecksumCommandIntegrationTest.java:32-46
    assert md5_many([k for k, _ in kats]) == [v for _, v in kats]


@pytest.mark.parametrize("misalign", [0, 1, 8])
def test_ragged_files_match_oracle(misalign):
    rng = np.random.default_rng(21 + misalign)
    lens = list(range(0, 130)) + [447, 448, 449, 511, 512, 513, 575, 576, 1023, 1024, 1025, 4096, 65537, 200000,
                                  (1 << 20) + 3]
    blobs = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in lens]
    got = md5_many(blobs, misalign)
    for n, b, g in zip(lens, blobs, got):
        assert g == orc.md5_hex(b) == hashlib.md5(b).hexdigest(), n


def test_many_files_of_synthetic_pages():
    """cfg 4 in miniature: 300 files of 1 MiB + ragged tails, generated on the device."""
    import torch
    from alluxio_b200 import capi
    from gpu_util import fill_pages
    n, flen = 300, (1 << 20)
    buf = torch.empty(n * flen, dtype=torch.uint8, device="cuda:0")
    fill_pages(buf, flen, flen, list(range(n)), [0] * n, 4)
    lens = [flen - (i % 7) * 13 for i in range(n)]
    ptrs = torch.tensor([buf.data_ptr() + i * flen for i in range(n)], dtype=torch.int64, device="cuda:0")
    d_len = torch.tensor(lens, dtype=torch.int64, device="cuda:0")
    d_out = torch.zeros(n * 16, dtype=torch.uint8, device="cuda:0")
    from gpu_util import cur_stream
    capi.check(capi.lib().aps_md5_batch_dev(0, n, ptrs.data_ptr(), d_len.data_ptr(), d_out.data_ptr(), cur_stream()))
    torch.cuda.synchronize()
    out = d_out.cpu().numpy().reshape(n, 16)
    host = buf.cpu().numpy()
    for i in range(n):
        assert out[i].tobytes().hex() == hashlib.md5(host[i * flen:i * flen + lens[i]].tobytes()).hexdigest(), i
