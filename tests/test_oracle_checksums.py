"""Pins the oracle's CRC32C and MD5 against published known-answer vectors."""
import hashlib
import os

import numpy as np
import pytest

from oracle import binding as orc

# RFC 3720 B.4 + the classic check value.  The reference has no CRC code: these
# vectors ARE the spec for the fused CRC.
CRC_KATS = [
    (b"123456789", 0xE3069283),
    (bytes(32), 0x8A9136AA),
    (b"\xff" * 32, 0x62A8AB43),
    (bytes(range(32)), 0x46DD794E),
    (bytes(range(31, -1, -1)), 0x113FDB5C),
    (b"", 0x00000000),
]


@pytest.mark.parametrize("impl", ["bitwise", "sb8", "hw"])
def test_crc32c_known_answers(impl):
    for data, want in CRC_KATS:
        got = orc.crc32c(data, impl)
        if got is None:
            pytest.skip("no SSE4.2 in this build")
        assert got == want, (impl, data[:8], hex(got), hex(want))


def test_crc32c_three_implementations_agree():
    rng = np.random.default_rng(7)
    for n in [1, 2, 3, 7, 8, 9, 15, 16, 17, 63, 64, 65, 511, 512, 513, 4095, 4096, 4097, 65536, 1 << 20]:
        for shift in (0, 1, 3):  # unaligned starts
            a = rng.integers(0, 256, n + shift, dtype=np.uint8)[shift:]
            ref = orc.crc32c(a, "bitwise") if n <= 65536 else orc.crc32c(a, "hw")
            assert orc.crc32c(a, "sb8") == ref
            hw = orc.crc32c(a, "hw")
            assert hw is None or hw == ref


def test_crc32c_combine_matches_concatenation():
    rng = np.random.default_rng(8)
    for la, lb in [(0, 5), (5, 0), (1, 1), (100, 37), (4096, 4096), (12345, 1 << 20)]:
        a = rng.integers(0, 256, la, dtype=np.uint8)
        b = rng.integers(0, 256, lb, dtype=np.uint8)
        whole = orc.crc32c(np.concatenate([a, b]))
        assert orc.crc32c_combine(orc.crc32c(a), orc.crc32c(b), lb) == whole


# RFC 1321 appendix A.5
MD5_KATS = [
    (b"", "d41d8cd98f00b204e9800998ecf8427e"),
    (b"a", "0cc175b9c0f1b6a831c399e269772661"),
    (b"abc", "900150983cd24fb0d6963f7d28e17f72"),
    (b"message digest", "f96b697d7cb7938d525a2f31aaf161d0"),
    (b"abcdefghijklmnopqrstuvwxyz", "c3fcd3d76192e4007dfb496cca67e13b"),
    (b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789", "d174ab98d277d9f5a5611c2c9f419d9f"),
    (b"1234567890" * 8, "57edf4a22be3c955ac49da2e2107b67a"),
]


def test_md5_rfc1321_suite():
    for data, want in MD5_KATS:
        assert orc.md5_hex(data) == want


def test_md5_reference_integration_vector():
    # ChecksumCommandIntegrationTest.java:32-46: a 10-byte file of bytes 0..9; the expected value
    # is computed there by DigestUtils.md5Hex at run time.  Digits from SURVEY.md §8c.
    assert orc.md5_hex(bytes(range(10))) == "c56bd5480f6e5413cb62a0ad9666613a"
    assert hashlib.md5(bytes(range(10))).hexdigest() == "c56bd5480f6e5413cb62a0ad9666613a"


def test_md5_matches_hashlib_on_padding_edges():
    rng = np.random.default_rng(9)
    for n in list(range(0, 130)) + [447, 448, 449, 511, 512, 513, 4095, 4096, 65537, (1 << 20) + 3]:
        a = rng.integers(0, 256, n, dtype=np.uint8)
        assert orc.md5_hex(a) == hashlib.md5(a.tobytes()).hexdigest(), n


def test_fill_page_is_counter_based_and_position_dependent():
    a = orc.fill_page(0xA11C10, 3, 5, 4096)
    b = orc.fill_page(0xA11C10, 3, 5, 1000)
    assert np.array_equal(a[:1000], b)
    assert not np.array_equal(a, orc.fill_page(0xA11C10, 3, 6, 4096))
    assert not np.array_equal(a, orc.fill_page(0xA11C10, 4, 5, 4096))
    # first word by hand: splitmix64 first output for the seed
    s = (0xA11C10 ^ (3 << 32) ^ 5) & (2**64 - 1)
    z = (s + 0x9E3779B97F4A7C15) & (2**64 - 1)
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & (2**64 - 1)
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & (2**64 - 1)
    z ^= z >> 31
    assert int.from_bytes(a[:8].tobytes(), "little") == z
