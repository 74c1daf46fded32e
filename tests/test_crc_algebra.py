"""CPU replay of the fused-CRC kernel's arithmetic (alluxio_b200/csrc/aps_kernels.cuh: warp_rows,
warp_bytes, copy_kernel's chunk combine) using the GF(2) constants libaps computes on the host.
Runs without a GPU: it proves the tables, the per-lane shift constants and the combine formulas
before any kernel is launched.  The device result itself is checked in tests/test_gpu_*.py."""
import ctypes as C

import numpy as np
import pytest

from alluxio_b200 import capi
from oracle import binding as orc

POLY = 0x82F63B78
CHUNK, ROW = 16384, 512  # CHUNK only sizes the test inputs


def gf_mul(a, b):
    r = 0
    for i in range(32):
        if (b >> (31 - i)) & 1:
            r ^= a
        a = (a >> 1) ^ (POLY if a & 1 else 0)
    return r


@pytest.fixture(scope="module")
def consts():
    fold = np.zeros((4, 256), np.uint32)
    z4 = np.zeros((4, 256), np.uint32)
    xp = np.zeros(32, np.uint32)
    ls = np.zeros(32, np.uint32)
    capi.check(capi.lib().aps_debug_crc_constants(fold.ctypes.data, z4.ctypes.data, xp.ctypes.data, ls.ctypes.data))
    return fold, z4, xp, ls


def xpow8(n, xp):
    r, k = 0x80000000, 0
    while n:
        if n & 1:
            r = gf_mul(r, int(xp[k]))
        n >>= 1
        k += 1
    return r


def apply4(tab, s):
    """vectorised T0[s&255]^T1[(s>>8)&255]^T2[(s>>16)&255]^T3[s>>24] on uint32 arrays"""
    return tab[0][s & 0xFF] ^ tab[1][(s >> 8) & 0xFF] ^ tab[2][(s >> 16) & 0xFF] ^ tab[3][s >> 24]


def warp_rows(data, consts):
    fold, z4, xp, ls = consts
    rows = data.size // ROW
    w = data[:rows * ROW].view("<u4").reshape(rows, 32, 4)  # [row][lane][k]
    S = np.zeros((32, 4), np.uint32)
    for r in range(rows):
        S = apply4(fold, S) ^ w[r]
    t = apply4(z4, apply4(z4, apply4(z4, S[:, 0]) ^ S[:, 1]) ^ S[:, 2]) ^ S[:, 3]
    acc = 0
    for lane in range(32):
        acc ^= gf_mul(int(t[lane]), int(ls[lane]))
    return acc


def crc_bitwise_zero_init(b):
    r = 0
    for x in b:
        r ^= int(x)
        for _ in range(8):
            r = (r >> 1) ^ (POLY if r & 1 else 0)
    return r


def warp_bytes(data, consts):
    xp = consts[2]
    n = data.size
    per = (n + 31) >> 5
    acc = 0
    for lane in range(32):
        lo = min(n, lane * per)
        hi = min(n, lo + per)
        acc ^= gf_mul(crc_bitwise_zero_init(data[lo:hi]), xpow8(n - hi, xp))
    return acc


def kernel_crc(data, consts, aligned=True, rows_per_task=32):
    """copy_kernel<true> for one job of len(data) bytes cut into tasks of rows_per_task rows: every
    task CRC is shifted by the bytes that follow it (x^(8*after)) and XORed into the accumulator; the
    finisher adds init/xorout."""
    xp = consts[2]
    n = data.size
    task_bytes = rows_per_task * ROW
    ntasks = (n + task_bytes - 1) // task_bytes
    acc = 0
    for c in range(ntasks):
        task = data[c * task_bytes:min(n, (c + 1) * task_bytes)]
        rows = task.size // ROW if aligned else 0
        tail = task.size - rows * ROW
        crc = warp_rows(task, consts) if rows else 0
        if tail:
            ct = warp_bytes(task[rows * ROW:], consts)
            crc = (gf_mul(crc, xpow8(tail, xp)) ^ ct) if rows else ct
        after = n - min(n, (c + 1) * task_bytes)
        acc ^= gf_mul(crc, xpow8(after, xp))
    return acc ^ gf_mul(0xFFFFFFFF, xpow8(n, xp)) ^ 0xFFFFFFFF


def test_constants_match_their_definitions(consts):
    fold, z4, xp, ls = consts
    assert int(xp[0]) == 0x00800000  # x^8 in the reflected register
    # advancing the register by n zero bytes == multiplying by x^(8n): check against the oracle
    for n in (1, 4, 512, 4096, 12345):
        for r in (1, 0x80000000, 0xDEADBEEF):
            raw = r
            for _ in range(n):
                raw = (raw >> 8) ^ crc_bitwise_zero_init(bytes([raw & 0xFF]))
            assert gf_mul(r, xpow8(n, xp)) == raw
    for j in range(4):
        for b in (0, 1, 0x80, 0xFF):
            assert int(fold[j][b]) == gf_mul((b << (8 * j)) & 0xFFFFFFFF, xpow8(512, xp))
            assert int(z4[j][b]) == gf_mul((b << (8 * j)) & 0xFFFFFFFF, xpow8(4, xp))
    for lane in (0, 1, 31):
        assert int(ls[lane]) == xpow8(500 - 16 * lane, xp)


@pytest.mark.parametrize("n", [0, 1, 15, 16, 511, 512, 513, 1024, 4096 + 7, CHUNK - 1, CHUNK, CHUNK + 1,
                               CHUNK + 512, 3 * CHUNK + 100, 65536])
def test_kernel_arithmetic_matches_oracle(consts, n):
    rng = np.random.default_rng(n)
    data = rng.integers(0, 256, n, dtype=np.uint8)
    assert kernel_crc(data, consts) == orc.crc32c(data)


def test_task_partitions_do_not_change_the_result(consts):
    rng = np.random.default_rng(77)
    data = rng.integers(0, 256, 100000, dtype=np.uint8)
    want = orc.crc32c(data)
    for rpt in (1, 7, 114, 228, 4096):
        assert kernel_crc(data, consts, rows_per_task=rpt) == want


def test_unaligned_path_matches_oracle(consts):
    rng = np.random.default_rng(5)
    for n in (33, 700, CHUNK + 9):
        data = rng.integers(0, 256, n, dtype=np.uint8)
        assert kernel_crc(data, consts, aligned=False) == orc.crc32c(data)


def test_known_answers_through_kernel_arithmetic(consts):
    assert kernel_crc(np.frombuffer(b"123456789", np.uint8), consts) == 0xE3069283
    assert kernel_crc(np.zeros(32, np.uint8), consts) == 0x8A9136AA
    assert kernel_crc(np.full(32, 0xFF, np.uint8), consts) == 0x62A8AB43
    assert kernel_crc(np.arange(32, dtype=np.uint8), consts) == 0x46DD794E


def warp_rows_independent(data, consts, fold4, group=4):
    """stream_rows with kIndep: row u of every group of four goes into its own state, carried forward by a whole group
    (tables x^(8*512*4)); the states are merged -- Horner with the one-row tables -- at the end of the segment, or just
    before a last, incomplete group, whose rows then go through the one-row tables."""
    fold, z4, xp, ls = consts
    rows = data.size // ROW
    w = data[:rows * ROW].view("<u4").reshape(rows, 32, 4)
    chains = [np.zeros((32, 4), np.uint32) for _ in range(group)]
    merged = None
    r = 0
    while r < rows:
        cnt = min(group, rows - r)
        if cnt == group:
            for u in range(group):
                chains[u] = apply4(fold4, chains[u]) ^ w[r + u]
        else:
            merged = chains[0]
            for u in range(1, group):
                merged = apply4(fold, merged) ^ chains[u]
            for u in range(cnt):
                merged = apply4(fold, merged) ^ w[r + u]
        r += cnt
    if merged is None:
        merged = chains[0]
        for u in range(1, group):
            merged = apply4(fold, merged) ^ chains[u]
    S = merged
    t = apply4(z4, apply4(z4, apply4(z4, S[:, 0]) ^ S[:, 1]) ^ S[:, 2]) ^ S[:, 3]
    acc = 0
    for lane in range(32):
        acc ^= gf_mul(int(t[lane]), int(ls[lane]))
    return acc


def test_four_independent_states_equal_the_serial_fold(consts):
    fold4 = np.zeros((4, 256), np.uint32)
    capi.check(capi.lib().aps_debug_crc_fold4(fold4.ctypes.data))
    xp = consts[2]
    for j in range(4):
        for b in (1, 0x5A, 0xFF):
            assert int(fold4[j][b]) == gf_mul((b << (8 * j)) & 0xFFFFFFFF, xpow8(4 * 512, xp))
    rng = np.random.default_rng(4)
    for rows in (1, 2, 3, 4, 5, 7, 8, 9, 31, 32, 33, 128):
        data = rng.integers(0, 256, rows * ROW, dtype=np.uint8)
        got = warp_rows_independent(data, consts, fold4)
        assert got == warp_rows(data, consts), rows
        # zero-init CRC of the rows -> the real CRC32C through init / xorout
        assert got ^ gf_mul(0xFFFFFFFF, xpow8(data.size, xp)) ^ 0xFFFFFFFF == orc.crc32c(data), rows
