"""CPU: the name-table reader of the snapshot file (aps_snapshot_names, alluxio_b200/csrc/aps_snapshot.cu) is host code and
runs without a GPU.  A snapshot is written here by hand in the documented layout (Header | payloads | Record[n] | name
table); the reader must return exactly the names, and every truncation or corruption of the file must end in a status
code -- never a crash, never a name that was not written."""
import ctypes as C
import os
import struct

import numpy as np

from alluxio_b200 import capi

MAGIC = 0x3150414E53535041  # "APSSNAP1"
NAME_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_char_p)  # aps_name_cb


def snapshot_bytes(names, n_pages=3, page_size=4096):
    payload = b"".join(bytes([i]) * 16 * (i + 1) for i in range(n_pages))
    records = b"".join(struct.pack("<QQQII", 0, 7, i, 16 * (i + 1), 0) for i in range(n_pages))
    table = b""
    if names is not None:
        table = struct.pack("<I", len(names)) + b"".join(struct.pack("<I", len(n)) + n for n in names)
    header = struct.pack("<QIIQQII", MAGIC, 1, page_size, n_pages, len(payload), 1, len(table))
    assert len(header) == 40
    return header + payload + records + table


def read_names(path):
    got = []

    def cb(user, index, name):
        got.append((index, name))
        return 0

    keep = NAME_CB(cb)
    rc = capi.lib().aps_snapshot_names(os.fsencode(path), C.cast(keep, C.c_void_p), None)
    return rc, got


def test_names_round_trip(tmp_path):
    names = [b"123e4567-e89b-12d3-a456-426614174000", b"/a/raw/path", b"", "é中".encode(), b"x" * 5000]
    p = tmp_path / "snap"
    p.write_bytes(snapshot_bytes(names))
    rc, got = read_names(p)
    assert rc == capi.APS_OK and got == list(enumerate(names))
    p.write_bytes(snapshot_bytes(None))                      # written by plain aps_snapshot_save: no table
    assert read_names(p) == (capi.APS_OK, [])
    p.write_bytes(snapshot_bytes([]))                        # an empty table
    assert read_names(p) == (capi.APS_OK, [])
    assert read_names(tmp_path / "absent")[0] == capi.APS_EIO


def test_damaged_files_end_in_a_status(tmp_path):
    names = [b"alpha", b"beta-gamma", b"", b"delta" * 40]
    good = snapshot_bytes(names)
    p = tmp_path / "snap"
    rng = np.random.default_rng(12)
    outcomes = set()
    for cut in list(range(0, 48)) + [len(good) - k for k in range(1, 40)]:          # every truncation near both ends
        p.write_bytes(good[:cut])
        rc, got = read_names(p)
        outcomes.add(rc)
        assert rc in (capi.APS_ECORRUPT, capi.APS_EIO), (cut, rc)
        assert not got
    for _ in range(400):                                                             # random byte and field damage
        b = bytearray(good)
        kind = int(rng.integers(4))
        if kind == 0:                                                                # anywhere in the header
            b[int(rng.integers(40))] = int(rng.integers(256))
        elif kind == 1:                                                              # inside the name table
            b[len(b) - 1 - int(rng.integers(len(b) - 40 - 3 * 48 - 3 * 32))] ^= 1 << int(rng.integers(8))
        elif kind == 2:                                                              # a header field set to an extreme
            off, fmt = [(16, "<Q"), (24, "<Q"), (36, "<I")][int(rng.integers(3))]
            struct.pack_into(fmt, b, off, [0, 1, 2**32 - 1, 2**63, 2**64 - 1][int(rng.integers(5))] & (2**(8 * struct.calcsize(fmt)) - 1))
        else:                                                                        # a name length set to an extreme
            table = len(b) - struct.unpack_from("<I", b, 36)[0]
            struct.pack_into("<I", b, table + 4, [2**32 - 1, 2**31, 10**6][int(rng.integers(3))])
        p.write_bytes(bytes(b))
        rc, got = read_names(p)
        outcomes.add(rc)
        assert rc in (capi.APS_OK, capi.APS_ECORRUPT, capi.APS_EIO), rc
        for index, name in got:                                                      # whatever is returned lies inside the file
            assert index < 4 and len(name) <= len(good)
    assert capi.APS_ECORRUPT in outcomes and capi.APS_OK in outcomes
