"""The committed golden vectors (tests/golden/) against the oracle AND the product's host-side policy code.
GPU kernels meet the same checksum vectors in tests/test_gpu_golden.py."""
import json
import os

import pytest

from alluxio_b200 import cache as apc
from alluxio_b200.page_store import PageId
from oracle import binding as orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KATS = json.load(open(os.path.join(GOLD, "checksum_kats.json")))
SEQS = json.load(open(os.path.join(GOLD, "evictor_sequences.json")))


def test_oracle_checksums_match_golden():
    for v in KATS["crc32c"]:
        data = bytes.fromhex(v["hex"])
        for impl in ("bitwise", "sb8"):
            assert "%08x" % orc.crc32c(data, impl) == v["crc32c"], v["name"]
    for v in KATS["md5"]:
        assert orc.md5_hex(bytes.fromhex(v["hex"])) == v["md5"], v["name"]


def _run(evictor, ops, pid):
    for op, arg in ops:
        if op == "e":
            got = evictor.evict()
            want = None if arg is None else (arg[0], arg[1])
            assert (None if got is None else (got.file_id, got.page_index)) == want
        else:
            {"g": evictor.update_on_get, "p": evictor.update_on_put, "d": evictor.update_on_delete}[op](pid(*arg))


@pytest.mark.parametrize("kind", ["LRU", "LFU"])
def test_evictors_replay_the_reference_sequences(kind):
    for name, ops in SEQS[kind].items():
        _run(orc.Evictor(kind, 2.0), ops, orc.PageId)          # the oracle
        _run(apc.CacheEvictor(kind, 2.0), ops, PageId)         # the product's native evictor
