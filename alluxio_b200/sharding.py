"""Routing for a cache sharded over several GPUs: which shard owns a PageId.

`shard_of_keys` is the vectorised (numpy) twin of aps_shard_of() in libaps (csrc/aps_device.cuh:
hash_key / shard_of); tests check the two against each other.  Hosts use it to send puts and deletes to
the owner and to lay out benchmark populations; gets need no routing on the host (the kernel reads the
owning shard itself)."""
import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z):
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash_keys(file_hi, file_lo, page_index):
    with np.errstate(over="ignore"):
        hi = np.asarray(file_hi, dtype=np.uint64)
        lo = np.asarray(file_lo, dtype=np.uint64)
        idx = np.asarray(page_index, dtype=np.uint64)
        return _mix64(hi ^ _mix64(lo ^ _mix64(idx + np.uint64(0x9E3779B97F4A7C15))))


def shard_of_keys(file_hi, file_lo, page_index, n_shards):
    return ((hash_keys(file_hi, file_lo, page_index) >> np.uint64(32)) % np.uint64(n_shards)).astype(np.int64)


def owned_by(rank, n_shards, file_lo, page_index):
    """Boolean mask of the (decimal file id, page index) pairs that shard `rank` owns."""
    return shard_of_keys(np.zeros_like(np.asarray(file_lo, dtype=np.uint64)), file_lo, page_index, n_shards) == rank
