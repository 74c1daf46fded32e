"""Host-side logic of the sharded cache on CPU, world_size 2 over gloo: every rank computes the same
owner for every PageId, the owners partition the id space, and the numpy router equals aps_shard_of()."""
import ctypes as C
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from alluxio_b200 import capi, sharding
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(1234)  # same ids on every rank
    files = rng.integers(0, 1 << 40, 5000).astype(np.uint64)
    pages = rng.integers(0, 1 << 20, 5000).astype(np.uint64)
    owners = sharding.shard_of_keys(np.zeros(5000, np.uint64), files, pages, world)
    # the C function agrees with the numpy router
    L = capi.lib()
    for i in range(0, 5000, 97):
        k = capi.aps_key(0, int(files[i]), int(pages[i]))
        assert L.aps_shard_of(C.byref(k), world) == owners[i]
    mine = torch.from_numpy((owners == rank).astype(np.int64))
    got = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(got, mine)
    total = sum(got)
    assert bool((total == 1).all()), "every id must have exactly one owner"
    all_owners = [torch.zeros(5000, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(all_owners, torch.from_numpy(owners))
    assert all(torch.equal(all_owners[0], o) for o in all_owners), "ranks disagree on ownership"
    share = float(mine.sum()) / 5000
    assert 0.4 < share < 0.6
    dist.barrier()
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")
    dist.destroy_process_group()


def test_two_ranks_agree_on_ownership(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]


def test_router_spreads_and_is_independent_of_bucket_bits():
    from alluxio_b200 import sharding
    ids = np.arange(1 << 16, dtype=np.uint64)
    for n in (2, 3, 4, 8):
        o = sharding.shard_of_keys(np.zeros_like(ids), ids // np.uint64(1024), ids % np.uint64(1024), n)
        counts = np.bincount(o, minlength=n)
        assert counts.min() > 0.9 * len(ids) / n and counts.max() < 1.1 * len(ids) / n
        # inside one shard the table's bucket bits (low hash bits) are still uniform
        h = sharding.hash_keys(np.zeros_like(ids), ids // np.uint64(1024), ids % np.uint64(1024))
        low = (h[o == 0] & np.uint64(255)).astype(np.int64)
        b = np.bincount(low, minlength=256)
        assert b.min() > 0.5 * b.mean() and b.max() < 1.6 * b.mean()
