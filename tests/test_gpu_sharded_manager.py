"""-m gpu: ONE cache over several GPU stores ("global policy on host driving all shards").
The reference has a single manager with a single byte budget (CACHE/LocalCacheManager.java:261); here that
manager drives a ShardedGpuPageStore, page -> shard hash(PageId) mod n.  Everything the reference defines --
hit/miss sequence, put results, victim order, final residency, returned bytes -- must equal the oracle's
single-store run.  The shards share device 0 on a one-GPU box; with two GPUs visible the same cases also
run across devices (puts and gets then cross NVLink)."""
import numpy as np
import pytest

from conftest import increasing
from oracle import binding as orc

pytestmark = pytest.mark.gpu

PS = 1024


def PID(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def _device_sets():
    import torch
    sets = [[0, 0], [0, 0, 0]]
    if torch.cuda.device_count() >= 2:
        sets.append([0, 1])
    return sets


def manager(devices, cache_size, evictor="LRU", page_size=PS, **kw):
    from alluxio_b200.cache import LocalCacheManager
    return LocalCacheManager(page_size, cache_size, evictor=evictor, shard_devices=devices, **kw)


@pytest.mark.parametrize("devices", _device_sets(), ids=lambda d: "dev" + "".join(map(str, d)))
def test_pages_live_on_their_owner_and_read_back(devices):
    m = manager(devices, 256 * PS, slab_slots=256)
    n = len(devices)
    assert m.shard_count() == n
    want, per_shard = {}, [0] * n
    for k in range(200):
        pid = PID(str(k % 9), k)
        data = increasing(k % 256, 1 + (k * 37) % PS)
        assert m.put(pid, data)
        want[pid] = data
        per_shard[m.shard_of(pid)] += 1
    assert all(c > 0 for c in per_shard)
    assert [m.shard_stats(i)["pages"] for i in range(n)] == per_shard
    assert m.pages() == 200 and m.bytes() == sum(len(d) for d in want.values())
    buf = bytearray(PS)
    for pid, data in want.items():
        assert m.get(pid, len(data), buf) == len(data) and bytes(buf[:len(data)]) == data
    for pid in list(want)[::2]:
        assert m.delete(pid)
        assert m.get(pid, 1, buf) == 0
    assert sum(m.shard_stats(i)["pages"] for i in range(n)) == m.pages() == 100
    m.close()


def test_eviction_and_faults_as_the_single_store_manager():
    m = manager([0, 0], PS, evictor="FIFO", slab_slots=8)           # budget of one page over two shards
    p1, p2 = PID("0L", 0), PID("1L", 1)
    a, b = increasing(PS), increasing(255, PS)
    buf = bytearray(PS)
    assert m.put(p1, a) and m.put(p2, b)                             # putEvict
    assert m.get(p1, PS, buf) == 0
    assert m.get(p2, PS, buf) == PS and bytes(buf) == b
    m.set_store_faulty(delete=True)                                  # failedPageStoreDeleteOnEviction
    assert not m.put(p1, a)
    m.set_store_faulty(put=True)                                     # recoverCacheFromFailedPut
    assert not m.put(p1, a)
    m.set_store_faulty()
    assert m.put(p1, a)
    assert m.get(p1, PS, buf) == PS and bytes(buf) == a
    m.close()


def _trace(n_ids, n_acc, ps, seed):
    rng = np.random.default_rng(seed)
    pz = np.arange(1, n_ids + 1, dtype=np.float64) ** -0.99
    pz /= pz.sum()
    ids = rng.permutation(n_ids)[rng.choice(n_ids, size=n_acc, p=pz)]
    files, pages = ids // 100, ids % 100
    lens = np.where(pages == 99, 1000 + files, ps).astype(np.int32)
    return files, pages, lens


@pytest.mark.parametrize("devices", _device_sets(), ids=lambda d: "dev" + "".join(map(str, d)))
@pytest.mark.parametrize("evictor", ["LRU", "LFU"])
def test_replays_equal_the_oracle(evictor, devices):
    ps = 65536
    n_ids, cap_pages, n_acc = 8000, 600, 20000
    files, pages, lens = _trace(n_ids, n_acc, ps, 5)
    cache_bytes = cap_pages * ps
    want = orc.replay(evictor, ps, cache_bytes, files, pages, lens)
    for batched in (False, True):
        m = manager(devices, cache_bytes, evictor=evictor, page_size=ps, slab_slots=cap_pages, max_batch=1024)
        if batched:
            got = m.replay_batched(files, pages, lens, seed=0xA11C10, window=512)
            assert got["bad_gets"] == 0 and got["gets"] == int(want["hit"].sum())
        else:
            got = m.replay(files, pages, lens, seed=0xA11C10, verify=True)
            assert got["verify_failures"] == 0
        assert np.array_equal(got["hit"], want["hit"])
        assert np.array_equal(got["put_ok"], want["put_ok"])
        assert got["n_victims"] == want["n_victims"] > 0
        assert np.array_equal(got["victims"], want["victims"])
        assert (got["final_pages"], got["final_bytes"]) == (want["final_pages"], want["final_bytes"])
        stats = [m.shard_stats(i) for i in range(len(devices))]
        assert sum(s["pages"] for s in stats) == got["final_pages"]
        assert sum(s["bytes"] for s in stats) == got["final_bytes"]
        assert all(s["crc_mismatches"] == 0 and s["pages"] > 0 for s in stats)
        m.close()


def test_batched_positioned_read_over_shards():
    ps = 4096
    file_len = 11 * ps + 333
    data = np.random.default_rng(18).integers(0, 256, file_len, dtype=np.uint8).tobytes()
    m = manager([0, 0], 64 * ps, page_size=ps, slab_slots=64)
    ext = lambda pos, n: data[pos:pos + n]  # noqa: E731
    rng = np.random.default_rng(19)
    for _ in range(30):
        pos = int(rng.integers(0, file_len))
        n = int(rng.integers(1, 5 * ps))
        got, _, _ = m.positioned_read("77", file_len, pos, n, ext, batched=True)
        assert got == data[pos:pos + n]
    got, fc, fe = m.positioned_read("77", file_len, 0, file_len, ext, batched=True)
    got, fc, fe = m.positioned_read("77", file_len, 0, file_len, ext, batched=True)
    assert got == data and fc == file_len and fe == 0
    m.close()
