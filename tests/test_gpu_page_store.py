"""-m gpu: GpuPageStore through the C ABI against the reference's PageStoreTest cases
(CTEST/store/PageStoreTest.java:79-183) and, op by op, against the oracle's LocalPageStore."""
import uuid

import numpy as np
import pytest

from conftest import increasing
from oracle import binding as orc

pytestmark = pytest.mark.gpu


@pytest.fixture
def store():
    from alluxio_b200.page_store import GpuPageStore
    s = GpuPageStore(page_size=1024, cache_size=65536, slab_slots=256)
    yield s
    s.close()


def PID(f, i):
    from alluxio_b200.page_store import PageId
    return PageId(f, i)


def test_hello_world(store):
    from alluxio_b200.page_store import PageNotFoundException
    msg = b"Hello, World!"
    pid = PID("0", 0)
    store.put(pid, msg)
    buf = bytearray(1024)
    assert store.get(pid, buf) == len(msg)
    assert bytes(buf[:len(msg)]) == msg
    assert store.last_crc == orc.crc32c(msg)
    store.delete(pid)
    with pytest.raises(PageNotFoundException):
        store.get(pid, buf)


def test_get_offset(store):
    n = 32
    pid = PID("0", 0)
    store.put(pid, increasing(n))
    buf = bytearray(n)
    for offset in range(1, n):
        got = store.get(pid, buf, page_offset=offset, bytes_to_read=n)
        assert got == n - offset
        assert bytes(buf[:got]) == increasing(offset, n - offset)
        assert store.last_crc == orc.crc32c(increasing(offset, n - offset))


def test_get_offset_overflow(store):
    pid = PID("0", 0)
    store.put(pid, increasing(32))
    with pytest.raises(ValueError):
        store.get(pid, bytearray(1024), page_offset=36, bytes_to_read=32)
    assert store.get(pid, bytearray(8), page_offset=32, bytes_to_read=8) == 0  # == length is legal


def test_get_pages(store):
    data = increasing(32)
    want = set()
    for i in range(16):
        store.put(PID("0", i), data)
        want.add((PID("0", i), 32))
    assert set(store.get_pages()) == want


def test_get_pages_uuid(store):
    data = increasing(32)
    want = set()
    for i in range(16):
        pid = PID(str(uuid.uuid4()), i)
        store.put(pid, data)
        want.add((pid, 32))
    assert set(store.get_pages()) == want


def test_get_small_len_and_small_buffer(store):
    pid = PID("0", 0)
    store.put(pid, increasing(32))
    buf = bytearray(1024)
    for b in range(1, 32):
        assert store.get(pid, buf, bytes_to_read=b) == b
        assert bytes(buf[:b]) == increasing(b)
        small = bytearray(b)
        assert store.get(pid, small, bytes_to_read=32) == b
        assert bytes(small) == increasing(b)


def test_put_overwrites_and_delete_missing(store):
    from alluxio_b200.page_store import PageNotFoundException
    pid = PID("0", 0)
    store.put(pid, increasing(32))
    store.put(pid, increasing(7, 8))
    buf = bytearray(64)
    assert store.get(pid, buf) == 8 and bytes(buf[:8]) == increasing(7, 8)
    assert store.stats()["pages"] == 1 and store.stats()["bytes"] == 8
    with pytest.raises(PageNotFoundException):
        store.delete(PID("0", 1))


def test_illegal_arguments(store):
    pid = PID("0", 0)
    store.put(pid, increasing(32))
    with pytest.raises(ValueError):
        store.get(pid, bytearray(8), page_offset=-1)
    with pytest.raises(ValueError):
        store.get(pid, bytearray(8), buffer_offset=9)
    with pytest.raises(ValueError):
        store.put(PID("0", 1), bytes(1025))  # longer than a slot


def test_extreme_ids_and_cache_size(store):
    # LocalCacheManagerTest.java:300-312 uses Long.MAX_VALUE page indices
    for f in ("0", str(2**63 - 1), str(-2**63), "ffffffffffffffffffffffffffffffff", "00000000000000000000000000000001",
              "d41d8cd98f00b204e9800998ecf8427e", "/a/raw/path", ""):
        for idx in (0, 2**63 - 2, 2**63 - 1):
            store.put(PID(f, idx), (f + str(idx)).encode())
    for f in ("0", str(2**63 - 1), str(-2**63), "ffffffffffffffffffffffffffffffff", "00000000000000000000000000000001",
              "d41d8cd98f00b204e9800998ecf8427e", "/a/raw/path", ""):
        for idx in (0, 2**63 - 2, 2**63 - 1):
            buf = bytearray(64)
            n = store.get(PID(f, idx), buf)
            assert bytes(buf[:n]) == (f + str(idx)).encode()
    assert store.get_cache_size() == 65536
    assert len(store.get_pages()) == 24


def test_fault_injection(store):
    from alluxio_b200.page_store import IOExceptionJ
    store.set_faulty(put=True)
    with pytest.raises(IOExceptionJ):
        store.put(PID("0", 0), b"x")
    store.set_faulty()
    store.put(PID("0", 0), b"x")
    store.set_faulty(delete=True)
    with pytest.raises(IOExceptionJ):
        store.delete(PID("0", 0))
    store.set_faulty()
    store.delete(PID("0", 0))


def test_crc_detects_a_flipped_byte(store):
    from alluxio_b200.page_store import IOExceptionJ
    pid = PID("9", 9)
    data = increasing(3, 1000)
    store.put(pid, data)
    buf = bytearray(1000)
    assert store.get(pid, buf) == 1000 and store.last_crc == orc.crc32c(data)
    store.debug_corrupt(pid, 500)
    with pytest.raises(IOExceptionJ):
        store.get(pid, buf)
    assert store.stats()["crc_mismatches"] == 1
    # a partial read cannot be verified and still succeeds
    assert store.get(pid, bytearray(10), bytes_to_read=10) == 10


def test_no_space_is_an_io_exception():
    from alluxio_b200.page_store import GpuPageStore, IOExceptionJ
    s = GpuPageStore(page_size=1024, cache_size=4096, slab_slots=4)
    for i in range(4):
        s.put(PID("1", i), b"a")
    with pytest.raises(IOExceptionJ):
        s.put(PID("1", 4), b"a")
    s.delete(PID("1", 0))
    s.put(PID("1", 4), b"a")
    assert s.stats()["pages"] == 4 and s.stats()["slots_free"] == 0
    s.close()


@pytest.mark.parametrize("page_size", [48, 100, 1024, 4099, 4112, 65536, 1000 * 16])
def test_random_ops_match_oracle_local_page_store(page_size, tmp_path):
    """Same random put/get/delete sequence into GpuPageStore and the oracle's LocalPageStore:
    every status, length, byte and CRC must agree."""
    from alluxio_b200.page_store import GpuPageStore, PageNotFoundException
    rng = np.random.default_rng(page_size)
    gpu = GpuPageStore(page_size=page_size, cache_size=64 * page_size, slab_slots=128)
    ref = orc.PageStore.create_local(tmp_path, page_size=page_size, cache_size=64 * page_size)
    ids = [(str(f), i) for f in range(4) for i in range(12)]
    lens = [n for n in (0, 1, 15, 16, 17, 511, 512, 513, page_size - 1, page_size) if n <= page_size]
    resident = 0
    for step in range(600):
        f, i = ids[rng.integers(len(ids))]
        g, r = PID(f, i), orc.PageId(f, i)
        op = rng.integers(10)
        if op < 4 and resident < 100:
            n = int(lens[rng.integers(len(lens))]) if rng.integers(2) else int(rng.integers(0, page_size + 1))
            data = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
            gpu.put(g, data)
            ref.put(r, data)
        elif op < 9:
            off = int(rng.integers(0, page_size + 2))
            want = int(rng.integers(0, page_size + 1))
            cap = int(rng.integers(0, page_size + 1))
            boff = int(rng.integers(0, 4))
            gb, rb = bytearray(cap + boff), bytearray(cap + boff)
            outcome = []
            for st, p, b in ((gpu, g, gb), (ref, r, rb)):
                try:
                    outcome.append(("ok", st.get(p, b, page_offset=off, bytes_to_read=want, buffer_offset=boff)))
                except (PageNotFoundException, orc.PageNotFoundException):
                    outcome.append(("notfound", 0))
                except ValueError:
                    outcome.append(("iae", 0))
            assert outcome[0] == outcome[1], (step, off, want, cap, boff)
            assert gb == rb
            if outcome[0][0] == "ok":
                n = outcome[0][1]
                assert gpu.last_crc == orc.crc32c(bytes(rb[boff:boff + n]))
        else:
            outcome = []
            for st, p in ((gpu, g), (ref, r)):
                try:
                    st.delete(p)
                    outcome.append("ok")
                except (PageNotFoundException, orc.PageNotFoundException):
                    outcome.append("notfound")
            assert outcome[0] == outcome[1]
        resident = gpu.stats()["pages"] if step % 50 == 0 else resident
    got = sorted((p.file_id, p.page_index, n) for p, n in gpu.get_pages())
    want = sorted((p.file_id, p.page_index, n) for p, n in ref.get_pages())
    assert got == want
    st = gpu.stats()
    assert st["pages"] == len(want) and st["bytes"] == sum(n for _, _, n in want)
    gpu.close()
