"""Ports of CTEST/store/PageStoreTest.java:79-183 and CTEST/store/LocalPageStoreTest.java:42-82
against the oracle's LocalPageStore restatement (a real directory) and its in-RAM twin."""
import os
import uuid

import pytest

from conftest import increasing
from oracle import binding as orc
from oracle.binding import PageId


@pytest.fixture(params=["local", "mem"])
def store(request, tmp_path):
    if request.param == "local":
        return orc.PageStore.create_local(tmp_path, page_size=1024, cache_size=65536)
    return orc.PageStore.open_mem(1024, 65536)


def test_hello_world(store):
    msg = b"Hello, World!"
    pid = PageId("0", 0)
    store.put(pid, msg)
    buf = bytearray(1024)
    assert store.get(pid, buf) == len(msg)
    assert bytes(buf[:len(msg)]) == msg
    store.delete(pid)
    with pytest.raises(orc.PageNotFoundException):
        store.get(pid, buf)


def test_get_offset(store):
    n = 32
    pid = PageId("0", 0)
    store.put(pid, increasing(n))
    buf = bytearray(n)
    for offset in range(1, n):
        got = store.get(pid, buf, page_offset=offset, bytes_to_read=n)
        assert got == n - offset
        assert bytes(buf[:got]) == increasing(offset, n - offset)


def test_get_offset_overflow(store):
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    with pytest.raises(ValueError):
        store.get(pid, bytearray(1024), page_offset=36, bytes_to_read=32)


def test_get_offset_equal_to_length_reads_zero(store):
    # LocalPageStore.java:99 allows pageOffset == pageLength
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    assert store.get(pid, bytearray(8), page_offset=32, bytes_to_read=8) == 0


def test_get_pages(store):
    data = increasing(32)
    want = set()
    for i in range(16):
        store.put(PageId("0", i), data)
        want.add((PageId("0", i), 32))
    assert set(store.get_pages()) == want


def test_get_pages_uuid(store):
    data = increasing(32)
    want = set()
    for i in range(16):
        pid = PageId(str(uuid.uuid4()), i)
        store.put(pid, data)
        want.add((pid, 32))
    assert set(store.get_pages()) == want


def test_get_small_len(store):
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    buf = bytearray(1024)
    for b in range(1, 32):
        assert store.get(pid, buf, bytes_to_read=b) == b
        assert bytes(buf[:b]) == increasing(b)


def test_get_small_buffer(store):
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    for b in range(1, 32):
        buf = bytearray(b)
        assert store.get(pid, buf, bytes_to_read=32) == b
        assert bytes(buf) == increasing(b)


def test_put_overwrites(store):
    # LocalPageStore.java:77 opens with append=false: the whole file is replaced
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    store.put(pid, increasing(7, 8))
    buf = bytearray(64)
    assert store.get(pid, buf) == 8 and bytes(buf[:8]) == increasing(7, 8)


def test_delete_missing(store):
    with pytest.raises(orc.PageNotFoundException):
        store.delete(PageId("0", 0))


def test_negative_offset_and_buffer_offset_beyond_buffer(store):
    pid = PageId("0", 0)
    store.put(pid, increasing(32))
    with pytest.raises(ValueError):
        store.get(pid, bytearray(8), page_offset=-1)
    with pytest.raises(ValueError):
        store.get(pid, bytearray(8), buffer_offset=9)


# ---- LocalPageStoreTest.java:48-73 (bucket layout) + path scheme LocalPageStore.java:139-147

def test_single_file_bucket(tmp_path):
    s = orc.PageStore.create_local(tmp_path, page_size=1 << 20, file_buckets=1)
    for i in range(100):
        s.put(PageId(str(i), 0), b"test")
    assert len(os.listdir(tmp_path / str(1 << 20))) == 1


def test_multi_file_bucket(tmp_path):
    s = orc.PageStore.create_local(tmp_path, page_size=1 << 20, file_buckets=10)
    for i in range(100):
        s.put(PageId(str(i), 0), b"test")
    assert len(os.listdir(tmp_path / str(1 << 20))) == 10


def test_path_scheme_and_java_hashcode(tmp_path):
    # "0".hashCode() == 48; floorMod(48, 1000) == 48
    s = orc.PageStore.create_local(tmp_path, page_size=1024)
    s.put(PageId("0", 7), b"x")
    assert os.path.isfile(tmp_path / "1024" / "48" / "0" / "7")
    # negative hash: "polygenelubricants".hashCode() == Integer.MIN_VALUE
    assert orc.java_string_hashcode("polygenelubricants") == -(2**31)
    assert orc.floor_mod(-(2**31), 1000) == 352
    assert orc.java_string_hashcode("hello") == 99162322
    assert orc.java_string_hashcode("") == 0
    # non-BMP code point hashes over its two UTF-16 surrogates: "\U0001F600" -> 0xD83D, 0xDE00
    assert orc.java_string_hashcode("\U0001F600") == ((0xD83D * 31 + 0xDE00) & 0xFFFFFFFF)


def test_delete_removes_empty_file_dir(tmp_path):
    s = orc.PageStore.create_local(tmp_path, page_size=1024)
    s.put(PageId("0", 1), b"x"); s.put(PageId("0", 2), b"y")
    s.delete(PageId("0", 1))
    assert os.path.isdir(tmp_path / "1024" / "48" / "0")
    s.delete(PageId("0", 2))
    assert not os.path.exists(tmp_path / "1024" / "48" / "0")


def test_get_pages_flags_unrecognised_files(tmp_path):
    s = orc.PageStore.create_local(tmp_path, page_size=1024)
    s.put(PageId("0", 1), b"x")
    (tmp_path / "invalidPageFile").write_bytes(b"")
    assert None in s.get_pages()
    # wrong bucket directory is rejected too (LocalPageStore.java:162-164)
    os.makedirs(tmp_path / "1024" / "49" / "0")
    (tmp_path / "1024" / "49" / "0" / "3").write_bytes(b"z")
    assert s.get_pages().count(None) == 2
