"""The product's native evictors and meta store (alluxio_b200/host, no GPU needed) against the
reference's scripted sequences (CTEST/LRUCacheEvictorTest.java, CTEST/LFUCacheEvictorTest.java,
CTEST/DefaultMetaStoreTest.java) and, on long random op sequences, against the oracle."""
import numpy as np
import pytest

from alluxio_b200 import cache as apc
from alluxio_b200.page_store import FileKeyCodec, PageId
from oracle import binding as orc

F, S, T = PageId("1L", 2), PageId("3L", 4), PageId("5L", 6)
ONE, TWO, THREE, FOUR = PageId("1L", 1), PageId("2L", 2), PageId("3L", 3), PageId("4L", 4)


def test_lru_scripted_sequences():
    e = apc.CacheEvictor("LRU")
    for p in (F, S, T, F):
        e.update_on_get(p)
    assert e.evict() == S; e.update_on_delete(S)
    assert e.evict() == T; e.update_on_delete(T)
    assert e.evict() == F and e.evict() == F  # evict() only peeks
    e.reset()
    assert e.evict() is None
    for p in (F, S, T):
        e.update_on_put(p)
    e.update_on_delete(S)
    assert e.evict() == F; e.update_on_delete(F)
    assert e.evict() == T; e.update_on_delete(T)
    assert e.evict() is None


def test_lfu_scripted_sequences():
    e = apc.CacheEvictor("LFU", 2.0)
    for p in (ONE, ONE, TWO, TWO, THREE, THREE, TWO, THREE, ONE):   # evictUpdatedGetRecentOrder
        e.update_on_get(p)
    assert e.evict() == TWO; e.update_on_delete(TWO)
    assert e.evict() == THREE; e.update_on_delete(THREE)
    assert e.evict() == ONE
    e.reset()
    for p in (ONE, FOUR, FOUR, FOUR, FOUR):                          # evictDeleteSkippedBucket
        e.update_on_get(p)
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == FOUR; e.update_on_delete(FOUR)
    assert e.evict() is None
    for p in (FOUR, TWO, FOUR, TWO, FOUR, ONE, FOUR):                # evictGetMixedOrder
        e.update_on_get(p)
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == TWO; e.update_on_delete(TWO)
    assert e.evict() == FOUR


@pytest.mark.parametrize("kind", ["LRU", "LFU", "FIFO"])
def test_evictors_match_oracle_on_random_sequences(kind):
    rng = np.random.default_rng({"LRU": 1, "LFU": 2, "FIFO": 3}[kind])
    ids = [PageId(str(f), i) for f in range(6) for i in range(10)] + [PageId("some/path", 3), PageId("x" * 40, 1)]
    a, b = apc.CacheEvictor(kind, 2.0), orc.Evictor(kind, 2.0)
    live = set()
    for step in range(6000):
        p = ids[rng.integers(len(ids))]
        o = orc.PageId(p.file_id, p.page_index)
        op = rng.integers(10)
        if op < 4:
            if kind == "FIFO" and p in live:
                continue  # the FIFO double is only ever fed new ids by the manager
            a.update_on_put(p); b.update_on_put(o); live.add(p)
        elif op < 7:
            if p in live or kind != "LFU":
                a.update_on_get(p); b.update_on_get(o)
                if kind != "FIFO":
                    live.add(p)
        elif op < 9:
            a.update_on_delete(p); b.update_on_delete(o); live.discard(p)
        va, vb = a.evict(), b.evict()
        assert (va is None) == (vb is None), step
        if va is not None:
            assert (va.file_id, va.page_index) == (vb.file_id, vb.page_index), (kind, step)


def test_meta_store_matches_reference_cases_and_oracle():
    m = apc.MetaStore("LRU")
    page = PageId("9L", 9)
    assert not m.has_page(page)
    with pytest.raises(apc.PageNotFoundException):
        m.remove_page(page)
    with pytest.raises(apc.PageNotFoundException):
        m.get_page_info(page)
    m.add_page(page, 1024)
    assert m.has_page(page) and m.get_page_info(page) == (page, 1024)
    m.add_page(F, 10); m.add_page(S, 20)
    assert (m.bytes(), m.pages()) == (1054, 3)
    m.has_page(page)
    assert m.evict() == (page, 1024)          # hasPage does not refresh recency ...
    m.get_page_info(page)
    assert m.evict() == (F, 10)               # ... getPageInfo does
    m.remove_page(F)
    assert (m.bytes(), m.pages()) == (1044, 2)
    m.reset()
    assert (m.bytes(), m.pages()) == (0, 0) and m.evict() is None


def test_java_string_hashcode_and_file_keys_agree_with_the_python_codec():
    L = apc.lib()
    for s in ("", "0", "hello", "polygenelubricants", "/a/b", "é中", "\U0001F600"):
        assert L.apc_java_string_hashcode(s.encode("utf-8")) == orc.java_string_hashcode(s)
    py = FileKeyCodec()
    assert py.encode("123") == (0, 123)
    assert py.encode("-1") == (0, 2**64 - 1)
    assert py.encode("0123")[0] == 2**64 - 1          # not canonical -> interned
    assert py.encode("d41d8cd98f00b204e9800998ecf8427e") == (0xd41d8cd98f00b204, 0xe9800998ecf8427e)
    assert py.encode("0000000000000000ffffffffffffffff")[0] == 2**64 - 1  # reserved high word -> interned


def test_hadoop_file_identifier_is_md5_hex_of_the_path_and_maps_exactly():
    """behind the Hadoop-compatible client the file id is md5(path) in hex
    (core/client/hdfs/.../HadoopUtils.java:225-238); its 32 hex characters are the 128-bit device key."""
    import hashlib
    from alluxio_b200.cache import file_key_of, hadoop_file_identifier
    paths = ["", "/", "hdfs://nn:8020/warehouse/t/part-00000.parquet", "s3a://bucket/kéy with spaces/ü", "x" * 55, "y" * 56,
             "z" * 64, "w" * 1000]
    for p in paths:
        want = hashlib.md5(p.encode("utf-8")).hexdigest()
        got = hadoop_file_identifier(p)
        assert got == want
        hi, lo = file_key_of(got)
        assert (hi, lo) == (int(want[:16], 16), int(want[16:], 16))
    assert file_key_of("12345") == (0, 12345)
    assert file_key_of("-1") == (0, (1 << 64) - 1)


def test_meta_store_random_ops_match_a_dict():
    """The meta store and the LRU evictor keep their state in open-addressing maps with backward-shift deletion
    (host/cache.h FlatMap): long random sequences of add / remove / lookup / evict against a plain model."""
    import random
    from collections import OrderedDict
    from alluxio_b200.cache import MetaStore, PageNotFoundException
    from alluxio_b200.page_store import PageId
    rnd = random.Random(5)
    m = MetaStore("LRU")
    model = OrderedDict()          # id -> size, in recency order (put and getPageInfo refresh, hasPage does not)
    ids = [PageId(str(f), i) for f in range(13) for i in range(40)]
    for step in range(30000):
        p = rnd.choice(ids)
        k = (p.file_id, p.page_index)
        op = rnd.random()
        if op < 0.40:
            if k not in model:
                size = rnd.randrange(1, 5000)
                m.add_page(p, size)
                model[k] = size
        elif op < 0.65:
            if k in model:
                m.remove_page(p)
                del model[k]
            else:
                with pytest.raises(PageNotFoundException):
                    m.remove_page(p)
        elif op < 0.85:
            assert m.has_page(p) == (k in model)
            if k in model:
                assert m.get_page_info(p)[1] == model[k]
                model.move_to_end(k)
        else:
            v = m.evict()
            if not model:
                assert v is None
            else:
                head = next(iter(model))
                assert (v[0].file_id, v[0].page_index) == head and v[1] == model[head]
        if step % 997 == 0:
            assert m.pages() == len(model) and m.bytes() == sum(model.values())
    assert m.pages() == len(model) and m.bytes() == sum(model.values())


@pytest.mark.parametrize("kind,base", [("LRU", 2.0), ("LFU", 2.0), ("LFU", 1.5), ("LFU", 10.0)])
def test_evictors_follow_the_oracle_through_random_sequences(kind, base):
    """Product evictor against the oracle's restatement of the reference's, step by step: same victim after every
    update, including deletes of unknown ids, repeated gets (LFU bucket moves) and resets."""
    import random
    from alluxio_b200.cache import CacheEvictor
    from alluxio_b200.page_store import PageId
    from oracle import binding as orc
    rnd = random.Random(int(base * 10) + len(kind))
    ours, ref = CacheEvictor(kind, base), orc.Evictor(kind, base)
    ids = [(str(f), i) for f in range(6) for i in range(25)]
    for step in range(20000):
        f, i = rnd.choice(ids)
        a, b = PageId(f, i), orc.PageId(f, i)
        op = rnd.random()
        if op < 0.35:
            ours.update_on_put(a); ref.update_on_put(b)
        elif op < 0.75:
            ours.update_on_get(a); ref.update_on_get(b)
        elif op < 0.995:
            ours.update_on_delete(a); ref.update_on_delete(b)
        else:
            ours.reset(); ref.reset()
        va, vb = ours.evict(), ref.evict()
        assert (va is None) == (vb is None), step
        if va is not None:
            assert (va.file_id, va.page_index) == (vb.file_id, vb.page_index), step


def test_three_file_key_codecs_agree_on_edge_cases():
    """PageId.fileId -> 128-bit key is written three times: Python (page_store.FileKeyCodec), C++ (host FileKeyCodec,
    apc_file_key_of) and Java (GpuPageStore.fileKey).  The Java one cannot run here, so its two regular expressions are
    read out of the source and applied with Java's String.matches semantics (the WHOLE string must match).  All three
    must put every string into the same class -- decimal long, 32-hex, or interned -- with the same key."""
    import os
    import re
    from alluxio_b200.cache import file_key_of
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "integration", "java", "alluxio",
                            "client", "file", "cache", "store", "GpuPageStore.java")).read()
    java_patterns = re.findall(r'fileId\.matches\("([^"]+)"\)', src)
    assert len(java_patterns) == 2
    java_dec, java_hex = (re.compile(p) for p in java_patterns)
    M = (1 << 64) - 1

    def java(s):
        if java_dec.fullmatch(s):
            v = int(s)
            if -(1 << 63) <= v < (1 << 63):          # Long.parseLong; NumberFormatException falls through to interning
                return 0, v & M
        if java_hex.fullmatch(s):
            hi, lo = int(s[:16], 16), int(s[16:], 16)
            if hi not in (0, M):                      # hi != 0L && hi != -1L
                return hi, lo
        return M, None

    corpus = ["0", "-0", "00", "+1", "1", "-1", "007", "9223372036854775807", "9223372036854775808", "-9223372036854775808",
              "-9223372036854775809", "1" * 19, "9" * 19, "1" * 20, "123\n", "\n123", " 123", "123 ", "12a", "１２３", "",
              "-", "d41d8cd98f00b204e9800998ecf8427e", "D41D8CD98F00B204E9800998ECF8427E", "d41d8cd98f00b204e9800998ecf8427",
              "d41d8cd98f00b204e9800998ecf8427e0", "0000000000000000ffffffffffffffff", "ffffffffffffffff0000000000000000",
              "0000000000000001ffffffffffffffff", "fffffffffffffffe0000000000000000", "d41d8cd98f00b204e9800998ecf8427e\n",
              "123e4567-e89b-12d3-a456-426614174000", "/a/raw/path", "g" * 32, "0" * 32, "f" * 32]
    py = FileKeyCodec()
    for s in corpus:
        want = java(s)
        got_py, got_cc = py.encode(s), file_key_of(s)
        if want[1] is None:
            assert got_py[0] == M and got_cc[0] == M, (s, got_py, got_cc)
        else:
            assert got_py == want and got_cc == want, (s, want, got_py, got_cc)
    # interned ids are exact: different strings never share a key, equal strings always do
    keys = [py.encode(s) for s in corpus]
    assert len(set(keys)) == len(set(corpus)) and [py.encode(s) for s in corpus] == keys
