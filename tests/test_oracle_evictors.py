"""Ports of CTEST/LRUCacheEvictorTest.java:38-107, CTEST/LFUCacheEvictorTest.java:39-154 and
CTEST/DefaultMetaStoreTest.java:45-89 against the oracle.  (CTEST/ = /root/reference/core/client/
fs/src/test/java/alluxio/client/file/cache/.)  These are the reference's golden victim sequences."""
import pytest

from oracle import binding as orc
from oracle.binding import PageId

F, S, T = PageId("1L", 2), PageId("3L", 4), PageId("5L", 6)


def lru():
    return orc.Evictor("LRU")


def test_lru_evict_get_order():
    e = lru(); e.update_on_get(F); e.update_on_get(S)
    assert e.evict() == F
    e.update_on_delete(F)
    assert e.evict() == S


def test_lru_evict_updated_get_order():
    e = lru()
    for p in (F, S, T, F):
        e.update_on_get(p)
    assert e.evict() == S; e.update_on_delete(S)
    assert e.evict() == T; e.update_on_delete(T)
    assert e.evict() == F


def test_lru_evict_put_order():
    e = lru(); e.update_on_put(F); e.update_on_put(S)
    assert e.evict() == F
    e.update_on_delete(F)
    assert e.evict() == S


def test_lru_evict_updated_put_order():
    e = lru()
    for p in (F, S, T, F):
        e.update_on_put(p)
    assert e.evict() == S; e.update_on_delete(S)
    assert e.evict() == T; e.update_on_delete(T)
    assert e.evict() == F


def test_lru_evict_after_delete():
    e = lru()
    for p in (F, S, T):
        e.update_on_put(p)
    e.update_on_delete(S)
    assert e.evict() == F; e.update_on_delete(F)
    assert e.evict() == T


def test_lru_evict_empty_and_all_gone():
    e = lru()
    assert e.evict() is None
    for p in (F, S, T):
        e.update_on_put(p)
    for p in (F, S, T):
        e.update_on_delete(p)
    assert e.evict() is None


def test_lru_evict_does_not_remove():
    # LRUCacheEvictor.java:65-69: evict() only peeks
    e = lru(); e.update_on_put(F); e.update_on_put(S)
    assert e.evict() == F and e.evict() == F


ONE, TWO, THREE, FOUR = PageId("1L", 1), PageId("2L", 2), PageId("3L", 3), PageId("4L", 4)


def lfu():
    return orc.Evictor("LFU", 2.0)


def gets(e, seq):
    for p in seq:
        e.update_on_get(p)


def test_lfu_evict_get_order():
    e = lfu(); gets(e, [ONE, TWO, TWO, FOUR, FOUR, FOUR, FOUR])
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == TWO; e.update_on_delete(TWO)
    assert e.evict() == FOUR; e.update_on_delete(FOUR)


def test_lfu_evict_get_mixed_order():
    e = lfu(); gets(e, [FOUR, TWO, FOUR, TWO, FOUR, ONE, FOUR])
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == TWO; e.update_on_delete(TWO)
    assert e.evict() == FOUR; e.update_on_delete(FOUR)


def test_lfu_evict_get_recent_order():
    e = lfu(); gets(e, [ONE, TWO])
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == TWO


def test_lfu_evict_updated_get_recent_order():
    e = lfu(); gets(e, [ONE, ONE, TWO, TWO, THREE, THREE, TWO, THREE, ONE])
    assert e.evict() == TWO; e.update_on_delete(TWO)
    assert e.evict() == THREE; e.update_on_delete(THREE)
    assert e.evict() == ONE


def test_lfu_evict_delete_skipped_bucket():
    e = lfu(); gets(e, [ONE, FOUR, FOUR, FOUR, FOUR])
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == FOUR; e.update_on_delete(FOUR)


def test_lfu_evict_delete_twice():
    e = lfu(); gets(e, [ONE, ONE])
    assert e.evict() == ONE
    e.update_on_delete(ONE)
    e.update_on_get(ONE)
    assert e.evict() == ONE


def test_lfu_evict_after_delete():
    e = lfu()
    for p in (ONE, TWO, TWO, THREE, THREE, THREE):
        e.update_on_put(p)
    e.update_on_delete(TWO)
    assert e.evict() == ONE; e.update_on_delete(ONE)
    assert e.evict() == THREE


def test_lfu_evict_empty_and_all_gone():
    e = lfu()
    assert e.evict() is None
    for p in (ONE, TWO, TWO, THREE, THREE, THREE):
        e.update_on_put(p)
    for p in (ONE, TWO, THREE):
        e.update_on_delete(p)
    assert e.evict() is None


# ---- DefaultMetaStoreTest.java
PAGE = PageId("1L", 2)


def test_meta_add_new_exist_has():
    m = orc.MetaStore("LRU")
    assert not m.has_page(PAGE)
    m.add_page(PAGE, 1024)
    assert m.has_page(PAGE)
    m.add_page(PAGE, 1024)
    assert m.has_page(PAGE)


def test_meta_remove():
    m = orc.MetaStore("LRU")
    m.add_page(PAGE, 1024)
    m.remove_page(PAGE)
    assert not m.has_page(PAGE)
    with pytest.raises(orc.PageNotFoundException):
        m.remove_page(PAGE)


def test_meta_get_page_info():
    m = orc.MetaStore("LRU")
    with pytest.raises(orc.PageNotFoundException):
        m.get_page_info(PAGE)
    m.add_page(PAGE, 1024)
    assert m.get_page_info(PAGE) == (PAGE, 1024)


def test_meta_counters_and_evict_heals_unknown_victim():
    m = orc.MetaStore("LRU")
    m.add_page(F, 10); m.add_page(S, 20)
    assert (m.bytes(), m.pages()) == (30, 2)
    assert m.evict() == (F, 10)
    m.remove_page(F)
    assert (m.bytes(), m.pages()) == (20, 1)
    m.reset()
    assert (m.bytes(), m.pages()) == (0, 0) and m.evict() is None


def test_get_page_info_refreshes_lru_but_has_page_does_not():
    # only getPageInfo calls updateOnGet (DefaultMetaStore.java:66-72)
    m = orc.MetaStore("LRU")
    m.add_page(F, 1); m.add_page(S, 1)
    m.has_page(F)
    assert m.evict()[0] == F
    m.get_page_info(F)
    assert m.evict()[0] == S
