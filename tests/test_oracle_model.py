"""A second, independent statement of the reference's cache manager -- forty lines of Python written from
CACHE/LocalCacheManager.java (put :206-347, get :360-399, delete :402-429), DefaultMetaStore.java and
LRUCacheEvictor.java, sharing no code with oracle/page_cache.cc -- driven by hypothesis against the oracle.  The oracle
is pinned by the reference's own tests; this guards the restatement against mistakes those tests do not reach (long
random histories, orphaned store pages, budget arithmetic with pages of every length)."""
from collections import OrderedDict

from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import binding as orc

PS = 64          # page size
CAP = 5 * PS     # cache size in bytes


class Model:
    """LocalCacheManager + DefaultMetaStore + LRUCacheEvictor + an in-memory PageStore, serial, sync write."""

    def __init__(self):
        self.meta = OrderedDict()   # PageId -> length, in LRU order (LRUCacheEvictor: a LinkedHashMap in access order;
        self.store = {}             # LocalCacheManager.get never calls updateOnGet, so the order is the put order)

    def bytes(self):
        return sum(self.meta.values())

    def put(self, pid, page):
        if pid in self.meta:                                   # :256-260 an existing page is not overwritten
            return True
        if self.bytes() + len(page) <= CAP:                    # :261-285
            self.meta[pid] = len(page)
            self.store[pid] = page
            return True
        if not self.meta:                                      # :265-271 nothing to evict
            return False
        victim = next(iter(self.meta))                         # LRUCacheEvictor.evict: the eldest entry
        del self.meta[victim]                                  # :301
        enough = self.bytes() + len(page) <= CAP               # :310
        if enough:
            self.meta[pid] = len(page)
        del self.store[victim]                                 # :318 whatever `enough` says
        if not enough:                                         # :331-334 one victim per put
            return False
        self.store[pid] = page
        return True

    def get(self, pid, page_offset, n, room):
        if page_offset > PS or n > room:                       # :362-366
            return "IAE", b""
        if pid not in self.meta:                               # :376-382
            return 0, b""
        page = self.store[pid]
        if page_offset > len(page):                            # LocalPageStore.get :97-99, not caught by getPage
            return "IAE", b""
        data = page[page_offset:page_offset + min(n, room)]
        if len(data) != n or n <= 0:                           # getPage :544-549, get :384-395
            del self.meta[pid]                                 # the store keeps its copy
            return -1, b""
        return n, data

    def delete(self, pid):
        if pid not in self.meta:                               # :411-419
            return False
        del self.meta[pid]
        del self.store[pid]
        return True


ids = st.tuples(st.integers(0, 2), st.integers(0, 3))
ops = st.lists(st.one_of(
    # mostly whole pages and whole-page reads, so that histories fill the cache and evict; the odd lengths and offsets
    # reach the short-read, zero-byte and illegal-argument branches
    st.tuples(st.just("put"), ids, st.sampled_from([PS, PS, PS, PS, PS - 1, PS // 2, 1, 0]), st.integers(0, 255)),
    st.tuples(st.just("put"), ids, st.sampled_from([PS, PS, PS, PS, PS - 1, PS // 2, 1, 0]), st.integers(0, 255)),
    st.tuples(st.just("get"), ids, st.sampled_from([0, 0, 0, 0, 1, PS // 2, PS, PS + 1]),
              st.sampled_from([PS, PS, PS, PS - 1, PS // 2, 1, 0]), st.sampled_from([0, 0, 0, 1])),
    st.tuples(st.just("delete"), ids)), min_size=10, max_size=200)


@settings(max_examples=400, deadline=None, derandomize=True)  # 4,000 random examples were run once when this was written
@given(ops)
def test_oracle_manager_equals_the_independent_model(history):
    model = Model()
    om = orc.LocalCacheManager(orc.PageStore.open_mem(PS, CAP), PS, "LRU")
    for op in history:
        pid = orc.PageId(str(op[1][0]), op[1][1])
        if op[0] == "put":
            page = bytes((op[3] + k) & 0xFF for k in range(op[2]))
            assert om.put(pid, page) == model.put(op[1], page), op
        elif op[0] == "get":
            _, _, off, n, short = op
            room = max(0, n - short)
            buf = bytearray(room)
            try:
                got = om.get(pid, n, buf, page_offset=off)
            except ValueError:
                got = "IAE"
            want, data = model.get(op[1], off, n, room)
            assert got == want, (op, got, want)
            if isinstance(want, int) and want > 0:
                assert bytes(buf[:want]) == data, op
        else:
            assert om.delete(pid) == model.delete(op[1]), op
        assert (om.pages(), om.bytes()) == (len(model.meta), model.bytes()), op


# ---- LFUCacheEvictor (CACHE/evictor/LFUCacheEvictor.java:61-153), stated again and independently ---------------------

import math


class LfuModel:
    """Counts per page, buckets by (int)(ln(count) / ln(base)), LRU inside a bucket, and mMinBucket maintained exactly
    as the Java does (a page that leaves the minimum bucket empty takes mMinBucket along to its new bucket, :117-119;
    a delete that empties it walks upwards, :143-149)."""

    def __init__(self, base):
        self.divisor = math.log(base)
        self.count, self.buckets, self.min_bucket = {}, {}, -1

    def bucket_of(self, count):
        return int(math.log(count) / self.divisor)

    def update(self, pid):                                     # updateOnGet = updateOnPut (:108-133)
        new_count = self.count[pid] = self.count.get(pid, 0) + 1
        new_bucket = self.bucket_of(new_count)
        if new_count > 1:
            old_bucket = self.bucket_of(new_count - 1)
            if new_bucket != old_bucket:
                pages = self.buckets.get(old_bucket)
                if pages is not None:
                    pages.pop(pid, None)
                    if not pages:
                        del self.buckets[old_bucket]
                        if old_bucket == self.min_bucket:
                            self.min_bucket = new_bucket
                self.buckets.setdefault(new_bucket, OrderedDict())[pid] = True
            elif pid in self.buckets.get(new_bucket, {}):
                self.buckets[new_bucket].move_to_end(pid)      # LinkedHashMap in access order: get() touches
        else:
            self.min_bucket = new_bucket
            self.buckets.setdefault(new_bucket, OrderedDict())[pid] = True

    def delete(self, pid):                                     # :135-151
        count = self.count.pop(pid, None)
        if count is None:
            return
        bucket = self.bucket_of(count)
        pages = self.buckets.get(bucket)
        if pages is not None:
            pages.pop(pid, None)
            if not pages:
                del self.buckets[bucket]
                if bucket == self.min_bucket and self.buckets:
                    while self.min_bucket not in self.buckets:
                        self.min_bucket += 1

    def evict(self):                                           # :153-164: the eldest page of bucket mMinBucket, or null
        pages = self.buckets.get(self.min_bucket)
        return next(iter(pages)) if pages else None


evictor_ops = st.lists(st.tuples(st.sampled_from(["get", "get", "put", "delete", "evict", "evict_and_delete"]), st.integers(0, 7)),
                       min_size=5, max_size=300)


@settings(max_examples=300, deadline=None, derandomize=True)
@given(evictor_ops, st.sampled_from([2.0, 1.1, 1.5, 3.0, 10.0]))
def test_oracle_lfu_evictor_equals_the_independent_model(history, base):
    model, oe = LfuModel(base), orc.Evictor("LFU", base)
    for op, k in history:
        pid = orc.PageId("f", k)
        if op in ("get", "put"):
            model.update(k)
            (oe.update_on_get if op == "get" else oe.update_on_put)(pid)
        elif op == "delete":
            model.delete(k)
            oe.update_on_delete(pid)
        else:
            want, got = model.evict(), oe.evict()
            assert (None if got is None else got.page_index) == want, (op, k, base)
            if op == "evict_and_delete" and want is not None:  # what DefaultMetaStore.evict's caller does next
                model.delete(want)
                oe.update_on_delete(orc.PageId("f", want))
