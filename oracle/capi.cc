// ORACLE -- TEST INFRASTRUCTURE ONLY (see checksums.c header).
//
// Flat C entry points over the restatement so tests/ and bench.py's CPU legs can drive it with
// ctypes.  A PageId crosses as (const char* fileId, int64 pageIndex).
#include <chrono>
#include <cstring>
#include <random>
#include <thread>

#include "page_cache.h"

using namespace orc;

extern "C" {

// ---------------------------------------------------------------- Java-isms
int32_t orc_java_string_hashcode(const char* utf8) { return JavaStringHashCode(utf8); }
int32_t orc_floor_mod(int32_t x, int32_t y) { return FloorMod(x, y); }

// ---------------------------------------------------------------- evictors
void* orc_evictor_create(const char* kind, double lfu_logbase) { return MakeEvictor(kind, lfu_logbase).release(); }
void orc_evictor_destroy(void* e) { delete (CacheEvictor*)e; }
void orc_evictor_update_on_get(void* e, const char* f, int64_t i) { ((CacheEvictor*)e)->UpdateOnGet(PageId{f, i}); }
void orc_evictor_update_on_put(void* e, const char* f, int64_t i) { ((CacheEvictor*)e)->UpdateOnPut(PageId{f, i}); }
void orc_evictor_update_on_delete(void* e, const char* f, int64_t i) { ((CacheEvictor*)e)->UpdateOnDelete(PageId{f, i}); }
void orc_evictor_reset(void* e) { ((CacheEvictor*)e)->Reset(); }
// 1 and (*file,*idx) filled, or 0 for Java null
int orc_evictor_evict(void* e, char* file_out, size_t cap, int64_t* idx_out) {
  PageId v;
  if (!((CacheEvictor*)e)->Evict(&v)) return 0;
  strncpy(file_out, v.fileId.c_str(), cap - 1);
  file_out[cap - 1] = 0;
  *idx_out = v.pageIndex;
  return 1;
}

// ---------------------------------------------------------------- meta store
void* orc_meta_create(const char* evictor_kind, double lfu_logbase) {
  return new DefaultMetaStore(MakeEvictor(evictor_kind, lfu_logbase));
}
void orc_meta_destroy(void* m) { delete (DefaultMetaStore*)m; }
int orc_meta_has_page(void* m, const char* f, int64_t i) { return ((DefaultMetaStore*)m)->HasPage(PageId{f, i}); }
void orc_meta_add_page(void* m, const char* f, int64_t i, int64_t size) {
  PageId id{f, i};
  ((DefaultMetaStore*)m)->AddPage(id, PageInfo{id, size});
}
// 1 ok (size filled), 0 == PageNotFoundException
int orc_meta_get_page_info(void* m, const char* f, int64_t i, int64_t* size) {
  PageInfo info;
  if (!((DefaultMetaStore*)m)->GetPageInfo(PageId{f, i}, &info)) return 0;
  *size = info.pageSize;
  return 1;
}
int orc_meta_remove_page(void* m, const char* f, int64_t i) { return ((DefaultMetaStore*)m)->RemovePage(PageId{f, i}); }
int64_t orc_meta_bytes(void* m) { return ((DefaultMetaStore*)m)->Bytes(); }
int64_t orc_meta_pages(void* m) { return ((DefaultMetaStore*)m)->Pages(); }
void orc_meta_reset(void* m) { ((DefaultMetaStore*)m)->Reset(); }
int orc_meta_evict(void* m, char* file_out, size_t cap, int64_t* idx_out, int64_t* size_out) {
  PageInfo v;
  if (!((DefaultMetaStore*)m)->Evict(&v)) return 0;
  strncpy(file_out, v.pageId.fileId.c_str(), cap - 1);
  file_out[cap - 1] = 0;
  *idx_out = v.pageId.pageIndex;
  *size_out = v.pageSize;
  return 1;
}

// ---------------------------------------------------------------- page stores
// wipe_first=1 is PageStore.create, 0 is PageStore.open.  NULL on failure (IOException).
void* orc_store_open_local(const char* root, int64_t page_size, int64_t cache_size, int32_t file_buckets,
                           int wipe_first) {
  return OpenLocalPageStore(root, page_size, cache_size, file_buckets, wipe_first != 0).release();
}
void* orc_store_open_mem(int64_t page_size, int64_t cache_size, int keep_bytes) {
  return OpenMemPageStore(page_size, cache_size, keep_bytes != 0).release();
}
// A store implemented by the caller (see ExternalPageStoreCalls): callbacks return 0 ok, 1 PageNotFoundException,
// 2 IllegalArgumentException, 3 IOException.
void* orc_store_open_external(void* user, int (*put)(void*, const char*, int64_t, const uint8_t*, int32_t),
                              int (*get)(void*, const char*, int64_t, int32_t, int32_t, uint8_t*, int32_t, int32_t, int32_t*),
                              int (*del)(void*, const char*, int64_t),
                              int (*get_pages)(void*, void (*)(void*, int, const char*, int64_t, int64_t), void*),
                              int64_t cache_size) {
  return OpenExternalPageStore(ExternalPageStoreCalls{user, put, get, del, get_pages}, cache_size).release();
}
void orc_store_destroy(void* s) { delete (PageStore*)s; }
int orc_store_put(void* s, const char* f, int64_t i, const uint8_t* page, int32_t len) {
  return ((PageStore*)s)->Put(PageId{f, i}, page, len);
}
int orc_store_get(void* s, const char* f, int64_t i, int32_t page_offset, int32_t bytes_to_read, uint8_t* buffer,
                  int32_t buf_len, int32_t buffer_offset, int32_t* bytes_read) {
  return ((PageStore*)s)->Get(PageId{f, i}, page_offset, bytes_to_read, buffer, buf_len, buffer_offset, bytes_read);
}
int orc_store_delete(void* s, const char* f, int64_t i) { return ((PageStore*)s)->Delete(PageId{f, i}); }
void orc_store_set_faulty(void* s, int put_faulty, int delete_faulty) {
  ((PageStore*)s)->putFaulty = put_faulty != 0;
  ((PageStore*)s)->deleteFaulty = delete_faulty != 0;
}
typedef void (*orc_page_cb)(void* user, int valid, const char* file, int64_t idx, int64_t size);
int orc_store_get_pages(void* s, orc_page_cb cb, void* user) {
  std::vector<std::pair<bool, PageInfo>> v;
  int st = ((PageStore*)s)->GetPages(&v);
  if (st != kOk) return st;
  for (auto& e : v) cb(user, e.first, e.second.pageId.fileId.c_str(), e.second.pageId.pageIndex, e.second.pageSize);
  return kOk;
}

// ---------------------------------------------------------------- manager
void* orc_manager_create(void* store, int64_t page_size, const char* evictor_kind, double lfu_logbase) {
  auto meta = std::unique_ptr<DefaultMetaStore>(new DefaultMetaStore(MakeEvictor(evictor_kind, lfu_logbase)));
  return new LocalCacheManager(page_size, std::move(meta), (PageStore*)store);
}
void orc_manager_destroy(void* m) { delete (LocalCacheManager*)m; }
int orc_manager_put(void* m, const char* f, int64_t i, const uint8_t* page, int32_t len) {
  return ((LocalCacheManager*)m)->Put(PageId{f, i}, page, len);
}
// bytes read; 0 miss; -1 error; -2 IllegalArgumentException
int32_t orc_manager_get(void* m, const char* f, int64_t i, int32_t page_offset, int32_t bytes_to_read, uint8_t* buffer,
                        int32_t buf_len, int32_t offset_in_buffer) {
  return ((LocalCacheManager*)m)->Get(PageId{f, i}, page_offset, bytes_to_read, buffer, buf_len, offset_in_buffer);
}
int orc_manager_delete(void* m, const char* f, int64_t i) { return ((LocalCacheManager*)m)->Delete(PageId{f, i}); }
int orc_manager_state(void* m) { return ((LocalCacheManager*)m)->state(); }
int32_t orc_manager_page_lock_id(void* m, const char* f, int64_t i) {
  return ((LocalCacheManager*)m)->GetPageLockId(PageId{f, i});
}
// out[18], in the order of LocalCacheManager::Metrics (page_cache.h)
void orc_manager_metrics(void* m, int64_t* out) {
  const LocalCacheManager::Metrics& x = ((LocalCacheManager*)m)->metrics;
  const int64_t v[18] = {x.bytesWrittenCache, x.bytesEvicted, x.pagesEvicted, x.getErrors, x.getNotReadyErrors,
                         x.getStoreReadErrors, x.cleanupGetErrors, x.putErrors, x.putNotReadyErrors, x.putEvictionErrors,
                         x.putStoreWriteErrors, x.putStoreDeleteErrors, x.putBenignRacingErrors, x.cleanupPutErrors,
                         x.deleteErrors, x.deleteNotReadyErrors, x.deleteNonExistingPageErrors, x.deleteStoreDeleteErrors};
  memcpy(out, v, sizeof(v));
}
int64_t orc_manager_bytes(void* m) { return ((LocalCacheManager*)m)->meta()->Bytes(); }
int64_t orc_manager_pages(void* m) { return ((LocalCacheManager*)m)->meta()->Pages(); }

// ---------------------------------------------------------------- stream (CACHE/LocalCacheFileInStream.java)
// external(user, pos, dst, len) -> bytes read stands for the external FileInStream.
typedef int32_t (*orc_external_read)(void* user, int64_t pos, uint8_t* dst, int32_t len);
void* orc_stream_open(void* m, const char* file_id, int64_t file_length, int64_t page_size, orc_external_read external,
                      void* user) {
  return new LocalCacheFileInStream(file_id, file_length, page_size, (LocalCacheManager*)m,
                                    [external, user](int64_t pos, uint8_t* dst, int32_t len) { return external(user, pos, dst, len); });
}
void orc_stream_destroy(void* s) { delete (LocalCacheFileInStream*)s; }
// bytes read; -1 end of file; -2 IllegalArgumentException; -3 IOException
int32_t orc_stream_read(void* s, uint8_t* b, int32_t b_len, int32_t off, int32_t len) {
  return ((LocalCacheFileInStream*)s)->Read(b, b_len, off, len);
}
int32_t orc_stream_positioned_read(void* s, int64_t pos, uint8_t* b, int32_t b_len, int32_t off, int32_t len) {
  return ((LocalCacheFileInStream*)s)->PositionedRead(pos, b, b_len, off, len);
}
int64_t orc_stream_skip(void* s, int64_t n) { return ((LocalCacheFileInStream*)s)->Skip(n); }  // -4 closed
int orc_stream_seek(void* s, int64_t pos) { return ((LocalCacheFileInStream*)s)->Seek(pos); }  // 0, -2, -4 closed
int64_t orc_stream_remaining(void* s) { return ((LocalCacheFileInStream*)s)->Remaining(); }
int64_t orc_stream_get_pos(void* s) { return ((LocalCacheFileInStream*)s)->GetPos(); }
void orc_stream_close(void* s) { ((LocalCacheFileInStream*)s)->Close(); }
// out[0..3] = bytes read from the cache, bytes requested from external, bytes read from external, external read calls
void orc_stream_counters(void* s, int64_t* out) {
  auto* st = (LocalCacheFileInStream*)s;
  out[0] = st->bytesReadCache;
  out[1] = st->bytesRequestedExternal;
  out[2] = st->bytesReadExternal;
  out[3] = st->externalReads;
}

// ---------------------------------------------------------------- trace replay (BASELINE config 3)
// Runs the reference read flow for each access: CacheManager.get; on a miss (0) the caller reads
// the page externally and calls CacheManager.put (CACHE/LocalCacheFileInStream.java:121-143).
// fileId = decimal file[k].  Page length = page_len[k] (last pages may be short).  Page payload is
// not needed to decide hits and victims, so the store keeps lengths only.
// Outputs: hit[k] (1 = get returned bytes), put_ok[k] (result of the put after a miss, else 255),
// victims (file, page) in eviction order, up to victims_cap; returns the number of victims.
int64_t orc_replay(const char* evictor_kind, double lfu_logbase, int64_t page_size, int64_t cache_size, int64_t n,
                   const int64_t* file, const int64_t* page, const int32_t* page_len, uint8_t* hit, uint8_t* put_ok,
                   int64_t* victim_file, int64_t* victim_page, int64_t victims_cap, int64_t* final_pages,
                   int64_t* final_bytes) {
  auto store = OpenMemPageStore(page_size, cache_size, false);
  auto meta = std::unique_ptr<DefaultMetaStore>(new DefaultMetaStore(MakeEvictor(evictor_kind, lfu_logbase)));
  LocalCacheManager mgr(page_size, std::move(meta), store.get());
  int64_t nv = 0;
  size_t seen = 0;
  for (int64_t k = 0; k < n; k++) {
    PageId id{std::to_string(file[k]), page[k]};
    int32_t got = mgr.Get(id, 0, page_len[k], nullptr, page_len[k], 0);
    hit[k] = got > 0;
    put_ok[k] = 255;
    if (got <= 0) put_ok[k] = mgr.Put(id, nullptr, page_len[k]) ? 1 : 0;
    for (; seen < mgr.evicted.size(); seen++, nv++) {
      if (nv < victims_cap) {
        victim_file[nv] = std::stoll(mgr.evicted[seen].fileId);
        victim_page[nv] = mgr.evicted[seen].pageIndex;
      }
    }
    if (mgr.evicted.size() > (1u << 20)) {
      mgr.evicted.clear();
      seen = 0;
    }
  }
  *final_pages = mgr.meta()->Pages();
  *final_bytes = mgr.meta()->Bytes();
  return nv;
}

// ---------------------------------------------------------------- synthetic pages
// Page payload used by every benchmark and parity test: 64-bit word k of page
// (file, index) is the k-th output of splitmix64 seeded with seed ^ (file << 32) ^ index, written
// little-endian.  Counter-based, so any byte range can be produced independently.
static inline uint64_t splitmix_word(uint64_t seed, uint64_t k) {
  uint64_t z = seed + (k + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
void orc_fill_page(uint64_t seed, uint64_t file, uint64_t index, uint8_t* out, int64_t len) {
  uint64_t s = seed ^ (file << 32) ^ index;
  int64_t k = 0;
  for (; (k + 1) * 8 <= len; k++) {
    uint64_t w = splitmix_word(s, (uint64_t)k);
    memcpy(out + 8 * k, &w, 8);
  }
  if (8 * k < len) {
    uint64_t w = splitmix_word(s, (uint64_t)k);
    memcpy(out + 8 * k, &w, (size_t)(len - 8 * k));
  }
}

// ---------------------------------------------------------------- CPU baseline (BASELINE config 1)
// LocalCacheManager + LocalPageStore on a real directory: populate n_files x pages_per_file pages
// of page_size bytes, then time n_gets uniform random whole-page gets on n_threads threads.  Each
// thread has its own mt19937_64(seed + t) and its own destination buffer, like independent client
// threads sharing one CacheManager.  Gets are read-only on the manager, so no lock is needed for
// a race-free run (the reference takes a striped read lock + the meta read lock; uncontended).
// Returns seconds for the timed gets, or a negative value on failure; *bytes_out = payload bytes.
// If check != 0 the first word of every returned page is compared with the generator.
double orc_bench_local_get(const char* root, int64_t page_size, int64_t n_files, int64_t pages_per_file, int populate,
                           int64_t n_gets, uint64_t data_seed, uint64_t access_seed, int32_t n_threads, int check,
                           int64_t* bytes_out) {
  int64_t total_pages = n_files * pages_per_file;
  auto store = OpenLocalPageStore(root, page_size, total_pages * page_size, 1000, populate != 0);
  if (!store) return -1.0;
  if (populate) {
    std::vector<uint8_t> pagebuf((size_t)page_size);
    for (int64_t f = 0; f < n_files; f++)
      for (int64_t p = 0; p < pages_per_file; p++) {
        orc_fill_page(data_seed, (uint64_t)f, (uint64_t)p, pagebuf.data(), page_size);
        if (store->Put(PageId{std::to_string(f), p}, pagebuf.data(), (int32_t)page_size) != kOk) return -2.0;
      }
  }
  auto meta = std::unique_ptr<DefaultMetaStore>(new DefaultMetaStore(MakeEvictor("LRU", 2.0)));
  LocalCacheManager mgr(page_size, std::move(meta), store.get());  // restore() re-scans the directory
  if (mgr.state() != READ_WRITE || mgr.meta()->Pages() != total_pages) return -3.0;
  std::vector<int> bad((size_t)n_threads, 0);
  std::vector<std::thread> th;
  auto t0 = std::chrono::steady_clock::now();
  for (int32_t t = 0; t < n_threads; t++) {
    th.emplace_back([&, t]() {
      std::mt19937_64 rng(access_seed + (uint64_t)t);
      std::vector<uint8_t> buf((size_t)page_size);
      int64_t mine = n_gets / n_threads + (t < n_gets % n_threads ? 1 : 0);
      for (int64_t k = 0; k < mine; k++) {
        uint64_t r = rng() % (uint64_t)total_pages;
        int64_t f = (int64_t)(r / (uint64_t)pages_per_file), p = (int64_t)(r % (uint64_t)pages_per_file);
        int32_t got = mgr.Get(PageId{std::to_string(f), p}, 0, (int32_t)page_size, buf.data(), (int32_t)page_size, 0);
        if (got != page_size) bad[(size_t)t]++;
        if (check) {
          uint64_t w, expect = splitmix_word(data_seed ^ ((uint64_t)f << 32) ^ (uint64_t)p, 0);
          memcpy(&w, buf.data(), 8);
          if (w != expect) bad[(size_t)t]++;
        }
      }
    });
  }
  for (auto& x : th) x.join();
  double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (int b : bad) if (b) return -4.0;
  *bytes_out = n_gets * page_size;
  return secs;
}

}  // extern "C"
