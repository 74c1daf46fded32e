// TEST INFRASTRUCTURE ONLY.  A multi-threaded driver of the host layer's C API (include/aps_cache.h) over the store
// double (fake_aps.cc), meant to be compiled with -fsanitize=thread or -fsanitize=address,undefined together with
// alluxio_b200/host/cache.cc (tests/test_host_sanitizers.py does that and runs it).  It checks what must hold for ANY
// interleaving, as the reference's own concurrency rules promise (CACHE/LocalCacheManager.java:172-203: striped page
// locks + the meta lock):
//   * a get that returns bytes returns the bytes of THAT page (contents are a function of the id);
//   * once the threads are done, the meta store and the page store hold the same pages and the same bytes, and the
//     byte budget was never exceeded at any sampling point;
//   * streams return the file's bytes whatever else happens to the cache;
//   * nothing races, leaks, overflows or touches freed memory (the sanitizers' part).
// Exit code 0 = all invariants held; the sanitizer runtime makes it non-zero by itself when it reports.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../../include/aps_cache.h"

namespace {

constexpr int kPage = 256;
std::atomic<int> g_failures{0};

#define CHECK(cond, ...)                       \
  do {                                         \
    if (!(cond)) {                             \
      fprintf(stderr, "FAILED %s:%d: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);            \
      fprintf(stderr, "\n");                   \
      g_failures++;                            \
    }                                          \
  } while (0)

// page (file, idx) holds len(file, idx) bytes of a pattern that depends on both
int page_len(int file, int idx) { return (file * 7 + idx * 13) % 5 == 0 ? 1 + (file * 31 + idx * 17) % (kPage - 1) : kPage; }
uint8_t page_byte(int file, int idx, int k) { return (uint8_t)(file * 131 + idx * 31 + k * 7 + (k >> 3)); }
void fill(int file, int idx, uint8_t* out, int len) {
  for (int k = 0; k < len; ++k) out[k] = page_byte(file, idx, k);
}

struct FileSource {  // the external file system of the streams: file f is n_pages pages of its pattern
  int file, n_pages;
  int64_t length() const {
    int64_t n = 0;
    for (int i = 0; i < n_pages; ++i) n += kPage;
    return n - kPage + 100;  // a short last page
  }
};
int external_read(void* user, int64_t pos, void* dst, int32_t len) {
  const FileSource* src = static_cast<const FileSource*>(user);
  const int64_t n = std::min<int64_t>(len, src->length() - pos);
  for (int64_t k = 0; k < n; ++k) ((uint8_t*)dst)[k] = (uint8_t)((pos + k) * 3 + src->file);
  return (int)n;
}

apc_manager* make(const char* evictor, int cache_pages, int async_threads, int64_t timeout_ms) {
  apc_options o{};
  o.store.device = 0;
  o.store.page_size = kPage;
  o.store.cache_bytes = (uint64_t)cache_pages * kPage;
  o.store.slab_slots = 4096;
  o.store.flags = APS_FLAG_CRC;
  o.evictor = evictor;
  o.lfu_logbase = 2.0;
  o.async_write_threads = async_threads;
  o.timeout_ms = timeout_ms;
  o.timeout_threads = 8;
  apc_manager* m = nullptr;
  const int rc = apc_manager_create_ex(&o, &m);
  CHECK(rc == APS_OK && m, "apc_manager_create_ex: %d", rc);
  return m;
}

void settle_and_compare(apc_manager* m, int cache_pages, const char* what) {
  apc_wait_async(m);
  aps_stats_t st{};
  CHECK(aps_stats(apc_store(m), &st) == APS_OK, "aps_stats");
  CHECK((int64_t)st.pages == apc_pages(m), "%s: store holds %llu pages, meta store %lld", what, (unsigned long long)st.pages,
        (long long)apc_pages(m));
  CHECK((int64_t)st.bytes == apc_bytes(m), "%s: store holds %llu bytes, meta store %lld", what, (unsigned long long)st.bytes,
        (long long)apc_bytes(m));
  CHECK(apc_bytes(m) <= (int64_t)cache_pages * kPage, "%s: %lld bytes in a cache of %d", what, (long long)apc_bytes(m),
        cache_pages * kPage);
}

// threads hammer one manager with puts, gets and deletes of ids from a shared space
void mixed_operations(const char* evictor, int async_threads, int64_t timeout_ms, int threads, int ops) {
  const int cache_pages = 40, files = 6, pages = 16;
  apc_manager* m = make(evictor, cache_pages, async_threads, timeout_ms);
  if (!m) return;
  std::atomic<int64_t> hits{0}, puts_ok{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] {
      std::mt19937 rng(1000 + t);
      uint8_t buf[kPage], page[kPage];
      for (int k = 0; k < ops; ++k) {
        const int f = rng() % files, i = rng() % pages, len = page_len(f, i);
        // decimal ids map straight onto the key; the others go through the intern table (its mutex is part of the test)
        const std::string fid = f % 2 ? std::to_string(f) : "file-" + std::to_string(f) + "-3f2b8c1e";
        const int op = rng() % 10;
        if (op < 4) {
          fill(f, i, page, len);
          puts_ok += apc_put(m, fid.c_str(), i, page, len);
        } else if (op < 9) {
          const int off = rng() % 3 == 0 ? rng() % len : 0;
          const int32_t got = apc_get(m, fid.c_str(), i, off, len - off, buf, kPage, 0);
          CHECK(got == 0 || got == -1 || got == len - off, "get returned %d of %d", got, len - off);
          if (got > 0) {
            ++hits;
            for (int b = 0; b < got; ++b)
              if (buf[b] != page_byte(f, i, off + b)) {
                CHECK(false, "page (%d, %d) byte %d is not its own", f, i, off + b);
                break;
              }
          }
        } else {
          apc_delete(m, fid.c_str(), i);
        }
        if (k % 64 == 0) CHECK(apc_bytes(m) <= (int64_t)cache_pages * kPage + kPage, "budget exceeded: %lld", (long long)apc_bytes(m));
      }
    });
  for (auto& th : pool) th.join();
  char what[96];
  snprintf(what, sizeof(what), "%s async=%d timeout=%lld", evictor, async_threads, (long long)timeout_ms);
  settle_and_compare(m, cache_pages, what);
  CHECK(hits > 0 && puts_ok > 0, "%s: %lld hits, %lld puts", what, (long long)hits.load(), (long long)puts_ok.load());
  CHECK(apc_manager_close(m) == APS_OK, "close");
}

// every thread reads files through its own streams over ONE shared, small cache, per page and batched
void concurrent_streams(int threads) {
  const int cache_pages = 24;
  apc_manager* m = make("LRU", cache_pages, 0, 0);
  if (!m) return;
  std::vector<FileSource> sources;
  for (int f = 0; f < 4; ++f) sources.push_back(FileSource{f, 9 + f});
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] {
      std::mt19937 rng(77 + t);
      for (int round = 0; round < 12; ++round) {
        FileSource* src = &sources[rng() % sources.size()];
        const std::string fid = std::to_string(500 + src->file);
        apc_stream* s = apc_file_open(m, fid.c_str(), src->length(), external_read, src);
        CHECK(s != nullptr, "apc_file_open");
        if (!s) return;
        apc_stream_set_batched(s, (int)(rng() % 2));
        std::vector<uint8_t> buf(3 * kPage + 9);
        for (int k = 0; k < 30; ++k) {
          if (rng() % 3 == 0) apc_stream_seek(s, rng() % (src->length() + 1));
          const int64_t pos = apc_stream_get_pos(s);
          const int32_t want = rng() % (3 * kPage), off = rng() % 9;
          const int32_t got = apc_stream_read(s, buf.data(), (int32_t)buf.size(), off, want);
          const int64_t expect = want == 0 ? 0 : pos >= src->length() ? -1 : std::min<int64_t>(want, src->length() - pos);
          CHECK(got == expect, "stream read %d at %lld gave %d, expected %lld", want, (long long)pos, got, (long long)expect);
          for (int b = 0; b < got; ++b)
            if (buf[off + b] != (uint8_t)((pos + b) * 3 + src->file)) {
              CHECK(false, "file %d byte %lld wrong", src->file, (long long)(pos + b));
              break;
            }
        }
        apc_stream_close(s);
        apc_stream_destroy(s);
      }
    });
  for (auto& th : pool) th.join();
  settle_and_compare(m, cache_pages, "streams");
  CHECK(apc_manager_close(m) == APS_OK, "close");
}

// CacheManager.Factory: many threads ask at once; exactly one manager is created, late callers get it or NULL
void factory_race(int threads) {
  apc_options o{};
  o.store.page_size = kPage;
  o.store.cache_bytes = 16 * kPage;
  o.evictor = "LRU";
  o.lfu_logbase = 2.0;
  std::vector<apc_manager*> seen(threads, nullptr);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] {
      apc_manager* m = nullptr;
      CHECK(apc_factory_get(&o, &m) == APS_OK, "apc_factory_get");
      seen[t] = m;
    });
  for (auto& th : pool) th.join();
  apc_manager* one = nullptr;
  for (apc_manager* m : seen)
    if (m) {
      CHECK(!one || one == m, "two managers out of one factory");
      one = m;
    }
  CHECK(one != nullptr, "nobody got the manager");
  apc_factory_clear();
}

}  // namespace

int main(int argc, char** argv) {
  const int threads = argc > 1 ? atoi(argv[1]) : 8;
  const int ops = argc > 2 ? atoi(argv[2]) : 4000;
  for (const char* evictor : {"LRU", "LFU"}) {
    mixed_operations(evictor, 0, 0, threads, ops);   // synchronous puts
    mixed_operations(evictor, 4, 0, threads, ops);   // asynchronous write pool
  }
  mixed_operations("LRU", 0, 2000, threads, ops / 4);  // behind TimeBoundPageStore
  concurrent_streams(threads);
  for (int k = 0; k < 20; ++k) factory_race(threads);
  if (g_failures) {
    fprintf(stderr, "%d invariant(s) violated\n", g_failures.load());
    return 1;
  }
  printf("ok\n");
  return 0;
}
