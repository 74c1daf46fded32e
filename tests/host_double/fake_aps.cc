// TEST DOUBLE -- never part of the product, never loaded by anything under alluxio_b200/.
//
// A host-memory stand-in for the entry points of include/aps.h that the host layer (alluxio_b200/host/cache.cc:
// LocalCacheManager, evictors, LocalCacheFileInStream, TimeBoundPageStore, the sharded store's routing, the replay
// drivers) calls.  tests/host_double/build.py links it with cache.cc into tests/_build/libaps_host_double.so, and ONLY
// there, so that the host layer's own logic -- everything above the PageStore boundary -- can be compared with the
// oracle on a machine without a GPU (`pytest -m "not gpu"`, tests/test_host_logic_cpu.py).  It says nothing about the
// CUDA store: the same comparisons run against the real libaps.so in the `-m gpu` tests, and every kernel is tested there.
// libaps.so and libaps_host.so do not contain this file and have no CPU path (alluxio_b200/capi.py raises without them;
// aps_open fails without a device).
//
// Contract followed: include/aps.h, entry by entry (status codes, the length rule of aps_get, distinct keys in put and
// delete batches, n <= max_batch, fault injection on the single-page calls only, idempotent close).  What it does not
// model: CRC32C VERIFICATION (results carry the CRC32C of the bytes moved, nothing is checked against a recorded one), device / pinned / pageable distinctions (all memory is host memory), the batch size of
// aps_iterate's walk order (here: insertion-independent map order; the real store walks its hash table).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "../../include/aps.h"

namespace {

using KeyT = std::tuple<uint64_t, uint64_t, uint64_t>;
KeyT K(const aps_key& k) { return KeyT(k.file_hi, k.file_lo, k.page_index); }

// CRC-32C (RFC 3720) of the bytes a call moves, as the library reports it in aps_result.crc (SSE4.2 crc32 instruction).
__attribute__((target("sse4.2"))) uint32_t crc32c(const uint8_t* p, size_t n) {
  uint64_t c = 0xFFFFFFFFu;
  size_t i = 0;
  for (; i + 8 <= n; i += 8) {
    uint64_t w;
    memcpy(&w, p + i, 8);
    c = __builtin_ia32_crc32di(c, w);
  }
  for (; i < n; ++i) c = __builtin_ia32_crc32qi((uint32_t)c, p[i]);
  return ~(uint32_t)c;
}

uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

}  // namespace

struct aps_store {
  aps_config cfg{};
  std::mutex mu;
  std::map<KeyT, std::vector<uint8_t>> pages;
  uint32_t faults = 0;
  bool closed = false;
  std::vector<aps_store*> shards;  // after aps_shard_attach_local
};

namespace {

int check_open(aps_store* s) { return !s ? APS_EINVAL : s->closed ? APS_ECLOSED : APS_OK; }

// one request of a get: the rule of aps.h (copies min(pageLen - page_off, nbytes); page_off > pageLen is EINVAL)
aps_result get_one(aps_store* s, const aps_get_req& r) {
  auto it = s->pages.find(K(r.key));
  if (it == s->pages.end()) return aps_result{-APS_ENOTFOUND, 0};
  const std::vector<uint8_t>& page = it->second;
  if (r.page_off > page.size()) return aps_result{-APS_EINVAL, 0};
  const uint32_t n = (uint32_t)std::min<uint64_t>(page.size() - r.page_off, r.nbytes);
  if (n) memcpy(r.dst, page.data() + r.page_off, n);
  return aps_result{(int32_t)n, (s->cfg.flags & APS_FLAG_CRC) ? crc32c(page.data() + r.page_off, n) : 0u};
}

aps_result put_one(aps_store* s, const aps_put_req& r) {
  if (r.len > s->cfg.page_size) return aps_result{-APS_EINVAL, 0};
  auto it = s->pages.find(K(r.key));
  if (it == s->pages.end() && s->pages.size() >= s->cfg.slab_slots) return aps_result{-APS_ENOSPC, 0};
  const uint8_t* src = static_cast<const uint8_t*>(r.src);
  std::vector<uint8_t> bytes(src, src + r.len);
  const uint32_t crc = (s->cfg.flags & APS_FLAG_CRC) ? crc32c(bytes.data(), bytes.size()) : 0u;
  s->pages[K(r.key)] = std::move(bytes);
  return aps_result{(int32_t)r.len, crc};
}

template <typename KeyOf>
bool distinct(uint32_t n, KeyOf keyOf) {
  std::vector<KeyT> keys;
  for (uint32_t i = 0; i < n; ++i) keys.push_back(K(keyOf(i)));
  std::sort(keys.begin(), keys.end());
  return std::adjacent_find(keys.begin(), keys.end()) == keys.end();
}

}  // namespace

extern "C" {

APS_API int aps_open(const aps_config* cfg, aps_store** out) {
  if (!cfg || !out || cfg->page_size == 0 || cfg->page_size > (1u << 30)) return APS_EINVAL;
  aps_store* s = new aps_store();
  s->cfg = *cfg;
  if (!s->cfg.slab_slots) {
    const uint64_t need = (cfg->cache_bytes + cfg->page_size - 1) / cfg->page_size;
    s->cfg.slab_slots = std::max<uint64_t>(64, need + need / 8);
  }
  if (!s->cfg.max_batch) s->cfg.max_batch = 4096;
  {  // handles live as long as the process (a closed one is a tombstone, as in the library): keep them reachable
    static std::mutex mu;
    static std::vector<aps_store*>* all = new std::vector<aps_store*>();
    std::lock_guard<std::mutex> lk(mu);
    all->push_back(s);
  }
  *out = s;
  return APS_OK;
}

APS_API int aps_close(aps_store* s) {
  if (!s) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  s->closed = true;  // the handle stays valid as a tombstone, as in the library
  s->pages.clear();
  return APS_OK;
}

APS_API uint64_t aps_cache_size(const aps_store* s) { return s ? s->cfg.cache_bytes : 0; }

APS_API int aps_set_faults(aps_store* s, uint32_t fault_mask) {
  if (int rc = check_open(s)) return rc;
  s->faults = fault_mask;
  return APS_OK;
}

APS_API int aps_put(aps_store* s, const aps_key* key, const void* src, uint32_t len) {
  if (int rc = check_open(s)) return rc;
  if (!key || (!src && len)) return APS_EINVAL;
  if (len > s->cfg.page_size) return APS_EINVAL;
  if (s->faults & APS_FAULT_PUT) return APS_EFAULT;
  std::lock_guard<std::mutex> lk(s->mu);
  const aps_result r = put_one(s, aps_put_req{*key, len, 0, src});
  return r.len < 0 ? -r.len : APS_OK;
}

APS_API int aps_get(aps_store* s, const aps_key* key, uint32_t page_off, uint32_t bytes_to_read, void* dst, uint32_t dst_cap,
                    uint32_t* out_len, uint32_t* out_crc) {
  if (int rc = check_open(s)) return rc;
  if (out_len) *out_len = 0;
  if (out_crc) *out_crc = 0;
  if (!key) return APS_EINVAL;
  const uint32_t want = std::min(bytes_to_read, dst_cap);
  if (!dst && want) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  const aps_result r = get_one(s, aps_get_req{*key, page_off, std::min(want, s->cfg.page_size), dst});
  if (r.len < 0) return -r.len;
  if (out_len) *out_len = (uint32_t)r.len;
  if (out_crc) *out_crc = r.crc;
  return APS_OK;
}

APS_API int aps_delete(aps_store* s, const aps_key* key) {
  if (int rc = check_open(s)) return rc;
  if (!key) return APS_EINVAL;
  if (s->faults & APS_FAULT_DELETE) return APS_EFAULT;
  std::lock_guard<std::mutex> lk(s->mu);
  return s->pages.erase(K(*key)) ? APS_OK : APS_ENOTFOUND;
}

APS_API int aps_iterate(aps_store* s, aps_iter_cb cb, void* user) {
  if (int rc = check_open(s)) return rc;
  if (!cb) return APS_EINVAL;
  std::vector<std::pair<aps_key, uint32_t>> all;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    for (const auto& e : s->pages)
      all.push_back({aps_key{std::get<0>(e.first), std::get<1>(e.first), std::get<2>(e.first)}, (uint32_t)e.second.size()});
  }
  // The real store walks its hash table: callers must not depend on the order.  APS_DOUBLE_ITERATE=reverse lets the
  // CPU suite run every case under a second order (tests/test_host_logic_cpu.py does).
  const char* order = getenv("APS_DOUBLE_ITERATE");
  if (order && !strcmp(order, "reverse")) std::reverse(all.begin(), all.end());
  for (const auto& e : all)
    if (cb(user, &e.first, e.second)) break;
  return APS_OK;
}

APS_API int aps_get_batch(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream) {
  if (int rc = check_open(s)) return rc;
  if (n == 0) return APS_OK;
  if (!reqs || !res) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  for (uint32_t i = 0; i < n; ++i) res[i] = get_one(s, reqs[i]);
  return APS_OK;
}

APS_API int aps_put_batch(aps_store* s, uint32_t n, const aps_put_req* reqs, aps_result* res, aps_stream) {
  if (int rc = check_open(s)) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !reqs || !res) return APS_EINVAL;
  if (!distinct(n, [&](uint32_t i) { return reqs[i].key; })) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  for (uint32_t i = 0; i < n; ++i) res[i] = put_one(s, reqs[i]);
  return APS_OK;
}

APS_API int aps_delete_batch(aps_store* s, uint32_t n, const aps_key* keys, aps_result* res, aps_stream) {
  if (int rc = check_open(s)) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !keys || !res) return APS_EINVAL;
  if (!distinct(n, [&](uint32_t i) { return keys[i]; })) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  for (uint32_t i = 0; i < n; ++i) res[i] = aps_result{s->pages.erase(K(keys[i])) ? 0 : -APS_ENOTFOUND, 0};
  return APS_OK;
}

// (hash >> 32) mod n with the library's key hash (alluxio_b200/csrc/aps_device.cuh: hash_key, shard_of); the CPU test
// compares it with libaps.so's own aps_shard_of, which is host code and runs without a GPU.
APS_API uint32_t aps_shard_of(const aps_key* key, uint32_t n_shards) {
  if (!key || !n_shards) return 0;
  const uint64_t h = mix64(key->file_hi ^ mix64(key->file_lo ^ mix64(key->page_index + 0x9E3779B97F4A7C15ull)));
  return (uint32_t)((h >> 32) % n_shards);
}

APS_API int aps_shard_attach_local(aps_store* s, uint32_t n_shards, uint32_t my_shard, aps_store* const* stores) {
  if (int rc = check_open(s)) return rc;
  if (!stores || !n_shards || my_shard >= n_shards || stores[my_shard] != s) return APS_EINVAL;
  for (uint32_t i = 0; i < n_shards; ++i)
    if (!stores[i] || stores[i]->cfg.page_size != s->cfg.page_size) return APS_EINVAL;
  s->shards.assign(stores, stores + n_shards);
  return APS_OK;
}

APS_API int aps_get_batch_sharded(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream) {
  if (int rc = check_open(s)) return rc;
  if (s->shards.empty()) return APS_EINVAL;
  if (n == 0) return APS_OK;
  if (!reqs || !res) return APS_EINVAL;
  for (uint32_t i = 0; i < n; ++i) {
    aps_store* owner = s->shards[aps_shard_of(&reqs[i].key, (uint32_t)s->shards.size())];
    std::lock_guard<std::mutex> lk(owner->mu);
    res[i] = owner->closed ? aps_result{-APS_ECLOSED, 0} : get_one(owner, reqs[i]);
  }
  return APS_OK;
}

// SURVEY.md 8d payload: 64-bit word k of page (file, index) = splitmix64 output k for seed ^ (file << 32) ^ index.
APS_API int aps_fill_pages(aps_store* s, void* d_dst, uint64_t stride, uint32_t page_len, uint32_t n_pages,
                           const uint64_t* files, const uint64_t* indices, uint64_t seed, aps_stream) {
  if (int rc = check_open(s)) return rc;
  if (!d_dst || !files || !indices || n_pages > s->cfg.max_batch) return APS_EINVAL;
  for (uint32_t p = 0; p < n_pages; ++p) {
    uint8_t* out = static_cast<uint8_t*>(d_dst) + (uint64_t)p * stride;
    const uint64_t base = seed ^ (files[p] << 32) ^ indices[p];
    for (uint64_t k = 0; k * 8 < page_len; ++k) {
      const uint64_t w = mix64(base + (k + 1) * 0x9E3779B97F4A7C15ull);
      memcpy(out + 8 * k, &w, (size_t)std::min<uint64_t>(8, page_len - 8 * k));
    }
  }
  return APS_OK;
}

APS_API int aps_device_alloc(int32_t, uint64_t bytes, void** out) {
  if (!out) return APS_EINVAL;
  *out = malloc(bytes ? bytes : 1);
  return *out ? APS_OK : APS_ENOMEM;
}
APS_API int aps_device_free(int32_t, void* p) {
  free(p);
  return APS_OK;
}
APS_API int aps_host_alloc_pinned(uint64_t bytes, void** out) { return aps_device_alloc(0, bytes, out); }
APS_API int aps_host_free_pinned(void* p) {
  free(p);
  return APS_OK;
}

APS_API int aps_stats(aps_store* s, aps_stats_t* out) {
  if (int rc = check_open(s)) return rc;
  if (!out) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  memset(out, 0, sizeof(*out));
  out->pages = s->pages.size();
  for (const auto& e : s->pages) out->bytes += e.second.size();
  out->slots_free = s->cfg.slab_slots - s->pages.size();
  return APS_OK;  // no table, no kernels: everything else stays zero
}

// Snapshots are files written by the CUDA library (aps_snapshot.cu); the double has none.
APS_API int aps_snapshot_save_names(aps_store*, const char*, uint32_t, const char* const*, uint64_t*) { return APS_EIO; }
APS_API int aps_snapshot_names(const char*, aps_name_cb, void*) { return APS_EIO; }
APS_API int aps_snapshot_load(aps_store*, const char*, uint64_t*, uint64_t*) { return APS_EIO; }

APS_API const char* aps_last_cuda_error(void) { return ""; }

APS_API const char* aps_strerror(int code) {
  static const char* const names[] = {"ok", "page not found", "invalid argument", "no space", "CUDA error", "corrupt page",
                                      "store closed", "out of memory", "injected fault", "limit", "I/O error", "timeout"};
  return code >= 0 && code < (int)(sizeof(names) / sizeof(names[0])) ? names[code] : "unknown";
}

}  // extern "C"
