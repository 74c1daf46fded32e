// Host-side cache layer (see cache.h).  Product code: links libaps.so, never the oracle.
#include "cache.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>

#include "../../include/aps_cache.h"

namespace alluxio_b200 {

// ================================================================================ Java-isms, keys

int32_t JavaStringHashCode(const std::string& s) {
  // 31-polynomial over UTF-16 code units with int32 wrap-around; the input is UTF-8.
  uint32_t h = 0;
  for (size_t i = 0; i < s.size();) {
    const uint8_t c = (uint8_t)s[i];
    uint32_t cp;
    int more;
    if (c < 0x80) { cp = c; more = 0; }
    else if (c < 0xE0) { cp = c & 0x1Fu; more = 1; }
    else if (c < 0xF0) { cp = c & 0x0Fu; more = 2; }
    else { cp = c & 0x07u; more = 3; }
    ++i;
    while (more-- > 0 && i < s.size()) cp = (cp << 6) | ((uint8_t)s[i++] & 0x3Fu);
    if (cp > 0xFFFF) {
      const uint32_t v = cp - 0x10000;
      h = h * 31u + (0xD800u | (v >> 10));
      h = h * 31u + (0xDC00u | (v & 0x3FFu));
    } else {
      h = h * 31u + cp;
    }
  }
  return (int32_t)h;
}

static int32_t FloorMod(int32_t x, int32_t y) {
  const int32_t r = x % y;
  return (r != 0 && ((r < 0) != (y < 0))) ? r + y : r;
}

static bool IsCanonicalDecimalLong(const std::string& s, int64_t* out) {
  if (s.empty() || s.size() > 20) return false;
  size_t i = 0;
  const bool neg = s[0] == '-';
  if (neg) i = 1;
  if (i >= s.size()) return false;
  if (s[i] == '0') {
    if (neg || s.size() != 1) return false;
    *out = 0;
    return true;
  }
  unsigned long long v = 0;
  const unsigned long long limit = neg ? (1ull << 63) : (1ull << 63) - 1;
  for (; i < s.size(); ++i) {
    if (s[i] < '0' || s[i] > '9') return false;
    const unsigned d = (unsigned)(s[i] - '0');
    if (v > (limit - d) / 10) return false;
    v = v * 10 + d;
  }
  *out = neg ? (int64_t)(0ull - v) : (int64_t)v;
  return true;
}

static bool ParseHex32(const std::string& s, uint64_t* hi, uint64_t* lo) {
  if (s.size() != 32) return false;
  uint64_t w[2] = {0, 0};
  for (int i = 0; i < 32; ++i) {
    const char c = s[(size_t)i];
    unsigned d;
    if (c >= '0' && c <= '9') d = (unsigned)(c - '0');
    else if (c >= 'a' && c <= 'f') d = (unsigned)(c - 'a' + 10);
    else return false;
    w[i / 16] = (w[i / 16] << 4) | d;
  }
  *hi = w[0];
  *lo = w[1];
  return true;
}

void FileKeyCodec::Encode(const std::string& fileId, uint64_t* hi, uint64_t* lo) {
  int64_t v;
  if (IsCanonicalDecimalLong(fileId, &v)) {
    *hi = 0;
    *lo = (uint64_t)v;
    return;
  }
  uint64_t h, l;
  if (ParseHex32(fileId, &h, &l) && h != 0 && h != ~0ull) {
    *hi = h;
    *lo = l;
    return;
  }
  std::lock_guard<std::mutex> g(mMu);
  auto it = mIds.find(fileId);
  if (it == mIds.end()) {
    it = mIds.emplace(fileId, (uint64_t)mNames.size()).first;
    mNames.push_back(fileId);
  }
  *hi = ~0ull;
  *lo = it->second;
}

std::vector<std::string> FileKeyCodec::Names() {
  std::lock_guard<std::mutex> g(mMu);
  return mNames;
}

void FileKeyCodec::AdoptNames(const std::vector<std::string>& names) {
  std::lock_guard<std::mutex> g(mMu);
  for (size_t i = mNames.size(); i < names.size(); ++i) {
    mIds.emplace(names[i], (uint64_t)i);
    mNames.push_back(names[i]);
  }
}

std::string FileKeyCodec::Decode(uint64_t hi, uint64_t lo) {
  if (hi == 0) return std::to_string((int64_t)lo);
  if (hi == ~0ull) {
    std::lock_guard<std::mutex> g(mMu);
    return lo < mNames.size() ? mNames[(size_t)lo] : std::string();
  }
  char buf[33];
  snprintf(buf, sizeof(buf), "%016llx%016llx", (unsigned long long)hi, (unsigned long long)lo);
  return buf;
}

// ---- HadoopUtils.toAlluxioUriStatus (core/client/hdfs/src/main/java/alluxio/hadoop/HadoopUtils.java:225-238):
// behind the Hadoop-compatible file system (how Presto reaches the cache) a file's identifier is the lower-case hex
// MD5 of its path string, which FileKeyCodec maps onto the 128-bit key exactly.  RFC 1321, one-shot, host side.
namespace {
inline uint32_t Rotl32(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }
void Md5(const uint8_t* data, size_t len, uint8_t out[16]) {
  static const int kShift[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9,  14, 20, 5, 9,
                                 14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
  uint32_t k[64];
  for (int i = 0; i < 64; ++i) k[i] = (uint32_t)(int64_t)std::floor(std::fabs(std::sin((double)(i + 1))) * 4294967296.0);
  uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
  std::vector<uint8_t> msg(data, data + len);
  msg.push_back(0x80);
  while (msg.size() % 64 != 56) msg.push_back(0);
  const uint64_t bits = (uint64_t)len * 8;
  for (int i = 0; i < 8; ++i) msg.push_back((uint8_t)(bits >> (8 * i)));
  for (size_t off = 0; off < msg.size(); off += 64) {
    uint32_t m[16];
    memcpy(m, msg.data() + off, 64);  // little endian host
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    for (int i = 0; i < 64; ++i) {
      uint32_t f;
      int g;
      if (i < 16) { f = (b & c) | (~b & d); g = i; }
      else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
      else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
      else { f = c ^ (b | ~d); g = (7 * i) & 15; }
      const uint32_t t = d;
      d = c;
      c = b;
      b = b + Rotl32(a + f + k[i] + m[g], kShift[i]);
      a = t;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
  }
  memcpy(out, h, 16);
}
}  // namespace

std::string HadoopFileIdentifier(const std::string& pathUtf8) {
  uint8_t d[16];
  Md5((const uint8_t*)pathUtf8.data(), pathUtf8.size(), d);
  char buf[33];
  for (int i = 0; i < 16; ++i) snprintf(buf + 2 * i, 3, "%02x", d[i]);
  return buf;
}

// ================================================================================ node pool

int32_t NodePool::Alloc(const Key& k) {
  int32_t n;
  if (!mFree.empty()) {
    n = mFree.back();
    mFree.pop_back();
  } else {
    n = (int32_t)mNodes.size();
    mNodes.push_back(ListNode());
  }
  mNodes[(size_t)n] = ListNode{k, -1, -1};
  return n;
}
void NodePool::Free(int32_t n) { mFree.push_back(n); }

void ListPushBack(NodePool& p, ListHead& l, int32_t n) {
  p[n].prev = l.tail;
  p[n].next = -1;
  if (l.tail >= 0) p[l.tail].next = n; else l.head = n;
  l.tail = n;
}
void ListUnlink(NodePool& p, ListHead& l, int32_t n) {
  const int32_t a = p[n].prev, b = p[n].next;
  if (a >= 0) p[a].next = b; else l.head = b;
  if (b >= 0) p[b].prev = a; else l.tail = a;
  p[n].prev = p[n].next = -1;
}

// ================================================================================ evictors

namespace {

// CACHE/evictor/LRUCacheEvictor.java:29-75: an access-ordered LinkedHashMap; put and get both move the
// id to the tail, evict() names the head without removing it.  `refreshOnPut` = false gives the FIFO
// double of the reference's tests (CTEST/LocalCacheManagerTest.java:731-761).
class RecencyEvictor : public CacheEvictor {
 public:
  RecencyEvictor(bool refreshOnGet, bool refreshOnPut) : mOnGet(refreshOnGet), mOnPut(refreshOnPut) {}
  void UpdateOnGet(const Key& k) override {
    if (mOnGet) Touch(k, false);
  }
  void UpdateOnPut(const Key& k) override { Touch(k, /*keepPlaceIfPresent=*/!mOnPut); }
  void UpdateOnDelete(const Key& k) override {
    std::lock_guard<std::mutex> g(mMu);
    const int32_t* at = mWhere.Find(k);
    if (!at) return;
    const int32_t n = *at;
    ListUnlink(mPool, mList, n);
    mPool.Free(n);
    mWhere.Erase(k);
  }
  bool Evict(Key* victim) override {
    std::lock_guard<std::mutex> g(mMu);
    if (mList.Empty()) return false;
    *victim = mPool[mList.head].key;
    return true;
  }
  void Reset() override {
    std::lock_guard<std::mutex> g(mMu);
    mWhere.Clear();
    mPool = NodePool();
    mList = ListHead();
  }
 private:
  void Touch(const Key& k, bool keepPlaceIfPresent) {
    std::lock_guard<std::mutex> g(mMu);
    bool inserted;
    int32_t* at = mWhere.Emplace(k, -1, &inserted);
    if (inserted) {
      *at = mPool.Alloc(k);
    } else {
      if (keepPlaceIfPresent) return;
      ListUnlink(mPool, mList, *at);
    }
    ListPushBack(mPool, mList, *at);
  }
  const bool mOnGet, mOnPut;
  std::mutex mMu;
  NodePool mPool;
  ListHead mList;
  FlatMap<int32_t> mWhere;
};

// CACHE/evictor/LFUCacheEvictor.java:35-161: ids grouped by bucket = (int)(ln(count)/ln(base)), LRU
// inside a bucket, the victim is the head of the lowest occupied bucket as tracked by mMinBucket.
class LFUEvictor : public CacheEvictor {
 public:
  explicit LFUEvictor(double logBase) : mDivisor(std::log(logBase)) {}
  void UpdateOnGet(const Key& k) override {
    std::lock_guard<std::mutex> g(mMu);
    bool inserted;
    Entry& e = *mEntries.Emplace(k, Entry(), &inserted);  // count starts at 0
    const int newCount = ++e.count;
    const int newBucket = Bucket(newCount);
    if (newCount > 1) {
      const int oldBucket = Bucket(newCount - 1);
      if (newBucket != oldBucket) {                          // :103-109
        const bool left = Detach(e.node, oldBucket);
        if (!left && oldBucket == mMinBucket) mMinBucket = newBucket;
        Attach(e.node, newBucket);
      } else {                                               // :110-112 touch = move to the bucket's tail
        auto it = mBuckets.find(newBucket);
        if (it != mBuckets.end()) {
          ListUnlink(mPool, it->second, e.node);
          ListPushBack(mPool, it->second, e.node);
        }
      }
    } else {                                                 // :113-116
      e.node = mPool.Alloc(k);
      mMinBucket = newBucket;
      Attach(e.node, newBucket);
    }
  }
  void UpdateOnPut(const Key& k) override { UpdateOnGet(k); }
  void UpdateOnDelete(const Key& k) override {
    std::lock_guard<std::mutex> g(mMu);
    const Entry* e = mEntries.Find(k);
    if (!e) return;
    const int bucket = Bucket(e->count);
    const int32_t node = e->node;
    mEntries.Erase(k);
    const bool left = Detach(node, bucket);
    mPool.Free(node);
    if (!left && bucket == mMinBucket && !mBuckets.empty())
      while (!mBuckets.count(mMinBucket)) ++mMinBucket;      // :133-137
  }
  bool Evict(Key* victim) override {
    std::lock_guard<std::mutex> g(mMu);
    auto it = mBuckets.find(mMinBucket);
    if (it == mBuckets.end()) return false;
    *victim = mPool[it->second.head].key;
    return true;
  }
  void Reset() override {
    std::lock_guard<std::mutex> g(mMu);
    mEntries.Clear();
    mBuckets.clear();
    mPool = NodePool();
    mMinBucket = -1;
  }
 private:
  struct Entry { int count = 0; int32_t node = -1; };
  int Bucket(int count) const { return (int)(std::log((double)count) / mDivisor); }
  void Attach(int32_t node, int bucket) { ListPushBack(mPool, mBuckets[bucket], node); }
  // true if the bucket still exists afterwards (computeIfPresent semantics, :76-85)
  bool Detach(int32_t node, int bucket) {
    auto it = mBuckets.find(bucket);
    if (it == mBuckets.end()) return false;
    ListUnlink(mPool, it->second, node);
    if (it->second.Empty()) {
      mBuckets.erase(it);
      return false;
    }
    return true;
  }
  std::mutex mMu;
  const double mDivisor;
  NodePool mPool;
  FlatMap<Entry> mEntries;
  std::map<int, ListHead> mBuckets;
  int mMinBucket = -1;
};

}  // namespace

std::unique_ptr<CacheEvictor> CacheEvictor::Create(const std::string& kind, double lfuLogBase) {
  if (kind == "LRU") return std::unique_ptr<CacheEvictor>(new RecencyEvictor(true, true));
  if (kind == "FIFO") return std::unique_ptr<CacheEvictor>(new RecencyEvictor(false, false));
  if (kind == "LFU") return std::unique_ptr<CacheEvictor>(new LFUEvictor(lfuLogBase));
  return nullptr;
}

// ================================================================================ meta store

void MetaStore::AddPage(const Key& k, const PageMeta& m) {  // DefaultMetaStore.java:56-63
  bool inserted;
  *mPages.Emplace(k, m, &inserted) = m;
  mBytes += m.size;
  mCount += 1;
  mEvictor->UpdateOnPut(k);
}
bool MetaStore::GetPageInfo(const Key& k, PageMeta* out) {  // :66-72
  const PageMeta* pm = mPages.Find(k);
  if (!pm) return false;
  *out = *pm;
  mEvictor->UpdateOnGet(k);
  return true;
}
bool MetaStore::RemovePage(const Key& k) {  // :75-85
  const PageMeta* pm = mPages.Find(k);
  if (!pm) return false;
  mBytes -= pm->size;
  mCount -= 1;
  mPages.Erase(k);
  mEvictor->UpdateOnDelete(k);
  return true;
}
void MetaStore::Reset() {  // :98-105
  mPages.Clear();
  mBytes = 0;
  mCount = 0;
  mEvictor->Reset();
}
bool MetaStore::Evict(Key* victim, PageMeta* meta) {  // :108-122
  if (!mEvictor->Evict(victim)) return false;
  const PageMeta* pm = mPages.Find(*victim);
  if (!pm) {
    mEvictor->UpdateOnDelete(*victim);  // heal an evictor that named an unknown page
    return false;
  }
  *meta = *pm;
  return true;
}

// ================================================================================ GpuPageStore

int GpuPageStore::Open(const aps_config& cfg, std::unique_ptr<GpuPageStore>* out) {
  aps_store* s = nullptr;
  const int rc = aps_open(&cfg, &s);
  if (rc != APS_OK) return rc;
  std::unique_ptr<GpuPageStore> g(new GpuPageStore());
  g->mStore = s;
  g->mCacheBytes = cfg.cache_bytes;
  g->mDevice = cfg.device;
  *out = std::move(g);
  return APS_OK;
}
std::unique_ptr<GpuPageStore> GpuPageStore::Adopt(aps_store* s) {
  std::unique_ptr<GpuPageStore> g(new GpuPageStore());
  g->mStore = s;
  g->mCacheBytes = aps_cache_size(s);
  g->mOwned = false;
  return g;
}
GpuPageStore::~GpuPageStore() {
  if (mStore && mOwned) aps_close(mStore);
}
static aps_key ToAps(const Key& k) { return aps_key{k.hi, k.lo, k.idx}; }
int GpuPageStore::Put(const Key& k, const void* page, int32_t len) {
  const aps_key a = ToAps(k);
  return aps_put(mStore, &a, page, (uint32_t)len);
}
int GpuPageStore::Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) {
  *bytesRead = 0;
  if (pageOffset < 0) return APS_EINVAL;  // LocalPageStore.java:91
  if (room < 0) return APS_EINVAL;        // :92-93 (bufferOffset beyond the buffer)
  const aps_key a = ToAps(k);
  uint32_t got = 0;
  const int rc = aps_get(mStore, &a, (uint32_t)pageOffset, (uint32_t)std::max(bytesToRead, 0), dst, (uint32_t)room, &got, nullptr);
  *bytesRead = (int32_t)got;
  return rc;
}
int GpuPageStore::Delete(const Key& k) {
  const aps_key a = ToAps(k);
  return aps_delete(mStore, &a);
}
int GpuPageStore::GetPages(std::vector<std::pair<Key, int64_t>>* out) {
  out->clear();
  return aps_iterate(
      mStore,
      [](void* user, const aps_key* key, uint32_t len) {
        static_cast<std::vector<std::pair<Key, int64_t>>*>(user)->push_back({Key{key->file_hi, key->file_lo, key->page_index}, (int64_t)len});
        return 0;
      },
      out);
}

// ================================================================================ ShardedGpuPageStore

int ShardedGpuPageStore::Open(const aps_config& cfg, uint32_t nShards, const int32_t* devices,
                              std::unique_ptr<ShardedGpuPageStore>* out) {
  if (nShards == 0 || nShards > 64 || !devices || cfg.page_size == 0) return APS_EINVAL;
  std::unique_ptr<ShardedGpuPageStore> g(new ShardedGpuPageStore());
  g->mCacheBytes = cfg.cache_bytes;
  const uint64_t pages = (cfg.cache_bytes + cfg.page_size - 1) / cfg.page_size;
  const uint64_t share = (pages + nShards - 1) / nShards;
  for (uint32_t i = 0; i < nShards; ++i) {
    aps_config c = cfg;
    c.device = devices[i];
    c.slab_slots = cfg.slab_slots ? cfg.slab_slots : share + share / 4 + 64;
    c.cache_bytes = c.slab_slots * (uint64_t)cfg.page_size;
    std::unique_ptr<GpuPageStore> s;
    const int rc = GpuPageStore::Open(c, &s);
    if (rc != APS_OK) return rc;  // shards opened so far close with g
    g->mShards.push_back(std::move(s));
  }
  std::vector<aps_store*> handles;
  for (auto& s : g->mShards) handles.push_back(s->handle());
  for (uint32_t i = 0; i < nShards; ++i) {  // every shard can serve a sharded get; shard 0 is the one used
    const int rc = aps_shard_attach_local(handles[i], nShards, i, handles.data());
    if (rc != APS_OK) return rc;
  }
  *out = std::move(g);
  return APS_OK;
}
ShardedGpuPageStore::~ShardedGpuPageStore() {}
uint32_t ShardedGpuPageStore::OwnerOf(const Key& k) const {
  const aps_key a = ToAps(k);
  return aps_shard_of(&a, (uint32_t)mShards.size());
}
int ShardedGpuPageStore::Put(const Key& k, const void* page, int32_t len) { return mShards[OwnerOf(k)]->Put(k, page, len); }
int ShardedGpuPageStore::Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) {
  return mShards[OwnerOf(k)]->Get(k, pageOffset, bytesToRead, dst, room, bytesRead);
}
int ShardedGpuPageStore::Delete(const Key& k) { return mShards[OwnerOf(k)]->Delete(k); }
int ShardedGpuPageStore::GetPages(std::vector<std::pair<Key, int64_t>>* out) {
  out->clear();
  std::vector<std::pair<Key, int64_t>> part;
  for (auto& s : mShards) {
    const int rc = s->GetPages(&part);
    if (rc != APS_OK) return rc;
    out->insert(out->end(), part.begin(), part.end());
  }
  return APS_OK;
}
int ShardedGpuPageStore::GetBatch(uint32_t n, const aps_get_req* reqs, aps_result* res) {
  return aps_get_batch_sharded(mShards[0]->handle(), n, reqs, res, nullptr);
}
namespace {
// Splits a batch by owner, runs `call(shard, count, sub-requests, sub-results)` per shard, scatters the results back.
template <typename Req, typename KeyOfReq, typename Call>
int ByOwner(uint32_t nShards, uint32_t n, const Req* reqs, aps_result* res, KeyOfReq keyOf, Call call) {
  std::vector<std::vector<Req>> sub(nShards);
  std::vector<std::vector<uint32_t>> where(nShards);
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t o = aps_shard_of(&keyOf(reqs[i]), nShards);
    sub[o].push_back(reqs[i]);
    where[o].push_back(i);
  }
  std::vector<aps_result> part;
  for (uint32_t o = 0; o < nShards; ++o) {
    if (sub[o].empty()) continue;
    part.resize(sub[o].size());
    const int rc = call(o, (uint32_t)sub[o].size(), sub[o].data(), part.data());
    if (rc != APS_OK) return rc;
    for (size_t j = 0; j < part.size(); ++j) res[where[o][j]] = part[j];
  }
  return APS_OK;
}
}  // namespace
int ShardedGpuPageStore::PutBatch(uint32_t n, const aps_put_req* reqs, aps_result* res) {
  return ByOwner(
      shards(), n, reqs, res, [](const aps_put_req& r) -> const aps_key& { return r.key; },
      [this](uint32_t o, uint32_t m, const aps_put_req* r, aps_result* out) { return mShards[o]->PutBatch(m, r, out); });
}
int ShardedGpuPageStore::DeleteBatch(uint32_t n, const aps_key* keys, aps_result* res) {
  return ByOwner(
      shards(), n, keys, res, [](const aps_key& k) -> const aps_key& { return k; },
      [this](uint32_t o, uint32_t m, const aps_key* k, aps_result* out) { return mShards[o]->DeleteBatch(m, k, out); });
}
int ShardedGpuPageStore::FillPages(void* d, uint64_t stride, uint32_t len, uint32_t n, const uint64_t* f, const uint64_t* i,
                                   uint64_t seed) {
  return mShards[0]->FillPages(d, stride, len, n, f, i, seed);
}
int ShardedGpuPageStore::SetFaults(uint32_t mask) {
  for (auto& s : mShards) {
    const int rc = s->SetFaults(mask);
    if (rc != APS_OK) return rc;
  }
  return APS_OK;
}

// ================================================================================ LocalCacheManager

HandOffPool::HandOffPool(int threads) {
  for (int i = 0; i < threads; ++i) mWorkers.emplace_back([this] { Work(); });
}
HandOffPool::~HandOffPool() {
  {
    std::lock_guard<std::mutex> lk(mMu);
    mStop = true;
  }
  mWake.notify_all();
  for (auto& t : mWorkers) t.join();
}
bool HandOffPool::TrySubmit(std::function<void()> task) {
  {
    std::lock_guard<std::mutex> lk(mMu);
    // every thread is running a task or already has one waiting for it: RejectedExecutionException
    if (mStop || mQueue.size() + (size_t)mRunning >= mWorkers.size()) return false;
    mQueue.push_back(std::move(task));
  }
  mWake.notify_one();
  return true;
}
void HandOffPool::WaitIdle() {
  std::unique_lock<std::mutex> lk(mMu);
  mQuiet.wait(lk, [this] { return mQueue.empty() && mRunning == 0; });
}
void HandOffPool::Work() {
  std::unique_lock<std::mutex> lk(mMu);
  for (;;) {
    mWake.wait(lk, [this] { return mStop || !mQueue.empty(); });
    if (mQueue.empty()) return;  // stopping
    std::function<void()> task = std::move(mQueue.front());
    mQueue.pop_front();
    ++mRunning;
    lk.unlock();
    task();
    lk.lock();
    --mRunning;
    if (mQueue.empty() && mRunning == 0) mQuiet.notify_all();
  }
}

// ================================================================================ TimeBoundPageStore

TimeBoundPageStore::TimeBoundPageStore(std::unique_ptr<PageStore> store, int64_t timeoutMs, int threads)
    : mStore(std::move(store)), mTimeoutMs(timeoutMs), mPool(new HandOffPool(std::max(1, threads))) {}

int TimeBoundPageStore::Bounded(std::function<int()> call, std::atomic<int64_t>* timeouts) {
  struct Outcome {
    std::mutex mu;
    std::condition_variable cv;
    bool done = false;
    int rc = APS_OK;
  };
  auto out = std::make_shared<Outcome>();
  if (!mPool->TrySubmit([out, call] {
        const int rc = call();
        std::lock_guard<std::mutex> lk(out->mu);
        out->rc = rc;
        out->done = true;
        out->cv.notify_all();
      })) {
    storeThreadsRejected++;
    return APS_ETIMEOUT;
  }
  std::unique_lock<std::mutex> lk(out->mu);
  if (!out->cv.wait_for(lk, std::chrono::milliseconds(mTimeoutMs), [&] { return out->done; })) {
    (*timeouts)++;
    return APS_ETIMEOUT;
  }
  return out->rc;
}

static void Nap(int32_t ms) {
  if (ms > 0) std::this_thread::sleep_for(std::chrono::milliseconds(ms));
}

int TimeBoundPageStore::Put(const Key& k, const void* page, int32_t len) {
  auto bytes = std::make_shared<std::vector<uint8_t>>((const uint8_t*)page, (const uint8_t*)page + std::max(len, 0));
  auto store = mStore;
  const int32_t delay = putDelayMs.load();
  return Bounded([store, k, bytes, len, delay] {
    Nap(delay);
    return store->Put(k, bytes->data(), len);
  }, &storePutTimeout);
}
int TimeBoundPageStore::Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) {
  *bytesRead = 0;
  if (room < 0) return APS_EINVAL;
  struct Landing {
    std::vector<uint8_t> bytes;
    int32_t got = 0;
  };
  auto land = std::make_shared<Landing>();
  land->bytes.resize((size_t)std::max(0, std::min(room, std::max(bytesToRead, 0))));
  auto store = mStore;
  const int32_t delay = getDelayMs.load();
  const int rc = Bounded([store, k, pageOffset, bytesToRead, land, delay] {
    Nap(delay);
    return store->Get(k, pageOffset, bytesToRead, land->bytes.data(), (int32_t)land->bytes.size(), &land->got);
  }, &storeGetTimeout);
  if (rc == APS_OK) {  // only a call that finished in time touches the caller's buffer
    memcpy(dst, land->bytes.data(), (size_t)land->got);
    *bytesRead = land->got;
  }
  return rc;
}
int TimeBoundPageStore::Delete(const Key& k) {
  auto store = mStore;
  const int32_t delay = deleteDelayMs.load();
  return Bounded([store, k, delay] {
    Nap(delay);
    return store->Delete(k);
  }, &storeDeleteTimeout);
}

LocalCacheManager::LocalCacheManager(int64_t pageSize, std::unique_ptr<MetaStore> meta, std::unique_ptr<PageStore> store,
                                     ManagerOptions options)
    : mOptions(options), mPageSize(pageSize), mCacheSize(store->GetCacheSize()), mMetaStore(std::move(meta)),
      mPageStore(std::move(store)), mState(READ_ONLY) {  // LocalCacheManager.java:145-166
  if (mOptions.asyncWriteThreads > 0) mAsyncCacheExecutor.reset(new HandOffPool(mOptions.asyncWriteThreads));
  mCodec.AdoptNames(mOptions.internedNames);  // before restore decodes the ids the store lists
  if (mOptions.asyncRestore) mInitService = std::thread([this] { RestoreOrInit(); });  // :126-133, READ_ONLY meanwhile
  else RestoreOrInit();                                                                // :134
}

LocalCacheManager::~LocalCacheManager() {
  if (mInitService.joinable()) mInitService.join();
  mAsyncCacheExecutor.reset();
}

void LocalCacheManager::WaitAsync() {
  if (mInitService.joinable()) mInitService.join();
  if (mAsyncCacheExecutor) mAsyncCacheExecutor->WaitIdle();
}

Key LocalCacheManager::MakeKey(const std::string& fileId, int64_t pageIndex) {
  Key k;
  mCodec.Encode(fileId, &k.hi, &k.lo);
  k.idx = (uint64_t)pageIndex;
  return k;
}

int32_t LocalCacheManager::GetPageLockId(const std::string& fileId, int64_t pageIndex) const {  // :172-174
  const uint64_t sum = (uint64_t)(int64_t)JavaStringHashCode(fileId) + (uint64_t)pageIndex;  // long add, wraps
  return FloorMod((int32_t)(uint32_t)sum, kLockSize);                                        // (int) narrowing
}

bool LocalCacheManager::Put(const std::string& fileId, int64_t pageIndex, const void* page, int32_t len) {  // :206-246
  if (mState.load() != READ_WRITE) {
    putNotReadyErrors++;
    metrics.putErrors++;
    return false;
  }
  const Key k = MakeKey(fileId, pageIndex);
  const int32_t lockId = GetPageLockId(fileId, pageIndex);
  if (!mAsyncCacheExecutor) {
    const bool ok = PutInternal(k, lockId, page, len);
    if (!ok) metrics.putErrors++;
    return ok;
  }
  {
    std::lock_guard<std::mutex> lk(mPendingMu);
    if (!mPendingRequests.insert(k).second) return false;  // already queued (:221-223)
  }
  // the Java caller hands over its array; a C caller keeps its buffer, so the bytes are copied for the worker
  auto bytes = std::make_shared<std::vector<uint8_t>>((const uint8_t*)page, (const uint8_t*)page + len);
  const bool accepted = mAsyncCacheExecutor->TrySubmit([this, k, lockId, bytes] {
    const bool ok = PutInternal(k, lockId, bytes->data(), (int32_t)bytes->size());
    if (!ok) metrics.putErrors++;
    std::lock_guard<std::mutex> lk(mPendingMu);
    mPendingRequests.erase(k);
  });
  if (!accepted) {  // every worker busy (:235-243)
    {
      std::lock_guard<std::mutex> lk(mPendingMu);
      mPendingRequests.erase(k);
    }
    putAsyncRejectionErrors++;
    metrics.putErrors++;
    return false;
  }
  return true;
}

bool LocalCacheManager::PutInternal(const Key& k, int32_t lockId, const void* page, int32_t len) {  // :249-347
  Key victim{};
  PageMeta victimMeta{};
  bool enoughSpace;
  {
    std::unique_lock<std::shared_mutex> pageLock(mPageLocks[lockId]);
    {
      std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
      if (mMetaStore->HasPage(k)) return true;  // :256-260 an existing page is NOT overwritten
      enoughSpace = mMetaStore->Bytes() + len <= mCacheSize;
      if (enoughSpace) {
        mMetaStore->AddPage(k, PageMeta{len, lockId});
      } else if (!mMetaStore->Evict(&victim, &victimMeta)) {
        metrics.putEvictionErrors++;
        return false;  // :265-271
      }
    }
    if (enoughSpace) {  // :274-285
      if (mPageStore->Put(k, page, len) == APS_OK) {
        metrics.bytesWrittenCache += len;  // :277 -- marked by the store write, so not for a page that was already there
        return true;
      }
      UndoAddPage(k);
      metrics.putStoreWriteErrors++;
      return false;
    }
  }
  // two page locks in id order (:197-203), then the two-phase commit (:288-346)
  const int32_t a = std::min(lockId, victimMeta.lockId), b = std::max(lockId, victimMeta.lockId);
  std::unique_lock<std::shared_mutex> first(mPageLocks[a]);
  std::unique_lock<std::shared_mutex> second;
  if (b != a) second = std::unique_lock<std::shared_mutex>(mPageLocks[b]);
  {
    std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
    if (mMetaStore->HasPage(k)) return true;  // :296-300 a racing thread inserted it
    if (!mMetaStore->RemovePage(victim)) {     // :301-309 benign race: victim already gone
      if (!mMetaStore->RemovePage(k)) metrics.cleanupPutErrors++;  // undoAddPage (:349-357)
      metrics.putBenignRacingErrors++;
      return false;
    }
    enoughSpace = mMetaStore->Bytes() + len <= mCacheSize;                // :310
    if (enoughSpace) mMetaStore->AddPage(k, PageMeta{len, lockId});      // :311-313
  }
  if (mPageStore->Delete(victim) != APS_OK) {  // :318-330 the victim is deleted whatever enoughSpace says
    if (enoughSpace) UndoAddPage(k);
    metrics.putStoreDeleteErrors++;
    return false;
  }
  if (logEvictions) evicted.push_back(victim);
  metrics.bytesEvicted += victimMeta.size;
  metrics.pagesEvicted++;
  if (!enoughSpace) {  // :331-334 one victim per put
    metrics.putEvictionErrors++;
    return false;
  }
  if (mPageStore->Put(k, page, len) == APS_OK) {
    metrics.bytesWrittenCache += len;  // :337
    return true;
  }
  UndoAddPage(k);  // :339-344
  metrics.putStoreWriteErrors++;
  return false;
}

void LocalCacheManager::UndoAddPage(const Key& k) {  // :349-357
  std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
  if (!mMetaStore->RemovePage(k)) metrics.cleanupPutErrors++;
}

int32_t LocalCacheManager::Get(const std::string& fileId, int64_t pageIndex, int32_t pageOffset, int32_t bytesToRead,
                               void* buffer, int32_t bufLen, int32_t offsetInBuffer) {  // :360-399
  if (!(pageOffset <= mPageSize)) return kGetIllegalArgument;
  if (!(bytesToRead <= bufLen - offsetInBuffer)) return kGetIllegalArgument;
  if (mState.load() == NOT_IN_USE) {  // :368-372
    metrics.getNotReadyErrors++;
    metrics.getErrors++;
    return -1;
  }
  const Key k = MakeKey(fileId, pageIndex);
  std::shared_lock<std::shared_mutex> pageLock(mPageLocks[GetPageLockId(fileId, pageIndex)]);
  bool hasPage;
  {
    std::shared_lock<std::shared_mutex> metaLock(mMetaLock);
    hasPage = mMetaStore->HasPage(k);  // hasPage does not touch the evictor
  }
  if (!hasPage) return 0;
  int32_t got = 0;
  const int rc = mPageStore->Get(k, pageOffset, bytesToRead, (uint8_t*)buffer + offsetInBuffer, bufLen - offsetInBuffer, &got);
  if (rc == APS_EINVAL) return kGetIllegalArgument;  // IllegalArgumentException is not caught by getPage (:550)
  // getPage (:540-555) turns a failed or short read into -1; get (:383-395) then treats ANY bytesRead <= 0 as an
  // error -- which includes a successful read of zero bytes -- removes the page from the meta store and returns -1
  const int32_t bytesRead = (rc != APS_OK || got != bytesToRead) ? -1 : bytesToRead;
  if (bytesRead <= 0) {
    std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
    if (!mMetaStore->RemovePage(k)) metrics.cleanupGetErrors++;  // :389-393
    metrics.getErrors++;
    metrics.getStoreReadErrors++;
    return -1;
  }
  return bytesToRead;
}

void LocalCacheManager::GetMany(const std::string& fileId, int64_t firstPage, int32_t n, const int32_t* offset,
                                const int32_t* len, void* const* dst, int32_t* got) {
  for (int32_t i = 0; i < n; ++i) got[i] = 0;
  if (mState.load() == NOT_IN_USE || n <= 0) return;
  // page read locks of all pages involved, in lock-id order (the reference's deadlock rule, :197-203)
  std::vector<int32_t> lockIds((size_t)n);
  for (int32_t i = 0; i < n; ++i) lockIds[(size_t)i] = GetPageLockId(fileId, firstPage + i);
  std::vector<int32_t> order(lockIds);
  std::sort(order.begin(), order.end());
  order.erase(std::unique(order.begin(), order.end()), order.end());
  std::vector<std::shared_lock<std::shared_mutex>> held;
  held.reserve(order.size());
  for (int32_t id : order) held.emplace_back(mPageLocks[id]);
  std::vector<aps_get_req> reqs;
  std::vector<int32_t> which;
  std::vector<Key> keys;
  {
    std::shared_lock<std::shared_mutex> metaLock(mMetaLock);
    for (int32_t i = 0; i < n; ++i) {
      const Key k = MakeKey(fileId, firstPage + i);
      if (!mMetaStore->HasPage(k)) continue;
      reqs.push_back(aps_get_req{aps_key{k.hi, k.lo, k.idx}, (uint32_t)offset[i], (uint32_t)len[i], dst[i]});
      which.push_back(i);
      keys.push_back(k);
    }
  }
  if (reqs.empty()) return;
  std::vector<aps_result> res(reqs.size());
  const int rc = mPageStore->GetBatch((uint32_t)reqs.size(), reqs.data(), res.data());
  for (size_t r = 0; r < reqs.size(); ++r) {
    const int32_t i = which[r];
    if (rc == APS_OK && res[r].len == len[i] && len[i] > 0) {
      got[i] = len[i];
    } else {  // inconsistent with the meta store, or a read of zero bytes (bytesRead <= 0, :384): drop the page, as get() does
      std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
      mMetaStore->RemovePage(keys[r]);
      metrics.getErrors++;
      metrics.getStoreReadErrors++;
      got[i] = -1;
    }
  }
}

bool LocalCacheManager::Delete(const std::string& fileId, int64_t pageIndex) {  // :402-429
  if (mState.load() != READ_WRITE) {
    metrics.deleteNotReadyErrors++;
    metrics.deleteErrors++;
    return false;
  }
  const Key k = MakeKey(fileId, pageIndex);
  std::unique_lock<std::shared_mutex> pageLock(mPageLocks[GetPageLockId(fileId, pageIndex)]);
  {
    std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
    if (!mMetaStore->RemovePage(k)) {
      metrics.deleteNonExistingPageErrors++;
      metrics.deleteErrors++;
      return false;
    }
  }
  const bool ok = mPageStore->Delete(k) == APS_OK;
  if (!ok) {  // :422-427
    metrics.deleteStoreDeleteErrors++;
    metrics.deleteErrors++;
  }
  return ok;
}

void LocalCacheManager::RestoreOrInit() {  // :441-464
  if (!Restore()) {
    {
      std::unique_lock<std::shared_mutex> metaLock(mMetaLock);
      mMetaStore->Reset();
    }
    // The reference closes the store and creates it anew (PageStore.create wipes the cache directory) and goes on
    // READ_WRITE with an empty cache; only if that fails is the cache NOT_IN_USE and the caller falls back
    // (FileSystem.java:156-158).  A GPU store is emptied in place.
    if (!WipeStore()) {
      mState = NOT_IN_USE;
      return;
    }
  }
  mState = READ_WRITE;
}

bool LocalCacheManager::WipeStore() {
  std::vector<std::pair<Key, int64_t>> pages;
  if (mPageStore->GetPages(&pages) != APS_OK) return false;
  constexpr size_t kChunk = 1024;  // well below any store's max_batch
  std::vector<aps_key> keys;
  std::vector<aps_result> res;
  for (size_t lo = 0; lo < pages.size(); lo += kChunk) {
    const size_t n = std::min(kChunk, pages.size() - lo);
    keys.clear();
    for (size_t i = 0; i < n; ++i) keys.push_back(aps_key{pages[lo + i].first.hi, pages[lo + i].first.lo, pages[lo + i].first.idx});
    res.assign(n, aps_result{0, 0});
    if (mPageStore->DeleteBatch((uint32_t)n, keys.data(), res.data()) != APS_OK) return false;
    for (const aps_result& r : res)
      if (r.len < 0 && r.len != -APS_ENOTFOUND) return false;
  }
  return true;
}

bool LocalCacheManager::Restore() {  // :466-509
  std::vector<std::pair<Key, int64_t>> pages;
  if (mPageStore->GetPages(&pages) != APS_OK) return false;
  for (auto& p : pages) {
    const bool enoughSpace = mMetaStore->Bytes() + p.second <= mPageStore->GetCacheSize();
    const std::string fileId = mCodec.Decode(p.first.hi, p.first.lo);
    if (enoughSpace) mMetaStore->AddPage(p.first, PageMeta{p.second, GetPageLockId(fileId, (int64_t)p.first.idx)});
    else if (mPageStore->Delete(p.first) != APS_OK) return false;
  }
  return true;
}

// ================================================================================ batched read

// One read of [pos, pos + len) of a file through the cache with ONE batched store call for all resident pages it
// touches: page read-locks in id order, residency from the meta store, a single GetBatch into a pinned staging
// area, the rest fetched whole from `external` and offered to the cache -- same bytes and cache effects as the
// per-page loop of LocalCacheFileInStream.positionedRead (:175-223).  *stage is a pinned buffer the caller keeps
// between calls (grown here when too small).
int32_t BatchedRead(LocalCacheManager* mgr, const std::string& fileId, int64_t fileLen, int64_t pos, uint8_t* dst, int32_t len,
                    ExternalRead external, void* user, void** stage, uint64_t* stageBytes, int64_t* fromCache,
                    int64_t* fromExternal, int64_t* readExternal) {
  if (len < 0 || pos < 0) return kGetIllegalArgument;
  if (len == 0) return 0;
  if (pos >= fileLen) return -1;
  const int64_t ps = mgr->pageSize();
  const int64_t toRead = std::min<int64_t>(len, fileLen - pos);
  const int64_t firstPage = pos / ps, lastPage = (pos + toRead - 1) / ps;
  const int32_t n = (int32_t)(lastPage - firstPage + 1);
  if (*stageBytes < (uint64_t)toRead) {
    if (*stage) aps_host_free_pinned(*stage);
    *stage = nullptr;
    *stageBytes = 0;
    const uint64_t want = std::max<uint64_t>((uint64_t)toRead, 1u << 20);
    if (aps_host_alloc_pinned(want, stage) != APS_OK) return -1;
    *stageBytes = want;
  }
  std::vector<int32_t> off((size_t)n), ln((size_t)n), got((size_t)n);
  std::vector<void*> dsts((size_t)n);
  int64_t cur = pos, done = 0;
  for (int32_t i = 0; i < n; ++i) {
    off[(size_t)i] = (int32_t)(cur % ps);
    ln[(size_t)i] = (int32_t)std::min<int64_t>(ps - off[(size_t)i], toRead - done);
    dsts[(size_t)i] = (uint8_t*)*stage + done;
    cur += ln[(size_t)i];
    done += ln[(size_t)i];
  }
  mgr->GetMany(fileId, firstPage, n, off.data(), ln.data(), dsts.data(), got.data());
  std::vector<uint8_t> page;
  int32_t result = (int32_t)toRead;
  done = 0;
  for (int32_t i = 0; i < n && result >= 0; ++i) {
    uint8_t* out = dst + done;
    if (got[(size_t)i] > 0) {
      memcpy(out, dsts[(size_t)i], (size_t)ln[(size_t)i]);
      mgr->metrics.bytesReadCache += ln[(size_t)i];
      *fromCache += ln[(size_t)i];
    } else {
      const int64_t pageStart = (firstPage + i) * ps;
      const int32_t plen = (int32_t)std::min<int64_t>(ps, fileLen - pageStart);
      page.resize((size_t)plen);
      int32_t have = 0;
      while (have < plen) {
        const int r = external(user, pageStart + have, page.data() + have, plen - have);
        if (r <= 0) break;
        have += r;
      }
      mgr->metrics.bytesReadExternal += have;
      *readExternal += have;
      if (have != plen) {
        result = -3;  // "Failed to read complete page from external storage"
        break;
      }
      memcpy(out, page.data() + off[(size_t)i], (size_t)ln[(size_t)i]);
      mgr->metrics.bytesRequestedExternal += ln[(size_t)i];
      *fromExternal += ln[(size_t)i];
      mgr->Put(fileId, firstPage + i, page.data(), plen);
    }
    done += ln[(size_t)i];
  }
  return result;
}

// ================================================================================ LocalCacheFileInStream

LocalCacheFileInStream::LocalCacheFileInStream(LocalCacheManager* manager, const std::string& fileId, int64_t fileLength,
                                               ExternalRead external, void* user)
    : mManager(manager), mFileId(fileId), mLength(fileLength), mPageSize(manager ? manager->pageSize() : 1),
      mExternal(external), mUser(user), mBypass(!manager || manager->state() == NOT_IN_USE) {}

LocalCacheFileInStream::~LocalCacheFileInStream() {
  if (mStage) aps_host_free_pinned(mStage);
}

bool LocalCacheFileInStream::ReadExternalPage(int64_t pos, std::vector<uint8_t>* page) {  // :291-310
  std::lock_guard<std::mutex> lk(mExternalMu);
  const int64_t pageStart = pos - (pos % mPageSize);
  const int32_t pageSize = (int32_t)std::min<int64_t>(mPageSize, mLength - pageStart);
  page->resize((size_t)pageSize);
  int32_t total = 0;
  while (total < pageSize) {
    const int r = mExternal(mUser, pageStart + total, page->data() + total, pageSize - total);
    if (r <= 0) break;
    total += r;
  }
  bytesReadExternal += total;
  if (mManager) mManager->metrics.bytesReadExternal += total;
  return total == pageSize;  // else "Failed to read complete page from external storage"
}

int32_t LocalCacheFileInStream::ReadAt(int64_t* position, void* b, int32_t bLen, int32_t off, int32_t len) {
  if (len < 0 || off < 0 || *position < 0) return kGetIllegalArgument;
  if (len == 0) return 0;
  if (*position >= mLength) return -1;
  // What has to fit into b is what the read delivers -- min(len, bytes left in the file) -- not len: the reference's only
  // bound is LocalCacheManager.get's bytesToRead <= buffer.length - offsetInBuffer (CACHE/LocalCacheManager.java:364-366),
  // applied page by page (CTEST/LocalCacheFileInStreamTest.java:215-238 reads with len = 2 * fileSize into a buffer with
  // 2 * fileSize - 2 bytes of room).  The per-page loop below leaves the check to Get, as the reference does; the two
  // single-call forms check the whole read here.
  const int64_t want = std::min<int64_t>(len, mLength - *position);
  if ((mBypass || mBatched) && want > (int64_t)bLen - off) return kGetIllegalArgument;
  if (mBypass) {  // the delegated file system's own stream
    int64_t total = 0;
    while (total < want) {
      const int r = mExternal(mUser, *position + total, (uint8_t*)b + off + total, (int32_t)(want - total));
      if (r <= 0) return -3;
      total += r;
    }
    *position += total;
    return (int32_t)total;
  }
  if (mBatched) {  // one store call for every resident page of the read
    const int32_t r = BatchedRead(mManager, mFileId, mLength, *position, (uint8_t*)b + off, len, mExternal, mUser, &mStage,
                                  &mStageBytes, &bytesReadCache, &bytesRequestedExternal, &bytesReadExternal);
    if (r > 0) *position += r;
    return r;
  }
  int32_t total = 0;
  const int64_t toRead = std::min<int64_t>(len, mLength - *position);
  std::vector<uint8_t> page;
  while (total < toRead) {
    const int64_t pageIdx = *position / mPageSize;
    const int32_t pageOff = (int32_t)(*position % mPageSize);
    const int32_t inPage = (int32_t)std::min<int64_t>(mPageSize - pageOff, toRead - total);
    const int32_t got = mManager->Get(mFileId, pageIdx, pageOff, inPage, b, bLen, off + total);
    if (got == kGetIllegalArgument) return kGetIllegalArgument;  // this piece does not fit into b: thrown out of get()
    if (got > 0) {
      total += got;
      *position += got;
      bytesReadCache += got;
      mManager->metrics.bytesReadCache += got;
    } else {  // miss (0) or error (-1): the whole page comes from outside and is offered to the cache
      if (!ReadExternalPage(*position, &page)) return -3;
      if (!page.empty()) {
        memcpy((uint8_t*)b + off + total, page.data() + pageOff, (size_t)inPage);
        total += inPage;
        *position += inPage;
        bytesRequestedExternal += inPage;
        mManager->metrics.bytesRequestedExternal += inPage;
        mManager->Put(mFileId, pageIdx, page.data(), (int32_t)page.size());
      }
    }
  }
  return total;
}

int32_t LocalCacheFileInStream::Read(void* b, int32_t bLen, int32_t off, int32_t len) {  // :108-152
  return ReadAt(&mPosition, b, bLen, off, len);
}
int32_t LocalCacheFileInStream::PositionedRead(int64_t pos, void* b, int32_t bLen, int32_t off, int32_t len) {  // :175-223
  int64_t p = pos;
  return ReadAt(&p, b, bLen, off, len);
}
int64_t LocalCacheFileInStream::Skip(int64_t n) {  // :154-162
  if (mClosed) return -4;
  if (n <= 0) return 0;
  const int64_t toSkip = std::min(Remaining(), n);
  mPosition += toSkip;
  return toSkip;
}
int LocalCacheFileInStream::Seek(int64_t pos) {  // :230-243
  if (mClosed) return -4;
  if (pos < 0 || pos > mLength) return kGetIllegalArgument;
  mPosition = pos;
  return 0;
}

void FillSyntheticPage(uint64_t seed, uint64_t file, uint64_t index, uint8_t* out, int64_t len) {
  const uint64_t s = seed ^ (file << 32) ^ index;
  for (int64_t k = 0; k * 8 < len; ++k) {
    uint64_t z = s + (uint64_t)(k + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    memcpy(out + 8 * k, &z, (size_t)std::min<int64_t>(8, len - 8 * k));
  }
}

}  // namespace alluxio_b200

// ==================================================================================== C view

using namespace alluxio_b200;

struct apc_manager {
  std::unique_ptr<LocalCacheManager> mgr;
  GpuPageStore* gpu = nullptr;             // single store, owned by mgr
  ShardedGpuPageStore* sharded = nullptr;  // or several, owned by mgr
  TimeBoundPageStore* bound = nullptr;     // the wrapper around either, when a timeout is configured
};
struct apc_evictor {
  std::unique_ptr<CacheEvictor> ev;
  FileKeyCodec codec;
};
struct apc_meta {
  std::unique_ptr<MetaStore> meta;
  FileKeyCodec codec;
};

static Key KeyOf(FileKeyCodec& c, const char* f, int64_t i) {
  Key k;
  c.Encode(f, &k.hi, &k.lo);
  k.idx = (uint64_t)i;
  return k;
}
static void CopyName(const std::string& s, char* out, size_t cap) {
  if (!cap) return;
  strncpy(out, s.c_str(), cap - 1);
  out[cap - 1] = 0;
}

extern "C" {

int apc_manager_create(const aps_config* cfg, const char* evictor, double lfu_logbase, apc_manager** out) {
  if (!cfg || !evictor || !out) return APS_EINVAL;
  *out = nullptr;
  auto ev = CacheEvictor::Create(evictor, lfu_logbase);
  if (!ev) return APS_EINVAL;
  std::unique_ptr<GpuPageStore> store;
  const int rc = GpuPageStore::Open(*cfg, &store);
  if (rc != APS_OK) return rc;
  apc_manager* m = new apc_manager();
  m->gpu = store.get();
  m->mgr.reset(new LocalCacheManager(cfg->page_size, std::unique_ptr<MetaStore>(new MetaStore(std::move(ev))), std::move(store)));
  *out = m;
  return APS_OK;
}
int apc_manager_create_sharded(const aps_config* cfg, uint32_t n_shards, const int32_t* devices, const char* evictor,
                               double lfu_logbase, apc_manager** out) {
  if (!cfg || !evictor || !out) return APS_EINVAL;
  *out = nullptr;
  auto ev = CacheEvictor::Create(evictor, lfu_logbase);
  if (!ev) return APS_EINVAL;
  std::unique_ptr<ShardedGpuPageStore> store;
  const int rc = ShardedGpuPageStore::Open(*cfg, n_shards, devices, &store);
  if (rc != APS_OK) return rc;
  apc_manager* m = new apc_manager();
  m->sharded = store.get();
  m->mgr.reset(new LocalCacheManager(cfg->page_size, std::unique_ptr<MetaStore>(new MetaStore(std::move(ev))), std::move(store)));
  *out = m;
  return APS_OK;
}
int apc_manager_create_ex(const apc_options* o, apc_manager** out) {
  if (!o || !o->evictor || !out) return APS_EINVAL;
  *out = nullptr;
  auto ev = CacheEvictor::Create(o->evictor, o->lfu_logbase);
  if (!ev) return APS_EINVAL;
  ManagerOptions mo;
  mo.asyncWriteThreads = o->async_write_threads;
  mo.asyncRestore = o->async_restore != 0;
  std::unique_ptr<apc_manager> m(new apc_manager());
  std::unique_ptr<PageStore> store;
  if (o->over) {
    std::unique_ptr<GpuPageStore> gs = GpuPageStore::Adopt(o->over);
    m->gpu = gs.get();
    store = std::move(gs);
  } else if (o->n_shards > 0) {
    std::unique_ptr<ShardedGpuPageStore> ss;
    const int rc = ShardedGpuPageStore::Open(o->store, o->n_shards, o->shard_devices, &ss);
    if (rc != APS_OK) return rc;
    m->sharded = ss.get();
    store = std::move(ss);
  } else {
    std::unique_ptr<GpuPageStore> gs;
    const int rc = GpuPageStore::Open(o->store, &gs);
    if (rc != APS_OK) return rc;
    m->gpu = gs.get();
    store = std::move(gs);
  }
  if (o->timeout_ms > 0) {  // PageStore.create wraps the store when a timeout is set (CACHE/PageStore.java:72-75)
    auto* tb = new TimeBoundPageStore(std::move(store), o->timeout_ms, o->timeout_threads > 0 ? o->timeout_threads : 32);
    m->bound = tb;
    store.reset(tb);
  }
  m->mgr.reset(new LocalCacheManager(o->store.page_size, std::unique_ptr<MetaStore>(new MetaStore(std::move(ev))),
                                     std::move(store), mo));
  *out = m.release();
  return APS_OK;
}
void apc_wait_async(apc_manager* m) { m->mgr->WaitAsync(); }
int apc_store_timeout_counters(apc_manager* m, int64_t* out4) {
  if (!m->bound) return APS_EINVAL;
  out4[0] = m->bound->storePutTimeout.load();
  out4[1] = m->bound->storeGetTimeout.load();
  out4[2] = m->bound->storeDeleteTimeout.load();
  out4[3] = m->bound->storeThreadsRejected.load();
  return APS_OK;
}
int apc_set_store_delays(apc_manager* m, int32_t put_ms, int32_t get_ms, int32_t delete_ms) {
  if (!m->bound) return APS_EINVAL;
  m->bound->putDelayMs = put_ms;
  m->bound->getDelayMs = get_ms;
  m->bound->deleteDelayMs = delete_ms;
  return APS_OK;
}
int64_t apc_put_async_rejections(apc_manager* m) { return m->mgr->putAsyncRejectionErrors.load(); }
int64_t apc_put_not_ready_errors(apc_manager* m) { return m->mgr->putNotReadyErrors.load(); }

// CacheManager.Factory (CACHE/CacheManager.java:69-127): one manager per process; a caller that arrives while
// another thread is still creating it gets null instead of waiting.
namespace {
std::mutex gFactoryLock;
std::atomic<apc_manager*> gFactoryManager{nullptr};
std::atomic<int64_t> gFactoryCreateErrors{0};
}  // namespace
int apc_factory_get(const apc_options* o, apc_manager** out) {
  if (!out) return APS_EINVAL;
  *out = nullptr;
  if (gFactoryManager.load() == nullptr) {
    std::unique_lock<std::mutex> lk(gFactoryLock, std::try_to_lock);
    if (lk.owns_lock() && gFactoryManager.load() == nullptr) {
      apc_manager* m = nullptr;
      const int rc = apc_manager_create_ex(o, &m);
      if (rc != APS_OK) {
        gFactoryCreateErrors++;
        return rc;  // IOException("Failed to create CacheManager")
      }
      gFactoryManager.store(m);
    }
  }
  *out = gFactoryManager.load();
  return APS_OK;
}
void apc_factory_clear(void) {
  std::lock_guard<std::mutex> lk(gFactoryLock);
  apc_manager* m = gFactoryManager.exchange(nullptr);
  if (m) apc_manager_close(m);
}
int64_t apc_factory_create_errors(void) { return gFactoryCreateErrors.load(); }

// LocalCacheFileSystem.openFile + LocalCacheFileInStream
struct apc_stream {
  std::unique_ptr<LocalCacheFileInStream> in;
};
apc_stream* apc_file_open(apc_manager* m, const char* file_id, int64_t file_len, apc_external_read external, void* user) {
  if (!file_id || !external || file_len < 0) return nullptr;
  apc_stream* s = new apc_stream();
  s->in.reset(new LocalCacheFileInStream(m ? m->mgr.get() : nullptr, file_id, file_len, external, user));
  return s;
}
int32_t apc_stream_read(apc_stream* s, void* b, int32_t b_len, int32_t off, int32_t len) { return s->in->Read(b, b_len, off, len); }
int32_t apc_stream_positioned_read(apc_stream* s, int64_t pos, void* b, int32_t b_len, int32_t off, int32_t len) {
  return s->in->PositionedRead(pos, b, b_len, off, len);
}
int64_t apc_stream_skip(apc_stream* s, int64_t n) { return s->in->Skip(n); }
int apc_stream_seek(apc_stream* s, int64_t pos) { return s->in->Seek(pos); }
int64_t apc_stream_remaining(apc_stream* s) { return s->in->Remaining(); }
int64_t apc_stream_get_pos(apc_stream* s) { return s->in->GetPos(); }
int apc_stream_bypass(apc_stream* s) { return s->in->bypass() ? 1 : 0; }
void apc_stream_counters(apc_stream* s, int64_t* out3) {
  out3[0] = s->in->bytesReadCache;
  out3[1] = s->in->bytesRequestedExternal;
  out3[2] = s->in->bytesReadExternal;
}
void apc_stream_set_batched(apc_stream* s, int on) { s->in->SetBatched(on != 0); }
void apc_stream_close(apc_stream* s) { s->in->Close(); }
void apc_stream_destroy(apc_stream* s) { delete s; }

uint32_t apc_shard_count(apc_manager* m) { return m->sharded ? m->sharded->shards() : 1; }
aps_store* apc_shard_store(apc_manager* m, uint32_t i) {
  if (m->sharded) return i < m->sharded->shards() ? m->sharded->shard(i)->handle() : nullptr;
  return i == 0 ? m->gpu->handle() : nullptr;
}
uint32_t apc_shard_of(apc_manager* m, const char* f, int64_t i) {
  if (!m->sharded) return 0;
  return m->sharded->OwnerOf(KeyOf(*m->mgr->codec(), f, i));
}
int apc_manager_create_over(aps_store* store, uint32_t page_size, const char* evictor, double lfu_logbase, apc_manager** out) {
  if (!store || !evictor || !out) return APS_EINVAL;
  *out = nullptr;
  auto ev = CacheEvictor::Create(evictor, lfu_logbase);
  if (!ev) return APS_EINVAL;
  std::unique_ptr<GpuPageStore> gs = GpuPageStore::Adopt(store);
  apc_manager* m = new apc_manager();
  m->gpu = gs.get();
  m->mgr.reset(new LocalCacheManager(page_size, std::unique_ptr<MetaStore>(new MetaStore(std::move(ev))), std::move(gs)));
  *out = m;
  return APS_OK;
}

// Warm restart with ids that are neither decimal nor 32-hex: they reach the store as indices into this process's intern
// table, which the snapshot therefore has to carry (aps_snapshot_save_names / aps_snapshot_names).
static int CollectName(void* user, uint32_t, const char* name) {
  static_cast<std::vector<std::string>*>(user)->push_back(name);
  return 0;
}
int apc_manager_create_over_snapshot(aps_store* store, uint32_t page_size, const char* evictor, double lfu_logbase,
                                     const char* snapshot_path, apc_manager** out) {
  if (!store || !evictor || !out || !snapshot_path) return APS_EINVAL;
  *out = nullptr;
  auto ev = CacheEvictor::Create(evictor, lfu_logbase);
  if (!ev) return APS_EINVAL;
  ManagerOptions mo;
  int rc = aps_snapshot_names(snapshot_path, CollectName, &mo.internedNames);
  if (rc != APS_OK) return rc;
  rc = aps_snapshot_load(store, snapshot_path, nullptr, nullptr);
  if (rc != APS_OK) return rc;
  std::unique_ptr<GpuPageStore> gs = GpuPageStore::Adopt(store);
  apc_manager* m = new apc_manager();
  m->gpu = gs.get();
  m->mgr.reset(new LocalCacheManager(page_size, std::unique_ptr<MetaStore>(new MetaStore(std::move(ev))), std::move(gs), mo));
  *out = m;
  return APS_OK;
}
int apc_manager_snapshot_save(apc_manager* m, const char* path, uint64_t* pages_out) {
  if (!m || !m->gpu || !path) return APS_EINVAL;
  const std::vector<std::string> names = m->mgr->codec()->Names();
  std::vector<const char*> ptrs;
  for (const std::string& n : names) ptrs.push_back(n.c_str());
  return aps_snapshot_save_names(m->gpu->handle(), path, (uint32_t)ptrs.size(), ptrs.data(), pages_out);
}

// NoExceptionCacheManager (CACHE/NoExceptionCacheManager.java:37-79): nothing escapes; errors are counted.
int32_t apc_get_noexcept(apc_manager* m, const char* f, int64_t i, int32_t page_offset, int32_t bytes_to_read, void* buffer,
                         int32_t buffer_len, int32_t offset_in_buffer) {
  const int32_t r = m->mgr->Get(f, i, page_offset, bytes_to_read, buffer, buffer_len, offset_in_buffer);
  if (r == kGetIllegalArgument) {
    m->mgr->metrics.getErrors++;
    return -1;
  }
  return r;
}

void apc_metrics_read(apc_manager* m, int64_t* out14) {
  const CacheMetrics& c = m->mgr->metrics;
  const int64_t v[14] = {c.bytesReadCache, c.bytesRequestedExternal, c.bytesReadExternal, c.bytesWrittenCache, c.bytesEvicted,
                         c.pagesEvicted, c.getErrors, c.putErrors, c.deleteErrors, c.putEvictionErrors, c.putStoreWriteErrors,
                         c.putStoreDeleteErrors, c.getStoreReadErrors, c.deleteNonExistingPageErrors};
  memcpy(out14, v, sizeof(v));
}

void apc_metrics_read_more(apc_manager* m, int64_t* out8) {
  const CacheMetrics& c = m->mgr->metrics;
  const int64_t v[8] = {c.getNotReadyErrors, c.deleteNotReadyErrors, c.deleteStoreDeleteErrors, c.putBenignRacingErrors,
                        c.cleanupGetErrors, c.cleanupPutErrors, m->mgr->putNotReadyErrors.load(),
                        m->mgr->putAsyncRejectionErrors.load()};
  memcpy(out8, v, sizeof(v));
}

// Multi-page read in one store call: LocalCacheFileInStream.positionedRead with the per-page
// CacheManager.get loop (CACHE/LocalCacheFileInStream.java:192-214) replaced by ONE batched get for all resident
// pages; pages that miss are fetched whole from `external` and put, as in the reference.  dst may be any host
// memory: a pageable dst is served through a pinned staging buffer.
int32_t apc_positioned_read_batched(apc_manager* m, const char* file_id, int64_t file_len, int64_t pos, void* dst, int32_t len,
                                    apc_external_read external, void* user, int64_t* from_cache, int64_t* from_external) {
  void* stage = nullptr;
  uint64_t stageBytes = 0;
  int64_t fc = 0, fe = 0, re = 0;
  const int32_t r = BatchedRead(m->mgr.get(), file_id, file_len, pos, (uint8_t*)dst, len, external, user, &stage, &stageBytes,
                                &fc, &fe, &re);
  if (stage) aps_host_free_pinned(stage);
  if (from_cache) *from_cache += fc;
  if (from_external) *from_external += fe;
  return r;
}

int apc_manager_close(apc_manager* m) {
  delete m;
  return APS_OK;
}
int apc_put(apc_manager* m, const char* f, int64_t i, const void* page, int32_t len) { return m->mgr->Put(f, i, page, len); }
int32_t apc_get(apc_manager* m, const char* f, int64_t i, int32_t page_offset, int32_t bytes_to_read, void* buffer,
                int32_t buffer_len, int32_t offset_in_buffer) {
  return m->mgr->Get(f, i, page_offset, bytes_to_read, buffer, buffer_len, offset_in_buffer);
}
int apc_delete(apc_manager* m, const char* f, int64_t i) { return m->mgr->Delete(f, i); }
int apc_state(apc_manager* m) { return m->mgr->state(); }
int64_t apc_bytes(apc_manager* m) { return m->mgr->meta()->Bytes(); }
int64_t apc_pages(apc_manager* m) { return m->mgr->meta()->Pages(); }
int32_t apc_page_lock_id(apc_manager* m, const char* f, int64_t i) { return m->mgr->GetPageLockId(f, i); }
int apc_set_store_faults(apc_manager* m, uint32_t mask) { return m->mgr->store()->SetFaults(mask); }
aps_store* apc_store(apc_manager* m) { return apc_shard_store(m, 0); }
void apc_file_key(apc_manager* m, const char* f, uint64_t* hi, uint64_t* lo) { m->mgr->codec()->Encode(f, hi, lo); }
int32_t apc_java_string_hashcode(const char* utf8) { return JavaStringHashCode(utf8); }
void apc_hadoop_file_identifier(const char* path_utf8, char* out33) {
  const std::string id = HadoopFileIdentifier(path_utf8 ? path_utf8 : "");
  memcpy(out33, id.c_str(), 33);
}
void apc_file_key_of(const char* file_id, uint64_t* hi, uint64_t* lo) {
  FileKeyCodec c;  // decimal and 32-hex ids need no state; anything else would be interned per codec
  c.Encode(file_id, hi, lo);
}

// CACHE/LocalCacheFileInStream.java:175-223 + readExternalPage :291-310
int32_t apc_positioned_read(apc_manager* m, const char* file_id, int64_t file_len, int64_t pos, void* dst, int32_t len,
                            apc_external_read external, void* user, int64_t* from_cache, int64_t* from_external) {
  if (len < 0 || pos < 0) return APC_GET_ILLEGAL_ARGUMENT;
  if (len == 0) return 0;
  if (pos >= file_len) return -1;
  LocalCacheManager* mgr = m->mgr.get();
  const int64_t ps = mgr->pageSize();
  int32_t total = 0;
  int64_t cur = pos;
  const int64_t toRead = std::min<int64_t>(len, file_len - pos);
  std::vector<uint8_t> page;
  while (total < toRead) {
    const int64_t pageIdx = cur / ps;
    const int32_t pageOff = (int32_t)(cur % ps);
    const int32_t inPage = (int32_t)std::min<int64_t>(ps - pageOff, toRead - total);
    const int32_t got = mgr->Get(file_id, pageIdx, pageOff, inPage, dst, len, total);
    if (got > 0) {
      total += got;
      cur += got;
      if (from_cache) *from_cache += got;
    } else {
      const int64_t pageStart = cur - (cur % ps);
      const int32_t plen = (int32_t)std::min<int64_t>(ps, file_len - pageStart);
      page.resize((size_t)plen);
      int32_t have = 0;
      while (have < plen) {
        const int r = external(user, pageStart + have, page.data() + have, plen - have);
        if (r <= 0) break;
        have += r;
      }
      if (have != plen) return -1;  // "Failed to read complete page from external storage"
      memcpy((uint8_t*)dst + total, page.data() + pageOff, (size_t)inPage);
      total += inPage;
      cur += inPage;
      if (from_external) *from_external += inPage;
      mgr->Put(file_id, pageIdx, page.data(), plen);
    }
  }
  return total;
}

int64_t apc_replay(apc_manager* m, int64_t n, const int64_t* file, const int64_t* page, const int32_t* page_len,
                   uint64_t seed, int verify, uint8_t* hit, uint8_t* put_ok, int64_t* victim_file, int64_t* victim_page,
                   int64_t victims_cap, int64_t* final_pages, int64_t* final_bytes, int64_t* verify_failures) {
  LocalCacheManager* mgr = m->mgr.get();
  mgr->logEvictions = true;
  mgr->evicted.clear();
  int64_t nv = 0, bad = 0;
  std::vector<uint8_t> buf, want;
  for (int64_t k = 0; k < n; ++k) {
    const std::string f = std::to_string(file[k]);
    const int32_t len = page_len[k];
    buf.resize((size_t)std::max(len, 1));
    const int32_t got = mgr->Get(f, page[k], 0, len, buf.data(), len, 0);
    hit[k] = got > 0;
    put_ok[k] = 255;
    if (got > 0) {
      if (verify) {
        want.resize((size_t)len);
        FillSyntheticPage(seed, (uint64_t)file[k], (uint64_t)page[k], want.data(), len);
        if (memcmp(want.data(), buf.data(), (size_t)len) != 0) ++bad;
      }
    } else {
      FillSyntheticPage(seed, (uint64_t)file[k], (uint64_t)page[k], buf.data(), len);
      put_ok[k] = mgr->Put(f, page[k], buf.data(), len) ? 1 : 0;
    }
    for (const Key& v : mgr->evicted) {
      if (nv < victims_cap) {
        victim_file[nv] = (int64_t)v.lo;  // decimal ids encode as (0, value)
        victim_page[nv] = (int64_t)v.idx;
      }
      ++nv;
    }
    mgr->evicted.clear();
  }
  mgr->logEvictions = false;
  if (final_pages) *final_pages = mgr->meta()->Pages();
  if (final_bytes) *final_bytes = mgr->meta()->Bytes();
  if (verify_failures) *verify_failures = bad;
  return nv;
}

// ---- policy objects on their own
apc_evictor* apc_evictor_create(const char* kind, double lfu_logbase) {
  auto ev = CacheEvictor::Create(kind, lfu_logbase);
  if (!ev) return nullptr;
  apc_evictor* e = new apc_evictor();
  e->ev = std::move(ev);
  return e;
}
void apc_evictor_destroy(apc_evictor* e) { delete e; }
void apc_evictor_update_on_get(apc_evictor* e, const char* f, int64_t i) { e->ev->UpdateOnGet(KeyOf(e->codec, f, i)); }
void apc_evictor_update_on_put(apc_evictor* e, const char* f, int64_t i) { e->ev->UpdateOnPut(KeyOf(e->codec, f, i)); }
void apc_evictor_update_on_delete(apc_evictor* e, const char* f, int64_t i) { e->ev->UpdateOnDelete(KeyOf(e->codec, f, i)); }
void apc_evictor_reset(apc_evictor* e) { e->ev->Reset(); }
int apc_evictor_evict(apc_evictor* e, char* file_out, size_t cap, int64_t* index_out) {
  Key v;
  if (!e->ev->Evict(&v)) return 0;
  CopyName(e->codec.Decode(v.hi, v.lo), file_out, cap);
  *index_out = (int64_t)v.idx;
  return 1;
}

apc_meta* apc_meta_create(const char* evictor, double lfu_logbase) {
  auto ev = CacheEvictor::Create(evictor, lfu_logbase);
  if (!ev) return nullptr;
  apc_meta* m = new apc_meta();
  m->meta.reset(new MetaStore(std::move(ev)));
  return m;
}
void apc_meta_destroy(apc_meta* m) { delete m; }
int apc_meta_has_page(apc_meta* m, const char* f, int64_t i) { return m->meta->HasPage(KeyOf(m->codec, f, i)); }
void apc_meta_add_page(apc_meta* m, const char* f, int64_t i, int64_t size) { m->meta->AddPage(KeyOf(m->codec, f, i), PageMeta{size, 0}); }
int apc_meta_get_page_info(apc_meta* m, const char* f, int64_t i, int64_t* size) {
  PageMeta pm;
  if (!m->meta->GetPageInfo(KeyOf(m->codec, f, i), &pm)) return 0;
  *size = pm.size;
  return 1;
}
int apc_meta_remove_page(apc_meta* m, const char* f, int64_t i) { return m->meta->RemovePage(KeyOf(m->codec, f, i)); }
int64_t apc_meta_bytes(apc_meta* m) { return m->meta->Bytes(); }
int64_t apc_meta_pages(apc_meta* m) { return m->meta->Pages(); }
void apc_meta_reset(apc_meta* m) { m->meta->Reset(); }
int apc_meta_evict(apc_meta* m, char* file_out, size_t cap, int64_t* index_out, int64_t* size_out) {
  Key v;
  PageMeta pm;
  if (!m->meta->Evict(&v, &pm)) return 0;
  CopyName(m->codec.Decode(v.hi, v.lo), file_out, cap);
  *index_out = (int64_t)v.idx;
  *size_out = pm.size;
  return 1;
}

}  // extern "C"

// ==================================================================================== batched replay
#include <chrono>
#include <unordered_set>

namespace {
using Clock = std::chrono::steady_clock;
inline double Secs(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double>(b - a).count(); }
}  // namespace

extern "C" int64_t apc_replay_batched(apc_manager* m, int64_t n, const int64_t* file, const int64_t* page,
                                      const int32_t* page_len, uint64_t seed, uint32_t window, uint8_t* hit, uint8_t* put_ok,
                                      int64_t* victim_file, int64_t* victim_page, int64_t victims_cap, int64_t* final_pages,
                                      int64_t* final_bytes, int64_t* store_ops, double* seconds_host, double* seconds_store) {
  LocalCacheManager* mgr = m->mgr.get();
  MetaStore* meta = mgr->meta();
  PageStore* store = mgr->store();
  const int64_t cacheSize = store->GetCacheSize();
  const uint32_t ps = (uint32_t)mgr->pageSize();
  if (window == 0) window = 1024;
  // device staging: put payloads and a sink for gets, on the store's (first) GPU
  void *d_put = nullptr, *d_sink = nullptr;
  const int32_t dev = store->StagingDevice();
  if (aps_device_alloc(dev, (uint64_t)window * ps, &d_put) != APS_OK) return -1;
  if (aps_device_alloc(dev, (uint64_t)window * ps, &d_sink) != APS_OK) {
    aps_device_free(dev, d_put);
    return -1;
  }
  // One window of store operations.  The main thread fills `cur` while the worker thread executes the previous one
  // (`run`): host policy and GPU batches overlap, windows still reach the store strictly in order.
  struct Window {
    std::vector<aps_key> dels;
    std::vector<aps_put_req> puts;
    std::vector<aps_get_req> gets;
    std::vector<uint64_t> put_files, put_pages;
    bool Empty() const { return dels.empty() && puts.empty() && gets.empty(); }
    void Clear() { dels.clear(); puts.clear(); gets.clear(); put_files.clear(); put_pages.clear(); }
  };
  Window cur, run;
  std::vector<aps_key>& dels = cur.dels;
  std::vector<aps_put_req>& puts = cur.puts;
  std::vector<aps_get_req>& gets = cur.gets;
  std::vector<uint64_t>&put_files = cur.put_files, &put_pages = cur.put_pages;
  // "is this id part of the window that has not been flushed yet?" is answered by the page's own meta entry: the
  // replay stamps it with the window number on every put and hit (no side table, no extra hash probe)
  uint32_t epoch = 1;
  int64_t ops[5] = {0, 0, 0, 0, 0};
  double tStore = 0, tWait = 0;
  std::atomic<bool> failed{false};
  auto execute = [&](Window& w) {  // worker thread
    const auto t0 = Clock::now();
    std::vector<aps_result> res;
    ops[3]++;
    if (!w.dels.empty()) {
      res.resize(w.dels.size());
      if (store->DeleteBatch((uint32_t)w.dels.size(), w.dels.data(), res.data()) != APS_OK) failed = true;
      for (auto& r : res) if (r.len != 0) failed = true;
      ops[0] += (int64_t)w.dels.size();
    }
    if (!w.puts.empty()) {
      // payload: all pages of this window generated at full page stride, then put with their real lengths
      if (store->FillPages(d_put, ps, ps, (uint32_t)w.puts.size(), w.put_files.data(), w.put_pages.data(), seed) != APS_OK)
        failed = true;
      res.resize(w.puts.size());
      if (store->PutBatch((uint32_t)w.puts.size(), w.puts.data(), res.data()) != APS_OK) failed = true;
      for (size_t i = 0; i < w.puts.size(); ++i) if (res[i].len != (int32_t)w.puts[i].len) failed = true;
      ops[1] += (int64_t)w.puts.size();
    }
    if (!w.gets.empty()) {
      res.resize(w.gets.size());
      if (store->GetBatch((uint32_t)w.gets.size(), w.gets.data(), res.data()) != APS_OK) failed = true;
      for (size_t i = 0; i < w.gets.size(); ++i) if (res[i].len != (int32_t)w.gets[i].nbytes) ops[4]++;
      ops[2] += (int64_t)w.gets.size();
    }
    w.Clear();
    tStore += Secs(t0, Clock::now());
  };
  std::mutex wmu;
  std::condition_variable wcv;
  bool busy = false, quit = false;
  std::thread worker([&] {
    std::unique_lock<std::mutex> lk(wmu);
    for (;;) {
      wcv.wait(lk, [&] { return busy || quit; });
      if (!busy) return;
      lk.unlock();
      execute(run);
      lk.lock();
      busy = false;
      wcv.notify_all();
    }
  });
  auto wait_idle = [&]() {
    const auto t0 = Clock::now();
    std::unique_lock<std::mutex> lk(wmu);
    wcv.wait(lk, [&] { return !busy; });
    tWait += Secs(t0, Clock::now());
  };
  auto flush = [&]() {
    if (cur.Empty()) return;
    wait_idle();  // the previous window is done: its vectors are free again
    {
      std::lock_guard<std::mutex> lk(wmu);
      std::swap(cur, run);  // (the references above keep naming `cur`)
      busy = true;
    }
    wcv.notify_all();
    ++epoch;
  };
  const auto tBegin = Clock::now();
  int64_t nv = 0;
  for (int64_t k = 0; k < n && !failed; ++k) {
    const Key key{0, (uint64_t)file[k], (uint64_t)page[k]};  // decimal file ids encode as (0, value)
    const int32_t len = page_len[k];
    put_ok[k] = 255;
    PageMeta* pm = meta->Find(key);
    if (pm && len == 0) {
      // get of zero bytes from a resident page: bytesRead <= 0 counts as a failed read (LocalCacheManager.java:383-395),
      // the page leaves the meta store (not the page store) and the caller falls through to the miss path
      if (pm->epoch == epoch) flush();  // its put is still queued in this window: keep per-id order in the store
      meta->RemovePage(key);
      pm = nullptr;
    }
    if (pm) {  // LocalCacheManager.get: hit
      hit[k] = 1;
      gets.push_back(aps_get_req{aps_key{key.hi, key.lo, key.idx}, 0, (uint32_t)len, (uint8_t*)d_sink + (size_t)gets.size() * ps});
      pm->epoch = epoch;
    } else {  // miss -> put (LocalCacheManager.putInternal, single-threaded reading)
      hit[k] = 0;
      bool enough = meta->Bytes() + len <= cacheSize, doPut = false;
      if (enough) {
        meta->AddPage(key, PageMeta{len, 0, epoch});
        doPut = true;
      } else {
        Key victim;
        PageMeta vm;
        if (!meta->Evict(&victim, &vm)) {
          put_ok[k] = 0;
        } else {
          meta->RemovePage(victim);
          if (vm.epoch == epoch) flush();  // the victim was put or read in this window: keep per-id order
          enough = meta->Bytes() + len <= cacheSize;
          if (enough) meta->AddPage(key, PageMeta{len, 0, epoch});
          dels.push_back(aps_key{victim.hi, victim.lo, victim.idx});
          if (nv < victims_cap) {
            victim_file[nv] = (int64_t)victim.lo;
            victim_page[nv] = (int64_t)victim.idx;
          }
          ++nv;
          doPut = enough;
          if (!enough) put_ok[k] = 0;
        }
      }
      if (doPut) {
        put_ok[k] = 1;
        puts.push_back(aps_put_req{aps_key{key.hi, key.lo, key.idx}, (uint32_t)len, 0, (uint8_t*)d_put + (size_t)puts.size() * ps});
        put_files.push_back((uint64_t)file[k]);
        put_pages.push_back((uint64_t)page[k]);
      }
    }
    if (dels.size() >= window || puts.size() >= window || gets.size() >= window) flush();
  }
  flush();
  wait_idle();
  {
    std::lock_guard<std::mutex> lk(wmu);
    quit = true;
  }
  wcv.notify_all();
  worker.join();
  const double tAll = Secs(tBegin, Clock::now());
  aps_device_free(dev, d_put);
  aps_device_free(dev, d_sink);
  if (final_pages) *final_pages = meta->Pages();
  if (final_bytes) *final_bytes = meta->Bytes();
  if (store_ops) memcpy(store_ops, ops, sizeof(ops));
  if (seconds_store) *seconds_store = tStore;
  if (seconds_host) *seconds_host = tAll - tWait;  // the policy's own time; tWait = what it spent waiting for the store
  return failed.load() ? -1 : nv;
}
