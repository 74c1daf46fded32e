// Host-side cache layer: a native mirror of the reference's Java classes that sit above the
// PageStore boundary, for embedders without a JVM.  Same names, argument meaning and results as
//   CACHE/PageId.java, CACHE/CacheEvictor.java, CACHE/evictor/LRUCacheEvictor.java,
//   CACHE/evictor/LFUCacheEvictor.java, CACHE/DefaultMetaStore.java, CACHE/LocalCacheManager.java,
//   CACHE/LocalCacheFileInStream.java (page arithmetic)
// (CACHE/ = /root/reference/core/client/fs/src/main/java/alluxio/client/file/cache/), but its own
// design: ids are turned into fixed 24-byte keys once at the API edge, policy state lives in index-
// linked node pools, and the manager is thread-safe with the reference's lock hierarchy.
// This is product code: it never touches oracle/.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/aps.h"

namespace alluxio_b200 {

// ---- keys -----------------------------------------------------------------------------------------
struct Key {
  uint64_t hi, lo, idx;
  bool operator==(const Key& o) const { return hi == o.hi && lo == o.lo && idx == o.idx; }
};
struct KeyHash {
  size_t operator()(const Key& k) const {
    uint64_t z = k.idx + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z ^= k.lo;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= k.hi;
    return (size_t)(z ^ (z >> 31));
  }
};

// Open-addressing map from Key to a small value: linear probing over a power-of-two array kept at most half
// full, deletion by backward shift (no tombstones).  The policy state of a cache with millions of pages is
// probed tens of millions of times per second; node-based std::unordered_map costs a cache miss per node.
template <typename V>
class FlatMap {
 public:
  FlatMap() { Rebuild(16); }
  V* Find(const Key& k) {
    for (size_t i = KeyHash()(k) & mMask;; i = (i + 1) & mMask) {
      Slot& s = mSlots[i];
      if (!s.used) return nullptr;
      if (s.key == k) return &s.val;
    }
  }
  const V* Find(const Key& k) const { return const_cast<FlatMap*>(this)->Find(k); }
  // Returns the value for k, inserting `init` if absent; *inserted tells which.
  V* Emplace(const Key& k, const V& init, bool* inserted) {
    if ((mSize + 1) * 2 > mSlots.size()) Rebuild(mSlots.size() * 2);
    for (size_t i = KeyHash()(k) & mMask;; i = (i + 1) & mMask) {
      Slot& s = mSlots[i];
      if (!s.used) {
        s.used = true;
        s.key = k;
        s.val = init;
        ++mSize;
        *inserted = true;
        return &s.val;
      }
      if (s.key == k) {
        *inserted = false;
        return &s.val;
      }
    }
  }
  bool Erase(const Key& k) {
    size_t i = KeyHash()(k) & mMask;
    for (;; i = (i + 1) & mMask) {
      if (!mSlots[i].used) return false;
      if (mSlots[i].key == k) break;
    }
    // shift the rest of the cluster back so that every key stays reachable from its home slot
    size_t hole = i;
    for (size_t j = (i + 1) & mMask; mSlots[j].used; j = (j + 1) & mMask) {
      const size_t home = KeyHash()(mSlots[j].key) & mMask;
      const bool movable = ((j - home) & mMask) >= ((j - hole) & mMask);
      if (movable) {
        mSlots[hole] = mSlots[j];
        hole = j;
      }
    }
    mSlots[hole].used = false;
    --mSize;
    return true;
  }
  void Clear() {
    mSlots.clear();
    Rebuild(16);
  }
  size_t Size() const { return mSize; }
 private:
  struct Slot {
    Key key;
    V val;
    bool used = false;
  };
  void Rebuild(size_t capacity) {
    std::vector<Slot> old;
    old.swap(mSlots);
    mSlots.assign(capacity, Slot());
    mMask = capacity - 1;
    mSize = 0;
    bool ins;
    for (const Slot& s : old)
      if (s.used) Emplace(s.key, s.val, &ins);
  }
  std::vector<Slot> mSlots;
  size_t mMask = 0, mSize = 0;
};

int32_t JavaStringHashCode(const std::string& utf8);  // String.hashCode over UTF-16 units
std::string HadoopFileIdentifier(const std::string& pathUtf8);  // md5 hex of the path, HadoopUtils.java:225-238

// String file id -> 128 bits, collision free: decimal longs -> (0, v); 32 lower-case hex
// chars -> (hi, lo) unless hi is 0 or all ones; anything else interned -> (all ones, n).
class FileKeyCodec {
 public:
  void Encode(const std::string& fileId, uint64_t* hi, uint64_t* lo);
  std::string Decode(uint64_t hi, uint64_t lo);
  // Warm restart: the interned ids in index order, and their adoption into a still empty table (aps_snapshot_names).
  std::vector<std::string> Names();
  void AdoptNames(const std::vector<std::string>& names);
 private:
  std::mutex mMu;
  std::unordered_map<std::string, uint64_t> mIds;
  std::vector<std::string> mNames;
};

// ---- index-linked list pool (shared by the evictors) ---------------------------------------------------
struct ListNode {
  Key key;
  int32_t prev, next;
};
class NodePool {
 public:
  int32_t Alloc(const Key& k);
  void Free(int32_t n);
  ListNode& operator[](int32_t n) { return mNodes[(size_t)n]; }
 private:
  std::vector<ListNode> mNodes;
  std::vector<int32_t> mFree;
};
struct ListHead {
  int32_t head = -1, tail = -1;
  bool Empty() const { return head < 0; }
};
void ListPushBack(NodePool& p, ListHead& l, int32_t n);
void ListUnlink(NodePool& p, ListHead& l, int32_t n);

// ---- CACHE/CacheEvictor.java:26-69 ---------------------------------------------------------------------
class CacheEvictor {
 public:
  virtual ~CacheEvictor() {}
  virtual void UpdateOnGet(const Key& k) = 0;
  virtual void UpdateOnPut(const Key& k) = 0;
  virtual void UpdateOnDelete(const Key& k) = 0;
  virtual bool Evict(Key* victim) = 0;  // false == null; the victim is NOT removed
  virtual void Reset() = 0;
  static std::unique_ptr<CacheEvictor> Create(const std::string& kind, double lfuLogBase);
};

// ---- CACHE/DefaultMetaStore.java:33-132 ----------------------------------------------------------------
struct PageMeta {
  int64_t size;
  int32_t lockId;      // LocalCacheManager.getPageLockId of this page, kept so a victim's lock is known
  uint32_t epoch = 0;  // batched replay: the flush window in which the page was last put or read
};
class MetaStore {
 public:
  explicit MetaStore(std::unique_ptr<CacheEvictor> e) : mEvictor(std::move(e)) {}
  bool HasPage(const Key& k) const { return mPages.Find(k) != nullptr; }
  PageMeta* Find(const Key& k) { return mPages.Find(k); }  // hasPage with the entry in hand: no evictor update
  void AddPage(const Key& k, const PageMeta& m);
  bool GetPageInfo(const Key& k, PageMeta* out);  // false == PageNotFoundException; refreshes the evictor
  bool RemovePage(const Key& k);                  // false == PageNotFoundException
  int64_t Bytes() const { return mBytes.load(); }
  int64_t Pages() const { return mCount.load(); }
  void Reset();
  bool Evict(Key* victim, PageMeta* meta);  // false == null
 private:
  FlatMap<PageMeta> mPages;
  std::atomic<int64_t> mBytes{0}, mCount{0};
  std::unique_ptr<CacheEvictor> mEvictor;
};

// ---- the boundary, CACHE/PageStore.java:37-181 -----------------------------------------------------------
class PageStore {
 public:
  virtual ~PageStore() {}
  virtual int Put(const Key& k, const void* page, int32_t len) = 0;  // aps status
  virtual int Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) = 0;
  virtual int Delete(const Key& k) = 0;
  virtual int GetPages(std::vector<std::pair<Key, int64_t>>* out) = 0;
  virtual int64_t GetCacheSize() const = 0;
  // Additive batched forms, aps_*_batch semantics: host request arrays, device or pinned
  // payload pointers, one result per request.  A store without a batched path answers APS_EINVAL.
  virtual int GetBatch(uint32_t, const aps_get_req*, aps_result*) { return APS_EINVAL; }
  virtual int PutBatch(uint32_t, const aps_put_req*, aps_result*) { return APS_EINVAL; }
  virtual int DeleteBatch(uint32_t, const aps_key*, aps_result*) { return APS_EINVAL; }
  // Synthetic payload generated on the store's device (replays): aps_fill_pages semantics.
  virtual int FillPages(void*, uint64_t, uint32_t, uint32_t, const uint64_t*, const uint64_t*, uint64_t) { return APS_EINVAL; }
  virtual int SetFaults(uint32_t) { return APS_EINVAL; }
  virtual int32_t StagingDevice() const { return 0; }
};

// PageStoreType.GPU: the HBM slab + table behind include/aps.h.
class GpuPageStore : public PageStore {
 public:
  static int Open(const aps_config& cfg, std::unique_ptr<GpuPageStore>* out);
  // Wraps a store somebody else opened (and will close): PageStore.open over existing contents.
  static std::unique_ptr<GpuPageStore> Adopt(aps_store* s);
  ~GpuPageStore() override;
  int Put(const Key& k, const void* page, int32_t len) override;
  int Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) override;
  int Delete(const Key& k) override;
  int GetPages(std::vector<std::pair<Key, int64_t>>* out) override;
  int64_t GetCacheSize() const override { return (int64_t)mCacheBytes; }
  int GetBatch(uint32_t n, const aps_get_req* reqs, aps_result* res) override { return aps_get_batch(mStore, n, reqs, res, nullptr); }
  int PutBatch(uint32_t n, const aps_put_req* reqs, aps_result* res) override { return aps_put_batch(mStore, n, reqs, res, nullptr); }
  int DeleteBatch(uint32_t n, const aps_key* keys, aps_result* res) override { return aps_delete_batch(mStore, n, keys, res, nullptr); }
  int FillPages(void* d, uint64_t stride, uint32_t len, uint32_t n, const uint64_t* f, const uint64_t* i, uint64_t seed) override {
    return aps_fill_pages(mStore, d, stride, len, n, f, i, seed, nullptr);
  }
  int SetFaults(uint32_t mask) override { return aps_set_faults(mStore, mask); }
  int32_t StagingDevice() const override { return mDevice; }
  aps_store* handle() { return mStore; }
 private:
  aps_store* mStore = nullptr;
  uint64_t mCacheBytes = 0;
  int32_t mDevice = 0;
  bool mOwned = true;
};

// One cache over several GPUs: a page lives in the store of GPU aps_shard_of(key, n).  Policy
// stays where the reference has it -- ONE LocalCacheManager, one meta store, one global byte budget
// (CACHE/LocalCacheManager.java:261) -- and this class is the PageStore it drives: put and delete go to the
// owner, a batched get is one kernel on shard 0 that reads every owner's table and slab in place.
class ShardedGpuPageStore : public PageStore {
 public:
  // cfg describes the WHOLE cache (cache_bytes = global budget); devices[i] = CUDA ordinal of shard i (several
  // shards may share a device).  cfg.slab_slots, when given, is per shard; otherwise each shard gets its share
  // of the budget plus a quarter (hash routing is only balanced on average) plus 64.
  static int Open(const aps_config& cfg, uint32_t nShards, const int32_t* devices, std::unique_ptr<ShardedGpuPageStore>* out);
  ~ShardedGpuPageStore() override;
  int Put(const Key& k, const void* page, int32_t len) override;
  int Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) override;
  int Delete(const Key& k) override;
  int GetPages(std::vector<std::pair<Key, int64_t>>* out) override;
  int64_t GetCacheSize() const override { return (int64_t)mCacheBytes; }
  int GetBatch(uint32_t n, const aps_get_req* reqs, aps_result* res) override;
  int PutBatch(uint32_t n, const aps_put_req* reqs, aps_result* res) override;
  int DeleteBatch(uint32_t n, const aps_key* keys, aps_result* res) override;
  int FillPages(void* d, uint64_t stride, uint32_t len, uint32_t n, const uint64_t* f, const uint64_t* i, uint64_t seed) override;
  int SetFaults(uint32_t mask) override;
  int32_t StagingDevice() const override { return mShards.empty() ? 0 : mShards[0]->StagingDevice(); }
  uint32_t shards() const { return (uint32_t)mShards.size(); }
  GpuPageStore* shard(uint32_t i) { return mShards[i].get(); }
  uint32_t OwnerOf(const Key& k) const;
 private:
  std::vector<std::unique_ptr<GpuPageStore>> mShards;
  uint64_t mCacheBytes = 0;
};

// new ThreadPoolExecutor(n, n, 60 s, new SynchronousQueue<>()) of CACHE/LocalCacheManager.java:156-161: a task
// is accepted only while a worker is free to take it, otherwise the submit is rejected at once.
class HandOffPool {
 public:
  explicit HandOffPool(int threads);
  ~HandOffPool();  // runs what was accepted, then joins
  bool TrySubmit(std::function<void()> task);
  void WaitIdle();
 private:
  void Work();
  std::mutex mMu;
  std::condition_variable mWake, mQuiet;
  std::deque<std::function<void()>> mQueue;
  std::vector<std::thread> mWorkers;
  int mRunning = 0;
  bool mStop = false;
};

// CACHE/TimeBoundPageStore.java:38-159: every put / get / delete of the wrapped store runs on a fixed pool over
// a synchronous hand-off and is abandoned after timeoutMs (IOException); a call that finds every thread busy is
// refused at once (IOException).  The reference cancels the abandoned call by interrupt; a CUDA call cannot be
// interrupted, so here it runs to its end on private copies of the caller's buffers and its result is dropped.
// The delays are the test double of CTEST/HangingPageStore.java, shortened from "forever".
class TimeBoundPageStore : public PageStore {
 public:
  TimeBoundPageStore(std::unique_ptr<PageStore> store, int64_t timeoutMs, int threads);
  int Put(const Key& k, const void* page, int32_t len) override;
  int Get(const Key& k, int32_t pageOffset, int32_t bytesToRead, void* dst, int32_t room, int32_t* bytesRead) override;
  int Delete(const Key& k) override;
  int GetPages(std::vector<std::pair<Key, int64_t>>* out) override { return mStore->GetPages(out); }
  int64_t GetCacheSize() const override { return mStore->GetCacheSize(); }
  int GetBatch(uint32_t n, const aps_get_req* r, aps_result* res) override { return mStore->GetBatch(n, r, res); }
  int PutBatch(uint32_t n, const aps_put_req* r, aps_result* res) override { return mStore->PutBatch(n, r, res); }
  int DeleteBatch(uint32_t n, const aps_key* k, aps_result* res) override { return mStore->DeleteBatch(n, k, res); }
  int FillPages(void* d, uint64_t stride, uint32_t len, uint32_t n, const uint64_t* f, const uint64_t* i, uint64_t seed) override {
    return mStore->FillPages(d, stride, len, n, f, i, seed);
  }
  int SetFaults(uint32_t mask) override { return mStore->SetFaults(mask); }
  int32_t StagingDevice() const override { return mStore->StagingDevice(); }
  std::atomic<int64_t> storePutTimeout{0}, storeGetTimeout{0}, storeDeleteTimeout{0}, storeThreadsRejected{0};
  std::atomic<int32_t> putDelayMs{0}, getDelayMs{0}, deleteDelayMs{0};  // HangingPageStore
 private:
  int Bounded(std::function<int()> call, std::atomic<int64_t>* timeouts);
  std::shared_ptr<PageStore> mStore;  // shared with calls that outlive their timeout
  int64_t mTimeoutMs;
  std::shared_ptr<HandOffPool> mPool;
};

// Counters of the reference's Client.Cache* metrics that live in this layer
// (core/common/src/main/java/alluxio/metrics/MetricKey.java:778-995; LocalCacheManager.java:557-627).
struct CacheMetrics {
  std::atomic<int64_t> bytesReadCache{0}, bytesRequestedExternal{0}, bytesReadExternal{0}, bytesWrittenCache{0},
      bytesEvicted{0}, pagesEvicted{0}, getErrors{0}, putErrors{0}, deleteErrors{0}, putEvictionErrors{0},
      putStoreWriteErrors{0}, putStoreDeleteErrors{0}, getStoreReadErrors{0}, deleteNonExistingPageErrors{0},
      getNotReadyErrors{0}, deleteNotReadyErrors{0}, deleteStoreDeleteErrors{0}, putBenignRacingErrors{0},
      cleanupGetErrors{0}, cleanupPutErrors{0};
};

// ---- CACHE/LocalCacheManager.java:75-627 -----------------------------------------------------------------
enum State { NOT_IN_USE = 0, READ_ONLY = 1, READ_WRITE = 2 };
constexpr int32_t kGetIllegalArgument = -2;

struct ManagerOptions {
  int asyncWriteThreads = 0;  // alluxio.user.client.cache.async.write.enabled / .threads (0 = synchronous puts)
  bool asyncRestore = false;  // alluxio.user.client.cache.async.restore.enabled
  std::vector<std::string> internedNames;  // intern table of the process that wrote the snapshot being restored
};

class LocalCacheManager {
 public:
  LocalCacheManager(int64_t pageSize, std::unique_ptr<MetaStore> meta, std::unique_ptr<PageStore> store,
                    ManagerOptions options = ManagerOptions());
  ~LocalCacheManager();  // close(): :511-521
  // Blocks until the restore thread has finished and no asynchronous put is queued or running (tests, shutdown).
  void WaitAsync();
  std::atomic<int64_t> putAsyncRejectionErrors{0}, putNotReadyErrors{0};
  bool Put(const std::string& fileId, int64_t pageIndex, const void* page, int32_t len);
  int32_t Get(const std::string& fileId, int64_t pageIndex, int32_t pageOffset, int32_t bytesToRead, void* buffer,
              int32_t bufLen, int32_t offsetInBuffer);
  bool Delete(const std::string& fileId, int64_t pageIndex);
  State state() const { return mState.load(); }
  int64_t pageSize() const { return mPageSize; }
  int32_t GetPageLockId(const std::string& fileId, int64_t pageIndex) const;
  MetaStore* meta() { return mMetaStore.get(); }
  PageStore* store() { return mPageStore.get(); }
  FileKeyCodec* codec() { return &mCodec; }
  // eviction log for replays (cleared by the caller)
  std::vector<Key> evicted;
  bool logEvictions = false;
  CacheMetrics metrics;
  // Several pages of one file in one store call: for pages [firstPage, firstPage + n) asks
  // the store for (offset[i], len[i]) of each page that is resident, into dst[i]; got[i] = bytes read, 0 for a
  // miss, -1 for an error (the page is then dropped, as in get()).  dst must be device or pinned memory.
  void GetMany(const std::string& fileId, int64_t firstPage, int32_t n, const int32_t* offset, const int32_t* len,
               void* const* dst, int32_t* got);
 private:
  static constexpr int kLockSize = 1024;
  Key MakeKey(const std::string& fileId, int64_t pageIndex);
  bool PutInternal(const Key& k, int32_t lockId, const void* page, int32_t len);
  void UndoAddPage(const Key& k);
  void RestoreOrInit();
  bool Restore();
  bool WipeStore();  // PageStore.create over a store that is emptied in place (the reference wipes the cache directory)
  ManagerOptions mOptions;
  std::unique_ptr<HandOffPool> mAsyncCacheExecutor;
  std::thread mInitService;
  std::mutex mPendingMu;
  std::unordered_set<Key, KeyHash> mPendingRequests;
  int64_t mPageSize, mCacheSize;
  std::unique_ptr<MetaStore> mMetaStore;
  std::unique_ptr<PageStore> mPageStore;
  FileKeyCodec mCodec;
  std::shared_mutex mPageLocks[kLockSize];
  std::shared_mutex mMetaLock;
  std::atomic<State> mState;
};

// ---- CACHE/LocalCacheFileInStream.java:43-330 behind CACHE/LocalCacheFileSystem.java:32-71 ---------------
// `external(user, filePos, dst, len) -> bytes read` stands for the FileInStream of the delegated file system.
typedef int (*ExternalRead)(void* user, int64_t filePos, void* dst, int32_t len);
class LocalCacheFileInStream {
 public:
  // LocalCacheFileSystem.openFile: with no manager, or one whose state is NOT_IN_USE, the stream is the
  // external one (bypass: the cache is neither read nor filled).
  LocalCacheFileInStream(LocalCacheManager* manager, const std::string& fileId, int64_t fileLength, ExternalRead external,
                         void* user);
  // read(byte[] b, int off, int len): bytes read, -1 at EOF, kGetIllegalArgument, or -3 for the IOException paths
  int32_t Read(void* b, int32_t bLen, int32_t off, int32_t len);
  int32_t PositionedRead(int64_t pos, void* b, int32_t bLen, int32_t off, int32_t len);
  int64_t Skip(int64_t n);       // -4 on a closed stream (IllegalStateException)
  int Seek(int64_t pos);         // 0, kGetIllegalArgument, or -4 on a closed stream
  int64_t Remaining() const { return mLength - mPosition; }
  int64_t GetPos() const { return mPosition; }
  void Close() { mClosed = true; }
  ~LocalCacheFileInStream();
  // serve every read with ONE batched store call for the resident pages it touches
  void SetBatched(bool on) { mBatched = on && !mBypass; }
  bool bypass() const { return mBypass; }
  int64_t bytesReadCache = 0, bytesRequestedExternal = 0, bytesReadExternal = 0;
 private:
  int32_t ReadAt(int64_t* position, void* b, int32_t bLen, int32_t off, int32_t len);
  bool ReadExternalPage(int64_t pos, std::vector<uint8_t>* page);
  LocalCacheManager* mManager;
  std::string mFileId;
  int64_t mLength, mPageSize, mPosition = 0;
  ExternalRead mExternal;
  void* mUser;
  bool mClosed = false, mBypass, mBatched = false;
  void* mStage = nullptr;  // pinned staging of the batched mode, kept between reads
  uint64_t mStageBytes = 0;
  std::mutex mExternalMu;  // readExternalPage is synchronized (:291)
};

int32_t BatchedRead(LocalCacheManager* mgr, const std::string& fileId, int64_t fileLen, int64_t pos, uint8_t* dst, int32_t len,
                    ExternalRead external, void* user, void** stage, uint64_t* stageBytes, int64_t* fromCache,
                    int64_t* fromExternal, int64_t* readExternal);

// Page payload of the benchmarks and replays: 64-bit word k of page (file, index) is the k-th
// splitmix64 output for seed ^ (file << 32) ^ index.
void FillSyntheticPage(uint64_t seed, uint64_t file, uint64_t index, uint8_t* out, int64_t len);

}  // namespace alluxio_b200
