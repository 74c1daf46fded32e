// ORACLE -- TEST INFRASTRUCTURE ONLY (see checksums.c header).
//
// CPU restatement, class by class, of the reference's client-side page cache.  "CACHE/" below is
// /root/reference/core/client/fs/src/main/java/alluxio/client/file/cache/.  The reference cannot
// run here (no JVM), so this restatement is pinned by porting the reference's own
// unit tests (tests/test_oracle_*.py cite each one).
//
// Single-threaded by design: bit-exact replay is defined for a serialised operation sequence
// (SURVEY.md §7 "Determinism under concurrency"); lock acquisition in the reference has no effect
// on the result of a serialised sequence, so locks are noted but not modelled.
#pragma once
#include <cstdint>
#include <functional>
#include <list>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

namespace orc {

// ---- Java-isms -----------------------------------------------------------------------------
int32_t JavaStringHashCode(const std::string& utf8);  // String.hashCode over UTF-16 code units
int32_t FloorMod(int32_t x, int32_t y);               // Math.floorMod(int,int)

// ---- CACHE/PageId.java:23-74, CACHE/PageInfo.java:24-75 --------------------------------------
struct PageId {
  std::string fileId;
  int64_t pageIndex;
  bool operator==(const PageId& o) const { return fileId == o.fileId && pageIndex == o.pageIndex; }
};
struct PageIdHash {
  size_t operator()(const PageId& p) const {
    return std::hash<std::string>()(p.fileId) * 1000003u ^ std::hash<int64_t>()(p.pageIndex);
  }
};
struct PageInfo {
  PageId pageId;
  int64_t pageSize;  // actual byte length of this page
};

// java.util.LinkedHashMap<PageId,Boolean>(accessOrder=true) reduced to the calls the evictors make.
class AccessOrderedSet {
 public:
  void Put(const PageId& k);          // insert at tail, or move existing to tail
  bool Get(const PageId& k);          // move to tail if present; false if absent
  bool Remove(const PageId& k);       // true if it was present
  bool Empty() const { return mOrder.empty(); }
  const PageId& First() const { return mOrder.front(); }
  void Clear();
 private:
  std::list<PageId> mOrder;
  std::unordered_map<PageId, std::list<PageId>::iterator, PageIdHash> mPos;
};

// ---- CACHE/CacheEvictor.java:26-69 -----------------------------------------------------------
class CacheEvictor {
 public:
  virtual ~CacheEvictor() {}
  virtual void UpdateOnGet(const PageId&) = 0;
  virtual void UpdateOnPut(const PageId&) = 0;
  virtual void UpdateOnDelete(const PageId&) = 0;
  virtual bool Evict(PageId* out) = 0;  // false == Java null
  virtual void Reset() = 0;
};
std::unique_ptr<CacheEvictor> MakeEvictor(const std::string& kind, double lfuLogBase);

// ---- CACHE/MetaStore.java:19-75, CACHE/DefaultMetaStore.java:33-132 ------------------------------
class DefaultMetaStore {
 public:
  explicit DefaultMetaStore(std::unique_ptr<CacheEvictor> e) : mEvictor(std::move(e)) {}
  bool HasPage(const PageId& id) const { return mPageMap.count(id) != 0; }
  void AddPage(const PageId& id, const PageInfo& info);
  bool GetPageInfo(const PageId& id, PageInfo* out);  // false == PageNotFoundException
  bool RemovePage(const PageId& id);                  // false == PageNotFoundException
  int64_t Bytes() const { return mBytes; }
  int64_t Pages() const { return mPages; }
  void Reset();
  bool Evict(PageInfo* out);  // false == null
  CacheEvictor* evictor() { return mEvictor.get(); }
 private:
  std::unordered_map<PageId, PageInfo, PageIdHash> mPageMap;
  int64_t mBytes = 0, mPages = 0;
  std::unique_ptr<CacheEvictor> mEvictor;
};

// ---- CACHE/PageStore.java:37-181 --------------------------------------------------------------
enum StoreStatus { kOk = 0, kNotFound = 1 /*PageNotFoundException*/, kIllegalArgument = 2, kIo = 3 };

class PageStore {
 public:
  virtual ~PageStore() {}
  virtual StoreStatus Put(const PageId& id, const uint8_t* page, int32_t len) = 0;
  // buffer has bufLen bytes in total (Java buffer.length).
  virtual StoreStatus Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer,
                          int32_t bufLen, int32_t bufferOffset, int32_t* bytesRead) = 0;
  virtual StoreStatus Delete(const PageId& id) = 0;
  // Each element: ok=false models a null PageInfo in the stream (unrecognised file).
  virtual StoreStatus GetPages(std::vector<std::pair<bool, PageInfo>>* out) = 0;
  virtual int64_t GetCacheSize() const = 0;
  virtual bool RootExists() const = 0;  // LocalCacheManager.restore: Files.exists(rootDir)
  virtual StoreStatus Recreate() = 0;   // close() + PageStore.create(options): wipe + reopen
  // fault injection used by the ported tests (FaultyPageStore, LocalCacheManagerTest.java:660-691)
  bool putFaulty = false, deleteFaulty = false;
};

// CACHE/store/LocalPageStore.java:44-212 on a real directory.
std::unique_ptr<PageStore> OpenLocalPageStore(const std::string& root, int64_t pageSize, int64_t cacheSize,
                                              int32_t fileBuckets, bool wipeFirst);
// Same contract, pages held in RAM (for replays too large for one-file-per-page on disk).
std::unique_ptr<PageStore> OpenMemPageStore(int64_t pageSize, int64_t cacheSize, bool keepBytes);

// A PageStore implemented OUTSIDE the oracle, reached through the five calls of the interface
// (CACHE/PageStore.java:119-171).  This is how the tests put the oracle's LocalCacheManager -- the restatement of the
// reference's Java -- on top of the store under test, exactly as the JVM puts the real LocalCacheManager on top of
// GpuPageStore: the manager's own logic is the checker's, only the PageStore behind it is the product.
// Each callback returns a StoreStatus; get_pages calls emit(emit_user, valid, fileId, pageIndex, size) per page.
struct ExternalPageStoreCalls {
  void* user;
  int (*put)(void* user, const char* fileId, int64_t pageIndex, const uint8_t* page, int32_t len);
  int (*get)(void* user, const char* fileId, int64_t pageIndex, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer,
             int32_t bufLen, int32_t bufferOffset, int32_t* bytesRead);
  int (*del)(void* user, const char* fileId, int64_t pageIndex);
  int (*get_pages)(void* user, void (*emit)(void* emitUser, int valid, const char* fileId, int64_t pageIndex, int64_t size),
                   void* emitUser);
};
std::unique_ptr<PageStore> OpenExternalPageStore(const ExternalPageStoreCalls& calls, int64_t cacheSize);

// ---- CACHE/CacheManager.java:34-188, CACHE/LocalCacheManager.java:75-627 ------------------------
enum State { NOT_IN_USE = 0, READ_ONLY = 1, READ_WRITE = 2 };
constexpr int kGetIllegalArgument = -2;  // stands for IllegalArgumentException out of get()

class LocalCacheManager {
 public:
  // = LocalCacheManager.create(conf, metaStore, pageStore) with sync restore (LocalCacheManager.java:118-138)
  LocalCacheManager(int64_t pageSize, std::unique_ptr<DefaultMetaStore> meta, PageStore* store);
  bool Put(const PageId& id, const uint8_t* page, int32_t len);
  int32_t Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer, int32_t bufLen,
              int32_t offsetInBuffer);
  bool Delete(const PageId& id);
  State state() const { return mState; }
  int32_t GetPageLockId(const PageId& id) const;  // LocalCacheManager.java:172-174
  DefaultMetaStore* meta() { return mMetaStore.get(); }
  // victims evicted so far (pageIndex + fileId), for replay comparison
  std::vector<PageId> evicted;
  // The Client.Cache* counters of LocalCacheManager.Metrics (:557-627), incremented where the reference increments them
  // (the line of each increment is cited in page_cache.cc).  Names: MetricKey.CLIENT_CACHE_<NAME>.
  struct Metrics {
    int64_t bytesWrittenCache = 0, bytesEvicted = 0, pagesEvicted = 0, getErrors = 0, getNotReadyErrors = 0,
            getStoreReadErrors = 0, cleanupGetErrors = 0, putErrors = 0, putNotReadyErrors = 0, putEvictionErrors = 0,
            putStoreWriteErrors = 0, putStoreDeleteErrors = 0, putBenignRacingErrors = 0, cleanupPutErrors = 0,
            deleteErrors = 0, deleteNotReadyErrors = 0, deleteNonExistingPageErrors = 0, deleteStoreDeleteErrors = 0;
  } metrics;
 private:
  bool PutInternal(const PageId& id, const uint8_t* page, int32_t len);
  void UndoAddPage(const PageId& id);
  void RestoreOrInit();
  bool Restore();
  int64_t mPageSize, mCacheSize;
  std::unique_ptr<DefaultMetaStore> mMetaStore;
  PageStore* mPageStore;
  State mState;
};

// ---- CACHE/LocalCacheFileInStream.java:37-336 ---------------------------------------------------
// The stream a client reads a file through: page by page out of the cache, whole pages from the external
// file system on a miss (and offered to the cache).  `ExternalRead(pos, dst, len) -> bytes read` stands for
// the external FileInStream after getExternalFileInStream() has positioned it (:256-270); <= 0 ends the read
// loop of readExternalPage (:299-303).  Results follow the Java ones with exceptions turned into codes:
// kGetIllegalArgument (Preconditions.checkArgument), kStreamIo (IOException), kStreamClosed
// (IllegalStateException from checkIfClosed).  The meters of :312-335 are process-wide statics in the
// reference; here they are per stream.
constexpr int kStreamIo = -3;
constexpr int kStreamClosed = -4;
using ExternalRead = std::function<int32_t(int64_t pos, uint8_t* dst, int32_t len)>;

class LocalCacheFileInStream {
 public:
  // pageSize = alluxio.user.client.cache.page.size (:86); fileId = URIStatus.getFileIdentifier() (:128).
  LocalCacheFileInStream(const std::string& fileId, int64_t fileLength, int64_t pageSize, LocalCacheManager* manager,
                         ExternalRead external);
  int32_t Read(uint8_t* b, int32_t bLen, int32_t off, int32_t len);                         // :108-152
  int32_t PositionedRead(int64_t pos, uint8_t* b, int32_t bLen, int32_t off, int32_t len);  // :175-223
  int64_t Skip(int64_t n);                                                                  // :154-162
  int Seek(int64_t pos);                                                                    // :230-243
  int64_t Remaining() const { return mEOF ? 0 : mLength - mPosition; }                      // :170-172
  int64_t GetPos() const { return mPosition; }                                              // :225-228
  void Close() { mClosed = true; }  // :164-167 closes the external stream; the tests here only see the flag
  int64_t bytesReadCache = 0, bytesRequestedExternal = 0, bytesReadExternal = 0;            // :312-320
  int64_t externalReads = 0;  // calls made on the external stream (not a reference metric; used by the tests)
 private:
  // the page loop both reads share; *position is mPosition (read) or a local copy of pos (positionedRead)
  int32_t ReadLoop(int64_t* position, uint8_t* b, int32_t bLen, int32_t off, int32_t len, bool positioned);
  bool ReadExternalPage(int64_t pos, std::vector<uint8_t>* page);                           // :291-310
  std::string mFileId;
  int64_t mLength, mPageSize;
  LocalCacheManager* mCacheManager;
  ExternalRead mExternal;
  int64_t mPosition = 0;
  bool mClosed = false, mEOF = false;
};

}  // namespace orc
