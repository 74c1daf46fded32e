// ORACLE -- TEST INFRASTRUCTURE ONLY (see checksums.c header and page_cache.h).
#include "page_cache.h"

#include <dirent.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstring>

namespace orc {

// ============================================================================ Java-isms

// java.lang.String.hashCode: s[0]*31^(n-1) + ... + s[n-1] over UTF-16 code units, int32 wraparound.
// Used at CACHE/LocalCacheManager.java:173 and CACHE/store/LocalPageStore.java:146.
int32_t JavaStringHashCode(const std::string& s) {
  uint32_t h = 0;
  size_t i = 0, n = s.size();
  while (i < n) {
    uint32_t c = (uint8_t)s[i], cp;
    int extra;
    if (c < 0x80) { cp = c; extra = 0; }
    else if ((c >> 5) == 0x6) { cp = c & 0x1F; extra = 1; }
    else if ((c >> 4) == 0xE) { cp = c & 0x0F; extra = 2; }
    else if ((c >> 3) == 0x1E) { cp = c & 0x07; extra = 3; }
    else { cp = 0xFFFD; extra = 0; }
    i++;
    for (int k = 0; k < extra && i < n; k++, i++) cp = (cp << 6) | ((uint8_t)s[i] & 0x3F);
    if (cp >= 0x10000) {  // surrogate pair
      cp -= 0x10000;
      h = 31u * h + (0xD800u + (cp >> 10));
      h = 31u * h + (0xDC00u + (cp & 0x3FF));
    } else {
      h = 31u * h + cp;
    }
  }
  return (int32_t)h;
}

// java.lang.Math.floorMod(int,int)
int32_t FloorMod(int32_t x, int32_t y) {
  int32_t m = x % y;
  if ((m ^ y) < 0 && m != 0) m += y;
  return m;
}

// ============================================================================ AccessOrderedSet

void AccessOrderedSet::Put(const PageId& k) {
  auto it = mPos.find(k);
  if (it != mPos.end()) {
    mOrder.splice(mOrder.end(), mOrder, it->second);
  } else {
    mOrder.push_back(k);
    mPos.emplace(k, std::prev(mOrder.end()));
  }
}
bool AccessOrderedSet::Get(const PageId& k) {
  auto it = mPos.find(k);
  if (it == mPos.end()) return false;
  mOrder.splice(mOrder.end(), mOrder, it->second);
  return true;
}
bool AccessOrderedSet::Remove(const PageId& k) {
  auto it = mPos.find(k);
  if (it == mPos.end()) return false;
  mOrder.erase(it->second);
  mPos.erase(it);
  return true;
}
void AccessOrderedSet::Clear() {
  mOrder.clear();
  mPos.clear();
}

// ============================================================================ evictors

// CACHE/evictor/LRUCacheEvictor.java:29-75
class LRUCacheEvictor : public CacheEvictor {
 public:
  void UpdateOnGet(const PageId& id) override { mLRUCache.Put(id); }     // :49-51
  void UpdateOnPut(const PageId& id) override { mLRUCache.Put(id); }     // :54-56
  void UpdateOnDelete(const PageId& id) override { mLRUCache.Remove(id); }  // :59-61
  bool Evict(PageId* out) override {                                      // :65-69 head, NOT removed
    if (mLRUCache.Empty()) return false;
    *out = mLRUCache.First();
    return true;
  }
  void Reset() override { mLRUCache.Clear(); }
 private:
  AccessOrderedSet mLRUCache;
};

// CACHE/evictor/LFUCacheEvictor.java:35-161
class LFUCacheEvictor : public CacheEvictor {
 public:
  explicit LFUCacheEvictor(double logBase) : mDivisor(std::log(logBase)) {}  // :56
  void UpdateOnGet(const PageId& id) override {                               // :98-116
    int newCount = ++mPageMap[id];
    int newBucket = GetBucket(newCount);
    if (newCount > 1) {
      int oldBucket = GetBucket(newCount - 1);
      if (newBucket != oldBucket) {
        bool pagesLeft = RemovePageFromBucket(id, oldBucket);
        if (!pagesLeft && oldBucket == mMinBucket) mMinBucket = newBucket;
        AddPageToBucket(id, newBucket);
      } else {
        TouchPageInBucket(id, newBucket);
      }
    } else {
      mMinBucket = newBucket;
      AddPageToBucket(id, newBucket);
    }
  }
  void UpdateOnPut(const PageId& id) override { UpdateOnGet(id); }  // :119-121
  void UpdateOnDelete(const PageId& id) override {                   // :124-138
    auto it = mPageMap.find(id);
    if (it == mPageMap.end()) return;
    int count = it->second;
    mPageMap.erase(it);
    int bucket = GetBucket(count);
    bool pagesLeft = RemovePageFromBucket(id, bucket);
    if (!pagesLeft && bucket == mMinBucket && !mBucketMap.empty()) {
      while (!mBucketMap.count(mMinBucket)) mMinBucket++;
    }
  }
  bool Evict(PageId* out) override {  // :142-152
    auto it = mBucketMap.find(mMinBucket);
    if (it == mBucketMap.end()) return false;
    *out = it->second.First();
    return true;
  }
  void Reset() override {  // :155-159
    mPageMap.clear();
    mBucketMap.clear();
    mMinBucket = -1;
  }
 private:
  int GetBucket(int count) const { return (int)(std::log((double)count) / mDivisor); }  // :61-63
  void AddPageToBucket(const PageId& id, int bucket) { mBucketMap[bucket].Put(id); }   // :65-74
  // Returns whether the bucket map still holds this bucket afterwards; computeIfPresent returns null
  // both when the bucket vanished and when it never existed (:76-85).
  bool RemovePageFromBucket(const PageId& id, int bucket) {
    auto it = mBucketMap.find(bucket);
    if (it == mBucketMap.end()) return false;
    it->second.Remove(id);
    if (it->second.Empty()) {
      mBucketMap.erase(it);
      return false;
    }
    return true;
  }
  void TouchPageInBucket(const PageId& id, int bucket) {  // :87-96 lruMap.get() refreshes recency
    auto it = mBucketMap.find(bucket);
    if (it != mBucketMap.end()) it->second.Get(id);
  }
  std::unordered_map<PageId, int, PageIdHash> mPageMap;
  std::map<int, AccessOrderedSet> mBucketMap;
  int mMinBucket = -1;
  double mDivisor;
};

// FIFOEvictor, the test double at CTEST/LocalCacheManagerTest.java:731-761.
class FIFOEvictor : public CacheEvictor {
 public:
  void UpdateOnGet(const PageId&) override {}
  void UpdateOnPut(const PageId& id) override { mQueue.push_back(id); }
  void UpdateOnDelete(const PageId& id) override {
    for (auto it = mQueue.begin(); it != mQueue.end(); ++it)
      if (*it == id) { mQueue.erase(it); return; }
  }
  bool Evict(PageId* out) override {
    if (mQueue.empty()) return false;
    *out = mQueue.front();
    return true;
  }
  void Reset() override { mQueue.clear(); }
 private:
  std::list<PageId> mQueue;
};

std::unique_ptr<CacheEvictor> MakeEvictor(const std::string& kind, double lfuLogBase) {
  if (kind == "LRU") return std::unique_ptr<CacheEvictor>(new LRUCacheEvictor());
  if (kind == "LFU") return std::unique_ptr<CacheEvictor>(new LFUCacheEvictor(lfuLogBase));
  if (kind == "FIFO") return std::unique_ptr<CacheEvictor>(new FIFOEvictor());
  return nullptr;
}

// ============================================================================ DefaultMetaStore

void DefaultMetaStore::AddPage(const PageId& id, const PageInfo& info) {  // DefaultMetaStore.java:56-63
  mPageMap[id] = info;
  mBytes += info.pageSize;
  mPages += 1;
  mEvictor->UpdateOnPut(id);
}
bool DefaultMetaStore::GetPageInfo(const PageId& id, PageInfo* out) {  // :66-72
  auto it = mPageMap.find(id);
  if (it == mPageMap.end()) return false;
  mEvictor->UpdateOnGet(id);
  *out = it->second;
  return true;
}
bool DefaultMetaStore::RemovePage(const PageId& id) {  // :75-85
  auto it = mPageMap.find(id);
  if (it == mPageMap.end()) return false;
  mBytes -= it->second.pageSize;
  mPages -= 1;
  mPageMap.erase(it);
  mEvictor->UpdateOnDelete(id);
  return true;
}
void DefaultMetaStore::Reset() {  // :98-105
  mPages = 0;
  mBytes = 0;
  mPageMap.clear();
  mEvictor->Reset();
}
bool DefaultMetaStore::Evict(PageInfo* out) {  // :108-122
  PageId victim;
  if (!mEvictor->Evict(&victim)) return false;
  auto it = mPageMap.find(victim);
  if (it == mPageMap.end()) {
    mEvictor->UpdateOnDelete(victim);
    return false;
  }
  *out = it->second;
  return true;
}

// ============================================================================ LocalPageStore

namespace {

std::string StripTrailingSlash(std::string s) {
  while (s.size() > 1 && s.back() == '/') s.pop_back();
  return s;
}
bool MkdirP(const std::string& path) {
  std::string cur;
  for (size_t i = 0; i <= path.size(); i++) {
    if (i == path.size() || path[i] == '/') {
      if (!cur.empty() && mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return false;
    }
    if (i < path.size()) cur.push_back(path[i]);
  }
  struct stat st;
  return stat(path.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}
bool RemoveRecursively(const std::string& path) {
  struct stat st;
  if (lstat(path.c_str(), &st) != 0) return errno == ENOENT;
  if (S_ISDIR(st.st_mode)) {
    DIR* d = opendir(path.c_str());
    if (!d) return false;
    bool ok = true;
    while (struct dirent* e = readdir(d)) {
      if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
      ok = RemoveRecursively(path + "/" + e->d_name) && ok;
    }
    closedir(d);
    return rmdir(path.c_str()) == 0 && ok;
  }
  return unlink(path.c_str()) == 0;
}
bool AllDigits(const std::string& s) {
  if (s.empty()) return false;
  for (char c : s) if (c < '0' || c > '9') return false;
  return true;
}
// Long.parseLong on a \d+ string; false == NumberFormatException (overflow)
bool ParseLong(const std::string& s, int64_t* out) {
  uint64_t v = 0;
  for (char c : s) {
    uint64_t d = (uint64_t)(c - '0');
    if (v > (0x7FFFFFFFFFFFFFFFull - d) / 10) return false;
    v = v * 10 + d;
  }
  *out = (int64_t)v;
  return true;
}

class LocalPageStore : public PageStore {
 public:
  LocalPageStore(const std::string& root, int64_t pageSize, int64_t cacheSize, int32_t fileBuckets)
      : mRoot(StripTrailingSlash(root)), mPageSize(pageSize), mCacheSize(cacheSize), mFileBuckets(fileBuckets) {}

  // PageStore.initialize (CACHE/PageStore.java:97-111): create root, delete everything inside it.
  bool Initialize() {
    if (!MkdirP(mRoot)) return false;
    DIR* d = opendir(mRoot.c_str());
    if (!d) return false;
    std::vector<std::string> names;
    while (struct dirent* e = readdir(d))
      if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) names.push_back(e->d_name);
    closedir(d);
    for (auto& n : names) RemoveRecursively(mRoot + "/" + n);  // failures only bump a metric (:104-107)
    return true;
  }

  // LocalPageStore.java:68-86
  StoreStatus Put(const PageId& id, const uint8_t* page, int32_t len) override {
    if (putFaulty) return kIo;
    std::string p = GetFilePath(id);
    struct stat st;
    if (stat(p.c_str(), &st) != 0) {
      std::string parent = p.substr(0, p.rfind('/'));
      if (!MkdirP(parent)) return kIo;
      int fd = open(p.c_str(), O_CREAT | O_EXCL | O_WRONLY, 0666);
      if (fd < 0) return kIo;
      close(fd);
    }
    int fd = open(p.c_str(), O_WRONLY | O_TRUNC | O_CREAT, 0666);
    bool ok = fd >= 0;
    int32_t done = 0;
    while (ok && done < len) {
      ssize_t w = write(fd, page + done, (size_t)(len - done));
      if (w <= 0) ok = false; else done += (int32_t)w;
    }
    if (fd >= 0) close(fd);
    if (!ok) {
      unlink(p.c_str());  // :82-85 leave no partial page
      return kIo;
    }
    return kOk;
  }

  // LocalPageStore.java:88-121
  StoreStatus Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer, int32_t bufLen,
                  int32_t bufferOffset, int32_t* bytesRead) override {
    *bytesRead = 0;
    if (pageOffset < 0) return kIllegalArgument;           // :91
    if (!(bufLen >= bufferOffset)) return kIllegalArgument;  // :92-93
    std::string p = GetFilePath(id);
    struct stat st;
    if (stat(p.c_str(), &st) != 0) return kNotFound;        // :95-97  (Files.exists)
    int64_t pageLength = st.st_size;                         // :98     (File.length)
    if (!(pageOffset <= pageLength)) return kIllegalArgument;  // :99-100
    int fd = open(p.c_str(), O_RDONLY);                      // :101    (new RandomAccessFile)
    if (fd < 0) return kIo;
    if (lseek(fd, pageOffset, SEEK_SET) != pageOffset) {     // :102-107 (skipBytes)
      close(fd);
      return kIo;
    }
    int64_t bytesLeft64 = std::min<int64_t>(pageLength - pageOffset, (int64_t)bufLen - bufferOffset);  // :109
    int32_t bytesLeft = (int32_t)bytesLeft64;
    bytesLeft = std::min(bytesLeft, bytesToRead);            // :110
    int32_t total = 0;
    while (bytesLeft >= 0) {                                 // :111-118
      ssize_t got = bytesLeft == 0 ? 0 : read(fd, buffer + bufferOffset + total, (size_t)bytesLeft);
      if (got <= 0) break;
      total += (int32_t)got;
      bytesLeft -= (int32_t)got;
    }
    close(fd);
    *bytesRead = total;
    return kOk;
  }

  // LocalPageStore.java:123-137
  StoreStatus Delete(const PageId& id) override {
    if (deleteFaulty) return kIo;
    std::string p = GetFilePath(id);
    struct stat st;
    if (stat(p.c_str(), &st) != 0) return kNotFound;
    if (unlink(p.c_str()) != 0) return kIo;
    std::string parent = p.substr(0, p.rfind('/'));
    DIR* d = opendir(parent.c_str());
    if (!d) return kIo;
    bool empty = true;
    while (struct dirent* e = readdir(d))
      if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) { empty = false; break; }
    closedir(d);
    if (empty && rmdir(parent.c_str()) != 0) return kIo;
    return kOk;
  }

  // LocalPageStore.java:149-205: walk, keep regular files, map each through getPageInfo.
  StoreStatus GetPages(std::vector<std::pair<bool, PageInfo>>* out) override {
    out->clear();
    return Walk(mRoot, out) ? kOk : kIo;
  }

  int64_t GetCacheSize() const override { return mCacheSize; }
  bool RootExists() const override {
    struct stat st;
    return stat(mRoot.c_str(), &st) == 0;
  }
  StoreStatus Recreate() override { return Initialize() ? kOk : kIo; }

 private:
  // :139-147  root/<pageSize>/<floorMod(fileId.hashCode(), buckets)>/<fileId>/<pageIndex>
  std::string GetFileBucket(const std::string& fileId) const {
    return std::to_string(FloorMod(JavaStringHashCode(fileId), mFileBuckets));
  }
  std::string GetFilePath(const PageId& id) const {
    return mRoot + "/" + std::to_string(mPageSize) + "/" + GetFileBucket(id.fileId) + "/" + id.fileId + "/" +
           std::to_string(id.pageIndex);
  }
  // :154-171 getPageId: regex root/<pageSize>/(\d+)/([^/]+)/(\d+) must match the WHOLE path
  bool GetPageId(const std::string& path, PageId* out) const {
    std::string prefix = mRoot + "/" + std::to_string(mPageSize) + "/";
    if (path.compare(0, prefix.size(), prefix) != 0) return false;
    std::string rest = path.substr(prefix.size());
    size_t a = rest.find('/');
    if (a == std::string::npos) return false;
    size_t b = rest.find('/', a + 1);
    if (b == std::string::npos) return false;
    if (rest.find('/', b + 1) != std::string::npos) return false;
    std::string bucket = rest.substr(0, a), fileId = rest.substr(a + 1, b - a - 1), name = rest.substr(b + 1);
    if (!AllDigits(bucket) || fileId.empty() || !AllDigits(name)) return false;
    if (bucket != GetFileBucket(fileId)) return false;  // :162-164
    int64_t idx;
    if (!ParseLong(name, &idx)) return false;           // :168-170
    out->fileId = fileId;
    out->pageIndex = idx;
    return true;
  }
  bool Walk(const std::string& dir, std::vector<std::pair<bool, PageInfo>>* out) const {
    DIR* d = opendir(dir.c_str());
    if (!d) return false;
    bool ok = true;
    while (struct dirent* e = readdir(d)) {
      if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
      std::string p = dir + "/" + e->d_name;
      struct stat st;
      if (stat(p.c_str(), &st) != 0) continue;
      if (S_ISDIR(st.st_mode)) {
        ok = Walk(p, out) && ok;
      } else if (S_ISREG(st.st_mode)) {
        PageInfo info;
        if (GetPageId(p, &info.pageId)) {  // :174-190 getPageInfo
          info.pageSize = st.st_size;
          out->push_back({true, info});
        } else {
          out->push_back({false, PageInfo()});
        }
      }
    }
    closedir(d);
    return ok;
  }

  std::string mRoot;
  int64_t mPageSize, mCacheSize;
  int32_t mFileBuckets;
};

// The LocalPageStore contract with pages in RAM, so a 10M-id replay (BASELINE config 3) does not
// need one file per page.  Same status codes and the same length arithmetic as LocalPageStore.get.
class MemPageStore : public PageStore {
 public:
  MemPageStore(int64_t pageSize, int64_t cacheSize, bool keepBytes)
      : mPageSize(pageSize), mCacheSize(cacheSize), mKeepBytes(keepBytes) {}
  StoreStatus Put(const PageId& id, const uint8_t* page, int32_t len) override {
    if (putFaulty) return kIo;
    Entry& e = mPages[id];
    e.len = len;
    if (mKeepBytes) e.bytes.assign(page, page + len);
    return kOk;
  }
  StoreStatus Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer, int32_t bufLen,
                  int32_t bufferOffset, int32_t* bytesRead) override {
    *bytesRead = 0;
    if (pageOffset < 0) return kIllegalArgument;
    if (!(bufLen >= bufferOffset)) return kIllegalArgument;
    auto it = mPages.find(id);
    if (it == mPages.end()) return kNotFound;
    int64_t pageLength = it->second.len;
    if (!(pageOffset <= pageLength)) return kIllegalArgument;
    int32_t bytesLeft = (int32_t)std::min<int64_t>(pageLength - pageOffset, (int64_t)bufLen - bufferOffset);
    bytesLeft = std::min(bytesLeft, bytesToRead);
    if (bytesLeft > 0) {
      if (mKeepBytes && buffer) memcpy(buffer + bufferOffset, it->second.bytes.data() + pageOffset, (size_t)bytesLeft);
      *bytesRead = bytesLeft;
    }
    return kOk;
  }
  StoreStatus Delete(const PageId& id) override {
    if (deleteFaulty) return kIo;
    return mPages.erase(id) ? kOk : kNotFound;
  }
  StoreStatus GetPages(std::vector<std::pair<bool, PageInfo>>* out) override {
    out->clear();
    for (auto& kv : mPages) out->push_back({true, PageInfo{kv.first, kv.second.len}});
    return kOk;
  }
  int64_t GetCacheSize() const override { return mCacheSize; }
  bool RootExists() const override { return true; }
  StoreStatus Recreate() override {
    mPages.clear();
    return kOk;
  }
 private:
  struct Entry { int64_t len = 0; std::vector<uint8_t> bytes; };
  std::unordered_map<PageId, Entry, PageIdHash> mPages;
  int64_t mPageSize, mCacheSize;
  bool mKeepBytes;
};

// See page_cache.h: the store lives outside; fault injection of the ported tests still works from this side.
class ExternalPageStore : public PageStore {
 public:
  ExternalPageStore(const ExternalPageStoreCalls& calls, int64_t cacheSize) : mCalls(calls), mCacheSize(cacheSize) {}
  StoreStatus Put(const PageId& id, const uint8_t* page, int32_t len) override {
    if (putFaulty) return kIo;
    return (StoreStatus)mCalls.put(mCalls.user, id.fileId.c_str(), id.pageIndex, page, len);
  }
  StoreStatus Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer, int32_t bufLen,
                  int32_t bufferOffset, int32_t* bytesRead) override {
    *bytesRead = 0;
    return (StoreStatus)mCalls.get(mCalls.user, id.fileId.c_str(), id.pageIndex, pageOffset, bytesToRead, buffer, bufLen,
                                   bufferOffset, bytesRead);
  }
  StoreStatus Delete(const PageId& id) override {
    if (deleteFaulty) return kIo;
    return (StoreStatus)mCalls.del(mCalls.user, id.fileId.c_str(), id.pageIndex);
  }
  StoreStatus GetPages(std::vector<std::pair<bool, PageInfo>>* out) override {
    out->clear();
    return (StoreStatus)mCalls.get_pages(
        mCalls.user,
        [](void* emitUser, int valid, const char* fileId, int64_t pageIndex, int64_t size) {
          auto* v = static_cast<std::vector<std::pair<bool, PageInfo>>*>(emitUser);
          v->push_back({valid != 0, PageInfo{PageId{fileId, pageIndex}, size}});
        },
        out);
  }
  int64_t GetCacheSize() const override { return mCacheSize; }
  bool RootExists() const override { return true; }
  StoreStatus Recreate() override {  // PageStore.create over a store that cannot be wiped in place: delete what it lists
    std::vector<std::pair<bool, PageInfo>> pages;
    if (GetPages(&pages) != kOk) return kIo;
    for (auto& e : pages)
      if (e.first && mCalls.del(mCalls.user, e.second.pageId.fileId.c_str(), e.second.pageId.pageIndex) != kOk) return kIo;
    return kOk;
  }
 private:
  ExternalPageStoreCalls mCalls;
  int64_t mCacheSize;
};

}  // namespace

std::unique_ptr<PageStore> OpenExternalPageStore(const ExternalPageStoreCalls& calls, int64_t cacheSize) {
  return std::unique_ptr<PageStore>(new ExternalPageStore(calls, cacheSize));
}

std::unique_ptr<PageStore> OpenLocalPageStore(const std::string& root, int64_t pageSize, int64_t cacheSize,
                                              int32_t fileBuckets, bool wipeFirst) {
  auto* s = new LocalPageStore(root, pageSize, cacheSize, fileBuckets);
  if (wipeFirst && !s->Initialize()) {  // PageStore.create = initialize + open (PageStore.java:49-52)
    delete s;
    return nullptr;
  }
  return std::unique_ptr<PageStore>(s);
}
std::unique_ptr<PageStore> OpenMemPageStore(int64_t pageSize, int64_t cacheSize, bool keepBytes) {
  return std::unique_ptr<PageStore>(new MemPageStore(pageSize, cacheSize, keepBytes));
}

// ============================================================================ LocalCacheManager

LocalCacheManager::LocalCacheManager(int64_t pageSize, std::unique_ptr<DefaultMetaStore> meta, PageStore* store)
    : mPageSize(pageSize), mCacheSize(store->GetCacheSize()), mMetaStore(std::move(meta)), mPageStore(store),
      mState(READ_ONLY) {  // LocalCacheManager.java:145-166
  RestoreOrInit();          // :134 (sync restore)
}

int32_t LocalCacheManager::GetPageLockId(const PageId& id) const {  // :172-174
  // (int)(hashCode + pageIndex): int widened to long, added, narrowed back to int.
  int64_t sum = (int64_t)JavaStringHashCode(id.fileId) + id.pageIndex;  // wraps like Java long
  return FloorMod((int32_t)(uint32_t)(uint64_t)sum, 1024);
}

bool LocalCacheManager::Put(const PageId& id, const uint8_t* page, int32_t len) {  // :206-246, sync write
  if (mState != READ_WRITE) {
    metrics.putNotReadyErrors++;                                       // :209
    metrics.putErrors++;                                               // :210
    return false;
  }
  const bool ok = PutInternal(id, page, len);
  if (!ok) metrics.putErrors++;                                        // :216-218
  return ok;
}

bool LocalCacheManager::PutInternal(const PageId& id, const uint8_t* page, int32_t len) {  // :249-347
  PageInfo victim;
  bool enoughSpace;
  if (mMetaStore->HasPage(id)) return true;                            // :256-260 no overwrite
  enoughSpace = mMetaStore->Bytes() + len <= mCacheSize;               // :261
  if (enoughSpace) {
    mMetaStore->AddPage(id, PageInfo{id, len});                        // :263
  } else {
    if (!mMetaStore->Evict(&victim)) {                                 // :265-271
      metrics.putEvictionErrors++;                                     // :269
      return false;
    }
  }
  if (enoughSpace) {                                                   // :274-285
    if (mPageStore->Put(id, page, len) == kOk) {
      metrics.bytesWrittenCache += len;                                // :277
      return true;
    }
    UndoAddPage(id);
    metrics.putStoreWriteErrors++;                                     // :282
    return false;
  }
  // two-phase commit :288-346
  if (mMetaStore->HasPage(id)) return true;                            // :296-300
  if (!mMetaStore->RemovePage(victim.pageId)) {                        // :301-309
    UndoAddPage(id);
    metrics.putBenignRacingErrors++;                                   // :307
    return false;
  }
  enoughSpace = mMetaStore->Bytes() + len <= mCacheSize;               // :310
  if (enoughSpace) mMetaStore->AddPage(id, PageInfo{id, len});        // :311-313
  StoreStatus del = mPageStore->Delete(victim.pageId);                 // :318-320 ALWAYS
  if (del != kOk) {
    if (enoughSpace) UndoAddPage(id);                                  // :323-326
    metrics.putStoreDeleteErrors++;                                    // :328
    return false;
  }
  metrics.bytesEvicted += victim.pageSize;                             // :320
  metrics.pagesEvicted++;                                              // :321
  evicted.push_back(victim.pageId);
  if (!enoughSpace) {                                                  // :331-334 one victim per put
    metrics.putEvictionErrors++;                                       // :332
    return false;
  }
  if (mPageStore->Put(id, page, len) == kOk) {                         // :336-338
    metrics.bytesWrittenCache += len;                                  // :337
    return true;
  }
  UndoAddPage(id);                                                     // :339-344
  metrics.putStoreWriteErrors++;                                       // :343
  return false;
}

void LocalCacheManager::UndoAddPage(const PageId& id) {  // :349-357
  if (!mMetaStore->RemovePage(id)) metrics.cleanupPutErrors++;         // :354 (removePage threw)
}

int32_t LocalCacheManager::Get(const PageId& id, int32_t pageOffset, int32_t bytesToRead, uint8_t* buffer,
                               int32_t bufLen, int32_t offsetInBuffer) {  // :360-399
  if (!(pageOffset <= mPageSize)) return kGetIllegalArgument;             // :362-363
  if (!(bytesToRead <= bufLen - offsetInBuffer)) return kGetIllegalArgument;  // :364-366
  if (mState == NOT_IN_USE) {                                             // :368-372
    metrics.getNotReadyErrors++;
    metrics.getErrors++;
    return -1;
  }
  if (!mMetaStore->HasPage(id)) return 0;                                 // :376-382
  // getPage :540-555
  int32_t ret = 0;
  StoreStatus st = mPageStore->Get(id, pageOffset, bytesToRead, buffer, bufLen, offsetInBuffer, &ret);
  int32_t bytesRead;
  if (st == kIllegalArgument) return kGetIllegalArgument;  // IAE is not caught by getPage's catch (:550)
  if (st != kOk) bytesRead = -1;
  else bytesRead = (ret != bytesToRead) ? -1 : bytesToRead;                // :544-549
  if (bytesRead <= 0) {                                                   // :384-395
    metrics.getErrors++;                                                  // :385
    metrics.getStoreReadErrors++;                                         // :386
    if (!mMetaStore->RemovePage(id)) metrics.cleanupGetErrors++;          // :389-393
    return -1;
  }
  return bytesRead;
}

bool LocalCacheManager::Delete(const PageId& id) {  // :402-429
  if (mState != READ_WRITE) {
    metrics.deleteNotReadyErrors++;                                       // :405
    metrics.deleteErrors++;                                               // :406
    return false;
  }
  if (!mMetaStore->RemovePage(id)) {                                      // :411-419
    metrics.deleteNonExistingPageErrors++;
    metrics.deleteErrors++;
    return false;
  }
  const bool ok = mPageStore->Delete(id) == kOk;  // deletePage :530-538
  if (!ok) {                                                              // :422-427
    metrics.deleteStoreDeleteErrors++;
    metrics.deleteErrors++;
  }
  return ok;
}

void LocalCacheManager::RestoreOrInit() {  // :441-464
  if (!Restore()) {
    mMetaStore->Reset();
    if (mPageStore->Recreate() != kOk) {
      mState = NOT_IN_USE;
      return;
    }
  }
  mState = READ_WRITE;
}

bool LocalCacheManager::Restore() {  // :466-509
  if (!mPageStore->RootExists()) return false;
  std::vector<std::pair<bool, PageInfo>> pages;
  if (mPageStore->GetPages(&pages) != kOk) return false;
  for (auto& e : pages) {
    if (!e.first) return false;  // null PageInfo :477-480
    const PageInfo& info = e.second;
    bool enoughSpace = mMetaStore->Bytes() + info.pageSize <= mPageStore->GetCacheSize();  // :486-487
    if (enoughSpace) mMetaStore->AddPage(info.pageId, info);
    else if (mPageStore->Delete(info.pageId) != kOk) return false;  // :493-497 (exception -> false)
  }
  return true;
}

// ================================================================================================
// CACHE/LocalCacheFileInStream.java
// ================================================================================================

LocalCacheFileInStream::LocalCacheFileInStream(const std::string& fileId, int64_t fileLength, int64_t pageSize,
                                               LocalCacheManager* manager, ExternalRead external)  // :83-93
    : mFileId(fileId), mLength(fileLength), mPageSize(pageSize), mCacheManager(manager), mExternal(std::move(external)) {}

// :291-310.  getExternalFileInStream (:256-270) opens the external stream once and seeks it to the start of
// the page when it is not there already; the callback takes the position instead.
bool LocalCacheFileInStream::ReadExternalPage(int64_t pos, std::vector<uint8_t>* page) {
  const int64_t pageStart = pos - (pos % mPageSize);                                     // :292
  const int32_t pageSize = (int32_t)std::min<int64_t>(mPageSize, mLength - pageStart);   // :294
  page->assign((size_t)pageSize, 0);                                                     // :295
  int32_t totalBytesRead = 0;
  while (totalBytesRead < pageSize) {                                                    // :297-303
    ++externalReads;
    const int32_t bytesRead = mExternal(pageStart + totalBytesRead, page->data() + totalBytesRead, pageSize - totalBytesRead);
    if (bytesRead <= 0) break;
    totalBytesRead += bytesRead;
  }
  bytesReadExternal += totalBytesRead;                                                   // :304
  return totalBytesRead == pageSize;                                                     // :305-308 IOException
}

// The loop of read (:120-144) and of positionedRead (:188-212): the same statements over mPosition and over
// a local currentPosition.
int32_t LocalCacheFileInStream::ReadLoop(int64_t* position, uint8_t* b, int32_t bLen, int32_t off, int32_t len,
                                         bool positioned) {
  int32_t totalBytesRead = 0;
  const int64_t lengthToRead = std::min<int64_t>(len, mLength - *position);              // :118 / :186
  std::vector<uint8_t> page;
  while (totalBytesRead < lengthToRead) {
    const int64_t currentPage = *position / mPageSize;
    const int32_t currentPageOffset = (int32_t)(*position % mPageSize);
    const int32_t bytesLeftInPage = (int32_t)std::min<int64_t>(mPageSize - currentPageOffset, lengthToRead - totalBytesRead);
    const PageId pageId{mFileId, currentPage};
    const int32_t bytesRead = mCacheManager->Get(pageId, currentPageOffset, bytesLeftInPage, b, bLen, off + totalBytesRead);
    if (bytesRead == kGetIllegalArgument) return kGetIllegalArgument;  // thrown out of get(), out of the loop
    if (bytesRead > 0) {                                                                 // :129-132
      totalBytesRead += bytesRead;
      *position += bytesRead;
      bytesReadCache += bytesRead;
    } else {                                                                             // :133-143
      if (!ReadExternalPage(*position, &page)) return kStreamIo;
      if (!page.empty()) {
        if (off + totalBytesRead + bytesLeftInPage > bLen) return kGetIllegalArgument;  // System.arraycopy bounds
        memcpy(b + off + totalBytesRead, page.data() + currentPageOffset, (size_t)bytesLeftInPage);
        totalBytesRead += bytesLeftInPage;
        *position += bytesLeftInPage;
        bytesRequestedExternal += bytesLeftInPage;
        mCacheManager->Put(pageId, page.data(), (int32_t)page.size());                   // result ignored (:141)
      }
    }
  }
  // :146-150 / :214-221
  const int64_t left = positioned ? mLength - *position : Remaining();
  if (totalBytesRead > len || (totalBytesRead < len && left > 0)) return kStreamIo;
  return totalBytesRead;
}

int32_t LocalCacheFileInStream::Read(uint8_t* b, int32_t bLen, int32_t off, int32_t len) {  // :108-152
  if (!(len >= 0)) return kGetIllegalArgument;                                           // :109
  if (!(off >= 0)) return kGetIllegalArgument;                                           // :110
  if (len == 0) return 0;                                                                // :111-113
  if (mPosition >= mLength) return -1;                                                   // :114-116
  return ReadLoop(&mPosition, b, bLen, off, len, false);
}

int32_t LocalCacheFileInStream::PositionedRead(int64_t pos, uint8_t* b, int32_t bLen, int32_t off, int32_t len) {  // :175-223
  if (!(len >= 0)) return kGetIllegalArgument;                                           // :176
  if (!(off >= 0)) return kGetIllegalArgument;                                           // :177
  if (!(pos >= 0)) return kGetIllegalArgument;                                           // :178
  if (len == 0) return 0;                                                                // :179-181
  if (pos < 0 || pos >= mLength) return -1;                                              // :182-184
  int64_t currentPosition = pos;                                                         // :186
  return ReadLoop(&currentPosition, b, bLen, off, len, true);
}

int64_t LocalCacheFileInStream::Skip(int64_t n) {  // :154-162
  if (mClosed) return kStreamClosed;
  if (n <= 0) return 0;
  const int64_t toSkip = std::min(Remaining(), n);
  mPosition += toSkip;
  return toSkip;
}

int LocalCacheFileInStream::Seek(int64_t pos) {  // :230-243
  if (mClosed) return kStreamClosed;
  if (!(pos >= 0)) return kGetIllegalArgument;
  if (!(pos <= mLength)) return kGetIllegalArgument;
  if (pos == mPosition) return 0;
  if (pos < mPosition) mEOF = false;
  mPosition = pos;
  return 0;
}

}  // namespace orc
