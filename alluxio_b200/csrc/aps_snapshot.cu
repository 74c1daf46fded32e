// Snapshot of a store to a local file and back.
//
// LocalPageStore is durable for free: its pages ARE files, and after a restart LocalCacheManager.restore
// walks them with getPages() (CACHE/LocalCacheManager.java:466-509, CACHE/store/LocalPageStore.java:193-199).
// HBM is not, so a GPU store that wants the same warm restart writes its resident pages to local NVMe before
// it closes and reads them back after it opens; the manager's restore then finds them through getPages()
// exactly as it finds LocalPageStore's files.
//
// Built on the public entry points only (aps_iterate, aps_get_batch, aps_put_batch): pages travel through one
// pinned staging buffer, a batch per kernel launch.  File layout, little endian:
//   Header | payload of page 0 | payload of page 1 | ... | Record[n] | name table (Header.names_bytes bytes)
// every payload padded to 16 bytes; the records come last so that the writer needs no seek and a truncated
// file is detected by its size.
//
// Name table.  A file id that is neither a decimal long nor 32 hex digits reaches the device as (2^64-1, n), n being
// its index in an intern table that lives in the calling process (FileKeyCodec in alluxio_b200/host, page_store.py and
// GpuPageStore.java).  Those indices mean nothing after a restart, so the caller hands its table to
// aps_snapshot_save_names and reads it back with aps_snapshot_names BEFORE it lists the loaded pages.  Pages whose
// interned index has no name in the table handed in (always the case for plain aps_snapshot_save) are left out of the
// file: restoring them under a wrong id would be worse than a miss.  Layout: u32 n_names, then per name u32 length +
// bytes (no terminator).
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <vector>

#include "aps_internal.h"

namespace {

constexpr uint64_t kSnapMagic = 0x3150414E53535041ull;  // "APSSNAP1"
constexpr uint64_t kStagingBytes = 64ull << 20;

struct Header {
  uint64_t magic;
  uint32_t version, page_size;
  uint64_t n_pages, payload_bytes;
  uint32_t has_crc, names_bytes;  // names_bytes: size of the name table that follows the records (0 = none)
};
struct Record {
  aps_key key;
  uint32_t len, crc;
};
static_assert(sizeof(Header) == 40 && sizeof(Record) == 32, "snapshot layout");

inline uint64_t pad16(uint64_t n) { return (n + 15u) & ~15ull; }

struct Pinned {
  void* p = nullptr;
  ~Pinned() { if (p) aps_host_free_pinned(p); }
};
struct File {
  FILE* f = nullptr;
  ~File() { if (f) fclose(f); }
};

int collect(void* user, const aps_key* key, uint32_t len) {
  static_cast<std::vector<Record>*>(user)->push_back(Record{*key, len, 0});
  return 0;
}

}  // namespace

extern "C" {

int aps_snapshot_save(aps_store* s, const char* path, uint64_t* pages_out) {
  return aps_snapshot_save_names(s, path, 0, nullptr, pages_out);
}

int aps_snapshot_save_names(aps_store* s, const char* path, uint32_t n_names, const char* const* names, uint64_t* pages_out) {
  if (!path || (n_names && !names)) return APS_EINVAL;
  aps_config cfg;
  int rc = aps::store_config(s, &cfg);
  if (rc) return rc;
  std::vector<Record> pages;
  rc = aps_iterate(s, collect, &pages);
  if (rc) return rc;
  // interned ids without a name in the caller's table cannot be restored under the right id: leave them out
  pages.erase(std::remove_if(pages.begin(), pages.end(),
                             [&](const Record& r) { return r.key.file_hi == ~0ull && r.key.file_lo >= n_names; }),
              pages.end());
  const uint64_t slot = pad16(cfg.page_size);
  const uint32_t per_batch = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(cfg.max_batch, kStagingBytes / slot));
  Pinned stage;
  rc = aps_host_alloc_pinned((uint64_t)per_batch * slot, &stage.p);
  if (rc) return rc;
  const std::string tmp = std::string(path) + ".tmp";
  File out;
  out.f = fopen(tmp.c_str(), "wb");
  if (!out.f) return APS_EIO;
  Header h{kSnapMagic, 1, cfg.page_size, 0, 0, (cfg.flags & APS_FLAG_CRC) ? 1u : 0u, 0};
  if (fwrite(&h, sizeof(h), 1, out.f) != 1) return APS_EIO;
  std::vector<Record> kept;
  kept.reserve(pages.size());
  std::vector<aps_get_req> reqs(per_batch);
  std::vector<aps_result> res(per_batch);
  for (size_t lo = 0; lo < pages.size(); lo += per_batch) {
    const uint32_t n = (uint32_t)std::min<size_t>(per_batch, pages.size() - lo);
    for (uint32_t i = 0; i < n; ++i)
      reqs[i] = aps_get_req{pages[lo + i].key, 0, cfg.page_size, (uint8_t*)stage.p + (uint64_t)i * slot};
    rc = aps_get_batch(s, n, reqs.data(), res.data(), nullptr);
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i) {
      if (res[i].len == -APS_ENOTFOUND) continue;  // deleted or evicted since the walk: not in the snapshot
      if (res[i].len < 0) return -res[i].len;      // e.g. APS_ECORRUPT: do not write a bad page out as good
      const uint64_t len = (uint64_t)res[i].len, padded = pad16(len);
      uint8_t* p = (uint8_t*)stage.p + (uint64_t)i * slot;
      memset(p + len, 0, padded - len);
      if (padded && fwrite(p, 1, padded, out.f) != padded) return APS_EIO;
      kept.push_back(Record{pages[lo + i].key, (uint32_t)len, res[i].crc});
      h.payload_bytes += padded;
    }
  }
  h.n_pages = kept.size();
  if (!kept.empty() && fwrite(kept.data(), sizeof(Record), kept.size(), out.f) != kept.size()) return APS_EIO;
  if (n_names) {
    uint64_t bytes = sizeof(uint32_t);
    if (fwrite(&n_names, sizeof(n_names), 1, out.f) != 1) return APS_EIO;
    for (uint32_t i = 0; i < n_names; ++i) {
      const uint32_t len = names[i] ? (uint32_t)strlen(names[i]) : 0u;
      if (fwrite(&len, sizeof(len), 1, out.f) != 1 || (len && fwrite(names[i], 1, len, out.f) != len)) return APS_EIO;
      bytes += sizeof(len) + len;
    }
    if (bytes > 0xFFFFFFFFull) return APS_EINVAL;
    h.names_bytes = (uint32_t)bytes;
  }
  if (fseek(out.f, 0, SEEK_SET) != 0 || fwrite(&h, sizeof(h), 1, out.f) != 1 || fflush(out.f) != 0) return APS_EIO;
  fclose(out.f);
  out.f = nullptr;
  if (rename(tmp.c_str(), path) != 0) return APS_EIO;
  if (pages_out) *pages_out = kept.size();
  return APS_OK;
}

int aps_snapshot_load(aps_store* s, const char* path, uint64_t* pages_out, uint64_t* rejected_out) {
  if (!path) return APS_EINVAL;
  if (pages_out) *pages_out = 0;
  if (rejected_out) *rejected_out = 0;
  aps_config cfg;
  int rc = aps::store_config(s, &cfg);
  if (rc) return rc;
  File in;
  in.f = fopen(path, "rb");
  if (!in.f) return APS_EIO;
  Header h;
  if (fread(&h, sizeof(h), 1, in.f) != 1 || h.magic != kSnapMagic || h.version != 1) return APS_ECORRUPT;
  if (h.page_size != cfg.page_size) return APS_EINVAL;  // as a LocalPageStore directory of another page size is not reused
  if (fseek(in.f, 0, SEEK_END) != 0) return APS_EIO;
  const uint64_t size = (uint64_t)ftell(in.f);
  if (h.n_pages > (size / sizeof(Record)) || size != sizeof(Header) + h.payload_bytes + h.n_pages * sizeof(Record) + h.names_bytes)
    return APS_ECORRUPT;  // truncated or overwritten
  std::vector<Record> recs(h.n_pages);
  if (fseek(in.f, (long)(sizeof(Header) + h.payload_bytes), SEEK_SET) != 0) return APS_EIO;
  if (h.n_pages && fread(recs.data(), sizeof(Record), recs.size(), in.f) != recs.size()) return APS_EIO;
  uint64_t total = 0;
  for (const Record& r : recs) {
    if (r.len > cfg.page_size) return APS_ECORRUPT;
    total += pad16(r.len);
  }
  if (total != h.payload_bytes) return APS_ECORRUPT;
  if (fseek(in.f, (long)sizeof(Header), SEEK_SET) != 0) return APS_EIO;
  const uint64_t slot = pad16(cfg.page_size);
  const uint32_t per_batch = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(cfg.max_batch, kStagingBytes / slot));
  Pinned stage;
  rc = aps_host_alloc_pinned((uint64_t)per_batch * slot, &stage.p);
  if (rc) return rc;
  const bool check = h.has_crc && (cfg.flags & APS_FLAG_CRC);
  std::vector<aps_put_req> reqs(per_batch);
  std::vector<aps_result> res(per_batch);
  uint64_t loaded = 0, rejected = 0;
  for (size_t lo = 0; lo < recs.size(); lo += per_batch) {
    const uint32_t n = (uint32_t)std::min<size_t>(per_batch, recs.size() - lo);
    uint64_t at = 0;
    for (uint32_t i = 0; i < n; ++i) {
      const Record& r = recs[lo + i];
      uint8_t* p = (uint8_t*)stage.p + at;
      const uint64_t padded = pad16(r.len);
      if (padded && fread(p, 1, padded, in.f) != padded) return APS_EIO;
      reqs[i] = aps_put_req{r.key, r.len, 0, p};
      at += padded;
    }
    rc = aps_put_batch(s, n, reqs.data(), res.data(), nullptr);
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i) {
      if (res[i].len < 0) {  // typically APS_ENOSPC: the new store is smaller than the snapshot
        ++rejected;
      } else if (check && res[i].crc != recs[lo + i].crc) {  // the file went bad on disk: do not serve it
        aps_delete(s, &recs[lo + i].key);
        ++rejected;
      } else {
        ++loaded;
      }
    }
  }
  if (pages_out) *pages_out = loaded;
  if (rejected_out) *rejected_out = rejected;
  return APS_OK;
}

int aps_snapshot_names(const char* path, aps_name_cb cb, void* user) {
  if (!path || !cb) return APS_EINVAL;
  File in;
  in.f = fopen(path, "rb");
  if (!in.f) return APS_EIO;
  Header h;
  if (fread(&h, sizeof(h), 1, in.f) != 1 || h.magic != kSnapMagic || h.version != 1) return APS_ECORRUPT;
  if (h.names_bytes == 0) return APS_OK;
  if (fseek(in.f, 0, SEEK_END) != 0) return APS_EIO;
  const uint64_t size = (uint64_t)ftell(in.f);
  if (h.n_pages > (size / sizeof(Record)) || size != sizeof(Header) + h.payload_bytes + h.n_pages * sizeof(Record) + h.names_bytes)
    return APS_ECORRUPT;
  if (fseek(in.f, (long)(size - h.names_bytes), SEEK_SET) != 0) return APS_EIO;
  std::vector<char> blob(h.names_bytes);
  if (fread(blob.data(), 1, blob.size(), in.f) != blob.size()) return APS_EIO;
  uint32_t n = 0;
  size_t at = sizeof(n);
  if (blob.size() < at) return APS_ECORRUPT;
  memcpy(&n, blob.data(), sizeof(n));
  std::string name;
  for (uint32_t i = 0; i < n; ++i) {
    uint32_t len = 0;
    if (at + sizeof(len) > blob.size()) return APS_ECORRUPT;
    memcpy(&len, blob.data() + at, sizeof(len));
    at += sizeof(len);
    if (at + len > blob.size()) return APS_ECORRUPT;
    name.assign(blob.data() + at, len);
    at += len;
    if (cb(user, i, name.c_str())) break;
  }
  return APS_OK;
}

}  // extern "C"
