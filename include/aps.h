/*
 * aps.h -- C ABI of the B200 page store ("Alluxio Page Store", libaps.so).
 *
 * This is the drop-in boundary for the client-side page-cache data path.  It is what a JNI shim
 * under a Java `GpuPageStore implements PageStore` binds (INTEGRATION.md shows that shim), and it
 * is what the JVM-free host mirror in alluxio_b200/host/ and the tests call.  Plain C: opaque
 * handle, pointers and sizes, integer status codes.  No C++ or torch types cross it.
 *
 * "CACHE/" = /root/reference/core/client/fs/src/main/java/alluxio/client/file/cache/.
 *
 * Status codes map to the reference's exceptions:
 *   APS_OK                         success
 *   APS_ENOTFOUND  -> alluxio.exception.PageNotFoundException
 *   APS_EINVAL     -> IllegalArgumentException (Preconditions.checkArgument in LocalPageStore.get)
 *   anything else  -> java.io.IOException(aps_strerror(code))
 *
 * Ownership: the caller owns every buffer.  Single-page calls do not touch caller memory after
 * they return.  Batched calls on a caller stream do not touch it after that stream has completed.
 *
 * Threading: every entry point may be called from any host thread.  Calls on one store are
 * serialised by a mutex inside the library while they enqueue work.  As in the reference
 * ("does not need to provide thread-safety", CACHE/PageStore.java:34-35), ordering between
 * operations on the SAME PageId is the caller's job (LocalCacheManager's striped page locks);
 * operations on different PageIds may be issued concurrently.
 */
#ifndef APS_H_
#define APS_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define APS_API __attribute__((visibility("default")))
#else
#define APS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define APS_OK 0
#define APS_ENOTFOUND 1  /* page not in the store */
#define APS_EINVAL 2     /* bad offset / length / argument */
#define APS_ENOSPC 3     /* no free slab slot or table entry */
#define APS_ECUDA 4      /* a CUDA call failed; aps_last_cuda_error() has the text */
#define APS_ECORRUPT 5   /* whole-page read whose CRC32C differs from the one recorded at put */
#define APS_ECLOSED 6    /* store already closed */
#define APS_ENOMEM 7     /* host or device allocation failed */
#define APS_EFAULT 8     /* injected fault (APS_FAULT_* below) */
#define APS_ELIMIT 9     /* request larger than the fixed scratch of aps_crc32c_noalloc */
#define APS_EIO 10       /* a snapshot file could not be read or written */
#define APS_ETIMEOUT 11  /* host layer: a store call outlived its time bound, or every store thread was busy */

/* PageId on the device: CACHE/PageId.java:23-74 is (String fileId, long pageIndex).  The file id
 * is carried as 128 bits; the host side maps strings to 128 bits without collisions (decimal,
 * 32-hex and UUID ids parse exactly, anything else is interned -- see alluxio_b200/host). */
typedef struct aps_key {
  uint64_t file_hi;
  uint64_t file_lo;
  uint64_t page_index;
} aps_key;

#define APS_FLAG_CRC 1u        /* compute CRC32C inside every copy, record it at put, verify whole-page gets */
/* (flag bit 2 is retired: it selected a bulk-async (TMA) form of the page copy, which measured slower than the 128-bit
 * load/store path in every configuration -- DESIGN.md 4.2.  The kernel stays in the library as an experiment, reachable
 * only through APS_COPY_VARIANT=tma; a store opened with the bit set ignores it.) */
#define APS_FLAG_STABLE_REQS 4u /* a promise by the caller of the device forms: a request array is complete in HBM before
                                 * the call that uses it is enqueued (it is not produced by earlier work on the same
                                 * stream).  Back-to-back gets on one stream always overlap: a get resolves its pages and
                                 * warms L2 with their first rows while the previous get is still finishing, waits for it,
                                 * then reads its requests AGAIN and starts over if they changed meanwhile.  With this
                                 * flag it trusts the first read, has its first loads in flight before it waits, and
                                 * skips the second read (0.92 instead of 0.91 of the HBM peak on 256 MiB batches;
                                 * without any overlap: 0.87). */
#define APS_FLAG_DEVICE_CLOCK 8u /* opt-in, SURVEY.md 8f-3: the store evicts by itself.  Every get and put sets a
                                 * reference bit on the page's slot; a put batch that finds fewer free slots than it has
                                 * new pages first runs a CLOCK sweep on the device (clear set bits, evict pages whose
                                 * bit is clear) for the difference.  Capacity is then counted in slab slots, not bytes,
                                 * and the victim order is CLOCK's, NOT the reference's LRU/LFU (CACHE/evictor/): never
                                 * combine it with a host-side LocalCacheManager, whose meta store would go stale.
                                 * For caches whose accesses stay on the GPU. */

/* Replaces LocalPageStoreOptions / PageStoreOptions (CACHE/store/PageStoreOptions.java:23-182). */
typedef struct aps_config {
  int32_t device;           /* CUDA ordinal */
  uint32_t page_size;       /* alluxio.user.client.cache.page.size: any byte count up to 1 GiB (slab slots are spaced at the next multiple of 16) */
  uint64_t cache_bytes;     /* alluxio.user.client.cache.size; what getCacheSize() reports */
  uint64_t slab_slots;      /* slots in the HBM slab; 0 = ceil(cache_bytes / page_size) + 1/8 slack (min 64) */
  uint64_t table_capacity;  /* hash-table entries; 0 = next power of two >= 2 * slab_slots (min 1024) */
  uint32_t max_batch;       /* largest n accepted by one batched call; 0 = 4096 */
  uint32_t flags;           /* APS_FLAG_* */
} aps_config;

typedef struct aps_store aps_store; /* opaque; one HBM slab + table on one GPU */
typedef void* aps_stream;           /* a cudaStream_t, or NULL for the store's own stream */

/* PageStore.open / PageStore.create (CACHE/PageStore.java:49-77).  HBM is volatile, so open and
 * create are the same thing: the store starts empty (getPages() yields nothing, which
 * LocalCacheManager.restore accepts, CACHE/LocalCacheManager.java:466-509). */
APS_API int aps_open(const aps_config* cfg, aps_store** out);
/* AutoCloseable.close (CACHE/PageStore.java:37).  Frees every device and pinned-host resource of the store.  Idempotent:
 * the handle itself stays valid as a small tombstone for the life of the process, so a second close returns APS_OK and
 * any other call on a closed handle returns APS_ECLOSED. */
APS_API int aps_close(aps_store* s);

/* ---- single-page operations: exactly the PageStore contract, host pointers ------------------ */

/* PageStore.put(PageId, byte[]) CACHE/PageStore.java:119 / LocalPageStore.java:68-86.
 * Stores exactly len bytes (0 <= len <= page_size); replaces the page if the id is present.
 * On failure no partial page is left behind. */
APS_API int aps_put(aps_store* s, const aps_key* key, const void* src, uint32_t len);

/* PageStore.get(PageId, pageOffset, bytesToRead, buffer, bufferOffset) CACHE/PageStore.java:147 /
 * LocalPageStore.java:88-121.  dst = buffer + bufferOffset and dst_cap = buffer.length -
 * bufferOffset (the shim checks bufferOffset <= buffer.length itself).  Copies
 * min(pageLen - page_off, dst_cap, bytes_to_read) bytes and reports that count in *out_len.
 * page_off > pageLen -> APS_EINVAL; unknown id -> APS_ENOTFOUND.  *out_crc (optional) receives the
 * CRC32C of the bytes returned when the store was opened with APS_FLAG_CRC. */
APS_API int aps_get(aps_store* s, const aps_key* key, uint32_t page_off, uint32_t bytes_to_read, void* dst, uint32_t dst_cap,
            uint32_t* out_len, uint32_t* out_crc);

/* PageStore.delete(PageId) CACHE/PageStore.java:157 / LocalPageStore.java:123-137. */
APS_API int aps_delete(aps_store* s, const aps_key* key);

/* PageStore.getCacheSize() CACHE/PageStore.java:171. */
APS_API uint64_t aps_cache_size(const aps_store* s);

/* PageStore.getPages() CACHE/PageStore.java:166 / LocalPageStore.java:193-199: calls cb once per
 * resident page with its id and stored length.  A non-zero return from cb stops the walk. */
typedef int (*aps_iter_cb)(void* user, const aps_key* key, uint32_t len);
APS_API int aps_iterate(aps_store* s, aps_iter_cb cb, void* user);

/* ---- batched operations: the benchmarked path (additive; the reference API is one page per
 *      call, CACHE/CacheManager.java:148-182) ------------------------------------- */

/* One request of a batched get.  dst may be device memory or pinned (page-locked) host memory. */
typedef struct aps_get_req {
  aps_key key;
  uint32_t page_off;
  uint32_t nbytes; /* min(bytes_to_read, room in dst) */
  void* dst;
} aps_get_req;

/* len >= 0: bytes copied.  len < 0: -(status code) for this request (-APS_ENOTFOUND, ...). */
typedef struct aps_result {
  int32_t len;
  uint32_t crc;
} aps_result;

typedef struct aps_put_req {
  aps_key key;
  uint32_t len;
  uint32_t reserved;
  const void* src; /* device memory or pinned host memory; host-array calls also take pageable host memory */
} aps_put_req;

/* Batched LocalPageStore.get.  reqs and res are HOST arrays; the call stages them, runs the
 * lookup + gather (+CRC) kernels on `stream`, copies the results back and waits for them.
 * Keys within one batch need not be distinct for gets.  n may exceed max_batch (served in pieces).
 * Destinations (all of one kind per call): device memory (written by the kernel), page-locked host memory
 * (batches of 8 MiB and more are gathered in an HBM staging area in chunks while the copy engine moves the
 * previous chunk; smaller ones are written by the kernel over PCIe), or plain pageable host memory -- a Java
 * ByteBuffer.allocateDirect buffer, a malloc'd array -- which always goes through the staging area and
 * cudaMemcpy.  The staging area (one slot per request of the largest batch, at most 1 GiB) is allocated on
 * first use and kept until close. */
APS_API int aps_get_batch(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream);

/* Batched LocalPageStore.put.  Keys within one batch must be distinct (else APS_EINVAL).  Sources (all of one
 * kind per call): device memory or page-locked host memory (read by the kernel in place), or plain pageable host
 * memory (copied into the staging area first). */
APS_API int aps_put_batch(aps_store* s, uint32_t n, const aps_put_req* reqs, aps_result* res, aps_stream stream);

/* Batched LocalPageStore.delete; res[i].len is 0 or -APS_ENOTFOUND.  Keys must be distinct. */
APS_API int aps_delete_batch(aps_store* s, uint32_t n, const aps_key* keys, aps_result* res, aps_stream stream);

/* Device-resident forms: d_reqs / d_res / d_keys are DEVICE arrays that are already in HBM.
 * Fully asynchronous on `stream`: nothing is copied to or from the host and the call does not
 * wait.  This is what the kernel-only benchmark times.
 * All calls on one store share that store's scratch (job list, CRC accumulators, staging): two of
 * them must never RUN on the GPU at the same time.  Host forms wait before they return and are
 * serialised by the store's mutex; device forms are ordered by their stream, so put them all on one
 * stream (or order the streams with events), and do not mix them with host forms from another
 * thread.  Use one store per thread, or one stream per store, when in doubt. */
APS_API int aps_get_batch_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res, aps_stream stream);
APS_API int aps_put_batch_dev(aps_store* s, uint32_t n, const aps_put_req* d_reqs, aps_result* d_res, aps_stream stream);
APS_API int aps_delete_batch_dev(aps_store* s, uint32_t n, const aps_key* d_keys, aps_result* d_res, aps_stream stream);

/* Membership only: the device form of DefaultMetaStore.hasPage (CACHE/DefaultMetaStore.java:51-53)
 * for a batch of ids.  d_len[i] = stored length, or -1 when absent. */
APS_API int aps_lookup_batch_dev(aps_store* s, uint32_t n, const aps_key* d_keys, int32_t* d_len, aps_stream stream);

/* ---- sharded cache over the GPUs of one box (additive; the reference cache is process-local) ------
 * Pages are owned by GPU aps_shard_of(key, n): puts and deletes go to the owner's store.  A batched
 * get may ask for any page: the requesting GPU probes the owner's table and reads the owner's slab in
 * place over NVLink (P2P), with lookup, gather and CRC check still one kernel -- no staging buffer and no
 * collective.  One process per GPU exchanges aps_shard_handle blobs (CUDA IPC) by any means it likes
 * (bench.py uses torch.distributed); one process that owns several stores passes them directly. */
typedef struct aps_shard_handle {
  unsigned char opaque[256];
} aps_shard_handle;
APS_API uint32_t aps_shard_of(const aps_key* key, uint32_t n_shards);
APS_API int aps_shard_export(aps_store* s, aps_shard_handle* out);
/* handles[i] = export of shard i (handles[my_shard] is ignored).  All shards must share page_size. */
APS_API int aps_shard_attach(aps_store* s, uint32_t n_shards, uint32_t my_shard, const aps_shard_handle* handles);
APS_API int aps_shard_attach_local(aps_store* s, uint32_t n_shards, uint32_t my_shard, aps_store* const* stores);
/* As aps_get_batch_dev, but every request is served from the shard that owns its key. */
APS_API int aps_get_batch_sharded_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res,
                                      aps_stream stream);
/* As aps_get_batch (host arrays, waits), served from the owning shards. */
APS_API int aps_get_batch_sharded(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream);

/* ---- push transport for the sharded multi-get (additive; no counterpart in the reference, whose cache is
 * process-local: CACHE/LocalCacheManager.java has no remote tier) --------------------------------------------------
 * Over NVLink with every GPU busy, stores into a peer run faster than loads from it (700 against 626 GB/s per GPU
 * measured, tools/micro/p2p_transports.cu), so here the OWNER moves a page: a multi-get posts its requests to the
 * owning GPUs, each owner's serve kernel looks the pages up in its own table and streams them from its own slab into
 * the requester's LANDING BUFFER with the CRC32C check fused in, and the requester's stream waits for the owners' done
 * words.  Destinations must therefore lie in memory the peers can write: the landing buffer allocated by
 * aps_push_setup (requests with a destination outside it are answered -APS_EINVAL).  The store must have APS_FLAG_CRC.
 *   set-up:  aps_push_setup on every store, exchange aps_push_handle blobs (CUDA IPC), aps_push_attach; or
 *            aps_push_attach_local for stores of one process.  depth = multi-gets a store may have in flight (1..8).
 *   use:     aps_get_batch_push_begin_dev posts batch k and starts this GPU's serve kernel for round k (on a stream of
 *            the store's own); aps_get_batch_push_end_dev makes `stream` wait until the oldest batch in flight has
 *            landed and writes its results.  aps_get_batch_push_dev = begin + end.  A caller that wants two batches in
 *            flight issues begin(k+1) before end(k), with different destination ranges.
 * COLLECTIVE: every store of the shard set must issue the same sequence of begin calls (an owner's round k waits for
 * round k's posting of every peer; a peer that never calls makes the others time out after 30 s with -APS_ETIMEOUT
 * results).  begin and end calls of one store must be ordered among themselves (one stream, or events).  Puts and
 * deletes on an owner are not ordered against the serve kernels: keep them apart, as for aps_get_batch_sharded_dev. */
typedef struct aps_push_handle {
  unsigned char opaque[192];
} aps_push_handle;
APS_API int aps_push_setup(aps_store* s, uint64_t landing_bytes, uint32_t depth);
APS_API void* aps_push_landing(aps_store* s, uint64_t* bytes_out);
APS_API int aps_push_export(aps_store* s, aps_push_handle* out);
APS_API int aps_push_attach(aps_store* s, uint32_t n_shards, uint32_t my_shard, const aps_push_handle* handles);
APS_API int aps_push_attach_local(aps_store* s, uint32_t n_shards, uint32_t my_shard, aps_store* const* stores);
APS_API int aps_get_batch_push_begin_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_stream stream);
APS_API int aps_get_batch_push_end_dev(aps_store* s, aps_result* d_res, aps_stream stream);
APS_API int aps_get_batch_push_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res, aps_stream stream);

/* ---- checksum path: `alluxio fs checksum` ------------------------------------------------------
 * ChecksumCommand.calculateChecksum (shell/src/main/java/alluxio/cli/fs/command/
 * ChecksumCommand.java:83-90) = DigestUtils.md5Hex over the file's bytes.  One digest per file,
 * files processed in parallel.  d_data[i] points at len[i] bytes in device memory (or pinned host
 * memory); d_data, d_len and d_out are DEVICE arrays.  Asynchronous on `stream`. */
APS_API int aps_md5_batch_dev(int32_t device, uint32_t n_files, const void* const* d_data, const uint64_t* d_len,
                      uint8_t (*d_out)[16], aps_stream stream);
/* Streaming form for file sets larger than HBM (4,096 x 64 MiB does not fit one GPU): the files are fed
 * in waves.  d_state[i] holds file i's four chaining words: initialise it to the MD5 IV (0x67452301,
 * 0xefcdab89, 0x98badcfe, 0x10325476) before the first wave.  With finalize == 0 a wave consumes d_len[i]
 * bytes (a multiple of 64) and updates d_state; with finalize != 0 it also pads for
 * d_prefix_len[i] + d_len[i] total bytes and writes the digests. */
APS_API int aps_md5_stream_dev(int32_t device, uint32_t n_files, const void* const* d_data, const uint64_t* d_len,
                               uint32_t (*d_state)[4], const uint64_t* d_prefix_len, int finalize, uint8_t (*d_out)[16],
                               aps_stream stream);
/* Host form: arrays on the host, file payloads wherever they are (device or pinned); waits. */
APS_API int aps_md5_batch(int32_t device, uint32_t n_files, const void* const* data, const uint64_t* len, uint8_t (*out)[16]);

/* CRC32C (RFC 3720, Castagnoli) of n device-resident byte ranges, without a copy.  Stand-alone
 * verify pass; the reference has no CRC at all. */
APS_API int aps_crc32c_batch_dev(aps_store* s, uint32_t n, const void* const* d_data, const uint32_t* d_len, uint32_t* d_crc,
                         aps_stream stream);

/* CRC32C of n device-resident byte ranges WITHOUT any allocation.  This entry point has no
 * counterpart in the reference; it exists for callers of this library that cannot ask the driver
 * for memory when they need a checksum: a scrubber re-checking pages while the slab already fills
 * HBM (cudaMalloc may fail then), read threads that must not be stalled by the device-wide
 * synchronisation cudaMalloc/cudaFree imply, or a host process whose device memory all belongs to
 * another component's pool.  It needs no aps_store.  All working memory beyond the caller's d_data,
 * d_len and d_crc is a fixed 128 KiB array inside the library (held by every GPU that loads the
 * library until the process exits), hence the limits: n <= APS_NOALLOC_MAX_RANGES and max_len <=
 * APS_NOALLOC_MAX_LEN, otherwise APS_ELIMIT is returned and nothing is launched.  max_len is the
 * caller's bound on every d_len[i]; longer ranges are cut to max_len.  d_data, d_len, d_crc are
 * DEVICE arrays.  Asynchronous on `stream` (NULL = the default stream).
 * The first call on a device sets that device's copy of the array up (a few synchronous copies): make
 * that call before any stream capture begins.  Every call reuses the same array, so calls that can
 * overlap on the GPU race with each other: issue them on one stream or order them yourself. */
#define APS_NOALLOC_MAX_RANGES 1024u
#define APS_NOALLOC_MAX_LEN (4u << 20)
APS_API int aps_crc32c_noalloc(int32_t device, uint32_t n, const void* const* d_data, const uint32_t* d_len,
                               uint32_t max_len, uint32_t* d_crc, aps_stream stream);

/* Explicit CLOCK sweep on a store opened with APS_FLAG_DEVICE_CLOCK: evicts n pages (fewer if the store holds
 * fewer).  d_victims (optional, DEVICE array of n keys) receives their ids, d_count (optional, DEVICE) how many.
 * No counterpart in the reference, whose eviction is decided by CacheEvictor on the host. */
APS_API int aps_clock_evict_dev(aps_store* s, uint32_t n, aps_key* d_victims, uint32_t* d_count, aps_stream stream);

/* ---- warm restart --------------------------------------------------------------
 * LocalPageStore's pages are files, so they survive a restart and LocalCacheManager.restore finds them
 * through getPages() (CACHE/LocalCacheManager.java:466-509; CACHE/store/LocalPageStore.java:193-199).
 * HBM does not survive; these two calls give a GPU store the same behaviour through a file on local NVMe.
 * aps_snapshot_save writes every resident page (id, length, CRC32C, bytes) to `path` (written as path.tmp,
 * then renamed); pages deleted while it runs are simply left out.  aps_snapshot_load puts the pages of a
 * snapshot into a freshly opened store of the same page size; a page that no longer fits (APS_ENOSPC) or
 * whose bytes no longer match the recorded CRC32C is counted in *rejected and not kept.  Afterwards
 * aps_iterate lists the loaded pages and the manager's restore adopts them.  Errors: APS_EIO (file),
 * APS_ECORRUPT (not a snapshot, truncated), APS_EINVAL (page size differs). */
APS_API int aps_snapshot_save(aps_store* s, const char* path, uint64_t* pages_out);
APS_API int aps_snapshot_load(aps_store* s, const char* path, uint64_t* pages_out, uint64_t* rejected_out);
/* File ids that are neither decimal longs nor 32 hex digits (UUIDs, raw paths) reach the store as (2^64-1, n), n
 * being an index into an intern table of the CALLING process; LocalPageStore keeps the id in the file path instead
 * (CACHE/store/LocalPageStore.java:139-147,177-191), so its getPages() returns the same ids after a restart.  To get
 * the same here the caller hands its table to aps_snapshot_save_names (names[i] = the id interned as n = i, UTF-8),
 * and after a restart reads it back with aps_snapshot_names into its empty table BEFORE it lists the loaded pages.
 * Pages whose interned index has no name (all of them with plain aps_snapshot_save) are left out of the file. */
typedef int (*aps_name_cb)(void* user, uint32_t index, const char* name); /* non-zero return stops the walk */
APS_API int aps_snapshot_save_names(aps_store* s, const char* path, uint32_t n_names, const char* const* names,
                                    uint64_t* pages_out);
APS_API int aps_snapshot_names(const char* path, aps_name_cb cb, void* user);

/* ---- synthetic data, stats, diagnostics ---------------------------------------------------------*/

/* Fills device memory with the benchmark payload: 64-bit word k of page (file, index) is the k-th
 * splitmix64 output for seed ^ (file << 32) ^ index.  Writes n_pages pages of
 * page_len bytes, page p at d_dst + p * stride, for ids files[p], indices[p] (DEVICE arrays). */
APS_API int aps_fill_pages_dev(int32_t device, void* d_dst, uint64_t stride, uint32_t page_len, uint32_t n_pages,
                       const uint64_t* d_files, const uint64_t* d_indices, uint64_t seed, aps_stream stream);

/* Host-array form of aps_fill_pages_dev (files / indices on the host, n_pages <= max_batch). */
APS_API int aps_fill_pages(aps_store* s, void* d_dst, uint64_t stride, uint32_t page_len, uint32_t n_pages,
                           const uint64_t* files, const uint64_t* indices, uint64_t seed, aps_stream stream);
/* Plain device buffers for embedders that have no CUDA runtime of their own (staging areas for batched
 * puts and gets).  Thin wrappers over cudaMalloc / cudaFree on `device`. */
APS_API int aps_device_alloc(int32_t device, uint64_t bytes, void** out);
/* Page-locked host memory (cudaHostAlloc / cudaFreeHost): kernels read and write it in place over PCIe. */
APS_API int aps_host_alloc_pinned(uint64_t bytes, void** out);
APS_API int aps_host_free_pinned(void* p);
APS_API int aps_device_free(int32_t device, void* p);

typedef struct aps_stats_t {
  uint64_t pages;           /* resident pages */
  uint64_t bytes;           /* sum of their stored lengths */
  uint64_t slots_free;      /* free slab slots */
  uint64_t table_capacity;  /* entries */
  uint64_t tombstones;      /* deleted entries still in probe chains */
  uint64_t lookups;         /* keys looked up since open */
  uint64_t probe_buckets;   /* 256-byte buckets read by those lookups */
  uint64_t bytes_copied;    /* payload bytes moved by get/put kernels */
  uint64_t crc_mismatches;  /* whole-page gets whose CRC differed from the one recorded at put */
  uint64_t kernel_launches; /* kernels this store has launched */
  uint64_t slow_sms;        /* SMs the open-time calibration classed as slow (they get one task less per launch) */
  uint64_t clock_evictions; /* pages evicted by the device-side CLOCK sweep (APS_FLAG_DEVICE_CLOCK) */
} aps_stats_t;
APS_API int aps_stats(aps_store* s, aps_stats_t* out);

/* Timing of the copy kernel (K2/K3) alone, for the roofline line of bench.py: between begin and end
 * every launch of it is bracketed by CUDA events on its own stream.  end() waits for them and
 * returns the summed device time and the number of launches timed. */
APS_API int aps_profile_begin(aps_store* s, uint32_t max_launches);
APS_API int aps_profile_end(aps_store* s, double* copy_ms_total, uint32_t* copy_launches);

/* Rebuilds the hash table without tombstones (they only arise when pages are deleted from buckets that were
 * completely full; aps_stats reports how many there are).  Waits; no other call may run
 * on the store meanwhile; not available once peers are attached (aps_shard_attach). */
APS_API int aps_compact(aps_store* s);

/* Fault injection for the reference's FaultyPageStore tests (CTEST/LocalCacheManagerTest.java:
 * 660-691): while set, aps_put / aps_delete fail with APS_EFAULT and change nothing. */
#define APS_FAULT_PUT 1u
#define APS_FAULT_DELETE 2u
APS_API int aps_set_faults(aps_store* s, uint32_t fault_mask);

/* Profiling hook: while enabled, every CTA of the copy kernel records four globaltimer stamps (ns):
 * start, CRC tables ready, first warp finished, end; every warp also records when it finished.  out
 * receives max_words values of the last launch: words [4*cta .. 4*cta+3] are the CTA stamps and
 * words [1024 + 32*cta + warp] the warp end stamps (all-ones = not written).  Costs a memset per
 * launch; off by default. */
APS_API int aps_debug_timeline(aps_store* s, int enable, uint64_t* out, uint32_t max_words);

/* Test hook: flips one byte of a resident page in the slab so the CRC check can be shown to fire. */
APS_API int aps_debug_corrupt(aps_store* s, const aps_key* key, uint32_t byte_offset);

/* Test hook, host arithmetic only (needs no GPU): the GF(2) constants of the fused CRC32C copy --
 * fold[4][256] (x^(8*512) per byte lane), z4[4][256] (x^32), xpow8[32] (x^(8*2^k)) and
 * lane_shift[32].  tests/test_crc_algebra.py replays the kernel's arithmetic with them on the CPU. */
APS_API int aps_debug_crc_constants(uint32_t* fold, uint32_t* z4, uint32_t* xpow8, uint32_t* lane_shift);
/* The same for fold4[4][256] (x^(8*512*4)): the tables of the kernels that fold the four rows of a group independently. */
APS_API int aps_debug_crc_fold4(uint32_t* fold4);

APS_API const char* aps_strerror(int code);
APS_API const char* aps_last_cuda_error(void); /* thread-local text of the last failed CUDA call */
APS_API const char* aps_version(void);
/* Total kernels launched by this library in this process (all stores + stateless entry points). */
APS_API uint64_t aps_total_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* APS_H_ */
