/*
 * aps_cache.h -- flat C view of the host-side cache layer (libaps_host.so).
 *
 * In an Alluxio deployment this layer IS the reference's Java code, unchanged: LocalCacheManager,
 * DefaultMetaStore and the evictors keep running in the JVM and drive the GPU store through the
 * PageStore interface.  Where there is no JVM (this repository's tests, bench, and
 * C/C++ embedders) the same logic is provided natively by alluxio_b200/host/: a C++ mirror with the
 * reference's names, argument meaning and results.  This header exposes it to C and ctypes.
 *
 * CACHE/ = /root/reference/core/client/fs/src/main/java/alluxio/client/file/cache/.
 */
#ifndef APS_CACHE_H_
#define APS_CACHE_H_

#include "aps.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct apc_manager apc_manager; /* LocalCacheManager over a GpuPageStore */
typedef struct apc_evictor apc_evictor; /* CacheEvictor */
typedef struct apc_meta apc_meta;       /* DefaultMetaStore */

/* CacheManager.State, CACHE/CacheManager.java:39-65 */
#define APC_NOT_IN_USE 0
#define APC_READ_ONLY 1
#define APC_READ_WRITE 2
/* get() result standing for the IllegalArgumentException of Preconditions.checkArgument,
 * CACHE/LocalCacheManager.java:362-366 */
#define APC_GET_ILLEGAL_ARGUMENT (-2)

/* LocalCacheManager.create(conf) with sync write + sync restore (CACHE/LocalCacheManager.java:102-138):
 * opens a GpuPageStore from cfg (cfg->cache_bytes = alluxio.user.client.cache.size, cfg->page_size =
 * ...page.size), restores (an HBM store starts empty) and goes READ_WRITE.
 * evictor: "LRU" (evictor/LRUCacheEvictor.java), "LFU" (evictor/LFUCacheEvictor.java) or "FIFO" (the
 * test double of CTEST/LocalCacheManagerTest.java:731-761). */
APS_API int apc_manager_create(const aps_config* cfg, const char* evictor, double lfu_logbase, apc_manager** out);
/* LocalCacheManager.create(conf, metaStore, pageStore) over a store that already holds pages: the restore
 * path (CACHE/LocalCacheManager.java:466-509) adopts them in getPages() order while they fit the cache size
 * and deletes the rest from the store.  If that restore fails (a delete that fails, a listing that fails) the store is
 * emptied and the cache goes on READ_WRITE and empty, as restoreOrInit does (:441-464); only if emptying fails too is the
 * manager NOT_IN_USE.  The store stays the caller's (apc_manager_close does not close it). */
APS_API int apc_manager_create_over(aps_store* store, uint32_t page_size, const char* evictor, double lfu_logbase,
                                    apc_manager** out);
/* Warm restart through the manager.  apc_manager_snapshot_save writes the store's pages together with
 * the manager's table of interned file ids (ids that are neither decimal longs nor 32 hex digits);
 * apc_manager_create_over_snapshot loads such a file into the freshly opened `store`, takes the table over and then runs
 * the reference's restore, so that getPages() yields the ids LocalPageStore's directory walk would
 * (CACHE/store/LocalPageStore.java:177-199, CACHE/LocalCacheManager.java:466-509). */
APS_API int apc_manager_snapshot_save(apc_manager* m, const char* path, uint64_t* pages_out);
APS_API int apc_manager_create_over_snapshot(aps_store* store, uint32_t page_size, const char* evictor, double lfu_logbase,
                                             const char* snapshot_path, apc_manager** out);
/* Everything LocalCacheManager.create reads from the configuration, in one place.  Exactly one of `over`,
 * n_shards > 0, or neither (a single new store on store.device) selects the page store. */
typedef struct apc_options {
  aps_config store;             /* GpuPageStoreOptions: page size, cache size, device, ... */
  const char* evictor;          /* alluxio.user.client.cache.evictor.class: "LRU" | "LFU" (| "FIFO", the test double) */
  double lfu_logbase;           /* alluxio.user.client.cache.evictor.lfu.logbase */
  uint32_t n_shards;            /* > 0: one cache over this many GPU stores (see apc_manager_create_sharded) */
  const int32_t* shard_devices; /* n_shards CUDA ordinals */
  aps_store* over;              /* non-NULL: adopt this store and restore from its pages (not closed by the manager) */
  int32_t async_write_threads;  /* alluxio.user.client.cache.async.write.enabled/.threads: 0 = synchronous puts.
                                 * Otherwise put() queues the page on a pool of this many threads that accepts a task
                                 * only while a thread is free (ThreadPoolExecutor over a SynchronousQueue,
                                 * CACHE/LocalCacheManager.java:156-161): a put of an id already queued, or one that
                                 * finds every thread busy, returns false (:221-243). */
  int32_t async_restore;        /* alluxio.user.client.cache.async.restore.enabled: create returns at once in state
                                 * READ_ONLY (gets served, puts refused, :208-212) and turns READ_WRITE when the
                                 * restore thread is done (:126-133) */
  int64_t timeout_ms;           /* alluxio.user.client.cache.timeout.duration: > 0 wraps the store in TimeBoundPageStore
                                 * (CACHE/TimeBoundPageStore.java:38-159; PageStore.create, CACHE/PageStore.java:72-75):
                                 * a put/get/delete that takes longer, or finds all timeout_threads busy, fails like an
                                 * IOException.  <= 0 (the reference's default -1): no wrapper. */
  int32_t timeout_threads;      /* alluxio.user.client.cache.timeout.threads (0 = the default 32) */
} apc_options;
APS_API int apc_manager_create_ex(const apc_options* options, apc_manager** out);
/* Waits for the restore thread and for every queued asynchronous put (tests; orderly shutdown). */
APS_API void apc_wait_async(apc_manager* m);
/* Client.CacheStorePutTimeout, ...GetTimeout, ...DeleteTimeout, ...ThreadsRejected; APS_EINVAL without a timeout. */
APS_API int apc_store_timeout_counters(apc_manager* m, int64_t* out4);
/* Test double of CTEST/HangingPageStore.java: the wrapped store's put / get / delete first sleep this long. */
APS_API int apc_set_store_delays(apc_manager* m, int32_t put_ms, int32_t get_ms, int32_t delete_ms);
APS_API int64_t apc_put_async_rejections(apc_manager* m); /* Client.CachePutAsyncRejectionErrors */
APS_API int64_t apc_put_not_ready_errors(apc_manager* m); /* Client.CachePutNotReadyErrors */

/* CacheManager.Factory.get / clear (CACHE/CacheManager.java:69-127): the process-wide manager, created by the
 * first caller.  A caller that arrives while another thread is still creating it receives *out = NULL (Java
 * null) and status APS_OK; a failed creation is counted (Client.CacheCreateErrors) and reported.  The manager
 * belongs to the factory: release it with apc_factory_clear, never apc_manager_close. */
APS_API int apc_factory_get(const apc_options* options, apc_manager** out);
APS_API void apc_factory_clear(void);
APS_API int64_t apc_factory_create_errors(void);

/* ONE cache over several GPUs: the manager, its meta store, its evictor and its byte budget
 * (cfg->cache_bytes) stay single and global, as in the reference (CACHE/LocalCacheManager.java:261); the pages
 * live in n_shards GPU stores, page -> shard aps_shard_of(key, n_shards) on device devices[shard].  put and
 * delete go to the owner; batched gets are one kernel that reads the owners in place.  cfg->slab_slots, when
 * non-zero, is per shard; by default every shard gets its share of the budget plus a quarter plus 64 slots,
 * because hash routing balances only on average (a put that finds its shard full fails like any store write
 * failure: the reference's undoAddPage path).  Victim order, hit/miss sequence and residency are those of the
 * single-store manager. */
APS_API int apc_manager_create_sharded(const aps_config* cfg, uint32_t n_shards, const int32_t* devices,
                                       const char* evictor, double lfu_logbase, apc_manager** out);
APS_API uint32_t apc_shard_count(apc_manager* m);
APS_API aps_store* apc_shard_store(apc_manager* m, uint32_t shard);
APS_API uint32_t apc_shard_of(apc_manager* m, const char* file_id, int64_t page_index);
APS_API int apc_manager_close(apc_manager* m);
/* CacheManager.put(PageId, byte[]) -> boolean, CACHE/LocalCacheManager.java:206-347.  put and delete never
 * throw in this mirror, so they are also their NoExceptionCacheManager forms (CACHE/NoExceptionCacheManager.java:37-46,81-90). */
APS_API int apc_put(apc_manager* m, const char* file_id, int64_t page_index, const void* page, int32_t len);
/* CacheManager.get(PageId, pageOffset, bytesToRead, buffer, offsetInBuffer) -> bytes read, 0 on a
 * miss, -1 on error, CACHE/LocalCacheManager.java:360-399 */
APS_API int32_t apc_get(apc_manager* m, const char* file_id, int64_t page_index, int32_t page_offset,
                        int32_t bytes_to_read, void* buffer, int32_t buffer_len, int32_t offset_in_buffer);
/* The same through NoExceptionCacheManager (CACHE/NoExceptionCacheManager.java:59-68): an illegal argument
 * becomes -1 and is counted as a get error. */
APS_API int32_t apc_get_noexcept(apc_manager* m, const char* file_id, int64_t page_index, int32_t page_offset,
                                 int32_t bytes_to_read, void* buffer, int32_t buffer_len, int32_t offset_in_buffer);
/* Client.Cache* counters kept by this layer, in this order: BytesReadCache, BytesRequestedExternal,
 * BytesReadExternal, BytesWrittenCache, BytesEvicted, PagesEvicted, GetErrors, PutErrors, DeleteErrors,
 * PutEvictionErrors, PutStoreWriteErrors, PutStoreDeleteErrors, GetStoreReadErrors, DeleteNonExistingPageErrors
 * (core/common/src/main/java/alluxio/metrics/MetricKey.java:778-995). */
APS_API void apc_metrics_read(apc_manager* m, int64_t* out14);
/* The rest of LocalCacheManager.Metrics (CACHE/LocalCacheManager.java:557-627), in this order: GetNotReadyErrors,
 * DeleteNotReadyErrors, DeleteStoreDeleteErrors, PutBenignRacingErrors, CleanupGetErrors, CleanupPutErrors,
 * PutNotReadyErrors, PutAsyncRejectionErrors. */
APS_API void apc_metrics_read_more(apc_manager* m, int64_t* out8);
/* CacheManager.delete(PageId) -> boolean, CACHE/LocalCacheManager.java:402-429 */
APS_API int apc_delete(apc_manager* m, const char* file_id, int64_t page_index);
APS_API int apc_state(apc_manager* m);
APS_API int64_t apc_bytes(apc_manager* m); /* MetaStore.bytes() */
APS_API int64_t apc_pages(apc_manager* m); /* MetaStore.pages() */
APS_API int32_t apc_page_lock_id(apc_manager* m, const char* file_id, int64_t page_index); /* :172-174 */
APS_API int apc_set_store_faults(apc_manager* m, uint32_t fault_mask);                     /* APS_FAULT_* */
APS_API aps_store* apc_store(apc_manager* m); /* the store, or shard 0 of a sharded manager */

/* LocalCacheFileInStream.positionedRead (CACHE/LocalCacheFileInStream.java:175-223): reads
 * [pos, pos+len) of a file of file_len bytes through the cache; pages that miss are fetched whole from
 * `external` (readExternalPage, :291-310) and put into the cache.  Returns bytes read or -1 at EOF /
 * on an external read failure.  Counters (optional): bytes served from the cache / from outside. */
typedef int (*apc_external_read)(void* user, int64_t file_pos, void* dst, int32_t len); /* returns bytes read */
APS_API int32_t apc_positioned_read(apc_manager* m, const char* file_id, int64_t file_len, int64_t pos, void* dst,
                                    int32_t len, apc_external_read external, void* user, int64_t* bytes_from_cache,
                                    int64_t* bytes_from_external);

/* LocalCacheFileSystem.openFile (CACHE/LocalCacheFileSystem.java:51-70) + LocalCacheFileInStream
 * (CACHE/LocalCacheFileInStream.java:43-330) as an object with a position.  m == NULL or a manager in state
 * NOT_IN_USE yields the delegated file system's own stream (every read goes to `external`, the cache is not
 * touched).  read / positioned_read return bytes read, -1 at end of file, APC_GET_ILLEGAL_ARGUMENT for negative
 * or out-of-buffer arguments, -3 where the reference throws IOException (external read came up short).
 * seek returns 0 or APC_GET_ILLEGAL_ARGUMENT (position negative or beyond the file); seek / skip on a closed
 * stream return -4 (IllegalStateException, :262-264). */
typedef struct apc_stream apc_stream;
APS_API apc_stream* apc_file_open(apc_manager* m, const char* file_id, int64_t file_len, apc_external_read external,
                                  void* user);
APS_API int32_t apc_stream_read(apc_stream* s, void* b, int32_t b_len, int32_t off, int32_t len);
APS_API int32_t apc_stream_positioned_read(apc_stream* s, int64_t pos, void* b, int32_t b_len, int32_t off, int32_t len);
APS_API int64_t apc_stream_skip(apc_stream* s, int64_t n);
APS_API int apc_stream_seek(apc_stream* s, int64_t pos);
APS_API int64_t apc_stream_remaining(apc_stream* s);
APS_API int64_t apc_stream_get_pos(apc_stream* s);
APS_API int apc_stream_bypass(apc_stream* s);
/* BytesReadCache, BytesRequestedExternal, BytesReadExternal of this stream (:312-330) */
APS_API void apc_stream_counters(apc_stream* s, int64_t* out3);
/* from now on every read of this stream makes ONE batched store call for all resident pages it
 * touches (and keeps a pinned staging buffer between reads) instead of one CacheManager.get per page. */
APS_API void apc_stream_set_batched(apc_stream* s, int on);
APS_API void apc_stream_close(apc_stream* s);
APS_API void apc_stream_destroy(apc_stream* s);

/* The same read with ONE batched store call for all resident pages it touches. */
APS_API int32_t apc_positioned_read_batched(apc_manager* m, const char* file_id, int64_t file_len, int64_t pos, void* dst,
                                            int32_t len, apc_external_read external, void* user,
                                            int64_t* bytes_from_cache, int64_t* bytes_from_external);

/* Trace replay, BASELINE config 3: for every access run the reference read flow -- get; on a miss the
 * page (page_len[k] bytes of the synthetic payload for (file, page), seed) is put.  file ids are the
 * decimal strings of file[k].  Outputs as in the oracle's replay so they can be compared element by
 * element: hit[k], put_ok[k] (255 = no put), victims in eviction order, final page/byte counts.
 * verify != 0 additionally checks every hit's bytes against the generator. Returns victims, or -1. */
APS_API int64_t apc_replay(apc_manager* m, int64_t n, const int64_t* file, const int64_t* page, const int32_t* page_len,
                           uint64_t seed, int verify, uint8_t* hit, uint8_t* put_ok, int64_t* victim_file,
                           int64_t* victim_page, int64_t victims_cap, int64_t* final_pages, int64_t* final_bytes,
                           int64_t* verify_failures);

/* The same replay with the store operations BATCHED: admission and eviction are decided on the host
 * access by access exactly as above (so hit/miss and victim sequences are identical), but the store
 * operations they imply are collected -- deletes of victims, puts of missed pages (payload generated on
 * the device), gets of hits -- and issued as aps_delete_batch / aps_put_batch / aps_get_batch in that order
 * every `window` accesses, or earlier when an operation would be reordered against a pending one on the
 * same id.  This is how the host-side evictor drives GPU hash-table inserts and erases at speed.
 * store_ops[0..2] receive the numbers of deletes, puts and gets issued, store_ops[3] the number of flushes,
 * store_ops[4] the gets whose length or status was not the expected one (must be 0).  The batches of one window
 * are issued by a worker thread while the caller's thread decides the next window; windows reach the store in
 * order.  seconds_host = time the policy itself took, seconds_store = time the worker spent in store calls. */
APS_API int64_t apc_replay_batched(apc_manager* m, int64_t n, const int64_t* file, const int64_t* page,
                                   const int32_t* page_len, uint64_t seed, uint32_t window, uint8_t* hit, uint8_t* put_ok,
                                   int64_t* victim_file, int64_t* victim_page, int64_t victims_cap, int64_t* final_pages,
                                   int64_t* final_bytes, int64_t* store_ops, double* seconds_host, double* seconds_store);

/* ---- policy objects on their own (no GPU needed) ------------------------------------------------ */
APS_API apc_evictor* apc_evictor_create(const char* kind, double lfu_logbase);
APS_API void apc_evictor_destroy(apc_evictor* e);
APS_API void apc_evictor_update_on_get(apc_evictor* e, const char* file_id, int64_t page_index);
APS_API void apc_evictor_update_on_put(apc_evictor* e, const char* file_id, int64_t page_index);
APS_API void apc_evictor_update_on_delete(apc_evictor* e, const char* file_id, int64_t page_index);
APS_API void apc_evictor_reset(apc_evictor* e);
/* 1 and the victim, or 0 for Java null */
APS_API int apc_evictor_evict(apc_evictor* e, char* file_out, size_t cap, int64_t* index_out);

APS_API apc_meta* apc_meta_create(const char* evictor, double lfu_logbase);
APS_API void apc_meta_destroy(apc_meta* m);
APS_API int apc_meta_has_page(apc_meta* m, const char* file_id, int64_t page_index);
APS_API void apc_meta_add_page(apc_meta* m, const char* file_id, int64_t page_index, int64_t size);
APS_API int apc_meta_get_page_info(apc_meta* m, const char* file_id, int64_t page_index, int64_t* size); /* 0 = not found */
APS_API int apc_meta_remove_page(apc_meta* m, const char* file_id, int64_t page_index);                   /* 0 = not found */
APS_API int64_t apc_meta_bytes(apc_meta* m);
APS_API int64_t apc_meta_pages(apc_meta* m);
APS_API void apc_meta_reset(apc_meta* m);
APS_API int apc_meta_evict(apc_meta* m, char* file_out, size_t cap, int64_t* index_out, int64_t* size_out);

/* String file id -> 128-bit key, and Java's String.hashCode for the lock striping. */
APS_API void apc_file_key(apc_manager* m, const char* file_id, uint64_t* hi, uint64_t* lo);
APS_API int32_t apc_java_string_hashcode(const char* utf8);
/* The file identifier the Hadoop-compatible client gives a file: lower-case hex MD5 of its path
 * string (core/client/hdfs/src/main/java/alluxio/hadoop/HadoopUtils.java:225-238).  out33 receives 32 characters
 * and a terminator.  Such ids map onto the 128-bit device key exactly: (hi, lo) are the digest's two halves. */
APS_API void apc_hadoop_file_identifier(const char* path_utf8, char* out33);
/* apc_file_key without a manager, for ids that need no interning (decimal longs, 32-hex). */
APS_API void apc_file_key_of(const char* file_id, uint64_t* hi, uint64_t* lo);

#ifdef __cplusplus
}
#endif
#endif /* APS_CACHE_H_ */
