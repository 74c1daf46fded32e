/*
 * Argument marshalling of the batched JNI calls, kept free of jni.h so that it can be compiled and run without a JVM
 * (integration/c/batch_example.c, tests/test_c_abi.py): flat primitive arrays, as a JNI shim gets them from
 * Get<Type>ArrayElements, plus the address of ONE direct ByteBuffer -> one aps_*_batch call.
 *   keys      3 longs per request: file_hi, file_lo, page_index (GpuPageStore.fileKey)
 *   base      GetDirectBufferAddress of the buffer the pages are read into / written from; pinned (nativeAllocPinned)
 *             or plain allocateDirect memory -- libaps stages the latter itself
 *   off       byte offset of request i inside that buffer
 * Returns the aps status of the call; per-request outcomes go to out_len (>= 0 bytes, < 0 = -(aps status)) and out_crc.
 */
#ifndef APS_MARSHAL_H_
#define APS_MARSHAL_H_
#include <stdint.h>
#include <stdlib.h>

#include "aps.h"

/* sharded != 0: the store is shard 0 of a set attached with aps_shard_attach_local; every request is served from the store
 * that owns its key (aps_get_batch_sharded). */
static inline int aps_marshal_get_batch_ex(aps_store* s, int sharded, int32_t n, const int64_t* keys, const int32_t* page_off,
                                    const int32_t* nbytes, void* base, const int64_t* off, int32_t* out_len, int32_t* out_crc) {
  aps_get_req* rq;
  aps_result* rs;
  int32_t i;
  int rc;
  if (n <= 0) return APS_OK;
  rq = (aps_get_req*)malloc((size_t)n * sizeof(*rq));
  rs = (aps_result*)malloc((size_t)n * sizeof(*rs));
  if (!rq || !rs) {
    free(rq);
    free(rs);
    return APS_ENOMEM;
  }
  for (i = 0; i < n; ++i) {
    rq[i].key.file_hi = (uint64_t)keys[3 * i];
    rq[i].key.file_lo = (uint64_t)keys[3 * i + 1];
    rq[i].key.page_index = (uint64_t)keys[3 * i + 2];
    rq[i].page_off = (uint32_t)page_off[i];
    rq[i].nbytes = (uint32_t)nbytes[i];
    rq[i].dst = (unsigned char*)base + off[i];
  }
  rc = sharded ? aps_get_batch_sharded(s, (uint32_t)n, rq, rs, NULL) : aps_get_batch(s, (uint32_t)n, rq, rs, NULL);
  for (i = 0; i < n && rc == APS_OK; ++i) {
    out_len[i] = rs[i].len;
    if (out_crc) out_crc[i] = (int32_t)rs[i].crc;
  }
  free(rq);
  free(rs);
  return rc;
}

static inline int aps_marshal_get_batch(aps_store* s, int32_t n, const int64_t* keys, const int32_t* page_off, const int32_t* nbytes,
                                 void* base, const int64_t* off, int32_t* out_len, int32_t* out_crc) {
  return aps_marshal_get_batch_ex(s, 0, n, keys, page_off, nbytes, base, off, out_len, out_crc);
}

static inline int aps_marshal_put_batch(aps_store* s, int32_t n, const int64_t* keys, const void* base, const int64_t* off,
                                 const int32_t* len, int32_t* out_status, int32_t* out_crc) {
  aps_put_req* rq;
  aps_result* rs;
  int32_t i;
  int rc;
  if (n <= 0) return APS_OK;
  rq = (aps_put_req*)malloc((size_t)n * sizeof(*rq));
  rs = (aps_result*)malloc((size_t)n * sizeof(*rs));
  if (!rq || !rs) {
    free(rq);
    free(rs);
    return APS_ENOMEM;
  }
  for (i = 0; i < n; ++i) {
    rq[i].key.file_hi = (uint64_t)keys[3 * i];
    rq[i].key.file_lo = (uint64_t)keys[3 * i + 1];
    rq[i].key.page_index = (uint64_t)keys[3 * i + 2];
    rq[i].len = (uint32_t)len[i];
    rq[i].reserved = 0;
    rq[i].src = (const unsigned char*)base + off[i];
  }
  rc = aps_put_batch(s, (uint32_t)n, rq, rs, NULL);
  for (i = 0; i < n && rc == APS_OK; ++i) {
    out_status[i] = rs[i].len < 0 ? -rs[i].len : APS_OK;
    if (out_crc) out_crc[i] = (int32_t)rs[i].crc;
  }
  free(rq);
  free(rs);
  return rc;
}

static inline int aps_marshal_delete_batch(aps_store* s, int32_t n, const int64_t* keys, int32_t* out_status) {
  aps_key* k;
  aps_result* rs;
  int32_t i;
  int rc;
  if (n <= 0) return APS_OK;
  k = (aps_key*)malloc((size_t)n * sizeof(*k));
  rs = (aps_result*)malloc((size_t)n * sizeof(*rs));
  if (!k || !rs) {
    free(k);
    free(rs);
    return APS_ENOMEM;
  }
  for (i = 0; i < n; ++i) {
    k[i].file_hi = (uint64_t)keys[3 * i];
    k[i].file_lo = (uint64_t)keys[3 * i + 1];
    k[i].page_index = (uint64_t)keys[3 * i + 2];
  }
  rc = aps_delete_batch(s, (uint32_t)n, k, rs, NULL);
  for (i = 0; i < n && rc == APS_OK; ++i) out_status[i] = rs[i].len < 0 ? -rs[i].len : APS_OK;
  free(k);
  free(rs);
  return rc;
}

/* `alluxio fs checksum` for n files whose bytes lie in one direct buffer: digests (16 bytes each) into out. */
static inline int aps_marshal_md5_batch(int32_t device, int32_t n, const void* base, const int64_t* off, const int64_t* len,
                                 unsigned char* out) {
  const void** data;
  uint64_t* l;
  int32_t i;
  int rc;
  if (n <= 0) return APS_OK;
  data = (const void**)malloc((size_t)n * sizeof(*data));
  l = (uint64_t*)malloc((size_t)n * sizeof(*l));
  if (!data || !l) {
    free(data);
    free(l);
    return APS_ENOMEM;
  }
  for (i = 0; i < n; ++i) {
    data[i] = (const unsigned char*)base + off[i];
    l[i] = (uint64_t)len[i];
  }
  rc = aps_md5_batch(device, (uint32_t)n, data, l, (uint8_t(*)[16])out);
  free(data);
  free(l);
  return rc;
}
#endif
