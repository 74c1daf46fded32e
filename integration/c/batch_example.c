/* The batched JNI path without a JVM: exactly what integration/jni/aps_jni.c does for GpuPageStore.putBatch / getBatch /
 * deleteBatch / md5Batch -- flat primitive arrays plus the address of ONE "direct buffer" -> one aps_*_batch call
 * (integration/jni/aps_marshal.h).  The direct buffer here is plain malloc'd memory, as ByteBuffer.allocateDirect gives.
 * Build:  gcc -std=c99 -Wall -I include -I integration/jni integration/c/batch_example.c -L alluxio_b200 -laps \
 *             -Wl,-rpath,$PWD/alluxio_b200 -o batch_example
 * Run on a machine with a B200:  ./batch_example        (prints "ok" and exits 0) */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "aps.h"
#include "aps_marshal.h"

#define N 48
#define PAGE 65536

static uint32_t crc32c_bitwise(const unsigned char* p, size_t n) {
  uint32_t r = 0xFFFFFFFFu;
  size_t i;
  int k;
  for (i = 0; i < n; ++i) {
    r ^= p[i];
    for (k = 0; k < 8; ++k) r = (r >> 1) ^ (0x82F63B78u & (0u - (r & 1u)));
  }
  return r ^ 0xFFFFFFFFu;
}

#define FAIL(msg) do { fprintf(stderr, "%s (line %d)\n", msg, __LINE__); return 1; } while (0)

int main(void) {
  aps_config cfg;
  aps_store* store = NULL;
  unsigned char* src = (unsigned char*)malloc((size_t)N * PAGE); /* the "direct buffers" */
  unsigned char* dst = (unsigned char*)malloc((size_t)(N + 1) * PAGE);
  int64_t keys[3 * (N + 1)], off[N + 1];
  int32_t len[N + 1], page_off[N + 1], nbytes[N + 1], status[N + 1], out_len[N + 1], crc_put[N + 1], crc_get[N + 1];
  int i, rc;
  if (!src || !dst) {
    fprintf(stderr, "out of memory\n");
    return 2;
  }
  memset(&cfg, 0, sizeof(cfg));
  cfg.device = 0;
  cfg.page_size = PAGE;
  cfg.cache_bytes = 256ull * PAGE;
  cfg.max_batch = 64;
  cfg.flags = APS_FLAG_CRC;
  rc = aps_open(&cfg, &store);
  if (rc != APS_OK) {
    fprintf(stderr, "aps_open: %s (%s)\n", aps_strerror(rc), aps_last_cuda_error());
    return 2;
  }
  for (i = 0; i < N * PAGE; ++i) src[i] = (unsigned char)((unsigned)i * 2654435761u >> 13);
  for (i = 0; i < N; ++i) {
    keys[3 * i] = 0;              /* decimal file id 17 + i / 8 */
    keys[3 * i + 1] = 17 + i / 8;
    keys[3 * i + 2] = i % 8;
    off[i] = (int64_t)i * PAGE;
    len[i] = i % 5 == 4 ? 1000 + i : PAGE; /* some short last pages */
  }
  if (aps_marshal_put_batch(store, N, keys, src, off, len, status, crc_put) != APS_OK) FAIL("put batch");
  for (i = 0; i < N; ++i)
    if (status[i] != APS_OK || (uint32_t)crc_put[i] != crc32c_bitwise(src + off[i], (size_t)len[i])) FAIL("put result");
  /* read everything back, plus one id nobody stored, one request at an offset */
  keys[3 * N] = 0; keys[3 * N + 1] = 999; keys[3 * N + 2] = 0;
  for (i = 0; i <= N; ++i) {
    page_off[i] = 0;
    nbytes[i] = PAGE;
    off[i] = (int64_t)i * PAGE;
  }
  page_off[3] = 100;
  memset(dst, 0xEE, (size_t)(N + 1) * PAGE);
  if (aps_marshal_get_batch(store, N + 1, keys, page_off, nbytes, dst, off, out_len, crc_get) != APS_OK) FAIL("get batch");
  if (out_len[N] != -APS_ENOTFOUND) FAIL("absent id");
  for (i = 0; i < N; ++i) {
    const int32_t want = len[i] - page_off[i];
    if (out_len[i] != want || memcmp(dst + off[i], src + (size_t)i * PAGE + page_off[i], (size_t)want) != 0) FAIL("get bytes");
    if ((uint32_t)crc_get[i] != crc32c_bitwise(dst + off[i], (size_t)want)) FAIL("get crc");
    if (want < PAGE && dst[off[i] + want] != 0xEE) FAIL("wrote past the returned length");
  }
  /* checksum path: MD5 of two byte ranges of a pinned buffer (RFC 1321 test vector "abc") */
  {
    void* pinned = NULL;
    int64_t moff[2] = {0, 64}, mlen[2] = {3, 0};
    unsigned char dig[32];
    static const unsigned char abc[16] = {0x90, 0x01, 0x50, 0x98, 0x3c, 0xd2, 0x4f, 0xb0, 0xd6, 0x96, 0x3f, 0x7d, 0x28, 0xe1, 0x7f, 0x72};
    static const unsigned char empty[16] = {0xd4, 0x1d, 0x8c, 0xd9, 0x8f, 0x00, 0xb2, 0x04, 0xe9, 0x80, 0x09, 0x98, 0xec, 0xf8, 0x42, 0x7e};
    if (aps_host_alloc_pinned(128, &pinned) != APS_OK) FAIL("pinned");
    memcpy(pinned, "abc", 3);
    if (aps_marshal_md5_batch(0, 2, pinned, moff, mlen, dig) != APS_OK) FAIL("md5 batch");
    if (memcmp(dig, abc, 16) != 0 || memcmp(dig + 16, empty, 16) != 0) FAIL("md5 digest");
    aps_host_free_pinned(pinned);
  }
  if (aps_marshal_delete_batch(store, N + 1, keys, status) != APS_OK) FAIL("delete batch");
  for (i = 0; i < N; ++i)
    if (status[i] != APS_OK) FAIL("delete result");
  if (status[N] != APS_ENOTFOUND) FAIL("delete of an absent id");
  if (aps_close(store) != APS_OK || aps_close(store) != APS_OK) FAIL("close twice"); /* idempotent */
  free(src);
  free(dst);
  printf("ok\n");
  return 0;
}
