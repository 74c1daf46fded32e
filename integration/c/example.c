/* Plain C user of the page-store ABI: what a JNI shim or any C embedder does, without Python or C++.
 * Build:  gcc -std=c99 -Wall -I include integration/c/example.c -L alluxio_b200 -laps -Wl,-rpath,$PWD/alluxio_b200 -o example
 * Run on a machine with a B200:  ./example        (prints "ok" and exits 0)
 * It stores a page, reads part of it back, checks the CRC32C the store reports, deletes the page. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "aps.h"

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

int main(void) {
  aps_config cfg;
  aps_store* store = NULL;
  aps_key key = {0, 42, 7}; /* file id "42", page index 7 */
  unsigned char *page, *back;
  uint32_t got = 0, crc = 0, i;
  int rc;
  memset(&cfg, 0, sizeof(cfg));
  cfg.device = 0;
  cfg.page_size = 1u << 20;
  cfg.cache_bytes = 64ull << 20;
  cfg.flags = APS_FLAG_CRC;
  rc = aps_open(&cfg, &store);
  if (rc != APS_OK) {
    fprintf(stderr, "aps_open: %s (%s)\n", aps_strerror(rc), aps_last_cuda_error());
    return 2;
  }
  page = (unsigned char*)malloc(cfg.page_size);
  back = (unsigned char*)malloc(cfg.page_size);
  if (!page || !back) {
    fprintf(stderr, "out of memory\n");
    return 2;
  }
  for (i = 0; i < cfg.page_size; ++i) page[i] = (unsigned char)(i * 31u + (i >> 8));
  rc = aps_put(store, &key, page, cfg.page_size);
  if (rc == APS_OK) rc = aps_get(store, &key, 0, cfg.page_size, back, cfg.page_size, &got, &crc);
  if (rc != APS_OK || got != cfg.page_size || memcmp(page, back, got) != 0 || crc != crc32c_bitwise(page, got)) {
    fprintf(stderr, "round trip failed: rc=%d got=%u\n", rc, got);
    return 1;
  }
  rc = aps_get(store, &key, 1000, 24, back, 24, &got, NULL); /* offset read */
  if (rc != APS_OK || got != 24 || memcmp(page + 1000, back, 24) != 0) return 1;
  if (aps_delete(store, &key) != APS_OK || aps_get(store, &key, 0, 1, back, 1, &got, NULL) != APS_ENOTFOUND) return 1;
  free(page);
  free(back);
  aps_close(store);
  printf("ok\n");
  return 0;
}
