/* TEST INFRASTRUCTURE ONLY: aps_md5_batch for the GPU-free runs of the plain-C callers and of the JNI shim, answered by
 * the oracle's RFC 1321 MD5 (oracle/checksums.c).  The MD5 kernel itself is tested in tests/test_gpu_md5.py. */
#include <stddef.h>
#include <stdint.h>

#include "../../include/aps.h"

void orc_md5(const void* data, size_t n, uint8_t out[16]); /* oracle/checksums.c */

int aps_md5_batch(int32_t device, uint32_t n_files, const void* const* data, const uint64_t* len, uint8_t (*out)[16]) {
  uint32_t i;
  (void)device;
  for (i = 0; i < n_files; ++i) orc_md5(data[i], (size_t)len[i], out[i]);
  return APS_OK;
}
