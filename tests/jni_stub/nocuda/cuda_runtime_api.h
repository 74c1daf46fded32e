/* TEST INFRASTRUCTURE ONLY: lets integration/jni/aps_jni.c build for the JVM-free, GPU-free run of tests/test_jni_shim_runs.py.
 * The shim's only CUDA calls are the pinned allocation behind GpuPageStore.allocPinned; here that is plain memory. */
#ifndef NOCUDA_RUNTIME_API_H_
#define NOCUDA_RUNTIME_API_H_
#include <stdlib.h>
typedef int cudaError_t;
#define cudaSuccess 0
#define cudaHostAllocDefault 0
static cudaError_t cudaHostAlloc(void** p, size_t bytes, unsigned flags) {
  (void)flags;
  *p = malloc(bytes ? bytes : 1);
  return *p ? cudaSuccess : 2;
}
static cudaError_t cudaGetDeviceCount(int* n) {
  *n = 2; /* two "GPUs", so that the sharded natives have something to route between */
  return cudaSuccess;
}
static cudaError_t cudaFreeHost(void* p) {
  free(p);
  return cudaSuccess;
}
#endif
