// Random-access read rate of HBM on one GPU: the roofline of the K1 hash-table kernels.
// Each group of G lanes reads one random, G*32-byte aligned granule (G = 1: one 32-byte sector per
// lane; G = 8: one 256-byte bucket per eight lanes, the shape of a table probe) of a buffer much
// larger than L2.  Prints granules/s and GB/s for both.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o random_sector_rate random_sector_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                    \
  do {                                                                           \
    cudaError_t e_ = (x);                                                        \
    if (e_ != cudaSuccess) {                                                     \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

template <int G>
__global__ void probe(const uint4* __restrict__ buf, uint64_t granule_mask, uint64_t n_granules, uint64_t seed,
                      unsigned long long* sink) {
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t g = gid / G, sub = gid % G;
  if (g >= n_granules) return;
  const uint64_t where = mix64(g + seed) & granule_mask;  // granule index
  const uint4* p = buf + (where * G + sub) * 2;           // 32 bytes per lane
  uint4 a = __ldcg(p), b = __ldcg(p + 1);
  uint32_t x = a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w;
  if (x == 0x12345678u) atomicAdd(sink, 1ull);  // never true for the zero buffer; keeps the loads alive
}

template <int G>
static int run(const uint4* buf, uint64_t bytes, unsigned long long* sink) {
  const uint64_t granule = 32ull * G, n = 1ull << 24;  // 16 M random granules per launch
  const uint64_t mask = bytes / granule - 1;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const int threads = 256;
  const uint64_t blocks = (n * G + threads - 1) / threads;
  for (int w = 0; w < 3; ++w) probe<G><<<(unsigned)blocks, threads>>>(buf, mask, n, 1000 + w, sink);
  CK(cudaDeviceSynchronize());
  const int reps = 10;
  CK(cudaEventRecord(e0));
  for (int r = 0; r < reps; ++r) probe<G><<<(unsigned)blocks, threads>>>(buf, mask, n, r, sink);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double per_s = (double)n * reps / (ms * 1e-3);
  printf("{\"granule_bytes\": %llu, \"buffer_GiB\": %.1f, \"Mgranules_per_s\": %.1f, \"GBps\": %.1f}\n",
         (unsigned long long)granule, bytes / 1073741824.0, per_s / 1e6, per_s * granule / 1e9);
  return 0;
}

int main(int argc, char** argv) {
  const uint64_t bytes = (argc > 1 ? strtoull(argv[1], nullptr, 10) : 8ull) << 30;  // power of two GiB
  uint4* buf;
  unsigned long long* sink;
  CK(cudaMalloc(&buf, bytes));
  CK(cudaMemset(buf, 0, bytes));
  CK(cudaMalloc(&sink, 8));
  CK(cudaMemset(sink, 0, 8));
  if (run<1>(buf, bytes, sink)) return 1;
  if (run<8>(buf, bytes, sink)) return 1;
  return 0;
}
