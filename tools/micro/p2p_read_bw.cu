// Micro-benchmark: how fast can a kernel on GPU 0 read GPU 1's memory over NVLink (and both at once),
// with 128-bit and with 256-bit loads?  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o p2p_read_bw p2p_read_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int W>  // W = 16 or 32 bytes per lane
__global__ void __launch_bounds__(512) copy_k(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t bytes) {
  const size_t per_warp = 4 * 32 * W;  // four accesses in flight per lane
  const size_t warps = (size_t)gridDim.x * (blockDim.x / 32), w = (size_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x & 31;
  for (size_t off = w * per_warp; off + per_warp <= bytes; off += warps * per_warp) {
    if (W == 16) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + off + (u * 32 + lane) * 16));
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + off + (u * 32 + lane) * 16), "r"(v[u].x), "r"(v[u].y), "r"(v[u].z), "r"(v[u].w) : "memory");
    } else {
      uint32_t v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]), "=r"(v[u][4]), "=r"(v[u][5]), "=r"(v[u][6]), "=r"(v[u][7]) : "l"(src + off + (u * 32 + lane) * 32));
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst + off + (u * 32 + lane) * 32), "r"(v[u][0]), "r"(v[u][1]), "r"(v[u][2]), "r"(v[u][3]), "r"(v[u][4]), "r"(v[u][5]), "r"(v[u][6]), "r"(v[u][7]) : "memory");
    }
  }
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  const size_t bytes = 1ull << 30;
  uint8_t *src[2] = {nullptr, nullptr}, *dst[2] = {nullptr, nullptr};
  cudaStream_t st[2];
  cudaEvent_t e0[2], e1[2];
  for (int d = 0; d < 2 && d < n; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaMalloc(&src[d], bytes));
    CK(cudaMalloc(&dst[d], bytes));
    CK(cudaMemset(src[d], d + 1, bytes));
    CK(cudaStreamCreate(&st[d]));
    CK(cudaEventCreate(&e0[d]));
    CK(cudaEventCreate(&e1[d]));
    if (n > 1) cudaDeviceEnablePeerAccess(1 - d, 0);
  }
  auto run = [&](int d, const uint8_t* s, int width) {
    cudaSetDevice(d);
    if (width == 16) copy_k<16><<<148 * 2, 512, 0, st[d]>>>(s, dst[d], bytes);
    else copy_k<32><<<148 * 2, 512, 0, st[d]>>>(s, dst[d], bytes);
  };
  auto push = [&](int d, int width) {  // read local, write into the peer
    cudaSetDevice(d);
    if (width == 16) copy_k<16><<<148 * 2, 512, 0, st[d]>>>(src[d], dst[1 - d], bytes);
    else copy_k<32><<<148 * 2, 512, 0, st[d]>>>(src[d], dst[1 - d], bytes);
  };
  for (int width : {16, 32}) {
    // local copy, for reference
    run(0, src[0], width);
    cudaSetDevice(0); cudaEventRecord(e0[0], st[0]);
    for (int i = 0; i < 5; ++i) run(0, src[0], width);
    cudaEventRecord(e1[0], st[0]); cudaEventSynchronize(e1[0]);
    float ms; cudaEventElapsedTime(&ms, e0[0], e1[0]);
    printf("width %2d  local copy          : %.1f GB/s read (+ same written)\n", width, 5.0 * bytes / ms / 1e6);
    if (n < 2) continue;
    // one direction: GPU0 reads GPU1
    run(0, src[1], width);
    cudaEventRecord(e0[0], st[0]);
    for (int i = 0; i < 5; ++i) run(0, src[1], width);
    cudaEventRecord(e1[0], st[0]); cudaEventSynchronize(e1[0]);
    cudaEventElapsedTime(&ms, e0[0], e1[0]);
    printf("width %2d  peer read, one way  : %.1f GB/s\n", width, 5.0 * bytes / ms / 1e6);
    // both directions at once
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventRecord(e0[d], st[d]); }
    for (int i = 0; i < 5; ++i) for (int d = 0; d < 2; ++d) run(d, src[1 - d], width);
    float worst = 0;
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventRecord(e1[d], st[d]); }
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventSynchronize(e1[d]); cudaEventElapsedTime(&ms, e0[d], e1[d]); worst = ms > worst ? ms : worst; }
    printf("width %2d  peer read, both ways: %.1f GB/s per GPU\n", width, 5.0 * bytes / worst / 1e6);
    // peer writes: one way, then both ways
    push(0, width);
    cudaSetDevice(0); cudaEventRecord(e0[0], st[0]);
    for (int i = 0; i < 5; ++i) push(0, width);
    cudaEventRecord(e1[0], st[0]); cudaEventSynchronize(e1[0]);
    cudaEventElapsedTime(&ms, e0[0], e1[0]);
    printf("width %2d  peer write, one way : %.1f GB/s\n", width, 5.0 * bytes / ms / 1e6);
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventRecord(e0[d], st[d]); }
    for (int i = 0; i < 5; ++i) for (int d = 0; d < 2; ++d) push(d, width);
    worst = 0;
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventRecord(e1[d], st[d]); }
    for (int d = 0; d < 2; ++d) { cudaSetDevice(d); cudaEventSynchronize(e1[d]); cudaEventElapsedTime(&ms, e0[d], e1[d]); worst = ms > worst ? ms : worst; }
    printf("width %2d  peer write, both ways: %.1f GB/s per GPU\n", width, 5.0 * bytes / worst / 1e6);
  }
  return 0;
}
