// Micro-benchmark: all-to-all peer traffic over NVLink / NVSwitch with plain SM loads and stores, every GPU of the
// box active at once -- the ceiling for the cross-shard multi-get.
//   pull: every GPU reads 2 KiB chunks round-robin from all its peers and writes them to local memory (what
//         aps_get_batch_sharded does: the requester reads the owner's slab in place)
//   push: every GPU reads local memory and writes the chunks round-robin into its peers' memory
//   both: every GPU does both at once on two streams (half the chunks each): do the two directions of use add up?
// Reports GB/s per GPU (bytes that crossed NVLink into / out of that GPU), min and mean over GPUs.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o p2p_all_bw p2p_all_bw.cu
#include <cstdint>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

struct Peers {
  uint8_t* p[8];
  int n;
};

// chunk c of this GPU's share: 2 KiB = 4 rows of 512 B, one 16-byte access per lane per row (the copy kernel's shape)
template <bool kPull>
__global__ void __launch_bounds__(512) xfer(Peers remote, uint8_t* local, size_t chunks) {
  const size_t warps = (size_t)gridDim.x * (blockDim.x / 32), w = (size_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x & 31;
  for (size_t c = w; c < chunks; c += warps) {
    uint8_t* far = remote.p[c % remote.n] + (c / remote.n) * 2048;
    uint8_t* near = local + c * 2048;
    const uint8_t* src = kPull ? far : near;
    uint8_t* dst = kPull ? near : far;
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + (u * 32 + lane) * 16));
#pragma unroll
    for (int u = 0; u < 4; ++u)
      asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + (u * 32 + lane) * 16), "r"(v[u].x), "r"(v[u].y), "r"(v[u].z), "r"(v[u].w) : "memory");
  }
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("{\"error\": \"needs at least 2 GPUs\"}\n"); return 0; }
  if (n > 8) n = 8;
  const size_t bytes = 1ull << 30;  // per GPU per run
  std::vector<uint8_t*> buf(n), loc(n);
  std::vector<cudaStream_t> st(n);
  std::vector<cudaEvent_t> e0(n), e1(n);
  for (int g = 0; g < n; ++g) {
    CK(cudaSetDevice(g));
    for (int p = 0; p < n; ++p)
      if (p != g) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { printf("peer access %d->%d: %s\n", g, p, cudaGetErrorString(e)); return 1; }
        cudaGetLastError();
      }
    CK(cudaMalloc(&buf[g], bytes));  // what peers read from / write into
    CK(cudaMalloc(&loc[g], bytes));
    CK(cudaMemset(buf[g], g + 1, bytes));
    CK(cudaMemset(loc[g], 0, bytes));
    CK(cudaStreamCreate(&st[g]));
    CK(cudaEventCreate(&e0[g]));
    CK(cudaEventCreate(&e1[g]));
  }
  std::vector<cudaStream_t> st2(n);
  std::vector<uint8_t*> loc2(n);
  for (int g = 0; g < n; ++g) {
    CK(cudaSetDevice(g));
    CK(cudaStreamCreate(&st2[g]));
    CK(cudaMalloc(&loc2[g], bytes));
    CK(cudaMemset(loc2[g], 0, bytes));
  }
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 3; ++rep) {  // the last repetition is reported
      for (int g = 0; g < n; ++g) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
      for (int g = 0; g < n; ++g) {
        CK(cudaSetDevice(g));
        Peers pr{};
        pr.n = n - 1;
        for (int p = 0, k = 0; p < n; ++p)
          if (p != g) pr.p[k++] = buf[p] + (size_t)g * (bytes / n);  // every reader/writer gets its own 1/n slice of a peer
        const size_t chunks = (bytes / n) * (n - 1) / 2048;          // so that slices never overlap
        CK(cudaEventRecord(e0[g], st[g]));
        if (mode == 2) {
          // pull from the first half of every peer slice on one stream, push into the second half on another;
          // 74 CTAs each so that the two kernels share the SMs
          Peers hi = pr;
          for (int k = 0; k < hi.n; ++k) hi.p[k] += (bytes / n) / 2;
          CK(cudaStreamWaitEvent(st2[g], e0[g], 0));
          for (int it = 0; it < 4; ++it) {
            xfer<true><<<74, 512, 0, st[g]>>>(pr, loc[g], chunks / 2);
            xfer<false><<<74, 512, 0, st2[g]>>>(hi, loc2[g], chunks / 2);
          }
          cudaEvent_t done;
          CK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
          CK(cudaEventRecord(done, st2[g]));
          CK(cudaStreamWaitEvent(st[g], done, 0));
        } else {
          for (int it = 0; it < 4; ++it) {
            if (mode == 0) xfer<true><<<148, 512, 0, st[g]>>>(pr, loc[g], chunks);
            else xfer<false><<<148, 512, 0, st[g]>>>(pr, loc[g], chunks);
          }
        }
        CK(cudaEventRecord(e1[g], st[g]));
      }
      double mn = 1e30, sum = 0;
      for (int g = 0; g < n; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaEventSynchronize(e1[g]));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0[g], e1[g]));
        const double gbps = 4.0 * (double)((bytes / n) * (n - 1)) / (ms * 1e-3) / 1e9;
        mn = gbps < mn ? gbps : mn;
        sum += gbps;
      }
      if (rep == 2)
        printf("{\"mode\": \"%s\", \"gpus\": %d, \"per_gpu_GBps_min\": %.1f, \"per_gpu_GBps_mean\": %.1f}\n", mode == 0 ? "pull" : mode == 1 ? "push" : "both", n,
               mn, sum / n);
    }
  }
  return 0;
}
