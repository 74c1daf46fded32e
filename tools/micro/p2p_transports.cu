// Micro-benchmark: which transport moves pages between GPUs fastest when EVERY GPU of the box is busy at once
// (the shape of the cross-shard multi-get)?  One process, all GPUs, peer access enabled.
//   pull128 / push128   SM 128-bit loads from / stores into all peers (2 KiB chunks round-robin over the peers)
//   pull256 / push256   the same with 256-bit accesses
//   tma_pull / tma_push bulk-async copies: peer -> shared -> local, local -> shared -> peer (one thread per CTA)
//   ce_pull / ce_push   cudaMemcpyAsync of 1 MiB (page-sized) and 4 MiB pieces on four streams per GPU, one host
//                       thread per GPU issuing them
//   pull128+ce / push128+ce  a fraction of the bytes moved by the SMs and the rest by the copy engines, at once
// Reports GB/s per GPU that crossed NVLink (min and mean over GPUs) as JSON lines.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o p2p_transports p2p_transports.cu -lpthread
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

struct Peers {
  uint8_t* p[8];
  int n;
};

template <bool kPull, int W>  // W = 16 or 32 bytes per lane; chunk = 4 accesses per lane
__global__ void __launch_bounds__(512) xfer(Peers remote, uint8_t* local, size_t chunks) {
  constexpr size_t kChunk = 4 * 32 * W;
  const size_t warps = (size_t)gridDim.x * (blockDim.x / 32), w = (size_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  const int lane = threadIdx.x & 31;
  for (size_t c = w; c < chunks; c += warps) {
    uint8_t* far = remote.p[c % remote.n] + (c / remote.n) * kChunk;
    uint8_t* near = local + c * kChunk;
    const uint8_t* src = kPull ? far : near;
    uint8_t* dst = kPull ? near : far;
    if (W == 16) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + (u * 32 + lane) * 16));
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + (u * 32 + lane) * 16), "r"(v[u].x), "r"(v[u].y), "r"(v[u].z), "r"(v[u].w) : "memory");
    } else {
      uint32_t v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]), "=r"(v[u][4]), "=r"(v[u][5]), "=r"(v[u][6]), "=r"(v[u][7]) : "l"(src + (u * 32 + lane) * 32));
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(dst + (u * 32 + lane) * 32), "r"(v[u][0]), "r"(v[u][1]), "r"(v[u][2]), "r"(v[u][3]), "r"(v[u][4]), "r"(v[u][5]), "r"(v[u][6]), "r"(v[u][7]) : "memory");
    }
  }
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}

// Bulk-async ring: one thread per CTA keeps NS stages of kStage bytes in flight; data never enters registers.
template <bool kPull, int kStage, int NS>
__global__ void __launch_bounds__(32) tma_xfer(Peers remote, uint8_t* local, size_t chunks) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) unsigned long long full[NS];
  if (threadIdx.x != 0) return;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem), bbase = (uint32_t)__cvta_generic_to_shared(full);
  for (int s = 0; s < NS; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bbase + s * 8));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const size_t my = chunks > blockIdx.x ? (chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto ends = [&](size_t i, const uint8_t*& src, uint8_t*& dst) {
    const size_t c = blockIdx.x + i * gridDim.x;
    uint8_t* far = remote.p[c % remote.n] + (c / remote.n) * (size_t)kStage;
    uint8_t* near = local + c * (size_t)kStage;
    src = kPull ? far : near;
    dst = kPull ? near : far;
  };
  auto load = [&](size_t i) {
    const uint8_t* src; uint8_t* dst;
    ends(i, src, dst);
    const uint32_t s = (uint32_t)(i % NS);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bbase + s * 8), "r"(kStage) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sbase + s * kStage), "l"(src), "r"(kStage), "r"(bbase + s * 8) : "memory");
  };
  for (size_t i = 0; i < my && i < NS; ++i) load(i);
  for (size_t i = 0; i < my; ++i) {
    const uint32_t s = (uint32_t)(i % NS);
    mbar_wait(bbase + s * 8, (uint32_t)(i / NS) & 1u);
    const uint8_t* src; uint8_t* dst;
    ends(i, src, dst);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(sbase + s * kStage), "r"(kStage) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (i >= 1 && i - 1 + NS < my) {  // the store before this one has finished reading its stage: refill it
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      load(i - 1 + NS);
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("{\"error\": \"needs at least 2 GPUs\"}\n"); return 0; }
  if (n > 8) n = 8;
  const size_t bytes = 1ull << 30;  // per GPU per pass
  constexpr int kStreams = 4;
  std::vector<uint8_t*> buf(n), loc(n);
  std::vector<std::vector<cudaStream_t>> st(n, std::vector<cudaStream_t>(kStreams + 1));
  std::vector<cudaEvent_t> e0(n), e1(n);
  std::vector<std::vector<cudaEvent_t>> join(n, std::vector<cudaEvent_t>(kStreams + 1));
  for (int g = 0; g < n; ++g) {
    CK(cudaSetDevice(g));
    for (int p = 0; p < n; ++p)
      if (p != g) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { printf("peer access %d->%d: %s\n", g, p, cudaGetErrorString(e)); return 1; }
        cudaGetLastError();
      }
    CK(cudaMalloc(&buf[g], bytes));
    CK(cudaMalloc(&loc[g], bytes));
    CK(cudaMemset(buf[g], g + 1, bytes));
    CK(cudaMemset(loc[g], 0, bytes));
    for (auto& s : st[g]) CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    for (auto& e : join[g]) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaEventCreate(&e0[g]));
    CK(cudaEventCreate(&e1[g]));
  }
  CK(cudaFuncSetAttribute(tma_xfer<true, 16384, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
  CK(cudaFuncSetAttribute(tma_xfer<false, 16384, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
  CK(cudaFuncSetAttribute(tma_xfer<true, 32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6));
  CK(cudaFuncSetAttribute(tma_xfer<false, 32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6));
  for (int g = 0; g < n; ++g) {  // the attribute is per device
    CK(cudaSetDevice(g));
    CK(cudaFuncSetAttribute(tma_xfer<true, 16384, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
    CK(cudaFuncSetAttribute(tma_xfer<false, 16384, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
    CK(cudaFuncSetAttribute(tma_xfer<true, 32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6));
    CK(cudaFuncSetAttribute(tma_xfer<false, 32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6));
  }

  // Every GPU moves `share` = (bytes / n) * (n - 1) bytes per pass: its own 1/n slice of every peer's buffer.
  const size_t slice = bytes / n, share = slice * (n - 1);
  struct Mode {
    const char* name;
    int sm;        // 0 none, 1 pull128, 2 push128, 3 pull256, 4 push256, 5 tma_pull16k, 6 tma_push16k, 7 tma_pull32k, 8 tma_push32k
    int ce;        // 0 none, 1 pull, 2 push
    size_t piece;  // CE piece size
    int sm_pct;    // share of the bytes the SM kernel takes when both are on
  };
  const Mode modes[] = {
      {"pull128", 1, 0, 0, 100},          {"push128", 2, 0, 0, 100},          {"pull256", 3, 0, 0, 100},
      {"push256", 4, 0, 0, 100},          {"tma_pull_16k", 5, 0, 0, 100},     {"tma_push_16k", 6, 0, 0, 100},
      {"tma_pull_32k", 7, 0, 0, 100},     {"tma_push_32k", 8, 0, 0, 100},     {"ce_pull_1m", 0, 1, 1 << 20, 0},
      {"ce_push_1m", 0, 2, 1 << 20, 0},   {"ce_pull_4m", 0, 1, 4 << 20, 0},   {"ce_push_4m", 0, 2, 4 << 20, 0},
      {"pull128+ce_pull_1m(75/25)", 1, 1, 1 << 20, 75}, {"pull128+ce_pull_1m(50/50)", 1, 1, 1 << 20, 50},
      {"push128+ce_push_1m(75/25)", 2, 2, 1 << 20, 75}, {"push128+ce_push_1m(50/50)", 2, 2, 1 << 20, 50},
      {"tma_push_16k+ce_push_1m(50/50)", 6, 2, 1 << 20, 50}, {"pull128+ce_push_1m(50/50)", 1, 2, 1 << 20, 50},
  };
  constexpr int kPasses = 4;
  for (const Mode& m : modes) {
    double mn = 0, mean = 0, issue_us = 0;
    for (int rep = 0; rep < 2; ++rep) {  // the second repetition is reported
      for (int g = 0; g < n; ++g) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
      std::atomic<int> ready{0};
      std::vector<double> host_us(n, 0.0);
      auto work = [&](int g) {
        CK(cudaSetDevice(g));
        Peers pr{};
        pr.n = n - 1;
        for (int p = 0, k = 0; p < n; ++p)
          if (p != g) pr.p[k++] = buf[p] + (size_t)g * slice;
        // SM part: the first sm_bytes of every peer slice; CE part: the rest, in pieces
        const size_t sm_slice = (slice * (size_t)m.sm_pct / 100) & ~((size_t)(4 << 20) - 1);
        const size_t sm_bytes = m.sm ? (m.ce ? sm_slice : slice) : 0;
        ready.fetch_add(1);
        while (ready.load() < n) {}
        CK(cudaEventRecord(e0[g], st[g][0]));
        for (int k = 1; k <= kStreams; ++k) CK(cudaStreamWaitEvent(st[g][k], e0[g], 0));
        cudaEvent_t t0, t1;
        (void)t0; (void)t1;
        const auto h0 = std::chrono::steady_clock::now();
        for (int pass = 0; pass < kPasses; ++pass) {
          if (sm_bytes) {
            const size_t total = sm_bytes * (n - 1);
            switch (m.sm) {
              case 1: xfer<true, 16><<<148, 512, 0, st[g][0]>>>(pr, loc[g], total / 2048); break;
              case 2: xfer<false, 16><<<148, 512, 0, st[g][0]>>>(pr, loc[g], total / 2048); break;
              case 3: xfer<true, 32><<<148, 512, 0, st[g][0]>>>(pr, loc[g], total / 4096); break;
              case 4: xfer<false, 32><<<148, 512, 0, st[g][0]>>>(pr, loc[g], total / 4096); break;
              case 5: tma_xfer<true, 16384, 8><<<148, 32, 16384 * 8, st[g][0]>>>(pr, loc[g], total / 16384); break;
              case 6: tma_xfer<false, 16384, 8><<<148, 32, 16384 * 8, st[g][0]>>>(pr, loc[g], total / 16384); break;
              case 7: tma_xfer<true, 32768, 6><<<148, 32, 32768 * 6, st[g][0]>>>(pr, loc[g], total / 32768); break;
              case 8: tma_xfer<false, 32768, 6><<<148, 32, 32768 * 6, st[g][0]>>>(pr, loc[g], total / 32768); break;
            }
          }
          if (m.ce) {
            int k = 0;
            for (size_t off = sm_bytes; off + m.piece <= slice; off += m.piece)
              for (int p = 0; p < n - 1; ++p, ++k) {
                uint8_t* far = pr.p[p] + off;
                uint8_t* near = loc[g] + (size_t)p * slice + off;
                if (m.ce == 1) CK(cudaMemcpyAsync(near, far, m.piece, cudaMemcpyDeviceToDevice, st[g][1 + k % kStreams]));
                else CK(cudaMemcpyAsync(far, near, m.piece, cudaMemcpyDeviceToDevice, st[g][1 + k % kStreams]));
              }
          }
        }
        host_us[g] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - h0).count();
        for (int k = 1; k <= kStreams; ++k) {
          CK(cudaEventRecord(join[g][k], st[g][k]));
          CK(cudaStreamWaitEvent(st[g][0], join[g][k], 0));
        }
        CK(cudaEventRecord(e1[g], st[g][0]));
        CK(cudaGetLastError());
      };
      std::vector<std::thread> th;
      for (int g = 0; g < n; ++g) th.emplace_back(work, g);
      for (auto& t : th) t.join();
      mn = 1e30; mean = 0; issue_us = 0;
      for (int g = 0; g < n; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaEventSynchronize(e1[g]));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0[g], e1[g]));
        const double gbps = (double)kPasses * (double)share / (ms * 1e-3) / 1e9;
        mn = gbps < mn ? gbps : mn;
        mean += gbps / n;
        issue_us = host_us[g] > issue_us ? host_us[g] : issue_us;
      }
    }
    printf("{\"mode\": \"%s\", \"gpus\": %d, \"per_gpu_GBps_min\": %.1f, \"per_gpu_GBps_mean\": %.1f, \"host_issue_us_per_pass\": %.0f}\n", m.name, n, mn,
           mean, issue_us / kPasses);
    fflush(stdout);
  }
  return 0;
}
