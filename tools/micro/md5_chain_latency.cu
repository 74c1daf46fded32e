// Micro-benchmark: how many cycles does ONE MD5 step cost a single warp when nothing but the step's own dependent chain
// limits it?  The MD5 kernel (aps_kernels.cuh, md5_block) keeps a + m + k off the chain, which leaves three dependent
// operations per step (SASS: LOP3 -> IMAD.IADD/IADD3 -> LEA.HI); with one warp per scheduler nothing hides their
// latency, so cycles per step = 3 x the dependent-issue latency of those operations on this GPU.  This program runs the
// same step macro on registers only (no shared or global memory inside the loop), one warp per block, and reports
// cycles per step for the four round functions -- the bound that `fs checksum` throughput = files x f_clk x 64 B /
// (64 x cycles per step) is quoted against in DESIGN.md 4.3.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o md5_chain_latency md5_chain_latency.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int s) { return __funnelshift_l(x, x, s); }
#define STEP(f, a, b, c, d, m, k, s)                                  \
  do {                                                                \
    uint32_t pre_;                                                    \
    asm("add.u32 %0, %1, %2;" : "=r"(pre_) : "r"(a), "r"((m) + (k))); \
    a = b + rotl32(f(b, c, d) + pre_, s);                             \
  } while (0)
#define F1(b, c, d) ((d) ^ ((b) & ((c) ^ (d))))
#define F2(b, c, d) ((c) ^ ((d) & ((b) ^ (c))))
#define F3(b, c, d) ((b) ^ (c) ^ (d))
#define F4(b, c, d) ((c) ^ ((b) | ~(d)))

template <int kRound>
__global__ void __launch_bounds__(32) chain(uint32_t* out, long long* cycles, int blocks16, uint32_t m0) {
  uint32_t a = 0x67452301u + threadIdx.x, b = 0xefcdab89u, c = 0x98badcfeu, d = 0x10325476u;
  uint32_t m[4] = {m0, m0 * 3u, m0 * 5u, m0 * 7u};
  const long long t0 = clock64();
  for (int i = 0; i < blocks16; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {  // 16 steps per iteration, the rotation amounts of the round
      if (kRound == 1) { STEP(F1, a, b, c, d, m[0], 0xd76aa478u, 7); STEP(F1, d, a, b, c, m[1], 0xe8c7b756u, 12); STEP(F1, c, d, a, b, m[2], 0x242070dbu, 17); STEP(F1, b, c, d, a, m[3], 0xc1bdceeeu, 22); }
      if (kRound == 2) { STEP(F2, a, b, c, d, m[0], 0xf61e2562u, 5); STEP(F2, d, a, b, c, m[1], 0xc040b340u, 9); STEP(F2, c, d, a, b, m[2], 0x265e5a51u, 14); STEP(F2, b, c, d, a, m[3], 0xe9b6c7aau, 20); }
      if (kRound == 3) { STEP(F3, a, b, c, d, m[0], 0xfffa3942u, 4); STEP(F3, d, a, b, c, m[1], 0x8771f681u, 11); STEP(F3, c, d, a, b, m[2], 0x6d9d6122u, 16); STEP(F3, b, c, d, a, m[3], 0xfde5380cu, 23); }
      if (kRound == 4) { STEP(F4, a, b, c, d, m[0], 0xf4292244u, 6); STEP(F4, d, a, b, c, m[1], 0x432aff97u, 10); STEP(F4, c, d, a, b, m[2], 0xab9423a7u, 15); STEP(F4, b, c, d, a, m[3], 0xfc93a039u, 21); }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * 32 + threadIdx.x] = a ^ b ^ c ^ d;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

int main() {
  uint32_t* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 32 * 4);
  cudaMalloc(&cyc, 148 * 8);
  const int iters = 20000;  // x 16 steps
  for (int round = 1; round <= 4; ++round) {
    for (int rep = 0; rep < 2; ++rep) {
      if (round == 1) chain<1><<<128, 32>>>(out, cyc, iters, 12345u);
      if (round == 2) chain<2><<<128, 32>>>(out, cyc, iters, 12345u);
      if (round == 3) chain<3><<<128, 32>>>(out, cyc, iters, 12345u);
      if (round == 4) chain<4><<<128, 32>>>(out, cyc, iters, 12345u);
      cudaDeviceSynchronize();
    }
    long long h[128];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double sum = 0;
    for (long long v : h) sum += (double)v;
    printf("{\"md5_round\": %d, \"cycles_per_step_bare_chain\": %.2f, \"warps\": 128, \"steps\": %d}\n", round, sum / 128 / (iters * 16.0), iters * 16);
  }
  return 0;
}
