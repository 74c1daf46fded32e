// aps_crc32c_noalloc: CRC32C of device-resident byte ranges for callers that must not allocate.
//
// Who cannot allocate.  In a client whose page cache is sized to fill HBM, the moment a background
// scrubber wants to re-check page checksums is exactly the moment there is no device memory left to
// ask for: cudaMalloc can fail, and even when it succeeds it (like cudaFree) takes a process-wide lock
// and synchronises the whole device, stalling every read thread that is in the middle of a batched
// get.  The same holds for a host process whose device memory is all owned by someone else's pool
// (a JVM embedding this library next to a framework that reserved the GPU up front).  Such callers
// get this entry point: everything it needs beyond the caller's own inputs and outputs lives in a
// fixed array inside the library image, so the call never asks the driver for memory.
//
// Price of that: the array is sized for the worst case below and is held by every GPU that loads
// the library until the process exits; all calls share it, so two calls that overlap on the GPU
// (different streams, or different host threads without ordering) corrupt each other's results.
//
// This translation unit is self-contained: its own scratch, its own constant set-up, its own job
// builder.  It reuses the fused-CRC streaming kernel (copy_kernel with no destination).
#include <cstddef>
#include <mutex>

#define APS_KERNELS_TEMPLATES_ONLY
#include "aps_kernels.cuh"

using namespace aps;

namespace {

// ---- fixed scratch ------------------------------------------------------------------------------
// Largest problem one call accepts.  Raising these grows the array below.
constexpr uint32_t kMaxRanges = APS_NOALLOC_MAX_RANGES;  // 1024 ranges per call
constexpr uint32_t kMaxLen = APS_NOALLOC_MAX_LEN;        // 4 MiB per range
constexpr uint32_t kMaxRows = kMaxLen / kRowBytes;

struct Scratch {
  Job jobs[kMaxRanges];          // one streaming job per range (written by the job builder each call)
  uint32_t acc[kMaxRanges];      // per-range XOR accumulator        -- must be zero when a call starts
  uint32_t done[kMaxRanges];     // per-range finished-row counter    -- must be zero when a call starts
  uint32_t sched[4];             // CTA rank counters of the kernel   -- must be zero when a call starts
  int error_flag;                // kernel could not place its tables (ruled out on the host at set-up) -- zeroed per call
  uint32_t pad[3];
  uint8_t sm_slow[256];          // all zero: no SM speed classes here (set once)
  CrcTables tables;              // GF(2) fold tables (set once)
  uint32_t rowshift[kMaxRows + 1];  // x^(8*512*k) (set once)
};
static_assert(sizeof(Scratch) <= 128 * 1024, "the scratch is meant to stay modest");

// The whole scratch: 128 KiB of device memory per GPU, part of the library image.
static __device__ __align__(256) unsigned char g_noalloc_scratch[128 * 1024];

constexpr size_t kZeroFrom = offsetof(Scratch, acc);
constexpr size_t kZeroBytes = offsetof(Scratch, sm_slow) - offsetof(Scratch, acc);

// Builds one checksum-only job per range.  Lengths above max_len are cut to max_len.
__global__ void noalloc_jobs_kernel(const void* const* data, const uint32_t* len, uint32_t n, uint32_t max_len, Job* jobs,
                                    uint32_t* crc_out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t l = min(len[i], max_len);
  jobs[i] = Job{(const uint8_t*)data[i], nullptr, l, 0, kJobOutCrc, 0};
  if (l == 0) crc_out[i] = 0;  // CRC32C of nothing
}

uint32_t host_xpow8(uint64_t nbytes) {
  uint32_t r = 0x80000000u, p = 0x00800000u;
  for (; nbytes; nbytes >>= 1) {
    if (nbytes & 1) r = gf_mul(r, p);
    p = gf_mul(p, p);
  }
  return r;
}

struct PerDevice {
  bool ready = false;
  Scratch* scratch = nullptr;  // DEVICE address of g_noalloc_scratch on this GPU
  int num_sms = 0;
};
std::mutex g_mu;
PerDevice g_dev[64];

constexpr int kThreads = 512, kGroup = 4;
const auto kKernel = copy_kernel<true, kThreads, kGroup, false>;

// Once per device, with that device current.  In host code the name g_noalloc_scratch is only a
// handle for the symbol, not something a kernel can dereference: the device address has to be asked
// for, here, in the file that defines the array.
int setup_device(int dev, PerDevice** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (dev < 0 || dev >= 64) return APS_EINVAL;
  PerDevice& d = g_dev[dev];
  if (!d.ready) {
    void* ptr = nullptr;
    if (cudaGetSymbolAddress(&ptr, g_noalloc_scratch) != cudaSuccess || ptr == nullptr) return APS_ECUDA;
    Scratch* sc = static_cast<Scratch*>(ptr);
    // host-side constants (stack / static host memory only)
    static CrcTables tables;
    static uint32_t rowshift[kMaxRows + 1];
    uint32_t xp[32], ls[32];
    const uint32_t xrow = host_xpow8(kRowBytes), x4 = host_xpow8(4), xrow4 = host_xpow8(4 * kRowBytes);
    for (uint32_t j = 0; j < 4; ++j)
      for (uint32_t b = 0; b < 256; ++b) {
        tables.fold[j][b] = gf_mul(b << (8 * j), xrow);
        tables.z4[j][b] = gf_mul(b << (8 * j), x4);
        tables.fold4[j][b] = gf_mul(b << (8 * j), xrow4);
      }
    rowshift[0] = 0x80000000u;
    for (uint32_t k = 1; k <= kMaxRows; ++k) rowshift[k] = gf_mul(rowshift[k - 1], xrow);
    for (int k = 0; k < 32; ++k) xp[k] = host_xpow8(1ull << k);
    for (int l = 0; l < 32; ++l) ls[l] = host_xpow8(500 - 16 * l);
    if (cudaMemset(sc, 0, sizeof(Scratch)) != cudaSuccess) return APS_ECUDA;
    if (cudaMemcpy(&sc->tables, &tables, sizeof(tables), cudaMemcpyHostToDevice) != cudaSuccess) return APS_ECUDA;
    if (cudaMemcpy(sc->rowshift, rowshift, sizeof(rowshift), cudaMemcpyHostToDevice) != cudaSuccess) return APS_ECUDA;
    if (cudaMemcpyToSymbol(c_xpow8, xp, sizeof(xp)) != cudaSuccess) return APS_ECUDA;
    if (cudaMemcpyToSymbol(c_lane_shift, ls, sizeof(ls)) != cudaSuccess) return APS_ECUDA;
    if (cudaFuncSetAttribute(kKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCopyDynSmem) != cudaSuccess)
      return APS_ECUDA;
    if (cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return APS_ECUDA;
    // The kernel keeps its CRC tables at fixed shared-memory addresses and gives up (sets Scratch::error_flag, writes
    // no checksum) if the dynamic window does not start below them.  This call cannot wait for the kernel to find
    // that out, so the condition is decided here, from what fixes it: the kernel's static shared memory plus the
    // per-block reserve.  If it does not hold, no call on this device is ever launched.
    cudaFuncAttributes fa{};
    int reserved = 0;
    if (cudaFuncGetAttributes(&fa, kKernel) != cudaSuccess) return APS_ECUDA;
    if (cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev) != cudaSuccess) return APS_ECUDA;
    if (fa.sharedSizeBytes + (size_t)reserved > kSmemP) return APS_ECUDA;
    // The set-up copies above ran on the legacy default stream; the caller's stream may be a non-blocking one that
    // does not order itself after it.  Wait here, once, so that the first call's kernels see the tables.
    if (cudaDeviceSynchronize() != cudaSuccess) return APS_ECUDA;
    d.scratch = sc;
    d.ready = true;
  }
  *out = &d;
  return APS_OK;
}

}  // namespace

// Declared in include/aps.h.
extern "C" APS_API int aps_crc32c_noalloc(int32_t device, uint32_t n, const void* const* d_data, const uint32_t* d_len,
                                          uint32_t max_len, uint32_t* d_crc, aps_stream stream) {
  if (n == 0) return APS_OK;
  if (!d_data || !d_len || !d_crc) return APS_EINVAL;
  if (n > kMaxRanges || max_len > kMaxLen) return APS_ELIMIT;  // beyond what the fixed scratch can hold
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  if (cudaSetDevice(device) != cudaSuccess) return APS_ECUDA;
  PerDevice* d = nullptr;
  int rc = setup_device(device, &d);
  if (rc == APS_OK) {
    cudaStream_t st = (cudaStream_t)stream;
    Scratch* sc = d->scratch;
    // the scratch still holds whatever the previous call left: clear the parts that must start at zero
    if (cudaMemsetAsync(reinterpret_cast<unsigned char*>(sc) + kZeroFrom, 0, kZeroBytes, st) != cudaSuccess) rc = APS_ECUDA;
    if (rc == APS_OK) {
      noalloc_jobs_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_data, d_len, n, max_len, sc->jobs, d_crc);
      if (cudaGetLastError() != cudaSuccess) rc = APS_ECUDA;
    }
    if (rc == APS_OK) {
      const uint32_t rows = (max_len + kRowBytes - 1) / kRowBytes;
      const uint32_t chip_warps = (uint32_t)d->num_sms * (kThreads / 32);
      uint32_t tpj = chip_warps / n;
      if (tpj < 1) tpj = 1;
      uint32_t cap = rows / 8 ? rows / 8 : 1;
      if (cap > kMaxTasksPerJob) cap = kMaxTasksPerJob;
      if (tpj > cap) tpj = cap;
      const uint32_t rpt = rows ? (rows + tpj - 1) / tpj : 1;
      tpj = rows ? (rows + rpt - 1) / rpt : 1;
      CopyArgs a{};
      a.jobs = sc->jobs;
      a.reqs = nullptr;
      a.crc_on = 1;
      a.n_jobs = n;
      a.tasks_per_job = tpj;
      a.rows_per_task = rpt;
      a.res = nullptr;
      a.crc_out = d_crc;
      a.acc = sc->acc;
      a.done = sc->done;
      a.tables = &sc->tables;
      a.rowshift = sc->rowshift;
      a.error_flag = &sc->error_flag;
      a.timeline = nullptr;
      a.steal = 1;
      a.sched = sc->sched;
      a.sm_slow = sc->sm_slow;
      const uint64_t items = (uint64_t)n * tpj;
      uint32_t grid = (uint32_t)(items < (uint64_t)d->num_sms ? items : (uint64_t)d->num_sms);
      if (grid == 0) grid = 1;
      kKernel<<<grid, kThreads, kCopyDynSmem, st>>>(a);
      if (cudaGetLastError() != cudaSuccess) rc = APS_ECUDA;
    }
  }
  if (prev >= 0) cudaSetDevice(prev);
  return rc;
}
