// libaps: host side of the C ABI in include/aps.h.  Owns the HBM slab, the hash table and the scratch
// buffers of one store, stages request arrays, launches the kernels in aps_kernels.cuh.
// There is NO CPU fallback: every data-path entry point needs a CUDA device and fails with APS_ECUDA
// when the runtime cannot provide one.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_set>
#include <vector>

#include "aps_internal.h"
#include "aps_kernels.cuh"

using namespace aps;

namespace {

thread_local char tl_cuda_error[512] = "";
std::atomic<uint64_t> g_launches{0};

#define CUDA_TRY(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess) {                                                                             \
      snprintf(tl_cuda_error, sizeof(tl_cuda_error), "%s:%d %s -> %s", __FILE__, __LINE__, #expr,        \
               cudaGetErrorString(_e));                                                                  \
      return APS_ECUDA;                                                                                  \
    }                                                                                                    \
  } while (0)

struct DeviceGuard {  // the library never leaves the caller's current device changed
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// ---- per-device constants: CRC tables and shift constants --------------------------------------
uint32_t host_xpow8(uint64_t nbytes) {  // x^(8*nbytes) mod P by square and multiply
  uint32_t r = 0x80000000u;  // x^0 (bit 31 is x^0)
  uint32_t p = 0x00800000u;  // x^8
  for (; nbytes; nbytes >>= 1) {
    if (nbytes & 1) r = gf_mul(r, p);
    p = gf_mul(p, p);
  }
  return r;
}

struct DeviceConsts {
  bool ready = false;
  CrcTables* d_tables = nullptr;
  uint8_t* d_table_image = nullptr;  // shared-memory image of the tables (bytes kSmemG..kSmemEnd of the window)
  uint8_t* d_table_image4 = nullptr; // the same with the replicated tables carrying four rows ahead (kernels with kIndep)
  int num_sms = 0;
  // SM speed classes (see calibrate_sms): d_sm_slow[smid] = 1 for SMs that stream measurably slower
  uint8_t* d_sm_slow = nullptr;
  int n_slow = 0;
  bool calibrated = false;
};
std::mutex g_dev_mu;
DeviceConsts g_dev[64];

// Launch shapes of the copy kernel (threads per CTA x rows per pipelined group).  The default was
// picked from measurements on B200 (profiles/); APS_COPY_VARIANT=<threads>x<group> overrides it.
typedef void (*CopyKernelFn)(CopyArgs);
struct CopyVariant {
  const char* name;
  uint32_t threads, group;
  CopyKernelFn crc, plain, get_crc, get_plain, put_crc, put_plain;
};
#define APS_VARIANT(T, G)                                                                                  \
  {#T "x" #G, T, G, copy_kernel<true, T, G, false>, copy_kernel<false, T, G, false>, copy_kernel<true, T, G, true>, \
   copy_kernel<false, T, G, true>, copy_kernel<true, T, G, false, true>, copy_kernel<false, T, G, false, true>}
const CopyVariant kCopyVariants[] = {
    APS_VARIANT(512, 4), APS_VARIANT(1024, 4), APS_VARIANT(1024, 2), APS_VARIANT(512, 8),
    APS_VARIANT(512, 2), APS_VARIANT(256, 8),
};
#undef APS_VARIANT
// The bulk-async (TMA) form of the copy kernel, selected with APS_COPY_VARIANT=tma.
bool use_tma_variant() {
  static const bool v = [] {
    const char* want = getenv("APS_COPY_VARIANT");
    return want && !strcmp(want, "tma");
  }();
  return v;
}
struct TmaKernels {
  CopyKernelFn crc, plain, get_crc, get_plain;
  int dyn_crc = 0, dyn_plain = 0;
};
TmaKernels g_tma = {tma_copy_kernel<true, false>, tma_copy_kernel<false, false>, tma_copy_kernel<true, true>,
                    tma_copy_kernel<false, true>};

const CopyVariant& copy_variant() {
  static const CopyVariant* v = [] {
    const char* want = getenv("APS_COPY_VARIANT");
    const CopyVariant* pick = &kCopyVariants[0];
    if (want)
      for (const CopyVariant& c : kCopyVariants)
        if (!strcmp(c.name, want)) pick = &c;
    return pick;
  }();
  return *v;
}

constexpr size_t kTimelineWords = 1024 + 256 * 32 + 256;  // 256 CTAs x 4 stamps, 256 x 32 warp end stamps, 256 SM ids

struct HostCrcConsts {
  CrcTables tables;
  uint32_t xpow8[32];
  uint32_t lane_shift[32];
};
// Pure host arithmetic: the GF(2) constants the copy kernel uses.
const HostCrcConsts& host_crc_consts() {
  static const HostCrcConsts* c = [] {
    auto* h = new HostCrcConsts();
    const uint32_t xrow = host_xpow8(kRowBytes), x4 = host_xpow8(4), xrow4 = host_xpow8(4 * kRowBytes);
    for (uint32_t j = 0; j < 4; ++j)
      for (uint32_t b = 0; b < 256; ++b) {
        h->tables.fold[j][b] = gf_mul(b << (8 * j), xrow);
        h->tables.z4[j][b] = gf_mul(b << (8 * j), x4);
        h->tables.fold4[j][b] = gf_mul(b << (8 * j), xrow4);
      }
    for (int k = 0; k < 32; ++k) h->xpow8[k] = host_xpow8(1ull << k);
    // lane l holds bytes 16l..16l+15 of a 512-byte row; after the in-lane reduction its value
    // still has 512 - 16l - 12 = 500 - 16l bytes of the row behind it
    for (int l = 0; l < 32; ++l) h->lane_shift[l] = host_xpow8(500 - 16 * l);
    return h;
  }();
  return *c;
}

int ensure_device_consts(int dev, DeviceConsts** out) {
  std::lock_guard<std::mutex> lk(g_dev_mu);
  if (dev < 0 || dev >= 64) return APS_EINVAL;
  DeviceConsts& d = g_dev[dev];
  if (!d.ready) {
    const HostCrcConsts& h = host_crc_consts();
    CUDA_TRY(cudaMalloc(&d.d_tables, sizeof(CrcTables)));
    CUDA_TRY(cudaMemcpy(d.d_tables, &h.tables, sizeof(CrcTables), cudaMemcpyHostToDevice));
    {
      // the image the copy kernel pulls into shared memory with bulk-async copies: plain x^32 tables, then
      // the four x^(8*512) tables with every entry replicated for the 32 lanes (see kSmemF's layout)
      std::vector<uint8_t> img(kSmemEnd - kSmemG, 0);
      memcpy(img.data(), h.tables.z4, sizeof(h.tables.z4));
      for (uint32_t j = 0; j < 4; ++j)
        for (uint32_t b = 0; b < 256; ++b)
          for (uint32_t l = 0; l < 32; ++l) {
            const uint32_t at = (kSmemF - kSmemG) + (j >> 1) * 0x10000u + b * 256u + (j & 1u) * 128u + l * 4u;
            memcpy(img.data() + at, &h.tables.fold[j][b], 4);
          }
      CUDA_TRY(cudaMalloc(&d.d_table_image, img.size()));
      CUDA_TRY(cudaMemcpy(d.d_table_image, img.data(), img.size(), cudaMemcpyHostToDevice));
      for (uint32_t j = 0; j < 4; ++j)
        for (uint32_t b = 0; b < 256; ++b)
          for (uint32_t l = 0; l < 32; ++l) {
            const uint32_t at = (kSmemF - kSmemG) + (j >> 1) * 0x10000u + b * 256u + (j & 1u) * 128u + l * 4u;
            memcpy(img.data() + at, &h.tables.fold4[j][b], 4);
          }
      CUDA_TRY(cudaMalloc(&d.d_table_image4, img.size()));
      CUDA_TRY(cudaMemcpy(d.d_table_image4, img.data(), img.size(), cudaMemcpyHostToDevice));
    }
    CUDA_TRY(cudaMemcpyToSymbol(c_xpow8, h.xpow8, sizeof(h.xpow8)));
    CUDA_TRY(cudaMemcpyToSymbol(c_lane_shift, h.lane_shift, sizeof(h.lane_shift)));
    for (const CopyVariant& v : kCopyVariants) {
      CUDA_TRY(cudaFuncSetAttribute(v.crc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCopyDynSmem));
      CUDA_TRY(cudaFuncSetAttribute(v.get_crc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCopyDynSmem));
      CUDA_TRY(cudaFuncSetAttribute(v.put_crc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCopyDynSmem));
    }
    {
      int optin = 0;
      CUDA_TRY(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
      for (CopyKernelFn fn : {g_tma.crc, g_tma.plain, g_tma.get_crc, g_tma.get_plain}) {
        cudaFuncAttributes fa{};
        CUDA_TRY(cudaFuncGetAttributes(&fa, fn));
        const int dyn = optin - (int)fa.sharedSizeBytes;
        CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn));
        g_tma.dyn_crc = g_tma.dyn_plain = dyn;
      }
    }
    CUDA_TRY(cudaFuncSetAttribute(copy_kernel<true, 512, 4, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)kCopyDynSmem));
    CUDA_TRY(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaMalloc(&d.d_sm_slow, 256));
    CUDA_TRY(cudaMemset(d.d_sm_slow, 0, 256));
    d.ready = true;
  }
  *out = &d;
  return APS_OK;
}

uint64_t next_pow2(uint64_t v) {
  uint64_t p = 1;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace

struct aps_store {
  std::mutex mu;
  bool closed = false;
  aps_config cfg{};
  cudaStream_t stream = nullptr;
  Table table{};
  uint64_t slots = 0, n_entries = 0;
  uint32_t page_rows = 0;  // 512-byte rows in a full page
  DeviceConsts* dc = nullptr;
  // scratch, sized for cfg.max_batch requests
  Job* d_jobs = nullptr;
  uint32_t* d_acc = nullptr;
  uint32_t* d_rowshift = nullptr;  // x^(8*512*k) for k = 0 .. page_rows
  uint32_t init_full = 0;          // 0xFFFFFFFF * x^(8 * page size)
  unsigned long long* d_timeline = nullptr;  // aps_debug_timeline
  uint32_t* d_sched = nullptr;               // CTA rank counters of the copy kernel (self-resetting)
  // sharded cache: tables of all shards (own + peers mapped over NVLink), see aps_shard_attach
  Table* d_shards = nullptr;
  uint32_t n_shards = 0, my_shard = 0;
  std::vector<void*> ipc_ptrs;  // peer allocations opened through CUDA IPC (closed in aps_close)
  uint32_t* d_done = nullptr;
  unsigned long long* d_jslot = nullptr;  // fused put: per-job hand-over word (see copy_kernel, kPut)
  uint32_t put_epoch = 0;
  bool fused_put = true;  // APS_NO_FUSED_PUT=1: resolve puts in a kernel of their own, as the bulk-async variant does
  uint32_t* d_clock = nullptr;  // APS_FLAG_DEVICE_CLOCK: {hand, cursor, need, found, CTAs finished}
  int* d_error = nullptr;
  int* h_error = nullptr;  // pinned and device-visible: the copy kernels' error flag, read after a sync without a copy
  void* d_reqs = nullptr;  // staging for host-array calls (max_batch * sizeof(aps_get_req))
  aps_result* d_res = nullptr;
  uint8_t* h_stage = nullptr;   // pinned: requests, then results
  uint8_t* h_bounce = nullptr;  // pinned: one page, for pageable single-page callers
  uint8_t* d_d2h_stage = nullptr;  // HBM staging for batched gets whose destinations are in host memory (grown on demand)
  uint64_t d2h_stage_bytes = 0;
  bool d2h_stage_failed = false;
  cudaStream_t copy_stream = nullptr;  // the copy engine's stream for staged host-destination gets
  cudaEvent_t chunk_ev[8] = {};
  bool use_d2h_stage = true;       // APS_NO_D2H_STAGE=1: let the kernel store into host memory directly
  uint32_t faults = 0;
  bool use_pdl = true;    // APS_NO_PDL=1 turns programmatic dependent launch off
  bool use_steal = true;  // APS_NO_STEAL=1 turns intra-CTA work stealing off
  bool host_form = false;       // set around launches made by the host-array entry points: their predecessor on the
                                // stream is a copy, not a kernel, so there is nothing to overlap with
  bool use_early = true;        // APS_NO_EARLY=1: every kernel waits for its predecessor before it reads anything
  bool use_table_image = true;  // APS_NO_TABLE_IMAGE=1: build the shared-memory CRC tables with plain loads
  uint32_t prefetch_rows = 48;  // rows a waiting warp prefetches into L2 (APS_PF_ROWS overrides; 0/16/32/48/64/96 rows
                                // measured 0.892/0.900/0.913/0.9175/0.9165/0.908 of the HBM peak)
  bool use_tma = false;         // (retired public flag) the bulk-async form of the copy kernel
  // what the store enqueued last: a get may start reading before its predecessor has finished only if that
  // predecessor was a get too (see launch_copy)
  bool last_was_get = false;
  cudaStream_t last_stream = nullptr;
  // push transport of the sharded multi-get (aps_push_*): everything is null until aps_push_setup
  struct Push {
    uint8_t* landing = nullptr;       // peers write pages here
    uint64_t landing_bytes = 0;
    uint8_t* ctrl = nullptr;          // inboxes | done words | result slots, one allocation (exported to the peers)
    uint64_t ctrl_bytes = 0, off_done = 0, off_results = 0;
    uint32_t cap = 0, depth = 0, n_shards = 0, me = 0;
    unsigned long long timeout_ns = 30ull * 1000 * 1000 * 1000;  // APS_PUSH_TIMEOUT_MS overrides
    uint32_t sharing = 1;             // shards whose serve kernels run on this device (they split the SMs)
    bool attached = false;
    PushServe* d_serve = nullptr;     // the owner-side maps, in device memory
    uint8_t* peer_ctrl[kPushMaxShards] = {};  // peers' control blocks as mapped here
    uint32_t* d_acc = nullptr;        // the serve kernel's own scratch: it runs next to this store's other kernels
    uint32_t* d_done = nullptr;
    uint32_t* d_sched = nullptr;
    uint32_t* d_served = nullptr;
    cudaStream_t serve_stream = nullptr;
    cudaStream_t hp_stream = nullptr;  // highest priority: the post and wait kernels (see aps_get_batch_push_begin_dev)
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;
    uint32_t seq_begun = 0, seq_ended = 0;
    uint32_t n_of[8] = {};            // batch size of the rounds in flight, by slot
    std::vector<void*> ipc_ptrs;
  } push;
  std::atomic<uint64_t> launches{0};
  // optional timing of the copy kernel (aps_profile_*): event pairs recorded around each launch
  bool prof_on = false;
  std::vector<cudaEvent_t> prof_events;
  size_t prof_used = 0;
};

namespace {

inline void count_launch(aps_store* s, uint64_t k = 1) {
  g_launches += k;
  if (s) s->launches += k;
}

int launch_copy(aps_store* s, uint32_t n, const aps_get_req* d_get_reqs, aps_result* d_res, uint32_t* d_crc_out,
                cudaStream_t st, bool sharded = false, const aps_put_req* d_put_reqs = nullptr) {
  const bool crc = (s->cfg.flags & APS_FLAG_CRC) != 0 || d_crc_out != nullptr;
  const CopyVariant& v = copy_variant();
  const uint32_t warps_per_block = v.threads / 32;
  const uint32_t chip_warps = (uint32_t)s->dc->num_sms * warps_per_block;
  // cut every job into equal warp-tasks so that one round of the persistent grid covers the batch
  // ... leaving one warp-slot free per slow SM, so that the CTAs on those SMs get one task less
  const uint32_t usable = chip_warps > (uint32_t)s->dc->n_slow + n ? chip_warps - (uint32_t)s->dc->n_slow : chip_warps;
  uint32_t tpj = std::max(1u, usable / n);
  tpj = std::min(tpj, std::min(kMaxTasksPerJob, std::max(1u, s->page_rows / 8)));
  const uint32_t rpt = (s->page_rows + tpj - 1) / tpj;
  tpj = (s->page_rows + rpt - 1) / rpt;
  CopyArgs a{};
  a.jobs = s->d_jobs;
  a.reqs = d_get_reqs;
  a.preqs = d_put_reqs;
  a.jslot = s->d_jslot;
  if (d_put_reqs) {
    if (++s->put_epoch == 0) s->put_epoch = 1;
    a.epoch = s->put_epoch;
  }
  a.table = s->table;
  a.crc_on = (s->cfg.flags & APS_FLAG_CRC) != 0;
  a.n_jobs = n;
  a.tasks_per_job = tpj;
  a.rows_per_task = rpt;
  a.res = d_res;
  a.crc_out = d_crc_out;
  a.acc = s->d_acc;
  a.done = s->d_done;
  a.tables = s->dc->d_tables;
  a.table_image = s->use_table_image ? (v.group == 4 ? s->dc->d_table_image4 : s->dc->d_table_image) : nullptr;
  a.rowshift = s->d_rowshift;
  a.init_full = s->init_full;
  static const uint32_t dbg_flags = [] { const char* e = getenv("APS_COPY_FLAGS"); return e ? (uint32_t)atoi(e) : 0u; }();
  a.serve_flags = dbg_flags;  // experiments: 4 = no short cut for a job done by one warp
  a.error_flag = s->h_error;
  a.timeline = s->d_timeline;
  a.steal = s->use_steal ? 1 : 0;
  a.sched = s->d_sched;
  a.sm_slow = s->dc->d_sm_slow;
  a.shards = sharded ? s->d_shards : nullptr;
  a.n_shards = sharded ? s->n_shards : 0;
  a.n_slow = (uint32_t)s->dc->n_slow;
  // Late dependency gate (see copy_kernel): only for a get that directly follows another get of this store on
  // the same stream -- then the table and the slab cannot change under it -- and only when the caller has
  // promised that request arrays are complete before the call is enqueued (APS_FLAG_STABLE_REQS).
  const bool is_get = d_get_reqs != nullptr;
  // (A sharded get also reads its peers' tables and slabs.  Those are not ordered against this stream in any case --
  // keeping puts on one GPU apart from gets on another is the caller's job -- so reading them early changes nothing.)
  a.early = (is_get && s->last_was_get && s->last_stream == st && s->use_pdl && !s->d_timeline && s->use_early && !s->host_form)
                ? 1 + (sharded ? 0 : (int)s->prefetch_rows) : 0;  // no L2 prefetch of peer memory: measured 18x slower
  a.trust_reqs = (s->cfg.flags & APS_FLAG_STABLE_REQS) != 0;
  // a fused put only prefetches its sources before it waits, which is safe after any predecessor
  if (d_put_reqs && s->use_pdl && !s->d_timeline && s->use_early && !s->host_form) a.early = 1 + (int)s->prefetch_rows;
  s->last_was_get = is_get;
  s->last_stream = st;
  if (use_tma_variant() || s->use_tma) {
    // bulk-async form: chunks of 16 rows, a contiguous weighted slice of the chunk space per CTA
    const uint32_t cpj = (s->page_rows + kStageRows - 1) / kStageRows;
    a.tasks_per_job = cpj;
    a.rows_per_task = kStageRows;
    const uint64_t chunks = (uint64_t)n * cpj;
    uint32_t tgrid = (uint32_t)std::min<uint64_t>((uint64_t)s->dc->num_sms, chunks);
    if (tgrid == 0) tgrid = 1;
    if ((uint64_t)n / tgrid + 2 > kJobCache) return APS_EINVAL;
    CopyKernelFn fn = d_get_reqs ? (crc ? g_tma.get_crc : g_tma.get_plain) : (crc ? g_tma.crc : g_tma.plain);
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(tgrid);
    lc.blockDim = dim3(kTmaThreads);
    lc.dynamicSmemBytes = (size_t)g_tma.dyn_crc;
    lc.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = s->use_pdl ? 1 : 0;
    lc.attrs = attr;
    lc.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&lc, fn, a));
    count_launch(s);
    return APS_OK;
  }
  if (s->d_timeline) CUDA_TRY(cudaMemsetAsync(s->d_timeline, 0xFF, kTimelineWords * sizeof(unsigned long long), st));
  const uint64_t items = (uint64_t)n * tpj;
  uint32_t grid = (uint32_t)std::min<uint64_t>((uint64_t)s->dc->num_sms, items);  // every SM takes part
  // A call for a page or two (the one-page-per-call PageStore interface) is latency, not bandwidth: four tasks per
  // CTA instead of one means a quarter of the CTAs fetch the 132 KB table image (64 KiB get: 25.4 -> 22.7 us).
  if (items <= 256) grid = (uint32_t)((items + 3) / 4);
  if (grid == 0) grid = 1;
  const bool timed = s->prof_on && s->prof_used + 2 <= s->prof_events.size();
  if (timed) CUDA_TRY(cudaEventRecord(s->prof_events[s->prof_used], st));
  {
    // Programmatic dependent launch: back-to-back copy kernels overlap launch latency; each kernel
    // still waits (griddepcontrol.wait) for all earlier work before it touches any data.
    CopyKernelFn fn = d_get_reqs ? (crc ? v.get_crc : v.get_plain)
                      : d_put_reqs ? (crc ? v.put_crc : v.put_plain) : (crc ? v.crc : v.plain);
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(v.threads);
    lc.dynamicSmemBytes = crc ? kCopyDynSmem : 0;
    lc.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = s->use_pdl ? 1 : 0;
    lc.attrs = attr;
    lc.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&lc, fn, a));
  }
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  if (timed) {
    CUDA_TRY(cudaEventRecord(s->prof_events[s->prof_used + 1], st));
    s->prof_used += 2;
  }
  return APS_OK;
}

inline dim3 tile_grid(uint32_t n) { return dim3((n * 8 + 255) / 256); }

// CLOCK sweep: evicts `fixed` pages, or (fixed == 0) as many as `incoming` new pages lack free slots.
int clock_evict(aps_store* s, uint32_t incoming, uint32_t fixed, aps_key* d_victims, uint32_t victims_cap, cudaStream_t st) {
  clock_need_kernel<<<1, 1, 0, st>>>(s->table, incoming, fixed, s->d_clock);
  // never more windows in flight than half a turn of the hand, so that a page keeps its second chance
  const unsigned grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(s->slots / 512, (uint64_t)s->dc->num_sms * 8));
  clock_evict_kernel<<<grid, 256, 0, st>>>(s->table, (uint32_t)s->slots, s->d_clock, d_victims, victims_cap);
  count_launch(s, 2);
  CUDA_TRY(cudaGetLastError());
  s->last_was_get = false;
  return APS_OK;
}

int check_open(aps_store* s) {
  if (!s) return APS_EINVAL;
  if (s->closed) return APS_ECLOSED;
  return APS_OK;
}

struct KeyHash {
  size_t operator()(const aps_key& k) const { return (size_t)hash_key(k); }
};
struct KeyEq {
  bool operator()(const aps_key& a, const aps_key& b) const {
    return a.file_hi == b.file_hi && a.file_lo == b.file_lo && a.page_index == b.page_index;
  }
};

template <class GetKey>
bool distinct_keys(uint32_t n, GetKey key_of) {
  std::unordered_set<aps_key, KeyHash, KeyEq> seen;
  seen.reserve(n * 2);
  for (uint32_t i = 0; i < n; ++i)
    if (!seen.insert(key_of(i)).second) return false;
  return true;
}

int check_device_error(aps_store* s, cudaStream_t st) {
  CUDA_TRY(cudaStreamSynchronize(st));
  const int err = *(volatile int*)s->h_error;
  if (err) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "copy kernel: shared-memory window not at the expected address");
    return APS_ECUDA;
  }
  return APS_OK;
}

void free_store(aps_store* s) {
  cudaFree(s->table.slab);
  cudaFree(s->table.entries);
  cudaFree(s->table.free_slots);
  cudaFree(s->table.free_top);
  cudaFree(s->table.slot_crc);
  cudaFree(s->table.counters);
  cudaFree(s->d_jobs);
  cudaFree(s->d_acc);
  cudaFree(s->d_rowshift);
  cudaFree(s->d_timeline);
  cudaFree(s->d_sched);
  cudaFree(s->d_shards);
  for (void* p : s->ipc_ptrs) cudaIpcCloseMemHandle(p);
  s->ipc_ptrs.clear();
  cudaFree(s->d_done);
  if (s->push.serve_stream) {
    cudaStreamSynchronize(s->push.serve_stream);
    cudaStreamDestroy(s->push.serve_stream);
  }
  if (s->push.hp_stream) {
    cudaStreamSynchronize(s->push.hp_stream);
    cudaStreamDestroy(s->push.hp_stream);
  }
  if (s->push.ev_in) cudaEventDestroy(s->push.ev_in);
  if (s->push.ev_out) cudaEventDestroy(s->push.ev_out);
  for (void* p : s->push.ipc_ptrs) cudaIpcCloseMemHandle(p);
  s->push.ipc_ptrs.clear();
  cudaFree(s->push.landing);
  cudaFree(s->push.ctrl);
  cudaFree(s->push.d_serve);
  cudaFree(s->push.d_acc);
  cudaFree(s->push.d_done);
  cudaFree(s->push.d_sched);
  cudaFree(s->push.d_served);
  cudaFree(s->d_jslot);
  cudaFree(s->d_clock);
  cudaFree(s->d_d2h_stage);
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  for (cudaEvent_t e : s->chunk_ev)
    if (e) cudaEventDestroy(e);
  cudaFree(s->table.slot_ref);
  cudaFree(s->table.slot_entry);
  cudaFree(s->d_error);
  if (s->h_error) cudaFreeHost(s->h_error);
  cudaFree(s->d_reqs);
  cudaFree(s->d_res);
  for (cudaEvent_t e : s->prof_events) cudaEventDestroy(e);
  s->prof_events.clear();
  if (s->h_stage) cudaFreeHost(s->h_stage);
  if (s->h_bounce) cudaFreeHost(s->h_bounce);
  if (s->stream) cudaStreamDestroy(s->stream);
}

// SM speed classes.  On a saturated memory system equal streams do not finish together: some SMs
// (larger GPCs share one path to L2) are consistently a few percent slower.  Measured once per
// device: copy a dummy batch inside the still empty slab a few times with per-CTA time stamps, and
// mark the SMs whose CTAs end later than the chip median by more than 1.5 % of the kernel time (median of seven runs).  The copy
// kernel then deals the surplus tasks to the other SMs (see copy_kernel's rank logic).
int calibrate_sms(aps_store* s) {
  DeviceConsts* dc = s->dc;
  if (dc->calibrated || getenv("APS_NO_SMWEIGHT")) return APS_OK;
  const uint64_t n64 = std::min<uint64_t>(std::min<uint64_t>(s->cfg.max_batch, 256), s->slots / 2);
  if (n64 * s->cfg.page_size < (64ull << 20)) return APS_OK;  // too small to say anything; all SMs equal
  const uint32_t n = (uint32_t)n64;
  std::vector<Job> jobs(n);
  for (uint32_t i = 0; i < n; ++i)
    jobs[i] = Job{s->table.slab + (uint64_t)i * s->table.page_size, s->table.slab + (uint64_t)(n + i) * s->table.page_size,
                  s->table.page_size, 0, 0, 0};
  CUDA_TRY(cudaMemcpy(s->d_jobs, jobs.data(), n * sizeof(Job), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMalloc(&s->d_timeline, kTimelineWords * sizeof(unsigned long long)));
  std::vector<unsigned long long> tl(kTimelineWords);
  std::vector<std::vector<double>> offsets(256);
  double dur_sum = 0;
  const int kWarm = 2, kRuns = 7;
  int rc = APS_OK;
  for (int run = 0; run < kWarm + kRuns && rc == APS_OK; ++run) {
    rc = launch_copy(s, n, nullptr, nullptr, nullptr, s->stream);
    if (rc != APS_OK) break;
    if (cudaStreamSynchronize(s->stream) != cudaSuccess ||
        cudaMemcpy(tl.data(), s->d_timeline, kTimelineWords * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) {
      rc = APS_ECUDA;
      break;
    }
    if (run < kWarm) continue;
    std::vector<std::pair<unsigned, double>> ends;
    unsigned long long t0 = ~0ull;
    for (int c = 0; c < 256; ++c)
      if (tl[c * 4] != ~0ull) t0 = std::min(t0, tl[c * 4]);
    for (int c = 0; c < 256; ++c)
      if (tl[c * 4] != ~0ull && tl[c * 4 + 3] != ~0ull)
        ends.push_back({(unsigned)(tl[1024 + 256 * 32 + c] & 255u), (double)(tl[c * 4 + 3] - t0)});
    if (ends.size() < 8) continue;
    std::vector<double> v;
    for (auto& e : ends) v.push_back(e.second);
    std::nth_element(v.begin(), v.begin() + v.size() / 2, v.end());
    const double med = v[v.size() / 2];
    dur_sum += med;
    for (auto& e : ends) offsets[e.first].push_back(e.second - med);
  }
  cudaFree(s->d_timeline);
  s->d_timeline = nullptr;
  if (rc != APS_OK) return rc;
  const double dur = dur_sum / kRuns;
  std::vector<uint8_t> slow(256, 0);
  int n_slow = 0;
  for (int i = 0; i < 256; ++i) {
    std::vector<double>& o = offsets[i];
    if ((int)o.size() < kRuns - 1) continue;  // SM not (always) part of the launch
    std::nth_element(o.begin(), o.begin() + o.size() / 2, o.end());
    // the two classes are well separated (fast: within +-0.5 %, slow: +2.5 % .. +6 % of the kernel time):
    // the median over the runs against a 1.5 % threshold
    if (o[o.size() / 2] > 0.015 * dur) {
      slow[i] = 1;
      ++n_slow;
    }
  }
  if (n_slow > dc->num_sms / 2) {  // not a few stragglers but a bimodal chip: leave the plain split
    std::fill(slow.begin(), slow.end(), 0);
    n_slow = 0;
  }
  CUDA_TRY(cudaMemcpy(dc->d_sm_slow, slow.data(), 256, cudaMemcpyHostToDevice));
  dc->n_slow = n_slow;
  dc->calibrated = true;
  return APS_OK;
}

// The MD5 kernel stages 2 x 32 x 2 KiB in dynamic shared memory: opt in once per device.
int md5_prepare(int device) {
  static std::mutex mu;
  static bool done[64] = {};
  std::lock_guard<std::mutex> lk(mu);
  if (device < 0 || device >= 64) return APS_EINVAL;
  if (!done[device]) {
    CUDA_TRY(cudaFuncSetAttribute(md5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMd5DynSmem));
    done[device] = true;
  }
  return APS_OK;
}

__global__ void corrupt_kernel(Table t, aps_key key, uint32_t off, int* found) {
  const uint32_t sub = threadIdx.x & 7u;
  Probe pr = tile_lookup(t, key, sub, 0);
  if (sub == 0) {
    *found = pr.meta != 0;
    if (pr.meta && off < meta_len(pr.meta)) t.slab[(uint64_t)meta_slot(pr.meta) * t.page_size + off] ^= 0x5Au;
  }
}

}  // namespace

namespace aps {
int store_config(aps_store* s, aps_config* out) {
  int rc = check_open(s);
  if (rc) return rc;
  *out = s->cfg;
  return APS_OK;
}
}  // namespace aps

extern "C" {

const char* aps_version(void) { return "libaps 0.1 (sm_100a)"; }

int aps_debug_crc_constants(uint32_t* fold, uint32_t* z4, uint32_t* xpow8_out, uint32_t* lane_shift) {
  const HostCrcConsts& h = host_crc_consts();
  if (fold) memcpy(fold, h.tables.fold, sizeof(h.tables.fold));
  if (z4) memcpy(z4, h.tables.z4, sizeof(h.tables.z4));
  if (xpow8_out) memcpy(xpow8_out, h.xpow8, sizeof(h.xpow8));
  if (lane_shift) memcpy(lane_shift, h.lane_shift, sizeof(h.lane_shift));
  return APS_OK;
}
int aps_debug_crc_fold4(uint32_t* fold4) {
  if (!fold4) return APS_EINVAL;
  memcpy(fold4, host_crc_consts().tables.fold4, sizeof(host_crc_consts().tables.fold4));
  return APS_OK;
}
const char* aps_last_cuda_error(void) { return tl_cuda_error; }
uint64_t aps_total_kernel_launches(void) { return g_launches.load(); }

const char* aps_strerror(int code) {
  switch (code) {
    case APS_OK: return "ok";
    case APS_ENOTFOUND: return "page not found";
    case APS_EINVAL: return "invalid argument";
    case APS_ENOSPC: return "no free slab slot or table entry";
    case APS_ECUDA: return "CUDA call failed";
    case APS_ECORRUPT: return "page CRC32C mismatch";
    case APS_ECLOSED: return "store is closed";
    case APS_ENOMEM: return "out of memory";
    case APS_EFAULT: return "injected fault";
    case APS_ELIMIT: return "request exceeds the fixed scratch of the no-allocation entry point";
    case APS_EIO: return "snapshot file could not be read or written";
    case APS_ETIMEOUT: return "store call timed out or all store threads busy";
    default: return "unknown error";
  }
}

int aps_open(const aps_config* cfg, aps_store** out) {
  if (!cfg || !out) return APS_EINVAL;
  *out = nullptr;
  // any page size the reference accepts (alluxio.user.client.cache.page.size is a free byte count, PropertyKey.java:3704);
  // slab slots are spaced at the next multiple of 16 so that whole-page copies stay on the 128-bit path
  if (cfg->page_size == 0 || cfg->page_size > (1u << 30)) return APS_EINVAL;
  const uint32_t stride = (cfg->page_size + 15u) & ~15u;
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (cfg->device < 0 || cfg->device >= ndev) return APS_EINVAL;
  DeviceGuard g(cfg->device);
  if (!g.ok) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "cudaSetDevice(%d) failed", cfg->device);
    return APS_ECUDA;
  }
  DeviceConsts* dc = nullptr;
  int rc = ensure_device_consts(cfg->device, &dc);
  if (rc != APS_OK) return rc;

  aps_store* s = new (std::nothrow) aps_store();
  if (!s) return APS_ENOMEM;
  s->cfg = *cfg;
  s->dc = dc;
  s->use_pdl = getenv("APS_NO_PDL") == nullptr;
  s->use_steal = getenv("APS_NO_STEAL") == nullptr;
  s->use_table_image = getenv("APS_NO_TABLE_IMAGE") == nullptr;
  s->use_d2h_stage = getenv("APS_NO_D2H_STAGE") == nullptr;
  s->use_early = getenv("APS_NO_EARLY") == nullptr;
  s->fused_put = getenv("APS_NO_FUSED_PUT") == nullptr;
  if (const char* pf = getenv("APS_PF_ROWS")) s->prefetch_rows = (uint32_t)std::min(256, std::max(0, atoi(pf)));
  s->use_tma = false;  // the public flag is retired (aps.h); APS_COPY_VARIANT=tma still selects the bulk-async kernel
  uint64_t slots = cfg->slab_slots;
  if (slots == 0) {
    uint64_t base = (cfg->cache_bytes + cfg->page_size - 1) / cfg->page_size;
    slots = base + std::max<uint64_t>(64, base / 8);
  }
  if (slots >= (1ull << 31)) { delete s; return APS_EINVAL; }
  uint64_t entries = cfg->table_capacity ? next_pow2(cfg->table_capacity) : std::max<uint64_t>(1024, next_pow2(2 * slots));
  if (entries < kBucketEntries) entries = kBucketEntries;
  if (s->cfg.max_batch == 0) s->cfg.max_batch = 4096;
  s->slots = slots;
  s->n_entries = entries;
  s->page_rows = (stride + kRowBytes - 1) / kRowBytes;
  s->init_full = gf_mul(0xFFFFFFFFu, host_xpow8(cfg->page_size));
  const uint32_t mb = s->cfg.max_batch;

#define OPEN_TRY(expr)                                                                                    \
  do {                                                                                                    \
    cudaError_t _e = (expr);                                                                              \
    if (_e != cudaSuccess) {                                                                              \
      snprintf(tl_cuda_error, sizeof(tl_cuda_error), "%s:%d %s -> %s", __FILE__, __LINE__, #expr,         \
               cudaGetErrorString(_e));                                                                   \
      free_store(s);                                                                                      \
      delete s;                                                                                           \
      return _e == cudaErrorMemoryAllocation ? APS_ENOMEM : APS_ECUDA;                                    \
    }                                                                                                     \
  } while (0)

  OPEN_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  OPEN_TRY(cudaMalloc(&s->table.slab, slots * (uint64_t)stride));
  OPEN_TRY(cudaMalloc(&s->table.entries, entries * sizeof(Entry)));
  OPEN_TRY(cudaMalloc(&s->table.free_slots, slots * sizeof(uint32_t)));
  OPEN_TRY(cudaMalloc(&s->table.free_top, sizeof(int32_t)));
  OPEN_TRY(cudaMalloc(&s->table.slot_crc, slots * sizeof(uint32_t)));
  OPEN_TRY(cudaMalloc(&s->table.counters, kNumCounters * sizeof(unsigned long long)));
  OPEN_TRY(cudaMalloc(&s->d_jobs, (uint64_t)mb * sizeof(Job)));
  OPEN_TRY(cudaMalloc(&s->d_acc, (uint64_t)mb * sizeof(uint32_t)));
  OPEN_TRY(cudaMalloc(&s->d_sched, 4 * sizeof(uint32_t)));
  OPEN_TRY(cudaMemset(s->d_sched, 0, 4 * sizeof(uint32_t)));
  {
    std::vector<uint32_t> rs(s->page_rows + 1);
    const uint32_t xrow = host_xpow8(kRowBytes);
    rs[0] = 0x80000000u;
    for (uint32_t k = 1; k <= s->page_rows; ++k) rs[k] = gf_mul(rs[k - 1], xrow);
    OPEN_TRY(cudaMalloc(&s->d_rowshift, rs.size() * sizeof(uint32_t)));
    OPEN_TRY(cudaMemcpy(s->d_rowshift, rs.data(), rs.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
  OPEN_TRY(cudaMalloc(&s->d_done, (uint64_t)mb * sizeof(uint32_t)));
  OPEN_TRY(cudaMalloc(&s->d_jslot, (uint64_t)mb * sizeof(unsigned long long)));
  OPEN_TRY(cudaMemset(s->d_jslot, 0, (uint64_t)mb * sizeof(unsigned long long)));
  OPEN_TRY(cudaMalloc(&s->d_error, sizeof(int)));
  OPEN_TRY(cudaHostAlloc(&s->h_error, sizeof(int), cudaHostAllocDefault));
  *s->h_error = 0;
  if (cfg->flags & APS_FLAG_DEVICE_CLOCK) {
    OPEN_TRY(cudaMalloc(&s->table.slot_ref, slots));
    OPEN_TRY(cudaMalloc(&s->table.slot_entry, slots * sizeof(uint64_t)));
    OPEN_TRY(cudaMalloc(&s->d_clock, 8 * sizeof(uint32_t)));
    OPEN_TRY(cudaMemset(s->table.slot_ref, 0, slots));
    OPEN_TRY(cudaMemset(s->table.slot_entry, 0xFF, slots * sizeof(uint64_t)));
    OPEN_TRY(cudaMemset(s->d_clock, 0, 8 * sizeof(uint32_t)));
  }
  OPEN_TRY(cudaMalloc(&s->d_reqs, (uint64_t)mb * sizeof(aps_get_req)));
  OPEN_TRY(cudaMalloc(&s->d_res, (uint64_t)mb * sizeof(aps_result)));
  OPEN_TRY(cudaHostAlloc(&s->h_stage, (uint64_t)mb * (sizeof(aps_get_req) + sizeof(aps_result)), cudaHostAllocDefault));
  OPEN_TRY(cudaHostAlloc(&s->h_bounce, stride, cudaHostAllocDefault));
  s->table.bucket_mask = entries / kBucketEntries - 1;
  s->table.page_size = stride;
  s->table.max_len = cfg->page_size;
  const int32_t top = (int32_t)slots;
  OPEN_TRY(cudaMemsetAsync(s->table.entries, 0, entries * sizeof(Entry), s->stream));
  OPEN_TRY(cudaMemsetAsync(s->table.slot_crc, 0, slots * sizeof(uint32_t), s->stream));
  OPEN_TRY(cudaMemsetAsync(s->table.counters, 0, kNumCounters * sizeof(unsigned long long), s->stream));
  OPEN_TRY(cudaMemsetAsync(s->d_done, 0, (uint64_t)mb * sizeof(uint32_t), s->stream));
  OPEN_TRY(cudaMemsetAsync(s->d_acc, 0, (uint64_t)mb * sizeof(uint32_t), s->stream));
  OPEN_TRY(cudaMemsetAsync(s->d_error, 0, sizeof(int), s->stream));
  OPEN_TRY(cudaMemcpyAsync(s->table.free_top, &top, sizeof(top), cudaMemcpyHostToDevice, s->stream));
  init_free_slots_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, s->stream>>>(s->table.free_slots, slots);
  count_launch(s);
  OPEN_TRY(cudaGetLastError());
  OPEN_TRY(cudaStreamSynchronize(s->stream));
#undef OPEN_TRY
  {
    std::lock_guard<std::mutex> lk(g_dev_mu);
    int crc_rc = calibrate_sms(s);
    if (crc_rc != APS_OK) {
      free_store(s);
      delete s;
      return crc_rc;
    }
  }
  *out = s;
  return APS_OK;
}

// PageStore is AutoCloseable and SURVEY 8b asks for an idempotent close.  The handle itself (a few hundred bytes of host
// memory, no device resources) is therefore never freed: after the first close it stays behind as a tombstone, so a
// second close returns APS_OK and any other call on it returns APS_ECLOSED instead of reading freed memory.
int aps_close(aps_store* s) {
  if (!s) return APS_EINVAL;
  std::lock_guard<std::mutex> lk(s->mu);
  if (s->closed) return APS_OK;
  s->closed = true;
  DeviceGuard g(s->cfg.device);
  cudaStreamSynchronize(s->stream);
  free_store(s);
  return APS_OK;
}

uint64_t aps_cache_size(const aps_store* s) { return s ? s->cfg.cache_bytes : 0; }

int aps_set_faults(aps_store* s, uint32_t fault_mask) {
  int rc = check_open(s);
  if (rc) return rc;
  s->faults = fault_mask;
  return APS_OK;
}

// ---------------------------------------------------------------- device-resident batched forms

int aps_get_batch_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !d_reqs || !d_res) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  return launch_copy(s, n, d_reqs, d_res, nullptr, st);  // lookup + gather + CRC in one launch
}

int aps_put_batch_dev(aps_store* s, uint32_t n, const aps_put_req* d_reqs, aps_result* d_res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !d_reqs || !d_res) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  s->last_was_get = false;
  if (s->d_clock) {  // make room on the device: evict as many pages as this batch lacks free slots
    rc = clock_evict(s, n, 0, nullptr, 0, st);
    if (rc) return rc;
  }
  // insert + scatter + CRC in one launch -- up to a few thousand pages.  The fused insert updates the store's counters
  // and free-slot stack once per page; a batch of a million tiny pages is better served by the resolve kernel, which
  // does that once per block (1.4 G against 0.47 G inserts/s).
  if (s->fused_put && n <= 8192 && !use_tma_variant() && !s->use_tma)
    return launch_copy(s, n, nullptr, d_res, nullptr, st, false, d_reqs);
  put_resolve_kernel<<<tile_grid(n), kK1Threads, 0, st>>>(s->table, d_reqs, n, s->d_jobs, d_res, (s->cfg.flags & APS_FLAG_CRC) != 0);
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  return launch_copy(s, n, nullptr, d_res, nullptr, st);
}

int aps_delete_batch_dev(aps_store* s, uint32_t n, const aps_key* d_keys, aps_result* d_res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (!d_keys || !d_res) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  s->last_was_get = false;
  delete_kernel<<<tile_grid(n), kK1Threads, 0, st>>>(s->table, d_keys, n, d_res);
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  return APS_OK;
}

int aps_lookup_batch_dev(aps_store* s, uint32_t n, const aps_key* d_keys, int32_t* d_len, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (!d_keys || !d_len) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  {
    static const uint32_t per_group = [] { const char* e = getenv("APS_LOOKUP_PER_GROUP"); return e ? (uint32_t)std::max(1, atoi(e)) : kLookupPerGroup; }();
    static const bool wide = [] { const char* e = getenv("APS_LOOKUP_256"); return e ? atoi(e) != 0 : true; }();
    const uint64_t groups = ((uint64_t)n + per_group - 1) / per_group;  // keys per group of four lanes
    const unsigned grid = (unsigned)std::max<uint64_t>(1, (groups * kLookupLanes + kK1Threads - 1) / kK1Threads);
    if (wide) lookup_kernel<true><<<grid, kK1Threads, 0, st>>>(s->table, d_keys, n, d_len);
    else lookup_kernel<false><<<grid, kK1Threads, 0, st>>>(s->table, d_keys, n, d_len);
  }
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  return APS_OK;
}

int aps_crc32c_batch_dev(aps_store* s, uint32_t n, const void* const* d_data, const uint32_t* d_len, uint32_t* d_crc,
                         aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !d_data || !d_len || !d_crc) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  crc_jobs_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_data, d_len, n, s->cfg.page_size, s->d_jobs, d_crc);
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  return launch_copy(s, n, nullptr, nullptr, d_crc, st);
}

// ---------------------------------------------------------------- host-array batched forms

static int get_batch_host(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream, bool sharded);

int aps_get_batch(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream) {
  return get_batch_host(s, n, reqs, res, stream, false);
}
int aps_get_batch_sharded(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream) {
  if (s && s->n_shards == 0) return APS_EINVAL;
  return get_batch_host(s, n, reqs, res, stream, true);
}

// What kind of memory is `p`?  Asked on every call: a memo keyed by address would go stale when the caller frees a
// pinned buffer and the allocator hands the same address out again as pageable memory (a kernel would then write to
// memory it cannot reach); the driver call costs about a microsecond of a call that moves megabytes.
enum PtrKind { kPtrDevice = 0, kPtrPinned = 1, kPtrPageable = 2 };  // device / managed; page-locked host; plain host
static int host_ptr_kind(aps_store*, const void* p) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) cudaGetLastError();
  return at.type == cudaMemoryTypeHost ? kPtrPinned
         : (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) ? kPtrDevice : kPtrPageable;
}

// The HBM staging area of the host-array batched calls (one slot per request of the largest batch one call serves),
// its copy stream and chunk events: allocated on first use and kept.
static bool ensure_stage(aps_store* s, uint64_t bytes) {
  if (s->d_d2h_stage) return true;
  if (s->d2h_stage_failed) return false;
  if (cudaMalloc(&s->d_d2h_stage, bytes) == cudaSuccess && cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking) == cudaSuccess) {
    s->d2h_stage_bytes = bytes;
    for (cudaEvent_t& e : s->chunk_ev)
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) s->d2h_stage_failed = true;
  } else {
    cudaGetLastError();
    s->d2h_stage_failed = true;  // no room for it; do not ask again
  }
  return s->d_d2h_stage && !s->d2h_stage_failed;
}

static int get_batch_host(aps_store* s, uint32_t n, const aps_get_req* reqs, aps_result* res, aps_stream stream, bool sharded) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (!reqs || !res) return APS_EINVAL;
  const uint64_t slot = ((uint64_t)s->table.page_size + 255u) & ~255ull;
  // one call serves at most max_batch requests and at most 1 GiB of staging slots; more is served in pieces
  const uint32_t limit = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(s->cfg.max_batch, (1ull << 30) / slot));
  if (n > limit) {
    for (uint32_t lo = 0; lo < n; lo += limit) {
      rc = get_batch_host(s, std::min(limit, n - lo), reqs + lo, res + lo, stream, sharded);
      if (rc) return rc;
    }
    return APS_OK;
  }
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  aps_get_req* h_reqs = reinterpret_cast<aps_get_req*>(s->h_stage);
  aps_result* h_res = reinterpret_cast<aps_result*>(s->h_stage + (size_t)s->cfg.max_batch * sizeof(aps_get_req));
  memcpy(h_reqs, reqs, (size_t)n * sizeof(aps_get_req));
  // Destinations in pinned HOST memory: stores issued by SMs cross PCIe at ~0.91 of what the copy engines reach
  // (52 against 57 GB/s measured), so large batches are gathered into an HBM staging area at HBM speed and handed
  // to the copy engine, which also coalesces neighbouring destinations into one transfer.  The staging area is
  // allocated once (for the largest batch one call serves) and kept.
  const uint64_t stage_want = (uint64_t)limit * slot;
  // Plain (pageable) host destinations -- a Java direct ByteBuffer from allocateDirect, a malloc'd array -- cannot be
  // written by a kernel at all: they always go through the staging area and cudaMemcpy.  All destinations of one call
  // must be of one kind.
  const int kind = host_ptr_kind(s, reqs[0].dst);
  bool via_stage = kind == kPtrPageable || (kind == kPtrPinned && s->use_d2h_stage && (uint64_t)n * s->cfg.page_size >= (8ull << 20));
  if (via_stage && !ensure_stage(s, stage_want)) {
    if (kind == kPtrPageable) return APS_ENOMEM;
    via_stage = false;  // pinned: store straight into host memory
  }
  if (!via_stage) {
    CUDA_TRY(cudaMemcpyAsync(s->d_reqs, h_reqs, (size_t)n * sizeof(aps_get_req), cudaMemcpyHostToDevice, st));
    s->host_form = true;
    rc = sharded ? aps_get_batch_sharded_dev(s, n, (const aps_get_req*)s->d_reqs, s->d_res, st)
                 : aps_get_batch_dev(s, n, (const aps_get_req*)s->d_reqs, s->d_res, st);
    s->host_form = false;
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(h_res, s->d_res, (size_t)n * sizeof(aps_result), cudaMemcpyDeviceToHost, st));
    rc = check_device_error(s, st);  // also waits for the stream
    if (rc) return rc;
    memcpy(res, h_res, (size_t)n * sizeof(aps_result));
    return APS_OK;
  }
  // Pipelined in two pieces: a small head (~8 MiB) whose only job is to get the link busy early, and the rest.  Both
  // gather kernels are enqueued at once and read their requests from, and write their results to, the pinned staging
  // area in place (no request or result copies).  The host follows them: as soon as a piece's results are in, its pages
  // are handed to the copy engine on a second stream; the rest is gathered (~90 us for 248 MiB) while the head crosses
  // PCIe (~150 us), so the link idles only while the head is gathered.  More pieces were measured and lose: every
  // separate device-to-host copy costs a start-up gap (8 pieces of 32 MiB: 0.957 of the link, one piece: 0.97).
  for (uint32_t i = 0; i < n; ++i) h_reqs[i].dst = s->d_d2h_stage + (uint64_t)i * slot;
  const uint32_t head = std::min<uint32_t>(n, std::max<uint32_t>(1, (uint32_t)((8ull << 20) / slot)));
  const uint32_t cut[3] = {0, head, n};
  const uint32_t n_chunks = head < n ? 2 : 1;
  s->host_form = true;
  for (uint32_t c = 0; c < n_chunks && rc == APS_OK; ++c) {
    const uint32_t lo = cut[c], cnt = cut[c + 1] - lo;
    rc = sharded ? aps_get_batch_sharded_dev(s, cnt, h_reqs + lo, h_res + lo, st) : aps_get_batch_dev(s, cnt, h_reqs + lo, h_res + lo, st);
    if (rc == APS_OK && cudaEventRecord(s->chunk_ev[c], st) != cudaSuccess) rc = APS_ECUDA;
  }
  s->host_form = false;
  if (rc) {
    cudaStreamSynchronize(st);
    return rc;
  }
  for (uint32_t c = 0; c < n_chunks; ++c) {
    const uint32_t lo = cut[c], hi = cut[c + 1];
    CUDA_TRY(cudaEventSynchronize(s->chunk_ev[c]));
    // exactly the bytes each get returned, runs of requests that are adjacent on both sides as one copy
    uint32_t i = lo;
    while (i < hi) {
      if (h_res[i].len <= 0) { ++i; continue; }
      uint64_t bytes = (uint64_t)h_res[i].len;
      uint32_t j = i + 1;
      while (j < hi && bytes == (uint64_t)(j - i) * slot && h_res[j].len > 0 &&
             (uint8_t*)reqs[j].dst == (uint8_t*)reqs[i].dst + bytes)
        bytes += (uint64_t)h_res[j++].len;
      CUDA_TRY(cudaMemcpyAsync(reqs[i].dst, s->d_d2h_stage + (uint64_t)i * slot, bytes, cudaMemcpyDefault, s->copy_stream));
      i = j;
    }
  }
  CUDA_TRY(cudaStreamSynchronize(s->copy_stream));
  if (*(volatile int*)s->h_error) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "copy kernel: shared-memory window not at the expected address");
    return APS_ECUDA;
  }
  memcpy(res, h_res, (size_t)n * sizeof(aps_result));
  return APS_OK;
}

int aps_put_batch(aps_store* s, uint32_t n, const aps_put_req* reqs, aps_result* res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !reqs || !res) return APS_EINVAL;
  if (!distinct_keys(n, [&](uint32_t i) { return reqs[i].key; })) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  // Sources in plain (pageable) host memory -- a Java direct ByteBuffer, a malloc'd array -- cannot be read by a kernel:
  // they are copied into the HBM staging area first.  All sources of one call must be of one kind.
  const uint64_t slot = ((uint64_t)s->table.page_size + 255u) & ~255ull;
  const uint32_t limit = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(s->cfg.max_batch, (1ull << 30) / slot));
  const void* first_src = nullptr;
  for (uint32_t i = 0; i < n && !first_src; ++i)
    if (reqs[i].len) first_src = reqs[i].src;
  bool pageable;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    pageable = first_src && host_ptr_kind(s, first_src) == kPtrPageable;
  }
  if (pageable && n > limit) {  // more than the staging area holds: in pieces
    for (uint32_t lo = 0; lo < n; lo += limit) {
      rc = aps_put_batch(s, std::min(limit, n - lo), reqs + lo, res + lo, stream);
      if (rc) return rc;
    }
    return APS_OK;
  }
  std::lock_guard<std::mutex> lk(s->mu);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  aps_put_req* h_reqs = reinterpret_cast<aps_put_req*>(s->h_stage);
  aps_result* h_res = reinterpret_cast<aps_result*>(s->h_stage + (size_t)s->cfg.max_batch * sizeof(aps_get_req));
  memcpy(h_reqs, reqs, (size_t)n * sizeof(aps_put_req));
  if (pageable) {
    if (!ensure_stage(s, (uint64_t)limit * slot)) return APS_ENOMEM;
    for (uint32_t i = 0; i < n; ++i) {
      if (reqs[i].len > s->cfg.page_size) continue;  // refused by the kernel; nothing to stage
      if (reqs[i].len)
        CUDA_TRY(cudaMemcpyAsync(s->d_d2h_stage + (uint64_t)i * slot, reqs[i].src, reqs[i].len, cudaMemcpyHostToDevice, st));
      h_reqs[i].src = s->d_d2h_stage + (uint64_t)i * slot;
    }
  }
  CUDA_TRY(cudaMemcpyAsync(s->d_reqs, h_reqs, (size_t)n * sizeof(aps_put_req), cudaMemcpyHostToDevice, st));
  s->host_form = true;
  rc = aps_put_batch_dev(s, n, (const aps_put_req*)s->d_reqs, s->d_res, st);
  s->host_form = false;
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(h_res, s->d_res, (size_t)n * sizeof(aps_result), cudaMemcpyDeviceToHost, st));
  rc = check_device_error(s, st);
  if (rc) return rc;
  memcpy(res, h_res, (size_t)n * sizeof(aps_result));
  return APS_OK;
}

int aps_delete_batch(aps_store* s, uint32_t n, const aps_key* keys, aps_result* res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !keys || !res) return APS_EINVAL;
  if (!distinct_keys(n, [&](uint32_t i) { return keys[i]; })) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  aps_key* h_keys = reinterpret_cast<aps_key*>(s->h_stage);
  aps_result* h_res = reinterpret_cast<aps_result*>(s->h_stage + (size_t)s->cfg.max_batch * sizeof(aps_get_req));
  memcpy(h_keys, keys, (size_t)n * sizeof(aps_key));
  if (n <= 8) {  // a handful of ids (PageStore.delete is one): let the kernel use the pinned staging area in place
    rc = aps_delete_batch_dev(s, n, h_keys, h_res, st);
    if (rc) return rc;
  } else {
    CUDA_TRY(cudaMemcpyAsync(s->d_reqs, h_keys, (size_t)n * sizeof(aps_key), cudaMemcpyHostToDevice, st));
    rc = aps_delete_batch_dev(s, n, (const aps_key*)s->d_reqs, s->d_res, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(h_res, s->d_res, (size_t)n * sizeof(aps_result), cudaMemcpyDeviceToHost, st));
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  memcpy(res, h_res, (size_t)n * sizeof(aps_result));
  return APS_OK;
}

// ---------------------------------------------------------------- single-page PageStore calls

static bool is_device_accessible(const void* p) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

int aps_put(aps_store* s, const aps_key* key, const void* src, uint32_t len) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!key || (!src && len)) return APS_EINVAL;
  if (len > s->cfg.page_size) return APS_EINVAL;
  if (s->faults & APS_FAULT_PUT) return APS_EFAULT;
  DeviceGuard g(s->cfg.device);
  aps_put_req rq{*key, len, 0, src};
  std::unique_lock<std::mutex> lk(s->mu);
  if (len && !is_device_accessible(src)) {  // pageable caller memory: stage through the pinned page
    memcpy(s->h_bounce, src, len);
    rq.src = s->h_bounce;
  }
  aps_put_req* h_req = reinterpret_cast<aps_put_req*>(s->h_stage);
  aps_result* h_res = reinterpret_cast<aps_result*>(s->h_stage + (size_t)s->cfg.max_batch * sizeof(aps_get_req));
  // One page, one launch, no staging copies: the kernel reads the request from, and writes the result to, the
  // pinned staging area itself (page-locked host memory is device-visible), so a call costs a launch and a sync.
  *h_req = rq;
  s->host_form = true;
  rc = s->fused_put ? aps_put_batch_dev(s, 1, h_req, h_res, s->stream) : APS_EINVAL;
  s->host_form = false;
  if (!s->fused_put) {
    CUDA_TRY(cudaMemcpyAsync(s->d_reqs, h_req, sizeof(rq), cudaMemcpyHostToDevice, s->stream));
    rc = aps_put_batch_dev(s, 1, (const aps_put_req*)s->d_reqs, s->d_res, s->stream);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(h_res, s->d_res, sizeof(aps_result), cudaMemcpyDeviceToHost, s->stream));
  }
  if (rc) return rc;
  rc = check_device_error(s, s->stream);
  if (rc) return rc;
  return h_res->len < 0 ? -h_res->len : APS_OK;
}

int aps_get(aps_store* s, const aps_key* key, uint32_t page_off, uint32_t bytes_to_read, void* dst, uint32_t dst_cap,
            uint32_t* out_len, uint32_t* out_crc) {
  int rc = check_open(s);
  if (rc) return rc;
  if (out_len) *out_len = 0;
  if (out_crc) *out_crc = 0;
  if (!key) return APS_EINVAL;
  const uint32_t want = std::min(bytes_to_read, dst_cap);
  if (!dst && want) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::unique_lock<std::mutex> lk(s->mu);
  const bool direct = want == 0 || is_device_accessible(dst);
  aps_get_req rq{*key, page_off, std::min(want, s->cfg.page_size), direct ? dst : (void*)s->h_bounce};
  aps_get_req* h_req = reinterpret_cast<aps_get_req*>(s->h_stage);
  aps_result* h_res = reinterpret_cast<aps_result*>(s->h_stage + (size_t)s->cfg.max_batch * sizeof(aps_get_req));
  *h_req = rq;
  s->host_form = true;
  rc = aps_get_batch_dev(s, 1, h_req, h_res, s->stream);  // request and result stay in the pinned staging area (see aps_put)
  s->host_form = false;
  if (rc) return rc;
  rc = check_device_error(s, s->stream);
  if (rc) return rc;
  if (h_res->len < 0) return -h_res->len;
  if (!direct && h_res->len > 0) memcpy(dst, s->h_bounce, (size_t)h_res->len);
  if (out_len) *out_len = (uint32_t)h_res->len;
  if (out_crc) *out_crc = h_res->crc;
  return APS_OK;
}

int aps_delete(aps_store* s, const aps_key* key) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!key) return APS_EINVAL;
  if (s->faults & APS_FAULT_DELETE) return APS_EFAULT;
  aps_result r{0, 0};
  rc = aps_delete_batch(s, 1, key, &r, nullptr);
  if (rc) return rc;
  return r.len < 0 ? -r.len : APS_OK;
}

int aps_iterate(aps_store* s, aps_iter_cb cb, void* user) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!cb) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::vector<Entry> host;
  {
    std::lock_guard<std::mutex> lk(s->mu);
    host.resize(s->n_entries);
    CUDA_TRY(cudaMemcpyAsync(host.data(), s->table.entries, s->n_entries * sizeof(Entry), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
  }
  for (const Entry& e : host) {
    if (!(e.meta >> 63)) continue;
    aps_key k{e.k0, e.k1, e.k2};
    if (cb(user, &k, meta_len(e.meta))) break;
  }
  return APS_OK;
}

int aps_compact(aps_store* s) {
  int rc = check_open(s);
  if (rc) return rc;
  if (s->n_shards > 0) return APS_EINVAL;  // attached shard tables (own copy included) point into the current table
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  Entry* fresh = nullptr;
  CUDA_TRY(cudaMalloc(&fresh, s->n_entries * sizeof(Entry)));
  cudaError_t e = cudaMemsetAsync(fresh, 0, s->n_entries * sizeof(Entry), s->stream);
  if (e == cudaSuccess) {
    Table nt = s->table;
    nt.entries = fresh;
    const uint64_t threads = s->n_entries * 8;
    s->last_was_get = false;
    rehash_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s->stream>>>(nt, s->table.entries, s->n_entries);
    count_launch(s);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemsetAsync(&s->table.counters[kCtrTombs], 0, sizeof(unsigned long long), s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
  }
  if (e != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "aps_compact: %s", cudaGetErrorString(e));
    cudaFree(fresh);
    return APS_ECUDA;
  }
  cudaFree(s->table.entries);
  s->table.entries = fresh;
  return APS_OK;
}

int aps_clock_evict_dev(aps_store* s, uint32_t n, aps_key* d_victims, uint32_t* d_count, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!s->d_clock) return APS_EINVAL;  // the store was opened without APS_FLAG_DEVICE_CLOCK
  if (n == 0) return APS_OK;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  rc = clock_evict(s, 0, n, d_victims, n, st);
  if (rc) return rc;
  if (d_count) CUDA_TRY(cudaMemcpyAsync(d_count, s->d_clock + 5, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  return APS_OK;
}

int aps_stats(aps_store* s, aps_stats_t* out) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!out) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  unsigned long long c[kNumCounters];
  int32_t top = 0;
  CUDA_TRY(cudaMemcpyAsync(c, s->table.counters, sizeof(c), cudaMemcpyDeviceToHost, s->stream));
  CUDA_TRY(cudaMemcpyAsync(&top, s->table.free_top, sizeof(top), cudaMemcpyDeviceToHost, s->stream));
  CUDA_TRY(cudaStreamSynchronize(s->stream));
  out->pages = c[kCtrPages];
  out->bytes = c[kCtrBytes];
  out->slots_free = (uint64_t)std::max(top, 0);
  out->table_capacity = s->n_entries;
  out->tombstones = c[kCtrTombs];
  out->lookups = c[kCtrLookups];
  out->probe_buckets = c[kCtrProbes];
  out->bytes_copied = c[kCtrBytesCopied];
  out->crc_mismatches = c[kCtrCrcMismatch];
  out->kernel_launches = s->launches.load();
  out->slow_sms = (uint64_t)s->dc->n_slow;
  out->clock_evictions = c[kCtrEvicted];
  return APS_OK;
}

int aps_profile_begin(aps_store* s, uint32_t max_launches) {
  int rc = check_open(s);
  if (rc) return rc;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  while (s->prof_events.size() < (size_t)max_launches * 2) {
    cudaEvent_t e;
    CUDA_TRY(cudaEventCreate(&e));
    s->prof_events.push_back(e);
  }
  s->prof_used = 0;
  s->prof_on = true;
  return APS_OK;
}

int aps_profile_end(aps_store* s, double* copy_ms_total, uint32_t* copy_launches) {
  int rc = check_open(s);
  if (rc) return rc;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  s->prof_on = false;
  double total = 0;
  for (size_t i = 0; i + 1 < s->prof_used; i += 2) {
    CUDA_TRY(cudaEventSynchronize(s->prof_events[i + 1]));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, s->prof_events[i], s->prof_events[i + 1]));
    total += ms;
  }
  if (copy_ms_total) *copy_ms_total = total;
  if (copy_launches) *copy_launches = (uint32_t)(s->prof_used / 2);
  s->prof_used = 0;
  return APS_OK;
}

// ---------------------------------------------------------------- sharded cache (several GPUs)

namespace {
struct ShardHandleBlob {
  cudaIpcMemHandle_t slab, entries, slot_crc;
  uint64_t bucket_mask;
  uint32_t page_size;
  int32_t device;
  uint32_t magic;
};
static_assert(sizeof(ShardHandleBlob) <= sizeof(aps_shard_handle), "handle blob must fit the opaque handle");
constexpr uint32_t kShardMagic = 0x41505348u;  // "APSH"

int install_shards(aps_store* s, const std::vector<Table>& tables, uint32_t my_shard) {
  cudaFree(s->d_shards);
  s->d_shards = nullptr;
  CUDA_TRY(cudaMalloc(&s->d_shards, tables.size() * sizeof(Table)));
  CUDA_TRY(cudaMemcpy(s->d_shards, tables.data(), tables.size() * sizeof(Table), cudaMemcpyHostToDevice));
  s->n_shards = (uint32_t)tables.size();
  s->my_shard = my_shard;
  return APS_OK;
}
}  // namespace

uint32_t aps_shard_of(const aps_key* key, uint32_t n_shards) { return (key && n_shards) ? shard_of(*key, n_shards) : 0; }

int aps_shard_export(aps_store* s, aps_shard_handle* out) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!out) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  ShardHandleBlob b{};
  CUDA_TRY(cudaIpcGetMemHandle(&b.slab, s->table.slab));
  CUDA_TRY(cudaIpcGetMemHandle(&b.entries, s->table.entries));
  CUDA_TRY(cudaIpcGetMemHandle(&b.slot_crc, s->table.slot_crc));
  b.bucket_mask = s->table.bucket_mask;
  b.page_size = s->table.page_size;
  b.device = s->cfg.device;
  b.magic = kShardMagic;
  memset(out, 0, sizeof(*out));
  memcpy(out, &b, sizeof(b));
  return APS_OK;
}

int aps_shard_attach(aps_store* s, uint32_t n_shards, uint32_t my_shard, const aps_shard_handle* handles) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!handles || n_shards == 0 || my_shard >= n_shards) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  std::vector<Table> tables(n_shards);
  for (uint32_t i = 0; i < n_shards; ++i) {
    if (i == my_shard) {
      tables[i] = s->table;
      continue;
    }
    ShardHandleBlob b;
    memcpy(&b, &handles[i], sizeof(b));
    if (b.magic != kShardMagic || b.page_size != s->table.page_size) return APS_EINVAL;
    void *slab = nullptr, *entries = nullptr, *crc = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&slab, b.slab, cudaIpcMemLazyEnablePeerAccess));
    s->ipc_ptrs.push_back(slab);
    CUDA_TRY(cudaIpcOpenMemHandle(&entries, b.entries, cudaIpcMemLazyEnablePeerAccess));
    s->ipc_ptrs.push_back(entries);
    CUDA_TRY(cudaIpcOpenMemHandle(&crc, b.slot_crc, cudaIpcMemLazyEnablePeerAccess));
    s->ipc_ptrs.push_back(crc);
    tables[i] = Table{(Entry*)entries, b.bucket_mask, nullptr, nullptr, (uint32_t*)crc, nullptr, (uint8_t*)slab, b.page_size};
  }
  return install_shards(s, tables, my_shard);
}

int aps_shard_attach_local(aps_store* s, uint32_t n_shards, uint32_t my_shard, aps_store* const* stores) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!stores || n_shards == 0 || my_shard >= n_shards) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  std::vector<Table> tables(n_shards);
  for (uint32_t i = 0; i < n_shards; ++i) {
    aps_store* p = i == my_shard ? s : stores[i];
    if (!p || p->closed || p->table.page_size != s->table.page_size) return APS_EINVAL;
    if (p->cfg.device != s->cfg.device) {
      cudaError_t e = cudaDeviceEnablePeerAccess(p->cfg.device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        snprintf(tl_cuda_error, sizeof(tl_cuda_error), "cudaDeviceEnablePeerAccess(%d): %s", p->cfg.device, cudaGetErrorString(e));
        return APS_ECUDA;
      }
      cudaGetLastError();
    }
    tables[i] = p->table;
    tables[i].free_slots = nullptr;
    tables[i].free_top = nullptr;
    if (i != my_shard) tables[i].counters = nullptr;
  }
  return install_shards(s, tables, my_shard);
}

int aps_get_batch_sharded_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n == 0) return APS_OK;
  if (n > s->cfg.max_batch || !d_reqs || !d_res || s->n_shards == 0) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  return launch_copy(s, n, d_reqs, d_res, nullptr, st, /*sharded=*/true);
}

// ---------------------------------------------------------------- push transport (owner-side serve kernels)

namespace {
struct PushHandleBlob {
  cudaIpcMemHandle_t landing, ctrl;
  uint64_t landing_bytes;
  uint32_t cap, depth;
  int32_t device;
  uint32_t magic;
  char pci[16];  // PCI bus id of the exporting store's GPU: ordinals mean nothing across processes
};
static_assert(sizeof(PushHandleBlob) <= sizeof(aps_push_handle), "handle blob must fit the opaque handle");
constexpr uint32_t kPushMagic = 0x41505055u;  // "APPU"

// Control block of one store (sized for kPushMaxShards so that it can be exported before the shard count is known):
//   [0, off_done)            inboxes [depth][n_shards] x (PushInboxHead + cap x PushReq), written by the requesters
//   [off_done, off_results)  done words, one per owner, written by the owners
//   [off_results, end)       result slots [depth][cap], written by the owners
void push_layout(aps_store::Push& p) {
  p.off_done = (uint64_t)p.depth * kPushMaxShards * push_inbox_stride(p.cap);
  p.off_done = (p.off_done + 255u) & ~255ull;
  p.off_results = p.off_done + 256;
  p.ctrl_bytes = p.off_results + (uint64_t)p.depth * p.cap * sizeof(aps_result);
}

int push_install(aps_store* s, uint32_t n_shards, uint32_t me, uint8_t* const* landing, uint8_t* const* ctrl, uint32_t sharing) {
  aps_store::Push& p = s->push;
  p.sharing = std::max(1u, sharing);
  PushServe h{};
  h.n_shards = n_shards;
  h.me = me;
  h.cap = p.cap;
  h.depth = p.depth;
  h.inbox = p.ctrl;
  h.served = p.d_served;
  h.timeout_ns = p.timeout_ns;
  for (uint32_t r = 0; r < n_shards; ++r) {
    h.landing[r] = landing[r];
    h.results[r] = reinterpret_cast<aps_result*>(ctrl[r] + p.off_results);
    h.done[r] = reinterpret_cast<unsigned long long*>(ctrl[r] + p.off_done) + me;
    p.peer_ctrl[r] = ctrl[r];
  }
  CUDA_TRY(cudaMemcpy(p.d_serve, &h, sizeof(h), cudaMemcpyHostToDevice));
  p.n_shards = n_shards;
  p.me = me;
  p.attached = true;
  return APS_OK;
}
}  // namespace

int aps_push_setup(aps_store* s, uint64_t landing_bytes, uint32_t depth) {
  int rc = check_open(s);
  if (rc) return rc;
  if (landing_bytes == 0 || depth == 0 || depth > 8 || !(s->cfg.flags & APS_FLAG_CRC)) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  aps_store::Push& p = s->push;
  if (p.ctrl) return APS_EINVAL;  // once per store
  {
    // The three kernels of a round wait for each other across streams and GPUs.  With lazy module loading the FIRST
    // launch of a kernel may have to wait for the device to go idle, which it never does while a serve kernel spins
    // for a posting that this very launch would lead to.  Load them now.
    cudaFuncAttributes fa{};
    CUDA_TRY(cudaFuncGetAttributes(&fa, push_post_kernel));
    CUDA_TRY(cudaFuncGetAttributes(&fa, push_wait_kernel));
    CUDA_TRY(cudaFuncGetAttributes(&fa, copy_kernel<true, 512, 4, true, false, true>));
    // The small kernels must be able to run BESIDE a serve CTA (serve kernels of consecutive rounds can hold every SM
    // between them).  An SM is configured for one split of its L1 / shared memory at a time, and the serve kernel needs
    // the largest shared part: ask for the same split, or the small kernels would wait for an SM to drain.
    CUDA_TRY(cudaFuncSetAttribute(push_post_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CUDA_TRY(cudaFuncSetAttribute(push_wait_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CUDA_TRY(cudaFuncSetAttribute(copy_kernel<true, 512, 4, true, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    if (const char* e = getenv("APS_PUSH_TIMEOUT_MS")) p.timeout_ns = (unsigned long long)std::max(1, atoi(e)) * 1000000ull;
  }
  p.cap = s->cfg.max_batch;
  p.depth = depth;
  p.landing_bytes = landing_bytes;
  push_layout(p);
  int prio_least = 0, prio_greatest = 0;
  const uint64_t jobs = (uint64_t)kPushMaxShards * p.cap;  // an owner may be asked for every request of every peer
  cudaError_t e;
  if ((e = cudaMalloc(&p.landing, landing_bytes)) != cudaSuccess || (e = cudaMalloc(&p.ctrl, p.ctrl_bytes)) != cudaSuccess ||
      (e = cudaMemset(p.ctrl, 0, p.ctrl_bytes)) != cudaSuccess || (e = cudaMalloc(&p.d_serve, sizeof(PushServe))) != cudaSuccess ||
      (e = cudaMalloc(&p.d_acc, jobs * sizeof(uint32_t))) != cudaSuccess || (e = cudaMemset(p.d_acc, 0, jobs * sizeof(uint32_t))) != cudaSuccess ||
      (e = cudaMalloc(&p.d_done, jobs * sizeof(uint32_t))) != cudaSuccess || (e = cudaMemset(p.d_done, 0, jobs * sizeof(uint32_t))) != cudaSuccess ||
      (e = cudaMalloc(&p.d_sched, 4 * sizeof(uint32_t))) != cudaSuccess || (e = cudaMemset(p.d_sched, 0, 4 * sizeof(uint32_t))) != cudaSuccess ||
      (e = cudaMalloc(&p.d_served, kPushMaxShards * sizeof(uint32_t))) != cudaSuccess ||
      (e = cudaMemset(p.d_served, 0, kPushMaxShards * sizeof(uint32_t))) != cudaSuccess ||
      (e = cudaStreamCreateWithFlags(&p.serve_stream, cudaStreamNonBlocking)) != cudaSuccess ||
      (e = cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest)) != cudaSuccess ||
      (e = cudaStreamCreateWithPriority(&p.hp_stream, cudaStreamNonBlocking, prio_greatest)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&p.ev_in, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaEventCreateWithFlags(&p.ev_out, cudaEventDisableTiming)) != cudaSuccess ||
      (e = cudaDeviceSynchronize()) != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "aps_push_setup: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? APS_ENOMEM : APS_ECUDA;
  }
  return APS_OK;
}

void* aps_push_landing(aps_store* s, uint64_t* bytes_out) {
  if (!s || s->closed) return nullptr;
  if (bytes_out) *bytes_out = s->push.landing_bytes;
  return s->push.landing;
}

int aps_push_export(aps_store* s, aps_push_handle* out) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!out || !s->push.ctrl) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  PushHandleBlob b{};
  CUDA_TRY(cudaIpcGetMemHandle(&b.landing, s->push.landing));
  CUDA_TRY(cudaIpcGetMemHandle(&b.ctrl, s->push.ctrl));
  b.landing_bytes = s->push.landing_bytes;
  b.cap = s->push.cap;
  b.depth = s->push.depth;
  b.device = s->cfg.device;
  b.magic = kPushMagic;
  CUDA_TRY(cudaDeviceGetPCIBusId(b.pci, sizeof(b.pci), s->cfg.device));
  memset(out, 0, sizeof(*out));
  memcpy(out, &b, sizeof(b));
  return APS_OK;
}

int aps_push_attach(aps_store* s, uint32_t n_shards, uint32_t my_shard, const aps_push_handle* handles) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!handles || n_shards == 0 || n_shards > kPushMaxShards || my_shard >= n_shards || !s->push.ctrl) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  uint8_t *landing[kPushMaxShards] = {}, *ctrl[kPushMaxShards] = {};
  uint32_t sharing = 1;
  char my_pci[16] = "";
  CUDA_TRY(cudaDeviceGetPCIBusId(my_pci, sizeof(my_pci), s->cfg.device));
  for (uint32_t i = 0; i < n_shards; ++i) {
    if (i == my_shard) {
      landing[i] = s->push.landing;
      ctrl[i] = s->push.ctrl;
      continue;
    }
    PushHandleBlob b;
    memcpy(&b, &handles[i], sizeof(b));
    if (b.magic != kPushMagic || b.cap != s->push.cap || b.depth != s->push.depth) return APS_EINVAL;
    void *l = nullptr, *c = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&l, b.landing, cudaIpcMemLazyEnablePeerAccess));
    s->push.ipc_ptrs.push_back(l);
    CUDA_TRY(cudaIpcOpenMemHandle(&c, b.ctrl, cudaIpcMemLazyEnablePeerAccess));
    s->push.ipc_ptrs.push_back(c);
    landing[i] = (uint8_t*)l;
    ctrl[i] = (uint8_t*)c;
    if (strncmp(b.pci, my_pci, sizeof(my_pci)) == 0) ++sharing;
  }
  return push_install(s, n_shards, my_shard, landing, ctrl, sharing);
}

int aps_push_attach_local(aps_store* s, uint32_t n_shards, uint32_t my_shard, aps_store* const* stores) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!stores || n_shards == 0 || n_shards > kPushMaxShards || my_shard >= n_shards || !s->push.ctrl) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  uint8_t *landing[kPushMaxShards] = {}, *ctrl[kPushMaxShards] = {};
  uint32_t sharing = 1;
  for (uint32_t i = 0; i < n_shards; ++i) {
    aps_store* p = i == my_shard ? s : stores[i];
    if (!p || p->closed || !p->push.ctrl || p->push.cap != s->push.cap || p->push.depth != s->push.depth) return APS_EINVAL;
    if (p->cfg.device != s->cfg.device) {
      cudaError_t e = cudaDeviceEnablePeerAccess(p->cfg.device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
        snprintf(tl_cuda_error, sizeof(tl_cuda_error), "cudaDeviceEnablePeerAccess(%d): %s", p->cfg.device, cudaGetErrorString(e));
        return APS_ECUDA;
      }
      cudaGetLastError();
    }
    landing[i] = p->push.landing;
    ctrl[i] = p->push.ctrl;
    if (i != my_shard && p->cfg.device == s->cfg.device) ++sharing;
  }
  return push_install(s, n_shards, my_shard, landing, ctrl, sharing);
}

int aps_get_batch_push_begin_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  aps_store::Push& p = s->push;
  if (!p.attached || n > p.cap || (n && !d_reqs)) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  if (p.seq_begun - p.seq_ended >= p.depth) return APS_EINVAL;  // more multi-gets in flight than the store was set up for
  const uint32_t seq = ++p.seq_begun;
  p.n_of[seq % p.depth] = n;
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  // requester side: route the batch to the owners' inboxes (a collective round even when n == 0)
  PushPostArgs pa{};
  pa.reqs = d_reqs;
  pa.n = n;
  pa.n_shards = p.n_shards;
  pa.me = p.me;
  pa.cap = p.cap;
  pa.depth = p.depth;
  pa.seq = seq;
  for (uint32_t o = 0; o < p.n_shards; ++o) pa.peer_inbox[o] = p.peer_ctrl[o];
  pa.landing = p.landing;
  pa.landing_bytes = p.landing_bytes;
  pa.my_results = reinterpret_cast<aps_result*>(p.ctrl + p.off_results);
  // The post and wait kernels run on a stream of the HIGHEST priority, tied into the caller's stream with events.  The
  // GPU hands out CTAs of equal priority in launch order: behind a serve kernel that still has CTAs to place (serve
  // kernels of consecutive rounds overlap and hold every SM between them) a post kernel of the caller's priority would
  // wait for the very serve kernels that wait for it.  With priority it is placed next, beside a serve CTA.
  CUDA_TRY(cudaEventRecord(p.ev_in, st));
  CUDA_TRY(cudaStreamWaitEvent(p.hp_stream, p.ev_in, 0));
  push_post_kernel<<<1, 256, 0, p.hp_stream>>>(pa);
  CUDA_TRY(cudaGetLastError());
  // owner side: serve what the peers post for this round.  One CTA per SM but one: the SM left over runs the small
  // post and wait kernels of this GPU, which the serve kernels of this and the other GPUs wait for (and a serve CTA is
  // compiled small enough for those kernels to fit beside it on any SM, see copy_kernel).
  CopyArgs a{};
  a.table = s->table;
  a.crc_on = 1;
  a.res = nullptr;
  a.acc = p.d_acc;
  a.done = p.d_done;
  a.tables = s->dc->d_tables;
  a.table_image = s->use_table_image ? s->dc->d_table_image : nullptr;
  a.rowshift = s->d_rowshift;
  a.init_full = s->init_full;
  a.error_flag = s->h_error;
  a.steal = s->use_steal ? 1 : 0;
  a.sched = p.d_sched;
  a.sm_slow = s->dc->d_sm_slow;
  a.n_slow = (uint32_t)s->dc->n_slow;
  a.timeline = s->d_timeline;  // diagnostics (aps_debug_timeline): stamps of the LAST round; launches are not separated
  a.push = p.d_serve;
  a.push_seq = seq;
  a.serve_rows = s->page_rows;
  // With four or more shards the tasks are dealt to the CTAs round robin (the local requester's cheap tasks are then
  // spread over all CTAs: 620 against 616 GB/s at 8 GPUs); with two, half of the tasks are local and whole CTAs of
  // them finish early and start on the next round, which measured faster (632 against 610 GB/s).
  static const int serve_flags = [] { const char* e = getenv("APS_SERVE_FLAGS"); return e ? atoi(e) : -1; }();
  a.serve_flags = serve_flags >= 0 ? (uint32_t)serve_flags : (p.n_shards >= 4 ? 2u : 0u);
  static const int spare = [] { const char* e = getenv("APS_SERVE_SPARE_SMS"); return e ? std::max(0, atoi(e)) : 1; }();
  // several shards on ONE device (tests, small deployments) split the SMs, so that their serve kernels together never
  // hold every SM while they wait for a posting
  const uint32_t grid = (uint32_t)std::max(1, (s->dc->num_sms - spare) / (int)p.sharing);
  const uint32_t warps = grid * 16u;
  a.serve_warps = warps > (uint32_t)s->dc->n_slow + p.cap ? warps - (uint32_t)s->dc->n_slow : warps;
  // Consecutive rounds overlap (programmatic dependent launch): the CTAs of round k+1 land on SMs as round k's retire, fetch
  // the CRC tables, pick up round k+1's postings, resolve their pages and warm L2 with their first rows, and only then
  // wait for round k to finish; everything they write stays behind that wait.  (Postings are final once announced.)
  a.early = s->use_pdl && s->use_early ? 1 + (int)s->prefetch_rows : 0;
  a.trust_reqs = 1;
  {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(grid);
    lc.blockDim = dim3(512);
    lc.dynamicSmemBytes = kCopyDynSmem;
    lc.stream = p.serve_stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = s->use_pdl ? 1 : 0;
    lc.attrs = attr;
    lc.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&lc, copy_kernel<true, 512, 4, true, false, true>, a));
  }
  CUDA_TRY(cudaGetLastError());
  count_launch(s, 2);
  return APS_OK;
}

int aps_get_batch_push_end_dev(aps_store* s, aps_result* d_res, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  aps_store::Push& p = s->push;
  if (!p.attached) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  if (p.seq_ended == p.seq_begun) return APS_EINVAL;  // nothing in flight
  const uint32_t seq = ++p.seq_ended;
  const uint32_t n = p.n_of[seq % p.depth];
  if (n && !d_res) return APS_EINVAL;
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  push_wait_kernel<<<1, 256, 0, p.hp_stream>>>(reinterpret_cast<const unsigned long long*>(p.ctrl + p.off_done), p.n_shards, seq,
                                      reinterpret_cast<const aps_result*>(p.ctrl + p.off_results) + (size_t)(seq % p.depth) * p.cap,
                                      d_res, n, s->h_error, p.timeout_ns);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaEventRecord(p.ev_out, p.hp_stream));
  CUDA_TRY(cudaStreamWaitEvent(st, p.ev_out, 0));  // `stream` continues once the batch has landed
  count_launch(s);
  return APS_OK;
}

int aps_get_batch_push_dev(aps_store* s, uint32_t n, const aps_get_req* d_reqs, aps_result* d_res, aps_stream stream) {
  int rc = aps_get_batch_push_begin_dev(s, n, d_reqs, stream);
  if (rc) return rc;
  return aps_get_batch_push_end_dev(s, d_res, stream);
}

int aps_debug_timeline(aps_store* s, int enable, uint64_t* out, uint32_t max_ctas) {
  int rc = check_open(s);
  if (rc) return rc;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  if (enable && !s->d_timeline) CUDA_TRY(cudaMalloc(&s->d_timeline, kTimelineWords * sizeof(unsigned long long)));
  if (out && s->d_timeline) {
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpy(out, s->d_timeline, std::min<size_t>(max_ctas, kTimelineWords) * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  }
  if (!enable && s->d_timeline) {
    CUDA_TRY(cudaDeviceSynchronize());
    cudaFree(s->d_timeline);
    s->d_timeline = nullptr;
  }
  return APS_OK;
}

int aps_debug_corrupt(aps_store* s, const aps_key* key, uint32_t byte_offset) {
  int rc = check_open(s);
  if (rc) return rc;
  if (!key) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  int found = 0;
  s->last_was_get = false;
  corrupt_kernel<<<1, 8, 0, s->stream>>>(s->table, *key, byte_offset, s->d_error + 0);
  count_launch(s);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpyAsync(&found, s->d_error, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  CUDA_TRY(cudaMemsetAsync(s->d_error, 0, sizeof(int), s->stream));
  CUDA_TRY(cudaStreamSynchronize(s->stream));
  return found ? APS_OK : APS_ENOTFOUND;
}

// ---------------------------------------------------------------- stateless entry points

int aps_fill_pages_dev(int32_t device, void* d_dst, uint64_t stride, uint32_t page_len, uint32_t n_pages,
                       const uint64_t* d_files, const uint64_t* d_indices, uint64_t seed, aps_stream stream) {
  if (!d_dst || !d_files || !d_indices) return APS_EINVAL;
  if (n_pages == 0 || page_len == 0) return APS_OK;
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  const uint64_t total = (uint64_t)((page_len + 7) / 8) * n_pages;
  const unsigned grid = (unsigned)std::min<uint64_t>((total + 255) / 256, 148ull * 32);
  fill_pages_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((uint8_t*)d_dst, stride, page_len, n_pages, d_files, d_indices, seed);
  count_launch(nullptr);
  CUDA_TRY(cudaGetLastError());
  return APS_OK;
}

int aps_fill_pages(aps_store* s, void* d_dst, uint64_t stride, uint32_t page_len, uint32_t n_pages, const uint64_t* files,
                   const uint64_t* indices, uint64_t seed, aps_stream stream) {
  int rc = check_open(s);
  if (rc) return rc;
  if (n_pages == 0) return APS_OK;
  if (!d_dst || !files || !indices || n_pages > s->cfg.max_batch) return APS_EINVAL;
  DeviceGuard g(s->cfg.device);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  uint64_t* h = reinterpret_cast<uint64_t*>(s->h_stage);  // 16 bytes per page <= sizeof(aps_get_req)
  memcpy(h, files, (size_t)n_pages * 8);
  memcpy(h + n_pages, indices, (size_t)n_pages * 8);
  CUDA_TRY(cudaMemcpyAsync(s->d_reqs, h, (size_t)n_pages * 16, cudaMemcpyHostToDevice, st));
  const uint64_t* d = reinterpret_cast<const uint64_t*>(s->d_reqs);
  rc = aps_fill_pages_dev(s->cfg.device, d_dst, stride, page_len, n_pages, d, d + n_pages, seed, st);
  if (rc) return rc;
  CUDA_TRY(cudaStreamSynchronize(st));  // the staging area is reused by the next call
  return APS_OK;
}

int aps_device_alloc(int32_t device, uint64_t bytes, void** out) {
  if (!out) return APS_EINVAL;
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "cudaMalloc(%llu): %s", (unsigned long long)bytes, cudaGetErrorString(e));
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? APS_ENOMEM : APS_ECUDA;
  }
  return APS_OK;
}
int aps_host_alloc_pinned(uint64_t bytes, void** out) {
  if (!out) return APS_EINVAL;
  cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocDefault);
  if (e != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "cudaHostAlloc(%llu): %s", (unsigned long long)bytes, cudaGetErrorString(e));
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? APS_ENOMEM : APS_ECUDA;
  }
  return APS_OK;
}
int aps_host_free_pinned(void* p) {
  CUDA_TRY(cudaFreeHost(p));
  return APS_OK;
}
int aps_device_free(int32_t device, void* p) {
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  CUDA_TRY(cudaFree(p));
  return APS_OK;
}

int aps_md5_batch_dev(int32_t device, uint32_t n_files, const void* const* d_data, const uint64_t* d_len,
                      uint8_t (*d_out)[16], aps_stream stream) {
  if (n_files == 0) return APS_OK;
  if (!d_data || !d_len || !d_out) return APS_EINVAL;
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  if (md5_prepare(device) != APS_OK) return APS_ECUDA;
  md5_kernel<<<(n_files + 31) / 32, 32, kMd5DynSmem, (cudaStream_t)stream>>>(n_files, d_data, d_len, d_out, nullptr, nullptr, 1);
  count_launch(nullptr);
  CUDA_TRY(cudaGetLastError());
  return APS_OK;
}

int aps_md5_stream_dev(int32_t device, uint32_t n_files, const void* const* d_data, const uint64_t* d_len,
                       uint32_t (*d_state)[4], const uint64_t* d_prefix_len, int finalize, uint8_t (*d_out)[16],
                       aps_stream stream) {
  if (n_files == 0) return APS_OK;
  if (!d_data || !d_len || !d_state || (finalize && !d_out)) return APS_EINVAL;
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  if (md5_prepare(device) != APS_OK) return APS_ECUDA;
  md5_kernel<<<(n_files + 31) / 32, 32, kMd5DynSmem, (cudaStream_t)stream>>>(n_files, d_data, d_len, d_out, d_state, d_prefix_len,
                                                                    finalize);
  count_launch(nullptr);
  CUDA_TRY(cudaGetLastError());
  return APS_OK;
}

int aps_md5_batch(int32_t device, uint32_t n_files, const void* const* data, const uint64_t* len, uint8_t (*out)[16]) {
  if (n_files == 0) return APS_OK;
  if (!data || !len || !out) return APS_EINVAL;
  DeviceGuard g(device);
  if (!g.ok) return APS_ECUDA;
  void** d_data = nullptr;
  uint64_t* d_len = nullptr;
  uint8_t(*d_out)[16] = nullptr;
  int rc = APS_OK;
  cudaError_t e;
  if ((e = cudaMalloc(&d_data, n_files * sizeof(void*))) != cudaSuccess ||
      (e = cudaMalloc(&d_len, n_files * sizeof(uint64_t))) != cudaSuccess ||
      (e = cudaMalloc(&d_out, n_files * 16)) != cudaSuccess ||
      (e = cudaMemcpy(d_data, data, n_files * sizeof(void*), cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(d_len, len, n_files * sizeof(uint64_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "aps_md5_batch staging: %s", cudaGetErrorString(e));
    rc = APS_ECUDA;
  }
  if (rc == APS_OK) rc = aps_md5_batch_dev(device, n_files, d_data, d_len, d_out, nullptr);
  if (rc == APS_OK && (e = cudaMemcpy(out, d_out, n_files * 16, cudaMemcpyDeviceToHost)) != cudaSuccess) {
    snprintf(tl_cuda_error, sizeof(tl_cuda_error), "aps_md5_batch readback: %s", cudaGetErrorString(e));
    rc = APS_ECUDA;
  }
  cudaFree(d_data);
  cudaFree(d_len);
  cudaFree(d_out);
  return rc;
}

}  // extern "C"
