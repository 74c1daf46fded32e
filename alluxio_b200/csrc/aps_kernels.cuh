// Kernels of libaps (sm_100a).  K1 table kernels, K2/K3 batched copy with fused CRC32C,
// K4 many-file MD5, and the synthetic-page generator.  Host launch code is in aps_api.cu.
#pragma once
#include "aps_device.cuh"

namespace aps {

// ------------------------------------------------------------------------------------------------
// Copy jobs: what a warp-task executes.  Gets and puts build theirs inside the copy kernel (lookup / insert fused);
// the stand-alone checksum pass and the bulk-async variant's puts prepare them in a small kernel of their own.
// ------------------------------------------------------------------------------------------------
struct Job {
  const uint8_t* src;
  uint8_t* dst;  // nullptr: checksum only
  uint32_t n;    // bytes to move; 0 = nothing to do
  uint32_t slot;
  uint32_t flags;
  uint32_t pad;
};
constexpr uint32_t kJobVerify = 1u;  // whole-page get: compare with the CRC recorded at put
constexpr uint32_t kJobRecord = 2u;  // put: record the CRC for the slot
constexpr uint32_t kJobOutCrc = 4u;  // write the CRC to crc_out[job] (stand-alone checksum pass)

constexpr uint32_t kRowBytes = 512;  // one warp-wide 128-bit access

// Shared-memory map of the CRC copy kernel, as absolute shared-window addresses:
//   0x0F000 .. 0x0FFFF  Z_4 tables  G[j][b], j = 0..3 (plain, 4 KiB)         -- row-state reduction
//   0x10000 .. 0x2FFFF  Z_512 tables F[j][b], each entry replicated for the 32 lanes so that lane l
//                       only ever touches bank l (conflict-free): address = 0x10000 + (j>>1)*0x10000
//                       + b*256 + (j&1)*128 + lane*4.  The 64 KiB alignment lets one PRMT build the
//                       address from a data byte (see fold_word).
constexpr uint32_t kSmemP = 0x0E000u;  // plain x^(8*512) tables (4 KiB) of the kernels that fold four rows at a time, see stream_rows
constexpr uint32_t kSmemG = 0x0F000u;
constexpr uint32_t kSmemF = 0x10000u;
constexpr uint32_t kSmemEnd = 0x30000u;
constexpr uint32_t kCopyDynSmem = kSmemEnd;  // requested size; the window starts at <= 1 KiB

struct CrcTables {      // device-global copy, built once on the host
  uint32_t fold[4][256];  // F_j[b] = (b << 8j) * x^(8*512)
  uint32_t z4[4][256];    // G_j[b] = (b << 8j) * x^32
  uint32_t fold4[4][256]; // (b << 8j) * x^(8*512*4): four rows ahead
};

__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// S * x^(8*512): four conflict-free table reads.  rb = 0x00010000 | lane*4.  PRMT picks byte j of S
// into bits 8..15, lane*4 into bits 0..7 and 0x0001 into the upper half: one instruction per address.
__device__ __forceinline__ uint32_t fold_word(uint32_t s, uint32_t rb) {
  uint32_t a0, a1, a2, a3, v0, v1, v2, v3;
  asm("prmt.b32 %0, %1, %2, 0x7604;" : "=r"(a0) : "r"(s), "r"(rb));
  asm("prmt.b32 %0, %1, %2, 0x7614;" : "=r"(a1) : "r"(s), "r"(rb));
  asm("prmt.b32 %0, %1, %2, 0x7624;" : "=r"(a2) : "r"(s), "r"(rb));
  asm("prmt.b32 %0, %1, %2, 0x7634;" : "=r"(a3) : "r"(s), "r"(rb));
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v0) : "r"(a0));
  asm volatile("ld.shared.u32 %0, [%1+128];" : "=r"(v1) : "r"(a1));
  asm volatile("ld.shared.u32 %0, [%1+65536];" : "=r"(v2) : "r"(a2));
  asm volatile("ld.shared.u32 %0, [%1+65664];" : "=r"(v3) : "r"(a3));
  return v0 ^ v1 ^ v2 ^ v3;
}

// s * x^(8*512) from the plain (not replicated, so bank-conflicted, 4 KiB) tables at kSmemP: used a few times per segment
__device__ __forceinline__ uint32_t fold1_plain(uint32_t r) {
  return lds_u32(kSmemP + (r & 0xFFu) * 4u) ^ lds_u32(kSmemP + 1024u + ((r >> 8) & 0xFFu) * 4u) ^
         lds_u32(kSmemP + 2048u + ((r >> 16) & 0xFFu) * 4u) ^ lds_u32(kSmemP + 3072u + (r >> 24) * 4u);
}

// r * x^32 (advance the register over four zero bytes) from the plain tables.
__device__ __forceinline__ uint32_t z4_word(uint32_t r) {
  return lds_u32(kSmemG + (r & 0xFFu) * 4u) ^ lds_u32(kSmemG + 1024u + ((r >> 8) & 0xFFu) * 4u) ^
         lds_u32(kSmemG + 2048u + ((r >> 16) & 0xFFu) * 4u) ^ lds_u32(kSmemG + 3072u + (r >> 24) * 4u);
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream(uint4* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ uint32_t warp_xor(uint32_t v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v ^= __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

// One 512-byte row per call: fold the four lane words forward by a row and add the new data.
#define APS_FOLD_ROW(v_)                                                  \
  do {                                                                    \
    S0 = fold_word(S0, rb) ^ (v_).x; S1 = fold_word(S1, rb) ^ (v_).y;     \
    S2 = fold_word(S2, rb) ^ (v_).z; S3 = fold_word(S3, rb) ^ (v_).w;     \
  } while (0)

// ---- mbarrier / bulk-async (TMA) primitives
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst),
               "l"(gsrc), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// ---- row ranges shared inside a CTA ------------------------------------------------------------
// Every warp publishes the rows it still has to do in one 64-bit shared-memory word:
//   bits  0..31  next unclaimed row (relative to the job)
//   bits 32..59  end row (exclusive)
//   bits 60..63  generation (bumped whenever the warp installs a new range)
// The owner claims G rows at a time with an atomic add on the low half; a warp that has run out of
// work takes the back half of the largest remaining range with a compare-and-swap on the whole word.
// Both are atomics on the same word, so a row is processed exactly once.
constexpr uint32_t kStealMinGroups = 2;  // do not split ranges shorter than this many groups

__device__ __forceinline__ unsigned long long range_pack(uint32_t lo, uint32_t hi, uint32_t gen) {
  return (unsigned long long)lo | ((unsigned long long)(hi & 0x0FFFFFFFu) << 32) | ((unsigned long long)(gen & 15u) << 60);
}
__device__ __forceinline__ uint32_t range_lo(unsigned long long w) { return (uint32_t)w; }
__device__ __forceinline__ uint32_t range_hi(unsigned long long w) { return (uint32_t)(w >> 32) & 0x0FFFFFFFu; }
__device__ __forceinline__ uint32_t range_gen(unsigned long long w) { return (uint32_t)(w >> 60); }

struct Claim {
  uint32_t pos, cnt;
};
template <int G>
__device__ __forceinline__ Claim claim_rows(unsigned long long* range, uint32_t lane) {
  unsigned long long old = 0;
  if (lane == 0) old = atomicAdd(range, (unsigned long long)G);
  old = __shfl_sync(0xFFFFFFFFu, old, 0);
  const uint32_t pos = range_lo(old), hi = range_hi(old);
  return Claim{pos, pos < hi ? min((uint32_t)G, hi - pos) : 0u};
}

// Aligned fast path: streams the rows this warp can claim from *range (512-byte rows, 16 bytes per
// lane per row).  Software pipelined: while the rows of one group are stored and folded, the loads of
// the next claimed group are already in flight.  Returns the zero-init CRC of the rows it processed
// (same value in every lane, when kCrc); *row_end receives the first row it did NOT process.
struct NoGate {
  __device__ __forceinline__ void operator()() const {}
};
// `gate` runs once the first loads are in flight and before anything is stored: the copy kernel uses it to
// wait for the previous kernel in the stream (programmatic dependent launch) as late as it safely can.
// kIndep (groups of four rows): the G rows of a group are folded into G INDEPENDENT states, each carried forward by a
// whole group (the replicated tables then hold x^(8*512*G)), so that the sixteen table reads of a group do not wait for
// each other (serially, row u+1 needs row u's result: four dependent shared-memory round trips per group and word).
// The states are merged -- Horner with the plain one-row tables -- once, at the end of the segment, or just before a
// last, incomplete group, whose rows then go through the one-row tables as well.
template <bool kCrc, int G, bool kIndep, typename Gate = NoGate>
__device__ __forceinline__ uint32_t stream_rows(const uint8_t* src, uint8_t* dst, unsigned long long* range, uint32_t row_lo,
                                                uint32_t lane, uint32_t* row_end, uint32_t tables_bar, Gate gate = Gate()) {
  const uint4* s = reinterpret_cast<const uint4*>(src) + lane;  // src/dst point at row 0 of the job
  uint4* d = reinterpret_cast<uint4*>(dst) + lane;
  const bool store = dst != nullptr;
  const uint32_t rb = 0x00010000u | (lane << 2);
  uint32_t S0 = 0, S1 = 0, S2 = 0, S3 = 0;
  constexpr int kChains = kIndep ? G : 1;
  uint4 C[kChains];  // kIndep: chain u holds rows u, u + G, ... of the segment; chain 0 doubles as the merged state
#pragma unroll
  for (int u = 0; u < kChains; ++u) C[u] = make_uint4(0u, 0u, 0u, 0u);
  bool merged = false;
  auto merge = [&]() {  // C[0] = (((C0 * X + C1) * X + C2) * X + C3), X = x^(8*512)
    if (merged) return;
    merged = true;
#pragma unroll
    for (int u = 1; u < kChains; ++u) {
      C[0].x = fold1_plain(C[0].x) ^ C[u].x; C[0].y = fold1_plain(C[0].y) ^ C[u].y;
      C[0].z = fold1_plain(C[0].z) ^ C[u].z; C[0].w = fold1_plain(C[0].w) ^ C[u].w;
    }
  };
  uint4 A[G], B[G];
  auto load = [&](uint4(&buf)[G], const Claim& c) {
#pragma unroll
    for (int u = 0; u < G; ++u)
      if ((uint32_t)u < c.cnt) buf[u] = ldg_stream(s + (size_t)(c.pos + u) * 32);
  };
  auto consume = [&](uint4(&buf)[G], const Claim& c) {
    if (store) {
#pragma unroll
      for (int u = 0; u < G; ++u)
        if ((uint32_t)u < c.cnt) stg_stream(d + (size_t)(c.pos + u) * 32, buf[u]);
    }
    if (kCrc && !kIndep) {
#pragma unroll
      for (int u = 0; u < G; ++u)
        if ((uint32_t)u < c.cnt) APS_FOLD_ROW(buf[u]);
    }
    if (kCrc && kIndep) {
      if (c.cnt == (uint32_t)G) {
#pragma unroll
        for (int u = 0; u < kChains; ++u) {
          C[u].x = fold_word(C[u].x, rb) ^ buf[u].x; C[u].y = fold_word(C[u].y, rb) ^ buf[u].y;
          C[u].z = fold_word(C[u].z, rb) ^ buf[u].z; C[u].w = fold_word(C[u].w, rb) ^ buf[u].w;
        }
      } else if (c.cnt) {  // the last, incomplete group of the segment
        merge();
        for (uint32_t u = 0; u < c.cnt; ++u) {
          uint4 v = buf[0];
#pragma unroll
          for (int q = 1; q < G; ++q)
            if (u == (uint32_t)q) v = buf[q];
          C[0].x = fold1_plain(C[0].x) ^ v.x; C[0].y = fold1_plain(C[0].y) ^ v.y;
          C[0].z = fold1_plain(C[0].z) ^ v.z; C[0].w = fold1_plain(C[0].w) ^ v.w;
        }
      }
    }
  };
  uint32_t end = row_lo;
  Claim ca = claim_rows<G>(range, lane);
  load(A, ca);
  // the first loads are in flight; only now make sure the CRC tables have landed in shared memory
  if (kCrc && tables_bar) mbar_wait(tables_bar, 0);
  gate();
  // (A third buffer -- two groups in flight while one is stored and folded -- was measured and loses: 2,885 against
  // 2,921 GB/s.  The warps do wait for loaded data at the first store of a group, but the memory system is what is busy.)
  while (ca.cnt) {
    Claim cb = claim_rows<G>(range, lane);
    load(B, cb);
    consume(A, ca);
    end = ca.pos + ca.cnt;
    if (!cb.cnt) break;
    ca = claim_rows<G>(range, lane);
    load(A, ca);
    consume(B, cb);
    end = cb.pos + cb.cnt;
  }
  *row_end = end;
  if (!kCrc) return 0;
  if (kIndep) {
    merge();
    S0 = C[0].x; S1 = C[0].y; S2 = C[0].z; S3 = C[0].w;
  }
  // lane words -> one register: t = Z4(Z4(Z4(S0)^S1)^S2)^S3, then shift to the end of the row
  uint32_t t = z4_word(z4_word(z4_word(S0) ^ S1) ^ S2) ^ S3;
  return warp_xor(gf_mul_call(t, c_lane_shift[lane]));
}

// Generic path (tails, unaligned requests): lane l takes the l-th contiguous slice, one byte at a
// time.  Returns the zero-init CRC of the n bytes (same value in every lane) when kCrc.
template <bool kCrc>
__device__ __forceinline__ uint32_t warp_bytes(const uint8_t* src, uint8_t* dst, uint32_t n, uint32_t lane) {
  uint32_t per = (n + 31u) >> 5;
  uint32_t lo = min(n, lane * per), hi = min(n, lo + per);
  uint32_t r = 0;
  for (uint32_t i = lo; i < hi; ++i) {
    uint32_t b = src[i];
    if (dst) dst[i] = (uint8_t)b;
    if (kCrc) r = crc_byte_bitwise(r, b);
  }
  if (!kCrc) return 0;
  return warp_xor(gf_mul_call(r, xpow8(n - hi)));
}

constexpr uint32_t kMaxTasksPerJob = 128;

// ------------------------------------------------------------------------------------------------
// Push transport of the sharded multi-get.  SM stores into peer memory run ~11 % faster over
// NVLink than SM loads from it when every GPU is busy (700 against 626 GB/s per GPU, tools/micro/p2p_transports.cu),
// so the OWNER of a page moves it: a requester posts its requests into each owner's inbox, the owner's serve kernel
// (this copy kernel with kServe) looks the pages up in its own table, streams them from its own slab into the
// requester's landing buffer with the CRC32C check fused in, writes each result into the requester's result array and,
// after the last page of that requester, raises the requester's done word.  All cross-GPU words are plain device
// memory mapped into the peers (CUDA IPC or peer access); ordering is fence.sys + one flag word per direction.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kPushMaxShards = 8;
struct PushReq {       // 48 bytes, three 16-byte stores
  aps_key key;
  uint32_t page_off, nbytes;
  uint64_t dst_off;    // destination, as an offset into the REQUESTER's landing buffer
  uint32_t res_idx;    // index of the request in the requester's batch
  uint32_t pad;
};
static_assert(sizeof(PushReq) == 48, "PushReq layout");
struct PushInboxHead {  // followed by `cap` PushReq
  unsigned long long post;  // sequence number << 32 | number of requests; written last by the requester's post kernel
  unsigned long long pad;
};
struct PushServe {     // one per store, in device memory; everything an owner needs to serve its peers
  uint32_t n_shards, me, cap, depth;
  uint8_t* inbox;                          // my inboxes: [depth][n_shards] x (PushInboxHead + cap * PushReq)
  uint8_t* landing[kPushMaxShards];        // requester r's landing buffer as mapped here
  aps_result* results[kPushMaxShards];     // requester r's result slots [depth][cap] as mapped here
  unsigned long long* done[kPushMaxShards];  // requester r's done word for owner `me` as mapped here
  uint32_t* served;                        // scratch [n_shards]: pages finished per requester in this launch (self-resetting)
  unsigned long long timeout_ns;           // how long a serve kernel waits for a peer's posting before it gives the peer up
};
__host__ __device__ __forceinline__ size_t push_inbox_stride(uint32_t cap) { return sizeof(PushInboxHead) + (size_t)cap * sizeof(PushReq); }

struct CopyArgs {
  const Job* jobs;           // neither kGet nor kPut: jobs prepared by an earlier kernel
  const aps_get_req* reqs;   // kGet == true: requests, resolved in this kernel
  const aps_put_req* preqs;  // kPut == true: put requests; the first task of a page to arrive inserts it, the others pick its slot up
  unsigned long long* jslot; // kPut scratch, one word per job: launch epoch << 32 | (0 = insert in progress, slot + 1, or 0x80000000 | status)
  uint32_t epoch;            // kPut: this launch's number (never 0), so the words above need no reset
  int trust_reqs;            // kGet with early: the caller promised that request arrays are final (APS_FLAG_STABLE_REQS)
  Table table;
  int crc_on;
  uint32_t n_jobs;
  uint32_t tasks_per_job;  // a job is cut into this many warp-tasks ...
  uint32_t rows_per_task;  // ... of this many 512-byte rows (the last one takes what is left)
  aps_result* res;         // may be null
  uint32_t* crc_out;       // for kJobOutCrc
  uint32_t* acc;           // scratch: per-job XOR of the shifted segment CRCs; zero on entry, reset by the finisher
  uint32_t* done;          // scratch: per-job count of finished 512-byte rows; zero on entry, reset by the finisher
  const CrcTables* tables;
  const uint8_t* table_image;  // optional: the shared-memory image of the CRC tables (bytes 0x0F000..0x30000 of the window),
                               // fetched with four bulk-async copies while the first pages are being looked up
  const uint32_t* rowshift;  // rowshift[k] = x^(8 * 512 * k), k = 0 .. rows of a full page
  int* error_flag;           // set if the shared-memory window is not where the kernel needs it
  unsigned long long* timeline;  // optional (aps_debug_timeline)
  int steal;                 // intra-CTA work stealing on/off
  const Table* shards;       // kGet with n_shards > 1: table + slab of every shard (peer memory mapped over NVLink)
  uint32_t n_shards;         // 0 or 1: only this GPU's table
  uint32_t* sched;           // scratch: {fast CTAs ranked, slow CTAs ranked, CTAs arrived}; zero on entry, self-resetting
  const uint8_t* sm_slow;    // per SM id: 1 = this SM streams measurably slower than the median (calibrated at open)
  uint32_t n_slow;           // how many SMs that is
  uint32_t init_full;        // 0xFFFFFFFF * x^(8 * page size): the init term of a full page's CRC (0 = compute it)
  const PushServe* push;     // kServe: the push transport's maps (device memory)
  uint32_t push_seq;         // kServe: which posting to serve
  uint32_t serve_rows;       // kServe: 512-byte rows of a full page ...
  uint32_t serve_flags;      // kServe experiments: 1 = no system fence per segment (UNSAFE, measurement only); 2 = tasks dealt round robin
  uint32_t serve_warps;      // ... and the warps of the grid: the job split is computed in the kernel, once the posts are in
  int early;                 // kGet, non-zero: requests, table and slab are known not to change under the previous kernel
                             // in the stream, so they may be read before waiting for it (see the kernel's dependency
                             // gate); early - 1 = rows to prefetch into L2 while waiting
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// x^(8 * nbytes) from the per-row table plus (ragged lengths only) the sub-row remainder
__device__ __forceinline__ uint32_t shift_for(const uint32_t* rowshift, uint32_t nbytes) {
  uint32_t sh = __ldg(&rowshift[nbytes >> 9]);
  if (nbytes & 511u) sh = gf_mul_call(sh, xpow8(nbytes & 511u));
  return sh;
}

// kServe: hand one page's result to its requester and count the page; whoever delivers the requester's last page raises
// the requester's done word.  `tag` is Job::pad (requester, index in the requester's batch); lane r of the calling warp
// holds requester r's request count (cnt).  Every warp that stored part of the page fenced at system scope before it
// added its rows to the page's row counter, and the caller observed that counter complete, so the fences below make the
// whole page visible to the requester before the result, and every result before the done word.
__device__ __forceinline__ void push_deliver(const PushServe& ps, uint32_t seq, uint32_t tag, aps_result r, uint32_t lane,
                                             uint32_t cnt) {
  const uint32_t who = (tag >> 24) & 0x7Fu, idx = tag & 0x00FFFFFFu;
  const uint32_t want = __shfl_sync(0xFFFFFFFFu, cnt, who);
  if (lane == 0) {
    ps.results[who][(size_t)(seq % ps.depth) * ps.cap + idx] = r;
    __threadfence_system();
    if (atomicAdd(&ps.served[who], 1u) + 1u == want) {
      ps.served[who] = 0;  // re-armed for the next launch (launches of one store's serve stream do not overlap)
      __threadfence_system();
      *(volatile unsigned long long*)ps.done[who] = seq;
    }
  }
}

// K2 + K3 (+ K1: lookup when kGet, insert when kPut).  Persistent grid: one CTA per SM.  A job (one page request) is
// cut into tasks_per_job warp-tasks of rows_per_task 512-byte rows, sized by the host so that one
// round of all warps of the chip covers the batch (no wave quantisation).  A task resolves its
// page (kGet: eight lanes probe the hash table; kPut: the first of the page's tasks to arrive inserts, the others pick the slot up), publishes its row range in shared memory and
// streams the rows it claims from it (128-bit loads and stores per lane, next group in flight while
// the current one is folded).  It reduces the four per-lane CRC words to the zero-init CRC of the
// segment, shifts it by the bytes that follow it in the page (GF(2) multiply) and XORs it into the
// job's accumulator.  Warps that run out of work take the back half of the longest remaining range
// in their CTA (finish times of equal streams differ by ~10 % on a saturated memory system).  The
// warp that completes a job's last row applies init/xorout and verifies or records the page CRC.
// (kServe is compiled for at most 96 registers per thread -- bounds of 640 threads, launched with 512 -- so that a serve
// CTA leaves room on its SM for the small post and wait kernels of the push transport, whatever SMs the serve kernels of
// two consecutive rounds occupy between them.)
template <bool kCrc, int kThreads, int G, bool kGet, bool kPut = false, bool kServe = false>
__global__ void __launch_bounds__(kServe ? 640 : kThreads, 1) copy_kernel(CopyArgs a) {
  static_assert(!(kGet && kPut), "a launch is a get, a put, or a list of prepared jobs");
  static_assert(!kServe || (kGet && kCrc), "the serve kernel is a get with the checksum on");
  constexpr bool kIndep = kCrc && G == 4 && !kServe;  // fold the four rows of a group independently (stream_rows)
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr uint32_t nwarps = kThreads / 32;
  __shared__ unsigned long long s_range[nwarps];  // published ranges: open to stealing
  __shared__ unsigned long long s_priv[nwarps];   // ranges of tasks that are not a warp's last: private
  __shared__ Job s_job[nwarps];
  __shared__ uint32_t s_jobidx[nwarps];
  __shared__ unsigned long long s_cnt[4];  // the CTA's share of the store's counters
  __shared__ uint32_t s_warps_out;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (a.timeline && threadIdx.x == 0) a.timeline[blockIdx.x * 4 + 0] = globaltimer_ns();
  // Programmatic dependent launch: let the next kernel in the stream be scheduled as soon as SMs free
  // up, and wait here until everything launched before this kernel has completed and is visible.
  asm volatile("griddepcontrol.launch_dependents;");
  // Dependency gate.  Normally a kernel waits right here.  When the host knows that the previous kernel in
  // the stream was another get of this store (a.early), nothing this kernel READS can change under it -- the
  // requests, the table and the slab -- so every warp resolves its first page and puts its first loads in
  // flight before it waits: the three dependent memory round trips of a task's start-up (request, bucket,
  // first rows) then overlap the previous kernel's tail instead of following it.  Everything this kernel
  // WRITES (pages, results, scratch, the rank counters) stays behind the gate.
  bool dep_ok = !((kGet || kPut) && a.early);
  if (dep_ok) asm volatile("griddepcontrol.wait;" ::: "memory");
  // kServe: the batch is what the peers have posted.  Lane r of every warp waits for requester r's posting of this
  // sequence number and keeps its request count and the exclusive prefix; the job split follows from the total.
  uint32_t srv_cnt = 0, srv_pre = 0;
  bool srv_lost = false;  // this lane's requester never posted: it gets no done word, so its own wait reports the time-out
  if (kServe) {
    const PushServe& ps = *a.push;
    if (lane < ps.n_shards) {
      const volatile unsigned long long* word = (const volatile unsigned long long*)(
          ps.inbox + ((size_t)(a.push_seq % ps.depth) * ps.n_shards + lane) * push_inbox_stride(ps.cap));
      const unsigned long long t0 = globaltimer_ns();
      unsigned long long w = *word;
      while ((uint32_t)(w >> 32) != a.push_seq) {
        if (globaltimer_ns() - t0 > ps.timeout_ns) {  // the peer never posted (it failed, or skipped a collective call)
          *a.error_flag = 3;
          w = (unsigned long long)a.push_seq << 32;
          srv_lost = true;
          break;
        }
        __nanosleep(100);
        w = *word;
      }
      srv_cnt = min((uint32_t)w, ps.cap);
    }
    __threadfence_system();  // acquire: the requests were written before the posting word
    uint32_t incl = srv_cnt;
#pragma unroll
    for (int o = 1; o < (int)kPushMaxShards; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= (uint32_t)o) incl += up;
    }
    srv_pre = incl - srv_cnt;
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, kPushMaxShards - 1);
    a.n_jobs = total;
    uint32_t tpj = max(1u, a.serve_warps / max(total, 1u));
    tpj = min(tpj, min(kMaxTasksPerJob, max(1u, a.serve_rows / 8u)));
    a.rows_per_task = (a.serve_rows + tpj - 1u) / tpj;
    a.tasks_per_job = (a.serve_rows + a.rows_per_task - 1u) / a.rows_per_task;
  }
  const uint32_t n_items = a.n_jobs * a.tasks_per_job;
  // Task assignment.  n_items / grid tasks per CTA are fixed by blockIdx, so warps can start at once.  The
  // n_items % grid surplus tasks go to the CTAs with the lowest RANK: CTAs on fast SMs count up from 0,
  // CTAs on slow SMs count down from the top (unique whatever the placement), so the surplus lands on the
  // fast SMs.  Only the warp that would take the surplus needs the rank; it asks for it now and looks at
  // the answer when it gets there.
  const uint32_t base_cnt = n_items / gridDim.x, extra = n_items % gridDim.x;
  const uint32_t surplus_warp = base_cnt % nwarps;
  uint32_t rank_reg = 0xFFFFFFFFu;
  auto ask_rank = [&]() {  // the counters are re-armed by the previous launch's last CTA: only behind the gate
    if (warp == surplus_warp && lane == 0) {
      uint32_t smid;
      asm("mov.u32 %0, %%smid;" : "=r"(smid));
      const bool slow = a.sm_slow[smid & 255u] != 0;
      rank_reg = slow ? (gridDim.x - 1u - atomicAdd(&a.sched[1], 1u)) : atomicAdd(&a.sched[0], 1u);
    }
  };
  const uint8_t* pf_ptr = nullptr;  // rows this warp will load right after the gate ...
  uint32_t pf_lines = 0;            // ... as a count of 128-byte lines
  auto gate = [&]() {
    if (!dep_ok) {
      // while this warp waits for its predecessor, pull its next rows into L2: DRAM is under-used during the
      // previous kernel's tail, and the loads that follow the gate then hit L2.  (Read-ahead of the same kind
      // INSIDE the streaming loop was measured too and only hurts: 0.80 at 32 rows ahead.)
      for (uint32_t l = lane; l < pf_lines; l += 32u) asm volatile("prefetch.global.L2 [%0];" ::"l"(pf_ptr + (size_t)l * 128u));
      asm volatile("griddepcontrol.wait;" ::: "memory");
      dep_ok = true;
      if (kServe && a.timeline && warp == 0 && lane == 0) a.timeline[blockIdx.x * 4 + 1] = globaltimer_ns();  // gate passed
      ask_rank();
    }
  };
  if (dep_ok) ask_rank();
  if (lane == 0) s_range[warp] = 0;
  if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0;  // (every path below reaches a __syncthreads before a warp can get to the end)
  if (threadIdx.x == 4) s_warps_out = 0;
  __shared__ __align__(8) unsigned long long s_tables_bar;
  uint32_t tables_bar = 0;
  if (kCrc) {
    uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_raw);
    uint32_t dyn;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
    if (base > kSmemP || base + dyn < kSmemEnd) {
      if (threadIdx.x == 0) *a.error_flag = 1;
      return;
    }
    if (kIndep) {  // the plain one-row tables (4 KiB, L2-resident), visible after the barrier both branches below end with
      const uint32_t* gfold1 = &a.tables->fold[0][0];
      for (uint32_t idx = threadIdx.x; idx < 1024; idx += kThreads) sts_u32(kSmemP + idx * 4u, __ldg(&gfold1[idx]));
    }
    if (a.table_image) {
      // bulk-async copies of the ready-made table image; nobody waits here
      tables_bar = (uint32_t)__cvta_generic_to_shared(&s_tables_bar);
      if (threadIdx.x == 0) {
        mbar_init(tables_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        constexpr uint32_t kImage = kSmemEnd - kSmemG, kPiece = kImage / 4;
        mbar_arrive_expect_tx(tables_bar, kImage);
#pragma unroll
        for (uint32_t q = 0; q < 4; ++q) bulk_load(kSmemG + q * kPiece, a.table_image + q * kPiece, kPiece, tables_bar);
      }
    } else {
      // Each thread fetches its table entries up front (coalesced, independent loads); the warp then
      // broadcasts entry by entry and every lane writes its own replica.
      const uint32_t* gfold = kIndep ? &a.tables->fold4[0][0] : &a.tables->fold[0][0];
      const uint32_t* gz4 = &a.tables->z4[0][0];
      constexpr uint32_t kPerWarp = 1024 / nwarps;  // fold entries per warp
      static_assert(1024 % nwarps == 0 && kPerWarp % 32 == 0, "fold table split");
      uint32_t mine[kPerWarp / 32], zmine[(1024 + kThreads - 1) / kThreads];
#pragma unroll
      for (uint32_t q = 0; q < kPerWarp / 32; ++q) mine[q] = __ldg(&gfold[warp * kPerWarp + q * 32 + lane]);
#pragma unroll
      for (uint32_t q = 0; q < (1024 + kThreads - 1) / kThreads; ++q) {
        const uint32_t idx = threadIdx.x + q * kThreads;
        zmine[q] = idx < 1024 ? __ldg(&gz4[idx]) : 0u;
      }
#pragma unroll
      for (uint32_t q = 0; q < kPerWarp / 32; ++q) {
#pragma unroll
        for (uint32_t k = 0; k < 32; ++k) {
          const uint32_t idx = warp * kPerWarp + q * 32 + k, j = idx >> 8, b = idx & 255u;
          sts_u32(kSmemF + (j >> 1) * 0x10000u + b * 256u + (j & 1u) * 128u + lane * 4u, __shfl_sync(0xFFFFFFFFu, mine[q], k));
        }
      }
#pragma unroll
      for (uint32_t q = 0; q < (1024 + kThreads - 1) / kThreads; ++q) {
        const uint32_t idx = threadIdx.x + q * kThreads;
        if (idx < 1024) sts_u32(kSmemG + idx * 4u, zmine[q]);
      }
      __syncthreads();
    }
  } else {
    __syncthreads();
  }
  if (!kServe && a.timeline && threadIdx.x == 0) a.timeline[blockIdx.x * 4 + 1] = globaltimer_ns();
  uint32_t put_pages = 0, put_tombs_gone = 0;  // kPut: this warp's share of the store's counters (lane 0)
  unsigned long long put_bytes = 0;             // (a two's complement sum: a replaced page may shrink)
  uint32_t rec_early = 0;  // kGet: the recorded CRC of the page this warp resolved last (lane 0)
  uint32_t cnt_lookups = 0, cnt_probes = 0;  // kGet: this warp's share of the store's counters (lane 0)
  unsigned long long cnt_bytes = 0;
  uint32_t local = warp;  // this warp takes the CTA's fixed tasks local, local + nwarps, ...
  bool surplus_pending = warp == surplus_warp && extra > 0;

  uint32_t gen = 0;
  for (;;) {
    // ---- pick the next segment: a static task, or (when those are exhausted) half of a neighbour's rows
    Job j;
    uint32_t job, row_lo, row_hi;
    bool tail_mine = false;  // task 0 of a job also owns the sub-row tail (n % 512 bytes)
    bool publish = false;
    uint32_t item = 0xFFFFFFFFu;
    bool last_item = false;
    bool report = false;  // kGet, task 0 of a page that needs no finisher: its result is written behind the gate
    int32_t report_status = 0;
    if (local < base_cnt) {
      // kServe with serve_flags & 2 deals the tasks out round robin: the job list is ordered by requester, and the requester that is this GPU
      // itself has local (fast) destinations -- in blocks, the CTAs holding those would finish in a fraction of the time
      // and leave NVLink to the others (measured at 2 GPUs: CTA end times from 96 to 219 us after the gate)
      item = (kServe && (a.serve_flags & 2u)) ? local * gridDim.x + blockIdx.x : blockIdx.x * base_cnt + local;
      local += nwarps;
      last_item = local >= base_cnt && !surplus_pending;
    } else if (surplus_pending) {
      surplus_pending = false;
      gate();
      const uint32_t rank = __shfl_sync(0xFFFFFFFFu, rank_reg, 0);
      if (rank < extra) {
        item = base_cnt * gridDim.x + rank;
        last_item = true;
      }
    }
    if (item != 0xFFFFFFFFu) {
      job = item / a.tasks_per_job;
      const uint32_t c = item - job * a.tasks_per_job;
      if (kGet) {
        // every task of a page probes for itself (the bucket is an L2 hit after the first); task 0 reports
        aps_get_req rq;
        uint32_t srv_tag = 0;  // kServe: 0x80000000 | requester << 24 | index in the requester's batch
        if (kServe) {
          const PushServe& ps = *a.push;
          const uint32_t owner_mask = __ballot_sync(0xFFFFFFFFu, lane < ps.n_shards && srv_pre <= job && srv_cnt != 0);
          const uint32_t r = 31u - (uint32_t)__clz((int)owner_mask);  // the last requester whose range starts at or before job
          const uint32_t i = job - __shfl_sync(0xFFFFFFFFu, srv_pre, r);
          const PushReq* q = reinterpret_cast<const PushReq*>(
              ps.inbox + ((size_t)(a.push_seq % ps.depth) * ps.n_shards + r) * push_inbox_stride(ps.cap) + sizeof(PushInboxHead)) + i;
          const uint4 q0 = __ldcg(reinterpret_cast<const uint4*>(q)), q1 = __ldcg(reinterpret_cast<const uint4*>(q) + 1),
                      q2 = __ldcg(reinterpret_cast<const uint4*>(q) + 2);
          rq.key.file_hi = (uint64_t)q0.x | ((uint64_t)q0.y << 32);
          rq.key.file_lo = (uint64_t)q0.z | ((uint64_t)q0.w << 32);
          rq.key.page_index = (uint64_t)q1.x | ((uint64_t)q1.y << 32);
          rq.page_off = q1.z;
          rq.nbytes = q1.w;
          rq.dst = ps.landing[r] + ((uint64_t)q2.x | ((uint64_t)q2.y << 32));
          srv_tag = 0x80000000u | (r << 24) | (q2.z & 0x00FFFFFFu);
        } else {
          rq = a.reqs[job];
        }
        Probe pr{0, 0, 0, 0};
        int32_t status = 0;
        for (;;) {
          // sharded cache: the page lives on GPU shard_of(key); its table and slab are read in place over
          // NVLink (P2P), so lookup, gather and CRC stay one kernel on the requesting GPU
          const Table tb = a.n_shards > 1 ? a.shards[shard_of(rq.key, a.n_shards)] : a.table;
          pr = Probe{0, 0, 0, 0};
          if (lane < 8) pr = tile_lookup(tb, rq.key, lane, 0);
          __syncwarp();
          const uint64_t meta = __shfl_sync(0xFFFFFFFFu, pr.meta, 0);
          status = 0;
          j = Job{nullptr, (uint8_t*)rq.dst, 0, 0, 0, kServe ? srv_tag : 0u};
          if (!meta) {
            status = -APS_ENOTFOUND;
          } else if (rq.page_off > meta_len(meta)) {  // LocalPageStore.java:99-100
            status = -APS_EINVAL;
          } else {
            const uint32_t plen = meta_len(meta);
            j.slot = meta_slot(meta);
            j.n = min(plen - rq.page_off, rq.nbytes);  // LocalPageStore.java:109-110
            j.src = tb.slab + (uint64_t)j.slot * tb.page_size + rq.page_off;
            j.pad = kServe ? srv_tag : (a.n_shards > 1 ? shard_of(rq.key, a.n_shards) + 1u : 0u);  // which shard's slot_crc to check
            j.flags = (a.crc_on && rq.page_off == 0 && j.n == plen && plen > 0) ? kJobVerify : 0u;
            status = (int32_t)j.n;
            if (c == 0 && lane == 0 && tb.slot_ref) tb.slot_ref[j.slot] = 1;  // CLOCK: referenced (a store that commutes)
            // the CRC recorded at put, asked for now and looked at when the page is done (same warp: see `whole`)
            if (kCrc && !kServe && lane == 0 && (j.flags & kJobVerify)) rec_early = __ldcg(&tb.slot_crc[j.slot]);
          }
          if (dep_ok || a.trust_reqs) break;
          // Early, but the request array is not promised to be final yet (no APS_FLAG_STABLE_REQS): use what was read
          // only to warm L2 with the rows this task will want, wait for the predecessor, then read the request again.
          // If it is unchanged the resolution above stands; otherwise it is redone, now behind the gate.
          if (j.n && a.n_shards <= 1) {
            const uint32_t rows_full = j.n / kRowBytes, lo = min(rows_full, c * a.rows_per_task);
            const uint32_t cnt = min(min(rows_full, lo + a.rows_per_task) - lo, (uint32_t)a.early - 1u + (uint32_t)G);
            const uint8_t* p0 = j.src + (size_t)lo * kRowBytes;
            for (uint32_t l = lane; l < cnt * (kRowBytes / 128u); l += 32u)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(p0 + (size_t)l * 128u));
          }
          gate();
          const aps_get_req again = a.reqs[job];
          const bool same = again.key.file_hi == rq.key.file_hi && again.key.file_lo == rq.key.file_lo &&
                            again.key.page_index == rq.key.page_index && again.page_off == rq.page_off &&
                            again.nbytes == rq.nbytes && again.dst == rq.dst;
          if (same) break;
          rq = again;
        }
        if (c == 0 && lane == 0) {
          // counted in registers and added to the store's counters once per warp, at the end: three atomics per PAGE on
          // the same three addresses serialise in L2 (65,536 pages of 4 KiB per step: ~60 us of a 265 us step)
          ++cnt_lookups;
          cnt_probes += pr.nbuckets;
          cnt_bytes += (unsigned long long)j.n;
          report = !kCrc || j.n == 0;  // otherwise the finisher reports
          report_status = status;
        }
        report = __shfl_sync(0xFFFFFFFFu, (int)report, 0) != 0;
      } else if (kPut) {
        // Fused insert (LocalPageStore.put's "create the file").  The page's tasks start at the same moment on different
        // warps, possibly in different CTAs.  Whichever of them ARRIVES FIRST takes the page's hand-over word
        // (a.jslot[job]: stale epoch -> this epoch, value 0) with a compare-and-swap, claims a table entry and a slab
        // slot with the first eight lanes and publishes the outcome; the others pick it up from the word.  A waiting warp
        // therefore only ever waits for a warp that is already running (the winner never waits for anybody), so the
        // kernel makes progress whatever part of its grid is resident -- a second kernel holding SMs (another store's put
        // on the same device) cannot deadlock it.
        aps_put_req rq = a.preqs[job];
        if (!dep_ok) {
          // A put cannot touch the table before its predecessor is done, but it can warm L2 with the source rows this
          // task will copy (a prefetch of whatever address the possibly unfinished request names is harmless), then
          // wait, then read the request for good.
          if (rq.src && rq.len <= a.table.page_size) {
            const uint32_t rows_full = rq.len / kRowBytes, lo = min(rows_full, c * a.rows_per_task);
            const uint32_t cnt = min(min(rows_full, lo + a.rows_per_task) - lo, (uint32_t)a.early - 1u + (uint32_t)G);
            const uint8_t* p0 = (const uint8_t*)rq.src + (size_t)lo * kRowBytes;
            for (uint32_t l = lane; l < cnt * (kRowBytes / 128u); l += 32u)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(p0 + (size_t)l * 128u));
          }
          gate();
          rq = a.preqs[job];
        }
        const Table& t = a.table;
        unsigned long long word = 0;
        int inserter = 0;
        if (lane == 0) {
          const unsigned long long mine = (unsigned long long)a.epoch << 32;  // this launch, outcome pending
          word = *(volatile unsigned long long*)&a.jslot[job];
          while ((uint32_t)(word >> 32) != a.epoch) {
            const unsigned long long seen = atomicCAS(&a.jslot[job], word, mine);
            if (seen == word) { inserter = 1; word = mine; break; }
            word = seen;
          }
        }
        inserter = __shfl_sync(0xFFFFFFFFu, inserter, 0);
        if (inserter) {
          int err = 0;
          uint32_t slot = kNoSlot;
          if (rq.len > t.max_len) {
            err = APS_EINVAL;
          } else {
            EntryClaim cl{kClaimFull, 0, 0};
            if (lane < 8) cl = tile_claim(t, rq.key, lane, 0);
            __syncwarp();
            cl.kind = __shfl_sync(0xFFFFFFFFu, cl.kind, 0);
            if (cl.kind == kClaimFull) {
              err = APS_ENOSPC;
            } else if (lane == 0) {
              Entry* e = t.entries + cl.entry_idx;
              if (cl.kind == kClaimReplace) {  // PageStore.put overwrites: same slot, new length
                slot = meta_slot(cl.was);
                e->meta = meta_make(slot, rq.len);
                if (t.slot_ref) t.slot_ref[slot] = 1;
                put_bytes += (unsigned long long)rq.len - (unsigned long long)meta_len(cl.was);
              } else {
                const int idx = atomicSub(t.free_top, 1) - 1;
                if (idx < 0) {
                  atomicAdd(t.free_top, 1);
                  e->meta = cl.was;  // no slab slot left: give the entry back
                } else {
                  slot = t.free_slots[idx];
                  e->k0 = rq.key.file_hi; e->k1 = rq.key.file_lo; e->k2 = rq.key.page_index;
                  __threadfence();
                  e->meta = meta_make(slot, rq.len);
                  if (t.slot_entry) {
                    t.slot_entry[slot] = cl.entry_idx;
                    t.slot_ref[slot] = 1;
                  }
                  ++put_pages;  // counted in registers, added once per warp at the end (see the get's counters)
                  put_bytes += (unsigned long long)rq.len;
                  if (cl.was == kMetaTomb) ++put_tombs_gone;
                }
              }
            }
            slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
            if (!err && slot == kNoSlot) err = APS_ENOSPC;
          }
          word = ((unsigned long long)a.epoch << 32) | (err ? (0x80000000u | (uint32_t)err) : slot + 1u);
          if (lane == 0) {
            if (!err) {
              if (rq.len == 0) t.slot_crc[slot] = 0;  // CRC32C of the empty string
              else cnt_bytes += (unsigned long long)rq.len;
            }
            __threadfence();
            *(volatile unsigned long long*)&a.jslot[job] = word;
            report = err != 0 || !kCrc || rq.len == 0;  // otherwise the finisher reports
            report_status = err ? -err : (int32_t)rq.len;
          }
          report = __shfl_sync(0xFFFFFFFFu, (int)report, 0) != 0;
        } else {
          if (lane == 0) {
            while ((uint32_t)word == 0u) word = *(volatile unsigned long long*)&a.jslot[job];  // the winner is running
            __threadfence();
          }
          word = __shfl_sync(0xFFFFFFFFu, word, 0);
        }
        const uint32_t v = (uint32_t)word;
        j = Job{nullptr, nullptr, 0, 0, 0, 0};
        if (!(v & 0x80000000u)) {
          j.src = (const uint8_t*)rq.src;
          j.slot = v - 1u;
          j.dst = t.slab + (uint64_t)j.slot * t.page_size;
          j.n = rq.len;
          j.flags = a.crc_on ? kJobRecord : 0u;
        }
      } else {
        j = a.jobs[job];
      }
      const uint32_t rows_full = j.n / kRowBytes;
      row_lo = min(rows_full, c * a.rows_per_task);
      row_hi = min(rows_full, row_lo + a.rows_per_task);
      tail_mine = c == 0 && (j.n % kRowBytes) != 0;
      if (report) {
        gate();
        if (kServe) {
          push_deliver(*a.push, a.push_seq, j.pad, aps_result{report_status, 0}, lane, srv_cnt);
        } else if (lane == 0) {
          a.res[job] = aps_result{report_status, 0};
        }
      }
      if (row_lo >= row_hi && !tail_mine) continue;
      publish = last_item;  // only a warp's final task is open to stealing (its descriptor stays valid)
    } else {
      if (!a.steal) break;
      // ---- steal: lanes look at one neighbour each, the warp agrees on the longest remaining range
      bool got = false;
      for (int attempt = 0; attempt < 8 && !got; ++attempt) {
        unsigned long long w = 0;
        if (lane < nwarps) w = *(volatile unsigned long long*)&s_range[lane];
        uint32_t rem = range_hi(w) > range_lo(w) ? range_hi(w) - range_lo(w) : 0u;
        uint32_t best = rem, who = lane;
#pragma unroll
        for (int o = 16; o; o >>= 1) {
          const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, best, o), ow = __shfl_xor_sync(0xFFFFFFFFu, who, o);
          if (ob > best || (ob == best && ow < who)) { best = ob; who = ow; }
        }
        if (best < kStealMinGroups * G) break;
        const unsigned long long vw = __shfl_sync(0xFFFFFFFFu, w, who);
        // the descriptor is read BEFORE the swap: if the swap succeeds the word (and its generation)
        // did not change in between, so the descriptor read here belongs to that range
        const Job vj = s_job[who];
        const uint32_t vjob = s_jobidx[who];
        __threadfence_block();
        const uint32_t take = max((uint32_t)G, ((best / 2) / G) * G);
        const uint32_t new_hi = range_hi(vw) - take;
        int ok = 0;
        if (lane == 0)
          ok = atomicCAS(&s_range[who], vw, range_pack(range_lo(vw), new_hi, range_gen(vw))) == vw;
        ok = __shfl_sync(0xFFFFFFFFu, ok, 0);
        if (ok) {
          j = vj;
          job = vjob;
          row_lo = new_hi;
          row_hi = range_hi(vw);
          got = true;
        }
      }
      if (!got) break;
      publish = true;
    }

    // ---- run the segment [row_lo, row_hi) of job `job` (+ its sub-row tail when tail_mine)
    const bool aligned = (((uintptr_t)j.src | (uintptr_t)j.dst) & 15u) == 0;
    uint32_t crc = 0, rows_done = 0, end_byte = row_lo * kRowBytes;
    if (row_lo < row_hi) {
      if (aligned) {
        ++gen;
        unsigned long long* rp = publish ? &s_range[warp] : &s_priv[warp];
        if (lane == 0) {
          if (publish) {
            s_job[warp] = j;
            s_jobidx[warp] = job;
            __threadfence_block();
          }
          *(volatile unsigned long long*)rp = range_pack(row_lo, row_hi, gen);
        }
        __syncwarp();
        uint32_t row_end = row_lo;
        if (!dep_ok) {
          const uint32_t ahead = row_hi - row_lo > (uint32_t)G ? min(row_hi - row_lo - (uint32_t)G, (uint32_t)a.early - 1u) : 0u;
          pf_ptr = j.src + (size_t)(row_lo + G) * kRowBytes;
          pf_lines = ahead * (kRowBytes / 128u);
        }
        // (Warming L2 with the NEXT page's hash bucket once this page's first loads are in flight -- so that small pages find
        // their bucket in L2 -- was measured and does not pay: 16 KiB pages 0.787 -> 0.779, 4 KiB 0.46 -> 0.44.)
        crc = stream_rows<kCrc, G, kIndep>(j.src, j.dst, rp, row_lo, lane, &row_end, tables_bar, gate);
        rows_done = row_end - row_lo;
        end_byte = row_end * kRowBytes;
      } else {
        gate();
        const uint32_t nb = (row_hi - row_lo) * kRowBytes;
        crc = warp_bytes<kCrc>(j.src + (size_t)row_lo * kRowBytes, j.dst ? j.dst + (size_t)row_lo * kRowBytes : nullptr, nb, lane);
        rows_done = row_hi - row_lo;
        end_byte = row_hi * kRowBytes;
      }
    }
    gate();  // a task made of a sub-row tail only has not passed it yet
    if (!kCrc) {
      if (tail_mine) {
        const uint32_t t0 = (j.n / kRowBytes) * kRowBytes;
        warp_bytes<false>(j.src + t0, j.dst ? j.dst + t0 : nullptr, j.n - t0, lane);
      }
      continue;
    }
    // shift the segment CRC to the end of the job and fold it into the job's accumulator
    const uint32_t after = j.n - end_byte;  // bytes of the job behind this segment (0: nothing to shift over)
    uint32_t contrib = rows_done ? (after ? gf_mul_call(crc, shift_for(a.rowshift, after)) : crc) : 0u;
    uint32_t units = rows_done;
    if (tail_mine) {  // the last n % 512 bytes: nothing follows them
      const uint32_t t0 = (j.n / kRowBytes) * kRowBytes;
      contrib ^= warp_bytes<true>(j.src + t0, j.dst ? j.dst + t0 : nullptr, j.n - t0, lane);
      units += 1;
    }
    const uint32_t target = (j.n + kRowBytes - 1) / kRowBytes;
    uint32_t prev = 0;
    // A warp that did the WHOLE job by itself (small pages: one task per page, nothing stolen) holds the job's zero-init
    // CRC already: no accumulator, no row count, none of their round trips to L2.  (Nobody else has rows of this job, so
    // the job's accumulator and counter stay untouched at zero.)
    // Measured (tools/bench_get_sizes.py, A/B on one box, fractions of the HBM peak): 64 KiB pages 0.845 -> 0.866, 16 KiB 0.70 ->
    // 0.77, 4 KiB 0.34 -> 0.45 -- once the per-page counter atomics were gone (with them, small pages got SLOWER with the
    // short cut: warps with nothing to wait for at the end of a page queued up behind the same three addresses in L2).
    const bool whole = !kServe && row_lo == 0 && units == target && !(a.serve_flags & 4u);
    if (kServe && !(a.serve_flags & 1u)) {  // every lane's stores into the requester's memory are ordered before the row count that announces them
      __threadfence_system();
      __syncwarp();
    }
    if (!whole) {
      if (lane == 0) {
        atomicXor(&a.acc[job], contrib);
        __threadfence();
        prev = atomicAdd(&a.done[job], units);
      }
      prev = __shfl_sync(0xFFFFFFFFu, prev, 0);
      if (prev + units != target) continue;
    }
    aps_result fin{0, 0};
    if (lane == 0) {  // this segment completed the job
      uint32_t x = contrib;
      if (!whole) {
        __threadfence();
        x = atomicExch(&a.acc[job], 0u);
        a.done[job] = 0;
      }
      const uint32_t init = (j.n == a.table.max_len && a.init_full) ? a.init_full : gf_mul_call(0xFFFFFFFFu, shift_for(a.rowshift, j.n));
      const uint32_t full = x ^ init ^ 0xFFFFFFFFu;
      bool corrupt = false;
      if (j.flags & kJobRecord) a.table.slot_crc[j.slot] = full;
      if (j.flags & kJobOutCrc) a.crc_out[job] = full;
      const uint32_t* recorded = (kGet && !kServe && j.pad) ? a.shards[j.pad - 1u].slot_crc : a.table.slot_crc;
      // (a page done by the warp that resolved it compares with the value that warp asked for at the start)
      if ((j.flags & kJobVerify) && ((kGet && !kServe && whole) ? rec_early : __ldcg(&recorded[j.slot])) != full) {
        corrupt = true;
        atomicAdd(&a.table.counters[kCtrCrcMismatch], 1ull);
      }
      fin = aps_result{corrupt ? -APS_ECORRUPT : (int32_t)j.n, full};
      if (!kServe && a.res) a.res[job] = fin;
    }
    if (kServe) push_deliver(*a.push, a.push_seq, j.pad, fin, lane, srv_cnt);
  }
  gate();  // a warp that found nothing to do arrives here without having waited
  // The warps' shares of the store's counters meet in shared memory; the last warp of the CTA to get here adds the CTA's
  // sums to the store: a few atomics per CTA and launch instead of a few per warp (2,368 warps reaching the same three
  // addresses at the very end of a launch cost the 8-GPU serve kernel 1.4 %).
  if ((kGet || kPut) && lane == 0) {
    if (kPut) {
      if (put_pages) atomicAdd(&s_cnt[0], (unsigned long long)put_pages);
      if (put_bytes) atomicAdd(&s_cnt[1], put_bytes);
      if (put_tombs_gone) atomicAdd(&s_cnt[2], (unsigned long long)put_tombs_gone);
    } else {
      if (cnt_lookups) atomicAdd(&s_cnt[0], (unsigned long long)cnt_lookups);
      if (cnt_probes) atomicAdd(&s_cnt[1], (unsigned long long)cnt_probes);
    }
    if (cnt_bytes) atomicAdd(&s_cnt[3], cnt_bytes);
    __threadfence_block();
    if (atomicAdd(&s_warps_out, 1u) == nwarps - 1u) {
      __threadfence_block();
      const unsigned long long c0 = s_cnt[0], c1 = s_cnt[1], c2 = s_cnt[2], c3 = s_cnt[3];
      if (kPut) {
        if (c0) atomicAdd(&a.table.counters[kCtrPages], c0);
        if (c1) atomicAdd(&a.table.counters[kCtrBytes], c1);
        if (c2) atomicAdd(&a.table.counters[kCtrTombs], 0ull - c2);
      } else {
        if (c0) atomicAdd(&a.table.counters[kCtrLookups], c0);
        if (c1) atomicAdd(&a.table.counters[kCtrProbes], c1);
      }
      if (c3) atomicAdd(&a.table.counters[kCtrBytesCopied], c3);
    }
  }
  if (kServe) {
    // a requester that asked this owner for nothing still waits for its done word (behind the gate: the previous round's
    // serve kernel has finished, so done words go up in round order)
    const PushServe& ps = *a.push;
    if (blockIdx.x == 0 && warp == 0 && lane < ps.n_shards && srv_cnt == 0 && !srv_lost) {
      __threadfence_system();
      *(volatile unsigned long long*)ps.done[lane] = a.push_seq;
    }
  }
  if (warp == surplus_warp && lane == 0) {
    // this CTA holds its rank (the answer was consumed above or is awaited here); the last CTA to say so
    // re-arms the counters for the next launch
    if (rank_reg != 0xFFFFFFFFu) __threadfence();
    if (atomicAdd(&a.sched[2], 1u) == gridDim.x - 1u) {
      a.sched[0] = 0;
      a.sched[1] = 0;
      a.sched[2] = 0;
    }
  }
  if (a.timeline) {
    if (lane == 0) a.timeline[1024 + blockIdx.x * 32 + warp] = globaltimer_ns();
    if (lane == 0 && (!kServe || warp == 0)) { if (kServe) a.timeline[blockIdx.x * 4 + 2] = globaltimer_ns(); else atomicMin(&a.timeline[blockIdx.x * 4 + 2], globaltimer_ns()); }
    __syncthreads();
    if (threadIdx.x == 0) {
      a.timeline[blockIdx.x * 4 + 3] = globaltimer_ns();
      uint32_t smid;
      asm("mov.u32 %0, %%smid;" : "=r"(smid));
      a.timeline[1024 + 256 * 32 + blockIdx.x] = smid;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K2 + K3, bulk-async (TMA) form.  Same job and CRC algebra as copy_kernel, different data movement:
// one producer thread per CTA walks the CTA's contiguous slice of the batch and issues
// cp.async.bulk global -> shared copies of 16-row (8 KiB) chunks into a ring of shared-memory stages,
// each guarded by a full/empty mbarrier pair; consumer warps take the stages round robin, start the
// cp.async.bulk shared -> global store of the chunk at once, fold its rows into the CRC straight from
// shared memory (conflict-free 128-bit reads), publish the chunk's shifted CRC, and hand the stage back
// when the store has finished reading it.  No data passes through registers on its way to the
// destination, and an SM keeps ~88 KiB of loads in flight whatever its warps are doing.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kStageRows = 16;
constexpr uint32_t kStageBytes = kStageRows * kRowBytes;
constexpr uint32_t kMaxStages = 27;
constexpr uint32_t kTmaThreads = 512;
constexpr uint32_t kTmaConsumers = 12;    // warps 1..12 consume; warp 0 produces
constexpr uint32_t kJobCache = 128;       // jobs a CTA's slice may span
constexpr uint32_t kDescTma = 1u, kDescTail = 2u, kDescEnd = 8u;

struct StageDesc {
  uint32_t job_local, row_lo, rows, flags;
};

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}

template <bool kCrc, bool kGet>
__global__ void __launch_bounds__(kTmaThreads, 1) tma_copy_kernel(CopyArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr uint32_t nwarps = kTmaThreads / 32;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  // ---- shared-memory map (absolute shared-window addresses)
  //   [base, +4 KiB)   job cache      [.., +NS*16) stage descriptors   [.., +NS*16) full/empty barriers
  //   stages of 8 KiB fill the rest below 0x0F000 (kCrc: the CRC tables sit at 0x0F000..0x30000) and
  //   everything above 0x30000; without CRC the whole window is stages.
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem_raw);
  uint32_t dyn;
  asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
  const uint32_t win_end = base + dyn;
  const uint32_t jobs_at = base, desc_at = jobs_at + kJobCache * sizeof(Job), bar_at = desc_at + kMaxStages * 16u;
  const uint32_t addr_at = bar_at + kMaxStages * 16u, stages_lo = (addr_at + kMaxStages * 4u + 127u) & ~127u;
  __shared__ uint32_t s_rank, s_ns;
  if (kCrc && (base > 0x2000u || win_end < kSmemEnd)) {
    if (threadIdx.x == 0) *a.error_flag = 1;
    return;
  }
  Job* s_jobs = reinterpret_cast<Job*>(smem_raw);
  StageDesc* s_desc = reinterpret_cast<StageDesc*>(smem_raw + (desc_at - base));
  uint32_t* s_stage_addr = reinterpret_cast<uint32_t*>(smem_raw + (addr_at - base));

  if (threadIdx.x == 0) {
    uint32_t smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    const bool slow = a.sm_slow[smid & 255u] != 0;
    s_rank = slow ? (gridDim.x - 1u - atomicAdd(&a.sched[1], 1u)) : atomicAdd(&a.sched[0], 1u);
    __threadfence();
    if (atomicAdd(&a.sched[2], 1u) == gridDim.x - 1u) {
      a.sched[0] = 0;
      a.sched[1] = 0;
      a.sched[2] = 0;
    }
    // lay the stages out
    uint32_t ns = 0;
    const uint32_t lo_end = kCrc ? kSmemG : win_end;
    for (uint32_t p = stages_lo; p + kStageBytes <= lo_end && ns < kMaxStages; p += kStageBytes) s_stage_addr[ns++] = p;
    if (kCrc)
      for (uint32_t p = kSmemEnd; p + kStageBytes <= win_end && ns < kMaxStages; p += kStageBytes) s_stage_addr[ns++] = p;
    // A stage must always come back to the SAME consumer warp (a warp that ran ahead to the stage's next
    // lap could otherwise pass a full-barrier whose current lap has not completed yet): keep the number of
    // stages a multiple of the number of consumers.
    {
      const uint32_t c = min(kTmaConsumers, ns);
      ns = c ? (ns / c) * c : 0;
    }
    s_ns = ns;
    for (uint32_t i = 0; i < ns; ++i) {
      mbar_init(bar_at + i * 16u, 1);       // full:  the producer's arrive (+ the bytes of the bulk load)
      mbar_init(bar_at + i * 16u + 8u, 1);  // empty: the consumer that drained the stage
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (kCrc) {
    const uint32_t* gfold = &a.tables->fold[0][0];
    const uint32_t* gz4 = &a.tables->z4[0][0];
    constexpr uint32_t kPerWarp = 1024 / nwarps;
    uint32_t mine[kPerWarp / 32], zmine[1024 / kTmaThreads];
#pragma unroll
    for (uint32_t q = 0; q < kPerWarp / 32; ++q) mine[q] = __ldg(&gfold[warp * kPerWarp + q * 32 + lane]);
#pragma unroll
    for (uint32_t q = 0; q < 1024 / kTmaThreads; ++q) zmine[q] = __ldg(&gz4[threadIdx.x + q * kTmaThreads]);
#pragma unroll
    for (uint32_t q = 0; q < kPerWarp / 32; ++q) {
#pragma unroll
      for (uint32_t k = 0; k < 32; ++k) {
        const uint32_t idx = warp * kPerWarp + q * 32 + k, j = idx >> 8, b = idx & 255u;
        sts_u32(kSmemF + (j >> 1) * 0x10000u + b * 256u + (j & 1u) * 128u + lane * 4u, __shfl_sync(0xFFFFFFFFu, mine[q], k));
      }
    }
#pragma unroll
    for (uint32_t q = 0; q < 1024 / kTmaThreads; ++q) sts_u32(kSmemG + (threadIdx.x + q * kTmaThreads) * 4u, zmine[q]);
  }
  __syncthreads();

  // ---- this CTA's contiguous slice of the chunk space, weighted by SM speed class
  const uint32_t cpj = a.tasks_per_job;  // chunks per job
  const uint64_t total = (uint64_t)a.n_jobs * cpj;
  const uint32_t rank = s_rank, ns = s_ns, grid = gridDim.x;
  const uint32_t n_slow = min(a.n_slow, grid), n_fast = grid - n_slow;
  const uint64_t units = 16ull * n_fast + 15ull * n_slow;
  auto cum = [&](uint32_t r) -> uint64_t {  // weight of the ranks below r: fast ranks first
    return r <= n_fast ? 16ull * r : 16ull * n_fast + 15ull * (r - n_fast);
  };
  const uint64_t lo = total * cum(rank) / units, hi = total * cum(rank + 1) / units;
  if (lo >= hi) return;
  const uint32_t jlo = (uint32_t)(lo / cpj), jhi = (uint32_t)((hi - 1) / cpj);
  if (jhi - jlo + 1 > kJobCache) {  // the host sizes batches so that this cannot happen
    if (threadIdx.x == 0) *a.error_flag = 2;
    return;
  }

  // ---- resolve the jobs of the slice (all warps, four jobs per warp at a time)
  for (uint32_t k = warp * 4 + (lane >> 3); jlo + (k & ~3u) <= jhi; k += nwarps * 4) {
    const uint32_t job = jlo + k;
    const uint32_t sub = lane & 7u, tshift = lane & 24u;
    if (job <= jhi) {  // tile-uniform
      Job j;
      if (kGet) {
        const aps_get_req rq = a.reqs[job];
        const Table tb = a.n_shards > 1 ? a.shards[shard_of(rq.key, a.n_shards)] : a.table;
        Probe pr = tile_lookup(tb, rq.key, sub, tshift);
        int32_t status = 0;
        j = Job{nullptr, (uint8_t*)rq.dst, 0, 0, 0, 0};
        if (!pr.meta) {
          status = -APS_ENOTFOUND;
        } else if (rq.page_off > meta_len(pr.meta)) {
          status = -APS_EINVAL;
        } else {
          const uint32_t plen = meta_len(pr.meta);
          j.slot = meta_slot(pr.meta);
          j.n = min(plen - rq.page_off, rq.nbytes);
          j.src = tb.slab + (uint64_t)j.slot * tb.page_size + rq.page_off;
          j.flags = (a.crc_on && rq.page_off == 0 && j.n == plen && plen > 0) ? kJobVerify : 0u;
          j.pad = a.n_shards > 1 ? shard_of(rq.key, a.n_shards) + 1u : 0u;
          status = (int32_t)j.n;
        }
        // the CTA whose slice holds the job's first chunk reports
        if (sub == 0 && (uint64_t)job * cpj >= lo) {
          atomicAdd(&a.table.counters[kCtrLookups], 1ull);
          atomicAdd(&a.table.counters[kCtrProbes], (unsigned long long)pr.nbuckets);
          if (j.n) atomicAdd(&a.table.counters[kCtrBytesCopied], (unsigned long long)j.n);
          if (!kCrc || j.n == 0) a.res[job] = aps_result{status, 0};
        }
      } else {
        j = a.jobs[job];
      }
      if (sub == 0) s_jobs[k] = j;
    }
  }
  __syncthreads();

  const uint32_t nc = min(kTmaConsumers, ns);
  if (warp == 0) {
    // ================================ producer: one thread
    if (lane == 0) {
      uint32_t q = 0;
      auto post = [&](const StageDesc& d, const Job* J) {
        const uint32_t st = q % ns, par = (q / ns) & 1u;
        mbar_wait(bar_at + st * 16u + 8u, par ^ 1u);  // stage drained (passes at once on the first lap)
        s_desc[st] = d;
        if (d.flags & kDescTma) {
          mbar_arrive_expect_tx(bar_at + st * 16u, d.rows * kRowBytes);
          bulk_load(s_stage_addr[st], J->src + (size_t)d.row_lo * kRowBytes, d.rows * kRowBytes, bar_at + st * 16u);
        } else {
          __threadfence_block();
          mbar_arrive(bar_at + st * 16u);
        }
        ++q;
      };
      for (uint64_t g = lo; g < hi; ++g) {
        const uint32_t job = (uint32_t)(g / cpj), c = (uint32_t)(g - (uint64_t)job * cpj);
        const Job* J = &s_jobs[job - jlo];
        const uint32_t rows_full = J->n / kRowBytes, tail = J->n % kRowBytes;
        const uint32_t job_rows = rows_full + (tail ? 1u : 0u);
        const uint32_t row_lo = c * kStageRows;
        if (row_lo >= job_rows) continue;
        const uint32_t rows = rows_full > row_lo ? min(kStageRows, rows_full - row_lo) : 0u;
        const bool aligned = (((uintptr_t)J->src | (uintptr_t)J->dst) & 15u) == 0;
        StageDesc d{job - jlo, row_lo, rows, 0u};
        if (aligned && rows) d.flags |= kDescTma;
        if (tail && row_lo + kStageRows >= job_rows) d.flags |= kDescTail;
        post(d, J);
      }
      for (uint32_t e = 0; e < nc; ++e) post(StageDesc{0, 0, 0, kDescEnd}, nullptr);
    }
    return;
  }
  if (warp > nc) return;

  // ================================ consumers: stage q goes to consumer q % nc
  const uint32_t rb = 0x00010000u | (lane << 2);
  for (uint32_t q = warp - 1;; q += nc) {
    const uint32_t st = q % ns, par = (q / ns) & 1u;
    mbar_wait(bar_at + st * 16u, par);
    const StageDesc d = s_desc[st];
    if (d.flags & kDescEnd) {
      if (lane == 0) mbar_arrive(bar_at + st * 16u + 8u);
      break;
    }
    const Job j = s_jobs[d.job_local];
    const uint32_t job = jlo + d.job_local;
    const uint32_t sbase = s_stage_addr[st];
    const size_t byte_lo = (size_t)d.row_lo * kRowBytes;
    uint32_t crc = 0;
    if (d.flags & kDescTma) {
      if (lane == 0 && j.dst) bulk_store(j.dst + byte_lo, sbase, d.rows * kRowBytes);  // starts at once; reads shared only
      if (kCrc) {
        uint32_t S0 = 0, S1 = 0, S2 = 0, S3 = 0;
        const uint32_t p = sbase + lane * 16u;
        uint32_t r = 0;
        for (; r + 4 <= d.rows; r += 4) {
          const uint4 w0 = lds_v4(p + (r + 0) * kRowBytes), w1 = lds_v4(p + (r + 1) * kRowBytes);
          const uint4 w2 = lds_v4(p + (r + 2) * kRowBytes), w3 = lds_v4(p + (r + 3) * kRowBytes);
          APS_FOLD_ROW(w0);
          APS_FOLD_ROW(w1);
          APS_FOLD_ROW(w2);
          APS_FOLD_ROW(w3);
        }
        for (; r < d.rows; ++r) {
          const uint4 w = lds_v4(p + r * kRowBytes);
          APS_FOLD_ROW(w);
        }
        const uint32_t t = z4_word(z4_word(z4_word(S0) ^ S1) ^ S2) ^ S3;
        crc = warp_xor(gf_mul_call(t, c_lane_shift[lane]));
      }
    } else if (d.rows) {  // unaligned job: bytes straight from global memory
      crc = warp_bytes<kCrc>(j.src + byte_lo, j.dst ? j.dst + byte_lo : nullptr, d.rows * kRowBytes, lane);
    }
    // the rows have been read (and the store has finished reading them): give the stage back before the
    // per-chunk arithmetic, so that its refill overlaps it
    if (lane == 0) {
      if (d.flags & kDescTma) bulk_store_wait_read();
      mbar_arrive(bar_at + st * 16u + 8u);
    }
    uint32_t contrib = 0, units = d.rows;
    if (kCrc && d.rows) contrib = gf_mul_call(crc, shift_for(a.rowshift, j.n - (uint32_t)(byte_lo + (size_t)d.rows * kRowBytes)));
    if (d.flags & kDescTail) {
      const uint32_t t0 = (j.n / kRowBytes) * kRowBytes;
      const uint32_t ct = warp_bytes<kCrc>(j.src + t0, j.dst ? j.dst + t0 : nullptr, j.n - t0, lane);
      if (kCrc) contrib ^= ct;
      units += 1;
    }
    if (!kCrc) continue;
    const uint32_t target = (j.n + kRowBytes - 1) / kRowBytes;
    if (lane == 0) {
      atomicXor(&a.acc[job], contrib);
      __threadfence();
      const uint32_t prev = atomicAdd(&a.done[job], units);
      if (prev + units == target) {
        __threadfence();
        const uint32_t x = atomicExch(&a.acc[job], 0u);
        const uint32_t full = x ^ gf_mul_call(0xFFFFFFFFu, shift_for(a.rowshift, j.n)) ^ 0xFFFFFFFFu;
        bool corrupt = false;
        if (j.flags & kJobRecord) a.table.slot_crc[j.slot] = full;
        if (j.flags & kJobOutCrc) a.crc_out[job] = full;
        const uint32_t* recorded = (kGet && j.pad) ? a.shards[j.pad - 1u].slot_crc : a.table.slot_crc;
        if ((j.flags & kJobVerify) && __ldcg(&recorded[j.slot]) != full) {
          corrupt = true;
          atomicAdd(&a.table.counters[kCtrCrcMismatch], 1ull);
        }
        if (a.res) a.res[job] = aps_result{corrupt ? -APS_ECORRUPT : (int32_t)j.n, full};
        a.done[job] = 0;
      }
    }
  }
  if (lane == 0) bulk_store_wait_all();  // the stores must have landed before the kernel is reported complete
}

#ifndef APS_KERNELS_TEMPLATES_ONLY  // a second translation unit wants the templates above, not these definitions
// ------------------------------------------------------------------------------------------------
// K1 kernels: eight lanes per request.
// ------------------------------------------------------------------------------------------------
constexpr int kK1Threads = 256;  // 32 tiles per block

__global__ void __launch_bounds__(kK1Threads) put_resolve_kernel(Table t, const aps_put_req* reqs, uint32_t n, Job* jobs,
                                                                 aps_result* res, int crc_on) {
  __shared__ BlockAcct acct;
  acct_init(acct);
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t i = gid >> 3, sub = gid & 7u, tshift = threadIdx.x & 24u, tmask = 0xFFu << tshift;
  const bool live = i < n;  // tile-uniform
  aps_put_req rq{};
  EntryClaim c{kClaimFull, 0, 0};
  int err = 0, rank = -1;
  if (live) {
    rq = reqs[i];
    if (rq.len > t.max_len) err = APS_EINVAL;
    else {
      c = tile_claim(t, rq.key, sub, tshift);
      if (c.kind == kClaimFull) err = APS_ENOSPC;
      if (c.kind == kClaimNew && sub == 0) rank = atomicAdd(&acct.need, 1);
    }
  }
  block_take_slots(acct, t);
  if (live) {
    Job jb{nullptr, nullptr, 0, 0, 0, 0};
    uint32_t slot = kNoSlot;
    if (!err && c.kind == kClaimReplace) {
      slot = meta_slot(c.was);
      if (sub == 0) {
        t.entries[c.entry_idx].meta = meta_make(slot, rq.len);
        if (t.slot_ref) t.slot_ref[slot] = 1;
        acct_add(acct, kCtrBytes, (unsigned long long)rq.len - (unsigned long long)meta_len(c.was));
      }
    } else if (!err) {
      if (sub == 0) {
        Entry* e = t.entries + c.entry_idx;
        if (rank < acct.avail) {
          slot = t.free_slots[acct.base - 1 - rank];
          e->k0 = rq.key.file_hi; e->k1 = rq.key.file_lo; e->k2 = rq.key.page_index;
          __threadfence();
          e->meta = meta_make(slot, rq.len);
          if (t.slot_entry) {
            t.slot_entry[slot] = c.entry_idx;
            t.slot_ref[slot] = 1;
          }
          acct_add(acct, kCtrPages, 1ull);
          acct_add(acct, kCtrBytes, (unsigned long long)rq.len);
          if (c.was == kMetaTomb) acct_add(acct, kCtrTombs, ~0ull);
        } else {
          e->meta = c.was;  // no slab slot left: give the entry back
        }
      }
      slot = __shfl_sync(tmask, slot, tshift);
      if (slot == kNoSlot) err = APS_ENOSPC;
    }
    if (sub == 0) {
      aps_result rs{err ? -err : (int32_t)rq.len, 0};
      if (!err) {
        jb = Job{(const uint8_t*)rq.src, t.slab + (uint64_t)slot * t.page_size, rq.len, slot, crc_on ? kJobRecord : 0u, 0};
        if (rq.len == 0) t.slot_crc[slot] = 0;  // CRC32C of the empty string
        else acct_add(acct, kCtrBytesCopied, (unsigned long long)rq.len);
      }
      jobs[i] = jb;
      res[i] = rs;
    }
  }
  acct_flush(acct, t);
}

// An erased entry in a bucket that still has an EMPTY entry goes straight back to EMPTY (no probe chain
// runs through such a bucket); otherwise it becomes a tombstone.
__global__ void __launch_bounds__(kK1Threads) delete_kernel(Table t, const aps_key* keys, uint32_t n, aps_result* res) {
  __shared__ BlockAcct acct;
  acct_init(acct);
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t i = gid >> 3, sub = gid & 7u, tshift = threadIdx.x & 24u;
  const bool live = i < n;
  Probe pr{0, 0, 0, 0};
  int rank = -1;
  if (live) {
    const aps_key key = keys[i];  // a copy: a reference into global memory is re-read on every probe
    pr = tile_lookup(t, key, sub, tshift);
    if (pr.meta && sub == 0) {
      const bool to_empty = pr.empties != 0;
      t.entries[pr.entry_idx].meta = to_empty ? kMetaEmpty : kMetaTomb;
      rank = atomicAdd(&acct.need, 1);
      acct_add(acct, kCtrPages, ~0ull);
      acct_add(acct, kCtrBytes, 0ull - (unsigned long long)meta_len(pr.meta));
      if (!to_empty) acct_add(acct, kCtrTombs, 1ull);
    }
  }
  block_return_slots(acct, t);
  if (live && sub == 0) {
    if (rank >= 0 && t.slot_entry) t.slot_entry[meta_slot(pr.meta)] = kNoEntry;
    if (rank >= 0) t.free_slots[acct.base + rank] = meta_slot(pr.meta);
    res[i] = aps_result{pr.meta ? 0 : -APS_ENOTFOUND, 0};
  }
  acct_flush(acct, t);
}

// Membership only.  A lookup needs no agreement on an entry (inserts and erases do, hence their eight-lane tiles),
// so a key gets kLookupLanes lanes that read 8 / kLookupLanes entries each: the per-warp instruction stream (hash,
// ballots, bookkeeping -- 65 % of the issue slots with eight lanes per key) is shared by more keys.
constexpr uint32_t kLookupLanes = 4;
constexpr uint32_t kLookupPerGroup = 8;  // keys a lane group handles one after the other (grid sizing)
// Each group of four lanes walks a strided list of keys.  A lookup is two dependent DRAM round trips -- the key, then its
// bucket -- so the next key is fetched while the current bucket is on its way (one exposed round trip per key instead of
// two), and the counters are summed in registers and reach shared memory once per lane group instead of once per key.
// (48 registers, 62 % occupancy; forcing 40 or 32 registers spills and measured 15.0 and 11.4 G lookups/s against 16.0.)
__device__ __forceinline__ void load_entry256(const Entry* e, uint64_t& k0, uint64_t& k1, uint64_t& k2, uint64_t& meta) {
  asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(k0), "=l"(k1), "=l"(k2), "=l"(meta) : "l"(e));  // one 256-bit access
}
template <bool k256>
__global__ void __launch_bounds__(kK1Threads) lookup_kernel(Table t, const aps_key* keys, uint32_t n, int32_t* len_out) {
  __shared__ BlockAcct acct;
  acct_init(acct);
  constexpr uint32_t L = kLookupLanes, E = kBucketEntries / L;
  const uint32_t sub = threadIdx.x % L, gshift = (threadIdx.x & 31u) & ~(L - 1u), gmask = ((1u << L) - 1u) << gshift;
  const uint32_t groups = gridDim.x * (kK1Threads / L), group = blockIdx.x * (kK1Threads / L) + threadIdx.x / L;
  uint32_t n_done = 0, n_buckets = 0;
  aps_key key = group < n ? keys[group] : aps_key{0, 0, 0};
  for (uint32_t i = group; i < n; i += groups) {
    uint64_t b = hash_key(key) & t.bucket_mask;
    uint64_t k0[E], k1[E], k2[E], m[E];
#pragma unroll
    for (uint32_t q = 0; q < E; ++q) {
      if (k256) load_entry256(t.entries + b * kBucketEntries + sub * E + q, k0[q], k1[q], k2[q], m[q]);
      else load_entry(t.entries + b * kBucketEntries + sub * E + q, k0[q], k1[q], k2[q], m[q]);
    }
    const aps_key cur = key;
    if (i + groups < n) key = keys[i + groups];  // in flight together with the bucket
    uint64_t found = 0;
    uint32_t nb = 0;
    for (uint64_t p = 0; p <= t.bucket_mask; ++p) {
      if (p) {  // beyond the home bucket (about one lookup in a hundred at load 0.5)
#pragma unroll
        for (uint32_t q = 0; q < E; ++q) load_entry(t.entries + b * kBucketEntries + sub * E + q, k0[q], k1[q], k2[q], m[q]);
      }
      ++nb;
      uint64_t mine = 0;
      bool empty = false;
#pragma unroll
      for (uint32_t q = 0; q < E; ++q) {
        if ((m[q] >> 63) && k0[q] == cur.file_hi && k1[q] == cur.file_lo && k2[q] == cur.page_index) mine = m[q];
        empty |= m[q] == kMetaEmpty;
      }
      const uint32_t bm = (__ballot_sync(gmask, mine != 0) >> gshift) & ((1u << L) - 1u);
      const uint32_t be = (__ballot_sync(gmask, empty) >> gshift) & ((1u << L) - 1u);
      if (bm) {
        found = __shfl_sync(gmask, mine, gshift + (__ffs(bm) - 1));
        break;
      }
      if (be) break;
      b = (b + 1) & t.bucket_mask;
    }
    if (sub == 0) {
      ++n_done;
      n_buckets += nb;
      len_out[i] = found ? (int32_t)meta_len(found) : -1;
    }
  }
  if (sub == 0 && n_done) {
    acct_add(acct, kCtrLookups, (unsigned long long)n_done);
    acct_add(acct, kCtrProbes, (unsigned long long)n_buckets);
  }
  acct_flush(acct, t);
}

// Push transport, requester side (one small CTA each; they run next to the serve kernel, which leaves one SM free).
// push_post_kernel routes a batch: request i goes into the inbox that requester `me` has on the GPU owning its page,
// at a position taken from a per-owner counter, then one posting word per owner (sequence number << 32 | count)
// announces them.  A destination outside the landing buffer cannot be written by a peer: that request is answered
// -APS_EINVAL on the spot and not posted.
struct PushPostArgs {
  const aps_get_req* reqs;
  uint32_t n, n_shards, me, cap, depth, seq;
  uint8_t* peer_inbox[kPushMaxShards];  // owner o's inbox array as mapped here
  const uint8_t* landing;               // my landing buffer
  uint64_t landing_bytes;
  aps_result* my_results;               // my result slots [depth][cap]
};
__global__ void __launch_bounds__(256) push_post_kernel(PushPostArgs p) {
  __shared__ uint32_t s_cnt[kPushMaxShards];
  if (threadIdx.x < kPushMaxShards) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t slot = p.seq % p.depth;
  for (uint32_t i = threadIdx.x; i < p.n; i += blockDim.x) {
    const aps_get_req rq = p.reqs[i];
    const uint8_t* d = (const uint8_t*)rq.dst;
    if (rq.nbytes && (d < p.landing || d + rq.nbytes > p.landing + p.landing_bytes)) {
      p.my_results[(size_t)slot * p.cap + i] = aps_result{-APS_EINVAL, 0};
      continue;
    }
    const uint32_t owner = shard_of(rq.key, p.n_shards);
    const uint32_t pos = atomicAdd(&s_cnt[owner], 1u);
    uint4* q = reinterpret_cast<uint4*>(p.peer_inbox[owner] + ((size_t)slot * p.n_shards + p.me) * push_inbox_stride(p.cap) +
                                        sizeof(PushInboxHead) + (size_t)pos * sizeof(PushReq));
    const uint64_t off = rq.nbytes ? (uint64_t)(d - p.landing) : 0;
    q[0] = make_uint4((uint32_t)rq.key.file_hi, (uint32_t)(rq.key.file_hi >> 32), (uint32_t)rq.key.file_lo, (uint32_t)(rq.key.file_lo >> 32));
    q[1] = make_uint4((uint32_t)rq.key.page_index, (uint32_t)(rq.key.page_index >> 32), rq.page_off, rq.nbytes);
    q[2] = make_uint4((uint32_t)off, (uint32_t)(off >> 32), i, 0u);
  }
  __threadfence_system();  // every thread: its requests are out before the posting words
  __syncthreads();
  if (threadIdx.x < p.n_shards) {
    volatile unsigned long long* word = (volatile unsigned long long*)(
        p.peer_inbox[threadIdx.x] + ((size_t)slot * p.n_shards + p.me) * push_inbox_stride(p.cap));
    *word = ((unsigned long long)p.seq << 32) | s_cnt[threadIdx.x];
  }
}

// push_wait_kernel: the stream-ordered end of a pushed multi-get.  Waits until every owner has raised this GPU's done
// word for `seq` (the pages are then in the landing buffer), then copies the results into the caller's array.
__global__ void __launch_bounds__(256) push_wait_kernel(const unsigned long long* done, uint32_t n_shards, uint32_t seq,
                                                        const aps_result* slot_results, aps_result* res, uint32_t n, int* error_flag,
                                                        unsigned long long timeout_ns) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < n_shards) {
    const volatile unsigned long long* word = (const volatile unsigned long long*)&done[threadIdx.x];
    const unsigned long long t0 = globaltimer_ns();
    while (*word < (unsigned long long)seq) {  // sequence numbers only grow
      if (globaltimer_ns() - t0 > timeout_ns) {
        *error_flag = 4;
        s_ok = 0;
        break;
      }
      __nanosleep(200);
    }
  }
  __threadfence_system();
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
    aps_result r{-APS_ETIMEOUT, 0};
    if (s_ok) {
      const unsigned long long v = __ldcg(reinterpret_cast<const unsigned long long*>(&slot_results[i]));  // written by a peer: from L2
      r.len = (int32_t)(uint32_t)v;
      r.crc = (uint32_t)(v >> 32);
    }
    res[i] = r;
  }
}

// jobs for the stand-alone checksum pass
__global__ void crc_jobs_kernel(const void* const* data, const uint32_t* len, uint32_t n, uint32_t max_len, Job* jobs,
                                uint32_t* crc_out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t l = min(len[i], max_len);
  jobs[i] = Job{(const uint8_t*)data[i], nullptr, l, 0, kJobOutCrc, 0};
  if (l == 0) crc_out[i] = 0;
}

// Rebuild of the table without tombstones: every VALID entry of `old` is re-inserted (same slot, same
// length) into the zeroed table `t`.  One tile of eight lanes per old entry.
__global__ void rehash_kernel(Table t, const Entry* old, uint64_t n_old) {
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t i = gid >> 3;
  const uint32_t sub = (uint32_t)gid & 7u, tshift = threadIdx.x & 24u, tmask = 0xFFu << tshift;
  if (i >= n_old) return;
  uint64_t k0, k1, k2, meta;
  load_entry(old + i, k0, k1, k2, meta);
  if (!(meta >> 63)) return;  // tile-uniform: all eight lanes read the same entry
  const aps_key key{k0, k1, k2};
  uint64_t b = hash_key(key) & t.bucket_mask;
  for (uint64_t visited = 0; visited <= t.bucket_mask;) {  // a lost race re-reads the same bucket and is not a visit
    uint64_t e0, e1, e2, em;
    load_entry(t.entries + b * kBucketEntries + sub, e0, e1, e2, em);
    uint32_t be = (__ballot_sync(tmask, em == kMetaEmpty) >> tshift) & 0xFFu;
    int ok = 0;
    if (be) {
      const uint64_t cand = b * kBucketEntries + (__ffs(be) - 1);
      if (sub == 0) ok = atomicCAS((unsigned long long*)&t.entries[cand].meta, 0ull, (unsigned long long)kMetaLocked) == 0ull;
      ok = __shfl_sync(tmask, ok, tshift);
      if (ok) {
        if (sub == 0) {
          Entry* e = t.entries + cand;
          e->k0 = k0; e->k1 = k1; e->k2 = k2;
          __threadfence();
          e->meta = meta;
          if (t.slot_entry) t.slot_entry[meta_slot(meta)] = cand;
        }
        return;
      }
      continue;  // lost the entry to another key: look at the same bucket again
    }
    b = (b + 1) & t.bucket_mask;
    ++visited;
  }
}

__global__ void init_free_slots_kernel(uint32_t* free_slots, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) free_slots[i] = (uint32_t)(n - 1 - i);  // slot 0 is handed out first
}

// ------------------------------------------------------------------------------------------------
// Opt-in eviction on the device: CLOCK, the classic one-bit approximation of LRU.
// A get or put sets the slot's reference bit; the hand sweeps the slots in index order, clears set
// bits and evicts pages whose bit is already clear.  It cannot reproduce the reference's exact LRU /
// LFU victim order, so it is never used for the parity runs -- it is for caches whose accesses stay
// on the GPU, where a host-side policy (3.5 M accesses/s) would be the bottleneck.
//   ctl[0] = hand (slot index)  ctl[1] = cursor (slots handed out to CTAs in this sweep)
//   ctl[2] = need               ctl[3] = candidates seen  ctl[4] = CTAs finished   ctl[5] = evicted (result)
// ------------------------------------------------------------------------------------------------
__global__ void clock_need_kernel(Table t, uint32_t n_incoming, uint32_t n_fixed, uint32_t* ctl) {
  const int free_now = max(*t.free_top, 0);
  ctl[2] = n_fixed ? n_fixed : ((int)n_incoming > free_now ? n_incoming - (uint32_t)free_now : 0u);
  ctl[3] = 0;
  ctl[1] = 0;
  ctl[4] = 0;
}

__global__ void __launch_bounds__(256) clock_evict_kernel(Table t, uint32_t n_slots, uint32_t* ctl, aps_key* victims,
                                                          uint32_t victims_cap) {
  __shared__ uint32_t s_base;
  const uint32_t need = ctl[2], hand = ctl[0];
  if (need) {
    for (;;) {
      if (threadIdx.x == 0) {
        // stop handing out windows once enough victims exist, or after two full turns of the hand (the
        // first turn clears every bit, so the second finds a victim in every occupied slot)
        s_base = (*(volatile uint32_t*)&ctl[3] >= need) ? 0xFFFFFFFFu : atomicAdd(&ctl[1], 256u);
      }
      __syncthreads();
      const uint32_t base = s_base;
      __syncthreads();
      if (base == 0xFFFFFFFFu || base >= 2u * n_slots + 256u) break;
      const uint32_t slot = (uint32_t)(((uint64_t)hand + base + threadIdx.x) % n_slots);
      const uint64_t e = t.slot_entry[slot];
      if (e == kNoEntry) continue;
      if (t.slot_ref[slot]) {
        t.slot_ref[slot] = 0;
        continue;
      }
      // On a small store the windows of the second turn can overlap those of the first that are still in
      // flight: whoever swaps the owner out first evicts the page, the other sees a free slot.
      if (atomicCAS((unsigned long long*)&t.slot_entry[slot], (unsigned long long)e, (unsigned long long)kNoEntry) != e) continue;
      const uint32_t idx = atomicAdd(&ctl[3], 1u);
      if (idx >= need) {  // enough already: this page stays (its bit is clear, it goes first next time)
        t.slot_entry[slot] = e;
        continue;
      }
      Entry* ent = t.entries + e;
      const uint64_t meta = ent->meta;
      if (victims && idx < victims_cap) victims[idx] = aps_key{ent->k0, ent->k1, ent->k2};
      bool has_empty = false;
      const Entry* bucket = t.entries + (e & ~(uint64_t)(kBucketEntries - 1));
      for (uint32_t q = 0; q < kBucketEntries; ++q) has_empty |= __ldcg(&bucket[q].meta) == kMetaEmpty;
      ent->meta = has_empty ? kMetaEmpty : kMetaTomb;  // same rule as tile_erase
      t.free_slots[atomicAdd(t.free_top, 1)] = slot;
      atomicAdd(&t.counters[kCtrPages], ~0ull);
      atomicAdd(&t.counters[kCtrBytes], 0ull - (unsigned long long)meta_len(meta));
      atomicAdd(&t.counters[kCtrEvicted], 1ull);
      if (!has_empty) atomicAdd(&t.counters[kCtrTombs], 1ull);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl[4], 1u) == gridDim.x - 1u) {  // last CTA out: the hand moves past what was swept
      const uint32_t swept = min(ctl[1], 2u * n_slots);
      ctl[0] = (uint32_t)(((uint64_t)hand + swept) % n_slots);
      ctl[5] = min(ctl[3], need);  // pages actually evicted (candidates beyond `need` were left alone)
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Synthetic pages: word k of page (file, index) = splitmix64 output k for the seed
// seed ^ (file << 32) ^ index.  One 8-byte word per thread, grid-stride.
// ------------------------------------------------------------------------------------------------
__global__ void fill_pages_kernel(uint8_t* dst, uint64_t stride, uint32_t page_len, uint32_t n_pages, const uint64_t* files,
                                  const uint64_t* indices, uint64_t seed) {
  const uint64_t words_per_page = (page_len + 7u) / 8u;
  const uint64_t total = words_per_page * n_pages;
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t p = g / words_per_page, k = g - p * words_per_page;
    const uint64_t s = seed ^ (files[p] << 32) ^ indices[p];
    const uint64_t w = mix64(s + (k + 1) * 0x9E3779B97F4A7C15ull);
    uint8_t* out = dst + p * stride + k * 8;
    if (k * 8 + 8 <= page_len && (((uintptr_t)out) & 7u) == 0) {
      *reinterpret_cast<uint64_t*>(out) = w;
    } else {
      for (uint32_t b = 0; b < 8 && k * 8 + b < page_len; ++b) out[b] = (uint8_t)(w >> (8 * b));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K4: MD5 of many files at once (RFC 1321).  One file per lane; the chain inside a file is serial,
// so parallelism is across files only (ChecksumCommand.java:83-90 replaced file by file).  A warp
// stages 512 bytes of each of its 32 files per step with 16-byte cp.async (coalesced per file) into
// a padded shared tile (row stride 132 words, read back 128 bits at a time without bank conflicts).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kMd5Seg = 2048;  // bytes of every file staged per step (round 1: 512; the staging loop is per step, so 4x fewer)
constexpr uint32_t kMd5DynSmem = 2u * 32u * (kMd5Seg / 4 + 4) * 4u;  // two buffers of 32 rows (132 KiB): dynamic shared memory
constexpr uint32_t kMd5RowWords = kMd5Seg / 4 + 4;  // rows of seg + 16 bytes: 16-byte aligned, and a quarter-warp of
                                                    // 128-bit reads (lane f at f*528) covers all 32 banks

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int s) { return __funnelshift_l(x, x, s); }

// The chain through a step is b -> F(b,c,d) -> (+) -> rotate -> (+b).  a + m + k does not depend on the
// previous step's result, so it is formed first and kept off that chain (3 dependent ops per step).
#define APS_MD5_STEP(f, a, b, c, d, m, k, s)                                                       \
  do {                                                                                             \
    uint32_t mk_, t_;                                                                              \
    asm("add.u32 %0, %1, %2;" : "=r"(mk_) : "r"(m), "r"(k)); /* m + k: off the chain, kept apart */  \
    asm("{\n\t.reg .u32 x;\n\tadd.u32 x, %1, %2;\n\tadd.u32 %0, x, %3;\n\t}" : "=r"(t_) : "r"(f(b, c, d)), "r"(a), "r"(mk_)); \
    a = b + rotl32(t_, s);                                                                         \
  } while (0)
#define APS_F1(b, c, d) ((d) ^ ((b) & ((c) ^ (d))))
#define APS_F2(b, c, d) ((c) ^ ((d) & ((b) ^ (c))))
#define APS_F3(b, c, d) ((b) ^ (c) ^ (d))
#define APS_F4(b, c, d) ((c) ^ ((b) | ~(d)))

__device__ __forceinline__ void md5_block(uint32_t h[4], const uint32_t m[16]) {
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
  APS_MD5_STEP(APS_F1, a, b, c, d, m[0], 0xd76aa478u, 7);  APS_MD5_STEP(APS_F1, d, a, b, c, m[1], 0xe8c7b756u, 12);
  APS_MD5_STEP(APS_F1, c, d, a, b, m[2], 0x242070dbu, 17); APS_MD5_STEP(APS_F1, b, c, d, a, m[3], 0xc1bdceeeu, 22);
  APS_MD5_STEP(APS_F1, a, b, c, d, m[4], 0xf57c0fafu, 7);  APS_MD5_STEP(APS_F1, d, a, b, c, m[5], 0x4787c62au, 12);
  APS_MD5_STEP(APS_F1, c, d, a, b, m[6], 0xa8304613u, 17); APS_MD5_STEP(APS_F1, b, c, d, a, m[7], 0xfd469501u, 22);
  APS_MD5_STEP(APS_F1, a, b, c, d, m[8], 0x698098d8u, 7);  APS_MD5_STEP(APS_F1, d, a, b, c, m[9], 0x8b44f7afu, 12);
  APS_MD5_STEP(APS_F1, c, d, a, b, m[10], 0xffff5bb1u, 17); APS_MD5_STEP(APS_F1, b, c, d, a, m[11], 0x895cd7beu, 22);
  APS_MD5_STEP(APS_F1, a, b, c, d, m[12], 0x6b901122u, 7);  APS_MD5_STEP(APS_F1, d, a, b, c, m[13], 0xfd987193u, 12);
  APS_MD5_STEP(APS_F1, c, d, a, b, m[14], 0xa679438eu, 17); APS_MD5_STEP(APS_F1, b, c, d, a, m[15], 0x49b40821u, 22);
  APS_MD5_STEP(APS_F2, a, b, c, d, m[1], 0xf61e2562u, 5);  APS_MD5_STEP(APS_F2, d, a, b, c, m[6], 0xc040b340u, 9);
  APS_MD5_STEP(APS_F2, c, d, a, b, m[11], 0x265e5a51u, 14); APS_MD5_STEP(APS_F2, b, c, d, a, m[0], 0xe9b6c7aau, 20);
  APS_MD5_STEP(APS_F2, a, b, c, d, m[5], 0xd62f105du, 5);  APS_MD5_STEP(APS_F2, d, a, b, c, m[10], 0x02441453u, 9);
  APS_MD5_STEP(APS_F2, c, d, a, b, m[15], 0xd8a1e681u, 14); APS_MD5_STEP(APS_F2, b, c, d, a, m[4], 0xe7d3fbc8u, 20);
  APS_MD5_STEP(APS_F2, a, b, c, d, m[9], 0x21e1cde6u, 5);  APS_MD5_STEP(APS_F2, d, a, b, c, m[14], 0xc33707d6u, 9);
  APS_MD5_STEP(APS_F2, c, d, a, b, m[3], 0xf4d50d87u, 14); APS_MD5_STEP(APS_F2, b, c, d, a, m[8], 0x455a14edu, 20);
  APS_MD5_STEP(APS_F2, a, b, c, d, m[13], 0xa9e3e905u, 5); APS_MD5_STEP(APS_F2, d, a, b, c, m[2], 0xfcefa3f8u, 9);
  APS_MD5_STEP(APS_F2, c, d, a, b, m[7], 0x676f02d9u, 14); APS_MD5_STEP(APS_F2, b, c, d, a, m[12], 0x8d2a4c8au, 20);
  APS_MD5_STEP(APS_F3, a, b, c, d, m[5], 0xfffa3942u, 4);  APS_MD5_STEP(APS_F3, d, a, b, c, m[8], 0x8771f681u, 11);
  APS_MD5_STEP(APS_F3, c, d, a, b, m[11], 0x6d9d6122u, 16); APS_MD5_STEP(APS_F3, b, c, d, a, m[14], 0xfde5380cu, 23);
  APS_MD5_STEP(APS_F3, a, b, c, d, m[1], 0xa4beea44u, 4);  APS_MD5_STEP(APS_F3, d, a, b, c, m[4], 0x4bdecfa9u, 11);
  APS_MD5_STEP(APS_F3, c, d, a, b, m[7], 0xf6bb4b60u, 16); APS_MD5_STEP(APS_F3, b, c, d, a, m[10], 0xbebfbc70u, 23);
  APS_MD5_STEP(APS_F3, a, b, c, d, m[13], 0x289b7ec6u, 4); APS_MD5_STEP(APS_F3, d, a, b, c, m[0], 0xeaa127fau, 11);
  APS_MD5_STEP(APS_F3, c, d, a, b, m[3], 0xd4ef3085u, 16); APS_MD5_STEP(APS_F3, b, c, d, a, m[6], 0x04881d05u, 23);
  APS_MD5_STEP(APS_F3, a, b, c, d, m[9], 0xd9d4d039u, 4);  APS_MD5_STEP(APS_F3, d, a, b, c, m[12], 0xe6db99e5u, 11);
  APS_MD5_STEP(APS_F3, c, d, a, b, m[15], 0x1fa27cf8u, 16); APS_MD5_STEP(APS_F3, b, c, d, a, m[2], 0xc4ac5665u, 23);
  APS_MD5_STEP(APS_F4, a, b, c, d, m[0], 0xf4292244u, 6);  APS_MD5_STEP(APS_F4, d, a, b, c, m[7], 0x432aff97u, 10);
  APS_MD5_STEP(APS_F4, c, d, a, b, m[14], 0xab9423a7u, 15); APS_MD5_STEP(APS_F4, b, c, d, a, m[5], 0xfc93a039u, 21);
  APS_MD5_STEP(APS_F4, a, b, c, d, m[12], 0x655b59c3u, 6); APS_MD5_STEP(APS_F4, d, a, b, c, m[3], 0x8f0ccc92u, 10);
  APS_MD5_STEP(APS_F4, c, d, a, b, m[10], 0xffeff47du, 15); APS_MD5_STEP(APS_F4, b, c, d, a, m[1], 0x85845dd1u, 21);
  APS_MD5_STEP(APS_F4, a, b, c, d, m[8], 0x6fa87e4fu, 6);  APS_MD5_STEP(APS_F4, d, a, b, c, m[15], 0xfe2ce6e0u, 10);
  APS_MD5_STEP(APS_F4, c, d, a, b, m[6], 0xa3014314u, 15); APS_MD5_STEP(APS_F4, b, c, d, a, m[13], 0x4e0811a1u, 21);
  APS_MD5_STEP(APS_F4, a, b, c, d, m[4], 0xf7537e82u, 6);  APS_MD5_STEP(APS_F4, d, a, b, c, m[11], 0xbd3af235u, 10);
  APS_MD5_STEP(APS_F4, c, d, a, b, m[2], 0x2ad7d2bbu, 15); APS_MD5_STEP(APS_F4, b, c, d, a, m[9], 0xeb86d391u, 21);
  h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One warp per block; lane f digests file blockIdx.x*32 + f.
// `state` (optional) carries the four chaining words between calls so that files larger than HBM can be
// digested in waves: with finalize == 0 the call consumes len (a multiple of 64) bytes and stores the
// words back; with finalize != 0 it also pads for prefix[file] + len total bytes and writes the digest.
__global__ void __launch_bounds__(32) md5_kernel(uint32_t n_files, const void* const* data, const uint64_t* len,
                                                 uint8_t (*out)[16], uint32_t (*state)[4], const uint64_t* prefix,
                                                 int finalize) {
  extern __shared__ __align__(16) uint32_t md5_tile[];
  uint32_t(*tile)[32 * kMd5RowWords] = reinterpret_cast<uint32_t(*)[32 * kMd5RowWords]>(md5_tile);
  const uint32_t lane = threadIdx.x;
  const uint32_t file = blockIdx.x * 32 + lane;
  const bool have = file < n_files;
  const uint8_t* my = have ? (const uint8_t*)data[file] : nullptr;
  const uint64_t my_len = have ? len[file] : 0;
  // staging needs a 16-byte aligned base; otherwise the lane reads its bytes straight from global
  const bool my_aligned = (((uintptr_t)my) & 15u) == 0;
  uint64_t max_len = my_len;
#pragma unroll
  for (int o = 16; o; o >>= 1) max_len = max(max_len, __shfl_xor_sync(0xFFFFFFFFu, max_len, o));
  const uint64_t nseg = max_len / kMd5Seg;  // segments that are full for at least one file

  uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
  if (state && have) {
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = state[file][i];
  }

  auto stage = [&](uint64_t seg, int buf) {
    // for each file of the warp: 32 lanes fetch its 512-byte segment, 16 bytes each
    for (uint32_t f = 0; f < 32; ++f) {
      const uint8_t* base = (const uint8_t*)__shfl_sync(0xFFFFFFFFu, (unsigned long long)my, f);
      const uint64_t flen = __shfl_sync(0xFFFFFFFFu, (unsigned long long)my_len, f);
      const bool fal = __shfl_sync(0xFFFFFFFFu, (int)my_aligned, f);
      if (fal && (seg + 1) * kMd5Seg <= flen) {
        const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(&tile[buf][f * kMd5RowWords]);
#pragma unroll
        for (uint32_t q = 0; q < kMd5Seg / 512u; ++q)
          cp_async16(sbase + (q * 32u + lane) * 16u, base + seg * kMd5Seg + (q * 32u + lane) * 16u);
      }
    }
    cp_async_commit();
  };

  if (nseg) stage(0, 0);
  for (uint64_t seg = 0; seg < nseg; ++seg) {
    const int buf = (int)(seg & 1);
    if (seg + 1 < nseg) { stage(seg + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncwarp();
    if (have && (seg + 1) * kMd5Seg <= my_len) {
      if (my_aligned) {
        const uint4* row = reinterpret_cast<const uint4*>(&tile[buf][lane * kMd5RowWords]);
        // the next block's sixteen words are read from shared memory while the current block's chain runs
        uint32_t m[16], mn[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 v = row[j];
          m[4 * j] = v.x; m[4 * j + 1] = v.y; m[4 * j + 2] = v.z; m[4 * j + 3] = v.w;
        }
#pragma unroll 1
        for (int blk = 0; blk < (int)(kMd5Seg / 64u); ++blk) {
          if (blk + 1 < (int)(kMd5Seg / 64u)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 v = row[(blk + 1) * 4 + j];
              mn[4 * j] = v.x; mn[4 * j + 1] = v.y; mn[4 * j + 2] = v.z; mn[4 * j + 3] = v.w;
            }
          }
          md5_block(h, m);
#pragma unroll
          for (int j = 0; j < 16; ++j) m[j] = mn[j];
        }
      } else {
        const uint8_t* g = my + seg * kMd5Seg;
        for (int blk = 0; blk < (int)(kMd5Seg / 64u); ++blk) {
          uint32_t m[16];
          for (int j = 0; j < 16; ++j) {
            const uint8_t* q = g + blk * 64 + j * 4;
            m[j] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
          }
          md5_block(h, m);
        }
      }
    }
    __syncwarp();
  }
  if (!have) return;
  // remainder (< one staged segment) + padding, straight from global memory
  uint64_t pos = (my_len / kMd5Seg) * kMd5Seg;
  while (pos + 64 <= my_len) {
    uint32_t m[16];
    for (int j = 0; j < 16; ++j) {
      const uint8_t* q = my + pos + j * 4;
      m[j] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    }
    md5_block(h, m);
    pos += 64;
  }
  if (!finalize) {  // intermediate wave: hand the chaining words to the next call
#pragma unroll
    for (int i = 0; i < 4; ++i) state[file][i] = h[i];
    return;
  }
  uint8_t last[128];
  uint32_t rem = (uint32_t)(my_len - pos);
  for (uint32_t i = 0; i < rem; ++i) last[i] = my[pos + i];
  last[rem] = 0x80;
  const uint32_t total = rem < 56 ? 64 : 128;
  for (uint32_t i = rem + 1; i < total - 8; ++i) last[i] = 0;
  const uint64_t bits = (my_len + (prefix ? prefix[file] : 0ull)) * 8;
  for (int i = 0; i < 8; ++i) last[total - 8 + i] = (uint8_t)(bits >> (8 * i));
  for (uint32_t o = 0; o < total; o += 64) {
    uint32_t m[16];
    for (int j = 0; j < 16; ++j)
      m[j] = (uint32_t)last[o + j * 4] | ((uint32_t)last[o + j * 4 + 1] << 8) | ((uint32_t)last[o + j * 4 + 2] << 16) |
             ((uint32_t)last[o + j * 4 + 3] << 24);
    md5_block(h, m);
  }
  for (int i = 0; i < 4; ++i)
    for (int k = 0; k < 4; ++k) out[file][4 * i + k] = (uint8_t)(h[i] >> (8 * k));
}

#endif  // APS_KERNELS_TEMPLATES_ONLY

}  // namespace aps
