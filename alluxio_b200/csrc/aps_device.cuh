// Device-side building blocks of libaps: hash-table probing (K1) and GF(2) CRC32C arithmetic (K3).
// sm_100a only.  See DESIGN.md for the layout and the roofline of each kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/aps.h"

namespace aps {

// ------------------------------------------------------------------------------------------------
// K1: hash table.  Open addressing over 256-byte buckets of eight 32-byte entries; one bucket is
// read by a tile of eight lanes (one 32-byte sector each), keys compared with a tile ballot.
// Replaces DefaultMetaStore.mPageMap (CACHE/DefaultMetaStore.java:36,52) and the path lookup in
// LocalPageStore.getFilePath (CACHE/store/LocalPageStore.java:139-143).
// ------------------------------------------------------------------------------------------------
struct __align__(32) Entry {
  uint64_t k0, k1, k2;  // file_hi, file_lo, page_index
  uint64_t meta;        // EMPTY / TOMB / LOCKED, or VALID | slot << 32 | len
};
constexpr uint32_t kBucketEntries = 8;
constexpr uint64_t kMetaEmpty = 0, kMetaTomb = 1, kMetaLocked = 2;
constexpr uint64_t kMetaValid = 1ull << 63;
constexpr uint32_t kNoSlot = 0xFFFFFFFFu;

__host__ __device__ __forceinline__ uint64_t meta_make(uint32_t slot, uint32_t len) {
  return kMetaValid | ((uint64_t)slot << 32) | len;
}
__host__ __device__ __forceinline__ uint32_t meta_slot(uint64_t m) { return (uint32_t)(m >> 32) & 0x7FFFFFFFu; }
__host__ __device__ __forceinline__ uint32_t meta_len(uint64_t m) { return (uint32_t)m; }

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t hash_key(const aps_key& k) {
  return mix64(k.file_hi ^ mix64(k.file_lo ^ mix64(k.page_index + 0x9E3779B97F4A7C15ull)));
}

// Owner shard of a page in a cache sharded over several GPUs.  Uses the high half of the key hash:
// the low bits pick the bucket inside the owner's table and must stay uniform there.
__host__ __device__ __forceinline__ uint32_t shard_of(const aps_key& k, uint32_t n_shards) {
  return (uint32_t)((hash_key(k) >> 32) % n_shards);
}

enum Counter { kCtrPages = 0, kCtrBytes, kCtrTombs, kCtrLookups, kCtrProbes, kCtrBytesCopied, kCtrCrcMismatch, kCtrEvicted, kNumCounters };

struct Table {
  Entry* entries;
  uint64_t bucket_mask;  // buckets - 1 (power of two)
  uint32_t* free_slots;  // stack of free slab slots
  int32_t* free_top;     // number of entries on the stack
  uint32_t* slot_crc;    // CRC32C recorded at put, per slab slot
  unsigned long long* counters;
  uint8_t* slab;
  uint32_t page_size;    // bytes between slab slots: the configured page size rounded up to a multiple of 16
  // opt-in device-side CLOCK eviction (APS_FLAG_DEVICE_CLOCK); both null otherwise
  uint8_t* slot_ref;     // per slab slot: 1 = referenced since the hand last passed
  uint64_t* slot_entry;  // per slab slot: index of the table entry that owns it, ~0 = free
  uint32_t max_len;      // longest page a put may store = alluxio.user.client.cache.page.size as configured (any value)
};
constexpr uint64_t kNoEntry = ~0ull;

__device__ __forceinline__ void load_entry(const Entry* e, uint64_t& k0, uint64_t& k1, uint64_t& k2, uint64_t& meta) {
  // .cg: always observe L2, entries are written by other SMs in earlier (or the same) launches
  const ulonglong2* p = reinterpret_cast<const ulonglong2*>(e);
  ulonglong2 a = __ldcg(p), b = __ldcg(p + 1);
  k0 = a.x; k1 = a.y; k2 = b.x; meta = b.y;
}

struct Probe {
  uint64_t meta;       // meta of the matching entry, 0 if absent
  uint64_t entry_idx;  // index of the matching entry
  uint32_t empties;    // bitmask of EMPTY entries in the bucket where the probe ended
  uint32_t nbuckets;   // buckets read
};

// Tile-cooperative lookup.  `sub` = lane within the tile (0..7), `tshift` = first lane of the tile in
// the warp.  All eight lanes of the tile must call it with the same key.
__device__ __forceinline__ Probe tile_lookup(const Table& t, const aps_key& key, uint32_t sub, uint32_t tshift) {
  const uint32_t tmask = 0xFFu << tshift;
  uint64_t b = hash_key(key) & t.bucket_mask;
  Probe r{0, 0, 0, 0};
  for (uint64_t p = 0; p <= t.bucket_mask; ++p) {
    uint64_t k0, k1, k2, meta;
    load_entry(t.entries + b * kBucketEntries + sub, k0, k1, k2, meta);
    r.nbuckets++;
    bool match = (meta >> 63) && k0 == key.file_hi && k1 == key.file_lo && k2 == key.page_index;
    uint32_t bm = (__ballot_sync(tmask, match) >> tshift) & 0xFFu;
    uint32_t be = (__ballot_sync(tmask, meta == kMetaEmpty) >> tshift) & 0xFFu;
    if (bm) {
      uint32_t src = __ffs(bm) - 1;
      r.meta = __shfl_sync(tmask, meta, tshift + src);
      r.entry_idx = b * kBucketEntries + src;
      r.empties = be;
      return r;
    }
    if (be) {
      r.empties = be;
      return r;
    }
    b = (b + 1) & t.bucket_mask;
  }
  return r;
}

// Block-scope accounting.  A batch kernel touches the store's counters and its free-slot stack once per
// request; done as global atomics these all land on a handful of addresses and serialise in L2 at about
// one per clock, which is what bounded the K1 kernels (1.1 G lookups/s against a random-read rate of 27 G
// buckets/s).  Instead each block adds into shared memory and issues one global atomic per counter, and
// takes or returns its free slots with one atomic for the whole block.
struct BlockAcct {
  uint32_t lo[kNumCounters], hi[kNumCounters];  // 64-bit sums kept as halves: shared memory adds natively in 32 bits only
  int need;   // tiles of this block that take (put) or return (delete) a slot
  int base;   // stack position the block's range starts at
  int avail;  // how many of `need` were granted
};
__device__ __forceinline__ void acct_init(BlockAcct& a) {
  if (threadIdx.x < kNumCounters) a.lo[threadIdx.x] = a.hi[threadIdx.x] = 0;
  if (threadIdx.x == 0) a.need = 0;
  __syncthreads();
}
__device__ __forceinline__ void acct_add(BlockAcct& a, Counter c, unsigned long long v) {  // v may be a two's complement negative
  const uint32_t vlo = (uint32_t)v, old = atomicAdd(&a.lo[c], vlo);
  const uint32_t vhi = (uint32_t)(v >> 32) + ((uint32_t)(old + vlo) < old ? 1u : 0u);  // carry out of the low half
  if (vhi) atomicAdd(&a.hi[c], vhi);
}
__device__ __forceinline__ void acct_flush(BlockAcct& a, const Table& t) {  // call with the whole block
  __syncthreads();
  if (threadIdx.x < kNumCounters) {
    const unsigned long long v = ((unsigned long long)a.hi[threadIdx.x] << 32) | a.lo[threadIdx.x];
    if (v) atomicAdd(&t.counters[threadIdx.x], v);
  }
}

// Tile-cooperative first half of insert-or-replace: finds the id or claims an entry for it (the entry is
// left LOCKED, which probes of other ids step over).  Keys in one launch must be distinct.
enum ClaimKind { kClaimFull = 0, kClaimReplace = 1, kClaimNew = 2 };
struct EntryClaim {
  int kind;
  uint64_t entry_idx;  // the id's entry (replace) or the claimed one (new)
  uint64_t was;        // replace: the old meta; new: what the claimed entry held (EMPTY or TOMB)
};
__device__ __forceinline__ EntryClaim tile_claim(const Table& t, const aps_key& key, uint32_t sub, uint32_t tshift) {
  const uint32_t tmask = 0xFFu << tshift;
  const uint64_t home = hash_key(key) & t.bucket_mask;
  for (int attempt = 0; attempt < (1 << 16); ++attempt) {  // a lost race means another key took one entry: bounded by the batch size
    uint64_t b = home;
    uint64_t cand = ~0ull, cand_seen = 0;
    bool ended = false;
    for (uint64_t p = 0; p <= t.bucket_mask && !ended; ++p) {
      uint64_t k0, k1, k2, meta;
      load_entry(t.entries + b * kBucketEntries + sub, k0, k1, k2, meta);
      bool match = (meta >> 63) && k0 == key.file_hi && k1 == key.file_lo && k2 == key.page_index;
      uint32_t bm = (__ballot_sync(tmask, match) >> tshift) & 0xFFu;
      uint32_t be = (__ballot_sync(tmask, meta == kMetaEmpty) >> tshift) & 0xFFu;
      uint32_t bt = (__ballot_sync(tmask, meta == kMetaTomb) >> tshift) & 0xFFu;
      if (bm) {  // present: PageStore.put overwrites, the slot is kept
        uint32_t src = __ffs(bm) - 1;
        return EntryClaim{kClaimReplace, b * kBucketEntries + src, __shfl_sync(tmask, meta, tshift + src)};
      }
      if (cand == ~0ull && (bt | be)) {  // keep EMPTY entries for chain termination: prefer a tombstone
        uint32_t pick = bt ? (__ffs(bt) - 1) : (__ffs(be) - 1);
        cand = b * kBucketEntries + pick;
        cand_seen = bt ? kMetaTomb : kMetaEmpty;
      }
      if (be) ended = true;
      b = (b + 1) & t.bucket_mask;
    }
    if (cand == ~0ull) return EntryClaim{kClaimFull, 0, 0};  // every bucket full
    int ok = 0;
    if (sub == 0)
      ok = atomicCAS((unsigned long long*)&t.entries[cand].meta, (unsigned long long)cand_seen,
                     (unsigned long long)kMetaLocked) == cand_seen;
    ok = __shfl_sync(tmask, ok, tshift);
    if (ok) return EntryClaim{kClaimNew, cand, cand_seen};
    // another key took the entry first: rescan
  }
  return EntryClaim{kClaimFull, 0, 0};
}

// Block-wide pop of `a.need` slots from the free stack with one atomic; fewer may be granted.
// Call with the whole block, after every tile that needs a slot has taken its rank from a.need.
__device__ __forceinline__ void block_take_slots(BlockAcct& a, const Table& t) {
  __syncthreads();
  if (threadIdx.x == 0 && a.need) {
    const int old = atomicSub(t.free_top, a.need);
    const int got = max(0, min(a.need, old));
    if (got < a.need) atomicAdd(t.free_top, a.need - got);
    a.base = old;
    a.avail = got;
  }
  __syncthreads();
}
// Block-wide push: reserves a.need positions; tile `rank` then writes its slot at a.base + rank.
__device__ __forceinline__ void block_return_slots(BlockAcct& a, const Table& t) {
  __syncthreads();
  if (threadIdx.x == 0 && a.need) a.base = atomicAdd(t.free_top, a.need);
  __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// K3: CRC32C arithmetic.  Reflected CRC-32C register algebra: bit 31 is x^0, bit 0 is x^31.
//   update(r, byte) = (r >> 8) ^ T[(r ^ byte) & 255];  Z_n = "advance the register by n zero bytes"
//   is multiplication by x^(8n) mod P, and crc(A||B) = Z_|B|(crc(A)) ^ crc(B) for zero-init CRCs.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kCrcPoly = 0x82F63B78u;

__host__ __device__ __forceinline__ uint32_t gf_mul(uint32_t a, uint32_t b) {  // a * b mod P
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    r ^= a & (0u - ((b >> (31 - i)) & 1u));
    a = (a >> 1) ^ (kCrcPoly & (0u - (a & 1u)));
  }
  return r;
}

// The same product as a CALL: inlined, its 32 unrolled steps are ~130 instructions at every one of a dozen sites, which
// made the copy kernel 135 KB of code -- more than the instruction cache holds, and a page-sized task walks through most
// of it.  None of the sites is in the streaming loop.
static __device__ __noinline__ uint32_t gf_mul_call(uint32_t a, uint32_t b) { return gf_mul(a, b); }

// x^(8 * 2^k) mod P for k = 0..31 (filled by the host at load time)
__constant__ uint32_t c_xpow8[32];
// per-lane constant of the 512-byte row reduction: x^(8 * (500 - 16 * lane))
__constant__ uint32_t c_lane_shift[32];

__device__ __forceinline__ uint32_t xpow8(uint32_t nbytes) {  // x^(8 * nbytes) mod P
  uint32_t r = 0x80000000u;                                    // x^0
  for (int k = 0; nbytes; ++k, nbytes >>= 1)
    if (nbytes & 1u) r = gf_mul_call(r, c_xpow8[k]);
  return r;
}

__device__ __forceinline__ uint32_t crc_byte_bitwise(uint32_t r, uint32_t byte) {
  r ^= byte;
#pragma unroll
  for (int k = 0; k < 8; ++k) r = (r >> 1) ^ (kCrcPoly & (0u - (r & 1u)));
  return r;
}

}  // namespace aps
