// hj_device.cuh — device-side data layout and small primitives shared by all libhj kernels (sm_100a).
//
// Data layout in HBM:
//   * input segments stay where the caller put them (ValueColumn arrays, dictionary + attribute vector);
//   * a partitioned tuple is (key, row index): 8 B for int32 keys, 16 B for int64 keys. The row index is the
//     position of the row in the concatenation of the relation's chunks; it is expanded back to
//     RowID{ChunkID, ChunkOffset} (types.hpp:97-99 of the reference) only when a result row is emitted. This is
//     the reference's PartitionedElement (join_hash.cpp:98-106: 16/24 B) without the stored hash and with a
//     4-byte position instead of the 8-byte RowID.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hj {

constexpr uint32_t kInvalid = 0xFFFFFFFFu;  // INVALID_CHUNK_ID / INVALID_CHUNK_OFFSET, types.hpp:94-95
constexpr int kMaxRadixBits = 15;            // full-resolution histogram must fit one CTA's shared memory (128 KB)

enum NullPolicy : int {
  kNullDrop = 0,    // NULL keys never enter the join (build side; probe side of Inner/Semi/Anti), join_hash.cpp:253
  kNullAsValue = 1, // probe side of Left/Right, literal reference behaviour (quirk Q4): key = stored value / T{}
  kNullRoute = 2    // probe side of Left/Right with HJ_FLAG_SQL_NULL_PROBE: routed to a side list, emitted unmatched
};

// Internal segment kind (outside hj_segment_kind 0..3; HJ_SEG_REFERENCE never reaches the device, it is gathered into a
// GATHERED segment first): tuples received from the multi-GPU exchange. `values` = keys,
// `nulls` = optional flags, `attr` = the rows' packed global RowIDs (uint64), which replace the position -> RowID mapping.
constexpr uint8_t kSegExchanged = 15;

// One chunk's key column on the device. 64 bytes.
struct DevSegment {
  const void* values;           // VALUE/GATHERED: T[n_rows]; DICT: dictionary
  const uint8_t* nulls;         // VALUE/GATHERED: 1 byte per row or nullptr
  const void* attr;             // DICT: attribute vector
  const uint32_t* row_offsets;  // GATHERED: ChunkOffset to report per row or nullptr
  uint32_t row_begin;           // position of the chunk's first row in the concatenated relation
  uint32_t n_rows;
  uint32_t chunk_id;
  uint32_t null_value_id;
  uint32_t tile_begin;          // index of the chunk's first tile
  uint32_t mask_word_begin;     // first word of the chunk's validity bitmap (DICT chunks only)
  uint8_t kind, key_type, attr_width, pad0;
  uint32_t pad1;
};
static_assert(sizeof(DevSegment) == 64, "DevSegment layout");

// Per-relation metadata handed to kernels by value.
struct RelMeta {
  const DevSegment* segs;
  const uint32_t* tile_seg;     // tile index -> segment index
  const uint32_t* seg_row_begin;  // n_segs+1 entries (for position -> RowID when chunks are irregular)
  uint32_t n_segs;
  uint32_t n_rows;
  uint32_t n_tiles;
  int32_t uniform_shift;        // >= 0: every chunk but the last has exactly 2^shift rows
  uint32_t identity;            // 1: chunk_id == segment index for all chunks and no row_offsets anywhere
  uint32_t pad;
};

// PAY: the tuple carries a 64-bit payload (idx, pad) — the row's global RowID after the multi-GPU exchange — instead of a
// 32-bit position. int64-key tuples have room for it anyway (16 B); int32-key tuples grow from 8 to 12 bytes.
template <typename K, bool PAY = false>
struct Tuple;
template <>
struct __align__(8) Tuple<int32_t, false> {
  int32_t key;
  uint32_t idx;
};
template <>
struct Tuple<int32_t, true> {
  int32_t key;
  uint32_t idx;
  uint32_t pad;
};
template <bool PAY>
struct __align__(16) Tuple<int64_t, PAY> {
  int64_t key;
  uint32_t idx;
  uint32_t pad;
};
static_assert(sizeof(Tuple<int32_t, false>) == 8 && sizeof(Tuple<int32_t, true>) == 12 && sizeof(Tuple<int64_t, false>) == 16 &&
                  sizeof(Tuple<int64_t, true>) == 16, "tuple layouts");

// vector loads of a partitioned tuple through the read-only path (evict-first .cs hints measured slower on B200: K1 +12 ms
// with st.cs, passes +0.1..0.4 ms with ld.cs)
// [p, p + bytes) requested into L2 by the TMA engine (one instruction; the range is widened to 16-byte bounds, which stay
// inside the allocation: device buffers are 256-byte aligned and padded)
__device__ __forceinline__ void bulk_prefetch_l2(const void* p, uint32_t bytes) {
  if (bytes == 0) return;
  const unsigned long long a0 = reinterpret_cast<unsigned long long>(p) & ~15ull;
  const unsigned long long a1 = (reinterpret_cast<unsigned long long>(p) + bytes + 15ull) & ~15ull;
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(static_cast<uint32_t>(a1 - a0)) : "memory");
}

__device__ __forceinline__ Tuple<int32_t, false> ld_stream(const Tuple<int32_t, false>* p) {
  const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
  Tuple<int32_t, false> t;
  t.key = static_cast<int32_t>(v.x);
  t.idx = v.y;
  return t;
}
__device__ __forceinline__ Tuple<int32_t, true> ld_stream(const Tuple<int32_t, true>* p) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(p);
  Tuple<int32_t, true> t;
  t.key = static_cast<int32_t>(__ldg(w));
  t.idx = __ldg(w + 1);
  t.pad = __ldg(w + 2);
  return t;
}
template <bool PAY>
__device__ __forceinline__ Tuple<int64_t, PAY> ld_stream(const Tuple<int64_t, PAY>* p) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
  Tuple<int64_t, PAY> t;
  t.key = static_cast<int64_t>(static_cast<uint64_t>(v.x) | (static_cast<uint64_t>(v.y) << 32));
  t.idx = v.z;
  t.pad = v.w;
  return t;
}

// ---------------------------------------------------------------------------------------------------------- hash
// The reference partitions on MurmurHash2(seed 17) (join_hash.cpp:254); the result multiset does not depend on the hash,
// so the device uses the cheaper murmur3 finaliser. Low `radix_bits` bits select the partition, the bits above them
// the hash-table slot, so the two are independent.
__host__ __device__ __forceinline__ uint32_t fmix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
__host__ __device__ __forceinline__ uint32_t hash_key(int32_t k) { return fmix32(static_cast<uint32_t>(k)); }
__host__ __device__ __forceinline__ uint32_t hash_key(int64_t k) {
  const uint32_t lo = static_cast<uint32_t>(k), hi = static_cast<uint32_t>(static_cast<uint64_t>(k) >> 32);
  return fmix32(lo ^ fmix32(hi + 0x9e3779b9u));
}

// Multi-GPU exchange: the rank that joins a key. Derived from a second mix of the key hash so that it is independent of
// both the local partition (low radix bits) and the hash-table slot (the bits above them); mulhi maps it to [0, n_ranks).
__host__ __device__ __forceinline__ uint32_t dest_rank_of_hash(uint32_t h, uint32_t n_ranks) {
  const uint32_t g = fmix32(h ^ 0x9e3779b9u);
  return static_cast<uint32_t>((static_cast<uint64_t>(g) * n_ranks) >> 32);
}

// ---------------------------------------------------------------------------------------------------------- RowID
// position in the concatenated relation -> RowID{chunk_id, chunk_offset} packed as chunk_id | offset << 32
// (little endian: chunk_id first, as in struct RowID).
__device__ __forceinline__ uint64_t pack_rowid(uint32_t chunk_id, uint32_t off) {
  return static_cast<uint64_t>(chunk_id) | (static_cast<uint64_t>(off) << 32);
}
constexpr uint64_t kNullRowId = 0xFFFFFFFFFFFFFFFFull;  // NULL_ROW_ID, types.hpp:150

__device__ __forceinline__ uint64_t make_rowid(const RelMeta& m, uint32_t idx) {
  uint32_t seg, off;
  if (m.uniform_shift >= 0) {
    seg = idx >> m.uniform_shift;
    off = idx & ((1u << m.uniform_shift) - 1u);
  } else {
    // largest seg with seg_row_begin[seg] <= idx (empty chunks share a begin; pick the non-empty one)
    uint32_t lo = 0, hi = m.n_segs;  // invariant: seg_row_begin[lo] <= idx < seg_row_begin[hi]
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (__ldg(m.seg_row_begin + mid) <= idx) lo = mid; else hi = mid;
    }
    seg = lo;
    off = idx - __ldg(m.seg_row_begin + lo);
  }
  if (m.identity) return pack_rowid(seg, off);
  const DevSegment* s = m.segs + seg;
  if (s->kind == kSegExchanged) return __ldg(static_cast<const uint64_t*>(s->attr) + off);
  const uint32_t* ro = s->row_offsets;
  return pack_rowid(s->chunk_id, ro ? __ldg(ro + off) : off);
}

// ---------------------------------------------------------------------------------------------------------- warp/block
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t warp_id() { return threadIdx.x >> 5; }

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, v, d);
    if (lane_id() >= static_cast<uint32_t>(d)) v += n;
  }
  return v;
}

// Exclusive scan of one value per thread over the whole CTA (blockDim.x a multiple of 32, <= 1024).
// `scratch` holds 33 words of shared memory. Returns the exclusive prefix; *total receives the CTA sum.
// Contains three __syncthreads() (the last one makes back-to-back calls on the same scratch safe).
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* scratch, uint32_t* total) {
  const uint32_t incl = warp_incl_scan(v);
  if (lane_id() == 31) scratch[warp_id()] = incl;
  __syncthreads();
  if (warp_id() == 0) {
    const uint32_t nw = blockDim.x >> 5;
    const uint32_t w = lane_id() < nw ? scratch[lane_id()] : 0u;
    const uint32_t wi = warp_incl_scan(w);
    scratch[lane_id()] = wi - w;
    if (lane_id() == 31) scratch[32] = wi;
  }
  __syncthreads();
  *total = scratch[32];
  const uint32_t r = scratch[warp_id()] + incl - v;
  __syncthreads();
  return r;
}

// ---- virtual CTAs: one physical CTA hosts several groups of VT threads that synchronise among themselves only
// barrier over the VT threads of virtual CTA v (hardware barriers 1..15; 0 is __syncthreads)
__device__ __forceinline__ void vbar(uint32_t v, uint32_t n_threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(v + 1u), "r"(n_threads) : "memory");
}

// block_excl_scan over one virtual CTA (vt = thread index inside it); scratch: 33 words per virtual CTA
template <int VT>
__device__ __forceinline__ uint32_t vblock_excl_scan(uint32_t val, uint32_t v, uint32_t vt, uint32_t* scratch, uint32_t* total) {
  const uint32_t incl = warp_incl_scan(val);
  const uint32_t w = vt >> 5;
  if ((vt & 31u) == 31u) scratch[w] = incl;
  vbar(v, VT);
  if (w == 0) {
    constexpr uint32_t nw = VT / 32;
    const uint32_t x = (vt & 31u) < nw ? scratch[vt & 31u] : 0u;
    const uint32_t xi = warp_incl_scan(x);
    scratch[vt & 31u] = xi - x;
    if ((vt & 31u) == 31u) scratch[32] = xi;
  }
  vbar(v, VT);
  *total = scratch[32];
  const uint32_t r = scratch[w] + incl - val;
  vbar(v, VT);
  return r;
}

}  // namespace hj
