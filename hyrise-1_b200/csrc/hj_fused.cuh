// hj_fused.cuh — the two-pass partitioning path for radix_bits > 8 without a separate histogram pass over the input.
//
//  P1 k_part1     replaces materialize_input + the first half of partition_radix_parallel (join_hash.cpp:203-355):
//                 decode (dictionary gather ONCE), full-resolution histogram, and a SLAB-LOCAL level-1 partitioning:
//                 every 4096-row tile is grouped by its level-1 digit in shared memory and written back to the tile's
//                 own place in the tuple buffer, plus a 130-entry offset table per tile ("fragments"). Grouping a tile
//                 needs no global offsets, hence no histogram BEFORE the scatter: K1's decoded-key round trip (4 B
//                 written + 4 B read per row) and pass 1's second sweep over the input disappear.
//  P2 k_scatter2f replaces the second half: a CTA gathers the fragments of ONE level-1 digit from a group of consecutive
//                 slabs (expected 4096 tuples), ranks them by the low digit, reserves the runs with one global atomic per
//                 final partition and writes them out — the final layout is the same contiguous partition array the
//                 one-histogram path produces, so K4/K5 are unchanged.
//  P3 k_emit_null_probe_frag  K7 for this layout (routed NULL probe rows live in each slab's last fragment).
//
// v1/v2 of the library (K1 + k_scatter1 + k_scatter2) spent 6.3 + 5.0 + 4.0 ms here on C2; see DESIGN.md §4.
#pragma once
#include <type_traits>

#include "hj_kernels.cuh"

#ifndef HJ_TUNE_P1_PREFETCH
#define HJ_TUNE_P1_PREFETCH 0
#endif
#ifndef HJ_TUNE_P1_DICT_PREFETCH
#define HJ_TUNE_P1_DICT_PREFETCH 0
#endif
#ifndef HJ_TUNE_P1_GATHER
// how P1 reads a dictionary entry: 0 = ld.global.nc (allocates an L1 line per gather; the default, measured), 1 = ld.global.cg
// (L2 only), 2 = ld.global.nc.L1::no_allocate. The gather never hits L1 (a 2048-row tile touches 2048 of ~1 M entries), and
// P1 is bound by the misses it can keep in flight. Measured on C2 (profiles/r02_sweep_variants5_c2_full.json): 0 = 7.88 ms,
// 1 = 7.75 ms, 2 = 19.96 ms.
#define HJ_TUNE_P1_GATHER 1
#endif
#ifndef HJ_TUNE_P2_SMEM
#define HJ_TUNE_P2_SMEM 1  // 1: pass 2's small shared arrays as members of one struct: 5.285 -> 5.175 ms on C2 (profiles/r02_sweep_variants5_c2_full.json)
#endif
#ifndef HJ_TUNE_P2_STEPS
#define HJ_TUNE_P2_STEPS 1
#endif
#ifndef HJ_TUNE_P2_OWN_SHIFT
#define HJ_TUNE_P2_OWN_SHIFT 3  // s_own[k] = fragment holding flat index k << shift
#endif

namespace hj {

__device__ __forceinline__ uint32_t smem_addr_of(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// Slab layout (pass 1 -> pass 2): tile t's slab starts at t * slab tuples (slab = tile rows + F1 + 2, even) and holds the
// tile's tuples grouped by level-1 digit; every fragment starts on a 16-byte boundary, so a digit with an odd number of
// 8-byte tuples is followed by a hole: a tuple whose position is kHoleIdx (no row has it: relations hold < 2^32 - 1 rows).
constexpr uint32_t kHoleIdx = 0xFFFFFFFFu;
constexpr uint32_t kSlabPad = kMaxDigits + 2;  // shared-memory staging per tile: tile rows + this
template <typename T>
__device__ __forceinline__ uint32_t slab_pad(uint32_t n) { return sizeof(T) == 8 ? (n + 1u) & ~1u : n; }
template <typename K>
__device__ __forceinline__ Tuple<K, false> hole_tuple() {
  Tuple<K, false> t;
  t.key = K(0);
  t.idx = kHoleIdx;
  if constexpr (sizeof(K) == 8) t.pad = 0;
  return t;
}

template <typename K>
__device__ __forceinline__ K dict_gather(const DevSegment& s, uint32_t vid) {
#if HJ_TUNE_P1_GATHER == 0
  return load_plain_key<K>(s, vid);
#elif HJ_TUNE_P1_GATHER == 1
  if (s.key_type == 0) return static_cast<K>(__ldcg(static_cast<const int32_t*>(s.values) + vid));
  return static_cast<K>(__ldcg(static_cast<const long long*>(s.values) + vid));
#else
  if (s.key_type == 0) {
    int32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(static_cast<const int32_t*>(s.values) + vid));
    return static_cast<K>(v);
  }
  long long v;
  asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(v) : "l"(static_cast<const long long*>(s.values) + vid));
  return static_cast<K>(v);
#endif
}

// One tile's keys into registers: row u * VT + vt of the tile for u = 0 .. ITEMS-1, bit u of vbits = the row is not NULL.
// The chunk's encoding (attribute-vector width, element width of the values / dictionary, NULL flags) is resolved ONCE per
// tile here and the unrolled loops below are instantiated per case; with the dispatch inside the loops (round 1) every row
// paid three branches on segment fields, ~15 % of pass 1's instructions.
template <typename K, typename A, typename D, int ITEMS, int VT>
__device__ __forceinline__ void load_dict_rows(const DevSegment& s, uint32_t r0, uint32_t cnt, uint32_t vt, K (&key)[ITEMS], uint32_t& vbits) {
  const A* attr = static_cast<const A*>(s.attr) + r0;
  const D* dict = static_cast<const D*>(s.values);
  const uint32_t nv = s.null_value_id;
  uint32_t vid[ITEMS];
#pragma unroll
  for (int u = 0; u < ITEMS; ++u) {
    const uint32_t i = u * VT + vt;
    vid[u] = i < cnt ? static_cast<uint32_t>(__ldg(attr + i)) : nv;  // rows past the end carry null_value_id
  }
#pragma unroll
  for (int u = 0; u < ITEMS; ++u) {
    const bool valid = vid[u] != nv;
    // NULL decodes to T{} (dictionary_column_iterable.hpp:80); ld.global.cg: the gather never hits L1 (HJ_TUNE_P1_GATHER)
    key[u] = valid ? static_cast<K>(__ldcg(dict + vid[u])) : K(0);
    vbits |= (valid ? 1u : 0u) << u;
  }
}
template <typename K, typename D, bool NULLS, int ITEMS, int VT>
__device__ __forceinline__ void load_value_rows(const DevSegment& s, uint32_t r0, uint32_t cnt, uint32_t vt, K (&key)[ITEMS], uint32_t& vbits) {
  const D* values = static_cast<const D*>(s.values) + r0;
  const uint8_t* nulls = s.nulls + r0;
  const bool zero_nulls = s.kind == 2;  // GATHERED NULL yields T{} (reference_column_iterable.hpp:60); a ValueColumn keeps its value
#pragma unroll
  for (int u = 0; u < ITEMS; ++u) {
    const uint32_t i = u * VT + vt;
    key[u] = K(0);
    if (i < cnt) {
      key[u] = static_cast<K>(__ldg(values + i));
      bool valid = true;
      if constexpr (NULLS) {
        valid = __ldg(nulls + i) == 0;
        if (zero_nulls && !valid) key[u] = K(0);
      }
      vbits |= (valid ? 1u : 0u) << u;
    }
  }
}
template <typename K, int ITEMS, int VT>
__device__ __forceinline__ void load_tile_keys(const DevSegment& s, uint32_t r0, uint32_t cnt, uint32_t vt, K (&key)[ITEMS], uint32_t& vbits) {
  vbits = 0;
  const bool narrow = sizeof(K) == 8 && s.key_type == 0;  // an int32 chunk of an int64 join
  if (s.kind == 1) {
    if constexpr (sizeof(K) == 8) {
      if (narrow) {
        if (s.attr_width == 4) load_dict_rows<K, uint32_t, int32_t, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
        else if (s.attr_width == 2) load_dict_rows<K, uint16_t, int32_t, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
        else load_dict_rows<K, uint8_t, int32_t, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
        return;
      }
    }
    using D = typename std::conditional<sizeof(K) == 8, long long, int32_t>::type;
    if (s.attr_width == 4) load_dict_rows<K, uint32_t, D, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
    else if (s.attr_width == 2) load_dict_rows<K, uint16_t, D, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
    else load_dict_rows<K, uint8_t, D, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
  } else {
    if constexpr (sizeof(K) == 8) {
      if (narrow) {
        if (s.nulls) load_value_rows<K, int32_t, true, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
        else load_value_rows<K, int32_t, false, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
        return;
      }
    }
    using D = typename std::conditional<sizeof(K) == 8, long long, int32_t>::type;
    if (s.nulls) load_value_rows<K, D, true, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
    else load_value_rows<K, D, false, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
  }
}

// =====================================================================================================================
// P1. One physical CTA of NT threads hosts VC virtual CTAs of VT = NT/VC threads (as K1 does) that share the
// full-resolution histogram; every virtual CTA has its own staging tile. Virtual CTA c owns tiles c, c+G, ...
// frag[tile * (F1 + 2) + d]: offset of digit d's fragment inside the slab; [F1] = routed NULL rows; [F1 + 1] = rows kept.
// A tile is staged as keys + 16-bit row numbers inside the tile (6 instead of 8 bytes per int32 row): shared memory is L1
// taken away from the dictionary gather, whose rate is set by the L1 lines that can be in flight — measured on C2 (14
// radix bits, 64 KB histogram): 4 virtual CTAs with 8-byte staging (228 KB carve-out) 28 ms, 2 virtual CTAs 8.9 ms.
// dynamic smem: (2^radix_bits + 1) * 4 (rounded up to 16) + VC * (TILE + kSlabPad) * sizeof(Tuple<K>).
// =====================================================================================================================
template <typename K, int NT, int VC, int TILE>
__global__ void __launch_bounds__(NT) k_part1(RelMeta rel, Tuple<K, false>* __restrict__ out1, uint16_t* __restrict__ frag,
                                              uint32_t* __restrict__ hist, int radix_bits, int bits2, int null_policy, uint32_t slab) {
  using T = Tuple<K, false>;
  constexpr int VT = NT / VC;
  constexpr int ITEMS = TILE / VT;
  constexpr int DPT = (kMaxDigits + VT - 1) / VT;  // level-1 digits per thread in the scan of the tile's counts
  static_assert(VT % 32 == 0 && TILE % VT == 0 && TILE <= 65535, "whole warps per virtual CTA; 16-bit fragment offsets");
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint32_t* s_hist = reinterpret_cast<uint32_t*>(s_raw);
  const uint32_t P = 1u << radix_bits;
  const uint32_t n_bins = P + 1u;
  unsigned char* stage_raw = s_raw + ((n_bins * 4u + 15u) & ~15u);
  __shared__ uint32_t s_cnt[VC][kMaxDigits + 1], s_loff[VC][kMaxDigits + 2], s_scratch[VC][33];
  const uint32_t pmask = P - 1u;
  const uint32_t F1 = P >> bits2, FS = F1 + 2u;
  const uint32_t v = threadIdx.x / VT, vt = threadIdx.x % VT;
  const uint32_t G = gridDim.x * VC, c = blockIdx.x * VC + v;
  T* stile = reinterpret_cast<T*>(stage_raw) + v * (TILE + kSlabPad);
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) s_hist[b] = 0;
  __syncthreads();

  for (uint32_t tile = c; tile < rel.n_tiles; tile += G) {
    const uint32_t seg_idx = __ldg(rel.tile_seg + tile);
    const DevSegment s = rel.segs[seg_idx];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
#if HJ_TUNE_P1_DICT_PREFETCH
    // A chunk's dictionary is first touched by that chunk's own gathers: random 32-byte DRAM reads that every warp of
    // the wave waits for. The tiles of chunk c therefore pull the dictionary of chunk c+1 into L2 with coalesced
    // prefetches, one slice per tile, a wave ahead of its first use.
    if (seg_idx + 1 < rel.n_segs) {
      const DevSegment* nx = rel.segs + seg_idx + 1;
      const uint32_t dict_n = nx->pad1;  // 0 unless the chunk is dictionary-encoded
      if (dict_n) {
        const unsigned long long bytes = static_cast<unsigned long long>(dict_n) * (nx->key_type == 0 ? 4u : 8u);
        const uint32_t nt = (s.n_rows + TILE - 1) / TILE, t = tile - s.tile_begin;
        const unsigned long long lo = (bytes * t / nt) & ~127ull, hi = bytes * (t + 1) / nt;
        const char* base = static_cast<const char*>(nx->values);
        for (unsigned long long off = lo + vt * 128ull; off < hi; off += VT * 128ull)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(base + off));
      }
    }
#endif
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
    for (uint32_t d = vt; d <= F1; d += VT) s_cnt[v][d] = 0;
    vbar(v, VT);  // also orders the previous tile's copy-out before this tile's staging

    K key[ITEMS];
    uint32_t vbits;  // bit u: row u*VT+vt is valid (non-NULL)
    load_tile_keys<K, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
#if HJ_TUNE_P1_PREFETCH
    // the virtual CTA's next tile: pull its attribute vector / value lines into L2 while this tile is ranked and staged
    // (no registers held across the tile, unlike a software-pipelined load)
    if (tile + G < rel.n_tiles && vt < TILE * sizeof(K) / 128) {
      const DevSegment* ns = rel.segs + __ldg(rel.tile_seg + tile + G);
      const uint32_t nr0 = (tile + G - ns->tile_begin) * TILE;
      const uint32_t width = ns->kind == 1 ? ns->attr_width : static_cast<uint32_t>(sizeof(K));
      const char* base = static_cast<const char*>(ns->kind == 1 ? ns->attr : ns->values) + static_cast<size_t>(nr0) * width;
      const uint32_t bytes = min(static_cast<uint32_t>(TILE), ns->n_rows - nr0) * width;
      if (vt * 128u < bytes) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + vt * 128u));
    }
#endif
    uint32_t packed[ITEMS];  // digit << 16 | rank inside the digit, or kNoDigit
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      packed[u] = kNoDigit;
      if (u * VT + vt < cnt) {
        uint32_t d = kNoDigit;
        if (((vbits >> u) & 1u) || null_policy == kNullAsValue) {
          const uint32_t p = hash_key(key[u]) & pmask;
          atomicAdd(&s_hist[p], 1u);
          d = p >> bits2;
        } else if (null_policy == kNullRoute) {
          atomicAdd(&s_hist[P], 1u);
          d = F1;
        }
        if (d != kNoDigit) packed[u] = (d << 16) | atomicAdd(&s_cnt[v][d], 1u);
      }
    }
    vbar(v, VT);
    {
      // fragment starts are 16-byte aligned inside the slab (pass 2 fetches fragments with TMA bulk copies): a digit with an
      // odd number of 8-byte tuples is followed by a hole, a tuple whose position is kHoleIdx
      uint32_t cd[DPT], mine = 0;  // thread vt scans digits vt * DPT .. + DPT
#pragma unroll
      for (int k = 0; k < DPT; ++k) {
        const uint32_t d = vt * DPT + k;
        cd[k] = d < F1 ? s_cnt[v][d] : 0u;
        mine += slab_pad<T>(cd[k]);
      }
      uint32_t total;
      uint32_t ex = vblock_excl_scan<VT>(mine, v, vt, s_scratch[v], &total);
#pragma unroll
      for (int k = 0; k < DPT; ++k) {
        const uint32_t d = vt * DPT + k;
        if (d < F1) {
          s_loff[v][d] = ex;
          if (slab_pad<T>(cd[k]) != cd[k]) stile[ex + cd[k]] = hole_tuple<K>();
          ex += slab_pad<T>(cd[k]);
        }
      }
      if (vt == 0) {
        const uint32_t cn = s_cnt[v][F1];
        s_loff[v][F1] = total;
        s_loff[v][F1 + 1] = total + slab_pad<T>(cn);
        if (slab_pad<T>(cn) != cn) stile[total + cn] = hole_tuple<K>();
      }
    }
    vbar(v, VT);
    {
      const uint32_t idx0 = s.row_begin + r0;
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        if (packed[u] != kNoDigit) {
          T t;
          t.key = key[u];
          t.idx = idx0 + u * VT + vt;
          if constexpr (sizeof(K) == 8) t.pad = 0;
          stile[s_loff[v][packed[u] >> 16] + (packed[u] & 0xFFFFu)] = t;
        }
      }
      for (uint32_t d = vt; d < FS; d += VT) frag[static_cast<size_t>(tile) * FS + d] = static_cast<uint16_t>(s_loff[v][d]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the staged tuples, for the TMA engine's reads
      vbar(v, VT);
      // the slab goes out with ONE bulk copy of the TMA engine (measured against per-thread stores: P1 7.77 -> 7.61 ms on C2,
      // profiles/r02_sweep_tma_variants.md); its size is a multiple of 16 bytes by construction
      const uint32_t n_out = s_loff[v][F1 + 1];
      if (vt == 0 && n_out) {
        T* dst = out1 + static_cast<size_t>(tile) * slab;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr_of(stile)),
                     "r"(n_out * static_cast<uint32_t>(sizeof(T))) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging may be overwritten after this
      }
    }
  }
  __syncthreads();
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) {
    const uint32_t cb = s_hist[b];
    if (cb) atomicAdd(hist + b, cb);
  }
}

// =====================================================================================================================
// P1, software-pipelined (HJ_TUNE_P1_PIPE=1; measured SLOWER than k_part1 for column inputs, see hj_api.cu). Same result as k_part1 — slab-local level-1 grouping + fragment
// table + full-resolution histogram — but a virtual CTA keeps THREE tiles in flight instead of one:
//   tile t+2: its attribute-vector entries (or NULL flags) are on their way into registers (`pre`);
//   tile t+1: its dictionary gathers are on their way into shared memory — cp.async (LDGSTS), 4/8 bytes per row, no
//             register and no scoreboard slot held while they fly;
//   tile t  : ranked, grouped and written back.
// k_part1 serialised "gather (wait) -> rank -> scan -> stage -> copy out" per virtual CTA, so the gather pipe idled
// while a tile was being ranked and copied and the other way round: 2.09 cycles per row and SM where the gather alone
// needs 1.03 (profiles/r02_microbench_gather.txt: LDG = LDGSTS = 1.03-1.09 cycles per row and SM for 4-byte entries of
// a 4 MB L2-resident table, whatever the occupancy; 16-byte bulk copies through the TMA engine 3.6-4.1).
// The gathered keys stay where cp.async put them (kbuf[row]); grouping a tile only writes a 16-bit permutation
// (perm[position in the grouped slab] = row) and the copy-out reads kbuf through it — one shared-memory store and one
// load per row fewer than staging (key, row) pairs.
// XCH: the relation consists of kSegExch8 segments (rows received from the multi-GPU exchange as finished tuples:
// int32 joins (hash, source position), int64 joins (key, source position, source rank)); an element is the whole tuple.
// dynamic smem: (2^radix_bits + 1) * 4 (rounded up to 16) + VC * (TILE * 2 * sizeof(E) + (TILE + kSlabPad) * 2).
// =====================================================================================================================
template <int BYTES>
__device__ __forceinline__ void cp_async_small(void* dst_smem, const void* src) {
  static_assert(BYTES == 4 || BYTES == 8 || BYTES == 16, "cp.async copies 4, 8 or 16 bytes");
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr_of(dst_smem)), "l"(src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(smem_addr_of(dst_smem)), "l"(src), "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr uint8_t kSegExch8 = 14;  // internal: a receive region of the push exchange (hj_xpush.cuh): values = tuples, chunk_id =
                                   // source rank, attr = per-chunk fill counts, row_offsets = &number of chunks in use

// rows of tile t (0-based inside its segment): a receive region is a row of 2048-tuple chunks, of which the first *nchunks
// (the word `row_offsets` points at) are in use, filled to fill[t] tuples (`attr`)
__device__ __forceinline__ uint32_t exch_tile_rows(const DevSegment& s, uint32_t t) {
  return t < __ldcg(s.row_offsets) ? __ldcg(static_cast<const uint32_t*>(s.attr) + t) : 0u;
}

template <typename K, bool XCH> struct P1Elem { using type = K; };
template <typename K> struct P1Elem<K, true> { using type = Tuple<K, false>; };

template <typename K, int NT, int VC, int TILE, bool XCH>
__global__ void __launch_bounds__(NT) k_part1p(RelMeta rel, Tuple<K, false>* __restrict__ out1, uint16_t* __restrict__ frag,
                                               uint32_t* __restrict__ hist, int radix_bits, int bits2, int null_policy, uint32_t slab) {
  using T = Tuple<K, false>;
  using E = typename P1Elem<K, XCH>::type;
  constexpr int VT = NT / VC;
  constexpr int ITEMS = TILE / VT;
  static_assert(VT >= kMaxDigits && TILE <= 65535 && ITEMS >= 1 && ITEMS <= 32 && TILE % VT == 0,
                "one thread per digit in the scan; 16-bit rows; validity bits in one word");
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint32_t* s_hist = reinterpret_cast<uint32_t*>(s_raw);
  const uint32_t P = 1u << radix_bits;
  const uint32_t n_bins = P + 1u;
  unsigned char* buf_raw = s_raw + ((n_bins * 4u + 15u) & ~15u);
  __shared__ uint32_t s_cnt[VC][kMaxDigits + 1], s_loff[VC][kMaxDigits + 2], s_scratch[VC][33];
  const uint32_t pmask = P - 1u;
  const uint32_t F1 = P >> bits2, FS = F1 + 2u;
  const uint32_t v = threadIdx.x / VT, vt = threadIdx.x % VT;
  const uint32_t G = gridDim.x * VC, c = blockIdx.x * VC + v;
  E* kbuf = reinterpret_cast<E*>(buf_raw) + static_cast<size_t>(v) * 2 * TILE;  // [2][TILE]
  uint16_t* perm = reinterpret_cast<uint16_t*>(buf_raw + static_cast<size_t>(VC) * 2 * TILE * sizeof(E)) + v * (TILE + kSlabPad);
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) s_hist[b] = 0;
  __syncthreads();

  uint32_t pre[ITEMS];  // tile t+2: value ids (DICT) or NULL flags (other kinds)
  // registers <- the tile's attribute-vector entries / NULL flags (plain loads; consumed one iteration later)
  auto preload = [&](uint32_t tile) {
    if (XCH || tile >= rel.n_tiles) return;
    const DevSegment s = rel.segs[__ldg(rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * VT + vt;
      if (s.kind == 1) pre[u] = i < cnt ? load_attr(s, r0 + i) : s.null_value_id;
      else pre[u] = (i < cnt && s.nulls && __ldg(s.nulls + r0 + i)) ? 1u : 0u;
    }
  };
  // shared memory <- the tile's keys, asynchronously; returns the validity bits of this thread's rows
  auto issue = [&](uint32_t tile, E* kb) -> uint32_t {
    uint32_t vbits = 0;
    if (tile < rel.n_tiles) {
      const DevSegment s = rel.segs[__ldg(rel.tile_seg + tile)];
      const uint32_t r0 = (tile - s.tile_begin) * TILE;
      const uint32_t cnt = XCH ? exch_tile_rows(s, tile - s.tile_begin) : min(static_cast<uint32_t>(TILE), s.n_rows - r0);
      if constexpr (XCH) {
        const E* src = static_cast<const E*>(s.values) + r0;
#pragma unroll
        for (int u = 0; u < ITEMS; ++u) {
          const uint32_t i = u * VT + vt;
          if (i < cnt) cp_async_small<sizeof(E)>(kb + i, src + i);
        }
        vbits = 0xFFFFFFFFu;  // NULL rows never travel (kNullDrop) or travel as their value (kNullAsValue)
      } else {
        const bool narrow = sizeof(K) == 8 && s.key_type == 0;  // int32 chunk of an int64 join: widened by hand
#pragma unroll
        for (int u = 0; u < ITEMS; ++u) {
          const uint32_t i = u * VT + vt;
          if (i >= cnt) continue;
          const bool valid = s.kind == 1 ? pre[u] != s.null_value_id : pre[u] == 0u;
          vbits |= (valid ? 1u : 0u) << u;
          if (s.kind == 1 && !valid) continue;             // no dictionary entry to fetch (the dictionary may be empty)
          const uint32_t at = s.kind == 1 ? pre[u] : r0 + i;
          if (narrow) kb[i] = static_cast<K>(__ldg(static_cast<const int32_t*>(s.values) + at));
          else cp_async_small<sizeof(K)>(kb + i, static_cast<const K*>(s.values) + at);
        }
      }
    }
    cp_async_commit();  // one group per tile and thread, also when it is empty: wait_group counts groups
    return vbits;
  };

  uint32_t tile = c;
  preload(tile);
  uint32_t vb_cur = issue(tile, kbuf);
  preload(tile + G);
  uint32_t b = 0;
  for (; tile < rel.n_tiles; tile += G, b ^= 1u) {
    for (uint32_t d = vt; d <= F1; d += VT) s_cnt[v][d] = 0;
    vbar(v, VT);  // the previous tile's copy-out has read kbuf[b ^ 1] and perm
    const uint32_t vb_nxt = issue(tile + G, kbuf + (b ^ 1u) * TILE);
    preload(tile + 2 * G);
    cp_async_wait<1>();  // everything but the group just committed has landed: this tile's keys (own rows)

    const DevSegment s = rel.segs[__ldg(rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = XCH ? exch_tile_rows(s, tile - s.tile_begin) : min(static_cast<uint32_t>(TILE), s.n_rows - r0);
    E* kb = kbuf + b * TILE;
    uint32_t packed[ITEMS];  // digit << 16 | rank inside the digit, or kNoDigit
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      packed[u] = kNoDigit;
      const uint32_t i = u * VT + vt;
      if (i < cnt) {
        bool valid = (vb_cur >> u) & 1u;
        if constexpr (XCH) valid = kb[i].idx != kHoleIdx;  // the exchange pads its runs to 16-byte units with holes
        uint32_t d = kNoDigit;
        if (valid || (!XCH && null_policy == kNullAsValue)) {
          uint32_t h;
          if constexpr (XCH) {
            const E e = kb[i];
            if constexpr (sizeof(K) == 4) h = static_cast<uint32_t>(e.key); else h = hash_key(e.key);
          } else {
            K key = kb[i];
            // NULL decodes to T{} for dictionary / gathered chunks (dictionary_column_iterable.hpp:80,
            // reference_column_iterable.hpp:60); a ValueColumn keeps its stored value (quirk Q4)
            if (!valid && s.kind != 0) { key = K(0); kb[i] = key; }
            h = hash_key(key);
          }
          const uint32_t p = h & pmask;
          atomicAdd(&s_hist[p], 1u);
          d = p >> bits2;
        } else if (null_policy == kNullRoute) {
          if constexpr (!XCH) kb[i] = K(0);
          atomicAdd(&s_hist[P], 1u);
          d = F1;
        }
        if (d != kNoDigit) packed[u] = (d << 16) | atomicAdd(&s_cnt[v][d], 1u);
      }
    }
    vbar(v, VT);
    {
      const uint32_t cd = vt < F1 ? s_cnt[v][vt] : 0u;  // fragments start on 16-byte boundaries: holes (see slab_pad)
      uint32_t total;
      const uint32_t ex = vblock_excl_scan<VT>(slab_pad<T>(cd), v, vt, s_scratch[v], &total);
      if (vt < F1) {
        s_loff[v][vt] = ex;
        if (slab_pad<T>(cd) != cd) perm[ex + cd] = 0xFFFFu;
      }
      if (vt == 0) {
        const uint32_t cn = s_cnt[v][F1];
        s_loff[v][F1] = total;
        s_loff[v][F1 + 1] = total + slab_pad<T>(cn);
        if (slab_pad<T>(cn) != cn) perm[total + cn] = 0xFFFFu;
      }
    }
    vbar(v, VT);
#pragma unroll
    for (int u = 0; u < ITEMS; ++u)
      if (packed[u] != kNoDigit) perm[s_loff[v][packed[u] >> 16] + (packed[u] & 0xFFFFu)] = static_cast<uint16_t>(u * VT + vt);
    for (uint32_t d = vt; d < FS; d += VT) frag[static_cast<size_t>(tile) * FS + d] = static_cast<uint16_t>(s_loff[v][d]);
    vbar(v, VT);  // perm complete; every thread's cp.async data and NULL fix-ups are visible to the others
    const uint32_t n_out = s_loff[v][F1 + 1];  // <= TILE + F1 + 1
    T* dst = out1 + static_cast<size_t>(tile) * slab;
    const uint32_t idx0 = s.row_begin + r0;
#pragma unroll
    for (int u = 0; u <= ITEMS; ++u) {
      const uint32_t i = u * VT + vt;
      if (i < n_out) {
        const uint32_t r = perm[i];
        T t;
        if (r == 0xFFFFu) {
          t = hole_tuple<K>();
        } else if constexpr (XCH) {
          t = kb[r];
          // the source rank (= the receive region, chunk_id): int64 tuples have a spare word; an int32 tuple's key field is
          // the key's hash, whose level-1 digit bits are redundant from here on (this fragment IS that digit)
          if constexpr (sizeof(K) == 8) t.pad = s.chunk_id;
          else t.key = static_cast<K>((static_cast<uint32_t>(t.key) & ~(((1u << (radix_bits - bits2)) - 1u) << bits2)) | (s.chunk_id << bits2));
        } else {
          t.key = kb[r];
          t.idx = idx0 + r;
          if constexpr (sizeof(K) == 8) t.pad = 0;
        }
        dst[i] = t;
      }
    }
    vb_cur = vb_nxt;
  }
  cp_async_wait<0>();
  __syncthreads();
  for (uint32_t bb = threadIdx.x; bb < n_bins; bb += NT) {
    const uint32_t cb = s_hist[bb];
    if (cb) atomicAdd(hist + bb, cb);
  }
}

// =====================================================================================================================
// P2. Work item = (group of `gs` consecutive slabs, level-1 digit d1); the grid is n_groups * n_digits with d1 running
// fastest, so the CTAs resident at one moment read the same slabs (frag table and slab lines are shared through L2) and
// add to different cursors. The item's tuples are addressed through a flat index over its fragments: s_pre[] is the prefix
// sum of the fragment lengths, s_own[k] the fragment holding flat index 32k, so a thread finds its fragment with one
// lookup and a step or two, and all ITEMS loads of a thread are independent. gs is chosen so that an item holds 14/16 or
// 15/16 of TILE tuples on average (Cfg<K>::kP2Fill16); an item that holds more (skew) is worked off in batches of TILE.
// d1 == F1 (only with kNullRoute): the routed NULL probe rows, copied behind the last partition (part_off[P]).
// dynamic smem: TILE * sizeof(Tuple<K>).
// =====================================================================================================================
template <typename K, int NT, int ITEMS, int MINB, bool PRE = false>
__global__ void __launch_bounds__(NT, MINB) k_scatter2f(const Tuple<K, false>* __restrict__ in, const uint16_t* __restrict__ frag,
                                                            uint32_t n_tiles, uint32_t slab, Tuple<K, false>* __restrict__ out,
                                                            uint32_t* __restrict__ cur2, int radix_bits, int bits2, uint32_t gs,
                                                            uint32_t n_digits) {
  using T = Tuple<K, false>;
  constexpr uint32_t TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  T* stage = reinterpret_cast<T*>(s_raw);
  constexpr int OS = HJ_TUNE_P2_OWN_SHIFT;
#if HJ_TUNE_P2_SMEM == 0
  __shared__ uint32_t s_cnt[kMaxDigits], s_loff[kMaxDigits], s_delta[kMaxDigits];
  __shared__ uint32_t s_pre[NT + 1], s_dsrc[NT];
  __shared__ uint8_t s_own[TILE >> OS];
  __shared__ uint32_t scratch[33];
#else
  // one shared-memory object: every array is the same base + a constant. With separate arrays the compiler rebuilds three
  // shared-window addresses (S2R SR_CgaCtaId + LEAs) inside each of the 16 unrolled items: 1992 -> 1776 SASS instructions for
  // <int, 256, 16, 4>, at 8 bytes of spill. (An explicit 32-bit base through inline ld.shared changed nothing: ptxas
  // rematerialises the base per item either way.)
  struct Sm {
    uint32_t cnt[kMaxDigits], loff[kMaxDigits], delta[kMaxDigits];
    uint32_t pre[NT + 1], dsrc[NT];
    uint32_t scratch[33];
    uint8_t own[TILE >> OS];
  };
  __shared__ Sm sm;
  uint32_t* const s_cnt = sm.cnt; uint32_t* const s_loff = sm.loff; uint32_t* const s_delta = sm.delta;
  uint32_t* const s_pre = sm.pre; uint32_t* const s_dsrc = sm.dsrc; uint32_t* const scratch = sm.scratch;
  uint8_t* const s_own = sm.own;
#endif

  const uint32_t P = 1u << radix_bits, F1 = P >> bits2, FS = F1 + 2u;
  const uint32_t group = blockIdx.x / n_digits, d1 = blockIdx.x % n_digits;
  const uint32_t s0 = group * gs, ns = min(gs, n_tiles - s0);  // gs < NT
  const bool null_item = d1 == F1;
  const uint32_t F2 = null_item ? 1u : 1u << bits2;
  uint32_t* cursors = cur2 + (null_item ? P : d1 << bits2);

  uint32_t len = 0, src = 0;
  if (threadIdx.x < ns) {
    const uint32_t tile = s0 + threadIdx.x;
    const uint16_t* f = frag + static_cast<size_t>(tile) * FS + d1;
    const uint32_t f0 = __ldg(f), f1 = __ldg(f + 1);
    src = tile * slab + f0;
    len = f1 - f0;  // holes included
  }
  uint32_t total;
  const uint32_t pre = block_excl_scan(len, scratch, &total);
  s_pre[threadIdx.x] = pre;  // threads >= ns hold `total`: s_pre[ns] closes the last fragment
  s_dsrc[threadIdx.x] = src - pre;
  if (threadIdx.x == NT - 1) s_pre[NT] = total;

  for (uint32_t lo = 0; lo < total; lo += TILE) {
    const uint32_t n_in = min(TILE, total - lo);
    for (uint32_t d = threadIdx.x; d < F2; d += NT) s_cnt[d] = 0;
    if (len) {
      const uint32_t a = max(pre, lo), b = min(pre + len, lo + TILE);
      for (uint32_t k = (a - lo + (1u << OS) - 1u) >> OS; lo + (k << OS) < b; ++k) s_own[k] = static_cast<uint8_t>(threadIdx.x);
    }
    __syncthreads();  // also orders the previous batch's copy-out before this batch's staging

    T tup[ITEMS];
    uint32_t packed[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      const uint32_t i = j * NT + threadIdx.x;
      if (i < n_in) {
        const uint32_t fi = lo + i;
        uint32_t s = s_own[i >> OS];
        // fragments hold ~32 tuples, so the owner of flat index 32k is a step or two behind. HJ_TUNE_P2_STEPS predicated
        // steps before the loop were measured SLOWER (C2 pass 2: 5.21 -> 5.34 ms with 2 steps) although the divergent loop
        // is 16 % of the kernel's instructions: the steps add two dependent shared-memory loads to every lane
#pragma unroll
        for (int step = 0; step < HJ_TUNE_P2_STEPS; ++step) s += fi >= s_pre[s + 1] ? 1u : 0u;
        while (fi >= s_pre[s + 1]) ++s;
        tup[j] = ld_stream(in + static_cast<uint32_t>(s_dsrc[s] + fi));
      }
    }
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      const uint32_t i = j * NT + threadIdx.x;
      packed[j] = kNoDigit;
      if (i < n_in && tup[j].idx != kHoleIdx) {
        const uint32_t d = tuple_hash<PRE>(tup[j].key) & (F2 - 1u);
        packed[j] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
      }
    }
    __syncthreads();
    uint32_t tot;
    {
      const uint32_t d = threadIdx.x;
      const uint32_t c = d < F2 ? s_cnt[d] : 0u;
      const uint32_t ex = block_excl_scan(c, scratch, &tot);
      if (d < F2) {
        s_loff[d] = ex;
        if (c) s_delta[d] = atomicAdd(cursors + d, c) - ex;
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      if (packed[j] != kNoDigit) stage[s_loff[packed[j] >> 16] + (packed[j] & 0xFFFFu)] = tup[j];
    }
    __syncthreads();
    copy_tile_out<K, false, NT, ITEMS, PRE>(stage, s_delta, out, tot, F2 - 1u, 0, 0xFFFFFFFFu, 0u);
  }
}

// =====================================================================================================================
// P2t. Pass 2 with the fragments fetched by the TMA engine (the default since round 2; k_scatter2f above is the same pass
// with per-thread loads, kept for the variant sweep). Same work items as k_scatter2f. The item's fragments — contiguous,
// 16-byte aligned runs of the slabs (see slab_pad) — are copied into the shared-memory tile with one cp.async.bulk per
// fragment, all tracked by ONE mbarrier (expect_tx = the batch's bytes); the threads then read their tuples from the tile
// with coalesced shared-memory loads. That removes what made k_scatter2f issue-bound (66 % issue slots, ~125 instructions
// per tuple): the fragment-owner table, the per-tuple owner search and flat-index arithmetic, and 16 scattered global loads
// per thread. Holes (odd fragments of 8-byte tuples) arrive with the data and are skipped by their position field.
// The tile is then reused as the staging area of the ranking/copy-out exactly as in k_scatter2f.
// dynamic smem: TILE * sizeof(Tuple<K>).
// =====================================================================================================================
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr_of(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr_of(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(smem_addr_of(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr_of(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_addr_of(bar)) : "memory");
}

template <typename K, int NT, int ITEMS, int MINB, bool PRE = false>
__global__ void __launch_bounds__(NT, MINB) k_scatter2t(const Tuple<K, false>* __restrict__ in, const uint16_t* __restrict__ frag,
                                                            uint32_t n_tiles, uint32_t slab, Tuple<K, false>* __restrict__ out,
                                                            uint32_t* __restrict__ cur2, int radix_bits, int bits2, uint32_t gs,
                                                            uint32_t n_digits) {
  using T = Tuple<K, false>;
  constexpr uint32_t TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  T* stage = reinterpret_cast<T*>(s_raw);
  struct Sm {
    uint32_t cnt[kMaxDigits], loff[kMaxDigits], delta[kMaxDigits];
    uint32_t scratch[33];
    uint32_t pad;
    unsigned long long bar;
  };
  __shared__ __align__(8) Sm sm;
  const uint32_t P = 1u << radix_bits, F1 = P >> bits2, FS = F1 + 2u;
  const uint32_t group = blockIdx.x / n_digits, d1 = blockIdx.x % n_digits;
  const uint32_t s0 = group * gs, ns = min(gs, n_tiles - s0);  // gs <= NT: one fragment per thread
  const bool null_item = d1 == F1;
  const uint32_t F2 = null_item ? 1u : 1u << bits2;
  uint32_t* cursors = cur2 + (null_item ? P : d1 << bits2);
  if (threadIdx.x == 0) mbar_init(&sm.bar, 1);

  uint32_t len = 0, src = 0;  // this thread's fragment: first tuple (index into `in`; the host checked that it fits 32 bits), tuples
  if (threadIdx.x < ns) {
    const uint32_t tile = s0 + threadIdx.x;
    const uint16_t* f = frag + static_cast<size_t>(tile) * FS + d1;
    const uint32_t f0 = __ldg(f), f1 = __ldg(f + 1);
    src = tile * slab + f0;
    len = f1 - f0;  // holes included: even for 8-byte tuples, like f0
  }
  uint32_t total;
  const uint32_t pre = block_excl_scan(len, sm.scratch, &total);  // its barriers also publish the mbarrier's initialisation
  uint32_t phase = 0;
  for (uint32_t lo = 0; lo < total; lo += TILE, phase ^= 1u) {
    const uint32_t n_in = min(TILE, total - lo);
    for (uint32_t d = threadIdx.x; d < F2; d += NT) sm.cnt[d] = 0;
    __syncthreads();  // the previous batch's copy-out has read the tile: the TMA engine may overwrite it
    if (threadIdx.x == 0) mbar_expect_tx(&sm.bar, n_in * static_cast<uint32_t>(sizeof(T)));
    if (len) {
      const uint32_t a = max(pre, lo), b = min(pre + len, lo + TILE);  // this fragment's part of the batch (even bounds)
      if (a < b) bulk_g2s(stage + (a - lo), in + (src + (a - pre)), (b - a) * static_cast<uint32_t>(sizeof(T)), &sm.bar);
    }
    mbar_wait(&sm.bar, phase);

    // the tile's tuples go to registers; a tuple's rank inside its digit is kept as 16 bits (two per register) and the digit
    // is re-derived from the key when the tuple is staged: 8 registers less than a (digit, rank) word per tuple, which is what
    // the kernel needs to stay close to the 64 registers that 1024 resident threads per SM (2 CTAs of 512 / 4 of 256) allow
    static_assert(ITEMS % 2 == 0, "ranks are packed in pairs");
    T tup[ITEMS];
    uint32_t rank2[ITEMS / 2];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      const uint32_t i = j * NT + threadIdx.x;
      uint32_t r = 0;
      tup[j].idx = kHoleIdx;
      if (i < n_in) {
        tup[j] = stage[i];
        if (tup[j].idx != kHoleIdx) r = atomicAdd(&sm.cnt[tuple_hash<PRE>(tup[j].key) & (F2 - 1u)], 1u);
      }
      if (j & 1) rank2[j / 2] |= r << 16; else rank2[j / 2] = r;
    }
    __syncthreads();  // every tuple of the tile is in registers; the counts are complete
    uint32_t tot;
    {
      const uint32_t d = threadIdx.x;
      const uint32_t c = d < F2 ? sm.cnt[d] : 0u;
      const uint32_t ex = block_excl_scan(c, sm.scratch, &tot);
      if (d < F2) {
        sm.loff[d] = ex;
        if (c) sm.delta[d] = atomicAdd(cursors + d, c) - ex;
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      if (tup[j].idx != kHoleIdx)
        stage[sm.loff[tuple_hash<PRE>(tup[j].key) & (F2 - 1u)] + ((rank2[j / 2] >> ((j & 1) * 16)) & 0xFFFFu)] = tup[j];
    }
    __syncthreads();
    copy_tile_out<K, false, NT, ITEMS, PRE>(stage, sm.delta, out, tot, F2 - 1u, 0, 0xFFFFFFFFu, 0u);
  }
}

}  // namespace hj
