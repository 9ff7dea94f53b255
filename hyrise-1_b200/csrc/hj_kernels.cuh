// hj_kernels.cuh — the hand-written sm_100a kernels of libhj. One translation unit (hj_api.cu) includes this.
//
//  K1 k_materialize_hist   replaces materialize_input            (reference join_hash.cpp:203-285)
//  K2 k_offsets            replaces the serial prefix sum         (join_hash.cpp:310-319)
//  K3 k_scatter<FromSegs>  replaces partition_radix_parallel      (join_hash.cpp:287-355), pass 1 from the segments,
//                          pass 2 from pass 1's output
//  K4 k_plan               work list for K5 (no reference counterpart: the reference schedules one JobTask per
//                          partition, join_hash.cpp:150,387,485)
//  K5 k_build_probe        replaces build + probe + probe_semi_anti (join_hash.cpp:127-185, 362-466, 468-527)
//  K6 k_finalize_multi     unmatched-row emission for build partitions larger than one shared-memory table
//  K7 k_emit_null_probe    (NULL, p) rows for NULL probe keys under HJ_FLAG_SQL_NULL_PROBE
//  K8 k_translate          PosList dereferencing of write_output_columns (join_hash.cpp:584-592)
#pragma once
#include "hj_device.cuh"

namespace hj {

// =====================================================================================================================
// Row loader shared by K1 and K3<FromSegs>: decodes row `i` of a segment into (key, valid).
// =====================================================================================================================
template <typename K>
__device__ __forceinline__ K load_plain_key(const DevSegment& s, uint32_t i) {
  if (s.key_type == 0) return static_cast<K>(__ldg(static_cast<const int32_t*>(s.values) + i));
  return static_cast<K>(__ldg(static_cast<const int64_t*>(s.values) + i));
}
__device__ __forceinline__ uint32_t load_attr(const DevSegment& s, uint32_t i) {
  if (s.attr_width == 4) return __ldg(static_cast<const uint32_t*>(s.attr) + i);
  if (s.attr_width == 2) return __ldg(static_cast<const uint16_t*>(s.attr) + i);
  return __ldg(static_cast<const uint8_t*>(s.attr) + i);
}

// =====================================================================================================================
// K1: materialize + histogram.
//   DICT chunks: attr -> dictionary gather (served by L2) -> decoded key written once to `keys_mat` (so that the
//   scatter pass does not repeat the gather) + one validity bit per row; VALUE/GATHERED chunks are only read.
//   Every kept row increments a full-resolution (2^radix_bits bins) shared-memory histogram; CTAs are persistent so
//   the histogram is flushed to global memory once per CTA.
// grid: min(n_tiles, SMs * occupancy), block: NT threads, dynamic smem: (2^radix_bits + 1) * 4 bytes.
// =====================================================================================================================
template <typename K, int NT, int TILE>
__global__ void __launch_bounds__(NT) k_materialize_hist(RelMeta rel, K* __restrict__ keys_mat,
                                                         uint32_t* __restrict__ valid_mask,
                                                         uint32_t* __restrict__ hist, int radix_bits, int null_policy) {
  extern __shared__ uint32_t s_hist[];
  const uint32_t n_bins = (1u << radix_bits) + 1u;  // last bin: NULL rows routed aside (kNullRoute)
  const uint32_t pmask = (1u << radix_bits) - 1u;
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) s_hist[b] = 0;
  __syncthreads();

  constexpr int ITEMS = TILE / NT;
  constexpr int U = ITEMS < 4 ? ITEMS : 4;
  for (uint32_t tile = blockIdx.x; tile < rel.n_tiles; tile += gridDim.x) {
    const uint32_t seg_idx = __ldg(rel.tile_seg + tile);
    const DevSegment s = rel.segs[seg_idx];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
    const bool is_dict = s.kind == 1;
#pragma unroll 1
    for (int j0 = 0; j0 < ITEMS; j0 += U) {
      K key[U];
      bool valid[U], active[U];
      if (is_dict) {
        uint32_t vid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t i = (j0 + u) * NT + threadIdx.x;
          active[u] = i < cnt;
          vid[u] = active[u] ? load_attr(s, r0 + i) : s.null_value_id;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          valid[u] = active[u] && vid[u] != s.null_value_id;
          key[u] = valid[u] ? load_plain_key<K>(s, vid[u]) : K(0);  // NULL decodes to T{} (dictionary_column_iterable.hpp:80)
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t i = (j0 + u) * NT + threadIdx.x;
          if (active[u]) keys_mat[s.row_begin + r0 + i] = key[u];
          const uint32_t bits = __ballot_sync(0xffffffffu, valid[u]);
          // rows r0+i for one warp instruction are 32 consecutive rows starting at a multiple of 32
          if (lane_id() == 0 && (i - lane_id()) < cnt) valid_mask[s.mask_word_begin + ((r0 + i) >> 5)] = bits;
        }
      } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t i = (j0 + u) * NT + threadIdx.x;
          active[u] = i < cnt;
          key[u] = active[u] ? load_plain_key<K>(s, r0 + i) : K(0);
          valid[u] = active[u] && !(s.nulls && __ldg(s.nulls + r0 + i));
        }
        if (s.kind == 2) {  // GATHERED: a NULL position yields T{} (reference_column_iterable.hpp:60)
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (!valid[u]) key[u] = K(0);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!active[u]) continue;
        if (valid[u] || null_policy == kNullAsValue) {
          atomicAdd(&s_hist[hash_key(key[u]) & pmask], 1u);
        } else if (null_policy == kNullRoute) {
          atomicAdd(&s_hist[n_bins - 1], 1u);
        }
      }
    }
  }
  __syncthreads();
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) {
    const uint32_t c = s_hist[b];
    if (c) atomicAdd(hist + b, c);
  }
}

// =====================================================================================================================
// K2: offsets. One CTA turns the histogram into
//   part_off[P+2] : start of every final partition, part_off[P] = start of the routed-NULL region, part_off[P+1] = total
//   cur1[F1+1]    : pass-1 write cursors (one per level-1 partition, + the NULL region)
//   cur2[P]       : pass-2 write cursors
//   l1_tiles[F1+1]: prefix sum of ceil(level-1 partition size / tile2) — tile -> partition map for pass 2
// =====================================================================================================================
template <int NT>
__global__ void __launch_bounds__(NT) k_offsets(const uint32_t* __restrict__ hist, uint32_t* __restrict__ part_off,
                                                uint32_t* __restrict__ cur1, uint32_t* __restrict__ cur2,
                                                uint32_t* __restrict__ l1_tiles, int radix_bits, int bits2,
                                                uint32_t tile2) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry;
  const uint32_t P = 1u << radix_bits;
  const uint32_t n = P + 1;  // + NULL bin
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n; base += NT) {
    const uint32_t b = base + threadIdx.x;
    const uint32_t c = b < n ? hist[b] : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(c, scratch, &total) + s_carry;
    if (b < n) {
      part_off[b] = ex;
      if (b < P) cur2[b] = ex;
      if (b == P) cur1[P >> bits2] = ex;
      else if ((b & ((1u << bits2) - 1u)) == 0) cur1[b >> bits2] = ex;
    }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) part_off[P + 1] = s_carry;
  __syncthreads();
  // level-1 tile prefix (F1 <= 512 entries)
  const uint32_t F1 = P >> bits2;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < F1; base += NT) {
    const uint32_t d = base + threadIdx.x;
    uint32_t t = 0;
    if (d < F1) {
      const uint32_t sz = part_off[(d + 1) << bits2] - part_off[d << bits2];
      t = (sz + tile2 - 1) / tile2;
    }
    uint32_t total;
    const uint32_t ex = block_excl_scan(t, scratch, &total) + s_carry;
    if (d < F1) l1_tiles[d] = ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) l1_tiles[F1] = s_carry;
}

// =====================================================================================================================
// K3: radix scatter, one tile per CTA.
//   rank every tuple inside its digit with a shared-memory atomic, scan the digit counts, reserve the tile's slice of
//   every destination partition with ONE global atomic per (tile, digit), reorder the tile in shared memory so that
//   each digit is a contiguous run, then copy the runs out with consecutive threads writing consecutive addresses.
//   FROM_SEGS = true : pass 1, reads the chunks (decoded keys of DICT chunks come from keys_mat), digit = high bits.
//   FROM_SEGS = false: pass 2, reads pass 1's output inside one level-1 partition, digit = low `bits2` bits.
// Within a partition the tuple order is arbitrary (the reference guarantees none: join_hash.hpp:19-20).
// block: NT threads, dynamic smem: NT*ITEMS*sizeof(Tuple<K>).
// =====================================================================================================================
template <typename K>
struct ScatterArgs {
  RelMeta rel;                  // pass 1
  const K* keys_mat;            // pass 1
  const uint32_t* valid_mask;   // pass 1
  const Tuple<K>* in;           // pass 2
  Tuple<K>* out;
  uint32_t* cursors;            // pass 1: cur1[F1+1]; pass 2: cur2[P]
  const uint32_t* part_off;     // pass 2
  const uint32_t* l1_tiles;     // pass 2
  int radix_bits;
  int bits2;                    // low bits handled by pass 2 (0: pass 1 is the only pass)
  int null_policy;
};

constexpr int kMaxDigits = 256;  // per pass; + 1 NULL bin in pass 1

template <typename K, int NT, int ITEMS, bool FROM_SEGS>
__global__ void __launch_bounds__(NT, 2) k_scatter(const ScatterArgs<K> a) {
  constexpr uint32_t TILE = NT * ITEMS;
  constexpr uint32_t NONE = 0xFFFFFFFFu;
  extern __shared__ __align__(16) unsigned char s_raw[];
  Tuple<K>* stage = reinterpret_cast<Tuple<K>*>(s_raw);
  __shared__ uint32_t s_cnt[kMaxDigits + 1], s_loff[kMaxDigits + 2], s_gdelta[kMaxDigits + 1];
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_l1[kMaxDigits + 1];

  const uint32_t pmask = (1u << a.radix_bits) - 1u;
  uint32_t n_digits, cur_base = 0, n_in = 0, in_base = 0;
  DevSegment seg;
  if (FROM_SEGS) {
    n_digits = (1u << (a.radix_bits - a.bits2)) + 1u;  // + NULL bin
    const uint32_t tile = blockIdx.x;
    seg = a.rel.segs[__ldg(a.rel.tile_seg + tile)];
    in_base = (tile - seg.tile_begin) * TILE;  // first row of the tile inside the chunk
    n_in = min(TILE, seg.n_rows - in_base);
  } else {
    n_digits = 1u << a.bits2;
    const uint32_t F1 = 1u << (a.radix_bits - a.bits2);
    for (uint32_t d = threadIdx.x; d <= F1; d += NT) s_l1[d] = a.l1_tiles[d];
    __syncthreads();
    if (blockIdx.x >= s_l1[F1]) return;  // grid is an upper bound
    uint32_t lo = 0, hi = F1;  // largest d1 with s_l1[d1] <= blockIdx.x
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (s_l1[mid] <= blockIdx.x) lo = mid; else hi = mid;
    }
    const uint32_t d1 = lo, k = blockIdx.x - s_l1[lo];
    const uint32_t p_lo = a.part_off[d1 << a.bits2], p_hi = a.part_off[(d1 + 1) << a.bits2];
    in_base = p_lo + k * TILE;
    n_in = min(TILE, p_hi - in_base);
    cur_base = d1 << a.bits2;
  }
  for (uint32_t d = threadIdx.x; d < n_digits; d += NT) s_cnt[d] = 0;
  __syncthreads();

  K key[ITEMS];
  uint32_t idx[ITEMS];
  uint32_t packed[ITEMS];  // digit << 16 | rank inside the digit, or NONE
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t i = j * NT + threadIdx.x;
    packed[j] = NONE;
    key[j] = K(0);
    idx[j] = 0;
    if (i < n_in) {
      bool valid = true;
      if (FROM_SEGS) {
        idx[j] = seg.row_begin + in_base + i;
        if (seg.kind == 1) {
          key[j] = a.keys_mat[idx[j]];
          valid = (__ldg(a.valid_mask + seg.mask_word_begin + ((in_base + i) >> 5)) >> (i & 31u)) & 1u;
        } else {
          key[j] = load_plain_key<K>(seg, in_base + i);
          valid = !(seg.nulls && __ldg(seg.nulls + in_base + i));
          if (seg.kind == 2 && !valid) key[j] = K(0);
        }
      } else {
        const Tuple<K> t = a.in[in_base + i];
        key[j] = t.key;
        idx[j] = t.idx;
      }
      uint32_t d = NONE;
      if (valid || a.null_policy == kNullAsValue) {
        const uint32_t p = hash_key(key[j]) & pmask;
        d = FROM_SEGS ? (p >> a.bits2) : (p & ((1u << a.bits2) - 1u));
      } else if (a.null_policy == kNullRoute) {
        d = n_digits - 1;
      }
      if (d != NONE) packed[j] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
    }
  }
  __syncthreads();

  // digit counts -> local offsets; one global atomic per non-empty digit reserves the destination slice
  uint32_t total;
  {
    const uint32_t d = threadIdx.x;
    const uint32_t c = d < n_digits ? s_cnt[d] : 0u;
    const uint32_t ex = block_excl_scan(c, scratch, &total);
    if (d < n_digits) {
      s_loff[d] = ex;
      if (c) s_gdelta[d] = atomicAdd(a.cursors + cur_base + d, c) - ex;
    }
    if (d == 0) s_loff[n_digits] = total;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    if (packed[j] != NONE) {
      Tuple<K> t;
      t.key = key[j];
      t.idx = idx[j];
      if (sizeof(K) == 8) reinterpret_cast<uint32_t*>(&t)[3] = 0;
      stage[s_loff[packed[j] >> 16] + (packed[j] & 0xFFFFu)] = t;
    }
  }
  __syncthreads();
  const uint32_t null_begin = FROM_SEGS ? s_loff[n_digits - 1] : NONE;
  for (uint32_t i = threadIdx.x; i < total; i += NT) {
    const Tuple<K> t = stage[i];
    uint32_t d;
    if (FROM_SEGS) {
      d = i >= null_begin ? n_digits - 1 : ((hash_key(t.key) & pmask) >> a.bits2);
    } else {
      d = hash_key(t.key) & ((1u << a.bits2) - 1u);
    }
    a.out[s_gdelta[d] + i] = t;
  }
}

// =====================================================================================================================
// K4: plan. One CTA turns the two partition-offset arrays into the work list of K5.
//   item = (partition p, build chunk c, probe slice s): a build partition that does not fit one shared-memory table
//   (skew, duplicates) is cut into chunks of <= cap tuples, a large probe partition into slices of <= slice tuples, and
//   every (chunk, slice) pair is joined independently. Partitions with > 1 chunk are "multi": for them unmatched-row
//   decisions (Left/Right/Semi/Anti) need all chunks, so K5 only records matches in a bitmap and K6 finishes them.
// =====================================================================================================================
struct WorkItem {
  uint32_t b_begin, b_cnt, p_begin, p_cnt;
  uint32_t flags, part, pad0, pad1;
};
constexpr uint32_t kItemMulti = 1u;

struct PlanOut {
  uint32_t n_items;
  uint32_t n_fin;            // multi partitions that need K6
  uint32_t n_overflow_parts; // build partitions with > 1 chunk
  uint32_t sliced;
};

template <int NT>
__global__ void __launch_bounds__(NT) k_plan(const uint32_t* __restrict__ off_b, const uint32_t* __restrict__ off_p,
                                             int radix_bits, int mode, uint32_t cap, uint32_t slice,
                                             uint32_t items_capacity, WorkItem* __restrict__ items,
                                             WorkItem* __restrict__ fin, PlanOut* __restrict__ out) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry, s_carry_fin, s_ovf;
  const uint32_t P = 1u << radix_bits;
  const bool keeps_unmatched = mode == 1 || mode == 2 || mode == 6;  // Left, Right, Anti emit probe rows without a match
  const bool needs_fin = mode != 0;                                  // everything but Inner

  // round 0: would slicing fit the item buffer? (without slicing sum(chunks) <= P + n_build/cap always fits)
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) {
    s_total = 0;
    s_carry = 0;
    s_carry_fin = 0;
    s_ovf = 0;
  }
  __syncthreads();
  unsigned long long local = 0;
  for (uint32_t p = threadIdx.x; p < P; p += NT) {
    const uint32_t nb = off_b[p + 1] - off_b[p], np = off_p[p + 1] - off_p[p];
    if (np == 0 || (nb == 0 && !keeps_unmatched)) continue;
    const uint32_t chunks = nb ? (nb + cap - 1) / cap : 1u;
    const uint32_t slices = (np + slice - 1) / slice;
    local += static_cast<unsigned long long>(chunks) * slices;
  }
  if (local) atomicAdd(&s_total, local);
  __syncthreads();
  const bool sliced = s_total <= items_capacity;

  for (uint32_t base = 0; base < P; base += NT) {
    const uint32_t p = base + threadIdx.x;
    uint32_t nb = 0, np = 0, chunks = 0, slices = 0, n = 0, multi = 0;
    if (p < P) {
      nb = off_b[p + 1] - off_b[p];
      np = off_p[p + 1] - off_p[p];
      if (np != 0 && (nb != 0 || keeps_unmatched)) {
        chunks = nb ? (nb + cap - 1) / cap : 1u;
        slices = sliced ? (np + slice - 1) / slice : 1u;
        n = chunks * slices;
        multi = chunks > 1;
      }
    }
    uint32_t total, total_fin;
    const uint32_t ex = block_excl_scan(n, scratch, &total) + s_carry;
    const uint32_t exf = block_excl_scan((multi && needs_fin) ? 1u : 0u, scratch, &total_fin) + s_carry_fin;
    if (n) {
      const uint32_t per_slice = sliced ? slice : np;
      for (uint32_t c = 0; c < chunks; ++c) {
        for (uint32_t s = 0; s < slices; ++s) {
          WorkItem it;
          it.b_begin = off_b[p] + c * cap;
          it.b_cnt = nb ? min(cap, nb - c * cap) : 0u;
          it.p_begin = off_p[p] + s * per_slice;
          it.p_cnt = min(per_slice, np - s * per_slice);
          it.flags = multi ? kItemMulti : 0u;
          it.part = p;
          it.pad0 = it.pad1 = 0;
          items[ex + c * slices + s] = it;
        }
      }
      if (multi) atomicAdd(&s_ovf, 1u);
      if (multi && needs_fin) {
        WorkItem f;
        f.b_begin = f.b_cnt = 0;
        f.p_begin = off_p[p];
        f.p_cnt = np;
        f.flags = kItemMulti;
        f.part = p;
        f.pad0 = f.pad1 = 0;
        fin[exf] = f;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      s_carry += total;
      s_carry_fin += total_fin;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out->n_items = s_carry;
    out->n_fin = s_carry_fin;
    out->n_overflow_parts = s_ovf;
    out->sliced = sliced ? 1u : 0u;
  }
}

// =====================================================================================================================
// K5: build + probe, persistent CTAs pulling work items from an atomic queue.
// Per item, in shared memory:
//   keys[ns]      open-addressing table of the chunk's DISTINCT build keys (linear probing, atomicCAS claim);
//   w[ns]         per slot: count of build rows with that key; after the scan (end_of_run << 16 | count);
//   grp[cap]      the chunk's build row positions grouped by key (counting sort through w) — a key's duplicates are one
//                 contiguous run, so a probe hit expands with plain indexed reads (the reference's
//                 variant<RowID, PosList> per key, join_hash.cpp:111-112,158-175, without the heap).
// The key equal to the EMPTY sentinel (INT_MIN) is diverted to a dedicated slot `NS`.
// Probe: each thread looks up V tuples; per-warp prefix sums of the match counts + one global atomic per CTA tile reserve a
// contiguous output range, so the two PosLists are written with consecutive lanes on consecutive addresses; a probe row
// with many matches is expanded by its whole warp.
// block: NT threads (1 CTA/SM), dynamic smem: (NS+1)*(sizeof(K)+4) + CAP*4.
// =====================================================================================================================
template <typename K>
struct JoinArgs {
  const Tuple<K>* build;
  const Tuple<K>* probe;
  const WorkItem* items;
  const PlanOut* plan;
  uint32_t* queue_head;
  RelMeta build_rel, probe_rel;
  uint64_t* out_build;                 // nullptr for Semi/Anti
  uint64_t* out_probe;
  unsigned long long* out_cursor;      // rows emitted so far (keeps counting past `capacity`)
  unsigned long long capacity;
  uint32_t* matched;                   // bit per probe tuple position; written for multi items only
  int radix_bits;
  int mode;                            // hj_join_mode
};

template <typename K> struct KeyTraits;
template <> struct KeyTraits<int32_t> {
  static constexpr int32_t kEmpty = INT32_MIN;
  using Atom = int;
};
template <> struct KeyTraits<int64_t> {
  static constexpr int64_t kEmpty = INT64_MIN;
  using Atom = unsigned long long;
};

template <typename K>
__device__ __forceinline__ K smem_cas(K* addr, K expected, K desired);
template <>
__device__ __forceinline__ int32_t smem_cas<int32_t>(int32_t* addr, int32_t expected, int32_t desired) {
  return atomicCAS(reinterpret_cast<int*>(addr), expected, desired);
}
template <>
__device__ __forceinline__ int64_t smem_cas<int64_t>(int64_t* addr, int64_t expected, int64_t desired) {
  return static_cast<int64_t>(atomicCAS(reinterpret_cast<unsigned long long*>(addr),
                                        static_cast<unsigned long long>(expected),
                                        static_cast<unsigned long long>(desired)));
}

// slot of `key`, claiming an empty one if absent. Bounded: the table always has empty slots (ns > distinct keys).
template <typename K, int NS>
__device__ __forceinline__ uint32_t table_insert(K* keys, uint32_t ns_mask, K key, uint32_t slot) {
  if (key == KeyTraits<K>::kEmpty) return NS;
  for (uint32_t it = 0; it <= ns_mask; ++it) {
    K cur = *reinterpret_cast<volatile K*>(keys + slot);
    if (cur == KeyTraits<K>::kEmpty) {
      cur = smem_cas<K>(keys + slot, KeyTraits<K>::kEmpty, key);
      if (cur == KeyTraits<K>::kEmpty) return slot;
    }
    if (cur == key) return slot;
    slot = (slot + 1) & ns_mask;
  }
  return NS + 1;  // unreachable
}

// slot of `key` or NS+1 if absent (read-only)
template <typename K, int NS>
__device__ __forceinline__ uint32_t table_find(const K* keys, uint32_t ns_mask, K key, uint32_t slot) {
  if (key == KeyTraits<K>::kEmpty) return NS;
  for (uint32_t it = 0; it <= ns_mask; ++it) {
    const K cur = keys[slot];
    if (cur == key) return slot;
    if (cur == KeyTraits<K>::kEmpty) return NS + 1;
    slot = (slot + 1) & ns_mask;
  }
  return NS + 1;
}

template <typename K, int NT, int NS, int CAP, int V>
__global__ void __launch_bounds__(NT, 1) k_build_probe(const JoinArgs<K> a) {
  static_assert(CAP < NS && CAP <= 65535, "table must keep empty slots; run ends are 16 bit");
  constexpr uint32_t NW = NT / 32;
  constexpr uint32_t MIN_NS = 32 * NW;  // every warp scans >= 32 slots
  constexpr uint32_t LIGHT = 4;         // matches per probe row expanded by the owning lane; more: by the warp
  extern __shared__ __align__(16) unsigned char s_raw[];
  K* s_keys = reinterpret_cast<K*>(s_raw);                          // NS + 1
  uint32_t* s_w = reinterpret_cast<uint32_t*>(s_keys + (NS + 1));   // NS + 1 (+1 pad)
  uint32_t* s_grp = s_w + (NS + 2);                                 // CAP
  __shared__ uint32_t s_wtot[NW + 1];
  __shared__ uint32_t s_item;
  __shared__ unsigned long long s_gbase;

  const uint32_t lane = lane_id(), warp = warp_id();
  const int mode = a.mode;
  const bool one_side = mode == 5 || mode == 6;
  const uint32_t n_items = a.plan->n_items;

  for (;;) {
    if (threadIdx.x == 0) s_item = atomicAdd(a.queue_head, 1u);
    __syncthreads();
    const uint32_t item_idx = s_item;
    if (item_idx >= n_items) break;
    const WorkItem it = a.items[item_idx];
    const bool multi = it.flags & kItemMulti;

    // table size for this chunk: >= 2x the tuples (load factor <= 0.5) up to NS (then <= CAP/NS)
    uint32_t ns = MIN_NS;
    while (ns < 2 * it.b_cnt && ns < NS) ns <<= 1;
    const uint32_t ns_mask = ns - 1;

    for (uint32_t s = threadIdx.x; s < ns; s += NT) {
      s_keys[s] = KeyTraits<K>::kEmpty;
      s_w[s] = 0;
    }
    if (threadIdx.x == 0) s_w[NS] = 0;
    __syncthreads();

    // ---- build pass A: distinct keys + counts
    for (uint32_t i = threadIdx.x; i < it.b_cnt; i += NT) {
      const K key = a.build[it.b_begin + i].key;
      const uint32_t slot = table_insert<K, NS>(s_keys, ns_mask, key, (hash_key(key) >> a.radix_bits) & ns_mask);
      atomicAdd(&s_w[slot], 1u);
    }
    __syncthreads();

    // ---- exclusive scan of the counts in slot order (warp-contiguous ranges), then the sentinel slot
    {
      const uint32_t per_warp = ns / NW;
      const uint32_t lo = warp * per_warp, hi = lo + per_warp;
      uint32_t sum = 0;
      for (uint32_t s = lo + lane; s < hi; s += 32) sum += s_w[s];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
      if (lane == 0) s_wtot[warp] = sum;
      __syncthreads();
      if (warp == 0) {
        const uint32_t v = lane < NW ? s_wtot[lane] : 0u;
        const uint32_t incl = warp_incl_scan(v);
        if (lane < NW) s_wtot[lane] = incl - v;
        if (lane == 31) s_wtot[NW] = incl;
      }
      __syncthreads();
      uint32_t carry = s_wtot[warp];
      for (uint32_t s0 = lo; s0 < hi; s0 += 32) {
        const uint32_t c = s_w[s0 + lane];
        const uint32_t incl = warp_incl_scan(c);
        s_w[s0 + lane] = ((carry + incl - c) << 16) | c;
        carry += __shfl_sync(0xffffffffu, incl, 31);
      }
      if (threadIdx.x == 0) s_w[NS] = (s_wtot[NW] << 16) | s_w[NS];
    }
    __syncthreads();

    // ---- build pass B: group the row positions by key
    for (uint32_t i = threadIdx.x; i < it.b_cnt; i += NT) {
      const Tuple<K> t = a.build[it.b_begin + i];
      const uint32_t slot = table_find<K, NS>(s_keys, ns_mask, t.key, (hash_key(t.key) >> a.radix_bits) & ns_mask);
      const uint32_t old = atomicAdd(&s_w[slot], 0x10000u);
      s_grp[old >> 16] = t.idx;
    }
    __syncthreads();  // s_w[slot] is now (end of run << 16) | count

    // ---- probe
    for (uint32_t base = 0; base < it.p_cnt; base += NT * V) {
      uint32_t pw[V], pidx[V], outc[V], ex[V];
      bool act[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const uint32_t i = base + warp * (32 * V) + j * 32 + lane;
        act[j] = i < it.p_cnt;
        pw[j] = 0;
        pidx[j] = 0;
        if (act[j]) {
          const Tuple<K> t = a.probe[it.p_begin + i];
          pidx[j] = t.idx;
          const uint32_t slot = table_find<K, NS>(s_keys, ns_mask, t.key, (hash_key(t.key) >> a.radix_bits) & ns_mask);
          pw[j] = slot <= NS ? s_w[slot] : 0u;
        }
      }
      uint32_t running = 0;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const uint32_t cnt = pw[j] & 0xFFFFu;
        uint32_t o;
        if (mode == 0) o = cnt;                                              // Inner
        else if (mode == 1 || mode == 2) o = multi ? cnt : (act[j] ? max(cnt, 1u) : 0u);  // Left/Right
        else if (mode == 5) o = (!multi && cnt) ? 1u : 0u;                    // Semi
        else o = (!multi && act[j] && cnt == 0) ? 1u : 0u;                    // Anti
        if (multi && cnt && mode != 0) {
          const uint32_t pos = it.p_begin + base + warp * (32 * V) + j * 32 + lane;
          atomicOr(a.matched + (pos >> 5), 1u << (pos & 31u));
        }
        outc[j] = o;
        const uint32_t incl = warp_incl_scan(o);
        ex[j] = running + incl - o;
        running += __shfl_sync(0xffffffffu, incl, 31);
      }
      if (lane == 0) s_wtot[warp] = running;
      __syncthreads();
      if (warp == 0) {
        const uint32_t v = lane < NW ? s_wtot[lane] : 0u;
        const uint32_t incl = warp_incl_scan(v);
        if (lane < NW) s_wtot[lane] = incl - v;
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        if (lane == 0) {
          s_wtot[NW] = total;
          s_gbase = total ? atomicAdd(a.out_cursor, static_cast<unsigned long long>(total)) : 0ull;
        }
      }
      __syncthreads();
      const unsigned long long gbase = s_gbase + s_wtot[warp];
      const bool fits = s_gbase + s_wtot[NW] <= a.capacity;  // CTA-uniform; on overflow only the count advances
      if (fits) {
        uint32_t heavy = 0;
#pragma unroll
        for (int j = 0; j < V; ++j) {
          if (outc[j] == 0) continue;
          const unsigned long long pos = gbase + ex[j];
          const uint64_t prid = make_rowid(a.probe_rel, pidx[j]);
          const uint32_t cnt = pw[j] & 0xFFFFu;
          if (one_side) {
            a.out_probe[pos] = prid;
          } else if (cnt == 0) {
            a.out_build[pos] = kNullRowId;  // join_hash.cpp:432-435
            a.out_probe[pos] = prid;
          } else if (cnt <= LIGHT) {
            const uint32_t start = (pw[j] >> 16) - cnt;
            for (uint32_t m = 0; m < cnt; ++m) {
              a.out_build[pos + m] = make_rowid(a.build_rel, s_grp[start + m]);
              a.out_probe[pos + m] = prid;
            }
          } else {
            heavy |= 1u << j;
          }
        }
        if (!one_side) {
#pragma unroll
          for (int j = 0; j < V; ++j) {
            uint32_t hmask = __ballot_sync(0xffffffffu, (heavy >> j) & 1u);
            while (hmask) {
              const int src = __ffs(hmask) - 1;
              hmask &= hmask - 1;
              const uint32_t w = __shfl_sync(0xffffffffu, pw[j], src);
              const uint32_t e = __shfl_sync(0xffffffffu, ex[j], src);
              const uint32_t pi = __shfl_sync(0xffffffffu, pidx[j], src);
              const uint32_t cnt = w & 0xFFFFu, start = (w >> 16) - cnt;
              const unsigned long long pos = gbase + e;
              const uint64_t prid = make_rowid(a.probe_rel, pi);
              for (uint32_t m = lane; m < cnt; m += 32) {
                a.out_build[pos + m] = make_rowid(a.build_rel, s_grp[start + m]);
                a.out_probe[pos + m] = prid;
              }
            }
          }
        }
      }
    }
  }
}

// =====================================================================================================================
// Block-wide compaction: every thread with `flag` emits one result row; one global atomic per call reserves the range.
// =====================================================================================================================
template <int NT>
__device__ __forceinline__ void emit_flagged(bool flag, uint64_t build_rid, uint64_t probe_rid, uint64_t* out_build,
                                             uint64_t* out_probe, unsigned long long* cursor,
                                             unsigned long long capacity, uint32_t* scratch,
                                             unsigned long long* s_base) {
  uint32_t total;
  const uint32_t ex = block_excl_scan(flag ? 1u : 0u, scratch, &total);
  if (threadIdx.x == 0) *s_base = total ? atomicAdd(cursor, static_cast<unsigned long long>(total)) : 0ull;
  __syncthreads();
  const unsigned long long base = *s_base;
  if (flag && base + total <= capacity) {
    if (out_build) out_build[base + ex] = build_rid;
    out_probe[base + ex] = probe_rid;
  }
  __syncthreads();
}

// K6: unmatched/matched decisions for multi partitions, from the bitmap K5 filled over all chunks.
template <typename K, int NT>
__global__ void __launch_bounds__(NT) k_finalize_multi(const JoinArgs<K> a, const WorkItem* __restrict__ fin) {
  __shared__ uint32_t scratch[33];
  __shared__ unsigned long long s_base;
  const uint32_t n_fin = a.plan->n_fin;
  const int mode = a.mode;
  for (uint32_t f = blockIdx.x; f < n_fin; f += gridDim.x) {
    const WorkItem it = fin[f];
    for (uint32_t base = 0; base < it.p_cnt; base += NT) {
      const uint32_t i = base + threadIdx.x;
      bool flag = false;
      uint64_t prid = 0;
      if (i < it.p_cnt) {
        const uint32_t pos = it.p_begin + i;
        const bool hit = (a.matched[pos >> 5] >> (pos & 31u)) & 1u;
        flag = (mode == 5) ? hit : !hit;  // Semi keeps matched rows; Left/Right/Anti the unmatched ones
        if (flag) prid = make_rowid(a.probe_rel, a.probe[pos].idx);
      }
      emit_flagged<NT>(flag, kNullRowId, prid, a.out_build, a.out_probe, a.out_cursor, a.capacity, scratch, &s_base);
    }
  }
}

// K7: HJ_FLAG_SQL_NULL_PROBE — NULL probe keys of a Left/Right join were routed aside by pass 1; each yields (NULL, p).
template <typename K, int NT>
__global__ void __launch_bounds__(NT) k_emit_null_probe(const JoinArgs<K> a, const Tuple<K>* __restrict__ pass1,
                                                        const uint32_t* __restrict__ off_p) {
  __shared__ uint32_t scratch[33];
  __shared__ unsigned long long s_base;
  const uint32_t P = 1u << a.radix_bits;
  const uint32_t lo = off_p[P], hi = off_p[P + 1];
  for (uint32_t base = lo + blockIdx.x * NT; base < hi; base += gridDim.x * NT) {
    const uint32_t i = base + threadIdx.x;
    const bool flag = i < hi;
    const uint64_t prid = flag ? make_rowid(a.probe_rel, pass1[i].idx) : 0;
    emit_flagged<NT>(flag, kNullRowId, prid, a.out_build, a.out_probe, a.out_cursor, a.capacity, scratch, &s_base);
  }
}

// K8: inout[i] = pos_lists[chunk_base[inout[i].chunk_id] + inout[i].chunk_offset]; NULL_ROW_ID stays (join_hash.cpp:584-592).
// Out-of-range references are reported through *error instead of faulting.
__global__ void k_translate(uint64_t* __restrict__ inout, uint64_t n, const uint64_t* __restrict__ pos_lists,
                            const uint64_t* __restrict__ chunk_base, const uint32_t* __restrict__ chunk_size,
                            uint32_t n_chunks, uint32_t* __restrict__ error) {
  for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    const uint64_t r = inout[i];
    const uint32_t chunk = static_cast<uint32_t>(r), off = static_cast<uint32_t>(r >> 32);
    if (off == kInvalid) continue;
    if (chunk >= n_chunks || off >= chunk_size[chunk]) {
      *error = 1;
      continue;
    }
    inout[i] = pos_lists[chunk_base[chunk] + off];
  }
}

}  // namespace hj
