// hj_scan.cuh — TableScan on the device: `column <predicate> literal` (int32 / int64 / float / double
// columns) over the same chunk descriptors the join reads, producing per-chunk PosLists that stay in HBM and feed hj_join
// as REFERENCE segments.
//
// Replaces TableScan::_on_execute (src/lib/operators/table_scan.cpp:84-171) with SingleColumnTableScanImpl
// (operators/table_scan/single_column_table_scan_impl.cpp:23-133) and IsNullTableScanImpl (is_null_table_scan_impl.cpp):
//   S0 k_scan_bounds  the literal's lower/upper bound in every dictionary chunk's dictionary (:135-150: _get_search_value_id)
//   S1 k_scan_mask    one match bit per row + match count per 4096-row tile. Dictionary chunks compare VALUE IDS against
//                     the bounds (:83-133) — the dictionary is never touched per row; value chunks compare the key.
//                     Streaming: reads w_in bytes per row once, writes 1 bit per row.
//   S2 k_scan_prefix  exclusive prefix sum of the tile counts (= every tile's place in the output) + chunk_begin[]
//   S3 k_scan_write   expands the match bits into RowIDs in ascending chunk offset (BaseTableScanImpl::_unary_scan order,
//                     base_table_scan_impl.hpp:36-47): reads 1 bit per row, writes 8 B per match. A REFERENCE chunk emits
//                     its own PosList entry instead (table_scan.cpp:124-131).
// All four are HBM-streaming kernels; algorithmic bytes per row: w_in + 8 B per match.
#pragma once
#include <type_traits>

#include "hj_kernels.cuh"

namespace hj {

constexpr int kScanTile = 4096;             // rows per tile (= tile of prepare_relation for a scan)
constexpr int kScanThreads = 256;
constexpr int kScanWordsPerTile = kScanTile / 32;

// PredicateCondition values (src/lib/types.hpp:166-179) the scan accepts
constexpr int kPredEq = 0, kPredNe = 1, kPredLt = 2, kPredLe = 3, kPredGt = 4, kPredGe = 5, kPredIsNull = 10, kPredIsNotNull = 11;

// S0: one thread per segment. bounds[2 s] = lower_bound, bounds[2 s + 1] = upper_bound of `value` in the dictionary of
// segment s (value ids; dictionary_column.cpp lower_bound/upper_bound); untouched for other kinds.
__global__ void k_scan_bounds(const DevSegment* __restrict__ segs, uint32_t n_segs, long long value, uint32_t* __restrict__ bounds) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_segs) return;
  const DevSegment sg = segs[s];
  if (sg.kind != 1) return;
  const uint32_t n = sg.pad1;  // dictionary entries
  auto entry = [&](uint32_t i) -> long long {
    return sg.key_type == 0 ? static_cast<long long>(__ldg(static_cast<const int32_t*>(sg.values) + i))
                            : static_cast<long long>(__ldg(static_cast<const int64_t*>(sg.values) + i));
  };
  uint32_t lo = 0, hi = n;  // first entry >= value
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (entry(mid) < value) lo = mid + 1; else hi = mid;
  }
  bounds[2 * s] = lo;
  hi = n;  // first entry > value (starts from lo: entries below it are < value)
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (entry(mid) <= value) lo = mid + 1; else hi = mid;
  }
  bounds[2 * s + 1] = lo;
}

template <typename V>
__device__ __forceinline__ bool scan_compare(int pred, V left, V right) {
  switch (pred) {
    case kPredEq: return left == right;
    case kPredNe: return left != right;
    case kPredLt: return left < right;
    case kPredLe: return left <= right;
    case kPredGt: return left > right;
    default: return left >= right;  // kPredGe
  }
}

// The stored key of a value-like chunk as the type the comparison runs in: V = K for integer columns; for float / double
// columns the chunks were rewritten by k_gather_refs into the IEEE bit patterns of the column type (K = int32 / int64
// carrying V = float / double), which are compared as the floating values they encode (NaN matches only !=, -0 == +0,
// as the reference's C++ operators on the column type do).
template <typename K, typename V>
__device__ __forceinline__ V scan_key(const DevSegment& s, uint32_t i) {
  const K raw = load_plain_key<K>(s, i);
  if constexpr (std::is_same<V, float>::value) return __int_as_float(static_cast<int>(raw));
  else if constexpr (std::is_same<V, double>::value) return __longlong_as_double(static_cast<long long>(raw));
  else return raw;
}

// Does row i of segment s match? NULL rows match IS NULL only (base_table_scan_impl.hpp:41, is_null_table_scan_impl.cpp).
template <typename K, typename V>
__device__ __forceinline__ bool scan_row(const DevSegment& s, uint32_t i, int pred, V value, uint32_t lb, uint32_t ub) {
  if (s.kind == 1) {
    const uint32_t vid = load_attr(s, i);
    const bool is_null = vid == s.null_value_id;
    if (pred == kPredIsNull) return is_null;
    if (is_null) return false;
    switch (pred) {  // single_column_table_scan_impl.cpp:89-99
      case kPredIsNotNull: return true;
      case kPredEq: return lb != ub && vid == lb;
      case kPredNe: return lb == ub || vid != lb;
      case kPredLt: return vid < lb;
      case kPredLe: return vid < ub;
      case kPredGt: return vid >= ub;
      default: return vid >= lb;  // kPredGe
    }
  }
  const bool is_null = s.nulls && __ldg(s.nulls + i);
  if (pred == kPredIsNull) return is_null;
  if (is_null) return false;
  if (pred == kPredIsNotNull) return true;
  return scan_compare<V>(pred, scan_key<K, V>(s, i), value);
}

// S1. A warp owns 32 consecutive rows at a time, so a ballot IS the mask word. mask: kScanWordsPerTile words per tile.
template <typename K, typename V>
__global__ void __launch_bounds__(kScanThreads) k_scan_mask(RelMeta rel, int pred, V value, const uint32_t* __restrict__ bounds,
                                                            uint32_t* __restrict__ mask, uint32_t* __restrict__ tile_count) {
  __shared__ uint32_t s_cnt[kScanThreads / 32];
  constexpr uint32_t NW = kScanThreads / 32;
  const uint32_t lane = lane_id(), warp = warp_id();
  for (uint32_t tile = blockIdx.x; tile < rel.n_tiles; tile += gridDim.x) {
    const uint32_t seg = __ldg(rel.tile_seg + tile);
    const DevSegment s = rel.segs[seg];
    const uint32_t r0 = (tile - s.tile_begin) * kScanTile;
    const uint32_t cnt = min(static_cast<uint32_t>(kScanTile), s.n_rows - r0);
    const uint32_t lb = s.kind == 1 ? __ldg(bounds + 2 * seg) : 0u, ub = s.kind == 1 ? __ldg(bounds + 2 * seg + 1) : 0u;
    uint32_t matches = 0;
#pragma unroll 4
    for (uint32_t w = warp; w < kScanWordsPerTile; w += NW) {
      const uint32_t i = w * 32 + lane;
      const bool m = i < cnt && scan_row<K, V>(s, r0 + i, pred, value, lb, ub);
      const uint32_t word = __ballot_sync(0xffffffffu, m);
      if (lane == 0) mask[static_cast<size_t>(tile) * kScanWordsPerTile + w] = word;
      matches += __popc(word);  // the same value in every lane
    }
    if (lane == 0) s_cnt[warp] = matches;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t t = 0;
#pragma unroll
      for (uint32_t k = 0; k < NW; ++k) t += s_cnt[k];
      tile_count[tile] = t;
    }
    __syncthreads();
  }
}

// S2 (one CTA): tile_base[t] = matches in tiles before t (tile_base[n_tiles] = all matches); chunk_begin[s] =
// tile_base[first tile of segment s], chunk_begin[n_segs] = all matches; *n_nonempty = chunks with a match.
template <int NT>
__global__ void __launch_bounds__(NT) k_scan_prefix(const uint32_t* __restrict__ tile_count, uint32_t n_tiles,
                                                    const DevSegment* __restrict__ segs, uint32_t n_segs,
                                                    uint32_t* __restrict__ tile_base, uint32_t* __restrict__ chunk_begin,
                                                    uint32_t* __restrict__ totals /* [0] rows, [1] non-empty chunks */) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry, s_nonempty;
  if (threadIdx.x == 0) {
    s_carry = 0;
    s_nonempty = 0;
  }
  __syncthreads();
  for (uint32_t base = 0; base < n_tiles; base += NT) {
    const uint32_t t = base + threadIdx.x;
    const uint32_t c = t < n_tiles ? tile_count[t] : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(c, scratch, &total);
    const uint32_t carry = s_carry;
    if (t < n_tiles) tile_base[t] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_base[n_tiles] = s_carry;
  __syncthreads();  // tile_base[] written by this CTA is visible to it after the barrier
  uint32_t nonempty = 0;
  for (uint32_t s = threadIdx.x; s < n_segs; s += NT) {
    const DevSegment sg = segs[s];
    const uint32_t nt = (sg.n_rows + kScanTile - 1) / kScanTile;
    const uint32_t b = tile_base[sg.tile_begin], e = tile_base[sg.tile_begin + nt];
    chunk_begin[s] = b;
    nonempty += e > b ? 1u : 0u;
  }
  if (threadIdx.x == 0) chunk_begin[n_segs] = s_carry;
  if (nonempty) atomicAdd(&s_nonempty, nonempty);
  __syncthreads();
  if (threadIdx.x == 0) {
    totals[0] = s_carry;
    totals[1] = s_nonempty;
  }
}

// S3. ref_pos[s]: the PosList of REFERENCE segment s (device memory) or nullptr for a data chunk.
__global__ void __launch_bounds__(kScanThreads) k_scan_write(RelMeta rel, const uint32_t* __restrict__ mask,
                                                             const uint32_t* __restrict__ tile_base,
                                                             const uint64_t* const* __restrict__ ref_pos, uint64_t* __restrict__ out) {
  __shared__ uint32_t s_pre[kScanWordsPerTile];
  __shared__ uint32_t scratch[33];
  static_assert(kScanWordsPerTile <= kScanThreads, "one thread per mask word in the prefix sum");
  constexpr uint32_t NW = kScanThreads / 32;
  const uint32_t lane = lane_id(), warp = warp_id();
  for (uint32_t tile = blockIdx.x; tile < rel.n_tiles; tile += gridDim.x) {
    const uint32_t base = tile_base[tile];
    if (tile_base[tile + 1] == base) continue;  // CTA-uniform: nothing matched in this tile
    const uint32_t seg = __ldg(rel.tile_seg + tile);
    const DevSegment* s = rel.segs + seg;
    const uint32_t r0 = (tile - s->tile_begin) * kScanTile;
    const uint32_t chunk_id = s->chunk_id;
    const uint64_t* pl = ref_pos ? ref_pos[seg] : nullptr;
    const uint32_t* words = mask + static_cast<size_t>(tile) * kScanWordsPerTile;
    const uint32_t c = threadIdx.x < kScanWordsPerTile ? __popc(words[threadIdx.x]) : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(c, scratch, &total);
    if (threadIdx.x < kScanWordsPerTile) s_pre[threadIdx.x] = ex;
    __syncthreads();
    for (uint32_t w = warp; w < kScanWordsPerTile; w += NW) {
      const uint32_t word = words[w];
      if ((word >> lane) & 1u) {
        const uint32_t off = r0 + w * 32 + lane;
        const uint32_t dst = base + s_pre[w] + __popc(word & ((1u << lane) - 1u));
        out[dst] = pl ? pl[off] : pack_rowid(chunk_id, off);
      }
    }
    __syncthreads();
  }
}

}  // namespace hj
