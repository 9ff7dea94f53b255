// hj_aggregate.cuh — ungrouped aggregates over a column that lies in device memory (the consumer after
// the join), typically the output of hj_gather_column: values T[n] + one NULL byte per row.
//
// Replaces, for an Aggregate without GROUP BY columns, Aggregate::_aggregate_column (src/lib/operators/aggregate.cpp:
// 203-252: NULL rows leave the aggregate untouched, every other row updates it and bumps aggregate_count) with the functors
// of :141-201 and the result types of operators/aggregate/aggregate_traits.hpp:16-66:
//   COUNT(col) = non-NULL rows (int64)          SUM(int32/int64) in int64, SUM(float/double) in double
//   MIN / MAX in the column's type              AVG = sum accumulated in double / count (:688-712)
// All of them come out of ONE streaming pass (4/8 B + 1 B read per row): A1 k_aggregate_partial leaves one partial per CTA,
// A2 k_aggregate_final folds the partials in CTA order, so the result does not depend on scheduling (the double sums are
// reproducible run to run; they differ from the reference's row-order sum by rounding only — DESIGN.md §8).
#pragma once
#include <cstdint>
#include <type_traits>

#include "hj_kernels.cuh"

namespace hj {

struct AggPartial {
  unsigned long long n_values;  // non-NULL rows
  long long sum_i;              // integer columns: wrap-around int64 sum like the reference's int64_t accumulator
  double sum_d;                 // floating columns: sum of the values widened to double
  long long min_i, max_i;
  double min_d, max_d;
};

constexpr int kAggThreads = 256;

__device__ __forceinline__ void agg_init(AggPartial& a) {
  a.n_values = 0;
  a.sum_i = 0;
  a.sum_d = 0.0;
  a.min_i = INT64_MAX;
  a.max_i = INT64_MIN;
  a.min_d = __longlong_as_double(0x7FF0000000000000ll);   // +inf
  a.max_d = __longlong_as_double(0xFFF0000000000000ull);  // -inf
}

// b folded into a; a's rows come first (the order matters for nothing but the rounding of sum_d)
__device__ __forceinline__ void agg_merge(AggPartial& a, const AggPartial& b) {
  a.n_values += b.n_values;
  a.sum_i = static_cast<long long>(static_cast<unsigned long long>(a.sum_i) + static_cast<unsigned long long>(b.sum_i));
  a.sum_d += b.sum_d;
  a.min_i = b.min_i < a.min_i ? b.min_i : a.min_i;
  a.max_i = b.max_i > a.max_i ? b.max_i : a.max_i;
  a.min_d = b.min_d < a.min_d ? b.min_d : a.min_d;
  a.max_d = b.max_d > a.max_d ? b.max_d : a.max_d;
}

__device__ __forceinline__ AggPartial agg_shfl_down(const AggPartial& a, int delta) {
  AggPartial b;
  b.n_values = __shfl_down_sync(0xffffffffu, a.n_values, delta);
  b.sum_i = __shfl_down_sync(0xffffffffu, a.sum_i, delta);
  b.sum_d = __shfl_down_sync(0xffffffffu, a.sum_d, delta);
  b.min_i = __shfl_down_sync(0xffffffffu, a.min_i, delta);
  b.max_i = __shfl_down_sync(0xffffffffu, a.max_i, delta);
  b.min_d = __shfl_down_sync(0xffffffffu, a.min_d, delta);
  b.max_d = __shfl_down_sync(0xffffffffu, a.max_d, delta);
  return b;
}

// T: int32_t / int64_t / float / double — the column's type; float / double columns arrive as their bit patterns in
// 4 / 8-byte integers (hj_gather_column), which is the same memory.
template <typename T>
__device__ __forceinline__ void agg_add(AggPartial& a, T v) {
  ++a.n_values;
  if constexpr (std::is_floating_point<T>::value) {
    const double d = static_cast<double>(v);
    a.sum_d += d;
    a.min_d = d < a.min_d ? d : a.min_d;  // value_smaller / value_greater: plain < and > (a NaN never replaces anything)
    a.max_d = d > a.max_d ? d : a.max_d;
  } else {
    const long long x = static_cast<long long>(v);
    a.sum_i = static_cast<long long>(static_cast<unsigned long long>(a.sum_i) + static_cast<unsigned long long>(x));
    a.min_i = x < a.min_i ? x : a.min_i;
    a.max_i = x > a.max_i ? x : a.max_i;
  }
}

// The CTA's partial: warp shuffles, then warp 0's lane 0 folds the warps in order.
__device__ __forceinline__ void agg_block_store(AggPartial a, AggPartial* s_part, AggPartial* __restrict__ dst) {
#pragma unroll
  for (int delta = 16; delta > 0; delta >>= 1) {
    const AggPartial b = agg_shfl_down(a, delta);
    agg_merge(a, b);
  }
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    AggPartial t = s_part[0];
#pragma unroll
    for (int w = 1; w < kAggThreads / 32; ++w) agg_merge(t, s_part[w]);
    *dst = t;
  }
}

// A1. Grid-stride over the rows, four independent loads in flight per thread; one partial per CTA.
template <typename T>
__global__ void __launch_bounds__(kAggThreads) k_aggregate_partial(const T* __restrict__ values, const uint8_t* __restrict__ nulls,
                                                                   unsigned long long n, AggPartial* __restrict__ partials) {
  __shared__ AggPartial s_part[kAggThreads / 32];
  AggPartial a;
  agg_init(a);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * kAggThreads;
  unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * kAggThreads + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    T v[4];
    uint8_t z[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u] = __ldg(values + i + u * stride);
      z[u] = nulls ? __ldg(nulls + i + u * stride) : uint8_t(0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (!z[u]) agg_add<T>(a, v[u]);
  }
  for (; i < n; i += stride)
    if (!(nulls && __ldg(nulls + i))) agg_add<T>(a, __ldg(values + i));
  agg_block_store(a, s_part, partials + blockIdx.x);
}

// A1'. The same over a column given as chunk descriptors (the join's and the scan's input format), without materialising
// it: VALUE / GATHERED chunks stream values + NULL bytes, DICT chunks stream the attribute vector and gather the
// dictionary (NULL = null_value_id). K = integer type the keys are stored in, T = the column's type (K itself, or float /
// double carried as bit patterns after k_gather_refs, as in hj_scan). TILE rows per tile as prepared by prepare_relation.
template <typename K, typename T, int TILE>
__global__ void __launch_bounds__(kAggThreads) k_aggregate_relation(RelMeta rel, AggPartial* __restrict__ partials) {
  __shared__ AggPartial s_part[kAggThreads / 32];
  AggPartial a;
  agg_init(a);
  for (uint32_t tile = blockIdx.x; tile < rel.n_tiles; tile += gridDim.x) {
    const DevSegment s = rel.segs[__ldg(rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
#pragma unroll 4
    for (uint32_t i = threadIdx.x; i < cnt; i += kAggThreads) {
      K raw;
      bool is_null;
      if (s.kind == 1) {
        const uint32_t vid = load_attr(s, r0 + i);
        is_null = vid == s.null_value_id;
        raw = is_null ? K(0) : load_plain_key<K>(s, vid);
      } else {
        is_null = s.nulls && __ldg(s.nulls + r0 + i);
        raw = load_plain_key<K>(s, r0 + i);
      }
      if (!is_null) {
        if constexpr (std::is_same<T, float>::value) agg_add<float>(a, __int_as_float(static_cast<int>(raw)));
        else if constexpr (std::is_same<T, double>::value) agg_add<double>(a, __longlong_as_double(static_cast<long long>(raw)));
        else agg_add<K>(a, raw);
      }
    }
  }
  agg_block_store(a, s_part, partials + blockIdx.x);
}

// A2 (one warp): partials folded in CTA order into out[0].
__global__ void __launch_bounds__(32) k_aggregate_final(const AggPartial* __restrict__ partials, uint32_t n_partials,
                                                        AggPartial* __restrict__ out) {
  // lane l folds partials l, l+32, ... in ascending order, then the lanes are folded in lane order: a fixed tree
  AggPartial a;
  agg_init(a);
  for (uint32_t p = threadIdx.x; p < n_partials; p += 32) agg_merge(a, partials[p]);
#pragma unroll
  for (int delta = 16; delta > 0; delta >>= 1) {
    const AggPartial b = agg_shfl_down(a, delta);
    agg_merge(a, b);
  }
  if (threadIdx.x == 0) out[0] = a;
}

}  // namespace hj
