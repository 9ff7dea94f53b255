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
#include <cstring>
#include <type_traits>

#include "hj_kernels.cuh"

namespace hj {

struct AggPartial {
  unsigned long long n_values;  // non-NULL rows
  long long sum_i;              // integer columns: wrap-around int64 sum like the reference's int64_t accumulator
  double sum_d;                 // floating columns: sum of the values widened to double
  long long min_i, max_i;
  double min_d, max_d;          // over the values that are not NaN
  unsigned long long first;     // floating columns: (row order of the first non-NULL value) << 1 | (that value is NaN)
};
// MIN / MAX and NaN (aggregate.cpp:141-165): the first non-NULL value seeds the aggregate and is then replaced on plain
// `<` / `>`, so a NaN that comes FIRST stays (nothing compares smaller or greater than it) and a NaN anywhere else replaces
// nothing. The streaming kernels therefore keep min / max over the non-NaN values plus `first`; agg_min_max() below applies
// the rule when the result is read.
constexpr unsigned long long kAggNoFirst = ~0ull;

constexpr int kAggThreads = 256;

__device__ __forceinline__ void agg_init(AggPartial& a) {
  a.n_values = 0;
  a.sum_i = 0;
  a.sum_d = 0.0;
  a.min_i = INT64_MAX;
  a.max_i = INT64_MIN;
  a.min_d = __longlong_as_double(0x7FF0000000000000ll);   // +inf
  a.max_d = __longlong_as_double(0xFFF0000000000000ull);  // -inf
  a.first = kAggNoFirst;
}

__host__ __device__ __forceinline__ bool agg_first_is_nan(unsigned long long first) { return first != kAggNoFirst && (first & 1ull); }
__host__ __device__ __forceinline__ double agg_quiet_nan() {
#ifdef __CUDA_ARCH__
  return __longlong_as_double(0x7FF8000000000000ll);
#else
  const long long b = 0x7FF8000000000000ll;
  double d;
  memcpy(&d, &b, 8);
  return d;
#endif
}
// MIN / MAX of a floating column as the reference reports them
__host__ __device__ __forceinline__ void agg_min_max(const AggPartial& a, double* mn, double* mx) {
  const bool nan_first = agg_first_is_nan(a.first);
  *mn = nan_first ? agg_quiet_nan() : a.min_d;
  *mx = nan_first ? agg_quiet_nan() : a.max_d;
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
  a.first = b.first < a.first ? b.first : a.first;
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
  b.first = __shfl_down_sync(0xffffffffu, a.first, delta);
  return b;
}

// T: int32_t / int64_t / float / double — the column's type; float / double columns arrive as their bit patterns in
// 4 / 8-byte integers (hj_gather_column), which is the same memory.
// `order`: position of the row in the column's row order (only its order matters; used for floating columns).
template <typename T>
__device__ __forceinline__ void agg_add(AggPartial& a, T v, unsigned long long order) {
  ++a.n_values;
  if constexpr (std::is_floating_point<T>::value) {
    const double d = static_cast<double>(v);
    a.sum_d += d;
    a.min_d = d < a.min_d ? d : a.min_d;  // value_smaller / value_greater: plain < and > (a NaN never replaces anything)
    a.max_d = d > a.max_d ? d : a.max_d;
    const unsigned long long f = order << 1 | (d != d ? 1ull : 0ull);
    a.first = f < a.first ? f : a.first;
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
      if (!z[u]) agg_add<T>(a, v[u], i + u * stride);
  }
  for (; i < n; i += stride)
    if (!(nulls && __ldg(nulls + i))) agg_add<T>(a, __ldg(values + i), i);
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
        const unsigned long long order = static_cast<unsigned long long>(tile) * TILE + i;  // tiles follow the chunk order
        if constexpr (std::is_same<T, float>::value) agg_add<float>(a, __int_as_float(static_cast<int>(raw)), order);
        else if constexpr (std::is_same<T, double>::value) agg_add<double>(a, __longlong_as_double(static_cast<long long>(raw)), order);
        else agg_add<K>(a, raw, order);
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


// =====================================================================================================================
// GROUP BY one integer column (Aggregate with one group-by column, aggregate.cpp:254-420): a hash table of the distinct
// group keys with the same accumulators as above per group.
//   G1 k_group_accumulate  every CTA pre-aggregates into a small shared-memory table (few groups: no global atomic per
//                          row — same-address global atomics serialise at ~21 ns), rows whose group finds no place there
//                          go to the global table directly; at the end the CTA's groups are merged into the global table
//   G2 k_group_count / k_group_emit  number of groups, then one output record per group (order unspecified, as in the
//                          reference: "we do not guarantee a stable operation with regards to positions")
// NULL is a group key of its own (slot `cap`), and so is the one key value that serves as the EMPTY marker (slot cap + 1).
// Every probe loop is bounded by the table size; a full global table raises *overflow instead of spinning.
// MIN / MAX of float / double columns are kept as order-preserving int64 images of the doubles (sortable()), over the values
// that are not NaN; `first` holds (row order of the group's first non-NULL value) << 1 | (it is NaN), and k_group_emit
// reports NaN for the groups whose first value is one (the reference's rule, see AggPartial).
// The global table is probed at most GroupTable::max_probes slots far; a row that finds no slot raises *overflow, stops its
// thread, and the other threads stop at their next look at the flag: the host then repeats the aggregate with a larger table.
// =====================================================================================================================
constexpr long long kGroupEmpty = static_cast<long long>(0x8000000000000000ull);
constexpr int kGroupSmemSlots = 256;  // power of two; + 2 special slots
constexpr int kGroupSmemProbes = 8;

struct GroupTable {  // structure of arrays, `cap` + 2 entries each (cap = power of two)
  long long* key;
  unsigned long long* rows;    // COUNT(*)
  unsigned long long* values;  // COUNT(column)
  long long* sum_i;
  double* sum_d;
  long long* mn;
  long long* mx;
  unsigned long long* first;
  unsigned long long cap;
  unsigned long long max_probes;  // <= cap
};
constexpr int kGroupArrays = 8;  // 8-byte arrays of cap + 2 entries behind one allocation

__device__ __forceinline__ bool group_overflowed(const uint32_t* overflow) {
  return *reinterpret_cast<const volatile uint32_t*>(overflow) != 0;
}

__host__ __device__ __forceinline__ long long sortable(double d) {
#ifdef __CUDA_ARCH__
  const long long b = __double_as_longlong(d);
#else
  long long b;
  memcpy(&b, &d, 8);
#endif
  return b >= 0 ? b : b ^ 0x7FFFFFFFFFFFFFFFll;
}
__host__ __device__ __forceinline__ double unsortable(long long s) {
  const long long b = s >= 0 ? s : s ^ 0x7FFFFFFFFFFFFFFFll;
#ifdef __CUDA_ARCH__
  return __longlong_as_double(b);
#else
  double d;
  memcpy(&d, &b, 8);
  return d;
#endif
}

__device__ __forceinline__ uint32_t group_hash(long long k) {
  return hash_key(static_cast<int64_t>(k));
}

// slot of `key` in an open-addressing table of `mask + 1` slots, inserting it if absent; at most `max_probes` slots are
// tried; returns mask + 1 ... never: returns 0xFFFFFFFFFFFFFFFF when no slot was found
__device__ __forceinline__ unsigned long long group_slot(long long* keys, unsigned long long mask, long long key, unsigned long long max_probes) {
  unsigned long long s = group_hash(key) & mask;
  for (unsigned long long p = 0; p < max_probes; ++p) {
    const long long prev = atomicCAS(reinterpret_cast<unsigned long long*>(keys + s), static_cast<unsigned long long>(kGroupEmpty),
                                     static_cast<unsigned long long>(key));
    if (prev == kGroupEmpty || prev == key) return s;
    s = (s + 1) & mask;
  }
  return ~0ull;
}

// one row's contribution (T = type of the aggregate column; has_value false: NULL value or COUNT(*) only)
template <typename T>
__device__ __forceinline__ void group_update(unsigned long long* rows, unsigned long long* values, long long* sum_i, double* sum_d,
                                             long long* mn, long long* mx, unsigned long long* first, unsigned long long s,
                                             bool has_value, T v, unsigned long long order) {
  atomicAdd(rows + s, 1ull);
  if (!has_value) return;
  atomicAdd(values + s, 1ull);
  long long x;
  if constexpr (std::is_floating_point<T>::value) {
    const double d = static_cast<double>(v);
    atomicAdd(sum_d + s, d);
    atomicMin(first + s, order << 1 | (d != d ? 1ull : 0ull));
    if (d != d) return;  // a NaN replaces nothing
    x = sortable(d);
  } else {
    x = static_cast<long long>(v);
    atomicAdd(reinterpret_cast<unsigned long long*>(sum_i + s), static_cast<unsigned long long>(x));
  }
  atomicMin(mn + s, x);
  atomicMax(mx + s, x);
}

__global__ void k_group_init(GroupTable t) {
  const unsigned long long n = t.cap + 2;
  for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
    t.key[i] = kGroupEmpty;
    t.rows[i] = 0;
    t.values[i] = 0;
    t.sum_i[i] = 0;
    t.sum_d[i] = 0.0;
    t.mn[i] = INT64_MAX;
    t.mx[i] = INT64_MIN;
    t.first[i] = kAggNoFirst;
  }
}

// G1. GK = type of the group column (int32_t / int64_t), T = type of the aggregate column (values may be nullptr: COUNT(*)).
template <typename GK, typename T>
__global__ void __launch_bounds__(kAggThreads) k_group_accumulate(const GK* __restrict__ group_keys, const uint8_t* __restrict__ group_nulls,
                                                                  const T* __restrict__ values, const uint8_t* __restrict__ value_nulls,
                                                                  unsigned long long n, unsigned long long row_base, GroupTable t,
                                                                  uint32_t* __restrict__ overflow) {
  constexpr int S = kGroupSmemSlots;
  __shared__ long long s_key[S + 2];
  __shared__ unsigned long long s_rows[S + 2], s_values[S + 2], s_first[S + 2];
  __shared__ long long s_sum_i[S + 2], s_mn[S + 2], s_mx[S + 2];
  __shared__ double s_sum_d[S + 2];
  for (int i = threadIdx.x; i < S + 2; i += kAggThreads) {
    s_key[i] = kGroupEmpty;
    s_rows[i] = 0;
    s_values[i] = 0;
    s_sum_i[i] = 0;
    s_sum_d[i] = 0.0;
    s_mn[i] = INT64_MAX;
    s_mx[i] = INT64_MIN;
    s_first[i] = kAggNoFirst;
  }
  __syncthreads();
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * kAggThreads;
  uint32_t trip = 0;
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * kAggThreads + threadIdx.x; i < n; i += stride) {
    if ((trip++ & 15u) == 0 && group_overflowed(overflow)) break;
    const bool key_null = group_nulls && __ldg(group_nulls + i);
    const long long key = key_null ? 0 : static_cast<long long>(__ldg(group_keys + i));
    const bool has_value = values && !(value_nulls && __ldg(value_nulls + i));
    const T v = has_value ? __ldg(values + i) : T(0);
    if (key_null || key == kGroupEmpty) {  // the two special groups always have their shared-memory slot
      group_update<T>(s_rows, s_values, s_sum_i, s_sum_d, s_mn, s_mx, s_first, key_null ? S : S + 1, has_value, v, row_base + i);
      continue;
    }
    const unsigned long long s = group_slot(s_key, S - 1, key, kGroupSmemProbes);
    if (s != ~0ull) {
      group_update<T>(s_rows, s_values, s_sum_i, s_sum_d, s_mn, s_mx, s_first, s, has_value, v, row_base + i);
    } else {  // the CTA's table has no room for this group: straight to the global table
      const unsigned long long g = group_slot(t.key, t.cap - 1, key, t.max_probes);
      if (g == ~0ull) {
        *overflow = 1;
        break;
      }
      group_update<T>(t.rows, t.values, t.sum_i, t.sum_d, t.mn, t.mx, t.first, g, has_value, v, row_base + i);
    }
  }
  __syncthreads();
  if (group_overflowed(overflow)) return;  // the table is discarded
  // merge this CTA's groups into the global table
  for (int i = threadIdx.x; i < S + 2; i += kAggThreads) {
    if (s_rows[i] == 0) continue;
    unsigned long long g;
    if (i >= S) {
      g = t.cap + (i - S);
    } else {
      g = group_slot(t.key, t.cap - 1, s_key[i], t.max_probes);
      if (g == ~0ull) {
        *overflow = 1;
        break;
      }
    }
    atomicAdd(t.rows + g, s_rows[i]);
    if (s_values[i]) {
      atomicAdd(t.values + g, s_values[i]);
      atomicAdd(reinterpret_cast<unsigned long long*>(t.sum_i + g), static_cast<unsigned long long>(s_sum_i[i]));
      atomicAdd(t.sum_d + g, s_sum_d[i]);
      atomicMin(t.mn + g, s_mn[i]);
      atomicMax(t.mx + g, s_mx[i]);
      atomicMin(t.first + g, s_first[i]);
    }
  }
}

struct GroupRecord {  // one per group, in the layout hj_group of include/hj.h
  long long key;
  unsigned long long rows, values;
  long long sum_i;
  double sum_d;
  long long mn, mx;
  unsigned long long key_is_null;
};

// G2. *n_groups must be zero before k_group_emit; out == nullptr: only count.
__global__ void k_group_emit(GroupTable t, GroupRecord* __restrict__ out, unsigned long long* __restrict__ n_groups) {
  const unsigned long long n = t.cap + 2;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * blockDim.x;
  const uint32_t lane = threadIdx.x & 31u;
  // whole warps run the loop together (blockDim is a multiple of 32): one atomic per warp and trip, not per group
  for (unsigned long long base = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + (threadIdx.x & ~31u); base < n; base += stride) {
    const unsigned long long i = base + lane;
    const bool live = i < n && t.rows[i] != 0;
    const uint32_t mask = __ballot_sync(0xffffffffu, live);
    if (mask == 0) continue;
    unsigned long long first = 0;
    if (lane == 0) first = atomicAdd(n_groups, static_cast<unsigned long long>(__popc(mask)));
    first = __shfl_sync(0xffffffffu, first, 0);
    if (!live || !out) continue;
    const unsigned long long dst = first + __popc(mask & ((1u << lane) - 1u));
    GroupRecord r;
    r.key = i == t.cap ? 0 : i == t.cap + 1 ? kGroupEmpty : t.key[i];
    r.rows = t.rows[i];
    r.values = t.values[i];
    r.sum_i = t.sum_i[i];
    r.sum_d = t.sum_d[i];
    const bool nan_first = agg_first_is_nan(t.first[i]);  // only floating aggregate columns set the bit
    r.mn = nan_first ? sortable(agg_quiet_nan()) : t.mn[i];
    r.mx = nan_first ? sortable(agg_quiet_nan()) : t.mx[i];
    r.key_is_null = i == t.cap ? 1 : 0;
    out[dst] = r;
  }
}

// =====================================================================================================================
// GROUP BY several columns: the tables are keyed by a REPRESENTATIVE ROW (the reference's hash_entry.row_id,
// aggregate.cpp:222): a slot holds the index of the row that claimed it; another row belongs to the group iff its group-by
// columns (raw value bits + NULL flags, NULL = NULL) equal that row's. Same accumulators, same two-level scheme, same bounded
// probing as above; GroupTable::key holds row indices (EMPTY = all ones), the two special slots are unused.
// =====================================================================================================================
constexpr int kMaxGroupColumns = 4;
constexpr unsigned long long kRowEmpty = ~0ull;

struct GroupColumns {
  const void* values[kMaxGroupColumns];
  const uint8_t* nulls[kMaxGroupColumns];
  uint32_t width[kMaxGroupColumns];  // 4 or 8 bytes per value
  uint32_t floating[kMaxGroupColumns];  // float / double column: -0.0 groups with +0.0 (the reference compares values with ==)
  uint32_t n;
};

// raw value bits of column c at row r (0 for NULL: a NULL's stored value does not take part in the comparison; the bits of
// +0.0 for a floating -0.0)
__device__ __forceinline__ unsigned long long group_cell(const GroupColumns& g, uint32_t c, unsigned long long r, bool* is_null) {
  *is_null = g.nulls[c] && __ldg(g.nulls[c] + r);
  if (*is_null) return 0;
  const unsigned long long v = g.width[c] == 4 ? static_cast<unsigned long long>(__ldg(static_cast<const uint32_t*>(g.values[c]) + r))
                                               : __ldg(static_cast<const unsigned long long*>(g.values[c]) + r);
  const unsigned long long minus_zero = g.width[c] == 4 ? 0x80000000ull : 0x8000000000000000ull;
  return g.floating[c] && v == minus_zero ? 0ull : v;
}

__device__ __forceinline__ bool group_rows_equal(const GroupColumns& g, unsigned long long a, unsigned long long b) {
  if (a == b) return true;
  for (uint32_t c = 0; c < g.n; ++c) {
    bool na, nb;
    const unsigned long long va = group_cell(g, c, a, &na), vb = group_cell(g, c, b, &nb);
    if (na != nb || va != vb) return false;
  }
  return true;
}

__device__ __forceinline__ uint32_t group_row_hash(const GroupColumns& g, unsigned long long r) {
  uint32_t h = 0x9e3779b9u;
  for (uint32_t c = 0; c < g.n; ++c) {
    bool is_null;
    const unsigned long long v = group_cell(g, c, r, &is_null);
    h = fmix32(h ^ hash_key(static_cast<int64_t>(v)) ^ (is_null ? 0x85ebca6bu : 0u));
  }
  return h;
}

// slot of row r's group in a table of representative rows (`mask` + 1 slots), claiming a slot if the group is new; at most
// max_probes slots are tried, ~0 if none was found
__device__ __forceinline__ unsigned long long group_row_slot(unsigned long long* reps, unsigned long long mask, const GroupColumns& g,
                                                             unsigned long long r, uint32_t h, unsigned long long max_probes) {
  unsigned long long s = h & mask;
  for (unsigned long long p = 0; p < max_probes; ++p) {
    const unsigned long long prev = atomicCAS(reps + s, kRowEmpty, r);
    if (prev == kRowEmpty || group_rows_equal(g, prev, r)) return s;
    s = (s + 1) & mask;
  }
  return ~0ull;
}

__global__ void k_group_init_rows(GroupTable t) {  // like k_group_init with the row-index EMPTY marker
  const unsigned long long n = t.cap + 2;
  for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
    t.key[i] = static_cast<long long>(kRowEmpty);
    t.rows[i] = 0;
    t.values[i] = 0;
    t.sum_i[i] = 0;
    t.sum_d[i] = 0.0;
    t.mn[i] = INT64_MAX;
    t.mx[i] = INT64_MIN;
    t.first[i] = kAggNoFirst;
  }
}

template <typename T>
__global__ void __launch_bounds__(kAggThreads) k_group_accumulate_rows(GroupColumns g, const T* __restrict__ values,
                                                                       const uint8_t* __restrict__ value_nulls, unsigned long long n,
                                                                       GroupTable t, uint32_t* __restrict__ overflow) {
  constexpr int S = kGroupSmemSlots;
  __shared__ unsigned long long s_rep[S];
  __shared__ unsigned long long s_rows[S], s_values[S], s_first[S];
  __shared__ long long s_sum_i[S], s_mn[S], s_mx[S];
  __shared__ double s_sum_d[S];
  for (int i = threadIdx.x; i < S; i += kAggThreads) {
    s_rep[i] = kRowEmpty;
    s_rows[i] = 0;
    s_values[i] = 0;
    s_sum_i[i] = 0;
    s_sum_d[i] = 0.0;
    s_mn[i] = INT64_MAX;
    s_mx[i] = INT64_MIN;
    s_first[i] = kAggNoFirst;
  }
  __syncthreads();
  auto* reps = reinterpret_cast<unsigned long long*>(t.key);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * kAggThreads;
  uint32_t trip = 0;
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * kAggThreads + threadIdx.x; i < n; i += stride) {
    if ((trip++ & 15u) == 0 && group_overflowed(overflow)) break;
    const bool has_value = values && !(value_nulls && __ldg(value_nulls + i));
    const T v = has_value ? __ldg(values + i) : T(0);
    const uint32_t h = group_row_hash(g, i);
    const unsigned long long s = group_row_slot(s_rep, S - 1, g, i, h, kGroupSmemProbes);
    if (s != ~0ull) {
      group_update<T>(s_rows, s_values, s_sum_i, s_sum_d, s_mn, s_mx, s_first, s, has_value, v, i);
    } else {
      const unsigned long long gs = group_row_slot(reps, t.cap - 1, g, i, h, t.max_probes);
      if (gs == ~0ull) {
        *overflow = 1;
        break;
      }
      group_update<T>(t.rows, t.values, t.sum_i, t.sum_d, t.mn, t.mx, t.first, gs, has_value, v, i);
    }
  }
  __syncthreads();
  if (group_overflowed(overflow)) return;  // the table is discarded
  for (int i = threadIdx.x; i < S; i += kAggThreads) {
    if (s_rows[i] == 0) continue;
    const unsigned long long r = s_rep[i];
    const unsigned long long gs = group_row_slot(reps, t.cap - 1, g, r, group_row_hash(g, r), t.max_probes);
    if (gs == ~0ull) {
      *overflow = 1;
      break;
    }
    atomicAdd(t.rows + gs, s_rows[i]);
    if (s_values[i]) {
      atomicAdd(t.values + gs, s_values[i]);
      atomicAdd(reinterpret_cast<unsigned long long*>(t.sum_i + gs), static_cast<unsigned long long>(s_sum_i[i]));
      atomicAdd(t.sum_d + gs, s_sum_d[i]);
      atomicMin(t.mn + gs, s_mn[i]);
      atomicMax(t.mx + gs, s_mx[i]);
      atomicMin(t.first + gs, s_first[i]);
    }
  }
}

}  // namespace hj
