// hj_kernels.cuh — the hand-written sm_100a kernels of libhj. One translation unit (hj_api.cu) includes this.
//
//  K1 k_materialize_hist   replaces materialize_input            (reference join_hash.cpp:203-285)
//  K2 k_offsets(+_l1)      replaces the serial prefix sum         (join_hash.cpp:310-319)
//  K3 k_scatter1/2         replaces partition_radix_parallel      (join_hash.cpp:287-355): pass 1 from the chunks, pass 2
//                          from pass 1's output inside each level-1 partition
//  K4 k_plan               work list for K5 (the reference schedules one JobTask per partition, join_hash.cpp:150,387,485)
//  K5 k_build_probe        replaces build + probe + probe_semi_anti (join_hash.cpp:127-185, 362-466, 468-527)
//  K6 k_finalize_multi     unmatched-row emission for build partitions larger than one shared-memory table
//  K7 k_emit_null_probe    (NULL, p) rows for NULL probe keys under HJ_FLAG_SQL_NULL_PROBE
//  K8 k_compact_plan/_move closes the holes the chunked output allocator leaves at CTA tails (dense PosLists)
//  K9 k_translate          PosList dereferencing of write_output_columns (join_hash.cpp:584-592)
//
// Measured facts behind the design (profiles/r01_microbench_*.txt, profiles/r01_ncu_v0_*.md):
//   * shared-memory atomics run at ~7.5 lanes/clk/SM — cheap enough to rank every tuple with one;
//   * atomics on ONE global address serialise at ~21 ns each, so nothing on a hot path may hit a single global
//     counter per tile: pass 1 gets private cursors from per-CTA histograms, pass 2 permutes its tile order so that
//     concurrent CTAs touch different cursors, the result writer reserves 128 Ki-row chunks per CTA;
//   * the kernels are issue/latency-bound, not DRAM-bound, so the hot loops are kept short (run-wise copies, ballot
//     prefix sums, register prefetch of the next tile).
#pragma once
#include "hj_device.cuh"

namespace hj {

constexpr int kMaxDigits = 256;  // fan-out per scatter pass (8 bits)

// =====================================================================================================================
// Row loaders shared by K1 and pass 1.
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

// K0: tile -> segment map, filled on the device from the segments' tile ranges (uploading it cost 0.7 ms per C2 join).
__global__ void k_fill_tile_seg(const DevSegment* __restrict__ segs, uint32_t n_segs, uint32_t tile_rows,
                                uint32_t* __restrict__ tile_seg) {
  for (uint32_t sidx = blockIdx.x; sidx < n_segs; sidx += gridDim.x) {
    const uint32_t t0 = segs[sidx].tile_begin, nt = (segs[sidx].n_rows + tile_rows - 1) / tile_rows;
    for (uint32_t t = threadIdx.x; t < nt; t += blockDim.x) tile_seg[t0 + t] = sidx;
  }
}

// =====================================================================================================================
// K1: materialize + histogram. Persistent "virtual CTA" c owns tiles c, c+G, c+2G, ... (G = gridDim.x); pass 1 uses the
// same ownership, which is what makes its write cursors private.
//   DICT chunks: attr -> dictionary gather (served by L2; this gather is the L1tex-bound part of the whole join) ->
//   decoded key written once to `keys_mat` + one validity bit per row; VALUE/GATHERED chunks are only read.
//   Every kept row increments a full-resolution (2^radix_bits bins) shared-memory histogram. At the end the CTA adds it
//   to the global histogram and stores its own level-1 digit counts to hist1[d * G + c].
// block: NT threads, dynamic smem: (2^radix_bits + 1) * 4 bytes.
// =====================================================================================================================
// One physical CTA of NT threads hosts VC virtual CTAs of NT/VC threads that share ONE full-resolution histogram (it is
// summed globally anyway) and keep their own level-1 digit counts. With VC = 4 and 1024 threads all 4 x #SM virtual CTAs
// are resident at once, so they sweep the chunks together and a chunk's dictionary is pulled into L2 once instead of
// once per wave (v1: 2 CTAs/SM because of the 64 KB histogram => two waves => 15.6 GB of DRAM reads for 1 G probe rows).
template <typename K, int NT, int VC, int TILE>
__global__ void __launch_bounds__(NT) k_materialize_hist(RelMeta rel, K* __restrict__ keys_mat,
                                                         uint32_t* __restrict__ valid_mask,
                                                         uint32_t* __restrict__ hist, uint32_t* __restrict__ hist1,
                                                         int radix_bits, int bits2, int null_policy) {
  constexpr int VT = NT / VC;         // threads per virtual CTA
  constexpr int ITEMS = TILE / VT;
  extern __shared__ uint32_t s_hist[];
  __shared__ uint32_t s_dcnt[VC > 1 ? VC : 1][kMaxDigits + 1];
  const uint32_t P = 1u << radix_bits;
  const uint32_t n_bins = P + 1u;  // last bin: NULL rows routed aside (kNullRoute)
  const uint32_t pmask = P - 1u;
  const uint32_t F1 = P >> bits2;
  const uint32_t v = threadIdx.x / VT, vt = threadIdx.x % VT;
  const uint32_t G = gridDim.x * VC, c = blockIdx.x * VC + v;  // virtual grid / this thread's virtual CTA
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) s_hist[b] = 0;
  if (VC > 1)
    for (uint32_t d = vt; d <= F1; d += VT) s_dcnt[v][d] = 0;
  __syncthreads();

  for (uint32_t tile = c; tile < rel.n_tiles; tile += G) {
    const uint32_t seg_idx = __ldg(rel.tile_seg + tile);
    const DevSegment s = rel.segs[seg_idx];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
    K key[ITEMS];
    uint32_t vbits = 0;  // bit u: row u*VT+vt is valid (non-NULL)
    if (s.kind == 1) {
      uint32_t vid[ITEMS];
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        const uint32_t i = u * VT + vt;
        vid[u] = i < cnt ? load_attr(s, r0 + i) : s.null_value_id;
      }
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        const bool valid = vid[u] != s.null_value_id;  // rows past the end carry null_value_id
        key[u] = valid ? load_plain_key<K>(s, vid[u]) : K(0);  // NULL decodes to T{} (dictionary_column_iterable.hpp:80)
        vbits |= (valid ? 1u : 0u) << u;
      }
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        const uint32_t i = u * VT + vt;
        if (i < cnt) keys_mat[s.row_begin + r0 + i] = key[u];
        const uint32_t bits = __ballot_sync(0xffffffffu, (vbits >> u) & 1u);
        // one warp instruction covers 32 consecutive rows starting at a multiple of 32
        if (lane_id() == 0 && i < cnt) valid_mask[s.mask_word_begin + ((r0 + i) >> 5)] = bits;
      }
    } else {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        const uint32_t i = u * VT + vt;
        key[u] = K(0);
        if (i < cnt) {
          key[u] = load_plain_key<K>(s, r0 + i);
          const bool valid = !(s.nulls && __ldg(s.nulls + r0 + i));
          if (s.kind == 2 && !valid) key[u] = K(0);  // GATHERED NULL yields T{} (reference_column_iterable.hpp:60)
          vbits |= (valid ? 1u : 0u) << u;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      if (u * VT + vt >= cnt) continue;
      if (((vbits >> u) & 1u) || null_policy == kNullAsValue) {
        const uint32_t p = hash_key(key[u]) & pmask;
        atomicAdd(&s_hist[p], 1u);
        if (VC > 1) atomicAdd(&s_dcnt[v][p >> bits2], 1u);
      } else if (null_policy == kNullRoute) {
        atomicAdd(&s_hist[P], 1u);
        if (VC > 1) atomicAdd(&s_dcnt[v][F1], 1u);
      }
    }
  }
  __syncthreads();
  for (uint32_t b = threadIdx.x; b < n_bins; b += NT) {
    const uint32_t cb = s_hist[b];
    if (cb) atomicAdd(hist + b, cb);
  }
  // the virtual CTAs' level-1 digit counts (digit F1 = routed NULLs)
  if (VC > 1) {
    for (uint32_t d = vt; d <= F1; d += VT) hist1[d * G + c] = s_dcnt[v][d];
  } else {
    const uint32_t F2 = 1u << bits2;
    for (uint32_t d = threadIdx.x; d <= F1; d += NT) {
      uint32_t sum = 0;
      if (d == F1) {
        sum = s_hist[P];
      } else {
        for (uint32_t q = 0; q < F2; ++q) sum += s_hist[(d << bits2) + ((q + d) & (F2 - 1u))];  // rotated: no bank conflicts
      }
      hist1[d * G + c] = sum;
    }
  }
}

// =====================================================================================================================
// K2: offsets. One CTA turns the global histogram into
//   part_off[P+2] : start of every final partition, part_off[P] = start of the routed-NULL region, part_off[P+1] = total
//   cur2[P]       : pass-2 write cursors
//   l1_tiles[F1+1]: prefix sum of ceil(level-1 partition size / tile2) — tile -> partition map for pass 2
// =====================================================================================================================
template <int NT>
__global__ void __launch_bounds__(NT) k_offsets(const uint32_t* __restrict__ hist, uint32_t* __restrict__ part_off,
                                                uint32_t* __restrict__ cur2, uint32_t* __restrict__ l1_tiles,
                                                int radix_bits, int bits2, uint32_t tile2) {
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
      if (b <= P) cur2[b] = ex;  // [P]: cursor of the routed-NULL region (fused pass 2)
    }
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) part_off[P + 1] = s_carry;
  __syncthreads();
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

// K2b: per level-1 digit d (one CTA each, digit F1 = routed NULLs), turns the per-CTA counts hist1[d*G + c] in place into
// pass 1's private write cursors: start of the digit's region + exclusive prefix over the virtual CTAs c.
template <int NT>
__global__ void __launch_bounds__(NT) k_offsets_l1(uint32_t* __restrict__ hist1, const uint32_t* __restrict__ part_off,
                                                   int radix_bits, int bits2, uint32_t G) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry;
  const uint32_t P = 1u << radix_bits, F1 = P >> bits2, d = blockIdx.x;
  if (threadIdx.x == 0) s_carry = d < F1 ? part_off[d << bits2] : part_off[P];
  __syncthreads();
  for (uint32_t base = 0; base < G; base += NT) {
    const uint32_t c = base + threadIdx.x;
    const uint32_t v = c < G ? hist1[d * G + c] : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(v, scratch, &total) + s_carry;
    if (c < G) hist1[d * G + c] = ex;
    __syncthreads();
    if (threadIdx.x == 0) s_carry += total;
    __syncthreads();
  }
}

// =====================================================================================================================
// K3: radix scatter. Per tile of NT*ITEMS tuples:
//   rank every tuple inside its digit with one shared-memory atomic, scan the digit counts, reorder the tile in shared
//   memory so that each digit is a contiguous run, then copy the runs out warp by warp (consecutive lanes on consecutive
//   addresses; no per-tuple address lookup in the copy loop).
// Within a partition the tuple order is arbitrary (the reference guarantees none: join_hash.hpp:19-20).
// dynamic smem: NT*ITEMS*sizeof(Tuple<K, PAY>).
// =====================================================================================================================
template <typename K, bool PAY>
struct ScatterArgs {
  RelMeta rel;                  // pass 1
  const K* keys_mat;            // pass 1
  const uint32_t* valid_mask;   // pass 1
  const uint32_t* base1;        // pass 1: private cursors [d * G + c]
  const Tuple<K, PAY>* in;           // pass 2
  Tuple<K, PAY>* out;
  uint32_t* cur2;               // pass 2: global cursors, one per final partition
  const uint32_t* part_off;     // pass 2
  const uint32_t* l1_tiles;     // pass 2
  int radix_bits;
  int bits2;                    // low bits handled by pass 2 (0: pass 1 is the only pass)
  int null_policy;
};

constexpr uint32_t kNoDigit = 0xFFFFFFFFu;

// Staging + copy-out shared by both passes. A tuple that got (digit d, rank r inside the digit) is staged at
// s_loff[d] + r; the copy-out is then one flat, fully unrolled loop: consecutive lanes read consecutive staged slots (one
// digit's run is contiguous, so the stores coalesce per run), re-derive the digit from the key and add the digit's
// (global - staged) delta, with no per-run control flow and no per-tuple destination array in shared memory. v1 copied run by run (a warp per digit) and spent ~120 thread-instructions per tuple on loop
// overhead for the ~32-tuple runs (profiles/r01_ncu_v1_c2_full_summary.md).
// PRE: the key field of an int32 tuple already IS the hash (tuples that came through the multi-GPU push exchange)
template <bool PRE, typename K>
__device__ __forceinline__ uint32_t tuple_hash(K key) {
  if constexpr (PRE && sizeof(K) == 4) return static_cast<uint32_t>(key);
  else return hash_key(key);
}

template <typename K, bool PAY, int NT, int ITEMS, bool PRE = false>
__device__ __forceinline__ void copy_tile_out(const Tuple<K, PAY>* stage, const uint32_t* s_delta, Tuple<K, PAY>* out, uint32_t n,
                                              uint32_t dmask, int shift, uint32_t null_start, uint32_t null_digit) {
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t i = j * NT + threadIdx.x;
    if (i < n) {
      const Tuple<K, PAY> t = stage[i];
      const uint32_t d = i >= null_start ? null_digit : (tuple_hash<PRE>(t.key) & dmask) >> shift;
      out[s_delta[d] + i] = t;  // s_delta[d] = global index of the digit's run - its offset in the staged tile (mod 2^32)
    }
  }
}

// Pass 1: persistent virtual CTA c = blockIdx.x owns tiles c, c+G, ... exactly as in K1, so the cursors computed from K1's
// per-CTA counts (base1) are private: no global atomics, deterministic placement. Digit = high (radix_bits - bits2) bits
// of the partition index; routed NULL rows form run number F1.
// dynamic smem: TILE * sizeof(Tuple<K, PAY>).
template <typename K, bool PAY, int NT, int ITEMS>
__global__ void __launch_bounds__(NT, 1024 / NT) k_scatter1(const ScatterArgs<K, PAY> a) {
  constexpr uint32_t TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  Tuple<K, PAY>* stage = reinterpret_cast<Tuple<K, PAY>*>(s_raw);
  __shared__ uint32_t s_cnt[kMaxDigits + 1], s_cur[kMaxDigits + 1];
  __shared__ uint32_t s_loff[kMaxDigits + 1], s_delta[kMaxDigits + 1];
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_total, s_null_start;

  const uint32_t pmask = (1u << a.radix_bits) - 1u;
  const uint32_t F1 = 1u << (a.radix_bits - a.bits2);  // <= 256 == regular digits; run F1 = routed NULLs
  const uint32_t G = gridDim.x;
  const int bits2 = a.bits2;
  const int null_policy = a.null_policy;
  for (uint32_t d = threadIdx.x; d <= F1; d += NT) s_cur[d] = a.base1[d * G + blockIdx.x];

  for (uint32_t tile = blockIdx.x; tile < a.rel.n_tiles; tile += G) {
    const DevSegment seg = a.rel.segs[__ldg(a.rel.tile_seg + tile)];
    const uint32_t r0 = (tile - seg.tile_begin) * TILE;
    const uint32_t n_in = min(TILE, seg.n_rows - r0);
    for (uint32_t d = threadIdx.x; d <= F1; d += NT) s_cnt[d] = 0;
    __syncthreads();  // also orders the previous tile's copy-out before this tile's staging

    K key[ITEMS];
    uint32_t vbits = 0;  // bit j: row j*NT+tid is a valid (non-NULL) row
    if (seg.kind == 1) {
      const K* km = a.keys_mat + seg.row_begin + r0;
      const uint32_t* vm = a.valid_mask + seg.mask_word_begin + (r0 >> 5);
#pragma unroll
      for (int j = 0; j < ITEMS; ++j) {
        const uint32_t i = j * NT + threadIdx.x;
        key[j] = K(0);
        if (i < n_in) {
          key[j] = __ldg(km + i);
          vbits |= ((__ldg(vm + (i >> 5)) >> (i & 31u)) & 1u) << j;
        }
      }
    } else {
      const uint8_t* nl = seg.nulls ? seg.nulls + r0 : nullptr;
#pragma unroll
      for (int j = 0; j < ITEMS; ++j) {
        const uint32_t i = j * NT + threadIdx.x;
        key[j] = K(0);
        if (i < n_in) {
          key[j] = load_plain_key<K>(seg, r0 + i);
          const bool valid = !(nl && __ldg(nl + i));
          if (seg.kind == 2 && !valid) key[j] = K(0);
          vbits |= (valid ? 1u : 0u) << j;
        }
      }
    }
    // rows that arrived through the multi-GPU exchange carry their global RowID in the tuple (idx = ChunkID, pad =
    // ChunkOffset) instead of a position: positions of received rows mean nothing to the RowID mapping. The RowIDs are
    // loaded together with the keys so that the tile pays the DRAM latency once.
    constexpr bool kHasPayload = sizeof(Tuple<K, PAY>) > 8;
    uint64_t rid[kHasPayload ? ITEMS : 1];
    const bool exchanged = kHasPayload && seg.kind == kSegExchanged;
    if (exchanged) {
      const uint64_t* rids = static_cast<const uint64_t*>(seg.attr) + r0 + threadIdx.x;
#pragma unroll
      for (int j = 0; j < (kHasPayload ? ITEMS : 1); ++j) rid[j] = j * NT + threadIdx.x < n_in ? __ldg(rids + j * NT) : 0ull;
    }
    uint32_t packed[ITEMS];  // digit << 16 | rank inside the digit, or kNoDigit
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      const uint32_t i = j * NT + threadIdx.x;
      packed[j] = kNoDigit;
      if (i < n_in) {
        uint32_t d = kNoDigit;
        if (((vbits >> j) & 1u) || null_policy == kNullAsValue) d = (hash_key(key[j]) & pmask) >> bits2;
        else if (null_policy == kNullRoute) d = F1;
        if (d != kNoDigit) packed[j] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
      }
    }
    __syncthreads();
    {
      // F1 <= 256 <= NT regular digits are scanned one per thread; the NULL run goes last
      const uint32_t d = threadIdx.x;
      const uint32_t c = d < F1 ? s_cnt[d] : 0u;
      uint32_t total;
      const uint32_t ex = block_excl_scan(c, scratch, &total);
      if (d < F1) {
        s_loff[d] = ex;
        s_delta[d] = s_cur[d] - ex;
        s_cur[d] += c;
      }
      if (d == 0) {
        const uint32_t cn = s_cnt[F1];
        s_loff[F1] = total;
        s_delta[F1] = s_cur[F1] - total;
        s_cur[F1] += cn;
        s_total = total + cn;
        s_null_start = total;
      }
    }
    __syncthreads();
    const uint32_t idx0 = seg.row_begin + r0 + threadIdx.x;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
      if (packed[j] != kNoDigit) {
        Tuple<K, PAY> t;
        t.key = key[j];
        t.idx = idx0 + j * NT;
        if constexpr (kHasPayload) {
          t.pad = 0;
          if (exchanged) {
            t.idx = static_cast<uint32_t>(rid[j]);
            t.pad = static_cast<uint32_t>(rid[j] >> 32);
          }
        }
        stage[s_loff[packed[j] >> 16] + (packed[j] & 0xFFFFu)] = t;
      }
    }
    __syncthreads();
    copy_tile_out<K, PAY, NT, ITEMS>(stage, s_delta, a.out, s_total, pmask, bits2, s_null_start, F1);
  }
}

// Pass 2: one tile per CTA inside one level-1 partition; digit = low bits2 bits; the tile's slice of each final partition
// is reserved with one global atomic per non-empty digit. blockIdx is mapped to the tile through a multiplicative
// permutation so that the CTAs resident at any moment work on different level-1 partitions, i.e. on different cursors
// (same-address atomics serialise at ~21 ns; profiles/r01_microbench_atomics_vs_match.txt).
// dynamic smem: TILE * sizeof(Tuple<K, PAY>).
template <typename K, bool PAY, int NT, int ITEMS>
__global__ void __launch_bounds__(NT, 1024 / NT) k_scatter2(const ScatterArgs<K, PAY> a) {
  constexpr uint32_t TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  Tuple<K, PAY>* stage = reinterpret_cast<Tuple<K, PAY>*>(s_raw);
  __shared__ uint32_t s_cnt[kMaxDigits], s_loff[kMaxDigits], s_delta[kMaxDigits];
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_l1[kMaxDigits + 1];

  const uint32_t F1 = 1u << (a.radix_bits - a.bits2), F2 = 1u << a.bits2;
  for (uint32_t d = threadIdx.x; d <= F1; d += NT) s_l1[d] = a.l1_tiles[d];
  for (uint32_t d = threadIdx.x; d < F2; d += NT) s_cnt[d] = 0;
  __syncthreads();
  const uint32_t total_tiles = s_l1[F1];
  if (blockIdx.x >= total_tiles) return;  // grid is an upper bound
  const uint32_t t = static_cast<uint32_t>((static_cast<uint64_t>(blockIdx.x) * 2654435761ull) % total_tiles);
  uint32_t lo = 0, hi = F1;  // largest d1 with s_l1[d1] <= t
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (s_l1[mid] <= t) lo = mid; else hi = mid;
  }
  const uint32_t d1 = lo;
  const uint32_t p_hi = a.part_off[(d1 + 1) << a.bits2];
  const uint32_t in_base = a.part_off[d1 << a.bits2] + (t - s_l1[d1]) * TILE;
  const uint32_t n_in = min(TILE, p_hi - in_base);
  uint32_t* cursors = a.cur2 + (d1 << a.bits2);
  const Tuple<K, PAY>* in = a.in + in_base;

  Tuple<K, PAY> tup[ITEMS];
  uint32_t packed[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t i = j * NT + threadIdx.x;
    if (i < n_in) tup[j] = ld_stream(in + i);
  }
#pragma unroll
  for (int j = 0; j < ITEMS; ++j) {
    const uint32_t i = j * NT + threadIdx.x;
    packed[j] = kNoDigit;
    if (i < n_in) {
      const uint32_t d = hash_key(tup[j].key) & (F2 - 1u);
      packed[j] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
    }
  }
  __syncthreads();
  {
    const uint32_t d = threadIdx.x;
    const uint32_t c = d < F2 ? s_cnt[d] : 0u;
    uint32_t total;
    const uint32_t ex = block_excl_scan(c, scratch, &total);
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
  copy_tile_out<K, PAY, NT, ITEMS>(stage, s_delta, a.out, n_in, F2 - 1u, 0, 0xFFFFFFFFu, 0u);
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
// Result writer: chunked allocation. A writer (one WARP of K5's probe loop, one CTA of K6/K7) owns one open chunk of the
// output arrays at a time and sub-allocates its batches from it with plain register/shared arithmetic; only when the
// chunk is exhausted does it touch the global chunk counter (one atomic per `chunk_rows` result rows per writer instead
// of one per batch; same-address atomics serialise at ~21 ns). A batch that does not fit fills the open chunk EXACTLY and
// continues in freshly reserved chunk(s), so every chunk except each writer's last one is dense; k_compact_* closes
// those last holes. The chunk size is chosen per join (2^chunk_shift rows, 256..16 Ki) from the expected result size so
// that the slack of the ~4.9 K writers stays small for small joins.
// =====================================================================================================================
constexpr uint32_t kMaxTails = 8192;  // >= writers: 148 SMs x 32 warps (K5) + 148 (K6) + 148 (K7)

struct OutTail {            // the unfilled end [pos, pos + left) of a writer's last chunk
  unsigned long long pos;
  unsigned long long left;
};

struct OutState {           // device-global result bookkeeping, zeroed before every probe pass
  unsigned long long n_chunks;     // chunks handed out
  unsigned long long n_rows;       // result rows emitted (exact even when the buffers overflow)
  uint32_t n_tails;
  uint32_t pad;
};

struct OutWriter {          // one writer's state (registers for a warp, shared memory for a CTA)
  unsigned long long base;  // next free position of the open chunk
  unsigned long long left;  // free rows in the open chunk
  unsigned long long rows;  // rows this writer emitted
  // mapping of the current batch: batch-relative index i -> i < first ? base0 + i : base1 + (i - first)
  unsigned long long base0, base1;
  uint32_t first;
  uint32_t pad;
};

__device__ __forceinline__ void out_writer_init(OutWriter* w) {
  w->base = 0; w->left = 0; w->rows = 0; w->base0 = 0; w->base1 = 0; w->first = 0;
}
// one thread: reserves `total` rows for the current batch
__device__ __forceinline__ void out_writer_reserve(OutWriter* w, OutState* g, unsigned long long total, int chunk_shift) {
  w->rows += total;
  w->base0 = w->base;
  if (total <= w->left) {
    w->first = 0xFFFFFFFFu;  // whole batch in the open chunk
    w->base += total;
    w->left -= total;
  } else {
    const unsigned long long first = w->left, rest = total - first;
    const unsigned long long k = (rest + (1ull << chunk_shift) - 1) >> chunk_shift;
    const unsigned long long nb = atomicAdd(&g->n_chunks, k) << chunk_shift;
    w->first = static_cast<uint32_t>(first);  // < chunk rows
    w->base1 = nb;
    w->base = nb + rest;
    w->left = (k << chunk_shift) - rest;
  }
}
// whole warp, writer state replicated in every lane: lane 0 takes fresh chunks, all lanes learn their position
__device__ __forceinline__ void warp_writer_reserve(OutWriter* w, OutState* g, unsigned long long total, int chunk_shift) {
  w->rows += total;
  w->base0 = w->base;
  if (total <= w->left) {  // warp-uniform
    w->first = 0xFFFFFFFFu;
    w->base += total;
    w->left -= total;
  } else {
    const unsigned long long first = w->left, rest = total - first;
    const unsigned long long k = (rest + (1ull << chunk_shift) - 1) >> chunk_shift;
    unsigned long long nb = 0;
    if (lane_id() == 0) nb = atomicAdd(&g->n_chunks, k) << chunk_shift;
    nb = __shfl_sync(0xffffffffu, nb, 0);
    w->first = static_cast<uint32_t>(first);
    w->base1 = nb;
    w->base = nb + rest;
    w->left = (k << chunk_shift) - rest;
  }
}
__device__ __forceinline__ unsigned long long out_pos(unsigned long long base0, unsigned long long base1, uint32_t first,
                                                      unsigned long long i) {
  return i < first ? base0 + i : base1 + (i - first);
}
// one thread, when the writer is done
__device__ __forceinline__ void out_writer_finish(const OutWriter* w, OutState* g, OutTail* tails) {
  if (w->rows) atomicAdd(&g->n_rows, w->rows);
  if (w->left) {
    const uint32_t t = atomicAdd(&g->n_tails, 1u);
    if (t < kMaxTails) tails[t] = OutTail{w->base, w->left};
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
// Probe: each thread looks up V tuples of the current tile while the next tile's tuples are already in flight; per-warp
// prefix sums of the match counts (ballot+popc when no row has > 1 match) place the rows, OutWriter reserves the range,
// and the two PosLists are written with consecutive lanes on consecutive addresses; a probe row with many matches is
// expanded by its whole warp.
// block: NT threads (1 CTA/SM), dynamic smem: (NS+1)*sizeof(K) + (NS+2)*4 + CAP*4.
// =====================================================================================================================
// Tuples that came through the multi-GPU push exchange (hj_xpush.cuh) name their row as (source rank, position in that
// rank's shard). Shards that take this path are regular — chunks of 2^shift rows (the last one may be shorter) with
// consecutive chunk ids — so the RowID is a shift, a mask and an add.
struct XchMap {
  uint32_t first_chunk[16];
  int32_t shift[16];
};
__device__ __forceinline__ uint64_t xch_rowid(const XchMap& m, uint32_t rank, uint32_t pos) {
  const int sh = m.shift[rank & 15u];
  return pack_rowid(m.first_chunk[rank & 15u] + (pos >> sh), pos & ((1u << sh) - 1u));
}

template <typename K, bool PAY>
struct JoinArgs {
  const Tuple<K, PAY>* build;
  const Tuple<K, PAY>* probe;
  const WorkItem* items;
  const PlanOut* plan;
  uint32_t* queue_head;
  RelMeta build_rel, probe_rel;
  uint64_t* out_build;                 // nullptr for Semi/Anti
  uint64_t* out_probe;
  OutState* out_state;
  OutTail* out_tails;
  unsigned long long capacity;         // rows the output arrays can hold (including chunk slack)
  uint32_t* matched;                   // bit per probe tuple position; written for multi items only
  int radix_bits;
  int mode;                            // hj_join_mode
  int chunk_shift;                     // result chunks are 2^chunk_shift rows
  int payload;                         // 1: tuples carry their RowID (idx, pad) — local join of hj_join_sharded; 2: (rank, position)
  int bits2;                           // payload 2, int32 keys: the source rank sits in hash bits [bits2, radix_bits)
  XchMap xmap[2];                      // payload 2: [0] build side, [1] probe side
};

// source rank of a push-exchanged tuple: int64 tuples carry it in `pad`, int32 tuples in the level-1 digit bits of their
// hash (redundant once pass 1 has grouped by them)
template <typename K, bool PAY>
__device__ __forceinline__ uint32_t xch_rank(const Tuple<K, PAY>& t, int bits2, int radix_bits) {
  if constexpr (sizeof(K) == 8) return t.pad;
  else return (static_cast<uint32_t>(t.key) >> bits2) & ((1u << (radix_bits - bits2)) - 1u);
}

template <typename K> struct KeyTraits;
template <> struct KeyTraits<int32_t> { static constexpr int32_t kEmpty = INT32_MIN; };
template <> struct KeyTraits<int64_t> { static constexpr int64_t kEmpty = INT64_MIN; };

__device__ __forceinline__ int32_t smem_cas(int32_t* addr, int32_t expected, int32_t desired) {
  return atomicCAS(reinterpret_cast<int*>(addr), expected, desired);
}
__device__ __forceinline__ int64_t smem_cas(int64_t* addr, int64_t expected, int64_t desired) {
  return static_cast<int64_t>(atomicCAS(reinterpret_cast<unsigned long long*>(addr),
                                        static_cast<unsigned long long>(expected),
                                        static_cast<unsigned long long>(desired)));
}

// slot of `key`, claiming an empty one if absent. Terminates: the table always keeps empty slots (CAP < NS).
template <typename K, int NS>
__device__ __forceinline__ uint32_t table_insert(K* keys, uint32_t ns_mask, K key, uint32_t slot) {
  if (key == KeyTraits<K>::kEmpty) return NS;
  for (;;) {
    K cur = *reinterpret_cast<volatile K*>(keys + slot);
    if (cur == KeyTraits<K>::kEmpty) {
      cur = smem_cas(keys + slot, KeyTraits<K>::kEmpty, key);
      if (cur == KeyTraits<K>::kEmpty) return slot;
    }
    if (cur == key) return slot;
    slot = (slot + 1) & ns_mask;
  }
}

// Linear probing starts at a 16-byte aligned slot (the low kBucketShift<K> bits of the start are zero), so that a lookup
// tests a whole 16-byte bucket (4 int32 / 2 int64 keys) per shared-memory load: with a load factor <= 0.75 almost every
// probe row is answered by its first bucket and the lanes of a warp stay together (with one-slot steps a warp made as
// many loop trips as the longest cluster walk among its 32 lanes: 2.8 extra trips per row on C2, ncu source counters in
// profiles/r01_ncu_reentry_c2_full_summary.md).
#ifndef HJ_TUNE_OUT_CS
#define HJ_TUNE_OUT_CS 0  // 1: result rows stored with st.global.cs (evict-first)
#endif
#ifndef HJ_TUNE_BP_NEXT
// 1: k_build_probe takes its NEXT work item while the current one is joined (queue pop during the build, descriptor into
// shared memory, the item's build tuples and first probe stripe requested into L2 by two bulk prefetches when the probe
// phase starts), so that an item starts with a shared-memory read instead of an atomic + a dependent load + DRAM latency.
// Measured on C2: build+probe 6.09 -> 6.07 ms (profiles/r02_sweep_bits_fill_bpnext_c2.json, r02_sweep_p2_occupancy_bpnext_c2.json),
// C3 unchanged, and the exchange instantiations spill 24-28 bytes with it: off — the start of an item is not what the kernel waits for
#define HJ_TUNE_BP_NEXT 0
#endif
#ifndef HJ_TUNE_FIND_ILP
#define HJ_TUNE_FIND_ILP 1  // 1: first-bucket lookups of a warp's V rows in straight-line code (needs HJ_TUNE_BUCKET_FIND)
#endif
#ifndef HJ_TUNE_BUCKET_FIND
#define HJ_TUNE_BUCKET_FIND 1
#endif
#if HJ_TUNE_FIND_ILP && !HJ_TUNE_BUCKET_FIND
#error "HJ_TUNE_FIND_ILP needs HJ_TUNE_BUCKET_FIND"
#endif
template <typename K> constexpr int kBucketShift = HJ_TUNE_BUCKET_FIND ? (sizeof(K) == 4 ? 2 : 1) : 0;
template <typename K>
__device__ __forceinline__ uint32_t table_start(uint32_t hash_above_radix, uint32_t ns_mask) {
  return (hash_above_radix << kBucketShift<K>) & ns_mask;
}

// slot of `key` or NS+1 if absent (read-only; slot NS+1 of w is a zero pad). `slot` comes from table_start().
template <typename K, int NS>
__device__ __forceinline__ uint32_t table_find(const K* keys, uint32_t ns_mask, K key, uint32_t slot) {
  if (key == KeyTraits<K>::kEmpty) return NS;
#if HJ_TUNE_BUCKET_FIND
  // a key sits before the first empty slot of its probe sequence (insertions claim the first empty slot, nothing is
  // ever removed), so "no match in this bucket and an empty slot in it" means absent
  constexpr K E = KeyTraits<K>::kEmpty;
  for (;;) {
    if constexpr (sizeof(K) == 4) {
      const int4 q = *reinterpret_cast<const int4*>(keys + slot);
      const bool m0 = q.x == key, m1 = q.y == key, m2 = q.z == key, m3 = q.w == key;
      if (m0 | m1 | m2 | m3) return slot + (m0 ? 0u : m1 ? 1u : m2 ? 2u : 3u);
      if ((q.x == E) | (q.y == E) | (q.z == E) | (q.w == E)) return NS + 1;
      slot = (slot + 4) & ns_mask;
    } else {
      const longlong2 q = *reinterpret_cast<const longlong2*>(keys + slot);
      const bool m0 = q.x == key, m1 = q.y == key;
      if (m0 | m1) return slot + (m0 ? 0u : 1u);
      if ((q.x == E) | (q.y == E)) return NS + 1;
      slot = (slot + 2) & ns_mask;
    }
  }
#else
  for (;;) {
    const K cur = keys[slot];
    if (cur == key) return slot;
    if (cur == KeyTraits<K>::kEmpty) return NS + 1;
    slot = (slot + 1) & ns_mask;
  }
#endif
}

// First bucket only: the slot of `key`, NS + 1 if the bucket proves it absent, kFindPending if the search must go on.
constexpr uint32_t kFindPending = 0xFFFFFFFFu;
template <typename K> constexpr uint32_t kBucketKeys = 16 / sizeof(K);
template <typename K, int NS>
__device__ __forceinline__ uint32_t table_find_first(const K* keys, K key, uint32_t slot) {
  constexpr K E = KeyTraits<K>::kEmpty;
  uint32_t r;
  if constexpr (sizeof(K) == 4) {
    const int4 q = *reinterpret_cast<const int4*>(keys + slot);
    const bool m0 = q.x == key, m1 = q.y == key, m2 = q.z == key, m3 = q.w == key;
    const bool e = (q.x == E) | (q.y == E) | (q.z == E) | (q.w == E);
    r = (m0 | m1 | m2 | m3) ? slot + (m0 ? 0u : m1 ? 1u : m2 ? 2u : 3u) : (e ? NS + 1u : kFindPending);
  } else {
    const longlong2 q = *reinterpret_cast<const longlong2*>(keys + slot);
    const bool m0 = q.x == key, m1 = q.y == key;
    const bool e = (q.x == E) | (q.y == E);
    r = (m0 | m1) ? slot + (m0 ? 0u : 1u) : (e ? NS + 1u : kFindPending);
  }
  return key == E ? static_cast<uint32_t>(NS) : r;
}

// How a tuple becomes a RowID when a result row is emitted:
//   kRidGeneric  position -> RowID through the relation's chunk table (make_rowid)
//   kRidFast     both relations have power-of-two chunks and identity chunk ids: shift/mask
//   kRidPayload  the tuple carries the RowID itself (idx = ChunkID, pad = ChunkOffset); int64-key tuples only
//   kRidXch      the tuple names its row as (source rank, position in that rank's shard): push exchange (xch_rowid)
constexpr int kRidGeneric = 0, kRidFast = 1, kRidPayload = 2, kRidXch = 3;
template <int RM>
__device__ __forceinline__ uint64_t rowid_of(const RelMeta& m, uint32_t idx) {
  if (RM == kRidFast) return pack_rowid(idx >> m.uniform_shift, idx & ((1u << m.uniform_shift) - 1u));
  return make_rowid(m, idx);
}
template <typename K, bool PAY>
__device__ __forceinline__ uint32_t tuple_hi(const Tuple<K, PAY>& t) {
  if constexpr (sizeof(Tuple<K, PAY>) > 8) return t.pad; else return 0u;
}
template <typename K, bool PAY>
__device__ __forceinline__ uint64_t tuple_rowid(const RelMeta& m, const Tuple<K, PAY>& t, int payload) {
  return payload ? pack_rowid(t.idx, tuple_hi(t)) : make_rowid(m, t.idx);
}
// probe-side RowID of a tuple in any of the three tuple flavours (K6, K7)
template <typename K, bool PAY>
__device__ __forceinline__ uint64_t probe_rowid(const JoinArgs<K, PAY>& a, const Tuple<K, PAY>& t) {
  if (a.payload == 2) return xch_rowid(a.xmap[1], xch_rank<K, PAY>(t, a.bits2, a.radix_bits), t.idx);
  return tuple_rowid<K, PAY>(a.probe_rel, t, a.payload);
}

// Mode classes the probe loop is specialised for (compile time: no per-row mode branches).
constexpr int kClsInner = 0, kClsOuter = 1, kClsSemi = 2, kClsAnti = 3;

// Writes the result rows of one probe tile. CONTIG: the tile's reserved range is one in-bounds contiguous piece of the
// output arrays (the common case), so a row's position is ob/op + a 32-bit index; otherwise every row goes through
// out_pos() and the capacity check (chunk boundary inside the tile, or the overflow pass that only counts).
template <typename K, bool PAY, int V, int CLS, int RM, bool CONTIG>
__device__ __forceinline__ void emit_tile(const JoinArgs<K, PAY>& a, const uint32_t* s_grp, const uint8_t* s_rk, const XchMap* s_xmap,
                                          const Tuple<K, PAY>* bchunk,
                                          const uint32_t (&pw)[V], const uint32_t (&pidx)[V], const uint32_t (&ppad)[V],
                                          const uint32_t (&outc)[V], const uint32_t (&ex)[V], uint32_t wbase,
                                          unsigned long long base0, unsigned long long base1, uint32_t first) {
  constexpr bool one_side = CLS == kClsSemi || CLS == kClsAnti;
  constexpr uint32_t LIGHT = 4;  // matches per probe row expanded by the owning lane; more: by the whole warp
  uint64_t* ob = one_side ? nullptr : a.out_build + (CONTIG ? base0 : 0ull);
  uint64_t* op = a.out_probe + (CONTIG ? base0 : 0ull);
  auto put = [&](unsigned long long e, uint64_t brid, uint64_t prid) {
    if (CONTIG) {
#if HJ_TUNE_OUT_CS
      if (!one_side) __stcs(reinterpret_cast<unsigned long long*>(ob + e), brid);
      __stcs(reinterpret_cast<unsigned long long*>(op + e), prid);
#else
      if (!one_side) ob[e] = brid;
      op[e] = prid;
#endif
    } else {
      const unsigned long long pos = out_pos(base0, base1, first, e);
      if (pos < a.capacity) {
        if (!one_side) ob[pos] = brid;
        op[pos] = prid;
      }
    }
  };
  // build-side RowID of grouped entry g: s_grp holds relation positions, or (payload mode) positions inside this build
  // chunk, whose tuples (<= CAP x 16 B, read by every probe row of the partition) stay cache-resident
  auto brid = [&](uint32_t g) -> uint64_t {
    if constexpr (RM == kRidPayload) {
      const Tuple<K, PAY> t = bchunk[s_grp[g]];
      return pack_rowid(t.idx, tuple_hi(t));
    } else if constexpr (RM == kRidXch) {
      return xch_rowid(s_xmap[0], s_rk[g], s_grp[g]);  // (source rank, position in that rank's shard), both grouped in smem
    } else {
      return rowid_of<RM>(a.build_rel, s_grp[g]);
    }
  };
  // pad: the probe tuple's ChunkOffset (kRidPayload) / source rank (kRidXch)
  auto prid_of = [&](uint32_t idx, uint32_t pad) -> uint64_t {
    if constexpr (RM == kRidPayload) return pack_rowid(idx, pad);
    else if constexpr (RM == kRidXch) return xch_rowid(s_xmap[1], pad, idx);
    else return rowid_of<RM>(a.probe_rel, idx);
  };
  const uint32_t lane = lane_id();
  uint32_t heavy = 0;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    if (outc[j] == 0) continue;
    const uint32_t e = wbase + ex[j];
    const uint64_t prid = prid_of(pidx[j], ppad[j]);
    const uint32_t cnt = pw[j] & 0xFFFFu;
    if (one_side || cnt == 0) {
      put(e, kNullRowId, prid);  // Semi/Anti row, or an unmatched outer row (join_hash.cpp:432-435)
    } else if (cnt == 1) {
      put(e, brid((pw[j] >> 16) - 1u), prid);
    } else if (cnt <= LIGHT) {
      const uint32_t start = (pw[j] >> 16) - cnt;
      for (uint32_t m = 0; m < cnt; ++m) put(e + m, brid(start + m), prid);
    } else {
      heavy |= 1u << j;
    }
  }
  if (one_side) return;
  if (__any_sync(0xffffffffu, heavy != 0)) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      uint32_t hmask = __ballot_sync(0xffffffffu, (heavy >> j) & 1u);
      while (hmask) {
        const int src = __ffs(hmask) - 1;
        hmask &= hmask - 1;
        const uint32_t w = __shfl_sync(0xffffffffu, pw[j], src);
        const uint32_t e = wbase + __shfl_sync(0xffffffffu, ex[j], src);
        const uint32_t pi = __shfl_sync(0xffffffffu, pidx[j], src);
        const uint32_t pp = __shfl_sync(0xffffffffu, ppad[j], src);
        const uint32_t cnt = w & 0xFFFFu, start = (w >> 16) - cnt;
        const uint64_t prid = prid_of(pi, pp);
        for (uint32_t m = lane; m < cnt; m += 32) put(static_cast<unsigned long long>(e) + m, brid(start + m), prid);
      }
    }
  }
}

template <typename K, bool PAY, int NT, int NS, int CAP, int V, int RM, int CLS>
__global__ void __launch_bounds__(NT, 1) k_build_probe(const JoinArgs<K, PAY> a) {
  static_assert(CAP < NS && CAP <= 65535, "table must keep empty slots; run ends are 16 bit");
  constexpr uint32_t NW = NT / 32;
  constexpr uint32_t MIN_NS = 32 * NW;  // every warp scans >= 32 slots
  constexpr int BU = 2 * (sizeof(K) == 4 ? 2 : 1) + 0;  // build tuples per thread in flight at a time
  constexpr int BR = ((CAP + NT - 1) / NT + BU - 1) / BU * BU;  // build tuples per thread, padded to whole rounds
  static_assert(BU % 2 == 0 && NS + 1 < 65536, "slots are kept as 16-bit halves");
  extern __shared__ __align__(16) unsigned char s_raw[];
  K* s_keys = reinterpret_cast<K*>(s_raw);                          // NS + 1
  uint32_t* s_w = reinterpret_cast<uint32_t*>(s_keys + (NS + 1));   // NS + 2 (slot NS: sentinel key, NS+1: always 0)
  uint32_t* s_grp = s_w + (NS + 2);                                 // CAP
  uint8_t* s_rk = reinterpret_cast<uint8_t*>(s_grp + CAP);          // CAP, kRidXch only: source rank of every grouped build row
  __shared__ uint32_t s_wtot[NW + 1];
  __shared__ uint32_t s_item;
#if HJ_TUNE_BP_NEXT
  __shared__ WorkItem s_next_item;
  uint32_t next_idx = 0;  // thread 0: the popped index on its way back from the atomic
  if (threadIdx.x == 0) {
    s_item = atomicAdd(a.queue_head, 1u);
    if (s_item < a.plan->n_items) s_next_item = a.items[s_item];
  }
#endif
  // kRidXch: the shard geometry by rank, indexed per lane (a divergent index into the parameter bank would be serialised)
  __shared__ XchMap s_xmap[RM == kRidXch ? 2 : 1];
  if constexpr (RM == kRidXch) {
    if (threadIdx.x < 32) {
      const int side = threadIdx.x >> 4, r = threadIdx.x & 15;
      s_xmap[side].first_chunk[r] = a.xmap[side].first_chunk[r];
      s_xmap[side].shift[r] = a.xmap[side].shift[r];
    }
  }

  const uint32_t lane = lane_id(), warp = warp_id();
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t n_items = a.plan->n_items;
  const int radix_bits = a.radix_bits;
  const int chunk_shift = a.chunk_shift;
  // push-exchanged int32 tuples: the key field is the key's hash (a bijection of the key) with the source rank written over
  // the level-1 digit bits; equal keys = equal fields once those bits are forced to one (the partition fixes them anyway)
  constexpr bool PRE = RM == kRidXch && sizeof(K) == 4;
  const uint32_t rank_bits = PRE ? ((1u << (radix_bits - a.bits2)) - 1u) << a.bits2 : 0u;
  auto canon = [&](K k) -> K {
    if constexpr (PRE) return static_cast<K>(static_cast<uint32_t>(k) | rank_bits); else return k;
  };
  OutWriter wr;  // every warp is its own result writer (state replicated in the lanes' registers, lane 0 allocates)
  out_writer_init(&wr);
  if (threadIdx.x == 0) s_w[NS + 1] = 0;

  for (;;) {
#if HJ_TUNE_BP_NEXT
    __syncthreads();
    const uint32_t item_idx = s_item;
    if (item_idx >= n_items) break;
    const WorkItem it = s_next_item;
#else
    if (threadIdx.x == 0) s_item = atomicAdd(a.queue_head, 1u);
    __syncthreads();
    const uint32_t item_idx = s_item;
    if (item_idx >= n_items) break;
    const WorkItem it = a.items[item_idx];
#endif
    const bool multi = it.flags & kItemMulti;

    // table size for this chunk: >= 2x the tuples (load factor <= 0.5) up to NS (then <= CAP/NS)
    uint32_t ns = MIN_NS;
    while (ns < 2 * it.b_cnt && ns < NS) ns <<= 1;
    const uint32_t ns_mask = ns - 1;
    const Tuple<K, PAY>* probe = a.probe + it.p_begin;
    const Tuple<K, PAY>* build = a.build + it.b_begin;

    // the warp's first probe rows go in flight before the table is built
    Tuple<K, PAY> nxt[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const uint32_t i = warp * (32 * V) + j * 32 + lane;
      if (i < it.p_cnt) nxt[j] = ld_stream(probe + i);
    }

    for (uint32_t s = threadIdx.x; s < ns; s += NT) {
      s_keys[s] = KeyTraits<K>::kEmpty;
      s_w[s] = 0;
    }
    if (threadIdx.x == 0) s_w[NS] = 0;
    __syncthreads();

    // ---- build pass A: distinct keys + counts, BU tuples per thread in flight at a time (v1 loaded one tuple per loop
    // iteration and paid the DRAM latency CAP/NT times); the slot of every tuple is kept (16 bit) for pass B
    uint32_t bslot[BR / 2];
#pragma unroll
    for (int r0 = 0; r0 < BR; r0 += BU) {
      if (r0 * NT >= it.b_cnt) break;  // CTA-uniform
      K bk[BU];
#pragma unroll
      for (int u = 0; u < BU; ++u) {
        const uint32_t i = (r0 + u) * NT + threadIdx.x;
        bk[u] = i < it.b_cnt ? canon(build[i].key) : K(0);
      }
#pragma unroll
      for (int u = 0; u < BU; ++u) {
        const uint32_t i = (r0 + u) * NT + threadIdx.x;
        uint32_t slot = NS + 1;
        if (i < it.b_cnt) {
          slot = table_insert<K, NS>(s_keys, ns_mask, bk[u], table_start<K>(tuple_hash<PRE>(bk[u]) >> radix_bits, ns_mask));
          atomicAdd(&s_w[slot], 1u);
        }
        if ((u & 1) == 0) bslot[(r0 + u) / 2] = slot; else bslot[(r0 + u) / 2] |= slot << 16;
      }
    }
    __syncthreads();
#if HJ_TUNE_BP_NEXT
    if (threadIdx.x == 0) next_idx = atomicAdd(a.queue_head, 1u);  // comes back while the counts are scanned
#endif

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

    // ---- build pass B: group the row positions by key (the tuples come from L2 this time)
#pragma unroll
    for (int r0 = 0; r0 < BR; r0 += BU) {
      if (r0 * NT >= it.b_cnt) break;  // CTA-uniform
      uint32_t bi[BU], brk[BU];
#pragma unroll
      for (int u = 0; u < BU; ++u) {
        const uint32_t i = (r0 + u) * NT + threadIdx.x;
        brk[u] = 0;
        if constexpr (RM == kRidXch) {
          if (i < it.b_cnt) {
            const Tuple<K, PAY> t = ld_stream(build + i);
            bi[u] = t.idx;
            brk[u] = xch_rank<K, PAY>(t, a.bits2, radix_bits);
          }
        } else {
          bi[u] = RM == kRidPayload ? i : (i < it.b_cnt ? build[i].idx : 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < BU; ++u) {
        const uint32_t i = (r0 + u) * NT + threadIdx.x;
        if (i < it.b_cnt) {
          const uint32_t slot = (bslot[(r0 + u) / 2] >> ((u & 1) * 16)) & 0xFFFFu;
          const uint32_t old = atomicAdd(&s_w[slot], 0x10000u);
          s_grp[old >> 16] = bi[u];
          if constexpr (RM == kRidXch) s_rk[old >> 16] = static_cast<uint8_t>(brk[u]);
        }
      }
    }
    __syncthreads();  // s_w[slot] is now (end of run << 16) | count
#if HJ_TUNE_BP_NEXT
    // every thread has read s_item / s_next_item (barriers above); the loop's top barrier publishes the new values
    if (threadIdx.x == 0) {
      s_item = next_idx;
      if (next_idx < n_items) {
        const WorkItem ni = a.items[next_idx];
        s_next_item = ni;
        bulk_prefetch_l2(a.build + ni.b_begin, ni.b_cnt * static_cast<uint32_t>(sizeof(Tuple<K, PAY>)));
        bulk_prefetch_l2(a.probe + ni.p_begin, min(ni.p_cnt, static_cast<uint32_t>(NT * V)) * static_cast<uint32_t>(sizeof(Tuple<K, PAY>)));
      }
    }
#endif

    // ---- probe: every warp walks its own rows (warp w: rows w*32V .. +32V of every NT*V-row stripe) without CTA barriers
    for (uint32_t base = warp * (32 * V); base < it.p_cnt; base += NT * V) {
      uint32_t pw[V], pidx[V], ppad[V], outc[V], ex[V];
      const uint32_t row0 = base + lane;
      bool any_multi = false;
#if HJ_TUNE_FIND_ILP
      // first buckets of all V rows in straight-line code (the loads are independent and can be in flight together);
      // only a row whose first bucket is full of other keys falls back to the loop
      uint32_t fstart[V], fslot[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        fstart[j] = table_start<K>(tuple_hash<PRE>(nxt[j].key) >> radix_bits, ns_mask);
        fslot[j] = table_find_first<K, NS>(s_keys, canon(nxt[j].key), fstart[j]);
      }
#endif
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const bool act = row0 + j * 32 < it.p_cnt;
        pw[j] = 0;
        pidx[j] = nxt[j].idx;
        ppad[j] = RM == kRidPayload ? tuple_hi(nxt[j]) : RM == kRidXch ? xch_rank<K, PAY>(nxt[j], a.bits2, radix_bits) : 0u;
#if HJ_TUNE_FIND_ILP
        if (act) pw[j] = s_w[fslot[j] != kFindPending ? fslot[j]
                                 : table_find<K, NS>(s_keys, ns_mask, canon(nxt[j].key), (fstart[j] + kBucketKeys<K>) & ns_mask)];
#else
        if (act) {
          const K key = canon(nxt[j].key);
          const uint32_t slot = table_find<K, NS>(s_keys, ns_mask, key, table_start<K>(tuple_hash<PRE>(nxt[j].key) >> radix_bits, ns_mask));
          pw[j] = s_w[slot];
        }
#endif
        const uint32_t cnt = pw[j] & 0xFFFFu;
        uint32_t o;
        if (CLS == kClsInner) o = cnt;
        else if (CLS == kClsOuter) o = multi ? cnt : (act ? max(cnt, 1u) : 0u);
        else if (CLS == kClsSemi) o = (!multi && cnt) ? 1u : 0u;
        else o = (!multi && act && cnt == 0) ? 1u : 0u;
        if (CLS != kClsInner && multi && cnt) {
          const uint32_t pos = it.p_begin + row0 + j * 32;
          atomicOr(a.matched + (pos >> 5), 1u << (pos & 31u));
        }
        outc[j] = o;
        any_multi |= o > 1;
      }
      // the warp's next rows are in flight before this batch's results are placed and written
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const uint32_t i = row0 + NT * V + j * 32;
        if (i < it.p_cnt) nxt[j] = ld_stream(probe + i);
      }
      uint32_t running = 0;
      if ((CLS == kClsSemi || CLS == kClsAnti) || !__any_sync(0xffffffffu, any_multi)) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const uint32_t bal = __ballot_sync(0xffffffffu, outc[j] != 0);
          ex[j] = running + __popc(bal & lt_mask);
          running += __popc(bal);
        }
      } else {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const uint32_t incl = warp_incl_scan(outc[j]);
          ex[j] = running + incl - outc[j];
          running += __shfl_sync(0xffffffffu, incl, 31);
        }
      }
      if (running == 0) continue;  // warp-uniform
      warp_writer_reserve(&wr, a.out_state, running, chunk_shift);
      if (wr.first == 0xFFFFFFFFu && wr.base0 + running <= a.capacity)
        emit_tile<K, PAY, V, CLS, RM, true>(a, s_grp, s_rk, s_xmap, build, pw, pidx, ppad, outc, ex, 0u, wr.base0, wr.base1, wr.first);
      else
        emit_tile<K, PAY, V, CLS, RM, false>(a, s_grp, s_rk, s_xmap, build, pw, pidx, ppad, outc, ex, 0u, wr.base0, wr.base1, wr.first);
    }
  }
  if (lane == 0) out_writer_finish(&wr, a.out_state, a.out_tails);
}

// =====================================================================================================================
// Block-wide compaction through the chunked writer: every thread with `flag` emits one result row.
// =====================================================================================================================
template <int NT>
__device__ __forceinline__ void emit_flagged(bool flag, uint64_t build_rid, uint64_t probe_rid, uint64_t* out_build,
                                             uint64_t* out_probe, unsigned long long capacity, OutWriter* w,
                                             OutState* g, uint32_t* scratch, int chunk_shift) {
  uint32_t total;
  const uint32_t ex = block_excl_scan(flag ? 1u : 0u, scratch, &total);
  if (threadIdx.x == 0 && total) out_writer_reserve(w, g, total, chunk_shift);
  __syncthreads();
  if (flag) {
    const unsigned long long pos = out_pos(w->base0, w->base1, w->first, ex);
    if (pos < capacity) {
      if (out_build) out_build[pos] = build_rid;
      out_probe[pos] = probe_rid;
    }
  }
  __syncthreads();
}

// K6: unmatched/matched decisions for multi partitions, from the bitmap K5 filled over all chunks.
template <typename K, bool PAY, int NT>
__global__ void __launch_bounds__(NT) k_finalize_multi(const JoinArgs<K, PAY> a, const WorkItem* __restrict__ fin) {
  __shared__ uint32_t scratch[33];
  __shared__ OutWriter s_out;
  const uint32_t n_fin = a.plan->n_fin;
  const int mode = a.mode;
  if (threadIdx.x == 0) out_writer_init(&s_out);
  __syncthreads();
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
        if (flag) prid = probe_rowid<K, PAY>(a, a.probe[pos]);
      }
      emit_flagged<NT>(flag, kNullRowId, prid, a.out_build, a.out_probe, a.capacity, &s_out, a.out_state, scratch, a.chunk_shift);
    }
  }
  if (threadIdx.x == 0) out_writer_finish(&s_out, a.out_state, a.out_tails);
}

// K7: HJ_FLAG_SQL_NULL_PROBE — NULL probe keys of a Left/Right join were routed aside by pass 1; each yields (NULL, p).
template <typename K, bool PAY, int NT>
__global__ void __launch_bounds__(NT) k_emit_null_probe(const JoinArgs<K, PAY> a, const Tuple<K, PAY>* __restrict__ pass1,
                                                        const uint32_t* __restrict__ off_p) {
  __shared__ uint32_t scratch[33];
  __shared__ OutWriter s_out;
  const uint32_t P = 1u << a.radix_bits;
  const uint32_t lo = off_p[P], hi = off_p[P + 1];
  if (threadIdx.x == 0) out_writer_init(&s_out);
  __syncthreads();
  for (uint32_t base = lo + blockIdx.x * NT; base < hi; base += gridDim.x * NT) {
    const uint32_t i = base + threadIdx.x;
    const bool flag = i < hi;
    const uint64_t prid = flag ? probe_rowid<K, PAY>(a, pass1[i]) : 0;
    emit_flagged<NT>(flag, kNullRowId, prid, a.out_build, a.out_probe, a.capacity, &s_out, a.out_state, scratch, a.chunk_shift);
  }
  if (threadIdx.x == 0) out_writer_finish(&s_out, a.out_state, a.out_tails);
}

// =====================================================================================================================
// K8: closes the holes at the writers' tails. With T = rows emitted, C = chunk rows and A = chunks handed out * C, the rows
// stored at positions >= T are moved into the holes below T (both sets have the same size), leaving the dense range
// [0, T). Every chunk is dense except the last chunk of each writer, and a chunk has at most one writer, so
// k_compact_plan (one CTA) first fills used[c] = C, lets every tail overwrite its chunk's entry, and then scans the
// per-chunk hole lengths below T and the per-chunk filled lengths above T; k_compact_move maps move k to its
// (source, destination) pair with two binary searches over those prefix arrays. No sorting of the tails.
// =====================================================================================================================
struct CompactHdr {
  uint32_t n_move;    // rows to move
  uint32_t n_chunks;  // entries of used/hole_pre/src_pre (+1 for the prefixes)
};

__global__ void __launch_bounds__(1024) k_compact_plan(const OutState* __restrict__ g, const OutTail* __restrict__ tails,
                                                       int chunk_shift, unsigned long long capacity,
                                                       uint32_t* __restrict__ used, uint32_t* __restrict__ hole_pre,
                                                       uint32_t* __restrict__ src_pre, CompactHdr* __restrict__ hdr) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry_h, s_carry_s;
  const unsigned long long T = g->n_rows, nc64 = g->n_chunks;
  const uint32_t C = 1u << chunk_shift;
  if ((nc64 << chunk_shift) > capacity) {  // the buffers overflowed: the host probes again, nothing to compact
    if (threadIdx.x == 0) { hdr->n_move = 0; hdr->n_chunks = 0; }
    return;
  }
  const uint32_t nc = static_cast<uint32_t>(nc64);
  for (uint32_t c = threadIdx.x; c < nc; c += blockDim.x) used[c] = C;
  if (threadIdx.x == 0) { s_carry_h = 0; s_carry_s = 0; }
  __syncthreads();
  const uint32_t nt = min(g->n_tails, kMaxTails);
  for (uint32_t t = threadIdx.x; t < nt; t += blockDim.x) {
    const unsigned long long pos = tails[t].pos;
    const unsigned long long c = pos >> chunk_shift;
    if (c < nc) used[c] = static_cast<uint32_t>(pos & (C - 1u));
  }
  __syncthreads();
  for (uint32_t base = 0; base < nc; base += blockDim.x) {
    const uint32_t c = base + threadIdx.x;
    uint32_t hole = 0, src = 0;
    if (c < nc) {
      const unsigned long long lo = static_cast<unsigned long long>(c) << chunk_shift, fe = lo + used[c], ce = lo + C;
      const unsigned long long he = ce < T ? ce : T;       // holes count below T only: [fe, min(ce, T))
      if (fe < he) hole = static_cast<uint32_t>(he - fe);
      const unsigned long long sb = lo > T ? lo : T;        // rows to move lie at or above T: [max(lo, T), fe)
      if (fe > sb) src = static_cast<uint32_t>(fe - sb);
    }
    uint32_t th, ts;
    const uint32_t eh = block_excl_scan(hole, scratch, &th) + s_carry_h;
    const uint32_t es = block_excl_scan(src, scratch, &ts) + s_carry_s;
    if (c < nc) {
      hole_pre[c] = eh;
      src_pre[c] = es;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      s_carry_h += th;
      s_carry_s += ts;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    hole_pre[nc] = s_carry_h;
    src_pre[nc] = s_carry_s;
    hdr->n_move = min(s_carry_h, s_carry_s);  // equal by construction
    hdr->n_chunks = nc;
  }
}

// largest c in [0, n) with pre[c] <= k (entries of zero length share their successor's prefix, so the result has pre[c+1] > k)
__device__ __forceinline__ uint32_t seg_of(const uint32_t* pre, uint32_t n, uint32_t k) {
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (pre[mid] <= k) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(256) k_compact_move(const CompactHdr* __restrict__ hdr, const OutState* __restrict__ g,
                                                      int chunk_shift, const uint32_t* __restrict__ used,
                                                      const uint32_t* __restrict__ hole_pre,
                                                      const uint32_t* __restrict__ src_pre, uint64_t* __restrict__ out_build,
                                                      uint64_t* __restrict__ out_probe) {
  const uint32_t n = hdr->n_move, nc = hdr->n_chunks;
  const unsigned long long T = g->n_rows;
  for (uint32_t k = blockIdx.x * 256u + threadIdx.x; k < n; k += gridDim.x * 256u) {
    const uint32_t cd = seg_of(hole_pre, nc, k), cs = seg_of(src_pre, nc, k);
    const unsigned long long dst = (static_cast<unsigned long long>(cd) << chunk_shift) + used[cd] + (k - hole_pre[cd]);
    const unsigned long long slo = static_cast<unsigned long long>(cs) << chunk_shift;
    const unsigned long long src = (slo > T ? slo : T) + (k - src_pre[cs]);
    if (out_build) out_build[dst] = out_build[src];
    out_probe[dst] = out_probe[src];
  }
}

// =====================================================================================================================
// K10: ReferenceColumn inputs (reference_column_iterable.hpp:59-90, what materialize_input sees through
// a ReferenceColumn): key = referenced_table.chunk[pos.chunk_id].column[pos.chunk_offset], NULL for NULL_ROW_ID and for
// NULL values; NULL rows carry T{} like the reference's iterator. One job per REFERENCE segment; rows of all jobs are
// numbered consecutively (row_begin). Out-of-range RowIDs are reported through *error and treated as NULL.
// =====================================================================================================================
// Point access into one chunk of a data table's key column: VALUE, DICT, FOR (frame of reference, 2048-row blocks) or
// RLE (binary search over the inclusive run ends; null_value_id holds the number of runs).
constexpr uint8_t kSegFor = 4, kSegRle = 5;  // == HJ_SEG_FOR / HJ_SEG_RLE
// SimdBp128 (simd_bp128_packing.cpp): data is a sequence of 16-byte words; meta block m (2048 values) starts at word
// meta_off[m] with one word holding the bit widths of its 16 blocks, block b then takes width[b] words; value j of a
// block lives in 32-bit lane j % 4 at bit (j / 4) * width of that lane's stream, possibly split over two words.
__device__ __forceinline__ uint32_t bp128_get(const uint32_t* data, const uint32_t* meta_off, uint32_t i) {
  const uint32_t m = i >> 11, b = (i >> 7) & 15u, j = i & 127u;
  const uint32_t base = __ldg(meta_off + m);
  const uint8_t* widths = reinterpret_cast<const uint8_t*>(data + static_cast<size_t>(base) * 4);
  uint32_t start = base + 1u;
  for (uint32_t q = 0; q < b; ++q) start += widths[q];
  const uint32_t width = widths[b];
  if (width == 0) return 0u;
  const uint32_t lane = j & 3u, bit = (j >> 2) * width, w = bit >> 5, o = bit & 31u;
  const uint32_t* p = data + (static_cast<size_t>(start) + w) * 4 + lane;
  uint32_t v = __ldg(p) >> o;
  if (o + width > 32u) v |= __ldg(p + 4) << (32u - o);
  return width == 32u ? v : v & ((1u << width) - 1u);
}

// start word of every meta block (one thread walks the stream: n_rows / 2048 steps)
__global__ void k_bp128_meta_offsets(const uint32_t* __restrict__ data, uint32_t n_rows, uint32_t* __restrict__ meta_off) {
  if (blockIdx.x || threadIdx.x) return;
  const uint32_t M = (n_rows + 2047u) >> 11;
  uint32_t pos = 0;
  for (uint32_t m = 0; m < M; ++m) {
    meta_off[m] = pos;
    const uint8_t* widths = reinterpret_cast<const uint8_t*>(data + static_cast<size_t>(pos) * 4);
    const uint32_t blocks = m + 1 < M ? 16u : (n_rows - (m << 11) + 127u) >> 7;
    uint32_t words = 1;
    for (uint32_t q = 0; q < blocks; ++q) words += widths[q];
    pos += words;
  }
}

// attribute-vector / offset entry i of a DICT or FOR chunk, FixedSizeByteAligned or SimdBp128 (attr_width 0; the meta
// block offsets hang on row_offsets, which these kinds do not use otherwise)
__device__ __forceinline__ uint32_t load_attr_any(const DevSegment& s, uint32_t i) {
  if (s.attr_width == 0) return bp128_get(static_cast<const uint32_t*>(s.attr), s.row_offsets, i);
  return load_attr(s, i);
}

// element i of a chunk's value array as a 64-bit container: integers sign-extended, float/double as their bit patterns
__device__ __forceinline__ int64_t load_raw(const DevSegment& s, uint32_t i) {
  if (s.key_type == 0) return __ldg(static_cast<const int32_t*>(s.values) + i);
  if (s.key_type == 2) return static_cast<int64_t>(__ldg(static_cast<const uint32_t*>(s.values) + i));
  return __ldg(static_cast<const int64_t*>(s.values) + i);
}

// key of source type `src` (hj_key_type) as the join's hashed type `hash_type`, returned as the integer the pipeline
// partitions and compares: the value itself for integer joins, the IEEE bit pattern for float/double joins. NaN equals
// nothing in the reference (operator== on the hashed type), so build-side and probe-side NaNs get two different
// canonical patterns.
__device__ __forceinline__ int64_t to_hashed_type(int64_t raw, uint32_t src, uint32_t hash_type, bool build_side) {
  if (hash_type < 2) return raw;
  if (hash_type == 2) {
    float f = src == 2 ? __uint_as_float(static_cast<uint32_t>(raw)) : src == 0 ? __int2float_rn(static_cast<int>(raw)) : __ll2float_rn(raw);
    uint32_t bits = __float_as_uint(f);
    if (f != f) bits = build_side ? 0x7FC00001u : 0x7FC00000u;
    return static_cast<int64_t>(static_cast<int32_t>(bits));
  }
  double d = src == 3 ? __longlong_as_double(raw)
           : src == 2 ? static_cast<double>(__uint_as_float(static_cast<uint32_t>(raw)))
           : src == 0 ? __int2double_rn(static_cast<int>(raw)) : __ll2double_rn(raw);
  long long bits = __double_as_longlong(d);
  if (d != d) bits = build_side ? 0x7FF8000000000001ll : 0x7FF8000000000000ll;
  return bits;
}

__device__ __forceinline__ void decode_point(const DevSegment& s, uint32_t off, int64_t* val, bool* is_null) {
  *val = 0;
  *is_null = false;
  if (s.kind == 1) {
    const uint32_t vid = load_attr_any(s, off);
    if (vid == s.null_value_id) *is_null = true; else *val = load_raw(s, vid);
  } else if (s.kind == kSegFor) {
    *is_null = s.nulls && __ldg(s.nulls + off);
    *val = load_raw(s, off >> 11) + static_cast<int64_t>(load_attr_any(s, off));
  } else if (s.kind == kSegRle) {
    const uint32_t* ends = static_cast<const uint32_t*>(s.attr);
    uint32_t lo = 0, hi = s.null_value_id;  // first run with ends[run] >= off
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (__ldg(ends + mid) < off) lo = mid + 1; else hi = mid;
    }
    *is_null = s.nulls && __ldg(s.nulls + lo);
    *val = load_raw(s, lo);
  } else {
    *is_null = s.nulls && __ldg(s.nulls + off);
    *val = load_raw(s, off);
  }
}

struct RefJob {
  const uint64_t* pos;     // the segment's PosList (packed RowIDs); nullptr: decode chunk ref[0] row by row (FOR/RLE)
  const DevSegment* ref;   // chunks of the referenced key column (VALUE or DICT)
  void* out_values;        // int32[n_rows] or int64[n_rows] (key_type)
  uint8_t* out_nulls;      // 1 byte per row
  uint32_t n_ref;
  uint32_t n_rows;
  uint32_t row_begin;
  uint32_t key_type;       // of out_values: 0 = 4-byte keys, 1 = 8-byte keys
  uint32_t hash_type;      // hj_key_type the join hashes in (to_hashed_type)
  uint32_t build_side;
};

__global__ void __launch_bounds__(256) k_gather_refs(const RefJob* __restrict__ jobs, uint32_t n_jobs, uint32_t total_rows,
                                                     uint32_t* __restrict__ error) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total_rows; i += gridDim.x * blockDim.x) {
    uint32_t lo = 0, hi = n_jobs;  // last job with row_begin <= i
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (jobs[mid].row_begin <= i) lo = mid; else hi = mid;
    }
    const RefJob& jb = jobs[lo];
    const uint32_t r = i - jb.row_begin;
    const uint64_t rid = jb.pos ? __ldg(jb.pos + r) : static_cast<uint64_t>(r) << 32;
    const uint32_t chunk = static_cast<uint32_t>(rid), off = static_cast<uint32_t>(rid >> 32);
    int64_t val = 0;
    bool is_null = off == kInvalid;
    if (!is_null) {
      if (chunk >= jb.n_ref || off >= jb.ref[chunk].n_rows) {
        *error = 1;
        is_null = true;
      } else {
        decode_point(jb.ref[chunk], off, &val, &is_null);
        // through a ReferenceColumn a NULL yields T{} (reference_column_iterable.hpp:60); a decoded FOR/RLE chunk keeps
        // the stored value like ValueColumnIterable does (it matters for quirk Q4 only)
        if (is_null && jb.pos) val = 0;
        else val = to_hashed_type(val, jb.ref[chunk].key_type, jb.hash_type, jb.build_side != 0);
      }
    }
    if (jb.key_type == 0) static_cast<int32_t*>(jb.out_values)[r] = static_cast<int32_t>(val);
    else static_cast<int64_t*>(jb.out_values)[r] = val;
    jb.out_nulls[r] = is_null ? 1 : 0;
  }
}

// K9': the same for a result that stays on the device, through the PosLists where they lie: seg_pos[id] = PosList of the
// REFERENCE segment with chunk id `id` (nullptr: a data chunk, its RowIDs are final), seg_rows[id] = its length
// (0xFFFFFFFF: no segment has this chunk id -> error).
__global__ void k_translate_segments(uint64_t* __restrict__ inout, uint64_t n, const uint64_t* const* __restrict__ seg_pos,
                                     const uint32_t* __restrict__ seg_rows, uint32_t n_ids, uint32_t* __restrict__ error) {
  for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    const uint64_t r = inout[i];
    const uint32_t chunk = static_cast<uint32_t>(r), off = static_cast<uint32_t>(r >> 32);
    if (off == kInvalid) continue;  // NULL_ROW_ID stays (join_hash.cpp:586-588)
    const uint32_t rows = chunk < n_ids ? seg_rows[chunk] : 0xFFFFFFFFu;
    if (rows == 0xFFFFFFFFu || off >= rows) {
      *error = 1;
      continue;
    }
    const uint64_t* pl = seg_pos[chunk];
    if (pl) inout[i] = pl[off];
  }
}

// K9: inout[i] = pos_lists[chunk_base[inout[i].chunk_id] + inout[i].chunk_offset]; NULL_ROW_ID stays (join_hash.cpp:584-592).
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
