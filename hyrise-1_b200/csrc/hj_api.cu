// hj_api.cu — host side of libhj: the C ABI of include/hj.h, device-memory management and kernel orchestration.
// No CPU fallback anywhere: every compute entry point launches the sm_100a kernels of hj_kernels.cuh.
#include <pthread.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <list>
#include <string>
#include <vector>

#include "hj.h"
#include "hj_kernels.cuh"
#include "hj_fused.cuh"
#include "hj_scan.cuh"
#include "hj_aggregate.cuh"

using namespace hj;

namespace {

// ---- tuning constants -----------------------------------------------------------------------------------
// The HJ_TUNE_* macros exist for tests/sweep_variants.py, which builds the library several times with -D overrides and
// times the variants side by side on the GPU; the defaults below are what ships.
#ifndef HJ_TUNE_ITEMS1_32
#define HJ_TUNE_ITEMS1_32 16
#endif
#ifndef HJ_TUNE_P1_TILE32
#define HJ_TUNE_P1_TILE32 2048
#endif
#ifndef HJ_TUNE_P1_VIRTUAL32
#define HJ_TUNE_P1_VIRTUAL32 4
#endif
#ifndef HJ_TUNE_P1_VIRTUAL64
#define HJ_TUNE_P1_VIRTUAL64 2
#endif
// Pass-2 geometry and the split of the radix bits between the passes, per key width (profiles/r02_sweep_p1_p2_geometry*_c{2,3}.json,
// r02_sweep_p2_occupancy_bpnext_c2.json; C2 = int32, C3 = int64, ms for P1 + pass 2):
//   int32, 14 bits: 7 + 7 bits, 256 threads x 16, fill 14/16 (round 1)   7.47 + 5.08
//                   6 + 8 bits                                            7.43 + 5.11   (P1: fewer digits to rank and scan)
//                   6 + 8 bits, 512 threads x 16 (2 CTAs per SM)           7.37 + 5.02   (32-row fragments fill an 8192-tuple batch)
//                   ... and fill 15/16                                     7.37 + 4.86   <- shipped
//                   ... fill 16/16 / 13/16: pass 2 6.2 / 5.2; 1024 threads: 8.4; 512 x 8 or x 12 tuples, 3-4 CTAs per SM: 5.15-6.3
//                   ... with 8 pass-1 streams of 1024 rows per SM: P1 7.16 but pass 2 5.06-5.30 (16-row fragments): no gain
//   int64, 15 bits: 8 + 7, 256 x 8 (round 1) 15.2 + 18.7; 7 + 8: 15.3 + 18.9; 512 threads: 15.3 + 23.4  -> unchanged
#ifndef HJ_TUNE_P2_THREADS32
#define HJ_TUNE_P2_THREADS32 512
#endif
#ifndef HJ_TUNE_P2_THREADS64
#define HJ_TUNE_P2_THREADS64 256
#endif
#ifndef HJ_TUNE_P2_ITEMS32
#define HJ_TUNE_P2_ITEMS32 16
#endif
#ifndef HJ_TUNE_P2_ITEMS64
#define HJ_TUNE_P2_ITEMS64 8
#endif
#ifndef HJ_TUNE_P2_FILL16_32
#define HJ_TUNE_P2_FILL16_32 15  // a pass-2 work item holds FILL16/16 of a batch on average
#endif
#ifndef HJ_TUNE_P2_FILL16_64
#define HJ_TUNE_P2_FILL16_64 14
#endif
#ifndef HJ_TUNE_BITS2_DELTA32
#define HJ_TUNE_BITS2_DELTA32 1  // two-pass split: pass 2 takes radix_bits / 2 + delta bits (both passes stay <= 8 bits)
#endif
#ifndef HJ_TUNE_BITS2_DELTA64
#define HJ_TUNE_BITS2_DELTA64 0
#endif
#ifndef HJ_TUNE_PROBE_V
#define HJ_TUNE_PROBE_V 4
#endif
template <typename K> struct Cfg;
template <> struct Cfg<int32_t> {
  static constexpr int kItems1 = HJ_TUNE_ITEMS1_32;  // pass 1: 256 threads x 16 = 4096 rows per tile (32 KB staged)
  static constexpr int kItems2 = 16;      // pass 2: 4096 tuples per tile
  static constexpr int kP2Items = HJ_TUNE_P2_ITEMS32;  // fused pass 2 (k_scatter2t/f): tuples per thread and batch
  static constexpr int kP2Threads = HJ_TUNE_P2_THREADS32, kP2Fill16 = HJ_TUNE_P2_FILL16_32, kBits2Delta = HJ_TUNE_BITS2_DELTA32;
  // fused pass 1 (k_part1): 4 virtual CTAs of 256 threads with 2048-row tiles. Same shared memory as 2 x 4096 (L1 left to
  // the dictionary gather), but four independent tile streams per SM keep the gather pipeline busier: C2 P1 8.54 -> 7.88 ms,
  // pass 2 +0.12 ms for the shorter fragments (profiles/r01_sweep_variants_c2_full.json)
  static constexpr int kP1Tile = HJ_TUNE_P1_TILE32, kP1Virtual = HJ_TUNE_P1_VIRTUAL32;
  static constexpr int kNS = 16384;       // hash-table slots per shared-memory table
  static constexpr int kCap = 12288;      // build tuples per table (load factor <= 0.75)
};
template <> struct Cfg<int64_t> {
  static constexpr int kItems1 = 8;       // 2048 rows per tile (32 KB staged)
  static constexpr int kItems2 = 8;
  static constexpr int kP2Items = HJ_TUNE_P2_ITEMS64;
  static constexpr int kP2Threads = HJ_TUNE_P2_THREADS64, kP2Fill16 = HJ_TUNE_P2_FILL16_64, kBits2Delta = HJ_TUNE_BITS2_DELTA64;
  static constexpr int kP1Tile = 2048, kP1Virtual = HJ_TUNE_P1_VIRTUAL64;
  static constexpr int kNS = 8192;
  static constexpr int kCap = 6144;
};
constexpr int kScatterThreads = 256;
constexpr int kHistThreads = 1024;        // K1: one physical CTA per SM ...
constexpr int kHistVirtual = 4;           // ... hosting 4 virtual CTAs of 256 threads (one wave of all 4 x #SM tile owners)
#ifndef HJ_TUNE_JOIN_THREADS
#define HJ_TUNE_JOIN_THREADS 1024
#endif
#ifndef HJ_TUNE_MAX_CHUNK_SHIFT
#define HJ_TUNE_MAX_CHUNK_SHIFT 14
#endif
constexpr int kJoinThreads = HJ_TUNE_JOIN_THREADS;
constexpr int kProbeV = HJ_TUNE_PROBE_V;
constexpr uint32_t kProbeSlice = 128 * 1024;
constexpr int kSinglePassMaxBits = 8;
constexpr uint64_t kMinSliceRows = 32u << 20;  // streamed join: probe slices of at least 32 Mi rows
constexpr int kMinChunkShift = 8, kMaxChunkShift = HJ_TUNE_MAX_CHUNK_SHIFT;  // result chunks of 256 .. 16 Ki rows
constexpr int kVirtualCtasPerSm = 4;       // K1 / pass 1 tile ownership: G = min(n_tiles, 4 * SMs) virtual CTAs
template <typename K> constexpr int tile1() { return kScatterThreads * Cfg<K>::kItems1; }
template <typename K> constexpr int tile2() { return kScatterThreads * Cfg<K>::kItems2; }
// fused pass 2: CTA size (per key width) and CTAs per SM (1024 threads per SM at 64 registers)
#ifndef HJ_TUNE_P2_MINBLOCKS32
#define HJ_TUNE_P2_MINBLOCKS32 (1024 / HJ_TUNE_P2_THREADS32)
#endif
template <typename K> constexpr int p2_min_blocks() { return sizeof(K) == 4 ? HJ_TUNE_P2_MINBLOCKS32 : 1024 / Cfg<K>::kP2Threads; }
template <typename K> constexpr int p2_tile() { return Cfg<K>::kP2Threads * Cfg<K>::kP2Items; }
template <typename K> constexpr int p1_tile() { return Cfg<K>::kP1Tile; }
// rows per tile of the fused pass 1: column inputs p1_tile(); tuples received through the push exchange come in chunks of
// kXpChunk (hj_xpush.cuh), and a chunk is a tile
constexpr int kXpTile = 2048;
template <typename K> constexpr uint32_t fused_tile(bool xch) { return xch ? uint32_t(kXpTile) : uint32_t(p1_tile<K>()); }
// 12-byte RowID-carrying tuples of int32 keys: same tile sizes with twice the threads and half the rows per thread, so that
// a thread's keys + RowIDs (pass 1) / three-word tuples (pass 2) fit the register file without spilling
template <typename K, bool PAY> constexpr int scatter_threads() { return (PAY && sizeof(K) == 4) ? 2 * kScatterThreads : kScatterThreads; }

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

thread_local std::string g_create_error;

}  // namespace

constexpr int kSides = 6;
// pinned readback area: Readback at 0, exchange count matrices at kPinnedCounts, IPC handle records at kPinnedHandles
constexpr size_t kPinnedBytes = 96 * 1024, kPinnedRefFlags = 3584, kPinnedCounts = 4096, kPinnedHandles = 64 * 1024;

struct hj_ctx {
  int device = 0;
  int n_sms = 0;
  hj_options opts{};
  cudaStream_t own_stream = nullptr;
  std::string err;
  std::vector<DevBuf*> all_bufs;
  // per relation (0 = build, 1 = probe)
  // per relation: 0 = build, 1 = probe, 2 = second probe slice set (streamed join);
  // sliced sharded join: 0/1 = the shards being exchanged, 2 = received build rows, 3/4 = received probe slices
  DevBuf stage_in[kSides], segs[kSides], tile_seg[kSides], seg_row_begin[kSides], keys_mat[kSides], valid_mask[kSides];
  DevBuf hist[kSides], hist1[kSides], part_off[kSides], cur2[kSides], l1_tiles[kSides], tup_a[kSides], tup_b[kSides], frag[kSides];
  DevBuf items, fin, plan, queue, out_state, out_tails, compact_plan, compact_used, compact_hole, compact_src, matched, out_build[2], out_probe[2], misc;
  void* pinned = nullptr;  // small pinned readback area
  cudaEvent_t ev[16] = {};
  // streamed host join (hj_bind_output): copy streams and per-slice-set events (staged / joined / downloaded)
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev_in[2] = {}, ev_done[2] = {}, ev_out[2] = {};
  hj_row_id* bound_build = nullptr;
  hj_row_id* bound_probe = nullptr;
  uint64_t bound_capacity = 0;
  bool bound = false;            // outputs bound for the next hj_join
  bool result_in_bound = false;  // the last join delivered its result into the bound host buffers
  // result of the last join
  bool has_result = false;
  const void* res_build = nullptr;  // device PosLists of the last join (out_build[0]/out_probe[0], or acc_* after a sliced
  const void* res_probe = nullptr;  // sharded join)
  uint64_t n_result = 0;
  int n_sides = 0;
  hj_timings timings{};
  // multi-GPU
  void* nccl_comm = nullptr;
  int n_ranks = 1, rank = 0;
  // receive sets: 0 = build, 1 = probe (or probe slices 0, 2, ..), 2 = probe slices 1, 3, ..
  DevBuf xchg_counts[2], xchg_cnt_dev, xchg_send_keys[2], xchg_send_rids[2], xchg_send_nulls[2], xchg_recv_keys[3],
      xchg_recv_rids[3], xchg_recv_nulls[3];
  DevBuf ref_arena[2 + HJ_GATHER_SLOTS], ref_out[2 + HJ_GATHER_SLOTS];  // REFERENCE segments: staged referenced columns +
                                    // PosLists / gathered keys + flags ([0] build, [1] probe / scan, [2 + slot] hj_gather_column)
  uint64_t ref_h2d_bytes = 0;
  DevBuf acc_build, acc_probe;      // sliced sharded join: the slices' PosLists, appended
  cudaEvent_t ev_x[2] = {};         // sliced sharded join: slice set exchanged (recorded on copy_in, the exchange stream)
  struct PeerMap { cudaIpcMemHandle_t handle; void* ptr = nullptr; };
  PeerMap xchg_peer[16][9];        // peers' receive buffers mapped into this process (direct exchange)
  std::vector<void*> xchg_graveyard;  // exported buffers that were outgrown; freed once no peer can have them mapped
  DevBuf xchg_hdl_dev;
  bool p2p_ok = false;
  // TableScan (N4): scratch (freed by hj_release_scratch) and the result slots (owned by the scan's caller)
  DevBuf scan_bounds, scan_mask, scan_tile_count, scan_tile_base, scan_ref_ptrs, scan_totals;
  DevBuf agg_partials;  // hj_aggregate_column: one AggPartial per CTA + the folded result
  DevBuf group_cols[kMaxGroupColumns + 1];  // hj_aggregate_grouped_rows_relation: the columns as contiguous device arrays
  DevBuf group_table, group_out, group_misc;  // hj_aggregate_grouped: hash table (SoA), one GroupRecord per group, count + overflow
  uint64_t group_n = 0, group_rows = 0;
  bool group_valid = false, group_floating = false;
  DevBuf scan_pos[HJ_SCAN_SLOTS], scan_chunk_begin[HJ_SCAN_SLOTS];
  uint64_t scan_rows[HJ_SCAN_SLOTS] = {};
  uint32_t scan_chunks[HJ_SCAN_SLOTS] = {};
  bool scan_valid[HJ_SCAN_SLOTS] = {};
  const uint64_t* xchg_rowids[kSides] = {};  // set while hj_join_sharded runs its local join
  // push exchange (hj_xpush.cuh): while set for a side, prepare_relation describes that relation's segments as receive
  // regions (segment i = the rows sent by rank i)
  struct XpRegion { const uint32_t* fill = nullptr; const uint32_t* nchunks = nullptr; uint32_t cap_chunks = 0; };
  XpRegion xp_region[kSides];
  XchMap xp_map[2] = {};
  DevBuf xp_cursor;
  uint64_t xchg_bytes_sent[2] = {0, 0};

  hj_ctx() {
    for (auto* arr : {stage_in, segs, tile_seg, seg_row_begin, keys_mat, valid_mask, hist, hist1, part_off, cur2, l1_tiles, tup_a,
                      tup_b, frag})
      for (int i = 0; i < kSides; ++i) all_bufs.push_back(&arr[i]);
    for (DevBuf* b : {&items, &fin, &plan, &queue, &out_state, &out_tails, &compact_plan, &compact_used, &compact_hole, &compact_src, &matched,
                      &out_build[0], &out_build[1], &out_probe[0], &out_probe[1], &misc, &xchg_counts[0], &xchg_counts[1], &xchg_cnt_dev, &xchg_send_keys[0],
                      &xchg_send_keys[1], &xchg_send_rids[0], &xchg_send_rids[1], &xchg_send_nulls[0], &xchg_send_nulls[1],
                      &xchg_recv_keys[0], &xchg_recv_keys[1], &xchg_recv_rids[0], &xchg_recv_rids[1], &xchg_recv_nulls[0],
                      &xchg_recv_nulls[1], &xchg_recv_keys[2], &xchg_recv_rids[2], &xchg_recv_nulls[2], &acc_build, &acc_probe, &ref_arena[0], &ref_arena[1], &ref_out[0], &ref_out[1],
                      &xchg_hdl_dev, &xp_cursor, &scan_bounds, &scan_mask, &scan_tile_count, &scan_tile_base, &scan_ref_ptrs, &scan_totals, &agg_partials, &group_table, &group_out, &group_misc})
      all_bufs.push_back(b);
    for (int i = 2; i < 2 + HJ_GATHER_SLOTS; ++i) {
      all_bufs.push_back(&ref_arena[i]);
      all_bufs.push_back(&ref_out[i]);
    }
    for (auto& b : group_cols) all_bufs.push_back(&b);
  }
};

namespace {

int fail(hj_ctx* ctx, int status, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf; else g_create_error = buf;
  return status;
}

#define HJ_CUDA(ctx, call)                                                                              \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess)                                                                              \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? HJ_ERR_OOM : HJ_ERR_CUDA, "%s failed: %s (%s:%d)", #call, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                          \
  } while (0)

#define HJ_TRY(...)           \
  do {                        \
    int st_ = (__VA_ARGS__);  \
    if (st_ != HJ_OK) return st_; \
  } while (0)

#define HJ_LAUNCH_CHECK(ctx, name)                                                                 \
  do {                                                                                             \
    cudaError_t e_ = cudaGetLastError();                                                           \
    if (e_ != cudaSuccess) return fail(ctx, HJ_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e_)); \
  } while (0)

// grow-only device buffer
int ensure(hj_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return HJ_OK;
  if (b.p) HJ_CUDA(ctx, cudaFree(b.p));
  b.p = nullptr;
  b.cap = 0;
  const size_t want = (bytes + 255) & ~size_t(255);
  HJ_CUDA(ctx, cudaMalloc(&b.p, want));
  b.cap = want;
  return HJ_OK;
}

struct PreparedRel {
  RelMeta meta{};
  std::vector<DevSegment> segs;
  bool has_dict = false;
  bool has_int64 = false;
  uint32_t mask_words = 0;
  uint64_t h2d_bytes = 0;
};

// FOR / RLE chunk descriptors
const char* encoded_segment_problem(const hj_segment& s) {
  if (s.kind == HJ_SEG_FOR) {
    if (s.key_type > HJ_KEY_INT64) return "FOR encodes int32/int64 columns only (encoding_type.hpp:34)";
    if (s.attr_width != 0 && s.attr_width != 1 && s.attr_width != 2 && s.attr_width != 4)
      return "FOR: offset vector width must be 1, 2 or 4 (FixedSizeByteAligned) or 0 (SimdBp128)";
    if (s.n_rows && (!s.values || !s.attr)) return "FOR without block minima / offsets";
  } else {
    if (s.n_rows && (!s.dict_n || !s.values || !s.attr)) return "RLE without runs";
  }
  return nullptr;
}

int validate_relation(hj_ctx* ctx, const hj_relation* r, const char* name) {
  if (!r) return fail(ctx, HJ_ERR_INVALID, "%s relation is NULL", name);
  if (r->n_segs && !r->segs) return fail(ctx, HJ_ERR_INVALID, "%s relation has no segment array", name);
  if (r->location != HJ_LOC_HOST && r->location != HJ_LOC_DEVICE)
    return fail(ctx, HJ_ERR_INVALID, "%s relation: bad location %u", name, r->location);
  uint64_t n = 0;
  for (uint32_t i = 0; i < r->n_segs; ++i) {
    const hj_segment& s = r->segs[i];
    if (s.key_type > HJ_KEY_DOUBLE)
      return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: key type %u unsupported (int32/int64/float/double keys; string keys: SURVEY 8f N3)",
                  name, i, s.key_type);
    if (s.kind > HJ_SEG_RLE) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: bad segment kind %u", name, i, s.kind);
    if (s.kind == HJ_SEG_FOR || s.kind == HJ_SEG_RLE) {
      if (const char* why = encoded_segment_problem(s)) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: %s", name, i, why);
      n += s.n_rows;
      continue;
    }
    if (s.reserved0) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: reserved0 must be 0", name, i);
    if (s.kind == HJ_SEG_REFERENCE) {
      const auto* ref = static_cast<const hj_relation*>(s.values);
      if (!ref) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: REFERENCE without the referenced column", name, i);
      if (s.n_rows && !s.attr) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: REFERENCE without a PosList", name, i);
      if (ref->location != r->location)
        return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced column and PosList must live in the same place", name, i);
      if (ref->n_segs && !ref->segs) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced column has no segment array", name, i);
      for (uint32_t c = 0; c < ref->n_segs; ++c) {
        const hj_segment& q = ref->segs[c];
        if (q.kind != HJ_SEG_VALUE && q.kind != HJ_SEG_DICT && q.kind != HJ_SEG_FOR && q.kind != HJ_SEG_RLE)
          return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u must be a data column (a ReferenceColumn never "
                      "points at another reference table)", name, i, c);
        if (q.kind == HJ_SEG_FOR || q.kind == HJ_SEG_RLE) {
          if (const char* why = encoded_segment_problem(q))
            return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u: %s", name, i, c, why);
          if (q.key_type != s.key_type) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u has another key type", name, i, c);
          continue;
        }
        if (q.key_type != s.key_type) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u has another key type", name, i, c);
        if (q.kind == HJ_SEG_DICT && q.attr_width != 0 && q.attr_width != 1 && q.attr_width != 2 && q.attr_width != 4)
          return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u: attribute vector width %u", name, i, c, q.attr_width);
        if (q.n_rows && (q.kind == HJ_SEG_DICT ? !q.attr || (q.dict_n && !q.values) : !q.values))
          return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: referenced chunk %u has no data", name, i, c);
      }
      n += s.n_rows;
      continue;
    }
    if (s.kind == HJ_SEG_DICT) {
      if (s.attr_width != 0 && s.attr_width != 1 && s.attr_width != 2 && s.attr_width != 4)
        return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: attribute vector width %u (FixedSizeByteAligned 1/2/4, or 0 = SimdBp128)", name, i,
                    s.attr_width);
      if (s.n_rows && !s.attr) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: DICT without attribute vector", name, i);
      if (s.dict_n && !s.values) return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: DICT without dictionary", name, i);
    } else if (s.n_rows && !s.values) {
      return fail(ctx, HJ_ERR_INVALID, "%s chunk %u: no values", name, i);
    }
    n += s.n_rows;
  }
  if (n != r->n_rows) return fail(ctx, HJ_ERR_INVALID, "%s relation: n_rows %llu != sum of chunks %llu", name,
                                  (unsigned long long)r->n_rows, (unsigned long long)n);
  if (n >= 0xFFFFFFFFull) return fail(ctx, HJ_ERR_TOO_LARGE, "%s relation has %llu rows (limit 2^32-2 per GPU)", name,
                                      (unsigned long long)n);
  return HJ_OK;
}

// Builds the device-side description of one relation; stages host-resident arrays into ctx->stage_in[side].
int prepare_relation(hj_ctx* ctx, const hj_relation* r, int side, uint32_t tile, cudaStream_t st, PreparedRel* out) {
  PreparedRel& pr = *out;
  const uint32_t n_segs = r->n_segs;
  pr.segs.resize(n_segs);
  std::vector<uint32_t> row_begin(n_segs + 1, 0);
  uint32_t rows = 0, tiles = 0, words = 0;
  bool identity = true;
  auto key_bytes = [](const hj_segment& s) { return s.key_type == HJ_KEY_INT64 ? 8u : 4u; };

  // staging plan for host-resident inputs: (src, bytes, offset)
  struct Copy { const void* src; size_t bytes; size_t off; };
  std::vector<Copy> copies;
  size_t stage_bytes = 0;
  auto plan_copy = [&](const void* src, size_t bytes) -> size_t {
    const size_t off = stage_bytes;
    copies.push_back({src, bytes, off});
    stage_bytes += (bytes + 255) & ~size_t(255);
    return off;
  };
  const bool host = r->location == HJ_LOC_HOST;
  std::vector<size_t> off_values(n_segs, 0), off_nulls(n_segs, 0), off_attr(n_segs, 0), off_ro(n_segs, 0);

  for (uint32_t i = 0; i < n_segs; ++i) {
    const hj_segment& s = r->segs[i];
    DevSegment& d = pr.segs[i];
    std::memset(&d, 0, sizeof(d));
    d.values = s.values;
    d.nulls = s.kind == HJ_SEG_DICT ? nullptr : s.nulls;
    d.attr = s.kind == HJ_SEG_DICT ? s.attr : nullptr;
    d.row_offsets = s.kind == HJ_SEG_GATHERED ? s.row_offsets : nullptr;
    d.row_begin = rows;
    d.n_rows = s.n_rows;
    d.chunk_id = s.chunk_id;
    d.null_value_id = s.null_value_id;
    d.tile_begin = tiles;
    d.mask_word_begin = words;
    d.kind = s.kind;
    if (ctx->xchg_rowids[side]) {  // local join of hj_join_sharded: the rows carry their RowIDs
      d.kind = kSegExchanged;
      d.attr = ctx->xchg_rowids[side];
    }
    if (ctx->xp_region[side].fill) {  // push exchange: segment i = the receive region of source rank i
      d.kind = kSegExch8;
      d.chunk_id = i;
      d.attr = ctx->xp_region[side].fill + size_t(i) * ctx->xp_region[side].cap_chunks;
      d.row_offsets = ctx->xp_region[side].nchunks + i;
    }
    d.key_type = s.key_type;
    d.attr_width = s.attr_width;
    d.pad1 = d.kind == HJ_SEG_DICT ? s.dict_n : 0u;  // dictionary entries (k_part1 prefetches the next chunk's dictionary)
    row_begin[i] = rows;
    if (s.key_type == HJ_KEY_INT64) pr.has_int64 = true;
    if (s.kind == HJ_SEG_DICT) {
      pr.has_dict = true;
      words += (s.n_rows + 31) / 32;
    }
    if (s.chunk_id != i || d.row_offsets || d.kind == kSegExchanged) identity = false;
    const uint32_t nt = (s.n_rows + tile - 1) / tile;
    tiles += nt;
    rows += s.n_rows;
    if (host && !s.reserved0) {  // reserved0 = 1: arrays already on the device (gathered REFERENCE segment)
      if (s.kind == HJ_SEG_DICT) {
        if (s.dict_n) off_values[i] = plan_copy(s.values, size_t(s.dict_n) * key_bytes(s));
        if (s.n_rows) off_attr[i] = plan_copy(s.attr, size_t(s.n_rows) * s.attr_width);
      } else {
        if (s.n_rows) off_values[i] = plan_copy(s.values, size_t(s.n_rows) * key_bytes(s));
        if (s.n_rows && d.nulls) off_nulls[i] = plan_copy(s.nulls, s.n_rows);
        if (s.n_rows && d.row_offsets) off_ro[i] = plan_copy(s.row_offsets, size_t(s.n_rows) * 4);
      }
    }
  }
  row_begin[n_segs] = rows;

  if (host && stage_bytes) {
    HJ_TRY(ensure(ctx, ctx->stage_in[side], stage_bytes));
    char* base = static_cast<char*>(ctx->stage_in[side].p);
    for (const Copy& c : copies) HJ_CUDA(ctx, cudaMemcpyAsync(base + c.off, c.src, c.bytes, cudaMemcpyHostToDevice, st));
    for (uint32_t i = 0; i < n_segs; ++i) {
      const hj_segment& s = r->segs[i];
      DevSegment& d = pr.segs[i];
      if (s.reserved0) continue;
      if (s.kind == HJ_SEG_DICT) {
        d.values = s.dict_n ? base + off_values[i] : nullptr;
        d.attr = s.n_rows ? base + off_attr[i] : nullptr;
      } else {
        d.values = s.n_rows ? base + off_values[i] : nullptr;
        if (d.nulls) d.nulls = s.n_rows ? reinterpret_cast<const uint8_t*>(base + off_nulls[i]) : nullptr;
        if (d.row_offsets) d.row_offsets = s.n_rows ? reinterpret_cast<const uint32_t*>(base + off_ro[i]) : nullptr;
      }
    }
    for (const Copy& c : copies) pr.h2d_bytes += c.bytes;
  }

  // every chunk but the last exactly 2^k rows, the last <= 2^k: position -> RowID is a shift
  int shift = -1;
  if (n_segs == 1) {
    shift = rows <= (1u << 31) ? 31 : -1;
  } else if (n_segs > 1) {
    const uint32_t first = r->segs[0].n_rows;
    if (first && (first & (first - 1)) == 0) {
      shift = 0;
      while ((1u << shift) != first) ++shift;
      for (uint32_t i = 1; i < n_segs; ++i) {
        const uint32_t n = r->segs[i].n_rows;
        if ((i + 1 < n_segs && n != first) || n > first) {
          shift = -1;
          break;
        }
      }
    }
  }

  HJ_TRY(ensure(ctx, ctx->segs[side], std::max<size_t>(1, n_segs) * sizeof(DevSegment)));
  HJ_TRY(ensure(ctx, ctx->tile_seg[side], std::max<size_t>(1, tiles) * 4));
  HJ_TRY(ensure(ctx, ctx->seg_row_begin[side], (size_t(n_segs) + 1) * 4));
  if (n_segs) HJ_CUDA(ctx, cudaMemcpyAsync(ctx->segs[side].p, pr.segs.data(), n_segs * sizeof(DevSegment), cudaMemcpyHostToDevice, st));
  if (tiles) {
    k_fill_tile_seg<<<std::min<uint32_t>(n_segs, 1024), 128, 0, st>>>(static_cast<const DevSegment*>(ctx->segs[side].p), n_segs, tile,
                                                                    static_cast<uint32_t*>(ctx->tile_seg[side].p));
    HJ_LAUNCH_CHECK(ctx, "k_fill_tile_seg");
  }
  HJ_CUDA(ctx, cudaMemcpyAsync(ctx->seg_row_begin[side].p, row_begin.data(), (size_t(n_segs) + 1) * 4, cudaMemcpyHostToDevice, st));
  // the three vectors above are pageable: cudaMemcpyAsync returns after staging them, so they may die with this frame

  pr.meta.segs = static_cast<const DevSegment*>(ctx->segs[side].p);
  pr.meta.tile_seg = static_cast<const uint32_t*>(ctx->tile_seg[side].p);
  pr.meta.seg_row_begin = static_cast<const uint32_t*>(ctx->seg_row_begin[side].p);
  pr.meta.n_segs = n_segs;
  pr.meta.n_rows = rows;
  pr.meta.n_tiles = tiles;
  pr.meta.uniform_shift = shift;
  pr.meta.identity = identity ? 1u : 0u;
  pr.mask_words = words;
  return HJ_OK;
}

// N1/N2: turns every HJ_SEG_REFERENCE segment of `in` (keys gathered through the PosList) and every HJ_SEG_FOR /
// HJ_SEG_RLE segment (decoded row by row) into a device-resident GATHERED / VALUE segment written by k_gather_refs.
// `out->rel` is the relation to join instead of `in`.
struct RefRewrite {
  std::vector<hj_segment> segs;
  hj_relation rel{};
  bool any = false;
};

// hash_type (hj_key_type): the type the join hashes in. For float/double joins (N3) EVERY chunk goes through the kernel,
// which converts the keys to the hashed type and hands the pipeline their bit patterns as int32 / int64 keys.
int gather_references(hj_ctx* ctx, const hj_relation* in, int side, cudaStream_t st, RefRewrite* out, uint32_t hash_type,
                      bool every_chunk = false) {
  out->rel = *in;
  out->any = false;
  const bool floating = hash_type >= HJ_KEY_FLOAT;
  std::vector<uint32_t> job_segs;
  for (uint32_t i = 0; i < in->n_segs; ++i) {
    const uint8_t kind = in->segs[i].kind;
    const bool bp128 = (kind == HJ_SEG_DICT || kind == HJ_SEG_FOR) && in->segs[i].attr_width == 0;  // decoded here, too
    if (every_chunk || floating || bp128 || kind == HJ_SEG_REFERENCE || kind == HJ_SEG_FOR || kind == HJ_SEG_RLE) job_segs.push_back(i);
  }
  if (job_segs.empty()) return HJ_OK;
  out->any = true;
  out->segs.assign(in->segs, in->segs + in->n_segs);
  const bool host = in->location == HJ_LOC_HOST;
  auto key_bytes = [](uint8_t key_type) { return (key_type == HJ_KEY_INT64 || key_type == HJ_KEY_DOUBLE) ? size_t(8) : size_t(4); };
  auto align = [](size_t x) { return (x + 255) & ~size_t(255); };
  // what the pipeline gets: the chunk's own integer type, or the bit pattern of the hashed floating type
  auto out_key_type = [&](const hj_segment& s) -> uint8_t {
    return !floating ? s.key_type : hash_type == HJ_KEY_FLOAT ? uint8_t(HJ_KEY_INT32) : uint8_t(HJ_KEY_INT64);
  };

  // chunk tables the jobs read from: one per distinct referenced column (normally one: all chunks of a reference table
  // point at the same data table) and a one-entry table per FOR/RLE segment (the segment itself)
  struct Table { const hj_segment* segs; uint32_t n; };
  std::vector<Table> tables;
  std::vector<uint32_t> table_of(job_segs.size());
  for (size_t j = 0; j < job_segs.size(); ++j) {
    const hj_segment& s = in->segs[job_segs[j]];
    Table t;
    if (s.kind == HJ_SEG_REFERENCE) {
      const auto* r = static_cast<const hj_relation*>(s.values);
      t = Table{r->segs, r->n_segs};
    } else {
      t = Table{&s, 1};
    }
    size_t k = 0;
    while (k < tables.size() && tables[k].segs != t.segs) ++k;
    if (k == tables.size()) tables.push_back(t);
    table_of[j] = uint32_t(k);
  }
  // arena: [error flag][DevSegment tables][RefJob table][host data: the tables' arrays, PosLists]
  struct Copy { const void* src; size_t bytes; size_t off; const void** fix; };
  std::vector<Copy> copies;
  size_t arena = 256;
  std::vector<size_t> tab_off(tables.size());
  for (size_t k = 0; k < tables.size(); ++k) {
    tab_off[k] = arena;
    arena += align(std::max<size_t>(1, tables[k].n) * sizeof(DevSegment));
  }
  const size_t job_off = arena;
  arena += align(job_segs.size() * sizeof(RefJob));
  auto plan = [&](const void* src, size_t bytes, const void** fix) {
    copies.push_back({src, bytes, arena, fix});
    arena += align(bytes);
  };
  std::vector<std::vector<DevSegment>> tabs(tables.size());
  std::list<std::vector<uint32_t>> bp_meta;  // host-walked SimdBp128 meta-block offsets, alive until they are copied
  struct BpWalk { const uint32_t* data; uint32_t n_rows; size_t off; const void** fix; };
  std::vector<BpWalk> bp_walks;              // device-resident streams: offsets computed by k_bp128_meta_offsets
  for (size_t k = 0; k < tables.size(); ++k) {
    tabs[k].resize(tables[k].n);
    for (uint32_t c = 0; c < tables[k].n; ++c) {
      const hj_segment& q = tables[k].segs[c];
      DevSegment& d = tabs[k][c];
      std::memset(&d, 0, sizeof(d));
      d.n_rows = q.n_rows;
      d.chunk_id = q.chunk_id;
      d.kind = q.kind;
      d.key_type = q.key_type;
      d.attr_width = q.kind == HJ_SEG_RLE ? 4 : q.attr_width;
      // array lengths per encoding: values / attr / nulls
      size_t n_val = q.n_rows, n_attr = 0, n_null = q.nulls ? q.n_rows : 0;
      if (q.kind == HJ_SEG_DICT) {
        n_val = q.dict_n; n_attr = q.n_rows; n_null = 0;
        d.null_value_id = q.null_value_id;
      } else if (q.kind == HJ_SEG_FOR) {
        n_val = (size_t(q.n_rows) + 2047) / 2048; n_attr = q.n_rows;
      } else if (q.kind == HJ_SEG_RLE) {
        n_val = q.dict_n; n_attr = q.dict_n; n_null = q.nulls ? q.dict_n : 0;
        d.null_value_id = q.dict_n;  // number of runs (decode_point)
      }
      d.values = n_val ? q.values : nullptr;
      d.attr = n_attr ? q.attr : nullptr;
      d.nulls = n_null ? q.nulls : nullptr;
      size_t attr_bytes = n_attr * d.attr_width;
      if (d.attr_width == 0 && n_attr) {  // SimdBp128 stream: the start of every 2048-value meta block is needed
        const uint32_t M = (q.n_rows + 2047u) >> 11;
        if (host) {
          bp_meta.emplace_back(M);
          std::vector<uint32_t>& off = bp_meta.back();
          const auto* words = static_cast<const uint8_t*>(q.attr);
          uint32_t pos = 0;
          for (uint32_t m = 0; m < M; ++m) {
            off[m] = pos;
            const uint8_t* widths = words + size_t(pos) * 16;
            const uint32_t blocks = m + 1 < M ? 16u : (q.n_rows - (m << 11) + 127u) >> 7;
            uint32_t len = 1;
            for (uint32_t b = 0; b < blocks; ++b) len += widths[b];
            pos += len;
          }
          attr_bytes = size_t(pos) * 16;
          plan(off.data(), size_t(M) * 4, reinterpret_cast<const void**>(&d.row_offsets));
        } else {
          bp_walks.push_back({static_cast<const uint32_t*>(q.attr), q.n_rows, arena, reinterpret_cast<const void**>(&d.row_offsets)});
          arena += align(size_t(M) * 4);
        }
      }
      if (host) {
        if (n_val) plan(q.values, n_val * key_bytes(q.key_type), &d.values);
        if (n_attr) plan(q.attr, attr_bytes, &d.attr);
        if (n_null) plan(q.nulls, n_null, reinterpret_cast<const void**>(&d.nulls));
      }
    }
  }
  std::vector<RefJob> jobs(job_segs.size());
  std::vector<const void*> row_offsets(job_segs.size(), nullptr);  // GATHERED chunks keep their reported offsets
  std::vector<size_t> o_out(job_segs.size(), 0), o_outn(job_segs.size(), 0);
  size_t out_bytes = 0;
  uint64_t total_rows = 0;
  for (size_t j = 0; j < job_segs.size(); ++j) {
    const hj_segment& s = in->segs[job_segs[j]];
    jobs[j].pos = nullptr;
    if (s.kind == HJ_SEG_REFERENCE) {
      jobs[j].pos = static_cast<const uint64_t*>(s.attr);
      if (host && s.n_rows) plan(s.attr, size_t(s.n_rows) * 8, reinterpret_cast<const void**>(&jobs[j].pos));
    }
    if (s.kind == HJ_SEG_GATHERED && s.row_offsets) {
      row_offsets[j] = s.row_offsets;
      if (host && s.n_rows) plan(s.row_offsets, size_t(s.n_rows) * 4, &row_offsets[j]);
    }
    o_out[j] = out_bytes;
    out_bytes += align(size_t(s.n_rows) * key_bytes(out_key_type(s)));
    o_outn[j] = out_bytes;
    out_bytes += align(s.n_rows);
    total_rows += s.n_rows;
  }
  if (total_rows >= 0xFFFFFFFFull) return fail(ctx, HJ_ERR_TOO_LARGE, "%llu rows behind REFERENCE/FOR/RLE chunks", (unsigned long long)total_rows);
  HJ_TRY(ensure(ctx, ctx->ref_arena[side], arena));
  HJ_TRY(ensure(ctx, ctx->ref_out[side], std::max<size_t>(256, out_bytes)));
  char* base = static_cast<char*>(ctx->ref_arena[side].p);
  char* obase = static_cast<char*>(ctx->ref_out[side].p);
  HJ_CUDA(ctx, cudaMemsetAsync(base, 0, 256, st));
  for (const Copy& c : copies) {
    HJ_CUDA(ctx, cudaMemcpyAsync(base + c.off, c.src, c.bytes, cudaMemcpyHostToDevice, st));
    *c.fix = base + c.off;  // the table entry / job now points at the staged copy
    ctx->ref_h2d_bytes += c.bytes;
  }
  for (const BpWalk& w : bp_walks) {
    k_bp128_meta_offsets<<<1, 32, 0, st>>>(w.data, w.n_rows, reinterpret_cast<uint32_t*>(base + w.off));
    HJ_LAUNCH_CHECK(ctx, "k_bp128_meta_offsets");
    *w.fix = base + w.off;
  }
  for (size_t k = 0; k < tables.size(); ++k)
    if (tables[k].n)
      HJ_CUDA(ctx, cudaMemcpyAsync(base + tab_off[k], tabs[k].data(), tables[k].n * sizeof(DevSegment), cudaMemcpyHostToDevice, st));
  uint32_t row_begin = 0;
  for (size_t j = 0; j < job_segs.size(); ++j) {
    hj_segment& s = out->segs[job_segs[j]];
    RefJob& jb = jobs[j];
    jb.ref = reinterpret_cast<const DevSegment*>(base + tab_off[table_of[j]]);
    jb.out_values = obase + o_out[j];
    jb.out_nulls = reinterpret_cast<uint8_t*>(obase + o_outn[j]);
    jb.n_ref = tables[table_of[j]].n;
    jb.n_rows = s.n_rows;
    jb.row_begin = row_begin;
    jb.key_type = out_key_type(s);
    jb.hash_type = hash_type;
    jb.build_side = side == 0 ? 1u : 0u;
    row_begin += s.n_rows;
    // the segment the join sees: keys + flags on the device; a gathered ReferenceColumn reports RowIDs (chunk_id, 0..n-1)
    s.kind = (s.kind == HJ_SEG_REFERENCE || s.kind == HJ_SEG_GATHERED) ? HJ_SEG_GATHERED : HJ_SEG_VALUE;
    s.key_type = jb.key_type;
    s.reserved0 = 1;
    s.values = jb.out_values;
    s.nulls = jb.out_nulls;
    s.attr = nullptr;
    s.row_offsets = static_cast<const uint32_t*>(row_offsets[j]);
    s.dict_n = 0;
  }
  HJ_CUDA(ctx, cudaMemcpyAsync(base + job_off, jobs.data(), jobs.size() * sizeof(RefJob), cudaMemcpyHostToDevice, st));
  if (total_rows) {
    const uint32_t grid = uint32_t(std::min<uint64_t>((total_rows + 255) / 256, uint64_t(ctx->n_sms) * 16));
    k_gather_refs<<<grid, 256, 0, st>>>(reinterpret_cast<const RefJob*>(base + job_off), uint32_t(jobs.size()),
                                        uint32_t(total_rows), reinterpret_cast<uint32_t*>(base));
    HJ_LAUNCH_CHECK(ctx, "k_gather_refs");
  }
  // jobs/tabs are pageable vectors: cudaMemcpyAsync has staged them when it returns, so they may die with this frame
  auto* flag = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->pinned) + kPinnedRefFlags) + side;
  HJ_CUDA(ctx, cudaMemcpyAsync(flag, base, 4, cudaMemcpyDeviceToHost, st));
  out->rel.segs = out->segs.data();
  return HJ_OK;
}

// REFERENCE gather for both relations of a join; the join then runs over rw[i].rel where rw[i].any.
// JoinHashTraits (hash_traits.hpp:14-35) over the key columns of both relations, as an hj_key_type
uint32_t hashed_type_of(const hj_relation* build, const hj_relation* probe) {
  uint32_t t = HJ_KEY_INT32;
  bool any_float = false, any_double = false, any_long = false;
  for (const hj_relation* r : {build, probe})
    for (uint32_t i = 0; i < r->n_segs; ++i) {
      const uint8_t k = r->segs[i].key_type;
      any_long = any_long || k == HJ_KEY_INT64;
      any_float = any_float || k == HJ_KEY_FLOAT;
      any_double = any_double || k == HJ_KEY_DOUBLE;
    }
  if (any_double) t = HJ_KEY_DOUBLE; else if (any_float) t = HJ_KEY_FLOAT; else if (any_long) t = HJ_KEY_INT64;
  return t;
}

int gather_both(hj_ctx* ctx, const hj_relation*& build, const hj_relation*& probe, cudaStream_t st, RefRewrite* rw,
                uint32_t hash_type) {
  ctx->ref_h2d_bytes = 0;
  std::memset(static_cast<char*>(ctx->pinned) + kPinnedRefFlags, 0, 8);
  HJ_TRY(gather_references(ctx, build, 0, st, &rw[0], hash_type));
  HJ_TRY(gather_references(ctx, probe, 1, st, &rw[1], hash_type));
  if (rw[0].any) build = &rw[0].rel;
  if (rw[1].any) probe = &rw[1].rel;
  return HJ_OK;
}

// after the join has synchronised the stream: did k_gather_refs meet a RowID outside the referenced table?
int check_reference_flags(hj_ctx* ctx, const RefRewrite* rw) {
  if (!rw[0].any && !rw[1].any) return HJ_OK;
  const auto* flag = reinterpret_cast<const uint32_t*>(static_cast<const char*>(ctx->pinned) + kPinnedRefFlags);
  ctx->timings.bytes_h2d += ctx->ref_h2d_bytes;
  if (flag[0] || flag[1]) {
    ctx->has_result = false;
    return fail(ctx, HJ_ERR_INVALID, "%s relation: a PosList entry points outside the referenced table", flag[0] ? "build" : "probe");
  }
  return HJ_OK;
}

int radix_bits_for(uint64_t n_build, int key_type) {
  const uint64_t cap = key_type == HJ_KEY_INT64 ? Cfg<int64_t>::kCap : Cfg<int32_t>::kCap;
  const uint64_t target = cap / 2;  // mean build partition = half a table: headroom for hash imbalance
  int bits = 0;
  while (bits < kMaxRadixBits && (n_build >> bits) > target) ++bits;
  return bits;
}

struct Launches {
  uint32_t n = 0;
};

template <typename KernelT>
int set_smem(hj_ctx* ctx, KernelT kernel, size_t bytes) {
  if (bytes > 32 * 1024) HJ_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return HJ_OK;
}

// ---- partitioning stages for one relation ------------------------------------------------------------------------------
// The fused path (hj_fused.cuh) and the two-sweep path cut a relation into tiles of different sizes, so the choice is made
// before prepare_relation builds the tile map. Rows received through the exchange carry RowIDs that only the two-sweep
// kernels load (k_scatter1).
// low bits resolved by pass 2 (0: one pass)
inline int pass2_bits(int radix_bits, bool wide) {
  if (radix_bits <= kSinglePassMaxBits) return 0;
  const int b2 = radix_bits / 2 + (wide ? Cfg<int64_t>::kBits2Delta : Cfg<int32_t>::kBits2Delta);
  return std::max(radix_bits - 8, std::min(8, b2));
}
template <typename K> int pass2_bits(int radix_bits) { return pass2_bits(radix_bits, sizeof(K) == 8); }

template <bool PAY>
bool use_fused(const hj_ctx* ctx, int bits2, int side) {
  return !PAY && bits2 > 0 && !(ctx->opts.flags & HJ_FLAG_TWO_SWEEP_PARTITION) && !ctx->xchg_rowids[side];
}
template <typename K, bool PAY>
uint32_t tile_rows(const hj_ctx* ctx, int bits2, int side) {
  return use_fused<PAY>(ctx, bits2, side) ? fused_tile<K>(ctx->xp_region[side].fill != nullptr) : uint32_t(tile1<K>());
}

struct PartState {
  uint32_t G = 0;      // virtual CTAs of K1 / pass 1
  bool fused = false;  // hj_fused.cuh: slab-local pass 1 inside the histogram kernel, fragment-gathering pass 2
  uint32_t gs = 0;     // fused: slabs per pass-2 work item
  bool xch = false;    // fused over push-exchanged tuples: k_part1p<XCH>, prehashed int32 tuples in pass 2
  uint32_t slab = 0;   // fused: tuples per pass-1 slab (tile rows + F1 + 2: room for the holes that align the fragments)
};

// P1's dynamic shared memory: the full-resolution histogram + per virtual CTA one staging tile of full tuples (k_part1) or
// two key buffers and a 16-bit permutation (k_part1p), each with room for the holes that align the fragments
// 1: software-pipelined pass 1 (k_part1p, cp.async gathers); 0: k_part1. Measured on C2 (profiles/r02_sweep_p1_pipelined.md):
// k_part1p 8.54 ms, k_part1 + ld.global.cg 7.75 ms — the pipelined kernel loses (more shared memory = less L1 for the gather's
// line allocations, 8-cycle LDGSTS issue), so k_part1 stays the default for column inputs; k_part1p<XCH> partitions the
// tuples received from the multi-GPU push exchange (hj_xpush.cuh).
#ifndef HJ_TUNE_P1_PIPE
#define HJ_TUNE_P1_PIPE 0
#endif
#ifndef HJ_TUNE_P2_TMA
#define HJ_TUNE_P2_TMA 1  // 1: pass 2 fetches its fragments with TMA bulk copies (k_scatter2t); 0: per-thread loads (k_scatter2f)
#endif
template <typename K>
size_t part1_smem(int radix_bits) {
  const size_t per_stream = HJ_TUNE_P1_PIPE ? size_t(p1_tile<K>()) * 2 * sizeof(K) + (p1_tile<K>() + kSlabPad) * 2
                                            : (p1_tile<K>() + kSlabPad) * sizeof(Tuple<K, false>);
  return ((((size_t(1) << radix_bits) + 1) * 4 + 15) & ~size_t(15)) + size_t(Cfg<K>::kP1Virtual) * per_stream;
}

template <typename K, bool PAY>
int stage_alloc(hj_ctx* ctx, const PreparedRel& pr, int side, int radix_bits, int bits2, PartState* ps) {
  const uint32_t P = 1u << radix_bits, F1 = P >> bits2;
  const uint32_t n = pr.meta.n_rows;
  ps->G = std::max<uint32_t>(1, std::min<uint32_t>(pr.meta.n_tiles, uint32_t(kVirtualCtasPerSm) * ctx->n_sms));
  ps->G = (ps->G + kHistVirtual - 1) / kHistVirtual * kHistVirtual;  // K1 hosts kHistVirtual virtual CTAs per physical CTA
  ps->fused = use_fused<PAY>(ctx, bits2, side);
  ps->xch = ps->fused && ctx->xp_region[side].fill != nullptr;
  if (ps->fused) {
    // an item gathers gs fragments of tile/F1 tuples on average: kP2Fill16/16 of a pass-2 batch. One fragment per thread of the
    // CTA at most (k_scatter2t issues one bulk copy per thread; k_scatter2f also keeps fragment owners in 8 bits)
    const uint32_t gs_max = HJ_TUNE_P2_TMA ? uint32_t(Cfg<K>::kP2Threads) : uint32_t(std::min(Cfg<K>::kP2Threads - 1, 255));
    ps->gs = std::min<uint32_t>(gs_max, std::max<uint64_t>(1, uint64_t(p2_tile<K>()) * Cfg<K>::kP2Fill16 / 16 * F1 / fused_tile<K>(ps->xch)));
    HJ_TRY(ensure(ctx, ctx->frag[side], std::max<size_t>(1, pr.meta.n_tiles) * (F1 + 2) * 2));
    ps->slab = tile_rows<K, PAY>(ctx, bits2, side) + F1 + 2;
    if (uint64_t(pr.meta.n_tiles) * ps->slab >= 0xFFFFFFFFull)
      return fail(ctx, HJ_ERR_TOO_LARGE, "%u tiles of %u slab tuples exceed the 32-bit tuple index of pass 2", pr.meta.n_tiles, ps->slab);
  }
  HJ_TRY(ensure(ctx, ctx->hist[side], (size_t(P) + 1) * 4));
  if (!ps->fused) HJ_TRY(ensure(ctx, ctx->hist1[side], size_t(ps->G) * (F1 + 1) * 4));
  HJ_TRY(ensure(ctx, ctx->part_off[side], (size_t(P) + 2) * 4));
  HJ_TRY(ensure(ctx, ctx->cur2[side], (size_t(P) + 1) * 4));
  HJ_TRY(ensure(ctx, ctx->l1_tiles[side], (size_t(F1) + 1) * 4));
  HJ_TRY(ensure(ctx, ctx->tup_a[side], std::max<size_t>(1, ps->fused ? size_t(pr.meta.n_tiles) * ps->slab : size_t(n)) * sizeof(Tuple<K, PAY>)));
  if (bits2) HJ_TRY(ensure(ctx, ctx->tup_b[side], std::max<size_t>(1, n) * sizeof(Tuple<K, PAY>)));
  if (pr.has_dict && !ps->fused) {
    HJ_TRY(ensure(ctx, ctx->keys_mat[side], std::max<size_t>(1, n) * sizeof(K)));
    HJ_TRY(ensure(ctx, ctx->valid_mask[side], std::max<size_t>(1, pr.mask_words) * 4));
  }
  return HJ_OK;
}

// K1 + K2 + K2b
template <typename K>
int stage_hist(hj_ctx* ctx, const PreparedRel& pr, int side, int radix_bits, int bits2, int null_policy,
               const PartState& ps, cudaStream_t st, Launches* nl) {
  const uint32_t P = 1u << radix_bits, F1 = P >> bits2;
  auto* hist = static_cast<uint32_t*>(ctx->hist[side].p);
  auto* hist1 = static_cast<uint32_t*>(ctx->hist1[side].p);
  HJ_CUDA(ctx, cudaMemsetAsync(hist, 0, (size_t(P) + 1) * 4, st));
  if (ps.fused && ps.xch) {
    if (pr.meta.n_tiles) {
      // per virtual CTA two tuple buffers + a permutation, next to a histogram of up to 2^15 bins, in 227 KB: as many tile
      // streams per SM as fit (2^14 bins: 4 streams of int32 tuples / 2 of int64 tuples; 2^15 bins: half as many)
      constexpr int VCW = sizeof(K) == 4 ? 4 : 2;
      auto smem_for = [&](int vc) {
        return ((((size_t(1) << radix_bits) + 1) * 4 + 15) & ~size_t(15)) +
               size_t(vc) * (size_t(kXpTile) * 2 * sizeof(Tuple<K, false>) + (kXpTile + kSlabPad) * 2);
      };
      const bool wide = smem_for(VCW) + 12 * 1024 <= 227 * 1024;
      auto k = wide ? k_part1p<K, kHistThreads, VCW, kXpTile, true> : k_part1p<K, kHistThreads, VCW / 2, kXpTile, true>;
      const int VC = wide ? VCW : VCW / 2;
      const size_t smem = smem_for(VC);
      HJ_TRY(set_smem(ctx, k, smem));
      const uint32_t grid = std::min<uint32_t>((pr.meta.n_tiles + VC - 1) / VC, ctx->n_sms);
      k<<<grid, kHistThreads, smem, st>>>(pr.meta, static_cast<Tuple<K, false>*>(ctx->tup_a[side].p),
                                          static_cast<uint16_t*>(ctx->frag[side].p), hist, radix_bits, bits2, null_policy, ps.slab);
      HJ_LAUNCH_CHECK(ctx, "k_part1p<XCH>");
      nl->n++;
    }
    k_offsets<1024><<<1, 1024, 0, st>>>(hist, static_cast<uint32_t*>(ctx->part_off[side].p),
                                        static_cast<uint32_t*>(ctx->cur2[side].p),
                                        static_cast<uint32_t*>(ctx->l1_tiles[side].p), radix_bits, bits2, tile2<K>());
    HJ_LAUNCH_CHECK(ctx, "k_offsets");
    nl->n++;
    return HJ_OK;
  }
  if (ps.fused) {
    if (pr.meta.n_tiles) {
      constexpr int VC = Cfg<K>::kP1Virtual;
#if HJ_TUNE_P1_PIPE
      auto k = k_part1p<K, kHistThreads, VC, p1_tile<K>(), false>;
#else
      auto k = k_part1<K, kHistThreads, VC, p1_tile<K>()>;
#endif
      const size_t smem = part1_smem<K>(radix_bits);
      HJ_TRY(set_smem(ctx, k, smem));
      const uint32_t grid = std::min<uint32_t>((pr.meta.n_tiles + VC - 1) / VC, ctx->n_sms);
      k<<<grid, kHistThreads, smem, st>>>(pr.meta, static_cast<Tuple<K, false>*>(ctx->tup_a[side].p),
                                          static_cast<uint16_t*>(ctx->frag[side].p), hist, radix_bits, bits2, null_policy, ps.slab);
      HJ_LAUNCH_CHECK(ctx, "k_part1");
      nl->n++;
    }
    k_offsets<1024><<<1, 1024, 0, st>>>(hist, static_cast<uint32_t*>(ctx->part_off[side].p),
                                        static_cast<uint32_t*>(ctx->cur2[side].p),
                                        static_cast<uint32_t*>(ctx->l1_tiles[side].p), radix_bits, bits2, tile2<K>());
    HJ_LAUNCH_CHECK(ctx, "k_offsets");
    nl->n++;
    return HJ_OK;
  }
  if (pr.meta.n_tiles) {
    auto k1 = k_materialize_hist<K, kHistThreads, kHistVirtual, tile1<K>()>;
    const size_t smem = (size_t(P) + 1) * 4;
    HJ_TRY(set_smem(ctx, k1, smem));
    k1<<<ps.G / kHistVirtual, kHistThreads, smem, st>>>(pr.meta, static_cast<K*>(ctx->keys_mat[side].p),
                                         static_cast<uint32_t*>(ctx->valid_mask[side].p), hist, hist1, radix_bits, bits2,
                                         null_policy);
    HJ_LAUNCH_CHECK(ctx, "k_materialize_hist");
    nl->n++;
  } else {
    HJ_CUDA(ctx, cudaMemsetAsync(hist1, 0, size_t(ps.G) * (F1 + 1) * 4, st));
  }
  k_offsets<1024><<<1, 1024, 0, st>>>(hist, static_cast<uint32_t*>(ctx->part_off[side].p),
                                      static_cast<uint32_t*>(ctx->cur2[side].p),
                                      static_cast<uint32_t*>(ctx->l1_tiles[side].p), radix_bits, bits2, tile2<K>());
  HJ_LAUNCH_CHECK(ctx, "k_offsets");
  k_offsets_l1<256><<<F1 + 1, 256, 0, st>>>(hist1, static_cast<const uint32_t*>(ctx->part_off[side].p), radix_bits, bits2, ps.G);
  HJ_LAUNCH_CHECK(ctx, "k_offsets_l1");
  nl->n += 2;
  return HJ_OK;
}

template <typename K, bool PAY>
ScatterArgs<K, PAY> scatter_args(hj_ctx* ctx, const PreparedRel& pr, int side, int radix_bits, int bits2, int null_policy) {
  ScatterArgs<K, PAY> a{};
  a.rel = pr.meta;
  a.keys_mat = static_cast<const K*>(ctx->keys_mat[side].p);
  a.valid_mask = static_cast<const uint32_t*>(ctx->valid_mask[side].p);
  a.base1 = static_cast<const uint32_t*>(ctx->hist1[side].p);
  a.in = static_cast<const Tuple<K, PAY>*>(ctx->tup_a[side].p);
  a.out = static_cast<Tuple<K, PAY>*>(ctx->tup_a[side].p);
  a.cur2 = static_cast<uint32_t*>(ctx->cur2[side].p);
  a.part_off = static_cast<const uint32_t*>(ctx->part_off[side].p);
  a.l1_tiles = static_cast<const uint32_t*>(ctx->l1_tiles[side].p);
  a.radix_bits = radix_bits;
  a.bits2 = bits2;
  a.null_policy = null_policy;
  return a;
}

template <typename K, bool PAY>
int stage_pass1(hj_ctx* ctx, const PreparedRel& pr, int side, int radix_bits, int bits2, int null_policy,
                const PartState& ps, cudaStream_t st, Launches* nl) {
  if (!pr.meta.n_tiles || ps.fused) return HJ_OK;
  ScatterArgs<K, PAY> a = scatter_args<K, PAY>(ctx, pr, side, radix_bits, bits2, null_policy);
  constexpr int NT1 = scatter_threads<K, PAY>();
  auto k = k_scatter1<K, PAY, NT1, tile1<K>() / NT1>;
  const size_t smem = size_t(tile1<K>()) * sizeof(Tuple<K, PAY>);
  HJ_TRY(set_smem(ctx, k, smem));
  k<<<ps.G, NT1, smem, st>>>(a);
  HJ_LAUNCH_CHECK(ctx, "k_scatter1");
  nl->n++;
  return HJ_OK;
}

template <typename K, bool PAY>
int stage_pass2(hj_ctx* ctx, const PreparedRel& pr, int side, int radix_bits, int bits2, int null_policy,
                const PartState& ps, cudaStream_t st, Launches* nl) {
  if (!pr.meta.n_tiles || !bits2) return HJ_OK;
  if constexpr (!PAY) {
    if (ps.fused) {
      constexpr int NT = Cfg<K>::kP2Threads, kP2MinBlocks = p2_min_blocks<K>();
      constexpr bool PREX = sizeof(K) == 4;  // push-exchanged int32 tuples carry the hash in the key field
#if HJ_TUNE_P2_TMA
      auto k = ps.xch ? k_scatter2t<K, NT, Cfg<K>::kP2Items, kP2MinBlocks, PREX> : k_scatter2t<K, NT, Cfg<K>::kP2Items, kP2MinBlocks, false>;
#else
      auto k = ps.xch ? k_scatter2f<K, NT, Cfg<K>::kP2Items, kP2MinBlocks, PREX> : k_scatter2f<K, NT, Cfg<K>::kP2Items, kP2MinBlocks, false>;
#endif
      const size_t smem = size_t(p2_tile<K>()) * sizeof(Tuple<K, false>);
      HJ_TRY(set_smem(ctx, k, smem));
      const uint32_t F1 = (1u << radix_bits) >> bits2;
      const uint32_t n_digits = F1 + (null_policy == kNullRoute ? 1u : 0u);
      const uint64_t grid = uint64_t((pr.meta.n_tiles + ps.gs - 1) / ps.gs) * n_digits;
      if (grid > 0x7FFFFFFFull) return fail(ctx, HJ_ERR_TOO_LARGE, "pass 2: %llu work items", (unsigned long long)grid);
      k<<<uint32_t(grid), NT, smem, st>>>(static_cast<const Tuple<K, false>*>(ctx->tup_a[side].p),
                                          static_cast<const uint16_t*>(ctx->frag[side].p), pr.meta.n_tiles, ps.slab,
                                          static_cast<Tuple<K, false>*>(ctx->tup_b[side].p),
                                          static_cast<uint32_t*>(ctx->cur2[side].p), radix_bits, bits2, ps.gs, n_digits);
      HJ_LAUNCH_CHECK(ctx, "k_scatter2f");
      nl->n++;
      return HJ_OK;
    }
  }
  ScatterArgs<K, PAY> a = scatter_args<K, PAY>(ctx, pr, side, radix_bits, bits2, kNullAsValue);
  a.out = static_cast<Tuple<K, PAY>*>(ctx->tup_b[side].p);
  constexpr int NT2 = scatter_threads<K, PAY>();
  auto k = k_scatter2<K, PAY, NT2, tile2<K>() / NT2>;
  const size_t smem = size_t(tile2<K>()) * sizeof(Tuple<K, PAY>);
  HJ_TRY(set_smem(ctx, k, smem));
  const uint32_t F1 = (1u << radix_bits) >> bits2;
  const uint32_t grid = (pr.meta.n_rows + tile2<K>() - 1) / tile2<K>() + F1;  // upper bound on sum(ceil(part / tile))
  k<<<grid, NT2, smem, st>>>(a);
  HJ_LAUNCH_CHECK(ctx, "k_scatter2");
  nl->n++;
  return HJ_OK;
}

struct Readback {  // layout of the pinned readback area (offset 0; REFERENCE error flags live at kPinnedRefFlags)
  OutState out;
  PlanOut plan;
  uint32_t n_build_tuples;
  uint32_t n_probe_tuples;
};

float ev_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0.f;
  }
  return ms;
}

template <typename K, bool PAY, int RM, int CLS>
int launch_build_probe_cls(hj_ctx* ctx, const JoinArgs<K, PAY>& a, cudaStream_t st) {
  auto k5 = k_build_probe<K, PAY, kJoinThreads, Cfg<K>::kNS, Cfg<K>::kCap, kProbeV, RM, CLS>;
  const size_t smem5 = (size_t(Cfg<K>::kNS) + 1) * sizeof(K) + (size_t(Cfg<K>::kNS) + 2) * 4 + size_t(Cfg<K>::kCap) * 4 +
                       (RM == kRidXch ? size_t(Cfg<K>::kCap) : 0);  // kRidXch: + one source-rank byte per grouped build row
  HJ_TRY(set_smem(ctx, k5, smem5));
  k5<<<ctx->n_sms, kJoinThreads, smem5, st>>>(a);
  HJ_LAUNCH_CHECK(ctx, "k_build_probe");
  return HJ_OK;
}

template <typename K, bool PAY, int RM>
int launch_build_probe(hj_ctx* ctx, const JoinArgs<K, PAY>& a, cudaStream_t st) {
  switch (a.mode) {
    case HJ_MODE_INNER: return launch_build_probe_cls<K, PAY, RM, kClsInner>(ctx, a, st);
    case HJ_MODE_LEFT:
    case HJ_MODE_RIGHT: return launch_build_probe_cls<K, PAY, RM, kClsOuter>(ctx, a, st);
    case HJ_MODE_SEMI: return launch_build_probe_cls<K, PAY, RM, kClsSemi>(ctx, a, st);
    default: return launch_build_probe_cls<K, PAY, RM, kClsAnti>(ctx, a, st);
  }
}

// K4..K8 for one (partitioned build, partitioned probe) pair: plan, build+probe, finalize, compaction; re-probes once with
// exact buffers when the first estimate was too small. `sb`/`sp` = which per-side buffer sets hold the build / probe partitions (sp: 1, or
// 1/2 alternating for the slices of a streamed join), `oset` = which result buffers to fill. Leaves the counters in the
// pinned Readback area.
template <typename K, bool PAY>
int probe_phase(hj_ctx* ctx, const PreparedRel& pb, const PreparedRel& pp, const PartState& pst, int sb, int sp, int oset,
                int radix_bits, int bits2, int mode, int probe_null_policy, int payload, uint64_t n_build_rows,
                uint64_t n_probe_rows, cudaStream_t st, Launches* nl, int* passes_out, cudaEvent_t ev_probe,
                cudaEvent_t ev_end) {
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const uint32_t P = 1u << radix_bits;
  const auto* tb = static_cast<const Tuple<K, PAY>*>(bits2 ? ctx->tup_b[sb].p : ctx->tup_a[sb].p);
  const auto* tp = static_cast<const Tuple<K, PAY>*>(bits2 ? ctx->tup_b[sp].p : ctx->tup_a[sp].p);
  // routed NULL rows stay in pass 1's output; the fused path's pass 2 moves them behind the last partition of ITS output
  const auto* tp1 = pst.fused ? tp : static_cast<const Tuple<K, PAY>*>(ctx->tup_a[sp].p);

  // ---- plan
  const uint32_t cap = Cfg<K>::kCap;
  const size_t items_capacity = size_t(P) + n_build_rows / cap + n_probe_rows / kProbeSlice + 64;
  HJ_TRY(ensure(ctx, ctx->items, items_capacity * sizeof(WorkItem)));
  HJ_TRY(ensure(ctx, ctx->fin, size_t(P) * sizeof(WorkItem)));
  HJ_TRY(ensure(ctx, ctx->plan, sizeof(PlanOut)));
  HJ_TRY(ensure(ctx, ctx->queue, 256));
  HJ_TRY(ensure(ctx, ctx->out_state, sizeof(OutState)));
  HJ_TRY(ensure(ctx, ctx->out_tails, sizeof(OutTail) * kMaxTails));
  HJ_TRY(ensure(ctx, ctx->compact_plan, sizeof(CompactHdr)));
  const size_t matched_bytes = (size_t(n_probe_rows) / 32 + 2) * 4;
  if (mode != HJ_MODE_INNER) HJ_TRY(ensure(ctx, ctx->matched, matched_bytes));
  auto* off_b = static_cast<uint32_t*>(ctx->part_off[sb].p);
  auto* off_p = static_cast<uint32_t*>(ctx->part_off[sp].p);
  k_plan<1024><<<1, 1024, 0, st>>>(off_b, off_p, radix_bits, mode, cap, kProbeSlice,
                                   (uint32_t)std::min<size_t>(items_capacity, 0xFFFFFFFFu),
                                   static_cast<WorkItem*>(ctx->items.p), static_cast<WorkItem*>(ctx->fin.p),
                                   static_cast<PlanOut*>(ctx->plan.p));
  HJ_LAUNCH_CHECK(ctx, "k_plan");
  nl->n++;

  // ---- output buffers: `rows` result rows + one chunk of slack per result writer (K5: every warp; K6/K7: every CTA).
  // Chunk size: a writer should go through >= ~4 chunks so that the slack stays a fraction of the result.
  const uint64_t n_writers = uint64_t(ctx->n_sms) * (kJoinThreads / 32) + (mode != HJ_MODE_INNER ? ctx->n_sms : 0) +
                             (probe_null_policy == kNullRoute ? ctx->n_sms : 0);
  auto chunk_shift_for = [&](uint64_t rows) {
    int sh = kMinChunkShift;
    while (sh < kMaxChunkShift && (uint64_t(1) << sh) * n_writers * 4 < rows) ++sh;
    return sh;
  };
  auto alloc_rows = [&](uint64_t rows, int sh) { return (((rows + (uint64_t(1) << sh) - 1) >> sh) + n_writers + 1) << sh; };
  uint64_t want_rows = ctx->opts.output_capacity ? ctx->opts.output_capacity : n_probe_rows;
  int chunk_shift = chunk_shift_for(want_rows);
  uint64_t capacity = alloc_rows(want_rows, chunk_shift);

  JoinArgs<K, PAY> a{};
  a.build = tb;
  a.probe = tp;
  a.items = static_cast<const WorkItem*>(ctx->items.p);
  a.plan = static_cast<const PlanOut*>(ctx->plan.p);
  a.queue_head = static_cast<uint32_t*>(ctx->queue.p);
  a.build_rel = pb.meta;
  a.probe_rel = pp.meta;
  a.out_state = static_cast<OutState*>(ctx->out_state.p);
  a.out_tails = static_cast<OutTail*>(ctx->out_tails.p);
  a.matched = static_cast<uint32_t*>(ctx->matched.p);
  a.radix_bits = radix_bits;
  a.mode = mode;
  a.payload = payload;  // 0: positions, 1: tuples carry RowIDs (general exchange), 2: (rank, position) (push exchange)
  a.bits2 = bits2;
  a.xmap[0] = ctx->xp_map[0];
  a.xmap[1] = ctx->xp_map[1];
  const bool fast_rowid = pb.meta.uniform_shift >= 0 && pb.meta.identity && pp.meta.uniform_shift >= 0 && pp.meta.identity;

  auto* rb = static_cast<Readback*>(ctx->pinned);
  int passes = 0;
  for (;;) {
    ++passes;
    if (!one_side) HJ_TRY(ensure(ctx, ctx->out_build[oset], capacity * 8));
    HJ_TRY(ensure(ctx, ctx->out_probe[oset], capacity * 8));
    a.out_build = one_side ? nullptr : static_cast<uint64_t*>(ctx->out_build[oset].p);
    a.out_probe = static_cast<uint64_t*>(ctx->out_probe[oset].p);
    a.capacity = capacity;
    a.chunk_shift = chunk_shift;
    const size_t max_chunks = size_t(capacity >> chunk_shift) + 1;
    HJ_TRY(ensure(ctx, ctx->compact_used, max_chunks * 4));
    HJ_TRY(ensure(ctx, ctx->compact_hole, (max_chunks + 1) * 4));
    HJ_TRY(ensure(ctx, ctx->compact_src, (max_chunks + 1) * 4));
    HJ_CUDA(ctx, cudaMemsetAsync(ctx->queue.p, 0, 4, st));
    HJ_CUDA(ctx, cudaMemsetAsync(ctx->out_state.p, 0, sizeof(OutState), st));
    if (mode != HJ_MODE_INNER) HJ_CUDA(ctx, cudaMemsetAsync(ctx->matched.p, 0, matched_bytes, st));
    if (payload == 2) {
      if constexpr (!PAY) HJ_TRY((launch_build_probe<K, PAY, kRidXch>(ctx, a, st)));
      else return fail(ctx, HJ_ERR_INVALID, "push-exchanged tuples are position tuples");
    } else if (payload) {
      if constexpr (sizeof(Tuple<K, PAY>) > 8) HJ_TRY((launch_build_probe<K, PAY, kRidPayload>(ctx, a, st)));
      else return fail(ctx, HJ_ERR_INVALID, "RowID-carrying tuples need a tuple type with a payload");
    } else if constexpr (PAY) {
      return fail(ctx, HJ_ERR_INVALID, "payload tuples are only used by the sharded local join");
    } else if (fast_rowid) {
      HJ_TRY((launch_build_probe<K, PAY, kRidFast>(ctx, a, st)));
    } else {
      HJ_TRY((launch_build_probe<K, PAY, kRidGeneric>(ctx, a, st)));
    }
    nl->n++;
    if (passes == 1 && ev_probe) HJ_CUDA(ctx, cudaEventRecord(ev_probe, st));
    if (mode != HJ_MODE_INNER) {
      k_finalize_multi<K, PAY, 256><<<ctx->n_sms, 256, 0, st>>>(a, static_cast<const WorkItem*>(ctx->fin.p));
      HJ_LAUNCH_CHECK(ctx, "k_finalize_multi");
      nl->n++;
    }
    if (probe_null_policy == kNullRoute) {
      k_emit_null_probe<K, PAY, 256><<<ctx->n_sms, 256, 0, st>>>(a, tp1, off_p);
      HJ_LAUNCH_CHECK(ctx, "k_emit_null_probe");
      nl->n++;
    }
    k_compact_plan<<<1, 1024, 0, st>>>(a.out_state, a.out_tails, chunk_shift, capacity, static_cast<uint32_t*>(ctx->compact_used.p),
                                       static_cast<uint32_t*>(ctx->compact_hole.p), static_cast<uint32_t*>(ctx->compact_src.p),
                                       static_cast<CompactHdr*>(ctx->compact_plan.p));
    HJ_LAUNCH_CHECK(ctx, "k_compact_plan");
    k_compact_move<<<ctx->n_sms * 8, 256, 0, st>>>(static_cast<const CompactHdr*>(ctx->compact_plan.p), a.out_state, chunk_shift,
                                                   static_cast<const uint32_t*>(ctx->compact_used.p),
                                                   static_cast<const uint32_t*>(ctx->compact_hole.p),
                                                   static_cast<const uint32_t*>(ctx->compact_src.p), a.out_build, a.out_probe);
    HJ_LAUNCH_CHECK(ctx, "k_compact_move");
    nl->n += 2;
    HJ_CUDA(ctx, cudaMemcpyAsync(&rb->out, ctx->out_state.p, sizeof(OutState), cudaMemcpyDeviceToHost, st));
    HJ_CUDA(ctx, cudaMemcpyAsync(&rb->plan, ctx->plan.p, sizeof(PlanOut), cudaMemcpyDeviceToHost, st));
    HJ_CUDA(ctx, cudaMemcpyAsync(&rb->n_build_tuples, off_b + P, 4, cudaMemcpyDeviceToHost, st));
    HJ_CUDA(ctx, cudaMemcpyAsync(&rb->n_probe_tuples, off_p + P + 1, 4, cudaMemcpyDeviceToHost, st));
    if (ev_end) HJ_CUDA(ctx, cudaEventRecord(ev_end, st));
    HJ_CUDA(ctx, cudaStreamSynchronize(st));
    if ((rb->out.n_chunks << chunk_shift) <= capacity) break;  // every reserved chunk lay inside the buffers
    if (passes >= 2) return fail(ctx, HJ_ERR_CUDA, "result size changed between probe passes (%llu chunks > capacity %llu rows)",
                                 rb->out.n_chunks, (unsigned long long)capacity);
    chunk_shift = chunk_shift_for(rb->out.n_rows);  // the exact row count is known now: re-probe into buffers that fit
    capacity = alloc_rows(rb->out.n_rows, chunk_shift);
  }

  *passes_out = passes;
  return HJ_OK;
}

template <typename K, bool PAY = false>
int run_join(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
             hj_result_info* info, bool payload = false) {
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  Launches nl;
  cudaEvent_t* ev = ctx->ev;
  // events: 0 start, 1 after staging, 2 after K1/K2, 3 after pass 1, 4 after pass 2, 5 after build+probe, 6 end
  HJ_CUDA(ctx, cudaEventRecord(ev[0], st));

  const int radix_bits = ctx->opts.radix_bits >= 0 ? std::min(ctx->opts.radix_bits, kMaxRadixBits)
                                                   : radix_bits_for(build->n_rows, sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32);
  const int bits2 = pass2_bits<K>(radix_bits);

  PreparedRel pb, pp;
  HJ_TRY(prepare_relation(ctx, build, 0, tile_rows<K, PAY>(ctx, bits2, 0), st, &pb));
  HJ_TRY(prepare_relation(ctx, probe, 1, tile_rows<K, PAY>(ctx, bits2, 1), st, &pp));
  HJ_CUDA(ctx, cudaEventRecord(ev[1], st));

  const uint32_t P = 1u << radix_bits;
  const int probe_null_policy = !outer ? kNullDrop : ((ctx->opts.flags & HJ_FLAG_SQL_NULL_PROBE) ? kNullRoute : kNullAsValue);

  PartState sb, sp;
  HJ_TRY(stage_alloc<K, PAY>(ctx, pb, 0, radix_bits, bits2, &sb));
  HJ_TRY(stage_alloc<K, PAY>(ctx, pp, 1, radix_bits, bits2, &sp));
  HJ_TRY(stage_hist<K>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl));
  HJ_TRY(stage_hist<K>(ctx, pp, 1, radix_bits, bits2, probe_null_policy, sp, st, &nl));
  HJ_CUDA(ctx, cudaEventRecord(ev[2], st));
  HJ_TRY(stage_pass1<K, PAY>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl));
  HJ_TRY(stage_pass1<K, PAY>(ctx, pp, 1, radix_bits, bits2, probe_null_policy, sp, st, &nl));
  HJ_CUDA(ctx, cudaEventRecord(ev[3], st));
  HJ_TRY(stage_pass2<K, PAY>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl));
  HJ_TRY(stage_pass2<K, PAY>(ctx, pp, 1, radix_bits, bits2, probe_null_policy, sp, st, &nl));
  HJ_CUDA(ctx, cudaEventRecord(ev[4], st));
  int passes = 0;
  HJ_TRY((probe_phase<K, PAY>(ctx, pb, pp, sp, 0, 1, 0, radix_bits, bits2, mode, probe_null_policy, payload ? 1 : 0, build->n_rows,
                              probe->n_rows, st, &nl, &passes, ev[5], ev[6])));
  auto* rb = static_cast<Readback*>(ctx->pinned);
  ctx->has_result = true;
  ctx->res_build = ctx->out_build[0].p;
  ctx->res_probe = ctx->out_probe[0].p;
  ctx->n_result = rb->out.n_rows;
  ctx->n_sides = one_side ? 1 : 2;
  info->n_rows = rb->out.n_rows;
  info->n_sides = ctx->n_sides;
  info->radix_bits = radix_bits;
  info->n_build_tuples = rb->n_build_tuples;
  info->n_probe_tuples = rb->n_probe_tuples;
  info->n_partitions = P;
  info->n_overflow_partitions = rb->plan.n_overflow_parts;
  info->n_probe_passes = passes;
  info->n_probe_slices = 0;

  hj_timings& t = ctx->timings;
  std::memset(&t, 0, sizeof(t));
  t.ms[HJ_PH_TOTAL] = ev_ms(ev[0], ev[6]);
  t.ms[HJ_PH_H2D] = ev_ms(ev[0], ev[1]);
  t.ms[HJ_PH_MATERIALIZE] = ev_ms(ev[1], ev[2]);
  t.ms[HJ_PH_SCATTER1] = ev_ms(ev[2], ev[3]);
  t.ms[HJ_PH_SCATTER2] = ev_ms(ev[3], ev[4]);
  t.ms[HJ_PH_BUILD_PROBE] = passes == 1 ? ev_ms(ev[4], ev[5]) : 0.f;
  t.ms[HJ_PH_FINALIZE] = passes == 1 ? ev_ms(ev[5], ev[6]) : 0.f;
  t.bytes_h2d = pb.h2d_bytes + pp.h2d_bytes;
  t.n_kernel_launches = nl.n;
  return HJ_OK;
}

// Streamed join for HOST-resident inputs with bound HOST outputs (hj_bind_output): the probe relation is cut into slices of
// whole chunks; while slice i is partitioned and probed, slice i+1 is copied in and the PosLists of slice i-1 are copied
// out, so the PCIe link runs in both directions at once and the join hides behind the result download. Valid for every
// mode: the build side is complete before the first probe, and every decision of Left/Right/Semi/Anti is per probe row.
template <typename K>
int run_join_streamed(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                      hj_result_info* info, uint32_t n_slices) {
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  Launches nl;
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
  const int radix_bits = ctx->opts.radix_bits >= 0 ? std::min(ctx->opts.radix_bits, kMaxRadixBits)
                                                   : radix_bits_for(build->n_rows, sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32);
  const int bits2 = pass2_bits<K>(radix_bits);
  const int probe_null_policy = !outer ? kNullDrop : ((ctx->opts.flags & HJ_FLAG_SQL_NULL_PROBE) ? kNullRoute : kNullAsValue);

  // slices: runs of whole chunks with about the same number of rows
  std::vector<uint32_t> cut{0};
  {
    const uint64_t per = (probe->n_rows + n_slices - 1) / n_slices;
    uint64_t acc = 0;
    for (uint32_t i = 0; i < probe->n_segs; ++i) {
      acc += probe->segs[i].n_rows;
      if (acc >= per && i + 1 < probe->n_segs) {
        cut.push_back(i + 1);
        acc = 0;
      }
    }
    cut.push_back(probe->n_segs);
  }
  const uint32_t n = uint32_t(cut.size()) - 1;
  auto slice_rel = [&](uint32_t i) {
    hj_relation r{};
    r.segs = probe->segs + cut[i];
    r.n_segs = cut[i + 1] - cut[i];
    r.location = probe->location;
    r.n_rows = 0;
    for (uint32_t k = 0; k < r.n_segs; ++k) r.n_rows += r.segs[k].n_rows;
    return r;
  };

  PreparedRel pb, pp[2];
  PartState sb, sp[2];
  hj_relation rel[2];
  // stage slice 0 while the build side is copied in and partitioned
  rel[0] = slice_rel(0);
  HJ_TRY(prepare_relation(ctx, &rel[0], 1, tile_rows<K, false>(ctx, bits2, 1), ctx->copy_in, &pp[0]));
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev_in[0], ctx->copy_in));
  HJ_TRY(prepare_relation(ctx, build, 0, tile_rows<K, false>(ctx, bits2, 0), st, &pb));
  HJ_TRY((stage_alloc<K, false>(ctx, pb, 0, radix_bits, bits2, &sb)));
  HJ_TRY(stage_hist<K>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl));
  HJ_TRY((stage_pass1<K, false>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl)));
  HJ_TRY((stage_pass2<K, false>(ctx, pb, 0, radix_bits, bits2, kNullDrop, sb, st, &nl)));

  auto* rb = static_cast<Readback*>(ctx->pinned);
  uint64_t total = 0, probe_tuples = 0, h2d = pb.h2d_bytes;
  uint32_t overflow_parts = 0;
  int max_passes = 0;
  int status = HJ_OK;
  for (uint32_t i = 0; i < n; ++i) {
    const int cur = i & 1, nxt = (i + 1) & 1;
    if (i + 1 < n) {  // stage the next slice into the other buffer set, once the slice that used it has been joined
      if (i >= 1) HJ_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, ctx->ev_done[nxt], 0));
      rel[nxt] = slice_rel(i + 1);
      pp[nxt] = PreparedRel{};
      HJ_TRY(prepare_relation(ctx, &rel[nxt], 1 + nxt, tile_rows<K, false>(ctx, bits2, 1 + nxt), ctx->copy_in, &pp[nxt]));
      HJ_CUDA(ctx, cudaEventRecord(ctx->ev_in[nxt], ctx->copy_in));
    }
    HJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_in[cur], 0));
    if (i >= 2) HJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_out[cur], 0));  // result buffers of this set are free again
    const int side = 1 + cur;
    HJ_TRY((stage_alloc<K, false>(ctx, pp[cur], side, radix_bits, bits2, &sp[cur])));
    HJ_TRY(stage_hist<K>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl));
    HJ_TRY((stage_pass1<K, false>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl)));
    HJ_TRY((stage_pass2<K, false>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl)));
    int passes = 0;
    HJ_TRY((probe_phase<K, false>(ctx, pb, pp[cur], sp[cur], 0, side, cur, radix_bits, bits2, mode, probe_null_policy, false,
                                  build->n_rows, rel[cur].n_rows, st, &nl, &passes, nullptr, nullptr)));
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_done[cur], st));
    const uint64_t rows = rb->out.n_rows;  // probe_phase synchronised the stream
    probe_tuples += rb->n_probe_tuples;
    overflow_parts = std::max(overflow_parts, rb->plan.n_overflow_parts);
    max_passes = std::max(max_passes, passes);
    h2d += pp[cur].h2d_bytes;
    if (total + rows > ctx->bound_capacity) {
      status = fail(ctx, HJ_ERR_CAPACITY, "bound output holds %llu rows, the result has more (>= %llu after %u of %u probe slices); "
                    "join again without hj_bind_output and size the buffers from hj_result_info.n_rows",
                    (unsigned long long)ctx->bound_capacity, (unsigned long long)(total + rows), i + 1, n);
      break;
    }
    if (rows) {
      HJ_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ctx->ev_done[cur], 0));
      if (!one_side)
        HJ_CUDA(ctx, cudaMemcpyAsync(ctx->bound_build + total, ctx->out_build[cur].p, rows * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
      HJ_CUDA(ctx, cudaMemcpyAsync(ctx->bound_probe + total, ctx->out_probe[cur].p, rows * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
    }
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_out[cur], ctx->copy_out));
    total += rows;
  }
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[6], ctx->copy_out));
  HJ_CUDA(ctx, cudaStreamSynchronize(ctx->copy_in));
  HJ_CUDA(ctx, cudaStreamSynchronize(ctx->copy_out));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (status != HJ_OK) return status;

  ctx->has_result = false;  // nothing left on the device: the PosLists are in the bound buffers
  ctx->result_in_bound = true;
  ctx->n_result = total;
  ctx->n_sides = one_side ? 1 : 2;
  info->n_rows = total;
  info->n_sides = ctx->n_sides;
  info->radix_bits = radix_bits;
  info->n_build_tuples = rb->n_build_tuples;
  info->n_probe_tuples = probe_tuples;
  info->n_partitions = 1u << radix_bits;
  info->n_overflow_partitions = overflow_parts;
  info->n_probe_passes = max_passes;
  info->n_probe_slices = n;
  hj_timings& t = ctx->timings;
  std::memset(&t, 0, sizeof(t));
  t.ms[HJ_PH_TOTAL] = ev_ms(ctx->ev[0], ctx->ev[6]);
  t.bytes_h2d = h2d;
  t.bytes_d2h = total * 8 * ctx->n_sides;
  t.n_kernel_launches = nl.n;
  return HJ_OK;
}

int check_ctx(hj_ctx* ctx) {
  if (!ctx) return HJ_ERR_INVALID;
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return fail(ctx, HJ_ERR_CUDA, "cudaSetDevice(%d): %s", ctx->device, cudaGetErrorString(e));
  return HJ_OK;
}

// ------------------------------------------------------------------------------------------------ TableScan (N4)

constexpr size_t kPinnedScanTotals = 3616, kPinnedAggregate = 3648;  // + one AggPartial (64 B), below kPinnedCounts  // two words inside the pinned readback area (behind kPinnedRefFlags)

// K: integer type of the stored keys; V: type the comparison runs in (= K, or float / double over their bit patterns);
// `value` the literal as V, `id_value` the literal the dictionary bounds are searched with (integer columns).
template <typename K, typename V>
int run_scan(hj_ctx* ctx, const hj_relation* column, const hj_relation* original, int predicate, V value, int64_t id_value,
             int slot, uint32_t scan_flags, cudaStream_t st, hj_scan_info* info, const uint32_t* given_bounds = nullptr) {
  PreparedRel pr;
  HJ_TRY(prepare_relation(ctx, column, 1, kScanTile, st, &pr));
  const uint32_t n_segs = pr.meta.n_segs, n_tiles = pr.meta.n_tiles;
  HJ_TRY(ensure(ctx, ctx->scan_bounds, std::max<size_t>(1, n_segs) * 8));
  HJ_TRY(ensure(ctx, ctx->scan_mask, std::max<size_t>(1, n_tiles) * kScanWordsPerTile * 4));
  HJ_TRY(ensure(ctx, ctx->scan_tile_count, std::max<size_t>(1, n_tiles) * 4));
  HJ_TRY(ensure(ctx, ctx->scan_tile_base, (size_t(n_tiles) + 1) * 4));
  HJ_TRY(ensure(ctx, ctx->scan_totals, 8));
  HJ_TRY(ensure(ctx, ctx->scan_chunk_begin[slot], (size_t(n_segs) + 1) * 4));
  auto* bounds = static_cast<uint32_t*>(ctx->scan_bounds.p);
  auto* mask = static_cast<uint32_t*>(ctx->scan_mask.p);
  auto* tile_count = static_cast<uint32_t*>(ctx->scan_tile_count.p);
  auto* tile_base = static_cast<uint32_t*>(ctx->scan_tile_base.p);
  auto* totals = static_cast<uint32_t*>(ctx->scan_totals.p);
  auto* chunk_begin = static_cast<uint32_t*>(ctx->scan_chunk_begin[slot].p);

  // HJ_SCAN_DEREFERENCE: REFERENCE chunks emit their own PosList entries — the device address of every such chunk's PosList
  const uint64_t* const* ref_ptrs = nullptr;
  if (scan_flags & HJ_SCAN_DEREFERENCE) {
    std::vector<const uint64_t*> ptrs(n_segs, nullptr);
    bool any = false;
    for (uint32_t i = 0; i < n_segs; ++i)
      if (original->segs[i].kind == HJ_SEG_REFERENCE) {
        ptrs[i] = static_cast<const uint64_t*>(original->segs[i].attr);
        any = true;
      }
    if (any) {
      HJ_TRY(ensure(ctx, ctx->scan_ref_ptrs, size_t(n_segs) * sizeof(void*)));
      HJ_CUDA(ctx, cudaMemcpyAsync(ctx->scan_ref_ptrs.p, ptrs.data(), size_t(n_segs) * sizeof(void*), cudaMemcpyHostToDevice, st));
      HJ_CUDA(ctx, cudaStreamSynchronize(st));  // `ptrs` is pageable and dies with this block
      ref_ptrs = static_cast<const uint64_t* const*>(ctx->scan_ref_ptrs.p);
    }
  }

  const uint32_t grid = std::max<uint32_t>(1, std::min<uint32_t>(n_tiles, uint32_t(ctx->n_sms) * 8));
  if (n_segs && given_bounds) {
    // hj_scan_value_ids: the caller searched the dictionaries (pageable memory: the copy is staged before the call returns)
    HJ_CUDA(ctx, cudaMemcpyAsync(bounds, given_bounds, size_t(n_segs) * 8, cudaMemcpyHostToDevice, st));
  } else if (n_segs) {
    k_scan_bounds<<<(n_segs + 127) / 128, 128, 0, st>>>(pr.meta.segs, n_segs, static_cast<long long>(id_value), bounds);
    HJ_LAUNCH_CHECK(ctx, "k_scan_bounds");
  }
  if (n_tiles) {
    k_scan_mask<K, V><<<grid, kScanThreads, 0, st>>>(pr.meta, predicate, value, bounds, mask, tile_count);
    HJ_LAUNCH_CHECK(ctx, "k_scan_mask");
  }
  k_scan_prefix<1024><<<1, 1024, 0, st>>>(tile_count, n_tiles, pr.meta.segs, n_segs, tile_base, chunk_begin, totals);
  HJ_LAUNCH_CHECK(ctx, "k_scan_prefix");
  auto* host_totals = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->pinned) + kPinnedScanTotals);
  HJ_CUDA(ctx, cudaMemcpyAsync(host_totals, totals, 8, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  const uint64_t n_rows = host_totals[0];
  HJ_TRY(ensure(ctx, ctx->scan_pos[slot], std::max<uint64_t>(1, n_rows) * 8));
  if (n_rows) {
    k_scan_write<<<grid, kScanThreads, 0, st>>>(pr.meta, mask, tile_base, ref_ptrs, static_cast<uint64_t*>(ctx->scan_pos[slot].p));
    HJ_LAUNCH_CHECK(ctx, "k_scan_write");
    HJ_CUDA(ctx, cudaStreamSynchronize(st));
  }
  ctx->scan_rows[slot] = n_rows;
  ctx->scan_chunks[slot] = n_segs;
  ctx->scan_valid[slot] = true;
  info->n_rows = n_rows;
  info->n_chunks = n_segs;
  info->n_nonempty_chunks = host_totals[1];
  return HJ_OK;
}


// hj_aggregate_grouped*: the hash table (power of two >= `want_slots`), the accumulate launches, then count + emit.
// Sizing (group_run): a caller's max_groups is taken at its word — 2 x max_groups slots, one attempt. Without one the number
// of groups is unknown until the pass is over, and 2 x n_rows slots of 64 bytes (round 1) cost 128 bytes of HBM per input row:
// the table starts at kGroupFirstSlots (or 2 x n_rows if that is less) and, when an accumulate launch reports it full
// (bounded probing, see GroupTable::max_probes), is rebuilt 8 x larger and the launches are repeated, up to 2 x n_rows slots,
// which cannot overflow. A retry costs one more pass over the group-by and aggregate columns; GROUP BYs with up to
// kGroupFirstSlots / 2 groups never retry.
constexpr uint64_t kGroupFirstSlots = uint64_t(1) << 20;  // 64 MB
constexpr uint64_t kGroupProbeLimit = 1024;                // probes per lookup before a non-final table counts as full
constexpr uint64_t kGroupProbeLimitFinal = 1 << 16;        // ... and the final one (half empty unless max_groups was wrong)
constexpr int kGroupTableFull = 1;                         // group_finish -> group_run: the overflow flag was set

int group_begin(hj_ctx* ctx, uint64_t want_slots, bool final_attempt, cudaStream_t st, GroupTable* out, bool row_keys = false) {
  const uint64_t want = std::max<uint64_t>(1024, want_slots);
  uint64_t cap = 1024;
  while (cap < want) cap <<= 1;
  if (cap > (uint64_t(1) << 31)) return fail(ctx, HJ_ERR_TOO_LARGE, "group table of %llu slots: pass max_groups", (unsigned long long)cap);
  const size_t entries = size_t(cap) + 2;
  HJ_TRY(ensure(ctx, ctx->group_table, entries * 8 * kGroupArrays));
  HJ_TRY(ensure(ctx, ctx->group_misc, 16));
  char* base = static_cast<char*>(ctx->group_table.p);
  GroupTable& t = *out;
  t.key = reinterpret_cast<long long*>(base);
  t.rows = reinterpret_cast<unsigned long long*>(base + entries * 8);
  t.values = reinterpret_cast<unsigned long long*>(base + entries * 16);
  t.sum_i = reinterpret_cast<long long*>(base + entries * 24);
  t.sum_d = reinterpret_cast<double*>(base + entries * 32);
  t.mn = reinterpret_cast<long long*>(base + entries * 40);
  t.mx = reinterpret_cast<long long*>(base + entries * 48);
  t.first = reinterpret_cast<unsigned long long*>(base + entries * 56);
  t.cap = cap;
  t.max_probes = std::min(cap, final_attempt ? kGroupProbeLimitFinal : kGroupProbeLimit);
  HJ_CUDA(ctx, cudaMemsetAsync(ctx->group_misc.p, 0, 16, st));
  const uint32_t grid = static_cast<uint32_t>(std::min<uint64_t>((entries + 255) / 256, uint64_t(ctx->n_sms) * 8));
  if (row_keys) k_group_init_rows<<<grid, 256, 0, st>>>(t); else k_group_init<<<grid, 256, 0, st>>>(t);
  HJ_LAUNCH_CHECK(ctx, "k_group_init");
  return HJ_OK;
}

int group_accumulate(hj_ctx* ctx, const GroupTable& t, const void* group_keys, const uint8_t* group_nulls, int group_key_type,
                     const void* values, const uint8_t* value_nulls, int value_type, uint64_t n_rows, uint64_t row_base, cudaStream_t st) {
  auto* overflow = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->group_misc.p) + 8);
  const uint32_t grid = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>((n_rows + kAggThreads - 1) / kAggThreads,
                                                                                      uint64_t(ctx->n_sms) * 4)));
#define HJ_GROUP_LAUNCH(GK, T)                                                                                                    \
  k_group_accumulate<GK, T><<<grid, kAggThreads, 0, st>>>(static_cast<const GK*>(group_keys), group_nulls, static_cast<const T*>(values), \
                                                          value_nulls, n_rows, row_base, t, overflow)
  const bool g64 = group_key_type == HJ_KEY_INT64;
  switch (value_type) {
    case HJ_KEY_INT32: if (g64) HJ_GROUP_LAUNCH(int64_t, int32_t); else HJ_GROUP_LAUNCH(int32_t, int32_t); break;
    case HJ_KEY_INT64: if (g64) HJ_GROUP_LAUNCH(int64_t, int64_t); else HJ_GROUP_LAUNCH(int32_t, int64_t); break;
    case HJ_KEY_FLOAT: if (g64) HJ_GROUP_LAUNCH(int64_t, float); else HJ_GROUP_LAUNCH(int32_t, float); break;
    default: if (g64) HJ_GROUP_LAUNCH(int64_t, double); else HJ_GROUP_LAUNCH(int32_t, double);
  }
#undef HJ_GROUP_LAUNCH
  HJ_LAUNCH_CHECK(ctx, "k_group_accumulate");
  return HJ_OK;
}

// returns kGroupTableFull (no error recorded) when an accumulate launch found the table full
int group_finish(hj_ctx* ctx, const GroupTable& t, uint64_t n_rows, bool floating, cudaStream_t st, hj_group_info* out_info) {
  auto* n_groups = static_cast<unsigned long long*>(ctx->group_misc.p);
  const uint32_t grid = static_cast<uint32_t>(std::min<uint64_t>((t.cap + 2 + 255) / 256, uint64_t(ctx->n_sms) * 8));
  k_group_emit<<<grid, 256, 0, st>>>(t, nullptr, n_groups);  // count
  HJ_LAUNCH_CHECK(ctx, "k_group_emit");
  auto* host = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedAggregate);
  HJ_CUDA(ctx, cudaMemcpyAsync(host, ctx->group_misc.p, 16, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (reinterpret_cast<const uint32_t*>(host)[2]) return kGroupTableFull;
  const uint64_t n = host[0];
  HJ_TRY(ensure(ctx, ctx->group_out, std::max<uint64_t>(1, n) * sizeof(GroupRecord)));
  HJ_CUDA(ctx, cudaMemsetAsync(n_groups, 0, 8, st));
  if (n) {
    k_group_emit<<<grid, 256, 0, st>>>(t, static_cast<GroupRecord*>(ctx->group_out.p), n_groups);
    HJ_LAUNCH_CHECK(ctx, "k_group_emit");
  }
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  ctx->group_n = n;
  ctx->group_rows = n_rows;
  ctx->group_floating = floating;
  ctx->group_valid = true;
  out_info->n_groups = n;
  out_info->n_rows = n_rows;
  return HJ_OK;
}

// table sizing + retry around `accumulate(const GroupTable&)`, which issues the accumulate launches of one attempt
template <typename Accumulate>
int group_run(hj_ctx* ctx, uint64_t n_rows, uint64_t max_groups, bool row_keys, bool floating, cudaStream_t st, hj_group_info* out_info,
              Accumulate&& accumulate) {
  const uint64_t full = 2 * (max_groups ? std::min(max_groups, n_rows) : n_rows);
  uint64_t want = max_groups ? full : std::min(full, kGroupFirstSlots);
  for (;;) {
    const bool final_attempt = want >= full;
    GroupTable t;
    HJ_TRY(group_begin(ctx, want, final_attempt, st, &t, row_keys));
    HJ_TRY(accumulate(t));
    const int rc = group_finish(ctx, t, n_rows, floating, st, out_info);
    if (rc != kGroupTableFull) return rc;
    if (final_attempt)
      return fail(ctx, HJ_ERR_CAPACITY, "more than %llu groups: the group table is full (max_groups was %llu)", (unsigned long long)t.cap,
                  (unsigned long long)max_groups);
    want = t.cap * 8;
  }
}

// A2 + read-back shared by hj_aggregate_column and hj_aggregate_relation: partials[1 .. grid] -> partials[0] -> *out
int finish_aggregate(hj_ctx* ctx, AggPartial* partials, uint32_t grid, uint64_t n_rows, int key_type, cudaStream_t st, hj_aggregate* out) {
  k_aggregate_final<<<1, 32, 0, st>>>(partials + 1, grid, partials);
  HJ_LAUNCH_CHECK(ctx, "k_aggregate_final");
  auto* host = reinterpret_cast<AggPartial*>(static_cast<char*>(ctx->pinned) + kPinnedAggregate);
  HJ_CUDA(ctx, cudaMemcpyAsync(host, partials, sizeof(AggPartial), cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  const bool floating = key_type >= HJ_KEY_FLOAT;
  out->n_rows = n_rows;
  out->n_values = host->n_values;
  out->is_floating = floating ? 1 : 0;
  out->sum_int = floating ? 0 : host->sum_i;
  // AVG accumulates in double for every column type (aggregate_traits.hpp:40-46); for an integer column the int64 sum is
  // that double sum exactly as long as it stays below 2^53
  out->sum_double = floating ? host->sum_d : static_cast<double>(host->sum_i);
  if (host->n_values) {
    out->min_int = floating ? 0 : host->min_i;
    out->max_int = floating ? 0 : host->max_i;
    out->min_double = static_cast<double>(host->min_i);
    out->max_double = static_cast<double>(host->max_i);
    if (floating) agg_min_max(*host, &out->min_double, &out->max_double);
  }
  return HJ_OK;
}

}  // namespace

// ============================================================================================================= C ABI
extern "C" {

int hj_abi_version(void) { return HJ_ABI_VERSION; }

size_t hj_abi_sizeof(int which) {
  switch (which) {
    case 0: return sizeof(hj_segment);
    case 1: return sizeof(hj_relation);
    case 2: return sizeof(hj_options);
    case 3: return sizeof(hj_result_info);
    case 4: return sizeof(hj_timings);
    case 5: return sizeof(hj_row_id);
    default: return 0;
  }
}

int hj_create(int device_id, const hj_options* opts, hj_ctx** out_ctx) {
  if (!out_ctx) return fail(nullptr, HJ_ERR_INVALID, "out_ctx is NULL");
  *out_ctx = nullptr;
  if (opts && opts->abi_version != HJ_ABI_VERSION)
    return fail(nullptr, HJ_ERR_INVALID, "hj_options.abi_version %u != %u", opts->abi_version, HJ_ABI_VERSION);
  if (opts && opts->radix_bits > kMaxRadixBits)
    return fail(nullptr, HJ_ERR_INVALID, "radix_bits %d > %d", opts->radix_bits, kMaxRadixBits);
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0) {
    (void)cudaGetLastError();
    return fail(nullptr, HJ_ERR_CUDA, "no CUDA device: %s (libhj has no CPU fallback)", cudaGetErrorString(e));
  }
  if (device_id < 0 || device_id >= n_dev) return fail(nullptr, HJ_ERR_INVALID, "device %d out of range [0,%d)", device_id, n_dev);
  auto* ctx = new hj_ctx();
  ctx->device = device_id;
  if (opts) ctx->opts = *opts; else { ctx->opts.abi_version = HJ_ABI_VERSION; ctx->opts.radix_bits = -1; }
  cudaDeviceProp prop{};
  bool ok = cudaSetDevice(device_id) == cudaSuccess && cudaGetDeviceProperties(&prop, device_id) == cudaSuccess;
  if (ok && prop.major < 10) {
    delete ctx;
    return fail(nullptr, HJ_ERR_CUDA, "device %d is sm_%d%d; libhj is built for sm_100a only", device_id, prop.major, prop.minor);
  }
  ok = ok && cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaHostAlloc(&ctx->pinned, kPinnedBytes, cudaHostAllocDefault) == cudaSuccess;
  for (auto& ev : ctx->ev) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking) == cudaSuccess;
  for (auto* arr : {ctx->ev_in, ctx->ev_done, ctx->ev_out, ctx->ev_x})
    for (int i = 0; i < 2; ++i) ok = ok && cudaEventCreateWithFlags(&arr[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    const char* msg = cudaGetErrorString(cudaGetLastError());
    hj_destroy(ctx);
    return fail(nullptr, HJ_ERR_CUDA, "context setup failed: %s", msg);
  }
  ctx->n_sms = prop.multiProcessorCount;
  *out_ctx = ctx;
  return HJ_OK;
}

int hj_release_scratch(hj_ctx* ctx) {
  if (check_ctx(ctx) != HJ_OK) return HJ_ERR_INVALID;
  for (DevBuf* b : ctx->all_bufs) {
    if (b->p) cudaFree(b->p);
    b->p = nullptr;
    b->cap = 0;
  }
  ctx->has_result = false;
  return HJ_OK;
}

void hj_comm_release(hj_ctx* ctx);  // hj_comm.cuh

void hj_destroy(hj_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  hj_comm_release(ctx);
  hj_release_scratch(ctx);
  hj_scan_release(ctx, -1);
  for (auto& ev : ctx->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto* arr : {ctx->ev_in, ctx->ev_done, ctx->ev_out, ctx->ev_x})
    for (int i = 0; i < 2; ++i)
      if (arr[i]) cudaEventDestroy(arr[i]);
  if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
  if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

const char* hj_last_error(const hj_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int hj_radix_bits(uint64_t n_build_rows, int key_type) { return radix_bits_for(n_build_rows, key_type); }

// ------------------------------------------------------------------------------------------------ TableScan (N4)
int hj_scan(hj_ctx* ctx, const hj_relation* column, int predicate, int64_t value, int value_is_null, int slot,
            uint32_t scan_flags, void* cuda_stream, hj_scan_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  if (slot < 0 || slot >= HJ_SCAN_SLOTS) return fail(ctx, HJ_ERR_INVALID, "scan slot %d outside 0..%d", slot, HJ_SCAN_SLOTS - 1);
  ctx->scan_valid[slot] = false;
  const bool null_test = predicate == HJ_PRED_IS_NULL || predicate == HJ_PRED_IS_NOT_NULL;
  if (!null_test && (predicate < HJ_PRED_EQUALS || predicate > HJ_PRED_GREATER_THAN_EQUALS))
    return fail(ctx, HJ_ERR_INVALID, "PredicateCondition %d is not a single-column comparison (Between/In/Like are refused, as "
                "SingleColumnTableScanImpl fails on them)", predicate);
  HJ_TRY(validate_relation(ctx, column, "scan"));
  bool wide = false, any_float = false, any_double = false;
  for (uint32_t i = 0; i < column->n_segs; ++i) {
    const hj_segment& sg = column->segs[i];
    any_float = any_float || sg.key_type == HJ_KEY_FLOAT;
    any_double = any_double || sg.key_type == HJ_KEY_DOUBLE;
    wide = wide || sg.key_type == HJ_KEY_INT64;
    if (sg.kind == HJ_SEG_REFERENCE && (scan_flags & HJ_SCAN_DEREFERENCE) && column->location != HJ_LOC_DEVICE)
      return fail(ctx, HJ_ERR_INVALID, "scan chunk %u: HJ_SCAN_DEREFERENCE needs a device-resident relation (the kernel reads the "
                  "chunk's PosList where it lies)", i);
  }
  // the literal in the column's type, as the reference's type_cast<ColumnDataType>(AllTypeVariant) converts it
  // (single_column_table_scan_impl.cpp:36, type_cast.hpp:39-58)
  const bool floating = any_float || any_double;
  double literal_d = 0.0;
  if (!null_test && !value_is_null) {
    if (scan_flags & HJ_SCAN_VALUE_IS_DOUBLE) {
      std::memcpy(&literal_d, &value, sizeof(double));
      if (!floating) {
        // integral column, floating literal: boost::numeric_cast<T>(double) truncates toward zero and throws outside T's range
        const double t = std::trunc(literal_d);
        const double lo = wide ? -9223372036854775808.0 : -2147483648.0, hi = wide ? 9223372036854775808.0 : 2147483648.0;
        if (!(t >= lo && t < hi))
          return fail(ctx, HJ_ERR_INVALID, "literal %g does not fit the %s column (boost::numeric_cast throws, type_cast.hpp:56)",
                      literal_d, wide ? "int64" : "int32");
        value = static_cast<int64_t>(t);
      }
    } else {
      literal_d = static_cast<double>(value);
    }
  }
  if (!floating && !null_test && !value_is_null && !wide && (value < INT32_MIN || value > INT32_MAX))
    return fail(ctx, HJ_ERR_INVALID, "literal %lld does not fit the int32 column (the reference's type_cast fails on it too)",
                (long long)value);
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  if (!null_test && value_is_null) {
    // comparing with NULL yields NULL: always the empty PosList (single_column_table_scan_impl.cpp:25-34)
    HJ_TRY(ensure(ctx, ctx->scan_chunk_begin[slot], (size_t(column->n_segs) + 1) * 4));
    HJ_TRY(ensure(ctx, ctx->scan_pos[slot], 8));
    HJ_CUDA(ctx, cudaMemsetAsync(ctx->scan_chunk_begin[slot].p, 0, (size_t(column->n_segs) + 1) * 4, st));
    HJ_CUDA(ctx, cudaStreamSynchronize(st));
    ctx->scan_rows[slot] = 0;
    ctx->scan_chunks[slot] = column->n_segs;
    ctx->scan_valid[slot] = true;
    out_info->n_chunks = column->n_segs;
    return HJ_OK;
  }
  // REFERENCE / FOR / RLE / SimdBp128 chunks become GATHERED / VALUE chunks on the device first (k_gather_refs)
  ctx->ref_h2d_bytes = 0;
  std::memset(static_cast<char*>(ctx->pinned) + kPinnedRefFlags, 0, 8);
  RefRewrite rw;
  // float / double columns: EVERY chunk is rewritten into the bit patterns of the column type (to_hashed_type; a column
  // with double chunks is compared in double), so the mask kernel sees value-like chunks only
  const uint32_t scan_type = any_double ? HJ_KEY_DOUBLE : any_float ? HJ_KEY_FLOAT : wide ? HJ_KEY_INT64 : HJ_KEY_INT32;
  HJ_TRY(gather_references(ctx, column, 1, st, &rw, scan_type));
  const hj_relation* rel = rw.any ? &rw.rel : column;
  switch (scan_type) {
    case HJ_KEY_DOUBLE: HJ_TRY((run_scan<int64_t, double>(ctx, rel, column, predicate, literal_d, 0, slot, scan_flags, st, out_info))); break;
    case HJ_KEY_FLOAT:
      HJ_TRY((run_scan<int32_t, float>(ctx, rel, column, predicate, static_cast<float>(literal_d), 0, slot, scan_flags, st, out_info)));
      break;
    case HJ_KEY_INT64: HJ_TRY((run_scan<int64_t, int64_t>(ctx, rel, column, predicate, value, value, slot, scan_flags, st, out_info))); break;
    default:
      HJ_TRY((run_scan<int32_t, int32_t>(ctx, rel, column, predicate, static_cast<int32_t>(value), value, slot, scan_flags, st, out_info)));
  }
  if (rw.any) {
    const auto* flag = reinterpret_cast<const uint32_t*>(static_cast<const char*>(ctx->pinned) + kPinnedRefFlags);
    if (flag[0] || flag[1]) {
      ctx->scan_valid[slot] = false;
      return fail(ctx, HJ_ERR_INVALID, "scan: a PosList entry points outside the referenced table");
    }
  }
  return HJ_OK;
}

int hj_scan_value_ids(hj_ctx* ctx, const hj_relation* column, int predicate, const uint32_t* bounds, int slot,
                      uint32_t scan_flags, void* cuda_stream, hj_scan_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  if (slot < 0 || slot >= HJ_SCAN_SLOTS) return fail(ctx, HJ_ERR_INVALID, "scan slot %d outside 0..%d", slot, HJ_SCAN_SLOTS - 1);
  ctx->scan_valid[slot] = false;
  const bool null_test = predicate == HJ_PRED_IS_NULL || predicate == HJ_PRED_IS_NOT_NULL;
  if (!null_test && (predicate < HJ_PRED_EQUALS || predicate > HJ_PRED_GREATER_THAN_EQUALS))
    return fail(ctx, HJ_ERR_INVALID, "PredicateCondition %d is not a single-column comparison", predicate);
  if (scan_flags) return fail(ctx, HJ_ERR_INVALID, "hj_scan_value_ids takes no scan flags (data chunks only, no literal)");
  if (!column || (column->n_segs && !column->segs)) return fail(ctx, HJ_ERR_INVALID, "scan relation is NULL or has no segment array");
  if (column->n_segs && !bounds && !null_test) return fail(ctx, HJ_ERR_INVALID, "bounds is NULL");
  // only the attribute vectors travel: the dictionaries (strings, or anything else sorted) stay with the caller
  std::vector<hj_segment> segs(column->segs, column->segs + column->n_segs);
  for (uint32_t i = 0; i < column->n_segs; ++i) {
    hj_segment& sg = segs[i];
    if (sg.kind != HJ_SEG_DICT || (sg.attr_width != 1 && sg.attr_width != 2 && sg.attr_width != 4))
      return fail(ctx, HJ_ERR_INVALID, "scan chunk %u: hj_scan_value_ids takes dictionary chunks with a FixedSizeByteAligned "
                  "attribute vector (kind %u, width %u)", i, sg.kind, sg.attr_width);
    sg.values = nullptr;
    sg.dict_n = 0;
    sg.key_type = HJ_KEY_INT32;
  }
  hj_relation rel = *column;
  rel.segs = segs.data();
  HJ_TRY(validate_relation(ctx, &rel, "scan"));  // after the rewrite: values / dict_n of the caller's segments are ignored
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  std::vector<uint32_t> zeros;
  if (!bounds) {
    zeros.assign(size_t(column->n_segs) * 2, 0u);
    bounds = zeros.data();
  }
  return run_scan<int32_t, int32_t>(ctx, &rel, &rel, predicate, 0, 0, slot, 0u, st, out_info, bounds);
}

// ------------------------------------------------------------------------------------------------ column values at a PosList
int hj_gather_column(hj_ctx* ctx, const hj_row_id* device_pos, uint64_t n_rows, const hj_relation* column, int slot,
                     void* cuda_stream, const void** out_values, const uint8_t** out_nulls, int* out_key_type) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_values || !out_nulls) return fail(ctx, HJ_ERR_INVALID, "out_values / out_nulls is NULL");
  *out_values = nullptr;
  *out_nulls = nullptr;
  if (slot < 0 || slot >= HJ_GATHER_SLOTS) return fail(ctx, HJ_ERR_INVALID, "gather slot %d outside 0..%d", slot, HJ_GATHER_SLOTS - 1);
  if (!column) return fail(ctx, HJ_ERR_INVALID, "column is NULL");
  if (column->location != HJ_LOC_DEVICE)
    return fail(ctx, HJ_ERR_INVALID, "hj_gather_column reads a device-resident column through a device-resident PosList");
  if (n_rows && !device_pos) return fail(ctx, HJ_ERR_INVALID, "device_pos is NULL");
  if (n_rows >= 0xFFFFFFFFull) return fail(ctx, HJ_ERR_TOO_LARGE, "PosList of %llu rows (limit 2^32-2)", (unsigned long long)n_rows);
  // the column's type: the widest of its chunks, like a join key column (JoinHashTraits order: double > float > long > int)
  uint32_t type = HJ_KEY_INT32;
  for (uint32_t i = 0; i < column->n_segs; ++i) {
    const uint8_t k = column->segs[i].key_type;
    if (k == HJ_KEY_DOUBLE) type = HJ_KEY_DOUBLE;
    else if (k == HJ_KEY_FLOAT && type != HJ_KEY_DOUBLE) type = HJ_KEY_FLOAT;
    else if (k == HJ_KEY_INT64 && type == HJ_KEY_INT32) type = HJ_KEY_INT64;
  }
  for (uint32_t i = 0; i < column->n_segs; ++i)
    if (column->segs[i].key_type != type)
      return fail(ctx, HJ_ERR_INVALID, "column chunk %u has key type %u, others %u: one column has one type", i, column->segs[i].key_type, type);
  // a one-chunk reference table over `column` whose PosList is device_pos: exactly what a join input looks like (N1)
  hj_segment ref{};
  ref.chunk_id = 0;
  ref.n_rows = static_cast<uint32_t>(n_rows);
  ref.kind = HJ_SEG_REFERENCE;
  ref.key_type = static_cast<uint8_t>(type);
  ref.values = column;
  ref.attr = device_pos;
  hj_relation rel{};
  rel.segs = &ref;
  rel.n_segs = 1;
  rel.location = HJ_LOC_DEVICE;
  rel.n_rows = n_rows;
  HJ_TRY(validate_relation(ctx, &rel, "gather"));
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  auto* flag = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->pinned) + kPinnedRefFlags) + 2 + slot;  // words 2..5 of 8
  *flag = 0;
  RefRewrite rw;
  HJ_TRY(gather_references(ctx, &rel, 2 + slot, st, &rw, type));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (*flag) return fail(ctx, HJ_ERR_INVALID, "gather: a PosList entry points outside the column");
  *out_values = rw.segs[0].values;
  *out_nulls = rw.segs[0].nulls;
  if (out_key_type) *out_key_type = static_cast<int>(type);
  return HJ_OK;
}

// ------------------------------------------------------------------------------------------------ ungrouped aggregates
int hj_aggregate_column(hj_ctx* ctx, const void* device_values, const uint8_t* device_nulls, uint64_t n_rows, int key_type,
                        void* cuda_stream, hj_aggregate* out) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out) return fail(ctx, HJ_ERR_INVALID, "out is NULL");
  std::memset(out, 0, sizeof(*out));
  if (key_type < HJ_KEY_INT32 || key_type > HJ_KEY_DOUBLE) return fail(ctx, HJ_ERR_INVALID, "key type %d", key_type);
  if (n_rows && !device_values) return fail(ctx, HJ_ERR_INVALID, "device_values is NULL");
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  const uint32_t grid = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>((n_rows + kAggThreads - 1) / kAggThreads,
                                                                                      uint64_t(ctx->n_sms) * 8)));
  HJ_TRY(ensure(ctx, ctx->agg_partials, (size_t(grid) + 1) * sizeof(AggPartial)));
  auto* partials = static_cast<AggPartial*>(ctx->agg_partials.p);
  switch (key_type) {
    case HJ_KEY_INT32:
      k_aggregate_partial<int32_t><<<grid, kAggThreads, 0, st>>>(static_cast<const int32_t*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    case HJ_KEY_INT64:
      k_aggregate_partial<int64_t><<<grid, kAggThreads, 0, st>>>(static_cast<const int64_t*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    case HJ_KEY_FLOAT:
      k_aggregate_partial<float><<<grid, kAggThreads, 0, st>>>(static_cast<const float*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    default:
      k_aggregate_partial<double><<<grid, kAggThreads, 0, st>>>(static_cast<const double*>(device_values), device_nulls, n_rows, partials + 1);
  }
  HJ_LAUNCH_CHECK(ctx, "k_aggregate_partial");
  return finish_aggregate(ctx, partials, grid, n_rows, key_type, st, out);
}

// ---- hj_aggregate_column without allocations (no reference counterpart) ----
}  // extern "C"

namespace {

// Scratch of hj_aggregate_column_static: [0] the folded result, [1..] one partial per CTA. File scope, so every GPU that
// loads libhj holds these 128 KiB until the process exits. In HOST code `g_static_agg_partials` is only a stand-in for the
// symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device (static_agg_scratch).
static __device__ __align__(256) AggPartial g_static_agg_partials[HJ_STATIC_AGG_MAX_PARTIALS + 1];

constexpr int kStaticMaxDevices = 64;
AggPartial* g_static_agg_ptr[kStaticMaxDevices] = {};
pthread_mutex_t g_static_agg_lock = PTHREAD_MUTEX_INITIALIZER;

// device address of the scratch array on ctx->device (which is current: check_ctx ran)
int static_agg_scratch(hj_ctx* ctx, AggPartial** out) {
  if (ctx->device < 0 || ctx->device >= kStaticMaxDevices) return fail(ctx, HJ_ERR_INVALID, "device %d: static scratch serves devices 0..%d", ctx->device, kStaticMaxDevices - 1);
  pthread_mutex_lock(&g_static_agg_lock);
  AggPartial* p = g_static_agg_ptr[ctx->device];
  cudaError_t e = cudaSuccess;
  if (!p) {
    void* addr = nullptr;
    e = cudaGetSymbolAddress(&addr, g_static_agg_partials);
    if (e == cudaSuccess) g_static_agg_ptr[ctx->device] = p = static_cast<AggPartial*>(addr);
  }
  pthread_mutex_unlock(&g_static_agg_lock);
  if (e != cudaSuccess) return fail(ctx, HJ_ERR_CUDA, "cudaGetSymbolAddress(g_static_agg_partials) failed: %s", cudaGetErrorString(e));
  *out = p;
  return HJ_OK;
}

// A2 of the static variant: the fold of k_aggregate_final (same order, same arithmetic) + finish_aggregate's conversion to
// the ABI record, written to the caller's device memory.
__global__ void __launch_bounds__(32) k_aggregate_final_out(const AggPartial* __restrict__ partials, uint32_t n_partials,
                                                            unsigned long long n_rows, int floating, hj_aggregate* __restrict__ out) {
  AggPartial a;
  agg_init(a);
  for (uint32_t p = threadIdx.x; p < n_partials; p += 32) agg_merge(a, partials[p]);
#pragma unroll
  for (int delta = 16; delta > 0; delta >>= 1) {
    const AggPartial b = agg_shfl_down(a, delta);
    agg_merge(a, b);
  }
  if (threadIdx.x == 0) {
    hj_aggregate r;
    r.n_rows = n_rows;
    r.n_values = a.n_values;
    r.is_floating = floating ? 1 : 0;
    r.reserved = 0;
    r.sum_int = floating ? 0 : a.sum_i;
    r.sum_double = floating ? a.sum_d : static_cast<double>(a.sum_i);
    r.min_int = r.max_int = 0;
    r.min_double = r.max_double = 0.0;
    if (a.n_values) {
      r.min_int = floating ? 0 : a.min_i;
      r.max_int = floating ? 0 : a.max_i;
      r.min_double = static_cast<double>(a.min_i);
      r.max_double = static_cast<double>(a.max_i);
      if (floating) agg_min_max(a, &r.min_double, &r.max_double);
    }
    *out = r;
  }
}

}  // namespace

extern "C" {

int hj_aggregate_column_static(hj_ctx* ctx, const void* device_values, const uint8_t* device_nulls, uint64_t n_rows, int key_type,
                               void* cuda_stream, hj_aggregate* device_out) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!device_out) return fail(ctx, HJ_ERR_INVALID, "device_out is NULL");
  if (key_type < HJ_KEY_INT32 || key_type > HJ_KEY_DOUBLE) return fail(ctx, HJ_ERR_INVALID, "key type %d", key_type);
  if (n_rows && !device_values) return fail(ctx, HJ_ERR_INVALID, "device_values is NULL");
  if (n_rows > HJ_STATIC_AGG_MAX_ROWS)
    return fail(ctx, HJ_ERR_TOO_LARGE, "hj_aggregate_column_static: %llu rows exceed HJ_STATIC_AGG_MAX_ROWS = %llu (use hj_aggregate_column)",
                (unsigned long long)n_rows, (unsigned long long)HJ_STATIC_AGG_MAX_ROWS);
  HJ_TRY(check_ctx(ctx));
  AggPartial* partials = nullptr;
  HJ_TRY(static_agg_scratch(ctx, &partials));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // the grid of hj_aggregate_column (so that double sums fold identically), capped by the scratch array
  const uint32_t grid = static_cast<uint32_t>(std::min<uint64_t>(
      HJ_STATIC_AGG_MAX_PARTIALS, std::max<uint64_t>(1, std::min<uint64_t>((n_rows + kAggThreads - 1) / kAggThreads, uint64_t(ctx->n_sms) * 8))));
  switch (key_type) {
    case HJ_KEY_INT32:
      k_aggregate_partial<int32_t><<<grid, kAggThreads, 0, st>>>(static_cast<const int32_t*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    case HJ_KEY_INT64:
      k_aggregate_partial<int64_t><<<grid, kAggThreads, 0, st>>>(static_cast<const int64_t*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    case HJ_KEY_FLOAT:
      k_aggregate_partial<float><<<grid, kAggThreads, 0, st>>>(static_cast<const float*>(device_values), device_nulls, n_rows, partials + 1);
      break;
    default:
      k_aggregate_partial<double><<<grid, kAggThreads, 0, st>>>(static_cast<const double*>(device_values), device_nulls, n_rows, partials + 1);
  }
  HJ_LAUNCH_CHECK(ctx, "k_aggregate_partial");
  k_aggregate_final_out<<<1, 32, 0, st>>>(partials + 1, grid, n_rows, key_type >= HJ_KEY_FLOAT ? 1 : 0, device_out);
  HJ_LAUNCH_CHECK(ctx, "k_aggregate_final_out");
  return HJ_OK;
}

int hj_aggregate_grouped(hj_ctx* ctx, const void* device_group_keys, const uint8_t* device_group_nulls, int group_key_type,
                         const void* device_values, const uint8_t* device_value_nulls, int value_type, uint64_t n_rows,
                         uint64_t max_groups, void* cuda_stream, hj_group_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  ctx->group_valid = false;
  if (group_key_type != HJ_KEY_INT32 && group_key_type != HJ_KEY_INT64)
    return fail(ctx, HJ_ERR_INVALID, "group-by column of key type %d: int32 / int64 only", group_key_type);
  if (device_values && (value_type < HJ_KEY_INT32 || value_type > HJ_KEY_DOUBLE)) return fail(ctx, HJ_ERR_INVALID, "value type %d", value_type);
  if (n_rows && !device_group_keys) return fail(ctx, HJ_ERR_INVALID, "device_group_keys is NULL");
  if (!device_values) value_type = HJ_KEY_INT32;  // COUNT(*) only
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return group_run(ctx, n_rows, max_groups, /*row_keys=*/false, device_values && value_type >= HJ_KEY_FLOAT, st, out_info,
                   [&](const GroupTable& t) -> int {
                     return group_accumulate(ctx, t, device_group_keys, device_group_nulls, group_key_type, device_values,
                                             device_value_nulls, value_type, n_rows, 0, st);
                   });
}

int hj_aggregate_grouped_rows(hj_ctx* ctx, const hj_group_column* group_columns, uint32_t n_group_columns, const void* device_values,
                              const uint8_t* device_value_nulls, int value_type, uint64_t n_rows, uint64_t max_groups, void* cuda_stream,
                              hj_group_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  ctx->group_valid = false;
  if (!group_columns || n_group_columns == 0 || n_group_columns > uint32_t(kMaxGroupColumns))
    return fail(ctx, HJ_ERR_INVALID, "%u group-by columns: 1..%d", n_group_columns, kMaxGroupColumns);
  GroupColumns g{};
  g.n = n_group_columns;
  for (uint32_t c = 0; c < n_group_columns; ++c) {
    const hj_group_column& col = group_columns[c];
    if (col.key_type < HJ_KEY_INT32 || col.key_type > HJ_KEY_DOUBLE) return fail(ctx, HJ_ERR_INVALID, "group-by column %u: key type %d", c, col.key_type);
    if (n_rows && !col.values) return fail(ctx, HJ_ERR_INVALID, "group-by column %u: values is NULL", c);
    g.values[c] = col.values;
    g.nulls[c] = col.nulls;
    g.width[c] = (col.key_type == HJ_KEY_INT64 || col.key_type == HJ_KEY_DOUBLE) ? 8u : 4u;
    g.floating[c] = col.key_type >= HJ_KEY_FLOAT ? 1u : 0u;
  }
  if (device_values && (value_type < HJ_KEY_INT32 || value_type > HJ_KEY_DOUBLE)) return fail(ctx, HJ_ERR_INVALID, "value type %d", value_type);
  if (!device_values) value_type = HJ_KEY_INT32;
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  const uint32_t grid = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>((n_rows + kAggThreads - 1) / kAggThreads,
                                                                                      uint64_t(ctx->n_sms) * 4)));
  return group_run(ctx, n_rows, max_groups, /*row_keys=*/true, device_values && value_type >= HJ_KEY_FLOAT, st, out_info,
                   [&](const GroupTable& t) -> int {
    auto* overflow = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->group_misc.p) + 8);
    switch (value_type) {
      case HJ_KEY_INT32:
        k_group_accumulate_rows<int32_t><<<grid, kAggThreads, 0, st>>>(g, static_cast<const int32_t*>(device_values), device_value_nulls, n_rows, t, overflow);
        break;
      case HJ_KEY_INT64:
        k_group_accumulate_rows<int64_t><<<grid, kAggThreads, 0, st>>>(g, static_cast<const int64_t*>(device_values), device_value_nulls, n_rows, t, overflow);
        break;
      case HJ_KEY_FLOAT:
        k_group_accumulate_rows<float><<<grid, kAggThreads, 0, st>>>(g, static_cast<const float*>(device_values), device_value_nulls, n_rows, t, overflow);
        break;
      default:
        k_group_accumulate_rows<double><<<grid, kAggThreads, 0, st>>>(g, static_cast<const double*>(device_values), device_value_nulls, n_rows, t, overflow);
    }
    HJ_LAUNCH_CHECK(ctx, "k_group_accumulate_rows");
    return int(HJ_OK);
  });
}

int hj_aggregate_grouped_rows_relation(hj_ctx* ctx, const hj_relation* const* group_columns, uint32_t n_group_columns,
                                       const hj_relation* value_column, uint64_t max_groups, void* cuda_stream, hj_group_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  ctx->group_valid = false;
  if (!group_columns || n_group_columns == 0 || n_group_columns > uint32_t(kMaxGroupColumns))
    return fail(ctx, HJ_ERR_INVALID, "%u group-by columns: 1..%d", n_group_columns, kMaxGroupColumns);
  std::vector<const hj_relation*> rels(group_columns, group_columns + n_group_columns);
  if (value_column) rels.push_back(value_column);
  for (size_t c = 0; c < rels.size(); ++c) {
    HJ_TRY(validate_relation(ctx, rels[c], c < n_group_columns ? "group-by" : "aggregate"));
    if (rels[c]->n_segs != rels[0]->n_segs || rels[c]->n_rows != rels[0]->n_rows)
      return fail(ctx, HJ_ERR_INVALID, "column %zu: the columns of one table have the same chunks", c);
    for (uint32_t i = 0; i < rels[0]->n_segs; ++i)
      if (rels[c]->segs[i].n_rows != rels[0]->segs[i].n_rows) return fail(ctx, HJ_ERR_INVALID, "column %zu: chunk %u has another row count", c, i);
  }
  const uint64_t n_rows = rels[0]->n_rows;
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // every column as ONE contiguous device array (values, then NULL bytes): K10 over every chunk into the buffers of gather
  // slot 0, then chunk by chunk device-to-device into the column's own buffer (representative rows index whole columns)
  auto* flags = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->pinned) + kPinnedRefFlags);
  std::vector<hj_group_column> cols(rels.size());
  uint32_t bad = 0;
  for (size_t c = 0; c < rels.size(); ++c) {
    uint32_t type = HJ_KEY_INT32;
    for (uint32_t i = 0; i < rels[c]->n_segs; ++i) type = std::max<uint32_t>(type, rels[c]->segs[i].key_type);
    for (uint32_t i = 0; i < rels[c]->n_segs; ++i)
      if (rels[c]->segs[i].key_type != type) return fail(ctx, HJ_ERR_INVALID, "column %zu chunk %u: a column has one type", c, i);
    const size_t width = (type == HJ_KEY_INT64 || type == HJ_KEY_DOUBLE) ? 8 : 4;
    const size_t nulls_off = (size_t(n_rows) * width + 255) & ~size_t(255);
    HJ_TRY(ensure(ctx, ctx->group_cols[c], nulls_off + std::max<uint64_t>(1, n_rows)));
    char* dst = static_cast<char*>(ctx->group_cols[c].p);
    flags[2] = 0;
    RefRewrite rw;
    HJ_TRY(gather_references(ctx, rels[c], 2, st, &rw, type, true));
    uint64_t row = 0;
    for (uint32_t i = 0; i < rels[c]->n_segs; ++i) {
      const hj_segment& sg = rw.segs[i];
      if (!sg.n_rows) continue;
      HJ_CUDA(ctx, cudaMemcpyAsync(dst + row * width, sg.values, size_t(sg.n_rows) * width, cudaMemcpyDeviceToDevice, st));
      HJ_CUDA(ctx, cudaMemcpyAsync(dst + nulls_off + row, sg.nulls, sg.n_rows, cudaMemcpyDeviceToDevice, st));
      row += sg.n_rows;
    }
    HJ_CUDA(ctx, cudaStreamSynchronize(st));  // the gather buffers are reused by the next column; the error flag has arrived
    bad |= flags[2];
    cols[c].values = dst;
    cols[c].nulls = reinterpret_cast<const uint8_t*>(dst + nulls_off);
    cols[c].key_type = static_cast<int32_t>(type);
    cols[c].reserved = 0;
  }
  if (bad) return fail(ctx, HJ_ERR_INVALID, "grouped aggregate: a PosList entry points outside the referenced table");
  const hj_group_column* value = value_column ? &cols[n_group_columns] : nullptr;
  return hj_aggregate_grouped_rows(ctx, cols.data(), n_group_columns, value ? value->values : nullptr, value ? value->nulls : nullptr,
                                   value ? value->key_type : HJ_KEY_INT32, n_rows, max_groups, cuda_stream, out_info);
}

int hj_aggregate_grouped_relation(hj_ctx* ctx, const hj_relation* group_column, const hj_relation* value_column, uint64_t max_groups,
                                  void* cuda_stream, hj_group_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  ctx->group_valid = false;
  HJ_TRY(validate_relation(ctx, group_column, "group-by"));
  if (value_column) {
    HJ_TRY(validate_relation(ctx, value_column, "aggregate"));
    if (value_column->n_segs != group_column->n_segs) return fail(ctx, HJ_ERR_INVALID, "group-by and aggregate column: different chunk counts");
    for (uint32_t i = 0; i < group_column->n_segs; ++i)
      if (value_column->segs[i].n_rows != group_column->segs[i].n_rows)
        return fail(ctx, HJ_ERR_INVALID, "group-by and aggregate column: chunk %u has %u vs %u rows", i, group_column->segs[i].n_rows,
                    value_column->segs[i].n_rows);
  }
  auto type_of = [](const hj_relation* r) {
    uint32_t t = HJ_KEY_INT32;
    for (uint32_t i = 0; i < r->n_segs; ++i) t = std::max<uint32_t>(t, r->segs[i].key_type);  // INT32 < INT64 < FLOAT < DOUBLE
    return t;
  };
  const uint32_t gtype = type_of(group_column), vtype = value_column ? type_of(value_column) : uint32_t(HJ_KEY_INT32);
  if (gtype != HJ_KEY_INT32 && gtype != HJ_KEY_INT64) return fail(ctx, HJ_ERR_INVALID, "group-by column of key type %u: int32 / int64 only", gtype);
  for (const hj_relation* r : {group_column, value_column})
    for (uint32_t i = 0; r && i < r->n_segs; ++i)
      if (r->segs[i].key_type != (r == group_column ? gtype : vtype)) return fail(ctx, HJ_ERR_INVALID, "chunk %u: a column has one type", i);
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // both columns as plain device arrays, chunk by chunk (K10 over every chunk; uses the buffers of gather slots 0 and 1)
  auto* flags = reinterpret_cast<uint32_t*>(static_cast<char*>(ctx->pinned) + kPinnedRefFlags);
  flags[2] = flags[3] = 0;
  RefRewrite rg, rv;
  HJ_TRY(gather_references(ctx, group_column, 2, st, &rg, gtype, true));
  if (value_column) HJ_TRY(gather_references(ctx, value_column, 3, st, &rv, vtype, true));
  HJ_TRY(group_run(ctx, group_column->n_rows, max_groups, /*row_keys=*/false, value_column && vtype >= HJ_KEY_FLOAT, st, out_info,
                   [&](const GroupTable& t) -> int {
    uint64_t row_base = 0;  // the chunks in order: the row order MIN / MAX's NaN rule refers to
    for (uint32_t i = 0; i < group_column->n_segs; ++i) {
      const hj_segment& g = rg.segs[i];
      if (!g.n_rows) continue;
      HJ_TRY(group_accumulate(ctx, t, g.values, g.nulls, static_cast<int>(gtype), value_column ? rv.segs[i].values : nullptr,
                              value_column ? rv.segs[i].nulls : nullptr, static_cast<int>(vtype), g.n_rows, row_base, st));
      row_base += g.n_rows;
    }
    return int(HJ_OK);
  }));
  if (flags[2] || flags[3]) {
    ctx->group_valid = false;
    return fail(ctx, HJ_ERR_INVALID, "grouped aggregate: a PosList entry points outside the referenced table");
  }
  return HJ_OK;
}

int hj_aggregate_grouped_fetch(hj_ctx* ctx, hj_group* groups, uint64_t capacity) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!ctx->group_valid) return fail(ctx, HJ_ERR_STATE, "no grouped aggregate result");
  const uint64_t n = ctx->group_n;
  if (capacity < n) return fail(ctx, HJ_ERR_CAPACITY, "capacity %llu < %llu groups", (unsigned long long)capacity, (unsigned long long)n);
  if (n && !groups) return fail(ctx, HJ_ERR_INVALID, "NULL output buffer");
  if (!n) return HJ_OK;
  HJ_TRY(check_ctx(ctx));
  std::vector<GroupRecord> rec(n);
  HJ_CUDA(ctx, cudaMemcpyAsync(rec.data(), ctx->group_out.p, n * sizeof(GroupRecord), cudaMemcpyDeviceToHost, ctx->own_stream));
  HJ_CUDA(ctx, cudaStreamSynchronize(ctx->own_stream));
  const bool floating = ctx->group_floating;
  for (uint64_t i = 0; i < n; ++i) {  // the device record in the ABI's layout (MIN / MAX of doubles back from their sortable images)
    const GroupRecord& r = rec[i];
    hj_group& g = groups[i];
    std::memset(&g, 0, sizeof(g));
    g.key = r.key;
    g.key_is_null = static_cast<int32_t>(r.key_is_null);
    g.agg.n_rows = r.rows;
    g.agg.n_values = r.values;
    g.agg.is_floating = floating ? 1 : 0;
    g.agg.sum_int = floating ? 0 : r.sum_i;
    g.agg.sum_double = floating ? r.sum_d : static_cast<double>(r.sum_i);
    if (r.values) {
      g.agg.min_int = floating ? 0 : r.mn;
      g.agg.max_int = floating ? 0 : r.mx;
      g.agg.min_double = floating ? unsortable(r.mn) : static_cast<double>(r.mn);
      g.agg.max_double = floating ? unsortable(r.mx) : static_cast<double>(r.mx);
    }
  }
  return HJ_OK;
}

int hj_aggregate_relation(hj_ctx* ctx, const hj_relation* column, void* cuda_stream, hj_aggregate* out) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!out) return fail(ctx, HJ_ERR_INVALID, "out is NULL");
  std::memset(out, 0, sizeof(*out));
  HJ_TRY(validate_relation(ctx, column, "aggregate"));
  bool wide = false, any_float = false, any_double = false;
  for (uint32_t i = 0; i < column->n_segs; ++i) {
    const uint8_t k = column->segs[i].key_type;
    wide = wide || k == HJ_KEY_INT64;
    any_float = any_float || k == HJ_KEY_FLOAT;
    any_double = any_double || k == HJ_KEY_DOUBLE;
  }
  const uint32_t type = any_double ? HJ_KEY_DOUBLE : any_float ? HJ_KEY_FLOAT : wide ? HJ_KEY_INT64 : HJ_KEY_INT32;
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // REFERENCE / FOR / RLE / SimdBp128 chunks (and every chunk of a float / double column) become value-like chunks on the
  // device first, exactly as for hj_scan; VALUE and DICT chunks of integer columns are read where they are
  ctx->ref_h2d_bytes = 0;
  std::memset(static_cast<char*>(ctx->pinned) + kPinnedRefFlags, 0, 8);
  RefRewrite rw;
  HJ_TRY(gather_references(ctx, column, 1, st, &rw, type));
  const hj_relation* rel = rw.any ? &rw.rel : column;
  PreparedRel pr;
  HJ_TRY(prepare_relation(ctx, rel, 1, kScanTile, st, &pr));
  const uint32_t grid = std::max<uint32_t>(1, std::min<uint32_t>(pr.meta.n_tiles, uint32_t(ctx->n_sms) * 8));
  HJ_TRY(ensure(ctx, ctx->agg_partials, (size_t(grid) + 1) * sizeof(AggPartial)));
  auto* partials = static_cast<AggPartial*>(ctx->agg_partials.p);
  switch (type) {
    case HJ_KEY_DOUBLE: k_aggregate_relation<int64_t, double, kScanTile><<<grid, kAggThreads, 0, st>>>(pr.meta, partials + 1); break;
    case HJ_KEY_FLOAT: k_aggregate_relation<int32_t, float, kScanTile><<<grid, kAggThreads, 0, st>>>(pr.meta, partials + 1); break;
    case HJ_KEY_INT64: k_aggregate_relation<int64_t, int64_t, kScanTile><<<grid, kAggThreads, 0, st>>>(pr.meta, partials + 1); break;
    default: k_aggregate_relation<int32_t, int32_t, kScanTile><<<grid, kAggThreads, 0, st>>>(pr.meta, partials + 1);
  }
  HJ_LAUNCH_CHECK(ctx, "k_aggregate_relation");
  HJ_TRY(finish_aggregate(ctx, partials, grid, column->n_rows, static_cast<int>(type), st, out));
  if (rw.any) {
    const auto* flag = reinterpret_cast<const uint32_t*>(static_cast<const char*>(ctx->pinned) + kPinnedRefFlags);
    if (flag[0] || flag[1]) {
      std::memset(out, 0, sizeof(*out));
      return fail(ctx, HJ_ERR_INVALID, "aggregate: a PosList entry points outside the referenced table");
    }
  }
  return HJ_OK;
}

int hj_scan_result_device(hj_ctx* ctx, int slot, const hj_row_id** pos, const uint32_t** chunk_begin, uint64_t* n_rows,
                          uint32_t* n_chunks) {
  if (!ctx) return HJ_ERR_INVALID;
  if (slot < 0 || slot >= HJ_SCAN_SLOTS || !ctx->scan_valid[slot]) return fail(ctx, HJ_ERR_STATE, "no scan result in slot %d", slot);
  if (pos) *pos = static_cast<const hj_row_id*>(ctx->scan_pos[slot].p);
  if (chunk_begin) *chunk_begin = static_cast<const uint32_t*>(ctx->scan_chunk_begin[slot].p);
  if (n_rows) *n_rows = ctx->scan_rows[slot];
  if (n_chunks) *n_chunks = ctx->scan_chunks[slot];
  return HJ_OK;
}

int hj_scan_fetch(hj_ctx* ctx, int slot, hj_row_id* pos, uint64_t capacity_rows, uint32_t* chunk_begin) {
  if (!ctx) return HJ_ERR_INVALID;
  if (slot < 0 || slot >= HJ_SCAN_SLOTS || !ctx->scan_valid[slot]) return fail(ctx, HJ_ERR_STATE, "no scan result in slot %d", slot);
  const uint64_t n = ctx->scan_rows[slot];
  // pos == NULL: only the chunk boundaries are wanted (the PosLists stay on the device for the next operator)
  if (pos && capacity_rows < n) return fail(ctx, HJ_ERR_CAPACITY, "capacity %llu < %llu scan result rows", (unsigned long long)capacity_rows,
                                            (unsigned long long)n);
  if (!pos && !chunk_begin) return fail(ctx, HJ_ERR_INVALID, "NULL output buffers");
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = ctx->own_stream;
  if (n && pos) HJ_CUDA(ctx, cudaMemcpyAsync(pos, ctx->scan_pos[slot].p, n * 8, cudaMemcpyDeviceToHost, st));
  if (chunk_begin)
    HJ_CUDA(ctx, cudaMemcpyAsync(chunk_begin, ctx->scan_chunk_begin[slot].p, (size_t(ctx->scan_chunks[slot]) + 1) * 4,
                                 cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  return HJ_OK;
}

int hj_scan_release(hj_ctx* ctx, int slot) {
  if (!ctx) return HJ_ERR_INVALID;
  if (slot >= HJ_SCAN_SLOTS) return fail(ctx, HJ_ERR_INVALID, "scan slot %d outside 0..%d", slot, HJ_SCAN_SLOTS - 1);
  if (check_ctx(ctx) != HJ_OK) return HJ_ERR_INVALID;
  for (int i = 0; i < HJ_SCAN_SLOTS; ++i) {
    if (slot >= 0 && i != slot) continue;
    for (DevBuf* b : {&ctx->scan_pos[i], &ctx->scan_chunk_begin[i]}) {
      if (b->p) cudaFree(b->p);
      b->p = nullptr;
      b->cap = 0;
    }
    ctx->scan_valid[i] = false;
    ctx->scan_rows[i] = 0;
  }
  return HJ_OK;
}

int join_dispatch(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                  hj_result_info* out_info, bool wide, bool bound);

int hj_join(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, void* cuda_stream,
            hj_result_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  ctx->has_result = false;
  // a binding (hj_bind_output) holds for exactly one join, whether that join succeeds or not: it is consumed before anything
  // can fail, so that a later unbound join never streams into buffers the caller may have freed
  const bool bound = ctx->bound;
  ctx->bound = false;
  ctx->result_in_bound = false;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  if (!(mode == HJ_MODE_INNER || mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT || mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI))
    return fail(ctx, HJ_ERR_INVALID, "JoinMode %d is not supported by JoinHash (Inner/Left/Right/Semi/Anti only)", mode);
  HJ_TRY(validate_relation(ctx, build, "build"));
  HJ_TRY(validate_relation(ctx, probe, "probe"));
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // JoinHashTraits: int32 x int64 hashes as int64, anything with a float/double column as that floating type
  const uint32_t hash_type = hashed_type_of(build, probe);
  const bool wide = hash_type == HJ_KEY_INT64 || hash_type == HJ_KEY_DOUBLE;  // 8-byte keys in the pipeline
  RefRewrite rw[2];  // REFERENCE segments -> gathered on the device (N1)
  HJ_TRY(gather_both(ctx, build, probe, st, rw, hash_type));
  const int status = join_dispatch(ctx, build, probe, mode, st, out_info, wide, bound);
  if (status != HJ_OK) return status;
  return check_reference_flags(ctx, rw);
}

// hj_join after argument checks and the REFERENCE gather
int join_dispatch(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                  hj_result_info* out_info, bool wide, bool bound) {
  if (!bound) return wide ? run_join<int64_t>(ctx, build, probe, mode, st, out_info)
                          : run_join<int32_t>(ctx, build, probe, mode, st, out_info);
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  if (!ctx->bound_probe || (!one_side && !ctx->bound_build)) return fail(ctx, HJ_ERR_INVALID, "hj_bind_output: NULL output buffer");
  // enough probe chunks to overlap the copies with the join? (slices of >= ~32 M rows, at most 24)
  const uint64_t min_slice = ctx->opts.stream_slice_rows ? ctx->opts.stream_slice_rows : kMinSliceRows;
  uint32_t n_slices = uint32_t(std::min<uint64_t>(24, probe->n_rows / min_slice));
  n_slices = std::min(n_slices, probe->n_segs);
  if (probe->location == HJ_LOC_HOST && n_slices >= 2)
    return wide ? run_join_streamed<int64_t>(ctx, build, probe, mode, st, out_info, n_slices)
                : run_join_streamed<int32_t>(ctx, build, probe, mode, st, out_info, n_slices);
  // small or device-resident probe side: plain join, then the download into the bound buffers
  HJ_TRY(wide ? run_join<int64_t>(ctx, build, probe, mode, st, out_info) : run_join<int32_t>(ctx, build, probe, mode, st, out_info));
  if (out_info->n_rows > ctx->bound_capacity)
    return fail(ctx, HJ_ERR_CAPACITY, "bound output holds %llu rows, the result has %llu (still on the device: hj_fetch it)",
                (unsigned long long)ctx->bound_capacity, (unsigned long long)out_info->n_rows);
  HJ_TRY(hj_fetch(ctx, one_side ? nullptr : ctx->bound_build, ctx->bound_probe, ctx->bound_capacity));
  ctx->result_in_bound = true;
  return HJ_OK;
}

int hj_bind_output(hj_ctx* ctx, hj_row_id* build_side, hj_row_id* probe_side, uint64_t capacity_rows) {
  if (!ctx) return HJ_ERR_INVALID;
  ctx->bound_build = build_side;
  ctx->bound_probe = probe_side;
  ctx->bound_capacity = capacity_rows;
  ctx->bound = probe_side != nullptr;
  return HJ_OK;
}

int hj_fetch(hj_ctx* ctx, hj_row_id* build_side, hj_row_id* probe_side, uint64_t capacity) {
  if (!ctx) return HJ_ERR_INVALID;
  if (ctx->result_in_bound && probe_side == ctx->bound_probe && (ctx->n_sides == 1 || build_side == ctx->bound_build))
    return HJ_OK;  // the join already delivered into these buffers (hj_bind_output)
  if (!ctx->has_result) return fail(ctx, HJ_ERR_STATE, "hj_fetch without a successful hj_join");
  if (capacity < ctx->n_result) return fail(ctx, HJ_ERR_CAPACITY, "capacity %llu < %llu result rows", (unsigned long long)capacity,
                                            (unsigned long long)ctx->n_result);
  if (ctx->n_sides == 1 && build_side) return fail(ctx, HJ_ERR_INVALID, "Semi/Anti joins have no build-side PosList");
  if (ctx->n_result && (!probe_side || (ctx->n_sides == 2 && !build_side))) return fail(ctx, HJ_ERR_INVALID, "NULL output buffer");
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = ctx->own_stream;
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[7], st));
  const size_t bytes = ctx->n_result * 8;
  if (bytes) {
    if (ctx->n_sides == 2) HJ_CUDA(ctx, cudaMemcpyAsync(build_side, ctx->res_build, bytes, cudaMemcpyDeviceToHost, st));
    HJ_CUDA(ctx, cudaMemcpyAsync(probe_side, ctx->res_probe, bytes, cudaMemcpyDeviceToHost, st));
  }
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[8], st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  ctx->timings.ms[HJ_PH_D2H] = ev_ms(ctx->ev[7], ctx->ev[8]);
  ctx->timings.bytes_d2h = bytes * ctx->n_sides;
  return HJ_OK;
}

int hj_result_device(hj_ctx* ctx, const hj_row_id** build_side, const hj_row_id** probe_side, uint64_t* n_rows) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!ctx->has_result) return fail(ctx, HJ_ERR_STATE, "hj_result_device without a successful hj_join");
  if (build_side) *build_side = ctx->n_sides == 2 ? static_cast<const hj_row_id*>(ctx->res_build) : nullptr;
  if (probe_side) *probe_side = static_cast<const hj_row_id*>(ctx->res_probe);
  if (n_rows) *n_rows = ctx->n_result;
  return HJ_OK;
}

int hj_translate(hj_ctx* ctx, hj_row_id* inout, uint64_t n, const hj_row_id* const* pos_lists,
                 const uint32_t* pos_list_sizes, uint32_t n_chunks) {
  if (!ctx) return HJ_ERR_INVALID;
  if (n == 0) return HJ_OK;
  if (!inout || (n_chunks && (!pos_lists || !pos_list_sizes))) return fail(ctx, HJ_ERR_INVALID, "NULL argument");
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = ctx->own_stream;
  std::vector<uint64_t> base(n_chunks + 1, 0);
  for (uint32_t c = 0; c < n_chunks; ++c) base[c + 1] = base[c] + pos_list_sizes[c];
  const size_t total = base[n_chunks];
  // misc layout: [inout n*8][pos lists total*8][chunk_base (n_chunks+1)*8][chunk_size n_chunks*4][error 4]
  const size_t o_pl = n * 8, o_base = o_pl + total * 8, o_size = o_base + (size_t(n_chunks) + 1) * 8,
               o_err = o_size + size_t(n_chunks) * 4;
  HJ_TRY(ensure(ctx, ctx->misc, o_err + 8));
  char* d = static_cast<char*>(ctx->misc.p);
  HJ_CUDA(ctx, cudaMemcpyAsync(d, inout, n * 8, cudaMemcpyHostToDevice, st));
  for (uint32_t c = 0; c < n_chunks; ++c)
    if (pos_list_sizes[c])
      HJ_CUDA(ctx, cudaMemcpyAsync(d + o_pl + base[c] * 8, pos_lists[c], size_t(pos_list_sizes[c]) * 8, cudaMemcpyHostToDevice, st));
  HJ_CUDA(ctx, cudaMemcpyAsync(d + o_base, base.data(), (size_t(n_chunks) + 1) * 8, cudaMemcpyHostToDevice, st));
  if (n_chunks) HJ_CUDA(ctx, cudaMemcpyAsync(d + o_size, pos_list_sizes, size_t(n_chunks) * 4, cudaMemcpyHostToDevice, st));
  HJ_CUDA(ctx, cudaMemsetAsync(d + o_err, 0, 4, st));
  const uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, uint64_t(ctx->n_sms) * 8);
  k_translate<<<grid, 256, 0, st>>>(reinterpret_cast<uint64_t*>(d), n, reinterpret_cast<const uint64_t*>(d + o_pl),
                                    reinterpret_cast<const uint64_t*>(d + o_base),
                                    reinterpret_cast<const uint32_t*>(d + o_size), n_chunks,
                                    reinterpret_cast<uint32_t*>(d + o_err));
  HJ_LAUNCH_CHECK(ctx, "k_translate");
  uint32_t err = 0;
  HJ_CUDA(ctx, cudaMemcpyAsync(inout, d, n * 8, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaMemcpyAsync(&err, d + o_err, 4, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (err) return fail(ctx, HJ_ERR_INVALID, "hj_translate: a RowID points outside its input PosList");
  return HJ_OK;
}

int hj_translate_device(hj_ctx* ctx, hj_row_id* device_inout, uint64_t n, const hj_relation* input, void* cuda_stream) {
  if (!ctx) return HJ_ERR_INVALID;
  if (!input) return fail(ctx, HJ_ERR_INVALID, "input relation is NULL");
  if (input->location != HJ_LOC_DEVICE)
    return fail(ctx, HJ_ERR_INVALID, "hj_translate_device reads the PosLists of a device-resident relation where they lie");
  HJ_TRY(validate_relation(ctx, input, "translate"));
  if (n == 0) return HJ_OK;
  if (!device_inout) return fail(ctx, HJ_ERR_INVALID, "device_inout is NULL");
  uint32_t n_ids = 0;
  for (uint32_t i = 0; i < input->n_segs; ++i) {
    if (input->segs[i].chunk_id >= (1u << 24))
      return fail(ctx, HJ_ERR_INVALID, "translate: chunk id %u (the table of PosLists is indexed by chunk id, limit 2^24)", input->segs[i].chunk_id);
    n_ids = std::max(n_ids, input->segs[i].chunk_id + 1);
  }
  std::vector<const uint64_t*> seg_pos(std::max<uint32_t>(1, n_ids), nullptr);
  std::vector<uint32_t> seg_rows(std::max<uint32_t>(1, n_ids), 0xFFFFFFFFu);
  for (uint32_t i = 0; i < input->n_segs; ++i) {
    const hj_segment& sg = input->segs[i];
    seg_rows[sg.chunk_id] = sg.n_rows;
    seg_pos[sg.chunk_id] = sg.kind == HJ_SEG_REFERENCE ? static_cast<const uint64_t*>(sg.attr) : nullptr;
  }
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  // misc layout: [PosList pointers n_ids*8][segment lengths n_ids*4, padded][error 4]
  const size_t o_rows = size_t(seg_pos.size()) * 8, o_err = (o_rows + size_t(seg_rows.size()) * 4 + 7) & ~size_t(7);
  HJ_TRY(ensure(ctx, ctx->misc, o_err + 8));
  char* d = static_cast<char*>(ctx->misc.p);
  HJ_CUDA(ctx, cudaMemcpyAsync(d, seg_pos.data(), seg_pos.size() * 8, cudaMemcpyHostToDevice, st));
  HJ_CUDA(ctx, cudaMemcpyAsync(d + o_rows, seg_rows.data(), seg_rows.size() * 4, cudaMemcpyHostToDevice, st));
  HJ_CUDA(ctx, cudaMemsetAsync(d + o_err, 0, 4, st));
  const uint32_t grid = (uint32_t)std::min<uint64_t>((n + 255) / 256, uint64_t(ctx->n_sms) * 8);
  k_translate_segments<<<grid, 256, 0, st>>>(reinterpret_cast<uint64_t*>(device_inout), n, reinterpret_cast<const uint64_t* const*>(d),
                                             reinterpret_cast<const uint32_t*>(d + o_rows), n_ids, reinterpret_cast<uint32_t*>(d + o_err));
  HJ_LAUNCH_CHECK(ctx, "k_translate_segments");
  uint32_t err = 0;
  HJ_CUDA(ctx, cudaMemcpyAsync(&err, d + o_err, 4, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));  // also: the pageable vectors above have been consumed
  if (err) return fail(ctx, HJ_ERR_INVALID, "hj_translate_device: a RowID points outside its input chunk");
  return HJ_OK;
}

int hj_get_timings(const hj_ctx* ctx, hj_timings* out) {
  if (!ctx || !out) return HJ_ERR_INVALID;
  *out = ctx->timings;
  return HJ_OK;
}

}  // extern "C"

// ============================================================================================================= multi-GPU
#include "hj_comm.cuh"  // includes hj_xpush.cuh
