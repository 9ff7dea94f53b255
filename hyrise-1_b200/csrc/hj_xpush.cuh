// hj_xpush.cuh — the fast multi-GPU path of hj_join_sharded ("push exchange"; no reference counterpart: the
// reference is single-process). Included by hj_api.cu behind hj_comm.cuh, whose IPC plumbing it uses.
//
// What round 1 did per rank (hj_comm.cuh): X1 count sweep (decode, decoded keys through HBM) -> host sync -> X3 scatter of
// (key, 8-byte RowID) into the peers -> local join over 12/16-byte RowID-carrying tuples on the two-sweep kernels: 62 ms per
// C2-sized shard on 8 GPUs against 19.6 ms on one. Here:
//   XP1 k_xchg_push   ONE sweep over the shard: decode (dictionary gather once), hash, destination rank, and the tile's rows
//                     grouped by destination are stored STRAIGHT into the destination GPUs' receive regions over NVLink as
//                     finished 8-byte tuples (int32 keys: (hash, position in the shard) — fmix32 is a bijection, so equal
//                     hashes = equal keys and keys are never output) or 16-byte tuples (int64 keys: (key, position, 0)).
//                     No count pass: a receive region is a row of 2048-tuple chunks that the sender's virtual CTAs take one
//                     at a time from a sender-local cursor (one global atomic per 16 KB sent); only a virtual CTA's last
//                     chunk per destination stays partly filled. The sender writes each chunk's fill count next to it.
//   XP2 k_xchg_publish  chunks used per destination -> the destinations' region headers; overflow flag.
//   barrier             1-word NCCL all-gather ("all stores have landed" + everybody's overflow flag).
//   local join          the unchanged single-GPU kernels on the fused path: k_part1p<XCH> reads the regions (a chunk = a
//                       tile, cp.async double-buffered: no gather here, the tuples stream), writes the source rank into the
//                       tuple (int64: the spare word; int32: over the level-1 digit bits of the hash, which the fragment it
//                       lands in makes redundant), k_scatter2f<PRE>, k_build_probe<kRidXch> emits RowIDs from (rank, position)
//                       through the all-gathered shard geometry (XchMap).
// Preconditions (else hj_join_sharded takes the general path of hj_comm.cuh): every rank maps every peer (CUDA IPC), two-pass
// radix partitioning (> 8 bits), no routed NULLs (HJ_FLAG_SQL_NULL_PROBE), every shard "regular" (chunks of 2^k rows with
// consecutive chunk ids, no reported row offsets). A destination region that would overflow (heavy skew) raises a flag that
// all ranks see at the barrier; they then repeat the join on the general path.
#pragma once

#ifndef HJ_TUNE_XP_BULK
#define HJ_TUNE_XP_BULK 1  // 1: k_xchg_push sends its runs with TMA bulk copies; 0: per-thread 8/16-byte stores
#endif

namespace hj {

constexpr uint32_t kXpChunk = kXpTile;  // tuples per receive chunk = one pass-1 tile of the local join (hj_api.cu)

struct XpushArgs {
  RelMeta rel;
  void* dst_tuples[kMaxRanks];     // destination d's receive region for MY rows (peer-mapped; own rank: local)
  uint32_t* dst_fill[kMaxRanks];   // fill count per chunk of that region
  uint32_t* cursor;                // [kMaxRanks] chunks taken per destination (sender-local)
  uint32_t* overflow;              // raised when a region is full
  uint32_t cap_chunks;             // chunks per region
  uint32_t pos_base;               // position of the relation's first row in the rank's shard (slices of a streamed join)
  uint32_t n_ranks;
  int null_policy;                 // kNullDrop or kNullAsValue
};

// One physical CTA = VC virtual CTAs of VT threads (named barriers), each streaming its own tiles: c, c + G, ...
// dynamic smem: VC * (TILE + kMaxRanks) * sizeof(Tuple<K>).
template <typename K, int NT, int VC, int TILE>
__global__ void __launch_bounds__(NT) k_xchg_push(const XpushArgs a) {
  using T = Tuple<K, false>;
  constexpr int VT = NT / VC;
  constexpr int ITEMS = TILE / VT;
  static_assert(TILE == kXpChunk && ITEMS >= 1 && ITEMS <= 32 && VT >= 32 && kMaxRanks * 2 <= 32, "a tile's run to one destination fits one chunk");
  extern __shared__ __align__(16) unsigned char s_raw[];
  T* stage = reinterpret_cast<T*>(s_raw) + static_cast<size_t>(threadIdx.x / VT) * (TILE + kMaxRanks);  // + one hole per run
  __shared__ uint32_t s_cnt[VC][32], s_loff[VC][32];
  __shared__ uint32_t s_run[VC][kMaxRanks], s_n[VC][kMaxRanks];                              // this tile's run per destination
  __shared__ uint32_t s_pos0[VC][kMaxRanks], s_first[VC][kMaxRanks], s_pos1[VC][kMaxRanks];  // ... and where it goes
  __shared__ uint32_t s_open[VC][kMaxRanks], s_left[VC][kMaxRanks];                          // open chunk (index, free tuples)
  const uint32_t v = threadIdx.x / VT, vt = threadIdx.x % VT;
  const uint32_t G = gridDim.x * VC, c = blockIdx.x * VC + v;
  const uint32_t R = a.n_ranks;
  const uint32_t nb = R > 1 ? 32u - __clz(R - 1u) : 0u;  // bits of a destination number
  const uint32_t lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  if (vt < kMaxRanks) {
    s_open[v][vt] = kInvalid;
    s_left[v][vt] = 0;
  }
  for (uint32_t tile = c; tile < a.rel.n_tiles; tile += G) {
    const DevSegment s = a.rel.segs[__ldg(a.rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t cnt = min(static_cast<uint32_t>(TILE), s.n_rows - r0);
    if (vt < 32) s_cnt[v][vt] = 0;
    K key[ITEMS];
    uint32_t vbits;
    load_tile_keys<K, ITEMS, VT>(s, r0, cnt, vt, key, vbits);
    // Destination + rank inside the tile's run for that destination, WITHOUT a shared-memory atomic per row: the lanes of a
    // warp that go to the same destination find each other with one ballot per destination bit (<= 4), the first of them
    // reserves the group's ranks with one atomic, the others derive theirs from their place in the group. Consecutive lanes
    // of a group get consecutive ranks, so the staging stores below stay bank-conflict free.
    // Measured alternatives for 1 B rows through ONE rank (tests/profile_join.py --sharded): a ballot per DESTINATION (loop
    // over R) 7.1 ms but 3 376 SASS instructions and ~80 more executed per row at R = 8; one atomic per row on R x 4
    // sub-counters 8.4 ms, on 32 sub-counters 10.6 ms (the sub-runs make the staging stores 4- / 32-way bank conflicts).
    vbar(v, VT);  // s_cnt zeroed; the previous tile's copy-out is done with `stage`
    uint32_t h[ITEMS], packed[ITEMS];  // packed: destination << 16 | rank
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const bool kept = u * VT + vt < cnt && (((vbits >> u) & 1u) || a.null_policy == kNullAsValue);
      h[u] = hash_key(key[u]);
      const uint32_t d = kept ? dest_rank_of_hash(h[u], R) : 0u;
      uint32_t peers = __ballot_sync(0xffffffffu, kept);
#pragma unroll
      for (uint32_t bit = 0; bit < 4; ++bit) {
        if (bit < nb) {  // uniform
          const uint32_t m = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
          peers &= ((d >> bit) & 1u) ? m : ~m;
        }
      }
      packed[u] = kNoDigit;
      const int leader = __ffs(peers) - 1;  // peers is only meaningful in kept lanes; it contains the lane itself there
      uint32_t base = 0;
      if (kept && static_cast<int>(lane) == leader) base = atomicAdd(&s_cnt[v][d], __popc(peers));
      base = __shfl_sync(0xffffffffu, base, kept ? leader : 0);
      if (kept) packed[u] = (d << 16) | (base + __popc(peers & lt));
    }
    vbar(v, VT);
    if (vt < 32) {
#if HJ_TUNE_XP_BULK
      // the previous tile's bulk copies (issued by threads of this warp) must have read the staging area before it is written
      // again; until here they had the next tile's loads, gathers and ranking to drain into NVLink
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
#endif
      // the virtual CTA's first warp, lane q = destination q: prefix sum -> where the tile's run for q is staged; then q's
      // run is placed: the rest of the open chunk, then a fresh one
      const uint32_t q = vt;
      const uint32_t real = q < R ? s_cnt[v][q] : 0u;
      const uint32_t n = slab_pad<T>(real);  // runs of 8-byte tuples are padded to whole 16-byte units: bulk copies below
      const uint32_t ex = warp_incl_scan(n) - n;
      if (q < R) {
        s_loff[v][q] = ex;
        s_run[v][q] = ex;
        s_n[v][q] = n;
        if (n != real) stage[ex + real] = hole_tuple<K>();  // skipped by the receiver's pass 1 (position = kHoleIdx)
        const uint32_t left = s_left[v][q], open = s_open[v][q];
        s_pos0[v][q] = open * kXpChunk + (kXpChunk - left);  // unused when left == 0
        s_first[v][q] = left;
        if (n > left) {
          if (open != kInvalid) a.dst_fill[q][open] = kXpChunk;  // the open chunk ends up full
          uint32_t fresh = atomicAdd(a.cursor + q, 1u);
          if (fresh >= a.cap_chunks) {  // the region is full (skew): flag it; the rows fall on the region's last chunk and are
            *a.overflow = 1u;           // never read, because every rank repeats the join on the general path
            fresh = a.cap_chunks - 1u;
          }
          s_pos1[v][q] = fresh * kXpChunk;
          s_open[v][q] = fresh;
          s_left[v][q] = kXpChunk - (n - left);
        } else {
          s_left[v][q] = left - n;
        }
      }
    }
    vbar(v, VT);
    const uint32_t pos0 = a.pos_base + s.row_begin + r0;
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      if (packed[u] != kNoDigit) {
        T t;
        if constexpr (sizeof(K) == 4) {
          t.key = static_cast<K>(h[u]);
        } else {
          t.key = key[u];
          t.pad = 0;
        }
        t.idx = pos0 + u * VT + vt;
        stage[s_loff[v][packed[u] >> 16] + (packed[u] & 0xFFFFu)] = t;
      }
    }
#if HJ_TUNE_XP_BULK
    // every run leaves with at most two bulk copies of the TMA engine (the part that fills the open chunk, the part that opens
    // a fresh one) straight into the destination GPU's memory: thread 2q + part issues it. Sources, destinations and sizes
    // are multiples of 16 bytes by construction (padded runs, 2048-tuple chunks).
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the staged tuples, for the TMA engine's reads
    vbar(v, VT);
    if (vt < 2 * R) {
      const uint32_t q = vt >> 1, part = vt & 1u;
      const uint32_t n = s_n[v][q], first = min(s_first[v][q], n);
      const uint32_t len = part ? n - first : first;
      if (len) {
        T* dst = static_cast<T*>(a.dst_tuples[q]) + (part ? s_pos1[v][q] : s_pos0[v][q]);
        const T* src = stage + s_run[v][q] + (part ? first : 0u);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_addr_of(src)),
                     "r"(len * static_cast<uint32_t>(sizeof(T))) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");  // waited for where the next tile needs the staging area
      }
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  vbar(v, VT);
#else
    vbar(v, VT);
    for (uint32_t q = 0; q < R; ++q) {
      const uint32_t n = s_n[v][q];
      if (n == 0) continue;  // uniform over the virtual CTA
      T* dst = static_cast<T*>(a.dst_tuples[q]);
      const T* src = stage + s_run[v][q];
      const uint32_t first = s_first[v][q], p0 = s_pos0[v][q], p1 = s_pos1[v][q];
      for (uint32_t i = vt; i < n; i += VT) dst[i < first ? p0 + i : p1 + (i - first)] = src[i];
    }
  }
  vbar(v, VT);
#endif
  if (vt < R && s_open[v][vt] != kInvalid) a.dst_fill[vt][s_open[v][vt]] = kXpChunk - s_left[v][vt];
  __threadfence_system();  // the stores went to peer memory; the ranks meet at a barrier after XP2
}

// XP2 (one CTA, after XP1 on the same stream): my chunk counts -> the header of my region on every destination.
struct XpushPublish {
  uint32_t* dst[kMaxRanks];
};
__global__ void k_xchg_publish(const uint32_t* __restrict__ cursor, uint32_t cap_chunks, uint32_t n_ranks, const XpushPublish p) {
  if (threadIdx.x < n_ranks) *p.dst[threadIdx.x] = min(cursor[threadIdx.x], cap_chunks);
  __threadfence_system();
}

}  // namespace hj

// ============================================================================================================= host side
namespace {

constexpr int kXpThreads = 1024;
template <typename K> constexpr int xp_virtual() { return sizeof(K) == 4 ? 4 : 2; }  // 64 KB of staging either way
constexpr uint32_t kXpSlackChunks = 1024;  // >= virtual CTAs of any sender: each may strand one partly filled chunk per region

// What every rank tells the others about its shards before the exchange (one all-gather of kXpDescWords words).
constexpr int kXpDescWords = 8;
struct ShardGeometry {
  bool regular = false;
  uint32_t first_chunk = 0;
  int32_t shift = 0;
};

// "regular": VALUE / DICT chunks of 2^k rows (the last one may be shorter) with consecutive chunk ids — the shape of a Hyrise
// data table (Chunk::MAX_SIZE-sized chunks appended in order, storage/table.cpp) and of a contiguous run of its chunks.
ShardGeometry shard_geometry(const hj_relation* r) {
  ShardGeometry g;
  if (r->n_segs == 0) {
    g.regular = true;
    return g;
  }
  for (uint32_t i = 0; i < r->n_segs; ++i) {
    const hj_segment& s = r->segs[i];
    if ((s.kind != HJ_SEG_VALUE && s.kind != HJ_SEG_DICT) || s.chunk_id != r->segs[0].chunk_id + i) return g;
  }
  g.first_chunk = r->segs[0].chunk_id;
  if (r->n_segs == 1) {
    if (r->segs[0].n_rows > (1u << 31)) return g;
    g.shift = 31;
    g.regular = true;
    return g;
  }
  const uint32_t first = r->segs[0].n_rows;
  if (!first || (first & (first - 1))) return g;
  while ((1u << g.shift) != first) ++g.shift;
  for (uint32_t i = 1; i < r->n_segs; ++i) {
    const uint32_t n = r->segs[i].n_rows;
    if ((i + 1 < r->n_segs && n != first) || n > first) return g;
  }
  g.regular = true;
  return g;
}

// The sharded join over the push exchange, for hashed key type K. `desc`: the all-gathered shard descriptors
// ([rank][kXpDescWords]: build rows, probe rows, build first chunk | shift << 32, probe first chunk | shift << 32, flags).
// *fell_back: a region overflowed somewhere — nothing was produced, the caller runs the general path (on every rank).
template <typename K>
int run_join_sharded_push(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                          hj_result_info* info, const unsigned long long* desc, int radix_bits, bool* fell_back) {
  using T = Tuple<K, false>;
  NcclApi* nc = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->nccl_comm);
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  const int probe_null_policy = outer ? kNullAsValue : kNullDrop;  // kNullRoute never comes here
  const int bits2 = pass2_bits<K>(radix_bits);
  *fell_back = false;
  Launches nl;
  cudaEvent_t* ev = ctx->ev;
  HJ_CUDA(ctx, cudaEventRecord(ev[9], st));

  // ---- geometry: regions of cap[side] chunks per (source, destination), the same number on every rank
  uint32_t cap[2];
  for (int side = 0; side < 2; ++side) {
    uint64_t max_rows = 0;
    for (uint32_t r = 0; r < R; ++r) {
      max_rows = std::max<uint64_t>(max_rows, desc[size_t(r) * kXpDescWords + side]);
      const unsigned long long g = desc[size_t(r) * kXpDescWords + 2 + side];
      ctx->xp_map[side].first_chunk[r] = static_cast<uint32_t>(g);
      ctx->xp_map[side].shift[r] = static_cast<int32_t>(g >> 32);
    }
    const uint64_t expect = (max_rows + R - 1) / R;  // rows one shard sends to one destination under a uniform hash
    cap[side] = static_cast<uint32_t>((expect + expect / 16 + kXpChunk - 1) / kXpChunk) + kXpSlackChunks;
  }

  // buffers outgrown during earlier sharded joins: their peers dropped the mappings when they saw the new handles
  for (void* p : ctx->xchg_graveyard) cudaFree(p);
  ctx->xchg_graveyard.clear();

  // ---- my shards as the kernels see them, receive regions, handles (collective; rank-local failures travel with it)
  PreparedRel ps_[2];
  int local = prepare_relation(ctx, build, 0, kXpChunk, st, &ps_[0]);
  if (local == HJ_OK) local = prepare_relation(ctx, probe, 1, kXpChunk, st, &ps_[1]);
  for (int side = 0; side < 2 && local == HJ_OK; ++side) {
    if (uint64_t(R) * cap[side] * kXpChunk >= 0xFFFFFFFFull)
      local = fail(ctx, HJ_ERR_TOO_LARGE, "rank %u: receive regions of %llu tuples (limit 2^32-2 per GPU)", me,
                   (unsigned long long)(uint64_t(R) * cap[side] * kXpChunk));
    if (local == HJ_OK) local = ensure_exported(ctx, ctx->xchg_recv_keys[side], size_t(R) * cap[side] * kXpChunk * sizeof(T));
    if (local == HJ_OK) local = ensure_exported(ctx, ctx->xchg_recv_rids[side], (size_t(kMaxRanks) + size_t(R) * cap[side]) * 4);
  }
  if (local == HJ_OK) local = ensure(ctx, ctx->xp_cursor, 64 * 4);
  HJ_TRY(exchange_handles(ctx, st, local));
  auto* cursor = static_cast<uint32_t*>(ctx->xp_cursor.p);  // [0..15] build cursors, [16..31] probe cursors, [32], [33] overflow flags
  HJ_CUDA(ctx, cudaMemsetAsync(cursor, 0, 64 * 4, st));
  // a region header nobody writes (a rank with an empty shard launches nothing) must read "0 chunks"
  for (int side = 0; side < 2; ++side) HJ_CUDA(ctx, cudaMemsetAsync(ctx->xchg_recv_rids[side].p, 0, kMaxRanks * 4, st));
  // every rank's headers are zeroed before anybody publishes into them
  auto* bar_dev = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p);
  HJ_NCCL(ctx, nc->AllGather(bar_dev, bar_dev + kMaxRanks, 1, ncclUint64, comm, st));

  // ---- XP1 + XP2 per side
  for (int side = 0; side < 2; ++side) {
    XpushArgs a{};
    XpushPublish pub{};
    a.rel = ps_[side].meta;
    a.cursor = cursor + side * kMaxRanks;
    a.overflow = cursor + 32 + side;
    a.cap_chunks = cap[side];
    a.n_ranks = R;
    a.null_policy = side == 0 ? kNullDrop : probe_null_policy;
    for (uint32_t d = 0; d < R; ++d) {
      char* tuples = static_cast<char*>(d == me ? ctx->xchg_recv_keys[side].p : ctx->xchg_peer[d][side * 3 + 0].ptr);
      auto* meta = static_cast<uint32_t*>(d == me ? ctx->xchg_recv_rids[side].p : ctx->xchg_peer[d][side * 3 + 1].ptr);
      if (!tuples || !meta) return fail(ctx, HJ_ERR_COMM, "rank %u has no mapping of rank %u's receive regions", me, d);
      a.dst_tuples[d] = tuples + size_t(me) * cap[side] * kXpChunk * sizeof(T);
      a.dst_fill[d] = meta + kMaxRanks + size_t(me) * cap[side];
      pub.dst[d] = meta + me;
    }
    if (ps_[side].meta.n_tiles) {
      constexpr int VC = xp_virtual<K>();
      auto k = k_xchg_push<K, kXpThreads, VC, int(kXpChunk)>;
      const size_t smem = size_t(VC) * (kXpChunk + kMaxRanks) * sizeof(T);
      HJ_TRY(set_smem(ctx, k, smem));
      const uint32_t grid = std::min<uint32_t>((ps_[side].meta.n_tiles + VC - 1) / VC, ctx->n_sms);
      k<<<grid, kXpThreads, smem, st>>>(a);
      HJ_LAUNCH_CHECK(ctx, "k_xchg_push");
      nl.n++;
    }
    k_xchg_publish<<<1, 32, 0, st>>>(a.cursor, cap[side], R, pub);
    HJ_LAUNCH_CHECK(ctx, "k_xchg_publish");
    nl.n++;
    if (side == 0) HJ_CUDA(ctx, cudaEventRecord(ev[10], st));
  }
  // barrier "all stores have landed" + everybody's overflow flags (one 64-bit word per rank: [32] | [33] << 32)
  auto* flag_dev = reinterpret_cast<unsigned long long*>(cursor + 32);
  HJ_NCCL(ctx, nc->AllGather(flag_dev, bar_dev + kMaxRanks, 1, ncclUint64, comm, st));
  auto* host_words = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedCounts);
  HJ_CUDA(ctx, cudaMemcpyAsync(host_words, bar_dev + kMaxRanks, size_t(R) * 8, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaMemcpyAsync(host_words + kMaxRanks, cursor, 32 * 4, cudaMemcpyDeviceToHost, st));  // my chunk counts (bytes sent)
  HJ_CUDA(ctx, cudaEventRecord(ev[11], st));

  // ---- local join over the regions: fused partitioning (k_part1p<XCH>, k_scatter2f) + k_build_probe<kRidXch>
  hj_segment seg[2][kMaxRanks];
  hj_relation rel[2];
  PreparedRel pr[2];
  PartState pst[2];
  struct RegionGuard {
    hj_ctx* c;
    ~RegionGuard() { for (int i = 0; i < kSides; ++i) c->xp_region[i] = hj_ctx::XpRegion{}; }
  } guard{ctx};
  for (int side = 0; side < 2; ++side) {
    const auto* meta = static_cast<const uint32_t*>(ctx->xchg_recv_rids[side].p);
    std::memset(seg[side], 0, sizeof(seg[side]));
    for (uint32_t r = 0; r < R; ++r) {
      seg[side][r].chunk_id = r;
      seg[side][r].n_rows = cap[side] * kXpChunk;
      seg[side][r].kind = HJ_SEG_VALUE;
      seg[side][r].key_type = sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32;
      seg[side][r].values = static_cast<const char*>(ctx->xchg_recv_keys[side].p) + size_t(r) * cap[side] * kXpChunk * sizeof(T);
    }
    rel[side].segs = seg[side];
    rel[side].n_segs = R;
    rel[side].location = HJ_LOC_DEVICE;
    rel[side].n_rows = uint64_t(R) * cap[side] * kXpChunk;
    ctx->xp_region[2 + side].fill = meta + kMaxRanks;
    ctx->xp_region[2 + side].nchunks = meta;
    ctx->xp_region[2 + side].cap_chunks = cap[side];
    HJ_TRY(prepare_relation(ctx, &rel[side], 2 + side, kXpTile, st, &pr[side]));
    HJ_TRY((stage_alloc<K, false>(ctx, pr[side], 2 + side, radix_bits, bits2, &pst[side])));
    if (!pst[side].xch) return fail(ctx, HJ_ERR_INVALID, "push exchange needs the fused partitioning path");
  }
  HJ_CUDA(ctx, cudaEventRecord(ev[1], st));
  HJ_TRY(stage_hist<K>(ctx, pr[0], 2, radix_bits, bits2, kNullAsValue, pst[0], st, &nl));
  HJ_TRY(stage_hist<K>(ctx, pr[1], 3, radix_bits, bits2, kNullAsValue, pst[1], st, &nl));
  HJ_CUDA(ctx, cudaEventRecord(ev[2], st));
  HJ_CUDA(ctx, cudaEventRecord(ev[3], st));
  HJ_TRY((stage_pass2<K, false>(ctx, pr[0], 2, radix_bits, bits2, kNullAsValue, pst[0], st, &nl)));
  HJ_TRY((stage_pass2<K, false>(ctx, pr[1], 3, radix_bits, bits2, kNullAsValue, pst[1], st, &nl)));
  HJ_CUDA(ctx, cudaEventRecord(ev[4], st));
  int passes = 0;
  // sizes for the planner and the first output estimate: what the regions can hold at most
  uint64_t est[2];
  for (int side = 0; side < 2; ++side) {
    uint64_t total = 0;
    for (uint32_t r = 0; r < R; ++r) total += desc[size_t(r) * kXpDescWords + side];
    est[side] = total / R + total / R / 16 + 1;
  }
  for (int side = 0; side < 2; ++side) est[side] = std::max<uint64_t>(est[side], rel[side].n_rows);  // the hard bound
  HJ_TRY((probe_phase<K, false>(ctx, pr[0], pr[1], pst[1], 2, 3, 0, radix_bits, bits2, mode, probe_null_policy, 2, est[0], est[1],
                                st, &nl, &passes, ev[5], ev[6])));
  // probe_phase synchronised the stream: the flags are here
  for (uint32_t r = 0; r < R; ++r)
    if (host_words[r]) *fell_back = true;
  if (*fell_back) return HJ_OK;

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
  info->n_partitions = 1u << radix_bits;
  info->n_overflow_partitions = rb->plan.n_overflow_parts;
  info->n_probe_passes = passes;
  info->n_probe_slices = 0;
  hj_timings& t = ctx->timings;
  std::memset(&t, 0, sizeof(t));
  t.ms[HJ_PH_XCHG_PARTITION] = ev_ms(ev[9], ev[10]);  // shard prepare, handles, build-side push
  t.ms[HJ_PH_XCHG_TRANSFER] = ev_ms(ev[10], ev[11]);  // probe-side push + barrier
  t.ms[HJ_PH_MATERIALIZE] = ev_ms(ev[1], ev[2]);      // pass 1 over the received tuples (both sides)
  t.ms[HJ_PH_SCATTER2] = ev_ms(ev[3], ev[4]);
  t.ms[HJ_PH_BUILD_PROBE] = passes == 1 ? ev_ms(ev[4], ev[5]) : 0.f;
  t.ms[HJ_PH_FINALIZE] = passes == 1 ? ev_ms(ev[5], ev[6]) : 0.f;
  t.ms[HJ_PH_TOTAL] = ev_ms(ev[9], ev[6]);
  t.n_kernel_launches = nl.n;
  t.xchg_path = 2;
  const auto* my_chunks = reinterpret_cast<const uint32_t*>(host_words + kMaxRanks);
  uint64_t sent = 0;
  for (int side = 0; side < 2; ++side)
    for (uint32_t d = 0; d < R; ++d)
      if (d != me) sent += uint64_t(my_chunks[side * kMaxRanks + d]) * kXpChunk * sizeof(T);
  t.bytes_nvlink_sent = sent;  // whole chunks: at most one partly filled chunk per virtual CTA and region too many
  return HJ_OK;
}


// ---- push exchange, streamed: host-resident shards, bound host outputs (hj_bind_output + hj_join_sharded) ---------------
// The sharded counterpart of run_join_streamed: every rank cuts its probe shard into the SAME number S of slices (runs of
// whole chunks). Per slice i: copy-in (stream copy_in) -> push into receive set i & 1 of every rank -> 1-word barrier -> local
// join of what arrived against the build side's partitions (exchanged and partitioned once, before the first slice) ->
// copy-out of the slice's PosLists into the bound host buffers (stream copy_out). Copy-in of slice i + 1 and copy-out of slice
// i - 1 overlap with the exchange + join of slice i, so both PCIe directions and the SMs work concurrently.
// Why two receive sets are enough: exchange and join of all slices run on ONE stream, so a rank enters barrier(i + 1) only
// after its join of slice i; a peer that has passed barrier(i + 1) may therefore overwrite set i & 1 (slice i + 2).
template <typename K>
int run_join_sharded_push_streamed(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                                   hj_result_info* info, const unsigned long long* desc, int radix_bits, uint32_t S, bool* fell_back) {
  using T = Tuple<K, false>;
  NcclApi* nc = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->nccl_comm);
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  const int probe_null_policy = outer ? kNullAsValue : kNullDrop;
  const int bits2 = pass2_bits<K>(radix_bits);
  *fell_back = false;
  Launches nl;
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));
  for (void* p : ctx->xchg_graveyard) cudaFree(p);
  ctx->xchg_graveyard.clear();

  // ---- my probe slices: runs of whole chunks with about the same number of rows; every rank has exactly S (possibly empty)
  std::vector<uint32_t> cut(S + 1, probe->n_segs);
  cut[0] = 0;
  {
    const uint64_t per = (probe->n_rows + S - 1) / S;
    uint64_t acc = 0;
    uint32_t k = 1;
    for (uint32_t i = 0; i < probe->n_segs && k < S; ++i) {
      acc += probe->segs[i].n_rows;
      if (acc >= per * k) cut[k++] = i + 1;
    }
  }
  auto slice_rel = [&](uint32_t i) {
    hj_relation r{};
    r.segs = probe->segs + cut[i];
    r.n_segs = cut[i + 1] - cut[i];
    r.location = probe->location;
    r.n_rows = 0;
    for (uint32_t k = 0; k < r.n_segs; ++k) r.n_rows += r.segs[k].n_rows;
    return r;
  };
  // position of a slice's first row in the shard (regular shards: every chunk before it has 2^shift rows)
  auto slice_pos = [&](uint32_t i) { return probe->n_segs > 1 ? cut[i] << ctx->xp_map[1].shift[me] : 0u; };

  // ---- geometry: region capacities (same on every rank). Build: as in the one-shot exchange; probe: per slice
  uint32_t cap_b, cap_p;
  uint64_t max_build = 0, max_probe = 0;
  for (uint32_t r = 0; r < R; ++r) {
    max_build = std::max<uint64_t>(max_build, desc[size_t(r) * kXpDescWords + 0]);
    max_probe = std::max<uint64_t>(max_probe, desc[size_t(r) * kXpDescWords + 1]);
    const unsigned long long gb = desc[size_t(r) * kXpDescWords + 2], gp = desc[size_t(r) * kXpDescWords + 3];
    ctx->xp_map[0].first_chunk[r] = static_cast<uint32_t>(gb);
    ctx->xp_map[0].shift[r] = static_cast<int32_t>(gb >> 32);
    ctx->xp_map[1].first_chunk[r] = static_cast<uint32_t>(gp);
    ctx->xp_map[1].shift[r] = static_cast<int32_t>(gp >> 32);
  }
  auto cap_for = [&](uint64_t rows) {
    const uint64_t expect = (rows + R - 1) / R;
    return static_cast<uint32_t>((expect + expect / 16 + kXpChunk - 1) / kXpChunk) + kXpSlackChunks;
  };
  cap_b = cap_for(max_build);
  // a slice holds at most one S-th of the largest shard plus one chunk of rounding (chunks are <= 2^shift rows)
  uint64_t max_chunk = 1;
  for (uint32_t r = 0; r < R; ++r) max_chunk = std::max<uint64_t>(max_chunk, desc[size_t(r) * kXpDescWords + 5]);
  cap_p = cap_for((max_probe + S - 1) / S + max_chunk);

  // ---- build shard + slice 0 staged; receive sets; handles
  PreparedRel psb, psp[2];
  hj_relation rel_s[2];
  int local = prepare_relation(ctx, build, 0, kXpChunk, st, &psb);
  rel_s[0] = slice_rel(0);
  if (local == HJ_OK) local = prepare_relation(ctx, &rel_s[0], 1, kXpChunk, ctx->copy_in, &psp[0]);
  if (local == HJ_OK && cudaEventRecord(ctx->ev_in[0], ctx->copy_in) != cudaSuccess) local = fail(ctx, HJ_ERR_CUDA, "cudaEventRecord failed");
  const uint32_t caps[3] = {cap_b, cap_p, cap_p};
  for (int set = 0; set < 3 && local == HJ_OK; ++set) {
    local = ensure_exported(ctx, ctx->xchg_recv_keys[set], size_t(R) * caps[set] * kXpChunk * sizeof(T));
    if (local == HJ_OK) local = ensure_exported(ctx, ctx->xchg_recv_rids[set], (size_t(kMaxRanks) + size_t(R) * caps[set]) * 4);
  }
  if (local == HJ_OK) local = ensure(ctx, ctx->xp_cursor, 64 * 4);
  HJ_TRY(exchange_handles(ctx, st, local));
  auto* cursor = static_cast<uint32_t*>(ctx->xp_cursor.p);  // [0..15] cursors of the push in flight, [32] overflow (sticky)
  auto* bar_dev = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p);
  HJ_CUDA(ctx, cudaMemsetAsync(cursor, 0, 64 * 4, st));
  // every rank has finished reading its receive sets of an earlier join before anybody stores into them again
  HJ_NCCL(ctx, nc->AllGather(bar_dev, bar_dev + kMaxRanks, 1, ncclUint64, comm, st));

  auto push = [&](const PreparedRel& pr, int set, int null_policy, uint32_t pos_base) -> int {
    XpushArgs a{};
    XpushPublish pub{};
    a.rel = pr.meta;
    a.pos_base = pos_base;
    a.cursor = cursor;
    a.overflow = cursor + 32;
    a.cap_chunks = caps[set];
    a.n_ranks = R;
    a.null_policy = null_policy;
    for (uint32_t d = 0; d < R; ++d) {
      char* tuples = static_cast<char*>(d == me ? ctx->xchg_recv_keys[set].p : ctx->xchg_peer[d][set * 3 + 0].ptr);
      auto* meta = static_cast<uint32_t*>(d == me ? ctx->xchg_recv_rids[set].p : ctx->xchg_peer[d][set * 3 + 1].ptr);
      if (!tuples || !meta) return fail(ctx, HJ_ERR_COMM, "rank %u has no mapping of rank %u's receive set %d", me, d, set);
      a.dst_tuples[d] = tuples + size_t(me) * caps[set] * kXpChunk * sizeof(T);
      a.dst_fill[d] = meta + kMaxRanks + size_t(me) * caps[set];
      pub.dst[d] = meta + me;
    }
    HJ_CUDA(ctx, cudaMemsetAsync(cursor, 0, kMaxRanks * 4, st));
    if (pr.meta.n_tiles) {
      constexpr int VC = xp_virtual<K>();
      auto k = k_xchg_push<K, kXpThreads, VC, int(kXpChunk)>;
      const size_t smem = size_t(VC) * (kXpChunk + kMaxRanks) * sizeof(T);
      HJ_TRY(set_smem(ctx, k, smem));
      const uint32_t grid = std::min<uint32_t>((pr.meta.n_tiles + VC - 1) / VC, ctx->n_sms);
      k<<<grid, kXpThreads, smem, st>>>(a);
      HJ_LAUNCH_CHECK(ctx, "k_xchg_push");
      nl.n++;
    }
    k_xchg_publish<<<1, 32, 0, st>>>(a.cursor, caps[set], R, pub);
    HJ_LAUNCH_CHECK(ctx, "k_xchg_publish");
    nl.n++;
    // barrier "all stores of this push have landed"
    HJ_NCCL(ctx, nc->AllGather(reinterpret_cast<unsigned long long*>(cursor + 32), bar_dev + kMaxRanks, 1, ncclUint64, comm, st));
    return HJ_OK;
  };
  hj_segment seg[3][kMaxRanks];
  hj_relation rrel[3];
  auto recv_relation = [&](int set, int side) {
    const auto* meta = static_cast<const uint32_t*>(ctx->xchg_recv_rids[set].p);
    std::memset(seg[set], 0, sizeof(seg[set]));
    for (uint32_t r = 0; r < R; ++r) {
      seg[set][r].chunk_id = r;
      seg[set][r].n_rows = caps[set] * kXpChunk;
      seg[set][r].kind = HJ_SEG_VALUE;
      seg[set][r].key_type = sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32;
      seg[set][r].values = static_cast<const char*>(ctx->xchg_recv_keys[set].p) + size_t(r) * caps[set] * kXpChunk * sizeof(T);
    }
    rrel[set].segs = seg[set];
    rrel[set].n_segs = R;
    rrel[set].location = HJ_LOC_DEVICE;
    rrel[set].n_rows = uint64_t(R) * caps[set] * kXpChunk;
    ctx->xp_region[side].fill = meta + kMaxRanks;
    ctx->xp_region[side].nchunks = meta;
    ctx->xp_region[side].cap_chunks = caps[set];
  };
  struct RegionGuard {
    hj_ctx* c;
    ~RegionGuard() { for (int i = 0; i < kSides; ++i) c->xp_region[i] = hj_ctx::XpRegion{}; }
  } guard{ctx};

  // ---- build side: exchange + partition once
  HJ_TRY(push(psb, 0, kNullDrop, 0));
  PreparedRel pb, pp[2];
  PartState sb, sp[2];
  recv_relation(0, 3);  // sides: 0 build shard, 1 / 2 probe slices being staged, 3 received build rows, 4 / 5 received probe slices
  HJ_TRY(prepare_relation(ctx, &rrel[0], 3, kXpTile, st, &pb));
  HJ_TRY((stage_alloc<K, false>(ctx, pb, 3, radix_bits, bits2, &sb)));
  HJ_TRY(stage_hist<K>(ctx, pb, 3, radix_bits, bits2, kNullAsValue, sb, st, &nl));
  HJ_TRY((stage_pass2<K, false>(ctx, pb, 3, radix_bits, bits2, kNullAsValue, sb, st, &nl)));

  auto* rb = static_cast<Readback*>(ctx->pinned);
  auto* host_words = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedCounts);
  uint64_t total = 0, probe_tuples = 0, h2d = psb.h2d_bytes;
  uint32_t overflow_parts = 0;
  int max_passes = 0, status = HJ_OK;
  for (uint32_t i = 0; i < S; ++i) {
    const int cur = i & 1, nxt = (i + 1) & 1;
    if (i + 1 < S) {  // stage the next slice into the other staging set, once the push that read it has run
      if (i >= 1) HJ_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, ctx->ev_x[nxt], 0));
      rel_s[nxt] = slice_rel(i + 1);
      psp[nxt] = PreparedRel{};
      HJ_TRY(prepare_relation(ctx, &rel_s[nxt], 1 + nxt, kXpChunk, ctx->copy_in, &psp[nxt]));
      HJ_CUDA(ctx, cudaEventRecord(ctx->ev_in[nxt], ctx->copy_in));
    }
    HJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_in[cur], 0));
    HJ_TRY(push(psp[cur], 1 + cur, probe_null_policy, slice_pos(i)));
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_x[cur], st));  // staging set `cur` may be overwritten
    if (i >= 2) HJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_out[cur], 0));  // result buffers of this set are free again
    const int side = 4 + cur;
    recv_relation(1 + cur, side);
    pp[cur] = PreparedRel{};
    HJ_TRY(prepare_relation(ctx, &rrel[1 + cur], side, kXpTile, st, &pp[cur]));
    HJ_TRY((stage_alloc<K, false>(ctx, pp[cur], side, radix_bits, bits2, &sp[cur])));
    HJ_TRY(stage_hist<K>(ctx, pp[cur], side, radix_bits, bits2, kNullAsValue, sp[cur], st, &nl));
    HJ_TRY((stage_pass2<K, false>(ctx, pp[cur], side, radix_bits, bits2, kNullAsValue, sp[cur], st, &nl)));
    HJ_CUDA(ctx, cudaMemcpyAsync(host_words, bar_dev + kMaxRanks, size_t(R) * 8, cudaMemcpyDeviceToHost, st));
    int passes = 0;
    HJ_TRY((probe_phase<K, false>(ctx, pb, pp[cur], sp[cur], 3, side, cur, radix_bits, bits2, mode, probe_null_policy, 2,
                                  rrel[0].n_rows, rrel[1 + cur].n_rows, st, &nl, &passes, nullptr, nullptr)));
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_done[cur], st));
    for (uint32_t r = 0; r < R; ++r)
      if (host_words[r]) *fell_back = true;  // a region overflowed somewhere: every rank sees it in the same slice
    if (*fell_back) break;
    const uint64_t rows = rb->out.n_rows;  // probe_phase synchronised the stream
    probe_tuples += rb->n_probe_tuples;
    overflow_parts = std::max(overflow_parts, rb->plan.n_overflow_parts);
    max_passes = std::max(max_passes, passes);
    h2d += psp[cur].h2d_bytes;
    if (total + rows > ctx->bound_capacity) {
      status = fail(ctx, HJ_ERR_CAPACITY, "bound output holds %llu rows, this rank's result has more (>= %llu after %u of %u probe slices)",
                    (unsigned long long)ctx->bound_capacity, (unsigned long long)(total + rows), i + 1, S);
      // (this rank keeps taking part in the remaining exchanges and barriers: the other ranks cannot know)
    } else if (rows) {
      HJ_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ctx->ev_done[cur], 0));
      if (!one_side)
        HJ_CUDA(ctx, cudaMemcpyAsync(ctx->bound_build + total, ctx->out_build[cur].p, rows * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
      HJ_CUDA(ctx, cudaMemcpyAsync(ctx->bound_probe + total, ctx->out_probe[cur].p, rows * 8, cudaMemcpyDeviceToHost, ctx->copy_out));
    }
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_out[cur], ctx->copy_out));
    if (status == HJ_OK) total += rows;
  }
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev[6], ctx->copy_out));
  HJ_CUDA(ctx, cudaStreamSynchronize(ctx->copy_in));
  HJ_CUDA(ctx, cudaStreamSynchronize(ctx->copy_out));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (*fell_back) return HJ_OK;
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
  info->n_probe_slices = S;
  hj_timings& t = ctx->timings;
  std::memset(&t, 0, sizeof(t));
  t.ms[HJ_PH_TOTAL] = ev_ms(ctx->ev[0], ctx->ev[6]);
  t.bytes_h2d = h2d;
  t.bytes_d2h = total * 8 * ctx->n_sides;
  t.n_kernel_launches = nl.n;
  t.xchg_path = 2;
  return HJ_OK;
}

}  // namespace
