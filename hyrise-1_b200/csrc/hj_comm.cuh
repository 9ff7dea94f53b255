// hj_comm.cuh — multi-GPU exchange (no reference counterpart: the reference is single-process).
//
// One process per GPU. Both relations are split by dest_rank_of_hash(hash(key)) so that every key's build and probe rows
// meet on exactly one rank; after the exchange each rank runs the unchanged single-GPU join on what it received. RowIDs
// travel with the keys (they are global), so per-rank results simply concatenate.
//
//   X1 k_xchg_count    decode (dictionary gather once, decoded keys kept like K1) + per-CTA counts per destination rank
//   X2 k_xchg_offsets  counts -> private send cursors per (destination, CTA) + totals per destination
//   X3 k_xchg_scatter  shared-memory-staged scatter of (key, RowID[, null flag]) into per-destination send regions (SoA)
//   all-to-all         counts: one ncclAllGather; data: one ncclGroup of ncclSend/ncclRecv per relation over NVLink
//
// NCCL is loaded with dlopen at hj_comm_init (the process usually has torch's libnccl.so.2 mapped already; linking a
// second copy at build time would pin a version).
#pragma once
#include <dlfcn.h>

namespace hj {

constexpr int kMaxRanks = 16;
constexpr uint64_t kMaxXchgSlices = 8;  // probe slices of the sliced sharded join

struct XchgArgs {
  RelMeta rel;
  void* keys_mat;            // decoded keys of DICT chunks (written by X1, read by X3)
  uint32_t* valid_mask;
  uint32_t* counts;          // [d * G + c]; X2 turns them into send cursors in place
  unsigned long long* send_cnt;  // [n_ranks] rows per destination
  // X3 destinations, one set per destination rank. Staged mode: all point into this rank's send buffers (adj = 0).
  // Direct mode: dst_*[d] is rank d's RECEIVE buffer, mapped into this process with CUDA IPC (own rank: the local
  // pointer), and adj[d] = (start of my region in d's buffer) - (start of region d in my send numbering), so that the
  // scatter kernel stores straight over NVLink and no send buffer, NCCL copy or staging pass exists.
  void* dst_keys[kMaxRanks];       // K per row
  uint64_t* dst_rids[kMaxRanks];   // packed global RowIDs
  uint8_t* dst_nulls[kMaxRanks];   // null flags (kNullRoute only) or nullptr
  uint32_t adj[kMaxRanks];
  uint32_t tile_begin, tile_end;   // the tiles of this launch (a slice of the shard; whole shard: 0 .. n_tiles)
  uint32_t n_ranks;
  uint32_t my_rank;
  int null_policy;
};

// Loads one tile's keys. DICT chunks go through the dictionary when `from_mat` is false (and the decoded key + validity
// bit are stored), else they are read back from keys_mat/valid_mask.
template <typename K, int NT, int ITEMS>
__device__ __forceinline__ void xchg_load_tile(const XchgArgs& a, const DevSegment& s, uint32_t r0, uint32_t cnt, bool from_mat,
                                               K (&key)[ITEMS], uint32_t& vbits) {
  K* keys_mat = static_cast<K*>(a.keys_mat);
  vbits = 0;
  if (s.kind == 1 && !from_mat) {
    uint32_t vid[ITEMS];
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * NT + threadIdx.x;
      vid[u] = i < cnt ? load_attr(s, r0 + i) : s.null_value_id;
    }
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const bool valid = vid[u] != s.null_value_id;  // inactive rows carry null_value_id
      key[u] = valid ? load_plain_key<K>(s, vid[u]) : K(0);
      vbits |= (valid ? 1u : 0u) << u;
    }
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * NT + threadIdx.x;
      if (i < cnt) keys_mat[s.row_begin + r0 + i] = key[u];
      const uint32_t bits = __ballot_sync(0xffffffffu, (vbits >> u) & 1u);
      if (lane_id() == 0 && i < cnt) a.valid_mask[s.mask_word_begin + ((r0 + i) >> 5)] = bits;
    }
  } else if (s.kind == 1) {
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * NT + threadIdx.x;
      key[u] = K(0);
      if (i < cnt) {
        key[u] = keys_mat[s.row_begin + r0 + i];
        vbits |= ((__ldg(a.valid_mask + s.mask_word_begin + ((r0 + i) >> 5)) >> (i & 31u)) & 1u) << u;
      }
    }
  } else {
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * NT + threadIdx.x;
      key[u] = K(0);
      if (i < cnt) {
        key[u] = load_plain_key<K>(s, r0 + i);
        const bool valid = !(s.nulls && __ldg(s.nulls + r0 + i));
        if (s.kind == 2 && !valid) key[u] = K(0);
        vbits |= (valid ? 1u : 0u) << u;
      }
    }
  }
}

// destination of one row, or kNoDigit when the row is dropped
template <typename K>
__device__ __forceinline__ uint32_t xchg_dest(const XchgArgs& a, K key, bool valid) {
  if (valid || a.null_policy == kNullAsValue) return dest_rank_of_hash(hash_key(key), a.n_ranks);
  if (a.null_policy == kNullRoute) return a.my_rank;  // stays here and is emitted as (NULL, p) by the local join
  return kNoDigit;
}

// X1. Persistent virtual CTA c owns tiles c, c+G, ... (X3 uses the same ownership: private cursors, no atomics).
template <typename K, int NT, int ITEMS>
__global__ void __launch_bounds__(NT) k_xchg_count(const XchgArgs a) {
  constexpr uint32_t TILE = NT * ITEMS;
  __shared__ uint32_t s_cnt[kMaxRanks];
  if (threadIdx.x < kMaxRanks) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  uint32_t cnt[kMaxRanks];
#pragma unroll
  for (int d = 0; d < kMaxRanks; ++d) cnt[d] = 0;
  for (uint32_t tile = a.tile_begin + blockIdx.x; tile < a.tile_end; tile += gridDim.x) {
    const DevSegment s = a.rel.segs[__ldg(a.rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t n_in = min(TILE, s.n_rows - r0);
    K key[ITEMS];
    uint32_t vbits;
    xchg_load_tile<K, NT, ITEMS>(a, s, r0, n_in, false, key, vbits);
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      if (u * NT + threadIdx.x < n_in) {
        const uint32_t d = xchg_dest<K>(a, key[u], (vbits >> u) & 1u);
#pragma unroll
        for (int q = 0; q < kMaxRanks; ++q) cnt[q] += (d == static_cast<uint32_t>(q)) ? 1u : 0u;
      }
    }
  }
#pragma unroll
  for (int q = 0; q < kMaxRanks; ++q) {
    uint32_t v = cnt[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane_id() == 0 && v) atomicAdd(&s_cnt[q], v);
  }
  __syncthreads();
  if (threadIdx.x < a.n_ranks) a.counts[threadIdx.x * gridDim.x + blockIdx.x] = s_cnt[threadIdx.x];
}

// X2. One CTA: totals per destination, exclusive scan over destinations (send region starts), then per destination the
// exclusive scan over the virtual CTAs (private cursors, in place).
template <int NT>
__global__ void __launch_bounds__(NT) k_xchg_offsets(uint32_t* __restrict__ counts, unsigned long long* __restrict__ send_cnt,
                                                     uint32_t n_ranks, uint32_t G) {
  __shared__ uint32_t scratch[33];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t d = 0; d < n_ranks; ++d) {
    const uint32_t start = s_carry;
    for (uint32_t base = 0; base < G; base += NT) {
      const uint32_t c = base + threadIdx.x;
      const uint32_t v = c < G ? counts[d * G + c] : 0u;
      uint32_t total;
      const uint32_t ex = block_excl_scan(v, scratch, &total) + s_carry;
      if (c < G) counts[d * G + c] = ex;
      __syncthreads();
      if (threadIdx.x == 0) s_carry += total;
      __syncthreads();
    }
    if (threadIdx.x == 0) send_cnt[d] = s_carry - start;
  }
}

// X3. Per tile: rank every row inside its destination with a shared-memory atomic, stage (key, RowID, flag) grouped by
// destination, copy the groups out with consecutive lanes on consecutive addresses.
// dynamic smem: TILE * (sizeof(K) + 8 + 1 + 1).
template <typename K, int NT, int ITEMS>
__global__ void __launch_bounds__(NT) k_xchg_scatter(const XchgArgs a) {
  constexpr uint32_t TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint64_t* s_rid = reinterpret_cast<uint64_t*>(s_raw);
  K* s_key = reinterpret_cast<K*>(s_rid + TILE);
  uint8_t* s_dig = reinterpret_cast<uint8_t*>(s_key + TILE);
  uint8_t* s_nul = s_dig + TILE;
  __shared__ uint32_t s_cnt[kMaxRanks], s_loff[kMaxRanks + 1], s_delta[kMaxRanks], s_cur[kMaxRanks];
  __shared__ K* s_dk[kMaxRanks];
  __shared__ uint64_t* s_dr[kMaxRanks];
  __shared__ uint8_t* s_dn[kMaxRanks];
  const uint32_t G = gridDim.x;
  if (threadIdx.x < a.n_ranks) {
    s_cur[threadIdx.x] = a.counts[threadIdx.x * G + blockIdx.x] + a.adj[threadIdx.x];
    s_dk[threadIdx.x] = static_cast<K*>(a.dst_keys[threadIdx.x]);
    s_dr[threadIdx.x] = a.dst_rids[threadIdx.x];
    s_dn[threadIdx.x] = a.dst_nulls[threadIdx.x];
  }

  for (uint32_t tile = a.tile_begin + blockIdx.x; tile < a.tile_end; tile += G) {
    const DevSegment s = a.rel.segs[__ldg(a.rel.tile_seg + tile)];
    const uint32_t r0 = (tile - s.tile_begin) * TILE;
    const uint32_t n_in = min(TILE, s.n_rows - r0);
    if (threadIdx.x < kMaxRanks) s_cnt[threadIdx.x] = 0;
    __syncthreads();  // also orders the previous tile's copy-out before this tile's staging
    K key[ITEMS];
    uint32_t vbits;
    xchg_load_tile<K, NT, ITEMS>(a, s, r0, n_in, true, key, vbits);
    uint32_t packed[ITEMS];
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      packed[u] = kNoDigit;
      if (u * NT + threadIdx.x < n_in) {
        const uint32_t d = xchg_dest<K>(a, key[u], (vbits >> u) & 1u);
        if (d != kNoDigit) packed[u] = (d << 16) | atomicAdd(&s_cnt[d], 1u);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t run = 0;
      for (uint32_t d = 0; d < a.n_ranks; ++d) {
        const uint32_t c = s_cnt[d];
        s_loff[d] = run;
        s_delta[d] = s_cur[d] - run;
        s_cur[d] += c;
        run += c;
      }
      s_loff[a.n_ranks] = run;
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      if (packed[u] != kNoDigit) {
        const uint32_t d = packed[u] >> 16, pos = s_loff[d] + (packed[u] & 0xFFFFu);
        s_key[pos] = key[u];
        s_rid[pos] = make_rowid(a.rel, s.row_begin + r0 + u * NT + threadIdx.x);
        s_dig[pos] = static_cast<uint8_t>(d);
        s_nul[pos] = ((vbits >> u) & 1u) ? 0 : 1;
      }
    }
    __syncthreads();
    const uint32_t n_out = s_loff[a.n_ranks];
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) {
      const uint32_t i = u * NT + threadIdx.x;
      if (i < n_out) {
        const uint32_t d = s_dig[i];
        const uint32_t g = s_delta[d] + i;
        s_dk[d][g] = s_key[i];
        s_dr[d][g] = s_rid[i];
        if (s_dn[d]) s_dn[d][g] = s_nul[i];
      }
    }
  }
  __threadfence_system();  // direct mode: the stores went to peer memory; the ranks meet at a barrier after this kernel
}

}  // namespace hj

// ============================================================================================================= host side
#include <nccl.h>

namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return &api;
  tried = true;
  for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
    api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle) {
    api.error = std::string("dlopen(libnccl.so.2) failed: ") + dlerror();
    return &api;
  }
  bool ok = true;
  auto sym = [&](const char* n) {
    void* p = dlsym(api.handle, n);
    if (!p) { ok = false; api.error = std::string("libnccl lacks ") + n; }
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
  api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
  api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
  api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  if (!ok) { dlclose(api.handle); api.handle = nullptr; }
  return &api;
}

#define HJ_NCCL(ctx, call)                                                                                   \
  do {                                                                                                       \
    ncclResult_t r_ = (call);                                                                                \
    if (r_ != ncclSuccess)                                                                                   \
      return fail(ctx, HJ_ERR_COMM, "%s failed: %s (%s:%d)", #call, nccl_api()->GetErrorString(r_), __FILE__, __LINE__); \
  } while (0)

static_assert(sizeof(ncclUniqueId) == HJ_NCCL_UNIQUE_ID_BYTES, "hj.h carries the NCCL unique id as 128 opaque bytes");

struct XchgSide {          // per relation, filled by the partition step
  PreparedRel pr;
  uint32_t G = 0;
  int null_policy = kNullDrop;
  bool send_nulls = false;
};

constexpr int kXchgThreads = 256;
template <typename K> constexpr int xchg_items() { return sizeof(K) == 4 ? 16 : 8; }
template <typename K> constexpr int xchg_tile() { return kXchgThreads * xchg_items<K>(); }

template <typename K>
XchgArgs xchg_args(hj_ctx* ctx, const XchgSide& xs, int side) {
  XchgArgs a{};
  a.rel = xs.pr.meta;
  a.keys_mat = ctx->keys_mat[side].p;
  a.valid_mask = static_cast<uint32_t*>(ctx->valid_mask[side].p);
  a.counts = static_cast<uint32_t*>(ctx->xchg_counts[side].p);
  a.send_cnt = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p) + size_t(side) * kMaxRanks;
  for (int d = 0; d < kMaxRanks; ++d) {  // staged mode; run_join_sharded overwrites these in direct mode
    a.dst_keys[d] = ctx->xchg_send_keys[side].p;
    a.dst_rids[d] = static_cast<uint64_t*>(ctx->xchg_send_rids[side].p);
    a.dst_nulls[d] = xs.send_nulls ? static_cast<uint8_t*>(ctx->xchg_send_nulls[side].p) : nullptr;
    a.adj[d] = 0;
  }
  a.tile_begin = 0;
  a.tile_end = xs.pr.meta.n_tiles;
  a.n_ranks = ctx->n_ranks;
  a.my_rank = ctx->rank;
  a.null_policy = xs.null_policy;
  return a;
}

// X1 + X2 for one relation shard
template <typename K>
int xchg_count(hj_ctx* ctx, const hj_relation* rel, int side, XchgSide* xs, cudaStream_t st, uint32_t* n_launch) {
  HJ_TRY(prepare_relation(ctx, rel, side, xchg_tile<K>(), st, &xs->pr));
  const uint32_t n = xs->pr.meta.n_rows;
  xs->G = std::max<uint32_t>(1, std::min<uint32_t>(xs->pr.meta.n_tiles, uint32_t(kVirtualCtasPerSm) * ctx->n_sms));
  HJ_TRY(ensure(ctx, ctx->xchg_counts[side], size_t(xs->G) * kMaxRanks * 4));
  if (xs->pr.has_dict) {
    HJ_TRY(ensure(ctx, ctx->keys_mat[side], std::max<size_t>(1, n) * sizeof(K)));
    HJ_TRY(ensure(ctx, ctx->valid_mask[side], std::max<size_t>(1, xs->pr.mask_words) * 4));
  }
  XchgArgs a = xchg_args<K>(ctx, *xs, side);
  HJ_CUDA(ctx, cudaMemsetAsync(a.counts, 0, size_t(xs->G) * kMaxRanks * 4, st));
  if (xs->pr.meta.n_tiles) {
    k_xchg_count<K, kXchgThreads, xchg_items<K>()><<<xs->G, kXchgThreads, 0, st>>>(a);
    HJ_LAUNCH_CHECK(ctx, "k_xchg_count");
    ++*n_launch;
  }
  k_xchg_offsets<256><<<1, 256, 0, st>>>(a.counts, a.send_cnt, ctx->n_ranks, xs->G);
  HJ_LAUNCH_CHECK(ctx, "k_xchg_offsets");
  ++*n_launch;
  return HJ_OK;
}

}  // namespace

namespace {

// ---- direct (peer-mapped) exchange: CUDA IPC plumbing ------------------------------------------------------------------
constexpr int kXchgBufKinds = 9;  // [receive set][keys, rids, nulls]
struct HandleRec {                // what every rank publishes before X3 in direct mode (1 KB)
  cudaIpcMemHandle_t h[kXchgBufKinds];
  uint8_t valid[kXchgBufKinds];
  uint8_t pad[1024 - kXchgBufKinds * sizeof(cudaIpcMemHandle_t) - kXchgBufKinds];
};
static_assert(sizeof(HandleRec) == 1024 && kPinnedHandles + (1 + kMaxRanks) * sizeof(HandleRec) <= kPinnedBytes, "HandleRec layout");

DevBuf* xchg_recv_buf(hj_ctx* ctx, int side, int kind) {
  return kind == 0 ? &ctx->xchg_recv_keys[side] : kind == 1 ? &ctx->xchg_recv_rids[side] : &ctx->xchg_recv_nulls[side];
}

// Grow-only like ensure(), but a buffer peers may still have mapped is parked (freed once every rank has entered the next
// sharded join, i.e. has dropped its mapping) and the new one gets 1/8 slack so that re-exports stay rare.
int ensure_exported(hj_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return HJ_OK;
  if (b.p) ctx->xchg_graveyard.push_back(b.p);
  b.p = nullptr;
  b.cap = 0;
  const size_t want = ((bytes + bytes / 8) + 255) & ~size_t(255);
  HJ_CUDA(ctx, cudaMalloc(&b.p, want));
  b.cap = want;
  return HJ_OK;
}

void close_peer_maps(hj_ctx* ctx) {
  for (auto& per_rank : ctx->xchg_peer)
    for (auto& m : per_rank) {
      if (m.ptr) cudaIpcCloseMemHandle(m.ptr);
      m.ptr = nullptr;
    }
  (void)cudaGetLastError();
}

// All ranks publish the IPC handles of their receive buffers; every rank maps what changed. Collective.
// local_status: what this rank's preparations (allocations) returned; every rank learns every rank's status in the same
// all-gather, so that all of them leave the join together instead of one returning early and the others hanging in the next
// collective. Returns HJ_ERR_COMM when a PEER failed (this rank's own error is already in hj_last_error).
int exchange_handles(hj_ctx* ctx, cudaStream_t st, int local_status = HJ_OK) {
  NcclApi* nc = nccl_api();
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  HJ_TRY(ensure(ctx, ctx->xchg_hdl_dev, sizeof(HandleRec) * (1 + kMaxRanks)));
  auto* dev = static_cast<HandleRec*>(ctx->xchg_hdl_dev.p);
  auto* host = reinterpret_cast<HandleRec*>(static_cast<char*>(ctx->pinned) + kPinnedHandles);  // [0] mine, [1..R] everybody's
  std::memset(&host[0], 0, sizeof(HandleRec));
  host[0].pad[0] = static_cast<uint8_t>(local_status);
  for (int side = 0; side < 3 && local_status == HJ_OK; ++side)
    for (int kind = 0; kind < 3; ++kind) {
      DevBuf* b = xchg_recv_buf(ctx, side, kind);
      if (!b->p) continue;
      if (cudaIpcGetMemHandle(&host[0].h[side * 3 + kind], b->p) != cudaSuccess) {
        (void)cudaGetLastError();
        host[0].pad[0] = static_cast<uint8_t>(local_status = fail(ctx, HJ_ERR_CUDA, "cudaIpcGetMemHandle failed for receive buffer %d/%d", side, kind));
        break;
      }
      host[0].valid[side * 3 + kind] = 1;
    }
  HJ_CUDA(ctx, cudaMemcpyAsync(dev, &host[0], sizeof(HandleRec), cudaMemcpyHostToDevice, st));
  HJ_NCCL(ctx, nc->AllGather(dev, dev + 1, sizeof(HandleRec), ncclUint8, static_cast<ncclComm_t>(ctx->nccl_comm), st));
  HJ_CUDA(ctx, cudaMemcpyAsync(&host[1], dev + 1, sizeof(HandleRec) * R, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (local_status != HJ_OK) return local_status;
  for (uint32_t p = 0; p < R; ++p)
    if (host[1 + p].pad[0] != HJ_OK) return fail(ctx, HJ_ERR_COMM, "rank %u could not prepare its receive buffers (status %u)", p, host[1 + p].pad[0]);
  for (uint32_t p = 0; p < R; ++p) {
    if (p == me) continue;
    for (int k = 0; k < kXchgBufKinds; ++k) {
      auto& m = ctx->xchg_peer[p][k];
      const HandleRec& rec = host[1 + p];
      if (!rec.valid[k]) continue;
      if (m.ptr && std::memcmp(&m.handle, &rec.h[k], sizeof(cudaIpcMemHandle_t)) == 0) continue;
      if (m.ptr) HJ_CUDA(ctx, cudaIpcCloseMemHandle(m.ptr));
      m.ptr = nullptr;
      HJ_CUDA(ctx, cudaIpcOpenMemHandle(&m.ptr, rec.h[k], cudaIpcMemLazyEnablePeerAccess));
      m.handle = rec.h[k];
    }
  }
  return HJ_OK;
}

// hj_comm_init: can every rank map every other rank's memory? (false e.g. for ranks that are threads of one process)
bool probe_p2p(hj_ctx* ctx) {
  NcclApi* nc = nccl_api();
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  cudaStream_t st = ctx->own_stream;
  void* test = nullptr;
  unsigned long long ok = 1;
  if (cudaMalloc(&test, 256) != cudaSuccess) ok = 0;
  if (ensure(ctx, ctx->xchg_hdl_dev, sizeof(HandleRec) * (1 + kMaxRanks)) != HJ_OK) return false;
  auto* dev = static_cast<HandleRec*>(ctx->xchg_hdl_dev.p);
  auto* host = reinterpret_cast<HandleRec*>(static_cast<char*>(ctx->pinned) + kPinnedHandles);
  std::memset(&host[0], 0, sizeof(HandleRec));
  if (ok && cudaIpcGetMemHandle(&host[0].h[0], test) != cudaSuccess) ok = 0;
  host[0].valid[0] = ok ? 1 : 0;
  bool comm_ok = cudaMemcpyAsync(dev, &host[0], sizeof(HandleRec), cudaMemcpyHostToDevice, st) == cudaSuccess &&
                 nc->AllGather(dev, dev + 1, sizeof(HandleRec), ncclUint8, static_cast<ncclComm_t>(ctx->nccl_comm), st) == ncclSuccess &&
                 cudaMemcpyAsync(&host[1], dev + 1, sizeof(HandleRec) * R, cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                 cudaStreamSynchronize(st) == cudaSuccess;
  if (!comm_ok) ok = 0;
  for (uint32_t p = 0; p < R && ok; ++p) {
    if (p == me) continue;
    void* mapped = nullptr;
    if (!host[1 + p].valid[0] || cudaIpcOpenMemHandle(&mapped, host[1 + p].h[0], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = 0;
    else cudaIpcCloseMemHandle(mapped);
  }
  (void)cudaGetLastError();
  // agree: direct mode only if EVERY rank can map every peer (the ranks must take the same branch later)
  auto* flag_dev = reinterpret_cast<unsigned long long*>(dev);
  auto* flag_host = reinterpret_cast<unsigned long long*>(host);
  bool all = comm_ok && cudaMemcpyAsync(flag_dev, &ok, 8, cudaMemcpyHostToDevice, st) == cudaSuccess &&
             nc->AllGather(flag_dev, flag_dev + 8, 1, ncclUint64, static_cast<ncclComm_t>(ctx->nccl_comm), st) == ncclSuccess &&
             cudaMemcpyAsync(flag_host, flag_dev + 8, 8 * R, cudaMemcpyDeviceToHost, st) == cudaSuccess &&
             cudaStreamSynchronize(st) == cudaSuccess;
  for (uint32_t p = 0; p < R && all; ++p) all = flag_host[p] != 0;
  // peers may still be opening `test`: keep it until the communicator goes away
  if (test) ctx->xchg_graveyard.push_back(test);
  return all;
}

// The whole sharded join for hashed key type K (agreed on by all ranks).
template <typename K>
int run_join_sharded(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                     hj_result_info* info) {
  NcclApi* nc = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->nccl_comm);
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  const int probe_null_policy = !outer ? kNullDrop : ((ctx->opts.flags & HJ_FLAG_SQL_NULL_PROBE) ? kNullRoute : kNullAsValue);
  uint32_t n_launch = 0;
  cudaEvent_t* ev = ctx->ev;
  HJ_CUDA(ctx, cudaEventRecord(ev[9], st));

  // ---- X1/X2 for both relations, then one all-gather of the 2 x R send counts
  XchgSide xs[2];
  xs[0].null_policy = kNullDrop;
  xs[1].null_policy = probe_null_policy;
  xs[1].send_nulls = probe_null_policy == kNullRoute;
  HJ_TRY(ensure(ctx, ctx->xchg_cnt_dev, size_t(2) * kMaxRanks * 8 * (1 + kMaxRanks)));
  auto* cnt_dev = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p);       // [2][kMaxRanks] mine
  auto* mat_dev = cnt_dev + 2 * kMaxRanks;                                     // [R][2][kMaxRanks] everybody's
  HJ_CUDA(ctx, cudaMemsetAsync(cnt_dev, 0, size_t(2) * kMaxRanks * 8, st));
  HJ_TRY(xchg_count<K>(ctx, build, 0, &xs[0], st, &n_launch));
  HJ_TRY(xchg_count<K>(ctx, probe, 1, &xs[1], st, &n_launch));
  HJ_NCCL(ctx, nc->AllGather(cnt_dev, mat_dev, 2 * kMaxRanks, ncclUint64, comm, st));
  auto* mat = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedCounts);
  HJ_CUDA(ctx, cudaMemcpyAsync(mat, mat_dev, size_t(R) * 2 * kMaxRanks * 8, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  auto cnt = [&](uint32_t src, int side, uint32_t dst) { return mat[(size_t(src) * 2 + side) * kMaxRanks + dst]; };

  // buffers outgrown during earlier sharded joins: every rank has entered this join, so nobody maps them any more
  for (void* p : ctx->xchg_graveyard) cudaFree(p);
  ctx->xchg_graveyard.clear();

  uint64_t n_recv[2] = {0, 0};
  const bool direct = ctx->p2p_ok && !(ctx->opts.flags & HJ_FLAG_XCHG_STAGED);
  if (direct) {
    // ---- X3 stores straight into the peers' receive buffers (CUDA IPC mappings over NVLink): no send buffers, no NCCL
    // copy; the only collectives are the count all-gather above, the handle all-gather and a 1-word barrier.
    // every rank holds the whole count matrix: the "too large" verdict is computed for ALL ranks, so all fail identically
    for (int side = 0; side < 2; ++side)
      for (uint32_t r = 0; r < R; ++r) {
        uint64_t n = 0;
        for (uint32_t p = 0; p < R; ++p) n += cnt(p, side, r);
        if (r == me) n_recv[side] = n;
        if (n >= 0xFFFFFFFFull)
          return fail(ctx, HJ_ERR_TOO_LARGE, "rank %u would receive %llu %s rows (limit 2^32-2 per GPU)", r,
                      (unsigned long long)n, side ? "probe" : "build");
      }
    // allocations are rank-local: their status travels with the handle all-gather, so that no rank is left waiting
    int local = HJ_OK;
    for (int side = 0; side < 2 && local == HJ_OK; ++side) {
      local = ensure_exported(ctx, ctx->xchg_recv_keys[side], std::max<uint64_t>(1, n_recv[side]) * sizeof(K));
      if (local == HJ_OK) local = ensure_exported(ctx, ctx->xchg_recv_rids[side], std::max<uint64_t>(1, n_recv[side]) * 8);
      if (local == HJ_OK && xs[side].send_nulls) local = ensure_exported(ctx, ctx->xchg_recv_nulls[side], std::max<uint64_t>(1, n_recv[side]));
    }
    HJ_TRY(exchange_handles(ctx, st, local));
    for (int side = 0; side < 2; ++side) {
      uint64_t n_send = 0;
      if (xs[side].pr.meta.n_tiles) {
        XchgArgs a = xchg_args<K>(ctx, xs[side], side);
        uint64_t send_off = 0;
        for (uint32_t d = 0; d < R; ++d) {
          uint64_t base_in_d = 0;  // where my rows start in rank d's receive buffer: after the rows of the ranks before me
          for (uint32_t s2 = 0; s2 < me; ++s2) base_in_d += cnt(s2, side, d);
          a.adj[d] = static_cast<uint32_t>(base_in_d - send_off);
          send_off += cnt(me, side, d);
          if (d == me) {
            a.dst_keys[d] = ctx->xchg_recv_keys[side].p;
            a.dst_rids[d] = static_cast<uint64_t*>(ctx->xchg_recv_rids[side].p);
            a.dst_nulls[d] = xs[side].send_nulls ? static_cast<uint8_t*>(ctx->xchg_recv_nulls[side].p) : nullptr;
          } else {
            a.dst_keys[d] = ctx->xchg_peer[d][side * 3 + 0].ptr;
            a.dst_rids[d] = static_cast<uint64_t*>(ctx->xchg_peer[d][side * 3 + 1].ptr);
            a.dst_nulls[d] = xs[side].send_nulls ? static_cast<uint8_t*>(ctx->xchg_peer[d][side * 3 + 2].ptr) : nullptr;
            if (cnt(me, side, d) && (!a.dst_keys[d] || !a.dst_rids[d] || (xs[side].send_nulls && !a.dst_nulls[d])))
              return fail(ctx, HJ_ERR_COMM, "rank %u has no mapping of rank %u's receive buffers", me, d);
          }
        }
        n_send = send_off;
        auto k = k_xchg_scatter<K, kXchgThreads, xchg_items<K>()>;
        const size_t smem = size_t(xchg_tile<K>()) * (sizeof(K) + 8 + 2);
        HJ_TRY(set_smem(ctx, k, smem));
        k<<<xs[side].G, kXchgThreads, smem, st>>>(a);
        HJ_LAUNCH_CHECK(ctx, "k_xchg_scatter");
        ++n_launch;
      }
      if (side == 0) HJ_CUDA(ctx, cudaEventRecord(ev[10], st));
      ctx->xchg_bytes_sent[side] = (n_send - cnt(me, side, me)) * (sizeof(K) + 8 + (xs[side].send_nulls ? 1 : 0));
    }
    // barrier: a rank may read its receive buffers only after EVERY rank's X3 has finished storing into them
    HJ_NCCL(ctx, nc->AllGather(cnt_dev, mat_dev, 1, ncclUint64, comm, st));
  } else {
    // ---- staged: X3 into send buffers, then one grouped ncclSend/ncclRecv (= all-to-all-v) per relation
  for (int side = 0; side < 2; ++side) {
    uint64_t send_off[kMaxRanks + 1] = {0}, recv_off[kMaxRanks + 1] = {0};
    for (uint32_t p = 0; p < R; ++p) {
      send_off[p + 1] = send_off[p] + cnt(me, side, p);
      recv_off[p + 1] = recv_off[p] + cnt(p, side, me);
    }
    const uint64_t n_send = send_off[R];
    n_recv[side] = recv_off[R];
    if (n_recv[side] >= 0xFFFFFFFFull)
      return fail(ctx, HJ_ERR_TOO_LARGE, "rank %u would receive %llu %s rows (limit 2^32-2 per GPU)", me,
                  (unsigned long long)n_recv[side], side ? "probe" : "build");
    const bool nulls = xs[side].send_nulls;
    HJ_TRY(ensure(ctx, ctx->xchg_send_keys[side], std::max<uint64_t>(1, n_send) * sizeof(K)));
    HJ_TRY(ensure(ctx, ctx->xchg_send_rids[side], std::max<uint64_t>(1, n_send) * 8));
    HJ_TRY(ensure(ctx, ctx->xchg_recv_keys[side], std::max<uint64_t>(1, n_recv[side]) * sizeof(K)));
    HJ_TRY(ensure(ctx, ctx->xchg_recv_rids[side], std::max<uint64_t>(1, n_recv[side]) * 8));
    if (nulls) {
      HJ_TRY(ensure(ctx, ctx->xchg_send_nulls[side], std::max<uint64_t>(1, n_send)));
      HJ_TRY(ensure(ctx, ctx->xchg_recv_nulls[side], std::max<uint64_t>(1, n_recv[side])));
    }
    if (xs[side].pr.meta.n_tiles) {
      XchgArgs a = xchg_args<K>(ctx, xs[side], side);
      auto k = k_xchg_scatter<K, kXchgThreads, xchg_items<K>()>;
      const size_t smem = size_t(xchg_tile<K>()) * (sizeof(K) + 8 + 2);
      HJ_TRY(set_smem(ctx, k, smem));
      k<<<xs[side].G, kXchgThreads, smem, st>>>(a);
      HJ_LAUNCH_CHECK(ctx, "k_xchg_scatter");
      ++n_launch;
    }
    if (side == 0) HJ_CUDA(ctx, cudaEventRecord(ev[10], st));  // (build partitioned; probe partition overlaps nothing yet)
    char* sk = static_cast<char*>(ctx->xchg_send_keys[side].p);
    char* sr = static_cast<char*>(ctx->xchg_send_rids[side].p);
    char* sn = static_cast<char*>(ctx->xchg_send_nulls[side].p);
    char* rk = static_cast<char*>(ctx->xchg_recv_keys[side].p);
    char* rr = static_cast<char*>(ctx->xchg_recv_rids[side].p);
    char* rn = static_cast<char*>(ctx->xchg_recv_nulls[side].p);
    // own share: device-to-device copy; everything else: one grouped send/recv (= all-to-all-v)
    {
      const uint64_t c = cnt(me, side, me);
      if (c) {
        HJ_CUDA(ctx, cudaMemcpyAsync(rk + recv_off[me] * sizeof(K), sk + send_off[me] * sizeof(K), c * sizeof(K), cudaMemcpyDeviceToDevice, st));
        HJ_CUDA(ctx, cudaMemcpyAsync(rr + recv_off[me] * 8, sr + send_off[me] * 8, c * 8, cudaMemcpyDeviceToDevice, st));
        if (nulls) HJ_CUDA(ctx, cudaMemcpyAsync(rn + recv_off[me], sn + send_off[me], c, cudaMemcpyDeviceToDevice, st));
      }
    }
    HJ_NCCL(ctx, nc->GroupStart());
    for (uint32_t p = 0; p < R; ++p) {
      if (p == me) continue;
      const uint64_t cs = cnt(me, side, p), cr = cnt(p, side, me);
      if (cs) {
        HJ_NCCL(ctx, nc->Send(sk + send_off[p] * sizeof(K), cs * sizeof(K), ncclUint8, (int)p, comm, st));
        HJ_NCCL(ctx, nc->Send(sr + send_off[p] * 8, cs * 8, ncclUint8, (int)p, comm, st));
        if (nulls) HJ_NCCL(ctx, nc->Send(sn + send_off[p], cs, ncclUint8, (int)p, comm, st));
      }
      if (cr) {
        HJ_NCCL(ctx, nc->Recv(rk + recv_off[p] * sizeof(K), cr * sizeof(K), ncclUint8, (int)p, comm, st));
        HJ_NCCL(ctx, nc->Recv(rr + recv_off[p] * 8, cr * 8, ncclUint8, (int)p, comm, st));
        if (nulls) HJ_NCCL(ctx, nc->Recv(rn + recv_off[p], cr, ncclUint8, (int)p, comm, st));
      }
    }
    HJ_NCCL(ctx, nc->GroupEnd());
    ctx->xchg_bytes_sent[side] = (n_send - cnt(me, side, me)) * (sizeof(K) + 8 + (nulls ? 1 : 0));
  }
  }
  HJ_CUDA(ctx, cudaEventRecord(ev[11], st));

  // ---- local join over what arrived: one device-resident VALUE segment per relation whose RowIDs come from the exchange
  hj_segment seg[2];
  hj_relation rel[2];
  for (int side = 0; side < 2; ++side) {
    std::memset(&seg[side], 0, sizeof(hj_segment));
    seg[side].chunk_id = 0;
    seg[side].n_rows = static_cast<uint32_t>(n_recv[side]);
    seg[side].kind = HJ_SEG_VALUE;
    seg[side].key_type = sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32;
    seg[side].values = ctx->xchg_recv_keys[side].p;
    seg[side].nulls = xs[side].send_nulls ? static_cast<const uint8_t*>(ctx->xchg_recv_nulls[side].p) : nullptr;
    rel[side].segs = &seg[side];
    rel[side].n_segs = 1;
    rel[side].location = HJ_LOC_DEVICE;
    rel[side].n_rows = n_recv[side];
    ctx->xchg_rowids[side] = static_cast<const uint64_t*>(ctx->xchg_recv_rids[side].p);
  }
  // the tuples carry the RowID next to the key: 12 bytes for int32 keys, the 16-byte int64 tuple has room anyway
  const int st_join = run_join<K, sizeof(K) == 4>(ctx, &rel[0], &rel[1], mode, st, info, /*payload=*/true);
  ctx->xchg_rowids[0] = ctx->xchg_rowids[1] = nullptr;
  HJ_TRY(st_join);
  hj_timings& t = ctx->timings;
  t.ms[HJ_PH_XCHG_PARTITION] = ev_ms(ev[9], ev[10]);   // shard prepare + X1/X2 (both) + build X3
  t.ms[HJ_PH_XCHG_TRANSFER] = ev_ms(ev[10], ev[11]);   // build all-to-all, probe X3, probe all-to-all
  t.ms[HJ_PH_TOTAL] += ev_ms(ev[9], ev[11]);
  t.n_kernel_launches += n_launch;
  t.xchg_path = 1;
  t.bytes_nvlink_sent = ctx->xchg_bytes_sent[0] + ctx->xchg_bytes_sent[1];
  return HJ_OK;
}

// Grow-only device buffer that keeps its first `keep` bytes (result accumulators).
int ensure_preserve(hj_ctx* ctx, DevBuf& b, size_t bytes, size_t keep, cudaStream_t st) {
  if (bytes <= b.cap) return HJ_OK;
  void* old = b.p;
  const size_t want = ((bytes + bytes / 2) + 255) & ~size_t(255);
  void* fresh = nullptr;
  HJ_CUDA(ctx, cudaMalloc(&fresh, want));
  if (old && keep) HJ_CUDA(ctx, cudaMemcpyAsync(fresh, old, keep, cudaMemcpyDeviceToDevice, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  if (old) HJ_CUDA(ctx, cudaFree(old));
  b.p = fresh;
  b.cap = want;
  return HJ_OK;
}

// The sharded join with the probe side cut into S slices (tile ranges of every rank's probe shard): while the rows of
// slice s are partitioned and probed locally (stream `st`), slice s+1 is stored into the peers' other receive set over
// NVLink (exchange stream `cs`). All counting (X1/X2 for the build shard and all S slices) happens up front with ONE
// count all-gather, so the loop has no host synchronisation on the exchange stream; the 1-word all-gather after each X3
// is the barrier "everybody's stores for this slice have landed", and because a rank only enters it after its local join
// of the slice that last used the receive set, it also protects that set from being overwritten too early.
// Direct (peer-mapped) exchange only; S is the same on every rank.
template <typename K>
int run_join_sharded_sliced(hj_ctx* ctx, const hj_relation* build, const hj_relation* probe, int mode, cudaStream_t st,
                            hj_result_info* info, uint32_t S) {
  constexpr bool PAY = sizeof(K) == 4;
  NcclApi* nc = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->nccl_comm);
  cudaStream_t cs = ctx->copy_in;
  const uint32_t R = ctx->n_ranks, me = ctx->rank;
  const bool one_side = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  const bool outer = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  const int probe_null_policy = !outer ? kNullDrop : ((ctx->opts.flags & HJ_FLAG_SQL_NULL_PROBE) ? kNullRoute : kNullAsValue);
  uint32_t n_launch = 0;
  cudaEvent_t* ev = ctx->ev;
  HJ_CUDA(ctx, cudaEventRecord(ev[9], st));
  HJ_CUDA(ctx, cudaStreamWaitEvent(cs, ev[9], 0));

  // ---- X1/X2: the build shard (slot 0) and the S probe slices (slots 1..S), then one all-gather of (S+1) x R counts
  const uint32_t n_slots = S + 1;
  XchgSide xs[2];
  xs[0].null_policy = kNullDrop;
  xs[1].null_policy = probe_null_policy;
  xs[1].send_nulls = probe_null_policy == kNullRoute;
  HJ_TRY(ensure(ctx, ctx->xchg_cnt_dev, size_t(n_slots) * kMaxRanks * 8 * (1 + kMaxRanks)));
  auto* cnt_dev = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p);   // [n_slots][kMaxRanks] mine
  auto* mat_dev = cnt_dev + size_t(n_slots) * kMaxRanks;                   // [R][n_slots][kMaxRanks] everybody's
  HJ_CUDA(ctx, cudaMemsetAsync(cnt_dev, 0, size_t(n_slots) * kMaxRanks * 8, cs));
  HJ_TRY(xchg_count<K>(ctx, build, 0, &xs[0], cs, &n_launch));  // send_cnt = cnt_dev + 0
  HJ_TRY(prepare_relation(ctx, probe, 1, xchg_tile<K>(), cs, &xs[1].pr));
  const uint32_t T = xs[1].pr.meta.n_tiles;
  if (xs[1].pr.has_dict) {
    HJ_TRY(ensure(ctx, ctx->keys_mat[1], std::max<size_t>(1, xs[1].pr.meta.n_rows) * sizeof(K)));
    HJ_TRY(ensure(ctx, ctx->valid_mask[1], std::max<size_t>(1, xs[1].pr.mask_words) * 4));
  }
  std::vector<uint32_t> t0(S + 1), Gs(S), coff(S + 1, 0);
  for (uint32_t q = 0; q <= S; ++q) t0[q] = uint32_t(uint64_t(T) * q / S);
  for (uint32_t q = 0; q < S; ++q) {
    Gs[q] = std::max<uint32_t>(1, std::min<uint32_t>(t0[q + 1] - t0[q], uint32_t(kVirtualCtasPerSm) * ctx->n_sms));
    coff[q + 1] = coff[q] + Gs[q] * kMaxRanks;
  }
  HJ_TRY(ensure(ctx, ctx->xchg_counts[1], size_t(coff[S]) * 4));
  HJ_CUDA(ctx, cudaMemsetAsync(ctx->xchg_counts[1].p, 0, size_t(coff[S]) * 4, cs));
  auto slice_args = [&](uint32_t q) {
    XchgArgs a = xchg_args<K>(ctx, xs[1], 1);
    a.counts = static_cast<uint32_t*>(ctx->xchg_counts[1].p) + coff[q];
    a.send_cnt = cnt_dev + size_t(1 + q) * kMaxRanks;
    a.tile_begin = t0[q];
    a.tile_end = t0[q + 1];
    return a;
  };
  for (uint32_t q = 0; q < S; ++q) {
    XchgArgs a = slice_args(q);
    if (a.tile_end > a.tile_begin) {
      k_xchg_count<K, kXchgThreads, xchg_items<K>()><<<Gs[q], kXchgThreads, 0, cs>>>(a);
      HJ_LAUNCH_CHECK(ctx, "k_xchg_count");
      ++n_launch;
    }
    k_xchg_offsets<256><<<1, 256, 0, cs>>>(a.counts, a.send_cnt, R, Gs[q]);
    HJ_LAUNCH_CHECK(ctx, "k_xchg_offsets");
    ++n_launch;
  }
  HJ_NCCL(ctx, nc->AllGather(cnt_dev, mat_dev, size_t(n_slots) * kMaxRanks, ncclUint64, comm, cs));
  auto* mat = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedCounts);
  if (size_t(R) * n_slots * kMaxRanks * 8 > kPinnedHandles - kPinnedCounts)
    return fail(ctx, HJ_ERR_INVALID, "%u probe slices x %u ranks exceed the count readback area", S, R);
  HJ_CUDA(ctx, cudaMemcpyAsync(mat, mat_dev, size_t(R) * n_slots * kMaxRanks * 8, cudaMemcpyDeviceToHost, cs));
  HJ_CUDA(ctx, cudaStreamSynchronize(cs));
  auto cnt = [&](uint32_t src, uint32_t slot, uint32_t dst) { return mat[(size_t(src) * n_slots + slot) * kMaxRanks + dst]; };

  for (void* p : ctx->xchg_graveyard) cudaFree(p);  // every rank has entered this join: nobody maps the outgrown buffers
  ctx->xchg_graveyard.clear();

  // ---- receive sets: 0 = build rows, 1/2 = probe slices (alternating), sized for the largest slice they will hold
  std::vector<uint64_t> n_recv(n_slots, 0);
  uint64_t set_rows[3] = {0, 0, 0}, probe_recv_total = 0;
  for (uint32_t slot = 0; slot < n_slots; ++slot) {
    for (uint32_t p = 0; p < R; ++p) n_recv[slot] += cnt(p, slot, me);
    if (n_recv[slot] >= 0xFFFFFFFFull)
      return fail(ctx, HJ_ERR_TOO_LARGE, "rank %u would receive %llu rows in one piece (limit 2^32-2 per GPU)", me,
                  (unsigned long long)n_recv[slot]);
    const int set = slot == 0 ? 0 : 1 + int((slot - 1) & 1u);
    set_rows[set] = std::max(set_rows[set], n_recv[slot]);
    if (slot) probe_recv_total += n_recv[slot];
  }
  for (int set = 0; set < 3; ++set) {
    HJ_TRY(ensure_exported(ctx, ctx->xchg_recv_keys[set], std::max<uint64_t>(1, set_rows[set]) * sizeof(K)));
    HJ_TRY(ensure_exported(ctx, ctx->xchg_recv_rids[set], std::max<uint64_t>(1, set_rows[set]) * 8));
    if (set && xs[1].send_nulls) HJ_TRY(ensure_exported(ctx, ctx->xchg_recv_nulls[set], std::max<uint64_t>(1, set_rows[set])));
  }
  HJ_TRY(exchange_handles(ctx, cs));

  uint64_t bytes_sent = 0;
  // X3 of one slot into receive set `set` of every rank, then the barrier
  auto scatter_slot = [&](uint32_t slot, int side, int set, XchgArgs a, uint32_t G) -> int {
    const bool nulls = xs[side].send_nulls;
    uint64_t send_off = 0;
    for (uint32_t d = 0; d < R; ++d) {
      uint64_t base_in_d = 0;  // where my rows start in rank d's receive set: after the rows of the ranks before me
      for (uint32_t s2 = 0; s2 < me; ++s2) base_in_d += cnt(s2, slot, d);
      a.adj[d] = static_cast<uint32_t>(base_in_d - send_off);
      send_off += cnt(me, slot, d);
      if (d == me) {
        a.dst_keys[d] = ctx->xchg_recv_keys[set].p;
        a.dst_rids[d] = static_cast<uint64_t*>(ctx->xchg_recv_rids[set].p);
        a.dst_nulls[d] = nulls ? static_cast<uint8_t*>(ctx->xchg_recv_nulls[set].p) : nullptr;
      } else {
        a.dst_keys[d] = ctx->xchg_peer[d][set * 3 + 0].ptr;
        a.dst_rids[d] = static_cast<uint64_t*>(ctx->xchg_peer[d][set * 3 + 1].ptr);
        a.dst_nulls[d] = nulls ? static_cast<uint8_t*>(ctx->xchg_peer[d][set * 3 + 2].ptr) : nullptr;
        if (cnt(me, slot, d) && (!a.dst_keys[d] || !a.dst_rids[d] || (nulls && !a.dst_nulls[d])))
          return fail(ctx, HJ_ERR_COMM, "rank %u has no mapping of rank %u's receive set %d", me, d, set);
      }
    }
    if (a.tile_end > a.tile_begin) {
      auto k = k_xchg_scatter<K, kXchgThreads, xchg_items<K>()>;
      const size_t smem = size_t(xchg_tile<K>()) * (sizeof(K) + 8 + 2);
      HJ_TRY(set_smem(ctx, k, smem));
      k<<<G, kXchgThreads, smem, cs>>>(a);
      HJ_LAUNCH_CHECK(ctx, "k_xchg_scatter");
      ++n_launch;
    }
    bytes_sent += (send_off - cnt(me, slot, me)) * (sizeof(K) + 8 + (nulls ? 1 : 0));
    HJ_NCCL(ctx, nc->AllGather(cnt_dev, mat_dev, 1, ncclUint64, comm, cs));  // barrier
    return HJ_OK;
  };

  // ---- build rows: exchange, then partition them locally (per-side set 2)
  HJ_TRY(scatter_slot(0, 0, 0, xchg_args<K>(ctx, xs[0], 0), xs[0].G));
  HJ_CUDA(ctx, cudaEventRecord(ev[10], cs));
  HJ_TRY(scatter_slot(1, 1, 1, slice_args(0), Gs[0]));  // first probe slice right behind it
  HJ_CUDA(ctx, cudaEventRecord(ctx->ev_x[0], cs));

  auto recv_relation = [&](int set, uint64_t rows, bool nulls, hj_segment* seg, hj_relation* rel) {
    std::memset(seg, 0, sizeof(hj_segment));
    seg->n_rows = static_cast<uint32_t>(rows);
    seg->kind = HJ_SEG_VALUE;
    seg->key_type = sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32;
    seg->values = ctx->xchg_recv_keys[set].p;
    seg->nulls = nulls ? static_cast<const uint8_t*>(ctx->xchg_recv_nulls[set].p) : nullptr;
    rel->segs = seg;
    rel->n_segs = 1;
    rel->location = HJ_LOC_DEVICE;
    rel->n_rows = rows;
  };
  struct RowidGuard {  // prepare_relation marks a relation "exchanged" through ctx->xchg_rowids[side]
    hj_ctx* c;
    ~RowidGuard() { for (auto& r : c->xchg_rowids) r = nullptr; }
  } guard{ctx};

  const int radix_bits = ctx->opts.radix_bits >= 0 ? std::min(ctx->opts.radix_bits, kMaxRadixBits)
                                                   : radix_bits_for(n_recv[0], sizeof(K) == 8 ? HJ_KEY_INT64 : HJ_KEY_INT32);
  const int bits2 = radix_bits > kSinglePassMaxBits ? radix_bits / 2 : 0;
  Launches nl;
  hj_segment seg_b, seg_p[2];
  hj_relation rel_b, rel_p[2];
  PreparedRel pb, pp[2];
  PartState sb, sp[2];
  HJ_CUDA(ctx, cudaStreamWaitEvent(st, ev[10], 0));
  recv_relation(0, n_recv[0], false, &seg_b, &rel_b);
  ctx->xchg_rowids[2] = static_cast<const uint64_t*>(ctx->xchg_recv_rids[0].p);
  HJ_TRY(prepare_relation(ctx, &rel_b, 2, tile1<K>(), st, &pb));
  HJ_TRY((stage_alloc<K, PAY>(ctx, pb, 2, radix_bits, bits2, &sb)));
  HJ_CUDA(ctx, cudaEventRecord(ev[12], st));
  HJ_TRY(stage_hist<K>(ctx, pb, 2, radix_bits, bits2, kNullDrop, sb, st, &nl));
  HJ_CUDA(ctx, cudaEventRecord(ev[13], st));
  HJ_TRY((stage_pass1<K, PAY>(ctx, pb, 2, radix_bits, bits2, kNullDrop, sb, st, &nl)));
  HJ_CUDA(ctx, cudaEventRecord(ev[14], st));
  HJ_TRY((stage_pass2<K, PAY>(ctx, pb, 2, radix_bits, bits2, kNullDrop, sb, st, &nl)));
  HJ_CUDA(ctx, cudaEventRecord(ev[15], st));
  float ph_ms[HJ_N_PHASES] = {};  // local-join phases, summed over the build side and all slices (they run while the next
                                  // slice is being exchanged, so they include that interference)

  // ---- probe slices: exchange slice q+1 while slice q is joined
  HJ_TRY(ensure_preserve(ctx, ctx->acc_probe, std::max<uint64_t>(1, probe_recv_total) * 8, 0, st));
  if (!one_side) HJ_TRY(ensure_preserve(ctx, ctx->acc_build, std::max<uint64_t>(1, probe_recv_total) * 8, 0, st));
  auto* rb = static_cast<Readback*>(ctx->pinned);
  uint64_t total = 0, probe_tuples = 0, build_tuples = 0;
  uint32_t overflow_parts = 0;
  int max_passes = 0;
  for (uint32_t q = 0; q < S; ++q) {
    const int cur = q & 1, nxt = (q + 1) & 1;
    if (q + 1 < S) {
      if (q >= 1) {
        // receive set `nxt` is about to be overwritten on EVERY rank: each rank enters this barrier only after its own
        // join of slice q-1 (the last reader of that set) has finished
        HJ_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_done[nxt], 0));
        HJ_NCCL(ctx, nc->AllGather(cnt_dev, mat_dev, 1, ncclUint64, comm, cs));
      }
      HJ_TRY(scatter_slot(q + 2, 1, 1 + nxt, slice_args(q + 1), Gs[q + 1]));
      HJ_CUDA(ctx, cudaEventRecord(ctx->ev_x[nxt], cs));
    }
    if (q + 1 == S) HJ_CUDA(ctx, cudaEventRecord(ev[11], cs));
    HJ_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_x[cur], 0));
    const int side = 3 + cur;
    recv_relation(1 + cur, n_recv[1 + q], xs[1].send_nulls, &seg_p[cur], &rel_p[cur]);
    ctx->xchg_rowids[side] = static_cast<const uint64_t*>(ctx->xchg_recv_rids[1 + cur].p);
    pp[cur] = PreparedRel{};
    HJ_TRY(prepare_relation(ctx, &rel_p[cur], side, tile1<K>(), st, &pp[cur]));
    HJ_TRY((stage_alloc<K, PAY>(ctx, pp[cur], side, radix_bits, bits2, &sp[cur])));
    HJ_CUDA(ctx, cudaEventRecord(ev[1], st));
    HJ_TRY(stage_hist<K>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl));
    HJ_CUDA(ctx, cudaEventRecord(ev[2], st));
    HJ_TRY((stage_pass1<K, PAY>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl)));
    HJ_CUDA(ctx, cudaEventRecord(ev[3], st));
    HJ_TRY((stage_pass2<K, PAY>(ctx, pp[cur], side, radix_bits, bits2, probe_null_policy, sp[cur], st, &nl)));
    HJ_CUDA(ctx, cudaEventRecord(ev[4], st));
    int passes = 0;
    HJ_TRY((probe_phase<K, PAY>(ctx, pb, pp[cur], sp[cur], 2, side, cur, radix_bits, bits2, mode, probe_null_policy, true,
                                n_recv[0], n_recv[1 + q], st, &nl, &passes, ev[5], ev[6])));
    if (q == 0) {
      ph_ms[HJ_PH_MATERIALIZE] += ev_ms(ev[12], ev[13]);
      ph_ms[HJ_PH_SCATTER1] += ev_ms(ev[13], ev[14]);
      ph_ms[HJ_PH_SCATTER2] += ev_ms(ev[14], ev[15]);
    }
    ph_ms[HJ_PH_MATERIALIZE] += ev_ms(ev[1], ev[2]);
    ph_ms[HJ_PH_SCATTER1] += ev_ms(ev[2], ev[3]);
    ph_ms[HJ_PH_SCATTER2] += ev_ms(ev[3], ev[4]);
    if (passes == 1) {
      ph_ms[HJ_PH_BUILD_PROBE] += ev_ms(ev[4], ev[5]);
      ph_ms[HJ_PH_FINALIZE] += ev_ms(ev[5], ev[6]);
    }
    const uint64_t rows = rb->out.n_rows;  // probe_phase synchronised `st`
    build_tuples = rb->n_build_tuples;
    probe_tuples += rb->n_probe_tuples;
    overflow_parts = std::max(overflow_parts, rb->plan.n_overflow_parts);
    max_passes = std::max(max_passes, passes);
    if (rows) {
      HJ_TRY(ensure_preserve(ctx, ctx->acc_probe, (total + rows) * 8, total * 8, st));
      HJ_CUDA(ctx, cudaMemcpyAsync(static_cast<char*>(ctx->acc_probe.p) + total * 8, ctx->out_probe[cur].p, rows * 8,
                                   cudaMemcpyDeviceToDevice, st));
      if (!one_side) {
        HJ_TRY(ensure_preserve(ctx, ctx->acc_build, (total + rows) * 8, total * 8, st));
        HJ_CUDA(ctx, cudaMemcpyAsync(static_cast<char*>(ctx->acc_build.p) + total * 8, ctx->out_build[cur].p, rows * 8,
                                     cudaMemcpyDeviceToDevice, st));
      }
    }
    HJ_CUDA(ctx, cudaEventRecord(ctx->ev_done[cur], st));
    total += rows;
  }
  HJ_CUDA(ctx, cudaEventRecord(ev[6], st));
  HJ_CUDA(ctx, cudaStreamSynchronize(cs));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));

  ctx->has_result = true;
  ctx->res_build = ctx->acc_build.p;
  ctx->res_probe = ctx->acc_probe.p;
  ctx->n_result = total;
  ctx->n_sides = one_side ? 1 : 2;
  info->n_rows = total;
  info->n_sides = ctx->n_sides;
  info->radix_bits = radix_bits;
  info->n_build_tuples = build_tuples;
  info->n_probe_tuples = probe_tuples;
  info->n_partitions = 1u << radix_bits;
  info->n_overflow_partitions = overflow_parts;
  info->n_probe_passes = max_passes;
  info->n_probe_slices = S;
  hj_timings& t = ctx->timings;
  std::memset(&t, 0, sizeof(t));
  for (int ph : {HJ_PH_MATERIALIZE, HJ_PH_SCATTER1, HJ_PH_SCATTER2, HJ_PH_BUILD_PROBE, HJ_PH_FINALIZE}) t.ms[ph] = ph_ms[ph];
  t.ms[HJ_PH_XCHG_PARTITION] = ev_ms(ev[9], ev[10]);  // shard prepare + all X1/X2 + build X3
  t.ms[HJ_PH_XCHG_TRANSFER] = ev_ms(ev[10], ev[11]);  // probe slices' X3 (overlaps the local joins of the earlier slices)
  t.ms[HJ_PH_TOTAL] = ev_ms(ev[9], ev[6]);
  t.n_kernel_launches = nl.n + n_launch;
  t.xchg_path = 1;
  t.bytes_nvlink_sent = bytes_sent;
  return HJ_OK;
}

}  // namespace

#include "hj_xpush.cuh"  // the push exchange: uses the IPC plumbing above, dispatched by hj_join_sharded below

extern "C" {

uint32_t hj_exchange_rank_of(int64_t key, int key_type, uint32_t n_ranks) {
  if (n_ranks == 0) return 0;
  const uint32_t h = key_type == HJ_KEY_INT64 ? hash_key(static_cast<int64_t>(key)) : hash_key(static_cast<int32_t>(key));
  return dest_rank_of_hash(h, n_ranks);
}

int hj_comm_unique_id(uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES]) {
  NcclApi* nc = nccl_api();
  if (!nc->handle) return fail(nullptr, HJ_ERR_COMM, "%s", nc->error.c_str());
  ncclUniqueId uid;
  if (nc->GetUniqueId(&uid) != ncclSuccess) return fail(nullptr, HJ_ERR_COMM, "ncclGetUniqueId failed");
  std::memcpy(id, &uid, HJ_NCCL_UNIQUE_ID_BYTES);
  return HJ_OK;
}

int hj_comm_init(hj_ctx* ctx, const uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES], int n_ranks, int rank) {
  if (!ctx) return HJ_ERR_INVALID;
  if (n_ranks < 1 || n_ranks > kMaxRanks || rank < 0 || rank >= n_ranks)
    return fail(ctx, HJ_ERR_INVALID, "hj_comm_init: rank %d of %d (1..%d ranks supported)", rank, n_ranks, kMaxRanks);
  HJ_TRY(check_ctx(ctx));
  NcclApi* nc = nccl_api();
  if (!nc->handle) return fail(ctx, HJ_ERR_COMM, "%s", nc->error.c_str());
  if (ctx->nccl_comm) {
    nc->CommDestroy(static_cast<ncclComm_t>(ctx->nccl_comm));
    ctx->nccl_comm = nullptr;
  }
  ncclUniqueId uid;
  std::memcpy(&uid, id, HJ_NCCL_UNIQUE_ID_BYTES);
  ncclComm_t comm = nullptr;
  HJ_NCCL(ctx, nc->CommInitRank(&comm, n_ranks, uid, rank));
  ctx->nccl_comm = comm;
  ctx->n_ranks = n_ranks;
  ctx->rank = rank;
  // one rank: the exchange degenerates to stores into the rank's own receive buffers — kept working because it is how the
  // exchange kernels are profiled on a single GPU (ncu follows one process; tests/profile_join.py --sharded)
  ctx->p2p_ok = n_ranks == 1 || probe_p2p(ctx);
  return HJ_OK;
}

int hj_join_sharded(hj_ctx* ctx, const hj_relation* build_shard, const hj_relation* probe_shard, int mode,
                    void* cuda_stream, hj_result_info* out_info) {
  if (!ctx) return HJ_ERR_INVALID;
  ctx->has_result = false;
  const bool bound = ctx->bound;  // hj_bind_output: consumed by this join whatever happens (see hj_join)
  ctx->bound = false;
  ctx->result_in_bound = false;
  if (!out_info) return fail(ctx, HJ_ERR_INVALID, "out_info is NULL");
  std::memset(out_info, 0, sizeof(*out_info));
  if (!(mode == HJ_MODE_INNER || mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT || mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI))
    return fail(ctx, HJ_ERR_INVALID, "JoinMode %d is not supported by JoinHash (Inner/Left/Right/Semi/Anti only)", mode);
  if (!ctx->nccl_comm) return fail(ctx, HJ_ERR_COMM, "hj_join_sharded without hj_comm_init");
  const bool one_side_mode = mode == HJ_MODE_SEMI || mode == HJ_MODE_ANTI;
  if (bound && (!ctx->bound_probe || (!one_side_mode && !ctx->bound_build))) return fail(ctx, HJ_ERR_INVALID, "hj_bind_output: NULL output buffer");
  // the join's result into the bound host buffers (joins that were not streamed)
  auto deliver_bound = [&]() -> int {
    if (!bound) return HJ_OK;
    if (out_info->n_rows > ctx->bound_capacity)
      return fail(ctx, HJ_ERR_CAPACITY, "bound output holds %llu rows, this rank's result has %llu (still on the device: hj_fetch it)",
                  (unsigned long long)ctx->bound_capacity, (unsigned long long)out_info->n_rows);
    HJ_TRY(hj_fetch(ctx, one_side_mode ? nullptr : ctx->bound_build, ctx->bound_probe, ctx->bound_capacity));
    ctx->result_in_bound = true;
    return HJ_OK;
  };
  HJ_TRY(validate_relation(ctx, build_shard, "build"));
  HJ_TRY(validate_relation(ctx, probe_shard, "probe"));
  HJ_TRY(check_ctx(ctx));
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  if (hashed_type_of(build_shard, probe_shard) >= HJ_KEY_FLOAT)
    return fail(ctx, HJ_ERR_INVALID, "hj_join_sharded: float/double keys are joined by hj_join only");
  RefRewrite rw[2];  // REFERENCE / FOR / RLE segments -> gathered or decoded on the device (N1, N2)
  HJ_TRY(gather_both(ctx, build_shard, probe_shard, st, rw, hashed_type_of(build_shard, probe_shard)));
  // One all-gather tells every rank what the others hold: the hashed key type must be the same everywhere (the destination
  // of a key depends on it), the probe rows decide about slicing, and the shard geometry about the push exchange.
  unsigned long long wide = 0;
  for (const hj_relation* r : {build_shard, probe_shard})
    for (uint32_t i = 0; i < r->n_segs; ++i) wide |= r->segs[i].key_type == HJ_KEY_INT64 ? 1ull : 0ull;
  HJ_TRY(ensure(ctx, ctx->xchg_cnt_dev, size_t(2) * kMaxRanks * 8 * (1 + kMaxRanks)));
  auto* dev = static_cast<unsigned long long*>(ctx->xchg_cnt_dev.p);
  auto* host = reinterpret_cast<unsigned long long*>(static_cast<char*>(ctx->pinned) + kPinnedCounts);
  NcclApi* nc = nccl_api();
  const bool direct = ctx->p2p_ok && !(ctx->opts.flags & HJ_FLAG_XCHG_STAGED);
  const bool outer_mode = mode == HJ_MODE_LEFT || mode == HJ_MODE_RIGHT;
  const ShardGeometry gb = shard_geometry(build_shard), gp = shard_geometry(probe_shard);
  const bool push_here = direct && gb.regular && gp.regular && !(ctx->opts.flags & (HJ_FLAG_XCHG_GENERAL | HJ_FLAG_TWO_SWEEP_PARTITION)) &&
                         !(outer_mode && (ctx->opts.flags & HJ_FLAG_SQL_NULL_PROBE)) && (bound || !ctx->opts.stream_slice_rows);
  unsigned long long word[kXpDescWords] = {build_shard->n_rows, probe_shard->n_rows,
                                           gb.first_chunk | (static_cast<unsigned long long>(static_cast<uint32_t>(gb.shift)) << 32),
                                           gp.first_chunk | (static_cast<unsigned long long>(static_cast<uint32_t>(gp.shift)) << 32),
                                           wide | (push_here ? 2ull : 0ull) | ((bound && probe_shard->location == HJ_LOC_HOST) ? 4ull : 0ull),
                                           probe_shard->n_segs ? probe_shard->segs[0].n_rows : 0ull, 0, 0};
  HJ_CUDA(ctx, cudaMemcpyAsync(dev, word, sizeof(word), cudaMemcpyHostToDevice, st));
  HJ_NCCL(ctx, nc->AllGather(dev, dev + kXpDescWords, kXpDescWords, ncclUint64, static_cast<ncclComm_t>(ctx->nccl_comm), st));
  std::vector<unsigned long long> desc(size_t(ctx->n_ranks) * kXpDescWords);
  HJ_CUDA(ctx, cudaMemcpyAsync(host, dev + kXpDescWords, desc.size() * 8, cudaMemcpyDeviceToHost, st));
  HJ_CUDA(ctx, cudaStreamSynchronize(st));
  std::memcpy(desc.data(), host, desc.size() * 8);  // the pinned area is reused by the joins below
  bool push = true, streamable = true;
  uint64_t max_probe_rows = 0, total_build = 0;
  for (int r = 0; r < ctx->n_ranks; ++r) {
    wide |= desc[size_t(r) * kXpDescWords + 4] & 1ull;
    push = push && (desc[size_t(r) * kXpDescWords + 4] & 2ull);
    streamable = streamable && (desc[size_t(r) * kXpDescWords + 4] & 4ull);
    max_probe_rows = std::max<uint64_t>(max_probe_rows, desc[size_t(r) * kXpDescWords + 1]);
    total_build += desc[size_t(r) * kXpDescWords + 0];
  }
  // Push exchange (hj_xpush.cuh): needs two partitioning passes in the local join — decided from the build rows a rank
  // expects, the same number on every rank
  const int xp_bits = ctx->opts.radix_bits >= 0 ? std::min(ctx->opts.radix_bits, kMaxRadixBits)
                                                : radix_bits_for(total_build / ctx->n_ranks + 1, wide ? HJ_KEY_INT64 : HJ_KEY_INT32);
  push = push && pass2_bits(xp_bits, wide) > 0 && (1 << (xp_bits - pass2_bits(xp_bits, wide))) >= ctx->n_ranks;
  if (push) {
    bool fell_back = false;
    // every rank bound host outputs for host-resident probe shards and the shards are long enough: streamed (copy-in, exchange
    // + join, copy-out of probe slices overlap), with the same number of slices on every rank
    const uint64_t min_slice = ctx->opts.stream_slice_rows ? ctx->opts.stream_slice_rows : kMinSliceRows;
    const uint32_t n_slices = streamable ? uint32_t(std::min<uint64_t>(24, max_probe_rows / min_slice)) : 0u;
    int st_push;
    if (n_slices >= 2)
      st_push = wide ? run_join_sharded_push_streamed<int64_t>(ctx, build_shard, probe_shard, mode, st, out_info, desc.data(), xp_bits, n_slices, &fell_back)
                     : run_join_sharded_push_streamed<int32_t>(ctx, build_shard, probe_shard, mode, st, out_info, desc.data(), xp_bits, n_slices, &fell_back);
    else
      st_push = wide ? run_join_sharded_push<int64_t>(ctx, build_shard, probe_shard, mode, st, out_info, desc.data(), xp_bits, &fell_back)
                     : run_join_sharded_push<int32_t>(ctx, build_shard, probe_shard, mode, st, out_info, desc.data(), xp_bits, &fell_back);
    if (st_push != HJ_OK) return st_push;
    if (!fell_back) {
      if (n_slices < 2) HJ_TRY(deliver_bound());
      return check_reference_flags(ctx, rw);
    }
    // a receive region overflowed on some rank (skew): every rank saw the flag and repeats the join on the general path
  }
  // Probe slices (same S on every rank): opt-in through hj_options.stream_slice_rows. Slicing bounds the receive buffers
  // to two slices, but it is NOT faster: measured on 2 x B200, C2 per rank, unsliced 54.8 ms, S=2 60.0 ms, S=4 66.4 ms —
  // the persistent X3 CTAs hold their SMs, so the 1024-thread join kernels of the previous slice wait instead of
  // overlapping, and every slice rebuilds the build-side hash tables.
  const uint32_t S = ctx->opts.stream_slice_rows
                         ? uint32_t(std::min<uint64_t>(kMaxXchgSlices, max_probe_rows / ctx->opts.stream_slice_rows)) : 1u;
  int status;
  if (direct && S >= 2)
    status = wide ? run_join_sharded_sliced<int64_t>(ctx, build_shard, probe_shard, mode, st, out_info, S)
                  : run_join_sharded_sliced<int32_t>(ctx, build_shard, probe_shard, mode, st, out_info, S);
  else
    status = wide ? run_join_sharded<int64_t>(ctx, build_shard, probe_shard, mode, st, out_info)
                  : run_join_sharded<int32_t>(ctx, build_shard, probe_shard, mode, st, out_info);
  if (status != HJ_OK) return status;
  HJ_TRY(deliver_bound());
  return check_reference_flags(ctx, rw);
}

}  // extern "C"

extern "C" void hj_comm_release(hj_ctx* ctx) {
  if (ctx && ctx->nccl_comm) {
    close_peer_maps(ctx);
    for (void* p : ctx->xchg_graveyard) cudaFree(p);
    ctx->xchg_graveyard.clear();
    ctx->p2p_ok = false;
    nccl_api()->CommDestroy(static_cast<ncclComm_t>(ctx->nccl_comm));
    ctx->nccl_comm = nullptr;
    ctx->n_ranks = 1;
    ctx->rank = 0;
  }
}
