// Microbenchmark behind P1's dictionary gather: how fast can one B200 read 4-byte entries at random
// positions of an L2-resident 4 MB table (one C2 chunk dictionary), and does the path the request takes matter?
//   mode 0  LDG        ld.global.nc           (what k_part1 does)
//   mode 1  LDG.cg     ld.global.cg           (no L1 allocation)
//   mode 2  LDGSTS     cp.async.ca 4 B        (global -> shared without a register / scoreboard slot per row)
//   mode 3  UBLKCP     cp.async.bulk 16 B     (TMA engine: the aligned 16-byte block holding the entry, mbarrier-tracked)
//   mode 4  LDG pairs  two adjacent lanes read the same 128-byte line (what sorting the value ids would buy)
//   mode 5  LDG.64     8-byte entries (is the cost per row or per byte?)
//   mode 6  stream     control: only the value ids are read (what the coalesced sweep alone costs)
// The GPU is warmed up first (a fresh process starts at idle clocks; 1-ms kernels do not ramp them).
// Every mode reads its value ids coalesced from an index array (like the attribute vector) and reduces the entries to a
// checksum so nothing is optimised away. Reported: rows/ns, cycles per row and SM at the clock nvidia-smi reports.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

template <int MODE, int NT, int ITEMS>
__global__ void __launch_bounds__(NT) k_gather(const uint32_t* __restrict__ idx, const uint32_t* __restrict__ table,
                                               unsigned long long n, unsigned long long* __restrict__ out) {
  constexpr int TILE = NT * ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  uint32_t* stage = reinterpret_cast<uint32_t*>(s_raw);            // modes 2: TILE words; mode 3: TILE * 4 words
  __shared__ __align__(8) unsigned long long s_bar;
  unsigned long long acc = 0;
  uint32_t phase = 0;
  if (MODE == 3) {
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&s_bar)), "r"(1));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
  const unsigned long long n_tiles = n / TILE;
  for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const uint32_t* ip = idx + t * TILE;
    uint32_t vid[ITEMS];
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) vid[u] = __ldg(ip + u * NT + threadIdx.x);
    if (MODE == 0 || MODE == 4) {
      uint32_t v[ITEMS];
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        const uint32_t a = MODE == 4 ? ((vid[u] & ~31u) | (threadIdx.x & 31u)) : vid[u];  // mode 4: a warp shares one line
        v[u] = __ldg(table + a);
      }
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += v[u];
    } else if (MODE == 1) {
      uint32_t v[ITEMS];
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) v[u] = __ldcg(table + vid[u]);
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += v[u];
    } else if (MODE == 5) {
      unsigned long long v[ITEMS];
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) v[u] = __ldg(reinterpret_cast<const unsigned long long*>(table) + (vid[u] >> 1));
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += v[u];
    } else if (MODE == 6) {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += vid[u];
    } else if (MODE == 2) {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(stage + u * NT + threadIdx.x)), "l"(table + vid[u]));
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += stage[u * NT + threadIdx.x];  // own slots: no barrier needed
    } else if (MODE == 3) {
      if (threadIdx.x == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_bar)), "r"(TILE * 16) : "memory");
      __syncthreads();
#pragma unroll
      for (int u = 0; u < ITEMS; ++u)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 16, [%2];" ::"r"(
                         smem_u32(stage + (u * NT + threadIdx.x) * 4)),
                     "l"(table + (vid[u] & ~3u)), "r"(smem_u32(&s_bar))
                     : "memory");
      uint32_t done = 0;
      while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(smem_u32(&s_bar)), "r"(phase) : "memory");
      phase ^= 1u;
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) acc += stage[(u * NT + threadIdx.x) * 4 + (vid[u] & 3u)];
      __syncthreads();  // everyone has read the tile before the next expect_tx / overwrite
    }
  }
  // warp reduce, one atomic per warp
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

static double g_mhz = 1965.0;

template <int MODE, int NT, int ITEMS>
void run(const char* name, int ctas_per_sm, int carve, const uint32_t* idx, const uint32_t* table, unsigned long long n,
         unsigned long long* out, int sms, unsigned long long want) {
  constexpr int TILE = NT * ITEMS;
  const size_t smem = MODE == 2 ? TILE * 4 : MODE == 3 ? TILE * 16 : 0;
  auto kern = k_gather<MODE, NT, ITEMS>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (carve >= 0) CK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  const int grid = sms * ctas_per_sm;
  float best = 1e9f;
  unsigned long long got = 0;
  for (int rep = 0; rep < 6; ++rep) {
    CK(cudaMemset(out, 0, 8));
    CK(cudaEventRecord(a));
    kern<<<grid, NT, smem>>>(idx, table, n, out);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
    CK(cudaMemcpy(&got, out, 8, cudaMemcpyDeviceToHost));
  }
  const unsigned long long rows = n / TILE * TILE;
  const char* ok = (MODE == 4 || MODE == 5 || MODE == 6) ? "n/a" : (got == want ? "ok" : "MISMATCH");
  printf("%-10s NT=%4d ITEMS=%2d ctas/sm=%d carve=%3d  %7.3f ms  %6.2f rows/ns  %5.2f cyc/row/SM  sum %s\n", name, NT, ITEMS,
         ctas_per_sm, carve, best, rows / best / 1e6, best * 1e-3 * g_mhz * 1e6 * sms / rows, ok);
  fflush(stdout);
}

int main(int argc, char** argv) {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  if (argc > 1) g_mhz = atof(argv[1]);
  printf("%s, %d SMs, cycles at %.0f MHz\n", p.name, sms, g_mhz);
  const uint32_t table_n = 1u << 20;                    // 4 MB: one C2 chunk dictionary
  const unsigned long long n = 1ull << 28;              // 268 M rows (1 GiB of value ids)
  std::vector<uint32_t> h_table(table_n);
  for (uint32_t i = 0; i < table_n; ++i) h_table[i] = i * 2654435761u;
  uint32_t *idx, *table; unsigned long long* out;
  CK(cudaMalloc(&idx, n * 4)); CK(cudaMalloc(&table, table_n * 4)); CK(cudaMalloc(&out, 8));
  CK(cudaMemcpy(table, h_table.data(), table_n * 4, cudaMemcpyHostToDevice));
  // random value ids: xorshift on the host for a 16 M block, replicated with an offset pattern on the device side is
  // overkill: fill on the host in one go (1 GiB, a few seconds)
  {
    std::vector<uint32_t> h_idx(1u << 24);
    uint64_t s = 0x9E3779B97F4A7C15ull;
    unsigned long long want_block = 0;
    for (auto& v : h_idx) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; v = static_cast<uint32_t>(s >> 20) & (table_n - 1); }
    for (unsigned long long off = 0; off < n; off += h_idx.size()) CK(cudaMemcpy(idx + off, h_idx.data(), h_idx.size() * 4, cudaMemcpyHostToDevice));
    for (auto v : h_idx) want_block += h_table[v];
    unsigned long long want = want_block * (n / h_idx.size());
    // every mode's TILE divides 2^28 only if TILE is a power of two: all configurations below are
    for (int w = 0; w < 300; ++w) k_gather<0, 256, 8><<<sms * 8, 256>>>(idx, table, n, out);  // ~0.3 s of warm-up
    CK(cudaDeviceSynchronize());
    run<6, 256, 8>("stream", 8, -1, idx, table, n, out, sms, want);
    run<6, 1024, 8>("stream", 2, -1, idx, table, n, out, sms, want);
    run<0, 256, 8>("LDG", 4, -1, idx, table, n, out, sms, want);
    run<0, 256, 8>("LDG", 8, -1, idx, table, n, out, sms, want);
    run<0, 256, 16>("LDG", 4, -1, idx, table, n, out, sms, want);
    run<0, 512, 8>("LDG", 4, -1, idx, table, n, out, sms, want);
    run<0, 1024, 8>("LDG", 2, -1, idx, table, n, out, sms, want);
    run<0, 1024, 8>("LDG", 1, -1, idx, table, n, out, sms, want);
    run<0, 1024, 8>("LDG", 1, 50, idx, table, n, out, sms, want);    // L1 as k_part1 leaves it (about half)
    run<0, 1024, 8>("LDG", 1, 100, idx, table, n, out, sms, want);   // almost no L1
    run<0, 256, 8>("LDG", 8, 0, idx, table, n, out, sms, want);
    run<1, 256, 8>("LDG.cg", 8, -1, idx, table, n, out, sms, want);
    run<1, 1024, 8>("LDG.cg", 2, -1, idx, table, n, out, sms, want);
    run<2, 256, 8>("LDGSTS", 8, -1, idx, table, n, out, sms, want);
    run<2, 256, 16>("LDGSTS", 4, -1, idx, table, n, out, sms, want);
    run<2, 1024, 8>("LDGSTS", 2, -1, idx, table, n, out, sms, want);
    run<3, 256, 8>("UBLKCP16", 4, -1, idx, table, n, out, sms, want);
    run<3, 256, 8>("UBLKCP16", 6, -1, idx, table, n, out, sms, want);
    run<3, 128, 8>("UBLKCP16", 8, -1, idx, table, n, out, sms, want);
    run<4, 256, 8>("LDG-line", 8, -1, idx, table, n, out, sms, want);
    run<5, 256, 8>("LDG.64", 8, -1, idx, table, n, out, sms, want);
  }
  return 0;
}
