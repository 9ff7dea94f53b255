// Microbenchmarks that decide the ranking/histogram design of libhj's partitioning kernels:
// shared-memory atomics vs warp match (hardware match.any vs ballot loop) vs global reductions, per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu ; run on the B200 box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t h) { h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h; }

constexpr int NT = 512;
constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(NT) bench(uint32_t* out, uint32_t* gbins, uint32_t nbins_mask) {
  extern __shared__ uint32_t s[];
  for (uint32_t i = threadIdx.x; i <= nbins_mask; i += NT) s[i] = 0;
  __syncthreads();
  uint32_t x = blockIdx.x * NT + threadIdx.x + 1, acc = 0;
  const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
  uint32_t* wcnt = s + (threadIdx.x >> 5) * 256;  // warp-private counters (MODE 3,4)
#pragma unroll 4
  for (int it = 0; it < ITERS; ++it) {
    x = x * 1664525u + 1013904223u;
    const uint32_t d = mix(x) & nbins_mask;
    if (MODE == 0) acc += atomicAdd(&s[d], 1u);                       // smem atomic with return
    else if (MODE == 1) atomicAdd(&s[d], 1u);                         // smem atomic, result unused
    else if (MODE == 2) acc += atomicCAS(&s[d], 0u, x | 1u);          // smem CAS
    else if (MODE == 3) {                                              // hardware match.any + warp-private counter
      const uint32_t peers = __match_any_sync(0xffffffffu, d);
      const int leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if ((int)lane == leader) { base = wcnt[d]; wcnt[d] = base + __popc(peers); }
      base = __shfl_sync(0xffffffffu, base, leader);
      acc += base + __popc(peers & lt);
    } else if (MODE == 4) {                                            // ballot-loop match (7 bits) + warp-private counter
      uint32_t peers = 0xffffffffu;
#pragma unroll
      for (int b = 0; b < 7; ++b) {
        const uint32_t m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
        peers &= ((d >> b) & 1u) ? m : ~m;
      }
      const int leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if ((int)lane == leader) { base = wcnt[d]; wcnt[d] = base + __popc(peers); }
      base = __shfl_sync(0xffffffffu, base, leader);
      acc += base + __popc(peers & lt);
    } else if (MODE == 5) atomicAdd(&gbins[d], 1u);                   // global reduction (L2), result unused
    else if (MODE == 6) { acc += s[d]; s[(d + lane) & nbins_mask] = acc; }  // plain LDS + STS
    else if (MODE == 7) {                                              // match only (hardware), no counters
      acc += __match_any_sync(0xffffffffu, d);
    }
  }
  if (acc == 0xdeadbeefu) out[0] = acc;
  if (MODE == 1 || MODE == 5) { __syncthreads(); if (threadIdx.x == 0) out[blockIdx.x + 1] = s[1]; }
}

template <int MODE>
int run(const char* name, uint32_t nbins, int ctas_per_sm, uint32_t* out, uint32_t* gbins, int sms) {
  const size_t smem = (MODE == 3 || MODE == 4) ? (NT / 32) * 256 * 4 : (size_t)nbins * 4;
  CK(cudaFuncSetAttribute(bench<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  const int grid = sms * ctas_per_sm;
  bench<MODE><<<grid, NT, smem>>>(out, gbins, nbins - 1);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  bench<MODE><<<grid, NT, smem>>>(out, gbins, nbins - 1);
  CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
  float ms; CK(cudaEventElapsedTime(&ms, a, b));
  const double ops = (double)grid * NT * ITERS;
  printf("%-44s bins=%6u ctas/sm=%d  %8.1f Gops/s  %6.3f lane-ops/ns/SM  (%.3f ms)\n", name, nbins, ctas_per_sm, ops / ms / 1e6,
         ops / ms / 1e6 / sms, ms);
  return 0;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  printf("%s, %d SMs, %d MHz max\n", p.name, p.multiProcessorCount, p.clockRate / 1000);
  uint32_t *out, *gbins; CK(cudaMalloc(&out, 1 << 20)); CK(cudaMalloc(&gbins, 1 << 20)); CK(cudaMemset(gbins, 0, 1 << 20));
  const int sms = p.multiProcessorCount;
  for (int c : {1, 2, 4}) {
    run<0>("smem atomicAdd (returning)", 128, c, out, gbins, sms);
    run<0>("smem atomicAdd (returning)", 16384, c, out, gbins, sms);
    run<1>("smem atomicAdd (no return)", 128, c, out, gbins, sms);
    run<1>("smem atomicAdd (no return)", 16384, c, out, gbins, sms);
    run<2>("smem atomicCAS", 16384, c, out, gbins, sms);
    run<3>("match.any + warp-private counters", 128, c, out, gbins, sms);
    run<4>("ballot-loop match (7b) + warp-private counters", 128, c, out, gbins, sms);
    run<7>("match.any only", 128, c, out, gbins, sms);
    run<5>("global atomicAdd (no return)", 16384, c, out, gbins, sms);
    run<5>("global atomicAdd (no return)", 128, c, out, gbins, sms);
    run<6>("plain LDS + STS", 16384, c, out, gbins, sms);
  }
  return 0;
}
