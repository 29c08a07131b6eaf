// Microbenchmark (development tool): throughput of warp-level digit ranking variants
// used by the radix-sort scatter kernel.  nvcc -arch=sm_100a -O3 -o rank rank.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr int BITS = 10, kBins = 1 << BITS, THREADS = 512, WARPS = THREADS / 32;

__device__ __forceinline__ uint32_t rank_tag(volatile uint16_t* wh, volatile uint32_t* peer_tbl,
                                             uint32_t d, uint32_t lane, uint32_t lt_mask) {
  const uint32_t c = (uint32_t)wh[d] >> 5;
  peer_tbl[lane] = 0u;
  __syncwarp();
  wh[d] = (uint16_t)(((c + 1u) << 5) | lane);
  __syncwarp();
  const uint32_t winner = (uint32_t)wh[d] & 31u;
  uint32_t peers;
  if (__all_sync(kFull, winner == __shfl_sync(kFull, winner, 0))) {
    peers = kFull;
  } else {
    atomicOr(const_cast<uint32_t*>(peer_tbl) + winner, 1u << lane);
    __syncwarp();
    peers = peer_tbl[winner];
  }
  if ((peers & lt_mask) == 0u) wh[d] = (uint16_t)((c + __popc(peers)) << 5);
  __syncwarp();
  return c + __popc(peers & lt_mask);
}
__device__ __forceinline__ uint32_t rank_match(volatile uint16_t* wh, uint32_t d, uint32_t lt_mask) {
  const uint32_t peers = __match_any_sync(kFull, d);
  const uint32_t c = (uint32_t)wh[d] >> 5;
  __syncwarp();
  if ((peers & lt_mask) == 0u) wh[d] = (uint16_t)((c + __popc(peers)) << 5);
  __syncwarp();
  return c + __popc(peers & lt_mask);
}
__device__ __forceinline__ uint32_t rank_ballot(volatile uint16_t* wh, uint32_t d, uint32_t lt_mask) {
  uint32_t peers = kFull;
#pragma unroll
  for (int b = 0; b < BITS; ++b) {
    const uint32_t m = __ballot_sync(kFull, (d >> b) & 1u);
    peers &= ((d >> b) & 1u) ? m : ~m;
  }
  const uint32_t c = (uint32_t)wh[d] >> 5;
  __syncwarp();
  if ((peers & lt_mask) == 0u) wh[d] = (uint16_t)((c + __popc(peers)) << 5);
  __syncwarp();
  return c + __popc(peers & lt_mask);
}

template <int V>
__global__ void __launch_bounds__(THREADS, 2) k(uint32_t* out, int iters, uint32_t mask) {
  __shared__ uint16_t wh[WARPS][kBins];
  __shared__ uint32_t ptbl[WARPS][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  for (int b = lane; b < kBins; b += 32) wh[warp][b] = 0;
  __syncwarp();
  uint32_t x = (blockIdx.x * THREADS + threadIdx.x) * 2654435761u + 12345u, acc = 0;
  for (int i = 0; i < iters; ++i) {
    x = x * 1664525u + 1013904223u;
    const uint32_t d = (x >> 12) & mask;
    if (V == 0) acc += rank_tag(wh[warp], ptbl[warp], d, lane, lt);
    if (V == 1) acc += rank_match(wh[warp], d, lt);
    if (V == 2) acc += rank_ballot(wh[warp], d, lt);
    if ((i & 31) == 31) {  // keep the 11-bit counters from overflowing
      for (int b = lane; b < kBins; b += 32) wh[warp][b] = 0;
      __syncwarp();
    }
  }
  out[blockIdx.x * THREADS + threadIdx.x] = acc;
}

template <int V>
void run(const char* name, uint32_t mask) {
  const int blocks = 148 * 2, iters = 4096;
  uint32_t* out;
  cudaMalloc(&out, blocks * THREADS * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<V><<<blocks, THREADS>>>(out, iters, mask);
  cudaEventRecord(e0);
  k<V><<<blocks, THREADS>>>(out, iters, mask);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double keys = (double)blocks * THREADS * iters;
  printf("%-8s mask=%4x  %.3f ms  %.1f Gkeys/s  (%s)\n", name, mask, ms, keys / ms / 1e6,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}
int main() {
  for (uint32_t mask : {0x3FFu, 0x3Fu, 0x3u, 0x0u}) {
    run<0>("tag", mask);
    run<1>("match", mask);
    run<2>("ballot", mask);
  }
  return 0;
}
