// karras.cu — A4: binary radix tree topology (Karras 2012), one thread per
// internal node.
//
// delta(i,j) = -1 outside [0,n); clz(k_i ^ k_j) when the keys differ; otherwise
// key_width + clz(i ^ j): duplicate Morton codes fall back to the sorted
// position, so the augmented keys are distinct and the tree is unique.
//
// Output: one parent word per node in canonical numbering (internal i in
// [0,n-1), leaf j at n-1+j): (parent << 2) | (local << 1) | slot (kernels.h).  The
// child ids themselves are written into the 64-byte nodes by the refit walk,
// which knows them (refit.cu), so this kernel stores 8 B per internal node.
//
// Bound: HBM/L2 latency; the probes of the two binary searches touch sorted
// neighbours, so most of them hit L1/L2.  ~4-8 B key read + 8 B written per tri.
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

// delta between sorted positions i and j = i + dir*off, all in 32-bit arithmetic
// (n < 2^30).  Returns -1 when j falls outside [0, n).
template <typename KeyT>
__device__ __forceinline__ int delta_off(const KeyT* __restrict__ keys, uint32_t n, uint32_t i,
                                         KeyT ki, int dir, uint32_t off) {
  uint32_t j;
  if (dir > 0) {
    if (off >= n - i) return -1;
    j = i + off;
  } else {
    if (off > i) return -1;
    j = i - off;
  }
  const KeyT kj = __ldg(keys + j);
  if (sizeof(KeyT) == 4) {
    const uint32_t x = (uint32_t)ki ^ (uint32_t)kj;
    return x ? __clz((int)x) : 32 + __clz((int)(i ^ j));
  } else {
    const uint64_t x = (uint64_t)ki ^ (uint64_t)kj;
    return x ? __clzll((long long)x) : 64 + __clz((int)(i ^ j));
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
karras_kernel(const KeyT* __restrict__ keys, uint32_t n, uint32_t* __restrict__ parent) {
  const uint32_t leaf0 = n - 1;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= leaf0) return;
  const KeyT ki = __ldg(keys + i);
  const int dir = (delta_off(keys, n, i, ki, 1, 1) - delta_off(keys, n, i, ki, -1, 1)) < 0 ? -1 : 1;
  const int dmin = delta_off(keys, n, i, ki, -dir, 1);
  uint32_t lmax = 2;
  while (delta_off(keys, n, i, ki, dir, lmax) > dmin) lmax <<= 1;
  uint32_t l = 0;
  for (uint32_t t = lmax >> 1; t >= 1; t >>= 1)
    if (delta_off(keys, n, i, ki, dir, l + t) > dmin) l += t;
  const uint32_t j = dir > 0 ? i + l : i - l;
  const int dnode = delta_off(keys, n, i, ki, dir, l);
  uint32_t s = 0, t = l;
  do {
    t = (t + 1) >> 1;
    if (delta_off(keys, n, i, ki, dir, s + t) > dnode) s += t;
  } while (t > 1);
  const uint32_t lo = min(i, j), hi = max(i, j);
  const uint32_t gamma = dir > 0 ? i + s : i - s - 1;  // i + s*dir + min(dir, 0)
  const uint32_t lc = (lo == gamma) ? leaf0 + gamma : gamma;
  const uint32_t rc = (hi == gamma + 1) ? leaf0 + gamma + 1 : gamma + 1;
  const uint32_t local = ((lo >> kRefitLeavesLog2) == (hi >> kRefitLeavesLog2)) ? 2u : 0u;
  parent[lc] = (i << 2) | local;
  parent[rc] = (i << 2) | local | 1u;
  if (i == 0) parent[0] = kNoParent;
}

__global__ void single_leaf_kernel(uint32_t* parent) { parent[0] = kNoParent; }

}  // namespace

int launch_karras(const void* keys, uint64_t n, uint32_t key_bits, uint32_t* parent,
                  cudaStream_t s) {
  if (n == 0) return 0;
  if (n == 1) {
    single_leaf_kernel<<<1, 1, 0, s>>>(parent);
    RT_LAUNCHED();
    return 0;
  }
  const uint64_t blocks = (n - 1 + 255) / 256;
  if (key_bits == 30)
    karras_kernel<uint32_t><<<(unsigned)blocks, 256, 0, s>>>((const uint32_t*)keys, (uint32_t)n,
                                                             parent);
  else
    karras_kernel<uint64_t><<<(unsigned)blocks, 256, 0, s>>>((const uint64_t*)keys, (uint32_t)n,
                                                             parent);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
