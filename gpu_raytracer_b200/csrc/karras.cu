// karras.cu — A4: binary radix tree topology (Karras 2012), one thread per
// internal node.
//
// delta(i,j) = -1 outside [0,n); clz(k_i ^ k_j) when the keys differ; otherwise
// key_width + clz(i ^ j): duplicate Morton codes fall back to the sorted
// position, so the augmented keys are distinct and the tree is unique.
//
// Output: one parent word per node in canonical numbering (internal i in
// [0,n-1), leaf j at n-1+j): (parent << 1) | slot, slot 0 = left child.  The
// child ids themselves are written into the 64-byte nodes by the refit walk,
// which knows them (refit.cu), so this kernel stores 8 B per internal node.
//
// Bound: HBM/L2 latency; the probes of the two binary searches touch sorted
// neighbours, so most of them hit L1/L2.  ~4-8 B key read + 8 B written per tri.
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

template <typename KeyT>
__device__ __forceinline__ int delta(const KeyT* __restrict__ keys, uint32_t n, int64_t i, KeyT ki,
                                     int64_t j) {
  if (j < 0 || j >= (int64_t)n) return -1;
  KeyT kj = __ldg(keys + j);
  if (sizeof(KeyT) == 4) {
    uint32_t x = (uint32_t)ki ^ (uint32_t)kj;
    if (x) return __clz((int)x);
    return 32 + __clz((int)((uint32_t)i ^ (uint32_t)j));
  } else {
    uint64_t x = (uint64_t)ki ^ (uint64_t)kj;
    if (x) return __clzll((long long)x);
    return 64 + __clz((int)((uint32_t)i ^ (uint32_t)j));
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
karras_kernel(const KeyT* __restrict__ keys, uint32_t n, uint32_t* __restrict__ parent) {
  const uint32_t leaf0 = n - 1;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (int64_t)n - 1;
       i += (int64_t)gridDim.x * blockDim.x) {
    const KeyT ki = __ldg(keys + i);
    const int d = (delta(keys, n, i, ki, i + 1) - delta(keys, n, i, ki, i - 1)) < 0 ? -1 : 1;
    const int dmin = delta(keys, n, i, ki, i - d);
    int64_t lmax = 2;
    while (delta(keys, n, i, ki, i + lmax * d) > dmin) lmax *= 2;
    int64_t l = 0;
    for (int64_t t = lmax / 2; t >= 1; t /= 2)
      if (delta(keys, n, i, ki, i + (l + t) * d) > dmin) l += t;
    const int64_t j = i + l * d;
    const int dnode = delta(keys, n, i, ki, j);
    int64_t s = 0;
    int64_t t = l;
    do {
      t = (t + 1) >> 1;
      if (delta(keys, n, i, ki, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int64_t gamma = i + s * d + min(d, 0);
    const uint32_t lc = (min(i, j) == gamma) ? (uint32_t)(leaf0 + gamma) : (uint32_t)gamma;
    const uint32_t rc =
        (max(i, j) == gamma + 1) ? (uint32_t)(leaf0 + gamma + 1) : (uint32_t)(gamma + 1);
    parent[lc] = ((uint32_t)i << 1);
    parent[rc] = ((uint32_t)i << 1) | 1u;
    if (i == 0) parent[0] = kNoParent;
  }
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
  uint64_t blocks = (n - 1 + 255) / 256;
  uint64_t cap = (uint64_t)num_sms() * 32;
  if (blocks > cap) blocks = cap;
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
