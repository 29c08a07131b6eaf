// morton.cu — A1 scene bounds and A2 Morton encoding.
//
// A1  bounds_kernel: one streaming pass over the 3N vertices (48 B/triangle).
//     Each thread keeps min/max as total-order integer keys (-0 < +0, exact and
//     order-free), warps reduce with redux.sync (__reduce_min/max_sync), blocks
//     reduce through shared memory and finish with 6 atomics per block.
// A2  morton_kernel: one thread per triangle: centroid of the triangle's own
//     box, quantised against the scene box, bit-interleaved (x highest).
//     Writes the key only; the payload of the first sort pass is the element
//     position (iota), which needs no memory.
//
// Roofline: both kernels are HBM-bound: 48 B/tri read (+4 or 8 B/tri written).
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

constexpr int kBoundsThreads = 256;

__global__ void __launch_bounds__(kBoundsThreads)
bounds_kernel(const float4* __restrict__ verts, uint64_t n_verts, uint32_t* __restrict__ out) {
  uint32_t mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
  const uint64_t stride = (uint64_t)gridDim.x * kBoundsThreads;
  uint64_t i = (uint64_t)blockIdx.x * kBoundsThreads + threadIdx.x;
  // 4 independent 16-byte loads in flight per thread
  for (; i + 3 * stride < n_verts; i += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = ld_stream(verts + i + k * stride);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint32_t kx = ord_key(v[k].x), ky = ord_key(v[k].y), kz = ord_key(v[k].z);
      mn[0] = min(mn[0], kx); mx[0] = max(mx[0], kx);
      mn[1] = min(mn[1], ky); mx[1] = max(mx[1], ky);
      mn[2] = min(mn[2], kz); mx[2] = max(mx[2], kz);
    }
  }
  for (; i < n_verts; i += stride) {
    float4 v = ld_stream(verts + i);
    uint32_t kx = ord_key(v.x), ky = ord_key(v.y), kz = ord_key(v.z);
    mn[0] = min(mn[0], kx); mx[0] = max(mx[0], kx);
    mn[1] = min(mn[1], ky); mx[1] = max(mx[1], ky);
    mn[2] = min(mn[2], kz); mx[2] = max(mx[2], kz);
  }
  __shared__ uint32_t s_mn[3], s_mx[3];
  if (threadIdx.x < 3) {
    s_mn[threadIdx.x] = 0xFFFFFFFFu;
    s_mx[threadIdx.x] = 0u;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    uint32_t a = __reduce_min_sync(0xFFFFFFFFu, mn[k]);
    uint32_t b = __reduce_max_sync(0xFFFFFFFFu, mx[k]);
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&s_mn[k], a);
      atomicMax(&s_mx[k], b);
    }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    atomicMin(&out[threadIdx.x], s_mn[threadIdx.x]);
    atomicMax(&out[3 + threadIdx.x], s_mx[threadIdx.x]);
  }
}

__global__ void init_bounds_kernel(uint32_t* out) {
  if (threadIdx.x < 3) out[threadIdx.x] = 0xFFFFFFFFu;
  else if (threadIdx.x < 6) out[threadIdx.x] = 0u;
}

__device__ __forceinline__ uint32_t expand10(uint32_t v) {
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__device__ __forceinline__ uint64_t expand21(uint64_t v) {
  v &= 0x1FFFFFull;
  v = (v | (v << 32)) & 0x001F00000000FFFFull;
  v = (v | (v << 16)) & 0x001F0000FF0000FFull;
  v = (v | (v << 8)) & 0x100F00F00F00F00Full;
  v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

__device__ __forceinline__ uint32_t quantize(float c, float smin, float scale, float qmax) {
  float x = (c - smin) * scale;
  x = fmaxf(x, 0.0f);  // NaN-ignoring, like the oracle's nmax/nmin
  x = fminf(x, qmax);
  return (uint32_t)x;
}

constexpr int kMortonThreads = 256;

template <int KEY_BITS>
__global__ void __launch_bounds__(kMortonThreads)
morton_kernel(const float4* __restrict__ tris, uint64_t n, const uint32_t* __restrict__ bkeys,
              void* __restrict__ keys_out) {
  constexpr float kCells = (KEY_BITS == 30) ? 1024.0f : 2097152.0f;

  float smin[3], scale[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    smin[k] = ord_unkey(bkeys[k]);
    float ext = ord_unkey(bkeys[3 + k]) - smin[k];
    scale[k] = (ext > 0.0f) ? (kCells / ext) : 0.0f;
  }
  const float qmax = kCells - 1.0f;

  for (uint64_t t = (uint64_t)blockIdx.x * kMortonThreads + threadIdx.x; t < n;
       t += (uint64_t)gridDim.x * kMortonThreads) {
    float4 a = ld_stream(tris + 3 * t), b = ld_stream(tris + 3 * t + 1),
           c = ld_stream(tris + 3 * t + 2);
    float cx = 0.5f * (omin(omin(a.x, b.x), c.x) + omax(omax(a.x, b.x), c.x));
    float cy = 0.5f * (omin(omin(a.y, b.y), c.y) + omax(omax(a.y, b.y), c.y));
    float cz = 0.5f * (omin(omin(a.z, b.z), c.z) + omax(omax(a.z, b.z), c.z));
    uint32_t qx = quantize(cx, smin[0], scale[0], qmax);
    uint32_t qy = quantize(cy, smin[1], scale[1], qmax);
    uint32_t qz = quantize(cz, smin[2], scale[2], qmax);
    if (KEY_BITS == 30) {
      uint32_t key = (expand10(qx) << 2) | (expand10(qy) << 1) | expand10(qz);
      reinterpret_cast<uint32_t*>(keys_out)[t] = key;
    } else {
      uint64_t key = (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz);
      reinterpret_cast<uint64_t*>(keys_out)[t] = key;
    }
  }
}

inline unsigned persistent_grid(uint64_t items, int threads, int blocks_per_sm) {
  uint64_t blocks = (items + threads - 1) / threads;
  uint64_t cap = (uint64_t)num_sms() * blocks_per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace

int launch_bounds(const float4* tris, uint64_t n, uint32_t* bounds_keys, cudaStream_t s) {
  init_bounds_kernel<<<1, 32, 0, s>>>(bounds_keys);
  RT_LAUNCHED();
  if (n == 0) return 0;
  uint64_t n_verts = 3 * n;
  // 4 loads per thread per trip: size the grid for the unrolled loop
  unsigned grid = persistent_grid((n_verts + 3) / 4, kBoundsThreads, 8);
  bounds_kernel<<<grid, kBoundsThreads, 0, s>>>(tris, n_verts, bounds_keys);
  RT_LAUNCHED();
  return 0;
}

int launch_morton(const float4* tris, uint64_t n, const uint32_t* bounds_keys, uint32_t key_bits,
                  void* keys, cudaStream_t s) {
  if (n == 0) return 0;
  unsigned grid = persistent_grid(n, kMortonThreads, 8);
  if (key_bits == 30)
    morton_kernel<30><<<grid, kMortonThreads, 0, s>>>(tris, n, bounds_keys, keys);
  else
    morton_kernel<63><<<grid, kMortonThreads, 0, s>>>(tris, n, bounds_keys, keys);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
