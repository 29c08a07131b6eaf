// morton.cu — A1 scene bounds and A2 Morton encoding.
//
// A1  bounds_kernel: one streaming pass over the triangles (48 B read).  Each
//     thread keeps min/max as total-order integer keys (-0 < +0, exact and
//     order-free), warps reduce with redux.sync (__reduce_min/max_sync), blocks
//     reduce through shared memory and finish with 6 atomics per block.  The pass
//     also stores the centroid of each triangle's own box (12 B written), because
//     that is all the Morton pass needs: the triangles are read once, not twice.
// A2  morton(_hist)_kernel: one thread per triangle: centroid, quantised against
//     the scene box, bit-interleaved (x highest).  Writes the key only; the
//     payload of the first sort pass is the element position (iota), which needs
//     no memory.  On the default digit widths the kernel runs on the sort's chunk
//     geometry and also counts the first digit (radix_sort.cu).
//
// Roofline: HBM-bound: 48 + 12 B/tri, then 12 + 4 (or 8) B/tri.
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

constexpr int kBoundsThreads = 256;

// One thread per triangle and trip, two triangles (6 x LDG.128) in flight per thread.
// The centroid of the triangle's own box goes to three planes of floats; the scene box is
// the min/max over the triangle boxes, which is the min/max over all vertices.
__global__ void __launch_bounds__(kBoundsThreads)
bounds_kernel(const float4* __restrict__ tris, uint64_t n, uint32_t* __restrict__ out,
              float* __restrict__ cx, float* __restrict__ cy, float* __restrict__ cz) {
  pdl_enter();
  uint32_t mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
  const uint64_t stride = (uint64_t)gridDim.x * kBoundsThreads;
  auto one = [&](uint64_t t, const float4 a, const float4 b, const float4 c) {
    const float lx = omin(omin(a.x, b.x), c.x), hx = omax(omax(a.x, b.x), c.x);
    const float ly = omin(omin(a.y, b.y), c.y), hy = omax(omax(a.y, b.y), c.y);
    const float lz = omin(omin(a.z, b.z), c.z), hz = omax(omax(a.z, b.z), c.z);
    cx[t] = 0.5f * (lx + hx);
    cy[t] = 0.5f * (ly + hy);
    cz[t] = 0.5f * (lz + hz);
    mn[0] = min(mn[0], ord_key(lx)); mx[0] = max(mx[0], ord_key(hx));
    mn[1] = min(mn[1], ord_key(ly)); mx[1] = max(mx[1], ord_key(hy));
    mn[2] = min(mn[2], ord_key(lz)); mx[2] = max(mx[2], ord_key(hz));
  };
  uint64_t t = (uint64_t)blockIdx.x * kBoundsThreads + threadIdx.x;
  for (; t + stride < n; t += 2 * stride) {
    const uint64_t u = t + stride;
    const float4 a0 = ld_stream(tris + 3 * t), b0 = ld_stream(tris + 3 * t + 1),
                 c0 = ld_stream(tris + 3 * t + 2);
    const float4 a1 = ld_stream(tris + 3 * u), b1 = ld_stream(tris + 3 * u + 1),
                 c1 = ld_stream(tris + 3 * u + 2);
    one(t, a0, b0, c0);
    one(u, a1, b1, c1);
  }
  if (t < n) one(t, ld_stream(tris + 3 * t), ld_stream(tris + 3 * t + 1), ld_stream(tris + 3 * t + 2));
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
    // the minimum is kept complemented, so both halves are maxima and start from zero:
    // one memset with the build's other memsets instead of an initialising kernel
    atomicMax(&out[threadIdx.x], ~s_mn[threadIdx.x]);
    atomicMax(&out[3 + threadIdx.x], s_mx[threadIdx.x]);
  }
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

struct Quant {
  float smin[3], scale[3], qmax;
};
template <int KEY_BITS>
__device__ __forceinline__ Quant load_quant(const uint32_t* __restrict__ bkeys) {
  constexpr float kCells = (KEY_BITS == 30) ? 1024.0f : 2097152.0f;
  Quant q;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    q.smin[k] = ord_unkey(~bkeys[k]);  // stored complemented (bounds_kernel)
    const float ext = ord_unkey(bkeys[3 + k]) - q.smin[k];
    q.scale[k] = (ext > 0.0f) ? (kCells / ext) : 0.0f;
  }
  q.qmax = kCells - 1.0f;
  return q;
}
template <int KEY_BITS>
__device__ __forceinline__ uint64_t encode(const Quant& q, float cx, float cy, float cz) {
  const uint32_t qx = quantize(cx, q.smin[0], q.scale[0], q.qmax);
  const uint32_t qy = quantize(cy, q.smin[1], q.scale[1], q.qmax);
  const uint32_t qz = quantize(cz, q.smin[2], q.scale[2], q.qmax);
  if (KEY_BITS == 30) return (expand10(qx) << 2) | (expand10(qy) << 1) | expand10(qz);
  return (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz);
}

template <int KEY_BITS>
__global__ void __launch_bounds__(kMortonThreads)
morton_kernel(const float* __restrict__ cx, const float* __restrict__ cy,
              const float* __restrict__ cz, uint64_t n, const uint32_t* __restrict__ bkeys,
              void* __restrict__ keys_out) {
  pdl_enter();
  const Quant q = load_quant<KEY_BITS>(bkeys);
  for (uint64_t t = (uint64_t)blockIdx.x * kMortonThreads + threadIdx.x; t < n;
       t += (uint64_t)gridDim.x * kMortonThreads) {
    const uint64_t key = encode<KEY_BITS>(q, cx[t], cy[t], cz[t]);
    if (KEY_BITS == 30) reinterpret_cast<uint32_t*>(keys_out)[t] = (uint32_t)key;
    else reinterpret_cast<uint64_t*>(keys_out)[t] = key;
  }
}

// The same encode on the radix sort's chunk geometry (one CTA per chunk), which also counts
// the lowest digit of the keys it produces: the sort's first chunk_hist_kernel, and its read
// of every key, disappear.
constexpr int kMortonHistThreads = 1024;
// HYB (63-bit keys, hybrid sort order): also stores the top 30 bits of every key as the
// 32-bit key of the packed passes and counts THEIR lowest digit, on the 30-bit geometry.
template <int KEY_BITS, int BITS, bool HYB = false>
__global__ void __launch_bounds__(kMortonHistThreads)
morton_hist_kernel(const float* __restrict__ cx, const float* __restrict__ cy,
                   const float* __restrict__ cz, ChunkGeom g, const uint32_t* __restrict__ bkeys,
                   void* __restrict__ keys_out, uint32_t* __restrict__ chunk_hist,
                   uint32_t* __restrict__ totals, uint32_t* __restrict__ k32) {
  pdl_enter();
  constexpr int kBins = 1 << BITS;
  __shared__ uint32_t sh[kBins];
  for (int b = threadIdx.x; b < kBins; b += kMortonHistThreads) sh[b] = 0u;
  __syncthreads();
  const Quant q = load_quant<KEY_BITS>(bkeys);
  const uint32_t c = blockIdx.x, end = g.end(c);
  constexpr int kUnroll = 4;  // 12 independent loads in flight per thread
  for (uint32_t base = g.begin(c); base < end; base += kUnroll * kMortonHistThreads) {  // CTA-uniform trips
    float x[kUnroll], y[kUnroll], z[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const uint32_t t = base + u * kMortonHistThreads + threadIdx.x;
      const bool valid = t < end;
      x[u] = valid ? cx[t] : 0.0f;
      y[u] = valid ? cy[t] : 0.0f;
      z[u] = valid ? cz[t] : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const uint32_t t = base + u * kMortonHistThreads + threadIdx.x;
      const bool valid = t < end;
      uint32_t d = 0;
      if (valid) {
        const uint64_t key = encode<KEY_BITS>(q, x[u], y[u], z[u]);
        if (KEY_BITS == 30) reinterpret_cast<uint32_t*>(keys_out)[t] = (uint32_t)key;
        else reinterpret_cast<uint64_t*>(keys_out)[t] = key;
        d = (uint32_t)key & (kBins - 1);
        if (HYB) {
          k32[t] = (uint32_t)(key >> 33);
          d = (uint32_t)(key >> 33) & (kBins - 1);
        }
      }
      // coherent meshes put one cell, hence one digit, in a whole warp: count it once
      const uint32_t d0 = __shfl_sync(0xFFFFFFFFu, d, 0);
      if (__all_sync(0xFFFFFFFFu, valid && d == d0)) {
        if ((threadIdx.x & 31) == 0) atomicAdd(&sh[d], 32u);
      } else if (valid) {
        atomicAdd(&sh[d], 1u);
      }
    }
  }
  __syncthreads();
  for (int b = threadIdx.x; b < kBins; b += kMortonHistThreads) {
    const uint32_t v = sh[b];
    chunk_hist[(size_t)c * kBins + b] = v;
    if (v) atomicAdd(&totals[b], v);
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

int launch_bounds(const float4* tris, uint64_t n, uint32_t* bounds_keys, float* centroids,
                  cudaStream_t s) {
  RT_CUDA(cudaMemsetAsync(bounds_keys, 0, 6 * sizeof(uint32_t), s));
  if (n == 0) return 0;
  // 2 triangles per thread per trip: size the grid for the unrolled loop
  unsigned grid = persistent_grid((n + 1) / 2, kBoundsThreads, 8);
  RT_CUDA(launch_pdl(bounds_kernel, grid, kBoundsThreads, 0, s, tris, n, bounds_keys, centroids,
                     centroids + n, centroids + 2 * n));
  RT_LAUNCHED();
  return 0;
}

int launch_morton(const float* centroids, uint64_t n, const uint32_t* bounds_keys, uint32_t key_bits,
                  void* keys, uint32_t* chunk_hist, uint32_t* totals, int* fused, cudaStream_t s,
                  uint32_t* k32_hybrid) {
  if (fused) *fused = 0;
  if (n == 0) return 0;
  const float *cx = centroids, *cy = centroids + n, *cz = centroids + 2 * n;
  if (k32_hybrid && key_bits == 63 && chunk_hist && totals && fused && sort_plan(30).bits == 10) {
    const ChunkGeom g = sort_geometry(n, 30);
    RT_CUDA(launch_pdl(morton_hist_kernel<63, 10, true>, g.chunks, kMortonHistThreads, 0, s, cx, cy,
                       cz, g, bounds_keys, keys, chunk_hist, totals, k32_hybrid));
    RT_LAUNCHED();
    *fused = 2;  // keys, their top 30 bits and the first histogram of the packed passes
    return 0;
  }
  const SortPlan plan = sort_plan(key_bits);
  const bool fuse = chunk_hist && totals && fused && plan.bits == (key_bits == 30 ? 10 : 9);
  if (fuse) {
    const ChunkGeom g = sort_geometry(n, key_bits);
    if (key_bits == 30)
      RT_CUDA(launch_pdl(morton_hist_kernel<30, 10>, g.chunks, kMortonHistThreads, 0, s, cx, cy, cz, g,
                         bounds_keys, keys, chunk_hist, totals, (uint32_t*)nullptr));
    else
      RT_CUDA(launch_pdl(morton_hist_kernel<63, 9>, g.chunks, kMortonHistThreads, 0, s, cx, cy, cz, g,
                         bounds_keys, keys, chunk_hist, totals, (uint32_t*)nullptr));
    RT_LAUNCHED();
    *fused = 1;
    return 0;
  }
  unsigned grid = persistent_grid(n, kMortonThreads, 8);
  if (key_bits == 30)
    RT_CUDA(launch_pdl(morton_kernel<30>, grid, kMortonThreads, 0, s, cx, cy, cz, n, bounds_keys, keys));
  else
    RT_CUDA(launch_pdl(morton_kernel<63>, grid, kMortonThreads, 0, s, cx, cy, cz, n, bounds_keys, keys));
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
