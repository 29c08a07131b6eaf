// ray_gen.cu — A8: on-device ray generation (primary, random, ambient occlusion)
// and the small shading kernel of SURVEY.md §8(f).1.
//
// Rays are produced from (seed, global ray index) only, so a shard of a batch
// is the same bytes whichever GPU generates it and 10^9 rays
// never cross PCIe.  One thread per ray, two coalesced float4 stores.
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

struct Cam {
  float v[12];
};
struct Box6 {
  float v[6];
};

__global__ void __launch_bounds__(256)
primary_kernel(Cam cam, uint32_t width, uint64_t first, uint64_t count, float4* __restrict__ out) {
  const float inf = __int_as_float(0x7f800000);
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t i = first + k;
    float fx = (float)(uint32_t)(i % width) + 0.5f, fy = (float)(uint32_t)(i / width) + 0.5f;
    float dx = (cam.v[3] + cam.v[6] * fx + cam.v[9] * fy) - cam.v[0];
    float dy = (cam.v[4] + cam.v[7] * fx + cam.v[10] * fy) - cam.v[1];
    float dz = (cam.v[5] + cam.v[8] * fx + cam.v[11] * fy) - cam.v[2];
    normalize3(dx, dy, dz);
    st_stream(out + 2 * k, make_float4(cam.v[0], cam.v[1], cam.v[2], 0.0f));
    st_stream(out + 2 * k + 1, make_float4(dx, dy, dz, inf));
  }
}

__global__ void __launch_bounds__(256)
random_kernel(Box6 box, uint64_t key, uint64_t first, uint64_t count, float4* __restrict__ out) {
  const float inf = __int_as_float(0x7f800000);
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t i = first + k;
    float ox = box.v[0] + u01(rnd64(key, i, 0)) * (box.v[3] - box.v[0]);
    float oy = box.v[1] + u01(rnd64(key, i, 1)) * (box.v[4] - box.v[1]);
    float oz = box.v[2] + u01(rnd64(key, i, 2)) * (box.v[5] - box.v[2]);
    float dx, dy, dz;
    sphere_sample(key, i, 4, dx, dy, dz);
    st_stream(out + 2 * k, make_float4(ox, oy, oz, 0.0f));
    st_stream(out + 2 * k + 1, make_float4(dx, dy, dz, inf));
  }
}

// geometric normal of the hit triangle, flipped against the ray direction
__device__ __forceinline__ void facing_normal(const float4* __restrict__ tris, int prim, float dx,
                                              float dy, float dz, float& nx, float& ny, float& nz) {
  float4 v0 = __ldg(tris + 3 * (uint64_t)prim), v1 = __ldg(tris + 3 * (uint64_t)prim + 1),
         v2 = __ldg(tris + 3 * (uint64_t)prim + 2);
  float e1x = v1.x - v0.x, e1y = v1.y - v0.y, e1z = v1.z - v0.z;
  float e2x = v2.x - v0.x, e2y = v2.y - v0.y, e2z = v2.z - v0.z;
  nx = fmaf(e1y, e2z, -(e1z * e2y));
  ny = fmaf(e1z, e2x, -(e1x * e2z));
  nz = fmaf(e1x, e2y, -(e1y * e2x));
  normalize3(nx, ny, nz);
  float dn = fmaf(nx, dx, fmaf(ny, dy, nz * dz));
  if (dn > 0.0f) {
    nx = -nx;
    ny = -ny;
    nz = -nz;
  }
}

__global__ void __launch_bounds__(256)
ao_kernel(const float4* __restrict__ tris, const float4* __restrict__ rays,
          const float4* __restrict__ hits, uint64_t n, uint32_t spp, uint64_t key, float radius,
          float4* __restrict__ out, uint64_t first) {
  const float eps = radius * 1e-3f;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    float4 h = hits[i];
    int prim = __float_as_int(h.w);
    if (prim < 0) {
      for (uint32_t s = 0; s < spp; ++s) {
        st_stream(out + 2 * (i * spp + s), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
        st_stream(out + 2 * (i * spp + s) + 1, make_float4(0.0f, 0.0f, 1.0f, -1.0f));
      }
      continue;
    }
    float4 ro = rays[2 * i], rd = rays[2 * i + 1];
    float nx, ny, nz;
    facing_normal(tris, prim, rd.x, rd.y, rd.z, nx, ny, nz);
    float ox = fmaf(h.x, rd.x, ro.x) + nx * eps;
    float oy = fmaf(h.x, rd.y, ro.y) + ny * eps;
    float oz = fmaf(h.x, rd.z, ro.z) + nz * eps;
    for (uint32_t s = 0; s < spp; ++s) {
      float sx, sy, sz;
      sphere_sample(key, (first + i) * spp + s, 0, sx, sy, sz);  // the sample of GLOBAL ray first + i
      float dx = nx + sx, dy = ny + sy, dz = nz + sz;
      float l2 = dx * dx + dy * dy + dz * dz;
      if (l2 < 1e-12f) {
        dx = nx;
        dy = ny;
        dz = nz;
      } else {
        normalize3(dx, dy, dz);
      }
      st_stream(out + 2 * (i * spp + s), make_float4(ox, oy, oz, 0.0f));
      st_stream(out + 2 * (i * spp + s) + 1, make_float4(dx, dy, dz, radius));
    }
  }
}

// rgb = |n.d| * (1 - occluded/spp); misses are black
__global__ void __launch_bounds__(256)
shade_kernel(const float4* __restrict__ tris, const float4* __restrict__ rays,
             const float4* __restrict__ hits, const float4* __restrict__ ao_hits, uint64_t n,
             uint32_t spp, float* __restrict__ rgb) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    float4 h = hits[i];
    int prim = __float_as_int(h.w);
    float c = 0.0f;
    if (prim >= 0) {
      float4 rd = rays[2 * i + 1];
      float nx, ny, nz;
      facing_normal(tris, prim, rd.x, rd.y, rd.z, nx, ny, nz);
      c = fabsf(fmaf(nx, rd.x, fmaf(ny, rd.y, nz * rd.z)));
      if (ao_hits != nullptr && spp > 0) {
        uint32_t occ = 0;
        for (uint32_t s = 0; s < spp; ++s)
          occ += (__float_as_int(ao_hits[i * spp + s].w) >= 0) ? 1u : 0u;
        c = c * (1.0f - (float)occ / (float)spp);
      }
    }
    rgb[3 * i + 0] = c;
    rgb[3 * i + 1] = c;
    rgb[3 * i + 2] = c;
  }
}

// hits with a primitive (closest hit) or the any-hit marker
__global__ void __launch_bounds__(256)
count_hits_kernel(const float4* __restrict__ hits, uint64_t n, unsigned long long* __restrict__ out) {
  uint32_t c = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x)
    c += (__float_as_int(ld_stream(hits + i).w) >= 0) ? 1u : 0u;
  c = __reduce_add_sync(0xFFFFFFFFu, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

inline unsigned grid_for(uint64_t count) {
  uint64_t blocks = (count + 255) / 256;
  uint64_t cap = (uint64_t)num_sms() * 16;
  return (unsigned)(blocks > cap ? cap : blocks);
}

}  // namespace

int launch_gen_primary(const float* cam12, uint32_t width, uint32_t height, uint64_t first,
                       uint64_t count, float4* out, cudaStream_t s) {
  (void)height;
  if (count == 0) return 0;
  Cam cam;
  for (int i = 0; i < 12; ++i) cam.v[i] = cam12[i];
  primary_kernel<<<grid_for(count), 256, 0, s>>>(cam, width, first, count, out);
  RT_LAUNCHED();
  return 0;
}

int launch_gen_random(const float* box6, uint64_t seed, uint64_t first, uint64_t count,
                      float4* out, cudaStream_t s) {
  if (count == 0) return 0;
  Box6 box;
  for (int i = 0; i < 6; ++i) box.v[i] = box6[i];
  random_kernel<<<grid_for(count), 256, 0, s>>>(box, stream_key(seed, kStreamRays), first, count,
                                                out);
  RT_LAUNCHED();
  return 0;
}

int launch_gen_ao(const float4* tris, const float4* rays, const float4* hits, uint64_t n,
                  uint32_t spp, uint64_t seed, float radius, float4* out, cudaStream_t s, uint64_t first) {
  if (n == 0 || spp == 0) return 0;
  ao_kernel<<<grid_for(n), 256, 0, s>>>(tris, rays, hits, n, spp, stream_key(seed, kStreamAO),
                                        radius, out, first);
  RT_LAUNCHED();
  return 0;
}

int launch_count_hits(const float4* hits, uint64_t n, unsigned long long* out, cudaStream_t s) {
  if (n == 0) return 0;
  count_hits_kernel<<<grid_for(n), 256, 0, s>>>(hits, n, out);
  RT_LAUNCHED();
  return 0;
}

int launch_shade(const float4* tris, const float4* rays, const float4* hits,
                 const float4* ao_hits, uint64_t n, uint32_t spp, float* rgb, cudaStream_t s) {
  if (n == 0) return 0;
  shade_kernel<<<grid_for(n), 256, 0, s>>>(tris, rays, hits, ao_hits, n, spp, rgb);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
