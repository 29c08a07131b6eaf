// scene_gen.cu — on-device synthetic scenes of SURVEY.md §8(d) / BASELINE.json:7-11.
//
// Every triangle is a pure function of (scene parameters, triangle index), so a
// scene of any size is produced at HBM write speed without touching the host:
// one thread per triangle, three coalesced float4 stores.  The arithmetic is
// integer hashes plus + - * / sqrt in a fixed order (compiled -fmad=false), so
// the bytes equal those of the CPU oracle's generator (tests/test_generators.py).
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }

__device__ __forceinline__ void store_tri(float4* out, uint64_t k, V3 a, V3 b, V3 c) {
  float4* o = out + k * 3;
  st_stream(o + 0, make_float4(a.x, a.y, a.z, 0.0f));
  st_stream(o + 1, make_float4(b.x, b.y, b.z, 0.0f));
  st_stream(o + 2, make_float4(c.x, c.y, c.z, 0.0f));
}

// ---- config 1: Cornell box ---------------------------------------------------
__constant__ float kWalls[5][9] = {
    {0, 0, 0, 1, 0, 0, 0, 0, 1}, {0, 1, 0, 1, 0, 0, 0, 0, 1}, {0, 0, 1, 1, 0, 0, 0, 1, 0},
    {0, 0, 0, 0, 0, 1, 0, 1, 0}, {1, 0, 0, 0, 0, 1, 0, 1, 0},
};
__constant__ float kBoxes[2][6] = {
    {0.13f, 0.0f, 0.13f, 0.3f, 0.3f, 0.3f},
    {0.55f, 0.0f, 0.5f, 0.3f, 0.6f, 0.3f},
};

__device__ __forceinline__ float grid_pt(float org, float U, float V, float s, float t) {
  return org + U * s + V * t;  // (org + U*s) + V*t, unfused
}

__device__ void grid_tri(const float* org, const float* U, const float* V, uint32_t n, uint32_t qi,
                         uint32_t qj, uint32_t half, V3& a, V3& b, V3& c) {
  float inv = 1.0f / (float)n;
  float s0 = (float)qi * inv, s1 = (float)(qi + 1) * inv;
  float t0 = (float)qj * inv, t1 = (float)(qj + 1) * inv;
  a = mk(grid_pt(org[0], U[0], V[0], s0, t0), grid_pt(org[1], U[1], V[1], s0, t0),
         grid_pt(org[2], U[2], V[2], s0, t0));
  V3 p11 = mk(grid_pt(org[0], U[0], V[0], s1, t1), grid_pt(org[1], U[1], V[1], s1, t1),
              grid_pt(org[2], U[2], V[2], s1, t1));
  if (half == 0) {
    b = mk(grid_pt(org[0], U[0], V[0], s1, t0), grid_pt(org[1], U[1], V[1], s1, t0),
           grid_pt(org[2], U[2], V[2], s1, t0));
    c = p11;
  } else {
    b = p11;
    c = mk(grid_pt(org[0], U[0], V[0], s0, t1), grid_pt(org[1], U[1], V[1], s0, t1),
           grid_pt(org[2], U[2], V[2], s0, t1));
  }
}

__device__ void cornell_tri(uint64_t t, V3& a, V3& b, V3& c) {
  if (t < 640) {
    uint32_t w = (uint32_t)(t / 128), r = (uint32_t)(t % 128);
    uint32_t q = r >> 1;
    grid_tri(&kWalls[w][0], &kWalls[w][3], &kWalls[w][6], 8, q & 7, q >> 3, r & 1, a, b, c);
    return;
  }
  t -= 640;
  uint32_t bx = (uint32_t)(t / 192), r = (uint32_t)(t % 192);
  uint32_t f = r / 32, rr = r % 32, q = rr >> 1;
  float org[3] = {kBoxes[bx][0], kBoxes[bx][1], kBoxes[bx][2]};
  float U[3] = {0.0f, 0.0f, 0.0f}, V[3] = {0.0f, 0.0f, 0.0f};
  uint32_t axis = f >> 1, side = f & 1;
  uint32_t a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
  if (side) org[axis] = kBoxes[bx][axis] + kBoxes[bx][3 + axis];
  U[a1] = kBoxes[bx][3 + a1];
  V[a2] = kBoxes[bx][3 + a2];
  grid_tri(org, U, V, 4, q & 3, q >> 2, rr & 1, a, b, c);
}

// ---- config 2: triangle soup ---------------------------------------------------
__device__ void soup_tri(uint64_t key, uint64_t t, V3& a, V3& b, V3& c) {
  float cx = u01(rnd64(key, t, 0)), cy = u01(rnd64(key, t, 1)), cz = u01(rnd64(key, t, 2));
  float v[9];
#pragma unroll
  for (uint32_t j = 0; j < 9; ++j) {
    float cen = (j % 3 == 0) ? cx : ((j % 3 == 1) ? cy : cz);
    v[j] = cen + (u01(rnd64(key, t, 3 + j)) - 0.5f) * 0.01f;
  }
  a = mk(v[0], v[1], v[2]);
  b = mk(v[3], v[4], v[5]);
  c = mk(v[6], v[7], v[8]);
}

// ---- config 3: sphereflake -----------------------------------------------------
__constant__ float kFlakeDirs[9][3] = {
    {1.0f, 0.0f, 0.0f},       {0.5f, 0.8660254f, 0.0f},   {-0.5f, 0.8660254f, 0.0f},
    {-1.0f, 0.0f, 0.0f},      {-0.5f, -0.8660254f, 0.0f}, {0.5f, -0.8660254f, 0.0f},
    {0.5196152f, 0.3f, 0.8f}, {-0.5196152f, 0.3f, 0.8f},  {0.0f, -0.6f, 0.8f},
};

__device__ void flake_sphere(uint64_t s, V3& c, float& r) {
  uint32_t level = 0;
  uint64_t off = 0, p = 1;
  while (s >= off + p) {
    off += p;
    p *= 9;
    ++level;
  }
  uint64_t w = s - off;
  // digits most-significant first: divide by 9^(level-1-l)
  c = mk(0.0f, 0.0f, 0.0f);
  r = 1.0f;
  V3 ax = mk(0.0f, 0.0f, 1.0f);
  uint64_t div = p / 9;  // 9^(level-1) when level > 0
  for (uint32_t l = 0; l < level; ++l) {
    uint32_t digit = (uint32_t)((w / div) % 9);
    div /= 9;
    // orthonormal basis around ax (Duff et al. 2017)
    float sign = copysignf(1.0f, ax.z);
    float a = -1.0f / (sign + ax.z);
    float b = ax.x * ax.y * a;
    V3 b1 = mk(1.0f + sign * ax.x * ax.x * a, sign * b, -sign * ax.x);
    V3 b2 = mk(b, sign + ax.y * ax.y * a, -ax.y);
    float Lx = kFlakeDirs[digit][0], Ly = kFlakeDirs[digit][1], Lz = kFlakeDirs[digit][2];
    float dx = Lx * b1.x + Ly * b2.x + Lz * ax.x;
    float dy = Lx * b1.y + Ly * b2.y + Lz * ax.y;
    float dz = Lx * b1.z + Ly * b2.z + Lz * ax.z;
    normalize3(dx, dy, dz);
    float rc = r * 0.33333334f;
    float dist = r + rc;
    c = mk(c.x + dx * dist, c.y + dy * dist, c.z + dz * dist);
    ax = mk(dx, dy, dz);
    r = rc;
  }
}

__device__ __forceinline__ V3 flake_vertex(V3 c, float r, V3 sgn, uint32_t m, uint32_t row,
                                           uint32_t col) {
  float fm = (float)m;
  float nx = sgn.x * ((float)col / fm);
  float ny = sgn.y * ((float)(row - col) / fm);
  float nz = sgn.z * ((float)(m - row) / fm);
  normalize3(nx, ny, nz);
  return mk(c.x + r * nx, c.y + r * ny, c.z + r * nz);
}

__device__ void flake_tri(uint32_t m, uint64_t t, V3& a, V3& b, V3& d) {
  uint64_t per = 8ull * m * m;
  uint64_t s = t / per;
  uint32_t rem = (uint32_t)(t % per);
  uint32_t face = rem / (m * m), sub = rem % (m * m);
  V3 c;
  float r;
  flake_sphere(s, c, r);
  V3 sgn = mk((face & 1) ? -1.0f : 1.0f, (face & 2) ? -1.0f : 1.0f, (face & 4) ? -1.0f : 1.0f);
  uint32_t row = (uint32_t)sqrtf((float)sub);
  while (row * row > sub) --row;
  while ((row + 1) * (row + 1) <= sub) ++row;
  uint32_t k = sub - row * row;
  if ((k & 1) == 0) {
    uint32_t col = k >> 1;
    a = flake_vertex(c, r, sgn, m, row, col);
    b = flake_vertex(c, r, sgn, m, row + 1, col);
    d = flake_vertex(c, r, sgn, m, row + 1, col + 1);
  } else {
    uint32_t col = (k - 1) >> 1;
    a = flake_vertex(c, r, sgn, m, row, col);
    b = flake_vertex(c, r, sgn, m, row + 1, col + 1);
    d = flake_vertex(c, r, sgn, m, row, col + 1);
  }
}

// ---- config 4: terrain ---------------------------------------------------------
__device__ float terrain_height(uint64_t key, uint32_t i, uint32_t j, uint32_t base) {
  float h = 0.0f, amp = 0.25f;
#pragma unroll 1
  for (uint32_t o = 0; o < 5; ++o) {
    uint32_t S = base >> o;
    if (S < 1) S = 1;
    uint32_t li = i / S, lj = j / S;
    float fi = (float)(i % S) / (float)S, fj = (float)(j % S) / (float)S;
    float wi = fi * fi * (3.0f - 2.0f * fi), wj = fj * fj * (3.0f - 2.0f * fj);
    uint64_t okey = key + o;
    float v00 = u01(rnd64(okey, ((uint64_t)lj << 24) + li, 0));
    float v10 = u01(rnd64(okey, ((uint64_t)lj << 24) + li + 1, 0));
    float v01 = u01(rnd64(okey, ((uint64_t)(lj + 1) << 24) + li, 0));
    float v11 = u01(rnd64(okey, ((uint64_t)(lj + 1) << 24) + li + 1, 0));
    float a = v00 + (v10 - v00) * wi;
    float b = v01 + (v11 - v01) * wi;
    float v = a + (b - a) * wj;
    h = h + amp * v;
    amp = amp * 0.5f;
  }
  return h;
}
__device__ __forceinline__ V3 terrain_vertex(uint64_t key, uint32_t g, uint32_t base, uint32_t i,
                                             uint32_t j) {
  return mk((float)i / (float)g, terrain_height(key, i, j, base), (float)j / (float)g);
}
__device__ void terrain_tri(uint64_t key, uint32_t g, uint32_t base, uint64_t t, V3& a, V3& b,
                            V3& c) {
  uint64_t cell = t >> 1;
  uint32_t ci = (uint32_t)(cell % g), cj = (uint32_t)(cell / g);
  a = terrain_vertex(key, g, base, ci, cj);
  V3 v11 = terrain_vertex(key, g, base, ci + 1, cj + 1);
  if ((t & 1) == 0) {
    b = terrain_vertex(key, g, base, ci + 1, cj);
    c = v11;
  } else {
    b = v11;
    c = terrain_vertex(key, g, base, ci, cj + 1);
  }
}

__global__ void __launch_bounds__(256)
gen_scene_kernel(int kind, uint64_t p0, uint64_t p1, uint64_t key, uint64_t first, uint64_t count,
                 float4* __restrict__ out) {
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < count;
       k += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t t = first + k;
    V3 a, b, c;
    switch (kind) {
      case 0: cornell_tri(t, a, b, c); break;
      case 1: soup_tri(key, t, a, b, c); break;
      case 2: flake_tri((uint32_t)p1, t, a, b, c); break;
      default: terrain_tri(key, (uint32_t)p0, (uint32_t)(p1 ? p1 : 1024), t, a, b, c); break;
    }
    store_tri(out, k, a, b, c);
  }
}

}  // namespace

int launch_gen_scene(int kind, uint64_t p0, uint64_t p1, uint64_t seed, uint64_t first,
                     uint64_t count, float4* out, cudaStream_t s) {
  if (count == 0) return 0;
  uint64_t key = stream_key(seed, kind == 1 ? kStreamSoup : kStreamTerrain);
  uint64_t blocks = (count + 255) / 256;
  uint64_t cap = (uint64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  gen_scene_kernel<<<(unsigned)blocks, 256, 0, s>>>(kind, p0, p1, key, first, count, out);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
