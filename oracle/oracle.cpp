// SPDX-License-Identifier: MIT
//
// oracle.cpp — spec-derived scalar C++ oracle, NOT derived from reference source.
//
// TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library.  The
// product (gpu_raytracer_b200/) never links, imports or calls it.
//
// PARITY UNPINNED: the mounted reference (/root/reference) holds README.md:1-2
// and test.txt:1 and no code, no golden vectors and no tests (SURVEY.md §0,
// §8c).  This file restates the algorithm named in BASELINE.json:5 with the
// arithmetic fixed in SURVEY.md §8(a) rows A1..A8.  It is pinned instead by
//   T0  a brute-force all-pairs intersector (orc_brute, no BVH at all),
//   T1  literal Morton known-answer vectors (tests/test_oracle.py),
//   T2  std::stable_sort from the standard library,
//   T3  structural tree invariants checked in the tests.
//
// This is its own translation unit and shares no header with the CUDA kernels,
// so a bug cannot be shared between the checker and the thing checked.
//
// Floating-point contract (must hold on the GPU side bit for bit):
//   * compiled with -ffp-contract=off: the only fused operations are the
//     explicit fmaf() calls below; every other + - * / sqrt is one IEEE
//     round-to-nearest-even operation;
//   * box min/max use a total order on the bit patterns (-0 < +0), so node
//     boxes are exact and independent of reduction order;
//   * slab-test min/max ignore NaN (return the other operand).
//
// Build: see oracle/Makefile  (g++ -O2 -ffp-contract=off -fopenmp -mfma).

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#include <parallel/algorithm>
#endif

#define ORC_API extern "C" __attribute__((visibility("default")))

namespace {

// ---------------------------------------------------------------------------
// counter-based random numbers (splitmix64 of (seed, index),
// 24-bit mantissa floats; only + - * / sqrt downstream)
// ---------------------------------------------------------------------------
inline uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline uint64_t stream_key(uint64_t seed, uint64_t stream) {
  return mix64(seed * 0xD6E8FEB86659FD93ull + stream);
}
// j < 256 draws per index
inline uint64_t rnd64(uint64_t key, uint64_t idx, uint32_t j) {
  return mix64(key ^ (idx * 256ull + j));
}
inline float u01(uint64_t h) { return (float)(uint32_t)(h >> 40) * 0x1p-24f; }

enum { STREAM_SOUP = 1, STREAM_TERRAIN = 2, STREAM_RAYS = 3, STREAM_AO = 4 };

// ---------------------------------------------------------------------------
// total-order float min/max for boxes (A1, A5): -0 < +0, exact, order-free
// ---------------------------------------------------------------------------
inline uint32_t fbits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float bitsf(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
inline uint32_t ord_key(float f) {
  uint32_t u = fbits(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
inline float ord_unkey(uint32_t k) {
  return bitsf((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}
inline float omin(float a, float b) { return ord_key(a) < ord_key(b) ? a : b; }
inline float omax(float a, float b) { return ord_key(a) > ord_key(b) ? a : b; }

// NaN-ignoring min/max for the slab test (A6)
inline float nmin(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a < b ? a : b)); }
inline float nmax(float a, float b) { return (a != a) ? b : ((b != b) ? a : (a > b ? a : b)); }

// ---------------------------------------------------------------------------
// A8 helper: uniform direction on the sphere, Marsaglia (1972) rejection.
// draws j0, j0+1 per attempt; at most 16 attempts, then +z.
// ---------------------------------------------------------------------------
inline void sphere_sample(uint64_t key, uint64_t idx, uint32_t j0, float out[3]) {
  for (uint32_t a = 0; a < 16; ++a) {
    float x1 = u01(rnd64(key, idx, j0 + 2 * a)) * 2.0f - 1.0f;
    float x2 = u01(rnd64(key, idx, j0 + 2 * a + 1)) * 2.0f - 1.0f;
    float s = x1 * x1 + x2 * x2;
    if (s < 1.0f) {
      float r = std::sqrt(1.0f - s);
      out[0] = 2.0f * x1 * r;
      out[1] = 2.0f * x2 * r;
      out[2] = 1.0f - 2.0f * s;
      return;
    }
  }
  out[0] = 0.0f; out[1] = 0.0f; out[2] = 1.0f;
}

inline void normalize3(float v[3]) {
  float l2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  float l = std::sqrt(l2);
  v[0] = v[0] / l; v[1] = v[1] / l; v[2] = v[2] / l;
}

// ---------------------------------------------------------------------------
// Scene generators.  Every triangle is a pure function of its index.
// Layout: 12 floats per triangle, v0.xyz,0, v1.xyz,0, v2.xyz,0 (3 x float4).
// ---------------------------------------------------------------------------
enum { SCENE_CORNELL = 0, SCENE_SOUP = 1, SCENE_SPHEREFLAKE = 2, SCENE_TERRAIN = 3 };

inline void put_tri(float* o, const float a[3], const float b[3], const float c[3]) {
  o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; o[3] = 0.0f;
  o[4] = b[0]; o[5] = b[1]; o[6] = b[2]; o[7] = 0.0f;
  o[8] = c[0]; o[9] = c[1]; o[10] = c[2]; o[11] = 0.0f;
}

// quad (qi,qj) of an n x n grid on the parallelogram org + s*U + t*V; half 0/1
inline void grid_tri(const float org[3], const float U[3], const float V[3], uint32_t n,
                     uint32_t qi, uint32_t qj, uint32_t half, float* out) {
  float inv = 1.0f / (float)n;
  float s0 = (float)qi * inv, s1 = (float)(qi + 1) * inv;
  float t0 = (float)qj * inv, t1 = (float)(qj + 1) * inv;
  float p00[3], p10[3], p11[3], p01[3];
  for (int k = 0; k < 3; ++k) {
    p00[k] = org[k] + U[k] * s0 + V[k] * t0;
    p10[k] = org[k] + U[k] * s1 + V[k] * t0;
    p11[k] = org[k] + U[k] * s1 + V[k] * t1;
    p01[k] = org[k] + U[k] * s0 + V[k] * t1;
  }
  if (half == 0) put_tri(out, p00, p10, p11);
  else put_tri(out, p00, p11, p01);
}

// config 1 (BASELINE.json:7): Cornell box in [0,1]^3, 5 walls x 8x8x2 = 640,
// two axis-aligned boxes x 6 faces x 4x4x2 = 384; 1,024 triangles.
void cornell_tri(uint64_t t, float* out) {
  static const float walls[5][9] = {
      // org, U, V
      {0, 0, 0, 1, 0, 0, 0, 0, 1},  // floor   y=0
      {0, 1, 0, 1, 0, 0, 0, 0, 1},  // ceiling y=1
      {0, 0, 1, 1, 0, 0, 0, 1, 0},  // back    z=1
      {0, 0, 0, 0, 0, 1, 0, 1, 0},  // left    x=0
      {1, 0, 0, 0, 0, 1, 0, 1, 0},  // right   x=1
  };
  static const float boxes[2][6] = {
      // min.xyz, size.xyz
      {0.13f, 0.0f, 0.13f, 0.3f, 0.3f, 0.3f},   // short box
      {0.55f, 0.0f, 0.5f, 0.3f, 0.6f, 0.3f},    // tall box
  };
  if (t < 640) {
    uint32_t w = (uint32_t)(t / 128), r = (uint32_t)(t % 128);
    uint32_t q = r >> 1;
    grid_tri(&walls[w][0], &walls[w][3], &walls[w][6], 8, q & 7, q >> 3, r & 1, out);
    return;
  }
  t -= 640;
  uint32_t b = (uint32_t)(t / 192), r = (uint32_t)(t % 192);
  uint32_t f = r / 32, rr = r % 32, q = rr >> 1;
  const float* B = boxes[b];
  float org[3] = {B[0], B[1], B[2]}, U[3] = {0, 0, 0}, V[3] = {0, 0, 0};
  uint32_t axis = f >> 1, side = f & 1;  // face normal axis, 0 = min side, 1 = max side
  uint32_t a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
  if (side) org[axis] = B[axis] + B[3 + axis];
  U[a1] = B[3 + a1];
  V[a2] = B[3 + a2];
  grid_tri(org, U, V, 4, q & 3, q >> 2, rr & 1, out);
}

// config 2 (BASELINE.json:8): centre uniform in [0,1)^3, vertices = centre +
// uniform offsets in [-0.005, 0.005)^3.
void soup_tri(uint64_t key, uint64_t t, float* out) {
  float c[3], v[3][3];
  for (uint32_t k = 0; k < 3; ++k) c[k] = u01(rnd64(key, t, k));
  for (uint32_t j = 0; j < 3; ++j)
    for (uint32_t k = 0; k < 3; ++k)
      v[j][k] = c[k] + (u01(rnd64(key, t, 3 + 3 * j + k)) - 0.5f) * 0.01f;
  put_tri(out, v[0], v[1], v[2]);
}

// config 3 (BASELINE.json:9): sphereflake.  Spheres form a 9-ary tree of the
// given depth (radius ratio 1/3).  Each sphere is an octahedron whose 8 faces
// are split into m*m triangles and pushed to the sphere.
const float kFlakeDirs[9][3] = {
    {1.0f, 0.0f, 0.0f},         {0.5f, 0.8660254f, 0.0f},  {-0.5f, 0.8660254f, 0.0f},
    {-1.0f, 0.0f, 0.0f},        {-0.5f, -0.8660254f, 0.0f}, {0.5f, -0.8660254f, 0.0f},
    {0.5196152f, 0.3f, 0.8f},   {-0.5196152f, 0.3f, 0.8f}, {0.0f, -0.6f, 0.8f},
};

inline uint64_t flake_spheres(uint32_t depth) {
  uint64_t n = 0, p = 1;
  for (uint32_t l = 0; l <= depth; ++l) { n += p; p *= 9; }
  return n;
}

// orthonormal basis around unit n (Duff et al. 2017), fixed operation order
inline void onb(const float n[3], float b1[3], float b2[3]) {
  float sign = std::copysign(1.0f, n[2]);
  float a = -1.0f / (sign + n[2]);
  float b = n[0] * n[1] * a;
  b1[0] = 1.0f + sign * n[0] * n[0] * a;
  b1[1] = sign * b;
  b1[2] = -sign * n[0];
  b2[0] = b;
  b2[1] = sign + n[1] * n[1] * a;
  b2[2] = -n[1];
}

void flake_sphere(uint64_t s, float c[3], float& r) {
  // level and index inside the level
  uint32_t level = 0;
  uint64_t off = 0, p = 1;
  while (s >= off + p) { off += p; p *= 9; ++level; }
  uint64_t w = s - off;
  uint32_t digits[16];
  for (uint32_t l = 0; l < level; ++l) { digits[level - 1 - l] = (uint32_t)(w % 9); w /= 9; }
  c[0] = 0.0f; c[1] = 0.0f; c[2] = 0.0f;
  r = 1.0f;
  float axis[3] = {0.0f, 0.0f, 1.0f};
  for (uint32_t l = 0; l < level; ++l) {
    float b1[3], b2[3];
    onb(axis, b1, b2);
    const float* L = kFlakeDirs[digits[l]];
    float d[3];
    for (int k = 0; k < 3; ++k) d[k] = L[0] * b1[k] + L[1] * b2[k] + L[2] * axis[k];
    normalize3(d);
    float rc = r * 0.33333334f;
    float dist = r + rc;
    for (int k = 0; k < 3; ++k) { c[k] = c[k] + d[k] * dist; axis[k] = d[k]; }
    r = rc;
  }
}

inline void flake_vertex(const float c[3], float r, const float sgn[3], uint32_t m,
                         uint32_t row, uint32_t col, float out[3]) {
  // barycentric grid point: i = col, j = row - col, k = m - row  (i + j + k = m)
  float fm = (float)m;
  float n[3] = {sgn[0] * ((float)col / fm), sgn[1] * ((float)(row - col) / fm),
                sgn[2] * ((float)(m - row) / fm)};
  normalize3(n);
  for (int k = 0; k < 3; ++k) out[k] = c[k] + r * n[k];
}

void flake_tri(uint32_t m, uint64_t t, float* out) {
  uint64_t per = 8ull * m * m;
  uint64_t s = t / per;
  uint32_t rem = (uint32_t)(t % per);
  uint32_t face = rem / (m * m), sub = rem % (m * m);
  float c[3], r;
  flake_sphere(s, c, r);
  float sgn[3] = {(face & 1) ? -1.0f : 1.0f, (face & 2) ? -1.0f : 1.0f, (face & 4) ? -1.0f : 1.0f};
  uint32_t row = (uint32_t)std::sqrt((float)sub);
  while (row * row > sub) --row;
  while ((row + 1) * (row + 1) <= sub) ++row;
  uint32_t k = sub - row * row;
  float a[3], b[3], d[3];
  if ((k & 1) == 0) {
    uint32_t col = k >> 1;
    flake_vertex(c, r, sgn, m, row, col, a);
    flake_vertex(c, r, sgn, m, row + 1, col, b);
    flake_vertex(c, r, sgn, m, row + 1, col + 1, d);
  } else {
    uint32_t col = (k - 1) >> 1;
    flake_vertex(c, r, sgn, m, row, col, a);
    flake_vertex(c, r, sgn, m, row + 1, col + 1, b);
    flake_vertex(c, r, sgn, m, row, col + 1, d);
  }
  put_tri(out, a, b, d);
}

// config 4 (BASELINE.json:10): terrain, g x g cells x 2 triangles over
// [0,1]^2 in x,z; height = 5 octaves of hash value noise, smoothstep weights.
inline float terrain_height(uint64_t key, uint32_t i, uint32_t j, uint32_t base) {
  float h = 0.0f, amp = 0.25f;
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
inline void terrain_vertex(uint64_t key, uint32_t g, uint32_t base, uint32_t i, uint32_t j,
                           float out[3]) {
  out[0] = (float)i / (float)g;
  out[1] = terrain_height(key, i, j, base);
  out[2] = (float)j / (float)g;
}
void terrain_tri(uint64_t key, uint32_t g, uint32_t base, uint64_t t, float* out) {
  uint64_t cell = t >> 1;
  uint32_t ci = (uint32_t)(cell % g), cj = (uint32_t)(cell / g);
  float v00[3], v10[3], v11[3], v01[3];
  terrain_vertex(key, g, base, ci, cj, v00);
  terrain_vertex(key, g, base, ci + 1, cj + 1, v11);
  if ((t & 1) == 0) {
    terrain_vertex(key, g, base, ci + 1, cj, v10);
    put_tri(out, v00, v10, v11);
  } else {
    terrain_vertex(key, g, base, ci, cj + 1, v01);
    put_tri(out, v00, v11, v01);
  }
}

// ---------------------------------------------------------------------------
// A2: Morton codes
// ---------------------------------------------------------------------------
inline uint32_t expand10(uint32_t v) {  // 10 bits -> every third bit
  v &= 0x3FFu;
  v = (v | (v << 16)) & 0x030000FFu;
  v = (v | (v << 8)) & 0x0300F00Fu;
  v = (v | (v << 4)) & 0x030C30C3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
inline uint64_t expand21(uint64_t v) {  // 21 bits -> every third bit
  v &= 0x1FFFFFull;
  v = (v | (v << 32)) & 0x001F00000000FFFFull;
  v = (v | (v << 16)) & 0x001F0000FF0000FFull;
  v = (v | (v << 8)) & 0x100F00F00F00F00Full;
  v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}
inline uint64_t morton_from_q(uint32_t bits, uint32_t qx, uint32_t qy, uint32_t qz) {
  if (bits == 10) return ((uint64_t)expand10(qx) << 2) | ((uint64_t)expand10(qy) << 1) | expand10(qz);
  return (expand21(qx) << 2) | (expand21(qy) << 1) | expand21(qz);
}
inline uint32_t quantize(float c, float smin, float scale, float qmax) {
  float x = (c - smin) * scale;
  x = nmax(x, 0.0f);
  x = nmin(x, qmax);
  return (uint32_t)x;  // truncation; x is finite and in [0, qmax]
}

struct Bvh {
  uint64_t n = 0;
  uint32_t key_bits = 30;  // 30 or 63
  float smin[3] = {0, 0, 0}, smax[3] = {0, 0, 0};
  std::vector<float> tris;       // copy of the input, 12 floats per triangle (empty when borrowed)
  const float* T = nullptr;      // the triangles: tris.data(), or the caller's array (orc_build_borrow)
  std::vector<uint64_t> keys;    // sorted Morton codes
  std::vector<uint32_t> idx;     // sorted primitive indices
  std::vector<uint32_t> left, right;  // per internal node, canonical ids
  std::vector<uint32_t> parent;  // per node (2n-1), 0xFFFFFFFF for the root
  std::vector<float> boxes;      // per node, 6 floats: min.xyz max.xyz
  double t_morton = 0, t_sort = 0, t_topo = 0, t_refit = 0;
};

double now_s() {
  using namespace std::chrono;
  return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// A4: delta.  kb = 32 (30-bit keys held in 32 bits) or 64.
inline int delta(const Bvh& b, int64_t i, int64_t j) {
  if (j < 0 || j >= (int64_t)b.n) return -1;
  uint64_t ki = b.keys[i], kj = b.keys[j];
  if (b.key_bits == 30) {
    uint32_t x = (uint32_t)ki ^ (uint32_t)kj;
    if (x) return __builtin_clz(x);
    return 32 + __builtin_clz((uint32_t)i ^ (uint32_t)j);
  }
  uint64_t x = ki ^ kj;
  if (x) return __builtin_clzll(x);
  return 64 + __builtin_clz((uint32_t)i ^ (uint32_t)j);
}

void build(Bvh& b) {
  const uint64_t n = b.n;
  const float* T = b.T;
  double t0 = now_s();
  // A1: scene bounds over all vertices
  if (n) {
    uint32_t kmin[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, kmax[3] = {0, 0, 0};
    for (uint64_t t = 0; t < n; ++t)
      for (int v = 0; v < 3; ++v)
        for (int k = 0; k < 3; ++k) {
          uint32_t key = ord_key(T[t * 12 + v * 4 + k]);
          kmin[k] = std::min(kmin[k], key);
          kmax[k] = std::max(kmax[k], key);
        }
    for (int k = 0; k < 3; ++k) { b.smin[k] = ord_unkey(kmin[k]); b.smax[k] = ord_unkey(kmax[k]); }
  }
  // A2: Morton codes
  const uint32_t bits = b.key_bits == 30 ? 10 : 21;
  const float cells = b.key_bits == 30 ? 1024.0f : 2097152.0f;
  const float qmax = cells - 1.0f;
  float scale[3];
  for (int k = 0; k < 3; ++k) {
    float ext = b.smax[k] - b.smin[k];
    scale[k] = (ext > 0.0f) ? (cells / ext) : 0.0f;
  }
  std::vector<std::pair<uint64_t, uint32_t>> kv(n);
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < (int64_t)n; ++t) {
    const float* p = T + t * 12;
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
      float lo = omin(omin(p[k], p[4 + k]), p[8 + k]);
      float hi = omax(omax(p[k], p[4 + k]), p[8 + k]);
      float c = 0.5f * (lo + hi);
      q[k] = quantize(c, b.smin[k], scale[k], qmax);
    }
    kv[t] = {morton_from_q(bits, q[0], q[1], q[2]), (uint32_t)t};
  }
  double t1 = now_s();
  // A3: ascending by (morton, index); indices start ascending so a stable
  // sort on the code alone is the same order.
  auto cmp = [](const std::pair<uint64_t, uint32_t>& a, const std::pair<uint64_t, uint32_t>& c) {
    return a.first < c.first;
  };
#ifdef _OPENMP
  if (omp_get_max_threads() > 1) __gnu_parallel::stable_sort(kv.begin(), kv.end(), cmp);
  else std::stable_sort(kv.begin(), kv.end(), cmp);
#else
  std::stable_sort(kv.begin(), kv.end(), cmp);
#endif
  b.keys.resize(n);
  b.idx.resize(n);
  for (uint64_t t = 0; t < n; ++t) { b.keys[t] = kv[t].first; b.idx[t] = kv[t].second; }
  kv.clear();
  kv.shrink_to_fit();
  double t2 = now_s();
  // A4: Karras 2012 topology
  const uint64_t nodes = n ? 2 * n - 1 : 0;
  b.left.assign(n ? n - 1 : 0, 0);
  b.right.assign(n ? n - 1 : 0, 0);
  b.parent.assign(nodes, 0xFFFFFFFFu);
  const uint32_t leaf0 = (uint32_t)(n ? n - 1 : 0);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n - 1; ++i) {
    int d = (delta(b, i, i + 1) - delta(b, i, i - 1)) < 0 ? -1 : 1;
    int dmin = delta(b, i, i - d);
    int64_t lmax = 2;
    while (delta(b, i, i + lmax * d) > dmin) lmax *= 2;
    int64_t l = 0;
    for (int64_t t = lmax / 2; t >= 1; t /= 2)
      if (delta(b, i, i + (l + t) * d) > dmin) l += t;
    int64_t j = i + l * d;
    int dnode = delta(b, i, j);
    int64_t s = 0;
    for (int64_t div = 2;; div *= 2) {
      int64_t t = (l + div - 1) / div;
      if (delta(b, i, i + (s + t) * d) > dnode) s += t;
      if (t <= 1) break;
    }
    int64_t gamma = i + s * d + std::min(d, 0);
    uint32_t lc = (std::min(i, j) == gamma) ? (uint32_t)(leaf0 + gamma) : (uint32_t)gamma;
    uint32_t rc = (std::max(i, j) == gamma + 1) ? (uint32_t)(leaf0 + gamma + 1) : (uint32_t)(gamma + 1);
    b.left[i] = lc;
    b.right[i] = rc;
    b.parent[lc] = (uint32_t)i;
    b.parent[rc] = (uint32_t)i;
  }
  double t3 = now_s();
  // A5: leaf boxes, then bottom-up unions.  Done here with a per-node counter
  // exactly like the GPU's atomic-flag walk, but sequential per leaf chain.
  b.boxes.assign(nodes * 6, 0.0f);
  std::vector<uint8_t> visits(n ? n - 1 : 0, 0);
#pragma omp parallel for schedule(static)
  for (int64_t j = 0; j < (int64_t)n; ++j) {
    const float* p = T + (uint64_t)b.idx[j] * 12;
    float* bx = &b.boxes[(uint64_t)(leaf0 + j) * 6];
    for (int k = 0; k < 3; ++k) {
      bx[k] = omin(omin(p[k], p[4 + k]), p[8 + k]);
      bx[3 + k] = omax(omax(p[k], p[4 + k]), p[8 + k]);
    }
    // walk up: the first arriver at a node stops, the second unions the two
    // child boxes (release/acquire on the counter orders the box stores)
    uint32_t node = b.parent[leaf0 + j];
    while (node != 0xFFFFFFFFu) {
      if (__atomic_fetch_add(&visits[node], (uint8_t)1, __ATOMIC_ACQ_REL) == 0) break;
      const float* a = &b.boxes[(uint64_t)b.left[node] * 6];
      const float* c = &b.boxes[(uint64_t)b.right[node] * 6];
      float* o = &b.boxes[(uint64_t)node * 6];
      for (int k = 0; k < 3; ++k) { o[k] = omin(a[k], c[k]); o[3 + k] = omax(a[3 + k], c[3 + k]); }
      node = b.parent[node];
    }
  }
  double t4 = now_s();
  b.t_morton = t1 - t0; b.t_sort = t2 - t1; b.t_topo = t3 - t2; b.t_refit = t4 - t3;
}

// ---------------------------------------------------------------------------
// A7: Moeller-Trumbore, two-sided, explicit fmaf chains in one fixed order
// ---------------------------------------------------------------------------
struct Ray { float o[3], tmin, d[3], tmax; };
struct Hit { float t, u, v; int32_t prim; };

inline bool tri_test(const float* p, const Ray& r, float& t, float& u, float& v) {
  float e1x = p[4] - p[0], e1y = p[5] - p[1], e1z = p[6] - p[2];
  float e2x = p[8] - p[0], e2y = p[9] - p[1], e2z = p[10] - p[2];
  float px = fmaf(r.d[1], e2z, -(r.d[2] * e2y));
  float py = fmaf(r.d[2], e2x, -(r.d[0] * e2z));
  float pz = fmaf(r.d[0], e2y, -(r.d[1] * e2x));
  float det = fmaf(e1x, px, fmaf(e1y, py, e1z * pz));
  if (det == 0.0f) return false;
  float inv = 1.0f / det;
  float sx = r.o[0] - p[0], sy = r.o[1] - p[1], sz = r.o[2] - p[2];
  u = fmaf(sx, px, fmaf(sy, py, sz * pz)) * inv;
  float qx = fmaf(sy, e1z, -(sz * e1y));
  float qy = fmaf(sz, e1x, -(sx * e1z));
  float qz = fmaf(sx, e1y, -(sy * e1x));
  v = fmaf(r.d[0], qx, fmaf(r.d[1], qy, r.d[2] * qz)) * inv;
  t = fmaf(e2x, qx, fmaf(e2y, qy, e2z * qz)) * inv;
  return (u >= 0.0f) && (v >= 0.0f) && ((u + v) <= 1.0f);
}

// ---------------------------------------------------------------------------
// A7 option: watertight ray/triangle test (Woop, Benthin & Wald 2013).  The ray
// is sheared so that its dominant axis becomes +z; the three scaled edge
// functions U, V, W decide the hit, recomputed in double precision when one of
// them is exactly zero (a ray through an edge or a vertex), so no ray slips
// between two triangles sharing that edge.  Plain * + - / in one fixed order
// (no contraction); barycentrics are returned in the Moeller-Trumbore
// convention (u weights v1, v weights v2).
// ---------------------------------------------------------------------------
struct WtRay { int kx, ky, kz; float Sx, Sy, Sz; };

inline WtRay wt_setup(const Ray& r) {
  WtRay w;
  w.kz = 0;
  if (std::fabs(r.d[1]) > std::fabs(r.d[w.kz])) w.kz = 1;
  if (std::fabs(r.d[2]) > std::fabs(r.d[w.kz])) w.kz = 2;
  w.kx = (w.kz + 1) % 3;
  w.ky = (w.kx + 1) % 3;
  if (r.d[w.kz] < 0.0f) std::swap(w.kx, w.ky);
  w.Sx = r.d[w.kx] / r.d[w.kz];
  w.Sy = r.d[w.ky] / r.d[w.kz];
  w.Sz = 1.0f / r.d[w.kz];
  return w;
}

inline bool tri_test_wt(const float* p, const Ray& r, const WtRay& w, float& t, float& u, float& v) {
  const float A[3] = {p[0] - r.o[0], p[1] - r.o[1], p[2] - r.o[2]};
  const float B[3] = {p[4] - r.o[0], p[5] - r.o[1], p[6] - r.o[2]};
  const float C[3] = {p[8] - r.o[0], p[9] - r.o[1], p[10] - r.o[2]};
  const float Ax = A[w.kx] - w.Sx * A[w.kz], Ay = A[w.ky] - w.Sy * A[w.kz];
  const float Bx = B[w.kx] - w.Sx * B[w.kz], By = B[w.ky] - w.Sy * B[w.kz];
  const float Cx = C[w.kx] - w.Sx * C[w.kz], Cy = C[w.ky] - w.Sy * C[w.kz];
  float U = Cx * By - Cy * Bx;
  float V = Ax * Cy - Ay * Cx;
  float W = Bx * Ay - By * Ax;
  if (U == 0.0f || V == 0.0f || W == 0.0f) {
    U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
    V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
    W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
  }
  if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
  const float det = (U + V) + W;
  if (det == 0.0f) return false;
  const float Az = w.Sz * A[w.kz], Bz = w.Sz * B[w.kz], Cz = w.Sz * C[w.kz];
  const float T = (U * Az + V * Bz) + W * Cz;
  const float inv = 1.0f / det;
  t = T * inv;
  u = V * inv;
  v = W * inv;
  return true;
}

// A6 acceptance rule: closest t in (tmin, tmax); on equal t the smaller
// primitive index wins.
inline bool accept(float t, int32_t prim, float tmin, float best_t, int32_t best_prim) {
  return (t > tmin) && ((t < best_t) || (t == best_t && best_prim >= 0 && prim < best_prim));
}

const float kSlabPad = 1.0000004f;  // 1 + 3 ulp: keeps the slab test conservative

// The two plane distances of an axis are combined with NaN-PROPAGATING min/max (pmin/pmax),
// the axes and tmin / best_t with NaN-ignoring ones (nmin/nmax).  idir[k] is +inf for a
// direction component that is exactly +-0: its plane distances are then -inf / +inf, or 0 * inf
// = NaN for an origin exactly on the plane, and a NaN drops the axis from tnear and tfar.  So a
// ray parallel to a slab passes when its origin lies in the CLOSED slab and fails otherwise
// ((-inf,-inf) -> tfar = -inf, (+inf,+inf) -> tnear = +inf).  (A clamped 1/0 = FLT_MAX, as the
// first version had it, gives 0 * FLT_MAX = 0 for an origin ON a face and culled nodes the ray
// does enter.)
inline float pmin(float a, float b) { return (a != a || b != b) ? std::numeric_limits<float>::quiet_NaN() : (a < b ? a : b); }
inline float pmax(float a, float b) { return (a != a || b != b) ? std::numeric_limits<float>::quiet_NaN() : (a > b ? a : b); }
inline bool slab(const float* bx, const Ray& r, const float idir[3], float best_t, float& tn) {
  float t0x = (bx[0] - r.o[0]) * idir[0], t1x = (bx[3] - r.o[0]) * idir[0];
  float t0y = (bx[1] - r.o[1]) * idir[1], t1y = (bx[4] - r.o[1]) * idir[1];
  float t0z = (bx[2] - r.o[2]) * idir[2], t1z = (bx[5] - r.o[2]) * idir[2];
  tn = nmax(nmax(pmin(t0x, t1x), pmin(t0y, t1y)), nmax(pmin(t0z, t1z), r.tmin));
  float tf = nmin(nmin(pmax(t0x, t1x), pmax(t0y, t1y)), nmin(pmax(t0z, t1z), best_t));
  tf = tf * kSlabPad;
  return tn <= tf;
}

// 1/d clamped to +-FLT_MAX (a denormal component then gives 0 * FLT_MAX = 0, not 0 * inf = NaN,
// for an origin on a box face); +inf for a component that is exactly +-0
inline float inv_dir(float d) {
  const float fmx = std::numeric_limits<float>::max();
  if (d == 0.0f) return std::numeric_limits<float>::infinity();
  return nmin(nmax(1.0f / d, -fmx), fmx);
}

// A6, second half of the acceptance rule: the hit must be consistent with the triangle's own
// box (the tree's leaf box: ordered min/max of the three vertices): slab() passes on it and its
// entry distance satisfies tnear <= t * kSlabPad.  A node is culled when tnear > min(tfar,
// best_t) * pad, ancestors contain the leaf box and slab() is monotone in the box, so a
// triangle with an acceptable hit is never below a culled node: the result is min over the
// triangles with a consistent hit by (t, prim), whatever the tree and the visiting order, and
// the all-pairs intersector below returns the same bytes as any BVH traversal.
inline bool own_box_admits(const float* p, const Ray& r, const float idir[3], float best_t, float t) {
  float bx[6];
  for (int k = 0; k < 3; ++k) {
    bx[k] = omin(omin(p[k], p[4 + k]), p[8 + k]);
    bx[3 + k] = omax(omax(p[k], p[4 + k]), p[8 + k]);
  }
  float tn;
  return slab(bx, r, idir, best_t, tn) && tn <= t * kSlabPad;
}

enum { TRACE_CLOSEST = 0, TRACE_ANY = 1, TRACE_WATERTIGHT = 2 };
const int32_t kAnyHitPrim = 0x7FFFFFFF;

inline Hit miss_hit() { Hit h; h.t = std::numeric_limits<float>::infinity(); h.u = 0; h.v = 0; h.prim = -1; return h; }

Hit trace_one(const Bvh& b, const Ray& r, uint32_t flags, uint64_t& nodes, uint64_t& tests) {
  Hit h = miss_hit();
  const uint64_t n = b.n;
  if (n == 0) return h;
  float best_t = r.tmax;
  int32_t best_prim = -1;
  float bu = 0, bv = 0;
  const float idir[3] = {inv_dir(r.d[0]), inv_dir(r.d[1]), inv_dir(r.d[2])};
  const uint32_t leaf0 = (uint32_t)(n - 1);
  uint32_t stack[128];
  int sp = 0;
  uint32_t node = 0;  // root: internal 0, or leaf 0 when n == 1 (leaf0 == 0)
  bool any = (flags & TRACE_ANY) != 0;
  const bool wt = (flags & TRACE_WATERTIGHT) != 0;
  WtRay wr = {0, 1, 2, 0, 0, 0};
  if (wt) wr = wt_setup(r);
  for (;;) {
    if (node >= leaf0) {
      uint32_t prim = b.idx[node - leaf0];
      float t, u, v;
      ++tests;
      const float* tp = b.T + (uint64_t)prim * 12;
      if ((wt ? tri_test_wt(tp, r, wr, t, u, v) : tri_test(tp, r, t, u, v)) &&
          accept(t, (int32_t)prim, r.tmin, best_t, best_prim) && own_box_admits(tp, r, idir, best_t, t)) {
        best_t = t; best_prim = (int32_t)prim; bu = u; bv = v;
        if (any) break;
      }
    } else {
      ++nodes;
      uint32_t c0 = b.left[node], c1 = b.right[node];
      float tn0, tn1;
      bool h0 = slab(&b.boxes[(uint64_t)c0 * 6], r, idir, best_t, tn0);
      bool h1 = slab(&b.boxes[(uint64_t)c1 * 6], r, idir, best_t, tn1);
      if (h0 && h1) {
        if (tn1 < tn0) { stack[sp++] = c0; node = c1; }
        else { stack[sp++] = c1; node = c0; }
        continue;
      }
      if (h0) { node = c0; continue; }
      if (h1) { node = c1; continue; }
    }
    if (sp == 0) break;
    node = stack[--sp];
  }
  if (best_prim >= 0) {
    if (any) { h.t = 0; h.u = 0; h.v = 0; h.prim = kAnyHitPrim; }
    else { h.t = best_t; h.u = bu; h.v = bv; h.prim = best_prim; }
  }
  return h;
}

}  // namespace

// ===========================================================================
// C ABI (ctypes).  Names are orc_*; none of them exist in the product library.
// ===========================================================================
ORC_API void orc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
ORC_API int orc_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

ORC_API uint32_t orc_morton30(uint32_t x, uint32_t y, uint32_t z) { return (uint32_t)morton_from_q(10, x, y, z); }
ORC_API uint64_t orc_morton63(uint32_t x, uint32_t y, uint32_t z) { return morton_from_q(21, x, y, z); }

ORC_API uint64_t orc_scene_size(int kind, uint64_t p0, uint64_t p1) {
  switch (kind) {
    case SCENE_CORNELL: return 1024;
    case SCENE_SOUP: return p0;
    case SCENE_SPHEREFLAKE: return flake_spheres((uint32_t)p0) * 8ull * p1 * p1;
    case SCENE_TERRAIN: return 2ull * p0 * p0;
  }
  return 0;
}

// triangles [first, first+count) of the scene into out (12 floats each).
// p0/p1: soup (n, -), sphereflake (depth, m), terrain (g, base spacing).
ORC_API int orc_gen_scene(int kind, uint64_t p0, uint64_t p1, uint64_t seed, uint64_t first,
                          uint64_t count, float* out) {
  if (kind < 0 || kind > 3) return -1;
  const uint64_t key = stream_key(seed, kind == SCENE_SOUP ? STREAM_SOUP : STREAM_TERRAIN);
#pragma omp parallel for schedule(static)
  for (int64_t k = 0; k < (int64_t)count; ++k) {
    uint64_t t = first + k;
    float* o = out + (uint64_t)k * 12;
    switch (kind) {
      case SCENE_CORNELL: cornell_tri(t, o); break;
      case SCENE_SOUP: soup_tri(key, t, o); break;
      case SCENE_SPHEREFLAKE: flake_tri((uint32_t)p1, t, o); break;
      case SCENE_TERRAIN: terrain_tri(key, (uint32_t)p0, (uint32_t)(p1 ? p1 : 1024), t, o); break;
    }
  }
  return 0;
}

// A8 primary rays.  cam = eye.xyz, corner.xyz, du.xyz, dv.xyz (12 floats);
// pixel (x,y) looks through corner + (x+0.5) du + (y+0.5) dv.  Row-major.
ORC_API void orc_gen_primary_rays(const float* cam, uint32_t width, uint32_t height, uint64_t first,
                                  uint64_t count, float* out) {
  (void)height;
#pragma omp parallel for schedule(static)
  for (int64_t k = 0; k < (int64_t)count; ++k) {
    uint64_t i = first + k;
    float fx = (float)(uint32_t)(i % width) + 0.5f, fy = (float)(uint32_t)(i / width) + 0.5f;
    float d[3];
    for (int a = 0; a < 3; ++a) d[a] = (cam[3 + a] + cam[6 + a] * fx + cam[9 + a] * fy) - cam[a];
    normalize3(d);
    float* o = out + (uint64_t)k * 8;
    o[0] = cam[0]; o[1] = cam[1]; o[2] = cam[2]; o[3] = 0.0f;
    o[4] = d[0]; o[5] = d[1]; o[6] = d[2]; o[7] = std::numeric_limits<float>::infinity();
  }
}

// A8 random rays (config 5): origin uniform in the box, direction uniform on
// the sphere.  box = min.xyz max.xyz.
ORC_API void orc_gen_random_rays(const float* box, uint64_t seed, uint64_t first, uint64_t count,
                                 float* out) {
  const uint64_t key = stream_key(seed, STREAM_RAYS);
#pragma omp parallel for schedule(static)
  for (int64_t k = 0; k < (int64_t)count; ++k) {
    uint64_t i = first + k;
    float* o = out + (uint64_t)k * 8;
    for (uint32_t a = 0; a < 3; ++a)
      o[a] = box[a] + u01(rnd64(key, i, a)) * (box[3 + a] - box[a]);
    o[3] = 0.0f;
    sphere_sample(key, i, 4, o + 4);
    o[7] = std::numeric_limits<float>::infinity();
  }
}

// A8 ambient-occlusion rays: spp rays per primary ray, from its hit point.
ORC_API void orc_gen_ao_rays(const float* tris, const float* rays, const float* hits, uint64_t n,
                             uint32_t spp, uint64_t seed, float radius, float* out) {
  const uint64_t key = stream_key(seed, STREAM_AO);
  const float eps = radius * 1e-3f;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    const float* r = rays + (uint64_t)i * 8;
    const float* h = hits + (uint64_t)i * 4;
    int32_t prim;
    std::memcpy(&prim, h + 3, 4);
    if (prim < 0) {
      for (uint32_t s = 0; s < spp; ++s) {
        float* o = out + ((uint64_t)i * spp + s) * 8;
        o[0] = 0; o[1] = 0; o[2] = 0; o[3] = 0; o[4] = 0; o[5] = 0; o[6] = 1.0f; o[7] = -1.0f;
      }
      continue;
    }
    const float* p = tris + (uint64_t)prim * 12;
    float e1[3] = {p[4] - p[0], p[5] - p[1], p[6] - p[2]};
    float e2[3] = {p[8] - p[0], p[9] - p[1], p[10] - p[2]};
    float nrm[3] = {fmaf(e1[1], e2[2], -(e1[2] * e2[1])), fmaf(e1[2], e2[0], -(e1[0] * e2[2])),
                    fmaf(e1[0], e2[1], -(e1[1] * e2[0]))};
    normalize3(nrm);
    float dn = fmaf(nrm[0], r[4], fmaf(nrm[1], r[5], nrm[2] * r[6]));
    if (dn > 0.0f) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
    float org[3];
    for (int a = 0; a < 3; ++a) org[a] = fmaf(h[0], r[4 + a], r[a]) + nrm[a] * eps;
    for (uint32_t s = 0; s < spp; ++s) {
      float sv[3], d[3];
      sphere_sample(key, (uint64_t)i * spp + s, 0, sv);
      for (int a = 0; a < 3; ++a) d[a] = nrm[a] + sv[a];
      float l2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
      if (l2 < 1e-12f) { d[0] = nrm[0]; d[1] = nrm[1]; d[2] = nrm[2]; }
      else normalize3(d);
      float* o = out + ((uint64_t)i * spp + s) * 8;
      o[0] = org[0]; o[1] = org[1]; o[2] = org[2]; o[3] = 0.0f;
      o[4] = d[0]; o[5] = d[1]; o[6] = d[2]; o[7] = radius;
    }
  }
}

// §8(f).1 shading: grey = |n.d| * (1 - occluded/spp) with the geometric normal of the hit
// triangle (ORIGINAL triangle order), misses black; ao_hits may be NULL (spp = 0).  Same
// operation order as the product's shade kernel, so the comparison is bit for bit.
ORC_API void orc_shade(const float* tris, const float* rays, const float* hits, const float* ao_hits,
                       uint64_t n, uint32_t spp, float* rgb) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    const float* r = rays + (uint64_t)i * 8;
    int32_t prim;
    std::memcpy(&prim, hits + (uint64_t)i * 4 + 3, 4);
    float c = 0.0f;
    if (prim >= 0) {
      const float* p = tris + (uint64_t)prim * 12;
      float e1[3] = {p[4] - p[0], p[5] - p[1], p[6] - p[2]};
      float e2[3] = {p[8] - p[0], p[9] - p[1], p[10] - p[2]};
      float nrm[3] = {fmaf(e1[1], e2[2], -(e1[2] * e2[1])), fmaf(e1[2], e2[0], -(e1[0] * e2[2])),
                      fmaf(e1[0], e2[1], -(e1[1] * e2[0]))};
      normalize3(nrm);
      float dn = fmaf(nrm[0], r[4], fmaf(nrm[1], r[5], nrm[2] * r[6]));
      if (dn > 0.0f) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
      c = std::fabs(fmaf(nrm[0], r[4], fmaf(nrm[1], r[5], nrm[2] * r[6])));
      if (ao_hits && spp > 0) {
        uint32_t occ = 0;
        for (uint32_t s = 0; s < spp; ++s) {
          int32_t ap;
          std::memcpy(&ap, ao_hits + ((uint64_t)i * spp + s) * 4 + 3, 4);
          occ += ap >= 0 ? 1u : 0u;
        }
        c = c * (1.0f - (float)occ / (float)spp);
      }
    }
    rgb[3 * i + 0] = c; rgb[3 * i + 1] = c; rgb[3 * i + 2] = c;
  }
}

ORC_API void* orc_build(const float* tris, uint64_t n, uint32_t key_bits) {
  if (key_bits != 30 && key_bits != 63) return nullptr;
  Bvh* b = new Bvh;
  b->n = n;
  b->key_bits = key_bits;
  b->tris.assign(tris, tris + n * 12);
  b->T = b->tris.data();
  build(*b);
  return b;
}
// the same build without the copy of the triangles: the caller keeps `tris` alive until
// orc_free (the 100M-triangle parity test: 4.8 GB less)
ORC_API void* orc_build_borrow(const float* tris, uint64_t n, uint32_t key_bits) {
  if (key_bits != 30 && key_bits != 63) return nullptr;
  Bvh* b = new Bvh;
  b->n = n;
  b->key_bits = key_bits;
  b->T = tris;
  build(*b);
  return b;
}
ORC_API void orc_free(void* h) { delete (Bvh*)h; }

// out arrays may be NULL.  keys: n u64; idx: n u32; left/right: n-1 u32;
// parent: 2n-1 u32; boxes: (2n-1) x 6 floats; bounds: 6 floats.
ORC_API void orc_read(const void* h, uint64_t* keys, uint32_t* idx, uint32_t* left, uint32_t* right,
                      uint32_t* parent, float* boxes, float* bounds) {
  const Bvh& b = *(const Bvh*)h;
  if (keys) std::copy(b.keys.begin(), b.keys.end(), keys);
  if (idx) std::copy(b.idx.begin(), b.idx.end(), idx);
  if (left) std::copy(b.left.begin(), b.left.end(), left);
  if (right) std::copy(b.right.begin(), b.right.end(), right);
  if (parent) std::copy(b.parent.begin(), b.parent.end(), parent);
  if (boxes) std::copy(b.boxes.begin(), b.boxes.end(), boxes);
  if (bounds) for (int k = 0; k < 3; ++k) { bounds[k] = b.smin[k]; bounds[3 + k] = b.smax[k]; }
}
// seconds: morton(+bounds), sort, topology, refit
ORC_API void orc_build_times(const void* h, double* out4) {
  const Bvh& b = *(const Bvh*)h;
  out4[0] = b.t_morton; out4[1] = b.t_sort; out4[2] = b.t_topo; out4[3] = b.t_refit;
}

// hits: n x 4 floats (t,u,v, prim as int bits).  counters (may be NULL):
// [0] internal nodes fetched (each = two 32 B child boxes), [1] triangles tested.
ORC_API void orc_trace(const void* h, const float* rays, uint64_t n, uint32_t flags, float* hits,
                       uint64_t* counters) {
  const Bvh& b = *(const Bvh*)h;
  uint64_t nodes = 0, tests = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : nodes, tests)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    Ray r;
    std::memcpy(&r, rays + (uint64_t)i * 8, 32);
    Hit hh = trace_one(b, r, flags, nodes, tests);
    std::memcpy(hits + (uint64_t)i * 4, &hh, 16);
  }
  if (counters) { counters[0] = nodes; counters[1] = tests; }
}

// ---------------------------------------------------------------------------
// A5w: collapse of the binary tree to a 4-wide BVH (the
// fixed rule is written out in DESIGN.md §4).  A wide node starts with the two
// children of its binary node; up to two times the slot chosen by
//   1. among slots with more than leaf_max triangles, the largest half area;
//      if there is none,
//   2. among slots with more than one triangle, the largest half area;
//   ties -> leftmost slot; half area = (dx*dy + dy*dz) + dz*dx in float
// is replaced in place by its two children.  Slots with <= leaf_max triangles
// are leaf ranges [first, first+cnt) of the sorted order, the others are wide
// nodes themselves.  Output: the wide nodes reachable from the root in
// pre-order (children in slot order), so two implementations can be compared
// array for array.
//   boxes: cap x 4 x 6 floats (zeros for empty slots)
//   desc:  cap x 4 x 3 u32: kind (0 empty, 1 leaf, 2 node), first, cnt for a
//          leaf; kind, pre-order index of the child, 0 for a node
// Returns the number of wide nodes (call with cap = 0 to size the arrays).
// ---------------------------------------------------------------------------
namespace {
struct WSlot { uint32_t node, first, cnt; };

float half_area(const float* bx) {
  float dx = bx[3] - bx[0], dy = bx[4] - bx[1], dz = bx[5] - bx[2];
  return (dx * dy + dy * dz) + dz * dx;
}

struct Collapser {
  const Bvh& b;
  uint32_t leaf_max;
  uint64_t cap;
  float* boxes;
  uint32_t* desc;
  std::vector<uint32_t> first, cnt;
  uint64_t count = 0;

  void ranges() {
    const uint64_t n = b.n, nodes = 2 * n - 1;
    const uint32_t leaf0 = (uint32_t)(n - 1);
    first.assign(nodes, 0);
    cnt.assign(nodes, 0);
    for (uint64_t j = 0; j < n; ++j) { first[leaf0 + j] = (uint32_t)j; cnt[leaf0 + j] = 1; }
    if (n < 2) return;
    std::vector<uint32_t> stack{0};
    std::vector<uint8_t> seen(n - 1, 0);
    while (!stack.empty()) {
      uint32_t i = stack.back();
      if (!seen[i]) {
        seen[i] = 1;
        if (b.left[i] < leaf0) stack.push_back(b.left[i]);
        if (b.right[i] < leaf0) stack.push_back(b.right[i]);
      } else {
        stack.pop_back();
        first[i] = first[b.left[i]];
        cnt[i] = cnt[b.left[i]] + cnt[b.right[i]];
      }
    }
  }
  WSlot slot_of(uint32_t node) const { return WSlot{node, first[node], cnt[node]}; }

  uint64_t emit(uint32_t i) {
    const uint64_t me = count++;
    WSlot s[4];
    int ns = 2;
    s[0] = slot_of(b.left[i]);
    s[1] = slot_of(b.right[i]);
    for (int it = 0; it < 2; ++it) {
      int best = -1;
      float best_area = 0.0f;
      for (int k = 0; k < ns; ++k)
        if (s[k].cnt > leaf_max) {
          float a = half_area(&b.boxes[(uint64_t)s[k].node * 6]);
          if (best < 0 || a > best_area) { best = k; best_area = a; }
        }
      if (best < 0)
        for (int k = 0; k < ns; ++k)
          if (s[k].cnt > 1) {
            float a = half_area(&b.boxes[(uint64_t)s[k].node * 6]);
            if (best < 0 || a > best_area) { best = k; best_area = a; }
          }
      if (best < 0) break;
      const uint32_t x = s[best].node;
      for (int k = ns; k > best + 1; --k) s[k] = s[k - 1];
      s[best] = slot_of(b.left[x]);
      s[best + 1] = slot_of(b.right[x]);
      ++ns;
    }
    for (int k = 0; k < 4; ++k) {
      uint32_t d[3] = {0, 0, 0};
      float bx[6] = {0, 0, 0, 0, 0, 0};
      if (k < ns) {
        std::memcpy(bx, &b.boxes[(uint64_t)s[k].node * 6], 24);
        if (s[k].cnt <= leaf_max) { d[0] = 1; d[1] = s[k].first; d[2] = s[k].cnt; }
        else { d[0] = 2; d[1] = (uint32_t)emit(s[k].node); }
      }
      if (me < cap) {
        std::memcpy(boxes + (me * 4 + k) * 6, bx, 24);
        std::memcpy(desc + (me * 4 + k) * 3, d, 12);
      }
    }
    return me;
  }
};
}  // namespace

ORC_API uint64_t orc_collapse4(const void* h, uint32_t leaf_max, uint64_t cap, float* boxes,
                               uint32_t* desc) {
  const Bvh& b = *(const Bvh*)h;
  if (b.n <= leaf_max) return 0;  // the whole scene is one leaf: no wide node
  Collapser c{b, leaf_max, cap, boxes, desc, {}, {}, 0};
  c.ranges();
  c.emit(0);
  return c.count;
}

// ---------------------------------------------------------------------------
// A5w, second wide layout: the compressed 8-wide BVH (80-byte nodes, 8-bit child boxes;
// Ylitie, Karras & Laine 2017).  Collapse rule, top-down from the binary tree: a wide node
// starts from the two children of its binary node and keeps replacing the child with the
// largest half surface area ((dx*dy + dy*dz) + dz*dx, no contraction; ties to the leftmost)
// among those holding more than 3 triangles by that child's own two children, in place, until
// it has eight children or none is left; then, while slots are free, it opens the largest child
// that holds 2 or 3 triangles (tighter leaf boxes at no cost in nodes).  Children of at most 3
// triangles are leaves.
// Slots: left to right, each child takes the free slot t maximising
// ((t&4 ? ox : -ox) + (t&2 ? oy : -oy)) + (t&1 ? oz : -oz), o = centre of the child's box -
// centre of the node's box (centre = 0.5f * (lo + hi)), ties to the lowest slot.  Child boxes
// are quantised outward on a grid of 2^e cells from the node's min corner (double-precision
// floor / ceil).  Numbering (canonical): breadth first; a node's internal children are
// consecutive from child_base in slot order, its leaf triangles consecutive from tri_base in
// slot order, both allocated in the order the nodes are visited.
//   nodes: cap x 80 bytes (layout in include/gpu_raytracer.h, bvh_read_wide8)
//   tris:  n x 12 floats: v0.xyz, prim bits, v1.xyz, 0, v2.xyz, 0 in wide-node order
// Returns the number of wide nodes (cap = 0: count only); 0 when n <= 3.
// ---------------------------------------------------------------------------
namespace {
struct Collapser8 {
  const Bvh& b;
  uint32_t leaf0 = 0;
  std::vector<uint32_t> first, cnt;

  void ranges() {
    const uint64_t n = b.n, nodes = 2 * n - 1;
    leaf0 = (uint32_t)(n - 1);
    first.assign(nodes, 0);
    cnt.assign(nodes, 0);
    for (uint64_t j = 0; j < n; ++j) { first[leaf0 + j] = (uint32_t)j; cnt[leaf0 + j] = 1; }
    std::vector<uint32_t> stack{0};
    std::vector<uint8_t> seen(n - 1, 0);
    while (!stack.empty()) {
      uint32_t i = stack.back();
      if (!seen[i]) {
        seen[i] = 1;
        if (b.left[i] < leaf0) stack.push_back(b.left[i]);
        if (b.right[i] < leaf0) stack.push_back(b.right[i]);
      } else {
        stack.pop_back();
        first[i] = first[b.left[i]];
        cnt[i] = cnt[b.left[i]] + cnt[b.right[i]];
      }
    }
  }

  static uint32_t quant_exponent(double ext) {
    if (!(ext > 0.0)) return 0u;
    int e;
    std::frexp(ext / 255.0, &e);
    while (std::ldexp(255.0, e - 1) >= ext) --e;
    while (std::ldexp(255.0, e) < ext) ++e;
    if (e < -126) e = -126;
    if (e > 127) e = 127;
    return (uint32_t)(e + 127);
  }
  static uint32_t quant_lo(float v, float p, uint32_t eb) {
    if (eb == 0u) return 0u;
    const double cell = std::ldexp(1.0, (int)eb - 127), d = (double)v - (double)p;
    double q = std::floor(d / cell);
    q = std::fmin(std::fmax(q, 0.0), 255.0);
    while (q > 0.0 && (double)p + q * cell > (double)v) q -= 1.0;
    return (uint32_t)q;
  }
  static uint32_t quant_hi(float v, float p, uint32_t eb) {
    if (eb == 0u) return 0u;
    const double cell = std::ldexp(1.0, (int)eb - 127), d = (double)v - (double)p;
    double q = std::ceil(d / cell);
    q = std::fmin(std::fmax(q, 0.0), 255.0);
    while (q < 255.0 && (double)p + q * cell < (double)v) q += 1.0;
    return (uint32_t)q;
  }

  // breadth-first; returns the node count; writes when nodes_out != nullptr
  uint64_t run(uint8_t* nodes_out, float* tris_out) {
    std::vector<uint32_t> queue{0};  // canonical binary node ids; wide node index = queue position
    uint32_t next_tri = 0;
    for (size_t at = 0; at < queue.size(); ++at) {
      const uint32_t node = queue[at];
      uint32_t s[8];
      int ns = 2;
      s[0] = b.left[node];
      s[1] = b.right[node];
      while (ns < 8) {
        int best = -1;
        float best_area = 0.0f;
        for (int k = 0; k < ns; ++k)
          if (cnt[s[k]] > 3) {
            const float a = half_area(&b.boxes[(uint64_t)s[k] * 6]);
            if (best < 0 || a > best_area) { best = k; best_area = a; }
          }
        if (best < 0)  // nothing left that must be opened: spend the free slots on tighter leaves
          for (int k = 0; k < ns; ++k)
            if (cnt[s[k]] > 1) {
              const float a = half_area(&b.boxes[(uint64_t)s[k] * 6]);
              if (best < 0 || a > best_area) { best = k; best_area = a; }
            }
        if (best < 0) break;
        const uint32_t x = s[best];
        for (int k = ns; k > best + 1; --k) s[k] = s[k - 1];
        s[best] = b.left[x];
        s[best + 1] = b.right[x];
        ++ns;
      }
      float lo[3], hi[3];
      for (int a = 0; a < 3; ++a) { lo[a] = b.boxes[(uint64_t)s[0] * 6 + a]; hi[a] = b.boxes[(uint64_t)s[0] * 6 + 3 + a]; }
      uint32_t n_int = 0, n_tri = 0;
      for (int k = 0; k < ns; ++k) {
        const float* bx = &b.boxes[(uint64_t)s[k] * 6];
        for (int a = 0; a < 3; ++a) { lo[a] = std::fmin(lo[a], bx[a]); hi[a] = std::fmax(hi[a], bx[3 + a]); }
        if (cnt[s[k]] > 3) ++n_int; else n_tri += cnt[s[k]];
      }
      uint32_t e[3];
      for (int a = 0; a < 3; ++a) e[a] = quant_exponent((double)hi[a] - (double)lo[a]);
      const uint32_t child_base = n_int ? (uint32_t)queue.size() : 0u;
      const uint32_t tri_base = n_tri ? next_tri : 0u;
      next_tri += n_tri;
      const float c[3] = {0.5f * (lo[0] + hi[0]), 0.5f * (lo[1] + hi[1]), 0.5f * (lo[2] + hi[2])};
      uint32_t used = 0, imask = 0, slot_of[8];
      for (int k = 0; k < ns; ++k) {
        const float* bx = &b.boxes[(uint64_t)s[k] * 6];
        const float ox = 0.5f * (bx[0] + bx[3]) - c[0], oy = 0.5f * (bx[1] + bx[4]) - c[1],
                    oz = 0.5f * (bx[2] + bx[5]) - c[2];
        int pick = -1;
        float pick_cost = 0.0f;
        for (int t = 0; t < 8; ++t)
          if (!((used >> t) & 1u)) {
            const float cost = (((t & 4) ? ox : -ox) + ((t & 2) ? oy : -oy)) + ((t & 1) ? oz : -oz);
            if (pick < 0 || cost > pick_cost) { pick = t; pick_cost = cost; }
          }
        used |= 1u << pick;
        slot_of[k] = (uint32_t)pick;
        if (cnt[s[k]] > 3) imask |= 1u << pick;
      }
      // internal children enter the queue in slot order
      uint32_t by_slot[8];
      for (int t = 0; t < 8; ++t) by_slot[t] = 0xFFFFFFFFu;
      for (int k = 0; k < ns; ++k)
        if (cnt[s[k]] > 3) by_slot[slot_of[k]] = s[k];
      for (int t = 0; t < 8; ++t)
        if (by_slot[t] != 0xFFFFFFFFu) queue.push_back(by_slot[t]);
      if (!nodes_out) continue;
      uint8_t rec[80];
      std::memset(rec, 0, sizeof(rec));
      std::memcpy(rec, lo, 12);
      rec[12] = (uint8_t)e[0]; rec[13] = (uint8_t)e[1]; rec[14] = (uint8_t)e[2]; rec[15] = (uint8_t)imask;
      std::memcpy(rec + 16, &child_base, 4);
      std::memcpy(rec + 20, &tri_base, 4);
      uint32_t tri_off = 0;
      for (int k = 0; k < ns; ++k) {
        const float* bx = &b.boxes[(uint64_t)s[k] * 6];
        const uint32_t slot = slot_of[k];
        for (int a = 0; a < 3; ++a) {
          rec[32 + 8 * a + slot] = (uint8_t)quant_lo(bx[a], lo[a], e[a]);
          rec[56 + 8 * a + slot] = (uint8_t)quant_hi(bx[3 + a], lo[a], e[a]);
        }
        if (cnt[s[k]] > 3) {
          rec[24 + slot] = (uint8_t)(0x20u | (24u + slot));
        } else {
          rec[24 + slot] = (uint8_t)((((1u << cnt[s[k]]) - 1u) << 5) | tri_off);
          for (uint32_t t = 0; t < cnt[s[k]]; ++t) {
            const uint32_t prim = b.idx[first[s[k]] + t];
            const float* src = b.T + (uint64_t)prim * 12;
            float* dst = tris_out + (uint64_t)(tri_base + tri_off + t) * 12;
            for (int v = 0; v < 3; ++v) { dst[4 * v] = src[4 * v]; dst[4 * v + 1] = src[4 * v + 1]; dst[4 * v + 2] = src[4 * v + 2]; dst[4 * v + 3] = 0.0f; }
            std::memcpy(dst + 3, &prim, 4);
          }
          tri_off += cnt[s[k]];
        }
      }
      std::memcpy(nodes_out + (uint64_t)at * 80, rec, 80);
    }
    return queue.size();
  }
};
}  // namespace

ORC_API uint64_t orc_collapse8(const void* h, uint64_t cap, uint8_t* nodes, float* tris) {
  const Bvh& b = *(const Bvh*)h;
  if (b.n <= 3) return 0;
  Collapser8 c{b, 0, {}, {}};
  c.ranges();
  const uint64_t m = c.run(nullptr, nullptr);
  if (cap >= m && nodes && tris) c.run(nodes, tris);
  return m;
}

// The same wide tree in canonical (breadth-first) numbering: the device allocates child blocks
// and triangle runs with atomics, so its numbers depend on arrival order; its TREE does not.
// in_nodes: m x 80 bytes, in_tris: n x 12 floats -> out_nodes, out_tris.  Returns the nodes
// reached from node 0 (== m for a well-formed tree), or 0 on a malformed tree.
ORC_API uint64_t orc_canon8(const uint8_t* in_nodes, uint64_t m, const float* in_tris, uint64_t n,
                            uint8_t* out_nodes, float* out_tris) {
  std::vector<uint32_t> queue{0};
  uint32_t next_tri = 0;
  for (size_t at = 0; at < queue.size(); ++at) {
    if (queue[at] >= m || queue.size() > m) return 0;
    uint8_t rec[80];
    std::memcpy(rec, in_nodes + (uint64_t)queue[at] * 80, 80);
    uint32_t child_base, tri_base;
    std::memcpy(&child_base, rec + 16, 4);
    std::memcpy(&tri_base, rec + 20, 4);
    const uint32_t imask = rec[15];
    uint32_t n_int = 0, n_tri = 0;
    for (int t = 0; t < 8; ++t) {
      const uint32_t mt = rec[24 + t];
      if (!mt) continue;
      if ((mt & 0x18u) == 0x18u && (mt >> 5) == 1u) ++n_int;
      else n_tri += (uint32_t)__builtin_popcount(mt >> 5);
    }
    if (n_int != (uint32_t)__builtin_popcount(imask) || next_tri + n_tri > n || (n_tri && tri_base + n_tri > n)) return 0;
    const uint32_t new_child_base = n_int ? (uint32_t)queue.size() : 0u;
    const uint32_t new_tri_base = n_tri ? next_tri : 0u;
    for (uint32_t r = 0; r < n_int; ++r) queue.push_back(child_base + r);
    if (n_tri) std::memcpy(out_tris + (uint64_t)new_tri_base * 12, in_tris + (uint64_t)tri_base * 12, (size_t)n_tri * 48);
    next_tri += n_tri;
    std::memcpy(rec + 16, &new_child_base, 4);
    std::memcpy(rec + 20, &new_tri_base, 4);
    std::memcpy(out_nodes + (uint64_t)at * 80, rec, 80);
  }
  return next_tri == n ? queue.size() : 0;
}

// T0: all-pairs intersector, no BVH.  Same triangle test and acceptance rule.
ORC_API void orc_brute(const float* tris, uint64_t n, const float* rays, uint64_t nr, uint32_t flags,
                       float* hits) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t i = 0; i < (int64_t)nr; ++i) {
    Ray r;
    std::memcpy(&r, rays + (uint64_t)i * 8, 32);
    float best_t = r.tmax, bu = 0, bv = 0;
    int32_t best_prim = -1;
    const bool wt = (flags & TRACE_WATERTIGHT) != 0;
    WtRay wr = {0, 1, 2, 0, 0, 0};
    if (wt) wr = wt_setup(r);
    const float idir[3] = {inv_dir(r.d[0]), inv_dir(r.d[1]), inv_dir(r.d[2])};
    for (uint64_t p = 0; p < n; ++p) {
      float t, u, v;
      if ((wt ? tri_test_wt(tris + p * 12, r, wr, t, u, v) : tri_test(tris + p * 12, r, t, u, v)) &&
          accept(t, (int32_t)p, r.tmin, best_t, best_prim) && own_box_admits(tris + p * 12, r, idir, best_t, t)) {
        best_t = t; best_prim = (int32_t)p; bu = u; bv = v;
      }
    }
    Hit h = miss_hit();
    if (best_prim >= 0) {
      if (flags & TRACE_ANY) { h.t = 0; h.u = 0; h.v = 0; h.prim = kAnyHitPrim; }
      else { h.t = best_t; h.u = bu; h.v = bv; h.prim = best_prim; }
    }
    std::memcpy(hits + (uint64_t)i * 4, &h, 16);
  }
}
