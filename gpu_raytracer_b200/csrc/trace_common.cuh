// trace_common.cuh — ray state, slab test and Moeller-Trumbore intersection shared by
// the BVH2 traversal (traverse.cu) and the wide BVH4 traversal (wide.cu).  One copy
// of the arithmetic, so both layouts return bit-identical hits.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace rt {

constexpr int kStack = kMaxDepth;  // >= 63 key bits + 32 tie-break bits of tree depth
constexpr int kEmpty = (int)0x80000000;
constexpr float kSlabPad = 1.0000004f;
constexpr int kTraceThreads = 256;

// 32 bytes in one request (LDG.E.256, sm_100): half the L1 tag lookups of two
// LDG.128 when every lane reads a different node, which is what bounds traversal
// (l1tex throughput 56-96 % in profiles/r01e_ncu_trace.md).  p must be 32-byte aligned.
struct F8 {
  float4 a, b;
};
__device__ __forceinline__ F8 ldg256(const void* p) {
  F8 r;
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z),
        "=f"(r.b.w)
      : "l"(p));
  return r;
}

struct RayState {
  float ox, oy, oz, tmin;
  float dx, dy, dz;
  float ix, iy, iz;  // 1/d clamped to +-FLT_MAX; +inf for a component that is exactly +-0
  float best_t, bu, bv;
  int best_prim;
};

// min / max that PROPAGATE a NaN operand (min.NaN.f32 / max.NaN.f32: FMNMX.NAN), where
// fminf / fmaxf return the other operand
__device__ __forceinline__ float min_nan(float a, float b) {
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// Slab test.  The two plane distances of an axis are combined with NaN-PROPAGATING min/max,
// the three axes (and tmin / best_t) with NaN-IGNORING ones.  A direction component that is
// exactly +-0 has ix = +inf, so its plane distances are -inf, +inf or, for an origin exactly ON
// the plane, 0 * inf = NaN:
//   origin strictly between the planes  (-inf, +inf): the axis allows every t
//   origin on a plane                   a NaN: the axis drops out of tnear and tfar, i.e. it
//                                       allows every t as well (the slab is closed)
//   origin outside                      (-inf, -inf) or (+inf, +inf): tfar = -inf or tnear = +inf
// which is the exact answer for a ray parallel to the slab, at the instruction count of the
// plain test and with no per-ray flag.  (The first version clamped 1/0 to FLT_MAX: the
// on-plane distance became 0 * FLT_MAX = 0 and the node was culled.)  Non-zero components
// never produce a NaN: ix is clamped to +-FLT_MAX and the inputs are finite.
__device__ __forceinline__ bool slab(float lx, float ly, float lz, float hx, float hy, float hz,
                                     const RayState& r, float& tn) {
  const float t0x = (lx - r.ox) * r.ix, t1x = (hx - r.ox) * r.ix;
  const float t0y = (ly - r.oy) * r.iy, t1y = (hy - r.oy) * r.iy;
  const float t0z = (lz - r.oz) * r.iz, t1z = (hz - r.oz) * r.iz;
  tn = fmaxf(fmaxf(min_nan(t0x, t1x), min_nan(t0y, t1y)), fmaxf(min_nan(t0z, t1z), r.tmin));
  float tf = fminf(fminf(max_nan(t0x, t1x), max_nan(t0y, t1y)), fminf(max_nan(t0z, t1z), r.best_t));
  tf = tf * kSlabPad;
  return tn <= tf;
}
__device__ __forceinline__ bool slab(const float4 lo, const float4 hi, const RayState& r, float& tn) {
  return slab(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, r, tn);
}

// Acceptance rule (A6): closest t in (tmin, best]; on equal t the smaller primitive wins; and the
// hit must be CONSISTENT WITH THE TRIANGLE'S OWN BOX: slab() passes on the box of its three
// vertices (ordered min/max, the tree's leaf box) and the box's entry distance satisfies
// tnear <= t * kSlabPad.  The last condition is what makes the result independent of the
// traversal: a node is culled when tnear(node) > min(tfar, best_t) * pad, every ancestor's box
// contains the leaf box and slab() is monotone in the box, so a triangle whose hit is
// acceptable (t <= best_t) can never sit below a culled node.  The accepted hit is therefore
// min over {triangles with a consistent hit} by (t, prim): a function of the ray and the
// triangles alone, which the binary tree, both wide layouts, any visiting order and the
// all-pairs intersector (oracle brute) all return bit for bit.  (Without it the answer
// depended on the order in marginal cases: a triangle whose computed t lies a few ulps before
// its own flat box could be culled, or not, depending on what had been found before.)
// The box is only evaluated for candidates that would otherwise be accepted (about one in three
// triangle tests), in a function of its own that reads the vertices again (L1 hits): inlined,
// the twelve vertex floats stayed live across the triangle test and cost 7-9 registers, i.e.
// one of five resident CTAs per SM.
static __device__ __noinline__ bool own_box_admits(const float4* __restrict__ tp, float ox, float oy, float oz,
                                            float ix, float iy, float iz, float tmin, float best_t, float t) {
  const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
  RayState r;
  r.ox = ox; r.oy = oy; r.oz = oz; r.ix = ix; r.iy = iy; r.iz = iz; r.tmin = tmin; r.best_t = best_t;
  float tn;
  return slab(omin(omin(v0.x, v1.x), v2.x), omin(omin(v0.y, v1.y), v2.y), omin(omin(v0.z, v1.z), v2.z),
              omax(omax(v0.x, v1.x), v2.x), omax(omax(v0.y, v1.y), v2.y), omax(omax(v0.z, v1.z), v2.z), r, tn) &&
         tn <= t * kSlabPad;
}
// SPILLED: the caller does not keep r.ix / iy / iz in registers (compact layout); `inv` points at
// this thread's copy of them in shared memory (inv[0], inv[kTraceThreads], inv[2 kTraceThreads]),
// read by the few candidates that get this far.  (Computing 1/d again instead cost three
// 20-instruction IEEE reciprocals per candidate: 5 % of the kernel's instructions.)
template <bool SPILLED = false>
__device__ __forceinline__ bool accept_hit(float t, float u, float v, int prim, const float4* __restrict__ tp,
                                           RayState& r, const float* inv = nullptr) {
  bool ok = (t > r.tmin) &&
            ((t < r.best_t) || (t == r.best_t && r.best_prim >= 0 && prim < r.best_prim));
  if (ok) {
    if (SPILLED) ok = own_box_admits(tp, r.ox, r.oy, r.oz, inv[0], inv[kTraceThreads], inv[2 * kTraceThreads], r.tmin, r.best_t, t);
    else ok = own_box_admits(tp, r.ox, r.oy, r.oz, r.ix, r.iy, r.iz, r.tmin, r.best_t, t);
  }
  if (ok) {
    r.best_t = t;
    r.bu = u;
    r.bv = v;
    r.best_prim = prim;
  }
  return ok;
}

// Moeller-Trumbore, two-sided.  The triangle record holds the raw vertices
// {v0 | prim, v1, v2}; the edges are one subtraction each, as in the oracle.
// Returns true when the hit was accepted.
template <bool SPILLED = false>
__device__ __forceinline__ bool intersect(const SortedTri* __restrict__ stris, uint32_t j,
                                          RayState& r, const float* sh_inv = nullptr) {
  const float4* tp = reinterpret_cast<const float4*>(stris + j);
  const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
  const float e1x = v1.x - v0.x, e1y = v1.y - v0.y, e1z = v1.z - v0.z;
  const float e2x = v2.x - v0.x, e2y = v2.y - v0.y, e2z = v2.z - v0.z;
  const float px = fmaf(r.dy, e2z, -(r.dz * e2y));
  const float py = fmaf(r.dz, e2x, -(r.dx * e2z));
  const float pz = fmaf(r.dx, e2y, -(r.dy * e2x));
  const float det = fmaf(e1x, px, fmaf(e1y, py, e1z * pz));
  if (det == 0.0f) return false;
  const float inv = 1.0f / det;
  const float sx = r.ox - v0.x, sy = r.oy - v0.y, sz = r.oz - v0.z;
  const float u = fmaf(sx, px, fmaf(sy, py, sz * pz)) * inv;
  const float qx = fmaf(sy, e1z, -(sz * e1y));
  const float qy = fmaf(sz, e1x, -(sx * e1z));
  const float qz = fmaf(sx, e1y, -(sy * e1x));
  const float v = fmaf(r.dx, qx, fmaf(r.dy, qy, r.dz * qz)) * inv;
  const float t = fmaf(e2x, qx, fmaf(e2y, qy, e2z * qz)) * inv;
  if (!((u >= 0.0f) && (v >= 0.0f) && ((u + v) <= 1.0f))) return false;
  return accept_hit<SPILLED>(t, u, v, __float_as_int(v0.w), tp, r, sh_inv);
}

// Watertight test (Woop, Benthin & Wald 2013), RT_TRACE_WATERTIGHT: the ray is
// sheared so its dominant axis is +z (kx, ky, kz, S* are per-ray constants); the
// scaled edge functions U, V, W decide the hit and are recomputed in double when one
// of them is exactly zero, so a ray through a shared edge or vertex hits one of the
// triangles around it.  Same operation order as the oracle's tri_test_wt.
struct WtRay {
  int kx, ky, kz;
  float Sx, Sy, Sz;
};
__device__ __forceinline__ float sel3(float x, float y, float z, int k) {
  return k == 0 ? x : (k == 1 ? y : z);
}
__device__ __forceinline__ WtRay wt_setup(const RayState& r) {
  WtRay w;
  w.kz = 0;
  float m = fabsf(r.dx);
  if (fabsf(r.dy) > m) { w.kz = 1; m = fabsf(r.dy); }
  if (fabsf(r.dz) > m) w.kz = 2;
  w.kx = (w.kz + 1) % 3;
  w.ky = (w.kx + 1) % 3;
  const float dz = sel3(r.dx, r.dy, r.dz, w.kz);
  if (dz < 0.0f) {
    const int t = w.kx;
    w.kx = w.ky;
    w.ky = t;
  }
  w.Sx = sel3(r.dx, r.dy, r.dz, w.kx) / dz;
  w.Sy = sel3(r.dx, r.dy, r.dz, w.ky) / dz;
  w.Sz = 1.0f / dz;
  return w;
}
__device__ __forceinline__ bool intersect_wt(const SortedTri* __restrict__ stris, uint32_t j,
                                             RayState& r, const WtRay& w) {
  const float4* tp = reinterpret_cast<const float4*>(stris + j);
  const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
  const float ax = v0.x - r.ox, ay = v0.y - r.oy, az = v0.z - r.oz;
  const float bx = v1.x - r.ox, by = v1.y - r.oy, bz = v1.z - r.oz;
  const float cx = v2.x - r.ox, cy = v2.y - r.oy, cz = v2.z - r.oz;
  const float Akz = sel3(ax, ay, az, w.kz), Bkz = sel3(bx, by, bz, w.kz), Ckz = sel3(cx, cy, cz, w.kz);
  const float Ax = sel3(ax, ay, az, w.kx) - w.Sx * Akz, Ay = sel3(ax, ay, az, w.ky) - w.Sy * Akz;
  const float Bx = sel3(bx, by, bz, w.kx) - w.Sx * Bkz, By = sel3(bx, by, bz, w.ky) - w.Sy * Bkz;
  const float Cx = sel3(cx, cy, cz, w.kx) - w.Sx * Ckz, Cy = sel3(cx, cy, cz, w.ky) - w.Sy * Ckz;
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
  const float Az = w.Sz * Akz, Bz = w.Sz * Bkz, Cz = w.Sz * Ckz;
  const float T = (U * Az + V * Bz) + W * Cz;
  const float inv = 1.0f / det;
  return accept_hit(T * inv, V * inv, W * inv, __float_as_int(v0.w), tp, r);
}

// 1/d of one direction component, clamped to +-FLT_MAX so that a denormal component gives
// 0 * FLT_MAX = 0, not 0 * inf = NaN, for an origin on a box face; +inf for a component that is
// exactly +-0, which slab() treats as a ray parallel to the slab
__device__ __forceinline__ float inv_dir(float d) {
  const float fmx = 3.402823466e+38f;
  return d == 0.0f ? __int_as_float(0x7F800000) : fminf(fmaxf(1.0f / d, -fmx), fmx);
}
// fills the ray state from the two float4 of a ray record
__device__ __forceinline__ void load_ray(RayState& r, const float4 ro, const float4 rd) {
  r.ox = ro.x; r.oy = ro.y; r.oz = ro.z; r.tmin = ro.w;
  r.dx = rd.x; r.dy = rd.y; r.dz = rd.z;
  r.ix = inv_dir(rd.x);
  r.iy = inv_dir(rd.y);
  r.iz = inv_dir(rd.z);
  r.best_t = rd.w; r.bu = 0.0f; r.bv = 0.0f; r.best_prim = -1;
}

// ---- compact layout (NodeQ): per-ray constants and the slab test on 16-bit planes ----------
// A plane q of axis a lies at origin + q * step; its distance along the ray is
//   T(q) = (origin - o) * c + q * (step * c),      c = r.ix (2^64 for a component that is exactly 0)
// computed as fma(m, A, B) with m = 2^23 + q (PRMT puts the 16 bits under the exponent of 2^23),
// A = step * c and B = (origin - o) * c - 2^23 * A.  The NEAR plane of an axis is its low plane
// when c >= 0, else its high plane: the ray holds the PRMT selector of each (sel, sel ^ 0x22), so
// no min / max per axis is needed, and the near planes use B - d, the far planes B + d,
// d = 2^-20 |(origin - o) * c|, which covers the roundings that grow with the distance of the
// origin; those that grow with the step (at most 0.53 |A|) are covered by the whole step that
// packing adds to every box (compact.cu).  NaNs (overflowing constants) drop out of the
// NaN-ignoring min / max: the test never culls what slab() passes on a box inside this one.
struct QRay {
  float ax, ay, az;  // A
  float nx, ny, nz;  // B - d
  float fx, fy, fz;  // B + d
  uint32_t sx, sy, sz;  // PRMT selector of the near plane: 0x7610 (low half) or 0x7632
};
__device__ __forceinline__ void q_axis(float origin, float step, float o, float d, float inv, float& A, float& Bn,
                                       float& Bf, uint32_t& sel) {
  const float c = d == 0.0f ? 18446744073709551616.0f : inv;  // inv = inv_dir(d), from load_ray
  A = step * c;
  const float B = (origin - o) * c;
  const float e = fabsf(B) * 9.5367431640625e-07f;  // 2^-20
  Bn = fmaf(-8388608.0f, A, B - e);
  Bf = fmaf(-8388608.0f, A, B + e);
  sel = c >= 0.0f ? 0x7610u : 0x7632u;
}
__device__ __forceinline__ QRay q_setup(const RayState& r, const QGrid& g) {
  QRay q;
  q_axis(g.ox, g.sx, r.ox, r.dx, r.ix, q.ax, q.nx, q.fx, q.sx);
  q_axis(g.oy, g.sy, r.oy, r.dy, r.iy, q.ay, q.ny, q.fy, q.sy);
  q_axis(g.oz, g.sz, r.oz, r.dz, r.iz, q.az, q.nz, q.fz, q.sz);
  return q;
}
__device__ __forceinline__ float q_plane(uint32_t w, uint32_t sel) {
  uint32_t r;  // prmt directly: __byte_perm masks a run-time selector with 0x7777 first (one LOP3 each)
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(0x4B000000u), "r"(sel));
  return __uint_as_float(r);
}
__device__ __forceinline__ bool slab_q(uint32_t wx, uint32_t wy, uint32_t wz, const QRay& q, const RayState& r,
                                       float& tn) {
  const float nx = fmaf(q_plane(wx, q.sx), q.ax, q.nx), fx = fmaf(q_plane(wx, q.sx ^ 0x22u), q.ax, q.fx);
  const float ny = fmaf(q_plane(wy, q.sy), q.ay, q.ny), fy = fmaf(q_plane(wy, q.sy ^ 0x22u), q.ay, q.fy);
  const float nz = fmaf(q_plane(wz, q.sz), q.az, q.nz), fz = fmaf(q_plane(wz, q.sz ^ 0x22u), q.az, q.fz);
  tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, r.tmin));
  float tf = fminf(fminf(fx, fy), fminf(fz, r.best_t));
  tf = tf * kSlabPad;
  return tn <= tf;
}
__device__ __forceinline__ F8 ldg256u(const void* p) { return ldg256(p); }

__device__ __forceinline__ float4 hit_record(const RayState& r, bool any) {
  if (r.best_prim < 0) return make_float4(__int_as_float(0x7f800000), 0.0f, 0.0f, __int_as_float(-1));
  if (any) return make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0x7FFFFFFF));
  return make_float4(r.best_t, r.bu, r.bv, __int_as_float(r.best_prim));
}

// expands M(any, stats, watertight) for the run-time values of the three flags
#define RT_DISPATCH3(M, a, b, c)              \
  do {                                        \
    if (a) {                                  \
      if (b) { if (c) M(true, true, true); else M(true, true, false); }       \
      else   { if (c) M(true, false, true); else M(true, false, false); }     \
    } else {                                  \
      if (b) { if (c) M(false, true, true); else M(false, true, false); }     \
      else   { if (c) M(false, false, true); else M(false, false, false); }   \
    }                                         \
  } while (0)

}  // namespace rt
