// trace_common.cuh — ray state, slab test and Moeller-Trumbore intersection shared by
// the BVH2 traversal (traverse.cu) and the wide BVH4 traversal (wide.cu).  One copy
// of the arithmetic, so both layouts return bit-identical hits.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace rt {

constexpr int kStack = 96;  // >= 63 key bits + 32 tie-break bits of tree depth
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
  float ix, iy, iz;
  float best_t, bu, bv;
  int best_prim;
};

__device__ __forceinline__ bool slab(const float4 lo, const float4 hi, const RayState& r,
                                     float& tn) {
  const float t0x = (lo.x - r.ox) * r.ix, t1x = (hi.x - r.ox) * r.ix;
  const float t0y = (lo.y - r.oy) * r.iy, t1y = (hi.y - r.oy) * r.iy;
  const float t0z = (lo.z - r.oz) * r.iz, t1z = (hi.z - r.oz) * r.iz;
  tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), r.tmin));
  float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), r.best_t));
  tf = tf * kSlabPad;
  return tn <= tf;
}

// returns true when the hit was accepted
__device__ __forceinline__ bool intersect(const SortedTri* __restrict__ stris, uint32_t j,
                                          RayState& r) {
  const float4* tp = reinterpret_cast<const float4*>(stris + j);
  const float4 v0 = __ldg(tp), e1 = __ldg(tp + 1), e2 = __ldg(tp + 2);
  const float px = fmaf(r.dy, e2.z, -(r.dz * e2.y));
  const float py = fmaf(r.dz, e2.x, -(r.dx * e2.z));
  const float pz = fmaf(r.dx, e2.y, -(r.dy * e2.x));
  const float det = fmaf(e1.x, px, fmaf(e1.y, py, e1.z * pz));
  if (det == 0.0f) return false;
  const float inv = 1.0f / det;
  const float sx = r.ox - v0.x, sy = r.oy - v0.y, sz = r.oz - v0.z;
  const float u = fmaf(sx, px, fmaf(sy, py, sz * pz)) * inv;
  const float qx = fmaf(sy, e1.z, -(sz * e1.y));
  const float qy = fmaf(sz, e1.x, -(sx * e1.z));
  const float qz = fmaf(sx, e1.y, -(sy * e1.x));
  const float v = fmaf(r.dx, qx, fmaf(r.dy, qy, r.dz * qz)) * inv;
  const float t = fmaf(e2.x, qx, fmaf(e2.y, qy, e2.z * qz)) * inv;
  if (!((u >= 0.0f) && (v >= 0.0f) && ((u + v) <= 1.0f))) return false;
  const int prim = __float_as_int(v0.w);
  const bool ok = (t > r.tmin) &&
                  ((t < r.best_t) || (t == r.best_t && r.best_prim >= 0 && prim < r.best_prim));
  if (ok) {
    r.best_t = t;
    r.bu = u;
    r.bv = v;
    r.best_prim = prim;
  }
  return ok;
}


// fills the ray state from the two float4 of a ray record
__device__ __forceinline__ void load_ray(RayState& r, const float4 ro, const float4 rd) {
  r.ox = ro.x; r.oy = ro.y; r.oz = ro.z; r.tmin = ro.w;
  r.dx = rd.x; r.dy = rd.y; r.dz = rd.z;
  // 1/d clamped to +-FLT_MAX: 0 * FLT_MAX = 0 instead of 0 * inf = NaN when
  // the origin lies on a box face and the direction component is zero
  const float fmx = 3.402823466e+38f;
  r.ix = fminf(fmaxf(1.0f / rd.x, -fmx), fmx);
  r.iy = fminf(fmaxf(1.0f / rd.y, -fmx), fmx);
  r.iz = fminf(fmaxf(1.0f / rd.z, -fmx), fmx);
  r.best_t = rd.w; r.bu = 0.0f; r.bv = 0.0f; r.best_prim = -1;
}

__device__ __forceinline__ float4 hit_record(const RayState& r, bool any) {
  if (r.best_prim < 0) return make_float4(__int_as_float(0x7f800000), 0.0f, 0.0f, __int_as_float(-1));
  if (any) return make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0x7FFFFFFF));
  return make_float4(r.best_t, r.bu, r.bv, __int_as_float(r.best_prim));
}

}  // namespace rt
