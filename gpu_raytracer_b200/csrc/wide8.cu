// wide8.cu — A5w, second wide layout: a compressed 8-wide BVH (80-byte nodes with 8-bit
// quantised child boxes, Ylitie, Karras & Laine 2017) collapsed from the binary LBVH, and its
// traversal (BASELINE.json:5 "optional collapse to a wide BVH4/8").
//
// Why: a BVH2 step moves 64 bytes for two boxes; this layout moves 80 bytes for eight, and the
// whole node array of a 10M-triangle scene is ~1/4 of the binary one.  For incoherent rays
// (config 5), whose BVH2 nodes come from DRAM, that is the lever; coherent frames are bound by
// instruction issue and gain less (profiles/r02*_wide8.md has the measurements).
//
// Collapse: top-down from the binary tree, one launch per level (see "collapse" below): a wide
// node takes the children of its binary node and keeps opening the one with the largest
// surface area until it has eight; subtrees of at most three triangles become leaves.  (A
// first version grouped the binary nodes by bit triples of the Morton codes, the octree the
// codes describe: purely local, no queue, slot = digit, but only 3.1-3.5 children per node
// on the test scenes, where this one gets 6-7.)
//
// Node (80 B, five 16-byte words):
//   w0  px py pz | ex ey ez imask      grid origin = min corner of the node's box; 2^(e-127) =
//                                      cell size per axis; imask bit s = slot s is an internal child
//   w1  child_base tri_base meta[8]    first internal child; first leaf triangle; per slot:
//                                      0 empty | 0x20 | 24 + s internal | unary(count) << 5 | offset leaf
//   w2..w4  qlo.x[8] qlo.y[8] qlo.z[8] qhi.x[8] qhi.y[8] qhi.z[8]
// Child box of slot s: [p + qlo*2^e, p + qhi*2^e], quantised outward in double precision, so it
// CONTAINS the exact box.
//
// Bit-identical hits.  In the binary tree a triangle is tested exactly when its own leaf box
// passes slab() (every ancestor's box contains it, and slab() is monotone in the box), so the
// closest accepted hit is a function of the per-triangle boxes alone.  Here the inner nodes
// cull with a CONSERVATIVE test (quantised boxes, one FMA per plane, an explicit rounding
// allowance: never rejects what slab() on the exact box accepts), and the acceptance rule
// (trace_common.cuh accept_hit) makes the accepted hit a function of the ray and the
// triangles alone: min over the triangles whose hit is consistent with their own box by (t,
// prim).  Conservative culling cannot lose such a triangle, so this tree returns the oracle's
// bytes whatever its shape and visiting order (tests/test_wide8.py).
#include "trace_common.cuh"

namespace rt {
namespace {

constexpr uint32_t kLeafMax8 = 3;

// ---- collapse -----------------------------------------------------------------------------------
// Top-down, one launch per level of the wide tree.  A work item is (BVH2 node, wide node
// index).  The item's thread gathers up to eight children: it starts from the BVH2 node's two
// children and keeps replacing the child with the largest half surface area among those that
// hold more than kLeafMax8 triangles by that child's own two children (left-to-right order
// is kept), until there are eight or nothing is left to open; then, while slots are free, it
// opens the largest child holding 2 or 3 triangles (tighter leaf boxes, no more nodes).  Children with at most
// kLeafMax8 triangles become leaves, the others internal children: their wide nodes are
// allocated as one consecutive block (one atomicAdd), the node's leaf triangles as one
// consecutive run of the second triangle array (another), and the internal children are
// queued for the next level.  Slots: every child prefers the octant its box centre lies in,
// seen from the centre of the node's box (slot bit 2 / 1 / 0 = upper half in x / y / z);
// children are placed in left-to-right order, each into the free slot that agrees best with
// its offset (sum over the axes of +-offset), ties to the lowest slot.  Node numbers depend
// on the order the atomics arrive in; the TREE (children, slots, boxes, triangles) does not,
// and is what the oracle's collapse8 is compared with (tests/test_wide8.py).
struct Item8 {
  uint32_t g, index;
};
struct Slot8 {
  float lx, ly, lz, hx, hy, hz;
  int enc;              // BVH2 child id: >= 0 split position, < 0 leaf (~sorted position)
  uint32_t first, cnt;  // leaf range of the subtree
};
__device__ __forceinline__ float half_area8(const Slot8& s) {
  const float dx = s.hx - s.lx, dy = s.hy - s.ly, dz = s.hz - s.lz;
  return (dx * dy + dy * dz) + dz * dx;  // -fmad=false: three products, two sums
}
__device__ __forceinline__ void load_children8(const Node* __restrict__ nodes, uint32_t x, Slot8& a, Slot8& b) {
  const float4* np = reinterpret_cast<const float4*>(nodes + x);
  const float4 lo0 = __ldg(np), hi0 = __ldg(np + 1), lo1 = __ldg(np + 2), hi1 = __ldg(np + 3);
  const uint32_t L = __float_as_uint(hi0.w), R = __float_as_uint(hi1.w);
  a.lx = lo0.x; a.ly = lo0.y; a.lz = lo0.z; a.hx = hi0.x; a.hy = hi0.y; a.hz = hi0.z;
  a.enc = __float_as_int(lo0.w); a.first = L; a.cnt = x - L + 1u;
  b.lx = lo1.x; b.ly = lo1.y; b.lz = lo1.z; b.hx = hi1.x; b.hy = hi1.y; b.hz = hi1.z;
  b.enc = __float_as_int(lo1.w); b.first = x + 1u; b.cnt = R - x;
}

__device__ __forceinline__ uint32_t quant_exponent(double ext) {
  // smallest e with 255 * 2^e >= ext, as the float exponent byte e + 127 (0: the axis is flat)
  if (!(ext > 0.0)) return 0u;
  int e;
  frexp(ext / 255.0, &e);  // ext / 255 = m * 2^e with m in [0.5, 1): 2^e >= ext / 255
  while (ldexp(255.0, e - 1) >= ext) --e;
  while (ldexp(255.0, e) < ext) ++e;
  if (e < -126) e = -126;
  if (e > 127) e = 127;
  return (uint32_t)(e + 127);
}
__device__ __forceinline__ uint32_t quant_lo(float v, float p, uint32_t eb) {
  if (eb == 0u) return 0u;
  const double cell = ldexp(1.0, (int)eb - 127), d = (double)v - (double)p;
  double q = floor(d / cell);
  q = fmin(fmax(q, 0.0), 255.0);
  while (q > 0.0 && (double)p + q * cell > (double)v) q -= 1.0;
  return (uint32_t)q;
}
__device__ __forceinline__ uint32_t quant_hi(float v, float p, uint32_t eb) {
  if (eb == 0u) return 0u;
  const double cell = ldexp(1.0, (int)eb - 127), d = (double)v - (double)p;
  double q = ceil(d / cell);
  q = fmin(fmax(q, 0.0), 255.0);
  while (q < 255.0 && (double)p + q * cell < (double)v) q += 1.0;
  return (uint32_t)q;
}

// counters: [0] wide nodes allocated, [1] leaf triangles placed, [2] overflow flag,
//           [4 + l] items of level l
__global__ void __launch_bounds__(128)
wide8_level_kernel(const Node* __restrict__ nodes, const SortedTri* __restrict__ stris,
                   const Item8* __restrict__ in, Item8* __restrict__ next, uint32_t* __restrict__ counters,
                   int level, uint32_t capacity, Node8* __restrict__ out, SortedTri* __restrict__ tris8) {
  const uint32_t n_items = counters[4 + level];
  for (uint32_t it = blockIdx.x * blockDim.x + threadIdx.x; it < n_items; it += gridDim.x * blockDim.x) {
    const Item8 item = in[it];
    Slot8 s[8];
    load_children8(nodes, item.g, s[0], s[1]);
    int ns = 2;
    while (ns < 8) {
      int best = -1;
      float best_area = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (k < ns && s[k].cnt > kLeafMax8) {
          const float a = half_area8(s[k]);
          if (best < 0 || a > best_area) { best = k; best_area = a; }
        }
      }
      if (best < 0) {  // nothing left that must be opened: spend the free slots on tighter leaves
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < ns && s[k].cnt > 1u) {
            const float a = half_area8(s[k]);
            if (best < 0 || a > best_area) { best = k; best_area = a; }
          }
        }
      }
      if (best < 0) break;
      int x = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k == best) x = s[k].enc;
      Slot8 a, b;
      load_children8(nodes, (uint32_t)x, a, b);
#pragma unroll
      for (int k = 7; k >= 1; --k)
        if (k > best + 1 && k <= ns) s[k] = s[k - 1];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (k == best) s[k] = a;
        else if (k == best + 1) s[k] = b;
      }
      ++ns;
    }
    // the node's box, its grid, the allocation of its children
    float bx0 = s[0].lx, by0 = s[0].ly, bz0 = s[0].lz, bx1 = s[0].hx, by1 = s[0].hy, bz1 = s[0].hz;
    uint32_t n_int = 0, n_tri = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < ns) {
        bx0 = fminf(bx0, s[k].lx); by0 = fminf(by0, s[k].ly); bz0 = fminf(bz0, s[k].lz);
        bx1 = fmaxf(bx1, s[k].hx); by1 = fmaxf(by1, s[k].hy); bz1 = fmaxf(bz1, s[k].hz);
        if (s[k].cnt > kLeafMax8) ++n_int; else n_tri += s[k].cnt;
      }
    }
    const uint32_t ex = quant_exponent((double)bx1 - (double)bx0), ey = quant_exponent((double)by1 - (double)by0),
                   ez = quant_exponent((double)bz1 - (double)bz0);
    uint32_t child_base = 0u, tri_base = 0u, queue_at = 0u;
    if (n_int) {
      child_base = atomicAdd(&counters[0], n_int);
      queue_at = atomicAdd(&counters[4 + level + 1], n_int);
      if (child_base + n_int > capacity) {
        counters[2] = 1u;  // more wide nodes than the array holds: the caller fails the collapse
        continue;
      }
    }
    if (n_tri) tri_base = atomicAdd(&counters[1], n_tri);
    // slots: left to right, each child into the free slot that agrees best with its offset
    const float cx = 0.5f * (bx0 + bx1), cy = 0.5f * (by0 + by1), cz = 0.5f * (bz0 + bz1);
    uint32_t used = 0u, imask = 0u;
    uint32_t q8[12] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t meta[2] = {0u, 0u};
    uint32_t slot_of[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      slot_of[k] = 0u;
      if (k < ns) {
        const float ox = 0.5f * (s[k].lx + s[k].hx) - cx, oy = 0.5f * (s[k].ly + s[k].hy) - cy,
                    oz = 0.5f * (s[k].lz + s[k].hz) - cz;
        int pick = -1;
        float pick_cost = 0.0f;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          if (!((used >> t) & 1u)) {
            const float cost = (((t & 4) ? ox : -ox) + ((t & 2) ? oy : -oy)) + ((t & 1) ? oz : -oz);
            if (pick < 0 || cost > pick_cost) { pick = t; pick_cost = cost; }
          }
        }
        used |= 1u << pick;
        slot_of[k] = (uint32_t)pick;
        if (s[k].cnt > kLeafMax8) imask |= 1u << pick;
      }
    }
    // internal children are stored in slot order: rank of the slot among the internal slots
    uint32_t tri_off = 0u;
    auto put = [&](int plane, uint32_t slot, uint32_t v) { q8[2 * plane + (slot >> 2)] |= v << (8u * (slot & 3u)); };
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < ns) {
        const uint32_t slot = slot_of[k];
        put(0, slot, quant_lo(s[k].lx, bx0, ex)); put(1, slot, quant_lo(s[k].ly, by0, ey)); put(2, slot, quant_lo(s[k].lz, bz0, ez));
        put(3, slot, quant_hi(s[k].hx, bx0, ex)); put(4, slot, quant_hi(s[k].hy, by0, ey)); put(5, slot, quant_hi(s[k].hz, bz0, ez));
        uint32_t m;
        if (s[k].cnt > kLeafMax8) {
          m = 0x20u | (24u + slot);
          const uint32_t rank = (uint32_t)__popc(imask & ((1u << slot) - 1u));
          next[queue_at + rank] = Item8{(uint32_t)s[k].enc, child_base + rank};
        } else {
          m = (((1u << s[k].cnt) - 1u) << 5) | tri_off;
          for (uint32_t t = 0; t < s[k].cnt; ++t) {
            const float4* src = reinterpret_cast<const float4*>(stris + s[k].first + t);
            float4* dst = reinterpret_cast<float4*>(tris8 + tri_base + tri_off + t);
            dst[0] = __ldg(src);
            dst[1] = __ldg(src + 1);
            dst[2] = __ldg(src + 2);
          }
          tri_off += s[k].cnt;
        }
        meta[slot >> 2] |= m << (8u * (slot & 3u));
      }
    }
    uint4* o = reinterpret_cast<uint4*>(out + item.index);
    o[0] = make_uint4(__float_as_uint(bx0), __float_as_uint(by0), __float_as_uint(bz0),
                      ex | (ey << 8) | (ez << 16) | (imask << 24));
    o[1] = make_uint4(child_base, tri_base, meta[0], meta[1]);
    o[2] = make_uint4(q8[0], q8[1], q8[2], q8[3]);
    o[3] = make_uint4(q8[4], q8[5], q8[6], q8[7]);
    o[4] = make_uint4(q8[8], q8[9], q8[10], q8[11]);
  }
}

// ---- traversal ---------------------------------------------------------------------------------------
constexpr int kStack8 = 34;  // one pending group per level: (63 key bits + 32 index bits) / 3 levels

__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x) {  // 0x80 bytes -> 0xFF, others -> 0x00
  uint32_t r;
  asm("prmt.b32 %0, %1, 0, 0xba98;" : "=r"(r) : "r"(x));
  return r;
}
// byte j of w as a float: one PRMT puts the byte into the mantissa of 2^23, one FADD takes 2^23
// away (exact for 0..255); I2F is a quarter-rate instruction and a node step converts 48 bytes
__device__ __forceinline__ float byte_f(uint32_t w, int j) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650u + (uint32_t)j)) - 8388608.0f;
}

template <bool ANY, bool STATS, bool WT>
__global__ void __launch_bounds__(kTraceThreads)
trace_wide8_kernel(const Node8* __restrict__ nodes8, const SortedTri* __restrict__ tris8,
                   const float4* __restrict__ rays, unsigned long long n_rays, float4* __restrict__ hits,
                   unsigned long long* __restrict__ ray_counter, unsigned long long* __restrict__ stats,
                   int chunk, int refill_min, int leaf_min, const uint32_t* __restrict__ perm) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t n_nodes = 0, n_tests = 0;

  uint2 stack[kStack8];
  int sp = 0;
  uint2 ng = make_uint2(0u, 0u);  // node group: x = first internal child, y = hit bits 24..31 | imask
  uint2 tg = make_uint2(0u, 0u);  // triangle group: x = first triangle, y = hit bits 0..23
  bool has_ray = false;
  unsigned long long my_ray = 0;
  RayState r;
  r.best_prim = -1;
  float cix = 0.0f, ciy = 0.0f, ciz = 0.0f;  // 1/d for the conservative test: 2^100 where d is exactly zero
  uint32_t octinv4 = 0u;
  WtRay wr = {0, 1, 2, 0.0f, 0.0f, 0.0f};
  unsigned long long chunk_next = 0, chunk_end = 0;  // warp-uniform
  bool exhausted = false;                            // warp-uniform

  while (true) {
    // ---- per-lane refill from the warp's chunk --------------------------------
    const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !has_ray);
    if (idle != 0u && !exhausted && (__popc(idle) >= refill_min || idle == 0xFFFFFFFFu)) {
      if (chunk_next >= chunk_end) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(ray_counter, (unsigned long long)chunk);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n_rays) {
          exhausted = true;
          chunk_next = chunk_end = 0;
        } else {
          chunk_next = base;
          chunk_end = min(base + (unsigned long long)chunk, n_rays);
        }
      }
      const unsigned long long avail = chunk_end - chunk_next;
      const uint32_t slot = __popc(idle & lt_mask);
      if (!has_ray && slot < avail) {
        my_ray = chunk_next + slot;
        if (perm) my_ray = perm[my_ray];
        const float4 rd = ld_stream(rays + 2 * my_ray + 1);
        load_ray(r, ld_stream(rays + 2 * my_ray), rd);
        // a component that is exactly zero: slab() works with +-inf / NaN plane distances there
        // (the ray is inside the closed slab for every t, or for none).  One FMA per plane
        // cannot: q * inf + b is NaN for every plane, the axis would never cull, and a ray in a
        // coordinate plane would visit a 2-D slice of the whole tree (one row of the 4K frame
        // has d.y == 0: 3,840 rays took 13 of the frame's 15 ms).  2^100 instead of inf keeps
        // every product finite and the sign of (plane - origin): strictly outside -> both
        // distances huge with one sign -> culled; inside or within the allowance -> passes.
        const float big = 1.2676506e30f;  // 2^100
        cix = rd.x == 0.0f ? big : r.ix;
        ciy = rd.y == 0.0f ? big : r.iy;
        ciz = rd.z == 0.0f ? big : r.iz;
        octinv4 = ((rd.x < 0.0f ? 0u : 4u) | (rd.y < 0.0f ? 0u : 2u) | (rd.z < 0.0f ? 0u : 1u)) * 0x01010101u;
        if (WT) wr = wt_setup(r);
        has_ray = true;
        sp = 0;
        ng = make_uint2(0u, 0x80000000u);  // a virtual group whose one child is node 0, the root
        tg = make_uint2(0u, 0u);
      }
      const unsigned long long need = (unsigned long long)__popc(idle);
      chunk_next += (need < avail) ? need : avail;
    }
    if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) {
      if (exhausted) break;
      continue;
    }

    // ---- node steps, until enough lanes hold triangles to test ------------------------------
    while (true) {
      const bool at_node = has_ray && tg.y == 0u && ng.y > 0x00FFFFFFu;
      if (__ballot_sync(0xFFFFFFFFu, at_node) == 0u) break;
      const uint32_t waiting = __ballot_sync(0xFFFFFFFFu, has_ray && !at_node);  // triangles pending, or finished
      if (__popc(waiting) >= leaf_min) break;
      if (at_node) {
        if (STATS) ++n_nodes;
        // next child of the group: the highest hit bit = the nearest octant first
        const uint32_t bit = 31u - (uint32_t)__clz((int)ng.y);
        const uint32_t imask_parent = ng.y & 0xFFu;
        ng.y &= ~(1u << bit);
        // (the root enters as the only child of a virtual group with block start 0 and no imask)
        const uint32_t slot = (bit - 24u) ^ (octinv4 & 7u);
        const uint32_t child = ng.x + __popc(imask_parent & ((1u << slot) - 1u));
        if (ng.y > 0x00FFFFFFu) stack[sp++] = ng;
        const uint4* np = reinterpret_cast<const uint4*>(nodes8 + child);
        const uint4 w0 = __ldg(np), w1 = __ldg(np + 1), w2 = __ldg(np + 2), w3 = __ldg(np + 3), w4 = __ldg(np + 4);
        // ray in the node's grid: t(q) = q * s + b per axis, one FMA per plane
        const float sx = __uint_as_float((w0.w & 0xFFu) << 23) * cix;
        const float sy = __uint_as_float(((w0.w >> 8) & 0xFFu) << 23) * ciy;
        const float sz = __uint_as_float(((w0.w >> 16) & 0xFFu) << 23) * ciz;
        const float bx = (__uint_as_float(w0.x) - r.ox) * cix;
        const float by = (__uint_as_float(w0.y) - r.oy) * ciy;
        const float bz = (__uint_as_float(w0.z) - r.oz) * ciz;
        // Rounding allowance, per axis (DESIGN.md §5 "wide8"): a plane distance computed here is
        // within d = 2^-21 (|b| + 256 |s|) of what slab() computes for any box the quantised one
        // contains (b: two roundings, the FMA one, slab()'s own two), so near planes are taken d
        // earlier and far planes d later.  Per axis, because a direction component close to zero
        // makes that axis's |b| huge: folded into one allowance for all axes it switched the
        // culling off for such rays.
        const float eps = 4.76837158203125e-07f;  // 2^-21
        const float dx8 = fmaf(256.0f, fabsf(sx), fabsf(bx)) * eps;
        const float dy8 = fmaf(256.0f, fabsf(sy), fabsf(by)) * eps;
        const float dz8 = fmaf(256.0f, fabsf(sz), fabsf(bz)) * eps;
        const float bnx = bx - dx8, bfx = bx + dx8, bny = by - dy8, bfy = by + dy8, bnz = bz - dz8, bfz = bz + dz8;
        // near / far planes by the sign of the direction
        const bool nx = r.dx < 0.0f, ny = r.dy < 0.0f, nz = r.dz < 0.0f;
        uint32_t hitmask = 0u;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const uint32_t meta4 = half ? w1.w : w1.z;
          const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
          const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
          const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
          const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
          const uint32_t qlx = half ? w2.y : w2.x, qly = half ? w2.w : w2.z, qlz = half ? w3.y : w3.x;
          const uint32_t qhx = half ? w3.w : w3.z, qhy = half ? w4.y : w4.x, qhz = half ? w4.w : w4.z;
          const uint32_t nqx = nx ? qhx : qlx, fqx = nx ? qlx : qhx;
          const uint32_t nqy = ny ? qhy : qly, fqy = ny ? qly : qhy;
          const uint32_t nqz = nz ? qhz : qlz, fqz = nz ? qlz : qhz;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float tnx = fmaf(byte_f(nqx, j), sx, bnx), tfx = fmaf(byte_f(fqx, j), sx, bfx);
            const float tny = fmaf(byte_f(nqy, j), sy, bny), tfy = fmaf(byte_f(fqy, j), sy, bfy);
            const float tnz = fmaf(byte_f(nqz, j), sz, bnz), tfz = fmaf(byte_f(fqz, j), sz, bfz);
            const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, r.tmin));
            const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, r.best_t));
            if (cmin <= cmax * kSlabPad)
              hitmask |= ((child_bits4 >> (8 * j)) & 0xFFu) << ((bit_index4 >> (8 * j)) & 0xFFu);
          }
        }
        ng = make_uint2(w1.x, (hitmask & 0xFF000000u) | (w0.w >> 24));
        tg = make_uint2(w1.y, hitmask & 0x00FFFFFFu);
        if (ng.y <= 0x00FFFFFFu) {  // no internal child was hit: back to a pending group
          if (sp > 0) ng = stack[--sp]; else ng.y = 0u;
        }
      }
    }

    // ---- one triangle per lane that holds some -------------------------------------------------
    if (has_ray && tg.y != 0u) {
      const uint32_t bit = 31u - (uint32_t)__clz((int)tg.y);
      tg.y &= ~(1u << bit);
      if (STATS) ++n_tests;
      const bool hit = WT ? intersect_wt(tris8, tg.x + bit, r, wr) : intersect(tris8, tg.x + bit, r);
      if (ANY && hit) {
        tg.y = 0u;
        ng.y = 0u;
        sp = 0;
      }
    }

    // ---- retire finished rays ------------------------------------------------------------------
    if (has_ray && tg.y == 0u && ng.y <= 0x00FFFFFFu) {
      st_stream(hits + my_ray, hit_record(r, ANY));
      has_ray = false;
    }
  }
  if (STATS) {
    n_nodes = __reduce_add_sync(0xFFFFFFFFu, n_nodes);
    n_tests = __reduce_add_sync(0xFFFFFFFFu, n_tests);
    if (lane == 0) {
      atomicAdd(stats, (unsigned long long)n_nodes);
      atomicAdd(stats + 1, (unsigned long long)n_tests);
    }
  }
}

}  // namespace

// ---- launchers -----------------------------------------------------------------------------------------
uint32_t wide8_leaf_max() { return kLeafMax8; }
// upper bound of the wide nodes of n triangles: every wide node but the root is an internal
// child holding more than kLeafMax8 triangles of its own, and disjoint ones hold disjoint sets
uint32_t wide8_node_capacity(uint64_t n) { return (uint32_t)(n / 2 + 64); }
constexpr int kLevels8 = 96;  // the wide tree is never deeper than the binary one
size_t wide8_scratch_bytes(uint64_t n) {
  // two item queues (ping-pong) + the counters
  return 2 * (size_t)wide8_node_capacity(n) * sizeof(Item8) + (4 + kLevels8 + 1) * sizeof(uint32_t) + 64;
}

// Collapses the BVH2 (split numbering) into `out` (wide8_node_capacity(n) slots) and tris8 (n
// records).  Launches the levels in batches and reads the queue length back after each batch;
// *n_nodes_out <- wide nodes used.  Synchronises the stream.
int launch_collapse8(const Node* nodes, const SortedTri* stris, uint64_t n, int root, void* scratch,
                     Node8* out, SortedTri* tris8, uint32_t* n_nodes_out, cudaStream_t s) {
  if (n <= kLeafMax8) return fail(-1, "wide8: needs more than %u triangles", kLeafMax8);
  const uint32_t capacity = wide8_node_capacity(n);
  Item8* q[2] = {reinterpret_cast<Item8*>(scratch), reinterpret_cast<Item8*>(scratch) + capacity};
  uint32_t* counters = reinterpret_cast<uint32_t*>(q[1] + capacity);
  uint32_t init[5] = {1u, 0u, 0u, 0u, 1u};  // node 0 = the root, already allocated; level 0 holds one item
  RT_CUDA(cudaMemsetAsync(counters, 0, (4 + kLevels8 + 1) * sizeof(uint32_t), s));
  RT_CUDA(cudaMemcpyAsync(counters, init, sizeof(init), cudaMemcpyHostToDevice, s));
  const Item8 first = {(uint32_t)root, 0u};
  RT_CUDA(cudaMemcpyAsync(q[0], &first, sizeof(first), cudaMemcpyHostToDevice, s));
  const unsigned grid = (unsigned)num_sms() * 8;
  int level = 0;
  uint32_t host[4 + kLevels8 + 1];
  while (true) {
    const int batch_end = level + 12 < kLevels8 ? level + 12 : kLevels8;
    for (; level < batch_end; ++level) {
      wide8_level_kernel<<<grid, 128, 0, s>>>(nodes, stris, q[level & 1], q[(level + 1) & 1], counters, level,
                                              capacity, out, tris8);
      RT_LAUNCHED();
    }
    RT_CUDA(cudaMemcpyAsync(host, counters, sizeof(host), cudaMemcpyDeviceToHost, s));
    RT_CUDA(cudaStreamSynchronize(s));
    if (host[2]) return fail(-2, "wide8: more than %u wide nodes", capacity);
    if (host[4 + level] == 0u) break;
    if (level >= kLevels8) return fail(-2, "wide8: the tree is deeper than %d levels", kLevels8);
  }
  if (host[1] != (uint32_t)n) return fail(-2, "wide8: %u of %llu triangles placed", host[1], (unsigned long long)n);
  *n_nodes_out = host[0];
  return 0;
}

int launch_trace_wide8(const Node8* nodes8, const SortedTri* tris8, const float4* rays, uint64_t n_rays,
                       uint32_t flags, float4* hits, unsigned long long* ray_counter, cudaStream_t s,
                       unsigned long long* stats, const uint32_t* perm) {
  if (n_rays == 0) return 0;
  RT_CUDA(cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), s));
  const Params& p = params();
  int bps = p.trace_blocks_per_sm;
  if (bps <= 0) {  // persistent grid: exactly the CTAs that fit (register-limited)
    static std::atomic<int> resident{0};
    int occ = resident.load(std::memory_order_relaxed);
    if (occ == 0) {
      RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_wide8_kernel<false, false, false>,
                                                            kTraceThreads, 0));
      if (occ < 1) occ = 1;
      resident.store(occ, std::memory_order_relaxed);
    }
    bps = occ;
  }
  uint64_t blocks = (uint64_t)num_sms() * bps;
  const uint64_t need = (n_rays + kTraceThreads - 1) / kTraceThreads;
  if (blocks > need) blocks = need;
  const bool any = (flags & 1u) != 0, wt = (flags & 2u) != 0;
  int chunk = p.trace_chunk > 0 ? (int)p.trace_chunk : 64;
  if (p.trace_chunk <= 0 && n_rays < blocks * (kTraceThreads / 32) * (uint64_t)chunk * 2) chunk = 32;
  const int refill_min = p.trace_refill_min > 0 ? (int)p.trace_refill_min : (any ? 16 : 8);
  const int leaf_min = p.trace_leaf_min > 0 ? (int)p.trace_leaf_min : 8;
#define RT_TRACE_WIDE8(A, S, W)                                              \
  trace_wide8_kernel<A, S, W><<<(unsigned)blocks, kTraceThreads, 0, s>>>(    \
      nodes8, tris8, rays, n_rays, hits, ray_counter, stats, chunk, refill_min, leaf_min, perm)
  RT_DISPATCH3(RT_TRACE_WIDE8, any, stats != nullptr, wt);
#undef RT_TRACE_WIDE8
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
