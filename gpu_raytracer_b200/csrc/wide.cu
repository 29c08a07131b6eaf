// wide.cu — A5w: collapse of the binary LBVH into a 4-wide BVH, and its traversal
// (Wald et al. 2008 / Ylitie et al. 2017 for the idea).
//
// Collapse.  One thread per BVH2 node g (split numbering, refit.cu).  The node's two
// children are the first two slots; up to two times the slot chosen by the rule
//     1. among slots holding more than leaf_max triangles (rt_set_param
//        "wide.leaf_max", 1..4, default 2), the one with the
//        largest half surface area; if there is none,
//     2. among slots holding more than one triangle, the one with the largest area;
//     ties go to the leftmost slot (smallest first leaf), areas are computed as
//     (dx*dy + dy*dz) + dz*dx in float with no contraction
// is replaced, in place, by its own two children, so the slots stay in left-to-right
// leaf order.  A slot that ends with at most leaf_max triangles becomes a leaf
// range [first, first+count) of the sorted triangle array; any other slot refers to
// the Node4 of its BVH2 node.  Node4[g] is written at the index of its BVH2 node, so
// there is no queue, no allocation, no level loop and no dependency between threads:
// the collapse is one streaming kernel (reads 64 B + up to 2 x 64 B of cached
// children, writes 128 B per node).  Node4s of subtrees with <= leaf_max
// triangles are never referenced and are not written.  The price is memory: the
// array has n-1 slots of 128 B, of which only those reachable from the root are
// ever read.
//
// Traversal.  The same persistent-warp scheme as traverse.cu (per-lane dynamic ray
// fetch, warp-voted switch between node steps and triangle steps).  A node step is
// one 128-byte fetch (3 x LDG.256 + LDG.128, four sectors of one line) and four slab tests;
// the children that are hit are ordered by entry distance with a 5-exchange sorting
// network on (distance bits | slot) keys, the nearest is visited next, the others
// are pushed farthest first.  A leaf step tests one triangle of the leaf's range.
// The slab and triangle arithmetic and the acceptance rule (closest t, equal t ->
// smaller primitive index) are the shared code of trace_common.cuh, so hits are
// bit-identical to the BVH2 traversal and to the oracle whatever the order.
//
// Stack bound: a wide node's children are BVH2 descendants one or two levels down,
// so the wide depth is at most the BVH2 depth (<= 95: 63 key bits + 32 tie-break
// bits) and at most three siblings are pending per level: 3 * 95 + 3 entries.
#include "trace_common.cuh"

namespace rt {
namespace {

struct Slot {
  float lx, ly, lz, hx, hy, hz;
  int enc;             // BVH2 child id: >= 0 split position, < 0 leaf (~sorted position)
  uint32_t first, cnt; // leaf range of the subtree
};

__device__ __forceinline__ float half_area(const Slot& s) {
  const float dx = s.hx - s.lx, dy = s.hy - s.ly, dz = s.hz - s.lz;
  return (dx * dy + dy * dz) + dz * dx;  // -fmad=false: three products, two sums
}

__device__ __forceinline__ void load_children(const Node* __restrict__ nodes, uint32_t x, Slot& a,
                                              Slot& b) {
  const float4* np = reinterpret_cast<const float4*>(nodes + x);
  const float4 lo0 = __ldg(np), hi0 = __ldg(np + 1), lo1 = __ldg(np + 2), hi1 = __ldg(np + 3);
  const uint32_t L = __float_as_uint(hi0.w), R = __float_as_uint(hi1.w);
  a.lx = lo0.x; a.ly = lo0.y; a.lz = lo0.z; a.hx = hi0.x; a.hy = hi0.y; a.hz = hi0.z;
  a.enc = __float_as_int(lo0.w); a.first = L; a.cnt = x - L + 1u;
  b.lx = lo1.x; b.ly = lo1.y; b.lz = lo1.z; b.hx = hi1.x; b.hy = hi1.y; b.hz = hi1.z;
  b.enc = __float_as_int(lo1.w); b.first = x + 1u; b.cnt = R - x;
}

__global__ void __launch_bounds__(256)
collapse4_kernel(const Node* __restrict__ nodes, uint32_t n, uint32_t leaf_max,
                 Node4* __restrict__ wide) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g + 1u >= n) return;
  Slot s[4];
  load_children(nodes, g, s[0], s[1]);
  if (s[0].cnt + s[1].cnt <= leaf_max) return;  // a leaf of its parent: never referenced
  int ns = 2;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    int best = -1;
    float best_area = 0.0f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (k < ns && s[k].cnt > leaf_max) {
        const float a = half_area(s[k]);
        if (best < 0 || a > best_area) { best = k; best_area = a; }
      }
    }
    if (best < 0) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (k < ns && s[k].cnt > 1u) {
          const float a = half_area(s[k]);
          if (best < 0 || a > best_area) { best = k; best_area = a; }
        }
      }
    }
    if (best < 0) break;
    int x = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (k == best) x = s[k].enc;
    Slot a, b;
    load_children(nodes, (uint32_t)x, a, b);
#pragma unroll
    for (int k = 3; k >= 1; --k)
      if (k > best + 1) s[k] = s[k - 1];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (k == best) s[k] = a;
      else if (k == best + 1) s[k] = b;
    }
    ++ns;
  }
  float lo[3][4], hi[3][4];
  uint32_t ref[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (k < ns) {
      lo[0][k] = s[k].lx; lo[1][k] = s[k].ly; lo[2][k] = s[k].lz;
      hi[0][k] = s[k].hx; hi[1][k] = s[k].hy; hi[2][k] = s[k].hz;
      ref[k] = (s[k].cnt <= leaf_max)
                   ? (kWideLeafBit | ((s[k].cnt - 1u) << 29) | s[k].first)
                   : (uint32_t)s[k].enc;
    } else {
      lo[0][k] = lo[1][k] = lo[2][k] = 0.0f;
      hi[0][k] = hi[1][k] = hi[2][k] = 0.0f;
      ref[k] = kWideEmpty;
    }
  }
  float4* wp = reinterpret_cast<float4*>(wide + g);
  wp[0] = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]);
  wp[1] = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
  wp[2] = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]);
  wp[3] = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
  wp[4] = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]);
  wp[5] = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
  reinterpret_cast<uint4*>(wp)[6] = make_uint4(ref[0], ref[1], ref[2], ref[3]);
  reinterpret_cast<uint4*>(wp)[7] = make_uint4((uint32_t)ns, 0u, 0u, 0u);
}

// ---- traversal ----------------------------------------------------------------------
constexpr int kWideStack = 3 * 95 + 3;

__device__ __forceinline__ void cswap(uint32_t& a, uint32_t& b) {
  const uint32_t lo = min(a, b), hi = max(a, b);
  a = lo;
  b = hi;
}
__device__ __forceinline__ uint32_t pick(const uint4 v, uint32_t k) {
  return (k & 2u) ? ((k & 1u) ? v.w : v.z) : ((k & 1u) ? v.y : v.x);
}

template <bool ANY, bool STATS, bool WT>
__global__ void __launch_bounds__(kTraceThreads)
trace_wide_kernel(const Node4* __restrict__ wide, const SortedTri* __restrict__ stris,
                  uint32_t root_ref, const float4* __restrict__ rays, unsigned long long n_rays,
                  float4* __restrict__ hits, unsigned long long* __restrict__ ray_counter,
                  unsigned long long* __restrict__ stats, int chunk, int refill_min, int leaf_min) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t n_nodes = 0, n_tests = 0;

  uint32_t stack[kWideStack];
  int sp = 0;
  uint32_t cur = kWideEmpty;
  bool has_ray = false;
  unsigned long long my_ray = 0;
  RayState r;
  r.best_prim = -1;
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
        load_ray(r, ld_stream(rays + 2 * my_ray), ld_stream(rays + 2 * my_ray + 1));
        if (WT) wr = wt_setup(r);
        has_ray = true;
        sp = 0;
        cur = root_ref;
      }
      const unsigned long long need = (unsigned long long)__popc(idle);
      chunk_next += (need < avail) ? need : avail;
    }
    if (__ballot_sync(0xFFFFFFFFu, has_ray) == 0u) {
      if (exhausted) break;
      continue;
    }

    // ---- node steps, until enough lanes wait at a leaf ---------------------------
    while (true) {
      const bool at_node = has_ray && cur < kWideEmpty;
      const uint32_t inner = __ballot_sync(0xFFFFFFFFu, at_node);
      if (inner == 0u) break;
      const uint32_t waiting = __ballot_sync(0xFFFFFFFFu, has_ray && !at_node);  // leaf or finished
      if (__popc(waiting) >= leaf_min) break;
      if (at_node) {
        if (STATS) ++n_nodes;
        const float4* np = reinterpret_cast<const float4*>(wide + cur);
        const F8 q0 = ldg256(np), q1 = ldg256(np + 2), q2 = ldg256(np + 4);
        const float4 lox = q0.a, loy = q0.b, loz = q1.a, hix = q1.b, hiy = q2.a, hiz = q2.b;
        const uint4 ref = __ldg(reinterpret_cast<const uint4*>(np) + 6);
        float t0, t1, t2, t3;
        const bool h0 = slab(lox.x, loy.x, loz.x, hix.x, hiy.x, hiz.x, r, t0) && ref.x != kWideEmpty;
        const bool h1 = slab(lox.y, loy.y, loz.y, hix.y, hiy.y, hiz.y, r, t1) && ref.y != kWideEmpty;
        const bool h2 = slab(lox.z, loy.z, loz.z, hix.z, hiy.z, hiz.z, r, t2) && ref.z != kWideEmpty;
        const bool h3 = slab(lox.w, loy.w, loz.w, hix.w, hiy.w, hiz.w, r, t3) && ref.w != kWideEmpty;
        // (entry distance | slot) keys; a miss sorts last
        uint32_t k0 = h0 ? ((__float_as_uint(t0) & ~3u) | 0u) : 0xFFFFFFFFu;
        uint32_t k1 = h1 ? ((__float_as_uint(t1) & ~3u) | 1u) : 0xFFFFFFFFu;
        uint32_t k2 = h2 ? ((__float_as_uint(t2) & ~3u) | 2u) : 0xFFFFFFFFu;
        uint32_t k3 = h3 ? ((__float_as_uint(t3) & ~3u) | 3u) : 0xFFFFFFFFu;
        cswap(k0, k1); cswap(k2, k3); cswap(k0, k2); cswap(k1, k3); cswap(k1, k2);
        if (k3 != 0xFFFFFFFFu) stack[sp++] = pick(ref, k3);
        if (k2 != 0xFFFFFFFFu) stack[sp++] = pick(ref, k2);
        if (k1 != 0xFFFFFFFFu) stack[sp++] = pick(ref, k1);
        if (k0 != 0xFFFFFFFFu) cur = pick(ref, k0);
        else cur = (sp > 0) ? stack[--sp] : kWideEmpty;
      }
    }

    // ---- one triangle per waiting lane ---------------------------------------------
    if (has_ray && (cur & kWideLeafBit)) {
      if (STATS) ++n_tests;
      const bool hit = WT ? intersect_wt(stris, cur & kWideFirstMask, r, wr)
                          : intersect(stris, cur & kWideFirstMask, r);
      if (ANY && hit) {
        cur = kWideEmpty;
        sp = 0;
      } else if (cur & (3u << 29)) {
        cur += 1u - (1u << 29);  // next triangle of the leaf: first + 1, count - 1
      } else {
        cur = (sp > 0) ? stack[--sp] : kWideEmpty;
      }
    }

    // ---- retire finished rays ------------------------------------------------------
    if (has_ray && cur == kWideEmpty) {
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

int launch_collapse4(const Node* nodes, uint64_t n, uint32_t leaf_max, Node4* wide, cudaStream_t s) {
  if (n > kWideMaxTris) return fail(-5, "wide BVH: %llu triangles exceed 2^29", (unsigned long long)n);
  if (n > (uint64_t)leaf_max) {
    collapse4_kernel<<<(unsigned)((n - 1 + 255) / 256), 256, 0, s>>>(nodes, (uint32_t)n, leaf_max, wide);
    RT_LAUNCHED();
  }
  return 0;
}

int launch_trace_wide(const Node4* wide, const SortedTri* stris, uint64_t n_tris, uint32_t root_ref,
                      const float4* rays, uint64_t n_rays, uint32_t flags, float4* hits,
                      unsigned long long* ray_counter, cudaStream_t s, unsigned long long* stats) {
  if (n_rays == 0) return 0;
  RT_CUDA(cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), s));
  const Params& p = params();
  int bps = p.trace_blocks_per_sm;
  if (bps <= 0) {  // persistent grid: exactly the CTAs that fit (register-limited)
    static std::atomic<int> resident{0};
    int occ = resident.load(std::memory_order_relaxed);
    if (occ == 0) {
      RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_wide_kernel<false, false, false>,
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
  int chunk = p.trace_chunk > 0 ? p.trace_chunk : 64;
  if (p.trace_chunk <= 0 && n_rays < blocks * (kTraceThreads / 32) * (uint64_t)chunk * 2) chunk = 32;
  const int refill_min = p.trace_refill_min > 0 ? p.trace_refill_min : 16;
  const int leaf_min = p.trace_leaf_min > 0 ? p.trace_leaf_min : 8;
  if (n_tris == 0) root_ref = kWideEmpty;
#define RT_TRACE_WIDE(A, S, W)                                              \
  trace_wide_kernel<A, S, W><<<(unsigned)blocks, kTraceThreads, 0, s>>>(    \
      wide, stris, root_ref, rays, n_rays, hits, ray_counter, stats, chunk, refill_min, leaf_min)
  RT_DISPATCH3(RT_TRACE_WIDE, any, stats != nullptr, wt);
#undef RT_TRACE_WIDE
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
