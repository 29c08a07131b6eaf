// refit.cu — A4 + A5 in one bottom-up pass: radix-tree topology, leaf boxes,
// triangle reorder and the AABB refit, plus the
// table-driven refit of refit_bvh and the kernels that turn the traversal layout
// back into the canonical arrays the parity tests compare.
//
// Topology without searches (Apetrei 2014).  A finished node covers the sorted
// leaves [L, R].  With delta(i) = common-prefix length of the (index-augmented)
// keys i and i+1, its parent merges it with the neighbour it shares the longer
// prefix with: delta(R) > delta(L-1) makes it the LEFT child of the node that
// splits after R, otherwise the RIGHT child of the node that splits after L-1.
// Two children meet at their split position g, so the rendezvous counter, and
// the 64-byte node itself, are indexed by g ("split numbering").  This is the
// same tree as Karras 2012: the Karras index of the child in slot s of node g is
// g + s, the root is Karras node 0, which is how export_kernel produces the
// canonical arrays.  A top-down Karras kernel (two binary searches per node, 11
// of 32 lanes active on average) used to take 0.28 ms for 10M leaves before the
// refit even started; here the topology costs two byte loads per climb step.
//
// One thread per leaf, one CTA per kRefitLeaves = 256 consecutive leaves (measured:
// 1024 -> 0.65 ms, 512 -> 0.62, 256 -> 0.59, 128 -> 0.59, 64 -> 0.92 ms for 10M).  The
// thread gathers its triangle through the sorted index, stores it in traversal
// form {v0 | prim, v1, v2}, then climbs: it writes the box it carries (child
// id, far end of its range) into its slot of node g — one full 32-byte sector —
// and meets its sibling.
//   phase 1  meetings whose split lies inside the CTA's leaf group happen in
//            shared memory with CTA-scope fences: deposit, atomicOr, first
//            arriver stops, second unions and climbs on.  No global atomics, no
//            MEMBAR.GPU (55 % of the stall samples of the first version).
//   phase 2  after a CTA barrier, a slot that saw only one child means the
//            sibling lives in another CTA: that child, and every walker whose
//            next split falls on a group boundary, is parked in a queue.
//   top      refit_top_kernel, one dense thread per parked walker, finishes
//            the upper ~3 % of the tree with the device-scope protocol (global
//            counter, __threadfence, sibling slot re-read from L2 with ld.cg).
// Second arrivers also record the parent word of both children (8 B per node),
// which is all refit_bvh needs to re-run the boxes later without keys.
// The walker that completes [0, n-1] stores the scene box and the root id in
// the header.
//
// min/max use the total order on bit patterns, so boxes are exact and
// independent of which thread arrives first.
#include "common.cuh"
#include "kernels.h"

namespace rt {
namespace {

__device__ __forceinline__ float4 ld_cg(const float4* p) {
  float4 r;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

struct Box {
  float lx, ly, lz, hx, hy, hz;
};

// CTA-scope release/acquire folded into the arrival atomic itself: the deposit
// written before it is visible to whoever sees the bit, and what the sibling
// deposited is visible after it.  (A separate __threadfence_block() compiles to
// MEMBAR.SC.CTA, which also waits for the walker's outstanding global stores.)
__device__ __forceinline__ uint32_t arrive_or_cta(uint32_t* flag, uint32_t bit) {
  uint32_t old;
  asm volatile("atom.acq_rel.cta.shared.or.b32 %0, [%1], %2;"
               : "=r"(old)
               : "r"((uint32_t)__cvta_generic_to_shared(flag)), "r"(bit)
               : "memory");
  return old;
}
__device__ __forceinline__ uint32_t arrive_add_cta(uint32_t* flag) {
  uint32_t old;
  asm volatile("atom.acq_rel.cta.shared.add.u32 %0, [%1], 1;"
               : "=r"(old)
               : "r"((uint32_t)__cvta_generic_to_shared(flag))
               : "memory");
  return old;
}
__device__ __forceinline__ float4 lds_volatile_v4(const void* p) {
  float4 r;
  asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "r"((uint32_t)__cvta_generic_to_shared(p))
               : "memory");
  return r;
}

// Shared state of one CTA = kRefitLeaves consecutive leaves: arrival bits and
// deposited children of the nodes whose split position lies inside the group
// (index = split position - first leaf of the CTA).
// Deposits are stored as planes of float4 indexed [slot][k]: neighbouring walkers
// deposit at neighbouring k, so a 16-byte stride keeps a warp's STS.128 / LDS.128
// free of bank conflicts (the array-of-structs layout was 4-way conflicted).
struct RefitSmem {
  float4 dep_lo[2][kRefitLeaves];  // min.xyz, child id
  float4 dep_hi[2][kRefitLeaves];  // max.xyz, far end of the child's leaf range
  uint32_t flag[kRefitLeaves];
  uint8_t delta[kRefitLeaves + 8];  // delta(base - 1 + t), t in [0, kRefitLeaves]
};

// A walker that leaves its CTA's group parks here instead of keeping a whole
// CTA resident while it climbs the (slow, device-scope) top of the tree.
struct __align__(16) Continuation {
  float lx, ly, lz;
  int child_enc;
  float hx, hy, hz;
  uint32_t pword;  // table-driven refit: parent word.  build: unused
  uint32_t L, R, pad0, pad1;
};
// queue capacity: one global bound, about twice what the scenes in BASELINE.json park
// with 256-leaf groups; a walker that finds the queue full climbs inline
static uint32_t queue_capacity(uint64_t n) { return (uint32_t)(n / 8 + 4096); }

__device__ __forceinline__ uint32_t node_slot_of(int child_enc, uint32_t n) {
  return child_enc < 0 ? (n - 1u) + (uint32_t)(~child_enc) : (uint32_t)child_enc;
}

// common-prefix length of the augmented keys i and i+1 (i + 1 < n)
template <typename KeyT>
__device__ __forceinline__ uint32_t delta_adj(KeyT a, KeyT b, uint32_t i) {
  if (sizeof(KeyT) == 4) {
    const uint32_t x = (uint32_t)a ^ (uint32_t)b;
    return x ? (uint32_t)__clz((int)x) : 32u + (uint32_t)__clz((int)(i ^ (i + 1u)));
  } else {
    const uint64_t x = (uint64_t)a ^ (uint64_t)b;
    return x ? (uint32_t)__clzll((long long)x) : 64u + (uint32_t)__clz((int)(i ^ (i + 1u)));
  }
}

__device__ __forceinline__ void store_root(const Box& bx, int enc, BvhHeader* header) {
  header->root_min[0] = bx.lx; header->root_min[1] = bx.ly; header->root_min[2] = bx.lz;
  header->root_max[0] = bx.hx; header->root_max[1] = bx.hy; header->root_max[2] = bx.hz;
  header->root_node = (uint32_t)enc;
}

// ---- build: device-scope climb with on-the-fly topology -------------------------
// One level costs two dependent L2 round trips: the arrival exchange, then (second
// arriver only) the sibling's box together with the four keys that decide the next
// parent.  The arrival word carries the first arriver's far end + 1, so the second one
// knows the merged range, and can ask for those keys, before it has seen the box.
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

template <typename KeyT>
__device__ __forceinline__ void climb_build_global(Box bx, int enc, uint32_t L, uint32_t R,
                                                   const KeyT* __restrict__ keys, uint32_t n,
                                                   uint32_t* __restrict__ flags,
                                                   Node* __restrict__ nodes,
                                                   BvhHeader* __restrict__ header) {
  // keys around both ends of [L, R]; out-of-range ends are never compared
  KeyT kl0 = 0, kl1 = 0, kr0 = 0, kr1 = 0;
  auto fetch = [&]() {
    if (L != 0u) { kl0 = __ldg(keys + L - 1); kl1 = __ldg(keys + L); }
    if (R != n - 1u) { kr0 = __ldg(keys + R); kr1 = __ldg(keys + R + 1); }
  };
  fetch();
  while (true) {
    if (L == 0u && R == n - 1u) {
      store_root(bx, enc, header);
      return;
    }
    bool left = (L == 0u);
    if (!left && R != n - 1u)
      left = delta_adj<KeyT>(kr0, kr1, R) > delta_adj<KeyT>(kl0, kl1, L - 1);
    const uint32_t g = left ? R : L - 1u, slot = left ? 0u : 1u;
    float4* np = reinterpret_cast<float4*>(nodes + g);
    st_v8(np + 2 * slot, make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(enc)),
          make_float4(bx.hx, bx.hy, bx.hz, __uint_as_float(left ? L : R)));
    fence_acq_rel_gpu();
    const uint32_t old = atomicExch(flags + g, (left ? L : R) + 1u);
    if (old == 0u) return;  // first arriver stops
    if (left) R = old - 1u; else L = old - 1u;
    fetch();  // next level's keys travel with the sibling's box
    // acquire side: the loads below are issued only after the exchange's result is
    // known (control dependency) and go to L2 (ld.cg), where the sibling's fenced
    // stores already are
    const float4 slo = ld_cg(np + 2 * (slot ^ 1u));
    const float4 shi = ld_cg(np + 2 * (slot ^ 1u) + 1);
    bx.lx = omin(bx.lx, slo.x); bx.ly = omin(bx.ly, slo.y); bx.lz = omin(bx.lz, slo.z);
    bx.hx = omax(bx.hx, shi.x); bx.hy = omax(bx.hy, shi.y); bx.hz = omax(bx.hz, shi.z);
    enc = (int)g;
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
build_top_kernel(const Continuation* __restrict__ queue, const uint32_t* __restrict__ queue_count,
                 uint32_t capacity, const KeyT* __restrict__ keys, uint32_t n,
                 uint32_t* __restrict__ flags, Node* __restrict__ nodes,
                 BvhHeader* __restrict__ header, const uint32_t* __restrict__ abort_flag) {
  pdl_enter();
  if (abort_flag != nullptr && *abort_flag != 0u) return;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t count = min(*queue_count, capacity);
  if (i >= count) return;
  const float4 a = reinterpret_cast<const float4*>(queue + i)[0];
  const float4 b = reinterpret_cast<const float4*>(queue + i)[1];
  const uint4 c = reinterpret_cast<const uint4*>(queue + i)[2];
  Box bx{a.x, a.y, a.z, b.x, b.y, b.z};
  climb_build_global<KeyT>(bx, __float_as_int(a.w), c.x, c.y, keys, n, flags, nodes, header);
}

template <typename KeyT>
__device__ __forceinline__ void park_or_climb(const Box& bx, int enc, uint32_t L, uint32_t R,
                                              Continuation* __restrict__ queue,
                                              uint32_t* __restrict__ queue_count, uint32_t capacity,
                                              const KeyT* __restrict__ keys, uint32_t n,
                                              uint32_t* __restrict__ flags, Node* __restrict__ nodes,
                                              BvhHeader* __restrict__ header) {
  const uint32_t q = atomicAdd(queue_count, 1u);
  if (q < capacity) {
    float4* dst = reinterpret_cast<float4*>(queue + q);
    dst[0] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(enc));
    dst[1] = make_float4(bx.hx, bx.hy, bx.hz, 0.0f);
    reinterpret_cast<uint4*>(dst)[2] = make_uint4(L, R, 0u, 0u);
  } else {
    climb_build_global<KeyT>(bx, enc, L, R, keys, n, flags, nodes, header);  // queue full
  }
}

// Shared state of one build CTA: G = THREADS x LPT consecutive leaves (the table-driven refit
// keeps its own group size, kRefitLeaves; the two are independent).
template <int G>
struct BuildSmem {
  float4 dep_lo[2][G];  // min.xyz, child id
  float4 dep_hi[2][G];  // max.xyz, far end of the child's leaf range
  uint32_t flag[G];
  uint8_t delta[G + 8];  // delta(base - 1 + t), t in [0, G]
};

// LPT consecutive leaves per thread, climbed one after the other through the same meeting
// protocol (a first arriver deposits and stops, whoever comes second — another thread, or
// the same thread with its next leaf — unions and climbs on).  The lowest levels of the
// tree, where most meetings happen, are then mostly meetings of a thread with itself, the
// work per thread is the sum of LPT climbs instead of one (less variance inside a warp: ncu
// showed 9.9 of 32 lanes active with one leaf per thread), a thread has 3 x LPT independent
// triangle loads in flight, and a group of G leaves parks 1/LPT of the walkers per leaf.
template <typename KeyT, int THREADS, int LPT>
__global__ void __launch_bounds__(THREADS)
build_walk_kernel(const float4* __restrict__ tris, const uint32_t* __restrict__ sorted_idx,
                  const KeyT* __restrict__ keys, uint32_t n, uint32_t* __restrict__ flags,
                  Node* __restrict__ nodes,
                  SortedTri* __restrict__ stris, BvhHeader* __restrict__ header,
                  Continuation* __restrict__ queue, uint32_t* __restrict__ queue_count,
                  uint32_t capacity, const uint32_t* __restrict__ abort_flag) {
  constexpr int G = THREADS * LPT;
  extern __shared__ __align__(16) unsigned char refit_smem_raw[];
  BuildSmem<G>& sm = *reinterpret_cast<BuildSmem<G>*>(refit_smem_raw);
  pdl_enter();
  if (abort_flag != nullptr && *abort_flag != 0u) return;  // the sort gave up (hybrid order, long run)
  const uint32_t tid = threadIdx.x;
  const uint32_t base = blockIdx.x * G;
  // delta(i) for i = base - 1 + t, t in [0, G]
  for (uint32_t t = tid; t <= (uint32_t)G; t += THREADS) {
    const uint32_t j = base + t;
    if (t < (uint32_t)G) sm.flag[t] = 0;
    if (j >= 1u && j < n) sm.delta[t] = (uint8_t)delta_adj<KeyT>(keys[j - 1u], keys[j], j - 1u);
  }
  // the thread's leaves: gather the triangles (all loads first), store them in traversal form
  const uint32_t j0 = base + tid * LPT;
  Box bx[LPT];
  uint32_t prim[LPT];
  float4 va[LPT], vb[LPT], vc[LPT];
#pragma unroll
  for (int i = 0; i < LPT; ++i) prim[i] = (j0 + i < n) ? sorted_idx[j0 + i] : 0u;
#pragma unroll
  for (int i = 0; i < LPT; ++i) {
    if (j0 + i < n) {
      va[i] = __ldg(tris + 3 * (uint64_t)prim[i]);
      vb[i] = __ldg(tris + 3 * (uint64_t)prim[i] + 1);
      vc[i] = __ldg(tris + 3 * (uint64_t)prim[i] + 2);
    }
  }
#pragma unroll
  for (int i = 0; i < LPT; ++i) {
    const uint32_t j = j0 + i;
    if (j < n) {
      const float4 a = va[i], b = vb[i], c = vc[i];
      // 48-byte records: the 32-byte aligned part goes out as one full sector
      float4* st = reinterpret_cast<float4*>(stris + j);
      const float4 r0 = make_float4(a.x, a.y, a.z, __uint_as_float(prim[i]));
      const float4 r1 = make_float4(b.x, b.y, b.z, 0.0f), r2 = make_float4(c.x, c.y, c.z, 0.0f);
      if ((j & 1u) == 0u) {
        st_v8(st, r0, r1);
        st[2] = r2;
      } else {
        st[0] = r0;
        st_v8(st + 1, r1, r2);
      }
      bx[i].lx = omin(omin(a.x, b.x), c.x); bx[i].ly = omin(omin(a.y, b.y), c.y); bx[i].lz = omin(omin(a.z, b.z), c.z);
      bx[i].hx = omax(omax(a.x, b.x), c.x); bx[i].hy = omax(omax(a.y, b.y), c.y); bx[i].hz = omax(omax(a.z, b.z), c.z);
    }
  }
  __syncthreads();

  // ---- phase 1: meetings inside the group, in shared memory -----------------------
#pragma unroll 1
  for (int i = 0; i < LPT; ++i) {
    const uint32_t j = j0 + i;
    Box b = bx[i];
    int enc = ~(int)j;
    uint32_t L = j, R = j;
    bool alive = j < n;
    while (alive) {
      if (L == 0u && R == n - 1u) {
        store_root(b, enc, header);
        break;
      }
      // L >= base and R < base + G hold for every walker in this loop
      bool left = (L == 0u);
      if (!left && R != n - 1u) left = sm.delta[R - base + 1u] > sm.delta[L - base];
      const uint32_t g = left ? R : L - 1u, slot = left ? 0u : 1u;
      const float4 my_lo = make_float4(b.lx, b.ly, b.lz, __int_as_float(enc));
      const float4 my_hi = make_float4(b.hx, b.hy, b.hz, __uint_as_float(left ? L : R));
      float4* np = reinterpret_cast<float4*>(nodes + g) + 2 * slot;
      if (g < base || g + 1u >= base + (uint32_t)G) {
        // the sibling starts or ends in another group: device-scope meeting
        park_or_climb<KeyT>(b, enc, L, R, queue, queue_count, capacity, keys, n, flags, nodes, header);
        break;
      }
      const uint32_t k = g - base;
      sm.dep_lo[slot][k] = my_lo;
      sm.dep_hi[slot][k] = my_hi;
      const uint32_t seen = arrive_or_cta(&sm.flag[k], 1u << slot);
      // the node half goes out after the meeting, so no fence ever waits for it
      st_v8(np, my_lo, my_hi);
      if (seen == 0u) break;  // first arriver stops
      const float4 slo = lds_volatile_v4(&sm.dep_lo[slot ^ 1u][k]);
      const float4 shi = lds_volatile_v4(&sm.dep_hi[slot ^ 1u][k]);
      b.lx = omin(b.lx, slo.x); b.ly = omin(b.ly, slo.y); b.lz = omin(b.lz, slo.z);
      b.hx = omax(b.hx, shi.x); b.hy = omax(b.hy, shi.y); b.hz = omax(b.hz, shi.z);
      if (left) R = __float_as_uint(shi.w); else L = __float_as_uint(shi.w);
      enc = (int)g;
    }
  }
  __syncthreads();

  // ---- phase 2: a slot with one child only waits for a sibling from elsewhere --------
  for (uint32_t k = tid; k < (uint32_t)G; k += THREADS) {
    const uint32_t f = sm.flag[k];
    if (f == 1u || f == 2u) {
      const uint32_t slot = f - 1u, g = base + k;
      const float4 dlo = sm.dep_lo[slot][k], dhi = sm.dep_hi[slot][k];
      Box b2{dlo.x, dlo.y, dlo.z, dhi.x, dhi.y, dhi.z};
      const uint32_t far_end = __float_as_uint(dhi.w);
      const uint32_t l2 = slot ? g + 1u : far_end, r2 = slot ? far_end : g;
      park_or_climb<KeyT>(b2, __float_as_int(dlo.w), l2, r2, queue, queue_count, capacity, keys, n, flags,
                          nodes, header);
    }
  }
}

// ---- refit_bvh: table-driven (parent words recorded by the build) -------------------
// the device-scope protocol, used for nodes whose children belong to different CTAs
__device__ __forceinline__ void climb_global(Box bx, int child_enc, uint32_t pword,
                                             const uint32_t* __restrict__ parent,
                                             uint32_t* __restrict__ flags, Node* __restrict__ nodes,
                                             BvhHeader* __restrict__ header) {
  while (pword != kNoParent) {
    const uint32_t p = pword >> 2, slot = pword & 1u;
    const uint32_t next_pword = parent[p];
    float4* np = reinterpret_cast<float4*>(nodes + p);
    const float keep = ld_cg(np + 2 * slot + 1).w;  // far end of the range: topology, kept
    np[2 * slot] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(child_enc));
    np[2 * slot + 1] = make_float4(bx.hx, bx.hy, bx.hz, keep);
    __threadfence();
    if (atomicAdd(flags + p, 1u) == 0u) return;
    const float4 slo = ld_cg(np + 2 * (slot ^ 1u));
    const float4 shi = ld_cg(np + 2 * (slot ^ 1u) + 1);
    bx.lx = omin(bx.lx, slo.x); bx.ly = omin(bx.ly, slo.y); bx.lz = omin(bx.lz, slo.z);
    bx.hx = omax(bx.hx, shi.x); bx.hy = omax(bx.hy, shi.y); bx.hz = omax(bx.hz, shi.z);
    child_enc = (int)p;
    pword = next_pword;
  }
  header->root_min[0] = bx.lx; header->root_min[1] = bx.ly; header->root_min[2] = bx.lz;
  header->root_max[0] = bx.hx; header->root_max[1] = bx.hy; header->root_max[2] = bx.hz;
}

__global__ void __launch_bounds__(256)
refit_top_kernel(const Continuation* __restrict__ queue, const uint32_t* __restrict__ queue_count,
                 uint32_t capacity, const uint32_t* __restrict__ parent,
                 uint32_t* __restrict__ flags, Node* __restrict__ nodes,
                 BvhHeader* __restrict__ header) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t count = min(*queue_count, capacity);
  if (i >= count) return;
  const float4 a = reinterpret_cast<const float4*>(queue + i)[0];
  const float4 b = reinterpret_cast<const float4*>(queue + i)[1];
  Box bx{a.x, a.y, a.z, b.x, b.y, b.z};
  climb_global(bx, __float_as_int(a.w), __float_as_uint(b.w), parent, flags, nodes, header);
}

// moved vertices, unchanged topology: prim comes from stris[j].v0p.w
__global__ void __launch_bounds__(kRefitLeaves)
refit_kernel(const float4* __restrict__ tris, uint32_t n, const uint32_t* __restrict__ parent,
             uint32_t* __restrict__ flags, Node* __restrict__ nodes, SortedTri* __restrict__ stris,
             BvhHeader* __restrict__ header, Continuation* __restrict__ queue,
             uint32_t* __restrict__ queue_count, uint32_t capacity) {
  extern __shared__ __align__(16) unsigned char refit_smem_raw[];
  RefitSmem& sm = *reinterpret_cast<RefitSmem*>(refit_smem_raw);
  const uint32_t base = blockIdx.x * kRefitLeaves;
  const uint32_t j = base + threadIdx.x;
  sm.flag[threadIdx.x] = 0;
  __syncthreads();
  if (j >= n) return;
  const uint32_t prim = __float_as_uint(stris[j].v0p.w);
  const float4 a = __ldg(tris + 3 * (uint64_t)prim), b = __ldg(tris + 3 * (uint64_t)prim + 1),
               c = __ldg(tris + 3 * (uint64_t)prim + 2);
  float4* st = reinterpret_cast<float4*>(stris + j);
  st[0] = make_float4(a.x, a.y, a.z, __uint_as_float(prim));
  st[1] = make_float4(b.x, b.y, b.z, 0.0f);
  st[2] = make_float4(c.x, c.y, c.z, 0.0f);
  Box bx;
  bx.lx = omin(omin(a.x, b.x), c.x); bx.ly = omin(omin(a.y, b.y), c.y); bx.lz = omin(omin(a.z, b.z), c.z);
  bx.hx = omax(omax(a.x, b.x), c.x); bx.hy = omax(omax(a.y, b.y), c.y); bx.hz = omax(omax(a.z, b.z), c.z);

  int child_enc = ~(int)j;
  uint32_t pword = parent[(n - 1) + j];
  while (pword != kNoParent && (pword & 2u)) {
    // both children were refitted by this CTA when the tree was built: meet in shared memory
    const uint32_t p = pword >> 2, slot = pword & 1u;
    const uint32_t next_pword = parent[p];  // independent of who wins below: overlap its latency
    float4* np = reinterpret_cast<float4*>(nodes + p);
    const float keep = np[2 * slot + 1].w;  // far end of the range: topology, kept
    const uint32_t k = p - base;
    sm.dep_lo[slot][k] = make_float4(bx.lx, bx.ly, bx.lz, 0.0f);
    sm.dep_hi[slot][k] = make_float4(bx.hx, bx.hy, bx.hz, 0.0f);
    const uint32_t seen = arrive_add_cta(&sm.flag[k]);
    np[2 * slot] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(child_enc));
    np[2 * slot + 1] = make_float4(bx.hx, bx.hy, bx.hz, keep);
    if (seen == 0u) return;  // first arriver stops
    const float4 slo = lds_volatile_v4(&sm.dep_lo[slot ^ 1u][k]);
    const float4 shi = lds_volatile_v4(&sm.dep_hi[slot ^ 1u][k]);
    bx.lx = omin(bx.lx, slo.x); bx.ly = omin(bx.ly, slo.y); bx.lz = omin(bx.lz, slo.z);
    bx.hx = omax(bx.hx, shi.x); bx.hy = omax(bx.hy, shi.y); bx.hz = omax(bx.hz, shi.z);
    child_enc = (int)p;
    pword = next_pword;
  }
  // the next parent (if any) has children in different CTAs: park the walker for
  // refit_top_kernel, so this CTA can retire
  const uint32_t q = atomicAdd(queue_count, 1u);
  if (q < capacity) {
    float4* dst = reinterpret_cast<float4*>(queue + q);
    dst[0] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(child_enc));
    dst[1] = make_float4(bx.hx, bx.hy, bx.hz, __uint_as_float(pword));
  } else {
    climb_global(bx, child_enc, pword, parent, flags, nodes, header);  // queue full: climb inline
  }
}

// parent words from the finished nodes (refit_bvh and the parity accessors need
// them; the build does not).  The far ends stored next to the child boxes give
// the node's leaf range [L, R]; the two children met in shared memory exactly
// when that range lies inside one group of kRefitLeaves leaves.
__global__ void __launch_bounds__(256)
derive_parents_kernel(const Node* __restrict__ nodes, uint32_t n, int root,
                      uint32_t* __restrict__ parent) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g == 0) parent[node_slot_of(root, n)] = kNoParent;
  if (g + 1u >= n) return;
  const float4* np = reinterpret_cast<const float4*>(nodes + g);
  const float4 lo0 = np[0], hi0 = np[1], lo1 = np[2], hi1 = np[3];
  const uint32_t L = __float_as_uint(hi0.w), R = __float_as_uint(hi1.w);
  const uint32_t local = ((L >> kRefitLeavesLog2) == (R >> kRefitLeavesLog2)) ? 2u : 0u;
  parent[node_slot_of(__float_as_int(lo0.w), n)] = (g << 2) | local;
  parent[node_slot_of(__float_as_int(lo1.w), n)] = (g << 2) | local | 1u;
}

// ---- validation of a BVH that did not come from build_bvh (bvh_load, bvh_commit) --------
// Everything traversal, refit_bvh and the accessors index with is checked, so a truncated
// or corrupt file fails with RT_ERR_INVALID instead of faulting the GPU: child ids and leaf
// positions in range, every node's leaf range [L, R] around its split position, each child's
// own range equal to its half of the parent's (ranges nest strictly, so the nodes under the
// root form a finite tree), the root covering [0, n-1], primitive indices below n.
__global__ void __launch_bounds__(256)
validate_nodes_kernel(const Node* __restrict__ nodes, const SortedTri* __restrict__ stris, uint32_t n,
                      int root, uint32_t* __restrict__ bad) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = true;
  if (g < n) ok = __float_as_uint(stris[g].v0p.w) < n;
  if (g == 0) {
    if (n == 1) {
      ok = ok && root == ~0;
    } else {
      ok = ok && root >= 0 && (uint32_t)root < n - 1u;
      if (ok) ok = __float_as_uint(nodes[root].hi0.w) == 0u && __float_as_uint(nodes[root].hi1.w) == n - 1u;
    }
  }
  if (g + 1u < n) {
    const float4* np = reinterpret_cast<const float4*>(nodes + g);
    const float4 lo0 = np[0], hi0 = np[1], lo1 = np[2], hi1 = np[3];
    const uint32_t L = __float_as_uint(hi0.w), R = __float_as_uint(hi1.w);
    const int c0 = __float_as_int(lo0.w), c1 = __float_as_int(lo1.w);
    ok = ok && L <= g && g < R && R < n;
    if (ok) {
      if (c0 < 0) {
        ok = (uint32_t)(~c0) == g && L == g;
      } else {
        ok = (uint32_t)c0 >= L && (uint32_t)c0 < g;
        if (ok) ok = __float_as_uint(nodes[c0].hi0.w) == L && __float_as_uint(nodes[c0].hi1.w) == g;
      }
    }
    if (ok) {
      if (c1 < 0) {
        ok = (uint32_t)(~c1) == g + 1u && R == g + 1u;
      } else {
        ok = (uint32_t)c1 > g && (uint32_t)c1 < R;
        if (ok) ok = __float_as_uint(nodes[c1].hi0.w) == g + 1u && __float_as_uint(nodes[c1].hi1.w) == R;
      }
    }
  }
  if (!ok) *bad = 1u;
}
// depth: the traversal stacks hold kMaxDepth entries; a tree the build made is at most 95
// deep (63 key bits + 32 tie-break bits), a file may claim anything
__global__ void __launch_bounds__(256)
validate_depth_kernel(const uint32_t* __restrict__ parent, uint32_t n, uint32_t* __restrict__ bad) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  uint32_t pw = parent[(n - 1u) + j];
  int depth = 0;
  while (pw != kNoParent && depth <= kMaxDepth) {
    pw = parent[pw >> 2];
    ++depth;
  }
  if (pw != kNoParent) *bad = 1u;
}

// canonical arrays from the traversal layout (parity accessors T3/T4).  Nodes are
// stored by split position g; the Karras index of the child in slot s of node g
// is g + s, the root is Karras node 0, leaves keep n-1+j.
__device__ __forceinline__ uint32_t karras_id(uint32_t x, uint32_t pw, uint32_t leaf0) {
  if (x >= leaf0) return x;
  return (pw == kNoParent) ? 0u : (pw >> 2) + (pw & 1u);
}

__global__ void __launch_bounds__(256)
export_kernel(const Node* __restrict__ nodes, const uint32_t* __restrict__ parent, uint32_t n,
              uint32_t* __restrict__ left, uint32_t* __restrict__ right,
              uint32_t* __restrict__ parent_out, float* __restrict__ boxes,
              const BvhHeader* __restrict__ header) {
  const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;  // storage slot: split or n-1+leaf
  const uint32_t n_nodes = 2 * n - 1, leaf0 = n - 1;
  if (x >= n_nodes) return;
  const uint32_t pw = parent[x];
  const uint32_t k = karras_id(x, pw, leaf0);
  if (parent_out)
    parent_out[k] = (pw == kNoParent) ? kNoParent : karras_id(pw >> 2, parent[pw >> 2], leaf0);
  if (boxes) {
    float* o = boxes + 6 * (uint64_t)k;
    if (pw == kNoParent) {
      for (int c = 0; c < 3; ++c) {
        o[c] = header->root_min[c];
        o[3 + c] = header->root_max[c];
      }
    } else {
      const float4* np = reinterpret_cast<const float4*>(nodes + (pw >> 2)) + 2 * (pw & 1u);
      float4 lo = np[0], hi = np[1];
      o[0] = lo.x; o[1] = lo.y; o[2] = lo.z; o[3] = hi.x; o[4] = hi.y; o[5] = hi.z;
    }
  }
  if (x < leaf0) {
    const int c0 = __float_as_int(nodes[x].lo0.w), c1 = __float_as_int(nodes[x].lo1.w);
    if (left) left[k] = (c0 < 0) ? (leaf0 + (uint32_t)(~c0)) : x;
    if (right) right[k] = (c1 < 0) ? (leaf0 + (uint32_t)(~c1)) : x + 1u;
  }
}

__global__ void unpack_sorted_kernel(const SortedTri* __restrict__ stris, uint32_t n,
                                     uint32_t* __restrict__ idx) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) idx[j] = __float_as_uint(stris[j].v0p.w);
}
__global__ void widen_keys_kernel(const uint32_t* __restrict__ in, uint32_t n,
                                  uint64_t* __restrict__ out) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = in[j];
}

}  // namespace

int launch_unpack_sorted(const SortedTri* stris, uint64_t n, uint32_t* idx, cudaStream_t s) {
  if (n == 0) return 0;
  unpack_sorted_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(stris, (uint32_t)n, idx);
  RT_LAUNCHED();
  return 0;
}
int launch_widen_keys(const uint32_t* in, uint64_t n, uint64_t* out, cudaStream_t s) {
  if (n == 0) return 0;
  widen_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, (uint32_t)n, out);
  RT_LAUNCHED();
  return 0;
}

static void queue_layout(void* queue_buf, uint32_t** count, Continuation** queue) {
  *count = reinterpret_cast<uint32_t*>(queue_buf);
  *queue = reinterpret_cast<Continuation*>(reinterpret_cast<char*>(queue_buf) + 32);
}

size_t refit_queue_bytes(uint64_t n) {
  return 32 + (size_t)queue_capacity(n) * sizeof(Continuation);
}

template <typename KeyT, int THREADS, int LPT>
static int launch_walk(const float4* tris, const uint32_t* sorted_idx, const KeyT* keys, uint64_t n,
                       uint32_t* flags, Node* nodes, SortedTri* stris, BvhHeader* header,
                       Continuation* queue, uint32_t* queue_count, uint32_t capacity, cudaStream_t s,
                       const uint32_t* abort_flag) {
  constexpr int G = THREADS * LPT;
  const unsigned blocks = (unsigned)((n + G - 1) / G);
  const size_t smem = sizeof(BuildSmem<G>);
  auto kern = build_walk_kernel<KeyT, THREADS, LPT>;
  RT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  RT_CUDA(launch_pdl(kern, blocks, THREADS, smem, s, tris, sorted_idx, keys, (uint32_t)n, flags, nodes, stris,
                     header, queue, queue_count, capacity, abort_flag));
  RT_LAUNCHED();
  return 0;
}

std::atomic<int> g_build_walk_shape{0};  // rt_set_param("build.walk"): THREADS * 16 + LPT; 0 = default

template <typename KeyT>
static int build_tree_impl(const float4* tris, const uint32_t* sorted_idx, const KeyT* keys,
                           uint64_t n, uint32_t* flags, Node* nodes, SortedTri* stris,
                           BvhHeader* header, void* queue_buf, cudaStream_t s,
                           const uint32_t* abort_flag) {
  const uint32_t capacity = queue_capacity(n);
  uint32_t* queue_count;
  Continuation* queue;
  queue_layout(queue_buf, &queue_count, &queue);
#define RT_WALK(T, L) launch_walk<KeyT, T, L>(tris, sorted_idx, keys, n, flags, nodes, stris, header, queue, \
                                               queue_count, capacity, s, abort_flag)
  switch (g_build_walk_shape.load(std::memory_order_relaxed)) {
    case 128 * 16 + 1: RT_TRY(RT_WALK(128, 1)); break;
    case 512 * 16 + 1: RT_TRY(RT_WALK(512, 1)); break;
    case 256 * 16 + 2: RT_TRY(RT_WALK(256, 2)); break;
    case 256 * 16 + 4: RT_TRY(RT_WALK(256, 4)); break;
    case 128 * 16 + 2: RT_TRY(RT_WALK(128, 2)); break;
    case 128 * 16 + 4: RT_TRY(RT_WALK(128, 4)); break;
    case 64 * 16 + 4: RT_TRY(RT_WALK(64, 4)); break;
    default: RT_TRY(RT_WALK(256, 1)); break;
  }
#undef RT_WALK
  RT_CUDA(launch_pdl(build_top_kernel<KeyT>, (capacity + 255) / 256, 256, 0, s,
                     (const Continuation*)queue, (const uint32_t*)queue_count, capacity, keys,
                     (uint32_t)n, flags, nodes, header, abort_flag));
  RT_LAUNCHED();
  return 0;
}

// Zeroes the meeting counters and the queue head.  A separate call so that build_bvh can
// issue it before the first kernel: the kernels of the build then follow each other with
// nothing in between, which is what programmatic dependent launch needs (common.cuh).
int prepare_build_tree(uint64_t n, uint32_t* flags, void* queue_buf, cudaStream_t s) {
  if (n == 0) return 0;
  uint32_t* queue_count;
  Continuation* queue;
  queue_layout(queue_buf, &queue_count, &queue);
  if (n > 1) RT_CUDA(cudaMemsetAsync(flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
  RT_CUDA(cudaMemsetAsync(queue_count, 0, 32, s));
  return 0;
}

int launch_build_tree(const float4* tris, const uint32_t* sorted_idx, const void* keys,
                      uint32_t key_bits, uint64_t n, uint32_t* flags, Node* nodes,
                      SortedTri* stris, BvhHeader* header, void* queue_buf, cudaStream_t s,
                      const uint32_t* abort_flag) {
  if (n == 0) return 0;
  if (key_bits == 30)
    return build_tree_impl<uint32_t>(tris, sorted_idx, (const uint32_t*)keys, n, flags, nodes,
                                     stris, header, queue_buf, s, abort_flag);
  return build_tree_impl<uint64_t>(tris, sorted_idx, (const uint64_t*)keys, n, flags, nodes, stris,
                                   header, queue_buf, s, abort_flag);
}

int launch_derive_parents(const Node* nodes, uint64_t n, int root, uint32_t* parent,
                          cudaStream_t s) {
  if (n == 0) return 0;
  derive_parents_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(nodes, (uint32_t)n, root, parent);
  RT_LAUNCHED();
  return 0;
}

int launch_validate(const Node* nodes, const SortedTri* stris, uint64_t n, int root, uint32_t* bad,
                    cudaStream_t s) {
  RT_CUDA(cudaMemsetAsync(bad, 0, sizeof(uint32_t), s));
  if (n == 0) return 0;
  validate_nodes_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(nodes, stris, (uint32_t)n, root, bad);
  RT_LAUNCHED();
  return 0;
}
int launch_validate_depth(const uint32_t* parent, uint64_t n, uint32_t* bad, cudaStream_t s) {
  if (n == 0) return 0;
  validate_depth_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(parent, (uint32_t)n, bad);
  RT_LAUNCHED();
  return 0;
}

int launch_refit_update(const float4* tris, uint64_t n, const uint32_t* parent, uint32_t* flags,
                        Node* nodes, SortedTri* stris, BvhHeader* header, void* queue_buf,
                        cudaStream_t s) {
  if (n == 0) return 0;
  const unsigned blocks = (unsigned)((n + kRefitLeaves - 1) / kRefitLeaves);
  const uint32_t capacity = queue_capacity(n);
  uint32_t* queue_count;
  Continuation* queue;
  queue_layout(queue_buf, &queue_count, &queue);
  if (n > 1) RT_CUDA(cudaMemsetAsync(flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
  RT_CUDA(cudaMemsetAsync(queue_count, 0, 32, s));
  const size_t smem = sizeof(RefitSmem);
  RT_CUDA(cudaFuncSetAttribute(refit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  refit_kernel<<<blocks, kRefitLeaves, smem, s>>>(tris, (uint32_t)n, parent, flags, nodes, stris,
                                                 header, queue, queue_count, capacity);
  RT_LAUNCHED();
  refit_top_kernel<<<(capacity + 255) / 256, 256, 0, s>>>(queue, queue_count, capacity, parent, flags,
                                                          nodes, header);
  RT_LAUNCHED();
  return 0;
}

int launch_export_topology(const Node* nodes, const uint32_t* parent, uint64_t n, uint32_t* left,
                           uint32_t* right, uint32_t* parent_out, float* boxes,
                           const BvhHeader* header, cudaStream_t s) {
  if (n == 0) return 0;
  const uint64_t n_nodes = 2 * n - 1;
  export_kernel<<<(unsigned)((n_nodes + 255) / 256), 256, 0, s>>>(
      nodes, parent, (uint32_t)n, left, right, parent_out, boxes, header);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
