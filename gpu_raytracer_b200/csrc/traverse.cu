// traverse.cu — A6/A7: persistent-thread while-while stack traversal with
// Moeller-Trumbore intersection (Aila & Laine 2009).
//
// Launch: a fixed grid of resident warps (SM count x CTAs per SM); each warp
// pulls 32 consecutive rays at a time with one atomicAdd by lane 0 (warp-level
// ray fetch), so the grid never depends on the batch size and long rays do not
// hold a whole CTA hostage.
//
// Per ray: the inner loop walks internal nodes — one 64-byte node fetch (2 x
// LDG.256, two sectors) carries both children's boxes and ids; near child
// first, far child pushed on a per-thread stack in local memory — until it
// reaches a leaf; the second loop intersects leaves (one triangle each, stored
// as v0|prim, v1, v2 = 48 B) and pops.  Box tests cull with tnear <= tfar*pad
// (<=, so equal-t candidates in other subtrees are still visited) and the
// acceptance rule breaks equal-t ties towards the smaller primitive index, so
// the result does not depend on traversal order.
//
// Bound: L2/HBM latency and divergence, not streaming bandwidth; algorithmic
// bytes/ray = 48 + 64 x nodes fetched + 48 x triangles tested.
#include "trace_common.cuh"

namespace rt {
namespace {

template <bool ANY, bool STATS, bool WT>
__device__ __forceinline__ void trace_one(const Node* __restrict__ nodes,
                                          const SortedTri* __restrict__ stris, int root,
                                          RayState& r, uint32_t& n_nodes, uint32_t& n_tests) {
  WtRay wr;
  if (WT) wr = wt_setup(r);
  int stack[kStack];
  int sp = 0;
  int node = root;
  while (node != kEmpty) {
    // ---- internal nodes until a leaf (or nothing) is at hand ----------------
    while (node >= 0) {
      if (STATS) ++n_nodes;
      const F8 c0h = ldg256(nodes + node), c1h = ldg256(&nodes[node].lo1);
      const float4 lo0 = c0h.a, hi0 = c0h.b, lo1 = c1h.a, hi1 = c1h.b;
      float tn0, tn1;
      const bool h0 = slab(lo0, hi0, r, tn0);
      const bool h1 = slab(lo1, hi1, r, tn1);
      const int c0 = __float_as_int(lo0.w), c1 = __float_as_int(lo1.w);
      if (h0 && h1) {
        const bool swap = tn1 < tn0;
        node = swap ? c1 : c0;
        stack[sp++] = swap ? c0 : c1;
      } else if (h0) {
        node = c0;
      } else if (h1) {
        node = c1;
      } else {
        node = (sp > 0) ? stack[--sp] : kEmpty;
      }
    }
    // ---- leaves ----------------------------------------------------------------
    while (node < 0 && node != kEmpty) {
      if (STATS) ++n_tests;
      const bool hit = WT ? intersect_wt(stris, (uint32_t)(~node), r, wr)
                          : intersect(stris, (uint32_t)(~node), r);
      if (ANY && hit) return;
      node = (sp > 0) ? stack[--sp] : kEmpty;
    }
  }
}

// STATS: also counts node fetches and triangle tests into stats[0..1] (the
// algorithmic-bytes figure of the roofline; never used on the timed path)
template <bool ANY, bool STATS, bool WT>
__global__ void __launch_bounds__(kTraceThreads)
trace_kernel(const Node* __restrict__ nodes, const SortedTri* __restrict__ stris, uint32_t n_tris,
             int root, const float4* __restrict__ rays, unsigned long long n_rays,
             float4* __restrict__ hits, unsigned long long* __restrict__ ray_counter,
             unsigned long long* __restrict__ stats) {
  uint32_t n_nodes = 0, n_tests = 0;
  const int lane = threadIdx.x & 31;
  const float inf = __int_as_float(0x7f800000);
  while (true) {
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(ray_counter, 32ull);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n_rays) break;
    const unsigned long long i = base + lane;
    if (i < n_rays) {
      const float4 ro = ld_stream(rays + 2 * i), rd = ld_stream(rays + 2 * i + 1);
      RayState r;
      load_ray(r, ro, rd);
      if (n_tris != 0) trace_one<ANY, STATS, WT>(nodes, stris, root, r, n_nodes, n_tests);
      float4 h;
      if (r.best_prim < 0) h = make_float4(inf, 0.0f, 0.0f, __int_as_float(-1));
      else if (ANY) h = make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0x7FFFFFFF));
      else h = make_float4(r.best_t, r.bu, r.bv, __int_as_float(r.best_prim));
      st_stream(hits + i, h);
    }
    __syncwarp();
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

// ---------------------------------------------------------------------------
// Variant 1 (default): the same while-while traversal with two changes that
// attack SIMD utilisation, which ncu shows is the limiter (6.4 of 32 lanes
// active per instruction for variant 0 on incoherent rays, DRAM < 1 %):
//   * per-lane dynamic ray fetch: a lane whose ray has terminated takes the
//     next ray of the warp's private chunk (kChunk consecutive rays, claimed
//     with one atomicAdd per chunk) as soon as `refill_min` lanes are idle, so a
//     warp never drains down to its slowest ray;
//   * the inner node loop is left, and the leaf loop entered, by warp vote: only
//     when `leaf_min` lanes hold an untested leaf (or nobody has a node left),
//     so triangle code runs for several lanes at once and node steps are not
//     held up by the first lane that reaches a leaf.
// Lanes keep their own stack and ray state; the arithmetic (slab, intersect,
// acceptance) is the shared code above, so results are bit-identical to
// variant 0 and to the oracle whatever the scheduling.
// ---------------------------------------------------------------------------
//
// PIPE (host-resident batches, api.cu trace_streamed): the same kernel runs while the
// copy engine is still delivering the rays.  A warp that has claimed rays beyond the
// upload watermark leaves its idle lanes idle and keeps traversing (or naps when it has
// nothing else), and every retired ray is counted into its segment; the thread that
// completes a segment raises that segment's flag in pinned host memory, which is the
// host's cue to start the device-to-host copy of those hits.  One launch, one tail,
// instead of one launch (and one ~0.25 ms tail of slowest rays) per staged chunk.
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 ld_cg_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

#ifndef RT_TRACE_PARK
#define RT_TRACE_PARK 1
#endif
#ifndef RT_TRACE_MINB
#define RT_TRACE_MINB 5  // resident CTAs per SM the register allocation is held to
#endif
// Q: the nodes are the 32-byte NodeQ records of the compact layout (qnodes, grid); one request
// per node step, the slab test on 16-bit planes (trace_common.cuh slab_q).
template <bool ANY, bool STATS, bool WT, bool PIPE, int SHS, bool Q = false>
__global__ void __launch_bounds__(kTraceThreads, (PIPE || WT || STATS) ? 1 : RT_TRACE_MINB)
trace_kernel_dyn(const Node* __restrict__ nodes, const SortedTri* __restrict__ stris,
                 uint32_t n_tris, int root, const float4* __restrict__ rays,
                 unsigned long long n_rays, float4* __restrict__ hits,
                 unsigned long long* __restrict__ ray_counter,
                 unsigned long long* __restrict__ stats, int chunk, int refill_min, int leaf_min,
                 TracePipe pipe, const uint32_t* __restrict__ perm,
                 const NodeQ* __restrict__ qnodes = nullptr, QGrid grid = QGrid()) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const float inf = __int_as_float(0x7f800000);
  uint32_t n_nodes = 0, n_tests = 0;

  // the first SHS stack levels live in shared memory ([level][thread]: bank = lane, one
  // wavefront whatever the lanes' depths), the rest in local memory
  __shared__ int sh_stack[SHS > 0 ? SHS : 1][kTraceThreads];
  // Q: 1/d of the thread's ray, for the acceptance rule's box test (the node loop does not use it)
  __shared__ float sh_inv[Q ? 3 : 1][kTraceThreads];
  // Compact layout: what only the leaf test and the retire need (direction, barycentrics, the
  // ray's index) waits in shared memory too; the per-ray plane constants take its registers
  // (62 -> 24 bytes of spills at 48 registers; measured -2 % on frames, -5..-9 % on incoherent
  // and reordered rays; the same parking costs the plain 64-byte kernel 1-5 %, which has no spills
  // to save, and gains its RT_TRACE_REORDER variant (stack top in shared memory, 16 bytes of spills) 2.7 %)
  constexpr bool PARK = RT_TRACE_PARK && (Q || SHS > 0) && !PIPE && !WT && !STATS;
  __shared__ float sh_park[PARK ? 5 : 1][kTraceThreads];
  __shared__ unsigned long long sh_idx[PARK ? kTraceThreads : 1];
  int stack[kStack - SHS];
  int sp = 0;
  auto push = [&](int v) {
    if (SHS > 0 && sp < SHS) sh_stack[sp][threadIdx.x] = v;
    else stack[sp - SHS] = v;
    ++sp;
  };
  auto pop = [&]() -> int {
    --sp;
    return (SHS > 0 && sp < SHS) ? sh_stack[sp][threadIdx.x] : stack[sp - SHS];
  };
  int node = kEmpty;
  bool has_ray = false;
  unsigned long long my_ray = 0;
  RayState r;
  r.best_prim = -1;
  WtRay wr = {0, 1, 2, 0.0f, 0.0f, 0.0f};
  QRay qr = {};
  unsigned long long chunk_next = 0, chunk_end = 0;  // warp-uniform
  bool exhausted = false;                            // warp-uniform
  unsigned long long uploaded = 0;                   // warp-uniform: last watermark seen (PIPE)
  unsigned long long next_poll = 0;                  // warp-uniform
  uint32_t polls = 0;
  // retired rays of the two segments the warp's live chunks belong to (warp-uniform); they
  // reach the segment counters once per chunk, not once per ray: 2M returning atomics on
  // one hot word would queue at its L2 slice and double the kernel's run time
  unsigned long long seg_a = ~0ull, seg_b = ~0ull;
  uint32_t cnt_a = 0, cnt_b = 0;    // retired, not yet added to the segment's counter
  uint32_t live_a = 0, live_b = 0;  // handed to lanes, not yet retired
  auto flush_segment = [&](unsigned long long seg, uint32_t cnt) {  // warp-uniform call
    if (cnt == 0u) return;
    asm volatile("fence.acq_rel.gpu;" ::: "memory");  // every lane: its hit stores first
    __syncwarp();
    if (lane == 0) {
      const unsigned long long seg_first = seg << pipe.seg_shift;
      const unsigned long long seg_rays = min(1ull << pipe.seg_shift, n_rays - seg_first);
      if (atomicAdd(pipe.seg_done + seg, cnt) + cnt == (uint32_t)seg_rays) {
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(pipe.host_flags + seg) = 1u;
      }
    }
  };
  constexpr unsigned long long kPollNs = 8000;
  constexpr uint32_t kPollLimit = 250000;  // x kPollNs = 2 s without progress

  while (true) {
    // ---- per-lane refill from the warp's chunk --------------------------------
    const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !has_ray);
    if (idle != 0u && !exhausted && (__popc(idle) >= refill_min || idle == 0xFFFFFFFFu)) {
      bool may_claim = chunk_next >= chunk_end;
      if (PIPE && may_claim && uploaded < n_rays) {
        // claim only rays that have arrived: a warp that holds a chunk of the future sits
        // out the next arrival, and the rays that did arrive are left to fewer warps
        const unsigned long long now = global_timer_ns();
        may_claim = false;
        if (now >= next_poll) {
          unsigned long long claimed = 0;
          const unsigned long long before = uploaded;
          if (lane == 0) {
            uploaded = ld_volatile_u64(pipe.watermark);
            claimed = ld_volatile_u64(ray_counter);
          }
          uploaded = __shfl_sync(0xFFFFFFFFu, uploaded, 0);
          claimed = __shfl_sync(0xFFFFFFFFu, claimed, 0);
          if (uploaded != before) polls = 0u;
          may_claim = uploaded >= n_rays || claimed + (unsigned long long)chunk <= uploaded;
          if (!may_claim) {
            next_poll = now + kPollNs;
            if (++polls > kPollLimit) {  // 2 s without progress: give up loudly
              if (lane == 0) atomicExch(pipe.error, 1u);
              exhausted = true;
            }
          }
        }
        if (!may_claim && idle == 0xFFFFFFFFu) __nanosleep(kPollNs / 2);  // nothing else to do
      }
      if (may_claim && !exhausted) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(ray_counter, (unsigned long long)chunk);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n_rays) {
          exhausted = true;
          chunk_next = chunk_end = 0;
        } else {
          chunk_next = base;
          chunk_end = min(base + (unsigned long long)chunk, n_rays);
          if (PIPE && (base >> pipe.seg_shift) != seg_b) {  // the oldest segment's count goes out
            flush_segment(seg_a, cnt_a);
            // rays of the old seg_a still in flight count themselves (retire path)
            seg_a = seg_b;
            cnt_a = cnt_b;
            live_a = live_b;
            seg_b = base >> pipe.seg_shift;
            cnt_b = live_b = 0u;
          }
        }
      }
      bool ready = true;
      if (PIPE && !exhausted && uploaded < chunk_end) {  // have the claimed rays arrived?
        // look at the watermark at most once per kPollNs and warp: thousands of warps
        // re-reading one word every loop trip would swamp its L2 slice
        ready = false;
        const unsigned long long now = global_timer_ns();
        if (now >= next_poll) {
          const unsigned long long before = uploaded;
          if (lane == 0) uploaded = ld_volatile_u64(pipe.watermark);
          uploaded = __shfl_sync(0xFFFFFFFFu, uploaded, 0);
          if (uploaded != before) polls = 0u;
          ready = uploaded >= chunk_end;
          next_poll = now + kPollNs;
          if (!ready && ++polls > kPollLimit) {  // 2 s without progress: give up loudly
            if (lane == 0) atomicExch(pipe.error, 1u);
            exhausted = true;
            chunk_next = chunk_end = 0;
          }
        }
        if (!ready && idle == 0xFFFFFFFFu) __nanosleep(kPollNs / 2);  // nothing else to do
      }
      const unsigned long long avail = ready ? chunk_end - chunk_next : 0ull;
      const uint32_t slot = __popc(idle & lt_mask);
      if (!has_ray && slot < avail) {
        my_ray = chunk_next + slot;
        if (!PIPE && perm) my_ray = perm[my_ray];  // RT_TRACE_REORDER: position in key order -> ray
        const float4 ro = PIPE ? ld_cg_f4(rays + 2 * my_ray) : ld_stream(rays + 2 * my_ray);
        const float4 rd = PIPE ? ld_cg_f4(rays + 2 * my_ray + 1) : ld_stream(rays + 2 * my_ray + 1);
        load_ray(r, ro, rd);
        if (WT) wr = wt_setup(r);
        if (Q) {
          qr = q_setup(r, grid);
          sh_inv[0][threadIdx.x] = r.ix;
          sh_inv[Q ? 1 : 0][threadIdx.x] = r.iy;
          sh_inv[Q ? 2 : 0][threadIdx.x] = r.iz;
        }
        if (PARK) {
          sh_park[0][threadIdx.x] = r.dx;
          sh_park[PARK ? 1 : 0][threadIdx.x] = r.dy;
          sh_park[PARK ? 2 : 0][threadIdx.x] = r.dz;
          sh_idx[threadIdx.x] = my_ray;
        }
        has_ray = true;
        sp = 0;
        node = (n_tris != 0) ? root : kEmpty;
      }
      const unsigned long long need = (unsigned long long)__popc(idle);
      const unsigned long long took = (need < avail) ? need : avail;
      chunk_next += took;
      if (PIPE) live_b += (uint32_t)took;
    }
    const uint32_t live = __ballot_sync(0xFFFFFFFFu, has_ray);  // does not change in the node loop
    if (live == 0u) {
      if (exhausted) break;
      continue;
    }

    // ---- internal nodes, until enough lanes wait at a leaf ----------------------
    // one vote per trip: the lanes with a ray and no inner node are the ones that wait (leaf or finished)
    const int inner_min = __popc(live) - leaf_min;  // leave when fewer lanes than this still walk
    while (true) {
      const uint32_t inner = __ballot_sync(0xFFFFFFFFu, has_ray && node >= 0);
      if (inner == 0u || __popc(inner) <= inner_min) break;
      if (has_ray && node >= 0) {
        if (STATS) ++n_nodes;
        float tn0, tn1;
        bool h0, h1;
        int c0, c1;
        if (Q) {
          const F8 w = ldg256(qnodes + node);
          h0 = slab_q(__float_as_uint(w.a.x), __float_as_uint(w.a.y), __float_as_uint(w.a.z), qr, r, tn0);
          h1 = slab_q(__float_as_uint(w.a.w), __float_as_uint(w.b.x), __float_as_uint(w.b.y), qr, r, tn1);
          c0 = __float_as_int(w.b.z);
          c1 = __float_as_int(w.b.w);
        } else {
          const F8 c0h = ldg256(nodes + node), c1h = ldg256(&nodes[node].lo1);
          const float4 lo0 = c0h.a, hi0 = c0h.b, lo1 = c1h.a, hi1 = c1h.b;
          h0 = slab(lo0, hi0, r, tn0);
          h1 = slab(lo1, hi1, r, tn1);
          c0 = __float_as_int(lo0.w);
          c1 = __float_as_int(lo1.w);
        }
        if (h0 && h1) {
          const bool swap = tn1 < tn0;
          node = swap ? c1 : c0;
          push(swap ? c0 : c1);
        } else if (h0) {
          node = c0;
        } else if (h1) {
          node = c1;
        } else {
          node = (sp > 0) ? pop() : kEmpty;
        }
      }
    }

    // ---- one leaf per waiting lane ---------------------------------------------
    if (has_ray && node < 0 && node != kEmpty) {
      if (STATS) ++n_tests;
      if (PARK) {
        r.dx = sh_park[0][threadIdx.x];
        r.dy = sh_park[PARK ? 1 : 0][threadIdx.x];
        r.dz = sh_park[PARK ? 2 : 0][threadIdx.x];
      }
      const bool hit = WT ? intersect_wt(stris, (uint32_t)(~node), r, wr)
                          : intersect<Q>(stris, (uint32_t)(~node), r, &sh_inv[0][threadIdx.x]);
      if (PARK && hit && !ANY) {
        sh_park[PARK ? 3 : 0][threadIdx.x] = r.bu;
        sh_park[PARK ? 4 : 0][threadIdx.x] = r.bv;
      }
      if (ANY && hit) {
        node = kEmpty;
        sp = 0;
      } else {
        node = (sp > 0) ? pop() : kEmpty;
      }
    }

    // ---- retire finished rays ------------------------------------------------------
    const bool retiring = has_ray && node == kEmpty;
    if (retiring) {
      float4 h;
      if (r.best_prim < 0) h = make_float4(inf, 0.0f, 0.0f, __int_as_float(-1));
      else if (ANY) h = make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0x7FFFFFFF));
      else if (PARK) h = make_float4(r.best_t, sh_park[PARK ? 3 : 0][threadIdx.x], sh_park[PARK ? 4 : 0][threadIdx.x], __int_as_float(r.best_prim));
      else h = make_float4(r.best_t, r.bu, r.bv, __int_as_float(r.best_prim));
      st_stream(hits + (PARK ? sh_idx[threadIdx.x] : my_ray), h);
      has_ray = false;
    }
    if (PIPE) {  // count this trip's retired rays into their segments
      const bool retired = retiring;
      const unsigned long long sgm = my_ray >> pipe.seg_shift;
      const uint32_t in_b = __ballot_sync(0xFFFFFFFFu, retired && sgm == seg_b);
      const uint32_t in_a = __ballot_sync(0xFFFFFFFFu, retired && sgm == seg_a && sgm != seg_b);
      cnt_b += __popc(in_b);
      live_b -= __popc(in_b);
      cnt_a += __popc(in_a);
      live_a -= __popc(in_a);
      // a segment's count goes out as soon as the warp holds no more of its rays, so the
      // host can start that segment's download while later segments are still traced
      if (cnt_a != 0u && live_a == 0u) {
        flush_segment(seg_a, cnt_a);
        cnt_a = 0u;
      }
      if (cnt_b != 0u && live_b == 0u && chunk_next >= chunk_end) {
        flush_segment(seg_b, cnt_b);
        cnt_b = 0u;
      }
      if (retired && sgm != seg_a && sgm != seg_b) {  // a straggler of an older chunk: rare
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        const unsigned long long seg_first = sgm << pipe.seg_shift;
        const unsigned long long seg_rays = min(1ull << pipe.seg_shift, n_rays - seg_first);
        if (atomicAdd(pipe.seg_done + sgm, 1u) + 1u == (uint32_t)seg_rays) {
          __threadfence_system();
          *reinterpret_cast<volatile uint32_t*>(pipe.host_flags + sgm) = 1u;
        }
      }
    }
  }
  if (PIPE) {
    flush_segment(seg_a, cnt_a);
    flush_segment(seg_b, cnt_b);
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

// RT_TRACE_REORDER: coherence key of a ray = Morton code of its origin in the scene box
// (4 bits per axis) above 8 interleaved bits of its direction on the unit cube (3, 3, 2
// bits): 20 bits, two passes of the radix sort, about one warp of rays per key at 32M rays.  Rays that start in the same cell and point the same way become neighbours in key
// order and share their node fetches in L1.  The key only orders the work: every ray's
// result is what it would be in any other order.
__device__ __forceinline__ uint32_t spread3_5(uint32_t x) {  // 5 bits -> every third bit
  x &= 0x1Fu;
  x = (x | (x << 8)) & 0x0000100Fu;
  x = (x | (x << 4)) & 0x000010C3u;
  x = (x | (x << 2)) & 0x00001249u;  // input bit b now at position 3b
  return x;
}
__global__ void __launch_bounds__(256)
ray_key_kernel(const float4* __restrict__ rays, uint32_t n, float lx, float ly, float lz, float sx,
               float sy, float sz, uint32_t* __restrict__ keys) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 o = ld_stream(rays + 2 * (size_t)i), d = ld_stream(rays + 2 * (size_t)i + 1);
  const uint32_t qx = (uint32_t)fminf(fmaxf((o.x - lx) * sx, 0.0f), 15.0f);
  const uint32_t qy = (uint32_t)fminf(fmaxf((o.y - ly) * sy, 0.0f), 15.0f);
  const uint32_t qz = (uint32_t)fminf(fmaxf((o.z - lz) * sz, 0.0f), 15.0f);
  const float m = fmaxf(fmaxf(fabsf(d.x), fabsf(d.y)), fmaxf(fabsf(d.z), 1e-30f));
  const float k = 4.0f / m;
  const uint32_t dx = (uint32_t)fminf(fmaxf(d.x * k + 4.0f, 0.0f), 7.0f);
  const uint32_t dy = (uint32_t)fminf(fmaxf(d.y * k + 4.0f, 0.0f), 7.0f);
  const uint32_t dz = (uint32_t)fminf(fmaxf(d.z * k + 4.0f, 0.0f), 7.0f) >> 1;
  const uint32_t om = (spread3_5(qx) << 2) | (spread3_5(qy) << 1) | spread3_5(qz);  // 12 bits
  const uint32_t dm = ((dx >> 2) << 7) | ((dy >> 2) << 6) | ((dz >> 1) << 5) | (((dx >> 1) & 1u) << 4) |
                      (((dy >> 1) & 1u) << 3) | ((dz & 1u) << 2) | ((dx & 1u) << 1) | (dy & 1u);
  keys[i] = (om << 8) | dm;
}

}  // namespace

int launch_ray_keys(const float4* rays, uint64_t n, const float* box, uint32_t* keys, cudaStream_t s) {
  if (n == 0) return 0;
  float sc[3];
  for (int a = 0; a < 3; ++a) {
    const float e = box[3 + a] - box[a];
    sc[a] = e > 0.0f ? 16.0f / e : 0.0f;
  }
  ray_key_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(rays, (uint32_t)n, box[0], box[1], box[2],
                                                            sc[0], sc[1], sc[2], keys);
  RT_LAUNCHED();
  return 0;
}

int launch_trace(const Node* nodes, const SortedTri* stris, uint64_t n_tris, int root,
                 const float4* rays,
                 uint64_t n_rays, uint32_t flags, float4* hits, unsigned long long* ray_counter,
                 cudaStream_t s, unsigned long long* stats, const TracePipe* pipe,
                 const uint32_t* perm, const Compact* compact) {
  if (n_rays == 0) return 0;
  if (perm && (pipe || params().trace_variant == 0))
    return fail(-1, "launch_trace: a ray permutation needs the dynamic-fetch kernel, device buffers");
  if (pipe && (stats || params().trace_variant == 0))
    return fail(-1, "launch_trace: the streamed mode needs the dynamic-fetch kernel without stats");
  RT_CUDA(cudaMemsetAsync(ray_counter, 0, sizeof(unsigned long long), s));
  int bps = params().trace_blocks_per_sm;
  if (bps <= 0) bps = 8;
  uint64_t blocks = (uint64_t)num_sms() * bps;
  const uint64_t need = (n_rays + kTraceThreads - 1) / kTraceThreads;
  if (blocks > need) blocks = need;
  const bool any = (flags & 1u) != 0, wt = (flags & 2u) != 0;
  const bool st = stats != nullptr;
  if (params().trace_variant != 0) {
    const Params& p = params();
    int chunk = p.trace_chunk > 0 ? p.trace_chunk : 64;
    // small batches: keep every resident warp busy rather than a few warps long
    if (p.trace_chunk <= 0 && n_rays < blocks * (kTraceThreads / 32) * (uint64_t)chunk * 2) chunk = 32;
    // closest-hit rays live longer than occlusion rays: refilling at 8 idle lanes instead of 16
    // measured +3.6 % on flake10m primary rays and within 1 % elsewhere; any-hit prefers 16
    const int refill_min = p.trace_refill_min > 0 ? p.trace_refill_min : (any ? 16 : 8);
    const int leaf_min = p.trace_leaf_min > 0 ? p.trace_leaf_min : 8;
    if (pipe) {
      // 32-ray chunks: when a piece arrives every resident warp gets a share of it.  The
      // segment accounting needs chunks that never straddle a segment: a power of two
      // not above the segment size (a caller-set "trace.chunk" is rounded down to one)
      if (p.trace_chunk <= 0) chunk = 32;
      int pow2 = 1;
      while (pow2 * 2 <= chunk && pow2 * 2 <= (1 << pipe->seg_shift)) pow2 *= 2;
      chunk = pow2;
#define RT_TRACE_PIPE(A, W)                                                          \
  trace_kernel_dyn<A, false, W, true, 0><<<(unsigned)blocks, kTraceThreads, 0, s>>>( \
      nodes, stris, (uint32_t)n_tris, root, rays, n_rays, hits, ray_counter, stats, chunk, refill_min, leaf_min, *pipe, nullptr)
      if (any) { if (wt) RT_TRACE_PIPE(true, true); else RT_TRACE_PIPE(true, false); }
      else     { if (wt) RT_TRACE_PIPE(false, true); else RT_TRACE_PIPE(false, false); }
#undef RT_TRACE_PIPE
      RT_LAUNCHED();
      return 0;
    }
    const TracePipe no_pipe = {nullptr, nullptr, nullptr, nullptr, 0};
#define RT_TRACE_DYN_K(A, S, W, K)                                                   \
  trace_kernel_dyn<A, S, W, false, K><<<(unsigned)blocks, kTraceThreads, 0, s>>>(    \
      nodes, stris, (uint32_t)n_tris, root, rays, n_rays, hits, ray_counter, stats, chunk, refill_min, leaf_min, no_pipe, perm)
#define RT_TRACE_DYN(A, S, W) RT_TRACE_DYN_K(A, S, W, 0)
    // auto: rays that come through a permutation are incoherent by the caller's word, and for
    // those the shared-memory stack top pays (config 5 with the flag: 4,050 -> 4,145 Mrays/s)
    const int shs = p.trace_sh_stack >= 0 ? p.trace_sh_stack : (perm ? 8 : 0);
    if (compact && compact->valid && !wt && p.trace_compact != 0) {
      // 32-byte nodes (watertight launches stay on the 64-byte layout)
      const NodeQ* qn = compact->nodes;
      const QGrid grid = compact->grid;
#define RT_TRACE_Q(A, S, K)                                                                \
  trace_kernel_dyn<A, S, false, false, K, true><<<(unsigned)blocks, kTraceThreads, 0, s>>>( \
      nodes, stris, (uint32_t)n_tris, root, rays, n_rays, hits, ray_counter, stats, chunk, refill_min, leaf_min, no_pipe, perm, qn, grid)
      if (st) { if (any) RT_TRACE_Q(true, true, 0); else RT_TRACE_Q(false, true, 0); }
      else if (shs > 0) { if (any) RT_TRACE_Q(true, false, 8); else RT_TRACE_Q(false, false, 8); }
      else { if (any) RT_TRACE_Q(true, false, 0); else RT_TRACE_Q(false, false, 0); }
#undef RT_TRACE_Q
    } else if (!st && !wt && shs > 0) {  // experiment: stack top in shared memory
      if (any) { if (shs == 8) RT_TRACE_DYN_K(true, false, false, 8); else if (shs == 16) RT_TRACE_DYN_K(true, false, false, 16); else RT_TRACE_DYN_K(true, false, false, 24); }
      else     { if (shs == 8) RT_TRACE_DYN_K(false, false, false, 8); else if (shs == 16) RT_TRACE_DYN_K(false, false, false, 16); else RT_TRACE_DYN_K(false, false, false, 24); }
    } else {
      RT_DISPATCH3(RT_TRACE_DYN, any, st, wt);
    }
#undef RT_TRACE_DYN
#undef RT_TRACE_DYN_K
    RT_LAUNCHED();
    return 0;
  }
#define RT_TRACE_LAUNCH(A, S, W)                                          \
  trace_kernel<A, S, W><<<(unsigned)blocks, kTraceThreads, 0, s>>>(       \
      nodes, stris, (uint32_t)n_tris, root, rays, n_rays, hits, ray_counter, stats)
  RT_DISPATCH3(RT_TRACE_LAUNCH, any, st, wt);
#undef RT_TRACE_LAUNCH
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
