// refit.cu — A5: leaf boxes, triangle reorder and the atomic-flag bottom-up
// AABB refit, plus the kernels that turn the traversal
// layout back into the canonical arrays the parity tests compare.
//
// One thread per leaf.  It gathers its triangle through the sorted index,
// stores it in traversal form {v0 | prim, e1, e2} at its sorted position, then
// walks up: it writes the box it carries (and its own child id) into its slot
// of the parent's 64-byte node — one full 32-byte sector — and bumps the
// parent's counter.  The first arriver stops; the second unions with the
// sibling's box and continues.  A CTA owns kRefitLeaves = 1024 consecutive
// leaves; ~97 % of all internal nodes have their whole leaf range inside one
// CTA (karras.cu marks them), and for those the counter and the deposited boxes
// live in shared memory with CTA-scope fences.  Only nodes that straddle CTAs
// use the device-scope protocol: global counter, __threadfence, sibling slot
// re-read from L2 (ld.cg).  ncu on the first version showed 55 % of the stall
// samples on MEMBAR.GPU; this removes them from all but the top of the tree.
// A walker that reaches such a node does not climb on inside its 1024-thread
// CTA (which would stay resident for one thread): it parks {box, id, parent
// word} in a queue and refit_top_kernel, one dense thread per parked walker,
// finishes the top ~3 % of the tree.  10M triangles: 1.02 ms -> 0.60 ms.  Every node slot is written exactly once and read at most
// once: ~96 B per internal node against the 112 B of the textbook layout.
// The walker that leaves the root stores the scene box in the header.
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

// Shared state of one CTA = kRefitLeaves consecutive leaves: arrival counters and
// deposited child boxes of the internal nodes whose leaf range lies inside the
// CTA (index = node id - first leaf of the CTA; a Karras node id is one end of
// its own range, so it falls in the same window).
struct RefitSmem {
  uint32_t flag[kRefitLeaves];
  float box[kRefitLeaves][2][6];
};

// A walker that leaves its CTA's group parks here instead of keeping a 1024-thread
// CTA resident while it climbs the (slow, device-scope) top of the tree.
struct __align__(16) Continuation {
  float lx, ly, lz;
  int child_enc;
  float hx, hy, hz;
  uint32_t pword;
};
constexpr int kContPerCta = 64;  // queue capacity per leaf CTA; overflow climbs inline

// the device-scope protocol, used for nodes whose children belong to different CTAs
__device__ __forceinline__ void climb_global(Box bx, int child_enc, uint32_t pword,
                                             const uint32_t* __restrict__ parent,
                                             uint32_t* __restrict__ flags, Node* __restrict__ nodes,
                                             BvhHeader* __restrict__ header) {
  while (pword != kNoParent) {
    const uint32_t p = pword >> 2, slot = pword & 1u;
    const uint32_t next_pword = parent[p];
    float4* np = reinterpret_cast<float4*>(nodes + p);
    np[2 * slot] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(child_enc));
    np[2 * slot + 1] = make_float4(bx.hx, bx.hy, bx.hz, 0.0f);
    __threadfence();
    if (atomicAdd(flags + p, 1u) == 0u) return;
    // acquire side: the loads below are issued only after the atomic's result is
    // known (control dependency) and go to L2 (ld.cg), where the sibling's
    // fenced stores already are
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

// FROM_INDEX: prim comes from sorted_idx (build); else from stris[j].v0p.w (refit_bvh)
template <bool FROM_INDEX>
__global__ void __launch_bounds__(kRefitLeaves)
refit_kernel(const float4* __restrict__ tris, const uint32_t* __restrict__ sorted_idx, uint32_t n,
             const uint32_t* __restrict__ parent, uint32_t* __restrict__ flags,
             Node* __restrict__ nodes, SortedTri* __restrict__ stris,
             BvhHeader* __restrict__ header, Continuation* __restrict__ queue,
             uint32_t* __restrict__ queue_count, uint32_t capacity) {
  extern __shared__ __align__(16) unsigned char refit_smem_raw[];
  RefitSmem& sm = *reinterpret_cast<RefitSmem*>(refit_smem_raw);
  const uint32_t base = blockIdx.x * kRefitLeaves;
  const uint32_t j = base + threadIdx.x;
  sm.flag[threadIdx.x] = 0;
  __syncthreads();
  if (j >= n) return;
  uint32_t prim;
  if (FROM_INDEX) prim = sorted_idx[j];
  else prim = __float_as_uint(stris[j].v0p.w);
  const float4 a = __ldg(tris + 3 * (uint64_t)prim), b = __ldg(tris + 3 * (uint64_t)prim + 1),
               c = __ldg(tris + 3 * (uint64_t)prim + 2);
  float4* st = reinterpret_cast<float4*>(stris + j);
  st[0] = make_float4(a.x, a.y, a.z, __uint_as_float(prim));
  st[1] = make_float4(b.x - a.x, b.y - a.y, b.z - a.z, 0.0f);
  st[2] = make_float4(c.x - a.x, c.y - a.y, c.z - a.z, 0.0f);
  Box bx;
  bx.lx = omin(omin(a.x, b.x), c.x); bx.ly = omin(omin(a.y, b.y), c.y); bx.lz = omin(omin(a.z, b.z), c.z);
  bx.hx = omax(omax(a.x, b.x), c.x); bx.hy = omax(omax(a.y, b.y), c.y); bx.hz = omax(omax(a.z, b.z), c.z);

  int child_enc = ~(int)j;
  uint32_t pword = parent[(n - 1) + j];
  while (pword != kNoParent && (pword & 2u)) {
    // both children are refitted by this CTA: meet in shared memory
    const uint32_t p = pword >> 2, slot = pword & 1u;
    const uint32_t next_pword = parent[p];  // independent of who wins below: overlap its latency
    float4* np = reinterpret_cast<float4*>(nodes + p);
    np[2 * slot] = make_float4(bx.lx, bx.ly, bx.lz, __int_as_float(child_enc));
    np[2 * slot + 1] = make_float4(bx.hx, bx.hy, bx.hz, 0.0f);
    const uint32_t k = p - base;
    float* mine = sm.box[k][slot];
    mine[0] = bx.lx; mine[1] = bx.ly; mine[2] = bx.lz;
    mine[3] = bx.hx; mine[4] = bx.hy; mine[5] = bx.hz;
    __threadfence_block();
    if (atomicAdd(&sm.flag[k], 1u) == 0u) return;  // first arriver stops
    __threadfence_block();
    const volatile float* sib = sm.box[k][slot ^ 1u];
    bx.lx = omin(bx.lx, sib[0]); bx.ly = omin(bx.ly, sib[1]); bx.lz = omin(bx.lz, sib[2]);
    bx.hx = omax(bx.hx, sib[3]); bx.hy = omax(bx.hy, sib[4]); bx.hz = omax(bx.hz, sib[5]);
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

// canonical arrays from the traversal layout (parity accessors T3/T4)
__global__ void __launch_bounds__(256)
export_kernel(const Node* __restrict__ nodes, const uint32_t* __restrict__ parent, uint32_t n,
              uint32_t* __restrict__ left, uint32_t* __restrict__ right,
              uint32_t* __restrict__ parent_out, float* __restrict__ boxes,
              const BvhHeader* __restrict__ header) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t n_nodes = 2 * n - 1, leaf0 = n - 1;
  if (i >= n_nodes) return;
  const uint32_t pw = parent[i];
  if (parent_out) parent_out[i] = (pw == kNoParent) ? kNoParent : (pw >> 2);
  if (boxes) {
    float* o = boxes + 6 * (uint64_t)i;
    if (pw == kNoParent) {
      for (int k = 0; k < 3; ++k) {
        o[k] = header->root_min[k];
        o[3 + k] = header->root_max[k];
      }
    } else {
      const float4* np = reinterpret_cast<const float4*>(nodes + (pw >> 2)) + 2 * (pw & 1u);
      float4 lo = np[0], hi = np[1];
      o[0] = lo.x; o[1] = lo.y; o[2] = lo.z; o[3] = hi.x; o[4] = hi.y; o[5] = hi.z;
    }
  }
  if (i < leaf0) {
    const int c0 = __float_as_int(nodes[i].lo0.w), c1 = __float_as_int(nodes[i].lo1.w);
    if (left) left[i] = (c0 < 0) ? (leaf0 + (uint32_t)(~c0)) : (uint32_t)c0;
    if (right) right[i] = (c1 < 0) ? (leaf0 + (uint32_t)(~c1)) : (uint32_t)c1;
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

static int refit_common(bool from_index, const float4* tris, const uint32_t* sorted_idx, uint64_t n,
                        const uint32_t* parent, uint32_t* flags, Node* nodes, SortedTri* stris,
                        BvhHeader* header, void* queue_buf, cudaStream_t s) {
  if (n == 0) return 0;
  const unsigned blocks = (unsigned)((n + kRefitLeaves - 1) / kRefitLeaves);
  const uint32_t capacity = blocks * kContPerCta;
  uint32_t* queue_count = reinterpret_cast<uint32_t*>(queue_buf);
  Continuation* queue = reinterpret_cast<Continuation*>(reinterpret_cast<char*>(queue_buf) + 32);
  if (n > 1) RT_CUDA(cudaMemsetAsync(flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
  RT_CUDA(cudaMemsetAsync(queue_count, 0, 32, s));
  const size_t smem = sizeof(RefitSmem);
  RT_CUDA(cudaFuncSetAttribute(refit_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  RT_CUDA(cudaFuncSetAttribute(refit_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (from_index)
    refit_kernel<true><<<blocks, kRefitLeaves, smem, s>>>(tris, sorted_idx, (uint32_t)n, parent, flags,
                                                         nodes, stris, header, queue, queue_count, capacity);
  else
    refit_kernel<false><<<blocks, kRefitLeaves, smem, s>>>(tris, sorted_idx, (uint32_t)n, parent, flags,
                                                          nodes, stris, header, queue, queue_count, capacity);
  RT_LAUNCHED();
  refit_top_kernel<<<(capacity + 255) / 256, 256, 0, s>>>(queue, queue_count, capacity, parent, flags,
                                                          nodes, header);
  RT_LAUNCHED();
  return 0;
}

size_t refit_queue_bytes(uint64_t n) {
  const uint64_t blocks = (n + kRefitLeaves - 1) / kRefitLeaves;
  return 32 + (blocks ? blocks : 1) * kContPerCta * sizeof(Continuation);
}

int launch_refit(const float4* tris, const uint32_t* sorted_idx, uint64_t n,
                 const uint32_t* parent, uint32_t* flags, Node* nodes, SortedTri* stris,
                 BvhHeader* header, void* queue_buf, cudaStream_t s) {
  return refit_common(true, tris, sorted_idx, n, parent, flags, nodes, stris, header, queue_buf, s);
}

int launch_refit_update(const float4* tris, uint64_t n, const uint32_t* parent, uint32_t* flags,
                        Node* nodes, SortedTri* stris, BvhHeader* header, void* queue_buf,
                        cudaStream_t s) {
  return refit_common(false, tris, nullptr, n, parent, flags, nodes, stris, header, queue_buf, s);
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
