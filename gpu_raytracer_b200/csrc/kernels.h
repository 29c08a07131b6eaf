// kernels.h — host-side launchers of the sm_100a kernels (internal header).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace rt {

// One BVH2 node, 64 bytes: the boxes of both children plus their ids, so that
// one node fetch (two 32-byte sectors) feeds two slab tests.
//   lo0 = child0.min.xyz, child0 id bits     hi0 = child0.max.xyz, first leaf of child0
//   lo1 = child1.min.xyz, child1 id bits     hi1 = child1.max.xyz, last leaf of child1
// child id: >= 0 internal node index, < 0 leaf (~sorted position).  A node's index
// is its split position g: child0 ends at sorted leaf g, child1 starts at g+1
// (refit.cu); the root's index is in the header.  hi0.w / hi1.w are the node's leaf
// range [L, R], ignored by traversal.
struct __align__(64) Node {
  float4 lo0, hi0, lo1, hi1;
};

// Reordered triangle, 48 bytes: v0.xyz + original primitive index, v1, v2 (raw
// vertices: both intersectors derive what they need with the oracle's operations).
struct __align__(16) SortedTri {
  float4 v0p, v1, v2;
};

// 16-float header living next to the node array (broadcast with it).
struct BvhHeader {
  float root_min[3];
  uint32_t n_tris_lo;
  float root_max[3];
  uint32_t n_tris_hi;
  uint32_t key_bits;
  uint32_t root_node;  // id of the root: a split position, or ~0 when the root is the one leaf
  uint32_t pad[6];
};

// deepest tree the traversal stacks hold: 63 key bits + 32 tie-break bits = 95 levels
constexpr int kMaxDepth = 96;

// parent word: (parent node id << 2) | (local << 1) | slot; kNoParent for the root.
// slot 0 = left child.  local = 1 when the two children met in the shared memory
// of one refit CTA while the tree was built (both leaf ranges inside one aligned
// group of kRefitLeaves leaves); refit_bvh then meets there again.
constexpr uint32_t kNoParent = 0xFFFFFFFFu;
#ifndef RT_REFIT_LOG2
#define RT_REFIT_LOG2 8
#endif
constexpr int kRefitLeavesLog2 = RT_REFIT_LOG2;
constexpr int kRefitLeaves = 1 << kRefitLeavesLog2;

// ---- generators (scene_gen.cu, ray_gen.cu) ---------------------------------
int launch_gen_scene(int kind, uint64_t p0, uint64_t p1, uint64_t seed, uint64_t first,
                     uint64_t count, float4* out, cudaStream_t s);
int launch_gen_primary(const float* cam12, uint32_t width, uint32_t height, uint64_t first,
                       uint64_t count, float4* out, cudaStream_t s);
int launch_gen_random(const float* box6, uint64_t seed, uint64_t first, uint64_t count,
                      float4* out, cudaStream_t s);
// first: global index of rays[0] (the samples of ray i are drawn for index first + i)
int launch_gen_ao(const float4* tris, const float4* rays, const float4* hits, uint64_t n,
                  uint32_t spp, uint64_t seed, float radius, float4* out, cudaStream_t s,
                  uint64_t first = 0);
int launch_shade(const float4* tris, const float4* rays, const float4* hits,
                 const float4* ao_hits, uint64_t n, uint32_t spp, float* rgb, cudaStream_t s);
// *out += number of hit records with prim >= 0 (closest hits, or occluded any-hit rays)
int launch_count_hits(const float4* hits, uint64_t n, unsigned long long* out, cudaStream_t s);

// ---- build (morton.cu, radix_sort.cu, karras.cu, refit.cu) -----------------
// bounds_keys: 6 ordered-int keys, min.xyz then max.xyz; must be pre-set to
// 0xFFFFFFFF x3, 0 x3 (launch_bounds does that itself).
// Also writes the centroid of every triangle's own box (3 planes of n floats: x, y, z), so
// the Morton pass reads 12 B per triangle instead of 48.
int launch_bounds(const float4* tris, uint64_t n, uint32_t* bounds_keys, float* centroids,
                  cudaStream_t s);

// where chunk c of a sort pass starts and ends (host and device agree on this)
struct ChunkGeom {
  uint32_t n, chunk_size, chunks;
  __host__ __device__ uint32_t begin(uint32_t c) const { return c * chunk_size; }
  __host__ __device__ uint32_t end(uint32_t c) const {
    const uint32_t e = (c + 1u) * chunk_size;
    return (e < n && c + 1u < chunks) ? e : n;
  }
};
ChunkGeom sort_geometry(uint64_t n, uint32_t key_bits);

// keys: n u32 (30-bit) or n u64 (63-bit).  When chunk_hist is not null the kernel runs on
// the sort's chunk geometry and also produces the first pass's per-chunk digit histograms
// (chunk_hist [chunks][bins]) and bin totals (totals [bins], zeroed by the caller); returns
// 1 in *fused when it did (default digit widths only), else 0 and the sort histograms itself.
int launch_morton(const float* centroids, uint64_t n, const uint32_t* bounds_keys, uint32_t key_bits,
                  void* keys, uint32_t* chunk_hist, uint32_t* totals, int* fused, cudaStream_t s,
                  uint32_t* k32_hybrid = nullptr);  // 63-bit hybrid sort: *fused = 2, see morton.cu

struct SortWorkspace {
  void* keys_alt = nullptr;     // 63-bit keys: n keys, ping-pong partner of the input
  uint32_t* vals_alt = nullptr; // 63-bit keys: n u32
  void* rec_a = nullptr;        // 30-bit keys: n packed 64-bit records (key << 32 | payload)
  void* rec_b = nullptr;        // 30-bit keys: ping-pong partner of rec_a
  uint32_t* vals = nullptr;     // n u32 (first pass writes it: payload starts as iota)
  uint32_t* hist = nullptr;       // [passes][bins] global bin totals (sort_hist_words)
  uint32_t* chunk_hist = nullptr; // [chunks][bins] per-chunk counts, then offsets
  uint64_t chunk_words = 0;       // sort_chunk_words
};
// digit schedule of the LSD sort: `passes` digits of `bits` bits, lowest first
// (30-bit keys: 3 x 10; 63-bit keys: 7 x 9; rt_set_param("sort.bits") overrides)
struct SortPlan {
  int passes, bits;
};
SortPlan sort_plan(uint32_t key_bits);
uint64_t sort_hist_words(uint32_t key_bits);
uint64_t sort_chunk_words(uint64_t n, uint32_t key_bits);
// Sorts (keys, iota) by key.  Result: *keys_out / *vals_out point at whichever
// ping-pong buffer holds the sorted data.
// first_hist_ready: ws.chunk_hist / ws.hist already hold the first pass's histograms
// (launch_morton produced them); ws.hist must then have been zeroed before that.
int launch_sort(void* keys, uint64_t n, uint32_t key_bits, const SortWorkspace& ws,
                bool iota_payload, void** keys_out, uint32_t** vals_out, cudaStream_t s,
                bool first_hist_ready = false, int max_passes = 0);

// 63-bit keys with an iota payload, hybrid order (radix_sort.cu): three packed passes on the
// top 30 bits, then ranks inside the runs of equal top bits.  ws must hold BOTH layouts
// (rec_a, rec_b, vals for the 30-bit passes; keys_alt, vals_alt for the result), chunk_hist
// sized for either key width; k32: n u32; long_run: one u32, zeroed by the caller.  No host
// synchronisation: when a run is longer than the ranking window the kernels set *long_run,
// the result buffers hold nothing useful, and the caller (who reads the flag at its next
// synchronisation) encodes the keys again and runs launch_sort on all 63 bits.
int launch_sort63_hybrid(uint64_t* k64, uint64_t n, const SortWorkspace& ws, uint32_t* k32,
                         uint32_t* long_run, void** keys_out, uint32_t** vals_out, cudaStream_t s,
                         bool k32_ready = false);

// A4 + A5 fused: topology, triangle reorder, boxes (refit.cu).  keys: the sorted Morton
// codes; flags: n-1 u32, zeroed by the launcher.  Writes nodes (indexed by split
// position), sorted tris, header box and root id.
int launch_build_tree(const float4* tris, const uint32_t* sorted_idx, const void* keys,
                      uint32_t key_bits, uint64_t n, uint32_t* flags, Node* nodes,
                      SortedTri* stris, BvhHeader* header, void* queue_buf, cudaStream_t s,
                      const uint32_t* abort_flag = nullptr);  // non-null and *abort_flag != 0: do nothing
// parent words (2n-1 u32) from the finished nodes; needed by launch_refit_update and
// launch_export_topology only, so the handle derives them on first use
int launch_derive_parents(const Node* nodes, uint64_t n, int root, uint32_t* parent,
                          cudaStream_t s);
// bvh_load / bvh_commit: *bad = 1 when the node array, the root or a primitive index is out
// of range or inconsistent (refit.cu); launch_validate_depth needs the parent words, which
// may only be derived once launch_validate has passed
int launch_validate(const Node* nodes, const SortedTri* stris, uint64_t n, int root, uint32_t* bad,
                    cudaStream_t s);
int launch_validate_depth(const uint32_t* parent, uint64_t n, uint32_t* bad, cudaStream_t s);
// must precede launch_build_tree on the same stream (zeroes flags and the queue head)
int prepare_build_tree(uint64_t n, uint32_t* flags, void* queue_buf, cudaStream_t s);
// bytes of the continuation queue launch_build_tree / launch_refit_update need (queue_buf)
size_t refit_queue_bytes(uint64_t n);
// refit with moved vertices: the primitive index comes from stris[j].v0p.w
int launch_refit_update(const float4* tris, uint64_t n, const uint32_t* parent, uint32_t* flags,
                        Node* nodes, SortedTri* stris, BvhHeader* header, void* queue_buf,
                        cudaStream_t s);
// canonical arrays for the parity accessors (device outputs, may be null)
int launch_export_topology(const Node* nodes, const uint32_t* parent, uint64_t n, uint32_t* left,
                           uint32_t* right, uint32_t* parent_out, float* boxes,
                           const BvhHeader* header, cudaStream_t s);

int launch_unpack_sorted(const SortedTri* stris, uint64_t n, uint32_t* idx, cudaStream_t s);
int launch_widen_keys(const uint32_t* in, uint64_t n, uint64_t* out, cudaStream_t s);

// ---- A5w: wide BVH4 (wide.cu) ---------------------------------------------------
// One BVH4 node, 128 bytes = one cache line: the boxes of up to four children in
// struct-of-arrays form (component k of every float4 belongs to child k) and their
// references.  A reference is
//   internal  the split position g of the BVH2 node the child was collapsed from
//             (Node4s live at the same index as their BVH2 node, so the collapse
//             needs no allocation and no queue),
//   leaf      kWideLeafBit | (count-1) << 29 | first : `count` (1..4) consecutive
//             sorted triangles starting at `first`,
//   empty     kWideEmpty.
struct __align__(128) Node4 {
  float4 lox, loy, loz, hix, hiy, hiz;
  uint4 ref;
  uint4 pad;
};
constexpr uint32_t kWideLeafBit = 0x80000000u;
constexpr uint32_t kWideEmpty = 0x7FFFFFFFu;
constexpr uint32_t kWideFirstMask = (1u << 29) - 1u;
constexpr uint32_t kWideMaxTris = 1u << 29;  // `first` has 29 bits
constexpr int kWideLeafMax = 4;      // upper bound of Params::wide_leaf_max (2 bits of a reference)
constexpr int kWideLeafDefault = 2;  // measured best on B200 (profiles/r01e_wide.md)

// Collapses the BVH2 (split numbering, refit.cu) into Node4s: wide[g] for every BVH2
// node g holding more than kWideLeafMax triangles (n-1 slots).
int launch_collapse4(const Node* nodes, uint64_t n, uint32_t leaf_max, Node4* wide, cudaStream_t s);
// the reference traversal starts from: the root's Node4, or one leaf when the whole
// scene fits a leaf (root = BvhHeader::root_node)
inline uint32_t wide_root_ref(uint64_t n, int root, uint32_t leaf_max) {
  if (n == 0) return kWideEmpty;
  if (n <= (uint64_t)leaf_max) return kWideLeafBit | ((uint32_t)(n - 1) << 29);
  return (uint32_t)root;
}
int launch_trace_wide(const Node4* wide, const SortedTri* stris, uint64_t n_tris, uint32_t root_ref,
                      const float4* rays, uint64_t n_rays, uint32_t flags, float4* hits,
                      unsigned long long* ray_counter, cudaStream_t s,
                      unsigned long long* stats = nullptr);

// ---- A5w: compressed 8-wide BVH (wide8.cu) -----------------------------------------
// 80-byte node: five 16-byte words (layout in wide8.cu and include/gpu_raytracer.h)
struct __align__(16) Node8 {
  uint4 w[5];
};
// the layout as a handle holds it
struct Wide8 {
  Node8* nodes = nullptr;
  SortedTri* tris = nullptr;  // the triangles in wide-node order: a node's leaf triangles are consecutive
  uint32_t n_nodes = 0;
  bool valid = false;
};
size_t wide8_scratch_bytes(uint64_t n);
uint32_t wide8_leaf_max();
uint32_t wide8_node_capacity(uint64_t n);
// out: wide8_node_capacity(n) nodes; tris8: n records; n > wide8_leaf_max().  Synchronises s.
int launch_collapse8(const Node* nodes, const SortedTri* stris, uint64_t n, int root, void* scratch,
                     Node8* out, SortedTri* tris8, uint32_t* n_nodes_out, cudaStream_t s);
int launch_trace_wide8(const Node8* nodes8, const SortedTri* tris8, const float4* rays, uint64_t n_rays,
                       uint32_t flags, float4* hits, unsigned long long* ray_counter, cudaStream_t s,
                       unsigned long long* stats = nullptr, const uint32_t* perm = nullptr);

// ---- compact binary layout (compact.cu; traversal in traverse.cu) ----------------
// The BVH2 node in 32 bytes, i.e. ONE 32-byte request per node step instead of two: the six
// planes of each child box as 16-bit coordinates on a grid that spans the scene box
// (plane = origin + q * step per axis), rounded outwards by at least one step, plus the two
// child ids of Node.  word 3c + a = lo | hi << 16 of child c on axis a.
struct __align__(32) NodeQ {
  uint32_t box[6];
  int32_t c0, c1;
};
struct QGrid {
  float ox, oy, oz;  // plane 0 of each axis (below the scene box by three steps)
  float sx, sy, sz;  // step, > 0
};
struct Compact {
  NodeQ* nodes = nullptr;  // n - 1 nodes, same ids as Node
  QGrid grid = {0, 0, 0, 1, 1, 1};
  bool valid = false;
};
QGrid make_qgrid(const float bounds[6]);
int launch_pack_nodes(const Node* nodes, uint64_t n_nodes, QGrid grid, NodeQ* out, cudaStream_t s);

// ---- traversal (traverse.cu) -------------------------------------------------
// Hand-shake of the streamed mode (host-resident batches): rays arrive while the kernel runs.
struct TracePipe {
  const unsigned long long* watermark;  // device: rays [0, *watermark) have been uploaded
  uint32_t* seg_done;                   // device: retired rays per segment, zeroed
  uint32_t* host_flags;                 // mapped pinned host memory: 1 = segment's hits complete
  uint32_t* error;                      // device: set when the upload stalled for seconds
  int seg_shift;                        // segment = 1 << seg_shift rays
};
int launch_trace(const Node* nodes, const SortedTri* stris, uint64_t n_tris, int root,
                 const float4* rays,
                 uint64_t n_rays, uint32_t flags, float4* hits, unsigned long long* ray_counter,
                 cudaStream_t s, unsigned long long* stats = nullptr,
                 const TracePipe* pipe = nullptr, const uint32_t* perm = nullptr,
                 const Compact* compact = nullptr);
// RT_TRACE_REORDER: 20-bit coherence keys (origin cell, direction bin) of n device rays;
// box = the scene bounds {min.xyz, max.xyz} (host values)
int launch_ray_keys(const float4* rays, uint64_t n, const float* box, uint32_t* keys, cudaStream_t s);

// tuning knobs (rt_set_param).  Process-wide; every field is an atomic, so a knob may be
// changed while other host threads trace or build (each call reads a knob once, when it
// chooses its launch): there is no torn read, only "this call or the next one".
struct Knob {
  std::atomic<int> v;
  explicit constexpr Knob(int x) : v(x) {}
  operator int() const { return v.load(std::memory_order_relaxed); }
  Knob& operator=(int x) {
    v.store(x, std::memory_order_relaxed);
    return *this;
  }
};
struct Params {
  Knob sort_variant{0};
  Knob sort_bits{0};            // 0 = default digit width (sort_plan)
  Knob sort_hybrid{1};          // 63-bit builds: 1 = top-30 LSD + ranks inside runs (default), 0 = 7 passes
  Knob trace_blocks_per_sm{0};  // 0 = kernel default
  Knob trace_variant{1};        // 0 = plain while-while, 1 = dynamic fetch + voted leaf loop
  Knob trace_chunk{0}, trace_refill_min{0}, trace_leaf_min{0};  // 0 = defaults
  Knob trace_sh_stack{-1};      // stack levels kept in shared memory (0, 8, 16, 24); -1 = 8 with RT_TRACE_REORDER, else 0
  Knob trace_compact{1};        // 1 = use the 32-byte binary layout of handles that hold one (RT_BUILD_COMPACT)
  Knob trace_wide{1};           // 1 = use the wide layout of handles that hold one (RT_BUILD_WIDE4 / RT_BUILD_WIDE8)
  Knob wide_leaf_max{kWideLeafDefault};  // triangles per BVH4 leaf (1..4), read when collapsing
  // bit 0: kernels read pinned host rays in place, bit 1: store hits to pinned host in
  // place.  Measured slower than the copy engines on B200/PCIe5 (profiles/r01b_e2e_modes.md),
  // so the default stages both.
  Knob zero_copy{0};
  Knob stage_chunk{0};          // rays per staged chunk (0 = 1M)
  Knob streamed{1};             // 1 = pinned host batches go through the single-launch streamed mode
  Knob render_band{0};          // rt_render_ao: log2 of the rays per band (16..20); 0 = 20
  Knob merge_fused{1};          // trace_rays_merged: 1 = every device merges its slice out of peer memory, 0 = gather on the first device
  Knob render_taper{3};         // rt_render_ao: bit 0 = the first two bands are 1/4 and 1/2 of a band, bit 1 = the last two
  Knob pool_keep_mb{0};         // build/trace temporaries kept cached per device after a call; 0 = all
};
Params& params();
int num_sms();  // of the current device (cached per device)

}  // namespace rt
