/* gpu_raytracer.h — C ABI of libgpu_raytracer_b200.so
 *
 * The drop-in boundary of the B200 ray-tracing core: scene in, hits out.
 *
 * What this replaces in the reference: nothing that exists.  The mounted
 * reference (/root/reference) holds README.md:1-2 ("A learning project about
 * GPU-accelerated raytracing") and test.txt:1 and no interface of any kind
 * (SURVEY.md §8b).  The three entry points build_bvh / trace_rays / free_bvh are
 * the ones BASELINE.json:5 names for the path ("a thin C-ABI (build_bvh /
 * trace_rays / free_bvh) that mirrors the reference's scene-in, hits-out
 * interface"); every other symbol here is an accessor, generator or multi-GPU
 * helper around them.  Each declaration cites the SURVEY.md §8 row it serves.
 *
 * Conventions
 *   - plain C types only; no exceptions cross the boundary; no exit().
 *   - every int-returning call returns RT_OK (0) or a negative RT_ERR_* code;
 *     bvh_last_error() gives the thread-local message of the last failure.
 *   - pointer arguments marked [host|device] are classified with
 *     cudaPointerGetAttributes; host data is staged through device memory by the
 *     library, device data (on the handle's device) is used in place.
 *   - all work runs on the CUDA device that is current when build_bvh is called
 *     (or opts->device) and on the legacy default stream unless a stream is
 *     given; calls are synchronous on return unless named *_async.
 *   - there is no CPU fallback: without a CUDA device every compute call fails
 *     with RT_ERR_NO_DEVICE.
 *   - threads: calls that take a `const rt_bvh*` (trace_rays, trace_camera, rt_render_ao,
 *     the readers) may run on one handle from several host threads at once; calls that take
 *     an `rt_bvh*` (refit_bvh, bvh_widen, bvh_widen8, bvh_compact, bvh_commit, free_bvh)
 *     change the handle and need it to themselves.  build_bvh calls are independent.
 *
 * Data layouts
 *   triangle  12 floats  v0.xyz,_, v1.xyz,_, v2.xyz,_      (3 x float4, w ignored)
 *   ray        8 floats  o.xyz, tmin, d.xyz, tmax          (2 x float4)
 *   hit        4 floats  t, u, v, prim                     (prim = int32 bits;
 *                        -1 on miss with t = +inf; any-hit mode: prim =
 *                        RT_ANY_HIT_PRIM, t = u = v = 0 when occluded)
 */
#ifndef GPU_RAYTRACER_H_
#define GPU_RAYTRACER_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RT_API __attribute__((visibility("default")))
#else
#define RT_API
#endif

#define RT_OK 0
#define RT_ERR_INVALID (-1)   /* bad argument */
#define RT_ERR_CUDA (-2)      /* CUDA runtime error, see bvh_last_error() */
#define RT_ERR_OOM (-3)       /* device or host allocation failed */
#define RT_ERR_NO_DEVICE (-4) /* no usable CUDA device */
#define RT_ERR_LIMIT (-5)     /* size beyond a documented limit */

/* trace_rays flags (A6) */
#define RT_TRACE_CLOSEST 0u
#define RT_TRACE_ANY 1u
#define RT_TRACE_WATERTIGHT 2u /* A7 option: Woop-Benthin-Wald watertight triangle test instead of
                                  Moeller-Trumbore; may be combined with RT_TRACE_ANY */
#define RT_TRACE_REORDER 4u    /* incoherent batches: trace the rays in the order of a coherence key
                                  (origin cell, direction bin; one radix sort of (key, index) per
                                  batch of <= 64M rays).  Hits are written at the rays' own
                                  positions and are bit-identical with and without the flag.
                                  Device buffers, BVH2, >= 64K rays; ignored otherwise.  Pays for
                                  random rays (1.2-1.3x), costs ~15 % on coherent ones. */
#define RT_ANY_HIT_PRIM 0x7FFFFFFF

/* build flags */
#define RT_BUILD_KEEP_KEYS 1u /* keep the sorted Morton codes for bvh_read_sorted */
#define RT_BUILD_TIMINGS 2u   /* record per-stage CUDA events (bvh_build_times) */
#define RT_BUILD_WIDE4 4u     /* A5w: also collapse the tree to a 4-wide BVH; trace_rays then
                                 walks that layout (same hits, bit for bit); n <= 2^29 */

#define RT_BUILD_WIDE8 8u     /* A5w: also collapse the tree to the compressed 8-wide layout (80-byte
                                 nodes, 8-bit child boxes); trace_rays then walks that layout: same
                                 hits, bit for bit.  More than 3 triangles; +48 n bytes (the triangles
                                 in wide-node order) + 80 bytes per wide node.  Wins for incoherent rays */

#define RT_BUILD_COMPACT 16u  /* also pack the binary tree into 32-byte nodes (16-bit child boxes on one
                                 grid per scene, rounded outwards); trace_rays then walks those with one
                                 32-byte request per node step instead of two: same hits, bit for bit.
                                 At least 2 triangles; +32 bytes per node.  Watertight launches and the
                                 wide layouts, when present, are not affected */

/* scene kinds for rt_gen_scene */
#define RT_SCENE_CORNELL 0     /* config 1: 1,024 triangles, no parameters       */
#define RT_SCENE_SOUP 1        /* config 2: p0 = triangle count                  */
#define RT_SCENE_SPHEREFLAKE 2 /* config 3/5: p0 = depth, p1 = segments per edge */
#define RT_SCENE_TERRAIN 3     /* config 4: p0 = cells per side, p1 = base noise spacing (0 = 1024) */

/* maximum triangle count of one BVH: node ids and look-back words are 30-bit */
#define RT_MAX_TRIS ((1ull << 30) - 1)

typedef struct rt_bvh rt_bvh; /* opaque; immutable after build_bvh returns */

typedef struct rt_build_opts {
  uint32_t key_bits; /* 30 (default when 0) or 63: Morton code width (A2) */
  uint32_t flags;    /* RT_BUILD_* */
  int32_t device;    /* CUDA device ordinal, -1 = current device */
  int32_t reserved;
  void* stream;      /* cudaStream_t, NULL = legacy default stream */
} rt_build_opts;

typedef struct rt_bvh_info {
  uint64_t n_tris;
  uint64_t n_nodes;      /* 2 n - 1 (0 when empty) */
  uint32_t key_bits;
  int32_t device;
  float bounds[6];       /* scene AABB min.xyz max.xyz (A1) */
  uint64_t device_bytes; /* memory held by the handle */
} rt_bvh_info;

/* build stages reported by bvh_build_times, in milliseconds */
enum {
  RT_STAGE_BOUNDS = 0, /* A1 */
  RT_STAGE_MORTON = 1, /* A2 + digit histograms */
  RT_STAGE_SORT = 2,   /* A3 */
  RT_STAGE_TOPOLOGY = 3, /* A4 */
  RT_STAGE_REFIT = 4,  /* A5 (leaf boxes, triangle reorder, bottom-up unions) */
  RT_STAGE_TOTAL = 5,  /* all stages, RT_STAGE_WIDE included */
  RT_STAGE_WIDE = 6,   /* A5w collapse to BVH4 (0 without RT_BUILD_WIDE4) */
  RT_STAGE_COUNT = 7
};

/* ---- the three entry points (BASELINE.json:5) ------------- */

/* Builds an LBVH over n_tris triangles.  tris: [host|device] n_tris x 12
 * floats; the caller keeps ownership, the handle holds its own reordered copy.
 * opts may be NULL.  n_tris == 0 gives a valid empty handle (every ray misses). */
RT_API int build_bvh(const float* tris, uint64_t n_tris, const rt_build_opts* opts, rt_bvh** out);

/* Traces n_rays rays.  rays: [host|device] n_rays x 8 floats; hits:
 * [host|device] n_rays x 4 floats.  Safe to call concurrently on one handle.
 * Host buffers: page-locked (cudaHostAlloc / cudaHostRegister) rays AND hits of at least
 * 512K rays take the streamed path (one kernel launch follows the upload, hits return per
 * 128K-ray segment); anything else is staged in 1M-ray chunks.  Same bytes either way. */
RT_API int trace_rays(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags, float* hits);

/* Frees every device allocation of the handle.  NULL is a no-op. */
RT_API void free_bvh(rt_bvh* h);

/* Thread-local message of the last failing call ("" if none). */
RT_API const char* bvh_last_error(void);

/* ---- stream variant ------------------------------------------------------ */

/* Device pointers only; enqueues on stream (cudaStream_t) and returns. */
RT_API int trace_rays_async(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags,
                     float* hits, void* stream);

/* Primary rays [first, first+count) of a width x height pinhole image (cam: host, 12 floats as
 * for rt_gen_primary_rays) are generated on the device and traced; only the hits cross the
 * bus.  hits: [host|device] count x 4 floats.  Same bytes as rt_gen_primary_rays followed by
 * trace_rays; what it saves is the 32 B/ray upload (rays are "generated
 * on device from a seed, never copied from host at scale"). */
RT_API int trace_camera(const rt_bvh* h, const float* cam, uint32_t width, uint32_t height,
                        uint64_t first, uint64_t count, uint32_t flags, float* hits);

/* Like trace_rays on device buffers, and also counts the work done: counters[0]
 * = internal nodes fetched (64 B each), counters[1] = triangles tested (48 B
 * each); these give the algorithmic bytes per ray of SURVEY.md §8(d).  Slower
 * than trace_rays; for measurement only.  counters: host, 2 x u64. */
RT_API int trace_rays_stats(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags,
                            float* hits, uint64_t* counters);

/* ---- accessors used by the parity tests (T2-T4) -------------------------- */

RT_API int bvh_info(const rt_bvh* h, rt_bvh_info* out);
/* keys: n u64 (needs RT_BUILD_KEEP_KEYS), prim_idx: n u32; host, may be NULL */
RT_API int bvh_read_sorted(const rt_bvh* h, uint64_t* keys, uint32_t* prim_idx);
/* canonical node ids: internal i in [0,n-1), leaf j at n-1+j; left/right: n-1
 * u32 each, parent: 2n-1 u32 (0xFFFFFFFF for the root); host, may be NULL */
RT_API int bvh_read_topology(const rt_bvh* h, uint32_t* left, uint32_t* right, uint32_t* parent);
/* boxes: (2n-1) x 6 floats min.xyz max.xyz in canonical node order; host */
RT_API int bvh_read_boxes(const rt_bvh* h, float* boxes);
/* ms: RT_STAGE_COUNT floats (needs RT_BUILD_TIMINGS) */
RT_API int bvh_build_times(const rt_bvh* h, float* ms);

/* ---- refit: same topology, moved vertices -------------- */
RT_API int refit_bvh(rt_bvh* h, const float* tris);

/* ---- A5w: wide BVH4 -------------------------------- */

/* Collapses the handle's binary tree to a 4-wide BVH (what RT_BUILD_WIDE4 does at
 * build time); also for handles filled through bvh_device_buffers + bvh_commit. */
RT_API int bvh_widen(rt_bvh* h);
/* nodes4: host, (n-1) x 128 bytes, may be NULL.  Slot g holds the wide node made from
 * the binary node that splits the sorted triangles after position g: 6 x float4
 * lo.x[4] lo.y[4] lo.z[4] hi.x[4] hi.y[4] hi.z[4], then u32 ref[4], u32 n_children,
 * 3 x u32 padding.  ref: 0x7FFFFFFF empty; bit 31 set = leaf, bits 29-30 = count-1,
 * bits 0-28 = first sorted triangle; otherwise the slot of the child node.  Slots of
 * subtrees with <= 4 triangles are unused (unspecified bytes).  root_ref: the
 * reference traversal starts from. */
RT_API int bvh_read_wide(const rt_bvh* h, void* nodes4, uint32_t* root_ref);

/* The 32-byte binary layout (what RT_BUILD_COMPACT does at build time) of any handle with at
 * least 2 triangles, also one filled through bvh_device_buffers + bvh_commit or bvh_load. */
RT_API int bvh_compact(rt_bvh* h);
/* nodes: host, (n - 1) x 32 bytes (may be NULL); grid6: 6 floats (may be NULL): plane 0 of the
 * grid on x, y, z, then the step on x, y, z.  Node i is the 64-byte node i of bvh_read_boxes:
 *   u32 box[6]; i32 child0, child1;     box[3 c + a] = lo | hi << 16 of child c on axis a
 * with plane q of axis a at grid6[a] + q * grid6[3 + a]; [lo, hi] contains the exact child box
 * and at least one whole step on either side of it. */
RT_API int bvh_read_compact(const rt_bvh* h, void* nodes, float* grid6);

/* Compressed 8-wide layout (what RT_BUILD_WIDE8 does at build time) of any handle with more than
 * 3 triangles, also one filled through bvh_device_buffers + bvh_commit or bvh_load. */
RT_API int bvh_widen8(rt_bvh* h);
/* nodes8: host, n_nodes x 80 bytes (call with nodes8 = tris8 = NULL for the count); tris8: host,
 * n x 12 floats, the triangles in wide-node order (v0.xyz, prim, v1.xyz, 0, v2.xyz, 0).  Node:
 *   f32 p[3]; u8 e[3]; u8 imask; u32 child_base; u32 tri_base; u8 meta[8];
 *   u8 qlo_x[8], qlo_y[8], qlo_z[8], qhi_x[8], qhi_y[8], qhi_z[8];
 * Slots follow octants (bit 2 / 1 / 0 = upper half of the node's box in x / y / z) as far as
 * the children allow.  Child box =
 * [p + qlo * 2^(e-127), p + qhi * 2^(e-127)] (contains the exact box).  meta[s]: 0 empty;
 * 0x20 | 24 + s internal child, stored at child_base + popcount(imask & ((1 << s) - 1));
 * ((1 << count) - 1) << 5 | offset: leaf child of count (1..3) triangles at tri_base + offset.
 * Node 0 is the root. */
RT_API int bvh_read_wide8(const rt_bvh* h, void* nodes8, float* tris8, uint32_t* n_nodes);

/* ---- multi-GPU helpers ----------------------------------- */

/* Device buffers that make up a built BVH, for a broadcast between ranks:
 * ptrs[0]/bytes[0] = node array, ptrs[1]/bytes[1] = reordered triangles,
 * ptrs[2]/bytes[2] = 16-float header (root box, counts). */
RT_API int bvh_device_buffers(const rt_bvh* h, void** ptrs, uint64_t* bytes);
/* Allocates a handle with uninitialised buffers of the right size for n_tris,
 * to be filled through bvh_device_buffers (receiver side of the broadcast). */
RT_API int bvh_alloc_empty(uint64_t n_tris, uint32_t key_bits, rt_bvh** out);
/* Re-reads the header buffer after it was filled by a broadcast, and validates what arrived
 * (child ids, leaf ranges, root, primitive indices, depth): RT_ERR_INVALID for buffers that
 * do not hold a tree the traversal can walk. */
RT_API int bvh_commit(rt_bvh* h);

/* Single-process multi-GPU (SURVEY.md §8b "Multi-GPU entry takes a device list", §8e): one
 * host thread per device inside the library, no collective on the data path.
 *
 * rt_shard_range: [first, first+count) of rank's contiguous share of n rays; shares are cut
 * at multiples of 32 rays, tile [0, n) exactly and differ by less than two granules.
 *
 * build_bvh_multi: one ordinary handle per entry of devices[] (out: n_devices pointers, all
 * NULL on failure).  RT_MULTI_REDUNDANT: every device runs build_bvh on tris.
 * RT_MULTI_BROADCAST: devices[0] builds, the others receive nodes + reordered triangles +
 * header by cudaMemcpyPeerAsync (NVLink where the devices are peers).  tris: host, or device
 * memory of any one device (peers get a temporary copy).  opts->device and opts->stream are
 * ignored.  The same device may appear twice (two shards on one GPU).
 *
 * trace_rays_sharded: rays and hits are HOST buffers holding the whole batch; handle i
 * traces share i (rt_shard_range(n_rays, i, n_devices)) through trace_rays, all devices at
 * once.  Byte-identical to one trace_rays call on the whole batch.
 *
 * trace_random_sharded: config 5: rays first..first+n_rays of the
 * rt_gen_random_rays sequence (box, seed) are generated on the devices, batch rays at a time
 * (0 = 32M), share i on handle i, and traced with flags (RT_TRACE_REORDER allowed).  hits:
 * host buffer for all n_rays hits, or NULL to keep them on the devices' scratch only;
 * n_hit[i]: rays of share i that hit; trace_ms[i]: device time of share i's traversal
 * (CUDA events, ray generation excluded).  n_hit and trace_ms may be NULL. */
#define RT_MULTI_REDUNDANT 0
#define RT_MULTI_BROADCAST 1
RT_API int rt_shard_range(uint64_t n, int rank, int world, uint64_t* first, uint64_t* count);
RT_API int build_bvh_multi(const float* tris, uint64_t n_tris, const rt_build_opts* opts,
                           const int* devices, int n_devices, int strategy, rt_bvh** out);
RT_API int trace_rays_sharded(rt_bvh* const* handles, int n_devices, const float* rays, uint64_t n_rays,
                              uint32_t flags, float* hits);
RT_API int trace_random_sharded(rt_bvh* const* handles, int n_devices, const float* box, uint64_t seed,
                                uint64_t first, uint64_t n_rays, uint32_t flags, uint64_t batch,
                                float* hits, uint64_t* n_hit, float* trace_ms);

/* Sharded BUILD: the scene is cut instead of the rays, for scenes beyond one
 * GPU's memory and for 1/N of the build time.  Part i is an ordinary BVH over a contiguous range
 * of the triangle array; every part traces all rays; the parts' hits are merged per ray by (t,
 * primitive index + the range's first triangle).  Because a ray's accepted hit is the minimum
 * over all triangles by (t, prim) whatever tree holds them, the merged hits equal one
 * build_bvh + trace_rays over the whole array bit for bit (any-hit: occluded if any part is).
 *
 * rt_merge_hits: parts: host array of n_parts (<= 16) DEVICE pointers to n x 4 hit records of
 * the same rays; prim_base[k] is added to part k's primitive indices (NULL = zeros); out:
 * device, n x 4; enqueued on stream.  This is the exchange step: between processes, move the
 * parts with an all-to-all or all-gather of hit records and call it (sharding.ShardedScene).
 * build_bvh_sharded: tris: HOST array; device i builds triangles [prim_base[i], prim_base[i+1]).
 * trace_rays_merged: host rays / host hits; each device uploads and traces all rays, the hits
 * are gathered on handles[0]'s device by peer copies and merged there. */
RT_API int rt_merge_hits(const float* const* parts, const uint32_t* prim_base, int n_parts, uint64_t n,
                         uint32_t flags, float* out, void* stream);
RT_API int build_bvh_sharded(const float* tris, uint64_t n_tris, const rt_build_opts* opts,
                             const int* devices, int n_devices, rt_bvh** out, uint32_t* prim_base);
RT_API int trace_rays_merged(rt_bvh* const* handles, const uint32_t* prim_base, int n_devices,
                             const float* rays, uint64_t n_rays, uint32_t flags, float* hits);

/* ---- scene input and BVH files --------------------- */

/* Reads a Wavefront OBJ file: "v x y z" and "f a b c ..." records (a, a/t, a//n,
 * a/t/n; negative = relative to the vertices read so far); polygons are fanned
 * around their first vertex; every other record is ignored.  *tris_out: host
 * memory, *n_out x 12 floats, to be released with rt_free_host.  Host code only. */
RT_API int rt_load_obj(const char* path, float** tris_out, uint64_t* n_out);
/* Flat binary triangle file: 8-byte magic "RTTRI1", u64 count, count x 12 floats. */
RT_API int rt_save_tris(const char* path, const float* tris, uint64_t n);
RT_API int rt_load_tris(const char* path, float** tris_out, uint64_t* n_out);
RT_API void rt_free_host(void* p);
/* A built BVH as a file (header, node array, reordered triangles) and back, on the
 * current device; a loaded handle traces exactly like the one that was saved.  The
 * optional BVH4 is not stored: call bvh_widen after bvh_load.  bvh_load checks the file's
 * length and validates the tree like bvh_commit: a truncated or corrupt file is
 * RT_ERR_INVALID, never a device fault. */
RT_API int bvh_save(const rt_bvh* h, const char* path);
RT_API int bvh_load(const char* path, rt_bvh** out);

/* ---- on-device generators (A8 and the §8d synthetic scenes) -------------- */

RT_API uint64_t rt_scene_size(int kind, uint64_t p0, uint64_t p1);
/* triangles [first, first+count) of the scene; out: [host|device] count x 12 */
RT_API int rt_gen_scene(int kind, uint64_t p0, uint64_t p1, uint64_t seed, uint64_t first,
                 uint64_t count, float* out);
/* cam: host, 12 floats eye.xyz corner.xyz du.xyz dv.xyz; row-major pixel order */
RT_API int rt_gen_primary_rays(const float* cam, uint32_t width, uint32_t height, uint64_t first,
                        uint64_t count, float* out);
/* box: host, 6 floats; ray i depends on (seed, first + i) only */
RT_API int rt_gen_random_rays(const float* box, uint64_t seed, uint64_t first, uint64_t count, float* out);
/* spp rays per primary ray from its hit point; tris in ORIGINAL order; all
 * pointers [host|device] but of one kind; out: n x spp x 8 floats */
RT_API int rt_gen_ao_rays(const float* tris, const float* rays, const float* hits, uint64_t n,
                   uint32_t spp, uint64_t seed, float radius, float* out);

/* ---- shading and image output --------------------------- */

/* rgb: [host|device] n x 3 floats.  Lambert-ish: |n.d| * (1 - occluded/spp);
 * ao_hits may be NULL (spp = 0).  Misses are black. */
RT_API int rt_shade(const float* tris, const float* rays, const float* hits, const float* ao_hits,
             uint64_t n, uint32_t spp, float* rgb);

/* One frame in one call (BASELINE.json config 3: primary + spp occlusion rays per pixel): the
 * closest-hit trace of n_rays primary rays, rt_gen_ao_rays from their hit points, the any-hit
 * trace of those, and rt_shade: hits (may be NULL) and rgb (n_rays x 3) hold the bytes the four
 * separate calls give.  tris: DEVICE memory, original order (what build_bvh was given);
 * rays / hits / rgb: [host|device]; flags: 0 or RT_TRACE_WATERTIGHT.  The frame flows through
 * three streams in bands of 1M rays, so with page-locked host buffers the upload of band k+1
 * and the download of band k-1 overlap the tracing of band k; the occlusion rays never leave
 * the device.  Synchronous on return. */
RT_API int rt_render_ao(const rt_bvh* h, const float* tris, const float* rays, uint64_t n_rays, uint32_t spp,
                        uint64_t seed, float radius, uint32_t flags, float* hits, float* rgb);
/* The same frame from the 12-float camera of rt_gen_primary_rays: the primary rays of the
 * width x height image are generated on the device band by band, so only the camera goes up and
 * hits (width * height x 4, may be NULL) and rgb (width * height x 3) come down: the bytes of
 * rt_gen_primary_rays followed by rt_render_ao.  cam: host. */
RT_API int rt_render_camera_ao(const rt_bvh* h, const float* tris, const float* cam, uint32_t width, uint32_t height,
                               uint32_t spp, uint64_t seed, float radius, uint32_t flags, float* hits, float* rgb);

/* ---- runtime --------------------------------------------------------------*/

RT_API int rt_device_count(void);
/* number of CUDA kernels this library has launched in this process */
RT_API uint64_t rt_kernel_launches(void);
/* Gives the memory the library keeps cached in its own stream-ordered pool (build and trace
 * temporaries, freed handles) back to the driver; device -1 = every device used so far.
 * Live handles are not touched.  Synchronises the device. */
RT_API int rt_trim(int device);
/* tuning knobs (process-wide, atomic: a knob may be changed while other host threads build
 * or trace; a call reads each knob once, so the change applies to that call or the next):
 * "pool.keep_mb" (memory the pool keeps cached per device between calls; 0, the default,
 * keeps everything, which makes repeated builds allocation-free), "sort.variant", "sort.bits",
 * "trace.variant", "trace.blocks_per_sm", "trace.chunk", "trace.refill_min", "trace.leaf_min",
 * "trace.wide", "trace.compact" (0 = ignore a handle's 32-byte nodes), "wide.leaf_max", "trace.streamed" (0 = always stage host batches),
 * "trace.stage_chunk", "trace.zero_copy", "render.band" (rt_render_ao: log2 of the rays per band,
 * 16..20; 0, the default, = 20), "merge.fused" (trace_rays_merged: 1, the default, = with peer access every device merges its slice of
 * the batch straight out of the other devices' memory; 0 = the parts' hits are copied to the first device),
 * "render.taper" (bit 0: the frame's first two bands are a quarter and
 * a half band, bit 1: so are its last two; default 3), "trace.sh_stack" (0 | 8 | 16 | 24 stack levels in
 * shared memory; -1, the default: 8 for RT_TRACE_REORDER launches, else 0), "build.pdl" (1 = build kernels launched as programmatic dependents, the
 * default), "sort.hybrid" (63-bit builds: 1 = packed passes on the top 30 bits, then ranks inside
 * runs of equal top bits, the default; 0 = seven passes over all 63 bits; same order either way).  Returns RT_ERR_INVALID for an unknown name. */
RT_API int rt_set_param(const char* name, int64_t value);
/* sorts n (key,value) pairs already on the device with the build's radix sort;
 * used by tests and the sort micro-benchmark.  key_bits 30 -> u32 keys, 63 ->
 * u64 keys.  keys/vals are overwritten with the sorted result. ms (may be
 * NULL) receives the device time of the sort alone. */
RT_API int rt_sort_pairs(void* keys, uint32_t* vals, uint64_t n, uint32_t key_bits, float* ms);
/* The same sort without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Sorts n pairs of 30-bit keys (u32, bits 30-31
 * zero) and u32 payloads, both device memory of the CURRENT device, in place, stably, on
 * `stream` (cudaStream_t, NULL = legacy default stream); it enqueues memsets and kernels and
 * returns without synchronising, and calls no cudaMalloc / cudaFree / allocating routine.  The
 * scratch memory is a static device array of 19.9 MB that every GPU which loads the library
 * holds until the process exits; hence the limit: n <= RT_SORT_STATIC_MAX_PAIRS, beyond which
 * the call fails with RT_ERR_LIMIT.  All calls share that array: calls that overlap on one GPU
 * (two streams, two host threads) race on it; calls on one stream are safe back to back.  Same
 * result as rt_sort_pairs(keys, vals, n, 30, NULL), bit for bit. */
#define RT_SORT_STATIC_MAX_PAIRS (1ull << 20)
RT_API int rt_sort_pairs_static(uint32_t* keys, uint32_t* vals, uint64_t n, void* stream);
RT_API const char* rt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GPU_RAYTRACER_H_ */
