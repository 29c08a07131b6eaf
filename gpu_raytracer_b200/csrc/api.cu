// api.cu — host runtime and C ABI of libgpu_raytracer_b200.so
// (include/gpu_raytracer.h).
//
// The runtime owns device buffers (stream-ordered pool allocations, so repeated
// builds do not pay cudaMalloc), sequences the build stages on one stream,
// pipelines host-resident ray batches through pinned-or-pageable copies on
// three streams, and keeps no global mutable state besides tuning knobs and a
// launch counter.  There is no CPU fallback: every entry point fails with
// RT_ERR_NO_DEVICE when no CUDA device is usable.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <mutex>
#include <new>
#include <vector>

#include "host.h"

namespace rt {

std::atomic<uint64_t> g_kernel_launches{0};
std::atomic<int> g_pdl{1};
extern std::atomic<int> g_build_walk_shape;  // refit.cu

namespace {
thread_local char t_error[512] = "";
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
  return code;
}

int fail_cuda(cudaError_t e, const char* what, const char* file, int line) {
  const char* base = strrchr(file, '/');
  snprintf(t_error, sizeof(t_error), "CUDA error %d (%s) at %s:%d: %s", (int)e,
           cudaGetErrorString(e), base ? base + 1 : file, line, what);
  cudaGetLastError();  // clear the sticky-less error state
  if (e == cudaErrorMemoryAllocation) return RT_ERR_OOM;
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return RT_ERR_NO_DEVICE;
  return RT_ERR_CUDA;
}

Params& params() {
  static Params p;
  return p;
}

namespace {
constexpr int kMaxDevices = 64;
std::atomic<int> g_sms[kMaxDevices];
}  // namespace

int num_sms() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    return kNumSMsB200;
  }
  int sms = g_sms[dev].load(std::memory_order_relaxed);
  if (sms == 0) {
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
      cudaGetLastError();
      sms = kNumSMsB200;
    }
    g_sms[dev].store(sms, std::memory_order_relaxed);
  }
  return sms;
}

int ensure_device() {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return fail(RT_ERR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU path",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  return 0;
}

// The library's own stream-ordered memory pool, one per device: handles and temporaries come
// from it, so what it keeps cached between calls (which is what makes repeated builds
// allocation-free) is the library's alone, can be bounded (rt_set_param "pool.keep_mb") and
// can be given back (rt_trim); the device's default pool, which other code in the process
// may share, is left as it was found.
namespace {
std::mutex g_pool_mu;
cudaMemPool_t g_pools[kMaxDevices] = {};
}  // namespace

int device_pool(int device, cudaMemPool_t* out) {
  if (device < 0 || device >= kMaxDevices) return fail(RT_ERR_INVALID, "device ordinal %d out of range", device);
  std::lock_guard<std::mutex> lock(g_pool_mu);
  if (!g_pools[device]) {
    cudaMemPoolProps props;
    memset(&props, 0, sizeof(props));
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    cudaMemPool_t pool;
    RT_CUDA(cudaMemPoolCreate(&pool, &props));
    const int keep = params().pool_keep_mb;
    uint64_t thr = keep > 0 ? (uint64_t)keep << 20 : UINT64_MAX;
    RT_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    g_pools[device] = pool;
  }
  *out = g_pools[device];
  return 0;
}

// stream-ordered allocation from the library's pool of the current device
int pool_alloc(void** out, size_t bytes, cudaStream_t s) {
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  cudaMemPool_t pool;
  RT_TRY(device_pool(dev, &pool));
  RT_CUDA(cudaMallocFromPoolAsync(out, bytes ? bytes : 16, pool, s));
  return 0;
}

// 0 = host (or unregistered), 1 = device/managed memory; *dev = owning device
int classify(const void* p, int* dev) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) {
    if (dev) *dev = a.device;
    return 1;
  }
  return 0;
}

namespace {

// Pinned (page-locked, mapped) host memory can be read and written by kernels
// directly over PCIe.  Returns the device-side alias, or nullptr.
void* mapped_alias(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (a.type == cudaMemoryTypeHost && a.devicePointer != nullptr) return a.devicePointer;
  return nullptr;
}

// page-locked host memory (cudaHostAlloc / cudaHostRegister / torch pin_memory)
bool is_pinned_host(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

}  // namespace
}  // namespace rt

using namespace rt;

namespace {

size_t key_size(uint32_t key_bits) { return key_bits == 30 ? 4 : 8; }

void release(rt_bvh* h) {
  if (!h) return;
  int prev = -1;
  cudaGetDevice(&prev);
  if (prev != h->device) cudaSetDevice(h->device);
  if (h->nodes) cudaFreeAsync(h->nodes, 0);
  if (h->stris) cudaFreeAsync(h->stris, 0);
  if (h->header) cudaFreeAsync(h->header, 0);
  if (h->parent) cudaFreeAsync(h->parent, 0);
  if (h->keys) cudaFreeAsync(h->keys, 0);
  if (h->wide) cudaFreeAsync(h->wide, 0);
  if (h->wide8.nodes) cudaFreeAsync(h->wide8.nodes, 0);
  if (h->wide8.tris) cudaFreeAsync(h->wide8.tris, 0);
  if (h->compact.nodes) cudaFreeAsync(h->compact.nodes, 0);
  if (prev != h->device && prev >= 0) cudaSetDevice(prev);
  delete h;
}

// parent words of a finished tree, derived once per handle (current device = h->device)
int ensure_parents(const rt_bvh* h, cudaStream_t s) {
  std::lock_guard<std::mutex> lock(h->parent_mu);
  if (h->parent || h->n == 0) return 0;
  uint32_t* p = nullptr;
  RT_TRY(pool_alloc((void**)&p, (2 * h->n - 1) * sizeof(uint32_t), s));
  int rc = launch_derive_parents(h->nodes, h->n, h->root, p, s);
  if (rc != 0) {
    cudaFreeAsync(p, s);
    return rc;
  }
  h->parent = p;
  return 0;
}

// (re)builds the BVH4 of a finished BVH2 (A5w); h->root must be current
int widen(rt_bvh* h, cudaStream_t s) {
  if (h->n > kWideMaxTris)
    return fail(RT_ERR_LIMIT, "wide BVH: %llu triangles exceed 2^29", (unsigned long long)h->n);
  int lm = params().wide_leaf_max;
  h->wide_leaf_max = (uint32_t)(lm < 1 ? 1 : (lm > kWideLeafMax ? kWideLeafMax : lm));
  if (!h->wide && h->n > 1) {
    const size_t bytes = (h->n - 1) * sizeof(Node4);
    RT_TRY(pool_alloc((void**)&h->wide, bytes, s));
    h->bytes += bytes;
  }
  RT_TRY(launch_collapse4(h->nodes, h->n, h->wide_leaf_max, h->wide, s));
  h->wide_root = wide_root_ref(h->n, h->root, h->wide_leaf_max);  // build_impl sets it again once the root is known
  h->is_wide = true;
  return 0;
}

// (re)builds the compressed 8-wide layout of a finished BVH2 (A5w, wide8.cu).  Host
// synchronisations: one per batch of levels (the queue length decides when to stop).
int widen8(rt_bvh* h, cudaStream_t s) {
  if (h->n <= wide8_leaf_max())
    return fail(RT_ERR_INVALID, "wide8: the layout needs more than %u triangles", wide8_leaf_max());
  TempPool tmp(s);
  void* scratch = nullptr;
  Node8* big = nullptr;
  const uint32_t capacity = wide8_node_capacity(h->n);
  RT_TRY(tmp.alloc(&scratch, wide8_scratch_bytes(h->n)));
  RT_TRY(tmp.alloc((void**)&big, (size_t)capacity * sizeof(Node8)));
  if (!h->wide8.tris) {
    RT_TRY(pool_alloc((void**)&h->wide8.tris, h->n * sizeof(SortedTri), s));
    h->bytes += h->n * sizeof(SortedTri);
  }
  h->wide8.valid = false;
  uint32_t n_nodes = 0;
  RT_TRY(launch_collapse8(h->nodes, h->stris, h->n, h->root, scratch, big, h->wide8.tris, &n_nodes, s));
  // the worst-case array goes back to the pool; the handle keeps an exact copy
  if (h->wide8.nodes) {
    cudaFreeAsync(h->wide8.nodes, s);
    h->bytes -= (uint64_t)h->wide8.n_nodes * sizeof(Node8);
    h->wide8.nodes = nullptr;
  }
  RT_TRY(pool_alloc((void**)&h->wide8.nodes, (size_t)n_nodes * sizeof(Node8), s));
  RT_CUDA(cudaMemcpyAsync(h->wide8.nodes, big, (size_t)n_nodes * sizeof(Node8), cudaMemcpyDeviceToDevice, s));
  h->bytes += (uint64_t)n_nodes * sizeof(Node8);
  h->wide8.n_nodes = n_nodes;
  h->wide8.valid = true;
  return 0;
}

// (re)packs the 32-byte binary layout of a finished BVH2 (compact.cu); needs h->bounds
int compact(rt_bvh* h, cudaStream_t s) {
  if (h->n < 2) return fail(RT_ERR_INVALID, "compact: the layout needs at least 2 triangles");
  h->compact.valid = false;
  if (!h->compact.nodes) {
    RT_TRY(pool_alloc((void**)&h->compact.nodes, (h->n - 1) * sizeof(NodeQ), s));
    h->bytes += (h->n - 1) * sizeof(NodeQ);
  }
  h->compact.grid = make_qgrid(h->bounds);
  RT_TRY(launch_pack_nodes(h->nodes, h->n - 1, h->compact.grid, h->compact.nodes, s));
  h->compact.valid = true;
  return 0;
}

}  // namespace
// traversal of whichever layout the handle holds
int rt::launch_trace_h(const rt_bvh* h, const float4* rays, uint64_t n_rays, uint32_t flags,
                       float4* hits, unsigned long long* counter, cudaStream_t s,
                       unsigned long long* stats) {
  if (h->wide8.valid && params().trace_wide != 0)
    return launch_trace_wide8(h->wide8.nodes, h->wide8.tris, rays, n_rays, flags, hits, counter, s, stats);
  if (h->is_wide && params().trace_wide != 0)
    return launch_trace_wide(h->wide, h->stris, h->n, h->wide_root, rays, n_rays, flags, hits,
                             counter, s, stats);
  return launch_trace(h->nodes, h->stris, h->n, h->root, rays, n_rays, flags, hits, counter, s,
                      stats, nullptr, nullptr, &h->compact);
}

namespace {
int alloc_handle_buffers(rt_bvh* h, bool keep_keys, cudaStream_t s) {
  const uint64_t n = h->n;
  const size_t node_bytes = (n > 1 ? (n - 1) : 1) * sizeof(Node);
  const size_t tri_bytes = (n ? n : 1) * sizeof(SortedTri);
  RT_TRY(pool_alloc((void**)&h->nodes, node_bytes, s));
  RT_TRY(pool_alloc((void**)&h->stris, tri_bytes, s));
  RT_TRY(pool_alloc((void**)&h->header, sizeof(BvhHeader), s));
  h->bytes = node_bytes + tri_bytes + sizeof(BvhHeader);
  if (keep_keys) {
    RT_TRY(pool_alloc(&h->keys, (n ? n : 1) * key_size(h->key_bits), s));
    h->bytes += (n ? n : 1) * key_size(h->key_bits);
  }
  return 0;
}

}  // namespace

// ===========================================================================
// build_bvh  (call stack 1 of §3)
// ===========================================================================
namespace {

int build_impl(rt_bvh* h, const float* tris_in, uint32_t flags, cudaStream_t s) {
  const uint64_t n = h->n;
  TempPool tmp(s);
  RT_TRY(alloc_handle_buffers(h, (flags & RT_BUILD_KEEP_KEYS) != 0, s));

  BvhHeader hdr;
  memset(&hdr, 0, sizeof(hdr));
  hdr.n_tris_lo = (uint32_t)n;
  hdr.n_tris_hi = (uint32_t)(n >> 32);
  hdr.key_bits = h->key_bits;
  RT_CUDA(cudaMemcpyAsync(h->header, &hdr, sizeof(hdr), cudaMemcpyHostToDevice, s));
  if (n == 0) {
    if (flags & RT_BUILD_WIDE4) RT_TRY(widen(h, s));
    RT_CUDA(cudaStreamSynchronize(s));
    return 0;
  }

  // triangles: use device memory in place, stage host memory
  const float4* tris = nullptr;
  int pdev = -1;
  if (classify(tris_in, &pdev)) {
    if (pdev != h->device)
      return fail(RT_ERR_INVALID, "build_bvh: tris live on device %d, build runs on device %d",
                  pdev, h->device);
    tris = reinterpret_cast<const float4*>(tris_in);
  } else {
    void* d = nullptr;
    RT_TRY(tmp.alloc(&d, n * 48));
    RT_CUDA(cudaMemcpyAsync(d, tris_in, n * 48, cudaMemcpyHostToDevice, s));
    tris = reinterpret_cast<const float4*>(d);
  }

  const size_t ksz = key_size(h->key_bits);
  void *keys = nullptr, *bkeys = nullptr, *flags_buf = nullptr;
  SortWorkspace ws;
  RT_TRY(tmp.alloc(&keys, n * ksz));
  if (h->key_bits == 30) {
    RT_TRY(tmp.alloc(&ws.rec_a, n * 8));
    RT_TRY(tmp.alloc(&ws.rec_b, n * 8));
  } else {
    RT_TRY(tmp.alloc(&ws.keys_alt, n * ksz));
    RT_TRY(tmp.alloc((void**)&ws.vals_alt, n * 4));
  }
  // 63-bit keys, hybrid order (launch_sort63_hybrid): the 30-bit passes' buffers as well
  const bool hybrid = h->key_bits == 63 && params().sort_hybrid != 0 && n >= 2;
  uint32_t *k32 = nullptr, *long_run = nullptr;
  if (hybrid) {
    RT_TRY(tmp.alloc(&ws.rec_a, n * 8));
    RT_TRY(tmp.alloc(&ws.rec_b, n * 8));
    RT_TRY(tmp.alloc((void**)&k32, n * 4));
    RT_TRY(tmp.alloc((void**)&long_run, 32));
  }
  RT_TRY(tmp.alloc((void**)&ws.vals, n * 4));
  RT_TRY(tmp.alloc((void**)&ws.hist, sort_hist_words(h->key_bits) * 4));
  ws.chunk_words = sort_chunk_words(n, h->key_bits);
  if (hybrid && sort_chunk_words(n, 30) > ws.chunk_words) ws.chunk_words = sort_chunk_words(n, 30);
  RT_TRY(tmp.alloc((void**)&ws.chunk_hist, ws.chunk_words * 4));
  RT_TRY(tmp.alloc(&bkeys, 32));
  float* centroids = nullptr;
  RT_TRY(tmp.alloc((void**)&centroids, n * 12));
  RT_TRY(tmp.alloc(&flags_buf, (n > 1 ? n - 1 : 1) * 4));
  void* queue_buf = nullptr;
  RT_TRY(tmp.alloc(&queue_buf, refit_queue_bytes(n)));

  const bool timed = (flags & RT_BUILD_TIMINGS) != 0;
  EventSet evs;  // destroyed on every return path
  if (timed) RT_TRY(evs.create(RT_STAGE_COUNT));
  cudaEvent_t* ev = evs.ev;
  auto mark = [&](int i) -> int {
    if (timed) RT_CUDA(cudaEventRecord(ev[i], s));
    return 0;
  };

  // every memset first: from here on the stream holds kernels only, launched as
  // programmatic dependents of each other (common.cuh pdl_enter)
  RT_CUDA(cudaMemsetAsync(ws.hist, 0, sort_hist_words(h->key_bits) * sizeof(uint32_t), s));
  RT_TRY(prepare_build_tree(n, (uint32_t*)flags_buf, queue_buf, s));
  if (hybrid) RT_CUDA(cudaMemsetAsync(long_run, 0, 4, s));
  RT_TRY(mark(0));
  RT_TRY(launch_bounds(tris, n, (uint32_t*)bkeys, centroids, s));
  RT_TRY(mark(1));
  int hist_ready = 0;
  RT_TRY(launch_morton(centroids, n, (const uint32_t*)bkeys, h->key_bits, keys, ws.chunk_hist,
                       ws.hist, &hist_ready, s, hybrid ? k32 : nullptr));
  RT_TRY(mark(2));
  void* skeys = nullptr;
  uint32_t* svals = nullptr;
  // 63-bit keys, hybrid order: launched on the assumption that no run of equal top bits is
  // longer than the ranking window.  The flag that says otherwise makes the kernels after it
  // return at once (they would read ranks that were never written) and is read with the
  // header at the end of the build: no host round trip in the middle of the stream.
  if (hybrid)
    RT_TRY(launch_sort63_hybrid((uint64_t*)keys, n, ws, k32, long_run, &skeys, &svals, s,
                                hist_ready == 2));
  else
    RT_TRY(launch_sort(keys, n, h->key_bits, ws, /*iota_payload=*/true, &skeys, &svals, s,
                       hist_ready == 1));
  RT_TRY(mark(3));
  RT_TRY(mark(4));  // A4 is fused into the bottom-up pass: the topology stage is empty
  RT_TRY(launch_build_tree(tris, svals, skeys, h->key_bits, n, (uint32_t*)flags_buf, h->nodes,
                           h->stris, h->header, queue_buf, s, hybrid ? long_run : nullptr));
  RT_TRY(mark(5));
  const bool want_wide = (flags & (RT_BUILD_WIDE4 | RT_BUILD_WIDE8 | RT_BUILD_COMPACT)) != 0;
  if (!want_wide) RT_TRY(mark(6));
  uint32_t long_flag = 0;
  if (hybrid) RT_CUDA(cudaMemcpyAsync(&long_flag, long_run, 4, cudaMemcpyDeviceToHost, s));
  if (h->keys) RT_CUDA(cudaMemcpyAsync(h->keys, skeys, n * ksz, cudaMemcpyDeviceToDevice, s));
  RT_CUDA(cudaMemcpyAsync(&hdr, h->header, sizeof(hdr), cudaMemcpyDeviceToHost, s));
  RT_CUDA(cudaStreamSynchronize(s));
  if (long_flag) {
    // a run longer than the window (e.g. thousands of equal centroids).  Encode again (the
    // hybrid pass works on the keys in place) and sort all 63 bits: the same order the hybrid
    // gives where it applies, bit for bit.
    RT_CUDA(cudaMemsetAsync(ws.hist, 0, sort_hist_words(h->key_bits) * sizeof(uint32_t), s));
    RT_TRY(prepare_build_tree(n, (uint32_t*)flags_buf, queue_buf, s));
    RT_TRY(launch_morton(centroids, n, (const uint32_t*)bkeys, h->key_bits, keys, ws.chunk_hist,
                         ws.hist, &hist_ready, s, nullptr));
    RT_TRY(launch_sort(keys, n, h->key_bits, ws, /*iota_payload=*/true, &skeys, &svals, s,
                       hist_ready == 1));
    RT_TRY(mark(3));
    RT_TRY(mark(4));
    RT_TRY(launch_build_tree(tris, svals, skeys, h->key_bits, n, (uint32_t*)flags_buf, h->nodes,
                             h->stris, h->header, queue_buf, s, nullptr));
    RT_TRY(mark(5));
    if (!want_wide) RT_TRY(mark(6));
    if (h->keys) RT_CUDA(cudaMemcpyAsync(h->keys, skeys, n * ksz, cudaMemcpyDeviceToDevice, s));
    RT_CUDA(cudaMemcpyAsync(&hdr, h->header, sizeof(hdr), cudaMemcpyDeviceToHost, s));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  for (int k = 0; k < 3; ++k) {
    h->bounds[k] = hdr.root_min[k];
    h->bounds[3 + k] = hdr.root_max[k];
  }
  h->root = (int)hdr.root_node;
  if (want_wide) {  // A5w, optional: needs the root
    if (flags & RT_BUILD_WIDE4) RT_TRY(widen(h, s));
    if ((flags & RT_BUILD_WIDE8) && n > wide8_leaf_max()) RT_TRY(widen8(h, s));
    if ((flags & RT_BUILD_COMPACT) && n >= 2) RT_TRY(compact(h, s));
    RT_TRY(mark(6));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  h->wide_root = wide_root_ref(n, h->root, h->wide_leaf_max);
  if (timed) {
    for (int i = 0; i < 5; ++i) RT_CUDA(cudaEventElapsedTime(&h->times[i], ev[i], ev[i + 1]));
    RT_CUDA(cudaEventElapsedTime(&h->times[RT_STAGE_WIDE], ev[5], ev[6]));
    RT_CUDA(cudaEventElapsedTime(&h->times[RT_STAGE_TOTAL], ev[0], ev[6]));
    h->has_times = true;
  }
  return 0;
}

}  // namespace

extern "C" int build_bvh(const float* tris, uint64_t n_tris, const rt_build_opts* opts,
                         rt_bvh** out) {
  if (!out) return fail(RT_ERR_INVALID, "build_bvh: out is NULL");
  *out = nullptr;
  if (n_tris && !tris) return fail(RT_ERR_INVALID, "build_bvh: tris is NULL");
  if (n_tris > RT_MAX_TRIS)
    return fail(RT_ERR_LIMIT, "build_bvh: %llu triangles exceed RT_MAX_TRIS",
                (unsigned long long)n_tris);
  uint32_t key_bits = (opts && opts->key_bits) ? opts->key_bits : 30;
  if (key_bits != 30 && key_bits != 63)
    return fail(RT_ERR_INVALID, "build_bvh: key_bits must be 30 or 63, got %u", key_bits);
  RT_TRY(ensure_device());
  DeviceScope scope;
  RT_TRY(scope.enter(opts ? opts->device : -1));
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  rt_bvh* h = new (std::nothrow) rt_bvh;
  if (!h) return fail(RT_ERR_OOM, "build_bvh: host allocation failed");
  h->n = n_tris;
  h->key_bits = key_bits;
  h->device = dev;
  cudaStream_t s = opts ? (cudaStream_t)opts->stream : (cudaStream_t)0;
  int rc = build_impl(h, tris, opts ? opts->flags : 0u, s);
  if (rc != 0) {
    release(h);
    return rc;
  }
  *out = h;
  return RT_OK;
}

extern "C" void free_bvh(rt_bvh* h) { release(h); }

extern "C" const char* bvh_last_error(void) { return rt::t_error; }

extern "C" int refit_bvh(rt_bvh* h, const float* tris_in) {
  if (!h) return fail(RT_ERR_INVALID, "refit_bvh: handle is NULL");
  if (h->n == 0) return RT_OK;
  if (!tris_in) return fail(RT_ERR_INVALID, "refit_bvh: tris is NULL");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  TempPool tmp(s);
  const float4* tris = nullptr;
  int pdev = -1;
  if (classify(tris_in, &pdev)) {
    tris = reinterpret_cast<const float4*>(tris_in);
  } else {
    void* d = nullptr;
    RT_TRY(tmp.alloc(&d, h->n * 48));
    RT_CUDA(cudaMemcpyAsync(d, tris_in, h->n * 48, cudaMemcpyHostToDevice, s));
    tris = reinterpret_cast<const float4*>(d);
  }
  void* flags_buf = nullptr;
  RT_TRY(tmp.alloc(&flags_buf, (h->n > 1 ? h->n - 1 : 1) * 4));
  void* queue_buf = nullptr;
  RT_TRY(tmp.alloc(&queue_buf, refit_queue_bytes(h->n)));
  RT_TRY(ensure_parents(h, s));
  RT_TRY(launch_refit_update(tris, h->n, h->parent, (uint32_t*)flags_buf, h->nodes, h->stris,
                             h->header, queue_buf, s));
  BvhHeader hdr;
  RT_CUDA(cudaMemcpyAsync(&hdr, h->header, sizeof(hdr), cudaMemcpyDeviceToHost, s));
  RT_CUDA(cudaStreamSynchronize(s));
  for (int k = 0; k < 3; ++k) {
    h->bounds[k] = hdr.root_min[k];
    h->bounds[3 + k] = hdr.root_max[k];
  }
  h->root = (int)hdr.root_node;
  if (h->is_wide) {  // same topology, new boxes: collapse again
    RT_TRY(widen(h, s));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  if (h->wide8.valid) {  // and the 8-wide layout
    RT_TRY(widen8(h, s));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  if (h->compact.valid) {  // and the 32-byte nodes, on the grid of the new scene box
    RT_TRY(compact(h, s));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  return RT_OK;
}

extern "C" int bvh_widen(rt_bvh* h) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_widen: handle is NULL");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  RT_TRY(widen(h, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}

extern "C" int bvh_read_wide(const rt_bvh* h, void* nodes4, uint32_t* root_ref) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_wide: handle is NULL");
  if (!h->is_wide) return fail(RT_ERR_INVALID, "bvh_read_wide: the handle holds no BVH4");
  if (root_ref) *root_ref = h->wide_root;
  if (nodes4 && h->wide) {
    DeviceScope scope;
    RT_TRY(scope.enter(h->device));
    RT_CUDA(cudaMemcpy(nodes4, h->wide, (h->n - 1) * sizeof(Node4), cudaMemcpyDeviceToHost));
  }
  return RT_OK;
}

extern "C" int bvh_widen8(rt_bvh* h) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_widen8: handle is NULL");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  RT_TRY(widen8(h, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}

extern "C" int bvh_compact(rt_bvh* h) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_compact: handle is NULL");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  RT_TRY(compact(h, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}

extern "C" int bvh_read_compact(const rt_bvh* h, void* nodes, float* grid6) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_compact: handle is NULL");
  if (!h->compact.valid) return fail(RT_ERR_INVALID, "bvh_read_compact: the handle holds no compact layout");
  if (grid6) memcpy(grid6, &h->compact.grid, 6 * sizeof(float));
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  if (nodes) RT_CUDA(cudaMemcpy(nodes, h->compact.nodes, (h->n - 1) * sizeof(NodeQ), cudaMemcpyDeviceToHost));
  return RT_OK;
}

extern "C" int bvh_read_wide8(const rt_bvh* h, void* nodes8, float* tris8, uint32_t* n_nodes) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_wide8: handle is NULL");
  if (!h->wide8.valid) return fail(RT_ERR_INVALID, "bvh_read_wide8: the handle holds no 8-wide layout");
  if (n_nodes) *n_nodes = h->wide8.n_nodes;
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  if (nodes8) RT_CUDA(cudaMemcpy(nodes8, h->wide8.nodes, (size_t)h->wide8.n_nodes * sizeof(Node8), cudaMemcpyDeviceToHost));
  if (tris8) RT_CUDA(cudaMemcpy(tris8, h->wide8.tris, h->n * sizeof(SortedTri), cudaMemcpyDeviceToHost));
  return RT_OK;
}

// ===========================================================================
// trace_rays  (call stack 2 of §3)
// ===========================================================================
namespace {

// per-thread, per-device resources of the host-buffer pipeline: three streams
// (H2D, compute, D2H), two staging slots, events.  Created on first use.
struct Pipeline {
  static constexpr uint64_t kChunk = 1ull << 20;  // staging slot capacity in rays (32 MiB in, 16 MiB out)
  int device = -1;
  cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
  cudaStream_t s_run2 = nullptr;  // rt_render_ao: the odd bands' compute stream
  float4* rays[2] = {nullptr, nullptr};
  float4* hits[2] = {nullptr, nullptr};
  unsigned long long* counters = nullptr;  // 4: one per trace launch that can be in flight
  cudaEvent_t ev_in[2], ev_run[2], ev_out[2];
  bool ready = false;
  // streamed mode (trace_streamed): one device control block = watermark (u64), error
  // (u32), pad, seg_done[kMaxSegs]; pinned host mirror = flags[kMaxSegs], marks[kMaxChunks]
  static constexpr int kSegShift = 17, kUpShift = 18;          // smallest: 128K-ray segments, 256K-ray uploads
  static constexpr uint64_t kMaxStream = 1ull << 26;           // rays per streamed super-batch
  static constexpr uint32_t kMaxSegs = (uint32_t)(kMaxStream >> kSegShift);
  static constexpr uint32_t kMaxChunks = (uint32_t)(kMaxStream >> kUpShift) + 8;  // + the tapered end
  unsigned char* d_ctl = nullptr;
  uint32_t* h_flags = nullptr;             // pinned, written by the kernel
  unsigned long long* h_marks = nullptr;   // pinned, source of the watermark copies
  static constexpr size_t kCtlBytes = 16 + (size_t)kMaxSegs * 4;
  // releases everything (on the device that owns it); also run when the host thread ends
  void destroy() {
    if (device < 0) return;
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != device) cudaSetDevice(device);
    for (int b = 0; b < 2; ++b) {
      if (rays[b]) cudaFree(rays[b]);
      if (hits[b]) cudaFree(hits[b]);
      rays[b] = hits[b] = nullptr;
      if (ready) {
        cudaEventDestroy(ev_in[b]);
        cudaEventDestroy(ev_run[b]);
        cudaEventDestroy(ev_out[b]);
      }
    }
    if (counters) cudaFree(counters);
    counters = nullptr;
    if (d_ctl) cudaFree(d_ctl);
    if (h_flags) cudaFreeHost(h_flags);
    if (h_marks) cudaFreeHost(h_marks);
    d_ctl = nullptr;
    h_flags = nullptr;
    h_marks = nullptr;
    if (s_in) cudaStreamDestroy(s_in);
    if (s_run) cudaStreamDestroy(s_run);
    if (s_out) cudaStreamDestroy(s_out);
    if (s_run2) cudaStreamDestroy(s_run2);
    s_in = s_run = s_out = s_run2 = nullptr;
    if (prev != device && prev >= 0) cudaSetDevice(prev);
    cudaGetLastError();  // a context torn down at process exit is not an error here
    ready = false;
    device = -1;
  }
  ~Pipeline() { destroy(); }
  int init(int dev) {
    if (ready && device == dev) return 0;
    destroy();  // the thread moved to another device: one pipeline per thread
    device = dev;
    RT_CUDA(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
    RT_CUDA(cudaStreamCreateWithFlags(&s_run, cudaStreamNonBlocking));
    RT_CUDA(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
    RT_CUDA(cudaStreamCreateWithFlags(&s_run2, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      RT_CUDA(cudaMalloc((void**)&rays[b], kChunk * 32));
      RT_CUDA(cudaMalloc((void**)&hits[b], kChunk * 16));
      RT_CUDA(cudaEventCreateWithFlags(&ev_in[b], cudaEventDisableTiming));
      RT_CUDA(cudaEventCreateWithFlags(&ev_run[b], cudaEventDisableTiming));
      RT_CUDA(cudaEventCreateWithFlags(&ev_out[b], cudaEventDisableTiming));
    }
    RT_CUDA(cudaMalloc((void**)&counters, 4 * sizeof(unsigned long long)));
    RT_CUDA(cudaMalloc((void**)&d_ctl, kCtlBytes));
    RT_CUDA(cudaHostAlloc((void**)&h_flags, (size_t)kMaxSegs * 4, cudaHostAllocMapped));
    RT_CUDA(cudaHostAlloc((void**)&h_marks, (size_t)kMaxChunks * 8, cudaHostAllocDefault));
    ready = true;
    return 0;
  }
};
thread_local Pipeline t_pipe;

// RT_TRACE_REORDER on device buffers: per batch of up to 64M rays, coherence keys (origin
// cell, direction bin), the build's radix sort on (key, ray index), and the traversal walks the
// rays in key order through the sorted index; hits land at the rays' own positions.
constexpr uint64_t kReorderBatch = 1ull << 26;
constexpr uint64_t kReorderMin = 1ull << 16;  // below this the sort's fixed cost outweighs the gain
int trace_reordered(const rt_bvh* h, const float4* rays, uint64_t n_rays, uint32_t flags,
                    float4* hits, unsigned long long* counter, cudaStream_t s) {
  TempPool tmp(s);
  const uint64_t cap = n_rays < kReorderBatch ? n_rays : kReorderBatch;
  void* keys = nullptr;
  SortWorkspace ws;
  RT_TRY(tmp.alloc(&keys, cap * 4));
  RT_TRY(tmp.alloc(&ws.rec_a, cap * 8));
  RT_TRY(tmp.alloc(&ws.rec_b, cap * 8));
  RT_TRY(tmp.alloc((void**)&ws.vals, cap * 4));
  RT_TRY(tmp.alloc((void**)&ws.hist, sort_hist_words(30) * 4));
  ws.chunk_words = sort_chunk_words(cap, 30);
  RT_TRY(tmp.alloc((void**)&ws.chunk_hist, ws.chunk_words * 4));
  for (uint64_t first = 0; first < n_rays; first += kReorderBatch) {
    const uint64_t n = (n_rays - first < kReorderBatch) ? n_rays - first : kReorderBatch;
    RT_TRY(launch_ray_keys(rays + 2 * first, n, h->bounds, (uint32_t*)keys, s));
    void* skeys = nullptr;
    uint32_t* perm = nullptr;
    RT_TRY(launch_sort(keys, n, 30, ws, /*iota_payload=*/true, &skeys, &perm, s, false,
                       /*max_passes=*/2));  // 20-bit keys
    if (h->wide8.valid && params().trace_wide != 0)
      RT_TRY(launch_trace_wide8(h->wide8.nodes, h->wide8.tris, rays + 2 * first, n, flags & 3u, hits + first,
                                counter, s, nullptr, perm));
    else
      RT_TRY(launch_trace(h->nodes, h->stris, h->n, h->root, rays + 2 * first, n, flags & 3u,
                          hits + first, counter, s, nullptr, nullptr, perm, &h->compact));
  }
  return 0;
}

}  // namespace
int rt::trace_device(const rt_bvh* h, const float4* rays, uint64_t n_rays, uint32_t flags,
                     float4* hits, cudaStream_t s) {
  unsigned long long* counter = nullptr;
  RT_TRY(pool_alloc((void**)&counter, sizeof(unsigned long long), s));
  int rc;
  if ((flags & RT_TRACE_REORDER) && n_rays >= kReorderMin && h->n != 0 &&
      params().trace_variant != 0 && !(h->is_wide && !h->wide8.valid && params().trace_wide != 0))
    rc = trace_reordered(h, rays, n_rays, flags, hits, counter, s);
  else
    rc = launch_trace_h(h, rays, n_rays, flags & 3u, hits, counter, s);
  cudaFreeAsync(counter, s);
  return rc;
}

namespace {
// host rays and/or host hits: chunked, copies overlap the traversal
int trace_staged(const rt_bvh* h, const float* rays, bool rays_dev, uint64_t n_rays, uint32_t flags,
                 float* hits, bool hits_dev) {
  Pipeline& p = t_pipe;
  RT_TRY(p.init(h->device));
  // the legacy default stream may still be producing the inputs
  RT_CUDA(cudaStreamSynchronize(0));
  uint64_t chunk = params().stage_chunk > 0 ? (uint64_t)params().stage_chunk : (1ull << 20);
  if (chunk > Pipeline::kChunk) chunk = Pipeline::kChunk;
  const uint64_t chunks = (n_rays + chunk - 1) / chunk;
  for (uint64_t c = 0; c < chunks; ++c) {
    const int b = (int)(c & 1);
    const uint64_t first = c * chunk;
    const uint64_t cnt = (n_rays - first < chunk) ? (n_rays - first) : chunk;
    const float4* rsrc;
    if (rays_dev) {
      rsrc = reinterpret_cast<const float4*>(rays) + 2 * first;
    } else {
      if (c >= 2) RT_CUDA(cudaStreamWaitEvent(p.s_in, p.ev_run[b], 0));  // slot's last reader
      RT_CUDA(cudaMemcpyAsync(p.rays[b], rays + 8 * first, cnt * 32, cudaMemcpyHostToDevice, p.s_in));
      RT_CUDA(cudaEventRecord(p.ev_in[b], p.s_in));
      RT_CUDA(cudaStreamWaitEvent(p.s_run, p.ev_in[b], 0));
      rsrc = p.rays[b];
    }
    float4* hdst;
    if (hits_dev) {
      hdst = reinterpret_cast<float4*>(hits) + first;
    } else {
      if (c >= 2) RT_CUDA(cudaStreamWaitEvent(p.s_run, p.ev_out[b], 0));  // slot drained
      hdst = p.hits[b];
    }
    RT_TRY(launch_trace_h(h, rsrc, cnt, flags, hdst, p.counters + b, p.s_run));
    RT_CUDA(cudaEventRecord(p.ev_run[b], p.s_run));
    if (!hits_dev) {
      RT_CUDA(cudaStreamWaitEvent(p.s_out, p.ev_run[b], 0));
      RT_CUDA(cudaMemcpyAsync(hits + 4 * first, p.hits[b], cnt * 16, cudaMemcpyDeviceToHost, p.s_out));
      RT_CUDA(cudaEventRecord(p.ev_out[b], p.s_out));
    }
  }
  RT_CUDA(cudaStreamSynchronize(p.s_run));
  RT_CUDA(cudaStreamSynchronize(p.s_out));
  return 0;
}

// Pinned host rays and hits, BVH2 dynamic-fetch kernel: ONE launch per super-batch.  The
// upload is queued in 256K-ray pieces, each followed by an 8-byte copy that advances the
// device watermark; the kernel starts at once and follows the watermark; the host polls the
// per-segment completion flags the kernel raises in pinned memory and queues each segment's
// device-to-host copy as soon as it is complete.  Against one launch per staged chunk this
// pays the tail of slow rays once, and the copies in both directions overlap the traversal
// at 128K/256K-ray granularity (profiles/r01h_e2e_streamed.md).
// The watermark's advance, ordered after the copy before it on s_in: a stream memory
// operation where the driver has one (a few us), else an 8-byte copy (15-25 us each).
typedef int (*StreamWriteValue64)(cudaStream_t, unsigned long long, unsigned long long, unsigned int);
static StreamWriteValue64 stream_write_value64() {
  static StreamWriteValue64 fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (getenv("RT_NO_STREAM_MEMOP") ||
        cudaGetDriverEntryPoint("cuStreamWriteValue64", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    cudaGetLastError();
    return (StreamWriteValue64)f;
  }();
  return fn;
}
static int advance_watermark(Pipeline& p, uint32_t c, uint64_t value) {
  if (StreamWriteValue64 wv = stream_write_value64()) {
    if (wv(p.s_in, (unsigned long long)(uintptr_t)p.d_ctl, value, 0u) == 0) return 0;
  }
  if (c >= Pipeline::kMaxChunks) return fail(RT_ERR_INVALID, "trace_rays: too many upload pieces");
  p.h_marks[c] = value;
  RT_CUDA(cudaMemcpyAsync(p.d_ctl, &p.h_marks[c], 8, cudaMemcpyHostToDevice, p.s_in));
  return 0;
}
// trace_streamed's verdict when the kernel saw no upload progress for seconds (a profiler
// that serialises or replays kernels against the copy stream does that): not an error, the
// batch is traced again through the staged pipeline and the mode stays off for the process.
constexpr int kStreamStalled = 1;
int trace_streamed(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags, float* hits) {
  Pipeline& p = t_pipe;
  RT_TRY(p.init(h->device));
  RT_CUDA(cudaStreamSynchronize(0));  // the legacy default stream may still be producing the inputs
  cudaEvent_t ev_ctl = p.ev_in[0], ev_k = p.ev_run[0];
  for (uint64_t done = 0; done < n_rays; done += Pipeline::kMaxStream) {
    const uint64_t n = (n_rays - done < Pipeline::kMaxStream) ? n_rays - done : Pipeline::kMaxStream;
    const float* src = rays + 8 * done;
    float* dst = hits + 4 * done;
    // upload pieces of about n/8 rays, 256K..1M: each copy costs the copy engine ~15 us of set-up
    // (10% of an 8 MiB piece), but the kernel can only trace what has arrived
    const int seg_shift = Pipeline::kSegShift;
    int up_shift = Pipeline::kUpShift;
    while (up_shift < 20 && (n >> (up_shift + 1)) >= 6) ++up_shift;
    const uint32_t nseg = (uint32_t)((n + (1ull << seg_shift) - 1) >> seg_shift);
    float4 *d_rays = nullptr, *d_hits = nullptr;
    TempPool batch(p.s_in);  // released on every return path, after the streams have drained
    struct Drain {
      Pipeline& p;
      ~Drain() {
        cudaStreamSynchronize(p.s_run);
        cudaStreamSynchronize(p.s_out);
        cudaStreamSynchronize(p.s_in);
      }
    } drain{p};
    RT_TRY(batch.alloc((void**)&d_rays, n * 32));
    RT_TRY(batch.alloc((void**)&d_hits, n * 16));
    RT_CUDA(cudaMemsetAsync(p.d_ctl, 0, 16 + (size_t)nseg * 4, p.s_in));
    memset(p.h_flags, 0, (size_t)nseg * 4);
    RT_CUDA(cudaEventRecord(ev_ctl, p.s_in));
    RT_CUDA(cudaStreamWaitEvent(p.s_run, ev_ctl, 0));
    RT_CUDA(cudaStreamWaitEvent(p.s_out, ev_ctl, 0));
    // upload pieces, each followed by the watermark's advance.  The last piece is cut small:
    // what is traced after the upload's end is the last piece, spread over all warps
    {
      const uint64_t piece = 1ull << up_shift, tail_piece = 1ull << 16;
      uint32_t c = 0;
      for (uint64_t first = 0; first < n; ++c) {
        const uint64_t left = n - first;
        const uint64_t cnt = left > piece ? piece : (left > tail_piece ? (left > 2 * tail_piece ? left / 2 : tail_piece) : left);
        RT_CUDA(cudaMemcpyAsync(d_rays + 2 * first, src + 8 * first, cnt * 32, cudaMemcpyHostToDevice, p.s_in));
        first += cnt;
        RT_TRY(advance_watermark(p, c, first));
      }
    }
    TracePipe pipe;
    pipe.watermark = reinterpret_cast<const unsigned long long*>(p.d_ctl);
    pipe.error = reinterpret_cast<uint32_t*>(p.d_ctl + 8);
    pipe.seg_done = reinterpret_cast<uint32_t*>(p.d_ctl + 16);
    RT_CUDA(cudaHostGetDevicePointer((void**)&pipe.host_flags, p.h_flags, 0));
    pipe.seg_shift = seg_shift;
    int rc = launch_trace(h->nodes, h->stris, h->n, h->root, d_rays, n, flags, d_hits, p.counters,
                          p.s_run, nullptr, &pipe);
    if (rc == 0 && cudaEventRecord(ev_k, p.s_run) != cudaSuccess) rc = RT_ERR_CUDA;
    // hits flow back segment by segment while the kernel runs
    volatile uint32_t* flag = p.h_flags;
    uint32_t next = 0;
    auto copy_seg = [&](uint32_t sgm) -> cudaError_t {
      const uint64_t first = (uint64_t)sgm << seg_shift;
      const uint64_t cnt = (n - first < (1ull << seg_shift)) ? n - first : (1ull << seg_shift);
      return cudaMemcpyAsync(dst + 4 * first, d_hits + first, cnt * 16, cudaMemcpyDeviceToHost, p.s_out);
    };
    bool kernel_done = false;
    while (rc == 0 && next < nseg) {
      if (flag[next]) {
        if (copy_seg(next) != cudaSuccess) rc = RT_ERR_CUDA;
        ++next;
        continue;
      }
      if (kernel_done) break;  // flags were all written before the kernel ended: a gap means an error
      const cudaError_t q = cudaEventQuery(ev_k);
      if (q == cudaSuccess) kernel_done = true;  // look at the flags once more
      else if (q != cudaErrorNotReady) rc = RT_ERR_CUDA;
    }
    cudaError_t e = cudaStreamSynchronize(p.s_run);
    uint32_t err = 0;
    if (e == cudaSuccess) e = cudaMemcpy(&err, p.d_ctl + 8, 4, cudaMemcpyDeviceToHost);
    if (rc == 0 && e == cudaSuccess && err == 0)
      for (; next < nseg; ++next)
        if (copy_seg(next) != cudaSuccess) rc = RT_ERR_CUDA;
    if (cudaStreamSynchronize(p.s_out) != cudaSuccess) rc = RT_ERR_CUDA;
    cudaStreamSynchronize(p.s_in);
    if (e != cudaSuccess) return fail_cuda(e, "trace_streamed", __FILE__, __LINE__);
    if (err != 0) return kStreamStalled;  // the caller runs the batch again through staged chunks
    if (rc != 0) return fail(rc, "trace_rays: streamed pipeline failed (%s)", cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}

}  // namespace

extern "C" int trace_rays_async(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags,
                                float* hits, void* stream) {
  if (!h) return fail(RT_ERR_INVALID, "trace_rays: handle is NULL");
  if (n_rays == 0) return RT_OK;
  if (!rays || !hits) return fail(RT_ERR_INVALID, "trace_rays: rays or hits is NULL");
  if (flags > 7u) return fail(RT_ERR_INVALID, "trace_rays: unknown flags 0x%x", flags);
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  return trace_device(h, reinterpret_cast<const float4*>(rays), n_rays, flags,
                      reinterpret_cast<float4*>(hits), (cudaStream_t)stream);
}

extern "C" int trace_rays(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags,
                          float* hits) {
  if (!h) return fail(RT_ERR_INVALID, "trace_rays: handle is NULL");
  if (n_rays == 0) return RT_OK;
  if (!rays || !hits) return fail(RT_ERR_INVALID, "trace_rays: rays or hits is NULL");
  if (flags > 7u) return fail(RT_ERR_INVALID, "trace_rays: unknown flags 0x%x", flags);
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  int d0 = -1, d1 = -1;
  const bool rays_dev = classify(rays, &d0) != 0, hits_dev = classify(hits, &d1) != 0;
  if ((rays_dev && d0 != h->device) || (hits_dev && d1 != h->device))
    return fail(RT_ERR_INVALID, "trace_rays: buffers are not on the handle's device %d", h->device);
  // Zero-copy: pinned host buffers are traversed in place.  The persistent
  // kernel pulls each ray over PCIe when a lane needs it and stores the hit
  // straight to host memory, so the transfers overlap the traversal at ray
  // granularity and nothing is staged.
  const int zc = params().zero_copy;
  const void* rays_k = rays_dev ? (const void*)rays : ((zc & 1) ? mapped_alias(rays) : nullptr);
  void* hits_k = hits_dev ? (void*)hits : ((zc & 2) ? mapped_alias(hits) : nullptr);
  if (rays_k && hits_k) {
    RT_TRY(trace_device(h, reinterpret_cast<const float4*>(rays_k), n_rays, flags,
                        reinterpret_cast<float4*>(hits_k), 0));
    RT_CUDA(cudaStreamSynchronize(0));
    return RT_OK;
  }
  flags &= 3u;  // RT_TRACE_REORDER applies to device-resident batches only
  // partial zero-copy: e.g. rays staged by the copy engine, hits stored by the
  // kernel straight into pinned host memory (posted PCIe writes need no staging)
  // both buffers pinned on the host: one kernel launch follows the upload (see trace_streamed)
  if (!rays_dev && !hits_dev && !rays_k && !hits_k && params().streamed != 0 &&
      params().trace_variant != 0 && !((h->is_wide || h->wide8.valid) && params().trace_wide != 0) &&
      n_rays >= (1ull << 19) && is_pinned_host(rays) && is_pinned_host(hits))
  {
    const int rc = trace_streamed(h, rays, n_rays, flags, hits);
    if (rc != kStreamStalled) return rc;
    params().streamed = 0;
  }
  if (hits_k) return trace_staged(h, rays, rays_dev, n_rays, flags, (float*)hits_k, true);
  if (rays_k) return trace_staged(h, (const float*)rays_k, true, n_rays, flags, hits, hits_dev);
  return trace_staged(h, rays, rays_dev, n_rays, flags, hits, hits_dev);
}

// One frame of the config-3 pipeline in one call: closest-hit trace of the primary rays, spp
// occlusion rays per pixel from the hit points (rt_gen_ao_rays), their any-hit trace, and the
// shaded frame (rt_shade) — the bytes the four separate calls give.  The frame is cut into bands
// of 1M rays that flow through three streams: band k+1 is uploaded while band k is traced, and
// band k's hits and pixels go down while band k+1 is traced; the occlusion rays and their hits
// never leave the device.  Host buffers should be page-locked for the copies to overlap.
namespace {
// cam != NULL: the primary rays are those of the width x height pinhole image (n_rays = width *
// height) and are generated on the device band by band; rays is ignored
int render_ao_impl(const rt_bvh* h, const float* tris, const float* rays, const float* cam, uint32_t width,
                   uint32_t height, uint64_t n_rays, uint32_t spp, uint64_t seed, float radius, uint32_t flags,
                   float* hits, float* rgb) {
  if (!h) return fail(RT_ERR_INVALID, "rt_render_ao: handle is NULL");
  if (n_rays == 0) return RT_OK;
  if (!tris || (!rays && !cam) || !rgb) return fail(RT_ERR_INVALID, "rt_render_ao: tris, rays or rgb is NULL");
  if (flags & ~(uint32_t)RT_TRACE_WATERTIGHT) return fail(RT_ERR_INVALID, "rt_render_ao: flags may only hold RT_TRACE_WATERTIGHT");
  if (spp > 64) return fail(RT_ERR_LIMIT, "rt_render_ao: at most 64 occlusion rays per pixel");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  int dt = -1, dr = -1, dh = -1, dg = -1;
  if (!classify(tris, &dt) || dt != h->device)
    return fail(RT_ERR_INVALID, "rt_render_ao: tris (original order) must be device memory of the handle's device");
  const bool gen = cam != nullptr;  // rays made on the device: nothing to upload, one staging slot per band
  const bool rays_dev = !gen && classify(rays, &dr) != 0, hits_dev = hits && classify(hits, &dh) != 0,
             rgb_dev = classify(rgb, &dg) != 0;
  if ((rays_dev && dr != h->device) || (hits_dev && dh != h->device) || (rgb_dev && dg != h->device))
    return fail(RT_ERR_INVALID, "rt_render_ao: device buffers are not on the handle's device %d", h->device);
  Pipeline& p = t_pipe;
  RT_TRY(p.init(h->device));
  RT_CUDA(cudaStreamSynchronize(0));  // the legacy default stream may still be producing the inputs
  // bands: 1M rays (render.band), halved until a band's occlusion rays fit 8M; four buffer slots so that
  // neither the uploads nor the launches wait for a slot to drain in the steady state
  constexpr int kSlots = 4;
  const int band_log2 = params().render_band;
  uint64_t band_max = band_log2 >= 16 && band_log2 <= 20 ? 1ull << band_log2 : Pipeline::kChunk;
  while (spp && band_max * spp > (8ull << 20) && band_max > (1ull << 16)) band_max >>= 1;
  const uint64_t band = n_rays < band_max ? n_rays : band_max;
  // The frame's first bands are a quarter and a half of a band (host rays: the first launch
  // waits for 8 MB, not 32) and so are its last ones (host results: what is still to go down
  // when the last launch ends is a quarter band).  render.taper: bit 0 = head, bit 1 = tail.
  const int taper_knob = params().render_taper;
  const bool big = n_rays >= 6 * band;
  const bool taper_in = big && !rays_dev && !gen && (taper_knob & 1) != 0;
  const bool taper_out = big && (!rgb_dev || (hits && !hits_dev)) && (taper_knob & 2) != 0;
  std::vector<uint64_t> cuts;  // band k = [cuts[k], cuts[k + 1])
  {
    const uint64_t head[2] = {band / 4, band / 2}, tail[2] = {band / 2, band / 4};
    const uint64_t tail_sum = taper_out ? tail[0] + tail[1] : 0;
    uint64_t at = 0;
    cuts.push_back(0);
    if (taper_in)
      for (uint64_t h : head) cuts.push_back(at += h);
    while (n_rays - at > band + tail_sum) cuts.push_back(at += band);
    if (taper_out) {
      // what is left (between tail_sum and band + tail_sum rays): one middle band, then the tail
      const uint64_t rest = n_rays - at;
      if (rest > tail_sum) cuts.push_back(at += rest - tail_sum);
      for (uint64_t t : tail) cuts.push_back(at += t);
    } else {
      cuts.push_back(n_rays);
    }
  }
  const uint64_t bands = cuts.size() - 1;
  const int slots = bands < (uint64_t)kSlots ? (int)bands : kSlots;
  TempPool tmp(p.s_run);
  float4 *d_rays[kSlots] = {}, *d_hits[kSlots] = {}, *d_ao[kSlots] = {}, *d_aoh[kSlots] = {};
  float* d_rgb[kSlots] = {};
  for (int b = 0; b < slots; ++b) {
    if (!rays_dev) RT_TRY(tmp.alloc((void**)&d_rays[b], band * 32));
    if (!hits_dev) RT_TRY(tmp.alloc((void**)&d_hits[b], band * 16));
    if (spp) {
      RT_TRY(tmp.alloc((void**)&d_ao[b], band * spp * 32));
      RT_TRY(tmp.alloc((void**)&d_aoh[b], band * spp * 16));
    }
    if (!rgb_dev) RT_TRY(tmp.alloc((void**)&d_rgb[b], band * 12));
  }
  RT_CUDA(cudaStreamSynchronize(p.s_run));  // the allocations are now usable from the other streams
  struct Drain {  // no stream may still touch the band buffers when they are freed
    Pipeline& p;
    ~Drain() {
      cudaStreamSynchronize(p.s_in);
      cudaStreamSynchronize(p.s_run);
      cudaStreamSynchronize(p.s_run2);
      cudaStreamSynchronize(p.s_out);
    }
  } drain{p};
  // per slot: rays uploaded, primary hits final (their download may start before the AO pass
  // ends), band computed, band downloaded
  struct BandEvents {
    cudaEvent_t in[kSlots], hit[kSlots], run[kSlots], out[kSlots];
    int made = 0;
    int create(int n) {
      for (; made < n; ++made) {
        RT_CUDA(cudaEventCreateWithFlags(&in[made], cudaEventDisableTiming));
        RT_CUDA(cudaEventCreateWithFlags(&hit[made], cudaEventDisableTiming));
        RT_CUDA(cudaEventCreateWithFlags(&run[made], cudaEventDisableTiming));
        RT_CUDA(cudaEventCreateWithFlags(&out[made], cudaEventDisableTiming));
      }
      return 0;
    }
    ~BandEvents() {
      for (int i = 0; i < made; ++i) {
        cudaEventDestroy(in[i]);
        cudaEventDestroy(hit[i]);
        cudaEventDestroy(run[i]);
        cudaEventDestroy(out[i]);
      }
    }
  } ev;
  RT_TRY(ev.create(slots));
  // development aid (RT_RENDER_TIMELINE=1): when each band's upload, first launch, last launch
  // and download ended, in ms from the start of the call, printed to stderr
  const bool timeline = getenv("RT_RENDER_TIMELINE") != nullptr;
  std::vector<cudaEvent_t> tl;
  cudaEvent_t tl0 = nullptr;
  if (timeline) {
    tl.resize(bands * 5);
    for (auto& e : tl) RT_CUDA(cudaEventCreate(&e));
    RT_CUDA(cudaEventCreate(&tl0));
    RT_CUDA(cudaEventRecord(tl0, p.s_in));
  }
  const float4* tris4 = reinterpret_cast<const float4*>(tris);
  // Even and odd bands compute on their own streams: every trace launch is a persistent grid
  // that fills the GPU, so the next band's blocks start as the blocks of this band's launch
  // retire — the drain of one launch (the last, longest rays) is covered by the other band.
  cudaStream_t s_cmp[2] = {p.s_run, p.s_run2};
  for (uint64_t k = 0; k < bands; ++k) {
    const int b = (int)(k % (uint64_t)slots), c = (int)(k & 1);
    cudaStream_t sc = s_cmp[c];
    const bool reused = k >= (uint64_t)slots;
    const uint64_t first = cuts[k], cnt = cuts[k + 1] - first;
    const float4* r4;
    if (rays_dev) {
      r4 = reinterpret_cast<const float4*>(rays) + 2 * first;
    } else if (gen) {
      // the slot's last reader (shade) ran on either compute stream
      if (reused) RT_CUDA(cudaStreamWaitEvent(sc, ev.run[b], 0));
      RT_TRY(launch_gen_primary(cam, width, height, first, cnt, d_rays[b], sc));
      r4 = d_rays[b];
    } else {
      if (reused) RT_CUDA(cudaStreamWaitEvent(p.s_in, ev.run[b], 0));  // the slot's last reader (shade)
      if (timeline) RT_CUDA(cudaEventRecord(tl[k * 5 + 0], p.s_in));
      RT_CUDA(cudaMemcpyAsync(d_rays[b], rays + 8 * first, cnt * 32, cudaMemcpyHostToDevice, p.s_in));
      if (timeline) RT_CUDA(cudaEventRecord(tl[k * 5 + 1], p.s_in));
      RT_CUDA(cudaEventRecord(ev.in[b], p.s_in));
      RT_CUDA(cudaStreamWaitEvent(sc, ev.in[b], 0));
      r4 = d_rays[b];
    }
    if (reused) {
      // the slot's buffers: its previous band ran on either compute stream and was then downloaded
      RT_CUDA(cudaStreamWaitEvent(sc, ev.run[b], 0));
      RT_CUDA(cudaStreamWaitEvent(sc, ev.out[b], 0));
    }
    float4* h4 = hits_dev ? reinterpret_cast<float4*>(hits) + first : d_hits[b];
    float* g3 = rgb_dev ? rgb + 3 * first : d_rgb[b];
    if (timeline) RT_CUDA(cudaEventRecord(tl[k * 5 + 2], sc));
    RT_TRY(launch_trace_h(h, r4, cnt, flags, h4, p.counters + 2 * c, sc));
    RT_CUDA(cudaEventRecord(ev.hit[b], sc));
    if (spp) {
      RT_TRY(launch_gen_ao(tris4, r4, h4, cnt, spp, seed, radius, d_ao[b], sc, first));
      RT_TRY(launch_trace_h(h, d_ao[b], cnt * spp, flags | RT_TRACE_ANY, d_aoh[b], p.counters + 2 * c + 1, sc));
    }
    RT_TRY(launch_shade(tris4, r4, h4, spp ? d_aoh[b] : nullptr, cnt, spp, g3, sc));
    RT_CUDA(cudaEventRecord(ev.run[b], sc));
    if (timeline) RT_CUDA(cudaEventRecord(tl[k * 5 + 3], sc));
    if (hits && !hits_dev) {
      RT_CUDA(cudaStreamWaitEvent(p.s_out, ev.hit[b], 0));
      RT_CUDA(cudaMemcpyAsync(hits + 4 * first, h4, cnt * 16, cudaMemcpyDeviceToHost, p.s_out));
    }
    if (!rgb_dev) {
      RT_CUDA(cudaStreamWaitEvent(p.s_out, ev.run[b], 0));
      RT_CUDA(cudaMemcpyAsync(rgb + 3 * first, g3, cnt * 12, cudaMemcpyDeviceToHost, p.s_out));
    }
    RT_CUDA(cudaEventRecord(ev.out[b], p.s_out));
    if (timeline) RT_CUDA(cudaEventRecord(tl[k * 5 + 4], p.s_out));
  }
  if (timeline) {
    cudaStreamSynchronize(p.s_in); cudaStreamSynchronize(p.s_run); cudaStreamSynchronize(p.s_run2); cudaStreamSynchronize(p.s_out);
    for (uint64_t k = 0; k < bands; ++k) {
      float t[5] = {0, 0, 0, 0, 0};
      for (int i = 0; i < 5; ++i)
        if (cudaEventElapsedTime(&t[i], tl0, tl[k * 5 + i]) != cudaSuccess) { cudaGetLastError(); t[i] = -1; }
      fprintf(stderr, "band %2llu rays %8llu  up %.3f..%.3f  run %.3f..%.3f  down %.3f\n", (unsigned long long)k,
              (unsigned long long)(cuts[k + 1] - cuts[k]), t[0], t[1], t[2], t[3], t[4]);
    }
    for (auto& e : tl) cudaEventDestroy(e);
    cudaEventDestroy(tl0);
  }
  RT_CUDA(cudaStreamSynchronize(p.s_run2));
  RT_CUDA(cudaStreamSynchronize(p.s_run));
  RT_CUDA(cudaStreamSynchronize(p.s_out));
  return RT_OK;
}
}  // namespace

extern "C" int rt_render_ao(const rt_bvh* h, const float* tris, const float* rays, uint64_t n_rays,
                            uint32_t spp, uint64_t seed, float radius, uint32_t flags, float* hits,
                            float* rgb) {
  if (n_rays != 0 && !rays) return fail(RT_ERR_INVALID, "rt_render_ao: rays is NULL");
  return render_ao_impl(h, tris, rays, nullptr, 0, 0, n_rays, spp, seed, radius, flags, hits, rgb);
}

// The same frame from the 12-float camera: the primary rays are generated on the device band by
// band (the bytes of rt_gen_primary_rays), so nothing but the camera goes up.
extern "C" int rt_render_camera_ao(const rt_bvh* h, const float* tris, const float* cam, uint32_t width,
                                   uint32_t height, uint32_t spp, uint64_t seed, float radius, uint32_t flags,
                                   float* hits, float* rgb) {
  if (!cam || width == 0 || height == 0) return fail(RT_ERR_INVALID, "rt_render_camera_ao: bad camera or image size");
  return render_ao_impl(h, tris, nullptr, cam, width, height, (uint64_t)width * height, spp, seed, radius, flags,
                        hits, rgb);
}

// Primary rays never cross PCIe: the 12-float camera goes in, the rays are generated on the
// device (the same bytes rt_gen_primary_rays gives), traced, and only the hits come back.
extern "C" int trace_camera(const rt_bvh* h, const float* cam, uint32_t width, uint32_t height,
                            uint64_t first, uint64_t count, uint32_t flags, float* hits) {
  if (!h) return fail(RT_ERR_INVALID, "trace_camera: handle is NULL");
  if (!cam || width == 0 || height == 0) return fail(RT_ERR_INVALID, "trace_camera: bad camera or image size");
  if (first + count > (uint64_t)width * height) return fail(RT_ERR_INVALID, "trace_camera: range beyond the image");
  if (count == 0) return RT_OK;
  if (!hits) return fail(RT_ERR_INVALID, "trace_camera: hits is NULL");
  if (flags > 7u) return fail(RT_ERR_INVALID, "trace_camera: unknown flags 0x%x", flags);
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  int hd = -1;
  const bool hits_dev = classify(hits, &hd) != 0;
  if (hits_dev && hd != h->device)
    return fail(RT_ERR_INVALID, "trace_camera: hits are not on the handle's device %d", h->device);
  cudaStream_t s = 0;
  TempPool tmp(s);
  float4* d_rays = nullptr;
  RT_TRY(tmp.alloc((void**)&d_rays, count * 32));
  RT_TRY(launch_gen_primary(cam, width, height, first, count, d_rays, s));
  if (hits_dev) {
    RT_TRY(trace_device(h, d_rays, count, flags, reinterpret_cast<float4*>(hits), s));
    RT_CUDA(cudaStreamSynchronize(s));
    return RT_OK;
  }
  // host hits: chunks of the traversal overlap the download of the chunk before
  return trace_staged(h, reinterpret_cast<const float*>(d_rays), true, count, flags & 3u, hits, false);
}

extern "C" int trace_rays_stats(const rt_bvh* h, const float* rays, uint64_t n_rays, uint32_t flags,
                                float* hits, uint64_t* counters) {
  if (!h || !rays || !hits || !counters)
    return fail(RT_ERR_INVALID, "trace_rays_stats: NULL argument");
  if (flags > 3u) return fail(RT_ERR_INVALID, "trace_rays_stats: unknown flags 0x%x", flags);
  counters[0] = counters[1] = 0;
  if (n_rays == 0) return RT_OK;
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  if (!classify(rays, nullptr) || !classify(hits, nullptr))
    return fail(RT_ERR_INVALID, "trace_rays_stats: rays and hits must be device memory");
  cudaStream_t s = 0;
  TempPool tmp(s);
  void* d = nullptr;
  RT_TRY(tmp.alloc(&d, 3 * sizeof(unsigned long long)));
  RT_CUDA(cudaMemsetAsync(d, 0, 3 * sizeof(unsigned long long), s));
  unsigned long long* dc = (unsigned long long*)d;
  RT_TRY(launch_trace_h(h, reinterpret_cast<const float4*>(rays), n_rays, flags,
                        reinterpret_cast<float4*>(hits), dc + 2, s, dc));
  unsigned long long host[2] = {0, 0};
  RT_CUDA(cudaMemcpyAsync(host, dc, sizeof(host), cudaMemcpyDeviceToHost, s));
  RT_CUDA(cudaStreamSynchronize(s));
  counters[0] = host[0];
  counters[1] = host[1];
  return RT_OK;
}

// ===========================================================================
// accessors (parity tests T2-T4)
// ===========================================================================
extern "C" int bvh_info(const rt_bvh* h, rt_bvh_info* out) {
  if (!h || !out) return fail(RT_ERR_INVALID, "bvh_info: NULL argument");
  out->n_tris = h->n;
  out->n_nodes = h->n ? 2 * h->n - 1 : 0;
  out->key_bits = h->key_bits;
  out->device = h->device;
  memcpy(out->bounds, h->bounds, sizeof(h->bounds));
  out->device_bytes = h->bytes;
  return RT_OK;
}

extern "C" int bvh_read_sorted(const rt_bvh* h, uint64_t* keys, uint32_t* prim_idx) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_sorted: handle is NULL");
  if (h->n == 0) return RT_OK;
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  TempPool tmp(s);
  if (keys) {
    if (!h->keys) return fail(RT_ERR_INVALID, "bvh_read_sorted: built without RT_BUILD_KEEP_KEYS");
    if (h->key_bits == 30) {
      void* wide = nullptr;
      RT_TRY(tmp.alloc(&wide, h->n * 8));
      RT_TRY(launch_widen_keys((const uint32_t*)h->keys, h->n, (uint64_t*)wide, s));
      RT_CUDA(cudaMemcpyAsync(keys, wide, h->n * 8, cudaMemcpyDeviceToHost, s));
    } else {
      RT_CUDA(cudaMemcpyAsync(keys, h->keys, h->n * 8, cudaMemcpyDeviceToHost, s));
    }
  }
  if (prim_idx) {
    void* idx = nullptr;
    RT_TRY(tmp.alloc(&idx, h->n * 4));
    RT_TRY(launch_unpack_sorted(h->stris, h->n, (uint32_t*)idx, s));
    RT_CUDA(cudaMemcpyAsync(prim_idx, idx, h->n * 4, cudaMemcpyDeviceToHost, s));
  }
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}

namespace {
int read_canonical(const rt_bvh* h, uint32_t* left, uint32_t* right, uint32_t* parent,
                   float* boxes) {
  if (h->n == 0) return RT_OK;
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  cudaStream_t s = 0;
  TempPool tmp(s);
  const uint64_t n = h->n, nodes = 2 * n - 1;
  void *dl = nullptr, *dr = nullptr, *dp = nullptr, *db = nullptr;
  if (left) RT_TRY(tmp.alloc(&dl, (n - 1) * 4));
  if (right) RT_TRY(tmp.alloc(&dr, (n - 1) * 4));
  if (parent) RT_TRY(tmp.alloc(&dp, nodes * 4));
  if (boxes) RT_TRY(tmp.alloc(&db, nodes * 24));
  RT_TRY(ensure_parents(h, s));
  RT_TRY(launch_export_topology(h->nodes, h->parent, n, (uint32_t*)dl, (uint32_t*)dr,
                                (uint32_t*)dp, (float*)db, h->header, s));
  if (left && n > 1) RT_CUDA(cudaMemcpyAsync(left, dl, (n - 1) * 4, cudaMemcpyDeviceToHost, s));
  if (right && n > 1) RT_CUDA(cudaMemcpyAsync(right, dr, (n - 1) * 4, cudaMemcpyDeviceToHost, s));
  if (parent) RT_CUDA(cudaMemcpyAsync(parent, dp, nodes * 4, cudaMemcpyDeviceToHost, s));
  if (boxes) RT_CUDA(cudaMemcpyAsync(boxes, db, nodes * 24, cudaMemcpyDeviceToHost, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}
}  // namespace

extern "C" int bvh_read_topology(const rt_bvh* h, uint32_t* left, uint32_t* right,
                                 uint32_t* parent) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_topology: handle is NULL");
  return read_canonical(h, left, right, parent, nullptr);
}

extern "C" int bvh_read_boxes(const rt_bvh* h, float* boxes) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_read_boxes: handle is NULL");
  return read_canonical(h, nullptr, nullptr, nullptr, boxes);
}

extern "C" int bvh_build_times(const rt_bvh* h, float* ms) {
  if (!h || !ms) return fail(RT_ERR_INVALID, "bvh_build_times: NULL argument");
  if (!h->has_times) return fail(RT_ERR_INVALID, "bvh_build_times: built without RT_BUILD_TIMINGS");
  memcpy(ms, h->times, sizeof(h->times));
  return RT_OK;
}

// ===========================================================================
// multi-GPU helpers: expose the buffers a rank broadcasts
// ===========================================================================
extern "C" int bvh_device_buffers(const rt_bvh* h, void** ptrs, uint64_t* bytes) {
  if (!h || !ptrs || !bytes) return fail(RT_ERR_INVALID, "bvh_device_buffers: NULL argument");
  ptrs[0] = h->nodes;
  bytes[0] = (h->n > 1 ? h->n - 1 : 0) * sizeof(Node);
  ptrs[1] = h->stris;
  bytes[1] = h->n * sizeof(SortedTri);
  ptrs[2] = h->header;
  bytes[2] = sizeof(BvhHeader);
  return RT_OK;
}

extern "C" int bvh_alloc_empty(uint64_t n_tris, uint32_t key_bits, rt_bvh** out) {
  if (!out) return fail(RT_ERR_INVALID, "bvh_alloc_empty: out is NULL");
  *out = nullptr;
  if (n_tris > RT_MAX_TRIS) return fail(RT_ERR_LIMIT, "bvh_alloc_empty: too many triangles");
  if (key_bits == 0) key_bits = 30;
  if (key_bits != 30 && key_bits != 63) return fail(RT_ERR_INVALID, "bvh_alloc_empty: key_bits");
  RT_TRY(ensure_device());
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  rt_bvh* h = new (std::nothrow) rt_bvh;
  if (!h) return fail(RT_ERR_OOM, "bvh_alloc_empty: host allocation failed");
  h->n = n_tris;
  h->key_bits = key_bits;
  h->device = dev;
  int rc = alloc_handle_buffers(h, false, 0);
  if (rc == 0 && cudaStreamSynchronize(0) != cudaSuccess) rc = RT_ERR_CUDA;
  if (rc != 0) {
    release(h);
    return rc;
  }
  *out = h;
  return RT_OK;
}

int rt::adopt_committed(rt_bvh* dst, const rt_bvh* src) {
  if (!dst || !src || dst->n != src->n) return fail(RT_ERR_INVALID, "adopt_committed: handles differ in size");
  dst->key_bits = src->key_bits;
  memcpy(dst->bounds, src->bounds, sizeof(dst->bounds));
  dst->root = src->root;
  // the layouts the source holds beside the binary tree are derived again on this device
  // (cheaper than copying them, and bit-identical: they are functions of the nodes)
  cudaStream_t s = 0;
  if (src->is_wide) {
    dst->wide_leaf_max = src->wide_leaf_max;
    RT_TRY(widen(dst, s));
  }
  dst->wide_root = wide_root_ref(dst->n, dst->root, dst->wide_leaf_max);
  if (src->wide8.valid) RT_TRY(widen8(dst, s));
  if (src->compact.valid) RT_TRY(compact(dst, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return 0;
}

extern "C" int bvh_commit(rt_bvh* h) {
  if (!h) return fail(RT_ERR_INVALID, "bvh_commit: handle is NULL");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  BvhHeader hdr;
  RT_CUDA(cudaMemcpy(&hdr, h->header, sizeof(hdr), cudaMemcpyDeviceToHost));
  const uint64_t n = ((uint64_t)hdr.n_tris_hi << 32) | hdr.n_tris_lo;
  if (n != h->n)
    return fail(RT_ERR_INVALID, "bvh_commit: header says %llu triangles, handle was sized for %llu",
                (unsigned long long)n, (unsigned long long)h->n);
  if (hdr.key_bits != 30 && hdr.key_bits != 63)
    return fail(RT_ERR_INVALID, "bvh_commit: header says %u key bits", hdr.key_bits);
  // The buffers came from a file or from another process: check everything the kernels index
  // with before anything walks it (refit.cu validate_*), so corrupt data is RT_ERR_INVALID,
  // not a device fault.
  if (n != 0) {
    cudaStream_t s = 0;
    TempPool tmp(s);
    uint32_t* bad = nullptr;
    RT_TRY(tmp.alloc((void**)&bad, sizeof(uint32_t)));
    uint32_t flag = 0;
    RT_TRY(launch_validate(h->nodes, h->stris, n, (int)hdr.root_node, bad, s));
    RT_CUDA(cudaMemcpyAsync(&flag, bad, 4, cudaMemcpyDeviceToHost, s));
    RT_CUDA(cudaStreamSynchronize(s));
    if (flag) return fail(RT_ERR_INVALID, "bvh_commit: node array, root or primitive indices are inconsistent");
    h->root = (int)hdr.root_node;
    {  // a tree re-committed after new data arrived: derive the parents again
      std::lock_guard<std::mutex> lock(h->parent_mu);
      if (h->parent) cudaFreeAsync(h->parent, s);
      h->parent = nullptr;
    }
    RT_TRY(ensure_parents(h, s));
    RT_TRY(launch_validate_depth(h->parent, n, bad, s));
    RT_CUDA(cudaMemcpyAsync(&flag, bad, 4, cudaMemcpyDeviceToHost, s));
    RT_CUDA(cudaStreamSynchronize(s));
    if (flag) return fail(RT_ERR_INVALID, "bvh_commit: the tree is deeper than %d levels", kMaxDepth);
  }
  h->key_bits = hdr.key_bits;
  for (int k = 0; k < 3; ++k) {
    h->bounds[k] = hdr.root_min[k];
    h->bounds[3 + k] = hdr.root_max[k];
  }
  h->root = (int)hdr.root_node;
  return RT_OK;
}

// ===========================================================================
// BVH serialisation: header, node array, reordered triangles
// ===========================================================================
namespace {
constexpr char kBvhMagic[8] = {'R', 'T', 'B', 'V', 'H', '1', 0, 0};
constexpr size_t kIoChunk = 64u << 20;

// device -> file or file -> device through one pinned bounce buffer
int stream_file(FILE* f, void* dev, size_t bytes, bool to_file) {
  if (bytes == 0) return 0;
  void* host = nullptr;
  const size_t chunk = bytes < kIoChunk ? bytes : kIoChunk;
  RT_CUDA(cudaMallocHost(&host, chunk));
  int rc = 0;
  for (size_t off = 0; off < bytes && rc == 0; off += chunk) {
    const size_t len = (bytes - off < chunk) ? bytes - off : chunk;
    if (to_file) {
      if (cudaMemcpy(host, (char*)dev + off, len, cudaMemcpyDeviceToHost) != cudaSuccess) rc = RT_ERR_CUDA;
      else if (fwrite(host, 1, len, f) != len) rc = RT_ERR_INVALID;
    } else {
      if (fread(host, 1, len, f) != len) rc = RT_ERR_INVALID;
      else if (cudaMemcpy((char*)dev + off, host, len, cudaMemcpyHostToDevice) != cudaSuccess) rc = RT_ERR_CUDA;
    }
  }
  cudaFreeHost(host);
  if (rc == RT_ERR_CUDA) return fail(RT_ERR_CUDA, "BVH file transfer: CUDA copy failed");
  if (rc != 0) return fail(RT_ERR_INVALID, "BVH file transfer: short %s", to_file ? "write" : "read");
  return 0;
}
}  // namespace

extern "C" int bvh_save(const rt_bvh* h, const char* path) {
  if (!h || !path) return fail(RT_ERR_INVALID, "bvh_save: NULL argument");
  DeviceScope scope;
  RT_TRY(scope.enter(h->device));
  FILE* f = fopen(path, "wb");
  if (!f) return fail(RT_ERR_INVALID, "bvh_save: cannot open %s", path);
  const uint64_t n = h->n;
  const uint32_t meta[2] = {h->key_bits, 0u};
  int rc = (fwrite(kBvhMagic, 1, 8, f) == 8 && fwrite(&n, 8, 1, f) == 1 && fwrite(meta, 4, 2, f) == 2)
               ? 0 : fail(RT_ERR_INVALID, "bvh_save: short write to %s", path);
  if (rc == 0) rc = stream_file(f, h->header, sizeof(BvhHeader), true);
  if (rc == 0) rc = stream_file(f, h->nodes, (n > 1 ? n - 1 : 0) * sizeof(Node), true);
  if (rc == 0) rc = stream_file(f, h->stris, n * sizeof(SortedTri), true);
  if (fclose(f) != 0 && rc == 0) rc = fail(RT_ERR_INVALID, "bvh_save: cannot close %s", path);
  return rc;
}

extern "C" int bvh_load(const char* path, rt_bvh** out) {
  if (!path || !out) return fail(RT_ERR_INVALID, "bvh_load: NULL argument");
  *out = nullptr;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(RT_ERR_INVALID, "bvh_load: cannot open %s", path);
  char magic[8];
  uint64_t n = 0;
  uint32_t meta[2] = {0, 0};
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, kBvhMagic, 8) != 0 || fread(&n, 8, 1, f) != 1 ||
      fread(meta, 4, 2, f) != 2) {
    fclose(f);
    return fail(RT_ERR_INVALID, "bvh_load: %s is not a BVH file", path);
  }
  if (n > RT_MAX_TRIS || (meta[0] != 30 && meta[0] != 63)) {
    fclose(f);
    return fail(RT_ERR_INVALID, "bvh_load: %s: bad triangle count or key width", path);
  }
  {  // the file must hold exactly header + nodes + triangles: a truncated file fails here
    const long long want = 24 + (long long)sizeof(BvhHeader) + (long long)(n > 1 ? n - 1 : 0) * (long long)sizeof(Node) +
                           (long long)n * (long long)sizeof(SortedTri);
    long long have = -1;
    if (fseek(f, 0, SEEK_END) == 0) have = ftell(f);
    if (have != want || fseek(f, 24, SEEK_SET) != 0) {
      fclose(f);
      return fail(RT_ERR_INVALID, "bvh_load: %s holds %lld bytes, a BVH of %llu triangles needs %lld", path,
                  have, (unsigned long long)n, want);
    }
  }
  rt_bvh* h = nullptr;
  int rc = bvh_alloc_empty(n, meta[0], &h);
  if (rc == 0) rc = stream_file(f, h->header, sizeof(BvhHeader), false);
  if (rc == 0) rc = stream_file(f, h->nodes, (n > 1 ? n - 1 : 0) * sizeof(Node), false);
  if (rc == 0) rc = stream_file(f, h->stris, n * sizeof(SortedTri), false);
  fclose(f);
  if (rc == 0) rc = bvh_commit(h);
  if (rc != 0) {
    release(h);
    return rc;
  }
  *out = h;
  return RT_OK;
}

// ===========================================================================
// generators and shading
// ===========================================================================
namespace {

// runs fn(device_out) and, when out is host memory, copies the result back
template <typename Fn>
int with_output(void* out, size_t bytes, Fn fn) {
  RT_TRY(ensure_device());
  if (bytes == 0) return RT_OK;
  if (!out) return fail(RT_ERR_INVALID, "output pointer is NULL");
  cudaStream_t s = 0;
  if (classify(out, nullptr)) {
    RT_TRY(fn(out, s));
    RT_CUDA(cudaStreamSynchronize(s));
    return RT_OK;
  }
  TempPool tmp(s);
  void* d = nullptr;
  RT_TRY(tmp.alloc(&d, bytes));
  RT_TRY(fn(d, s));
  RT_CUDA(cudaMemcpyAsync(out, d, bytes, cudaMemcpyDeviceToHost, s));
  RT_CUDA(cudaStreamSynchronize(s));
  return RT_OK;
}

// device view of an input: in place when already on the device, staged otherwise
int stage_input(TempPool& tmp, const void* p, size_t bytes, const void** out) {
  if (classify(p, nullptr)) {
    *out = p;
    return 0;
  }
  void* d = nullptr;
  RT_TRY(tmp.alloc(&d, bytes));
  RT_CUDA(cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, tmp.s));
  *out = d;
  return 0;
}

}  // namespace

extern "C" uint64_t rt_scene_size(int kind, uint64_t p0, uint64_t p1) {
  switch (kind) {
    case RT_SCENE_CORNELL: return 1024;
    case RT_SCENE_SOUP: return p0;
    case RT_SCENE_SPHEREFLAKE: {
      uint64_t spheres = 0, p = 1;
      for (uint64_t l = 0; l <= p0 && l < 16; ++l) {
        spheres += p;
        p *= 9;
      }
      return spheres * 8ull * p1 * p1;
    }
    case RT_SCENE_TERRAIN: return 2ull * p0 * p0;
  }
  return 0;
}

extern "C" int rt_gen_scene(int kind, uint64_t p0, uint64_t p1, uint64_t seed, uint64_t first,
                            uint64_t count, float* out) {
  if (kind < 0 || kind > 3) return fail(RT_ERR_INVALID, "rt_gen_scene: unknown scene kind %d", kind);
  if (first + count > rt_scene_size(kind, p0, p1))
    return fail(RT_ERR_INVALID, "rt_gen_scene: range beyond the scene's %llu triangles",
                (unsigned long long)rt_scene_size(kind, p0, p1));
  if (kind == RT_SCENE_SPHEREFLAKE && (p0 > 8 || p1 == 0 || p1 > 1024))
    return fail(RT_ERR_INVALID, "rt_gen_scene: sphereflake depth <= 8, 1 <= segments <= 1024");
  if (kind == RT_SCENE_TERRAIN && p0 >= (1u << 23))
    return fail(RT_ERR_INVALID, "rt_gen_scene: terrain side too large");
  return with_output(out, count * 48, [&](void* d, cudaStream_t s) {
    return launch_gen_scene(kind, p0, p1, seed, first, count, (float4*)d, s);
  });
}

extern "C" int rt_gen_primary_rays(const float* cam, uint32_t width, uint32_t height,
                                   uint64_t first, uint64_t count, float* out) {
  if (!cam || width == 0 || height == 0)
    return fail(RT_ERR_INVALID, "rt_gen_primary_rays: bad camera or image size");
  if (first + count > (uint64_t)width * height)
    return fail(RT_ERR_INVALID, "rt_gen_primary_rays: range beyond the image");
  return with_output(out, count * 32, [&](void* d, cudaStream_t s) {
    return launch_gen_primary(cam, width, height, first, count, (float4*)d, s);
  });
}

extern "C" int rt_gen_random_rays(const float* box, uint64_t seed, uint64_t first, uint64_t count,
                                  float* out) {
  if (!box) return fail(RT_ERR_INVALID, "rt_gen_random_rays: box is NULL");
  return with_output(out, count * 32, [&](void* d, cudaStream_t s) {
    return launch_gen_random(box, seed, first, count, (float4*)d, s);
  });
}

extern "C" int rt_gen_ao_rays(const float* tris, const float* rays, const float* hits, uint64_t n,
                              uint32_t spp, uint64_t seed, float radius, float* out) {
  if (!tris || !rays || !hits) return fail(RT_ERR_INVALID, "rt_gen_ao_rays: NULL input");
  return with_output(out, n * spp * 32, [&](void* d, cudaStream_t s) {
    // tris is indexed by primitive id: its extent is unknown here, so it must
    // already be on the device unless rays/hits are host too (small cases):
    TempPool tmp(s);
    const void *dr = nullptr, *dh = nullptr;
    if (!classify(tris, nullptr))
      return fail(RT_ERR_INVALID, "rt_gen_ao_rays: tris must be device memory");
    RT_TRY(stage_input(tmp, rays, n * 32, &dr));
    RT_TRY(stage_input(tmp, hits, n * 16, &dh));
    return launch_gen_ao((const float4*)tris, (const float4*)dr, (const float4*)dh, n, spp, seed,
                         radius, (float4*)d, s);
  });
}

extern "C" int rt_shade(const float* tris, const float* rays, const float* hits,
                        const float* ao_hits, uint64_t n, uint32_t spp, float* rgb) {
  if (!tris || !rays || !hits) return fail(RT_ERR_INVALID, "rt_shade: NULL input");
  return with_output(rgb, n * 12, [&](void* d, cudaStream_t s) {
    TempPool tmp(s);
    const void *dr = nullptr, *dh = nullptr, *da = nullptr;
    if (!classify(tris, nullptr)) return fail(RT_ERR_INVALID, "rt_shade: tris must be device memory");
    RT_TRY(stage_input(tmp, rays, n * 32, &dr));
    RT_TRY(stage_input(tmp, hits, n * 16, &dh));
    if (ao_hits && spp) RT_TRY(stage_input(tmp, ao_hits, n * spp * 16, &da));
    return launch_shade((const float4*)tris, (const float4*)dr, (const float4*)dh,
                        (const float4*)da, n, spp, (float*)d, s);
  });
}

// ===========================================================================
// runtime
// ===========================================================================
extern "C" int rt_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return count;
}

extern "C" uint64_t rt_kernel_launches(void) { return g_kernel_launches.load(); }

extern "C" int rt_set_param(const char* name, int64_t value) {
  if (!name) return fail(RT_ERR_INVALID, "rt_set_param: name is NULL");
  Params& p = params();
  if (!strcmp(name, "sort.variant")) p.sort_variant = (int)value;
  else if (!strcmp(name, "sort.bits")) p.sort_bits = (int)value;
  else if (!strcmp(name, "sort.hybrid")) p.sort_hybrid = (int)value;
  else if (!strcmp(name, "trace.blocks_per_sm")) p.trace_blocks_per_sm = (int)value;
  else if (!strcmp(name, "trace.variant")) p.trace_variant = (int)value;
  else if (!strcmp(name, "trace.wide")) p.trace_wide = (int)value;
  else if (!strcmp(name, "trace.compact")) p.trace_compact = (int)value;
  else if (!strcmp(name, "trace.streamed")) p.streamed = (int)value;
  else if (!strcmp(name, "wide.leaf_max")) p.wide_leaf_max = (int)value;
  else if (!strcmp(name, "trace.zero_copy")) p.zero_copy = (int)value;
  else if (!strcmp(name, "trace.stage_chunk")) p.stage_chunk = (int)value;
  else if (!strcmp(name, "trace.chunk")) p.trace_chunk = (int)value;
  else if (!strcmp(name, "trace.refill_min")) p.trace_refill_min = (int)value;
  else if (!strcmp(name, "trace.leaf_min")) p.trace_leaf_min = (int)value;
  else if (!strcmp(name, "trace.sh_stack")) p.trace_sh_stack = (int)value;
  else if (!strcmp(name, "build.pdl")) g_pdl = value != 0;
  else if (!strcmp(name, "build.walk")) g_build_walk_shape = (int)value;
  else if (!strcmp(name, "render.band")) p.render_band = (int)value;
  else if (!strcmp(name, "render.taper")) p.render_taper = (int)value;
  else if (!strcmp(name, "merge.fused")) p.merge_fused = (int)value;
  else if (!strcmp(name, "pool.keep_mb")) {
    p.pool_keep_mb = (int)value;
    std::lock_guard<std::mutex> lock(g_pool_mu);
    const uint64_t thr = value > 0 ? (uint64_t)value << 20 : UINT64_MAX;
    for (int d = 0; d < kMaxDevices; ++d)
      if (g_pools[d]) cudaMemPoolSetAttribute(g_pools[d], cudaMemPoolAttrReleaseThreshold, (void*)&thr);
    cudaGetLastError();
  }
  else return fail(RT_ERR_INVALID, "rt_set_param: unknown parameter '%s'", name);
  return RT_OK;
}

extern "C" int rt_trim(int device) {
  RT_TRY(ensure_device());
  int first = device, last = device;
  if (device < 0) {
    first = 0;
    last = kMaxDevices - 1;
  }
  if (last >= kMaxDevices) return fail(RT_ERR_INVALID, "rt_trim: device ordinal %d out of range", device);
  for (int d = first; d <= last; ++d) {
    cudaMemPool_t pool = nullptr;
    {
      std::lock_guard<std::mutex> lock(g_pool_mu);
      pool = g_pools[d];
    }
    if (!pool) continue;
    DeviceScope scope;
    RT_TRY(scope.enter(d));
    RT_CUDA(cudaDeviceSynchronize());  // frees queued on any stream have happened
    RT_CUDA(cudaMemPoolTrimTo(pool, 0));
  }
  return RT_OK;
}

extern "C" int rt_sort_pairs(void* keys, uint32_t* vals, uint64_t n, uint32_t key_bits, float* ms) {
  if (key_bits != 30 && key_bits != 63) return fail(RT_ERR_INVALID, "rt_sort_pairs: key_bits");
  if (n == 0) return RT_OK;
  if (!keys || !vals) return fail(RT_ERR_INVALID, "rt_sort_pairs: NULL buffer");
  RT_TRY(ensure_device());
  if (!classify(keys, nullptr) || !classify(vals, nullptr))
    return fail(RT_ERR_INVALID, "rt_sort_pairs: keys and vals must be device memory");
  int dev = 0;
  RT_CUDA(cudaGetDevice(&dev));
  cudaStream_t s = 0;
  TempPool tmp(s);
  const size_t ksz = key_size(key_bits);
  SortWorkspace ws;
  ws.vals = vals;
  if (key_bits == 30) {
    RT_TRY(tmp.alloc(&ws.rec_a, n * 8));
    RT_TRY(tmp.alloc(&ws.rec_b, n * 8));
  } else {
    RT_TRY(tmp.alloc(&ws.keys_alt, n * ksz));
    RT_TRY(tmp.alloc((void**)&ws.vals_alt, n * 4));
  }
  RT_TRY(tmp.alloc((void**)&ws.hist, sort_hist_words(key_bits) * 4));
  ws.chunk_words = sort_chunk_words(n, key_bits);
  RT_TRY(tmp.alloc((void**)&ws.chunk_hist, ws.chunk_words * 4));
  cudaEvent_t e0, e1;
  RT_CUDA(cudaEventCreate(&e0));
  RT_CUDA(cudaEventCreate(&e1));
  RT_CUDA(cudaEventRecord(e0, s));
  void* ko = nullptr;
  uint32_t* vo = nullptr;
  int rc = launch_sort(keys, n, key_bits, ws, /*iota_payload=*/false, &ko, &vo, s);
  cudaEventRecord(e1, s);
  cudaError_t e = cudaStreamSynchronize(s);
  float t = 0.0f;
  cudaEventElapsedTime(&t, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (rc != 0) return rc;
  RT_CUDA(e);
  if (ko != keys) {  // odd number of passes: bring the result back to the caller's buffers
    RT_CUDA(cudaMemcpyAsync(keys, ko, n * ksz, cudaMemcpyDeviceToDevice, s));
    RT_CUDA(cudaMemcpyAsync(vals, vo, n * 4, cudaMemcpyDeviceToDevice, s));
    RT_CUDA(cudaStreamSynchronize(s));
  }
  if (ms) *ms = t;
  return RT_OK;
}

extern "C" const char* rt_version(void) { return "gpu_raytracer_b200 0.1.0 (sm_100a)"; }
