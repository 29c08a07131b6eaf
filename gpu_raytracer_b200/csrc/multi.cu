// multi.cu — multi-GPU entry points of the C ABI (SURVEY.md §8e and §8b "Multi-GPU entry
// takes a device list"): one process, one host thread per device.
//
// The path shards by rays: a batch is cut into contiguous ranges on 32-ray boundaries
// (rt_shard_range, the rule gpu_raytracer_b200/sharding.py uses for one process per GPU),
// every device traces its range against its own copy of the BVH, and no device talks to
// another while tracing.  The copies are either built redundantly (every device runs
// build_bvh on the triangles) or built once and copied: nodes, reordered triangles and
// header go from the first device to the others with cudaMemcpyPeerAsync, i.e. over NVLink
// where the devices are peers.  There is no collective because there is no exchange step.
//
// The handles are ordinary rt_bvh handles: everything else in the ABI works on them.
#include <cstring>
#include <string>
#include <thread>

#include "host.h"

using namespace rt;

extern "C" int rt_shard_range(uint64_t n, int rank, int world, uint64_t* first, uint64_t* count) {
  if (world <= 0 || rank < 0 || rank >= world || !first || !count)
    return fail(RT_ERR_INVALID, "rt_shard_range: bad rank/world or NULL output");
  const uint64_t align = 32;
  const uint64_t units = (n + align - 1) / align;
  const uint64_t base = units / (uint64_t)world, rem = units % (uint64_t)world;
  const uint64_t r = (uint64_t)rank;
  const uint64_t u0 = r * base + (r < rem ? r : rem);
  const uint64_t u1 = u0 + base + (r < rem ? 1 : 0);
  const uint64_t f = u0 * align < n ? u0 * align : n;
  const uint64_t l = u1 * align < n ? u1 * align : n;
  *first = f;
  *count = l - f;
  return RT_OK;
}

namespace {

// runs fn(i) on one host thread per device; the first failure's code and message are
// re-raised on the calling thread (bvh_last_error is thread-local)
template <typename Fn>
int for_each_device(int n, Fn fn) {
  std::vector<int> rc((size_t)n, 0);
  std::vector<std::string> msg((size_t)n);
  auto body = [&](int i) {
    rc[(size_t)i] = fn(i);
    if (rc[(size_t)i] != 0) msg[(size_t)i] = bvh_last_error();
  };
  if (n == 1) {
    body(0);
  } else {
    std::vector<std::thread> th;
    th.reserve((size_t)n);
    for (int i = 0; i < n; ++i) th.emplace_back(body, i);
    for (auto& t : th) t.join();
  }
  for (int i = 0; i < n; ++i)
    if (rc[(size_t)i] != 0) return fail(rc[(size_t)i], "device slot %d: %s", i, msg[(size_t)i].c_str());
  return RT_OK;
}

template <typename Fn, typename Pre>
int for_each_device(int n, Fn fn, Pre pre) {  // pre(i) runs for every slot on the calling thread first
  for (int i = 0; i < n; ++i) RT_TRY(pre(i));
  return for_each_device(n, fn);
}

// lets `device` read memory of the library's pool on `owner` directly (peer copies then go
// over NVLink instead of bouncing through the host); failure only costs bandwidth
void allow_peer(int owner, int device) {
  if (owner == device) return;
  int can = 0;
  if (cudaDeviceCanAccessPeer(&can, device, owner) != cudaSuccess || !can) {
    cudaGetLastError();
    return;
  }
  cudaMemPool_t pool;
  if (device_pool(owner, &pool) == 0) {
    cudaMemAccessDesc d;
    memset(&d, 0, sizeof(d));
    d.location.type = cudaMemLocationTypeDevice;
    d.location.id = device;
    d.flags = cudaMemAccessFlagsProtReadWrite;
    cudaMemPoolSetAccess(pool, &d, 1);
  }
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(device);
  cudaDeviceEnablePeerAccess(owner, 0);
  if (prev >= 0) cudaSetDevice(prev);
  cudaGetLastError();
}

int check_devices(const int* devices, int n_devices, const char* who) {
  if (!devices || n_devices <= 0) return fail(RT_ERR_INVALID, "%s: empty device list", who);
  RT_TRY(ensure_device());
  const int have = rt_device_count();
  for (int i = 0; i < n_devices; ++i)
    if (devices[i] < 0 || devices[i] >= have)
      return fail(RT_ERR_INVALID, "%s: device %d (slot %d) does not exist (%d visible)", who, devices[i], i,
                  have);
  return 0;
}

}  // namespace

namespace {

// device buffers of several devices that do not outlive an early return
struct PartBuffers {
  std::vector<std::pair<int, void*>> held;  // (device, pointer)
  int alloc(int device, void** out, size_t bytes, cudaStream_t s) {
    RT_TRY(pool_alloc(out, bytes, s));  // the caller has made `device` current
    held.emplace_back(device, *out);
    return 0;
  }
  ~PartBuffers() {
    int prev = -1;
    cudaGetDevice(&prev);
    for (auto& e : held) {
      cudaSetDevice(e.first);
      cudaFreeAsync(e.second, 0);
    }
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// can every device of the list read every other one's memory directly?
bool all_peers(const int* devices, int n) {
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      const int a = devices[i], b = devices[j];
      if (a == b) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess || !can) {
        cudaGetLastError();
        return false;
      }
    }
  return true;
}
bool all_peers(rt_bvh* const* handles, int n) {
  std::vector<int> d((size_t)n);
  for (int i = 0; i < n; ++i) d[(size_t)i] = handles[i]->device;
  return all_peers(d.data(), n);
}

}  // namespace

extern "C" int build_bvh_multi(const float* tris, uint64_t n_tris, const rt_build_opts* opts,
                               const int* devices, int n_devices, int strategy, rt_bvh** out) {
  if (!out) return fail(RT_ERR_INVALID, "build_bvh_multi: out is NULL");
  for (int i = 0; i < (n_devices > 0 ? n_devices : 0); ++i) out[i] = nullptr;
  RT_TRY(check_devices(devices, n_devices, "build_bvh_multi"));
  if (strategy != RT_MULTI_REDUNDANT && strategy != RT_MULTI_BROADCAST)
    return fail(RT_ERR_INVALID, "build_bvh_multi: unknown strategy %d", strategy);
  if (opts && opts->stream)
    return fail(RT_ERR_INVALID, "build_bvh_multi: a stream belongs to one device; leave opts->stream NULL");
  int src_dev = -1;
  const bool tris_on_device = n_tris != 0 && classify(tris, &src_dev) != 0;

  auto build_on = [&](int slot, const float* t) -> int {
    rt_build_opts o;
    memset(&o, 0, sizeof(o));
    if (opts) o = *opts;
    o.device = devices[slot];
    o.stream = nullptr;
    return build_bvh(t, n_tris, &o, &out[slot]);
  };
  // triangles that live on one device, as seen from another: a peer copy into a temporary
  auto build_from_peer = [&](int slot) -> int {
    const int dev = devices[slot];
    if (!tris_on_device || dev == src_dev) return build_on(slot, tris);
    DeviceScope scope;
    RT_TRY(scope.enter(dev));
    allow_peer(src_dev, dev);
    cudaStream_t s = 0;
    TempPool tmp(s);
    void* local = nullptr;
    RT_TRY(tmp.alloc(&local, n_tris * 48));
    RT_CUDA(cudaMemcpyPeerAsync(local, dev, tris, src_dev, n_tris * 48, s));
    const int rc = build_on(slot, (const float*)local);
    RT_CUDA(cudaStreamSynchronize(s));
    return rc;
  };

  int rc;
  if (strategy == RT_MULTI_REDUNDANT && n_devices > 1 && !tris_on_device && n_tris >= (1ull << 16) &&
      all_peers(devices, n_devices) && params().merge_fused != 0) {
    // host triangles, devices that see each other: every device takes 1/n of the array over its
    // own PCIe link and pulls the rest from its peers over NVLink (an all-gather over peer
    // memory), then builds from its device copy.  The array crosses PCIe once, not n times.
    PartBuffers bufs;
    std::vector<float*> d_tris((size_t)n_devices, nullptr);
    auto lo = [&](int i) { return n_tris * (uint64_t)i / (uint64_t)n_devices; };
    for (int i = 0; i < n_devices; ++i)
      for (int j = 0; j < n_devices; ++j) allow_peer(devices[i], devices[j]);
    rc = for_each_device(n_devices, [&](int i) -> int {
      DeviceScope scope;
      RT_TRY(scope.enter(devices[i]));
      const uint64_t f = lo(i), c = lo(i + 1) - f;
      if (c) RT_CUDA(cudaMemcpyAsync(d_tris[(size_t)i] + 12 * f, tris + 12 * f, c * 48, cudaMemcpyHostToDevice, 0));
      RT_CUDA(cudaStreamSynchronize(0));
      return 0;
    }, [&](int i) -> int {
      DeviceScope scope;
      RT_TRY(scope.enter(devices[i]));
      RT_TRY(bufs.alloc(devices[i], (void**)&d_tris[(size_t)i], n_tris * 48, 0));
      RT_CUDA(cudaStreamSynchronize(0));
      return 0;
    });
    if (rc == 0)
      rc = for_each_device(n_devices, [&](int i) -> int {
        DeviceScope scope;
        RT_TRY(scope.enter(devices[i]));
        for (int k = 0; k < n_devices; ++k) {
          const uint64_t f = lo(k), c = lo(k + 1) - f;
          if (k != i && c)
            RT_CUDA(cudaMemcpyPeerAsync(d_tris[(size_t)i] + 12 * f, devices[i], d_tris[(size_t)k] + 12 * f, devices[k],
                                        c * 48, 0));
        }
        RT_CUDA(cudaStreamSynchronize(0));
        return build_on(i, d_tris[(size_t)i]);
      });
  } else if (strategy == RT_MULTI_REDUNDANT || n_devices == 1) {
    rc = for_each_device(n_devices, build_from_peer);
  } else {
    rc = build_from_peer(0);
    if (rc == 0) {
      rt_bvh* src = out[0];
      rc = for_each_device(n_devices - 1, [&](int k) -> int {
        const int slot = k + 1, dev = devices[slot];
        DeviceScope scope;
        RT_TRY(scope.enter(dev));
        allow_peer(src->device, dev);
        RT_TRY(bvh_alloc_empty(src->n, src->key_bits, &out[slot]));
        void *sp[3], *dp[3];
        uint64_t sb[3], db[3];
        RT_TRY(bvh_device_buffers(src, sp, sb));
        RT_TRY(bvh_device_buffers(out[slot], dp, db));
        cudaStream_t s = 0;
        for (int b = 0; b < 3; ++b)
          if (sb[b]) RT_CUDA(cudaMemcpyPeerAsync(dp[b], dev, sp[b], src->device, sb[b], s));
        RT_CUDA(cudaStreamSynchronize(s));
        return adopt_committed(out[slot], src);
      });
    }
  }
  if (rc != 0) {
    std::string keep = bvh_last_error();
    for (int i = 0; i < n_devices; ++i) {
      free_bvh(out[i]);
      out[i] = nullptr;
    }
    return fail(rc, "%s", keep.c_str());
  }
  return RT_OK;
}

extern "C" int trace_rays_sharded(rt_bvh* const* handles, int n_devices, const float* rays,
                                  uint64_t n_rays, uint32_t flags, float* hits) {
  if (!handles || n_devices <= 0) return fail(RT_ERR_INVALID, "trace_rays_sharded: empty handle list");
  for (int i = 0; i < n_devices; ++i)
    if (!handles[i]) return fail(RT_ERR_INVALID, "trace_rays_sharded: handle %d is NULL", i);
  if (n_rays == 0) return RT_OK;
  if (!rays || !hits) return fail(RT_ERR_INVALID, "trace_rays_sharded: rays or hits is NULL");
  if (classify(rays, nullptr) || classify(hits, nullptr))
    return fail(RT_ERR_INVALID, "trace_rays_sharded: rays and hits are host buffers (one batch for all devices)");
  return for_each_device(n_devices, [&](int i) -> int {
    uint64_t first = 0, count = 0;
    RT_TRY(rt_shard_range(n_rays, i, n_devices, &first, &count));
    if (count == 0) return 0;
    return trace_rays(handles[i], rays + 8 * first, count, flags, hits + 4 * first);
  });
}

extern "C" int trace_random_sharded(rt_bvh* const* handles, int n_devices, const float* box,
                                    uint64_t seed, uint64_t first, uint64_t n_rays, uint32_t flags,
                                    uint64_t batch, float* hits, uint64_t* n_hit, float* trace_ms) {
  if (!handles || n_devices <= 0) return fail(RT_ERR_INVALID, "trace_random_sharded: empty handle list");
  for (int i = 0; i < n_devices; ++i)
    if (!handles[i]) return fail(RT_ERR_INVALID, "trace_random_sharded: handle %d is NULL", i);
  if (!box) return fail(RT_ERR_INVALID, "trace_random_sharded: box is NULL");
  if (flags > 7u) return fail(RT_ERR_INVALID, "trace_random_sharded: unknown flags 0x%x", flags);
  if (hits && classify(hits, nullptr))
    return fail(RT_ERR_INVALID, "trace_random_sharded: hits is a host buffer (or NULL)");
  if (batch == 0) batch = 1ull << 25;
  if (batch > (1ull << 28)) batch = 1ull << 28;
  return for_each_device(n_devices, [&](int i) -> int {
    const rt_bvh* h = handles[i];
    uint64_t f0 = 0, count = 0;
    RT_TRY(rt_shard_range(n_rays, i, n_devices, &f0, &count));
    if (n_hit) n_hit[i] = 0;
    if (trace_ms) trace_ms[i] = 0.0f;
    if (count == 0) return 0;
    DeviceScope scope;
    RT_TRY(scope.enter(h->device));
    cudaStream_t s = nullptr, s_out = nullptr;
    RT_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    struct Streams {
      cudaStream_t a, b;
      ~Streams() {
        if (a) cudaStreamDestroy(a);
        if (b) cudaStreamDestroy(b);
      }
    } streams{s, nullptr};
    if (hits) {
      RT_CUDA(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
      streams.b = s_out;
    }
    const uint64_t cap = count < batch ? count : batch;
    TempPool tmp(s);
    float4 *d_rays = nullptr, *d_hits[2] = {nullptr, nullptr};
    unsigned long long* d_count = nullptr;
    RT_TRY(tmp.alloc((void**)&d_rays, cap * 32));
    RT_TRY(tmp.alloc((void**)&d_hits[0], cap * 16));
    if (hits) RT_TRY(tmp.alloc((void**)&d_hits[1], cap * 16));
    RT_TRY(tmp.alloc((void**)&d_count, sizeof(unsigned long long)));
    RT_CUDA(cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), s));
    EventSet ev;
    RT_TRY(ev.create(6));  // 0/1: trace bracket; 2,3: slot traced; 4,5: slot drained
    float ms_total = 0.0f;
    uint64_t it = 0;
    for (uint64_t done = 0; done < count; done += cap, ++it) {
      const uint64_t c = count - done < cap ? count - done : cap;
      const int slot = hits ? (int)(it & 1) : 0;
      RT_TRY(launch_gen_random(box, seed, first + f0 + done, c, d_rays, s));
      if (hits && it >= 2) RT_CUDA(cudaStreamWaitEvent(s, ev.ev[4 + slot], 0));  // slot drained
      RT_CUDA(cudaEventRecord(ev.ev[0], s));
      RT_TRY(trace_device(h, d_rays, c, flags, d_hits[slot], s));
      RT_CUDA(cudaEventRecord(ev.ev[1], s));
      RT_TRY(launch_count_hits(d_hits[slot], c, d_count, s));
      if (hits) {
        RT_CUDA(cudaEventRecord(ev.ev[2 + slot], s));
        RT_CUDA(cudaStreamWaitEvent(s_out, ev.ev[2 + slot], 0));
        RT_CUDA(cudaMemcpyAsync(hits + 4 * (f0 + done), d_hits[slot], c * 16, cudaMemcpyDeviceToHost, s_out));
        RT_CUDA(cudaEventRecord(ev.ev[4 + slot], s_out));
      }
      // one batch at a time: the events are reused, and the generator must not overwrite rays
      // a running traversal still reads
      RT_CUDA(cudaStreamSynchronize(s));
      float ms = 0.0f;
      RT_CUDA(cudaEventElapsedTime(&ms, ev.ev[0], ev.ev[1]));
      ms_total += ms;
    }
    unsigned long long total = 0;
    RT_CUDA(cudaMemcpyAsync(&total, d_count, sizeof(total), cudaMemcpyDeviceToHost, s));
    RT_CUDA(cudaStreamSynchronize(s));
    if (s_out) RT_CUDA(cudaStreamSynchronize(s_out));
    if (n_hit) n_hit[i] = total;
    if (trace_ms) trace_ms[i] = ms_total;
    return 0;
  });
}

// ===========================================================================
// Sharded build: the SCENE is cut, not the rays.
//
// Device i builds a BVH over a contiguous range of the triangle array (1/N of the build work
// and of the memory per GPU: the way to scenes beyond one GPU's memory), every device traces
// all rays against its part, and the parts' hits are merged: per ray the minimum by (t,
// primitive index), the primitive index made global by adding the range's first triangle.
// This is exact, not approximate: the acceptance rule (trace_common.cuh accept_hit) makes the
// accepted hit of a ray min over {triangles with a consistent hit} by (t, prim), a function
// of the ray and the triangles alone, so the minimum over the parts' minima is what one BVH
// over all triangles returns, bit for bit (tests/test_round2_gpu.py).  Any-hit: occluded when
// any part is.  The merge is the path's one exchange step; between processes it is an
// all-to-all of hit records (gpu_raytracer_b200/sharding.py ShardedScene), inside one process
// peer copies to the first device.
// ===========================================================================
namespace {

constexpr int kMaxParts = 16;
struct MergeParts {
  const float4* part[kMaxParts];
  uint32_t base[kMaxParts];
  int n;
};

__global__ void __launch_bounds__(256)
merge_hits_kernel(MergeParts parts, uint64_t n, bool any, float4* __restrict__ out) {
  const float inf = __int_as_float(0x7f800000);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    float4 best = make_float4(inf, 0.0f, 0.0f, __int_as_float(-1));
    int best_prim = -1;
    for (int k = 0; k < parts.n; ++k) {
      const float4 h = parts.part[k][i];
      const int prim = __float_as_int(h.w);
      if (prim < 0) continue;
      if (any) {
        best = make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0x7FFFFFFF));
        best_prim = 0x7FFFFFFF;
        break;
      }
      const int gp = prim + (int)parts.base[k];
      if (best_prim < 0 || h.x < best.x || (h.x == best.x && gp < best_prim)) {
        best = make_float4(h.x, h.y, h.z, __int_as_float(gp));
        best_prim = gp;
      }
    }
    out[i] = best;
  }
}

}  // namespace

extern "C" int rt_merge_hits(const float* const* parts, const uint32_t* prim_base, int n_parts, uint64_t n,
                             uint32_t flags, float* out, void* stream) {
  if (!parts || n_parts <= 0 || n_parts > kMaxParts)
    return fail(RT_ERR_INVALID, "rt_merge_hits: 1..%d parts", kMaxParts);
  if (n == 0) return RT_OK;
  if (!out) return fail(RT_ERR_INVALID, "rt_merge_hits: out is NULL");
  RT_TRY(ensure_device());
  MergeParts mp;
  mp.n = n_parts;
  for (int k = 0; k < n_parts; ++k) {
    if (!parts[k] || !classify(parts[k], nullptr)) return fail(RT_ERR_INVALID, "rt_merge_hits: part %d is not device memory", k);
    mp.part[k] = reinterpret_cast<const float4*>(parts[k]);
    mp.base[k] = prim_base ? prim_base[k] : 0u;
  }
  if (!classify(out, nullptr)) return fail(RT_ERR_INVALID, "rt_merge_hits: out is not device memory");
  uint64_t blocks = (n + 255) / 256;
  const uint64_t cap = (uint64_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  merge_hits_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(mp, n, (flags & RT_TRACE_ANY) != 0,
                                                                       reinterpret_cast<float4*>(out));
  RT_LAUNCHED();
  return RT_OK;
}

extern "C" int build_bvh_sharded(const float* tris, uint64_t n_tris, const rt_build_opts* opts,
                                 const int* devices, int n_devices, rt_bvh** out, uint32_t* prim_base) {
  if (!out || !prim_base) return fail(RT_ERR_INVALID, "build_bvh_sharded: out or prim_base is NULL");
  for (int i = 0; i < (n_devices > 0 ? n_devices : 0); ++i) out[i] = nullptr;
  RT_TRY(check_devices(devices, n_devices, "build_bvh_sharded"));
  if (n_devices > kMaxParts) return fail(RT_ERR_LIMIT, "build_bvh_sharded: at most %d parts", kMaxParts);
  if (n_tris > 0x7FFFFFFEull) return fail(RT_ERR_LIMIT, "build_bvh_sharded: global primitive indices are 31-bit");
  if (n_tris && !tris) return fail(RT_ERR_INVALID, "build_bvh_sharded: tris is NULL");
  if (n_tris && classify(tris, nullptr))
    return fail(RT_ERR_INVALID, "build_bvh_sharded: tris is a host array (each device uploads its own range)");
  if (opts && opts->stream) return fail(RT_ERR_INVALID, "build_bvh_sharded: leave opts->stream NULL");
  for (int i = 0; i < n_devices; ++i) prim_base[i] = (uint32_t)(n_tris * (uint64_t)i / (uint64_t)n_devices);
  const int rc = for_each_device(n_devices, [&](int i) -> int {
    const uint64_t first = prim_base[i];
    const uint64_t last = n_tris * (uint64_t)(i + 1) / (uint64_t)n_devices;
    rt_build_opts o;
    memset(&o, 0, sizeof(o));
    if (opts) o = *opts;
    o.device = devices[i];
    o.stream = nullptr;
    return build_bvh(tris + 12 * first, last - first, &o, &out[i]);
  });
  if (rc != 0) {
    std::string keep = bvh_last_error();
    for (int i = 0; i < n_devices; ++i) {
      free_bvh(out[i]);
      out[i] = nullptr;
    }
    return fail(rc, "%s", keep.c_str());
  }
  return RT_OK;
}

extern "C" int trace_rays_merged(rt_bvh* const* handles, const uint32_t* prim_base, int n_devices,
                                 const float* rays, uint64_t n_rays, uint32_t flags, float* hits) {
  if (!handles || n_devices <= 0 || n_devices > kMaxParts)
    return fail(RT_ERR_INVALID, "trace_rays_merged: 1..%d handles", kMaxParts);
  for (int i = 0; i < n_devices; ++i)
    if (!handles[i]) return fail(RT_ERR_INVALID, "trace_rays_merged: handle %d is NULL", i);
  if (n_rays == 0) return RT_OK;
  if (!rays || !hits) return fail(RT_ERR_INVALID, "trace_rays_merged: rays or hits is NULL");
  if (flags > 7u) return fail(RT_ERR_INVALID, "trace_rays_merged: unknown flags 0x%x", flags);
  if (classify(rays, nullptr) || classify(hits, nullptr))
    return fail(RT_ERR_INVALID, "trace_rays_merged: rays and hits are host buffers");
  const int dev0 = handles[0]->device;
  const uint64_t kBatch = 1ull << 22;  // rays per round: 128 MiB up, 64 MiB of hits per part
  // With peer access between all the devices nothing crosses PCIe twice and nothing is staged:
  // device j owns slice j of the batch (rt_shard_range).  It uploads that slice of the rays, all
  // devices pull each other's slices over NVLink (an all-gather over peer memory), every part
  // traces the whole batch, and device j's merge kernel LOADS the candidates of its slice from
  // every part's memory (a reduce-scatter by (t, prim) over peer memory, exchange and merge in one
  // kernel), then sends its slice to the host over its own PCIe link.
  // Without peer access: the parts' hits are copied to the first device and merged there.
  const bool fused = all_peers(handles, n_devices) && params().merge_fused != 0;
  if (fused)
    for (int i = 0; i < n_devices; ++i)
      for (int j = 0; j < n_devices; ++j) allow_peer(handles[i]->device, handles[j]->device);
  for (uint64_t first = 0; first < n_rays; first += kBatch) {
    const uint64_t cnt = n_rays - first < kBatch ? n_rays - first : kBatch;
    if (fused) {
      PartBuffers bufs;
      float4* part_hits[kMaxParts] = {};
      float4* part_rays[kMaxParts] = {};
      uint64_t sf[kMaxParts], sc[kMaxParts];  // slice j of the batch
      for (int j = 0; j < n_devices; ++j) RT_TRY(rt_shard_range(cnt, j, n_devices, &sf[j], &sc[j]));
      // phase 0: device i takes slice i of the rays from the host (1/n of the batch over its own
      // PCIe link, not the whole batch over every link)
      RT_TRY(for_each_device(n_devices, [&](int i) -> int {
        if (sc[i] == 0) return 0;
        DeviceScope scope;
        RT_TRY(scope.enter(handles[i]->device));
        RT_CUDA(cudaMemcpyAsync(part_rays[i] + 2 * sf[i], rays + 8 * (first + sf[i]), sc[i] * 32, cudaMemcpyHostToDevice, 0));
        RT_CUDA(cudaStreamSynchronize(0));
        return 0;
      }, [&](int i) -> int {  // before the threads start: the buffers (PartBuffers is not thread-safe)
        DeviceScope scope;
        RT_TRY(scope.enter(handles[i]->device));
        RT_TRY(bufs.alloc(handles[i]->device, (void**)&part_hits[i], cnt * 16, 0));
        RT_TRY(bufs.alloc(handles[i]->device, (void**)&part_rays[i], cnt * 32, 0));
        RT_CUDA(cudaStreamSynchronize(0));
        return 0;
      }));
      // phase 1: every part pulls the other slices from their owners (NVLink: an all-gather over
      // peer memory) and traces the whole batch into its own memory
      RT_TRY(for_each_device(n_devices, [&](int i) -> int {
        const rt_bvh* h = handles[i];
        DeviceScope scope;
        RT_TRY(scope.enter(h->device));
        cudaStream_t s = 0;
        for (int k = 0; k < n_devices; ++k)
          if (k != i && sc[k] != 0)
            RT_CUDA(cudaMemcpyPeerAsync(part_rays[i] + 2 * sf[k], h->device, part_rays[k] + 2 * sf[k],
                                        handles[k]->device, sc[k] * 32, s));
        RT_TRY(trace_device(h, part_rays[i], cnt, flags, part_hits[i], s));
        RT_CUDA(cudaStreamSynchronize(s));
        return 0;
      }));
      // phase 2: device j merges slice j out of everybody's memory and sends it down
      RT_TRY(for_each_device(n_devices, [&](int j) -> int {
        const uint64_t f = sf[j], c = sc[j];
        if (c == 0) return 0;
        DeviceScope scope;
        RT_TRY(scope.enter(handles[j]->device));
        cudaStream_t s = 0;
        TempPool tmp(s);
        float4* merged = nullptr;
        RT_TRY(tmp.alloc((void**)&merged, c * 16));
        const float* parts[kMaxParts];
        for (int k = 0; k < n_devices; ++k) parts[k] = reinterpret_cast<const float*>(part_hits[k] + f);
        RT_TRY(rt_merge_hits(parts, prim_base, n_devices, c, flags, reinterpret_cast<float*>(merged), s));
        RT_CUDA(cudaMemcpyAsync(hits + 4 * (first + f), merged, c * 16, cudaMemcpyDeviceToHost, s));
        RT_CUDA(cudaStreamSynchronize(s));
        return 0;
      }));
      continue;
    }
    // the parts' hits, gathered on the first handle's device
    DeviceScope scope0;
    RT_TRY(scope0.enter(dev0));
    cudaStream_t s0 = 0;
    TempPool tmp0(s0);
    float4* gathered = nullptr;
    float4* merged = nullptr;
    RT_TRY(tmp0.alloc((void**)&gathered, (size_t)n_devices * cnt * 16));
    RT_TRY(tmp0.alloc((void**)&merged, cnt * 16));
    RT_CUDA(cudaStreamSynchronize(s0));
    RT_TRY(for_each_device(n_devices, [&](int i) -> int {
      const rt_bvh* h = handles[i];
      DeviceScope scope;
      RT_TRY(scope.enter(h->device));
      cudaStream_t s = 0;
      TempPool tmp(s);
      float4 *d_rays = nullptr, *d_hits = nullptr;
      RT_TRY(tmp.alloc((void**)&d_rays, cnt * 32));
      RT_CUDA(cudaMemcpyAsync(d_rays, rays + 8 * first, cnt * 32, cudaMemcpyHostToDevice, s));
      float4* dst = gathered + (size_t)i * cnt;
      if (h->device == dev0) {
        RT_TRY(trace_device(h, d_rays, cnt, flags, dst, s));
      } else {
        RT_TRY(tmp.alloc((void**)&d_hits, cnt * 16));
        RT_TRY(trace_device(h, d_rays, cnt, flags, d_hits, s));
        allow_peer(dev0, h->device);
        RT_CUDA(cudaMemcpyPeerAsync(dst, dev0, d_hits, h->device, cnt * 16, s));
      }
      RT_CUDA(cudaStreamSynchronize(s));
      return 0;
    }));
    const float* parts[kMaxParts];
    for (int i = 0; i < n_devices; ++i) parts[i] = reinterpret_cast<const float*>(gathered + (size_t)i * cnt);
    RT_TRY(rt_merge_hits(parts, prim_base, n_devices, cnt, flags, reinterpret_cast<float*>(merged), s0));
    RT_CUDA(cudaMemcpyAsync(hits + 4 * first, merged, cnt * 16, cudaMemcpyDeviceToHost, s0));
    RT_CUDA(cudaStreamSynchronize(s0));
  }
  return RT_OK;
}
