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
  if (strategy == RT_MULTI_REDUNDANT || n_devices == 1) {
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
