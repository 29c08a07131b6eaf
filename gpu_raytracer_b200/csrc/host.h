// host.h — host-runtime internals shared by api.cu and multi.cu (not part of the ABI).
#pragma once

#include <mutex>
#include <vector>

#include "../../include/gpu_raytracer.h"
#include "common.cuh"
#include "kernels.h"

namespace rt {

int ensure_device();
// the library's own stream-ordered pool of `device` (created on first use)
int device_pool(int device, cudaMemPool_t* out);
// stream-ordered allocation from the library's pool of the CURRENT device
int pool_alloc(void** out, size_t bytes, cudaStream_t s);
// 0 = host (or unregistered), 1 = device/managed memory; *dev = owning device
int classify(const void* p, int* dev);

struct DeviceScope {
  int prev = -1;
  bool switched = false;
  int enter(int device) {
    RT_CUDA(cudaGetDevice(&prev));
    if (device >= 0 && device != prev) {
      RT_CUDA(cudaSetDevice(device));
      switched = true;
    }
    return 0;
  }
  ~DeviceScope() {
    if (switched) cudaSetDevice(prev);
  }
};

// stream-ordered temporary allocations, released on scope exit
struct TempPool {
  cudaStream_t s;
  std::vector<void*> ptrs;
  explicit TempPool(cudaStream_t stream) : s(stream) {}
  int alloc(void** out, size_t bytes) {
    RT_TRY(pool_alloc(out, bytes, s));
    ptrs.push_back(*out);
    return 0;
  }
  ~TempPool() {
    for (void* p : ptrs) cudaFreeAsync(p, s);
  }
};

// timing events that do not outlive an early return
struct EventSet {
  cudaEvent_t ev[RT_STAGE_COUNT + 1];
  int count = 0;
  int create(int n) {
    for (; count < n; ++count) RT_CUDA(cudaEventCreate(&ev[count]));
    return 0;
  }
  ~EventSet() {
    for (int i = 0; i < count; ++i) cudaEventDestroy(ev[i]);
  }
};

}  // namespace rt

struct rt_bvh {
  uint64_t n = 0;
  uint32_t key_bits = 30;
  int device = 0;
  rt::Node* nodes = nullptr;
  rt::SortedTri* stris = nullptr;
  rt::BvhHeader* header = nullptr;
  // parent words: derived from the nodes on first use (refit_bvh, bvh_read_topology/boxes)
  mutable uint32_t* parent = nullptr;
  mutable std::mutex parent_mu;
  void* keys = nullptr;  // only with RT_BUILD_KEEP_KEYS
  rt::Node4* wide = nullptr;  // only with RT_BUILD_WIDE4 / bvh_widen: n-1 slots
  uint32_t wide_root = 0;     // reference the BVH4 traversal starts from
  bool is_wide = false;
  uint32_t wide_leaf_max = rt::kWideLeafDefault;
  rt::Wide8 wide8;            // only with RT_BUILD_WIDE8 / bvh_widen8: compressed 8-wide layout
  rt::Compact compact;        // only with RT_BUILD_COMPACT / bvh_compact: 32-byte binary nodes
  float bounds[6] = {0, 0, 0, 0, 0, 0};
  int root = 0;  // node id where traversal starts (header.root_node)
  float times[RT_STAGE_COUNT] = {0, 0, 0, 0, 0, 0, 0};
  bool has_times = false;
  uint64_t bytes = 0;
};

namespace rt {
// traversal of whichever layout the handle holds (api.cu)
int launch_trace_h(const rt_bvh* h, const float4* rays, uint64_t n_rays, uint32_t flags,
                   float4* hits, unsigned long long* counter, cudaStream_t s,
                   unsigned long long* stats = nullptr);
// receiver side of a copy made inside this process (build_bvh_multi): takes the host-side
// fields (root, bounds, key width) from the handle whose buffers were copied; no validation
int adopt_committed(rt_bvh* dst, const rt_bvh* src);
// device-buffer trace on stream s, RT_TRACE_REORDER honoured (api.cu)
int trace_device(const rt_bvh* h, const float4* rays, uint64_t n_rays, uint32_t flags, float4* hits,
                 cudaStream_t s);
}  // namespace rt
