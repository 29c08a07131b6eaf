// common.cuh — shared device/host helpers of the B200 LBVH ray-tracing core.
//
// Floating-point contract: every translation unit is compiled
// with -fmad=false, so nvcc never contracts a*b+c; the only fused operations are
// the explicit fmaf() calls.  Division and sqrt are IEEE (nvcc defaults
// -prec-div=true -prec-sqrt=true, -ftz=false).  This is what makes the kernels
// bit-comparable with a -ffp-contract=off CPU build of the same arithmetic.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace rt {

constexpr int kNumSMsB200 = 148;

// ---- error plumbing (host) -------------------------------------------------
int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what, const char* file, int line);
extern std::atomic<uint64_t> g_kernel_launches;

#define RT_CUDA(expr)                                                         \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return ::rt::fail_cuda(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define RT_TRY(expr)          \
  do {                        \
    int _rc = (expr);         \
    if (_rc != 0) return _rc; \
  } while (0)

// counts the launch (rt_kernel_launches) and checks the launch status
#define RT_LAUNCHED()                                         \
  do {                                                        \
    ::rt::g_kernel_launches.fetch_add(1, std::memory_order_relaxed); \
    RT_CUDA(cudaGetLastError());                              \
  } while (0)

// ---- programmatic dependent launch (kernel -> kernel edges of the build) -----
// Every kernel of the build chain starts with pdl_enter() and is launched with the
// programmatic-stream-serialization attribute: the launch of kernel k+1 is staged while
// kernel k still runs, and its CTAs start as soon as k has completed instead of one
// launch latency later.  griddepcontrol.wait returns when the previous kernel has
// completed and its writes are visible, and it precedes the first memory access, so the
// data flow is that of plain stream order.  Measured (tools/tune_pdl.py, B200): 1M
// triangles 0.239 -> 0.213 ms, 10M 1.017 -> 0.994 ms, 1K 0.112 -> 0.088 ms.  Triggering
// the dependents early (griddepcontrol.launch_dependents at kernel entry) was measured
// too and is not used: the waiting CTAs of the next kernels take SM slots from the
// running one (1M +6 %, 10M +12 %).  Without the launch attribute the wait is a no-op.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
extern std::atomic<int> g_pdl;  // rt_set_param("build.pdl"): 1 = launch the build chain with the attribute
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = g_pdl.load(std::memory_order_relaxed) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- counter-based random numbers -----------------------------------------
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint64_t stream) {
  return mix64(seed * 0xD6E8FEB86659FD93ull + stream);
}
__host__ __device__ __forceinline__ uint64_t rnd64(uint64_t key, uint64_t idx, uint32_t j) {
  return mix64(key ^ (idx * 256ull + j));
}
__host__ __device__ __forceinline__ float u01(uint64_t h) {
  return (float)(uint32_t)(h >> 40) * 0x1p-24f;
}
enum : uint64_t { kStreamSoup = 1, kStreamTerrain = 2, kStreamRays = 3, kStreamAO = 4 };

#ifdef __CUDACC__
// ---- total-order float keys: -0 < +0, exact, order-free min/max -----------
__device__ __forceinline__ uint32_t ord_key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord_unkey(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}
__device__ __forceinline__ float omin(float a, float b) { return ord_key(a) < ord_key(b) ? a : b; }
__device__ __forceinline__ float omax(float a, float b) { return ord_key(a) > ord_key(b) ? a : b; }

// 32 bytes in one store (STG.E.256, sm_100): a full sector per request where two
// 16-byte stores are two partial-sector writes.  p must be 32-byte aligned.
__device__ __forceinline__ void st_v8(void* p, const float4& a, const float4& b) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
               : "memory");
}

__device__ __forceinline__ void normalize3(float& x, float& y, float& z) {
  float l2 = x * x + y * y + z * z;  // -fmad=false: ((x*x + y*y) + z*z)
  float l = sqrtf(l2);
  x = x / l;
  y = y / l;
  z = z / l;
}

// uniform direction on the sphere, Marsaglia rejection; draws j0.. per attempt
__device__ __forceinline__ void sphere_sample(uint64_t key, uint64_t idx, uint32_t j0, float& ox,
                                              float& oy, float& oz) {
  for (uint32_t a = 0; a < 16; ++a) {
    float x1 = u01(rnd64(key, idx, j0 + 2 * a)) * 2.0f - 1.0f;
    float x2 = u01(rnd64(key, idx, j0 + 2 * a + 1)) * 2.0f - 1.0f;
    float s = x1 * x1 + x2 * x2;
    if (s < 1.0f) {
      float r = sqrtf(1.0f - s);
      ox = 2.0f * x1 * r;
      oy = 2.0f * x2 * r;
      oz = 1.0f - 2.0f * s;
      return;
    }
  }
  ox = 0.0f;
  oy = 0.0f;
  oz = 1.0f;
}

// streaming loads/stores that do not pollute L1
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
#endif  // __CUDACC__

}  // namespace rt
