// bench_main.cpp — the "small C++ bench main" of SURVEY.md §8(b): a host program with no
// Python and no PyTorch in it that drives the C ABI of include/gpu_raytracer.h the way a C++
// renderer would: scene -> build_bvh -> rays -> trace_rays -> free_bvh, everything through
// host (malloc'd, then page-locked by the CUDA runtime) or device pointers of its own.
//
//   bench_main [soup N | flake DEPTH SEG | cornell] [width height] [reps] [cam=12 floats, comma separated]
//
// Prints one JSON line: triangles, build ms (bvh_build_times) and Mtri/s, rays, trace ms
// (host clock around trace_rays on DEVICE buffers, i.e. the kernels plus one synchronise) and
// Mrays/s, hit count, and a checksum of the hit records that tests/ compare with the Python
// path's.  Build: g++ -std=c++17 -I include bench_main.cpp libgpu_raytracer_b200.so -lcudart
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "gpu_raytracer.h"

#define CHECK(call)                                                          \
  do {                                                                       \
    const int rc_ = (call);                                                  \
    if (rc_ != RT_OK) {                                                      \
      std::fprintf(stderr, "%s: %d %s\n", #call, rc_, bvh_last_error());     \
      return 1;                                                              \
    }                                                                        \
  } while (0)

static void normalise(float* v) {
  const float l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  for (int i = 0; i < 3; ++i) v[i] /= l;
}

// the 12-float pinhole camera of rt_gen_primary_rays looking at the centre of `box` (any
// camera will do here; the Python package's camera.frame_box is the reference one)
static void frame_box(const float* box, uint32_t w, uint32_t h, float* cam) {
  float c[3], d[3] = {0.3f, 0.25f, -1.0f}, up[3] = {0, 1, 0};
  float r = 0;
  for (int i = 0; i < 3; ++i) {
    c[i] = 0.5f * (box[i] + box[3 + i]);
    r += (box[3 + i] - box[i]) * (box[3 + i] - box[i]);
  }
  r = 0.5f * std::sqrt(r);
  normalise(d);
  const float half = 20.0f * 3.14159265f / 180.0f, dist = r / std::sin(half);
  float eye[3], f[3], s[3], u[3];
  for (int i = 0; i < 3; ++i) {
    eye[i] = c[i] + d[i] * dist;
    f[i] = -d[i];
  }
  s[0] = f[1] * up[2] - f[2] * up[1]; s[1] = f[2] * up[0] - f[0] * up[2]; s[2] = f[0] * up[1] - f[1] * up[0];
  normalise(s);
  u[0] = s[1] * f[2] - s[2] * f[1]; u[1] = s[2] * f[0] - s[0] * f[2]; u[2] = s[0] * f[1] - s[1] * f[0];
  const float th = std::tan(half), tw = th * (float)w / (float)h;
  for (int i = 0; i < 3; ++i) {
    cam[i] = eye[i];
    cam[3 + i] = eye[i] + f[i] - tw * s[i] + th * u[i];  // top-left corner of the image plane at distance 1
    cam[6 + i] = 2.0f * tw * s[i] / (float)w;       // per pixel to the right
    cam[9 + i] = -2.0f * th * u[i] / (float)h;      // per pixel down
  }
}

int main(int argc, char** argv) {
  int kind = RT_SCENE_SOUP;
  uint64_t p0 = 1000000, p1 = 0, seed = 1;
  int a = 1;
  if (a < argc && !std::strcmp(argv[a], "soup")) { kind = RT_SCENE_SOUP; p0 = std::strtoull(argv[a + 1], nullptr, 10); a += 2; }
  else if (a < argc && !std::strcmp(argv[a], "flake")) { kind = RT_SCENE_SPHEREFLAKE; p0 = std::strtoull(argv[a + 1], nullptr, 10); p1 = std::strtoull(argv[a + 2], nullptr, 10); seed = 2; a += 3; }
  else if (a < argc && !std::strcmp(argv[a], "cornell")) { kind = RT_SCENE_CORNELL; p0 = p1 = seed = 0; a += 1; }
  const uint32_t width = a < argc ? (uint32_t)std::atoi(argv[a]) : 1920, height = a + 1 < argc ? (uint32_t)std::atoi(argv[a + 1]) : 1080;
  const int reps = a + 2 < argc ? std::atoi(argv[a + 2]) : 5;
  float cam[12];
  bool have_cam = false;
  for (int i = 1; i < argc; ++i)
    if (!std::strncmp(argv[i], "cam=", 4)) {
      const char* q = argv[i] + 4;
      for (int k = 0; k < 12; ++k) {
        char* end = nullptr;
        cam[k] = std::strtof(q, &end);
        q = (*end == ',') ? end + 1 : end;
      }
      have_cam = true;
    }

  if (rt_device_count() == 0) {
    std::printf("{\"error\": \"no CUDA device\"}\n");
    return 0;
  }
  const uint64_t n_tris = rt_scene_size(kind, p0, p1), n_rays = (uint64_t)width * height;
  float *d_tris = nullptr, *d_rays = nullptr, *d_hits = nullptr;
  if (cudaMalloc((void**)&d_tris, n_tris * 48) != cudaSuccess || cudaMalloc((void**)&d_rays, n_rays * 32) != cudaSuccess ||
      cudaMalloc((void**)&d_hits, n_rays * 16) != cudaSuccess) {
    std::fprintf(stderr, "cudaMalloc failed\n");
    return 1;
  }
  CHECK(rt_gen_scene(kind, p0, p1, seed, 0, n_tris, d_tris));

  rt_build_opts opts;
  std::memset(&opts, 0, sizeof(opts));
  opts.key_bits = 30;
  opts.flags = RT_BUILD_TIMINGS;
  opts.device = -1;
  float best_build = 1e30f, ms[RT_STAGE_COUNT];
  rt_bvh* h = nullptr;
  for (int i = 0; i < reps + 2; ++i) {  // the first builds fill the library's memory pool
    if (h) free_bvh(h);
    h = nullptr;
    CHECK(build_bvh(d_tris, n_tris, &opts, &h));
    CHECK(bvh_build_times(h, ms));
    if (i >= 2 && ms[RT_STAGE_TOTAL] < best_build) best_build = ms[RT_STAGE_TOTAL];
  }
  rt_bvh_info info;
  CHECK(bvh_info(h, &info));
  if (!have_cam) frame_box(info.bounds, width, height, cam);
  CHECK(rt_gen_primary_rays(cam, width, height, 0, n_rays, d_rays));

  double best_trace = 1e30;
  for (int i = 0; i < reps + 2; ++i) {
    cudaDeviceSynchronize();
    const auto t0 = std::chrono::steady_clock::now();
    CHECK(trace_rays(h, d_rays, n_rays, RT_TRACE_CLOSEST, d_hits));  // synchronous on return
    const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (i >= 2 && t < best_trace) best_trace = t;
  }
  std::vector<float> hits(n_rays * 4);
  cudaMemcpy(hits.data(), d_hits, n_rays * 16, cudaMemcpyDeviceToHost);
  uint64_t n_hit = 0, sum = 1469598103934665603ull;  // FNV-1a over the hit records' bytes
  const unsigned char* bytes = reinterpret_cast<const unsigned char*>(hits.data());
  for (uint64_t i = 0; i < n_rays * 16; ++i) sum = (sum ^ bytes[i]) * 1099511628211ull;
  for (uint64_t i = 0; i < n_rays; ++i) {
    int32_t prim;
    std::memcpy(&prim, &hits[4 * i + 3], 4);
    n_hit += prim >= 0;
  }
  // the same batch from pageable host memory: the library stages it (H2D, trace, D2H)
  std::vector<float> h_rays(n_rays * 8), h_hits(n_rays * 4);
  cudaMemcpy(h_rays.data(), d_rays, n_rays * 32, cudaMemcpyDeviceToHost);
  CHECK(trace_rays(h, h_rays.data(), n_rays, RT_TRACE_CLOSEST, h_hits.data()));
  const bool same = std::memcmp(h_hits.data(), hits.data(), n_rays * 16) == 0;

  std::printf("{\"n_tris\": %llu, \"build_ms\": %.4f, \"build_mtri_s\": %.1f, \"n_rays\": %llu, \"trace_ms\": %.4f, "
              "\"trace_mrays_s\": %.1f, \"hits\": %llu, \"hits_fnv1a\": \"%016llx\", \"host_path_equal\": %s, "
              "\"kernel_launches\": %llu, \"version\": \"%s\"}\n",
              (unsigned long long)n_tris, best_build, n_tris / best_build / 1e3, (unsigned long long)n_rays, best_trace,
              n_rays / best_trace / 1e3, (unsigned long long)n_hit, (unsigned long long)sum, same ? "true" : "false",
              (unsigned long long)rt_kernel_launches(), rt_version());
  free_bvh(h);
  cudaFree(d_tris);
  cudaFree(d_rays);
  cudaFree(d_hits);
  return same ? 0 : 1;
}
