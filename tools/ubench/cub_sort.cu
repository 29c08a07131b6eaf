// Third-party yardstick for the sort stage (development tool, not part
// of the product): cub::DeviceRadixSort::SortPairs on the same (key, u32 payload) pairs the
// build's rt_sort_pairs sorts.  CUB ships with the CUDA toolkit; nothing of it is linked into
// libgpu_raytracer_b200.so.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cub_sort cub_sort.cu
//   ./cub_sort            one JSON line per (n, key width): median ms of 9 runs after 3 warm-ups
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>

#include <cub/cub.cuh>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__host__ __device__ inline uint64_t mix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
template <typename K>
__global__ void fill(K* keys, uint32_t* vals, uint64_t n, int bits) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    keys[i] = (K)(mix64(i) >> (64 - bits));
    vals[i] = (uint32_t)i;
  }
}

template <typename K>
int run(uint64_t n, int bits) {
  K *k0, *k1;
  uint32_t *v0, *v1;
  CK(cudaMalloc(&k0, n * sizeof(K)));
  CK(cudaMalloc(&k1, n * sizeof(K)));
  CK(cudaMalloc(&v0, n * 4));
  CK(cudaMalloc(&v1, n * 4));
  size_t tmp_bytes = 0;
  CK((cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, v1, (int64_t)n, 0, bits)));
  void* tmp;
  CK(cudaMalloc(&tmp, tmp_bytes));
  void* flush;
  CK(cudaMalloc(&flush, 256u << 20));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  std::vector<float> ms;
  for (int it = 0; it < 12; ++it) {
    fill<K><<<148 * 8, 256>>>(k0, v0, n, bits);
    CK(cudaMemsetAsync(flush, it, 256u << 20));  // L2 flush
    CK(cudaEventRecord(e0));
    CK((cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k0, k1, v0, v1, (int64_t)n, 0, bits)));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float t;
    CK(cudaEventElapsedTime(&t, e0, e1));
    if (it >= 3) ms.push_back(t);
  }
  // sortedness of the result
  std::vector<K> h(std::min<uint64_t>(n, 1u << 20));
  CK(cudaMemcpy(h.data(), k1, h.size() * sizeof(K), cudaMemcpyDeviceToHost));
  const bool sorted = std::is_sorted(h.begin(), h.end());
  std::sort(ms.begin(), ms.end());
  const double med = ms[ms.size() / 2];
  const double alg = (bits == 30 ? 68.0 : 200.0) * (double)n;  // SURVEY.md §8(d) bytes of the stage
  printf("{\"impl\": \"cub::DeviceRadixSort::SortPairs\", \"n\": %llu, \"key_bits\": %d, \"ms\": %.4f, \"min_ms\": %.4f, "
         "\"gpairs_s\": %.2f, \"alg_gbs\": %.1f, \"sorted\": %s, \"temp_bytes\": %zu}\n",
         (unsigned long long)n, bits, med, ms[0], n / med / 1e6, alg / med / 1e6, sorted ? "true" : "false", tmp_bytes);
  fflush(stdout);
  cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1); cudaFree(tmp); cudaFree(flush);
  return 0;
}

int main() {
  const uint64_t sizes[] = {1000000ull, 9979112ull, 100026368ull};
  for (uint64_t n : sizes) {
    if (run<uint32_t>(n, 30)) return 1;
    if (run<uint64_t>(n, 63)) return 1;
  }
  return 0;
}
