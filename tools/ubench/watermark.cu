// Microbenchmark (development tool): when does a running kernel see small host-to-device
// copies that are queued between large ones on another stream?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void watcher(const volatile unsigned long long* wm, unsigned long long* seen_ns, int n) {
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  int next = 1;
  while (next <= n) {
    if (*wm >= (unsigned long long)next) {
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      seen_ns[next - 1] = t - t0;
      ++next;
    } else {
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      if (t - t0 > 2000000000ull) break;
      __nanosleep(500);
    }
  }
}
int main() {
  const int n = 8;
  const size_t big = 8u << 20;
  char *h_big, *d_big;
  unsigned long long *h_marks, *d_wm, *d_seen, h_seen[n];
  cudaMallocHost(&h_big, big * n);
  cudaMallocHost(&h_marks, n * 8);
  cudaMalloc(&d_big, big * n);
  cudaMalloc(&d_wm, 8);
  cudaMalloc(&d_seen, n * 8);
  cudaStream_t s_in, s_run;
  cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&s_run, cudaStreamNonBlocking);
  for (int mode = 0; mode < 2; ++mode) {
    cudaMemset(d_wm, 0, 8);
    cudaMemset(d_seen, 0, n * 8);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    if (mode == 0) watcher<<<1, 1, 0, s_run>>>(d_wm, d_seen, n);  // kernel first, copies after
    cudaEventRecord(e0, s_in);
    for (int c = 0; c < n; ++c) {
      cudaMemcpyAsync(d_big + c * big, h_big + c * big, big, cudaMemcpyHostToDevice, s_in);
      h_marks[c] = c + 1;
      cudaMemcpyAsync(d_wm, &h_marks[c], 8, cudaMemcpyHostToDevice, s_in);
    }
    cudaEventRecord(e1, s_in);
    if (mode == 1) watcher<<<1, 1, 0, s_run>>>(d_wm, d_seen, n);  // copies first, kernel after
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(h_seen, d_seen, n * 8, cudaMemcpyDeviceToHost);
    printf("mode %d (%s): upload %.3f ms; watermark k seen at (us):", mode,
           mode == 0 ? "kernel launched first" : "copies queued first", ms);
    for (int c = 0; c < n; ++c) printf(" %.0f", h_seen[c] / 1e3);
    printf("  (%s)\n", cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
