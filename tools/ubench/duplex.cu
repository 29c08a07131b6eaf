// Microbenchmark (development tool): host<->device copy bandwidth, each direction alone and
// both at once, at the e2e bench's sizes (rays up : hits down = 2 : 1).
#include <cstdio>
#include <cuda_runtime.h>
int main() {
  const size_t up = 256u << 20, down = 128u << 20;
  char *h_up, *h_down, *d_up, *d_down;
  cudaMallocHost(&h_up, up); cudaMallocHost(&h_down, down);
  cudaMalloc(&d_up, up); cudaMalloc(&d_down, down);
  cudaStream_t a, b;
  cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
  cudaEvent_t e0, e1, f0, f1;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&f0); cudaEventCreate(&f1);
  for (int rep = 0; rep < 3; ++rep) {
    float ms_u, ms_d, ms_u2, ms_d2;
    cudaEventRecord(e0, a); cudaMemcpyAsync(d_up, h_up, up, cudaMemcpyHostToDevice, a); cudaEventRecord(e1, a);
    cudaDeviceSynchronize(); cudaEventElapsedTime(&ms_u, e0, e1);
    cudaEventRecord(f0, b); cudaMemcpyAsync(h_down, d_down, down, cudaMemcpyDeviceToHost, b); cudaEventRecord(f1, b);
    cudaDeviceSynchronize(); cudaEventElapsedTime(&ms_d, f0, f1);
    cudaEventRecord(e0, a); cudaEventRecord(f0, b);
    cudaMemcpyAsync(d_up, h_up, up, cudaMemcpyHostToDevice, a);
    cudaMemcpyAsync(h_down, d_down, down, cudaMemcpyDeviceToHost, b);
    cudaEventRecord(e1, a); cudaEventRecord(f1, b);
    cudaDeviceSynchronize(); cudaEventElapsedTime(&ms_u2, e0, e1); cudaEventElapsedTime(&ms_d2, f0, f1);
    printf("alone: up %.2f GB/s (%.3f ms) down %.2f GB/s (%.3f ms) | together: up %.2f GB/s (%.3f ms) down %.2f GB/s (%.3f ms)\n",
           up / ms_u / 1e6, ms_u, down / ms_d / 1e6, ms_d, up / ms_u2 / 1e6, ms_u2, down / ms_d2 / 1e6, ms_d2);
  }
  return 0;
}
