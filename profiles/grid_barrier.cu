// Profiling tool: cost of a software grid barrier across one CTA per SM (cooperative launch),
// the building block of a persistent decode-step kernel.
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ void grid_barrier(unsigned int* counter, unsigned int target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    unsigned int v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
    } while (v < target);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(384, 1) bar_kernel(unsigned int* counter, int n, float* sink, int mode) {
  float acc = 0.f;
  if (mode == 0) {
    for (int i = 0; i < n; ++i) grid_barrier(counter, (i + 1) * gridDim.x);
  } else {
    cg::grid_group g = cg::this_grid();
    for (int i = 0; i < n; ++i) g.sync();
  }
  if (threadIdx.x == 0 && blockIdx.x == 0) sink[0] = acc;
}

int main() {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  unsigned int* counter; float* sink;
  cudaMalloc(&counter, 4); cudaMalloc(&sink, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) for (int n : {1, 101, 1001}) {
    cudaMemset(counter, 0, 4);
    void* args[] = {&counter, &n, &sink, &mode};
    cudaEventRecord(e0);
    cudaError_t err = cudaLaunchCooperativeKernel((void*)bar_kernel, dim3(sms), dim3(384), args, 0, 0);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%s n=%4d: %.2f us total, %s\n", mode ? "cg::grid.sync" : "atomic+spin  ", n, ms * 1e3, cudaGetErrorString(err));
  }
  return 0;
}
