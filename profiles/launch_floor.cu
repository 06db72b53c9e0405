// Profiling tool: per-kernel floor of a dependent chain of tiny kernels on this box,
// plain stream launches vs CUDA graph, with and without programmatic dependent launch.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void tiny_smem(float* x, int pdl) {
  __shared__ float buf[35 * 256];  // 35 KB like decode attention
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  buf[threadIdx.x] = x[blockIdx.x * blockDim.x + threadIdx.x];
  __syncthreads();
  x[blockIdx.x * blockDim.x + threadIdx.x] = buf[threadIdx.x ^ 1] + 1.0f;
}

__global__ void tiny(float* x, int pdl) {
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  x[blockIdx.x * blockDim.x + threadIdx.x] += 1.0f;
}

static void launch(float* x, int ctas, int pdl, cudaStream_t s) {
  if (ctas >= 256) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(128); cfg.stream = s;
    cudaLaunchAttribute a[1];
    a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    a[0].val.programmaticStreamSerializationAllowed = pdl;
    cfg.attrs = a; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, tiny_smem, x, pdl);
    return;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(128); cfg.stream = s;
  cudaLaunchAttribute a[1];
  a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  a[0].val.programmaticStreamSerializationAllowed = pdl;
  cfg.attrs = a; cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, tiny, x, pdl);
}

int main() {
  float* x; cudaMalloc(&x, 1 << 20); cudaMemset(x, 0, 1 << 20);
  cudaStream_t s; cudaStreamCreate(&s);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int N = 88;
  for (int ctas : {64, 148, 384, 768, 1536}) for (int pdl : {0, 1}) {
    for (int i = 0; i < N; ++i) launch(x, ctas, pdl, s);
    cudaStreamSynchronize(s);
    cudaEventRecord(e0, s);
    for (int r = 0; r < 10; ++r) for (int i = 0; i < N; ++i) launch(x, ctas, pdl, s);
    cudaEventRecord(e1, s); cudaStreamSynchronize(s);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("stream  ctas=%3d pdl=%d: %.2f us/kernel\n", ctas, pdl, ms * 1e3 / (10 * N));
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < N; ++i) launch(x, ctas, pdl, s);
    cudaStreamEndCapture(s, &g);
    cudaGraphInstantiate(&ge, g, 0);
    cudaGraphLaunch(ge, s); cudaStreamSynchronize(s);
    cudaEventRecord(e0, s);
    for (int r = 0; r < 20; ++r) cudaGraphLaunch(ge, s);
    cudaEventRecord(e1, s); cudaStreamSynchronize(s);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("graph   ctas=%3d pdl=%d: %.2f us/kernel\n", ctas, pdl, ms * 1e3 / (20 * N));
  }
  return 0;
}
