// Profiling tool: how long does the FIRST global load after griddepcontrol.wait take when it reads
// what the previous kernel of a CUDA-graph + PDL chain just wrote -- on a quiet GPU, and while
// another stream keeps HBM busy the way weight / KV prefetch does inside a decode step?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 profiles/dep_load_latency.cu -o profiles/dep_load_latency
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// 64 CTAs x 192 threads: one float4 column per thread of a [64][768] fp32 matrix (LayerNorm-like)
__global__ void producer(const float4* in, float4* out, unsigned long long* stamp) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const unsigned long long t0 = gtime();
  float4 v = in[blockIdx.x * 192 + threadIdx.x];
  v.x += 1.f;
  const unsigned long long t1 = gtime();
  out[blockIdx.x * 192 + threadIdx.x] = v;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    stamp[0] += t1 - t0;
    stamp[1] += 1;
  }
}

// G CTAs x 128 threads: every thread reads `per_thread` float4 of the producer's output (GEMM-like:
// each CTA needs a slab of all rows)
__global__ void consumer(const float4* in, float4* out, int per_thread, unsigned long long* stamp) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const unsigned long long t0 = gtime();
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const int n = 64 * 192;
  int idx = (blockIdx.x * 37 + threadIdx.x) % n;
  for (int i = 0; i < per_thread; ++i) {
    const float4 v = __ldcg(in + idx);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    idx += 128;
    if (idx >= n) idx -= n;
  }
  const unsigned long long t1 = gtime();
  if (blockIdx.x < 64 && threadIdx.x < 192) out[blockIdx.x * 192 + threadIdx.x] = acc;  // feeds the next producer
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    stamp[2] += t1 - t0;
    stamp[3] += 1;
  }
}

__global__ void hbm_stream(const float4* src, float4* dst, size_t n, int rounds) {
  for (int r = 0; r < rounds; ++r) {
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
      dst[i] = src[i];
    }
  }
}

template <typename K, typename... A>
static void launch_pdl(K k, int grid, int block, cudaStream_t s, A... a) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, k, a...);
}

int main() {
  float4 *a, *b, *big0, *big1;
  unsigned long long* stamp;
  cudaMalloc(&a, 64 * 192 * 16); cudaMalloc(&b, 64 * 192 * 16);
  cudaMemset(a, 0, 64 * 192 * 16); cudaMemset(b, 0, 64 * 192 * 16);
  const size_t big_n = (size_t)(512 << 20) / 16;
  cudaMalloc(&big0, big_n * 16); cudaMalloc(&big1, big_n * 16);
  cudaMemset(big0, 0, big_n * 16);
  cudaMalloc(&stamp, 4 * sizeof(unsigned long long));
  cudaStream_t s, bg; cudaStreamCreate(&s); cudaStreamCreateWithFlags(&bg, cudaStreamNonBlocking);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int pairs = 48;
  for (int bg_ctas : {0, 16, 64}) for (int g : {24, 72, 144}) for (int per_thread : {1, 4, 12}) {
    cudaGraph_t gr; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
    for (int i = 0; i < pairs; ++i) {
      launch_pdl(producer, 64, 192, s, (const float4*)a, b, stamp);
      launch_pdl(consumer, g, 128, s, (const float4*)b, a, per_thread, stamp);
    }
    cudaStreamEndCapture(s, &gr);
    cudaGraphInstantiate(&ge, gr, 0);
    cudaGraphLaunch(ge, s); cudaStreamSynchronize(s);
    cudaMemset(stamp, 0, 4 * sizeof(unsigned long long));
    if (bg_ctas > 0) hbm_stream<<<bg_ctas, 256, 0, bg>>>(big0, big1, big_n, 40);
    cudaEventRecord(e0, s);
    for (int r = 0; r < 10; ++r) cudaGraphLaunch(ge, s);
    cudaEventRecord(e1, s); cudaStreamSynchronize(s);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaDeviceSynchronize();
    unsigned long long h[4];
    cudaMemcpy(h, stamp, sizeof(h), cudaMemcpyDeviceToHost);
    printf("background CTAs %2d | consumer grid %3d x %2d float4/thread: %.2f us per kernel; first load after the wait: "
           "producer %.2f us, consumer %.2f us\n",
           bg_ctas, g, per_thread, ms * 1e3 / (10 * 2 * pairs), h[0] / (double)h[1] / 1e3, h[2] / (double)h[3] / 1e3);
    cudaGraphExecDestroy(ge); cudaGraphDestroy(gr);
  }
  return 0;
}
