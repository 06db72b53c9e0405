// Profiling tool: what bandwidth can a grid shaped like decode attention (768 CTAs x 128 threads,
// 6 per SM, 32 KB of shared memory each) pull from HBM when it reads K/V the way the paged cache
// is laid out -- 2 KB slabs (16 tokens x 128 B per page and head) scattered over a large pool --
// compared with one contiguous range per CTA? Loads go through cp.async into shared memory in
// 32 KB rounds, like the kernel's staging, with no arithmetic at all.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 profiles/kv_stream_bw.cu -o profiles/kv_stream_bw
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>

__device__ __forceinline__ void cp_async_16(void* s, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(s)), "l"(g) : "memory");
}

// slabs: per CTA `n_slabs` slab indices (2 KB units) into pool; mode 0: one 32 KB round at a time
// (issue, wait, next); mode 1: two 16 KB halves, the next half requested before waiting.
__global__ void __launch_bounds__(128, 6) stream_kernel(const uint4* pool, const int* slabs, int n_slabs, int mode,
                                                        float* sink) {
  __shared__ uint4 buf[2048];  // 32 KB
  const int* my = slabs + (size_t)blockIdx.x * n_slabs;
  float acc = 0.f;
  if (mode == 0) {
    for (int s0 = 0; s0 < n_slabs; s0 += 16) {  // 16 slabs = 32 KB
      for (int i = threadIdx.x; i < 2048; i += 128) {
        const int sl = s0 + (i >> 7);
        if (sl < n_slabs) cp_async_16(&buf[i], pool + (size_t)my[sl] * 128 + (i & 127));
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      acc += reinterpret_cast<float*>(buf)[threadIdx.x];
      __syncthreads();
    }
  } else {
    auto issue = [&](int half) {
      const int s0 = half * 8;
      for (int i = threadIdx.x; i < 1024; i += 128) {
        const int sl = s0 + (i >> 7);
        if (sl < n_slabs) cp_async_16(&buf[(half & 1) * 1024 + i], pool + (size_t)my[sl] * 128 + (i & 127));
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const int halves = (n_slabs + 7) / 8;
    issue(0);
    issue(1);
    for (int h = 0; h < halves; ++h) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      __syncthreads();
      acc += reinterpret_cast<float*>(buf)[(h & 1) * 4096 + threadIdx.x];
      __syncthreads();
      issue(h + 2);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  if (acc == 123.456f) sink[0] = acc;
}

int main() {
  const size_t pool_slabs = (size_t)(4096ull << 20) / 2048;  // 4 GB pool
  uint4* pool; cudaMalloc(&pool, pool_slabs * 2048); cudaMemset(pool, 0, pool_slabs * 2048);
  float* sink; cudaMalloc(&sink, 4);
  const int ctas = 768;
  std::mt19937_64 rng(1);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int n_slabs : {18, 68, 136}) {          // 144, 544, 1088 tokens of K+V per (row, head)
    for (int layout = 0; layout < 3; ++layout) {
      // 0: contiguous per CTA; 1: 2 KB slabs scattered; 2: 4 KB pairs (K and V slab adjacent) scattered
      const int reps = 8;  // different data every rep so nothing is served from L2
      std::vector<int> h((size_t)reps * ctas * n_slabs);
      std::vector<int> perm(pool_slabs / 2);
      for (size_t i = 0; i < perm.size(); ++i) perm[i] = (int)i;
      std::shuffle(perm.begin(), perm.end(), rng);
      size_t next = 0;
      for (int r = 0; r < reps; ++r) for (int c = 0; c < ctas; ++c) for (int s = 0; s < n_slabs; ++s) {
        int v;
        if (layout == 0) v = (int)(((size_t)(r * ctas + c) * n_slabs + s) % pool_slabs);
        else if (layout == 1) v = perm[next++ % perm.size()] * 2;
        else { v = perm[(next / 2) % perm.size()] * 2 + (int)(next & 1); next++; }
        h[((size_t)r * ctas + c) * n_slabs + s] = v;
      }
      int* d; cudaMalloc(&d, h.size() * 4); cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
      for (int mode = 0; mode < 2; ++mode) {
        stream_kernel<<<ctas, 128>>>(pool, d, n_slabs, mode, sink);
        cudaDeviceSynchronize();
        float best = 1e9;
        for (int r = 0; r < reps; ++r) {
          cudaEventRecord(e0);
          stream_kernel<<<ctas, 128>>>(pool, d + (size_t)r * ctas * n_slabs, n_slabs, mode, sink);
          cudaEventRecord(e1); cudaEventSynchronize(e1);
          float ms; cudaEventElapsedTime(&ms, e0, e1);
          best = std::min(best, ms);
        }
        const double bytes = (double)ctas * n_slabs * 2048;
        printf("%4d tokens/CTA  layout %s  %s: %6.1f MB in %6.2f us = %5.2f TB/s\n", n_slabs * 8,
               layout == 0 ? "contiguous   " : layout == 1 ? "2 KB scattered" : "4 KB scattered",
               mode == 0 ? "32 KB rounds    " : "2 x 16 KB ring  ", bytes / 1e6, best * 1e3, bytes / best / 1e9);
      }
      cudaFree(d);
    }
  }
  return 0;
}
