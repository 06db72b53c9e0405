// Profiling tool (not part of the library): cycles per tcgen05.mma (bf16, M=128, K=16, operands in
// shared memory) as a function of N, for a chain of accumulations into one TMEM accumulator.
// Decode GEMMs run with N = 16..64 tokens; this shows what one k-block (4 instructions) costs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I<pkg>/csrc profiles/umma_rate.cu -o profiles/umma_rate
#include <cstdio>
#include <cuda.h>
#include <cuda_bf16.h>
#include <string>

#include "common.cuh"

namespace b200llm {
static thread_local std::string g_err;
void set_last_error(const std::string& m) { g_err = m; }
int fail(int code, const std::string& m) { g_err = m; return code; }
}  // namespace b200llm
using namespace b200llm;

template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(unsigned long long* out, int reps, int kblocks, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~static_cast<uintptr_t>(1023u));
  __shared__ uint64_t bar;
  __shared__ uint64_t stage_bar[16], ready_bar[16];
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 16; ++i) {
      mbar_init(&stage_bar[i], 1);
      mbar_init(&ready_bar[i], 1);
      mbar_arrive(&ready_bar[i]);  // phase 0 already complete: what a landed k-block looks like
    }
    fence_barrier_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(&tmem_slot, 256);
    tmem_relinquish();
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t da = umma_desc_k_sw128(smem_u32(smem));
    const uint64_t db = umma_desc_k_sw128(smem_u32(smem + 16384));
    uint32_t phase = 0;
    for (int r = 0; r < reps; ++r) {
      const unsigned long long t0 = clock64();
      for (int kb = 0; kb < kblocks; ++kb) {
        if (mode & 2) {  // wait on an (already complete) full barrier + fence, like the GEMM's k-loop
          mbar_wait(&ready_bar[kb & 15], 0);
          if (!(mode & 4)) tc_fence_after();
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, da + k * 2, db + k * 2, idesc, (kb > 0 || k > 0) ? 1u : 0u);
        if (mode & 1) umma_commit(&stage_bar[kb & 15]);  // per-k-block "slot free" commit
      }
      const unsigned long long t1 = clock64();
      umma_commit(&bar);
      mbar_wait(&bar, phase);
      phase ^= 1u;
      const unsigned long long t2 = clock64();
      out[2 * r] = t1 - t0;
      out[2 * r + 1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

template <int N>
__global__ void __launch_bounds__(128, 1) rate_kernel_converged(unsigned long long* out, int reps, int kblocks, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~static_cast<uintptr_t>(1023u));
  __shared__ uint64_t bar;
  __shared__ uint64_t stage_bar[16], ready_bar[16];
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    for (int i = 0; i < 16; ++i) {
      mbar_init(&stage_bar[i], 1);
      mbar_init(&ready_bar[i], 1);
      mbar_arrive(&ready_bar[i]);  // phase 0 already complete: what a landed k-block looks like
    }
    fence_barrier_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(&tmem_slot, 256);
    tmem_relinquish();
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x < 32) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t da = umma_desc_k_sw128(smem_u32(smem));
    const uint64_t db = umma_desc_k_sw128(smem_u32(smem + 16384));
    uint32_t phase = 0;
    for (int r = 0; r < reps; ++r) {
      const unsigned long long t0 = clock64();
      for (int kb = 0; kb < kblocks; ++kb) {
        if (mode & 2) {  // wait on an (already complete) full barrier + fence, like the GEMM's k-loop
          mbar_wait(&ready_bar[kb & 15], 0);
          if (!(mode & 4)) tc_fence_after();
        }
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, da + k * 2, db + k * 2, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          if (mode & 1) umma_commit(&stage_bar[kb & 15]);  // per-k-block "slot free" commit
        }
        __syncwarp();
      }
      const unsigned long long t1 = clock64();
      if (elect_one()) umma_commit(&bar);
      __syncwarp();
      mbar_wait(&bar, phase);
      phase ^= 1u;
      const unsigned long long t2 = clock64();
      if (threadIdx.x == 0) {
        out[2 * r] = t1 - t0;
        out[2 * r + 1] = t2 - t0;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

template <int N>
void run(unsigned long long* d) {
  cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int mode : {0, 3, 8, 11})
  for (int kblocks : {3, 8}) {
    if (mode & 8) {
      cudaFuncSetAttribute(rate_kernel_converged<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
      rate_kernel_converged<N><<<1, 128, 64 * 1024>>>(d, 4, kblocks, mode & 7);
    } else {
      rate_kernel<N><<<1, 128, 64 * 1024>>>(d, 4, kblocks, mode);
    }
    unsigned long long h[8];
    cudaError_t err = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    if (err != cudaSuccess) { printf("N=%d: %s\n", N, cudaGetErrorString(err)); return; }
    printf("mode %d N=%3d k-blocks=%2d (%2d mma): issue %5llu cyc, issue+complete %5llu cyc -> %.1f cyc per mma\n", mode, N, kblocks,
           4 * kblocks, h[6], h[7], (double)h[7] / (4 * kblocks));
  }
}

int main() {
  unsigned long long* d;
  cudaMalloc(&d, 64 * sizeof(unsigned long long));
  printf("mode bit 0: tcgen05.commit after every k-block; bit 1: mbarrier wait + fence before every k-block; bit 2: wait without the fence; bit 3: the whole warp runs the loop, an elected lane issues\n");
  run<64>(d);
  run<256>(d);
  return 0;
}
