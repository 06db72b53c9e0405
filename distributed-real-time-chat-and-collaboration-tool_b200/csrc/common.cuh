// Shared device/host helpers for the B200 (sm_100a) generation engine.
//
// Everything here is plain CUDA + inline PTX: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (UMMA + TMEM) wrappers, warp reductions, bf16 packing, and the
// status/last-error plumbing the C ABI uses (include/b200llm.h).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace b200llm {

// ---------------------------------------------------------------------------------------
// Host-side status plumbing. The C ABI never throws: every entry point returns an int and
// stores a human-readable reason retrievable through b200llm_last_error().
// ---------------------------------------------------------------------------------------
enum Status : int {
  kOk = 0,
  kErrInvalidArg = 1,
  kErrCuda = 2,
  kErrOutOfMemory = 3,
  kErrCapacity = 4,
  kErrUnsupported = 5,
};

void set_last_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define B200_CUDA_OK(expr)                                                                   \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      return ::b200llm::fail(::b200llm::kErrCuda, std::string(#expr) + ": " +                \
                                                      cudaGetErrorString(_e));               \
    }                                                                                        \
  } while (0)

#define B200_CHECK_ARG(cond, msg)                                                            \
  do {                                                                                       \
    if (!(cond)) return ::b200llm::fail(::b200llm::kErrInvalidArg, std::string(msg));        \
  } while (0)

#define B200_PROPAGATE(expr)                                                                 \
  do {                                                                                       \
    int _s = (expr);                                                                         \
    if (_s != 0) return _s;                                                                  \
  } while (0)

#ifdef __CUDACC__

// ---------------------------------------------------------------------------------------
// Small device utilities
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// gelu_new: 0.5 x (1 + tanh( sqrt(2/pi) (x + 0.044715 x^3) ))
// (transformers/activations.py:59-66, NewGELUActivation). tanh via exp so that the result
// stays within ~1e-6 of torch.tanh in fp32.
__device__ __forceinline__ float gelu_new(float x) {
  const float k0 = 0.7978845608028654f;  // sqrt(2/pi)
  const float k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  // tanh(u) = 1 - 2/(exp(2u)+1); clamp keeps exp finite.
  float uu = fminf(fmaxf(u, -15.0f), 15.0f);
  float e = __expf(2.0f * uu);
  float t = 1.0f - __fdividef(2.0f, e + 1.0f);
  return 0.5f * x * (1.0f + t);
}

// Same function through the hardware tanh (MUFU.TANH, rel. error 2^-11): used where the result is
// rounded to bf16 anyway (2^-9) and the exact version made the epilogue issue-bound.
__device__ __forceinline__ float gelu_new_fast(float x) {
  const float k0 = 0.7978845608028654f;
  const float k1 = 0.044715f * 0.7978845608028654f;
  const float u = x * fmaf(k1, x * x, k0);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
}

// Streaming 16-byte global load that does not allocate in L1 (KV pages, weights read once).
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// ---------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t"
      "}\n"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}

// Bounded wait: a protocol bug turns into a trap (a CUDA error the host reports) instead of
// a hung GPU. The bound is far beyond any legitimate wait (seconds).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > 200000000u) {
      __trap();
    }
  }
}

// ---------------------------------------------------------------------------------------
// TMA (bulk tensor copy global -> shared, completion on an mbarrier)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(desc)) : "memory");
}

// L2 cache-policy constants (same encodings CUTLASS uses for TMA::CacheHintSm90).
constexpr uint64_t kEvictNormal = 0x1000000000000000ull;
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull;
constexpr uint64_t kEvictLast = 0x14F0000000000000ull;

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* desc,
                                            uint64_t* bar, int32_t c0, int32_t c1,
                                            uint64_t cache_hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      :
      : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)),
        "r"(c0), "r"(c1), "l"(cache_hint)
      : "memory");
}

// ---------------------------------------------------------------------------------------
// tcgen05: TMEM allocation, UMMA issue, commit, TMEM loads
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate, one CTA.
__device__ __forceinline__ void umma_bf16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                             uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      :
      : "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has retired.
// (Implies tcgen05.fence::before_thread_sync.)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}

// Shared-memory matrix descriptor for a K-major bf16 tile written by TMA with SWIZZLE_128B:
// rows of 64 elements (128 B), 8-row swizzle atoms 1024 B apart (SBO), version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);  // start address   [0,14)
  d |= static_cast<uint64_t>(1) << 16;                      // LBO (unused for SW128 K-major)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;              // SBO = 1024 B    [32,46)
  d |= static_cast<uint64_t>(1) << 46;                      // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;                      // layout = SWIZZLE_128B
  return d;
}

// Instruction descriptor: kind::f16, A=B=bf16, D=fp32, both operands K-major, M x N tile.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(uint32_t m, uint32_t n) {
  return (1u << 4)            // c_format = F32
         | (1u << 7)          // a_format = BF16
         | (1u << 10)         // b_format = BF16
         | (0u << 15)         // a_major = K
         | (0u << 16)         // b_major = K
         | ((n >> 3) << 17)   // N / 8
         | ((m >> 4) << 24);  // M / 16
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Programmatic dependent launch: wait for the upstream grid's memory to be visible /
// allow the downstream grid to start its prologue.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---------------------------------------------------------------------------------------
// Profiling builds only (-DB200_STEP_TRACE, build.py --trace): block 0 / thread 0 of every
// kernel records %globaltimer at entry and exit into a shared buffer, in launch order, so
// the real timeline of a CUDA-graph + PDL decode step can be read back without a profiler.
// ---------------------------------------------------------------------------------------
#ifdef B200_STEP_TRACE
struct StepTraceSink {
  unsigned long long* buf;  // [16 * capacity] records
  unsigned int* idx;
  unsigned int capacity;
};
static __device__ StepTraceSink g_step_trace;  // one copy per translation unit, same target
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Record layout (16 x u64): [0] entry, [1] exit, [2] tag, [4 + i] milestone i (B200_STEP_MARK).
#define B200_STEP_BEGIN(tag)                                                                  \
  __shared__ unsigned int _st_slot_sh;                                                        \
  unsigned int _st_slot = 0xffffffffu;                                                        \
  if (threadIdx.x == 0) {                                                                     \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && g_step_trace.buf != nullptr) { \
      _st_slot = atomicAdd(g_step_trace.idx, 1u);                                             \
      if (_st_slot < g_step_trace.capacity) {                                                 \
        g_step_trace.buf[16 * _st_slot] = global_timer_ns();                                  \
        g_step_trace.buf[16 * _st_slot + 2] = (tag);                                          \
      }                                                                                       \
    }                                                                                         \
    _st_slot_sh = _st_slot;                                                                   \
  }
#define B200_STEP_END()                                                                       \
  if (_st_slot < g_step_trace.capacity) g_step_trace.buf[16 * _st_slot + 1] = global_timer_ns();
// any thread of block 0, after a __syncthreads() that follows B200_STEP_BEGIN
#define B200_STEP_MARK(i)                                                                     \
  do {                                                                                        \
    const unsigned int _s = _st_slot_sh;                                                      \
    if (_s < g_step_trace.capacity) g_step_trace.buf[16 * _s + 4 + (i)] = global_timer_ns();  \
  } while (0)
#define B200_DEFINE_STEP_TRACE_SETTER(fn)                                                     \
  void fn(unsigned long long* buf, unsigned int* idx, unsigned int capacity) {                \
    StepTraceSink s{buf, idx, capacity};                                                      \
    cudaMemcpyToSymbol(g_step_trace, &s, sizeof(s));                                          \
  }
#else
#define B200_STEP_BEGIN(tag)
#define B200_STEP_END()
#define B200_STEP_MARK(i) \
  do {                    \
  } while (0)
#define B200_DEFINE_STEP_TRACE_SETTER(fn)
#endif

#endif  // __CUDACC__

}  // namespace b200llm
