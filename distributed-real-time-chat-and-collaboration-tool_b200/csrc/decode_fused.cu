// Decode-step GEMMs with their neighbours fused in, for batches of up to 64 live rows.
//
// A decode layer used to be eight dependent kernels (LayerNorm, QKV GEMM, attention, out-proj
// GEMM, LayerNorm, MLP-up GEMM, gelu-reduce, MLP-down GEMM) and each dependency hop costs ~1 us
// however little the kernel does. Two kernels here cut it to five:
//
//   ln_gemm_kernel     out = epilogue( LayerNorm(x) . W^T + b )       (QKV; MLP-up with gelu_new)
//       modeling_gpt2.py:273 + :185 (ln_1, c_attn) and :304 + :239-240 (ln_2, c_fc, act)
//       One CTA = 128 output features x 16 tokens. Sixteen warps first normalise one residual row
//       each (fp32 statistics, shifted second moment) and write it as the
//       bf16 B operand straight into shared memory in the SWIZZLE_128B K-major layout, so the
//       normalised activations never exist in global memory; the weight tiles stream through a
//       TMA ring that starts before the grid dependency resolves.
//
//   gemm_redadd_kernel x += act . W^T + b        (attention out-proj, MLP down-proj; :282, :307)
//       Split-K over gridDim.y with the partial tiles ADDED into the fp32 residual stream in L2
//       by TMA (cp.reduce.async.bulk.tensor ... add) instead of written as partial planes that the
//       next kernel reduces. Without planes the consumer (ln_gemm_kernel) reads one finished row
//       per token, and more K splits cost nothing downstream. The additions of one element arrive
//       in no fixed order: results can differ in the last fp32 bit from run to run
//       (decode_reduce = 1 keeps the ordered partial-plane path).
//
// Both: swap-AB orientation (weights on the 128 UMMA rows, tokens on UMMA N), tcgen05.mma with the
// accumulator in TMEM, weights by TMA; warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warps 4-7 = LayerNorm prologue / epilogue.
#include "decode_fused.cuh"

#include "common.cuh"
#include "gemm_tcgen05.cuh"

#ifdef B200_STEP_TRACE
#define B200_FTRACE(slot)                                             \
  do {                                                                \
    if (blockIdx.x == 0 && blockIdx.y == 0) B200_STEP_MARK(slot);     \
  } while (0)
#else
#define B200_FTRACE(slot) \
  do {                    \
  } while (0)
#endif

namespace b200llm {

namespace {

constexpr int kM = 128;        // features per CTA (UMMA M)
constexpr int kKB = 64;        // K elements per k-block (one 128-byte swizzle row)
constexpr int kABytes = kM * kKB * 2;
constexpr int kThreads = 256;
constexpr int kNT = 16;        // tokens per CTA of ln_gemm_kernel (smallest UMMA N for M = 128)
constexpr int kMaxF4 = 10;     // float4 per LayerNorm lane: d / 4 / 32 lanes per row -> d <= 1280

// ------------------------------------------------------------------------------------------
// ln_gemm_kernel
// ------------------------------------------------------------------------------------------
struct LnGemmKernelArgs {
  const float* x;       // [rows, d] fp32 residual stream (complete: nothing pending)
  const float* gamma;   // [d]
  const float* beta;    // [d]
  const float* bias;    // [features]
  __nv_bfloat16* out;   // [rows, ld_out]
  int rows;             // valid token rows
  int d;                // K
  int features;
  int ld_out;
  int stages;           // weight ring depth
  float eps;
  // K split over gridDim.z: this CTA multiplies k-blocks [z * kb_per_split, (z + 1) * kb_per_split) and
  // writes an fp32 partial plane (no bias, no activation) that the consumer reduces in a fixed order
  int kb_per_split;     // = d / 64 without a split
  float* partial_out;   // [gridDim.z][plane_rows, ld_out] or null
  int plane_rows;
};

// 4 role warps + 16 LayerNorm warps (one token row each; the first 4 also own a TMEM lane quarter
// in the epilogue, where the 4 warps of a quarter take 4 token columns each).
constexpr int kLnWarps = kNT;
constexpr int kLnThreads = 128 + kLnWarps * 32;

template <bool GELU>
__global__ void __launch_bounds__(kLnThreads, 1)
ln_gemm_kernel(const __grid_constant__ CUtensorMap map_w, const LnGemmKernelArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int n_kb = a.kb_per_split;        // k-blocks of this CTA
  const int kb0 = blockIdx.z * n_kb;      // its first k-block
  uint8_t* ring = smem;                                   // stages x 16 KB weight tiles
  uint8_t* bop = smem + a.stages * kABytes;               // n_kb x [16 tokens x 64] bf16 (2 KB each)
  float* gb = reinterpret_cast<float*>(bop + n_kb * kNT * 128);  // gamma[d] | beta[d]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(gb + 2 * a.d);
  uint64_t* empty_bar = full_bar + a.stages;
  uint64_t* b_ready = empty_bar + a.stages;
  uint64_t* tmem_full = b_ready + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_blk = blockIdx.x;           // feature tile
  const int tok0 = blockIdx.y * kNT;      // first token of this CTA
  // every k-block of the CTA resident at once: one barrier, one uninterrupted MMA chain (a wait on a
  // complete barrier costs the issuing thread ~200 cycles, a per-slot commit ~80)
  const bool single = a.stages == n_kb;
  B200_STEP_BEGIN(300 + (GELU ? 1 : 0));
  if ((smem_u32(smem) & 1023u) != 0) __trap();

  if (warp == 0 && lane == 0) tma_prefetch_desc(&map_w);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(b_ready, kLnWarps);  // one arrival per LayerNorm warp
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 32);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();  // the next kernel may start its own weight prefetch; it still waits for this grid

  if (warp == 0) {
    // ---- weight tiles: nothing here depends on the upstream kernel ----
    if (lane == 0) {
      if (single) mbar_arrive_expect_tx(&full_bar[0], static_cast<uint32_t>(n_kb) * kABytes);
      for (int kb = 0; kb < n_kb; ++kb) {
        const int s = kb % a.stages;
        if (!single) {
          if (kb >= a.stages) mbar_wait(&empty_bar[s], static_cast<uint32_t>(kb / a.stages - 1) & 1u);
          mbar_arrive_expect_tx(&full_bar[s], kABytes);
        }
        tma_load_2d(ring + s * kABytes, &map_w, &full_bar[single ? 0 : s], (kb0 + kb) * kKB, m_blk * kM, kEvictFirst);
      }
    }
  } else if (warp == 1) {
    // ---- MMA issue: 128 features x 16 tokens, K = d ----
    constexpr uint32_t idesc = umma_idesc_bf16(kM, kNT);
    mbar_wait(b_ready, 0);
    if (lane == 0) B200_FTRACE(8);
    tc_fence_after();
    if (single) {
      mbar_wait(&full_bar[0], 0);
      tc_fence_after();
      if (elect_one()) {
        uint64_t da = umma_desc_k_sw128(smem_u32(ring));
        uint64_t db = umma_desc_k_sw128(smem_u32(bop));
        for (int kb = 0; kb < n_kb; ++kb) {
#pragma unroll
          for (int k = 0; k < kKB / 16; ++k) {
            umma_bf16_ss(tmem_base, da + static_cast<uint64_t>(k * 2), db + static_cast<uint64_t>(k * 2), idesc,
                         (kb > 0 || k > 0) ? 1u : 0u);
          }
          da += static_cast<uint64_t>(kABytes >> 4);
          db += static_cast<uint64_t>((kNT * 128) >> 4);
        }
        umma_commit(tmem_full);
      }
      __syncwarp();
    } else {
      for (int kb = 0; kb < n_kb; ++kb) {
        const int s = kb % a.stages;
        mbar_wait(&full_bar[s], static_cast<uint32_t>(kb / a.stages) & 1u);
        tc_fence_after();
        const uint64_t da = umma_desc_k_sw128(smem_u32(ring + s * kABytes));
        const uint64_t db = umma_desc_k_sw128(smem_u32(bop + kb * kNT * 128));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kKB / 16; ++k) {
            umma_bf16_ss(tmem_base, da + static_cast<uint64_t>(k * 2), db + static_cast<uint64_t>(k * 2), idesc,
                         (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (kb == n_kb - 1) umma_commit(tmem_full);
        }
        __syncwarp();
      }
    }
    if (lane == 0) B200_FTRACE(9);
  } else if (warp >= 4) {
    // ---- LayerNorm: warp g normalises token row g -> bf16 B operand in shared memory ----
    const int g = warp - 4;                  // 0..15
    const int te = threadIdx.x - 128;        // 0..511
    const int n4 = a.d >> 2;
    // gamma / beta do not depend on the upstream kernel: stage them before the grid dependency
    {
      const float4* g4 = reinterpret_cast<const float4*>(a.gamma);
      const float4* b4 = reinterpret_cast<const float4*>(a.beta);
      float4* sg = reinterpret_cast<float4*>(gb);
      for (int i = te; i < n4; i += kLnWarps * 32) {
        sg[i] = __ldg(g4 + i);
        sg[n4 + i] = __ldg(b4 + i);
      }
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");  // staged by all LayerNorm threads; off the critical path
    pdl_wait();                              // x is complete once the upstream grid has finished
    if (te == 0) B200_FTRACE(3);
    const int row = tok0 + g;
    const bool row_ok = row < a.rows;
    const float4* xr = reinterpret_cast<const float4*>(a.x + static_cast<size_t>(row_ok ? row : 0) * a.d);
    // The lane's whole share of the row is fetched with independent loads before anything is used.
    float4 v[kMaxF4];
#pragma unroll
    for (int j = 0; j < kMaxF4; ++j) {
      const int i = lane + 32 * j;
      v[j] = (row_ok && i < n4) ? xr[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (te == 0) B200_FTRACE(4);
    // both moments about the row's first element (cancellation-safe in fp32)
    const float shift = __shfl_sync(0xffffffffu, v[0].x, 0);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxF4; ++j) {
      if (lane + 32 * j < n4) {
        const float a0 = v[j].x - shift, a1 = v[j].y - shift, a2 = v[j].z - shift, a3 = v[j].w - shift;
        s1 += (a0 + a1) + (a2 + a3);
        s2 += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
      }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    const float inv_d = 1.0f / static_cast<float>(a.d);
    const float m1 = s1 * inv_d;
    const float mean = shift + m1;
    const float rstd = rsqrtf(fmaxf(s2 * inv_d - m1 * m1, 0.f) + a.eps);
    const float4* sg = reinterpret_cast<const float4*>(gb);
    const uint32_t row_base = smem_u32(bop) + static_cast<uint32_t>(g * 128);
#pragma unroll
    for (int j = 0; j < kMaxF4; ++j) {
      const int i = lane + 32 * j;
      if (i < n4) {
        uint32_t lo = 0u, hi = 0u;  // rows beyond the batch feed zeros (their outputs are never stored)
        if (row_ok) {
          const float4 gm = sg[i], bb = sg[n4 + i];
          lo = pack_bf16x2((v[j].x - mean) * rstd * gm.x + bb.x, (v[j].y - mean) * rstd * gm.y + bb.y);
          hi = pack_bf16x2((v[j].z - mean) * rstd * gm.z + bb.z, (v[j].w - mean) * rstd * gm.w + bb.w);
        }
        const int col = i << 2;                         // K index of the first of the 4 elements
        const int kb = (col >> 6) - kb0;                // k-block within this CTA's K range
        if (kb >= 0 && kb < n_kb) {
          const uint32_t chunk = static_cast<uint32_t>((col & 63) >> 3) ^ static_cast<uint32_t>(g & 7);
          const uint32_t addr = row_base + static_cast<uint32_t>(kb) * (kNT * 128) + (chunk << 4) +
                                static_cast<uint32_t>((col & 4) << 1);
          asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(lo), "r"(hi) : "memory");
        }
      }
    }
    fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's reads
    __syncwarp();
    if (lane == 0) mbar_arrive(b_ready);
    if (te == 0) B200_FTRACE(5);

    // ---- epilogue: lane = feature; the 4 warps of a TMEM lane quarter take 4 token columns each ----
    const int q = warp & 3;                  // TMEM lanes [32 q, 32 q + 32): the quarter this warp may read
    const int c0 = (g >> 2) * 4;             // its token columns
    const int feature = m_blk * kM + q * 32 + lane;
    const bool f_ok = feature < a.features;
    const float bias = (f_ok && a.bias != nullptr) ? __ldg(a.bias + feature) : 0.f;
    mbar_wait(tmem_full, 0);
    if (te == 0) B200_FTRACE(6);
    tc_fence_after();
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0))
                 : "memory");
    tmem_ld_wait();
    const int n_valid = f_ok ? (a.rows - tok0 - c0) : 0;  // token columns of this warp that exist
    if (a.partial_out != nullptr) {
      float* o = a.partial_out + (static_cast<size_t>(blockIdx.z) * a.plane_rows + tok0 + c0) * a.ld_out + feature;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < n_valid) *o = __uint_as_float(r[j]);  // 32 lanes -> 128 contiguous bytes of one token row
        o += a.ld_out;
      }
    } else {
      __nv_bfloat16* o = a.out + static_cast<size_t>(tok0 + c0) * a.ld_out + feature;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float y = __uint_as_float(r[j]) + bias;
        if (GELU) y = gelu_new(y);
        if (j < n_valid) *o = __float2bfloat16(y);  // 32 lanes -> 64 contiguous bytes of one token row
        o += a.ld_out;
      }
    }
    if (te == 0) B200_FTRACE(7);
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 32);
  }
  B200_STEP_END();
}

// ------------------------------------------------------------------------------------------
// gemm_redadd_kernel
// ------------------------------------------------------------------------------------------
struct RedAddKernelArgs {
  const float* bias;  // [features] added by split 0
  int features;
  int kb_per_split;   // k-blocks of one CTA: all resident in shared memory
};

template <int NB>  // token columns (UMMA N): 16, 32 or 64
__global__ void __launch_bounds__(kThreads, 1)
gemm_redadd_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_act,
                   const __grid_constant__ CUtensorMap map_x, const RedAddKernelArgs a) {
  constexpr int kBBytes = NB * kKB * 2;
  constexpr int kStage = kABytes + kBBytes;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* ring = smem;  // kb_per_split stages; reused as the epilogue's staging area (NB x 512 B)
  const int ring_bytes = a.kb_per_split * kStage > NB * 512 ? a.kb_per_split * kStage : NB * 512;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + ring_bytes);
  uint64_t* tmem_full = full_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_blk = blockIdx.x, split = blockIdx.y;
  const int kb0 = split * a.kb_per_split;
  B200_STEP_BEGIN(310);
  if ((smem_u32(smem) & 1023u) != 0) __trap();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_w);
    tma_prefetch_desc(&map_act);
    tma_prefetch_desc(&map_x);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(full_bar, 1);
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, NB < 32 ? 32 : NB);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(full_bar, static_cast<uint32_t>(a.kb_per_split) * kStage);
      for (int i = 0; i < a.kb_per_split; ++i) {  // weights first: independent of the upstream kernel
        tma_load_2d(ring + i * kStage, &map_w, full_bar, (kb0 + i) * kKB, m_blk * kM, kEvictFirst);
      }
      pdl_wait();
      B200_FTRACE(3);
      for (int i = 0; i < a.kb_per_split; ++i) {
        tma_load_2d(ring + i * kStage + kABytes, &map_act, full_bar, (kb0 + i) * kKB, 0, kEvictLast);
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(kM, NB);
    mbar_wait(full_bar, 0);
    if (lane == 0) B200_FTRACE(4);
    tc_fence_after();
    if (elect_one()) {
      for (int i = 0; i < a.kb_per_split; ++i) {
        const uint64_t da = umma_desc_k_sw128(smem_u32(ring + i * kStage));
        const uint64_t db = umma_desc_k_sw128(smem_u32(ring + i * kStage + kABytes));
#pragma unroll
        for (int k = 0; k < kKB / 16; ++k) {
          umma_bf16_ss(tmem_base, da + static_cast<uint64_t>(k * 2), db + static_cast<uint64_t>(k * 2), idesc,
                       (i > 0 || k > 0) ? 1u : 0u);
        }
      }
      umma_commit(tmem_full);
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ---- epilogue: [128 features x NB tokens] in TMEM -> [NB tokens x 128 features] fp32 in
    // shared memory as four SWIZZLE_128B boxes of 32 features -> TMA reduce-add into x ----
    const int q = warp - 4;
    const int feature = m_blk * kM + q * 32 + lane;
    const float bias = (split == 0 && a.bias != nullptr && feature < a.features) ? __ldg(a.bias + feature) : 0.f;
    pdl_wait();  // x must not be touched before the upstream grid (its last reader) has finished
    mbar_wait(tmem_full, 0);  // every MMA has retired: the operand ring is free to be overwritten
    if (threadIdx.x == 128) B200_FTRACE(6);
    tc_fence_after();
    uint8_t* box = ring + q * (NB * 128);  // this warp's 32 features: rows = tokens, 128 B per row
    constexpr int kChunk = NB >= 32 ? 32 : 16;
#pragma unroll
    for (int c0 = 0; c0 < NB; c0 += kChunk) {
      uint32_t r[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0);
      if (kChunk == 32) {
        tmem_ld_32x32(taddr, r);
      } else {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(taddr)
            : "memory");
      }
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < kChunk; ++j) {
        const int t = c0 + j;  // token row of the box; lane = column (feature)
        const uint32_t off = static_cast<uint32_t>(t * 128) +
                             ((static_cast<uint32_t>(lane >> 2) ^ static_cast<uint32_t>(t & 7)) << 4) +
                             static_cast<uint32_t>((lane & 3) << 2);
        *reinterpret_cast<float*>(box + off) = __uint_as_float(r[j]) + bias;
      }
    }
    tc_fence_before();
    fence_proxy_async();
    __syncwarp();
    if (threadIdx.x == 128) B200_FTRACE(5);
    if (lane == 0) {  // each warp hands its own box of 32 features to the TMA unit as soon as it is staged
      const int f0 = m_blk * kM + q * 32;
      if (f0 < a.features) {
        // rows beyond the tensor map's extent (tokens past the batch) are clipped by the TMA unit
        asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];"
                     :
                     : "l"(reinterpret_cast<uint64_t>(&map_x)), "r"(smem_u32(box)), "r"(f0), "r"(0)
                     : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // complete before the CTA exits
      if (threadIdx.x == 128) B200_FTRACE(7);
    }
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, NB < 32 ? 32 : NB);
  }
  B200_STEP_END();
}

template <typename K>
int set_smem(K kernel, int bytes) {
  // per device, and cheap: set on every launch rather than remembering which devices have it
  B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return kOk;
}

template <typename K, typename... Args>
int launch_pdl(K kernel, dim3 grid, int threads, int smem_bytes, cudaStream_t stream, int pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, args...));
  return kOk;
}

constexpr int kSmemLimit = 227 * 1024;

}  // namespace

bool ln_gemm_supported(int rows, int d, int features) {
  if (rows <= 0 || rows > 64 || d % kKB != 0 || d > kMaxF4 * 128 || features % 8 != 0) return false;
  const int n_kb = d / kKB;
  return n_kb * kNT * 128 + 2 * d * 4 + 2 * kABytes + 512 <= kSmemLimit;
}

// True when a CTA can keep its whole [128 features x d] weight tile in shared memory next to the
// B operand (d <= 768): one TMA barrier, one uninterrupted MMA chain. That is the shape for which
// the fused decode layer was measured faster than the partial-plane path (GPT-2 small, 1..64
// rows: 228 / 265 / 314 / 348 us per step against 270 / 301 / 355 / 374); with a recycling ring
// (d = 1024, 1280) a CTA's serial pass over K is longer than what the three saved hops give back
// (medium 16 rows 1043 vs 1002 us, large 32 rows 1321 vs 1089 us).
bool ln_gemm_weights_resident(int d) {
  const int n_kb = d / kKB;
  return d % kKB == 0 && n_kb * (kABytes + kNT * 128) + 2 * d * 4 + 512 <= kSmemLimit;
}

int ln_gemm(const LnGemmArgs& g, cudaStream_t stream) {
  const int split = g.split > 1 ? g.split : 1;
  B200_CHECK_ARG(g.x && g.gamma && g.beta && g.w && (split > 1 ? g.partial_out != nullptr : g.out != nullptr),
                 "ln_gemm: null argument");
  B200_CHECK_ARG(ln_gemm_supported(g.rows, g.d, g.features), "ln_gemm: shape not supported (rows <= 64, d % 64 == 0)");
  B200_CHECK_ARG((g.d / kKB) % split == 0, "ln_gemm: split must divide d / 64");
  B200_CHECK_ARG(split == 1 || !g.gelu, "ln_gemm: the gelu epilogue needs the whole K range in one CTA");
  const int n_kb = g.d / kKB / split;
  const int bop_bytes = n_kb * kNT * 128 + 2 * g.d * 4;  // B operand + staged gamma / beta
  int stages = (kSmemLimit - bop_bytes - 512) / kABytes;
  if (stages > n_kb) stages = n_kb;
  CUtensorMap map_w;
  B200_PROPAGATE(gemm_tensor_map(g.w, g.features, g.d, kM, &map_w));
  LnGemmKernelArgs a = {};
  a.x = g.x;
  a.gamma = g.gamma;
  a.beta = g.beta;
  a.bias = g.bias;
  a.out = g.out;
  a.rows = g.rows;
  a.d = g.d;
  a.features = g.features;
  a.ld_out = g.ld_out;
  a.stages = stages;
  a.eps = g.eps;
  a.kb_per_split = n_kb;
  a.partial_out = split > 1 ? g.partial_out : nullptr;
  a.plane_rows = g.plane_rows > 0 ? g.plane_rows : g.rows;
  const int smem_bytes = stages * kABytes + bop_bytes + (2 * stages + 2) * 8 + 16;
  const dim3 grid((g.features + kM - 1) / kM, (g.rows + kNT - 1) / kNT, split);
  if (g.gelu) {
    B200_PROPAGATE(set_smem(ln_gemm_kernel<true>, smem_bytes));
    return launch_pdl(ln_gemm_kernel<true>, grid, kLnThreads, smem_bytes, stream, g.pdl, map_w, a);
  }
  B200_PROPAGATE(set_smem(ln_gemm_kernel<false>, smem_bytes));
  return launch_pdl(ln_gemm_kernel<false>, grid, kLnThreads, smem_bytes, stream, g.pdl, map_w, a);
}

// K splits of a reduce-add GEMM. Every split adds a [tokens x 128] fp32 tile into the same lines of x,
// so unlike partial planes more splits do not cost the consumer anything, but the L2 adders see
// `split`-way contention: measured on GPT-2 small at 64 rows, (out-proj, MLP-down) = (4, 8) gave
// 348 us per step against 352 (3, 6), 356 (2, 6), 358 (6, 12) and 370 (12, 24). So: the divisor of
// the k-block count nearest to 4 (K = d) or 8 (K = 4d) whose k-blocks fit one CTA's shared memory.
int gemm_redadd_pick_split(int features, int k, int rows, int sms) {
  (void)sms;
  const int n_kb = k / kKB;
  const int nb = rows <= 16 ? 16 : rows <= 32 ? 32 : 64;
  const int max_resident = (kSmemLimit - 512) / (kABytes + nb * kKB * 2);
  const int target = k > features ? 8 : 4;
  int best = 0;
  for (int s = 1; s <= n_kb; ++s) {
    if (n_kb % s != 0 || n_kb / s > max_resident) continue;
    if (best == 0 || (s > target ? s - target : target - s) < (best > target ? best - target : target - best)) best = s;
  }
  return best;
}

int gemm_redadd(const GemmRedAddArgs& g, cudaStream_t stream) {
  B200_CHECK_ARG(g.act && g.w && g.x, "gemm_redadd: null argument");
  B200_CHECK_ARG(g.rows > 0 && g.rows <= 64 && g.k % kKB == 0 && g.features % 32 == 0,
                 "gemm_redadd: shape not supported (rows <= 64, k % 64 == 0, features % 32 == 0)");
  const int n_kb = g.k / kKB;
  const int split = g.split > 0 ? g.split : 1;
  B200_CHECK_ARG(n_kb % split == 0, "gemm_redadd: split must divide k / 64");
  const int nb = g.rows <= 16 ? 16 : g.rows <= 32 ? 32 : 64;
  const int kb_per_split = n_kb / split;
  const int stage = kABytes + nb * kKB * 2;
  B200_CHECK_ARG(kb_per_split * stage + 512 <= kSmemLimit, "gemm_redadd: too few splits: the CTA's k-blocks must fit shared memory");
  CUtensorMap map_w, map_act, map_x;
  B200_PROPAGATE(gemm_tensor_map(g.w, g.features, g.k, kM, &map_w));
  B200_PROPAGATE(gemm_tensor_map(g.act, g.act_rows_alloc > 0 ? g.act_rows_alloc : g.rows, g.k, nb, &map_act));
  // fp32 [rows, features]: box = [nb rows, 32 floats]; the row extent is the live batch, so the
  // reduction never touches rows beyond it
  B200_PROPAGATE(gemm_tensor_map_f32(g.x, g.rows, g.features, nb, &map_x));
  RedAddKernelArgs a = {};
  a.bias = g.bias;
  a.features = g.features;
  a.kb_per_split = kb_per_split;
  const int smem_bytes = (kb_per_split * stage > nb * 512 ? kb_per_split * stage : nb * 512) + 64;
  const dim3 grid((g.features + kM - 1) / kM, split, 1);
  switch (nb) {
    case 16:
      B200_PROPAGATE(set_smem(gemm_redadd_kernel<16>, smem_bytes));
      return launch_pdl(gemm_redadd_kernel<16>, grid, kThreads, smem_bytes, stream, g.pdl, map_w, map_act, map_x, a);
    case 32:
      B200_PROPAGATE(set_smem(gemm_redadd_kernel<32>, smem_bytes));
      return launch_pdl(gemm_redadd_kernel<32>, grid, kThreads, smem_bytes, stream, g.pdl, map_w, map_act, map_x, a);
    default:
      B200_PROPAGATE(set_smem(gemm_redadd_kernel<64>, smem_bytes));
      return launch_pdl(gemm_redadd_kernel<64>, grid, kThreads, smem_bytes, stream, g.pdl, map_w, map_act, map_x, a);
  }
}

B200_DEFINE_STEP_TRACE_SETTER(step_trace_set_fused)

}  // namespace b200llm
