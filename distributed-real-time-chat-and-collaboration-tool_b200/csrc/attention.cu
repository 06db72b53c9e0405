// Attention kernels (see attention.cuh for the oracle citations).
//
// prefill_attn_kernel : flash-style causal attention over packed ragged prompts. One CTA =
//   64 query rows of one (sequence, head); 4 warps x 16 rows; K/V tiles of 64 keys staged in
//   padded shared memory; QK^T and PV on the tensor cores (mma.sync m16n8k16 bf16, fp32
//   accumulate); online softmax in fp32 registers. This is <= 8 % of prefill FLOPs for the
//   BASELINE configs; the tcgen05 path is used for the GEMMs that dominate.
//
// decode_attn_kernel : one new token per sequence against the paged KV cache. HBM-bound:
//   one CTA (4 warps) per (row, head, KV split) stages its cached K/V rows in shared memory with
//   cp.async, 128 tokens per pass (the first pass before the grid dependency resolves), every
//   K/V byte read once; both products on the tensor cores (mma.sync m16n8k16, query split into
//   two bf16 parts), online softmax per warp. Optional split-KV for long contexts, merged
//   deterministically by a combine pass or, for small grids, inside a thread-block cluster.
//
// decode_attn_stream_kernel : the same step when the KV volume is large (many (row, head) units per SM,
//   long contexts): persistent warps that each walk whole units in 32-token tiles fetched by TMA from
//   the paged pools into a per-warp mbarrier ring; see the comment above the kernel.
#include "attention.cuh"

#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "ops.cuh"

#ifdef B200_ATTN_TRACE
// Profiling build only (profiles/attn_trace.cu): CTA 0 records SM-clock milestones.
__device__ unsigned long long g_attn_trace[16];
// ... and every CTA of the last launch records %globaltimer at entry / after the grid
// dependency wait / at exit, plus the SM it ran on.
__device__ unsigned long long g_attn_cta_trace[4096][4];
#define B200_ATRACE(slot)                                                                  \
  do {                                                                                     \
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_attn_trace[slot] = clock64(); \
  } while (0)
#define B200_ATRACE_CTA(which)                                                             \
  do {                                                                                     \
    if (threadIdx.x == 0 && blockIdx.y == 0 && blockIdx.x < 4096) {                        \
      unsigned long long _t;                                                               \
      unsigned int _sm;                                                                    \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t));                               \
      asm volatile("mov.u32 %0, %%smid;" : "=r"(_sm));                                     \
      g_attn_cta_trace[blockIdx.x][which] = _t;                                            \
      g_attn_cta_trace[blockIdx.x][3] = _sm;                                               \
    }                                                                                      \
  } while (0)
#define B200_STRACE_DECL                                                                     \
  unsigned long long st_acc[6] = {0, 0, 0, 0, 0, 0}, st_t = clock64(), st_t0 = st_t, st_g0; \
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(st_g0));
#define B200_STRACE(i)                      \
  do {                                      \
    const unsigned long long _n = clock64(); \
    st_acc[i] += _n - st_t;                 \
    st_t = _n;                              \
  } while (0)
#define B200_STRACE_END                                                             \
  do {                                                                              \
    if (blockIdx.x == gridDim.x / 2 && threadIdx.x == 0) {                          \
      for (int _i = 0; _i < 6; ++_i) g_attn_trace[8 + _i] = st_acc[_i];             \
      g_attn_trace[14] = clock64() - st_t0;                                         \
      unsigned long long _g;                                                        \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_g));                        \
      g_attn_trace[15] = _g - st_g0;                                                \
    }                                                                               \
  } while (0)
#else
#define B200_STRACE_DECL
#define B200_STRACE(i) \
  do {                 \
  } while (0)
#define B200_STRACE_END \
  do {                  \
  } while (0)
#define B200_ATRACE(slot) \
  do {                    \
  } while (0)
#define B200_ATRACE_CTA(which) \
  do {                         \
  } while (0)
#endif

namespace b200llm {

namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kScaleLog2 = 0.125f * kLog2e;  // 1/sqrt(64) folded into the exp2 argument

// ---------------------------------------------------------------------------------------
// Prefill
// ---------------------------------------------------------------------------------------
constexpr int kBQ = 64;
constexpr int kBKV = 64;
constexpr int kLds = 72;  // padded smem row stride (bf16 elements): conflict-free ldmatrix

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem_ptr)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(smem_ptr)));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, "
      "{%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 64 rows x 8 chunks of 16 B, 128 threads -> 4 cp.async each; rows beyond `rows_valid` are
// zero-filled (src-size 0), so the copy never reads outside the sequence.
__device__ __forceinline__ void load_tile_64x64_async(__nv_bfloat16 (*dst)[kLds], const __nv_bfloat16* src,
                                                      int ld, int rows_valid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = threadIdx.x + 128 * i;
    const int row = c >> 3;
    const int col = (c & 7) * 8;
    const bool ok = row < rows_valid;
    const __nv_bfloat16* g = src + static_cast<size_t>(ok ? row : 0) * ld + col;
    const int bytes = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(&dst[row][col])), "l"(g),
                 "r"(bytes)
                 : "memory");
  }
}

// The same tile, gathered from KV pages ([page][head][16 tokens][64]): key position pos0 + row of
// one head; rows at or beyond `total` are zero-filled.
__device__ __forceinline__ void load_tile_64x64_paged(__nv_bfloat16 (*dst)[kLds], const __nv_bfloat16* pool,
                                                      const int32_t* pt, int n_head, int h, int pos0,
                                                      int total) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = threadIdx.x + 128 * i;
    const int row = c >> 3;
    const int col = (c & 7) * 8;
    const int pos = pos0 + row;
    const bool ok = pos < total;
    const int page = ok ? pt[pos / kPageTokens] : 0;
    const __nv_bfloat16* g =
        pool + ((static_cast<size_t>(page) * n_head + h) * kPageTokens + (ok ? pos % kPageTokens : 0)) * kHeadDim + col;
    const int bytes = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(&dst[row][col])), "l"(g),
                 "r"(bytes)
                 : "memory");
  }
}

// kPaged = false: the prompt's own K/V rows come from the packed qkv buffer (and the diagonal
// CTA fills the pages). kPaged = true (shared-prefix reuse): the rows are prompt suffixes at
// positions prefix + 0, 1, ...; every key is read from the pages.
template <bool kPaged>
__global__ void __launch_bounds__(128)
prefill_attn_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ cu_seqlens,
                    int n_head, __nv_bfloat16* __restrict__ out, PrefillCache cache,
                    const int32_t* __restrict__ prefix_len) {
  __shared__ __align__(16) __nv_bfloat16 sQ[kBQ][kLds];
  __shared__ __align__(16) __nv_bfloat16 sKb[2][kBKV][kLds];  // double-buffered K/V tiles
  __shared__ __align__(16) __nv_bfloat16 sVb[2][kBKV][kLds];

  const int b = blockIdx.z;
  const int h = blockIdx.y;
  const int qt = blockIdx.x;
  const int row0 = cu_seqlens[b];
  const int len_q = cu_seqlens[b + 1] - row0;       // rows of this sequence in qkv
  if (qt * kBQ >= len_q) return;
  const int prefix = kPaged ? prefix_len[b] : 0;    // keys in front of them, already in the pages
  const int len = prefix + len_q;                   // keys in all
  const int32_t* pt = (kPaged || cache.slots != nullptr)
                          ? cache.page_table + static_cast<size_t>(cache.slots[b]) * cache.max_pages_per_seq
                          : nullptr;
  // last key tile this CTA's queries can see
  const int kt_last = kPaged ? (prefix + min(len_q, (qt + 1) * kBQ) - 1) / kBKV : qt;
  const int d = n_head * kHeadDim;
  const int ld = 3 * d;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;  // row within the 8-row group
  const int t = lane & 3;   // column pair

  const __nv_bfloat16* q_base = qkv + static_cast<size_t>(row0 + qt * kBQ) * ld + h * kHeadDim;
  const __nv_bfloat16* kv_base = qkv + static_cast<size_t>(row0) * ld + d + h * kHeadDim;
  // Q and the first K/V tile in one cp.async group
  load_tile_64x64_async(sQ, q_base, ld, len_q - qt * kBQ);
  if constexpr (kPaged) {
    load_tile_64x64_paged(sKb[0], cache.k_pool, pt, n_head, h, 0, len);
    load_tile_64x64_paged(sVb[0], cache.v_pool, pt, n_head, h, 0, len);
  } else {
    load_tile_64x64_async(sKb[0], kv_base, ld, len);
    load_tile_64x64_async(sVb[0], kv_base + d, ld, len);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");

  uint32_t qf[4][4];

  float o[8][4];
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  }
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const int qrow_lo = prefix + qt * kBQ + warp * 16 + g;  // this thread's key-space rows: qrow_lo, qrow_lo + 8

  for (int kt = 0; kt <= kt_last; ++kt) {
    __nv_bfloat16 (*sK)[kLds] = sKb[kt & 1];
    __nv_bfloat16 (*sV)[kLds] = sVb[kt & 1];
    // prefetch the next tile into the other buffer (its last readers finished at the barrier
    // that closed the previous iteration), then wait for the current one
    if (kt + 1 <= kt_last) {
      if constexpr (kPaged) {
        load_tile_64x64_paged(sKb[(kt + 1) & 1], cache.k_pool, pt, n_head, h, (kt + 1) * kBKV, len);
        load_tile_64x64_paged(sVb[(kt + 1) & 1], cache.v_pool, pt, n_head, h, (kt + 1) * kBKV, len);
      } else {
        const __nv_bfloat16* nk = kv_base + static_cast<size_t>((kt + 1) * kBKV) * ld;
        load_tile_64x64_async(sKb[(kt + 1) & 1], nk, ld, len - (kt + 1) * kBKV);
        load_tile_64x64_async(sVb[(kt + 1) & 1], nk + d, ld, len - (kt + 1) * kBKV);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    if (kt == 0) {
      const int i = lane >> 3, r = lane & 7;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        ldmatrix_x4(qf[ks], &sQ[warp * 16 + (i & 1) * 8 + r][ks * 16 + (i >> 1) * 8]);
      }
    }
    if (!kPaged && kt == qt && cache.slots != nullptr) {
      // diagonal tile: this CTA fills the KV pages ([page][head][16 tokens][64]) from smem
      const int valid = min(kBKV, len - kt * kBKV);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = threadIdx.x + 128 * i;
        const int row = c >> 3, col = (c & 7) * 8;
        if (row < valid) {
          const int pos = kt * kBKV + row;
          const size_t dst = ((static_cast<size_t>(pt[pos / kPageTokens]) * n_head + h) * kPageTokens +
                              pos % kPageTokens) * kHeadDim + col;
          *reinterpret_cast<uint4*>(cache.k_pool + dst) = *reinterpret_cast<const uint4*>(&sK[row][col]);
          *reinterpret_cast<uint4*>(cache.v_pool + dst) = *reinterpret_cast<const uint4*>(&sV[row][col]);
        }
      }
    }

    float s[8][4];
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
    }
    {
      const int i = lane >> 3, r = lane & 7;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
        for (int np = 0; np < 4; ++np) {
          uint32_t kf[4];
          ldmatrix_x4(kf, &sK[np * 16 + (i >> 1) * 8 + r][ks * 16 + (i & 1) * 8]);
          mma_bf16_16816(s[2 * np], qf[ks], kf[0], kf[1]);
          mma_bf16_16816(s[2 * np + 1], qf[ks], kf[2], kf[3]);
        }
      }
    }

    // causal + length mask (only the diagonal tile and a ragged last tile need it), row maxima
    const bool need_mask = ((kt + 1) * kBKV - 1 > prefix + qt * kBQ) || ((kt + 1) * kBKV > len);
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int n = 0; n < 8; ++n) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (need_mask) {
          const int col = kt * kBKV + n * 8 + 2 * t + (c & 1);
          const int row = qrow_lo + (c >> 1) * 8;
          if (col > row || col >= len) s[n][c] = -INFINITY;
        }
        mx[c >> 1] = fmaxf(mx[c >> 1], s[n][c]);
      }
    }
    float alpha[2], m_use[2];
#pragma unroll
    for (int r2 = 0; r2 < 2; ++r2) {
      mx[r2] = fmaxf(mx[r2], __shfl_xor_sync(0xffffffffu, mx[r2], 1));
      mx[r2] = fmaxf(mx[r2], __shfl_xor_sync(0xffffffffu, mx[r2], 2));
      const float m_new = fmaxf(m_run[r2], mx[r2]);
      m_use[r2] = (m_new == -INFINITY) ? 0.f : m_new;  // fully masked row (beyond len)
      alpha[r2] = (m_run[r2] == -INFINITY) ? 0.f : exp2f((m_run[r2] - m_use[r2]) * kScaleLog2);
      m_run[r2] = m_new;
      l_run[r2] *= alpha[r2];
    }
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      o[n][0] *= alpha[0];
      o[n][1] *= alpha[0];
      o[n][2] *= alpha[1];
      o[n][3] *= alpha[1];
    }
#pragma unroll
    for (int n = 0; n < 8; ++n) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float p = exp2f((s[n][c] - m_use[c >> 1]) * kScaleLog2);  // exp2(-inf) = 0
        s[n][c] = p;
        l_run[c >> 1] += p;
      }
    }

    // O += P V
    {
      const int i = lane >> 3, r = lane & 7;
#pragma unroll
      for (int ks2 = 0; ks2 < 4; ++ks2) {
        uint32_t pf[4];
        pf[0] = pack_bf16x2(s[2 * ks2][0], s[2 * ks2][1]);
        pf[1] = pack_bf16x2(s[2 * ks2][2], s[2 * ks2][3]);
        pf[2] = pack_bf16x2(s[2 * ks2 + 1][0], s[2 * ks2 + 1][1]);
        pf[3] = pack_bf16x2(s[2 * ks2 + 1][2], s[2 * ks2 + 1][3]);
#pragma unroll
        for (int dp = 0; dp < 4; ++dp) {
          uint32_t vf[4];
          ldmatrix_x4_trans(vf, &sV[ks2 * 16 + (i & 1) * 8 + r][dp * 16 + (i >> 1) * 8]);
          mma_bf16_16816(o[2 * dp], pf, vf[0], vf[1]);
          mma_bf16_16816(o[2 * dp + 1], pf, vf[2], vf[3]);
        }
      }
    }
    __syncthreads();  // this tile's buffer may be overwritten by the prefetch two tiles ahead
  }

  // finalize
#pragma unroll
  for (int r2 = 0; r2 < 2; ++r2) {
    l_run[r2] += __shfl_xor_sync(0xffffffffu, l_run[r2], 1);
    l_run[r2] += __shfl_xor_sync(0xffffffffu, l_run[r2], 2);
  }
#pragma unroll
  for (int r2 = 0; r2 < 2; ++r2) {
    const int row = qrow_lo + r2 * 8;
    if (row < len) {
      const float inv = 1.0f / l_run[r2];
      __nv_bfloat16* orow = out + static_cast<size_t>(row0 + row - prefix) * d + h * kHeadDim;
#pragma unroll
      for (int n = 0; n < 8; ++n) {
        const uint32_t pk = pack_bf16x2(o[n][2 * r2] * inv, o[n][2 * r2 + 1] * inv);
        *reinterpret_cast<uint32_t*>(orow + n * 8 + 2 * t) = pk;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Decode
// ---------------------------------------------------------------------------------------
// One CTA per (row, head, KV split), 4 warps. The CTA's cached K/V rows are staged in shared
// memory 128 tokens at a time (the first 128 before the grid dependency resolves) and both
// products run on the tensor cores: a single query only fills one row of an m16n8k16 tile, but
// the kernel is bound by instruction issue, not by math, and the MMA form needs ~5x fewer
// instructions per cached token than a CUDA-core dot product per lane group.
//   scores:  D[16 x 8 tokens] = A[16 x 16 dims] * B[16 dims x 8 tokens]
//            A row 0 = q (bf16 high part), row 8 = q - high part (low part): lanes 0..3 hold both
//            rows, so c0 + c2 is the score with q at ~fp32 precision;  B = K rows via ldmatrix.
//   output:  D[16 dims x 8] = A[16 dims x 16 tokens] * B[16 tokens x 8]
//            A = V^T via ldmatrix.trans;  B column 0 = softmax weights, which sit in lanes 0..3 in
//            exactly the layout the score tile left them in.  Column 0 of D is the result.
constexpr int kScratchStride = kHeadDim + 2;  // o[64], m, l
constexpr int kMaxSmemPages = 256;            // page ids staged in smem (4096 tokens)
constexpr int kStageTokens = 128;             // K/V rows staged per pass (32 KB)
constexpr int kWarpTokens = kStageTokens / 4; // rows of a pass one warp attends
constexpr int kTailRows = 32;                 // rows of the second pass prefetched into registers

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

__device__ __forceinline__ uint32_t attn_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void attn_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ float ld_peer_smem(const float* local, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local)), "r"(rank));
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
  return v;
}

// Ordering argument for everything this kernel does BEFORE griddepcontrol.wait.
// In a decode step the GEMM, LayerNorm and gelu-reduce kernels signal launch_dependents at entry;
// the embed kernel (first of every step), this kernel and the selector signal only AFTER their own
// wait has returned, i.e. after their predecessor has completely finished. A CTA of this kernel
// can therefore only be running if this step's embed kernel has passed its wait, which means the
// previous step's selector -- and with it the whole previous step -- is complete and visible:
// seq_len / finished are final, and KV positions < seq_len (written by earlier steps) are
// immutable. Only the QKV projection of THIS step is not ready. So the cached KV rows are
// fetched before the wait and overlap the QKV GEMM.
//
// Stage layout: row = token (64 bf16 = 8 chunks of 16 B); chunk c of row r is stored at chunk
// position c ^ (r & 7), which makes every ldmatrix phase (8 rows x 16 B) conflict-free.
// kCluster: the KV splits of one (row, head) form a thread-block cluster and merge in the kernel
// (a separate instantiation, so the one-split kernel of the large-batch step is not touched).
template <bool kCluster>
__global__ void __launch_bounds__(128, 6) decode_attn_kernel(DecodeAttnArgs a) {
  __shared__ float sm_o[4][kHeadDim];
  __shared__ float sm_ml[4][2];
  __shared__ float sm_part[kCluster ? kHeadDim + 2 : 1];  // this CTA's merged (o, m, l) for the in-cluster combine
  __shared__ int32_t sm_pages[kMaxSmemPages];
  __shared__ __align__(16) __nv_bfloat16 sm_q[2][kHeadDim];   // this token's q: high / low bf16 parts
  __shared__ __align__(16) __nv_bfloat16 sm_kv[2][kHeadDim];  // this token's k, v (cache dtype)
  __shared__ __align__(16) uint4 sm_k[kStageTokens][8];       // staged K rows (swizzled chunks)
  __shared__ __align__(16) uint4 sm_v[kStageTokens][8];       // staged V rows (swizzled chunks)

  const int bh = blockIdx.x;
  const int b = bh / a.n_head;
  const int h = bh % a.n_head;
  const int split = blockIdx.y;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int d = a.n_head * kHeadDim;
  B200_ATRACE(0);
  B200_ATRACE_CTA(0);
  B200_STEP_BEGIN(5);

  const int slot = a.active_slots ? a.active_slots[b] : b;
  if (slot < 0) {
    pdl_wait();
    pdl_launch_dependents();
    return;
  }
  const int32_t* pt = a.page_table + static_cast<size_t>(slot) * a.max_pages_per_seq;
  const int n_stage_pages = a.max_pages_per_seq < kMaxSmemPages ? a.max_pages_per_seq : kMaxSmemPages;
  for (int i = threadIdx.x; i < n_stage_pages; i += blockDim.x) sm_pages[i] = pt[i];
  const int pos = a.seq_len[slot];
  const int ctx = pos + 1;
  const bool fin = a.finished != nullptr && a.finished[slot] != 0;
  float rbias[2] = {0.f, 0.f};
  if (a.qkv_partials != nullptr) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int vi = threadIdx.x + r * 128;
      if (vi < 3 * kHeadDim) rbias[r] = __ldg(a.qkv_bias + (vi >> 6) * d + h * kHeadDim + (vi & 63));
    }
  }
  __syncthreads();

  int chunk = (ctx + a.n_splits - 1) / a.n_splits;
  chunk = ((chunk + kPageTokens - 1) / kPageTokens) * kPageTokens;
  const int t_begin = split * chunk;
  const int t_end = min(ctx, t_begin + chunk);           // may include `pos` (the new token)
  const int n_tok = max(0, t_end - t_begin);             // rows this CTA attends
  const int n_cached = max(0, min(t_end, pos) - t_begin);  // of which already in the cache
  const bool own_new = pos >= t_begin && pos < t_end;    // then n_tok == n_cached + 1
  const int new_pass = (n_cached / kStageTokens) * kStageTokens;  // pass holding the new token's row
  const int new_row = n_cached - new_pass;

  auto page_of = [&](int tpos) -> int {
    const int pi = tpos / kPageTokens;
    return pi < n_stage_pages ? sm_pages[pi] : pt[pi];
  };
  auto kv_offset = [&](int tpos) -> size_t {
    return ((static_cast<size_t>(page_of(tpos)) * a.n_head + h) * kPageTokens + tpos % kPageTokens) * kHeadDim;
  };
  // Stage the rows of the pass that starts at CTA-relative token p0: cached rows by cp.async, and
  // zeros into the V rows of the last 16-row tile that hold no token (0 * garbage must not be NaN).
  // The new token's row is written by nobody here; it is filled from sm_kv after the wait.
  uint4 tail_k[2], tail_v[2];  // rows 128..159 of the CTA's range, fetched before the wait (below)
  auto stage_pass = [&](int p0) {
    const int rows_cached = max(0, min(n_cached - p0, kStageTokens));
    int first_async = 0;
    if (p0 == kStageTokens) {  // these rows are already in registers
      first_async = kTailRows;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const int row = (threadIdx.x >> 3) + 16 * k, c = threadIdx.x & 7;
        if (row < rows_cached) {
          sm_k[row][c ^ (row & 7)] = tail_k[k];
          sm_v[row][c ^ (row & 7)] = tail_v[k];
        }
      }
    }
    for (int i = first_async * 16 + threadIdx.x; i < rows_cached * 16; i += blockDim.x) {
      const int row = i >> 4, part = i & 15, c = part & 7;
      const size_t off = kv_offset(t_begin + p0 + row) + c * 8;
      if (part < 8) {
        cp_async_16(&sm_k[row][c ^ (row & 7)], a.k_pool + off);
      } else {
        cp_async_16(&sm_v[row][c ^ (row & 7)], a.v_pool + off);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    const int rows_pass = min(n_tok - p0, kStageTokens);
    const int first_empty = rows_cached + ((own_new && new_pass == p0) ? 1 : 0);
    const int rows_tiles = ((rows_pass + 15) / 16) * 16;
    for (int i = first_empty * 8 + threadIdx.x; i < rows_tiles * 8; i += blockDim.x) {
      sm_v[i >> 3][i & 7] = make_uint4(0u, 0u, 0u, 0u);
    }
  };
  stage_pass(0);
  // The first rows of the second pass wait in registers (a context a little over 128 tokens is the
  // common case for short prompts); anything beyond is pulled into L2 now and staged later.
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int r = kStageTokens + (threadIdx.x >> 3) + 16 * k;
    tail_k[k] = make_uint4(0u, 0u, 0u, 0u);
    tail_v[k] = make_uint4(0u, 0u, 0u, 0u);
    if (r < n_cached) {
      const size_t off = kv_offset(t_begin + r) + (threadIdx.x & 7) * 8;
      tail_k[k] = ld_nc_v4(a.k_pool + off);
      tail_v[k] = ld_nc_v4(a.v_pool + off);
    }
  }
  for (int r = kStageTokens + kTailRows + (threadIdx.x >> 1); r < n_cached; r += 64) {
    const __nv_bfloat16* pool = (threadIdx.x & 1) ? a.v_pool : a.k_pool;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(pool + kv_offset(t_begin + r)));
  }

  B200_ATRACE(1);
  pdl_wait();
  pdl_launch_dependents();
  B200_ATRACE(2);
  B200_ATRACE_CTA(1);

  // This token's q/k/v. With a split-K projection the CTA reduces the fp32 partial planes
  // (fixed summation order -> deterministic) and adds the bias. q (scaled for exp2) is split into
  // two bf16 parts, k/v are rounded to the cache dtype; everyone reads them from shared memory.
  {
    float val[2] = {rbias[0], rbias[1]};
    if (a.qkv_partials != nullptr) {
      const size_t plane = static_cast<size_t>(a.partial_rows) * 3 * d;
      for (int s0 = 0; s0 < a.n_qkv_partials; s0 += 8) {
        float x[2][8];
#pragma unroll
        for (int r = 0; r < 2; ++r) {  // both values' planes requested before anything is summed
          const int vi = threadIdx.x + r * 128;
          const float* src = a.qkv_partials + static_cast<size_t>(b) * 3 * d + (vi >> 6) * d + h * kHeadDim + (vi & 63);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            x[r][j] = (vi < 3 * kHeadDim && s0 + j < a.n_qkv_partials)
                          ? __ldcg(src + static_cast<size_t>(s0 + j) * plane)
                          : 0.f;
          }
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
          for (int j = 0; j < 8; ++j) val[r] += x[r][j];
        }
      }
    } else {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int vi = threadIdx.x + r * 128;
        if (vi < 3 * kHeadDim) {
          val[r] = __bfloat162float(a.qkv[static_cast<size_t>(b) * 3 * d + (vi >> 6) * d + h * kHeadDim + (vi & 63)]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int vi = threadIdx.x + r * 128;
      if (vi < 3 * kHeadDim) {
        const int part = vi >> 6, dim = vi & 63;
        if (part == 0) {
          const float x = val[r] * kScaleLog2;
          const __nv_bfloat16 hi = __float2bfloat16(x);
          sm_q[0][dim] = hi;
          sm_q[1][dim] = __float2bfloat16(x - __bfloat162float(hi));
        } else {
          const __nv_bfloat16 v16 = __float2bfloat16(val[r]);
          sm_kv[part - 1][dim] = v16;
          if (own_new && new_pass == 0) {  // single-pass case: straight into its stage row
            uint4* rowp = part == 1 ? &sm_k[new_row][(dim >> 3) ^ (new_row & 7)] : &sm_v[new_row][(dim >> 3) ^ (new_row & 7)];
            reinterpret_cast<__nv_bfloat16*>(rowp)[dim & 7] = v16;
          }
        }
      }
    }
  }
  B200_ATRACE(7);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  B200_ATRACE(8);
  __syncthreads();
  B200_ATRACE(4);

  // append this token to its page (the split that owns position `pos` does it)
  if (own_new && !fin && threadIdx.x < 16) {
    const int c = threadIdx.x & 7;
    const size_t dst = kv_offset(pos) + c * 8;
    __nv_bfloat16* pool = threadIdx.x < 8 ? a.k_pool : a.v_pool;
    *reinterpret_cast<uint4*>(pool + dst) = *reinterpret_cast<const uint4*>(&sm_kv[threadIdx.x >> 3][c * 8]);
  }

  // q as the A operand: lanes 0..3 carry row 0 (high part) and row 8 (low part), all others zero
  auto load_q = [&](uint32_t (&qa)[4], int s) {
    const int j2 = 2 * (lane & 3);
    const bool on = lane < 4;
    qa[0] = on ? *reinterpret_cast<const uint32_t*>(&sm_q[0][16 * s + j2]) : 0u;
    qa[1] = on ? *reinterpret_cast<const uint32_t*>(&sm_q[1][16 * s + j2]) : 0u;
    qa[2] = on ? *reinterpret_cast<const uint32_t*>(&sm_q[0][16 * s + 8 + j2]) : 0u;
    qa[3] = on ? *reinterpret_cast<const uint32_t*>(&sm_q[1][16 * s + 8 + j2]) : 0u;
  };

  float m_run = -INFINITY;  // warp-uniform running max (log2 domain)
  float l_part = 0.f;       // this lane's share of the running sum (lanes 0..3 matter)
  float oacc[4][4];
#pragma unroll
  for (int mt = 0; mt < 4; ++mt) {
#pragma unroll
    for (int e = 0; e < 4; ++e) oacc[mt][e] = 0.f;
  }

  for (int p0 = 0; p0 < n_tok; p0 += kStageTokens) {  // CTA-uniform trip count
    if (p0 > 0) {
      __syncthreads();  // every warp is done with the previous pass's rows
      stage_pass(p0);
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      if (own_new && new_pass == p0 && threadIdx.x < 16) {
        const int c = threadIdx.x & 7;
        uint4* dstp = threadIdx.x < 8 ? &sm_k[new_row][c ^ (new_row & 7)] : &sm_v[new_row][c ^ (new_row & 7)];
        *dstp = *reinterpret_cast<const uint4*>(&sm_kv[threadIdx.x >> 3][c * 8]);
      }
      __syncthreads();
    }
    const int rows_pass = min(n_tok - p0, kStageTokens);
    const int w0 = warp * kWarpTokens;
    if (w0 < rows_pass) {  // warp-uniform
      float sacc[4][4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
#pragma unroll
        for (int e = 0; e < 4; ++e) sacc[t][e] = 0.f;
      }
#pragma unroll
      for (int sp = 0; sp < 2; ++sp) {  // 32 dims per ldmatrix.x4
        uint32_t qa0[4], qa1[4];
        load_q(qa0, 2 * sp);
        load_q(qa1, 2 * sp + 1);
#pragma unroll
        for (int t = 0; t < 4; ++t) {   // 8 tokens per score tile
          uint32_t kb[4];
          const int row = w0 + 8 * t + (lane & 7);
          const int c = 4 * sp + (lane >> 3);
          ldmatrix_x4(kb, &sm_k[row][c ^ (row & 7)]);
          mma_bf16_16816(sacc[t], qa0, kb[0], kb[1]);
          mma_bf16_16816(sacc[t], qa1, kb[2], kb[3]);
        }
      }
      float s[4][2];
      bool ok[4][2];
      float m_loc = -INFINITY;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          ok[t][e] = lane < 4 && (w0 + 8 * t + 2 * (lane & 3) + e) < rows_pass;
          s[t][e] = sacc[t][e] + sacc[t][e + 2];
          m_loc = ok[t][e] ? fmaxf(m_loc, s[t][e]) : m_loc;
        }
      }
      m_loc = fmaxf(m_loc, __shfl_xor_sync(0xffffffffu, m_loc, 1));
      m_loc = fmaxf(m_loc, __shfl_xor_sync(0xffffffffu, m_loc, 2));
      m_loc = __shfl_sync(0xffffffffu, m_loc, 0);  // lanes 0..3 own the real scores
      const float m_new = fmaxf(m_run, m_loc);     // finite: row w0 of this pass is a real token
      if (m_run != -INFINITY) {                    // warp-uniform
        const float f = exp2f(m_run - m_new);
        l_part *= f;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
#pragma unroll
          for (int e = 0; e < 4; ++e) oacc[mt][e] *= f;
        }
      }
      m_run = m_new;
      float p[4][2];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          p[t][e] = ok[t][e] ? exp2f(s[t][e] - m_new) : 0.f;
          l_part += p[t][e];
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {  // 16 tokens per output tile
        if (w0 + 16 * u < rows_pass) {  // warp-uniform; rows past the last 16-row tile are never read
          const uint32_t b0 = pack_bf16x2(p[2 * u][0], p[2 * u][1]);
          const uint32_t b1 = pack_bf16x2(p[2 * u + 1][0], p[2 * u + 1][1]);
#pragma unroll
          for (int mt = 0; mt < 4; ++mt) {  // 16 dims per tile
            uint32_t va[4];
            const int row = w0 + 16 * u + ((lane >> 4) << 3) + (lane & 7);
            const int c = 2 * mt + ((lane >> 3) & 1);
            ldmatrix_x4_trans(va, &sm_v[row][c ^ (row & 7)]);
            mma_bf16_16816(oacc[mt], va, b0, b1);
          }
        }
      }
    }
  }
  B200_ATRACE(5);

  // per-warp result -> shared: column 0 of the output tiles lives in lanes 0, 4, ..., 28
  float l_warp = l_part + __shfl_xor_sync(0xffffffffu, l_part, 1);
  l_warp += __shfl_xor_sync(0xffffffffu, l_warp, 2);
  if ((lane & 3) == 0) {
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      sm_o[warp][16 * mt + (lane >> 2)] = oacc[mt][0];
      sm_o[warp][16 * mt + 8 + (lane >> 2)] = oacc[mt][2];
    }
    if (lane == 0) {
      sm_ml[warp][0] = m_run;
      sm_ml[warp][1] = l_warp;
    }
  }
  __syncthreads();
  if (threadIdx.x < kHeadDim) {
    const int dim = threadIdx.x;
    float m = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) m = fmaxf(m, sm_ml[w][0]);
    float l = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float f = (sm_ml[w][0] == -INFINITY) ? 0.f : exp2f(sm_ml[w][0] - m);
      l += sm_ml[w][1] * f;
      o += sm_o[w][dim] * f;
    }
    if (a.n_splits == 1) {
      a.out[static_cast<size_t>(b) * d + h * kHeadDim + dim] = __float2bfloat16(o / l);
    } else if (kCluster) {
      sm_part[dim] = o;
      if (dim == 0) {
        sm_part[kHeadDim] = m;
        sm_part[kHeadDim + 1] = l;
      }
    } else {
      float* sc = a.scratch + (static_cast<size_t>(bh) * a.n_splits + split) * kScratchStride;
      sc[dim] = o;
      if (dim == 0) {
        sc[kHeadDim] = m;
        sc[kHeadDim + 1] = l;
      }
    }
  }
  if (kCluster && a.n_splits > 1) {
    // The splits of this (row, head) are the CTAs of one cluster (same blockIdx.x). After a cluster
    // barrier the rank-0 CTA reads every rank's (o, m, l) through distributed shared memory and
    // merges them in rank order (deterministic); the second barrier keeps the other CTAs'
    // shared memory alive until it has been read.
    attn_cluster_sync();
    if (attn_cluster_rank() == 0 && threadIdx.x < kHeadDim) {
      const int dim = threadIdx.x;
      float m = -INFINITY;
      for (int r = 0; r < a.n_splits; ++r) m = fmaxf(m, ld_peer_smem(&sm_part[kHeadDim], r));
      float l = 0.f, o = 0.f;
      for (int r = 0; r < a.n_splits; ++r) {
        const float mr = ld_peer_smem(&sm_part[kHeadDim], r);
        const float f = (mr == -INFINITY) ? 0.f : exp2f(mr - m);
        l += ld_peer_smem(&sm_part[kHeadDim + 1], r) * f;
        o += ld_peer_smem(&sm_part[dim], r) * f;
      }
      a.out[static_cast<size_t>(b) * d + h * kHeadDim + dim] = __float2bfloat16(o / l);
    }
    attn_cluster_sync();
  }
  B200_ATRACE(6);
  B200_ATRACE_CTA(2);
  B200_STEP_END();
}

__global__ void __launch_bounds__(64) decode_combine_kernel(DecodeAttnArgs a) {
  const int bh = blockIdx.x;
  const int b = bh / a.n_head;
  const int h = bh % a.n_head;
  const int dim = threadIdx.x;
  pdl_wait();
  pdl_launch_dependents();
  const int slot = a.active_slots ? a.active_slots[b] : b;
  if (slot < 0) return;
  const float* sc = a.scratch + static_cast<size_t>(bh) * a.n_splits * kScratchStride;
  float m = -INFINITY;
  for (int s = 0; s < a.n_splits; ++s) m = fmaxf(m, sc[s * kScratchStride + kHeadDim]);
  float l = 0.f, o = 0.f;
  for (int s = 0; s < a.n_splits; ++s) {
    const float ms = sc[s * kScratchStride + kHeadDim];
    const float f = (ms == -INFINITY) ? 0.f : exp2f(ms - m);
    l += sc[s * kScratchStride + kHeadDim + 1] * f;
    o += sc[s * kScratchStride + dim] * f;
  }
  const int d = a.n_head * kHeadDim;
  a.out[static_cast<size_t>(b) * d + h * kHeadDim + dim] = __float2bfloat16(o / l);
}

// ---------------------------------------------------------------------------------------
// Decode, streaming form: many (row, head) units per SM and / or long contexts
// ---------------------------------------------------------------------------------------
// decode_attn_kernel above is built for the single-wave step (64 rows x 12 heads at context ~150: every
// CTA resident at once, KV rows fetched under the QKV projection). With more units than CTA slots, or
// several 128-row passes per CTA, its load -> wait -> compute chain repeats serially per CTA and the
// 16 cp.async per thread and pass cost more instructions than the math. This kernel is the
// bandwidth-oriented form for those shapes:
//   * persistent warps (3 per CTA, 3 CTAs per SM); a WARP owns a contiguous range of
//     (row, split, head) units and walks each unit's tokens in tiles of 32 (two KV pages) -- the very
//     per-warp MMA / online-softmax body of the kernel above, so no cross-warp merge and no CTA barrier;
//   * every tile arrives by TMA: one 16 x 64 box per page straight out of the paged pools, SWIZZLE_128B
//     (= the chunk ^ (row & 7) layout ldmatrix wants), completion on an mbarrier; a warp keeps
//     kStreamStages tiles in flight in its own ring and issues across unit boundaries, so the loads of
//     the next unit run under the math of this one. The first tiles are issued before the grid
//     dependency resolves (same ordering argument as above);
//   * split-KV partials go to the same scratch as above and are merged by decode_combine_kernel.
template <int kWarps, int kStages, int kTile>
struct StreamCfg {
  static constexpr int kTileBytes = 2 * kTile * 128;  // K rows + V rows
  static constexpr int kMiscBytes = 640;              // per warp: q parts, result staging, append pages
  static constexpr int kSmemBytes = kWarps * (kStages * kTileBytes + kMiscBytes) + kWarps * kStages * 8 + 1024;
};

// Warp-uniform cursor over (row, split, head) units. Moving to the next unit costs a few integer
// operations; only a new ROW loads anything (its slot, position, finished flag) and divides.
struct StreamCursor {
  int u, b, split, h;
  int slot;      // < 0: padding row
  int pos;       // position of the row's new token
  int chunk;     // tokens per split of this row (page-aligned)
  int fin;       // row is finished: no append
  // derived for (row, split):
  int t_begin, n_tok, n_cached, own_new, n_tiles;
};

__device__ __forceinline__ void stream_load_row(const DecodeAttnArgs& a, StreamCursor& c) {
  c.slot = a.active_slots ? a.active_slots[c.b] : c.b;
  c.pos = 0, c.chunk = kPageTokens, c.fin = 0;
  if (c.slot < 0) return;
  c.pos = a.seq_len[c.slot];
  c.fin = a.finished != nullptr ? a.finished[c.slot] : 0;
  const int chunk = (c.pos + a.n_splits) / a.n_splits;  // ceil((pos + 1) / n_splits)
  c.chunk = ((chunk + kPageTokens - 1) / kPageTokens) * kPageTokens;
}
__device__ __forceinline__ void stream_derive(StreamCursor& c) {
  c.t_begin = c.split * c.chunk;
  const int ctx = c.pos + 1;
  const int t_end = min(ctx, c.t_begin + c.chunk);
  c.n_tok = c.slot < 0 ? 0 : max(0, t_end - c.t_begin);
  c.n_cached = max(0, min(t_end, c.pos) - c.t_begin);
  c.own_new = (c.slot >= 0 && c.pos >= c.t_begin && c.pos < t_end) ? 1 : 0;
  c.n_tiles = (c.n_tok + 31) >> 5;
}
__device__ __forceinline__ void stream_seek(const DecodeAttnArgs& a, StreamCursor& c, int u) {
  c.u = u;
  c.h = u % a.n_head;
  const int bs = u / a.n_head;
  c.split = bs % a.n_splits;
  c.b = bs / a.n_splits;
  stream_load_row(a, c);
  stream_derive(c);
}
__device__ __forceinline__ void stream_next(const DecodeAttnArgs& a, StreamCursor& c, int u_end) {
  ++c.u;
  if (++c.h == a.n_head) {
    c.h = 0;
    if (++c.split == a.n_splits) {
      c.split = 0;
      ++c.b;
      if (c.u < u_end) stream_load_row(a, c);
    }
    stream_derive(c);
  }
}

template <int kStreamWarps, int kStreamStages, int kStreamTile, int kStreamCtasPerSm>
__global__ void __launch_bounds__(kStreamWarps * 32, kStreamCtasPerSm)
decode_attn_stream_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                          const DecodeAttnArgs a) {
  using Cfg = StreamCfg<kStreamWarps, kStreamStages, kStreamTile>;
  constexpr int kStreamTileBytes = Cfg::kTileBytes;
  constexpr int kStreamMiscBytes = Cfg::kMiscBytes;
  constexpr int kTilePages = kStreamTile / kPageTokens;
  static_assert(kStreamTile == 32, "two pages per tile");
  // (only dynamic shared memory in this kernel: the array starts at offset 0 of the CTA's window, so the
  // SWIZZLE_128B tiles are 1024-byte aligned; no pointer arithmetic through integers, which would turn
  // every shared-memory access below into a generic one)
  extern __shared__ __align__(1024) uint8_t stream_smem_raw[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  uint8_t* base = stream_smem_raw;
  uint8_t* ring = base + warp * (kStreamStages * kStreamTileBytes);  // [stage]: K rows 0..31, then V rows 0..31
  uint8_t* misc = base + kStreamWarps * (kStreamStages * kStreamTileBytes) + warp * kStreamMiscBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(base + kStreamWarps * (kStreamStages * kStreamTileBytes + kStreamMiscBytes)) +
                   warp * kStreamStages;
  __nv_bfloat16* sq = reinterpret_cast<__nv_bfloat16*>(misc);  // [2][64] q high / low parts
  float* so = reinterpret_cast<float*>(misc + 256);            // [64] result staging
  int* apg = reinterpret_cast<int*>(misc + 512);               // [stage] page that takes the new token (set with the unit's last tile)
  const int d = a.n_head * kHeadDim;
  B200_STRACE_DECL
  B200_ATRACE_CTA(0);

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStreamStages; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  __syncwarp();

  // this warp's units: a contiguous range, ordered (row, split, head) so neighbours share the row's data
  const int live = a.n_live != nullptr ? min(*a.n_live, a.batch) : a.batch;
  const long long total = static_cast<long long>(live) * a.n_splits * a.n_head;
  const long long gw = static_cast<long long>(blockIdx.x) * kStreamWarps + warp;
  const long long n_gw = static_cast<long long>(gridDim.x) * kStreamWarps;
  const int u_begin = static_cast<int>(gw * total / n_gw);
  const int u_end = static_cast<int>((gw + 1) * total / n_gw);
  if (u_begin >= u_end) {
    pdl_wait();
    pdl_launch_dependents();
    return;
  }

  // ---- producer cursor: which tile is loaded next ----
  StreamCursor P;
  stream_seek(a, P, u_begin);
  int p_tile = 0;
  uint32_t p_count = 0;        // tiles issued so far
  int p_pg = 0;                // lane l: page id of page (first page of the range + 32 * refill + l)
  int p_pg_slot = -1, p_pg_first = -1;
  auto p_skip_empty = [&]() {  // move to the next unit that has tiles
    while (P.u < u_end && P.n_tiles == 0) stream_next(a, P, u_end);
    p_tile = 0;
  };
  auto p_issue = [&]() {       // load tile p_tile of unit P into the next ring slot, advance
    const int first_page = P.t_begin / kPageTokens + ((p_tile * kTilePages) & ~31);
    if (p_pg_slot != P.slot || p_pg_first != first_page) {
      const int pi = first_page + lane;
      p_pg = pi < a.max_pages_per_seq ? a.page_table[static_cast<size_t>(P.slot) * a.max_pages_per_seq + pi] : 0;
      p_pg_slot = P.slot;
      p_pg_first = first_page;
    }
    const int pgi = lane >> 1;  // lane 2j loads K of the tile's page j, lane 2j + 1 its V
    const int page = __shfl_sync(0xffffffffu, p_pg, (p_tile * kTilePages + pgi) & 31);
    const int rows = min(P.n_tok - p_tile * kStreamTile, kStreamTile);
    const int n_pg = (rows + kPageTokens - 1) / kPageTokens;
    const uint32_t stage = p_count % kStreamStages;
    uint8_t* dst = ring + stage * kStreamTileBytes;
    if (lane == 0) mbar_arrive_expect_tx(&full[stage], static_cast<uint32_t>(n_pg) * 2u * kPageTokens * 128u);
    if (lane < 2 * n_pg) {
      const int which = lane & 1;   // 0: K, 1: V
      tma_load_2d(dst + which * (kStreamTile * 128) + pgi * (kPageTokens * 128), which ? &map_v : &map_k,
                  &full[stage], 0, (page * a.n_head + P.h) * kPageTokens, kEvictNormal);
      // the unit's last tile of the split that owns the new token: remember the page it is appended to
      if (which == 0 && P.own_new && p_tile == P.n_tiles - 1 && pgi == n_pg - 1) apg[stage] = page;
    }
    ++p_count;
    if (++p_tile == P.n_tiles) {
      stream_next(a, P, u_end);
      p_skip_empty();
    }
  };
  p_skip_empty();
#pragma unroll 1
  for (int i = 0; i < kStreamStages - 1 && P.u < u_end; ++i) p_issue();

  pdl_wait();
  pdl_launch_dependents();
  B200_STRACE(0);  // start-up: first units, first tiles issued, grid dependency

  // ---- consumer: unit by unit, tile by tile ----
  // The projection of the NEXT unit is requested before this unit's tiles are walked and used after
  // them: under the KV stream an L2 access takes microseconds, so nothing on the unit path waits for a
  // load that could have been asked for earlier.
  constexpr int kPre = 4;   // planes of the next unit's projection held in registers
  float2 nx0[kPre], nx1[kPre], nx2[kPre], nb0, nb1, nb2;
  auto fetch = [&](int b, int h) {  // issue the loads of a unit's q / k / v (lane l: dims 2l, 2l + 1); no use here
    const size_t col = static_cast<size_t>(h) * kHeadDim + 2 * lane;
    if (a.qkv_partials != nullptr) {
      const size_t plane = static_cast<size_t>(a.partial_rows) * 3 * d;
      const float* src = a.qkv_partials + static_cast<size_t>(b) * 3 * d + col;
      nb0 = __ldg(reinterpret_cast<const float2*>(a.qkv_bias + col));
      nb1 = __ldg(reinterpret_cast<const float2*>(a.qkv_bias + d + col));
      nb2 = __ldg(reinterpret_cast<const float2*>(a.qkv_bias + 2 * d + col));
#pragma unroll
      for (int j = 0; j < kPre; ++j) {
        const bool on = j < a.n_qkv_partials;
        const float* p = src + static_cast<size_t>(j) * plane;
        nx0[j] = on ? __ldcg(reinterpret_cast<const float2*>(p)) : make_float2(0.f, 0.f);
        nx1[j] = on ? __ldcg(reinterpret_cast<const float2*>(p + d)) : make_float2(0.f, 0.f);
        nx2[j] = on ? __ldcg(reinterpret_cast<const float2*>(p + 2 * d)) : make_float2(0.f, 0.f);
      }
    } else {
      const __nv_bfloat16* src = a.qkv + static_cast<size_t>(b) * 3 * d + col;
      nx0[0].x = __uint_as_float(__ldcg(reinterpret_cast<const unsigned int*>(src)));
      nx1[0].x = __uint_as_float(__ldcg(reinterpret_cast<const unsigned int*>(src + d)));
      nx2[0].x = __uint_as_float(__ldcg(reinterpret_cast<const unsigned int*>(src + 2 * d)));
    }
  };
  StreamCursor C;
  stream_seek(a, C, u_begin);
  fetch(C.b, C.h);

  uint32_t c_count = 0;  // tiles consumed so far
#pragma unroll 1
  for (; C.u < u_end; stream_next(a, C, u_end)) {
    // this unit's q, k, v from the prefetched values (split-K planes summed in plane order, bias added)
    float2 vq, vk, vv;
    if (a.qkv_partials != nullptr) {
      vq = nb0, vk = nb1, vv = nb2;
#pragma unroll
      for (int j = 0; j < kPre; ++j) {
        vq.x += nx0[j].x, vq.y += nx0[j].y;
        vk.x += nx1[j].x, vk.y += nx1[j].y;
        vv.x += nx2[j].x, vv.y += nx2[j].y;
      }
      if (a.n_qkv_partials > kPre) {  // more planes than the prefetch holds: the rest on demand
        const size_t plane = static_cast<size_t>(a.partial_rows) * 3 * d;
        const float* src = a.qkv_partials + static_cast<size_t>(C.b) * 3 * d + static_cast<size_t>(C.h) * kHeadDim + 2 * lane;
        for (int j = kPre; j < a.n_qkv_partials; ++j) {
          const float* p = src + static_cast<size_t>(j) * plane;
          const float2 x0 = __ldcg(reinterpret_cast<const float2*>(p));
          const float2 x1 = __ldcg(reinterpret_cast<const float2*>(p + d));
          const float2 x2 = __ldcg(reinterpret_cast<const float2*>(p + 2 * d));
          vq.x += x0.x, vq.y += x0.y, vk.x += x1.x, vk.y += x1.y, vv.x += x2.x, vv.y += x2.y;
        }
      }
    } else {
      const uint32_t w0 = __float_as_uint(nx0[0].x), w1 = __float_as_uint(nx1[0].x), w2 = __float_as_uint(nx2[0].x);
      vq = make_float2(__uint_as_float(w0 << 16), __uint_as_float(w0 & 0xffff0000u));
      vk = make_float2(__uint_as_float(w1 << 16), __uint_as_float(w1 & 0xffff0000u));
      vv = make_float2(__uint_as_float(w2 << 16), __uint_as_float(w2 & 0xffff0000u));
    }
    if (C.u + 1 < u_end) {  // the next unit: next head of this row, or head 0 of the next (row, split)
      const bool wrap = C.h + 1 == a.n_head;
      fetch((wrap && C.split + 1 == a.n_splits) ? C.b + 1 : C.b, wrap ? 0 : C.h + 1);
    }
    if (C.slot < 0) continue;
    const int bh = C.b * a.n_head + C.h;
    const uint32_t k_new = pack_bf16x2(vk.x, vk.y);  // rounded to the cache dtype
    const uint32_t v_new = pack_bf16x2(vv.x, vv.y);
    // q (scaled for exp2) as the A operand of the score MMAs: lanes 0..3 carry row 0 (bf16 high part) and
    // row 8 (low part) of the four 16-dim steps, so c0 + c2 is the score at ~fp32 query precision
    uint32_t qa[4][4];
    {
      const float x0 = vq.x * kScaleLog2, x1 = vq.y * kScaleLog2;
      const __nv_bfloat16 h0 = __float2bfloat16(x0), h1 = __float2bfloat16(x1);
      __nv_bfloat162 hi, lo;
      hi.x = h0, hi.y = h1;
      lo.x = __float2bfloat16(x0 - __bfloat162float(h0));
      lo.y = __float2bfloat16(x1 - __bfloat162float(h1));
      __syncwarp();  // the previous unit's readers of the staging area are done
      reinterpret_cast<__nv_bfloat162*>(sq)[lane] = hi;
      reinterpret_cast<__nv_bfloat162*>(sq + kHeadDim)[lane] = lo;
      __syncwarp();
      const int j2 = 2 * (lane & 3);
      const bool on = lane < 4;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        qa[s][0] = on ? *reinterpret_cast<const uint32_t*>(&sq[16 * s + j2]) : 0u;
        qa[s][1] = on ? *reinterpret_cast<const uint32_t*>(&sq[kHeadDim + 16 * s + j2]) : 0u;
        qa[s][2] = on ? *reinterpret_cast<const uint32_t*>(&sq[16 * s + 8 + j2]) : 0u;
        qa[s][3] = on ? *reinterpret_cast<const uint32_t*>(&sq[kHeadDim + 16 * s + 8 + j2]) : 0u;
      }
    }
    B200_STRACE(1);  // unit set-up

    float m_run = -INFINITY;  // warp-uniform running max (log2 domain)
    float l_part = 0.f;       // this lane's share of the running sum (lanes 0..3 matter)
    float oacc[4][4];
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) oacc[mt][e] = 0.f;
    }
#pragma unroll 1
    for (int t = 0; t < C.n_tiles; ++t) {
      // keep the ring full: the slot this load goes to was consumed in the previous iteration
      __syncwarp();
      if (P.u < u_end) p_issue();
      B200_STRACE(2);  // issuing the next tile
      const uint32_t stage = c_count % kStreamStages;
      mbar_wait(&full[stage], (c_count / kStreamStages) & 1u);
      ++c_count;
      B200_STRACE(3);  // waiting for the tile
      uint4 (*sk)[8] = reinterpret_cast<uint4 (*)[8]>(ring + stage * kStreamTileBytes);
      uint4 (*sv)[8] = reinterpret_cast<uint4 (*)[8]>(ring + stage * kStreamTileBytes + kStreamTile * 128);
      const int rows = min(C.n_tok - t * kStreamTile, kStreamTile);
      if (t == C.n_tiles - 1) {  // the unit's last tile: the new token's row, rows past the last token
        if (C.own_new) {
          const int row = C.n_cached - t * kStreamTile;  // (the page holds no data there yet)
          const int c = (lane >> 2) ^ (row & 7);
          reinterpret_cast<uint32_t*>(&sk[row][c])[lane & 3] = k_new;
          reinterpret_cast<uint32_t*>(&sv[row][c])[lane & 3] = v_new;
          if (!C.fin) {  // ... and it is appended to its page (nothing here reads it back from global memory)
            __syncwarp();
            const size_t dst = ((static_cast<size_t>(apg[stage]) * a.n_head + C.h) * kPageTokens + C.pos % kPageTokens) * kHeadDim + 2 * lane;
            *reinterpret_cast<uint32_t*>(a.k_pool + dst) = k_new;
            *reinterpret_cast<uint32_t*>(a.v_pool + dst) = v_new;
          }
        }
        const int rows_pad = ((rows + kPageTokens - 1) / kPageTokens) * kPageTokens;
        // V rows of the last page that hold no token: 0 * stale must not be NaN
        for (int i = rows * 8 + lane; i < rows_pad * 8; i += 32) sv[i >> 3][i & 7] = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async();  // these writes are ordered before the TMA load that reuses the slot
        __syncwarp();
      }
      // scores = K rows (ldmatrix) x q: 8 tokens per tile; the two 32-dim halves accumulate separately
      // (two dependent MMAs per chain instead of four). Rows without a token are masked below.
      float sacc[2][4][4];
#pragma unroll
      for (int sp = 0; sp < 2; ++sp) {
#pragma unroll
        for (int tt = 0; tt < 4; ++tt) {
#pragma unroll
          for (int e = 0; e < 4; ++e) sacc[sp][tt][e] = 0.f;
          uint32_t kb[4];
          const int row = 8 * tt + (lane & 7);
          const int c = 4 * sp + (lane >> 3);
          ldmatrix_x4(kb, &sk[row][c ^ (row & 7)]);
          mma_bf16_16816(sacc[sp][tt], qa[2 * sp], kb[0], kb[1]);
          mma_bf16_16816(sacc[sp][tt], qa[2 * sp + 1], kb[2], kb[3]);
        }
      }
      float sc[4][2];
      bool ok[4][2];
      float m_loc = -INFINITY;
#pragma unroll
      for (int tt = 0; tt < 4; ++tt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          ok[tt][e] = lane < 4 && (8 * tt + 2 * (lane & 3) + e) < rows;  // lanes 0..3 own the real scores
          sc[tt][e] = (sacc[0][tt][e] + sacc[0][tt][e + 2]) + (sacc[1][tt][e] + sacc[1][tt][e + 2]);
          m_loc = ok[tt][e] ? fmaxf(m_loc, sc[tt][e]) : m_loc;
        }
      }
      // warp maximum with one redux: floats ordered as signed integers
      int key = __float_as_int(m_loc);
      key = key >= 0 ? key : key ^ 0x7fffffff;
      key = __reduce_max_sync(0xffffffffu, key);
      m_loc = __int_as_float(key >= 0 ? key : key ^ 0x7fffffff);
      const float m_new = fmaxf(m_run, m_loc);  // finite: row 0 of a tile is a real token
      const float f = exp2f(m_run - m_new);     // 0 on the first tile
      m_run = m_new;
      l_part *= f;
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[mt][e] *= f;
      }
      float p[4][2];
#pragma unroll
      for (int tt = 0; tt < 4; ++tt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          p[tt][e] = ok[tt][e] ? exp2f(sc[tt][e] - m_new) : 0.f;
          l_part += p[tt][e];
        }
      }
      // output = V^T (ldmatrix.trans) x weights, which sit in the B-fragment layout the score tiles left them in
#pragma unroll
      for (int uu = 0; uu < kTilePages; ++uu) {
        if (16 * uu < rows) {  // warp-uniform: a page that was not loaded is never read
          const uint32_t b0 = pack_bf16x2(p[2 * uu][0], p[2 * uu][1]);
          const uint32_t b1 = pack_bf16x2(p[2 * uu + 1][0], p[2 * uu + 1][1]);
#pragma unroll
          for (int mt = 0; mt < 4; ++mt) {  // 16 dims per tile
            uint32_t va[4];
            const int row = 16 * uu + ((lane >> 4) << 3) + (lane & 7);
            const int c = 2 * mt + ((lane >> 3) & 1);
            ldmatrix_x4_trans(va, &sv[row][c ^ (row & 7)]);
            mma_bf16_16816(oacc[mt], va, b0, b1);
          }
        }
      }
      B200_STRACE(4);  // math
    }

    // the unit's result: column 0 of the output tiles lives in lanes 0, 4, ..., 28; through shared memory
    // to lane l = dims 2l, 2l + 1
    float l_unit = l_part + __shfl_xor_sync(0xffffffffu, l_part, 1);
    l_unit += __shfl_xor_sync(0xffffffffu, l_unit, 2);
    l_unit = __shfl_sync(0xffffffffu, l_unit, 0);
    __syncwarp();
    if ((lane & 3) == 0) {
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        so[16 * mt + (lane >> 2)] = oacc[mt][0];
        so[16 * mt + 8 + (lane >> 2)] = oacc[mt][2];
      }
    }
    __syncwarp();
    const float2 o2 = *reinterpret_cast<const float2*>(&so[2 * lane]);
    if (a.n_splits == 1) {
      const float inv = 1.f / l_unit;  // a unit of the only split always holds the new token
      *reinterpret_cast<uint32_t*>(a.out + static_cast<size_t>(C.b) * d + C.h * kHeadDim + 2 * lane) =
          pack_bf16x2(o2.x * inv, o2.y * inv);
    } else {
      float* sc0 = a.scratch + static_cast<size_t>(bh) * a.n_splits * kScratchStride;
      float* sc = sc0 + C.split * kScratchStride;
      *reinterpret_cast<float2*>(&sc[2 * lane]) = o2;  // an empty split leaves (0, -inf, 0)
      if (lane == 0) {
        sc[kHeadDim] = m_run;
        sc[kHeadDim + 1] = l_unit;
      }
      if (a.split_counters != nullptr) {
        // The split that finishes last merges all of them, in split order (so the result does not depend
        // on who that is), and leaves the counter at zero for the next launch: no combine kernel.
        __threadfence();
        __syncwarp();
        int prev = 0;
        if (lane == 0) prev = atomicAdd(a.split_counters + bh, 1);
        prev = __shfl_sync(0xffffffffu, prev, 0);
        if (prev == a.n_splits - 1) {
          __threadfence();
          float m = -INFINITY;
          for (int s = 0; s < a.n_splits; ++s) m = fmaxf(m, __ldcg(sc0 + s * kScratchStride + kHeadDim));
          float l = 0.f;
          float2 o = make_float2(0.f, 0.f);
          for (int s = 0; s < a.n_splits; ++s) {
            const float ms = __ldcg(sc0 + s * kScratchStride + kHeadDim);
            const float fs = (ms == -INFINITY) ? 0.f : exp2f(ms - m);
            const float2 os = __ldcg(reinterpret_cast<const float2*>(sc0 + s * kScratchStride + 2 * lane));
            l = fmaf(__ldcg(sc0 + s * kScratchStride + kHeadDim + 1), fs, l);
            o.x = fmaf(os.x, fs, o.x);
            o.y = fmaf(os.y, fs, o.y);
          }
          const float inv = 1.f / l;
          *reinterpret_cast<uint32_t*>(a.out + static_cast<size_t>(C.b) * d + C.h * kHeadDim + 2 * lane) =
              pack_bf16x2(o.x * inv, o.y * inv);
          if (lane == 0) a.split_counters[bh] = 0;
        }
      }
    }
    B200_STRACE(5);  // result
  }
  B200_STRACE_END;
  B200_ATRACE_CTA(2);
}

int launch_with_pdl(void (*kernel)(DecodeAttnArgs), dim3 grid, dim3 block, cudaStream_t stream,
                    int pdl, const DecodeAttnArgs& a) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, a));
  return kOk;
}

}  // namespace

int prefill_attention(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int batch,
                      int max_seq_len, int n_head, __nv_bfloat16* out, cudaStream_t stream,
                      const PrefillCache* cache) {
  B200_CHECK_ARG(batch > 0 && max_seq_len > 0 && n_head > 0, "prefill_attention: bad shape");
  dim3 grid((max_seq_len + kBQ - 1) / kBQ, n_head, batch);
  prefill_attn_kernel<false><<<grid, 128, 0, stream>>>(qkv, cu_seqlens, n_head, out,
                                                       cache ? *cache : PrefillCache{}, nullptr);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

int prefill_attention_paged(const PagedPrefillArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.max_suffix_len > 0 && a.n_head > 0, "prefill_attention_paged: bad shape");
  PrefillCache pc;
  pc.slots = a.slots;
  pc.page_table = a.page_table;
  pc.max_pages_per_seq = a.max_pages_per_seq;
  pc.k_pool = const_cast<__nv_bfloat16*>(a.k_pool);
  pc.v_pool = const_cast<__nv_bfloat16*>(a.v_pool);
  dim3 grid((a.max_suffix_len + kBQ - 1) / kBQ, a.n_head, a.batch);
  prefill_attn_kernel<true><<<grid, 128, 0, stream>>>(a.qkv, a.cu_seqlens, a.n_head, a.out, pc, a.prefix_len);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

namespace {
// One thread per 16-byte chunk of a row's K and V (2 * d / 8 chunks per row).
__global__ void kv_scatter_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ row_pos,
                                  const int32_t* __restrict__ row_slot, int n_tok, int n_head,
                                  const int32_t* __restrict__ page_table, int max_pages_per_seq,
                                  __nv_bfloat16* __restrict__ k_pool, __nv_bfloat16* __restrict__ v_pool) {
  const int d = n_head * kHeadDim;
  const int chunks_per_row = 2 * d / 8;
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<long long>(n_tok) * chunks_per_row) return;
  const int t = static_cast<int>(idx / chunks_per_row);
  const int c = static_cast<int>(idx % chunks_per_row);
  const int which = c / (d / 8);          // 0 = K, 1 = V
  const int col = (c % (d / 8)) * 8;      // feature within d
  const int h = col / kHeadDim, dim = col % kHeadDim;
  const int pos = row_pos[t];
  const int page = page_table[static_cast<size_t>(row_slot[t]) * max_pages_per_seq + pos / kPageTokens];
  const size_t dst = ((static_cast<size_t>(page) * n_head + h) * kPageTokens + pos % kPageTokens) * kHeadDim + dim;
  const uint4 v = *reinterpret_cast<const uint4*>(qkv + static_cast<size_t>(t) * 3 * d + (1 + which) * d + col);
  *reinterpret_cast<uint4*>((which ? v_pool : k_pool) + dst) = v;
}
}  // namespace

int kv_scatter(const __nv_bfloat16* qkv, const int32_t* row_pos, const int32_t* row_slot, int n_tok, int n_head,
               const int32_t* page_table, int max_pages_per_seq, __nv_bfloat16* k_pool, __nv_bfloat16* v_pool,
               cudaStream_t stream) {
  B200_CHECK_ARG(n_tok > 0 && n_head > 0, "kv_scatter: bad shape");
  const long long n = static_cast<long long>(n_tok) * (2 * n_head * kHeadDim / 8);
  kv_scatter_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(
      qkv, row_pos, row_slot, n_tok, n_head, page_table, max_pages_per_seq, k_pool, v_pool);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

size_t decode_attention_scratch_floats(int batch, int n_head, int n_splits) {
  if (n_splits <= 1) return 0;
  return static_cast<size_t>(batch) * n_head * n_splits * kScratchStride;
}

// Split-KV partials for callers that must not allocate. The array is part of the library's
// device image: every GPU that loads libb200llm.so carries these 4 MiB until the process exits.
// In host code the name below only stands for the symbol; the device address comes from
// cudaGetSymbolAddress (done once per device, with that device current).
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

int decode_attention_noalloc(DecodeAttnArgs a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.n_head > 0 && a.n_splits >= 1, "decode_attention_noalloc: bad shape");
  const size_t need = decode_attention_scratch_floats(a.batch, a.n_head, a.n_splits);
  if (need > kNoAllocScratchFloats) {
    return fail(kErrCapacity,
                "decode_attention_noalloc: batch*n_head*n_splits*66 = " + std::to_string(need) +
                    " floats of split-KV partials exceed the built-in scratch of " +
                    std::to_string(kNoAllocScratchFloats) +
                    " floats; use fewer splits or the entry point that takes a scratch buffer");
  }
  static std::mutex mu;
  static float* resolved[64] = {};  // per device ordinal
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < 64, "decode_attention_noalloc: unsupported device ordinal");
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (resolved[dev] == nullptr) {
      void* p = nullptr;
      const cudaError_t err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (err != cudaSuccess || p == nullptr) {
        return fail(kErrCuda, std::string("decode_attention_noalloc: cudaGetSymbolAddress: ") +
                                  cudaGetErrorString(err));
      }
      resolved[dev] = static_cast<float*>(p);
    }
    scratch = resolved[dev];
  }
  // Nothing in the scratch has to start at zero: every (row, head, split) slot that the
  // combine pass reads is written by its own CTA first, empty splits included.
  a.scratch = scratch;
  return decode_attention(a, stream);
}

B200_DEFINE_STEP_TRACE_SETTER(step_trace_set_attention)

#ifdef B200_ATTN_TRACE
}  // namespace b200llm
// Profiling builds only: read CTA 0's milestone timestamps of the last decode_attn_kernel launch.
extern "C" int b200llm_debug_attn_trace(unsigned long long* out16) {
  return cudaMemcpyFromSymbol(out16, g_attn_trace, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : 2;
}
extern "C" int b200llm_debug_attn_cta_trace(unsigned long long* out, int n_ctas) {
  if (n_ctas > 4096) n_ctas = 4096;
  return cudaMemcpyFromSymbol(out, g_attn_cta_trace, static_cast<size_t>(n_ctas) * 4 * sizeof(unsigned long long)) ==
                 cudaSuccess
             ? 0
             : 2;
}
namespace b200llm {
#endif

template <int kWarps, int kStages, int kTile, int kCtasPerSm>
int launch_stream(const CUtensorMap& map_k, const CUtensorMap& map_v, const DecodeAttnArgs& a, cudaStream_t stream) {
  auto kernel = decode_attn_stream_kernel<kWarps, kStages, kTile, kCtasPerSm>;
  constexpr int smem = StreamCfg<kWarps, kStages, kTile>::kSmemBytes;
  B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  // Persistent: every SM gets its full set of CTAs when there are units enough (warps of one SM share its
  // bandwidth and issue slots, so what has to be even is the work per SM, not per warp).
  const long long units = static_cast<long long>(a.batch) * a.n_head * a.n_splits;
  const long long want = (units + kWarps - 1) / kWarps;
  const long long cap = static_cast<long long>(device_sm_count()) * kCtasPerSm;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(want < cap ? want : cap));
  cfg.blockDim = dim3(kWarps * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = a.pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, map_k, map_v, a));
  return kOk;
}

int decode_attention_stream(const DecodeAttnArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.n_pages > 0, "decode_attention: the streaming kernel needs n_pages (extent of its tensor maps)");
  const long long pool_rows = static_cast<long long>(a.n_pages) * a.n_head * kPageTokens;
  B200_CHECK_ARG(pool_rows < (1ll << 31), "decode_attention: KV pool too large for one tensor map");
  CUtensorMap map_k, map_v;
  B200_PROPAGATE(gemm_tensor_map(a.k_pool, static_cast<int>(pool_rows), kHeadDim, kPageTokens, &map_k));
  B200_PROPAGATE(gemm_tensor_map(a.v_pool, static_cast<int>(pool_rows), kHeadDim, kPageTokens, &map_v));
  // stream_kernel selects the instance (the engine's rule is next to its plan): 1 = four warps x two-tile
  // rings, 2 = three warps x three-tile rings; the others exist in tune / profiling builds only
  switch (a.stream_kernel) {
    case 2: B200_PROPAGATE((launch_stream<3, 3, 32, 3>(map_k, map_v, a, stream))); break;
#ifdef B200_TUNE_KNOBS
    case 3: B200_PROPAGATE((launch_stream<4, 3, 32, 2>(map_k, map_v, a, stream))); break;
    case 4: B200_PROPAGATE((launch_stream<2, 2, 32, 6>(map_k, map_v, a, stream))); break;
    case 5: B200_PROPAGATE((launch_stream<2, 4, 32, 3>(map_k, map_v, a, stream))); break;
#endif
    default: B200_PROPAGATE((launch_stream<4, 2, 32, 3>(map_k, map_v, a, stream))); break;
  }
  if (a.n_splits > 1 && a.split_counters == nullptr) {
    B200_PROPAGATE(
        launch_with_pdl(decode_combine_kernel, dim3(a.batch * a.n_head), dim3(64), stream, a.pdl, a));
  }
  return kOk;
}

int decode_attention(const DecodeAttnArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.n_head > 0 && a.n_splits >= 1, "decode_attention: bad shape");
  B200_CHECK_ARG(a.n_splits == 1 || a.scratch != nullptr,
                 "decode_attention: split-KV needs a scratch buffer");
  dim3 grid(a.batch * a.n_head, a.n_splits);
  DecodeAttnArgs args = a;
  if (args.stream_kernel != 0) return decode_attention_stream(args, stream);
  if (args.n_splits < 2 || args.n_splits > 8) args.cluster_combine = 0;
  if (args.cluster_combine != 0) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(128);
    cfg.stream = stream;
    cudaLaunchAttribute attrs[2];
    attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[0].val.programmaticStreamSerializationAllowed = a.pdl ? 1 : 0;
    attrs[1].id = cudaLaunchAttributeClusterDimension;
    attrs[1].val.clusterDim.x = 1;
    attrs[1].val.clusterDim.y = static_cast<unsigned>(args.n_splits);
    attrs[1].val.clusterDim.z = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = 2;
    B200_CUDA_OK(cudaLaunchKernelEx(&cfg, decode_attn_kernel<true>, args));
    return kOk;
  }
  B200_PROPAGATE(launch_with_pdl(decode_attn_kernel<false>, grid, dim3(128), stream, a.pdl, args));
  if (a.n_splits > 1) {
    B200_PROPAGATE(
        launch_with_pdl(decode_combine_kernel, dim3(a.batch * a.n_head), dim3(64), stream, a.pdl, a));
  }
  return kOk;
}

}  // namespace b200llm
