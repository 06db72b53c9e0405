// Attention kernels (see attention.cuh for the oracle citations).
//
// prefill_attn_kernel : flash-style causal attention over packed ragged prompts. One CTA =
//   64 query rows of one (sequence, head); 4 warps x 16 rows; K/V tiles of 64 keys staged in
//   padded shared memory; QK^T and PV on the tensor cores (mma.sync m16n8k16 bf16, fp32
//   accumulate); online softmax in fp32 registers. This is <= 8 % of prefill FLOPs for the
//   BASELINE configs; the tcgen05 path is used for the GEMMs that dominate.
//
// decode_attn_kernel : one new token per sequence against the paged KV cache. HBM-bound:
//   every K/V byte is read exactly once with 16-byte no-allocate loads, 8 lanes per cached
//   token (one 128-byte row), 16 tokens (= one 2 KB page of a head) per CTA iteration, 4-way
//   unrolled so each thread keeps 8 independent 16-byte loads in flight; softmax statistics
//   by warp shuffles. Optional split-KV for long contexts with a deterministic combine pass.
#include "attention.cuh"

#include <mutex>

#include "common.cuh"
#include "ops.cuh"

#ifdef B200_ATTN_TRACE
// Profiling build only (profiles/attn_trace.cu): CTA 0 records SM-clock milestones.
__device__ unsigned long long g_attn_trace[16];
#define B200_ATRACE(slot)                                                                  \
  do {                                                                                     \
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_attn_trace[slot] = clock64(); \
  } while (0)
#else
#define B200_ATRACE(slot) \
  do {                    \
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

__global__ void __launch_bounds__(128)
prefill_attn_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ cu_seqlens,
                    int n_head, __nv_bfloat16* __restrict__ out, PrefillCache cache) {
  __shared__ __align__(16) __nv_bfloat16 sQ[kBQ][kLds];
  __shared__ __align__(16) __nv_bfloat16 sKb[2][kBKV][kLds];  // double-buffered K/V tiles
  __shared__ __align__(16) __nv_bfloat16 sVb[2][kBKV][kLds];

  const int b = blockIdx.z;
  const int h = blockIdx.y;
  const int qt = blockIdx.x;
  const int row0 = cu_seqlens[b];
  const int len = cu_seqlens[b + 1] - row0;
  if (qt * kBQ >= len) return;
  const int d = n_head * kHeadDim;
  const int ld = 3 * d;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;  // row within the 8-row group
  const int t = lane & 3;   // column pair

  const __nv_bfloat16* q_base = qkv + static_cast<size_t>(row0 + qt * kBQ) * ld + h * kHeadDim;
  const __nv_bfloat16* kv_base = qkv + static_cast<size_t>(row0) * ld + d + h * kHeadDim;
  // Q and the first K/V tile in one cp.async group
  load_tile_64x64_async(sQ, q_base, ld, len - qt * kBQ);
  load_tile_64x64_async(sKb[0], kv_base, ld, len);
  load_tile_64x64_async(sVb[0], kv_base + d, ld, len);
  asm volatile("cp.async.commit_group;" ::: "memory");

  uint32_t qf[4][4];

  float o[8][4];
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  }
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const int qrow_lo = qt * kBQ + warp * 16 + g;  // this thread's rows: qrow_lo, qrow_lo + 8

  for (int kt = 0; kt <= qt; ++kt) {
    __nv_bfloat16 (*sK)[kLds] = sKb[kt & 1];
    __nv_bfloat16 (*sV)[kLds] = sVb[kt & 1];
    // prefetch the next tile into the other buffer (its last readers finished at the barrier
    // that closed the previous iteration), then wait for the current one
    if (kt + 1 <= qt) {
      const __nv_bfloat16* nk = kv_base + static_cast<size_t>((kt + 1) * kBKV) * ld;
      load_tile_64x64_async(sKb[(kt + 1) & 1], nk, ld, len - (kt + 1) * kBKV);
      load_tile_64x64_async(sVb[(kt + 1) & 1], nk + d, ld, len - (kt + 1) * kBKV);
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
    if (kt == qt && cache.slots != nullptr) {
      // diagonal tile: this CTA fills the KV pages ([page][head][16 tokens][64]) from smem
      const int32_t* pt = cache.page_table + static_cast<size_t>(cache.slots[b]) * cache.max_pages_per_seq;
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
    const bool need_mask = (kt == qt) || ((kt + 1) * kBKV > len);
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
      __nv_bfloat16* orow = out + static_cast<size_t>(row0 + row) * d + h * kHeadDim;
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
struct Softmax8 {
  float m;     // running max (log2 domain)
  float l;     // running sum
  float o[8];  // running weighted V for this lane's 8 dims
};

__device__ __forceinline__ void merge_into(Softmax8& a, float m2, float l2, const float (&o2)[8]) {
  const float m = fmaxf(a.m, m2);
  const float f1 = (a.m == -INFINITY) ? 0.f : exp2f(a.m - m);
  const float f2 = (m2 == -INFINITY) ? 0.f : exp2f(m2 - m);
  a.l = a.l * f1 + l2 * f2;
#pragma unroll
  for (int i = 0; i < 8; ++i) a.o[i] = a.o[i] * f1 + o2[i] * f2;
  a.m = m;
}

__device__ __forceinline__ void bf16x8_to_float(const uint4& v, float (&f)[8]) {
  float2 a = unpack_bf16x2(v.x), b = unpack_bf16x2(v.y), c = unpack_bf16x2(v.z),
         d = unpack_bf16x2(v.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
  f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

constexpr int kDecUnroll = 3;
constexpr int kScratchStride = kHeadDim + 2;  // o[64], m, l
constexpr int kMaxSmemPages = 256;            // page ids staged in smem (4096 tokens)
constexpr int kStageTokens = 128;             // cached tokens staged in smem per CTA (32 KB)

struct KvBatch {
  uint4 k[kDecUnroll];
  uint4 v[kDecUnroll];
};

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// Online-softmax update with N cached tokens at once: the N dot products, their shuffle
// reductions and exponentials are independent, so one pass through the dependent chain
// (max -> rescale -> accumulate) serves N tokens instead of one.
template <int N>
__device__ __forceinline__ void attend_tokens(Softmax8& acc, const float (&q)[8], const uint4 (&kraw)[N],
                                              const uint4 (&vraw)[N], const bool (&valid)[N]) {
  float s[N];
#pragma unroll
  for (int n = 0; n < N; ++n) {
    float kf[8];
    bf16x8_to_float(kraw[n], kf);
    float d = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) d = fmaf(q[i], kf[i], d);
    s[n] = d;
  }
#pragma unroll
  for (int o = 1; o <= 4; o <<= 1) {
#pragma unroll
    for (int n = 0; n < N; ++n) s[n] += __shfl_xor_sync(0xffffffffu, s[n], o);
  }
  float m = acc.m;
#pragma unroll
  for (int n = 0; n < N; ++n) m = valid[n] ? fmaxf(m, s[n]) : m;
  if (m == -INFINITY) return;  // nothing valid yet (uniform within the 8-lane group)
  const float f = (acc.m == -INFINITY) ? 0.f : exp2f(acc.m - m);
  float p[N];
  float psum = 0.f;
#pragma unroll
  for (int n = 0; n < N; ++n) {
    p[n] = valid[n] ? exp2f(s[n] - m) : 0.f;
    psum += p[n];
  }
  acc.l = acc.l * f + psum;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc.o[i] *= f;
#pragma unroll
  for (int n = 0; n < N; ++n) {
    float vf[8];
    bf16x8_to_float(vraw[n], vf);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc.o[i] = fmaf(p[n], vf[i], acc.o[i]);
  }
  acc.m = m;
}

// Ordering argument for everything this kernel does BEFORE griddepcontrol.wait.
// In a decode step the GEMM, LayerNorm and gelu-reduce kernels signal launch_dependents at entry;
// the embed kernel (first of every step), this kernel and the selector signal only AFTER their own
// wait has returned, i.e. after their predecessor has completely finished. A CTA of this kernel
// can therefore only be running if this step's embed kernel has passed its wait, which means the
// previous step's selector -- and with it the whole previous step -- is complete and visible:
// seq_len / finished are final, and KV positions < seq_len (written by earlier steps) are
// immutable. Only the QKV projection of THIS step is not ready. So the whole cached KV stream is
// fetched before the wait and overlaps the QKV GEMM.
__global__ void __launch_bounds__(128, 6) decode_attn_kernel(DecodeAttnArgs a) {
  __shared__ float sm_o[4][kHeadDim];
  __shared__ float sm_ml[4][2];
  __shared__ int32_t sm_pages[kMaxSmemPages];
  __shared__ __align__(16) float sm_q[kHeadDim];              // this token's q (pre-scaled)
  __shared__ __align__(16) __nv_bfloat16 sm_kv[2][kHeadDim];  // this token's k, v (cache dtype)
  __shared__ __align__(16) uint4 sm_k[kStageTokens][8];       // staged cached K rows
  __shared__ __align__(16) uint4 sm_v[kStageTokens][8];       // staged cached V rows

  const int bh = blockIdx.x;
  const int b = bh / a.n_head;
  const int h = bh % a.n_head;
  const int split = blockIdx.y;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int sub = lane & 7;                // which 16-byte chunk (8 dims) of the 128-byte row
  const int grp = warp * 4 + (lane >> 3);  // 0..15: token within the 16-token page
  const int d = a.n_head * kHeadDim;
  B200_ATRACE(0);
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
  const int t_end = min(ctx, t_begin + chunk);        // may include `pos` (the new token)
  const int t_cached_end = min(t_end, pos);           // cached tokens only
  const int n_staged = max(0, min(t_cached_end - t_begin, kStageTokens));

  auto page_of = [&](int tpos) -> int {
    const int pi = tpos / kPageTokens;
    return pi < n_stage_pages ? sm_pages[pi] : pt[pi];
  };
  auto kv_offset = [&](int tpos) -> size_t {
    return ((static_cast<size_t>(page_of(tpos)) * a.n_head + h) * kPageTokens + tpos % kPageTokens) * kHeadDim;
  };
  // (1) the first kStageTokens cached tokens: global -> shared with cp.async, no registers held
  for (int i = threadIdx.x; i < n_staged * 16; i += blockDim.x) {
    const int tok = i >> 4, part = i & 15;
    const size_t off = kv_offset(t_begin + tok) + (part & 7) * 8;
    if (part < 8) {
      cp_async_16(&sm_k[tok][part], a.k_pool + off);
    } else {
      cp_async_16(&sm_v[tok][part - 8], a.v_pool + off);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  // (2) the remaining tokens stream through registers; their first batch is requested now and
  //     the rest is pulled into L2
  const int t_rest = t_begin + n_staged;
  auto load_batch = [&](int base, KvBatch& kb) {
#pragma unroll
    for (int u = 0; u < kDecUnroll; ++u) {
      const int tpos = base + u * kPageTokens + grp;
      kb.k[u] = make_uint4(0u, 0u, 0u, 0u);
      kb.v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (tpos < t_cached_end) {
        const size_t off = kv_offset(tpos) + sub * 8;
        kb.k[u] = ld_nc_v4(a.k_pool + off);
        kb.v[u] = ld_nc_v4(a.v_pool + off);
      }
    }
  };
  KvBatch cur;
  load_batch(t_rest, cur);
  if (sub < 2) {
    const __nv_bfloat16* pool = sub == 0 ? a.k_pool : a.v_pool;
    for (int tpos = t_rest + kPageTokens * kDecUnroll + grp; tpos < t_cached_end; tpos += kPageTokens) {
      asm volatile("prefetch.global.L2 [%0];" ::"l"(pool + kv_offset(tpos)));
    }
  }

  B200_ATRACE(1);
  pdl_wait();
  pdl_launch_dependents();
  B200_ATRACE(2);

  // This token's q/k/v. With a split-K projection the CTA reduces the fp32 partial planes
  // (fixed summation order -> deterministic), adds the bias, keeps q in fp32 and rounds k/v to
  // the cache dtype; everyone then reads them from shared memory.
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
          sm_q[dim] = val[r] * kScaleLog2;
        } else {
          sm_kv[part - 1][dim] = __float2bfloat16(val[r]);
        }
      }
    }
  }
  B200_ATRACE(7);
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  B200_ATRACE(8);
  __syncthreads();
  B200_ATRACE(4);
  float q[8];
  {
    const float4 q0 = *reinterpret_cast<const float4*>(&sm_q[sub * 8]);
    const float4 q1 = *reinterpret_cast<const float4*>(&sm_q[sub * 8 + 4]);
    q[0] = q0.x; q[1] = q0.y; q[2] = q0.z; q[3] = q0.w;
    q[4] = q1.x; q[5] = q1.y; q[6] = q1.z; q[7] = q1.w;
  }
  const uint4 k_new = *reinterpret_cast<const uint4*>(&sm_kv[0][sub * 8]);
  const uint4 v_new = *reinterpret_cast<const uint4*>(&sm_kv[1][sub * 8]);

  // append this token to its page (the split that owns position `pos` does it)
  if (pos >= t_begin && pos < t_end && !fin && warp == 0 && lane < 8) {
    const size_t dst = kv_offset(pos) + sub * 8;
    *reinterpret_cast<uint4*>(a.k_pool + dst) = k_new;
    *reinterpret_cast<uint4*>(a.v_pool + dst) = v_new;
  }

  Softmax8 acc;
  acc.m = -INFINITY;
  acc.l = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc.o[i] = 0.f;

  // staged tokens (shared memory; a warp's 4 groups read 4 consecutive 128-byte rows), 4 at a time
  // (trip count is CTA-uniform: the shuffles inside need every lane of the warp)
  for (int it0 = 0; it0 < n_staged; it0 += 4 * kPageTokens) {
    uint4 kk[4], vv[4];
    bool ok[4];
#pragma unroll
    for (int n = 0; n < 4; ++n) {
      const int i = it0 + grp + n * kPageTokens;
      ok[n] = i < n_staged;
      const int ii = ok[n] ? i : 0;
      kk[n] = sm_k[ii][sub];
      vv[n] = sm_v[ii][sub];
    }
    attend_tokens<4>(acc, q, kk, vv, ok);
  }
  // streamed tokens; the new token rides along with the last batch
  const bool own_new = pos >= t_begin && pos < t_end;
  bool new_done = !own_new;
  for (int base = t_rest; base < t_cached_end || !new_done; base += kPageTokens * kDecUnroll) {
    if (base != t_rest && base < t_cached_end) load_batch(base, cur);
    uint4 kk[kDecUnroll + 1], vv[kDecUnroll + 1];
    bool ok[kDecUnroll + 1];
#pragma unroll
    for (int u = 0; u < kDecUnroll; ++u) {
      kk[u] = cur.k[u];
      vv[u] = cur.v[u];
      ok[u] = base + u * kPageTokens + grp < t_cached_end;
    }
    const bool last = base + kPageTokens * kDecUnroll >= t_cached_end;
    kk[kDecUnroll] = k_new;
    vv[kDecUnroll] = v_new;
    ok[kDecUnroll] = last && own_new && grp == 0;
    attend_tokens<kDecUnroll + 1>(acc, q, kk, vv, ok);
    if (last) new_done = true;
  }

  B200_ATRACE(5);
  // merge the 4 token groups of this warp (lanes differing in bits 3 and 4)
#pragma unroll
  for (int off = 8; off <= 16; off <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, acc.m, off);
    const float l2 = __shfl_xor_sync(0xffffffffu, acc.l, off);
    float o2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o2[i] = __shfl_xor_sync(0xffffffffu, acc.o[i], off);
    merge_into(acc, m2, l2, o2);
  }
  if (lane < 8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) sm_o[warp][sub * 8 + i] = acc.o[i];
    if (lane == 0) {
      sm_ml[warp][0] = acc.m;
      sm_ml[warp][1] = acc.l;
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
    } else {
      float* sc = a.scratch + (static_cast<size_t>(bh) * a.n_splits + split) * kScratchStride;
      sc[dim] = o;
      if (dim == 0) {
        sc[kHeadDim] = m;
        sc[kHeadDim + 1] = l;
      }
    }
  }
  B200_ATRACE(6);
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
  prefill_attn_kernel<<<grid, 128, 0, stream>>>(qkv, cu_seqlens, n_head, out,
                                                cache ? *cache : PrefillCache{});
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
namespace b200llm {
#endif

int decode_attention(const DecodeAttnArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.n_head > 0 && a.n_splits >= 1, "decode_attention: bad shape");
  B200_CHECK_ARG(a.n_splits == 1 || a.scratch != nullptr,
                 "decode_attention: split-KV needs a scratch buffer");
  dim3 grid(a.batch * a.n_head, a.n_splits);
  B200_PROPAGATE(launch_with_pdl(decode_attn_kernel, grid, dim3(128), stream, a.pdl, a));
  if (a.n_splits > 1) {
    B200_PROPAGATE(
        launch_with_pdl(decode_combine_kernel, dim3(a.batch * a.n_head), dim3(64), stream, a.pdl, a));
  }
  return kOk;
}

}  // namespace b200llm
