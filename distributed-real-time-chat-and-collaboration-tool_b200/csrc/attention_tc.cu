// Prefill attention on the 5th-generation tensor cores.
//
// Oracle: eager_attention_forward (modeling_gpt2.py:54-72): softmax(Q K^T / sqrt(64) + causal
// mask) V with the softmax in fp32; the past keys/values come first (cache append, :194-196).
//
// One CTA = 128 query rows of one (sequence, head). Warps 0-3 (128 threads) own one query row
// each (= one TMEM lane: a row's softmax needs no shuffles); warp 4 moves data and issues the
// MMAs. The keys of a row are, in order,
//   (A) the sequence's cached prefix, read from the KV pages ([page][head][16 tokens][64] bf16;
//       shared-prefix reuse) in tiles of 128 keys = 8 page boxes of 16 x 64, lanes 0-7 of warp 4
//       fetching one page each, and
//   (B) this prefill's own rows, read from the packed qkv buffer in tiles of 128 rows (one TMA
//       box for K, one for V), causal within the tile on the diagonal.
// A cold prompt has only (B). The CTA on the diagonal also writes its K/V tile into the
// sequence's KV pages, so prefill needs no scatter pass.
//
//   S = Q K^T : tcgen05.mma 128 x 128 x 16 (x4)   -> TMEM columns [0, 128)  fp32
//   softmax   : tcgen05.ld of the thread's own row, running max / sum in registers,
//               P -> bf16 -> shared memory in the K-major SWIZZLE_128B layout the next MMA reads
//   O_t = P V : tcgen05.mma 128 x 64 x 16 (x8), V as an MN-major B operand -> TMEM [128, 192)
//   O = O * alpha + O_t in registers (64 fp32 per thread), so no TMEM rescale pass.
//
// Pipeline (mbarriers only, no __syncthreads in the loop): K/V tiles are double-buffered, the
// next tile's TMA is issued as soon as the MMAs that read its buffer have retired, and S of tile
// t+1 is issued right behind P V of tile t, so it is computed while the row threads fold O_t
// into their accumulators. One CTA needs 112 KB of shared memory and 256 TMEM columns: two CTAs
// share an SM, and one's softmax overlaps the other's MMAs.
#include <mutex>
#include <type_traits>

#include "attention.cuh"
#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "ops.cuh"

namespace b200llm {

namespace {

constexpr int kTQ = 128;                 // query rows per CTA
constexpr int kTK = 128;                 // keys per tile
constexpr int kTileBytes = 128 * 128;    // 128 rows x 64 bf16
constexpr int kPageBytes = kPageTokens * kHeadDim * 2;  // 2 KB: one (page, head)
constexpr int kPagesPerTile = kTK / kPageTokens;
constexpr uint32_t kTmemCols = 256;      // S: 128, O tile: 64 (power of two >= 192)
constexpr float kScaleLog2 = 0.125f * 1.4426950408889634f;  // 1/sqrt(64) folded into exp2

struct TcSmem {
  uint8_t q[kTileBytes];
  uint8_t k[2][kTileBytes];
  uint8_t v[2][kTileBytes];
  uint8_t p[2 * kTileBytes];  // [128 q x 128 keys] bf16 as two 64-key column blocks
  uint64_t q_bar;
  uint64_t kv_bar[2];  // TMA -> MMA issuer (and the diagonal page fill)
  uint64_t s_bar;      // S = Q K^T complete
  uint64_t p_bar;      // 128 row threads: P written, S and O_t of the previous tile consumed
  uint64_t o_bar;      // O_t = P V complete (K/V buffer and P reusable)
  uint32_t tmem_slot;
};

// Sequences of at most one key tile and no cached prefix (prompts of <= 128 tokens): one K/V buffer,
// P overwrites Q and K once S is complete, and O_t reuses the TMEM columns of S once every row has
// read it. 48 KB + 128 TMEM columns + no 64-float accumulator per thread: four CTAs per SM instead of
// two, which is what a CTA that is all set-up, one tile and tear-down needs.
struct TcSmemSingle {
  uint8_t q[kTileBytes];   // then P, first half
  uint8_t k[kTileBytes];   // then P, second half
  uint8_t v[kTileBytes];
  uint64_t q_bar;
  uint64_t kv_bar[2];
  uint64_t s_bar;
  uint64_t p_bar;
  uint64_t o_bar;
  uint32_t tmem_slot;
};

struct TcArgs {
  const int32_t* cu_seqlens;
  const int32_t* prefix_len;  // may be null (no cached prefixes)
  const int32_t* slots;       // may be null with page_table null: nothing cached, nothing filled
  const int32_t* page_table;
  int max_pages_per_seq;
  int n_head;
  __nv_bfloat16* k_pool;
  __nv_bfloat16* v_pool;
  __nv_bfloat16* out;
};

// MN-major bf16 operand written by TMA with SWIZZLE_128B as rows of 64 elements along MN (128 B),
// one row per K index: atoms of 8 K-rows (1024 B) follow each other along K (SBO); a second 64-wide
// column block along MN would sit LBO away (not used: N = 64).
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(kTileBytes >> 4) << 16;  // LBO (no second MN block exists)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;        // SBO = 1024 B between 8-row groups along K
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int kRowThreads = 128;
constexpr int kTcThreads = kRowThreads + 32;

template <bool kSingle>
__global__ void __launch_bounds__(kTcThreads, kSingle ? 4 : 2)
prefill_attn_tc_kernel(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_k,
                       const __grid_constant__ CUtensorMap map_v, const TcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  using Smem = typename std::conditional<kSingle, TcSmemSingle, TcSmem>::type;
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  uint8_t* const q_s = sm.q;
  uint8_t* const p_s = kSingle ? sm.q : reinterpret_cast<TcSmem*>(smem_raw)->p;
  auto k_s = [&](int buf) -> uint8_t* {
    if constexpr (kSingle) return reinterpret_cast<TcSmemSingle*>(smem_raw)->k;
    else return reinterpret_cast<TcSmem*>(smem_raw)->k[buf];
  };
  auto v_s = [&](int buf) -> uint8_t* {
    if constexpr (kSingle) return reinterpret_cast<TcSmemSingle*>(smem_raw)->v;
    else return reinterpret_cast<TcSmem*>(smem_raw)->v[buf];
  };
  constexpr uint32_t kCols = kSingle ? 128u : kTmemCols;

  const int b = blockIdx.z, h = blockIdx.y, qt = blockIdx.x;
  const int row0 = a.cu_seqlens[b];
  const int len_q = a.cu_seqlens[b + 1] - row0;
  if (qt * kTQ >= len_q) return;  // whole CTA, before anything is allocated
  const int prefix = a.prefix_len ? a.prefix_len[b] : 0;
  const int d = a.n_head * kHeadDim;
  const int32_t* pt = a.page_table ? a.page_table + static_cast<size_t>(a.slots[b]) * a.max_pages_per_seq : nullptr;
  // key tiles: n_pre over the cached prefix (pages), then qt + 1 over this prefill's own rows
  const int n_pre = (prefix + kTK - 1) / kTK;
  const int n_tiles = n_pre + qt + 1;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if ((smem_u32(smem_raw) & 1023u) != 0) __trap();  // SWIZZLE_128B tiles need 1024 B alignment

  // Warp 4 initialises the barriers and puts the first loads (Q, key tile 0) in flight at once; the
  // TMEM allocation (warp 1) and the CTA-wide barrier run under their latency.
  constexpr uint32_t idesc_s = umma_idesc_bf16(kTQ, kTK);                    // both operands K-major
  constexpr uint32_t idesc_o = umma_idesc_bf16(kTQ, kHeadDim) | (1u << 16);  // B (= V) is MN-major
  auto load_kv = [&](int t, int buf) {  // whole warp 4
    if (t < n_pre) {
      const int n_keys = min(kTK, prefix - t * kTK);
      const int pages = (n_keys + kPageTokens - 1) / kPageTokens;
      if (lane == 0) mbar_arrive_expect_tx(&sm.kv_bar[buf], static_cast<uint32_t>(pages) * 2u * kPageBytes);
      __syncwarp();
      if (lane < pages) {
        const int row = (pt[t * kPagesPerTile + lane] * a.n_head + h) * kPageTokens;
        tma_load_2d(k_s(buf) + lane * kPageBytes, &map_k, &sm.kv_bar[buf], 0, row, kEvictNormal);
        tma_load_2d(v_s(buf) + lane * kPageBytes, &map_v, &sm.kv_bar[buf], 0, row, kEvictNormal);
      }
    } else if (lane == 0) {
      const int r = row0 + (t - n_pre) * kTK;
      mbar_arrive_expect_tx(&sm.kv_bar[buf], 2u * kTileBytes);
      tma_load_2d(k_s(buf), &map_qkv, &sm.kv_bar[buf], d + h * kHeadDim, r, kEvictNormal);
      tma_load_2d(v_s(buf), &map_qkv, &sm.kv_bar[buf], 2 * d + h * kHeadDim, r, kEvictNormal);
    }
  };
  if (warp == 4) {
    if (lane == 0) {
      mbar_init(&sm.q_bar, 1);
      mbar_init(&sm.kv_bar[0], 1);
      mbar_init(&sm.kv_bar[1], 1);
      mbar_init(&sm.s_bar, 1);
      mbar_init(&sm.p_bar, kRowThreads);
      mbar_init(&sm.o_bar, 1);
      fence_barrier_init();
      mbar_arrive_expect_tx(&sm.q_bar, kTileBytes);
      tma_load_2d(q_s, &map_qkv, &sm.q_bar, h * kHeadDim, row0 + qt * kTQ, kEvictFirst);
    }
    __syncwarp();
    load_kv(0, 0);
  }
  if (warp == 1) {
    tmem_alloc(&sm.tmem_slot, kCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_slot;

  if (warp == 4) {
    // ================= data movement + MMA issue =================
    auto issue_s = [&](int t) {  // lane 0
      mbar_wait(&sm.kv_bar[t & 1], static_cast<uint32_t>(t >> 1) & 1u);
      tc_fence_after();
      const uint64_t dq = umma_desc_k_sw128(smem_u32(q_s));
      const uint64_t dk = umma_desc_k_sw128(smem_u32(k_s(t & 1)));
#pragma unroll
      for (int k = 0; k < kHeadDim / 16; ++k) {  // 16 dims = 32 B inside the 128 B swizzle row
        umma_bf16_ss(tmem_base, dq + static_cast<uint64_t>(k * 2), dk + static_cast<uint64_t>(k * 2), idesc_s,
                     k > 0 ? 1u : 0u);
      }
      umma_commit(&sm.s_bar);
    };
    if (lane == 0) {
      mbar_wait(&sm.q_bar, 0);
      issue_s(0);
    }
    for (int t = 0; t < n_tiles; ++t) {
      // the other buffer was last read by the MMAs of tile t - 1
      if (t + 1 < n_tiles) {
        if (t >= 1) mbar_wait(&sm.o_bar, static_cast<uint32_t>(t - 1) & 1u);
        load_kv(t + 1, (t + 1) & 1);
      }
      if (lane == 0) {
        mbar_wait(&sm.p_bar, static_cast<uint32_t>(t) & 1u);  // P(t) in shared memory, S(t) consumed
        tc_fence_after();
        const uint64_t dv = umma_desc_mn_sw128(smem_u32(v_s(t & 1)));
#pragma unroll
        for (int s = 0; s < kTK / 16; ++s) {
          // A: 16 keys = 32 B inside the 128 B row of column block s / 4; B: 16 key rows = 2048 B further
          const uint64_t dp =
              umma_desc_k_sw128(smem_u32(p_s) + static_cast<uint32_t>((s >> 2) * kTileBytes + (s & 3) * 32));
          umma_bf16_ss(tmem_base + (kSingle ? 0u : static_cast<uint32_t>(kTK)), dp, dv + static_cast<uint64_t>((s * 16 * 128) >> 4), idesc_o, s > 0 ? 1u : 0u);
        }
        umma_commit(&sm.o_bar);
        if (t + 1 < n_tiles) issue_s(t + 1);  // computed while the row threads fold O_t(t)
      }
      __syncwarp();
    }
  } else {
    // ================= 128 row threads =================
    const uint32_t tmem_s = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
    const uint32_t tmem_o = tmem_s + (kSingle ? 0u : static_cast<uint32_t>(kTK));  // single tile: O_t over the consumed S
    float o_acc[kSingle ? 1 : kHeadDim];  // single tile: O_t goes from TMEM straight to the output
#pragma unroll
    for (int j = 0; j < (kSingle ? 1 : kHeadDim); ++j) o_acc[j] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;  // m_run in units of raw scores
    const int q_local = qt * kTQ + tid;    // this thread's row within the sequence's rows
    const uint32_t p_row = smem_u32(p_s) + static_cast<uint32_t>((tid >> 3) * 1024 + (tid & 7) * 128);

    for (int t = 0; t < n_tiles; ++t) {
      const int buf = t & 1;
      const bool pre = t < n_pre;
      // keys of this tile this row may see: [0, lim)
      //   prefix tile: the cached keys it holds (all of them precede every query)
      //   own-rows tile: causal, and nothing beyond the sequence
      const int tile_lo = pre ? t * kTK : (t - n_pre) * kTK;
      const int lim = pre ? min(kTK, prefix - tile_lo) : min(min(q_local, len_q - 1) - tile_lo + 1, kTK);
      const bool need_mask = pre ? (tile_lo + kTK > prefix) : (t == n_tiles - 1 || tile_lo + kTK > len_q);

      mbar_wait(&sm.s_bar, static_cast<uint32_t>(t) & 1u);
      tc_fence_after();
      if constexpr (kSingle) {
        // P is about to overwrite Q and K: this row's K goes to its page first
        if (pt != nullptr && q_local < len_q) {
          mbar_wait(&sm.kv_bar[buf], 0);
          const int pos = prefix + q_local;
          const size_t dst = ((static_cast<size_t>(pt[pos / kPageTokens]) * a.n_head + h) * kPageTokens +
                              pos % kPageTokens) * kHeadDim;
          const uint8_t* krow = k_s(buf) + (tid >> 3) * 1024 + (tid & 7) * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            *reinterpret_cast<uint4*>(a.k_pool + dst + c * 8) =
                *reinterpret_cast<const uint4*>(krow + ((c ^ (tid & 7)) * 16));
          }
        }
        // (row r of P's second half lands exactly on row r of K, which only thread r reads: no barrier)
      }
      // 32-key chunks of the tile that at least one row of this warp can see (lim grows with the
      // row, so lane 31 has the warp's largest): on a diagonal tile warp w sees w + 1 of the 4 chunks,
      // the others are all-masked and cost neither TMEM reads nor exponentials
      const int warp_lim = need_mask ? __shfl_sync(0xffffffffu, lim, 31) : kTK;
      // ---- softmax of this thread's row over the tile's 128 keys (two passes over TMEM) ----
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c * 32 >= warp_lim) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(tmem_s + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float s = __uint_as_float(r[j]);
          if (need_mask && c * 32 + j >= lim) s = -INFINITY;
          mx = fmaxf(mx, s);
        }
      }
      const float m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;  // nothing visible yet: all weights 0
      const float alpha = (m_run == -INFINITY) ? 0.f : ex2_approx((m_run - m_use) * kScaleLog2);
      m_run = m_new;
      const float neg_m = -m_use * kScaleLog2;
      float row_sum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t pk[16];
        const uint32_t blk0 = p_row + static_cast<uint32_t>((c >> 1) * kTileBytes);
        if (c * 32 >= warp_lim) {  // warp-uniform: nothing visible in this chunk, P = 0
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t chunk = static_cast<uint32_t>((c & 1) * 4 + g) ^ static_cast<uint32_t>(tid & 7);
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(blk0 + (chunk << 4)), "r"(0u) : "memory");
          }
          continue;
        }
        uint32_t r[32];
        tmem_ld_32x32(tmem_s + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float s0 = __uint_as_float(r[j]), s1 = __uint_as_float(r[j + 1]);
          if (need_mask) {
            if (c * 32 + j >= lim) s0 = -INFINITY;
            if (c * 32 + j + 1 >= lim) s1 = -INFINITY;
          }
          const float p0 = ex2_approx(fmaf(s0, kScaleLog2, neg_m));  // ex2(-inf) = 0
          const float p1 = ex2_approx(fmaf(s1, kScaleLog2, neg_m));
          row_sum += p0 + p1;
          pk[j >> 1] = pack_bf16x2(p0, p1);
        }
        // keys c*32 .. c*32+31 of row tid: 64-key column block c >> 1, 16-byte chunks (c & 1) * 4 + 0..3
        const uint32_t blk = p_row + static_cast<uint32_t>((c >> 1) * kTileBytes);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t chunk = static_cast<uint32_t>((c & 1) * 4 + g) ^ static_cast<uint32_t>(tid & 7);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(blk + (chunk << 4)), "r"(pk[4 * g]),
                       "r"(pk[4 * g + 1]), "r"(pk[4 * g + 2]), "r"(pk[4 * g + 3])
                       : "memory");
        }
      }
      l_run = l_run * alpha + row_sum;
      if (pre && tile_lo + kTK > prefix) {
        // last, partial prefix tile: the V rows of pages that were not loaded hold whatever the
        // buffer had before; their weights are exactly 0, but 0 x NaN is NaN, so clear them.
        // (TMA never touches these rows, and the page rows it did load are complete: s_bar fired.)
        const int loaded = ((prefix - tile_lo + kPageTokens - 1) / kPageTokens) * kPageTokens;
        uint4* vz = reinterpret_cast<uint4*>(v_s(buf) + loaded * 128);
        for (int i = tid; i < (kTK - loaded) * 8; i += kRowThreads) vz[i] = make_uint4(0u, 0u, 0u, 0u);
      }
      fence_proxy_async();  // P (and cleared V rows) must be visible to the tensor core's proxy
      tc_fence_before();
      mbar_arrive(&sm.p_bar);

      if (!pre && t - n_pre == qt && pt != nullptr) {
        // diagonal tile: this CTA fills the KV pages with its 128 rows of K and V while P V runs (the
        // tile landed before S was computed). Row tid of the swizzled tile: chunk c lives at chunk
        // position c ^ (tid & 7). (Single-tile instance: K went out before P overwrote it, see above.)
        mbar_wait(&sm.kv_bar[buf], static_cast<uint32_t>(t >> 1) & 1u);  // complete long ago: orders the reads below
        if (q_local < len_q) {
          const int pos = prefix + q_local;
          const size_t dst = ((static_cast<size_t>(pt[pos / kPageTokens]) * a.n_head + h) * kPageTokens +
                              pos % kPageTokens) * kHeadDim;
          const uint8_t* krow = k_s(buf) + (tid >> 3) * 1024 + (tid & 7) * 128;
          const uint8_t* vrow = v_s(buf) + (tid >> 3) * 1024 + (tid & 7) * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int sc = (c ^ (tid & 7)) * 16;
            if (!kSingle) *reinterpret_cast<uint4*>(a.k_pool + dst + c * 8) = *reinterpret_cast<const uint4*>(krow + sc);
            *reinterpret_cast<uint4*>(a.v_pool + dst + c * 8) = *reinterpret_cast<const uint4*>(vrow + sc);
          }
        }
      }

      mbar_wait(&sm.o_bar, static_cast<uint32_t>(t) & 1u);
      tc_fence_after();
      if constexpr (kSingle) {
        // the only tile: normalise O_t and store it, 16 columns at a time (few registers)
        const float inv = 1.0f / l_run;
        __nv_bfloat16* orow = a.out + static_cast<size_t>(row0 + q_local) * d + h * kHeadDim;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t r[16];
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
              "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
              : "r"(tmem_o + c * 16)
              : "memory");
          tmem_ld_wait();
          if (q_local < len_q) {
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              uint4 v;
              v.x = pack_bf16x2(__uint_as_float(r[8 * g]) * inv, __uint_as_float(r[8 * g + 1]) * inv);
              v.y = pack_bf16x2(__uint_as_float(r[8 * g + 2]) * inv, __uint_as_float(r[8 * g + 3]) * inv);
              v.z = pack_bf16x2(__uint_as_float(r[8 * g + 4]) * inv, __uint_as_float(r[8 * g + 5]) * inv);
              v.w = pack_bf16x2(__uint_as_float(r[8 * g + 6]) * inv, __uint_as_float(r[8 * g + 7]) * inv);
              *reinterpret_cast<uint4*>(orow + 16 * c + 8 * g) = v;
            }
          }
        }
        (void)alpha;
      } else {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_o + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) o_acc[c * 32 + j] = fmaf(o_acc[c * 32 + j], alpha, __uint_as_float(r[j]));
        }
      }
      tc_fence_before();
    }

    if constexpr (!kSingle) {
      if (q_local < len_q) {
        const float inv = 1.0f / l_run;
        __nv_bfloat16* orow = a.out + static_cast<size_t>(row0 + q_local) * d + h * kHeadDim;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint4 v;
          v.x = pack_bf16x2(o_acc[8 * g] * inv, o_acc[8 * g + 1] * inv);
          v.y = pack_bf16x2(o_acc[8 * g + 2] * inv, o_acc[8 * g + 3] * inv);
          v.z = pack_bf16x2(o_acc[8 * g + 4] * inv, o_acc[8 * g + 5] * inv);
          v.w = pack_bf16x2(o_acc[8 * g + 6] * inv, o_acc[8 * g + 7] * inv);
          *reinterpret_cast<uint4*>(orow + 8 * g) = v;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kCols);
}

}  // namespace

int prefill_attention_tc(const PagedPrefillArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.max_suffix_len > 0 && a.n_head > 0, "prefill_attention_tc: bad shape");
  B200_CHECK_ARG(a.qkv_rows > 0, "prefill_attention_tc: qkv_rows is needed for the tensor map");
  const bool paged = a.page_table != nullptr;
  B200_CHECK_ARG(paged || a.prefix_len == nullptr, "prefill_attention_tc: cached prefixes need the page table");
  B200_CHECK_ARG(!paged || (a.slots && a.k_pool && a.v_pool && a.n_pages > 0),
                 "prefill_attention_tc: page table without pools / slots / n_pages");
  const int d = a.n_head * kHeadDim;
  CUtensorMap map_qkv, map_k, map_v;
  B200_PROPAGATE(gemm_tensor_map(a.qkv, a.qkv_rows, 3 * d, kTQ, &map_qkv));
  if (paged) {
    const long long pool_rows = static_cast<long long>(a.n_pages) * a.n_head * kPageTokens;
    B200_CHECK_ARG(pool_rows < (1ll << 31), "prefill_attention_tc: KV pool too large for one tensor map");
    B200_PROPAGATE(gemm_tensor_map(a.k_pool, static_cast<int>(pool_rows), kHeadDim, kPageTokens, &map_k));
    B200_PROPAGATE(gemm_tensor_map(a.v_pool, static_cast<int>(pool_rows), kHeadDim, kPageTokens, &map_v));
  } else {
    map_k = map_qkv;  // never dereferenced without a prefix
    map_v = map_qkv;
  }
  TcArgs k = {};
  k.cu_seqlens = a.cu_seqlens;
  k.prefix_len = a.prefix_len;
  k.slots = a.slots;
  k.page_table = a.page_table;
  k.max_pages_per_seq = a.max_pages_per_seq;
  k.n_head = a.n_head;
  k.k_pool = const_cast<__nv_bfloat16*>(a.k_pool);
  k.v_pool = const_cast<__nv_bfloat16*>(a.v_pool);
  k.out = a.out;
  dim3 grid((a.max_suffix_len + kTQ - 1) / kTQ, a.n_head, a.batch);
  // (the opt-in shared-memory size is per device, and cheap: set on every launch)
  if (a.prefix_len == nullptr && a.max_suffix_len <= kTK) {
    // every sequence is one key tile: the light instance (four CTAs per SM)
    B200_CUDA_OK(cudaFuncSetAttribute(prefill_attn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(sizeof(TcSmemSingle))));
    prefill_attn_tc_kernel<true><<<grid, kTcThreads, sizeof(TcSmemSingle), stream>>>(map_qkv, map_k, map_v, k);
  } else {
    B200_CUDA_OK(cudaFuncSetAttribute(prefill_attn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(sizeof(TcSmem))));
    prefill_attn_tc_kernel<false><<<grid, kTcThreads, sizeof(TcSmem), stream>>>(map_qkv, map_k, map_v, k);
  }
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

}  // namespace b200llm
