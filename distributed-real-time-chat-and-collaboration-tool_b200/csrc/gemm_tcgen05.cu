// bf16 GEMM for the dense contractions of GPT-2 (QKV, attention out-proj, MLP up/down, LM
// head) on B200: TMA-fed shared-memory ring, tcgen05.mma (UMMA 128xN, fp32 accumulators in
// TMEM), warp-specialised persistent CTAs, fused epilogues.
//
// Oracle being reproduced: transformers Conv1D = addmm(bias, x, W[in,out])
// (pytorch_utils.py:97-123) used at modeling_gpt2.py:106-107 (c_attn / c_proj),
// :233-241 (MLP c_fc + gelu_new + c_proj), residual adds at :282 and :307, and the tied
// LM head at :646-651 / :705-706.
//
// Logical problem:  out[token, feature] = sum_k act[token, k] * w[feature, k]   (+ epilogue)
// Both operands are K-major (row-major [rows, k]); weights are transposed once at load.
//
// Two orientations share one kernel body:
//   TRANS = false : UMMA M (128 TMEM lanes) = tokens,   UMMA N = features   (prefill)
//   TRANS = true  : UMMA M (128 TMEM lanes) = features, UMMA N = tokens     (decode: "swap
//                   AB" so a batch of 1..256 tokens does not waste the 128-row MMA and the
//                   weight matrix is streamed exactly once)
//
// Warp roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
// allocator, warps 4-7 = epilogue (TMEM -> registers -> global).
// Pipelines: smem ring full/empty (TMA <-> MMA), TMEM accumulator full/empty (MMA <->
// epilogue, two accumulator stages so the epilogue of tile i overlaps the MMAs of tile i+1).
#include "gemm_tcgen05.cuh"

#include <atomic>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"

#ifdef B200_GEMM_TRACE
// Profiling build only (profiles/gemm_trace.cu): CTA 0 records SM-clock timestamps of its
// pipeline milestones so the per-kernel latency chain can be read without a profiler.
__device__ unsigned long long g_gemm_trace[16];
#define B200_TRACE(slot)                                                   \
  do {                                                                     \
    if (blockIdx.x == 0 && blockIdx.y == 0) g_gemm_trace[slot] = clock64(); \
  } while (0)
#elif defined(B200_STEP_TRACE)
// in-pipeline timeline build: the same milestones go into this launch's step-trace record
#define B200_TRACE(slot)                                              \
  do {                                                                \
    if (blockIdx.x == 0 && blockIdx.y == 0) B200_STEP_MARK(slot);     \
  } while (0)
#else
#define B200_TRACE(slot) \
  do {                   \
  } while (0)
#endif

namespace b200llm {

namespace {

constexpr int kBlockM = 128;  // UMMA M
constexpr int kBlockK = 64;   // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kUmmaK = 16;
constexpr int kThreads = 256;
constexpr int kAccumStages = 2;

template <int BLOCK_N>
struct TileCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kMaxStages = (BLOCK_N >= 256) ? 4 : (BLOCK_N >= 128) ? 6 : 8;
  static constexpr int kTmemCols = (kAccumStages * BLOCK_N < 32) ? 32 : kAccumStages * BLOCK_N;
  // ring + barriers + tmem base slot, plus 1024 B slack for manual alignment. The ring depth
  // is a launch-time choice: short split-K CTAs take a small ring so that two CTAs (this
  // kernel's and the next kernel's, started early by programmatic dependent launch) fit on
  // one SM; long weight streams take the full ring.
  static constexpr int smem_bytes(int stages) { return stages * kStageBytes + 1024 + 256; }
};

struct KernelArgs {
  const float* bias;
  void* out;
  int a_rows;  // real rows of the 128-row operand
  int b_rows;  // real rows of the N operand
  int k_blocks_total;
  int k_blocks_per_split;
  int ld_out;
  int tokens;  // logical token count (partials stride)
  float* amax_val;
  int32_t* amax_idx;
  float* eos_logit;
  int ld_amax;
  int eos_id;
  int stages;  // smem ring depth actually allocated (<= TileCfg::kMaxStages)
  // Swapped orientation with every k-block of the CTA resident in the ring (decode GEMMs): all
  // operand tiles complete on full_bar[0], the MMA thread waits once and issues one uninterrupted
  // chain (the per-k-block wait + fence + commit costs more than the MMAs of a 16..64-token tile).
  int single_pass;
  // BAL instantiation (swapped orientation, one token tile, more weight tiles than CTAs -- the LM
  // head: 393 tiles on 148 SMs): instead of striding whole 128-row tiles (97 CTAs get three, 51 get
  // two, and the launch takes as long as three everywhere), CTA c owns rows
  // [c * bal_rows, (c + 1) * bal_rows) and cuts them into 128-row tiles plus a tail of
  // bal_tail_rows, loaded with a smaller TMA box (the third tensor map). bal_rows is a multiple of
  // 32 so a warp's 32 rows stay an aligned group.
  int bal_rows;
  int bal_tail_rows;  // bal_rows % 128 (0 = no tail tile)
};

template <int N>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&r)[32]);

template <>
__device__ __forceinline__ void tmem_ld_cols<32>(uint32_t taddr, uint32_t (&r)[32]) {
  tmem_ld_32x32(taddr, r);
}

template <>
__device__ __forceinline__ void tmem_ld_cols<16>(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// TMA_EPI (tokens-on-M orientation only): the epilogue stages 128-row x 128-byte boxes of the
// output tile in shared memory (SWIZZLE_128B) and hands them to the TMA store unit --
// cp.async.bulk.tensor for bf16 outputs, cp.reduce.async.bulk.tensor ... .add for the fp32
// residual stream (x += tile, performed in L2). Direct per-thread stores from a 128-lane TMEM
// tile hit 32 different cache lines per instruction and made the prefill GEMMs epilogue-bound.
constexpr int kEpiBoxBytes = kBlockM * 128;  // one staging box: 128 rows x 128 B

template <int BLOCK_N, int EPI, bool TRANS, bool TMA_EPI, bool BAL = false>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
            const __grid_constant__ CUtensorMap map_out, const KernelArgs args) {
  using Cfg = TileCfg<BLOCK_N>;
  const int kStages = args.stages;
  constexpr int kChunk = (BLOCK_N >= 32) ? 32 : 16;  // columns per tcgen05.ld

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024 B alignment.
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~static_cast<uintptr_t>(1023u));
  uint8_t* ring = smem;
  uint8_t* epi_stage = smem + kStages * Cfg::kStageBytes;  // 2 boxes, 1024 B aligned (TMA_EPI only)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + (TMA_EPI ? 2 * kEpiBoxBytes : 0));
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint64_t* tmem_empty_bar = tmem_full_bar + kAccumStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + kAccumStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  B200_STEP_BEGIN(100 + EPI);
  if (threadIdx.x == 0) B200_TRACE(0);

  const int num_m_blocks = (args.a_rows + kBlockM - 1) / kBlockM;
  const int num_n_blocks = (args.b_rows + BLOCK_N - 1) / BLOCK_N;
  // BAL: this CTA's row range and tile count; `num_tiles` is then chosen so that the strided loops
  // below (tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) run exactly bal_my_tiles times
  const int bal_begin = BAL ? static_cast<int>(blockIdx.x) * args.bal_rows : 0;
  const int bal_end = BAL ? min(bal_begin + args.bal_rows, args.a_rows) : 0;
  const int bal_my_tiles = (BAL && bal_end > bal_begin) ? (bal_end - bal_begin + kBlockM - 1) / kBlockM : 0;
  const int bal_tiles_per_cta = BAL ? (args.bal_rows + kBlockM - 1) / kBlockM : 0;
  const int num_tiles =
      BAL ? (bal_my_tiles > 0 ? static_cast<int>(blockIdx.x) + (bal_my_tiles - 1) * static_cast<int>(gridDim.x) + 1 : 0)
          : num_m_blocks * num_n_blocks;
  const int split = blockIdx.y;
  const int kb_begin = split * args.k_blocks_per_split;
  int kb_end = kb_begin + args.k_blocks_per_split;
  if (kb_end > args.k_blocks_total) kb_end = args.k_blocks_total;
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < kAccumStages; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Let the next kernel of the step start its own prologue / weight prefetch now; it still
  // waits (griddepcontrol.wait) for this grid to finish before touching activations.
  pdl_launch_dependents();
  if (threadIdx.x == 0) B200_TRACE(1);

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      const int my_tiles =
          (static_cast<int>(blockIdx.x) < num_tiles)
              ? (num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                    static_cast<int>(gridDim.x)
              : 0;
      const int total_iters = my_tiles * num_kb;
      // In the swapped orientation operand A is the weight matrix, which does not depend on
      // the upstream kernel: fill the whole ring with weight tiles BEFORE the grid
      // dependency resolves (programmatic dependent launch), so the weight stream of this
      // GEMM overlaps the tail of the previous kernel. The activation operand must wait.
      const int pre = TRANS ? (total_iters < kStages ? total_iters : kStages) : 0;
      const bool single = TRANS && args.single_pass != 0;  // host guarantees total_iters <= kStages
      if (single) mbar_arrive_expect_tx(&full_bar[0], static_cast<uint32_t>(total_iters) * Cfg::kStageBytes);
      for (int it = 0; it < pre; ++it) {
        const int tile = blockIdx.x + (it / num_kb) * gridDim.x;
        const int kb = kb_begin + it % num_kb;
        const int m_blk = tile % num_m_blocks;
        uint8_t* sa = ring + it * Cfg::kStageBytes;
        if constexpr (BAL) {
          const int row0 = bal_begin + (it / num_kb) * kBlockM;
          const bool tail = args.bal_tail_rows > 0 && bal_end - row0 <= args.bal_tail_rows;
          mbar_arrive_expect_tx(&full_bar[it], (tail ? static_cast<uint32_t>(args.bal_tail_rows) * 128u
                                                     : static_cast<uint32_t>(Cfg::kABytes)) +
                                                   static_cast<uint32_t>(Cfg::kStageBytes - Cfg::kABytes));
          tma_load_2d(sa, tail ? &map_out : &map_a, &full_bar[it], kb * kBlockK, row0, kEvictFirst);
          continue;
        }
        if (!single) mbar_arrive_expect_tx(&full_bar[it], Cfg::kStageBytes);
        tma_load_2d(sa, &map_a, &full_bar[single ? 0 : it], kb * kBlockK, m_blk * kBlockM, kEvictFirst);
      }
      B200_TRACE(2);
      pdl_wait();
      B200_TRACE(3);
      for (int it = 0; it < pre; ++it) {
        const int tile = blockIdx.x + (it / num_kb) * gridDim.x;
        const int kb = kb_begin + it % num_kb;
        const int n_blk = BAL ? 0 : tile / num_m_blocks;
        uint8_t* sb = ring + it * Cfg::kStageBytes + Cfg::kABytes;
        tma_load_2d(sb, &map_b, &full_bar[single ? 0 : it], kb * kBlockK, n_blk * BLOCK_N, kEvictLast);
      }
      for (int it = pre; it < total_iters; ++it) {
        const int stage = it % kStages;
        const uint32_t phase = static_cast<uint32_t>(it / kStages) & 1u;
        const int tile = blockIdx.x + (it / num_kb) * gridDim.x;
        const int kb = kb_begin + it % num_kb;
        const int m_blk = tile % num_m_blocks;
        const int n_blk = BAL ? 0 : tile / num_m_blocks;
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        uint8_t* sa = ring + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        if constexpr (BAL) {
          const int row0 = bal_begin + (it / num_kb) * kBlockM;
          const bool tail = args.bal_tail_rows > 0 && bal_end - row0 <= args.bal_tail_rows;
          mbar_arrive_expect_tx(&full_bar[stage], (tail ? static_cast<uint32_t>(args.bal_tail_rows) * 128u
                                                        : static_cast<uint32_t>(Cfg::kABytes)) +
                                                      static_cast<uint32_t>(Cfg::kStageBytes - Cfg::kABytes));
          tma_load_2d(sa, tail ? &map_out : &map_a, &full_bar[stage], kb * kBlockK, row0, kEvictFirst);
          tma_load_2d(sb, &map_b, &full_bar[stage], kb * kBlockK, 0, kEvictLast);
          continue;
        }
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        tma_load_2d(sa, &map_a, &full_bar[stage], kb * kBlockK, m_blk * kBlockM,
                    TRANS ? kEvictFirst : kEvictNormal);
        tma_load_2d(sb, &map_b, &full_bar[stage], kb * kBlockK, n_blk * BLOCK_N, kEvictLast);
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    // The whole warp runs the loop (waits, descriptor arithmetic: all warp-uniform) and one elected
    // lane issues. Issuing from inside an `if (lane == 0)` region instead makes the compiler wrap
    // every tcgen05 instruction in its own elect/branch sequence: ~25 % more cycles per k-block
    // (profiles/umma_rate.cu), and the issue loop, not the tensor pipe, is what a 16..64-token
    // tile waits for.
    {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      int astage = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty_bar[astage], aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(astage * BLOCK_N);
        if (TRANS && args.single_pass != 0) {
          // (a CTA has at most one tile per ring slot group here; `it` counts its k-blocks so far)
          const int it0 = ((tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x)) * num_kb;
          if (it0 == 0) {
            mbar_wait(&full_bar[0], 0);
            if (lane == 0) {
              B200_TRACE(4);
              B200_TRACE(11);
            }
            tc_fence_after();
          }
          if (elect_one()) {
            uint64_t desc_a = umma_desc_k_sw128(smem_u32(ring + it0 * Cfg::kStageBytes));
            uint64_t desc_b = umma_desc_k_sw128(smem_u32(ring + it0 * Cfg::kStageBytes + Cfg::kABytes));
            for (int kb = 0; kb < num_kb; ++kb) {
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                umma_bf16_ss(tmem_d, desc_a + static_cast<uint64_t>(k * 2), desc_b + static_cast<uint64_t>(k * 2),
                             idesc, (kb > 0 || k > 0) ? 1u : 0u);
              }
              desc_a += static_cast<uint64_t>(Cfg::kStageBytes >> 4);
              desc_b += static_cast<uint64_t>(Cfg::kStageBytes >> 4);
            }
            umma_commit(&tmem_full_bar[astage]);
          }
          __syncwarp();
          if (lane == 0) B200_TRACE(5);
          if (++astage == kAccumStages) {
            astage = 0;
            aphase ^= 1u;
          }
          continue;
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          if (lane == 0) {
            if (kb == 0) B200_TRACE(4);
            if (kb == num_kb - 1) B200_TRACE(11);
          }
          tc_fence_after();
          const uint32_t sa = smem_u32(ring + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
          const uint64_t desc_a = umma_desc_k_sw128(sa);
          const uint64_t desc_b = umma_desc_k_sw128(sb);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
              // advance 16 bf16 = 32 B inside the 128 B swizzle row: +2 in 16-byte units
              umma_bf16_ss(tmem_d, desc_a + static_cast<uint64_t>(k * 2),
                           desc_b + static_cast<uint64_t>(k * 2), idesc,
                           (kb > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
            // accumulator complete (same lane as the MMAs: a commit covers its own thread's work)
            if (kb == num_kb - 1) umma_commit(&tmem_full_bar[astage]);
          }
          __syncwarp();
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        if (lane == 0) B200_TRACE(5);
        if (++astage == kAccumStages) {
          astage = 0;
          aphase ^= 1u;
        }
      }
    }
  } else if (warp >= 4) {
    // ================= Epilogue (128 threads, one TMEM lane each) =================
    const int q = warp - 4;  // == warp % 4: TMEM lane quarter this warp may access
    int astage = 0;
    uint32_t aphase = 0;
    int epi_box = 0;  // staging-buffer rotation (TMA_EPI)
    pdl_wait();  // residual / output buffers may be read by the upstream kernel
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile % num_m_blocks;
      const int n_blk = BAL ? 0 : tile / num_m_blocks;
      mbar_wait(&tmem_full_bar[astage], aphase);
      if (threadIdx.x == 128) B200_TRACE(6);
      tc_fence_after();
      const int bal_i = BAL ? (tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) : 0;
      const int a_row = (BAL ? bal_begin + bal_i * kBlockM : m_blk * kBlockM) + q * 32 + lane;  // row of the 128-row operand
      const bool a_ok = BAL ? (a_row < bal_end) : (a_row < args.a_rows);
      const int group0 = BAL ? (static_cast<int>(blockIdx.x) * bal_tiles_per_cta + bal_i) * 4 : m_blk * 4;

      float bias_lane = 0.0f;
      if (TRANS && EPI != kEpiF32 && EPI != kEpiPartialF32 && EPI != kEpiLogitsArgmax && a_ok &&
          args.bias != nullptr) {
        bias_lane = args.bias[a_row];
      }

      if constexpr (TMA_EPI) {
        // ---- staged epilogue: TMEM -> registers -> swizzled smem box -> TMA store / reduce-add
        constexpr bool kF32Out = (EPI == kEpiBiasResidF32);
        constexpr int kBoxCols = kF32Out ? 32 : 64;  // 128 bytes per row either way
        const int row_in_tile = q * 32 + lane;
        const uint32_t sw = static_cast<uint32_t>(row_in_tile & 7);
#pragma unroll 1
        for (int c0 = 0; c0 < BLOCK_N; c0 += kBoxCols) {
          const int f_box = n_blk * BLOCK_N + c0;
          if (f_box >= args.b_rows) break;  // uniform across the epilogue warps
          uint8_t* box = epi_stage + (epi_box & 1) * kEpiBoxBytes;
          // the TMA store issued two boxes ago must have finished reading this buffer
          if (threadIdx.x == 128) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          asm volatile("bar.sync 1, 128;" ::: "memory");
          uint8_t* row_ptr = box + row_in_tile * 128;
#pragma unroll
          for (int cc = 0; cc < kBoxCols; cc += 32) {
            uint32_t r[32];
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                   static_cast<uint32_t>(astage * BLOCK_N + c0 + cc);
            tmem_ld_cols<32>(taddr, r);
            tmem_ld_wait();
            if (threadIdx.x == 128 && c0 == 0 && cc == 0) B200_TRACE(9);
            const int f0 = f_box + cc;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              float v[8];
              float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
              if (args.bias != nullptr && f0 + j + 7 < args.b_rows) {
                b0 = __ldg(reinterpret_cast<const float4*>(args.bias + f0 + j));
                b1 = __ldg(reinterpret_cast<const float4*>(args.bias + f0 + j + 4));
              }
              v[0] = __uint_as_float(r[j + 0]) + b0.x;
              v[1] = __uint_as_float(r[j + 1]) + b0.y;
              v[2] = __uint_as_float(r[j + 2]) + b0.z;
              v[3] = __uint_as_float(r[j + 3]) + b0.w;
              v[4] = __uint_as_float(r[j + 4]) + b1.x;
              v[5] = __uint_as_float(r[j + 5]) + b1.y;
              v[6] = __uint_as_float(r[j + 6]) + b1.z;
              v[7] = __uint_as_float(r[j + 7]) + b1.w;
              if (EPI == kEpiBiasGeluBf16) {
#pragma unroll
                for (int t = 0; t < 8; ++t) v[t] = gelu_new_fast(v[t]);
              }
              if (kF32Out) {
                // 8 floats = two 16-byte chunks of the 128-byte row (chunk index = col / 4)
                const uint32_t ch = static_cast<uint32_t>((cc + j) >> 2);
                *reinterpret_cast<float4*>(row_ptr + (((ch + 0) ^ sw) << 4)) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(row_ptr + (((ch + 1) ^ sw) << 4)) = make_float4(v[4], v[5], v[6], v[7]);
              } else {
                // 8 bf16 = one 16-byte chunk (chunk index = col / 8)
                const uint32_t ch = static_cast<uint32_t>((cc + j) >> 3);
                uint4 pk;
                pk.x = pack_bf16x2(v[0], v[1]);
                pk.y = pack_bf16x2(v[2], v[3]);
                pk.z = pack_bf16x2(v[4], v[5]);
                pk.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(row_ptr + ((ch ^ sw) << 4)) = pk;
              }
            }
          }
          fence_proxy_async();  // make the generic-proxy smem writes visible to the TMA unit
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (threadIdx.x == 128) {
            const uint64_t desc = reinterpret_cast<uint64_t>(&map_out);
            const uint32_t src = smem_u32(box);
            const int crd0 = f_box, crd1 = m_blk * kBlockM;
            if (kF32Out) {
              asm volatile(
                  "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];"
                  :
                  : "l"(desc), "r"(src), "r"(crd0), "r"(crd1)
                  : "memory");
            } else {
              asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                           :
                           : "l"(desc), "r"(src), "r"(crd0), "r"(crd1)
                           : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          ++epi_box;
        }
        if (threadIdx.x == 128) B200_TRACE(7);
      } else {
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += kChunk) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(astage * BLOCK_N + c0);
        tmem_ld_cols<kChunk>(taddr, r);
        tmem_ld_wait();
        if (threadIdx.x == 128 && c0 == 0) B200_TRACE(9);
        if (threadIdx.x == 128 && c0 == kChunk) B200_TRACE(10);
        const int b_row0 = n_blk * BLOCK_N + c0;  // first row of the N operand in this chunk

        if (TRANS && EPI == kEpiLogitsArgmax) {
          // LM head with the greedy selection fused in: lane = vocabulary entry, column =
          // token. Each warp reduces its 32 entries to one (max, argmax) candidate per token
          // with shuffles (ties -> lowest index); lane j keeps the candidate of column j so
          // the 32 results leave as one coalesced store. EOS is excluded from the candidates
          // and reported separately so the selector can apply min_new_tokens exactly.
          const int feature = a_row;
          const int n_valid = args.b_rows - b_row0;  // token columns of this chunk that exist
          float v[kChunk];
#pragma unroll
          for (int j = 0; j < kChunk; ++j) {
            v[j] = a_ok ? __uint_as_float(r[j]) : -INFINITY;
          }
          if (a_ok && args.out != nullptr) {
            float* o = reinterpret_cast<float*>(args.out) + static_cast<size_t>(b_row0) * args.ld_out + feature;
#pragma unroll
            for (int j = 0; j < kChunk; ++j) {
              if (j < n_valid) *o = v[j];
              o += args.ld_out;
            }
          }
          if (a_ok && feature == args.eos_id) {
#pragma unroll
            for (int j = 0; j < kChunk; ++j) {
              if (j < n_valid) args.eos_logit[b_row0 + j] = v[j];
              v[j] = -INFINITY;
            }
          }
          // Butterfly transpose-reduce over the warp's 32 vocabulary entries: at each level a
          // lane keeps half of its columns and trades the other half with lane^offset, so 31
          // (value, index) exchanges finish all kChunk columns and lane j ends with column j.
          int idx[kChunk / 2];
          {
            constexpr int n = kChunk / 2;
            const bool upper = (lane & n) != 0;  // offset == n at the first level (kChunk 32 or 16)
#pragma unroll
            for (int j = 0; j < n; ++j) {
              const float send = upper ? v[j] : v[j + n];
              const float keep = upper ? v[j + n] : v[j];
              const float recv = __shfl_xor_sync(0xffffffffu, send, n);
              const int other = feature ^ n;  // partner lane's feature (32-aligned groups)
              const bool take = recv > keep || (recv == keep && other < feature);
              v[j] = take ? recv : keep;
              idx[j] = take ? other : feature;
            }
          }
#pragma unroll
          for (int n = kChunk / 4; n >= 1; n >>= 1) {
            const bool upper = (lane & n) != 0;
#pragma unroll
            for (int j = 0; j < n; ++j) {
              const float send_v = upper ? v[j] : v[j + n];
              const int send_i = upper ? idx[j] : idx[j + n];
              const float keep_v = upper ? v[j + n] : v[j];
              const int keep_i = upper ? idx[j + n] : idx[j];
              const float recv_v = __shfl_xor_sync(0xffffffffu, send_v, n);
              const int recv_i = __shfl_xor_sync(0xffffffffu, send_i, n);
              const bool take = recv_v > keep_v || (recv_v == keep_v && recv_i < keep_i);
              v[j] = take ? recv_v : keep_v;
              idx[j] = take ? recv_i : keep_i;
            }
          }
          if (kChunk == 16) {
            // 16 columns leave one column per lane PAIR (lane, lane^16), each reduced over its
            // own 16 lanes: one plain exchange merges the two halves.
            const float recv_v = __shfl_xor_sync(0xffffffffu, v[0], 16);
            const int recv_i = __shfl_xor_sync(0xffffffffu, idx[0], 16);
            const bool take = recv_v > v[0] || (recv_v == v[0] && recv_i < idx[0]);
            v[0] = take ? recv_v : v[0];
            idx[0] = take ? recv_i : idx[0];
          }
          const int col = (kChunk == 32) ? lane : (lane & 15);
          const int token = b_row0 + col;
          if ((kChunk == 32 || lane < 16) && token < args.b_rows) {
            const size_t pidx = static_cast<size_t>(group0 + q) * args.ld_amax + token;
            args.amax_val[pidx] = v[0];
            args.amax_idx[pidx] = idx[0];
          }
        } else if (TRANS) {
          // lane = feature (a_row), columns = tokens. For each token the warp's 32 lanes
          // write 32 consecutive features -> one coalesced 64/128-byte store. Addresses advance
          // by a running pointer and the tail is predicated (no per-store branches / 64-bit
          // multiplies: that version took ~60 cycles per store).
          const int feature = a_row;
          const int n_valid = a_ok ? (args.b_rows - b_row0) : 0;  // valid token columns here
          const size_t ld = static_cast<size_t>(args.ld_out);
          if (EPI == kEpiBiasBf16 || EPI == kEpiBiasGeluBf16) {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out) +
                               static_cast<size_t>(b_row0) * ld + feature;
#pragma unroll
            for (int j = 0; j < kChunk; ++j) {
              float x = __uint_as_float(r[j]) + bias_lane;
              if (EPI == kEpiBiasGeluBf16) x = gelu_new(x);
              const __nv_bfloat16 y = __float2bfloat16(x);
              if (j < n_valid) *o = y;
              o += ld;
            }
          } else {
            float* o = reinterpret_cast<float*>(args.out) + feature;
            if (EPI == kEpiPartialF32) {
              o += (static_cast<size_t>(split) * args.tokens + b_row0) * ld;
            } else {
              o += static_cast<size_t>(b_row0) * ld;
            }
            if (EPI == kEpiBiasResidF32) {
              float old[kChunk];
              {
                const float* oi = o;
#pragma unroll
                for (int j = 0; j < kChunk; ++j) {
                  old[j] = (j < n_valid) ? *oi : 0.f;
                  oi += ld;
                }
              }
#pragma unroll
              for (int j = 0; j < kChunk; ++j) {
                const float x = old[j] + (__uint_as_float(r[j]) + bias_lane);
                if (j < n_valid) *o = x;
                o += ld;
              }
            } else {
#pragma unroll
              for (int j = 0; j < kChunk; ++j) {
                if (j < n_valid) *o = __uint_as_float(r[j]);
                o += ld;
              }
            }
          }
        } else {
          // lane = token (a_row), columns = features: each thread owns a contiguous run of
          // kChunk features of its token row.
          const int token = a_row;
          const int f0 = b_row0;
          if (a_ok && f0 < args.b_rows) {
            const bool full = (f0 + kChunk <= args.b_rows);
            if (EPI == kEpiBiasBf16 || EPI == kEpiBiasGeluBf16) {
              __nv_bfloat16* o =
                  reinterpret_cast<__nv_bfloat16*>(args.out) + static_cast<size_t>(token) * args.ld_out + f0;
              if (full && ((reinterpret_cast<uintptr_t>(o) & 15u) == 0)) {
#pragma unroll
                for (int j = 0; j < kChunk; j += 8) {
                  float v[8];
#pragma unroll
                  for (int t = 0; t < 8; ++t) {
                    float b = args.bias ? __ldg(args.bias + f0 + j + t) : 0.0f;
                    float x = __uint_as_float(r[j + t]) + b;
                    v[t] = (EPI == kEpiBiasGeluBf16) ? gelu_new(x) : x;
                  }
                  uint4 pk;
                  pk.x = pack_bf16x2(v[0], v[1]);
                  pk.y = pack_bf16x2(v[2], v[3]);
                  pk.z = pack_bf16x2(v[4], v[5]);
                  pk.w = pack_bf16x2(v[6], v[7]);
                  *reinterpret_cast<uint4*>(o + j) = pk;
                }
              } else {
#pragma unroll
                for (int j = 0; j < kChunk; ++j) {
                  if (f0 + j < args.b_rows) {
                    float b = args.bias ? __ldg(args.bias + f0 + j) : 0.0f;
                    float x = __uint_as_float(r[j]) + b;
                    o[j] = __float2bfloat16((EPI == kEpiBiasGeluBf16) ? gelu_new(x) : x);
                  }
                }
              }
            } else {
              float* o;
              if (EPI == kEpiPartialF32) {
                o = reinterpret_cast<float*>(args.out) +
                    (static_cast<size_t>(split) * args.tokens + token) * args.ld_out + f0;
              } else {
                o = reinterpret_cast<float*>(args.out) + static_cast<size_t>(token) * args.ld_out + f0;
              }
              if (full && ((reinterpret_cast<uintptr_t>(o) & 15u) == 0)) {
#pragma unroll
                for (int j = 0; j < kChunk; j += 4) {
                  float4 v;
                  v.x = __uint_as_float(r[j + 0]);
                  v.y = __uint_as_float(r[j + 1]);
                  v.z = __uint_as_float(r[j + 2]);
                  v.w = __uint_as_float(r[j + 3]);
                  if (EPI == kEpiBiasResidF32) {
                    const float4 old = *reinterpret_cast<const float4*>(o + j);
                    const float4 b = args.bias
                                         ? __ldg(reinterpret_cast<const float4*>(args.bias + f0 + j))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
                    v.x = old.x + (v.x + b.x);
                    v.y = old.y + (v.y + b.y);
                    v.z = old.z + (v.z + b.z);
                    v.w = old.w + (v.w + b.w);
                  }
                  *reinterpret_cast<float4*>(o + j) = v;
                }
              } else {
#pragma unroll
                for (int j = 0; j < kChunk; ++j) {
                  if (f0 + j < args.b_rows) {
                    float x = __uint_as_float(r[j]);
                    if (EPI == kEpiBiasResidF32) {
                      float b = args.bias ? __ldg(args.bias + f0 + j) : 0.0f;
                      x = o[j] + (x + b);
                    }
                    o[j] = x;
                  }
                }
              }
            }
          }
        }
      }
      }  // !TMA_EPI
      // All TMEM reads of this accumulator stage are complete (tcgen05.wait::ld above).
      tc_fence_before();
      if (!TMA_EPI && threadIdx.x == 128) B200_TRACE(7);
      mbar_arrive(&tmem_empty_bar[astage]);
      if (++astage == kAccumStages) {
        astage = 0;
        aphase ^= 1u;
      }
    }
    // every bulk store / reduction of this CTA must be complete before it exits
    if (TMA_EPI && threadIdx.x == 128) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    (void)epi_box;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
  if (threadIdx.x == 64) B200_TRACE(8);
  B200_STEP_END();
}


// ---------------------------------------------------------------------------------------
// Cluster split-K variant (swap-AB orientation, one 128-feature tile per cluster).
//
// Decode GEMMs have only 6..40 weight tiles, far fewer than 148 SMs, and one SM can ingest only
// ~56 B/clk through TMA, so K is split over the CTAs of a thread-block cluster: CTA `rank`
// multiplies its K slice into TMEM, spills the fp32 partial tile to its own shared memory, and
// after a cluster barrier CTA `rank` reduces token columns [rank*cpr, (rank+1)*cpr) of ALL ranks
// through distributed shared memory (fixed rank order -> deterministic) and applies the final
// epilogue (bias / gelu / residual). No global partial planes, no consumer-side reduction.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(remote)
               : "memory");
  return v;
}

template <int BLOCK_N, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm_cluster_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const KernelArgs args) {
  using Cfg = TileCfg<BLOCK_N>;
  const int kStages = args.stages;
  constexpr int kChunk = (BLOCK_N >= 32) ? 32 : 16;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~static_cast<uintptr_t>(1023u));
  uint8_t* ring = smem;
  // barriers live after max(ring, staging) so the staging tile never overwrites them
  const int ring_bytes = kStages * Cfg::kStageBytes;
  const int stage_tile_bytes = BLOCK_N * kBlockM * 4;
  const int body_bytes = ring_bytes > stage_tile_bytes ? ring_bytes : stage_tile_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + body_bytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);
  float* stage_tile = reinterpret_cast<float*>(ring);  // [BLOCK_N tokens][128 features] fp32

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  B200_STEP_BEGIN(110 + EPI);
  if (threadIdx.x == 0) B200_TRACE(0);
  const uint32_t rank = cluster_ctarank();
  const uint32_t csize = cluster_nctarank();
  const int m_blk = blockIdx.x;
  const int kb_begin = static_cast<int>(rank) * args.k_blocks_per_split;
  int kb_end = kb_begin + args.k_blocks_per_split;
  if (kb_end > args.k_blocks_total) kb_end = args.k_blocks_total;
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) B200_TRACE(1);

  if (warp == 0) {
    if (lane == 0) {
      // weights first (independent of the upstream kernel), activations after the dependency
      const int pre = num_kb < kStages ? num_kb : kStages;
      for (int it = 0; it < pre; ++it) {
        uint8_t* sa = ring + it * Cfg::kStageBytes;
        mbar_arrive_expect_tx(&full_bar[it], Cfg::kStageBytes);
        tma_load_2d(sa, &map_a, &full_bar[it], (kb_begin + it) * kBlockK, m_blk * kBlockM, kEvictFirst);
      }
      pdl_wait();
      for (int it = 0; it < pre; ++it) {
        uint8_t* sb = ring + it * Cfg::kStageBytes + Cfg::kABytes;
        tma_load_2d(sb, &map_b, &full_bar[it], (kb_begin + it) * kBlockK, 0, kEvictLast);
      }
      for (int it = pre; it < num_kb; ++it) {
        const int stage = it % kStages;
        const uint32_t phase = static_cast<uint32_t>(it / kStages) & 1u;
        mbar_wait(&empty_bar[stage], phase ^ 1u);
        uint8_t* sa = ring + stage * Cfg::kStageBytes;
        mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        tma_load_2d(sa, &map_a, &full_bar[stage], (kb_begin + it) * kBlockK, m_blk * kBlockM, kEvictFirst);
        tma_load_2d(sa + Cfg::kABytes, &map_b, &full_bar[stage], (kb_begin + it) * kBlockK, 0, kEvictLast);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, BLOCK_N);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int stage = kb % kStages;
        const uint32_t phase = static_cast<uint32_t>(kb / kStages) & 1u;
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(ring + stage * Cfg::kStageBytes);
        const uint64_t desc_a = umma_desc_k_sw128(sa);
        const uint64_t desc_b = umma_desc_k_sw128(sa + Cfg::kABytes);
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          umma_bf16_ss(tmem_base, desc_a + static_cast<uint64_t>(k * 2), desc_b + static_cast<uint64_t>(k * 2),
                       idesc, (kb > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
      }
      umma_commit(tmem_full_bar);
      B200_TRACE(5);
    }
  } else if (warp >= 4) {
    // phase A: this CTA's partial tile TMEM -> own shared memory, [token][feature] fp32
    const int q = warp - 4;
    mbar_wait(tmem_full_bar, 0);
    if (threadIdx.x == 128) B200_TRACE(6);
    tc_fence_after();
    const int f_local = q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < BLOCK_N; c0 += kChunk) {
      uint32_t r[32];
      tmem_ld_cols<kChunk>(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c0), r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < kChunk; ++j) {
        stage_tile[(c0 + j) * kBlockM + f_local] = (num_kb > 0) ? __uint_as_float(r[j]) : 0.f;
      }
    }
    tc_fence_before();
    if (threadIdx.x == 128) B200_TRACE(9);
  }

  cluster_sync_all();  // every rank's partial tile is visible cluster-wide
  if (threadIdx.x == 128) B200_TRACE(10);

  if (warp >= 4) {
    // phase B: reduce my token columns over all ranks through DSMEM and finish the epilogue
    const int q = warp - 4;
    const int cpr = (BLOCK_N + static_cast<int>(csize) - 1) / static_cast<int>(csize);
    const int t_begin = static_cast<int>(rank) * cpr;
    int t_end = t_begin + cpr;
    if (t_end > BLOCK_N) t_end = BLOCK_N;
    if (t_end > args.b_rows) t_end = args.b_rows;
    const int fb = m_blk * kBlockM + lane * 4;  // first of this lane's 4 features
    const bool f_ok = fb < args.a_rows;         // a_rows % 4 == 0 is checked on the host
    float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f_ok && args.bias != nullptr) bias4 = __ldg(reinterpret_cast<const float4*>(args.bias + fb));
    const uint32_t tile_addr = smem_u32(stage_tile);
    pdl_wait();
    for (int t = t_begin + q; t < t_end; t += 4) {
      const uint32_t laddr = tile_addr + static_cast<uint32_t>((t * kBlockM + lane * 4) * 4);
      // all remote loads of this column in flight together; summed in rank order (deterministic)
      float4 v[8];
#pragma unroll
      for (uint32_t s = 0; s < 8; ++s) {
        v[s] = (s < csize) ? ld_dsmem_f4(laddr, s) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      float4 acc = v[0];
#pragma unroll
      for (uint32_t s = 1; s < 8; ++s) {
        acc.x += v[s].x;
        acc.y += v[s].y;
        acc.z += v[s].z;
        acc.w += v[s].w;
      }
      if (!f_ok) continue;
      acc.x += bias4.x;
      acc.y += bias4.y;
      acc.z += bias4.z;
      acc.w += bias4.w;
      const size_t idx = static_cast<size_t>(t) * args.ld_out + fb;
      if (EPI == kEpiBiasBf16 || EPI == kEpiBiasGeluBf16) {
        if (EPI == kEpiBiasGeluBf16) {
          acc.x = gelu_new(acc.x);
          acc.y = gelu_new(acc.y);
          acc.z = gelu_new(acc.z);
          acc.w = gelu_new(acc.w);
        }
        uint2 pk;
        pk.x = pack_bf16x2(acc.x, acc.y);
        pk.y = pack_bf16x2(acc.z, acc.w);
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(args.out) + idx) = pk;
      } else {  // kEpiBiasResidF32
        float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(args.out) + idx);
        const float4 old = *o;
        *o = make_float4(old.x + acc.x, old.y + acc.y, old.z + acc.z, old.w + acc.w);
      }
    }
  }

  if (threadIdx.x == 128) B200_TRACE(7);
  cluster_sync_all();  // nobody leaves while a peer may still read its shared memory
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
  if (threadIdx.x == 64) B200_TRACE(8);
  B200_STEP_END();
}


// ---------------------------------------------------------------------------------------
// CTA-pair variant for prefill (tokens on UMMA-M): tcgen05.mma.cta_group::2, M = 256.
//
// Two CTAs of a cluster (one TPC) own a 256 x 256 output tile: CTA `rank` loads its own 128
// token rows of A and HALF of the weight tile (128 of the 256 feature rows), the leader issues
// one UMMA that reads both halves, and each CTA's TMEM receives its 128 rows x 256 columns.
// Per SM that is 32 KB of operands per 64-deep k block instead of 48 KB, which matters because
// these GEMMs are bound by per-SM operand ingest through TMA, not by the tensor pipe.
// Barriers: TMA of both CTAs completes on the LEADER's full barrier; the leader's
// tcgen05.commit multicasts to the empty / tmem_full barriers of both CTAs; both epilogues
// arrive (locally / remotely) on the leader's tmem_empty barrier.
// ---------------------------------------------------------------------------------------
constexpr int kPairStageBytes = 2 * kBlockM * kBlockK * 2;  // A 16 KB + half of B 16 KB
constexpr int kPairStages = 6;
constexpr int kPairN = 256;
constexpr int kPairSmemBytes = kPairStages * kPairStageBytes + 2 * kEpiBoxBytes + 1024 + 256;

__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* desc, uint32_t leader_bar,
                                                 int32_t c0, int32_t c1, uint64_t cache_hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      :
      : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(leader_bar), "r"(c0), "r"(c1),
        "l"(cache_hint)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                                  uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n"
      :
      : "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once the leader's previously issued MMAs retire) on the barrier at this smem offset in
// BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  const uint16_t mask = 0x3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta_rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(cta_rank));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_out, const KernelArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~static_cast<uintptr_t>(1023u));
  uint8_t* ring = smem;
  uint8_t* epi_stage = smem + kPairStages * kPairStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + 2 * kEpiBoxBytes);
  uint64_t* empty_bar = full_bar + kPairStages;
  uint64_t* tmem_full_bar = empty_bar + kPairStages;
  uint64_t* tmem_empty_bar = tmem_full_bar + kAccumStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + kAccumStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int n_clusters = gridDim.x >> 1;
  const int num_m2 = (args.a_rows + 2 * kBlockM - 1) / (2 * kBlockM);
  const int num_n = (args.b_rows + kPairN - 1) / kPairN;
  const int num_tiles = num_m2 * num_n;
  const int num_kb = args.k_blocks_total;
  // Work item = (tile, K split). The residual epilogue adds into the output (TMA reduce-add), so a tile's
  // K range can be cut into several items at no extra pass: with 96 tiles on 74 CTA pairs, three items per
  // tile make four short waves out of two long ones. Items are ordered split-major, so the adds into one
  // output tile are far apart in time. Other epilogues run with one split.
  const int kb_per = args.k_blocks_per_split;
  const int n_split = (num_kb + kb_per - 1) / kb_per;
  const int num_items = num_tiles * n_split;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPairStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < kAccumStages; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], 8);  // 4 epilogue warps of each CTA
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers exist before anyone signals across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      pdl_wait();
      int it = 0;
      for (int item = cluster_id; item < num_items; item += n_clusters) {
        const int tile = item % num_tiles, ks = item / num_tiles;
        const int m2 = tile % num_m2, n_blk = tile / num_m2;
        const int kb_end = min(num_kb, (ks + 1) * kb_per);
        for (int kb = ks * kb_per; kb < kb_end; ++kb, ++it) {
          const int stage = it % kPairStages;
          const uint32_t phase = static_cast<uint32_t>(it / kPairStages) & 1u;
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = ring + stage * kPairStageBytes;
          uint8_t* sb = sa + kBlockM * kBlockK * 2;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kPairStageBytes);
          // completion bytes of BOTH CTAs go to the leader's barrier (peer bit cleared)
          const uint32_t leader_bar = smem_u32(&full_bar[stage]) & 0xFEFFFFFFu;
          tma_load_2d_pair(sa, &map_a, leader_bar, kb * kBlockK, m2 * 2 * kBlockM + static_cast<int>(rank) * kBlockM,
                           kEvictNormal);
          tma_load_2d_pair(sb, &map_b, leader_bar, kb * kBlockK, n_blk * kPairN + static_cast<int>(rank) * kBlockM,
                           kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * kBlockM, kPairN);
      int it = 0, astage = 0;
      uint32_t aphase = 0;
      for (int item = cluster_id; item < num_items; item += n_clusters) {
        const int kb_begin = (item / num_tiles) * kb_per;
        const int kb_end = min(num_kb, kb_begin + kb_per);
        mbar_wait(&tmem_empty_bar[astage], aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(astage * kPairN);
        for (int kb = kb_begin; kb < kb_end; ++kb, ++it) {
          const int stage = it % kPairStages;
          const uint32_t phase = static_cast<uint32_t>(it / kPairStages) & 1u;
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(ring + stage * kPairStageBytes);
          const uint64_t desc_a = umma_desc_k_sw128(sa);
          const uint64_t desc_b = umma_desc_k_sw128(sa + kBlockM * kBlockK * 2);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            umma_bf16_ss_pair(tmem_d, desc_a + static_cast<uint64_t>(k * 2), desc_b + static_cast<uint64_t>(k * 2),
                              idesc, (kb > kb_begin || k > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);
        }
        umma_commit_pair(&tmem_full_bar[astage]);
        if (++astage == kAccumStages) {
          astage = 0;
          aphase ^= 1u;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp - 4;
    constexpr bool kF32Out = (EPI == kEpiBiasResidF32);
    constexpr int kBoxCols = kF32Out ? 32 : 64;
    const int row_in_tile = q * 32 + lane;
    const uint32_t sw = static_cast<uint32_t>(row_in_tile & 7);
    int astage = 0, epi_box = 0;
    uint32_t aphase = 0;
    pdl_wait();
    for (int item = cluster_id; item < num_items; item += n_clusters) {
      const int tile = item % num_tiles;
      const bool with_bias = args.bias != nullptr && item < num_tiles;  // the first K split adds the bias
      const int m2 = tile % num_m2, n_blk = tile / num_m2;
      mbar_wait(&tmem_full_bar[astage], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < kPairN; c0 += kBoxCols) {
        const int f_box = n_blk * kPairN + c0;
        if (f_box >= args.b_rows) break;
        uint8_t* box = epi_stage + (epi_box & 1) * kEpiBoxBytes;
        if (threadIdx.x == 128) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        uint8_t* row_ptr = box + row_in_tile * 128;
#pragma unroll
        for (int cc = 0; cc < kBoxCols; cc += 32) {
          uint32_t r[32];
          tmem_ld_cols<32>(tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(astage * kPairN + c0 + cc), r);
          tmem_ld_wait();
          const int f0 = f_box + cc;
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            float v[8];
            float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
            if (with_bias && f0 + j + 7 < args.b_rows) {
              b0 = __ldg(reinterpret_cast<const float4*>(args.bias + f0 + j));
              b1 = __ldg(reinterpret_cast<const float4*>(args.bias + f0 + j + 4));
            }
            v[0] = __uint_as_float(r[j + 0]) + b0.x;
            v[1] = __uint_as_float(r[j + 1]) + b0.y;
            v[2] = __uint_as_float(r[j + 2]) + b0.z;
            v[3] = __uint_as_float(r[j + 3]) + b0.w;
            v[4] = __uint_as_float(r[j + 4]) + b1.x;
            v[5] = __uint_as_float(r[j + 5]) + b1.y;
            v[6] = __uint_as_float(r[j + 6]) + b1.z;
            v[7] = __uint_as_float(r[j + 7]) + b1.w;
            if (EPI == kEpiBiasGeluBf16) {
#pragma unroll
              for (int t = 0; t < 8; ++t) v[t] = gelu_new_fast(v[t]);
            }
            if (kF32Out) {
              const uint32_t ch = static_cast<uint32_t>((cc + j) >> 2);
              *reinterpret_cast<float4*>(row_ptr + (((ch + 0) ^ sw) << 4)) = make_float4(v[0], v[1], v[2], v[3]);
              *reinterpret_cast<float4*>(row_ptr + (((ch + 1) ^ sw) << 4)) = make_float4(v[4], v[5], v[6], v[7]);
            } else {
              const uint32_t ch = static_cast<uint32_t>((cc + j) >> 3);
              uint4 pk;
              pk.x = pack_bf16x2(v[0], v[1]);
              pk.y = pack_bf16x2(v[2], v[3]);
              pk.z = pack_bf16x2(v[4], v[5]);
              pk.w = pack_bf16x2(v[6], v[7]);
              *reinterpret_cast<uint4*>(row_ptr + ((ch ^ sw) << 4)) = pk;
            }
          }
        }
        fence_proxy_async();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 128) {
          const uint64_t desc = reinterpret_cast<uint64_t>(&map_out);
          const uint32_t src = smem_u32(box);
          const int crd0 = f_box, crd1 = m2 * 2 * kBlockM + static_cast<int>(rank) * kBlockM;
          if (kF32Out) {
            asm volatile(
                "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];"
                :
                : "l"(desc), "r"(src), "r"(crd0), "r"(crd1)
                : "memory");
          } else {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                         :
                         : "l"(desc), "r"(src), "r"(crd0), "r"(crd1)
                         : "memory");
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        ++epi_box;
      }
      // this CTA's half of the accumulator is drained: tell the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) {
          mbar_arrive(&tmem_empty_bar[astage]);
        } else {
          mbar_arrive_remote(&tmem_empty_bar[astage], 0);
        }
      }
      if (++astage == kAccumStages) {
        astage = 0;
        aphase ^= 1u;
      }
    }
    if (threadIdx.x == 128) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer may still be signalling into this CTA's barriers / reading TMEM
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---------------------------------------------------------------------------------------
// Host side: tensor maps + dispatch
// ---------------------------------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

using MapKey = std::tuple<const void*, int, int, int, int>;  // ptr, rows, cols, box_rows, elem bytes
std::map<MapKey, CUtensorMap>& map_cache() {
  static std::map<MapKey, CUtensorMap> cache;
  return cache;
}
std::mutex& map_mutex() {
  static std::mutex m;
  return m;
}

// Row-major bf16 [rows, cols] matrix; box = [box_rows, 64 cols] with 128 B swizzle.
// elem_bytes 2: bf16 operand / output (box 64 columns); 4: fp32 output (box 32 columns). Either
// way a box row is 128 bytes, the SWIZZLE_128B span.
int get_tensor_map(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out, int elem_bytes = 2) {
  std::lock_guard<std::mutex> lock(map_mutex());
  MapKey key{ptr, rows, cols, box_rows, elem_bytes};
  auto it = map_cache().find(key);
  if (it != map_cache().end()) {
    *out = it->second;
    return kOk;
  }
  EncodeTiledFn encode = get_encode_fn();
  if (encode == nullptr) return fail(kErrCuda, "cuTensorMapEncodeTiled entry point unavailable");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15u) != 0) {
    return fail(kErrInvalidArg, "GEMM operand must be 16-byte aligned for TMA");
  }
  CUtensorMap map;
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * elem_bytes};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / elem_bytes), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t elem_strides[2] = {1, 1};
  CUresult r = encode(&map, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                      2, const_cast<void*>(ptr), dims,
                      strides, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return fail(kErrCuda, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string(r));
  }
  map_cache()[key] = map;
  *out = map;
  return kOk;
}

// SM count of the CURRENT device (one process may drive several GPUs, one worker thread each).
int num_sms() {
  static std::atomic<int> cached[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// Rows per CTA for the balanced swap-AB layout, or 0 when whole tiles are as good: one token tile,
// no split-K, more 128-row weight tiles than SMs, and a smaller per-CTA maximum than striding gives.
int balanced_rows(int a_rows, int b_rows, int block_n, int split) {
  const int sms = num_sms();
  const int num_m = (a_rows + kBlockM - 1) / kBlockM;
  if (split > 1 || b_rows > block_n || num_m <= sms) return 0;
  const int rows = (((a_rows + sms - 1) / sms) + 31) / 32 * 32;
  const int strided_max = ((num_m + sms - 1) / sms) * kBlockM;
  return rows < strided_max ? rows : 0;
}
int balanced_groups(int bal_rows) { return num_sms() * ((bal_rows + kBlockM - 1) / kBlockM) * 4; }

template <int BLOCK_N, int EPI, bool TRANS, bool TMA_EPI = false>
int launch(const GemmArgs& g, cudaStream_t stream) {
  using Cfg = TileCfg<BLOCK_N>;
  const int a_rows = TRANS ? g.features : g.tokens;
  const int b_rows = TRANS ? g.tokens : g.features;
  const void* a_ptr = TRANS ? static_cast<const void*>(g.w) : static_cast<const void*>(g.act);
  const void* b_ptr = TRANS ? static_cast<const void*>(g.act) : static_cast<const void*>(g.w);
  const int act_rows = g.act_rows_alloc > 0 ? g.act_rows_alloc : g.tokens;
  const int a_rows_map = TRANS ? g.features : act_rows;
  const int b_rows_map = TRANS ? act_rows : g.features;

  CUtensorMap map_a, map_b, map_out;
  B200_PROPAGATE(get_tensor_map(a_ptr, a_rows_map, g.k, kBlockM, &map_a));
  B200_PROPAGATE(get_tensor_map(b_ptr, b_rows_map, g.k, BLOCK_N, &map_b));
  if (TMA_EPI) {
    B200_PROPAGATE(get_tensor_map(g.out, g.tokens, g.ld_out, kBlockM, &map_out,
                                  EPI == kEpiBiasResidF32 ? 4 : 2));
  } else {
    map_out = map_a;  // unused
  }

  KernelArgs ka;
  ka.bias = g.bias;
  ka.out = g.out;
  ka.a_rows = a_rows;
  ka.b_rows = b_rows;
  ka.k_blocks_total = g.k / kBlockK;
  const int split = g.split_k > 0 ? g.split_k : 1;
  ka.k_blocks_per_split = (ka.k_blocks_total + split - 1) / split;
  ka.ld_out = g.ld_out;
  ka.tokens = g.tokens;
  ka.amax_val = g.amax_val;
  ka.amax_idx = g.amax_idx;
  ka.eos_logit = g.eos_logit;
  ka.ld_amax = g.ld_amax;
  ka.eos_id = g.eos_id;

  const int num_tiles = ((a_rows + kBlockM - 1) / kBlockM) * ((b_rows + BLOCK_N - 1) / BLOCK_N);
  int grid_x = num_tiles < num_sms() ? num_tiles : num_sms();
  if (split > 1) {
    int cap = num_sms() / split;
    if (cap < 1) cap = 1;
    if (grid_x > cap) grid_x = cap;
  }
  // ring depth: enough for every k-block this CTA will ever load, at least 2, at most the max
  const int iters_per_cta = ((num_tiles + grid_x - 1) / grid_x) * ka.k_blocks_per_split;
  int stages = iters_per_cta < Cfg::kMaxStages ? iters_per_cta : Cfg::kMaxStages;
  if (stages < 2) stages = 2;
  ka.stages = stages;
  ka.single_pass = (TRANS && iters_per_cta <= stages && !g.act_via_tma) ? 1 : 0;

  auto kernel = gemm_kernel<BLOCK_N, EPI, TRANS, TMA_EPI>;
  constexpr int kEpiBytes = TMA_EPI ? 2 * kEpiBoxBytes : 0;
  // the opt-in shared-memory limit is a per-device function attribute
  static bool attr_set[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::smem_bytes(Cfg::kMaxStages) + kEpiBytes));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  // balanced row ranges (KernelArgs::bal_rows): LM-head shaped launches only, own instantiation
  ka.bal_rows = 0;
  ka.bal_tail_rows = 0;
  if constexpr (TRANS && !TMA_EPI && (EPI == kEpiF32 || EPI == kEpiLogitsArgmax)) {
    const int bal = g.act_via_tma ? 0 : balanced_rows(a_rows, b_rows, BLOCK_N, split);
    if (bal > 0 && !ka.single_pass &&
        (EPI != kEpiLogitsArgmax || balanced_groups(bal) <= g.amax_groups_alloc)) {
      ka.bal_rows = bal;
      ka.bal_tail_rows = bal % kBlockM;
      if (ka.bal_tail_rows > 0) {
        B200_PROPAGATE(get_tensor_map(a_ptr, a_rows_map, g.k, ka.bal_tail_rows, &map_out));
      }
      kernel = gemm_kernel<BLOCK_N, EPI, TRANS, TMA_EPI, true>;
      static bool bal_attr_set[64] = {};
      if (dev < 0 || dev >= 64 || !bal_attr_set[dev]) {
        B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          Cfg::smem_bytes(Cfg::kMaxStages) + kEpiBytes));
        if (dev >= 0 && dev < 64) bal_attr_set[dev] = true;
      }
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid_x, split, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::smem_bytes(stages) + kEpiBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = g.pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, map_a, map_b, map_out, ka));
  return kOk;
}

template <int BLOCK_N, bool TRANS>
int dispatch_epi(const GemmArgs& g, cudaStream_t stream) {
  if constexpr (!TRANS) {
    // staged TMA-store epilogue when the output satisfies TMA's alignment rules
    const int eb = g.epilogue == kEpiBiasResidF32 ? 4 : 2;
    const bool tma_ok = (reinterpret_cast<uintptr_t>(g.out) & 15u) == 0 &&
                        (static_cast<size_t>(g.ld_out) * eb) % 16 == 0 && g.features % 8 == 0 &&
                        (g.bias == nullptr || (reinterpret_cast<uintptr_t>(g.bias) & 15u) == 0);
    if (tma_ok) {
      switch (g.epilogue) {
        case kEpiBiasBf16:
          return launch<BLOCK_N, kEpiBiasBf16, false, true>(g, stream);
        case kEpiBiasGeluBf16:
          return launch<BLOCK_N, kEpiBiasGeluBf16, false, true>(g, stream);
        case kEpiBiasResidF32:
          return launch<BLOCK_N, kEpiBiasResidF32, false, true>(g, stream);
        default:
          break;
      }
    }
  }
  switch (g.epilogue) {
    case kEpiBiasBf16:
      return launch<BLOCK_N, kEpiBiasBf16, TRANS>(g, stream);
    case kEpiBiasGeluBf16:
      return launch<BLOCK_N, kEpiBiasGeluBf16, TRANS>(g, stream);
    case kEpiBiasResidF32:
      return launch<BLOCK_N, kEpiBiasResidF32, TRANS>(g, stream);
    case kEpiF32:
      return launch<BLOCK_N, kEpiF32, TRANS>(g, stream);
    case kEpiPartialF32:
      return launch<BLOCK_N, kEpiPartialF32, TRANS>(g, stream);
    case kEpiLogitsArgmax:
      if constexpr (TRANS) {
        return launch<BLOCK_N, kEpiLogitsArgmax, true>(g, stream);
      } else {
        return fail(kErrUnsupported, "gemm: the fused-argmax epilogue needs the swap-AB orientation");
      }
    default:
      return fail(kErrInvalidArg, "unknown GEMM epilogue");
  }
}


template <int BLOCK_N, int EPI>
int launch_cluster(const GemmArgs& g, cudaStream_t stream) {
  using Cfg = TileCfg<BLOCK_N>;
  CUtensorMap map_a, map_b;
  const int act_rows = g.act_rows_alloc > 0 ? g.act_rows_alloc : g.tokens;
  B200_PROPAGATE(get_tensor_map(g.w, g.features, g.k, kBlockM, &map_a));
  B200_PROPAGATE(get_tensor_map(g.act, act_rows, g.k, BLOCK_N, &map_b));
  const int split = g.cluster_split;
  KernelArgs ka = {};
  ka.bias = g.bias;
  ka.out = g.out;
  ka.a_rows = g.features;
  ka.b_rows = g.tokens;
  ka.k_blocks_total = g.k / kBlockK;
  ka.k_blocks_per_split = ka.k_blocks_total / split;
  ka.ld_out = g.ld_out;
  ka.tokens = g.tokens;
  int stages = ka.k_blocks_per_split < Cfg::kMaxStages ? ka.k_blocks_per_split : Cfg::kMaxStages;
  if (stages < 2) stages = 2;
  ka.stages = stages;
  const int ring_bytes = stages * Cfg::kStageBytes;
  const int tile_bytes = BLOCK_N * kBlockM * 4;
  const int smem_bytes = (ring_bytes > tile_bytes ? ring_bytes : tile_bytes) + 1024 + 256;
  const int max_bytes = (Cfg::kMaxStages * Cfg::kStageBytes > tile_bytes ? Cfg::kMaxStages * Cfg::kStageBytes : tile_bytes) + 1024 + 256;

  auto kernel = gemm_cluster_kernel<BLOCK_N, EPI>;
  static bool attr_set[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_bytes));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((g.features + kBlockM - 1) / kBlockM, split, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = 1;
  attrs[0].val.clusterDim.y = split;
  attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[1].val.programmaticStreamSerializationAllowed = g.pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 2;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, map_a, map_b, ka));
  return kOk;
}

template <int BLOCK_N>
int dispatch_cluster(const GemmArgs& g, cudaStream_t stream) {
  switch (g.epilogue) {
    case kEpiBiasBf16:
      return launch_cluster<BLOCK_N, kEpiBiasBf16>(g, stream);
    case kEpiBiasGeluBf16:
      return launch_cluster<BLOCK_N, kEpiBiasGeluBf16>(g, stream);
    case kEpiBiasResidF32:
      return launch_cluster<BLOCK_N, kEpiBiasResidF32>(g, stream);
    default:
      return fail(kErrUnsupported, "gemm: the cluster split-K path supports the bias / gelu / residual epilogues");
  }
}


template <int EPI>
int launch_pair(const GemmArgs& g, cudaStream_t stream) {
  CUtensorMap map_a, map_b, map_out;
  const int act_rows = g.act_rows_alloc > 0 ? g.act_rows_alloc : g.tokens;
  B200_PROPAGATE(get_tensor_map(g.act, act_rows, g.k, kBlockM, &map_a));
  B200_PROPAGATE(get_tensor_map(g.w, g.features, g.k, kBlockM, &map_b));
  B200_PROPAGATE(get_tensor_map(g.out, g.tokens, g.ld_out, kBlockM, &map_out, EPI == kEpiBiasResidF32 ? 4 : 2));
  KernelArgs ka = {};
  ka.bias = g.bias;
  ka.out = g.out;
  ka.a_rows = g.tokens;
  ka.b_rows = g.features;
  ka.k_blocks_total = g.k / kBlockK;
  ka.ld_out = g.ld_out;
  ka.tokens = g.tokens;
  ka.stages = kPairStages;
  const int num_tiles = ((g.tokens + 2 * kBlockM - 1) / (2 * kBlockM)) * ((g.features + kPairN - 1) / kPairN);
  const int max_clusters = num_sms() / 2;
  // K splits of a tile (residual epilogue only: its reduce-add needs no extra pass). Cost model: waves of
  // work items x (k-blocks per item + 12: the fill, and every item sends its whole fp32 tile through the
  // L2 adders again); an item keeps at least 12 k-blocks. GPT-2 small, 8192 tokens, MLP down-projection
  // (96 tiles on 74 CTA pairs, 48 k-blocks): two waves become three of half the length; GPT-2 medium
  // (256 tiles: 3.46 waves) stays unsplit -- measured 12.32 vs 12.37 ms per prefill with two splits.
  int split = 1;
  if (EPI == kEpiBiasResidF32 && g.pair_split > 1) {
    split = g.pair_split < ka.k_blocks_total ? g.pair_split : ka.k_blocks_total;  // forced (tests, sweeps)
  } else if (EPI == kEpiBiasResidF32 && g.pair_split < 0) {
    int best = 1 << 30;
    for (int s = 1; s <= 8; ++s) {
      if (s > 1 && ka.k_blocks_total / s < 12) break;
      const int per = (ka.k_blocks_total + s - 1) / s;
      const int waves = (num_tiles * s + max_clusters - 1) / max_clusters;
      const int cost = waves * (per + 12);
      if (cost < best) {
        best = cost;
        split = s;
      }
    }
  }
  ka.k_blocks_per_split = (ka.k_blocks_total + split - 1) / split;
  const int num_items = num_tiles * ((ka.k_blocks_total + ka.k_blocks_per_split - 1) / ka.k_blocks_per_split);
  const int clusters = num_items < max_clusters ? num_items : max_clusters;

  auto kernel = gemm_pair_kernel<EPI>;
  static bool attr_set[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPairSmemBytes));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = kPairSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = g.pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, map_a, map_b, map_out, ka));
  return kOk;
}

}  // namespace

B200_DEFINE_STEP_TRACE_SETTER(step_trace_set_gemm)

int device_sm_count() { return num_sms(); }

int gemm_tensor_map(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out) {
  return get_tensor_map(ptr, rows, cols, box_rows, out);
}

int gemm_tensor_map_f32(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out) {
  return get_tensor_map(ptr, rows, cols, box_rows, out, 4);
}

void gemm_clear_tensor_map_cache() {
  std::lock_guard<std::mutex> lock(map_mutex());
  map_cache().clear();
}

int gemm_argmax_groups(int features, int tokens) {
  int block_n = 256;
  if (tokens <= 16) block_n = 16;
  else if (tokens <= 32) block_n = 32;
  else if (tokens <= 64) block_n = 64;
  else if (tokens <= 128) block_n = 128;
  const int bal = balanced_rows(features, tokens, block_n, 1);
  return bal > 0 ? balanced_groups(bal) : ((features + kBlockM - 1) / kBlockM) * 4;
}

int gemm_bf16(const GemmArgs& g, cudaStream_t stream) {
  B200_CHECK_ARG(g.act && g.w && (g.out || g.epilogue == kEpiLogitsArgmax), "gemm: null operand");
  B200_CHECK_ARG(g.epilogue != kEpiLogitsArgmax || (g.amax_val && g.amax_idx && g.eos_logit),
                 "gemm: fused-argmax epilogue needs its output arrays");
  B200_CHECK_ARG(g.tokens > 0 && g.features > 0 && g.k > 0, "gemm: empty problem");
  B200_CHECK_ARG(g.k % kBlockK == 0, "gemm: k must be a multiple of 64");
  B200_CHECK_ARG(g.split_k <= 1 || g.epilogue == kEpiPartialF32,
                 "gemm: split_k > 1 requires the partial epilogue");
  B200_CHECK_ARG(g.split_k <= 1 || (g.k / kBlockK) % g.split_k == 0,
                 "gemm: split_k must divide the number of 64-wide k blocks");

  if (g.cluster_split > 1) {
    // split-K across a thread-block cluster with an on-chip (DSMEM) reduction
    B200_CHECK_ARG(g.cluster_split <= 8, "gemm: cluster split-K is limited to the portable cluster size 8");
    B200_CHECK_ARG((g.k / kBlockK) % g.cluster_split == 0, "gemm: cluster_split must divide the k blocks");
    B200_CHECK_ARG(g.tokens <= 256 && g.features % 4 == 0 && g.ld_out % 4 == 0 && g.split_k <= 1,
                   "gemm: cluster split-K needs <= 256 tokens and 4-aligned features");
    if (g.tokens <= 16) return dispatch_cluster<16>(g, stream);
    if (g.tokens <= 32) return dispatch_cluster<32>(g, stream);
    if (g.tokens <= 64) return dispatch_cluster<64>(g, stream);
    if (g.tokens <= 128) return dispatch_cluster<128>(g, stream);
    return dispatch_cluster<256>(g, stream);
  }

  bool trans;
  if (g.force_orientation == 0) {
    trans = false;
  } else if (g.force_orientation == 1) {
    trans = true;
  } else {
    // Up to 256 tokens fit one UMMA N tile: stream the weights once, tokens on N.
    trans = g.tokens <= 256;
  }

  if (trans) {
    if (g.tokens <= 16) return dispatch_epi<16, true>(g, stream);
    if (g.tokens <= 32) return dispatch_epi<32, true>(g, stream);
    if (g.tokens <= 64) return dispatch_epi<64, true>(g, stream);
    if (g.tokens <= 128) return dispatch_epi<128, true>(g, stream);
    return dispatch_epi<256, true>(g, stream);
  }
  // tokens on M. Large problems whose output can go through TMA use CTA pairs (UMMA M = 256).
  {
    const int eb = g.epilogue == kEpiBiasResidF32 ? 4 : 2;
    const bool tma_ok = (reinterpret_cast<uintptr_t>(g.out) & 15u) == 0 &&
                        (static_cast<size_t>(g.ld_out) * eb) % 16 == 0 &&
                        (g.bias == nullptr || (reinterpret_cast<uintptr_t>(g.bias) & 15u) == 0);
    const long pair_tiles = static_cast<long>((g.tokens + 255) / 256) * (g.features / 256);
    const bool pair_epi = g.epilogue == kEpiBiasBf16 || g.epilogue == kEpiBiasGeluBf16 ||
                          g.epilogue == kEpiBiasResidF32;
    if (g.use_pair != 0 && tma_ok && pair_epi && g.features % 256 == 0 && g.split_k <= 1 &&
        (g.use_pair > 0 || pair_tiles >= num_sms() / 2)) {
      switch (g.epilogue) {
        case kEpiBiasBf16:
          return launch_pair<kEpiBiasBf16>(g, stream);
        case kEpiBiasGeluBf16:
          return launch_pair<kEpiBiasGeluBf16>(g, stream);
        default:
          return launch_pair<kEpiBiasResidF32>(g, stream);
      }
    }
  }
  // Otherwise single-CTA tiles; wide N tiles when there is enough work to fill the machine.
  const long tiles256 = static_cast<long>((g.tokens + 127) / 128) * ((g.features + 255) / 256);
  if (g.features % 256 == 0 && tiles256 >= num_sms()) return dispatch_epi<256, false>(g, stream);
  return dispatch_epi<128, false>(g, stream);
}

}  // namespace b200llm
