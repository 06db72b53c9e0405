// Host-visible interface of the tcgen05/TMA GEMM (gemm_tcgen05.cu).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200llm {

// Epilogue selector. "token" / "feature" below are logical output coordinates: the GEMM
// always produces out[token, feature] = sum_k act[token, k] * w[feature, k].
enum GemmEpilogue : int {
  kEpiBiasBf16 = 0,      // out_bf16 = acc + bias[feature]
  kEpiBiasGeluBf16 = 1,  // out_bf16 = gelu_new(acc + bias[feature])
  kEpiBiasResidF32 = 2,  // out_f32 += acc + bias[feature]   (in-place residual stream)
  kEpiF32 = 3,           // out_f32 = acc                    (LM-head logits, no bias)
  kEpiPartialF32 = 4,    // split-K partial: out_f32[split][token, feature] = acc
  kEpiLogitsArgmax = 5,  // LM head, swap-AB only: per-(32-feature group, token) running argmax
                         // written to amax_val/amax_idx (EOS excluded, its logit goes to
                         // eos_logit); fp32 logits also stored when out != null
};

struct GemmArgs {
  const __nv_bfloat16* act;  // [tokens, k] row-major
  const __nv_bfloat16* w;    // [features, k] row-major (pre-transposed Conv1D weight)
  const float* bias;         // [features] or null
  void* out;                 // [tokens, ld_out] bf16 or fp32 depending on epilogue
  int tokens;
  int features;
  int k;
  int ld_out;
  int epilogue;
  int split_k;         // >= 1; > 1 only with kEpiPartialF32 (partials reduced by the consumer)
  int act_rows_alloc;  // rows backing `act` (>= tokens) for the tensor map
  int force_orientation;  // -1 auto, 0 tokens-on-M, 1 weights-on-M (swap-AB)
  int pdl;                // launch with programmatic stream serialization
  // kEpiLogitsArgmax only: [ceil(features/128)*4, ld_amax] group maxima / indices, [tokens] EOS
  float* amax_val;
  int32_t* amax_idx;
  float* eos_logit;
  int ld_amax;
  int eos_id;
  // > 1: swap-AB split-K over a thread-block cluster of this many CTAs, reduced through
  // distributed shared memory inside the kernel (deterministic; final epilogue applied)
  int cluster_split;
  // tokens-on-M orientation: -1 = CTA pairs (cta_group::2, 256 x 256 tiles) when the problem is
  // large enough, 0 = never, 1 = always when legal. Zero-initialised structs get "never"; the
  // engine and the op-level ABI pass -1.
  int use_pair;
  // CTA pairs with the residual epilogue (out += act . w^T + bias by TMA reduce-add): K splits per output
  // tile. 0 / 1 = one (every output element receives ONE add: bit-reproducible), -1 = chosen by the wave
  // cost model, n > 1 = forced. With splits the fp32 adds of a tile land in no fixed order.
  int pair_split;
  // swap-AB launches whose CTAs keep all their k-blocks in the ring copy the activation operand
  // with ordinary loads; 1 forces the TMA path for it (A/B comparisons, tests)
  int act_via_tma;
  int amax_groups_alloc;  // kEpiLogitsArgmax: group rows the amax arrays can hold (see gemm_argmax_groups)
};

// Enqueue one GEMM on `stream`. Returns 0 or an error code (see common.cuh Status).
int gemm_bf16(const GemmArgs& args, cudaStream_t stream);

// kEpiLogitsArgmax: number of candidate groups (rows of amax_val / amax_idx) a launch with this
// shape writes or leaves untouched; untouched ones must hold -inf (the arrays start that way and
// no launch shape ever writes a group another shape leaves out).
int gemm_argmax_groups(int features, int tokens);

// Cached SWIZZLE_128B tensor map of a row-major bf16 [rows, cols] matrix with boxes of
// [box_rows, 64 columns] (the operand layout every UMMA descriptor in this library assumes).
int gemm_tensor_map(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out);
// The same for an fp32 matrix: boxes of [box_rows, 32 columns] (a box row is again 128 bytes).
int gemm_tensor_map_f32(const void* ptr, int rows, int cols, int box_rows, CUtensorMap* out);

// SM count of the current device (cached per device).
int device_sm_count();

// Drop cached tensor maps (engine teardown / tests).
void gemm_clear_tensor_map_cache();

}  // namespace b200llm
