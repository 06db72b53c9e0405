// Fused decode-step GEMMs (decode_fused.cu): LayerNorm folded into the consuming GEMM's operand
// path, and split-K GEMMs that add their partial tiles into the fp32 residual stream in L2.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200llm {

// out[row, f] = epilogue( sum_k LayerNorm(x[row, :])[k] * w[f, k] + bias[f] ), bf16 out.
// modeling_gpt2.py:273 + :185 (ln_1 -> c_attn), :304 + :239-240 (ln_2 -> c_fc -> gelu_new).
struct LnGemmArgs {
  const float* x = nullptr;        // [rows, d] fp32 residual stream, complete (no pending partials)
  const float* gamma = nullptr;    // [d]
  const float* beta = nullptr;     // [d]
  float eps = 1e-5f;
  const __nv_bfloat16* w = nullptr;  // [features, d] K-major
  const float* bias = nullptr;     // [features] or null
  __nv_bfloat16* out = nullptr;    // [rows, ld_out]
  int rows = 0;                    // 1..64
  int d = 0;                       // multiple of 64
  int features = 0;
  int ld_out = 0;
  int gelu = 0;                    // 1: gelu_new after the bias
  int pdl = 0;
  // split > 1: K split over `split` CTAs per tile; each writes an fp32 partial plane
  // partial_out[split][plane_rows, ld_out] (no bias, no activation) for the consumer to reduce in a
  // fixed order (decode attention does, for the QKV projection). Light CTAs: d / split of the weights.
  int split = 1;
  float* partial_out = nullptr;
  int plane_rows = 0;
};
bool ln_gemm_supported(int rows, int d, int features);
bool ln_gemm_weights_resident(int d);  // the shape class for which the engine picks the fused path by itself
int ln_gemm(const LnGemmArgs& a, cudaStream_t stream);

// x[row, f] += sum_k act[row, k] * w[f, k] + bias[f]   (fp32, added in L2 by TMA reduce-add; the
// order of the `split` additions per element is not fixed). modeling_gpt2.py:223 + :282, :241 + :307.
struct GemmRedAddArgs {
  const __nv_bfloat16* act = nullptr;  // [rows, k]
  const __nv_bfloat16* w = nullptr;    // [features, k] K-major
  const float* bias = nullptr;         // [features] or null
  float* x = nullptr;                  // [rows, features] fp32, 16-byte aligned, features % 32 == 0
  int rows = 0;                        // 1..64
  int features = 0;
  int k = 0;
  int split = 1;                       // must divide k / 64; see gemm_redadd_pick_split
  int act_rows_alloc = 0;              // rows backing act (>= 16/32/64, the token tile) or 0 = rows
  int pdl = 0;
};
// Split factor that gives the most CTAs within one wave while a CTA's k-blocks fit shared memory.
int gemm_redadd_pick_split(int features, int k, int rows, int sms);
int gemm_redadd(const GemmRedAddArgs& a, cudaStream_t stream);

}  // namespace b200llm
