// Memory-bound kernels around the GEMMs: embedding gather-add, LayerNorm (optionally fused
// with the residual add + split-K reduction that precedes it), greedy / sampled token
// selection with EOS bookkeeping, KV page scatter, weight conversion.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200llm {

constexpr int kPageTokens = 16;  // tokens per KV page
constexpr int kHeadDim = 64;     // all GPT-2 sizes

// x[t, :] = wte[ids[t], :] + wpe[pos[t], :]      (modeling_gpt2.py:577-585)
// Prefill form: ids/pos are per packed token row.
int embed_tokens(const float* wte, const float* wpe, const int32_t* ids, const int32_t* pos,
                 float* x, int n_tok, int d, int vocab, int n_positions, cudaStream_t stream);

// Decode form: row b reads its token and position from per-slot generation state.
int embed_decode(const float* wte, const float* wpe, const int32_t* active_slots,
                 const int32_t* last_tok, const int32_t* seq_len, float* x, int batch, int d,
                 int vocab, int n_positions, cudaStream_t stream, int pdl = 0);

struct LayerNormArgs {
  float* x;                  // [rows, d] fp32 residual stream (updated in place if partials)
  const float* partials;     // [n_partials, rows, d] fp32 split-K partial sums or null
  int n_partials;
  int partial_rows;          // row stride between partial planes (0 -> rows)
  const float* add_bias;     // [d] bias added with the partials (or null)
  const float* gamma;        // [d]
  const float* beta;         // [d]
  const int32_t* row_index;  // optional gather: out row i normalises x[row_index[i]]
  __nv_bfloat16* out;        // [rows_out, d] bf16
  int rows;                  // rows_out
  int d;
  float eps;
  int pdl;
};
// out = LayerNorm(x (+ bias + sum partials)) with biased variance, eps inside the sqrt
// (modeling_gpt2.py:252-254, 273, 304, 628; torch.nn.functional.layer_norm).
int layernorm(const LayerNormArgs& a, cudaStream_t stream);

struct SelectArgs {
  const float* logits;           // [batch, ld] fp32 (may be null when every row is greedy and
                                 // the LM-head GEMM supplied group candidates)
  int ld;
  int vocab;
  // candidates from the LM-head GEMM's fused argmax (gemm_tcgen05.cuh kEpiLogitsArgmax)
  const float* amax_val;         // [n_groups, ld_amax] or null
  const int32_t* amax_idx;
  const float* eos_logit;        // [batch]
  int n_groups;
  int ld_amax;
  const int32_t* active_slots;   // [batch] slot per row (-1 = padding row)
  int batch;
  // per-slot generation state
  int32_t* last_tok;
  int32_t* seq_len;
  int32_t* finished;
  int32_t* n_gen;
  const int32_t* min_new;
  const int32_t* max_new;
  int32_t* out_tokens;           // [n_slots, out_stride]
  int out_stride;
  int eos_id;
  int pad_id;
  // per-slot sampling controls; null or temperature <= 0 or top_k <= 1 -> greedy argmax
  const float* temperature;
  const int32_t* top_k;
  const uint64_t* seed;
  int advance_seq;               // 1: decode step (cache grew by one token), 0: after prefill
  int pdl;                       // launch with programmatic stream serialization
};
// Greedy: argmax over fp32 logits, ties to the lowest index (generation/utils.py:2762,2793);
// finished rows emit pad (:2796-2797); EOS marks a row finished (stopping_criteria.py:450-471).
int select_tokens(const SelectArgs& a, cudaStream_t stream);

// out_bf16[r, c] = gelu_new(bias[c] + sum_s partials[s][r, c]): finishes a split-K MLP
// up-projection (modeling_gpt2.py:239-240) so the down-projection can read it through TMA.
int reduce_bias_gelu(const float* partials, int n_partials, int partial_rows, const float* bias,
                     __nv_bfloat16* out, int rows, int cols, int pdl, cudaStream_t stream);

// Plain row-wise argmax (unit tests / teacher-forced parity).
int argmax_rows(const float* logits, int rows, int ld, int vocab, int32_t* out,
                cudaStream_t stream);



// Weight preparation (load time): fp32 [rows, cols] -> bf16 [cols, rows] (Conv1D [in,out] to
// K-major [out,in]) and plain fp32 -> bf16.
int convert_transpose_bf16(const float* src, __nv_bfloat16* dst, int rows, int cols,
                           cudaStream_t stream);
int convert_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t stream);

}  // namespace b200llm
