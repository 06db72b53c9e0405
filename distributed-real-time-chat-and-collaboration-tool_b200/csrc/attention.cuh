// Causal self-attention for GPT-2 (head_dim 64): a tensor-core prefill kernel over packed,
// ragged prompts and a bandwidth-oriented decode kernel over the paged KV cache.
//
// Oracle: eager_attention_forward (modeling_gpt2.py:54-72) / sdpa_attention_forward
// (integrations/sdpa_attention.py:92-101): softmax(Q K^T / sqrt(64) + causal mask) V with the
// softmax evaluated in fp32; KV cache growth at cache_utils.py:119-120.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200llm {

// Prefill: qkv is the packed [n_tok, 3*d] bf16 output of the QKV GEMM (q | k | v along the
// feature axis, each [n_head, 64]); sequence b owns rows [cu_seqlens[b], cu_seqlens[b+1]).
// out is [n_tok, d] bf16 with heads merged (modeling_gpt2.py:222).
// When `cache` is given, the CTA on the diagonal (query tile == key tile) also writes that
// K/V tile into the sequence's KV pages, so prefill needs no separate scatter pass.
struct PrefillCache {
  const int32_t* slots = nullptr;       // [batch] slot per sequence; null = do not fill pages
  const int32_t* page_table = nullptr;  // [n_slots, max_pages_per_seq]
  int max_pages_per_seq = 0;
  __nv_bfloat16* k_pool = nullptr;
  __nv_bfloat16* v_pool = nullptr;
};
int prefill_attention(const __nv_bfloat16* qkv, const int32_t* cu_seqlens, int batch,
                      int max_seq_len, int n_head, __nv_bfloat16* out, cudaStream_t stream,
                      const PrefillCache* cache = nullptr);

// Shared-prefix reuse: prefill over prompt SUFFIXES. Sequence b already has prefix_len[b] tokens in
// its KV pages; rows [cu_seqlens[b], cu_seqlens[b+1]) of qkv are its remaining tokens (positions
// prefix_len[b] + 0, 1, ...). kv_scatter writes their K/V to the pages, then the attention kernel
// reads every key -- cached or just written -- from the pages (modeling_gpt2.py:194-196: the past
// is concatenated in front of the new keys; here it is already in place).
int kv_scatter(const __nv_bfloat16* qkv, const int32_t* row_pos, const int32_t* row_slot, int n_tok,
               int n_head, const int32_t* page_table, int max_pages_per_seq, __nv_bfloat16* k_pool,
               __nv_bfloat16* v_pool, cudaStream_t stream);
struct PagedPrefillArgs {
  const __nv_bfloat16* qkv = nullptr;     // [n_rows, 3*d]; only the q third is read
  const int32_t* cu_seqlens = nullptr;    // [batch+1] prefix sums of SUFFIX lengths
  const int32_t* prefix_len = nullptr;    // [batch]
  const int32_t* slots = nullptr;         // [batch]
  int batch = 0;
  int max_suffix_len = 0;
  int n_head = 0;
  const int32_t* page_table = nullptr;
  int max_pages_per_seq = 0;
  const __nv_bfloat16* k_pool = nullptr;
  const __nv_bfloat16* v_pool = nullptr;
  __nv_bfloat16* out = nullptr;           // [n_rows, d]
  int qkv_rows = 0;                       // rows backing qkv   } only the tcgen05 kernel needs these
  int n_pages = 0;                        // pages in each pool } (extents of its TMA tensor maps)
};
// The engine's prefill attention (attention_tc.cu): tcgen05.mma for Q K^T and P V, accumulators
// in TMEM, Q / K / V tiles by TMA.
int prefill_attention_tc(const PagedPrefillArgs& a, cudaStream_t stream);
// The same computation on the legacy tensor path (mma.sync + cp.async); kept as the A/B partner of
// the kernel above (b200llm_op_prefill_attention_paged impl = 1).
int prefill_attention_paged(const PagedPrefillArgs& a, cudaStream_t stream);

struct DecodeAttnArgs {
  const __nv_bfloat16* qkv;      // [batch, 3*d] current token (bias already added), or null when
                                 // the projection arrives as split-K partial planes:
  const float* qkv_partials;     // [n_qkv_partials][partial_rows, 3*d] fp32 (no bias)
  int n_qkv_partials;
  int partial_rows;
  const float* qkv_bias;         // [3*d] added while reducing the partials
  const int32_t* active_slots;   // [batch] slot per row (-1 = padding)
  const int32_t* seq_len;        // [n_slots] tokens already cached == position of this token
  const int32_t* finished;       // [n_slots] or null; finished rows do not append
  const int32_t* page_table;     // [n_slots, max_pages_per_seq]
  int max_pages_per_seq;
  __nv_bfloat16* k_pool;         // [n_pages, n_head, 16, 64]
  __nv_bfloat16* v_pool;
  __nv_bfloat16* out;            // [batch, d]
  int batch;
  int n_head;
  int n_splits;                  // KV splits per (row, head); > 1 needs scratch
  float* scratch;                // n_splits > 1: [batch*n_head*n_splits*(64+2)] fp32
  int pdl;
  // 2..8 splits only: launch the splits of one (row, head) as a thread-block cluster and merge
  // their partial results through distributed shared memory inside the kernel (no scratch, no
  // combine launch). The engine sets it; the operator entry points keep the scratch path.
  int cluster_combine;
  // Streaming form (persistent warps, one unit per warp at a time, K/V tiles by TMA): the engine picks
  // it when the (row, head, split) units outnumber the resident CTAs of the kernel above or contexts
  // run to several 128-row passes. Needs n_pages (extent of the tensor maps over the pools).
  int stream_kernel;
  int n_pages;
  // Streaming form with n_splits > 1: [batch * n_head] counters, zero before the first launch and left
  // at zero by every launch; the split of a (row, head) that finishes last merges them in the kernel.
  // Null: decode_combine_kernel is launched instead.
  int32_t* split_counters;
  // Streaming form: device int with the number of rows that hold a sequence (they are the first rows of
  // the batch); null = all `batch` rows. The units of the padding rows are then not handed out at all,
  // so a partly filled bucket still spreads over every SM.
  const int32_t* n_live;
};
// One decode step of attention: appends this token's K/V to its pages, then attends over
// positions 0..seq_len (inclusive).
int decode_attention(const DecodeAttnArgs& a, cudaStream_t stream);

// Floats of scratch decode_attention needs for (batch, n_head, n_splits).
size_t decode_attention_scratch_floats(int batch, int n_head, int n_splits);

// Same step, but the split-KV partials live in a fixed array that is part of the loaded
// library image instead of a caller-provided buffer (a.scratch is ignored). See the comment
// on b200llm_op_decode_attention_noalloc in include/b200llm.h for who needs this and its rules.
constexpr size_t kNoAllocScratchFloats = 1u << 20;  // 4 MiB per GPU that loads the library
int decode_attention_noalloc(DecodeAttnArgs a, cudaStream_t stream);

}  // namespace b200llm
