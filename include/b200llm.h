/*
 * b200llm.h -- C ABI of the B200-native text generator that replaces the remote
 * `self.model.generate_content(...)` call behind the reference's llm.LLMService
 * (reference: llm_server/llm_server.py:32-33 creates the generator; :167, :231, :287, :403
 * are the four call sites).  The reference ships no native code and no local model; the
 * arithmetic this library reproduces is GPT-2 as implemented by Hugging Face transformers
 * (the dependency pinned at reference requirements.txt:6-7), cited per entry point below
 * with $TF = site-packages/transformers.
 *
 * Conventions
 *   - Every function returns 0 on success or a positive error code; b200llm_last_error()
 *     returns the reason for the calling thread's last failure. Nothing throws across the ABI.
 *   - No torch types: plain pointers, sizes and an opaque engine handle.
 *   - "host" / "dev" in a parameter name says where the pointer must live.
 *   - `stream` arguments are `cudaStream_t` passed as `void*` (NULL = the legacy default
 *     stream). Engine calls run on the engine's own stream unless b200llm_engine_set_stream
 *     installed another one.
 *   - One engine belongs to one GPU and must be driven from one thread at a time (the
 *     scheduler thread of that GPU); op-level calls are thread safe.
 */
#ifndef B200LLM_H_
#define B200LLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: b200llm_config starts with `struct_size`; batched token read, cancel, prefix-cache entry
 *    points added. A library and a binding that disagree on the struct are refused at
 *    b200llm_engine_create instead of reading each other's fields at the wrong offsets. */
#define B200LLM_ABI_VERSION 2

enum b200llm_status {
  B200LLM_OK = 0,
  B200LLM_ERR_INVALID_ARG = 1,
  B200LLM_ERR_CUDA = 2,
  B200LLM_ERR_OUT_OF_MEMORY = 3,
  B200LLM_ERR_CAPACITY = 4,
  B200LLM_ERR_UNSUPPORTED = 5
};

/* GEMM epilogues (b200llm_op_gemm). */
enum b200llm_epilogue {
  B200LLM_EPI_BIAS_BF16 = 0,      /* out_bf16 = acc + bias                                  */
  B200LLM_EPI_BIAS_GELU_BF16 = 1, /* out_bf16 = gelu_new(acc + bias)  ($TF/activations.py:59-66) */
  B200LLM_EPI_BIAS_RESID_F32 = 2, /* out_f32 += acc + bias            (modeling_gpt2.py:282,307) */
  B200LLM_EPI_F32 = 3,            /* out_f32 = acc                    (LM head, :705-706)        */
  B200LLM_EPI_PARTIAL_F32 = 4     /* split-K partial planes [split][tokens][ld_out]          */
};

typedef struct b200llm_engine b200llm_engine;

/* Model + capacity description. Mirrors transformers GPT2Config (n_layer, n_head, n_embd,
 * vocab_size, n_positions, layer_norm_epsilon, eos_token_id) plus engine capacities. */
typedef struct b200llm_config {
  uint32_t struct_size;     /* = sizeof(b200llm_config) of the header the caller was built against;
                               anything else is rejected with B200LLM_ERR_INVALID_ARG             */
  int32_t n_layer;
  int32_t n_head;
  int32_t d_model;          /* n_embd; head_dim must be 64 (all GPT-2 sizes)               */
  int32_t vocab_size;       /* 50257                                                       */
  int32_t n_positions;      /* rows of the learned position table                          */
  float ln_eps;             /* 1e-5                                                        */
  int32_t eos_token_id;     /* 50256                                                       */
  int32_t pad_token_id;     /* value finished rows emit ($TF/generation/utils.py:2796-2797) */
  int32_t max_seqs;         /* concurrent sequences (KV slots)                             */
  int32_t max_seq_len;      /* per-sequence KV capacity in tokens (prompt + generated);
                               must not exceed n_positions (there is no position row beyond it) */
  int32_t max_batch_tokens; /* packed prompt tokens per prefill call                       */
  int32_t max_new_tokens;   /* per-sequence cap on generated tokens kept on device         */
  int64_t kv_pool_tokens;   /* total KV tokens in the paged pool; 0 = max_seqs*max_seq_len */
  int32_t device;           /* CUDA device ordinal                                         */
  int32_t use_cuda_graphs;  /* 1 = replay the decode step from a captured CUDA graph       */
  int32_t use_pdl;          /* 1 = programmatic dependent launch between step kernels      */
  int32_t decode_split_k;   /* 0 = auto; 1 = off; >1 = split-K factor cap for decode GEMMs */
  int32_t keep_logits;      /* 1 = decode steps also store the fp32 logits (greedy rows do not
                               need them: the LM-head GEMM fuses the argmax)                 */
  int32_t decode_reduce;    /* how decode GEMMs combine their K splits:
                               0 = automatic: the fused path (3) for up to 64 rows when d_model <= 768
                                   (where it was measured faster), else planes (1);
                               1 = fp32 partial planes in L2 reduced in a fixed order by the consumer
                                   kernel: 8 kernels per layer, bit-reproducible;
                               2 = inside the GEMM, over a thread-block cluster through distributed
                                   shared memory (DSMEM moves ~20 B/clk per SM: measured slower);
                               3 = fused (<= 64 rows): LayerNorm inside the QKV / MLP-up GEMM, out-proj /
                                   MLP-down add their splits into the residual stream in L2 (TMA
                                   reduce-add): 5 kernels per layer; the fp32 additions of one element
                                   arrive in no fixed order (last-bit differences run to run)      */
  int32_t prefix_cache;     /* 1 = shared-prefix KV reuse: the full 16-token pages of every prompt
                               stay in a ref-counted, LRU-evicted cache keyed by their token chain;
                               b200llm_seq_alloc_with_prefix maps a new prompt's longest cached page
                               prefix into its slot and prefill then runs over the suffix only.
                               The reference's prompts that share a head: llm_server.py:273-285
                               (summaries of a growing channel, raft_node.py:2032-2084), :221-229 */
} b200llm_config;

/* Number of fp32 values in the weight blob for `cfg` (layout below). */
size_t b200llm_weight_blob_floats(const b200llm_config* cfg);

/*
 * Weight blob: fp32, host memory, Hugging Face GPT2LMHeadModel state-dict tensors
 * concatenated in this order (Conv1D weights keep HF's [in, out] layout,
 * $TF/pytorch_utils.py:112):
 *   wte [vocab, d], wpe [n_positions, d],
 *   per layer: ln_1.weight [d], ln_1.bias [d], attn.c_attn.weight [d, 3d],
 *              attn.c_attn.bias [3d], attn.c_proj.weight [d, d], attn.c_proj.bias [d],
 *              ln_2.weight [d], ln_2.bias [d], mlp.c_fc.weight [d, 4d], mlp.c_fc.bias [4d],
 *              mlp.c_proj.weight [4d, d], mlp.c_proj.bias [d],
 *   ln_f.weight [d], ln_f.bias [d].
 * The LM head is tied to wte (modeling_gpt2.py:646-651). The engine keeps GEMM weights in
 * bf16 (transposed to [out, in]), everything else in fp32.
 */
int b200llm_engine_create(const b200llm_config* cfg, const float* weights_host, size_t n_floats,
                          b200llm_engine** out_engine);
int b200llm_engine_destroy(b200llm_engine* engine);

/* Run all engine work on `stream` (a cudaStream_t) instead of the engine's own stream. */
int b200llm_engine_set_stream(b200llm_engine* engine, void* stream);

/* ---- KV-cache manager: paged pool, 16-token pages, one slot per live sequence --------- */
/* Replaces the cat-and-reallocate cache of $TF/cache_utils.py:119-120. */
int b200llm_seq_alloc(b200llm_engine* engine, int32_t capacity_tokens, int32_t* out_slot);
int b200llm_seq_free(b200llm_engine* engine, int32_t slot);
/* seq_alloc for a prompt whose token ids are known (host memory). With cfg.prefix_cache the
 * longest run of leading full pages that an earlier prompt already filled is shared with this
 * slot (at least one prompt token always stays uncached: its logits select the first new token);
 * *out_cached_tokens says how many prompt tokens the next b200llm_prefill of this slot will skip.
 * That prefill must pass exactly these ids from HOST memory. Without cfg.prefix_cache this is
 * seq_alloc with *out_cached_tokens = 0. */
int b200llm_seq_alloc_with_prefix(b200llm_engine* engine, const int32_t* ids_host, int32_t n_ids,
                                  int32_t capacity_tokens, int32_t* out_slot,
                                  int32_t* out_cached_tokens);
/* Prefix-cache counters since engine creation: prompt tokens looked up, prompt tokens served
 * from cached pages, pages currently held by the cache. Any pointer may be NULL. */
int b200llm_prefix_cache_stats(b200llm_engine* engine, int64_t* out_lookup_tokens,
                               int64_t* out_hit_tokens, int64_t* out_cached_pages);
/* Stop generating for a slot whose caller gave up (llm_server.py's callers time out after
 * 10-20 s, raft_node.py:2018,2084,2126,2187): the row is marked finished on the device, in
 * stream order, so later decode steps emit pad for it and append nothing. Free the slot with
 * b200llm_seq_free once the steps already enqueued for it have drained (after the next
 * b200llm_read_status / b200llm_sync). */
int b200llm_seq_cancel(b200llm_engine* engine, int32_t slot);
int b200llm_kv_pages_free(b200llm_engine* engine, int64_t* out_free_pages, int64_t* out_total_pages);

/* Per-sequence generation controls (defaults: greedy, min_new 0, max_new = engine cap). */
typedef struct b200llm_sampling {
  int32_t max_new_tokens;
  int32_t min_new_tokens; /* EOS is masked until this many tokens exist (MinNewTokensLength) */
  float temperature;      /* <= 0 -> greedy argmax ($TF/generation/utils.py:2793)            */
  int32_t top_k;          /* <= 1 -> greedy; capped at 64                                     */
  uint64_t seed;
} b200llm_sampling;

/*
 * Prefill: one full forward over the packed prompts ($TF/generation/utils.py:2736,
 * GPT2Model.forward modeling_gpt2.py:522-636), filling the KV pages of each slot and
 * selecting each sequence's first new token (utils.py:2762-2805).
 *   ids            packed token ids, n_tok = cu_seqlens_host[batch]; device memory when
 *                  ids_on_device != 0, else host memory (copied on the engine stream)
 *   cu_seqlens_host [batch+1] prefix sums of prompt lengths (positions = 0..len-1,
 *                  generation/utils.py:719-721 for an all-ones mask)
 *   slots_host     [batch] slots from b200llm_seq_alloc
 *   sampling_host  [batch] or NULL (greedy, engine defaults)
 *   logits_all_dev NULL, or [n_tok, vocab] fp32 device buffer receiving the logits of every
 *                  prompt position (parity tests; GPT2LMHeadModel.forward :658-725)
 * Asynchronous: returns after enqueueing unless host ids were passed from pageable memory.
 */
int b200llm_prefill(b200llm_engine* engine, const int32_t* ids, int32_t ids_on_device,
                    const int32_t* cu_seqlens_host, const int32_t* slots_host,
                    const b200llm_sampling* sampling_host, int32_t batch, float* logits_all_dev);

/*
 * Decode: `n_steps` autoregressive steps for the given live slots (the while-loop body at
 * $TF/generation/utils.py:2743-2805: forward of the last token against the cache, fp32
 * logits, argmax / sampling, EOS -> pad bookkeeping). All state stays on the device; the
 * step is replayed from a CUDA graph when enabled. Asynchronous.
 */
int b200llm_decode(b200llm_engine* engine, const int32_t* slots_host, int32_t batch,
                   int32_t n_steps);

/* Block until the engine stream is idle. */
int b200llm_sync(b200llm_engine* engine);

/* Copy out a slot's generated tokens (synchronises the engine stream).
 *   out_tokens_host [max_tokens]; *out_count = number generated; *out_finished = 1 when the
 *   sequence hit EOS or its max_new_tokens. */
int b200llm_read_tokens(b200llm_engine* engine, int32_t slot, int32_t* out_tokens_host,
                        int32_t max_tokens, int32_t* out_count, int32_t* out_finished);

/* Batched read of (n_gen, finished) for several slots with one device->host copy. */
int b200llm_read_status(b200llm_engine* engine, const int32_t* slots_host, int32_t batch,
                        int32_t* out_counts_host, int32_t* out_finished_host);

/* Status and generated tokens of several slots with ONE stream synchronisation: a gather of
 * (n_gen, finished), then the token rows of the finished slots (or of every slot when
 * only_finished == 0) into pinned staging, then one wait. Replaces read_status + one
 * read_tokens per finished request in a serving loop.
 *   out_tokens_host [batch, out_stride]: row b holds min(n_gen, out_stride) tokens when it was
 *   copied, untouched otherwise; out_counts_host / out_finished_host [batch]. */
int b200llm_read_tokens_batch(b200llm_engine* engine, const int32_t* slots_host, int32_t batch,
                              int32_t only_finished, int32_t* out_tokens_host, int32_t out_stride,
                              int32_t* out_counts_host, int32_t* out_finished_host);

/* Last-position logits of the most recent prefill/decode call, row b = b-th slot passed.
 * [batch, vocab] fp32 to host (synchronises). Parity tests. */
int b200llm_read_last_logits(b200llm_engine* engine, int32_t batch, float* out_logits_host);

/*
 * End-to-end generation with HOST buffers: the call LocalGenerativeModel makes for a batch
 * of prompts ($TF/generation/utils.py:2131 GenerationMixin.generate, greedy/sampled).
 * Allocates slots, copies ids to the device, prefills, decodes until every sequence is
 * finished or reached max_new_tokens, copies the tokens back, frees the slots.
 *   out_tokens_host [batch, out_stride] (rows padded with pad_token_id), out_counts_host [batch]
 */
int b200llm_generate(b200llm_engine* engine, const int32_t* ids_host,
                     const int32_t* cu_seqlens_host, const b200llm_sampling* sampling_host,
                     int32_t batch, int32_t* out_tokens_host, int32_t out_stride,
                     int32_t* out_counts_host);

/* Number of kernels the engine has launched (graph replays count their captured kernels). */
int64_t b200llm_kernel_launches(const b200llm_engine* engine);

/* ---- Operator-level entry points (unit tests, profiling). Device pointers. ------------- */

/* out[token, feature] = sum_k act[token,k] * w[feature,k] (+ epilogue).  act [tokens,k] bf16,
 * w [features,k] bf16 (K-major), bias [features] fp32 or NULL. Conv1D: $TF/pytorch_utils.py:119-122.
 * orientation: -1 auto, 0 tokens on UMMA-M with single-CTA tiles, 1 weights on UMMA-M (swap-AB,
 *          small batches), 2 tokens on UMMA-M with CTA pairs (tcgen05 cta_group::2, 256x256 tiles).
 * split_k: 1 = none; n > 1 = n partial planes (B200LLM_EPI_PARTIAL_F32 only, the consumer reduces);
 *          -n = split K over a thread-block cluster of n CTAs (n <= 8) reduced on chip through
 *          distributed shared memory, with the normal bias / gelu / residual epilogue.
 *          With CTA pairs and B200LLM_EPI_BIAS_RESID_F32: n > 1 = a tile's K range cut into n work
 *          items whose products meet in the TMA reduce-add, -1 = that count by the wave cost model. */
int b200llm_op_gemm(const void* act_dev, const void* w_dev, const float* bias_dev, void* out_dev,
                    int32_t tokens, int32_t features, int32_t k, int32_t ld_out, int32_t epilogue,
                    int32_t split_k, int32_t orientation, void* stream);

/* out_bf16[row, f] = epilogue( LayerNorm(x_f32[row, :]) . w[f, :] + bias[f] ), epilogue = identity or
 * gelu_new: LayerNorm folded into the consuming GEMM's operand path (the decode step's ln_1 + c_attn
 * and ln_2 + c_fc + act; modeling_gpt2.py:273,185 / :304,239-240). rows <= 64, d % 64 == 0.
 * split > 1: K split over `split` CTAs per tile, each writing an fp32 partial plane
 * partial_out_dev[split][rows, ld_out] (no bias, no activation; out_bf16_dev unused) that the consumer
 * reduces in a fixed order -- the form decode attention takes the QKV projection in. */
int b200llm_op_ln_gemm(const float* x_dev, const float* gamma_dev, const float* beta_dev, float eps,
                       const void* w_dev, const float* bias_dev, void* out_bf16_dev, int32_t rows,
                       int32_t features, int32_t d, int32_t ld_out, int32_t gelu, int32_t split,
                       float* partial_out_dev, void* stream);
/* x_f32[row, f] += act[row, :] . w[f, :] + bias[f], split-K over `split` CTAs per 128 features whose
 * partial tiles are added into x in L2 by TMA reduce-add (the decode step's attention out-proj and
 * MLP down-proj with their residual adds, modeling_gpt2.py:223,282 / :241,307). split 0 = automatic.
 * act must be backed by 16 / 32 / 64 rows (the token tile that holds `rows`). */
int b200llm_op_gemm_redadd(const void* act_dev, const void* w_dev, const float* bias_dev, float* x_dev,
                           int32_t rows, int32_t features, int32_t k, int32_t split, void* stream);

/* out_bf16 = LayerNorm(x_f32) ; optional fused "x += bias + sum partials" first.
 * modeling_gpt2.py:252-254 (nn.LayerNorm, eps inside sqrt, biased variance). */
int b200llm_op_layernorm(float* x_dev, const float* partials_dev, int32_t n_partials,
                         const float* add_bias_dev, const float* gamma_dev, const float* beta_dev,
                         void* out_bf16_dev, int32_t rows, int32_t d, float eps, void* stream);

/* x[t] = wte[ids[t]] + wpe[pos[t]]   (modeling_gpt2.py:577-585) */
int b200llm_op_embed(const float* wte_dev, const float* wpe_dev, const int32_t* ids_dev,
                     const int32_t* pos_dev, float* x_dev, int32_t n_tok, int32_t d,
                     int32_t vocab, int32_t n_positions, void* stream);

/* Causal attention over packed prompts; qkv [n_tok, 3*d] bf16 -> out [n_tok, d] bf16.
 * modeling_gpt2.py:54-72, 185-226. */
int b200llm_op_prefill_attention(const void* qkv_dev, const int32_t* cu_seqlens_dev, int32_t batch,
                                 int32_t max_seq_len, int32_t n_head, void* out_dev, void* stream);

/* The LM head of a decode step exactly as the engine launches it: logits[token, v] =
 * act[token, :] . wte[v, :] (tied head, modeling_gpt2.py:646-651, :705-706) with the greedy
 * argmax ($TF/generation/utils.py:2793) fused into the GEMM epilogue. Every group of 32
 * vocabulary rows leaves its best (value, index) per token in amax_val/amax_idx
 * [b200llm_lm_head_argmax_groups(vocab, tokens), ld_amax]; entry eos_id is left out of the groups
 * and its logit goes to eos_logit[token] (MinNewTokensLength masks it later). Both arrays must
 * start at -inf / 0. logits_dev may be NULL (the decode step of greedy rows stores no logits).
 * Algorithmic bytes: vocab*k*2 + tokens*k*2 (+ tokens*vocab*4 with logits). tokens <= 256. */
int b200llm_op_lm_head_argmax(const void* act_dev, const void* wte_dev, int32_t tokens, int32_t vocab,
                              int32_t k, int32_t eos_id, float* logits_dev, float* amax_val_dev,
                              int32_t* amax_idx_dev, float* eos_logit_dev, int32_t ld_amax,
                              void* stream);
int32_t b200llm_lm_head_argmax_groups(int32_t vocab, int32_t tokens);

/* Causal attention of packed prompt SUFFIXES against the paged KV cache (shared-prefix reuse):
 * sequence b has prefix_len_dev[b] tokens already cached in its pages and contributes rows
 * [cu_seqlens_dev[b], cu_seqlens_dev[b+1]) of qkv, whose positions are prefix_len + 0, 1, ...
 * The suffix K/V must already be in the pages (b200llm_op_kv_scatter). slots_dev[b] selects the
 * page-table row. modeling_gpt2.py:54-72 with the past taken from the cache (:194-196).
 * This is the engine's prefill attention for EVERY prompt (prefix_len 0 for a cold one).
 * n_rows = rows of qkv, n_pages = pages in each pool (extents of the TMA tensor maps).
 * impl: 0 = tcgen05 (S = Q K^T and O = P V by tcgen05.mma with TMEM accumulators, Q / K / V tiles
 * by TMA; what the engine runs), 1 = the same computation on the legacy mma.sync path (A/B partner).
 * Pages must hold finite values behind a sequence's last token (the engine zeroes its pools). */
int b200llm_op_prefill_attention_paged(const void* qkv_dev, int32_t n_rows, const int32_t* cu_seqlens_dev,
                                       const int32_t* prefix_len_dev, const int32_t* slots_dev,
                                       int32_t batch, int32_t max_suffix_len, int32_t n_head,
                                       const int32_t* page_table_dev, int32_t max_pages_per_seq,
                                       const void* k_pool_dev, const void* v_pool_dev, int32_t n_pages,
                                       void* out_dev, int32_t impl, void* stream);
/* K and V of packed rows -> KV pages: row t of qkv goes to position row_pos_dev[t] of slot
 * row_slot_dev[t] ($TF/cache_utils.py:119-120). */
int b200llm_op_kv_scatter(const void* qkv_dev, const int32_t* row_pos_dev, const int32_t* row_slot_dev,
                          int32_t n_tok, int32_t n_head, const int32_t* page_table_dev,
                          int32_t max_pages_per_seq, void* k_pool_dev, void* v_pool_dev, void* stream);

/* One decode step of attention over paged KV (appends the new token's K/V first).
 * k_pool/v_pool: [n_pages, n_head, 16, 64] bf16; page_table [n_slots, max_pages_per_seq];
 * row b uses slot b. |n_splits| > 1 needs scratch_dev of b200llm_decode_attention_scratch_floats()
 * floats; a NEGATIVE n_splits (-2..-8) merges the splits of a (row, head) inside the kernel over a
 * thread-block cluster (the engine's choice for one or a few live rows) instead of through the
 * scratch and a combine launch. modeling_gpt2.py:54-72 with q_len = 1; cache append
 * $TF/cache_utils.py:119-120. */
int b200llm_op_decode_attention(const void* qkv_dev, const int32_t* seq_len_dev,
                                const int32_t* page_table_dev, int32_t max_pages_per_seq,
                                void* k_pool_dev, void* v_pool_dev, void* out_dev, int32_t batch,
                                int32_t n_head, int32_t n_splits, float* scratch_dev, void* stream);
size_t b200llm_decode_attention_scratch_floats(int32_t batch, int32_t n_head, int32_t n_splits);
/* The same step on the streaming form of the kernel: persistent warps that each walk whole
 * (row, head, split) units, K/V tiles of two pages fetched by TMA into a per-warp ring. The engine
 * runs it when the units outnumber the resident CTAs of the kernel above or contexts run to several
 * 128-row passes (large batches, long conversations). n_pages = pages in each pool (extent of the
 * tensor maps); n_splits >= 1, scratch as above. split_counters_dev: null, or batch * n_head int32
 * that are zero before the first call (every call leaves them at zero): the split of a (row, head)
 * that finishes last then merges them inside the kernel (in split order) and no combine kernel is
 * launched -- the engine's setting. Same results up to fp32 summation order. */
int b200llm_op_decode_attention_stream(const void* qkv_dev, const int32_t* seq_len_dev,
                                       const int32_t* page_table_dev, int32_t max_pages_per_seq,
                                       void* k_pool_dev, void* v_pool_dev, int32_t n_pages, void* out_dev,
                                       int32_t batch, int32_t n_head, int32_t n_splits, float* scratch_dev,
                                       int32_t* split_counters_dev, void* stream);

/*
 * b200llm_op_decode_attention without a scratch argument. (No counterpart in the reference: it
 * has no attention code at all; this exists only for embedders of this library.)
 *
 * Who it is for: a host that cannot obtain device memory at the moment it wants to run a
 * split-KV decode step -- for example a serving loop whose memory was handed out once at start-up
 * to its own pool so that a later cudaMalloc would fail or would stall every stream of the device
 * while the driver serves it, or a plug-in that is only given input and output buffers by its
 * caller and has nowhere to keep a workspace of its own. Such a caller still passes the inputs and
 * the output exactly as above; the partial results of the KV splits go to a fixed array inside the
 * library image instead.
 *
 * Rules
 *   - Limit: batch * n_head * n_splits * 66 floats must not exceed
 *     B200LLM_NOALLOC_SCRATCH_FLOATS (4 MiB). Larger requests return B200LLM_ERR_CAPACITY with an
 *     explanatory b200llm_last_error(); nothing is launched.
 *   - Cost: every GPU on which the library is loaded holds those 4 MiB until the process exits.
 *   - The call performs no cudaMalloc / cudaFree and uses no allocating library routine.
 *   - The first call on each device looks up the array's device address (with that device
 *     current). Make that first call before starting any stream capture on the device -- and at
 *     start-up, BEFORE the device's memory is handed to your own pool: CUDA loads a module (and
 *     reserves the memory of its global arrays, this one included) the first time something in
 *     it is used, so a first call made after all device memory is taken can fail or stall.
 *   - All calls on a device share the one array, which keeps whatever the previous call left in
 *     it. Nothing in it needs clearing (each slot is written before it is read within a call),
 *     but two calls whose kernels can run at the same time on the GPU -- different streams without
 *     an ordering between them -- race on it. Order them.
 */
#define B200LLM_NOALLOC_SCRATCH_FLOATS (1u << 20)
int b200llm_op_decode_attention_noalloc(const void* qkv_dev, const int32_t* seq_len_dev,
                                        const int32_t* page_table_dev, int32_t max_pages_per_seq,
                                        void* k_pool_dev, void* v_pool_dev, void* out_dev,
                                        int32_t batch, int32_t n_head, int32_t n_splits,
                                        void* stream);

/* Row-wise argmax of fp32 logits, ties to the lowest index ($TF/generation/utils.py:2793). */
int b200llm_op_argmax(const float* logits_dev, int32_t rows, int32_t ld, int32_t vocab,
                      int32_t* out_dev, void* stream);

/* fp32 [rows, cols] -> bf16 [cols, rows]: Conv1D [in,out] weight to the K-major [out,in] the
 * GEMM consumes. */
int b200llm_op_convert_transpose(const float* src_dev, void* dst_bf16_dev, int32_t rows,
                                 int32_t cols, void* stream);

const char* b200llm_last_error(void);
int b200llm_abi_version(void);

#ifdef __cplusplus
}
#endif

#endif /* B200LLM_H_ */
