// Generation engine: model executor + paged KV-cache manager + sampler state, exported
// through the C ABI in include/b200llm.h.
//
// It replaces the remote generator behind the reference's LLMService
// (llm_server/llm_server.py:32-33, call sites :167, :231, :287, :403) with local GPT-2
// generation whose arithmetic follows transformers' GPT2LMHeadModel.generate
// (generation/utils.py:2131, :2658-2841; models/gpt2/modeling_gpt2.py:522-725).
//
// HBM layout (one engine = one GPU):
//   weights   wte fp32 [V,d] (embedding gather) + wte bf16 [V,d] (tied LM head, K-major),
//             wpe fp32 [P,d]; per layer bf16 K-major [out,in] GEMM weights, fp32 biases and
//             LayerNorm parameters.
//   KV pool   per layer K and V: [n_pages][n_head][16 tokens][64] bf16; page table int32
//             [max_seqs][max_pages_per_seq]; the host owns the free list.
//   slots     per-sequence generation state (seq_len, last token, finished, counters,
//             sampling controls, generated tokens) so a decode step needs no host input.
//   workspace residual stream x fp32 [rows,d]; h / qkv / attn / ffn bf16; logits fp32
//             [max_batch_rows, V]; split-K partial planes; split-KV attention scratch.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/b200llm.h"
#include "attention.cuh"
#include "common.cuh"
#include "decode_fused.cuh"
#include "gemm_tcgen05.cuh"
#include "ops.cuh"

namespace b200llm {

static thread_local std::string g_last_error;

void set_last_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

namespace {

constexpr size_t kAlign = 256;
inline size_t align_up(size_t v, size_t a = kAlign) { return (v + a - 1) / a * a; }

// `cu` counts the rows that are actually prefilled; with shared-prefix reuse sequence b has
// prefix[b] more tokens in front of them that are already in its KV pages.
__global__ void prefill_meta_kernel(const int32_t* __restrict__ cu, const int32_t* __restrict__ slots,
                                    const int32_t* __restrict__ prefix, int batch,
                                    int32_t* __restrict__ row_pos, int32_t* __restrict__ row_slot,
                                    int32_t* __restrict__ last_rows, int32_t* __restrict__ seq_len) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_tok = cu[batch];
  if (t < n_tok) {
    int lo = 0, hi = batch;  // largest b with cu[b] <= t
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cu[mid] <= t) lo = mid; else hi = mid;
    }
    row_pos[t] = t - cu[lo] + (prefix ? prefix[lo] : 0);
    row_slot[t] = slots[lo];
  }
  if (t < batch) {
    last_rows[t] = cu[t + 1] - 1;
    seq_len[slots[t]] = cu[t + 1] - cu[t] + (prefix ? prefix[t] : 0);
  }
}

__global__ void cancel_slot_kernel(int32_t* finished, int slot) { finished[slot] = 1; }

// Copies the generated-token rows of the selected slots into a packed staging buffer
// [batch, stride] (one D2H copy afterwards instead of one per slot).
__global__ void gather_tokens_kernel(const int32_t* __restrict__ slots, int batch,
                                     const int32_t* __restrict__ finished, int only_finished,
                                     const int32_t* __restrict__ out_tokens, int stride,
                                     int32_t* __restrict__ packed) {
  const int b = blockIdx.x;
  if (b >= batch) return;
  const int s = slots[b];
  if (only_finished && !finished[s]) return;
  for (int t = threadIdx.x; t < stride; t += blockDim.x) {
    packed[static_cast<size_t>(b) * stride + t] = out_tokens[static_cast<size_t>(s) * stride + t];
  }
}

// Shared-prefix cache key: the physical page holding the previous 16 tokens (-1 for the first
// page of a prompt) plus this page's 16 token ids. By induction the parent page identifies the
// whole token prefix, so equal keys mean equal KV content; no hash collision can alias prompts.
struct PrefixKey {
  int32_t parent;
  int32_t tok[kPageTokens];
  bool operator==(const PrefixKey& o) const {
    return parent == o.parent && std::memcmp(tok, o.tok, sizeof(tok)) == 0;
  }
};
struct PrefixKeyHash {
  size_t operator()(const PrefixKey& k) const {
    uint64_t h = 1469598103934665603ull ^ static_cast<uint32_t>(k.parent);
    for (int i = 0; i < kPageTokens; ++i) h = (h ^ static_cast<uint32_t>(k.tok[i])) * 1099511628211ull;
    return static_cast<size_t>(h ^ (h >> 29));
  }
};
struct PrefixEntry {
  int32_t page;
  int32_t children = 0;   // cached pages whose key names this page as parent
  uint64_t last_use = 0;
};

struct SlotInit {
  int32_t slot;
  int32_t max_new;
  int32_t min_new;
  int32_t top_k;
  float temperature;
  uint32_t pad;
  uint64_t seed;
};

__global__ void init_slots_kernel(const SlotInit* __restrict__ init, int batch, int32_t* last_tok,
                                  int32_t* finished, int32_t* n_gen, int32_t* min_new,
                                  int32_t* max_new, float* temperature, int32_t* top_k,
                                  uint64_t* seed) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const SlotInit s = init[b];
  last_tok[s.slot] = 0;
  finished[s.slot] = 0;
  n_gen[s.slot] = 0;
  min_new[s.slot] = s.min_new;
  max_new[s.slot] = s.max_new;
  temperature[s.slot] = s.temperature;
  top_k[s.slot] = s.top_k;
  seed[s.slot] = s.seed;
}

__global__ void gather_status_kernel(const int32_t* __restrict__ slots, int batch,
                                     const int32_t* __restrict__ n_gen,
                                     const int32_t* __restrict__ finished, int32_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  out[b] = n_gen[slots[b]];
  out[batch + b] = finished[slots[b]];
}

struct LayerWeights {
  float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  __nv_bfloat16 *w_qkv, *w_o, *w_fc, *w_proj;
  float *b_qkv, *b_o, *b_fc, *b_proj;
};

}  // namespace

}  // namespace b200llm

using namespace b200llm;

struct b200llm_engine {
  b200llm_config cfg;
  int d, H, L, V, P;
  int max_pages_per_seq;
  int64_t n_pages;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;

  std::vector<void*> allocations;

  // weights
  float* wte_f32 = nullptr;
  __nv_bfloat16* wte_bf16 = nullptr;
  float* wpe = nullptr;
  float *lnf_g = nullptr, *lnf_b = nullptr;
  std::vector<LayerWeights> layers;

  // KV
  __nv_bfloat16* k_pool = nullptr;  // [L][n_pages][H][16][64]
  __nv_bfloat16* v_pool = nullptr;
  size_t pool_layer_stride = 0;  // elements
  int32_t* page_table = nullptr;  // device
  std::vector<int32_t> free_pages;
  std::vector<std::vector<int32_t>> slot_pages;
  std::vector<char> slot_used;
  std::vector<int32_t> slot_capacity;
  std::vector<int32_t> slot_len_host;  // host mirror of seq_len upper bound
  std::vector<int32_t> page_table_host;  // host mirror [max_seqs][max_pages_per_seq]
  std::vector<char> slot_pt_dirty;       // row not yet uploaded (uploaded in bulk at prefill)
  // shared-prefix reuse (cfg.prefix_cache): pages are reference counted -- one reference per
  // slot that maps the page plus one while the cache itself holds it
  std::vector<int32_t> page_ref;
  std::vector<int32_t> slot_prefix;               // cached leading tokens the next prefill skips
  std::vector<std::vector<int32_t>> slot_prompt;  // ids given to seq_alloc_with_prefix
  std::unordered_map<PrefixKey, PrefixEntry, PrefixKeyHash> prefix_map;
  std::vector<PrefixKey> page_key;                // key of a cached page (valid while page_cached)
  std::vector<char> page_cached;
  uint64_t prefix_clock = 0;
  int64_t prefix_lookup_tokens = 0, prefix_hit_tokens = 0, prefix_cached_pages = 0;
  int32_t* prefix_dev = nullptr;                  // [max_seqs] prefix length per sequence of a prefill
  int32_t* tok_stage = nullptr;                   // [max_seqs, max_new] packed rows for read_tokens_batch

  // slot state (device)
  int32_t *seq_len = nullptr, *last_tok = nullptr, *finished = nullptr, *n_gen = nullptr;
  int32_t *min_new = nullptr, *max_new = nullptr, *top_k = nullptr, *out_tokens = nullptr;
  float* temperature = nullptr;
  uint64_t* seed = nullptr;

  // workspace
  int max_rows = 0;
  float* x = nullptr;
  __nv_bfloat16 *h = nullptr, *qkv = nullptr, *attn = nullptr, *ffn = nullptr;
  float* logits = nullptr;
  int logits_rows = 0;
  float* amax_val = nullptr;   // [n_groups, logits_rows] LM-head argmax candidates
  int32_t* amax_idx = nullptr;
  float* eos_logit = nullptr;  // [logits_rows]
  int n_groups = 0;
  std::vector<char> slot_samples;  // host: slot uses temperature/top-k sampling
  float* partials = nullptr;
  size_t partials_floats = 0;
  float* attn_scratch = nullptr;
  int32_t* attn_counters = nullptr;
  size_t attn_scratch_floats = 0;
  int32_t *ids_dev = nullptr, *row_pos = nullptr, *row_slot = nullptr, *cu_dev = nullptr,
          *slots_dev = nullptr, *last_rows = nullptr, *active_dev = nullptr, *status_dev = nullptr;
  int32_t* live_count_dev = nullptr;  // rows of the active set that hold a sequence (they come first)
  SlotInit* init_dev = nullptr;

  // pinned staging
  int32_t* pin_i32 = nullptr;
  size_t pin_i32_count = 0;
  cudaEvent_t pin_event = nullptr;  // recorded after the last copy that read the pinned staging buffers
  bool pin_busy = false;
  SlotInit* pin_init = nullptr;
  int32_t* pin_pt = nullptr;    // pinned mirror of the page table (bulk uploads)
  int32_t* pin_tok = nullptr;   // pinned landing zone for generated tokens [max_seqs][max_new]

  // decode graphs
  struct GraphKey {
    int bucket, splits, logits, steps, attn_cluster;
    bool operator<(const GraphKey& o) const {
      if (bucket != o.bucket) return bucket < o.bucket;
      if (splits != o.splits) return splits < o.splits;
      if (logits != o.logits) return logits < o.logits;
      if (steps != o.steps) return steps < o.steps;
      return attn_cluster < o.attn_cluster;
    }
  };
  struct GraphEntry {
    cudaGraphExec_t exec = nullptr;
    int kernels = 0;
  };
  std::map<GraphKey, GraphEntry> graphs;
  std::vector<int32_t> active_host;  // last active set uploaded (padded)
  int64_t launches = 0;
  int launches_in_capture = 0;

  int last_batch = 0;
  // Profiling aid (libb200llm_tune.so only, env B200LLM_SKIP, bit mask): drop a class of kernels
  // from the decode step to read its in-graph cost by difference. Results are garbage when set.
  int skip_mask = 0;
  int sm_count = 148;
  int prefill_pdl = 1;  // B200LLM_PREFILL_PDL=0 disables
  int prefill_split = 1;  // B200LLM_PREFILL_SPLIT=0: no split-K planes for short prompts
  int steps_per_graph = 4;  // decode steps captured per CUDA graph (B200LLM_STEPS_PER_GRAPH overrides)
  int act_via_tma = 0;  // B200LLM_ACT_TMA=1: decode GEMMs fetch activations through TMA (comparison aid)
  int pair_mode = -1;  // GemmArgs::use_pair for the prefill GEMMs (env B200LLM_PAIR overrides)
  int pair_split = -1; // GemmArgs::pair_split: K splits of the residual prefill GEMMs (-1 by the wave model; 1 when
                       // cfg.decode_reduce == 1 asks for the ordered, bit-reproducible path; tune build: B200LLM_PAIR_SPLIT)
  int tune_splits[4] = {0, 0, 0, 0};  // tune build: split-K overrides for qkv, o, fc, proj
  int tune_attn_splits = 0;           // tune build: KV split override
  int tune_attn_stream = -1;          // tune build: -1 auto, 0 never, >= 1 always the streaming decode attention, instance n (B200LLM_ATTN_STREAM)
  int tune_attn_merge = 0;            // tune build: 1 = in-kernel split merge instead of the combine kernel (B200LLM_ATTN_MERGE; measured slower)
  int prefill_attn_mode = -1;         // -1 auto, 0 tcgen05 always, 1 mma.sync always (tune build: B200LLM_PREFILL_ATTN)
  int tune_fused_qkv = -1;            // tune build: 0 / 1 force ln_1 + c_attn as LayerNorm + planes / as the fused kernel
  int tune_lnqkv_split = 0;           // tune build: K split of the light ln_1 + c_attn kernel (0 = LayerNorm + planes GEMM)
  int tune_fused_splits[2] = {0, 0};  // tune build: K splits of the fused path's out-proj / MLP-down
};

namespace b200llm {
namespace {

int dev_alloc(b200llm_engine* e, void** p, size_t bytes) {
  void* q = nullptr;
  cudaError_t err = cudaMalloc(&q, align_up(std::max<size_t>(bytes, 1)));
  if (err != cudaSuccess) {
    return fail(kErrOutOfMemory, std::string("cudaMalloc failed: ") + cudaGetErrorString(err));
  }
  e->allocations.push_back(q);
  *p = q;
  return kOk;
}

template <typename T>
int dev_alloc_t(b200llm_engine* e, T** p, size_t count) {
  return dev_alloc(e, reinterpret_cast<void**>(p), count * sizeof(T));
}

int bucket_for(int batch) {
  if (batch <= 4) return 4;  // one or a few live requests: fewer padding CTAs in every kernel of the step
  if (batch <= 16) return 16;
  if (batch <= 32) return 32;
  if (batch <= 64) return 64;
  if (batch <= 128) return 128;
  return (batch + 255) / 256 * 256;
}

// Cluster split: largest divisor s <= 8 of k_blocks with tiles * s <= sms.
int pick_cluster(int features, int k, int sms) {
  const int tiles = (features + 127) / 128;
  const int kb = k / 64;
  int best = 1;
  for (int s = 1; s <= 8; ++s) {
    if (kb % s == 0 && tiles * s <= sms) best = s;
  }
  return best;
}

// Largest divisor s of k_blocks with tiles * s <= sms and s <= cap.
// Split-K factor of a decode GEMM. Every plane costs the consuming kernel (LayerNorm, attention,
// gelu-reduce) one more fp32 read per element on the step's critical path, while the weight
// stream of a CTA is prefetched before its grid dependency resolves; so a CTA keeps at least
// max(3, kb/8) k-blocks (sweeps in profiles/README.md), and the split never exceeds `cap` planes
// or the SM count.
int pick_split(int features, int k, int sms, int cap) {
  const int tiles = (features + 127) / 128;
  const int kb = k / 64;
  const int min_kb = kb / 8 > 3 ? kb / 8 : 3;
  int best = 1;
  for (int s = 1; s <= kb && s <= cap; ++s) {
    if (kb % s == 0 && tiles * s <= sms && kb / s >= min_kb) best = s;
  }
  return best;
}

struct StepPlan {
  int rows;           // activation rows fed to the GEMMs
  bool decode;        // decode step (slot-driven) or prefill
  int split_qkv = 1;  // split-K of c_attn (partials reduced inside decode attention)
  int split_fc = 1;   // split-K of the MLP up-proj (partials reduced by reduce_bias_gelu)
  int split_o = 1;    // split-K of the attention out-proj (partials reduced by the next LayerNorm)
  int split_p = 1;    // split-K of the MLP down-proj
  // cluster (DSMEM) split-K factors; when > 1 they take precedence over the plane splits
  int cl_qkv = 1, cl_o = 1, cl_fc = 1, cl_p = 1;
  int attn_splits = 1;
  bool attn_stream = false;  // decode attention on the streaming kernel (many units per SM / long contexts)
  int attn_stream_instance = 1;
  int live_rows = 1 << 30;  // decode: sequences in the call (the bucket's other rows are padding)
  int pdl = 0;
  int batch = 0;       // sequences (prefill) / bucket rows (decode)
  int max_len = 0;     // prefill: longest prompt
  float* logits_all = nullptr;
  bool store_logits = true;  // materialise last-position logits (sampling rows / parity reads)
  // decode, <= 64 rows: LayerNorm folded into the QKV / MLP-up GEMMs, out-proj / MLP-down add
  // their K splits into x in L2 (decode_fused.cu): 5 kernels per layer instead of 8
  bool fused = false;
  bool fused_qkv = true;   // false: ln_1 and c_attn stay LayerNorm + split-K planes (light CTAs: the
                           // attention kernel of the same layer can start beside them and prefetch its KV)
  int lnqkv_split = 0;     // > 1 (with fused_qkv false): ln_1 + c_attn as ONE K-split kernel writing planes
  int rs_o = 1, rs_p = 1;  // K splits of the two reduce-add GEMMs
  int max_keys = 0;            // prefill: most keys (cached prefix + own rows) any sequence of the call has
  bool paged_prefill = false;  // some prefill rows are prompt suffixes behind a cached prefix (prefix_dev is valid)
};

enum SkipBits { kSkipLn = 1, kSkipGemm = 2, kSkipAttn = 4, kSkipHead = 8, kSkipSelect = 16, kSkipGelu = 32,
                kSkipLnGemm = 64, kSkipRedAdd = 128 };  // the two fused-path kernel classes

int count(b200llm_engine* e, int rc) {
  if (rc == kOk) {
    e->launches += 1;
    e->launches_in_capture += 1;
  }
  return rc;
}

int gemm(b200llm_engine* e, const __nv_bfloat16* act, const __nv_bfloat16* w, const float* bias,
         void* out, int tokens, int features, int k, int ld_out, int epi, int split, int pdl,
         int cluster_split = 0) {
  GemmArgs g = {};
  g.cluster_split = cluster_split;
  g.use_pair = e->pair_mode;
  g.pair_split = e->pair_split;
  g.act = act;
  g.w = w;
  g.bias = bias;
  g.out = out;
  g.tokens = tokens;
  g.features = features;
  g.k = k;
  g.ld_out = ld_out;
  g.epilogue = epi;
  g.split_k = split;
  g.act_rows_alloc = e->max_rows;
  g.force_orientation = -1;
  g.pdl = pdl;
  g.act_via_tma = e->act_via_tma;
  if (e->skip_mask & kSkipGemm) return kOk;
  return count(e, gemm_bf16(g, e->stream));
}

int ln(b200llm_engine* e, const float* partials, int n_partials, int partial_rows,
       const float* add_bias, const float* gamma, const float* beta, const int32_t* row_index,
       int rows, int pdl) {
  LayerNormArgs a = {};
  a.x = e->x;
  a.partials = partials;
  a.n_partials = n_partials;
  a.partial_rows = partial_rows;
  a.add_bias = add_bias;
  a.gamma = gamma;
  a.beta = beta;
  a.row_index = row_index;
  a.out = e->h;
  a.rows = rows;
  a.d = e->d;
  a.eps = e->cfg.ln_eps;
  a.pdl = pdl;
  if (e->skip_mask & kSkipLn) return kOk;
  return count(e, layernorm(a, e->stream));
}

// The transformer stack on `rows` activation rows already embedded in e->x, followed by the
// final LayerNorm, the LM head and token selection.
int run_stack(b200llm_engine* e, const StepPlan& p) {
  const int d = e->d, rows = p.rows;
  const int pdl = p.pdl;
  const float* pending_partials = nullptr;  // split-K planes not yet folded into x
  int pending_n = 0;
  const float* pending_bias = nullptr;

  for (int l = 0; l < e->L; ++l) {
    const LayerWeights& w = e->layers[l];
    __nv_bfloat16* kp = e->k_pool + static_cast<size_t>(l) * e->pool_layer_stride;
    __nv_bfloat16* vp = e->v_pool + static_cast<size_t>(l) * e->pool_layer_stride;

    if (p.decode && p.fused) {
      const bool skip_gemm = (e->skip_mask & kSkipGemm) != 0;
      // ln_1 + c_attn (modeling_gpt2.py:273, :185) -> bf16 qkv
      LnGemmArgs lg;
      lg.x = e->x;
      lg.gamma = w.ln1_g;
      lg.beta = w.ln1_b;
      lg.eps = e->cfg.ln_eps;
      lg.w = w.w_qkv;
      lg.bias = w.b_qkv;
      lg.out = e->qkv;
      lg.rows = rows;
      lg.d = d;
      lg.features = 3 * d;
      lg.ld_out = 3 * d;
      lg.gelu = 0;
      lg.pdl = pdl;
      const bool skip_lg = skip_gemm || (e->skip_mask & kSkipLnGemm) != 0;
      const bool skip_ra = skip_gemm || (e->skip_mask & kSkipRedAdd) != 0;
      const bool qkv_planes = !p.fused_qkv && (p.lnqkv_split > 1 || p.split_qkv > 1);
      const int qkv_plane_count = p.lnqkv_split > 1 ? p.lnqkv_split : p.split_qkv;
      const bool lnqkv_planes = !p.fused_qkv && p.lnqkv_split > 1;
      if (p.fused_qkv) {
        if (!skip_lg) B200_PROPAGATE(count(e, ln_gemm(lg, e->stream)));
      } else if (lnqkv_planes) {
        // ln_1 inside a K-split c_attn with light CTAs (half of the weight tile each): the attention
        // CTAs of the layer start beside them; it reduces the partial planes as on the ordered path
        LnGemmArgs ls = lg;
        ls.bias = nullptr;
        ls.out = nullptr;
        ls.split = p.lnqkv_split;
        ls.partial_out = e->partials;
        ls.plane_rows = rows;
        if (!skip_lg) B200_PROPAGATE(count(e, ln_gemm(ls, e->stream)));
      } else {
        B200_PROPAGATE(ln(e, nullptr, 0, rows, nullptr, w.ln1_g, w.ln1_b, nullptr, rows, pdl));
        if (p.split_qkv > 1) {
          B200_PROPAGATE(gemm(e, e->h, w.w_qkv, nullptr, e->partials, rows, 3 * d, d, 3 * d, kEpiPartialF32,
                              p.split_qkv, pdl));
        } else {
          B200_PROPAGATE(gemm(e, e->h, w.w_qkv, w.b_qkv, e->qkv, rows, 3 * d, d, 3 * d, kEpiBiasBf16, 1, pdl));
        }
      }
      // attention (:186-222) over the paged cache
      DecodeAttnArgs a = {};
      a.qkv = qkv_planes ? nullptr : e->qkv;
      a.qkv_partials = qkv_planes ? e->partials : nullptr;
      a.n_qkv_partials = qkv_planes ? qkv_plane_count : 0;
      a.partial_rows = rows;
      a.qkv_bias = w.b_qkv;
      a.active_slots = e->active_dev;
      a.seq_len = e->seq_len;
      a.finished = e->finished;
      a.page_table = e->page_table;
      a.max_pages_per_seq = e->max_pages_per_seq;
      a.k_pool = kp;
      a.v_pool = vp;
      a.out = e->attn;
      a.batch = rows;
      a.n_head = e->H;
      a.n_splits = p.attn_splits;
      a.scratch = e->attn_scratch;
      a.pdl = pdl;
      a.cluster_combine = (!p.attn_stream && p.live_rows * e->H * p.attn_splits <= 2 * e->sm_count) ? 1 : 0;
      a.stream_kernel = p.attn_stream ? p.attn_stream_instance : 0;
      a.n_pages = static_cast<int>(e->n_pages);
      a.n_live = e->live_count_dev;
      a.split_counters = (p.attn_stream && e->tune_attn_merge) ? e->attn_counters : nullptr;
      if (!(e->skip_mask & kSkipAttn)) B200_PROPAGATE(count(e, decode_attention(a, e->stream)));
      if (p.attn_splits > 1 && a.split_counters == nullptr && !(a.cluster_combine && p.attn_splits <= 8)) {
        e->launches += 1, e->launches_in_capture += 1;  // separate combine kernel
      }
      // attn c_proj + residual (:223, :282): x += attn . W_o + b_o, K splits added in L2
      GemmRedAddArgs ra;
      ra.act = e->attn;
      ra.w = w.w_o;
      ra.bias = w.b_o;
      ra.x = e->x;
      ra.rows = rows;
      ra.features = d;
      ra.k = d;
      ra.split = p.rs_o;
      ra.act_rows_alloc = e->max_rows;
      ra.pdl = pdl;
      if (!skip_ra) B200_PROPAGATE(count(e, gemm_redadd(ra, e->stream)));
      // ln_2 + c_fc + gelu_new (:304, :239-240) -> bf16 ffn
      lg.gamma = w.ln2_g;
      lg.beta = w.ln2_b;
      lg.w = w.w_fc;
      lg.bias = w.b_fc;
      lg.out = e->ffn;
      lg.features = 4 * d;
      lg.ld_out = 4 * d;
      lg.gelu = 1;
      if (!skip_lg) B200_PROPAGATE(count(e, ln_gemm(lg, e->stream)));
      // mlp c_proj + residual (:241, :307)
      ra.act = e->ffn;
      ra.w = w.w_proj;
      ra.bias = w.b_proj;
      ra.k = 4 * d;
      ra.split = p.rs_p;
      if (!skip_ra) B200_PROPAGATE(count(e, gemm_redadd(ra, e->stream)));
      continue;
    }
    // ln_1 (modeling_gpt2.py:273), folding the previous layer's MLP partials into x first
    B200_PROPAGATE(ln(e, pending_partials, pending_n, rows, pending_bias, w.ln1_g, w.ln1_b,
                      nullptr, rows, pdl));
    pending_partials = nullptr;
    pending_n = 0;
    // c_attn (:185)
    const bool qkv_split = p.decode && p.split_qkv > 1 && p.cl_qkv <= 1;
    if (p.decode && p.cl_qkv > 1) {
      B200_PROPAGATE(gemm(e, e->h, w.w_qkv, w.b_qkv, e->qkv, rows, 3 * d, d, 3 * d, kEpiBiasBf16, 1, pdl, p.cl_qkv));
    } else if (qkv_split) {
      B200_PROPAGATE(gemm(e, e->h, w.w_qkv, nullptr, e->partials, rows, 3 * d, d, 3 * d,
                          kEpiPartialF32, p.split_qkv, pdl));
    } else {
      B200_PROPAGATE(gemm(e, e->h, w.w_qkv, w.b_qkv, e->qkv, rows, 3 * d, d, 3 * d, kEpiBiasBf16, 1, pdl));
    }
    // attention (:186-222)
    if (p.decode) {
      DecodeAttnArgs a = {};
      a.qkv = qkv_split ? nullptr : e->qkv;
      a.qkv_partials = qkv_split ? e->partials : nullptr;
      a.n_qkv_partials = qkv_split ? p.split_qkv : 0;
      a.partial_rows = rows;
      a.qkv_bias = w.b_qkv;
      a.active_slots = e->active_dev;
      a.seq_len = e->seq_len;
      a.finished = e->finished;
      a.page_table = e->page_table;
      a.max_pages_per_seq = e->max_pages_per_seq;
      a.k_pool = kp;
      a.v_pool = vp;
      a.out = e->attn;
      a.batch = rows;
      a.n_head = e->H;
      a.n_splits = p.attn_splits;
      a.scratch = e->attn_scratch;
      a.pdl = pdl;
      // small grids (a handful of rows): 2..8 splits are merged inside the kernel, no combine launch.
      // With a thousand CTAs the co-scheduling of whole clusters costs more than the launch it saves
      // (GPT-2 medium, 16 rows x 16 heads x 4 splits: 1148 vs 1014 us per step).
      a.cluster_combine = (!p.attn_stream && p.live_rows * e->H * p.attn_splits <= 2 * e->sm_count) ? 1 : 0;
      a.stream_kernel = p.attn_stream ? p.attn_stream_instance : 0;
      a.n_pages = static_cast<int>(e->n_pages);
      a.n_live = e->live_count_dev;
      a.split_counters = (p.attn_stream && e->tune_attn_merge) ? e->attn_counters : nullptr;
      if (!(e->skip_mask & kSkipAttn)) B200_PROPAGATE(count(e, decode_attention(a, e->stream)));
      if (p.attn_splits > 1 && a.split_counters == nullptr && !(a.cluster_combine && p.attn_splits <= 8)) {
        e->launches += 1, e->launches_in_capture += 1;  // separate combine kernel
      }
    } else {
      // Prefill: every query attends over its sequence's cached prefix (shared-prefix reuse; read
      // from the pages) followed by the rows of this call (read from qkv); the CTA on the diagonal
      // writes its K/V tile into the pages, so there is no scatter pass.
      // tcgen05 for every prompt longer than 64 tokens (a light single-tile instance for <= 128 keys,
      // the pipelined one beyond; 64 x 128 tokens: 2.18 vs 2.20 ms per prefill against the mma.sync
      // kernel, 16 x 1024: 12.3 vs 13.7 ms). Prompts of at most 64 tokens would fill less than half
      // of a 128-row tcgen05 tile and stay on the 64-row mma.sync kernel (32 x 64 tokens, GPT-2 large:
      // 6.02 vs 6.08 ms).
      const bool tc = e->prefill_attn_mode == 0 || (e->prefill_attn_mode < 0 && p.max_keys > 64);
      if (!tc && !p.paged_prefill) {
        PrefillCache pc;
        pc.slots = e->slots_dev;
        pc.page_table = e->page_table;
        pc.max_pages_per_seq = e->max_pages_per_seq;
        pc.k_pool = kp;
        pc.v_pool = vp;
        B200_PROPAGATE(count(e, prefill_attention(e->qkv, e->cu_dev, p.batch, p.max_len, e->H, e->attn, e->stream, &pc)));
      } else {
      if (!tc) {  // short prompts behind a cached prefix: rows to the pages first, then every key from the pages
        B200_PROPAGATE(count(e, kv_scatter(e->qkv, e->row_pos, e->row_slot, rows, e->H, e->page_table,
                                           e->max_pages_per_seq, kp, vp, e->stream)));
      }
      PagedPrefillArgs pa;
      pa.qkv = e->qkv;
      pa.cu_seqlens = e->cu_dev;
      pa.prefix_len = p.paged_prefill ? e->prefix_dev : nullptr;
      pa.slots = e->slots_dev;
      pa.batch = p.batch;
      pa.max_suffix_len = p.max_len;
      pa.n_head = e->H;
      pa.page_table = e->page_table;
      pa.max_pages_per_seq = e->max_pages_per_seq;
      pa.k_pool = kp;
      pa.v_pool = vp;
      pa.out = e->attn;
      pa.qkv_rows = e->max_rows;
      pa.n_pages = static_cast<int>(e->n_pages);
      if (tc) {
        B200_PROPAGATE(count(e, prefill_attention_tc(pa, e->stream)));
      } else {
        B200_PROPAGATE(count(e, prefill_attention_paged(pa, e->stream)));
      }
      }
    }
    // attn c_proj + residual (:223, :282)
    if (p.decode && p.cl_o > 1) {
      B200_PROPAGATE(gemm(e, e->attn, w.w_o, w.b_o, e->x, rows, d, d, d, kEpiBiasResidF32, 1, pdl, p.cl_o));
      B200_PROPAGATE(ln(e, nullptr, 0, rows, nullptr, w.ln2_g, w.ln2_b, nullptr, rows, pdl));
    } else if (p.split_o > 1) {
      B200_PROPAGATE(gemm(e, e->attn, w.w_o, nullptr, e->partials, rows, d, d, d, kEpiPartialF32,
                          p.split_o, pdl));
      B200_PROPAGATE(ln(e, e->partials, p.split_o, rows, w.b_o, w.ln2_g, w.ln2_b, nullptr, rows, pdl));
    } else {
      B200_PROPAGATE(gemm(e, e->attn, w.w_o, w.b_o, e->x, rows, d, d, d, kEpiBiasResidF32, 1, pdl));
      B200_PROPAGATE(ln(e, nullptr, 0, rows, nullptr, w.ln2_g, w.ln2_b, nullptr, rows, pdl));
    }
    // mlp c_fc + gelu_new (:239-240)
    if (p.decode && p.cl_fc > 1) {
      B200_PROPAGATE(gemm(e, e->h, w.w_fc, w.b_fc, e->ffn, rows, 4 * d, d, 4 * d, kEpiBiasGeluBf16, 1, pdl, p.cl_fc));
    } else if (p.decode && p.split_fc > 1) {
      B200_PROPAGATE(gemm(e, e->h, w.w_fc, nullptr, e->partials, rows, 4 * d, d, 4 * d,
                          kEpiPartialF32, p.split_fc, pdl));
      if (!(e->skip_mask & kSkipGelu)) {
        B200_PROPAGATE(count(e, reduce_bias_gelu(e->partials, p.split_fc, rows, w.b_fc, e->ffn, rows,
                                                 4 * d, pdl, e->stream)));
      }
    } else {
      B200_PROPAGATE(gemm(e, e->h, w.w_fc, w.b_fc, e->ffn, rows, 4 * d, d, 4 * d, kEpiBiasGeluBf16, 1, pdl));
    }
    // mlp c_proj + residual (:241, :307)
    if (p.decode && p.cl_p > 1) {
      B200_PROPAGATE(gemm(e, e->ffn, w.w_proj, w.b_proj, e->x, rows, d, 4 * d, d, kEpiBiasResidF32, 1, pdl, p.cl_p));
    } else if (p.split_p > 1) {
      B200_PROPAGATE(gemm(e, e->ffn, w.w_proj, nullptr, e->partials, rows, d, 4 * d, d,
                          kEpiPartialF32, p.split_p, pdl));
      pending_partials = e->partials;
      pending_n = p.split_p;
      pending_bias = w.b_proj;
    } else {
      B200_PROPAGATE(gemm(e, e->ffn, w.w_proj, w.b_proj, e->x, rows, d, 4 * d, d, kEpiBiasResidF32, 1, pdl));
    }
  }

  // ln_f (:628) and the tied LM head (:705-706)
  if (p.logits_all != nullptr) {
    B200_PROPAGATE(ln(e, pending_partials, pending_n, rows, pending_bias, e->lnf_g, e->lnf_b,
                      nullptr, rows, pdl));
    B200_PROPAGATE(gemm(e, e->h, e->wte_bf16, nullptr, p.logits_all, rows, e->V, d, e->V, kEpiF32, 1, pdl));
    pending_partials = nullptr;
    pending_n = 0;
  }
  const int head_rows = p.decode ? rows : p.batch;
  B200_PROPAGATE(ln(e, pending_partials, pending_n, rows, pending_bias, e->lnf_g, e->lnf_b,
                    p.decode ? nullptr : e->last_rows, head_rows, pdl));
  {
    // LM head with the greedy argmax fused into the epilogue; logits stored only on request
    GemmArgs g = {};
    g.act = e->h;
    g.w = e->wte_bf16;
    g.bias = nullptr;
    g.out = p.store_logits ? e->logits : nullptr;
    g.tokens = head_rows;
    g.features = e->V;
    g.k = d;
    g.ld_out = e->V;
    g.epilogue = kEpiLogitsArgmax;
    g.split_k = 1;
    g.act_rows_alloc = e->max_rows;
    g.force_orientation = 1;
    g.pdl = pdl;
    g.amax_groups_alloc = e->n_groups;
    g.amax_val = e->amax_val;
    g.amax_idx = e->amax_idx;
    g.eos_logit = e->eos_logit;
    g.ld_amax = e->logits_rows;
    g.eos_id = e->cfg.eos_token_id;
    if (!(e->skip_mask & kSkipHead)) B200_PROPAGATE(count(e, gemm_bf16(g, e->stream)));
  }

  SelectArgs s = {};
  s.logits = p.store_logits ? e->logits : nullptr;
  s.ld = e->V;
  s.vocab = e->V;
  s.amax_val = e->amax_val;
  s.amax_idx = e->amax_idx;
  s.eos_logit = e->eos_logit;
  s.n_groups = std::min(e->n_groups, gemm_argmax_groups(e->V, head_rows));
  s.ld_amax = e->logits_rows;
  s.active_slots = p.decode ? e->active_dev : e->slots_dev;
  s.batch = head_rows;
  s.last_tok = e->last_tok;
  s.seq_len = e->seq_len;
  s.finished = e->finished;
  s.n_gen = e->n_gen;
  s.min_new = e->min_new;
  s.max_new = e->max_new;
  s.out_tokens = e->out_tokens;
  s.out_stride = e->cfg.max_new_tokens;
  s.eos_id = e->cfg.eos_token_id;
  s.pad_id = e->cfg.pad_token_id;
  s.temperature = e->temperature;
  s.top_k = e->top_k;
  s.seed = e->seed;
  s.advance_seq = p.decode ? 1 : 0;
  s.pdl = pdl;
  if (!(e->skip_mask & kSkipSelect)) B200_PROPAGATE(count(e, select_tokens(s, e->stream)));
  return kOk;
}

int decode_step_launches(b200llm_engine* e, const StepPlan& p) {
  B200_PROPAGATE(count(e, embed_decode(e->wte_f32, e->wpe, e->active_dev, e->last_tok, e->seq_len,
                                       e->x, p.rows, e->d, e->V, e->P, e->stream, p.pdl)));
  return run_stack(e, p);
}

// The pinned staging buffers are reused call to call. Instead of draining the whole stream before
// a reuse (which would make every prefill / active-set upload wait for the decode burst queued
// ahead of it), an event marks the last copy that read them; only that copy has to be done.
int wait_pinned_free(b200llm_engine* e) {
  if (e->pin_busy) {
    B200_CUDA_OK(cudaEventSynchronize(e->pin_event));
    e->pin_busy = false;
  }
  return kOk;
}
int mark_pinned_used(b200llm_engine* e) {
  if (e->pin_event == nullptr) B200_CUDA_OK(cudaEventCreateWithFlags(&e->pin_event, cudaEventDisableTiming));
  B200_CUDA_OK(cudaEventRecord(e->pin_event, e->stream));
  e->pin_busy = true;
  return kOk;
}

// ---- shared-prefix cache (cfg.prefix_cache) ---------------------------------------------------
// Make room: drop least-recently-used cached pages that nothing else references and that no other
// cached page hangs off (leaves first, so a chain is never cut in the middle). One scan of the
// cache serves a whole batch of evictions -- at least `want` pages and a few percent of the pool on
// top -- so a pool that is permanently full of cached prompts does not pay a scan per page.
int evict_cached_pages(b200llm_engine* e, int want) {
  int freed = 0;
  const int goal = std::max(want, static_cast<int>(e->n_pages / 16));
  while (freed < want) {
    std::vector<std::pair<uint64_t, int32_t>> leaves;  // (last_use, page)
    for (const auto& kv : e->prefix_map) {
      const PrefixEntry& en = kv.second;
      if (en.children == 0 && e->page_ref[en.page] == 1) leaves.emplace_back(en.last_use, en.page);
    }
    if (leaves.empty()) break;
    std::sort(leaves.begin(), leaves.end());
    for (const auto& lf : leaves) {
      if (freed >= goal) break;
      const int32_t page = lf.second;
      const PrefixKey key = e->page_key[page];
      if (key.parent >= 0 && e->page_cached[key.parent]) {
        auto it = e->prefix_map.find(e->page_key[key.parent]);
        if (it != e->prefix_map.end()) it->second.children -= 1;
      }
      e->prefix_map.erase(key);
      e->page_cached[page] = 0;
      e->prefix_cached_pages -= 1;
      if (--e->page_ref[page] == 0) e->free_pages.push_back(page);
      ++freed;
    }
    // parents of the evicted leaves may have become leaves themselves: next round
  }
  return freed;
}

int alloc_slot(b200llm_engine* e, int capacity_tokens, const int32_t* ids, int n_ids, int32_t* out_slot,
               int32_t* out_cached) {
  B200_CHECK_ARG(capacity_tokens > 0 && capacity_tokens <= e->cfg.max_seq_len,
                 "seq_alloc: capacity exceeds max_seq_len");
  int slot = -1;
  for (int s = 0; s < e->cfg.max_seqs; ++s) {
    if (!e->slot_used[s]) {
      slot = s;
      break;
    }
  }
  if (slot < 0) return fail(kErrCapacity, "seq_alloc: no free sequence slot");
  const int need = (capacity_tokens + kPageTokens - 1) / kPageTokens;
  std::vector<int32_t>& pages = e->slot_pages[slot];
  pages.clear();
  // longest cached run of leading full pages; the last prompt token is never served from the
  // cache (its logits select the first new token)
  int shared = 0;
  if (e->cfg.prefix_cache && ids != nullptr) {
    const int max_shared = (n_ids - 1) / kPageTokens;
    PrefixKey key;
    key.parent = -1;
    for (int i = 0; i < max_shared; ++i) {
      std::memcpy(key.tok, ids + static_cast<size_t>(i) * kPageTokens, sizeof(key.tok));
      auto it = e->prefix_map.find(key);
      if (it == e->prefix_map.end()) break;
      it->second.last_use = ++e->prefix_clock;
      pages.push_back(it->second.page);
      e->page_ref[it->second.page] += 1;  // pinned from here on: eviction below cannot take it
      key.parent = it->second.page;
      ++shared;
    }
    e->prefix_lookup_tokens += n_ids;
  }
  const int fresh = need - shared;
  if (static_cast<int64_t>(e->free_pages.size()) < fresh) {
    if (e->cfg.prefix_cache) evict_cached_pages(e, fresh - static_cast<int>(e->free_pages.size()));
    if (static_cast<int64_t>(e->free_pages.size()) < fresh) {
      for (int32_t p : pages) e->page_ref[p] -= 1;  // cached pages keep the cache's own reference
      pages.clear();
      return fail(kErrCapacity, "seq_alloc: KV pool exhausted");
    }
  }
  for (int i = 0; i < fresh; ++i) {
    const int32_t p = e->free_pages.back();
    e->free_pages.pop_back();
    e->page_ref[p] = 1;
    pages.push_back(p);
  }
  // The row goes to the device in bulk with the next prefill that uses the slot (one copy per
  // run of consecutive slots instead of one per sequence).
  std::copy(pages.begin(), pages.end(),
            e->page_table_host.begin() + static_cast<size_t>(slot) * e->max_pages_per_seq);
  e->slot_pt_dirty[slot] = 1;
  e->slot_used[slot] = 1;
  e->slot_capacity[slot] = capacity_tokens;
  e->slot_len_host[slot] = 0;
  e->slot_prefix[slot] = shared * kPageTokens;
  e->prefix_hit_tokens += shared * kPageTokens;
  if (e->cfg.prefix_cache && ids != nullptr) {
    e->slot_prompt[slot].assign(ids, ids + n_ids);
  } else {
    e->slot_prompt[slot].clear();
  }
  *out_slot = slot;
  if (out_cached) *out_cached = shared * kPageTokens;
  return kOk;
}

// After a prefill has been enqueued: the slot's pages that the prompt fills completely join the
// cache (the stream orders any later prefill that maps them after the one that writes them).
void register_prompt_pages(b200llm_engine* e, int slot, const int32_t* ids, int n_ids) {
  const int full = n_ids / kPageTokens;
  const std::vector<int32_t>& pages = e->slot_pages[slot];
  PrefixKey key;
  key.parent = -1;
  for (int i = 0; i < full && i < static_cast<int>(pages.size()); ++i) {
    std::memcpy(key.tok, ids + static_cast<size_t>(i) * kPageTokens, sizeof(key.tok));
    auto it = e->prefix_map.find(key);
    if (it == e->prefix_map.end()) {
      const int32_t page = pages[i];
      if (e->page_cached[page]) break;  // already cached under another chain (cannot happen: keys are exact)
      PrefixEntry en;
      en.page = page;
      en.last_use = ++e->prefix_clock;
      e->prefix_map.emplace(key, en);
      e->page_key[page] = key;
      e->page_cached[page] = 1;
      e->page_ref[page] += 1;
      e->prefix_cached_pages += 1;
      if (key.parent >= 0 && e->page_cached[key.parent]) {
        auto pit = e->prefix_map.find(e->page_key[key.parent]);
        if (pit != e->prefix_map.end()) pit->second.children += 1;
      }
      key.parent = page;
    } else {
      it->second.last_use = ++e->prefix_clock;
      key.parent = it->second.page;  // same tokens already cached (possibly in another physical page)
    }
  }
}

int ensure_pinned(b200llm_engine* e, size_t count_i32) {
  if (count_i32 <= e->pin_i32_count) return kOk;
  if (e->pin_i32) cudaFreeHost(e->pin_i32);
  e->pin_i32 = nullptr;
  const size_t n = std::max<size_t>(count_i32, 1 << 16);
  B200_CUDA_OK(cudaMallocHost(reinterpret_cast<void**>(&e->pin_i32), n * sizeof(int32_t)));
  e->pin_i32_count = n;
  return kOk;
}

}  // namespace
}  // namespace b200llm

#ifdef B200_STEP_TRACE
namespace b200llm {
void step_trace_set_gemm(unsigned long long*, unsigned int*, unsigned int);
void step_trace_set_ops(unsigned long long*, unsigned int*, unsigned int);
void step_trace_set_attention(unsigned long long*, unsigned int*, unsigned int);
void step_trace_set_fused(unsigned long long*, unsigned int*, unsigned int);
}  // namespace b200llm
#endif

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" {

const char* b200llm_last_error(void) { return g_last_error.c_str(); }
int b200llm_abi_version(void) { return B200LLM_ABI_VERSION; }

size_t b200llm_weight_blob_floats(const b200llm_config* c) {
  if (c == nullptr || c->struct_size != sizeof(b200llm_config)) return 0;
  const size_t d = c->d_model, V = c->vocab_size, P = c->n_positions, L = c->n_layer;
  const size_t per_layer = 2 * d + d * 3 * d + 3 * d + d * d + d + 2 * d + d * 4 * d + 4 * d + 4 * d * d + d;
  return V * d + P * d + L * per_layer + 2 * d;
}

int b200llm_engine_create(const b200llm_config* cfg, const float* weights_host, size_t n_floats,
                          b200llm_engine** out_engine) {
  B200_CHECK_ARG(cfg && weights_host && out_engine, "engine_create: null argument");
  B200_CHECK_ARG(cfg->struct_size == sizeof(b200llm_config),
                 "engine_create: b200llm_config.struct_size does not match this library's struct (ABI " +
                     std::to_string(B200LLM_ABI_VERSION) + ", " + std::to_string(sizeof(b200llm_config)) +
                     " bytes): the binding was written against another b200llm.h");
  B200_CHECK_ARG(cfg->n_head > 0 && cfg->d_model == cfg->n_head * kHeadDim,
                 "engine_create: head_dim must be 64 (d_model == 64 * n_head)");
  B200_CHECK_ARG(cfg->d_model % 128 == 0, "engine_create: d_model must be a multiple of 128");
  B200_CHECK_ARG(cfg->n_layer > 0 && cfg->vocab_size > 0 && cfg->n_positions > 0,
                 "engine_create: bad model shape");
  B200_CHECK_ARG(cfg->max_seqs > 0 && cfg->max_seq_len > 0 && cfg->max_batch_tokens > 0 &&
                     cfg->max_new_tokens > 0,
                 "engine_create: capacities must be positive");
  B200_CHECK_ARG(cfg->max_seq_len <= cfg->n_positions,
                 "engine_create: max_seq_len exceeds n_positions (no position embedding beyond it)");
  B200_CHECK_ARG(n_floats == b200llm_weight_blob_floats(cfg),
                 "engine_create: weight blob size does not match the config");
  B200_CUDA_OK(cudaSetDevice(cfg->device));

  std::unique_ptr<b200llm_engine> e(new b200llm_engine());
  e->cfg = *cfg;
  cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, e->cfg.device);
  if (e->sm_count <= 0) e->sm_count = 148;
#ifdef B200_TUNE_KNOBS
  // Tuning / ablation switches exist only in libb200llm_tune.so (build.py --tune); the product
  // library reads no environment variable.
  if (const char* skip = std::getenv("B200LLM_SKIP")) e->skip_mask = std::atoi(skip);
  if (const char* pp = std::getenv("B200LLM_PREFILL_PDL")) e->prefill_pdl = std::atoi(pp);
  if (const char* ps = std::getenv("B200LLM_PREFILL_SPLIT")) e->prefill_split = std::atoi(ps);
  if (const char* at = std::getenv("B200LLM_ACT_TMA")) e->act_via_tma = std::atoi(at);
  if (const char* sg = std::getenv("B200LLM_STEPS_PER_GRAPH")) e->steps_per_graph = std::max(1, std::min(32, std::atoi(sg)));
  if (const char* pair = std::getenv("B200LLM_PAIR")) e->pair_mode = std::atoi(pair);
  if (const char* env = std::getenv("B200LLM_SPLITS")) {  // "qkv,o,fc,proj" (0 keeps the pick)
    std::sscanf(env, "%d,%d,%d,%d", &e->tune_splits[0], &e->tune_splits[1], &e->tune_splits[2], &e->tune_splits[3]);
  }
  if (const char* env = std::getenv("B200LLM_PREFILL_ATTN")) e->prefill_attn_mode = std::atoi(env);
  if (const char* env = std::getenv("B200LLM_LNQKV_SPLIT")) e->tune_lnqkv_split = std::atoi(env);
  if (const char* env = std::getenv("B200LLM_FUSED_QKV")) e->tune_fused_qkv = std::atoi(env);
  if (const char* env = std::getenv("B200LLM_FUSED_SPLITS")) {  // "o,p"
    std::sscanf(env, "%d,%d", &e->tune_fused_splits[0], &e->tune_fused_splits[1]);
  }
  if (const char* env = std::getenv("B200LLM_ATTN_SPLITS")) e->tune_attn_splits = std::max(1, std::min(16, std::atoi(env)));
  if (const char* env = std::getenv("B200LLM_ATTN_STREAM")) e->tune_attn_stream = std::atoi(env);
  if (const char* env = std::getenv("B200LLM_ATTN_MERGE")) e->tune_attn_merge = std::atoi(env);
  if (const char* env = std::getenv("B200LLM_PAIR_SPLIT")) e->pair_split = std::atoi(env);
#endif
  // the ordered reductions (planes, cluster) promise bit-reproducible results: one add per output element
  if (cfg->decode_reduce == 1 || cfg->decode_reduce == 2) e->pair_split = 1;
  e->d = cfg->d_model;
  e->H = cfg->n_head;
  e->L = cfg->n_layer;
  e->V = cfg->vocab_size;
  e->P = cfg->n_positions;
  const int d = e->d, V = e->V, P = e->P, L = e->L, H = e->H;
  B200_CUDA_OK(cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking));
  e->stream = e->own_stream;
  cudaStream_t st = e->stream;

  auto cleanup = [&](int rc) {
    b200llm_engine_destroy(e.release());
    return rc;
  };
#define E_TRY(expr)                         \
  do {                                      \
    int _rc = (expr);                       \
    if (_rc != 0) return cleanup(_rc);      \
  } while (0)
#define E_CUDA(expr)                                                                         \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return cleanup(fail(kErrCuda, std::string(#expr) + ": " + cudaGetErrorString(_e)));    \
  } while (0)

  // ---- weights: stage fp32 pieces through one device buffer, convert on the GPU ----------
  const size_t biggest = std::max<size_t>(static_cast<size_t>(V) * d, static_cast<size_t>(4) * d * d);
  float* stage = nullptr;
  E_CUDA(cudaMalloc(reinterpret_cast<void**>(&stage), biggest * sizeof(float)));
  const float* src = weights_host;
  auto upload_f32 = [&](float** dst, size_t n) -> int {
    B200_PROPAGATE(dev_alloc_t(e.get(), dst, n));
    B200_CUDA_OK(cudaMemcpyAsync(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice, st));
    src += n;
    return kOk;
  };
  auto upload_conv1d = [&](__nv_bfloat16** dst, int in, int out) -> int {
    const size_t n = static_cast<size_t>(in) * out;
    B200_PROPAGATE(dev_alloc_t(e.get(), dst, n));
    B200_CUDA_OK(cudaMemcpyAsync(stage, src, n * sizeof(float), cudaMemcpyHostToDevice, st));
    B200_PROPAGATE(convert_transpose_bf16(stage, *dst, in, out, st));
    B200_CUDA_OK(cudaStreamSynchronize(st));  // `stage` is reused by the next tensor
    src += n;
    return kOk;
  };
  auto free_stage = [&](int rc) {
    cudaFree(stage);
    return rc;
  };
#define W_TRY(expr)                                    \
  do {                                                 \
    int _rc = (expr);                                  \
    if (_rc != 0) return cleanup(free_stage(_rc));     \
  } while (0)

  W_TRY(upload_f32(&e->wte_f32, static_cast<size_t>(V) * d));
  W_TRY(dev_alloc_t(e.get(), &e->wte_bf16, static_cast<size_t>(V) * d));
  W_TRY(convert_bf16(e->wte_f32, e->wte_bf16, static_cast<size_t>(V) * d, st));
  W_TRY(upload_f32(&e->wpe, static_cast<size_t>(P) * d));
  e->layers.resize(L);
  for (int l = 0; l < L; ++l) {
    LayerWeights& w = e->layers[l];
    W_TRY(upload_f32(&w.ln1_g, d));
    W_TRY(upload_f32(&w.ln1_b, d));
    W_TRY(upload_conv1d(&w.w_qkv, d, 3 * d));
    W_TRY(upload_f32(&w.b_qkv, 3 * d));
    W_TRY(upload_conv1d(&w.w_o, d, d));
    W_TRY(upload_f32(&w.b_o, d));
    W_TRY(upload_f32(&w.ln2_g, d));
    W_TRY(upload_f32(&w.ln2_b, d));
    W_TRY(upload_conv1d(&w.w_fc, d, 4 * d));
    W_TRY(upload_f32(&w.b_fc, 4 * d));
    W_TRY(upload_conv1d(&w.w_proj, 4 * d, d));
    W_TRY(upload_f32(&w.b_proj, d));
  }
  W_TRY(upload_f32(&e->lnf_g, d));
  W_TRY(upload_f32(&e->lnf_b, d));
  cudaStreamSynchronize(st);
  cudaFree(stage);

  // ---- KV pool ---------------------------------------------------------------------------
  e->max_pages_per_seq = (cfg->max_seq_len + kPageTokens - 1) / kPageTokens;
  int64_t pool_tokens = cfg->kv_pool_tokens > 0
                            ? cfg->kv_pool_tokens
                            : static_cast<int64_t>(cfg->max_seqs) * e->max_pages_per_seq * kPageTokens;
  e->n_pages = (pool_tokens + kPageTokens - 1) / kPageTokens;
  e->pool_layer_stride = static_cast<size_t>(e->n_pages) * H * kPageTokens * kHeadDim;
  E_TRY(dev_alloc_t(e.get(), &e->k_pool, e->pool_layer_stride * L));
  E_TRY(dev_alloc_t(e.get(), &e->v_pool, e->pool_layer_stride * L));
  // Pages start at zero: the prefill attention multiplies exact-zero softmax weights with whatever a
  // partly filled page holds behind its last token, and 0 x (uninitialised NaN pattern) is NaN.
  E_CUDA(cudaMemsetAsync(e->k_pool, 0, e->pool_layer_stride * L * sizeof(__nv_bfloat16), st));
  E_CUDA(cudaMemsetAsync(e->v_pool, 0, e->pool_layer_stride * L * sizeof(__nv_bfloat16), st));
  E_TRY(dev_alloc_t(e.get(), &e->page_table, static_cast<size_t>(cfg->max_seqs) * e->max_pages_per_seq));
  E_CUDA(cudaMemsetAsync(e->page_table, 0,
                         static_cast<size_t>(cfg->max_seqs) * e->max_pages_per_seq * sizeof(int32_t), st));
  e->free_pages.resize(e->n_pages);
  for (int64_t i = 0; i < e->n_pages; ++i) e->free_pages[i] = static_cast<int32_t>(e->n_pages - 1 - i);
  e->slot_pages.resize(cfg->max_seqs);
  e->slot_used.assign(cfg->max_seqs, 0);
  e->slot_capacity.assign(cfg->max_seqs, 0);
  e->slot_len_host.assign(cfg->max_seqs, 0);
  e->page_table_host.assign(static_cast<size_t>(cfg->max_seqs) * e->max_pages_per_seq, 0);
  e->slot_pt_dirty.assign(cfg->max_seqs, 0);
  e->page_ref.assign(e->n_pages, 0);
  e->page_cached.assign(e->n_pages, 0);
  e->page_key.resize(e->n_pages);
  e->slot_prefix.assign(cfg->max_seqs, 0);
  e->slot_prompt.resize(cfg->max_seqs);

  // ---- slot state -------------------------------------------------------------------------
  const size_t S = cfg->max_seqs;
  E_TRY(dev_alloc_t(e.get(), &e->seq_len, S));
  E_TRY(dev_alloc_t(e.get(), &e->last_tok, S));
  E_TRY(dev_alloc_t(e.get(), &e->finished, S));
  E_TRY(dev_alloc_t(e.get(), &e->n_gen, S));
  E_TRY(dev_alloc_t(e.get(), &e->min_new, S));
  E_TRY(dev_alloc_t(e.get(), &e->max_new, S));
  E_TRY(dev_alloc_t(e.get(), &e->top_k, S));
  E_TRY(dev_alloc_t(e.get(), &e->temperature, S));
  E_TRY(dev_alloc_t(e.get(), &e->seed, S));
  E_TRY(dev_alloc_t(e.get(), &e->out_tokens, S * cfg->max_new_tokens));
  E_CUDA(cudaMemsetAsync(e->seq_len, 0, S * sizeof(int32_t), st));
  E_CUDA(cudaMemsetAsync(e->last_tok, 0, S * sizeof(int32_t), st));
  E_CUDA(cudaMemsetAsync(e->finished, 0, S * sizeof(int32_t), st));
  E_CUDA(cudaMemsetAsync(e->n_gen, 0, S * sizeof(int32_t), st));

  // ---- workspace --------------------------------------------------------------------------
  const int dec_rows = bucket_for(cfg->max_seqs);
  e->max_rows = std::max(cfg->max_batch_tokens, dec_rows);
  e->max_rows = (e->max_rows + 127) / 128 * 128;
  const size_t R = e->max_rows;
  E_TRY(dev_alloc_t(e.get(), &e->x, R * d));
  E_TRY(dev_alloc_t(e.get(), &e->h, R * d));
  E_TRY(dev_alloc_t(e.get(), &e->qkv, R * 3 * d));
  E_TRY(dev_alloc_t(e.get(), &e->attn, R * d));
  E_TRY(dev_alloc_t(e.get(), &e->ffn, R * 4 * d));
  E_CUDA(cudaMemsetAsync(e->h, 0, R * d * sizeof(__nv_bfloat16), st));
  E_CUDA(cudaMemsetAsync(e->qkv, 0, R * 3 * d * sizeof(__nv_bfloat16), st));
  E_CUDA(cudaMemsetAsync(e->attn, 0, R * d * sizeof(__nv_bfloat16), st));
  E_CUDA(cudaMemsetAsync(e->ffn, 0, R * 4 * d * sizeof(__nv_bfloat16), st));
  E_CUDA(cudaMemsetAsync(e->x, 0, R * d * sizeof(float), st));
  e->logits_rows = dec_rows;
  E_TRY(dev_alloc_t(e.get(), &e->logits, static_cast<size_t>(e->logits_rows) * V));
  // candidate groups of the LM head's fused argmax: whole-tile layout or balanced row ranges,
  // depending on the row count of a launch (gemm_argmax_groups); the arrays hold the larger and
  // start at -inf, which is what a group that a layout leaves out must contain
  e->n_groups = std::max(((V + 127) / 128) * 4, gemm_argmax_groups(V, 1));
  E_TRY(dev_alloc_t(e.get(), &e->amax_val, static_cast<size_t>(e->n_groups) * e->logits_rows));
  E_TRY(dev_alloc_t(e.get(), &e->amax_idx, static_cast<size_t>(e->n_groups) * e->logits_rows));
  {
    std::vector<float> ninf(static_cast<size_t>(e->n_groups) * e->logits_rows, -INFINITY);
    E_CUDA(cudaMemcpy(e->amax_val, ninf.data(), ninf.size() * sizeof(float), cudaMemcpyHostToDevice));
    E_CUDA(cudaMemset(e->amax_idx, 0, ninf.size() * sizeof(int32_t)));
  }
  E_TRY(dev_alloc_t(e.get(), &e->eos_logit, static_cast<size_t>(e->logits_rows)));
  e->slot_samples.assign(cfg->max_seqs, 0);
  // split-K planes: at most 32 splits of [decode rows, d]
  // decode planes (up to 32 per row of the largest bucket), and room for the 2..4 planes of a
  // short-prompt prefill (<= 1024 rows) whatever max_seqs is
  e->partials_floats = std::max(static_cast<size_t>(32) * dec_rows * d,
                                static_cast<size_t>(4) * std::min(1024, e->cfg.max_batch_tokens) * d);
  E_TRY(dev_alloc_t(e.get(), &e->partials, e->partials_floats));
  e->attn_scratch_floats = decode_attention_scratch_floats(dec_rows, H, 16);
  E_TRY(dev_alloc_t(e.get(), &e->attn_scratch, e->attn_scratch_floats));
  // counters of the in-kernel split merge: zero now, and every launch leaves them at zero
  E_TRY(dev_alloc_t(e.get(), &e->attn_counters, static_cast<size_t>(dec_rows) * H));
  E_CUDA(cudaMemset(e->attn_counters, 0, static_cast<size_t>(dec_rows) * H * sizeof(int32_t)));
  E_TRY(dev_alloc_t(e.get(), &e->ids_dev, R));
  E_TRY(dev_alloc_t(e.get(), &e->row_pos, R));
  E_TRY(dev_alloc_t(e.get(), &e->row_slot, R));
  E_TRY(dev_alloc_t(e.get(), &e->cu_dev, S + 1));
  E_TRY(dev_alloc_t(e.get(), &e->slots_dev, S));
  E_TRY(dev_alloc_t(e.get(), &e->last_rows, S));
  E_TRY(dev_alloc_t(e.get(), &e->active_dev, static_cast<size_t>(dec_rows) + 1));  // [dec_rows]: live row count
  e->live_count_dev = e->active_dev + dec_rows;
  E_TRY(dev_alloc_t(e.get(), &e->status_dev, 2 * S));
  E_TRY(dev_alloc_t(e.get(), &e->init_dev, S));
  E_TRY(dev_alloc_t(e.get(), &e->prefix_dev, S));
  E_TRY(dev_alloc_t(e.get(), &e->tok_stage, S * cfg->max_new_tokens));
  E_CUDA(cudaMallocHost(reinterpret_cast<void**>(&e->pin_init), S * sizeof(SlotInit)));
  E_CUDA(cudaMallocHost(reinterpret_cast<void**>(&e->pin_pt), S * e->max_pages_per_seq * sizeof(int32_t)));
  E_CUDA(cudaMallocHost(reinterpret_cast<void**>(&e->pin_tok), S * cfg->max_new_tokens * sizeof(int32_t)));
  E_TRY(ensure_pinned(e.get(), R + 4 * S + 16));
  E_CUDA(cudaStreamSynchronize(st));
#undef E_TRY
#undef E_CUDA
#undef W_TRY
  *out_engine = e.release();
  return kOk;
}

int b200llm_engine_destroy(b200llm_engine* e) {
  if (e == nullptr) return kOk;
  cudaSetDevice(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& kv : e->graphs) {
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  }
  for (void* p : e->allocations) cudaFree(p);
  if (e->pin_event) cudaEventDestroy(e->pin_event);
  if (e->pin_i32) cudaFreeHost(e->pin_i32);
  if (e->pin_init) cudaFreeHost(e->pin_init);
  if (e->pin_pt) cudaFreeHost(e->pin_pt);
  if (e->pin_tok) cudaFreeHost(e->pin_tok);
  if (e->own_stream) cudaStreamDestroy(e->own_stream);
  gemm_clear_tensor_map_cache();
  delete e;
  return kOk;
}

int b200llm_engine_set_stream(b200llm_engine* e, void* stream) {
  B200_CHECK_ARG(e != nullptr, "set_stream: null engine");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  B200_CUDA_OK(cudaStreamSynchronize(e->stream));
  e->stream = stream ? reinterpret_cast<cudaStream_t>(stream) : e->own_stream;
  // graphs were captured on the previous stream's topology; they replay on any stream
  return kOk;
}

int b200llm_seq_alloc(b200llm_engine* e, int32_t capacity_tokens, int32_t* out_slot) {
  B200_CHECK_ARG(e && out_slot, "seq_alloc: null argument");
  return alloc_slot(e, capacity_tokens, nullptr, 0, out_slot, nullptr);
}

int b200llm_seq_alloc_with_prefix(b200llm_engine* e, const int32_t* ids_host, int32_t n_ids,
                                  int32_t capacity_tokens, int32_t* out_slot, int32_t* out_cached) {
  B200_CHECK_ARG(e && ids_host && out_slot, "seq_alloc_with_prefix: null argument");
  B200_CHECK_ARG(n_ids > 0 && n_ids <= capacity_tokens, "seq_alloc_with_prefix: prompt exceeds the capacity");
  return alloc_slot(e, capacity_tokens, ids_host, n_ids, out_slot, out_cached);
}

int b200llm_prefix_cache_stats(b200llm_engine* e, int64_t* lookups, int64_t* hits, int64_t* pages) {
  B200_CHECK_ARG(e != nullptr, "prefix_cache_stats: null engine");
  if (lookups) *lookups = e->prefix_lookup_tokens;
  if (hits) *hits = e->prefix_hit_tokens;
  if (pages) *pages = e->prefix_cached_pages;
  return kOk;
}

int b200llm_seq_free(b200llm_engine* e, int32_t slot) {
  B200_CHECK_ARG(e != nullptr, "seq_free: null engine");
  B200_CHECK_ARG(slot >= 0 && slot < e->cfg.max_seqs && e->slot_used[slot], "seq_free: bad slot");
  for (int32_t p : e->slot_pages[slot]) {
    if (--e->page_ref[p] == 0) e->free_pages.push_back(p);
  }
  e->slot_pages[slot].clear();
  e->slot_prompt[slot].clear();
  e->slot_prefix[slot] = 0;
  e->slot_used[slot] = 0;
  e->slot_capacity[slot] = 0;
  e->slot_len_host[slot] = 0;
  return kOk;
}

int b200llm_seq_cancel(b200llm_engine* e, int32_t slot) {
  B200_CHECK_ARG(e != nullptr, "seq_cancel: null engine");
  B200_CHECK_ARG(slot >= 0 && slot < e->cfg.max_seqs && e->slot_used[slot], "seq_cancel: bad slot");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  cancel_slot_kernel<<<1, 1, 0, e->stream>>>(e->finished, slot);
  B200_CUDA_OK(cudaGetLastError());
  e->launches += 1;
  return kOk;
}

int b200llm_kv_pages_free(b200llm_engine* e, int64_t* out_free, int64_t* out_total) {
  B200_CHECK_ARG(e != nullptr, "kv_pages_free: null engine");
  if (out_free) *out_free = static_cast<int64_t>(e->free_pages.size());
  if (out_total) *out_total = e->n_pages;
  return kOk;
}

int b200llm_prefill(b200llm_engine* e, const int32_t* ids, int32_t ids_on_device,
                    const int32_t* cu, const int32_t* slots, const b200llm_sampling* sampling,
                    int32_t batch, float* logits_all_dev) {
  B200_CHECK_ARG(e && ids && cu && slots, "prefill: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->cfg.max_seqs, "prefill: bad batch");
  B200_CHECK_ARG(batch <= e->logits_rows, "prefill: batch exceeds logits workspace");
  B200_CHECK_ARG(cu[0] == 0, "prefill: cu_seqlens must start at 0");
  // Rows that are computed: the whole prompt, or -- for a slot whose leading pages came from the
  // prefix cache -- only the tokens behind them. `cu` always describes the full prompts.
  std::vector<int32_t> prefix(batch, 0), cu_rows(batch + 1, 0);
  bool any_prefix = false;
  int max_len = 0, max_keys = 0;
  for (int b = 0; b < batch; ++b) {
    const int len = cu[b + 1] - cu[b];
    B200_CHECK_ARG(len > 0, "prefill: empty prompt");
    B200_CHECK_ARG(len <= e->P, "prefill: prompt longer than the position table");
    const int s = slots[b];
    B200_CHECK_ARG(s >= 0 && s < e->cfg.max_seqs && e->slot_used[s], "prefill: slot not allocated");
    B200_CHECK_ARG(len <= e->slot_capacity[s], "prefill: prompt exceeds the slot's KV capacity");
    prefix[b] = e->slot_prefix[s];
    if (prefix[b] > 0) {
      B200_CHECK_ARG(!ids_on_device, "prefill: a slot with a cached prefix needs its ids in host memory");
      B200_CHECK_ARG(static_cast<int>(e->slot_prompt[s].size()) == len &&
                         std::memcmp(e->slot_prompt[s].data(), ids + cu[b], sizeof(int32_t) * len) == 0,
                     "prefill: ids differ from the prompt given to seq_alloc_with_prefix");
      B200_CHECK_ARG(logits_all_dev == nullptr, "prefill: all-position logits need an uncached prompt");
      any_prefix = true;
    }
    cu_rows[b + 1] = cu_rows[b] + len - prefix[b];
    max_len = std::max(max_len, len - prefix[b]);
    max_keys = std::max(max_keys, len);
  }
  const int n_tok = cu_rows[batch];
  if (n_tok > e->cfg.max_batch_tokens) {
    return fail(kErrCapacity, "prefill: packed tokens exceed max_batch_tokens");
  }
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = e->stream;

  // staging: [cu (batch+1) | slots (batch) | prefix (batch) | ids (n_tok, host case)]
  // pinned buffers are reused call to call: the previous call's copies must have drained
  B200_PROPAGATE(wait_pinned_free(e));
  B200_PROPAGATE(ensure_pinned(e, static_cast<size_t>(3 * batch + 1 + n_tok)));
  int32_t* pin = e->pin_i32;
  std::memcpy(pin, cu_rows.data(), (batch + 1) * sizeof(int32_t));
  std::memcpy(pin + batch + 1, slots, batch * sizeof(int32_t));
  std::memcpy(pin + 2 * batch + 1, prefix.data(), batch * sizeof(int32_t));
  B200_CUDA_OK(cudaMemcpyAsync(e->cu_dev, pin, (batch + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  B200_CUDA_OK(cudaMemcpyAsync(e->slots_dev, pin + batch + 1, batch * sizeof(int32_t),
                               cudaMemcpyHostToDevice, st));
  if (any_prefix) {
    B200_CUDA_OK(cudaMemcpyAsync(e->prefix_dev, pin + 2 * batch + 1, batch * sizeof(int32_t),
                                 cudaMemcpyHostToDevice, st));
  }
  // page-table rows of freshly allocated slots: stage in pinned memory, one copy per run of
  // consecutive slots
  {
    std::vector<int32_t> dirty;
    for (int b = 0; b < batch; ++b) {
      if (e->slot_pt_dirty[slots[b]]) dirty.push_back(slots[b]);
    }
    std::sort(dirty.begin(), dirty.end());
    dirty.erase(std::unique(dirty.begin(), dirty.end()), dirty.end());
    const size_t mp = e->max_pages_per_seq;
    size_t i = 0;
    int32_t* pt_pin = e->pin_pt;
    while (i < dirty.size()) {
      size_t j = i + 1;
      while (j < dirty.size() && dirty[j] == dirty[j - 1] + 1) ++j;
      const size_t first = dirty[i], n_rows = j - i;
      std::memcpy(pt_pin + first * mp, e->page_table_host.data() + first * mp, n_rows * mp * sizeof(int32_t));
      B200_CUDA_OK(cudaMemcpyAsync(e->page_table + first * mp, pt_pin + first * mp,
                                   n_rows * mp * sizeof(int32_t), cudaMemcpyHostToDevice, st));
      i = j;
    }
    for (int32_t sidx : dirty) e->slot_pt_dirty[sidx] = 0;
  }
  const int32_t* ids_dev = ids;
  if (!ids_on_device) {
    const int n_all = cu[batch];
    for (int i = 0; i < n_all; ++i) {
      if (ids[i] < 0 || ids[i] >= e->V) return fail(kErrInvalidArg, "prefill: token id out of range");
    }
    int32_t* dst = pin + 3 * batch + 1;
    for (int b = 0; b < batch; ++b) {  // only the rows that are computed travel
      const int n = cu[b + 1] - cu[b] - prefix[b];
      std::memcpy(dst + cu_rows[b], ids + cu[b] + prefix[b], n * sizeof(int32_t));
    }
    B200_CUDA_OK(cudaMemcpyAsync(e->ids_dev, dst, n_tok * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    ids_dev = e->ids_dev;
  }
  for (int b = 0; b < batch; ++b) {
    SlotInit& si = e->pin_init[b];
    si.slot = slots[b];
    const int cap_gen = e->cfg.max_new_tokens;
    int mx = sampling ? sampling[b].max_new_tokens : cap_gen;
    if (mx <= 0 || mx > cap_gen) mx = cap_gen;
    const int room = e->slot_capacity[slots[b]] - (cu[b + 1] - cu[b]) + 1;  // last token needs no KV
    if (mx > room) mx = room;
    if (mx < 1) mx = 1;
    si.max_new = mx;
    si.min_new = sampling ? std::min(sampling[b].min_new_tokens, mx) : 0;
    si.temperature = sampling ? sampling[b].temperature : 0.0f;
    si.top_k = sampling ? sampling[b].top_k : 0;
    si.seed = sampling ? sampling[b].seed : 0;
    si.pad = 0;
    e->slot_len_host[slots[b]] = cu[b + 1] - cu[b];
    e->slot_samples[slots[b]] = (si.temperature > 0.0f && si.top_k > 1) ? 1 : 0;
  }
  B200_CUDA_OK(cudaMemcpyAsync(e->init_dev, e->pin_init, batch * sizeof(SlotInit),
                               cudaMemcpyHostToDevice, st));
  B200_PROPAGATE(mark_pinned_used(e));
  init_slots_kernel<<<(batch + 127) / 128, 128, 0, st>>>(e->init_dev, batch, e->last_tok, e->finished,
                                                         e->n_gen, e->min_new, e->max_new,
                                                         e->temperature, e->top_k, e->seed);
  B200_CUDA_OK(cudaGetLastError());
  const int meta_threads = std::max(n_tok, batch);
  prefill_meta_kernel<<<(meta_threads + 255) / 256, 256, 0, st>>>(
      e->cu_dev, e->slots_dev, any_prefix ? e->prefix_dev : nullptr, batch, e->row_pos, e->row_slot,
      e->last_rows, e->seq_len);
  B200_CUDA_OK(cudaGetLastError());
  e->launches += 2;

  B200_PROPAGATE(count(e, embed_tokens(e->wte_f32, e->wpe, ids_dev, e->row_pos, e->x, n_tok, e->d,
                                       e->V, e->P, st)));
  StepPlan p;
  p.rows = n_tok;
  p.decode = false;
  p.batch = batch;
  p.max_len = max_len;
  p.max_keys = max_keys;
  p.logits_all = logits_all_dev;
  p.store_logits = true;
  p.paged_prefill = any_prefix;
  // Programmatic launch for the kernels that wait on their grid dependency themselves (GEMMs,
  // LayerNorm, select); prefill attention is a plain launch and so still waits for the whole QKV
  // GEMM. A one-request prompt is ~90 tiny kernels: the launch gaps are most of its time.
  p.pdl = (e->cfg.use_pdl && e->prefill_pdl) ? 1 : 0;
  // Short prompts (a single chat request): the two projections back to d have only
  // ceil(n_tok / 128) * d / 128 tiles, a fraction of the SMs, and the MLP one runs 4d / 64
  // k-blocks per tile. Split K into partial planes (reduced by the LayerNorm that follows, like
  // in a decode step) when the decode-sized plane buffer can hold them.
  if (n_tok <= 1024 && e->prefill_split) {
    const int tiles = ((n_tok + 127) / 128) * ((e->d + 127) / 128);
    auto pick = [&](int kb) {
      int best = 1;
      for (int sp = 2; sp <= 4; ++sp) {
        if (kb % sp == 0 && tiles * sp <= e->sm_count &&
            static_cast<size_t>(sp) * n_tok * e->d <= e->partials_floats) {
          best = sp;
        }
      }
      return best;
    };
    p.split_o = pick(e->d / 64);
    p.split_p = pick(4 * e->d / 64);
  }
  B200_PROPAGATE(run_stack(e, p));
  // the cached prefixes are consumed; full prompt pages join the cache
  for (int b = 0; b < batch; ++b) {
    e->slot_prefix[slots[b]] = 0;
    if (e->cfg.prefix_cache && !ids_on_device) {
      register_prompt_pages(e, slots[b], ids + cu[b], cu[b + 1] - cu[b]);
    }
  }
  e->last_batch = batch;
  return kOk;
}

int b200llm_decode(b200llm_engine* e, const int32_t* slots, int32_t batch, int32_t n_steps) {
  B200_CHECK_ARG(e && slots, "decode: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->cfg.max_seqs, "decode: bad batch");
  B200_CHECK_ARG(n_steps > 0, "decode: n_steps must be positive");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = e->stream;
  const int bucket = bucket_for(batch);
  B200_CHECK_ARG(bucket <= e->logits_rows, "decode: batch exceeds workspace");

  int max_ctx = 1;
  bool need_logits = e->cfg.keep_logits != 0;
  for (int b = 0; b < batch; ++b) {
    const int s = slots[b];
    B200_CHECK_ARG(s >= 0 && s < e->cfg.max_seqs && e->slot_used[s], "decode: slot not allocated");
    max_ctx = std::max(max_ctx, e->slot_len_host[s] + n_steps);
    need_logits = need_logits || e->slot_samples[s] != 0;
  }

  // upload the active set when it changed
  std::vector<int32_t> active(bucket, -1);
  std::copy(slots, slots + batch, active.begin());
  if (active != e->active_host) {
    B200_PROPAGATE(wait_pinned_free(e));
    B200_PROPAGATE(ensure_pinned(e, static_cast<size_t>(bucket) + 1));
    std::memcpy(e->pin_i32, active.data(), bucket * sizeof(int32_t));
    e->pin_i32[bucket] = batch;
    B200_CUDA_OK(cudaMemcpyAsync(e->active_dev, e->pin_i32, bucket * sizeof(int32_t),
                                 cudaMemcpyHostToDevice, st));
    B200_CUDA_OK(cudaMemcpyAsync(e->live_count_dev, e->pin_i32 + bucket, sizeof(int32_t),
                                 cudaMemcpyHostToDevice, st));
    B200_PROPAGATE(mark_pinned_used(e));
    e->active_host = active;
  }

  StepPlan p;
  p.rows = bucket;
  p.decode = true;
  p.batch = bucket;
  p.pdl = e->cfg.use_pdl ? 1 : 0;
  p.store_logits = need_logits;
  p.live_rows = batch;
  const int sms = e->sm_count;
  // auto: pick_split's rule, never more than 12 planes
  const int cap = e->cfg.decode_split_k == 0 ? 12 : std::min(e->cfg.decode_split_k, 32);
  // automatic: the fused layer where it was measured faster (weights of a CTA resident: d <= 768);
  // decode_reduce = 3 forces it wherever the kernels support the shape
  if (bucket <= 64 && ln_gemm_supported(bucket, e->d, 3 * e->d) && e->d % 128 == 0 &&
      (e->cfg.decode_reduce == 3 || (e->cfg.decode_reduce == 0 && ln_gemm_weights_resident(e->d)))) {
    p.fused = true;
    p.rs_o = gemm_redadd_pick_split(e->d, e->d, bucket, sms);
    p.rs_p = gemm_redadd_pick_split(e->d, 4 * e->d, bucket, sms);
    if (e->tune_fused_splits[0] > 0) p.rs_o = e->tune_fused_splits[0];
    if (e->tune_fused_splits[1] > 0) p.rs_p = e->tune_fused_splits[1];
    if (p.rs_o <= 0 || p.rs_p <= 0) p.fused = false;
  }
  if (p.fused) {
    // no split-K planes anywhere in the step
  } else if (bucket <= 256 && cap > 1 && e->cfg.decode_reduce == 2) {
    p.cl_qkv = pick_cluster(3 * e->d, e->d, sms);
    p.cl_o = pick_cluster(e->d, e->d, sms);
    p.cl_fc = pick_cluster(4 * e->d, e->d, sms);
    p.cl_p = pick_cluster(e->d, 4 * e->d, sms);
  } else if (bucket <= 256 && cap > 1) {
    p.split_o = pick_split(e->d, e->d, sms, cap);
    p.split_p = pick_split(e->d, 4 * e->d, sms, cap);
    p.split_qkv = pick_split(3 * e->d, e->d, sms, cap);
    p.split_fc = pick_split(4 * e->d, e->d, sms, cap);
  }
  if (e->tune_splits[0] > 0) p.split_qkv = e->tune_splits[0];
  if (e->tune_splits[1] > 0) p.split_o = e->tune_splits[1];
  if (e->tune_splits[2] > 0) p.split_fc = e->tune_splits[2];
  if (e->tune_splits[3] > 0) p.split_p = e->tune_splits[3];
  // split-KV when (rows x heads) alone cannot fill the machine and contexts are long
  {
    // One CTA per (row, head, split), six of them per SM, each walking its share of the context in
    // 128-row passes. Cost model: waves of CTAs x (passes per CTA + a fixed half pass of set-up);
    // fewer splits win ties. Doubling alone picked 4 splits for 16 rows x 16 heads at context 1100:
    // 1024 CTAs on 888 slots = two waves; 3 splits (768 CTAs, one wave) is the better cut.
    const int ctas = bucket * e->H;
    const int slots = 6 * sms;
    int splits = 1;
    float best = 1e30f;
    for (int s = 1; s <= 16; ++s) {
      const int per = (max_ctx + s - 1) / s;
      if (s > 1 && per < 64) break;
      const int waves = (ctas * s + slots - 1) / slots;
      const float cost = static_cast<float>(waves) * (static_cast<float>((per + 127) / 128) + 0.5f);
      if (cost < best - 1e-3f) {
        best = cost;
        splits = s;
      }
    }
    p.attn_splits = splits;
    // More units than resident CTAs, or several passes per CTA: the streaming kernel (persistent warps,
    // twelve per SM, each walking whole units in 32-token tiles). Its cut: rounds of units per warp x
    // (tiles per unit + two tiles' worth of per-unit set-up).
    // It takes over by KV volume: from ~4096 tiles of 32 tokens per layer (~33 MB). Measured, us per step,
    // kernel above -> streaming: 64 x 12 heads at context 144 (3456 tiles) 343 -> 345, at 192 387 -> 380, at
    // 256 445 -> 408, at 512 641 -> 515; 16 x 12 at 512 (3072 tiles) 321 -> 335; 32 x 20 at 96 1094 -> 1139,
    // at 256 1417 -> 1302; 16 x 16 at 1100 968 -> 835; 256 x 20 at 96 3300 -> 2675; one request at context
    // 500 (192 tiles; the kernel above merges its splits inside a cluster) 233 -> 273.
    p.attn_stream = static_cast<long long>(p.live_rows) * e->H * ((max_ctx + 31) / 32) >= 4096;
    if (e->tune_attn_stream >= 0) p.attn_stream = e->tune_attn_stream != 0;
    if (p.attn_stream) {
      // Two instances: nine warps per SM with three tiles in their rings when that seats every unit at
      // once (long contexts: GPT-2 medium, 16 rows x 16 heads x 5 splits at context 1100: 835 vs 853 us
      // per step), twelve warps with two tiles otherwise (GPT-2 large, 256 rows: 2690 vs 2748 us).
      auto cut = [&](int warps, int* out_splits) {
        float best_cost = 1e30f;
        for (int s = 1; s <= 16; ++s) {
          const int per = (max_ctx + s - 1) / s;
          if (s > 1 && per < 64) break;
          const int rounds = (ctas * s + warps - 1) / warps;
          const float cost = static_cast<float>(rounds) * (static_cast<float>((per + 31) / 32) + 2.f);
          if (cost < best_cost - 1e-3f) {
            best_cost = cost;
            *out_splits = s;
          }
        }
      };
      int s9 = 1, s12 = 1;
      cut(9 * sms, &s9);
      cut(12 * sms, &s12);
      if (ctas * s9 <= 9 * sms) {
        p.attn_stream_instance = 2;
        p.attn_splits = s9;
      } else {
        p.attn_stream_instance = 1;
        p.attn_splits = s12;
      }
      if (e->tune_attn_stream >= 1 && e->tune_attn_stream <= 5) p.attn_stream_instance = e->tune_attn_stream;  // (9: forced on, instance by the rule)
    }
    if (e->tune_attn_splits > 0) p.attn_splits = e->tune_attn_splits;
  }
  if (p.fused) {
    // ln_1 + c_attn as one kernel holds whole SMs (its CTAs keep their 196 KB weight tile resident), so
    // the attention CTAs of the layer can only start early -- and prefetch their KV rows under the
    // projection -- on the SMs it leaves free. When they all fit there the fused kernel wins (16 rows:
    // 265 vs 276 us per step); with 64 rows x 12 heads they do not, and LayerNorm + split-K planes
    // (light CTAs) is the faster producer for the attention kernel (340 vs 347 us).
    {
      const int qkv_ctas = ((3 * e->d + 127) / 128) * ((bucket + 15) / 16);
      const int early_slots = (sms - qkv_ctas) * 6;  // six attention CTAs share an SM
      p.fused_qkv = bucket * e->H * p.attn_splits <= early_slots;
    }
    if (e->tune_fused_qkv >= 0) p.fused_qkv = e->tune_fused_qkv != 0;
    p.lnqkv_split = e->tune_lnqkv_split;
    if (!p.fused_qkv && cap > 1) p.split_qkv = pick_split(3 * e->d, e->d, sms, cap);
    if (e->tune_splits[0] > 0) p.split_qkv = e->tune_splits[0];
  }

  if (!e->cfg.use_cuda_graphs) {
    for (int i = 0; i < n_steps; ++i) B200_PROPAGATE(decode_step_launches(e, p));
  } else {
    // Steps are captured in groups: inside a graph the selector -> embed edge of two consecutive
    // steps is a programmatic edge (~1 us), between two graph launches it is a full drain + launch
    // (~4 us). The remainder runs on the single-step graph.
    auto get_graph = [&](int steps, b200llm_engine::GraphEntry** out) -> int {
      // (the in-kernel split combine depends on the live row count, so it is part of the key)
      const int attn_cluster = p.attn_stream ? 1 + p.attn_stream_instance : (p.live_rows * e->H * p.attn_splits <= 2 * e->sm_count) ? 1 : 0;
      b200llm_engine::GraphKey key{bucket, p.attn_splits, need_logits ? 1 : 0, steps, attn_cluster};
      auto it = e->graphs.find(key);
      if (it == e->graphs.end()) {
        cudaGraph_t graph = nullptr;
        const int64_t launches_before = e->launches;
        e->launches_in_capture = 0;
        B200_CUDA_OK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        int rc = kOk;
        for (int i = 0; i < steps && rc == kOk; ++i) rc = decode_step_launches(e, p);
        cudaError_t cerr = cudaStreamEndCapture(st, &graph);
        e->launches = launches_before;  // captured, not executed
        if (rc != kOk) {
          if (graph) cudaGraphDestroy(graph);
          return rc;
        }
        if (cerr != cudaSuccess) {
          return fail(kErrCuda, std::string("decode: graph capture failed: ") + cudaGetErrorString(cerr));
        }
        b200llm_engine::GraphEntry entry;
        entry.kernels = e->launches_in_capture;
        cerr = cudaGraphInstantiate(&entry.exec, graph, 0);
        cudaGraphDestroy(graph);
        if (cerr != cudaSuccess) {
          return fail(kErrCuda, std::string("decode: graph instantiate failed: ") + cudaGetErrorString(cerr));
        }
        it = e->graphs.emplace(key, entry).first;
      }
      *out = &it->second;
      return kOk;
    };
    const int group = e->steps_per_graph;
    int done = 0;
    if (group > 1 && n_steps >= group) {
      b200llm_engine::GraphEntry* g = nullptr;
      B200_PROPAGATE(get_graph(group, &g));
      for (; done + group <= n_steps; done += group) {
        B200_CUDA_OK(cudaGraphLaunch(g->exec, st));
        e->launches += g->kernels;
      }
    }
    if (done < n_steps) {
      b200llm_engine::GraphEntry* g = nullptr;
      B200_PROPAGATE(get_graph(1, &g));
      for (; done < n_steps; ++done) {
        B200_CUDA_OK(cudaGraphLaunch(g->exec, st));
        e->launches += g->kernels;
      }
    }
  }
  for (int b = 0; b < batch; ++b) {
    int& len = e->slot_len_host[slots[b]];
    len = std::min(len + n_steps, e->slot_capacity[slots[b]]);
  }
  e->last_batch = bucket;
  return kOk;
}

int b200llm_sync(b200llm_engine* e) {
  B200_CHECK_ARG(e != nullptr, "sync: null engine");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  B200_CUDA_OK(cudaStreamSynchronize(e->stream));
  return kOk;
}

int b200llm_read_tokens(b200llm_engine* e, int32_t slot, int32_t* out, int32_t max_tokens,
                        int32_t* out_count, int32_t* out_finished) {
  B200_CHECK_ARG(e && out && out_count, "read_tokens: null argument");
  B200_CHECK_ARG(slot >= 0 && slot < e->cfg.max_seqs, "read_tokens: bad slot");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  int32_t st2[2] = {0, 0};
  B200_CUDA_OK(cudaMemcpyAsync(&st2[0], e->n_gen + slot, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
  B200_CUDA_OK(cudaMemcpyAsync(&st2[1], e->finished + slot, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
  B200_CUDA_OK(cudaStreamSynchronize(e->stream));
  const int n = std::min(st2[0], max_tokens);
  if (n > 0) {
    B200_CUDA_OK(cudaMemcpyAsync(out, e->out_tokens + static_cast<size_t>(slot) * e->cfg.max_new_tokens,
                                 n * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    B200_CUDA_OK(cudaStreamSynchronize(e->stream));
  }
  *out_count = n;
  if (out_finished) *out_finished = st2[1];
  return kOk;
}

int b200llm_read_status(b200llm_engine* e, const int32_t* slots, int32_t batch, int32_t* out_counts,
                        int32_t* out_finished) {
  B200_CHECK_ARG(e && slots && out_counts && out_finished, "read_status: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->cfg.max_seqs, "read_status: bad batch");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = e->stream;
  B200_CUDA_OK(cudaStreamSynchronize(st));
  e->pin_busy = false;
  B200_PROPAGATE(ensure_pinned(e, static_cast<size_t>(3 * batch)));
  std::memcpy(e->pin_i32, slots, batch * sizeof(int32_t));
  B200_CUDA_OK(cudaMemcpyAsync(e->last_rows, e->pin_i32, batch * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  gather_status_kernel<<<(batch + 127) / 128, 128, 0, st>>>(e->last_rows, batch, e->n_gen, e->finished,
                                                            e->status_dev);
  B200_CUDA_OK(cudaGetLastError());
  B200_CUDA_OK(cudaMemcpyAsync(e->pin_i32 + batch, e->status_dev, 2 * batch * sizeof(int32_t),
                               cudaMemcpyDeviceToHost, st));
  B200_CUDA_OK(cudaStreamSynchronize(st));
  std::memcpy(out_counts, e->pin_i32 + batch, batch * sizeof(int32_t));
  std::memcpy(out_finished, e->pin_i32 + 2 * batch, batch * sizeof(int32_t));
  return kOk;
}

int b200llm_read_tokens_batch(b200llm_engine* e, const int32_t* slots, int32_t batch, int32_t only_finished,
                              int32_t* out_tokens, int32_t out_stride, int32_t* out_counts,
                              int32_t* out_finished) {
  B200_CHECK_ARG(e && slots && out_tokens && out_counts && out_finished, "read_tokens_batch: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->cfg.max_seqs, "read_tokens_batch: bad batch");
  B200_CHECK_ARG(out_stride > 0, "read_tokens_batch: bad stride");
  for (int b = 0; b < batch; ++b) {
    B200_CHECK_ARG(slots[b] >= 0 && slots[b] < e->cfg.max_seqs, "read_tokens_batch: bad slot");
  }
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  cudaStream_t st = e->stream;
  const int mn = e->cfg.max_new_tokens;
  // everything is enqueued behind the decode steps already on the stream; ONE wait at the end
  B200_PROPAGATE(wait_pinned_free(e));
  B200_PROPAGATE(ensure_pinned(e, static_cast<size_t>(3 * batch)));
  std::memcpy(e->pin_i32, slots, batch * sizeof(int32_t));
  B200_CUDA_OK(cudaMemcpyAsync(e->last_rows, e->pin_i32, batch * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  gather_status_kernel<<<(batch + 127) / 128, 128, 0, st>>>(e->last_rows, batch, e->n_gen, e->finished,
                                                            e->status_dev);
  B200_CUDA_OK(cudaGetLastError());
  gather_tokens_kernel<<<batch, 64, 0, st>>>(e->last_rows, batch, e->finished, only_finished ? 1 : 0,
                                             e->out_tokens, mn, e->tok_stage);
  B200_CUDA_OK(cudaGetLastError());
  e->launches += 2;
  B200_CUDA_OK(cudaMemcpyAsync(e->pin_i32 + batch, e->status_dev, 2 * batch * sizeof(int32_t),
                               cudaMemcpyDeviceToHost, st));
  B200_CUDA_OK(cudaMemcpyAsync(e->pin_tok, e->tok_stage, static_cast<size_t>(batch) * mn * sizeof(int32_t),
                               cudaMemcpyDeviceToHost, st));
  B200_CUDA_OK(cudaStreamSynchronize(st));
  e->pin_busy = false;
  for (int b = 0; b < batch; ++b) {
    const int n = e->pin_i32[batch + b], fin = e->pin_i32[2 * batch + b];
    out_counts[b] = n;
    out_finished[b] = fin;
    if (only_finished && !fin) continue;
    const int take = std::min(std::min(n, out_stride), mn);
    std::memcpy(out_tokens + static_cast<size_t>(b) * out_stride, e->pin_tok + static_cast<size_t>(b) * mn,
                take * sizeof(int32_t));
  }
  return kOk;
}

int b200llm_read_last_logits(b200llm_engine* e, int32_t batch, float* out) {
  B200_CHECK_ARG(e && out, "read_last_logits: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->logits_rows, "read_last_logits: bad batch");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  B200_CUDA_OK(cudaMemcpyAsync(out, e->logits, static_cast<size_t>(batch) * e->V * sizeof(float),
                               cudaMemcpyDeviceToHost, e->stream));
  B200_CUDA_OK(cudaStreamSynchronize(e->stream));
  return kOk;
}

int b200llm_generate(b200llm_engine* e, const int32_t* ids_host, const int32_t* cu,
                     const b200llm_sampling* sampling, int32_t batch, int32_t* out_tokens,
                     int32_t out_stride, int32_t* out_counts) {
  B200_CHECK_ARG(e && ids_host && cu && out_tokens && out_counts, "generate: null argument");
  B200_CHECK_ARG(batch > 0 && batch <= e->cfg.max_seqs, "generate: bad batch");
  B200_CUDA_OK(cudaSetDevice(e->cfg.device));
  std::vector<int32_t> slots(batch, -1);
  int max_new = 1;
  auto release = [&](int rc) {
    for (int s : slots) {
      if (s >= 0) b200llm_seq_free(e, s);
    }
    return rc;
  };
  for (int b = 0; b < batch; ++b) {
    int mx = sampling ? sampling[b].max_new_tokens : e->cfg.max_new_tokens;
    if (mx <= 0 || mx > e->cfg.max_new_tokens) mx = e->cfg.max_new_tokens;
    const int len = cu[b + 1] - cu[b];
    int cap = std::min(len + mx, e->cfg.max_seq_len);
    if (len > cap) return release(fail(kErrCapacity, "generate: prompt exceeds max_seq_len"));
    int rc = alloc_slot(e, cap, ids_host + cu[b], len, &slots[b], nullptr);
    if (rc != kOk) return release(rc);
    max_new = std::max(max_new, mx);
  }
  int rc = b200llm_prefill(e, ids_host, 0, cu, slots.data(), sampling, batch, nullptr);
  if (rc != kOk) return release(rc);

  // Decode in chunks; between chunks one small device->host status read decides whether
  // every sequence is finished (EOS) so the loop can stop early.
  std::vector<int32_t> counts(batch), fin(batch);
  int done_steps = 1;  // prefill produced the first token
  bool can_finish_early = false;
  for (int b = 0; b < batch; ++b) {
    const int mn = sampling ? sampling[b].min_new_tokens : 0;
    int mx = sampling ? sampling[b].max_new_tokens : e->cfg.max_new_tokens;
    if (mx <= 0 || mx > e->cfg.max_new_tokens) mx = e->cfg.max_new_tokens;
    if (mn < mx) can_finish_early = true;
  }
  while (done_steps < max_new) {
    const int chunk = can_finish_early ? std::min(16, max_new - done_steps) : (max_new - done_steps);
    rc = b200llm_decode(e, slots.data(), batch, chunk);
    if (rc != kOk) return release(rc);
    done_steps += chunk;
    if (can_finish_early && done_steps < max_new) {
      rc = b200llm_read_status(e, slots.data(), batch, counts.data(), fin.data());
      if (rc != kOk) return release(rc);
      bool all = true;
      for (int b = 0; b < batch; ++b) all = all && fin[b] != 0;
      if (all) break;
    }
  }
  rc = b200llm_read_status(e, slots.data(), batch, counts.data(), fin.data());
  if (rc != kOk) return release(rc);
  // generated tokens: one device->host copy per run of consecutive slots into pinned memory
  {
    const size_t mn = e->cfg.max_new_tokens;
    std::vector<int32_t> order(slots);
    std::sort(order.begin(), order.end());
    size_t i = 0;
    while (i < order.size()) {
      size_t j = i + 1;
      while (j < order.size() && order[j] == order[j - 1] + 1) ++j;
      const size_t first = order[i], n_rows = j - i;
      cudaError_t err = cudaMemcpyAsync(e->pin_tok + first * mn, e->out_tokens + first * mn,
                                        n_rows * mn * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream);
      if (err != cudaSuccess) return release(fail(kErrCuda, cudaGetErrorString(err)));
      i = j;
    }
    cudaError_t err = cudaStreamSynchronize(e->stream);
    if (err != cudaSuccess) return release(fail(kErrCuda, cudaGetErrorString(err)));
    for (int b = 0; b < batch; ++b) {
      const int n = std::min(counts[b], out_stride);
      const int32_t* row = e->pin_tok + static_cast<size_t>(slots[b]) * mn;
      for (int t = 0; t < out_stride; ++t) {
        out_tokens[static_cast<size_t>(b) * out_stride + t] = t < n ? row[t] : e->cfg.pad_token_id;
      }
      out_counts[b] = n;
    }
  }
  return release(kOk);
}

int64_t b200llm_kernel_launches(const b200llm_engine* e) { return e ? e->launches : 0; }

#ifdef B200_STEP_TRACE
// Profiling builds only. op 0: arm (allocates a buffer of `capacity` records, resets the counter);
// op 1: read back up to `capacity` 16-word records (entry ns, exit ns, tag, -, milestones...) and return the count.
int b200llm_debug_step_trace(int op, unsigned long long* out, int capacity) {
  static unsigned long long* buf = nullptr;
  static unsigned int* idx = nullptr;
  static int cap = 0;
  if (op == 0) {
    if (buf == nullptr) {
      cap = capacity;
      cudaMalloc(&buf, static_cast<size_t>(cap) * 16 * sizeof(unsigned long long));
      cudaMalloc(&idx, sizeof(unsigned int));
    }
    cudaMemset(idx, 0, sizeof(unsigned int));
    cudaMemset(buf, 0, static_cast<size_t>(cap) * 16 * sizeof(unsigned long long));
    b200llm::step_trace_set_gemm(buf, idx, cap);
    b200llm::step_trace_set_ops(buf, idx, cap);
    b200llm::step_trace_set_attention(buf, idx, cap);
    b200llm::step_trace_set_fused(buf, idx, cap);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 2;
  }
  unsigned int n = 0;
  cudaMemcpy(&n, idx, sizeof(n), cudaMemcpyDeviceToHost);
  if (static_cast<int>(n) > capacity) n = capacity;
  if (static_cast<int>(n) > cap) n = cap;
  cudaMemcpy(out, buf, static_cast<size_t>(n) * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  return static_cast<int>(n);
}
#endif

// ---- operator-level entry points ----------------------------------------------------------
int b200llm_op_gemm(const void* act, const void* w, const float* bias, void* out, int32_t tokens,
                    int32_t features, int32_t k, int32_t ld_out, int32_t epilogue, int32_t split_k,
                    int32_t orientation, void* stream) {
  GemmArgs g = {};
  g.act = static_cast<const __nv_bfloat16*>(act);
  g.w = static_cast<const __nv_bfloat16*>(w);
  g.bias = bias;
  g.out = out;
  g.tokens = tokens;
  g.features = features;
  g.k = k;
  g.ld_out = ld_out;
  g.epilogue = epilogue;
  g.split_k = split_k > 0 ? split_k : 1;
  g.cluster_split = split_k < 0 ? -split_k : 0;  // negative split_k selects the cluster (DSMEM) reduction
  g.act_rows_alloc = tokens;
  g.force_orientation = orientation == 2 ? 0 : orientation;
  g.use_pair = orientation == 2 ? 1 : (orientation == 0 ? 0 : -1);  // 2: CTA pairs, 0: single-CTA tiles
  if (orientation == 2 && epilogue == kEpiBiasResidF32 && (split_k > 1 || split_k == -1)) {
    g.pair_split = split_k;  // K splits of a pair tile (reduce-add): a count, or -1 = by the wave model
    g.split_k = 1;
    g.cluster_split = 0;
  }
  g.pdl = 0;
  return gemm_bf16(g, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_layernorm(float* x, const float* partials, int32_t n_partials, const float* add_bias,
                         const float* gamma, const float* beta, void* out, int32_t rows, int32_t d,
                         float eps, void* stream) {
  B200_CHECK_ARG(x && gamma && beta && out, "op_layernorm: null argument");
  LayerNormArgs a = {};
  a.x = x;
  a.partials = partials;
  a.n_partials = n_partials;
  a.partial_rows = rows;
  a.add_bias = add_bias;
  a.gamma = gamma;
  a.beta = beta;
  a.row_index = nullptr;
  a.out = static_cast<__nv_bfloat16*>(out);
  a.rows = rows;
  a.d = d;
  a.eps = eps;
  a.pdl = 0;
  return layernorm(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_embed(const float* wte, const float* wpe, const int32_t* ids, const int32_t* pos,
                     float* x, int32_t n_tok, int32_t d, int32_t vocab, int32_t n_positions,
                     void* stream) {
  B200_CHECK_ARG(wte && wpe && ids && pos && x, "op_embed: null argument");
  return embed_tokens(wte, wpe, ids, pos, x, n_tok, d, vocab, n_positions,
                      reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_prefill_attention(const void* qkv, const int32_t* cu_seqlens_dev, int32_t batch,
                                 int32_t max_seq_len, int32_t n_head, void* out, void* stream) {
  B200_CHECK_ARG(qkv && cu_seqlens_dev && out, "op_prefill_attention: null argument");
  return prefill_attention(static_cast<const __nv_bfloat16*>(qkv), cu_seqlens_dev, batch, max_seq_len,
                           n_head, static_cast<__nv_bfloat16*>(out),
                           reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_decode_attention(const void* qkv, const int32_t* seq_len, const int32_t* page_table,
                                int32_t max_pages_per_seq, void* k_pool, void* v_pool, void* out,
                                int32_t batch, int32_t n_head, int32_t n_splits, float* scratch,
                                void* stream) {
  B200_CHECK_ARG(qkv && seq_len && page_table && k_pool && v_pool && out,
                 "op_decode_attention: null argument");
  B200_CHECK_ARG(n_splits >= -8 && n_splits != 0 && n_splits != -1,
                 "op_decode_attention: n_splits must be >= 1, or -2..-8 for the in-cluster merge");
  DecodeAttnArgs a = {};
  a.qkv = static_cast<const __nv_bfloat16*>(qkv);
  a.qkv_partials = nullptr;
  a.n_qkv_partials = 0;
  a.partial_rows = 0;
  a.qkv_bias = nullptr;
  a.active_slots = nullptr;
  a.seq_len = seq_len;
  a.finished = nullptr;
  a.page_table = page_table;
  a.max_pages_per_seq = max_pages_per_seq;
  a.k_pool = static_cast<__nv_bfloat16*>(k_pool);
  a.v_pool = static_cast<__nv_bfloat16*>(v_pool);
  a.out = static_cast<__nv_bfloat16*>(out);
  a.batch = batch;
  a.n_head = n_head;
  a.n_splits = n_splits < 0 ? -n_splits : n_splits;
  a.cluster_combine = n_splits < 0 ? 1 : 0;
  a.scratch = scratch;
  a.pdl = 0;
  return decode_attention(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_decode_attention_stream(const void* qkv, const int32_t* seq_len, const int32_t* page_table,
                                       int32_t max_pages_per_seq, void* k_pool, void* v_pool, int32_t n_pages,
                                       void* out, int32_t batch, int32_t n_head, int32_t n_splits,
                                       float* scratch, int32_t* split_counters, void* stream) {
  B200_CHECK_ARG(qkv && seq_len && page_table && k_pool && v_pool && out,
                 "op_decode_attention_stream: null argument");
  B200_CHECK_ARG(n_splits >= 1 && n_pages > 0, "op_decode_attention_stream: n_splits and n_pages must be positive");
  DecodeAttnArgs a = {};
  a.qkv = static_cast<const __nv_bfloat16*>(qkv);
  a.seq_len = seq_len;
  a.page_table = page_table;
  a.max_pages_per_seq = max_pages_per_seq;
  a.k_pool = static_cast<__nv_bfloat16*>(k_pool);
  a.v_pool = static_cast<__nv_bfloat16*>(v_pool);
  a.out = static_cast<__nv_bfloat16*>(out);
  a.batch = batch;
  a.n_head = n_head;
  a.n_splits = n_splits;
  a.scratch = scratch;
  a.stream_kernel = 1;
  a.n_pages = n_pages;
  a.split_counters = split_counters;
  return decode_attention(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_decode_attention_noalloc(const void* qkv, const int32_t* seq_len,
                                        const int32_t* page_table, int32_t max_pages_per_seq,
                                        void* k_pool, void* v_pool, void* out, int32_t batch,
                                        int32_t n_head, int32_t n_splits, void* stream) {
  B200_CHECK_ARG(qkv && seq_len && page_table && k_pool && v_pool && out,
                 "op_decode_attention_noalloc: null argument");
  DecodeAttnArgs a = {};
  a.qkv = static_cast<const __nv_bfloat16*>(qkv);
  a.seq_len = seq_len;
  a.page_table = page_table;
  a.max_pages_per_seq = max_pages_per_seq;
  a.k_pool = static_cast<__nv_bfloat16*>(k_pool);
  a.v_pool = static_cast<__nv_bfloat16*>(v_pool);
  a.out = static_cast<__nv_bfloat16*>(out);
  a.batch = batch;
  a.n_head = n_head;
  a.n_splits = n_splits;
  return decode_attention_noalloc(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_lm_head_argmax(const void* act, const void* wte, int32_t tokens, int32_t vocab, int32_t k,
                              int32_t eos_id, float* logits, float* amax_val, int32_t* amax_idx,
                              float* eos_logit, int32_t ld_amax, void* stream) {
  B200_CHECK_ARG(act && wte && amax_val && amax_idx && eos_logit, "op_lm_head_argmax: null argument");
  B200_CHECK_ARG(tokens > 0 && tokens <= 256 && ld_amax >= tokens, "op_lm_head_argmax: 1..256 tokens, ld_amax >= tokens");
  GemmArgs g = {};
  g.act = static_cast<const __nv_bfloat16*>(act);
  g.w = static_cast<const __nv_bfloat16*>(wte);
  g.bias = nullptr;
  g.out = logits;
  g.tokens = tokens;
  g.features = vocab;
  g.k = k;
  g.ld_out = vocab;
  g.epilogue = kEpiLogitsArgmax;
  g.split_k = 1;
  g.act_rows_alloc = tokens;
  g.force_orientation = 1;
  g.pdl = 0;
  g.amax_groups_alloc = gemm_argmax_groups(vocab, tokens);
  g.amax_val = amax_val;
  g.amax_idx = amax_idx;
  g.eos_logit = eos_logit;
  g.ld_amax = ld_amax;
  g.eos_id = eos_id;
  return gemm_bf16(g, reinterpret_cast<cudaStream_t>(stream));
}

int32_t b200llm_lm_head_argmax_groups(int32_t vocab, int32_t tokens) { return gemm_argmax_groups(vocab, tokens); }

int b200llm_op_prefill_attention_paged(const void* qkv, int32_t n_rows, const int32_t* cu_seqlens,
                                       const int32_t* prefix_len, const int32_t* slots, int32_t batch,
                                       int32_t max_suffix_len, int32_t n_head, const int32_t* page_table,
                                       int32_t max_pages_per_seq, const void* k_pool, const void* v_pool,
                                       int32_t n_pages, void* out, int32_t impl, void* stream) {
  B200_CHECK_ARG(qkv && cu_seqlens && prefix_len && slots && page_table && k_pool && v_pool && out,
                 "op_prefill_attention_paged: null argument");
  PagedPrefillArgs a;
  a.qkv = static_cast<const __nv_bfloat16*>(qkv);
  a.cu_seqlens = cu_seqlens;
  a.prefix_len = prefix_len;
  a.slots = slots;
  a.batch = batch;
  a.max_suffix_len = max_suffix_len;
  a.n_head = n_head;
  a.page_table = page_table;
  a.max_pages_per_seq = max_pages_per_seq;
  a.k_pool = static_cast<const __nv_bfloat16*>(k_pool);
  a.v_pool = static_cast<const __nv_bfloat16*>(v_pool);
  a.out = static_cast<__nv_bfloat16*>(out);
  a.qkv_rows = n_rows;
  a.n_pages = n_pages;
  return impl == 1 ? prefill_attention_paged(a, reinterpret_cast<cudaStream_t>(stream))
                   : prefill_attention_tc(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_kv_scatter(const void* qkv, const int32_t* row_pos, const int32_t* row_slot, int32_t n_tok,
                          int32_t n_head, const int32_t* page_table, int32_t max_pages_per_seq, void* k_pool,
                          void* v_pool, void* stream) {
  B200_CHECK_ARG(qkv && row_pos && row_slot && page_table && k_pool && v_pool, "op_kv_scatter: null argument");
  return kv_scatter(static_cast<const __nv_bfloat16*>(qkv), row_pos, row_slot, n_tok, n_head, page_table,
                    max_pages_per_seq, static_cast<__nv_bfloat16*>(k_pool), static_cast<__nv_bfloat16*>(v_pool),
                    reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_ln_gemm(const float* x, const float* gamma, const float* beta, float eps, const void* w,
                       const float* bias, void* out, int32_t rows, int32_t features, int32_t d, int32_t ld_out,
                       int32_t gelu, int32_t split, float* partial_out, void* stream) {
  LnGemmArgs a;
  a.x = x;
  a.gamma = gamma;
  a.beta = beta;
  a.eps = eps;
  a.w = static_cast<const __nv_bfloat16*>(w);
  a.bias = bias;
  a.out = static_cast<__nv_bfloat16*>(out);
  a.rows = rows;
  a.d = d;
  a.features = features;
  a.ld_out = ld_out;
  a.gelu = gelu;
  a.split = split;
  a.partial_out = partial_out;
  a.plane_rows = rows;
  return ln_gemm(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_gemm_redadd(const void* act, const void* w, const float* bias, float* x, int32_t rows,
                           int32_t features, int32_t k, int32_t split, void* stream) {
  GemmRedAddArgs a;
  a.act = static_cast<const __nv_bfloat16*>(act);
  a.w = static_cast<const __nv_bfloat16*>(w);
  a.bias = bias;
  a.x = x;
  a.rows = rows;
  a.features = features;
  a.k = k;
  int sms = 148;
  if (split <= 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  a.split = split > 0 ? split : gemm_redadd_pick_split(features, k, rows, sms);
  a.act_rows_alloc = rows <= 16 ? 16 : rows <= 32 ? 32 : 64;  // the caller's buffer holds the whole token tile
  return gemm_redadd(a, reinterpret_cast<cudaStream_t>(stream));
}

size_t b200llm_decode_attention_scratch_floats(int32_t batch, int32_t n_head, int32_t n_splits) {
  return decode_attention_scratch_floats(batch, n_head, n_splits);
}

int b200llm_op_argmax(const float* logits, int32_t rows, int32_t ld, int32_t vocab, int32_t* out,
                      void* stream) {
  B200_CHECK_ARG(logits && out, "op_argmax: null argument");
  return argmax_rows(logits, rows, ld, vocab, out, reinterpret_cast<cudaStream_t>(stream));
}

int b200llm_op_convert_transpose(const float* src, void* dst, int32_t rows, int32_t cols,
                                 void* stream) {
  B200_CHECK_ARG(src && dst, "op_convert_transpose: null argument");
  return convert_transpose_bf16(src, static_cast<__nv_bfloat16*>(dst), rows, cols,
                                reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
