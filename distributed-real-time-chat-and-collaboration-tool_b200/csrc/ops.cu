// Memory-bound kernels of the GPT-2 step (see ops.cuh for the oracle citations).
// All are HBM/L2 streaming kernels: 16-byte vector accesses, one warp per activation row,
// warp-shuffle reductions, no shared-memory staging unless a cross-warp reduction needs it.
#include "ops.cuh"

#include "common.cuh"

namespace b200llm {

namespace {

// ---------------------------------------------------------------------------------------
// Embedding
// ---------------------------------------------------------------------------------------
__global__ void embed_kernel(const float* __restrict__ wte, const float* __restrict__ wpe,
                             const int32_t* __restrict__ ids, const int32_t* __restrict__ pos,
                             const int32_t* __restrict__ active_slots,
                             const int32_t* __restrict__ last_tok,
                             const int32_t* __restrict__ seq_len, float* __restrict__ x,
                             int n_rows, int d, int vocab, int n_positions) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x;
  if (row >= n_rows) return;
  int tok, p;
  if (active_slots != nullptr) {
    const int slot = active_slots[row];
    if (slot < 0) {
      tok = 0;
      p = 0;
    } else {
      tok = last_tok[slot];
      p = seq_len[slot];
    }
  } else {
    tok = ids[row];
    p = pos[row];
  }
  // Clamp instead of faulting on a corrupt id; the host validates ids up front.
  tok = min(max(tok, 0), vocab - 1);
  p = min(max(p, 0), n_positions - 1);
  const float4* e = reinterpret_cast<const float4*>(wte + static_cast<size_t>(tok) * d);
  const float4* q = reinterpret_cast<const float4*>(wpe + static_cast<size_t>(p) * d);
  float4* o = reinterpret_cast<float4*>(x + static_cast<size_t>(row) * d);
  for (int i = threadIdx.x; i < d / 4; i += blockDim.x) {
    const float4 a = __ldg(e + i);
    const float4 b = __ldg(q + i);
    o[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
}

// ---------------------------------------------------------------------------------------
// LayerNorm (+ residual / split-K reduction)
// ---------------------------------------------------------------------------------------
// One CTA per row, one float4 column per thread (d/4 threads). With split-K partials every
// thread has `n_partials` independent 16-byte loads in flight, so the reduction is bound by
// L2 bandwidth instead of one warp's load latency; the fixed summation order keeps it
// deterministic.
__device__ __forceinline__ float2 block_sum2(float a, float b, float2* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = (blockDim.x + 31) >> 5;
  if (lane == 0) red[warp] = make_float2(a, b);
  __syncthreads();
  float2 t = (lane < nwarps) ? red[lane] : make_float2(0.f, 0.f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    t.x += __shfl_xor_sync(0xffffffffu, t.x, o);
    t.y += __shfl_xor_sync(0xffffffffu, t.y, o);
  }
  return t;
}

__global__ void __launch_bounds__(512)
layernorm_kernel(float* __restrict__ x, const float* __restrict__ partials, int n_partials,
                 const float* __restrict__ add_bias, const float* __restrict__ gamma,
                 const float* __restrict__ beta, const int32_t* __restrict__ row_index,
                 __nv_bfloat16* __restrict__ out, int d, int rows_total, float eps) {
  __shared__ float2 red2[32];
  B200_STEP_BEGIN(1);
  // This kernel reads nothing mutable before its dependency wait, so the next kernel (a GEMM
  // that starts by prefetching weights) may be launched as soon as this one is resident.
  // (embed / attention / select signal only AFTER their wait: they are the ordering fences the
  // attention kernel's early KV reads rely on, see attention.cu.)
  pdl_launch_dependents();
  const int orow = blockIdx.x;
  const int c = threadIdx.x;  // float4 column
  const int ncol = d >> 2;
  const bool active = c < ncol;
  // parameters do not depend on the upstream kernel: fetch them before the dependency wait
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f), b = g, ab = g;
  if (active) {
    g = __ldg(reinterpret_cast<const float4*>(gamma) + c);
    b = __ldg(reinterpret_cast<const float4*>(beta) + c);
    if (n_partials > 0 && add_bias != nullptr) ab = __ldg(reinterpret_cast<const float4*>(add_bias) + c);
  }
  pdl_wait();
  const int irow = row_index ? row_index[orow] : orow;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  float4* xr = reinterpret_cast<float4*>(x + static_cast<size_t>(irow) * d);
  if (active) {
    v = xr[c];
    if (n_partials > 0) {
      float4 acc = ab;
      const float4* p = reinterpret_cast<const float4*>(partials + static_cast<size_t>(irow) * d) + c;
      const size_t plane = static_cast<size_t>(rows_total) * d / 4;
      int s = 0;
      for (; s + 4 <= n_partials; s += 4) {
        const float4 p0 = p[(s + 0) * plane], p1 = p[(s + 1) * plane], p2 = p[(s + 2) * plane],
                     p3 = p[(s + 3) * plane];
        acc.x += p0.x; acc.y += p0.y; acc.z += p0.z; acc.w += p0.w;
        acc.x += p1.x; acc.y += p1.y; acc.z += p1.z; acc.w += p1.w;
        acc.x += p2.x; acc.y += p2.y; acc.z += p2.z; acc.w += p2.w;
        acc.x += p3.x; acc.y += p3.y; acc.z += p3.z; acc.w += p3.w;
      }
      for (; s < n_partials; ++s) {
        const float4 p0 = p[s * plane];
        acc.x += p0.x; acc.y += p0.y; acc.z += p0.z; acc.w += p0.w;
      }
      v.x += acc.x; v.y += acc.y; v.z += acc.z; v.w += acc.w;
      xr[c] = v;
    }
  }
  // One block reduction for both moments. The second moment is taken about the row's first
  // element (every thread reads the same word), so a large common offset cannot cancel
  // catastrophically in E[(x-c)^2] - (E[x-c])^2.
  const float inv_d = 1.0f / static_cast<float>(d);
  __shared__ float shift_s;
  if (threadIdx.x == 0) shift_s = v.x;
  __syncthreads();
  const float shift = shift_s;
  float s1 = 0.f, s2 = 0.f;
  if (active) {
    const float a0 = v.x - shift, a1 = v.y - shift, a2 = v.z - shift, a3 = v.w - shift;
    s1 = (a0 + a1) + (a2 + a3);
    s2 = (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
  }
  const float2 tot = block_sum2(s1, s2, red2);
  const float dm = tot.x * inv_d;  // mean - shift
  const float mean = shift + dm;
  const float var = fmaxf(tot.y * inv_d - dm * dm, 0.f);
  const float rstd = 1.0f / sqrtf(var + eps);
  if (active) {
    uint2 pk;
    pk.x = pack_bf16x2((v.x - mean) * rstd * g.x + b.x, (v.y - mean) * rstd * g.y + b.y);
    pk.y = pack_bf16x2((v.z - mean) * rstd * g.z + b.z, (v.w - mean) * rstd * g.w + b.w);
    reinterpret_cast<uint2*>(out + static_cast<size_t>(orow) * d)[c] = pk;
  }
  B200_STEP_END();
}

// Many rows, no partial planes (prefill): one warp per row, 8 rows per CTA, the whole row in
// registers (VEC float4 per lane) -- one CTA per row is launch-bound at thousands of rows.
template <int VEC>  // d = VEC * 128
__global__ void __launch_bounds__(256)
layernorm_rows_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                      const float* __restrict__ beta, const int32_t* __restrict__ row_index,
                      __nv_bfloat16* __restrict__ out, int rows, float eps) {
  constexpr int D = VEC * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int orow = blockIdx.x * 8 + warp;
  pdl_wait();
  pdl_launch_dependents();
  if (orow >= rows) return;
  const int irow = row_index ? row_index[orow] : orow;
  const float4* xr = reinterpret_cast<const float4*>(x + static_cast<size_t>(irow) * D);
  float4 v[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) v[i] = xr[lane + 32 * i];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(sum) * (1.0f / D);
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const float a0 = v[i].x - mean, a1 = v[i].y - mean, a2 = v[i].z - mean, a3 = v[i].w - mean;
    sq += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
  }
  const float rstd = 1.0f / sqrtf(warp_sum(sq) * (1.0f / D) + eps);
  uint2* o = reinterpret_cast<uint2*>(out + static_cast<size_t>(orow) * D);
#pragma unroll
  for (int i = 0; i < VEC; ++i) {
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + lane + 32 * i);
    uint2 pk;
    pk.x = pack_bf16x2((v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y);
    pk.y = pack_bf16x2((v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w);
    o[lane + 32 * i] = pk;
  }
}

// ---------------------------------------------------------------------------------------
// Split-K reduce + bias + gelu_new -> bf16 (one float4 column group per thread)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
reduce_bias_gelu_kernel(const float* __restrict__ partials, int n_partials, size_t plane_f4,
                        const float* __restrict__ bias, __nv_bfloat16* __restrict__ out, int rows,
                        int cols4) {
  B200_STEP_BEGIN(2);
  pdl_launch_dependents();  // nothing mutable is read before the wait (see layernorm_kernel)
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = rows * cols4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const int c4 = idx % cols4;
  if (idx < total) acc = __ldg(reinterpret_cast<const float4*>(bias) + c4);
  pdl_wait();
  if (idx >= total) {
    B200_STEP_END();
    return;
  }
  const float4* p = reinterpret_cast<const float4*>(partials) + idx;
  int s = 0;
  for (; s + 2 <= n_partials; s += 2) {
    const float4 p0 = p[static_cast<size_t>(s) * plane_f4];
    const float4 p1 = p[static_cast<size_t>(s + 1) * plane_f4];
    acc.x += p0.x; acc.y += p0.y; acc.z += p0.z; acc.w += p0.w;
    acc.x += p1.x; acc.y += p1.y; acc.z += p1.z; acc.w += p1.w;
  }
  for (; s < n_partials; ++s) {
    const float4 p0 = p[static_cast<size_t>(s) * plane_f4];
    acc.x += p0.x; acc.y += p0.y; acc.z += p0.z; acc.w += p0.w;
  }
  uint2 pk;
  pk.x = pack_bf16x2(gelu_new(acc.x), gelu_new(acc.y));
  pk.y = pack_bf16x2(gelu_new(acc.z), gelu_new(acc.w));
  reinterpret_cast<uint2*>(out)[idx] = pk;
  B200_STEP_END();
}

// ---------------------------------------------------------------------------------------
// Token selection
// ---------------------------------------------------------------------------------------
struct MaxIdx {
  float v;
  int i;
};

__device__ __forceinline__ MaxIdx better(MaxIdx a, MaxIdx b) {
  // larger value wins; on equal values the lower index wins (torch.argmax on ties)
  if (b.v > a.v || (b.v == a.v && b.i < a.i)) return b;
  return a;
}

__device__ __forceinline__ MaxIdx block_argmax(MaxIdx m, MaxIdx* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxIdx other;
    other.v = __shfl_xor_sync(0xffffffffu, m.v, o);
    other.i = __shfl_xor_sync(0xffffffffu, m.i, o);
    m = better(m, other);
  }
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  __syncthreads();
  if (lane == 0) smem[warp] = m;
  __syncthreads();
  if (warp == 0) {
    MaxIdx t;
    t.v = -INFINITY;
    t.i = 0x7fffffff;
    if (lane < nwarps) t = smem[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      MaxIdx other;
      other.v = __shfl_xor_sync(0xffffffffu, t.v, o);
      other.i = __shfl_xor_sync(0xffffffffu, t.i, o);
      t = better(t, other);
    }
    if (lane == 0) smem[0] = t;
  }
  __syncthreads();
  return smem[0];
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

constexpr int kMaxTopK = 64;

// ---------------------------------------------------------------------------------------
// Top-k of one logits row by radix select (sampling path).
// k successive arg-max passes over the 50k-entry row cost ~14 us each; here the k-th largest
// value is located with histogram passes over an order-preserving 32-bit key (11 + 11 + 10 bits;
// normally the first pass leaves a short list that is finished in shared memory), a last pass
// collects everything above it plus the lowest-index ties, and the <= 64
// survivors are ranked by (value desc, index asc) -- the same order the pass-per-entry version
// produced, so the draw that follows is bit-identical.
// ---------------------------------------------------------------------------------------
constexpr int kTopkBins = 2048;
constexpr int kTopkEqCap = 4096;  // also the capacity of the short list after the first pass

struct TopkSmem {
  unsigned int hist[kTopkBins];
  int eq_idx[kTopkEqCap];             // indices of the ties (long path) / of the short list
  unsigned int list_key[kTopkEqCap];  // keys of the short list
  float cand_v[kMaxTopK];
  int cand_i[kMaxTopK];
  unsigned int n_cand, n_eq;
  unsigned int sel_bin, sel_above;
};

__device__ __forceinline__ unsigned int topk_key(float v) {
  unsigned int b = __float_as_uint(v);
  if (b == 0x80000000u) b = 0u;  // -0 == +0
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Warp 0: walk the histogram from the highest bin down and find the bin in which the running
// count reaches `need`; publishes the bin and the count strictly above it.
__device__ __forceinline__ void topk_find_bin(TopkSmem& sm, int nbins, unsigned int need) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const int per = nbins / 32;
    const int hi = nbins - 1 - lane * per;  // this lane's segment: bins hi, hi-1, ..., hi-per+1
    unsigned int seg = 0;
    for (int j = 0; j < per; ++j) seg += sm.hist[hi - j];
    unsigned int incl = seg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned int up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    const unsigned int excl = incl - seg;
    if (excl < need && incl >= need) {  // exactly one lane (need >= 1 and need <= total)
      unsigned int run = excl;
      for (int j = 0; j < per; ++j) {
        const unsigned int c = sm.hist[hi - j];
        if (run + c >= need) {
          sm.sel_bin = static_cast<unsigned int>(hi - j);
          sm.sel_above = run;
          break;
        }
        run += c;
      }
    }
  }
}

// All threads of the CTA. Returns false (nothing written) when more than kTopkEqCap entries tie
// with the k-th value; the caller then uses the pass-per-entry selection. On success top_v/top_i
// hold the k entries in (value desc, index asc) order.
__device__ bool topk_radix(const float* __restrict__ lg, int vocab, int skip_id, int k, TopkSmem& sm, float* top_v,
                           int* top_i) {
  unsigned int prefix = 0, prefix_mask = 0;
  unsigned int need = static_cast<unsigned int>(k);
  bool short_list = false;
  const int shifts[3] = {21, 10, 0};
  const int widths[3] = {11, 11, 10};
#pragma unroll
  for (int pass = 0; pass < 3; ++pass) {
    const int nbins = 1 << widths[pass];
    if (pass == 1 && short_list) break;  // CTA-uniform
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) sm.hist[i] = 0u;
    __syncthreads();
    // eight independent loads in flight per thread: a load -> atomic loop runs at one L2 round trip
    // per element
    for (int i0 = threadIdx.x; i0 < vocab; i0 += 8 * blockDim.x) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * blockDim.x;
        v[u] = (i < vocab) ? __ldcg(lg + i) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * blockDim.x;
        if (i < vocab && i != skip_id) {
          const unsigned int key = topk_key(v[u]);
          if ((key & prefix_mask) == prefix) atomicAdd(&sm.hist[(key >> shifts[pass]) & (nbins - 1)], 1u);
        }
      }
    }
    __syncthreads();
    topk_find_bin(sm, nbins, need);
    __syncthreads();
    prefix |= sm.sel_bin << shifts[pass];
    prefix_mask |= static_cast<unsigned int>(nbins - 1) << shifts[pass];
    need -= sm.sel_above;
    if (pass == 0) short_list = sm.hist[sm.sel_bin] <= static_cast<unsigned int>(kTopkEqCap);
    __syncthreads();  // sel_* and hist are rewritten by the next pass
  }
  if (short_list) {
    // Usual case: few entries share the leading 11 key bits of the k-th largest one. One more pass
    // over the row moves everything above that bin straight to the survivors and the bin itself
    // to a short list in shared memory, where the remaining `need` entries are found by rank.
    const unsigned int bin0 = prefix >> 21;
    if (threadIdx.x == 0) {
      sm.n_cand = 0u;
      sm.n_eq = 0u;
    }
    __syncthreads();
    for (int i0 = threadIdx.x; i0 < vocab; i0 += 8 * blockDim.x) {
      float vv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * blockDim.x;
        vv[u] = (i < vocab) ? __ldcg(lg + i) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * blockDim.x;
        if (i >= vocab || i == skip_id) continue;
        const unsigned int key = topk_key(vv[u]);
        const unsigned int top = key >> 21;
        if (top > bin0) {
          const unsigned int p = atomicAdd(&sm.n_cand, 1u);
          if (p < static_cast<unsigned int>(kMaxTopK)) {
            sm.cand_v[p] = vv[u];
            sm.cand_i[p] = i;
          }
        } else if (top == bin0) {
          const unsigned int p = atomicAdd(&sm.n_eq, 1u);
          if (p < static_cast<unsigned int>(kTopkEqCap)) {
            sm.list_key[p] = key;
            sm.eq_idx[p] = i;
          }
        }
      }
    }
    __syncthreads();
    const unsigned int n_list = sm.n_eq;
    const unsigned int n_gt = sm.n_cand;  // == k - need
    for (unsigned int t = threadIdx.x; t < n_list; t += blockDim.x) {
      const unsigned int key = sm.list_key[t];
      const int idx = sm.eq_idx[t];
      unsigned int rank = 0;
      for (unsigned int j = 0; j < n_list; ++j) {
        const unsigned int kj = sm.list_key[j];
        rank += (kj > key || (kj == key && sm.eq_idx[j] < idx)) ? 1u : 0u;
      }
      if (rank < need) {
        const unsigned int b = (key & 0x80000000u) ? (key & 0x7fffffffu) : ~key;
        sm.cand_v[n_gt + rank] = __uint_as_float(b);
        sm.cand_i[n_gt + rank] = idx;
      }
    }
    __syncthreads();
  } else {
  // prefix is now the key of the k-th largest entry; `need` of the entries equal to it are taken
  const unsigned int thr = prefix;
  if (threadIdx.x == 0) {
    sm.n_cand = 0u;
    sm.n_eq = 0u;
  }
  __syncthreads();
  for (int i0 = threadIdx.x; i0 < vocab; i0 += 8 * blockDim.x) {
    float vv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * blockDim.x;
      vv[u] = (i < vocab) ? __ldcg(lg + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * blockDim.x;
      if (i >= vocab || i == skip_id) continue;
      const float v = vv[u];
      const unsigned int key = topk_key(v);
      if (key > thr) {
        const unsigned int p = atomicAdd(&sm.n_cand, 1u);
        if (p < static_cast<unsigned int>(kMaxTopK)) {
          sm.cand_v[p] = v;
          sm.cand_i[p] = i;
        }
      } else if (key == thr) {
        const unsigned int p = atomicAdd(&sm.n_eq, 1u);
        if (p < static_cast<unsigned int>(kTopkEqCap)) sm.eq_idx[p] = i;
      }
    }
  }
  __syncthreads();
  const unsigned int n_eq = sm.n_eq;
  const unsigned int n_gt = sm.n_cand;  // == k - need
  if (n_eq > static_cast<unsigned int>(kTopkEqCap)) return false;
  // the `need` lowest indices among the ties
  float thr_v = 0.f;
  {
    const unsigned int b = (thr & 0x80000000u) ? (thr & 0x7fffffffu) : ~thr;
    thr_v = __uint_as_float(b);
  }
  for (unsigned int t = threadIdx.x; t < n_eq; t += blockDim.x) {
    const int mine = sm.eq_idx[t];
    unsigned int rank = 0;
    for (unsigned int j = 0; j < n_eq; ++j) rank += sm.eq_idx[j] < mine ? 1u : 0u;
    if (rank < need) {
      sm.cand_v[n_gt + rank] = thr_v;
      sm.cand_i[n_gt + rank] = mine;
    }
  }
  __syncthreads();
  }  // long-list path
  // rank the k survivors
  if (threadIdx.x < static_cast<unsigned int>(k)) {
    const float v = sm.cand_v[threadIdx.x];
    const int idx = sm.cand_i[threadIdx.x];
    int rank = 0;
    for (int j = 0; j < k; ++j) {
      const float vj = sm.cand_v[j];
      const int ij = sm.cand_i[j];
      rank += (vj > v || (vj == v && ij < idx)) ? 1 : 0;
    }
    top_v[rank] = v;
    top_i[rank] = idx;
  }
  __syncthreads();
  return true;
}

__global__ void __launch_bounds__(1024) select_kernel(SelectArgs a) {
  __shared__ MaxIdx red[32];
  __shared__ float top_v[kMaxTopK];
  __shared__ int top_i[kMaxTopK];
  // The sampler's 41 KB of scratch is dynamic shared memory that only launches with sampling rows
  // request: as a static array it kept this kernel's CTAs from becoming resident next to the LM
  // head's (which fill their SMs' shared memory), and the greedy step paid ~8 us for it.
  extern __shared__ __align__(16) unsigned char topk_dyn[];
  TopkSmem& topk_sm = *reinterpret_cast<TopkSmem*>(topk_dyn);
  B200_STEP_BEGIN(3);
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x;
  const int slot = a.active_slots ? a.active_slots[row] : row;
  if (slot < 0) return;
  const float* lg = a.logits + static_cast<size_t>(row) * a.ld;
  const int g = a.n_gen[slot];
  const bool ban_eos = g < a.min_new[slot];  // MinNewTokensLength: EOS logit = -inf
  const int was_finished = a.finished[slot];

  const float temperature = a.temperature ? a.temperature[slot] : 0.0f;
  const int top_k_req = a.top_k ? a.top_k[slot] : 0;
  const uint64_t seed = a.seed ? a.seed[slot] : 0ull;

  int chosen;
  if (temperature <= 0.0f || top_k_req <= 1 || a.logits == nullptr) {
    MaxIdx m;
    m.v = -INFINITY;
    m.i = 0x7fffffff;
    if (a.amax_val != nullptr) {
      // reduce the per-group candidates the LM-head epilogue produced (EOS handled apart). The
      // thread's candidates are fetched with independent loads first: a load -> compare loop pays
      // one L2 round trip per candidate (7 of them for GPT-2's 1572 groups on 256 threads).
      constexpr int kBatch = 8;
      for (int i0 = threadIdx.x; i0 < a.n_groups; i0 += blockDim.x * kBatch) {
        float cv[kBatch];
        int ci[kBatch];
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
          const int i = i0 + j * static_cast<int>(blockDim.x);
          const bool ok = i < a.n_groups;
          cv[j] = ok ? a.amax_val[static_cast<size_t>(i) * a.ld_amax + row] : -INFINITY;
          ci[j] = ok ? a.amax_idx[static_cast<size_t>(i) * a.ld_amax + row] : 0x7fffffff;
        }
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
          MaxIdx c;
          c.v = cv[j];
          c.i = ci[j];
          m = better(m, c);
        }
      }
      m = block_argmax(m, red);
      if (!ban_eos) {
        MaxIdx e;
        e.v = a.eos_logit[row];
        e.i = a.eos_id;
        m = better(m, e);
      }
    } else {
      for (int i = threadIdx.x; i < a.vocab; i += blockDim.x) {
        if (ban_eos && i == a.eos_id) continue;
        MaxIdx c;
        c.v = lg[i];
        c.i = i;
        m = better(m, c);
      }
      m = block_argmax(m, red);
    }
    chosen = m.i;
  } else {
    // top-k by k successive block-wide argmax passes over the (L2-resident) logits row,
    // then a temperature softmax over the k survivors and one inverse-CDF draw.
    int k = top_k_req < kMaxTopK ? top_k_req : kMaxTopK;
    const int n_valid = a.vocab - ((ban_eos && a.eos_id >= 0 && a.eos_id < a.vocab) ? 1 : 0);
    if (k > n_valid) k = n_valid;
    const bool have_topk = topk_radix(lg, a.vocab, ban_eos ? a.eos_id : -1, k, topk_sm, top_v, top_i);
    float prev_v = INFINITY;
    int prev_i = -1;
    for (int r = 0; r < (have_topk ? 0 : k); ++r) {  // fallback: one arg-max pass per entry
      MaxIdx m;
      m.v = -INFINITY;
      m.i = 0x7fffffff;
      for (int i = threadIdx.x; i < a.vocab; i += blockDim.x) {
        if (ban_eos && i == a.eos_id) continue;
        const float v = lg[i];
        // strictly after (prev_v, prev_i) in (value desc, index asc) order
        if (v > prev_v || (v == prev_v && i <= prev_i)) continue;
        MaxIdx c;
        c.v = v;
        c.i = i;
        m = better(m, c);
      }
      m = block_argmax(m, red);
      prev_v = m.v;
      prev_i = m.i;
      if (threadIdx.x == 0) {
        top_v[r] = m.v;
        top_i[r] = m.i;
      }
    }
    __syncthreads();
    chosen = top_i[0];
    if (threadIdx.x == 0) {
      const float inv_t = 1.0f / temperature;
      const float mx = top_v[0];
      float z = 0.f;
      for (int r = 0; r < k; ++r) z += expf((top_v[r] - mx) * inv_t);
      const uint64_t h = splitmix64(seed ^ splitmix64(static_cast<uint64_t>(slot) << 32 | static_cast<uint32_t>(g)));
      const float u = static_cast<float>(h >> 40) * (1.0f / 16777216.0f) * z;
      float c = 0.f;
      int pick = top_i[k - 1];
      for (int r = 0; r < k; ++r) {
        c += expf((top_v[r] - mx) * inv_t);
        if (u < c) {
          pick = top_i[r];
          break;
        }
      }
      red[1].i = pick;
    }
    __syncthreads();
    chosen = red[1].i;
  }

  if (threadIdx.x == 0) {
    const int tok = was_finished ? a.pad_id : chosen;
    const int cap = a.max_new[slot];
    if (g < cap) {
      a.out_tokens[static_cast<size_t>(slot) * a.out_stride + g] = tok;
      a.n_gen[slot] = g + 1;
    }
    a.last_tok[slot] = tok;
    if (!was_finished && (tok == a.eos_id || g + 1 >= cap)) a.finished[slot] = 1;
    if (a.advance_seq && !was_finished) a.seq_len[slot] = a.seq_len[slot] + 1;
    B200_STEP_END();
  }
  pdl_launch_dependents();
}

__global__ void __launch_bounds__(1024)
argmax_rows_kernel(const float* __restrict__ logits, int ld, int vocab, int32_t* __restrict__ out) {
  __shared__ MaxIdx red[32];
  const float* lg = logits + static_cast<size_t>(blockIdx.x) * ld;
  MaxIdx m;
  m.v = -INFINITY;
  m.i = 0x7fffffff;
  for (int i = threadIdx.x; i < vocab; i += blockDim.x) {
    MaxIdx c;
    c.v = lg[i];
    c.i = i;
    m = better(m, c);
  }
  m = block_argmax(m, red);
  if (threadIdx.x == 0) out[blockIdx.x] = m.i;
}

// ---------------------------------------------------------------------------------------
// Weight conversion
// ---------------------------------------------------------------------------------------
__global__ void convert_transpose_kernel(const float* __restrict__ src,
                                         __nv_bfloat16* __restrict__ dst, int rows, int cols) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    tile[j][threadIdx.x] = (r < rows && c < cols) ? src[static_cast<size_t>(r) * cols + c] : 0.f;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;  // dst[c][r]
    if (c < cols && r < rows) dst[static_cast<size_t>(c) * rows + r] = __float2bfloat16(tile[threadIdx.x][j]);
  }
}

__global__ void convert_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                               size_t n) {
  size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < n; i += stride) dst[i] = __float2bfloat16(src[i]);
}

template <typename Kernel, typename... Args>
int launch_pdl(Kernel kernel, dim3 grid, dim3 block, cudaStream_t stream, int pdl, Args... args) {
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
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, args...));
  return kOk;
}

}  // namespace

B200_DEFINE_STEP_TRACE_SETTER(step_trace_set_ops)

int embed_tokens(const float* wte, const float* wpe, const int32_t* ids, const int32_t* pos,
                 float* x, int n_tok, int d, int vocab, int n_positions, cudaStream_t stream) {
  B200_CHECK_ARG(n_tok > 0 && d % 4 == 0, "embed: bad shape");
  embed_kernel<<<n_tok, 128, 0, stream>>>(wte, wpe, ids, pos, nullptr, nullptr, nullptr, x, n_tok,
                                          d, vocab, n_positions);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

int embed_decode(const float* wte, const float* wpe, const int32_t* active_slots,
                 const int32_t* last_tok, const int32_t* seq_len, float* x, int batch, int d,
                 int vocab, int n_positions, cudaStream_t stream, int pdl) {
  B200_CHECK_ARG(batch > 0 && d % 4 == 0, "embed: bad shape");
  const int32_t* no_ids = nullptr;
  return launch_pdl(embed_kernel, dim3(batch), dim3(128), stream, pdl, wte, wpe, no_ids, no_ids,
                    active_slots, last_tok, seq_len, x, batch, d, vocab, n_positions);
}

int layernorm(const LayerNormArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.rows > 0 && a.d % 4 == 0 && a.d / 4 <= 512,
                 "layernorm: d must be a multiple of 4 and at most 2048");
  if (a.n_partials == 0 && a.rows >= 512 && a.d % 128 == 0) {
    const dim3 grid((a.rows + 7) / 8), block(256);
    const float* xc = a.x;
#define B200_LN_ROWS(V)                                                                          \
  case V:                                                                                        \
    return launch_pdl(layernorm_rows_kernel<V>, grid, block, stream, a.pdl, xc, a.gamma, a.beta, \
                      a.row_index, a.out, a.rows, a.eps);
    switch (a.d / 128) {
      B200_LN_ROWS(1)
      B200_LN_ROWS(2)
      B200_LN_ROWS(4)
      B200_LN_ROWS(6)
      B200_LN_ROWS(8)
      B200_LN_ROWS(10)
      default:
        break;  // fall through to the generic kernel
    }
#undef B200_LN_ROWS
  }
  const int threads = ((a.d / 4 + 31) / 32) * 32;
  const int rows_total = a.partial_rows > 0 ? a.partial_rows : a.rows;
  return launch_pdl(layernorm_kernel, dim3(a.rows), dim3(threads), stream, a.pdl, a.x, a.partials,
                    a.n_partials, a.add_bias, a.gamma, a.beta, a.row_index, a.out, a.d, rows_total,
                    a.eps);
}

int reduce_bias_gelu(const float* partials, int n_partials, int partial_rows, const float* bias,
                     __nv_bfloat16* out, int rows, int cols, int pdl, cudaStream_t stream) {
  B200_CHECK_ARG(partials && bias && out && rows > 0 && cols % 4 == 0 && n_partials > 0,
                 "reduce_bias_gelu: bad arguments");
  const int total = rows * (cols / 4);
  const size_t plane_f4 = static_cast<size_t>(partial_rows > 0 ? partial_rows : rows) * cols / 4;
  return launch_pdl(reduce_bias_gelu_kernel, dim3((total + 255) / 256), dim3(256), stream, pdl,
                    partials, n_partials, plane_f4, bias, out, rows, cols / 4);
}

int select_tokens(const SelectArgs& a, cudaStream_t stream) {
  B200_CHECK_ARG(a.batch > 0 && a.vocab > 0, "select: bad shape");
  // all-greedy steps only reduce the LM head's group candidates: a small CTA has the shorter
  // reduction tree; sampling scans the whole logits row and wants every thread it can get
  const bool greedy_only = a.logits == nullptr && a.amax_val != nullptr;
  const int threads = greedy_only ? 256 : 1024;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a.batch);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = greedy_only ? 0 : sizeof(TopkSmem);
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = a.pdl ? 1 : 0;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, select_kernel, a));
  return kOk;
}

int argmax_rows(const float* logits, int rows, int ld, int vocab, int32_t* out,
                cudaStream_t stream) {
  B200_CHECK_ARG(rows > 0 && vocab > 0, "argmax: bad shape");
  argmax_rows_kernel<<<rows, 1024, 0, stream>>>(logits, ld, vocab, out);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

int convert_transpose_bf16(const float* src, __nv_bfloat16* dst, int rows, int cols,
                           cudaStream_t stream) {
  dim3 grid((cols + 31) / 32, (rows + 31) / 32), block(32, 8);
  convert_transpose_kernel<<<grid, block, 0, stream>>>(src, dst, rows, cols);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

int convert_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t stream) {
  convert_kernel<<<1024, 256, 0, stream>>>(src, dst, n);
  B200_CUDA_OK(cudaGetLastError());
  return kOk;
}

}  // namespace b200llm
