// Profiling tool: decode attention at the BASELINE config-2 shape (64 seqs x 12 heads,
// ctx ~144, scattered pages), cold L2, CTA-0 milestone trace + event timings.
#include <cstdio>
#include <string>
#include <vector>
#include <numeric>
#include <algorithm>
#include <random>

#include "attention.cu"

namespace b200llm {
static thread_local std::string g_err;
void set_last_error(const std::string& m) { g_err = m; }
int fail(int code, const std::string& m) { g_err = m; return code; }
}  // namespace b200llm
using namespace b200llm;

int main(int argc, char** argv) {
  const int B = 64, H = 12, d = H * 64, ctx = argc > 1 ? atoi(argv[1]) : 144, splits = argc > 2 ? atoi(argv[2]) : 1;
  const int max_pages = (ctx + 1 + 15) / 16 + 1;
  const int n_pages = B * max_pages;
  std::vector<int> perm(n_pages); std::iota(perm.begin(), perm.end(), 0);
  std::shuffle(perm.begin(), perm.end(), std::mt19937(1));
  int *pt, *seq; __nv_bfloat16 *kp, *vp, *qkv, *out; float* scratch;
  cudaMalloc(&pt, n_pages * 4); cudaMemcpy(pt, perm.data(), n_pages * 4, cudaMemcpyHostToDevice);
  std::vector<int> sl(B, ctx);
  cudaMalloc(&seq, B * 4); cudaMemcpy(seq, sl.data(), B * 4, cudaMemcpyHostToDevice);
  size_t pool = (size_t)n_pages * H * 16 * 64;
  cudaMalloc(&kp, pool * 2); cudaMalloc(&vp, pool * 2); cudaMemset(kp, 0, pool * 2); cudaMemset(vp, 0, pool * 2);
  cudaMalloc(&qkv, B * 3 * d * 2); cudaMemset(qkv, 0, B * 3 * d * 2);
  cudaMalloc(&out, B * d * 2);
  cudaMalloc(&scratch, decode_attention_scratch_floats(B, H, 16) * 4 + 16);
  DecodeAttnArgs a = {};
  a.qkv = qkv; a.seq_len = seq; a.page_table = pt; a.max_pages_per_seq = max_pages; a.k_pool = kp; a.v_pool = vp;
  a.out = out; a.batch = B; a.n_head = H; a.n_splits = splits; a.scratch = scratch;
  void* flush; cudaMalloc(&flush, 256 << 20);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) decode_attention(a, 0);
  cudaDeviceSynchronize();
  float best = 1e9;
  for (int r = 0; r < 5; ++r) {
    cudaMemset(flush, r, 256 << 20); cudaDeviceSynchronize();
    cudaEventRecord(e0); decode_attention(a, 0); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); best = std::min(best, ms);
  }
  unsigned long long t[16]; cudaMemcpyFromSymbol(t, g_attn_trace, sizeof(t));
  double bytes = (double)B * H * ctx * 256;
  printf("ctx %d splits %d: cold event %.2f us (%.0f GB/s algorithmic) | CTA0 cycles: prewait %llu, pos known %llu, "
         "first loads issued %llu, qkv staged %llu, loop done %llu, end %llu\n",
         ctx, splits, best * 1e3, bytes / (best * 1e-3) / 1e9, t[1] - t[0], t[2] - t[0], t[3] - t[0], t[4] - t[0],
         t[5] - t[0], t[6] - t[0]);
  cudaEventRecord(e0); for (int i = 0; i < 20; ++i) decode_attention(a, 0); cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaMemcpyFromSymbol(t, g_attn_trace, sizeof(t));
  printf("  back-to-back (L2-warm, %0.1f MB) avg %.2f us | CTA0 cycles: prewait %llu, pos known %llu, first loads issued %llu, "
         "qkv staged %llu, loop done %llu, end %llu\n", bytes / 1e6, ms * 1e3 / 20, t[1] - t[0], t[2] - t[0], t[3] - t[0],
         t[4] - t[0], t[5] - t[0], t[6] - t[0]);
  return 0;
}
