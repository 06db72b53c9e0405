// Profiling tool (not part of the library): runs decode-shaped GEMMs with the kernel's
// B200_GEMM_TRACE timestamps enabled and prints CTA 0's milestone timeline in SM cycles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DB200_GEMM_TRACE \
//        -I<pkg>/csrc profiles/gemm_trace.cu -o profiles/gemm_trace -lcudart
#include <cstdio>
#include <string>
#include <vector>

#include "gemm_tcgen05.cu"

namespace b200llm {
static thread_local std::string g_err;
void set_last_error(const std::string& m) { g_err = m; }
int fail(int code, const std::string& m) { g_err = m; return code; }
}  // namespace b200llm

using namespace b200llm;

static void run(const char* name, int tokens, int features, int k, int epi, int split, int cluster = 0, int orient = 1) {
  __nv_bfloat16 *act, *w;
  float *bias, *out;
  const int act_rows = tokens > 256 ? tokens : 256;
  cudaMalloc(&act, (size_t)act_rows * k * 2);
  cudaMalloc(&w, (size_t)features * k * 2);
  cudaMalloc(&bias, features * 4);
  cudaMalloc(&out, (size_t)(tokens > 256 ? tokens : 32 * 256) * features * 4);
  cudaMemset(act, 0, (size_t)act_rows * k * 2);
  cudaMemset(w, 0, (size_t)features * k * 2);
  cudaMemset(bias, 0, features * 4);
  GemmArgs g = {};
  g.act = act; g.w = w; g.bias = bias; g.out = out; g.tokens = tokens; g.features = features; g.k = k;
  g.ld_out = features; g.epilogue = epi; g.split_k = split; g.act_rows_alloc = act_rows;
  g.force_orientation = orient == 2 ? 0 : orient; g.use_pair = orient == 2 ? 1 : 0; g.pdl = 0; g.cluster_split = cluster;
  float* amv = nullptr; int* ami = nullptr; float* eosl = nullptr;
  if (epi == kEpiLogitsArgmax) {
    const int groups = ((features + 127) / 128) * 4;
    cudaMalloc(&amv, (size_t)groups * 256 * 4); cudaMalloc(&ami, (size_t)groups * 256 * 4); cudaMalloc(&eosl, 256 * 4);
    g.amax_val = amv; g.amax_idx = ami; g.eos_logit = eosl; g.ld_amax = 256; g.eos_id = 50256; g.out = nullptr;
  }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) {
    if (gemm_bf16(g, 0) != 0) { printf("%s: %s\n", name, g_err.c_str()); return; }
  }
  cudaDeviceSynchronize();
  // flush L2 so the weights come from HBM like in a real step
  void* flush; cudaMalloc(&flush, 256 << 20); cudaMemset(flush, 1, 256 << 20); cudaDeviceSynchronize();
  cudaEventRecord(e0);
  gemm_bf16(g, 0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long t[16];
  cudaMemcpyFromSymbol(t, g_gemm_trace, sizeof(t));
  printf("%-28s event %.2f us | cycles from CTA start: setup %llu, A issued %llu, pdl ok %llu, first full %llu, "
         "mma done-issue %llu, tmem_full %llu, ld0 done %llu, ld1 done %llu, epilogue end %llu, exit %llu\n",
         name, ms * 1e3, t[1] - t[0], t[2] - t[0], t[3] - t[0], t[4] - t[0], t[5] - t[0], t[6] - t[0],
         t[9] - t[0], t[10] - t[0], t[7] - t[0], t[8] - t[0]);
  // warm L2 timing of 20 back-to-back launches
  cudaEventRecord(e0);
  for (int i = 0; i < 20; ++i) gemm_bf16(g, 0);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  cudaEventElapsedTime(&ms, e0, e1);
  printf("%-28s back-to-back avg %.2f us  (%.0f TFLOP/s)\n", name, ms * 1e3 / 20,
         2.0 * tokens * features * k / (ms * 1e-3 / 20) / 1e12);
  cudaFree(act); cudaFree(w); cudaFree(bias); cudaFree(out); cudaFree(flush);
}

int main() {
  run("qkv 64x2304x768 nosplit", 64, 2304, 768, kEpiBiasBf16, 1);
  run("qkv 64x2304x768 split6", 64, 2304, 768, kEpiPartialF32, 6);
  run("proj 64x768x768 split12", 64, 768, 768, kEpiPartialF32, 12);
  run("fc 64x3072x768 nosplit", 64, 3072, 768, kEpiBiasGeluBf16, 1);
  run("fc 64x3072x768 split6", 64, 3072, 768, kEpiPartialF32, 6);
  run("proj2 64x768x3072 split24", 64, 768, 3072, kEpiPartialF32, 24);
  run("head 64x50257x768", 64, 50257, 768, kEpiF32, 1);
  run("head fused argmax", 64, 50257, 768, kEpiLogitsArgmax, 1);
  printf("-- prefill orientation, 8192 tokens (first tile of CTA 0; flops in TF/s from back-to-back)\n");
  run("pf qkv 8192x2304x768", 8192, 2304, 768, kEpiBiasBf16, 1, 0, 0);
  run("pf fc 8192x3072x768 gelu", 8192, 3072, 768, kEpiBiasGeluBf16, 1, 0, 0);
  run("pf proj 8192x768x768 resid", 8192, 768, 768, kEpiBiasResidF32, 1, 0, 0);
  run("pf proj2 8192x768x3072 resid", 8192, 768, 3072, kEpiBiasResidF32, 1, 0, 0);
  printf("-- prefill orientation with CTA pairs (cta_group::2)\n");
  run("pair qkv 8192x2304x768", 8192, 2304, 768, kEpiBiasBf16, 1, 0, 2);
  run("pair fc 8192x3072x768 gelu", 8192, 3072, 768, kEpiBiasGeluBf16, 1, 0, 2);
  run("pair proj 8192x768x768 resid", 8192, 768, 768, kEpiBiasResidF32, 1, 0, 2);
  run("pair proj2 8192x768x3072", 8192, 768, 3072, kEpiBiasResidF32, 1, 0, 2);
  printf("-- cluster (DSMEM) split-K: ld0=staged in smem, ld1=after cluster barrier 1, epilogue end=before barrier 2\n");
  run("qkv cluster6", 64, 2304, 768, kEpiBiasBf16, 1, 6);
  run("proj cluster6", 64, 768, 768, kEpiBiasResidF32, 1, 6);
  run("fc cluster6", 64, 3072, 768, kEpiBiasGeluBf16, 1, 6);
  run("proj2 cluster8", 64, 768, 3072, kEpiBiasResidF32, 1, 8);
  run("proj2 cluster4", 64, 768, 3072, kEpiBiasResidF32, 1, 4);
  return 0;
}
