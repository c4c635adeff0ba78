// Developer probe (not part of the product library): times one afb_gemm_bf16 shape and, built with -DAFB_GEMM_PROFILE,
// prints where the MMA issuer and two epilogue warps of the first CTAs spend their cycles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo [-DAFB_GEMM_PROFILE] tools/gemm_probe.cu -o tools/bin/gemm_probe -lcuda
//   tools/bin/gemm_probe m n k [geglu 0/1] [ln 0/1] [residual 0/1] [stats 0/1]
#include "../adaface_b200/csrc/common.cu"
#include "../adaface_b200/csrc/gemm.cu"
#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char** argv) {
  const long long m = argc > 1 ? atoll(argv[1]) : 65536, n = argc > 2 ? atoll(argv[2]) : 2560, k = argc > 3 ? atoll(argv[3]) : 320;
  const int geglu = argc > 4 ? atoi(argv[4]) : 1, ln = argc > 5 ? atoi(argv[5]) : 1, res = argc > 6 ? atoi(argv[6]) : 0;
  const int stats = argc > 7 ? atoi(argv[7]) : 0;
  const long long n_out = geglu ? n / 2 : n;
  std::vector<__nv_bfloat16> h((size_t)std::max(m * k, n * k));
  unsigned s = 12345u;
  for (auto& v : h) {
    s = s * 1664525u + 1013904223u;
    v = __float2bfloat16((((s >> 8) & 0xffff) / 65536.0f * 2.0f - 1.0f) * 0.25f);
  }
  __nv_bfloat16 *a, *w, *o, *r;
  float *bias, *cs, *st, *so;
  cudaMalloc(&a, m * k * 2); cudaMalloc(&w, n * k * 2); cudaMalloc(&o, m * n_out * 2); cudaMalloc(&r, m * n_out * 2);
  cudaMalloc(&bias, n * 4); cudaMalloc(&cs, n * 4); cudaMalloc(&st, m * 64 * 2 * 4); cudaMalloc(&so, m * 64 * 2 * 4);
  cudaMemcpy(a, h.data(), m * k * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(w, h.data(), n * k * 2, cudaMemcpyHostToDevice);
  cudaMemset(r, 0, m * n_out * 2); cudaMemset(bias, 0, n * 4); cudaMemset(cs, 0, n * 4);
  std::vector<float> hst((size_t)m * 8 * 2, 1.0f);
  cudaMemcpy(st, hst.data(), hst.size() * 4, cudaMemcpyHostToDevice);
#ifdef AFB_GEMM_PROFILE
  long long* prof;
  cudaMalloc(&prof, 4 * 4 * 8 * sizeof(long long));
  cudaMemset(prof, 0, 4 * 4 * 8 * sizeof(long long));
  cudaMemcpyToSymbol(afb::g_gemm_prof, &prof, sizeof(prof));
#endif
  afb_gemm_args g{};
  g.a = (const afb_bf16*)a; g.w = (const afb_bf16*)w; g.out = o; g.bias = bias;
  g.residual = res ? r : nullptr;
  g.m = m; g.n = n; g.k = k; g.lda = k; g.ldw = k; g.ldo = n_out; g.ldr = n_out;
  g.epilogue = geglu ? AFB_EPI_GEGLU : AFB_EPI_STORE;
  g.out_dtype = AFB_BF16; g.residual_dtype = AFB_BF16;
  if (ln) { g.ln_stats = st; g.ln_colsum = cs; g.ln_parts = 8; g.ln_eps = 1e-5f; }
  if (stats) g.row_stats_out = so;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i)
    if (int rc = afb_gemm_bf16(&g, nullptr)) { printf("launch failed rc=%d: %s\n", rc, afb_last_error()); return 1; }
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
  const int iters = 10;
  cudaEventRecord(e0);
  for (int i = 0; i < iters; ++i) afb_gemm_bf16(&g, nullptr);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(err)); return 1; }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double us = ms / iters * 1e3;
  printf("m=%lld n=%lld k=%lld geglu=%d ln=%d res=%d stats=%d: %.1f us / launch, %.1f TF/s\n", m, n, k, geglu, ln, res, stats, us,
         2.0 * m * n * k / (us * 1e-6) / 1e12);
#ifdef AFB_GEMM_PROFILE
  std::vector<long long> pr(4 * 4 * 8);
  cudaMemcpy(pr.data(), prof, pr.size() * sizeof(long long), cudaMemcpyDeviceToHost);
  const char* mma_names[8] = {"wait acc free", "wait operands", "issue", "", "", "", "", ""};
  const char* epi_names[8] = {"loop top", "wait acc full", "store drain", "wait staging", "compute", "bar+store", "tmem ld", ""};
  for (int b = 0; b < 4; ++b) {
    printf("cta %d MMA issuer   :", b);
    for (int i = 0; i < 3; ++i) printf(" %s=%lld", mma_names[i], pr[(b * 4 + 0) * 8 + i]);
    printf("\n");
    for (int sl = 1; sl <= 2; ++sl) {
      printf("cta %d epilogue w%d :", b, sl == 1 ? 4 : 7);
      for (int i = 0; i < 7; ++i) printf(" %s=%lld", epi_names[i], pr[(b * 4 + sl) * 8 + i]);
      printf("\n");
    }
  }
#endif
  return 0;
}
