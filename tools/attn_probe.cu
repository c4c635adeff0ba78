// Developer probe (not part of the product library): builds the attention kernel with AFB_ATTN_PROFILE and prints
// (a) the average launch time and (b) where the softmax warps of the first CTAs spend their cycles.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -DAFB_ATTN_PROFILE tools/attn_probe.cu -o tools/bin/attn_probe -lcuda
//   tools/bin/attn_probe [head_dim] [q_len] [kv_len] [batch]
#include "../adaface_b200/csrc/common.cu"
#include "../adaface_b200/csrc/attention.cu"
#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char** argv) {
  const int d = argc > 1 ? atoi(argv[1]) : 40;
  const int nq = argc > 2 ? atoi(argv[2]) : 4096;
  const int nk = argc > 3 ? atoi(argv[3]) : 4096;
  const int batch = argc > 4 ? atoi(argv[4]) : 16;
  const int heads = 8, width = heads * d;
  const size_t nelem_q = (size_t)batch * nq * width, nelem_k = (size_t)batch * nk * width;
  std::vector<__nv_bfloat16> h(std::max(nelem_q, nelem_k));
  unsigned s = 12345u;
  for (auto& v : h) {
    s = s * 1664525u + 1013904223u;
    v = __float2bfloat16(((s >> 8) & 0xffff) / 65536.0f * 2.0f - 1.0f);
  }
  __nv_bfloat16 *q, *k, *v, *o;
  cudaMalloc(&q, nelem_q * 2); cudaMalloc(&k, nelem_k * 2); cudaMalloc(&v, nelem_k * 2); cudaMalloc(&o, nelem_q * 2);
  cudaMemcpy(q, h.data(), nelem_q * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(k, h.data(), nelem_k * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(v, h.data(), nelem_k * 2, cudaMemcpyHostToDevice);
  const int prof_blocks = 4;
#ifdef AFB_ATTN_PROFILE
  cudaMalloc(&afb::g_attn_prof, prof_blocks * 136 * sizeof(long long));
  cudaMemset(afb::g_attn_prof, 0, prof_blocks * 136 * sizeof(long long));
  afb::g_attn_prof_blocks = prof_blocks;
#endif
  afb_attn_args a{};
  a.q = (const afb_bf16*)q; a.k = (const afb_bf16*)k; a.v = (const afb_bf16*)v; a.out = (afb_bf16*)o; a.key_mask = nullptr;
  a.batch = batch; a.heads = heads; a.q_len = nq; a.kv_len = nk; a.head_dim = d;
  a.ldq = a.ldk = a.ldv = a.ldo = width;
  a.scale = 1.0f / sqrtf((float)d);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i)
    if (int rc = afb_attention_fwd(&a, nullptr)) { printf("launch failed rc=%d: %s\n", rc, afb_last_error()); return 1; }
  cudaDeviceSynchronize();
  const int iters = 10;
  cudaEventRecord(e0);
  for (int i = 0; i < iters; ++i) afb_attention_fwd(&a, nullptr);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(err)); return 1; }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double flops = 4.0 * batch * heads * (double)nq * nk * d;
  printf("d=%d q=%d k=%d b=%d: %.1f us / launch, %.1f TF/s\n", d, nq, nk, batch, ms / iters * 1e3, flops / (ms / iters * 1e-3) / 1e12);
#ifndef AFB_ATTN_PROFILE
  return 0;
#else
  std::vector<long long> prof(prof_blocks * 136);
  cudaMemcpy(prof.data(), afb::g_attn_prof, prof.size() * sizeof(long long), cudaMemcpyDeviceToHost);
  const char* names[8] = {"wait S", "ldtm", "max", "token", "exp", "wait P buf", "P store", "loop"};
  const int kv_tiles = (nk + 63) / 64;
  const int nwarps = d <= 40 ? 16 : 8;
  for (int b = 0; b < 2; ++b)
    for (int w = 0; w < nwarps; w += 5) {
      printf("block %d warp %d (cycles per KV tile):", b, w);
      long long tot = 0;
      for (int i = 0; i < 8; ++i) { printf(" %s=%lld", names[i], prof[(b * 17 + w) * 8 + i] / kv_tiles); tot += prof[(b * 17 + w) * 8 + i]; }
      printf(" | total=%lld\n", tot / kv_tiles);
    }
  for (int b = 0; b < 2; ++b) {
    const long long* m = &prof[(b * 17 + 16) * 8];
    printf("block %d MMA thread (cycles per KV tile): wait K/V=%lld poll=%lld issue=%lld loop=%lld\n", b, m[0] / kv_tiles,
           m[1] / kv_tiles, m[2] / kv_tiles, m[3] / kv_tiles);
  }
  return 0;
#endif
}
