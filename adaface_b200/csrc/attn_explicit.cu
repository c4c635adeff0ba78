// Explicit (materialised) cross-attention for the optional modes of CrossAttention.forward that need the score
// matrix itself: `save_attn_vars` (attention.py:245-255 caches attn and attnscore as [B, 8, N, 77]) and the
// convolutional-attention row replacement (attention.py:208-216 -> ldm/util.py:701-879).  Only the 77-token
// cross-attention ever runs here (the score tensor is B*8*N*77 fp32, 161 MB at 64x64, B = 16: the kernels are bound by
// writing and re-reading it); the default path stays on the fused tcgen05 kernel of attention.cu.  The two contractions
// (Q K^T and P V) run on warp-level tensor-core MMAs (mma.sync.m16n8k16 bf16, fp32 accumulate); the score matrix has to
// exist in global memory between them (the conv-attention pass rewrites rows of it), which is why this is not a mode of
// the fused kernel.
//
//   cross_scores_kernel     sim[b,h,n,t] = scale * q[b,n,h,:] . k[b,t,h,:]
//   conv_attn_conv_kernel   conv[i,h,p]  = scale / ks^1.5 * sum_{dy,dx} q[p + (dy,dx) - pad] . k[tok_i[dy*ks+dx]]
//                           (F.pad + grouped F.conv2d of the query image with the subject keys as the filter)
//   conv_attn_scatter_kernel  sim[b_i,h,p,tok_i[m]] = conv[i,h,p - offset_m]  (zero outside the image)
//   softmax_pv_kernel       attn = softmax_t(sim), out[b,n,h,:] = attn . v[b,:,h,:]
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

constexpr int kRows = 128;      // query rows per block (one per thread)

__device__ __forceinline__ float dot8(const uint4& a, const uint4& b, float acc) {
  const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
  const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z), b3 = unpack_bf16x2(b.w);
  acc = fmaf(a0.x, b0.x, acc); acc = fmaf(a0.y, b0.y, acc);
  acc = fmaf(a1.x, b1.x, acc); acc = fmaf(a1.y, b1.y, acc);
  acc = fmaf(a2.x, b2.x, acc); acc = fmaf(a2.y, b2.y, acc);
  acc = fmaf(a3.x, b3.x, acc); acc = fmaf(a3.y, b3.y, acc);
  return acc;
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// shared-memory row pitches (elements): + 8 keeps every row 16-byte aligned and the 8 rows of an ldmatrix / the 8 key
// rows of a B fragment on distinct banks for D = 40 (padded to 48), 80, 160 and for the 80-wide probability rows
template <int D>
struct ExCfg {
  static constexpr int kD16 = (D + 15) / 16 * 16;      // head dim padded to whole k16 steps (zero columns)
  static constexpr int kPitch = kD16 + 8;
};
constexpr int kTMax = 256;

// grid (ceil(N / 128), heads, B); block = 4 warps, 32 query rows (two m16 tiles) each.
// dynamic smem: K head slice [Tp][kPitch], Q tile [128][kPitch] bf16 (Tp = T rounded up to 8; padding rows / columns zero)
template <int D>
__global__ void __launch_bounds__(kRows) cross_scores_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                                                         float* __restrict__ sim, int N, int T, int heads, long long ldq,
                                                         long long ldk, float scale) {
  using X = ExCfg<D>;
  extern __shared__ __align__(16) uint8_t sh_raw[];
  const int Tp = (T + 7) & ~7;
  __nv_bfloat16* sk = reinterpret_cast<__nv_bfloat16*>(sh_raw);           // [Tp][kPitch]
  __nv_bfloat16* sq = sk + Tp * X::kPitch;                                // [128][kPitch]
  constexpr int V = X::kD16 / 8;                                          // 16-byte vectors per padded row
  const int h = blockIdx.y, b = blockIdx.z, n0 = blockIdx.x * kRows;
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  for (int i = threadIdx.x; i < Tp * V; i += blockDim.x) {
    const int t = i / V, c = i % V;
    uint4 val = zero;
    if (t < T && c * 8 < D) val = *reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * T + t) * ldk + h * D + c * 8);
    *reinterpret_cast<uint4*>(sk + t * X::kPitch + c * 8) = val;
  }
  for (int i = threadIdx.x; i < kRows * V; i += blockDim.x) {
    const int r = i / V, c = i % V;
    uint4 val = zero;
    if (n0 + r < N && c * 8 < D) val = *reinterpret_cast<const uint4*>(q + (static_cast<long long>(b) * N + n0 + r) * ldq + h * D + c * 8);
    *reinterpret_cast<uint4*>(sq + r * X::kPitch + c * 8) = val;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int lm_row = ((lane >> 3) & 1) * 8 + (lane & 7), lm_k = (lane >> 4) * 8;
#pragma unroll 1
  for (int mt = 0; mt < 2; ++mt) {
    const int r0 = warp * 32 + mt * 16;
    uint32_t a[X::kD16 / 16][4];
#pragma unroll
    for (int kk = 0; kk < X::kD16 / 16; ++kk)
      ldmatrix_x4(a[kk], smem_u32(sq + (r0 + lm_row) * X::kPitch + kk * 16 + lm_k));
    const int row_a = n0 + r0 + g, row_b = row_a + 8;
    float* out_a = sim + ((static_cast<long long>(b) * heads + h) * N + row_a) * T;
    float* out_b = out_a + 8ll * T;
    for (int nt = 0; nt < Tp / 8; ++nt) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const __nv_bfloat16* kb = sk + (nt * 8 + g) * X::kPitch + 2 * t4;
#pragma unroll
      for (int kk = 0; kk < X::kD16 / 16; ++kk)
        mma_16816(acc, a[kk], *reinterpret_cast<const uint32_t*>(kb + kk * 16), *reinterpret_cast<const uint32_t*>(kb + kk * 16 + 8));
      const int c0 = nt * 8 + 2 * t4;
      if (row_a < N) {
        if (c0 < T) out_a[c0] = acc[0] * scale;
        if (c0 + 1 < T) out_a[c0 + 1] = acc[1] * scale;
      }
      if (row_b < N) {
        if (c0 < T) out_b[c0] = acc[2] * scale;
        if (c0 + 1 < T) out_b[c0 + 1] = acc[3] * scale;
      }
    }
  }
}

// grid (ceil(N / 128), heads, n_inst)
template <int D>
__global__ void __launch_bounds__(kRows) conv_attn_conv_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                                                           const int* __restrict__ inst_batch, const int* __restrict__ tok,
                                                           float* __restrict__ conv, int H, int W, int T, int heads,
                                                           long long ldq, long long ldk, int ks, float scale_conv) {
  constexpr int V = D / 8;
  const int N = H * W;
  const int h = blockIdx.y, ii = blockIdx.z;
  const int b = inst_batch[ii];
  const int p = blockIdx.x * kRows + threadIdx.x;
  if (p >= N) return;
  const int y = p / W, x = p % W;
  const int pad = (ks - 1) / 2;                                  // left / top zero padding: ks 2 -> 0, 3 -> 1, 4 -> 1
  float acc = 0.f;
  for (int m = 0; m < ks * ks; ++m) {
    const int yy = y + m / ks - pad, xx = x + m % ks - pad;
    if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;        // zero padding
    const uint4* qp = reinterpret_cast<const uint4*>(q + (static_cast<long long>(b) * N + yy * W + xx) * ldq + h * D);
    const uint4* kp = reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * T + tok[ii * ks * ks + m]) * ldk + h * D);
#pragma unroll
    for (int c = 0; c < V; ++c) acc = dot8(qp[c], __ldg(kp + c), acc);
  }
  conv[(static_cast<long long>(ii) * heads + h) * N + p] = acc * scale_conv;
}

// one thread per (instance, head, pixel, subject token)
__global__ void conv_attn_scatter_kernel(const float* __restrict__ conv, const int* __restrict__ inst_batch,
                                         const int* __restrict__ tok, float* __restrict__ sim, int H, int W, int T,
                                         int heads, int ks, long long total) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int M = ks * ks, N = H * W;
  const int m = static_cast<int>(idx % M);
  const int p = static_cast<int>((idx / M) % N);
  const int h = static_cast<int>((idx / M / N) % heads);
  const int ii = static_cast<int>(idx / M / N / heads);
  const int pad = (ks - 1) / 2;
  const int oy = m / ks - pad, ox = m % ks - pad;
  const int ys = p / W - oy, xs = p % W - ox;                    // shifted[y, x] = conv[y - oy, x - ox], zero outside
  float v = 0.f;
  if (ys >= 0 && ys < H && xs >= 0 && xs < W) v = conv[(static_cast<long long>(ii) * heads + h) * N + ys * W + xs];
  sim[((static_cast<long long>(inst_batch[ii]) * heads + h) * N + p) * T + tok[ii * M + m]] = v;
}

// grid (ceil(N / 128), heads, B); block = 4 warps.  Phase 1, one thread per query row: row max, exponentials (kept
// unnormalised, <= 1) as bf16 into shared memory, 1 / sum per row, the optional normalised fp32 attention matrix.
// Phase 2: out = (E V) / sum on the tensor cores, 32 rows (two m16 tiles) per warp, V transposed in shared memory so
// that a B fragment's token pair is one 32-bit load.
// dynamic smem: E [128][kEP], V^T [kD16][kEP] bf16 (kEP = Tp16 + 8, Tp16 = T rounded up to 16; padding zero), 1/sum [128] fp32
template <int D>
__global__ void __launch_bounds__(kRows) softmax_pv_kernel(const float* __restrict__ sim, const __nv_bfloat16* __restrict__ v,
                                                       float* __restrict__ attn, __nv_bfloat16* __restrict__ out, int N, int T,
                                                       int heads, long long ldv, long long ldo) {
  using X = ExCfg<D>;
  extern __shared__ __align__(16) uint8_t sh_raw[];
  const int Tp16 = (T + 15) & ~15, EP = Tp16 + 8;
  __nv_bfloat16* se = reinterpret_cast<__nv_bfloat16*>(sh_raw);           // [128][EP]
  __nv_bfloat16* svt = se + kRows * EP;                                   // [kD16][EP]
  float* s_inv = reinterpret_cast<float*>(svt + X::kD16 * EP);            // [128]
  const int h = blockIdx.y, b = blockIdx.z, n0 = blockIdx.x * kRows;
  for (int i = threadIdx.x; i < X::kD16 * EP / 2; i += blockDim.x) reinterpret_cast<uint32_t*>(svt)[i] = 0u;
  __syncthreads();
  for (int i = threadIdx.x; i < T * (D / 8); i += blockDim.x) {
    const int t = i / (D / 8), c = i % (D / 8);
    const uint4 u = *reinterpret_cast<const uint4*>(v + (static_cast<long long>(b) * T + t) * ldv + h * D + c * 8);
    const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&u);
#pragma unroll
    for (int j = 0; j < 8; ++j) svt[(c * 8 + j) * EP + t] = e[j];
  }
  {
    const int n = n0 + threadIdx.x;
    __nv_bfloat16* erow = se + threadIdx.x * EP;
    float inv = 0.f;
    if (n < N) {
      const long long row = ((static_cast<long long>(b) * heads + h) * N + n) * T;
      float mx = -INFINITY;
      for (int t = 0; t < T; ++t) mx = fmaxf(mx, sim[row + t]);
      float sum = 0.f;
      for (int t = 0; t < T; ++t) {
        const float e = __expf(sim[row + t] - mx);          // (second read of the row: L1-resident)
        sum += e;
        erow[t] = __float2bfloat16_rn(e);
      }
      inv = 1.0f / sum;
      if (attn != nullptr)
        for (int t = 0; t < T; ++t) attn[row + t] = __expf(sim[row + t] - mx) * inv;
      for (int t = T; t < Tp16; ++t) erow[t] = __float2bfloat16_rn(0.f);
    } else {
      for (int t = 0; t < Tp16; ++t) erow[t] = __float2bfloat16_rn(0.f);
    }
    s_inv[threadIdx.x] = inv;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int lm_row = ((lane >> 3) & 1) * 8 + (lane & 7), lm_k = (lane >> 4) * 8;
  const int ksteps = Tp16 / 16;
#pragma unroll 1
  for (int mt = 0; mt < 2; ++mt) {
    const int r0 = warp * 32 + mt * 16;
    uint32_t a[kTMax / 16][4];
#pragma unroll
    for (int kk = 0; kk < kTMax / 16; ++kk)
      if (kk < ksteps) ldmatrix_x4(a[kk], smem_u32(se + (r0 + lm_row) * EP + kk * 16 + lm_k));
    const int row_a = n0 + r0 + g, row_b = row_a + 8;
    const float inv_a = s_inv[r0 + g], inv_b = s_inv[r0 + g + 8];
    for (int nt = 0; nt < D / 8; ++nt) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      const __nv_bfloat16* vb = svt + (nt * 8 + g) * EP + 2 * t4;
#pragma unroll
      for (int kk = 0; kk < kTMax / 16; ++kk)
        if (kk < ksteps)
          mma_16816(acc, a[kk], *reinterpret_cast<const uint32_t*>(vb + kk * 16), *reinterpret_cast<const uint32_t*>(vb + kk * 16 + 8));
      const int c0 = h * D + nt * 8 + 2 * t4;
      if (row_a < N)
        *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(b) * N + row_a) * ldo + c0) = pack_bf16x2(acc[0] * inv_a, acc[1] * inv_a);
      if (row_b < N)
        *reinterpret_cast<uint32_t*>(out + (static_cast<long long>(b) * N + row_b) * ldo + c0) = pack_bf16x2(acc[2] * inv_b, acc[3] * inv_b);
    }
  }
}

template <typename Kern>
int set_smem(Kern kern, size_t bytes) {
  if (bytes > 48 * 1024) AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
  return 0;
}

int check_common(const void* a, const void* b, int batch, int heads, int head_dim, int n, int t, long long lda, long long ldb,
                 const char* who) {
  AFB_REQUIRE(a && b, "%s: NULL operand", who);
  AFB_REQUIRE(batch > 0 && heads > 0 && n > 0 && t > 0 && batch <= 65535 && heads <= 65535, "%s: bad sizes", who);
  AFB_REQUIRE(head_dim == 40 || head_dim == 80 || head_dim == 160, "%s: head_dim %d not supported (40, 80, 160)", who, head_dim);
  AFB_REQUIRE(lda >= heads * head_dim && ldb >= heads * head_dim && lda % 8 == 0 && ldb % 8 == 0,
              "%s: leading dimensions must cover heads*head_dim and be multiples of 8", who);
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0, "%s: operands must be 16-byte aligned", who);
  AFB_REQUIRE(t <= kTMax, "%s: context length %d > %d (this path is for the 77-token cross-attention)", who, t, kTMax);
  return 0;
}

}  // namespace
}  // namespace afb

#define AFB_DISPATCH_D(D_, ...)                               \
  switch (D_) {                                               \
    case 40: { constexpr int kD = 40; __VA_ARGS__ } break;    \
    case 80: { constexpr int kD = 80; __VA_ARGS__ } break;    \
    default: { constexpr int kD = 160; __VA_ARGS__ } break;   \
  }

extern "C" int afb_cross_scores(const afb_bf16* q, const afb_bf16* k, float* sim, int batch, int heads, int head_dim, int q_len,
                                int kv_len, int64_t ldq, int64_t ldk, float scale, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_common(q, k, batch, heads, head_dim, q_len, kv_len, ldq, ldk, "afb_cross_scores")) return rc;
  AFB_REQUIRE(sim != nullptr, "afb_cross_scores: NULL output");
  const dim3 grid((q_len + kRows - 1) / kRows, heads, batch);
  const int d16 = (head_dim + 15) / 16 * 16;
  const size_t sh = (static_cast<size_t>((kv_len + 7) & ~7) + kRows) * (d16 + 8) * 2;
  const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(q);
  const __nv_bfloat16* kb = reinterpret_cast<const __nv_bfloat16*>(k);
  AFB_DISPATCH_D(head_dim, {
    if (int rc = set_smem(cross_scores_kernel<kD>, sh)) return rc;
    cross_scores_kernel<kD><<<grid, kRows, sh, stream>>>(qb, kb, sim, q_len, kv_len, heads, ldq, ldk, scale);
  })
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_conv_attn_rows(float* sim, const afb_bf16* q, const afb_bf16* k, const int* inst_batch, const int* tok,
                                  float* conv_ws, int batch, int heads, int head_dim, int h, int w, int kv_len, int64_t ldq,
                                  int64_t ldk, int n_inst, int ks, float scale, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_common(q, k, batch, heads, head_dim, h * w, kv_len, ldq, ldk, "afb_conv_attn_rows")) return rc;
  AFB_REQUIRE(sim && inst_batch && tok && conv_ws, "afb_conv_attn_rows: NULL argument");
  AFB_REQUIRE(ks >= 2 && ks <= 4, "afb_conv_attn_rows: kernel size %d not in 2..4 (ldm/util.py:735-744)", ks);
  AFB_REQUIRE(n_inst > 0 && n_inst <= 65535, "afb_conv_attn_rows: bad instance count %d", n_inst);
  const int n = h * w;
  const dim3 grid((n + kRows - 1) / kRows, heads, n_inst);
  const float scale_conv = scale / (static_cast<float>(ks) * sqrtf(static_cast<float>(ks)));
  const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(q);
  const __nv_bfloat16* kb = reinterpret_cast<const __nv_bfloat16*>(k);
  AFB_DISPATCH_D(head_dim, {
    conv_attn_conv_kernel<kD><<<grid, kRows, 0, stream>>>(qb, kb, inst_batch, tok, conv_ws, h, w, kv_len, heads, ldq, ldk, ks, scale_conv);
  })
  count_launch();
  const long long total = static_cast<long long>(n_inst) * heads * n * ks * ks;
  conv_attn_scatter_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(conv_ws, inst_batch, tok, sim, h, w, kv_len,
                                                                                           heads, ks, total);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_softmax_pv(const float* sim, const afb_bf16* v, float* attn, afb_bf16* out, int batch, int heads, int head_dim,
                              int q_len, int kv_len, int64_t ldv, int64_t ldo, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_common(v, out, batch, heads, head_dim, q_len, kv_len, ldv, ldo, "afb_softmax_pv")) return rc;
  AFB_REQUIRE(sim != nullptr, "afb_softmax_pv: NULL scores");
  const dim3 grid((q_len + kRows - 1) / kRows, heads, batch);
  const int d16 = (head_dim + 15) / 16 * 16, ep = ((kv_len + 15) & ~15) + 8;
  const size_t sh = (static_cast<size_t>(kRows) + d16) * ep * 2 + kRows * sizeof(float);
  AFB_REQUIRE(sh <= 200 * 1024, "afb_softmax_pv: context too long for shared memory");
  const __nv_bfloat16* vb = reinterpret_cast<const __nv_bfloat16*>(v);
  __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(out);
  AFB_DISPATCH_D(head_dim, {
    if (int rc = set_smem(softmax_pv_kernel<kD>, sh)) return rc;
    softmax_pv_kernel<kD><<<grid, kRows, sh, stream>>>(sim, vb, attn, ob, q_len, kv_len, heads, ldv, ldo);
  })
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
