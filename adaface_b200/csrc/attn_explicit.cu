// Explicit (materialised) cross-attention for the optional modes of CrossAttention.forward that need the score
// matrix itself: `save_attn_vars` (attention.py:245-255 caches attn and attnscore as [B, 8, N, 77]) and the
// convolutional-attention row replacement (attention.py:208-216 -> ldm/util.py:701-879).  Only the 77-token
// cross-attention ever runs here (the score tensor is B*8*N*77 fp32, 161 MB at 64x64, B = 16); the default path
// stays on the fused tcgen05 kernel of attention.cu.  These are plain CUDA-core kernels: 4*N*77*C FLOPs per sample.
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

// grid (ceil(N / 128), heads, B); dynamic smem: K head slice [T][D] bf16
template <int D>
__global__ void __launch_bounds__(kRows) cross_scores_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k,
                                                         float* __restrict__ sim, int N, int T, int heads, long long ldq,
                                                         long long ldk, float scale) {
  extern __shared__ __align__(16) uint8_t sh_raw[];
  uint4* sk = reinterpret_cast<uint4*>(sh_raw);                 // [T][D / 8]
  constexpr int V = D / 8;
  const int h = blockIdx.y, b = blockIdx.z;
  for (int i = threadIdx.x; i < T * V; i += blockDim.x) {
    const int t = i / V, c = i % V;
    sk[i] = *reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * T + t) * ldk + h * D + c * 8);
  }
  __syncthreads();
  const int n = blockIdx.x * kRows + threadIdx.x;
  if (n >= N) return;
  uint4 qr[V];
  const uint4* qp = reinterpret_cast<const uint4*>(q + (static_cast<long long>(b) * N + n) * ldq + h * D);
#pragma unroll
  for (int c = 0; c < V; ++c) qr[c] = qp[c];
  float* out = sim + ((static_cast<long long>(b) * heads + h) * N + n) * T;
  for (int t = 0; t < T; ++t) {
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < V; ++c) acc = dot8(qr[c], sk[t * V + c], acc);
    out[t] = acc * scale;
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

// grid (ceil(N / 128), heads, B); dynamic smem: V head slice [T][D] bf16 + probabilities [T][128] fp32
template <int D>
__global__ void __launch_bounds__(kRows) softmax_pv_kernel(const float* __restrict__ sim, const __nv_bfloat16* __restrict__ v,
                                                       float* __restrict__ attn, __nv_bfloat16* __restrict__ out, int N, int T,
                                                       int heads, long long ldv, long long ldo) {
  extern __shared__ __align__(16) uint8_t sh_raw[];
  constexpr int V = D / 8;
  uint4* sv = reinterpret_cast<uint4*>(sh_raw);                  // [T][V]
  float* sp = reinterpret_cast<float*>(sh_raw + static_cast<size_t>(T) * V * 16);   // [T][kRows]
  const int h = blockIdx.y, b = blockIdx.z;
  for (int i = threadIdx.x; i < T * V; i += blockDim.x) {
    const int t = i / V, c = i % V;
    sv[i] = *reinterpret_cast<const uint4*>(v + (static_cast<long long>(b) * T + t) * ldv + h * D + c * 8);
  }
  __syncthreads();
  const int n = blockIdx.x * kRows + threadIdx.x;
  if (n >= N) return;
  const long long row = ((static_cast<long long>(b) * heads + h) * N + n) * T;
  float mx = -INFINITY;
  for (int t = 0; t < T; ++t) {
    const float s = sim[row + t];
    sp[t * kRows + threadIdx.x] = s;
    mx = fmaxf(mx, s);
  }
  float sum = 0.f;
  for (int t = 0; t < T; ++t) {
    const float e = __expf(sp[t * kRows + threadIdx.x] - mx);
    sp[t * kRows + threadIdx.x] = e;
    sum += e;
  }
  const float inv = 1.0f / sum;
  if (attn != nullptr)
    for (int t = 0; t < T; ++t) attn[row + t] = sp[t * kRows + threadIdx.x] * inv;
  __nv_bfloat16* orow = out + (static_cast<long long>(b) * N + n) * ldo + h * D;
#pragma unroll 1
  for (int c = 0; c < V; ++c) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int t = 0; t < T; ++t) {
      const float pr = sp[t * kRows + threadIdx.x];
      const uint4 u = sv[t * V + c];
      const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
      acc[0] = fmaf(pr, a0.x, acc[0]); acc[1] = fmaf(pr, a0.y, acc[1]);
      acc[2] = fmaf(pr, a1.x, acc[2]); acc[3] = fmaf(pr, a1.y, acc[3]);
      acc[4] = fmaf(pr, a2.x, acc[4]); acc[5] = fmaf(pr, a2.y, acc[5]);
      acc[6] = fmaf(pr, a3.x, acc[6]); acc[7] = fmaf(pr, a3.y, acc[7]);
    }
    *reinterpret_cast<uint4*>(orow + c * 8) =
        make_uint4(pack_bf16x2(acc[0] * inv, acc[1] * inv), pack_bf16x2(acc[2] * inv, acc[3] * inv),
                   pack_bf16x2(acc[4] * inv, acc[5] * inv), pack_bf16x2(acc[6] * inv, acc[7] * inv));
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
  AFB_REQUIRE(t <= 256, "%s: context length %d > 256 (this path is for the 77-token cross-attention)", who, t);
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
  const size_t sh = static_cast<size_t>(kv_len) * head_dim * 2;
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
  const size_t sh = static_cast<size_t>(kv_len) * head_dim * 2 + static_cast<size_t>(kv_len) * kRows * 4;
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
