// Backward kernels of the inversion step (BASELINE.json configs[3]): gradients flow from the block output back through
// the GEGLU feed-forward, the 77-token cross-attention and its key / value projections into the AdaEmbedding head and
// the attentional pooler, whose parameters are the only trainable tensors (the U-Net is frozen).
//
// The reference has no backward code of its own: torch autograd differentiates the modules of
// ldm/modules/attention.py (:32-59 GEGLU, :195-243 attention, :269-283 LayerNorm) and
// ldm/modules/embedding_manager.py (:146-321 AttentionalPooler, :847-893 AdaEmbedding head) under
// ddpm.py:599 (manual_backward).  Each kernel below is the analytic derivative of the forward kernel it pairs with
// (gemm.cu GEGLU / LayerNorm-fold epilogues, attention.cu, ada.cu); matrix products against frozen weights reuse the
// forward GEMM (dA = dC W), so no weight-gradient GEMM exists on the path.
#include "common.cuh"
#include <stdlib.h>
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

constexpr int kHeads = 8;
constexpr float kInvSqrt8 = 0.35355339059327373f;

__device__ __forceinline__ void unpack8f(const uint4& u, float (&v)[8]) {
  const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ uint4 pack8f(const float (&v)[8]) {
  return make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
}

__device__ __forceinline__ float block_sum_t(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// ---- GEGLU backward (attention.py:32-39): out = a * gelu(g) with (a, g) the halves of h = proj(x).
// h is given in the INTERLEAVED column order of the forward GEMM (32 value columns, their 32 gate columns, ...), and dh
// is produced in the same order, so the dX GEMM that follows multiplies by the same interleaved weight.
// Thread = 8 output columns.  gelu'(g) = Phi(g) + g phi(g) (exact erf form, like the forward).
__global__ void __launch_bounds__(256) geglu_bwd_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ dout,
                                                        __nv_bfloat16* __restrict__ dh, long long rows, int H) {
  const long long vecs_per_row = H >> 3;
  const long long total = rows * vecs_per_row;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long m = i / vecs_per_row;
    const int j0 = static_cast<int>(i % vecs_per_row) * 8;       // first output column
    const int blk = j0 >> 5, off = j0 & 31;
    const long long hrow = m * 2 * H + static_cast<long long>(blk) * 64;
    float a[8], g[8], d[8], da[8], dg[8];
    unpack8f(*reinterpret_cast<const uint4*>(h + hrow + off), a);
    unpack8f(*reinterpret_cast<const uint4*>(h + hrow + 32 + off), g);
    unpack8f(*reinterpret_cast<const uint4*>(dout + m * H + j0), d);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float x = g[k];
      const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
      const float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
      da[k] = d[k] * x * cdf;
      dg[k] = d[k] * a[k] * (cdf + x * pdf);
    }
    *reinterpret_cast<uint4*>(dh + hrow + off) = pack8f(da);
    *reinterpret_cast<uint4*>(dh + hrow + 32 + off) = pack8f(dg);
  }
}

// ---- LayerNorm backward without affine (the affine part is folded into the consumer's weight, attention.py:269-283):
// a_hat = (a - mean) * rstd; da = rstd * (g - mean(g) - a_hat * mean(g * a_hat)) (+ add: the gradient that reaches `a`
// along the residual stream).  One warp per row, the row in registers (C <= 8 * 32 * kMaxChunks).
template <int kMaxChunks>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ g,
                                                            const __nv_bfloat16* __restrict__ add, __nv_bfloat16* __restrict__ da,
                                                            long long rows, int C, float eps) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunks = C >> 3;
  const uint4* ar = reinterpret_cast<const uint4*>(a + row * C);
  const uint4* gr = reinterpret_cast<const uint4*>(g + row * C);
  float av[kMaxChunks][8], gv[kMaxChunks][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks) {
      unpack8f(ar[ch], av[i]);
      unpack8f(gr[ch], gv[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += av[i][j];
    }
  }
  const float mean = warp_sum(s) / C;
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i)
    if (lane + 32 * i < nchunks)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = av[i][j] - mean;
        m2 += d * d;
      }
  const float rstd = rsqrtf(warp_sum(m2) / C + eps);
  float sg = 0.f, sga = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i)
    if (lane + 32 * i < nchunks)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        av[i][j] = (av[i][j] - mean) * rstd;
        sg += gv[i][j];
        sga += gv[i][j] * av[i][j];
      }
  const float mg = warp_sum(sg) / C, mga = warp_sum(sga) / C;
  uint4* outr = reinterpret_cast<uint4*>(da + row * C);
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks) {
      float o[8], r[8];
      if (add != nullptr) unpack8f(reinterpret_cast<const uint4*>(add + row * C)[ch], r);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = rstd * (gv[i][j] - mg - av[i][j] * mga) + (add != nullptr ? r[j] : 0.f);
      outr[ch] = pack8f(o);
    }
  }
}

// ---- attention backward for a short key sequence (the 77-token context, attention.py:195-243):
// S = scale q k^T, P = softmax(S), O = P v;  dV = P^T dO, dP = dO v^T, dS = P * (dP - rowsum(P * dP)),
// dQ = scale dS k, dK = scale dS^T q.  Block (row chunk, head, sample): K and V of the head live in shared memory as
// fp32, the block walks its query rows in tiles of 32 rows: scores and probabilities of the tile in shared memory, dQ
// written per tile, dK / dV accumulated in registers (thread = fixed (key, channel) outputs) over all tiles of the
// block and added to global memory once at the end (fp32 atomics: dk / dv must be zero on entry).
constexpr int kAbRows = 32;          // query rows per tile
constexpr int kAbThreads = 256;
constexpr int kAbMaxKeys = 128;

template <int D>
__global__ void __launch_bounds__(kAbThreads) attn_bwd_short_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
    const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dq, float* __restrict__ dk, float* __restrict__ dv,
    int q_len, int kv_len, long long ldq, long long ldk, long long ldv, long long ldo, long long lddq, long long lddk,
    float scale, int tiles_per_block) {
  extern __shared__ float sh[];
  constexpr int DP = D + 1;                                  // padded row: conflict-free column walks
  const int T = kv_len;
  float* sK = sh;                                            // [T][DP]
  float* sV = sK + T * DP;                                   // [T][DP]
  float* sQ = sV + T * DP;                                   // [32][DP]
  float* sO = sQ + kAbRows * DP;                             // [32][DP]  dO tile
  float* sP = sO + kAbRows * DP;                             // [32][T+1] probabilities
  float* sS = sP + kAbRows * (T + 1);                        // [32][T+1] dS (scaled)
  const int head = blockIdx.y, b = blockIdx.z;
  const int col0 = head * D;
  const int tid = threadIdx.x;
  for (int idx = tid; idx < T * D; idx += kAbThreads) {
    const int t = idx / D, i = idx - t * D;
    sK[t * DP + i] = __bfloat162float(k[(static_cast<long long>(b) * T + t) * ldk + col0 + i]);
    sV[t * DP + i] = __bfloat162float(v[(static_cast<long long>(b) * T + t) * ldv + col0 + i]);
  }
  // outputs owned by this thread for dK / dV: (t, i) pairs idx = tid + n * kAbThreads
  constexpr int kMaxT = D >= 160 ? 80 : kAbMaxKeys;            // bounds the per-thread accumulators (d = 160: 77 keys)
  constexpr int kOwnMax = (kMaxT * D + kAbThreads - 1) / kAbThreads;
  float acc_k[kOwnMax], acc_v[kOwnMax];
#pragma unroll
  for (int n = 0; n < kOwnMax; ++n) acc_k[n] = acc_v[n] = 0.f;
  const int n_own = (T * D + kAbThreads - 1) / kAbThreads;

  const int tile0 = blockIdx.x * tiles_per_block;
  for (int tile = tile0; tile < tile0 + tiles_per_block; ++tile) {
    const int r0 = tile * kAbRows;
    if (r0 >= q_len) break;
    __syncthreads();
    for (int idx = tid; idx < kAbRows * D; idx += kAbThreads) {
      const int r = idx / D, i = idx - r * D;
      const int row = r0 + r;
      float qv = 0.f, ov = 0.f;
      if (row < q_len) {
        qv = __bfloat162float(q[(static_cast<long long>(b) * q_len + row) * ldq + col0 + i]);
        ov = __bfloat162float(dout[(static_cast<long long>(b) * q_len + row) * ldo + col0 + i]);
      }
      sQ[r * DP + i] = qv;
      sO[r * DP + i] = ov;
    }
    __syncthreads();
    // scores and dP: thread (r = tid % 32, keys t = tid / 32 + 8 n)
    {
      const int r = tid & 31;
      for (int t = tid >> 5; t < T; t += kAbThreads / 32) {
        float s = 0.f, dp = 0.f;
#pragma unroll 8
        for (int i = 0; i < D; ++i) {
          s = fmaf(sQ[r * DP + i], sK[t * DP + i], s);
          dp = fmaf(sO[r * DP + i], sV[t * DP + i], dp);
        }
        sP[r * (T + 1) + t] = s * scale;
        sS[r * (T + 1) + t] = dp;
      }
    }
    __syncthreads();
    // softmax and dS per row: one warp handles 4 rows
    {
      const int warp = tid >> 5, lane = tid & 31;
      for (int r = warp; r < kAbRows; r += kAbThreads / 32) {
        float mx = -INFINITY;
        for (int t = lane; t < T; t += 32) mx = fmaxf(mx, sP[r * (T + 1) + t]);
        mx = warp_max(mx);
        float l = 0.f;
        for (int t = lane; t < T; t += 32) {
          const float e = __expf(sP[r * (T + 1) + t] - mx);
          sP[r * (T + 1) + t] = e;
          l += e;
        }
        l = warp_sum(l);
        const float inv = 1.0f / l;
        float delta = 0.f;
        for (int t = lane; t < T; t += 32) {
          const float pr = sP[r * (T + 1) + t] * inv;
          sP[r * (T + 1) + t] = pr;
          delta += pr * sS[r * (T + 1) + t];
        }
        delta = warp_sum(delta);
        const bool live = r0 + r < q_len;
        for (int t = lane; t < T; t += 32) {
          const float pr = sP[r * (T + 1) + t];
          sS[r * (T + 1) + t] = live ? pr * (sS[r * (T + 1) + t] - delta) * scale : 0.f;
          if (!live) sP[r * (T + 1) + t] = 0.f;
        }
      }
    }
    __syncthreads();
    // dQ[r][i] = sum_t dS[r][t] k[t][i]
    if (dq != nullptr) {
      for (int idx = tid; idx < kAbRows * D; idx += kAbThreads) {
        const int r = idx / D, i = idx - r * D;
        if (r0 + r < q_len) {
          float a = 0.f;
          for (int t = 0; t < T; ++t) a = fmaf(sS[r * (T + 1) + t], sK[t * DP + i], a);
          dq[(static_cast<long long>(b) * q_len + r0 + r) * lddq + col0 + i] = __float2bfloat16_rn(a);
        }
      }
    }
    // dK[t][i] += sum_r dS[r][t] q[r][i];  dV[t][i] += sum_r P[r][t] dO[r][i]
#pragma unroll
    for (int n = 0; n < kOwnMax; ++n) {
      if (n < n_own) {
        const int idx = tid + n * kAbThreads;
        if (idx < T * D) {
          const int t = idx / D, i = idx - t * D;
          float ak = 0.f, av = 0.f;
#pragma unroll 8
          for (int r = 0; r < kAbRows; ++r) {
            ak = fmaf(sS[r * (T + 1) + t], sQ[r * DP + i], ak);
            av = fmaf(sP[r * (T + 1) + t], sO[r * DP + i], av);
          }
          acc_k[n] += ak;
          acc_v[n] += av;
        }
      }
    }
  }
#pragma unroll
  for (int n = 0; n < kOwnMax; ++n) {
    if (n < n_own) {
      const int idx = tid + n * kAbThreads;
      if (idx < T * D) {
        const int t = idx / D, i = idx - t * D;
        const long long o = (static_cast<long long>(b) * T + t) * lddk + col0 + i;
        atomicAdd(dk + o, acc_k[n]);
        atomicAdd(dv + o, acc_v[n]);
      }
    }
  }
}

// ---- AdaEmbedding head backward (embedding_manager.py:847-893), block (k, b) like the forward kernel:
// recomputes feat = LNCat3(fg, bg, time), coeffs, vec and its LayerNorm, then
//   dbias[k] += g;  dvec = LN'(g / sqrt(D));  dbasis[k, r] += coeffs[r] dvec;  dcoeffs[r] = <basis[k, r], dvec>;
//   dcoeff_b[k, r] += dcoeffs[r];  dcoeff_w[k, r, :] += dcoeffs[r] feat;  dfeat[b, :] += sum_r dcoeffs[r] coeff_w[k, r, :]
// (all sums over the batch and, for dfeat, over k go through fp32 atomics: the gradient buffers must be zero on entry).
__global__ void __launch_bounds__(256) ada_head_bwd_kernel(
    const float* __restrict__ infeat_fg, const float* __restrict__ infeat_bg, const float* __restrict__ time_emb, int time_ld,
    const float* __restrict__ coeff_w, const float* __restrict__ coeff_b, const float* __restrict__ basis,
    const float* __restrict__ gout, float* __restrict__ d_coeff_w, float* __restrict__ d_coeff_b, float* __restrict__ d_basis,
    float* __restrict__ d_bias, float* __restrict__ d_feat, int C, int TD, int K, int R, int D) {
  extern __shared__ float sh[];    // feat [F], coeffs [R4], dcoeffs [R4], vec [D], dvec [D]
  __shared__ float red[32];
  const int k = blockIdx.x, b = blockIdx.y;
  const int F = 2 * C + TD, R4 = (R + 3) & ~3;
  float* feat = sh;
  float* coeffs = feat + F;
  float* dcoeffs = coeffs + R4;
  float* vec = dcoeffs + R4;
  float* dvec = vec + D;
  for (int i = threadIdx.x; i < F; i += blockDim.x) {
    float v;
    if (i < C) v = infeat_fg[static_cast<long long>(b) * C + i];
    else if (i < 2 * C) v = infeat_bg[static_cast<long long>(b) * C + i - C];
    else v = time_emb[static_cast<long long>(b) * time_ld + i - 2 * C];
    feat[i] = v;
  }
  __syncthreads();
  for (int part = 0; part < 2; ++part) {
    const int lo = part == 0 ? 0 : 2 * C, n = part == 0 ? 2 * C : TD;
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += feat[lo + i];
    const float mean = block_sum_t(s, red) / n;
    float m2 = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float d = feat[lo + i] - mean;
      m2 += d * d;
    }
    const float rstd = rsqrtf(block_sum_t(m2, red) / n + 1e-5f);
    for (int i = threadIdx.x; i < n; i += blockDim.x) feat[lo + i] = (feat[lo + i] - mean) * rstd;
    __syncthreads();
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int o = warp; o < R; o += nwarps) {
    const float* w = coeff_w + static_cast<long long>(k * R + o) * F;
    float a = 0.f;
    for (int i = lane; i < F; i += 32) a += w[i] * feat[i];
    a = warp_sum(a);
    if (lane == 0) coeffs[o] = a + coeff_b[k * R + o];
  }
  __syncthreads();
  const float* bk = basis + static_cast<long long>(k) * R * D;
  float s = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    for (int r = 0; r < R; ++r) a += coeffs[r] * bk[static_cast<long long>(r) * D + d];
    vec[d] = a;
    s += a;
  }
  const float mean = block_sum_t(s, red) / D;
  float m2 = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float dd = vec[d] - mean;
    m2 += dd * dd;
  }
  const float rstd = rsqrtf(block_sum_t(m2, red) / D + 1e-5f);
  const float inv_sqrt_d = rsqrtf(static_cast<float>(D));
  const float* g = gout + (static_cast<long long>(b) * K + k) * D;
  // LayerNorm backward: y = (vec - mean) rstd, out = y / sqrt(D) + bias
  float sg = 0.f, sgy = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float gy = g[d] * inv_sqrt_d;
    const float y = (vec[d] - mean) * rstd;
    sg += gy;
    sgy += gy * y;
    if (d_bias != nullptr) atomicAdd(d_bias + static_cast<long long>(k) * D + d, g[d]);
  }
  const float mg = block_sum_t(sg, red) / D;
  const float mgy = block_sum_t(sgy, red) / D;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float gy = g[d] * inv_sqrt_d;
    const float y = (vec[d] - mean) * rstd;
    dvec[d] = rstd * (gy - mg - y * mgy);
  }
  __syncthreads();
  for (int r = warp; r < R; r += nwarps) {
    float a = 0.f;
    for (int d = lane; d < D; d += 32) {
      a += bk[static_cast<long long>(r) * D + d] * dvec[d];
      atomicAdd(d_basis + (static_cast<long long>(k) * R + r) * D + d, coeffs[r] * dvec[d]);
    }
    a = warp_sum(a);
    if (lane == 0) {
      dcoeffs[r] = a;
      atomicAdd(d_coeff_b + k * R + r, a);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < F; i += blockDim.x) {
    float df = 0.f;
    for (int r = 0; r < R; ++r) {
      const long long wi = static_cast<long long>(k * R + r) * F + i;
      df += dcoeffs[r] * coeff_w[wi];
      atomicAdd(d_coeff_w + wi, dcoeffs[r] * feat[i]);
    }
    atomicAdd(d_feat + static_cast<long long>(b) * F + i, df);
  }
}

// LNCat3 backward for the pooled-feature part (the time slice carries no trainable parameter): block b,
// dinfeat = rstd (df - mean(df) - a_hat mean(df a_hat)) over the 2C pooled features -> d_fg [B, C], d_bg [B, C]
__global__ void __launch_bounds__(256) ada_lncat_bwd_kernel(const float* __restrict__ infeat_fg, const float* __restrict__ infeat_bg,
                                                            const float* __restrict__ d_feat, float* __restrict__ d_fg,
                                                            float* __restrict__ d_bg, int C, int F) {
  __shared__ float red[32];
  const int b = blockIdx.x, n = 2 * C;
  auto in = [&](int i) { return i < C ? infeat_fg[static_cast<long long>(b) * C + i] : infeat_bg[static_cast<long long>(b) * C + i - C]; };
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += in(i);
  const float mean = block_sum_t(s, red) / n;
  float m2 = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float d = in(i) - mean;
    m2 += d * d;
  }
  const float rstd = rsqrtf(block_sum_t(m2, red) / n + 1e-5f);
  float sg = 0.f, sga = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float g = d_feat[static_cast<long long>(b) * F + i], a = (in(i) - mean) * rstd;
    sg += g;
    sga += g * a;
  }
  const float mg = block_sum_t(sg, red) / n, mga = block_sum_t(sga, red) / n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float g = d_feat[static_cast<long long>(b) * F + i], a = (in(i) - mean) * rstd;
    const float o = rstd * (g - mg - a * mga);
    if (i < C) d_fg[static_cast<long long>(b) * C + i] = o;
    else d_bg[static_cast<long long>(b) * C + i - C] = o;
  }
}

// ---- AttentionalPooler backward w.r.t. the folded pooling queries u [B, 2, C] (embedding_manager.py:146-321, eval
// semantics).  With v_t = (x_t + LN(q_t)) / sqrt(8), raw scores s_i,g[t] = <u_i, v_t>_g, smoothed
// s' = s / 2 + mean_g(s) / 2, joint softmax p_i,g[t] over (i, t) with the forward's (max, sum) per (sample, head):
//   dp_i,g[t] = <dout_i, v_t>_g;  ds'_i,g[t] = p (dp - Delta_g),  Delta_g = sum_{c in g} (dfg out_fg + dbg out_bg)[c];
//   ds_i,g = ds'_i,g / 2 + mean_g'(ds'_i,g') / 2;   du_i[c] += sum_t ds_i,g(c)[t] v_t[c].
// One warp per token, per-lane channel accumulators, block partials added to du with atomics (du zero on entry).
// The gradients towards x and q (the layer features) are not produced: the trainable tensors are reached through u.
constexpr int kPbThreads = 256;
constexpr int kPbWarps = kPbThreads / 32;

template <int kChunks>
__global__ void __launch_bounds__(kPbThreads) ada_pool_bwd_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ q, const float* __restrict__ u,
    const uint8_t* __restrict__ key_mask, const float* __restrict__ stats, const float* __restrict__ out_fg,
    const float* __restrict__ out_bg, const float* __restrict__ d_fg, const float* __restrict__ d_bg, float* __restrict__ du,
    int N, int C, int tokens_per_block, float drop_p, uint32_t seed) {
  extern __shared__ float sh[];
  float* s_u = sh;                       // [2][C]
  float* s_d = s_u + 2 * C;              // [2][C] dout
  float* s_part = s_d + 2 * C;           // [8 warps][4][kChunks * 32]
  __shared__ float s_delta[kHeads], s_m[kHeads], s_l[kHeads];
  const int dg = C / kHeads;
  const int b = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int idx = threadIdx.x; idx < 2 * C; idx += blockDim.x) {
    s_u[idx] = u[static_cast<long long>(b) * 2 * C + idx];
    s_d[idx] = idx < C ? d_fg[static_cast<long long>(b) * C + idx] : d_bg[static_cast<long long>(b) * C + idx - C];
  }
  if (threadIdx.x < kHeads) {
    s_m[threadIdx.x] = stats[(static_cast<long long>(b) * kHeads + threadIdx.x) * 2];
    s_l[threadIdx.x] = stats[(static_cast<long long>(b) * kHeads + threadIdx.x) * 2 + 1];
  }
  __syncthreads();
  if (warp < kHeads) {                   // Delta_g
    float a = 0.f;
    for (int c = lane; c < dg; c += 32) {
      const int cc = warp * dg + c;
      a += s_d[cc] * out_fg[static_cast<long long>(b) * C + cc] + s_d[C + cc] * out_bg[static_cast<long long>(b) * C + cc];
    }
    a = warp_sum(a);
    if (lane == 0) s_delta[warp] = a;
  }
  __syncthreads();
  const int nchunks = C >> 3;
  const int chunks_per_group = dg >> 3;
  float acc_fg[kChunks][8], acc_bg[kChunks][8];
#pragma unroll
  for (int i = 0; i < kChunks; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_fg[i][j] = acc_bg[i][j] = 0.f;
  float* my_part = s_part + warp * 4 * (kChunks * 32);
  const int t_begin = blk * tokens_per_block;
  const int t_end = min(N, t_begin + tokens_per_block);
  for (int t = t_begin + warp; t < t_end; t += kPbWarps) {
    const long long rowoff = (static_cast<long long>(b) * N + t) * C;
    const uint4* xr = reinterpret_cast<const uint4*>(x + rowoff);
    const uint4* qr = reinterpret_cast<const uint4*>(q + rowoff);
    float xv[kChunks][8], qv[kChunks][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        unpack8f(xr[ch], xv[i]);
        unpack8f(qr[ch], qv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += qv[i][j];
      }
    }
    const float mean = warp_sum(s) / C;
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i)
      if (lane + 32 * i < nchunks)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = qv[i][j] - mean;
          m2 += d * d;
        }
    const float rstd = rsqrtf(warp_sum(m2) / C + 1e-5f);
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        float a_fg = 0.f, a_bg = 0.f, p_fg = 0.f, p_bg = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xv[i][j] = (xv[i][j] + (qv[i][j] - mean) * rstd) * kInvSqrt8;     // v
          a_fg += s_u[ch * 8 + j] * xv[i][j];
          a_bg += s_u[C + ch * 8 + j] * xv[i][j];
          p_fg += s_d[ch * 8 + j] * xv[i][j];
          p_bg += s_d[C + ch * 8 + j] * xv[i][j];
        }
        my_part[ch] = a_fg;
        my_part[kChunks * 32 + ch] = a_bg;
        my_part[2 * kChunks * 32 + ch] = p_fg;
        my_part[3 * kChunks * 32 + ch] = p_bg;
      }
    }
    __syncwarp();
    // lanes 0..15: (i = lane >> 3, g = lane & 7): raw score and dp of head g
    float sc = 0.f, dp = 0.f;
    if (lane < 16) {
      const float* pp = my_part + (lane >> 3) * (kChunks * 32) + (lane & 7) * chunks_per_group;
      const float* pd = pp + 2 * (kChunks * 32);
      for (int kk = 0; kk < chunks_per_group; ++kk) {
        sc += pp[kk];
        dp += pd[kk];
      }
    }
    __syncwarp();
    float tot = sc;
    tot += __shfl_xor_sync(0xffffffffu, tot, 1);
    tot += __shfl_xor_sync(0xffffffffu, tot, 2);
    tot += __shfl_xor_sync(0xffffffffu, tot, 4);
    sc = 0.5f * sc + 0.5f * (tot * 0.125f);
    const bool masked = key_mask != nullptr && key_mask[static_cast<long long>(b) * N + t] == 0;
    const int g = lane & 7;
    float dsp = 0.f;                                              // d s' of (i, g) held by lanes 0..15
    // (train mode: out = sum_t k_t p_t v_t with the dropout keep-scale k_t, so d s'_t = p_t (k_t dp_t - Delta))
    if (lane < 16 && !masked)
      dsp = __expf(sc - s_m[g]) / s_l[g] * (pool_drop_scale(seed, b, g, lane >> 3, t, N, drop_p) * dp - s_delta[g]);
    float dtot = dsp;
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 1);
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 2);
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 4);
    const float ds = 0.5f * dsp + 0.5f * (dtot * 0.125f);         // d s (raw score) of (i, g)
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      const bool ok = ch < nchunks;
      const int gg = ok ? ch / chunks_per_group : 0;
      const float f = __shfl_sync(0xffffffffu, ds, gg);
      const float bgv = __shfl_sync(0xffffffffu, ds, 8 + gg);
      if (ok) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc_fg[i][j] = fmaf(f, xv[i][j], acc_fg[i][j]);
          acc_bg[i][j] = fmaf(bgv, xv[i][j], acc_bg[i][j]);
        }
      }
    }
  }
  float* dub = du + static_cast<long long>(b) * 2 * C;
#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(dub + ch * 8 + j, acc_fg[i][j]);
        atomicAdd(dub + C + ch * 8 + j, acc_bg[i][j]);
      }
  }
}

// ---- backward of the pooling-query preparation (ada.cu ada_qfinish: LayerNorm of the projected query, grouped conv
// lora_to_{fg,bg}_q, fold through lora_to_k) w.r.t. the three grouped-conv weights [C/8, C/8] (8 groups of lg x dg):
// block (b, i): n = LN(proj_i[b]);  lq[g, o] = <W_i[g, o, :], n_g> / 8;  u[c] = 8^-1/2 sum_o lq[g, o] W_k[g, o, c - g dg]
//   dW_k[g, o, k'] += 8^-1/2 lq[g, o] du[g dg + k'];  dlq[g, o] = 8^-1/2 <W_k[g, o, :], du_g>;  dW_i[g, o, k] += dlq[g, o] n_g[k] / 8
// proj [B + 1, C] is the forward's workspace (row 0: the shared fg probe, row 1 + b: sample b's bg query).
__global__ void __launch_bounds__(256) ada_prep_bwd_kernel(const float* __restrict__ proj, const float* __restrict__ w_fg,
                                                           const float* __restrict__ w_bg, const float* __restrict__ w_k,
                                                           const float* __restrict__ du, float* __restrict__ d_w_fg,
                                                           float* __restrict__ d_w_bg, float* __restrict__ d_w_k, int C) {
  extern __shared__ float sh[];      // n [C], lq [dg], dlq [dg]
  __shared__ float red[32];
  const int b = blockIdx.x, i = blockIdx.y;
  const int dg = C / kHeads, lg = dg / kHeads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float* s_n = sh;
  float* s_lq = sh + C;
  float* s_dlq = s_lq + dg;
  const float* src = proj + static_cast<long long>(i == 0 ? 0 : 1 + b) * C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    s_n[c] = src[c];
    s += src[c];
  }
  const float mean = block_sum_t(s, red) / C;
  float m2 = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float d = s_n[c] - mean;
    m2 += d * d;
  }
  const float rstd = rsqrtf(block_sum_t(m2, red) / C + 1e-5f);
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_n[c] = (s_n[c] - mean) * rstd;
  __syncthreads();
  const float* wsel = (i == 0) ? w_fg : w_bg;
  float* d_wsel = (i == 0) ? d_w_fg : d_w_bg;
  const float* dub = du + (static_cast<long long>(b) * 2 + i) * C;
  for (int j = warp; j < dg; j += nwarps) {      // j = g * lg + o
    const int g = j / lg;
    float a = 0.f, d = 0.f;
    for (int k = lane; k < dg; k += 32) {
      a += wsel[static_cast<long long>(j) * dg + k] * s_n[g * dg + k];
      d += w_k[static_cast<long long>(j) * dg + k] * dub[g * dg + k];
    }
    a = warp_sum(a);
    d = warp_sum(d);
    if (lane == 0) {
      s_lq[j] = a * kInvSqrt8 * kInvSqrt8;
      s_dlq[j] = d * kInvSqrt8;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < dg * dg; idx += blockDim.x) {
    const int j = idx / dg, k = idx - j * dg;
    const int g = j / lg;
    atomicAdd(d_w_k + idx, kInvSqrt8 * s_lq[j] * dub[g * dg + k]);
    atomicAdd(d_wsel + idx, s_dlq[j] * kInvSqrt8 * kInvSqrt8 * s_n[g * dg + k]);
  }
}

}  // namespace
}  // namespace afb

extern "C" int afb_geglu_bwd(const afb_bf16* h, const afb_bf16* dout, afb_bf16* dh, int64_t rows, int hidden, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(h && dout && dh, "afb_geglu_bwd: NULL argument");
  AFB_REQUIRE(rows > 0 && hidden > 0 && hidden % 32 == 0, "afb_geglu_bwd: hidden %d must be a positive multiple of 32", hidden);
  const uintptr_t all = reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(dh);
  AFB_REQUIRE((all & 15) == 0, "afb_geglu_bwd: buffers must be 16-byte aligned");
  const long long total = rows * (hidden / 8);
  const long long blocks = (total + 255) / 256;
  const int grid = static_cast<int>(blocks < 8LL * num_sms() ? blocks : 8LL * num_sms());
  geglu_bwd_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(h), reinterpret_cast<const __nv_bfloat16*>(dout),
                                            reinterpret_cast<__nv_bfloat16*>(dh), rows, hidden);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_layernorm_bwd(const afb_bf16* a, const afb_bf16* g, const afb_bf16* add, afb_bf16* da, int64_t rows,
                                 int channels, float eps, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(a && g && da, "afb_layernorm_bwd: NULL argument");
  AFB_REQUIRE(rows > 0 && channels > 0 && channels % 8 == 0 && channels <= 8 * 32 * 5,
              "afb_layernorm_bwd: channels %d must be a multiple of 8, <= 1280", channels);
  const uintptr_t all = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(add) |
                        reinterpret_cast<uintptr_t>(da);
  AFB_REQUIRE((all & 15) == 0, "afb_layernorm_bwd: buffers must be 16-byte aligned");
  const __nv_bfloat16* ab = reinterpret_cast<const __nv_bfloat16*>(a);
  const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(g);
  const __nv_bfloat16* addb = reinterpret_cast<const __nv_bfloat16*>(add);
  __nv_bfloat16* db = reinterpret_cast<__nv_bfloat16*>(da);
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  const int cpl = (channels / 8 + 31) / 32;
  if (cpl <= 2) layernorm_bwd_kernel<2><<<grid, 256, 0, stream>>>(ab, gb, addb, db, rows, channels, eps);
  else if (cpl <= 3) layernorm_bwd_kernel<3><<<grid, 256, 0, stream>>>(ab, gb, addb, db, rows, channels, eps);
  else layernorm_bwd_kernel<5><<<grid, 256, 0, stream>>>(ab, gb, addb, db, rows, channels, eps);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

namespace afb {
namespace {
template <int D>
int launch_attn_bwd(const afb_attn_args& a, const __nv_bfloat16* dout, __nv_bfloat16* dq, float* dk, float* dv, long long lddq,
                    long long lddk, cudaStream_t stream) {
  const int T = a.kv_len;
  const size_t sh = (2 * static_cast<size_t>(T) * (D + 1) + 2 * static_cast<size_t>(kAbRows) * (D + 1) +
                     2 * static_cast<size_t>(kAbRows) * (T + 1)) * sizeof(float);
  auto kern = attn_bwd_short_kernel<D>;
  static size_t set_for = 0;
  if (set_for < sh) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sh)));
    set_for = sh;
  }
  const int tiles = (a.q_len + kAbRows - 1) / kAbRows;
  // about two waves of blocks; every block keeps its dK / dV partial in registers over its tiles
  int chunks = (2 * num_sms() + a.batch * a.heads - 1) / (a.batch * a.heads);
  if (chunks > tiles) chunks = tiles;
  if (chunks < 1) chunks = 1;
  const int tiles_per_block = (tiles + chunks - 1) / chunks;
  dim3 grid((tiles + tiles_per_block - 1) / tiles_per_block, a.heads, a.batch);
  kern<<<grid, kAbThreads, sh, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.q), reinterpret_cast<const __nv_bfloat16*>(a.k),
                                         reinterpret_cast<const __nv_bfloat16*>(a.v), dout, dq, dk, dv, a.q_len, a.kv_len, a.ldq,
                                         a.ldk, a.ldv, a.ldo, lddq, lddk, a.scale, tiles_per_block);
  return 0;
}
}  // namespace
}  // namespace afb

extern "C" int afb_ada_head_bwd(const float* infeat_fg, const float* infeat_bg, const float* time_emb, int time_ld,
                                const float* coeff_w, const float* coeff_b, const float* basis, const float* grad_out,
                                float* d_coeff_w, float* d_coeff_b, float* d_basis, float* d_bias, float* d_feat_ws,
                                float* d_infeat_fg, float* d_infeat_bg, int batch, int channels, int td, int num_vectors, int r,
                                int out_dim, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(infeat_fg && infeat_bg && time_emb && coeff_w && coeff_b && basis && grad_out && d_coeff_w && d_coeff_b && d_basis &&
                  d_feat_ws && d_infeat_fg && d_infeat_bg,
              "afb_ada_head_bwd: NULL argument");
  AFB_REQUIRE(batch > 0 && channels > 0 && td > 0 && td <= time_ld && num_vectors > 0 && r > 0 && out_dim > 0, "afb_ada_head_bwd: bad sizes");
  const int F = 2 * channels + td;
  const size_t sh = (static_cast<size_t>(F) + 2 * static_cast<size_t>((r + 3) & ~3) + 2 * static_cast<size_t>(out_dim)) * sizeof(float);
  AFB_REQUIRE(sh <= 48 * 1024, "afb_ada_head_bwd: feature vector too large for one block (%zu bytes)", sh);
  AFB_CHECK_CUDA(cudaMemsetAsync(d_feat_ws, 0, static_cast<size_t>(batch) * F * sizeof(float), stream));
  ada_head_bwd_kernel<<<dim3(num_vectors, batch), 256, sh, stream>>>(infeat_fg, infeat_bg, time_emb, time_ld, coeff_w, coeff_b, basis,
                                                                      grad_out, d_coeff_w, d_coeff_b, d_basis, d_bias, d_feat_ws,
                                                                      channels, td, num_vectors, r, out_dim);
  count_launch();
  ada_lncat_bwd_kernel<<<batch, 256, 0, stream>>>(infeat_fg, infeat_bg, d_feat_ws, d_infeat_fg, d_infeat_bg, channels, F);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_pool_bwd(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask,
                                const float* stats, const float* fg_out, const float* bg_out, const float* d_fg, const float* d_bg,
                                float* d_pool_q, int batch, int n_tokens, int channels, float drop_p, uint32_t seed,
                                afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && q && pool_q && stats && fg_out && bg_out && d_fg && d_bg && d_pool_q, "afb_ada_pool_bwd: NULL argument");
  AFB_REQUIRE(batch > 0 && n_tokens > 0 && channels % 64 == 0 && channels <= 1280, "afb_ada_pool_bwd: bad sizes");
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) & 15) == 0, "afb_ada_pool_bwd: x/q must be 16-byte aligned");
  const int C = channels;
  int blocks = 2 * num_sms() / batch;
  // at least AFB_ADA_BWD_MIN_TOKENS (default 32) tokens per block: every block ends with 2 C atomics on du, which is what
  // a 16x16 layer's launch consisted of at one token per warp
  static const int min_tokens = [] {
    const char* e = getenv("AFB_ADA_BWD_MIN_TOKENS");
    const int v = e ? atoi(e) : 32;
    return v < 8 ? 8 : v;
  }();
  const int max_blocks = (n_tokens + min_tokens - 1) / min_tokens;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  const int tokens_per_block = (n_tokens + blocks - 1) / blocks;
  dim3 grid((n_tokens + tokens_per_block - 1) / tokens_per_block, batch);
  const int cpl = (C / 8 + 31) / 32;
  AFB_CHECK_CUDA(cudaMemsetAsync(d_pool_q, 0, static_cast<size_t>(batch) * 2 * C * sizeof(float), stream));
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(q);
#define AFB_LAUNCH_PB(KCH)                                                                                          \
  {                                                                                                                 \
    auto kern = ada_pool_bwd_kernel<KCH>;                                                                           \
    const size_t shb = (4 * static_cast<size_t>(C) + kPbWarps * 4 * (KCH * 32)) * sizeof(float);                    \
    static size_t set_for = 0;                                                                                      \
    if (set_for < shb) {                                                                                            \
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(shb))); \
      set_for = shb;                                                                                                \
    }                                                                                                               \
    kern<<<grid, kPbThreads, shb, stream>>>(xb, qb, pool_q, key_mask, stats, fg_out, bg_out, d_fg, d_bg, d_pool_q, n_tokens, C, \
                                            tokens_per_block, drop_p, seed);                                        \
  }
  if (cpl <= 2) AFB_LAUNCH_PB(2)
  else if (cpl <= 3) AFB_LAUNCH_PB(3)
  else AFB_LAUNCH_PB(5)
#undef AFB_LAUNCH_PB
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_pool_prep_bwd(const float* proj_ws, const float* lora_to_fg_q, const float* lora_to_bg_q,
                                     const float* lora_to_k, const float* d_pool_q, float* d_lora_to_fg_q, float* d_lora_to_bg_q,
                                     float* d_lora_to_k, int batch, int channels, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(proj_ws && lora_to_fg_q && lora_to_bg_q && lora_to_k && d_pool_q && d_lora_to_fg_q && d_lora_to_bg_q && d_lora_to_k,
              "afb_ada_pool_prep_bwd: NULL argument");
  AFB_REQUIRE(batch > 0 && channels % 64 == 0 && channels <= 1280, "afb_ada_pool_prep_bwd: bad sizes");
  const size_t sh = (static_cast<size_t>(channels) + 2 * static_cast<size_t>(channels / 8)) * sizeof(float);
  ada_prep_bwd_kernel<<<dim3(batch, 2), 256, sh, stream>>>(proj_ws, lora_to_fg_q, lora_to_bg_q, lora_to_k, d_pool_q, d_lora_to_fg_q,
                                                            d_lora_to_bg_q, d_lora_to_k, channels);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ================================================================================================================
// Second stage: the pieces that carry a layer's gradient further down the U-Net - self-attention (long key sequences),
// GroupNorm, and the pooler's gradient towards the layer features.
// ================================================================================================================
namespace afb {
namespace {

// delta[b, h, q] = sum_i dO[b, q, h d + i] * O[b, q, h d + i]  (= rowsum(P * dP)); one thread per (row, head)
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                                                         float* __restrict__ delta, long long rows, int q_len, int heads, int d,
                                                         long long ldo, long long lddo) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= rows * heads) return;
  const long long row = idx / heads;
  const int h = static_cast<int>(idx % heads);
  const __nv_bfloat16* po = o + row * ldo + h * d;
  const __nv_bfloat16* pd = dout + row * lddo + h * d;
  float a = 0.f;
  for (int i = 0; i < d; i += 8) {
    float x[8], y[8];
    unpack8f(*reinterpret_cast<const uint4*>(po + i), x);
    unpack8f(*reinterpret_cast<const uint4*>(pd + i), y);
#pragma unroll
    for (int j = 0; j < 8; ++j) a = fmaf(x[j], y[j], a);
  }
  const long long b = row / q_len, qi = row % q_len;
  delta[(b * heads + h) * q_len + qi] = a;
}

// ---- attention backward for long key sequences (self-attention, attention.py:195-243), flash style:
// block (key tile of 64, head, sample) keeps K, V of its tile in shared memory and walks the query tiles of 64 rows:
//   P = 2^(s scale log2e - lse), dP = dO V^T, dS = P (dP - delta) scale,
//   dV += P^T dO, dK += dS^T Q (register accumulators over all query tiles, written once),
//   dQ += dS K (fp32 atomics: every key tile adds its share; dq must be zero on entry).
// fp32 CUDA-core arithmetic with 4 x 4 register tiles: a first, correct version of this kernel - the tensor-core
// (tcgen05) form of the forward kernel is not built for the backward yet.
constexpr int kAlT = 64;             // keys per block tile and query rows per step
constexpr int kAlThreads = 256;

template <int D>
__global__ void __launch_bounds__(kAlThreads) attn_bwd_long_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
    const __nv_bfloat16* __restrict__ dout, const float* __restrict__ lse, const float* __restrict__ delta,
    float* __restrict__ dq, float* __restrict__ dk, float* __restrict__ dv, int q_len, int kv_len, int heads, long long ldq,
    long long ldk, long long ldv, long long ldo, long long ldg, float scale) {
  extern __shared__ float sh[];
  constexpr int DP = D + 1;
  constexpr int TP = kAlT + 1;
  float* sK = sh;                         // [64][DP]
  float* sV = sK + kAlT * DP;
  float* sQ = sV + kAlT * DP;
  float* sO = sQ + kAlT * DP;             // dO tile
  float* sP = sO + kAlT * DP;             // [64 rows][TP]
  float* sS = sP + kAlT * TP;             // dS
  float* sL = sS + kAlT * TP;             // [64] lse, [64] delta
  const int head = blockIdx.y, b = blockIdx.z;
  const int k0 = blockIdx.x * kAlT;
  const int col0 = head * D;
  const int tid = threadIdx.x;
  const float scale_log2 = scale * 1.4426950408889634f;
  for (int idx = tid; idx < kAlT * D; idx += kAlThreads) {
    const int t = idx / D, i = idx - t * D;
    const bool ok = k0 + t < kv_len;
    sK[t * DP + i] = ok ? __bfloat162float(k[(static_cast<long long>(b) * kv_len + k0 + t) * ldk + col0 + i]) : 0.f;
    sV[t * DP + i] = ok ? __bfloat162float(v[(static_cast<long long>(b) * kv_len + k0 + t) * ldv + col0 + i]) : 0.f;
  }
  // dK / dV outputs of this thread: key t = tid % 64, channels i = (tid / 64) + 4 n
  constexpr int kOwn = (D + 3) / 4;
  float acc_k[kOwn], acc_v[kOwn];
#pragma unroll
  for (int n = 0; n < kOwn; ++n) acc_k[n] = acc_v[n] = 0.f;
  const int ty = tid >> 4, tx = tid & 15;       // 4 x 4 micro tile: rows 4 ty.., keys 4 tx..
  for (int r0 = 0; r0 < q_len; r0 += kAlT) {
    __syncthreads();
    for (int idx = tid; idx < kAlT * D; idx += kAlThreads) {
      const int r = idx / D, i = idx - r * D;
      const bool ok = r0 + r < q_len;
      sQ[r * DP + i] = ok ? __bfloat162float(q[(static_cast<long long>(b) * q_len + r0 + r) * ldq + col0 + i]) : 0.f;
      sO[r * DP + i] = ok ? __bfloat162float(dout[(static_cast<long long>(b) * q_len + r0 + r) * ldo + col0 + i]) : 0.f;
    }
    if (tid < kAlT) {
      const bool ok = r0 + tid < q_len;
      const long long li = (static_cast<long long>(b) * heads + head) * q_len + r0 + tid;
      sL[tid] = ok ? lse[li] : 0.f;
      sL[kAlT + tid] = ok ? delta[li] : 0.f;
    }
    __syncthreads();
    {
      float s[4][4], dp[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) s[a][c] = dp[a][c] = 0.f;
#pragma unroll 4
      for (int i = 0; i < D; ++i) {
        float qa[4], oa[4], kc[4], vc[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          qa[a] = sQ[(4 * ty + a) * DP + i];
          oa[a] = sO[(4 * ty + a) * DP + i];
          kc[a] = sK[(4 * tx + a) * DP + i];
          vc[a] = sV[(4 * tx + a) * DP + i];
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            s[a][c] = fmaf(qa[a], kc[c], s[a][c]);
            dp[a][c] = fmaf(oa[a], vc[c], dp[a][c]);
          }
      }
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int r = 4 * ty + a;
        const bool row_ok = r0 + r < q_len;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int t = 4 * tx + c;
          const bool ok = row_ok && k0 + t < kv_len;
          const float pr = ok ? exp2f(s[a][c] * scale_log2 - sL[r]) : 0.f;
          sP[r * TP + t] = pr;
          sS[r * TP + t] = pr * (dp[a][c] - sL[kAlT + r]) * scale;
        }
      }
    }
    __syncthreads();
    // dQ[r][i] += sum_t dS[r][t] K[t][i]: thread (r = tid % 64, channels (tid / 64) + 4 n)
    {
      const int r = tid & 63, i0 = tid >> 6;
      if (r0 + r < q_len) {
        float* dst = dq + (static_cast<long long>(b) * q_len + r0 + r) * ldg + col0;
        for (int i = i0; i < D; i += 4) {
          float a = 0.f;
#pragma unroll 8
          for (int t = 0; t < kAlT; ++t) a = fmaf(sS[r * TP + t], sK[t * DP + i], a);
          atomicAdd(dst + i, a);
        }
      }
    }
    // dK[t][i] += sum_r dS[r][t] Q[r][i];  dV[t][i] += sum_r P[r][t] dO[r][i]
    {
      const int t = tid & 63, i0 = tid >> 6;
#pragma unroll
      for (int n = 0; n < kOwn; ++n) {
        const int i = i0 + 4 * n;
        if (i < D) {
          float ak = 0.f, av = 0.f;
#pragma unroll 8
          for (int r = 0; r < kAlT; ++r) {
            ak = fmaf(sS[r * TP + t], sQ[r * DP + i], ak);
            av = fmaf(sP[r * TP + t], sO[r * DP + i], av);
          }
          acc_k[n] += ak;
          acc_v[n] += av;
        }
      }
    }
  }
  {
    const int t = tid & 63, i0 = tid >> 6;
    if (k0 + t < kv_len) {
      const long long o = (static_cast<long long>(b) * kv_len + k0 + t) * ldg + col0;
#pragma unroll
      for (int n = 0; n < kOwn; ++n) {
        const int i = i0 + 4 * n;
        if (i < D) {
          dk[o + i] = acc_k[n];
          dv[o + i] = acc_v[n];
        }
      }
    }
  }
}

// ---- the same backward on the tensor cores (warp-level mma.sync m16n8k16, bf16 operands, fp32 accumulation) ------
// (The first version of this - 64-key tiles, 4 warps, a 32-bit shared-memory load per fragment register and 24 scalar
//  atomics per thread and tile for dQ - ran a 64x64 layer at batch 8 in 3.78 ms; it is gone, the numbers are in
//  profiles/r02_attention_bwd_probe.txt.)
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ---- 128-key tiles, 8 warps, every fragment through ldmatrix (AFB_ATTN_BWD=mma2; the default at d = 160) -----------
// The first version spent its issue slots on shared-memory loads (a 32-bit load per A / B register, two 16-bit loads per
// transposed one: ~490 loads for 120 MMAs per warp and tile) and sent 24 scalar fp32 atomics per thread and tile to dQ.
// Here: K / V of 128 keys stay resident, the 64-row Q / dO tiles (+ lse, delta) arrive through a cp.async double buffer
// while the previous tile computes, all operand fragments come from ldmatrix / ldmatrix.trans (78 loads per warp and
// tile), P = 2^(s - lse) on ex2.approx, and dQ leaves as red.global.add.v2.f32 - 8x fewer atomic instructions per key.
// Warp w = (rw = w & 3, kh = w >> 2): S / dP for rows 16 rw.. x keys 64 kh.. (two 32-key halves to bound the registers),
// dQ for rows 16 rw.. x column half kh over all 128 keys, dK / dV for keys 16 w.. over the 64 rows.
constexpr int kA2Keys = 128;
constexpr int kA2Rows = 64;
constexpr int kA2Threads = 256;
constexpr int kA2PS = kA2Keys + 8;     // row pitch of the P / dS tiles (bf16): 272 B, conflict-free ldmatrix rows

template <int D, int kSplit>
struct A2Cfg {
  static constexpr int DPAD = (D + 15) / 16 * 16;
  static constexpr int LD = DPAD + 8;              // operand row pitch (bf16): 16-byte aligned rows, conflict-free ldmatrix
  static constexpr int CW = DPAD / kSplit;         // output columns of this block
  static constexpr int NT = CW / 8;                // dK / dV accumulator tiles of a warp
  static constexpr int NTQ = CW / 16;              // dQ accumulator tiles of a warp (half of the block's columns)
  static constexpr int kStages = DPAD <= 80 ? 2 : 1;
  static constexpr int kMinBlocks = DPAD <= 48 ? 2 : 1;
  static constexpr size_t kSmem = (static_cast<size_t>(2 * kA2Keys + kStages * 2 * kA2Rows) * LD + 2 * kA2Rows * kA2PS) * 2 +
                                  static_cast<size_t>(kStages) * 2 * kA2Rows * sizeof(float);
  static_assert(CW % 16 == 0 && NT % 2 == 0, "column slice");
  static_assert(kSmem <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ void bw_ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void bw_ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void bw_ldsm_x2_t(uint32_t& r0, uint32_t& r1, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void bw_cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;                            // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void bw_cp_async4(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void bw_red_add_v2(float* dst, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(a), "f"(b) : "memory");
}

template <int D, int kSplit>
__global__ void __launch_bounds__(kA2Threads, A2Cfg<D, kSplit>::kMinBlocks) attn_bwd_mma2_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
    const __nv_bfloat16* __restrict__ dout, const float* __restrict__ lse, const float* __restrict__ delta,
    float* __restrict__ dq, float* __restrict__ dk, float* __restrict__ dv, int q_len, int kv_len, int heads, long long ldq,
    long long ldk, long long ldv, long long ldo, long long ldg, float scale, const uint8_t* __restrict__ key_mask) {
  using C = A2Cfg<D, kSplit>;
  constexpr int DPAD = C::DPAD, LD = C::LD, CW = C::CW, NT = C::NT, NTQ = C::NTQ, NS = C::kStages;
  constexpr int PS = kA2PS;
  __shared__ uint8_t s_keep[kA2Keys];            // key mask of this block's keys (1 = attended)
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_raw);      // [128][LD]
  __nv_bfloat16* sV = sK + kA2Keys * LD;
  __nv_bfloat16* sQ = sV + kA2Keys * LD;                               // [NS][64][LD]
  __nv_bfloat16* sO = sQ + NS * kA2Rows * LD;                          // [NS][64][LD]  dO tiles
  __nv_bfloat16* sP = sO + NS * kA2Rows * LD;                          // [64][PS]
  __nv_bfloat16* sS = sP + kA2Rows * PS;                               // [64][PS]      dS
  float* sL = reinterpret_cast<float*>(sS + kA2Rows * PS);             // [NS][lse 64 | delta 64]
  const int head = blockIdx.y, b = blockIdx.z;
  const int k0 = (blockIdx.x / kSplit) * kA2Keys;
  const int cs = (blockIdx.x % kSplit) * CW;     // first output column of this block
  const int col0 = head * D;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int rw = warp & 3, kh = warp >> 2;
  const float scale_log2 = scale * 1.4426950408889634f;
  const bool has_mask = key_mask != nullptr;
  // ldmatrix lane offsets.  A operand, rows x k (plain): matrices (rows 0-7, k 0-7), (rows 8-15, k 0-7), (rows 0-7, k 8-15), ..
  const int la_row = (lane & 7) + ((lane >> 3) & 1) * 8, la_col = (lane >> 4) * 8;
  // B operand from a [n][k] tile (plain): (n 0-7, k 0-7), (n 0-7, k 8-15), (n 8-15, k 0-7), (n 8-15, k 8-15)
  const int lb_row = (lane & 7) + (lane >> 4) * 8, lb_col = ((lane >> 3) & 1) * 8;
  // transposed loads read a [k][m or n] tile: same lane pattern as the A operand with rows = k
  //   B: (k 0-7, n 0-7), (k 8-15, n 0-7), (k 0-7, n 8-15), (k 8-15, n 8-15)        -> la_row / la_col
  //   A: (k 0-7, m 0-7), (k 0-7, m 8-15), (k 8-15, m 0-7), (k 8-15, m 8-15)        -> lb_row / lb_col
  auto issue_tile = [&](int it, int stage) {
    const int r0 = it * kA2Rows;
    const uint32_t uq = smem_u32(sQ + stage * kA2Rows * LD), uo = smem_u32(sO + stage * kA2Rows * LD);
    for (int idx = tid; idx < kA2Rows * (DPAD / 8); idx += kA2Threads) {
      const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
      const bool ok = r0 + r < q_len && c8 < D;
      const long long row = static_cast<long long>(b) * q_len + (ok ? r0 + r : 0);
      const int cc = ok ? c8 : 0;
      bw_cp_async16(uq + (r * LD + c8) * 2, q + row * ldq + col0 + cc, ok);
      bw_cp_async16(uo + (r * LD + c8) * 2, dout + row * ldo + col0 + cc, ok);
    }
    if (tid < 2 * kA2Rows) {
      const int r = tid & (kA2Rows - 1);
      const bool ok = r0 + r < q_len;
      const long long li = (static_cast<long long>(b) * heads + head) * q_len + (ok ? r0 + r : 0);
      bw_cp_async4(smem_u32(sL + stage * 2 * kA2Rows + tid), (tid < kA2Rows ? lse : delta) + li, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (NS == 2) issue_tile(0, 0);
  // K, V tiles (rows past kv_len and the padding columns are zero)
  for (int idx = tid; idx < kA2Keys * (DPAD / 8); idx += kA2Threads) {
    const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
    uint4 kk4 = make_uint4(0, 0, 0, 0), vv4 = kk4;
    if (k0 + r < kv_len && c8 < D) {
      kk4 = *reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * kv_len + k0 + r) * ldk + col0 + c8);
      vv4 = *reinterpret_cast<const uint4*>(v + (static_cast<long long>(b) * kv_len + k0 + r) * ldv + col0 + c8);
    }
    *reinterpret_cast<uint4*>(sK + r * LD + c8) = kk4;
    *reinterpret_cast<uint4*>(sV + r * LD + c8) = vv4;
  }
  if (tid < kA2Keys)
    s_keep[tid] = (!has_mask || k0 + tid >= kv_len) ? 1 : key_mask[static_cast<long long>(b) * kv_len + k0 + tid];
  float acc_k[NT][4], acc_v[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc_k[n][e] = acc_v[n][e] = 0.f;
  const uint32_t uK = smem_u32(sK), uV = smem_u32(sV), uP = smem_u32(sP), uS = smem_u32(sS);
  const bool keys_full = k0 + kA2Keys <= kv_len && !has_mask;

  const int n_tiles = (q_len + kA2Rows - 1) / kA2Rows;
  for (int it = 0; it < n_tiles; ++it) {
    const int st = NS == 2 ? (it & 1) : 0;
    const int r0 = it * kA2Rows;
    if (NS == 1) {
      __syncthreads();                            // the single stage is consumed
      issue_tile(it, 0);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                              // tile `it` (and, first time, K / V) visible; tile it-1 fully consumed
    if (NS == 2 && it + 1 < n_tiles) issue_tile(it + 1, st ^ 1);
    const uint32_t uQ = smem_u32(sQ + st * kA2Rows * LD), uO = smem_u32(sO + st * kA2Rows * LD);
    const float* tl = sL + st * 2 * kA2Rows;
    const bool full = keys_full && r0 + kA2Rows <= q_len;
    // ---- S = Q K^T and dP = dO V^T: rows 16 rw .., keys 64 kh .. in two halves of 32
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int key_base = 64 * kh + 32 * hh;
      float s_acc[4][4], p_acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) s_acc[j][e] = p_acc[j][e] = 0.f;
#pragma unroll
      for (int kk = 0; kk < DPAD / 16; ++kk) {
        uint32_t aq[4], ao[4];
        const uint32_t a_off = ((16 * rw + la_row) * LD + kk * 16 + la_col) * 2;
        bw_ldsm_x4(aq, uQ + a_off);
        bw_ldsm_x4(ao, uO + a_off);
#pragma unroll
        for (int jp = 0; jp < 2; ++jp) {
          uint32_t bk[4], bv[4];
          const uint32_t b_off = ((key_base + jp * 16 + lb_row) * LD + kk * 16 + lb_col) * 2;
          bw_ldsm_x4(bk, uK + b_off);
          bw_ldsm_x4(bv, uV + b_off);
          mma_bf16_16816(s_acc[2 * jp], aq, bk[0], bk[1]);
          mma_bf16_16816(s_acc[2 * jp + 1], aq, bk[2], bk[3]);
          mma_bf16_16816(p_acc[2 * jp], ao, bv[0], bv[1]);
          mma_bf16_16816(p_acc[2 * jp + 1], ao, bv[2], bv[3]);
        }
      }
      // P and dS of rows g and g + 8 of this warp's 16, keys key_base + 8 j + 2 t, + 1
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = 16 * rw + g + 8 * half;
        const bool row_ok = r0 + r < q_len;
        const float l2 = tl[r], dl = tl[kA2Rows + r];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = key_base + 8 * j + 2 * t;
          float pr[2], ds[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const bool ok = full || (row_ok && k0 + c + e < kv_len);
            pr[e] = ok ? ex2_approx(fmaf(s_acc[j][2 * half + e], scale_log2, -l2)) : 0.f;
            ds[e] = pr[e] * (p_acc[j][2 * half + e] - dl) * scale;
            if (has_mask && s_keep[c + e] == 0) {
              // masked key (attention.py:223-232, masked_fill after the scaling): its score is a constant, so no gradient
              // reaches q / k through it; its probability is 0 unless the whole row is masked (the forward's lse is then
              // the fill value itself and the row averages V uniformly)
              pr[e] = (ok && l2 < -1.0e29f) ? 1.0f / static_cast<float>(kv_len) : 0.f;
              ds[e] = 0.f;
            }
          }
          *reinterpret_cast<uint32_t*>(sP + r * PS + c) = pack_bf16x2(pr[0], pr[1]);
          *reinterpret_cast<uint32_t*>(sS + r * PS + c) = pack_bf16x2(ds[0], ds[1]);
        }
      }
    }
    __syncthreads();                              // every warp's P and dS are in shared memory
    // ---- dQ[16 rows][CW / 2] = dS[16 x 128] K[128 x CW / 2]
    {
      float d_acc[NTQ][4];
#pragma unroll
      for (int n = 0; n < NTQ; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) d_acc[n][e] = 0.f;
      const int cq = cs + kh * (CW / 2);
#pragma unroll
      for (int kk = 0; kk < kA2Keys / 16; ++kk) {
        uint32_t a[4];
        bw_ldsm_x4(a, uS + ((16 * rw + la_row) * PS + kk * 16 + la_col) * 2);
#pragma unroll
        for (int n = 0; n < NTQ; n += 2) {
          if (n + 1 < NTQ) {
            uint32_t bb[4];
            bw_ldsm_x4_t(bb, uK + ((kk * 16 + la_row) * LD + cq + 8 * n + la_col) * 2);
            mma_bf16_16816(d_acc[n], a, bb[0], bb[1]);
            mma_bf16_16816(d_acc[n + 1 < NTQ ? n + 1 : n], a, bb[2], bb[3]);
          } else {
            uint32_t b0, b1;
            bw_ldsm_x2_t(b0, b1, uK + ((kk * 16 + la_row) * LD + cq + 8 * n) * 2);
            mma_bf16_16816(d_acc[n], a, b0, b1);
          }
        }
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = r0 + 16 * rw + g + 8 * half;
        if (r < q_len) {
          float* dst = dq + (static_cast<long long>(b) * q_len + r) * ldg + col0;
#pragma unroll
          for (int n = 0; n < NTQ; ++n) {
            const int c = cq + 8 * n + 2 * t;          // even; D is a multiple of 8, so c < D covers c + 1
            if (c < D) bw_red_add_v2(dst + c, d_acc[n][2 * half], d_acc[n][2 * half + 1]);
          }
        }
      }
    }
    // ---- dV[16 keys][CW] += P^T dO,  dK[16 keys][CW] += dS^T Q  (keys 16 w .. of the tile; A = the tiles read transposed)
#pragma unroll
    for (int kk = 0; kk < kA2Rows / 16; ++kk) {
      uint32_t ap[4], as[4];
      const uint32_t a_off = ((kk * 16 + lb_row) * PS + 16 * warp + lb_col) * 2;
      bw_ldsm_x4_t(ap, uP + a_off);
      bw_ldsm_x4_t(as, uS + a_off);
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        uint32_t bo[4], bq[4];
        const uint32_t b_off = ((kk * 16 + la_row) * LD + cs + 8 * n + la_col) * 2;
        bw_ldsm_x4_t(bo, uO + b_off);
        bw_ldsm_x4_t(bq, uQ + b_off);
        mma_bf16_16816(acc_v[n], ap, bo[0], bo[1]);
        mma_bf16_16816(acc_v[n + 1], ap, bo[2], bo[3]);
        mma_bf16_16816(acc_k[n], as, bq[0], bq[1]);
        mma_bf16_16816(acc_k[n + 1], as, bq[2], bq[3]);
      }
    }
  }
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int key = k0 + 16 * warp + g + 8 * half;
    if (key < kv_len) {
      const long long o = (static_cast<long long>(b) * kv_len + key) * ldg + col0;
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        const int c = cs + 8 * n + 2 * t;
        if (c < D) {
          *reinterpret_cast<float2*>(dk + o + c) = make_float2(acc_k[n][2 * half], acc_k[n][2 * half + 1]);
          *reinterpret_cast<float2*>(dv + o + c) = make_float2(acc_v[n][2 * half], acc_v[n][2 * half + 1]);
        }
      }
    }
  }
}

template <int D, int kSplit>
int launch_attn_bwd_mma2(const afb_attn_args& a, const __nv_bfloat16* dout, const float* delta, float* dq, float* dk, float* dv,
                         long long ldg, cudaStream_t stream) {
  using C = A2Cfg<D, kSplit>;
  auto kern = attn_bwd_mma2_kernel<D, kSplit>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(C::kSmem)));
    attr_set = true;
  }
  dim3 grid(((a.kv_len + kA2Keys - 1) / kA2Keys) * kSplit, a.heads, a.batch);
  kern<<<grid, kA2Threads, C::kSmem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.q), reinterpret_cast<const __nv_bfloat16*>(a.k),
                                               reinterpret_cast<const __nv_bfloat16*>(a.v), dout, a.lse, delta, dq, dk, dv, a.q_len,
                                               a.kv_len, a.heads, a.ldq, a.ldk, a.ldv, a.ldo, ldg, a.scale, a.key_mask);
  return 0;
}

// ---- third version (the default): the scores are computed TRANSPOSED, so P^T and dS^T never leave the registers -------
// ncu on the version above: the shared-memory data pipe is 88 % busy with (almost conflict-free) ldmatrix wavefronts -
// 352 per warp and tile - while the HMMA pipe idles at 42 %.  Here warp w owns keys 16 w .. for everything but dQ:
// S^T = K Q^T and dP^T = V dO^T take their A operand (the warp's 16 keys of K / V) from fragments that are loaded ONCE
// per block and stay in registers, and their accumulators (m = key, n = query row) have exactly the layout of the A
// operand of dV += P^T dO and dK += dS^T Q, so P^T and dS^T go from accumulators to MMA operands through a bf16 pack -
// no P tile in shared memory, no transposed re-read.  Only dS^T is written out ([key][row], 16 stores per thread) for
// dQ = dS K, which reads it back transposed as before.  The 64-row tile is walked in 16-row steps (one k16 step of the
// dV / dK MMAs) to bound the registers.  Per warp and tile: ~300 shared-memory wavefronts instead of ~390.
constexpr int kA3TS = kA2Rows + 8;       // row pitch of the dS^T tile (bf16): 144 B, conflict-free ldmatrix rows

template <int D, int kSplit>
struct A3Cfg {
  static constexpr int DPAD = (D + 15) / 16 * 16;
  static constexpr int LD = DPAD + 8;
  static constexpr int CW = DPAD / kSplit;
  static constexpr int NT = CW / 8;
  static constexpr int NTQ = CW / 16;
  static constexpr int KK = DPAD / 16;
  static constexpr bool kKeepKV = DPAD <= 80;       // K / V fragments of the warp's keys stay in registers
  static constexpr int kUnroll = DPAD <= 48 ? 4 : 1;  // 16-row steps unrolled where the registers allow (d = 40: 128)
  static constexpr int kStages = DPAD <= 80 ? 2 : 1;
  static constexpr int kMinBlocks = DPAD <= 48 ? 2 : 1;
  static constexpr size_t kSmem = (static_cast<size_t>(2 * kA2Keys + kStages * 2 * kA2Rows) * LD + kA2Keys * kA3TS) * 2 +
                                  static_cast<size_t>(kStages) * 2 * kA2Rows * sizeof(float);
  static_assert(CW % 16 == 0 && NT % 2 == 0, "column slice");
  static_assert(kSmem <= 227 * 1024, "shared memory budget");
};

template <int D, int kSplit>
__global__ void __launch_bounds__(kA2Threads, A3Cfg<D, kSplit>::kMinBlocks) attn_bwd_mma3_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
    const __nv_bfloat16* __restrict__ dout, const float* __restrict__ lse, const float* __restrict__ delta,
    float* __restrict__ dq, float* __restrict__ dk, float* __restrict__ dv, int q_len, int kv_len, int heads, long long ldq,
    long long ldk, long long ldv, long long ldo, long long ldg, float scale, const uint8_t* __restrict__ key_mask) {
  using C = A3Cfg<D, kSplit>;
  constexpr int DPAD = C::DPAD, LD = C::LD, CW = C::CW, NT = C::NT, NTQ = C::NTQ, NS = C::kStages, KK = C::KK;
  constexpr int TS = kA3TS;
  __shared__ uint8_t s_keep[kA2Keys];
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_raw);      // [128][LD]
  __nv_bfloat16* sV = sK + kA2Keys * LD;
  __nv_bfloat16* sQ = sV + kA2Keys * LD;                               // [NS][64][LD]
  __nv_bfloat16* sO = sQ + NS * kA2Rows * LD;                          // [NS][64][LD]  dO tiles
  __nv_bfloat16* sT = sO + NS * kA2Rows * LD;                          // [128][TS]     dS^T: [key][row]
  float* sL = reinterpret_cast<float*>(sT + kA2Keys * TS);             // [NS][lse 64 | delta 64]
  const int head = blockIdx.y, b = blockIdx.z;
  const int k0 = (blockIdx.x / kSplit) * kA2Keys;
  const int cs = (blockIdx.x % kSplit) * CW;
  const int col0 = head * D;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int rw = warp & 3, kh = warp >> 2;
  const float scale_log2 = scale * 1.4426950408889634f;
  const bool has_mask = key_mask != nullptr;
  const int la_row = (lane & 7) + ((lane >> 3) & 1) * 8, la_col = (lane >> 4) * 8;     // see attn_bwd_mma2_kernel
  const int lb_row = (lane & 7) + (lane >> 4) * 8, lb_col = ((lane >> 3) & 1) * 8;
  auto issue_tile = [&](int it, int stage) {
    const int r0 = it * kA2Rows;
    const uint32_t uq = smem_u32(sQ + stage * kA2Rows * LD), uo = smem_u32(sO + stage * kA2Rows * LD);
    for (int idx = tid; idx < kA2Rows * (DPAD / 8); idx += kA2Threads) {
      const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
      const bool ok = r0 + r < q_len && c8 < D;
      const long long row = static_cast<long long>(b) * q_len + (ok ? r0 + r : 0);
      const int cc = ok ? c8 : 0;
      bw_cp_async16(uq + (r * LD + c8) * 2, q + row * ldq + col0 + cc, ok);
      bw_cp_async16(uo + (r * LD + c8) * 2, dout + row * ldo + col0 + cc, ok);
    }
    if (tid < 2 * kA2Rows) {
      const int r = tid & (kA2Rows - 1);
      const bool ok = r0 + r < q_len;
      const long long li = (static_cast<long long>(b) * heads + head) * q_len + (ok ? r0 + r : 0);
      bw_cp_async4(smem_u32(sL + stage * 2 * kA2Rows + tid), (tid < kA2Rows ? lse : delta) + li, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (NS == 2) issue_tile(0, 0);
  for (int idx = tid; idx < kA2Keys * (DPAD / 8); idx += kA2Threads) {
    const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
    uint4 kk4 = make_uint4(0, 0, 0, 0), vv4 = kk4;
    if (k0 + r < kv_len && c8 < D) {
      kk4 = *reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * kv_len + k0 + r) * ldk + col0 + c8);
      vv4 = *reinterpret_cast<const uint4*>(v + (static_cast<long long>(b) * kv_len + k0 + r) * ldv + col0 + c8);
    }
    *reinterpret_cast<uint4*>(sK + r * LD + c8) = kk4;
    *reinterpret_cast<uint4*>(sV + r * LD + c8) = vv4;
  }
  if (tid < kA2Keys)
    s_keep[tid] = (!has_mask || k0 + tid >= kv_len) ? 1 : key_mask[static_cast<long long>(b) * kv_len + k0 + tid];
  float acc_k[NT][4], acc_v[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc_k[n][e] = acc_v[n][e] = 0.f;
  const uint32_t uK = smem_u32(sK), uV = smem_u32(sV), uT = smem_u32(sT);
  __syncthreads();                                // K / V / s_keep visible
  // this warp's 16 keys of K and V as A fragments (m = key, k = channel)
  uint32_t fk[C::kKeepKV ? KK : 1][4], fv[C::kKeepKV ? KK : 1][4];
  if constexpr (C::kKeepKV) {
#pragma unroll
    for (int kk = 0; kk < KK; ++kk) {
      const uint32_t a_off = ((16 * warp + la_row) * LD + kk * 16 + la_col) * 2;
      bw_ldsm_x4(fk[kk], uK + a_off);
      bw_ldsm_x4(fv[kk], uV + a_off);
    }
  }
  const int key_lo = 16 * warp + g;               // this thread's accumulator rows: keys key_lo and key_lo + 8 of the tile
  const bool kvalid[2] = {k0 + key_lo < kv_len, k0 + key_lo + 8 < kv_len};
  const bool kkeep[2] = {s_keep[key_lo] != 0, s_keep[key_lo + 8] != 0};
  const bool keys_full = k0 + kA2Keys <= kv_len && !has_mask;
  const float inv_kv = 1.0f / static_cast<float>(kv_len);

  const int n_tiles = (q_len + kA2Rows - 1) / kA2Rows;
  for (int it = 0; it < n_tiles; ++it) {
    const int st = NS == 2 ? (it & 1) : 0;
    const int r0 = it * kA2Rows;
    if (NS == 1) {
      __syncthreads();                            // the single stage is consumed
      issue_tile(it, 0);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                              // tile `it` visible; tile it-1 (and its dS^T) fully consumed
    if (NS == 2 && it + 1 < n_tiles) issue_tile(it + 1, st ^ 1);
    const uint32_t uQ = smem_u32(sQ + st * kA2Rows * LD), uO = smem_u32(sO + st * kA2Rows * LD);
    const float* tl = sL + st * 2 * kA2Rows;
    const bool full = keys_full && r0 + kA2Rows <= q_len;
#pragma unroll(C::kUnroll)
    for (int s16 = 0; s16 < kA2Rows / 16; ++s16) {
      // ---- S^T = K Q^T, dP^T = V dO^T: keys 16 w .. x rows 16 s16 ..
      float st_acc[2][4], dp_acc[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) st_acc[j][e] = dp_acc[j][e] = 0.f;
#pragma unroll
      for (int kk = 0; kk < KK; ++kk) {
        uint32_t bq[4], bo[4];
        const uint32_t b_off = ((16 * s16 + lb_row) * LD + kk * 16 + lb_col) * 2;       // B: n = row, k = channel
        bw_ldsm_x4(bq, uQ + b_off);
        bw_ldsm_x4(bo, uO + b_off);
        if constexpr (C::kKeepKV) {
          mma_bf16_16816(st_acc[0], fk[kk], bq[0], bq[1]);
          mma_bf16_16816(st_acc[1], fk[kk], bq[2], bq[3]);
          mma_bf16_16816(dp_acc[0], fv[kk], bo[0], bo[1]);
          mma_bf16_16816(dp_acc[1], fv[kk], bo[2], bo[3]);
        } else {
          uint32_t ak[4], av[4];
          const uint32_t a_off = ((16 * warp + la_row) * LD + kk * 16 + la_col) * 2;
          bw_ldsm_x4(ak, uK + a_off);
          bw_ldsm_x4(av, uV + a_off);
          mma_bf16_16816(st_acc[0], ak, bq[0], bq[1]);
          mma_bf16_16816(st_acc[1], ak, bq[2], bq[3]);
          mma_bf16_16816(dp_acc[0], av, bo[0], bo[1]);
          mma_bf16_16816(dp_acc[1], av, bo[2], bo[3]);
        }
      }
      // ---- P^T, dS^T: accumulator element (j, 2 half + e) = key key_lo + 8 half, row 16 s16 + 8 j + 2 t + e
      uint32_t ap[4], as[4];                      // A fragments of the dV / dK MMAs: index 2 j + half
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int r = 16 * s16 + 8 * j + 2 * t;
        const float2 l2 = *reinterpret_cast<const float2*>(tl + r);
        const float2 dl = *reinterpret_cast<const float2*>(tl + kA2Rows + r);
        const bool rok[2] = {r0 + r < q_len, r0 + r + 1 < q_len};
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          float pr[2], ds[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const float le = e ? l2.y : l2.x, de = e ? dl.y : dl.x;
            const bool ok = full || (rok[e] && kvalid[half]);
            pr[e] = ok ? ex2_approx(fmaf(st_acc[j][2 * half + e], scale_log2, -le)) : 0.f;
            ds[e] = pr[e] * (dp_acc[j][2 * half + e] - de) * scale;
            if (has_mask && !kkeep[half]) {
              // masked key (attention.py:223-232, masked_fill after the scaling): its score is a constant, so no gradient
              // reaches q / k through it; its probability is 0 unless the whole row is masked (the forward's lse is then
              // the fill value itself and the row averages V uniformly)
              pr[e] = (ok && le < -1.0e29f) ? inv_kv : 0.f;
              ds[e] = 0.f;
            }
          }
          ap[2 * j + half] = pack_bf16x2(pr[0], pr[1]);
          as[2 * j + half] = pack_bf16x2(ds[0], ds[1]);
          *reinterpret_cast<uint32_t*>(sT + (key_lo + 8 * half) * TS + r) = as[2 * j + half];
        }
      }
      // ---- dV[16 keys][CW] += P^T dO,  dK[16 keys][CW] += dS^T Q over these 16 rows (one k16 step)
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        uint32_t bo[4], bq[4];
        const uint32_t b_off = ((16 * s16 + la_row) * LD + cs + 8 * n + la_col) * 2;    // B: k = row, n = channel
        bw_ldsm_x4_t(bo, uO + b_off);
        bw_ldsm_x4_t(bq, uQ + b_off);
        mma_bf16_16816(acc_v[n], ap, bo[0], bo[1]);
        mma_bf16_16816(acc_v[n + 1], ap, bo[2], bo[3]);
        mma_bf16_16816(acc_k[n], as, bq[0], bq[1]);
        mma_bf16_16816(acc_k[n + 1], as, bq[2], bq[3]);
      }
    }
    __syncthreads();                              // dS^T of the whole tile is in shared memory
    // ---- dQ[16 rows][CW / 2] = dS[16 x 128] K[128 x CW / 2]: rows 16 rw .., column half kh
    {
      float d_acc[NTQ][4];
#pragma unroll
      for (int n = 0; n < NTQ; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) d_acc[n][e] = 0.f;
      const int cq = cs + kh * (CW / 2);
#pragma unroll
      for (int kk = 0; kk < kA2Keys / 16; ++kk) {
        uint32_t a[4];
        bw_ldsm_x4_t(a, uT + ((kk * 16 + lb_row) * TS + 16 * rw + lb_col) * 2);           // A: k = key, m = row
#pragma unroll
        for (int n = 0; n < NTQ; n += 2) {
          if (n + 1 < NTQ) {
            uint32_t bb[4];
            bw_ldsm_x4_t(bb, uK + ((kk * 16 + la_row) * LD + cq + 8 * n + la_col) * 2);
            mma_bf16_16816(d_acc[n], a, bb[0], bb[1]);
            mma_bf16_16816(d_acc[n + 1 < NTQ ? n + 1 : n], a, bb[2], bb[3]);
          } else {
            uint32_t b0, b1;
            bw_ldsm_x2_t(b0, b1, uK + ((kk * 16 + la_row) * LD + cq + 8 * n) * 2);
            mma_bf16_16816(d_acc[n], a, b0, b1);
          }
        }
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = r0 + 16 * rw + g + 8 * half;
        if (r < q_len) {
          float* dst = dq + (static_cast<long long>(b) * q_len + r) * ldg + col0;
#pragma unroll
          for (int n = 0; n < NTQ; ++n) {
            const int c = cq + 8 * n + 2 * t;
            if (c < D) bw_red_add_v2(dst + c, d_acc[n][2 * half], d_acc[n][2 * half + 1]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int key = k0 + key_lo + 8 * half;
    if (key < kv_len) {
      const long long o = (static_cast<long long>(b) * kv_len + key) * ldg + col0;
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        const int c = cs + 8 * n + 2 * t;
        if (c < D) {
          *reinterpret_cast<float2*>(dk + o + c) = make_float2(acc_k[n][2 * half], acc_k[n][2 * half + 1]);
          *reinterpret_cast<float2*>(dv + o + c) = make_float2(acc_v[n][2 * half], acc_v[n][2 * half + 1]);
        }
      }
    }
  }
}

template <int D, int kSplit>
int launch_attn_bwd_mma3(const afb_attn_args& a, const __nv_bfloat16* dout, const float* delta, float* dq, float* dk, float* dv,
                         long long ldg, cudaStream_t stream) {
  using C = A3Cfg<D, kSplit>;
  auto kern = attn_bwd_mma3_kernel<D, kSplit>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(C::kSmem)));
    attr_set = true;
  }
  dim3 grid(((a.kv_len + kA2Keys - 1) / kA2Keys) * kSplit, a.heads, a.batch);
  kern<<<grid, kA2Threads, C::kSmem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.q), reinterpret_cast<const __nv_bfloat16*>(a.k),
                                               reinterpret_cast<const __nv_bfloat16*>(a.v), dout, a.lse, delta, dq, dk, dv, a.q_len,
                                               a.kv_len, a.heads, a.ldq, a.ldk, a.ldv, a.ldo, ldg, a.scale, a.key_mask);
  return 0;
}

// ---- the short-key backward (cross-attention, 77 keys) on tensor cores -------------------------------------------------
// All keys of a head fit one tile (80 = 10 n8 tiles), so the softmax of a row is complete inside the warp that owns it:
// block (row chunk, head, sample) = 5 warps walks 80-row query tiles; warp w computes S = Q K^T and dP = dO V^T of rows
// 16 w .. (accumulators in registers), the row softmax / delta / dS with quad shuffles, dQ = dS K with dS as the A operand
// straight from those registers (bf16 rows written once, no atomics), leaves P and dS in shared memory and then owns keys
// 16 w .. for dV += P^T dO and dK += dS^T Q over the tile's 80 rows (ldmatrix.trans operands; accumulators in registers
// over all tiles of the block, added to global memory once at the end as vector atomics: dk / dv zero on entry).
// The first version (attn_bwd_short_kernel: fp32 CUDA cores, scalar bf16 loads) took 1.62 ms for one 64x64 layer at
// batch 8 - 14 ms of the 78 ms inversion step for 8 GFLOP per layer.
constexpr int kAsKeys = 80;
constexpr int kAsRows = 80;
constexpr int kAsThreads = 160;
constexpr int kAsPS = kAsKeys + 8;     // row pitch of the P / dS tiles (bf16): 176 B, conflict-free ldmatrix rows

template <int D, int kSplit>
struct AsCfg {
  static constexpr int DPAD = (D + 15) / 16 * 16;
  static constexpr int LD = DPAD + 8;
  static constexpr int CW = DPAD / kSplit;         // output columns of this block (dQ, dK, dV)
  static constexpr int NT = CW / 8;
  static constexpr int kStages = DPAD <= 80 ? 2 : 1;
  static constexpr size_t kSmem = (static_cast<size_t>(2 * kAsKeys + kStages * 2 * kAsRows) * LD + 2 * kAsRows * kAsPS) * 2;
  static_assert(NT % 2 == 0, "column slice");
  static_assert(kSmem <= 227 * 1024, "shared memory budget");
};

template <int D, int kSplit>
__global__ void __launch_bounds__(kAsThreads) attn_bwd_short_mma_kernel(
    const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ k, const __nv_bfloat16* __restrict__ v,
    const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dq, float* __restrict__ dk, float* __restrict__ dv,
    int q_len, int kv_len, long long ldq, long long ldk, long long ldv, long long ldo, long long lddq, long long lddk,
    float scale, int tiles_per_block) {
  using C = AsCfg<D, kSplit>;
  constexpr int DPAD = C::DPAD, LD = C::LD, CW = C::CW, NT = C::NT, NS = C::kStages, PS = kAsPS;
  constexpr int NJ = kAsKeys / 8;                // key tiles of the score accumulators
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_raw);      // [80][LD]
  __nv_bfloat16* sV = sK + kAsKeys * LD;
  __nv_bfloat16* sQ = sV + kAsKeys * LD;                               // [NS][80][LD]
  __nv_bfloat16* sO = sQ + NS * kAsRows * LD;                          // [NS][80][LD]  dO tiles
  __nv_bfloat16* sP = sO + NS * kAsRows * LD;                          // [80][PS]
  __nv_bfloat16* sS = sP + kAsRows * PS;                               // [80][PS]      dS
  const int head = blockIdx.y, b = blockIdx.z, T = kv_len;
  const int cs = (blockIdx.x % kSplit) * CW;
  const int col0 = head * D;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int n_tiles = (q_len + kAsRows - 1) / kAsRows;
  const int tile0 = (blockIdx.x / kSplit) * tiles_per_block;
  const int tile_end = tile0 + tiles_per_block < n_tiles ? tile0 + tiles_per_block : n_tiles;
  if (tile0 >= tile_end) return;
  const float scale_log2 = scale * 1.4426950408889634f;
  const int la_row = (lane & 7) + ((lane >> 3) & 1) * 8, la_col = (lane >> 4) * 8;     // see attn_bwd_mma2_kernel
  const int lb_row = (lane & 7) + (lane >> 4) * 8, lb_col = ((lane >> 3) & 1) * 8;
  auto issue_tile = [&](int it, int stage) {
    const int r0 = it * kAsRows;
    const uint32_t uq = smem_u32(sQ + stage * kAsRows * LD), uo = smem_u32(sO + stage * kAsRows * LD);
    for (int idx = tid; idx < kAsRows * (DPAD / 8); idx += kAsThreads) {
      const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
      const bool ok = r0 + r < q_len && c8 < D;
      const long long row = static_cast<long long>(b) * q_len + (ok ? r0 + r : 0);
      const int cc = ok ? c8 : 0;
      bw_cp_async16(uq + (r * LD + c8) * 2, q + row * ldq + col0 + cc, ok);
      bw_cp_async16(uo + (r * LD + c8) * 2, dout + row * ldo + col0 + cc, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (NS == 2) issue_tile(tile0, 0);
  for (int idx = tid; idx < kAsKeys * (DPAD / 8); idx += kAsThreads) {
    const int r = idx / (DPAD / 8), c8 = (idx % (DPAD / 8)) * 8;
    uint4 kk4 = make_uint4(0, 0, 0, 0), vv4 = kk4;
    if (r < T && c8 < D) {
      kk4 = *reinterpret_cast<const uint4*>(k + (static_cast<long long>(b) * T + r) * ldk + col0 + c8);
      vv4 = *reinterpret_cast<const uint4*>(v + (static_cast<long long>(b) * T + r) * ldv + col0 + c8);
    }
    *reinterpret_cast<uint4*>(sK + r * LD + c8) = kk4;
    *reinterpret_cast<uint4*>(sV + r * LD + c8) = vv4;
  }
  float acc_k[NT][4], acc_v[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc_k[n][e] = acc_v[n][e] = 0.f;
  const uint32_t uK = smem_u32(sK), uV = smem_u32(sV), uP = smem_u32(sP), uS = smem_u32(sS);

  for (int it = tile0; it < tile_end; ++it) {
    const int st = NS == 2 ? ((it - tile0) & 1) : 0;
    const int r0 = it * kAsRows;
    if (NS == 1) {
      __syncthreads();
      issue_tile(it, 0);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                              // tile `it` (first time: K / V) visible; the previous tile is consumed
    if (NS == 2 && it + 1 < tile_end) issue_tile(it + 1, st ^ 1);
    const uint32_t uQ = smem_u32(sQ + st * kAsRows * LD), uO = smem_u32(sO + st * kAsRows * LD);
    // ---- S = Q K^T, dP = dO V^T: rows 16 w .. x all keys
    float s_acc[NJ][4], p_acc[NJ][4];
#pragma unroll
    for (int j = 0; j < NJ; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) s_acc[j][e] = p_acc[j][e] = 0.f;
#pragma unroll
    for (int kk = 0; kk < DPAD / 16; ++kk) {
      uint32_t aq[4], ao[4];
      const uint32_t a_off = ((16 * warp + la_row) * LD + kk * 16 + la_col) * 2;
      bw_ldsm_x4(aq, uQ + a_off);
      bw_ldsm_x4(ao, uO + a_off);
#pragma unroll
      for (int jp = 0; jp < NJ / 2; ++jp) {
        uint32_t bk[4], bv[4];
        const uint32_t b_off = ((jp * 16 + lb_row) * LD + kk * 16 + lb_col) * 2;
        bw_ldsm_x4(bk, uK + b_off);
        bw_ldsm_x4(bv, uV + b_off);
        mma_bf16_16816(s_acc[2 * jp], aq, bk[0], bk[1]);
        mma_bf16_16816(s_acc[2 * jp + 1], aq, bk[2], bk[3]);
        mma_bf16_16816(p_acc[2 * jp], ao, bv[0], bv[1]);
        mma_bf16_16816(p_acc[2 * jp + 1], ao, bv[2], bv[3]);
      }
    }
    // ---- softmax, delta = sum_t P dP, dS = P (dP - delta) scale for rows g and g + 8 (a row lives in one quad)
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const bool live = r0 + 16 * warp + g + 8 * half < q_len;
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < NJ; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e)
          if (8 * j + 2 * t + e < T) mx = fmaxf(mx, s_acc[j][2 * half + e]);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float off = mx * scale_log2;
      float l = 0.f;
#pragma unroll
      for (int j = 0; j < NJ; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pr = (8 * j + 2 * t + e < T) ? ex2_approx(fmaf(s_acc[j][2 * half + e], scale_log2, -off)) : 0.f;
          s_acc[j][2 * half + e] = pr;
          l += pr;
        }
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const float inv = live ? 1.0f / l : 0.f;       // rows past q_len contribute nothing
      float dl = 0.f;
#pragma unroll
      for (int j = 0; j < NJ; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pr = s_acc[j][2 * half + e] * inv;
          s_acc[j][2 * half + e] = pr;
          dl = fmaf(pr, p_acc[j][2 * half + e], dl);
        }
      dl += __shfl_xor_sync(0xffffffffu, dl, 1);
      dl += __shfl_xor_sync(0xffffffffu, dl, 2);
#pragma unroll
      for (int j = 0; j < NJ; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) p_acc[j][2 * half + e] = s_acc[j][2 * half + e] * (p_acc[j][2 * half + e] - dl) * scale;
    }
    uint32_t ds_pk[NJ][2];
#pragma unroll
    for (int j = 0; j < NJ; ++j)
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = 16 * warp + g + 8 * half, c = 8 * j + 2 * t;
        ds_pk[j][half] = pack_bf16x2(p_acc[j][2 * half], p_acc[j][2 * half + 1]);
        *reinterpret_cast<uint32_t*>(sP + r * PS + c) = pack_bf16x2(s_acc[j][2 * half], s_acc[j][2 * half + 1]);
        *reinterpret_cast<uint32_t*>(sS + r * PS + c) = ds_pk[j][half];
      }
    // ---- dQ[16 rows][CW] = dS[16 x 80] K[80 x CW]: A from the registers above
    if (dq != nullptr) {
      float d_acc[NT][4];
#pragma unroll
      for (int n = 0; n < NT; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) d_acc[n][e] = 0.f;
#pragma unroll
      for (int kk = 0; kk < NJ / 2; ++kk) {
        const uint32_t a[4] = {ds_pk[2 * kk][0], ds_pk[2 * kk][1], ds_pk[2 * kk + 1][0], ds_pk[2 * kk + 1][1]};
#pragma unroll
        for (int n = 0; n < NT; n += 2) {
          uint32_t bb[4];
          bw_ldsm_x4_t(bb, uK + ((kk * 16 + la_row) * LD + cs + 8 * n + la_col) * 2);
          mma_bf16_16816(d_acc[n], a, bb[0], bb[1]);
          mma_bf16_16816(d_acc[n + 1], a, bb[2], bb[3]);
        }
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = r0 + 16 * warp + g + 8 * half;
        if (r < q_len) {
          __nv_bfloat16* dst = dq + (static_cast<long long>(b) * q_len + r) * lddq + col0;
#pragma unroll
          for (int n = 0; n < NT; ++n) {
            const int c = cs + 8 * n + 2 * t;
            if (c < D) *reinterpret_cast<uint32_t*>(dst + c) = pack_bf16x2(d_acc[n][2 * half], d_acc[n][2 * half + 1]);
          }
        }
      }
    }
    __syncthreads();                              // every warp's P and dS rows are in shared memory
    // ---- dV[16 keys][CW] += P^T dO,  dK[16 keys][CW] += dS^T Q  (keys 16 w ..; A = the tiles read transposed)
#pragma unroll
    for (int kk = 0; kk < kAsRows / 16; ++kk) {
      uint32_t ap[4], as[4];
      const uint32_t a_off = ((kk * 16 + lb_row) * PS + 16 * warp + lb_col) * 2;
      bw_ldsm_x4_t(ap, uP + a_off);
      bw_ldsm_x4_t(as, uS + a_off);
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        uint32_t bo[4], bq[4];
        const uint32_t b_off = ((kk * 16 + la_row) * LD + cs + 8 * n + la_col) * 2;
        bw_ldsm_x4_t(bo, uO + b_off);
        bw_ldsm_x4_t(bq, uQ + b_off);
        mma_bf16_16816(acc_v[n], ap, bo[0], bo[1]);
        mma_bf16_16816(acc_v[n + 1], ap, bo[2], bo[3]);
        mma_bf16_16816(acc_k[n], as, bq[0], bq[1]);
        mma_bf16_16816(acc_k[n + 1], as, bq[2], bq[3]);
      }
    }
  }
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int key = 16 * warp + g + 8 * half;
    if (key < T) {
      const long long o = (static_cast<long long>(b) * T + key) * lddk + col0;
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        const int c = cs + 8 * n + 2 * t;
        if (c < D) {
          bw_red_add_v2(dk + o + c, acc_k[n][2 * half], acc_k[n][2 * half + 1]);
          bw_red_add_v2(dv + o + c, acc_v[n][2 * half], acc_v[n][2 * half + 1]);
        }
      }
    }
  }
}

template <int D, int kSplit>
int launch_attn_bwd_short_mma(const afb_attn_args& a, const __nv_bfloat16* dout, __nv_bfloat16* dq, float* dk, float* dv,
                              long long lddq, long long lddk, cudaStream_t stream) {
  using C = AsCfg<D, kSplit>;
  auto kern = attn_bwd_short_mma_kernel<D, kSplit>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(C::kSmem)));
    attr_set = true;
  }
  const int tiles = (a.q_len + kAsRows - 1) / kAsRows;
  // about two resident sets of blocks; every block keeps its dK / dV partial in registers over its tiles
  const int per_sm = C::kSmem + 1024 <= 113 * 1024 ? 2 : 1;
  int chunks = (2 * per_sm * num_sms() + a.batch * a.heads * kSplit - 1) / (a.batch * a.heads * kSplit);
  if (chunks > tiles) chunks = tiles;
  if (chunks < 1) chunks = 1;
  const int tiles_per_block = (tiles + chunks - 1) / chunks;
  dim3 grid(((tiles + tiles_per_block - 1) / tiles_per_block) * kSplit, a.heads, a.batch);
  kern<<<grid, kAsThreads, C::kSmem, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.q), reinterpret_cast<const __nv_bfloat16*>(a.k),
                                               reinterpret_cast<const __nv_bfloat16*>(a.v), dout, dq, dk, dv, a.q_len, a.kv_len, a.ldq,
                                               a.ldk, a.ldv, a.ldo, lddq, lddk, a.scale, tiles_per_block);
  return 0;
}

// AFB_ATTN_BWD: unset = the transposed-score version (mma3); "mma2" = the 128-key ldmatrix version, "simt" = the CUDA-core
// version (A/B measurement and cross-check); read once
int attn_bwd_mode() {      // 3 = mma3 (default), 2 = mma2, 0 = simt
  static const int mode = [] {
    const char* e = getenv("AFB_ATTN_BWD");
    if (e == nullptr) return 3;
    if (e[0] == 's') return 0;
    return (e[0] == 'm' && e[1] == 'm' && e[2] == 'a' && e[3] == '2') ? 2 : 3;
  }();
  return mode;
}


template <int D>
int launch_attn_bwd_long(const afb_attn_args& a, const __nv_bfloat16* dout, const float* delta, float* dq, float* dk, float* dv,
                         long long ldg, cudaStream_t stream) {
  const size_t sh = (4 * static_cast<size_t>(kAlT) * (D + 1) + 2 * static_cast<size_t>(kAlT) * (kAlT + 1) + 2 * kAlT) * sizeof(float);
  auto kern = attn_bwd_long_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sh)));
    attr_set = true;
  }
  dim3 grid((a.kv_len + kAlT - 1) / kAlT, a.heads, a.batch);
  kern<<<grid, kAlThreads, sh, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.q), reinterpret_cast<const __nv_bfloat16*>(a.k),
                                         reinterpret_cast<const __nv_bfloat16*>(a.v), dout, a.lse, delta, dq, dk, dv, a.q_len,
                                         a.kv_len, a.heads, a.ldq, a.ldk, a.ldv, a.ldo, ldg, a.scale);
  return 0;
}

}  // namespace
}  // namespace afb

extern "C" int afb_attention_bwd_short(const afb_attn_args* args, const afb_bf16* dout, afb_bf16* dq, float* dk, float* dv,
                                       int64_t lddq, int64_t lddk, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(args != nullptr && dout && dk && dv, "afb_attention_bwd_short: NULL argument");
  const afb_attn_args& a = *args;
  AFB_REQUIRE(a.q && a.k && a.v, "afb_attention_bwd_short: NULL operand");
  AFB_REQUIRE(a.key_mask == nullptr, "afb_attention_bwd_short: key masks are a self-attention feature (not built for the backward)");
  AFB_REQUIRE(a.batch > 0 && a.heads > 0 && a.q_len > 0 && a.kv_len > 0 && a.batch <= 65535 && a.heads <= 65535,
              "afb_attention_bwd_short: bad sizes");
  const int max_keys = a.head_dim >= 160 ? 80 : kAbMaxKeys;
  AFB_REQUIRE(a.kv_len <= max_keys, "afb_attention_bwd_short: kv_len %d > %d: only the short-key (cross-attention) backward is "
              "built; the self-attention backward is not", a.kv_len, max_keys);
  const int width = a.heads * a.head_dim;
  AFB_REQUIRE(a.ldq >= width && a.ldk >= width && a.ldv >= width && a.ldo >= width && lddk >= width && (dq == nullptr || lddq >= width),
              "afb_attention_bwd_short: leading dimension smaller than heads*head_dim");
  const __nv_bfloat16* db = reinterpret_cast<const __nv_bfloat16*>(dout);
  __nv_bfloat16* dqb = reinterpret_cast<__nv_bfloat16*>(dq);
  int rc;
  // the tensor-core version: up to 80 keys, 16-byte vector loads, bf16x2 dQ stores, 8-byte vector atomics for dK / dV
  // (AFB_ATTN_BWD=simt keeps the CUDA-core version for the A/B measurement)
  const bool mma = attn_bwd_mode() >= 2 && a.kv_len <= kAsKeys && a.head_dim % 8 == 0 &&
                   ((reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) | reinterpret_cast<uintptr_t>(a.v) |
                     reinterpret_cast<uintptr_t>(dout)) & 15) == 0 &&
                   a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0 && lddk % 2 == 0 &&
                   ((reinterpret_cast<uintptr_t>(dk) | reinterpret_cast<uintptr_t>(dv)) & 7) == 0 &&
                   (dq == nullptr || (lddq % 2 == 0 && (reinterpret_cast<uintptr_t>(dq) & 3) == 0));
#define AFB_BWD_CASE(DIM, SPLIT)                                                                      \
  case DIM:                                                                                           \
    rc = mma ? launch_attn_bwd_short_mma<DIM, SPLIT>(a, db, dqb, dk, dv, lddq, lddk, stream)          \
             : launch_attn_bwd<DIM>(a, db, dqb, dk, dv, lddq, lddk, stream);                          \
    break;
  switch (a.head_dim) {
    AFB_BWD_CASE(40, 1)
    AFB_BWD_CASE(80, 1)
    AFB_BWD_CASE(160, 2)
    default: AFB_FAIL(2, "afb_attention_bwd_short: head_dim %d not supported (SD-1.5 uses 40, 80, 160)", a.head_dim);
  }
#undef AFB_BWD_CASE
  if (rc) return rc;
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_attention_bwd(const afb_attn_args* args, const afb_bf16* dout, float* delta_ws, float* dq, float* dk,
                                 float* dv, int64_t ldg, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(args != nullptr && dout && delta_ws && dq && dk && dv, "afb_attention_bwd: NULL argument");
  const afb_attn_args& a = *args;
  AFB_REQUIRE(a.q && a.k && a.v && a.out && a.lse, "afb_attention_bwd: q, k, v, the forward output and its lse are needed");

  AFB_REQUIRE(a.batch > 0 && a.heads > 0 && a.q_len > 0 && a.kv_len > 0 && a.batch <= 65535 && a.heads <= 65535,
              "afb_attention_bwd: bad sizes");
  const int width = a.heads * a.head_dim;
  AFB_REQUIRE(a.ldq >= width && a.ldk >= width && a.ldv >= width && a.ldo >= width && ldg >= width,
              "afb_attention_bwd: leading dimension smaller than heads*head_dim");
  AFB_REQUIRE(a.head_dim % 8 == 0 && a.ldo % 8 == 0, "afb_attention_bwd: head_dim and ldo must be multiples of 8");
  const long long rows = static_cast<long long>(a.batch) * a.q_len;
  const __nv_bfloat16* db = reinterpret_cast<const __nv_bfloat16*>(dout);
  const long long n = rows * a.heads;
  // `dout` shares the forward output's row pitch (ldo)
  attn_delta_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(a.out), db,
                                                                                delta_ws, rows, a.q_len, a.heads, a.head_dim,
                                                                                a.ldo, a.ldo);
  count_launch();
  // dQ is accumulated with atomics: zero its `width` columns of every row (dq / dk / dv may be column blocks of one
  // [rows, 3 C] gradient buffer with ldg = 3 C: only the dQ block is cleared)
  if (ldg == width) AFB_CHECK_CUDA(cudaMemsetAsync(dq, 0, static_cast<size_t>(rows) * ldg * sizeof(float), stream));
  else AFB_CHECK_CUDA(cudaMemset2DAsync(dq, static_cast<size_t>(ldg) * sizeof(float), 0, static_cast<size_t>(width) * sizeof(float),
                                        static_cast<size_t>(rows), stream));
  int rc;
  // the tensor-core version reads its operands with 16-byte vector loads
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) | reinterpret_cast<uintptr_t>(a.v) |
                        reinterpret_cast<uintptr_t>(dout)) & 15) == 0 &&
                      a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0;
  // the tensor-core versions read 16-byte vectors, send dQ as 8-byte vector atomics and store dK / dV as float2
  const bool mma2 = (attn_bwd_mode() != 0 || a.key_mask != nullptr) && vec_ok && ldg % 2 == 0 &&
                    ((reinterpret_cast<uintptr_t>(dq) | reinterpret_cast<uintptr_t>(dk) | reinterpret_cast<uintptr_t>(dv)) & 7) == 0;
  AFB_REQUIRE(a.key_mask == nullptr || mma2, "afb_attention_bwd: the key-masked backward needs 16-byte aligned operands, "
              "8-byte aligned gradients and an even ldg");
  const bool mma3 = mma2 && attn_bwd_mode() == 3 && a.head_dim <= 80;      // (d = 160 reloads its K / V fragments: 72 vs 68 us)
#define AFB_BWD_CASE(DIM, SPLIT)                                                                                \
  case DIM:                                                                                                     \
    rc = mma3 ? launch_attn_bwd_mma3<DIM, SPLIT>(a, db, delta_ws, dq, dk, dv, ldg, stream)                      \
       : mma2 ? launch_attn_bwd_mma2<DIM, SPLIT>(a, db, delta_ws, dq, dk, dv, ldg, stream)                      \
              : launch_attn_bwd_long<DIM>(a, db, delta_ws, dq, dk, dv, ldg, stream);                            \
    break;
  switch (a.head_dim) {
    AFB_BWD_CASE(40, 1)
    AFB_BWD_CASE(80, 1)
    AFB_BWD_CASE(160, 2)
    default: AFB_FAIL(2, "afb_attention_bwd: head_dim %d not supported (SD-1.5 uses 40, 80, 160)", a.head_dim);
  }
#undef AFB_BWD_CASE
  if (rc) return rc;
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

namespace afb {
namespace {

constexpr int kGnBwdMaxCpg = 128;      // channels per group the backward stages (SD-1.5: 10 .. 80)

// ---- GroupNorm backward (attention.py:71-72,325: GroupNorm(32, C, eps), token-major bf16 in and out; kSilu: the
// ResBlocks' GroupNorm32 -> SiLU with the conv-bias / time-embedding pre-add, openaimodel.py:259-279): block (group,
// sample) walks the group's cpg channels of all tokens three times (the slab is L2 resident) with 4-byte loads:
// statistics of u = x + pre_add, then the two means of the LayerNorm-style backward, then
//   dx = rstd (gg - mean(gg) - x_hat mean(gg x_hat)),  gg = gamma g  (times silu'(gamma x_hat + beta) when fused).
template <bool kSilu>
__global__ void __launch_bounds__(512) gn_bwd_tok_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ g,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         const float* __restrict__ pre_add, long long add_ld,
                                                         __nv_bfloat16* __restrict__ dx, int C, int HW, int cpg, float eps) {
  __shared__ float red[32];
  __shared__ float s_gamma[kGnBwdMaxCpg], s_beta[kGnBwdMaxCpg], s_pre[kGnBwdMaxCpg];
  const int grp = blockIdx.x, b = blockIdx.y;
  const long long base = static_cast<long long>(b) * HW * C + grp * cpg;
  if (threadIdx.x < cpg) {
    s_gamma[threadIdx.x] = gamma[grp * cpg + threadIdx.x];
    s_beta[threadIdx.x] = beta != nullptr ? beta[grp * cpg + threadIdx.x] : 0.f;
    s_pre[threadIdx.x] = pre_add != nullptr ? pre_add[b * add_ld + grp * cpg + threadIdx.x] : 0.f;
  }
  __syncthreads();
  const int half = cpg >> 1;                     // channel pairs per token (cpg is even: 4-byte loads)
  const int n2 = HW * half;
  const int n = HW * cpg;
  float s = 0.f, s2 = 0.f;
  for (int idx = threadIdx.x; idx < n2; idx += blockDim.x) {
    const int t = idx / half, j = 2 * (idx - t * half);
    const float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + base + static_cast<long long>(t) * C + j));
    const float a0 = v.x + s_pre[j], a1 = v.y + s_pre[j + 1];
    s += a0 + a1;
    s2 = fmaf(a0, a0, fmaf(a1, a1, s2));
  }
  const float mean = block_sum_t(s, red) / n;
  const float var = fmaxf(block_sum_t(s2, red) / n - mean * mean, 0.f);
  const float rstd = rsqrtf(var + eps);
  // gamma * dL/dz of one element: through the SiLU when fused (z = gamma x_hat + beta, silu'(z) = s (1 + z (1 - s)))
  auto gz = [&](float xh, float gy, int j) {
    float d = gy;
    if (kSilu) {
      const float z = fmaf(s_gamma[j], xh, s_beta[j]);
      const float sg = 1.0f / (1.0f + __expf(-z));
      d *= sg * fmaf(z, 1.0f - sg, 1.0f);
    }
    return d * s_gamma[j];
  };
  float sg = 0.f, sgx = 0.f;
  for (int idx = threadIdx.x; idx < n2; idx += blockDim.x) {
    const int t = idx / half, j = 2 * (idx - t * half);
    const long long o = base + static_cast<long long>(t) * C + j;
    const float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + o));
    const float2 gy = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(g + o));
    const float xh0 = (v.x + s_pre[j] - mean) * rstd, xh1 = (v.y + s_pre[j + 1] - mean) * rstd;
    const float g0 = gz(xh0, gy.x, j), g1 = gz(xh1, gy.y, j + 1);
    sg += g0 + g1;
    sgx = fmaf(g0, xh0, fmaf(g1, xh1, sgx));
  }
  const float mg = block_sum_t(sg, red) / n, mgx = block_sum_t(sgx, red) / n;
  for (int idx = threadIdx.x; idx < n2; idx += blockDim.x) {
    const int t = idx / half, j = 2 * (idx - t * half);
    const long long o = base + static_cast<long long>(t) * C + j;
    const float2 v = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(x + o));
    const float2 gy = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(g + o));
    const float xh0 = (v.x + s_pre[j] - mean) * rstd, xh1 = (v.y + s_pre[j + 1] - mean) * rstd;
    const float g0 = gz(xh0, gy.x, j), g1 = gz(xh1, gy.y, j + 1);
    *reinterpret_cast<uint32_t*>(dx + o) = pack_bf16x2(rstd * (g0 - mg - xh0 * mgx), rstd * (g1 - mg - xh1 * mgx));
  }
}

// ---- streaming version for large slabs (the ResBlocks' and SpatialTransformers' GroupNorms at 32x32 and above) --------
// The slab kernel above reads 20..160-byte channel runs of a (sample, group) with 4-byte loads, 256 blocks per launch:
// 86 us for a 21 MB tensor (b8 n4096 C320).  Here every block streams whole token rows (16-byte loads, all channels) of
// a token range; a thread owns one 8-channel column over its rows, so the per-channel constants live in registers.
// Three passes: per-group sums of u = x + pre_add and u^2; per-group sums of gg and gg x_hat; dx.  Per-group partials go
// through shared memory into fp32 atomics on a [B, G, 4] workspace (zeroed by the entry point).
template <int kPass, bool kSilu>
__global__ void __launch_bounds__(320) gn_bwd_stream_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ g,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            const float* __restrict__ pre_add, long long add_ld,
                                                            float* __restrict__ ws, __nv_bfloat16* __restrict__ dx, int C, int HW,
                                                            int cpg, float eps, int tokens_per_block, int rpi) {
  extern __shared__ float sh[];                 // [rpi][C][2] per-channel partials of this block's rows
  __shared__ float sg[2 * 64];                  // per-group partials
  const int b = blockIdx.y, G = C / cpg;
  const int nvec = C >> 3;
  const int cv = threadIdx.x % nvec, r = threadIdx.x / nvec;
  const int c0 = cv * 8;
  const int t0 = blockIdx.x * tokens_per_block;
  const int t1 = min(HW, t0 + tokens_per_block);
  const float inv_n = 1.0f / (static_cast<float>(HW) * cpg);
  float ga[8], be[8], pa[8], mean[8], rstd[8], mg[8], mgx[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int ch = c0 + j, grp = ch / cpg;
    pa[j] = pre_add != nullptr ? pre_add[b * add_ld + ch] : 0.f;
    ga[j] = be[j] = mean[j] = rstd[j] = mg[j] = mgx[j] = 0.f;
    if (kPass >= 1) {
      const float* w = ws + (static_cast<long long>(b) * G + grp) * 4;
      ga[j] = gamma[ch];
      be[j] = beta != nullptr ? beta[ch] : 0.f;
      mean[j] = w[0] * inv_n;
      rstd[j] = rsqrtf(fmaxf(w[1] * inv_n - mean[j] * mean[j], 0.f) + eps);
      if (kPass == 2) {
        mg[j] = w[2] * inv_n;
        mgx[j] = w[3] * inv_n;
      }
    }
  }
  float a0[8], a1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a0[j] = a1[j] = 0.f;
  // two rows per iteration: both rows' 16-byte loads are in flight before either is consumed
  for (int t = t0 + r; t < t1; t += 2 * rpi) {
    const bool two = t + rpi < t1;
    const long long o0 = (static_cast<long long>(b) * HW + t) * C + c0;
    const long long o1 = two ? o0 + static_cast<long long>(rpi) * C : o0;
    uint4 xr[2], gr[2];
    xr[0] = *reinterpret_cast<const uint4*>(x + o0);
    xr[1] = *reinterpret_cast<const uint4*>(x + o1);
    if (kPass >= 1) {
      gr[0] = *reinterpret_cast<const uint4*>(g + o0);
      gr[1] = *reinterpret_cast<const uint4*>(g + o1);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !two) break;
      float xv[8], gv[8], out[8];
      unpack8f(xr[h], xv);
      if (kPass >= 1) unpack8f(gr[h], gv);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float u = xv[j] + pa[j];
        if (kPass == 0) {
          a0[j] += u;
          a1[j] = fmaf(u, u, a1[j]);
        } else {
          const float xh = (u - mean[j]) * rstd[j];
          float d = gv[j];
          if (kSilu) {
            const float z = fmaf(ga[j], xh, be[j]);
            const float s = 1.0f / (1.0f + __expf(-z));
            d *= s * fmaf(z, 1.0f - s, 1.0f);
          }
          const float gg = d * ga[j];
          if (kPass == 1) {
            a0[j] += gg;
            a1[j] = fmaf(gg, xh, a1[j]);
          } else {
            out[j] = rstd[j] * (gg - mg[j] - xh * mgx[j]);
          }
        }
      }
      if (kPass == 2) *reinterpret_cast<uint4*>(dx + (h == 0 ? o0 : o1)) = pack8f(out);
    }
  }
  if (kPass == 2) return;
  if (threadIdx.x < 2 * G) sg[threadIdx.x] = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sh[(r * C + c0 + j) * 2] = a0[j];
    sh[(r * C + c0 + j) * 2 + 1] = a1[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int rr = 0; rr < rpi; ++rr) {
      s0 += sh[(rr * C + c) * 2];
      s1 += sh[(rr * C + c) * 2 + 1];
    }
    atomicAdd(&sg[2 * (c / cpg)], s0);
    atomicAdd(&sg[2 * (c / cpg) + 1], s1);
  }
  __syncthreads();
  if (threadIdx.x < 2 * G)
    atomicAdd(ws + (static_cast<long long>(b) * G + (threadIdx.x >> 1)) * 4 + 2 * kPass + (threadIdx.x & 1), sg[threadIdx.x]);
}

// AFB_GN_BWD=slab keeps the (sample, group) slab kernel for every shape (A/B)
bool gn_bwd_stream_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_GN_BWD");
    return e == nullptr || !(e[0] == 's' && e[1] == 'l');
  }();
  return on;
}

template <bool kSilu>
int launch_gn_bwd_stream(const __nv_bfloat16* x, const __nv_bfloat16* g, const float* gamma, const float* beta, const float* pre_add,
                         long long add_ld, float* ws, __nv_bfloat16* dx, int batch, int C, int HW, int groups, float eps,
                         cudaStream_t stream) {
  const int nvec = C / 8;
  const int rpi = nvec >= 256 ? 1 : 256 / nvec;
  const int threads = nvec * rpi;
  // about two blocks per SM: a thread sets up the constants of its 8 channels once and should then walk >= 16 rows
  long long tpb = (static_cast<long long>(batch) * HW + 2 * num_sms() - 1) / (2 * num_sms());
  if (tpb < rpi) tpb = rpi;
  tpb = (tpb + rpi - 1) / rpi * rpi;
  const dim3 grid(static_cast<unsigned>((HW + tpb - 1) / tpb), batch);
  const size_t sh = static_cast<size_t>(rpi) * C * 2 * sizeof(float);
  AFB_CHECK_CUDA(cudaMemsetAsync(ws, 0, static_cast<size_t>(batch) * groups * 4 * sizeof(float), stream));
  const int cpg = C / groups, t = static_cast<int>(tpb);
  gn_bwd_stream_kernel<0, kSilu><<<grid, threads, sh, stream>>>(x, g, gamma, beta, pre_add, add_ld, ws, dx, C, HW, cpg, eps, t, rpi);
  gn_bwd_stream_kernel<1, kSilu><<<grid, threads, sh, stream>>>(x, g, gamma, beta, pre_add, add_ld, ws, dx, C, HW, cpg, eps, t, rpi);
  gn_bwd_stream_kernel<2, kSilu><<<grid, threads, 0, stream>>>(x, g, gamma, beta, pre_add, add_ld, ws, dx, C, HW, cpg, eps, t, rpi);
  return 0;
}

}  // namespace
}  // namespace afb

extern "C" int afb_groupnorm_tokens_bwd(const afb_bf16* x, const afb_bf16* grad_y, const float* gamma, const float* beta,
                                        const float* pre_add, int64_t add_ld, int act, afb_bf16* dx, float* ws, int batch,
                                        int channels, int hw, int groups, float eps, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && grad_y && gamma && dx, "afb_groupnorm_tokens_bwd: NULL argument");
  AFB_REQUIRE(batch > 0 && batch <= 65535 && channels > 0 && hw > 0 && groups > 0 && channels % groups == 0,
              "afb_groupnorm_tokens_bwd: bad sizes");
  const int cpg = channels / groups;
  AFB_REQUIRE(cpg % 2 == 0 && cpg <= kGnBwdMaxCpg, "afb_groupnorm_tokens_bwd: channels per group %d must be even and <= %d", cpg,
              kGnBwdMaxCpg);
  AFB_REQUIRE(act == 0 || act == 1, "afb_groupnorm_tokens_bwd: act must be 0 (none) or 1 (SiLU)");
  AFB_REQUIRE(act == 0 || beta != nullptr, "afb_groupnorm_tokens_bwd: the SiLU backward needs beta");
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(grad_y) | reinterpret_cast<uintptr_t>(dx)) & 3) == 0,
              "afb_groupnorm_tokens_bwd: buffers must be 4-byte aligned");
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(grad_y);
  __nv_bfloat16* db = reinterpret_cast<__nv_bfloat16*>(dx);
  // large slabs: three streaming passes over whole token rows (needs the [B, G, 4] workspace and 16-byte rows)
  const bool stream_ok = ws != nullptr && gn_bwd_stream_enabled() && channels % 8 == 0 && channels / 8 <= 320 && groups <= 64 &&
                         cpg >= 8 && hw >= 2048 &&      // (measured: below ~2k tokens per sample the slab kernel is faster)
                         ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(grad_y) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  if (stream_ok) {
    const int rc = act == 1 ? launch_gn_bwd_stream<true>(xb, gb, gamma, beta, pre_add, add_ld, ws, db, batch, channels, hw, groups, eps, stream)
                            : launch_gn_bwd_stream<false>(xb, gb, gamma, beta, pre_add, add_ld, ws, db, batch, channels, hw, groups, eps, stream);
    if (rc) return rc;
  } else if (act == 1)
    gn_bwd_tok_kernel<true><<<dim3(groups, batch), 512, 0, stream>>>(xb, gb, gamma, beta, pre_add, add_ld, db, channels, hw, cpg, eps);
  else
    gn_bwd_tok_kernel<false><<<dim3(groups, batch), 512, 0, stream>>>(xb, gb, gamma, beta, pre_add, add_ld, db, channels, hw, cpg, eps);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

namespace afb {
namespace {

// ---- AttentionalPooler backward towards the layer features (embedding_manager.py:146-321): with the quantities of
// ada_pool_bwd_kernel, dv_t[c] = sum_i (ds_i,g(c)[t] u_i[c] + p_i,g(c)[t] dout_i[c]); v = (x + LN(q)) / sqrt(8), so
// dx_t = dv_t / sqrt(8) and dq_t = LayerNorm'(dv_t / sqrt(8)) (no affine).  One warp per token, rows written as bf16.
// (AdaEmbedding builds its poolers with infeat_grad_scale = 1, embedding_manager.py:667: the gradient is not scaled.)
template <int kChunks>
__global__ void __launch_bounds__(kPbThreads) ada_pool_bwd_feat_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ q, const float* __restrict__ u,
    const uint8_t* __restrict__ key_mask, const float* __restrict__ stats, const float* __restrict__ out_fg,
    const float* __restrict__ out_bg, const float* __restrict__ d_fg, const float* __restrict__ d_bg,
    __nv_bfloat16* __restrict__ dx, __nv_bfloat16* __restrict__ dqo, int N, int C, int tokens_per_block, float drop_p,
    uint32_t seed) {
  extern __shared__ float sh[];
  float* s_u = sh;                       // [2][C]
  float* s_d = s_u + 2 * C;              // [2][C] dout
  float* s_part = s_d + 2 * C;           // [8 warps][4][kChunks * 32]
  __shared__ float s_delta[kHeads], s_m[kHeads], s_l[kHeads];
  const int dg = C / kHeads;
  const int b = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int idx = threadIdx.x; idx < 2 * C; idx += blockDim.x) {
    s_u[idx] = u[static_cast<long long>(b) * 2 * C + idx];
    s_d[idx] = idx < C ? d_fg[static_cast<long long>(b) * C + idx] : d_bg[static_cast<long long>(b) * C + idx - C];
  }
  if (threadIdx.x < kHeads) {
    s_m[threadIdx.x] = stats[(static_cast<long long>(b) * kHeads + threadIdx.x) * 2];
    s_l[threadIdx.x] = stats[(static_cast<long long>(b) * kHeads + threadIdx.x) * 2 + 1];
  }
  __syncthreads();
  if (warp < kHeads) {
    float a = 0.f;
    for (int c = lane; c < dg; c += 32) {
      const int cc = warp * dg + c;
      a += s_d[cc] * out_fg[static_cast<long long>(b) * C + cc] + s_d[C + cc] * out_bg[static_cast<long long>(b) * C + cc];
    }
    a = warp_sum(a);
    if (lane == 0) s_delta[warp] = a;
  }
  __syncthreads();
  const int nchunks = C >> 3;
  const int chunks_per_group = dg >> 3;
  float* my_part = s_part + warp * 4 * (kChunks * 32);
  const int t_begin = blk * tokens_per_block;
  const int t_end = min(N, t_begin + tokens_per_block);
  for (int t = t_begin + warp; t < t_end; t += kPbWarps) {
    const long long rowoff = (static_cast<long long>(b) * N + t) * C;
    const uint4* xr = reinterpret_cast<const uint4*>(x + rowoff);
    const uint4* qr = reinterpret_cast<const uint4*>(q + rowoff);
    float vv[kChunks][8], kh[kChunks][8];        // v, then dv / sqrt(8);  q, then LN(q)
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        unpack8f(xr[ch], vv[i]);
        unpack8f(qr[ch], kh[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += kh[i][j];
      }
    }
    const float mean = warp_sum(s) / C;
    float m2 = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i)
      if (lane + 32 * i < nchunks)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = kh[i][j] - mean;
          m2 += d * d;
        }
    const float rstd = rsqrtf(warp_sum(m2) / C + 1e-5f);
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        float a_fg = 0.f, a_bg = 0.f, p_fg = 0.f, p_bg = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          kh[i][j] = (kh[i][j] - mean) * rstd;
          vv[i][j] = (vv[i][j] + kh[i][j]) * kInvSqrt8;
          a_fg += s_u[ch * 8 + j] * vv[i][j];
          a_bg += s_u[C + ch * 8 + j] * vv[i][j];
          p_fg += s_d[ch * 8 + j] * vv[i][j];
          p_bg += s_d[C + ch * 8 + j] * vv[i][j];
        }
        my_part[ch] = a_fg;
        my_part[kChunks * 32 + ch] = a_bg;
        my_part[2 * kChunks * 32 + ch] = p_fg;
        my_part[3 * kChunks * 32 + ch] = p_bg;
      }
    }
    __syncwarp();
    float sc = 0.f, dp = 0.f;
    if (lane < 16) {
      const float* pp = my_part + (lane >> 3) * (kChunks * 32) + (lane & 7) * chunks_per_group;
      const float* pd = pp + 2 * (kChunks * 32);
      for (int kk = 0; kk < chunks_per_group; ++kk) {
        sc += pp[kk];
        dp += pd[kk];
      }
    }
    __syncwarp();
    float tot = sc;
    tot += __shfl_xor_sync(0xffffffffu, tot, 1);
    tot += __shfl_xor_sync(0xffffffffu, tot, 2);
    tot += __shfl_xor_sync(0xffffffffu, tot, 4);
    sc = 0.5f * sc + 0.5f * (tot * 0.125f);
    const bool masked = key_mask != nullptr && key_mask[static_cast<long long>(b) * N + t] == 0;
    const int g = lane & 7;
    float pr = 0.f;
    if (lane < 16 && !masked) pr = __expf(sc - s_m[g]) / s_l[g];
    const float prk = pr * pool_drop_scale(seed, b, g, (lane >> 3) & 1, t, N, drop_p);      // kept probability (train-mode dropout)
    const float dsp = prk * dp - pr * s_delta[g];
    float dtot = dsp;
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 1);
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 2);
    dtot += __shfl_xor_sync(0xffffffffu, dtot, 4);
    const float ds = 0.5f * dsp + 0.5f * (dtot * 0.125f);
    float sg = 0.f, sga = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      const bool ok = ch < nchunks;
      const int gg = ok ? ch / chunks_per_group : 0;
      const float f_ds = __shfl_sync(0xffffffffu, ds, gg), b_ds = __shfl_sync(0xffffffffu, ds, 8 + gg);
      const float f_p = __shfl_sync(0xffffffffu, prk, gg), b_p = __shfl_sync(0xffffffffu, prk, 8 + gg);
      if (ok) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = ch * 8 + j;
          const float gx = (f_ds * s_u[c] + b_ds * s_u[C + c] + f_p * s_d[c] + b_p * s_d[C + c]) * kInvSqrt8;
          vv[i][j] = gx;
          sg += gx;
          sga = fmaf(gx, kh[i][j], sga);
        }
      }
    }
    const float mg = warp_sum(sg) / C, mga = warp_sum(sga) / C;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        float o[8];
        reinterpret_cast<uint4*>(dx + rowoff)[ch] = pack8f(vv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (vv[i][j] - mg - kh[i][j] * mga);
        reinterpret_cast<uint4*>(dqo + rowoff)[ch] = pack8f(o);
      }
    }
  }
}

}  // namespace
}  // namespace afb

extern "C" int afb_ada_pool_bwd_features(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask,
                                         const float* stats, const float* fg_out, const float* bg_out, const float* d_fg,
                                         const float* d_bg, afb_bf16* dx, afb_bf16* dq, int batch, int n_tokens, int channels,
                                         float drop_p, uint32_t seed, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && q && pool_q && stats && fg_out && bg_out && d_fg && d_bg && dx && dq, "afb_ada_pool_bwd_features: NULL argument");
  AFB_REQUIRE(batch > 0 && n_tokens > 0 && channels % 64 == 0 && channels <= 1280, "afb_ada_pool_bwd_features: bad sizes");
  const uintptr_t all = reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(dx) |
                        reinterpret_cast<uintptr_t>(dq);
  AFB_REQUIRE((all & 15) == 0, "afb_ada_pool_bwd_features: x / q / dx / dq must be 16-byte aligned");
  const int C = channels;
  int blocks = 4 * num_sms() / batch;
  const int max_blocks = (n_tokens + 7) / 8;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  const int tokens_per_block = (n_tokens + blocks - 1) / blocks;
  dim3 grid((n_tokens + tokens_per_block - 1) / tokens_per_block, batch);
  const int cpl = (C / 8 + 31) / 32;
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(q);
  __nv_bfloat16* dxb = reinterpret_cast<__nv_bfloat16*>(dx);
  __nv_bfloat16* dqb = reinterpret_cast<__nv_bfloat16*>(dq);
#define AFB_LAUNCH_PBF(KCH)                                                                                         \
  {                                                                                                                 \
    auto kern = ada_pool_bwd_feat_kernel<KCH>;                                                                      \
    const size_t shb = (4 * static_cast<size_t>(C) + kPbWarps * 4 * (KCH * 32)) * sizeof(float);                    \
    static size_t set_for = 0;                                                                                      \
    if (set_for < shb) {                                                                                            \
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(shb))); \
      set_for = shb;                                                                                                \
    }                                                                                                               \
    kern<<<grid, kPbThreads, shb, stream>>>(xb, qb, pool_q, key_mask, stats, fg_out, bg_out, d_fg, d_bg, dxb, dqb, n_tokens, C, \
                                            tokens_per_block, drop_p, seed);                                        \
  }
  if (cpl <= 2) AFB_LAUNCH_PBF(2)
  else if (cpl <= 3) AFB_LAUNCH_PBF(3)
  else AFB_LAUNCH_PBF(5)
#undef AFB_LAUNCH_PBF
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
