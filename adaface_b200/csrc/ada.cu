// K12/K13: AttentionalPooler + AdaEmbedding head (ldm/modules/embedding_manager.py:100-321, 539-910).
//
// K12 streams the layer's x and q ([B,N,C] bf16) exactly once (HBM-bound: 2*N*C*2 bytes per sample).
//
// Algebra used (exact, eval mode).  With v_t = (x_t + LN(q_t)) / sqrt(8) (:160-178), the reference computes
// lora_k_t = conv_g(v_t) / sqrt(8) (:227, grouped 1x1 conv, 8 groups, C/8 -> C/64 per group) and the per-head
// scores <lora_q[i][g], lora_k_t[g]> (:247).  The conv is linear, so the score of query i and head g is
//      s[i][g][t] = sum_{c in head g} u[i][c] * v_t[c],   u[i][c] = sum_o lora_q[i][g][o] * W_k[g][o][c] / sqrt(8)
// i.e. TWO DOT PRODUCTS per head per token against a per-sample vector u that the step-invariant prep computes once.
//
//   prep    ada_qproj:   proj[s] = to_k(query embedding s), one warp per output channel, weight read once
//           ada_qfinish: LayerNorm, grouped conv of the queries (:208-235), then the fold through lora_to_k -> u [B,2,C]
//   main    one warp per token: LN(q), v, the 16 head dots, head-average smoothing (:251-257), then a flash-style
//           online softmax over the joint (fg|bg) x token axis (:273-286) accumulating the fg- and bg-weighted
//           sums of v; per-block partials
//   combine exact log-sum-exp merge of the block partials -> fg_out, bg_out [B, C] (+ softmax stats)
//   probs   optional: attn_fg/attn_bg [B, 8, N] from the stored scores and the merged stats
// K13: block (k, b): LNCat, the r coefficients of vector k, basis combination, per-vector LN, bias; optionally
// writes the vector as bf16 straight into the layer's context rows (the subject-token splice).
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

constexpr int kHeads = 8;
constexpr int kPoolThreads = 256;          // 8 warps
constexpr int kPoolWarps = kPoolThreads / 32;
constexpr int kMaxChunksPerLane = 5;       // C <= 1280: C/8 16-byte chunks spread over 32 lanes
constexpr float kInvSqrt8 = 0.35355339059327373f;
constexpr float kPoolMasked = -3.0e38f;

__device__ __forceinline__ float block_sum_256(float v, float* red) {
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

// ---- prep, kernel 1: proj[s][c] = to_k(query embedding s)[c] (:193, :203); s = 0 is the shared fg probe, s = 1 + b
// the bg query of sample b.  One warp per output channel: the weight row is read once (float4, all loads in flight)
// and dotted with all query vectors (staged in shared memory).
constexpr int kPrepMaxSrc = 9;     // sources per pass (B + 1 = 9 for 8 samples); more -> extra passes
__device__ __forceinline__ void ada_qproj_body(
    const float* __restrict__ to_k_w, const float* __restrict__ fg_q_emb, const float* __restrict__ bg_q_emb,
    float* __restrict__ proj, int C, int ctx_dim, int n_src) {
  extern __shared__ float sh[];      // [min(n_src, kPrepMaxSrc)][ctx_dim]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * kPoolWarps + warp;
  const int n4 = ctx_dim >> 2;
  for (int s0 = 0; s0 < n_src; s0 += kPrepMaxSrc) {
    const int ns = min(kPrepMaxSrc, n_src - s0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < ns * ctx_dim; idx += blockDim.x) {
      const int s = s0 + idx / ctx_dim, k = idx % ctx_dim;
      sh[idx] = (s == 0) ? fg_q_emb[k] : bg_q_emb[static_cast<long long>(s - 1) * ctx_dim + k];
    }
    __syncthreads();
    if (c < C) {
      float acc[kPrepMaxSrc];
#pragma unroll
      for (int i = 0; i < kPrepMaxSrc; ++i) acc[i] = 0.f;
      const float4* wr = reinterpret_cast<const float4*>(to_k_w + static_cast<long long>(c) * ctx_dim);
#pragma unroll 2
      for (int k4 = lane; k4 < n4; k4 += 32) {
        const float4 w = __ldg(wr + k4);
#pragma unroll
        for (int i = 0; i < kPrepMaxSrc; ++i) {
          if (i < ns) {
            const float4 e = *reinterpret_cast<const float4*>(sh + i * ctx_dim + 4 * k4);
            acc[i] += w.x * e.x + w.y * e.y + w.z * e.z + w.w * e.w;
          }
        }
      }
#pragma unroll
      for (int i = 0; i < kPrepMaxSrc; ++i) {
        if (i < ns) {
          const float a = warp_sum(acc[i]);
          if (lane == 0) proj[static_cast<long long>(s0 + i) * C + c] = a;
        }
      }
    }
  }
}

__global__ void __launch_bounds__(kPoolThreads) ada_qproj_kernel(
    const float* __restrict__ to_k_w, const float* __restrict__ fg_q_emb, const float* __restrict__ bg_q_emb,
    float* __restrict__ proj, int C, int ctx_dim, int n_src) {
  ada_qproj_body(to_k_w, fg_q_emb, bg_q_emb, proj, C, ctx_dim, n_src);
}
// all (layer, subject) items of one U-Net call in one launch: blockIdx.y = item (channels differ per item)
__global__ void __launch_bounds__(kPoolThreads) ada_qproj_batched_kernel(const afb_ada_prep_item* __restrict__ items,
                                                                        int ctx_dim, int n_src) {
  const afb_ada_prep_item it = items[blockIdx.y];
  if (static_cast<int>(blockIdx.x) * kPoolWarps >= it.channels) return;
  ada_qproj_body(it.to_k_weight, it.fg_q_emb, it.bg_q_emb, it.ws, it.channels, ctx_dim, n_src);
}

// ---- prep, kernel 2: block (b, i): LayerNorm (no affine, :208-209) of the projected query, grouped 1x1 conv
// (:219, :222) with the 8^-1/2 * 8^-1/2 factor (:235) -> lora_q [dg], then the fold through lora_to_k:
// u[b][i][c] = 8^-1/2 * sum_o lora_q[g*lg + o] * W_k[g*lg + o][c - g*dg]     (g = c / dg)
__device__ __forceinline__ void ada_qfinish_body(const float* __restrict__ proj,
                                                 const float* __restrict__ w_fg,
                                                 const float* __restrict__ w_bg,
                                                 const float* __restrict__ w_k,
                                                 float* __restrict__ u, int C) {
  extern __shared__ float sh[];      // [C] normalised query, [dg] lora_q
  __shared__ float red[32];
  const int b = blockIdx.x, i = blockIdx.y;
  const int dg = C / kHeads, lg = dg / kHeads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* s_q = sh;
  float* s_lq = sh + C;
  const float* src = proj + static_cast<long long>(i == 0 ? 0 : 1 + b) * C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float v = src[c];
    s_q[c] = v;
    s += v;
  }
  const float mean = block_sum_256(s, red) / C;
  float m2 = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float d = s_q[c] - mean;
    m2 += d * d;
  }
  const float rstd = rsqrtf(block_sum_256(m2, red) / C + 1e-5f);
  for (int c = threadIdx.x; c < C; c += blockDim.x) s_q[c] = (s_q[c] - mean) * rstd;
  __syncthreads();
  const float* wsel = (i == 0) ? w_fg : w_bg;
  for (int j = warp; j < dg; j += kPoolWarps) {      // j = g*lg + o; one warp per output, coalesced weight row
    const int g = j / lg;
    const float* w = wsel + static_cast<long long>(j) * dg;
    float a = 0.f;
    for (int k = lane; k < dg; k += 32) a += w[k] * s_q[g * dg + k];
    a = warp_sum(a);
    if (lane == 0) s_lq[j] = a * kInvSqrt8 * kInvSqrt8;
  }
  __syncthreads();
  float* out = u + (static_cast<long long>(b) * 2 + i) * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const int g = c / dg, k = c - g * dg;
    float a = 0.f;
    for (int o = 0; o < lg; ++o) a += s_lq[g * lg + o] * w_k[static_cast<long long>(g * lg + o) * dg + k];
    out[c] = a * kInvSqrt8;
  }
}

__global__ void __launch_bounds__(kPoolThreads) ada_qfinish_kernel(const float* __restrict__ proj,
                                                                  const float* __restrict__ w_fg,
                                                                  const float* __restrict__ w_bg,
                                                                  const float* __restrict__ w_k,
                                                                  float* __restrict__ u, int C) {
  ada_qfinish_body(proj, w_fg, w_bg, w_k, u, C);
}
__global__ void __launch_bounds__(kPoolThreads) ada_qfinish_batched_kernel(const afb_ada_prep_item* __restrict__ items) {
  const afb_ada_prep_item it = items[blockIdx.z];
  ada_qfinish_body(it.ws, it.lora_to_fg_q, it.lora_to_bg_q, it.lora_to_k, it.pool_q, it.channels);
}

// partial record per block: m[8], l[8], acc_fg[C], acc_bg[C]
__host__ __device__ inline long long pool_partial_floats(int C) { return 16 + 2ll * C; }

template <int kChunks>
__global__ void __launch_bounds__(kPoolThreads, (kChunks <= 2 ? 2 : 1)) ada_pool_main_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ q, const float* __restrict__ u,
    const uint8_t* __restrict__ key_mask, float* __restrict__ scores, float* __restrict__ partials, int N, int C,
    int tokens_per_block, float drop_p, uint32_t seed) {
  extern __shared__ float sh[];
  float* s_u = sh;                                      // [2][C] folded queries of this sample
  float* s_part = s_u + 2 * C;                          // [8 warps][2][kChunks * 32] per-chunk dots
  float* s_red = sh;                                    // [8 warps][16 + 2C], reuses the buffers after the token loop
  pdl_wait();
  pdl_trigger();

  const int dg = C / kHeads;
  const int b = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int idx = threadIdx.x; idx < 2 * C; idx += blockDim.x) s_u[idx] = u[static_cast<long long>(b) * 2 * C + idx];
  __syncthreads();

  const int nchunks = C >> 3;
  const int chunks_per_group = dg >> 3;
  float acc_fg[kChunks][8], acc_bg[kChunks][8];
#pragma unroll
  for (int i = 0; i < kChunks; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_fg[i][j] = acc_bg[i][j] = 0.f;
  // every lane tracks the running max / sum of head lane & 7 (lanes 0..7 are the ones written out)
  float m_run = -INFINITY, l_run = 0.f;

  float* my_part = s_part + warp * 2 * (kChunks * 32);
  const int t_begin = blk * tokens_per_block;
  const int t_end = min(N, t_begin + tokens_per_block);
  // The per-token chain (two warp reductions, the head dots through shared memory, the softmax update) is long and
  // serial, and a warp has one token in flight: the next token's rows are requested before the current one is
  // processed, or the loop runs at one DRAM latency per token (kChunks == 5 has no registers left for it).
  constexpr bool kPrefetch = kChunks <= 3;
  uint4 nx[kChunks], nq[kChunks];
  auto load_rows = [&](int t, uint4 (&dx)[kChunks], uint4 (&dq)[kChunks]) {
    const long long rowoff = (static_cast<long long>(b) * N + t) * C;
    const uint4* xr = reinterpret_cast<const uint4*>(x + rowoff);
    const uint4* qr = reinterpret_cast<const uint4*>(q + rowoff);
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      if (ch < nchunks) {
        dx[i] = xr[ch];
        dq[i] = qr[ch];
      }
    }
  };
  if (kPrefetch && t_begin + warp < t_end) load_rows(t_begin + warp, nx, nq);
  for (int t = t_begin + warp; t < t_end; t += kPoolWarps) {
    uint4 ux[kChunks], uq[kChunks];
    if constexpr (kPrefetch) {
#pragma unroll
      for (int i = 0; i < kChunks; ++i) {
        ux[i] = nx[i];
        uq[i] = nq[i];
      }
      if (t + kPoolWarps < t_end) load_rows(t + kPoolWarps, nx, nq);
    } else {
      load_rows(t, ux, uq);
    }
    float xv[kChunks][8], qv[kChunks][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      if (lane + 32 * i < nchunks) {
        float2 a;
        a = unpack_bf16x2(ux[i].x); xv[i][0] = a.x; xv[i][1] = a.y;
        a = unpack_bf16x2(ux[i].y); xv[i][2] = a.x; xv[i][3] = a.y;
        a = unpack_bf16x2(ux[i].z); xv[i][4] = a.x; xv[i][5] = a.y;
        a = unpack_bf16x2(ux[i].w); xv[i][6] = a.x; xv[i][7] = a.y;
        a = unpack_bf16x2(uq[i].x); qv[i][0] = a.x; qv[i][1] = a.y;
        a = unpack_bf16x2(uq[i].y); qv[i][2] = a.x; qv[i][3] = a.y;
        a = unpack_bf16x2(uq[i].z); qv[i][4] = a.x; qv[i][5] = a.y;
        a = unpack_bf16x2(uq[i].w); qv[i][6] = a.x; qv[i][7] = a.y;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += qv[i][j];
      }
    }
    const float mean = warp_sum(s) / C;                                        // k_ln = LN(q): :160
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
        const float4* uf = reinterpret_cast<const float4*>(s_u + ch * 8);
        const float4* ub = reinterpret_cast<const float4*>(s_u + C + ch * 8);
        const float4 f0 = uf[0], f1 = uf[1], g0 = ub[0], g1 = ub[1];
        const float fu[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
        const float bu[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        float d_fg = 0.f, d_bg = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xv[i][j] = (xv[i][j] + (qv[i][j] - mean) * rstd) * kInvSqrt8;       // v = (x + k_ln) * 8^-1/2 : :178
          d_fg += fu[j] * xv[i][j];
          d_bg += bu[j] * xv[i][j];
        }
        my_part[ch] = d_fg;
        my_part[kChunks * 32 + ch] = d_bg;
      }
    }
    __syncwarp();
    // lanes 0..15: (i = lane >> 3, g = lane & 7): score = sum of the chunk dots of head g  (:247 after the fold)
    float sc = 0.f;
    if (lane < 16) {
      const float* pp = my_part + (lane >> 3) * (kChunks * 32) + (lane & 7) * chunks_per_group;
      for (int k = 0; k < chunks_per_group; ++k) sc += pp[k];
    }
    __syncwarp();
    // smooth with the head average (:251-257): mean over the 8 lanes of each i
    float tot = sc;
    tot += __shfl_xor_sync(0xffffffffu, tot, 1);
    tot += __shfl_xor_sync(0xffffffffu, tot, 2);
    tot += __shfl_xor_sync(0xffffffffu, tot, 4);
    sc = 0.5f * sc + 0.5f * (tot * 0.125f);
    if (key_mask != nullptr && key_mask[static_cast<long long>(b) * N + t] == 0) sc = kPoolMasked;   // :259-269
    if (scores != nullptr && lane < 16)
      scores[((static_cast<long long>(b) * kHeads + (lane & 7)) * 2 + (lane >> 3)) * N + t] = sc;
    // online joint softmax over (fg|bg) x tokens, per head g = lane & 7
    const float s_fg = __shfl_sync(0xffffffffu, sc, lane & 7);
    const float s_bg = __shfl_sync(0xffffffffu, sc, 8 + (lane & 7));
    const float m_new = fmaxf(m_run, fmaxf(s_fg, s_bg));
    const float corr = (m_run == -INFINITY) ? 0.f : __expf(m_run - m_new);
    const float e_fg = __expf(s_fg - m_new), e_bg = __expf(s_bg - m_new);
    l_run = l_run * corr + e_fg + e_bg;
    m_run = m_new;
    // train mode: dropout on the attention probabilities (:296) - the normaliser keeps every entry, the pooled sums use
    // the kept ones scaled by 1 / (1 - p)
    const float e_fg_k = e_fg * pool_drop_scale(seed, b, lane & 7, 0, t, N, drop_p);
    const float e_bg_k = e_bg * pool_drop_scale(seed, b, lane & 7, 1, t, N, drop_p);
    // every lane scales its channel chunks by their head's factors
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      const int ch = lane + 32 * i;
      const bool ok = ch < nchunks;
      const int g = ok ? ch / chunks_per_group : 0;
      const float cg = __shfl_sync(0xffffffffu, corr, g);
      const float fg = __shfl_sync(0xffffffffu, e_fg_k, g);
      const float bg = __shfl_sync(0xffffffffu, e_bg_k, g);
      if (ok) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc_fg[i][j] = acc_fg[i][j] * cg + fg * xv[i][j];
          acc_bg[i][j] = acc_bg[i][j] * cg + bg * xv[i][j];
        }
      }
    }
  }

  // ---- merge the 8 warps of the block (exact log-sum-exp merge), then write the block partial
  __syncthreads();
  float* rec = s_red + warp * (16 + 2 * C);
  if (lane < 8) {
    rec[lane] = m_run;
    rec[8 + lane] = l_run;
  }
#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        rec[16 + ch * 8 + j] = acc_fg[i][j];
        rec[16 + C + ch * 8 + j] = acc_bg[i][j];
      }
  }
  __syncthreads();
  float* out = partials + (static_cast<long long>(b) * gridDim.x + blk) * pool_partial_floats(C);
  __shared__ float s_scale[8][8];
  if (threadIdx.x < 8) {
    const int g = threadIdx.x;
    float m = -INFINITY;
    for (int w = 0; w < 8; ++w) m = fmaxf(m, s_red[w * (16 + 2 * C) + g]);
    float l = 0.f;
    for (int w = 0; w < 8; ++w) {
      const float mw = s_red[w * (16 + 2 * C) + g];
      const float scl = (mw == -INFINITY) ? 0.f : __expf(mw - m);
      s_scale[w][g] = scl;
      l += scl * s_red[w * (16 + 2 * C) + 8 + g];
    }
    out[g] = m;
    out[8 + g] = l;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 2 * C; idx += blockDim.x) {
    const int c = idx % C;
    const int g = c / dg;
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) a += s_scale[w][g] * s_red[w * (16 + 2 * C) + 16 + idx];
    out[16 + idx] = a;
  }
}

// ============================================================================================
// Tensor-core pooling kernel (round 2).  The SIMT kernel above spends ~250 warp instructions per token on a serial
// chain of warp reductions (ncu: issue-bound at 61 %, 1.2 TB/s); here the three contractions run on warp-level MMAs:
//   phase A  per token t and (stream i, head g):  Dx = sum_{c in g} u_i[c] x_t[c],  Dq = sum_{c in g} u_i[c] q_t[c],
//            plus sum_c q_t[c] and sum_c q_t[c]^2:  [tokens x C] x [C x 16] products with the folded queries as a
//            block-structured B operand (u split into bf16 hi + lo parts: fp32-accurate scores)
//   phase B  (one warp, lane = token)  LayerNorm statistics of q_t, raw score s8 (Dx + rstd (Dq - mean U_i,g)), head
//            smoothing, key mask, online joint softmax bookkeeping, probabilities p and p * rstd as bf16 hi + lo tables
//   phase C  acc^T[c][(i, g(c))] += sum_t x_t[c] p_t + q_t[c] (p_t rstd_t):  [C x tokens] x [tokens x 8] products, A =
//            the staged tile read transposed (ldmatrix.trans); out_i[c] = s8 (acc - sum_t p_t rstd_t mean_t)
// v = (x + LN(q)) / sqrt(8) is never formed.  Block = 4 warps, tile = kTok tokens of x and q in shared memory.
constexpr int kPmThreads = 128;

__device__ __forceinline__ void pm_ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void pm_ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void pm_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void pm_cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ uint32_t pm_bf16x2_mul(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// hi + lo bf16 split of an fp32 value (hi = rn(v), lo = rn(v - hi))
__device__ __forceinline__ void pm_split(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

template <int kC>
struct PmCfg {
  static constexpr int kDg = kC / kHeads;                       // channels per head: 40 / 80 / 160
  static constexpr int kTok = kC <= 320 ? 64 : (kC <= 640 ? 32 : 16);   // tokens per tile (~82 KB of x + q)
  static constexpr int kMT = kTok / 16;                          // token m16-tiles per tile
  static constexpr int kKS = 4 / kMT;                            // K slices of phase A (warps per token m-tile)
  static constexpr int kPitch = kC + 8;                          // tile row pitch (bf16): 16-byte aligned, conflict-free ldmatrix
  static constexpr int kKSteps = kC / 16;
  static constexpr int kStepsPerSlice = kKSteps / kKS;
  static constexpr int kCT = kC / 16;                            // channel m16-tiles of phase C
  static constexpr int kCTPerWarp = kCT / 4;
  static constexpr int kPP = kTok + 8;                           // pitch of the transposed probability tables
  static constexpr int kPartCols = 36;                           // Dx[16] Dq[16] sum sq (+2 pad)
  static constexpr size_t kSmem = 2ull * kTok * kPitch * 2 + 4ull * kC * 2 + 4ull * 16 * kPP * 2 +
                                  static_cast<size_t>(kKS) * kTok * kPartCols * 4 + 224 * 4;
  static_assert(kKSteps % kKS == 0 && kCT % 4 == 0 && kDg % 8 == 0, "geometry");
};

template <int kC>
__global__ void __launch_bounds__(kPmThreads, 2) ada_pool_mma_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ q, const float* __restrict__ u,
    const uint8_t* __restrict__ key_mask, float* __restrict__ scores, float* __restrict__ partials, int N,
    int tokens_per_block, float drop_p, uint32_t seed) {
  using P = PmCfg<kC>;
  extern __shared__ __align__(16) uint8_t pm_smem[];
  __nv_bfloat16* sx = reinterpret_cast<__nv_bfloat16*>(pm_smem);          // [kTok][kPitch]
  __nv_bfloat16* sq = sx + P::kTok * P::kPitch;                           // [kTok][kPitch]
  __nv_bfloat16* su = sq + P::kTok * P::kPitch;                           // [hi fg | hi bg | lo fg | lo bg][kC]
  __nv_bfloat16* sp = su + 4 * kC;                                        // [p_hi | p_lo | pr_hi | pr_lo][16][kPP]
  float* s_part = reinterpret_cast<float*>(sp + 4 * 16 * P::kPP);         // [kKS][kTok][kPartCols]
  float* s_misc = s_part + P::kKS * P::kTok * P::kPartCols;               // [0..8) corr, [8] any rescale, [16..32) U_i,g, [32..48) sum p rstd mean, [48..80) tile max per warp, [80..88) l
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.y, blk = blockIdx.x;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gl = lane >> 2, t4 = lane & 3;
  // folded queries of this sample as bf16 hi + lo; U_i,g = sum of u_i over head g (fp32)
  for (int idx = tid; idx < 2 * kC; idx += kPmThreads) {
    __nv_bfloat16 hi, lo;
    pm_split(u[static_cast<long long>(b) * 2 * kC + idx], hi, lo);
    su[idx] = hi;
    su[2 * kC + idx] = lo;
  }
  {
    // eight lanes per (stream, head): strided partial sums, then three shuffles
    const int ig = tid >> 3, sub = tid & 7;
    const float* up = u + static_cast<long long>(b) * 2 * kC + (ig >> 3) * kC + (ig & 7) * P::kDg;
    float a = 0.f;
#pragma unroll 5
    for (int c = sub; c < P::kDg; c += 8) a += up[c];
    a += __shfl_xor_sync(0xffffffffu, a, 1);
    a += __shfl_xor_sync(0xffffffffu, a, 2);
    a += __shfl_xor_sync(0xffffffffu, a, 4);
    if (sub == 0) s_misc[16 + ig] = a;
  }
  const int t_begin = blk * tokens_per_block;
  const int t_end = min(N, t_begin + tokens_per_block);
  // phase A role: token m-tile and K slice of this warp; phase C role: channel m-tiles warp, warp + 4, ...
  const int a_mt = warp % P::kMT, a_ks = warp / P::kMT;
  const int lm_row = ((lane >> 3) & 1) * 8 + (lane & 7), lm_k = (lane >> 4) * 8;            // ldmatrix (A = tokens x channels)
  const int lt_tok = (lane & 7) + (lane >> 4) * 8, lt_ch = ((lane >> 3) & 1) * 8;          // ldmatrix.trans (A = channels x tokens)
  float acc[P::kCTPerWarp][4];
#pragma unroll
  for (int i = 0; i < P::kCTPerWarp; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
  // softmax state of this thread's (stream, heads) - see phase B: the running max (equal in every thread that owns the
  // head), its partial sum of exponentials and of p * rstd * mean over the tokens it has handled
  float m_run[kHeads], l_part[kHeads], c_part[kHeads];
#pragma unroll
  for (int g = 0; g < kHeads; ++g) {
    m_run[g] = -INFINITY;
    l_part[g] = c_part[g] = 0.f;
  }
  const uint32_t sx_u = smem_u32(sx), sq_u = smem_u32(sq);

  for (int t0 = t_begin; t0 < t_end; t0 += P::kTok) {
    __syncthreads();                                      // the previous tile (and the u tables, first time) is done with
    // ---- stage the tile: rows past t_end are zero-filled
    for (int i = tid; i < P::kTok * (kC / 8); i += kPmThreads) {
      const int r = i / (kC / 8), v = i - r * (kC / 8);
      const bool ok = t0 + r < t_end;
      const long long off = (static_cast<long long>(b) * N + (ok ? t0 + r : t_begin)) * kC + v * 8;
      pm_cp_async16(sx_u + (r * P::kPitch + v * 8) * 2, x + off, ok);
      pm_cp_async16(sq_u + (r * P::kPitch + v * 8) * 2, q + off, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    // ---- phase A: this warp's 16 tokens x its K slice
    {
      float dx[2][4], dq[2][4], dsum[4], dsq[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) dx[0][e] = dx[1][e] = dq[0][e] = dq[1][e] = dsum[e] = dsq[e] = 0.f;
      const uint32_t ax = sx_u + ((a_mt * 16 + lm_row) * P::kPitch + lm_k) * 2;
      const uint32_t aq = sq_u + ((a_mt * 16 + lm_row) * P::kPitch + lm_k) * 2;
      const uint32_t ones = 0x3F803F80u;                  // bf16 (1, 1)
#pragma unroll 2
      for (int s = 0; s < P::kStepsPerSlice; ++s) {
        const int ks = a_ks * P::kStepsPerSlice + s;
        uint32_t fx[4], fq[4], fq2[4];
        pm_ldmatrix_x4(fx, ax + ks * 32);
        pm_ldmatrix_x4(fq, aq + ks * 32);
#pragma unroll
        for (int e = 0; e < 4; ++e) fq2[e] = pm_bf16x2_mul(fq[e], fq[e]);
        pm_mma(dsum, fq, ones, ones);
        pm_mma(dsq, fq2, ones, ones);
        // B fragments: column n = gl is head gl; rows k = 2 t4, 2 t4 + 1 (b0) and + 8 (b1) are channels of this step
        const int c0 = ks * 16 + 2 * t4, c1 = c0 + 8;
        const bool in0 = c0 / P::kDg == gl, in1 = c1 / P::kDg == gl;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint32_t h0 = in0 ? *reinterpret_cast<const uint32_t*>(su + i * kC + c0) : 0u;
          const uint32_t h1 = in1 ? *reinterpret_cast<const uint32_t*>(su + i * kC + c1) : 0u;
          const uint32_t l0 = in0 ? *reinterpret_cast<const uint32_t*>(su + (2 + i) * kC + c0) : 0u;
          const uint32_t l1 = in1 ? *reinterpret_cast<const uint32_t*>(su + (2 + i) * kC + c1) : 0u;
          pm_mma(dx[i], fx, h0, h1);
          pm_mma(dx[i], fx, l0, l1);
          pm_mma(dq[i], fq, h0, h1);
          pm_mma(dq[i], fq, l0, l1);
        }
      }
      // D fragment: rows gl, gl + 8 (tokens of this m-tile), columns 2 t4, 2 t4 + 1 (heads)
      float* pa = s_part + (a_ks * P::kTok + a_mt * 16 + gl) * P::kPartCols;
      float* pb = pa + 8 * P::kPartCols;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        pa[i * 8 + 2 * t4] = dx[i][0]; pa[i * 8 + 2 * t4 + 1] = dx[i][1];
        pb[i * 8 + 2 * t4] = dx[i][2]; pb[i * 8 + 2 * t4 + 1] = dx[i][3];
        pa[16 + i * 8 + 2 * t4] = dq[i][0]; pa[16 + i * 8 + 2 * t4 + 1] = dq[i][1];
        pb[16 + i * 8 + 2 * t4] = dq[i][2]; pb[16 + i * 8 + 2 * t4 + 1] = dq[i][3];
      }
      if (t4 == 0) {
        pa[32] = dsum[0]; pa[33] = dsq[0];
        pb[32] = dsum[2]; pb[33] = dsq[2];
      }
    }
    __syncthreads();
    // ---- phase B: all 128 threads; the 128 / kTok adjacent lanes of a token split its 2 x 8 (stream, head) scores
    {
      constexpr int kLanesPerTok = kPmThreads / P::kTok;          // 2, 4, 8
      constexpr int kHeadsPerThr = 16 / kLanesPerTok;             // 8, 4, 2 heads of ONE stream
      const int r = tid / kLanesPerTok, part = tid % kLanesPerTok;
      const int i = part / (kLanesPerTok / 2);                    // stream: 0 = fg, 1 = bg
      const int g_lo = (part % (kLanesPerTok / 2)) * kHeadsPerThr;
      const bool live = t0 + r < t_end;
      float s_sum = 0.f, s_sq = 0.f;
#pragma unroll
      for (int s = 0; s < P::kKS; ++s) {
        const float2 f = *reinterpret_cast<const float2*>(s_part + (s * P::kTok + r) * P::kPartCols + 32);
        s_sum += f.x;
        s_sq += f.y;
      }
      const float mean = s_sum * (1.0f / kC);
      const float rstd = rsqrtf(fmaxf(s_sq * (1.0f / kC) - mean * mean, 0.f) + 1e-5f);          // k_ln = LN(q): :160
      float sc[kHeadsPerThr];
      float tot = 0.f;
#pragma unroll
      for (int j = 0; j < kHeadsPerThr; ++j) {
        const int g = g_lo + j;
        float vx = 0.f, vq = 0.f;
#pragma unroll
        for (int s = 0; s < P::kKS; ++s) {
          const float* pp = s_part + (s * P::kTok + r) * P::kPartCols;
          vx += pp[i * 8 + g];
          vq += pp[16 + i * 8 + g];
        }
        // s = sum_c u_i[c] v[c], v = (x + (q - mean) rstd) / sqrt(8)   (:178, :247 after the fold)
        sc[j] = kInvSqrt8 * (vx + rstd * (vq - mean * s_misc[16 + i * 8 + g]));
        tot += sc[j];
      }
      // the other head groups of this (token, stream) sit in the adjacent lanes
#pragma unroll
      for (int o = 1; o < kLanesPerTok / 2; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
      const bool masked = live && key_mask != nullptr && key_mask[static_cast<long long>(b) * N + t0 + r] == 0;
#pragma unroll
      for (int j = 0; j < kHeadsPerThr; ++j) {
        float s = 0.5f * sc[j] + 0.5f * (tot * 0.125f);                                          // :251-257
        if (masked) s = kPoolMasked;                                                            // :259-269
        if (!live) s = -INFINITY;
        sc[j] = s;
        if (scores != nullptr && live)
          scores[((static_cast<long long>(b) * kHeads + g_lo + j) * 2 + i) * N + t0 + r] = s;
        // tile max per head over tokens and both streams: lanes with equal head group are kLanesPerTok / 2 apart... reduce
        // over everything except the head-group bits, then across the four warps through shared memory
        float mx = s;
#pragma unroll
        for (int o = kLanesPerTok / 2; o < 32; o <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane < kLanesPerTok / 2) s_misc[48 + warp * 8 + g_lo + j] = mx;
      }
      __syncthreads();
      __nv_bfloat16* t_hi = sp;
      __nv_bfloat16* t_lo = sp + 16 * P::kPP;
      __nv_bfloat16* q_hi = sp + 2 * 16 * P::kPP;
      __nv_bfloat16* q_lo = sp + 3 * 16 * P::kPP;
      bool any = false;
#pragma unroll
      for (int j = 0; j < kHeadsPerThr; ++j) {
        const int g = g_lo + j;
        const float tm = fmaxf(fmaxf(s_misc[48 + g], s_misc[56 + g]), fmaxf(s_misc[64 + g], s_misc[72 + g]));
        const float m_new = fmaxf(m_run[j], tm);
        const float corr = (m_run[j] == -INFINITY) ? 0.f : __expf(m_run[j] - m_new);
        m_run[j] = m_new;
        l_part[j] *= corr;
        c_part[j] *= corr;
        any = any || corr != 1.0f;
        if (r == 0 && i == 0) s_misc[g] = corr;           // (one writer per head)
        const float e = live ? __expf(sc[j] - m_new) : 0.f;
        l_part[j] += e;
        // train mode: dropout on the probabilities (:296): the normaliser keeps every entry
        const float ek = e * (live ? pool_drop_scale(seed, b, g, i, t0 + r, N, drop_p) : 0.f);
        const float er = ek * rstd;
        c_part[j] = fmaf(er, mean, c_part[j]);
        __nv_bfloat16 hi, lo;
        pm_split(ek, hi, lo);
        t_hi[(i * 8 + g) * P::kPP + r] = hi;
        t_lo[(i * 8 + g) * P::kPP + r] = lo;
        pm_split(er, hi, lo);
        q_hi[(i * 8 + g) * P::kPP + r] = hi;
        q_lo[(i * 8 + g) * P::kPP + r] = lo;
      }
      if (__syncthreads_or(any ? 1 : 0) != 0) {
        if (tid == 0) s_misc[8] = 1.f;
      } else if (tid == 0) {
        s_misc[8] = 0.f;
      }
    }
    __syncthreads();
    // ---- phase C: acc^T[channel][(i, head of the channel)] += x^T p + q^T (p rstd) over this tile's tokens
    {
      const bool rescale = s_misc[8] != 0.f;
      const __nv_bfloat16* tabs = sp;
#pragma unroll
      for (int ci = 0; ci < P::kCTPerWarp; ++ci) {
        const int c0 = (warp + 4 * ci) * 16;              // first channel of this m16 tile
        const int h0 = c0 / P::kDg, h1 = (c0 + 15) / P::kDg;
        if (rescale) {
          const float fa = s_misc[(c0 + gl) / P::kDg], fb = s_misc[(c0 + gl + 8) / P::kDg];
          acc[ci][0] *= fa; acc[ci][1] *= fa;
          acc[ci][2] *= fb; acc[ci][3] *= fb;
        }
        // B columns: 0 = (fg, h0), 1 = (bg, h0), 2 = (fg, h1), 3 = (bg, h1), 4..7 = 0; column n = gl of this lane
        const int col = (gl & 1) * 8 + ((gl & 2) ? h1 : h0);
        const __nv_bfloat16* bp = tabs + col * P::kPP + 2 * t4;
        const uint32_t ax = sx_u + (lt_tok * P::kPitch + c0 + lt_ch) * 2;
        const uint32_t aq = sq_u + (lt_tok * P::kPitch + c0 + lt_ch) * 2;
#pragma unroll
        for (int kk = 0; kk < P::kTok / 16; ++kk) {
          uint32_t fx[4], fq[4];
          pm_ldmatrix_x4_trans(fx, ax + kk * 16 * P::kPitch * 2);
          pm_ldmatrix_x4_trans(fq, aq + kk * 16 * P::kPitch * 2);
          uint32_t bh0 = 0u, bh1 = 0u, bl0 = 0u, bl1 = 0u, rh0 = 0u, rh1 = 0u, rl0 = 0u, rl1 = 0u;
          if (gl < 4) {
            const __nv_bfloat16* b = bp + kk * 16;
            bh0 = *reinterpret_cast<const uint32_t*>(b);
            bh1 = *reinterpret_cast<const uint32_t*>(b + 8);
            bl0 = *reinterpret_cast<const uint32_t*>(b + 16 * P::kPP);
            bl1 = *reinterpret_cast<const uint32_t*>(b + 16 * P::kPP + 8);
            rh0 = *reinterpret_cast<const uint32_t*>(b + 2 * 16 * P::kPP);
            rh1 = *reinterpret_cast<const uint32_t*>(b + 2 * 16 * P::kPP + 8);
            rl0 = *reinterpret_cast<const uint32_t*>(b + 3 * 16 * P::kPP);
            rl1 = *reinterpret_cast<const uint32_t*>(b + 3 * 16 * P::kPP + 8);
          }
          pm_mma(acc[ci], fx, bh0, bh1);
          pm_mma(acc[ci], fx, bl0, bl1);
          pm_mma(acc[ci], fq, rh0, rh1);
          pm_mma(acc[ci], fq, rl0, rl1);
        }
      }
    }
  }

  // ---- block partial: (max, sum) per head, then the two pooled streams  out_i[c] = s8 (acc - sum_t p rstd mean)
  float* out = partials + (static_cast<long long>(b) * gridDim.x + blk) * pool_partial_floats(kC);
  __syncthreads();
  {
    // per (stream, head) sums over the tokens: shuffles over the token bits of the lane index, one slot per warp, then
    // four adds (shared-memory float atomics with 64 contenders per address cost more than the whole token loop)
    constexpr int kLanesPerTok = kPmThreads / P::kTok;
    constexpr int kHeadsPerThr = 16 / kLanesPerTok;
    const int r = tid / kLanesPerTok, part = tid % kLanesPerTok;
    const int i = part / (kLanesPerTok / 2), g_lo = (part % (kLanesPerTok / 2)) * kHeadsPerThr;
    float* s_red = s_misc + 96;                           // [4 warps][l | c][16 (stream, head)]
#pragma unroll
    for (int j = 0; j < kHeadsPerThr; ++j) {
      float ls = l_part[j], cs = c_part[j];
#pragma unroll
      for (int o = kLanesPerTok; o < 32; o <<= 1) {
        ls += __shfl_xor_sync(0xffffffffu, ls, o);
        cs += __shfl_xor_sync(0xffffffffu, cs, o);
      }
      if (lane < kLanesPerTok) {
        s_red[(warp * 2 + 0) * 16 + i * 8 + g_lo + j] = ls;
        s_red[(warp * 2 + 1) * 16 + i * 8 + g_lo + j] = cs;
      }
      if (r == 0 && i == 0) out[g_lo + j] = m_run[j];
    }
    __syncthreads();
    if (tid < 16) s_misc[32 + tid] = s_red[16 + tid] + s_red[48 + tid] + s_red[80 + tid] + s_red[112 + tid];
    if (tid < 8)
      out[8 + tid] = (s_red[tid] + s_red[8 + tid]) + (s_red[32 + tid] + s_red[40 + tid]) + (s_red[64 + tid] + s_red[72 + tid]) +
                     (s_red[96 + tid] + s_red[104 + tid]);
  }
  __syncthreads();
#pragma unroll
  for (int ci = 0; ci < P::kCTPerWarp; ++ci) {
    const int c0 = (warp + 4 * ci) * 16;
    const int h0 = c0 / P::kDg, h1 = (c0 + 15) / P::kDg;
    // D fragment: rows gl, gl + 8 (channels), columns 2 t4, 2 t4 + 1: t4 = 0 holds (fg, bg) of head h0, t4 = 1 of head h1
    if (t4 < 2) {
      const int hh = t4 == 0 ? h0 : h1;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int c = c0 + gl + 8 * half;
        if (c / P::kDg == hh && (t4 == 0 || h1 != h0)) {
          out[16 + c] = kInvSqrt8 * (acc[ci][2 * half] - s_misc[32 + hh]);
          out[16 + kC + c] = kInvSqrt8 * (acc[ci][2 * half + 1] - s_misc[40 + hh]);
        }
      }
    }
  }
}

// grid (B, ceil(2C / 256)): every block recomputes the 8 heads' merge factors (nblocks exps each), then one thread
// per pooled channel sums its nblocks partials
__global__ void __launch_bounds__(256) ada_pool_combine_kernel(const float* __restrict__ partials, int nblocks, int C,
                                                               float* __restrict__ fg_out, float* __restrict__ bg_out,
                                                               float* __restrict__ stats) {
  extern __shared__ float sh[];     // [nblocks][8] scales
  __shared__ float s_l[8];
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.x;
  const int dg = C / kHeads;
  const long long rec = pool_partial_floats(C);
  const float* base = partials + static_cast<long long>(b) * nblocks * rec;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  {
    const int g = warp;               // 8 warps = 8 heads
    float m = -INFINITY;
    for (int k = lane; k < nblocks; k += 32) m = fmaxf(m, base[k * rec + g]);
    m = warp_max(m);
    float l = 0.f;
    for (int k = lane; k < nblocks; k += 32) {
      const float mk = base[k * rec + g];
      const float scl = (mk == -INFINITY) ? 0.f : __expf(mk - m);
      sh[k * 8 + g] = scl;
      l += scl * base[k * rec + 8 + g];
    }
    l = warp_sum(l);
    if (lane == 0) {
      s_l[g] = l;
      if (blockIdx.y == 0) {
        stats[(static_cast<long long>(b) * 8 + g) * 2] = m;
        stats[(static_cast<long long>(b) * 8 + g) * 2 + 1] = l;
      }
    }
  }
  __syncthreads();
  const int idx = blockIdx.y * blockDim.x + threadIdx.x;
  if (idx < 2 * C) {
    const int c = idx % C, g = c / dg;
    float a = 0.f;
    for (int k = 0; k < nblocks; ++k) a += sh[k * 8 + g] * base[k * rec + 16 + idx];
    a /= s_l[g];
    if (idx < C) fg_out[static_cast<long long>(b) * C + c] = a;
    else bg_out[static_cast<long long>(b) * C + c] = a;
  }
}

__global__ void __launch_bounds__(256) ada_pool_probs_kernel(const float* __restrict__ scores, const float* __restrict__ stats,
                                                             float* __restrict__ attn_fg, float* __restrict__ attn_bg, int N,
                                                             long long total, float drop_p, uint32_t seed) {
  // scores: [B][8][2][N]; attn_*: [B][8][N]
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long bg_idx = i / N;            // (b*8 + g)*2 + which
    const int n = static_cast<int>(i % N);
    const long long bgi = bg_idx >> 1;
    const int which = static_cast<int>(bg_idx & 1);
    const float m = stats[bgi * 2], l = stats[bgi * 2 + 1];
    const float pr = __expf(scores[i] - m) / l *
                     pool_drop_scale(seed, static_cast<int>(bgi >> 3), static_cast<int>(bgi & 7), which, n, N, drop_p);
    (which == 0 ? attn_fg : attn_bg)[bgi * N + n] = pr;
  }
}

// ---- K13 ------------------------------------------------------------------------------------------------
// Block (k, b): LNCat of the pooled features and the time slice (recomputed per block, 2C + TD values), the r
// coefficients of vector k (r dot products of length 2C + TD, one warp each, float4 loads), the basis combination,
// the per-vector LayerNorm, 1/sqrt(D) and the bias.  When ctx_out is given the vector is also written as bf16 into
// row token_pos[k] of sample b's context (the splice of the subject rows).
__global__ void __launch_bounds__(256) ada_head_kernel(const float* __restrict__ infeat_fg, const float* __restrict__ infeat_bg,
                                                       const float* __restrict__ time_emb, int time_ld,
                                                       const float* __restrict__ coeff_w, const float* __restrict__ coeff_b,
                                                       const float* __restrict__ basis, const float* __restrict__ bias,
                                                       float* __restrict__ out, __nv_bfloat16* __restrict__ ctx_out,
                                                       const int* __restrict__ token_pos, long long ctx_sample_stride,
                                                       long long basis_sample_stride, int C, int TD, int K, int R, int D,
                                                       const float* __restrict__ partials, int nblocks,
                                                       float* __restrict__ fg_out, float* __restrict__ bg_out,
                                                       float* __restrict__ stats_out) {
  extern __shared__ float sh[];    // feat [2C + TD], coeffs [R], vec [D], (fused combine) merge factors [nblocks][8]
  __shared__ float red[32];
  __shared__ float s_lsum[kHeads];
  const int k = blockIdx.x, b = blockIdx.y;
  const int F = 2 * C + TD;
  {
    // The coefficient rows and basis vectors of this k are parameters, not outputs of the pooling kernels: pull them
    // into L2 while the predecessor is still running (programmatic dependent launch), so that the dependent part of
    // this latency-bound kernel finds them there.
    auto prefetch = [&](const float* ptr, long long nfloats) {
      const char* pc = reinterpret_cast<const char*>(ptr);
      for (long long off = static_cast<long long>(threadIdx.x) * 128; off < nfloats * 4; off += static_cast<long long>(blockDim.x) * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pc + off));
    };
    prefetch(coeff_w + static_cast<long long>(k) * R * F, static_cast<long long>(R) * F);
    prefetch(basis + static_cast<long long>(b) * basis_sample_stride + static_cast<long long>(k) * R * D, static_cast<long long>(R) * D);
    if (bias != nullptr) prefetch(bias + static_cast<long long>(k) * D, D);
  }
  pdl_wait();
  pdl_trigger();
  float* feat = sh;
  float* coeffs = feat + F;
  float* vec = coeffs + ((R + 3) & ~3);
  if (partials != nullptr) {
    // the pooling kernel's chunk partials are merged HERE (the separate combine launch, 6.6 us on the critical path of
    // every cross-attention layer, disappears): every block recomputes the 8 heads' merge factors, then each thread sums
    // its pooled channels over the chunks; the k = 0 blocks also publish the pooled features and the softmax statistics
    float* s_scl = vec + D;
    const int dg = C / kHeads;
    const long long rec = pool_partial_floats(C);
    const float* base = partials + static_cast<long long>(b) * nblocks * rec;
    const int warp_ = threadIdx.x >> 5, lane_ = threadIdx.x & 31;
    {
      const int g = warp_;              // 8 warps = 8 heads
      float m = -INFINITY;
      for (int j = lane_; j < nblocks; j += 32) m = fmaxf(m, base[j * rec + g]);
      m = warp_max(m);
      float l = 0.f;
      for (int j = lane_; j < nblocks; j += 32) {
        const float mj = base[j * rec + g];
        const float scl = (mj == -INFINITY) ? 0.f : __expf(mj - m);
        s_scl[j * 8 + g] = scl;
        l += scl * base[j * rec + 8 + g];
      }
      l = warp_sum(l);
      if (lane_ == 0) {
        s_lsum[g] = l;
        if (k == 0 && stats_out != nullptr) {
          stats_out[(static_cast<long long>(b) * 8 + g) * 2] = m;
          stats_out[(static_cast<long long>(b) * 8 + g) * 2 + 1] = l;
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
      const int c = i % C, g = c / dg;
      float a = 0.f;
      for (int j = 0; j < nblocks; ++j) a += s_scl[j * 8 + g] * base[j * rec + 16 + i];
      a /= s_lsum[g];
      if (k == 0) {
        if (i < C) { if (fg_out != nullptr) fg_out[static_cast<long long>(b) * C + c] = a; }
        else if (bg_out != nullptr) bg_out[static_cast<long long>(b) * C + c] = a;
      }
      if (i >= C && infeat_bg != nullptr) a = infeat_bg[static_cast<long long>(b) * C + c];   // cached bg pooling (:789-792)
      feat[i] = a;
    }
    for (int i = 2 * C + threadIdx.x; i < F; i += blockDim.x) feat[i] = time_emb[static_cast<long long>(b) * time_ld + i - 2 * C];
  } else {
    for (int i = threadIdx.x; i < F; i += blockDim.x) {
      float v;
      if (i < C) v = infeat_fg[static_cast<long long>(b) * C + i];
      else if (i < 2 * C) v = infeat_bg[static_cast<long long>(b) * C + i - C];
      else v = time_emb[static_cast<long long>(b) * time_ld + i - 2 * C];        // time_emb[:, :TD]  :849
      feat[i] = v;
    }
  }
  __syncthreads();
  // LNCat3 (:38-58): LayerNorm (no affine) on the pooled features and on the time slice separately
  for (int part = 0; part < 2; ++part) {
    const int lo = part == 0 ? 0 : 2 * C, n = part == 0 ? 2 * C : TD;
    float s = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += feat[lo + i];
    const float mean = block_sum_256(s, red) / n;
    float m2 = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float d = feat[lo + i] - mean;
      m2 += d * d;
    }
    const float rstd = rsqrtf(block_sum_256(m2, red) / n + 1e-5f);
    for (int i = threadIdx.x; i < n; i += blockDim.x) feat[lo + i] = (feat[lo + i] - mean) * rstd;
    __syncthreads();
  }
  // basis_dyn_coeffs[k] = Linear(feat)[k*R : (k+1)*R] (:858), one warp per output
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int F4 = F >> 2;
  for (int o = warp; o < R; o += (blockDim.x >> 5)) {
    const float4* w = reinterpret_cast<const float4*>(coeff_w + static_cast<long long>(k * R + o) * F);
    float a = 0.f;
#pragma unroll 4
    for (int i = lane; i < F4; i += 32) {
      const float4 ww = __ldg(w + i);
      const float4 ff = *reinterpret_cast<const float4*>(feat + 4 * i);
      a += ww.x * ff.x + ww.y * ff.y + ww.z * ff.z + ww.w * ff.w;
    }
    a = warp_sum(a);
    if (lane == 0) coeffs[o] = a + coeff_b[k * R + o];
  }
  __syncthreads();
  // out[k] = LN(coeffs[k] @ basis[k]) / sqrt(D) + bias[k]  (:881-893)
  const float inv_sqrt_d = rsqrtf(static_cast<float>(D));
  float s = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    // basis_sample_stride = 0: shared basis [K, R, D]; > 0: per-instance zero-shot basis [B, K, R, D] (:860-868)
    const float* bk = basis + static_cast<long long>(b) * basis_sample_stride + static_cast<long long>(k) * R * D;
    for (int r = 0; r < R; ++r) a += coeffs[r] * __ldg(bk + static_cast<long long>(r) * D + d);
    vec[d] = a;
    s += a;
  }
  const float mean = block_sum_256(s, red) / D;
  float m2 = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float dd = vec[d] - mean;
    m2 += dd * dd;
  }
  const float rstd = rsqrtf(block_sum_256(m2, red) / D + 1e-5f);
  __nv_bfloat16* crow = ctx_out == nullptr ? nullptr
                                           : ctx_out + static_cast<long long>(b) * ctx_sample_stride +
                                                 static_cast<long long>(token_pos[k]) * D;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float v = (vec[d] - mean) * rstd * inv_sqrt_d;
    if (bias != nullptr) v += bias[static_cast<long long>(k) * D + d];
    out[(static_cast<long long>(b) * K + k) * D + d] = v;
    if (crow != nullptr) crow[d] = __float2bfloat16_rn(v);
  }
}

// ---- N2: StaticLayerwiseEmbedding + splice ------------------------------------------------------------------
// Block (k, l): coefficients basis_rand_weights[l, k] + basis_comm_weights[0, k] (embedding_manager.py:501), the
// combination of vector k's basis (pre_vecs | basis_vecs, :516-522), LayerNorm without affine (:529), 1/sqrt(D)
// (:530) and the layer/vector bias (:534).
__global__ void __launch_bounds__(256) static_layerwise_kernel(const float* __restrict__ rand_w, const float* __restrict__ comm_w,
                                                               const float* __restrict__ pre_vecs, int n_pre,
                                                               const float* __restrict__ basis_vecs,
                                                               const float* __restrict__ bias, float* __restrict__ out, int K,
                                                               int R, int D) {
  extern __shared__ float sh[];    // coeffs [R rounded up to 4], vec [D]
  __shared__ float red[32];
  pdl_wait();
  pdl_trigger();
  const int k = blockIdx.x, l = blockIdx.y;
  float* coeffs = sh;
  float* vec = sh + ((R + 3) & ~3);
  for (int r = threadIdx.x; r < R; r += blockDim.x)
    coeffs[r] = rand_w[(static_cast<long long>(l) * K + k) * R + r] + comm_w[static_cast<long long>(k) * R + r];
  __syncthreads();
  const float* pre = pre_vecs == nullptr ? nullptr : pre_vecs + static_cast<long long>(k) * n_pre * D;
  const float* bas = basis_vecs + static_cast<long long>(k) * (R - n_pre) * D;
  float s = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    for (int r = 0; r < n_pre; ++r) a += coeffs[r] * __ldg(pre + static_cast<long long>(r) * D + d);
    for (int r = n_pre; r < R; ++r) a += coeffs[r] * __ldg(bas + static_cast<long long>(r - n_pre) * D + d);
    vec[d] = a;
    s += a;
  }
  const float mean = block_sum_256(s, red) / D;
  float m2 = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float dd = vec[d] - mean;
    m2 += dd * dd;
  }
  const float rstd = rsqrtf(block_sum_256(m2, red) / D + 1e-5f);
  const float inv_sqrt_d = rsqrtf(static_cast<float>(D));
  const long long row = (static_cast<long long>(l) * K + k) * D;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float v = (vec[d] - mean) * rstd * inv_sqrt_d;
    if (bias != nullptr) v += bias[row + d];
    out[row + d] = v;
  }
}

// one warp per prompt: position of the first occurrence of `token` (-1: none)  (ldm/util.py:2114-2124)
__global__ void __launch_bounds__(32) first_token_pos_kernel(const long long* __restrict__ tokens, long long token,
                                                             int* __restrict__ first_pos, int N) {
  pdl_wait();
  pdl_trigger();
  const long long* row = tokens + static_cast<long long>(blockIdx.x) * N;
  int best = 0x7fffffff;
  for (int n = threadIdx.x; n < N; n += 32)
    if (row[n] == token) best = min(best, n);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (threadIdx.x == 0) first_pos[blockIdx.x] = best == 0x7fffffff ? -1 : best;
}

// Block (n, l, b) writes row n of prompt copy b * L + l (instance-major / layer-minor, :1341-1352): the token
// embedding, or vector n - first of layer l's static embedding for the K positions that start at the placeholder
// (:1510-1552).  One float4 per thread and pass.
__global__ void __launch_bounds__(256) splice_static_kernel(const float* __restrict__ embedded_text, const int* __restrict__ first_pos,
                                                            const float* __restrict__ static_emb, float* __restrict__ out_f32,
                                                            __nv_bfloat16* __restrict__ out_bf16, int N, int L, int K, int D) {
  pdl_wait();
  pdl_trigger();
  const int n = blockIdx.x, l = blockIdx.y, b = blockIdx.z;
  const int first = first_pos[b];
  const bool subj = first >= 0 && n >= first && n < first + K;
  const float* src = subj ? static_emb + (static_cast<long long>(l) * K + (n - first)) * D
                          : embedded_text + (static_cast<long long>(b) * N + n) * D;
  const long long dst = ((static_cast<long long>(b) * L + l) * N + n) * D;
  for (int i = threadIdx.x * 4; i < D; i += blockDim.x * 4) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    if (out_f32 != nullptr) *reinterpret_cast<float4*>(out_f32 + dst + i) = v;
    if (out_bf16 != nullptr) {
      uint2 pk;
      pk.x = pack_bf16x2(v.x, v.y);
      pk.y = pack_bf16x2(v.z, v.w);
      *reinterpret_cast<uint2*>(out_bf16 + dst + i) = pk;
    }
  }
}

int pool_blocks_per_sample(int batch, int n_tokens, int channels) {
  // ONE wave of blocks over the GPU (the main kernel fits two blocks per SM up to 512 channels, one above: its
  // registers hold the running sums of all channels), at least 8 tokens (one per warp) per block.  A second wave
  // would only add block prologues and warp merges: the token loop is latency-bound, not occupancy-bound.
  const int per_sm = channels <= 512 ? 2 : 1;
  int want = per_sm * num_sms() / batch;
  int max_blocks = (n_tokens + 7) / 8;
  if (want > max_blocks) want = max_blocks;
  if (want < 1) want = 1;
  return want;
}

}  // namespace
}  // namespace afb

extern "C" int64_t afb_ada_pool_ws_floats(int batch, int n_tokens, int channels) {
  using namespace afb;
  const int64_t nblocks = (n_tokens + 7) / 8 < 2 * 160 ? (n_tokens + 7) / 8 : 2 * 160;   // upper bound on blocks/sample
  return static_cast<int64_t>(batch) * kHeads * 2 * n_tokens        // scores
         + static_cast<int64_t>(batch) * nblocks * pool_partial_floats(channels)
         + static_cast<int64_t>(batch) * kHeads * 2;                // merged (max, sum)
}

extern "C" int64_t afb_ada_pool_stats_offset(int batch, int n_tokens, int channels) {
  // where afb_ada_pool leaves the merged softmax statistics [B, 8, 2] inside its workspace (same geometry as the launch)
  using namespace afb;
  const int nblocks = pool_blocks_per_sample(batch, n_tokens, channels);
  const int tokens_per_block = (n_tokens + nblocks - 1) / nblocks;
  const int grid_x = (n_tokens + tokens_per_block - 1) / tokens_per_block;
  return static_cast<int64_t>(batch) * kHeads * 2 * n_tokens + static_cast<int64_t>(batch) * grid_x * pool_partial_floats(channels);
}

extern "C" int64_t afb_ada_pool_prep_ws_floats(int batch, int channels) {
  return (static_cast<int64_t>(batch) + 1) * channels;
}

extern "C" int afb_ada_pool_prep(const float* to_k_weight, int ctx_dim, const float* fg_q_emb, const float* bg_q_emb,
                                 const float* lora_to_fg_q, const float* lora_to_bg_q, const float* lora_to_k, float* pool_q,
                                 float* ws, int batch, int channels, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(to_k_weight && fg_q_emb && bg_q_emb && lora_to_fg_q && lora_to_bg_q && lora_to_k && pool_q && ws,
              "afb_ada_pool_prep: NULL argument");
  AFB_REQUIRE(batch > 0 && ctx_dim > 0 && ctx_dim % 4 == 0, "afb_ada_pool_prep: ctx_dim must be a positive multiple of 4");
  AFB_REQUIRE(channels % 64 == 0 && channels <= 1280, "afb_ada_pool_prep: channels %d must be a multiple of 64, <= 1280", channels);
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(to_k_weight) & 15) == 0, "afb_ada_pool_prep: to_k weight must be 16-byte aligned");
  const int n_src = batch + 1;
  const int src_in_smem = n_src < kPrepMaxSrc ? n_src : kPrepMaxSrc;
  const size_t sh1 = static_cast<size_t>(src_in_smem) * ctx_dim * sizeof(float);
  AFB_REQUIRE(sh1 <= 96 * 1024, "afb_ada_pool_prep: ctx_dim %d too large", ctx_dim);
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(ada_qproj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set = true;
  }
  ada_qproj_kernel<<<(channels + kPoolWarps - 1) / kPoolWarps, kPoolThreads, sh1, stream>>>(to_k_weight, fg_q_emb, bg_q_emb, ws,
                                                                                           channels, ctx_dim, n_src);
  count_launch();
  const size_t sh2 = (static_cast<size_t>(channels) + channels / kHeads) * sizeof(float);
  ada_qfinish_kernel<<<dim3(batch, 2), kPoolThreads, sh2, stream>>>(ws, lora_to_fg_q, lora_to_bg_q, lora_to_k, pool_q, channels);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_pool_prep_batched(const afb_ada_prep_item* items, int n_items, int max_channels, int ctx_dim,
                                         int batch, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(items != nullptr && n_items > 0 && n_items <= 65535, "afb_ada_pool_prep_batched: 1..65535 items expected");
  AFB_REQUIRE(batch > 0 && ctx_dim > 0 && ctx_dim % 4 == 0, "afb_ada_pool_prep_batched: ctx_dim must be a positive multiple of 4");
  AFB_REQUIRE(max_channels % 64 == 0 && max_channels > 0 && max_channels <= 1280,
              "afb_ada_pool_prep_batched: max_channels %d must be a multiple of 64, <= 1280", max_channels);
  const int n_src = batch + 1;
  const int src_in_smem = n_src < kPrepMaxSrc ? n_src : kPrepMaxSrc;
  const size_t sh1 = static_cast<size_t>(src_in_smem) * ctx_dim * sizeof(float);
  AFB_REQUIRE(sh1 <= 96 * 1024, "afb_ada_pool_prep_batched: ctx_dim %d too large", ctx_dim);
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(ada_qproj_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set = true;
  }
  ada_qproj_batched_kernel<<<dim3((max_channels + kPoolWarps - 1) / kPoolWarps, n_items), kPoolThreads, sh1, stream>>>(items, ctx_dim,
                                                                                                                    n_src);
  count_launch();
  const size_t sh2 = (static_cast<size_t>(max_channels) + max_channels / kHeads) * sizeof(float);
  ada_qfinish_batched_kernel<<<dim3(batch, 2, n_items), kPoolThreads, sh2, stream>>>(items);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_pool(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask, float* fg_out,
                            float* bg_out, float* attn_fg, float* attn_bg, float* ws, int batch, int n_tokens, int channels,
                            float drop_p, uint32_t seed, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && q && pool_q && ws, "afb_ada_pool: NULL argument");
  // fg_out == bg_out == NULL: the chunk partials stay in ws and afb_ada_head_pooled merges them (no combine launch)
  const bool deferred = fg_out == nullptr && bg_out == nullptr;
  AFB_REQUIRE(deferred || (fg_out && bg_out), "afb_ada_pool: fg_out and bg_out must both be given or both NULL");
  AFB_REQUIRE(!deferred || attn_fg == nullptr, "afb_ada_pool: attention maps need the merged statistics (give fg_out / bg_out)");
  AFB_REQUIRE(batch > 0 && n_tokens > 0, "afb_ada_pool: empty input");
  AFB_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "afb_ada_pool: drop_p must be in [0, 1)");
  AFB_REQUIRE(channels % 64 == 0 && channels <= 1280, "afb_ada_pool: channels %d must be a multiple of 64, <= 1280", channels);
  AFB_REQUIRE((attn_fg == nullptr) == (attn_bg == nullptr), "afb_ada_pool: attn_fg and attn_bg must both be given or both NULL");
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) & 15) == 0, "afb_ada_pool: x/q must be 16-byte aligned");
  const int C = channels;
  const int nblocks = pool_blocks_per_sample(batch, n_tokens, channels);
  const int tokens_per_block = (n_tokens + nblocks - 1) / nblocks;
  const int grid_x = (n_tokens + tokens_per_block - 1) / tokens_per_block;
  float* scores = ws;
  float* partials = scores + static_cast<long long>(batch) * kHeads * 2 * n_tokens;
  float* stats = partials + static_cast<long long>(batch) * grid_x * pool_partial_floats(C);
  float* scores_arg = attn_fg != nullptr ? scores : nullptr;

  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(q);
  dim3 grid(grid_x, batch);
  const int chunks_per_lane = (C / 8 + 31) / 32;
  const size_t sh_b = kPoolWarps * (16 + 2 * static_cast<size_t>(C)) * sizeof(float);
#define AFB_LAUNCH_POOL(KCH)                                                                                              \
  {                                                                                                                       \
    auto kern = ada_pool_main_kernel<KCH>;                                                                                \
    const size_t sh_a = (2 * static_cast<size_t>(C) + kPoolWarps * 2 * (KCH * 32)) * sizeof(float);                       \
    const size_t sh_main = sh_a > sh_b ? sh_a : sh_b;                                                                     \
    static size_t set_for = 0;                                                                                            \
    if (set_for < sh_main) {                                                                                              \
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sh_main))); \
      set_for = sh_main;                                                                                                  \
    }                                                                                                                     \
    AFB_CHECK_CUDA(launch_pdl(kern, grid, dim3(kPoolThreads), sh_main, stream, 1, xb, qb, pool_q, key_mask, scores_arg,     \
                              partials, n_tokens, C, tokens_per_block, drop_p, seed));                                    \
  }
#define AFB_LAUNCH_POOL_MMA(KC)                                                                                           \
  {                                                                                                                       \
    auto kern = ada_pool_mma_kernel<KC>;                                                                                  \
    static bool set = false;                                                                                              \
    if (!set) {                                                                                                           \
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(PmCfg<KC>::kSmem))); \
      set = true;                                                                                                         \
    }                                                                                                                     \
    AFB_CHECK_CUDA(launch_pdl(kern, grid, dim3(kPmThreads), PmCfg<KC>::kSmem, stream, 1, xb, qb, pool_q, key_mask, scores_arg, \
                              partials, n_tokens, tokens_per_block, drop_p, seed));                                       \
  }
  // AFB_ADA_POOL=simt keeps the round-1 CUDA-core kernel (A/B measurement and cross-check); read once
  static const bool pool_mma = [] { const char* e = getenv("AFB_ADA_POOL"); return e == nullptr || e[0] != 's'; }();
  if (pool_mma && C == 320) AFB_LAUNCH_POOL_MMA(320)
  else if (pool_mma && C == 640) AFB_LAUNCH_POOL_MMA(640)
  else if (pool_mma && C == 1280) AFB_LAUNCH_POOL_MMA(1280)
  else if (chunks_per_lane <= 2) AFB_LAUNCH_POOL(2)
  else if (chunks_per_lane <= 3) AFB_LAUNCH_POOL(3)
  else AFB_LAUNCH_POOL(kMaxChunksPerLane)
#undef AFB_LAUNCH_POOL
#undef AFB_LAUNCH_POOL_MMA
  count_launch();
  if (!deferred) {
    AFB_CHECK_CUDA(launch_pdl(ada_pool_combine_kernel, dim3(batch, (2 * C + 255) / 256), dim3(256), grid_x * 8 * sizeof(float), stream, 1,
                              partials, grid_x, C, fg_out, bg_out, stats));
    count_launch();
  }
  if (attn_fg != nullptr) {
    const long long total = static_cast<long long>(batch) * kHeads * 2 * n_tokens;
    const long long blocks = (total + 255) / 256;
    ada_pool_probs_kernel<<<static_cast<unsigned>(blocks < 2048 ? blocks : 2048), 256, 0, stream>>>(scores, stats, attn_fg, attn_bg,
                                                                                                 n_tokens, total, drop_p, seed);
    count_launch();
  }
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_head(const float* infeat_fg, const float* infeat_bg, const float* time_emb, int time_ld,
                            const float* coeff_w, const float* coeff_b, const float* basis, const float* bias, float* out,
                            afb_bf16* ctx_out, const int* token_pos, int64_t ctx_sample_stride,
                            int64_t basis_sample_stride, int batch, int channels, int td, int num_vectors, int r,
                            int out_dim, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(infeat_fg && infeat_bg && time_emb && coeff_w && coeff_b && basis && out, "afb_ada_head: NULL argument");
  AFB_REQUIRE(batch > 0 && channels > 0 && td > 0 && td <= time_ld && num_vectors > 0 && r > 0 && out_dim > 0,
              "afb_ada_head: bad sizes");
  AFB_REQUIRE((2 * channels + td) % 4 == 0 && (reinterpret_cast<uintptr_t>(coeff_w) & 15) == 0,
              "afb_ada_head: coefficient map rows must be 16-byte aligned (2C + TD multiple of 4)");
  AFB_REQUIRE((ctx_out == nullptr) == (token_pos == nullptr), "afb_ada_head: ctx_out and token_pos go together");
  const size_t sh = (2 * static_cast<size_t>(channels) + td + static_cast<size_t>((r + 3) & ~3) + out_dim) * sizeof(float);
  AFB_REQUIRE(sh <= 48 * 1024, "afb_ada_head: feature vector too large for one block (%zu bytes)", sh);
  AFB_CHECK_CUDA(launch_pdl(ada_head_kernel, dim3(num_vectors, batch), dim3(256), sh, stream, 1, infeat_fg, infeat_bg, time_emb,
                            time_ld, coeff_w, coeff_b, basis, bias, out, reinterpret_cast<__nv_bfloat16*>(ctx_out), token_pos,
                            ctx_sample_stride, basis_sample_stride, channels, td, num_vectors, r, out_dim,
                            static_cast<const float*>(nullptr), 0, static_cast<float*>(nullptr), static_cast<float*>(nullptr),
                            static_cast<float*>(nullptr)));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_ada_head_pooled(const float* pool_ws, int n_tokens, float* fg_out, float* bg_out, const float* cached_bg,
                                   const float* time_emb, int time_ld, const float* coeff_w, const float* coeff_b,
                                   const float* basis, const float* bias, float* out, afb_bf16* ctx_out, const int* token_pos,
                                   int64_t ctx_sample_stride, int64_t basis_sample_stride, int batch, int channels, int td,
                                   int num_vectors, int r, int out_dim, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(pool_ws && time_emb && coeff_w && coeff_b && basis && out, "afb_ada_head_pooled: NULL argument");
  AFB_REQUIRE(batch > 0 && n_tokens > 0 && channels > 0 && td > 0 && td <= time_ld && num_vectors > 0 && r > 0 && out_dim > 0,
              "afb_ada_head_pooled: bad sizes");
  AFB_REQUIRE(channels % 64 == 0 && channels <= 1280, "afb_ada_head_pooled: channels %d must be a multiple of 64, <= 1280", channels);
  AFB_REQUIRE((2 * channels + td) % 4 == 0 && (reinterpret_cast<uintptr_t>(coeff_w) & 15) == 0,
              "afb_ada_head_pooled: coefficient map rows must be 16-byte aligned (2C + TD multiple of 4)");
  AFB_REQUIRE((ctx_out == nullptr) == (token_pos == nullptr), "afb_ada_head_pooled: ctx_out and token_pos go together");
  // the geometry afb_ada_pool used for the same (batch, n_tokens, channels)
  const int nblocks0 = pool_blocks_per_sample(batch, n_tokens, channels);
  const int tokens_per_block = (n_tokens + nblocks0 - 1) / nblocks0;
  const int grid_x = (n_tokens + tokens_per_block - 1) / tokens_per_block;
  const float* partials = pool_ws + static_cast<long long>(batch) * kHeads * 2 * n_tokens;
  float* stats = const_cast<float*>(partials) + static_cast<long long>(batch) * grid_x * pool_partial_floats(channels);
  const size_t sh = (2 * static_cast<size_t>(channels) + td + static_cast<size_t>((r + 3) & ~3) + out_dim +
                     static_cast<size_t>(grid_x) * 8) * sizeof(float);
  AFB_REQUIRE(sh <= 48 * 1024, "afb_ada_head_pooled: feature vector too large for one block (%zu bytes)", sh);
  AFB_CHECK_CUDA(launch_pdl(ada_head_kernel, dim3(num_vectors, batch), dim3(256), sh, stream, 1, static_cast<const float*>(nullptr),
                            cached_bg, time_emb, time_ld, coeff_w, coeff_b, basis, bias, out,
                            reinterpret_cast<__nv_bfloat16*>(ctx_out), token_pos, ctx_sample_stride, basis_sample_stride, channels,
                            td, num_vectors, r, out_dim, partials, grid_x, fg_out, bg_out, stats));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_static_layerwise_embedding(const float* basis_rand_weights, const float* basis_comm_weights,
                                              const float* pre_vecs, int n_pre, const float* basis_vecs, const float* bias,
                                              float* out, int num_layers, int num_vectors, int r, int out_dim,
                                              afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(basis_rand_weights && basis_comm_weights && basis_vecs && out, "afb_static_layerwise_embedding: NULL argument");
  AFB_REQUIRE(num_layers > 0 && num_layers <= 65535 && num_vectors > 0 && r > 0 && out_dim > 0,
              "afb_static_layerwise_embedding: bad sizes");
  AFB_REQUIRE(n_pre >= 0 && n_pre < r && (n_pre == 0) == (pre_vecs == nullptr),
              "afb_static_layerwise_embedding: pre_vecs and n_pre go together, and n_pre < r");
  const size_t sh = (static_cast<size_t>((r + 3) & ~3) + out_dim) * sizeof(float);
  AFB_REQUIRE(sh <= 48 * 1024, "afb_static_layerwise_embedding: vector too long for one block (%zu bytes)", sh);
  AFB_CHECK_CUDA(launch_pdl(static_layerwise_kernel, dim3(num_vectors, num_layers), dim3(256), sh, stream, 1, basis_rand_weights,
                            basis_comm_weights, pre_vecs, n_pre, basis_vecs, bias, out, num_vectors, r, out_dim));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_first_token_pos(const int64_t* tokens, int64_t token, int* first_pos, int batch, int n_tokens,
                                   afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(tokens && first_pos, "afb_first_token_pos: NULL argument");
  AFB_REQUIRE(batch > 0 && n_tokens > 0, "afb_first_token_pos: empty problem");
  AFB_CHECK_CUDA(launch_pdl(first_token_pos_kernel, dim3(batch), dim3(32), 0, stream, 1,
                            reinterpret_cast<const long long*>(tokens), static_cast<long long>(token), first_pos, n_tokens));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_splice_static_embedding(const float* embedded_text, const int* first_pos, const float* static_emb,
                                           float* out_f32, afb_bf16* out_bf16, int batch, int n_tokens, int num_layers,
                                           int num_vectors, int dim, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(embedded_text && first_pos && static_emb, "afb_splice_static_embedding: NULL argument");
  AFB_REQUIRE(out_f32 || out_bf16, "afb_splice_static_embedding: no output given");
  AFB_REQUIRE(batch > 0 && batch <= 65535 && num_layers > 0 && num_layers <= 65535 && n_tokens > 0 && num_vectors > 0,
              "afb_splice_static_embedding: bad sizes");
  AFB_REQUIRE(dim > 0 && dim % 4 == 0, "afb_splice_static_embedding: dim must be a multiple of 4");
  const uintptr_t all = reinterpret_cast<uintptr_t>(embedded_text) | reinterpret_cast<uintptr_t>(static_emb) |
                        reinterpret_cast<uintptr_t>(out_f32) | reinterpret_cast<uintptr_t>(out_bf16);
  AFB_REQUIRE((all & 15) == 0, "afb_splice_static_embedding: buffers must be 16-byte aligned");
  const int threads = dim / 4 < 256 ? ((dim / 4 + 31) / 32) * 32 : 256;
  AFB_CHECK_CUDA(launch_pdl(splice_static_kernel, dim3(n_tokens, num_layers, batch), dim3(threads), 0, stream, 1, embedded_text,
                            first_pos, static_emb, out_f32, reinterpret_cast<__nv_bfloat16*>(out_bf16), n_tokens, num_layers,
                            num_vectors, dim));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
