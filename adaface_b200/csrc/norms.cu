// GroupNorm (K1) and LayerNorm (K3) for the token-major bf16 activation layout.  HBM-bound kernels:
// fp32 statistics (two-pass, exact merge of per-chunk partials), vectorised coalesced streams.
// Replaces attention.py:71-72,325 (GroupNorm(32, C, eps=1e-6) + 'b c h w -> b (h w) c') and
// attention.py:267-269,277-283 (LayerNorm, eps=1e-5).
#include "common.cuh"
#include "../../include/adaface_b200.h"

#include <cooperative_groups.h>

#include <mutex>


namespace afb {
void count_launch();
namespace {


__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (threadIdx.x < nwarps) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum(t);
    if (lane == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// ---- layout 0: x fp32 [B, C, HW]; one block per (b, group): contiguous cpg*HW floats ------------------
__global__ void __launch_bounds__(512) gn_stats_nchw_kernel(const float* __restrict__ x, float* __restrict__ stats,
                                                            int cpg_hw, float eps) {
  __shared__ float red[32];
  const float* base = x + static_cast<long long>(blockIdx.x) * cpg_hw;
  const float4* b4 = reinterpret_cast<const float4*>(base);
  const int n4 = cpg_hw >> 2;
  float s = 0.f;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    const float4 v = b4[i];
    s += (v.x + v.y) + (v.z + v.w);
  }
  for (int i = (n4 << 2) + threadIdx.x; i < cpg_hw; i += blockDim.x) s += base[i];
  const float mean = block_sum(s, red) / static_cast<float>(cpg_hw);
  float m2 = 0.f;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    const float4 v = b4[i];
    const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
    m2 += (a * a + b * b) + (c * c + d * d);
  }
  for (int i = (n4 << 2) + threadIdx.x; i < cpg_hw; i += blockDim.x) {
    const float a = base[i] - mean;
    m2 += a * a;
  }
  const float var = block_sum(m2, red) / static_cast<float>(cpg_hw);
  if (threadIdx.x == 0) {
    stats[2 * blockIdx.x] = mean;
    stats[2 * blockIdx.x + 1] = rsqrtf(var + eps);
  }
}

// y[b, t, c] = (x[b, c, t] - mean) * rstd * gamma[c] + beta[c]; 32 tokens x 64 channels per block
__global__ void __launch_bounds__(256) gn_apply_nchw_kernel(const float* __restrict__ x, const float* __restrict__ stats,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            __nv_bfloat16* __restrict__ y, int C, int HW, int cpg, int act) {
  __shared__ float tile[64][33];
  const int b = blockIdx.z, c0 = blockIdx.y * 64, t0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const int groups = C / cpg;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = c0 + ty + 8 * i;
    const int t = t0 + tx;
    float v = 0.f;
    if (c < C && t < HW) {
      const int g = c / cpg;
      const float mean = stats[2 * (b * groups + g)], rstd = stats[2 * (b * groups + g) + 1];
      v = (x[(static_cast<long long>(b) * C + c) * HW + t] - mean) * rstd * gamma[c] + beta[c];
      if (act == 1) v = __fdividef(v, 1.0f + __expf(-v));
    }
    tile[ty + 8 * i][tx] = v;
  }
  __syncthreads();
  // write: each thread packs 2 channels; 32 channel pairs x 8 tokens per pass
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = t0 + ty + 8 * i;
    const int c = c0 + 2 * tx;
    if (t < HW && c < C) {
      const uint32_t pk = pack_bf16x2(tile[2 * tx][ty + 8 * i], tile[2 * tx + 1][ty + 8 * i]);
      *reinterpret_cast<uint32_t*>(y + (static_cast<long long>(b) * HW + t) * C + c) = pk;
    }
  }
}

// ---- layout 1: x bf16 [B, HW, C] (token-major = channels-last) -------------------------------------------------
// A chunk of T consecutive tokens is one contiguous T*C*2-byte span.  Thread = 8 adjacent channels (one 16-byte
// vector); a block covers its T tokens in passes of R = blockDim / (C/8) token rows.  Per-group reductions go
// through shared-memory atomics (32 groups), once per block.
constexpr int kGnMaxGroups = 64;
constexpr int kGnUnroll = 4;        // token rows in flight per thread (16-byte loads)
constexpr int kGnMaxChunks = 32;    // chunk partials per sample: merged by one warp lane each in the apply kernel

__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
  const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}

// adds the 8 per-channel values of this thread into their groups' shared-memory slots (runs of equal group merged)
__device__ __forceinline__ void group_accumulate(float* slots, const float (&s)[8], int c0, int cpg) {
  int cur = c0 / cpg;
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    if (g != cur) {
      atomicAdd(&slots[cur], acc);
      acc = 0.f;
      cur = g;
    }
    acc += s[j];
  }
  atomicAdd(&slots[cur], acc);
}

// Token-major GroupNorm, two kernels, each block streaming a long run of tokens (T = HW / nchunks, nchunks <= 32 per
// sample) with kGnUnroll 16-byte loads in flight per thread and no block-wide synchronisation inside the stream:
//   gn_partial_tok   every thread owns 8 channels and walks its token rows accumulating SHIFTED sums
//                    sum(x - p), sum((x - p)^2) with p = its first value of that channel (one pass, no cancellation
//                    problem: p is a sample of the distribution); at the end per-channel (n, mean, M2) are merged into
//                    per-(sample, group, chunk) (mean, M2) with Chan's formula through two shared-memory reductions.
//   gn_apply_tok     prologue: one warp per group merges the <= 32 chunk partials of its sample (Chan) -> mean, rstd;
//                    then normalise + affine (+ SiLU) over the same token run.
// pre_add (optional, fp32 [B, C] with row stride pre_add_ld): a per-(sample, channel) vector added to x BEFORE the
// statistics and the normalisation - the conv bias + time-embedding add that precedes out_layers' GroupNorm in a
// ResBlock (openaimodel.py:259-279), folded in so that no separate elementwise pass is needed.
__global__ void __launch_bounds__(320, 2) gn_partial_tok_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ x2, int C1,
                                                             const float* __restrict__ pre_add,
                                                             int pre_add_ld, float* __restrict__ partial, int C, int HW,
                                                             int cpg, int T, int nchunks) {
  __shared__ float g_sum[kGnMaxGroups], g_m2[kGnMaxGroups];
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int t0 = chunk * T;
  const int nt = min(T, HW - t0);
  const int vpr = C >> 3;                       // 16-byte vectors per token row
  const int R = blockDim.x / vpr;               // token rows covered per pass
  const int vec = threadIdx.x % vpr, row = threadIdx.x / vpr;
  const int c0 = vec * 8;
  const int groups = C / cpg;
  for (int i = threadIdx.x; i < groups; i += blockDim.x) g_sum[i] = g_m2[i] = 0.f;
  // virtual channel concatenation (x2 != NULL): channels [0, C1) live in x [B, HW, C1], channels [C1, C) in
  // x2 [B, HW, C - C1] -- a thread owns one 8-channel vector, so it reads one of the two sources throughout
  const bool second = x2 != nullptr && c0 >= C1;
  const int src_c = x2 == nullptr ? C : (second ? C - C1 : C1);
  const int src_vpr = src_c >> 3;
  const uint4* xp = reinterpret_cast<const uint4*>((second ? x2 : x) + (static_cast<long long>(b) * HW + t0) * src_c) +
                    (second ? vec - (C1 >> 3) : vec);
  float piv[8], s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) piv[j] = s1[j] = s2[j] = 0.f;
  int n_mine = 0;
  if (row < nt) unpack8(xp[static_cast<long long>(row) * src_vpr], piv);
  // the next batch of rows is requested before the current one is reduced: the loads never drain
  uint4 data[kGnUnroll];
#pragma unroll
  for (int u = 0; u < kGnUnroll; ++u) {
    const int tt = row + u * R;
    if (tt < nt) data[u] = xp[static_cast<long long>(tt) * src_vpr];
  }
  for (int t = row; t < nt; t += R * kGnUnroll) {
    uint4 next[kGnUnroll];
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int tt = t + (kGnUnroll + u) * R;
      if (tt < nt) next[u] = xp[static_cast<long long>(tt) * src_vpr];
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      if (t + u * R < nt) {
        float v[8];
        unpack8(data[u], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[j] - piv[j];
          s1[j] += d;
          s2[j] = fmaf(d, d, s2[j]);
        }
        ++n_mine;
      }
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) data[u] = next[u];
  }
  // per-channel (n, mean, M2) of this thread's rows; pre_add shifts the mean only
  float add[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) add[j] = 0.f;
  if (pre_add != nullptr) {
    const float4* ap = reinterpret_cast<const float4*>(pre_add + static_cast<long long>(b) * pre_add_ld + c0);
    const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
    add[0] = a0.x; add[1] = a0.y; add[2] = a0.z; add[3] = a0.w;
    add[4] = a1.x; add[5] = a1.y; add[6] = a1.z; add[7] = a1.w;
  }
  const float n = static_cast<float>(n_mine);
  const float inv = n_mine > 0 ? 1.0f / n : 0.f;
  float mean_c[8], wsum[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float dm = s1[j] * inv;
    mean_c[j] = piv[j] + dm + add[j];
    s2[j] = fmaxf(s2[j] - s1[j] * dm, 0.f);     // M2 about this thread's own channel mean
    wsum[j] = n * mean_c[j];
  }
  __syncthreads();
  group_accumulate(g_sum, wsum, c0, cpg);
  __syncthreads();
  const float inv_n = 1.0f / static_cast<float>(nt * cpg);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float d = mean_c[j] - g_sum[(c0 + j) / cpg] * inv_n;
    s2[j] = fmaf(n * d, d, s2[j]);
  }
  group_accumulate(g_m2, s2, c0, cpg);
  __syncthreads();
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float* out = partial + (static_cast<long long>(b * groups + g) * nchunks + chunk) * 2;
    out[0] = g_sum[g] * inv_n;
    out[1] = g_m2[g];
  }
}

// x * sigmoid(x) = h + h * tanh(h), h = x / 2: ONE MUFU (tanh.approx, relative error 2^-11, below bf16 rounding)
// instead of ex2 + rcp -- the SiLU apply pass is otherwise bound by the 16-lane XU pipe, not by memory.
__device__ __forceinline__ float silu(float v) {
  const float h = 0.5f * v;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}

template <bool kSilu>
__global__ void __launch_bounds__(320, 2) gn_apply_tok_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ x2, int C1,
                                                           const float* __restrict__ partial,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const float* __restrict__ pre_add, int pre_add_ld,
                                                           __nv_bfloat16* __restrict__ y, int C, int HW, int cpg, int T,
                                                           int nchunks, float eps) {
  __shared__ float s_mean[kGnMaxGroups], s_rstd[kGnMaxGroups];
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * T;
  const int nt = min(T, HW - t0);
  const int vpr = C >> 3;
  const int R = blockDim.x / vpr;
  const int vec = threadIdx.x % vpr, row = threadIdx.x / vpr;
  const int c0 = vec * 8;
  const int groups = C / cpg;
  const long long base = (static_cast<long long>(b) * HW + t0) * vpr + vec;
  const bool second = x2 != nullptr && c0 >= C1;                     // virtual concatenation, see gn_partial_tok_kernel
  const int src_vpr = x2 == nullptr ? vpr : ((second ? C - C1 : C1) >> 3);
  const uint4* xp = reinterpret_cast<const uint4*>(second ? x2 : x) + (static_cast<long long>(b) * HW + t0) * src_vpr +
                    (second ? vec - (C1 >> 3) : vec);
  uint4* yp = reinterpret_cast<uint4*>(y) + base;
  // first loads go out before the statistics merge so their latency overlaps it
  uint4 data[kGnUnroll];
#pragma unroll
  for (int u = 0; u < kGnUnroll; ++u) {
    const int tt = row + u * R;
    if (tt < nt) data[u] = xp[static_cast<long long>(tt) * src_vpr];
  }
  {
    // merge the chunk partials of sample b: lane c holds chunk c (nchunks <= 32), exact Chan merge
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int g = warp; g < groups; g += nwarps) {
      const float* pp = partial + static_cast<long long>(b * groups + g) * nchunks * 2;
      float nk = 0.f, mk = 0.f, m2k = 0.f;
      if (lane < nchunks) {
        nk = static_cast<float>(min(T, HW - lane * T) * cpg);
        mk = pp[2 * lane];
        m2k = pp[2 * lane + 1];
      }
      const float tot = warp_sum(nk);
      const float gmean = warp_sum(nk * mk) / tot;
      const float d = mk - gmean;
      const float m2 = warp_sum(m2k + nk * d * d);
      if (lane == 0) {
        s_mean[g] = gmean;
        s_rstd[g] = rsqrtf(m2 / tot + eps);
      }
    }
  }
  __syncthreads();
  float sc[8], of[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    const float add = pre_add != nullptr ? pre_add[static_cast<long long>(b) * pre_add_ld + c0 + j] : 0.f;
    sc[j] = s_rstd[g] * gamma[c0 + j];
    of[j] = beta[c0 + j] + (add - s_mean[g]) * sc[j];
  }
  for (int t = row; t < nt; t += R * kGnUnroll) {
    uint4 next[kGnUnroll];
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int tt = t + (kGnUnroll + u) * R;
      if (tt < nt) next[u] = xp[static_cast<long long>(tt) * src_vpr];
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) {
      const int tt = t + u * R;
      if (tt < nt) {
        float v[8];
        unpack8(data[u], v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[j] = v[j] * sc[j] + of[j];
          if (kSilu) v[j] = silu(v[j]);
        }
        yp[static_cast<long long>(tt) * vpr] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                           pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
      }
    }
#pragma unroll
    for (int u = 0; u < kGnUnroll; ++u) data[u] = next[u];
  }
}

// ---- register-resident slab variant --------------------------------------------------------------------------
// One CTA = (sample, slab of S = lcm(C/groups, 16) channels = S / cpg WHOLE groups), all HW tokens of the slab held in
// registers (kItems <= 16 16-byte vectors per thread, <= 320 threads, HW * S * 2 <= 80 KB): x is read once, the variance is the exact two-pass
// form Sum (x - mean)^2 over the registers, no atomics (deterministic), no second launch, no inter-CTA exchange -- the
// groups of a slab live in this CTA alone.  A token contributes S * 2 = 160..480 contiguous bytes: whole 32-byte
// sectors (with S = lcm(cpg, 8) = 40 channels every other slab boundary fell inside a sector, the partial-sector
// stores made L2 fetch the output lines from DRAM first and the kernel ran at 1.2 TB/s, measured), and the activations
// of these sizes (<= 21 MB) sit in L2 anyway.
// Replaces the round-1 8-CTA cluster kernel (span staged in shared memory, partials over DSMEM, two cluster barriers:
// 10-30 us for 2.6-16 MB tensors against 4.6-11 us here) and the two-kernel path for every tensor whose slab fits.
constexpr int kGnSlabMaxThreads = 320;
constexpr int kGnSlabMaxGroups = 8;

template <int kItems, bool kSilu>
__global__ void __launch_bounds__(kGnSlabMaxThreads, 1)
gn_slab_tok_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ x2, int C1,
                   const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ pre_add,
                   int pre_add_ld, __nv_bfloat16* __restrict__ y, int C, int HW, int cpg, int S, int R, float eps) {
  __shared__ float s_pa[kGnSlabMaxThreads], s_pb[kGnSlabMaxThreads];      // per-thread partials of its (first, second) group
  __shared__ float s_tot[kGnSlabMaxGroups];
  pdl_wait();
  const int b = blockIdx.y;
  const int vps = S >> 3;                               // 16-byte vectors per token of the slab
  const int vec = threadIdx.x % vps, row = threadIdx.x / vps;
  const bool valid = row < R;                           // the block is rounded up to whole warps
  const int c0 = blockIdx.x * S + vec * 8;
  const int g0 = (blockIdx.x * S) / cpg;                // first group of the slab
  const int ngroups = S / cpg;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool second = x2 != nullptr && c0 >= C1;        // virtual concatenation: a slab lies in one source (C1 % S == 0)
  const int src_c = x2 == nullptr ? C : (second ? C - C1 : C1);
  const __nv_bfloat16* src = (second ? x2 : x) + static_cast<long long>(b) * HW * src_c + (second ? c0 - C1 : c0);
  uint4 data[kItems];
#pragma unroll
  for (int u = 0; u < kItems; ++u) {
    const int t = row + u * R;
    if (valid && t < HW) data[u] = *reinterpret_cast<const uint4*>(src + static_cast<long long>(t) * src_c);
  }
  pdl_trigger();
  // parameters: requested ahead of the two reduction rounds
  float add[8], gm[8], bt[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    add[j] = 0.f;
    gm[j] = 1.f;
    bt[j] = 0.f;
  }
  if (valid) {
    const float4* gp = reinterpret_cast<const float4*>(gamma + c0);
    const float4* bp = reinterpret_cast<const float4*>(beta + c0);
    const float4 ga = __ldg(gp), gb = __ldg(gp + 1), ba = __ldg(bp), bb = __ldg(bp + 1);
    gm[0] = ga.x; gm[1] = ga.y; gm[2] = ga.z; gm[3] = ga.w; gm[4] = gb.x; gm[5] = gb.y; gm[6] = gb.z; gm[7] = gb.w;
    bt[0] = ba.x; bt[1] = ba.y; bt[2] = ba.z; bt[3] = ba.w; bt[4] = bb.x; bt[5] = bb.y; bt[6] = bb.z; bt[7] = bb.w;
    if (pre_add != nullptr) {
      const float4* ap = reinterpret_cast<const float4*>(pre_add + static_cast<long long>(b) * pre_add_ld + c0);
      const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
      add[0] = a0.x; add[1] = a0.y; add[2] = a0.z; add[3] = a0.w;
      add[4] = a1.x; add[5] = a1.y; add[6] = a1.z; add[7] = a1.w;
    }
  }
  // a thread's 8 channels lie in at most two groups (cpg >= 8): channels [0, split) in group ga, the rest in ga + 1
  const int ga_ = c0 / cpg - g0;
  const int split = min(8, (g0 + ga_ + 1) * cpg - c0);
  // block-wide per-group sum of the threads' (first-group, second-group) partials: warp g scans all threads' partials
  // of group g (thread i's first group follows from i % vps), then one shuffle reduction
  auto block_group_sum = [&](float pa, float pb) {
    __syncthreads();                                    // s_tot of the previous round has been read
    s_pa[threadIdx.x] = valid ? pa : 0.f;
    s_pb[threadIdx.x] = valid ? pb : 0.f;
    __syncthreads();
    for (int g = warp; g < ngroups; g += (blockDim.x >> 5)) {
      float acc = 0.f;
      for (int i = lane; i < blockDim.x; i += 32) {
        const int gi = (blockIdx.x * S + (i % vps) * 8) / cpg - g0;
        acc += (gi == g ? s_pa[i] : 0.f) + (gi + 1 == g ? s_pb[i] : 0.f);
      }
      acc = warp_sum(acc);
      if (lane == 0) s_tot[g] = acc;
    }
    __syncthreads();
  };
  // ---- mean
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  int n_mine = 0;
#pragma unroll
  for (int u = 0; u < kItems; ++u) {
    if (valid && row + u * R < HW) {
      float v[8];
      unpack8(data[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += v[j];
      ++n_mine;
    }
  }
  float pa = 0.f, pb = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float v = s[j] + static_cast<float>(n_mine) * add[j];
    if (j < split) pa += v;
    else pb += v;
  }
  block_group_sum(pa, pb);
  const float inv_n = 1.0f / (static_cast<float>(HW) * static_cast<float>(cpg));
  const float mean_a = s_tot[ga_] * inv_n, mean_b = split < 8 ? s_tot[ga_ + 1] * inv_n : 0.f;
  float off[8];                                         // add - mean of the channel's group
#pragma unroll
  for (int j = 0; j < 8; ++j) off[j] = add[j] - (j < split ? mean_a : mean_b);
  // ---- variance about the mean
  pa = 0.f;
  pb = 0.f;
#pragma unroll
  for (int u = 0; u < kItems; ++u) {
    if (valid && row + u * R < HW) {
      float v[8];
      unpack8(data[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[j] + off[j];
        if (j < split) pa = fmaf(d, d, pa);
        else pb = fmaf(d, d, pb);
      }
    }
  }
  block_group_sum(pa, pb);
  if (!valid) return;
  const float rstd_a = rsqrtf(s_tot[ga_] * inv_n + eps), rstd_b = split < 8 ? rsqrtf(s_tot[ga_ + 1] * inv_n + eps) : 0.f;
  float sc[8], of[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = (j < split ? rstd_a : rstd_b) * gm[j];
    of[j] = fmaf(off[j], sc[j], bt[j]);
  }
  __nv_bfloat16* dst = y + static_cast<long long>(b) * HW * C + c0;
#pragma unroll
  for (int u = 0; u < kItems; ++u) {
    const int t = row + u * R;
    if (t < HW) {
      float v[8];
      unpack8(data[u], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = fmaf(v[j], sc[j], of[j]);
        if (kSilu) v[j] = silu(v[j]);
      }
      *reinterpret_cast<uint4*>(dst + static_cast<long long>(t) * C) =
          make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    }
  }
}

// AFB_GN_SLAB=0 keeps every tensor on the two-kernel path (A/B measurement); read once
bool gn_slab_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_GN_SLAB");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

struct GnSlabGeom {
  int S, items, R, threads;     // items == 0: the tensor does not fit the slab kernel
};
inline GnSlabGeom gn_slab_geom(int batch, int channels, int channels1, bool cat, int hw, int cpg) {
  GnSlabGeom g{0, 0, 0, 0};
  int S = cpg;
  while (S % 16 != 0) S += cpg;                         // lcm(cpg, 16): slabs begin and end on 32-byte sectors
  if (S > channels || channels % S != 0 || S / cpg > kGnSlabMaxGroups || cpg < 8) return g;
  if (cat && channels1 % S != 0) return g;
  if (static_cast<long long>(hw) * S > 40960) return g; // 80 KB of registers per CTA
  const int vps = S / 8;
  for (int items = 1; items <= 16; items *= 2) {
    const int R = (hw + items - 1) / items;
    const int threads = R * vps;
    if (threads <= kGnSlabMaxThreads) {
      // eight and more vectors per thread leave one CTA per SM: beyond two waves the two-kernel path is faster
      // (measured: [16, 256, 2560] 21.0 vs 19.0 us)
      if (items >= 8 && static_cast<long long>(channels / S) * batch > 2 * num_sms()) return g;
      g.S = S;
      g.items = items;
      g.R = R;
      g.threads = (threads + 31) / 32 * 32;
      return g;
    }
  }
  return g;
}

template <int kItems>
int launch_gn_slab(const GnSlabGeom& g, const __nv_bfloat16* x, const __nv_bfloat16* x2, int C1, const float* gamma,
                   const float* beta, const float* pre_add, int pre_add_ld, __nv_bfloat16* y, int batch, int C, int HW,
                   int cpg, float eps, int act, cudaStream_t stream) {
  auto kern = act ? gn_slab_tok_kernel<kItems, true> : gn_slab_tok_kernel<kItems, false>;
  AFB_CHECK_CUDA(launch_pdl(kern, dim3(C / g.S, batch, 1), dim3(g.threads, 1, 1), 0, stream, 1, x, x2, C1, gamma, beta, pre_add,
                            pre_add_ld, y, C, HW, cpg, g.S, g.R, eps));
  return 0;
}

// launch geometry of the token-major kernels: threads per block, token rows per pass, tokens per block
// resident blocks per SM of gn_apply_tok_kernel at a given block size (register-limited), cached per size
int gn_apply_blocks_per_sm(int threads) {
  static int cache[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};      // index threads / 32, threads <= 320
  int& slot = cache[threads / 32];
  if (slot == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, gn_apply_tok_kernel<true>, threads, 0) != cudaSuccess || n < 1) n = 2;
    slot = n;
  }
  return slot;
}

struct GnTokGeom {
  int threads, rows, T, nchunks;
};
inline GnTokGeom gn_tok_geom(int batch, int channels, int hw) {
  GnTokGeom g;
  const int vpr = channels / 8;
  g.rows = 256 / vpr > 0 ? 256 / vpr : 1;
  g.threads = g.rows * vpr;
  // ONE wave of the apply kernel (the heavier of the two passes: 90 registers) over the whole launch -- a second,
  // partly filled wave costs up to 40 % --, at most kGnMaxChunks chunks per sample, at least one pass of rows
  int want = gn_apply_blocks_per_sm(g.threads) * num_sms() / batch;
  if (want > kGnMaxChunks) want = kGnMaxChunks;
  if (want < 1) want = 1;
  g.T = (hw + want - 1) / want;
  g.T = ((g.T + g.rows - 1) / g.rows) * g.rows;
  g.nchunks = (hw + g.T - 1) / g.T;
  return g;
}

// ---- LayerNorm: one warp per row, row kept in registers (C <= 8*32*kMaxChunks) -------------------------
template <int kMaxChunks>
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const __nv_bfloat16* __restrict__ x,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta,
                                                             __nv_bfloat16* __restrict__ y, long long rows, int C,
                                                             float eps) {
  const long long row = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunks = C >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * C);
  float v[kMaxChunks][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks) {
      const uint4 u = xr[ch];
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y;
      v[i][4] = c.x; v[i][5] = c.y; v[i][6] = d.x; v[i][7] = d.y;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[i][j];
    }
  }
  const float mean = warp_sum(s) / static_cast<float>(C);
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    if (lane + 32 * i < nchunks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[i][j] - mean;
        m2 += d * d;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(m2) / static_cast<float>(C) + eps);
  uint4* yr = reinterpret_cast<uint4*>(y + row * C);
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int ch = lane + 32 * i;
    if (ch < nchunks) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd;
      if (gamma != nullptr) {
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + 8 * ch));
        const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + 8 * ch + 4));
        o[0] *= g0.x; o[1] *= g0.y; o[2] *= g0.z; o[3] *= g0.w;
        o[4] *= g1.x; o[5] *= g1.y; o[6] *= g1.z; o[7] *= g1.w;
      }
      if (beta != nullptr) {
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + 8 * ch));
        const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta + 8 * ch + 4));
        o[0] += b0.x; o[1] += b0.y; o[2] += b0.z; o[3] += b0.w;
        o[4] += b1.x; o[5] += b1.y; o[6] += b1.z; o[7] += b1.w;
      }
      yr[ch] = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                          pack_bf16x2(o[6], o[7]));
    }
  }
}

}  // namespace
}  // namespace afb

extern "C" int64_t afb_groupnorm_ws_floats(int batch, int hw, int groups) {
  // worst case of gn_tok_geom: one token row per chunk
  return 2ll * batch * groups * (static_cast<int64_t>(hw) + 1);
}

extern "C" int afb_groupnorm_tokens(const void* x, int x_layout, const float* gamma, const float* beta, afb_bf16* y,
                                    float* stats_ws, int batch, int channels, int hw, int groups, float eps, int act,
                                    const float* pre_add, int pre_add_ld, afb_stream_t stream_) {
  return afb_groupnorm_tokens_cat(x, 0, nullptr, x_layout, gamma, beta, y, stats_ws, batch, channels, hw, groups, eps, act,
                                  pre_add, pre_add_ld, stream_);
}

extern "C" int afb_groupnorm_tokens_cat(const void* x, int channels1, const afb_bf16* x2, int x_layout, const float* gamma,
                                        const float* beta, afb_bf16* y, float* stats_ws, int batch, int channels, int hw,
                                        int groups, float eps, int act, const float* pre_add, int pre_add_ld,
                                        afb_stream_t stream_) {
  using namespace afb;
  const __nv_bfloat16* xb2 = reinterpret_cast<const __nv_bfloat16*>(x2);
  if (x2 != nullptr) {
    AFB_REQUIRE(x_layout == 1 && channels1 > 0 && channels1 < channels && channels1 % 8 == 0 && (channels - channels1) % 8 == 0 &&
                    (reinterpret_cast<uintptr_t>(x2) & 15) == 0,
                "afb_groupnorm_tokens_cat: two token-major bf16 sources with channel counts that are multiples of 8");
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && gamma && beta && y && stats_ws, "afb_groupnorm_tokens: NULL argument");
  AFB_REQUIRE(batch > 0 && channels > 0 && hw > 0 && groups > 0, "afb_groupnorm_tokens: empty input");
  AFB_REQUIRE(channels % groups == 0, "afb_groupnorm_tokens: channels %d not divisible by groups %d", channels, groups);
  const int cpg = channels / groups;
  AFB_REQUIRE(cpg % 2 == 0 && channels % 64 == 0, "afb_groupnorm_tokens: need channels %% 64 == 0 and even channels/group");
  AFB_REQUIRE(act == 0 || act == 1, "afb_groupnorm_tokens: act must be 0 (none) or 1 (SiLU)");
  __nv_bfloat16* yb = reinterpret_cast<__nv_bfloat16*>(y);
  if (x_layout == 0) {
    AFB_REQUIRE(pre_add == nullptr, "afb_groupnorm_tokens: pre_add is only supported for the token-major layout");
    AFB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (static_cast<long long>(cpg) * hw) % 4 == 0,
                "afb_groupnorm_tokens: NCHW input must be 16-byte aligned with (C/groups)*HW %% 4 == 0");
    gn_stats_nchw_kernel<<<batch * groups, 512, 0, stream>>>(static_cast<const float*>(x), stats_ws, cpg * hw, eps);
    count_launch();
    dim3 grid((hw + 31) / 32, channels / 64, batch);
    gn_apply_nchw_kernel<<<grid, 256, 0, stream>>>(static_cast<const float*>(x), stats_ws, gamma, beta, yb, channels, hw, cpg, act);
    count_launch();
  } else if (x_layout == 1) {
    AFB_REQUIRE(channels / 8 <= 320 && groups <= kGnMaxGroups,
                "afb_groupnorm_tokens: channels %d / groups %d not supported by the token-major kernel", channels, groups);
    AFB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0,
                "afb_groupnorm_tokens: buffers must be 16-byte aligned");
    const GnTokGeom geo = gn_tok_geom(batch, channels, hw);
    float* partial = stats_ws + 2ll * batch * groups;
    const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(x);
    dim3 grid(geo.nchunks, batch);
    AFB_REQUIRE(batch <= 65535, "afb_groupnorm_tokens: batch %d exceeds the grid limit", batch);
    AFB_REQUIRE(pre_add == nullptr || (pre_add_ld >= channels && pre_add_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(pre_add) & 15) == 0),
                "afb_groupnorm_tokens: pre_add must be 16-byte aligned fp32 [B, C] with a row stride that is a multiple of 4");
    // tensors whose (sample, group-slab) fits the registers of one CTA: one launch, x read once, deterministic
    if (gn_slab_enabled()) {
      const GnSlabGeom sg = gn_slab_geom(batch, channels, channels1, x2 != nullptr, hw, cpg);
      if (sg.items != 0) {
        int rc = 0;
        switch (sg.items) {
          case 1: rc = launch_gn_slab<1>(sg, xb, xb2, channels1, gamma, beta, pre_add, pre_add_ld, yb, batch, channels, hw, cpg, eps, act, stream); break;
          case 2: rc = launch_gn_slab<2>(sg, xb, xb2, channels1, gamma, beta, pre_add, pre_add_ld, yb, batch, channels, hw, cpg, eps, act, stream); break;
          case 4: rc = launch_gn_slab<4>(sg, xb, xb2, channels1, gamma, beta, pre_add, pre_add_ld, yb, batch, channels, hw, cpg, eps, act, stream); break;
          case 8: rc = launch_gn_slab<8>(sg, xb, xb2, channels1, gamma, beta, pre_add, pre_add_ld, yb, batch, channels, hw, cpg, eps, act, stream); break;
          default: rc = launch_gn_slab<16>(sg, xb, xb2, channels1, gamma, beta, pre_add, pre_add_ld, yb, batch, channels, hw, cpg, eps, act, stream); break;
        }
        if (rc) return rc;
        count_launch();
        return 0;
      }
    }
    AFB_CHECK_CUDA(launch_pdl(gn_partial_tok_kernel, grid, dim3(geo.threads), 0, stream, 1, xb, xb2, channels1, pre_add, pre_add_ld, partial, channels, hw,
                              cpg, geo.T, geo.nchunks));
    count_launch();
    AFB_CHECK_CUDA(launch_pdl(act ? gn_apply_tok_kernel<true> : gn_apply_tok_kernel<false>, grid, dim3(geo.threads), 0, stream, 1, xb, xb2, channels1,
                              partial, gamma, beta, pre_add, pre_add_ld, yb, channels, hw, cpg, geo.T, geo.nchunks, eps));
    count_launch();
  } else {
    AFB_FAIL(2, "afb_groupnorm_tokens: unknown x_layout %d", x_layout);
  }
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_layernorm_bf16(const afb_bf16* x, const float* gamma, const float* beta, afb_bf16* y, int64_t rows,
                                  int channels, float eps, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && y, "afb_layernorm_bf16: NULL argument");
  AFB_REQUIRE(rows > 0 && channels > 0, "afb_layernorm_bf16: empty input");
  AFB_REQUIRE(channels % 8 == 0 && channels <= 8 * 32 * 12, "afb_layernorm_bf16: channels %d must be a multiple of 8, <= 3072",
              channels);
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0,
              "afb_layernorm_bf16: buffers must be 16-byte aligned");
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yb = reinterpret_cast<__nv_bfloat16*>(y);
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  const int chunks_per_lane = (channels / 8 + 31) / 32;
  if (chunks_per_lane <= 2)
    layernorm_bf16_kernel<2><<<grid, 256, 0, stream>>>(xb, gamma, beta, yb, rows, channels, eps);
  else if (chunks_per_lane <= 5)
    layernorm_bf16_kernel<5><<<grid, 256, 0, stream>>>(xb, gamma, beta, yb, rows, channels, eps);
  else
    layernorm_bf16_kernel<12><<<grid, 256, 0, stream>>>(xb, gamma, beta, yb, rows, channels, eps);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- allocation-free variant: the statistics workspace lives in the library ------------------------------------
// afb_groupnorm_tokens_static_ws == afb_groupnorm_tokens with `stats_ws` taken from a file-scope device array instead
// of a caller buffer, for callers that cannot allocate (CUDA graph capture, tight serving loops).  No counterpart in
// the reference.  Every GPU that loads the library holds the array
// (kGnStaticWsFloats * 4 bytes = 1 MiB) until the process exits.
namespace afb {
namespace {
constexpr int64_t kGnStaticWsFloats = 256 * 1024;
static __device__ __align__(256) float g_gn_static_ws[kGnStaticWsFloats];

// In host code `g_gn_static_ws` is only the host-side shadow of the symbol, not a device address: the address a
// kernel may use is resolved once per device (with that device current) through cudaGetSymbolAddress and cached.
constexpr int kGnStaticMaxDevices = 64;
float* g_gn_static_ws_addr[kGnStaticMaxDevices] = {};
std::mutex g_gn_static_ws_mu;

int gn_static_ws(float** out) {
  int dev = -1;
  AFB_CHECK_CUDA(cudaGetDevice(&dev));
  AFB_REQUIRE(dev >= 0 && dev < kGnStaticMaxDevices, "afb_groupnorm_tokens_static_ws: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_gn_static_ws_mu);
  if (g_gn_static_ws_addr[dev] == nullptr) {
    void* p = nullptr;
    AFB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_gn_static_ws));
    AFB_REQUIRE(p != nullptr, "afb_groupnorm_tokens_static_ws: cudaGetSymbolAddress returned NULL");
    g_gn_static_ws_addr[dev] = static_cast<float*>(p);
  }
  *out = g_gn_static_ws_addr[dev];
  return 0;
}
}  // namespace
}  // namespace afb

extern "C" int64_t afb_groupnorm_static_ws_max_floats(void) { return afb::kGnStaticWsFloats; }

extern "C" int afb_groupnorm_tokens_static_ws(const void* x, int x_layout, const float* gamma, const float* beta, afb_bf16* y,
                                              int batch, int channels, int hw, int groups, float eps, int act,
                                              const float* pre_add, int pre_add_ld, afb_stream_t stream) {
  using namespace afb;
  AFB_REQUIRE(batch > 0 && groups > 0, "afb_groupnorm_tokens_static_ws: empty input");
  // what the kernels touch: 2*B*G floats of (mean, rstd) + 2*B*G*nchunks chunk partials, nchunks <= kGnMaxChunks
  const int64_t need = 2ll * batch * groups * (1 + kGnMaxChunks);
  AFB_REQUIRE(need <= kGnStaticWsFloats,
              "afb_groupnorm_tokens_static_ws: batch %d x groups %d needs %lld workspace floats, the built-in scratch holds "
              "%lld (batch * groups <= %lld); use afb_groupnorm_tokens with a caller workspace",
              batch, groups, static_cast<long long>(need), static_cast<long long>(kGnStaticWsFloats),
              static_cast<long long>(kGnStaticWsFloats / (2 * (1 + kGnMaxChunks))));
  float* ws = nullptr;
  if (const int rc = gn_static_ws(&ws)) return rc;
  // Nothing in the workspace has to start at zero: the statistics kernel writes every (sample, group, chunk) slot the
  // apply kernel reads in the same call, so the previous call's contents are never observed.
  return afb_groupnorm_tokens_cat(x, 0, nullptr, x_layout, gamma, beta, y, ws, batch, channels, hw, groups, eps, act,
                                  pre_add, pre_add_ld, stream);
}
