// GroupNorm (K1) and LayerNorm (K3) for the token-major bf16 activation layout.  HBM-bound kernels:
// fp32 statistics (two-pass, exact merge of per-chunk partials), vectorised coalesced streams.
// Replaces attention.py:71-72,325 (GroupNorm(32, C, eps=1e-6) + 'b c h w -> b (h w) c') and
// attention.py:267-269,277-283 (LayerNorm, eps=1e-5).
#include "common.cuh"
#include "../../include/adaface_b200.h"

#include <atomic>

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
// vector) x up to kGnTpt tokens held in registers (all loads in flight at once); a block covers T tokens with
// R = blockDim / (C/8) token rows.  Per-group reductions go through shared-memory atomics (32 groups).
//   gn_partial_tok   chunk (mean, M2) per (b, group, chunk): two passes over the REGISTER copy, one read of x; the
//                    last block of each sample merges the chunk partials (exact Chan merge) -> (mean, rstd)
//   gn_apply_tok     normalise + affine (+ SiLU) with 16-byte loads / stores
constexpr int kGnTpt = 8;           // tokens per thread
constexpr int kGnMaxGroups = 64;

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

// Tickets for the "last block of a sample finalises its statistics" pattern: self-resetting, one row of kGnMaxBatch
// counters per in-flight launch slot (the host rotates over kGnTicketSlots slots, so launches that overlap on
// different streams do not share counters unless more than kGnTicketSlots are in flight at once).
constexpr int kGnMaxBatch = 1024;
constexpr int kGnTicketSlots = 16;
__device__ unsigned int g_gn_tickets[kGnTicketSlots * kGnMaxBatch];

// pre_add (optional, fp32 [B, C] with row stride pre_add_ld): a per-(sample, channel) vector added to x BEFORE the
// statistics and the normalisation - the conv bias + time-embedding add that precedes out_layers' GroupNorm in a
// ResBlock (openaimodel.py:259-279), folded in so that no separate elementwise pass is needed.
__global__ void __launch_bounds__(320, 2) gn_partial_tok_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ pre_add,
                                                             int pre_add_ld, float* __restrict__ partial,
                                                             float* __restrict__ stats, int C, int HW, int cpg, int T,
                                                             int nchunks, float eps, int ticket_slot) {
  __shared__ float g_sum[kGnMaxGroups], g_m2[kGnMaxGroups];
  __shared__ unsigned int s_ticket;
  const int b = blockIdx.y, chunk = blockIdx.x;
  const int t0 = chunk * T;
  const int nt = min(T, HW - t0);
  const int vpr = C >> 3;                       // 16-byte vectors per token row
  const int R = blockDim.x / vpr;               // token rows covered per pass
  const int vec = threadIdx.x % vpr, row = threadIdx.x / vpr;
  const int c0 = vec * 8;
  const int groups = C / cpg;
  for (int i = threadIdx.x; i < groups; i += blockDim.x) g_sum[i] = g_m2[i] = 0.f;
  __syncthreads();
  const uint4* xp = reinterpret_cast<const uint4*>(x + (static_cast<long long>(b) * HW + t0) * C) + vec;
  uint4 data[kGnTpt];
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) {
    const int t = row + i * R;
    data[i] = (t < nt) ? xp[static_cast<long long>(t) * vpr] : make_uint4(0, 0, 0, 0);
  }
  float add[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) add[j] = 0.f;
  int my_rows = 0;                               // token rows this thread really holds
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) my_rows += (row + i * R < nt) ? 1 : 0;
  if (pre_add != nullptr) {
    const float4* ap = reinterpret_cast<const float4*>(pre_add + static_cast<long long>(b) * pre_add_ld + c0);
    const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
    add[0] = a0.x; add[1] = a0.y; add[2] = a0.z; add[3] = a0.w;
    add[4] = a1.x; add[5] = a1.y; add[6] = a1.z; add[7] = a1.w;
  }
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = static_cast<float>(my_rows) * add[j];
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) {
    float v[8];
    unpack8(data[i], v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j];   // rows beyond nt hold zeros
  }
  group_accumulate(g_sum, s, c0, cpg);
  __syncthreads();
  const float inv_n = 1.0f / static_cast<float>(nt * cpg);
  float mean[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) mean[j] = g_sum[(c0 + j) / cpg] * inv_n - add[j];   // (x + add) - mean = x - (mean - add)
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) {
    if (row + i * R < nt) {
      float v[8];
      unpack8(data[i], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = v[j] - mean[j];
        s[j] += d * d;
      }
    }
  }
  group_accumulate(g_m2, s, c0, cpg);
  __syncthreads();
  for (int g = threadIdx.x; g < groups; g += blockDim.x) {
    float* out = partial + (static_cast<long long>(b * groups + g) * nchunks + chunk) * 2;
    out[0] = g_sum[g] * inv_n;
    out[1] = g_m2[g];
  }
  // ---- the last block of sample b to get here merges the chunk partials (exact Chan merge) into (mean, rstd)
  __threadfence();
  __syncthreads();
  unsigned int* ticket = &g_gn_tickets[ticket_slot * kGnMaxBatch + b];
  if (threadIdx.x == 0) s_ticket = atomicAdd(ticket, 1u);
  __syncthreads();
  if (s_ticket != static_cast<unsigned int>(nchunks - 1)) return;
  __threadfence();
  if (threadIdx.x == 0) *ticket = 0u;            // self-reset for the next launch that uses this slot
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int g = warp; g < groups; g += nwarps) {
    const volatile float* p = partial + static_cast<long long>(b * groups + g) * nchunks * 2;
    float tot = 0.f, wsum = 0.f;
    for (int c = lane; c < nchunks; c += 32) {
      const float n = static_cast<float>(min(T, HW - c * T) * cpg);
      wsum += n * p[2 * c];
      tot += n;
    }
    tot = warp_sum(tot);
    const float gmean = warp_sum(wsum) / tot;
    float m2 = 0.f;
    for (int c = lane; c < nchunks; c += 32) {
      const float n = static_cast<float>(min(T, HW - c * T) * cpg);
      const float d = p[2 * c] - gmean;
      m2 += p[2 * c + 1] + n * d * d;
    }
    m2 = warp_sum(m2);
    if (lane == 0) {
      stats[2 * (b * groups + g)] = gmean;
      stats[2 * (b * groups + g) + 1] = rsqrtf(m2 / tot + eps);
    }
  }
}

__device__ __forceinline__ float silu(float v) { return __fdividef(v, 1.0f + __expf(-v)); }

template <bool kSilu>
__global__ void __launch_bounds__(320, 3) gn_apply_tok_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ stats,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const float* __restrict__ pre_add, int pre_add_ld,
                                                           __nv_bfloat16* __restrict__ y, int C, int HW, int cpg, int T) {
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * T;
  const int nt = min(T, HW - t0);
  const int vpr = C >> 3;
  const int R = blockDim.x / vpr;
  const int vec = threadIdx.x % vpr, row = threadIdx.x / vpr;
  const int c0 = vec * 8;
  const int groups = C / cpg;
  float sc[8], of[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    const float mean = stats[2 * (b * groups + g)], rstd = stats[2 * (b * groups + g) + 1];
    const float add = pre_add != nullptr ? pre_add[static_cast<long long>(b) * pre_add_ld + c0 + j] : 0.f;
    sc[j] = rstd * gamma[c0 + j];
    of[j] = beta[c0 + j] + (add - mean) * sc[j];
  }
  const long long base = (static_cast<long long>(b) * HW + t0) * vpr + vec;
  const uint4* xp = reinterpret_cast<const uint4*>(x) + base;
  uint4* yp = reinterpret_cast<uint4*>(y) + base;
  uint4 data[kGnTpt];
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) {
    const int t = row + i * R;
    if (t < nt) data[i] = xp[static_cast<long long>(t) * vpr];
  }
#pragma unroll
  for (int i = 0; i < kGnTpt; ++i) {
    const int t = row + i * R;
    if (t < nt) {
      float v[8];
      unpack8(data[i], v);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = v[j] * sc[j] + of[j];
        if (kSilu) v[j] = silu(v[j]);
      }
      yp[static_cast<long long>(t) * vpr] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                        pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
    }
  }
}

// launch geometry of the token-major kernels: threads per block, token rows per pass, tokens per block
struct GnTokGeom {
  int threads, rows, T, nchunks;
};
inline GnTokGeom gn_tok_geom(int batch, int channels, int hw) {
  GnTokGeom g;
  const int vpr = channels / 8;
  g.rows = 256 / vpr > 0 ? 256 / vpr : 1;
  g.threads = g.rows * vpr;
  g.T = g.rows * kGnTpt;
  // keep at least ~2 waves of blocks on the GPU when the tensor is small
  while (g.T > g.rows && static_cast<long long>(batch) * ((hw + g.T - 1) / g.T) < 2ll * num_sms()) g.T -= g.rows;
  if (g.T > hw) g.T = ((hw + g.rows - 1) / g.rows) * g.rows;
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
  using namespace afb;
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
    AFB_REQUIRE(batch <= kGnMaxBatch, "afb_groupnorm_tokens: batch %d > %d", batch, kGnMaxBatch);
    AFB_REQUIRE(pre_add == nullptr || (pre_add_ld >= channels && pre_add_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(pre_add) & 15) == 0),
                "afb_groupnorm_tokens: pre_add must be 16-byte aligned fp32 [B, C] with a row stride that is a multiple of 4");
    static std::atomic<unsigned int> launch_seq{0};
    const int slot = static_cast<int>(launch_seq.fetch_add(1, std::memory_order_relaxed) % kGnTicketSlots);
    gn_partial_tok_kernel<<<grid, geo.threads, 0, stream>>>(xb, pre_add, pre_add_ld, partial, stats_ws, channels, hw, cpg, geo.T,
                                                           geo.nchunks, eps, slot);
    count_launch();
    if (act) gn_apply_tok_kernel<true><<<grid, geo.threads, 0, stream>>>(xb, stats_ws, gamma, beta, pre_add, pre_add_ld, yb, channels, hw, cpg, geo.T);
    else gn_apply_tok_kernel<false><<<grid, geo.threads, 0, stream>>>(xb, stats_ws, gamma, beta, pre_add, pre_add_ld, yb, channels, hw, cpg, geo.T);
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
