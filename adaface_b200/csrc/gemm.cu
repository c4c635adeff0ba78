// Persistent, warp-specialised BF16 GEMM for sm_100a:  Y = epilogue(A[M,K] * W[N,K]^T).
//
// Hot variant (bf16 out, optional bf16 residual, plain or GEGLU epilogue) - gemm_bf16_tma_kernel, 384 threads:
//   warp 0 (one lane)  TMA producer: A tile 128x64 and W tile BNx64 (both K-major, SWIZZLE_128B) into a
//                      kStages-deep shared-memory ring, mbarrier complete_tx signalling
//   warp 1 (one lane)  tcgen05.mma issuer: 128 x BN x 16 UMMAs, fp32 accumulators in TMEM, two
//                      accumulator buffers so the epilogue of tile i overlaps the main loop of tile i+1
//   warp 2             TMEM allocator
//   warp 3 (one lane)  staging manager: when the output staging buffer of tile i is free again, TMA-loads the
//                      RESIDUAL tile of tile i into it (one tile ahead of the epilogue, so DRAM latency is hidden)
//   warps 4..11        epilogue, 2 warps per TMEM lane quarter: tcgen05.ld 32 columns -> + bias (+ residual read
//                      in place from the staging box) [GEGLU: value * gelu(gate)] -> bf16 -> staging box
//                      ([128 rows x 32 cols], SWIZZLE_64B, conflict-free 16-byte accesses) -> TMA store (fully
//                      coalesced writes, clipped at the M / N edges by the tensor map)
// Cold variants (fp32 out, fp32 residual, NCHW-transposing epilogue of the fp32 drop-in path) -
// gemm_bf16_kernel, 256 threads, registers -> global directly.
//
// Replaces the nn.Linear / 1x1 Conv2d calls of ldm/modules/attention.py (:35, :55, :157-165, :302-317).
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();

namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kThreads = 256;

struct GemmDev {
  void* out;
  const float* bias;
  const void* residual;
  int M, N, K;
  long long ldo, ldr;
  int out_f32, res_f32, hw;
  int num_m_blocks, num_n_blocks, num_k_blocks;
};

template <int BN>
struct Cfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN <= 128) ? 6 : 5;
  static constexpr int kTmemCols = (2 * BN <= 256) ? 256 : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  static_assert(kBBytes % 1024 == 0, "W tile must keep 1024B alignment of the next stage");
};

template <int BN, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                 const GemmDev p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::kStages;
  uint64_t* tmem_full = bars + 2 * C::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(&tm_w);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  pdl_trigger();

  const int num_tiles = p.num_m_blocks * p.num_n_blocks;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / p.num_n_blocks;
        const int n_blk = tile % p.num_n_blocks;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
          tma_load_2d(&tm_a, &full_bar[stage], sa, kb * kBK, m_blk * kBM);
          tma_load_2d(&tm_w, &full_bar[stage], sb, kb * kBK, n_blk * BN, kEvictLast);
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    {   // converged warp, one elected issuer (see gemm_bf16_tma_kernel)
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, 0, 0);
      const bool leader = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (leader) {
            const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
            const uint64_t da = umma_desc_kmajor_sw128(sa);
            const uint64_t db = umma_desc_kmajor_sw128(sa + C::kABytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the 128B swizzle atom: +2 in the >>4 address field
              umma_ss(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
            umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (leader) umma_commit(&tmem_full[acc]);
        __syncwarp();
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m_blk = tile / p.num_n_blocks;
      const int n_blk = tile % p.num_n_blocks;
      const int m = m_blk * kBM + q * 32 + lane;
      const bool row_ok = m < p.M;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int ch = 0; ch < BN / 32; ++ch) {
        uint32_t r[32];
        tmem_ld32(taddr + ch * 32, r);
        tmem_ld_wait();
        const int n0 = n_blk * BN + ch * 32;
        if (n0 >= p.N) continue;
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (p.bias != nullptr) {
          const float4* b4 = reinterpret_cast<const float4*>(p.bias + n0);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(b4 + j);
            v[4 * j + 0] += b.x;
            v[4 * j + 1] += b.y;
            v[4 * j + 2] += b.z;
            v[4 * j + 3] += b.w;
          }
        }
        if constexpr (EPI == AFB_EPI_STORE) {
          if (row_ok) {
            if (p.residual != nullptr) {
              if (p.res_f32) {
                const float4* r4 = reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) +
                                                                   static_cast<long long>(m) * p.ldr + n0);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float4 t = r4[j];
                  v[4 * j + 0] += t.x;
                  v[4 * j + 1] += t.y;
                  v[4 * j + 2] += t.z;
                  v[4 * j + 3] += t.w;
                }
              } else {
                const uint4* r4 = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.residual) +
                                                                 static_cast<long long>(m) * p.ldr + n0);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint4 t = r4[j];
                  const float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y), c = unpack_bf16x2(t.z),
                               d = unpack_bf16x2(t.w);
                  v[8 * j + 0] += a.x;
                  v[8 * j + 1] += a.y;
                  v[8 * j + 2] += b.x;
                  v[8 * j + 3] += b.y;
                  v[8 * j + 4] += c.x;
                  v[8 * j + 5] += c.y;
                  v[8 * j + 6] += d.x;
                  v[8 * j + 7] += d.y;
                }
              }
            }
            if (p.out_f32) {
              float4* o4 = reinterpret_cast<float4*>(static_cast<float*>(p.out) + static_cast<long long>(m) * p.ldo + n0);
#pragma unroll
              for (int j = 0; j < 8; ++j) o4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
              uint4* o4 = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + static_cast<long long>(m) * p.ldo + n0);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                o4[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                                   pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
            }
          }
        } else {  // AFB_EPI_NCHW: out[b][n][t] fp32, lanes run along t (coalesced)
          if (row_ok) {
            const int b = m / p.hw;
            const int t = m - b * p.hw;
            const long long base = (static_cast<long long>(b) * p.N + n0) * p.hw + t;
            float* o = static_cast<float*>(p.out) + base;
            const float* rs = static_cast<const float*>(p.residual);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              float val = v[j];
              if (rs != nullptr) val += rs[base + static_cast<long long>(j) * p.hw];
              o[static_cast<long long>(j) * p.hw] = val;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<C::kTmemCols>(tmem_base);
}

// ============================================================================================
// Hot variant: TMA-staged epilogue
// ============================================================================================
constexpr int kBoxBytes = kBM * 32 * 2;      // one staging box: 128 rows x 32 bf16 columns

// kPair: two CTAs of a cluster (one TPC) compute a 256 x BN tile with cta_group::2 MMAs; each stages its own 128
// rows of A and HALF of the W tile, so the operand bytes an SM pulls from L2 per output drop by 25-30%.
// kWide (pair mode, STORE epilogues): a unit's tile is 2 * BN columns wide -- every A stage feeds TWO MMAs, one per BN-wide
// half, so the activation tile crosses the L2 -> SM fabric once per 2 * BN output columns instead of once per BN
// (operand bytes per CTA and k-block for 2 x 128 x 160 outputs: 16 + 2 * 10 KB instead of 2 * (16 + 10) KB).  The halves
// are separate accumulators of BN columns rotating through THREE TMEM slots (3 * 160 = 480 of 512 columns): the epilogue
// and the staging manager keep working on BN-wide half tiles, one slot drains while the other two fill.
template <int BN, bool GEGLU, bool kPair = false, bool kWide = false>
struct TCfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBRows = kPair ? BN / 2 : BN;                  // rows of one W box
  static constexpr int kBBoxBytes = kBRows * kBK * 2;
  static constexpr int kHalves = kWide ? 2 : 1;
  static constexpr int kBBytes = kHalves * kBBoxBytes;
  static constexpr int kAccSlots = kWide ? 3 : 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kOutChunks = GEGLU ? BN / 64 : BN / 32;        // 32-column output boxes per tile
  // warps 0-3: TMA producer, MMA issuer, TMEM allocator, staging manager; then the epilogue warps, four (one per TMEM
  // lane quarter) per column group.  The GEGLU epilogue is ~35 issue slots per output element: it needs four warps
  // per scheduler to keep up with the tensor pipe at K = 320.
  // (measured: sixteen epilogue warps for the STORE epilogues change nothing -- those GEMMs are bound by the SM <-> L2
  //  traffic of their operands and outputs, not by the epilogue arithmetic; -DAFB_GEMM_EPI_WARPS=16 rebuilds that way)
#ifndef AFB_GEMM_EPI_WARPS
#define AFB_GEMM_EPI_WARPS 8
#endif
  // (wide tiles, measured: the mainloop's operand wait drops by a third, then the MMA issuer waits for a free accumulator
  //  instead -- a wide tile needs TWO of the three slots, so only one half tile of epilogue overlaps the next mainloop; a
  //  second team of eight epilogue warps, one per half, changed nothing for that reason: 20.9 vs 19.5 us at
  //  16384 x 640 x 640, and was removed)
#ifndef AFB_GEMM_WIDE_EPI_WARPS
#define AFB_GEMM_WIDE_EPI_WARPS 8
#endif
  static constexpr int kEpiWarps = GEGLU ? 16 : (kWide ? AFB_GEMM_WIDE_EPI_WARPS : AFB_GEMM_EPI_WARPS);
  static constexpr int kGroups = kEpiWarps / 4;
  static constexpr int kThreads = 128 + 32 * kEpiWarps;
  static constexpr int kStagingBytes = 2 * kOutChunks * kBoxBytes;     // double-buffered
  static constexpr int kColParamBytes = 2 * 2 * BN * 4;                // {bias, colsum} of a tile's columns, double-buffered
  static constexpr int kRowParamBytes = 2 * kBM * 8;                   // {rstd, -mean * rstd} of a tile's rows, double-buffered
  static constexpr int kParamBytes = kColParamBytes + kRowParamBytes;
  static constexpr int kBarrierBytes = 512;
  static constexpr int kStages = (227 * 1024 - 1024 - kBarrierBytes - kStagingBytes - kParamBytes) / kStageBytes;
  static constexpr int kTmemCols = (kAccSlots * BN <= 256) ? 256 : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kParamBytes + 1024 + kBarrierBytes;
  static_assert(kAccSlots * BN <= 512, "accumulator slots exceed TMEM");
  static_assert(!kWide || (kPair && !GEGLU), "wide tiles: pair mode, STORE epilogues");
  static_assert(kBBoxBytes % 1024 == 0, "W tile must keep 1024B alignment of the next stage");
  static_assert(kStages >= 3 && kStages <= 16, "pipeline depth");
  static_assert(!kPair || BN % 32 == 0, "pair mode: each CTA stages BN/2 rows of W, MMA N must be a multiple of 16");
  static_assert(GEGLU ? BN % 64 == 0 : BN % 32 == 0, "tile width vs output boxes");
};

#ifdef AFB_GEMM_PROFILE
// developer probe (tools/gemm_probe.cu): cycles the MMA issuer and one epilogue warp of the first CTAs spend per phase
__device__ long long* g_gemm_prof = nullptr;
#define AFB_GPROF_DECL long long gp_t = clock64(); long long gp_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define AFB_GPROF(i) { const long long gp_n = clock64(); gp_acc[i] += gp_n - gp_t; gp_t = gp_n; }
#define AFB_GPROF_FLUSH(slot) if (lane == 0 && g_gemm_prof != nullptr && blockIdx.x < 4) { for (int gi = 0; gi < 8; ++gi) g_gemm_prof[(blockIdx.x * 4 + (slot)) * 8 + gi] = gp_acc[gi]; }
#else
#define AFB_GPROF_DECL
#define AFB_GPROF(i)
#define AFB_GPROF_FLUSH(slot)
#endif

// tile = unit + i * num_units walked as (m_unit, n_blk): one division pair at the start instead of one per tile in
// every warp role (~300 cycles per tile of the epilogue's serial path)
struct TileWalk {
  int tile, m_unit, n_blk, dq, dr, nnb, step;
  __device__ TileWalk(int unit, int num_units, int nnb_)
      : tile(unit), m_unit(unit / nnb_), n_blk(unit % nnb_), dq(num_units / nnb_), dr(num_units % nnb_), nnb(nnb_), step(num_units) {}
  __device__ void next() {
    tile += step;
    m_unit += dq;
    n_blk += dr;
    if (n_blk >= nnb) {
      n_blk -= nnb;
      ++m_unit;
    }
  }
};

struct GemmTmaDev {
  const float* bias;
  int M, N, K;
  int has_res;
  int num_m_blocks, num_n_blocks, num_k_blocks;
  float* stats_out;          // [M][2 * num_n_blocks][2] or NULL
  const float* ln_stats;     // [M][ln_parts][2] or NULL
  const float* ln_colsum;    // [N]
  int ln_parts;
  float ln_eps;
  // 3x3 convolution as an implicit GEMM: A is the NHWC image [B, H, W, Cin] behind a 4-D tensor map, the K loop
  // walks (tap, 64-channel block); conv_cblocks = Cin / 64 (0: plain GEMM)
  int conv_cblocks, conv_h, conv_w;
  int conv_stride;           // 1, or 2: conv_h / conv_w are the OUTPUT size, tap coordinates are 2 * pixel + (dy, dx)
  // conv_boxed != 0: the tile's 128 rows are a (bb, bh, bw) sub-box of the image that is NOT a run of consecutive pixels
  // (image widths like 96 / 48 / 24 / 12: bw = the largest power of two dividing W, ...): row block m_blk is box
  // (m_blk % conv_tw, (m_blk / conv_tw) % conv_th, m_blk / (conv_tw * conv_th)), and the output / residual tiles move
  // through 4-D tensor maps over [B, H, W, N] (5-D with one store per image for the folded upsampling)
  int conv_boxed, conv_bw, conv_bh, conv_bb, conv_tw, conv_th;
  // nearest-2x upsampling folded into the convolution (conv_fold != 0): output phase (py, px) = n_blk / nb_per_phase
  // is a 2x2 convolution of the LOW-resolution image with taps (py - 1 + a, px - 1 + c); its BN-column blocks store to
  // the pixels (2i + py, 2j + px) of the high-resolution output through a 5-D tensor map; fold_cout = N / 4
  int conv_fold, nb_per_phase, fold_cout;
  // A virtually concatenated along K (plain GEMM only): k-blocks [0, k1_blocks) come from tm_a, the rest from tm_a2
  int k1_blocks;
  // K split (k_splits > 1): an output tile is computed by k_splits units, each over num_k_blocks / k_splits k-blocks,
  // and their bf16 partial tiles are ADDED into the pre-zeroed output by TMA reduce-stores; split 0 carries the bias and
  // the residual.  For launches with fewer tiles than SM pairs and a deep K (the 3x3 convolutions at 8x8: 32 tiles).
  int k_splits;
};

template <int BN, bool GEGLU, bool kLn, bool kStats, bool kPair, bool kWide = false>
__global__ void __launch_bounds__((TCfg<BN, GEGLU, kPair, kWide>::kThreads), 1)
gemm_bf16_tma_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_w,
                     const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_res,
                     const __grid_constant__ CUtensorMap tm_a2,
                     const GemmTmaDev p) {
  using C = TCfg<BN, GEGLU, kPair, kWide>;
  constexpr int NC = C::kOutChunks;
  constexpr int kSlots = C::kAccSlots;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + C::kStages * C::kStageBytes;
  float* params = reinterpret_cast<float*>(staging + C::kStagingBytes);      // [2 buffers][bias BN | colsum BN]
  float2* row_params = reinterpret_cast<float2*>(staging + C::kStagingBytes + C::kColParamBytes);   // [2 buffers][kBM]
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + C::kStagingBytes + C::kParamBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::kStages;
  uint64_t* tmem_full = bars + 2 * C::kStages;
  uint64_t* tmem_empty = tmem_full + 3;
  uint64_t* out_free = tmem_empty + 3;     // staging buffer b: its TMA stores have finished reading shared memory
  uint64_t* res_full = out_free + 2;       // staging buffer b: residual landed (or: buffer handed to the epilogue)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // work unit: one CTA (kPair: one CTA pair); a unit's tile is kBM (kPair: 2 * kBM) rows by BN columns
  const int cta_rank = kPair ? static_cast<int>(cluster_ctarank()) : 0;
  const int unit = kPair ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int num_units = kPair ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int num_m_units = kPair ? (p.num_m_blocks + 1) >> 1 : p.num_m_blocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a);
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_out);
    if (p.has_res) tma_prefetch_desc(&tm_res);
    if (p.k1_blocks < p.num_k_blocks) tma_prefetch_desc(&tm_a2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < kSlots; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], (kPair ? 2 : 1) * C::kEpiWarps);      // pair: the leader collects both CTAs' epilogue warps
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&out_free[i], C::kGroups);
      mbar_init(&res_full[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (kPair) tmem_alloc_pair<C::kTmemCols>(tmem_slot);
    else tmem_alloc<C::kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (kPair) cluster_sync_relaxed();  // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // everything above touched only this CTA's shared memory / TMEM and overlapped the previous kernel's tail
  pdl_wait();
  pdl_trigger();

  // tiles of the producer / MMA walk: BN (kWide: 2 * BN) columns wide
  const int nnb_t = (kWide ? p.num_n_blocks >> 1 : p.num_n_blocks) * p.k_splits;     // the split index is the fastest one
  const int num_tiles = num_m_units * nnb_t;
  const int kb_per_split = p.num_k_blocks / p.k_splits;
  auto walk_m_blk = [&](const TileWalk& t) { return kPair ? 2 * t.m_unit + cta_rank : t.m_unit; };

  if (warp == 0) {
    // ------------------------------------------------------------------ operand producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (TileWalk t(unit, num_units, nnb_t); t.tile < num_tiles; t.next()) {
        const int m_blk = walk_m_blk(t);
        const int split = t.n_blk % p.k_splits, nt_blk = t.n_blk / p.k_splits;
        const int n_blk = kWide ? 2 * nt_blk : nt_blk;
        const int kb0 = split * kb_per_split;
        // convolution: the tile's 128 rows are the pixel box (b0.., h0.., w0..) of the image; tap (dy, dx) of the
        // filter reads the same box shifted by (dy, dx), out-of-image pixels zero-filled by TMA (= zero padding)
        int cw0 = 0, ch0 = 0, cb0 = 0, cblk = 0, dy = -1, dx = -1, dx0 = -1, dx_end = 2;
        if (p.conv_cblocks != 0) {
          if (p.conv_boxed != 0) {
            const int q = m_blk / p.conv_tw;
            cw0 = (m_blk - q * p.conv_tw) * p.conv_bw;
            cb0 = q / p.conv_th;
            ch0 = (q - cb0 * p.conv_th) * p.conv_bh;
            cb0 *= p.conv_bb;
          } else {
            const int m0 = m_blk * kBM;
            const int r = m0 / p.conv_w;
            cw0 = m0 - r * p.conv_w;
            cb0 = r / p.conv_h;
            ch0 = r - cb0 * p.conv_h;
          }
          if (p.conv_fold != 0) {
            const int phase = n_blk / p.nb_per_phase;
            dy = (phase >> 1) - 1;
            dx0 = (phase & 1) - 1;
            dx = dx0;
            dx_end = dx0 + 2;
          } else if (kb0 != 0) {                      // a later K split starts inside the tap walk
            const int tap = kb0 / p.conv_cblocks;
            cblk = kb0 - tap * p.conv_cblocks;
            dy = tap / 3 - 1;
            dx = tap - (tap / 3) * 3 - 1;
          }
        }
        for (int kb = kb0; kb < kb0 + kb_per_split; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::kStageBytes;
          uint8_t* sb = sa + C::kABytes;
          if constexpr (kPair) {
            // both CTAs' bytes are counted on the leader's barrier; the peer's may land before the leader arms the
            // phase (the pending arrival keeps it open), never before the previous phase closed (empty_bar)
            if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * C::kStageBytes);
            if (p.conv_cblocks != 0) tma_load_4d_pair(&tm_a, &full_bar[stage], sa, cblk * kBK, p.conv_stride * cw0 + dx, p.conv_stride * ch0 + dy, cb0);
            else if (kb < p.k1_blocks) tma_load_2d_pair(&tm_a, &full_bar[stage], sa, kb * kBK, m_blk * kBM);
            else tma_load_2d_pair(&tm_a2, &full_bar[stage], sa, (kb - p.k1_blocks) * kBK, m_blk * kBM);
            tma_load_2d_pair(&tm_w, &full_bar[stage], sb, kb * kBK, n_blk * BN + cta_rank * C::kBRows, kEvictLast);
            if constexpr (kWide)
              tma_load_2d_pair(&tm_w, &full_bar[stage], sb + C::kBBoxBytes, kb * kBK, (n_blk + 1) * BN + cta_rank * C::kBRows, kEvictLast);
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
            if (p.conv_cblocks != 0) tma_load_4d(&tm_a, &full_bar[stage], sa, cblk * kBK, p.conv_stride * cw0 + dx, p.conv_stride * ch0 + dy, cb0);
            else if (kb < p.k1_blocks) tma_load_2d(&tm_a, &full_bar[stage], sa, kb * kBK, m_blk * kBM);
            else tma_load_2d(&tm_a2, &full_bar[stage], sa, (kb - p.k1_blocks) * kBK, m_blk * kBM);
            tma_load_2d(&tm_w, &full_bar[stage], sb, kb * kBK, n_blk * BN, kEvictLast);
          }
          if (p.conv_cblocks != 0 && ++cblk == p.conv_cblocks) {
            cblk = 0;
            if (++dx == dx_end) {
              dx = dx0;
              ++dy;
            }
          }
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // The whole warp walks the loop converged and one elected lane issues: in divergent code (if (lane == 0)) the
    // compiler wraps every tcgen05 instruction in an election loop and keeps the descriptors in vector registers,
    // ~10 extra instructions per MMA on the one thread that must never fall behind the tensor pipe.
    if (cta_rank == 0) {      // pair: the even CTA issues for both
      constexpr uint32_t idesc = umma_idesc_bf16(kPair ? 2 * kBM : kBM, BN, 0, 0);
      const bool leader = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      AFB_GPROF_DECL
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        const uint32_t tmem_d = tmem_base + acc * BN;
        // wide tile: the second half accumulates in the next slot of the ring
        int acc1 = acc;
        uint32_t acc1_phase = acc_phase;
        if constexpr (kWide) {
          if (++acc1 == kSlots) {
            acc1 = 0;
            acc1_phase ^= 1;
          }
          mbar_wait(&tmem_empty[acc1], acc1_phase ^ 1);
        }
        const uint32_t tmem_d1 = tmem_base + acc1 * BN;
        tc_fence_after();
        AFB_GPROF(0)
        for (int kb = 0; kb < kb_per_split; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          AFB_GPROF(1)
          if (leader) {
            const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
            const uint64_t da = umma_desc_kmajor_sw128(sa);
            const uint64_t db = umma_desc_kmajor_sw128(sa + C::kABytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k) {
              if constexpr (kPair) umma_ss_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
              else umma_ss(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
            if constexpr (kWide) {
              const uint64_t db1 = umma_desc_kmajor_sw128(sa + C::kABytes + C::kBBoxBytes);
#pragma unroll
              for (int k = 0; k < kBK / 16; ++k) umma_ss_pair(tmem_d1, da + 2 * k, db1 + 2 * k, idesc, (kb | k) != 0);
            }
            if constexpr (kPair) umma_commit_pair(&empty_bar[stage]);
            else umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          AFB_GPROF(2)
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (leader) {
          if constexpr (kPair) umma_commit_pair(&tmem_full[acc]);
          else umma_commit(&tmem_full[acc]);
          if constexpr (kWide) umma_commit_pair(&tmem_full[acc1]);
        }
        __syncwarp();
        if constexpr (kWide) {
          acc = acc1;
          acc_phase = acc1_phase;
        }
        if (++acc == kSlots) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
      AFB_GPROF_FLUSH(0)
    }
  } else if (warp == 3) {
    // ------------------------------------------------------------------ staging manager
    // One tile ahead of the epilogue: copies the tile's column parameters (bias, LayerNorm column sums) into shared
    // memory -- every epilogue thread needs all of them, and 16-32 uniform LDG.128 per thread per box were its top
    // stall --, reduces the producer GEMM's partial row sums to the LayerNorm (rstd, -mean * rstd) of the tile's rows
    // (a dependent-load chain of ~2000 cycles per tile when the epilogue warps did it), and prefetches the residual
    // boxes by TMA.
    int it = 0;
    for (TileWalk t(unit, num_units, nnb_t); t.tile < num_tiles; t.next())
    for (int half = 0; half < C::kHalves; ++half, ++it) {      // BN-wide half tiles of a wide tile, in accumulator-slot order
      const int sb = it & 1;
      const uint32_t use = static_cast<uint32_t>(it >> 1);
      // global loads first -- they do not depend on the staging buffer and run ahead of the wait for it
      constexpr int kColVecs = (BN / 4 + 31) / 32;           // float4 per lane covering the tile's BN columns
      const int split = t.n_blk % p.k_splits, nt_blk = t.n_blk / p.k_splits;
      const int n_blk = kWide ? 2 * nt_blk + half : nt_blk;
      const bool tile_res = p.has_res && split == 0;      // K split: split 0 carries bias and residual
      float4 cb[kColVecs], cc[kColVecs];
#pragma unroll
      for (int u = 0; u < kColVecs; ++u) {
        const int i = lane + 32 * u;
        const int n0 = n_blk * BN + 4 * i;
        cb[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        cc[u] = cb[u];
        if (i < BN / 4 && n0 < p.N) {
          if (p.bias != nullptr && split == 0) cb[u] = __ldg(reinterpret_cast<const float4*>(p.bias + n0));
          if constexpr (kLn) cc[u] = __ldg(reinterpret_cast<const float4*>(p.ln_colsum + n0));
        }
      }
      float2 rp[4];                                            // four rows per lane
      if constexpr (kLn) {
        const int m0 = walk_m_blk(t) * kBM;
        const int nq = p.ln_parts >> 1;                        // parts is even: two (sum, sum of squares) partials per 16 bytes
        const float inv_k = 1.0f / static_cast<float>(p.K);
        float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
        for (int i0 = 0; i0 < nq; i0 += 4) {
          float4 t[4][4];
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const int m_row = min(m0 + r * 32 + lane, p.M - 1);
            const float4* sp = reinterpret_cast<const float4*>(p.ln_stats + static_cast<long long>(m_row) * p.ln_parts * 2);
#pragma unroll
            for (int i = 0; i < 4; ++i) t[r][i] = (i0 + i < nq) ? __ldg(sp + i0 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              s1[r] += t[r][i].x + t[r][i].z;
              s2[r] += t[r][i].y + t[r][i].w;
            }
          }
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float mean = s1[r] * inv_k;
          const float rstd = rsqrtf(fmaxf(s2[r] * inv_k - mean * mean, 0.f) + p.ln_eps);
          rp[r] = make_float2(rstd, -mean * rstd);
        }
      }
      mbar_wait(&out_free[sb], (use & 1) ^ 1);
      {
        float4* pb = reinterpret_cast<float4*>(params + sb * 2 * BN);
#pragma unroll
        for (int u = 0; u < kColVecs; ++u) {
          const int i = lane + 32 * u;
          if (i < BN / 4) {
            pb[i] = cb[u];
            if constexpr (kLn) pb[BN / 4 + i] = cc[u];
          }
        }
        if constexpr (kLn) {
#pragma unroll
          for (int r = 0; r < 4; ++r) row_params[sb * kBM + r * 32 + lane] = rp[r];
        }
        __syncwarp();      // lane 0's arrive below publishes the whole warp's writes
      }
      if (lane == 0) {
        if (tile_res) {
          const int m_blk = walk_m_blk(t);
          int valid = 0;
#pragma unroll
          for (int ch = 0; ch < NC; ++ch)
            if (n_blk * BN + ch * 32 < p.N) ++valid;
          mbar_arrive_expect_tx(&res_full[sb], valid * kBoxBytes);
#pragma unroll
          for (int ch = 0; ch < NC; ++ch) {
            const int n0 = n_blk * BN + ch * 32;
            if (n0 < p.N) {
              if (p.conv_boxed != 0) {
                const int q = m_blk / p.conv_tw, bq = q / p.conv_th;
                tma_load_4d(&tm_res, &res_full[sb], staging + (sb * NC + ch) * kBoxBytes, n0, (m_blk - q * p.conv_tw) * p.conv_bw,
                            (q - bq * p.conv_th) * p.conv_bh, bq * p.conv_bb, kEvictFirst);
              } else {
                tma_load_2d(&tm_res, &res_full[sb], staging + (sb * NC + ch) * kBoxBytes, n0, m_blk * kBM, kEvictFirst);
              }
            }
          }
        } else {
          mbar_arrive(&res_full[sb]);
        }
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue
    const int q = warp & 3;                         // TMEM lane quarter
    const int grp = (warp - 4) >> 2;                // column group
    const int ch_lo = (grp * NC) / C::kGroups;
    const int ch_hi = ((grp + 1) * NC) / C::kGroups;
    const bool leader = (q == 0) && (lane == 0);    // issues this group's TMA stores
    const int row = q * 32 + lane;
    const uint32_t sw = static_cast<uint32_t>((row >> 1) & 3);
    int acc = 0;
    uint32_t acc_phase = 0;
    int it = 0;
    auto release_acc = [&](int a) {      // this warp has drained accumulator a
      if constexpr (kPair) mbar_arrive_cluster(&tmem_empty[a], 0);
      else mbar_arrive(&tmem_empty[a]);
    };
    AFB_GPROF_DECL
    for (TileWalk t(unit, num_units, nnb_t); t.tile < num_tiles; t.next())
    for (int half = 0; half < C::kHalves; ++half, ++it) {      // BN-wide half tiles of a wide tile, in accumulator-slot order
      const int m_blk = walk_m_blk(t);
      const int split = t.n_blk % p.k_splits, nt_blk = t.n_blk / p.k_splits;
      const int n_blk = kWide ? 2 * nt_blk + half : nt_blk;
      const bool tile_res = p.has_res && split == 0;      // K split: split 0 carries bias and residual
      const int sb = it & 1;
      const uint32_t use = static_cast<uint32_t>(it >> 1);
      const int m_row = m_blk * kBM + row;
      AFB_GPROF(0)
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      AFB_GPROF(1)
      if (leader) {
        // the stores of the previous tile (other staging buffer) have certainly been issued: once they have
        // finished reading shared memory, hand that buffer back to the staging manager
        bulk_wait_group_read<0>();
        if (it > 0) mbar_arrive(&out_free[sb ^ 1]);
      }
      AFB_GPROF(2)
      mbar_wait(&res_full[sb], use & 1);
      AFB_GPROF(3)
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      bool released = false;
      // LayerNorm fold: rstd * (acc - mean * colsum) = rstd * acc + ln_nm * colsum
      float ln_rstd = 1.f, ln_nm = 0.f;
      if constexpr (kLn) {
        const float2 rp = row_params[sb * kBM + row];
        ln_rstd = rp.x;
        ln_nm = rp.y;
      }
      float st_s1 = 0.f, st_s2 = 0.f;
#pragma unroll 1
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        const int n_out0 = (GEGLU ? n_blk * (BN / 2) : n_blk * BN) + ch * 32;     // first output column of this box
        const int n_acc0 = GEGLU ? n_blk * BN + ch * 64 : n_out0;                  // first accumulator column
        if (n_acc0 >= p.N) break;
        // this box's column parameters (shared memory, same address in every lane: broadcast reads)
        const uint32_t pbias = smem_u32(params + sb * 2 * BN + (n_acc0 - n_blk * BN));
        const uint32_t pcs = pbias + BN * 4;
        // staging box [128 rows][32 cols] bf16, 64-byte rows, SWIZZLE_64B: 16-byte chunk c of row r sits at c ^ ((r >> 1) & 3)
        uint8_t* brow = staging + (sb * NC + ch) * kBoxBytes + row * 64;
        uint32_t pk[16];                                // the box row: 32 bf16
        float v[32];
        if constexpr (GEGLU) {
          const uint64_t rstd2 = pack_f32x2(ln_rstd, ln_rstd), nm2 = pack_f32x2(ln_nm, ln_nm);
          // Eight output columns at a time, the next eight already in flight: the TMEM read port moves 64 B per clock
          // per SM (2048 cycles for a 128 x 256 fp32 tile), so the loads of all sixteen warps have to overlap the
          // arithmetic instead of alternating with it.  Two register stages of 8 + 8 fit the 96-register budget.
          uint32_t ra[2][8], rg[2][8];
          tmem_ld8(taddr + ch * 64, ra[0]);
          tmem_ld8(taddr + ch * 64 + 32, rg[0]);
#pragma unroll
          for (int qt = 0; qt < 4; ++qt) {
            const int cur = qt & 1;
            tmem_ld_wait_on(ra[cur], rg[cur]);
            AFB_GPROF(6)
            if (qt < 3) {
              tmem_ld8(taddr + ch * 64 + (qt + 1) * 8, ra[cur ^ 1]);
              tmem_ld8(taddr + ch * 64 + 32 + (qt + 1) * 8, rg[cur ^ 1]);
            } else if (ch == ch_hi - 1 || n_acc0 + 64 >= p.N) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) release_acc(acc);
              released = true;
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const int jj = qt * 2 + j;
              const float4 ba = lds128(pbias + 16 * jj), bg = lds128(pbias + 16 * (8 + jj));
              uint64_t a01 = pack_f32x2(__uint_as_float(ra[cur][4 * j + 0]), __uint_as_float(ra[cur][4 * j + 1]));
              uint64_t a23 = pack_f32x2(__uint_as_float(ra[cur][4 * j + 2]), __uint_as_float(ra[cur][4 * j + 3]));
              uint64_t g01 = pack_f32x2(__uint_as_float(rg[cur][4 * j + 0]), __uint_as_float(rg[cur][4 * j + 1]));
              uint64_t g23 = pack_f32x2(__uint_as_float(rg[cur][4 * j + 2]), __uint_as_float(rg[cur][4 * j + 3]));
              if constexpr (kLn) {
                // rstd * (acc - mean * colsum) + bias = rstd * acc + (ln_nm * colsum + bias)
                const float4 ca = lds128(pcs + 16 * jj), cg = lds128(pcs + 16 * (8 + jj));
                a01 = ffma2(rstd2, a01, ffma2(nm2, pack_f32x2(ca.x, ca.y), pack_f32x2(ba.x, ba.y)));
                a23 = ffma2(rstd2, a23, ffma2(nm2, pack_f32x2(ca.z, ca.w), pack_f32x2(ba.z, ba.w)));
                g01 = ffma2(rstd2, g01, ffma2(nm2, pack_f32x2(cg.x, cg.y), pack_f32x2(bg.x, bg.y)));
                g23 = ffma2(rstd2, g23, ffma2(nm2, pack_f32x2(cg.z, cg.w), pack_f32x2(bg.z, bg.w)));
              } else {
                a01 = fadd2(a01, pack_f32x2(ba.x, ba.y));
                a23 = fadd2(a23, pack_f32x2(ba.z, ba.w));
                g01 = fadd2(g01, pack_f32x2(bg.x, bg.y));
                g23 = fadd2(g23, pack_f32x2(bg.z, bg.w));
              }
              float o0, o1, o2, o3;
              unpack_f32x2(geglu_poly_x2(a01, g01), o0, o1);
              unpack_f32x2(geglu_poly_x2(a23, g23), o2, o3);
              pk[2 * j] = cvt_bf16x2(o0, o1);
              pk[2 * j + 1] = cvt_bf16x2(o2, o3);
            }
            *reinterpret_cast<uint4*>(brow + ((static_cast<uint32_t>(qt) ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          }
        } else {
          uint32_t r[32];
          tmem_ld32(taddr + ch * 32, r);
          tmem_ld_wait();
          if (ch == ch_hi - 1 || n_acc0 + 32 >= p.N) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) release_acc(acc);
            released = true;
          }
          const uint64_t rstd2 = pack_f32x2(ln_rstd, ln_rstd), nm2 = pack_f32x2(ln_nm, ln_nm);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint64_t v01 = pack_f32x2(__uint_as_float(r[4 * j + 0]), __uint_as_float(r[4 * j + 1]));
            uint64_t v23 = pack_f32x2(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
            if constexpr (kLn) {
              const float4 b = lds128(pbias + 16 * j), cs = lds128(pcs + 16 * j);
              v01 = ffma2(rstd2, v01, ffma2(nm2, pack_f32x2(cs.x, cs.y), pack_f32x2(b.x, b.y)));
              v23 = ffma2(rstd2, v23, ffma2(nm2, pack_f32x2(cs.z, cs.w), pack_f32x2(b.z, b.w)));
            } else if (p.bias != nullptr) {
              const float4 b = lds128(pbias + 16 * j);
              v01 = fadd2(v01, pack_f32x2(b.x, b.y));
              v23 = fadd2(v23, pack_f32x2(b.z, b.w));
            }
            unpack_f32x2(v01, v[4 * j + 0], v[4 * j + 1]);
            unpack_f32x2(v23, v[4 * j + 2], v[4 * j + 3]);
          }
        }
        if (!GEGLU && tile_res) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const uint4 t = *reinterpret_cast<const uint4*>(brow + ((static_cast<uint32_t>(c) ^ sw) << 4));
            const float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y), cc = unpack_bf16x2(t.z), d = unpack_bf16x2(t.w);
            v[8 * c + 0] += a.x;
            v[8 * c + 1] += a.y;
            v[8 * c + 2] += b.x;
            v[8 * c + 3] += b.y;
            v[8 * c + 4] += cc.x;
            v[8 * c + 5] += cc.y;
            v[8 * c + 6] += d.x;
            v[8 * c + 7] += d.y;
          }
        }
        if constexpr (!GEGLU) {
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = cvt_bf16x2(v[2 * j], v[2 * j + 1]);
        }
#pragma unroll
        for (int c = 0; c < (GEGLU ? 0 : 4); ++c) {
          const uint4 pq = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
          *reinterpret_cast<uint4*>(brow + ((static_cast<uint32_t>(c) ^ sw) << 4)) = pq;
          if constexpr (!GEGLU && kStats) {
            // statistics of exactly what the consumer will read: the bf16-rounded values
            const float2 a = unpack_bf16x2(pq.x), b = unpack_bf16x2(pq.y), cc = unpack_bf16x2(pq.z), d = unpack_bf16x2(pq.w);
            st_s1 += ((a.x + a.y) + (b.x + b.y)) + ((cc.x + cc.y) + (d.x + d.y));
            st_s2 = fmaf(a.x, a.x, st_s2); st_s2 = fmaf(a.y, a.y, st_s2); st_s2 = fmaf(b.x, b.x, st_s2); st_s2 = fmaf(b.y, b.y, st_s2);
            st_s2 = fmaf(cc.x, cc.x, st_s2); st_s2 = fmaf(cc.y, cc.y, st_s2); st_s2 = fmaf(d.x, d.x, st_s2); st_s2 = fmaf(d.y, d.y, st_s2);
          }
        }
        fence_proxy_async_smem();
        AFB_GPROF(4)
        named_bar_sync(1 + grp, 128);
        if (leader) {
          if (p.conv_boxed != 0) {
            const int q = m_blk / p.conv_tw, bq = q / p.conv_th;
            const int w0 = (m_blk - q * p.conv_tw) * p.conv_bw, h0 = (q - bq * p.conv_th) * p.conv_bh, b0 = bq * p.conv_bb;
            if (p.conv_fold != 0) {
              // one store per image of the box: its bh low-res rows are consecutive in the merged (sample, row) dimension
              const int phase = n_blk / p.nb_per_phase;
              for (int ib = 0; ib < p.conv_bb; ++ib)
                tma_store_5d(&tm_out, staging + (sb * NC + ch) * kBoxBytes + ib * (p.conv_bh * p.conv_bw * 64),
                             n_out0 - phase * p.fold_cout, phase & 1, w0, phase >> 1, (b0 + ib) * p.conv_h + h0);
            } else if (p.k_splits > 1) {
              tma_reduce_add_4d(&tm_out, staging + (sb * NC + ch) * kBoxBytes, n_out0, w0, h0, b0);
            } else {
              tma_store_4d(&tm_out, staging + (sb * NC + ch) * kBoxBytes, n_out0, w0, h0, b0);
            }
          } else if (p.conv_fold != 0) {
            const int phase = n_blk / p.nb_per_phase;
            const int m0 = m_blk * kBM;
            const int r = m0 / p.conv_w;              // merged (sample, low-res row) index of the tile's first pixel
            tma_store_5d(&tm_out, staging + (sb * NC + ch) * kBoxBytes, n_out0 - phase * p.fold_cout, phase & 1,
                         m0 - r * p.conv_w, phase >> 1, r);
          } else if (p.k_splits > 1) {
            tma_reduce_add_2d(&tm_out, staging + (sb * NC + ch) * kBoxBytes, n_out0, m_blk * kBM);
          } else {
            tma_store_2d(&tm_out, staging + (sb * NC + ch) * kBoxBytes, n_out0, m_blk * kBM);
          }
        }
        AFB_GPROF(5)
      }
      if (!released) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(acc);
      }
      if (!GEGLU && kStats && m_row < p.M)
        reinterpret_cast<float2*>(p.stats_out)[static_cast<long long>(m_row) * (C::kGroups * p.num_n_blocks) + C::kGroups * n_blk + grp] =
            make_float2(st_s1, st_s2);
      if (leader) bulk_commit_group();
      if (++acc == kSlots) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (leader) bulk_wait_group_read<0>();      // shared memory must outlive the last stores' reads
    if (warp == 4) { AFB_GPROF_FLUSH(1) }
    if (warp == 7) { AFB_GPROF_FLUSH(2) }
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (kPair) cluster_sync_all();      // neither CTA may exit (or free TMEM) while the other can still signal it
  if (warp == 2) {
    if constexpr (kPair) tmem_dealloc_pair<C::kTmemCols>(tmem_base);
    else tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}

// Pixel box of a conv-mode row block.  Images whose rows tile into runs of 128 CONSECUTIVE pixels (W a power of two up to
// 128 or a multiple of 128, H likewise) keep the plain [M, N] view of the output; every other size (96, 48, 24, 12 ...)
// takes (bb, bh, bw) sub-boxes with bw / bh the largest powers of two dividing W / H.
struct ConvGeom {
  bool boxed;
  uint32_t bw, bh, bb, tw, th, tb;
};
inline ConvGeom conv_geom(int W, int H, long long batch) {
  ConvGeom g{};
  const int bw0 = W < kBM ? W : kBM, rl = kBM / bw0;
  if ((W < kBM ? kBM % W == 0 : W % kBM == 0) && (H < rl ? rl % H == 0 : H % rl == 0)) {
    g.boxed = false;
    g.bw = bw0;
    g.bh = H < rl ? H : rl;
  } else {
    g.boxed = true;
    g.bw = 1;
    while (g.bw < static_cast<uint32_t>(kBM) && W % (2 * g.bw) == 0) g.bw *= 2;
    g.bh = 1;
    while (g.bw * g.bh < static_cast<uint32_t>(kBM) && H % (2 * g.bh) == 0) g.bh *= 2;
  }
  g.bb = kBM / (g.bw * g.bh);
  g.tw = W / g.bw;
  g.th = H / g.bh;
  g.tb = static_cast<uint32_t>((batch + g.bb - 1) / g.bb);
  return g;
}

int make_box_map(CUtensorMap* map, const void* base, int64_t cols, int64_t rows, int64_t ld) {
  const uint64_t dims[2] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(rows)};
  const uint64_t strides[1] = {static_cast<uint64_t>(ld) * 2};
  const uint32_t box[2] = {32, kBM};
  return make_tensor_map(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, Swizzle::B64);
}

// AFB_GEMM_SPLITK=0 never splits K (A/B measurement); read once
bool splitk_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_GEMM_SPLITK");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

__global__ void __launch_bounds__(256) zero_bf16_kernel(uint4* __restrict__ p, long long n) {
  pdl_wait();
  pdl_trigger();
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

template <int BN, bool GEGLU, bool kLn, bool kStats, bool kPair, bool kWide = false>
int launch_tma_impl(const afb_gemm_args& a, cudaStream_t stream) {
  using C = TCfg<BN, GEGLU, kPair, kWide>;
  CUtensorMap tm_a, tm_w, tm_out, tm_res, tm_a2;
  const int64_t k1 = a.a2 != nullptr ? a.k_split : a.k;      // columns of A that come from a.a
  if (a.conv_h > 0) {
    // NHWC image [B, H, W, Cin], pixel stride lda; the box is 128 consecutive pixels in (b, h, w) order
    const uint64_t cin = static_cast<uint64_t>(a.k / (a.conv_fold != 0 ? 4 : 9)), W = static_cast<uint64_t>(a.conv_w),
                   H = static_cast<uint64_t>(a.conv_h);
    const ConvGeom cg = conv_geom(a.conv_w, a.conv_h, a.m / (static_cast<int64_t>(a.conv_w) * a.conv_h));
    const uint32_t bw = cg.bw, bh = cg.bh;
    // stride 2: W, H are the output size; the input image is twice as large and the box walks it with element stride 2
    const uint32_t cs = a.conv_stride == 2 ? 2u : 1u;
    const uint64_t Wi = W * cs, Hi = H * cs;
    const uint64_t dims[4] = {cin, Wi, Hi, static_cast<uint64_t>(a.m) / (W * H)};
    const uint64_t strides[3] = {static_cast<uint64_t>(a.lda) * 2, static_cast<uint64_t>(a.lda) * 2 * Wi,
                                 static_cast<uint64_t>(a.lda) * 2 * Wi * Hi};
    const uint32_t box[4] = {kBK, bw * cs, bh * cs, cg.bb};
    const uint32_t estr[4] = {1, cs, cs, 1};
    if (int rc = make_tensor_map(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.a, dims, strides, box, Swizzle::B128,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, estr))
      return rc;
  } else {
    const uint64_t dims[2] = {static_cast<uint64_t>(k1), static_cast<uint64_t>(a.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.lda) * 2};
    const uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tensor_map(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.a, dims, strides, box, Swizzle::B128))
      return rc;
  }
  tm_a2 = tm_a;
  if (a.a2 != nullptr) {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.k - k1), static_cast<uint64_t>(a.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.lda2) * 2};
    const uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tensor_map(&tm_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.a2, dims, strides, box, Swizzle::B128))
      return rc;
  }
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.k), static_cast<uint64_t>(a.n)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.ldw) * 2};
    const uint32_t box[2] = {kBK, static_cast<uint32_t>(C::kBRows)};
    if (int rc = make_tensor_map(&tm_w, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.w, dims, strides, box, Swizzle::B128))
      return rc;
  }
  const int64_t n_out = GEGLU ? a.n / 2 : a.n;
  if (a.conv_fold != 0) {
    // high-resolution NHWC output [B, 2H, 2W, Cout] seen as [B*H][py][W][px][Cout]; a tile stores its 128 low-res pixels
    // of one phase: box {32 channels, 1, bw, 1, bh*bb}
    const uint64_t W = static_cast<uint64_t>(a.conv_w), H = static_cast<uint64_t>(a.conv_h), cout = static_cast<uint64_t>(a.n / 4);
    const uint64_t ld = static_cast<uint64_t>(a.ldo) * 2;
    const ConvGeom cg = conv_geom(a.conv_w, a.conv_h, a.m / (static_cast<int64_t>(a.conv_w) * a.conv_h));
    const uint64_t dims[5] = {cout, 2, W, 2, static_cast<uint64_t>(a.m) / W};
    const uint64_t strides[4] = {ld, 2 * ld, 2 * W * ld, 4 * W * ld};
    // (boxed: one store per image of the box, bh consecutive rows of the merged (sample, low-res row) dimension each)
    const uint32_t box[5] = {32, 1, cg.bw, 1, cg.boxed ? cg.bh : kBM / cg.bw};
    (void)H;
    if (int rc = make_tensor_map(&tm_out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, a.out, dims, strides, box, Swizzle::B64)) return rc;
  } else if (a.conv_h > 0 && conv_geom(a.conv_w, a.conv_h, 1).boxed) {
    // sub-box tiles: the output (and the residual) as [B, H, W, N] behind a 4-D map, box {32 channels, bw, bh, bb}
    const ConvGeom cg = conv_geom(a.conv_w, a.conv_h, a.m / (static_cast<int64_t>(a.conv_w) * a.conv_h));
    const uint64_t W = static_cast<uint64_t>(a.conv_w), H = static_cast<uint64_t>(a.conv_h);
    const uint32_t box[4] = {32, cg.bw, cg.bh, cg.bb};
    auto map4 = [&](CUtensorMap* map, const void* base, int64_t ld) {
      const uint64_t dims[4] = {static_cast<uint64_t>(n_out), W, H, static_cast<uint64_t>(a.m) / (W * H)};
      const uint64_t strides[3] = {static_cast<uint64_t>(ld) * 2, static_cast<uint64_t>(ld) * 2 * W, static_cast<uint64_t>(ld) * 2 * W * H};
      return make_tensor_map(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box, Swizzle::B64);
    };
    if (int rc = map4(&tm_out, a.out, a.ldo)) return rc;
    if (a.residual != nullptr) {
      if (int rc = map4(&tm_res, a.residual, a.ldr)) return rc;
    }
  } else if (int rc = make_box_map(&tm_out, a.out, n_out, a.m, a.ldo)) {
    return rc;
  }
  const bool boxed = a.conv_h > 0 && conv_geom(a.conv_w, a.conv_h, 1).boxed;
  if (a.residual != nullptr && !boxed) {
    if (int rc = make_box_map(&tm_res, a.residual, n_out, a.m, a.ldr)) return rc;
  } else if (a.residual == nullptr) {
    tm_res = tm_out;
  }
  GemmTmaDev p;
  p.bias = a.bias;
  p.M = static_cast<int>(a.m);
  p.N = static_cast<int>(a.n);
  p.K = static_cast<int>(a.k);
  p.has_res = a.residual != nullptr;
  p.num_m_blocks = static_cast<int>((a.m + kBM - 1) / kBM);
  p.num_n_blocks = static_cast<int>((a.n + BN - 1) / BN);
  p.num_k_blocks = static_cast<int>((a.k + kBK - 1) / kBK);
  p.stats_out = a.row_stats_out;
  p.ln_stats = a.ln_stats;
  p.ln_colsum = a.ln_colsum;
  p.ln_parts = a.ln_parts;
  p.ln_eps = a.ln_eps;
  const int taps = a.conv_fold != 0 ? 4 : 9;
  p.conv_cblocks = a.conv_h > 0 ? static_cast<int>(a.k / taps / kBK) : 0;
  p.conv_fold = a.conv_fold;
  p.k1_blocks = a.a2 != nullptr ? static_cast<int>(k1 / kBK) : p.num_k_blocks;
  p.conv_boxed = 0;
  p.conv_bw = p.conv_bh = p.conv_bb = p.conv_tw = p.conv_th = 1;
  if (a.conv_h > 0) {
    const ConvGeom cg = conv_geom(a.conv_w, a.conv_h, a.m / (static_cast<int64_t>(a.conv_w) * a.conv_h));
    p.conv_boxed = cg.boxed ? 1 : 0;
    p.conv_bw = static_cast<int>(cg.bw);
    p.conv_bh = static_cast<int>(cg.bh);
    p.conv_bb = static_cast<int>(cg.bb);
    p.conv_tw = static_cast<int>(cg.tw);
    p.conv_th = static_cast<int>(cg.th);
    if (cg.boxed) p.num_m_blocks = static_cast<int>(cg.tw * cg.th * cg.tb);
  }
  p.k_splits = 1;
  if constexpr (kPair && !GEGLU && !kLn && !kStats) {
    // fewer tiles than SM pairs and a deep K: split K so that every pair has work (measured on the 3x3 convolutions at
    // 8x8, 32 tiles of 180-360 k-blocks on 74 pairs).  The output is zeroed first; partial tiles meet by bf16 reduce-adds.
    const int tiles = ((p.num_m_blocks + 1) / 2) * (kWide ? p.num_n_blocks / 2 : p.num_n_blocks);
    const int pairs = num_sms() / 2;
    if (splitk_enabled() && a.conv_fold == 0 && a.a2 == nullptr && a.residual != a.out && 2 * tiles <= pairs + pairs / 4 && p.num_k_blocks >= 64 &&
        a.ldo == a.n) {
      int ks = (pairs + pairs / 4) / tiles;
      if (ks > 4) ks = 4;
      while (ks > 1 && p.num_k_blocks % ks != 0) --ks;
      if (ks > 1) {
        p.k_splits = ks;
        AFB_CHECK_CUDA(launch_pdl(zero_bf16_kernel, dim3(static_cast<unsigned>((a.m * a.n / 8 + 255) / 256)), dim3(256), 0, stream, 1,
                                  reinterpret_cast<uint4*>(a.out), static_cast<long long>(a.m * a.n / 8)));
        count_launch();
      }
    }
  }
  p.fold_cout = static_cast<int>(a.n / 4);
  p.nb_per_phase = a.conv_fold != 0 ? static_cast<int>(a.n / 4 / BN) : 1;
  p.conv_h = a.conv_h;
  p.conv_w = a.conv_w;
  p.conv_stride = a.conv_stride == 2 ? 2 : 1;
  auto kern = gemm_bf16_tma_kernel<BN, GEGLU, kLn, kStats, kPair, kWide>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  int grid;
  if constexpr (kPair) {
    const int tiles = ((p.num_m_blocks + 1) / 2) * (kWide ? p.num_n_blocks / 2 : p.num_n_blocks) * p.k_splits;
    const int pairs = tiles < num_sms() / 2 ? tiles : num_sms() / 2;
    grid = 2 * pairs;
  } else {
    const int tiles = p.num_m_blocks * p.num_n_blocks;
    grid = tiles < num_sms() ? tiles : num_sms();
  }
  AFB_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(C::kThreads), C::kSmemBytes, stream, kPair ? 2 : 1, tm_a, tm_w, tm_out, tm_res, tm_a2, p));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// AFB_GEMM_PAIR=0 keeps every GEMM on single-CTA tiles (A/B measurement); read once
bool pair_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_GEMM_PAIR");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

// AFB_GEMM_WIDE=0 keeps the BN-wide tiles everywhere (A/B measurement); read once
bool wide_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_GEMM_WIDE");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

template <int BN, bool GEGLU, bool kPair, bool kWide = false>
int launch_tma_mode(const afb_gemm_args& a, cudaStream_t stream) {
  const bool ln = a.ln_stats != nullptr, st = a.row_stats_out != nullptr;
  if constexpr (GEGLU) {
    return ln ? launch_tma_impl<BN, true, true, false, kPair>(a, stream)
              : launch_tma_impl<BN, true, false, false, kPair>(a, stream);
  } else {
    if (ln && st) AFB_FAIL(2, "afb_gemm_bf16: folded LayerNorm and row statistics in one GEMM are not built");
    if (ln) return launch_tma_impl<BN, false, true, false, kPair, kWide>(a, stream);
    if (st) return launch_tma_impl<BN, false, false, true, kPair, kWide>(a, stream);
    return launch_tma_impl<BN, false, false, false, kPair, kWide>(a, stream);
  }
}

template <int BN, bool GEGLU>
int launch_tma(const afb_gemm_args& a, cudaStream_t stream) {
  // a CTA pair needs two row blocks to share one W tile
  if (a.m > kBM && pair_enabled()) {
    // 2 * BN-wide tiles when the columns pair up and there are enough tiles left to fill the pairs of SMs
    if constexpr (!GEGLU && BN == 160) {
      // (measured per shape, B200: K >= 1280 gains 2-10 % -- the 3x3 convolutions, FF-out, the C = 1280 projections --,
      //  K = 640 is even, K = 320 loses 10-15 %: its [65536, 320] activation streams from HBM and the three-stage ring of
      //  the wide tile is too shallow for that)
      static const int wide_min_k = [] { const char* e = getenv("AFB_GEMM_WIDE_MIN_K"); return e != nullptr ? atoi(e) : 1280; }();
      if (a.n % (2 * BN) == 0 && a.conv_fold == 0 && a.k >= wide_min_k && wide_enabled()) {
        // waves x operand KB per CTA and k-block: 16 + 2 * 10 for a wide tile, 16 + 10 for a BN-wide one
        const int64_t pairs = num_sms() / 2, m_units = (a.m + 2 * kBM - 1) / (2 * kBM);
        const int64_t wide = (m_units * (a.n / (2 * BN)) + pairs - 1) / pairs * 36, narrow = (m_units * (a.n / BN) + pairs - 1) / pairs * 26;
        if (wide <= narrow) return launch_tma_mode<BN, GEGLU, true, true>(a, stream);
      }
    }
    return launch_tma_mode<BN, GEGLU, true>(a, stream);
  }
  return launch_tma_mode<BN, GEGLU, false>(a, stream);
}

template <int BN, int EPI>
int launch(const afb_gemm_args& a, cudaStream_t stream) {
  using C = Cfg<BN>;
  CUtensorMap tm_a, tm_w;
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.k), static_cast<uint64_t>(a.m)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.lda) * 2};
    const uint32_t box[2] = {kBK, kBM};
    if (int rc = make_tensor_map(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.a, dims, strides, box, Swizzle::B128))
      return rc;
  }
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.k), static_cast<uint64_t>(a.n)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.ldw) * 2};
    const uint32_t box[2] = {kBK, BN};
    if (int rc = make_tensor_map(&tm_w, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.w, dims, strides, box, Swizzle::B128))
      return rc;
  }
  GemmDev p;
  p.out = a.out;
  p.bias = a.bias;
  p.residual = a.residual;
  p.M = static_cast<int>(a.m);
  p.N = static_cast<int>(a.n);
  p.K = static_cast<int>(a.k);
  p.ldo = a.ldo;
  p.ldr = a.ldr;
  p.out_f32 = a.out_dtype == AFB_F32;
  p.res_f32 = a.residual_dtype == AFB_F32;
  p.hw = a.hw;
  p.num_m_blocks = static_cast<int>((a.m + kBM - 1) / kBM);
  p.num_n_blocks = static_cast<int>((a.n + BN - 1) / BN);
  p.num_k_blocks = static_cast<int>((a.k + kBK - 1) / kBK);
  auto kern = gemm_bf16_kernel<BN, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  const int tiles = p.num_m_blocks * p.num_n_blocks;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  AFB_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), C::kSmemBytes, stream, 1, tm_a, tm_w, p));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int EPI>
int dispatch_bn(const afb_gemm_args& a, cudaStream_t stream) {
  if (a.n % 160 == 0) return launch<160, EPI>(a, stream);
  return launch<128, EPI>(a, stream);
}

}  // namespace
}  // namespace afb

extern "C" int afb_gemm_bf16(const afb_gemm_args* args, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(args != nullptr, "afb_gemm_bf16: args is NULL");
  const afb_gemm_args& a = *args;
  AFB_REQUIRE(a.a && a.w && a.out, "afb_gemm_bf16: NULL operand");
  AFB_REQUIRE(a.m > 0 && a.n > 0 && a.k > 0, "afb_gemm_bf16: empty problem m=%lld n=%lld k=%lld", (long long)a.m,
              (long long)a.n, (long long)a.k);
  AFB_REQUIRE(a.m < (1ll << 31) && a.n < (1ll << 31) && a.k < (1ll << 31), "afb_gemm_bf16: dimension too large");
  AFB_REQUIRE(a.lda % 8 == 0 && a.ldw % 8 == 0, "afb_gemm_bf16: lda/ldw must be multiples of 8 elements (TMA 16B strides)");
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(a.a) & 15) == 0 && (reinterpret_cast<uintptr_t>(a.w) & 15) == 0,
              "afb_gemm_bf16: operands must be 16-byte aligned");
  if (a.conv_h > 0 || a.conv_w > 0) {
    // 3x3 / stride 1 / zero padding 1 convolution as an implicit GEMM (openaimodel.py:203-205,236-238)
    const int64_t hw = static_cast<int64_t>(a.conv_h) * a.conv_w;
    AFB_REQUIRE(a.conv_h > 0 && a.conv_w > 0 && a.m % hw == 0, "afb_gemm_bf16: conv needs conv_h, conv_w > 0 and M = B*H*W");
    const int taps = a.conv_fold != 0 ? 4 : 9;
    AFB_REQUIRE(a.conv_stride == 0 || a.conv_stride == 1 || (a.conv_stride == 2 && a.conv_fold == 0),
                "afb_gemm_bf16: conv_stride must be 1 or 2 (2 not together with conv_fold)");
    AFB_REQUIRE(a.k % (taps * 64) == 0, "afb_gemm_bf16: conv needs K = %d * Cin with Cin %% 64 == 0 (K=%lld)", taps, (long long)a.k);
    AFB_REQUIRE(a.lda >= a.k / taps && a.ldw >= a.k, "afb_gemm_bf16: conv leading dimension smaller than Cin / K");
    if (a.conv_fold != 0) {
      const int64_t cout = a.n / 4;
      AFB_REQUIRE(a.n % 4 == 0 && cout % (cout % 160 == 0 ? 160 : 128) == 0 && a.residual == nullptr && a.ldo == cout,
                  "afb_gemm_bf16: folded upsampling needs N = 4 * Cout with Cout a multiple of the tile width (160 or 128), "
                  "no residual and a contiguous [B, 2H, 2W, Cout] output");
    }
    // any image size tiles into 128-pixel boxes (conv_geom); widths / heights with a small power-of-two factor just get
    // narrow boxes: require at least 4-pixel runs so that a TMA box row is not a single pixel
    AFB_REQUIRE(a.conv_w % 4 == 0 || kBM % a.conv_w == 0, "afb_gemm_bf16: conv width %d must be a multiple of 4 (or divide 128)", a.conv_w);
    AFB_REQUIRE(a.epilogue == AFB_EPI_STORE && a.out_dtype == AFB_BF16 && (a.residual == nullptr || a.residual_dtype == AFB_BF16) &&
                    a.ln_stats == nullptr && a.row_stats_out == nullptr,
                "afb_gemm_bf16: conv mode takes the bf16 STORE epilogue (bias, bf16 residual) only");
  } else if (a.a2 != nullptr) {
    // A = [a | a2] along K: the decoder's 1x1 skip convolution of th.cat([h, hs.pop()], 1) (openaimodel.py:1004-1010, 277)
    AFB_REQUIRE(a.k_split > 0 && a.k_split < a.k && a.k_split % 64 == 0, "afb_gemm_bf16: k_split must be a multiple of 64 inside (0, K)");
    AFB_REQUIRE(a.lda >= a.k_split && a.lda2 >= a.k - a.k_split && a.lda2 % 8 == 0 && a.ldw >= a.k &&
                    (reinterpret_cast<uintptr_t>(a.a2) & 15) == 0,
                "afb_gemm_bf16: a2 must be 16-byte aligned with lda2 %% 8 == 0 and the leading dimensions must cover their halves");
    AFB_REQUIRE(a.epilogue == AFB_EPI_STORE && a.out_dtype == AFB_BF16 && (a.residual == nullptr || a.residual_dtype == AFB_BF16),
                "afb_gemm_bf16: the two-source A operand takes the bf16 STORE epilogue only");
  } else {
    AFB_REQUIRE(a.lda >= a.k && a.ldw >= a.k, "afb_gemm_bf16: leading dimension smaller than K");
  }
  AFB_REQUIRE(a.n % 32 == 0, "afb_gemm_bf16: N=%lld must be a multiple of 32", (long long)a.n);
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(a.out) & 15) == 0, "afb_gemm_bf16: out must be 16-byte aligned");
  AFB_REQUIRE(a.bias == nullptr || (reinterpret_cast<uintptr_t>(a.bias) & 15) == 0, "afb_gemm_bf16: bias must be 16-byte aligned");
  if (a.ln_stats != nullptr) {
    AFB_REQUIRE(a.ln_colsum != nullptr && a.ln_parts > 0 && a.ln_parts <= 64, "afb_gemm_bf16: folded LayerNorm needs ln_colsum and 1..64 ln_parts");
    AFB_REQUIRE(a.ln_parts % 2 == 0 && (reinterpret_cast<uintptr_t>(a.ln_stats) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(a.ln_colsum) & 15) == 0,
                "afb_gemm_bf16: ln_parts must be even, ln_stats and ln_colsum 16-byte aligned");
  }
  AFB_REQUIRE(a.row_stats_out == nullptr || (reinterpret_cast<uintptr_t>(a.row_stats_out) & 7) == 0,
              "afb_gemm_bf16: row_stats_out must be 8-byte aligned");
  switch (a.epilogue) {
    case AFB_EPI_STORE:
      AFB_REQUIRE((a.conv_fold != 0 || a.ldo >= a.n) && a.ldo % 8 == 0, "afb_gemm_bf16: ldo must be >= N and a multiple of 8");
      if (a.residual) {
        AFB_REQUIRE(a.ldr >= a.n && a.ldr % 8 == 0, "afb_gemm_bf16: ldr must be >= N and a multiple of 8");
        AFB_REQUIRE((reinterpret_cast<uintptr_t>(a.residual) & 15) == 0, "afb_gemm_bf16: residual must be 16-byte aligned");
      }
      if (a.out_dtype == AFB_BF16 && (a.residual == nullptr || a.residual_dtype == AFB_BF16)) {
        if (a.conv_fold != 0 ? (a.n / 4) % 160 == 0 : a.n % 160 == 0) return launch_tma<160, false>(a, stream);
        return launch_tma<128, false>(a, stream);
      }
      AFB_REQUIRE(a.row_stats_out == nullptr && a.ln_stats == nullptr,
                  "afb_gemm_bf16: row statistics / folded LayerNorm need bf16 output and a bf16 (or no) residual");
      return dispatch_bn<AFB_EPI_STORE>(a, stream);
    case AFB_EPI_GEGLU:
      AFB_REQUIRE(a.ldo >= a.n / 2 && a.ldo % 8 == 0, "afb_gemm_bf16: GEGLU ldo must be >= N/2 and a multiple of 8");
      AFB_REQUIRE(a.residual == nullptr, "afb_gemm_bf16: GEGLU epilogue takes no residual");
      AFB_REQUIRE(a.n % 64 == 0 && a.bias != nullptr, "afb_gemm_bf16: GEGLU needs N %% 64 == 0 and a bias");
      AFB_REQUIRE(a.row_stats_out == nullptr, "afb_gemm_bf16: the GEGLU epilogue does not emit row statistics");
      // pair mode takes 256-wide tiles: per SM and k-block 32 KB staged for 128 x 256 outputs instead of 32 KB for
      // 128 x 128, which is what the shared-memory port (TMA fill + UMMA operand reads) can sustain
      if (a.m > kBM && a.n % 256 == 0 && pair_enabled()) return launch_tma_mode<256, true, true>(a, stream);
      return launch_tma<128, true>(a, stream);
    case AFB_EPI_NCHW:
      AFB_REQUIRE(a.hw > 0 && a.m % a.hw == 0, "afb_gemm_bf16: NCHW epilogue needs hw > 0 dividing M");
      AFB_REQUIRE(a.row_stats_out == nullptr && a.ln_stats == nullptr,
                  "afb_gemm_bf16: row statistics / folded LayerNorm are not available with the NCHW epilogue");
      return dispatch_bn<AFB_EPI_NCHW>(a, stream);
    default:
      AFB_FAIL(2, "afb_gemm_bf16: unknown epilogue %d", a.epilogue);
  }
}

extern "C" int afb_gemm_stats_parts(int64_t n) {
  const int64_t bn = (n % 160 == 0) ? 160 : 128;      // the tile width afb_gemm_bf16 picks for a STORE epilogue
  return static_cast<int>((AFB_GEMM_EPI_WARPS / 4) * ((n + bn - 1) / bn));   // one partial per epilogue column group
}
