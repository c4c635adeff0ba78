// Persistent, warp-specialised BF16 GEMM for sm_100a:  Y = epilogue(A[M,K] * W[N,K]^T).
//
//   warp 0 (one lane)  TMA producer: A tile 128x64 and W tile BNx64 (both K-major, SWIZZLE_128B) into a
//                      kStages-deep shared-memory ring, mbarrier complete_tx signalling
//   warp 1 (one lane)  tcgen05.mma issuer: 128 x BN x 16 UMMAs, fp32 accumulators in TMEM, two
//                      accumulator buffers so the epilogue of tile i overlaps the main loop of tile i+1
//   warp 2             TMEM allocator
//   warps 4..7         epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> bias / residual / GEGLU /
//                      NCHW transposition -> global
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
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, 0, 0);
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
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint64_t da = umma_desc_kmajor_sw128(sa);
          const uint64_t db = umma_desc_kmajor_sw128(sa + C::kABytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128B swizzle atom: +2 in the >>4 address field
            umma_ss(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == C::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[acc]);
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
        } else if constexpr (EPI == AFB_EPI_GEGLU) {
          // columns [0,16) of the chunk are the value rows, [16,32) the gate rows of the same 16 outputs
          if (row_ok) {
            float o[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) o[j] = v[j] * gelu_erf(v[16 + j]);
            uint4* o4 = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + static_cast<long long>(m) * p.ldo + (n0 >> 1));
#pragma unroll
            for (int j = 0; j < 2; ++j)
              o4[j] = make_uint4(pack_bf16x2(o[8 * j], o[8 * j + 1]), pack_bf16x2(o[8 * j + 2], o[8 * j + 3]),
                                 pack_bf16x2(o[8 * j + 4], o[8 * j + 5]), pack_bf16x2(o[8 * j + 6], o[8 * j + 7]));
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
  kern<<<grid, kThreads, C::kSmemBytes, stream>>>(tm_a, tm_w, p);
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
  AFB_REQUIRE(a.lda >= a.k && a.ldw >= a.k, "afb_gemm_bf16: leading dimension smaller than K");
  AFB_REQUIRE(a.n % 32 == 0, "afb_gemm_bf16: N=%lld must be a multiple of 32", (long long)a.n);
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(a.out) & 15) == 0, "afb_gemm_bf16: out must be 16-byte aligned");
  AFB_REQUIRE(a.bias == nullptr || (reinterpret_cast<uintptr_t>(a.bias) & 15) == 0, "afb_gemm_bf16: bias must be 16-byte aligned");
  switch (a.epilogue) {
    case AFB_EPI_STORE:
      AFB_REQUIRE(a.ldo >= a.n && a.ldo % 8 == 0, "afb_gemm_bf16: ldo must be >= N and a multiple of 8");
      if (a.residual) {
        AFB_REQUIRE(a.ldr >= a.n && a.ldr % 8 == 0, "afb_gemm_bf16: ldr must be >= N and a multiple of 8");
        AFB_REQUIRE((reinterpret_cast<uintptr_t>(a.residual) & 15) == 0, "afb_gemm_bf16: residual must be 16-byte aligned");
      }
      return dispatch_bn<AFB_EPI_STORE>(a, stream);
    case AFB_EPI_GEGLU:
      AFB_REQUIRE(a.ldo >= a.n / 2 && a.ldo % 8 == 0, "afb_gemm_bf16: GEGLU ldo must be >= N/2 and a multiple of 8");
      AFB_REQUIRE(a.residual == nullptr, "afb_gemm_bf16: GEGLU epilogue takes no residual");
      return dispatch_bn<AFB_EPI_GEGLU>(a, stream);
    case AFB_EPI_NCHW:
      AFB_REQUIRE(a.hw > 0 && a.m % a.hw == 0, "afb_gemm_bf16: NCHW epilogue needs hw > 0 dividing M");
      return dispatch_bn<AFB_EPI_NCHW>(a, stream);
    default:
      AFB_FAIL(2, "afb_gemm_bf16: unknown epilogue %d", a.epilogue);
  }
}
