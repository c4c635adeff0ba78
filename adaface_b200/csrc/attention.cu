// Fused attention forward for sm_100a:  O = softmax(Q K^T * scale [+ key mask]) V, online softmax.
//
// One CTA per (128-query tile, head, sample); 192 threads:
//   warp 0 (one lane)  TMA producer: Q once, then K/V tiles of kBlockKV keys into a kStages ring
//                      (64-column boxes, SWIZZLE_128B; the head is a column slice of the [B, N, ld] buffer)
//   warp 1 (one lane)  tcgen05.mma issuer: S = Q K^T (128 x kBlockKV, fp32 in TMEM), then O += P V
//                      (128 x d, fp32 in TMEM; V is the MN-major B operand straight from its [key][d] layout)
//   warps 2..5         softmax: one query row per thread. tcgen05.ld S -> registers, scale/mask, running max
//                      with lazy rescaling of O in TMEM (only when the max grows by > 2^8), exp2, row sum,
//                      P -> bf16 -> shared memory in the canonical K-major 128B-swizzled layout; final
//                      O / l -> bf16 -> global.
// Two CTAs are resident per SM for head_dim 40/80 (112 KB smem, 256 TMEM columns each), so one CTA's
// softmax overlaps the other's MMAs.
//
// Replaces attention.py:198-242 (einsum QK^T, masked_fill, softmax, einsum PV, head merge) without ever
// materialising the [8B, N, N] score tensor.
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

constexpr int kBlockQ = 128;
constexpr int kAttnThreads = 192;
constexpr float kMaskedScore = -1.0e30f;   // finite: a fully masked row degrades to a uniform average,
                                           // like masked_fill(-finfo.max) + softmax in the reference
constexpr float kRescaleThreshold = 8.0f;  // log2 units

struct AttnDev {
  __nv_bfloat16* out;
  const uint8_t* key_mask;
  int q_len, kv_len, heads;
  long long ldo;
  float scale_log2;
};

template <int D>
struct AttnCfg {
  static constexpr int kAtoms = (D + 63) / 64;                 // 64-column (128-byte) swizzle atoms per row
  static constexpr int kBlockKV = (D <= 40) ? 128 : 64;
  static constexpr int kStages = 2;
  static constexpr int kDPad = (D + 15) / 16 * 16;             // UMMA N / K granularity
  static constexpr int kQBytes = kAtoms * kBlockQ * 128;
  static constexpr int kKVTileBytes = kAtoms * kBlockKV * 128;  // one of K or V
  static constexpr int kPAtoms = kBlockKV / 64;
  static constexpr int kPBytes = kPAtoms * kBlockQ * 128;
  static constexpr int kSmemBytes = kQBytes + kStages * 2 * kKVTileBytes + kPBytes;   // barriers are static smem
  static constexpr int kTmemCols = 256;
  static constexpr int kTmemS = 0;
  static constexpr int kTmemO = kBlockKV;                       // O accumulator starts after S
  // 228 KB per SM, 1 KB reserved per CTA, dynamic smem starts at a 1 KB boundary after the static barriers
  static constexpr int kMinBlocks = (kSmemBytes <= 112 * 1024) ? 2 : 1;
  static_assert(kBlockKV + kDPad <= kTmemCols, "TMEM budget");
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int D, bool kHasMask>
__global__ void __launch_bounds__(kAttnThreads, AttnCfg<D>::kMinBlocks)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                     const __grid_constant__ CUtensorMap tm_v, const AttnDev p) {
  using C = AttnCfg<D>;
  constexpr int BKV = C::kBlockKV;
  extern __shared__ __align__(1024) uint8_t smem[];     // SWIZZLE_128B atoms need 1024-byte alignment
  __shared__ __align__(8) uint64_t bars[5 + 2 * C::kStages + 1];
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + C::kQBytes;                       // stage s: K at sKV + s*2*tile, V right after
  uint8_t* sP = sKV + C::kStages * 2 * C::kKVTileBytes;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* q_full = bars + 0;
  uint64_t* q_ready = bars + 1;
  uint64_t* s_full = bars + 2;
  uint64_t* p_ready = bars + 3;
  uint64_t* pv_done = bars + 4;
  uint64_t* kv_full = bars + 5;                          // [kStages]
  uint64_t* kv_empty = kv_full + C::kStages;             // [kStages]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(kv_empty + C::kStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q_tile = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int num_kv_tiles = (p.kv_len + BKV - 1) / BKV;
  const int col0 = head * D;                             // first column of this head in the row

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(q_full, 1);
    mbar_init(q_ready, 128);
    mbar_init(s_full, 1);
    mbar_init(p_ready, 128);
    mbar_init(pv_done, 1);
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<C::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, C::kQBytes);
#pragma unroll
      for (int a = 0; a < C::kAtoms; ++a)
        tma_load_3d(&tm_q, q_full, sQ + a * kBlockQ * 128, col0 + a * 64, q_tile * kBlockQ, b, kEvictFirst);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < num_kv_tiles; ++j) {
        mbar_wait(&kv_empty[stage], phase ^ 1);
        uint8_t* sk = sKV + stage * 2 * C::kKVTileBytes;
        uint8_t* sv = sk + C::kKVTileBytes;
        mbar_arrive_expect_tx(&kv_full[stage], 2 * C::kKVTileBytes);
#pragma unroll
        for (int a = 0; a < C::kAtoms; ++a) {
          tma_load_3d(&tm_k, &kv_full[stage], sk + a * BKV * 128, col0 + a * 64, j * BKV, b, kEvictLast);
          tma_load_3d(&tm_v, &kv_full[stage], sv + a * BKV * 128, col0 + a * 64, j * BKV, b, kEvictLast);
        }
        if (++stage == C::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(kBlockQ, BKV, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(kBlockQ, C::kDPad, 0, 1);   // B (= V) is MN-major
      const uint32_t tmem_s = tmem_base + C::kTmemS;
      const uint32_t tmem_o = tmem_base + C::kTmemO;
      const uint32_t q_addr = smem_u32(sQ);
      const uint32_t p_addr = smem_u32(sP);
      mbar_wait(q_full, 0);
      mbar_wait(q_ready, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int j = 0; j < num_kv_tiles; ++j) {
        mbar_wait(&kv_full[stage], phase);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sKV + stage * 2 * C::kKVTileBytes);
        const uint32_t v_addr = k_addr + C::kKVTileBytes;
        // S = Q K^T : K-steps of 16 over the (padded) head dim
#pragma unroll
        for (int kk = 0; kk < C::kDPad / 16; ++kk) {
          const uint32_t qa = q_addr + (kk >> 2) * (kBlockQ * 128) + (kk & 3) * 32;
          const uint32_t ka = k_addr + (kk >> 2) * (BKV * 128) + (kk & 3) * 32;
          umma_ss(tmem_s, umma_desc_kmajor_sw128(qa), umma_desc_kmajor_sw128(ka), idesc_qk, kk != 0);
        }
        umma_commit(s_full);
        // wait for P_j (and the O rescale) from the softmax warps
        mbar_wait(p_ready, j & 1);
        tc_fence_after();
        // O += P V : K-steps of 16 keys; V tile is [key][64 cols] per atom -> MN-major, LBO = atom stride
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
          const uint32_t pa = p_addr + (kk >> 2) * (kBlockQ * 128) + (kk & 3) * 32;
          const uint32_t va = v_addr + kk * (16 * 128);
          umma_ss(tmem_o, umma_desc_kmajor_sw128(pa), umma_desc_mnmajor_sw128(va, BKV * 128), idesc_pv,
                  (j | kk) != 0);
        }
        umma_commit(&kv_empty[stage]);
        umma_commit(pv_done);
        if (++stage == C::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax / correction / epilogue
    const int quarter = warp & 3;                         // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;                  // query row inside the tile
    const uint32_t lane_addr = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tmem_s = tmem_base + lane_addr + C::kTmemS;
    const uint32_t tmem_o = tmem_base + lane_addr + C::kTmemO;

    mbar_wait(q_full, 0);
    if constexpr (D % 16 != 0) {
      // head_dim 40: K-step 2 covers columns 32..47; zero this row's columns 40..47 of Q (one 16-byte chunk,
      // swizzled position) so that whatever K holds there (the next head's data) contributes nothing.
      constexpr int chunk = D / 8;                        // 16-byte chunk index inside the 128-byte row
      uint8_t* dst = sQ + (C::kAtoms - 1) * kBlockQ * 128 + row * 128 + (((chunk & 7) ^ (row & 7)) << 4);
      *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
      fence_proxy_async_smem();
    }
    mbar_arrive(q_ready);

    float m_used = 0.f;      // max (log2 domain) the exponentials are currently referenced to
    float l_run = 0.f;       // running row sum
    const uint8_t* mask_row = kHasMask ? p.key_mask + static_cast<long long>(b) * p.kv_len : nullptr;

    for (int j = 0; j < num_kv_tiles; ++j) {
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      float s[BKV];
#pragma unroll
      for (int c = 0; c < BKV / 32; ++c) {
        uint32_t r[32];
        tmem_ld32(tmem_s + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(r[i]) * p.scale_log2;
      }
      const int kv0 = j * BKV;
      if (kHasMask) {
#pragma unroll
        for (int i = 0; i < BKV; ++i)
          if (kv0 + i < p.kv_len && mask_row[kv0 + i] == 0) s[i] = kMaskedScore;
      }
      if (kv0 + BKV > p.kv_len) {
#pragma unroll
        for (int i = 0; i < BKV; ++i)
          if (kv0 + i >= p.kv_len) s[i] = -INFINITY;
      }
      float m_tile = s[0];
#pragma unroll
      for (int i = 1; i < BKV; ++i) m_tile = fmaxf(m_tile, s[i]);

      // P buffer and O are free once PV(j-1) has completed
      if (j > 0) {
        mbar_wait(pv_done, (j - 1) & 1);
        tc_fence_after();
      }
      bool need = (j == 0) || (m_tile > m_used + kRescaleThreshold);
      if (j == 0) {
        m_used = m_tile;
      } else if (__any_sync(0xffffffffu, need)) {
        const float m_new = need ? m_tile : m_used;
        const float alpha = fast_exp2(m_used - m_new);   // 1 for rows that keep their reference max
        m_used = m_new;
        l_run *= alpha;
#pragma unroll
        for (int c = 0; c < C::kDPad / 16; ++c) {
          uint32_t r[16];
          tmem_ld16(tmem_o + c * 16, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st16(tmem_o + c * 16, r);
        }
        tmem_st_wait();
      }
      float l_tile = 0.f;
#pragma unroll
      for (int i = 0; i < BKV; ++i) {
        s[i] = fast_exp2(s[i] - m_used);
        l_tile += s[i];
      }
      l_run += l_tile;
      // P -> bf16, canonical K-major SW128 layout: atom a = 64 keys, row pitch 128 B, 16-byte chunk c at c ^ (row & 7)
#pragma unroll
      for (int a = 0; a < C::kPAtoms; ++a) {
        uint8_t* prow = sP + a * (kBlockQ * 128) + row * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float* sv = &s[a * 64 + c * 8];
          const uint4 pk = make_uint4(pack_bf16x2(sv[0], sv[1]), pack_bf16x2(sv[2], sv[3]),
                                      pack_bf16x2(sv[4], sv[5]), pack_bf16x2(sv[6], sv[7]));
          *reinterpret_cast<uint4*>(prow + ((c ^ (row & 7)) << 4)) = pk;
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_ready);
    }

    // ---- epilogue: O / l -> bf16 -> out[b, q, head*D : head*D + D]
    mbar_wait(pv_done, (num_kv_tiles - 1) & 1);
    tc_fence_after();
    const int q_idx = q_tile * kBlockQ + row;
    const float inv_l = 1.0f / l_run;
    __nv_bfloat16* orow = p.out + (static_cast<long long>(b) * p.q_len + q_idx) * p.ldo + col0;
#pragma unroll
    for (int c = 0; c < C::kDPad / 16; ++c) {
      uint32_t r[16];
      tmem_ld16(tmem_o + c * 16, r);
      tmem_ld_wait();
      if (q_idx < p.q_len) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (c * 16 + h * 8 < D) {
            const uint4 pk = make_uint4(
                pack_bf16x2(__uint_as_float(r[8 * h + 0]) * inv_l, __uint_as_float(r[8 * h + 1]) * inv_l),
                pack_bf16x2(__uint_as_float(r[8 * h + 2]) * inv_l, __uint_as_float(r[8 * h + 3]) * inv_l),
                pack_bf16x2(__uint_as_float(r[8 * h + 4]) * inv_l, __uint_as_float(r[8 * h + 5]) * inv_l),
                pack_bf16x2(__uint_as_float(r[8 * h + 6]) * inv_l, __uint_as_float(r[8 * h + 7]) * inv_l));
            *reinterpret_cast<uint4*>(orow + c * 16 + h * 8) = pk;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<C::kTmemCols>(tmem_base);
}

int make_qkv_map(CUtensorMap* map, const void* ptr, int width, int len, int batch, long long ld, int box_rows) {
  const uint64_t dims[3] = {static_cast<uint64_t>(width), static_cast<uint64_t>(len), static_cast<uint64_t>(batch)};
  const uint64_t strides[2] = {static_cast<uint64_t>(ld) * 2, static_cast<uint64_t>(ld) * 2 * static_cast<uint64_t>(len)};
  const uint32_t box[3] = {64, static_cast<uint32_t>(box_rows), 1};
  return make_tensor_map(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box, Swizzle::B128);
}

template <int D>
int launch_attn(const afb_attn_args& a, cudaStream_t stream) {
  using C = AttnCfg<D>;
  CUtensorMap tm_q, tm_k, tm_v;
  const int width = a.heads * a.head_dim;
  if (int rc = make_qkv_map(&tm_q, a.q, width, a.q_len, a.batch, a.ldq, kBlockQ)) return rc;
  if (int rc = make_qkv_map(&tm_k, a.k, width, a.kv_len, a.batch, a.ldk, C::kBlockKV)) return rc;
  if (int rc = make_qkv_map(&tm_v, a.v, width, a.kv_len, a.batch, a.ldv, C::kBlockKV)) return rc;
  AttnDev p;
  p.out = reinterpret_cast<__nv_bfloat16*>(a.out);
  p.key_mask = a.key_mask;
  p.q_len = a.q_len;
  p.kv_len = a.kv_len;
  p.heads = a.heads;
  p.ldo = a.ldo;
  p.scale_log2 = a.scale * 1.4426950408889634f;
  dim3 grid((a.q_len + kBlockQ - 1) / kBlockQ, a.heads, a.batch);
  if (a.key_mask != nullptr) {
    auto kern = attention_fwd_kernel<D, true>;
    static bool attr_set = false;
    if (!attr_set) {
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
      attr_set = true;
    }
    kern<<<grid, kAttnThreads, C::kSmemBytes, stream>>>(tm_q, tm_k, tm_v, p);
  } else {
    auto kern = attention_fwd_kernel<D, false>;
    static bool attr_set = false;
    if (!attr_set) {
      AFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
      attr_set = true;
    }
    kern<<<grid, kAttnThreads, C::kSmemBytes, stream>>>(tm_q, tm_k, tm_v, p);
  }
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace afb

extern "C" int afb_attention_fwd(const afb_attn_args* args, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(args != nullptr, "afb_attention_fwd: args is NULL");
  const afb_attn_args& a = *args;
  AFB_REQUIRE(a.q && a.k && a.v && a.out, "afb_attention_fwd: NULL operand");
  AFB_REQUIRE(a.batch > 0 && a.heads > 0 && a.q_len > 0 && a.kv_len > 0, "afb_attention_fwd: empty problem");
  AFB_REQUIRE(a.batch <= 65535 && a.heads <= 65535, "afb_attention_fwd: batch/heads exceed the grid limits");
  const int width = a.heads * a.head_dim;
  AFB_REQUIRE(a.ldq >= width && a.ldk >= width && a.ldv >= width && a.ldo >= width,
              "afb_attention_fwd: leading dimension smaller than heads*head_dim");
  AFB_REQUIRE(a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0,
              "afb_attention_fwd: leading dimensions must be multiples of 8 elements");
  const uintptr_t all = reinterpret_cast<uintptr_t>(a.q) | reinterpret_cast<uintptr_t>(a.k) |
                        reinterpret_cast<uintptr_t>(a.v) | reinterpret_cast<uintptr_t>(a.out);
  AFB_REQUIRE((all & 15) == 0, "afb_attention_fwd: q/k/v/out must be 16-byte aligned");
  switch (a.head_dim) {
    case 40: return launch_attn<40>(a, stream);
    case 80: return launch_attn<80>(a, stream);
    case 160: return launch_attn<160>(a, stream);
    default:
      AFB_FAIL(2, "afb_attention_fwd: head_dim %d not supported (SD-1.5 uses 40, 80, 160)", a.head_dim);
  }
}
