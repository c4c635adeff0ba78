// K14: classifier-free-guidance combine + DDIM update (ldm/models/diffusion/ddim.py:252-296) and the
// dtype casts on the path's edges.  HBM-bound streams: 128-bit loads/stores, grid sized in SM multiples.
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

struct DdimCoef {
  float scale, sqrt_one_minus_at, sqrt_at, dir_coef, sqrt_a_prev, sigma_t, temperature;
};

// Same operation order as the reference and no FMA contraction (__fmul_rn/__fadd_rn are never fused), so
// the result is bit-identical to the eager fp32 torch sequence.
__device__ __forceinline__ void ddim_one(float x, float ec, float eu, float nz, const DdimCoef& c, float& x_prev,
                                         float& pred_x0) {
  const float e = __fadd_rn(eu, __fmul_rn(c.scale, __fsub_rn(ec, eu)));                 // :260
  pred_x0 = __fdiv_rn(__fsub_rn(x, __fmul_rn(c.sqrt_one_minus_at, e)), c.sqrt_at);      // :279
  const float dir = __fmul_rn(c.dir_coef, e);                                           // :283
  const float noise = __fmul_rn(__fmul_rn(c.sigma_t, nz), c.temperature);               // :287
  x_prev = __fadd_rn(__fadd_rn(__fmul_rn(c.sqrt_a_prev, pred_x0), dir), noise);         // :295
}

template <bool kHasNoise>
__global__ void __launch_bounds__(256) cfg_ddim_kernel(const float* __restrict__ x, const float* __restrict__ ec,
                                                       const float* __restrict__ eu, const float* __restrict__ nz,
                                                       float* __restrict__ x_prev, float* __restrict__ pred_x0,
                                                       long long n, DdimCoef c) {
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 vx = reinterpret_cast<const float4*>(x)[i];
    const float4 vc = reinterpret_cast<const float4*>(ec)[i];
    const float4 vu = reinterpret_cast<const float4*>(eu)[i];
    float4 vn = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kHasNoise) vn = reinterpret_cast<const float4*>(nz)[i];
    float4 xp, p0;
    ddim_one(vx.x, vc.x, vu.x, vn.x, c, xp.x, p0.x);
    ddim_one(vx.y, vc.y, vu.y, vn.y, c, xp.y, p0.y);
    ddim_one(vx.z, vc.z, vu.z, vn.z, c, xp.z, p0.z);
    ddim_one(vx.w, vc.w, vu.w, vn.w, c, xp.w, p0.w);
    reinterpret_cast<float4*>(x_prev)[i] = xp;
    reinterpret_cast<float4*>(pred_x0)[i] = p0;
  }
  // tail (n % 4)
  for (long long i = (n4 << 2) + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float xp, p0;
    ddim_one(x[i], ec[i], eu[i], kHasNoise ? nz[i] : 0.f, c, xp, p0);
    x_prev[i] = xp;
    pred_x0[i] = p0;
  }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                            long long n) {
  const long long n8 = n >> 3;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    reinterpret_cast<uint4*>(dst)[i] =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
  for (long long i = (n8 << 3) + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                                                            long long n) {
  const long long n8 = n >> 3;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const uint4 u = reinterpret_cast<const uint4*>(src)[i];
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    reinterpret_cast<float4*>(dst)[2 * i] = make_float4(a.x, a.y, b.x, b.y);
    reinterpret_cast<float4*>(dst)[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
  }
  for (long long i = (n8 << 3) + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = __bfloat162float(src[i]);
}

// out[r][c] = a[r][c] + b[r][c] + bias[c]; 8 channels (16 bytes) per thread
__global__ void __launch_bounds__(256) add_bias_residual_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b,
                                                                const float* __restrict__ bias, uint4* __restrict__ out,
                                                                long long nvec, int vpr) {
  pdl_wait();
  pdl_trigger();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 ua = a[i];
    float v[8];
    float2 t;
    t = unpack_bf16x2(ua.x); v[0] = t.x; v[1] = t.y;
    t = unpack_bf16x2(ua.y); v[2] = t.x; v[3] = t.y;
    t = unpack_bf16x2(ua.z); v[4] = t.x; v[5] = t.y;
    t = unpack_bf16x2(ua.w); v[6] = t.x; v[7] = t.y;
    if (b != nullptr) {
      const uint4 ub = b[i];
      t = unpack_bf16x2(ub.x); v[0] += t.x; v[1] += t.y;
      t = unpack_bf16x2(ub.y); v[2] += t.x; v[3] += t.y;
      t = unpack_bf16x2(ub.z); v[4] += t.x; v[5] += t.y;
      t = unpack_bf16x2(ub.w); v[6] += t.x; v[7] += t.y;
    }
    if (bias != nullptr) {
      const int c0 = static_cast<int>(i % vpr) * 8;
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + c0));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + c0 + 4));
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
      v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    out[i] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}

inline unsigned stream_grid(long long work_items) {
  const long long blocks = (work_items + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  return static_cast<unsigned>(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace
}  // namespace afb

extern "C" int afb_cfg_ddim_step(const float* x, const float* e_cond, const float* e_uncond, const float* noise,
                                 float* x_prev, float* pred_x0, int64_t n, float guidance_scale, float a_t, float a_prev,
                                 float sigma_t, float sqrt_one_minus_at, float temperature, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && e_cond && e_uncond && x_prev && pred_x0, "afb_cfg_ddim_step: NULL argument");
  AFB_REQUIRE(n >= 0, "afb_cfg_ddim_step: negative size");
  AFB_REQUIRE(noise != nullptr || sigma_t == 0.f, "afb_cfg_ddim_step: noise is required when sigma_t != 0");
  if (n == 0) return 0;
  const uintptr_t all = reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(e_cond) |
                        reinterpret_cast<uintptr_t>(e_uncond) | reinterpret_cast<uintptr_t>(x_prev) |
                        reinterpret_cast<uintptr_t>(pred_x0) | reinterpret_cast<uintptr_t>(noise);
  AFB_REQUIRE((all & 15) == 0, "afb_cfg_ddim_step: buffers must be 16-byte aligned");
  DdimCoef c;
  c.scale = guidance_scale;
  c.sqrt_one_minus_at = sqrt_one_minus_at;
  c.sqrt_at = sqrtf(a_t);                                   // a_t.sqrt()                      ddim.py:279
  c.dir_coef = sqrtf((1.0f - a_prev) - sigma_t * sigma_t);  // (1. - a_prev - sigma_t**2).sqrt() ddim.py:283
  c.sqrt_a_prev = sqrtf(a_prev);                            // a_prev.sqrt()                   ddim.py:295
  c.sigma_t = sigma_t;
  c.temperature = temperature;
  const unsigned grid = stream_grid((n + 3) / 4);
  if (noise != nullptr)
    cfg_ddim_kernel<true><<<grid, 256, 0, stream>>>(x, e_cond, e_uncond, noise, x_prev, pred_x0, n, c);
  else
    cfg_ddim_kernel<false><<<grid, 256, 0, stream>>>(x, e_cond, e_uncond, nullptr, x_prev, pred_x0, n, c);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_cast_f32_to_bf16(const float* src, afb_bf16* dst, int64_t n, afb_stream_t stream_) {
  using namespace afb;
  AFB_REQUIRE(src && dst && n >= 0, "afb_cast_f32_to_bf16: bad argument");
  if (n == 0) return 0;
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0,
              "afb_cast_f32_to_bf16: buffers must be 16-byte aligned");
  cast_f32_bf16_kernel<<<stream_grid((n + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_cast_bf16_to_f32(const afb_bf16* src, float* dst, int64_t n, afb_stream_t stream_) {
  using namespace afb;
  AFB_REQUIRE(src && dst && n >= 0, "afb_cast_bf16_to_f32: bad argument");
  if (n == 0) return 0;
  AFB_REQUIRE(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0,
              "afb_cast_bf16_to_f32: buffers must be 16-byte aligned");
  cast_bf16_f32_kernel<<<stream_grid((n + 7) / 8), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), dst, n);
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_add_bias_residual(const afb_bf16* a, const afb_bf16* b, const float* bias, afb_bf16* out, int64_t rows,
                                     int channels, afb_stream_t stream_) {
  using namespace afb;
  AFB_REQUIRE(a && out && rows >= 0 && channels > 0, "afb_add_bias_residual: bad argument");
  AFB_REQUIRE(channels % 8 == 0, "afb_add_bias_residual: channels %d must be a multiple of 8", channels);
  if (rows == 0) return 0;
  const uintptr_t all = reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(out) |
                        reinterpret_cast<uintptr_t>(bias);
  AFB_REQUIRE((all & 15) == 0, "afb_add_bias_residual: buffers must be 16-byte aligned");
  const long long nvec = static_cast<long long>(rows) * (channels / 8);
  AFB_CHECK_CUDA(launch_pdl(add_bias_residual_kernel, dim3(stream_grid(nvec)), dim3(256), 0, static_cast<cudaStream_t>(stream_), 1,
                            reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), bias,
                            reinterpret_cast<uint4*>(out), nvec, channels / 8));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
