// The two convolutions of the U-Net that have no 64-channel side for the implicit-GEMM kernel:
//   conv_in   openaimodel.py:575-579  Conv2d(4, 320, 3, padding=1) on the latent: an im2col into 64-wide rows (36 used)
//             that the tcgen05 GEMM consumes as a plain [M, 64] x [320, 64]^T product (bias in its epilogue);
//   conv_out  openaimodel.py:716-720  Conv2d(320, 4, 3, padding=1) after GroupNorm + SiLU: reads the 42 MB activation
//             once and writes fp32 NCHW directly (warp-level mma.sync m16n8k16: N = 4 outputs padded to 8 cannot feed a
//             128-row tcgen05 tile; the kernel is bound by the read of its input).
#include "common.cuh"
#include "../../include/adaface_b200.h"

namespace afb {
void count_launch();
namespace {

// ---- im2col for a 4-channel NHWC image: cols[p][tap * 4 + c] = x[pixel p shifted by tap][c], zero outside / beyond 36
__global__ void __launch_bounds__(256) im2col3x3_c4_kernel(const uint2* __restrict__ x, uint4* __restrict__ cols, int H, int W,
                                                           long long total_vecs) {
  pdl_wait();
  pdl_trigger();
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total_vecs) return;
  const int v = static_cast<int>(i & 7);                  // 16-byte vector of the 128-byte row: taps 2v, 2v + 1
  const long long p = i >> 3;
  const int xw = static_cast<int>(p % W);
  const long long r = p / W;
  const int yh = static_cast<int>(r % H);
  uint2 t0 = make_uint2(0u, 0u), t1 = make_uint2(0u, 0u);
  auto tap = [&](int k) {
    const int dy = k / 3 - 1, dx = k % 3 - 1;
    const int yy = yh + dy, xx = xw + dx;
    if (yy < 0 || yy >= H || xx < 0 || xx >= W) return make_uint2(0u, 0u);
    return x[p + static_cast<long long>(dy) * W + dx];
  };
  if (2 * v < 9) t0 = tap(2 * v);
  if (2 * v + 1 < 9) t1 = tap(2 * v + 1);
  cols[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
}

// ---- conv_out: [B, H, W, C] bf16 (C % 64 == 0) -> fp32 NCHW [B, Cout <= 8, H, W]
constexpr int kCoTileH = 8, kCoTileW = 16;                // output pixels per block: one 16-pixel row per warp
constexpr int kCoChunk = 64;                              // channels staged per round
constexpr int kCoPitch = kCoChunk * 2 + 16;               // bytes per halo pixel (padded: conflict-free ldmatrix)
constexpr int kCoHaloPix = (kCoTileH + 2) * (kCoTileW + 2);

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

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;                            // src-size 0: the 16 bytes are zero-filled (the padding)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

// Block = 8 warps = an 8 x 16 tile of output pixels (one 16-pixel row per warp = one m16 tile); the (8+2) x (16+2) halo is
// staged 64 channels at a time through a double-buffered cp.async ring (the next chunk is in flight while the nine taps of
// the current one are multiplied); the cout <= 8 weight rows sit in shared memory for the whole block.
__global__ void __launch_bounds__(kCoTileH * 32) conv3x3_few_out_kernel(const __nv_bfloat16* __restrict__ x,
                                                                       const __nv_bfloat16* __restrict__ wq,   // [8][9 * C] bf16
                                                                       const float* __restrict__ bias, float* __restrict__ out,
                                                                       int H, int W, int C, int cout) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int wpitch = 9 * C + 8;                            // elements per weight row (padded: conflict-free fragment loads)
  __nv_bfloat16* w_s = reinterpret_cast<__nv_bfloat16*>(smem);
  uint8_t* halo = smem + ((static_cast<size_t>(8) * wpitch * 2 + 15) & ~static_cast<size_t>(15));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * kCoTileW, y0 = blockIdx.y * kCoTileH, b = blockIdx.z;
  const uint32_t halo_u = smem_u32(halo);
  // weights do not depend on the previous kernel: requested before the dependency wait (rows >= cout are never read)
  {
    const int vecs_per_row = 9 * C / 8;
    const uint32_t w_u = smem_u32(w_s);
    for (int i = threadIdx.x; i < cout * vecs_per_row; i += blockDim.x) {
      const int row = i / vecs_per_row, col = i - row * vecs_per_row;
      cp_async16(w_u + (row * wpitch + col * 8) * 2, reinterpret_cast<const uint4*>(wq) + i, true);
    }
  }
  pdl_wait();
  pdl_trigger();
  // this thread's share of a halo chunk: pixel / vector indices do not change from chunk to chunk
  constexpr int kVecs = kCoHaloPix * (kCoChunk / 8);                       // 1440
  constexpr int kPerThread = (kVecs + kCoTileH * 32 - 1) / (kCoTileH * 32);   // 6
  auto stage_chunk = [&](int c0, int buf) {
#pragma unroll
    for (int u = 0; u < kPerThread; ++u) {
      const int i = threadIdx.x + u * (kCoTileH * 32);
      if (i < kVecs) {
        const int pix = i >> 3, v = i & 7;
        const int hy = pix / (kCoTileW + 2), hx = pix - hy * (kCoTileW + 2);
        const int yy = y0 + hy - 1, xx = x0 + hx - 1;
        const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
        const __nv_bfloat16* src = ok ? x + ((static_cast<long long>(b) * H + yy) * W + xx) * C + c0 + v * 8 : x;
        cp_async16(halo_u + buf * (kCoHaloPix * kCoPitch) + pix * kCoPitch + v * 16, src, ok);
      }
    }
    cp_async_commit();
  };
  stage_chunk(0, 0);                                       // (the weights ride in this first group)
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  const int g = lane >> 2, t = lane & 3;
  // ldmatrix row of this lane: matrix id = lane / 8 -> pixel rows (id & 1) * 8 + lane % 8, channels (id >> 1) * 8
  const int lm_px = ((lane >> 3) & 1) * 8 + (lane & 7), lm_k = (lane >> 4) * 8;
  const bool wrow_ok = g < cout;
  const int nchunks = C / kCoChunk;
  for (int ci = 0; ci < nchunks; ++ci) {
    const int buf = ci & 1;
    if (ci + 1 < nchunks) {
      stage_chunk((ci + 1) * kCoChunk, buf ^ 1);           // (that buffer was consumed before the barrier that ended round ci - 1)
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t hb = halo_u + buf * (kCoHaloPix * kCoPitch);
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int dy = tap / 3, dx = tap % 3;                // halo coordinates are shifted by +1
      const uint32_t arow = hb + ((warp + dy) * (kCoTileW + 2) + lm_px + dx) * kCoPitch + lm_k * 2;
      const __nv_bfloat16* wrow = w_s + g * wpitch + tap * C + ci * kCoChunk + 2 * t;
#pragma unroll
      for (int kk = 0; kk < kCoChunk / 16; ++kk) {
        uint32_t a[4];
        ldmatrix_x4(a, arow + kk * 32);
        const uint32_t b0 = wrow_ok ? *reinterpret_cast<const uint32_t*>(wrow + kk * 16) : 0u;
        const uint32_t b1 = wrow_ok ? *reinterpret_cast<const uint32_t*>(wrow + kk * 16 + 8) : 0u;
        mma_16816(acc, a, b0, b1);
      }
    }
    __syncthreads();                                       // this buffer may be refilled by the next round's prefetch
  }
  // C fragment: rows g, g + 8 (pixels of the warp's row), columns 2t, 2t + 1 (output channels)
  const int y = y0 + warp;
  if (y < H) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int xx = x0 + g + 8 * h;
      if (xx < W) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int co = 2 * t + j;
          if (co < cout)
            out[((static_cast<long long>(b) * cout + co) * H + y) * W + xx] = acc[2 * h + j] + (bias != nullptr ? bias[co] : 0.f);
        }
      }
    }
  }
}

}  // namespace
}  // namespace afb

extern "C" int afb_im2col3x3_c4(const afb_bf16* x, afb_bf16* cols, int batch, int h, int w, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && cols && batch > 0 && h > 0 && w > 0, "afb_im2col3x3_c4: bad arguments");
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 7) == 0 && (reinterpret_cast<uintptr_t>(cols) & 15) == 0,
              "afb_im2col3x3_c4: x must be 8-byte and cols 16-byte aligned");
  const long long total = static_cast<long long>(batch) * h * w * 8;
  AFB_CHECK_CUDA(launch_pdl(im2col3x3_c4_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0, stream, 1,
                            reinterpret_cast<const uint2*>(x), reinterpret_cast<uint4*>(cols), h, w, total));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int afb_conv3x3_few_out(const afb_bf16* x, const afb_bf16* w8, const float* bias, float* out, int batch, int h,
                                   int w, int channels, int cout, afb_stream_t stream_) {
  using namespace afb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AFB_REQUIRE(x && w8 && out && batch > 0 && h > 0 && w > 0, "afb_conv3x3_few_out: bad arguments");
  AFB_REQUIRE(channels % 64 == 0 && cout >= 1 && cout <= 8, "afb_conv3x3_few_out: channels %% 64 == 0 and 1..8 outputs expected");
  AFB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(w8) & 15) == 0,
              "afb_conv3x3_few_out: x and the weights must be 16-byte aligned");
  AFB_REQUIRE(batch <= 65535, "afb_conv3x3_few_out: batch %d exceeds the grid limit", batch);
  const size_t wbytes = (static_cast<size_t>(8) * (9 * channels + 8) * 2 + 15) & ~static_cast<size_t>(15);
  const size_t smem = wbytes + 2 * static_cast<size_t>(kCoHaloPix) * kCoPitch;      // two halo buffers
  AFB_REQUIRE(smem <= 200 * 1024, "afb_conv3x3_few_out: %d channels do not fit the shared memory", channels);
  static size_t set_for = 0;
  if (set_for < smem) {
    AFB_CHECK_CUDA(cudaFuncSetAttribute(conv3x3_few_out_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    set_for = smem;
  }
  dim3 grid((w + kCoTileW - 1) / kCoTileW, (h + kCoTileH - 1) / kCoTileH, batch);
  AFB_CHECK_CUDA(launch_pdl(conv3x3_few_out_kernel, grid, dim3(kCoTileH * 32), smem, stream, 1,
                            reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<const __nv_bfloat16*>(w8), bias, out, h, w,
                            channels, cout));
  count_launch();
  AFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
