// Developer probe (not part of the product library): per-SM throughput of tcgen05.ld / tcgen05.st, MUFU.EX2 and the
// fp32 -> bf16x2 conversion, the four candidates for what bounds the d = 40 attention softmax.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/tmem_probe.cu -o tools/bin/tmem_probe
#include "../adaface_b200/csrc/common.cuh"
#include <cstdio>
#include <vector>

using namespace afb;

__device__ __forceinline__ void ld64(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]),
        "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]),
        "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]),
        "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]),
        "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr)
      : "memory");
}

// mode 0: tcgen05.ld x32 back to back; 1: x64; 2: x8; 3: tcgen05.st x32; 4: MUFU.EX2; 5: cvt bf16x2; 6: ld x32 + 32 ex2
template <int MODE>
__global__ void __launch_bounds__(512, 1) probe(long long* out, int iters, float seed) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = slot + (static_cast<uint32_t>((warp & 3) * 32) << 16) + ((warp >> 2) * 64) % 448;
  uint32_t acc = 0;
  float facc = seed * lane;
  float fa[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) fa[i] = -seed * (lane + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 6) {
      uint32_t r[32];
      tmem_ld32p(base, r);
      tmem_ld_wait();
      if (MODE == 6) {
#pragma unroll
        for (int i = 0; i < 32; ++i) facc += ex2_approx(__uint_as_float(r[i]) * seed);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= r[i];
      }
    } else if (MODE == 1) {
      uint32_t r[64];
      ld64(base, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 64; ++i) acc ^= r[i];
    } else if (MODE == 2) {
      uint32_t r[8];
      tmem_ld8(base, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 8; ++i) acc ^= r[i];
    } else if (MODE == 3) {
      uint32_t r[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) r[i] = acc + i + it;
      tmem_st32p(base, r);
      tmem_st_wait();
    } else if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < 32; ++i) fa[i & 7] = ex2_approx(fa[i & 7]);
    } else if (MODE == 5) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const uint32_t c = cvt_bf16x2(fa[i & 7], fa[(i + 1) & 7]);
        fa[i & 7] = __uint_as_float(c);
      }
    }
  }
  const long long t1 = clock64();
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(slot);
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
#pragma unroll
  for (int i = 0; i < 8; ++i) facc += fa[i];
  if (acc == 0xdeadbeefu || facc == 1234.5f) out[blockIdx.x + gridDim.x] = acc;
}

template <int MODE>
void run(const char* name, double units_per_warp_iter, const char* unit) {
  long long* d;
  cudaMalloc(&d, 2 * 148 * sizeof(long long));
  const int iters = 2000;
  for (int nw : {4, 8, 16}) {
    probe<MODE><<<148, nw * 32>>>(d, iters, 0.37f);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s failed: %s\n", name, cudaGetErrorString(e)); return; }
    std::vector<long long> h(148);
    cudaMemcpy(h.data(), d, 148 * sizeof(long long), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (auto v : h) avg += v;
    avg /= 148;
    printf("%-28s warps=%2d: %.1f clk / iteration, %.1f %s per clk per SM\n", name, nw, avg / iters,
           units_per_warp_iter * nw * iters / avg, unit);
  }
  cudaFree(d);
}

int main() {
  run<0>("tcgen05.ld 32x32b.x32", 32 * 32 * 4.0, "B");
  run<1>("tcgen05.ld 32x32b.x64", 32 * 64 * 4.0, "B");
  run<2>("tcgen05.ld 32x32b.x8", 32 * 8 * 4.0, "B");
  run<3>("tcgen05.st 32x32b.x32", 32 * 32 * 4.0, "B");
  run<4>("MUFU.EX2 (8 chains/thread)", 32 * 32.0, "ex2");
  run<5>("cvt.rn.bf16x2.f32", 32 * 32.0, "cvt");
  run<6>("ld x32 + 32 ex2", 32 * 32.0, "elem");
  return 0;
}
