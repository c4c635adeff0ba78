// Host-side plumbing of the C ABI: thread-local error message, device query, launch counter and TMA
// tensor-map construction through the driver entry point (no link-time dependency on libcuda).
#include "common.cuh"
#include <stdlib.h>
#include "../../include/adaface_b200.h"

#include <atomic>
#include <cstdarg>

namespace afb {

namespace {
thread_local char g_error[512] = "";
std::atomic<int64_t> g_launches{0};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
  }
  return fn;
}
}  // namespace

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("AFB_PDL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
}

int make_tensor_map(CUtensorMap* map, CUtensorMapDataType dtype, int rank, const void* base, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, Swizzle swizzle, CUtensorMapL2promotion l2,
                    const uint32_t* element_strides) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) AFB_FAIL(3, "cuTensorMapEncodeTiled is not available (no CUDA driver / device?)");
  cuuint64_t gdims[5];
  cuuint64_t gstrides[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = element_strides != nullptr ? element_strides[i] : 1;
    if (i + 1 < rank) gstrides[i] = strides_bytes[i];
  }
  CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_NONE;
  switch (swizzle) {
    case Swizzle::B32: sw = CU_TENSOR_MAP_SWIZZLE_32B; break;
    case Swizzle::B64: sw = CU_TENSOR_MAP_SWIZZLE_64B; break;
    case Swizzle::B128: sw = CU_TENSOR_MAP_SWIZZLE_128B; break;
    default: break;
  }
  CUresult r = fn(map, dtype, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, sw, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    AFB_FAIL(3, "cuTensorMapEncodeTiled failed (CUresult %d; rank %d dims %llu,%llu box %u,%u stride0 %llu base %p)",
             static_cast<int>(r), rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0],
             rank > 1 ? box[1] : 0, (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), base);
  }
  return 0;
}

}  // namespace afb

extern "C" const char* afb_last_error(void) { return afb::g_error; }
extern "C" int afb_abi_version(void) { return 13; }
#ifndef AFB_HEADER_DIGEST
#define AFB_HEADER_DIGEST "unknown"
#endif
extern "C" const char* afb_header_digest(void) { return AFB_HEADER_DIGEST; }
extern "C" int64_t afb_launch_count(void) { return afb::g_launches.load(std::memory_order_relaxed); }
