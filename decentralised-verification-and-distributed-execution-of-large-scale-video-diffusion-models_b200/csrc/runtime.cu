// Library runtime: error state, device check, launch counter, TMA descriptor encoding.
#include <stdlib.h>
#include <atomic>
#include <cstdarg>
#include <mutex>
#include <string>

#include "common.cuh"
#include "tmap.h"

namespace b2v {

static thread_local std::string g_last_error;
static std::atomic<uint64_t> g_launches{0};

void set_error(const std::string& msg) { g_last_error = msg; }

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  return code;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("B2V_PDL");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- driver entry points for tensor-map encoding (no link-time dependency on libcuda) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn g_encode_tiled = nullptr;
static EncodeIm2colFn g_encode_im2col = nullptr;
static std::once_flag g_entry_once;

static void load_entry_points() {
  std::call_once(g_entry_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      g_encode_tiled = (EncodeTiledFn)fn;
    fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      g_encode_im2col = (EncodeIm2colFn)fn;
  });
}

int tmap_tiled_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box) {
  return tmap_tiled_bf16_sw(out, base, rank, dims, strides_bytes, box, true);
}

int tmap_tiled_bf16_sw(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                       const uint32_t* box, bool swizzle128) {
  load_entry_points();
  if (!g_encode_tiled) return fail(B2V_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = g_encode_tiled(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                              (const cuuint64_t*)dims, (const cuuint64_t*)strides_bytes, (const cuuint32_t*)box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(B2V_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d): rank %d dims %llu,%llu box %u,%u base %p", (int)r,
                rank, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1], base);
  return 0;
}

int tmap_im2col_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                     const int* lower, const int* upper, uint32_t channels_per_pixel, uint32_t pixels_per_column,
                     const uint32_t* elem_strides) {
  load_entry_points();
  if (!g_encode_im2col) return fail(B2V_ERR_CUDA, "cuTensorMapEncodeIm2col entry point unavailable");
  CUresult r = g_encode_im2col(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base),
                               (const cuuint64_t*)dims, (const cuuint64_t*)strides_bytes, lower, upper,
                               channels_per_pixel, pixels_per_column, (const cuuint32_t*)elem_strides,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(B2V_ERR_CUDA, "cuTensorMapEncodeIm2col failed (%d): dims %llu,%llu,%llu,%llu", (int)r,
                (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                (unsigned long long)dims[3]);
  return 0;
}

}  // namespace b2v

extern "C" int b2v_version(void) { return B2V_VERSION; }
extern "C" const char* b2v_last_error(void) { return b2v::g_last_error.c_str(); }
extern "C" uint64_t b2v_launch_count(void) { return b2v::g_launches.load(std::memory_order_relaxed); }

extern "C" int b2v_device_check(void) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return b2v::fail(B2V_ERR_DEVICE, "no CUDA device (%s); libb2v has no CPU fallback", cudaGetErrorString(e));
  static int ok[64] = {0};
  if (dev >= 0 && dev < 64 && ok[dev] == 1) return 0;
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return b2v::fail(B2V_ERR_DEVICE, "cudaDeviceGetAttribute: %s", cudaGetErrorString(e));
  if (major != 10)
    return b2v::fail(B2V_ERR_DEVICE, "device %d is sm_%d; libb2v is built for sm_100a only (no fallback)", dev,
                     major * 10);
  if (dev >= 0 && dev < 64) ok[dev] = 1;
  return 0;
}
