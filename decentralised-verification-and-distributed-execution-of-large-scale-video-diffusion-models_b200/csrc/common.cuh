// Shared helpers for the b2v sm_100a kernels: error plumbing, dtype tags, small PTX wrappers.
// Everything in csrc/ is compiled with -gencode arch=compute_100a,code=sm_100a.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/b2v.h"

namespace b2v {

// ---- error plumbing (thread-local last error, negative status codes; see include/b2v.h) ----
void set_error(const std::string& msg);
int fail(int code, const char* fmt, ...);

#define B2V_CUDA_OK(expr)                                                                  \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return ::b2v::fail(B2V_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,        \
                         cudaGetErrorString(_e));                                          \
  } while (0)

#define B2V_CHECK(cond, ...)                                                               \
  do {                                                                                     \
    if (!(cond)) return ::b2v::fail(B2V_ERR_ARG, __VA_ARGS__);                              \
  } while (0)

#define B2V_TRY(expr)                                                                      \
  do {                                                                                     \
    int _s = (expr);                                                                       \
    if (_s != 0) return _s;                                                                \
  } while (0)

static inline cudaError_t launch_check() { return cudaPeekAtLastError(); }

// count of kernels this library launched (bench.py reports it as gpu_launches)
void count_launch(int n = 1);

// ---- programmatic dependent launch (PDL) ----
// The forward is ~1000 back-to-back kernels replayed from a CUDA graph; between two of them the GPU drains, flushes and
// launches (~5 us: 5 % of a forward). Kernels launched through launch_pdl() may start as soon as the CTAs of the
// previous kernel have exited and run their prologue (barrier init, TMEM allocation, tensor-map prefetch) there; every
// such kernel executes pdl_wait() before it touches global memory, which blocks until the previous kernel has fully
// completed and its writes are visible — so the order of memory effects is exactly that of plain stream order.
// (griddepcontrol.wait is a no-op for a kernel launched without the attribute.) B2V_PDL=0 disables the attribute.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

static inline size_t dtype_size(int dt) {
  switch (dt) {
    case B2V_F32: return 4;
    case B2V_F16: return 2;
    case B2V_BF16: return 2;
    default: return 0;
  }
}

template <typename T> struct DT;
template <> struct DT<float> {
  static __device__ __forceinline__ float ld(const float* p) { return *p; }
  static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
};
template <> struct DT<__half> {
  static __device__ __forceinline__ float ld(const __half* p) { return __half2float(*p); }
  static __device__ __forceinline__ void st(__half* p, float v) { *p = __float2half_rn(v); }
};
template <> struct DT<__nv_bfloat16> {
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

// dispatch a generic lambda on a dtype tag
#define B2V_DISPATCH_DTYPE(dt, T, ...)                                                     \
  switch (dt) {                                                                            \
    case B2V_F32: { using T = float; __VA_ARGS__; break; }                                 \
    case B2V_F16: { using T = __half; __VA_ARGS__; break; }                                \
    case B2V_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }                        \
    default: return ::b2v::fail(B2V_ERR_ARG, "unsupported dtype %d", (int)(dt));           \
  }

// ---- device helpers ----
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }
__device__ __forceinline__ float gelu_erf_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// ---- 2-byte activation element: bf16 (default) or fp16 (the reference's torch_dtype=float16, chunk_only.py:44-47).
// Kernels are templated on `bool F16`; both formats run at the same tensor-core rate (tcgen05 kind::f16). ----
template <bool F16>
__device__ __forceinline__ uint32_t pack_act2(float lo, float hi) {
  if constexpr (F16) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    return pack_bf16x2(lo, hi);
  }
}
template <bool F16>
__device__ __forceinline__ float2 unpack_act2(uint32_t u) {
  if constexpr (F16) {
    __half2 v = *reinterpret_cast<__half2*>(&u);
    return __half22float2(v);
  } else {
    return unpack_bf16x2(u);
  }
}
template <bool F16>
__device__ __forceinline__ uint16_t to_act(float x) {
  if constexpr (F16) return __half_as_ushort(__float2half_rn(x));
  else return __bfloat16_as_ushort(__float2bfloat16_rn(x));
}
template <bool F16>
__device__ __forceinline__ float act_to_float(uint16_t v) {
  if constexpr (F16) return __half2float(__ushort_as_half(v));
  else return __bfloat162float(__ushort_as_bfloat16(v));
}
static inline bool act_ok(int adt) { return adt == B2V_BF16 || adt == B2V_F16; }
// run `...` with a constexpr bool F16 selected by the activation dtype tag
#define B2V_DISPATCH_ACT(adt, F16, ...)                                                    \
  do {                                                                                     \
    if ((adt) == B2V_F16) { constexpr bool F16 = true; __VA_ARGS__; }                      \
    else if ((adt) == B2V_BF16) { constexpr bool F16 = false; __VA_ARGS__; }               \
    else return ::b2v::fail(B2V_ERR_ARG, "activation dtype %d is neither bf16 nor fp16", (int)(adt)); \
  } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

int sm_count();

}  // namespace b2v
