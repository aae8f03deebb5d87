// Shared device helpers for the B200 paged-attention kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200pa.h"

namespace b200pa {

// Set by every launcher on failure (thread-local; read via b200pa_last_cuda_error()).
extern thread_local int g_last_cuda_error;
extern thread_local const char* g_last_attention_path;
extern thread_local int g_last_attention_partitions;

#define B200PA_CUDA_CHECK(expr)                          \
  do {                                                   \
    cudaError_t _e = (expr);                             \
    if (_e != cudaSuccess) {                             \
      ::b200pa::g_last_cuda_error = static_cast<int>(_e); \
      return B200PA_ERR_CUDA;                            \
    }                                                    \
  } while (0)

// ---- scalar conversions -------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ float to_float(T v);
template <>
__device__ __forceinline__ float to_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_float<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}

template <typename T>
__device__ __forceinline__ T from_float(float v);
template <>
__device__ __forceinline__ float from_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half from_float<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// ---- fp8 KV conversions ---------------------------------------------------------------------
// Semantics of the reference's fp8::scaled_convert (quant_utils.cuh:293-300, 349-359, 406-415
// for loads; :456-488 for stores): quantise through an IEEE fp32 divide + RN-even
// saturate-to-finite convert, dequantise fp8 -> half -> float * scale -> T.
template <int KV>
__device__ __forceinline__ constexpr __nv_fp8_interpretation_t fp8_interp() {
  return KV == B200PA_KV_FP8_E5M2 ? __NV_E5M2 : __NV_E4M3;
}

template <int KV, typename T>
__device__ __forceinline__ uint8_t fp8_quant(T v, float scale) {
  return static_cast<uint8_t>(
      __nv_cvt_float_to_fp8(__fdiv_rn(to_float<T>(v), scale), __NV_SATFINITE, fp8_interp<KV>()));
}

template <int KV>
__device__ __forceinline__ float fp8_to_float_unscaled(uint8_t b) {
  __half_raw h = __nv_cvt_fp8_to_halfraw(static_cast<__nv_fp8_storage_t>(b), fp8_interp<KV>());
  return __half2float(__half(h));
}

template <int KV, typename T>
__device__ __forceinline__ T fp8_dequant(uint8_t b, float scale) {
  return from_float<T>(__fmul_rn(fp8_to_float_unscaled<KV>(b), scale));
}

// Load one cache element as float (KV == AUTO: cache_t == T).
template <typename T, int KV>
struct CacheElem {
  using type = uint8_t;
  static __device__ __forceinline__ float load(const uint8_t* p, float scale) {
    // value the reference kernel would see after scaled_convert<T, uint8_t>
    return to_float<T>(fp8_dequant<KV, T>(*p, scale));
  }
  static __device__ __forceinline__ float val(uint8_t b, float scale) { return to_float<T>(fp8_dequant<KV, T>(b, scale)); }
};
template <typename T>
struct CacheElem<T, B200PA_KV_AUTO> {
  using type = T;
  static __device__ __forceinline__ float load(const T* p, float) { return to_float<T>(*p); }
  static __device__ __forceinline__ float val(T v, float) { return to_float<T>(v); }
};

// Store side: element written to the cache for a source element of type T (KV == AUTO: unchanged).
template <typename T, int KV>
struct CacheStore {
  using type = uint8_t;
  static __device__ __forceinline__ uint8_t cvt(T v, float scale) { return fp8_quant<KV, T>(v, scale); }
};
template <typename T>
struct CacheStore<T, B200PA_KV_AUTO> {
  using type = T;
  static __device__ __forceinline__ T cvt(T v, float) { return v; }
};

// ---- 128-bit streaming accesses ----------------------------------------------------------------
__device__ __forceinline__ int4 ld_nc_v4(const void* ptr) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(ptr));
  return r;
}
__device__ __forceinline__ void st_na_v4(void* ptr, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace b200pa
