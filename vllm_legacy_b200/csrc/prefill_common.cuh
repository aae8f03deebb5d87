// Kernel-side parameter block shared by the two prefix-prefill kernel families
// (prefill.cu: mma.sync; prefill_tc.cu: tcgen05 + TMEM).
#pragma once

#include <type_traits>

#include "attention_common.cuh"

namespace b200pa {

struct PrefillParams {
  void* out;
  const void* q;
  const void* k;
  const void* v;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  const int* start_loc;
  const int* seq_lens;
  const int* ctx_lens;
  const float* alibi_slopes;
  int64_t q_stride_t, q_stride_h, k_stride_t, k_stride_h, v_stride_t, v_stride_h, o_stride_t, o_stride_h;
  int64_t kc_block_stride, kc_head_stride, vc_block_stride, vc_head_stride;  // cache elements (bytes for fp8)
  int64_t block_table_stride;
  int num_heads, num_kv_heads, head_size, block_shift, sliding_window;
  float scale_log2;  // scale * log2(e)
  int kv_dtype;      // B200PA_KV_AUTO: 16-bit cache (x = 8); fp8: uint8 cache (x = 16), dequantised while staging
  float kv_scale;
  int use_tma;       // prefill_tc.cu: full cached-prefix tiles staged by TMA (set by the launcher when a tensor map exists)
};

// ---- fp8 cache bytes -> T while staging a cached K/V tile -------------------------------------------
// Same arithmetic as the decode kernels' scaled_convert (quant_utils.cuh:293-454): fp8 -> half (exact) -> float
// x kv_scale -> T (round to nearest even).  8 bytes in, one 16-byte shared-memory row piece out.
// Two bytes at a time (cvt.rn.f16x2.e?m?x2, two f16 -> f32, two fmul.rn, one packed round): the same operations per
// element as fp8_dequant, so the results are bit-identical, in about half the instructions.
template <typename T, int KV>
__device__ __forceinline__ uint32_t fp8x2_to_t(uint32_t two_bytes, float scale) {
  const __half2_raw hr = __nv_cvt_fp8x2_to_halfraw2(static_cast<__nv_fp8x2_storage_t>(two_bytes & 0xffffu), fp8_interp<KV>());
  const float2 f = __half22float2(__half2(hr));
  const float lo = __fmul_rn(f.x, scale), hi = __fmul_rn(f.y, scale);
  if constexpr (std::is_same<T, __half>::value) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
  } else {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
}
template <typename T, int KV>
__device__ __forceinline__ uint4 fp8x8_to_t(uint2 b, float scale) {
  return make_uint4(fp8x2_to_t<T, KV>(b.x, scale), fp8x2_to_t<T, KV>(b.x >> 16, scale),
                    fp8x2_to_t<T, KV>(b.y, scale), fp8x2_to_t<T, KV>(b.y >> 16, scale));
}
template <typename T>
__device__ __forceinline__ uint4 fp8x8_to_t(int kv, uint2 b, float scale) {
  return kv == B200PA_KV_FP8_E5M2 ? fp8x8_to_t<T, B200PA_KV_FP8_E5M2>(b, scale)
                                  : fp8x8_to_t<T, B200PA_KV_FP8_E4M3>(b, scale);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint2 ld_nc_v2(const void* ptr) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(ptr));
  return r;
}
// keep the first n (0..8) of 8 consecutive 16-bit elements, zero the rest (tokens behind ctx_len)
__device__ __forceinline__ uint4 keep_first_elems(uint4 v, int n) {
  uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = n - 2 * i;  // elements of word i that stay
    w[i] = k >= 2 ? w[i] : (k == 1 ? (w[i] & 0xffffu) : 0u);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// prefill_tc.cu — returns B200PA_OK, 1 when the shape has no tcgen05 instantiation, or a negative status.
int launch_prefill_tc(const PrefillParams& kp, int batch, int max_query_len, int dtype, cudaStream_t stream);

}  // namespace b200pa
