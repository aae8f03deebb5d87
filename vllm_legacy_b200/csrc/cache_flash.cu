// reshape_and_cache_flash for B200 (sm_100a): scatter new K/V rows into the flash-attention cache layout
// [num_blocks, block_size, num_heads, head_size].  C-ABI entry point declared in include/b200pa.h.
//
// Semantics follow the reference (csrc/cache_kernels.cu:204-235 kernel, :289-345 host, csrc/cache.h:27-39):
//   dst = block_idx * block_stride + block_offset * num_heads * head_size + i      for i in [0, H*D)
// i.e. one token = one contiguous row of H*D elements on both sides ("auto" KV dtype only; the reference
// rejects fp8 here).  Structure is ours: a flat grid over 16-byte vectors of all tokens (128-bit
// ld.global.nc / st.global), several tokens per CTA, instead of one CTA per token moving 2-byte elements.
#include <algorithm>

#include "common.cuh"

namespace b200pa {

constexpr int kFlashThreads = 256;

// One thread = one 16-byte vector of K and the matching vector of V.  vecs_per_tok = H*D*sizeof(T)/16.
__global__ void __launch_bounds__(kFlashThreads)
    reshape_and_cache_flash_vec_kernel(const char* __restrict__ key, const char* __restrict__ value,
                                       char* __restrict__ k_cache, char* __restrict__ v_cache,
                                       const int64_t* __restrict__ slot_mapping, int64_t key_stride_bytes,
                                       int64_t value_stride_bytes, int64_t block_stride_bytes, int64_t row_bytes,
                                       int vecs_per_tok, int block_size, int64_t total_vecs) {
  for (int64_t idx = blockIdx.x * (int64_t)kFlashThreads + threadIdx.x; idx < total_vecs;
       idx += (int64_t)gridDim.x * kFlashThreads) {
    const int64_t token = idx / vecs_per_tok;
    const int c = (int)(idx - token * vecs_per_tok);
    const int64_t slot = slot_mapping[token];
    if (slot < 0) continue;  // padding token (cache_kernels.cu:218-221)
    const int64_t block_idx = slot / block_size;
    const int64_t block_off = slot - block_idx * block_size;
    const int64_t dst = block_idx * block_stride_bytes + block_off * row_bytes + (int64_t)c * 16;
    const int4 kv = ld_nc_v4(key + token * key_stride_bytes + (int64_t)c * 16);
    const int4 vv = ld_nc_v4(value + token * value_stride_bytes + (int64_t)c * 16);
    *reinterpret_cast<int4*>(k_cache + dst) = kv;
    *reinterpret_cast<int4*>(v_cache + dst) = vv;
  }
}

// Any alignment / any element size: one CTA per token, element granular (index math as the reference).
template <typename T>
__global__ void reshape_and_cache_flash_scalar_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                                      T* __restrict__ k_cache, T* __restrict__ v_cache,
                                                      const int64_t* __restrict__ slot_mapping, int64_t key_stride,
                                                      int64_t value_stride, int64_t block_stride, int n,
                                                      int block_size) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int64_t dst = block_idx * block_stride + block_off * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    k_cache[dst + i] = key[token * key_stride + i];
    v_cache[dst + i] = value[token * value_stride + i];
  }
}

template <typename T>
static int launch_flash_scalar(const b200pa_reshape_flash_params_t* p, cudaStream_t stream) {
  const int n = p->num_heads * p->head_size;
  reshape_and_cache_flash_scalar_kernel<T><<<p->num_tokens, std::min(n, 512), 0, stream>>>(
      static_cast<const T*>(p->key), static_cast<const T*>(p->value), static_cast<T*>(p->key_cache),
      static_cast<T*>(p->value_cache), p->slot_mapping, p->key_stride, p->value_stride, p->block_stride, n,
      p->block_size);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

}  // namespace b200pa

using namespace b200pa;

extern "C" int b200pa_reshape_and_cache_flash(const b200pa_reshape_flash_params_t* p, b200pa_stream_t stream) {
  if (!p) return B200PA_ERR_INVALID_ARG;
  if (p->num_tokens < 0 || p->num_heads <= 0 || p->head_size <= 0 || p->block_size <= 0) return B200PA_ERR_INVALID_ARG;
  if (p->num_tokens == 0) return B200PA_OK;
  if (!p->key || !p->value || !p->key_cache || !p->value_cache || !p->slot_mapping) return B200PA_ERR_INVALID_ARG;
  int esz;
  switch (p->dtype) {
    case B200PA_F16:
    case B200PA_BF16: esz = 2; break;
    case B200PA_F32: esz = 4; break;
    default: return B200PA_ERR_INPUT_DTYPE;
  }
  const int64_t row_bytes = (int64_t)p->num_heads * p->head_size * esz;
  auto a16 = [](const void* q) { return reinterpret_cast<uintptr_t>(q) % 16 == 0; };
  const bool vec_ok = row_bytes % 16 == 0 && a16(p->key) && a16(p->value) && a16(p->key_cache) &&
                      a16(p->value_cache) && (p->key_stride * esz) % 16 == 0 && (p->value_stride * esz) % 16 == 0 &&
                      (p->block_stride * esz) % 16 == 0;
  if (!vec_ok) {
    if (esz == 2) return launch_flash_scalar<uint16_t>(p, stream);
    return launch_flash_scalar<uint32_t>(p, stream);
  }
  const int vecs_per_tok = (int)(row_bytes / 16);
  const int64_t total = (int64_t)p->num_tokens * vecs_per_tok;
  // enough CTAs to cover the vectors once, capped at 8 CTAs per SM (grid-stride beyond that)
  const int blocks = (int)std::min<int64_t>((total + kFlashThreads - 1) / kFlashThreads, 148 * 8);
  reshape_and_cache_flash_vec_kernel<<<blocks, kFlashThreads, 0, stream>>>(
      static_cast<const char*>(p->key), static_cast<const char*>(p->value), static_cast<char*>(p->key_cache),
      static_cast<char*>(p->value_cache), p->slot_mapping, p->key_stride * esz, p->value_stride * esz,
      p->block_stride * esz, row_bytes, vecs_per_tok, p->block_size, total);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}
