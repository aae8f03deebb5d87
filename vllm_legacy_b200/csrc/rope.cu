// rotary_embedding / batched_rotary_embedding for B200 (sm_100a), optionally fused with the
// reshape_and_cache that follows it in the decode step.  C-ABI: b200pa_rotary_embedding
// (include/b200pa.h).
//
// Reference semantics: csrc/pos_encoding_kernels.cu:10-131 (one CTA per token, NeoX pairs
// (i, i + rot_dim/2), GPT-J pairs (2i, 2i+1), cos/sin read from cos_sin_cache[pos (+offset)]),
// followed for the fused form by csrc/cache_kernels.cu:150-202.  Not its structure: each thread
// rotates one 16-byte vector of x and of y (NeoX) or one 16-byte vector of interleaved pairs
// (GPT-J), and the fused form writes the rotated K group straight into the paged K layout
// [nb, H, D/x, B, x] with one vector store, so q/k/v of a decode token are touched once instead of
// by two launches.
//
// Bit-exactness: the reference computes `x*cos - y*sin` and `y*cos + x*sin` with c10::Half /
// c10::BFloat16 operators, i.e. every product and every sum is evaluated in fp32 and rounded (RN-even)
// back to the 16-bit type before the next operation (torch/headeronly/util/Half.h:501-511,
// BFloat16.h:185-198).  For fp32 nvcc contracts the expressions to
// fma(x, cos, -(y*sin)) and fma(x, sin, y*cos) (SASS of the reference build: FMUL; FMUL; FFMA; FFMA).
// Both forms are written out explicitly below with non-contractable intrinsics.
#include <algorithm>

#include "common.cuh"

namespace b200pa {

template <typename T>
__device__ __forceinline__ void rotate_pair(T x, T y, T c, T s, T& nx, T& ny) {
  const float xf = to_float<T>(x), yf = to_float<T>(y), cf = to_float<T>(c), sf = to_float<T>(s);
  const float xc = to_float<T>(from_float<T>(__fmul_rn(xf, cf)));
  const float ys = to_float<T>(from_float<T>(__fmul_rn(yf, sf)));
  const float yc = to_float<T>(from_float<T>(__fmul_rn(yf, cf)));
  const float xs = to_float<T>(from_float<T>(__fmul_rn(xf, sf)));
  nx = from_float<T>(__fsub_rn(xc, ys));
  ny = from_float<T>(__fadd_rn(yc, xs));
}
template <>
__device__ __forceinline__ void rotate_pair<float>(float x, float y, float c, float s, float& nx, float& ny) {
  nx = __fmaf_rn(x, c, -__fmul_rn(y, s));
  ny = __fmaf_rn(x, s, __fmul_rn(y, c));
}

template <typename T, int VEC>
struct alignas(sizeof(T) * VEC) VecT {
  T e[VEC];
};

template <typename T, int VEC>
__device__ __forceinline__ VecT<T, VEC> ld_vec(const T* p) {
  return *reinterpret_cast<const VecT<T, VEC>*>(p);
}
template <typename T, int VEC>
__device__ __forceinline__ void st_vec(T* p, const VecT<T, VEC>& v) {
  *reinterpret_cast<VecT<T, VEC>*>(p) = v;
}

struct RopeKernelParams {
  const int64_t* positions;
  void* query;
  void* key;
  const void* cos_sin_cache;
  const int64_t* offsets;
  int64_t query_stride, key_stride, value_stride;
  const void* value;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  int num_heads, num_kv_heads, head_size, rot_dim, block_size, x;
  float kv_scale;
};

// Write VEC consecutive head elements [d0, d0+VEC) of (kv head h) of one token into the paged K
// layout (they sit inside one x-group when x % VEC == 0; otherwise element by element).
template <typename T, int KV, int VEC>
__device__ __forceinline__ void store_k_cache(typename CacheStore<T, KV>::type* kblk, const VecT<T, VEC>& v, int h,
                                              int d0, int block_off, const RopeKernelParams& p) {
  using CT = typename CacheStore<T, KV>::type;
  const int x = p.x;
  VecT<CT, VEC> o;
#pragma unroll
  for (int j = 0; j < VEC; ++j) o.e[j] = CacheStore<T, KV>::cvt(v.e[j], p.kv_scale);
  if (VEC > 1 && x % VEC == 0) {
    CT* dst = kblk + (((int64_t)h * (p.head_size / x) + d0 / x) * p.block_size + block_off) * x + d0 % x;
    st_vec<CT, VEC>(dst, o);
  } else {
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      const int d = d0 + j;
      kblk[(((int64_t)h * (p.head_size / x) + d / x) * p.block_size + block_off) * x + d % x] = o.e[j];
    }
  }
}

// One CTA per token.  Work items of a token, strided over the CTA's threads:
//   [0, nq)            rotation units of the query heads
//   [nq, nq + nkr)     rotation units of the key heads (+ K-cache store when FUSE)
//   then (FUSE only)   pass-through units of the key heads (d >= rot_dim) and the value vectors.
// NEOX : unit u of a head = x vector at d0 = u*VEC and y vector at rot_dim/2 + d0.
// GPT-J: unit u of a head = GV = max(VEC, 2) interleaved elements (GV/2 pairs) at d0 = u*GV.
template <typename T, int KV, bool NEOX, bool FUSE, int VEC>
__global__ void __launch_bounds__(256)
    rope_kernel(const __grid_constant__ RopeKernelParams p) {
  using CT = typename CacheStore<T, KV>::type;
  constexpr int GV = NEOX ? VEC : (VEC < 2 ? 2 : VEC);
  const int64_t token = blockIdx.x;
  int64_t pos = p.positions[token];
  if (p.offsets) pos += p.offsets[token];
  const int embed = p.rot_dim / 2;
  const T* cos_ptr = static_cast<const T*>(p.cos_sin_cache) + pos * p.rot_dim;
  const T* sin_ptr = cos_ptr + embed;
  T* qtok = static_cast<T*>(p.query) + token * p.query_stride;
  T* ktok = static_cast<T*>(p.key) + token * p.key_stride;

  const int RU = NEOX ? embed / VEC : p.rot_dim / GV;
  const int nq = p.num_heads * RU;
  const int nkr = p.num_kv_heads * RU;
  int PU = 0, VU = 0, block_off = 0;
  bool write_cache = false;
  CT* kblk = nullptr;
  CT* vblk = nullptr;
  const T* vtok = nullptr;
  if constexpr (FUSE) {
    const int64_t slot = p.slot_mapping[token];
    write_cache = slot >= 0;  // padding tokens are rotated but not cached (cache_kernels.cu:163-167)
    if (write_cache) {
      PU = (p.head_size - p.rot_dim) / VEC;
      VU = p.head_size / VEC;
      const int64_t block_idx = slot / p.block_size;
      block_off = (int)(slot - block_idx * p.block_size);
      const int64_t blk_elems = (int64_t)p.num_kv_heads * p.head_size * p.block_size;
      kblk = static_cast<CT*>(p.key_cache) + block_idx * blk_elems;
      vblk = static_cast<CT*>(p.value_cache) + block_idx * blk_elems;
      vtok = static_cast<const T*>(p.value) + token * p.value_stride;
    }
  }
  const int total = nq + nkr + p.num_kv_heads * (PU + VU);

  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    if (i < nq + nkr) {
      const bool is_k = i >= nq;
      const int j = is_k ? i - nq : i;
      const int h = j / RU;
      const int u = j - h * RU;
      T* arr = (is_k ? ktok : qtok) + (int64_t)h * p.head_size;
      if constexpr (NEOX) {
        const int d0 = u * VEC;
        VecT<T, VEC> xv = ld_vec<T, VEC>(arr + d0), yv = ld_vec<T, VEC>(arr + embed + d0);
        const VecT<T, VEC> cv = ld_vec<T, VEC>(cos_ptr + d0), sv = ld_vec<T, VEC>(sin_ptr + d0);
#pragma unroll
        for (int e = 0; e < VEC; ++e) rotate_pair<T>(xv.e[e], yv.e[e], cv.e[e], sv.e[e], xv.e[e], yv.e[e]);
        st_vec<T, VEC>(arr + d0, xv);
        st_vec<T, VEC>(arr + embed + d0, yv);
        if constexpr (FUSE) {
          if (is_k && write_cache) {
            store_k_cache<T, KV, VEC>(kblk, xv, h, d0, block_off, p);
            store_k_cache<T, KV, VEC>(kblk, yv, h, embed + d0, block_off, p);
          }
        }
      } else {
        const int d0 = u * GV;
        VecT<T, GV> v;
        if constexpr (VEC > 1) {
          v = ld_vec<T, GV>(arr + d0);
        } else {
          v.e[0] = arr[d0];
          v.e[1] = arr[d0 + 1];
        }
        VecT<T, GV / 2> cv, sv;
        if constexpr (VEC > 1) {
          cv = ld_vec<T, GV / 2>(cos_ptr + d0 / 2);
          sv = ld_vec<T, GV / 2>(sin_ptr + d0 / 2);
        } else {
          cv.e[0] = cos_ptr[d0 / 2];
          sv.e[0] = sin_ptr[d0 / 2];
        }
#pragma unroll
        for (int e = 0; e < GV / 2; ++e)
          rotate_pair<T>(v.e[2 * e], v.e[2 * e + 1], cv.e[e], sv.e[e], v.e[2 * e], v.e[2 * e + 1]);
        if constexpr (VEC > 1) {
          st_vec<T, GV>(arr + d0, v);
        } else {
          arr[d0] = v.e[0];
          arr[d0 + 1] = v.e[1];
        }
        if constexpr (FUSE) {
          if (is_k && write_cache) {
            if constexpr (VEC > 1) {
              store_k_cache<T, KV, GV>(kblk, v, h, d0, block_off, p);
            } else {
              VecT<T, 1> a{{v.e[0]}}, b{{v.e[1]}};
              store_k_cache<T, KV, 1>(kblk, a, h, d0, block_off, p);
              store_k_cache<T, KV, 1>(kblk, b, h, d0 + 1, block_off, p);
            }
          }
        }
      }
    } else if constexpr (FUSE) {
      int j = i - nq - nkr;
      const int npass = p.num_kv_heads * PU;
      if (j < npass) {  // key elements beyond rot_dim: cached unchanged
        const int h = j / PU;
        const int d0 = p.rot_dim + (j - h * PU) * VEC;
        VecT<T, VEC> v;
        if constexpr (VEC > 1) v = ld_vec<T, VEC>(ktok + (int64_t)h * p.head_size + d0);
        else v.e[0] = ktok[(int64_t)h * p.head_size + d0];
        store_k_cache<T, KV, VEC>(kblk, v, h, d0, block_off, p);
      } else {          // value: token-minor layout [nb, H, D, B], one element every block_size
        j -= npass;
        const int h = j / VU;
        const int d0 = (j - h * VU) * VEC;
        VecT<T, VEC> v;
        if constexpr (VEC > 1) v = ld_vec<T, VEC>(vtok + (int64_t)h * p.head_size + d0);
        else v.e[0] = vtok[(int64_t)h * p.head_size + d0];
        CT* dst = vblk + ((int64_t)h * p.head_size + d0) * p.block_size + block_off;
#pragma unroll
        for (int e = 0; e < VEC; ++e) dst[(int64_t)e * p.block_size] = CacheStore<T, KV>::cvt(v.e[e], p.kv_scale);
      }
    }
  }
}

template <typename T, int KV, bool FUSE>
static int launch_rope(const b200pa_rope_params_t* a, cudaStream_t stream) {
  using CT = typename CacheStore<T, KV>::type;
  constexpr int VEC = 16 / sizeof(T);
  RopeKernelParams kp{};
  kp.positions = a->positions;
  kp.query = a->query;
  kp.key = a->key;
  kp.cos_sin_cache = a->cos_sin_cache;
  kp.offsets = a->cos_sin_cache_offsets;
  kp.query_stride = a->query_stride;
  kp.key_stride = a->key_stride;
  kp.value_stride = a->value_stride;
  kp.value = a->value;
  kp.key_cache = a->key_cache;
  kp.value_cache = a->value_cache;
  kp.slot_mapping = a->slot_mapping;
  kp.num_heads = a->num_heads;
  kp.num_kv_heads = a->num_kv_heads;
  kp.head_size = a->head_size;
  kp.rot_dim = a->rot_dim;
  kp.block_size = a->block_size;
  kp.x = a->x;
  kp.kv_scale = a->kv_scale;

  auto al16 = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  const int embed = a->rot_dim / 2;
  bool vec_ok = al16(a->query) && al16(a->key) && al16(a->cos_sin_cache) &&
                (a->query_stride * sizeof(T)) % 16 == 0 && (a->key_stride * sizeof(T)) % 16 == 0 &&
                a->head_size % VEC == 0 && (a->is_neox ? embed % VEC == 0 : a->rot_dim % VEC == 0) &&
                (a->rot_dim * sizeof(T)) % 16 == 0;
  if (FUSE)
    vec_ok = vec_ok && al16(a->value) && al16(a->key_cache) && (a->value_stride * sizeof(T)) % 16 == 0 &&
             (a->head_size - a->rot_dim) % VEC == 0;  // x % VEC != 0 only costs the vector K store

  const int ru = a->is_neox ? (vec_ok ? embed / VEC : embed) : (vec_ok ? a->rot_dim / VEC : embed);
  int items = (a->num_heads + a->num_kv_heads) * ru;
  if (FUSE) {
    const int v = vec_ok ? VEC : 1;
    items += a->num_kv_heads * ((a->head_size - a->rot_dim) / v + a->head_size / v);
  }
  int threads = std::min(std::max(items, 32), 256);
  threads = (threads + 31) / 32 * 32;
  dim3 grid(a->num_tokens);
#define B200PA_ROPE_LAUNCH(NEOX, V) rope_kernel<T, KV, NEOX, FUSE, V><<<grid, threads, 0, stream>>>(kp)
  if (vec_ok) {
    if (a->is_neox) B200PA_ROPE_LAUNCH(true, VEC);
    else B200PA_ROPE_LAUNCH(false, VEC);
  } else {
    if (a->is_neox) B200PA_ROPE_LAUNCH(true, 1);
    else B200PA_ROPE_LAUNCH(false, 1);
  }
#undef B200PA_ROPE_LAUNCH
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

template <typename T>
static int dispatch_rope_kv(const b200pa_rope_params_t* a, cudaStream_t stream) {
  if (!a->key_cache) return launch_rope<T, B200PA_KV_AUTO, false>(a, stream);
  switch (a->kv_dtype) {
    case B200PA_KV_AUTO: return launch_rope<T, B200PA_KV_AUTO, true>(a, stream);
    case B200PA_KV_FP8_E4M3: return launch_rope<T, B200PA_KV_FP8_E4M3, true>(a, stream);
    case B200PA_KV_FP8_E5M2: return launch_rope<T, B200PA_KV_FP8_E5M2, true>(a, stream);
    default: return B200PA_ERR_KV_DTYPE;
  }
}

}  // namespace b200pa

extern "C" int b200pa_rotary_embedding(const b200pa_rope_params_t* a, b200pa_stream_t stream) {
  using namespace b200pa;
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->num_tokens < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->head_size <= 0 || a->rot_dim <= 0 ||
      a->rot_dim % 2 != 0 || a->rot_dim > a->head_size)
    return B200PA_ERR_INVALID_ARG;
  if (a->num_tokens == 0) return B200PA_OK;
  if (!a->positions || !a->query || !a->key || !a->cos_sin_cache) return B200PA_ERR_INVALID_ARG;
  if (a->key_cache) {
    if (!a->value || !a->value_cache || !a->slot_mapping || a->block_size <= 0 || a->x <= 0 ||
        a->head_size % a->x != 0)
      return B200PA_ERR_INVALID_ARG;
  }
  switch (a->dtype) {
    case B200PA_F16: return dispatch_rope_kv<__half>(a, stream);
    case B200PA_BF16: return dispatch_rope_kv<__nv_bfloat16>(a, stream);
    case B200PA_F32: return dispatch_rope_kv<float>(a, stream);
    default: return B200PA_ERR_INPUT_DTYPE;
  }
}
