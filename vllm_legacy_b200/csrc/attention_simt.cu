// General SIMT paged-attention kernel (any head size in the reference's list, block 8/16/32,
// fp16/bf16/fp32 queries, auto/fp8 KV, ALiBi, block-sparse) and the split-K reduce kernel.
//
// This family covers what the tensor-core/TMA family (attention_mma.cu) does not instantiate:
// fp32 queries, block size 8, block-sparse masks.  Same math as the reference kernel
// (csrc/attention/attention_kernels.cu:86-496): logit = scale * q.k + alibi * (j - len + 1),
// p = exp(logit - max) / (sum + 1e-6), out = sum_j p_j v_j — but computed with an online softmax
// per warp (no full-sequence logits in shared memory, so no max_seq_len limit for v1), lanes
// mapped to consecutive tokens for K (16-byte loads) and to consecutive head-dim rows for V.
#include <cfloat>
#include <type_traits>

#include "attention_common.cuh"

namespace b200pa {

// Element j (a compile-time constant after unrolling) of a 16-byte group, without a union (a union of an int4 and a
// byte / float array is spilled to local memory by the compiler).
__device__ __forceinline__ int word_of(const int4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }
template <typename CT>
__device__ __forceinline__ CT elem_of(const int4& v, int j);
template <>
__device__ __forceinline__ float elem_of<float>(const int4& v, int j) { return __int_as_float(word_of(v, j)); }
template <>
__device__ __forceinline__ __half elem_of<__half>(const int4& v, int j) {
  return __ushort_as_half((unsigned short)((unsigned)word_of(v, j >> 1) >> (16 * (j & 1))));
}
template <>
__device__ __forceinline__ __nv_bfloat16 elem_of<__nv_bfloat16>(const int4& v, int j) {
  return __ushort_as_bfloat16((unsigned short)((unsigned)word_of(v, j >> 1) >> (16 * (j & 1))));
}
template <>
__device__ __forceinline__ uint8_t elem_of<uint8_t>(const int4& v, int j) {
  return (uint8_t)((unsigned)word_of(v, j >> 2) >> (8 * (j & 3)));
}

// Elements j, j + 1 (j even, a compile-time constant after unrolling) of a 16-byte group as floats, i.e. the value the
// reference kernel sees after its load-time conversion.  16-bit caches convert the packed pair straight from its word
// (bf16: one shift / one mask; fp16: one packed convert) instead of extracting and converting element by element.
template <typename T, int KV, typename CT>
__device__ __forceinline__ float2 pair_of(const int4& v, int j, float kv_scale) {
  if constexpr (KV == B200PA_KV_AUTO && std::is_same<CT, __nv_bfloat16>::value) {
    const unsigned w = (unsigned)word_of(v, j >> 1);
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
  } else if constexpr (KV == B200PA_KV_AUTO && std::is_same<CT, __half>::value) {
    const unsigned w = (unsigned)word_of(v, j >> 1);
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
  } else {
    return make_float2(CacheElem<T, KV>::val(elem_of<CT>(v, j), kv_scale), CacheElem<T, KV>::val(elem_of<CT>(v, j + 1), kv_scale));
  }
}

constexpr int kSimtWarps = 4;
constexpr int kSimtThreads = kSimtWarps * 32;
constexpr int kMaxHeadSize = 256;
constexpr int kAccPerLane = kMaxHeadSize / 32;

template <typename T, int KV, bool BLOCK_SPARSE>
__global__ void __launch_bounds__(kSimtThreads, 5)  // <= 102 registers: 5 CTAs (20 warps) per SM, no spills (6 CTAs spill)
    paged_attention_simt_kernel(const __grid_constant__ AttnKernelParams p) {
  using CT = typename CacheElem<T, KV>::type;
  constexpr int X = 16 / sizeof(CT);

  const int head = blockIdx.x;
  const int seq = blockIdx.y;
  const int part = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int D = p.head_size;
  const int B = p.block_size;
  const int seq_len = p.seq_lens[seq];
  const int t_begin = p.num_parts > 1 ? part * p.partition_size : 0;
  if (p.num_parts > 1 && t_begin >= seq_len) return;  // attention_kernels.cu:118-121
  const int t_end = p.num_parts > 1 ? min(seq_len, t_begin + p.partition_size) : seq_len;
  const int kv_head = head / (p.num_heads / p.num_kv_heads);
  const float alibi = p.alibi_slopes ? p.alibi_slopes[head] : 0.f;
  const int* block_table = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;

  __shared__ __align__(16) float q_s[kMaxHeadSize];
  __shared__ __align__(16) float p_s[kSimtWarps][32];
  __shared__ float red_m[kSimtWarps], red_l[kSimtWarps];
  __shared__ float red_acc[kSimtWarps][kMaxHeadSize];

  const T* q = static_cast<const T*>(p.q) + (int64_t)seq * p.q_stride + (int64_t)head * D;
  for (int d = threadIdx.x; d < D; d += kSimtThreads) q_s[d] = to_float<T>(q[d]);
  __syncthreads();

  BlockSparse bs;
  if constexpr (BLOCK_SPARSE) bs.init(p, seq_len, head, kv_head);

  const CT* kc = static_cast<const CT*>(p.k_cache) + (int64_t)kv_head * p.kv_head_stride;
  const CT* vc = static_cast<const CT*>(p.v_cache) + (int64_t)kv_head * p.kv_head_stride;

  float m = -FLT_MAX, l = 0.f;
  float acc[kAccPerLane];
#pragma unroll
  for (int k = 0; k < kAccPerLane; ++k) acc[k] = 0.f;

  for (int t0 = t_begin + warp * 32; t0 < t_end; t0 += kSimtThreads) {
    const int tok = t0 + lane;
    bool valid = tok < t_end;
    if constexpr (BLOCK_SPARSE) valid = valid && bs.attended(tok / B, B);
    float logit = -FLT_MAX;
    int64_t my_phys = 0;  // physical block of this lane's token (the P.V phase takes it by shuffle)
    if (valid) {
      my_phys = block_table[tok / B];
      const CT* kp = kc + my_phys * p.kv_block_stride + (tok % B) * X;
      // 16-byte K groups, four loads in flight before their products; q comes from shared memory as float4
      // (broadcast), and the products go to two accumulators by turns so that the FMA chain is half as long
      float dot0 = 0.f, dot1 = 0.f;
      const int ndh = D / X;
      for (int dh0 = 0; dh0 < ndh; dh0 += 4) {
        int4 kv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (dh0 + i < ndh) kv[i] = *reinterpret_cast<const int4*>(kp + (int64_t)(dh0 + i) * B * X);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (dh0 + i < ndh) {
            const float4* q4 = reinterpret_cast<const float4*>(q_s + (dh0 + i) * X);
#pragma unroll
            for (int j4 = 0; j4 < X / 4; ++j4) {
              const float4 qv = q4[j4];
              const float2 k01 = pair_of<T, KV, CT>(kv[i], 4 * j4, p.kv_scale);
              const float2 k23 = pair_of<T, KV, CT>(kv[i], 4 * j4 + 2, p.kv_scale);
              dot0 = fmaf(qv.x, k01.x, dot0);
              dot1 = fmaf(qv.y, k01.y, dot1);
              dot0 = fmaf(qv.z, k23.x, dot0);
              dot1 = fmaf(qv.w, k23.y, dot1);
            }
          }
        }
      }
      logit = (dot0 + dot1) * p.scale + alibi * (float)(tok - seq_len + 1);
    }
    const float m_new = fmaxf(m, warp_max(logit));
    const float alpha = __expf(m - m_new);
    const float pj = valid ? __expf(logit - m_new) : 0.f;
    l = l * alpha + warp_sum(pj);
    m = m_new;
#pragma unroll
    for (int k = 0; k < kAccPerLane; ++k) acc[k] *= alpha;
    p_s[warp][lane] = pj;
    __syncwarp();
    // P.V for the (up to 32/B) KV blocks this token group spans.
    const int t_stop = min(t0 + 32, t_end);
    for (int tb = t0; tb < t_stop; tb += B) {
      const int blk = tb / B;
      if constexpr (BLOCK_SPARSE) {
        if (!bs.attended(blk, B)) continue;
      }
      // the block's first token belongs to lane tb - t0, which is valid (and attended) here
      const int64_t phys = __shfl_sync(0xffffffffu, my_phys, tb - t0);
      const int n_here = min(B, t_stop - tb);  // never uses tokens >= seq_len (stale cache bits)
      const CT* vb = vc + phys * p.kv_block_stride;
      const float* pp = &p_s[warp][tb - t0];
      if (B % X == 0) {
        // a row's B tokens are whole 16-byte groups: one 128-bit load per (row, group), all of the lane's rows in
        // flight before their products; p as float4 (tb - t0 is a multiple of B).  Tokens >= n_here are loaded (the
        // bytes exist) but never used.
        for (int j0 = 0; j0 < n_here; j0 += X) {
          int4 rv[kAccPerLane];
#pragma unroll
          for (int k = 0; k < kAccPerLane; ++k)
            if (lane + 32 * k < D) rv[k] = *reinterpret_cast<const int4*>(vb + (int64_t)(lane + 32 * k) * B + j0);
          float4 pv[X / 4];
#pragma unroll
          for (int j4 = 0; j4 < X / 4; ++j4) pv[j4] = *reinterpret_cast<const float4*>(pp + j0 + 4 * j4);
          const int left = n_here - j0;
          if (left >= X) {
#pragma unroll
            for (int k = 0; k < kAccPerLane; ++k) {
              if (lane + 32 * k < D) {
                float a = acc[k];
#pragma unroll
                for (int j4 = 0; j4 < X / 4; ++j4) {
                  const float2 v01 = pair_of<T, KV, CT>(rv[k], 4 * j4, p.kv_scale);
                  const float2 v23 = pair_of<T, KV, CT>(rv[k], 4 * j4 + 2, p.kv_scale);
                  a = fmaf(pv[j4].x, v01.x, a);
                  a = fmaf(pv[j4].y, v01.y, a);
                  a = fmaf(pv[j4].z, v23.x, a);
                  a = fmaf(pv[j4].w, v23.y, a);
                }
                acc[k] = a;
              }
            }
          } else {  // the sequence ends inside this group
#pragma unroll
            for (int k = 0; k < kAccPerLane; ++k) {
              if (lane + 32 * k < D) {
                float a = acc[k];
#pragma unroll
                for (int j4 = 0; j4 < X / 4; ++j4) {
                  const int l4 = left - 4 * j4;
                  if (l4 > 0) a = fmaf(pv[j4].x, CacheElem<T, KV>::val(elem_of<CT>(rv[k], 4 * j4 + 0), p.kv_scale), a);
                  if (l4 > 1) a = fmaf(pv[j4].y, CacheElem<T, KV>::val(elem_of<CT>(rv[k], 4 * j4 + 1), p.kv_scale), a);
                  if (l4 > 2) a = fmaf(pv[j4].z, CacheElem<T, KV>::val(elem_of<CT>(rv[k], 4 * j4 + 2), p.kv_scale), a);
                  if (l4 > 3) a = fmaf(pv[j4].w, CacheElem<T, KV>::val(elem_of<CT>(rv[k], 4 * j4 + 3), p.kv_scale), a);
                }
                acc[k] = a;
              }
            }
          }
        }
      } else {  // rows shorter than 16 bytes (fp8 cache with block size 8): element loads
#pragma unroll
        for (int k = 0; k < kAccPerLane; ++k) {
          const int d = lane + 32 * k;
          if (d < D) {
            const CT* row = vb + (int64_t)d * B;
            float a = acc[k];
            for (int j = 0; j < n_here; ++j) a = fmaf(pp[j], CacheElem<T, KV>::load(row + j, p.kv_scale), a);
            acc[k] = a;
          }
        }
      }
    }
    __syncwarp();
  }

  // Merge the warps' partial softmax states.
  if (lane == 0) {
    red_m[warp] = m;
    red_l[warp] = l;
  }
#pragma unroll
  for (int k = 0; k < kAccPerLane; ++k) {
    const int d = lane + 32 * k;
    if (d < D) red_acc[warp][d] = acc[k];
  }
  __syncthreads();
  float m_all = red_m[0];
#pragma unroll
  for (int w = 1; w < kSimtWarps; ++w) m_all = fmaxf(m_all, red_m[w]);
  float f[kSimtWarps];
  float l_all = 0.f;
#pragma unroll
  for (int w = 0; w < kSimtWarps; ++w) {
    f[w] = __expf(red_m[w] - m_all);
    l_all += red_l[w] * f[w];
  }
  const float inv = __fdividef(1.f, l_all + 1e-6f);  // attention_kernels.cu:343
  const int64_t row = (int64_t)seq * p.num_heads + head;
  T* dst = p.num_parts > 1 ? static_cast<T*>(p.tmp_out) + (row * p.scratch_parts + part) * D
                           : static_cast<T*>(p.out) + row * D;
  for (int d = threadIdx.x; d < D; d += kSimtThreads) {
    float o = 0.f;
#pragma unroll
    for (int w = 0; w < kSimtWarps; ++w) o += red_acc[w][d] * f[w];
    dst[d] = from_float<T>(o * inv);
  }
  if (p.num_parts > 1 && threadIdx.x == 0) {
    p.max_logits[row * p.scratch_parts + part] = m_all;
    p.exp_sums[row * p.scratch_parts + part] = l_all;
  }
}

// Split-K reduce (reference: paged_attention_v2_reduce_kernel, attention_kernels.cu:565-669):
//   w_i = exp_sums_i * exp(max_i - max); out = sum_i tmp_out_i * w_i / (sum_i w_i + 1e-6)
template <typename T>
__global__ void __launch_bounds__(128) paged_attention_reduce_kernel(const __grid_constant__ AttnKernelParams p) {
  extern __shared__ float w_s[];  // [num_parts]
  __shared__ float red[4];
  const int head = blockIdx.x;
  const int seq = blockIdx.y;
  const int D = p.head_size;
  const int seq_len = p.seq_lens[seq];
  const int n = min(p.num_parts, (seq_len + p.partition_size - 1) / p.partition_size);
  const int64_t row = (int64_t)seq * p.num_heads + head;
  const float* ml = p.max_logits + row * p.scratch_parts;
  const float* es = p.exp_sums + row * p.scratch_parts;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  float mx = -FLT_MAX;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = ml[i];
    w_s[i] = v;
    mx = fmaxf(mx, v);
  }
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float sum = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float w = es[i] * expf(w_s[i] - mx);
    w_s[i] = w;
    sum += w;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = red[0] + red[1] + red[2] + red[3];
  const float inv = __fdividef(1.f, sum + 1e-6f);
  const T* tmp = static_cast<const T*>(p.tmp_out) + row * p.scratch_parts * D;
  T* out = static_cast<T*>(p.out) + row * D;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    for (int i = 0; i < n; ++i) a += to_float<T>(tmp[(int64_t)i * D + d]) * w_s[i];
    out[d] = from_float<T>(a * inv);
  }
}

template <typename T, int KV>
static int launch_simt_t(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  dim3 grid(kp.num_heads, num_seqs, kp.num_parts);
  if (kp.bs_vert_stride > 1)
    paged_attention_simt_kernel<T, KV, true><<<grid, kSimtThreads, 0, stream>>>(kp);
  else
    paged_attention_simt_kernel<T, KV, false><<<grid, kSimtThreads, 0, stream>>>(kp);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

template <int KV>
static int launch_simt_kv(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream) {
  switch (dtype) {
    case B200PA_F16: return launch_simt_t<__half, KV>(kp, num_seqs, stream);
    case B200PA_BF16: return launch_simt_t<__nv_bfloat16, KV>(kp, num_seqs, stream);
    case B200PA_F32: return launch_simt_t<float, KV>(kp, num_seqs, stream);
    default: return B200PA_ERR_INPUT_DTYPE;
  }
}

int launch_attention_simt(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream) {
  switch (kv_dtype) {
    case B200PA_KV_AUTO: return launch_simt_kv<B200PA_KV_AUTO>(kp, num_seqs, dtype, stream);
    case B200PA_KV_FP8_E4M3: return launch_simt_kv<B200PA_KV_FP8_E4M3>(kp, num_seqs, dtype, stream);
    case B200PA_KV_FP8_E5M2: return launch_simt_kv<B200PA_KV_FP8_E5M2>(kp, num_seqs, dtype, stream);
    default: return B200PA_ERR_KV_DTYPE;
  }
}

int launch_attention_reduce(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream) {
  dim3 grid(kp.num_heads, num_seqs);
  const size_t smem = sizeof(float) * kp.num_parts;
  switch (dtype) {
    case B200PA_F16: paged_attention_reduce_kernel<__half><<<grid, 128, smem, stream>>>(kp); break;
    case B200PA_BF16: paged_attention_reduce_kernel<__nv_bfloat16><<<grid, 128, smem, stream>>>(kp); break;
    case B200PA_F32: paged_attention_reduce_kernel<float><<<grid, 128, smem, stream>>>(kp); break;
    default: return B200PA_ERR_INPUT_DTYPE;
  }
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

}  // namespace b200pa
