// Paged-attention decode kernel for B200 (sm_100a): TMA bulk-copy staged KV blocks + mma.sync
// tensor-core contractions over the whole GQA group + online softmax.
//
// Work decomposition (B200-first, not the reference's):
//   * one CTA per (KV head [x tile of 8/16 query heads], sequence, split) — the KV bytes of a
//     KV head are read from HBM ONCE for all query heads that share it (the reference launches
//     one CTA per query head and re-reads them Hq/Hkv times, attention_kernels.cu:734-766);
//   * 4 warps, each an independent pipeline: warp w owns KV blocks w, w+4, ... of the CTA's
//     range, its own ring of `num_stages` shared-memory stages and its own mbarriers.  Lane 0
//     follows the block table and issues two `cp.async.bulk` (TMA) copies per block — the K chunk
//     [D/x][B][x] and the V chunk [D][B] of one (block, KV head) are each one contiguous run of
//     B*D*sizeof(cache_t) bytes in the reference's layout — completing on the stage's mbarrier.
//     No CTA-wide barrier in the main loop;
//   * QK^T: S[16 q rows x 8 tokens] += Q[16 x 16] K^T via mma.sync.m16n8k16 (Q fragments live in
//     registers for the whole kernel; K fragments come from `ldmatrix` straight out of the
//     reference's K layout, whose 16-byte x-groups are exactly ldmatrix rows);
//   * online softmax in registers (quad shuffles), running max/sum per query row;
//   * P.V computed transposed: O^T[16 d rows x 8 q] += V[16 x 16 tokens] P^T, so the S accumulator
//     fragment is re-used directly as the B operand (no shuffle / shared-memory round trip) and the
//     token-minor V layout is read with conflict-free 8/16-byte shared loads;
//   * token order inside a block is permuted consistently between the two contractions
//     (`tokmap`) so that both the ldmatrix rows and the V loads are bank-conflict free;
//   * fp8 KV: bytes are converted fp8->fp16 in registers (exact), both contractions run in fp16,
//     kv_scale is folded into the logit scale and the output scale;
//   * the four warps' partial (max, sum, O) are merged through shared memory at the end; split
//     launches write (tmp_out, max_logits, exp_sums) in the reference's v2 scratch format.
//
// Math parity with the reference kernel (attention_kernels.cu:86-496): fp32 accumulation of
// QK, logit = scale*qk + alibi*(j - len + 1), P rounded to the 16-bit type before P.V
// (attention_kernels.cu:398-400), out = sum(p v) / (sum(p) + 1e-6), tokens >= seq_len masked in
// both contractions (V lanes zeroed like :420-430).
#pragma once

#include <type_traits>

#include "attention_common.cuh"

namespace b200pa {

constexpr int kMmaWarps = 4;
constexpr int kMmaThreads = kMmaWarps * 32;
constexpr int kMmaMaxParts = kMmaMaxSplits;  // most context splits per (sequence, KV head): the launch plan grows the split beyond

// 16-/32-bit load that bypasses L1 (data written by other SMs during this launch)
__device__ __forceinline__ float ldcg_t(const float* p) { return __ldcg(p); }
__device__ __forceinline__ __half ldcg_t(const __half* p) {
  const unsigned short v = __ldcg(reinterpret_cast<const unsigned short*>(p));
  return __ushort_as_half(v);
}
__device__ __forceinline__ __nv_bfloat16 ldcg_t(const __nv_bfloat16* p) {
  const unsigned short v = __ldcg(reinterpret_cast<const unsigned short*>(p));
  return __ushort_as_bfloat16(v);
}

// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_evict_normal_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar,
                                             uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

template <typename MT>
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                  uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// m16n8k8 (block size 8: one P.V k-step is the block's 8 tokens)
template <typename MT>
__device__ __forceinline__ void mma_1688(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0);
template <>
__device__ __forceinline__ void mma_1688<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(b0));
}
template <>
__device__ __forceinline__ void mma_1688<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(b0));
}

template <typename MT>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// four fp8 bytes -> two packed half2 (bytes 0,1 -> lo; bytes 2,3 -> hi), exact.  Written in PTX so that
// ptxas folds the 16-bit half selection into the convert (SASS: F2FP.F16.E4M3.UNPACK_B Rd, Rs.H0/.H1)
// instead of the byte shuffles the cuda_fp8.h struct round trip produces.
template <int KV>
__device__ __forceinline__ void fp8x4_to_half2x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  if constexpr (KV == B200PA_KV_FP8_E5M2) {
    asm("{\n.reg .b16 l, h;\nmov.b32 {l, h}, %2;\ncvt.rn.f16x2.e5m2x2 %0, l;\ncvt.rn.f16x2.e5m2x2 %1, h;\n}"
        : "=r"(lo), "=r"(hi)
        : "r"(w));
  } else {
    asm("{\n.reg .b16 l, h;\nmov.b32 {l, h}, %2;\ncvt.rn.f16x2.e4m3x2 %0, l;\ncvt.rn.f16x2.e4m3x2 %1, h;\n}"
        : "=r"(lo), "=r"(hi)
        : "r"(w));
  }
}

// 2^x, one MUFU.EX2 (inputs here are <= 0 or the finite kNegBig sentinel; ftz is harmless)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T>
__device__ __forceinline__ float2 unpack2(uint32_t w);
template <>
__device__ __forceinline__ float2 unpack2<__half>(uint32_t w) {
  return __half22float2(*reinterpret_cast<__half2*>(&w));
}
template <>
__device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
}

// Per-warp shared-memory region: max(ring of stages, merge scratch), 128-byte granular.
__host__ __device__ constexpr int mma_merge_floats(int D, int HPT) { return 32 + HPT * (D + 4); }
__host__ __device__ constexpr int mma_region_bytes(int stage_bytes, int stages, int D, int HPT) {
  int a = stage_bytes * stages, b = mma_merge_floats(D, HPT) * 4;
  int m = a > b ? a : b;
  return (m + 127) / 128 * 128;
}

// ---- shared epilogue of the tensor-core decode kernels (paged layout and flash layout) ---------------------
// On entry every warp has left its partial state for the CTA's HPT query heads in its own shared-memory region
// (floats): [0,16) running maxima (log2 domain), [16,32) sums, then O as [HPT][D + 4].  Merges the four warps,
// writes the output row (one split) or the split's partial in the reference's v2 scratch format — and, for split
// launches, lets the LAST split of this (sequence, KV head) to finish do the split-K reduce.
template <typename T, int D, int HPT>
__device__ __forceinline__ void decode_epilogue(const AttnKernelParams& p, uint8_t* smem, int region, int seq,
                                                int seq_len, int kv_head, int qt, int G, int part, float out_scale) {
  constexpr int OROW = D + 4;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  __syncthreads();

  for (int idx = threadIdx.x; idx < HPT * D; idx += kMmaThreads) {
    const int g = idx / D;
    const int d = idx - g * D;
    const int hr = qt * HPT + g;
    if (hr >= G) break;  // g is non-decreasing in idx
    float mw[kMmaWarps], m_all = kNegBig;
#pragma unroll
    for (int w = 0; w < kMmaWarps; ++w) {
      mw[w] = reinterpret_cast<const float*>(smem + w * region)[g];
      m_all = fmaxf(m_all, mw[w]);
    }
    float l_all = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < kMmaWarps; ++w) {
      const float* mow = reinterpret_cast<const float*>(smem + w * region);
      const float f = exp2f(mw[w] - m_all);
      l_all = fmaf(mow[16 + g], f, l_all);
      o = fmaf(mow[32 + g * OROW + d], f, o);
    }
    const float res = o * out_scale * __fdividef(1.f, l_all + 1e-6f);
    const int64_t row = (int64_t)seq * p.num_heads + kv_head * G + hr;
    if (p.num_parts > 1) {
      static_cast<T*>(p.tmp_out)[(row * p.scratch_parts + part) * D + d] = from_float<T>(res);
      if (d == 0) {
        p.max_logits[row * p.scratch_parts + part] = m_all * kLn2;
        p.exp_sums[row * p.scratch_parts + part] = l_all;
      }
    } else {
      static_cast<T*>(p.out)[row * D + d] = from_float<T>(res);
    }
  }
  if (p.num_parts == 1) return;

  // ---- split-K reduce, fused: the LAST split of this (sequence, KV head) to finish merges them all ------
  // (reference: a second kernel, paged_attention_v2_reduce_kernel, attention_kernels.cu:565-669; same formula:
  //  w_i = exp_sums_i * exp(max_i - max); out = sum_i tmp_out_i * w_i / (sum_i w_i + 1e-6)).  The arrival
  // counter lives in the per-device workspace and is reset by the split that uses it, so every launch starts
  // from zero; the merge order is the split order, whoever arrives last — results do not depend on timing.
  const int n_parts = min(p.num_parts, (seq_len + p.partition_size - 1) / p.partition_size);  // non-empty splits
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  unsigned* counter = p.sk_counters + ((int64_t)seq * gridDim.x + blockIdx.x);
  if (threadIdx.x == 0) s_last = atomicAdd(counter, 1u) == (unsigned)(n_parts - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) *counter = 0u;
  // weights per (head of this CTA, split) into shared memory: warp w takes heads w, w + 4, ...
  float* w_s = reinterpret_cast<float*>(smem);  // [HPT][n_parts], n_parts <= kMmaMaxParts
  for (int g = warp; g < HPT; g += kMmaWarps) {
    const int hr = qt * HPT + g;
    if (hr >= G) break;
    const int64_t row = (int64_t)seq * p.num_heads + kv_head * G + hr;
    const float* ml = p.max_logits + row * p.scratch_parts;
    const float* es = p.exp_sums + row * p.scratch_parts;
    float mv[kMmaMaxParts / 32], ev[kMmaMaxParts / 32], mx = -3.0e38f;
#pragma unroll
    for (int k = 0; k < kMmaMaxParts / 32; ++k) {
      const int i = lane + 32 * k;
      mv[k] = i < n_parts ? __ldcg(ml + i) : -3.0e38f;
      ev[k] = i < n_parts ? __ldcg(es + i) : 0.f;
      mx = fmaxf(mx, mv[k]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < kMmaMaxParts / 32; ++k) {
      ev[k] *= expf(mv[k] - mx);
      sum += ev[k];
    }
    sum = warp_sum(sum);
    const float inv = __fdividef(1.f, sum + 1e-6f);
#pragma unroll
    for (int k = 0; k < kMmaMaxParts / 32; ++k) {
      const int i = lane + 32 * k;
      if (i < n_parts) w_s[g * n_parts + i] = ev[k] * inv;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < HPT * D; idx += kMmaThreads) {
    const int g = idx / D;
    const int d = idx - g * D;
    const int hr = qt * HPT + g;
    if (hr >= G) break;
    const int64_t row = (int64_t)seq * p.num_heads + kv_head * G + hr;
    const T* tmp = static_cast<const T*>(p.tmp_out) + row * p.scratch_parts * D + d;
    const float* w = w_s + g * n_parts;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;  // four loads in flight
    int i = 0;
    for (; i + 4 <= n_parts; i += 4) {
      const T t0 = ldcg_t(tmp + (int64_t)i * D), t1 = ldcg_t(tmp + (int64_t)(i + 1) * D);
      const T t2 = ldcg_t(tmp + (int64_t)(i + 2) * D), t3 = ldcg_t(tmp + (int64_t)(i + 3) * D);
      a0 = fmaf(to_float<T>(t0), w[i], a0);
      a1 = fmaf(to_float<T>(t1), w[i + 1], a1);
      a2 = fmaf(to_float<T>(t2), w[i + 2], a2);
      a3 = fmaf(to_float<T>(t3), w[i + 3], a3);
    }
    for (; i < n_parts; ++i) a0 = fmaf(to_float<T>(ldcg_t(tmp + (int64_t)i * D)), w[i], a0);
    static_cast<T*>(p.out)[row * D + d] = from_float<T>((a0 + a1) + (a2 + a3));
  }
}

// T: query/out type (half | bf16).  KV: b200pa_kv_dtype_t.  D: head size.  B: block size (16 | 32; 8 for 16-bit caches:
// one 8-token n-tile per block, one ldmatrix.x4 = two k-steps of QK^T, P.V with m16n8k8 whose B operand is the S fragment).
// NG: 8-row query-head tiles per CTA (1: group <= 8, 2: group <= 16 per CTA).
// BS: block-sparse masks (attention_kernels.cu:212-256, 387-393).  A warp then walks only the KV blocks that at least
// one head of the CTA's group attends: consumer and TMA producer step through the same filtered sequence (`nxt`), the
// ring's stage / phase count attended blocks, and where the heads' offsets differ (head_sliding_step != 0) a row that
// does not attend a block gets p = 0 for it.  The dense instantiations (BS = false) compile to the same code as before.
template <typename T, int KV, int D, int B, int NG, bool BS = false>
__global__ void __launch_bounds__(kMmaThreads, (D * NG <= 256) ? 2 : 1)
    paged_attention_mma_kernel(const __grid_constant__ AttnKernelParams p) {
  using MT = typename std::conditional<KV == B200PA_KV_AUTO, T, __half>::type;  // tensor-core input type
  constexpr bool FP8 = KV != B200PA_KV_AUTO;
  constexpr int CTB = FP8 ? 1 : 2;         // bytes per cache element
  constexpr int NT = B / 8;                // 8-token n-tiles of the QK product
  constexpr int TPT = B / 4;               // tokens each lane of a quad owns
  constexpr int NP = B / 8;                // token pairs per lane
  constexpr int KS = D / 16;               // k-steps of QK == 16-row d-tiles of PV
  constexpr int KK = B / 16;               // k-steps of PV
  constexpr int UNIT = B * 16;             // bytes between consecutive 16-byte d-groups of K
  constexpr int CHUNK = B * D * CTB;       // bytes of one K (or V) chunk
  constexpr int STAGE = 2 * CHUNK;
  constexpr int HPT = 8 * NG;              // query heads per CTA
  constexpr int OROW = D + 4;              // padded row of the merge scratch (floats)
  static_assert(B == 16 || B == 32 || (B == 8 && !FP8), "block size");
  static_assert(D % 16 == 0, "head size");

  extern __shared__ __align__(128) uint8_t smem[];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int gid = lane >> 2;  // mma "groupID": fragment row
  const int tig = lane & 3;   // thread in quad

  const int G = p.num_heads / p.num_kv_heads;
  const int q_tiles = (G + HPT - 1) / HPT;
  const int kv_head = blockIdx.x / q_tiles;
  const int qt = blockIdx.x - kv_head * q_tiles;
  const int seq = blockIdx.y;
  const int part = blockIdx.z;
  const int seq_len = p.seq_lens[seq];
  const int NST = p.num_stages;

  const int t_begin = p.num_parts > 1 ? part * p.partition_size : 0;
  if (p.num_parts > 1 && t_begin >= seq_len) {
    // an empty split does not take part in the merge; for an empty context split 0 writes the zeros
    if (part == 0) {
      const int G0 = p.num_heads / p.num_kv_heads;
      const int HPT0 = 8 * NG;
      const int qt0 = blockIdx.x % ((G0 + HPT0 - 1) / HPT0), kvh0 = blockIdx.x / ((G0 + HPT0 - 1) / HPT0);
      for (int idx = threadIdx.x; idx < HPT0 * D; idx += kMmaThreads) {
        const int hr = qt0 * HPT0 + idx / D;
        if (hr < G0)
          static_cast<T*>(p.out)[((int64_t)seq * p.num_heads + kvh0 * G0 + hr) * D + idx % D] = from_float<T>(0.f);
      }
    }
    return;
  }
  const int t_end = p.num_parts > 1 ? min(seq_len, t_begin + p.partition_size) : seq_len;
  const int blk_begin = t_begin / B;
  const int nblk = (t_end - t_begin + B - 1) / B;
  const int my_n = warp < nblk ? (nblk - warp + kMmaWarps - 1) / kMmaWarps : 0;

  const int region = mma_region_bytes(STAGE, NST, D, HPT);
  uint8_t* my_region = smem + warp * region;
  const uint32_t stage_base = smem_u32(my_region);
  const uint32_t bar_base = smem_u32(smem + kMmaWarps * region) + warp * NST * 8;

  const int* block_table = p.block_tables + (int64_t)seq * p.max_blocks_per_seq + blk_begin + warp;
  const char* kbase = static_cast<const char*>(p.k_cache) + (int64_t)kv_head * p.kv_head_stride * CTB;
  const char* vbase = static_cast<const char*>(p.v_cache) + (int64_t)kv_head * p.kv_head_stride * CTB;
  const int64_t blk_stride_bytes = p.kv_block_stride * CTB;

  // ---- block-sparse: which of this warp's blocks does any head of the group attend? ---------------
  BlockSparse bs_lane;     // lane l < HPT: head l of the CTA's q-tile (the any-head test is a warp vote)
  BlockSparse bs_row[NG];  // the heads of this lane's S-fragment rows
  bool lane_head = false;
  if constexpr (BS) {
    const int hl = qt * HPT + lane;
    lane_head = lane < HPT && hl < G;
    bs_lane.init(p, seq_len, kv_head * G + (lane_head ? hl : 0), kv_head);
#pragma unroll
    for (int r = 0; r < NG; ++r) bs_row[r].init(p, seq_len, kv_head * G + min(qt * HPT + gid + 8 * r, G - 1), kv_head);
  }
  // first ordinal >= i of this warp's blocks that is attended (warp-uniform; identity when dense)
  auto nxt = [&](int i) -> int {
    if constexpr (BS) {
      while (i < my_n) {
        const bool a = lane_head && bs_lane.attended(blk_begin + warp + i * kMmaWarps, B);
        if (__any_sync(0xffffffffu, a)) break;
        ++i;
      }
    }
    return i;
  };

  // ---- pipeline prologue: barriers + first NST blocks in flight --------------------------------
  uint64_t policy = 0;
  int next_phys = 0;
  int pi = 0;  // block-sparse: the producer's cursor into the filtered sequence (warp-uniform)
  if (lane == 0) {
    for (int s = 0; s < NST; ++s) mbar_init(bar_base + s * 8, 1);
    fence_proxy_async_smem();
    policy = (p.flags & 1) ? l2_evict_normal_policy() : l2_evict_first_policy();
    if constexpr (!BS) {
      const int n0 = min(NST, my_n);
      for (int i = 0; i < n0; ++i) {
        const int64_t phys = block_table[i * kMmaWarps];
        const uint32_t bar = bar_base + i * 8, dst = stage_base + i * STAGE;
        mbar_arrive_expect_tx(bar, STAGE);
        tma_bulk_g2s(dst, kbase + phys * blk_stride_bytes, CHUNK, bar, policy);
        tma_bulk_g2s(dst + CHUNK, vbase + phys * blk_stride_bytes, CHUNK, bar, policy);
      }
      if (NST < my_n) next_phys = block_table[NST * kMmaWarps];
    }
  }
  __syncwarp();
  if constexpr (BS) {
    for (int s = 0; s < NST; ++s) {
      pi = nxt(pi);
      if (pi >= my_n) break;
      if (lane == 0) {
        const int64_t phys = block_table[pi * kMmaWarps];
        const uint32_t bar = bar_base + s * 8, dst = stage_base + s * STAGE;
        mbar_arrive_expect_tx(bar, STAGE);
        tma_bulk_g2s(dst, kbase + phys * blk_stride_bytes, CHUNK, bar, policy);
        tma_bulk_g2s(dst + CHUNK, vbase + phys * blk_stride_bytes, CHUNK, bar, policy);
      }
      ++pi;
    }
    __syncwarp();
  }

  // ---- Q fragments (A operand of QK^T), resident for the whole kernel --------------------------
  // 16-bit cache: k index of step ks <-> d = 16ks + {2tig, 2tig+1 | +8};
  // fp8 cache   : k index of step ks <-> d = 16ks + {4tig, 4tig+1 | 4tig+2, 4tig+3}
  // (the ldmatrix row of an fp8 K group holds 16 d-values, 4 per lane).
  uint32_t qa[KS][2 * NG];
  {
    const T* qbase = static_cast<const T*>(p.q) + (int64_t)seq * p.q_stride;
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      const int hr = qt * HPT + gid + 8 * r;  // head index inside the GQA group
      const bool valid = hr < G;
      const T* qrow = qbase + (int64_t)(kv_head * G + (valid ? hr : 0)) * D;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const int d0 = FP8 ? 16 * ks + 4 * tig : 16 * ks + 2 * tig;
        const int d1 = FP8 ? d0 + 2 : d0 + 8;
        uint32_t w0 = valid ? *reinterpret_cast<const uint32_t*>(qrow + d0) : 0u;
        uint32_t w1 = valid ? *reinterpret_cast<const uint32_t*>(qrow + d1) : 0u;
        if constexpr (FP8) {  // q -> fp16, pre-multiplied by the softmax scale
          const float2 f0 = unpack2<T>(w0), f1 = unpack2<T>(w1);
          w0 = pack2<__half>(f0.x * p.scale, f0.y * p.scale);
          w1 = pack2<__half>(f1.x * p.scale, f1.y * p.scale);
        }
        qa[ks][r] = w0;           // a0 (r=0) / a1 (r=1)
        qa[ks][NG + r] = w1;      // a2 (r=0) / a3 (r=1)
      }
    }
  }
  const float z_mult = (FP8 ? p.kv_scale : p.scale) * kLog2e;  // S_raw -> log2-domain logit
  float alibi_l2[NG];
#pragma unroll
  for (int r = 0; r < NG; ++r) {
    const int hr = qt * HPT + gid + 8 * r;
    alibi_l2[r] = (p.alibi_slopes && hr < G) ? p.alibi_slopes[kv_head * G + hr] * kLog2e : 0.f;
  }
  const bool has_alibi = p.alibi_slopes != nullptr;

  // ---- per-lane constant offsets ----------------------------------------------------------------
  // tokmap(nt, j): token (inside the block) that row j of n-tile nt stands for.  Chosen so that
  //  (a) lane (gid, tig) ends up owning tokens TPT*tig .. TPT*tig+TPT-1 in its S fragments, which is
  //      what one 8/16-byte shared load of a V row gives it, and
  //  (b) the 8 rows of every ldmatrix matrix are distinct modulo 8 tokens => 8 distinct 16-byte
  //      bank groups => conflict-free.
  auto rot = [](int q) { return q * NP / 4; };
  auto tokmap = [&](int nt, int j) { return TPT * (j >> 1) + 2 * ((nt + rot(j >> 1)) % NP) + (j & 1); };
  const int mi = lane >> 3, mj = lane & 7;
  uint32_t koff[NT / 2 > 0 ? (FP8 ? 2 : NT / 2) : 1];
  if constexpr (B == 8) {
    // one call per PAIR of k-steps: matrix mi = d-group 4kp + mi of the block's single n-tile
    koff[0] = mi * UNIT + tokmap(0, mj) * 16;
  } else if constexpr (!FP8) {
    // call c of k-step ks: matrices (d-group 2ks + (mi&1), n-tile 2c + (mi>>1))
#pragma unroll
    for (int c = 0; c < NT / 2; ++c) koff[c] = (mi & 1) * UNIT + tokmap(2 * c + (mi >> 1), mj) * 16;
  } else if constexpr (B == 16) {
    // one call per PAIR of k-steps: matrices (d-group 2kp + (mi>>1), n-tile mi&1)
    koff[0] = (mi >> 1) * UNIT + tokmap(mi & 1, mj) * 16;
    koff[1] = tokmap(mi & 1, mj) * 16;  // odd tail: both halves read d-group KS-1
  } else {
    // one call per k-step: matrices (d-group ks, n-tile mi)
    koff[0] = tokmap(mi, mj) * 16;
    koff[1] = 0;
  }
  const uint32_t voff = gid * (B * CTB) + tig * (TPT * CTB);
  int tpos[NT];  // first token of this lane's S fragment pair in n-tile nt
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) tpos[nt] = TPT * tig + 2 * ((nt + rot(tig)) % NP);

  // ---- running state ------------------------------------------------------------------------------
  float m_run[NG], l_run[NG];
  float oacc[NG][KS][4];
#pragma unroll
  for (int r = 0; r < NG; ++r) {
    m_run[r] = kNegBig;
    l_run[r] = 0.f;
#pragma unroll
    for (int dt = 0; dt < KS; ++dt)
#pragma unroll
      for (int e = 0; e < 4; ++e) oacc[r][dt][e] = 0.f;
  }

  int stage = 0;
  uint32_t phase = 0;
  for (int i = nxt(0); i < my_n; i = nxt(i + 1)) {
    const uint32_t bar = bar_base + stage * 8;
    const uint32_t kb = stage_base + stage * STAGE;
    const uint32_t vb = kb + CHUNK;
    const int blk_tok0 = (blk_begin + warp + i * kMmaWarps) * B;  // first token of this block
    const int n_valid = t_end - blk_tok0;                         // >= 1; < B only in the last block
    bool att[NG];  // block-sparse: does the head of fragment row r attend this block?
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      att[r] = true;
      if constexpr (BS) att[r] = bs_row[r].attended(blk_begin + warp + i * kMmaWarps, B);
    }
    mbar_wait(bar, phase);

    // ================= S = Q K^T =================
    float sacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) sacc[nt][e] = 0.f;

    if constexpr (B == 8) {
#pragma unroll
      for (int kp = 0; kp < (KS + 1) / 2; ++kp) {
        // (for an odd KS the last call also reads two d-groups behind the K chunk — the V chunk, same stage — unused)
        uint32_t r0, r1, r2, r3;
        ldmatrix_x4(kb + 4 * kp * UNIT + koff[0], r0, r1, r2, r3);
        if constexpr (NG == 2) {
          mma_16816<MT>(sacc[0], qa[2 * kp][0], qa[2 * kp][1], qa[2 * kp][2], qa[2 * kp][3], r0, r1);
          if (2 * kp + 1 < KS)
            mma_16816<MT>(sacc[0], qa[2 * kp + 1][0], qa[2 * kp + 1][1], qa[2 * kp + 1][2], qa[2 * kp + 1][3], r2, r3);
        } else {
          mma_16816<MT>(sacc[0], qa[2 * kp][0], 0u, qa[2 * kp][1], 0u, r0, r1);
          if (2 * kp + 1 < KS) mma_16816<MT>(sacc[0], qa[2 * kp + 1][0], 0u, qa[2 * kp + 1][1], 0u, r2, r3);
        }
      }
    } else if constexpr (!FP8) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
        for (int c = 0; c < NT / 2; ++c) {
          uint32_t r0, r1, r2, r3;
          ldmatrix_x4(kb + ks * 2 * UNIT + koff[c], r0, r1, r2, r3);
          if constexpr (NG == 2) {
            mma_16816<MT>(sacc[2 * c], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], r0, r1);
            mma_16816<MT>(sacc[2 * c + 1], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], r2, r3);
          } else {
            mma_16816<MT>(sacc[2 * c], qa[ks][0], 0u, qa[ks][1], 0u, r0, r1);
            mma_16816<MT>(sacc[2 * c + 1], qa[ks][0], 0u, qa[ks][1], 0u, r2, r3);
          }
        }
      }
    } else if constexpr (B == 16) {
#pragma unroll
      for (int kp = 0; kp < (KS + 1) / 2; ++kp) {
        uint32_t r[4];
        const bool tail = (2 * kp + 1 >= KS);
        ldmatrix_x4(kb + 2 * kp * UNIT + (tail ? koff[1] : koff[0]), r[0], r[1], r[2], r[3]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ks = 2 * kp + h;
          if (ks < KS) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
              uint32_t b0, b1;
              fp8x4_to_half2x2<KV>(r[2 * h + nt], b0, b1);
              if constexpr (NG == 2)
                mma_16816<MT>(sacc[nt], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
              else
                mma_16816<MT>(sacc[nt], qa[ks][0], 0u, qa[ks][1], 0u, b0, b1);
            }
          }
        }
      }
    } else {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t r[4];
        ldmatrix_x4(kb + ks * UNIT + koff[0], r[0], r[1], r[2], r[3]);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          uint32_t b0, b1;
          fp8x4_to_half2x2<KV>(r[nt], b0, b1);
          if constexpr (NG == 2)
            mma_16816<MT>(sacc[nt], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], b0, b1);
          else
            mma_16816<MT>(sacc[nt], qa[ks][0], 0u, qa[ks][1], 0u, b0, b1);
        }
      }
    }

    // ================= online softmax =================
    uint32_t pk[NG][NT];  // P (rounded to the 16-bit MMA type), packed token pairs
    float alpha[NG];
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      float z[NT][2];
      float mx = kNegBig;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          float v = sacc[nt][2 * r + e] * z_mult;
          const int tin = tpos[nt] + e;
          if (has_alibi) v = fmaf(alibi_l2[r], (float)(blk_tok0 + tin - seq_len + 1), v);
          if (tin >= n_valid) v = kNegBig;
          if constexpr (BS) {
            if (!att[r]) v = kNegBig;  // (the running max must not move for a row that skips this block)
          }
          z[nt][e] = v;
          mx = fmaxf(mx, v);
        }
      }
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[r], mx);
      alpha[r] = fast_exp2(m_run[r] - m_new);
      m_run[r] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float p0 = fast_exp2(z[nt][0] - m_new), p1 = fast_exp2(z[nt][1] - m_new);
        if constexpr (BS) {
          if (!att[r]) p0 = p1 = 0.f;  // exp2(-big - (-big)) would be 1 while the row has seen no token yet
        }
        sum += p0 + p1;
        pk[r][nt] = pack2<MT>(p0, p1);
      }
      l_run[r] = l_run[r] * alpha[r] + sum;
    }

    // P pairs in token order: pair q of this lane's tokens sits in n-tile (q - rot(tig)) mod NP.
    uint32_t pp[NG][NP];
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      if constexpr (NP == 1) {
        pp[r][0] = pk[r][0];
      } else if constexpr (NP == 2) {
        const bool sw = (tig >> 1) != 0;
        pp[r][0] = sw ? pk[r][1] : pk[r][0];
        pp[r][1] = sw ? pk[r][0] : pk[r][1];
      } else {
        uint32_t t[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) t[q] = (tig & 1) ? pk[r][(q + 3) & 3] : pk[r][q];
#pragma unroll
        for (int q = 0; q < 4; ++q) pp[r][q] = (tig & 2) ? t[(q + 2) & 3] : t[q];
      }
    }

    // rescale the running output when some row's max moved (warp-uniform test)
    {
      bool moved = false;
#pragma unroll
      for (int r = 0; r < NG; ++r) moved |= (alpha[r] != 1.f);
      if (__any_sync(0xffffffffu, moved)) {
#pragma unroll
        for (int r = 0; r < NG; ++r) {
          // O^T columns of this lane are query rows 2tig, 2tig+1, owned by quads 2tig, 2tig+1
          const float a0 = __shfl_sync(0xffffffffu, alpha[r], (2 * tig) * 4);
          const float a1 = __shfl_sync(0xffffffffu, alpha[r], (2 * tig + 1) * 4);
#pragma unroll
          for (int dt = 0; dt < KS; ++dt) {
            oacc[r][dt][0] *= a0;
            oacc[r][dt][1] *= a1;
            oacc[r][dt][2] *= a0;
            oacc[r][dt][3] *= a1;
          }
        }
      }
    }

    // ================= O^T += V P^T =================
    uint32_t vmask[NP];
    const bool partial = n_valid < B;  // warp-uniform
    if (partial) {
#pragma unroll
      for (int q = 0; q < NP; ++q) {
        const int t0 = TPT * tig + 2 * q;
        vmask[q] = (t0 < n_valid ? 0x0000ffffu : 0u) | (t0 + 1 < n_valid ? 0xffff0000u : 0u);
      }
    }
#pragma unroll
    for (int dt = 0; dt < KS; ++dt) {
      uint32_t va[2][NP];  // [row gid | gid+8][token pair]
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t addr = vb + (16 * dt + 8 * h) * (B * CTB) + voff;
        if constexpr (B == 8) {
          va[h][0] = lds32(addr);  // tokens 2 tig, 2 tig + 1 of row gid (+ 8 h): 128 contiguous bytes per warp
        } else if constexpr (!FP8 && B == 16) {
          const uint2 v = lds64(addr);
          va[h][0] = v.x; va[h][1] = v.y;
        } else if constexpr (!FP8 && B == 32) {
          const uint4 v = lds128(addr);
          va[h][0] = v.x; va[h][1] = v.y; va[h][2] = v.z; va[h][3] = v.w;
        } else if constexpr (FP8 && B == 16) {
          const uint32_t v = lds32(addr);
          fp8x4_to_half2x2<KV>(v, va[h][0], va[h][1]);
        } else {
          const uint2 v = lds64(addr);
          fp8x4_to_half2x2<KV>(v.x, va[h][0], va[h][1]);
          fp8x4_to_half2x2<KV>(v.y, va[h][2], va[h][3]);
        }
        if (partial) {
#pragma unroll
          for (int q = 0; q < NP; ++q) va[h][q] &= vmask[q];
        }
      }
      if constexpr (B == 8) {
#pragma unroll
        for (int r = 0; r < NG; ++r) mma_1688<MT>(oacc[r][dt], va[0][0], va[1][0], pp[r][0]);
      }
#pragma unroll
      for (int kk = 0; kk < KK; ++kk)
#pragma unroll
        for (int r = 0; r < NG; ++r)
          mma_16816<MT>(oacc[r][dt], va[0][2 * kk], va[1][2 * kk], va[0][2 * kk + 1], va[1][2 * kk + 1],
                        pp[r][2 * kk], pp[r][2 * kk + 1]);
    }

    // ================= refill this stage with block i + NST =================
    __syncwarp();
    if constexpr (BS) {
      pi = nxt(pi);  // the next attended block nobody has issued yet
      if (pi < my_n) {
        if (lane == 0) {
          fence_proxy_async_smem();
          const int64_t phys = block_table[pi * kMmaWarps];
          mbar_arrive_expect_tx(bar, STAGE);
          tma_bulk_g2s(kb, kbase + phys * blk_stride_bytes, CHUNK, bar, policy);
          tma_bulk_g2s(vb, vbase + phys * blk_stride_bytes, CHUNK, bar, policy);
        }
        ++pi;
      }
    } else if (lane == 0 && i + NST < my_n) {
      fence_proxy_async_smem();
      const int64_t phys = next_phys;
      mbar_arrive_expect_tx(bar, STAGE);
      tma_bulk_g2s(kb, kbase + phys * blk_stride_bytes, CHUNK, bar, policy);
      tma_bulk_g2s(vb, vbase + phys * blk_stride_bytes, CHUNK, bar, policy);
      if (i + NST + 1 < my_n) next_phys = block_table[(i + NST + 1) * kMmaWarps];
    }
    if (++stage == NST) {
      stage = 0;
      phase ^= 1;
    }
  }

  // ---- merge the four warps' partial states -------------------------------------------------------
  // (this warp's ring is idle now: every copy it issued has been waited for)
  float* mo = reinterpret_cast<float*>(my_region);
#pragma unroll
  for (int r = 0; r < NG; ++r) {
    float l = l_run[r];
    l += __shfl_xor_sync(0xffffffffu, l, 1);
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    if (tig == 0) {
      mo[8 * r + gid] = m_run[r];
      mo[16 + 8 * r + gid] = l;
    }
#pragma unroll
    for (int dt = 0; dt < KS; ++dt) {
      float* o0 = mo + 32 + (8 * r + 2 * tig) * OROW + 16 * dt + gid;
      o0[0] = oacc[r][dt][0];
      o0[OROW] = oacc[r][dt][1];
      o0[8] = oacc[r][dt][2];
      o0[OROW + 8] = oacc[r][dt][3];
    }
  }
  decode_epilogue<T, D, HPT>(p, smem, region, seq, seq_len, kv_head, qt, G, part, FP8 ? p.kv_scale : 1.f);
}

}  // namespace b200pa
