// context_attention_fwd for B200 (sm_100a): prefill attention of new tokens over (cached prefix in the
// paged KV cache) + (the new tokens themselves, causal).  C-ABI entry b200pa_context_attention_fwd.
//
// Replaces the reference's Triton kernels `_fwd_kernel` / `_fwd_kernel_alibi` and their host wrapper
// `context_attention_fwd` (vllm/attention/ops/prefix_prefill.py:10-245, 442-681, 684-817), reached through
// PagedAttention.forward_prefix (vllm/attention/ops/paged_attn.py:185-216).
//
// Design (not the reference's):
//   * one CTA = 64 query rows of one (sequence, query head); 4 warps x 16 rows; grid x = head so the
//     CTAs that share a KV head (GQA group) are adjacent in launch order and hit the same K/V tiles in L2;
//     heavy (late, long-causal) row tiles are scheduled first;
//   * K/V tiles of BN tokens are staged by 16-byte cp.async into a double-buffered shared-memory ring and
//     RE-LAID-OUT on the way: the paged K chunk [D/8][B][8] becomes token-major rows, the paged V chunk
//     [D][B] becomes d-major rows, both with a 16-byte pad per row so every ldmatrix phase touches 8
//     distinct bank groups.  The copy is block-size agnostic (any power of two >= 8);
//   * QK^T and PV run on tensor cores (mma.sync.m16n8k16, fp32 accumulate); the S accumulator fragments
//     are re-used in registers as the A operand of PV (no shared-memory round trip); V fragments come from
//     ldmatrix (cached prefix, d-major) or ldmatrix.trans (new tokens, token-major);
//   * online softmax in the log2 domain, finite -1e30 sentinel instead of -inf/-10000 (a fully masked
//     tile contributes exactly 0 once a real logit is seen; every stored row sees its own diagonal);
//   * KV tiles that the sliding window masks for all 64 rows are skipped, not computed and discarded.
#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "prefill_common.cuh"

namespace b200pa {

namespace prefill {

constexpr int kThreads = 128;
constexpr int kBM = 64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// 16-byte async copy global -> shared; src_bytes == 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (std::is_same<T, __half>::value) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}
template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (std::is_same<T, __half>::value) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  } else {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DP, int BN, int MT>
struct Geometry {
  static constexpr int BM = 64 * MT;           // query rows per CTA (4 warps x MT m16 tiles)
  static constexpr int KPITCH = (DP + 8) * 2;  // bytes per token-major row (Q, K, new V)
  static constexpr int VPITCH = (BN + 8) * 2;  // bytes per d-major row (cached V)
  static constexpr int QBYTES = BM * KPITCH;
  static constexpr int KBYTES = BN * KPITCH;
  static constexpr int VBYTES = (BN * KPITCH > DP * VPITCH) ? BN * KPITCH : DP * VPITCH;
  static constexpr int STAGE = KBYTES + VBYTES;
  static constexpr int SMEM = QBYTES + 2 * STAGE;
};

// T: half | bf16 (query, new k/v, 16-bit cache, output).  DP: head size padded to a multiple of 16
// (columns >= head_size are zero-filled).  BN: KV tokens per tile.  MT: m16 row tiles per warp — with
// MT = 2 every K/V fragment read from shared memory feeds two MMAs and every staged tile serves 128 rows.
template <typename T, int DP, int BN, int MT>
__global__ void __launch_bounds__(kThreads, (DP <= 128 && MT == 1) ? 3 : (DP <= 128 ? 2 : 1))
    context_attention_kernel(const __grid_constant__ PrefillParams p) {
  using G = Geometry<DP, BN, MT>;
  constexpr int BM = G::BM;
  constexpr int KS = DP / 16;  // k-steps of QK^T
  constexpr int NT = BN / 8;   // 8-token n-tiles of S
  constexpr int ND = DP / 8;   // 8-column n-tiles of O
  constexpr int KK = BN / 16;  // k-steps of PV
  constexpr int DG = DP / 8;   // 16-byte d-groups per token row
  static_assert(DP % 16 == 0 && BN % 16 == 0 && kThreads % BN == 0, "tile geometry");

  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int gid = lane >> 2, tig = lane & 3;
  const int mi = lane >> 3, mj = lane & 7;

  const int head = blockIdx.x;
  const int mt_idx = gridDim.y - 1 - blockIdx.y;  // heavy row tiles first
  const int seq = blockIdx.z;
  const int ctx_len = p.ctx_lens[seq];
  const int q_len = p.seq_lens[seq] - ctx_len;
  const int m0 = mt_idx * BM;
  if (m0 >= q_len) return;
  const int64_t tok0 = p.start_loc[seq];
  const int kv_head = head / (p.num_heads / p.num_kv_heads);
  const int D = p.head_size;
  const int W = p.alibi_slopes ? 0 : p.sliding_window;  // prefix_prefill.py:717-760: the ALiBi kernel has no window
  const int bshift = p.block_shift;
  const int B = 1 << bshift;

  // ---- tile ranges ---------------------------------------------------------------------------------
  // cached-prefix tiles [c_first, c_end), then new-token tiles [j_first, j_end) (causal upper bound)
  int c_first = 0, j_first = 0;
  if (W > 0) {
    c_first = max(0, ctx_len + m0 - W + 1) / BN;  // keys below ctx_len + m0 - W + 1 are masked for every row
    j_first = max(0, m0 - W + 1) / BN;
  }
  const int c_end = (ctx_len + BN - 1) / BN;
  const int j_end = (min(q_len, m0 + BM) + BN - 1) / BN;
  const int n_ctx = max(0, c_end - c_first);
  const int n_tiles = n_ctx + (j_end - j_first);

  const T* kc = static_cast<const T*>(p.k_cache) + (int64_t)kv_head * p.kc_head_stride;
  const T* vc = static_cast<const T*>(p.v_cache) + (int64_t)kv_head * p.vc_head_stride;
  const T* kn = static_cast<const T*>(p.k) + tok0 * p.k_stride_t + (int64_t)kv_head * p.k_stride_h;
  const T* vn = static_cast<const T*>(p.v) + tok0 * p.v_stride_t + (int64_t)kv_head * p.v_stride_h;
  const int* table = p.block_tables + (int64_t)seq * p.block_table_stride;
  const uint32_t q_smem = smem_u32(smem);
  const uint32_t kv_smem = q_smem + G::QBYTES;

  // ---- stage one KV tile (all threads) -----------------------------------------------------------
  auto issue = [&](int it, int buf) {
    const uint32_t kb = kv_smem + buf * G::STAGE;
    const uint32_t vb = kb + G::KBYTES;
    if (it < n_ctx && p.kv_dtype != B200PA_KV_AUTO) {
      // fp8 cache (x = 16): K chunk [D/16][B][16] bytes, V chunk [D][B] bytes.  Same destinations as below, but
      // through registers: 128-bit / 64-bit loads, fp8 -> T (x kv_scale, the decode kernels' conversion), 128-bit
      // shared stores.  Plain stores: the __syncthreads that publishes the cp.async tiles publishes these too.
      const int t0 = (c_first + it) * BN;
      const uint8_t* kc8 = static_cast<const uint8_t*>(p.k_cache) + (int64_t)kv_head * p.kc_head_stride;
      const uint8_t* vc8 = static_cast<const uint8_t*>(p.v_cache) + (int64_t)kv_head * p.vc_head_stride;
      {
        const int t = tid % BN;
        const int kpos = t0 + t;
        const bool tv = kpos < ctx_len;
        const int64_t phys = tv ? table[kpos >> bshift] : 0;
        const uint8_t* src = kc8 + phys * p.kc_block_stride + (kpos & (B - 1)) * 16;
        const uint32_t dst = kb + t * G::KPITCH;
#pragma unroll
        for (int c = tid / BN; c < DP / 16; c += kThreads / BN) {
          const bool ok = tv && c * 16 < D;
          int4 raw = make_int4(0, 0, 0, 0);
          if (ok) raw = ld_nc_v4(src + (int64_t)c * B * 16);
          const uint4 zero = make_uint4(0, 0, 0, 0);
          st_shared_v4(dst + c * 32, ok ? fp8x8_to_t<T>(p.kv_dtype, make_uint2(raw.x, raw.y), p.kv_scale) : zero);
          st_shared_v4(dst + c * 32 + 16, ok ? fp8x8_to_t<T>(p.kv_dtype, make_uint2(raw.z, raw.w), p.kv_scale) : zero);
        }
      }
#pragma unroll 4
      for (int u = tid; u < DP * BN / 8; u += kThreads) {
        const int d = u / (BN / 8);
        const int tt = (u - d * (BN / 8)) * 8;
        const int kpos = t0 + tt;
        const bool ok = kpos < ctx_len && d < D;
        uint4 w = make_uint4(0, 0, 0, 0);
        if (ok) {
          const int64_t phys = table[kpos >> bshift];
          w = fp8x8_to_t<T>(p.kv_dtype, ld_nc_v2(vc8 + phys * p.vc_block_stride + (int64_t)d * B + (kpos & (B - 1))), p.kv_scale);
        }
        st_shared_v4(vb + d * G::VPITCH + tt * 2, w);
      }
    } else if (it < n_ctx) {
      const int t0 = (c_first + it) * BN;
      {  // K: unit (d-group g, token t) of the paged chunk [D/8][B][8] -> row t, column group g
        const int t = tid % BN;
        const int kpos = t0 + t;
        const bool tv = kpos < ctx_len;
        const int64_t phys = tv ? table[kpos >> bshift] : 0;
        const T* src = kc + phys * p.kc_block_stride + (kpos & (B - 1)) * 8;
        const uint32_t dst = kb + t * G::KPITCH;
#pragma unroll
        for (int g = tid / BN; g < DG; g += kThreads / BN) {
          const bool ok = tv && g * 8 < D;
          cp_async16(dst + g * 16, ok ? (const void*)(src + (int64_t)g * B * 8) : (const void*)kc, ok ? 16 : 0);
        }
      }
      {  // V: 8 consecutive tokens of one d row of the paged chunk [D][B] -> d-major row
        const int rshift = min(bshift, (BN == 64 ? 6 : 5));  // tokens per contiguous run R = min(B, BN)
        const int R = 1 << rshift;
        const int upr_shift = rshift - 3;                 // 16-byte units per (d row, run)
        const int units_per_run = DP << upr_shift;
#pragma unroll 4
        for (int u = tid; u < DP * BN / 8; u += kThreads) {
          const int run = u / units_per_run;
          const int rem = u - run * units_per_run;
          const int d = rem >> upr_shift;
          const int tt = run * R + ((rem & ((1 << upr_shift) - 1)) << 3);  // token inside the tile
          const int kpos = t0 + tt;
          const bool ok = kpos < ctx_len && d < D;
          const int64_t phys = ok ? table[kpos >> bshift] : 0;
          const T* src = vc + phys * p.vc_block_stride + (int64_t)d * B + (kpos & (B - 1));
          cp_async16(vb + d * G::VPITCH + tt * 2, ok ? (const void*)src : (const void*)vc, ok ? 16 : 0);
        }
      }
    } else {
      const int t0 = (j_first + it - n_ctx) * BN;
#pragma unroll 4
      for (int u = tid; u < BN * DG; u += kThreads) {
        const int r = u / DG, c = u - r * DG;
        const int j = t0 + r;
        const bool ok = j < q_len && c * 8 < D;
        const uint32_t off = r * G::KPITCH + c * 16;
        cp_async16(kb + off, ok ? (const void*)(kn + (int64_t)j * p.k_stride_t + c * 8) : (const void*)kn, ok ? 16 : 0);
        cp_async16(vb + off, ok ? (const void*)(vn + (int64_t)j * p.v_stride_t + c * 8) : (const void*)vn, ok ? 16 : 0);
      }
    }
  };

  // ---- Q tile -> shared (rows >= q_len and columns >= head_size zero-filled), with the first KV tile ----
  {
    const T* qb = static_cast<const T*>(p.q) + tok0 * p.q_stride_t + (int64_t)head * p.q_stride_h;
    const bool q16 = (reinterpret_cast<uintptr_t>(p.q) % 16 == 0) && (p.q_stride_t % 8 == 0) && (p.q_stride_h % 8 == 0);
    if (q16) {
#pragma unroll 4
      for (int u = tid; u < BM * DG; u += kThreads) {
        const int r = u / DG, c = u - r * DG;
        const bool ok = m0 + r < q_len && c * 8 < D;
        cp_async16(q_smem + r * G::KPITCH + c * 16, ok ? (const void*)(qb + (int64_t)(m0 + r) * p.q_stride_t + c * 8) : (const void*)qb,
                   ok ? 16 : 0);
      }
    } else {  // 4-byte aligned query rows: plain 32-bit copies
      for (int u = tid; u < BM * (DP / 2); u += kThreads) {
        const int r = u / (DP / 2), c = u - r * (DP / 2);
        const bool ok = m0 + r < q_len && c * 2 < D;
        const uint32_t w = ok ? *reinterpret_cast<const uint32_t*>(qb + (int64_t)(m0 + r) * p.q_stride_t + c * 2) : 0u;
        *reinterpret_cast<uint32_t*>(smem + r * G::KPITCH + c * 4) = w;
      }
    }
  }
  issue(0, 0);
  cp_async_commit();

  const float slope_l2 = p.alibi_slopes ? p.alibi_slopes[head] * kLog2e : 0.f;
  const bool has_alibi = p.alibi_slopes != nullptr;
  const int wrow0 = m0 + warp * (16 * MT);  // first query row of this warp

  float m_run[MT][2], l_run[MT][2];
  float oacc[MT][ND][4];
#pragma unroll
  for (int i = 0; i < MT; ++i) {
    m_run[i][0] = m_run[i][1] = kNegBig;
    l_run[i][0] = l_run[i][1] = 0.f;
#pragma unroll
    for (int nd = 0; nd < ND; ++nd)
#pragma unroll
      for (int e = 0; e < 4; ++e) oacc[i][nd][e] = 0.f;
  }

  // per-lane ldmatrix offsets
  const uint32_t q_lane = q_smem + (warp * (16 * MT) + (mi & 1) * 8 + mj) * G::KPITCH + (mi >> 1) * 16;  // A: (row half, k half)
  const uint32_t k_lane = ((mi >> 1) * 8 + mj) * G::KPITCH + (mi & 1) * 16;   // (n-tile pair, d half)
  const uint32_t vc_lane = ((mi >> 1) * 8 + mj) * G::VPITCH + (mi & 1) * 16;  // d-major: (d tile pair, token half)
  const uint32_t vn_lane = ((mi & 1) * 8 + mj) * G::KPITCH + (mi >> 1) * 16;  // token-major, .trans

  for (int it = 0; it < n_tiles; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_tiles) {
      issue(it + 1, buf ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t kb = kv_smem + buf * G::STAGE;
    const uint32_t vb = kb + G::KBYTES;
    const bool is_ctx = it < n_ctx;
    const int t0 = is_ctx ? (c_first + it) * BN : (j_first + it - n_ctx) * BN;

    if (is_ctx && t0 + BN > ctx_len) {
      // last prefix tile: zero the V columns of tokens >= ctx_len (stale cache bits there could be NaN,
      // and 0 * NaN would poison the row; the reference masks the load, prefix_prefill.py:163-167)
      const int nvalid = ctx_len - t0, cnt = BN - nvalid;
      T* vrow = reinterpret_cast<T*>(smem + G::QBYTES + buf * G::STAGE + G::KBYTES);
      for (int idx = tid; idx < DP * cnt; idx += kThreads) {
        const int d = idx / cnt, c = nvalid + (idx - d * cnt);
        vrow[d * (G::VPITCH / 2) + c] = from_float<T>(0.f);
      }
      __syncthreads();
    }

    // ================= S = Q K^T =================
    float sacc[MT][NT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) sacc[i][nt][e] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t qa[MT][4];
#pragma unroll
      for (int i = 0; i < MT; ++i) ldsm_x4(q_lane + i * 16 * G::KPITCH + ks * 32, qa[i][0], qa[i][1], qa[i][2], qa[i][3]);
#pragma unroll
      for (int np = 0; np < NT / 2; ++np) {
        uint32_t r0, r1, r2, r3;
        ldsm_x4(kb + np * 16 * G::KPITCH + ks * 32 + k_lane, r0, r1, r2, r3);
#pragma unroll
        for (int i = 0; i < MT; ++i) {
          mma16816<T>(sacc[i][2 * np], qa[i], r0, r1);
          mma16816<T>(sacc[i][2 * np + 1], qa[i], r2, r3);
        }
      }
    }

    // ================= mask + online softmax =================
    // Only tiles that a mask edge crosses pay for the per-element tests: the last prefix tile, the causal
    // diagonal, and tiles the sliding-window edge passes through (all CTA-uniform conditions).
    const int base_rel = is_ctx ? (t0 - ctx_len) : t0;    // rel = base_rel + col - m  (key pos - query pos)
    const int col_limit = is_ctx ? (ctx_len - t0) : BN;   // columns >= col_limit are past the cached prefix
    bool need_mask;
    if (is_ctx)
      need_mask = (t0 + BN > ctx_len) || (W > 0 && ctx_len + m0 + BM - 1 - t0 >= W);
    else
      need_mask = (t0 + BN - 1 > m0) || (W > 0 && m0 + BM - 1 - t0 >= W);

    uint32_t pa[MT][KK][4];
    bool rescale = false;
    float alpha[MT][2];
#pragma unroll
    for (int i = 0; i < MT; ++i) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int m = wrow0 + 16 * i + gid + 8 * h;
        float mx = kNegBig;
        if (need_mask || has_alibi) {
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int col = 8 * nt + 2 * tig + e;
              const int rel = base_rel + col - m;
              float z = sacc[i][nt][2 * h + e] * p.scale_log2;
              if (has_alibi) z = fmaf(slope_l2, (float)rel, z);
              const bool vis = !need_mask || (col < col_limit && rel <= 0 && (W <= 0 || -rel < W));
              z = vis ? z : kNegBig;
              sacc[i][nt][2 * h + e] = z;
              mx = fmaxf(mx, z);
            }
          }
        } else {
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const float z0 = sacc[i][nt][2 * h] * p.scale_log2, z1 = sacc[i][nt][2 * h + 1] * p.scale_log2;
            sacc[i][nt][2 * h] = z0;
            sacc[i][nt][2 * h + 1] = z1;
            mx = fmaxf(mx, fmaxf(z0, z1));
          }
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run[i][h], mx);
        alpha[i][h] = ex2(m_run[i][h] - m_new);
        rescale |= alpha[i][h] != 1.f;
        m_run[i][h] = m_new;
        float sum = 0.f;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          const float p0 = ex2(sacc[i][nt][2 * h] - m_new), p1 = ex2(sacc[i][nt][2 * h + 1] - m_new);
          sum += p0 + p1;
          // A fragment of PV: k-step nt/2; a0/a1 (rows gid/gid+8, k 2tig..) from even nt, a2/a3 from odd nt
          pa[i][nt >> 1][(nt & 1) * 2 + h] = pack2<T>(p0, p1);
        }
        l_run[i][h] = l_run[i][h] * alpha[i][h] + sum;  // per-lane partial; quad-reduced at the end
      }
    }
    if (__any_sync(0xffffffffu, rescale)) {
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int nd = 0; nd < ND; ++nd) {
          oacc[i][nd][0] *= alpha[i][0];
          oacc[i][nd][1] *= alpha[i][0];
          oacc[i][nd][2] *= alpha[i][1];
          oacc[i][nd][3] *= alpha[i][1];
        }
    }

    // ================= O += P V =================
    if (is_ctx) {
#pragma unroll
      for (int kk = 0; kk < KK; ++kk)
#pragma unroll
        for (int dp = 0; dp < ND / 2; ++dp) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4(vb + dp * 16 * G::VPITCH + kk * 32 + vc_lane, r0, r1, r2, r3);
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            mma16816<T>(oacc[i][2 * dp], pa[i][kk], r0, r1);
            mma16816<T>(oacc[i][2 * dp + 1], pa[i][kk], r2, r3);
          }
        }
    } else {
#pragma unroll
      for (int kk = 0; kk < KK; ++kk)
#pragma unroll
        for (int dp = 0; dp < ND / 2; ++dp) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4_t(vb + kk * 16 * G::KPITCH + dp * 32 + vn_lane, r0, r1, r2, r3);
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            mma16816<T>(oacc[i][2 * dp], pa[i][kk], r0, r1);
            mma16816<T>(oacc[i][2 * dp + 1], pa[i][kk], r2, r3);
          }
        }
    }
    __syncthreads();  // everyone is done with `buf` before the next iteration refills it
  }

  // ---- epilogue ---------------------------------------------------------------------------------------
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float l = l_run[i][h];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const int m = wrow0 + 16 * i + gid + 8 * h;
      if (m >= q_len) continue;
      const float inv = 1.f / l;  // l > 0: every stored row sees at least its own key
      T* orow = static_cast<T*>(p.out) + (tok0 + m) * p.o_stride_t + (int64_t)head * p.o_stride_h;
#pragma unroll
      for (int nd = 0; nd < ND; ++nd) {
        const int d = 8 * nd + 2 * tig;
        if (d < D) *reinterpret_cast<uint32_t*>(orow + d) = pack2<T>(oacc[i][nd][2 * h] * inv, oacc[i][nd][2 * h + 1] * inv);
      }
    }
}

template <typename T, int DP, int BN, int MT>
static int launch(const PrefillParams& kp, int batch, int max_query_len, cudaStream_t stream) {
  using G = Geometry<DP, BN, MT>;
  auto kernel = context_attention_kernel<T, DP, BN, MT>;
  static bool configured[64] = {false};
  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid(kp.num_heads, (max_query_len + G::BM - 1) / G::BM, batch);
  kernel<<<grid, kThreads, G::SMEM, stream>>>(kp);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

template <typename T, int DP>
static int launch_m(const PrefillParams& kp, int batch, int max_query_len, cudaStream_t stream) {
  // 128-row CTAs (two m16 tiles per warp) once the queries are long enough to fill them
  static const int force_mt = [] { const char* s = std::getenv("B200PA_PREFILL_MT"); return s ? std::atoi(s) : 0; }();
  const bool wide = force_mt ? force_mt == 2 : max_query_len > 64;
  if (wide) return launch<T, DP, 64, 2>(kp, batch, max_query_len, stream);
  return launch<T, DP, 64, 1>(kp, batch, max_query_len, stream);
}

template <typename T>
static int launch_t(const PrefillParams& kp, int batch, int max_query_len, cudaStream_t stream) {
  const int d = kp.head_size;
  if (d <= 32) return launch_m<T, 32>(kp, batch, max_query_len, stream);
  if (d <= 64) return launch_m<T, 64>(kp, batch, max_query_len, stream);
  if (d <= 96) return launch_m<T, 96>(kp, batch, max_query_len, stream);
  if (d <= 128) return launch_m<T, 128>(kp, batch, max_query_len, stream);
  if (d <= 192) return launch<T, 192, 32, 1>(kp, batch, max_query_len, stream);
  return launch<T, 256, 32, 1>(kp, batch, max_query_len, stream);
}

}  // namespace prefill
}  // namespace b200pa

static constexpr bool kTcDefault = true;

extern "C" int b200pa_context_attention_fwd(const b200pa_prefill_params_t* a, b200pa_stream_t stream) {
  using namespace b200pa;
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->dtype != B200PA_F16 && a->dtype != B200PA_BF16) return B200PA_ERR_INPUT_DTYPE;
  const bool fp8 = a->kv_dtype == B200PA_KV_FP8_E4M3 || a->kv_dtype == B200PA_KV_FP8_E5M2;
  if (a->kv_dtype != B200PA_KV_AUTO && !fp8) return B200PA_ERR_KV_DTYPE;
  if (a->head_size <= 0 || a->head_size > 256 || a->head_size % 8 != 0) return B200PA_ERR_HEAD_SIZE;
  if (fp8 && a->head_size % 16 != 0) return B200PA_ERR_HEAD_SIZE;  // x = 16 bytes of one token per K run
  int bshift = 0;
  while ((1 << bshift) < a->block_size) ++bshift;
  if (a->block_size < 8 || (1 << bshift) != a->block_size) return B200PA_ERR_BLOCK_SIZE;
  if (a->batch < 0 || a->max_query_len < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 ||
      a->num_heads % a->num_kv_heads != 0)
    return B200PA_ERR_INVALID_ARG;
  if (a->batch == 0 || a->max_query_len == 0) return B200PA_OK;
  if (!a->out || !a->query || !a->key || !a->value || !a->key_cache || !a->value_cache || !a->block_tables ||
      !a->start_loc || !a->seq_lens || !a->ctx_lens)
    return B200PA_ERR_INVALID_ARG;
  auto mis = [](const void* ptr, int n) { return reinterpret_cast<uintptr_t>(ptr) % n != 0; };
  // 16-byte cp.async rows for k / v / caches, 32-bit accesses for q / out
  if (mis(a->key, 16) || mis(a->value, 16) || mis(a->key_cache, 16) || mis(a->value_cache, 16) || mis(a->query, 4) ||
      mis(a->out, 4) || a->k_stride_t % 8 || a->k_stride_h % 8 || a->v_stride_t % 8 || a->v_stride_h % 8 ||
      a->kc_block_stride % (fp8 ? 16 : 8) || a->kc_head_stride % (fp8 ? 16 : 8) || a->vc_block_stride % 8 ||
      a->vc_head_stride % 8 ||
      a->q_stride_t % 2 || a->q_stride_h % 2 || a->o_stride_t % 2 || a->o_stride_h % 2)
    return B200PA_ERR_INVALID_ARG;

  PrefillParams kp{};
  kp.out = a->out;
  kp.q = a->query;
  kp.k = a->key;
  kp.v = a->value;
  kp.k_cache = a->key_cache;
  kp.v_cache = a->value_cache;
  kp.block_tables = a->block_tables;
  kp.start_loc = a->start_loc;
  kp.seq_lens = a->seq_lens;
  kp.ctx_lens = a->ctx_lens;
  kp.alibi_slopes = a->alibi_slopes;
  kp.q_stride_t = a->q_stride_t;
  kp.q_stride_h = a->q_stride_h;
  kp.k_stride_t = a->k_stride_t;
  kp.k_stride_h = a->k_stride_h;
  kp.v_stride_t = a->v_stride_t;
  kp.v_stride_h = a->v_stride_h;
  kp.o_stride_t = a->o_stride_t;
  kp.o_stride_h = a->o_stride_h;
  kp.kc_block_stride = a->kc_block_stride;
  kp.kc_head_stride = a->kc_head_stride;
  kp.vc_block_stride = a->vc_block_stride;
  kp.vc_head_stride = a->vc_head_stride;
  kp.block_table_stride = a->block_table_stride;
  kp.num_heads = a->num_heads;
  kp.num_kv_heads = a->num_kv_heads;
  kp.head_size = a->head_size;
  kp.block_shift = bshift;
  kp.sliding_window = a->sliding_window > 0 ? a->sliding_window : 0;
  kp.scale_log2 = a->scale * kLog2e;
  kp.kv_dtype = a->kv_dtype;
  kp.kv_scale = fp8 ? a->kv_scale : 1.0f;
  // tcgen05 + TMEM family first (128-row CTAs; head sizes <= 128 that are multiples of 16); the mma.sync family
  // takes short queries and every other shape.  B200PA_PREFILL_TC=0/1 forces the choice (tuning / tests).
  static const int tc_knob = [] { const char* e = std::getenv("B200PA_PREFILL_TC"); return e ? std::atoi(e) : -1; }();
  const bool want_tc = tc_knob >= 0 ? tc_knob != 0 : kTcDefault && a->max_query_len > 64;
  if (want_tc) {
    const int rc = launch_prefill_tc(kp, a->batch, a->max_query_len, a->dtype, stream);
    if (rc != 1) return rc;
  }
  if (a->dtype == B200PA_F16) return prefill::launch_t<__half>(kp, a->batch, a->max_query_len, stream);
  return prefill::launch_t<__nv_bfloat16>(kp, a->batch, a->max_query_len, stream);
}
