// Decode attention over the FLASH-ATTENTION cache layout — key_cache / value_cache
// [num_blocks, block_size, num_kv_heads, head_size], both token-major — for B200 (sm_100a).
//
// Reference: the FLASH_ATTN / FLASHINFER backends keep this layout (vllm/attention/backends/flash_attn.py:37-44,
// written by reshape_and_cache_flash, csrc/cache_kernels.cu:204-235) and decode it with the external
// flash_attn_with_kvcache (flash_attn.py:354-367).  This kernel is that call on the same engine as the paged-layout
// kernel (attention_mma_kernel.cuh): one CTA per (KV head [x tile of 8/16 query heads], sequence, split) so the KV
// bytes are read once for the whole GQA group, four warps each streaming blocks w, w+4, ... through their own ring,
// online softmax in registers, split-K with the reduce fused into the launch (decode_epilogue).
//
// What the layout changes: one (block, head) is block_size rows of head_size elements, token stride H*D — not one
// contiguous chunk — so the ring is filled with 16-byte cp.async (LDGSTS; half a warp reads one 256-byte row,
// L2 evict-first like the paged kernel's bulk copies) into rows padded by 16 bytes, which makes every ldmatrix
// phase conflict-free for all head sizes.  Both contractions are the plain m16n8k16 forms: S = Q K^T takes K rows
// with ldmatrix, O += P V takes the S accumulators as its A operand and V with ldmatrix.trans.
#include <algorithm>

#include "attention_mma_kernel.cuh"

namespace b200pa {

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, int src_bytes, uint64_t policy) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2, %3;" ::"r"(dst), "l"(src), "r"(src_bytes),
               "l"(policy)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                                  uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

__host__ __device__ constexpr int flash_row_bytes(int D) { return D * 2 + 16; }
__host__ __device__ constexpr int flash_stage_bytes(int D, int B) { return 2 * B * flash_row_bytes(D); }
constexpr int kFlashSmemPerSM = 227 * 1024;
constexpr int kFlashReserved = 1024;
__host__ __device__ constexpr int flash_smem_bytes(int D, int B, int NG, int NST) {
  return kMmaWarps * mma_region_bytes(flash_stage_bytes(D, B), NST, D, 8 * NG);
}
// CTAs per SM the kernel is compiled for: what two ring stages allow in shared memory, capped by the register file
// (three CTAs of 128 threads leave 168 registers per thread: enough up to D * NG = 128)
__host__ __device__ constexpr int flash_ctas(int D, int B, int NG) {
  return (flash_smem_bytes(D, B, NG, 2) <= kFlashSmemPerSM / 3 - kFlashReserved && D * NG <= 128) ? 3
         : (flash_smem_bytes(D, B, NG, 2) <= kFlashSmemPerSM / 2 - kFlashReserved && D * NG <= 256) ? 2 : 1;
}

// T: half | bf16.  D: head size (multiple of 16).  B: block size (16 | 32).  NG: 8-row query-head tiles per CTA.
// NST: ring depth per warp.
template <typename T, int D, int B, int NG, int NST>
__global__ void __launch_bounds__(kMmaThreads, flash_ctas(D, B, NG))
    paged_attention_flash_kernel(const __grid_constant__ AttnKernelParams p) {
  constexpr int NT = B / 8;            // 8-token n-tiles of the QK product
  constexpr int KS = D / 16;           // k-steps of QK
  constexpr int ND = D / 8;            // 8-column n-tiles of the PV product
  constexpr int KK = B / 16;           // k-steps of PV
  constexpr int ROW = flash_row_bytes(D);
  constexpr int TILE = B * ROW;        // one K (or V) tile in shared memory
  constexpr int STAGE = 2 * TILE;
  constexpr int HPT = 8 * NG;
  constexpr int OROW = D + 4;
  constexpr int CPR = D / 8;           // 16-byte chunks per row
  constexpr int CHUNKS = B * CPR;      // per tile
  constexpr uint32_t FULL = 0xffffffffu;
  static_assert(D % 16 == 0 && (B == 16 || B == 32), "shape");

  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gid = lane >> 2, tig = lane & 3;

  const int G = p.num_heads / p.num_kv_heads;
  const int q_tiles = (G + HPT - 1) / HPT;
  const int kv_head = blockIdx.x / q_tiles;
  const int qt = blockIdx.x - kv_head * q_tiles;
  const int seq = blockIdx.y;
  const int part = blockIdx.z;
  const int seq_len = p.seq_lens[seq];

  const int t_begin = p.num_parts > 1 ? part * p.partition_size : 0;
  if (p.num_parts > 1 && t_begin >= seq_len) {
    if (part == 0) {  // empty context: zeros (nobody else writes this row)
      for (int idx = threadIdx.x; idx < HPT * D; idx += kMmaThreads) {
        const int hr = qt * HPT + idx / D;
        if (hr < G) static_cast<T*>(p.out)[((int64_t)seq * p.num_heads + kv_head * G + hr) * D + idx % D] = from_float<T>(0.f);
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
  const int* block_table = p.block_tables + (int64_t)seq * p.max_blocks_per_seq + blk_begin + warp;
  const T* kbase = static_cast<const T*>(p.k_cache) + (int64_t)kv_head * p.kv_head_stride;
  const T* vbase = static_cast<const T*>(p.v_cache) + (int64_t)kv_head * p.kv_head_stride;
  const uint64_t policy = (p.flags & 1) ? l2_evict_normal_policy() : l2_evict_first_policy();

  // one block = rows (tokens) of D elements, token stride kv_token_stride; rows at or beyond t_end are zero-filled
  auto issue = [&](int i, int stage, int64_t phys) {
    const int tok0 = (blk_begin + warp + i * kMmaWarps) * B;
    const T* kb = kbase + phys * p.kv_block_stride;
    const T* vb = vbase + phys * p.kv_block_stride;
    const uint32_t dst = stage_base + stage * STAGE;
#pragma unroll
    for (int q = lane; q < CHUNKS; q += 32) {
      const int r = q / CPR, c = q - r * CPR;
      const int bytes = tok0 + r < t_end ? 16 : 0;
      const int64_t off = (int64_t)r * p.kv_token_stride + c * 8;
      cp_async_16(dst + r * ROW + c * 16, kb + off, bytes, policy);
      cp_async_16(dst + TILE + r * ROW + c * 16, vb + off, bytes, policy);
    }
  };
#pragma unroll
  for (int s = 0; s < NST; ++s) {
    if (s < my_n) issue(s, s, block_table[s * kMmaWarps]);
    cp_async_commit();
  }
  int next_phys = NST < my_n ? block_table[NST * kMmaWarps] : 0;  // table entry one refill ahead

  // ---- Q fragments (A operand of QK^T): rows = query heads of the group, k = d ------------------------------
  uint32_t qa[KS][2 * NG];
  float alibi_l2[NG];
  {
    const T* qbase = static_cast<const T*>(p.q) + (int64_t)seq * p.q_stride;
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      const int hr = qt * HPT + gid + 8 * r;
      const bool valid = hr < G;
      const T* qrow = qbase + (int64_t)(kv_head * G + (valid ? hr : 0)) * D;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        qa[ks][r] = valid ? *reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 2 * tig) : 0u;           // a0 / a1
        qa[ks][NG + r] = valid ? *reinterpret_cast<const uint32_t*>(qrow + 16 * ks + 2 * tig + 8) : 0u;  // a2 / a3
      }
      alibi_l2[r] = (p.alibi_slopes && valid) ? p.alibi_slopes[kv_head * G + hr] * kLog2e : 0.f;
    }
  }
  const float z_mult = p.scale * kLog2e;
  const bool has_alibi = p.alibi_slopes != nullptr;

  // ldmatrix lane addresses (bytes inside a tile)
  //   K (non-transposed, one call = 16 tokens x 16 d): matrix (lane>>3): token half (lane>>4), d half ((lane>>3)&1)
  const uint32_t k_lane = (uint32_t)(((lane & 7) + ((lane >> 4) << 3)) * ROW + (((lane >> 3) & 1) << 4));
  //   V (transposed, one call = 16 tokens x 16 d): matrix (lane>>3): token half ((lane>>3)&1), d half (lane>>4)
  const uint32_t v_lane = (uint32_t)(((lane & 7) + (((lane >> 3) & 1) << 3)) * ROW + ((lane >> 4) << 4));

  float m_run[NG], l_run[NG];
  float oacc[ND][4];  // [.][0,1]: row gid, [.][2,3]: row gid + 8 (NG == 2 only)
#pragma unroll
  for (int r = 0; r < NG; ++r) {
    m_run[r] = kNegBig;
    l_run[r] = 0.f;
  }
#pragma unroll
  for (int nd = 0; nd < ND; ++nd)
#pragma unroll
    for (int e = 0; e < 4; ++e) oacc[nd][e] = 0.f;

  int stage = 0;
  for (int i = 0; i < my_n; ++i) {
    cp_async_wait<NST - 1>();
    __syncwarp();
    const uint32_t kb = stage_base + stage * STAGE;
    const uint32_t vb = kb + TILE;
    const int blk_tok0 = (blk_begin + warp + i * kMmaWarps) * B;
    const int n_valid = t_end - blk_tok0;

    // ================= S = Q K^T =================
    float sacc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) sacc[nt][e] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int c = 0; c < NT / 2; ++c) {
        uint32_t r0, r1, r2, r3;
        ldmatrix_x4(kb + c * 16 * ROW + ks * 32 + k_lane, r0, r1, r2, r3);
        if constexpr (NG == 2) {
          mma_16816<T>(sacc[2 * c], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], r0, r1);
          mma_16816<T>(sacc[2 * c + 1], qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], r2, r3);
        } else {
          mma_16816<T>(sacc[2 * c], qa[ks][0], 0u, qa[ks][1], 0u, r0, r1);
          mma_16816<T>(sacc[2 * c + 1], qa[ks][0], 0u, qa[ks][1], 0u, r2, r3);
        }
      }
    }

    // ================= online softmax (row gid; row gid + 8 when NG == 2) =================
    uint32_t pk[NG][NT];
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
          const int tin = 8 * nt + 2 * tig + e;
          if (has_alibi) v = fmaf(alibi_l2[r], (float)(blk_tok0 + tin - seq_len + 1), v);
          if (tin >= n_valid) v = kNegBig;
          z[nt][e] = v;
          mx = fmaxf(mx, v);
        }
      }
      mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 2));
      const float m_new = fmaxf(m_run[r], mx);
      alpha[r] = fast_exp2(m_run[r] - m_new);
      m_run[r] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const float p0 = fast_exp2(z[nt][0] - m_new), p1 = fast_exp2(z[nt][1] - m_new);
        sum += p0 + p1;
        pk[r][nt] = pack2<T>(p0, p1);  // P rounded to the 16-bit type before P.V (attention_kernels.cu:398-400)
      }
      l_run[r] = l_run[r] * alpha[r] + sum;
    }
    {
      bool moved = false;
#pragma unroll
      for (int r = 0; r < NG; ++r) moved |= (alpha[r] != 1.f);
      if (__any_sync(FULL, moved)) {
#pragma unroll
        for (int nd = 0; nd < ND; ++nd) {
          oacc[nd][0] *= alpha[0];
          oacc[nd][1] *= alpha[0];
          if constexpr (NG == 2) {
            oacc[nd][2] *= alpha[1];
            oacc[nd][3] *= alpha[1];
          }
        }
      }
    }

    // ================= O += P V  (A = P from the S accumulators, B = V via ldmatrix.trans) =================
#pragma unroll
    for (int kk = 0; kk < KK; ++kk) {
      const uint32_t a0 = pk[0][2 * kk], a2 = pk[0][2 * kk + 1];
      const uint32_t a1 = NG == 2 ? pk[NG - 1][2 * kk] : 0u, a3 = NG == 2 ? pk[NG - 1][2 * kk + 1] : 0u;
#pragma unroll
      for (int dp = 0; dp < ND / 2; ++dp) {
        uint32_t r0, r1, r2, r3;
        ldmatrix_x4_trans(vb + kk * 16 * ROW + dp * 32 + v_lane, r0, r1, r2, r3);
        mma_16816<T>(oacc[2 * dp], a0, a1, a2, a3, r0, r1);
        mma_16816<T>(oacc[2 * dp + 1], a0, a1, a2, a3, r2, r3);
      }
    }

    // ================= refill this stage with block i + NST =================
    __syncwarp();
    if (i + NST < my_n) {
      issue(i + NST, stage, next_phys);
      if (i + NST + 1 < my_n) next_phys = block_table[(i + NST + 1) * kMmaWarps];
    }
    cp_async_commit();
    if (++stage == NST) stage = 0;
  }
  cp_async_wait<0>();
  __syncwarp();

  // ---- leave this warp's partial state in its region for the shared epilogue ---------------------------------------
  float* mo = reinterpret_cast<float*>(my_region);
#pragma unroll
  for (int r = 0; r < NG; ++r) {
    float l = l_run[r];
    l += __shfl_xor_sync(FULL, l, 1);
    l += __shfl_xor_sync(FULL, l, 2);
    if (tig == 0) {
      mo[8 * r + gid] = m_run[r];
      mo[16 + 8 * r + gid] = l;
    }
#pragma unroll
    for (int nd = 0; nd < ND; ++nd) {
      float* o0 = mo + 32 + (8 * r + gid) * OROW + 8 * nd + 2 * tig;
      o0[0] = oacc[nd][2 * r];
      o0[1] = oacc[nd][2 * r + 1];
    }
  }
  decode_epilogue<T, D, HPT>(p, smem, region, seq, seq_len, kv_head, qt, G, part, 1.f);
}

template <typename T, int D, int B, int NG, int NST>
static int launch_flash_one(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  constexpr int HPT = 8 * NG;
  const int smem = flash_smem_bytes(D, B, NG, NST);
  auto kernel = paged_attention_flash_kernel<T, D, B, NG, NST>;
  static int configured[64] = {0};
  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || configured[dev] < smem) {
    B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev >= 0 && dev < 64) configured[dev] = smem;
  }
  const int G = kp.num_heads / kp.num_kv_heads;
  const int q_tiles = (G + HPT - 1) / HPT;
  dim3 grid(kp.num_kv_heads * q_tiles, num_seqs, kp.num_parts);
  kernel<<<grid, kMmaThreads, smem, stream>>>(kp);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

// Ring depth: 3 stages when they fit the CTAs-per-SM the kernel is compiled for, else 2 (B200PA_STAGES overrides).
template <typename T, int D, int B, int NG>
static int launch_flash_stages(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  constexpr int budget = kFlashSmemPerSM / flash_ctas(D, B, NG) - kFlashReserved;
  constexpr bool three_fit = flash_smem_bytes(D, B, NG, 3) <= budget;
  constexpr bool three_possible = flash_smem_bytes(D, B, NG, 3) <= kFlashSmemPerSM - kFlashReserved;
  constexpr bool two_possible = flash_smem_bytes(D, B, NG, 2) <= kFlashSmemPerSM - kFlashReserved;
  if constexpr (three_possible) {
    if (kp.num_stages == 3 || (three_fit && kp.num_stages != 2)) return launch_flash_one<T, D, B, NG, 3>(kp, num_seqs, stream);
  }
  if constexpr (two_possible) return launch_flash_one<T, D, B, NG, 2>(kp, num_seqs, stream);
  // block 32 with head sizes 224 / 256: two 2 x 32-token stages per warp exceed 227 KiB; a single stage per warp
  // still overlaps across the four warps of the CTA
  if constexpr (!two_possible && flash_smem_bytes(D, B, NG, 1) <= kFlashSmemPerSM - kFlashReserved)
    return launch_flash_one<T, D, B, NG, 1>(kp, num_seqs, stream);
  return 1;
}

template <typename T, int D>
static int launch_flash_d(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  const int G = kp.num_heads / kp.num_kv_heads;
  if (kp.block_size == 16)
    return G <= 8 ? launch_flash_stages<T, D, 16, 1>(kp, num_seqs, stream) : launch_flash_stages<T, D, 16, 2>(kp, num_seqs, stream);
  if (kp.block_size == 32)
    return G <= 8 ? launch_flash_stages<T, D, 32, 1>(kp, num_seqs, stream) : launch_flash_stages<T, D, 32, 2>(kp, num_seqs, stream);
  return 1;
}

template <typename T>
static int launch_flash_t(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  switch (kp.head_size) {  // flash_attn.py:24-25 (get_supported_head_sizes)
    case 32: return launch_flash_d<T, 32>(kp, num_seqs, stream);
    case 64: return launch_flash_d<T, 64>(kp, num_seqs, stream);
    case 96: return launch_flash_d<T, 96>(kp, num_seqs, stream);
    case 128: return launch_flash_d<T, 128>(kp, num_seqs, stream);
    case 160: return launch_flash_d<T, 160>(kp, num_seqs, stream);
    case 192: return launch_flash_d<T, 192>(kp, num_seqs, stream);
    case 224: return launch_flash_d<T, 224>(kp, num_seqs, stream);
    case 256: return launch_flash_d<T, 256>(kp, num_seqs, stream);
    default: return 1;
  }
}

bool flash_head_size_supported(int d) { return d >= 32 && d <= 256 && d % 32 == 0; }

int launch_attention_flash(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream) {
  if (dtype == B200PA_F16) return launch_flash_t<__half>(kp, num_seqs, stream);
  if (dtype == B200PA_BF16) return launch_flash_t<__nv_bfloat16>(kp, num_seqs, stream);
  return 1;
}

}  // namespace b200pa
