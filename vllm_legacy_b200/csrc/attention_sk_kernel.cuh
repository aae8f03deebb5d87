// Persistent ("stream-K") paged-attention decode kernel for B200 (sm_100a).
//
// Same per-block math as the grid kernel it replaces (TMA bulk-copy staged KV blocks, mma.sync
// contractions over the whole GQA group, online softmax in registers — see the notes in
// attention_mma_kernel.cuh), but the launch is ONE resident wave of CTAs and the work is the
// flattened list of KV blocks of all (sequence, KV head) items, cut into equal contiguous ranges
// (attention_sched.cuh):
//
//   * every CTA streams the same number of blocks whatever the batch shape — no wave tail for
//     mid-size shards (256 sequences x 1 KV head at TP8), no long-sequence stragglers in ragged
//     batches, and a single sequence is spread over all SMs (v1 needs no caller scratch for that);
//   * each warp's TMA ring runs ACROSS item boundaries: while the warps merge one item's partial
//     results the copies of the next item's blocks are already in flight, so the bytes in flight
//     per SM stay constant for the whole kernel (the grid kernel drained its ring at the end of
//     every CTA — ~10 % of a 512 KiB CTA's life);
//   * an item cut by a CTA boundary leaves un-normalised (max, sum, O) partials in a workspace slot
//     (<= 2 per CTA); the last part to finish — found with a self-resetting arrival counter — merges
//     the slots in a fixed order (deterministic) and writes the output row.  No second launch.
//
// Math parity with the reference kernel (attention_kernels.cu:86-496) is unchanged: fp32 QK
// accumulation, logit = scale*qk + alibi*(j - len + 1), P rounded to the 16-bit type before P.V,
// out = sum(p v) / (sum(p) + 1e-6), tokens >= seq_len masked in both contractions.
#pragma once

#include "attention_mma_kernel.cuh"
#include "attention_sched.cuh"

namespace b200pa {

// named barriers of the intra-CTA merge tree (id 0 is __syncthreads)
constexpr int kSkBarFullA = 1, kSkBarEmptyA = 2, kSkBarFullB = 3, kSkBarEmptyB = 4;

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// A partial result in "fragment layout": register i of lane l at [i * 32 + l]; registers are the
// O^T accumulators (NG * D/16 * 4), then the running maxima (NG) and the quad-reduced sums (NG).
__host__ __device__ constexpr int sk_slot_floats(int D, int NG) { return 32 * (NG * (D / 16) * 4 + 2 * NG); }
// shared memory next to the rings: three merge buffers + 64 bytes of scan scratch
__host__ __device__ constexpr int sk_fixed_smem_bytes(int D, int NG) { return 3 * sk_slot_floats(D, NG) * 4 + 64; }

template <typename T, int KV, int D, int B, int NG>
__global__ void __launch_bounds__(kMmaThreads, (D * NG <= 256) ? 2 : 1)
    paged_attention_sk_kernel(const __grid_constant__ AttnKernelParams p) {
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
  constexpr int NO = NG * KS * 4;          // O^T accumulator registers per lane
  constexpr int SLOT = sk_slot_floats(D, NG);
  constexpr uint32_t FULL = 0xffffffffu;
  static_assert(B == 16 || B == 32, "block size");
  static_assert(D % 16 == 0, "head size");

  extern __shared__ __align__(128) uint8_t smem[];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int gid = lane >> 2;  // mma "groupID": fragment row
  const int tig = lane & 3;   // thread in quad

  const int G = p.num_heads / p.num_kv_heads;
  const int q_tiles = (G + HPT - 1) / HPT;
  const int HK = p.num_kv_heads * q_tiles;  // work units per sequence
  const int NST = p.num_stages;
  const int S = p.num_seqs;

  // shared memory: [4 warps x NST stages][bufA0 | bufA1 | bufB][64 B scan scratch][4 x NST mbarriers]
  const uint32_t stage_base = smem_u32(smem) + warp * NST * STAGE;
  float* bufA = reinterpret_cast<float*>(smem + kMmaWarps * NST * STAGE);
  float* bufB = bufA + 2 * SLOT;
  uint8_t* misc = reinterpret_cast<uint8_t*>(bufA + 3 * SLOT);
  long long* scan_tot = reinterpret_cast<long long*>(misc);          // [4]
  long long* start_info = reinterpret_cast<long long*>(misc + 32);   // [0] first sequence, [1] blocks before it
  const uint32_t bar_base = smem_u32(misc + 64) + warp * NST * 8;

  auto blocks_of = [](int len) { return len > 0 ? (len + B - 1) / B : 0; };

  // ---- phase 0: where does this CTA's range start?  (block-count scan over seq_lens) ---------------
  const int per = (S + kMmaThreads - 1) / kMmaThreads;
  const int lo = min((int)threadIdx.x * per, S), hi = min(lo + per, S);
  long long mine = 0;
  for (int s = lo; s < hi; ++s) {
    const int len = p.seq_lens[s];
    mine += blocks_of(len);
    if (len <= 0 && blockIdx.x == 0) {  // empty context => zeros (nobody else touches this row)
      T* o = static_cast<T*>(p.out) + (int64_t)s * p.num_heads * D;
      for (int i = 0; i < p.num_heads * D; ++i) o[i] = from_float<T>(0.f);
    }
  }
  long long incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long v = __shfl_up_sync(FULL, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) scan_tot[warp] = incl;
  __syncthreads();
  long long total = 0;
#pragma unroll
  for (int w2 = 0; w2 < kMmaWarps; ++w2) {
    const long long t = scan_tot[w2];
    if (w2 < warp) incl += t;
    total += t;
  }
  const long long W = total * HK;
  const int Geff = sk_effective_ctas(W, (int)gridDim.x);
  const int cta = blockIdx.x;
  if (cta >= Geff) return;  // (whole CTA; also W == 0)
  long long t_begin = 0;
  if (p.sk_trace && threadIdx.x == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_begin));
  const long long f0 = sk_range_begin(W, Geff, cta);
  const int len = (int)(sk_range_begin(W, Geff, cta + 1) - f0);  // >= 1 blocks of this CTA
  {
    const long long x = f0 / HK;  // the start in units of "blocks of one unit"
    const long long excl = incl - mine;
    if (x >= excl && x < incl) {  // exactly one thread: the one whose slice holds the start
      long long N = excl;
      int s = lo;
      for (;; ++s) {
        const int n = blocks_of(p.seq_lens[s]);
        if (x < N + n) break;
        N += n;
      }
      start_info[0] = s;
      start_info[1] = N;
    }
  }
  __syncthreads();

  // ---- cursors over the item list -------------------------------------------------------------------
  // `base` = position of the item's block 0 relative to f0 (negative when the item began earlier).
  struct Cur {
    int seq, hk, n, slen, base;
  };
  auto next_item = [&](Cur& c) {
    c.base += c.n;
    if (++c.hk == HK) {
      c.hk = 0;
      do {
        ++c.seq;
        c.slen = p.seq_lens[c.seq];
        c.n = blocks_of(c.slen);
      } while (c.n == 0 && c.seq + 1 < S);
    }
  };
  Cur c0;
  {
    c0.seq = (int)start_info[0];
    c0.slen = p.seq_lens[c0.seq];
    c0.n = blocks_of(c0.slen);
    const long long r0 = f0 - start_info[1] * HK;  // in [0, n * HK)
    c0.hk = (int)(r0 / c0.n);
    c0.base = -(int)(r0 - (long long)c0.hk * c0.n);
  }

  const char* kcache = static_cast<const char*>(p.k_cache);
  const char* vcache = static_cast<const char*>(p.v_cache);
  const int64_t head_stride_bytes = p.kv_head_stride * CTB;
  const int64_t blk_stride_bytes = p.kv_block_stride * CTB;
  uint64_t policy = 0;
  if (lane == 0) policy = (p.flags & 1) ? l2_evict_normal_policy() : l2_evict_first_policy();
  policy = __shfl_sync(FULL, policy, 0);

  auto issue = [&](const Cur& c, int rel, int phys, uint32_t bar, uint32_t dst) {
    (void)rel;
    const int kvh = q_tiles == 1 ? c.hk : c.hk / q_tiles;
    const int64_t off = (int64_t)kvh * head_stride_bytes + (int64_t)phys * blk_stride_bytes;
    mbar_arrive_expect_tx(bar, STAGE);
    tma_bulk_g2s(dst, kcache + off, CHUNK, bar, policy);
    tma_bulk_g2s(dst + CHUNK, vcache + off, CHUNK, bar, policy);
  };

  // ---- pipeline prologue: lane i issues this warp's block i (positions warp + 4 i) ---------------
  if (lane == 0) {
    for (int s = 0; s < NST; ++s) mbar_init(bar_base + s * 8, 1);
    fence_proxy_async_smem();
  }
  __syncwarp();
  {
    const int lrel = warp + kMmaWarps * lane;
    if (lane < NST && lrel < len) {
      Cur lc = c0;
      while (lrel >= lc.base + lc.n) next_item(lc);
      const int phys = p.block_tables[(int64_t)lc.seq * p.max_blocks_per_seq + (lrel - lc.base)];
      issue(lc, lrel, phys, bar_base + lane * 8, stage_base + lane * STAGE);
    }
    __syncwarp();
  }
  // producer state (warp-uniform): the NEXT block to issue, its item and its physical block number
  Cur pc = c0;
  int prel = warp + kMmaWarps * NST;
  int pphys = 0;
  auto prod_seek = [&]() {
    while (prel >= pc.base + pc.n) next_item(pc);
    pphys = p.block_tables[(int64_t)pc.seq * p.max_blocks_per_seq + (prel - pc.base)];
  };
  if (prel < len) prod_seek();

  // ---- per-lane constant offsets (see attention_mma_kernel.cuh for the token permutation) ---------
  auto rot = [](int q) { return q * NP / 4; };
  auto tokmap = [&](int nt, int j) { return TPT * (j >> 1) + 2 * ((nt + rot(j >> 1)) % NP) + (j & 1); };
  const int mi = lane >> 3, mj = lane & 7;
  uint32_t koff[NT / 2 > 0 ? (FP8 ? 2 : NT / 2) : 1];
  if constexpr (!FP8) {
#pragma unroll
    for (int c = 0; c < NT / 2; ++c) koff[c] = (mi & 1) * UNIT + tokmap(2 * c + (mi >> 1), mj) * 16;
  } else if constexpr (B == 16) {
    koff[0] = (mi >> 1) * UNIT + tokmap(mi & 1, mj) * 16;
    koff[1] = tokmap(mi & 1, mj) * 16;  // odd tail: both halves read d-group KS-1
  } else {
    koff[0] = tokmap(mi, mj) * 16;
    koff[1] = 0;
  }
  const uint32_t voff = gid * (B * CTB) + tig * (TPT * CTB);
  int tpos[NT];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) tpos[nt] = TPT * tig + 2 * ((nt + rot(tig)) % NP);

  const float z_mult = (FP8 ? p.kv_scale : p.scale) * kLog2e;  // S_raw -> log2-domain logit
  const float out_scale = FP8 ? p.kv_scale : 1.f;
  const bool has_alibi = p.alibi_slopes != nullptr;

  // ---- running state of the current item ----------------------------------------------------------
  float m_run[NG], l_run[NG];
  float oacc[NG][KS][4];

  // fragment-layout store / merge of (m, l, O); `l_run` must already be quad-reduced
  auto store_frag = [&](float* buf) {
#pragma unroll
    for (int r = 0; r < NG; ++r) {
#pragma unroll
      for (int dt = 0; dt < KS; ++dt)
#pragma unroll
        for (int e = 0; e < 4; ++e) buf[((r * KS + dt) * 4 + e) * 32 + lane] = oacc[r][dt][e];
      buf[(NO + r) * 32 + lane] = m_run[r];
      buf[(NO + NG + r) * 32 + lane] = l_run[r];
    }
  };
  auto merge_frag = [&](const float* buf, auto ld) {
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      // O^T columns of this lane are query rows 2tig, 2tig+1 (owned by quads 2tig, 2tig+1)
      const float ms0 = __shfl_sync(FULL, m_run[r], 8 * tig), ms1 = __shfl_sync(FULL, m_run[r], 8 * tig + 4);
      const float mo0 = ld(buf + (NO + r) * 32 + 8 * tig), mo1 = ld(buf + (NO + r) * 32 + 8 * tig + 4);
      const float M0 = fmaxf(ms0, mo0), M1 = fmaxf(ms1, mo1);
      const float fs0 = exp2f(ms0 - M0), fo0 = exp2f(mo0 - M0);
      const float fs1 = exp2f(ms1 - M1), fo1 = exp2f(mo1 - M1);
#pragma unroll
      for (int dt = 0; dt < KS; ++dt) {
        const float* b = buf + ((r * KS + dt) * 4) * 32 + lane;
        oacc[r][dt][0] = oacc[r][dt][0] * fs0 + ld(b) * fo0;
        oacc[r][dt][1] = oacc[r][dt][1] * fs1 + ld(b + 32) * fo1;
        oacc[r][dt][2] = oacc[r][dt][2] * fs0 + ld(b + 64) * fo0;
        oacc[r][dt][3] = oacc[r][dt][3] * fs1 + ld(b + 96) * fo1;
      }
      // row state of this lane's own row (gid)
      const float mo = ld(buf + (NO + r) * 32 + lane), lo_ = ld(buf + (NO + NG + r) * 32 + lane);
      const float M = fmaxf(m_run[r], mo);
      l_run[r] = l_run[r] * exp2f(m_run[r] - M) + lo_ * exp2f(mo - M);
      m_run[r] = M;
    }
  };
  auto ld_smem = [](const float* q) { return *q; };
  auto ld_glob = [](const float* q) { return ld_cg_f32(q); };

  // ---- consumer loop over items -----------------------------------------------------------------------
  Cur cc = c0;
  int crel = warp;  // position (relative to f0) of this warp's next block
  int stage = 0;
  uint32_t phase = 0;
  bool first_item = true;
  for (;;) {
    const int seq = cc.seq;
    const int seq_len = cc.slen;
    const int kv_head = q_tiles == 1 ? cc.hk : cc.hk / q_tiles;
    const int qt = cc.hk - kv_head * q_tiles;
    const int jb = max(0, -cc.base);            // this CTA's part of the item: blocks [jb, je)
    const int je = min(cc.n, len - cc.base);
    const int part_end = cc.base + je;
    const bool last_item = cc.base + cc.n >= len;

    // Q fragments (A operand of QK^T) of this item
    // 16-bit cache: k index of step ks <-> d = 16ks + {2tig, 2tig+1 | +8};
    // fp8 cache   : k index of step ks <-> d = 16ks + {4tig, 4tig+1 | 4tig+2, 4tig+3}
    uint32_t qa[KS][2 * NG];
    float alibi_l2[NG];
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
            const float2 f0_ = unpack2<T>(w0), f1_ = unpack2<T>(w1);
            w0 = pack2<__half>(f0_.x * p.scale, f0_.y * p.scale);
            w1 = pack2<__half>(f1_.x * p.scale, f1_.y * p.scale);
          }
          qa[ks][r] = w0;       // a0 (r=0) / a1 (r=1)
          qa[ks][NG + r] = w1;  // a2 (r=0) / a3 (r=1)
        }
        alibi_l2[r] = (has_alibi && valid) ? p.alibi_slopes[kv_head * G + hr] * kLog2e : 0.f;
      }
    }
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      m_run[r] = kNegBig;
      l_run[r] = 0.f;
#pragma unroll
      for (int dt = 0; dt < KS; ++dt)
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[r][dt][e] = 0.f;
    }

    for (; crel < part_end; crel += kMmaWarps) {
      const uint32_t bar = bar_base + stage * 8;
      const uint32_t kb = stage_base + stage * STAGE;
      const uint32_t vb = kb + CHUNK;
      const int blk_tok0 = (crel - cc.base) * B;  // first token of this block
      const int n_valid = seq_len - blk_tok0;     // >= 1; < B only in the last block
      mbar_wait(bar, phase);

      // ================= S = Q K^T =================
      float sacc[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) sacc[nt][e] = 0.f;

      if constexpr (!FP8) {
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
          pk[r][nt] = pack2<MT>(p0, p1);
        }
        l_run[r] = l_run[r] * alpha[r] + sum;
      }

      // P pairs in token order: pair q of this lane's tokens sits in n-tile (q - rot(tig)) mod NP.
      uint32_t pp[NG][NP];
#pragma unroll
      for (int r = 0; r < NG; ++r) {
        if constexpr (NP == 2) {
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
        if (__any_sync(FULL, moved)) {
#pragma unroll
          for (int r = 0; r < NG; ++r) {
            const float a0 = __shfl_sync(FULL, alpha[r], 8 * tig);
            const float a1 = __shfl_sync(FULL, alpha[r], 8 * tig + 4);
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
          if constexpr (!FP8 && B == 16) {
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
#pragma unroll
        for (int kk = 0; kk < KK; ++kk)
#pragma unroll
          for (int r = 0; r < NG; ++r)
            mma_16816<MT>(oacc[r][dt], va[0][2 * kk], va[1][2 * kk], va[0][2 * kk + 1], va[1][2 * kk + 1],
                          pp[r][2 * kk], pp[r][2 * kk + 1]);
      }

      // ================= refill this stage with the producer's next block =================
      __syncwarp();
      if (prel < len) {
        if (lane == 0) {
          fence_proxy_async_smem();
          issue(pc, prel, pphys, bar, kb);
        }
        prel += kMmaWarps;
        if (prel < len) prod_seek();
      }
      if (++stage == NST) {
        stage = 0;
        phase ^= 1;
      }
    }

    // ---- merge the four warps' partial states of this item: (2,3) -> (0,1), 1 -> 0 -------------------
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      l_run[r] += __shfl_xor_sync(FULL, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(FULL, l_run[r], 2);
    }
    if (warp >= 2) {
      if (!first_item) named_bar_sync(kSkBarEmptyA, 128);
      store_frag(bufA + (warp - 2) * SLOT);
      __threadfence_block();
      named_bar_arrive(kSkBarFullA, 128);
    } else {
      named_bar_sync(kSkBarFullA, 128);
      merge_frag(bufA + warp * SLOT, ld_smem);
      if (!last_item) named_bar_arrive(kSkBarEmptyA, 128);
      if (warp == 1) {
        if (!first_item) named_bar_sync(kSkBarEmptyB, 64);
        store_frag(bufB);
        __threadfence_block();
        named_bar_arrive(kSkBarFullB, 64);
      } else {
        named_bar_sync(kSkBarFullB, 64);
        merge_frag(bufB, ld_smem);
        if (!last_item) named_bar_arrive(kSkBarEmptyB, 64);

        // ---- warp 0 owns the merged part: write the row, or leave a partial for the last arriver ----
        auto write_out = [&]() {
#pragma unroll
          for (int r = 0; r < NG; ++r) {
            const float l0 = __shfl_sync(FULL, l_run[r], 8 * tig), l1 = __shfl_sync(FULL, l_run[r], 8 * tig + 4);
            const float inv0 = out_scale * __fdividef(1.f, l0 + 1e-6f);
            const float inv1 = out_scale * __fdividef(1.f, l1 + 1e-6f);
            const int hr0 = qt * HPT + 8 * r + 2 * tig;
            T* o0 = static_cast<T*>(p.out) + ((int64_t)seq * p.num_heads + kv_head * G + hr0) * D + gid;
#pragma unroll
            for (int dt = 0; dt < KS; ++dt) {
              if (hr0 < G) {
                o0[16 * dt] = from_float<T>(oacc[r][dt][0] * inv0);
                o0[16 * dt + 8] = from_float<T>(oacc[r][dt][2] * inv0);
              }
              if (hr0 + 1 < G) {
                o0[D + 16 * dt] = from_float<T>(oacc[r][dt][1] * inv1);
                o0[D + 16 * dt + 8] = from_float<T>(oacc[r][dt][3] * inv1);
              }
            }
          }
        };
        if (jb == 0 && je == cc.n) {
          write_out();
        } else {
          float* slots = p.sk_slots;
          store_frag(slots + (size_t)sk_slot_index(cta, jb > 0) * SLOT);
          __threadfence();
          __syncwarp();
          const long long a = f0 + cc.base;  // flat index of the item's block 0
          const int ch = sk_cta_of(W, Geff, a), cl = sk_cta_of(W, Geff, a + cc.n - 1);
          int last = 0;
          if (lane == 0) last = atomicAdd(p.sk_counters + ch, 1u) == (unsigned)(cl - ch);
          last = __shfl_sync(FULL, last, 0);
          if (last) {  // every part of the item is in its slot: merge them in CTA order
            __threadfence();
#pragma unroll
            for (int r = 0; r < NG; ++r) {
              m_run[r] = kNegBig;
              l_run[r] = 0.f;
#pragma unroll
              for (int dt = 0; dt < KS; ++dt)
#pragma unroll
                for (int e = 0; e < 4; ++e) oacc[r][dt][e] = 0.f;
            }
            for (int c2 = ch; c2 <= cl; ++c2)
              merge_frag(slots + (size_t)sk_slot_index(c2, c2 != ch) * SLOT, ld_glob);
            write_out();
            if (lane == 0) p.sk_counters[ch] = 0u;  // self-resetting: the next launch starts from zero
          }
        }
      }
    }

    if (last_item) break;
    next_item(cc);
    first_item = false;
  }
  if (p.sk_trace && threadIdx.x == 0) {
    long long t_end;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_end));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    long long* tr = p.sk_trace + (size_t)cta * 4;
    tr[0] = smid; tr[1] = t_begin; tr[2] = t_end; tr[3] = len;
  }
}

}  // namespace b200pa
