// Persistent paged-attention decode kernel for B200 (sm_100a): warp-specialised producer /
// consumers over a dynamically scheduled, flattened KV-block list.
//
// Same per-block math as the round-1 grid kernel (TMA bulk-copy staged KV blocks, mma.sync
// contractions over the whole GQA group, online softmax in registers — see the notes in
// attention_mma_kernel.cuh), but:
//
//   * ONE resident wave of CTAs (SMs x CTAs/SM) streams the flattened list of KV blocks of all
//     (sequence, KV head) items.  The list is cut into a fixed set of chunks (attention_sched.cuh):
//     one big static chunk per CTA + a pool of small chunks taken with an atomic counter, so every
//     SM streams at its own rate until the work is gone (B200's SMs differ by +-20 % on equal
//     ranges) and the tail is one small chunk.  A single sequence is spread over all SMs; v1 needs
//     no caller scratch for that.
//   * warp 4 is the SCHEDULER: it scans seq_lens, takes chunks, walks the items of a chunk,
//     publishes one "part" descriptor per (item x chunk) intersection in a shared-memory queue,
//     looks the blocks up in the block table (32 entries per coalesced load, one lane per block)
//     and appends each block's cache byte offset to the ADDRESS QUEUE of the consumer warp that
//     owns it (blocks of a part go round-robin to warps 0-3).  It runs up to 32 blocks per warp
//     ahead, across item and chunk boundaries, so its own latencies (atomics, block-table loads,
//     64-bit divisions) are never seen by the memory pipeline.
//   * warps 0-3 are CONSUMERS, each with its own TMA ring: lane 0 pops addresses and issues the
//     two `cp.async.bulk` copies per block as soon as a stage is free — ACROSS item and chunk
//     boundaries, so the bytes in flight per SM stay constant for the life of the kernel — then
//     the warp waits for a stage and runs QK^T / softmax / PV on it.  At the
//     end of a part the four partial (max, sum, O) states are merged through shared memory
//     ((2,3) -> (0,1), 1 -> 0; named barriers, no CTA-wide barrier) and warp 0 writes the output
//     row — or, when the item is cut by a chunk boundary, an un-normalised partial into the
//     chunk's workspace slot.  The last part of an item to finish (self-resetting arrival
//     counter) merges all its slots in chunk order — deterministic — with all four warps.
//
// Math parity with the reference kernel (attention_kernels.cu:86-496) is unchanged: fp32 QK
// accumulation, logit = scale*qk + alibi*(j - len + 1), P rounded to the 16-bit type before P.V,
// out = sum(p v) / (sum(p) + 1e-6), tokens >= seq_len masked in both contractions.
#pragma once

#include "attention_mma_kernel.cuh"
#include "attention_sched.cuh"

namespace b200pa {

constexpr int kSkThreads = kMmaThreads + 32;  // 4 consumer warps + 1 producer warp
constexpr int kSkPartQueue = 64;              // part descriptors in flight (scheduler waits for room)
constexpr int kSkAddrQueue = 32;              // block addresses in flight per consumer warp

// named barriers (id 0 is __syncthreads): merge tree of the consumer warps + the fix-up decision
constexpr int kSkBarFullA = 1, kSkBarEmptyA = 2, kSkBarFullB = 3, kSkBarEmptyB = 4, kSkBarFix = 5;

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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ int ld_acquire_smem(uint32_t addr) {
  int v;
  asm volatile("ld.acquire.cta.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_smem(uint32_t addr, int v) {
  asm volatile("st.release.cta.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// A partial result in "fragment layout": register i of lane l at [i * 32 + l]; registers are the
// O^T accumulators (NG * D/16 * 4), then the running maxima (NG) and the quad-reduced sums (NG).
__host__ __device__ constexpr int sk_slot_floats(int D, int NG) { return 32 * (NG * (D / 16) * 4 + 2 * NG); }
// shared memory next to the rings: three merge buffers, the part queue, four address queues, 64 bytes of flags
__host__ __device__ constexpr int sk_fixed_smem_bytes(int D, int NG) {
  return 3 * sk_slot_floats(D, NG) * 4 + kSkPartQueue * 32 + kMmaWarps * kSkAddrQueue * 8 + 64;
}
// flags (ints): parts published | parts consumed | fix-up decision | scheduler done | addresses published [4] |
// addresses taken [4]
enum { kSkFlTail = 0, kSkFlHead = 1, kSkFlFix = 2, kSkFlDone = 3, kSkFlAvail = 4, kSkFlTaken = 8 };

// Two CTAs per SM when 2 x 160 threads fit the register file at this head size / group tile.
__host__ __device__ constexpr bool sk_two_ctas(int D, int NG) { return NG == 1 && D <= 128; }

template <typename T, int KV, int D, int B, int NG>
__global__ void __launch_bounds__(kSkThreads, sk_two_ctas(D, NG) ? 2 : 1)
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
  const int cta = blockIdx.x;

  // shared memory: [4 warps x NST stages][bufA0 | bufA1 | bufB][part queue][address queues][flags][full bars]
  const uint32_t ring_base = smem_u32(smem);
  float* bufA = reinterpret_cast<float*>(smem + kMmaWarps * NST * STAGE);
  float* bufB = bufA + 2 * SLOT;
  int4* pq = reinterpret_cast<int4*>(bufA + 3 * SLOT);                         // [kSkPartQueue][2]
  long long* aq = reinterpret_cast<long long*>(pq + kSkPartQueue * 2);         // [4][kSkAddrQueue]
  int* flags = reinterpret_cast<int*>(aq + kMmaWarps * kSkAddrQueue);          // [16]
  const uint32_t flags_addr = smem_u32(flags);
  const uint32_t tail_addr = flags_addr + 4 * kSkFlTail;
  const uint32_t full_base = flags_addr + 64;

  if (threadIdx.x == kMmaThreads) {  // scheduler lane 0 sets the pipeline up
    for (int i = 0; i < kMmaWarps * NST; ++i) mbar_init(full_base + i * 8, 1);
    for (int i = 0; i < 16; ++i) flags[i] = 0;
    fence_proxy_async_smem();
  }
  long long t_begin = 0;
  if (p.sk_trace && threadIdx.x == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_begin));
  __syncthreads();

  if (warp == kMmaWarps) {
    // =============================================== SCHEDULER ===============================================
    auto blocks_of = [](int len) { return len > 0 ? (len + B - 1) / B : 0; };
    // ---- block-count scan over seq_lens: lane l owns the slice [lo, hi) -------------------------------------
    const int per = (S + 31) / 32;
    const int lo = min(lane * per, S), hi = min(lo + per, S);
    long long mine = 0;
    for (int s = lo; s < hi; ++s) {
      const int len = p.seq_lens[s];
      mine += blocks_of(len);
      if (len <= 0 && cta == 0) {  // empty context => zeros (nobody else touches this row)
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
    const long long excl = incl - mine;
    const long long W = __shfl_sync(FULL, incl, 31) * HK;
    const SkGeom geo = sk_geometry(W, (int)gridDim.x, p.sk_pool_pct, p.sk_chunks_per_cta, p.sk_cs_min, p.sk_max_chunks);
    const int NC = sk_num_chunks(geo);

    // `base` = position of the item's block 0 relative to the chunk start (negative: the item began earlier)
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
    auto locate = [&](long long f, Cur& c) {  // all lanes; the result is warp-uniform
      const long long x = f / HK;             // in units of "blocks of one unit"
      const int src = __ffs(__ballot_sync(FULL, x >= excl && x < incl)) - 1;  // the slice that holds it
      int s = 0;
      long long N = 0;
      if (lane == src) {
        N = excl;
        for (s = lo;; ++s) {
          const int n = blocks_of(p.seq_lens[s]);
          if (x < N + n) break;
          N += n;
        }
      }
      s = __shfl_sync(FULL, s, src);
      N = __shfl_sync(FULL, N, src);
      c.seq = s;
      c.slen = p.seq_lens[s];
      c.n = blocks_of(c.slen);
      const long long r0 = f - N * HK;  // in [0, n * HK)
      c.hk = (int)(r0 / c.n);
      c.base = -(int)(r0 - (long long)c.hk * c.n);
    };

    const int64_t head_stride_bytes = p.kv_head_stride * CTB;
    const int64_t blk_stride_bytes = p.kv_block_stride * CTB;

    int nq0 = 0, nq1 = 0, nq2 = 0, nq3 = 0;  // addresses appended to the queue of consumer warp 0..3
    int npart = 0;                           // parts published
    int nxt = 0;        // (lane 0) next pool chunk, taken a little before the current chunk ends
    int chunk = cta;
    if (cta >= geo.Gs) {
      if (lane == 0) nxt = geo.Gs + (int)atomicAdd(p.sk_pool, 1u);
      chunk = __shfl_sync(FULL, nxt, 0);
    }
    while (chunk < NC) {
      const long long f0c = sk_chunk_begin(geo, chunk);
      const int lenc = (int)(sk_chunk_begin(geo, chunk + 1) - f0c);
      bool have_next = false;
      Cur cur;
      locate(f0c, cur);
      for (int rel = 0;;) {
        // ---- one part: the blocks [jb, je) of item `cur` that lie in this chunk ------------------------------
        const int jb = rel - cur.base;
        const int je = min(cur.n, lenc - cur.base);
        const int cnt = je - jb;
        int slot = -1, chx = 0;
        if (!(jb == 0 && je == cur.n)) {  // the item is cut: who else holds a part of it?
          slot = sk_slot_index(chunk, jb > 0);
          const long long a = f0c + cur.base;  // flat index of the item's block 0
          const int ch = sk_chunk_of(geo, a), cl = sk_chunk_of(geo, a + cur.n - 1);
          chx = ch | ((cl - ch) << 16);
        }
        // Block i of the part goes to consumer warp (i + rot) & 3.  rot depends on the part only (not on what
        // this CTA streamed before), so the grouping of the partial sums — hence every output bit — is the same
        // whichever CTA takes the chunk; the seq + hk term spreads the first blocks of short items over the warps.
        const int rot = (jb + cur.seq + cur.hk) & 3;
        if (lane == 0) {
          // room in the part queue?  (the consumers are within two parts of each other)
          while (npart - ld_acquire_smem(flags_addr + 4 * kSkFlHead) > kSkPartQueue - 8) {
          }
          int4* e = pq + (npart & (kSkPartQueue - 1)) * 2;
          e[0] = make_int4(cur.seq, cur.hk, cur.slen, jb);
          e[1] = make_int4(cnt, rot, slot, chx);
          st_release_smem(tail_addr, npart + 1);
        }
        ++npart;
        const int kvh = q_tiles == 1 ? cur.hk : cur.hk / q_tiles;
        const int64_t head_off = (int64_t)kvh * head_stride_bytes;
        const int* bt = p.block_tables + (int64_t)cur.seq * p.max_blocks_per_seq + jb;
        for (int i0 = 0; i0 < cnt; i0 += 32) {
          if (!have_next && lenc - (rel + i0) <= 64) {  // close to the end of the chunk: take the next one
            if (lane == 0) nxt = geo.Gs + (int)atomicAdd(p.sk_pool, 1u);
            have_next = true;
          }
          const int nb = min(32, cnt - i0);
          // lane t owns block i0 + t of the part; lanes t, t + 4, ... feed the same consumer warp
          const int64_t off = head_off + (lane < nb ? (int64_t)bt[i0 + lane] * blk_stride_bytes : 0);
          const int w = (i0 + lane + rot) & 3;
          const int first = (w - i0 - rot) & 3;                  // the first lane of this batch that feeds warp w
          const int nw = nb > first ? (nb - first + 3) >> 2 : 0;  // blocks of this batch for warp w
          const int base = w == 0 ? nq0 : w == 1 ? nq1 : w == 2 ? nq2 : nq3;
          if (lane < kMmaWarps && nw > 0) {  // (lanes 0-3 feed four different warps) room in that warp's queue?
            while (base + nw - ld_acquire_smem(flags_addr + 4 * (kSkFlTaken + w)) > kSkAddrQueue) {
            }
          }
          __syncwarp();
          if (lane < nb) aq[w * kSkAddrQueue + ((base + ((lane - first) >> 2)) & (kSkAddrQueue - 1))] = off;
          __syncwarp();
          if (lane < kMmaWarps && nw > 0) st_release_smem(flags_addr + 4 * (kSkFlAvail + w), base + nw);
          // every lane advances all four counters
          {
            const int f0_ = (0 - i0 - rot) & 3, f1_ = (1 - i0 - rot) & 3, f2_ = (2 - i0 - rot) & 3, f3_ = (3 - i0 - rot) & 3;
            nq0 += nb > f0_ ? (nb - f0_ + 3) >> 2 : 0;
            nq1 += nb > f1_ ? (nb - f1_ + 3) >> 2 : 0;
            nq2 += nb > f2_ ? (nb - f2_ + 3) >> 2 : 0;
            nq3 += nb > f3_ ? (nb - f3_ + 3) >> 2 : 0;
          }
        }
        rel += cnt;
        if (rel >= lenc) break;
        next_item(cur);
      }
      if (!have_next && lane == 0) nxt = geo.Gs + (int)atomicAdd(p.sk_pool, 1u);
      chunk = __shfl_sync(FULL, nxt, 0);
    }
    if (lane == 0) {
      // end of stream: a part with no blocks
      while (npart - ld_acquire_smem(flags_addr + 4 * kSkFlHead) > kSkPartQueue - 8) {
      }
      int4* e = pq + (npart & (kSkPartQueue - 1)) * 2;
      e[0] = make_int4(0, 0, 0, 0);
      e[1] = make_int4(0, 0, -1, 0);
      st_release_smem(tail_addr, npart + 1);
      st_release_smem(flags_addr + 4 * kSkFlDone, 1);
      // the last CTA to get here re-arms the pool counter for the next launch
      __threadfence();
      if (atomicAdd(p.sk_pool + 1, 1u) == gridDim.x - 1) {
        __threadfence();
        p.sk_pool[0] = 0u;
        p.sk_pool[1] = 0u;
      }
    }
    return;
  }

  // ================================================= CONSUMERS =================================================
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

  // ---- running state of the current part ----------------------------------------------------------
  float m_run[NG], l_run[NG];
  float oacc[NG][KS][4];
  auto reset_state = [&]() {
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      m_run[r] = kNegBig;
      l_run[r] = 0.f;
#pragma unroll
      for (int dt = 0; dt < KS; ++dt)
#pragma unroll
        for (int e = 0; e < 4; ++e) oacc[r][dt][e] = 0.f;
    }
  };

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
  // Two workspace slots at once (all loads issued before the first use).
  auto merge_frag2 = [&](const float* x, const float* y) {
    float ox[NO + 3 * NG], oy[NO + 3 * NG];
#pragma unroll
    for (int r = 0; r < NG; ++r) {
#pragma unroll
      for (int i = 0; i < KS * 4; ++i) {
        ox[r * KS * 4 + i] = ld_cg_f32(x + (r * KS * 4 + i) * 32 + lane);
        oy[r * KS * 4 + i] = ld_cg_f32(y + (r * KS * 4 + i) * 32 + lane);
      }
      ox[NO + 3 * r] = ld_cg_f32(x + (NO + r) * 32 + 8 * tig);
      ox[NO + 3 * r + 1] = ld_cg_f32(x + (NO + r) * 32 + 8 * tig + 4);
      ox[NO + 3 * r + 2] = ld_cg_f32(x + (NO + r) * 32 + lane);
      oy[NO + 3 * r] = ld_cg_f32(y + (NO + r) * 32 + 8 * tig);
      oy[NO + 3 * r + 1] = ld_cg_f32(y + (NO + r) * 32 + 8 * tig + 4);
      oy[NO + 3 * r + 2] = ld_cg_f32(y + (NO + r) * 32 + lane);
    }
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      const float lx = ld_cg_f32(x + (NO + NG + r) * 32 + lane), ly = ld_cg_f32(y + (NO + NG + r) * 32 + lane);
      const float ms0 = __shfl_sync(FULL, m_run[r], 8 * tig), ms1 = __shfl_sync(FULL, m_run[r], 8 * tig + 4);
      const float M0 = fmaxf(ms0, fmaxf(ox[NO + 3 * r], oy[NO + 3 * r]));
      const float M1 = fmaxf(ms1, fmaxf(ox[NO + 3 * r + 1], oy[NO + 3 * r + 1]));
      const float fs0 = exp2f(ms0 - M0), fx0 = exp2f(ox[NO + 3 * r] - M0), fy0 = exp2f(oy[NO + 3 * r] - M0);
      const float fs1 = exp2f(ms1 - M1), fx1 = exp2f(ox[NO + 3 * r + 1] - M1), fy1 = exp2f(oy[NO + 3 * r + 1] - M1);
#pragma unroll
      for (int dt = 0; dt < KS; ++dt) {
        const int i = (r * KS + dt) * 4;
        oacc[r][dt][0] = (oacc[r][dt][0] * fs0 + ox[i] * fx0) + oy[i] * fy0;
        oacc[r][dt][1] = (oacc[r][dt][1] * fs1 + ox[i + 1] * fx1) + oy[i + 1] * fy1;
        oacc[r][dt][2] = (oacc[r][dt][2] * fs0 + ox[i + 2] * fx0) + oy[i + 2] * fy0;
        oacc[r][dt][3] = (oacc[r][dt][3] * fs1 + ox[i + 3] * fx1) + oy[i + 3] * fy1;
      }
      const float mx = ox[NO + 3 * r + 2], my = oy[NO + 3 * r + 2];
      const float M = fmaxf(m_run[r], fmaxf(mx, my));
      l_run[r] = (l_run[r] * exp2f(m_run[r] - M) + lx * exp2f(mx - M)) + ly * exp2f(my - M);
      m_run[r] = M;
    }
  };
  auto ld_smem = [](const float* q) { return *q; };
  auto ld_glob = [](const float* q) { return ld_cg_f32(q); };


  // Merge the four warps' (m, l, O) into warp 0: (2,3) -> (0,1) through bufA, 1 -> 0 through bufB.
  // A buffer is handed back ("empty") at the start of the NEXT merge, so no arrival is left dangling.
  int merges = 0;
  auto tree_merge = [&]() {
    if (warp >= 2) {
      if (merges) named_bar_sync(kSkBarEmptyA, 128);
      store_frag(bufA + (warp - 2) * SLOT);
      __threadfence_block();
      named_bar_arrive(kSkBarFullA, 128);
    } else {
      if (merges) named_bar_arrive(kSkBarEmptyA, 128);
      named_bar_sync(kSkBarFullA, 128);
      merge_frag(bufA + warp * SLOT, ld_smem);
      if (warp == 1) {
        if (merges) named_bar_sync(kSkBarEmptyB, 64);
        store_frag(bufB);
        __threadfence_block();
        named_bar_arrive(kSkBarFullB, 64);
      } else {
        if (merges) named_bar_arrive(kSkBarEmptyB, 64);
        named_bar_sync(kSkBarFullB, 64);
        merge_frag(bufB, ld_smem);
      }
    }
    ++merges;
  };

  // ---- this warp's TMA ring: lane 0 pops block addresses from the warp's queue and issues the copies ----------
  const uint32_t my_ring = ring_base + warp * NST * STAGE;
  const uint32_t my_full = full_base + warp * NST * 8;
  const long long* my_aq = aq + warp * kSkAddrQueue;
  const uint32_t avail_addr = flags_addr + 4 * (kSkFlAvail + warp), taken_addr = flags_addr + 4 * (kSkFlTaken + warp);
  const char* kcache = static_cast<const char*>(p.k_cache);
  const char* vcache = static_cast<const char*>(p.v_cache);
  uint64_t policy = 0;
  if (lane == 0) policy = (p.flags & 1) ? l2_evict_normal_policy() : l2_evict_first_policy();
  int consumed = 0;           // blocks this warp has finished
  int issued = 0, istage = 0;  // (lane 0) blocks whose copies were issued; the stage the next one goes to
  // Issue every block whose address is there and whose stage is free (non-blocking unless `need`: then block
  // until block `consumed` itself has been issued — its part is published, so its address is on the way).
  auto top_up = [&](bool need) {
    if (lane == 0) {
      for (;;) {
        if (issued >= consumed + NST) break;  // ring full
        int av = ld_acquire_smem(avail_addr);
        if (av <= issued) {
          if (!need || issued > consumed) break;
          continue;  // spin: the scheduler is about to publish it
        }
        const int take = min(av, consumed + NST);
        fence_proxy_async_smem();  // the stages below were read through the generic proxy
        for (; issued < take; ++issued) {
          const long long off = my_aq[issued & (kSkAddrQueue - 1)];
          const uint32_t bar = my_full + istage * 8, dst = my_ring + istage * STAGE;
          mbar_arrive_expect_tx(bar, STAGE);
          tma_bulk_g2s(dst, kcache + off, CHUNK, bar, policy);
          tma_bulk_g2s(dst + CHUNK, vcache + off, CHUNK, bar, policy);
          if (++istage == NST) istage = 0;
        }
        st_release_smem(taken_addr, issued);
      }
    }
    __syncwarp();
  };

  int stage = 0;
  uint32_t phase = 0;
  int traced_blocks = 0;
  for (int pi = 0;; ++pi) {
    top_up(false);
    while (ld_acquire_smem(tail_addr) <= pi) {
    }
    const int4 e0 = pq[(pi & (kSkPartQueue - 1)) * 2], e1 = pq[(pi & (kSkPartQueue - 1)) * 2 + 1];
    const int seq = e0.x, hk = e0.y, seq_len = e0.z, jb = e0.w;
    const int cnt = e1.x, rot = e1.y, slot = e1.z, chx = e1.w;
    if (cnt == 0) break;
    if (threadIdx.x == 0) st_release_smem(flags_addr + 4 * kSkFlHead, pi + 1);
    traced_blocks += cnt;
    const int kv_head = q_tiles == 1 ? hk : hk / q_tiles;
    const int qt = hk - kv_head * q_tiles;

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
    reset_state();


    // this warp's blocks of the part: i with (i + rot) % 4 == warp
    for (int i = (warp - rot) & 3; i < cnt; i += kMmaWarps) {
      const uint32_t kb = my_ring + stage * STAGE;
      const uint32_t vb = kb + CHUNK;
      const int blk_tok0 = (jb + i) * B;       // first token of this block
      const int n_valid = seq_len - blk_tok0;  // >= 1; < B only in the last block
      top_up(true);
      mbar_wait(my_full + stage * 8, phase);

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

      // ================= the stage is free: refill it with the next block of this warp's stream =================
      __syncwarp();
      ++consumed;
      top_up(false);
      if (++stage == NST) {
        stage = 0;
        phase ^= 1;
      }
    }

    // ---- end of the part: merge the four warps' states; warp 0 owns the result ------------------------
#pragma unroll
    for (int r = 0; r < NG; ++r) {
      l_run[r] += __shfl_xor_sync(FULL, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(FULL, l_run[r], 2);
    }
    tree_merge();

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
    if (slot < 0) {  // the whole item was in this chunk
      if (warp == 0) write_out();
      continue;
    }
    // ---- the item is cut by a chunk boundary: leave the partial; the last part to arrive merges them all ----
    const int ch = chx & 0xffff, span = chx >> 16;  // parts live in chunks ch .. ch + span
    const bool coop = span >= 3;                    // many parts: all four warps merge (uniform: from the descriptor)
    int last = 0;
    if (warp == 0) {
      store_frag(p.sk_slots + (size_t)slot * SLOT);
      __threadfence();
      __syncwarp();
      if (lane == 0) last = atomicAdd(p.sk_counters + ch, 1u) == (unsigned)span;
      last = __shfl_sync(FULL, last, 0);
      if (coop && lane == 0) flags[kSkFlFix] = last;
    }
    if (coop) {
      named_bar_sync(kSkBarFix, 128);  // (also publishes the decision)
      last = *reinterpret_cast<volatile int*>(flags + kSkFlFix);
    }
    if (!last) continue;  // (warps 1-3 of a non-cooperative part always continue)
    __threadfence();
    reset_state();
    {
      const int step = coop ? kMmaWarps : 1;
      int c2 = ch + (coop ? warp : 0);
      if constexpr (NO <= 32) {  // two slots per round: twice the loads in flight
        for (; c2 + step <= ch + span; c2 += 2 * step)
          merge_frag2(p.sk_slots + (size_t)sk_slot_index(c2, c2 != ch) * SLOT,
                      p.sk_slots + (size_t)sk_slot_index(c2 + step, true) * SLOT);
      }
      for (; c2 <= ch + span; c2 += step)
        merge_frag(p.sk_slots + (size_t)sk_slot_index(c2, c2 != ch) * SLOT, ld_glob);
    }
    if (coop) tree_merge();
    if (warp == 0) {
      write_out();
      if (lane == 0) p.sk_counters[ch] = 0u;  // self-resetting: the next launch starts from zero
    }
  }
  if (p.sk_trace && threadIdx.x == 0) {
    long long t_end;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_end));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    long long* tr = p.sk_trace + (size_t)cta * 4;
    tr[0] = smid;
    tr[1] = t_begin;
    tr[2] = t_end;
    tr[3] = traced_blocks;
  }
}

}  // namespace b200pa
