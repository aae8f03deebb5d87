// context_attention_fwd on the 5th-generation tensor cores (sm_100a): tcgen05.mma with the S, P and O operands in
// Tensor Memory.  Same contract as prefill.cu (which stays the family for shapes this file does not instantiate);
// replaces vllm/attention/ops/prefix_prefill.py:10-245, 442-681.
//
//   * one CTA = 128 query rows of one (sequence, query head) = the UMMA M; softmax thread t owns row t (TMEM lane t),
//     so row maxima and sums are thread-local — no shuffles;
//   * S[128 x 64] = Q K^T : D/16 tcgen05.mma (M128 N64 K16, kind::f16, fp32 accumulate) issued by ONE thread,
//     A = Q tile and B = K tile in shared memory in the canonical no-swizzle K-major core-matrix layout
//     (8 rows x 16 B contiguous) — the paged K chunk [D/8][B][8] already IS that layout per block, the staging
//     copy only re-bases the 8-token groups of several blocks to one uniform stride;
//   * softmax: tcgen05.ld the row (32x32b), p = 2^(s * scale * log2e - m_ref) with one FFMA + MUFU per element, P
//     rounded to the 16-bit type and written with tcgen05.st OVER ITS OWN S COLUMNS (two tokens per 32-bit cell):
//     the second MMA reads its A operand straight from Tensor Memory.  The reference maximum m_ref is lazy: it only
//     moves when the row maximum grows by more than 2^8, so O (in TMEM) is rescaled a few times per row at most;
//   * O[128 x D] += P V : 4 tcgen05.mma, B = V tile K-major for the cached prefix (the paged V chunk [D][B] is
//     token-minor) and MN-major for the new tokens ([token][d] rows) — nothing is ever transposed;
//   * K/V tiles are staged with 16-byte cp.async straight into the core-matrix layouts (tokens behind ctx_len are
//     zero-filled by the copy's src-size); MMA completion is signalled with tcgen05.commit -> mbarrier.
//
// Two kernels share these pieces:
//   context_attention_ws_kernel (default)  warp-specialised: 4 softmax warps + 1 warp that stages Q/K and issues
//                                          every MMA + 1 warp that stages V; the roles only meet on mbarriers;
//   context_attention_tc_kernel            the same pipeline with all four warps doing everything (two CTA-wide
//                                          barriers per tile); kept selectable with B200PA_PREFILL_WS=0.
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched from the driver at run time)

#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "prefill_common.cuh"

namespace b200pa {
namespace prefill_tc {

// Developer-only phase timestamps of CTA (0, 0, 0) (tools/prefill_trace.py): compiled out of the product build.
#ifdef B200PA_PREFILL_TRACE
__device__ long long g_trace[128 * 16];
#define B200PA_TRACE(it, slot)                                                                   \
  do {                                                                                           \
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x & 31) == 0 && (it) < 128) \
      g_trace[(it) * 16 + (slot)] = clock64();                                                   \
  } while (0)
#else
#define B200PA_TRACE(it, slot) do { } while (0)
#endif

constexpr int kThreads = 128;
constexpr int kBM = 128;
constexpr int kBN = 64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "TC_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra TC_DONE;\n"
      "bra TC_WAIT;\n"
      "TC_DONE:\n"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// Same wait for the copy / MMA-issue warps, which spend most of their time waiting: back off between polls so that
// the spin does not take issue slots from the softmax warps that share their scheduler.
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "TC_RWAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra TC_RDONE;\n"
      "nanosleep.u32 64;\n"
      "bra TC_RWAIT;\n"
      "TC_RDONE:\n"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// The MMA-issue warp is the critical path of the warp-specialised kernel (every tensor-pipe instruction goes through
// it): a tight poll (B200PA_PREFILL_MMA_SPIN=1 at build time) measured no gain over the relaxed wait (period 2208 vs 2202 clocks), so the back-off stays.
// Alternating the two heads' (independent) MMA chains instead of issuing each head's chain back to back measured
// slightly SLOWER (871 vs 890 TFLOP/s, period 2288 vs 2208 clocks; profiles/r02_prefill_experiments.log): the ~30
// clocks every tcgen05.mma costs on top of its math time (QK N64: 62-70, PV N128: ~94) are per instruction, not a
// dependent-issue latency.  Kept as a build-time knob, default off.
#ifndef B200PA_PREFILL_INTERLEAVE
#define B200PA_PREFILL_INTERLEAVE 0
#endif
#ifndef B200PA_PREFILL_MMA_SPIN
#define B200PA_PREFILL_MMA_SPIN 0
#endif
__device__ __forceinline__ void mbar_wait_mma(uint32_t bar, uint32_t parity) {
#if B200PA_PREFILL_MMA_SPIN
  mbar_wait(bar, parity);
#else
  mbar_wait_relaxed(bar, parity);
#endif
}
// MMA completion -> mbarrier (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// One leader lane of a converged warp.  A branch on elect.sync tells the compiler exactly one thread is active, so
// the uniform-datapath tcgen05 instructions under it are emitted straight; under `if (lane == 0)` every one of them
// is wrapped in an ELECT / BRA.U.ANY loop (~13 SASS instructions, ~60 clocks per MMA, measured).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}"
      : "=r"(pred));
  return pred != 0;
}
// D[tmem] (+)= A[smem desc] . B[smem desc]
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem desc]: A = 128 lanes x (K/2) columns, each 32-bit cell two consecutive K elements
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread i of the warp gets columns [c, c+32) of lane (base + i)
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
// ---- TMA staging of full cached-prefix tiles ----------------------------------------------------------
// expect_tx without an arrival: the loader still arrives once per tile at the top of that tile's iteration (as it
// does for cp.async tiles), the phase completes when the arrivals AND the announced bytes are in.
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// 1-D bulk copy (SASS UBLKCP): one contiguous run of the paged K chunk
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// 4-D tensor-map copy (SASS UTMALDG): box {8 tokens, DP rows} of the paged V chunk [nb][Hkv][D][B] lands as DP rows
// of 16 bytes = one K-major core-matrix column of the P.V B operand; rows >= head_size are zero-filled by the TMA.
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}

__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor, no swizzle (cute/arch/mma_sm100_desc.hpp SmemDescriptor):
// start address >> 4 in [0,14), leading byte offset >> 4 in [16,30), stride byte offset >> 4 in [32,46),
// version = 1 in [46,48), layout type 0 (SWIZZLE_NONE) in [61,64).
// K-major : SBO = bytes between 8-row groups (M/N direction), LBO = bytes between the two 16-byte K chunks.
// MN-major: SBO = bytes between 8-element groups along M/N, LBO = bytes between groups of 8 K rows.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// Instruction descriptor (UMMA::InstrDescriptor): c_format F32 = 1 at [4,6), a/b format (F16 0 | BF16 1) at
// [7,10) / [10,13), a_major [15], b_major [16] (0 = K, 1 = MN), N >> 3 at [17,23), M >> 4 at [24,29).
template <typename T>
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int b_mn_major) {
  const uint32_t fmt = std::is_same<T, __half>::value ? 0u : 1u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
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

// Packed fp32x2 arithmetic (sm_100 FFMA2 / FADD2): two softmax elements per instruction.
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{\n.reg .b64 ra, rb, rc, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rb, {%4, %4};\nmov.b64 rc, {%5, %5};\n"
      "fma.rn.f32x2 rd, ra, rb, rc;\nmov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1) {
  asm("{\n.reg .b64 ra, rd;\nmov.b64 ra, {%2, %3};\nmov.b64 rd, {%0, %1};\nadd.rn.f32x2 rd, rd, ra;\n"
      "mov.b64 {%0, %1}, rd;\n}"
      : "+f"(d0), "+f"(d1)
      : "f"(a0), "f"(a1));
}

template <int DP>
struct Geometry {
  static constexpr int QBYTES = kBM * DP * 2;  // [D/8][128 rows][16 B]
  static constexpr int KBYTES = kBN * DP * 2;  // [D/8][BN tokens][16 B]
  static constexpr int VBYTES = kBN * DP * 2;  // prefix: [BN/8][D rows][16 B]; new: [BN/8][D/8][8 tokens][16 B]
  static constexpr int VSTAGES = 3;            // V(i) in use by PV(i), PV(i-1) possibly still running, V(i+1) landing
  static constexpr int SMEM = QBYTES + 2 * KBYTES + VSTAGES * VBYTES + 64;  // + barriers and the TMEM base slot
  static constexpr int TMEM_COLS = 256;        // S0 [0,64) | S1 [64,128) | O [128, 128 + DP)
};

// Software pipeline of one CTA (tile i = 64 KV tokens):
//   iteration i:  wait K(i+1), V(i) landed -> issue S(i+1) = Q K(i+1)^T   (tensor pipe works while we do softmax(i))
//                 wait S(i), wait PV(i-1)  -> issue cp.async K(i+2), V(i+1)
//                 softmax(i) from TMEM, P(i) -> shared                    -> issue O += P(i) V(i), no wait
// so the tensor pipe always has the next QK^T queued behind the current PV, and the copies of a tile have a
// whole iteration to land.  S is double buffered in TMEM, K and V each double buffered in shared memory.
template <typename T, int DP>
__global__ void __launch_bounds__(kThreads, 2) context_attention_tc_kernel(const __grid_constant__ PrefillParams p) {
  using G = Geometry<DP>;
  constexpr int BM = kBM, BN = kBN;
  constexpr int DG = DP / 8;
  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;

  const int head = blockIdx.x;
  const int mt_idx = gridDim.y - 1 - blockIdx.y;  // heavy row tiles first
  const int seq = blockIdx.z;
  const int ctx_len = p.ctx_lens[seq];
  const int q_len = p.seq_lens[seq] - ctx_len;
  const int m0 = mt_idx * BM;
  if (m0 >= q_len) return;  // CTA-uniform, before any allocation
  const int64_t tok0 = p.start_loc[seq];
  const int kv_head = head / (p.num_heads / p.num_kv_heads);
  const int D = p.head_size;
  const int W = p.alibi_slopes ? 0 : p.sliding_window;
  const int bshift = p.block_shift;
  const int B = 1 << bshift;

  int c_first = 0, j_first = 0;
  if (W > 0) {
    c_first = max(0, ctx_len + m0 - W + 1) / BN;
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
  const uint32_t k_smem = q_smem + G::QBYTES;
  const uint32_t v_smem = k_smem + 2 * G::KBYTES;
  const uint32_t bar_s = v_smem + G::VSTAGES * G::VBYTES;  // two barriers (S0, S1); region sizes are multiples of 128
  const uint32_t bar_v = bar_s + 16;                       // one per V stage: "the PV that read this stage is done"
  constexpr int kSlotOff = G::QBYTES + 2 * G::KBYTES + G::VSTAGES * G::VBYTES + 48;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kSlotOff);

  // ---- one-time setup: TMEM allocation (warp 0), mbarriers (thread 32) ----------------------------------
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "n"(G::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else if (tid == 32) {
    mbar_init(bar_s, 1);
    mbar_init(bar_s + 8, 1);
    for (int s = 0; s < G::VSTAGES; ++s) mbar_init(bar_v + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  // ---- staging: 16-byte cp.async straight into the UMMA core-matrix layouts ---------------------------
  // Every thread always copies the same token (K: tid % 64; cached V: the 8-token group tid % 8), so it needs
  // ONE block-table entry per tile and copy kind; those are fetched one pipeline step ahead (phys_k / phys_v).
  const int kt = tid % BN;        // K / new-V token inside the tile
  const int vg = (tid & 7) * 8;   // first token of this thread's 8-token group (cached V)
  auto k_entry = [&](int it) -> int {  // physical block of this thread's K token in tile `it` (prefix tiles only)
    const int kpos = (c_first + it) * BN + kt;
    return (it < n_ctx && kpos < ctx_len) ? table[kpos >> bshift] : -1;
  };
  auto v_entry = [&](int it) -> int {
    const int kpos = (c_first + it) * BN + vg;
    return (it < n_ctx && kpos < ctx_len) ? table[kpos >> bshift] : -1;
  };
  auto load_k = [&](int it, int phys) {
    if (it >= n_tiles) return;
    const uint32_t dst = k_smem + (it & 1) * G::KBYTES + kt * 16;  // (token t, d-group g) -> g * (BN * 16) + t * 16
    if (it < n_ctx) {
      const int kpos = (c_first + it) * BN + kt;
      const T* src = kc + (int64_t)max(phys, 0) * p.kc_block_stride + (kpos & (B - 1)) * 8;
#pragma unroll
      for (int g = tid / BN; g < DG; g += kThreads / BN) {
        const bool ok = phys >= 0 && g * 8 < D;
        cp_async16(dst + g * (BN * 16), ok ? (const void*)(src + (int64_t)g * B * 8) : (const void*)kc, ok ? 16 : 0);
      }
    } else {
      const int j = (j_first + it - n_ctx) * BN + kt;
      const T* src = kn + (int64_t)j * p.k_stride_t;
#pragma unroll
      for (int g = tid / BN; g < DG; g += kThreads / BN) {
        const bool ok = j < q_len && g * 8 < D;
        cp_async16(dst + g * (BN * 16), ok ? (const void*)(src + g * 8) : (const void*)kn, ok ? 16 : 0);
      }
    }
  };
  auto load_v = [&](int it, int phys) {
    if (it >= n_tiles) return;
    const uint32_t vb = v_smem + (it % G::VSTAGES) * G::VBYTES;
    if (it < n_ctx) {
      // (row d, 8 tokens starting at vg) -> (vg / 8) * (DP * 16) + d * 16     [K-major B operand, N = d]
      const int kpos = (c_first + it) * BN + vg;
      const T* src = vc + (int64_t)max(phys, 0) * p.vc_block_stride + (kpos & (B - 1));
      const uint32_t dst = vb + (vg >> 3) * (DP * 16);
#pragma unroll
      for (int d = tid >> 3; d < DP; d += kThreads / 8) {
        const bool ok = phys >= 0 && d < D;
        cp_async16(dst + d * 16, ok ? (const void*)(src + (int64_t)d * B) : (const void*)vc, ok ? 16 : 0);
      }
    } else {
      // MN-major core matrices: 8 tokens x 16 B (8 d values); d-groups 128 B apart, token groups DP * 16 apart
      const int j = (j_first + it - n_ctx) * BN + kt;
      const T* src = vn + (int64_t)j * p.v_stride_t;
      const uint32_t dst = vb + (kt >> 3) * (DP * 16) + (kt & 7) * 16;
#pragma unroll
      for (int g = tid / BN; g < DG; g += kThreads / BN) {
        const bool ok = j < q_len && g * 8 < D;
        cp_async16(dst + g * 128, ok ? (const void*)(src + g * 8) : (const void*)vn, ok ? 16 : 0);
      }
    }
  };

  {  // Q: (row r, d-group c) -> c * (BM * 16) + r * 16; rows >= q_len and columns >= head_size zero-filled
    const T* qb = static_cast<const T*>(p.q) + tok0 * p.q_stride_t + (int64_t)head * p.q_stride_h;
    const bool rv = m0 + tid < q_len;
    const T* src = qb + (int64_t)(m0 + tid) * p.q_stride_t;
#pragma unroll
    for (int c = 0; c < DG; ++c) {
      const bool ok = rv && c * 8 < D;
      cp_async16(q_smem + c * (BM * 16) + tid * 16, ok ? (const void*)(src + c * 8) : (const void*)qb, ok ? 16 : 0);
    }
  }
  load_k(0, k_entry(0));
  cp_async_commit();               // group: Q, K(0)
  load_k(1, k_entry(1));
  load_v(0, v_entry(0));
  cp_async_commit();               // group: K(1), V(0)
  int phys_k = k_entry(2);         // for the copies issued in iteration 0
  int phys_v = v_entry(1);

  cp_async_wait<1>();  // Q, K(0)
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_o = tmem_base + 2 * BN;              // O: columns [128, 128 + DP)
  const uint32_t lane_off = (uint32_t)(warp * 32) << 16;   // this warp's TMEM lanes

  constexpr uint32_t idesc_qk = make_idesc<T>(BM, BN, 0);
  constexpr uint32_t idesc_pv_k = make_idesc<T>(BM, DP, 0);
  constexpr uint32_t idesc_pv_mn = make_idesc<T>(BM, DP, 1);

  auto issue_qk = [&](int it) {  // S(it & 1) = Q K(it)^T   (one thread)
    const uint32_t kb = k_smem + (it & 1) * G::KBYTES;
#pragma unroll
    for (int ks = 0; ks < DP / 16; ++ks) {
      const uint64_t a = make_desc(q_smem + ks * 2 * (BM * 16), BM * 16, 128);
      const uint64_t b = make_desc(kb + ks * 2 * (BN * 16), BN * 16, 128);
      tc_mma(tmem_base + (it & 1) * BN, a, b, idesc_qk, ks > 0 ? 1u : 0u);
    }
    tc_commit(bar_s + 8 * (it & 1));
  };
  if (tid == 0) issue_qk(0);

  const float slope_l2 = p.alibi_slopes ? p.alibi_slopes[head] * kLog2e : 0.f;
  const bool has_alibi = p.alibi_slopes != nullptr;
  const int m = m0 + tid;  // this thread's query row
  float m_ref = kNegBig, l_run = 0.f;

  for (int it = 0; it < n_tiles; ++it) {
    const bool is_ctx = it < n_ctx;
    const int t0 = is_ctx ? (c_first + it) * BN : (j_first + it - n_ctx) * BN;
    const uint32_t vb = v_smem + (it % G::VSTAGES) * G::VBYTES;

    // ---- (a) K(it+1) and V(it) have landed ----------------------------------------------------------
    cp_async_wait<0>();
    if (is_ctx && t0 + BN > ctx_len) {
      // last prefix tile: zero the V columns of tokens >= ctx_len (stale cache bits could be NaN; 0 * NaN = NaN)
      __syncthreads();
      const int nvalid = ctx_len - t0, cnt = BN - nvalid;
      uint8_t* vrow = smem + G::QBYTES + 2 * G::KBYTES + (it % G::VSTAGES) * G::VBYTES;
      for (int idx = tid; idx < DP * cnt; idx += kThreads) {
        const int d = idx / cnt, c = nvalid + (idx - d * cnt);
        *reinterpret_cast<uint16_t*>(vrow + (c >> 3) * (DP * 16) + d * 16 + (c & 7) * 2) = 0;
      }
    }
    fence_proxy_async();  // this thread's staged bytes -> visible to the tensor-core (async) proxy
    __syncthreads();

    // ---- (b) queue the NEXT tile's QK^T behind the previous PV ------------------------------------------
    if (tid == 0 && it + 1 < n_tiles) {
      tc_fence_after();
      issue_qk(it + 1);
    }
    // ---- (c, d) S(it) ready (K(it)'s stage is free); PV(it-2) done long ago (its V stage is free) --------
    mbar_wait(bar_s + 8 * (it & 1), (it >> 1) & 1);
    if (it >= 2) mbar_wait(bar_v + 8 * ((it - 2) % G::VSTAGES), ((it - 2) / G::VSTAGES) & 1);
    tc_fence_after();

    // ---- (e) copies for the following iteration --------------------------------------------------------
    load_k(it + 2, phys_k);
    load_v(it + 1, phys_v);
    cp_async_commit();
    phys_k = k_entry(it + 3);
    phys_v = v_entry(it + 2);

    // ---- (f) softmax on this thread's row ------------------------------------------------------------
    uint32_t sr[BN];
    {
      uint32_t (&lo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[0]);
      uint32_t (&hi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[32]);
      const uint32_t ts = tmem_base + (it & 1) * BN + lane_off;
      tc_ld32(ts, lo);
      tc_ld32(ts + 32, hi);
      tc_wait_ld();
    }
    const int base_rel = is_ctx ? (t0 - ctx_len) : t0;
    const int col_limit = is_ctx ? (ctx_len - t0) : BN;
    bool need_mask;
    if (is_ctx)
      need_mask = (t0 + BN > ctx_len) || (W > 0 && ctx_len + m0 + BM - 1 - t0 >= W);
    else
      need_mask = (t0 + BN - 1 > m0) || (W > 0 && m0 + BM - 1 - t0 >= W);
    float mx = kNegBig;
    float mult;  // p = 2^(sr * mult - m_ref)
    if (need_mask || has_alibi) {
#pragma unroll
      for (int c = 0; c < BN; ++c) {
        const int rel = base_rel + c - m;
        float z = __uint_as_float(sr[c]) * p.scale_log2;
        if (has_alibi) z = fmaf(slope_l2, (float)rel, z);
        const bool vis = !need_mask || (c < col_limit && rel <= 0 && (W <= 0 || -rel < W));
        z = vis ? z : kNegBig;
        sr[c] = __float_as_uint(z);
        mx = fmaxf(mx, z);
      }
      mult = 1.f;
    } else {  // interior tile: the maximum commutes with the (positive) scale, which then rides in the FFMA
#pragma unroll
      for (int c = 0; c < BN; ++c) mx = fmaxf(mx, __uint_as_float(sr[c]));
      mx *= p.scale_log2;
      mult = p.scale_log2;
    }
    // Lazy reference maximum: only move it (and rescale O in TMEM) when the row maximum grew by more than 2^8;
    // until then p = 2^(z - m_ref) <= 256, harmless for the 16-bit P and the fp32 sums.
    float alpha = 1.f;
    if (mx > m_ref + 8.f) {
      alpha = ex2(m_ref - mx);
      m_ref = mx;
    }
    if (it > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
      // the only place that needs PV(it-1) finished before the next PV is issued: O is about to be rewritten
      mbar_wait(bar_v + 8 * ((it - 1) % G::VSTAGES), ((it - 1) / G::VSTAGES) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < DP; c0 += 32) {
        uint32_t o[32];
        tc_ld32(tmem_o + lane_off + c0, o);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
        tc_st32(tmem_o + lane_off + c0, o);
      }
      tc_wait_st();
    }
    float sum = 0.f;
    const float neg_ref = -m_ref;
    // P (rounded to the 16-bit type) goes back into the TMEM columns its S row came from — two tokens per 32-bit
    // cell — and is the A operand of the second MMA straight from Tensor Memory: no shared-memory round trip,
    // and P is double buffered for free because S is.
    uint32_t pw[BN / 2];
#pragma unroll
    for (int c = 0; c < BN / 2; ++c) {
      const float p0 = ex2(fmaf(__uint_as_float(sr[2 * c]), mult, neg_ref));
      const float p1 = ex2(fmaf(__uint_as_float(sr[2 * c + 1]), mult, neg_ref));
      sum += p0 + p1;
      pw[c] = pack2<T>(p0, p1);
    }
    tc_st32(tmem_base + (it & 1) * BN + lane_off, pw);
    tc_wait_st();
    l_run = l_run * alpha + sum;
    tc_fence_before();
    __syncthreads();

    // ---- (g) O += P V (one thread issues; nobody waits here) -------------------------------------------
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < BN / 16; ++kk) {
        const uint64_t b = make_desc(vb + kk * 2 * (DP * 16), DP * 16, 128);
        tc_mma_ts(tmem_o, tmem_base + (it & 1) * BN + kk * 8, b, is_ctx ? idesc_pv_k : idesc_pv_mn,
                  (it > 0 || kk > 0) ? 1u : 0u);
      }
      tc_commit(bar_v + 8 * (it % G::VSTAGES));
    }
  }
  mbar_wait(bar_v + 8 * ((n_tiles - 1) % G::VSTAGES), ((n_tiles - 1) / G::VSTAGES) & 1);
  tc_fence_after();

  // ---- epilogue: O / l -> out -------------------------------------------------------------------------
  {
    const float inv = 1.f / l_run;
    T* orow = static_cast<T*>(p.out) + (tok0 + m) * p.o_stride_t + (int64_t)head * p.o_stride_h;
#pragma unroll 1
    for (int c0 = 0; c0 < DP; c0 += 32) {
      uint32_t o[32];
      tc_ld32(tmem_o + lane_off + c0, o);
      tc_wait_ld();
      if (m < q_len) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          if (c0 + j < D) {
            uint4 w;
            w.x = pack2<T>(__uint_as_float(o[j]) * inv, __uint_as_float(o[j + 1]) * inv);
            w.y = pack2<T>(__uint_as_float(o[j + 2]) * inv, __uint_as_float(o[j + 3]) * inv);
            w.z = pack2<T>(__uint_as_float(o[j + 4]) * inv, __uint_as_float(o[j + 5]) * inv);
            w.w = pack2<T>(__uint_as_float(o[j + 6]) * inv, __uint_as_float(o[j + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c0 + j) = w;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(G::TMEM_COLS) : "memory");
  }
}

// ====================================================================================================
// Warp-specialised variant: warps 0-3 do nothing but softmax (thread t = query row t = TMEM lane t), warp 4
// stages K/V (all 32 lanes, 16-byte cp.async into the core-matrix layouts) and issues every MMA (lane 0).
// The roles meet only on mbarriers:
//   s_full[2]  : S(i) = Q K(i)^T finished                      (tcgen05.commit)     softmax waits, producer waits (K stage free)
//   p_full[2]  : all 128 softmax threads stored P(i) to TMEM   (128 arrivals)       producer waits before issuing P(i) V(i)
//   pv_done[3] : P(i) V(i) finished, per V stage               (tcgen05.commit)     producer waits (V stage free); softmax waits
//                                                                                    only before rescaling O and in the epilogue
// No __syncthreads in the main loop.  Tensor-pipe order: QK(0), QK(1), PV(0), QK(2), PV(1), ... so the next S is
// always being computed while the softmax warps work on the current one.
// ====================================================================================================
constexpr int kWsProducerWarps = 3;  // warp 4 NQ: MMA issue; 4 NQ + 1: V copies; 4 NQ + 2: Q / K copies

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Asynchronous arrive: fires when every cp.async this thread has issued so far has completed (counts as one of the
// barrier's expected arrivals).  The loader never waits for a tile to land: it issues, posts this, and moves on.
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

// NQ = query heads per CTA.  With NQ = 2 the CTA serves two query heads of one GQA group against the SAME staged
// K/V tiles (the L2 -> shared-memory copies, not the math, bound the NQ = 1 kernel: removing them took it from
// 640 to 1025 TFLOP/s), with 8 softmax warps, all 512 TMEM columns and one CTA per SM.
// ROWS = false: the NQ tiles are the same 128 rows of NQ adjacent heads (GQA groups with an even head count);
// ROWS = true: they are NQ consecutive 128-row tiles of ONE head (MHA / odd groups): the CTA walks the K/V range of
// its last tile and the earlier tile sees the up to two extra causal tiles fully masked (p = 0).
template <typename T, int DP, int NQ, bool ROWS>
__global__ void __launch_bounds__(NQ * 128 + 32 + 64 * NQ, NQ == 1 ? 2 : 1)
    context_attention_ws_kernel(const __grid_constant__ PrefillParams p, const __grid_constant__ CUtensorMap vmap) {
  using G = Geometry<DP>;
  constexpr int QBYTES = NQ * G::QBYTES;
  constexpr int TMEM_COLS = NQ * 256;
  constexpr int LW = NQ;  // loader warps per operand
  constexpr int MMA_WARP = 4 * NQ, V_WARP = 4 * NQ + 1, K_WARP = V_WARP + LW;
  constexpr int BM = kBM, BN = kBN;
  constexpr int DG = DP / 8;
  extern __shared__ __align__(128) uint8_t smem[];

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  const int head = blockIdx.x * (ROWS ? 1 : NQ);  // first of the NQ adjacent query heads (same KV head)
  constexpr int CTA_ROWS = ROWS ? NQ * kBM : kBM;
  const int mt_idx = gridDim.y - 1 - blockIdx.y;
  const int seq = blockIdx.z;
  const int ctx_len = p.ctx_lens[seq];
  const int q_len = p.seq_lens[seq] - ctx_len;
  const int m0 = mt_idx * CTA_ROWS;  // first query row of the CTA
  if (m0 >= q_len) return;
  const int64_t tok0 = p.start_loc[seq];
  const int kv_head = head / (p.num_heads / p.num_kv_heads);
  const int D = p.head_size;
  const int W = p.alibi_slopes ? 0 : p.sliding_window;
  const int bshift = p.block_shift;
  const int B = 1 << bshift;

  int c_first = 0, j_first = 0;
  if (W > 0) {
    c_first = max(0, ctx_len + m0 - W + 1) / BN;
    j_first = max(0, m0 - W + 1) / BN;
  }
  const int c_end = (ctx_len + BN - 1) / BN;
  const int j_end = (min(q_len, m0 + CTA_ROWS) + BN - 1) / BN;
  const int n_ctx = max(0, c_end - c_first);
  const int n_tiles = n_ctx + (j_end - j_first);
  // Cached-prefix tiles that lie completely below ctx_len are staged by TMA (bulk copies for K, tensor-map boxes for
  // V); the last, partial prefix tile and the new-token tiles keep the cp.async path (it zero-fills behind ctx_len).
  const bool use_tma = p.use_tma != 0;
  auto tma_tile = [&](int it) -> bool { return use_tma && it < n_ctx && (c_first + it + 1) * BN <= ctx_len; };

  const uint32_t q_smem = smem_u32(smem);
  const uint32_t k_smem = q_smem + QBYTES;
  const uint32_t v_smem = k_smem + 2 * G::KBYTES;
  const uint32_t bars = v_smem + G::VSTAGES * G::VBYTES;
  // s_full[NQ][2] | p_full[NQ][2] | pv_done[3] | v_full[3] | k_full[2]   (8 bytes each)
  const uint32_t s_full = bars, p_full = bars + 16 * NQ, pv_done = bars + 32 * NQ, v_full = pv_done + 24;
  const uint32_t k_full = v_full + 24;
  constexpr int kSlotOff = QBYTES + 2 * G::KBYTES + G::VSTAGES * G::VBYTES + 32 * NQ + 64 + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kSlotOff);

  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else if (tid == 0) {
    for (int i = 0; i < 2 * NQ; ++i) {
      mbar_init(s_full + 8 * i, 1);
      mbar_init(p_full + 8 * i, 128);
    }
    for (int s = 0; s < G::VSTAGES; ++s) mbar_init(pv_done + 8 * s, 1);
    // every lane of every loader warp arrives once per tile (asynchronously for cp.async tiles)
    for (int s = 0; s < G::VSTAGES; ++s) mbar_init(v_full + 8 * s, 32 * LW);
    mbar_init(k_full, 32 * LW);
    mbar_init(k_full + 8, 32 * LW);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= K_WARP) {
    const int lw = warp - K_WARP;  // this loader's share: token halves / Q row groups lw, lw + LW, ...
    // ================================ Q / K loader ================================
    // (A copy warp blocks in the cp.async issue while the L2 -> shared-memory path is busy -- measured ~1000 clocks
    // per 16 KB K tile -- so the copies live in warps of their own and the MMA warp below only issues MMAs.)
    const T* kc = static_cast<const T*>(p.k_cache) + (int64_t)kv_head * p.kc_head_stride;
    const T* kn = static_cast<const T*>(p.k) + tok0 * p.k_stride_t + (int64_t)kv_head * p.k_stride_h;
    const int* table = p.block_tables + (int64_t)seq * p.block_table_stride;

    // lane l copies tokens l and l + 32 of a K tile, so it needs two block-table entries per tile; they are fetched
    // one refill ahead (ek0 / ek1) so that no copy ever waits on a dependent global load.
    auto k_entry = [&](int it, int h) -> int {
      const int kpos = (c_first + it) * BN + lane + 32 * h;
      return (it < n_ctx && kpos < ctx_len) ? table[kpos >> bshift] : -1;
    };
    auto load_k = [&](int it, int e0, int e1) {
      if (it >= n_tiles) return;
      const uint32_t kb = k_smem + (it & 1) * G::KBYTES;
#pragma unroll
      for (int h = lw; h < 2; h += LW) {
        const int kt = lane + 32 * h;
        const uint32_t dst = kb + kt * 16;  // (token t, d-group g) -> g * (BN * 16) + t * 16
        if (it < n_ctx && p.kv_dtype != B200PA_KV_AUTO) {
          // fp8 cache: the token's 16-byte runs [D/16][B][16] are dequantised in registers (the decode kernels'
          // conversion) and stored as two 16-byte core-matrix rows each.  Plain stores: the fence.proxy.async +
          // mbarrier arrive that publish the cp.async tiles publish these too.
          const int phys = h ? e1 : e0;
          const int kpos = (c_first + it) * BN + kt;
          const uint8_t* src = static_cast<const uint8_t*>(p.k_cache) + (int64_t)kv_head * p.kc_head_stride +
                               (int64_t)max(phys, 0) * p.kc_block_stride + (kpos & (B - 1)) * 16;
          int4 raw[DP / 16];
#pragma unroll
          for (int c = 0; c < DP / 16; ++c) {
            raw[c] = make_int4(0, 0, 0, 0);
            if (phys >= 0 && c * 16 < D) raw[c] = ld_nc_v4(src + (int64_t)c * B * 16);
          }
#pragma unroll
          for (int c = 0; c < DP / 16; ++c) {
            const bool ok = phys >= 0 && c * 16 < D;
            const uint4 zero = make_uint4(0, 0, 0, 0);
            st_shared_v4(dst + (2 * c) * (BN * 16),
                         ok ? fp8x8_to_t<T>(p.kv_dtype, make_uint2(raw[c].x, raw[c].y), p.kv_scale) : zero);
            st_shared_v4(dst + (2 * c + 1) * (BN * 16),
                         ok ? fp8x8_to_t<T>(p.kv_dtype, make_uint2(raw[c].z, raw[c].w), p.kv_scale) : zero);
          }
        } else if (it < n_ctx) {
          const int phys = h ? e1 : e0;
          const int kpos = (c_first + it) * BN + kt;
          const T* src = kc + (int64_t)max(phys, 0) * p.kc_block_stride + (kpos & (B - 1)) * 8;
#pragma unroll
          for (int g = 0; g < DG; ++g) {
            const bool ok = phys >= 0 && g * 8 < D;
            cp_async16(dst + g * (BN * 16), ok ? (const void*)(src + (int64_t)g * B * 8) : (const void*)kc, ok ? 16 : 0);
          }
        } else {
          const int j = (j_first + it - n_ctx) * BN + kt;
          const T* src = kn + (int64_t)j * p.k_stride_t;
#pragma unroll
          for (int g = 0; g < DG; ++g) {
            const bool ok = j < q_len && g * 8 < D;
            cp_async16(dst + g * (BN * 16), ok ? (const void*)(src + g * 8) : (const void*)kn, ok ? 16 : 0);
          }
        }
      }
    };
    {  // Q: rows lane, lane + 32, lane + 64, lane + 96
      const T* qb = static_cast<const T*>(p.q) + tok0 * p.q_stride_t + (int64_t)head * p.q_stride_h;
#pragma unroll
      for (int h = lw; h < 4 * NQ; h += LW) {
        const int r = lane + 32 * (h & 3);
        const int jq = h >> 2;  // which of the CTA's heads
        const int qm = m0 + r + (ROWS ? jq * BM : 0);
        const bool rv = qm < q_len;
        const T* src = qb + (int64_t)qm * p.q_stride_t + (int64_t)(ROWS ? 0 : jq) * p.q_stride_h;
#pragma unroll
        for (int c = 0; c < DG; ++c) {
          const bool ok = rv && c * 8 < D;
          cp_async16(q_smem + jq * G::QBYTES + c * (BM * 16) + r * 16, ok ? (const void*)(src + c * 8) : (const void*)qb,
                     ok ? 16 : 0);
        }
      }
    }
    // ---- TMA path: a full prefix tile is (BN / run) x (D / 8) contiguous runs of run x 16 bytes, run = min(B, BN)
    // tokens of one d-group of one block ([D/8][B][8] chunk): one bulk copy each, this warp issues its 1/LW of them.
    const int rshift = min(bshift, 6);
    const int run_tok = 1 << rshift;
    const int DGv = D >> 3;                       // d-groups that exist (D % 16 == 0)
    const int Cw = ((BN >> rshift) * DGv) / LW;   // copies per tile and loader warp
    const uint32_t run_bytes = (uint32_t)run_tok * 16u;
    constexpr int KJ = 4;                         // ceil(max Cw / 32): B = 8, D = 128, LW = 1
    auto k_tma_entries = [&](int it, int (&ent)[KJ]) {
      const int t0 = (c_first + it) * BN;
#pragma unroll
      for (int j = 0; j < KJ; ++j) {
        const int i = lw * Cw + lane + 32 * j;
        ent[j] = (lane + 32 * j < Cw) ? table[(t0 + (i / DGv) * run_tok) >> bshift] : 0;
      }
    };
    auto issue_k_tma = [&](int it, const int (&ent)[KJ]) {
      const uint32_t kb = k_smem + (it & 1) * G::KBYTES;
      const uint32_t bar = k_full + 8 * (it & 1);
      const int t0 = (c_first + it) * BN;
      if (lane == 0) mbar_expect_tx(bar, (uint32_t)Cw * run_bytes);
      __syncwarp();
#pragma unroll
      for (int j = 0; j < KJ; ++j) {
        if (lane + 32 * j < Cw) {
          const int i = lw * Cw + lane + 32 * j;
          const int r = i / DGv, g = i - r * DGv;
          const int tok = t0 + r * run_tok;
          const T* src = kc + (int64_t)ent[j] * p.kc_block_stride + (int64_t)g * B * 8 + (tok & (B - 1)) * 8;
          bulk_g2s(kb + g * (BN * 16) + (r * run_tok) * 16, src, run_bytes, bar);
        }
      }
    };
    if (use_tma && DGv < DG) {
      // d-groups behind head_size are never written by the bulk copies: zero them once in both K stages (Q's padded
      // columns are zero too, but 0 x stale NaN bits would still poison S)
      for (int u = lane + 32 * lw; u < 2 * (DG - DGv) * BN; u += 32 * LW) {
        const int st = u / ((DG - DGv) * BN), rem = u - st * ((DG - DGv) * BN);
        st_shared_v4(k_smem + st * G::KBYTES + DGv * (BN * 16) + rem * 16, make_uint4(0, 0, 0, 0));
      }
      fence_proxy_async();
    }
    auto stage_k = [&](int it, int e0, int e1) {  // prologue tiles: entries fetched inline
      if (it < n_tiles && tma_tile(it)) {
        int ent[KJ];
        k_tma_entries(it, ent);
        issue_k_tma(it, ent);
      } else {
        load_k(it, e0, e1);
      }
    };
    // A tile is announced by an arrive of every lane on its k_full barrier, posted right after the lane has issued
    // its share and never waited for: cp.async tiles use the asynchronous arrive (fires when the lane's copies have
    // landed), fp8 tiles (plain shared stores) a proxy fence + arrive, TMA tiles an arrive next to the expect_tx.
    // (With a synchronous wait + arrive per tile the announce of a landed tile sat behind the ~1000-clock issue of the
    // next one and the MMA warp waited ~590 of every 2450 clocks for it: profiles/r02_prefill_phase_trace*.log.)
    const bool fp8_cache = p.kv_dtype != B200PA_KV_AUTO;
    auto announce_k = [&](int it) {
      if (it >= n_tiles) return;
      const uint32_t bar = k_full + 8 * (it & 1);
      if (tma_tile(it) || (fp8_cache && it < n_ctx)) {
        if (it == 0) {  // Q's cp.async copies ride on tile 0: a plain arrive must not overtake them
          cp_async_commit();
          cp_async_wait<0>();
        }
        fence_proxy_async();
        mbar_arrive(bar);
      } else {
        cp_async_arrive(bar);
      }
    };
    stage_k(0, k_entry(0, 0), k_entry(0, 1));  // (Q rides on K(0)'s announce)
    announce_k(0);
    stage_k(1, k_entry(1, 0), k_entry(1, 1));
    announce_k(1);
    int ek0 = k_entry(2, 0), ek1 = k_entry(2, 1);  // for the refill of iteration 0

    for (int it = 0; it + 2 < n_tiles; ++it) {
      // K(it+2) goes into the stage QK(it) reads: wait for the last head's QK(it).  That wait also implies every
      // loader lane has arrived for tile it, so an expect_tx below lands in tile it+2's phase of the barrier.
      const bool t2 = tma_tile(it + 2);
      int ent[KJ] = {0, 0, 0, 0};
      if (t2) k_tma_entries(it + 2, ent);  // block-table loads fly during the wait below
      mbar_wait_relaxed(s_full + 16 * (NQ - 1) + 8 * (it & 1), (it >> 1) & 1);
      if (t2) issue_k_tma(it + 2, ent);
      else load_k(it + 2, ek0, ek1);
      announce_k(it + 2);
      ek0 = k_entry(it + 3, 0);
      ek1 = k_entry(it + 3, 1);
    }
    cp_async_commit();
    cp_async_wait<0>();
  } else if (warp == MMA_WARP) {
    // ================================ MMA issuer ================================
    constexpr uint32_t idesc_qk = make_idesc<T>(BM, BN, 0);
    constexpr uint32_t idesc_pv_k = make_idesc<T>(BM, DP, 0);
    constexpr uint32_t idesc_pv_mn = make_idesc<T>(BM, DP, 1);
    auto issue_qk = [&](int it) {  // S_j(it & 1) = Q_j K(it)^T for every head j of the CTA
      const uint32_t kb = k_smem + (it & 1) * G::KBYTES;
#if B200PA_PREFILL_INTERLEAVE
      // the heads' accumulator chains are independent: alternate them, so that an MMA never queues right behind the
      // one it accumulates onto (and both heads consume one K slice back to back)
#pragma unroll
      for (int ks = 0; ks < DP / 16; ++ks) {
        const uint64_t b = make_desc(kb + ks * 2 * (BN * 16), BN * 16, 128);
#pragma unroll
        for (int jq = 0; jq < NQ; ++jq) {
          const uint64_t a = make_desc(q_smem + jq * G::QBYTES + ks * 2 * (BM * 16), BM * 16, 128);
          tc_mma(tmem_base + jq * 256 + (it & 1) * BN, a, b, idesc_qk, ks > 0 ? 1u : 0u);
        }
      }
#pragma unroll
      for (int jq = 0; jq < NQ; ++jq) tc_commit(s_full + 16 * jq + 8 * (it & 1));
#else
#pragma unroll
      for (int jq = 0; jq < NQ; ++jq) {
#pragma unroll
        for (int ks = 0; ks < DP / 16; ++ks) {
          const uint64_t a = make_desc(q_smem + jq * G::QBYTES + ks * 2 * (BM * 16), BM * 16, 128);
          const uint64_t b = make_desc(kb + ks * 2 * (BN * 16), BN * 16, 128);
          tc_mma(tmem_base + jq * 256 + (it & 1) * BN, a, b, idesc_qk, ks > 0 ? 1u : 0u);
        }
        tc_commit(s_full + 16 * jq + 8 * (it & 1));
      }
#endif
    };


    // (fence.proxy.async on the consumer side too: the loaders' asynchronous arrives carry no fence of their own)
    mbar_wait_mma(k_full, 0);  // Q, K(0)
    fence_proxy_async();
    tc_fence_after();
    if (elect_one()) issue_qk(0);
    __syncwarp();

    for (int it = 0; it < n_tiles; ++it) {
      B200PA_TRACE(it, 4);
      const bool is_ctx = it < n_ctx;
      const uint32_t vb = v_smem + (it % G::VSTAGES) * G::VBYTES;
      if (it + 1 < n_tiles) {
        mbar_wait_mma(k_full + 8 * ((it + 1) & 1), ((it + 1) >> 1) & 1);
        B200PA_TRACE(it, 5);
        fence_proxy_async();
        tc_fence_after();
        if (elect_one()) issue_qk(it + 1);
        __syncwarp();
      }
      B200PA_TRACE(it, 6);
      // V(it) has landed (V loader) and P(it) is in TMEM (softmax warps): O += P(it) V(it)
      mbar_wait_mma(v_full + 8 * (it % G::VSTAGES), (it / G::VSTAGES) & 1);
      fence_proxy_async();
      B200PA_TRACE(it, 8);
#if B200PA_PREFILL_INTERLEAVE
#pragma unroll
      for (int jq = 0; jq < NQ; ++jq) mbar_wait_mma(p_full + 16 * jq + 8 * (it & 1), (it >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint64_t b = make_desc(vb + kk * 2 * (DP * 16), DP * 16, 128);
#pragma unroll
          for (int jq = 0; jq < NQ; ++jq)
            tc_mma_ts(tmem_base + jq * 256 + 2 * BN, tmem_base + jq * 256 + (it & 1) * BN + kk * 8, b,
                      is_ctx ? idesc_pv_k : idesc_pv_mn, (it > 0 || kk > 0) ? 1u : 0u);
        }
        tc_commit(pv_done + 8 * (it % G::VSTAGES));  // covers every head's PV(it)
      }
      __syncwarp();
      B200PA_TRACE(it, 9);
      B200PA_TRACE(it, 10);
#else
#pragma unroll
      for (int jq = 0; jq < NQ; ++jq) {
        mbar_wait_mma(p_full + 16 * jq + 8 * (it & 1), (it >> 1) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < BN / 16; ++kk) {
            const uint64_t b = make_desc(vb + kk * 2 * (DP * 16), DP * 16, 128);
            tc_mma_ts(tmem_base + jq * 256 + 2 * BN, tmem_base + jq * 256 + (it & 1) * BN + kk * 8, b,
                      is_ctx ? idesc_pv_k : idesc_pv_mn, (it > 0 || kk > 0) ? 1u : 0u);
          }
          if (jq == NQ - 1) tc_commit(pv_done + 8 * (it % G::VSTAGES));  // covers every head's PV(it)
        }
        __syncwarp();
        B200PA_TRACE(it, 9 + jq);
      }
#endif
    }
  } else if (warp >= V_WARP) {
    const int lw = warp - V_WARP;
    // ================================ V loader ================================
    // Lane l always copies the 8-token group l % 8 (cached V: rows l / 8, l / 8 + 4, ...) or tokens l and l + 32
    // (new V), so it needs one block-table entry per tile, fetched a tile ahead.  Tokens >= ctx_len inside the last
    // group are zero-filled by the copy itself (src-size < 16): stale NaN bits behind the prefix never reach the
    // tensor core.
    const T* vc = static_cast<const T*>(p.v_cache) + (int64_t)kv_head * p.vc_head_stride;
    const T* vn = static_cast<const T*>(p.v) + tok0 * p.v_stride_t + (int64_t)kv_head * p.v_stride_h;
    const int* table = p.block_tables + (int64_t)seq * p.block_table_stride;
    const int vg = (lane & 7) * 8;
    auto v_entry = [&](int it) -> int {
      const int kpos = (c_first + it) * BN + vg;
      return (it < n_ctx && kpos < ctx_len) ? table[kpos >> bshift] : -1;
    };
    auto load_v = [&](int it, int phys) {
      if (it >= n_tiles) return;
      const uint32_t vb = v_smem + (it % G::VSTAGES) * G::VBYTES;
      if (it < n_ctx && p.kv_dtype != B200PA_KV_AUTO) {
        // fp8 cache: 8 tokens of one d row are 8 bytes; dequantise in registers, zero the tokens behind ctx_len
        const int kpos = (c_first + it) * BN + vg;
        const int nvalid = min(8, max(0, ctx_len - kpos));
        const uint8_t* src = static_cast<const uint8_t*>(p.v_cache) + (int64_t)kv_head * p.vc_head_stride +
                             (int64_t)max(phys, 0) * p.vc_block_stride + (kpos & (B - 1));
        const uint32_t dst = vb + (vg >> 3) * (DP * 16);
        constexpr int kRows = DP / (4 * LW);  // rows of this lane
        constexpr int kBatch = 8;
        const int d0 = (lane >> 3) + 4 * lw;
#pragma unroll 1
        for (int r0 = 0; r0 < kRows; r0 += kBatch) {
          uint2 raw[kBatch];
#pragma unroll
          for (int r = 0; r < kBatch; ++r) {
            const int d = d0 + (r0 + r) * 4 * LW;
            raw[r] = make_uint2(0, 0);
            if (phys >= 0 && d < D && nvalid > 0) raw[r] = ld_nc_v2(src + (int64_t)d * B);
          }
#pragma unroll
          for (int r = 0; r < kBatch; ++r) {
            const int d = d0 + (r0 + r) * 4 * LW;
            const bool ok = phys >= 0 && d < D && nvalid > 0;
            uint4 w = make_uint4(0, 0, 0, 0);
            if (ok) w = keep_first_elems(fp8x8_to_t<T>(p.kv_dtype, raw[r], p.kv_scale), nvalid);
            st_shared_v4(dst + d * 16, w);
          }
        }
      } else if (it < n_ctx) {
        const int kpos = (c_first + it) * BN + vg;
        const int nbytes = 2 * min(8, max(0, ctx_len - kpos));
        const T* src = vc + (int64_t)max(phys, 0) * p.vc_block_stride + (kpos & (B - 1));
        const uint32_t dst = vb + (vg >> 3) * (DP * 16);  // (row d, 8 tokens at vg) -> (vg / 8) * (DP * 16) + d * 16
#pragma unroll 8
        for (int d = (lane >> 3) + 4 * lw; d < DP; d += 4 * LW) {
          const bool ok = phys >= 0 && d < D;
          cp_async16(dst + d * 16, ok ? (const void*)(src + (int64_t)d * B) : (const void*)vc, ok ? nbytes : 0);
        }
      } else {
#pragma unroll
        for (int h = lw; h < 2; h += LW) {
          const int kt = lane + 32 * h;
          const int j = (j_first + it - n_ctx) * BN + kt;
          const T* src = vn + (int64_t)j * p.v_stride_t;
          const uint32_t dst = vb + (kt >> 3) * (DP * 16) + (kt & 7) * 16;  // MN-major core matrices
#pragma unroll
          for (int g = 0; g < DG; ++g) {
            const bool ok = j < q_len && g * 8 < D;
            cp_async16(dst + g * 128, ok ? (const void*)(src + g * 8) : (const void*)vn, ok ? 16 : 0);
          }
        }
      }
    };
    // ---- TMA path: a full prefix tile is BN / 8 boxes {8 tokens, DP rows} of the paged V chunk, one per lane
    // (this warp issues boxes lw, lw + LW, ...); each lands as one 16-byte column block [DP][16 B] of the stage.
    constexpr int VBOX = (BN / 8) / LW;  // boxes per tile and loader warp
    auto v_tma_entry = [&](int it) -> int {
      const int kpos = (c_first + it) * BN + (lw + LW * lane) * 8;
      return lane < VBOX ? table[kpos >> bshift] : 0;
    };
    auto issue_v_tma = [&](int it, int phys) {
      const uint32_t vb = v_smem + (it % G::VSTAGES) * G::VBYTES;
      const uint32_t bar = v_full + 8 * (it % G::VSTAGES);
      if (lane == 0) mbar_expect_tx(bar, (uint32_t)(VBOX * DP * 16));
      __syncwarp();
      if (lane < VBOX) {
        const int tg = lw + LW * lane;
        const int kpos = (c_first + it) * BN + tg * 8;
        tma_load_4d(vb + tg * (DP * 16), &vmap, kpos & (B - 1), 0, kv_head, phys, bar);
      }
    };
    auto stage_v = [&](int it) {  // prologue tiles
      if (it < n_tiles && tma_tile(it)) issue_v_tma(it, v_tma_entry(it));
      else load_v(it, v_entry(it));
    };
    const bool fp8_cache = p.kv_dtype != B200PA_KV_AUTO;
    auto announce_v = [&](int it) {  // see the K loader
      if (it >= n_tiles) return;
      const uint32_t bar = v_full + 8 * (it % G::VSTAGES);
      if (tma_tile(it)) {
        mbar_arrive(bar);
      } else if (fp8_cache && it < n_ctx) {
        fence_proxy_async();
        mbar_arrive(bar);
      } else {
        cp_async_arrive(bar);
      }
    };
    stage_v(0);
    announce_v(0);
    stage_v(1);
    announce_v(1);
    int ev = v_entry(2);
    for (int it = 0; it + 2 < n_tiles; ++it) {
      // V(it+2) goes into the stage PV(it-1) read (its completion implies every loader lane arrived for tile it-1)
      const bool t2 = tma_tile(it + 2);
      const int et = t2 ? v_tma_entry(it + 2) : 0;  // the block-table load flies during the wait below
      if (it >= 1) mbar_wait_relaxed(pv_done + 8 * ((it - 1) % G::VSTAGES), ((it - 1) / G::VSTAGES) & 1);
      if (t2) issue_v_tma(it + 2, et);
      else load_v(it + 2, ev);
      announce_v(it + 2);
      ev = v_entry(it + 3);
    }
    cp_async_commit();
    cp_async_wait<0>();
  } else {
    // ================================ softmax warps: thread = query row ================================
    const int jq = warp >> 2;                 // this warp's head inside the CTA
    const int row = tid & 127;                // query row = TMEM lane
    const int my_head = head + (ROWS ? 0 : jq);
    const int m0_q = m0 + (ROWS ? jq * BM : 0);  // first row of this warp's tile
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tmem_q = tmem_base + jq * 256;   // S0 | S1 | O of this head
    const uint32_t tmem_o = tmem_q + 2 * BN;
    const uint32_t s_full_q = s_full + 16 * jq, p_full_q = p_full + 16 * jq;
    const float slope_l2 = p.alibi_slopes ? p.alibi_slopes[my_head] * kLog2e : 0.f;
    const bool has_alibi = p.alibi_slopes != nullptr;
    const int m = m0_q + row;
    float m_ref = kNegBig, l_run = 0.f;

    for (int it = 0; it < n_tiles; ++it) {
      const bool is_ctx = it < n_ctx;
      const int t0 = is_ctx ? (c_first + it) * BN : (j_first + it - n_ctx) * BN;
      if (warp == 0) B200PA_TRACE(it, 0);
      mbar_wait(s_full_q + 8 * (it & 1), (it >> 1) & 1);
      if (warp == 0) B200PA_TRACE(it, 1);
      tc_fence_after();
      const uint32_t ts = tmem_q + (it & 1) * BN + lane_off;
      uint32_t sr[BN];
      {
        uint32_t (&lo)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[0]);
        uint32_t (&hi)[32] = *reinterpret_cast<uint32_t(*)[32]>(&sr[32]);
        tc_ld32(ts, lo);
        tc_ld32(ts + 32, hi);
        tc_wait_ld();
      }
      const int base_rel = is_ctx ? (t0 - ctx_len) : t0;
      const int col_limit = is_ctx ? (ctx_len - t0) : BN;
      bool need_mask;
      if (is_ctx)
        need_mask = (t0 + BN > ctx_len) || (W > 0 && ctx_len + m0_q + BM - 1 - t0 >= W);
      else
        need_mask = (t0 + BN - 1 > m0_q) || (W > 0 && m0_q + BM - 1 - t0 >= W);
      float mx = kNegBig;
      float mult;  // p = 2^(sr * mult - m_ref)
      if (need_mask || has_alibi) {
#pragma unroll
        for (int c = 0; c < BN; ++c) {
          const int rel = base_rel + c - m;
          float z = __uint_as_float(sr[c]) * p.scale_log2;
          if (has_alibi) z = fmaf(slope_l2, (float)rel, z);
          const bool vis = !need_mask || (c < col_limit && rel <= 0 && (W <= 0 || -rel < W));
          z = vis ? z : kNegBig;
          sr[c] = __float_as_uint(z);
          mx = fmaxf(mx, z);
        }
        mult = 1.f;
      } else {
#pragma unroll
        for (int c = 0; c < BN; ++c) mx = fmaxf(mx, __uint_as_float(sr[c]));
        mx *= p.scale_log2;
        mult = p.scale_log2;
      }
      float alpha = 1.f;
      if (mx > m_ref + 8.f) {  // lazy reference maximum (see the kernel above)
        alpha = ex2(m_ref - mx);
        m_ref = mx;
      }
      if (it > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
        mbar_wait(pv_done + 8 * ((it - 1) % G::VSTAGES), ((it - 1) / G::VSTAGES) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < DP; c0 += 32) {
          uint32_t o[32];
          tc_ld32(tmem_o + lane_off + c0, o);
          tc_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
          tc_st32(tmem_o + lane_off + c0, o);
        }
        tc_wait_st();
      }
      if (warp == 0) B200PA_TRACE(it, 2);
      float sum = 0.f, sum1 = 0.f;
      const float neg_ref = -m_ref;
      uint32_t pw[BN / 2];
#pragma unroll
      for (int c = 0; c < BN / 2; ++c) {
        float a0, a1;
        fma2(a0, a1, __uint_as_float(sr[2 * c]), __uint_as_float(sr[2 * c + 1]), mult, neg_ref);
        const float p0 = ex2(a0), p1 = ex2(a1);
        add2(sum, sum1, p0, p1);
        pw[c] = pack2<T>(p0, p1);
      }
      sum += sum1;
      tc_st32(ts, pw);  // P over its own S columns: the A operand of P.V, read from Tensor Memory
      tc_wait_st();
      l_run = l_run * alpha + sum;
      tc_fence_before();
      mbar_arrive(p_full_q + 8 * (it & 1));
      if (warp == 0) B200PA_TRACE(it, 3);
    }
    mbar_wait(pv_done + 8 * ((n_tiles - 1) % G::VSTAGES), ((n_tiles - 1) / G::VSTAGES) & 1);
    tc_fence_after();
    {
      const float inv = 1.f / l_run;
      T* orow = static_cast<T*>(p.out) + (tok0 + m) * p.o_stride_t + (int64_t)my_head * p.o_stride_h;
#pragma unroll 1
      for (int c0 = 0; c0 < DP; c0 += 32) {
        uint32_t o[32];
        tc_ld32(tmem_o + lane_off + c0, o);
        tc_wait_ld();
        if (m < q_len) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            if (c0 + j < D) {
              uint4 w;
              w.x = pack2<T>(__uint_as_float(o[j]) * inv, __uint_as_float(o[j + 1]) * inv);
              w.y = pack2<T>(__uint_as_float(o[j + 2]) * inv, __uint_as_float(o[j + 3]) * inv);
              w.z = pack2<T>(__uint_as_float(o[j + 4]) * inv, __uint_as_float(o[j + 5]) * inv);
              w.w = pack2<T>(__uint_as_float(o[j + 6]) * inv, __uint_as_float(o[j + 7]) * inv);
              *reinterpret_cast<uint4*>(orow + c0 + j) = w;
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
}

// Tensor map of the paged V cache for the TMA loader: dims (innermost first) {B tokens, D rows, Hkv, blocks}, box
// {8, DP, 1, 1}, no swizzle, zero fill for rows >= D.  The encoder is a driver entry point fetched once through the
// runtime (no link-time dependency on libcuda).  Returns false when the driver has no encoder or rejects the layout
// (strides that are not 16-byte multiples): the kernel then keeps its cp.async loaders.
typedef CUresult (*TensorMapEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                           const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                           CUtensorMapFloatOOBfill);
static TensorMapEncodeTiledFn tensor_map_encoder() {
  static const TensorMapEncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    (void)cudaGetLastError();
    return reinterpret_cast<TensorMapEncodeTiledFn>(f);
  }();
  return fn;
}
static bool make_v_tensor_map(const PrefillParams& kp, int dtype, int dp, CUtensorMap* map) {
  // Developer knob, default OFF: measured on B200 (profiles/r02_prefill_tma_vs_cpasync.log) this staging is SLOWER
  // than the cp.async loader warps (513 vs 833 TFLOP/s on llama8b_chunk2k_ctx4k): 72 TMA instructions per tile, and
  // the {8 tokens x DP rows} V boxes fetch one 32-byte sector per 16 bytes used (ncu: 6.44 GB through TMA for 4.29 GB
  // of K+V).  Parity is green with it on (396 prefill tests).  B200PA_PREFILL_TMA=1 enables it.
  static const bool off = [] { const char* e = std::getenv("B200PA_PREFILL_TMA"); return !(e && std::atoi(e) == 1); }();
  if (off || kp.kv_dtype != B200PA_KV_AUTO) return false;
  TensorMapEncodeTiledFn enc = tensor_map_encoder();
  if (!enc) return false;
  const cuuint64_t B = 1ull << kp.block_shift;
  // one-entry cache: a model calls this once per layer with the same geometry and a different base pointer
  struct Key { const void* ptr; int64_t hs, bs; int d, hkv, bshift, dtype, dp; };
  static thread_local Key last_key{};
  static thread_local CUtensorMap last_map;
  static thread_local bool last_ok = false;
  Key key;
  std::memset(&key, 0, sizeof(key));
  key.ptr = kp.v_cache; key.hs = kp.vc_head_stride; key.bs = kp.vc_block_stride; key.d = kp.head_size;
  key.hkv = kp.num_kv_heads; key.bshift = kp.block_shift; key.dtype = dtype; key.dp = dp;
  if (last_ok && std::memcmp(&key, &last_key, sizeof(Key)) == 0) {
    *map = last_map;
    return true;
  }
  const cuuint64_t dims[4] = {B, (cuuint64_t)kp.head_size, (cuuint64_t)kp.num_kv_heads, 1ull << 24};
  const cuuint64_t strides[3] = {B * 2, (cuuint64_t)kp.vc_head_stride * 2, (cuuint64_t)kp.vc_block_stride * 2};
  const cuuint32_t box[4] = {8, (cuuint32_t)dp, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  if (kp.vc_head_stride <= 0 || kp.vc_block_stride <= 0 || strides[0] % 16 || strides[1] % 16 || strides[2] % 16)
    return false;
  const CUresult rc = enc(map, dtype == B200PA_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                          const_cast<void*>(kp.v_cache), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) return false;
  last_key = key;
  last_map = *map;
  last_ok = true;
  return true;
}

template <typename T, int DP, int NQ, bool ROWS>
static int launch_ws_nq(const PrefillParams& kp_in, int batch, int max_query_len, int dtype, cudaStream_t stream) {
  using G = Geometry<DP>;
  PrefillParams kp = kp_in;
  alignas(64) CUtensorMap vmap;
  std::memset(&vmap, 0, sizeof(vmap));
  kp.use_tma = make_v_tensor_map(kp, dtype, DP, &vmap) ? 1 : 0;
  auto kernel = context_attention_ws_kernel<T, DP, NQ, ROWS>;
  constexpr int kCtaRows = ROWS ? NQ * kBM : kBM;
  constexpr int kSmem = NQ * G::QBYTES + 2 * G::KBYTES + G::VSTAGES * G::VBYTES + 256;
  static bool configured[64] = {false};
  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid(kp.num_heads / (ROWS ? 1 : NQ), (max_query_len + kCtaRows - 1) / kCtaRows, batch);
  kernel<<<grid, NQ * 128 + 32 + 64 * NQ, kSmem, stream>>>(kp, vmap);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

// Two Q tiles per CTA share every staged K/V tile: two heads of a GQA group with an even head count, else two
// consecutive row tiles of one head when the prompts are longer than one tile.  B200PA_PREFILL_NQ (developer knob):
// 1 = one tile per CTA, 3 = row pairs even where head pairs apply.
template <typename T, int DP>
static int launch_ws(const PrefillParams& kp, int batch, int max_query_len, int dtype, cudaStream_t stream) {
  static const int nq_knob = [] { const char* e = std::getenv("B200PA_PREFILL_NQ"); return e ? std::atoi(e) : 0; }();
  const int group = kp.num_heads / kp.num_kv_heads;
  if (nq_knob == 1) return launch_ws_nq<T, DP, 1, false>(kp, batch, max_query_len, dtype, stream);
  if (group % 2 == 0 && nq_knob != 3) return launch_ws_nq<T, DP, 2, false>(kp, batch, max_query_len, dtype, stream);
  if (max_query_len > kBM || nq_knob == 3) return launch_ws_nq<T, DP, 2, true>(kp, batch, max_query_len, dtype, stream);
  return launch_ws_nq<T, DP, 1, false>(kp, batch, max_query_len, dtype, stream);
}

template <typename T, int DP>
static int launch(const PrefillParams& kp, int batch, int max_query_len, cudaStream_t stream) {
  using G = Geometry<DP>;
  auto kernel = context_attention_tc_kernel<T, DP>;
  static bool configured[64] = {false};
  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid(kp.num_heads, (max_query_len + kBM - 1) / kBM, batch);
  kernel<<<grid, kThreads, G::SMEM, stream>>>(kp);
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

}  // namespace prefill_tc

static constexpr int kWsDefault = 1;

int launch_prefill_tc(const PrefillParams& kp, int batch, int max_query_len, int dtype, cudaStream_t stream) {
  auto mis16 = [](const void* ptr) { return reinterpret_cast<uintptr_t>(ptr) % 16 != 0; };
  // 16-byte rows for the cp.async of Q and the 128-bit stores of O
  if (mis16(kp.q) || mis16(kp.out) || kp.q_stride_t % 8 || kp.q_stride_h % 8 || kp.o_stride_t % 8 || kp.o_stride_h % 8) return 1;
  const int d = kp.head_size;
  if (d % 16 != 0) return 1;
  static const int ws_knob = [] { const char* e = std::getenv("B200PA_PREFILL_WS"); return e ? std::atoi(e) : kWsDefault; }();
  if (!ws_knob && kp.kv_dtype != B200PA_KV_AUTO) return 1;  // fp8 staging lives in the warp-specialised kernel's loaders
  if (ws_knob) {
    if (dtype == B200PA_F16) {
      if (d <= 64) return prefill_tc::launch_ws<__half, 64>(kp, batch, max_query_len, dtype, stream);
      if (d <= 128) return prefill_tc::launch_ws<__half, 128>(kp, batch, max_query_len, dtype, stream);
    } else if (dtype == B200PA_BF16) {
      if (d <= 64) return prefill_tc::launch_ws<__nv_bfloat16, 64>(kp, batch, max_query_len, dtype, stream);
      if (d <= 128) return prefill_tc::launch_ws<__nv_bfloat16, 128>(kp, batch, max_query_len, dtype, stream);
    }
    return 1;
  }
  if (dtype == B200PA_F16) {
    if (d <= 64) return prefill_tc::launch<__half, 64>(kp, batch, max_query_len, stream);
    if (d <= 128) return prefill_tc::launch<__half, 128>(kp, batch, max_query_len, stream);
  } else if (dtype == B200PA_BF16) {
    if (d <= 64) return prefill_tc::launch<__nv_bfloat16, 64>(kp, batch, max_query_len, stream);
    if (d <= 128) return prefill_tc::launch<__nv_bfloat16, 128>(kp, batch, max_query_len, stream);
  }
  return 1;
}

}  // namespace b200pa

#ifdef B200PA_PREFILL_TRACE
extern "C" int b200pa_debug_prefill_trace(long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, b200pa::prefill_tc::g_trace, sizeof(long long) * (size_t)n);
}
#endif
