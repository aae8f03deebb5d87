// Instantiations of the TMA + mma.sync paged-attention kernel for ONE head size
// (compile with -DB200PA_INST_D=<head size>; one object per head size so they build in parallel).
#include <algorithm>

#include "attention_mma_kernel.cuh"

#ifndef B200PA_INST_D
#error "compile with -DB200PA_INST_D=<head size>"
#endif

namespace b200pa {

constexpr int kSmemPerSM = 227 * 1024;       // usable shared memory per SM / max per CTA (opt-in)
constexpr int kSmemReservedPerCta = 1024;    // system-reserved shared memory per resident CTA

template <typename T, int KV, int D, int B, int NG, bool BS = false>
static int launch_one(AttnKernelParams kp, int num_seqs, cudaStream_t stream) {
  constexpr int CTB = KV == B200PA_KV_AUTO ? 2 : 1;
  constexpr int STAGE = 2 * B * D * CTB;
  constexpr int HPT = 8 * NG;
  constexpr bool kTwoCtas = (D * NG <= 256);
  auto total_bytes = [&](int nst) { return kMmaWarps * mma_region_bytes(STAGE, nst, D, HPT) + kMmaWarps * nst * 8; };
  int nst = kp.num_stages;
  if (nst <= 0) {
    // Bytes in flight per SM are what hide the HBM latency (~128 KiB suffice, profiles/r02_tune_grid_fused.log);
    // more, smaller CTAs per SM smooth over the start-up and drain bubbles of each CTA.  So: the deepest ring
    // (>= 2 stages) that lets THREE CTAs share an SM (bf16 / D128 / B16: 3 CTAs x 4 warps x 2 stages x 8 KiB =
    // 192 KiB), else the deepest (>= 3) that lets two, else the whole SM for one CTA.  fp8 stages are half the
    // size: 4 of them measured better than the 6-8 that would fit (more only delay the first use of a block).
    constexpr int kMaxStages = KV == B200PA_KV_AUTO ? 8 : 4;
    constexpr bool kThreeCtas = (D * NG <= 128);  // register file: 3 x 128 threads x <= 168 registers
    nst = 0;
    if (kThreeCtas)
      for (int n = kMaxStages; n >= 2 && nst == 0; --n)
        if (total_bytes(n) <= kSmemPerSM / 3 - kSmemReservedPerCta) nst = n;
    if (nst == 0 && kTwoCtas)
      for (int n = kMaxStages; n >= 3 && nst == 0; --n)
        if (total_bytes(n) <= kSmemPerSM / 2 - kSmemReservedPerCta) nst = n;
    if (nst == 0)
      for (int n = 8; n >= 1 && nst == 0; --n)
        if (total_bytes(n) <= kSmemPerSM - kSmemReservedPerCta) nst = n;
    if (nst == 0) return 1;  // does not fit: the SIMT family handles it
  } else {
    nst = std::min(nst, 8);
    while (nst > 1 && total_bytes(nst) > kSmemPerSM - kSmemReservedPerCta) --nst;
    if (total_bytes(nst) > kSmemPerSM - kSmemReservedPerCta) return 1;
  }
  kp.num_stages = nst;
  const int smem = total_bytes(nst);
  auto kernel = paged_attention_mma_kernel<T, KV, D, B, NG, BS>;
  // cudaFuncSetAttribute is per device; remember the largest value configured on each.
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

template <typename T, int KV, int D>
static int launch_bg(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  const int G = kp.num_heads / kp.num_kv_heads;
  if (kp.bs_vert_stride > 1) {
    // block-sparse masks: instantiated for the head sizes the block-sparse models use (phi-3-small: 128) and the
    // reference's test grid (tests/kernels/test_blocksparse_attention.py:29: 64, 112); the general (SIMT) family takes
    // the rest
    if constexpr (D == 64 || D == 112 || D == 128) {
      if (kp.block_size == 16)
        return G <= 8 ? launch_one<T, KV, D, 16, 1, true>(kp, num_seqs, stream)
                      : launch_one<T, KV, D, 16, 2, true>(kp, num_seqs, stream);
      if (kp.block_size == 32)
        return G <= 8 ? launch_one<T, KV, D, 32, 1, true>(kp, num_seqs, stream)
                      : launch_one<T, KV, D, 32, 2, true>(kp, num_seqs, stream);
    }
    return 1;
  }
  if (kp.block_size == 8) {  // 16-bit caches only (an fp8 row of 8 tokens is 8 bytes: the SIMT family keeps those)
    if constexpr (KV == B200PA_KV_AUTO)
      return G <= 8 ? launch_one<T, KV, D, 8, 1>(kp, num_seqs, stream) : launch_one<T, KV, D, 8, 2>(kp, num_seqs, stream);
    return 1;
  }
  if (kp.block_size == 16)
    return G <= 8 ? launch_one<T, KV, D, 16, 1>(kp, num_seqs, stream) : launch_one<T, KV, D, 16, 2>(kp, num_seqs, stream);
  if (kp.block_size == 32)
    return G <= 8 ? launch_one<T, KV, D, 32, 1>(kp, num_seqs, stream) : launch_one<T, KV, D, 32, 2>(kp, num_seqs, stream);
  return 1;
}

template <int D>
static int launch_d(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream) {
  if (dtype == B200PA_F16) {
    switch (kv_dtype) {
      case B200PA_KV_AUTO: return launch_bg<__half, B200PA_KV_AUTO, D>(kp, num_seqs, stream);
      case B200PA_KV_FP8_E4M3: return launch_bg<__half, B200PA_KV_FP8_E4M3, D>(kp, num_seqs, stream);
      case B200PA_KV_FP8_E5M2: return launch_bg<__half, B200PA_KV_FP8_E5M2, D>(kp, num_seqs, stream);
    }
  } else if (dtype == B200PA_BF16) {
    switch (kv_dtype) {
      case B200PA_KV_AUTO: return launch_bg<__nv_bfloat16, B200PA_KV_AUTO, D>(kp, num_seqs, stream);
      case B200PA_KV_FP8_E4M3: return launch_bg<__nv_bfloat16, B200PA_KV_FP8_E4M3, D>(kp, num_seqs, stream);
      case B200PA_KV_FP8_E5M2: return launch_bg<__nv_bfloat16, B200PA_KV_FP8_E5M2, D>(kp, num_seqs, stream);
    }
  }
  return 1;
}

#define B200PA_CAT2(a, b) a##b
#define B200PA_CAT(a, b) B200PA_CAT2(a, b)
int B200PA_CAT(launch_attention_mma_d, B200PA_INST_D)(const AttnKernelParams& kp, int num_seqs, int dtype,
                                                     int kv_dtype, cudaStream_t stream) {
  return launch_d<B200PA_INST_D>(kp, num_seqs, dtype, kv_dtype, stream);
}

}  // namespace b200pa
