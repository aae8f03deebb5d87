// Instantiations of the TMA + mma.sync paged-attention kernel for ONE head size
// (compile with -DB200PA_INST_D=<head size>; one object per head size so they build in parallel).
#include <algorithm>

#include "attention_sk_kernel.cuh"

#ifndef B200PA_INST_D
#error "compile with -DB200PA_INST_D=<head size>"
#endif

namespace b200pa {

extern long long* g_sk_trace;

constexpr int kSmemPerSM = 227 * 1024;       // usable shared memory per SM / max per CTA (opt-in)
constexpr int kSmemReservedPerCta = 1024;    // system-reserved shared memory per resident CTA

template <typename T, int KV, int D, int B, int NG>
static int launch_one(AttnKernelParams kp, int num_seqs, cudaStream_t stream) {
  constexpr int CTB = KV == B200PA_KV_AUTO ? 2 : 1;
  constexpr int STAGE = 2 * B * D * CTB;
  constexpr int HPT = 8 * NG;
  constexpr bool kTwoCtas = (D * NG <= 256);
  auto total_bytes = [&](int nst) { return kMmaWarps * mma_region_bytes(STAGE, nst, D, HPT) + kMmaWarps * nst * 8; };
  int nst = kp.num_stages;
  if (nst <= 0) {
    // Deepest ring (<= 8 stages) that still lets two CTAs share an SM; if that is shallower than
    // 3 stages take the whole SM for one CTA instead.  Bytes in flight per SM are what hide the
    // HBM latency: 2 CTAs x 4 warps x 3 stages x 8 KiB = 192 KiB for bf16 / D128 / B16.
    nst = 0;
    if (kTwoCtas)
      for (int n = 8; n >= 3 && nst == 0; --n)
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
  auto kernel = paged_attention_mma_kernel<T, KV, D, B, NG>;
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

// Persistent launch: one resident wave of CTAs (SMs x CTAs per SM), fewer for tiny problems.
template <typename T, int KV, int D, int B, int NG>
static int launch_sk_one(AttnKernelParams kp, int num_seqs, cudaStream_t stream, int* grid_out) {
  constexpr int CTB = KV == B200PA_KV_AUTO ? 2 : 1;
  constexpr int STAGE = 2 * B * D * CTB;
  constexpr int HPT = 8 * NG;
  constexpr bool kTwoCtas = sk_two_ctas(D, NG);
  const Tuning& tn = tuning();
  // rings + merge buffers / part queue / address queues / flags + one mbarrier per stage
  auto total_bytes = [&](int nst) { return kMmaWarps * nst * STAGE + sk_fixed_smem_bytes(D, NG) + kMmaWarps * nst * 8; };
  const int full = kSmemPerSM - kSmemReservedPerCta, half = kSmemPerSM / 2 - kSmemReservedPerCta;
  int nst = tn.stages;
  if (nst <= 0) {
    // Deepest ring (<= 8 stages) that still lets two CTAs share an SM; if that is shallower than
    // 3 stages take the whole SM for one CTA instead.  Bytes in flight per SM are what hide the
    // HBM latency: 2 CTAs x 4 warps x 3 stages x 8 KiB = 192 KiB for bf16 / D128 / B16.
    nst = 0;
    if (kTwoCtas)
      for (int n = 8; n >= 3 && nst == 0; --n)
        if (total_bytes(n) <= half) nst = n;
    if (nst == 0)
      for (int n = 8; n >= 1 && nst == 0; --n)
        if (total_bytes(n) <= full) nst = n;
    if (nst == 0) return 1;  // does not fit: the SIMT family handles it
  } else {
    nst = std::min(nst, 8);
    while (nst > 1 && total_bytes(nst) > full) --nst;
    if (total_bytes(nst) > full) return 1;
  }
  kp.num_stages = nst;
  if (tn.evict_normal < 0) kp.flags = 0;  // table default: evict_first for the once-read KV stream
  const int smem = total_bytes(nst);
  const int ctas_per_sm = (kTwoCtas && smem <= half) ? 2 : 1;
  auto kernel = paged_attention_sk_kernel<T, KV, D, B, NG>;
  static int configured[64] = {0};
  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || configured[dev] < smem) {
    B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev >= 0 && dev < 64) configured[dev] = smem;
  }
  SkWorkspace ws;
  const int rc = sk_workspace(dev, &ws);
  if (rc != B200PA_OK) return rc;
  kp.sk_slots = ws.slots;
  kp.sk_counters = ws.counters;
  kp.sk_pool = ws.pool;
  kp.sk_max_chunks = (int)std::min<int64_t>(ws.max_counters, ws.slot_bytes / (2 * sk_slot_floats(D, NG) * 4));
  kp.sk_pool_pct = tn.pool_pct;
  kp.sk_chunks_per_cta = tn.chunks_per_cta;
  kp.sk_cs_min = tn.cs_min;
  kp.sk_trace = g_sk_trace;
  const int G = kp.num_heads / kp.num_kv_heads;
  const int q_tiles = (G + HPT - 1) / HPT;
  const int64_t upper = (int64_t)num_seqs * kp.num_kv_heads * q_tiles * std::max(1, (kp.max_seq_len + B - 1) / B);
  int64_t grid = tn.grid_ctas > 0 ? tn.grid_ctas : (int64_t)ws.num_sms * ctas_per_sm;
  grid = std::min<int64_t>(grid, std::max<int64_t>(1, upper / tn.min_blocks_per_cta));
  grid = std::min<int64_t>(grid, kp.sk_max_chunks / 2);
  kernel<<<(unsigned)grid, kSkThreads, smem, stream>>>(kp);
  B200PA_CUDA_CHECK(cudaGetLastError());
  if (grid_out) *grid_out = (int)grid;
  return B200PA_OK;
}

template <typename T, int KV, int D>
static int launch_sk_bg(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream, int* grid_out) {
  const int G = kp.num_heads / kp.num_kv_heads;
  if (kp.block_size == 16)
    return G <= 8 ? launch_sk_one<T, KV, D, 16, 1>(kp, num_seqs, stream, grid_out)
                  : launch_sk_one<T, KV, D, 16, 2>(kp, num_seqs, stream, grid_out);
  if (kp.block_size == 32)
    return G <= 8 ? launch_sk_one<T, KV, D, 32, 1>(kp, num_seqs, stream, grid_out)
                  : launch_sk_one<T, KV, D, 32, 2>(kp, num_seqs, stream, grid_out);
  return 1;
}

template <int D>
static int launch_sk_d(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream,
                       int* grid_out) {
  if (dtype == B200PA_F16) {
    switch (kv_dtype) {
      case B200PA_KV_AUTO: return launch_sk_bg<__half, B200PA_KV_AUTO, D>(kp, num_seqs, stream, grid_out);
      case B200PA_KV_FP8_E4M3: return launch_sk_bg<__half, B200PA_KV_FP8_E4M3, D>(kp, num_seqs, stream, grid_out);
      case B200PA_KV_FP8_E5M2: return launch_sk_bg<__half, B200PA_KV_FP8_E5M2, D>(kp, num_seqs, stream, grid_out);
    }
  } else if (dtype == B200PA_BF16) {
    switch (kv_dtype) {
      case B200PA_KV_AUTO: return launch_sk_bg<__nv_bfloat16, B200PA_KV_AUTO, D>(kp, num_seqs, stream, grid_out);
      case B200PA_KV_FP8_E4M3: return launch_sk_bg<__nv_bfloat16, B200PA_KV_FP8_E4M3, D>(kp, num_seqs, stream, grid_out);
      case B200PA_KV_FP8_E5M2: return launch_sk_bg<__nv_bfloat16, B200PA_KV_FP8_E5M2, D>(kp, num_seqs, stream, grid_out);
    }
  }
  return 1;
}

template <typename T, int KV, int D>
static int launch_bg(const AttnKernelParams& kp, int num_seqs, cudaStream_t stream) {
  const int G = kp.num_heads / kp.num_kv_heads;
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
int B200PA_CAT(launch_attention_sk_d, B200PA_INST_D)(const AttnKernelParams& kp, int num_seqs, int dtype,
                                                    int kv_dtype, cudaStream_t stream, int* grid_out) {
  return launch_sk_d<B200PA_INST_D>(kp, num_seqs, dtype, kv_dtype, stream, grid_out);
}

}  // namespace b200pa
