// Kernel-side parameter block and launch plan shared by the two paged-attention kernel
// families (attention_mma.cu: TMA-bulk staged + mma.sync GQA kernel; attention_simt.cu: general
// SIMT kernel + the split-K reduce).
#pragma once

#include "common.cuh"

namespace b200pa {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kNegBig = -1.0e30f;  // finite "minus infinity" (survives * log2e, differences stay finite)

struct AttnKernelParams {
  void* out;           // [S, Hq, D] T
  float* exp_sums;     // [S, Hq, scratch_parts]     (split only)
  float* max_logits;   // [S, Hq, scratch_parts]     (split only)
  void* tmp_out;       // [S, Hq, scratch_parts, D]  (split only)
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  const int* seq_lens;
  const float* alibi_slopes;
  int64_t q_stride;
  int64_t kv_block_stride;  // elements
  int64_t kv_head_stride;   // elements
  int num_heads;
  int num_kv_heads;
  int head_size;
  int block_size;
  int max_blocks_per_seq;
  int partition_size;  // tokens per split (multiple of block_size); ignored when num_parts == 1
  int num_parts;       // gridDim.z; 1 => write `out` directly
  int scratch_parts;   // partition stride of the scratch buffers (caller's max_num_partitions)
  int num_stages;      // mma kernel: TMA ring depth per warp
  int flags;           // bit 0: L2 evict_normal instead of evict_first for the KV stream (tuning knob)
  float scale;
  float kv_scale;
  int tp_rank;
  int bs_local_blocks;
  int bs_vert_stride;
  int bs_block_size;
  int bs_head_sliding_step;
  // persistent (stream-K) mma kernel only
  int num_seqs;
  int max_seq_len;          // caller's bound; sizes the grid, never trusted for correctness
  float* sk_slots;          // per-device workspace: 2 partial-result slots per CTA
  unsigned* sk_counters;    // per-device workspace: one self-resetting arrival counter per CTA
  unsigned* sk_pool;        // per-device workspace: [0] next pool chunk, [1] CTAs finished (both self-resetting)
  int sk_max_chunks;        // workspace capacity in chunks (2 slots + 1 counter each)
  int sk_pool_pct;          // share of the blocks handed out dynamically, percent
  int sk_chunks_per_cta;    // pool chunks aimed at per CTA
  int sk_cs_min;            // smallest pool chunk, blocks
  long long* sk_trace;      // developer trace (b200pa_set_trace): [cta][4] = {smid, t_begin, t_end, blocks}, or null
};

// Tuning knobs: read from the environment ONCE, when the library is first used (never on the
// launch path), and adjustable afterwards only through b200pa_set_tuning() (developer tools).
struct Tuning {
  int stages = 0;           // B200PA_STAGES: TMA ring depth per warp (0 = table / deepest that fits)
  int evict_normal = -1;    // B200PA_EVICT_NORMAL: -1 = table, 0 = evict_first, 1 = evict_normal
  int force_simt = 0;       // B200PA_FORCE_SIMT: tests only — run the general SIMT family
  int partition_size = 0;   // B200PA_PARTITION_SIZE: SIMT split size override (multiple of 512)
  int max_partition = 4096; // B200PA_MAX_PARTITION: SIMT split size cap
  int grid_ctas = 0;        // B200PA_GRID_CTAS: persistent kernel grid override (0 = SMs x occupancy)
  int min_blocks_per_cta = 8;  // B200PA_MIN_BLOCKS_PER_CTA: small problems use fewer CTAs
  int pool_pct = 25;        // B200PA_POOL_PCT: share of the KV blocks scheduled dynamically (0 = static split)
  int chunks_per_cta = 6;   // B200PA_CHUNKS_PER_CTA: pool chunks per CTA aimed at (smaller chunks, smaller tail)
  int cs_min = 24;          // B200PA_CS_MIN: smallest pool chunk in KV blocks
  int grid_kernel = 0;      // B200PA_GRID_KERNEL: A/B only — the round-1 one-CTA-per-(seq, KV head, split) kernel
};
Tuning& tuning();

// Per-device workspace of the persistent kernel (allocated once per device, on first use).
struct SkWorkspace {
  float* slots = nullptr;
  unsigned* counters = nullptr;
  unsigned* pool = nullptr;  // [2], right behind the counters
  size_t slot_bytes = 0;   // bytes behind `slots`: 2 partials per CTA fit for grid <= slot_bytes / (2 * slot size)
  int max_counters = 0;
  int num_sms = 0;
};
constexpr size_t kSkWorkspaceSlotBytes = 48u << 20;
constexpr int kSkMaxCounters = 16384;
int sk_workspace(int device, SkWorkspace* ws);  // B200PA_OK or a negative status

// attention_mma.cu — returns B200PA_OK, or 1 when (dtype, kv, block, head) has no MMA
// instantiation (caller falls through to the SIMT family), or a negative status.
int launch_attention_mma(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream);
// persistent stream-K variant (attention_sk_kernel.cuh); *grid_out = CTAs launched
int launch_attention_sk(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream,
                        int* grid_out);
// Shared memory the mma kernel needs for (head_size, block_size, cache bytes, stages, g-tiles).
size_t mma_smem_bytes(int head_size, int block_size, int cache_elem_bytes, int stages, int ng);

// attention_simt.cu
int launch_attention_simt(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream);
int launch_attention_reduce(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream);

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-sparse predicate of the reference (attention_kernels.cu:212-224, 236-241):
// a KV block is attended when it is "remote" (on the vertical stride) or "local".
struct BlockSparse {
  int offset;      // bs_block_offset
  int q_block_id;  // q_bs_block_id
  int local_blocks, vert_stride, bs_block_size;
  __device__ __forceinline__ void init(const AttnKernelParams& p, int seq_len, int head_idx, int kv_head_idx) {
    q_block_id = (seq_len - 1) / p.bs_block_size;
    if (p.bs_head_sliding_step >= 0)
      offset = (p.tp_rank * p.num_heads + head_idx) * p.bs_head_sliding_step + 1;
    else
      offset = (p.tp_rank * p.num_kv_heads + kv_head_idx) * (-p.bs_head_sliding_step) + 1;
    local_blocks = p.bs_local_blocks;
    vert_stride = p.bs_vert_stride;
    bs_block_size = p.bs_block_size;
  }
  __device__ __forceinline__ bool attended(int block_idx, int block_size) const {
    const int k_bs_block_id = block_idx * block_size / bs_block_size;
    const bool is_remote = ((k_bs_block_id + offset) % vert_stride == 0);
    const bool is_local = (k_bs_block_id > q_block_id - local_blocks);
    return is_remote || is_local;
  }
};

}  // namespace b200pa
