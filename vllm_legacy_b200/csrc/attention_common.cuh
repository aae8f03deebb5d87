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
  int64_t kv_token_stride;  // elements; flash layout [nb, B, H, D] only
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
  unsigned* sk_counters;    // mma family, num_parts > 1: one self-resetting arrival counter per (sequence, KV-head unit)
};

// Tuning knobs: read from the environment ONCE, when the library is first used (never on the
// launch path), and adjustable afterwards only through b200pa_set_tuning() (developer tools).
struct Tuning {
  int stages = 0;           // B200PA_STAGES: TMA ring depth per warp (0 = table / deepest that fits)
  int evict_normal = -1;    // B200PA_EVICT_NORMAL: -1 = table, 0 = evict_first, 1 = evict_normal
  int force_simt = 0;       // B200PA_FORCE_SIMT: tests only — run the general SIMT family
  int partition_size = 0;   // B200PA_PARTITION_SIZE: split size override (multiple of 512)
  int max_partition = 4096; // B200PA_MAX_PARTITION: split size cap
  int split_cost = 512;     // B200PA_SPLIT_COST: fixed cost of one more CTA in tokens of streaming (split plan)
};
Tuning& tuning();

// Per-device workspace (allocated once per device, on first use): arrival counters of the fused
// split-K reduce and the scratch that paged_attention_v1 splits through (its ABI carries none).
struct AttnWorkspace {
  char* scratch = nullptr;
  size_t scratch_bytes = 0;
  unsigned* counters = nullptr;
  int max_counters = 0;
  int num_sms = 0;
};
constexpr int kMmaMaxSplits = 64;  // most context splits per (sequence, KV head) in the mma family
constexpr size_t kWorkspaceScratchBytes = 64u << 20;
constexpr int kWorkspaceCounters = 1 << 20;
int attn_workspace(int device, AttnWorkspace* ws);  // B200PA_OK or a negative status

// attention_mma.cu — returns B200PA_OK, or 1 when (dtype, kv, block, head) has no MMA
// instantiation (caller falls through to the SIMT family), or a negative status.
int launch_attention_mma(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream);

// Shared memory the mma kernel needs for (head_size, block_size, cache bytes, stages, g-tiles).
size_t mma_smem_bytes(int head_size, int block_size, int cache_elem_bytes, int stages, int ng);

// attention_simt.cu
int launch_attention_simt(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream);
int launch_attention_reduce(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream);

// attention_flash.cu — decode over the flash-attention cache layout [num_blocks, block_size, num_kv_heads, head_size];
// returns B200PA_OK, 1 when (dtype, head size, block size) has no instantiation, or a negative status.
int launch_attention_flash(const AttnKernelParams& kp, int num_seqs, int dtype, cudaStream_t stream);
bool flash_head_size_supported(int d);

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
