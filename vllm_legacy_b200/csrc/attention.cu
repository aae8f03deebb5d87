// C-ABI entry points b200pa_paged_attention_v1 / _v2 (include/b200pa.h): argument checks with
// the reference's error classes, launch planning (split count sized for 148 SMs), dispatch to the
// TMA + mma.sync family or the general SIMT family, and the split-K reduce.
//
// Reference host side: csrc/attention/attention_kernels.cu:690-825 (v1), :844-1014 (v2).
#include <algorithm>
#include <cstdlib>

#include "attention_common.cuh"

namespace b200pa {

static int env_int(const char* name, int dflt) {
  const char* s = std::getenv(name);
  return (s && *s) ? std::atoi(s) : dflt;
}

static bool head_size_supported(int d) {  // attention_kernels.cu:738-766
  return d == 64 || d == 80 || d == 96 || d == 112 || d == 128 || d == 192 || d == 256;
}
static bool block_size_supported(int b) {  // attention_kernels.cu:789-803
  return b == 8 || b == 16 || b == 32;
}

constexpr int kRefPartition = 512;  // granularity the callers size the v2 scratch with
constexpr int kNumSMs = 148;

static int run_attention(const b200pa_attn_params_t* a, bool allow_split, cudaStream_t stream) {
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->kv_dtype != B200PA_KV_AUTO && a->kv_dtype != B200PA_KV_FP8_E4M3 && a->kv_dtype != B200PA_KV_FP8_E5M2)
    return B200PA_ERR_KV_DTYPE;
  if (a->dtype != B200PA_F16 && a->dtype != B200PA_BF16 && a->dtype != B200PA_F32) return B200PA_ERR_INPUT_DTYPE;
  if (!block_size_supported(a->block_size)) return B200PA_ERR_BLOCK_SIZE;
  if (!head_size_supported(a->head_size)) return B200PA_ERR_HEAD_SIZE;
  if (a->num_seqs < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0 ||
      a->max_seq_len < 0 || a->max_num_blocks_per_seq < 0)
    return B200PA_ERR_INVALID_ARG;
  if (a->num_seqs == 0) return B200PA_OK;
  if (!a->out || !a->query || !a->key_cache || !a->value_cache || !a->block_tables || !a->seq_lens)
    return B200PA_ERR_INVALID_ARG;
  const int cache_bytes = a->kv_dtype == B200PA_KV_AUTO ? (a->dtype == B200PA_F32 ? 4 : 2) : 1;
  // 16-byte granularity of the paged layout (x = 16 / sizeof(cache_t)): the reference's vector
  // loads assume it too (attention_kernels.cu:200, 272-290).
  if (reinterpret_cast<uintptr_t>(a->key_cache) % 16 || reinterpret_cast<uintptr_t>(a->value_cache) % 16 ||
      (a->kv_block_stride * cache_bytes) % 16 || (a->kv_head_stride * cache_bytes) % 16)
    return B200PA_ERR_INVALID_ARG;

  AttnKernelParams kp{};
  kp.out = a->out;
  kp.exp_sums = a->exp_sums;
  kp.max_logits = a->max_logits;
  kp.tmp_out = a->tmp_out;
  kp.q = a->query;
  kp.k_cache = a->key_cache;
  kp.v_cache = a->value_cache;
  kp.block_tables = a->block_tables;
  kp.seq_lens = a->seq_lens;
  kp.alibi_slopes = a->alibi_slopes;
  kp.q_stride = a->q_stride;
  kp.kv_block_stride = a->kv_block_stride;
  kp.kv_head_stride = a->kv_head_stride;
  kp.num_heads = a->num_heads;
  kp.num_kv_heads = a->num_kv_heads;
  kp.head_size = a->head_size;
  kp.block_size = a->block_size;
  kp.max_blocks_per_seq = a->max_num_blocks_per_seq;
  kp.scale = a->scale;
  kp.kv_scale = a->kv_scale;
  kp.tp_rank = a->tp_rank;
  kp.bs_local_blocks = a->blocksparse_local_blocks;
  kp.bs_vert_stride = a->blocksparse_vert_stride;
  kp.bs_block_size = a->blocksparse_block_size;
  kp.bs_head_sliding_step = a->blocksparse_head_sliding_step;
  kp.num_stages = env_int("B200PA_STAGES", 0);
  kp.flags = env_int("B200PA_EVICT_NORMAL", 0) ? 1 : 0;

  const bool blocksparse = a->blocksparse_vert_stride > 1;
  if (blocksparse && (a->blocksparse_block_size <= 0 || a->blocksparse_block_size % a->block_size != 0))
    return B200PA_ERR_INVALID_ARG;
  const bool mma_ok = !blocksparse && a->dtype != B200PA_F32 && a->block_size != 8 &&
                      env_int("B200PA_FORCE_SIMT", 0) == 0;

  // ---- split plan -----------------------------------------------------------------------
  // One CTA streams the KV of one (sequence, KV head[, split]) for the whole GQA group (mma
  // family) or of one (sequence, query head[, split]) (SIMT family).  Split the context when
  // those units alone cannot give every SM several CTAs; never below the reference's 512-token
  // partition so the caller-sized scratch (ceil(max_seq_len/512) slots) always fits.
  kp.num_parts = 1;
  kp.partition_size = std::max(a->max_seq_len, 1);
  kp.scratch_parts = std::max(a->max_num_partitions, 1);
  if (allow_split) {
    if (!a->exp_sums || !a->max_logits || !a->tmp_out || a->max_num_partitions <= 0) return B200PA_ERR_INVALID_ARG;
    const int64_t units = (int64_t)a->num_seqs * (mma_ok ? a->num_kv_heads : a->num_heads);
    const int64_t target = (int64_t)kNumSMs * 8;
    int want = (int)std::min<int64_t>((target + units - 1) / units, a->max_num_partitions);
    int psize = (a->max_seq_len + want - 1) / std::max(want, 1);
    psize = std::min(psize, env_int("B200PA_MAX_PARTITION", 4096));
    psize = (psize + kRefPartition - 1) / kRefPartition * kRefPartition;
    const int forced = env_int("B200PA_PARTITION_SIZE", 0);
    if (forced > 0) psize = (forced + kRefPartition - 1) / kRefPartition * kRefPartition;
    const int parts = (a->max_seq_len + psize - 1) / psize;
    if (parts > 1) {
      kp.num_parts = parts;  // parts <= max_num_partitions because psize >= 512
      kp.partition_size = psize;
    }
  }

  int rc = 1;
  if (mma_ok) {
    rc = launch_attention_mma(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
    if (rc == B200PA_OK) g_last_attention_path = "mma_tma";
  }
  if (rc == 1) {
    rc = launch_attention_simt(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
    if (rc == B200PA_OK) g_last_attention_path = "simt";
  }
  if (rc != B200PA_OK) return rc;
  g_last_attention_partitions = kp.num_parts;
  if (kp.num_parts > 1) return launch_attention_reduce(kp, a->num_seqs, a->dtype, stream);
  return B200PA_OK;
}

}  // namespace b200pa

extern "C" int b200pa_paged_attention_v1(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*allow_split=*/false, stream);
}

extern "C" int b200pa_paged_attention_v2(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*allow_split=*/true, stream);
}
