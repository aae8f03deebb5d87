// Kernel-side parameter block shared by the two prefix-prefill kernel families
// (prefill.cu: mma.sync; prefill_tc.cu: tcgen05 + TMEM).
#pragma once

#include "attention_common.cuh"

namespace b200pa {

struct PrefillParams {
  void* out;
  const void* q;
  const void* k;
  const void* v;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  const int* start_loc;
  const int* seq_lens;
  const int* ctx_lens;
  const float* alibi_slopes;
  int64_t q_stride_t, q_stride_h, k_stride_t, k_stride_h, v_stride_t, v_stride_h, o_stride_t, o_stride_h;
  int64_t kc_block_stride, kc_head_stride, vc_block_stride, vc_head_stride;  // elements
  int64_t block_table_stride;
  int num_heads, num_kv_heads, head_size, block_shift, sliding_window;
  float scale_log2;  // scale * log2(e)
};

// prefill_tc.cu — returns B200PA_OK, 1 when the shape has no tcgen05 instantiation, or a negative status.
int launch_prefill_tc(const PrefillParams& kp, int batch, int max_query_len, int dtype, cudaStream_t stream);

}  // namespace b200pa
