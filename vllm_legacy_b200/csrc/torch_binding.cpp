// Thin PyTorch binding over the C-ABI (include/b200pa.h): a drop-in for the hot-path subset of the
// reference's `vllm._C` pybind module — submodules `ops`, `cache_ops`, `cuda_utils` with the
// fork's exact positional signatures (csrc/pybind.cpp:8-14, 74-95; csrc/ops.h:5-30;
// csrc/cache.h:8-42), including its pointer-string arguments for paged_attention_v2 and
// reshape_and_cache.  Everything here is marshalling: tensors -> raw pointers/strides -> b200pa_*.
// Error texts follow the reference's TORCH_CHECKs.  No CPU fallback: non-CUDA tensors are rejected.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <cmath>
#include <string>
#include <vector>

#include "b200pa.h"

namespace {

b200pa_stream_t current_stream() {
  return reinterpret_cast<b200pa_stream_t>(at::cuda::getCurrentCUDAStream().stream());
}

int scalar_dtype(const torch::Tensor& t) {
  switch (t.scalar_type()) {
    case at::ScalarType::Half: return B200PA_F16;
    case at::ScalarType::BFloat16: return B200PA_BF16;
    case at::ScalarType::Float: return B200PA_F32;
    case at::ScalarType::Byte: return B200PA_U8;
    default: return -1;
  }
}

// quant_utils.cuh:529-566 (DISPATCH_BY_KV_CACHE_DTYPE)
int parse_kv_dtype(const std::string& s) {
  if (s == "auto") return B200PA_KV_AUTO;
  if (s == "fp8" || s == "fp8_e4m3") return B200PA_KV_FP8_E4M3;
  if (s == "fp8_e5m2") return B200PA_KV_FP8_E5M2;
  TORCH_CHECK(false, "Unsupported data type of kv cache: ", s);
}

void* parse_ptr(const std::string& s) { return reinterpret_cast<void*>(std::stoull(s)); }

void check_status(int rc, const b200pa_attn_params_t* a, const torch::Tensor& src, const std::string& kv) {
  if (rc == B200PA_OK) return;
  switch (rc) {
    case B200PA_ERR_HEAD_SIZE: TORCH_CHECK(false, "Unsupported head size: ", a ? a->head_size : -1);
    case B200PA_ERR_BLOCK_SIZE: TORCH_CHECK(false, "Unsupported block size: ", a ? a->block_size : -1);
    case B200PA_ERR_KV_DTYPE: TORCH_CHECK(false, "Unsupported data type of kv cache: ", kv);
    case B200PA_ERR_INPUT_DTYPE: TORCH_CHECK(false, "Unsupported input type of kv cache: ", src.scalar_type());
    case B200PA_ERR_DEVICE_COMBINATION: TORCH_CHECK(false, "Invalid device combination");
    case B200PA_ERR_CUDA:
      TORCH_CHECK(false, "b200pa: CUDA error ", b200pa_last_cuda_error(), " (",
                  cudaGetErrorString(static_cast<cudaError_t>(b200pa_last_cuda_error())), ")");
    default: TORCH_CHECK(false, "b200pa: ", b200pa_status_string(rc));
  }
}

void fill_attention(b200pa_attn_params_t& a, torch::Tensor& out, torch::Tensor& query, int num_kv_heads, float scale,
                    torch::Tensor& block_tables, torch::Tensor& seq_lens, int block_size, int max_seq_len,
                    const c10::optional<torch::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                    float kv_scale, int tp_rank, int bs_local_blocks, int bs_vert_stride, int bs_block_size,
                    int bs_head_sliding_step) {
  TORCH_CHECK(query.is_cuda() && out.is_cuda() && block_tables.is_cuda() && seq_lens.is_cuda(),
              "b200pa paged_attention: tensors must be on a CUDA device (there is no CPU path)");
  TORCH_CHECK(query.dim() == 3 && query.stride(2) == 1 && query.stride(1) == query.size(2),
              "query must be [num_seqs, num_heads, head_size] with contiguous heads");
  TORCH_CHECK(out.is_contiguous(), "out must be contiguous");
  TORCH_CHECK(out.scalar_type() == query.scalar_type() && out.numel() == query.size(0) * query.size(1) * query.size(2),
              "out must have the query's dtype and [num_seqs, num_heads, head_size] elements");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && seq_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and seq_lens must be int32");
  TORCH_CHECK(block_tables.dim() == 2 && block_tables.stride(1) == 1 && block_tables.stride(0) == block_tables.size(1),
              "block_tables must be contiguous [num_seqs, max_num_blocks_per_seq]");
  a.out = out.data_ptr();
  a.query = query.data_ptr();
  a.block_tables = block_tables.data_ptr<int>();
  a.seq_lens = seq_lens.data_ptr<int>();
  a.alibi_slopes = nullptr;
  if (alibi_slopes.has_value()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_cuda(),
                "alibi_slopes must be a float32 CUDA tensor");
    a.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  a.q_stride = query.stride(0);
  a.num_seqs = query.size(0);
  a.num_heads = query.size(1);
  a.head_size = query.size(2);
  a.num_kv_heads = num_kv_heads;
  a.block_size = block_size;
  a.max_seq_len = max_seq_len;
  a.max_num_blocks_per_seq = block_tables.size(1);
  a.scale = scale;
  a.kv_scale = kv_scale;
  a.kv_dtype = parse_kv_dtype(kv_cache_dtype);
  a.dtype = scalar_dtype(query);
  if (a.dtype < 0 || a.dtype == B200PA_U8)
    TORCH_CHECK(false, "Unsupported input type of kv cache: ", query.scalar_type());
  a.tp_rank = tp_rank;
  a.blocksparse_local_blocks = bs_local_blocks;
  a.blocksparse_vert_stride = bs_vert_stride;
  a.blocksparse_block_size = bs_block_size;
  a.blocksparse_head_sliding_step = bs_head_sliding_step;
}

// csrc/ops.h:5-12
void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int num_kv_heads, float scale, torch::Tensor& block_tables,
                        torch::Tensor& seq_lens, int block_size, int max_seq_len,
                        const c10::optional<torch::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                        float kv_scale, const int tp_rank, const int blocksparse_local_blocks,
                        const int blocksparse_vert_stride, const int blocksparse_block_size,
                        const int blocksparse_head_sliding_step) {
  b200pa_attn_params_t a{};
  fill_attention(a, out, query, num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                 kv_cache_dtype, kv_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                 blocksparse_block_size, blocksparse_head_sliding_step);
  TORCH_CHECK(key_cache.is_cuda() && value_cache.is_cuda(), "kv caches must be CUDA tensors");
  // a cache of the wrong element size would be reinterpreted silently (and bulk-copied with the wrong chunk size)
  const int64_t cache_elt = kv_cache_dtype == "auto" ? query.element_size() : 1;
  TORCH_CHECK(key_cache.element_size() == cache_elt && value_cache.element_size() == cache_elt,
              "kv cache element size ", key_cache.element_size(), " does not match kv_cache_dtype '", kv_cache_dtype, "'");
  TORCH_CHECK(key_cache.dim() >= 2 && value_cache.dim() >= 2 && key_cache.stride(0) == value_cache.stride(0) &&
                  key_cache.stride(1) == value_cache.stride(1),
              "key_cache and value_cache must share their block and head strides");
  a.key_cache = key_cache.data_ptr();
  a.value_cache = value_cache.data_ptr();
  a.kv_block_stride = key_cache.stride(0);
  a.kv_head_stride = key_cache.stride(1);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_paged_attention_v1(&a, current_stream()), &a, query, kv_cache_dtype);
}

// decode over the flash-attention cache layout [num_blocks, block_size, num_kv_heads, head_size]
// (what flash_attn_with_kvcache does for the FLASH_ATTN backend, vllm/attention/backends/flash_attn.py:354-367)
void paged_attention_flash(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                           torch::Tensor& value_cache, float scale, torch::Tensor& block_tables,
                           torch::Tensor& seq_lens, int max_seq_len,
                           const c10::optional<torch::Tensor>& alibi_slopes) {
  TORCH_CHECK(key_cache.is_cuda() && value_cache.is_cuda(), "kv caches must be CUDA tensors");
  TORCH_CHECK(key_cache.dim() == 4 && value_cache.dim() == 4 && key_cache.stride(3) == 1 && value_cache.stride(3) == 1,
              "flash layout: key_cache / value_cache [num_blocks, block_size, num_kv_heads, head_size]");
  TORCH_CHECK(key_cache.scalar_type() == query.scalar_type() && value_cache.scalar_type() == query.scalar_type(),
              "flash layout: the cache has the query dtype");
  TORCH_CHECK(key_cache.strides() == value_cache.strides(), "key_cache and value_cache must share their strides");
  b200pa_attn_params_t a{};
  fill_attention(a, out, query, key_cache.size(2), scale, block_tables, seq_lens, key_cache.size(1), max_seq_len,
                 alibi_slopes, "auto", 1.0f, 0, 0, 0, 64, 0);
  a.key_cache = key_cache.data_ptr();
  a.value_cache = value_cache.data_ptr();
  a.kv_block_stride = key_cache.stride(0);
  a.kv_head_stride = key_cache.stride(2);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_paged_attention_flash(&a, key_cache.stride(1), current_stream()), &a, query, "auto");
}

// csrc/ops.h:14-30 (fork ABI: cache pointers as decimal strings + key strides)
void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
                        torch::Tensor& tmp_out, torch::Tensor& query, const std::string& key_cache_data_ptr_str,
                        int key_cache_stride_0, int key_cache_stride_1,
                        const std::string& value_cache_data_ptr_str, int num_kv_heads, float scale,
                        torch::Tensor& block_tables, torch::Tensor& seq_lens, int block_size, int max_seq_len,
                        const c10::optional<torch::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                        float kv_scale, const int tp_rank, const int blocksparse_local_blocks,
                        const int blocksparse_vert_stride, const int blocksparse_block_size,
                        const int blocksparse_head_sliding_step) {
  b200pa_attn_params_t a{};
  fill_attention(a, out, query, num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                 kv_cache_dtype, kv_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
                 blocksparse_block_size, blocksparse_head_sliding_step);
  TORCH_CHECK(exp_sums.is_cuda() && max_logits.is_cuda() && tmp_out.is_cuda(), "v2 scratch must be CUDA tensors");
  TORCH_CHECK(exp_sums.scalar_type() == at::ScalarType::Float && max_logits.scalar_type() == at::ScalarType::Float,
              "exp_sums and max_logits must be float32");
  a.key_cache = parse_ptr(key_cache_data_ptr_str);
  a.value_cache = parse_ptr(value_cache_data_ptr_str);
  a.kv_block_stride = key_cache_stride_0;
  a.kv_head_stride = key_cache_stride_1;
  a.exp_sums = exp_sums.data_ptr<float>();
  a.max_logits = max_logits.data_ptr<float>();
  a.tmp_out = tmp_out.data_ptr();
  a.max_num_partitions = (max_seq_len + 511) / 512;  // attention_kernels.cu:869 (PARTITION_SIZE = 512)
  TORCH_CHECK(exp_sums.numel() >= (int64_t)a.num_seqs * a.num_heads * a.max_num_partitions &&
                  max_logits.numel() >= exp_sums.numel() - 0 &&
                  tmp_out.numel() >= (int64_t)a.num_seqs * a.num_heads * a.max_num_partitions * a.head_size,
              "v2 scratch buffers are smaller than [num_seqs, num_heads, ceil(max_seq_len/512)(, head_size)]");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_paged_attention_v2(&a, current_stream()), &a, query, kv_cache_dtype);
}

// csrc/cache.h:15-25 (fork ABI)
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, const std::string& key_cache_data_ptr_str,
                       int key_cache_block_size, int key_cache_x, const std::string& value_cache_data_ptr_str,
                       torch::Tensor& slot_mapping, const std::string& kv_cache_dtype, const float kv_scale) {
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && slot_mapping.is_cuda(),
              "b200pa reshape_and_cache: tensors must be on a CUDA device (there is no CPU path)");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be contiguous int64");
  TORCH_CHECK(key.dim() == 3 && key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value must be [num_tokens, num_heads, head_size] with contiguous heads");
  b200pa_reshape_params_t p{};
  p.key = key.data_ptr();
  p.value = value.data_ptr();
  p.key_cache = parse_ptr(key_cache_data_ptr_str);
  p.value_cache = parse_ptr(value_cache_data_ptr_str);
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.key_stride = key.stride(0);
  p.value_stride = value.stride(0);
  p.num_tokens = key.size(0);
  p.num_heads = key.size(1);
  p.head_size = key.size(2);
  p.block_size = key_cache_block_size;
  p.x = key_cache_x;
  p.kv_dtype = parse_kv_dtype(kv_cache_dtype);
  p.kv_scale = kv_scale;
  p.dtype = scalar_dtype(key);
  if (p.dtype < 0 || p.dtype == B200PA_U8) TORCH_CHECK(false, "Unsupported input type of kv cache: ", key.scalar_type());
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  check_status(b200pa_reshape_and_cache(&p, current_stream()), nullptr, key, kv_cache_dtype);
}

// csrc/cache.h:27-39 (fork ABI: cache pointers as decimal strings + block size / block strides)
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, const std::string& key_cache_data_ptr_str,
                             int key_cache_size_1, int key_cache_stride_0, const std::string& value_cache_data_ptr_str,
                             int value_cache_stride_0, torch::Tensor& slot_mapping, const std::string& kv_cache_dtype) {
  // cache_kernels.cu:303-306: only the "auto" KV dtype
  if (kv_cache_dtype != "auto") TORCH_CHECK(false, "Unsupported data type of kv cache: ", kv_cache_dtype);
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && slot_mapping.is_cuda(),
              "b200pa reshape_and_cache_flash: tensors must be on a CUDA device (there is no CPU path)");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be contiguous int64");
  TORCH_CHECK(key.dim() == 3 && key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              "key/value must be [num_tokens, num_heads, head_size] with contiguous heads");
  TORCH_CHECK(key_cache_stride_0 == value_cache_stride_0, "k_cache and v_cache must have the same block stride");
  b200pa_reshape_flash_params_t p{};
  p.key = key.data_ptr();
  p.value = value.data_ptr();
  p.key_cache = parse_ptr(key_cache_data_ptr_str);
  p.value_cache = parse_ptr(value_cache_data_ptr_str);
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.key_stride = key.stride(0);
  p.value_stride = value.stride(0);
  p.block_stride = key_cache_stride_0;
  p.num_tokens = key.size(0);
  p.num_heads = key.size(1);
  p.head_size = key.size(2);
  p.block_size = key_cache_size_1;
  p.dtype = scalar_dtype(key);
  if (p.dtype < 0 || p.dtype == B200PA_U8) TORCH_CHECK(false, "Unsupported input type of kv cache: ", key.scalar_type());
  const at::cuda::OptionalCUDAGuard device_guard(device_of(key));
  check_status(b200pa_reshape_and_cache_flash(&p, current_stream()), nullptr, key, kv_cache_dtype);
}

// csrc/ops.h:38-47 (rotary_embedding / batched_rotary_embedding) and the fused
// rotary + reshape_and_cache decode-step op (not in the reference; bit-identical to the two calls).
static void rope_common(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                        torch::Tensor& cos_sin_cache, bool is_neox, int rot_dim, const int64_t* offsets,
                        int64_t num_tokens, b200pa_rope_params_t& p) {
  TORCH_CHECK(query.is_cuda() && key.is_cuda() && positions.is_cuda() && cos_sin_cache.is_cuda(),
              "b200pa rotary_embedding: tensors must be on a CUDA device (there is no CPU path)");
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long && positions.is_contiguous(),
              "positions must be contiguous int64");
  TORCH_CHECK(cos_sin_cache.is_contiguous() && cos_sin_cache.scalar_type() == query.scalar_type() &&
                  key.scalar_type() == query.scalar_type(),
              "query, key and cos_sin_cache must share one dtype");
  TORCH_CHECK(query.stride(-1) == 1 && key.stride(-1) == 1, "query/key must have a contiguous last dimension");
  p.positions = positions.data_ptr<int64_t>();
  p.query = query.data_ptr();
  p.key = key.data_ptr();
  p.cos_sin_cache = cos_sin_cache.data_ptr();
  p.cos_sin_cache_offsets = offsets;
  p.query_stride = query.stride(-2);
  p.key_stride = key.stride(-2);
  p.num_tokens = num_tokens;
  p.num_heads = query.size(-1) / head_size;
  p.num_kv_heads = key.size(-1) / head_size;
  p.head_size = head_size;
  p.rot_dim = rot_dim;
  p.is_neox = is_neox ? 1 : 0;
  p.dtype = scalar_dtype(query);
  if (p.dtype < 0 || p.dtype == B200PA_U8) TORCH_CHECK(false, "rotary_embedding: unsupported dtype ", query.scalar_type());
}

void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox) {
  b200pa_rope_params_t p{};
  rope_common(positions, query, key, head_size, cos_sin_cache, is_neox, cos_sin_cache.size(1), nullptr,
              query.numel() / query.size(-1), p);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_rotary_embedding(&p, current_stream()), nullptr, query, "auto");
}

void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox, int rot_dim,
                              torch::Tensor& cos_sin_cache_offsets) {
  TORCH_CHECK(cos_sin_cache_offsets.is_cuda() && cos_sin_cache_offsets.scalar_type() == at::ScalarType::Long &&
                  cos_sin_cache_offsets.is_contiguous(),
              "cos_sin_cache_offsets must be a contiguous CUDA int64 tensor");
  b200pa_rope_params_t p{};
  rope_common(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
              cos_sin_cache_offsets.data_ptr<int64_t>(), cos_sin_cache_offsets.size(0), p);
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_rotary_embedding(&p, current_stream()), nullptr, query, "auto");
}

void rotary_embedding_reshape_and_cache(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key,
                                        torch::Tensor& value, int head_size, torch::Tensor& cos_sin_cache,
                                        bool is_neox, const std::string& key_cache_data_ptr_str,
                                        int key_cache_block_size, int key_cache_x,
                                        const std::string& value_cache_data_ptr_str, torch::Tensor& slot_mapping,
                                        const std::string& kv_cache_dtype, float kv_scale) {
  const int kv = parse_kv_dtype(kv_cache_dtype);
  b200pa_rope_params_t p{};
  rope_common(positions, query, key, head_size, cos_sin_cache, is_neox, cos_sin_cache.size(1), nullptr,
              query.numel() / query.size(-1), p);
  TORCH_CHECK(value.is_cuda() && slot_mapping.is_cuda() && value.scalar_type() == query.scalar_type() &&
                  value.stride(-1) == 1,
              "value must be a CUDA tensor of the query dtype with a contiguous last dimension");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous() &&
                  slot_mapping.numel() == p.num_tokens,
              "slot_mapping must be contiguous int64 [num_tokens]");
  p.value = value.data_ptr();
  p.value_stride = value.dim() == 3 ? value.stride(0) : value.stride(-2);
  p.key_cache = parse_ptr(key_cache_data_ptr_str);
  p.value_cache = parse_ptr(value_cache_data_ptr_str);
  p.slot_mapping = slot_mapping.data_ptr<int64_t>();
  p.block_size = key_cache_block_size;
  p.x = key_cache_x;
  p.kv_dtype = kv;
  p.kv_scale = kv_scale;
  const at::cuda::OptionalCUDAGuard device_guard(device_of(query));
  check_status(b200pa_rotary_embedding(&p, current_stream()), nullptr, query, kv_cache_dtype);
}

// csrc/cache.h:11-13
void copy_blocks(std::vector<torch::Tensor>& key_caches, std::vector<torch::Tensor>& value_caches,
                 const torch::Tensor& block_mapping) {
  const int num_layers = key_caches.size();
  TORCH_CHECK(num_layers == (int)value_caches.size());
  if (num_layers == 0) return;
  torch::Device cache_device = key_caches[0].device();
  TORCH_CHECK(cache_device.is_cuda());
  TORCH_CHECK(block_mapping.device() == cache_device, "block_mapping must be on the same device as the caches");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.is_contiguous(),
              "block_mapping must be contiguous int64 [num_pairs, 2]");
  std::vector<void*> kptrs(num_layers), vptrs(num_layers);
  for (int l = 0; l < num_layers; ++l) {
    kptrs[l] = key_caches[l].data_ptr();
    vptrs[l] = value_caches[l].data_ptr();
  }
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  const int64_t num_pairs = block_mapping.size(0);
  const at::cuda::OptionalCUDAGuard device_guard(cache_device);
  check_status(b200pa_copy_blocks(kptrs.data(), vptrs.data(), num_layers, block_mapping.data_ptr<int64_t>(), num_pairs,
                                  block_bytes, current_stream()),
               nullptr, key_caches[0], "auto");
}

// csrc/cache.h:8-9, device rules of cache_kernels.cu:27-45
void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping) {
  torch::Device src_device = src.device();
  torch::Device dst_device = dst.device();
  int kind;
  if (src_device.is_cuda() && dst_device.is_cuda()) {
    TORCH_CHECK(src_device.index() == dst_device.index(), "src and dst must be on the same GPU");
    kind = B200PA_D2D;
  } else if (src_device.is_cuda() && dst_device.is_cpu()) {
    kind = B200PA_D2H;
  } else if (src_device.is_cpu() && dst_device.is_cuda()) {
    kind = B200PA_H2D;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long, "block_mapping must be int64");
  const torch::Tensor mapping = block_mapping.contiguous();
  const int64_t block_bytes = src.element_size() * src[0].numel();
  const at::cuda::OptionalCUDAGuard device_guard(src_device.is_cuda() ? src_device : dst_device);
  check_status(b200pa_swap_blocks(src.data_ptr(), dst.data_ptr(), mapping.data_ptr<int64_t>(), mapping.size(0),
                                  block_bytes, kind, current_stream()),
               nullptr, src, "auto");
}

// csrc/cache.h:41-42
void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const float scale,
                 const std::string& kv_cache_dtype) {
  torch::Device src_device = src_cache.device();
  torch::Device dst_device = dst_cache.device();
  TORCH_CHECK(src_device.is_cuda(), "src must be on a GPU");
  TORCH_CHECK(dst_device.is_cuda(), "dst must be on a GPU");
  TORCH_CHECK(src_device.index() == dst_device.index(), "src and dst must be on the same GPU");
  int kv;
  if (kv_cache_dtype == "fp8" || kv_cache_dtype == "fp8_e4m3") {
    kv = B200PA_KV_FP8_E4M3;
  } else {
    // "auto" reaches assert(false) in the reference's scaled_convert (quant_utils.cuh:513-523)
    TORCH_CHECK(false, "Unsupported data type: ", kv_cache_dtype);
  }
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous() && src_cache.numel() == dst_cache.numel(),
              "convert_fp8: caches must be contiguous and of equal numel");
  const int sd = scalar_dtype(src_cache), dd = scalar_dtype(dst_cache);
  const at::cuda::OptionalCUDAGuard device_guard(src_device);
  check_status(b200pa_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(), dd, sd, scale, kv,
                                  current_stream()),
               nullptr, sd == B200PA_U8 ? dst_cache : src_cache, kv_cache_dtype);
}

int get_max_shared_memory_per_block_device_attribute(int device_id) {
  const int v = b200pa_get_max_shared_memory_per_block_device_attribute(device_id);
  TORCH_CHECK(v >= 0, "cudaDeviceGetAttribute failed for device ", device_id);
  return v;
}

// csrc/cuda_utils.h / cuda_utils_kernels.cu:5-15
int get_device_attribute(int attribute, int device_id) {
  const int v = b200pa_get_device_attribute(attribute, device_id);
  TORCH_CHECK(v >= 0, "cudaDeviceGetAttribute(", attribute, ") failed for device ", device_id);
  return v;
}

// vllm/attention/ops/prefix_prefill.py:684-696 (context_attention_fwd), same argument order.  The
// index tensors may be int32 (vLLM's metadata) or int64 (the reference's kernel test); int64 ones are
// narrowed here.  scale = 1/sqrt(head_size) as the reference's wrapper fixes it (:706).
void context_attention_fwd(torch::Tensor& q, torch::Tensor& k, torch::Tensor& v, torch::Tensor& o,
                           torch::Tensor& k_cache, torch::Tensor& v_cache, torch::Tensor& b_loc,
                           torch::Tensor& b_start_loc, torch::Tensor& b_seq_len, torch::Tensor& b_ctx_len,
                           int max_input_len, const c10::optional<torch::Tensor>& alibi_slopes,
                           const c10::optional<int64_t>& sliding_window, const std::string& kv_cache_dtype,
                           float kv_scale) {
  TORCH_CHECK(q.is_cuda() && k.is_cuda() && v.is_cuda() && o.is_cuda() && k_cache.is_cuda() && v_cache.is_cuda() &&
                  b_loc.is_cuda() && b_start_loc.is_cuda() && b_seq_len.is_cuda() && b_ctx_len.is_cuda(),
              "b200pa context_attention_fwd: tensors must be on a CUDA device (there is no CPU path)");
  TORCH_CHECK(q.dim() == 3 && k.dim() == 3 && v.dim() == 3 && o.dim() == 3, "q, k, v, o must be [tokens, heads, head_size]");
  TORCH_CHECK(q.size(2) == k.size(2) && k.size(2) == v.size(2), "q, k, v head sizes differ");  // :702-703
  TORCH_CHECK(q.stride(2) == 1 && k.stride(2) == 1 && v.stride(2) == 1 && o.stride(2) == 1,
              "q, k, v, o must be contiguous in head_size");
  TORCH_CHECK(k.scalar_type() == q.scalar_type() && v.scalar_type() == q.scalar_type() &&
                  o.scalar_type() == q.scalar_type(), "q, k, v, o dtypes differ");
  TORCH_CHECK(k_cache.dim() == 5 && v_cache.dim() == 4, "k_cache must be [nb, H, D/x, B, x], v_cache [nb, H, D, B]");
  TORCH_CHECK(k_cache.stride(4) == 1 && k_cache.stride(3) == k_cache.size(4) &&
                  k_cache.stride(2) == k_cache.size(3) * k_cache.size(4) && v_cache.stride(3) == 1 &&
                  v_cache.stride(2) == v_cache.size(3),
              "cache blocks must be contiguous per (block, head)");
  TORCH_CHECK(b_loc.dim() == 2, "b_loc must be [batch, max_blocks]");
  const at::cuda::OptionalCUDAGuard device_guard(device_of(q));
  auto as_i32 = [](torch::Tensor& t) {
    TORCH_CHECK(t.scalar_type() == at::ScalarType::Int || t.scalar_type() == at::ScalarType::Long,
                "index tensors must be int32 or int64");
    torch::Tensor r = t.scalar_type() == at::ScalarType::Int ? t : t.to(at::ScalarType::Int);
    return r.stride(-1) == 1 ? r : r.contiguous();
  };
  torch::Tensor loc = as_i32(b_loc), start = as_i32(b_start_loc), seql = as_i32(b_seq_len), ctxl = as_i32(b_ctx_len);
  b200pa_prefill_params_t p{};
  p.out = o.data_ptr();
  p.query = q.data_ptr();
  p.key = k.data_ptr();
  p.value = v.data_ptr();
  p.key_cache = k_cache.data_ptr();
  p.value_cache = v_cache.data_ptr();
  p.block_tables = loc.data_ptr<int>();
  p.start_loc = start.data_ptr<int>();
  p.seq_lens = seql.data_ptr<int>();
  p.ctx_lens = ctxl.data_ptr<int>();
  if (alibi_slopes.has_value()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_cuda() &&
                    alibi_slopes->is_contiguous(), "alibi_slopes must be a contiguous float32 CUDA tensor");
    p.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  p.q_stride_t = q.stride(0); p.q_stride_h = q.stride(1);
  p.k_stride_t = k.stride(0); p.k_stride_h = k.stride(1);
  p.v_stride_t = v.stride(0); p.v_stride_h = v.stride(1);
  p.o_stride_t = o.stride(0); p.o_stride_h = o.stride(1);
  p.kc_block_stride = k_cache.stride(0); p.kc_head_stride = k_cache.stride(1);
  p.vc_block_stride = v_cache.stride(0); p.vc_head_stride = v_cache.stride(1);
  p.block_table_stride = loc.stride(0);
  p.batch = seql.size(0);
  p.max_query_len = max_input_len;
  p.num_heads = q.size(1);
  p.num_kv_heads = k.size(1);
  p.head_size = q.size(2);
  p.block_size = v_cache.size(3);
  p.sliding_window = sliding_window.has_value() ? static_cast<int>(*sliding_window) : 0;
  p.dtype = scalar_dtype(q);
  const int cdt = scalar_dtype(k_cache);
  TORCH_CHECK(k_cache.size(4) * k_cache.element_size() == 16 && v_cache.scalar_type() == k_cache.scalar_type(),
              "context_attention_fwd: k_cache must be [nb, H, D/x, B, x] with x = 16 bytes, v_cache of the same dtype");
  if (cdt == B200PA_U8) {
    // uint8 cache = fp8.  With kv_cache_dtype "auto" (the reference wrapper has no such argument) the bytes are
    // read as e5m2 without a scale, which is what the reference's Triton kernels do (prefix_prefill.py:122-123).
    p.kv_dtype = kv_cache_dtype == "auto" ? B200PA_KV_FP8_E5M2 : parse_kv_dtype(kv_cache_dtype);
    p.kv_scale = kv_cache_dtype == "auto" ? 1.0f : kv_scale;
    TORCH_CHECK(p.kv_dtype != B200PA_KV_AUTO, "Unsupported data type of kv cache: ", kv_cache_dtype);
  } else {
    TORCH_CHECK(cdt == p.dtype && kv_cache_dtype == "auto",
                "context_attention_fwd: a 16-bit paged cache must have the query dtype and kv_cache_dtype=\"auto\"");
    p.kv_dtype = B200PA_KV_AUTO;
    p.kv_scale = 1.0f;
  }
  p.scale = 1.0f / std::sqrt(static_cast<float>(p.head_size));
  const int rc = b200pa_context_attention_fwd(&p, current_stream());
  if (rc == B200PA_ERR_HEAD_SIZE) TORCH_CHECK(false, "Unsupported head size: ", p.head_size);
  if (rc == B200PA_ERR_BLOCK_SIZE) TORCH_CHECK(false, "Unsupported block size: ", p.block_size);
  check_status(rc, nullptr, q, "auto");
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  // csrc/pybind.cpp:8-14
  pybind11::module ops = m.def_submodule("ops", "vLLM custom operators (B200 paged-attention hot path)");
  ops.def("paged_attention_v1", &paged_attention_v1,
          "Compute the attention between an input query and the cached keys/values using PagedAttention.");
  ops.def("paged_attention_v2", &paged_attention_v2, "PagedAttention V2.");
  ops.def("paged_attention_flash", &paged_attention_flash,
          "Decode attention over the flash-attention cache layout [num_blocks, block_size, num_kv_heads, head_size]");
  // csrc/pybind.cpp:33-38
  ops.def("context_attention_fwd", &context_attention_fwd,
          "Prefill attention of new tokens over the cached prefix (paged cache) + themselves (causal)",
          pybind11::arg("q"), pybind11::arg("k"), pybind11::arg("v"), pybind11::arg("o"), pybind11::arg("k_cache"),
          pybind11::arg("v_cache"), pybind11::arg("b_loc"), pybind11::arg("b_start_loc"), pybind11::arg("b_seq_len"),
          pybind11::arg("b_ctx_len"), pybind11::arg("max_input_len"), pybind11::arg("alibi_slopes") = pybind11::none(),
          pybind11::arg("sliding_window") = pybind11::none(), pybind11::arg("kv_cache_dtype") = "auto",
          pybind11::arg("kv_scale") = 1.0f);
  ops.def("rotary_embedding", &rotary_embedding, "Apply GPT-NeoX or GPT-J style rotary embedding to query and key");
  ops.def("batched_rotary_embedding", &batched_rotary_embedding,
          "Apply GPT-NeoX or GPT-J style rotary embedding to query and key (batched cos/sin offsets)");
  ops.def("rotary_embedding_reshape_and_cache", &rotary_embedding_reshape_and_cache,
          "rotary_embedding on query/key fused with reshape_and_cache of the rotated key and the value");
  // csrc/pybind.cpp:74-84
  pybind11::module cache_ops = m.def_submodule("cache_ops", "vLLM cache ops (B200)");
  cache_ops.def("swap_blocks", &swap_blocks, "Swap in (out) the cache blocks from src to dst");
  cache_ops.def("copy_blocks", &copy_blocks, "Copy the cache blocks from src to dst");
  cache_ops.def("reshape_and_cache", &reshape_and_cache, "Reshape the key and value tensors and cache them");
  cache_ops.def("reshape_and_cache_flash", &reshape_and_cache_flash,
                "Reshape the key and value tensors and cache them (flash-attention layout)");
  cache_ops.def("convert_fp8", &convert_fp8, "Convert the key and value cache to fp8 data type");
  // csrc/pybind.cpp:87-95
  pybind11::module cuda_utils = m.def_submodule("cuda_utils", "vLLM cuda utils (B200)");
  cuda_utils.def("get_device_attribute", &get_device_attribute, "Gets the specified device attribute.");
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute", &get_max_shared_memory_per_block_device_attribute,
                 "Gets the maximum shared memory per block device attribute.");
  // diagnostics (not in the reference)
  m.def("last_attention_path", []() { return std::string(b200pa_last_attention_path()); });
  m.def("last_attention_partitions", []() { return b200pa_last_attention_partitions(); });
  m.def("abi_version", []() { return b200pa_abi_version(); });
  m.def("read_stream_probe", [](torch::Tensor& src, torch::Tensor& sink) {
    TORCH_CHECK(src.is_cuda() && src.is_contiguous() && sink.is_cuda() && sink.numel() >= 1 && sink.element_size() == 4,
                "read_stream_probe: contiguous CUDA tensor + a 4-byte CUDA sink");
    const at::cuda::OptionalCUDAGuard device_guard(device_of(src));
    check_status(b200pa_read_stream_probe(src.data_ptr(), src.numel() * src.element_size(), sink.data_ptr(),
                                          current_stream()), nullptr, src, "");
  }, "Measurement aid: stream a tensor once with read-only 128-bit loads");
  m.def("prepare", [](int device) { check_status(b200pa_prepare(device), nullptr, torch::Tensor(), ""); },
        "Allocate the per-device attention workspace now (e.g. before a CUDA-graph capture)");
  m.def("set_tuning", [](const std::string& key, int value) {
    TORCH_CHECK(b200pa_set_tuning(key.c_str(), value) == B200PA_OK, "b200pa: unknown tuning key ", key);
  }, "Developer hook: override a tuning knob (tools/tune.py)");
}
