// TEST INFRASTRUCTURE — not part of the product path.
//
// pybind11 shim exposing the reference's OWN CUDA kernels for the hot path
// (compiled unchanged for sm_100a from /root/reference/csrc/attention/
// attention_kernels.cu, csrc/cache_kernels.cu and csrc/pos_encoding_kernels.cu by oracle/build_ref.py) as the
// module `ref_cuda`.  It is the GPU-side oracle and "the kernel to beat"; the
// fork's csrc/pybind.cpp is not used because it pulls in every other kernel.
// Declarations follow csrc/ops.h:5-30 and csrc/cache.h:8-42.
#include <torch/extension.h>

#include <string>
#include <vector>

void paged_attention_v1(
    torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int num_kv_heads, float scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int block_size,
    int max_seq_len, const c10::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, float kv_scale, const int tp_rank,
    const int blocksparse_local_blocks, const int blocksparse_vert_stride,
    const int blocksparse_block_size, const int blocksparse_head_sliding_step);

void paged_attention_v2(
    torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
    torch::Tensor& tmp_out, torch::Tensor& query,
    const std::string& key_cache_data_ptr_str, int key_cache_stride_0,
    int key_cache_stride_1, const std::string& value_cache_data_ptr_str,
    int num_kv_heads, float scale, torch::Tensor& block_tables,
    torch::Tensor& seq_lens, int block_size, int max_seq_len,
    const c10::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, float kv_scale, const int tp_rank,
    const int blocksparse_local_blocks, const int blocksparse_vert_stride,
    const int blocksparse_block_size, const int blocksparse_head_sliding_step);

void swap_blocks(torch::Tensor& src, torch::Tensor& dst,
                 const torch::Tensor& block_mapping);

void copy_blocks(std::vector<torch::Tensor>& key_caches,
                 std::vector<torch::Tensor>& value_caches,
                 const torch::Tensor& block_mapping);

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value,
                       const std::string& key_cache_data_ptr_str,
                       int key_cache_block_size, int key_cache_x,
                       const std::string& value_cache_data_ptr_str,
                       torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, const float kv_scale);

void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache,
                 const float scale, const std::string& kv_cache_dtype);

// csrc/cache.h:27-39
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value,
                             const std::string& key_cache_data_ptr_str,
                             int key_cache_size_1, int key_cache_stride_0,
                             const std::string& value_cache_data_ptr_str,
                             int value_cache_stride_0,
                             torch::Tensor& slot_mapping,
                             const std::string& kv_cache_dtype);

// csrc/ops.h (pos_encoding_kernels.cu:125-199)
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query,
                      torch::Tensor& key, int head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox);

void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query,
                              torch::Tensor& key, int head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox,
                              int rot_dim,
                              torch::Tensor& cos_sin_cache_offsets);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("reshape_and_cache_flash", &reshape_and_cache_flash);
  m.def("rotary_embedding", &rotary_embedding);
  m.def("batched_rotary_embedding", &batched_rotary_embedding);
  m.def("paged_attention_v1", &paged_attention_v1);
  m.def("paged_attention_v2", &paged_attention_v2);
  m.def("swap_blocks", &swap_blocks);
  m.def("copy_blocks", &copy_blocks);
  m.def("reshape_and_cache", &reshape_and_cache);
  m.def("convert_fp8", &convert_fp8);
}
