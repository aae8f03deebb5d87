// TEST INFRASTRUCTURE — not part of the product path.
//
// pybind11 shim that exposes the reference's OWN CPU implementation of the
// hot path (compiled unchanged from /root/reference/csrc/cpu/attention.cpp and
// cache.cpp by oracle/build_ref.py) as the module `ref_cpu`.
//
// The fork's own csrc/cpu/pybind.cpp cannot be used: it includes csrc/ops.h /
// csrc/cache.h whose paged_attention_v2 / reshape_and_cache declarations take
// pointer *strings*, while csrc/cpu/attention.cpp:738 and csrc/cpu/cache.cpp:104
// define the *tensor* overloads.  We declare the tensor overloads here
// (signatures: csrc/cpu/attention.cpp:421-428, 738-746; csrc/cpu/cache.cpp:85-107).
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
    torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
    torch::Tensor& value_cache, int num_kv_heads, float scale,
    torch::Tensor& block_tables, torch::Tensor& seq_lens, int block_size,
    int max_seq_len, const c10::optional<torch::Tensor>& alibi_slopes,
    const std::string& kv_cache_dtype, float kv_scale, const int tp_rank,
    const int blocksparse_local_blocks, const int blocksparse_vert_stride,
    const int blocksparse_block_size, const int blocksparse_head_sliding_step);

void copy_blocks(std::vector<torch::Tensor>& key_caches,
                 std::vector<torch::Tensor>& value_caches,
                 const torch::Tensor& block_mapping);

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value,
                       torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, float kv_scale);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("paged_attention_v1", &paged_attention_v1);
  m.def("paged_attention_v2", &paged_attention_v2);
  m.def("copy_blocks", &copy_blocks);
  m.def("reshape_and_cache", &reshape_and_cache);
}
