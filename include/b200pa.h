/*
 * b200pa.h — C-ABI of the B200-native (sm_100a) paged-attention hot path.
 *
 * This is the drop-in boundary underneath the reference's `vllm._C` operator
 * module (DeepAuto-AI/vllm-legacy, vLLM 0.4.3 fork).  Plain pointers and sizes,
 * no torch types.  Every entry point names the reference interface it replaces.
 * All device entry points are asynchronous on `stream`, allocate nothing, do no
 * host synchronisation and are CUDA-graph capturable (swap_blocks too when both
 * sides are device-visible: it is then a gather kernel, see below).
 *
 * Return value: B200PA_OK (0) or a negative b200pa_status_t.  The PyTorch
 * binding (vllm_legacy_b200/csrc/torch_binding.cpp) maps the codes to the
 * reference's TORCH_CHECK texts (attention_kernels.cu:764,801;
 * quant_utils.cuh:538,564; cache_kernels.cu:31-45).
 */
#ifndef B200PA_H_
#define B200PA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PA_ABI_VERSION 2 /* 2: b200pa_prefill_params_t gained kv_scale (fp8 KV prefix) */

/* CUDA stream handle (== cudaStream_t == CUstream). */
typedef struct CUstream_st* b200pa_stream_t;

/* scalar_t of query/out/key/value (reference: VLLM_DISPATCH_FLOATING_TYPES,
 * csrc/dispatch_utils.h) and raw-byte cache element (uint8). */
typedef enum {
  B200PA_F16 = 0,
  B200PA_BF16 = 1,
  B200PA_F32 = 2,
  B200PA_U8 = 3
} b200pa_dtype_t;

/* kv_cache_dtype string of the reference ("auto" | "fp8"/"fp8_e4m3" | "fp8_e5m2"),
 * csrc/attention/dtype_fp8.cuh:14-18, quant_utils.cuh:529-566. */
typedef enum {
  B200PA_KV_AUTO = 0,
  B200PA_KV_FP8_E4M3 = 1,
  B200PA_KV_FP8_E5M2 = 2
} b200pa_kv_dtype_t;

typedef enum {
  B200PA_OK = 0,
  B200PA_ERR_HEAD_SIZE = -1,   /* "Unsupported head size: "            attention_kernels.cu:764 */
  B200PA_ERR_BLOCK_SIZE = -2,  /* "Unsupported block size: "           attention_kernels.cu:801 */
  B200PA_ERR_KV_DTYPE = -3,    /* "Unsupported data type of kv cache: " quant_utils.cuh:564     */
  B200PA_ERR_INPUT_DTYPE = -4, /* "Unsupported input type of kv cache: " quant_utils.cuh:538    */
  B200PA_ERR_INVALID_ARG = -5, /* null pointer / negative size / misaligned cache               */
  B200PA_ERR_CUDA = -6,        /* a CUDA runtime call failed: see b200pa_last_cuda_error()      */
  B200PA_ERR_DEVICE_COMBINATION = -7, /* "Invalid device combination"   cache_kernels.cu:41-43   */
  B200PA_ERR_SCRATCH_LIMIT = -8 /* b200pa_paged_attention_static_scratch: the problem exceeds the static scratch */
} b200pa_status_t;

/*
 * paged_attention_v1 / paged_attention_v2
 * replaces: csrc/ops.h:5-30, csrc/attention/attention_kernels.cu:805-825 (v1),
 *           :984-1014 (v2, the fork's pointer-string ABI) and the kernels at
 *           :86-496, :565-669.
 *
 * Layouts (x = 16 / sizeof(cache element)):
 *   out, query      [num_seqs, num_heads, head_size]     dtype; query row stride q_stride
 *   key_cache       [num_blocks, num_kv_heads, head_size/x, block_size, x]
 *   value_cache     [num_blocks, num_kv_heads, head_size, block_size]
 *                   both with strides kv_block_stride / kv_head_stride (elements),
 *                   innermost three/two dims contiguous (as the reference assumes)
 *   block_tables    [num_seqs, max_num_blocks_per_seq] int32
 *   seq_lens        [num_seqs] int32
 *   alibi_slopes    [num_heads] float or NULL
 *   v2 scratch      exp_sums, max_logits [num_seqs, num_heads, max_num_partitions] float
 *                   tmp_out [num_seqs, num_heads, max_num_partitions, head_size] dtype
 *                   with max_num_partitions = ceil(max_seq_len / 512) as the
 *                   reference's callers size it (vllm/attention/ops/paged_attn.py:111-158).
 */
typedef struct {
  void* out;
  const void* query;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes; /* nullable */
  float* exp_sums;           /* v2 only */
  float* max_logits;         /* v2 only */
  void* tmp_out;             /* v2 only */
  int64_t q_stride;          /* elements between query rows (query.stride(0)) */
  int64_t kv_block_stride;   /* key_cache.stride(0), elements */
  int64_t kv_head_stride;    /* key_cache.stride(1), elements */
  int32_t num_seqs;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t max_seq_len;
  int32_t max_num_blocks_per_seq;
  int32_t max_num_partitions; /* v2 only: partition dim of the scratch buffers */
  int32_t dtype;              /* b200pa_dtype_t of query/out */
  int32_t kv_dtype;           /* b200pa_kv_dtype_t */
  float scale;
  float kv_scale;
  int32_t tp_rank;
  int32_t blocksparse_local_blocks;
  int32_t blocksparse_vert_stride; /* > 1 enables block-sparse */
  int32_t blocksparse_block_size;
  int32_t blocksparse_head_sliding_step;
} b200pa_attn_params_t;

int b200pa_paged_attention_v1(const b200pa_attn_params_t* p, b200pa_stream_t stream);
int b200pa_paged_attention_v2(const b200pa_attn_params_t* p, b200pa_stream_t stream);

/*
 * paged_attention_static_scratch — split-K decode attention that allocates nothing
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and arithmetic as b200pa_paged_attention_v1 on the TMA + mma.sync kernel family (dtype
 * f16 | bf16, block_size 16 | 32, dense, kv_dtype auto | fp8); exp_sums / max_logits / tmp_out /
 * max_num_partitions of the parameter block are ignored.  Where v1 keeps its split-K partials and arrival
 * counters in a per-device workspace that its first call cudaMallocs, this entry point keeps them in ONE
 * file-scope `static __device__` array inside the library (csrc/attention_noalloc.cu): it calls no cudaMalloc,
 * cudaFree or allocating library routine, on any call.  Every GPU that loads the library holds the array
 * (B200PA_STATIC_SCRATCH_BYTES + 4 * B200PA_STATIC_SCRATCH_COUNTERS bytes) until exit.
 * Limits — beyond them the call returns B200PA_ERR_SCRATCH_LIMIT and launches nothing:
 *   num_seqs * num_kv_heads * ceil(group / (group <= 8 ? 8 : 16))   <= B200PA_STATIC_SCRATCH_COUNTERS
 *   2 * num_seqs * num_heads * (head_size * 2 + 8)                  <= B200PA_STATIC_SCRATCH_BYTES
 * (group = num_heads / num_kv_heads; i.e. two context splits of the whole batch must fit.  Larger batches fill
 * the GPU without a split: use b200pa_paged_attention_v1 / _v2 for them.)
 * All calls share the array and it keeps the last call's contents: the counters are cleared on `stream` at the
 * start of every call (a memset node under capture), the partials are written before they are read.  Calls that
 * OVERLAP on the GPU (different streams, concurrently replayed graphs) race on it — serialise them.
 */
#define B200PA_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
#define B200PA_STATIC_SCRATCH_COUNTERS 4096
int b200pa_paged_attention_static_scratch(const b200pa_attn_params_t* p, b200pa_stream_t stream);

/*
 * paged_attention_flash — decode over the flash-attention cache layout
 * replaces: flash_attn_with_kvcache(q.unsqueeze(1), key_cache, value_cache, block_table=..., cache_seqlens=...,
 *           softmax_scale=..., causal=True, alibi_slopes=...) as called for decode by the FLASH_ATTN backend
 *           (vllm/attention/backends/flash_attn.py:354-367); the cache is written by reshape_and_cache_flash.
 *   key_cache, value_cache  [num_blocks, block_size, num_kv_heads, head_size] dtype (f16 | bf16), with
 *                           kv_block_stride = stride(0), kv_token_stride = stride(1), kv_head_stride = stride(2)
 *                           in elements, head_size contiguous, rows 16-byte aligned
 *   head_size in {32, 64, ..., 256} (multiples of 32), block_size 16 | 32, kv_dtype AUTO, no block-sparse.
 * Everything else as paged_attention_v1 (no caller scratch: long contexts split through the per-device workspace).
 */
int b200pa_paged_attention_flash(const b200pa_attn_params_t* p, int64_t kv_token_stride, b200pa_stream_t stream);

/*
 * reshape_and_cache
 * replaces: csrc/cache.h:15-25, csrc/cache_kernels.cu:150-202, 241-287.
 *   key, value   [num_tokens, num_heads, head_size] dtype, row strides key_stride/value_stride
 *   key_cache    [num_blocks, num_heads, head_size/x, block_size, x]   (contiguous)
 *   value_cache  [num_blocks, num_heads, head_size, block_size]        (contiguous)
 *   slot_mapping [num_tokens] int64, slot < 0 = padding token (skipped)
 * kv_dtype AUTO stores dtype elements; FP8_* stores uint8 = cvt(float(v) / kv_scale).
 */
typedef struct {
  const void* key;
  const void* value;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  int64_t key_stride;
  int64_t value_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t x;
  int32_t dtype;    /* b200pa_dtype_t of key/value */
  int32_t kv_dtype; /* b200pa_kv_dtype_t */
  float kv_scale;
} b200pa_reshape_params_t;

int b200pa_reshape_and_cache(const b200pa_reshape_params_t* p, b200pa_stream_t stream);

/*
 * reshape_and_cache_flash  (flash-attention cache layout; "auto" KV dtype only, as the reference)
 * replaces: csrc/cache.h:27-39, csrc/cache_kernels.cu:204-235 (kernel), :289-345 (host).
 *   key, value   [num_tokens, num_heads, head_size] dtype, row strides key_stride/value_stride
 *   key_cache    [num_blocks, block_size, num_heads, head_size], block stride block_stride (elements)
 *   value_cache  same shape and block stride
 *   slot_mapping [num_tokens] int64, slot < 0 = padding token (skipped)
 */
typedef struct {
  const void* key;
  const void* value;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  int64_t key_stride;
  int64_t value_stride;
  int64_t block_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t dtype; /* b200pa_dtype_t of key/value/cache */
} b200pa_reshape_flash_params_t;

int b200pa_reshape_and_cache_flash(const b200pa_reshape_flash_params_t* p, b200pa_stream_t stream);

/*
 * rotary_embedding / batched_rotary_embedding, optionally FUSED with reshape_and_cache
 * replaces: csrc/ops.h:38-47, csrc/pos_encoding_kernels.cu:10-201 (GPT-NeoX and GPT-J styles,
 *           in place on query and key) and, when key_cache != NULL, additionally the
 *           reshape_and_cache call that follows it in the decode step
 *           (vllm/model_executor/models/llama.py:164-171 -> vllm/attention/ops/paged_attn.py:62-79,
 *           csrc/cache_kernels.cu:150-202): the rotated key and the value of every token with
 *           slot >= 0 are written to the paged cache by the same launch.  Results are bit-identical
 *           to the two reference launches run back to back.
 *   positions             : [num_tokens] int64
 *   query / key           : [num_tokens, num_heads|num_kv_heads, head_size], token stride in elements
 *   cos_sin_cache         : [max_position, rot_dim] (cos half, sin half), same dtype as query
 *   cos_sin_cache_offsets : NULL (rotary_embedding) or [num_tokens] int64 (batched variant)
 *   value/key_cache/value_cache/slot_mapping/block_size/x/kv_dtype/kv_scale : as reshape_and_cache;
 *                           key_cache == NULL selects the plain rotary op.
 * Arithmetic follows the reference exactly: every product and sum is rounded to the tensor dtype
 * (c10::Half / c10::BFloat16 operators), fp32 uses fma(x, cos, -(y*sin)) / fma(y, cos, x*sin) as
 * nvcc contracts the reference expression.
 */
typedef struct {
  const int64_t* positions;
  void* query;
  void* key;
  const void* cos_sin_cache;
  const int64_t* cos_sin_cache_offsets;
  int64_t query_stride;
  int64_t key_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t rot_dim;
  int32_t is_neox;
  int32_t dtype; /* b200pa_dtype_t of query/key/value/cos_sin_cache */
  /* fused cache write (all NULL / 0 for the plain op) */
  const void* value;
  int64_t value_stride;
  void* key_cache;
  void* value_cache;
  const int64_t* slot_mapping;
  int32_t block_size;
  int32_t x;
  int32_t kv_dtype; /* b200pa_kv_dtype_t */
  float kv_scale;
} b200pa_rope_params_t;

int b200pa_rotary_embedding(const b200pa_rope_params_t* p, b200pa_stream_t stream);

/*
 * context_attention_fwd (prefix-prefill / chunked-prefill attention over the paged cache)
 * replaces: vllm/attention/ops/prefix_prefill.py:684-817 (host wrapper) and the Triton kernels
 *           _fwd_kernel (:10-245) and _fwd_kernel_alibi (:442-681); called from
 *           vllm/attention/ops/paged_attn.py:185-216 (PagedAttention.forward_prefix).
 *
 * For sequence b with ctx = ctx_lens[b] cached tokens and q_len = seq_lens[b] - ctx new tokens whose
 * rows start at start_loc[b] in query/key/value/out, new token m attends to the cached keys 0..ctx-1
 * (read through block_tables[b] from the paged cache) and to the new keys 0..m (causal):
 *   out[m] = softmax_j(scale * q[m].k[j] + bias) . v[j]
 * bias: ALiBi slope_h * (kpos - qpos) when alibi_slopes != NULL (the window is then ignored, as in the
 * reference, :717-760); else keys with qpos - kpos >= sliding_window are masked when sliding_window > 0.
 *
 * Layouts:
 *   query, out     [num_tokens, num_heads, head_size]      strides (q|o)_stride_t / _h, innermost contiguous
 *   key, value     [num_tokens, num_kv_heads, head_size]   strides (k|v)_stride_t / _h, 16-byte aligned rows
 *   key_cache      [num_blocks, num_kv_heads, head_size/x, block_size, x]   x = 8 (16-bit cache) | 16 (fp8 cache)
 *   value_cache    [num_blocks, num_kv_heads, head_size, block_size]
 *   block_tables   [batch, block_table_stride] int32; start_loc, seq_lens, ctx_lens [batch] int32
 * head_size: any multiple of 8 up to 256 (16 with an fp8 cache); block_size: any power of two >= 8;
 * dtype f16 | bf16.
 * kv_dtype: B200PA_KV_AUTO (cache has the query dtype) or fp8 e4m3 / e5m2 (uint8 cache).  fp8 bytes are
 * dequantised while the tile is staged, with the decode kernels' conversion (fp8 -> half -> float x kv_scale
 * -> dtype, quant_utils.cuh:293-454), so prefill and decode see the same K/V values.  The reference's own
 * fp8 branch here is an e5m2 bit-cast without a scale (prefix_prefill.py:122-123, 168-169) that its callers
 * document as broken (vllm/attention/backends/xformers.py:301-304); kv_dtype = e5m2 with kv_scale = 1 is
 * that behaviour, everything else is what the fork's fp8 e4m3 caches need (BASELINE config 4).
 */
typedef struct {
  void* out;
  const void* query;
  const void* key;
  const void* value;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* start_loc;
  const int32_t* seq_lens;
  const int32_t* ctx_lens;
  const float* alibi_slopes; /* nullable, [num_heads] */
  int64_t q_stride_t, q_stride_h;
  int64_t k_stride_t, k_stride_h;
  int64_t v_stride_t, v_stride_h;
  int64_t o_stride_t, o_stride_h;
  int64_t kc_block_stride, kc_head_stride; /* key_cache.stride(0), .stride(1), elements */
  int64_t vc_block_stride, vc_head_stride; /* value_cache.stride(0), .stride(1), elements */
  int64_t block_table_stride;
  int32_t batch;
  int32_t max_query_len; /* the reference's max_input_len: upper bound of q_len over the batch */
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t sliding_window; /* <= 0 disables */
  int32_t dtype;          /* b200pa_dtype_t */
  int32_t kv_dtype;       /* b200pa_kv_dtype_t */
  float scale;            /* the reference uses 1/sqrt(head_size) (:706) */
  float kv_scale;         /* fp8 caches only (ignored for B200PA_KV_AUTO) */
} b200pa_prefill_params_t;

int b200pa_context_attention_fwd(const b200pa_prefill_params_t* p, b200pa_stream_t stream);

/*
 * copy_blocks
 * replaces: csrc/cache.h:11-13, csrc/cache_kernels.cu:68-146.
 *   key_cache_ptrs / value_cache_ptrs : HOST arrays of num_layers device pointers
 *   block_mapping                     : DEVICE int64 [num_pairs, 2] (src, dst)
 *   block_bytes                       : bytes of one block of one layer's K (== V) cache
 * The pointer arrays travel as kernel parameters: no H2D copy, no sync.
 */
int b200pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs,
                       int32_t num_layers, const int64_t* block_mapping,
                       int64_t num_pairs, int64_t block_bytes, b200pa_stream_t stream);

/*
 * swap_blocks
 * replaces: csrc/cache.h:8-9, csrc/cache_kernels.cu:25-64.
 *   block_mapping : HOST int64 [num_pairs, 2]
 *   kind          : 0 device->device, 1 device->host, 2 host->device
 * When both sides are visible to the device (device memory, pinned or mapped host
 * memory) and block_bytes % 16 == 0, the copy is ONE gather kernel per 1024 pairs:
 * the host-resident mapping travels as kernel parameters and every (pair, slice)
 * is copied by its own CTA with 128-bit loads/stores over NVLink-C2C / PCIe.
 * PRECONDITION of that path: destination blocks are pairwise distinct (the block
 * manager guarantees it, vllm/core/block_manager_v1.py swap_in/swap_out).  The
 * reference's sequential loop gives last-writer-wins for duplicate destinations;
 * the parallel gather leaves such a block undefined.
 * Otherwise (pageable host memory, odd sizes, B200PA_SWAP_MEMCPY set at load):
 * one cudaMemcpyAsync per run of consecutive (src+1, dst+1) pairs, in mapping
 * order, i.e. the reference's semantics including duplicates.
 */
typedef enum { B200PA_D2D = 0, B200PA_D2H = 1, B200PA_H2D = 2 } b200pa_copy_kind_t;

int b200pa_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host,
                       int64_t num_pairs, int64_t block_bytes, int32_t kind,
                       b200pa_stream_t stream);

/*
 * convert_fp8 ("only for testing" in the reference)
 * replaces: csrc/cache.h:41-42, csrc/cache_kernels.cu:347-420.
 * Exactly one of src_dtype / dst_dtype is B200PA_U8; the other is F16/BF16/F32.
 *   to fp8  : dst = cvt_satfinite(float(src) / scale)
 *   from fp8: dst = T(float(half(src)) * scale)
 */
int b200pa_convert_fp8(void* dst, const void* src, int64_t numel, int32_t dst_dtype,
                       int32_t src_dtype, float scale, int32_t kv_dtype,
                       b200pa_stream_t stream);

/* replaces: csrc/cuda_utils.h, csrc/cuda_utils_kernels.cu:17-29 */
int b200pa_get_max_shared_memory_per_block_device_attribute(int32_t device_id);

/* replaces: csrc/cuda_utils.h, csrc/cuda_utils_kernels.cu:5-15 (device_id < 0 = current device) */
int b200pa_get_device_attribute(int32_t attribute, int32_t device_id);

/*
 * Attention workspace.  The tensor-core decode kernel splits long contexts over several CTAs and
 * fuses the split-K reduce into the same launch (the last split of a (sequence, KV head) to finish
 * merges them).  That needs one arrival counter per (sequence, KV head) — and, for
 * paged_attention_v1 whose ABI carries no scratch, somewhere to keep the partial results — in a
 * per-device workspace (64 MB scratch + 4 MB counters), allocated by the first attention call on a
 * device and kept for the life of the process.  b200pa_prepare() allocates it explicitly (hosts
 * that capture their very first call into a CUDA graph may prefer that, although the lazy path is
 * capture-safe too).  Attention calls on ONE device must not run concurrently on different streams
 * (the reference's worker issues them on one stream).
 */
int b200pa_prepare(int32_t device);

/* Developer hook (tools/tune.py, tests): override a tuning knob at run time.  The same knobs are
 * read from the environment (B200PA_<KEY>) once, when the library is first used — never per call.
 * Keys: stages, evict_normal, force_simt, partition_size, max_partition, split_cost. */
int b200pa_set_tuning(const char* key, int32_t value);

/* Measurement aid (bench.py): read `bytes` of device memory once with streaming 128-bit loads — the
 * read-only HBM stream rate a KV decode can be held against (MEASURED_PEAKS' figure is a read+write
 * copy).  sink_u32: one device uint32 the kernel may increment. */
int b200pa_read_stream_probe(const void* src, int64_t bytes, void* sink_u32, b200pa_stream_t stream);

/* Diagnostics */
int b200pa_abi_version(void);
const char* b200pa_status_string(int status);
int b200pa_last_cuda_error(void); /* cudaError_t of the last B200PA_ERR_CUDA on this thread */
/* Name of the kernel family the last attention call on this thread dispatched to — "mma_tma" (paged layout: dense,
 * block-sparse and block-8 instantiations of the TMA + mma.sync kernel), "mma_flash" (flash layout), "mma_tma_static"
 * (static-scratch entry point) or "simt" (general family) — and the number of context splits it used; for tests and
 * the bench. */
const char* b200pa_last_attention_path(void);
int b200pa_last_attention_partitions(void);

#ifdef __cplusplus
}
#endif
#endif /* B200PA_H_ */
