"""Python wrappers of the hot-path custom ops — same names, argument order, defaults and
marshalling as the reference's `vllm/_custom_ops.py:37-99, 304-354` (including the fork's
habit of passing the KV-cache pointers of paged_attention_v2 / reshape_and_cache as decimal strings
so duck-typed managed-memory caches work: vllm/worker/cache_engine/map_cache_engine.py:35-66).

The ops come from the compiled sm_100a extension; importing this module fails loudly without it.
"""
from typing import List, Optional

import torch

from . import load_extension

_C = load_extension()
vllm_ops = _C.ops
vllm_cache_ops = _C.cache_ops
vllm_cuda_utils = _C.cuda_utils


# page attention ops
def paged_attention_v1(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    kv_scale: float,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    vllm_ops.paged_attention_v1(
        out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
        seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
        kv_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
        blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_v2(
    out: torch.Tensor,
    exp_sum: torch.Tensor,
    max_logits: torch.Tensor,
    tmp_out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    kv_scale: float,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    # key_cache / value_cache only need data_ptr() and stride(): tensors or ManagedTensor-likes.
    vllm_ops.paged_attention_v2(
        out, exp_sum, max_logits, tmp_out, query,
        str(key_cache.data_ptr()), key_cache.stride(0), key_cache.stride(1),
        str(value_cache.data_ptr()),
        num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
        alibi_slopes, kv_cache_dtype, kv_scale, tp_rank,
        blocksparse_local_blocks, blocksparse_vert_stride,
        blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_flash(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                          scale: float, block_tables: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int,
                          alibi_slopes: Optional[torch.Tensor] = None) -> None:
    """Decode attention over the flash-attention cache layout `[num_blocks, block_size, num_kv_heads, head_size]`
    (the FLASH_ATTN / FLASHINFER backends' layout; the reference decodes it with the external
    flash_attn_with_kvcache, vllm/attention/backends/flash_attn.py:354-367).  Not in the reference's op list."""
    vllm_ops.paged_attention_flash(out, query, key_cache, value_cache, scale, block_tables, seq_lens, max_seq_len,
                                   alibi_slopes)


# pos encoding ops (vllm/_custom_ops.py:103-124); in place on query and key
def rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                     cos_sin_cache: torch.Tensor, is_neox: bool) -> None:
    vllm_ops.rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox)


def batched_rotary_embedding(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                             cos_sin_cache: torch.Tensor, is_neox: bool, rot_dim: int,
                             cos_sin_cache_offsets: torch.Tensor) -> None:
    vllm_ops.batched_rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim,
                                      cos_sin_cache_offsets)


def rotary_embedding_reshape_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                       value: torch.Tensor, head_size: int, cos_sin_cache: torch.Tensor,
                                       is_neox: bool, key_cache: torch.Tensor, value_cache: torch.Tensor,
                                       slot_mapping: torch.Tensor, kv_cache_dtype: str, kv_scale: float) -> None:
    """rotary_embedding(positions, query, key, ...) followed by reshape_and_cache(key, value, ...) in ONE launch
    (SURVEY 8f rank 2); bit-identical to the two reference calls.  Not in the reference."""
    vllm_ops.rotary_embedding_reshape_and_cache(
        positions, query, key, value, head_size, cos_sin_cache, is_neox,
        str(key_cache.data_ptr()), key_cache.size(3), key_cache.size(4), str(value_cache.data_ptr()),
        slot_mapping, kv_cache_dtype, kv_scale)


# prefix-prefill attention (vllm/attention/ops/prefix_prefill.py:684-696)
def context_attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, o: torch.Tensor,
                          k_cache: torch.Tensor, v_cache: torch.Tensor, b_loc: torch.Tensor,
                          b_start_loc: torch.Tensor, b_seq_len: torch.Tensor, b_ctx_len: torch.Tensor,
                          max_input_len: int, alibi_slopes: Optional[torch.Tensor] = None,
                          sliding_window: Optional[int] = None, kv_cache_dtype: str = "auto",
                          kv_scale: float = 1.0) -> None:
    # kv_cache_dtype / kv_scale are additions (the reference wrapper has neither): an fp8 (uint8) cache is
    # dequantised while it is staged, like the decode kernels do; "auto" with a uint8 cache keeps the
    # reference kernels' e5m2 bit-cast (prefix_prefill.py:122-123).
    vllm_ops.context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len,
                                   max_input_len, alibi_slopes, sliding_window, kv_cache_dtype, kv_scale)


# cache ops
def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    kv_scale: float,
) -> None:
    vllm_cache_ops.reshape_and_cache(
        key, value,
        str(key_cache.data_ptr()), key_cache.size(3), key_cache.size(4),
        str(value_cache.data_ptr()),
        slot_mapping, kv_cache_dtype, kv_scale)


def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
) -> None:
    # Same tensor-taking signature as vllm/_custom_ops.py:326-335; the fork's wrapper forwards the tensors
    # to a pybind function that expects pointer strings (csrc/cache.h:27-39) and cannot work as shipped —
    # here the marshalling matches the C++ side (as vllm/attention/backends/flash_attn.py:288-298 does by hand).
    vllm_cache_ops.reshape_and_cache_flash(
        key, value,
        str(key_cache.data_ptr()), key_cache.size(1), key_cache.stride(0),
        str(value_cache.data_ptr()), value_cache.stride(0),
        slot_mapping, kv_cache_dtype)


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: torch.Tensor) -> None:
    vllm_cache_ops.copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor,
                block_mapping: torch.Tensor) -> None:
    vllm_cache_ops.swap_blocks(src, dst, block_mapping)


def convert_fp8(output: torch.Tensor,
                input: torch.Tensor,
                scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    vllm_cache_ops.convert_fp8(output, input, scale, kv_dtype)


# cuda utils (vllm/utils.py:149-160 calls this through vllm._C.cuda_utils)
def get_max_shared_memory_per_block_device_attribute(device_id: int) -> int:
    return vllm_cuda_utils.get_max_shared_memory_per_block_device_attribute(device_id)


# diagnostics (not in the reference)
def last_attention_path() -> str:
    return _C.last_attention_path()


def last_attention_partitions() -> int:
    return _C.last_attention_partitions()


def prepare(device: int = 0) -> None:
    """Allocate the per-device attention workspace now (optional; it is otherwise allocated by the first call)."""
    _C.prepare(device)


def set_tuning(key: str, value: int) -> None:
    """Developer hook (tools/tune.py, tests): override a launch-tuning knob of libb200pa at run time."""
    _C.set_tuning(key, value)


def read_stream_probe(src: torch.Tensor, sink: torch.Tensor) -> None:
    """Measurement aid (bench.py): read `src` once with streaming loads; `sink` is a 1-element int32 CUDA tensor."""
    _C.read_stream_probe(src, sink)
