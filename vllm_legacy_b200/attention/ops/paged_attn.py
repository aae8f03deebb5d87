"""`PagedAttention` static helpers — the interface of the reference's
`vllm/attention/ops/paged_attn.py:30-239` for the decode path, on top of the B200 ops.

Same method names, arguments, KV-cache shape `(2, num_blocks, block_size*num_kv_heads*head_size)`,
K view `[nb, H, D/x, B, x]` / V view `[nb, H, D, B]`, V1/V2 selection rule and scratch shapes, so the
reference's backends (xformers.py:269-335, hip.py:166-319, blocksparse_attn.py:357-390,
torch_sdpa.py:162-216) and cache engine (cache_engine.py:89-100) can call it unchanged.
`forward_prefix` runs the sm_100a prefix-prefill kernel (csrc/prefill.cu) instead of the reference's Triton one.
"""
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from vllm_legacy_b200 import _custom_ops as ops

# Should be the same as PARTITION_SIZE in `paged_attention_v2_launcher` (paged_attn.py:10).
_PARTITION_SIZE = 512


@dataclass
class PagedAttentionMetadata:
    """Metadata for PagedAttention (paged_attn.py:13-27)."""
    seq_lens_tensor: Optional[torch.Tensor]
    max_decode_seq_len: int
    block_tables: Optional[torch.Tensor]


class PagedAttention:

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [64, 80, 96, 112, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int,
                       head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        x = 16 // kv_cache.element_size()
        num_blocks = kv_cache.shape[1]
        key_cache = kv_cache[0]
        key_cache = key_cache.view(num_blocks, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1]
        value_cache = value_cache.view(num_blocks, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                             value_cache: torch.Tensor, slot_mapping: torch.Tensor,
                             kv_cache_dtype: str, kv_scale: float) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping.flatten(),
                              kv_cache_dtype, kv_scale)

    @staticmethod
    def _partitions(max_seq_len: int) -> int:
        """Scratch partitions the callers size for v2: ceil(max_seq_len / 512) (paged_attn.py:111)."""
        return -(-max_seq_len // _PARTITION_SIZE)

    @staticmethod
    def _pick_v1(num_seqs: int, num_heads: int, max_seq_len: int) -> bool:
        """The reference's V1/V2 rule (paged_attn.py:112-121), same outcome for every input: V2 whenever the
        context exceeds 8192 tokens; below that V1 if a single partition covers it or if the batch already has
        more than 512 (sequence, head) pairs to fill the GPU."""
        if max_seq_len > 8192:
            return False
        return PagedAttention._partitions(max_seq_len) == 1 or num_seqs * num_heads > 512

    @staticmethod
    def forward_decode(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                       block_tables: torch.Tensor, seq_lens: torch.Tensor, max_seq_len: int, kv_cache_dtype: str,
                       num_kv_heads: int, scale: float, alibi_slopes: Optional[torch.Tensor], kv_scale: float,
                       tp_rank: int = 0, blocksparse_local_blocks: int = 0, blocksparse_vert_stride: int = 0,
                       blocksparse_block_size: int = 64, blocksparse_head_sliding_step: int = 0) -> torch.Tensor:
        """One decode step over the paged cache (paged_attn.py:87-183): query [S, Hq, D] -> output [S, Hq, D]."""
        num_seqs, num_heads, head_size = query.shape
        block_size = value_cache.shape[3]
        sparse = blocksparse_vert_stride is not None and blocksparse_vert_stride > 1
        if sparse and not (blocksparse_block_size > 0 and blocksparse_block_size % block_size == 0):
            raise AssertionError(f"{blocksparse_block_size=} needs to be a multiple of {block_size=} used in block_tables.")
        tail = (num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
                kv_scale, tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                blocksparse_head_sliding_step)
        out = torch.empty_like(query)
        if PagedAttention._pick_v1(num_seqs, num_heads, max_seq_len):
            ops.paged_attention_v1(out, query, key_cache, value_cache, *tail)
            return out
        assert _PARTITION_SIZE % block_size == 0
        parts = PagedAttention._partitions(max_seq_len)
        stats_shape = (num_seqs, num_heads, parts)
        partial_out = torch.empty(size=(*stats_shape, head_size), dtype=out.dtype, device=out.device)
        exp_sums = torch.empty(size=stats_shape, dtype=torch.float32, device=out.device)
        max_logits = torch.empty_like(exp_sums)
        ops.paged_attention_v2(out, exp_sums, max_logits, partial_out, query, key_cache, value_cache, *tail)
        return out

    @staticmethod
    def forward_prefix(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor,
                       value_cache: torch.Tensor, block_tables: torch.Tensor, query_start_loc: torch.Tensor,
                       seq_lens_tensor: torch.Tensor, context_lens: torch.Tensor, max_query_len: int,
                       alibi_slopes: Optional[torch.Tensor], sliding_window: Optional[int],
                       kv_cache_dtype: str = "auto", kv_scale: float = 1.0) -> torch.Tensor:
        """Prompt tokens against their cached prefix + themselves (paged_attn.py:185-216): one launch of the
        sm_100a prefix-prefill kernel (csrc/prefill_tc.cu / prefill.cu).  `query_start_loc` has batch + 1 entries;
        the kernel wants each sequence's first row, i.e. all but the last."""
        out = torch.empty_like(query)
        first_rows = query_start_loc[:-1]
        ops.context_attention_fwd(query, key, value, out, key_cache, value_cache, block_tables, first_rows,
                                  seq_lens_tensor, context_lens, max_query_len, alibi_slopes, sliding_window,
                                  kv_cache_dtype, kv_scale)
        return out

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None:
        ops.swap_blocks(src_kv_cache[0], dst_kv_cache[0], src_to_dst)
        ops.swap_blocks(src_kv_cache[1], dst_kv_cache[1], src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        key_caches = [kv_cache[0] for kv_cache in kv_caches]
        value_caches = [kv_cache[1] for kv_cache in kv_caches]
        ops.copy_blocks(key_caches, value_caches, src_to_dists)
