"""`HiPAttentionBackend` — the fork's own backend (`vllm/attention/backends/hip.py`), dense paths only.

The fork routes attention through `PagedAttention` (this repo's kernels) in these cases
(`hip.py:206-209, 306-319`): decode in the first `HIP_DENSE_LAYERS` layers (`decode_method == 'vllm'`), decode in ANY
layer while `max_decode_seq_len < 3 * HIP_K`, and prompts with a cached prefix in batches of more than one sequence
(`hip.py:283-296`).  Those paths — KV-cache write included (`hip.py:151-159`) — are what this class implements, with the
reference's constructor (`layer_index` is required), environment knobs and metadata class.

The remaining branches call `paged_hip_attention` / `hip_attention` of the external `hip` sparse-attention package,
which is neither in /root/reference nor pinned by it (parity unpinned, out of scope).  They raise
`HiPSparsePathUnavailable`; nothing is silently replaced by another algorithm.
"""
import os
from dataclasses import dataclass
from typing import Any, Dict, List, Literal, Optional, Tuple, Type

import torch

from vllm_legacy_b200.attention.backends.abstract import AttentionBackend, AttentionImpl
from vllm_legacy_b200.attention.backends.b200_paged import (B200PagedAttentionImpl,
                                                             B200PagedAttentionMetadata)
from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention


class HiPSparsePathUnavailable(NotImplementedError):
    pass


class HiPAttentionBackend(AttentionBackend):

    @staticmethod
    def get_name() -> str:
        return "hip-attn"

    @staticmethod
    def get_impl_cls() -> Type["HiPAttentionImpl"]:
        return HiPAttentionImpl

    @staticmethod
    def make_metadata(*args, **kwargs) -> "HiPAttentionMetadata":
        return HiPAttentionMetadata(*args, **kwargs)

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)  # hip.py:341-349

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)


@dataclass
class HiPAttentionMetadata(B200PagedAttentionMetadata):
    """hip.py:31-33: the XFormers metadata, unchanged."""


class HiPAttentionImpl(AttentionImpl[HiPAttentionMetadata]):

    def __init__(
        self,
        num_heads: int,
        head_size: int,
        scale: float,
        num_kv_heads: int,
        alibi_slopes: Optional[List[float]],
        sliding_window: Optional[int],
        kv_cache_dtype: str,
        blocksparse_params: Optional[Dict[str, Any]] = None,
        layer_index: Optional[int] = None,
    ) -> None:
        assert blocksparse_params is None, ValueError("HiP does not support block-sparse attention.")  # hip.py:74-75
        assert layer_index is not None, 'layer index should be not none'                              # hip.py:103
        self._dense = B200PagedAttentionImpl(num_heads, head_size, scale, num_kv_heads, alibi_slopes, sliding_window,
                                             kv_cache_dtype)
        self.layer_index = layer_index
        self.hip_k = int(os.environ.get('HIP_K', '1024'))                                              # hip.py:105-108
        self.hip_dense_layers = list(range(int(os.environ.get('HIP_DENSE_LAYERS', '3'))))

    def forward(
        self,
        query: torch.Tensor,
        key: torch.Tensor,
        value: torch.Tensor,
        kv_cache: Optional[torch.Tensor],
        attn_metadata: HiPAttentionMetadata,
        kv_scale: float = 1.0,
        rope_method: Literal['none', 'self_extend'] = 'none',
        rope_cos: Optional[torch.Tensor] = None,
        rope_sin: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        assert kv_scale == 1.0, "kv_scale is not supported in HiPAttention."                           # hip.py:143
        decode_method = 'vllm' if self.layer_index in self.hip_dense_layers else 'hip'                # hip.py:206-209
        if (meta := attn_metadata.decode_metadata) is not None:
            if not (decode_method == 'vllm' or meta.max_decode_seq_len < self.hip_k * 3):             # hip.py:306
                raise HiPSparsePathUnavailable(
                    f"layer {self.layer_index}, max_decode_seq_len {meta.max_decode_seq_len} >= 3 * HIP_K: the fork "
                    "decodes this case with the external `hip` package (paged_hip_attention), which is out of scope")
        if (meta := attn_metadata.prefill_metadata) is not None:
            has_prefix = kv_cache is not None and meta.block_tables is not None and meta.block_tables.numel() > 0
            if not has_prefix or len(meta.context_lens_tensor) == 1:                                   # hip.py:214-282
                raise HiPSparsePathUnavailable(
                    "the fork runs prompts without a cached prefix, and single-sequence prefix prompts, through the "
                    "external `hip` package (hip_attention / paged_hip_attention), which is out of scope")
        assert rope_method == 'none', "self-extend RoPE is applied inside the external `hip` kernels only"
        return self._dense.forward(query, key, value, kv_cache, attn_metadata, kv_scale)
