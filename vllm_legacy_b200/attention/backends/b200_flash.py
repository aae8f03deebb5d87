"""`B200FlashAttentionBackend` — the reference's FLASH_ATTN backend interface (`vllm/attention/backends/flash_attn.py`)
on this repo's kernels: KV cache `[2, num_blocks, block_size, num_kv_heads, head_size]`, written by
`reshape_and_cache_flash`, decoded by `paged_attention_flash` (the reference calls the external
`flash_attn_with_kvcache`, flash_attn.py:354-367).

Prompts without a cached prefix run through `context_attention_fwd` with every context length 0 (the reference calls
`flash_attn_varlen_func`, flash_attn.py:318-335).  Prefix-enabled prompts over THIS layout (flash_attn.py:336-352)
have no kernel here and raise: the paged-layout backend (`b200_paged.py`) serves prefix caching.
"""
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Tuple, Type

import torch

from vllm_legacy_b200 import _custom_ops as ops
from vllm_legacy_b200.attention.backends.abstract import AttentionBackend, AttentionImpl
from vllm_legacy_b200.attention.backends.b200_paged import B200PagedAttentionMetadata
from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention


class B200FlashAttentionBackend(AttentionBackend):

    @staticmethod
    def get_name() -> str:
        return "b200-flash"

    @staticmethod
    def get_impl_cls() -> Type["B200FlashAttentionImpl"]:
        return B200FlashAttentionImpl

    @staticmethod
    def make_metadata(*args, **kwargs) -> "B200FlashAttentionMetadata":
        return B200FlashAttentionMetadata(*args, **kwargs)

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [32, 64, 96, 128, 160, 192, 224, 256]  # flash_attn.py:24-25

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> Tuple[int, ...]:
        if block_size % 16 != 0:
            raise ValueError("Block size must be a multiple of 16.")  # flash_attn.py:42-43
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        ops.swap_blocks(src_kv_cache[0], dst_kv_cache[0], src_to_dst)  # flash_attn.py:46-58
        ops.swap_blocks(src_kv_cache[1], dst_kv_cache[1], src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        ops.copy_blocks([kv[0] for kv in kv_caches], [kv[1] for kv in kv_caches], src_to_dists)  # flash_attn.py:60-67


@dataclass
class B200FlashAttentionMetadata(B200PagedAttentionMetadata):
    """Field-compatible with FlashAttentionMetadata (flash_attn.py:70-190): the same fields as the xformers one."""


class B200FlashAttentionImpl(AttentionImpl[B200FlashAttentionMetadata]):

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", blocksparse_params: Optional[Dict[str, Any]] = None) -> None:
        assert blocksparse_params is None, ValueError("FlashAttention does not support block-sparse attention.")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = torch.tensor(alibi_slopes, dtype=torch.float32) if alibi_slopes is not None else None
        self.sliding_window = sliding_window
        self.kv_cache_dtype = kv_cache_dtype
        assert self.num_heads % self.num_kv_heads == 0
        if sliding_window is not None:
            raise ValueError("Sliding window is not supported in FlashAttention.")  # flash_attn.py:236-239
        supported = B200FlashAttentionBackend.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by FlashAttention. "
                             f"Supported head sizes are: {supported}.")

    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: Optional[torch.Tensor],
                attn_metadata: B200FlashAttentionMetadata, kv_scale: float = 1.0) -> torch.Tensor:
        """query [T, Hq*D], key/value [T, Hkv*D], kv_cache [2, nb, B, Hkv, D] -> [T, Hq*D] (flash_attn.py:249-370)."""
        assert kv_scale == 1.0, "kv_scale is not supported in FlashAttention."
        num_tokens, hidden_size = query.shape
        query = query.view(-1, self.num_heads, self.head_size)
        key = key.view(-1, self.num_kv_heads, self.head_size)
        value = value.view(-1, self.num_kv_heads, self.head_size)
        key_cache = value_cache = None
        if kv_cache is not None:
            key_cache, value_cache = kv_cache[0], kv_cache[1]
            ops.reshape_and_cache_flash(key, value, key_cache, value_cache, attn_metadata.slot_mapping.flatten(),
                                        self.kv_cache_dtype)
        num_prefill_tokens = attn_metadata.num_prefill_tokens
        assert key.shape[0] == num_prefill_tokens + attn_metadata.num_decode_tokens
        output = torch.empty_like(query)
        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)

        if prefill_meta := attn_metadata.prefill_metadata:
            if not (kv_cache is None or prefill_meta.block_tables is None or prefill_meta.block_tables.numel() == 0):
                raise NotImplementedError("prefix-enabled prompts over the flash cache layout (flash_attn.py:336-352) "
                                          "have no kernel here; the paged-layout backend serves prefix caching")
            q, k, v = query[:num_prefill_tokens], key[:num_prefill_tokens], value[:num_prefill_tokens]
            dev, nseq = query.device, prefill_meta.num_prefills
            dummy_k = torch.empty(1, self.num_kv_heads, self.head_size // 8, 8, 8, dtype=query.dtype, device=dev)
            dummy_v = torch.empty(1, self.num_kv_heads, self.head_size, 8, dtype=query.dtype, device=dev)
            qsl = prefill_meta.query_start_loc
            output[:num_prefill_tokens] = PagedAttention.forward_prefix(
                q, k, v, dummy_k, dummy_v, torch.zeros(nseq, 1, dtype=torch.int32, device=dev), qsl,
                qsl[1:nseq + 1] - qsl[:nseq], torch.zeros(nseq, dtype=torch.int32, device=dev),
                prefill_meta.max_query_len, self.alibi_slopes, None)

        if decode_meta := attn_metadata.decode_metadata:
            assert kv_cache is not None, "decode needs the KV cache"
            dq = query[num_prefill_tokens:]
            out = torch.empty_like(dq)
            ops.paged_attention_flash(out, dq, key_cache, value_cache, self.scale, decode_meta.block_tables,
                                      decode_meta.seq_lens_tensor, decode_meta.max_decode_seq_len, self.alibi_slopes)
            output[num_prefill_tokens:] = out
        return output.view(num_tokens, hidden_size)
