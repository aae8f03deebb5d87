"""Attention-backend selector for the paged-attention path (`vllm/attention/selector.py:15-167`).

Same entry points and the same `VLLM_ATTENTION_BACKEND` override as the reference.  Backends whose decode runs on
this repo's kernels are served: `XFORMERS` and `TORCH_SDPA` map to `B200PagedAttentionBackend` (both decode through
`PagedAttention.forward_decode` in the reference, `xformers.py:269-335`, `torch_sdpa.py:162-216`), `HIP_ATTN` to the
fork's `HiPAttentionBackend` (dense paths), `FLASH_ATTN` / `FLASHINFER` to the flash-layout backend
(`[2, nb, B, H, D]` cache, `flash_attn.py:281-367`).  `ROCM_FLASH` is refused: this library is sm_100a only.
"""
import enum
import os
from functools import lru_cache
from typing import Optional, Type

import torch

from vllm_legacy_b200.attention.backends.abstract import AttentionBackend


class _Backend(enum.Enum):  # selector.py:15-21
    FLASH_ATTN = enum.auto()
    XFORMERS = enum.auto()
    ROCM_FLASH = enum.auto()
    TORCH_SDPA = enum.auto()
    FLASHINFER = enum.auto()
    HIP_ATTN = enum.auto()


def which_attn_to_use(num_heads: int, head_size: int, num_kv_heads: int, sliding_window: Optional[int],
                      dtype: torch.dtype, kv_cache_dtype: Optional[str], block_size: int) -> _Backend:
    """selector.py:80-167 for an NVIDIA sm_100a device: FLASH_ATTN unless the environment says otherwise or the
    configuration is one the flash layout does not serve (the reference's own exclusions, :122-147)."""
    selected = _Backend.FLASH_ATTN
    by_env = os.getenv("VLLM_ATTENTION_BACKEND", None)
    if by_env is not None:
        if by_env not in _Backend.__members__:
            raise ValueError(f"Invalid attention backend '{by_env}'. "
                             f"Available backends: {', '.join(_Backend.__members__)} (case-sensitive).")
        selected = _Backend[by_env]
    if selected == _Backend.FLASH_ATTN:
        if (dtype not in (torch.float16, torch.bfloat16)
                or (kv_cache_dtype is not None and kv_cache_dtype.startswith("fp8"))
                or block_size % 16 != 0 or sliding_window is not None
                or head_size not in (32, 64, 96, 128, 160, 192, 224, 256)):  # flash_attn.py:24-25
            selected = _Backend.XFORMERS
    return selected


@lru_cache(maxsize=None)
def get_attn_backend(num_heads: int, head_size: int, num_kv_heads: int, sliding_window: Optional[int],
                     dtype: torch.dtype, kv_cache_dtype: Optional[str], block_size: int,
                     is_blocksparse: bool = False) -> Type[AttentionBackend]:
    """selector.py:24-77.  Block-sparse models keep the paged layout and decode through the same ops
    (`blocksparse_attn.py:357-390`), so they get the paged backend with their `blocksparse_params`."""
    from vllm_legacy_b200.attention.backends.b200_paged import B200PagedAttentionBackend
    if is_blocksparse:
        return B200PagedAttentionBackend
    backend = which_attn_to_use(num_heads, head_size, num_kv_heads, sliding_window, dtype, kv_cache_dtype, block_size)
    if backend in (_Backend.XFORMERS, _Backend.TORCH_SDPA):
        return B200PagedAttentionBackend
    if backend == _Backend.HIP_ATTN:
        from vllm_legacy_b200.attention.backends.hip import HiPAttentionBackend
        return HiPAttentionBackend
    if backend in (_Backend.FLASH_ATTN, _Backend.FLASHINFER):
        from vllm_legacy_b200.attention.backends.b200_flash import B200FlashAttentionBackend
        return B200FlashAttentionBackend
    raise ValueError(f"{backend.name} is not available: this library targets NVIDIA sm_100a only.")
