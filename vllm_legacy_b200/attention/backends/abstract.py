"""Attention backend interface of this package.

Same three contracts, names and call signatures as the reference's `vllm/attention/backends/abstract.py:9-133`
(including the fork's extra `rope_method` / `rope_cos` / `rope_sin` / `position_ids` keywords on
`AttentionImpl.forward`, :119-133), so a backend written against the reference also satisfies these ABCs.
Written for this package: the interface is what is shared, not the text.
"""
import abc
import dataclasses
from typing import Any, Dict, Generic, List, Literal, Optional, Set, Tuple, Type, TypeVar

import torch

KVCacheShape = Tuple[int, ...]


def _required(fn):
    """Abstract static hook: subclasses must override it; calling the base raises."""
    return staticmethod(abc.abstractmethod(fn))


@dataclasses.dataclass
class AttentionMetadata:
    """Per-step metadata of a batch laid out as [prefill tokens | decode tokens]."""
    num_prefills: int          # prompt sequences in the batch
    num_prefill_tokens: int    # their tokens (rows 0 .. num_prefill_tokens - 1)
    num_decode_tokens: int     # one row per decoding sequence, after the prefill rows
    slot_mapping: torch.Tensor  # [num_tokens] paged-cache slot of every row; negative = padding row

    @property
    @abc.abstractmethod
    def prefill_metadata(self) -> Optional["AttentionMetadata"]:
        """View restricted to the prefill rows, or None when the batch has none."""

    @property
    @abc.abstractmethod
    def decode_metadata(self) -> Optional["AttentionMetadata"]:
        """View restricted to the decode rows, or None when the batch has none."""

    def asdict_zerocopy(self, skip_fields: Optional[Set[str]] = None) -> Dict[str, Any]:
        """Field name -> the very object stored (no deep copy as dataclasses.asdict would make)."""
        skip = skip_fields or set()
        return {f.name: getattr(self, f.name) for f in dataclasses.fields(self) if f.name not in skip}


T = TypeVar("T", bound=AttentionMetadata)


class AttentionImpl(abc.ABC, Generic[T]):
    """One attention layer's kernel-facing object: built once per layer, called every step."""

    @abc.abstractmethod
    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", blocksparse_params: Optional[Dict[str, Any]] = None) -> None:
        raise NotImplementedError

    @abc.abstractmethod
    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                attn_metadata: T, kv_scale: float = 1.0,
                rope_method: Literal['none', 'self_extend'] = 'none', rope_cos: Optional[torch.Tensor] = None,
                rope_sin: Optional[torch.Tensor] = None,
                position_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError


class AttentionBackend(abc.ABC):
    """Factory + cache-layout owner a model runner talks to; every hook is a static method."""

    @_required
    def get_name() -> str:
        raise NotImplementedError

    @_required
    def get_impl_cls() -> Type[AttentionImpl]:
        raise NotImplementedError

    @_required
    def make_metadata(*args, **kwargs) -> AttentionMetadata:
        raise NotImplementedError

    @_required
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int) -> KVCacheShape:
        raise NotImplementedError

    @_required
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor, src_to_dst: torch.Tensor) -> None:
        raise NotImplementedError

    @_required
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        raise NotImplementedError
