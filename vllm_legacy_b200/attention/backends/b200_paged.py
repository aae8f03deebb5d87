"""`B200PagedAttentionBackend` — a decode backend with the shape of the reference's backends that
route decode through `PagedAttention` (`vllm/attention/backends/xformers.py:19-58, 196-335`;
the fork's `hip.py:166-174, 306-319` dense path; `blocksparse_attn.py:357-390`).

It writes the step's K/V into the paged cache (`reshape_and_cache`), runs decode attention
(`paged_attention_v1/v2`) and prefill attention — with or without a cached prefix — on the B200 kernels
(`context_attention_fwd`, csrc/prefill.cu; the reference uses xformers for prompts without a prefix and a
Triton kernel for prompts with one, xformers.py:291-318).
"""
from dataclasses import dataclass
from typing import Any, Dict, List, Literal, Optional, Tuple, Type

import torch

from vllm_legacy_b200.attention.backends.abstract import (AttentionBackend, AttentionImpl,
                                                           AttentionMetadata)
from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention, PagedAttentionMetadata


class B200PagedAttentionBackend(AttentionBackend):

    @staticmethod
    def get_name() -> str:
        return "b200-paged"

    @staticmethod
    def get_impl_cls() -> Type["B200PagedAttentionImpl"]:
        return B200PagedAttentionImpl

    @staticmethod
    def make_metadata(*args, **kwargs) -> "B200PagedAttentionMetadata":
        return B200PagedAttentionMetadata(*args, **kwargs)

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None:
        PagedAttention.swap_blocks(src_kv_cache, dst_kv_cache, src_to_dst)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        PagedAttention.copy_blocks(kv_caches, src_to_dists)


@dataclass
class B200PagedAttentionMetadata(AttentionMetadata, PagedAttentionMetadata):
    """Field-compatible with XFormersMetadata (xformers.py:60-110) for the decode path."""
    seq_lens: Optional[List[int]] = None
    max_query_len: Optional[int] = None
    max_prefill_seq_len: int = 0
    query_start_loc: Optional[torch.Tensor] = None
    seq_start_loc: Optional[torch.Tensor] = None
    context_lens_tensor: Optional[torch.Tensor] = None
    use_cuda_graph: bool = False

    _cached_prefill_metadata: Optional["B200PagedAttentionMetadata"] = None
    _cached_decode_metadata: Optional["B200PagedAttentionMetadata"] = None

    def _view(self, prefill: bool) -> "B200PagedAttentionMetadata":
        """The prefill rows come first, then the decode rows (xformers.py:120-177): slice every per-sequence and
        per-token field accordingly."""
        npf, npt = self.num_prefills, self.num_prefill_tokens
        seqs = slice(0, npf) if prefill else slice(npf, None)
        toks = slice(0, npt) if prefill else slice(npt, None)
        cut = lambda t, sl: None if t is None else t[sl]  # noqa: E731
        return B200PagedAttentionMetadata(
            num_prefills=npf if prefill else 0,
            num_prefill_tokens=npt if prefill else 0,
            num_decode_tokens=0 if prefill else self.num_decode_tokens,
            slot_mapping=cut(self.slot_mapping, toks),
            seq_lens=cut(self.seq_lens, seqs) if prefill else None,
            seq_lens_tensor=cut(self.seq_lens_tensor, seqs),
            max_query_len=self.max_query_len if prefill else None,
            max_prefill_seq_len=self.max_prefill_seq_len if prefill else 0,
            max_decode_seq_len=0 if prefill else self.max_decode_seq_len,
            query_start_loc=cut(self.query_start_loc, slice(0, npf + 1)) if prefill else None,
            seq_start_loc=None,
            context_lens_tensor=cut(self.context_lens_tensor, seqs) if prefill else None,
            block_tables=cut(self.block_tables, seqs),
            use_cuda_graph=False if prefill else self.use_cuda_graph,
        )

    @property
    def prefill_metadata(self) -> Optional["B200PagedAttentionMetadata"]:
        if self.num_prefills == 0:
            return None
        if self._cached_prefill_metadata is None:
            self._cached_prefill_metadata = self._view(prefill=True)
        return self._cached_prefill_metadata

    @property
    def decode_metadata(self) -> Optional["B200PagedAttentionMetadata"]:
        if self.num_decode_tokens == 0:
            return None
        if self.num_prefills == 0:
            return self
        if self._cached_decode_metadata is None:
            self._cached_decode_metadata = self._view(prefill=False)
        return self._cached_decode_metadata


class B200PagedAttentionImpl(AttentionImpl[B200PagedAttentionMetadata]):

    def __init__(
        self,
        num_heads: int,
        head_size: int,
        scale: float,
        num_kv_heads: Optional[int] = None,
        alibi_slopes: Optional[List[float]] = None,
        sliding_window: Optional[int] = None,
        kv_cache_dtype: str = "auto",
        blocksparse_params: Optional[Dict[str, Any]] = None,
    ) -> None:
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.alibi_slopes = (torch.tensor(alibi_slopes, dtype=torch.float32)
                             if alibi_slopes is not None else None)
        self.sliding_window = sliding_window
        self.kv_cache_dtype = kv_cache_dtype
        self.blocksparse_params = blocksparse_params or {}
        assert self.num_heads % self.num_kv_heads == 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        supported = PagedAttention.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by PagedAttention. "
                             f"Supported head sizes are: {supported}.")

    def forward(
        self,
        query: torch.Tensor,
        key: torch.Tensor,
        value: torch.Tensor,
        kv_cache: Optional[torch.Tensor],
        attn_metadata: B200PagedAttentionMetadata,
        kv_scale: float = 1.0,
        rope_method: Literal['none', 'self_extend'] = 'none',
        rope_cos: Optional[torch.Tensor] = None,
        rope_sin: Optional[torch.Tensor] = None,
        position_ids: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        """query [T, Hq*D], key/value [T, Hkv*D], kv_cache [2, nb, B*Hkv*D] -> [T, Hq*D]."""
        assert rope_method == 'none', "self-extend RoPE belongs to the HiP backend, not this path"
        query = query.view(-1, self.num_heads, self.head_size)
        key = key.view(-1, self.num_kv_heads, self.head_size)
        value = value.view(-1, self.num_kv_heads, self.head_size)

        key_cache = value_cache = None
        if kv_cache is not None:
            key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads,
                                                                   self.head_size)
            PagedAttention.write_to_paged_cache(key, value, key_cache, value_cache,
                                                attn_metadata.slot_mapping, self.kv_cache_dtype,
                                                kv_scale)

        num_prefill_tokens = attn_metadata.num_prefill_tokens
        num_decode_tokens = attn_metadata.num_decode_tokens
        assert key.shape[0] == num_prefill_tokens + num_decode_tokens
        output = torch.empty_like(query)
        decode_query = query[num_prefill_tokens:]

        if self.alibi_slopes is not None and self.alibi_slopes.device != query.device:
            self.alibi_slopes = self.alibi_slopes.to(query.device)

        if prefill_meta := attn_metadata.prefill_metadata:
            output[:num_prefill_tokens] = self._forward_prefill(
                query[:num_prefill_tokens], key[:num_prefill_tokens], value[:num_prefill_tokens],
                key_cache, value_cache, prefill_meta, kv_scale)

        if decode_meta := attn_metadata.decode_metadata:
            assert kv_cache is not None, "decode needs the paged KV cache"
            bs = self.blocksparse_params
            output[num_prefill_tokens:] = PagedAttention.forward_decode(
                decode_query, key_cache, value_cache, decode_meta.block_tables,
                decode_meta.seq_lens_tensor, decode_meta.max_decode_seq_len, self.kv_cache_dtype,
                self.num_kv_heads, self.scale, self.alibi_slopes, kv_scale,
                tp_rank=bs.get("tp_rank", 0),
                blocksparse_local_blocks=bs.get("local_blocks", 0),
                blocksparse_vert_stride=bs.get("vert_stride", 0),
                blocksparse_block_size=bs.get("block_size", 64),
                blocksparse_head_sliding_step=bs.get("head_sliding_step", 0))
        return output.view(-1, self.num_heads * self.head_size)

    def _forward_prefill(self, query, key, value, key_cache, value_cache, meta, kv_scale=1.0) -> torch.Tensor:
        """Prompt run (xformers.py:291-318).  With a cached prefix the reference calls
        PagedAttention.forward_prefix; without one (no cache yet, or empty block tables) it runs plain causal
        attention through xformers — here that is the same kernel with every context length 0."""
        no_prefix = key_cache is None or meta.block_tables is None or meta.block_tables.numel() == 0
        kv_cache_dtype = self.kv_cache_dtype
        if no_prefix:
            kv_cache_dtype = "auto"  # the dummy cache below has the query dtype
            # plain causal attention over the new tokens only: the (possibly fp8) cache is not read (xformers.py:291-300)
            dev, nseq = query.device, meta.num_prefills
            key_cache = torch.empty(1, self.num_kv_heads, self.head_size // 8, 8, 8, dtype=query.dtype, device=dev)
            value_cache = torch.empty(1, self.num_kv_heads, self.head_size, 8, dtype=query.dtype, device=dev)
            block_tables = torch.zeros(nseq, 1, dtype=torch.int32, device=dev)
            context_lens = torch.zeros(nseq, dtype=torch.int32, device=dev)
            seq_lens = meta.query_start_loc[1:nseq + 1] - meta.query_start_loc[:nseq]
        else:
            block_tables, context_lens, seq_lens = meta.block_tables, meta.context_lens_tensor, meta.seq_lens_tensor
        return PagedAttention.forward_prefix(query, key, value, key_cache, value_cache, block_tables,
                                             meta.query_start_loc, seq_lens, context_lens, meta.max_query_len,
                                             self.alibi_slopes, self.sliding_window, kv_cache_dtype, kv_scale)
