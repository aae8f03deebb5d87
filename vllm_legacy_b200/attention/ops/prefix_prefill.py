"""`context_attention_fwd` — prefill attention of new tokens over (cached prefix in the paged KV cache)
+ (the new tokens, causal).  Same name, argument order and defaults as the reference's Triton wrapper
(vllm/attention/ops/prefix_prefill.py:684-696); the work is done by the hand-written sm_100a kernel in
csrc/prefill.cu behind the C-ABI entry `b200pa_context_attention_fwd`.  No Triton, no fallback.
"""
from typing import Optional

import torch

from vllm_legacy_b200 import _custom_ops as ops


@torch.inference_mode()
def context_attention_fwd(q: torch.Tensor,
                          k: torch.Tensor,
                          v: torch.Tensor,
                          o: torch.Tensor,
                          k_cache: torch.Tensor,
                          v_cache: torch.Tensor,
                          b_loc: torch.Tensor,
                          b_start_loc: torch.Tensor,
                          b_seq_len: torch.Tensor,
                          b_ctx_len: torch.Tensor,
                          max_input_len: int,
                          alibi_slopes: Optional[torch.Tensor] = None,
                          sliding_window: Optional[int] = None,
                          kv_cache_dtype: str = "auto",
                          kv_scale: float = 1.0) -> None:
    ops.context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len,
                              max_input_len, alibi_slopes, sliding_window, kv_cache_dtype, kv_scale)
