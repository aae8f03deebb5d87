"""TEST INFRASTRUCTURE — restatement of the reference's host-side metadata loops, line by line, for
tests/test_attn_metadata.py and bench.py's host-metadata timing.  Never imported by the product.

  slot mapping   vllm/worker/model_runner.py:464-502 (one Python iteration per token)
  block tables   vllm/utils.py:505-521 make_tensor_with_pad (one per sequence), model_runner.py:543-556
"""
from typing import List, Optional, Sequence

import numpy as np

_PAD_SLOT_ID = -1


def slot_mapping_loops(block_tables: Sequence[Optional[Sequence[int]]], context_lens: Sequence[int],
                       seq_lens: Sequence[int], block_size: int, sliding_window: Optional[int] = None) -> List[int]:
    slot_mapping: List[int] = []
    for block_table, context_len, seq_len in zip(block_tables, context_lens, seq_lens):
        query_len = seq_len - context_len
        if not block_table:                                   # model_runner.py:465-471
            slot_mapping.extend([_PAD_SLOT_ID] * seq_len)
            continue
        start_idx = 0
        if sliding_window is not None:                        # :483-493
            start_idx = max(0, query_len - sliding_window)
        for i in range(context_len, seq_len):                 # :494-502
            if i < start_idx:
                slot_mapping.append(_PAD_SLOT_ID)
                continue
            block_number = block_table[i // block_size]
            block_offset = i % block_size
            slot_mapping.append(block_number * block_size + block_offset)
    return slot_mapping


def make_tensor_with_pad_loops(x: Sequence[Optional[Sequence[int]]], max_len: int, pad: int) -> np.ndarray:
    padded_x = np.zeros([len(x), max_len], dtype=np.int32) + pad   # vllm/utils.py:517-521
    for ind, blocktb in enumerate(x):
        blocktb = blocktb or []
        assert len(blocktb) <= max_len
        padded_x[ind, :len(blocktb)] = blocktb
    return padded_x
