"""Host-side metadata of the paged-attention path, vectorised.

The reference builds the slot mapping with one Python iteration PER TOKEN and the block-table tensor with one
per sequence (`vllm/worker/model_runner.py:494-502, 543-556`, `vllm/utils.py:505-521`): at batch 256 that is the
largest host cost of a decode step once the kernels take 0.6 ms, and a 16k-token prefill chunk spends milliseconds
there.  These builders return the same arrays — bit-identical, see tests/test_attn_metadata.py — from a handful of
numpy operations over the whole batch, written into pinned host tensors ready for one non-blocking H2D copy each.
"""
import itertools
from typing import Dict, Hashable, List, Optional, Sequence, Tuple

import numpy as np
import torch

_PAD_SLOT_ID = -1  # vllm/worker/model_runner.py:34


def _pinned(shape, dtype) -> torch.Tensor:
    try:
        return torch.empty(shape, dtype=dtype, pin_memory=torch.cuda.is_available())
    except RuntimeError:  # no CUDA runtime on this host
        return torch.empty(shape, dtype=dtype)


def pad_block_tables(block_tables: Sequence[Optional[Sequence[int]]], max_len: Optional[int] = None,
                     pad: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`make_tensor_with_pad(block_tables, max_len, pad, torch.int, device)` (vllm/utils.py:505-521) without the
    per-row Python loop: int32 `[num_seqs, max_len]`, rows right-padded with `pad`; `None` / empty rows are all
    padding.  `out` (e.g. the CUDA-graph table, model_runner.py:543-549) is filled in place when given."""
    n = len(block_tables)
    lens = np.fromiter((len(bt) if bt else 0 for bt in block_tables), dtype=np.int64, count=n)
    width = int(lens.max(initial=0)) if max_len is None else int(max_len)
    if out is None:
        out = _pinned((n, width), torch.int32)
        arr = out.numpy()
        arr.fill(pad)
    else:
        arr = out.numpy()[:n]
        assert arr.shape[1] >= int(lens.max(initial=0))
    total = int(lens.sum())
    if total:
        flat = np.fromiter(itertools.chain.from_iterable(bt for bt in block_tables if bt), dtype=np.int32, count=total)
        rows = np.repeat(np.arange(n), lens)
        cols = np.arange(total) - np.repeat(np.cumsum(lens) - lens, lens)
        arr[rows, cols] = flat
    return out


def build_slot_mapping(block_tables: Sequence[Optional[Sequence[int]]], context_lens: Sequence[int],
                       seq_lens: Sequence[int], block_size: int, sliding_window: Optional[int] = None,
                       is_prompt: Optional[Sequence[bool]] = None) -> torch.Tensor:
    """The slot of every token this step writes to the KV cache (model_runner.py:464-502), int64 `[num_tokens]`:
    for sequence s and position i in [context_len, seq_len): `block_table[i // B] * B + i % B`; positions below
    `start_idx = max(0, query_len - sliding_window)` and sequences without a block table get `_PAD_SLOT_ID`."""
    n = len(seq_lens)
    ctx = np.asarray(context_lens, dtype=np.int64)
    end = np.asarray(seq_lens, dtype=np.int64)
    qlen = end - ctx
    total = int(qlen.sum())
    out = _pinned((total,), torch.int64)
    if total == 0:
        return out
    has_table = np.fromiter((bool(bt) for bt in block_tables), dtype=bool, count=n)
    # Empty block tables (memory profiling, embeddings) pad seq_len slots, not query_len (model_runner.py:465-471)
    counts = np.where(has_table, qlen, end)
    if not has_table.all():
        total = int(counts.sum())
        out = _pinned((total,), torch.int64)
    seq_id = np.repeat(np.arange(n), counts)
    first = np.cumsum(counts) - counts
    pos = np.arange(total) - np.repeat(first, counts) + np.repeat(np.where(has_table, ctx, 0), counts)
    # Only the table entries the new tokens fall into are converted from Python ints (~45 ns each): blocks
    # ctx // B .. (seq_len - 1) // B of every sequence — one entry per sequence in a decode step.
    b0 = np.where(has_table, ctx // block_size, 0)
    b1 = np.where(has_table, (end - 1) // block_size + 1, 0)
    need = b1 - b0
    flat = np.fromiter(itertools.chain.from_iterable(
        bt[lo:hi] for bt, lo, hi in zip(block_tables, b0.tolist(), b1.tolist()) if hi > lo),
        dtype=np.int64, count=int(need.sum()))
    if flat.size == 0:
        flat = np.zeros(1, dtype=np.int64)
    off = np.cumsum(need) - need
    idx = np.repeat(off - b0, counts) + pos // block_size
    blk = flat[np.clip(idx, 0, flat.size - 1)]
    slots = blk * block_size + pos % block_size
    masked = ~np.repeat(has_table, counts)
    if sliding_window is not None:
        start_idx = np.maximum(0, qlen - sliding_window)   # 0 for decode rows (query_len 1)
        masked |= pos < np.repeat(start_idx, counts)
    slots[masked] = _PAD_SLOT_ID
    out.numpy()[:] = slots
    return out


def build_decode_metadata(block_tables: Sequence[Sequence[int]], seq_lens: Sequence[int], block_size: int,
                          graph_batch_size: Optional[int] = None,
                          graph_block_tables: Optional[torch.Tensor] = None
                          ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, int]:
    """One decode step (model_runner.py:517-556 for the decode-only case): returns pinned
    (slot_mapping int64 [S], seq_lens int32 [S], block_tables int32 [S, max_blocks], max_decode_seq_len).
    With `graph_batch_size` the batch is padded for CUDA-graph replay exactly as the reference does: pad rows have
    slot `_PAD_SLOT_ID`, seq_len 1 and an all-zero block table taken from `graph_block_tables`."""
    n = len(seq_lens)
    lens = np.asarray(seq_lens, dtype=np.int64)
    slots = build_slot_mapping(block_tables, lens - 1, lens, block_size)
    S = n if graph_batch_size is None else int(graph_batch_size)
    assert S >= n
    slot_t = _pinned((S,), torch.int64)
    slot_t.numpy()[:n] = slots.numpy()
    slot_t.numpy()[n:] = _PAD_SLOT_ID
    len_t = _pinned((S,), torch.int32)
    len_t.numpy()[:n] = lens
    len_t.numpy()[n:] = 1
    if graph_block_tables is not None:
        bt = graph_block_tables[:S]
        pad_block_tables(block_tables, out=bt)
    else:
        bt = pad_block_tables(list(block_tables) + [[]] * (S - n))
    return slot_t, len_t, bt, int(lens.max(initial=0))


class BlockTableCache:
    """Block tables kept as ONE pinned int32 array across steps, updated incrementally.

    Converting the scheduler's Python lists to a tensor costs ~45 ns per entry whatever the method
    (`make_tensor_with_pad`, vllm/utils.py:505-521: 3-4 ms per decode step at batch 256 x 4096 tokens) although a
    decode step appends at most one block per sequence.  This cache keeps a row per live sequence and copies only
    the entries that are new, so a step costs one dict lookup and two comparisons per sequence plus one gather.

    A row is rewritten in full whenever its head or its last known entry changed or it shrank (copy-on-write of the
    last block, sliding-window rotation, swap-in with other physical blocks); `invalidate` / `free` let the caller
    say so explicitly (swap-out, preemption, finished sequence).  `strict=True` re-verifies every row against the
    list it was given (tests)."""

    def __init__(self, max_blocks_per_seq: int, rows: int = 256, strict: bool = False) -> None:
        self.width = int(max_blocks_per_seq)
        self.strict = strict
        self._arr = _pinned((rows, self.width), torch.int32)
        self._np = self._arr.numpy()
        self._np.fill(0)
        self._len = np.zeros(rows, dtype=np.int64)
        self._row_of: Dict[Hashable, int] = {}
        self._free: List[int] = list(range(rows - 1, -1, -1))

    def _grow(self) -> None:
        old = self._np
        rows = old.shape[0] * 2
        self._arr = _pinned((rows, self.width), torch.int32)
        self._np = self._arr.numpy()
        self._np[:old.shape[0]] = old
        self._np[old.shape[0]:] = 0
        self._len = np.concatenate([self._len, np.zeros(rows - old.shape[0], dtype=np.int64)])
        self._free.extend(range(rows - 1, old.shape[0] - 1, -1))

    def free(self, seq_id: Hashable) -> None:
        row = self._row_of.pop(seq_id, None)
        if row is not None:
            self._len[row] = 0
            self._free.append(row)

    def invalidate(self, seq_id: Hashable) -> None:
        row = self._row_of.get(seq_id)
        if row is not None:
            self._len[row] = 0

    def update(self, seq_id: Hashable, block_table: Optional[Sequence[int]]) -> int:
        """Bring the row of `seq_id` up to date with `block_table`; returns the row index."""
        row = self._row_of.get(seq_id)
        if row is None:
            if not self._free:
                self._grow()
            row = self._free.pop()
            self._row_of[seq_id] = row
            self._len[row] = 0
        bt = block_table or ()
        m, k = len(bt), int(self._len[row])
        assert m <= self.width, "block table longer than max_blocks_per_seq"
        r = self._np[row]
        if k and (m < k or r[0] != bt[0] or r[k - 1] != bt[k - 1]):
            k = 0  # rotated / copied-on-write / swapped: rewrite
        if m > k:
            r[k:m] = bt[k:m]
        self._len[row] = m
        if self.strict:
            assert r[:m].tolist() == list(bt), "stale block-table row (missing invalidate()?)"
        return row

    def gather(self, seq_ids: Sequence[Hashable], block_tables: Sequence[Optional[Sequence[int]]],
               out: Optional[torch.Tensor] = None, pad: int = 0) -> Tuple[torch.Tensor, np.ndarray]:
        """Rows of the given sequences in batch order: pinned int32 `[n, max_len]` (or `out[:n]`, e.g. the CUDA-graph
        table) padded with `pad` — what make_tensor_with_pad returns — and the row indices used."""
        rows = np.fromiter((self.update(s, bt) for s, bt in zip(seq_ids, block_tables)), dtype=np.int64,
                           count=len(seq_ids))
        lens = self._len[rows]
        width = int(lens.max(initial=0))
        if out is None:
            out = _pinned((len(rows), width), torch.int32)
            dst = out.numpy()
        else:
            dst = out.numpy()[:len(rows)]
            width = dst.shape[1]
        np.take(self._np[:, :width], rows, axis=0, out=dst)
        if pad != 0 or True:  # entries behind a row's length may hold blocks of an earlier, longer table
            dst[np.arange(width)[None, :] >= lens[:, None]] = pad
        return out, rows

    def decode_step(self, seq_ids: Sequence[Hashable], block_tables: Sequence[Sequence[int]], seq_lens: Sequence[int],
                    block_size: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, int]:
        """`build_decode_metadata` for a decode-only batch, from the cache: (slot_mapping, seq_lens, block_tables,
        max_decode_seq_len), all pinned."""
        bt, rows = self.gather(seq_ids, block_tables)
        lens = np.asarray(seq_lens, dtype=np.int64)
        last = lens - 1
        slot_t = _pinned((len(rows),), torch.int64)
        slot_t.numpy()[:] = self._np[rows, last // block_size].astype(np.int64) * block_size + last % block_size
        len_t = _pinned((len(rows),), torch.int32)
        len_t.numpy()[:] = lens
        return slot_t, len_t, bt, int(lens.max(initial=0))
