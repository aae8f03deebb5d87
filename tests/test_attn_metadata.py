"""Host-side metadata builders are bit-identical to the reference's per-token / per-sequence loops
(restated in oracle/metadata_ref.py from vllm/worker/model_runner.py:464-502, 543-556 and vllm/utils.py:505-521)."""
import random

import numpy as np
import pytest
import torch

from oracle import metadata_ref as ref
from vllm_legacy_b200.worker import attn_metadata as am


def _batch(rng, n, block_size, max_len, prompt_frac=0.5, empty_frac=0.0):
    bts, ctxs, lens = [], [], []
    for _ in range(n):
        seq_len = rng.randint(1, max_len)
        ctx = rng.randint(0, seq_len - 1) if rng.random() < prompt_frac else seq_len - 1
        nblk = (seq_len + block_size - 1) // block_size
        bt = [rng.randrange(10 ** 6) for _ in range(nblk + rng.randint(0, 2))]
        if rng.random() < empty_frac:
            bt = rng.choice([None, []])
        bts.append(bt)
        ctxs.append(ctx)
        lens.append(seq_len)
    return bts, ctxs, lens


@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("sliding_window", [None, 7, 4096])
@pytest.mark.parametrize("seed", range(4))
def test_slot_mapping_equals_reference_loops(block_size, sliding_window, seed):
    rng = random.Random(seed * 31 + block_size)
    bts, ctxs, lens = _batch(rng, rng.choice([1, 5, 64]), block_size, rng.choice([1, 50, 3000]),
                             empty_frac=0.2 if seed % 2 else 0.0)
    want = ref.slot_mapping_loops(bts, ctxs, lens, block_size, sliding_window)
    got = am.build_slot_mapping(bts, ctxs, lens, block_size, sliding_window)
    assert got.dtype == torch.int64 and got.tolist() == want


def test_slot_mapping_examples_from_the_reference_comment():
    # model_runner.py:476-481: prompt len 10, window 8, block size 4 -> [-1, -1, 2, 3, 4, 5, 6, 7, 0, 1]
    assert am.build_slot_mapping([[0, 1, 0]], [0], [10], 4, 8).tolist() == [-1, -1, 2, 3, 4, 5, 6, 7, 0, 1]
    assert am.build_slot_mapping([], [], [], 16).numel() == 0
    assert am.build_slot_mapping([None], [0], [3], 16).tolist() == [-1, -1, -1]


@pytest.mark.parametrize("seed", range(4))
def test_block_tables_equal_make_tensor_with_pad(seed):
    rng = random.Random(seed)
    bts, _, _ = _batch(rng, rng.choice([1, 7, 256]), 16, 5000, empty_frac=0.1)
    width = max(len(b) if b else 0 for b in bts)
    for pad, max_len in ((0, width), (-1, width + 3)):
        want = ref.make_tensor_with_pad_loops(bts, max_len, pad)
        got = am.pad_block_tables(bts, max_len, pad)
        assert got.dtype == torch.int32 and np.array_equal(got.numpy(), want)


def test_decode_metadata_with_cuda_graph_padding():
    """model_runner.py:517-556: pad rows get slot -1, seq_len 1 and a zero block table in the graph's table."""
    rng = random.Random(3)
    B = 16
    lens = [rng.randint(1, 900) for _ in range(5)]
    bts = [[rng.randrange(5000) for _ in range((n + B - 1) // B)] for n in lens]
    graph_tables = torch.zeros(8, 64, dtype=torch.int32)
    slots, seq_lens, bt, max_len = am.build_decode_metadata(bts, lens, B, graph_batch_size=8,
                                                            graph_block_tables=graph_tables)
    assert slots.tolist() == ref.slot_mapping_loops(bts, [n - 1 for n in lens], lens, B) + [-1] * 3
    assert seq_lens.tolist() == lens + [1, 1, 1] and max_len == max(lens)
    assert bt.data_ptr() == graph_tables.data_ptr() and bt.shape == (8, 64)
    want = ref.make_tensor_with_pad_loops(bts + [[]] * 3, 64, 0)
    assert np.array_equal(bt.numpy(), want)
    slots2, seq_lens2, bt2, _ = am.build_decode_metadata(bts, lens, B)
    assert slots2.tolist() == slots.tolist()[:5] and seq_lens2.tolist() == lens
    assert np.array_equal(bt2.numpy(), ref.make_tensor_with_pad_loops(bts, max(map(len, bts)), 0))


def test_block_table_cache_follows_the_scheduler():
    """Appends, copy-on-write of the last block, sliding-window rotation, swap-in with new blocks, finished and new
    sequences: after every step the cached rows equal make_tensor_with_pad of the lists, and the decode metadata
    equals the reference loops."""
    rng = random.Random(0)
    B, W = 16, 64
    cache = am.BlockTableCache(W, rows=4, strict=False)
    live = {}
    next_id = 0
    for step in range(300):
        for sid in list(live):                               # one decode step for every live sequence
            st = live[sid]
            st["len"] += 1
            if (st["len"] + B - 1) // B > len(st["bt"]):
                st["bt"].append(rng.randrange(10 ** 6))
            ev = rng.random()
            if ev < 0.03:                                    # copy-on-write of the last block
                st["bt"][-1] = rng.randrange(10 ** 6)
            elif ev < 0.05 and len(st["bt"]) > 2:            # sliding-window style rotation
                st["bt"] = st["bt"][1:] + st["bt"][:1]
            elif ev < 0.07:                                  # swapped out and in: all new physical blocks
                st["bt"] = [rng.randrange(10 ** 6) for _ in st["bt"]]
            elif ev < 0.09 or st["len"] >= B * W - 2:        # finished
                cache.free(sid)
                del live[sid]
        while len(live) < rng.randint(1, 9):                 # new arrivals (with a prompt already cached)
            n = rng.randint(1, 300)
            live[next_id] = {"len": n, "bt": [rng.randrange(10 ** 6) for _ in range((n + B - 1) // B)]}
            next_id += 1
        ids = list(live)
        rng.shuffle(ids)
        bts = [live[i]["bt"] for i in ids]
        lens = [live[i]["len"] for i in ids]
        slots, seq_lens, bt, max_len = cache.decode_step(ids, bts, lens, B)
        width = max(len(b) for b in bts)
        assert np.array_equal(bt.numpy(), ref.make_tensor_with_pad_loops(bts, width, 0))
        assert slots.tolist() == ref.slot_mapping_loops(bts, [n - 1 for n in lens], lens, B)
        assert seq_lens.tolist() == lens and max_len == max(lens)
