#!/usr/bin/env python3
"""GPU-box tool: time context_attention_fwd (ours, csrc/prefill.cu) and — when staged under oracle/_ref/ — the
reference's Triton kernels on the same inputs.  Reports algorithmic TFLOP/s:
flops = 4 * D * Hq * (visible (query, key) pairs): per sequence q_len * ctx_len + q_len * (q_len + 1) / 2 (QK^T + PV,
causal part counted once); with a sliding window W only the min(W, ctx_len + m + 1) keys query row m can see count.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import helpers  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402

CASES = {
    # name: (Hq, q_per_kv, D, block, batch x (q_len, ctx_len), dtype)
    "llama8b_chunk2k_ctx4k": (32, 4, 128, 16, [(2048, 4096)] * 8, torch.bfloat16),
    "llama8b_chunk512_ctx8k": (32, 4, 128, 16, [(512, 8192)] * 16, torch.bfloat16),
    "llama8b_prefix1k_new1k": (32, 4, 128, 16, [(1024, 1024)] * 16, torch.bfloat16),
    "llama70b_tp8_chunk2k_ctx4k": (8, 8, 128, 16, [(2048, 4096)] * 8, torch.bfloat16),
    "llama7b_mha_chunk2k_ctx4k": (32, 1, 128, 16, [(2048, 4096)] * 8, torch.bfloat16),
    "ref_test_shape_fp16": (64, 8, 128, 32, None, torch.float16),
    "window256_chunk2k_ctx4k": (32, 4, 128, 16, [(2048, 4096)] * 8, torch.bfloat16),
    "llama8b_chunk2k_ctx4k_fp8kv": (32, 4, 128, 16, [(2048, 4096)] * 8, torch.bfloat16),
}


def time_fn(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def visible_pairs(q_len: int, ctx_len: int, window: int) -> int:
    """(query row, key) pairs that are not masked: row m sees keys 0 .. ctx_len + m, the last `window` of them if a
    sliding window is set."""
    if window <= 0:
        return q_len * ctx_len + q_len * (q_len + 1) // 2
    # rows whose visible range is still shorter than the window: ctx_len + m + 1 < window
    short = max(0, min(q_len, window - 1 - ctx_len))
    return short * ctx_len + short * (short + 1) // 2 + (q_len - short) * window


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default=",".join(CASES))
    args = ap.parse_args()
    ref_mod = helpers.load_ref_prefix_prefill()
    for name in args.cases.split(","):
        Hq, qpk, D, B, lens, dtype = CASES[name]
        kw = dict(amplitude=1.0)
        if lens is not None:
            kw.update(query_lens=[a for a, _ in lens], ctx_lens=[b for _, b in lens])
        case = helpers.make_prefill_case(Hq, qpk, D, dtype, "cuda:0", block_size=B, **kw)
        window = 256 if name.startswith("window") else 0
        flops = 4.0 * D * Hq * sum(visible_pairs(q, c, window) for q, c in zip(case["query_lens"], case["ctx_lens"]))
        out = torch.empty_like(case["query"])
        a = (case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"], case["block_table"],
             case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"], None, window)
        if name.endswith("_fp8kv"):   # same shape over an fp8 e4m3 cache (dequantised while staged); no Triton counterpart
            from oracle import oracle as _oracle
            q8 = helpers.quantize_prefill_case(case, _oracle, "fp8", 0.5)
            a8 = a[:4] + (q8["k_cache"], q8["v_cache"]) + a[6:] + ("fp8", 0.5)
            ms = time_fn(lambda: ops.context_attention_fwd(*a8))
            print(f"{name:32s} ours {ms:8.3f} ms {flops / ms / 1e9:8.1f} TFLOP/s  (fp8 e4m3 KV prefix, kv_scale 0.5)", flush=True)
            continue
        ms = time_fn(lambda: ops.context_attention_fwd(*a))
        ref = helpers.ref_context_attention(case, None, window) if flops < 3e12 else None
        err = (out.float() - ref).abs().max().item() if ref is not None else float("nan")
        line = f"{name:32s} ours {ms:8.3f} ms {flops / ms / 1e9:8.1f} TFLOP/s  max|err| vs fp32 {err:.2e}"
        if ref_mod is not None:
            out2 = torch.empty_like(case["query"])
            a2 = a[:3] + (out2,) + a[4:]
            try:
                ms2 = time_fn(lambda: ref_mod.context_attention_fwd(*a2))
                err2 = (out2.float() - out.float()).abs().max().item()
                line += f" | reference Triton {ms2:8.3f} ms {flops / ms2 / 1e9:8.1f} TFLOP/s ({ms2 / ms:4.1f}x)  max|ours-ref| {err2:.2e}"
            except Exception as e:  # noqa: BLE001
                line += f" | reference Triton failed: {type(e).__name__}: {str(e)[:80]}"
        print(line, flush=True)


if __name__ == "__main__":
    main()
