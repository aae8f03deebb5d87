#!/usr/bin/env python3
"""GPU-box tool: staged checks of the tcgen05 prefill kernel against fp32 dense attention
(prefix only / new tokens only / both), printing max errors instead of asserting."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import helpers  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402


def run(name, dtype, D, q, c, B=16, Hq=4, qpk=2, window=0):
    case = helpers.make_prefill_case(Hq, qpk, D, dtype, "cuda:0", query_lens=q, ctx_lens=c, block_size=B, amplitude=1.0)
    out = torch.full_like(case["query"], float("nan"))
    ops.context_attention_fwd(case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"],
                              case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                              case["max_input_len"], None, window)
    torch.cuda.synchronize()
    ref = helpers.ref_context_attention(case, None, window)
    err = (out.float() - ref).abs()
    print(f"{name:40s} max|err| {err.max().item():.3e}  nan {int(torch.isnan(out).sum())}  "
          f"rows>1e-2: {int((err.amax(dim=(1, 2)) > 1e-2).sum())}/{out.shape[0]}", flush=True)


run("prefix only-ish (q=128 exact, ctx=256)", torch.bfloat16, 128, [128], [256])
run("prefix heavy (q=128, ctx=1000)", torch.bfloat16, 128, [128], [1000])
run("new only (q=128, ctx=0)", torch.bfloat16, 128, [128], [0])
run("new only (q=300, ctx=0)", torch.bfloat16, 128, [300], [0])
run("both ragged", torch.bfloat16, 128, [70, 200, 129], [130, 0, 31])
run("fp16 both", torch.float16, 128, [333, 100], [500, 77])
run("D=64 both", torch.bfloat16, 64, [333, 100], [500, 77])
run("D=96 both", torch.bfloat16, 96, [333, 100], [500, 77])
run("block 32 window 100", torch.bfloat16, 128, [333, 100], [500, 77], B=32, window=100)
