#!/usr/bin/env python3
"""GPU-box developer tool: per-phase clock64 timestamps of one CTA of the warp-specialised tcgen05 prefill kernel.

Needs a library built with B200PA_EXTRA_NVCC_FLAGS=-DB200PA_PREFILL_TRACE (python -m vllm_legacy_b200.build); the
product build compiles the trace points out and does not export b200pa_debug_prefill_trace.
Slots per iteration: softmax warp 0: 0 before S wait, 1 S ready, 2 max/rescale done, 3 P stored + arrived;
MMA warp: 4 loop top, 5 K(it+1) landed, 6 QK(it+1) issued, 7 K(it+2) copy issued, 8 V(it) ready, 9/10 PV issued.
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import helpers  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402


def main():
    Hq, qpk, D, B = 32, 4, 128, 16
    lens = [(2048, 4096)] * 8
    case = helpers.make_prefill_case(Hq, qpk, D, torch.bfloat16, "cuda:0", block_size=B, amplitude=1.0,
                                     query_lens=[a for a, _ in lens], ctx_lens=[b for _, b in lens])
    out = torch.empty_like(case["query"])
    a = (case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"], case["block_table"],
         case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"], None, 0)
    for _ in range(3):
        ops.context_attention_fwd(*a)
    torch.cuda.synchronize()
    lib = ctypes.CDLL(os.path.join(ROOT, "vllm_legacy_b200", "libb200pa.so"))
    buf = np.zeros(128 * 16, dtype=np.int64)
    rc = lib.b200pa_debug_prefill_trace(buf.ctypes.data_as(ctypes.c_void_p), buf.size)
    assert rc == 0, rc
    t = buf.reshape(128, 16)
    n = 96
    t0 = t[0, 4]
    print("NQ env:", os.environ.get("B200PA_PREFILL_NQ", "default"))
    print("iteration period (MMA loop top to top), it 10..90: mean %.0f clk" % np.diff(t[10:90, 4]).mean())
    names = ["sm:wait_S", "sm:S_ready", "sm:max_done", "sm:P_arrived", "mma:top", "mma:K_landed", "mma:QK_issued",
             "mma:K_copy_issued", "mma:V_ready", "mma:PV0_issued", "mma:PV1_issued"]
    print("per-iteration phase offsets relative to the MMA loop top (mean over it 10..90):")
    for s, nm in enumerate(names):
        if t[20, s] == 0:
            continue
        print("  %-20s %8.0f" % (nm, (t[10:90, s] - t[10:90, 4]).mean()))
    print("softmax: wait_S->S_ready %.0f | S_ready->max_done %.0f | max_done->P_arrived %.0f | P_arrived->next wait_S %.0f"
          % ((t[10:90, 1] - t[10:90, 0]).mean(), (t[10:90, 2] - t[10:90, 1]).mean(), (t[10:90, 3] - t[10:90, 2]).mean(),
             (t[11:91, 0] - t[10:90, 3]).mean()))
    print("raw (it 40..43), relative to kernel start:")
    for it in range(40, 44):
        print("  it %d: " % it + " ".join("%d:%d" % (s, t[it, s] - t0) for s in range(11) if t[it, s]))


if __name__ == "__main__":
    main()
