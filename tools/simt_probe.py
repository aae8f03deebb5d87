#!/usr/bin/env python3
"""GPU-box developer tool: run the general (SIMT) decode kernel on one shape a few times (for ncu captures).
Usage: python tools/simt_probe.py [block8|fp32|blocksparse]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402


def sweep():
    """Time the three SIMT cases over context split sizes (partition_size knob; 0 = the library's plan)."""
    dev = torch.device("cuda", 0)
    for kind in ("fp32", "block8", "blocksparse"):
        B = 8 if kind == "block8" else 16
        dtype = torch.float32 if kind == "fp32" else torch.bfloat16
        c = bench.make_decode_case(dev, [4096] * 128, 32, 8, 128, B, dtype=dtype)
        fn = bench.attention_call(ops, c, "v2", (16, 8, 64, 0) if kind == "blocksparse" else None)
        for ps in (0, 512, 1024, 2048):
            ops.set_tuning("partition_size", ps)
            ms = bench.time_device(fn, 10)
            print(f"{kind:12s} partition_size={ps:5d}  {ms:8.3f} ms  path={ops.last_attention_path()}/p{ops.last_attention_partitions()}",
                  flush=True)
        ops.set_tuning("partition_size", 0)
        del c
        torch.cuda.empty_cache()


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "sweep":
        return sweep()
    kind = sys.argv[1] if len(sys.argv) > 1 else "block8"
    dev = torch.device("cuda", 0)
    B = 8 if kind == "block8" else 16
    dtype = torch.float32 if kind == "fp32" else torch.bfloat16
    c = bench.make_decode_case(dev, [4096] * 128, 32, 8, 128, B, dtype=dtype)
    fn = bench.attention_call(ops, c, "v2", (16, 8, 64, 0) if kind == "blocksparse" else None)
    for _ in range(6):
        fn()
    torch.cuda.synchronize()
    print(kind, ops.last_attention_path(), ops.last_attention_partitions())


if __name__ == "__main__":
    main()
