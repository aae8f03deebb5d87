#!/usr/bin/env python3
"""Developer tool (GPU box): per-CTA timeline of the persistent decode kernel (b200pa_set_trace) — do all resident
CTAs stream their equal block ranges at the same rate, or do some SMs lag?"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from vllm_legacy_b200 import _C  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402
import tune  # noqa: E402


def main():
    for name in sys.argv[1:] or ["c2", "c3tp8"]:
        v1, v2, nbytes, S = tune.build(name)
        for grid in (0,):
            ops.set_tuning("grid_ctas", grid)
            for _ in range(3):
                v2()
            n = ops.last_attention_partitions()
            buf = torch.zeros(n, 4, dtype=torch.int64, device="cuda")
            _C.set_trace(buf)
            v2()
            torch.cuda.synchronize()
            _C.set_trace(None)
            t = buf.cpu()
            t0 = int(t[:, 1].min())
            beg, end, dur = (t[:, 1] - t0).float() / 1e3, (t[:, 2] - t0).float() / 1e3, (t[:, 2] - t[:, 1]).float() / 1e3
            q = lambda x: [round(float(v), 1) for v in torch.quantile(x, torch.tensor([0., .1, .5, .9, 1.]))]  # noqa: E731
            print(f"{name} grid={n}: total {float(end.max()):.1f} us = {nbytes / float(end.max()) / 1e3:.0f} GB/s | "
                  f"begin us {q(beg)} | end us {q(end)} | duration us {q(dur)} | blocks/CTA {int(t[:, 3].min())}..{int(t[:, 3].max())}")
            if grid == 0:
                # rate per SM: blocks / duration, grouped by SM id
                sm = t[:, 0]
                rate = t[:, 3].float() / dur
                per_sm = {}
                for s, r in zip(sm.tolist(), rate.tolist()):
                    per_sm.setdefault(s, []).append(r)
                rs = torch.tensor([sum(v) for v in per_sm.values()])
                print(f"   SMs used {len(per_sm)}; blocks/us per SM quantiles {q(rs)}")
                slow = sorted(per_sm.items(), key=lambda kv: sum(kv[1]))[:10]
                print("   slowest SMs (smid: blocks/us):", [(k, round(sum(v), 2)) for k, v in slow])
                fast = sorted(per_sm.items(), key=lambda kv: -sum(kv[1]))[:10]
                print("   fastest SMs (smid: blocks/us):", [(k, round(sum(v), 2)) for k, v in fast])
        ops.set_tuning("grid_ctas", 0)
        del v1, v2
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
