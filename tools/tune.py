#!/usr/bin/env python3
"""Developer tool (GPU box): time the attention ops on the BASELINE configs under different launch knobs
(b200pa_set_tuning keys; the library reads its environment only once, at load).  One line per (config, knob set).

  python tools/tune.py --configs c2,c4 --knobs "" "stages=2" "stages=4,evict_normal=1" "split_cost=256"
"""
import argparse
import math
import os
import random
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda"


def ragged_lens(n=64, lo=128, hi=131072, seed=0):
    rng = random.Random(seed)
    lens = [int(round(math.exp(rng.uniform(math.log(lo), math.log(hi))))) for _ in range(n)]
    lens[0], lens[1] = hi, lo
    return lens


CONFIGS = {
    # name: (lens, Hq, Hkv, D, B, kv)
    "c1": ([1024] * 16, 32, 32, 128, 16, "auto"),
    "c2": ([4096] * 256, 32, 8, 128, 16, "auto"),
    "c3tp2": ([8192] * 256, 32, 4, 128, 16, "auto"),
    "c3tp8": ([8192] * 256, 8, 1, 128, 16, "auto"),
    "c4": ([32768] * 128, 32, 8, 128, 16, "fp8"),
    "c5b16": (ragged_lens(), 32, 8, 128, 16, "auto"),
    "c5b32": (ragged_lens(), 32, 8, 128, 32, "auto"),
    "small": ([2048] * 8, 32, 8, 128, 16, "auto"),
    "one": ([32768] * 1, 32, 8, 128, 16, "auto"),
    "mid": ([8192] * 32, 32, 8, 128, 16, "auto"),          # 17-64 sequences: the reference's rule still picks v1
    "c3tp4": ([8192] * 256, 16, 2, 128, 16, "auto"),
    "short": ([200] * 1024, 32, 8, 128, 16, "auto"),       # many short sequences: item switches dominate
}
DEFAULTS = {"stages": 0, "evict_normal": -1, "force_simt": 0, "partition_size": 0, "max_partition": 4096,
            "split_cost": 512}


def build_flash(name):
    """The same shapes over the flash-attention cache layout [nb, B, H, D] (paged_attention_flash)."""
    lens, Hq, Hkv, D, B, kvd = CONFIGS[name[:-len("_flash")]]
    S, L = len(lens), max(lens)
    scale = D ** -0.5
    nblk = [(l + B - 1) // B for l in lens]
    NB = sum(nblk) + 1
    kv = torch.empty(2, NB, B, Hkv, D, dtype=torch.bfloat16, device=DEV).uniform_(-scale, scale)
    q = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV).uniform_(-scale, scale)
    perm = torch.randperm(NB - 1, device=DEV).to(torch.int32)
    bt = torch.zeros(S, max(nblk), dtype=torch.int32, device=DEV)
    o = 0
    for i, n in enumerate(nblk):
        bt[i, :n] = perm[o:o + n]
        o += n
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = torch.empty_like(q)
    nbytes = sum(lens) * Hkv * D * 2 * 2 + 2 * S * Hq * D * 2
    fn = lambda: ops.paged_attention_flash(out, q, kv[0], kv[1], scale, bt, sl, L, None)
    return fn, fn, nbytes, S


def build(name):
    if name.endswith("_flash"):
        return build_flash(name)
    lens, Hq, Hkv, D, B, kvd = CONFIGS[name]
    S, L = len(lens), max(lens)
    scale = D ** -0.5
    nblk = [(l + B - 1) // B for l in lens]
    NB = sum(nblk) + 1
    esz = 1 if kvd == "fp8" else 2
    x = 16 // esz
    cdt = torch.uint8 if kvd == "fp8" else torch.bfloat16
    kc = torch.empty(NB, Hkv, D // x, B, x, dtype=cdt, device=DEV)
    vc = torch.empty(NB, Hkv, D, B, dtype=cdt, device=DEV)
    if kvd == "fp8":
        kc.random_(0, 120)
        vc.random_(0, 120)
    else:
        kc.uniform_(-scale, scale)
        vc.uniform_(-scale, scale)
    q = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV).uniform_(-scale, scale)
    perm = torch.randperm(NB - 1, device=DEV).to(torch.int32)
    bt = torch.zeros(S, max(nblk), dtype=torch.int32, device=DEV)
    o = 0
    for i, n in enumerate(nblk):
        bt[i, :n] = perm[o:o + n]
        o += n
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = torch.empty_like(q)
    P = (L + 511) // 512
    tmp = torch.empty(S, Hq, P, D, dtype=q.dtype, device=DEV)
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    nbytes = sum(lens) * Hkv * D * 2 * esz + 2 * S * Hq * D * 2
    v1 = lambda: ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0)
    v2 = lambda: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0)
    return v1, v2, nbytes, S


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    # back to back (what bench.py times): 20 launches between two events, no flush (inputs > L2 except c1/small/one)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return ts[len(ts) // 2], ts[0], e0.elapsed_time(e1) / 20


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c2")
    ap.add_argument("--knobs", nargs="*", default=[""])
    ap.add_argument("--versions", default="v2")
    ap.add_argument("--iters", type=int, default=15)
    args = ap.parse_args()
    for name in args.configs.split(","):
        v1, v2, nbytes, S = build(name)
        for knob in args.knobs:
            for k, v in DEFAULTS.items():
                ops.set_tuning(k, v)
            for kv in filter(None, knob.split(",")):
                k, v = kv.split("=")
                ops.set_tuning(k.strip().lower().replace("b200pa_", ""), int(v))
            for ver in args.versions.split(","):
                fn = v1 if ver == "v1" else v2
                med, best, b2b = timed(fn, args.iters)
                print(f"{name:6s} {ver} [{knob or 'default':30s}] med {med:8.4f} ms best {best:8.4f} ms  "
                      f"{nbytes / med / 1e6:7.0f} GB/s  b2b {b2b:8.4f} ms {nbytes / b2b / 1e6:7.0f} GB/s  "
                      f"tok/s {S / med * 1e3:10.0f}  "
                      f"path={ops.last_attention_path()}/p{ops.last_attention_partitions()}", flush=True)
        del v1, v2
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
