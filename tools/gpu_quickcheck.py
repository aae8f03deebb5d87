#!/usr/bin/env python3
"""Developer tool (GPU box): quick numerics + timing sweep of the B200 ops against a torch
reference and, when oracle/_ref/ref_cuda.so is present, the reference's own CUDA kernels.
Not part of the test suite; prints a table.  Usage: python tools/gpu_quickcheck.py [--time]"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import helpers  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402


def ref_v2_call(ref, out, es, ml, tmp, q, kc, vc, nkv, scale, bt, sl, B, maxlen, alibi, kvd, kvs, bs=(0, 0, 64, 0)):
    ref.paged_attention_v2(out, es, ml, tmp, q, str(kc.data_ptr()), kc.stride(0), kc.stride(1),
                           str(vc.data_ptr()), nkv, scale, bt, sl, B, maxlen, alibi, kvd, kvs, 0, *bs)


def attention_case(ref, S, Hq, Hkv, D, B, dtype, kvd, maxlen, alibi=False, seed=0, version="v2", kv_scale=1.0):
    torch.manual_seed(seed)
    dev = "cuda"
    scale = D ** -0.5
    q = torch.empty(S, Hq, D, dtype=dtype, device=dev).uniform_(-scale, scale)
    sl = torch.randint(1, maxlen + 1, (S,), dtype=torch.int32)
    sl[-1] = maxlen
    nblk = (maxlen + B - 1) // B
    NB = max(1024, min(S * nblk, 8192))
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, seed), dtype=torch.int32, device=dev)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 1, Hkv, D, kvd, dtype, seed, dev, ops.convert_fp8)
    kc, vc = kcs[0], vcs[0]
    sl = sl.to(dev)
    al = torch.randn(Hq, dtype=torch.float32, device=dev) if alibi else None
    out = torch.empty_like(q)
    P = (maxlen + 511) // 512
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    tmp = torch.empty(S, Hq, P, D, dtype=dtype, device=dev)
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, maxlen, al, kvd, kv_scale)
    else:
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, maxlen, al, kvd, kv_scale)
    torch.cuda.synchronize()
    path = ops.last_attention_path() + f"/p{ops.last_attention_partitions()}"
    # torch reference (dequantised cache for fp8, as tests/kernels/test_attention.py:232-248)
    if kvd == "fp8":
        kd = torch.empty(kc.shape, dtype=dtype, device=dev)
        vd = torch.empty(vc.shape, dtype=dtype, device=dev)
        ops.convert_fp8(kd, kc, kv_scale)
        ops.convert_fp8(vd, vc, kv_scale)
    else:
        kd, vd = kc, vc
    ref_out = torch.empty_like(q)
    helpers.ref_single_query_cached_kv_attention(ref_out, q, Hq // Hkv, kd, vd, bt, sl, scale, al)
    d_torch = (out.float() - ref_out.float()).abs().max().item()
    d_ref = float("nan")
    if ref is not None:
        out2 = torch.empty_like(q)
        if version == "v1":
            ref.paged_attention_v1(out2, q, kc, vc, Hkv, scale, bt, sl, B, maxlen, al, kvd, kv_scale, 0, 0, 0, 64, 0)
        else:
            ref_v2_call(ref, out2, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, maxlen, al, kvd, kv_scale)
        torch.cuda.synchronize()
        d_ref = (out.float() - out2.float()).abs().max().item()
    print(f"{version} S={S:4d} Hq={Hq:3d} Hkv={Hkv:2d} D={D:3d} B={B:2d} {str(dtype)[6:]:9s} kv={kvd:4s} "
          f"L={maxlen:6d} alibi={int(alibi)} path={path:12s} |d_torch|={d_torch:.2e} |d_refcuda|={d_ref:.2e}",
          flush=True)
    return d_torch


def timed(fn, iters=10, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def timing_case(ref, S, Hq, Hkv, D, B, dtype, kvd, L, label, bs=None, visible_frac=1.0):
    """bs = (local_blocks, vert_stride, block_size, head_sliding_step) runs the block-sparse kernels (v2 only);
    visible_frac scales the algorithmic bytes to the blocks the mask lets through."""
    dev = "cuda"
    scale = D ** -0.5
    nblk = (L + B - 1) // B
    NB = S * nblk + 16
    esz = 1 if kvd == "fp8" else torch.empty(0, dtype=dtype).element_size()
    x = 16 // esz
    cdt = torch.uint8 if kvd == "fp8" else dtype
    kc = torch.empty(NB, Hkv, D // x, B, x, dtype=cdt, device=dev)
    vc = torch.empty(NB, Hkv, D, B, dtype=cdt, device=dev)
    if kvd == "fp8":
        kc.random_(0, 120)
        vc.random_(0, 120)
    else:
        kc.uniform_(-scale, scale)
        vc.uniform_(-scale, scale)
    q = torch.empty(S, Hq, D, dtype=dtype, device=dev).uniform_(-scale, scale)
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 0, distinct=True), dtype=torch.int32, device=dev)
    sl = torch.full((S,), L, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    P = (L + 511) // 512
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=dev)
    ml = torch.empty_like(es)
    tmp = torch.empty(S, Hq, P, D, dtype=dtype, device=dev)
    nbytes = int(S * L * visible_frac) * Hkv * D * 2 * esz + 2 * S * Hq * D * q.element_size()
    rows = []
    bsa = bs if bs is not None else (0, 0, 64, 0)
    f1 = lambda: ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0, 0, *bsa)
    f2 = lambda: ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0, 0, *bsa)
    for name, fn in (("ours v1", f1), ("ours v2", f2)):
        ms = timed(fn)
        rows.append((name, ms, ops.last_attention_path(), ops.last_attention_partitions()))
    if ref is not None:
        if L <= 8192 * 4 and bs is None:
            try:
                g1 = lambda: ref.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0, 0, 0, 0, 64, 0)
                rows.append(("ref  v1", timed(g1, iters=5, warmup=2), "ref", 1))
            except Exception as e:  # noqa: BLE001
                print("ref v1 failed:", e)
        g2 = lambda: ref_v2_call(ref, out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, 1.0, bsa)
        rows.append(("ref  v2", timed(g2, iters=5, warmup=2), "ref", P))
    for name, ms, path, parts in rows:
        print(f"[{label}] {name}: {ms:8.3f} ms  {nbytes / ms / 1e6:8.1f} GB/s  ({nbytes / ms / 1e6 / 6553 * 100:5.1f}% of 6553)  "
              f"tok/s={S / ms * 1e3:10.0f} path={path}/p{parts}", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    ref = helpers.load_ref_cuda()
    print("ref_cuda:", "loaded" if ref is not None else "absent", "| device:", torch.cuda.get_device_name(0))
    bf, hf, f32 = torch.bfloat16, torch.float16, torch.float32
    if not args.no_check:
        worst = 0.0
        for version in ("v1", "v2"):
            for (S, Hq, Hkv, D, B, dt, kvd, L, al) in [
                (7, 32, 8, 128, 16, bf, "auto", 700, False),
                (7, 32, 8, 128, 16, hf, "auto", 3000, True),
                (7, 32, 32, 128, 16, bf, "auto", 1024, False),
                (5, 64, 8, 128, 16, bf, "auto", 2111, False),
                (5, 40, 40, 80, 16, hf, "auto", 900, True),
                (5, 64, 8, 256, 32, bf, "auto", 1500, False),
                (5, 64, 8, 64, 32, hf, "auto", 1500, True),
                (5, 32, 8, 128, 32, bf, "auto", 4099, False),
                (5, 32, 8, 112, 16, bf, "fp8", 1300, False),
                (5, 32, 8, 128, 16, hf, "fp8", 2300, True),
                (5, 64, 8, 128, 32, bf, "fp8", 2300, False),
                (5, 64, 8, 192, 32, hf, "fp8", 777, False),
                (3, 32, 8, 128, 16, f32, "auto", 1000, True),
                (3, 32, 8, 128, 8, bf, "auto", 1000, False),
                (2, 32, 8, 128, 16, bf, "auto", 20000, False),
            ]:
                worst = max(worst, attention_case(ref, S, Hq, Hkv, D, B, dt, kvd, L, al, version=version))
        print("worst |d_torch| =", worst)
    if args.time:
        timing_case(ref, 16, 32, 32, 128, 16, bf, "auto", 1024, "C1 16x1024 mha")
        timing_case(ref, 256, 32, 8, 128, 16, bf, "auto", 4096, "C2 256x4096 gqa4")
        timing_case(ref, 256, 8, 1, 128, 16, bf, "auto", 8192, "C3/tp8 256x8192 g8")
        timing_case(ref, 128, 32, 8, 128, 16, bf, "fp8", 32768, "C4 128x32k fp8")
        # fp32 queries (the general SIMT family) and the two shapes that moved to the mma family in round 2: block 8, block-sparse
        timing_case(ref, 128, 32, 8, 128, 16, f32, "auto", 4096, "SIMT fp32 128x4096")
        timing_case(ref, 128, 32, 8, 128, 8, bf, "auto", 4096, "block8 128x4096")
        nb, qb = 4096 // 64, (4096 - 1) // 64
        vis = sum(1 for k in range(nb) if k > qb - 16 or (k + 1) % 8 == 0)
        timing_case(ref, 128, 32, 8, 128, 16, bf, "auto", 4096, "blocksparse 128x4096", bs=(16, 8, 64, 0),
                    visible_frac=vis / nb)


if __name__ == "__main__":
    main()
