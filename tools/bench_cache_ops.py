#!/usr/bin/env python3
"""Developer tool (GPU box): throughput of reshape_and_cache / copy_blocks / swap_blocks / convert_fp8 on the
BASELINE config-4/5 shapes, next to the reference's own CUDA kernels (oracle/_ref/ref_cuda.so) when present.
Algorithmic bytes per SURVEY.md §8(d)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import helpers  # noqa: E402
from vllm_legacy_b200 import _custom_ops as ops  # noqa: E402

DEV = "cuda"


def timed(fn, iters=20, warmup=3, flush=True):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    fl = torch.empty(256 << 20, dtype=torch.uint8, device=DEV) if flush else None
    ts = []
    for _ in range(iters):
        if fl is not None:
            fl.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def report(name, ms, nbytes, ref_ms=None):
    s = f"{name:58s} {ms * 1e3:9.1f} us  {nbytes / ms / 1e6:8.1f} GB/s"
    if ref_ms is not None:
        s += f"   | reference CUDA {ref_ms * 1e3:9.1f} us  {nbytes / ref_ms / 1e6:8.1f} GB/s  ({ref_ms / ms:4.1f}x)"
    print(s, flush=True)


def main():
    ref = helpers.load_ref_cuda()
    H, D, B = 8, 128, 16
    dt = torch.bfloat16
    # ---- reshape_and_cache ---------------------------------------------------------------------------------
    for T, consecutive in ((128, False), (256, False), (4096, True), (32768, True), (32768, False)):
        for kvd in ("auto", "fp8"):
            NB = T // B + 1024
            esz = 1 if kvd == "fp8" else 2
            x = 16 // esz
            cdt = torch.uint8 if kvd == "fp8" else dt
            kc = torch.zeros(NB, H, D // x, B, x, dtype=cdt, device=DEV)
            vc = torch.zeros(NB, H, D, B, dtype=cdt, device=DEV)
            qkv = torch.randn(T, 3, H, D, dtype=dt, device=DEV)
            _, key, value = qkv.unbind(dim=1)
            if consecutive:  # a prefill chunk: each sequence's tokens fill whole blocks in order
                blocks = torch.randperm(NB, device=DEV)[:T // B]
                slots = (blocks[:, None] * B + torch.arange(B, device=DEV)[None, :]).flatten()
            else:
                slots = torch.randperm(NB * B, device=DEV)[:T]
            nbytes = T * H * D * 2 * (2 + esz)
            ms = timed(lambda: ops.reshape_and_cache(key, value, kc, vc, slots, kvd, 0.5))
            rms = None
            if ref is not None:
                rms = timed(lambda: ref.reshape_and_cache(key, value, str(kc.data_ptr()), B, x, str(vc.data_ptr()),
                                                          slots, kvd, 0.5))
            report(f"reshape_and_cache T={T} {kvd} {'consecutive' if consecutive else 'random'} slots", ms, nbytes, rms)
    # ---- copy_blocks: 256 pairs x 32 layers (config 5) ---------------------------------------------------------
    L, NB = 32, 2048
    kcs = [torch.randn(NB, H, D // 8, B, 8, device=DEV).to(dt) for _ in range(L)]
    vcs = [torch.randn(NB, H, D, B, device=DEV).to(dt) for _ in range(L)]
    perm = torch.randperm(NB)
    mapping = torch.stack([perm[:256], perm[256:512]], dim=1).to(DEV)
    block_bytes = kcs[0][0].numel() * 2
    nbytes = L * 256 * 2 * block_bytes * 2
    ms = timed(lambda: ops.copy_blocks(kcs, vcs, mapping))
    rms = timed(lambda: ref.copy_blocks(kcs, vcs, mapping)) if ref is not None else None
    report("copy_blocks 32 layers x 256 pairs (32 KiB K + 32 KiB V each)", ms, nbytes, rms)
    import time
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        ops.copy_blocks(kcs, vcs, mapping)
    torch.cuda.synchronize()
    print(f"  copy_blocks wall per call incl. host: {(time.perf_counter() - t0) / 20 * 1e6:.1f} us", flush=True)
    if ref is not None:
        t0 = time.perf_counter()
        for _ in range(20):
            ref.copy_blocks(kcs, vcs, mapping)
        torch.cuda.synchronize()
        print(f"  reference wall per call incl. host (2 H2D pointer copies): {(time.perf_counter() - t0) / 20 * 1e6:.1f} us")
    # ---- swap_blocks: 256 pairs each direction, pinned host ---------------------------------------------------------
    host = torch.zeros(512, H, D // 8, B, 8, dtype=dt).pin_memory()
    for name, runs in (("scattered", False), ("contiguous run", True)):
        src = torch.arange(256) if runs else perm[:256]
        m_out = torch.stack([src, torch.arange(256)], dim=1)
        m_in = torch.stack([torch.arange(256), src], dim=1)
        nb = 256 * block_bytes
        for label, a, b, m in (("D2H", kcs[0], host, m_out), ("H2D", host, kcs[0], m_in)):
            def run(fn=ops.swap_blocks):
                fn(a, b, m)
            ms = timed(run, flush=False)
            rms = timed(lambda: ref.swap_blocks(a, b, m), flush=False) if ref is not None else None
            report(f"swap_blocks {label} 256 x 32 KiB, {name}", ms, nb, rms)
    # ---- convert_fp8 ---------------------------------------------------------------------------------------------
    x = torch.randn(64 << 20, dtype=dt, device=DEV)
    q8 = torch.empty(x.shape, dtype=torch.uint8, device=DEV)
    ms = timed(lambda: ops.convert_fp8(q8, x, 1.0, "fp8"))
    rms = timed(lambda: ref.convert_fp8(q8.view(1024, -1), x.view(1024, -1), 1.0, "fp8")) if ref is not None else None
    report("convert_fp8 bf16 -> e4m3, 64 Mi elements", ms, x.numel() * 3, rms)


if __name__ == "__main__":
    main()
