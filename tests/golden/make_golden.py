#!/usr/bin/env python3
"""Generates tests/golden/ref_cpu_golden.npz by running the REFERENCE's own CPU implementation
(csrc/cpu/attention.cpp, csrc/cpu/cache.cpp compiled unchanged into oracle/_ref/ref_cpu*.so by
oracle/build_ref.py) on small seeded inputs.  The reference ships no golden vectors for this path
(SURVEY.md §8c), so these are the committed known-answer fixtures: inputs AND reference outputs.

Run in the build container (needs /root/reference to have been compiled):
    python oracle/build_ref.py cpu && python tests/golden/make_golden.py
bf16 tensors are stored as int16 bit patterns.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests import helpers  # noqa: E402


def to_np(t: torch.Tensor) -> np.ndarray:
    if t.dtype == torch.bfloat16:
        return t.contiguous().view(torch.int16).numpy()
    return t.contiguous().numpy()


def attention_case(ref, name, out, S, Hq, Hkv, D, NB, seq_lens, dtype, alibi, seed):
    B = 16  # the reference CPU extension supports block size 16 only (csrc/cpu/attention.cpp:412-418)
    g = torch.Generator().manual_seed(seed)
    scale = float(D ** -0.5)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).mul(scale).to(dtype)
    kc = (torch.rand(NB, Hkv, D // x, B, x, generator=g) * 2 - 1).mul(scale).to(dtype)
    vc = (torch.rand(NB, Hkv, D, B, generator=g) * 2 - 1).mul(scale).to(dtype)
    max_len = max(seq_lens)
    nblk = (max_len + B - 1) // B
    bt = torch.randint(0, NB, (S, nblk), generator=g, dtype=torch.int32)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    al = torch.randn(Hq, generator=g, dtype=torch.float32) if alibi else None
    o1 = torch.empty_like(q)
    ref.paged_attention_v1(o1, q, kc, vc, Hkv, scale, bt, sl, B, max_len, al, "auto", 1.0, 0, 0, 0, 64, 0)
    P = (max_len + 511) // 512
    o2 = torch.empty_like(q)
    es = torch.zeros(S, Hq, P, dtype=torch.float32)
    ml = torch.zeros(S, Hq, P, dtype=torch.float32)
    tmp = torch.zeros(S, Hq, P, D, dtype=dtype)
    ref.paged_attention_v2(o2, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, max_len, al, "auto", 1.0, 0, 0, 0, 64, 0)
    out.update({
        f"{name}.q": to_np(q), f"{name}.key_cache": to_np(kc), f"{name}.value_cache": to_np(vc),
        f"{name}.block_tables": to_np(bt), f"{name}.seq_lens": to_np(sl),
        f"{name}.meta": np.array([S, Hq, Hkv, D, B, NB, max_len, int(dtype == torch.bfloat16), int(alibi)]),
        f"{name}.scale": np.array([scale], dtype=np.float32),
        f"{name}.out_v1": to_np(o1), f"{name}.out_v2": to_np(o2),
    })
    if al is not None:
        out[f"{name}.alibi"] = to_np(al)
    print(name, "v1-v2 max diff", (o1.float() - o2.float()).abs().max().item())


def reshape_case(ref, name, out, T, H, D, NB, dtype, seed):
    B = 16
    g = torch.Generator().manual_seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    qkv = torch.randn(T, 3, H, D, generator=g).to(dtype)
    _, key, value = qkv.unbind(dim=1)  # non-contiguous rows, as tests/kernels/test_cache.py:137-138
    kc = torch.randn(NB, H, D // x, B, x, generator=g).to(dtype)
    vc = torch.randn(NB, H, D, B, generator=g).to(dtype)
    slots = torch.randperm(NB * B, generator=g)[:T].to(torch.int64)
    slots[T // 2] = -1  # a padding token
    kc0, vc0 = kc.clone(), vc.clone()
    ref.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
    out.update({
        f"{name}.qkv": to_np(qkv), f"{name}.slots": to_np(slots), f"{name}.key_cache_in": to_np(kc0),
        f"{name}.value_cache_in": to_np(vc0), f"{name}.key_cache_out": to_np(kc), f"{name}.value_cache_out": to_np(vc),
        f"{name}.meta": np.array([T, H, D, B, NB, int(dtype == torch.bfloat16)]),
    })
    print(name, "changed K elems", (kc != kc0).sum().item())


def copy_case(ref, name, out, L, NB, H, D, dtype, seed):
    B = 16
    g = torch.Generator().manual_seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    kcs = [torch.randn(NB, H, D // x, B, x, generator=g).to(dtype) for _ in range(L)]
    vcs = [torch.randn(NB, H, D, B, generator=g).to(dtype) for _ in range(L)]
    mapping = torch.tensor([[0, 3], [0, 5], [2, 7], [6, 1]], dtype=torch.int64)
    for i in range(L):
        out[f"{name}.key_cache_in.{i}"] = to_np(kcs[i])
        out[f"{name}.value_cache_in.{i}"] = to_np(vcs[i])
    ref.copy_blocks(kcs, vcs, mapping)
    for i in range(L):
        out[f"{name}.key_cache_out.{i}"] = to_np(kcs[i])
        out[f"{name}.value_cache_out.{i}"] = to_np(vcs[i])
    out[f"{name}.mapping"] = to_np(mapping)
    out[f"{name}.meta"] = np.array([L, NB, H, D, B, int(dtype == torch.bfloat16)])


def main():
    ref = helpers.load_ref_cpu()
    assert ref is not None, "build oracle/_ref first: python oracle/build_ref.py cpu"
    torch.set_num_threads(1)
    out = {}
    attention_case(ref, "attn_bf16_gqa", out, 3, 8, 2, 128, 24, [1, 37, 300], torch.bfloat16, False, 1)
    attention_case(ref, "attn_f32_alibi", out, 2, 4, 2, 64, 44, [17, 700], torch.float32, True, 2)
    attention_case(ref, "attn_bf16_long", out, 1, 4, 1, 128, 80, [1200], torch.bfloat16, False, 3)
    reshape_case(ref, "reshape_bf16", out, 9, 2, 64, 8, torch.bfloat16, 4)
    reshape_case(ref, "reshape_f32", out, 5, 3, 80, 4, torch.float32, 5)
    copy_case(ref, "copy_bf16", out, 2, 8, 2, 64, torch.bfloat16, 6)
    path = os.path.join(HERE, "ref_cpu_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
