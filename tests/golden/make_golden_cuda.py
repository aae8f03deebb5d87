#!/usr/bin/env python3
"""Generates tests/golden/ref_cuda_golden.npz ON THE B200 BOX by running the REFERENCE's own GPU code on small
seeded inputs and storing inputs + outputs:

  * oracle/_ref/ref_cuda.so      the reference's csrc/attention/attention_kernels.cu, csrc/cache_kernels.cu and
                                 csrc/pos_encoding_kernels.cu compiled unchanged for sm_100a (oracle/build_ref.py cuda)
  * oracle/_ref/prefix_prefill_ref.py  the reference's Triton prefix-prefill module, staged verbatim

The reference ships no golden vectors for this path; these fixtures pin the oracle (CPU suite) and the
product (GPU suite) to the reference's CUDA / Triton arithmetic on every box, whether or not oracle/_ref travelled.

    gpurun -- 'python tests/golden/make_golden_cuda.py gpurun_out/ref_cuda_golden.npz'
    cp gpurun_out/ref_cuda_golden.npz tests/golden/

bf16 / fp16 tensors are stored as int16 bit patterns; `<case>.dtype` names the element type.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests import helpers  # noqa: E402

DEV = "cuda:0"
DT = {"f16": torch.float16, "bf16": torch.bfloat16, "f32": torch.float32}


def to_np(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).numpy()
    return t.numpy()


def attention_case(ref, out, name, dt, kvd, S, Hq, Hkv, D, B, NB, lens, alibi, kv_scale, seed):
    dtype = DT[dt]
    g = torch.Generator().manual_seed(seed)
    scale = float(D ** -0.5)
    esz = 1 if kvd != "auto" else torch.tensor([], dtype=dtype).element_size()
    x = 16 // esz
    q = (torch.rand(S, Hq, D, generator=g) * 2 - 1).mul(scale).to(dtype).to(DEV)
    kc = (torch.rand(NB, Hkv, D // x, B, x, generator=g) * 2 - 1).mul(scale)
    vc = (torch.rand(NB, Hkv, D, B, generator=g) * 2 - 1).mul(scale)
    if kvd == "auto":
        kc, vc = kc.to(dtype).to(DEV), vc.to(dtype).to(DEV)
    else:  # as vllm/utils.py:324-341: random fp16 values quantised by the reference's own convert_fp8
        kh, vh = kc.to(torch.float16).to(DEV), vc.to(torch.float16).to(DEV)
        kc, vc = torch.empty_like(kh, dtype=torch.uint8), torch.empty_like(vh, dtype=torch.uint8)
        ref.convert_fp8(kc, kh, kv_scale, kvd)
        ref.convert_fp8(vc, vh, kv_scale, kvd)
    L = max(lens)
    nblk = (L + B - 1) // B
    bt = torch.randint(0, NB, (S, nblk), generator=g, dtype=torch.int32).to(DEV)
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    al = torch.randn(Hq, generator=g, dtype=torch.float32).to(DEV) if alibi else None
    o1 = torch.empty_like(q)
    ref.paged_attention_v1(o1, q, kc, vc, Hkv, scale, bt, sl, B, L, al, kvd, kv_scale, 0, 0, 0, 64, 0)
    P = (L + 511) // 512
    o2 = torch.empty_like(q)
    es = torch.zeros(S, Hq, P, dtype=torch.float32, device=DEV)
    ml = torch.zeros_like(es)
    tmp = torch.zeros(S, Hq, P, D, dtype=dtype, device=DEV)
    ref.paged_attention_v2(o2, es, ml, tmp, q, str(kc.data_ptr()), kc.stride(0), kc.stride(1), str(vc.data_ptr()), Hkv,
                           scale, bt, sl, B, L, al, kvd, kv_scale, 0, 0, 0, 64, 0)
    torch.cuda.synchronize()
    out.update({f"{name}.q": to_np(q), f"{name}.key_cache": to_np(kc), f"{name}.value_cache": to_np(vc),
                f"{name}.block_tables": to_np(bt), f"{name}.seq_lens": to_np(sl),
                f"{name}.meta": np.array([S, Hq, Hkv, D, B, NB, L]), f"{name}.dtype": np.array(dt),
                f"{name}.kv_dtype": np.array(kvd), f"{name}.scales": np.array([scale, kv_scale], dtype=np.float32),
                f"{name}.out_v1": to_np(o1), f"{name}.out_v2": to_np(o2)})
    if al is not None:
        out[f"{name}.alibi"] = to_np(al)
    print(name, "v1-v2 max diff", (o1.float() - o2.float()).abs().max().item())


def reshape_case(ref, out, name, dt, kvd, T, H, D, B, NB, kv_scale, seed):
    dtype = DT[dt]
    g = torch.Generator().manual_seed(seed)
    esz = 1 if kvd != "auto" else torch.tensor([], dtype=dtype).element_size()
    x = 16 // esz
    cdt = torch.uint8 if kvd != "auto" else dtype
    qkv = torch.randn(T, 3, H, D, generator=g).mul(2.0).to(dtype).to(DEV)
    _, key, value = qkv.unbind(dim=1)  # non-contiguous rows, as tests/kernels/test_cache.py:137-138
    kc = torch.zeros(NB, H, D // x, B, x, dtype=cdt, device=DEV)
    vc = torch.zeros(NB, H, D, B, dtype=cdt, device=DEV)
    slots = torch.randperm(NB * B, generator=g)[:T].to(torch.int64)
    slots[T // 2] = -1  # a padding token
    slots = slots.to(DEV)
    ref.reshape_and_cache(key, value, str(kc.data_ptr()), B, x, str(vc.data_ptr()), slots, kvd, kv_scale)
    torch.cuda.synchronize()
    out.update({f"{name}.qkv": to_np(qkv), f"{name}.slots": to_np(slots), f"{name}.key_cache_out": to_np(kc),
                f"{name}.value_cache_out": to_np(vc), f"{name}.meta": np.array([T, H, D, B, NB]),
                f"{name}.dtype": np.array(dt), f"{name}.kv_dtype": np.array(kvd),
                f"{name}.scales": np.array([kv_scale], dtype=np.float32)})


def convert_case(ref, out, name, dt, seed):
    dtype = DT[dt]
    g = torch.Generator().manual_seed(seed)
    edge = torch.tensor([0.0, -0.0, 448.0, -448.0, 464.0, 465.0, 500.0, -1e4, 2 ** -6, 2 ** -9, 2 ** -10, 1.5 * 2 ** -9,
                         0.0625, 17.0, 19.0, 239.9, 240.1, 1e-8, float("inf"), -float("inf")])
    x = torch.cat([edge, torch.randn(1000, generator=g) * 100, torch.rand(1000, generator=g) * 0.05]).to(dtype).to(DEV)
    x = x.view(4, -1)
    q8 = torch.empty_like(x, dtype=torch.uint8)
    ref.convert_fp8(q8, x, 1.0, "fp8")
    q8s = torch.empty_like(q8)
    ref.convert_fp8(q8s, x, 0.37, "fp8")
    codes = torch.arange(256, dtype=torch.uint8, device=DEV).view(4, 64)
    back = torch.empty(4, 64, dtype=dtype, device=DEV)
    ref.convert_fp8(back, codes, 1.0, "fp8")
    back_s = torch.empty_like(back)
    ref.convert_fp8(back_s, codes, 0.37, "fp8")
    torch.cuda.synchronize()
    out.update({f"{name}.x": to_np(x), f"{name}.q8": to_np(q8), f"{name}.q8_scaled": to_np(q8s),
                f"{name}.codes": to_np(codes), f"{name}.back": to_np(back), f"{name}.back_scaled": to_np(back_s),
                f"{name}.dtype": np.array(dt), f"{name}.scales": np.array([0.37], dtype=np.float32)})


def rope_case(ref, out, name, dt, is_neox, T, Hq, Hkv, D, rot, seed):
    dtype = DT[dt]
    g = torch.Generator().manual_seed(seed)
    max_pos = 4096
    inv_freq = 1.0 / (10000 ** (torch.arange(0, rot, 2, dtype=torch.float) / rot))
    freqs = torch.einsum("i,j -> ij", torch.arange(max_pos, dtype=torch.float), inv_freq)
    cache = torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype).to(DEV)
    pos = torch.randint(0, max_pos, (T,), generator=g).to(DEV)
    q = torch.randn(T, Hq * D, generator=g).to(dtype).to(DEV)
    k = torch.randn(T, Hkv * D, generator=g).to(dtype).to(DEV)
    q0, k0 = q.clone(), k.clone()
    ref.rotary_embedding(pos, q, k, D, cache, is_neox)
    torch.cuda.synchronize()
    out.update({f"{name}.positions": to_np(pos), f"{name}.query_in": to_np(q0), f"{name}.key_in": to_np(k0),
                f"{name}.query_out": to_np(q), f"{name}.key_out": to_np(k), f"{name}.cos_sin_cache": to_np(cache),
                f"{name}.meta": np.array([T, Hq, Hkv, D, rot, int(is_neox)]), f"{name}.dtype": np.array(dt)})


def prefill_case(mod, out, name, dt, mode, seed):
    dtype = DT[dt]
    Hq, qpk, D, B = 8, 4, 64, 16
    case = helpers.make_prefill_case(Hq, qpk, D, dtype, DEV, query_lens=[70, 5, 33], ctx_lens=[100, 48, 0], block_size=B,
                                     amplitude=1.0, seed=seed)
    slopes = None
    window = 0
    if mode == "alibi":
        slopes = torch.tensor([2.0 ** (-(i + 1)) for i in range(Hq)], dtype=torch.float32, device=DEV)
    if mode == "window":
        window = 40
    o = torch.empty_like(case["query"])
    mod.context_attention_fwd(case["query"], case["k"], case["v"], o, case["k_cache"], case["v_cache"],
                              case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                              case["max_input_len"], slopes, window)
    torch.cuda.synchronize()
    for f in ("query", "k", "v", "k_cache", "v_cache", "block_table", "b_start_loc", "b_seq_len", "b_ctx_len"):
        out[f"{name}.{f}"] = to_np(case[f])
    out[f"{name}.out"] = to_np(o)
    out[f"{name}.meta"] = np.array([Hq, Hq // qpk, D, B, case["max_input_len"], window])
    out[f"{name}.dtype"] = np.array(dt)
    if slopes is not None:
        out[f"{name}.alibi"] = to_np(slopes)


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "ref_cuda_golden.npz")
    ref = helpers.load_ref_cuda()
    assert ref is not None, "oracle/_ref/ref_cuda.so missing: python oracle/build_ref.py cuda (in the build container)"
    out = {}
    attention_case(ref, out, "attn_bf16_gqa", "bf16", "auto", 4, 8, 2, 128, 16, 96, [1, 37, 300, 1200], False, 1.0, 1)
    attention_case(ref, out, "attn_f16_alibi_b32", "f16", "auto", 3, 4, 4, 64, 32, 40, [5, 64, 700], True, 1.0, 2)
    attention_case(ref, out, "attn_bf16_fp8", "bf16", "fp8", 3, 8, 2, 128, 16, 80, [17, 513, 1000], False, 0.25, 3)
    attention_case(ref, out, "attn_f16_fp8_alibi", "f16", "fp8", 2, 4, 1, 128, 32, 30, [33, 600], True, 1.0, 4)
    attention_case(ref, out, "attn_f32_b8", "f32", "auto", 2, 4, 2, 80, 8, 40, [9, 130], False, 1.0, 5)
    reshape_case(ref, out, "reshape_bf16_fp8", "bf16", "fp8", 11, 2, 64, 16, 6, 0.5, 6)
    reshape_case(ref, out, "reshape_f16_b32", "f16", "auto", 9, 3, 80, 32, 4, 1.0, 7)
    reshape_case(ref, out, "reshape_f32_fp8", "f32", "fp8", 7, 2, 64, 8, 9, 2.0, 8)
    for dt in ("f16", "bf16", "f32"):
        convert_case(ref, out, f"convert_{dt}", dt, 9)
        rope_case(ref, out, f"rope_neox_{dt}", dt, True, 7, 4, 2, 64, 64, 10)
        rope_case(ref, out, f"rope_gptj_{dt}", dt, False, 5, 6, 2, 80, 32, 11)
    mod = helpers.load_ref_prefix_prefill()
    if mod is not None:
        for mode in ("plain", "window", "alibi"):
            try:
                prefill_case(mod, out, f"prefill_{mode}_f16", "f16", mode, 12)
                prefill_case(mod, out, f"prefill_{mode}_bf16", "bf16", mode, 13)
            except Exception as e:  # noqa: BLE001
                print("reference Triton kernel failed:", mode, type(e).__name__, e)
    else:
        print("oracle/_ref/prefix_prefill_ref.py not staged: no prefill fixtures")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB;", len(out), "arrays")


if __name__ == "__main__":
    main()
