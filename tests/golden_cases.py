"""Shared drivers for the reference-CUDA / reference-Triton fixtures (tests/golden/ref_cuda_golden.npz): each runs one
fixture through an implementation with the `vllm._custom_ops` surface (the CPU oracle in the CPU suite, the product ops
on the GPU) and compares with the reference's stored output.  Bit-exact for cache writes, fp8 conversion and RoPE;
attention within the north-star tolerance (atol 1e-3 / rtol 1e-2; fp8 KV: atol 2e-3 because both sides requantise
products differently; fp32: 1e-5-class)."""
import torch


def _bits(t):
    return t.contiguous().view(torch.uint8)


def _dev(t, device):
    return t.to(device) if torch.is_tensor(t) else t


def check_attention(impl, c, device):
    S, Hq, Hkv, D, B, NB, L = (int(v) for v in c["meta"])
    scale, kv_scale = (float(v) for v in c["scales"])
    kvd = c["kv_dtype"]
    q, kc, vc = _dev(c["q"], device), _dev(c["key_cache"], device), _dev(c["value_cache"], device)
    bt, sl = _dev(c["block_tables"], device), _dev(c["seq_lens"], device)
    al = _dev(c["alibi"], device) if "alibi" in c else None
    o1 = torch.empty_like(q)
    impl.paged_attention_v1(o1, q, kc, vc, Hkv, scale, bt, sl, B, L, al, kvd, kv_scale)
    P = (L + 511) // 512
    o2 = torch.empty_like(q)
    es = torch.zeros(S, Hq, P, dtype=torch.float32, device=device)
    ml = torch.zeros_like(es)
    tmp = torch.zeros(S, Hq, P, D, dtype=q.dtype, device=device)
    impl.paged_attention_v2(o2, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, al, kvd, kv_scale)
    if q.dtype == torch.float32:
        atol, rtol = 2e-5, 1e-4
    else:
        atol, rtol = (2e-3 if kvd != "auto" else 1e-3), 1e-2
    for got, want in ((o1, c["out_v1"]), (o2, c["out_v2"])):
        torch.testing.assert_close(got.cpu().float(), want.float(), atol=atol, rtol=rtol)


def check_reshape(impl, c, device):
    T, H, D, B, NB = (int(v) for v in c["meta"])
    kvd, kv_scale = c["kv_dtype"], float(c["scales"][0])
    qkv = _dev(c["qkv"], device)
    _, key, value = qkv.unbind(dim=1)
    kc = torch.zeros_like(_dev(c["key_cache_out"], device))
    vc = torch.zeros_like(_dev(c["value_cache_out"], device))
    impl.reshape_and_cache(key, value, kc, vc, _dev(c["slots"], device), kvd, kv_scale)
    assert torch.equal(_bits(kc.cpu()), _bits(c["key_cache_out"]))
    assert torch.equal(_bits(vc.cpu()), _bits(c["value_cache_out"]))


def check_convert(impl, c, device):
    x, codes = _dev(c["x"], device), _dev(c["codes"], device)
    s = float(c["scales"][0])
    for scale, want in ((1.0, c["q8"]), (s, c["q8_scaled"])):
        got = torch.empty_like(x, dtype=torch.uint8)
        impl.convert_fp8(got, x, scale, "fp8")
        assert torch.equal(got.cpu(), want), f"quantise, scale {scale}"
    for scale, want in ((1.0, c["back"]), (s, c["back_scaled"])):
        got = torch.empty(codes.shape, dtype=x.dtype, device=device)
        impl.convert_fp8(got, codes, scale, "fp8")
        g, w = got.cpu(), want
        nan = torch.isnan(w.float())
        assert torch.equal(torch.isnan(g.float()), nan)                     # 0x7f / 0xff are NaN in e4m3
        assert torch.equal(_bits(g)[~nan.repeat_interleave(g.element_size(), -1).view(_bits(g).shape)],
                           _bits(w)[~nan.repeat_interleave(w.element_size(), -1).view(_bits(w).shape)]), \
            f"dequantise, scale {scale}"


def check_rope(impl, c, device):
    T, Hq, Hkv, D, rot, is_neox = (int(v) for v in c["meta"])
    q, k = _dev(c["query_in"].clone(), device), _dev(c["key_in"].clone(), device)
    impl.rotary_embedding(_dev(c["positions"], device), q, k, D, _dev(c["cos_sin_cache"], device), bool(is_neox))
    assert torch.equal(_bits(q.cpu()), _bits(c["query_out"]))
    assert torch.equal(_bits(k.cpu()), _bits(c["key_out"]))


def check_prefill(impl, c, device):
    Hq, Hkv, D, B, max_input_len, window = (int(v) for v in c["meta"])
    g = {n: _dev(c[n], device) for n in ("query", "k", "v", "k_cache", "v_cache", "block_table", "b_start_loc",
                                          "b_seq_len", "b_ctx_len")}
    al = _dev(c["alibi"], device) if "alibi" in c else None
    o = torch.empty_like(g["query"])
    impl.context_attention_fwd(g["query"], g["k"], g["v"], o, g["k_cache"], g["v_cache"], g["block_table"],
                               g["b_start_loc"], g["b_seq_len"], g["b_ctx_len"], max_input_len, al, window)
    torch.testing.assert_close(o.cpu().float(), c["out"].float(), atol=2e-3, rtol=1e-2)


CHECKS = (("attn_", check_attention), ("reshape_", check_reshape), ("convert_", check_convert), ("rope_", check_rope),
          ("prefill_", check_prefill))


def check_case(impl, name, case, device):
    for prefix, fn in CHECKS:
        if name.startswith(prefix):
            return fn(impl, case, device)
    raise KeyError(name)
