"""Pins the CPU oracle (oracle/paged_attn_oracle.c) BEFORE it is trusted as the checker:
  - against the committed golden fixtures produced by the reference's own CPU code,
  - against that reference CPU code run live (oracle/_ref/ref_cpu*.so) when it is present,
  - against CUDA 12.9's cuda_fp8.h host conversions (oracle/fp8_pin) and torch.float8 casts,
  - against the torch restatement of the reference's test reference (test_attention.py:43-111).
CPU only."""
import os
import random
import subprocess

import pytest
import torch

from tests import helpers

BF, HF, F32 = torch.bfloat16, torch.float16, torch.float32


# ------------------------------------------------------------------------------------------ fp8
def test_fp8_pin_against_cuda_fp8_header(oracle):
    """Exhaustive fp16/bf16 + 16M-point fp32 sweep vs __nv_cvt_float_to_fp8 / __nv_cvt_fp8_to_halfraw."""
    here = os.path.dirname(oracle.LIB)
    if not os.path.exists("/usr/local/cuda/include/cuda_fp8.h"):
        pytest.skip("CUDA toolkit headers absent")
    subprocess.check_call(["make", "-C", here, "fp8_pin"], stdout=subprocess.DEVNULL)
    r = subprocess.run([os.path.join(here, "fp8_pin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    assert " 0 mismatches" in r.stdout


@pytest.mark.parametrize("dtype", [HF, BF, F32])
def test_fp8_e4m3_quant_matches_torch(oracle, dtype):
    """SURVEY §8c probe: SATFINITE e4m3 == clamp(+-448) then torch float8_e4m3fn cast."""
    g = torch.Generator().manual_seed(0)
    x = (torch.rand(1 << 16, generator=g) * 2 - 1).mul(600).to(dtype)
    x[:8] = torch.tensor([0.0, -0.0, 448.0, -448.0, 464.0, 465.0, 1e-4, -3e-3]).to(dtype)
    for scale in (1.0, 0.0418, 2.5):
        out = torch.empty(x.shape, dtype=torch.uint8)
        oracle.convert_fp8(out, x, scale, "fp8")
        want = (x.float() / scale).clamp(-448, 448).to(torch.float8_e4m3fn).view(torch.uint8)
        assert torch.equal(out, want)


def test_fp8_e4m3_dequant_all_codes(oracle):
    codes = torch.arange(256, dtype=torch.uint8)
    for dtype in (HF, BF, F32):
        out = torch.empty(256, dtype=dtype)
        oracle.convert_fp8(out, codes, 1.0, "fp8")
        want = codes.view(torch.float8_e4m3fn).float().to(dtype)
        ok = (out == want) | (out.isnan() & want.isnan())
        assert ok.all()
    out = torch.empty(256, dtype=BF)
    oracle.convert_fp8(out, codes, 0.0418, "fp8")
    want = (codes.view(torch.float8_e4m3fn).to(HF).float() * 0.0418).to(BF)
    assert ((out == want) | (out.isnan() & want.isnan())).all()


def test_fp8_roundtrip_tolerance(oracle):
    """tests/kernels/test_cache.py:347-380: uniform(-224, 224) -> fp8 -> back, atol 1e-3 rtol 0.1."""
    x = torch.empty(4, 64, 16, dtype=HF).uniform_(-224, 224)
    q = torch.empty(x.shape, dtype=torch.uint8)
    oracle.convert_fp8(q, x)
    y = torch.empty_like(x)
    oracle.convert_fp8(y, q)
    assert torch.allclose(x, y, atol=1e-3, rtol=0.1)


def test_convert_fp8_rejects_auto(oracle):
    with pytest.raises(RuntimeError):
        oracle.convert_fp8(torch.empty(4, dtype=torch.uint8), torch.zeros(4, dtype=HF), 1.0, "auto")


# ------------------------------------------------------------------------------- cache ops vs golden
@pytest.mark.parametrize("case", ["reshape_bf16", "reshape_f32"])
def test_reshape_and_cache_golden(oracle, golden, case):
    c = golden[case]
    _, key, value = c["qkv"].unbind(dim=1)
    kc, vc = c["key_cache_in"].clone(), c["value_cache_in"].clone()
    oracle.reshape_and_cache(key, value, kc, vc, c["slots"], "auto", 1.0)
    assert torch.equal(kc.view(torch.uint8), c["key_cache_out"].view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), c["value_cache_out"].view(torch.uint8))


def test_copy_blocks_golden(oracle, golden):
    c = golden["copy_bf16"]
    L = int(c["meta"][0])
    kcs = [c[f"key_cache_in.{i}"].clone() for i in range(L)]
    vcs = [c[f"value_cache_in.{i}"].clone() for i in range(L)]
    oracle.copy_blocks(kcs, vcs, c["mapping"])
    for i in range(L):
        assert torch.equal(kcs[i].view(torch.int16), c[f"key_cache_out.{i}"].view(torch.int16))
        assert torch.equal(vcs[i].view(torch.int16), c[f"value_cache_out.{i}"].view(torch.int16))


# ----------------------------------------------------------------------- cache ops vs live reference
@pytest.mark.parametrize("dtype", [BF, F32])
@pytest.mark.parametrize("num_heads,head_size", [(8, 128), (3, 80)])
def test_reshape_and_cache_vs_ref_cpu(oracle, ref_cpu, dtype, num_heads, head_size):
    torch.manual_seed(0)
    T, B, NB = 42, 16, 20
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    qkv = torch.randn(T, 3, num_heads, head_size).to(dtype)
    _, key, value = qkv.unbind(dim=1)
    slots = torch.tensor(random.Random(0).sample(range(NB * B), T), dtype=torch.int64)
    slots[5] = -1
    kc = torch.randn(NB, num_heads, head_size // x, B, x).to(dtype)
    vc = torch.randn(NB, num_heads, head_size, B).to(dtype)
    kc2, vc2 = kc.clone(), vc.clone()
    ref_cpu.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
    oracle.reshape_and_cache(key, value, kc2, vc2, slots, "auto", 1.0)
    assert torch.equal(kc.view(torch.uint8), kc2.view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), vc2.view(torch.uint8))


def test_copy_blocks_vs_ref_cpu(oracle, ref_cpu):
    torch.manual_seed(0)
    L, NB, H, D, B = 3, 32, 4, 64, 16
    kcs = [torch.randn(NB, H, D // 8, B, 8).to(BF) for _ in range(L)]
    vcs = [torch.randn(NB, H, D, B).to(BF) for _ in range(L)]
    k2, v2 = [t.clone() for t in kcs], [t.clone() for t in vcs]
    rng = random.Random(1)
    src = rng.sample(range(NB // 2), 6)
    dst = rng.sample(range(NB // 2, NB), 12)
    mapping = torch.tensor([[s, dst[2 * i + j]] for i, s in enumerate(src) for j in range(2)], dtype=torch.int64)
    ref_cpu.copy_blocks(kcs, vcs, mapping)
    oracle.copy_blocks(k2, v2, mapping)
    for a, b in zip(kcs + vcs, k2 + v2):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))


def test_swap_blocks_oracle():
    from oracle import oracle as o
    src = torch.randn(8, 2, 16, 16).to(BF)
    dst = torch.zeros(8, 2, 16, 16).to(BF)
    m = torch.tensor([[0, 5], [3, 1], [7, 7]], dtype=torch.int64)
    o.swap_blocks(src, dst, m)
    for s, d in m.tolist():
        assert torch.equal(dst[d], src[s])
    assert dst[0].abs().sum() == 0


# ---------------------------------------------------------------------------------------- attention
def _run_oracle_attention(oracle, c, version):
    S, Hq, Hkv, D, B, NB, max_len, _, has_alibi = [int(v) for v in c["meta"]]
    q = c["q"]
    out = torch.empty_like(q)
    al = c["alibi"] if has_alibi else None
    scale = float(c["scale"][0])
    if version == "v1":
        oracle.paged_attention_v1(out, q, c["key_cache"], c["value_cache"], Hkv, scale, c["block_tables"],
                                  c["seq_lens"], B, max_len, al, "auto", 1.0)
    else:
        P = (max_len + 511) // 512
        es = torch.zeros(S, Hq, P, dtype=F32)
        ml = torch.zeros(S, Hq, P, dtype=F32)
        tmp = torch.zeros(S, Hq, P, D, dtype=q.dtype)
        oracle.paged_attention_v2(out, es, ml, tmp, q, c["key_cache"], c["value_cache"], Hkv, scale,
                                  c["block_tables"], c["seq_lens"], B, max_len, al, "auto", 1.0)
    return out


@pytest.mark.parametrize("case", ["attn_bf16_gqa", "attn_f32_alibi", "attn_bf16_long"])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_attention_golden(oracle, golden, case, version):
    """Oracle (restating the CUDA kernel's arithmetic) vs the reference CPU kernel's output.  The two
    reference implementations differ in rounding (CPU keeps P in fp32 and has no +1e-6), so the bar is
    the reference tests' own tolerance (test_attention.py:271-274: atol 1e-3)."""
    c = golden[case]
    out = _run_oracle_attention(oracle, c, version)
    want = c["out_" + version]
    assert torch.allclose(out.float(), want.float(), atol=1e-3, rtol=1e-2)
    # and in practice much closer
    assert (out.float() - want.float()).abs().max() < 3e-4


@pytest.mark.parametrize("dtype", [BF, HF, F32])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_attention_vs_torch_reference(oracle, dtype, kv, block_size):
    """Oracle vs the torch restatement of tests/kernels/test_attention.py:58-111 (all block sizes,
    fp16, fp8 KV — configurations the reference CPU extension cannot run)."""
    torch.manual_seed(0)
    S, Hq, Hkv, D, NB, max_len = 3, 8, 2, 64, 40, 200
    scale = D ** -0.5
    q = torch.empty(S, Hq, D, dtype=dtype).uniform_(-scale, scale)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, block_size, 1, Hkv, D, kv, dtype, 0, "cpu",
                                                    oracle.convert_fp8)
    kc, vc = kcs[0], vcs[0]
    nblk = (max_len + block_size - 1) // block_size
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 0), dtype=torch.int32)
    sl = torch.tensor([1, 77, max_len], dtype=torch.int32)
    al = torch.randn(Hq, dtype=F32)
    out = torch.empty_like(q)
    oracle.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, block_size, max_len, al, kv, 1.0)
    if kv == "fp8":
        kd, vd = torch.empty(kc.shape, dtype=dtype), torch.empty(vc.shape, dtype=dtype)
        oracle.convert_fp8(kd, kc)
        oracle.convert_fp8(vd, vc)
    else:
        kd, vd = kc, vc
    want = torch.empty_like(q)
    helpers.ref_single_query_cached_kv_attention(want, q.float(), Hq // Hkv, kd.float(), vd.float(), bt, sl, scale, al)
    assert torch.allclose(out.float(), want.float(), atol=1e-3, rtol=1e-5 if dtype == F32 else 1e-2)


@pytest.mark.parametrize("dtype", [BF, F32])
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_attention_vs_ref_cpu(oracle, ref_cpu, dtype, version):
    """BASELINE config-1 shape, shrunk: oracle vs the reference CPU extension run live."""
    torch.manual_seed(0)
    S, Hq, Hkv, D, B, NB, max_len = 4, 8, 8, 128, 16, 128, 1024
    scale = D ** -0.5
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    q = torch.empty(S, Hq, D, dtype=dtype).uniform_(-scale, scale)
    kc = torch.empty(NB, Hkv, D // x, B, x, dtype=dtype).uniform_(-scale, scale)
    vc = torch.empty(NB, Hkv, D, B, dtype=dtype).uniform_(-scale, scale)
    bt = torch.tensor(helpers.make_block_tables(S, max_len // B, NB, 0), dtype=torch.int32)
    sl = torch.tensor([1, 513, 700, max_len], dtype=torch.int32)
    P = (max_len + 511) // 512
    outs = []
    for impl in (ref_cpu, oracle):
        out = torch.empty_like(q)
        if version == "v1":
            impl.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, max_len, None, "auto", 1.0, 0, 0, 0, 64, 0)
        else:
            es, ml = torch.zeros(S, Hq, P, dtype=F32), torch.zeros(S, Hq, P, dtype=F32)
            tmp = torch.zeros(S, Hq, P, D, dtype=dtype)
            impl.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, max_len, None, "auto", 1.0,
                                    0, 0, 0, 64, 0)
        outs.append(out.float())
    assert torch.allclose(outs[0], outs[1], atol=1e-3, rtol=1e-2)


def test_attention_blocksparse_masks_blocks(oracle):
    """Block-sparse predicate (attention_kernels.cu:212-256): equals dense attention over the attended
    tokens only (the mask of tests/kernels/test_blocksparse_attention.py:114-136)."""
    torch.manual_seed(0)
    S, Hq, Hkv, D, B, NB, L = 1, 4, 2, 64, 16, 64, 600
    bs_block, local, vert = 64, 2, 4
    scale = D ** -0.5
    q = torch.empty(S, Hq, D, dtype=F32).uniform_(-scale, scale)
    kc = torch.empty(NB, Hkv, D // 4, B, 4, dtype=F32).uniform_(-scale, scale)
    vc = torch.empty(NB, Hkv, D, B, dtype=F32).uniform_(-scale, scale)
    bt = torch.tensor(helpers.make_block_tables(S, (L + B - 1) // B, NB, 0, distinct=True), dtype=torch.int32)
    sl = torch.tensor([L], dtype=torch.int32)
    for slide in (0, 2, -1):
        out = torch.empty_like(q)
        oracle.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0, 0, local, vert,
                                  bs_block, slide)
        idx = torch.arange(L)
        keys = kc[bt[0, idx // B].long(), :, :, idx % B, :].reshape(L, Hkv, D)
        vals = vc[bt[0, idx // B].long(), :, :, idx % B]
        q_blk = (L - 1) // bs_block
        for h in range(Hq):
            kvh = h // (Hq // Hkv)
            off = (h * slide + 1) if slide >= 0 else (kvh * (-slide) + 1)
            kb = idx // bs_block
            mask = ((kb + off) % vert == 0) | (kb > q_blk - local)
            w = torch.softmax((scale * keys[:, kvh] @ q[0, h]).masked_fill(~mask, float("-inf")), 0)
            want = w @ vals[:, kvh]
            assert torch.allclose(out[0, h], want, atol=1e-5, rtol=1e-4), (slide, h)


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16, torch.float])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_oracle_reshape_and_cache_flash_matches_reference_test_loop(oracle, dtype, block_size):
    """oracle.reshape_and_cache_flash == the index assignment of tests/kernels/test_cache.py:265-279."""
    torch.manual_seed(0)
    T, H, D, NB = 42, 8, 80, 32
    slots = torch.randperm(NB * block_size)[:T]
    slots[7] = -1
    qkv = torch.randn(T, 3, H, D, dtype=dtype)
    _, key, value = qkv.unbind(dim=1)
    kc = torch.randn(NB, block_size, H, D, dtype=dtype)
    vc = torch.randn(NB, block_size, H, D, dtype=dtype)
    ek, ev = kc.clone(), vc.clone()
    oracle.reshape_and_cache_flash(key, value, kc, vc, slots, "auto")
    for i, s in enumerate(slots.tolist()):
        if s >= 0:
            ek[s // block_size, s % block_size] = key[i]
            ev[s // block_size, s % block_size] = value[i]
    assert torch.equal(kc.view(torch.uint8), ek.view(torch.uint8))
    assert torch.equal(vc.view(torch.uint8), ev.view(torch.uint8))
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        oracle.reshape_and_cache_flash(key, value, kc, vc, slots, "fp8")


# ------------------------------------------------------------------------------------- rotary_embedding
def _rope_native(positions, query, key, head_size, cache, is_neox):
    """vllm/model_executor/layers/rotary_embedding.py:105-147 (_forward_native), the checker the reference's
    own tests/kernels/test_pos_encoding.py uses; evaluated per-op in the tensor dtype like the CUDA kernel."""
    rot = cache.shape[1]
    cos, sin = cache[positions.flatten()].chunk(2, dim=-1)
    cos, sin = cos[:, None, :], sin[:, None, :]
    outs = []
    for a in (query, key):
        a = a.reshape(positions.numel(), -1, head_size).clone()
        x, y = (a[..., :rot // 2], a[..., rot // 2:rot]) if is_neox else (a[..., 0:rot:2], a[..., 1:rot:2])
        nx, ny = x * cos - y * sin, y * cos + x * sin
        if is_neox:
            a[..., :rot // 2], a[..., rot // 2:rot] = nx, ny
        else:
            a[..., 0:rot:2], a[..., 1:rot:2] = nx, ny
        outs.append(a)
    return outs


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("head_size,rot", [(64, 64), (80, 32), (128, 128), (256, 32)])
def test_oracle_rotary_embedding_vs_native(oracle, is_neox, dtype, head_size, rot):
    torch.manual_seed(0)
    T, H, Hkv = 23, 7, 3
    inv = 1.0 / (10000 ** (torch.arange(0, rot, 2, dtype=torch.float) / rot))
    fr = torch.einsum("i,j -> ij", torch.arange(512, dtype=torch.float), inv)
    cache = torch.cat((fr.cos(), fr.sin()), dim=-1).to(dtype)
    positions = torch.randint(0, 512, (T,))
    q = torch.randn(T, H * head_size, dtype=dtype)
    k = torch.randn(T, Hkv * head_size, dtype=dtype)
    want_q, want_k = _rope_native(positions, q, k, head_size, cache, is_neox)
    oracle.rotary_embedding(positions, q, k, head_size, cache, is_neox)
    if dtype == torch.float32:  # the kernel's fma contraction differs from torch's separate ops by <= 1 ulp
        torch.testing.assert_close(q.view(T, H, head_size), want_q, atol=1e-6, rtol=1e-6)
        torch.testing.assert_close(k.view(T, Hkv, head_size), want_k, atol=1e-6, rtol=1e-6)
    else:
        assert torch.equal(q.view(T, H, head_size), want_q) and torch.equal(k.view(T, Hkv, head_size), want_k)


# ---------------------------------------------------------------------------------------------------
# context_attention_fwd oracle (prefix_prefill.py restatement) vs independent fp32 dense attention
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("mode", ["plain", "window", "alibi"])
def test_oracle_context_attention_vs_dense(oracle, dtype, mode):
    from tests import helpers
    case = helpers.make_prefill_case(num_heads=4, num_queries_per_kv=2, head_size=32, dtype=dtype, device="cpu",
                                     query_lens=[5, 150, 37], ctx_lens=[200, 17, 0], block_size=16, amplitude=1.0)
    slopes = torch.tensor([0.5, 0.25, 0.125, 0.0625]) if mode == "alibi" else None
    window = 48 if mode == "window" else 0
    out = torch.empty_like(case["query"])
    oracle.context_attention_fwd(case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"],
                                 case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                                 case["max_input_len"], slopes, window)
    ref = helpers.ref_context_attention(case, slopes, window)
    tol = {torch.float16: 2e-3, torch.bfloat16: 1e-2, torch.float32: 1e-5}[dtype]
    torch.testing.assert_close(out.float(), ref, atol=tol, rtol=1e-2)


def test_fp8_prefix_cache_dequant_layout(oracle):
    """quantize_prefill_case (x = 8 -> fp8 x = 16 layout) and oracle.dequant_prefix_caches (back) are inverse
    re-layouts, and the values are exactly convert_fp8(convert_fp8(x)) element by element."""
    from tests import helpers
    case = helpers.make_prefill_case(4, 2, 32, torch.bfloat16, "cpu", query_lens=[3, 5], ctx_lens=[20, 9],
                                     block_size=8, amplitude=1.0)
    q8 = helpers.quantize_prefill_case(case, oracle, "fp8", 0.5)
    assert q8["k_cache"].dtype == torch.uint8 and q8["k_cache"].shape[-1] == 16
    kd, vd = oracle.dequant_prefix_caches(q8["k_cache"], q8["v_cache"], torch.bfloat16, 0.5, "fp8")
    assert kd.shape == case["k_cache"].shape and vd.shape == case["v_cache"].shape
    for orig, deq in ((case["k_cache"], kd), (case["v_cache"], vd)):
        tmp = torch.empty(orig.shape, dtype=torch.uint8)
        oracle.convert_fp8(tmp, orig.contiguous(), 0.5, "fp8")
        want = torch.empty_like(orig)
        oracle.convert_fp8(want, tmp, 0.5, "fp8")
        assert torch.equal(want.view(torch.int16), deq.view(torch.int16))
    # and the oracle accepts the uint8 caches directly
    out = torch.empty_like(case["query"])
    oracle.context_attention_fwd(case["query"], case["k"], case["v"], out, q8["k_cache"], q8["v_cache"],
                                 case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                                 case["max_input_len"], None, None, "fp8", 0.5)
    ref = torch.empty_like(case["query"])
    oracle.context_attention_fwd(case["query"], case["k"], case["v"], ref, kd, vd, case["block_table"],
                                 case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"])
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))
