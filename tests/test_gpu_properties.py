"""Size-independent properties at BASELINE.json's FULL sizes (where no CPU oracle finishes in time):

  attention : physical-block permutation invariance (bit-exact), linearity in V (x2 is exact in binary
              floating point -> bit-exact), convexity (V == c -> out == c), v1 == v2 within tolerance, and
              spot-checked sequences against the torch restatement of the reference's test reference;
  cache ops : reshape_and_cache -> gather round trip, fp8 quantise idempotence, copy_blocks idempotence,
              swap out -> swap in round trip (all bit-exact).
Configs: 2 (Llama-3-8B 32q/8kv, 256 x 4096, bf16), 3 per-GPU shard at TP8 (8q/1kv, 256 x 8192),
4 (fp8 KV, 128 x 32768), 5 (ragged 128..128k, batch 64, block 16/32)."""
import math
import random

import pytest
import torch

from tests import helpers

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _setup(ops, S, Hq, Hkv, D, B, lens, dtype, kvd, seed=0):
    torch.manual_seed(seed)
    scale = float(D ** -0.5)
    L = max(lens)
    nblk_per_seq = [(l + B - 1) // B for l in lens]
    NB = sum(nblk_per_seq) + 8
    esz = 1 if kvd == "fp8" else 2
    x = 16 // esz
    cdt = torch.uint8 if kvd == "fp8" else dtype
    kc = torch.empty(NB, Hkv, D // x, B, x, dtype=cdt, device=DEV)
    vc = torch.empty(NB, Hkv, D, B, dtype=cdt, device=DEV)
    if kvd == "fp8":
        # random e4m3 codes with |value| < 2 and no NaN
        kc.random_(0, 256).bitwise_and_(0xB7)
        vc.random_(0, 256).bitwise_and_(0xB7)
    else:
        kc.uniform_(-scale, scale)
        vc.uniform_(-scale, scale)
    q = torch.empty(S, Hq, D, dtype=dtype, device=DEV).uniform_(-scale, scale)
    perm = torch.randperm(NB - 8, device=DEV).to(torch.int32)
    bt = torch.zeros(S, (L + B - 1) // B, dtype=torch.int32, device=DEV)
    o = 0
    for i, n in enumerate(nblk_per_seq):
        bt[i, :n] = perm[o:o + n]
        o += n
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    return q, kc, vc, bt, sl, scale, L


def _attn(ops, version, q, kc, vc, Hkv, scale, bt, sl, B, L, kvd, kv_scale=1.0):
    S, Hq, D = q.shape
    out = torch.empty_like(q)
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, kv_scale)
    else:
        P = (L + 511) // 512
        tmp = torch.empty(S, Hq, P, D, dtype=q.dtype, device=DEV)
        es = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, kvd, kv_scale)
    return out


def _spot_check(ops, out, q, kc, vc, bt, sl, scale, Hkv, kvd, rows, atol, kv_scale=1.0):
    S, Hq, D = q.shape
    rows = torch.tensor(rows, device=DEV)
    bt_rows = bt[rows]
    if kvd == "fp8":
        # dequantise only the blocks these rows touch, compacted, and remap the table onto them
        used, inv = torch.unique(bt_rows.flatten().long(), return_inverse=True)
        kd = torch.empty(kc[used].shape, dtype=q.dtype, device=DEV)
        vd = torch.empty(vc[used].shape, dtype=q.dtype, device=DEV)
        ops.convert_fp8(kd, kc[used].contiguous(), kv_scale)
        ops.convert_fp8(vd, vc[used].contiguous(), kv_scale)
        bt_rows = inv.view(bt_rows.shape).to(torch.int32)
    else:
        kd, vd = kc, vc
    want = torch.empty(len(rows), Hq, D, dtype=q.dtype, device=DEV)
    helpers.ref_single_query_cached_kv_attention(want, q[rows], Hq // Hkv, kd, vd, bt_rows, sl[rows], scale, None)
    got = out[rows]
    assert torch.allclose(got.float(), want.float(), atol=atol, rtol=1e-2), (got.float() - want.float()).abs().max()


CONFIGS = {
    # name: (S, Hq, Hkv, D, B, lens, dtype, kv)
    "config2_llama3_8b_256x4096": (256, 32, 8, 128, 16, [4096] * 256, torch.bfloat16, "auto"),
    "config3_tp8_shard_256x8192": (256, 8, 1, 128, 16, [8192] * 256, torch.bfloat16, "auto"),
    "config3_tp2_shard_64x8192": (64, 32, 4, 128, 16, [8192] * 64, torch.bfloat16, "auto"),
    "config4_fp8_128x32768": (128, 32, 8, 128, 16, [32768] * 128, torch.bfloat16, "fp8"),
}


@pytest.mark.parametrize("name", list(CONFIGS))
@torch.inference_mode()
def test_attention_properties_full_size(ops, name):
    S, Hq, Hkv, D, B, lens, dtype, kvd = CONFIGS[name]
    q, kc, vc, bt, sl, scale, L = _setup(ops, S, Hq, Hkv, D, B, lens, dtype, kvd)
    out1 = _attn(ops, "v1", q, kc, vc, Hkv, scale, bt, sl, B, L, kvd)
    assert ops.last_attention_path() == "mma_tma"
    out2 = _attn(ops, "v2", q, kc, vc, Hkv, scale, bt, sl, B, L, kvd)
    assert torch.isfinite(out1).all()
    # v1 and v2 (different split of the context) agree
    assert torch.allclose(out1.float(), out2.float(), atol=1e-3, rtol=1e-2)
    # spot-check rows against the torch reference
    atol = 1e-3 if kvd == "auto" else 1e-2
    _spot_check(ops, out1, q, kc, vc, bt, sl, scale, Hkv, kvd, [0, S // 2, S - 1], atol)
    # (1) paged indirection: moving every block to another physical slot (and updating the table) changes nothing
    NB = kc.shape[0]
    sigma = torch.randperm(NB, device=DEV)
    kc2 = torch.empty_like(kc)
    vc2 = torch.empty_like(vc)
    kc2[sigma] = kc
    vc2[sigma] = vc
    bt2 = sigma[bt.long()].to(torch.int32)
    out_p = _attn(ops, "v1", q, kc2, vc2, Hkv, scale, bt2, sl, B, L, kvd)
    assert torch.equal(out_p.view(torch.int16), out1.view(torch.int16))
    del kc2, vc2
    # (2) linearity in V: doubling V doubles the output exactly (power-of-two scaling commutes with rounding)
    if kvd == "auto":
        out_d = _attn(ops, "v1", q, kc, vc * 2, Hkv, scale, bt, sl, B, L, kvd)
        assert torch.equal((out1 * 2).view(torch.int16), out_d.view(torch.int16))
    else:
        out_d = _attn(ops, "v1", q, kc, vc, Hkv, scale, bt, sl, B, L, kvd, kv_scale=2.0)
        # kv_scale scales K too; only check it stays finite and differs
        assert torch.isfinite(out_d).all()
    # (3) convexity: a constant V comes back (sum of softmax weights == 1 up to the reference's 1e-6 epsilon)
    if kvd == "auto":
        vconst = torch.full_like(vc, 0.5)
        out_c = _attn(ops, "v2", q, kc, vconst, Hkv, scale, bt, sl, B, L, kvd)
        assert torch.allclose(out_c.float(), torch.full_like(out_c, 0.5).float(), atol=4e-3, rtol=0)


@pytest.mark.parametrize("block_size", [16, 32])
@torch.inference_mode()
def test_ragged_long_context_mix(ops, block_size):
    """Config 5: batch 64, lengths log-uniform in [128, 131072] with both extremes present."""
    rng = random.Random(0)
    lens = [int(round(math.exp(rng.uniform(math.log(128), math.log(131072))))) for _ in range(64)]
    lens[0], lens[1] = 131072, 128
    S, Hq, Hkv, D = 64, 32, 8, 128
    q, kc, vc, bt, sl, scale, L = _setup(ops, S, Hq, Hkv, D, block_size, lens, torch.bfloat16, "auto", seed=5)
    out2 = _attn(ops, "v2", q, kc, vc, Hkv, scale, bt, sl, block_size, L, "auto")
    assert ops.last_attention_partitions() > 1
    out1 = _attn(ops, "v1", q, kc, vc, Hkv, scale, bt, sl, block_size, L, "auto")
    assert torch.allclose(out1.float(), out2.float(), atol=1e-3, rtol=1e-2)
    order = sorted(range(S), key=lambda i: lens[i])
    _spot_check(ops, out2, q, kc, vc, bt, sl, scale, Hkv, "auto", [0, 1, order[S // 2], order[S // 4]], 1e-3)
    # prefix sharing / beam fork on the same cache: copy the first 8 blocks of sequence 2 to fresh blocks,
    # point sequence 3's table at the copies -> identical logits for a query attending to that prefix only
    NB = kc.shape[0]
    src = bt[2, :8].long()
    dst = torch.arange(NB - 8, NB, device=DEV)
    mapping = torch.stack([src, dst], dim=1).contiguous()
    ops.copy_blocks([kc], [vc], mapping)
    bt_f = bt.clone()
    bt_f[3, :8] = dst.to(torch.int32)
    bt_f[2, :8] = src.to(torch.int32)
    sl_f = sl.clone()
    sl_f[2] = sl_f[3] = 8 * block_size
    q_f = q.clone()
    q_f[3] = q_f[2]
    o = _attn(ops, "v1", q_f, kc, vc, Hkv, scale, bt_f, sl_f, block_size, L, "auto")
    # The persistent kernel cuts the flattened block list at positions that depend on where a sequence sits in the
    # batch, so two rows with identical content may group their fp32 partial sums differently: equal up to the
    # rounding of the bf16 output (1 ulp at |o| < 0.125 is 4.9e-4), and bit-identical from launch to launch.
    assert torch.allclose(o[2].float(), o[3].float(), atol=5e-4, rtol=0)
    o_again = _attn(ops, "v1", q_f, kc, vc, Hkv, scale, bt_f, sl_f, block_size, L, "auto")
    assert torch.equal(o.view(torch.int16), o_again.view(torch.int16))


@torch.inference_mode()
def test_reshape_and_cache_round_trip_full_size(ops):
    """Config 4 write side: T = 32768 tokens (a prefill chunk) and T = 128 (one decode step), bf16 -> fp8 and auto.
    Reading the slots back through the layout's index math returns exactly what was written; quantising an
    already-quantised tensor is idempotent."""
    torch.manual_seed(0)
    H, D, B = 8, 128, 16
    for T in (128, 32768):
        NB = (T + B - 1) // B + 64
        key = torch.randn(T, H, D, dtype=torch.bfloat16, device=DEV)
        value = torch.randn(T, H, D, dtype=torch.bfloat16, device=DEV)
        slots = torch.randperm(NB * B, device=DEV)[:T]
        b, o = (slots // B), (slots % B)
        # auto
        kc = torch.zeros(NB, H, D // 8, B, 8, dtype=torch.bfloat16, device=DEV)
        vc = torch.zeros(NB, H, D, B, dtype=torch.bfloat16, device=DEV)
        ops.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
        assert torch.equal(kc[b, :, :, o, :].reshape(T, H, D), key)
        assert torch.equal(vc[b, :, :, o], value)
        assert kc.count_nonzero() <= T * H * D  # nothing else was touched
        # fp8
        kc8 = torch.zeros(NB, H, D // 16, B, 16, dtype=torch.uint8, device=DEV)
        vc8 = torch.zeros(NB, H, D, B, dtype=torch.uint8, device=DEV)
        ops.reshape_and_cache(key, value, kc8, vc8, slots, "fp8", 0.05)
        want_k = torch.empty(T, H, D, dtype=torch.uint8, device=DEV)
        ops.convert_fp8(want_k, key, 0.05, "fp8")
        assert torch.equal(kc8[b, :, :, o, :].reshape(T, H, D), want_k)
        deq = torch.empty(T, H, D, dtype=torch.bfloat16, device=DEV)
        ops.convert_fp8(deq, want_k, 0.05, "fp8")
        again = torch.empty_like(want_k)
        ops.convert_fp8(again, deq, 0.05, "fp8")
        # re-quantising the dequantised values gives the same codes (|x| well inside the range)
        assert (again == want_k).float().mean() > 0.999


@torch.inference_mode()
def test_copy_and_swap_round_trips_full_size(ops):
    """Config 5 maintenance ops: copy_blocks with 256 pairs x 32 layers (idempotent, exact) and swap_blocks
    256 pairs out to pinned host memory and back in (exact round trip)."""
    torch.manual_seed(0)
    L, NB, H, D, B = 32, 1024, 8, 128, 16
    kcs = [torch.randn(NB, H, D // 8, B, 8, device=DEV).to(torch.bfloat16) for _ in range(L)]
    vcs = [torch.randn(NB, H, D, B, device=DEV).to(torch.bfloat16) for _ in range(L)]
    perm = torch.randperm(NB)
    src, dst = perm[:256], perm[256:512]
    mapping = torch.stack([src, dst], dim=1).to(DEV)
    before = [t[src.to(DEV)].clone() for t in kcs]
    ops.copy_blocks(kcs, vcs, mapping)
    for t, bsrc in zip(kcs, before):
        assert torch.equal(t[dst.to(DEV)], bsrc)
    snapshot = kcs[7].clone()
    ops.copy_blocks(kcs, vcs, mapping)  # idempotent
    assert torch.equal(kcs[7], snapshot)
    # swap out -> wipe -> swap in
    host = torch.zeros(512, H, D // 8, B, 8, dtype=torch.bfloat16).pin_memory()
    out_map = torch.stack([src, torch.arange(256)], dim=1)
    in_map = torch.stack([torch.arange(256), src], dim=1)
    ops.swap_blocks(kcs[0], host, out_map)
    torch.cuda.synchronize()
    saved = kcs[0][src.to(DEV)].clone()
    kcs[0][src.to(DEV)] = 0
    ops.swap_blocks(host, kcs[0], in_map)
    torch.cuda.synchronize()
    assert torch.equal(kcs[0][src.to(DEV)], saved)
