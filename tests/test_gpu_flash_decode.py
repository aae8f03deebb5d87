"""GPU parity of `paged_attention_flash` — decode over the flash-attention cache layout
`[num_blocks, block_size, num_kv_heads, head_size]` (reference: flash_attn_with_kvcache as called in
vllm/attention/backends/flash_attn.py:354-367) — and of the backend built on it.

Checker: the CPU oracle's paged attention on the SAME cache contents re-laid-out into the reference's paged layout
(a pure permutation, done here in torch), plus a dense fp32 torch reference; tolerance atol 1e-3 / rtol 1e-2."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _to_paged(kf, vf):
    """flash [nb, B, H, D] -> paged K [nb, H, D/8, B, 8], V [nb, H, D, B] (16-bit elements: x = 8)."""
    nb, B, H, D = kf.shape
    k = kf.view(nb, B, H, D // 8, 8).permute(0, 2, 3, 1, 4).contiguous()
    v = vf.permute(0, 2, 3, 1).contiguous()
    return k, v


def _case(S, Hq, Hkv, D, B, lens, dtype, seed, alibi=False, distinct=False):
    g = torch.Generator().manual_seed(seed)
    L = max(lens)
    nblk = (L + B - 1) // B
    NB = max(S * nblk + 3, 64) if distinct else 97
    scale = float(D ** -0.5)
    kv = torch.empty(2, NB, B, Hkv, D).uniform_(-scale, scale, generator=g).to(dtype)
    q = torch.empty(S, Hq, D).uniform_(-scale, scale, generator=g).to(dtype)
    if distinct:
        bt = torch.randperm(NB, generator=g)[:S * nblk].view(S, nblk).to(torch.int32)
    else:
        bt = torch.randint(0, NB, (S, nblk), generator=g, dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    al = torch.randn(Hq, generator=g, dtype=torch.float32) if alibi else None
    return q, kv, bt, sl, al, scale, L


def _run(ops, q, kv, bt, sl, al, scale, L):
    kvd = kv.to(DEV)
    out = torch.empty_like(q, device=DEV)
    ops.paged_attention_flash(out, q.to(DEV), kvd[0], kvd[1], scale, bt.to(DEV), sl.to(DEV), L,
                              al.to(DEV) if al is not None else None)
    assert ops.last_attention_path() == "mma_flash"
    return out.cpu()


def _oracle(oracle, q, kv, bt, sl, al, scale, L, Hkv, B):
    kp, vp = _to_paged(kv[0], kv[1])
    want = torch.empty_like(q)
    oracle.paged_attention_v1(want, q, kp, vp, Hkv, scale, bt, sl, B, L, al, "auto", 1.0)
    return want


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("head_size", [32, 64, 96, 128, 160, 192, 224, 256])
@pytest.mark.parametrize("heads", [(8, 8), (32, 8), (16, 1), (24, 1)])
@pytest.mark.parametrize("alibi", [False, True])
@torch.inference_mode()
def test_flash_decode_vs_oracle(ops, oracle, dtype, block_size, head_size, heads, alibi):
    Hq, Hkv = heads
    rng = random.Random(head_size * 7 + block_size + Hq)
    lens = [rng.randint(1, 700) for _ in range(5)] + [1, block_size, 1500]
    q, kv, bt, sl, al, scale, L = _case(len(lens), Hq, Hkv, head_size, block_size, lens, dtype, Hq + head_size, alibi)
    got = _run(ops, q, kv, bt, sl, al, scale, L)
    want = _oracle(oracle, q, kv, bt, sl, al, scale, L, Hkv, block_size)
    torch.testing.assert_close(got.float(), want.float(), atol=1e-3, rtol=1e-2)


@pytest.mark.parametrize("lens", [[40000], [9000, 3, 17000], [0, 5, 0, 2048]])
@torch.inference_mode()
def test_flash_decode_long_split_and_empty(ops, oracle, lens):
    """Long contexts split across CTAs (fused reduce through the per-device workspace); empty contexts give zeros."""
    Hq, Hkv, D, B = 16, 2, 128, 16
    q, kv, bt, sl, al, scale, L = _case(len(lens), Hq, Hkv, D, B, lens, torch.bfloat16, 3, distinct=False)
    got = _run(ops, q, kv, bt, sl, al, scale, L)
    if L > 4096:
        assert ops.last_attention_partitions() > 1
    want = _oracle(oracle, q, kv, bt, sl, al, scale, L, Hkv, B)
    torch.testing.assert_close(got.float(), want.float(), atol=1e-3, rtol=1e-2)
    for i, n in enumerate(lens):
        if n == 0:
            assert torch.count_nonzero(got[i]) == 0


@torch.inference_mode()
def test_flash_decode_stale_nan_tail_and_strided_query(ops, oracle):
    """NaN bits behind seq_len in the last block never reach the output; the query may be a slice of a fused qkv."""
    Hq, Hkv, D, B = 8, 2, 128, 16
    lens = [5, 33, 100]
    q, kv, bt, sl, al, scale, L = _case(3, Hq, Hkv, D, B, lens, torch.half, 11, distinct=True)
    kv_nan = kv.clone()
    for i, n in enumerate(lens):
        blk, off = int(bt[i, (n - 1) // B]), (n - 1) % B + 1
        kv_nan[:, blk, off:] = float("nan")
    want = _oracle(oracle, q, kv, bt, sl, al, scale, L, Hkv, B)
    qkv = torch.zeros(3, Hq + 4, D, dtype=torch.half, device=DEV)
    qkv[:, :Hq] = q.to(DEV)
    qs = qkv[:, :Hq]   # row stride (Hq + 4) * D
    out = torch.empty(3, Hq, D, dtype=torch.half, device=DEV)
    kvd = kv_nan.to(DEV)
    ops.paged_attention_flash(out, qs, kvd[0], kvd[1], scale, bt.to(DEV), sl.to(DEV), L, None)
    assert torch.isfinite(out).all()
    torch.testing.assert_close(out.cpu().float(), want.float(), atol=1e-3, rtol=1e-2)


@torch.inference_mode()
def test_flash_decode_error_classes(ops):
    q = torch.zeros(1, 8, 48, dtype=torch.half, device=DEV)
    kv = torch.zeros(2, 4, 16, 2, 48, dtype=torch.half, device=DEV)
    bt = torch.zeros(1, 4, dtype=torch.int32, device=DEV)
    sl = torch.ones(1, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size: 48"):
        ops.paged_attention_flash(torch.empty_like(q), q, kv[0], kv[1], 1.0, bt, sl, 64, None)
    q = torch.zeros(1, 8, 64, dtype=torch.half, device=DEV)
    kv = torch.zeros(2, 4, 8, 2, 64, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported block size: 8"):
        ops.paged_attention_flash(torch.empty_like(q), q, kv[0], kv[1], 1.0, bt, sl, 64, None)
    kv = torch.zeros(2, 4, 16, 2, 64, dtype=torch.float, device=DEV)
    with pytest.raises(RuntimeError, match="query dtype"):
        ops.paged_attention_flash(torch.empty_like(q), q, kv[0], kv[1], 1.0, bt, sl, 64, None)


@torch.inference_mode()
def test_flash_backend_prefill_then_decode(ops, oracle):
    """B200FlashAttentionBackend end to end: a prompt without prefix (cache written by reshape_and_cache_flash, causal
    attention over the new tokens), then a decode step over the flash-layout cache; checked against dense fp32."""
    from vllm_legacy_b200.attention.backends.b200_flash import B200FlashAttentionBackend as Backend
    Hq, Hkv, D, B, NB = 8, 2, 128, 16, 32
    dt = torch.bfloat16
    impl = Backend.get_impl_cls()(Hq, D, D ** -0.5, Hkv, None, None, "auto")
    kv_cache = torch.zeros(Backend.get_kv_cache_shape(NB, B, Hkv, D), dtype=dt, device=DEV)
    torch.manual_seed(0)
    lens = [37, 80]
    T = sum(lens)
    q = torch.randn(T, Hq * D, device=DEV).mul(0.3).to(dt)
    k = torch.randn(T, Hkv * D, device=DEV).mul(0.3).to(dt)
    v = torch.randn(T, Hkv * D, device=DEV).mul(0.3).to(dt)
    bt = torch.tensor([[3, 9, 1, 0, 0, 0], [5, 2, 7, 11, 4, 6]], dtype=torch.int32, device=DEV)
    slots = torch.cat([torch.stack([bt[i, t // B].long() * B + t % B for t in range(n)]) for i, n in enumerate(lens)])
    qsl = torch.tensor([0, 37, 117], dtype=torch.int32, device=DEV)
    meta = Backend.make_metadata(num_prefills=2, num_prefill_tokens=T, num_decode_tokens=0, slot_mapping=slots,
                                 seq_lens=lens, seq_lens_tensor=torch.tensor(lens, dtype=torch.int32, device=DEV),
                                 max_query_len=80, max_prefill_seq_len=80, max_decode_seq_len=0, query_start_loc=qsl,
                                 seq_start_loc=qsl, context_lens_tensor=torch.zeros(2, dtype=torch.int32, device=DEV),
                                 block_tables=torch.empty(0, dtype=torch.int32, device=DEV), use_cuda_graph=False)
    out = impl.forward(q, k, v, kv_cache, meta, 1.0).float().cpu()

    def dense(qi, ki, vi, causal):
        qh = qi.float().view(-1, Hq, D).transpose(0, 1)
        kh = ki.float().view(-1, Hkv, D).repeat_interleave(Hq // Hkv, 1).transpose(0, 1)
        vh = vi.float().view(-1, Hkv, D).repeat_interleave(Hq // Hkv, 1).transpose(0, 1)
        s = qh @ kh.transpose(1, 2) * D ** -0.5
        if causal:
            n, m = s.shape[1:]
            s = s.masked_fill(torch.ones(n, m, dtype=torch.bool, device=s.device).triu(m - n + 1), float("-inf"))
        return (s.softmax(-1) @ vh).transpose(0, 1).reshape(-1, Hq * D)

    o = 0
    for n in lens:
        want = dense(q[o:o + n], k[o:o + n], v[o:o + n], True).cpu()
        torch.testing.assert_close(out[o:o + n], want, atol=8e-3, rtol=2e-2)
        o += n
    # decode step: one new token per sequence
    qd = torch.randn(2, Hq * D, device=DEV).mul(0.3).to(dt)
    kd = torch.randn(2, Hkv * D, device=DEV).mul(0.3).to(dt)
    vd = torch.randn(2, Hkv * D, device=DEV).mul(0.3).to(dt)
    new_lens = [n + 1 for n in lens]
    dslots = torch.stack([bt[i, lens[i] // B].long() * B + lens[i] % B for i in range(2)])
    dmeta = Backend.make_metadata(num_prefills=0, num_prefill_tokens=0, num_decode_tokens=2, slot_mapping=dslots,
                                  seq_lens_tensor=torch.tensor(new_lens, dtype=torch.int32, device=DEV),
                                  max_decode_seq_len=max(new_lens), block_tables=bt)
    od = impl.forward(qd, kd, vd, kv_cache, dmeta, 1.0).float().cpu()
    assert ops.last_attention_path() == "mma_flash"
    o = 0
    for i, n in enumerate(lens):
        kk = torch.cat([k[o:o + n], kd[i:i + 1]])
        vv = torch.cat([v[o:o + n], vd[i:i + 1]])
        want = dense(qd[i:i + 1], kk, vv, False).cpu()
        torch.testing.assert_close(od[i:i + 1], want, atol=2e-3, rtol=1e-2)
        o += n
