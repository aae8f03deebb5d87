"""GPU parity of rotary_embedding / batched_rotary_embedding and the fused rotary + reshape_and_cache op
(SURVEY 8f rank 2).  Reference: csrc/pos_encoding_kernels.cu:10-201, tests/kernels/test_pos_encoding.py.
Checker: the CPU oracle (bit-exact) and, when built, the reference's own CUDA kernels (bit-exact)."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
DTYPES = [torch.half, torch.bfloat16, torch.float]


def _bits(t):
    return t.contiguous().view(torch.uint8)


def _cos_sin_cache(max_position, rot_dim, base, dtype):
    # vllm/model_executor/layers/rotary_embedding.py:79-103 (_compute_inv_freq / _compute_cos_sin_cache)
    inv_freq = 1.0 / (base ** (torch.arange(0, rot_dim, 2, dtype=torch.float) / rot_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("batch_seq", [(1, 11), (5, 11), (2, 1024)])
@pytest.mark.parametrize("num_heads", [7, 17])
@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("rotary_dim", [None, 32])
@pytest.mark.parametrize("dtype", DTYPES)
@torch.inference_mode()
def test_rotary_embedding(ops, oracle, ref_cuda_optional, is_neox, batch_seq, num_heads, head_size, rotary_dim, dtype):
    """tests/kernels/test_pos_encoding.py:24-79 grid; bit-exact vs the oracle and the reference CUDA kernel."""
    batch, seq_len = batch_seq
    rot = head_size if rotary_dim is None else rotary_dim
    torch.manual_seed(0)
    max_position = 8192
    cache = _cos_sin_cache(max_position, rot, 10000, dtype)
    positions = torch.randint(0, max_position, (batch, seq_len))
    query = torch.randn(batch, seq_len, num_heads * head_size, dtype=dtype)
    key = torch.randn_like(query)
    q_d, k_d = query.to(DEV), key.to(DEV)
    ops.rotary_embedding(positions.to(DEV), q_d, k_d, head_size, cache.to(DEV), is_neox)
    if ref_cuda_optional is not None:
        q_r, k_r = query.to(DEV), key.to(DEV)
        ref_cuda_optional.rotary_embedding(positions.to(DEV), q_r, k_r, head_size, cache.to(DEV), is_neox)
        assert torch.equal(_bits(q_d), _bits(q_r)) and torch.equal(_bits(k_d), _bits(k_r))
    oracle.rotary_embedding(positions, query, key, head_size, cache, is_neox)
    assert torch.equal(_bits(q_d.cpu()), _bits(query))
    assert torch.equal(_bits(k_d.cpu()), _bits(key))


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("head_size,rot", [(128, 128), (80, 32), (64, 20)])
@torch.inference_mode()
def test_rotary_embedding_gqa_strided_and_misaligned(ops, oracle, ref_cuda_optional, is_neox, dtype, head_size, rot):
    """q/k as views of one fused qkv projection (token stride = (Hq+2Hkv)*D, as llama.py:164-166 produces),
    GQA head counts, and a misaligned / odd-rot_dim case that takes the scalar kernel."""
    torch.manual_seed(1)
    T, Hq, Hkv = 37, 8, 2
    cache = _cos_sin_cache(512, rot, 10000, dtype)
    positions = torch.randint(0, 512, (T,))
    qkv = torch.randn(T, (Hq + 2 * Hkv) * head_size + 3, dtype=dtype)
    for off in (0, 1):  # off=1: every row misaligned for 16-byte vectors
        q, k = qkv[:, off:off + Hq * head_size], qkv[:, off + Hq * head_size: off + (Hq + Hkv) * head_size]
        qkv_d = qkv.to(DEV)
        q_d = qkv_d[:, off:off + Hq * head_size]
        k_d = qkv_d[:, off + Hq * head_size: off + (Hq + Hkv) * head_size]
        ops.rotary_embedding(positions.to(DEV), q_d, k_d, head_size, cache.to(DEV), is_neox)
        if ref_cuda_optional is not None:
            qkv_r = qkv.to(DEV)
            ref_cuda_optional.rotary_embedding(positions.to(DEV), qkv_r[:, off:off + Hq * head_size],
                                               qkv_r[:, off + Hq * head_size: off + (Hq + Hkv) * head_size],
                                               head_size, cache.to(DEV), is_neox)
            assert torch.equal(_bits(qkv_d), _bits(qkv_r))
        oracle.rotary_embedding(positions, q, k, head_size, cache, is_neox)
        assert torch.equal(_bits(qkv_d.cpu()), _bits(qkv))  # includes the untouched V / padding columns


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", DTYPES)
@torch.inference_mode()
def test_batched_rotary_embedding(ops, oracle, ref_cuda_optional, is_neox, dtype):
    """tests/kernels/test_pos_encoding.py:82-207: several cos/sin tables concatenated, per-token row offsets."""
    torch.manual_seed(2)
    T, H, D, rot = 53, 7, 128, 64
    tables = [_cos_sin_cache(256, rot, b, dtype) for b in (10000, 500000, 37)]
    cache = torch.cat(tables, dim=0)
    positions = torch.randint(0, 256, (T,))
    offsets = torch.tensor([256 * random.Random(i).randrange(3) for i in range(T)], dtype=torch.long)
    query = torch.randn(T, H * D, dtype=dtype)
    key = torch.randn(T, H * D, dtype=dtype)
    q_d, k_d = query.to(DEV), key.to(DEV)
    ops.batched_rotary_embedding(positions.to(DEV), q_d, k_d, D, cache.to(DEV), is_neox, rot, offsets.to(DEV))
    if ref_cuda_optional is not None:
        q_r, k_r = query.to(DEV), key.to(DEV)
        ref_cuda_optional.batched_rotary_embedding(positions.to(DEV), q_r, k_r, D, cache.to(DEV), is_neox, rot,
                                                   offsets.to(DEV))
        assert torch.equal(_bits(q_d), _bits(q_r)) and torch.equal(_bits(k_d), _bits(k_r))
    oracle.rotary_embedding(positions, query, key, D, cache, is_neox, rot, offsets)
    assert torch.equal(_bits(q_d.cpu()), _bits(query)) and torch.equal(_bits(k_d.cpu()), _bits(key))


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype,kv_scale", [("auto", 1.0), ("fp8", 0.37), ("fp8_e5m2", 1.0)])
@pytest.mark.parametrize("head_size,rot,block_size", [(128, 128, 16), (128, 128, 32), (80, 32, 16), (256, 64, 8),
                                                      (64, 20, 16)])
@torch.inference_mode()
def test_fused_rotary_reshape_and_cache(ops, oracle, is_neox, dtype, kv_cache_dtype, kv_scale, head_size, rot,
                                        block_size):
    """The fused launch == rotary_embedding followed by reshape_and_cache (the reference's two launches in
    llama.py:164-171 / paged_attn.py:62-79), bit for bit: q, k in place, K/V caches, padding slot untouched."""
    torch.manual_seed(3)
    random.seed(3)
    T, Hq, Hkv, NB = 42, 8, 2, 16
    cache = _cos_sin_cache(1024, rot, 10000, dtype)
    positions = torch.randint(0, 1024, (T,))
    qkv = torch.randn(T, (Hq + 2 * Hkv) * head_size, dtype=dtype)
    q, k, v = qkv.split([Hq * head_size, Hkv * head_size, Hkv * head_size], dim=-1)
    slots = torch.tensor(random.sample(range(NB * block_size), T), dtype=torch.long)
    slots[7] = -1
    cdt = dtype if kv_cache_dtype == "auto" else torch.uint8
    x = 16 // torch.tensor([], dtype=cdt).element_size()
    kc = torch.randint(0, 255, (NB, Hkv, head_size // x, block_size, x * torch.tensor([], dtype=cdt).element_size()),
                       dtype=torch.uint8).view(cdt).view(NB, Hkv, head_size // x, block_size, x)
    vc = torch.randint(0, 255, (NB, Hkv, head_size, block_size * torch.tensor([], dtype=cdt).element_size()),
                       dtype=torch.uint8).view(cdt).view(NB, Hkv, head_size, block_size)
    qkv_d, kc_d, vc_d = qkv.to(DEV), kc.to(DEV), vc.to(DEV)
    q_d, k_d, v_d = qkv_d.split([Hq * head_size, Hkv * head_size, Hkv * head_size], dim=-1)
    ops.rotary_embedding_reshape_and_cache(positions.to(DEV), q_d, k_d, v_d, head_size, cache.to(DEV), is_neox,
                                           kc_d, vc_d, slots.to(DEV), kv_cache_dtype, kv_scale)
    # oracle: the two reference steps
    oracle.rotary_embedding(positions, q, k, head_size, cache, is_neox)
    oracle.reshape_and_cache(k.view(T, Hkv, head_size), v.view(T, Hkv, head_size), kc, vc, slots, kv_cache_dtype,
                             kv_scale)
    assert torch.equal(_bits(qkv_d.cpu()), _bits(qkv))
    assert torch.equal(_bits(kc_d.cpu()), _bits(kc))
    assert torch.equal(_bits(vc_d.cpu()), _bits(vc))
    # and the product's own two-launch sequence gives the same bits
    qkv3 = torch.randn(T, (Hq + 2 * Hkv) * head_size, dtype=dtype).to(DEV)
    q3, k3, v3 = qkv3.split([Hq * head_size, Hkv * head_size, Hkv * head_size], dim=-1)
    qkv4 = qkv3.clone()
    q4, k4, v4 = qkv4.split([Hq * head_size, Hkv * head_size, Hkv * head_size], dim=-1)
    kc3, vc3, kc4, vc4 = kc_d.clone(), vc_d.clone(), kc_d.clone(), vc_d.clone()
    ops.rotary_embedding_reshape_and_cache(positions.to(DEV), q3, k3, v3, head_size, cache.to(DEV), is_neox, kc3, vc3,
                                           slots.to(DEV), kv_cache_dtype, kv_scale)
    ops.rotary_embedding(positions.to(DEV), q4, k4, head_size, cache.to(DEV), is_neox)
    ops.reshape_and_cache(k4.view(T, Hkv, head_size), v4.view(T, Hkv, head_size), kc4, vc4, slots.to(DEV),
                          kv_cache_dtype, kv_scale)
    assert torch.equal(_bits(qkv3), _bits(qkv4)) and torch.equal(_bits(kc3), _bits(kc4))
    assert torch.equal(_bits(vc3), _bits(vc4))


@torch.inference_mode()
def test_fused_rotary_cuda_graph_and_errors(ops):
    """The fused op is CUDA-graph capturable (decode step, model_runner.py:1097-1218) and rejects bad input."""
    torch.manual_seed(4)
    T, Hq, Hkv, D, B, NB = 16, 32, 8, 128, 16, 8
    dt = torch.bfloat16
    cache = _cos_sin_cache(4096, D, 10000, dt).to(DEV)
    pos = torch.randint(0, 4096, (T,), device=DEV)
    qkv0 = torch.randn(T, (Hq + 2 * Hkv) * D, dtype=dt, device=DEV)
    slots = torch.randperm(NB * B, device=DEV)[:T]
    kc = torch.zeros(NB, Hkv, D // 8, B, 8, dtype=dt, device=DEV)
    vc = torch.zeros(NB, Hkv, D, B, dtype=dt, device=DEV)

    def step(buf, kcx, vcx):
        q, k, v = buf.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        ops.rotary_embedding_reshape_and_cache(pos, q, k, v, D, cache, True, kcx, vcx, slots, "auto", 1.0)

    eager, kce, vce = qkv0.clone(), kc.clone(), vc.clone()
    step(eager, kce, vce)
    static = qkv0.clone()
    step(static.clone(), kc.clone(), vc.clone())  # warm-up outside capture
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            step(static, kc, vc)
    static.copy_(qkv0)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(_bits(static), _bits(eager)) and torch.equal(_bits(kc), _bits(kce))
    assert torch.equal(_bits(vc), _bits(vce))
    with pytest.raises(RuntimeError):
        q, k, v = qkv0.cpu().split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        ops.rotary_embedding(pos.cpu(), q, k, D, cache.cpu(), True)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        step2 = qkv0.clone().split([Hq * D, Hkv * D, Hkv * D], dim=-1)
        ops.rotary_embedding_reshape_and_cache(pos, *step2, D, cache, True, kc, vc, slots, "int4", 1.0)
