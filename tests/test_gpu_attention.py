"""GPU parity of paged_attention_v1 / v2 (+reduce).

 * test_paged_attention            — the reference's own grid and sizes (tests/kernels/test_attention.py:
                                      114-274) against the torch restatement of its reference
 * test_paged_attention_vs_oracle  — small cases against the pinned CPU oracle (all block sizes, e5m2, kv_scale)
 * test_blocksparse_*              — tests/kernels/test_blocksparse_attention.py semantics vs the oracle
 * test_*_golden                   — fixtures produced by the reference's own CPU kernels
 * test_*_vs_reference_cuda        — the reference's CUDA kernels compiled for sm_100a, same inputs
 * edge cases: empty / length-1 / CUDA-graph padded rows, stale NaN bits behind seq_len, strided queries,
   CUDA-graph capture, error classes.
Tolerances are the reference's: atol 1e-3, rtol 1e-5 (fp8 KV: atol 1e-2) — test_attention.py:265-274;
north_star: atol 1e-3 / rtol 1e-2 against the reference implementation."""
import random

import pytest
import torch

from tests import helpers

pytestmark = pytest.mark.gpu

DEV = "cuda"
FLOAT32_BYTES = 4
NUM_BLOCKS = 4321        # test_attention.py:20
PARTITION_SIZE = 512     # test_attention.py:21
DTYPES = [torch.half, torch.bfloat16, torch.float]
NUM_GEN_SEQS = [7]
NUM_HEADS = [(40, 40), (64, 8)]
HEAD_SIZES = [64, 80, 96, 112, 128, 192, 256]
BLOCK_SIZES = [16, 32]
KV_CACHE_DTYPE = ["auto", "fp8"]


def _max_seq_len(ops):
    # test_attention.py:13-17 — on B200: 232448 / 4 - 512 = 57600
    return ops.get_max_shared_memory_per_block_device_attribute(0) // FLOAT32_BYTES - 512


def _run(ops, version, out, q, kc, vc, Hkv, scale, bt, sl, B, max_len, alibi, kvd, kv_scale=1.0, bs=(0, 0, 64, 0)):
    S, Hq, D = q.shape
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, max_len, alibi, kvd, kv_scale, 0, *bs)
        return None
    P = (max_len + PARTITION_SIZE - 1) // PARTITION_SIZE
    tmp = torch.empty(S, Hq, P, D, dtype=out.dtype, device=DEV)
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, max_len, alibi, kvd, kv_scale, 0, *bs)
    return es, ml, tmp


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("num_seqs", NUM_GEN_SEQS)
@pytest.mark.parametrize("num_heads", NUM_HEADS)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
@pytest.mark.parametrize("use_alibi", [False, True])
@pytest.mark.parametrize("block_size", BLOCK_SIZES)
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", KV_CACHE_DTYPE)
@pytest.mark.parametrize("seed", [0])
@torch.inference_mode()
def test_paged_attention(ops, version, num_seqs, num_heads, head_size, use_alibi, block_size, dtype, kv_cache_dtype,
                         seed):
    """Restates tests/kernels/test_attention.py:114-274 (same grid, NUM_BLOCKS, MAX_SEQ_LEN, seeds, tolerances)."""
    random.seed(seed)
    torch.random.manual_seed(seed)
    torch.cuda.manual_seed(seed)
    MAX_SEQ_LEN = _max_seq_len(ops)
    scale = float(1.0 / (head_size ** 0.5))
    num_query_heads, num_kv_heads = num_heads
    query = torch.empty(num_seqs, num_query_heads, head_size, dtype=dtype, device=DEV)
    query.uniform_(-scale, scale)
    assert num_query_heads % num_kv_heads == 0
    num_queries_per_kv = num_query_heads // num_kv_heads
    alibi_slopes = torch.randn(num_query_heads, dtype=torch.float, device=DEV) if use_alibi else None
    seq_lens = [random.randint(1, MAX_SEQ_LEN) for _ in range(num_seqs)]
    seq_lens[-1] = MAX_SEQ_LEN
    max_seq_len = max(seq_lens)
    seq_lens = torch.tensor(seq_lens, dtype=torch.int, device=DEV)
    max_num_blocks_per_seq = (max_seq_len + block_size - 1) // block_size
    block_tables = [[random.randint(0, NUM_BLOCKS - 1) for _ in range(max_num_blocks_per_seq)] for _ in range(num_seqs)]
    block_tables = torch.tensor(block_tables, dtype=torch.int, device=DEV)
    key_caches, value_caches = helpers.create_kv_caches_with_random(NUM_BLOCKS, block_size, 1, num_kv_heads, head_size,
                                                                    kv_cache_dtype, dtype, seed, DEV, ops.convert_fp8)
    key_cache, value_cache = key_caches[0], value_caches[0]
    kv_scale = 1.0
    output = torch.empty_like(query)
    _run(ops, version, output, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
         max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale)

    if kv_cache_dtype == "fp8":  # test_attention.py:232-248
        dequantized_key_cache = torch.empty(key_cache.shape, dtype=dtype, device=DEV)
        ops.convert_fp8(dequantized_key_cache, key_cache)
        key_cache = dequantized_key_cache
        dequantized_value_cache = torch.empty(value_cache.shape, dtype=dtype, device=DEV)
        ops.convert_fp8(dequantized_value_cache, value_cache)
        value_cache = dequantized_value_cache
    ref_output = torch.empty_like(query)
    helpers.ref_single_query_cached_kv_attention(ref_output, query, num_queries_per_kv, key_cache, value_cache,
                                                 block_tables, seq_lens, scale, alibi_slopes)
    atol, rtol = 1e-3, 1e-5
    if kv_cache_dtype == "fp8":
        atol, rtol = 1e-2, 1e-5
    assert torch.allclose(output, ref_output, atol=atol, rtol=rtol)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("num_heads", [(8, 8), (16, 2), (32, 2), (24, 1)])  # group 1, 8, 16, 24 (two q tiles)
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv", [("auto", 1.0), ("fp8", 1.0), ("fp8", 0.37), ("fp8_e5m2", 2.0)])
@torch.inference_mode()
def test_paged_attention_vs_oracle(ops, oracle, version, num_heads, head_size, block_size, dtype, kv):
    """Small cases, checker = pinned CPU oracle (restates the reference CUDA arithmetic incl. fp8 dequant
    rounding and kv_scale).  Covers block size 8, e5m2, non-unit kv_scale, group sizes up to 24."""
    kv_cache_dtype, kv_scale = kv
    torch.manual_seed(1)
    random.seed(1)
    Hq, Hkv = num_heads
    S, NB = 4, 96
    lens = [1, block_size + 3, 517, 600]
    max_len = max(lens)
    scale = float(head_size ** -0.5)
    q = torch.empty(S, Hq, head_size, dtype=dtype).uniform_(-scale, scale)
    cdt = "fp8" if kv_cache_dtype.startswith("fp8") else "auto"
    if cdt == "fp8":
        x = 16
        kc = torch.randint(0, 256, (NB, Hkv, head_size // x, block_size, x), dtype=torch.uint8)
        vc = torch.randint(0, 256, (NB, Hkv, head_size, block_size), dtype=torch.uint8)
        for c in (kc, vc):  # keep magnitudes sane: clear the top exponent bit, avoid NaN codes
            c &= 0xB7 if kv_cache_dtype != "fp8_e5m2" else 0xBB
    else:
        kcs, vcs = helpers.create_kv_caches_with_random(NB, block_size, 1, Hkv, head_size, "auto", dtype, 1, "cpu")
        kc, vc = kcs[0], vcs[0]
    nblk = (max_len + block_size - 1) // block_size
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 1), dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    al = torch.randn(Hq, dtype=torch.float32) * 0.1
    want = torch.empty_like(q)
    oracle.paged_attention_v1(want, q, kc, vc, Hkv, scale, bt, sl, block_size, max_len, al, kv_cache_dtype, kv_scale)
    out = torch.empty_like(q, device=DEV)
    _run(ops, version, out, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), block_size, max_len,
         al.to(DEV), kv_cache_dtype, kv_scale)
    o, w = out.cpu().float(), want.float()
    tol = 1e-3 if cdt == "auto" else 1e-2
    ref_mag = w.abs().max().item() + 1e-6
    assert torch.allclose(o, w, atol=max(tol, tol * ref_mag), rtol=1e-2), (o - w).abs().max()


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("head_sliding_step", [0, 2, -1])
@pytest.mark.parametrize("num_heads", [(40, 40), (64, 8)])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16, torch.float])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("block_size", [16, 32])
@torch.inference_mode()
def test_blocksparse_paged_attention(ops, oracle, version, head_sliding_step, num_heads, dtype, kv_cache_dtype,
                                     block_size):
    """tests/kernels/test_blocksparse_attention.py:143-328 semantics (local 16, vert stride 8, sparse block 64)
    at a size the CPU oracle finishes quickly."""
    torch.manual_seed(2)
    Hq, Hkv = num_heads
    D, S, NB = 128, 3, 200
    lens = [5, 1700, 2300]
    max_len = max(lens)
    scale = float(D ** -0.5)
    q = torch.empty(S, Hq, D, dtype=dtype).uniform_(-scale, scale)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, block_size, 1, Hkv, D, kv_cache_dtype, dtype, 2, "cpu",
                                                    oracle.convert_fp8)
    kc, vc = kcs[0], vcs[0]
    nblk = (max_len + block_size - 1) // block_size
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 2), dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    bs = (16, 8, 64, head_sliding_step)
    want = torch.empty_like(q)
    oracle.paged_attention_v1(want, q, kc, vc, Hkv, scale, bt, sl, block_size, max_len, None, kv_cache_dtype, 1.0, 0,
                              *bs)
    out = torch.empty_like(q, device=DEV)
    _run(ops, version, out, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), block_size, max_len,
         None, kv_cache_dtype, 1.0, bs)
    # 16-bit queries with head size 64 / 128 run the block-sparse variant of the TMA + mma.sync kernel, fp32 the SIMT one
    assert ops.last_attention_path() == ("simt" if dtype == torch.float else "mma_tma")
    tol = 1e-3 if kv_cache_dtype == "auto" else 1e-2
    assert torch.allclose(out.cpu().float(), want.float(), atol=tol, rtol=1e-2)
    # and it really differs from dense attention
    dense = torch.empty_like(q, device=DEV)
    _run(ops, version, dense, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), block_size,
         max_len, None, kv_cache_dtype, 1.0)
    assert (dense[2] - out[2]).abs().max() > 1e-4


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("heads", [(8, 2), (32, 2), (24, 1)])   # groups 4, 16, 24 (2 q-tiles); 1 and 8: the test above
@pytest.mark.parametrize("head_size", [64, 112, 128])
@pytest.mark.parametrize("sparse", [(2, 4, 32, 1), (1, 16, 64, 3), (4, 2, 128, -2)])
@pytest.mark.parametrize("block_size,kv_cache_dtype,alibi", [(16, "auto", False), (32, "auto", True), (16, "fp8", False),
                                                             (32, "fp8_e5m2", False)])
@torch.inference_mode()
def test_blocksparse_mma_family(ops, oracle, version, heads, head_size, sparse, block_size, kv_cache_dtype, alibi):
    """The block-sparse variant of the TMA + mma.sync kernel (filtered block iterator shared by consumer and TMA
    producer, per-row masks where the heads' offsets differ) vs the oracle: (local_blocks, vert_stride, sparse block,
    head_sliding_step) incl. per-head and per-KV-head offsets, GQA groups up to 24, contexts long enough to split,
    lengths ending inside a block / a sparse block, ALiBi, fp8."""
    torch.manual_seed(7)
    Hq, Hkv = heads
    D, NB = head_size, 400
    lens = [1, 63, 700, 5000]   # (sized for the CPU oracle: the whole GPU suite has to fit the driver's 20-minute step)
    S, max_len = len(lens), max(lens)
    scale = float(D ** -0.5)
    dtype = torch.bfloat16 if kv_cache_dtype != "fp8" else torch.half
    q = torch.empty(S, Hq, D, dtype=dtype).uniform_(-scale, scale)
    # (the helper builds e4m3 bytes; read as e5m2 they are other, still finite, values — fine for a parity test)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, block_size, 1, Hkv, D,
                                                    "fp8" if kv_cache_dtype != "auto" else "auto", dtype, 3, "cpu",
                                                    oracle.convert_fp8)
    kc, vc = kcs[0], vcs[0]
    nblk = (max_len + block_size - 1) // block_size
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 3), dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    al = torch.randn(Hq, dtype=torch.float32) * 0.1 if alibi else None
    kv_scale = 1.0 if kv_cache_dtype == "auto" else 0.5
    want = torch.empty_like(q)
    oracle.paged_attention_v1(want, q, kc, vc, Hkv, scale, bt, sl, block_size, max_len, al, kv_cache_dtype, kv_scale, 0,
                              *sparse)
    out = torch.full_like(q, float("nan"), device=DEV)
    _run(ops, version, out, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), block_size, max_len,
         al.to(DEV) if al is not None else None, kv_cache_dtype, kv_scale, sparse)
    assert ops.last_attention_path() == "mma_tma"
    assert not torch.isnan(out).any()
    tol = 1e-3 if kv_cache_dtype == "auto" else 1e-2
    ref_mag = want.float().abs().max().item() + 1e-6
    assert torch.allclose(out.cpu().float(), want.float(), atol=max(tol, tol * ref_mag), rtol=1e-2), \
        (out.cpu().float() - want.float()).abs().max()


@pytest.mark.parametrize("case", ["attn_bf16_gqa", "attn_f32_alibi", "attn_bf16_long"])
@pytest.mark.parametrize("version", ["v1", "v2"])
@torch.inference_mode()
def test_paged_attention_golden(ops, golden, case, version):
    """Fixtures computed by the reference's own CPU kernels (csrc/cpu/attention.cpp) — tests/golden."""
    c = golden[case]
    S, Hq, Hkv, D, B, NB, max_len, _, has_alibi = [int(v) for v in c["meta"]]
    q = c["q"].to(DEV)
    al = c["alibi"].to(DEV) if has_alibi else None
    out = torch.empty_like(q)
    _run(ops, version, out, q, c["key_cache"].to(DEV), c["value_cache"].to(DEV), Hkv, float(c["scale"][0]),
         c["block_tables"].to(DEV), c["seq_lens"].to(DEV), B, max_len, al, "auto", 1.0)
    want = c["out_" + version].float()
    assert torch.allclose(out.cpu().float(), want, atol=1e-3, rtol=1e-2)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("shape", [(16, 32, 32, 128, 16, 1024), (24, 32, 8, 128, 16, 4096), (8, 64, 8, 128, 32, 8192),
                                   (6, 40, 40, 80, 16, 3000)])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@torch.inference_mode()
def test_paged_attention_vs_reference_cuda(ops, ref_cuda, version, shape, dtype, kv_cache_dtype):
    """Same inputs into the reference's own paged_attention_v{1,2} kernels (sm_100a build): atol 1e-3 / rtol 1e-2
    (north_star); fp8 KV bound: atol 2e-3 (our fp8 path multiplies kv_scale after the contraction instead of
    rounding every dequantised element to the 16-bit type first)."""
    torch.manual_seed(3)
    random.seed(3)
    S, Hq, Hkv, D, B, L = shape
    scale = float(D ** -0.5)
    q = torch.empty(S, Hq, D, dtype=dtype, device=DEV).uniform_(-scale, scale)
    NB = 2048
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 1, Hkv, D, kv_cache_dtype, dtype, 3, DEV, ops.convert_fp8)
    kc, vc = kcs[0], vcs[0]
    lens = [random.randint(1, L) for _ in range(S)]
    lens[-1] = L
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    bt = torch.tensor(helpers.make_block_tables(S, (L + B - 1) // B, NB, 3), dtype=torch.int32, device=DEV)
    al = torch.randn(Hq, dtype=torch.float32, device=DEV)
    kv_scale = 1.0 if kv_cache_dtype == "auto" else 0.75
    out = torch.empty_like(q)
    _run(ops, version, out, q, kc, vc, Hkv, scale, bt, sl, B, L, al, kv_cache_dtype, kv_scale)
    want = torch.empty_like(q)
    if version == "v1":
        ref_cuda.paged_attention_v1(want, q, kc, vc, Hkv, scale, bt, sl, B, L, al, kv_cache_dtype, kv_scale, 0, 0, 0, 64, 0)
    else:
        P = (L + 511) // 512
        tmp = torch.empty(S, Hq, P, D, dtype=dtype, device=DEV)
        es = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        ref_cuda.paged_attention_v2(want, es, ml, tmp, q, str(kc.data_ptr()), kc.stride(0), kc.stride(1),
                                    str(vc.data_ptr()), Hkv, scale, bt, sl, B, L, al, kv_cache_dtype, kv_scale, 0, 0, 0,
                                    64, 0)
    atol = 1e-3 if kv_cache_dtype == "auto" else 2e-3
    assert torch.allclose(out.float(), want.float(), atol=atol, rtol=1e-2), (out.float() - want.float()).abs().max()


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float])
@torch.inference_mode()
def test_edge_rows_and_stale_cache_bits(ops, version, dtype):
    """seq_len 0 / 1 / block-boundary rows, CUDA-graph style padded rows (seq_len 1, block table of zeros,
    model_runner.py:526-546), and NaN bit patterns behind seq_len in the last block must not leak
    (the reference zeroes those V lanes, attention_kernels.cu:420-430)."""
    torch.manual_seed(4)
    Hq, Hkv, D, B, NB = 16, 4, 128, 16, 64
    scale = float(D ** -0.5)
    lens = [0, 1, 16, 17, 1, 1000]
    S, L = len(lens), max(lens)
    q = torch.empty(S, Hq, D, dtype=dtype, device=DEV).uniform_(-scale, scale)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 1, Hkv, D, "auto", dtype, 4, DEV)
    kc, vc = kcs[0], vcs[0]
    bt = torch.tensor(helpers.make_block_tables(S, (L + B - 1) // B, NB - 1, 4, distinct=False), dtype=torch.int32,
                      device=DEV)
    bt[4] = 0  # padded row
    # poison everything behind each sequence's last token inside its last block
    poisoned_k, poisoned_v = kc.clone(), vc.clone()
    special = NB - 1  # private last block for sequence 3 (len 17 -> 1 token in its 2nd block)
    bt[3, 1] = special
    poisoned_k[special, :, :, 1:, :] = float("nan")
    poisoned_v[special, :, :, 1:] = float("nan")
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = torch.full_like(q, 7.0)
    _run(ops, version, out, q, poisoned_k, poisoned_v, Hkv, scale, bt, sl, B, L, None, "auto")
    assert torch.isfinite(out).all()
    assert out[0].abs().max() == 0  # empty context -> zeros (reference: 0 / (0 + 1e-6))
    want = torch.empty_like(q)
    helpers.ref_single_query_cached_kv_attention(want, q, Hq // Hkv, kc, vc, bt, sl, scale, None)
    assert torch.allclose(out, want, atol=1e-3, rtol=1e-5)
    # seq_len == 1: output is exactly V of the single token
    v0 = vc[bt[1, 0].long(), :, :, 0]  # [Hkv, D]
    assert torch.allclose(out[1].view(Hkv, Hq // Hkv, D), v0.unsqueeze(1).expand(-1, Hq // Hkv, -1), atol=2e-3, rtol=1e-2)


@pytest.mark.parametrize("version", ["v1", "v2"])
@torch.inference_mode()
def test_strided_query_and_pointer_string_abi(ops, version):
    """query rows with a larger stride (q_stride, attention_kernels.cu:706) as produced by a fused qkv output, and
    caches that are views into the flat [2, nb, B*H*D] tensor (paged_attn.py:46-59)."""
    from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention
    torch.manual_seed(5)
    S, Hq, Hkv, D, B, NB, L = 5, 32, 8, 128, 16, 128, 900
    scale = float(D ** -0.5)
    qkv = torch.empty(S, (Hq + 2 * Hkv) * D, dtype=torch.bfloat16, device=DEV).uniform_(-scale, scale)
    q = qkv[:, :Hq * D].view(S, Hq, D)
    assert not q.is_contiguous()
    kv_cache = torch.empty(PagedAttention.get_kv_cache_shape(NB, B, Hkv, D), dtype=torch.bfloat16, device=DEV)
    kv_cache.uniform_(-scale, scale)
    kc, vc = PagedAttention.split_kv_cache(kv_cache, Hkv, D)
    bt = torch.tensor(helpers.make_block_tables(S, (L + B - 1) // B, NB, 5), dtype=torch.int32, device=DEV)
    sl = torch.tensor([900, 3, 64, 511, 513], dtype=torch.int32, device=DEV)
    out = torch.empty(S, Hq, D, dtype=torch.bfloat16, device=DEV)
    _run(ops, version, out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto")
    want = torch.empty_like(out)
    helpers.ref_single_query_cached_kv_attention(want, q.contiguous(), Hq // Hkv, kc, vc, bt, sl, scale, None)
    assert torch.allclose(out, want, atol=1e-3, rtol=1e-5)
    fd = PagedAttention.forward_decode(q, kc, vc, bt, sl, L, "auto", Hkv, scale, None, 1.0)
    assert torch.allclose(fd, want, atol=1e-3, rtol=1e-5)


@torch.inference_mode()
def test_cuda_graph_capture_decode_step(ops):
    """§3.5: reshape_and_cache + paged_attention_v1/v2 must be capturable (no sync, no allocation, no H2D) and
    read seq_lens / block_tables / slot_mapping from device memory at replay time."""
    torch.manual_seed(6)
    S, Hq, Hkv, D, B, NB, L = 8, 32, 8, 128, 16, 256, 2048
    scale = float(D ** -0.5)
    dt = torch.bfloat16
    q = torch.empty(S, Hq, D, dtype=dt, device=DEV).uniform_(-scale, scale)
    k = torch.empty(S, Hkv, D, dtype=dt, device=DEV).uniform_(-scale, scale)
    v = torch.empty_like(k).uniform_(-scale, scale)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 1, Hkv, D, "auto", dt, 6, DEV)
    kc, vc = kcs[0], vcs[0]
    nblk = L // B
    bt = torch.zeros(S, nblk, dtype=torch.int32, device=DEV)
    sl = torch.ones(S, dtype=torch.int32, device=DEV)
    slots = torch.full((S,), -1, dtype=torch.long, device=DEV)  # capture-time padding (model_runner.py:1122-1158)
    out1, out2 = torch.empty_like(q), torch.empty_like(q)
    P = L // 512
    tmp = torch.empty(S, Hq, P, D, dtype=dt, device=DEV)
    es = torch.empty(S, Hq, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)

    def step():
        ops.reshape_and_cache(k, v, kc, vc, slots, "auto", 1.0)
        ops.paged_attention_v1(out1, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0)
        ops.paged_attention_v2(out2, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0)

    step()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    # now fill in the real step inputs and replay
    real_bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 6, distinct=False), dtype=torch.int32, device=DEV)
    real_sl = torch.tensor([1, 100, 512, 513, 2048, 7, 1999, 1024], dtype=torch.int32, device=DEV)
    bt.copy_(real_bt)
    sl.copy_(real_sl)
    new_slots = (real_bt[torch.arange(S), ((real_sl - 1) // B).long()].long() * B + ((real_sl - 1) % B).long())
    slots.copy_(new_slots)
    g.replay()
    torch.cuda.synchronize()
    want = torch.empty_like(q)
    helpers.ref_single_query_cached_kv_attention(want, q, Hq // Hkv, kc, vc, bt, sl, scale, None)
    assert torch.allclose(out1, want, atol=1e-3, rtol=1e-5)
    assert torch.allclose(out2, want, atol=1e-3, rtol=1e-5)
    # the new token really is in the cache (last position of each sequence == k/v of this step)
    b = (new_slots // B).long()
    o = (new_slots % B).long()
    assert torch.equal(vc[b, :, :, o], v)


def test_error_classes(ops):
    """TORCH_CHECK texts of attention_kernels.cu:764, :801 and quant_utils.cuh:564."""
    q = torch.zeros(1, 8, 120, dtype=torch.half, device=DEV)
    kc = torch.zeros(4, 2, 15, 16, 8, dtype=torch.half, device=DEV)
    vc = torch.zeros(4, 2, 120, 16, dtype=torch.half, device=DEV)
    bt = torch.zeros(1, 4, dtype=torch.int32, device=DEV)
    sl = torch.ones(1, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size: 120"):
        ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, 2, 1.0, bt, sl, 16, 64, None, "auto", 1.0)
    q = torch.zeros(1, 8, 128, dtype=torch.half, device=DEV)
    kc = torch.zeros(4, 2, 16, 64, 8, dtype=torch.half, device=DEV)
    vc = torch.zeros(4, 2, 128, 64, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported block size: 64"):
        ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, 2, 1.0, bt, sl, 64, 64, None, "auto", 1.0)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache: int8"):
        ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, 2, 1.0, bt, sl, 16, 64, None, "int8", 1.0)


@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("heads", [(16, 1), (5, 5)])
@pytest.mark.parametrize("head_size", [64, 112, 128])
@pytest.mark.parametrize("dtype", [torch.bfloat16])   # (fp16 / fp32 / fp8 with block 8: test_paged_attention_vs_oracle)
@pytest.mark.parametrize("alibi", [False, True])
@torch.inference_mode()
def test_block8_long_contexts(ops, oracle, version, heads, head_size, dtype, alibi):
    """Block size 8 on the TMA + mma.sync family (one n-tile per block, P.V with m16n8k8): contexts long enough to
    wrap every warp's ring many times and to split, lengths ending inside a block, odd numbers of k-steps (head 112)."""
    torch.manual_seed(11)
    Hq, Hkv = heads
    B, NB = 8, 1400
    lens = [9001, 1, 8, 9, 1027]
    S, max_len = len(lens), max(lens)
    scale = float(head_size ** -0.5)
    q = torch.empty(S, Hq, head_size, dtype=dtype).uniform_(-scale, scale)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 1, Hkv, head_size, "auto", dtype, 5, "cpu")
    kc, vc = kcs[0], vcs[0]
    nblk = (max_len + B - 1) // B
    bt = torch.tensor(helpers.make_block_tables(S, nblk, NB, 5), dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    al = torch.randn(Hq, dtype=torch.float32) * 0.1 if alibi else None
    want = torch.empty_like(q)
    oracle.paged_attention_v1(want, q, kc, vc, Hkv, scale, bt, sl, B, max_len, al, "auto", 1.0)
    out = torch.full_like(q, float("nan"), device=DEV)
    _run(ops, version, out, q.to(DEV), kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV), sl.to(DEV), B, max_len,
         al.to(DEV) if al is not None else None, "auto", 1.0)
    assert ops.last_attention_path() == "mma_tma" and ops.last_attention_partitions() > 1
    assert not torch.isnan(out).any()
    ref_mag = want.float().abs().max().item() + 1e-6
    assert torch.allclose(out.cpu().float(), want.float(), atol=max(1e-3, 1e-3 * ref_mag), rtol=1e-2), \
        (out.cpu().float() - want.float()).abs().max()


def test_mma_family_is_the_one_that_runs(ops):
    """The 16-bit configurations (block 16 | 32, and block 8 over a 16-bit cache) must dispatch to the TMA +
    tensor-core family, not silently to the SIMT one; fp32 queries and fp8 caches with block 8 stay on the SIMT one."""
    torch.manual_seed(0)
    for (dtype, kvd, B, want) in [(torch.bfloat16, "auto", 16, "mma_tma"), (torch.half, "fp8", 32, "mma_tma"),
                                  (torch.float, "auto", 16, "simt"), (torch.bfloat16, "auto", 8, "mma_tma"),
                                  (torch.half, "fp8", 8, "simt")]:
        q = torch.zeros(2, 32, 128, dtype=dtype, device=DEV)
        kcs, vcs = helpers.create_kv_caches_with_random(8, B, 1, 8, 128, kvd, dtype, 0, DEV, ops.convert_fp8)
        bt = torch.zeros(2, (50 + B - 1) // B, dtype=torch.int32, device=DEV)  # enough entries for 50 tokens
        sl = torch.tensor([5, 50], dtype=torch.int32, device=DEV)
        ops.paged_attention_v1(torch.empty_like(q), q, kcs[0], vcs[0], 8, 0.1, bt, sl, B, 50, None, kvd, 1.0)
        torch.cuda.synchronize()
        assert ops.last_attention_path() == want
