"""GPU parity of context_attention_fwd (csrc/prefill.cu via the C-ABI b200pa_context_attention_fwd) on the
reference's own test grid (tests/kernels/test_prefix_prefill.py:13-60, 214-262) plus what it does not cover:
bf16, block sizes 8/16/32/128, realistic magnitudes, empty prefix, stale NaN bits behind ctx_len, strided q/k/v.
Checkers: fp32 dense attention (the xformers computation of the reference test), the CPU oracle restating the
Triton kernels, and — when staged under oracle/_ref/ — the reference's Triton kernels themselves.
"""
import math

import pytest
import torch

from tests import helpers

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

NUM_HEADS = [64]
NUM_QUERIES_PER_KV = [1, 8, 64]
HEAD_SIZES = [128, 96, 24]
SLIDING_WINDOW = [0, 16, 64, 128, 256, 512, 2048]


def _run(ops, case, slopes=None, window=None):
    out = torch.full_like(case["query"], float("nan"))
    ops.context_attention_fwd(case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"],
                              case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                              case["max_input_len"], slopes, window)
    torch.cuda.synchronize()
    return out


def _alibi_slopes(total_num_heads: int) -> torch.Tensor:
    """bloom-style slopes, as the reference test builds them (test_prefix_prefill.py:239-262)."""
    p2 = 2 ** math.floor(math.log2(total_num_heads))
    base = 2 ** (-(2 ** -(math.log2(p2) - 3)))
    slopes = [base ** i for i in range(1, p2 + 1)]
    if p2 != total_num_heads:
        extra = 2 ** (-(2 ** -(math.log2(2 * p2) - 3)))
        slopes += [extra ** i for i in range(1, 1 + 2 * min(p2, total_num_heads - p2), 2)]
    return torch.tensor(slopes, dtype=torch.float32)


@pytest.mark.parametrize("num_heads", NUM_HEADS)
@pytest.mark.parametrize("num_queries_per_kv", NUM_QUERIES_PER_KV)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
@pytest.mark.parametrize("sliding_window", SLIDING_WINDOW)
@torch.inference_mode()
def test_contexted_kv_attention_reference_grid(ops, num_heads, num_queries_per_kv, head_size, sliding_window):
    """The reference's test verbatim: fp16, U(-1e-3, 1e-3), block 32, atol 1e-6 (:211)."""
    case = helpers.make_prefill_case(num_heads, num_queries_per_kv, head_size, torch.float16, DEV)
    out = _run(ops, case, None, sliding_window)
    ref = helpers.ref_context_attention(case, None, sliding_window)
    assert torch.allclose(ref, out.float(), atol=1e-6, rtol=0)


@pytest.mark.parametrize("num_queries_per_kv", NUM_QUERIES_PER_KV)
@pytest.mark.parametrize("head_size", HEAD_SIZES)
@torch.inference_mode()
def test_contexted_kv_attention_alibi_reference_grid(ops, num_queries_per_kv, head_size):
    case = helpers.make_prefill_case(64, num_queries_per_kv, head_size, torch.float16, DEV)
    slopes = _alibi_slopes(64).to(DEV)
    out = _run(ops, case, slopes, None)
    ref = helpers.ref_context_attention(case, slopes, 0)
    assert torch.allclose(ref, out.float(), atol=1e-6, rtol=0)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256, 24])
@pytest.mark.parametrize("block_size", [8, 16, 32, 128])
@pytest.mark.parametrize("mode", ["plain", "window", "alibi"])
@torch.inference_mode()
def test_realistic_magnitudes(ops, dtype, head_size, block_size, mode):
    """U(-1, 1) inputs: logits O(1), so masks, softmax and the p rounding actually matter."""
    case = helpers.make_prefill_case(16, 4, head_size, dtype, DEV, batch=6, max_q=300, max_ctx=400,
                                     block_size=block_size, amplitude=1.0, min_len=1, seed=block_size + head_size,
                                     stale_nan=True)
    slopes = _alibi_slopes(16).to(DEV) if mode == "alibi" else None
    window = 100 if mode == "window" else 0
    out = _run(ops, case, slopes, window)
    ref = helpers.ref_context_attention(case, slopes, window)
    assert not torch.isnan(out).any()
    atol = 1e-3 if dtype == torch.float16 else 4e-3   # north_star: atol 1e-3 / rtol 1e-2 for 16-bit types
    torch.testing.assert_close(out.float(), ref, atol=atol, rtol=1e-2)


@pytest.mark.parametrize("mode", ["plain", "window", "alibi"])
@torch.inference_mode()
def test_vs_cpu_oracle(ops, oracle, mode):
    case = helpers.make_prefill_case(8, 2, 64, torch.float16, DEV, query_lens=[70, 1, 200, 129],
                                     ctx_lens=[130, 300, 0, 31], block_size=16, amplitude=1.0)
    slopes = _alibi_slopes(8) if mode == "alibi" else None
    window = 77 if mode == "window" else 0
    out = _run(ops, case, slopes.to(DEV) if slopes is not None else None, window)
    cpu = {n: (t.cpu() if torch.is_tensor(t) else t) for n, t in case.items()}
    exp = torch.empty_like(cpu["query"])
    oracle.context_attention_fwd(cpu["query"], cpu["k"], cpu["v"], exp, cpu["k_cache"], cpu["v_cache"],
                                 cpu["block_table"], cpu["b_start_loc"], cpu["b_seq_len"], cpu["b_ctx_len"],
                                 cpu["max_input_len"], slopes, window)
    torch.testing.assert_close(out.cpu().float(), exp.float(), atol=1e-3, rtol=1e-2)


@pytest.mark.parametrize("sliding_window", [0, 64, 2048])
@pytest.mark.parametrize("alibi", [False, True])
@torch.inference_mode()
def test_vs_reference_triton(ops, sliding_window, alibi):
    """The reference's Triton kernels run beside ours (only when oracle/build_ref.py staged the module)."""
    ref_mod = helpers.load_ref_prefix_prefill()
    if ref_mod is None:
        pytest.skip("oracle/_ref/prefix_prefill_ref.py not staged or Triton unusable")
    if alibi and sliding_window:
        pytest.skip("the reference's ALiBi kernel has no window")
    case = helpers.make_prefill_case(32, 4, 128, torch.float16, DEV, amplitude=1.0)
    slopes = _alibi_slopes(32).to(DEV) if alibi else None
    out = _run(ops, case, slopes, sliding_window)
    exp = torch.empty_like(case["query"])
    try:
        ref_mod.context_attention_fwd(case["query"], case["k"], case["v"], exp, case["k_cache"], case["v_cache"],
                                      case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                                      case["max_input_len"], slopes, sliding_window)
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference Triton kernel does not run with this Triton: {type(e).__name__}")
    torch.testing.assert_close(out.float(), exp.float(), atol=1e-3, rtol=1e-2)


@torch.inference_mode()
def test_int64_index_tensors_and_strided_inputs(ops):
    """The reference test passes int64 block tables / lengths (:83-90); vLLM passes q/k/v as slices of a fused qkv."""
    case = helpers.make_prefill_case(16, 4, 128, torch.float16, DEV, batch=4, max_q=200, max_ctx=200, amplitude=1.0)
    ref = helpers.ref_context_attention(case)
    T = case["query"].shape[0]
    qkv = torch.empty(T, (16 + 4 + 4) * 128, dtype=torch.float16, device=DEV)
    qkv[:, :16 * 128] = case["query"].reshape(T, -1)
    qkv[:, 16 * 128:20 * 128] = case["k"].reshape(T, -1)
    qkv[:, 20 * 128:] = case["v"].reshape(T, -1)
    q, k, v = qkv.split([16 * 128, 4 * 128, 4 * 128], dim=-1)
    out = torch.empty(T, 16, 128, dtype=torch.float16, device=DEV)
    ops.context_attention_fwd(q.view(T, 16, 128), k.view(T, 4, 128), v.view(T, 4, 128), out, case["k_cache"],
                              case["v_cache"], case["block_table"].long(), case["b_start_loc"].long(),
                              case["b_seq_len"].long(), case["b_ctx_len"].long(), case["max_input_len"])
    torch.testing.assert_close(out.float(), ref, atol=1e-3, rtol=1e-2)


@torch.inference_mode()
def test_forward_prefix_helper(ops):
    """PagedAttention.forward_prefix (paged_attn.py:185-216) end to end, query_start_loc of batch + 1 entries."""
    from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention
    case = helpers.make_prefill_case(8, 1, 128, torch.bfloat16, DEV, batch=3, max_q=150, max_ctx=150, amplitude=1.0,
                                     block_size=16)
    qsl = torch.cat([case["b_start_loc"], torch.tensor([case["query"].shape[0]], dtype=torch.int32, device=DEV)])
    out = PagedAttention.forward_prefix(case["query"], case["k"], case["v"], case["k_cache"], case["v_cache"],
                                        case["block_table"], qsl, case["b_seq_len"], case["b_ctx_len"],
                                        case["max_input_len"], None, None)
    torch.testing.assert_close(out.float(), helpers.ref_context_attention(case), atol=4e-3, rtol=1e-2)


@torch.inference_mode()
def test_errors(ops):
    case = helpers.make_prefill_case(8, 1, 128, torch.float16, DEV, batch=2, max_q=32, max_ctx=32)
    out = torch.empty_like(case["query"])
    with pytest.raises(RuntimeError):   # fp32 is not a tensor-core input type of this kernel
        ops.context_attention_fwd(case["query"].float(), case["k"].float(), case["v"].float(), out.float(),
                                  case["k_cache"].float(), case["v_cache"].float(), case["block_table"],
                                  case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], 32)
    with pytest.raises(RuntimeError):   # CPU tensors: no CPU path
        ops.context_attention_fwd(case["query"].cpu(), case["k"].cpu(), case["v"].cpu(), out.cpu(),
                                  case["k_cache"].cpu(), case["v_cache"].cpu(), case["block_table"].cpu(),
                                  case["b_start_loc"].cpu(), case["b_seq_len"].cpu(), case["b_ctx_len"].cpu(), 32)


@pytest.mark.parametrize("seed", range(24))
@torch.inference_mode()
def test_random_ragged_batches(ops, seed):
    """Randomised shapes over both kernel families: ragged query/context lengths (including 0-token prefixes and
    1-token queries), GQA ratios, head sizes, block sizes, windows, ALiBi."""
    import random
    rng = random.Random(1000 + seed)
    dtype = rng.choice([torch.float16, torch.bfloat16])
    head_size = rng.choice([64, 128, 128, 128, 80, 96, 256])
    qpk = rng.choice([1, 2, 4, 8])
    num_heads = qpk * rng.choice([1, 2, 3])
    block_size = rng.choice([8, 16, 16, 32, 64])
    batch = rng.randint(1, 6)
    qmax = rng.choice([1, 17, 64, 65, 129, 400, 700])
    query_lens = [rng.randint(1, qmax) for _ in range(batch)]
    ctx_lens = [rng.choice([0, rng.randint(1, 900)]) for _ in range(batch)]
    mode = rng.choice(["plain", "plain", "window", "alibi"])
    case = helpers.make_prefill_case(num_heads, qpk, head_size, dtype, DEV, query_lens=query_lens, ctx_lens=ctx_lens,
                                     block_size=block_size, amplitude=1.0, seed=seed, stale_nan=True)
    slopes = _alibi_slopes(num_heads).to(DEV) if mode == "alibi" else None
    window = rng.choice([1, 7, 64, 130, 1000]) if mode == "window" else 0
    out = _run(ops, case, slopes, window)
    ref = helpers.ref_context_attention(case, slopes, window)
    assert not torch.isnan(out).any()
    atol = 1e-3 if dtype == torch.float16 else 4e-3
    torch.testing.assert_close(out.float(), ref, atol=atol, rtol=1e-2)


@pytest.mark.parametrize("qpk", [1, 2, 3])
@pytest.mark.parametrize("mode", ["plain", "window", "alibi"])
@pytest.mark.parametrize("head_size", [64, 128])
@torch.inference_mode()
def test_q_tile_pairing(ops, qpk, mode, head_size):
    """The tcgen05 kernel pairs two Q tiles per CTA: two heads of an even GQA group (qpk 2), else two consecutive
    128-row tiles of one head (qpk 1 and 3).  Query lengths sit on every side of the 128 / 256-row edges so that the
    second tile of a pair is empty, partial or full, with and without a prefix; the 130-token window makes the first
    tiles of the upper tile fully masked before it has seen a visible key."""
    query_lens = [1, 127, 128, 129, 255, 256, 257, 385, 600]
    ctx_lens = [300, 0, 64, 5, 0, 1000, 33, 129, 0]
    num_heads = 2 * qpk
    case = helpers.make_prefill_case(num_heads, qpk, head_size, torch.bfloat16, DEV, query_lens=query_lens,
                                     ctx_lens=ctx_lens, block_size=16, amplitude=1.0, seed=7 + qpk, stale_nan=True)
    slopes = _alibi_slopes(num_heads).to(DEV) if mode == "alibi" else None
    window = 130 if mode == "window" else 0
    out = _run(ops, case, slopes, window)
    ref = helpers.ref_context_attention(case, slopes, window)
    assert not torch.isnan(out).any()
    torch.testing.assert_close(out.float(), ref, atol=4e-3, rtol=1e-2)
    # one-tile prompts with an odd group take the single-tile kernel
    short = helpers.make_prefill_case(num_heads, qpk, head_size, torch.bfloat16, DEV, query_lens=[128, 70, 3],
                                      ctx_lens=[200, 0, 17], block_size=16, amplitude=1.0, seed=11, stale_nan=True)
    out = _run(ops, short, slopes, window)
    torch.testing.assert_close(out.float(), helpers.ref_context_attention(short, slopes, window), atol=4e-3, rtol=1e-2)


@torch.inference_mode()
def test_deterministic_and_graph_capturable(ops):
    """Same inputs -> bit-identical output on every launch; the op allocates nothing and never synchronises when the
    index tensors are int32, so a prefill chunk can be captured in a CUDA graph and replayed."""
    case = helpers.make_prefill_case(16, 4, 128, torch.bfloat16, DEV, batch=4, max_q=300, max_ctx=500, amplitude=1.0,
                                     block_size=16)
    a = _run(ops, case)
    b = _run(ops, case)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    out = torch.zeros_like(case["query"])
    args = (case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"], case["block_table"],
            case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"], None, 0)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        ops.context_attention_fwd(*args)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ops.context_attention_fwd(*args)
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int16), a.view(torch.int16))


@torch.inference_mode()
def test_long_context_chunk(ops):
    """BASELINE-sized shape through a size-independent property: a 512-token chunk over a 32k-token prefix with
    constant V must return that constant for every row and head (softmax weights sum to 1), and permuting the
    physical blocks must not change a bit."""
    H, qpk, D, B = 8, 4, 128, 16
    case = helpers.make_prefill_case(H, qpk, D, torch.bfloat16, DEV, query_lens=[512], ctx_lens=[32768], block_size=B,
                                     amplitude=1.0)
    a = _run(ops, case)
    perm = torch.randperm(case["k_cache"].shape[0], device=DEV)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(perm.numel(), device=DEV)
    case2 = dict(case)
    case2["k_cache"] = case["k_cache"][perm].contiguous()
    case2["v_cache"] = case["v_cache"][perm].contiguous()
    case2["block_table"] = inv[case["block_table"].long()].to(torch.int32)
    b = _run(ops, case2)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    case3 = dict(case)
    case3["v_cache"] = torch.full_like(case["v_cache"], 0.5)
    case3["v"] = torch.full_like(case["v"], 0.5)
    c = _run(ops, case3)
    torch.testing.assert_close(c.float(), torch.full_like(c, 0.5).float(), atol=4e-3, rtol=0)


# ---- prefix-prefill over an fp8 KV cache (dequantised while staged; SURVEY §8 f1, BASELINE config 4) -----------
def _run_fp8(ops, case, kv_dtype, kv_scale, slopes=None, window=None):
    out = torch.full_like(case["query"], float("nan"))
    ops.context_attention_fwd(case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"],
                              case["block_table"], case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"],
                              case["max_input_len"], slopes, window, kv_dtype, kv_scale)
    torch.cuda.synchronize()
    return out


def _oracle_fp8(oracle, case, kv_dtype, kv_scale, slopes=None, window=None):
    cpu = {n: (t.cpu() if torch.is_tensor(t) else t) for n, t in case.items()}
    exp = torch.empty_like(cpu["query"])
    oracle.context_attention_fwd(cpu["query"], cpu["k"], cpu["v"], exp, cpu["k_cache"], cpu["v_cache"],
                                 cpu["block_table"], cpu["b_start_loc"], cpu["b_seq_len"], cpu["b_ctx_len"],
                                 cpu["max_input_len"], slopes, window, kv_dtype, kv_scale)
    return exp


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("head_size", [128, 64, 96, 192])   # 128/64: tcgen05 loaders; 96/192: mma.sync staging
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("kv", [("fp8", 1.0), ("fp8_e4m3", 0.37), ("fp8_e5m2", 2.0)])
@torch.inference_mode()
def test_fp8_kv_prefix_vs_oracle(ops, oracle, dtype, head_size, block_size, kv):
    """fp8 cache bytes -> T with the decode kernels' conversion, then the usual prefix attention: compare with the
    oracle run over the dequantised cache.  Stale slots hold fp8 NaN bytes."""
    kv_dtype, kv_scale = kv
    case = helpers.make_prefill_case(8, 4, head_size, dtype, DEV, query_lens=[150, 1, 260, 129, 40],
                                     ctx_lens=[130, 300, 0, 31, 7], block_size=block_size, amplitude=1.0,
                                     seed=head_size + block_size, stale_nan=True)
    q8 = helpers.quantize_prefill_case(case, oracle, kv_dtype, kv_scale)
    out = _run_fp8(ops, q8, kv_dtype, kv_scale)
    assert not torch.isnan(out).any()
    exp = _oracle_fp8(oracle, q8, kv_dtype, kv_scale)
    atol = 1e-3 if dtype == torch.float16 else 4e-3
    torch.testing.assert_close(out.cpu().float(), exp.float(), atol=atol, rtol=1e-2)


@pytest.mark.parametrize("mode", ["window", "alibi"])
@pytest.mark.parametrize("max_q", [48, 400])   # <= 64 rows: mma.sync family; longer: tcgen05
@torch.inference_mode()
def test_fp8_kv_prefix_masks(ops, oracle, mode, max_q):
    case = helpers.make_prefill_case(16, 8, 128, torch.bfloat16, DEV, batch=5, max_q=max_q, max_ctx=500,
                                     block_size=16, amplitude=1.0, min_len=1, seed=max_q, stale_nan=True)
    q8 = helpers.quantize_prefill_case(case, oracle, "fp8", 0.5)
    slopes = _alibi_slopes(16) if mode == "alibi" else None
    window = 100 if mode == "window" else 0
    out = _run_fp8(ops, q8, "fp8", 0.5, slopes.to(DEV) if slopes is not None else None, window)
    assert not torch.isnan(out).any()
    exp = _oracle_fp8(oracle, q8, "fp8", 0.5, slopes, window)
    torch.testing.assert_close(out.cpu().float(), exp.float(), atol=4e-3, rtol=1e-2)


@torch.inference_mode()
def test_fp8_kv_prefix_auto_is_e5m2_bitcast(ops, oracle):
    """A uint8 cache with kv_cache_dtype "auto" is read as e5m2 without a scale — the reference Triton kernels' own
    rule (prefix_prefill.py:122-123, 168-169); an explicit fp8 dtype + kv_scale gives the decode kernels' values."""
    H, Hkv, D, B = 8, 2, 128, 16
    case = helpers.make_prefill_case(H, H // Hkv, D, torch.bfloat16, DEV, query_lens=[1, 1, 140], ctx_lens=[100, 257, 31],
                                     block_size=B, amplitude=1.0, seed=5)
    q8 = helpers.quantize_prefill_case(case, oracle, "fp8", 0.5)
    out = _run_fp8(ops, q8, "fp8", 0.5)
    torch.testing.assert_close(out.cpu().float(), _oracle_fp8(oracle, q8, "fp8", 0.5).float(), atol=4e-3, rtol=1e-2)
    q5 = helpers.quantize_prefill_case(case, oracle, "fp8_e5m2", 1.0)
    out5 = _run_fp8(ops, q5, "auto", 1.0)
    torch.testing.assert_close(out5.cpu().float(), _oracle_fp8(oracle, q5, "fp8_e5m2", 1.0).float(), atol=4e-3, rtol=1e-2)


@torch.inference_mode()
def test_fp8_kv_prefix_rejects_bad_shapes(ops):
    case = helpers.make_prefill_case(4, 1, 24, torch.float16, DEV, query_lens=[5], ctx_lens=[9], block_size=16)
    k8 = torch.zeros(case["k_cache"].shape[0], 4, 24 // 8, 16, 8, dtype=torch.uint8, device=DEV)  # x = 8 bytes: not an fp8 layout
    v8 = torch.zeros_like(case["v_cache"], dtype=torch.uint8)
    out = torch.empty_like(case["query"])
    with pytest.raises(RuntimeError):
        ops.context_attention_fwd(case["query"], case["k"], case["v"], out, k8, v8, case["block_table"],
                                  case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"],
                                  None, None, "fp8", 1.0)
