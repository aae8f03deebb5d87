"""Shared test fixtures: restatements of the reference's test helpers for this path.

  create_kv_caches_with_random  <- vllm/utils.py:398-446 (+ _generate_random_fp8, :324-341)
  ref_single_query_cached_kv_attention / ref_masked_attention
                                <- tests/kernels/test_attention.py:43-111
  load_ref_cpu / load_ref_cuda  -> the reference's own sources compiled into oracle/_ref/
"""
import importlib.util
import os
import random
import sys
from typing import List, Optional, Tuple

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def _load_so(name: str, path: str):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def load_ref_cpu():
    """The reference's csrc/cpu/attention.cpp + cache.cpp (unchanged) — or None if not built /
    the host lacks AVX512."""
    flags = cpu_flags()
    if not {"avx512f", "avx512bw", "avx512vl", "avx512dq"} <= flags:
        return None
    name = "ref_cpu_bf16" if "avx512_bf16" in flags else "ref_cpu"
    path = os.path.join(REF_DIR, name + ".so")
    if not os.path.exists(path):
        return None
    return _load_so(name, path)


def load_ref_cuda():
    """The reference's csrc/attention/attention_kernels.cu + cache_kernels.cu built for sm_100a."""
    path = os.path.join(REF_DIR, "ref_cuda.so")
    if not os.path.exists(path) or not torch.cuda.is_available():
        return None
    return _load_so("ref_cuda", path)


STR_DTYPE_TO_TORCH_DTYPE = {"half": torch.half, "bfloat16": torch.bfloat16, "float": torch.float}


def create_kv_caches_with_random(num_blocks: int, block_size: int, num_layers: int, num_heads: int,
                                 head_size: int, cache_dtype: str, model_dtype, seed: int = 0,
                                 device: str = "cuda", convert_fp8=None
                                 ) -> Tuple[List[torch.Tensor], List[torch.Tensor]]:
    """vllm/utils.py:398-446: K [nb, H, D/x, B, x], V [nb, H, D, B], uniform(-D^-0.5, D^-0.5);
    fp8 caches are generated in fp16 and converted with convert_fp8 (vllm/utils.py:324-341)."""
    torch.random.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
    if isinstance(model_dtype, str):
        model_dtype = STR_DTYPE_TO_TORCH_DTYPE[model_dtype]
    if cache_dtype == "auto":
        torch_dtype = model_dtype
    elif cache_dtype in ("half", "bfloat16", "float"):
        torch_dtype = STR_DTYPE_TO_TORCH_DTYPE[cache_dtype]
    elif cache_dtype == "fp8":
        torch_dtype = torch.uint8
    else:
        raise ValueError(f"Invalid kv cache dtype: {cache_dtype}")
    scale = head_size ** -0.5
    x = 16 // torch.tensor([], dtype=torch_dtype).element_size()
    key_cache_shape = (num_blocks, num_heads, head_size // x, block_size, x)
    value_cache_shape = (num_blocks, num_heads, head_size, block_size)

    def _fill(shape):
        t = torch.empty(size=shape, dtype=torch_dtype, device=device)
        if cache_dtype == "fp8":
            tmp = torch.empty_like(t, dtype=torch.float16)
            tmp.uniform_(-scale, scale)
            assert convert_fp8 is not None, "fp8 caches need a convert_fp8 op"
            convert_fp8(t, tmp)
        else:
            t.uniform_(-scale, scale)
        return t

    key_caches = [_fill(key_cache_shape) for _ in range(num_layers)]
    value_caches = [_fill(value_cache_shape) for _ in range(num_layers)]
    return key_caches, value_caches


def ref_masked_attention(query, key, value, scale, attn_mask=None):
    """tests/kernels/test_attention.py:43-55"""
    attn_weights = scale * torch.einsum("qhd,khd->hqk", query, key).float()
    if attn_mask is not None:
        attn_weights = attn_weights + attn_mask.float()
    attn_weights = torch.softmax(attn_weights, dim=-1).to(value.dtype)
    return torch.einsum("hqk,khd->qhd", attn_weights, value)


def ref_single_query_cached_kv_attention(output, query, num_queries_per_kv, key_cache, value_cache,
                                         block_tables, seq_lens, scale, alibi_slopes) -> None:
    """tests/kernels/test_attention.py:58-111 (gathers token by token; small cases only)."""
    num_query_heads = query.shape[1]
    num_kv_heads = value_cache.shape[1]
    head_size = value_cache.shape[2]
    block_size = value_cache.shape[3]
    num_seqs = query.shape[0]
    block_tables_lst = block_tables.cpu().tolist()
    seq_lens_lst = seq_lens.cpu().tolist()
    for i in range(num_seqs):
        q = query[i].unsqueeze(0)
        block_table = block_tables_lst[i]
        seq_len = int(seq_lens_lst[i])
        if seq_len == 0:
            output[i].zero_()
            continue
        idx = torch.arange(seq_len, device=query.device)
        blocks = torch.tensor(block_table, device=query.device, dtype=torch.long)[idx // block_size]
        offs = idx % block_size
        # key_cache[b, :, :, off, :] -> [H, D/x, x] -> [H, D]
        keys = key_cache[blocks, :, :, offs, :].reshape(seq_len, num_kv_heads, head_size)
        values = value_cache[blocks, :, :, offs]
        if num_queries_per_kv > 1:
            keys = torch.repeat_interleave(keys, num_queries_per_kv, dim=1)
            values = torch.repeat_interleave(values, num_queries_per_kv, dim=1)
        alibi_bias = None
        if alibi_slopes is not None:
            position_ids = torch.arange(seq_len, device=query.device).int()
            alibi_bias = (position_ids - seq_len + 1).float()
            alibi_bias = alibi_slopes.view(-1, 1, 1) * alibi_bias.view(1, 1, -1)
        out = ref_masked_attention(q, keys, values, scale, alibi_bias)
        output[i].copy_(out.view(num_query_heads, head_size), non_blocking=True)


def make_block_tables(num_seqs: int, max_num_blocks_per_seq: int, num_blocks: int, seed: int = 0,
                      distinct: bool = False) -> List[List[int]]:
    """Random block tables (test_attention.py:159-167, with replacement) or, for bandwidth
    measurements, a permutation of distinct physical blocks."""
    rng = random.Random(seed)
    if distinct:
        assert num_seqs * max_num_blocks_per_seq <= num_blocks
        perm = list(range(num_blocks))
        rng.shuffle(perm)
        return [perm[i * max_num_blocks_per_seq:(i + 1) * max_num_blocks_per_seq]
                for i in range(num_seqs)]
    return [[rng.randint(0, num_blocks - 1) for _ in range(max_num_blocks_per_seq)]
            for _ in range(num_seqs)]


# ---------------------------------------------------------------------------------------------------
# prefix-prefill (context_attention_fwd) fixtures — tests/kernels/test_prefix_prefill.py:50-124
# ---------------------------------------------------------------------------------------------------
def make_prefill_case(num_heads: int, num_queries_per_kv: int, head_size: int, dtype: torch.dtype, device: str,
                      batch: int = 10, max_q: int = 1024, max_ctx: int = 1024, block_size: int = 32, seed: int = 0,
                      amplitude: float = 1e-3, min_len: int = 16, query_lens=None, ctx_lens=None, stale_nan=False):
    """Inputs of the reference's test: ragged query/context lengths, the context scattered into a paged cache
    through a random block table, K cache [nb, H, D/8, B, 8], V cache [nb, H, D, B]."""
    rng = random.Random(seed)
    g = torch.Generator().manual_seed(seed)
    query_lens = query_lens or [rng.randint(min_len, max_q) for _ in range(batch)]
    ctx_lens = ctx_lens if ctx_lens is not None else [rng.randint(min_len, max_ctx) for _ in range(batch)]
    batch = len(query_lens)
    seq_lens = [a + b for a, b in zip(query_lens, ctx_lens)]
    num_kv_heads = num_heads // num_queries_per_kv
    max_block_per_request = max(1, (max(ctx_lens) + block_size - 1) // block_size)
    cache_size = batch * max_block_per_request
    num_tokens = sum(query_lens)
    query = torch.empty(num_tokens, num_heads, head_size, dtype=torch.float32).uniform_(-amplitude, amplitude, generator=g)
    kv = torch.empty(sum(seq_lens), 2, num_kv_heads, head_size, dtype=torch.float32).uniform_(-amplitude, amplitude, generator=g)
    query, kv = query.to(dtype), kv.to(dtype)
    key, value = kv.unbind(dim=1)
    fill = float("nan") if stale_nan else 0.0
    k_cache = torch.full((cache_size, block_size, num_kv_heads, head_size), fill, dtype=dtype)
    v_cache = torch.full((cache_size, block_size, num_kv_heads, head_size), fill, dtype=dtype)
    k = torch.zeros(num_tokens, num_kv_heads, head_size, dtype=dtype)
    v = torch.zeros(num_tokens, num_kv_heads, head_size, dtype=dtype)
    block_table = torch.randperm(cache_size, generator=g).view(batch, max_block_per_request)
    b_start_loc = torch.cumsum(torch.tensor([0] + query_lens[:-1]), dim=0)
    b_seq_start_loc = torch.cumsum(torch.tensor([0] + seq_lens[:-1]), dim=0)
    for i in range(batch):
        s, c, ql = int(b_seq_start_loc[i]), ctx_lens[i], query_lens[i]
        k[int(b_start_loc[i]):int(b_start_loc[i]) + ql] = key[s + c:s + c + ql]
        v[int(b_start_loc[i]):int(b_start_loc[i]) + ql] = value[s + c:s + c + ql]
        for blk in range((c + block_size - 1) // block_size):
            n = min(block_size, c - blk * block_size)
            phys = int(block_table[i, blk])
            k_cache[phys, :n] = key[s + blk * block_size:s + blk * block_size + n]
            v_cache[phys, :n] = value[s + blk * block_size:s + blk * block_size + n]
    k_cache = k_cache.view(-1, block_size, num_kv_heads, head_size // 8, 8).permute(0, 2, 3, 1, 4).contiguous()
    v_cache = v_cache.view(-1, block_size, num_kv_heads, head_size).permute(0, 2, 3, 1).contiguous()
    case = dict(query=query, k=k, v=v, k_cache=k_cache, v_cache=v_cache, block_table=block_table.to(torch.int32),
                b_start_loc=b_start_loc.to(torch.int32), b_seq_len=torch.tensor(seq_lens, dtype=torch.int32),
                b_ctx_len=torch.tensor(ctx_lens, dtype=torch.int32), key_full=key.contiguous(),
                value_full=value.contiguous())
    case = {n: t.to(device) for n, t in case.items()}
    case.update(query_lens=query_lens, ctx_lens=ctx_lens, seq_lens=seq_lens, max_input_len=max(query_lens),
                num_kv_heads=num_kv_heads)
    return case


def ref_context_attention(case, alibi_slopes: Optional[torch.Tensor] = None, sliding_window: int = 0,
                          scale: Optional[float] = None) -> torch.Tensor:
    """fp32 dense masked attention per sequence over the un-paged keys/values (what xformers'
    BlockDiagonalCausalFromBottomRightMask [+ local attention] computes in the reference's test, :160-186)."""
    query, key, value = case["query"], case["key_full"], case["value_full"]
    T, H, D = query.shape
    Hkv = key.shape[1]
    scale = scale if scale is not None else 1.0 / (D ** 0.5)
    out = torch.empty(T, H, D, dtype=torch.float32, device=query.device)
    qs = ss = 0
    for ql, sl in zip(case["query_lens"], case["seq_lens"]):
        q = query[qs:qs + ql].float().transpose(0, 1)                               # [H, ql, D]
        kk = key[ss:ss + sl].float().repeat_interleave(H // Hkv, dim=1).transpose(0, 1)   # [H, sl, D]
        vv = value[ss:ss + sl].float().repeat_interleave(H // Hkv, dim=1).transpose(0, 1)
        logits = torch.matmul(q, kk.transpose(1, 2)) * scale                        # [H, ql, sl]
        qpos = torch.arange(sl - ql, sl, device=query.device)[:, None]
        kpos = torch.arange(sl, device=query.device)[None, :]
        mask = kpos <= qpos
        if alibi_slopes is not None:
            logits = logits + alibi_slopes.float()[:, None, None] * (kpos - qpos).float()[None]
        elif sliding_window and sliding_window > 0:
            mask = mask & (qpos - kpos < sliding_window)
        logits = logits.masked_fill(~mask[None], float("-inf"))
        out[qs:qs + ql] = torch.matmul(torch.softmax(logits, dim=-1), vv).transpose(0, 1)
        qs += ql
        ss += sl
    return out


def load_ref_prefix_prefill():
    """The reference's own Triton prefix-prefill module, when oracle/build_ref.py staged it under oracle/_ref/
    (git-ignored; test infrastructure only) and Triton + a GPU are available; else None."""
    path = os.path.join(REF_DIR, "prefix_prefill_ref.py")
    if not os.path.exists(path) or not torch.cuda.is_available():
        return None
    try:
        spec = importlib.util.spec_from_file_location("prefix_prefill_ref", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:  # noqa: BLE001 - an incompatible Triton just means "no reference run"
        return None


def quantize_prefill_case(case, oracle, kv_dtype: str = "fp8", kv_scale: float = 1.0):
    """Same prefill case over an fp8 KV cache: the 16-bit caches are quantised with the reference's
    reshape_and_cache conversion (oracle.convert_fp8 = cvt_satfinite(x / kv_scale), quant_utils.cuh:456-488) and K is
    re-laid-out to the fp8 layout [nb, H, D/16, B, 16] (x = 16).  NaN fill (stale slots) stays NaN in fp8."""
    out = dict(case)
    kc, vc = case["k_cache"].cpu().contiguous(), case["v_cache"].cpu().contiguous()
    dev = case["k_cache"].device
    nb, h, dg, b, x = kc.shape
    assert x == 8 and dg % 2 == 0, "fp8 caches need head_size % 16 == 0"
    k_tok = kc.permute(0, 1, 3, 2, 4).reshape(nb, h, b, dg // 2, 16).permute(0, 1, 3, 2, 4).contiguous()
    k8 = torch.empty(k_tok.shape, dtype=torch.uint8)
    v8 = torch.empty(vc.shape, dtype=torch.uint8)
    oracle.convert_fp8(k8, k_tok, kv_scale, kv_dtype)
    oracle.convert_fp8(v8, vc, kv_scale, kv_dtype)
    out["k_cache"], out["v_cache"] = k8.to(dev), v8.to(dev)
    return out
