"""TEST INFRASTRUCTURE — ctypes front-end of the CPU oracle (oracle/paged_attn_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package never does.  Arguments are CPU torch tensors (used purely as
typed buffers); argument names follow the reference's ops (vllm/_custom_ops.py).
"""
import ctypes
import os
import subprocess
from typing import List, Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2, torch.uint8: 3}
KV = {"auto": 0, "fp8": 1, "fp8_e4m3": 1, "fp8_e5m2": 2}


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "paged_attn_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return LIB


class _AttnArgs(ctypes.Structure):
    _fields_ = [
        ("q", ctypes.c_void_p), ("k_cache", ctypes.c_void_p), ("v_cache", ctypes.c_void_p),
        ("block_tables", ctypes.c_void_p), ("seq_lens", ctypes.c_void_p), ("alibi_slopes", ctypes.c_void_p),
        ("q_stride", ctypes.c_int64), ("kv_block_stride", ctypes.c_int64), ("kv_head_stride", ctypes.c_int64),
        ("num_seqs", ctypes.c_int), ("num_heads", ctypes.c_int), ("num_kv_heads", ctypes.c_int),
        ("head_size", ctypes.c_int), ("block_size", ctypes.c_int), ("max_blocks_per_seq", ctypes.c_int),
        ("dt", ctypes.c_int), ("kv", ctypes.c_int),
        ("scale", ctypes.c_float), ("kv_scale", ctypes.c_float),
        ("tp_rank", ctypes.c_int), ("bs_local_blocks", ctypes.c_int), ("bs_vert_stride", ctypes.c_int),
        ("bs_block_size", ctypes.c_int), ("bs_head_sliding_step", ctypes.c_int),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
        _lib.oracle_float_to_fp8.restype = ctypes.c_uint8
        _lib.oracle_float_to_fp8.argtypes = [ctypes.c_float, ctypes.c_int]
        _lib.oracle_fp8_to_half.restype = ctypes.c_uint16
        _lib.oracle_fp8_to_half.argtypes = [ctypes.c_uint8, ctypes.c_int]
    return _lib


def _cpu(t: torch.Tensor) -> torch.Tensor:
    assert t.device.type == "cpu", "the oracle runs on host buffers"
    return t


def _attn_args(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
               alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, bs_local, bs_vert, bs_block, bs_slide):
    for t in (query, key_cache, value_cache, block_tables, seq_lens):
        _cpu(t)
    assert block_tables.dtype == torch.int32 and seq_lens.dtype == torch.int32
    assert block_tables.is_contiguous() and seq_lens.is_contiguous()
    a = _AttnArgs()
    a.q, a.k_cache, a.v_cache = query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr()
    a.block_tables, a.seq_lens = block_tables.data_ptr(), seq_lens.data_ptr()
    a.alibi_slopes = alibi_slopes.data_ptr() if alibi_slopes is not None else None
    a.q_stride, a.kv_block_stride, a.kv_head_stride = query.stride(0), key_cache.stride(0), key_cache.stride(1)
    a.num_seqs, a.num_heads, a.head_size = query.shape
    a.num_kv_heads, a.block_size, a.max_blocks_per_seq = num_kv_heads, block_size, block_tables.shape[1]
    a.dt, a.kv = DT[query.dtype], KV[kv_cache_dtype]
    a.scale, a.kv_scale = scale, kv_scale
    a.tp_rank, a.bs_local_blocks, a.bs_vert_stride = tp_rank, bs_local, bs_vert
    a.bs_block_size, a.bs_head_sliding_step = bs_block, bs_slide
    return a


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens,
                       block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale, tp_rank=0,
                       blocksparse_local_blocks=0, blocksparse_vert_stride=0, blocksparse_block_size=64,
                       blocksparse_head_sliding_step=0) -> None:
    a = _attn_args(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                   alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, blocksparse_local_blocks,
                   blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)
    assert out.is_contiguous()
    rc = lib().oracle_paged_attention_v1(ctypes.c_void_p(_cpu(out).data_ptr()), ctypes.byref(a),
                                         ctypes.c_int(max_seq_len))
    assert rc == 0


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, kv_scale,
                       tp_rank=0, blocksparse_local_blocks=0, blocksparse_vert_stride=0,
                       blocksparse_block_size=64, blocksparse_head_sliding_step=0) -> None:
    a = _attn_args(query, key_cache, value_cache, num_kv_heads, scale, block_tables, seq_lens, block_size,
                   alibi_slopes, kv_cache_dtype, kv_scale, tp_rank, blocksparse_local_blocks,
                   blocksparse_vert_stride, blocksparse_block_size, blocksparse_head_sliding_step)
    for t in (out, exp_sums, max_logits, tmp_out):
        assert _cpu(t).is_contiguous()
    rc = lib().oracle_paged_attention_v2(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(exp_sums.data_ptr()),
                                         ctypes.c_void_p(max_logits.data_ptr()), ctypes.c_void_p(tmp_out.data_ptr()),
                                         ctypes.byref(a), ctypes.c_int(max_seq_len))
    assert rc == 0


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype, kv_scale) -> None:
    for t in (key, value, key_cache, value_cache, slot_mapping):
        _cpu(t)
    assert slot_mapping.dtype == torch.int64 and slot_mapping.is_contiguous()
    assert key_cache.is_contiguous() and value_cache.is_contiguous()
    T, H, D = key.shape
    rc = lib().oracle_reshape_and_cache(
        ctypes.c_void_p(key.data_ptr()), ctypes.c_void_p(value.data_ptr()), ctypes.c_void_p(key_cache.data_ptr()),
        ctypes.c_void_p(value_cache.data_ptr()), ctypes.c_void_p(slot_mapping.data_ptr()), ctypes.c_int64(T),
        ctypes.c_int64(key.stride(0)), ctypes.c_int64(value.stride(0)), ctypes.c_int(H), ctypes.c_int(D),
        ctypes.c_int(key_cache.shape[3]), ctypes.c_int(key_cache.shape[4]), ctypes.c_int(DT[key.dtype]),
        ctypes.c_int(KV[kv_cache_dtype]), ctypes.c_float(kv_scale))
    assert rc == 0


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping, kv_cache_dtype) -> None:
    """csrc/cache_kernels.cu:204-235, 289-345; caches [nb, block_size, H, D] (only "auto")."""
    if kv_cache_dtype != "auto":
        raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")
    for t in (key, value, key_cache, value_cache, slot_mapping):
        _cpu(t)
    assert slot_mapping.dtype == torch.int64 and slot_mapping.is_contiguous()
    assert key_cache.stride(0) == value_cache.stride(0) and key_cache[0].is_contiguous()
    T, H, D = key.shape
    rc = lib().oracle_reshape_and_cache_flash(
        ctypes.c_void_p(key.data_ptr()), ctypes.c_void_p(value.data_ptr()), ctypes.c_void_p(key_cache.data_ptr()),
        ctypes.c_void_p(value_cache.data_ptr()), ctypes.c_void_p(slot_mapping.data_ptr()), ctypes.c_int64(T),
        ctypes.c_int64(key.stride(0)), ctypes.c_int64(value.stride(0)), ctypes.c_int64(key_cache.stride(0)),
        ctypes.c_int(H), ctypes.c_int(D), ctypes.c_int(key_cache.shape[1]), ctypes.c_int(DT[key.dtype]))
    assert rc == 0


def rotary_embedding(positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim=None,
                     cos_sin_cache_offsets=None) -> None:
    """csrc/pos_encoding_kernels.cu:10-201, in place on query and key; with cos_sin_cache_offsets it is the
    batched form (:95-131, :166-201)."""
    for t in (positions, query, key, cos_sin_cache):
        _cpu(t)
    assert positions.dtype == torch.int64 and positions.is_contiguous() and cos_sin_cache.is_contiguous()
    assert query.stride(-1) == 1 and key.stride(-1) == 1 and query.dtype == key.dtype == cos_sin_cache.dtype
    if rot_dim is None:
        rot_dim = cos_sin_cache.shape[1]
    if cos_sin_cache_offsets is not None:
        _cpu(cos_sin_cache_offsets)
        assert cos_sin_cache_offsets.dtype == torch.int64 and cos_sin_cache_offsets.is_contiguous()
        num_tokens = cos_sin_cache_offsets.shape[0]
        off = ctypes.c_void_p(cos_sin_cache_offsets.data_ptr())
    else:
        num_tokens = query.numel() // query.shape[-1]
        off = ctypes.c_void_p(None)
    rc = lib().oracle_rotary_embedding(
        ctypes.c_void_p(positions.data_ptr()), ctypes.c_void_p(query.data_ptr()), ctypes.c_void_p(key.data_ptr()),
        ctypes.c_void_p(cos_sin_cache.data_ptr()), off, ctypes.c_int64(num_tokens),
        ctypes.c_int64(query.stride(-2)), ctypes.c_int64(key.stride(-2)),
        ctypes.c_int(query.shape[-1] // head_size), ctypes.c_int(key.shape[-1] // head_size),
        ctypes.c_int(head_size), ctypes.c_int(rot_dim), ctypes.c_int(1 if is_neox else 0),
        ctypes.c_int(DT[query.dtype]))
    assert rc == 0


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor], block_mapping) -> None:
    n = len(key_caches)
    if n == 0:
        return
    bm = _cpu(block_mapping).contiguous()
    assert bm.dtype == torch.int64
    kp = (ctypes.c_void_p * n)(*[_cpu(t).data_ptr() for t in key_caches])
    vp = (ctypes.c_void_p * n)(*[_cpu(t).data_ptr() for t in value_caches])
    block_bytes = key_caches[0][0].numel() * key_caches[0].element_size()
    rc = lib().oracle_copy_blocks(kp, vp, ctypes.c_int(n), ctypes.c_void_p(bm.data_ptr()),
                                  ctypes.c_int64(bm.shape[0]), ctypes.c_int64(block_bytes))
    assert rc == 0


def swap_blocks(src, dst, block_mapping) -> None:
    bm = _cpu(block_mapping).contiguous()
    block_bytes = src[0].numel() * src.element_size()
    rc = lib().oracle_swap_blocks(ctypes.c_void_p(_cpu(src).data_ptr()), ctypes.c_void_p(_cpu(dst).data_ptr()),
                                  ctypes.c_void_p(bm.data_ptr()), ctypes.c_int64(bm.shape[0]),
                                  ctypes.c_int64(block_bytes))
    assert rc == 0


def convert_fp8(output, input, scale: float = 1.0, kv_dtype: str = "fp8") -> None:
    assert _cpu(output).is_contiguous() and _cpu(input).is_contiguous()
    rc = lib().oracle_convert_fp8(ctypes.c_void_p(output.data_ptr()), ctypes.c_void_p(input.data_ptr()),
                                  ctypes.c_int64(input.numel()), ctypes.c_int(DT[output.dtype]),
                                  ctypes.c_int(DT[input.dtype]), ctypes.c_float(scale), ctypes.c_int(KV[kv_dtype]))
    if rc != 0:
        raise RuntimeError(f"oracle_convert_fp8 rc={rc}")


def float_to_fp8(x: float, kv_dtype: str = "fp8") -> int:
    return int(lib().oracle_float_to_fp8(ctypes.c_float(x), KV[kv_dtype]))


def dequant_prefix_caches(k_cache_u8, v_cache_u8, dtype, kv_scale: float = 1.0, kv_dtype: str = "fp8"):
    """fp8 paged caches (K [nb, H, D/16, B, 16], V [nb, H, D, B], uint8) -> the 16-bit caches the kernels see
    after `scaled_convert` on load (quant_utils.cuh:293-454: fp8 -> half -> float x kv_scale -> dtype), K
    re-laid-out to the 16-bit layout [nb, H, D/8, B, 8].  This IS the fp8 semantics of the decode kernels
    (attention_kernels.cu:284-289, 413-419), applied to prefix-prefill."""
    k8, v8 = _cpu(k_cache_u8).contiguous(), _cpu(v_cache_u8).contiguous()
    assert k8.dtype == torch.uint8 and v8.dtype == torch.uint8 and k8.shape[-1] == 16
    kd = torch.empty(k8.shape, dtype=dtype)
    vd = torch.empty(v8.shape, dtype=dtype)
    convert_fp8(kd, k8, kv_scale, kv_dtype)
    convert_fp8(vd, v8, kv_scale, kv_dtype)
    nb, h, dg, b, _ = k8.shape
    # [nb, H, D/16, B, 16] -> token-major [nb, H, B, D] -> [nb, H, D/8, B, 8]
    kd = kd.permute(0, 1, 3, 2, 4).reshape(nb, h, b, dg * 2, 8).permute(0, 1, 3, 2, 4).contiguous()
    return kd, vd


def context_attention_fwd(q, k, v, o, k_cache, v_cache, b_loc, b_start_loc, b_seq_len, b_ctx_len, max_input_len,
                          alibi_slopes=None, sliding_window=None, kv_cache_dtype: str = "auto",
                          kv_scale: float = 1.0) -> None:
    """Restatement of the reference's Triton prefix-prefill kernels (prefix_prefill.py:10-245, 442-681).
    uint8 caches are dequantised first (`dequant_prefix_caches`); with kv_cache_dtype "auto" they are read as
    e5m2 without a scale, the reference kernels' own bit-cast (prefix_prefill.py:122-123, 168-169)."""
    if k_cache.dtype == torch.uint8:
        kvd, sc = ("fp8_e5m2", 1.0) if kv_cache_dtype == "auto" else (kv_cache_dtype, kv_scale)
        k_cache, v_cache = dequant_prefix_caches(k_cache, v_cache, q.dtype, sc, kvd)
    for t in (q, k, v, o, k_cache, v_cache):
        assert _cpu(t).is_contiguous() and t.dtype == q.dtype
    assert k_cache.shape[-1] == 8, "16-bit cache layout (x = 8)"
    loc = _cpu(b_loc).to(torch.int32).contiguous()
    start = _cpu(b_start_loc).to(torch.int32).contiguous()
    seql = _cpu(b_seq_len).to(torch.int32).contiguous()
    ctxl = _cpu(b_ctx_len).to(torch.int32).contiguous()
    slopes = _cpu(alibi_slopes).to(torch.float32).contiguous() if alibi_slopes is not None else None
    f = lib().oracle_context_attention_fwd
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_void_p] * 7 + [ctypes.c_int64] + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 5 + \
        [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    rc = f(o.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
           loc.data_ptr(), loc.stride(0), start.data_ptr(), seql.data_ptr(), ctxl.data_ptr(), seql.shape[0],
           q.shape[1], k.shape[1], q.shape[2], v_cache.shape[3],
           slopes.data_ptr() if slopes is not None else None,
           int(sliding_window) if sliding_window else 0, DT[q.dtype])
    assert rc == 0
