"""The C-ABI shared library loads, exports every symbol include/b200pa.h declares, and its argument
validation answers with the reference's error classes — without touching a GPU (validation happens
before any CUDA call).  Also checks the torch extension's module surface (csrc/pybind.cpp:8-95)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200pa.h")
LIB = os.path.join(ROOT, "vllm_legacy_b200", "libb200pa.so")


def _declared_symbols():
    with open(HEADER) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pa_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} is not built: run `python -m vllm_legacy_b200.build` (there is no fallback)")
    return ctypes.CDLL(LIB)


def test_header_declares_the_hot_path():
    syms = _declared_symbols()
    for want in ("b200pa_paged_attention_v1", "b200pa_paged_attention_v2", "b200pa_reshape_and_cache",
                 "b200pa_copy_blocks", "b200pa_swap_blocks", "b200pa_convert_fp8",
                 "b200pa_get_max_shared_memory_per_block_device_attribute"):
        assert want in syms


def test_library_exports_every_declared_symbol(lib):
    for s in _declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/b200pa.h but not exported"
    assert lib.b200pa_abi_version() == 2


class AttnParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("out", "query", "key_cache", "value_cache", "block_tables",
                                               "seq_lens", "alibi_slopes", "exp_sums", "max_logits", "tmp_out")]
    _fields_ += [(n, ctypes.c_int64) for n in ("q_stride", "kv_block_stride", "kv_head_stride")]
    _fields_ += [(n, ctypes.c_int32) for n in ("num_seqs", "num_heads", "num_kv_heads", "head_size", "block_size",
                                               "max_seq_len", "max_num_blocks_per_seq", "max_num_partitions",
                                               "dtype", "kv_dtype")]
    _fields_ += [("scale", ctypes.c_float), ("kv_scale", ctypes.c_float)]
    _fields_ += [(n, ctypes.c_int32) for n in ("tp_rank", "blocksparse_local_blocks", "blocksparse_vert_stride",
                                               "blocksparse_block_size", "blocksparse_head_sliding_step")]


def _params(**kw):
    p = AttnParams()
    p.num_seqs, p.num_heads, p.num_kv_heads, p.head_size, p.block_size = 1, 8, 2, 128, 16
    p.max_seq_len, p.max_num_blocks_per_seq, p.dtype, p.kv_dtype = 64, 4, 1, 0
    p.scale = p.kv_scale = 1.0
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def test_static_scratch_entry_point_validation(lib):
    """b200pa_paged_attention_static_scratch (no reference counterpart): argument classes and the static-scratch
    limits are checked before any CUDA call, so they answer without a GPU."""
    f = lib.b200pa_paged_attention_static_scratch
    f.restype = ctypes.c_int
    assert f(None, None) == -5
    assert f(ctypes.byref(_params()), None) == -5           # null pointers
    assert f(ctypes.byref(_params(block_size=8)), None) == -2
    assert f(ctypes.byref(_params(dtype=2)), None) == -4    # fp32 queries: not in the mma family
    assert f(ctypes.byref(_params(kv_dtype=7)), None) == -3
    fake = {n: 4096 for n in ("out", "query", "key_cache", "value_cache", "block_tables", "seq_lens")}
    strides = dict(q_stride=8 * 128, kv_block_stride=2 * 128 * 16, kv_head_stride=128 * 16)
    # 4097 sequences x 1 KV head: one arrival counter too many
    assert f(ctypes.byref(_params(num_seqs=4097, num_heads=8, num_kv_heads=1, max_seq_len=4096, **fake, **strides)), None) == -8
    # 2 x [2048 seqs, 8 heads] x (128 x 2 + 8) bytes = 8.25 MiB of partials: over the 8 MiB array
    assert f(ctypes.byref(_params(num_seqs=2048, num_heads=8, num_kv_heads=1, max_seq_len=4096, **fake, **strides)), None) == -8
    lib.b200pa_status_string.restype = ctypes.c_char_p
    assert b"static scratch" in lib.b200pa_status_string(-8)
    # dense only: block-sparse parameters are an invalid argument here (b200pa_paged_attention_v1 / _v2 serve them)
    assert f(ctypes.byref(_params(blocksparse_vert_stride=8, blocksparse_block_size=64, **fake, **strides)), None) == -5


def test_attention_argument_validation(lib):
    """Error classes of attention_kernels.cu:764 (head size), :801 (block size), quant_utils.cuh:564/538."""
    f = lib.b200pa_paged_attention_v1
    f.restype = ctypes.c_int
    assert f(ctypes.byref(_params(head_size=100)), None) == -1
    assert f(ctypes.byref(_params(block_size=64)), None) == -2
    assert f(ctypes.byref(_params(kv_dtype=7)), None) == -3
    assert f(ctypes.byref(_params(dtype=3)), None) == -4
    assert f(ctypes.byref(_params()), None) == -5          # null pointers
    assert f(None, None) == -5
    assert f(ctypes.byref(_params(num_seqs=0)), None) == 0  # empty batch is a no-op
    lib.b200pa_status_string.restype = ctypes.c_char_p
    assert lib.b200pa_status_string(-1) == b"Unsupported head size: "
    assert lib.b200pa_status_string(-2) == b"Unsupported block size: "


def test_cache_ops_argument_validation(lib):
    assert lib.b200pa_copy_blocks(None, None, 0, None, ctypes.c_int64(0), ctypes.c_int64(0), None) == 0
    assert lib.b200pa_swap_blocks(None, None, None, ctypes.c_int64(0), ctypes.c_int64(0), 0, None) == 0
    assert lib.b200pa_swap_blocks(None, None, None, ctypes.c_int64(2), ctypes.c_int64(16), 0, None) == -5
    assert lib.b200pa_convert_fp8(None, None, ctypes.c_int64(0), 3, 0, ctypes.c_float(1.0), 1, None) == 0
    buf = (ctypes.c_char * 16)()
    assert lib.b200pa_convert_fp8(buf, buf, ctypes.c_int64(4), 3, 0, ctypes.c_float(1.0), 0, None) == -3
    assert lib.b200pa_reshape_and_cache(None, None) == -5


def test_torch_extension_surface():
    """Same submodules / names as csrc/pybind.cpp registers for this path."""
    from vllm_legacy_b200 import load_extension
    C = load_extension()
    for name in ("paged_attention_v1", "paged_attention_v2"):
        assert hasattr(C.ops, name)
    for name in ("swap_blocks", "copy_blocks", "reshape_and_cache", "convert_fp8"):
        assert hasattr(C.cache_ops, name)
    assert hasattr(C.cuda_utils, "get_max_shared_memory_per_block_device_attribute")


def test_no_cpu_fallback():
    """CPU tensors are rejected loudly (the reference would dispatch to its CPU extension; we have none)."""
    from vllm_legacy_b200 import _custom_ops as ops
    q = torch.zeros(1, 8, 128, dtype=torch.bfloat16)
    kc = torch.zeros(4, 2, 16, 16, 8, dtype=torch.bfloat16)
    vc = torch.zeros(4, 2, 128, 16, dtype=torch.bfloat16)
    bt = torch.zeros(1, 4, dtype=torch.int32)
    sl = torch.ones(1, dtype=torch.int32)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, 2, 1.0, bt, sl, 16, 64, None, "auto", 1.0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.reshape_and_cache(q[:, :2], q[:, :2], kc, vc, torch.zeros(1, dtype=torch.int64), "auto", 1.0)
    with pytest.raises(RuntimeError, match="Invalid device combination"):
        ops.swap_blocks(kc, kc.clone(), torch.zeros(1, 2, dtype=torch.int64))


def test_product_never_imports_oracle():
    """The product package must not reference oracle/ (a routed-through oracle voids parity)."""
    pkg = os.path.join(ROOT, "vllm_legacy_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                with open(os.path.join(dp, fn)) as f:
                    src = f.read()
                assert "oracle" not in src.lower(), f"{fn} mentions the oracle"
