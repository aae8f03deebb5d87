"""b200pa_paged_attention_static_scratch — the decode attention entry point that allocates nothing (split-K partials
and arrival counters in a file-scope `static __device__` array, csrc/attention_noalloc.cu).  It has no counterpart in
the reference; the existing entry point b200pa_paged_attention_v1 is the checker here.  Both are called through the
C-ABI with ctypes (plain pointers, no torch types in the signatures).

This file is named to be collected LAST: a failure here must not hide the
parity tests.  It passes alone in a fresh process and after the other tests (it depends on no state of theirs: the
static array is cleared where it must be by the entry point itself).
"""
import ctypes
import os

import pytest
import torch

from tests.test_abi import AttnParams

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ATOL, ATOL_FP8, RTOL = 1e-3, 1e-2, 1e-5   # the attention tests' tolerances (tests/kernels/test_attention.py:271-274)
ERR_SCRATCH_LIMIT = -8
STATIC_SCRATCH_BYTES = 8 * 1024 * 1024
STATIC_SCRATCH_COUNTERS = 4096


@pytest.fixture(scope="module")
def lib():
    path = os.path.join(ROOT, "vllm_legacy_b200", "libb200pa.so")
    if not os.path.exists(path):
        pytest.fail(f"{path} is not built: run `python -m vllm_legacy_b200.build` (there is no fallback)")
    lib = ctypes.CDLL(path)
    for name in ("b200pa_paged_attention_v1", "b200pa_paged_attention_static_scratch"):
        getattr(lib, name).restype = ctypes.c_int
        getattr(lib, name).argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.b200pa_last_attention_partitions.restype = ctypes.c_int
    lib.b200pa_last_attention_path.restype = ctypes.c_char_p
    lib.b200pa_status_string.restype = ctypes.c_char_p
    lib.b200pa_status_string.argtypes = [ctypes.c_int]
    return lib


def _case(seed, lens, Hq=32, Hkv=8, D=128, B=16, dtype=torch.bfloat16, fp8=False):
    g = torch.Generator(device=DEV).manual_seed(seed)
    S, L = len(lens), max(lens)
    nblk = (L + B - 1) // B
    NB = S * nblk + 3
    scale = float(D) ** -0.5
    x = 16 if fp8 else 8
    cdt = torch.uint8 if fp8 else dtype
    if fp8:   # random finite e4m3 bytes of small magnitude (|x| < 0.5) with random signs
        def rnd(shape):
            mag = torch.randint(0, 0x30, shape, dtype=torch.uint8, device=DEV, generator=g)
            sign = torch.randint(0, 2, shape, dtype=torch.uint8, device=DEV, generator=g)
            return mag | (sign << 7)
        kc, vc = rnd((NB, Hkv, D // x, B, x)), rnd((NB, Hkv, D, B))
    else:
        kc = torch.empty(NB, Hkv, D // x, B, x, dtype=cdt, device=DEV).uniform_(-scale, scale, generator=g)
        vc = torch.empty(NB, Hkv, D, B, dtype=cdt, device=DEV).uniform_(-scale, scale, generator=g)
    q = torch.empty(S, Hq, D, dtype=dtype, device=DEV).uniform_(-scale, scale, generator=g)
    bt = torch.randperm(NB, device=DEV, generator=g)[:S * nblk].view(S, nblk).to(torch.int32).contiguous()
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    return dict(q=q, kc=kc, vc=vc, bt=bt, sl=sl, L=L, scale=scale, Hkv=Hkv, B=B, fp8=fp8)


def _params(c, out):
    p = AttnParams()
    p.out, p.query = out.data_ptr(), c["q"].data_ptr()
    p.key_cache, p.value_cache = c["kc"].data_ptr(), c["vc"].data_ptr()
    p.block_tables, p.seq_lens = c["bt"].data_ptr(), c["sl"].data_ptr()
    p.q_stride = c["q"].stride(0)
    p.kv_block_stride, p.kv_head_stride = c["kc"].stride(0), c["kc"].stride(1)
    p.num_seqs, p.num_heads, p.head_size = c["q"].shape
    p.num_kv_heads, p.block_size = c["Hkv"], c["B"]
    p.max_seq_len, p.max_num_blocks_per_seq = c["L"], c["bt"].shape[1]
    p.dtype = 1 if c["q"].dtype == torch.bfloat16 else 0
    p.kv_dtype = 1 if c["fp8"] else 0
    p.scale = c["scale"]
    p.kv_scale = 0.5 if c["fp8"] else 1.0
    p.blocksparse_block_size = 64
    return p


def _call(lib, fn, c, out=None):
    if out is None:
        out = torch.full_like(c["q"], float("nan"))
    p = _params(c, out)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = getattr(lib, fn)(ctypes.byref(p), stream)
    assert rc == 0, f"{fn}: {rc} ({lib.b200pa_status_string(rc).decode()})"
    parts = lib.b200pa_last_attention_partitions()
    torch.cuda.synchronize()
    return out, parts


@pytest.mark.parametrize("fp8", [False, True])
@torch.inference_mode()
def test_static_scratch_matches_v1_on_two_calls_in_a_row(lib, fp8):
    """Two different problems back to back through the no-alloc entry point (the second one finds the first one's
    partials and counters in the shared array), each equal to the existing entry point within the attention tests'
    tolerance; and the two calls do not take device memory."""
    first = _case(1, [3000, 17, 1500, 2049, 1, 777], fp8=fp8)
    second = _case(2, [4096, 4000, 513, 64, 2500, 3100, 1023, 9], fp8=fp8)
    want1, _ = _call(lib, "b200pa_paged_attention_v1", first)     # (also loads the kernels: module loading is not
    want2, _ = _call(lib, "b200pa_paged_attention_v1", second)    #  what the memory check below is about)
    _call(lib, "b200pa_paged_attention_static_scratch", first)
    # outputs allocated BEFORE the measurement window: what is bracketed below is the entry point alone, not torch's
    # caching allocator
    out1 = torch.full_like(first["q"], float("nan"))
    out2 = torch.full_like(second["q"], float("nan"))
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    got1, parts1 = _call(lib, "b200pa_paged_attention_static_scratch", first, out1)
    assert lib.b200pa_last_attention_path() == b"mma_tma_static"
    got2, parts2 = _call(lib, "b200pa_paged_attention_static_scratch", second, out2)
    free_after, _ = torch.cuda.mem_get_info()
    assert parts1 > 1 and parts2 > 1, "the cases must exercise the split path (that is what uses the scratch)"
    assert free_after >= free_before, "the no-alloc entry point took device memory"
    for got, want in ((got1, want1), (got2, want2)):
        assert not torch.isnan(got).any()
        torch.testing.assert_close(got.float(), want.float(), atol=ATOL_FP8 if fp8 else ATOL, rtol=RTOL)
    # and once more the first problem: same bits as its first run (nothing leaks from call to call)
    again, _ = _call(lib, "b200pa_paged_attention_static_scratch", first)
    assert torch.equal(again.view(torch.int16), got1.view(torch.int16))


@torch.inference_mode()
def test_static_scratch_rejects_what_does_not_fit(lib):
    """Beyond the stated maximum the call answers B200PA_ERR_SCRATCH_LIMIT and launches nothing."""
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # counters: num_seqs * num_kv_heads units > 4096
    many = _case(3, [40] * 600)
    assert 600 * 8 > STATIC_SCRATCH_COUNTERS
    out = torch.full_like(many["q"], 7.0)
    rc = lib.b200pa_paged_attention_static_scratch(ctypes.byref(_params(many, out)), stream)
    assert rc == ERR_SCRATCH_LIMIT
    assert b"static scratch" in lib.b200pa_status_string(rc)
    # partials: two splits of [num_seqs, num_heads, head_size] rows do not fit 8 MiB
    wide = _case(4, [600] * 300, Hq=64, Hkv=8)
    assert 2 * 300 * 64 * (128 * 2 + 8) > STATIC_SCRATCH_BYTES and 300 * 8 <= STATIC_SCRATCH_COUNTERS
    out2 = torch.full_like(wide["q"], 7.0)
    assert lib.b200pa_paged_attention_static_scratch(ctypes.byref(_params(wide, out2)), stream) == ERR_SCRATCH_LIMIT
    torch.cuda.synchronize()
    assert bool((out == 7.0).all()) and bool((out2 == 7.0).all()), "a rejected call must not write its output"
    # the existing entry point takes both
    assert lib.b200pa_paged_attention_v1(ctypes.byref(_params(many, out)), stream) == 0
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
