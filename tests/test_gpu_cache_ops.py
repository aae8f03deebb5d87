"""GPU parity of the cache ops (reshape_and_cache incl. fp8, copy_blocks, swap_blocks, convert_fp8).
Bit-exact bar.  Parametrisation follows tests/kernels/test_cache.py of the reference; the checker
is the CPU oracle (and the reference's own CUDA kernels when oracle/_ref/ref_cuda.so is present).
Calls go through the product's public ops (torch extension -> C-ABI) and, in one test, through the
C-ABI directly with ctypes."""
import ctypes
import os
import random

import pytest
import torch

from tests import helpers

pytestmark = pytest.mark.gpu

DTYPES = [torch.half, torch.bfloat16, torch.float]
DEV = "cuda"


def _bits(t):
    return t.contiguous().view(torch.uint8)


# -------------------------------------------------------------------------------- reshape_and_cache
@pytest.mark.parametrize("num_tokens", [42])
@pytest.mark.parametrize("num_heads", [8])
@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@torch.inference_mode()
def test_reshape_and_cache(ops, oracle, num_tokens, num_heads, head_size, block_size, dtype, kv_cache_dtype):
    """tests/kernels/test_cache.py:105-193 — non-contiguous K/V rows (qkv.unbind), random unique slots."""
    random.seed(0)
    torch.manual_seed(0)
    num_blocks = 64
    slot_mapping = torch.tensor(random.sample(range(num_blocks * block_size), num_tokens), dtype=torch.long)
    slot_mapping[3] = -1  # padding token (model_runner.py:1122-1158 captures graphs with slot -1)
    qkv = torch.randn(num_tokens, 3, num_heads, head_size, dtype=dtype)
    _, key, value = qkv.unbind(dim=1)
    kcs, vcs = helpers.create_kv_caches_with_random(num_blocks, block_size, 1, num_heads, head_size, kv_cache_dtype,
                                                    dtype, 0, "cpu", oracle.convert_fp8)
    kc_ref, vc_ref = kcs[0], vcs[0]
    kc, vc = kc_ref.to(DEV), vc_ref.to(DEV)
    qkv_d = qkv.to(DEV)
    _, key_d, value_d = qkv_d.unbind(dim=1)
    for kv_scale in ((1.0,) if kv_cache_dtype == "auto" else (1.0, 0.0418)):
        ops.reshape_and_cache(key_d, value_d, kc, vc, slot_mapping.to(DEV), kv_cache_dtype, kv_scale)
        oracle.reshape_and_cache(key, value, kc_ref, vc_ref, slot_mapping, kv_cache_dtype, kv_scale)
        assert torch.equal(_bits(kc.cpu()), _bits(kc_ref))
        assert torch.equal(_bits(vc.cpu()), _bits(vc_ref))


@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@torch.inference_mode()
def test_reshape_and_cache_vs_reference_cuda(ops, ref_cuda, dtype, kv_cache_dtype):
    """Bit-exact against the reference's own reshape_and_cache_kernel (cache_kernels.cu:150-202) on sm_100a."""
    torch.manual_seed(0)
    T, H, D, B, NB = 333, 8, 128, 16, 64
    slots = torch.randperm(NB * B)[:T].to(DEV)
    qkv = torch.randn(T, 3, H, D, dtype=dtype, device=DEV)
    _, key, value = qkv.unbind(dim=1)
    kcs, vcs = helpers.create_kv_caches_with_random(NB, B, 2, H, D, kv_cache_dtype, dtype, 0, DEV, ops.convert_fp8)
    kcs[1].copy_(kcs[0])
    vcs[1].copy_(vcs[0])
    ops.reshape_and_cache(key, value, kcs[0], vcs[0], slots, kv_cache_dtype, 0.5)
    ref_cuda.reshape_and_cache(key, value, str(kcs[1].data_ptr()), kcs[1].size(3), kcs[1].size(4),
                               str(vcs[1].data_ptr()), slots, kv_cache_dtype, 0.5)
    assert torch.equal(_bits(kcs[0]), _bits(kcs[1])) and torch.equal(_bits(vcs[0]), _bits(vcs[1]))


@torch.inference_mode()
def test_reshape_and_cache_scalar_path_and_empty(ops, oracle):
    """Misaligned rows (odd element offset) take the scalar kernel; zero tokens is a no-op."""
    torch.manual_seed(0)
    T, H, D, B, NB = 7, 3, 64, 16, 4
    flat = torch.randn(T * H * D * 2 + 8, dtype=torch.bfloat16)
    key = flat[1:1 + T * H * D].view(T, H, D)          # 2-byte aligned only
    value = flat[T * H * D + 3:2 * T * H * D + 3].view(T, H, D)
    kc = torch.zeros(NB, H, D // 8, B, 8, dtype=torch.bfloat16)
    vc = torch.zeros(NB, H, D, B, dtype=torch.bfloat16)
    slots = torch.tensor([5, 0, 63, 17, 18, 31, 32], dtype=torch.long)
    kc_d, vc_d = kc.to(DEV), vc.to(DEV)
    flat_d = flat.to(DEV)
    key_d = flat_d[1:1 + T * H * D].view(T, H, D)
    value_d = flat_d[T * H * D + 3:2 * T * H * D + 3].view(T, H, D)
    ops.reshape_and_cache(key_d, value_d, kc_d, vc_d, slots.to(DEV), "auto", 1.0)
    oracle.reshape_and_cache(key, value, kc, vc, slots, "auto", 1.0)
    assert torch.equal(_bits(kc_d.cpu()), _bits(kc)) and torch.equal(_bits(vc_d.cpu()), _bits(vc))
    ops.reshape_and_cache(key_d[:0], value_d[:0], kc_d, vc_d, slots[:0].to(DEV), "auto", 1.0)
    assert torch.equal(_bits(kc_d.cpu()), _bits(kc))


@pytest.mark.parametrize("shape", [(8, 128, 16), (3, 80, 32), (40, 64, 8), (5, 256, 16), (12, 192, 16)])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@pytest.mark.parametrize("consecutive", [True, False])
@torch.inference_mode()
def test_reshape_and_cache_tile_path(ops, oracle, shape, dtype, kv_cache_dtype, consecutive):
    """num_tokens >= 512 takes the shared-memory tile kernel (prefill): prompt-like consecutive slots starting
    at an arbitrary offset inside a block, or fully random slots, with padding tokens and a ragged last tile."""
    num_heads, head_size, block_size = shape
    random.seed(1)
    torch.manual_seed(1)
    T, num_blocks = 1000 + 7, 160
    if consecutive:
        start = 5  # not block aligned
        slots = torch.arange(start, start + T, dtype=torch.long)
    else:
        slots = torch.tensor(random.sample(range(num_blocks * block_size), T), dtype=torch.long)
    slots[17] = -1
    slots[T - 1] = -1
    qkv = torch.randn(T, 3, num_heads, head_size, dtype=dtype)
    _, key, value = qkv.unbind(dim=1)
    nb = max(num_blocks, (T + 5) // block_size + 2)
    kcs, vcs = helpers.create_kv_caches_with_random(nb, block_size, 1, num_heads, head_size, kv_cache_dtype, dtype, 1,
                                                    "cpu", oracle.convert_fp8)
    kc_ref, vc_ref = kcs[0], vcs[0]
    kc, vc = kc_ref.to(DEV), vc_ref.to(DEV)
    qkv_d = qkv.to(DEV)
    _, key_d, value_d = qkv_d.unbind(dim=1)
    kv_scale = 1.0 if kv_cache_dtype == "auto" else 0.25
    ops.reshape_and_cache(key_d, value_d, kc, vc, slots.to(DEV), kv_cache_dtype, kv_scale)
    oracle.reshape_and_cache(key, value, kc_ref, vc_ref, slots, kv_cache_dtype, kv_scale)
    assert torch.equal(_bits(kc.cpu()), _bits(kc_ref))
    assert torch.equal(_bits(vc.cpu()), _bits(vc_ref))


def test_reshape_and_cache_bad_kv_dtype(ops):
    k = torch.zeros(1, 2, 64, dtype=torch.half, device=DEV)
    kc = torch.zeros(2, 2, 8, 16, 8, dtype=torch.half, device=DEV)
    vc = torch.zeros(2, 2, 64, 16, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.reshape_and_cache(k, k, kc, vc, torch.zeros(1, dtype=torch.long, device=DEV), "int4", 1.0)


# ---------------------------------------------------------------------------------------- copy_blocks
@pytest.mark.parametrize("num_mappings", [256])
@pytest.mark.parametrize("num_layers", [1, 3])
@pytest.mark.parametrize("num_heads", [8])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("num_blocks", [1024])
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@torch.inference_mode()
def test_copy_blocks(ops, num_mappings, num_layers, num_heads, head_size, block_size, num_blocks, dtype,
                     kv_cache_dtype):
    """tests/kernels/test_cache.py:31-102 — each source copied to two destinations, mapping on device."""
    random.seed(0)
    torch.manual_seed(0)
    assert 2 * num_mappings <= num_blocks
    src_blocks = random.sample(range(num_blocks), num_mappings)
    remaining = list(set(range(num_blocks)) - set(src_blocks))
    dst_blocks = random.sample(remaining, 2 * num_mappings)
    block_mapping = []
    for i in range(num_mappings):
        block_mapping.append((src_blocks[i], dst_blocks[2 * i]))
        block_mapping.append((src_blocks[i], dst_blocks[2 * i + 1]))
    key_caches, value_caches = helpers.create_kv_caches_with_random(num_blocks, block_size, num_layers, num_heads,
                                                                    head_size, kv_cache_dtype, dtype, 0, DEV,
                                                                    ops.convert_fp8)
    cloned_k = [t.clone() for t in key_caches]
    cloned_v = [t.clone() for t in value_caches]
    bm = torch.tensor(block_mapping, dtype=torch.int64, device=DEV).view(-1, 2)
    ops.copy_blocks(key_caches, value_caches, bm)
    for src, dst in block_mapping:
        for c in cloned_k + cloned_v:
            c[dst].copy_(c[src])
    for a, b in zip(key_caches + value_caches, cloned_k + cloned_v):
        assert torch.equal(_bits(a), _bits(b))


@torch.inference_mode()
def test_copy_blocks_many_layers_and_vs_reference_cuda(ops, ref_cuda):
    """130 layers (more than one parameter pack of 120) and bit-equality with copy_blocks_kernel."""
    torch.manual_seed(0)
    L, NB = 130, 16
    kcs = [torch.randn(NB, 2, 8, 16, 8, device=DEV).to(torch.bfloat16) for _ in range(L)]
    vcs = [torch.randn(NB, 2, 64, 16, device=DEV).to(torch.bfloat16) for _ in range(L)]
    k2, v2 = [t.clone() for t in kcs], [t.clone() for t in vcs]
    bm = torch.tensor([[0, 9], [0, 10], [3, 11], [5, 15]], dtype=torch.int64, device=DEV)
    ops.copy_blocks(kcs, vcs, bm)
    ref_cuda.copy_blocks(k2, v2, bm)
    for a, b in zip(kcs + vcs, k2 + v2):
        assert torch.equal(_bits(a), _bits(b))


# ---------------------------------------------------------------------------------------- swap_blocks
@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cuda", "cuda"), ("cpu", "cuda")])
@pytest.mark.parametrize("pinned", [True, False])  # pinned host memory -> gather kernel, pageable -> memcpy runs
@pytest.mark.parametrize("num_mappings", [256])
@pytest.mark.parametrize("head_size", [64, 128])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", [torch.half, torch.float])
@pytest.mark.parametrize("kv_cache_dtype", ["auto", "fp8"])
@torch.inference_mode()
def test_swap_blocks(ops, oracle, direction, pinned, num_mappings, head_size, block_size, dtype, kv_cache_dtype):
    """tests/kernels/test_cache.py:274-344 — mapping on CPU, all three directions, exact."""
    random.seed(0)
    torch.manual_seed(0)
    num_blocks, num_heads = 1024, 8
    src_dev, dst_dev = direction
    # 40 pairs form one contiguous run (exercises run merging), the rest are random and disjoint from it
    run = 40
    src_pool = [b for b in range(num_blocks) if not 100 <= b < 100 + run]
    dst_pool = [b for b in range(num_blocks) if not 500 <= b < 500 + run]
    src_blocks = list(range(100, 100 + run)) + random.sample(src_pool, num_mappings - run)
    dst_blocks = list(range(500, 500 + run)) + random.sample(dst_pool, num_mappings - run)
    block_mapping = torch.tensor(list(zip(src_blocks, dst_blocks)), dtype=torch.int64, device="cpu").view(-1, 2)
    conv = ops.convert_fp8 if src_dev == "cuda" else oracle.convert_fp8
    src_k, src_v = helpers.create_kv_caches_with_random(num_blocks, block_size, 1, num_heads, head_size,
                                                        kv_cache_dtype, dtype, 0, src_dev, conv)
    conv = ops.convert_fp8 if dst_dev == "cuda" else oracle.convert_fp8
    dst_k, dst_v = helpers.create_kv_caches_with_random(num_blocks, block_size, 1, num_heads, head_size,
                                                        kv_cache_dtype, dtype, 1, dst_dev, conv)
    if pinned:
        if src_dev == "cpu":
            src_k, src_v = [src_k[0].pin_memory()], [src_v[0].pin_memory()]
        if dst_dev == "cpu":
            dst_k, dst_v = [dst_k[0].pin_memory()], [dst_v[0].pin_memory()]
    src_k_clone, src_v_clone = src_k[0].clone(), src_v[0].clone()
    dst_k_before = dst_k[0].clone()
    ops.swap_blocks(src_k[0], dst_k[0], block_mapping)
    ops.swap_blocks(src_v[0], dst_v[0], block_mapping)
    torch.cuda.synchronize()
    touched = set()
    for src, dst in block_mapping.tolist():
        if dst in touched:
            continue
        assert torch.equal(_bits(src_k_clone[src].cpu()), _bits(dst_k[0][dst].cpu()))
        assert torch.equal(_bits(src_v_clone[src].cpu()), _bits(dst_v[0][dst].cpu()))
        touched.add(dst)
    untouched = [b for b in range(num_blocks) if b not in set(dst_blocks)][:50]
    for b in untouched:
        assert torch.equal(_bits(dst_k_before[b].cpu()), _bits(dst_k[0][b].cpu()))


def test_swap_blocks_device_rules(ops):
    """cache_kernels.cu:27-45."""
    a = torch.zeros(4, 8, device=DEV)
    with pytest.raises(RuntimeError, match="block_mapping must be on CPU"):
        ops.swap_blocks(a, a.clone(), torch.zeros(1, 2, dtype=torch.int64, device=DEV))
    with pytest.raises(RuntimeError, match="Invalid device combination"):
        ops.swap_blocks(a.cpu(), a.cpu(), torch.zeros(1, 2, dtype=torch.int64))


# ---------------------------------------------------------------------------------------- convert_fp8
@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("scale", [1.0, 0.0418, 3.0])
@torch.inference_mode()
def test_fp8_conversion_bit_exact(ops, oracle, dtype, scale):
    """Both directions, bit-exact vs the pinned oracle; includes saturation, subnormals, -0, all codes."""
    torch.manual_seed(0)
    x = torch.empty(1 << 16, dtype=dtype).uniform_(-600, 600)
    x[:10] = torch.tensor([0.0, -0.0, 448.0, -448.0, 464.0, 480.0, 1e-3, -2e-3, 1.7e-2, -1e-9]).to(dtype)
    q_d = torch.empty(x.shape, dtype=torch.uint8, device=DEV)
    ops.convert_fp8(q_d, x.to(DEV), scale, "fp8")
    q = torch.empty(x.shape, dtype=torch.uint8)
    oracle.convert_fp8(q, x, scale, "fp8")
    assert torch.equal(q_d.cpu(), q)
    codes = torch.arange(256, dtype=torch.uint8).repeat(4)
    y_d = torch.empty(codes.shape, dtype=dtype, device=DEV)
    ops.convert_fp8(y_d, codes.to(DEV), scale, "fp8")
    y = torch.empty(codes.shape, dtype=dtype)
    oracle.convert_fp8(y, codes, scale, "fp8")
    yd = y_d.cpu()
    assert ((yd == y) | (yd.isnan() & y.isnan())).all()
    ok = ~y.isnan()
    assert torch.equal(_bits(yd[ok]), _bits(y[ok]))


@pytest.mark.parametrize("head_size", [64, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", [torch.half, torch.bfloat16, torch.float])
@torch.inference_mode()
def test_fp8_e4m3_conversion(ops, head_size, block_size, dtype):
    """tests/kernels/test_cache.py:347-380 — uniform(-224, 224) round trip, atol 1e-3 rtol 0.1."""
    torch.manual_seed(0)
    cache = torch.empty(128, 8, head_size, block_size, dtype=dtype, device=DEV).uniform_(-224.0, 224.0)
    cache_fp8 = torch.empty_like(cache, dtype=torch.uint8)
    ops.convert_fp8(cache_fp8, cache)
    converted = torch.empty_like(cache)
    ops.convert_fp8(converted, cache_fp8)
    assert torch.allclose(cache, converted, atol=0.001, rtol=0.1)


def test_convert_fp8_vs_reference_cuda(ops, ref_cuda):
    x = torch.empty(1 << 18, dtype=torch.bfloat16, device=DEV).uniform_(-500, 500)
    a = torch.empty(x.shape, dtype=torch.uint8, device=DEV)
    b = torch.empty_like(a)
    ops.convert_fp8(a, x, 0.7, "fp8")
    ref_cuda.convert_fp8(b, x, 0.7, "fp8")
    assert torch.equal(a, b)
    ya, yb = torch.empty_like(x), torch.empty_like(x)
    ops.convert_fp8(ya, a, 0.7, "fp8")
    ref_cuda.convert_fp8(yb, a, 0.7, "fp8")
    assert torch.equal(_bits(ya), _bits(yb))


def test_convert_fp8_rejects_auto(ops):
    x = torch.zeros(8, dtype=torch.half, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported data type"):
        ops.convert_fp8(torch.empty(8, dtype=torch.uint8, device=DEV), x, 1.0, "auto")


# ------------------------------------------------------------------------------- golden + direct C-ABI
@pytest.mark.parametrize("case", ["reshape_bf16", "reshape_f32"])
def test_reshape_and_cache_golden(ops, golden, case):
    """Against fixtures produced by the reference's own CPU reshape_and_cache (tests/golden)."""
    c = golden[case]
    qkv = c["qkv"].to(DEV)
    _, key, value = qkv.unbind(dim=1)
    kc, vc = c["key_cache_in"].to(DEV), c["value_cache_in"].to(DEV)
    ops.reshape_and_cache(key, value, kc, vc, c["slots"].to(DEV), "auto", 1.0)
    assert torch.equal(_bits(kc.cpu()), _bits(c["key_cache_out"]))
    assert torch.equal(_bits(vc.cpu()), _bits(c["value_cache_out"]))


def test_copy_blocks_golden_through_c_abi(golden):
    """Calls b200pa_copy_blocks directly through ctypes (no torch types in the signature)."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = ctypes.CDLL(os.path.join(root, "vllm_legacy_b200", "libb200pa.so"))
    c = golden["copy_bf16"]
    L = int(c["meta"][0])
    kcs = [c[f"key_cache_in.{i}"].to(DEV) for i in range(L)]
    vcs = [c[f"value_cache_in.{i}"].to(DEV) for i in range(L)]
    bm = c["mapping"].to(DEV)
    kp = (ctypes.c_void_p * L)(*[t.data_ptr() for t in kcs])
    vp = (ctypes.c_void_p * L)(*[t.data_ptr() for t in vcs])
    block_bytes = kcs[0][0].numel() * kcs[0].element_size()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b200pa_copy_blocks(kp, vp, ctypes.c_int32(L), ctypes.c_void_p(bm.data_ptr()),
                                ctypes.c_int64(bm.shape[0]), ctypes.c_int64(block_bytes), stream)
    assert rc == 0
    torch.cuda.synchronize()
    for i in range(L):
        assert torch.equal(_bits(kcs[i].cpu()), _bits(c[f"key_cache_out.{i}"]))
        assert torch.equal(_bits(vcs[i].cpu()), _bits(c[f"value_cache_out.{i}"]))


def test_get_max_shared_memory(ops):
    v = ops.get_max_shared_memory_per_block_device_attribute(0)
    assert v >= 200 * 1024  # B200: 227 KiB opt-in


# -------------------------------------------------------------------------- reshape_and_cache_flash
@pytest.mark.parametrize("num_tokens", [42])
@pytest.mark.parametrize("num_heads", [8])
@pytest.mark.parametrize("head_size", [64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("dtype", DTYPES)
@torch.inference_mode()
def test_reshape_and_cache_flash(ops, oracle, num_tokens, num_heads, head_size, block_size, dtype):
    """tests/kernels/test_cache.py:196-282 — flash layout [nb, B, H, D], strided K/V rows, random unique slots,
    plus a padding slot.  Bit-exact vs the oracle and vs the test's own index assignment."""
    random.seed(0)
    torch.manual_seed(0)
    num_blocks = 64
    slot_mapping = torch.tensor(random.sample(range(num_blocks * block_size), num_tokens), dtype=torch.long)
    slot_mapping[5] = -1
    qkv = torch.randn(num_tokens, 3, num_heads, head_size, dtype=dtype)
    _, key, value = qkv.unbind(dim=1)
    scale = head_size ** -0.5
    kc_ref = torch.empty(num_blocks, block_size, num_heads, head_size, dtype=dtype).uniform_(-scale, scale)
    vc_ref = torch.empty_like(kc_ref).uniform_(-scale, scale)
    kc, vc = kc_ref.to(DEV), vc_ref.to(DEV)
    _, key_d, value_d = qkv.to(DEV).unbind(dim=1)
    ops.reshape_and_cache_flash(key_d, value_d, kc, vc, slot_mapping.to(DEV), "auto")
    expect_k, expect_v = kc_ref.clone(), vc_ref.clone()
    oracle.reshape_and_cache_flash(key, value, kc_ref, vc_ref, slot_mapping, "auto")
    for i, s in enumerate(slot_mapping.tolist()):
        if s >= 0:
            expect_k[s // block_size, s % block_size] = key[i]
            expect_v[s // block_size, s % block_size] = value[i]
    assert torch.equal(_bits(kc_ref), _bits(expect_k)) and torch.equal(_bits(vc_ref), _bits(expect_v))
    assert torch.equal(_bits(kc.cpu()), _bits(kc_ref))
    assert torch.equal(_bits(vc.cpu()), _bits(vc_ref))


@torch.inference_mode()
def test_reshape_and_cache_flash_variants(ops, oracle, ref_cuda_optional):
    """Large prefill-sized call (grid-stride path), a block-strided cache view (block_stride > B*H*D), a
    misaligned source (scalar kernel), empty input and the fp8 rejection (cache_kernels.cu:303-306)."""
    torch.manual_seed(0)
    T, H, D, B, NB = 5000, 8, 128, 16, 400
    dtype = torch.bfloat16
    slots = torch.randperm(NB * B)[:T]
    key = torch.randn(T, H, D, dtype=dtype)
    value = torch.randn(T, H, D, dtype=dtype)
    big = torch.zeros(NB, 2, B, H, D, dtype=dtype)      # K and V interleaved per block: stride(0) = 2*B*H*D
    kc_ref, vc_ref = big[:, 0], big[:, 1]
    big_d = big.to(DEV)
    kc, vc = big_d[:, 0], big_d[:, 1]
    ops.reshape_and_cache_flash(key.to(DEV), value.to(DEV), kc, vc, slots.to(DEV), "auto")
    oracle.reshape_and_cache_flash(key, value, kc_ref, vc_ref, slots, "auto")
    assert torch.equal(_bits(big_d.cpu()), _bits(big))
    if ref_cuda_optional is not None:
        big2 = torch.zeros_like(big_d)
        ref_cuda_optional.reshape_and_cache_flash(key.to(DEV), value.to(DEV), str(big2[:, 0].data_ptr()), B,
                                                  big2.stride(0), str(big2[:, 1].data_ptr()), big2.stride(0),
                                                  slots.to(DEV), "auto")
        assert torch.equal(_bits(big2), _bits(big_d))
    # misaligned rows -> scalar kernel
    flat = torch.randn(2 * 7 * 3 * 64 + 8, dtype=dtype)
    k2, v2 = flat[1:1 + 7 * 192].view(7, 3, 64), flat[7 * 192 + 3:2 * 7 * 192 + 3].view(7, 3, 64)
    flat_d = flat.to(DEV)
    k2d, v2d = flat_d[1:1 + 7 * 192].view(7, 3, 64), flat_d[7 * 192 + 3:2 * 7 * 192 + 3].view(7, 3, 64)
    c_ref = torch.zeros(2, 4, 16, 3, 64, dtype=dtype)
    c_d = c_ref.to(DEV)
    s2 = torch.tensor([5, 0, 63, 17, -1, 31, 32], dtype=torch.long)
    ops.reshape_and_cache_flash(k2d, v2d, c_d[0], c_d[1], s2.to(DEV), "auto")
    oracle.reshape_and_cache_flash(k2, v2, c_ref[0], c_ref[1], s2, "auto")
    assert torch.equal(_bits(c_d.cpu()), _bits(c_ref))
    ops.reshape_and_cache_flash(k2d[:0], v2d[:0], c_d[0], c_d[1], s2[:0].to(DEV), "auto")
    assert torch.equal(_bits(c_d.cpu()), _bits(c_ref))
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.reshape_and_cache_flash(k2d, v2d, c_d[0], c_d[1], s2.to(DEV), "fp8")
