"""Host-side mirror of the reference interface (vllm/attention/ops/paged_attn.py, backends) and the
KV-head sharding used for N > 1 GPUs.  CPU only; device ops are replaced by a recorder."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vllm_legacy_b200.attention.backends.b200_paged import (B200PagedAttentionBackend,
                                                             B200PagedAttentionMetadata)
from vllm_legacy_b200.attention.ops import paged_attn as pa_mod
from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention
from vllm_legacy_b200.parallel import shard_heads


def test_kv_cache_shape_and_views():
    """paged_attn.py:37-59: (2, nb, B*H*D); K view [nb,H,D/x,B,x], V view [nb,H,D,B] share storage."""
    nb, B, H, D = 6, 16, 8, 128
    assert PagedAttention.get_kv_cache_shape(nb, B, H, D) == (2, nb, B * H * D)
    assert B200PagedAttentionBackend.get_kv_cache_shape(nb, B, H, D) == (2, nb, B * H * D)
    for dtype, x in ((torch.bfloat16, 8), (torch.float32, 4), (torch.uint8, 16)):
        kv = torch.zeros(2, nb, B * H * D, dtype=dtype)
        k, v = PagedAttention.split_kv_cache(kv, H, D)
        assert k.shape == (nb, H, D // x, B, x) and v.shape == (nb, H, D, B)
        assert k.data_ptr() == kv[0].data_ptr() and v.data_ptr() == kv[1].data_ptr()
        assert k.stride(0) == B * H * D and k.stride(1) == B * D
    assert PagedAttention.get_supported_head_sizes() == [64, 80, 96, 112, 128, 192, 256]


class _Recorder:
    def __init__(self):
        self.calls = []

    def __getattr__(self, name):
        def f(*args, **kw):
            self.calls.append((name, args, kw))
        return f


@pytest.fixture
def recorder(monkeypatch):
    r = _Recorder()
    monkeypatch.setattr(pa_mod, "ops", r)
    return r


@pytest.mark.parametrize("num_seqs,num_heads,max_len,want", [
    (16, 32, 1024, "paged_attention_v2"),    # BASELINE config 1: S*H == 512 is not > 512
    (17, 32, 1024, "paged_attention_v1"),    # S*H > 512
    (256, 32, 4096, "paged_attention_v1"),   # BASELINE config 2: the heuristic picks V1
    (4, 8, 4096, "paged_attention_v2"),      # few (seq, head) pairs, several partitions
    (4, 8, 512, "paged_attention_v1"),       # one partition
    (256, 32, 8193, "paged_attention_v2"),   # > 8192 always V2
])
def test_forward_decode_v1_v2_rule(recorder, num_seqs, num_heads, max_len, want):
    """paged_attn.py:111-121 heuristic and :148-158 scratch shapes."""
    D, B, Hkv = 128, 16, 8
    q = torch.zeros(num_seqs, num_heads, D)
    kc = torch.zeros(4, Hkv, D // 4, B, 4)
    vc = torch.zeros(4, Hkv, D, B)
    bt = torch.zeros(num_seqs, 4, dtype=torch.int32)
    sl = torch.ones(num_seqs, dtype=torch.int32)
    out = PagedAttention.forward_decode(q, kc, vc, bt, sl, max_len, "auto", Hkv, 0.1, None, 1.0)
    assert out.shape == q.shape
    (name, args, _), = recorder.calls
    assert name == want
    if want == "paged_attention_v2":
        P = (max_len + 511) // 512
        _, exp_sums, max_logits, tmp_out = args[:4]
        assert exp_sums.shape == (num_seqs, num_heads, P) and exp_sums.dtype == torch.float32
        assert max_logits.shape == exp_sums.shape
        assert tmp_out.shape == (num_seqs, num_heads, P, D)
        assert args[12] == max_len and args[11] == B
    else:
        assert args[9] == max_len and args[8] == B


def test_write_swap_copy_route_to_ops(recorder):
    H, D, B = 2, 64, 16
    kv = [torch.zeros(2, 4, B * H * D) for _ in range(3)]
    PagedAttention.copy_blocks(kv, torch.zeros(1, 2, dtype=torch.int64))
    PagedAttention.swap_blocks(kv[0], kv[1], torch.zeros(1, 2, dtype=torch.int64))
    k, v = PagedAttention.split_kv_cache(kv[0], H, D)
    PagedAttention.write_to_paged_cache(torch.zeros(5, H, D), torch.zeros(5, H, D), k, v,
                                        torch.arange(5).view(5, 1), "auto", 1.0)
    names = [c[0] for c in recorder.calls]
    assert names == ["copy_blocks", "swap_blocks", "swap_blocks", "reshape_and_cache"]
    key_caches, value_caches, _ = recorder.calls[0][1]
    assert len(key_caches) == 3 and key_caches[0].data_ptr() == kv[0][0].data_ptr()
    assert value_caches[2].data_ptr() == kv[2][1].data_ptr()
    assert recorder.calls[3][1][4].shape == (5,)  # slot_mapping flattened (paged_attn.py:76)


def test_backend_impl_decode_path(recorder):
    """Backend shim = reshape_and_cache + forward_decode, like xformers.py:269-335."""
    Hq, Hkv, D, B, nb, S = 8, 2, 64, 16, 4, 3
    impl = B200PagedAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto")
    meta = B200PagedAttentionBackend.make_metadata(
        seq_lens_tensor=torch.tensor([3, 5, 9], dtype=torch.int32), max_decode_seq_len=9,
        block_tables=torch.zeros(S, 1, dtype=torch.int32), num_prefills=0, num_prefill_tokens=0,
        num_decode_tokens=S, slot_mapping=torch.tensor([2, 4, 8]))
    assert isinstance(meta, B200PagedAttentionMetadata)
    assert meta.prefill_metadata is None and meta.decode_metadata is meta
    kv_cache = torch.zeros(*B200PagedAttentionBackend.get_kv_cache_shape(nb, B, Hkv, D))
    out = impl.forward(torch.zeros(S, Hq * D), torch.zeros(S, Hkv * D), torch.zeros(S, Hkv * D), kv_cache, meta, 1.0)
    assert out.shape == (S, Hq * D)
    assert [c[0] for c in recorder.calls] == ["reshape_and_cache", "paged_attention_v1"]
    with pytest.raises(ValueError):
        B200PagedAttentionBackend.get_impl_cls()(8, 100, 1.0)


def test_backend_impl_mixed_prefill_decode(recorder):
    """Prefill rows first, then decode rows; metadata split like XFormersMetadata (xformers.py:120-177);
    prompts with a cached prefix go to context_attention_fwd, decode rows to paged attention."""
    Hq, Hkv, D, B, nb = 8, 2, 64, 16, 8
    impl = B200PagedAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto")
    # 2 prompts (5 and 3 new tokens, contexts 16 and 0) + 2 decode rows
    meta = B200PagedAttentionBackend.make_metadata(
        num_prefills=2, num_prefill_tokens=8, num_decode_tokens=2, slot_mapping=torch.arange(10),
        seq_lens=[21, 3, 7, 9], seq_lens_tensor=torch.tensor([21, 3, 7, 9], dtype=torch.int32),
        max_query_len=5, max_prefill_seq_len=21, max_decode_seq_len=9,
        query_start_loc=torch.tensor([0, 5, 8, 9, 10], dtype=torch.int32), seq_start_loc=None,
        context_lens_tensor=torch.tensor([16, 0, 6, 8], dtype=torch.int32),
        block_tables=torch.arange(8, dtype=torch.int32).view(4, 2), use_cuda_graph=False)
    pm, dm = meta.prefill_metadata, meta.decode_metadata
    assert pm.num_prefills == 2 and pm.num_decode_tokens == 0 and pm.seq_lens == [21, 3]
    assert pm.query_start_loc.tolist() == [0, 5, 8] and pm.context_lens_tensor.tolist() == [16, 0]
    assert pm.block_tables.tolist() == [[0, 1], [2, 3]] and pm.slot_mapping.tolist() == list(range(8))
    assert dm.num_prefills == 0 and dm.seq_lens_tensor.tolist() == [7, 9] and dm.block_tables.tolist() == [[4, 5], [6, 7]]
    assert dm.slot_mapping.tolist() == [8, 9] and dm.max_decode_seq_len == 9
    assert meta.prefill_metadata is pm and meta.decode_metadata is dm   # cached
    kv_cache = torch.zeros(*B200PagedAttentionBackend.get_kv_cache_shape(nb, B, Hkv, D))
    out = impl.forward(torch.zeros(10, Hq * D), torch.zeros(10, Hkv * D), torch.zeros(10, Hkv * D), kv_cache, meta, 1.0)
    assert out.shape == (10, Hq * D)
    assert [c[0] for c in recorder.calls] == ["reshape_and_cache", "context_attention_fwd", "paged_attention_v1"]
    pf = recorder.calls[1][1]
    assert pf[0].shape == (8, Hq, D) and pf[1].shape == (8, Hkv, D) and pf[3].shape == (8, Hq, D)
    assert pf[7].tolist() == [0, 5] and pf[8].tolist() == [21, 3] and pf[9].tolist() == [16, 0] and pf[10] == 5
    assert recorder.calls[2][1][1].shape == (2, Hq, D)
    # prompt run without any cache (profiling run): plain causal attention = the same kernel with context 0
    recorder.calls.clear()
    meta2 = B200PagedAttentionBackend.make_metadata(
        num_prefills=2, num_prefill_tokens=8, num_decode_tokens=0, slot_mapping=torch.arange(8),
        seq_lens=[5, 3], seq_lens_tensor=torch.tensor([5, 3], dtype=torch.int32), max_query_len=5,
        max_prefill_seq_len=5, max_decode_seq_len=0, query_start_loc=torch.tensor([0, 5, 8], dtype=torch.int32),
        context_lens_tensor=torch.zeros(2, dtype=torch.int32), block_tables=torch.zeros(2, 0, dtype=torch.int32))
    impl.forward(torch.zeros(8, Hq * D), torch.zeros(8, Hkv * D), torch.zeros(8, Hkv * D), None, meta2, 1.0)
    (name, args, _), = recorder.calls
    assert name == "context_attention_fwd" and args[8].tolist() == [5, 3] and args[9].tolist() == [0, 0]


def test_backend_fp8_cache_prefill_paths(recorder):
    """fp8 KV cache: a prompt with a cached prefix hands the cache dtype and kv_scale to the prefix kernel (which
    dequantises while staging); a prompt without one never reads the cache (xformers.py:291-300) and runs on a dummy
    16-bit cache with kv_cache_dtype "auto"."""
    Hq, Hkv, D, B, nb = 8, 2, 64, 16, 8
    impl = B200PagedAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "fp8")
    kv_cache = torch.zeros(*B200PagedAttentionBackend.get_kv_cache_shape(nb, B, Hkv, D), dtype=torch.uint8)
    meta = B200PagedAttentionBackend.make_metadata(
        num_prefills=1, num_prefill_tokens=5, num_decode_tokens=1, slot_mapping=torch.arange(6),
        seq_lens=[21, 7], seq_lens_tensor=torch.tensor([21, 7], dtype=torch.int32), max_query_len=5,
        max_prefill_seq_len=21, max_decode_seq_len=7, query_start_loc=torch.tensor([0, 5, 6], dtype=torch.int32),
        context_lens_tensor=torch.tensor([16, 6], dtype=torch.int32),
        block_tables=torch.arange(4, dtype=torch.int32).view(2, 2), use_cuda_graph=False)
    q = torch.zeros(6, Hq * D, dtype=torch.bfloat16)
    kv = torch.zeros(6, Hkv * D, dtype=torch.bfloat16)
    impl.forward(q, kv, kv.clone(), kv_cache, meta, 0.5)
    names = [c[0] for c in recorder.calls]
    assert names == ["reshape_and_cache", "context_attention_fwd", "paged_attention_v1"]
    pf = recorder.calls[1][1]
    assert pf[4].dtype == torch.uint8 and pf[4].shape[-1] == 16   # K cache view with x = 16
    assert pf[13] == "fp8" and pf[14] == 0.5
    recorder.calls.clear()
    meta2 = B200PagedAttentionBackend.make_metadata(
        num_prefills=1, num_prefill_tokens=5, num_decode_tokens=0, slot_mapping=torch.arange(5),
        seq_lens=[5], seq_lens_tensor=torch.tensor([5], dtype=torch.int32), max_query_len=5, max_prefill_seq_len=5,
        max_decode_seq_len=0, query_start_loc=torch.tensor([0, 5], dtype=torch.int32),
        context_lens_tensor=torch.zeros(1, dtype=torch.int32), block_tables=torch.zeros(1, 0, dtype=torch.int32))
    impl.forward(q[:5], kv[:5], kv[:5].clone(), kv_cache, meta2, 0.5)
    assert [c[0] for c in recorder.calls] == ["reshape_and_cache", "context_attention_fwd"]
    pf = recorder.calls[1][1]
    assert pf[4].dtype == torch.bfloat16 and pf[13] == "auto"


def test_shard_heads_matches_reference_rules():
    """vllm/config.py:320-333."""
    s = [shard_heads(64, 8, 8, r) for r in range(8)]
    assert all(x.num_heads == 8 and x.num_kv_heads == 1 for x in s)
    assert [x.kv_head_start for x in s] == list(range(8))
    assert [x.q_head_start for x in s] == list(range(0, 64, 8))
    s = [shard_heads(64, 8, 2, r) for r in range(2)]
    assert [(x.num_heads, x.num_kv_heads, x.kv_head_start) for x in s] == [(32, 4, 0), (32, 4, 4)]
    s = [shard_heads(32, 2, 4, r) for r in range(4)]  # fewer KV heads than ranks: replicate
    assert [(x.num_kv_heads, x.kv_head_start, x.kv_replicas) for x in s] == [(1, 0, 2), (1, 0, 2), (1, 1, 2), (1, 1, 2)]
    with pytest.raises(ValueError):
        shard_heads(30, 6, 4, 0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard_worker(rank, world, port, q):
    """Each rank attends over its KV-head shard with NO data-path collective; the gloo all_gather here
    only collects results for the check, like the e2e TP step would after o_proj."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import oracle
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)  # replicated inputs, as the driver broadcasts them
        S, Hq, Hkv, D, B, NB, L = 2, 8, 4, 64, 16, 16, 100
        scale = D ** -0.5
        qf = torch.empty(S, Hq, D).uniform_(-scale, scale)
        kc = torch.empty(NB, Hkv, D // 4, B, 4).uniform_(-scale, scale)
        vc = torch.empty(NB, Hkv, D, B).uniform_(-scale, scale)
        bt = torch.randint(0, NB, (S, (L + B - 1) // B), dtype=torch.int32)
        sl = torch.tensor([L, 33], dtype=torch.int32)
        sh = shard_heads(Hq, Hkv, world, rank)
        q_loc = qf[:, sh.q_head_start:sh.q_head_start + sh.num_heads].contiguous()
        kc_loc = kc[:, sh.kv_head_start:sh.kv_head_start + sh.num_kv_heads].contiguous()
        vc_loc = vc[:, sh.kv_head_start:sh.kv_head_start + sh.num_kv_heads].contiguous()
        out_loc = torch.empty_like(q_loc)
        oracle.paged_attention_v1(out_loc, q_loc, kc_loc, vc_loc, sh.num_kv_heads, scale, bt, sl, B, L, None, "auto", 1.0)
        parts = [torch.empty_like(out_loc) for _ in range(world)]
        dist.all_gather(parts, out_loc)
        if rank == 0:
            full = torch.empty_like(qf)
            oracle.paged_attention_v1(full, qf, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0)
            q.put(bool(torch.equal(torch.cat(parts, dim=1), full)))
    finally:
        dist.destroy_process_group()


def test_kv_head_sharding_world2_gloo():
    """world_size 2 on CPU/gloo: concatenating the per-rank outputs over heads == unsharded result,
    bit for bit, with no exchange on the data path."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True


def test_hip_backend_dense_paths(recorder, monkeypatch):
    """The fork's own backend (hip.py:206-209, 306-319): dense layers and short contexts decode through
    PagedAttention; everything that needs the external `hip` package raises instead of being replaced."""
    from vllm_legacy_b200.attention.backends.hip import (HiPAttentionBackend, HiPAttentionMetadata,
                                                        HiPSparsePathUnavailable)
    Hq, Hkv, D, B, nb, S = 8, 2, 64, 16, 4, 3
    assert HiPAttentionBackend.get_kv_cache_shape(nb, B, Hkv, D) == (2, nb, B * Hkv * D)
    with pytest.raises(AssertionError):  # layer_index is mandatory in the fork (hip.py:103)
        HiPAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto")

    def meta(max_len):
        return HiPAttentionBackend.make_metadata(
            seq_lens_tensor=torch.tensor([3, 5, 9], dtype=torch.int32), max_decode_seq_len=max_len,
            block_tables=torch.zeros(S, 1, dtype=torch.int32), num_prefills=0, num_prefill_tokens=0,
            num_decode_tokens=S, slot_mapping=torch.tensor([2, 4, 8]))
    assert isinstance(meta(9), HiPAttentionMetadata)
    kv_cache = torch.zeros(*HiPAttentionBackend.get_kv_cache_shape(nb, B, Hkv, D))
    args = (torch.zeros(S, Hq * D), torch.zeros(S, Hkv * D), torch.zeros(S, Hkv * D), kv_cache)
    # a dense layer (index < HIP_DENSE_LAYERS = 3): any context length
    dense = HiPAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto", None, layer_index=1)
    assert dense.forward(*args, meta(100000)).shape == (S, Hq * D)
    assert [c[0] for c in recorder.calls] == ["reshape_and_cache", "paged_attention_v2"]
    # a sparse layer: PagedAttention while max_decode_seq_len < 3 * HIP_K, the external package beyond
    recorder.calls.clear()
    sparse = HiPAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto", None, layer_index=7)
    assert sparse.forward(*args, meta(3 * 1024 - 1)).shape == (S, Hq * D)
    assert [c[0] for c in recorder.calls] == ["reshape_and_cache", "paged_attention_v2"]
    with pytest.raises(HiPSparsePathUnavailable):
        sparse.forward(*args, meta(3 * 1024))
    monkeypatch.setenv("HIP_K", "64")
    assert HiPAttentionBackend.get_impl_cls()(Hq, D, 0.125, Hkv, None, None, "auto", None, layer_index=7).hip_k == 64
    with pytest.raises(AssertionError):
        dense.forward(*args, meta(9), 0.5)  # kv_scale != 1 is rejected by the fork (hip.py:143)


def test_selector(monkeypatch):
    """vllm/attention/selector.py:24-167: env override, the fork's HIP_ATTN entry, the reference's FLASH_ATTN exclusions."""
    from vllm_legacy_b200.attention import selector
    from vllm_legacy_b200.attention.backends.hip import HiPAttentionBackend
    assert [m.name for m in selector._Backend] == ["FLASH_ATTN", "XFORMERS", "ROCM_FLASH", "TORCH_SDPA", "FLASHINFER",
                                                   "HIP_ATTN"]
    a = (32, 128, 8, None, torch.bfloat16, "auto", 16)
    monkeypatch.delenv("VLLM_ATTENTION_BACKEND", raising=False)
    assert selector.which_attn_to_use(*a) == selector._Backend.FLASH_ATTN
    assert selector.which_attn_to_use(32, 128, 8, None, torch.bfloat16, "fp8", 16) == selector._Backend.XFORMERS
    assert selector.which_attn_to_use(32, 128, 8, None, torch.float32, "auto", 16) == selector._Backend.XFORMERS
    assert selector.which_attn_to_use(32, 128, 8, 4096, torch.bfloat16, "auto", 16) == selector._Backend.XFORMERS
    assert selector.which_attn_to_use(32, 128, 8, None, torch.bfloat16, "auto", 8) == selector._Backend.XFORMERS
    assert selector.which_attn_to_use(32, 80, 8, None, torch.bfloat16, "auto", 16) == selector._Backend.XFORMERS
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "HIP_ATTN")
    selector.get_attn_backend.cache_clear()
    assert selector.get_attn_backend(*a) is HiPAttentionBackend
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "XFORMERS")
    selector.get_attn_backend.cache_clear()
    assert selector.get_attn_backend(*a) is B200PagedAttentionBackend
    assert selector.get_attn_backend(*a, is_blocksparse=True) is B200PagedAttentionBackend
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "ROCM_FLASH")
    selector.get_attn_backend.cache_clear()
    with pytest.raises(ValueError, match="sm_100a"):
        selector.get_attn_backend(*a)
    monkeypatch.setenv("VLLM_ATTENTION_BACKEND", "NOPE")
    with pytest.raises(ValueError, match="Invalid attention backend"):
        selector.which_attn_to_use(*a)
    selector.get_attn_backend.cache_clear()


def test_flash_backend_interface(monkeypatch):
    """flash_attn.py:18-67: cache shape, block-size rule, head sizes; the selector hands it out for FLASH_ATTN."""
    from vllm_legacy_b200.attention import selector
    from vllm_legacy_b200.attention.backends.b200_flash import B200FlashAttentionBackend as FB
    assert FB.get_kv_cache_shape(7, 16, 8, 128) == (2, 7, 16, 8, 128)
    with pytest.raises(ValueError, match="multiple of 16"):
        FB.get_kv_cache_shape(7, 8, 8, 128)
    assert FB.get_supported_head_sizes() == [32, 64, 96, 128, 160, 192, 224, 256]
    with pytest.raises(ValueError, match="Head size 80"):
        FB.get_impl_cls()(8, 80, 1.0, 2, None, None, "auto")
    with pytest.raises(ValueError, match="Sliding window"):
        FB.get_impl_cls()(8, 64, 1.0, 2, None, 128, "auto")
    monkeypatch.delenv("VLLM_ATTENTION_BACKEND", raising=False)
    selector.get_attn_backend.cache_clear()
    assert selector.get_attn_backend(32, 128, 8, None, torch.bfloat16, "auto", 16) is FB
    selector.get_attn_backend.cache_clear()
