#!/usr/bin/env python3
"""Benchmark of the B200 paged-attention decode hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one decode-attention layer-call (paged_attention_v2) over the whole synthetic batch.

  N = 1   BASELINE.json configs[1]: Llama-3-8B GQA decode, 32 q / 8 kv heads, head 128, block 16, batch 256 x 4096
          context, bf16 KV (the configuration the metric is quoted on).  `extra.configs` carries the other BASELINE
          configs (1, 3-shards, 4, 5 + cache ops), each with algorithmic bytes, ms and roofline fraction.
  N > 1   BASELINE.json configs[2]: Llama-3-70B layout, 64 q / 8 kv heads, batch 256 x 8192 context, sharded by KV head
          over the N ranks (vllm/config.py:320-333 -> (32,4) / (16,2) / (8,1) heads per GPU at 2 / 4 / 8): the total
          work is fixed, STRONG scaling, `value` = 8.59 GB of KV (all ranks) / max-over-ranks device time.  No collective
          on the data path.  `extra.weak` keeps the replica run (config 2 on every rank) and `extra.tp_step` times the
          end-to-end TP decode step: fused rope + reshape_and_cache -> attention on the shard -> NCCL all-reduce of the
          [256, 8192] bf16 activations (vllm/model_executor/layers/linear.py:764-769), all-reduce time broken out.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for how each field is produced.
`--impl reference` times the reference's own CPU implementation (oracle/_ref, csrc/cpu/attention.cpp
compiled unchanged) on the host cores on a bounded sample of the same workload.
"""
import argparse
import ctypes
import importlib.util
import json
import math
import os
import random
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decode_attn_kv_gbps"
UNIT = "GB/s"

# BASELINE.json configs[1] and configs[2] (whole model; sharded by KV head for N > 1)
CFG2 = dict(num_seqs=256, ctx=4096, num_heads=32, num_kv_heads=8, head_size=128, block_size=16)
CFG3 = dict(num_seqs=256, ctx=8192, num_heads=64, num_kv_heads=8, head_size=128, block_size=16)
WORKLOAD2 = ("Llama-3-8B GQA decode attention (32q/8kv, head 128), block 16, batch 256 x 4096 ctx, bf16, "
             "paged_attention_v2, distinct physical blocks (random permutation)")
WORKLOAD3 = ("Llama-3-70B layout decode attention (64q/8kv, head 128) sharded by KV head, block 16, batch 256 x 8192 "
             "ctx, bf16, paged_attention_v2, distinct physical blocks (random permutation)")


def bench_config(world):
    """The `config` object of the JSON line — identical for both arms (--impl ours / reference) at a given N."""
    if world == 1:
        return {"workload": WORKLOAD2, "per_gpu": dict(CFG2), "sharding": "single GPU",
                "l2": "inputs larger than L2 (4.3 GB KV streamed once per step)"}
    per = dict(CFG3)
    per["num_heads"] //= world
    per["num_kv_heads"] = max(1, per["num_kv_heads"] // world)
    return {"workload": WORKLOAD3, "per_gpu": per,
            "sharding": f"kv-head shards of the 64q/8kv model, tp={world}, strong scaling, no collective",
            "l2": f"inputs larger than L2 ({8.59 / world:.2f} GB KV per GPU streamed once per step)"}


def algorithmic_bytes(num_seqs, ctx, num_heads, num_kv_heads, head_size, cache_elem_bytes=2, io_elem_bytes=2,
                      total_tokens=None):
    """sum(seq_len) * Hkv * D * 2 (K+V) * sizeof(cache_t) + q read + out write."""
    tokens = num_seqs * ctx if total_tokens is None else total_tokens
    kv = tokens * num_kv_heads * head_size * 2 * cache_elem_bytes
    qo = 2 * num_seqs * num_heads * head_size * io_elem_bytes
    return kv + qo


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
        return {}


def measured_peak():
    p = measured_peaks()
    if "hbm_gbs" in p:
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes (read+write) per launch of the dominant kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get("dram_bytes_per_launch")
    except Exception:  # noqa: BLE001
        return None


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "applications_clocks_setting": 0x2, "sync_boost": 0x10}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.004)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(1.0)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# workload construction (product arm: nothing here touches tests/ or oracle/)
# ------------------------------------------------------------------------------------------------
def ragged_lens(n=64, lo=128, hi=131072, seed=0):
    """BASELINE config 5: log-uniform lengths with both extremes present."""
    rng = random.Random(seed)
    lens = [int(round(math.exp(rng.uniform(math.log(lo), math.log(hi))))) for _ in range(n)]
    lens[0], lens[1] = hi, lo
    return lens


def make_decode_case(dev, lens, Hq, Hkv, D, B, kvd="auto", seed=0, spare_blocks=0, dtype=torch.bfloat16):
    """Paged KV cache resident in HBM whose block tables are a random permutation of DISTINCT physical blocks (the
    working set equals the algorithmic bytes; the reference benchmark's 1024 reused blocks would sit in L2)."""
    from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention
    g = torch.Generator(device=dev).manual_seed(seed)
    S, L = len(lens), max(lens)
    scale = float(D ** -0.5)
    nblk = [(l + B - 1) // B for l in lens]
    NB = sum(nblk) + S + spare_blocks
    cdt = torch.uint8 if kvd != "auto" else dtype
    eb = torch.empty(0, dtype=dtype).element_size()
    kv_cache = torch.empty(PagedAttention.get_kv_cache_shape(NB, B, Hkv, D), dtype=cdt, device=dev)
    if kvd != "auto":
        kv_cache.random_(0, 120, generator=g)  # finite e4m3 codes
    else:
        kv_cache.uniform_(-scale, scale, generator=g)
    perm = torch.randperm(NB, generator=g, device=dev).to(torch.int32)
    if len(set(nblk)) == 1:
        block_tables = perm[:S * nblk[0]].view(S, nblk[0]).contiguous()
    else:
        block_tables = torch.zeros(S, max(nblk), dtype=torch.int32, device=dev)
        o = 0
        for i, n in enumerate(nblk):
            block_tables[i, :n] = perm[o:o + n]
            o += n
    seq_lens = torch.tensor(lens, dtype=torch.int32, device=dev)
    q = torch.empty(S, Hq, D, dtype=dtype, device=dev).uniform_(-scale, scale, generator=g)
    key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, Hkv, D)
    P = (L + 511) // 512
    return dict(kv_cache=kv_cache, key_cache=key_cache, value_cache=value_cache, block_tables=block_tables,
                seq_lens=seq_lens, q=q, out=torch.empty_like(q), scale=scale, S=S, L=L, Hq=Hq, Hkv=Hkv, D=D, B=B,
                kvd=kvd, NB=NB, lens=lens,
                tmp_out=torch.empty(S, Hq, P, D, dtype=q.dtype, device=dev),
                exp_sums=torch.empty(S, Hq, P, dtype=torch.float32, device=dev),
                max_logits=torch.empty(S, Hq, P, dtype=torch.float32, device=dev),
                nbytes=algorithmic_bytes(S, L, Hq, Hkv, D, 1 if kvd != "auto" else eb, eb, total_tokens=sum(lens)))


def attention_call(ops, c, version="v2", blocksparse=None):
    if blocksparse is not None:  # (local_blocks, vert_stride, block_size, head_sliding_step), tp_rank 0
        return lambda: ops.paged_attention_v2(c["out"], c["exp_sums"], c["max_logits"], c["tmp_out"], c["q"],
                                              c["key_cache"], c["value_cache"], c["Hkv"], c["scale"], c["block_tables"],
                                              c["seq_lens"], c["B"], c["L"], None, c["kvd"], 1.0, 0, *blocksparse)
    if version == "v1":
        return lambda: ops.paged_attention_v1(c["out"], c["q"], c["key_cache"], c["value_cache"], c["Hkv"], c["scale"],
                                              c["block_tables"], c["seq_lens"], c["B"], c["L"], None, c["kvd"], 1.0)
    return lambda: ops.paged_attention_v2(c["out"], c["exp_sums"], c["max_logits"], c["tmp_out"], c["q"],
                                          c["key_cache"], c["value_cache"], c["Hkv"], c["scale"], c["block_tables"],
                                          c["seq_lens"], c["B"], c["L"], None, c["kvd"], 1.0)


def time_device(fn, iters, warmup=3, flush=None):
    """Mean device time per call in ms: CUDA events on the current stream around `iters` back-to-back calls (inputs
    larger than L2), or per call with an L2 flush in between (`flush` = a buffer larger than L2) for small inputs."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if flush is None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def host_step_inputs(c, dev):
    """Pinned host buffers of one decode step (q, new k/v, slot of the new token, block tables, seq_lens)."""
    S, L, Hq, Hkv, D, B = c["S"], c["L"], c["Hq"], c["Hkv"], c["D"], c["B"]
    scale = c["scale"]
    h = {
        "query": torch.empty(S, Hq * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "key": torch.empty(S, Hkv * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "value": torch.empty(S, Hkv * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "block_tables": c["block_tables"].cpu().pin_memory(),
        "seq_lens": c["seq_lens"].cpu().pin_memory(),
    }
    last_blk = c["block_tables"][:, (L - 1) // B].long().cpu()
    h["slot_mapping"] = (last_blk * B + (L - 1) % B).to(torch.int64).pin_memory()
    h["out"] = torch.empty(S, Hq * D, dtype=torch.bfloat16).pin_memory()
    return h


def read_stream_gbps(ops, dev, nbytes=4 << 30, iters=10):
    """Read-only HBM stream rate of this GPU, measured now with a plain 128-bit streaming-load kernel."""
    buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    buf.zero_()
    sink = torch.zeros(1, dtype=torch.int32, device=dev)
    ms = time_device(lambda: ops.read_stream_probe(buf, sink), iters)
    return nbytes / ms / 1e6


# ------------------------------------------------------------------------------------------------
# product arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from vllm_legacy_b200 import _custom_ops as ops
    from vllm_legacy_b200.attention.backends.b200_paged import B200PagedAttentionBackend as Backend
    from vllm_legacy_b200.parallel import shard_heads

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world == 1:
        cfg = dict(CFG2)
        total_bytes = algorithmic_bytes(**{k: cfg[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads",
                                                                "head_size")})
    else:
        sh = shard_heads(CFG3["num_heads"], CFG3["num_kv_heads"], world, rank)
        cfg = dict(CFG3, num_heads=sh.num_heads, num_kv_heads=sh.num_kv_heads)
        total_bytes = algorithmic_bytes(**{k: CFG3[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads",
                                                                 "head_size")})
    S, L, Hq, Hkv, D, B = (cfg[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads", "head_size", "block_size"))
    c = make_decode_case(dev, [L] * S, Hq, Hkv, D, B, seed=rank)
    step = attention_call(ops, c, "v2")
    bytes_per_launch = c["nbytes"]  # this rank's shard

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- kernel-only arm: inputs resident in HBM -------------------------------------------------
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        step()
    path = ops.last_attention_path()
    launches_per_step = 1 if path == "mma_tma" else (1 + (1 if ops.last_attention_partitions() > 1 else 0))
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local_rank) as clocks:
        ev[0].record()
        for i in range(args.steps):
            step()
            ev[i + 1].record()
        barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    per_step = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps))
    avg_kernel_ms = total_ms / args.steps  # step == one launch of the dominant kernel, back to back

    # ---- end-to-end arm: host buffers -> backend API -> host result ---------------------------------
    impl = Backend.get_impl_cls()(Hq, D, c["scale"], Hkv, None, None, "auto")
    h = host_step_inputs(c, dev)
    d_in = {k: torch.empty_like(h[k], device=dev) for k in ("query", "key", "value", "block_tables", "seq_lens",
                                                            "slot_mapping")}
    h2d = sum(h[k].numel() * h[k].element_size() for k in d_in)
    d2h = h["out"].numel() * h["out"].element_size()

    def e2e_step():
        for k in d_in:
            d_in[k].copy_(h[k], non_blocking=True)
        meta = Backend.make_metadata(seq_lens_tensor=d_in["seq_lens"], max_decode_seq_len=L,
                                     block_tables=d_in["block_tables"], num_prefills=0, num_prefill_tokens=0,
                                     num_decode_tokens=S, slot_mapping=d_in["slot_mapping"])
        o = impl.forward(d_in["query"], d_in["key"], d_in["value"], c["kv_cache"], meta, 1.0)
        h["out"].copy_(o, non_blocking=True)

    def timed_e2e(one_step):
        for _ in range(3):
            one_step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            one_step()
        e1.record()
        barrier()
        return e0.elapsed_time(e1)

    e2e_eager_ms = timed_e2e(e2e_step)
    e2e_path = ops.last_attention_path()
    # The same step — H2D copies from the pinned buffers, backend forward, D2H copy — captured ONCE in a CUDA graph and
    # replayed per step, which is how the reference runs decode (CUDAGraphRunner, vllm/worker/model_runner.py:1066-1158:
    # persistent input buffers, one replay per step).  Every step still performs every copy and every kernel; what goes
    # away is the Python / launch overhead of the eager loop, which on a slow or busy host (not on the GPU) decides
    # whether the GPU is kept fed: eager measured 0.75 ms/step on one box and 0.99 on another for identical GPU work.
    e2e_mode, e2e_ms = "eager", e2e_eager_ms
    try:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            e2e_step()
        e2e_ms = timed_e2e(graph.replay)
        e2e_mode = "cuda_graph"
    except Exception as exc:  # noqa: BLE001 - keep the eager figure
        sys.stderr.write(f"bench: CUDA-graph capture of the e2e step failed ({type(exc).__name__}: {exc}); eager loop kept\n")

    extra = {}
    if world > 1:
        extra["tp_step"] = tp_step_extra(ops, dist, c, cfg, dev, rank, world, min(args.steps, 50))
        c.clear(), h.clear(), d_in.clear()  # the closures above keep the dicts, not the tensors
        torch.cuda.empty_cache()
        extra["weak"] = weak_extra(ops, dist, dev, rank, world, min(args.steps, 50))

    # ---- max over ranks ------------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_ms, e2e_eager_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, e2e_eager_ms = t.tolist()
    if rank != 0:
        return None
    gbps = total_bytes * args.steps / (total_ms * 1e-3) / 1e9
    e2e_gbps = total_bytes * args.steps / (e2e_ms * 1e-3) / 1e9
    peak, peak_src = measured_peak()
    achieved = bytes_per_launch / (avg_kernel_ms * 1e-3) / 1e9
    try:
        read_peak = round(read_stream_gbps(ops, dev), 1)
    except Exception:  # noqa: BLE001
        read_peak = None
    line = {
        "metric": METRIC, "value": round(gbps, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": round(total_ms / args.steps, 5), "higher_is_better": True,
        "scaling": "weak" if world == 1 else "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": bench_config(world),
        "kernel": f"{path} x{launches_per_step}/step",
        "tokens_per_s": round(S / (total_ms / args.steps * 1e-3), 1),
        "pct_of_8000_gbps_spec": round(gbps / world / 8000 * 100, 1),
        "e2e": {"value": round(e2e_gbps, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(e2e_ms / args.steps, 5),
                "api": "B200PagedAttentionImpl.forward (reshape_and_cache + forward_decode) from pinned host buffers",
                "mode": e2e_mode + (" (the step, with its H2D / D2H copies, captured once and replayed per step, as the "
                                    "reference's CUDAGraphRunner runs decode)" if e2e_mode == "cuda_graph" else ""),
                "eager_ms_per_step": round(e2e_eager_ms / args.steps, 5),
                "kernel": e2e_path},
        "gpu_launches": args.steps * launches_per_step,
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": UNIT,
                     "frac": round(achieved / peak, 4), "traffic": ncu_traffic(), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": bytes_per_launch,
                     "read_stream_peak": read_peak,
                     "frac_of_read_stream_peak": round(achieved / read_peak, 4) if read_peak else None,
                     "read_stream_peak_source": "measured in this run: b200pa_read_stream_probe, 4 GiB, 128-bit "
                                                "streaming loads (a KV decode is a pure read stream; the contract "
                                                "peak above is a read+write copy)",
                     "kernel_ms_avg": round(avg_kernel_ms, 5), "kernel_ms_median": round(per_step[len(per_step) // 2], 5)},
        "clocks": clocks.summary(),
    }
    if extra:
        line["extra"] = extra
    return line


def weak_extra(ops, dist, dev, rank, world, steps):
    """The replica run of round 1: every rank streams its own config-2 cache (weak scaling, no exchange)."""
    cfg = CFG2
    c = make_decode_case(dev, [cfg["ctx"]] * cfg["num_seqs"], cfg["num_heads"], cfg["num_kv_heads"], cfg["head_size"],
                         cfg["block_size"], seed=100 + rank)
    fn = attention_call(ops, c, "v2")
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    ms = time_device(fn, steps, warmup=0)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    return {"workload": "config 2 replicated on every rank (weak scaling)", "ms_per_step": round(ms, 5),
            "value": round(world * c["nbytes"] / ms / 1e6, 1), "unit": UNIT, "scaling": "weak"}


def tp_step_extra(ops, dist, c, cfg, dev, rank, world, steps):
    """End-to-end TP decode step around attention (no GEMMs): fused rotary_embedding + reshape_and_cache on this
    rank's q/k/v -> paged attention on the shard -> NCCL all-reduce of the [S, hidden] bf16 activations that follows
    o_proj in the reference (vllm/model_executor/layers/linear.py:764-769 -> communication_op.py:73-100)."""
    S, L, Hq, Hkv, D, B = (cfg[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads", "head_size", "block_size"))
    hidden = CFG3["num_heads"] * D  # 8192 for the 70B layout
    g = torch.Generator(device=dev).manual_seed(7 + rank)
    q = torch.empty(S, Hq * D, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
    k = torch.empty(S, Hkv * D, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
    v = torch.empty(S, Hkv * D, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
    pos = torch.full((S,), L - 1, dtype=torch.int64, device=dev)
    cos_sin = torch.empty(L, D, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
    last_blk = c["block_tables"][:, (L - 1) // B].long()
    slots = (last_blk * B + (L - 1) % B).to(torch.int64)
    act = torch.empty(S, hidden, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
    attn = attention_call(ops, c, "v2")

    def front():
        ops.rotary_embedding_reshape_and_cache(pos, q, k, v, D, cos_sin, True, c["key_cache"], c["value_cache"], slots,
                                               "auto", 1.0)
        attn()

    def full():
        front()
        dist.all_reduce(act)

    res = {}
    for name, fn in (("attention_front_ms", front), ("step_ms", full), ("all_reduce_ms", lambda: dist.all_reduce(act))):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        ms = time_device(fn, steps, warmup=0)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = round(t.item(), 5)
    res.update({"workload": f"rope+reshape_and_cache (fused) -> paged_attention_v2 on the ({Hq},{Hkv})-head shard -> "
                            f"NCCL all-reduce [{S}, {hidden}] bf16 ({S * hidden * 2 >> 20} MiB)",
                "tokens_per_s": round(S / (res["step_ms"] * 1e-3), 1),
                "all_reduce_share": round(res["all_reduce_ms"] / res["step_ms"], 3)})
    return res


# ------------------------------------------------------------------------------------------------
# extra (N = 1): the other BASELINE configs, cache ops, prefill
# ------------------------------------------------------------------------------------------------
def configs_extra(ops, dev, peak):
    from vllm_legacy_b200.parallel import shard_heads
    out = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def rec(name, workload, nbytes, ms, **kw):
        gbps = nbytes / ms / 1e6
        d = {"name": name, "workload": workload, "algorithmic_bytes": int(nbytes), "ms": round(ms, 5),
             "gbps": round(gbps, 1), "frac": round(gbps / peak, 4), "pct_of_8000_gbps_spec": round(gbps / 80, 1)}
        d.update(kw)
        out.append(d)

    def attn(name, workload, lens, Hq, Hkv, D, B, kvd="auto", versions=("v2",), small=False, iters=20,
             dtype=torch.bfloat16, blocksparse=None, nbytes=None):
        c = make_decode_case(dev, lens, Hq, Hkv, D, B, kvd, dtype=dtype)
        for ver in versions:
            ms = time_device(attention_call(ops, c, ver, blocksparse), iters, flush=flush if small else None)
            rec(f"{name} {ver}", workload, nbytes if nbytes is not None else c["nbytes"], ms,
                kernel=ops.last_attention_path(), tokens_per_s=round(len(lens) / ms * 1e3, 1))
        return c

    # config 1 (the reference's CPU-runnable case): v1 + reshape_and_cache, MHA 32 heads, 16 x 1024
    c1 = attn("config1", "paged_attention_v1, 32 q/kv heads, head 128, block 16, 16 seqs x 1024 ctx, bf16 "
              "(L2 flushed between calls)", [1024] * 16, 32, 32, 128, 16, versions=("v1", "v2"), small=True)
    key = torch.randn(16, 32, 128, device=dev).to(torch.bfloat16)
    val = torch.randn(16, 32, 128, device=dev).to(torch.bfloat16)
    slots = (c1["block_tables"][:, -1].long() * 16 + 15).to(torch.int64)
    ms = time_device(lambda: ops.reshape_and_cache(key, val, c1["key_cache"], c1["value_cache"], slots, "auto", 1.0),
                     20, flush=flush)
    rec("config1 reshape_and_cache", "16 tokens x 32 heads x 128, bf16 (launch-latency bound)", 16 * 32 * 128 * 2 * 4, ms)
    del c1
    # config 3: the per-GPU shards of the 70B layout, each on this one GPU
    for tp in (2, 4, 8):
        sh = shard_heads(64, 8, tp, 0)
        attn(f"config3 tp{tp} shard", f"64q/8kv model sharded {tp}-way: ({sh.num_heads},{sh.num_kv_heads}) heads, "
             "256 x 8192 ctx, bf16, block 16", [8192] * 256, sh.num_heads, sh.num_kv_heads, 128, 16,
             versions=("v1", "v2"))
        torch.cuda.empty_cache()
    # config 4: fp8 e4m3 KV, 128 x 32k, + the quantising cache write at decode and prefill-chunk sizes
    c4 = attn("config4 fp8", "32q/8kv, head 128, block 16, 128 x 32768 ctx, fp8 e4m3 KV (in-kernel dequant)",
              [32768] * 128, 32, 8, 128, 16, kvd="fp8", versions=("v1", "v2"), iters=10)
    for T in (128, 32768):
        kv_new = torch.randn(T, 2, 8, 128, device=dev).to(torch.bfloat16)
        k_new, v_new = kv_new.unbind(dim=1)
        if T == 128:   # decode: one new token per sequence, the slot after its last block
            slots = (c4["block_tables"][:, -1].long() * 16 + 15).to(torch.int64)
        else:          # a prefill chunk: whole blocks, in order
            blocks = c4["block_tables"][0, :T // 16].long()
            slots = (blocks[:, None] * 16 + torch.arange(16, device=dev)[None, :]).flatten()
        ms = time_device(lambda: ops.reshape_and_cache(k_new, v_new, c4["key_cache"], c4["value_cache"], slots, "fp8",
                                                       0.5), 20, flush=flush)
        rec(f"config4 reshape_and_cache fp8 T={T}", f"{T} tokens x 8 kv heads x 128, bf16 -> e4m3 (kv_scale 0.5)",
            T * 8 * 128 * 2 * (2 + 1), ms)
    del c4
    torch.cuda.empty_cache()
    # the general (SIMT) kernel family: what the TMA + mma.sync family does not instantiate (fp32 queries; fp8 caches
    # with block size 8) — a coverage kernel, measured so that its distance from the roofline is on record — and the two
    # shapes that moved to the mma family in round 2: block size 8 over a 16-bit cache, and block-sparse masks (head size
    # 64 / 112 / 128, filtered block iterator)
    S_, L_ = 128, 4096
    attn("simt fp32", "fp32 queries + fp32 cache (x = 4), 32q/8kv, head 128, block 16, 128 x 4096 ctx", [L_] * S_, 32, 8,
         128, 16, dtype=torch.float32, iters=10)
    torch.cuda.empty_cache()
    attn("block8", "block size 8, bf16, 32q/8kv, head 128, 128 x 4096 ctx (mma family: one n-tile per block, P.V with "
         "m16n8k8)", [L_] * S_, 32, 8, 128, 8, iters=10)
    torch.cuda.empty_cache()
    # block-sparse (attention_kernels.cu:212-256): with head_sliding_step 0 a 64-token block k is read when
    # k > q_block - local_blocks or (k + 1) % vert_stride == 0; only those blocks are algorithmic bytes
    local_blocks, vert_stride, bs_block = 16, 8, 64
    nb_ = L_ // bs_block
    q_blk = (L_ - 1) // bs_block
    visible = sum(1 for k in range(nb_) if k > q_blk - local_blocks or (k + 1) % vert_stride == 0)
    attn("blocksparse", f"block-sparse (local {local_blocks} x {bs_block}-token blocks + every {vert_stride}th), "
         f"bf16, 32q/8kv, head 128, block 16, 128 x 4096 ctx: {visible} of {nb_} blocks visible", [L_] * S_, 32, 8, 128,
         16, iters=10, blocksparse=(local_blocks, vert_stride, bs_block, 0),
         nbytes=algorithmic_bytes(S_, visible * bs_block, 32, 8, 128))
    torch.cuda.empty_cache()
    # config 5: ragged 128..128k mix, block 16 / 32, + copy_blocks / swap_blocks for prefix / beam sharing
    lens = ragged_lens()
    for B in (16, 32):
        attn(f"config5 ragged b{B}", f"32q/8kv, head 128, block {B}, 64 seqs log-uniform 128..131072 ctx "
             f"({sum(lens)} tokens), bf16", lens, 32, 8, 128, B, versions=("v1", "v2"), iters=10)
        torch.cuda.empty_cache()
    layers, NB, H, D, B = 32, 1024, 8, 128, 16
    kcs = [torch.empty(NB, H, D // 8, B, 8, dtype=torch.bfloat16, device=dev).normal_() for _ in range(layers)]
    vcs = [torch.empty(NB, H, D, B, dtype=torch.bfloat16, device=dev).normal_() for _ in range(layers)]
    perm = torch.randperm(NB)
    mapping = torch.stack([perm[:256], perm[256:512]], dim=1).to(dev)
    block_bytes = kcs[0][0].numel() * 2
    ms = time_device(lambda: ops.copy_blocks(kcs, vcs, mapping), 20, flush=flush)
    rec("config5 copy_blocks", "256 (src,dst) pairs x 32 layers, 32 KiB K + 32 KiB V blocks (read + write)",
        layers * 256 * 2 * block_bytes * 2, ms)
    host = torch.zeros(512, H, D // 8, B, 8, dtype=torch.bfloat16).pin_memory()
    m_out = torch.stack([perm[:256], torch.arange(256)], dim=1)
    m_in = torch.stack([torch.arange(256), perm[:256]], dim=1)
    m_dd = torch.stack([perm[:256], perm[256:512]], dim=1)
    for label, a, b, m in (("D2H", kcs[0], host, m_out), ("H2D", host, kcs[0], m_in), ("D2D", kcs[0], kcs[1], m_dd)):
        ms = time_device(lambda: ops.swap_blocks(a, b, m), 20)
        rec(f"config5 swap_blocks {label}", "256 scattered 32 KiB blocks, pinned host memory"
            if label != "D2D" else "256 scattered 32 KiB blocks, device to device", 256 * block_bytes, ms,
            bound="pcie" if label != "D2D" else "hbm")
    return out


def prefill_extra(ops, dev, iters=10):
    """Extra, outside the headline metric: the tcgen05 prefix-prefill kernel on a Llama-3-8B
    chunked-prefill shape (8 sequences x 2048 new tokens over 4096 cached tokens, 32q/8kv, head 128, block 16, bf16),
    timed with CUDA events, reported against the MEASURED cuBLAS bf16 peak (tensor roofline)."""
    Hq, Hkv, D, B, nseq, ql, cl = 32, 8, 128, 16, 8, 2048, 4096
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)  # noqa: E731
    nblk = cl // B
    q, k, v = rnd(nseq * ql, Hq, D), rnd(nseq * ql, Hkv, D), rnd(nseq * ql, Hkv, D)
    k_cache, v_cache = rnd(nseq * nblk, Hkv, D // 8, B, 8), rnd(nseq * nblk, Hkv, D, B)
    block_table = torch.randperm(nseq * nblk, generator=g, device=dev).to(torch.int32).view(nseq, nblk)
    start = torch.arange(nseq, dtype=torch.int32, device=dev) * ql
    seq_len = torch.full((nseq,), ql + cl, dtype=torch.int32, device=dev)
    ctx_len = torch.full((nseq,), cl, dtype=torch.int32, device=dev)
    flops = 4.0 * D * Hq * nseq * (ql * cl + ql * (ql + 1) / 2)
    out = torch.empty_like(q)
    a = (q, k, v, out, k_cache, v_cache, block_table, start, seq_len, ctx_len, ql, None, 0)
    ms = time_device(lambda: ops.context_attention_fwd(*a), iters)
    tflops = flops / ms / 1e9
    peak = measured_peaks().get("bf16_tflops")
    # the same step over an fp8 e4m3 KV cache (K [nb, H, D/16, B, 16], V [nb, H, D, B], uint8; random finite codes of
    # small magnitude): the loader warps dequantise while staging (the reference's kernel has no working fp8 path)
    def fp8_bytes(*shape):
        mag = torch.randint(0, 0x38, shape, dtype=torch.uint8, device=dev, generator=g)
        return mag | (torch.randint(0, 2, shape, dtype=torch.uint8, device=dev, generator=g) << 7)
    a8 = a[:4] + (fp8_bytes(nseq * nblk, Hkv, D // 16, B, 16), fp8_bytes(nseq * nblk, Hkv, D, B)) + a[6:] + ("fp8", 0.5)
    ms8 = time_device(lambda: ops.context_attention_fwd(*a8), iters)
    return {"op": "context_attention_fwd (tcgen05 + TMEM prefix-prefill kernel)",
            "fp8_kv_prefix": {"workload": "same step over an fp8 e4m3 KV cache (kv_scale 0.5), dequantised while staged",
                              "ms": round(ms8, 4), "algorithmic_tflops": round(flops / ms8 / 1e9, 1)},
            "workload": "8 seqs x 2048 new tokens over 4096 cached, 32q/8kv, head 128, block 16, bf16",
            "ms": round(ms, 4), "algorithmic_tflops": round(tflops, 1), "launches": iters,
            "roofline": {"bound": "tensor", "achieved": round(tflops, 1), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(tflops / peak, 4) if peak else None,
                         "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops, cuBLAS burst)"}}


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU kernels (oracle/_ref) on the host cores
# ------------------------------------------------------------------------------------------------
def _load_ref_cpu():
    """The reference's csrc/cpu/attention.cpp + cache.cpp compiled unchanged into oracle/_ref (oracle/build_ref.py);
    needs AVX512 on the host.  None when absent."""
    try:
        with open("/proc/cpuinfo") as f:
            flags = f.read()
    except OSError:
        flags = ""
    if "avx512f" not in flags:
        return None
    name = "ref_cpu_bf16" if "avx512_bf16" in flags else "ref_cpu"
    path = os.path.join(ROOT, "oracle", "_ref", name + ".so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _cpu_reference_impl():
    """('reference', module) when the reference's csrc/cpu sources were compiled into oracle/_ref and the
    host has AVX512; else ('port', oracle module) — the plain-C restatement."""
    mod = _load_ref_cpu()
    if mod is not None:
        return "reference", mod
    from oracle import oracle
    oracle.build()
    return "port", oracle


def cpu_reference_run(cfg, total_budget_s, steps, warmup, version="v2"):
    """Times paged_attention of the CPU reference on a bounded sample of `cfg` (fewer sequences, same heads /
    context / block size / dtype, distinct physical blocks).  Returns (GB/s, description dict)."""
    kind, impl = _cpu_reference_impl()
    # all host threads this process may use (torchrun exports OMP_NUM_THREADS=1; undo that for the CPU arm)
    try:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    except Exception:  # noqa: BLE001
        pass
    L, Hq, Hkv, D, B = (cfg[k] for k in ("ctx", "num_heads", "num_kv_heads", "head_size", "block_size"))
    cores = torch.get_num_threads() if kind == "reference" else 1
    scale = float(D ** -0.5)
    nblk = L // B

    def build(S):
        g = torch.Generator().manual_seed(0)
        NB = S * nblk
        kc = torch.empty(NB, Hkv, D // 8, B, 8, dtype=torch.bfloat16).uniform_(-scale, scale, generator=g)
        vc = torch.empty(NB, Hkv, D, B, dtype=torch.bfloat16).uniform_(-scale, scale, generator=g)
        q = torch.empty(S, Hq, D, dtype=torch.bfloat16).uniform_(-scale, scale, generator=g)
        bt = torch.randperm(NB, generator=g).to(torch.int32).view(S, nblk).contiguous()
        sl = torch.full((S,), L, dtype=torch.int32)
        P = (L + 511) // 512
        out = torch.empty_like(q)
        es = torch.empty(S, Hq, P, dtype=torch.float32)
        ml = torch.empty_like(es)
        tmp = torch.empty(S, Hq, P, D, dtype=torch.bfloat16)

        def call():
            if version == "v1":
                impl.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0, 0, 0, 0, 64, 0)
            else:
                impl.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0, 0, 0,
                                        0, 64, 0)
        return call

    # calibrate on one sequence, then size the sample to the time budget
    call = build(1)
    call()
    t0 = time.perf_counter()
    call()
    per_seq = max(time.perf_counter() - t0, 1e-4)
    if kind == "reference":
        per_seq /= min(cores, Hq)  # one sequence only feeds `num_heads` parallel work items
    budget_per_step = total_budget_s / max(steps + warmup, 1)
    S = int(max(1, min(cfg["num_seqs"], budget_per_step / per_seq)))
    call = build(S)
    for _ in range(warmup):
        call()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    dt = (time.perf_counter() - t0) / steps
    nbytes = algorithmic_bytes(S, L, Hq, Hkv, D)
    desc = {"kind": kind, "cores": cores,
            "sample": f"{S} of {cfg['num_seqs']} sequences x {L} ctx ({Hq}q/{Hkv}kv heads, same block/dtype), "
                      f"paged_attention_{version} of "
                      + ("csrc/cpu/attention.cpp compiled unchanged (oracle/_ref)" if kind == "reference"
                         else "oracle/paged_attn_oracle.c (scalar port)")
                      + f", {steps} timed calls after {warmup} warm-up",
            "ms_per_step": round(dt * 1e3, 3), "tokens_per_s": round(S / dt, 1)}
    return nbytes / dt / 1e9, desc


def host_metadata_extra():
    """SURVEY §8 f4, host only: the reference's per-token / per-sequence Python loops that build slot_mapping and
    block_tables (restated in oracle/metadata_ref.py from vllm/worker/model_runner.py:464-502, 543-556) timed beside
    the vectorised builders of vllm_legacy_b200.worker.attn_metadata, on a config-2 decode step and a prefill chunk."""
    from oracle import metadata_ref as ref
    from vllm_legacy_b200.worker import attn_metadata as am
    rng = random.Random(0)
    B = 16
    out = []

    def best(fn, reps=5):
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        return min(ts) * 1e3

    for name, lens, ctxs in (("config 2 decode step (256 seqs x 4096 ctx)", [4096] * 256, [4095] * 256),
                             ("prefill chunk (8 seqs x 2048 new tokens over 4096 cached)", [6144] * 8, [4096] * 8)):
        bts = [[rng.randrange(10 ** 6) for _ in range((n + B - 1) // B)] for n in lens]
        width = max(len(b) for b in bts)
        ms_ref = best(lambda: (ref.slot_mapping_loops(bts, ctxs, lens, B), ref.make_tensor_with_pad_loops(bts, width, 0)))
        ms_new = best(lambda: (am.build_slot_mapping(bts, ctxs, lens, B), am.pad_block_tables(bts, width, 0)))
        out.append({"step": name, "reference_loops_ms": round(ms_ref, 3), "vectorised_ms": round(ms_new, 3),
                    "speedup": round(ms_ref / ms_new, 1)})
    # steady-state decode: every step each sequence grows by one token (a new block every 16 steps); the reference
    # rebuilds both arrays from the Python lists each step, the cache copies only what is new
    lens = [4096 - 64] * 256
    bts = [[rng.randrange(10 ** 6) for _ in range((n + B - 1) // B)] for n in lens]
    cache = am.BlockTableCache(4096 // B)
    ids = list(range(256))
    cache.decode_step(ids, bts, lens, B)
    t_ref = t_new = 0.0
    for _ in range(48):
        lens = [n + 1 for n in lens]
        for bt, n in zip(bts, lens):
            if (n + B - 1) // B > len(bt):
                bt.append(rng.randrange(10 ** 6))
        t0 = time.perf_counter()
        ref.slot_mapping_loops(bts, [n - 1 for n in lens], lens, B)
        ref.make_tensor_with_pad_loops(bts, max(len(b) for b in bts), 0)
        t1 = time.perf_counter()
        cache.decode_step(ids, bts, lens, B)
        t2 = time.perf_counter()
        t_ref += t1 - t0
        t_new += t2 - t1
    out.append({"step": "48 consecutive config-2 decode steps, per step (BlockTableCache: incremental rows)",
                "reference_loops_ms": round(t_ref / 48 * 1e3, 3), "vectorised_ms": round(t_new / 48 * 1e3, 3),
                "speedup": round(t_ref / t_new, 1)})
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return None
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    cfg = CFG2 if world == 1 else CFG3  # the CPU is not sharded: it samples the whole model's heads
    gbps, desc = cpu_reference_run(cfg, total_budget_s=90.0, steps=min(steps, 50), warmup=min(warmup, 5))
    cb = dict(desc)
    cb.update({"value": round(gbps, 3), "unit": UNIT})
    return {
        "impl": "reference", "metric": METRIC, "value": round(gbps, 3), "unit": UNIT, "n_gpus": world,
        "steps": min(steps, 50), "warmup": min(warmup, 5), "ms_per_step": desc["ms_per_step"],
        "higher_is_better": True, "scaling": "weak" if world == 1 else "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "config": bench_config(world),
        "note": "CPU reference on a bounded sample of the workload, all host cores (rank 0 only)",
        "tokens_per_s": desc["tokens_per_s"],
        "cpu_baseline": cb,
        "e2e": {"value": round(gbps, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


_JSON_FD = None


def _stdout_json_only():
    """stdout carries the one JSON line and nothing else: file descriptor 1 is pointed at stderr for the rest of the
    run (NCCL prints its version banner on stdout when NCCL_DEBUG is set on the box) and the line is written to the
    saved descriptor."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    _stdout_json_only()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU baseline leg (profiling runs)")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra (non-headline) measurements")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        line = run_reference(args, rank, world)
        if line is not None:
            _emit(line)
        return 0

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU reference arm)")
    import torch.distributed as dist
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        line = run_ours(args, rank, world, local_rank)
        if rank == 0:
            if world == 1 and not args.no_extra:
                from vllm_legacy_b200 import _custom_ops as ops
                dev = torch.device("cuda", local_rank)
                extra = line.setdefault("extra", {})
                for name, fn in (("configs", lambda: configs_extra(ops, dev, line["roofline"]["peak"])),
                                 ("prefill", lambda: prefill_extra(ops, dev))):
                    try:
                        extra[name] = fn()
                    except Exception as e:  # noqa: BLE001
                        extra[name] = {"failed": f"{type(e).__name__}: {e}"[:300]}
                    torch.cuda.empty_cache()
            if world == 1 and not args.no_cpu_baseline:
                try:
                    gbps, desc = cpu_reference_run(CFG2, total_budget_s=15.0, steps=3, warmup=1)
                    cb = dict(desc)
                    cb.update({"value": round(gbps, 3), "unit": UNIT})
                    line["cpu_baseline"] = cb
                    line.setdefault("extra", {})["host_metadata"] = host_metadata_extra()
                    if "extra" in line and isinstance(line["extra"].get("configs"), list):
                        # BASELINE configs[0] as worded: paged_attention_v1 on the vLLM CPU extension, 16 x 1024 MHA
                        c1 = dict(num_seqs=16, ctx=1024, num_heads=32, num_kv_heads=32, head_size=128, block_size=16)
                        g1, d1 = cpu_reference_run(c1, total_budget_s=5.0, steps=3, warmup=1, version="v1")
                        line["extra"]["configs"].append({"name": "config1 v1 CPU reference", "workload": d1["sample"],
                                                         "gbps": round(g1, 3), "ms": d1["ms_per_step"],
                                                         "cores": d1["cores"], "kind": d1["kind"]})
                except Exception as e:  # noqa: BLE001
                    line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port",
                                            "sample": f"failed: {e}"}
            _emit(line)
    finally:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
