#!/usr/bin/env python3
"""Benchmark of the B200 paged-attention decode hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--extra]

One "step" = one decode-attention layer-call (paged_attention_v2) over the whole synthetic batch.
N = 1 workload = BASELINE.json configs[1]: Llama-3-8B GQA decode, 32 q / 8 kv heads, head 128, block 16,
batch 256 x 4096 context, bf16 KV.  For N > 1 (launched with torchrun, one rank per GPU) every rank owns
a KV-head shard of that same per-GPU shape (tensor-parallel layout, vllm/config.py:320-333): no collective
on the data path, weak scaling, `value` = KV bytes streamed by all ranks / max-over-ranks device time.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for how each field is produced.
`--impl reference` times the reference's own CPU implementation (oracle/_ref, csrc/cpu/attention.cpp
compiled unchanged) on the host cores on a bounded sample of the same workload.
"""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decode_attn_kv_gbps"
UNIT = "GB/s"

# BASELINE.json configs[1]
CFG = dict(num_seqs=256, ctx=4096, num_heads=32, num_kv_heads=8, head_size=128, block_size=16)
WORKLOAD = ("Llama-3-8B GQA decode attention (32q/8kv, head 128), block 16, batch 256 x 4096 ctx, bf16, "
            "paged_attention_v2, distinct physical blocks (random permutation)")


def algorithmic_bytes(num_seqs, ctx, num_heads, num_kv_heads, head_size, cache_elem_bytes=2, io_elem_bytes=2):
    """sum(seq_len) * Hkv * D * 2 (K+V) * sizeof(cache_t) + q read + out write."""
    kv = num_seqs * ctx * num_kv_heads * head_size * 2 * cache_elem_bytes
    qo = 2 * num_seqs * num_heads * head_size * io_elem_bytes
    return kv + qo


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes (read+write) per launch of the dominant kernel from the committed ncu capture, or None."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        with open(p) as f:
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
# workload construction
# ------------------------------------------------------------------------------------------------
def make_device_workload(dev, cfg, seed):
    """Synthetic Llama-shaped paged KV cache resident in HBM + per-step host inputs (pinned)."""
    S, L, Hq, Hkv, D, B = (cfg[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads", "head_size", "block_size"))
    g = torch.Generator(device=dev).manual_seed(seed)
    scale = float(D ** -0.5)
    nblk = (L + B - 1) // B
    NB = S * nblk + S  # one spare block per sequence
    from vllm_legacy_b200.attention.backends.b200_paged import B200PagedAttentionBackend as Backend
    kv_cache = torch.empty(Backend.get_kv_cache_shape(NB, B, Hkv, D), dtype=torch.bfloat16, device=dev)
    kv_cache.uniform_(-scale, scale, generator=g)
    perm = torch.randperm(NB, generator=g, device=dev).to(torch.int32)
    block_tables = perm[:S * nblk].view(S, nblk).contiguous()
    seq_lens = torch.full((S,), L, dtype=torch.int32, device=dev)
    # host side of one decode step: q, new k/v, slot of the new token (the last position of each sequence)
    host = {
        "query": torch.empty(S, Hq * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "key": torch.empty(S, Hkv * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "value": torch.empty(S, Hkv * D, dtype=torch.bfloat16).uniform_(-scale, scale).pin_memory(),
        "block_tables": block_tables.cpu().pin_memory(),
        "seq_lens": seq_lens.cpu().pin_memory(),
    }
    last_blk = block_tables[:, (L - 1) // B].long().cpu()
    host["slot_mapping"] = (last_blk * B + (L - 1) % B).to(torch.int64).pin_memory()
    host["out"] = torch.empty(S, Hq * D, dtype=torch.bfloat16).pin_memory()
    return dict(kv_cache=kv_cache, block_tables=block_tables, seq_lens=seq_lens, scale=scale, host=host, NB=NB)


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from vllm_legacy_b200 import _custom_ops as ops
    from vllm_legacy_b200.attention.backends.b200_paged import B200PagedAttentionBackend as Backend
    from vllm_legacy_b200.attention.ops.paged_attn import PagedAttention

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    cfg = dict(CFG)
    S, L, Hq, Hkv, D, B = (cfg[k] for k in ("num_seqs", "ctx", "num_heads", "num_kv_heads", "head_size", "block_size"))
    w = make_device_workload(dev, cfg, seed=rank)
    key_cache, value_cache = PagedAttention.split_kv_cache(w["kv_cache"], Hkv, D)
    q = w["host"]["query"].to(dev).view(S, Hq, D)
    out = torch.empty_like(q)
    P = (L + 511) // 512
    tmp_out = torch.empty(S, Hq, P, D, dtype=q.dtype, device=dev)
    exp_sums = torch.empty(S, Hq, P, dtype=torch.float32, device=dev)
    max_logits = torch.empty_like(exp_sums)
    bytes_per_step = algorithmic_bytes(S, L, Hq, Hkv, D)

    def step():
        ops.paged_attention_v2(out, exp_sums, max_logits, tmp_out, q, key_cache, value_cache, Hkv, w["scale"],
                               w["block_tables"], w["seq_lens"], B, L, None, "auto", 1.0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- kernel-only arm: inputs resident in HBM -------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    launches_per_step = 1 + (1 if ops.last_attention_partitions() > 1 else 0)
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
    impl = Backend.get_impl_cls()(Hq, D, w["scale"], Hkv, None, None, "auto")
    h = w["host"]
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
        o = impl.forward(d_in["query"], d_in["key"], d_in["value"], w["kv_cache"], meta, 1.0)
        h["out"].copy_(o, non_blocking=True)

    for _ in range(3):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    e2e_path = ops.last_attention_path()

    # ---- max over ranks ------------------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = t.tolist()
    if rank != 0:
        return None
    gbps = world * bytes_per_step * args.steps / (total_ms * 1e-3) / 1e9
    e2e_gbps = world * bytes_per_step * args.steps / (e2e_ms * 1e-3) / 1e9
    peak, peak_src = measured_peak()
    achieved = bytes_per_step / (avg_kernel_ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": round(gbps, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(total_ms / args.steps, 5), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "per_gpu": cfg, "sharding": f"kv-head shards, tp={world}, no collective",
                   "l2": "inputs larger than L2 (4.3 GB KV per GPU streamed once per step)",
                   "kernel": f"{ops.last_attention_path()} x{launches_per_step}/step"},
        "tokens_per_s": round(S / (total_ms / args.steps * 1e-3), 1),
        "pct_of_8000_gbps_spec": round(gbps / world / 8000 * 100, 1),
        "e2e": {"value": round(e2e_gbps, 1), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(e2e_ms / args.steps, 5),
                "api": "B200PagedAttentionImpl.forward (reshape_and_cache + forward_decode) from pinned host buffers",
                "kernel": e2e_path},
        "gpu_launches": args.steps * launches_per_step,
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": UNIT,
                     "frac": round(achieved / peak, 4), "traffic": ncu_traffic(), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": bytes_per_step,
                     "kernel_ms_avg": round(avg_kernel_ms, 5), "kernel_ms_median": round(per_step[len(per_step) // 2], 5)},
        "clocks": clocks.summary(),
    }
    return line


def prefill_extra(dev, iters=10):
    """Extra, outside the headline metric: the tcgen05 prefix-prefill kernel on a Llama-3-8B
    chunked-prefill shape (8 sequences x 2048 new tokens over 4096 cached tokens, 32q/8kv, head 128, block 16, bf16),
    timed with CUDA events, reported against the MEASURED cuBLAS bf16 peak (tensor roofline)."""
    from tests import helpers
    from vllm_legacy_b200 import _custom_ops as ops
    Hq, qpk, D, B = 32, 4, 128, 16
    case = helpers.make_prefill_case(Hq, qpk, D, torch.bfloat16, dev, block_size=B, amplitude=1.0,
                                     query_lens=[2048] * 8, ctx_lens=[4096] * 8)
    flops = 4.0 * D * Hq * sum(q * c + q * (q + 1) / 2 for q, c in zip(case["query_lens"], case["ctx_lens"]))
    out = torch.empty_like(case["query"])
    a = (case["query"], case["k"], case["v"], out, case["k_cache"], case["v_cache"], case["block_table"],
         case["b_start_loc"], case["b_seq_len"], case["b_ctx_len"], case["max_input_len"], None, 0)
    for _ in range(3):
        ops.context_attention_fwd(*a)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.context_attention_fwd(*a)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    tflops = flops / ms / 1e9
    peak = None
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["bf16_tflops"])
    except Exception:  # noqa: BLE001
        pass
    return {"op": "context_attention_fwd (tcgen05 + TMEM prefix-prefill kernel)",
            "workload": "8 seqs x 2048 new tokens over 4096 cached, 32q/8kv, head 128, block 16, bf16",
            "ms": round(ms, 4), "algorithmic_tflops": round(tflops, 1), "launches": iters,
            "roofline": {"bound": "tensor", "achieved": round(tflops, 1), "peak": peak, "unit": "TFLOP/s",
                         "frac": round(tflops / peak, 4) if peak else None,
                         "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops, cuBLAS burst)"}}


# ------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU kernels (oracle/_ref) on the host cores
# ------------------------------------------------------------------------------------------------
def _cpu_reference_impl():
    """('reference', module) when the reference's csrc/cpu sources were compiled into oracle/_ref and the
    host has AVX512; else ('port', oracle module) — the plain-C restatement."""
    from tests import helpers
    mod = helpers.load_ref_cpu()
    if mod is not None:
        return "reference", mod
    from oracle import oracle
    oracle.build()
    return "port", oracle


def cpu_reference_run(total_budget_s, steps, warmup):
    """Times paged_attention_v2 of the CPU reference on a bounded sample of CFG (fewer sequences, same heads /
    context / block size / dtype, distinct physical blocks).  Returns (GB/s, description dict)."""
    kind, impl = _cpu_reference_impl()
    # all host threads this process may use (torchrun exports OMP_NUM_THREADS=1; undo that for the CPU arm)
    try:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    except Exception:  # noqa: BLE001
        pass
    cfg = dict(CFG)
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
            impl.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, B, L, None, "auto", 1.0, 0, 0, 0,
                                    64, 0)
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
            "sample": f"{S} of {cfg['num_seqs']} sequences x {L} ctx (same heads/block/dtype), paged_attention_v2 of "
                      + ("csrc/cpu/attention.cpp compiled unchanged (oracle/_ref)" if kind == "reference"
                         else "oracle/paged_attn_oracle.c (scalar port)")
                      + f", {steps} timed calls after {warmup} warm-up",
            "ms_per_step": round(dt * 1e3, 3), "tokens_per_s": round(S / dt, 1)}
    return nbytes / dt / 1e9, desc


def run_reference(args, rank, world):
    if rank != 0:
        return None
    steps, warmup = max(args.steps, 1), max(args.warmup, 0)
    gbps, desc = cpu_reference_run(total_budget_s=90.0, steps=min(steps, 50), warmup=min(warmup, 5))
    cb = dict(desc)
    cb.update({"value": round(gbps, 3), "unit": UNIT})
    return {
        "impl": "reference", "metric": METRIC, "value": round(gbps, 3), "unit": UNIT, "n_gpus": world,
        "steps": min(steps, 50), "warmup": min(warmup, 5), "ms_per_step": desc["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "per_gpu": dict(CFG), "note": "CPU reference on a bounded sample, host cores"},
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
    ap.add_argument("--no-extra", action="store_true", help="skip the extra (non-headline) prefill measurement")
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
                try:
                    line["extra"] = {"prefill": prefill_extra(torch.device("cuda", local_rank))}
                except Exception as e:  # noqa: BLE001
                    line["extra"] = {"prefill": {"failed": str(e)[:200]}}
            if world == 1 and not args.no_cpu_baseline:
                try:
                    gbps, desc = cpu_reference_run(total_budget_s=15.0, steps=3, warmup=1)
                    cb = dict(desc)
                    cb.update({"value": round(gbps, 3), "unit": UNIT})
                    line["cpu_baseline"] = cb
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
