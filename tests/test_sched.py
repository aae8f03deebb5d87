"""Work split of the persistent decode kernel (csrc/attention_sched.cuh, csrc/attention_sk_kernel.cuh).

CPU only.  The index arithmetic is called through the C-ABI (b200pa_sched_*: the very functions the
kernel compiles); the kernel's control flow around it — range per CTA, cursor walk over the items,
block interleave over the four warps, slot / arrival-counter protocol of split items — is restated
here step by step and checked for the properties the kernel relies on:

  * every (sequence, unit, block) is consumed exactly once, by exactly one (CTA, warp);
  * a warp's ring sees its blocks in stream order, producer NST blocks ahead of the consumer;
  * an item is either complete inside one CTA (direct write) or split into parts whose
    [first CTA, last CTA] = [sk_cta_of(first block), sk_cta_of(last block)], one part per CTA in
    between, the head part in slot 1 of the first CTA and every other part in slot 0 of its CTA;
  * no slot is written twice in a launch and every arrival counter returns to zero.
"""
import ctypes
import os
import random

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "vllm_legacy_b200", "libb200pa.so")
WARPS = 4


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} is not built: run `python -m vllm_legacy_b200.build`")
    L = ctypes.CDLL(LIB)
    L.b200pa_sched_range_begin.restype = ctypes.c_int64
    L.b200pa_sched_range_begin.argtypes = [ctypes.c_int64, ctypes.c_int32, ctypes.c_int32]
    L.b200pa_sched_cta_of.restype = ctypes.c_int32
    L.b200pa_sched_cta_of.argtypes = [ctypes.c_int64, ctypes.c_int32, ctypes.c_int64]
    L.b200pa_sched_effective_ctas.restype = ctypes.c_int32
    L.b200pa_sched_effective_ctas.argtypes = [ctypes.c_int64, ctypes.c_int32]
    return L


def test_cta_of_inverts_range_begin(lib):
    rng = random.Random(0)
    for _ in range(300):
        W = rng.choice([1, 2, 3, 7, 100, 296, 297, 1000, 131072, 524288, rng.randrange(1, 10 ** 6)])
        G = lib.b200pa_sched_effective_ctas(W, rng.choice([1, 2, 3, 148, 296]))
        assert 1 <= G <= W
        begins = [lib.b200pa_sched_range_begin(W, G, c) for c in range(G + 1)]
        assert begins[0] == 0 and begins[G] == W
        sizes = [b - a for a, b in zip(begins, begins[1:])]
        assert min(sizes) >= 1 and max(sizes) - min(sizes) <= 1  # equal shares, never empty
        probes = set(begins[:-1]) | {b - 1 for b in begins[1:]} | {rng.randrange(W) for _ in range(50)}
        for f in probes:
            c = lib.b200pa_sched_cta_of(W, G, f)
            assert begins[c] <= f < begins[c + 1], (W, G, f, c)


def test_large_totals_do_not_overflow(lib):
    W = (1 << 40) + 12345
    G = 296
    prev = 0
    for c in range(1, G + 1):
        b = lib.b200pa_sched_range_begin(W, G, c)
        assert b > prev
        assert lib.b200pa_sched_cta_of(W, G, b - 1) == c - 1
        if c < G:
            assert lib.b200pa_sched_cta_of(W, G, b) == c
        prev = b
    assert prev == W


def _simulate(lib, seq_lens, HK, grid, block, nst):
    """Restatement of paged_attention_sk_kernel's control flow (no arithmetic on KV data)."""
    blocks_of = lambda l: (l + block - 1) // block if l > 0 else 0  # noqa: E731
    n = [blocks_of(l) for l in seq_lens]
    S = len(seq_lens)
    total = sum(n)
    W = total * HK
    G = lib.b200pa_sched_effective_ctas(W, grid)
    consumed = {}            # (seq, hk, j) -> (cta, warp)
    slots = {}               # (cta, slot) -> (seq, hk)
    counters = [0] * max(G, 1)
    direct = set()           # items written directly
    merged = {}              # item -> list of (cta, slot) merged by the last arriver
    arrivals = {}            # item -> [(cta, ch, cl)]
    prefix = [0]
    for v in n:
        prefix.append(prefix[-1] + v)

    for cta in range(G):
        f0 = lib.b200pa_sched_range_begin(W, G, cta)
        ln = lib.b200pa_sched_range_begin(W, G, cta + 1) - f0
        assert ln >= 1
        x = f0 // HK
        s0 = next(s for s in range(S) if prefix[s] <= x < prefix[s + 1])  # the scan's result
        r0 = f0 - prefix[s0] * HK
        c0 = dict(seq=s0, hk=r0 // n[s0], n=n[s0], base=-(r0 % n[s0]))

        def next_item(c):
            c["base"] += c["n"]
            c["hk"] += 1
            if c["hk"] == HK:
                c["hk"] = 0
                while True:
                    c["seq"] += 1
                    c["n"] = n[c["seq"]]
                    if not (c["n"] == 0 and c["seq"] + 1 < S):
                        break

        # producer stream per warp (prologue lanes + refills) and consumer stream must agree
        for warp in range(WARPS):
            issued = []
            pc = dict(c0)
            rel = warp
            while rel < ln:
                while rel >= pc["base"] + pc["n"]:
                    next_item(pc)
                issued.append((pc["seq"], pc["hk"], rel - pc["base"]))
                rel += WARPS
            got = []
            cc = dict(c0)
            crel = warp
            while True:
                je = min(cc["n"], ln - cc["base"])
                while crel < cc["base"] + je:
                    got.append((cc["seq"], cc["hk"], crel - cc["base"]))
                    crel += WARPS
                if cc["base"] + cc["n"] >= ln:
                    break
                next_item(cc)
            assert got == issued, "consumer and producer walk different block streams"
            for k, key in enumerate(got):
                assert key not in consumed, f"block {key} consumed twice"
                consumed[key] = (cta, warp)
            # ring bookkeeping: block k uses stage k % nst; the producer is exactly nst ahead
            assert len(issued[:nst]) == min(nst, len(got))

        # item loop of the CTA (what warp 0 does with the merged part)
        cc = dict(c0)
        while True:
            jb = max(0, -cc["base"])
            je = min(cc["n"], ln - cc["base"])
            item = (cc["seq"], cc["hk"])
            assert 0 <= jb < je <= cc["n"]
            if jb == 0 and je == cc["n"]:
                assert item not in direct
                direct.add(item)
            else:
                slot = (cta, 0 if jb > 0 else 1)
                assert slot not in slots, "a workspace slot is written twice in one launch"
                slots[slot] = item
                a = f0 + cc["base"]
                ch = lib.b200pa_sched_cta_of(W, G, a)
                cl = lib.b200pa_sched_cta_of(W, G, a + cc["n"] - 1)
                assert ch <= cta <= cl and ch < cl
                arrivals.setdefault(item, []).append((cta, ch, cl))
                old = counters[ch]
                counters[ch] += 1
                if old == cl - ch:  # last arriver
                    merged[item] = [(c2, 1 if c2 == ch else 0) for c2 in range(ch, cl + 1)]
                    counters[ch] = 0
            if cc["base"] + cc["n"] >= ln:
                break
            next_item(cc)

    # ---- global properties -------------------------------------------------------------------------
    want = {(s, hk, j) for s in range(S) for hk in range(HK) for j in range(n[s])}
    assert set(consumed) == want
    assert all(c == 0 for c in counters), "arrival counters must return to zero"
    items = {(s, hk) for s in range(S) for hk in range(HK) if n[s] > 0}
    assert direct | set(merged) == items and not (direct & set(merged))
    for item, parts in merged.items():
        for part in parts:
            assert slots[part] == item, "the last arriver would merge a slot of another item"
        ctas = sorted(c for c, _, _ in arrivals[item])
        assert ctas == list(range(ctas[0], ctas[-1] + 1)) and len(ctas) == len(parts)
    return G


@pytest.mark.parametrize("seed", range(12))
def test_split_random_batches(lib, seed):
    rng = random.Random(seed)
    S = rng.choice([1, 2, 3, 7, 16, 64, 300])
    hi = rng.choice([1, 17, 100, 1000, 9000])
    seq_lens = [rng.choice([0, 1, rng.randrange(1, hi + 1)]) for _ in range(S)]
    if sum(seq_lens) == 0:
        seq_lens[-1] = 5
    HK = rng.choice([1, 2, 8, 12])
    grid = rng.choice([1, 2, 5, 37, 148, 296])
    _simulate(lib, seq_lens, HK, grid, rng.choice([16, 32]), rng.choice([1, 3, 6]))


def test_split_baseline_shapes(lib):
    # config 2 / config 3 at TP8 / one long sequence / ragged mix (scaled down 8x in length to stay fast)
    assert _simulate(lib, [512] * 256, 8, 296, 16, 3) == 296
    assert _simulate(lib, [1024] * 256, 1, 296, 16, 3) == 296
    assert _simulate(lib, [4096], 8, 296, 16, 3) == 296
    rng = random.Random(0)
    lens = [int(round(2 ** rng.uniform(4, 14))) for _ in range(64)]
    _simulate(lib, lens, 8, 296, 32, 3)
    # fewer blocks than CTAs: one block per CTA, nothing idle-spins
    assert _simulate(lib, [16, 1, 0, 33], 2, 296, 16, 3) == (1 + 1 + 0 + 3) * 2
