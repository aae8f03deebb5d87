"""Work split of the persistent decode kernel (csrc/attention_sched.cuh, csrc/attention_sk_kernel.cuh).

CPU only.  The index arithmetic is called through the C-ABI (b200pa_sched_*: the very functions the
kernel compiles); the kernel's control flow around it — chunk list (static region + dynamic pool), the
producer's walk over the items of a chunk, block round-robin over the four consumer warps, part queue,
slot / arrival-counter protocol of items cut by chunk boundaries — is restated here step by step, with
the pool chunks handed to CTAs in a random order, and checked for the properties the kernel relies on:

  * every (sequence, unit, block) is issued exactly once, to exactly one (CTA, warp), in ring order;
  * the part queue never holds more parts than its 64 entries between producer and slowest consumer;
  * an item is either complete inside one chunk (direct write) or split into parts whose
    [first chunk, last chunk] = [sk_chunk_of(first block), sk_chunk_of(last block)], one part per chunk in
    between, the head part in slot 1 of the first chunk and every other part in slot 0 of its chunk;
  * no slot is written twice in a launch, every arrival counter returns to zero, and the merge order
    (hence the result) does not depend on which CTA took which chunk.
"""
import ctypes
import os
import random

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "vllm_legacy_b200", "libb200pa.so")
WARPS = 4
MAX_CHUNKS = 16384


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} is not built: run `python -m vllm_legacy_b200.build`")
    L = ctypes.CDLL(LIB)
    i32, i64 = ctypes.c_int32, ctypes.c_int64
    L.b200pa_sched_range_begin.restype = i64
    L.b200pa_sched_range_begin.argtypes = [i64, i32, i32]
    L.b200pa_sched_cta_of.restype = i32
    L.b200pa_sched_cta_of.argtypes = [i64, i32, i64]
    L.b200pa_sched_effective_ctas.restype = i32
    L.b200pa_sched_effective_ctas.argtypes = [i64, i32]
    L.b200pa_sched_geometry.restype = None
    L.b200pa_sched_geometry.argtypes = [i64, i32, i32, i32, i32, i32, ctypes.POINTER(i64)]
    L.b200pa_sched_chunk_begin.restype = i64
    L.b200pa_sched_chunk_begin.argtypes = [i64, i32, i32, i32, i32, i32, i32]
    L.b200pa_sched_chunk_of.restype = i32
    L.b200pa_sched_chunk_of.argtypes = [i64, i32, i32, i32, i32, i32, i64]
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


def _geometry(lib, W, grid, knobs):
    out = (ctypes.c_int64 * 5)()
    lib.b200pa_sched_geometry(W, grid, *knobs, MAX_CHUNKS, out)
    return dict(W1=out[0], Gs=out[1], cs=out[2], PC=out[3], NC=out[4])


@pytest.mark.parametrize("W,grid,knobs", [
    (524288, 296, (25, 8, 8)), (131072, 296, (25, 8, 8)), (16384, 296, (25, 8, 8)), (8, 296, (25, 8, 8)),
    (1, 296, (25, 8, 8)), (5000, 148, (0, 8, 8)), (5000, 148, (100, 8, 8)), (10 ** 7, 296, (25, 64, 8)),
    (777, 3, (50, 2, 1)),
])
def test_chunk_list_tiles_the_work(lib, W, grid, knobs):
    g = _geometry(lib, W, grid, knobs)
    assert g["Gs"] <= grid and g["NC"] == g["Gs"] + g["PC"] <= MAX_CHUNKS
    begins = [lib.b200pa_sched_chunk_begin(W, grid, *knobs, MAX_CHUNKS, c) for c in range(g["NC"] + 1)]
    assert begins[0] == 0 and begins[-1] == W
    assert all(b > a for a, b in zip(begins, begins[1:])), "every chunk is non-empty"
    assert begins[g["Gs"]] == g["W1"]
    if g["PC"]:
        assert all(b - a == g["cs"] for a, b in zip(begins[g["Gs"]:-2], begins[g["Gs"] + 1:-1]))
    rng = random.Random(1)
    for f in set(begins[:-1]) | {b - 1 for b in begins[1:]} | {rng.randrange(W) for _ in range(200)}:
        c = lib.b200pa_sched_chunk_of(W, grid, *knobs, MAX_CHUNKS, f)
        assert begins[c] <= f < begins[c + 1]


def _simulate(lib, seq_lens, HK, grid, block, nst, knobs=(25, 8, 8), seed=0):
    """Restatement of paged_attention_sk_kernel's control flow (no arithmetic on KV data).  Returns the merge plan
    {item: [slots in merge order]} so that callers can compare runs with different chunk-to-CTA assignments."""
    blocks_of = lambda l: (l + block - 1) // block if l > 0 else 0  # noqa: E731
    n = [blocks_of(l) for l in seq_lens]
    S = len(seq_lens)
    W = sum(n) * HK
    g = _geometry(lib, W, grid, knobs)
    NC, Gs = g["NC"], g["Gs"]
    begin = [lib.b200pa_sched_chunk_begin(W, grid, *knobs, MAX_CHUNKS, c) for c in range(NC + 1)]
    chunk_of = lambda f: lib.b200pa_sched_chunk_of(W, grid, *knobs, MAX_CHUNKS, f)  # noqa: E731
    prefix = [0]
    for v in n:
        prefix.append(prefix[-1] + v)

    # which CTA streams which chunks: chunk c < Gs -> CTA c; pool chunks in the order of an atomic counter, taken by
    # whichever CTA "finishes first" (random here)
    rng = random.Random(seed)
    streams = {c: ([c] if c < Gs else []) for c in range(grid)}
    for c in range(Gs, NC):
        streams[rng.randrange(grid)].append(c)

    consumed = {}            # (seq, hk, j) -> (cta, warp)
    slots = {}               # slot index -> item
    counters = {}
    direct = set()
    merged = {}
    arrivals = {}
    events = []              # (cta, order) per part with a slot, to replay arrivals in a random interleaving

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

    for cta, chunks in streams.items():
        gpos = 0
        per_warp = [[] for _ in range(WARPS)]   # producer's issue order per consumer ring
        parts = []                               # part queue
        for chunk in chunks:
            f0c, lenc = begin[chunk], begin[chunk + 1] - begin[chunk]
            x = f0c // HK
            s0 = next(s for s in range(S) if prefix[s] <= x < prefix[s + 1])   # locate()
            r0 = f0c - prefix[s0] * HK
            cur = dict(seq=s0, hk=r0 // n[s0], n=n[s0], base=-(r0 % n[s0]))
            rel = 0
            while True:
                jb = rel - cur["base"]
                je = min(cur["n"], lenc - cur["base"])
                cnt = je - jb
                assert 0 <= jb < je <= cur["n"]
                slot, ch, span = -1, 0, 0
                if not (jb == 0 and je == cur["n"]):
                    slot = chunk * 2 + (0 if jb > 0 else 1)
                    a = f0c + cur["base"]
                    ch, cl = chunk_of(a), chunk_of(a + cur["n"] - 1)
                    span = cl - ch
                    assert ch <= chunk <= cl and span >= 1 and ch < 65536 and span < 32768
                rot = (jb + cur["seq"] + cur["hk"]) % WARPS   # depends on the part only: bit-reproducible results
                parts.append(dict(seq=cur["seq"], hk=cur["hk"], jb=jb, cnt=cnt, rot=rot, slot=slot, ch=ch, span=span,
                                  first_block_index=gpos))
                for i in range(cnt):
                    per_warp[(i + rot) % WARPS].append((cur["seq"], cur["hk"], jb + i))
                gpos += cnt
                rel += cnt
                if rel >= lenc:
                    break
                next_item(cur)
        # consumers: walk the part queue, take the blocks (gpos + i) % 4 == warp in ring order
        for warp in range(WARPS):
            got = []
            for part in parts:
                i = (warp - part["rot"]) % WARPS
                while i < part["cnt"]:
                    got.append((part["seq"], part["hk"], part["jb"] + i))
                    i += WARPS
            assert got == per_warp[warp], "a consumer warp reads its ring in a different order than it was filled"
            for key in got:
                assert key not in consumed, f"block {key} consumed twice"
                consumed[key] = (cta, warp)
        # part-queue depth: when the producer publishes part k + 64, every consumer warp must have read part k.  The
        # producer is at most nst blocks ahead of the warp that owns them, and the merge barriers keep the four
        # warps within two parts of each other; so it suffices that in any 64 consecutive parts SOME warp owns more
        # than nst + 2 blocks (then the producer blocks on that warp's ring before it can lap the queue).
        for k in range(len(parts) - 64):
            own = [0] * WARPS
            for pt in parts[k:k + 64]:
                for i in range(pt["cnt"]):
                    own[(i + pt["rot"]) % WARPS] += 1
            assert max(own) > nst + 2, "the 64-entry part queue could be overrun"
        for k, part in enumerate(parts):
            item = (part["seq"], part["hk"])
            if part["slot"] < 0:
                assert item not in direct
                direct.add(item)
            else:
                assert part["slot"] not in slots, "a workspace slot is written twice in one launch"
                slots[part["slot"]] = item
                events.append((cta, k, item, part["slot"], part["ch"], part["span"]))

    # arrivals at the counters in a random global order (per-CTA order preserved)
    by_cta = {}
    for ev in events:
        by_cta.setdefault(ev[0], []).append(ev)
    queues = [sorted(v, key=lambda e: e[1]) for v in by_cta.values()]
    while queues:
        q = rng.choice(queues)
        _, _, item, slot, ch, span = q.pop(0)
        if not q:
            queues.remove(q)
        arrivals.setdefault(item, []).append(slot // 2)
        old = counters.get(ch, 0)
        counters[ch] = old + 1
        if old == span:   # last arriver merges chunk ch (slot 1) then ch+1 .. ch+span (slot 0)
            assert item not in merged
            merged[item] = [2 * c2 + (1 if c2 == ch else 0) for c2 in range(ch, ch + span + 1)]
            counters[ch] = 0

    want = {(s, hk, j) for s in range(S) for hk in range(HK) for j in range(n[s])}
    assert set(consumed) == want
    assert all(c == 0 for c in counters.values()), "arrival counters must return to zero"
    items = {(s, hk) for s in range(S) for hk in range(HK) if n[s] > 0}
    assert direct | set(merged) == items and not (direct & set(merged))
    for item, order in merged.items():
        for sl in order:
            assert slots[sl] == item, "the last arriver would merge a slot of another item"
        chunks = sorted(arrivals[item])
        assert chunks == list(range(chunks[0], chunks[-1] + 1)) and len(chunks) == len(order)
    return merged, NC


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
    knobs = rng.choice([(25, 8, 8), (0, 8, 8), (100, 4, 1), (50, 16, 2)])
    block, nst = rng.choice([16, 32]), rng.choice([1, 3, 6, 8])
    plan_a, _ = _simulate(lib, seq_lens, HK, grid, block, nst, knobs, seed=1)
    plan_b, _ = _simulate(lib, seq_lens, HK, grid, block, nst, knobs, seed=2)
    assert plan_a == plan_b, "the merge plan must not depend on which CTA took which pool chunk"


def test_split_baseline_shapes(lib):
    # config 2 / config 3 at TP8 / one long sequence / ragged mix (scaled down 8x in length to stay fast)
    _simulate(lib, [512] * 256, 8, 296, 16, 3)
    _simulate(lib, [1024] * 256, 1, 296, 16, 3)
    _simulate(lib, [4096], 8, 296, 16, 3)
    rng = random.Random(0)
    lens = [int(round(2 ** rng.uniform(4, 14))) for _ in range(64)]
    _simulate(lib, lens, 8, 296, 32, 3)
    # fewer blocks than CTAs
    _, nc = _simulate(lib, [16, 1, 0, 33], 2, 296, 16, 3)
    assert nc <= (1 + 1 + 0 + 3) * 2
