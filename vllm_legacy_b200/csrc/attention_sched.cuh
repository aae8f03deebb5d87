// Work decomposition of the persistent paged-attention decode kernel (attention_sk_kernel.cuh).
//
// The work of one launch is the list of KV blocks of every (sequence, unit) pair, where a unit is
// one (KV head, query-head tile) and a sequence of length L has n_s = ceil(L / block_size) blocks:
//
//     flat index f  =  N_s * HK  +  unit * n_s  +  j        N_s = sum_{s' < s} n_s'   (blocks before s)
//                                                            HK  = units per sequence, j in [0, n_s)
//
// i.e. sequence-major, then unit, then block.  W = HK * sum_s n_s is the total (SURVEY §7 "flattened
// work queue"; the reference's grid is one CTA per (head, seq, 512-token partition),
// attention_kernels.cu:844-946).
//
// [0, W) is cut into a FIXED list of chunks — a function of (W, grid size) only, so results do not
// depend on timing:
//
//   * a static region [0, W1): Gs equal contiguous chunks, chunk c is streamed by CTA c.  Big
//     chunks mean few cut items (<= 2 partial results per chunk);
//   * a pool region [W1, W) of small chunks of `cs` blocks that the CTAs take dynamically (one
//     atomic counter) once their static chunk is done.  On B200 the SMs do not stream at the same
//     rate (measured 5.5 - 7.7 blocks/us per SM on equal ranges, profiles/r02_sk_trace_static_split.log),
//     so an equal static split waits for the slowest SM; with the pool every SM runs at its own
//     rate until the work is gone and the tail is one small chunk.
//
// An (s, unit) item that straddles chunk boundaries is computed in parts; each part leaves an
// un-normalised (max, sum, O) partial in a workspace slot and the last part to finish merges them
// in chunk order.  The helpers below are the index arithmetic both sides of that hand-off agree
// on; they are __host__ __device__ so the CPU suite can check them exhaustively
// (tests/test_sched.py through b200pa_sched_* in attention.cu).
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define B200PA_HD __host__ __device__ __forceinline__
#else
#define B200PA_HD inline
#endif

namespace b200pa {

// First flat index of part c when [0, W) is cut into G equal parts (c in [0, G]); f0(G) == W.
B200PA_HD int64_t sk_range_begin(int64_t W, int G, int c) { return (W * (int64_t)c) / G; }

// The equal part that contains flat index f (0 <= f < W): the largest c with f0(c) <= f.
//   floor(W c / G) <= f  <=>  W c < (f + 1) G  <=>  c <= ceil((f + 1) G / W) - 1 = floor(((f + 1) G - 1) / W)
B200PA_HD int sk_cta_of(int64_t W, int G, int64_t f) { return (int)((((f + 1) * (int64_t)G) - 1) / W); }

// Never more parts than blocks, so that every part is non-empty.
B200PA_HD int sk_effective_ctas(int64_t W, int grid) { return W < (int64_t)grid ? (int)W : grid; }

// Workspace slot of a part: a chunk leaves at most two partials — one for the item it enters in
// the middle of (the item began in an earlier chunk: slot 0) and one for the item it leaves in
// the middle of having started it (slot 1).
B200PA_HD int sk_slot_index(int chunk, bool item_began_earlier) { return chunk * 2 + (item_began_earlier ? 0 : 1); }

struct SkGeom {
  int64_t W;   // total blocks
  int64_t W1;  // the static region is [0, W1)
  int Gs;      // static chunks (chunk c < Gs belongs to CTA c)
  int cs;      // blocks per pool chunk
  int PC;      // pool chunks
};

// pool_pct: share of the work handed out dynamically; chunks_per_cta: pool chunks aimed at per CTA
// (more = smaller tail, more partials); cs_min: smallest pool chunk; max_chunks: workspace capacity.
B200PA_HD SkGeom sk_geometry(int64_t W, int grid, int pool_pct, int chunks_per_cta, int cs_min, int max_chunks) {
  SkGeom g;
  g.W = W;
  int64_t pool = W * pool_pct / 100;
  // Small launches stay static: a pool chunk costs a part hand-off (queue entry, merge, workspace slot,
  // arrival counter), which only pays when the static chunks are several pool chunks long.
  if (W < (int64_t)grid * 6 * cs_min) pool = 0;
  g.W1 = W - pool;
  g.Gs = g.W1 > 0 ? sk_effective_ctas(g.W1, grid) : 0;
  int64_t cap = (int64_t)chunks_per_cta * grid;
  if (cap > max_chunks - g.Gs) cap = max_chunks - g.Gs;
  if (cap < 1) cap = 1;
  int64_t cs = (pool + cap - 1) / cap;
  if (cs < cs_min) cs = cs_min;
  g.cs = (int)cs;
  g.PC = (int)((pool + cs - 1) / cs);
  return g;
}

B200PA_HD int sk_num_chunks(const SkGeom& g) { return g.Gs + g.PC; }

B200PA_HD int64_t sk_chunk_begin(const SkGeom& g, int c) {
  if (c < g.Gs) return sk_range_begin(g.W1, g.Gs, c);
  const int64_t b = g.W1 + (int64_t)(c - g.Gs) * g.cs;
  return b < g.W ? b : g.W;
}

// The chunk that contains flat index f (0 <= f < W).
B200PA_HD int sk_chunk_of(const SkGeom& g, int64_t f) {
  return f < g.W1 ? sk_cta_of(g.W1, g.Gs, f) : g.Gs + (int)((f - g.W1) / g.cs);
}

}  // namespace b200pa
