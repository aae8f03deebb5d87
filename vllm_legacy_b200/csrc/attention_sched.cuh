// Work decomposition of the persistent paged-attention decode kernel (attention_mma_kernel.cuh):
// a "stream-K" split of the flattened KV-block list over the resident CTAs.
//
// The work of one launch is the list of KV blocks of every (sequence, unit) pair, where a unit is
// one (KV head, query-head tile) and a sequence of length L has n_s = ceil(L / block_size) blocks:
//
//     flat index f  =  N_s * HK  +  unit * n_s  +  j        N_s = sum_{s' < s} n_s'   (blocks before s)
//                                                            HK  = units per sequence, j in [0, n_s)
//
// i.e. sequence-major, then unit, then block.  W = HK * sum_s n_s is the total.  CTA c of G owns the
// contiguous range [f0(c), f0(c+1)), f0(c) = floor(W c / G): every CTA streams the same number of
// blocks (+-1) whatever the length distribution, so there is no wave tail and no per-sequence
// imbalance (SURVEY §7 "flattened work queue"; the reference's grid is one CTA per (head, seq,
// 512-token partition), attention_kernels.cu:844-946).
//
// An (s, unit) item that straddles CTA boundaries is computed in parts; each part leaves an
// un-normalised (max, sum, O) partial in a workspace slot and the last part to finish merges them
// (attention_mma_kernel.cuh).  The helpers below are the index arithmetic both sides of that
// hand-off agree on; they are __host__ __device__ so the CPU suite can check them exhaustively
// (tests/test_host_logic.py through b200pa_sched_* in attention.cu).
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define B200PA_HD __host__ __device__ __forceinline__
#else
#define B200PA_HD inline
#endif

namespace b200pa {

// First flat index of CTA c (c in [0, G]); f0(G) == W.
B200PA_HD int64_t sk_range_begin(int64_t W, int G, int c) { return (W * (int64_t)c) / G; }

// The CTA whose range contains flat index f (0 <= f < W): the largest c with f0(c) <= f.
//   floor(W c / G) <= f  <=>  W c < (f + 1) G  <=>  c <= ceil((f + 1) G / W) - 1 = floor(((f + 1) G - 1) / W)
B200PA_HD int sk_cta_of(int64_t W, int G, int64_t f) { return (int)((((f + 1) * (int64_t)G) - 1) / W); }

// Number of CTAs actually used for W blocks on a grid of `grid` CTAs: never more CTAs than blocks,
// so that every used CTA has a non-empty range (the part count of a split item is then simply
// sk_cta_of(last) - sk_cta_of(first) + 1).
B200PA_HD int sk_effective_ctas(int64_t W, int grid) { return W < (int64_t)grid ? (int)W : grid; }

// Workspace slot of a part: a CTA writes at most two partials per launch — one for the item it
// enters in the middle of (the item began in an earlier CTA: slot 0) and one for the item it
// leaves in the middle of having started it (slot 1).
B200PA_HD int sk_slot_index(int cta, bool item_began_earlier) { return cta * 2 + (item_began_earlier ? 0 : 1); }

}  // namespace b200pa
