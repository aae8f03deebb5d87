// C-ABI entry points b200pa_paged_attention_v1 / _v2 (include/b200pa.h): argument checks with
// the reference's error classes, launch planning (context splits sized for 148 SMs), dispatch to
// the TMA + mma.sync family (split-K reduce fused into the kernel) or the general SIMT family
// (+ its reduce kernel), the per-device workspace, and the tuning knobs (environment read once,
// never on the launch path).
//
// Reference host side: csrc/attention/attention_kernels.cu:690-825 (v1), :844-1014 (v2).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "attention_common.cuh"

namespace b200pa {

static int env_int(const char* name, int dflt) {
  const char* s = std::getenv(name);
  return (s && *s) ? std::atoi(s) : dflt;
}

Tuning& tuning() {
  static Tuning t = [] {
    Tuning v;
    v.stages = env_int("B200PA_STAGES", v.stages);
    v.evict_normal = env_int("B200PA_EVICT_NORMAL", v.evict_normal);
    v.force_simt = env_int("B200PA_FORCE_SIMT", v.force_simt);
    v.partition_size = env_int("B200PA_PARTITION_SIZE", v.partition_size);
    v.max_partition = env_int("B200PA_MAX_PARTITION", v.max_partition);
    v.split_cost = std::max(1, env_int("B200PA_SPLIT_COST", v.split_cost));
    return v;
  }();
  return t;
}

// ---- per-device workspace -----------------------------------------------------------------------------
// Arrival counters of the fused split-K reduce (4 MB) + the scratch paged_attention_v1 splits through
// (64 MB; v2 uses the caller's exp_sums / max_logits / tmp_out).  Allocated by the first attention call on
// a device (or by b200pa_prepare()) and kept for the life of the process.  Calls on one device must not
// overlap in time on different streams (they never do in the reference's worker: one stream, one step).
static int alloc_workspace(int device, AttnWorkspace* ws) {
  int sms = 0;
  B200PA_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  ws->num_sms = sms;
  ws->max_counters = kWorkspaceCounters;
  ws->scratch_bytes = kWorkspaceScratchBytes;
  const size_t counter_bytes = (size_t)ws->max_counters * sizeof(unsigned);
  // The first call may sit inside a stream capture (global mode forbids cudaMalloc): allocate in
  // relaxed mode and zero the counters on a private stream, so nothing is recorded in the graph.
  cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
  B200PA_CUDA_CHECK(cudaThreadExchangeStreamCaptureMode(&mode));
  int prev = 0;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess) e = cudaSetDevice(device);
  char* base = nullptr;
  cudaStream_t s = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&base, ws->scratch_bytes + counter_bytes);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMemsetAsync(base + ws->scratch_bytes, 0, counter_bytes, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (s) cudaStreamDestroy(s);
  cudaSetDevice(prev);
  cudaThreadExchangeStreamCaptureMode(&mode);
  if (e != cudaSuccess) {
    g_last_cuda_error = static_cast<int>(e);
    return B200PA_ERR_CUDA;
  }
  ws->scratch = base;
  ws->counters = reinterpret_cast<unsigned*>(base + ws->scratch_bytes);
  return B200PA_OK;
}

int attn_workspace(int device, AttnWorkspace* out) {
  constexpr int kMaxDevices = 64;
  static AttnWorkspace table[kMaxDevices];
  static std::mutex mu;
  if (device < 0 || device >= kMaxDevices) return B200PA_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!table[device].scratch) {
    const int rc = alloc_workspace(device, &table[device]);
    if (rc != B200PA_OK) return rc;
  }
  *out = table[device];
  return B200PA_OK;
}

static bool head_size_supported(int d) {  // attention_kernels.cu:738-766
  return d == 64 || d == 80 || d == 96 || d == 112 || d == 128 || d == 192 || d == 256;
}
static bool block_size_supported(int b) {  // attention_kernels.cu:789-803
  return b == 8 || b == 16 || b == 32;
}

constexpr int kRefPartition = 512;  // granularity the callers size the v2 scratch with

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

// layout: 0 = the reference's paged layout (K [nb,H,D/x,B,x], V [nb,H,D,B]); 1 = flash layout (K, V [nb,B,H,D])
static int run_attention(const b200pa_attn_params_t* a, bool is_v2, cudaStream_t stream, int layout = 0,
                         int64_t kv_token_stride = 0) {
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->kv_dtype != B200PA_KV_AUTO && a->kv_dtype != B200PA_KV_FP8_E4M3 && a->kv_dtype != B200PA_KV_FP8_E5M2)
    return B200PA_ERR_KV_DTYPE;
  if (a->dtype != B200PA_F16 && a->dtype != B200PA_BF16 && a->dtype != B200PA_F32) return B200PA_ERR_INPUT_DTYPE;
  if (layout == 1) {  // flash_attn.py:24-25, :37-44 (block size a multiple of 16, 16-bit only, no fp8)
    if (a->block_size != 16 && a->block_size != 32) return B200PA_ERR_BLOCK_SIZE;
    if (!flash_head_size_supported(a->head_size)) return B200PA_ERR_HEAD_SIZE;
    if (a->kv_dtype != B200PA_KV_AUTO) return B200PA_ERR_KV_DTYPE;
    if (a->dtype == B200PA_F32) return B200PA_ERR_INPUT_DTYPE;
    if (a->blocksparse_vert_stride > 1 || reinterpret_cast<uintptr_t>(a->query) % 4 != 0 || a->q_stride % 2 != 0 ||
        (kv_token_stride * 2) % 16 != 0)
      return B200PA_ERR_INVALID_ARG;
  } else {
    if (!block_size_supported(a->block_size)) return B200PA_ERR_BLOCK_SIZE;
    if (!head_size_supported(a->head_size)) return B200PA_ERR_HEAD_SIZE;
  }
  if (a->num_seqs < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0 ||
      a->max_seq_len < 0 || a->max_num_blocks_per_seq < 0)
    return B200PA_ERR_INVALID_ARG;
  if (a->num_seqs == 0) return B200PA_OK;
  if (!a->out || !a->query || !a->key_cache || !a->value_cache || !a->block_tables || !a->seq_lens)
    return B200PA_ERR_INVALID_ARG;
  if (is_v2 && (!a->exp_sums || !a->max_logits || !a->tmp_out || a->max_num_partitions <= 0))
    return B200PA_ERR_INVALID_ARG;
  const int elem = a->dtype == B200PA_F32 ? 4 : 2;
  const int cache_bytes = a->kv_dtype == B200PA_KV_AUTO ? elem : 1;
  // 16-byte granularity of the paged layout (x = 16 / sizeof(cache_t)): the reference's vector
  // loads assume it too (attention_kernels.cu:200, 272-290).
  if (reinterpret_cast<uintptr_t>(a->key_cache) % 16 || reinterpret_cast<uintptr_t>(a->value_cache) % 16 ||
      (a->kv_block_stride * cache_bytes) % 16 || (a->kv_head_stride * cache_bytes) % 16)
    return B200PA_ERR_INVALID_ARG;

  AttnKernelParams kp{};
  kp.out = a->out;
  kp.q = a->query;
  kp.k_cache = a->key_cache;
  kp.v_cache = a->value_cache;
  kp.block_tables = a->block_tables;
  kp.seq_lens = a->seq_lens;
  kp.alibi_slopes = a->alibi_slopes;
  kp.q_stride = a->q_stride;
  kp.kv_block_stride = a->kv_block_stride;
  kp.kv_head_stride = a->kv_head_stride;
  kp.kv_token_stride = kv_token_stride;
  kp.num_heads = a->num_heads;
  kp.num_kv_heads = a->num_kv_heads;
  kp.head_size = a->head_size;
  kp.block_size = a->block_size;
  kp.max_blocks_per_seq = a->max_num_blocks_per_seq;
  kp.scale = a->scale;
  kp.kv_scale = a->kv_scale;
  kp.tp_rank = a->tp_rank;
  kp.bs_local_blocks = a->blocksparse_local_blocks;
  kp.bs_vert_stride = a->blocksparse_vert_stride;
  kp.bs_block_size = a->blocksparse_block_size;
  kp.bs_head_sliding_step = a->blocksparse_head_sliding_step;
  const Tuning& tn = tuning();
  kp.num_stages = tn.stages;
  kp.flags = tn.evict_normal > 0 ? 1 : 0;

  const bool blocksparse = a->blocksparse_vert_stride > 1;
  if (blocksparse && (a->blocksparse_block_size <= 0 || a->blocksparse_block_size % a->block_size != 0))
    return B200PA_ERR_INVALID_ARG;
  // TMA + mma.sync family: 16-bit queries, block size 16 / 32 (8 for a 16-bit cache), 4-byte aligned query rows; dense,
  // or block-sparse for the head sizes that variant is instantiated for (64, 112, 128; block 16 / 32)
  const bool bs_mma = blocksparse && a->block_size != 8 &&
                      (a->head_size == 64 || a->head_size == 112 || a->head_size == 128);
  const bool b8_mma = a->block_size != 8 || (a->kv_dtype == B200PA_KV_AUTO && !blocksparse);
  const bool mma_ok = layout == 1 ||
                      ((!blocksparse || bs_mma) && a->dtype != B200PA_F32 && b8_mma && tn.force_simt == 0 &&
                       reinterpret_cast<uintptr_t>(a->query) % 4 == 0 && a->q_stride % 2 == 0);

  // ---- split plan -----------------------------------------------------------------------
  // One CTA streams the KV of one (sequence, KV head[, split]) for the whole GQA group (mma family)
  // or of one (sequence, query head[, split]) (SIMT family).  The context is split when those units
  // alone cannot give every SM several CTAs.  Splits are multiples of the reference's 512-token
  // partition, so that v2's caller-sized scratch (ceil(max_seq_len/512) slots) always fits.
  //   * mma family: v1 and v2 split alike.  v2 keeps its partials in the caller's scratch, v1 (whose ABI
  //     has none) in the per-device workspace; the reduce is fused into the kernel (the last split of a
  //     (sequence, KV head) to finish merges them), so both are ONE launch.
  //   * SIMT family: v2 only, with the separate reduce kernel.
  AttnWorkspace ws;
  if (mma_ok) {
    int dev = 0;
    B200PA_CUDA_CHECK(cudaGetDevice(&dev));
    const int rc = attn_workspace(dev, &ws);
    if (rc != B200PA_OK) return rc;
  }
  kp.num_parts = 1;
  kp.partition_size = std::max(a->max_seq_len, 1);
  kp.scratch_parts = 1;
  const int hpt = (a->num_heads / a->num_kv_heads) <= 8 ? 8 : 16;
  const int q_tiles = mma_ok ? (a->num_heads / a->num_kv_heads + hpt - 1) / hpt : 1;
  const int64_t units = (int64_t)a->num_seqs * (mma_ok ? a->num_kv_heads * q_tiles : a->num_heads);
  const bool may_split = mma_ok ? units <= ws.max_counters : is_v2;
  if (may_split && a->max_seq_len > kRefPartition) {
    int psize;
    if (mma_ok) {
      // Splits p trade the tail of the launch (the last CTAs to finish leave SMs idle for about half a CTA's
      // duration, L / p tokens) against the fixed cost of every CTA (start-up, drain, merge: worth ~split_cost
      // tokens of streaming), units * p / slots times: p* = sqrt(L * slots / (units * split_cost)).
      // Measured optimum on the BASELINE shapes (profiles/r02_tune_grid_fused.log): config 2 -> 1, config 3 at
      // TP2/4/8 -> 2/3/4, config 1 -> 1, one 32k sequence -> 32+.
      const double slots = 2.0 * ws.num_sms;
      const double pstar = std::sqrt((double)a->max_seq_len * slots / ((double)units * tn.split_cost));
      const int want = std::max(1, (int)(pstar + 0.5));
      psize = (a->max_seq_len + want - 1) / want;
    } else {
      const int64_t target = (int64_t)148 * 8;  // SIMT family: ~8 CTAs per SM
      const int want = (int)std::max<int64_t>(1, (target + units - 1) / units);
      psize = (a->max_seq_len + want - 1) / want;
    }
    psize = std::min(psize, tn.max_partition);
    if (tn.partition_size > 0) psize = tn.partition_size;
    psize = round_up(std::max(psize, 1), kRefPartition);
    if (mma_ok) psize = std::max(psize, round_up((a->max_seq_len + kMmaMaxSplits - 1) / kMmaMaxSplits, kRefPartition));
    if (mma_ok && !is_v2) {
      // v1: the partials must fit the workspace — [S, Hq, parts, D] T + 2 x [S, Hq, parts] float
      const int64_t per_part = (int64_t)a->num_seqs * a->num_heads * ((int64_t)a->head_size * elem + 8);
      const int64_t max_parts = std::max<int64_t>(1, (int64_t)ws.scratch_bytes / std::max<int64_t>(per_part, 1));
      psize = std::max(psize, round_up((int)((a->max_seq_len + max_parts - 1) / max_parts), kRefPartition));
    }
    const int parts = (a->max_seq_len + psize - 1) / psize;
    if (parts > 1) {
      kp.num_parts = parts;  // v2: parts <= max_num_partitions because psize is a multiple of 512
      kp.partition_size = psize;
    }
  }
  if (kp.num_parts > 1) {
    if (is_v2) {
      kp.exp_sums = a->exp_sums;
      kp.max_logits = a->max_logits;
      kp.tmp_out = a->tmp_out;
      kp.scratch_parts = a->max_num_partitions;
    } else {
      const int64_t rows = (int64_t)a->num_seqs * a->num_heads * kp.num_parts;
      kp.tmp_out = ws.scratch;
      kp.max_logits = reinterpret_cast<float*>(ws.scratch + rows * a->head_size * elem);
      kp.exp_sums = kp.max_logits + rows;
      kp.scratch_parts = kp.num_parts;
    }
    kp.sk_counters = ws.counters;
  }

  int rc = 1;
  if (mma_ok) {
    rc = layout == 1 ? launch_attention_flash(kp, a->num_seqs, a->dtype, stream)
                     : launch_attention_mma(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
    if (layout == 1 && rc == 1) return B200PA_ERR_HEAD_SIZE;  // (no other family reads this layout)
    if (rc == B200PA_OK) {
      g_last_attention_path = layout == 1 ? "mma_flash" : "mma_tma";
      g_last_attention_partitions = kp.num_parts;
      return B200PA_OK;  // (the reduce is part of the kernel)
    }
    if (rc != 1) return rc;
    if (!is_v2) {  // no instantiation for this shape: the SIMT family runs v1 unsplit
      kp.num_parts = 1;
      kp.partition_size = std::max(a->max_seq_len, 1);
    }
  }
  rc = launch_attention_simt(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
  if (rc != B200PA_OK) return rc;
  g_last_attention_path = "simt";
  g_last_attention_partitions = kp.num_parts;
  if (kp.num_parts > 1) return launch_attention_reduce(kp, a->num_seqs, a->dtype, stream);
  return B200PA_OK;
}

}  // namespace b200pa

extern "C" int b200pa_paged_attention_v1(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*is_v2=*/false, stream);
}

extern "C" int b200pa_paged_attention_v2(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*is_v2=*/true, stream);
}

extern "C" int b200pa_paged_attention_flash(const b200pa_attn_params_t* p, int64_t kv_token_stride,
                                            b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*is_v2=*/false, stream, /*layout=*/1, kv_token_stride);
}

// Allocate the per-device attention workspace now (e.g. before a CUDA-graph capture).
extern "C" int b200pa_prepare(int32_t device) {
  b200pa::AttnWorkspace ws;
  return b200pa::attn_workspace(device, &ws);
}

// Developer hook (tools/tune.py, tests): override one tuning knob at run time.  Keys are the
// environment names without the B200PA_ prefix, lower case.  Returns 0, or -5 for an unknown key.
extern "C" int b200pa_set_tuning(const char* key, int32_t value) {
  if (!key) return B200PA_ERR_INVALID_ARG;
  b200pa::Tuning& t = b200pa::tuning();
  struct {
    const char* k;
    int* v;
  } map[] = {{"stages", &t.stages},
             {"evict_normal", &t.evict_normal},
             {"force_simt", &t.force_simt},
             {"partition_size", &t.partition_size},
             {"max_partition", &t.max_partition},
             {"split_cost", &t.split_cost}};
  for (auto& m : map)
    if (std::strcmp(m.k, key) == 0) {
      *m.v = value;
      return B200PA_OK;
    }
  return B200PA_ERR_INVALID_ARG;
}
