// C-ABI entry points b200pa_paged_attention_v1 / _v2 (include/b200pa.h): argument checks with
// the reference's error classes, launch planning, dispatch to the persistent TMA + mma.sync
// kernel or the general SIMT family (+ its split-K reduce), the per-device workspace of the
// persistent kernel, and the tuning knobs (environment read once, never on the launch path).
//
// Reference host side: csrc/attention/attention_kernels.cu:690-825 (v1), :844-1014 (v2).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "attention_common.cuh"
#include "attention_sched.cuh"

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
    v.grid_ctas = env_int("B200PA_GRID_CTAS", v.grid_ctas);
    v.min_blocks_per_cta = std::max(1, env_int("B200PA_MIN_BLOCKS_PER_CTA", v.min_blocks_per_cta));
    v.grid_kernel = env_int("B200PA_GRID_KERNEL", v.grid_kernel);
    v.pool_pct = std::min(100, std::max(0, env_int("B200PA_POOL_PCT", v.pool_pct)));
    v.chunks_per_cta = std::max(1, env_int("B200PA_CHUNKS_PER_CTA", v.chunks_per_cta));
    v.cs_min = std::max(1, env_int("B200PA_CS_MIN", v.cs_min));
    return v;
  }();
  return t;
}

// ---- per-device workspace of the persistent kernel -------------------------------------------------
// 2 partial-result slots per resident CTA + one arrival counter per CTA (~10 MB per device),
// allocated on the first attention call on a device (or by b200pa_prepare()) and kept for the
// life of the process.  Calls on one device must not overlap in time on different streams (they
// never do in the reference's worker: one stream, one step at a time).
static int alloc_workspace(int device, SkWorkspace* ws) {
  int sms = 0;
  B200PA_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  ws->num_sms = sms;
  ws->max_counters = kSkMaxCounters;
  ws->slot_bytes = kSkWorkspaceSlotBytes;
  const size_t slot_bytes = ws->slot_bytes;
  const size_t counter_bytes = ((size_t)ws->max_counters + 2) * sizeof(unsigned);
  // The first call may sit inside a stream capture (global mode forbids cudaMalloc): allocate in
  // relaxed mode and zero the counters on a private stream, so nothing is recorded in the graph.
  cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
  B200PA_CUDA_CHECK(cudaThreadExchangeStreamCaptureMode(&mode));
  int prev = 0;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess) e = cudaSetDevice(device);
  char* base = nullptr;
  cudaStream_t s = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&base, slot_bytes + counter_bytes);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMemsetAsync(base + slot_bytes, 0, counter_bytes, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (s) cudaStreamDestroy(s);
  cudaSetDevice(prev);
  cudaThreadExchangeStreamCaptureMode(&mode);
  if (e != cudaSuccess) {
    g_last_cuda_error = static_cast<int>(e);
    return B200PA_ERR_CUDA;
  }
  ws->slots = reinterpret_cast<float*>(base);
  ws->counters = reinterpret_cast<unsigned*>(base + slot_bytes);
  ws->pool = ws->counters + ws->max_counters;
  return B200PA_OK;
}

int sk_workspace(int device, SkWorkspace* out) {
  constexpr int kMaxDevices = 64;
  static SkWorkspace table[kMaxDevices];
  static std::mutex mu;
  if (device < 0 || device >= kMaxDevices) return B200PA_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!table[device].slots) {
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
constexpr int kNumSMs = 148;

static int run_attention(const b200pa_attn_params_t* a, bool allow_split, cudaStream_t stream) {
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->kv_dtype != B200PA_KV_AUTO && a->kv_dtype != B200PA_KV_FP8_E4M3 && a->kv_dtype != B200PA_KV_FP8_E5M2)
    return B200PA_ERR_KV_DTYPE;
  if (a->dtype != B200PA_F16 && a->dtype != B200PA_BF16 && a->dtype != B200PA_F32) return B200PA_ERR_INPUT_DTYPE;
  if (!block_size_supported(a->block_size)) return B200PA_ERR_BLOCK_SIZE;
  if (!head_size_supported(a->head_size)) return B200PA_ERR_HEAD_SIZE;
  if (a->num_seqs < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0 ||
      a->max_seq_len < 0 || a->max_num_blocks_per_seq < 0)
    return B200PA_ERR_INVALID_ARG;
  if (a->num_seqs == 0) return B200PA_OK;
  if (!a->out || !a->query || !a->key_cache || !a->value_cache || !a->block_tables || !a->seq_lens)
    return B200PA_ERR_INVALID_ARG;
  const int cache_bytes = a->kv_dtype == B200PA_KV_AUTO ? (a->dtype == B200PA_F32 ? 4 : 2) : 1;
  // 16-byte granularity of the paged layout (x = 16 / sizeof(cache_t)): the reference's vector
  // loads assume it too (attention_kernels.cu:200, 272-290).
  if (reinterpret_cast<uintptr_t>(a->key_cache) % 16 || reinterpret_cast<uintptr_t>(a->value_cache) % 16 ||
      (a->kv_block_stride * cache_bytes) % 16 || (a->kv_head_stride * cache_bytes) % 16)
    return B200PA_ERR_INVALID_ARG;

  AttnKernelParams kp{};
  kp.out = a->out;
  kp.exp_sums = a->exp_sums;
  kp.max_logits = a->max_logits;
  kp.tmp_out = a->tmp_out;
  kp.q = a->query;
  kp.k_cache = a->key_cache;
  kp.v_cache = a->value_cache;
  kp.block_tables = a->block_tables;
  kp.seq_lens = a->seq_lens;
  kp.alibi_slopes = a->alibi_slopes;
  kp.q_stride = a->q_stride;
  kp.kv_block_stride = a->kv_block_stride;
  kp.kv_head_stride = a->kv_head_stride;
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
  kp.num_seqs = a->num_seqs;
  kp.max_seq_len = a->max_seq_len;
  const Tuning& tn = tuning();
  kp.num_stages = tn.stages;
  kp.flags = tn.evict_normal > 0 ? 1 : 0;

  const bool blocksparse = a->blocksparse_vert_stride > 1;
  if (blocksparse && (a->blocksparse_block_size <= 0 || a->blocksparse_block_size % a->block_size != 0))
    return B200PA_ERR_INVALID_ARG;
  const bool mma_ok = !blocksparse && a->dtype != B200PA_F32 && a->block_size != 8 &&
                      tn.force_simt == 0;

  // ---- persistent TMA + mma.sync kernel: one launch, stream-K split over the resident CTAs --------
  // (v1 and v2 are the same launch: the split needs no caller scratch, so v2's exp_sums /
  // max_logits / tmp_out stay untouched.)
  int rc = 1;
  if (mma_ok && !tn.grid_kernel) {
    int grid = 0;
    rc = launch_attention_sk(kp, a->num_seqs, a->dtype, a->kv_dtype, stream, &grid);
    if (rc == B200PA_OK) {
      g_last_attention_path = "mma_tma";
      g_last_attention_partitions = grid;
      return B200PA_OK;
    }
    if (rc != 1) return rc;
  }

  // ---- split plan of the grid kernels (general SIMT family) ----------------------------------------
  // One CTA streams the KV of one (sequence, query head[, split]).  Split the context when those
  // units alone cannot give every SM several CTAs; never below the reference's 512-token
  // partition so the caller-sized scratch (ceil(max_seq_len/512) slots) always fits.
  kp.num_parts = 1;
  kp.partition_size = std::max(a->max_seq_len, 1);
  kp.scratch_parts = std::max(a->max_num_partitions, 1);
  const bool grid_mma = mma_ok && tn.grid_kernel;
  if (allow_split) {
    if (!a->exp_sums || !a->max_logits || !a->tmp_out || a->max_num_partitions <= 0) return B200PA_ERR_INVALID_ARG;
    const int64_t units = (int64_t)a->num_seqs * (grid_mma ? a->num_kv_heads : a->num_heads);
    const int64_t target = (int64_t)kNumSMs * 8;
    int want = (int)std::min<int64_t>((target + units - 1) / units, a->max_num_partitions);
    int psize = (a->max_seq_len + want - 1) / std::max(want, 1);
    psize = std::min(psize, tn.max_partition);
    psize = (psize + kRefPartition - 1) / kRefPartition * kRefPartition;
    const int forced = tn.partition_size;
    if (forced > 0) psize = (forced + kRefPartition - 1) / kRefPartition * kRefPartition;
    const int parts = (a->max_seq_len + psize - 1) / psize;
    if (parts > 1) {
      kp.num_parts = parts;  // parts <= max_num_partitions because psize >= 512
      kp.partition_size = psize;
    }
  }

  rc = 1;
  if (grid_mma) {
    rc = launch_attention_mma(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
    if (rc == B200PA_OK) g_last_attention_path = "mma_grid";
  }
  if (rc == 1) {
    rc = launch_attention_simt(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
    if (rc == B200PA_OK) g_last_attention_path = "simt";
  }
  if (rc != B200PA_OK) return rc;
  g_last_attention_partitions = kp.num_parts;
  if (kp.num_parts > 1) return launch_attention_reduce(kp, a->num_seqs, a->dtype, stream);
  return B200PA_OK;
}

}  // namespace b200pa

extern "C" int b200pa_paged_attention_v1(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*allow_split=*/false, stream);
}

extern "C" int b200pa_paged_attention_v2(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention(p, /*allow_split=*/true, stream);
}

namespace b200pa { long long* g_sk_trace = nullptr; }
// Developer hook: device buffer of [grid][4] int64 the persistent kernel fills with {smid, t_begin, t_end, blocks}
// per CTA (globaltimer ns); NULL switches the trace off.
extern "C" void b200pa_set_trace(void* device_buffer) { b200pa::g_sk_trace = static_cast<long long*>(device_buffer); }

// Allocate the per-device workspace of the persistent kernel now (e.g. before a CUDA-graph capture).
extern "C" int b200pa_prepare(int32_t device) {
  b200pa::SkWorkspace ws;
  return b200pa::sk_workspace(device, &ws);
}

// Developer hook (tools/tune.py, tests): override one tuning knob at run time.  Keys are the
// environment names without the B200PA_ prefix, lower case.  Returns 0, or -5 for an unknown key.
extern "C" int b200pa_set_tuning(const char* key, int32_t value) {
  if (!key) return B200PA_ERR_INVALID_ARG;
  b200pa::Tuning& t = b200pa::tuning();
  struct { const char* k; int* v; } map[] = {
      {"stages", &t.stages}, {"evict_normal", &t.evict_normal}, {"force_simt", &t.force_simt},
      {"partition_size", &t.partition_size}, {"max_partition", &t.max_partition}, {"grid_ctas", &t.grid_ctas},
      {"min_blocks_per_cta", &t.min_blocks_per_cta}, {"grid_kernel", &t.grid_kernel}, {"pool_pct", &t.pool_pct},
      {"chunks_per_cta", &t.chunks_per_cta}, {"cs_min", &t.cs_min}};
  for (auto& m : map)
    if (std::strcmp(m.k, key) == 0) {
      *m.v = value;
      return B200PA_OK;
    }
  return B200PA_ERR_INVALID_ARG;
}

// Index arithmetic of the persistent kernel's work split, exported for the CPU test suite.
extern "C" int64_t b200pa_sched_range_begin(int64_t W, int32_t G, int32_t c) { return b200pa::sk_range_begin(W, G, c); }
extern "C" int32_t b200pa_sched_cta_of(int64_t W, int32_t G, int64_t f) { return b200pa::sk_cta_of(W, G, f); }
extern "C" int32_t b200pa_sched_effective_ctas(int64_t W, int32_t grid) { return b200pa::sk_effective_ctas(W, grid); }
// geometry of the chunk list: out[0..4] = {W1, Gs, cs, PC, number of chunks}
extern "C" void b200pa_sched_geometry(int64_t W, int32_t grid, int32_t pool_pct, int32_t chunks_per_cta, int32_t cs_min,
                                      int32_t max_chunks, int64_t* out) {
  const b200pa::SkGeom g = b200pa::sk_geometry(W, grid, pool_pct, chunks_per_cta, cs_min, max_chunks);
  out[0] = g.W1; out[1] = g.Gs; out[2] = g.cs; out[3] = g.PC; out[4] = b200pa::sk_num_chunks(g);
}
extern "C" int64_t b200pa_sched_chunk_begin(int64_t W, int32_t grid, int32_t pool_pct, int32_t chunks_per_cta,
                                            int32_t cs_min, int32_t max_chunks, int32_t chunk) {
  return b200pa::sk_chunk_begin(b200pa::sk_geometry(W, grid, pool_pct, chunks_per_cta, cs_min, max_chunks), chunk);
}
extern "C" int32_t b200pa_sched_chunk_of(int64_t W, int32_t grid, int32_t pool_pct, int32_t chunks_per_cta,
                                         int32_t cs_min, int32_t max_chunks, int64_t f) {
  return b200pa::sk_chunk_of(b200pa::sk_geometry(W, grid, pool_pct, chunks_per_cta, cs_min, max_chunks), f);
}
