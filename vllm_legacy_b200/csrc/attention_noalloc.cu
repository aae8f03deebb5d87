// b200pa_paged_attention_static_scratch (include/b200pa.h): the split-K decode attention of the TMA + mma.sync
// family for callers that must not allocate device memory — code under CUDA-graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  It has NO counterpart in the
// reference.
//
// The regular entry points keep the split-K partials in the caller's v2 scratch or in a per-device workspace that the
// first call cudaMallocs (attention.cu).  Here the partials AND the arrival counters of the fused reduce live in one
// file-scope `static __device__` array of a fixed, modest size, so this entry point calls no cudaMalloc / cudaFree and
// no allocating library routine.  Every GPU that loads the library holds the array (8 MiB + 16 KiB) until exit.
//
// Rules that follow from the array being shared and persistent:
//   * every call uses the same memory: two calls that overlap on the GPU (different streams, or graphs replayed
//     concurrently) RACE on it — serialise them;
//   * it keeps the last call's contents: the partials are fully overwritten before they are read, the arrival
//     counters must start at zero and are cleared on `stream` at the start of every call (a memset node under capture);
//   * problems whose partials or counters do not fit are rejected with B200PA_ERR_SCRATCH_LIMIT, never truncated.
//
// In host code the array's NAME is a host-side stand-in, not a device address: its device address is fetched once per
// device, with that device current, through cudaGetSymbolAddress (return code checked) and that pointer is what the
// kernel gets.
#include <algorithm>
#include <cmath>
#include <mutex>

#include "attention_common.cuh"

namespace b200pa {
namespace {

constexpr size_t kPartialBytes = (size_t)B200PA_STATIC_SCRATCH_BYTES;  // tmp_out + max_logits + exp_sums
constexpr int kCounters = B200PA_STATIC_SCRATCH_COUNTERS;              // one per (sequence, KV head x q-tile)
constexpr size_t kArrayBytes = kPartialBytes + (size_t)kCounters * sizeof(unsigned);
static_assert(kPartialBytes % 256 == 0, "the counters start 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kArrayBytes];

// Device address of g_static_scratch on `device` (which must be the current device), resolved once per device.
int static_scratch_base(int device, unsigned char** out) {
  constexpr int kMaxDevices = 64;
  static unsigned char* table[kMaxDevices] = {nullptr};
  static std::mutex mu;
  if (device < 0 || device >= kMaxDevices) return B200PA_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!table[device]) {
    unsigned char* ptr = nullptr;
    B200PA_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_scratch));
    if (!ptr) return B200PA_ERR_INVALID_ARG;
    table[device] = ptr;
  }
  *out = table[device];
  return B200PA_OK;
}

int round_up(int v, int m) { return (v + m - 1) / m * m; }

}  // namespace

int run_attention_static_scratch(const b200pa_attn_params_t* a, cudaStream_t stream) {
  if (!a) return B200PA_ERR_INVALID_ARG;
  if (a->kv_dtype != B200PA_KV_AUTO && a->kv_dtype != B200PA_KV_FP8_E4M3 && a->kv_dtype != B200PA_KV_FP8_E5M2)
    return B200PA_ERR_KV_DTYPE;
  // the TMA + mma.sync family only: 16-bit queries, block 16 / 32, dense (what every BASELINE config uses)
  if (a->dtype != B200PA_F16 && a->dtype != B200PA_BF16) return B200PA_ERR_INPUT_DTYPE;
  if (a->block_size != 16 && a->block_size != 32) return B200PA_ERR_BLOCK_SIZE;
  if (a->num_seqs < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0 ||
      a->max_seq_len < 0 || a->max_num_blocks_per_seq < 0 || a->head_size <= 0 || a->blocksparse_vert_stride > 1)
    return B200PA_ERR_INVALID_ARG;
  if (a->num_seqs == 0) return B200PA_OK;
  if (!a->out || !a->query || !a->key_cache || !a->value_cache || !a->block_tables || !a->seq_lens)
    return B200PA_ERR_INVALID_ARG;
  const int elem = 2;
  const int cache_bytes = a->kv_dtype == B200PA_KV_AUTO ? elem : 1;
  if (reinterpret_cast<uintptr_t>(a->key_cache) % 16 || reinterpret_cast<uintptr_t>(a->value_cache) % 16 ||
      (a->kv_block_stride * cache_bytes) % 16 || (a->kv_head_stride * cache_bytes) % 16 ||
      reinterpret_cast<uintptr_t>(a->query) % 4 || a->q_stride % 2)
    return B200PA_ERR_INVALID_ARG;

  // ---- limits of the static array: rejected, never truncated ---------------------------------------------------
  const int group = a->num_heads / a->num_kv_heads;
  const int hpt = group <= 8 ? 8 : 16;
  const int q_tiles = (group + hpt - 1) / hpt;
  const int64_t units = (int64_t)a->num_seqs * a->num_kv_heads * q_tiles;
  const int64_t per_part = (int64_t)a->num_seqs * a->num_heads * ((int64_t)a->head_size * elem + 8);
  if (units > kCounters || 2 * per_part > (int64_t)kPartialBytes) return B200PA_ERR_SCRATCH_LIMIT;

  int dev = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  unsigned char* base = nullptr;
  int rc = static_scratch_base(dev, &base);
  if (rc != B200PA_OK) return rc;
  int sms = 0;
  B200PA_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));

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
  kp.num_heads = a->num_heads;
  kp.num_kv_heads = a->num_kv_heads;
  kp.head_size = a->head_size;
  kp.block_size = a->block_size;
  kp.max_blocks_per_seq = a->max_num_blocks_per_seq;
  kp.scale = a->scale;
  kp.kv_scale = a->kv_scale;
  kp.tp_rank = a->tp_rank;
  kp.bs_block_size = a->blocksparse_block_size;
  const Tuning& tn = tuning();
  kp.num_stages = tn.stages;
  kp.flags = tn.evict_normal > 0 ? 1 : 0;

  // ---- split plan: the one of attention.cu (tail of the launch against the fixed cost of a CTA), then capped by
  // what the static partial region holds; splits are multiples of 512 tokens like everywhere else.
  kp.num_parts = 1;
  kp.partition_size = std::max(a->max_seq_len, 1);
  kp.scratch_parts = 1;
  if (a->max_seq_len > 512) {
    const double slots = 2.0 * sms;
    const double pstar = std::sqrt((double)a->max_seq_len * slots / ((double)units * std::max(1, tn.split_cost)));
    const int want = std::max(1, (int)(pstar + 0.5));
    int psize = round_up(std::max((a->max_seq_len + want - 1) / want, 1), 512);
    psize = std::max(psize, round_up((a->max_seq_len + kMmaMaxSplits - 1) / kMmaMaxSplits, 512));
    const int64_t max_parts = (int64_t)kPartialBytes / per_part;  // >= 2 (checked above)
    psize = std::max(psize, round_up((int)((a->max_seq_len + max_parts - 1) / max_parts), 512));
    const int parts = (a->max_seq_len + psize - 1) / psize;
    if (parts > 1) {
      kp.num_parts = parts;
      kp.partition_size = psize;
    }
  }
  if (kp.num_parts > 1) {
    const int64_t rows = (int64_t)a->num_seqs * a->num_heads * kp.num_parts;
    kp.tmp_out = base;
    kp.max_logits = reinterpret_cast<float*>(base + rows * a->head_size * elem);
    kp.exp_sums = kp.max_logits + rows;
    kp.scratch_parts = kp.num_parts;
    kp.sk_counters = reinterpret_cast<unsigned*>(base + kPartialBytes);
    // the array keeps whatever the last call (possibly an aborted one) left: the arrival counters must start at zero
    B200PA_CUDA_CHECK(cudaMemsetAsync(kp.sk_counters, 0, (size_t)units * sizeof(unsigned), stream));
  }
  rc = launch_attention_mma(kp, a->num_seqs, a->dtype, a->kv_dtype, stream);
  if (rc == 1) return B200PA_ERR_HEAD_SIZE;  // no instantiation of the mma family for this head size
  if (rc != B200PA_OK) return rc;
  g_last_attention_path = "mma_tma_static";
  g_last_attention_partitions = kp.num_parts;
  return B200PA_OK;
}

}  // namespace b200pa

extern "C" int b200pa_paged_attention_static_scratch(const b200pa_attn_params_t* p, b200pa_stream_t stream) {
  return b200pa::run_attention_static_scratch(p, stream);
}
