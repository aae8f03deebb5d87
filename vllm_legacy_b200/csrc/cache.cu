// KV-cache maintenance kernels for B200 (sm_100a): reshape_and_cache (+fp8 quantisation),
// copy_blocks, swap_blocks, convert_fp8.  C-ABI entry points are declared in include/b200pa.h.
//
// Semantics follow the reference (DeepAuto-AI/vllm-legacy):
//   reshape_and_cache  csrc/cache_kernels.cu:150-202, 241-287
//   copy_blocks        csrc/cache_kernels.cu:68-146
//   swap_blocks        csrc/cache_kernels.cu:25-64
//   convert_fp8        csrc/cache_kernels.cu:347-420
// but not their structure: 128-bit accesses where the layout allows, pointer tables passed as
// kernel parameters (no per-call H2D copy / sync), swap runs merged into large async copies.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace b200pa {

thread_local int g_last_cuda_error = 0;
thread_local const char* g_last_attention_path = "";
thread_local int g_last_attention_partitions = 0;

// =============================================================================================
// reshape_and_cache
// =============================================================================================

// Generic path: any x, any alignment.  One CTA per token, one element per thread-iteration.
// Index math as cache_kernels.cu:169-200.
template <typename T, int KV>
__global__ void reshape_and_cache_scalar_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                                typename CacheStore<T, KV>::type* __restrict__ key_cache,
                                                typename CacheStore<T, KV>::type* __restrict__ value_cache,
                                                const int64_t* __restrict__ slot_mapping,
                                                int64_t key_stride, int64_t value_stride, int num_heads,
                                                int head_size, int block_size, int x, float kv_scale) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding token
  const int64_t block_idx = slot / block_size;
  const int64_t block_off = slot % block_size;
  const int n = num_heads * head_size;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int h = i / head_size;
    const int d = i % head_size;
    const int64_t kdst = block_idx * num_heads * (head_size / x) * block_size * x +
                         (int64_t)h * (head_size / x) * block_size * x + (int64_t)(d / x) * block_size * x +
                         block_off * x + (d % x);
    const int64_t vdst = block_idx * num_heads * head_size * block_size + (int64_t)h * head_size * block_size +
                         (int64_t)d * block_size + block_off;
    key_cache[kdst] = CacheStore<T, KV>::cvt(key[token * key_stride + i], kv_scale);
    value_cache[vdst] = CacheStore<T, KV>::cvt(value[token * value_stride + i], kv_scale);
  }
}

// Vector path: x * sizeof(cache_t) == 16 and 16-byte aligned sources.  Each thread-iteration
// moves one x-group of K with 128-bit loads and ONE 128-bit store (the K layout keeps the x
// elements of a token contiguous), and x elements of V with 128-bit loads and x element stores
// (the V layout is token-minor, so one token's elements are block_size apart by construction).
template <typename T, int KV>
__global__ void reshape_and_cache_vec_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                             typename CacheStore<T, KV>::type* __restrict__ key_cache,
                                             typename CacheStore<T, KV>::type* __restrict__ value_cache,
                                             const int64_t* __restrict__ slot_mapping, int64_t key_stride,
                                             int64_t value_stride, int num_heads, int head_size,
                                             int block_size, float kv_scale) {
  using CT = typename CacheStore<T, KV>::type;
  constexpr int X = 16 / sizeof(CT);                // elements per K group
  constexpr int SRC_VECS = X * sizeof(T) / 16;      // int4 loads per group
  constexpr int EPV = 16 / sizeof(T);               // source elements per int4
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t block_idx = slot / block_size;
  const int block_off = (int)(slot % block_size);
  const int groups_per_head = head_size / X;
  const int n_groups = num_heads * groups_per_head;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  CT* kblk = key_cache + block_idx * (int64_t)num_heads * head_size * block_size;
  CT* vblk = value_cache + block_idx * (int64_t)num_heads * head_size * block_size;

  // 2 * n_groups work items: the K groups, then the V groups (a decode step is latency-bound: more threads
  // per token, each with one load and its stores, finish sooner than fewer threads doing K and V in turn)
  for (int w = threadIdx.x; w < 2 * n_groups; w += blockDim.x) {
    const bool is_v = w >= n_groups;
    const int g = is_v ? w - n_groups : w;
    const int h = g / groups_per_head;
    const int dx = g % groups_per_head;
    const int e0 = h * head_size + dx * X;  // first source element of the group
    union { int4 v[SRC_VECS]; T e[X]; } in;
    const T* src = (is_v ? vsrc : ksrc) + e0;
#pragma unroll
    for (int j = 0; j < SRC_VECS; ++j) in.v[j] = ld_nc_v4(src + j * EPV);
    if (!is_v) {
      union { int4 v; CT e[X]; } kout;
#pragma unroll
      for (int j = 0; j < X; ++j) kout.e[j] = CacheStore<T, KV>::cvt(in.e[j], kv_scale);
      CT* kdst = kblk + ((int64_t)(h * groups_per_head + dx) * block_size + block_off) * X;
      *reinterpret_cast<int4*>(kdst) = kout.v;
    } else {
      CT* vdst = vblk + ((int64_t)h * head_size + dx * X) * block_size + block_off;
#pragma unroll
      for (int j = 0; j < X; ++j) vdst[(int64_t)j * block_size] = CacheStore<T, KV>::cvt(in.e[j], kv_scale);
    }
  }
}


// Tile path (prefill / chunked prefill, num_tokens >= kTileMinTokens): one CTA stages TT = 16
// consecutive tokens x (heads_per_cta * D) elements of K and V in shared memory (coalesced 128-bit row
// reads, converted to the cache type on the way in), then writes them out with the TOKEN index fastest
// across lanes.  When the tile's slots are consecutive inside a block — the normal case for a prompt —
// a warp's K stores cover 512 contiguous bytes ([dx][o][x]: 16 tokens x 16 B, next dx right behind) and
// its V stores fill whole 32-byte sectors ([d][o] rows), instead of one 16-B / 2-B piece per sector.
// Arbitrary slot mappings stay correct (the mapping only changes which addresses coalesce).
// Shared rows are padded (K: +16 B, V: +4 B) so both transposed reads are bank-conflict free.
constexpr int kTileTokens = 16;
constexpr int kTileMinTokens = 512;
constexpr int kTileThreads = 256;

template <typename T, int KV>
__global__ void __launch_bounds__(kTileThreads)
    reshape_and_cache_tile_kernel(const T* __restrict__ key, const T* __restrict__ value,
                                  typename CacheStore<T, KV>::type* __restrict__ key_cache,
                                  typename CacheStore<T, KV>::type* __restrict__ value_cache,
                                  const int64_t* __restrict__ slot_mapping, int64_t key_stride, int64_t value_stride,
                                  int num_tokens, int num_heads, int head_size, int block_size, int heads_per_cta,
                                  float kv_scale) {
  using CT = typename CacheStore<T, KV>::type;
  constexpr int TT = kTileTokens;
  constexpr int X = 16 / sizeof(CT);        // cache elements per 16-byte K group
  constexpr int EPV = 16 / sizeof(T);       // source elements per 128-bit load
  constexpr int VPAD = 4 / sizeof(CT);      // one 32-bit word of padding per V row
  extern __shared__ __align__(16) unsigned char tile_smem[];
  __shared__ int64_t slot_s[TT];

  const int t0 = blockIdx.x * TT;
  const int h0 = blockIdx.y * heads_per_cta;
  const int nh = min(heads_per_cta, num_heads - h0);
  const int n = nh * head_size;             // elements per token handled here
  const int rsk = heads_per_cta * head_size + X;     // K row stride (elements): 16-byte aligned rows
  const int rsv = heads_per_cta * head_size + VPAD;  // V row stride (elements): odd number of words
  CT* ks = reinterpret_cast<CT*>(tile_smem);
  CT* vs = ks + TT * rsk;

  if (threadIdx.x < TT) {
    const int t = t0 + threadIdx.x;
    slot_s[threadIdx.x] = t < num_tokens ? slot_mapping[t] : -1;
  }
  __syncthreads();

  // ---- stage: coalesced 128-bit row reads -> cache type -> shared ------------------------------------
  const int vec_per_tok = n / EPV;
  for (int idx = threadIdx.x; idx < TT * vec_per_tok; idx += kTileThreads) {
    const int j = idx / vec_per_tok;
    const int c = idx - j * vec_per_tok;
    if (slot_s[j] < 0) continue;
    const int64_t t = t0 + j;
    union { int4 v; T e[EPV]; } kin, vin;
    kin.v = ld_nc_v4(key + t * key_stride + (int64_t)h0 * head_size + c * EPV);
    vin.v = ld_nc_v4(value + t * value_stride + (int64_t)h0 * head_size + c * EPV);
    union { CT e[EPV]; uint32_t w[EPV * sizeof(CT) / 4]; } ko, vo;
#pragma unroll
    for (int e = 0; e < EPV; ++e) {
      ko.e[e] = CacheStore<T, KV>::cvt(kin.e[e], kv_scale);
      vo.e[e] = CacheStore<T, KV>::cvt(vin.e[e], kv_scale);
    }
    uint32_t* kd = reinterpret_cast<uint32_t*>(ks + j * rsk + c * EPV);
    uint32_t* vd = reinterpret_cast<uint32_t*>(vs + j * rsv + c * EPV);
#pragma unroll
    for (int w = 0; w < (int)(EPV * sizeof(CT) / 4); ++w) {
      kd[w] = ko.w[w];
      vd[w] = vo.w[w];
    }
  }
  __syncthreads();

  // ---- write out ---------------------------------------------------------------------------------------------
  // dst(K) = kc + b*H*D*B + ((h*G + dx)*B + o)*X with h*G + dx = h0*G + u   (G = D/X groups per head)
  // dst(V) = vc + b*H*D*B + (h*D + d)*B + o      with h*D + d  = h0*D + e
  const int64_t blk_elems = (int64_t)num_heads * head_size * block_size;
  const int groups_per_head = head_size / X;
  const int n_groups = n / X;
  const int64_t kstep = (int64_t)block_size * X;
  // Is the tile one run of consecutive slots (a prompt filling its blocks in order)?  Then the TOKEN index goes
  // fastest across lanes: a warp's K stores cover 512 contiguous bytes and its V stores fill whole 32-byte
  // sectors.  Scattered slots (a decode batch, a shuffled mapping) gain nothing from that order — every token
  // lands in another block — so they are written token by token with the row index fastest, which keeps the 256
  // threads on one token's rows (the reference's order, cache_kernels.cu:169-200) and is what the L2 merges best.
  bool run = true;
#pragma unroll
  for (int t = 1; t < TT; ++t) run &= (slot_s[t] == slot_s[0] + t);
  run &= slot_s[0] >= 0;
  if (run) {
    // thread = (token j = tid % TT, row lane = tid / TT)
    const int j = threadIdx.x % TT;
    const int r0 = threadIdx.x / TT;
    constexpr int RL = kTileThreads / TT;  // row lanes
    const int64_t slot = slot_s[j];
    const int64_t b = slot / block_size;
    const int o = (int)(slot - b * block_size);
    CT* kdst = key_cache + b * blk_elems + ((int64_t)h0 * groups_per_head * block_size + o) * X;
    CT* vdst = value_cache + b * blk_elems + (int64_t)h0 * head_size * block_size + o;
    const CT* ksrc = ks + j * rsk;
    const CT* vsrc = vs + j * rsv;
#pragma unroll 4
    for (int u = r0; u < n_groups; u += RL)
      *reinterpret_cast<int4*>(kdst + u * kstep) = *reinterpret_cast<const int4*>(ksrc + u * X);
#pragma unroll 8
    for (int e = r0; e < n; e += RL) vdst[(int64_t)e * block_size] = vsrc[e];
    return;
  }
  for (int j = 0; j < TT; ++j) {
    const int64_t slot = slot_s[j];
    if (slot < 0) continue;
    const int64_t b = slot / block_size;
    const int o = (int)(slot - b * block_size);
    CT* kdst = key_cache + b * blk_elems + ((int64_t)h0 * groups_per_head * block_size + o) * X;
    CT* vdst = value_cache + b * blk_elems + (int64_t)h0 * head_size * block_size + o;
    const CT* ksrc = ks + j * rsk;
    const CT* vsrc = vs + j * rsv;
    for (int u = threadIdx.x; u < n_groups; u += kTileThreads)
      *reinterpret_cast<int4*>(kdst + u * kstep) = *reinterpret_cast<const int4*>(ksrc + u * X);
#pragma unroll 4
    for (int e = threadIdx.x; e < n; e += kTileThreads) vdst[(int64_t)e * block_size] = vsrc[e];
  }
}

template <typename T, int KV>
static int launch_reshape(const b200pa_reshape_params_t* p, cudaStream_t stream) {
  using CT = typename CacheStore<T, KV>::type;
  if (p->num_tokens == 0) return B200PA_OK;
  const int n = p->num_heads * p->head_size;
  constexpr int X = 16 / sizeof(CT);
  const bool vec_ok = p->x == X && p->head_size % X == 0 &&
                      (reinterpret_cast<uintptr_t>(p->key) % 16 == 0) &&
                      (reinterpret_cast<uintptr_t>(p->value) % 16 == 0) &&
                      (reinterpret_cast<uintptr_t>(p->key_cache) % 16 == 0) &&
                      (p->key_stride * sizeof(T)) % 16 == 0 && (p->value_stride * sizeof(T)) % 16 == 0;
  dim3 grid(p->num_tokens);
  if (vec_ok && p->num_tokens >= kTileMinTokens && p->head_size % (16 / (int)sizeof(T)) == 0) {
    const int n_max = 2048 / (int)sizeof(CT);  // elements of one token staged per CTA (K and V each)
    if (p->head_size <= n_max) {
      const int hpc = std::min(p->num_heads, std::max(1, n_max / p->head_size));
      const size_t smem = (size_t)kTileTokens * ((size_t)hpc * p->head_size * 2 + X + 4 / sizeof(CT)) * sizeof(CT);
      auto kernel = reshape_and_cache_tile_kernel<T, KV>;
      static int configured[64] = {0};
      int dev = 0;
      B200PA_CUDA_CHECK(cudaGetDevice(&dev));
      if (dev < 0 || dev >= 64 || configured[dev] < (int)smem) {
        B200PA_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev >= 0 && dev < 64) configured[dev] = (int)smem;
      }
      dim3 tgrid((p->num_tokens + kTileTokens - 1) / kTileTokens, (p->num_heads + hpc - 1) / hpc);
      kernel<<<tgrid, kTileThreads, smem, stream>>>(
          static_cast<const T*>(p->key), static_cast<const T*>(p->value), static_cast<CT*>(p->key_cache),
          static_cast<CT*>(p->value_cache), p->slot_mapping, p->key_stride, p->value_stride, p->num_tokens,
          p->num_heads, p->head_size, p->block_size, hpc, p->kv_scale);
      B200PA_CUDA_CHECK(cudaGetLastError());
      return B200PA_OK;
    }
  }
  if (vec_ok) {
    int threads = std::min(std::max(2 * n / X, 32), 512);
    threads = (threads + 31) / 32 * 32;
    reshape_and_cache_vec_kernel<T, KV><<<grid, threads, 0, stream>>>(
        static_cast<const T*>(p->key), static_cast<const T*>(p->value), static_cast<CT*>(p->key_cache),
        static_cast<CT*>(p->value_cache), p->slot_mapping, p->key_stride, p->value_stride, p->num_heads,
        p->head_size, p->block_size, p->kv_scale);
  } else {
    int threads = std::min(n, 512);
    reshape_and_cache_scalar_kernel<T, KV><<<grid, threads, 0, stream>>>(
        static_cast<const T*>(p->key), static_cast<const T*>(p->value), static_cast<CT*>(p->key_cache),
        static_cast<CT*>(p->value_cache), p->slot_mapping, p->key_stride, p->value_stride, p->num_heads,
        p->head_size, p->block_size, p->x, p->kv_scale);
  }
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

template <int KV>
static int dispatch_reshape_dtype(const b200pa_reshape_params_t* p, cudaStream_t stream) {
  switch (p->dtype) {
    case B200PA_F16: return launch_reshape<__half, KV>(p, stream);
    case B200PA_BF16: return launch_reshape<__nv_bfloat16, KV>(p, stream);
    case B200PA_F32: return launch_reshape<float, KV>(p, stream);
    default: return B200PA_ERR_INPUT_DTYPE;
  }
}

// =============================================================================================
// copy_blocks
// =============================================================================================

constexpr int kMaxLayersPerLaunch = 120;  // 2 * 120 * 8 B = 1920 B of kernel parameters

struct LayerPtrs {
  char* k[kMaxLayersPerLaunch];
  char* v[kMaxLayersPerLaunch];
};

// One CTA per (layer, pair): K block then V block, 128-bit accesses, 4 loads in flight per thread.
__global__ void copy_blocks_vec_kernel(const __grid_constant__ LayerPtrs ptrs,
                                       const int64_t* __restrict__ block_mapping, int64_t vecs_per_block) {
  const int layer = blockIdx.x;
  const int64_t pair = blockIdx.y;
  const int64_t src = block_mapping[2 * pair];
  const int64_t dst = block_mapping[2 * pair + 1];
#pragma unroll
  for (int kv = 0; kv < 2; ++kv) {
    char* base = kv == 0 ? ptrs.k[layer] : ptrs.v[layer];
    const int4* s = reinterpret_cast<const int4*>(base) + src * vecs_per_block;
    int4* d = reinterpret_cast<int4*>(base) + dst * vecs_per_block;
    int64_t i = threadIdx.x;
    for (; i + 3 * (int64_t)blockDim.x < vecs_per_block; i += 4 * (int64_t)blockDim.x) {
      int4 a = ld_nc_v4(s + i), b = ld_nc_v4(s + i + blockDim.x), c = ld_nc_v4(s + i + 2 * blockDim.x),
           e = ld_nc_v4(s + i + 3 * blockDim.x);
      st_na_v4(d + i, a);
      st_na_v4(d + i + blockDim.x, b);
      st_na_v4(d + i + 2 * blockDim.x, c);
      st_na_v4(d + i + 3 * blockDim.x, e);
    }
    for (; i < vecs_per_block; i += blockDim.x) st_na_v4(d + i, ld_nc_v4(s + i));
  }
}

__global__ void copy_blocks_byte_kernel(const __grid_constant__ LayerPtrs ptrs,
                                        const int64_t* __restrict__ block_mapping, int64_t block_bytes) {
  const int layer = blockIdx.x;
  const int64_t pair = blockIdx.y;
  const int64_t src = block_mapping[2 * pair];
  const int64_t dst = block_mapping[2 * pair + 1];
  for (int kv = 0; kv < 2; ++kv) {
    char* base = kv == 0 ? ptrs.k[layer] : ptrs.v[layer];
    const char* s = base + src * block_bytes;
    char* d = base + dst * block_bytes;
    for (int64_t i = threadIdx.x; i < block_bytes; i += blockDim.x) d[i] = s[i];
  }
}

// =============================================================================================
// swap_blocks: gather/scatter kernel over device-visible memory (HBM or mapped pinned host memory)
// =============================================================================================

constexpr int kSwapPairsPerLaunch = 1024;  // 16 KiB of kernel parameters (large-parameter launches, CUDA >= 12.1)

struct SwapPairs {
  int64_t src[kSwapPairsPerLaunch];
  int64_t dst[kSwapPairsPerLaunch];
};

// grid (pairs, slices): every CTA copies one slice of one block with 128-bit accesses, 4 in flight per
// thread.  Over PCIe this keeps thousands of reads/writes outstanding in ONE launch instead of one
// cudaMemcpyAsync (~4 us of driver time each) per 32 KiB block.
__global__ void __launch_bounds__(256)
    swap_blocks_kernel(const __grid_constant__ SwapPairs pairs, const char* __restrict__ src, char* __restrict__ dst,
                       int64_t vecs_per_block, int64_t vecs_per_slice) {
  const int64_t s = pairs.src[blockIdx.x], d = pairs.dst[blockIdx.x];
  const int4* sp = reinterpret_cast<const int4*>(src) + s * vecs_per_block;
  int4* dp = reinterpret_cast<int4*>(dst) + d * vecs_per_block;
  const int64_t begin = blockIdx.y * vecs_per_slice;
  const int64_t end = min(begin + vecs_per_slice, vecs_per_block);
  int64_t i = begin + threadIdx.x;
  for (; i + 3 * 256 < end; i += 4 * 256) {
    const int4 a = sp[i], b = sp[i + 256], c = sp[i + 512], e = sp[i + 768];
    dp[i] = a;
    dp[i + 256] = b;
    dp[i + 512] = c;
    dp[i + 768] = e;
  }
  for (; i < end; i += 256) dp[i] = sp[i];
}

// Device-visible alias of a pointer: device memory as is; pinned (page-locked, mapped) host memory through
// its device pointer; pageable host memory -> nullptr (caller falls back to cudaMemcpyAsync).
static const void* device_visible(const void* p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    (void)cudaGetLastError();
    return nullptr;
  }
  if (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged) return attr.devicePointer;
  if (attr.type == cudaMemoryTypeHost) return attr.devicePointer;  // null when not mapped
  return nullptr;
}

// =============================================================================================
// convert_fp8
// =============================================================================================

// Vector bodies: one thread converts 16 elements per step — one 128-bit fp8 vector against
// 16 * sizeof(T) bytes of T (2 or 4 128-bit accesses), streaming loads/stores; the scalar loop takes the
// < 16-element tail and misaligned buffers.  Element arithmetic is fp8_quant / fp8_dequant unchanged.
template <typename T, int KV>
__global__ void fp8_quant_kernel(uint8_t* __restrict__ dst, const T* __restrict__ src, int64_t n, float scale,
                                 int vec_ok) {
  constexpr int NV = 16 * sizeof(T) / 16;  // int4 loads per 16 elements
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int64_t nvec = vec_ok ? n / 16 : 0;
  for (int64_t i = tid; i < nvec; i += nthreads) {
    union { int4 v[NV]; T e[16]; } in;
#pragma unroll
    for (int j = 0; j < NV; ++j) in.v[j] = ld_nc_v4(src + i * 16 + j * (16 / (int)sizeof(T)));
    union { int4 v; uint8_t e[16]; } o;
#pragma unroll
    for (int j = 0; j < 16; ++j) o.e[j] = fp8_quant<KV, T>(in.e[j], scale);
    st_na_v4(dst + i * 16, o.v);
  }
  for (int64_t i = nvec * 16 + tid; i < n; i += nthreads) dst[i] = fp8_quant<KV, T>(src[i], scale);
}
template <typename T, int KV>
__global__ void fp8_dequant_kernel(T* __restrict__ dst, const uint8_t* __restrict__ src, int64_t n, float scale,
                                   int vec_ok) {
  constexpr int NV = 16 * sizeof(T) / 16;
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
  const int64_t nvec = vec_ok ? n / 16 : 0;
  for (int64_t i = tid; i < nvec; i += nthreads) {
    union { int4 v; uint8_t e[16]; } in;
    in.v = ld_nc_v4(src + i * 16);
    union { int4 v[NV]; T e[16]; } o;
#pragma unroll
    for (int j = 0; j < 16; ++j) o.e[j] = fp8_dequant<KV, T>(in.e[j], scale);
#pragma unroll
    for (int j = 0; j < NV; ++j) st_na_v4(dst + i * 16 + j * (16 / (int)sizeof(T)), o.v[j]);
  }
  for (int64_t i = nvec * 16 + tid; i < n; i += nthreads) dst[i] = fp8_dequant<KV, T>(src[i], scale);
}

template <int KV>
static int launch_convert(void* dst, const void* src, int64_t n, int dst_dtype, int src_dtype, float scale,
                          cudaStream_t stream) {
  const int threads = 256;
  const int vec_ok = (reinterpret_cast<uintptr_t>(dst) % 16 == 0 && reinterpret_cast<uintptr_t>(src) % 16 == 0) ? 1 : 0;
  const int64_t items = vec_ok ? (n + 15) / 16 : n;
  const int64_t want = (items + threads - 1) / threads;
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(want, 148 * 16));
  if (dst_dtype == B200PA_U8) {
    uint8_t* d = static_cast<uint8_t*>(dst);
    switch (src_dtype) {
      case B200PA_F16: fp8_quant_kernel<__half, KV><<<blocks, threads, 0, stream>>>(d, (const __half*)src, n, scale, vec_ok); break;
      case B200PA_BF16: fp8_quant_kernel<__nv_bfloat16, KV><<<blocks, threads, 0, stream>>>(d, (const __nv_bfloat16*)src, n, scale, vec_ok); break;
      case B200PA_F32: fp8_quant_kernel<float, KV><<<blocks, threads, 0, stream>>>(d, (const float*)src, n, scale, vec_ok); break;
      default: return B200PA_ERR_INPUT_DTYPE;
    }
  } else if (src_dtype == B200PA_U8) {
    const uint8_t* s = static_cast<const uint8_t*>(src);
    switch (dst_dtype) {
      case B200PA_F16: fp8_dequant_kernel<__half, KV><<<blocks, threads, 0, stream>>>((__half*)dst, s, n, scale, vec_ok); break;
      case B200PA_BF16: fp8_dequant_kernel<__nv_bfloat16, KV><<<blocks, threads, 0, stream>>>((__nv_bfloat16*)dst, s, n, scale, vec_ok); break;
      case B200PA_F32: fp8_dequant_kernel<float, KV><<<blocks, threads, 0, stream>>>((float*)dst, s, n, scale, vec_ok); break;
      default: return B200PA_ERR_INPUT_DTYPE;
    }
  } else {
    return B200PA_ERR_INPUT_DTYPE;
  }
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

}  // namespace b200pa

// =============================================================================================
// C-ABI
// =============================================================================================
using namespace b200pa;

extern "C" int b200pa_reshape_and_cache(const b200pa_reshape_params_t* p, b200pa_stream_t stream) {
  if (!p) return B200PA_ERR_INVALID_ARG;
  if (p->num_tokens < 0 || p->num_heads <= 0 || p->head_size <= 0 || p->block_size <= 0 || p->x <= 0 ||
      p->head_size % p->x != 0)
    return B200PA_ERR_INVALID_ARG;
  if (p->num_tokens > 0 && (!p->key || !p->value || !p->key_cache || !p->value_cache || !p->slot_mapping))
    return B200PA_ERR_INVALID_ARG;
  switch (p->kv_dtype) {
    case B200PA_KV_AUTO: return dispatch_reshape_dtype<B200PA_KV_AUTO>(p, stream);
    case B200PA_KV_FP8_E4M3: return dispatch_reshape_dtype<B200PA_KV_FP8_E4M3>(p, stream);
    case B200PA_KV_FP8_E5M2: return dispatch_reshape_dtype<B200PA_KV_FP8_E5M2>(p, stream);
    default: return B200PA_ERR_KV_DTYPE;
  }
}

extern "C" int b200pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int32_t num_layers,
                                  const int64_t* block_mapping, int64_t num_pairs, int64_t block_bytes,
                                  b200pa_stream_t stream) {
  if (num_layers < 0 || num_pairs < 0 || block_bytes < 0) return B200PA_ERR_INVALID_ARG;
  if (num_layers == 0 || num_pairs == 0 || block_bytes == 0) return B200PA_OK;
  if (!key_cache_ptrs || !value_cache_ptrs || !block_mapping) return B200PA_ERR_INVALID_ARG;
  bool aligned = block_bytes % 16 == 0;
  for (int l = 0; l < num_layers && aligned; ++l)
    aligned = reinterpret_cast<uintptr_t>(key_cache_ptrs[l]) % 16 == 0 &&
              reinterpret_cast<uintptr_t>(value_cache_ptrs[l]) % 16 == 0;
  for (int l0 = 0; l0 < num_layers; l0 += kMaxLayersPerLaunch) {
    const int nl = std::min(kMaxLayersPerLaunch, num_layers - l0);
    LayerPtrs ptrs;
    std::memset(&ptrs, 0, sizeof(ptrs));
    for (int l = 0; l < nl; ++l) {
      ptrs.k[l] = static_cast<char*>(key_cache_ptrs[l0 + l]);
      ptrs.v[l] = static_cast<char*>(value_cache_ptrs[l0 + l]);
    }
    // gridDim.y is limited to 65535
    for (int64_t p0 = 0; p0 < num_pairs; p0 += 65535) {
      const int np = (int)std::min<int64_t>(65535, num_pairs - p0);
      dim3 grid(nl, np);
      if (aligned) {
        const int64_t vecs = block_bytes / 16;
        const int threads = (int)std::min<int64_t>(256, std::max<int64_t>(32, (vecs + 3) / 4 / 32 * 32));
        copy_blocks_vec_kernel<<<grid, threads, 0, stream>>>(ptrs, block_mapping + 2 * p0, vecs);
      } else {
        copy_blocks_byte_kernel<<<grid, 256, 0, stream>>>(ptrs, block_mapping + 2 * p0, block_bytes);
      }
      B200PA_CUDA_CHECK(cudaGetLastError());
    }
  }
  return B200PA_OK;
}

extern "C" int b200pa_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int64_t num_pairs,
                                  int64_t block_bytes, int32_t kind, b200pa_stream_t stream) {
  if (num_pairs < 0 || block_bytes < 0) return B200PA_ERR_INVALID_ARG;
  if (num_pairs == 0 || block_bytes == 0) return B200PA_OK;
  if (!src || !dst || !block_mapping_host) return B200PA_ERR_INVALID_ARG;
  cudaMemcpyKind k;
  switch (kind) {
    case B200PA_D2D: k = cudaMemcpyDeviceToDevice; break;
    case B200PA_D2H: k = cudaMemcpyDeviceToHost; break;
    case B200PA_H2D: k = cudaMemcpyHostToDevice; break;
    default: return B200PA_ERR_DEVICE_COMBINATION;
  }
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  // Fast path: both sides visible to the device and 16-byte granular -> one gather kernel per 1024 pairs, the
  // (host-resident) mapping travels as kernel parameters.  Destination blocks must be distinct (they are: the
  // scheduler allocates them), as with any parallel copy.
  static const bool no_kernel = std::getenv("B200PA_SWAP_MEMCPY") != nullptr;
  if (!no_kernel && block_bytes % 16 == 0) {
    const char* sv = static_cast<const char*>(device_visible(src));
    char* dv = const_cast<char*>(static_cast<const char*>(device_visible(dst)));
    if (sv && dv && reinterpret_cast<uintptr_t>(sv) % 16 == 0 && reinterpret_cast<uintptr_t>(dv) % 16 == 0) {
      const int64_t vecs = block_bytes / 16;
      // >= 2 slices per block keeps enough CTAs in flight for small mappings; 1024 vectors (16 KiB) per slice
      const int64_t per_slice = std::max<int64_t>(256, std::min<int64_t>(1024, (vecs + 1) / 2));
      const int slices = (int)((vecs + per_slice - 1) / per_slice);
      for (int64_t p0 = 0; p0 < num_pairs; p0 += kSwapPairsPerLaunch) {
        const int np = (int)std::min<int64_t>(kSwapPairsPerLaunch, num_pairs - p0);
        SwapPairs pk;
        for (int i = 0; i < np; ++i) {
          pk.src[i] = block_mapping_host[2 * (p0 + i)];
          pk.dst[i] = block_mapping_host[2 * (p0 + i) + 1];
        }
        swap_blocks_kernel<<<dim3(np, slices), 256, 0, stream>>>(pk, sv, dv, vecs, per_slice);
        B200PA_CUDA_CHECK(cudaGetLastError());
      }
      return B200PA_OK;
    }
  }
  // Fallback (pageable host memory / odd sizes): cudaMemcpyAsync per run of consecutive (src+1 -> dst+1)
  // pairs — the scheduler hands out mostly-contiguous ranges, and one large transfer beats many small ones.
  int64_t i = 0;
  while (i < num_pairs) {
    const int64_t s0 = block_mapping_host[2 * i], d0 = block_mapping_host[2 * i + 1];
    int64_t run = 1;
    while (i + run < num_pairs && block_mapping_host[2 * (i + run)] == s0 + run &&
           block_mapping_host[2 * (i + run) + 1] == d0 + run)
      ++run;
    B200PA_CUDA_CHECK(cudaMemcpyAsync(d + d0 * block_bytes, s + s0 * block_bytes, (size_t)(run * block_bytes), k, stream));
    i += run;
  }
  return B200PA_OK;
}

extern "C" int b200pa_convert_fp8(void* dst, const void* src, int64_t numel, int32_t dst_dtype, int32_t src_dtype,
                                  float scale, int32_t kv_dtype, b200pa_stream_t stream) {
  if (numel < 0) return B200PA_ERR_INVALID_ARG;
  if (numel == 0) return B200PA_OK;
  if (!dst || !src) return B200PA_ERR_INVALID_ARG;
  switch (kv_dtype) {
    case B200PA_KV_FP8_E4M3: return launch_convert<B200PA_KV_FP8_E4M3>(dst, src, numel, dst_dtype, src_dtype, scale, stream);
    case B200PA_KV_FP8_E5M2: return launch_convert<B200PA_KV_FP8_E5M2>(dst, src, numel, dst_dtype, src_dtype, scale, stream);
    default: return B200PA_ERR_KV_DTYPE;  // "auto" is not convertible (reference: assert(false) path)
  }
}

// ---- measurement aid: read-only HBM stream (bench.py's read-stream roofline denominator) ------------
// MEASURED_PEAKS.json's hbm_gbs is a read+write copy; a KV decode is a pure read stream, which HBM3e
// serves faster (no read/write turnarounds).  This kernel reads `bytes` once with 128-bit streaming
// loads, 8 in flight per thread, and folds them into one word per CTA so nothing is optimised away.
namespace b200pa {
__global__ void __launch_bounds__(256) read_stream_probe_kernel(const int4* __restrict__ src, int64_t n_vec,
                                                                 unsigned* __restrict__ sink) {
  constexpr int U = 8;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned acc = 0;
  for (; i + (U - 1) * stride < n_vec; i += U * stride) {
    int4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld_nc_v4(src + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  for (; i < n_vec; i += stride) {
    const int4 v = ld_nc_v4(src + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc == 0x9e3779b9u) atomicAdd(sink, 1u);  // practically never; keeps the loads live
}
}  // namespace b200pa

extern "C" int b200pa_read_stream_probe(const void* src, int64_t bytes, void* sink_u32, b200pa_stream_t stream) {
  if (!src || !sink_u32 || bytes < 16 || reinterpret_cast<uintptr_t>(src) % 16) return B200PA_ERR_INVALID_ARG;
  int dev = 0, sms = 0;
  B200PA_CUDA_CHECK(cudaGetDevice(&dev));
  B200PA_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  b200pa::read_stream_probe_kernel<<<sms * 8, 256, 0, stream>>>(static_cast<const int4*>(src), bytes / 16,
                                                                static_cast<unsigned*>(sink_u32));
  B200PA_CUDA_CHECK(cudaGetLastError());
  return B200PA_OK;
}

extern "C" int b200pa_get_max_shared_memory_per_block_device_attribute(int32_t device_id) {
  int v = 0;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device_id) != cudaSuccess) return -1;
  return v;
}

// csrc/cuda_utils_kernels.cu:5-15 (device_id < 0 = current device)
extern "C" int b200pa_get_device_attribute(int32_t attribute, int32_t device_id) {
  int device = device_id, v = 0;
  if (device_id < 0 && cudaGetDevice(&device) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&v, static_cast<cudaDeviceAttr>(attribute), device) != cudaSuccess) return -1;
  return v;
}

extern "C" int b200pa_abi_version(void) { return B200PA_ABI_VERSION; }
extern "C" int b200pa_last_cuda_error(void) { return g_last_cuda_error; }
extern "C" const char* b200pa_last_attention_path(void) { return g_last_attention_path; }
extern "C" int b200pa_last_attention_partitions(void) { return g_last_attention_partitions; }

extern "C" const char* b200pa_status_string(int status) {
  switch (status) {
    case B200PA_OK: return "ok";
    case B200PA_ERR_HEAD_SIZE: return "Unsupported head size: ";
    case B200PA_ERR_BLOCK_SIZE: return "Unsupported block size: ";
    case B200PA_ERR_KV_DTYPE: return "Unsupported data type of kv cache: ";
    case B200PA_ERR_INPUT_DTYPE: return "Unsupported input type of kv cache: ";
    case B200PA_ERR_INVALID_ARG: return "invalid argument";
    case B200PA_ERR_CUDA: return "CUDA error";
    case B200PA_ERR_DEVICE_COMBINATION: return "Invalid device combination";
    case B200PA_ERR_SCRATCH_LIMIT:
      return "problem exceeds the static scratch of b200pa_paged_attention_static_scratch (limits in include/b200pa.h)";
    default: return "unknown status";
  }
}
