/*
 * TEST INFRASTRUCTURE — CPU oracle for the paged-attention hot path.  NOT a product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library (the product, vllm_legacy_b200, has no CPU path and never links it).
 *
 * Plain-C restatement of the *algorithm* of the reference's CUDA implementation
 * (DeepAuto-AI/vllm-legacy).  Every function cites the reference lines it follows.
 * Parity pinning (see DESIGN.md "Oracle"): this file is checked by tests/test_oracle.py against
 *   - the reference's own CPU extension compiled unchanged (oracle/_ref/ref_cpu*.so) and fixtures
 *     generated with it (tests/golden/, tests/golden/make_golden.py),
 *   - CUDA 12.9's cuda_fp8.h host conversions (oracle/fp8_pin.cpp) and torch.float8 casts,
 * and on the GPU box against the reference's CUDA kernels built for sm_100a (oracle/_ref/ref_cuda.so).
 *
 * Element types are passed as codes: 0 = fp16, 1 = bf16, 2 = fp32, 3 = uint8 (fp8 storage).
 * kv codes: 0 = auto (cache holds the model dtype), 1 = fp8 e4m3, 2 = fp8 e5m2.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { DT_F16 = 0, DT_BF16 = 1, DT_F32 = 2, DT_U8 = 3 };
enum { KV_AUTO = 0, KV_E4M3 = 1, KV_E5M2 = 2 };

/* ---------------------------------------------------------------------------------------------
 * scalar conversions (IEEE, round-to-nearest-even) — what cvt.rn.f16.f32 / __float2bfloat16_rn do
 * --------------------------------------------------------------------------------------------- */
static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float oracle_half_to_float(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
  if (exp == 0) {
    if (man == 0) return u2f(sign);
    float v = (float)man * 5.9604644775390625e-8f; /* 2^-24 */
    return (sign ? -v : v);
  }
  if (exp == 31) return u2f(sign | 0x7f800000u | (man << 13));
  return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t oracle_float_to_half(float f) {
  uint32_t x = f2u(f);
  uint32_t sign = (x >> 16) & 0x8000u;
  uint32_t ax = x & 0x7fffffffu;
  if (ax > 0x7f800000u) return (uint16_t)(sign | 0x7fffu);                 /* NaN */
  if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);               /* >= 65520 -> inf */
  if (ax < 0x33000001u) return (uint16_t)sign;                            /* <= 2^-25 -> 0 */
  int e = (int)(ax >> 23) - 127;
  uint32_t man = (ax & 0x7fffffu) | 0x800000u;
  int shift = (e < -14) ? (13 + (-14 - e)) : 13; /* bits dropped */
  uint32_t kept = man >> shift;
  uint32_t rem = man & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (kept & 1u))) kept++;
  uint32_t out;
  if (e < -14) out = kept;                                  /* subnormal (may carry into normal) */
  else out = ((uint32_t)(e + 15) << 10) + (kept - 0x400u);  /* carry propagates into exponent */
  return (uint16_t)(sign | out);
}

float oracle_bf16_to_float(uint16_t b) { return u2f((uint32_t)b << 16); }

uint16_t oracle_float_to_bf16(float f) {
  uint32_t x = f2u(f);
  if ((x & 0x7fffffffu) > 0x7f800000u) return 0x7fffu; /* __float2bfloat16_rn: canonical NaN */
  uint32_t lsb = (x >> 16) & 1u;
  x += 0x7fffu + lsb;
  return (uint16_t)(x >> 16);
}

static float load_elem(const void* p, int64_t i, int dt) {
  switch (dt) {
    case DT_F16: return oracle_half_to_float(((const uint16_t*)p)[i]);
    case DT_BF16: return oracle_bf16_to_float(((const uint16_t*)p)[i]);
    default: return ((const float*)p)[i];
  }
}
static void store_elem(void* p, int64_t i, int dt, float v) {
  switch (dt) {
    case DT_F16: ((uint16_t*)p)[i] = oracle_float_to_half(v); break;
    case DT_BF16: ((uint16_t*)p)[i] = oracle_float_to_bf16(v); break;
    default: ((float*)p)[i] = v; break;
  }
}
static float round_to(float v, int dt) {
  if (dt == DT_F16) return oracle_half_to_float(oracle_float_to_half(v));
  if (dt == DT_BF16) return oracle_bf16_to_float(oracle_float_to_bf16(v));
  return v;
}
static int elem_size(int dt) { return dt == DT_F32 ? 4 : (dt == DT_U8 ? 1 : 2); }

/* ---------------------------------------------------------------------------------------------
 * fp8 (third-party arithmetic: CUDA 12.9 cuda_fp8.h, __nv_cvt_float_to_fp8 with __NV_SATFINITE and
 * __nv_cvt_fp8_to_halfraw; call sites quant_utils.cuh:298,311,355,410,461,474,486).
 * Restated from the format definition: RN-even to {e4m3: 4 exp bits bias 7, 3 mantissa bits, no
 * inf, max 448; e5m2: 5 exp bits bias 15, 2 mantissa bits, max finite 57344}, saturate to max finite.
 * --------------------------------------------------------------------------------------------- */
uint8_t oracle_float_to_fp8(float f, int kv) {
  const int mbits = (kv == KV_E5M2) ? 2 : 3;
  const int bias = (kv == KV_E5M2) ? 15 : 7;
  const uint8_t maxnorm = (kv == KV_E5M2) ? 0x7b : 0x7e;
  const float maxval = (kv == KV_E5M2) ? 57344.0f : 448.0f;
  uint32_t x = f2u(f);
  uint8_t sign = (uint8_t)((x >> 24) & 0x80u);
  uint32_t ax = x & 0x7fffffffu;
  if (ax > 0x7f800000u) return 0x7f; /* any NaN -> canonical +NaN first (cuda_fp8.hpp:271-275) -> 0x7f in both formats */
  float a = u2f(ax);
  /* overflow threshold = maxnorm + half ulp (ties go to even = maxnorm's neighbour is out of range) */
  const float half_ulp_at_max = (kv == KV_E5M2) ? 4096.0f : 16.0f;
  if (a > maxval + half_ulp_at_max || (a == maxval + half_ulp_at_max)) {
    /* cuda: absx > OVERFLOW_THRESHOLD saturates; e4m3 threshold is exactly max+half ulp (tie -> rounds
       to even = 448 since mantissa 0b110 is even), e5m2 threshold is (max+half ulp) - 1 double ulp, i.e.
       the tie itself saturates too.  Both end at maxnorm. */
    return (uint8_t)(sign | maxnorm);
  }
  if (a == 0.0f) return sign ? 0x80 : 0x00;
  int e;
  float m = frexpf(a, &e); /* a = m * 2^e, m in [0.5, 1) */
  int exp_unb = e - 1;     /* a = (2m) * 2^(e-1), 2m in [1, 2) */
  int min_exp = 1 - bias;  /* smallest normal exponent */
  float scaled;            /* value in units of the target's ulp at this exponent */
  int target_exp = exp_unb < min_exp ? min_exp : exp_unb;
  scaled = ldexpf(a, mbits - target_exp); /* exact: power-of-two scaling */
  /* round to nearest even integer */
  float r = nearbyintf(scaled); /* default rounding mode = RN-even */
  uint32_t q = (uint32_t)r;
  if (q == 0) return sign; /* underflow keeps the sign (pinned against cuda_fp8.h: -tiny -> 0x80) */
  uint8_t out;
  if (exp_unb < min_exp) {
    out = (uint8_t)q; /* subnormal; q == 2^mbits carries into the first normal, same encoding */
  } else {
    /* q in [2^mbits, 2^(mbits+1)] */
    out = (uint8_t)(((uint32_t)(target_exp + bias) << mbits) + (q - (1u << mbits)));
  }
  if ((out & 0x7f) > maxnorm) out = maxnorm;
  return (uint8_t)(sign | out);
}

/* fp8 -> half bits, exact (__nv_cvt_fp8_to_halfraw) */
uint16_t oracle_fp8_to_half(uint8_t b, int kv) {
  if (kv == KV_E5M2) return (uint16_t)((uint16_t)b << 8);
  uint16_t sign = (uint16_t)((b & 0x80u) << 8);
  uint32_t exp = (b >> 3) & 0xfu, man = b & 0x7u;
  if (exp == 0xf && man == 0x7) return (uint16_t)(sign | 0x7fffu); /* NaN */
  if (exp == 0) {
    if (man == 0) return sign;
    /* subnormal: man * 2^-9 ; normalise into half */
    int sh = 0;
    while (!(man & 0x8u)) { man <<= 1; sh++; }
    /* value = (man/8) * 2^(-6 - sh + ... ) ; man now in [8,16): 1.xxx * 2^(-6 - sh) with sh>=1 */
    uint32_t hexp = (uint32_t)(15 - 6 - sh);
    return (uint16_t)(sign | (hexp << 10) | ((man & 0x7u) << 7));
  }
  return (uint16_t)(sign | ((exp + 8u) << 10) | (man << 7));
}

/* quant_utils.cuh:456-488: fp8 = cvt(float(x) / scale) ;  :293-300, 349-359, 406-415: x = T(float(half(fp8)) * scale) */
static uint8_t quant_fp8(float v, float scale, int kv) { return oracle_float_to_fp8(v / scale, kv); }
static float dequant_fp8(uint8_t b, float scale, int kv, int dt) {
  return round_to(oracle_half_to_float(oracle_fp8_to_half(b, kv)) * scale, dt);
}

/* csrc/cache_kernels.cu:347-420 (convert_fp8_kernel + host dispatch) */
int oracle_convert_fp8(void* dst, const void* src, int64_t n, int dst_dt, int src_dt, float scale, int kv) {
  if (kv != KV_E4M3 && kv != KV_E5M2) return -3;
  if (dst_dt == DT_U8) {
    for (int64_t i = 0; i < n; ++i) ((uint8_t*)dst)[i] = quant_fp8(load_elem(src, i, src_dt), scale, kv);
  } else if (src_dt == DT_U8) {
    for (int64_t i = 0; i < n; ++i) store_elem(dst, i, dst_dt, dequant_fp8(((const uint8_t*)src)[i], scale, kv, dst_dt));
  } else {
    return -4;
  }
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * reshape_and_cache — csrc/cache_kernels.cu:150-202 (index math :169-200)
 * --------------------------------------------------------------------------------------------- */
int oracle_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                             const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                             int64_t value_stride, int num_heads, int head_size, int block_size, int x, int dt,
                             int kv, float kv_scale) {
  const int64_t n = (int64_t)num_heads * head_size;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue; /* padding token, :161-164 */
    const int64_t b = slot / block_size, o = slot % block_size;
    for (int64_t i = 0; i < n; ++i) {
      const int h = (int)(i / head_size), d = (int)(i % head_size);
      const int64_t kdst = b * num_heads * (head_size / x) * block_size * x + (int64_t)h * (head_size / x) * block_size * x +
                           (int64_t)(d / x) * block_size * x + o * x + (d % x);
      const int64_t vdst = b * num_heads * head_size * block_size + (int64_t)h * head_size * block_size +
                           (int64_t)d * block_size + o;
      if (kv == KV_AUTO) {
        const int es = elem_size(dt);
        memcpy((char*)key_cache + kdst * es, (const char*)key + (t * key_stride + i) * es, es);
        memcpy((char*)value_cache + vdst * es, (const char*)value + (t * value_stride + i) * es, es);
      } else {
        ((uint8_t*)key_cache)[kdst] = quant_fp8(load_elem(key, t * key_stride + i, dt), kv_scale, kv);
        ((uint8_t*)value_cache)[vdst] = quant_fp8(load_elem(value, t * value_stride + i, dt), kv_scale, kv);
      }
    }
  }
  return 0;
}

/* reshape_and_cache_flash — csrc/cache_kernels.cu:204-235: cache [nb, block_size, H, D] with block stride
 * `block_stride` elements; tgt = block_idx*block_stride + block_offset*H*D + head_idx*D + head_offset (:226-230) */
int oracle_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                   const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                                   int64_t value_stride, int64_t block_stride, int num_heads, int head_size,
                                   int block_size, int dt) {
  const int64_t n = (int64_t)num_heads * head_size;
  const int es = elem_size(dt);
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue; /* :218-221 */
    const int64_t b = slot / block_size, o = slot % block_size;
    for (int64_t i = 0; i < n; ++i) {
      const int64_t h = i / head_size, d = i % head_size;
      const int64_t tgt = b * block_stride + o * num_heads * head_size + h * head_size + d;
      memcpy((char*)k_cache + tgt * es, (const char*)key + (t * key_stride + i) * es, es);
      memcpy((char*)v_cache + tgt * es, (const char*)value + (t * value_stride + i) * es, es);
    }
  }
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * rotary_embedding / batched_rotary_embedding — csrc/pos_encoding_kernels.cu:10-131.
 * apply_token_rotary_embedding (:10-36): NeoX pairs (i, embed+i) with cos[i], sin[i]; GPT-J pairs
 * (2i, 2i+1) with cos[i], sin[i]; arr[x] = x*cos - y*sin; arr[y] = y*cos + x*sin evaluated with
 * scalar_t operators: for c10::Half / c10::BFloat16 every product and sum is computed in fp32 and
 * rounded back to the 16-bit type (torch/headeronly/util/Half.h:501-511, BFloat16.h:185-198); for
 * float nvcc contracts to fma(x, cos, -(y*sin)) and fma(x, sin, y*cos) (SASS of the reference
 * build: FMUL, FMUL, FFMA, FFMA).  offsets == NULL: rotary_embedding (:70-93), else the batched
 * form (:95-131) with cache row pos + offset.
 * --------------------------------------------------------------------------------------------- */
static void rotate_pair(void* arr, int64_t xi, int64_t yi, float c, float s, int dt) {
  const float x = load_elem(arr, xi, dt), y = load_elem(arr, yi, dt);
  float nx, ny;
  if (dt == DT_F32) {
    nx = fmaf(x, c, -(y * s));
    ny = fmaf(x, s, y * c);
  } else {
    nx = round_to(x * c, dt) - round_to(y * s, dt);
    ny = round_to(y * c, dt) + round_to(x * s, dt);
  }
  store_elem(arr, xi, dt, nx);
  store_elem(arr, yi, dt, ny);
}

int oracle_rotary_embedding(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                            const int64_t* offsets, int64_t num_tokens, int64_t query_stride, int64_t key_stride,
                            int num_heads, int num_kv_heads, int head_size, int rot_dim, int is_neox, int dt) {
  const int embed = rot_dim / 2;
  for (int64_t t = 0; t < num_tokens; ++t) {
    const int64_t pos = positions[t] + (offsets ? offsets[t] : 0);
    const int64_t cbase = pos * rot_dim;
    for (int pass = 0; pass < 2; ++pass) {
      void* base = pass == 0 ? query : key;
      const int nh = pass == 0 ? num_heads : num_kv_heads;
      const int64_t stride = pass == 0 ? query_stride : key_stride;
      for (int h = 0; h < nh; ++h) {
        const int64_t th = t * stride + (int64_t)h * head_size;
        for (int i = 0; i < embed; ++i) {
          const float c = load_elem(cos_sin_cache, cbase + i, dt);
          const float s = load_elem(cos_sin_cache, cbase + embed + i, dt);
          if (is_neox) rotate_pair(base, th + i, th + embed + i, c, s, dt);
          else rotate_pair(base, th + 2 * i, th + 2 * i + 1, c, s, dt);
        }
      }
    }
  }
  return 0;
}

/* copy_blocks — csrc/cache_kernels.cu:68-95: K_l[dst] = K_l[src], V_l[dst] = V_l[src] for every layer */
int oracle_copy_blocks(void* const* key_caches, void* const* value_caches, int num_layers,
                       const int64_t* block_mapping, int64_t num_pairs, int64_t block_bytes) {
  for (int l = 0; l < num_layers; ++l)
    for (int64_t p = 0; p < num_pairs; ++p) {
      const int64_t s = block_mapping[2 * p], d = block_mapping[2 * p + 1];
      memmove((char*)key_caches[l] + d * block_bytes, (char*)key_caches[l] + s * block_bytes, (size_t)block_bytes);
      memmove((char*)value_caches[l] + d * block_bytes, (char*)value_caches[l] + s * block_bytes, (size_t)block_bytes);
    }
  return 0;
}

/* swap_blocks — csrc/cache_kernels.cu:47-63: dst[d] = src[s] per pair */
int oracle_swap_blocks(const void* src, void* dst, const int64_t* block_mapping, int64_t num_pairs,
                       int64_t block_bytes) {
  for (int64_t p = 0; p < num_pairs; ++p)
    memcpy((char*)dst + block_mapping[2 * p + 1] * block_bytes, (const char*)src + block_mapping[2 * p] * block_bytes,
           (size_t)block_bytes);
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * paged attention — csrc/attention/attention_kernels.cu:86-496 (one (seq, head, partition) at a time)
 * --------------------------------------------------------------------------------------------- */
typedef struct {
  const void* q; const void* k_cache; const void* v_cache;
  const int32_t* block_tables; const int32_t* seq_lens; const float* alibi_slopes;
  int64_t q_stride, kv_block_stride, kv_head_stride;
  int num_seqs, num_heads, num_kv_heads, head_size, block_size, max_blocks_per_seq;
  int dt, kv;
  float scale, kv_scale;
  int tp_rank, bs_local_blocks, bs_vert_stride, bs_block_size, bs_head_sliding_step;
} attn_args_t;

static float cache_load(const void* cache, int64_t idx, const attn_args_t* a) {
  if (a->kv == KV_AUTO) return load_elem(cache, idx, a->dt);
  return dequant_fp8(((const uint8_t*)cache)[idx], a->kv_scale, a->kv, a->dt); /* :284-289, 413-419 */
}

/* One partition [t0, t1) of one (seq, head).  Writes the normalised partial output (rounded to T as
 * the kernel stores it, :484-495), returns max logit and exp sum (:349-357). */
static void attend_range(const attn_args_t* a, int seq, int head, int t0, int t1, float* logits /* [t1-t0] */,
                         float* out /* [D] */, float* max_logit, float* exp_sum) {
  const int D = a->head_size, B = a->block_size;
  const int x = 16 / (a->kv == KV_AUTO ? elem_size(a->dt) : 1);
  const int seq_len = a->seq_lens[seq];
  const int kv_head = head / (a->num_heads / a->num_kv_heads);
  const float alibi = a->alibi_slopes ? a->alibi_slopes[head] : 0.f;
  const int32_t* bt = a->block_tables + (int64_t)seq * a->max_blocks_per_seq;
  const int sparse = a->bs_vert_stride > 1;
  int bs_offset = 0, q_bs_block = 0;
  if (sparse) { /* :212-224 */
    q_bs_block = (seq_len - 1) / a->bs_block_size;
    if (a->bs_head_sliding_step >= 0) bs_offset = (a->tp_rank * a->num_heads + head) * a->bs_head_sliding_step + 1;
    else bs_offset = (a->tp_rank * a->num_kv_heads + kv_head) * (-a->bs_head_sliding_step) + 1;
  }
  float qmax = -3.402823466e+38f;
  for (int t = t0; t < t1; ++t) {
    const int blk = t / B, off = t % B;
    int skip = 0;
    if (sparse) { /* :236-241 */
      const int kb = blk * B / a->bs_block_size;
      const int is_remote = ((kb + bs_offset) % a->bs_vert_stride == 0);
      const int is_local = (kb > q_bs_block - a->bs_local_blocks);
      skip = !is_remote && !is_local;
    }
    if (skip) { logits[t - t0] = -3.402823466e+38f; continue; } /* :250 */
    const int64_t base = (int64_t)bt[blk] * a->kv_block_stride + (int64_t)kv_head * a->kv_head_stride;
    float qk = 0.f;
    for (int d = 0; d < D; ++d) { /* Qk_dot: fp32 products + fp32 accumulation, attention_utils.cuh:30-56 */
      const float qv = load_elem(a->q, (int64_t)seq * a->q_stride + (int64_t)head * D + d, a->dt);
      const float kvv = cache_load(a->k_cache, base + (int64_t)(d / x) * B * x + (int64_t)off * x + (d % x), a);
      qk += qv * kvv;
    }
    qk = a->scale * qk;
    if (alibi != 0.f) qk += alibi * (float)(t - seq_len + 1); /* :297 */
    logits[t - t0] = qk;
    if (qk > qmax) qmax = qk;
  }
  float sum = 0.f;
  for (int t = t0; t < t1; ++t) { /* :333-340 */
    const float v = expf(logits[t - t0] - qmax);
    logits[t - t0] = v;
    sum += v;
  }
  const float inv = 1.f / (sum + 1e-6f); /* :343 */
  for (int d = 0; d < D; ++d) out[d] = 0.f;
  for (int t = t0; t < t1; ++t) {
    const int blk = t / B, off = t % B;
    if (sparse) { /* :387-393 */
      const int vb = blk * B / a->bs_block_size;
      if (!((vb + bs_offset) % a->bs_vert_stride == 0) && !(vb > q_bs_block - a->bs_local_blocks)) continue;
    }
    /* P rounded to T before the product (from_float(logits_vec), :398-400); product in T precision then
       fp32 accumulate (dot(), attention_generic.cuh:41-43 + dtype_*.cuh mul/sum) */
    const float p = round_to(logits[t - t0] * inv, a->dt);
    const int64_t base = (int64_t)bt[blk] * a->kv_block_stride + (int64_t)kv_head * a->kv_head_stride;
    for (int d = 0; d < D; ++d) {
      const float vv = cache_load(a->v_cache, base + (int64_t)d * B + off, a);
      out[d] += round_to(p * vv, a->dt);
    }
  }
  for (int d = 0; d < D; ++d) out[d] = round_to(out[d], a->dt);
  *max_logit = qmax;
  *exp_sum = sum;
}

/* paged_attention_v1 — attention_kernels.cu:499-527, 805-825 */
int oracle_paged_attention_v1(void* out, const attn_args_t* a, int max_seq_len) {
  const int D = a->head_size;
  float* logits = (float*)malloc(sizeof(float) * (size_t)(max_seq_len > 0 ? max_seq_len : 1));
  float* o = (float*)malloc(sizeof(float) * (size_t)D);
  if (!logits || !o) return -1;
  for (int s = 0; s < a->num_seqs; ++s)
    for (int h = 0; h < a->num_heads; ++h) {
      float m, l;
      attend_range(a, s, h, 0, a->seq_lens[s], logits, o, &m, &l);
      for (int d = 0; d < D; ++d) store_elem(out, ((int64_t)s * a->num_heads + h) * D + d, a->dt, o[d]);
    }
  free(logits);
  free(o);
  return 0;
}

/* paged_attention_v2 + reduce — attention_kernels.cu:530-562, 565-669 (PARTITION_SIZE = 512, :846) */
int oracle_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out, const attn_args_t* a,
                              int max_seq_len) {
  const int D = a->head_size, PS = 512;
  const int max_parts = (max_seq_len + PS - 1) / PS;
  float logits[512];
  float* o = (float*)malloc(sizeof(float) * (size_t)D);
  if (!o) return -1;
  for (int s = 0; s < a->num_seqs; ++s)
    for (int h = 0; h < a->num_heads; ++h) {
      const int seq_len = a->seq_lens[s];
      const int parts = (seq_len + PS - 1) / PS;
      const int64_t row = (int64_t)s * a->num_heads + h;
      for (int p = 0; p < parts; ++p) {
        const int t0 = p * PS, t1 = (t0 + PS < seq_len) ? t0 + PS : seq_len;
        float m, l;
        attend_range(a, s, h, t0, t1, logits, o, &m, &l);
        for (int d = 0; d < D; ++d) store_elem(tmp_out, (row * max_parts + p) * D + d, a->dt, o[d]);
        max_logits[row * max_parts + p] = m;
        exp_sums[row * max_parts + p] = l;
      }
      if (parts == 1) { /* :582-594 */
        for (int d = 0; d < D; ++d)
          store_elem(out, row * D + d, a->dt, load_elem(tmp_out, (row * max_parts) * D + d, a->dt));
        continue;
      }
      float mx = -3.402823466e+38f; /* :606-625 */
      for (int p = 0; p < parts; ++p) if (max_logits[row * max_parts + p] > mx) mx = max_logits[row * max_parts + p];
      float gsum = 0.f; /* :637-652 */
      for (int p = 0; p < parts; ++p) gsum += exp_sums[row * max_parts + p] * expf(max_logits[row * max_parts + p] - mx);
      const float inv = 1.f / (gsum + 1e-6f);
      for (int d = 0; d < D; ++d) { /* :655-668 */
        float acc = 0.f;
        for (int p = 0; p < parts; ++p) {
          const float w = exp_sums[row * max_parts + p] * expf(max_logits[row * max_parts + p] - mx);
          acc += load_elem(tmp_out, (row * max_parts + p) * D + d, a->dt) * w * inv;
        }
        store_elem(out, row * D + d, a->dt, acc);
      }
    }
  free(o);
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * context_attention_fwd (prefix-prefill attention over the paged cache) — restates the reference's
 * Triton kernels row by row with their own tile order and rounding points:
 *   vllm/attention/ops/prefix_prefill.py:10-245  (_fwd_kernel: normalised accumulator, window -> -10000)
 *   vllm/attention/ops/prefix_prefill.py:442-681 (_fwd_kernel_alibi: un-normalised accumulator, no window)
 *   BLOCK_M = BLOCK_N = 128 (:698-699), sm_scale = 1/sqrt(D) (:706), p rounded to the value dtype before
 *   tl.dot (:170, :236, :596, :667), fp32 accumulation.
 * k_cache [nb, Hkv, D/8, B, 8], v_cache [nb, Hkv, D, B] of dtype dt; q/k/v/out [tokens, heads, D] contiguous.
 * --------------------------------------------------------------------------------------------- */
int oracle_context_attention_fwd(void* out, const void* q, const void* k, const void* v, const void* k_cache,
                                 const void* v_cache, const int32_t* b_loc, int64_t b_loc_stride,
                                 const int32_t* b_start_loc, const int32_t* b_seq_len, const int32_t* b_ctx_len,
                                 int batch, int num_heads, int num_kv_heads, int D, int block_size,
                                 const float* alibi_slopes, int sliding_window, int dt) {
  enum { BLK = 128 };
  if (dt != DT_F16 && dt != DT_BF16 && dt != DT_F32) return -1;
  const float sm_scale = 1.0f / sqrtf((float)D);
  const int qpk = num_heads / num_kv_heads;
  const int x = 8;
  float* kt = (float*)malloc(sizeof(float) * BLK * D);
  float* vt = (float*)malloc(sizeof(float) * BLK * D);
  float* acc = (float*)malloc(sizeof(float) * D);
  float* qrow = (float*)malloc(sizeof(float) * D);
  float qk[BLK], pr[BLK];
  const int W = alibi_slopes ? 0 : (sliding_window > 0 ? sliding_window : 0);
  for (int b = 0; b < batch; ++b) {
    const int ctx = b_ctx_len[b], qlen = b_seq_len[b] - ctx;
    const int64_t t0 = b_start_loc[b];
    for (int h = 0; h < num_heads; ++h) {
      const int kvh = h / qpk;
      for (int m = 0; m < qlen; ++m) {
        for (int d = 0; d < D; ++d) { qrow[d] = load_elem(q, ((t0 + m) * num_heads + h) * D + d, dt); acc[d] = 0.f; }
        float m_i = -INFINITY, l_i = 0.f;
        const int start_m = m / BLK;
        const int n_ctx_tiles = (ctx + BLK - 1) / BLK;
        const int n_new_tiles = start_m + 1; /* range(0, (start_m + 1) * BLOCK_M, BLOCK_N) (:193, :615) */
        for (int tile = 0; tile < n_ctx_tiles + n_new_tiles; ++tile) {
          const int is_ctx = tile < n_ctx_tiles;
          const int start_n = (is_ctx ? tile : tile - n_ctx_tiles) * BLK;
          /* gather the K / V tile as the masked tl.load would (other = 0) */
          for (int n = 0; n < BLK; ++n) {
            const int pos = start_n + n;
            const int valid = is_ctx ? (pos < ctx) : (pos < qlen);
            for (int d = 0; d < D; ++d) {
              float kv = 0.f, vv = 0.f;
              if (valid && is_ctx) {
                const int64_t blk = b_loc[b * b_loc_stride + pos / block_size];
                const int o = pos % block_size;
                kv = load_elem(k_cache, (((blk * num_kv_heads + kvh) * (D / x) + d / x) * block_size + o) * x + d % x, dt);
                vv = load_elem(v_cache, ((blk * num_kv_heads + kvh) * D + d) * block_size + o, dt);
              } else if (valid) {
                kv = load_elem(k, ((t0 + pos) * num_kv_heads + kvh) * D + d, dt);
                vv = load_elem(v, ((t0 + pos) * num_kv_heads + kvh) * D + d, dt);
              }
              kt[n * D + d] = kv;
              vt[n * D + d] = vv;
            }
          }
          float m_ij = -INFINITY;
          for (int n = 0; n < BLK; ++n) {
            const int pos = start_n + n;
            float s = 0.f;
            for (int d = 0; d < D; ++d) s += qrow[d] * kt[n * D + d];
            if (is_ctx) {
              s = (pos < ctx) ? s : -INFINITY;
              s *= sm_scale;
              if (W > 0 && !((ctx + m) - pos < W)) s = -10000.f;                      /* :141-143 */
            } else {
              s *= sm_scale;
              s = (m >= pos) ? s : -INFINITY;                                          /* :207-208 */
              if (W > 0 && !(m - pos < W)) s = -10000.f;                              /* :209-212 */
            }
            if (alibi_slopes) {                                                        /* :560-565, :633-638 */
              const float qpos = (float)(ctx + m), kpos = (float)((is_ctx ? 0 : ctx) + pos);
              float a = (kpos - qpos) * alibi_slopes[h];
              if (!(a <= 0.f)) a = -INFINITY;
              s += a;
            }
            qk[n] = s;
            if (s > m_ij) m_ij = s;
          }
          if (alibi_slopes) {
            const float m_new = fmaxf(m_i, m_ij);
            float l_ij = 0.f;
            for (int n = 0; n < BLK; ++n) { pr[n] = expf(qk[n] - m_new); l_ij += pr[n]; }
            const float alpha = expf(m_i - m_new);
            for (int d = 0; d < D; ++d) acc[d] *= alpha;
            for (int n = 0; n < BLK; ++n) {
              const float pn = round_to(pr[n], dt);
              if (pn != 0.f) for (int d = 0; d < D; ++d) acc[d] += pn * vt[n * D + d];
            }
            l_i = alpha * l_i + l_ij;
            m_i = m_new;
          } else {
            float l_ij = 0.f;
            for (int n = 0; n < BLK; ++n) { pr[n] = expf(qk[n] - m_ij); l_ij += pr[n]; }
            const float m_new = fmaxf(m_i, m_ij);
            const float alpha = expf(m_i - m_new), beta = expf(m_ij - m_new);
            const float l_new = alpha * l_i + beta * l_ij;
            const float p_scale = beta / l_new;
            const float acc_scale = l_i / l_new * alpha;
            for (int d = 0; d < D; ++d) acc[d] *= acc_scale;
            for (int n = 0; n < BLK; ++n) {
              const float pn = round_to(pr[n] * p_scale, dt);
              if (pn != 0.f) for (int d = 0; d < D; ++d) acc[d] += pn * vt[n * D + d];
            }
            l_i = l_new;
            m_i = m_new;
          }
        }
        for (int d = 0; d < D; ++d)
          store_elem(out, ((t0 + m) * num_heads + h) * D + d, dt, alibi_slopes ? acc[d] / l_i : acc[d]);
      }
    }
  }
  free(kt); free(vt); free(acc); free(qrow);
  return 0;
}
