// TEST INFRASTRUCTURE: pins the oracle's fp8 restatement against the third-party arithmetic the
// reference calls — CUDA 12.9 cuda_fp8.h (__nv_cvt_float_to_fp8 with __NV_SATFINITE,
// __nv_cvt_fp8_to_halfraw; host-callable) — call sites quant_utils.cuh:298,355,410,461,474,486.
// Exhaustive over all fp16 and bf16 inputs, all 256 fp8 codes, and a 16M-point sweep of fp32 bit
// patterns, for both e4m3 and e5m2.  Exit code 0 = identical.
#include <cuda_fp8.h>

#include <cstdint>
#include <cstdio>
#include <cstring>

extern "C" {
uint8_t oracle_float_to_fp8(float f, int kv);
uint16_t oracle_fp8_to_half(uint8_t b, int kv);
float oracle_half_to_float(uint16_t h);
uint16_t oracle_float_to_half(float f);
float oracle_bf16_to_float(uint16_t b);
uint16_t oracle_float_to_bf16(float f);
}

static float bits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

int main() {
  long bad = 0, n = 0;
  for (int kv = 1; kv <= 2; ++kv) {
    const __nv_fp8_interpretation_t it = kv == 1 ? __NV_E4M3 : __NV_E5M2;
    auto check = [&](float f) {
      const uint8_t a = oracle_float_to_fp8(f, kv);
      const uint8_t b = (uint8_t)__nv_cvt_float_to_fp8(f, __NV_SATFINITE, it);
      ++n;
      if (a != b && bad++ < 10) { uint32_t u; memcpy(&u, &f, 4); printf("kv=%d f=%g (0x%08x): oracle 0x%02x cuda 0x%02x\n", kv, f, u, a, b); }
    };
    for (uint32_t h = 0; h < 65536; ++h) {
      check(oracle_half_to_float((uint16_t)h));
      check(oracle_bf16_to_float((uint16_t)h));
    }
    for (uint64_t u = 0; u < (1ull << 32); u += 251) check(bits((uint32_t)u));
    for (int b = 0; b < 256; ++b) {
      const uint16_t a = oracle_fp8_to_half((uint8_t)b, kv);
      const __half_raw r = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, it);
      ++n;
      const bool both_nan = ((a & 0x7fff) > 0x7c00) && ((r.x & 0x7fff) > 0x7c00);
      if (a != r.x && !both_nan && bad++ < 10) printf("kv=%d code 0x%02x: oracle 0x%04x cuda 0x%04x\n", kv, b, a, r.x);
    }
  }
  // fp16 / bf16 scalar conversions against the toolkit's host implementations
  for (uint64_t u = 0; u < (1ull << 32); u += 509) {
    const float f = bits((uint32_t)u);
    const __half_raw hr = __float2half_rn(f);
    const uint16_t a = oracle_float_to_half(f);
    const bool both_nan = ((a & 0x7fff) > 0x7c00) && ((hr.x & 0x7fff) > 0x7c00);
    ++n;
    if (a != hr.x && !both_nan && bad++ < 10) printf("f2h 0x%08x: oracle 0x%04x cuda 0x%04x\n", (uint32_t)u, a, hr.x);
  }
  printf("fp8_pin: %ld comparisons, %ld mismatches\n", n, bad);
  return bad ? 1 : 0;
}
