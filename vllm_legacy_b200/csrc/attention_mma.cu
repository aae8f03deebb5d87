// Head-size dispatch for the TMA + mma.sync paged-attention family (kernels are instantiated per
// head size in attention_mma_inst.cu so that the objects build in parallel).
#include "attention_common.cuh"

namespace b200pa {

#define B200PA_DECL(D) \
  int launch_attention_mma_d##D(const AttnKernelParams&, int, int, int, cudaStream_t);
B200PA_DECL(64)
B200PA_DECL(80)
B200PA_DECL(96)
B200PA_DECL(112)
B200PA_DECL(128)
B200PA_DECL(192)
B200PA_DECL(256)
#undef B200PA_DECL

int launch_attention_mma(const AttnKernelParams& kp, int num_seqs, int dtype, int kv_dtype, cudaStream_t stream) {
  // 32-bit loads of query pairs
  if (reinterpret_cast<uintptr_t>(kp.q) % 4 != 0 || kp.q_stride % 2 != 0) return 1;
  switch (kp.head_size) {
    case 64: return launch_attention_mma_d64(kp, num_seqs, dtype, kv_dtype, stream);
    case 80: return launch_attention_mma_d80(kp, num_seqs, dtype, kv_dtype, stream);
    case 96: return launch_attention_mma_d96(kp, num_seqs, dtype, kv_dtype, stream);
    case 112: return launch_attention_mma_d112(kp, num_seqs, dtype, kv_dtype, stream);
    case 128: return launch_attention_mma_d128(kp, num_seqs, dtype, kv_dtype, stream);
    case 192: return launch_attention_mma_d192(kp, num_seqs, dtype, kv_dtype, stream);
    case 256: return launch_attention_mma_d256(kp, num_seqs, dtype, kv_dtype, stream);
    default: return 1;
  }
}

}  // namespace b200pa
