// K6: C-ABI front of causal flash attention for head_dim 128 with GQA on tcgen05.
//
// The kernels live in attention_v3.cu (forward: S / O in TMEM, double-buffered P, lazy rescale) and
// attention_bwd_v3.cu (backward: five GEMMs on transposed score tiles pipelined across query tiles, dK / dV
// resident in TMEM, dQ through fp32 staging + cp.reduce.async.bulk). The first-generation kernels that used to
// be in this file (single P buffer; one-after-another backward schedule) measured 686 / 508 TFLOP/s against
// 836 / 711 for their successors on B200 and were removed once the successors had passed their tests on
// hardware; bz_flash_attn_fwd / bz_flash_attn_bwd keep their signatures and run the current kernels.
//
// The reference has no attention code (it ships no model: SURVEY.md §2.3 F2).
#include "common.cuh"
#include "../include/b200zero.h"

extern "C" int bz_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int Hq, int Hk, int head_dim,
                                 float softmax_scale, void* stream) {
  return bz_flash_attn_fwd_v3(qkv, out, lse, /*error_flag=*/nullptr, B, S, Hq, Hk, head_dim, softmax_scale,
                              /*poly_stride=*/0, stream);
}

// fp32 dQ accumulator [B*S*Hq*128] followed by delta [B*S*Hq]
extern "C" long long bz_flash_attn_bwd_workspace_floats(int B, int S, int Hq) {
  return (long long)B * S * Hq * 128 + (long long)B * S * Hq;
}

extern "C" int bz_flash_attn_bwd(const void* qkv, const void* out, const float* lse, const void* dout, void* dqkv,
                                 float* workspace, int B, int S, int Hq, int Hk, int head_dim, float softmax_scale,
                                 void* stream) {
  return bz_flash_attn_bwd_v3(qkv, out, lse, dout, dqkv, workspace, /*error_flag=*/nullptr, B, S, Hq, Hk, head_dim,
                              softmax_scale, stream);
}
