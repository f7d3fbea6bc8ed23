// Pre-norm feed-forward sublayer of one decode step as a single cluster kernel; see ffn_fused.cu.
#pragma once
#include "common.cuh"

namespace retr {

struct FfnFusedParams {
  float* Y = nullptr;                                  // [M, 256] fp32 residual stream, updated in place
  const float* gamma = nullptr; const float* beta = nullptr; float eps = 1e-5f;
  const bf16* W1 = nullptr; const float* b1 = nullptr; // [d_ff, 256], [d_ff]
  const bf16* W2 = nullptr; const float* b2 = nullptr; // [256, d_ff], [256]
  int M = 0, d_ff = 0;
};

void ffn_fused_configure();                            // opt-in shared memory; call outside stream capture
bool ffn_fused_supported(int d_model, int d_ff);
void ffn_fused(const FfnFusedParams& p, cudaStream_t s);

}  // namespace retr
