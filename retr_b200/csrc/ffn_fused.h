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
  // Optional L2 prefetch issued at kernel entry: the head-major cross-attention K/V blocks ([M][8][pf_rows][pf_dh] bf16,
  // pf_nvalid[m] valid rows each) of the NEXT decoder layer for this kernel's rows.  HBM is idle during the FFN /
  // projection phase of a decode step; the next cross-attention then reads L2 instead of DRAM.
  const bf16* pf_k = nullptr; const bf16* pf_v = nullptr; const int* pf_nvalid = nullptr;
  int pf_rows = 0, pf_dh = 0;
};

// y <- y + Wo att + bo ; q = Wq LN(y) + bq + rowbias[*rowbias_dyn]   (d_model = 256; see ffn_fused.cu)
struct AttnOutProjParams {
  const bf16* att = nullptr; int64_t ld_att = 0;       // [M, 256] attention output (heads concatenated)
  const bf16* Wo = nullptr; const float* bo = nullptr; // [256, 256], [256]
  float* Y = nullptr;                                  // [M, 256] fp32 residual stream, updated in place
  const float* gamma = nullptr; const float* beta = nullptr; float eps = 1e-5f;
  const bf16* Wq = nullptr; const float* bq = nullptr; // [256, 256], [256] (may be null)
  const float* rowbias = nullptr; const int* rowbias_dyn = nullptr;   // [rows, 256] table, row picked on the device
  bf16* Q = nullptr; int64_t ld_q = 0;                 // [M, 256]
  int M = 0;
};
void attn_out_ln_proj(const AttnOutProjParams& p, cudaStream_t s);

void ffn_fused_configure();                            // opt-in shared memory; call outside stream capture
bool ffn_fused_supported(int d_model, int d_ff);
void ffn_fused(const FfnFusedParams& p, cudaStream_t s);

}  // namespace retr
