// GEMM / implicit-GEMM convolution with fused epilogue — shared parameter block.
//   C[M,N] = epi( A[M,K] · W[N,K]^T )      A, W K-major (row-major, K contiguous)
//   epi(acc) = relu?( acc*scale[n] + bias[n] + rowbias[m % period][n] ) + residual[m][n]
// (ReLU is applied AFTER the residual add when relu == 2: the Bottleneck tail
//  relu(bn3(conv3) + identity), torchvision/models/resnet.py:160-163.)
#pragma once
#include "common.cuh"

namespace retr {

struct ConvView {        // implicit-GEMM view of an NHWC activation as A; kh == 0 -> plain matrix
  int kh = 0, kw = 0, stride = 1, pad = 0, dil = 1;
  int H = 0, W = 0, C = 0, OH = 0, OW = 0;
};

struct GemmParams {
  const void* A = nullptr;      // fp32 or bf16
  const void* W = nullptr;      // same type as A, [N,K]
  void* C = nullptr;            // out_fp32 ? float : A's type; may be null with argmax_out
  int64_t M = 0;
  int N = 0, K = 0;
  int64_t lda = 0, ldc = 0;
  const float* scale = nullptr;
  const float* bias = nullptr;
  const float* rowbias = nullptr;
  int rowbias_period = 1;
  const void* residual = nullptr;
  int64_t ldr = 0;
  bool res_fp32 = false;
  int relu = 0;                 // 0 none, 1 before residual, 2 after residual, 3: ReLU BACKWARD — `residual` is the
                                // forward activation, the result is kept where it was > 0 (bf16 / tcgen05 only)
  bool out_fp32 = false;
  ConvView conv;
  unsigned long long* argmax_out = nullptr;
  // C row remap: out_row = (m / c_group) * c_group_stride + (m % c_group)   (0 = identity).
  // Lets input_proj write image tokens straight into the concatenated [B,S,d] buffer.
  int c_group = 0, c_group_stride = 0;
  // If set, every row uses rowbias[*rowbias_dyn] (decode: the device-side step counter).
  const int* rowbias_dyn = nullptr;
  // > 0: only columns [0, rowbias_cols) carry a per-row bias (a multiple of 32); the columns after it take `bias` alone.
  // Saves the epilogue's gather of table entries that are the same for every row (the V third of a q|k|v projection).
  int rowbias_cols = 0;
  // If set, row m uses rowbias[rowbias_idx[m]] (packed encoder: the token's position in its sequence).
  const int* rowbias_idx = nullptr;
  // If set, only the first *m_dyn rows (<= M) are computed: M sizes the tensor maps and the grid, the persistent loop
  // stops at ceil(*m_dyn / 128) row blocks (packed encoder: the number of unpadded tokens is known on the device only).
  // bf16 / tcgen05 only.
  const int* m_dyn = nullptr;
  // Weight-gradient shape (small M x N, very long K): split the contraction over split_k CTAs per output
  // tile (0 = choose from the SM count), fp32 partial tiles in split_ws ([split_k*M, N] floats), summed into C
  // (C += sum when accumulate).  Requires a plain fp32 C with ldc == N and no epilogue operands.
  bool pdl = true;             // programmatic dependent launch (off where a parked 148-CTA grid would block parallel
                               // graph branches: the vocabulary head of the decode step)
  int split_k = 1;
  float* split_ws = nullptr;
  int64_t split_ws_floats = 0;
  bool accumulate = false;
  // Both operands MN-major (bf16 / tcgen05 only): C[M,N] = A^T . B with A given as [K, M] (row pitch lda) and W as
  // [K, N] (row pitch ldb) — the weight-gradient contraction dW = dY^T . X over tokens/pixels, read in place, no
  // transposed copies.  With a conv view, W is the NHWC layer input and B its im2col [pixels, kh*kw*C] (TMA im2col),
  // K = images*OH*OW.
  bool mn_major = false;
  int64_t ldb = 0;
};

// Vocabulary head of one greedy step (gemm_sm100.cu: vocab_argmax_kernel): keys[m] = max over n of
// pack_argmax(A[m,:] . W[n,:] + bias[n], n); the logits are never written.  With `ticket` set (a zeroed device word) the
// last CTA also does greedy_update's token bookkeeping (models/caption.py:161-174 + eval_utils/decode.py:70-79).
struct VocabArgmaxParams {
  const void* A = nullptr;      // bf16 [M, K], row pitch lda
  const void* W = nullptr;      // bf16 [N, K]
  const float* bias = nullptr;
  int64_t M = 0, lda = 0;
  int N = 0, K = 0;
  unsigned long long* keys = nullptr;
  int64_t* caption = nullptr;
  int max_len = 0, eos = 0;
  int* step = nullptr;
  int* done = nullptr;
  uint8_t* finished = nullptr;
  unsigned int* ticket = nullptr;
};
bool vocab_argmax_supported(int64_t M, int N, int K);
void vocab_argmax(const VocabArgmaxParams& p, cudaStream_t stream);

void gemm_fp32(const GemmParams& p, cudaStream_t stream);   // gemm_simt.cu  (A, W fp32)
void gemm_bf16(const GemmParams& p, cudaStream_t stream);   // gemm_sm100.cu (A, W bf16, tcgen05)
int  sm_count();
// Per-thread cap on the persistent grid of gemm_bf16 (0 = one CTA per SM); set around a phase that shares the GPU.
void gemm_set_cta_cap(int ctas);

}  // namespace retr
