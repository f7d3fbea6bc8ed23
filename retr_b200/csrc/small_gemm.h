// Decode-step GEMM (small M) with optional LayerNorm prologue and deterministic split-K; see small_gemm.cu.
#pragma once
#include "common.cuh"

namespace retr {

struct SmallGemmParams {
  const bf16* A = nullptr; int64_t lda = 0;            // bf16 input, used when X == nullptr
  const float* X = nullptr; int64_t ldx = 0;           // fp32 input normalised in the prologue (K == d_model)
  const float* gamma = nullptr; const float* beta = nullptr; float eps = 1e-5f;
  const bf16* W = nullptr;                             // [N, K]
  void* C = nullptr; int64_t ldc = 0; int out_fp32 = 0;
  int M = 0, N = 0, K = 0;
  const float* bias = nullptr;
  const float* rowbias = nullptr; const int* rowbias_dyn = nullptr;   // rowbias[*rowbias_dyn][N]
  const float* residual = nullptr; int64_t ldr = 0;    // fp32, may alias C
  int relu = 0;
  int split_k = 1; float* scratch = nullptr; unsigned* tickets = nullptr;
};

void small_gemm_configure();                           // opt-in shared memory; call outside stream capture
bool small_gemm_supported(int M, int N, int K, bool ln);
void small_gemm(const SmallGemmParams& p, cudaStream_t s);
// C[M,N] (fp32) = A[M,K] . W[N,K]^T + bias, deterministic split-K on file-scope static device scratch: allocates
// nothing (see small_gemm.cu); split_k * M * N <= 1 Mi floats.
void small_gemm_splitk_static(const bf16* A, const bf16* W, const float* bias, float* C, int M, int N, int K,
                              int split_k, cudaStream_t s);

}  // namespace retr
