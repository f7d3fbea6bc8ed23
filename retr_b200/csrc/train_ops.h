// Kernels of the training step (engine.py:52-87 train_one_epoch: teacher-forced forward with
// dropout, cross-entropy, backward, global-norm clip, AdamW) that are not GEMMs; see train_ops.cu.
// Activations are T (float in the fp32 parity mode, bf16 in the tcgen05 mode); gradients of the
// residual streams, LayerNorm statistics, the loss and every parameter gradient are fp32.
#pragma once
#include "common.cuh"

namespace retr {

// ---- dropout (Philox4x32-10 counter RNG: element idx of site `site` under `seed`) ---------------
struct DropCfg {
  float p = 0.f;                  // drop probability (0 = identity)
  unsigned long long seed = 0;
  unsigned site = 0;              // one id per dropout call site per layer
};

// ---- layout helpers for the wgrad GEMMs (contraction over rows needs row-contiguous operands) --------
// out[n, m] = in[m, n] for m < M, 0 for M <= m < ldo   (in: TI [M,N] pitch ld_in; out: TO [N, ldo])
template <typename TI, typename TO>
void transpose_pad(const TI* in, int64_t ld_in, TO* out, int64_t ldo, int64_t M, int N, cudaStream_t s);
// colT[(r*kw+s)*C + c, m] = x[n, oh*stride-pad+r*dil, ow*stride-pad+s*dil, c] (0 outside / for m >= M)
template <typename T>
void im2col_T(const T* x, T* colT, int64_t ldm, int64_t nimg, int H, int W, int C, int k, int stride, int pad, int dil,
              int OH, int OW, cudaStream_t s);
// dx[n,ih,iw,c] (+)= sum over taps of dcol[m(oh,ow), (r*k+s)*C + c]     (dcol: [M, k*k*C])
template <typename T>
void col2im(const T* dcol, T* dx, int64_t nimg, int H, int W, int C, int k, int stride, int pad, int dil, int OH, int OW,
            const T* add, cudaStream_t s);
// Filter of the input-gradient convolution (stride 1): out[c, ((kh-1-r)*kw + (kw-1-s))*O + o] = w[o, c, r, s]
// (OIHW fp32 master -> [C_in, (kh,kw,C_out)] with the taps flipped), so dx = conv(dy, out) with pad' = dil*(k-1) - pad.
template <typename T>
void repack_filter_dgrad(const float* oihw, T* out, int O, int I, int kh, int kw, cudaStream_t s);
// dW[o, c, r, s] = dWk[o, (r*kw+s)*C + c]      (fp32; kpad = row pitch of dWk)
void filter_grad_unpack(const float* dwk, float* dw, int O, int C, int kh, int kw, int kpad, cudaStream_t s);
// sum of partial C tiles written by a split-K GEMM: out[i] (+)= sum_s part[s*n + i]
void splitk_reduce(const float* part, float* out, int64_t n, int splits, int accumulate, cudaStream_t s);

// ---- elementwise backward pieces ---------------------------------------------------------------
// g = dy * (out > 0) [* scale[c]]; optionally also the unscaled masked gradient (identity branch)
template <typename T>
void relu_bn_bwd(const T* dy, const T* out, const float* scale, T* g_scaled, T* g_plain, int64_t M, int C,
                 cudaStream_t s);
// out[n] (+)= sum_m dy[m, n]
template <typename T>
void colsum(const T* dy, int64_t ld, float* out, int64_t M, int N, int accumulate, cudaStream_t s);
// out = x + dropout(branch)  (fp32 streams)
void dropout_add(const float* x, const float* branch, float* out, int64_t n, DropCfg dc, cudaStream_t s);
// g = T( dropout_mask * dx )   (backward of the residual-branch dropout; p = 0 -> a cast)
template <typename T>
void dropout_bwd_cast(const float* dx, T* g, int64_t n, DropCfg dc, cudaStream_t s);
// ReLU backward in place on a T gradient: g *= (h > 0)
template <typename T>
void relu_bwd(T* g, const T* h, int64_t n, cudaStream_t s);

// ---- LayerNorm --------------------------------------------------------------------------------------
// y = LN(x); y2 = y + pos[r % period] (optional second output: the q/k projection input)
template <typename T>
void layernorm_pos(const float* x, const float* gamma, const float* beta, T* y, T* y2, const float* pos, int period,
                   int64_t rows, int d, float eps, cudaStream_t s);
// dx (+)= LN'(x)·(dy1 + dy2); dgamma, dbeta += ...   (dy fp32; dy2 optional)
void layernorm_bwd(const float* x, const float* gamma, const float* dy1, const float* dy2, float* dx, int accumulate_dx,
                   float* dgamma, float* dbeta, int64_t rows, int d, float eps, cudaStream_t s);

// ---- attention with stored probabilities (training) -----------------------------------------------------
struct AttnTrainParams {
  const void* q; const void* k; const void* v;      // T, rows [b*L + i], heads side by side
  int64_t ldq, ldk, ldv;
  void* out; int64_t ldo;                           // T [B*Lq, d]
  void* probs;                                      // T [B,H,Lq,Lk]  softmax probabilities BEFORE dropout (SIMT path)
  float* lse;                                       // fp32 [B,H,Lq] log2-sum-exp per query row (tensor-core path)
  int B, Lq, Lk, H, dh;
  const uint8_t* key_pad;                           // [B,Lk] non-zero = masked
  int causal;
  DropCfg drop;
};
template <typename T> void attention_train_fwd(const AttnTrainParams& p, cudaStream_t s);
struct AttnBwdParams {
  AttnTrainParams f;                                // forward operands (out unused)
  const void* dout; int64_t lddo;                   // T [B*Lq, d]
  void* ds;                                         // T [B,H,Lq,Lk] scratch: softmax-backward scores (SIMT path)
  float* delta;                                     // fp32 [B,H,Lq] scratch: rowsum(dO o O) (tensor-core path)
  void* dq; void* dk; void* dv;                     // T, same pitches as q, k, v
  int64_t lddq, lddk, lddv;
};
template <typename T> void attention_train_bwd(const AttnBwdParams& p, cudaStream_t s);
// bf16 tensor-core (mma.sync) flash variants, attention_train_mma.cu: nothing of size [Lq,Lk] reaches HBM.
// f.out must hold the forward output (delta = rowsum(dO o O)).
void flash_train_fwd_bf16(const AttnTrainParams& p, cudaStream_t s);
void flash_train_bwd_bf16(const AttnBwdParams& p, cudaStream_t s);

// ---- embeddings ------------------------------------------------------------------------------------------
// y = dropout(LN_eps(E_word[ids] + E_pos[t]));  pre = the LN input kept for the backward pass
void embed_ln_train(const int64_t* ids, const float* e_word, const float* e_pos, const float* gamma, const float* beta,
                    float* pre, float* y, int B, int Tn, int d, float eps, DropCfg dc, cudaStream_t s);
// backward of the above: scatters into dE_word / dE_pos (atomic fp32), accumulates dgamma / dbeta
void embed_ln_bwd(const int64_t* ids, const float* pre, const float* gamma, const float* dy, float* d_word, float* d_pos,
                  float* dgamma, float* dbeta, float* tmp /* [2*B*Tn*d] */, int B, int Tn, int d, float eps, DropCfg dc,
                  int pad_id, cudaStream_t s);
// dE_pos[t, :] += sum_b g[b*Tn + t, :]   (query_pos added to projection inputs)
void pos_grad_accumulate(const float* g, float* d_pos, int B, int Tn, int d, cudaStream_t s);

// ---- location tokens ----------------------------------------------------------------------------------------
void loc_tokens_bwd(const float* dx, const float* loc, float* dw, float* db, int B, int n_loc, int per_scalar, int d,
                    int S, int off, cudaStream_t s);

// ---- loss -------------------------------------------------------------------------------------------------------
// Mean cross-entropy over ALL rows (pads included, models/caption.py:210) and its gradient in place:
// logits[r, :] <- (softmax(logits[r,:]) - onehot(target[r])) * grad_scale ; loss_sum += -log p[target]
template <typename T>
void ce_loss_fwd_bwd(T* logits, int64_t ld, const int64_t* targets, int64_t rows, int V, float grad_scale,
                     float* loss_sum, cudaStream_t s);

// ---- optimiser ----------------------------------------------------------------------------------------------------
// sumsq[0] += sum g^2 over a flat buffer
void grad_sumsq(const float* g, int64_t n, float* sumsq, cudaStream_t s);
// clip_grad_norm_ (engine.py:82) + AdamW (main.py:38-39) fused over a flat range:
//   coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)); g *= coef; decoupled weight decay; bias-corrected update
void adamw_step(float* p, const float* g, float* m, float* v, int64_t n, const float* sumsq, float max_norm, float lr,
                float beta1, float beta2, float eps, float weight_decay, int step, cudaStream_t s);

}  // namespace retr
