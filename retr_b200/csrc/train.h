// Trainer: the teacher-forced training step of the reference (engine.py:52-87 train_one_epoch,
// main.py:30-40 optimiser recipe) — train-mode forward (dropout, activations kept), backward into
// bound fp32 gradient buffers, and the fused clip + AdamW update.  See train.cu.
#pragma once
#include <memory>
#include <string>

#include "../../include/retr_b200.h"
#include "common.cuh"

namespace retr {

class TrainerBase {
 public:
  virtual ~TrainerBase() {}
  // bind a reference state_dict entry: `param` fp32 device memory read at every step (never copied),
  // `grad` (may be null = frozen / buffer) fp32 device memory the backward pass writes
  virtual void bind(const std::string& name, float* param, float* grad, const int64_t* shape, int ndim) = 0;
  // optional: all trainable parameters / gradients live in two flat fp32 arenas (same element order), the first
  // n_backbone elements being the backbone group -> one memset, one norm kernel, two AdamW launches per step
  virtual void set_arena(float* params, float* grads, int64_t n_total, int64_t n_backbone) = 0;
  virtual void finalize() = 0;
  // train-mode forward; logits_out (optional) fp32 [B,T,V]
  virtual void forward(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                       const float* loc, const int64_t* ids, const uint8_t* id_mask, float* logits_out, float dropout_p,
                       unsigned long long seed, cudaStream_t s) = 0;
  // backward of the last forward from dL/dlogits (fp32 [B,T,V]); every bound grad is overwritten
  virtual void backward(const float* dlogits, cudaStream_t s) = 0;
  // forward + mean cross-entropy against targets [B,T] + backward in one call; the loss stays on the device
  // (loss_dev, fp32) so the caller decides when to synchronise
  virtual void loss_and_grads(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                              const uint8_t* g_mask, const float* loc, const int64_t* ids, const uint8_t* id_mask,
                              const int64_t* targets, float dropout_p, unsigned long long seed, float* loss_dev,
                              cudaStream_t s) = 0;
  // clip_grad_norm_(all grads, max_norm) + AdamW over every trainable tensor; names containing "backbone" use
  // lr_backbone (main.py:31-36).  `step` is the 1-based update count (bias correction).
  virtual void clip_adamw(float max_norm, float lr, float lr_backbone, float beta1, float beta2, float eps,
                          float weight_decay, int step, float* grad_norm_dev, cudaStream_t s) = 0;
  // Data parallel.  set_comm: from now on every backward pass all-reduces (mean over `world` ranks) the
  // gradient arena bucket by bucket, each bucket on a side stream as soon as the backward pass has finished it, and
  // the caller's stream waits for the last one before backward returns control to it (comm == null switches it off).
  // allreduce_grads: the same buckets, issued at once on `s` (for callers that drive the exchange themselves).
  virtual void set_comm(void* nccl_comm, int world) = 0;
  virtual void allreduce_grads(void* nccl_comm, cudaStream_t s) = 0;
  virtual int bucket_info(int64_t* lo_hi, int max_buckets) const = 0;   // -> number of buckets; lo_hi[2*i], [2*i+1]
  long long launches = 0;
};

// pad_token_id: nn.Embedding(padding_idx=...) row that receives no gradient (models/transformer_modules.py:103-105)
std::unique_ptr<TrainerBase> make_trainer(const retr_model_desc& d, int pad_token_id);

}  // namespace retr
