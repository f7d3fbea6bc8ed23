// TrainerT<T>: schedule of one training step of the reference (engine.py:64-83):
//   outputs = model(*samples, caps[:, :-1], cap_masks[:, :-1])   train mode, dropout p      -> forward()
//   loss = CrossEntropyLoss()(outputs.permute(0,2,1), caps[:, 1:])   mean over B*T, pads included
//   loss.backward()                                                                         -> backward()
//   clip_grad_norm_(model.parameters(), max_norm); optimizer.step()   (AdamW, main.py:30-40)  -> clip_adamw()
// T = float: fp32 SIMT GEMMs (parity mode);  T = bf16: tcgen05 GEMMs, fp32 residual streams / LayerNorm /
// softmax / loss / gradients of parameters.  Parameters are READ from the bound fp32 tensors at every step
// (the optimiser updates them in place); gradients are written to the bound fp32 buffers.
//
// Backward contractions reuse the forward GEMM kernel: dX = dY·W needs W^T (refreshed per step), dW = dY^T·X
// contracts over tokens/pixels, so both operands are transposed (transpose_pad / im2col_T) and the GEMM runs
// split-K.  Frozen parts (conv1, layer1 — models/backbone.py:58-60 — and all FrozenBatchNorm2d buffers) get no
// gradient and nothing below them is back-propagated.
#include "train.h"
#include "dp.h"
#include "dropout.cuh"

#include <algorithm>
#include <map>
#include <type_traits>
#include <vector>

#include "gemm.h"
#include "ops.h"
#include "train_ops.h"

namespace retr {
void gemm_bf16_prepare();
namespace {

template <typename T> void gemmT(const GemmParams& p, cudaStream_t s);
template <> void gemmT<float>(const GemmParams& p, cudaStream_t s) { gemm_fp32(p, s); }
template <> void gemmT<bf16>(const GemmParams& p, cudaStream_t s) { gemm_bf16(p, s); }

inline int64_t rup(int64_t a, int64_t b) { return cdiv(a, b) * b; }

struct PBind {
  float* p = nullptr;
  float* g = nullptr;
  std::vector<int64_t> shape;
  int64_t numel = 0;
  float *m = nullptr, *v = nullptr;     // AdamW state (lazily allocated)
};

template <typename T> struct TLin {
  const float* w = nullptr; const float* b = nullptr;
  float* gw = nullptr; float* gb = nullptr;
  int out = 0, in = 0, out_pad = 0;
  T* wT = nullptr;        // [out, in]
  T* wtT = nullptr;       // [in, out_pad]
};
template <typename T> struct TConv {
  const float* w = nullptr; float* gw = nullptr;
  int cin = 0, cout = 0, k = 1, stride = 1, pad = 0, dil = 1, kpad = 0;
  T* wk = nullptr;        // [cout, kpad]
  T* wkt = nullptr;       // [kpad, cout]
  T* wkd = nullptr;       // [cin, k*k*cout], taps flipped: filter of the input-gradient convolution (stride-1 k>1, bf16)
  float *scale = nullptr, *shift = nullptr;
};
template <typename T> struct TBlock {
  TConv<T> c1, c2, c3, ds;
  bool has_ds = false, trainable = false;
  bool basic = false;     // torchvision BasicBlock (ResNet-18/34): c1 = 3x3 (stride), c2 = 3x3, no c3
  int out_channels() const { return basic ? c2.cout : c3.cout; }
  int Hin = 0, Hout = 0;
  T *o1 = nullptr, *o2 = nullptr, *idn = nullptr, *y = nullptr;     // kept activations (trainable blocks)
};
template <typename T> struct TEnc {
  float *g1, *b1, *g2, *b2, *gg1, *gb1, *gg2, *gb2;                 // LayerNorm params / grads
  TLin<T> qk, v, o, ff1, ff2;
  float *gin_b = nullptr;                                            // in_proj_bias grad [3d]
  float *xa, *xb;                                                    // fp32 stream before each sublayer
  T *n1, *n1p, *qkv, *P, *att, *n2, *h;
  float* lse = nullptr;
};
template <typename T> struct TDec {
  float *g1, *b1, *g2, *b2, *g3, *b3, *gg1, *gb1, *gg2, *gb2, *gg3, *gb3;
  TLin<T> s_qk, s_v, s_o, c_q, c_k, c_v, c_o, ff1, ff2;
  float *gs_b = nullptr, *gc_b = nullptr;                            // in_proj_bias grads [3d]
  float *ya, *yb, *yc;
  T *n1, *n1p, *qkv, *Ps, *att_s, *n2p, *qc, *ckv, *Pc, *att_c, *n3, *h;
  float *lse_s = nullptr, *lse_c = nullptr;
};

__global__ void cast_rows_kernel(const float* __restrict__ in, int64_t ld_in, bf16* __restrict__ out, int64_t ld_out,
                                 int64_t M, int N) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M * N; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / N;
    const int c = (int)(i - r * N);
    out[r * ld_out + c] = __float2bfloat16_rn(in[r * ld_in + c]);
  }
}
__global__ void cast_rows_kernel(const float* __restrict__ in, int64_t ld_in, float* __restrict__ out, int64_t ld_out,
                                 int64_t M, int N) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M * N; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / N;
    const int c = (int)(i - r * N);
    out[r * ld_out + c] = in[r * ld_in + c];
  }
}
// image-token rows of the concatenated [B,S,d] gradient -> compact [n_images*B*g2, d] in T
template <typename T>
__global__ void gather_tok_kernel(const float* __restrict__ dx, T* __restrict__ out, int B, int g2, int S, int off_g,
                                  int n_images, int d) {
  const int64_t n = (int64_t)n_images * B * g2 * d;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % d);
    int64_t r = i / d;
    const int j = (int)(r % g2); r /= g2;
    const int b = (int)(r % B);
    const int set = (int)(r / B);
    out[i] = from_f<T>(dx[((int64_t)b * S + (set ? off_g : 0) + j) * d + c]);
  }
}
// memory = x, memory + pos (pre_norm=False: no final encoder LayerNorm)
template <typename T>
__global__ void cast_pos_kernel(const float* __restrict__ x, T* __restrict__ y, T* __restrict__ y2,
                                const float* __restrict__ pos, int period, int64_t n, int d) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = x[i];
    y[i] = from_f<T>(v);
    y2[i] = from_f<T>(v + pos[((i / d) % period) * d + i % d]);
  }
}
// learned positions in train mode (models/position_encoding.py:38-63): the LayerNormed table is repeated per sample and
// dropped out per element; backward sums the (masked) gradients of every use over the batch
__global__ void pos_expand_kernel(const float* __restrict__ pos_ln, float* __restrict__ posD, int64_t n, int64_t sd,
                                  DropCfg dc) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    posD[i] = pos_ln[i % sd] * drop_scale(dc, (uint64_t)i);
}
__global__ void pos_reduce_kernel(const float* __restrict__ dPos, float* __restrict__ out, int B, int64_t sd, DropCfg dc) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < sd; j += (int64_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int b = 0; b < B; ++b) acc += dPos[b * sd + j] * drop_scale(dc, (uint64_t)(b * sd + j));
    out[j] = acc;
  }
}
__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] += x[i];
}
inline unsigned ewb(int64_t n) { return (unsigned)std::min<int64_t>(cdiv(n, 256), 148 * 16); }

template <typename T>
class TrainerT final : public TrainerBase {
 public:
  TrainerT(const retr_model_desc& d, int pad_token_id) : desc(d), pad_id_(pad_token_id) {
    RETR_CHECK(d.n_images == 1 || d.n_images == 2, "n_images must be 1 or 2");
    CUDA_OK(cudaSetDevice(d.device));
    int side = d.image_size;
    side = (side + 2 * 3 - 7) / 2 + 1;
    side = (side + 2 * 1 - 3) / 2 + 1;
    for (int st = 1; st < 4; ++st)
      if (!(st == 3 && d.dilation)) side = (side + 2 * 1 - 3) / 2 + 1;
    grid_hw = side;
    n_loc_tok = d.loc_in_dim == 0 ? 0 : (d.loc_in_dim == 1 ? d.n_loc : 1);
    S = grid_hw * grid_hw * d.n_images + n_loc_tok;
    Vp = (int)rup(d.vocab, 64);
  }
  ~TrainerT() override {
    for (void* p : owned_) cudaFree(p);
    if (dp_stream_) cudaStreamDestroy(dp_stream_);
    for (auto e : dp_ready_) cudaEventDestroy(e);
    if (dp_done_) cudaEventDestroy(dp_done_);
  }
  void set_comm(void* nccl_comm, int world) override {
    RETR_CHECK(finalized_, "set_comm before finalize");
    RETR_CHECK(!nccl_comm || arena_g_, "data-parallel training needs the flat gradient arena (retr_train_set_arena)");
    if (nccl_comm) nccl();                                 // load the library now, not in the middle of a step
    comm_ = static_cast<nccl_comm_t>(nccl_comm); world_ = world;
    if (comm_ && !dp_stream_) {
      CUDA_OK(cudaStreamCreateWithFlags(&dp_stream_, cudaStreamNonBlocking));
      for (size_t i = 0; i < buckets_.size(); ++i) {
        cudaEvent_t e;
        CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        dp_ready_.push_back(e);
      }
      CUDA_OK(cudaEventCreateWithFlags(&dp_done_, cudaEventDisableTiming));
    }
  }
  void allreduce_grads(void* nccl_comm, cudaStream_t s) override {
    RETR_CHECK(nccl_comm && arena_g_, "allreduce_grads needs a communicator and the flat gradient arena");
    nccl_ok(nccl().GroupStart(), "ncclGroupStart");
    for (const GradBucket& b : buckets_)
      nccl_ok(nccl().AllReduce(arena_g_ + b.lo, arena_g_ + b.lo, (size_t)(b.hi - b.lo), NCCL_FLOAT32, NCCL_AVG,
                               static_cast<nccl_comm_t>(nccl_comm), s), "ncclAllReduce");
    nccl_ok(nccl().GroupEnd(), "ncclGroupEnd");
  }
  int bucket_info(int64_t* lo_hi, int max_buckets) const override {
    for (int i = 0; i < (int)buckets_.size() && i < max_buckets; ++i) { lo_hi[2 * i] = buckets_[i].lo; lo_hi[2 * i + 1] = buckets_[i].hi; }
    return (int)buckets_.size();
  }
  void bind(const std::string& name, float* param, float* grad, const int64_t* shape, int ndim) override {
    PBind b;
    b.p = param; b.g = grad; b.numel = 1;
    for (int i = 0; i < ndim; ++i) { b.shape.push_back(shape[i]); b.numel *= shape[i]; }
    auto it = params_.find(name);
    if (it != params_.end()) { b.m = it->second.m; b.v = it->second.v; }
    RETR_CHECK(!finalized_, "bind after finalize");
    params_[name] = b;
  }
  void set_arena(float* params, float* grads, int64_t n_total, int64_t n_backbone) override {
    arena_p_ = params; arena_g_ = grads; arena_n_ = n_total; arena_nb_ = n_backbone;
  }
  void finalize() override;
  void forward(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
               const float* loc, const int64_t* ids, const uint8_t* id_mask, float* logits_out, float dropout_p,
               unsigned long long seed, cudaStream_t s) override;
  void backward(const float* dlogits, cudaStream_t s) override;
  void loss_and_grads(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                      const float* loc, const int64_t* ids, const uint8_t* id_mask, const int64_t* targets,
                      float dropout_p, unsigned long long seed, float* loss_dev, cudaStream_t s) override;
  void clip_adamw(float max_norm, float lr, float lr_backbone, float beta1, float beta2, float eps, float weight_decay,
                  int step, float* grad_norm_dev, cudaStream_t s) override;

 private:
  template <typename U> U* alloc(int64_t n) {
    void* p = nullptr;
    CUDA_OK(cudaMalloc(&p, std::max<int64_t>(n, 4) * sizeof(U)));
    owned_.push_back(p);
    return static_cast<U*>(p);
  }
  const PBind& P(const std::string& n) const {
    auto it = params_.find(n);
    RETR_CHECK(it != params_.end(), "missing parameter '" + n + "' (bind every state_dict entry)");
    return it->second;
  }
  TLin<T> lin(const std::string& wname, const std::string& bname, int row0 = 0, int rows = -1);
  TConv<T> conv(const std::string& wname, const std::string& bn, int stride, int pad, int dil, bool trainable);
  void refresh(const TLin<T>& l, cudaStream_t s);
  void refresh(const TConv<T>& c, cudaStream_t s);
  void refresh_all(cudaStream_t s);
  GemmParams lp(const TLin<T>& l, const void* A, int64_t lda, int64_t M, void* C, int64_t ldc) {
    GemmParams p;
    p.A = A; p.lda = lda; p.W = l.wT; p.C = C; p.M = M; p.N = l.out; p.K = l.in; p.ldc = ldc; p.bias = l.b;
    return p;
  }
  void run_conv(const TConv<T>& c, const T* in, int64_t n, int H, int W, T* out, const T* residual, int relu, int* OH,
                cudaStream_t s);
  void branch_out(const TLin<T>& l, const T* A, int64_t M, const float* x_in, float* x_out, DropCfg dc, cudaStream_t s);
  void wgrad(const T* dy, int64_t ld_dy, int n_out, const T* x, int64_t ld_x, int k_in, int64_t M, float* gw,
             bool accumulate, cudaStream_t s);
  void lin_bwd(const TLin<T>& l, const T* dy, int64_t ld_dy, const T* x, int64_t ld_x, int64_t M, void* dx, bool dx_fp32,
               float* bias_grad, cudaStream_t s);
  // returns true when the ReLU/FrozenBN backward of the layer BELOW (dx <- dx * (relu_out > 0) * relu_scale[c]) was
  // folded into the input-gradient GEMM's epilogue; otherwise the caller runs relu_bn_bwd on dx
  bool conv_bwd(const TConv<T>& c, const T* g, const T* x, int64_t nimg, int H, int OH, T* dx, const T* dx_add,
                bool need_dx, cudaStream_t s, const T* relu_out = nullptr, const float* relu_scale = nullptr);
  void attn_fwd(const T* q, int64_t ldq, const T* k, const T* v, int64_t ldkv, T* out, T* probs, float* lse, int Lq,
                int Lk, const uint8_t* pad, int causal, unsigned site, cudaStream_t s);
  void attn_bwd(const T* q, int64_t ldq, const T* k, const T* v, int64_t ldkv, const T* probs, const float* lse,
                const T* out, const T* dout, T* dq,
                int64_t lddq, T* dk, T* dv, int64_t lddkv, int Lq, int Lk, const uint8_t* pad, int causal, unsigned site,
                cudaStream_t s);
  void forward_impl(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                    const float* loc, const int64_t* ids, const uint8_t* id_mask, float* logits_out, bool logits_T,
                    float dropout_p, unsigned long long seed, cudaStream_t s);
  void backward_impl(cudaStream_t s);
  DropCfg drop(unsigned site) const { DropCfg dc; dc.p = drop_p_; dc.seed = seed_; dc.site = site; return dc; }

  retr_model_desc desc;
  int pad_id_ = 0;
  int S = 0, grid_hw = 0, n_loc_tok = 0, Vp = 0;
  bool frozen_ready_ = false;
  T* block_in_ = nullptr;                              // output of the last frozen block (input of layer2.0)
  std::map<std::string, PBind> params_;
  std::vector<void*> owned_;
  bool finalized_ = false;

  TConv<T> stem_;
  std::vector<TBlock<T>> blocks_;
  int first_trainable_ = 0;
  bool train_backbone_ = false;
  TLin<T> input_proj_;
  std::vector<TEnc<T>> enc_;
  std::vector<TDec<T>> dec_;
  TLin<T> mlp_[3];
  // learned position table ('v3' / 'learned'): LN(pos_embed.weight[:S]) -> per-sample dropout (train mode)
  bool learned_ = false;
  const float *pos_w_ = nullptr, *pos_g_ = nullptr, *pos_b_ = nullptr;
  float *g_pos_w_ = nullptr, *g_pos_g_ = nullptr, *g_pos_b_ = nullptr;
  float *pos_ln_ = nullptr, *posD_ = nullptr, *dPos_ = nullptr, *dposln_ = nullptr;
  // the reference hard-codes this dropout to 0.1 (build_position_encoding); it follows the step's dropout switch here
  DropCfg pos_drop() const { DropCfg dc; dc.p = drop_p_ > 0.f ? 0.1f : 0.f; dc.seed = seed_; dc.site = 2; return dc; }
  const float* pos_ptr() const { return learned_ ? posD_ : pe_; }
  int pos_period(int64_t Me) const { return learned_ ? (int)Me : S; }
  const float *pe_ = nullptr, *e_word_ = nullptr, *e_pos_ = nullptr, *emb_g_ = nullptr, *emb_b_ = nullptr;
  float *g_e_word_ = nullptr, *g_e_pos_ = nullptr, *g_emb_g_ = nullptr, *g_emb_b_ = nullptr;
  const float *enc_ng_ = nullptr, *enc_nb_ = nullptr, *dec_ng_ = nullptr, *dec_nb_ = nullptr;
  float *g_enc_ng_ = nullptr, *g_enc_nb_ = nullptr, *g_dec_ng_ = nullptr, *g_dec_nb_ = nullptr;
  const float *loc_w_ = nullptr, *loc_b_ = nullptr;
  float *g_loc_w_ = nullptr, *g_loc_b_ = nullptr;

  // activations
  T *im2col_ = nullptr, *a0_ = nullptr, *a1_ = nullptr, *fz_[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  T* feat_ = nullptr;                                  // == last block's y
  float *x_final_ = nullptr, *y_final_ = nullptr, *emb_pre_ = nullptr;
  T *memory_ = nullptr, *memp_ = nullptr, *hs_ = nullptr, *m1_ = nullptr, *m2_ = nullptr, *dlg_ = nullptr;
  uint8_t *tok_mask_ = nullptr, *id_mask_ = nullptr;
  int64_t* ids_ = nullptr;
  float* loc_ = nullptr;
  // backward scratch
  float *dX_ = nullptr, *dY_ = nullptr, *dMem_ = nullptr, *tmpf_ = nullptr, *tmpf2_ = nullptr, *embtmp_ = nullptr;
  T *gb_ = nullptr, *dh_ = nullptr, *datt_ = nullptr, *dqkv_ = nullptr, *dckv_ = nullptr, *dS_ = nullptr, *dtok_ = nullptr;
  T *tA_ = nullptr, *tB_ = nullptr, *dcol_ = nullptr;
  int64_t tA_cap_ = 0, tB_cap_ = 0;
  T *bb_[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // backbone gradient buffers
  float *dwk_ = nullptr, *ws_ = nullptr, *sumsq_ = nullptr, *delta_ = nullptr;
  // bf16: tensor-core flash attention (log-sum-exp kept, probabilities recomputed in the backward pass)
  bool flash_ = std::is_same<T, bf16>::value && getenv("RETR_TRAIN_SIMT_ATTENTION") == nullptr;
  int64_t ws_floats_ = 0;
  float *arena_p_ = nullptr, *arena_g_ = nullptr, *arena_m_ = nullptr, *arena_v_ = nullptr;
  int64_t arena_n_ = 0, arena_nb_ = 0;
  int cur_B_ = 0;
  float drop_p_ = 0.f;
  unsigned long long seed_ = 0;
  // data parallel: gradient buckets in REVERSE-EXECUTION order (head, decoder + embeddings, encoder + token
  // projections, layer4, layer3, layer2), each a contiguous range of the arena
  std::vector<GradBucket> buckets_;
  void build_buckets();
  void bucket_ready(int k, cudaStream_t s);              // gradients of bucket k are final on stream s
  void buckets_join(cudaStream_t s);                     // s waits for every bucket's all-reduce
  nccl_comm_t comm_ = nullptr;
  int world_ = 1;
  cudaStream_t dp_stream_ = nullptr;
  std::vector<cudaEvent_t> dp_ready_;
  cudaEvent_t dp_done_ = nullptr;
};

static const char* const kFallbackBucket = "all";
extern const char* kBucketNamesFwd[6];
template <typename T>
void TrainerT<T>::build_buckets() {
  const char* const* kBucketNames = kBucketNamesFwd;
  buckets_.clear();
  if (!arena_g_) return;
  auto bucket_of = [](const std::string& n) {
    if (n.rfind("mlp.", 0) == 0) return 0;
    if (n.rfind("transformer.decoder.", 0) == 0 || n.rfind("transformer.embeddings.", 0) == 0) return 1;
    if (n.find("backbone") != std::string::npos) {
      if (n.find("layer4") != std::string::npos) return 3;
      if (n.find("layer3") != std::string::npos) return 4;
      return 5;
    }
    return 2;
  };
  int64_t lo[6], hi[6], cnt[6];
  for (int i = 0; i < 6; ++i) { lo[i] = arena_n_; hi[i] = 0; cnt[i] = 0; }
  for (auto& kv : params_) {
    if (!kv.second.g) continue;
    const int64_t off = kv.second.g - arena_g_;
    RETR_CHECK(off >= 0 && off + kv.second.numel <= arena_n_, "gradient of '" + kv.first + "' is outside the arena");
    const int k = bucket_of(kv.first);
    lo[k] = std::min(lo[k], off); hi[k] = std::max(hi[k], off + kv.second.numel); cnt[k] += kv.second.numel;
  }
  bool tiled = true;
  int64_t total = 0;
  for (int i = 0; i < 6; ++i) {
    if (cnt[i] == 0) continue;
    if (hi[i] - lo[i] != cnt[i]) tiled = false;            // a bucket's tensors must be contiguous in the arena
    total += cnt[i];
  }
  if (!tiled || total != arena_n_) {                       // unexpected arena order: one bucket, after the backward pass
    GradBucket b; b.name = kFallbackBucket; b.lo = 0; b.hi = arena_n_;
    buckets_.push_back(b);
    return;
  }
  for (int i = 0; i < 6; ++i)
    if (cnt[i]) { GradBucket b; b.name = kBucketNames[i]; b.lo = lo[i]; b.hi = hi[i]; buckets_.push_back(b); }
}

const char* kBucketNamesFwd[6] = {"mlp head", "decoder + embeddings", "encoder + input/loc projections",
                                  "backbone layer4", "backbone layer3", "backbone layer2"};
static const char* const* kBucketNames = kBucketNamesFwd;

template <typename T>
void TrainerT<T>::bucket_ready(int k, cudaStream_t s) {
  if (!comm_) return;
  for (size_t i = 0; i < buckets_.size(); ++i) {
    if (buckets_[i].name != kBucketNames[k]) continue;     // stages without trainable tensors have no bucket
    CUDA_OK(cudaEventRecord(dp_ready_[i], s));
    CUDA_OK(cudaStreamWaitEvent(dp_stream_, dp_ready_[i], 0));
    nccl_ok(nccl().AllReduce(arena_g_ + buckets_[i].lo, arena_g_ + buckets_[i].lo, (size_t)(buckets_[i].hi - buckets_[i].lo),
                             NCCL_FLOAT32, NCCL_AVG, comm_, dp_stream_), "ncclAllReduce");
  }
}

template <typename T>
void TrainerT<T>::buckets_join(cudaStream_t s) {
  if (!comm_) return;
  if (buckets_.size() == 1 && buckets_[0].name == kFallbackBucket) {   // unexpected arena order: one exchange, now
    CUDA_OK(cudaEventRecord(dp_ready_[0], s));
    CUDA_OK(cudaStreamWaitEvent(dp_stream_, dp_ready_[0], 0));
    nccl_ok(nccl().AllReduce(arena_g_, arena_g_, (size_t)arena_n_, NCCL_FLOAT32, NCCL_AVG, comm_, dp_stream_), "ncclAllReduce");
  }
  CUDA_OK(cudaEventRecord(dp_done_, dp_stream_));
  CUDA_OK(cudaStreamWaitEvent(s, dp_done_, 0));
}

// ---- construction ------------------------------------------------------------------------------------------
template <typename T>
TLin<T> TrainerT<T>::lin(const std::string& wname, const std::string& bname, int row0, int rows) {
  const PBind& w = P(wname);
  TLin<T> l;
  l.in = (int)(w.numel / w.shape[0]);
  l.out = rows < 0 ? (int)w.shape[0] : rows;
  l.out_pad = (int)rup(l.out, 64);
  l.w = w.p + (int64_t)row0 * l.in;
  l.gw = w.g ? w.g + (int64_t)row0 * l.in : nullptr;
  if (!bname.empty()) {
    const PBind& b = P(bname);
    l.b = b.p + row0;
    l.gb = b.g ? b.g + row0 : nullptr;
  }
  if (std::is_same<T, float>::value) l.wT = reinterpret_cast<T*>(const_cast<float*>(l.w));
  else l.wT = alloc<T>((int64_t)l.out * l.in);
  l.wtT = alloc<T>((int64_t)l.in * l.out_pad);
  return l;
}

template <typename T>
TConv<T> TrainerT<T>::conv(const std::string& wname, const std::string& bn, int stride, int pad, int dil, bool trainable) {
  const PBind& w = P(wname);
  RETR_CHECK(w.shape.size() == 4, "conv weight must be OIHW");
  TConv<T> c;
  c.w = w.p; c.gw = trainable ? w.g : nullptr;
  c.cout = (int)w.shape[0]; c.cin = (int)w.shape[1]; c.k = (int)w.shape[2];
  c.stride = stride; c.pad = pad; c.dil = dil;
  c.kpad = (int)rup((int64_t)c.k * c.k * c.cin, 16);
  c.wk = alloc<T>((int64_t)c.cout * c.kpad);
  if (trainable) c.wkt = alloc<T>((int64_t)c.kpad * c.cout);
  if (trainable && std::is_same<T, bf16>::value && c.k > 1 && c.stride == 1 && c.cout % 64 == 0 && c.cin % 8 == 0 &&
      c.dil * (c.k - 1) - c.pad >= 0 && !(getenv("RETR_DGRAD_IMPLICIT") && atoi(getenv("RETR_DGRAD_IMPLICIT")) == 0))
    c.wkd = alloc<T>((int64_t)c.cin * c.k * c.k * c.cout);
  // FrozenBatchNorm2d (models/backbone.py:41-51): buffers, never trained -> folded once
  const int n = c.cout;
  std::vector<float> bw(n), bb(n), rm(n), rv(n), sc(n), sh(n);
  CUDA_OK(cudaMemcpy(bw.data(), P(bn + ".weight").p, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(bb.data(), P(bn + ".bias").p, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(rm.data(), P(bn + ".running_mean").p, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(rv.data(), P(bn + ".running_var").p, n * 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < n; ++i) {
    sc[i] = bw[i] * (1.0f / sqrtf(rv[i] + 1e-5f));
    sh[i] = bb[i] - rm[i] * sc[i];
  }
  c.scale = alloc<float>(n);
  c.shift = alloc<float>(n);
  CUDA_OK(cudaMemcpy(c.scale, sc.data(), n * 4, cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(c.shift, sh.data(), n * 4, cudaMemcpyHostToDevice));
  return c;
}

template <typename T>
void TrainerT<T>::finalize() {
  RETR_CHECK(!finalized_, "finalize called twice");
  const int d = desc.d_model, Tn = desc.max_pos, H = desc.n_heads;
  const int64_t B = desc.max_batch, NI = B * desc.n_images, ff = desc.d_ff;
  const std::string bb = "backbone.body.";
  stem_ = conv(bb + "conv1.weight", bb + "bn1", 2, 3, 1, false);
  int dil = 1, Hc = ((desc.image_size + 6 - 7) / 2 + 1 + 2 - 3) / 2 + 1;
  int64_t maxact = 0, max_col = 0, max_tA = 0, max_tB = 0, max_dwk = 0;
  for (int st = 0; st < 4; ++st) {
    int stride = st == 0 ? 1 : 2;
    const int first_dil = dil;
    if (st == 3 && desc.dilation) { dil *= stride; stride = 1; }
    for (int bi = 0; bi < desc.blocks[st]; ++bi) {
      const std::string p = bb + "layer" + std::to_string(st + 1) + "." + std::to_string(bi) + ".";
      const int s_ = bi == 0 ? stride : 1, dl = bi == 0 ? first_dil : dil;
      TBlock<T> b;
      // conv1 + layer1 frozen (models/backbone.py:58-60); a parameter with no bound grad is frozen too
      b.trainable = st >= 1 && P(p + "conv1.weight").g != nullptr;
      b.basic = params_.count(p + "conv3.weight") == 0;
      if (b.basic) {          // torchvision/models/resnet.py:59-105
        RETR_CHECK(dl == 1, "BasicBlock backbones (ResNet-18/34) need dilation = False");
        b.c1 = conv(p + "conv1.weight", p + "bn1", s_, 1, 1, b.trainable);
        b.c2 = conv(p + "conv2.weight", p + "bn2", 1, 1, 1, b.trainable);
      } else {
        b.c1 = conv(p + "conv1.weight", p + "bn1", 1, 0, 1, b.trainable);
        b.c2 = conv(p + "conv2.weight", p + "bn2", s_, dl, dl, b.trainable);
        b.c3 = conv(p + "conv3.weight", p + "bn3", 1, 0, 1, b.trainable);
      }
      b.has_ds = params_.count(p + "downsample.0.weight") > 0;
      if (b.has_ds) b.ds = conv(p + "downsample.0.weight", p + "downsample.1", s_, 0, 1, b.trainable);
      b.Hin = Hc;
      b.Hout = (Hc + 2 * dl - dl * 2 - 1) / s_ + 1;
      Hc = b.Hout;
      if (b.trainable) {
        if (!train_backbone_) first_trainable_ = (int)blocks_.size();
        train_backbone_ = true;
        const int64_t Min = NI * b.Hin * b.Hin, Mout = NI * b.Hout * b.Hout, co = b.out_channels();
        const int64_t M1 = b.basic ? Mout : Min;                     // rows of conv1's output
        b.o1 = alloc<T>(M1 * b.c1.cout);
        if (!b.basic) b.o2 = alloc<T>(Mout * b.c2.cout);
        if (b.has_ds) b.idn = alloc<T>(Mout * co);
        b.y = alloc<T>(Mout * co);
        maxact = std::max({maxact, M1 * b.c1.cout, Mout * co, Min * b.c1.cin});
        max_col = std::max(max_col, rup(Mout, 16) * (b.basic ? std::max(b.c1.kpad, b.c2.kpad) : b.c2.kpad));
        max_tA = std::max({max_tA, rup(Mout, 16) * co, rup(M1, 16) * b.c1.cout});
        max_tB = std::max({max_tB, rup(Min, 16) * b.c1.cin, rup(Mout, 16) * b.c2.kpad, rup(Mout, 16) * (b.basic ? b.c1.kpad : b.c3.cin)});
        max_dwk = std::max({max_dwk, (int64_t)b.c2.cout * b.c2.kpad, b.basic ? (int64_t)b.c1.cout * b.c1.kpad : (int64_t)0});
      }
      blocks_.push_back(b);
    }
  }
  RETR_CHECK(Hc == grid_hw, "backbone output side mismatch");
  const int64_t g2 = (int64_t)grid_hw * grid_hw, cfeat = blocks_.back().out_channels();
  // frozen-part buffers (ping-pong, nothing kept)
  const int64_t h1 = (desc.image_size + 6 - 7) / 2 + 1, hp = (h1 + 2 - 3) / 2 + 1;
  im2col_ = alloc<T>(NI * h1 * h1 * stem_.kpad);
  a0_ = alloc<T>(NI * h1 * h1 * 64);
  a1_ = alloc<T>(NI * hp * hp * 64);
  for (auto& f : fz_) f = alloc<T>(NI * hp * hp * 256);
  if (!blocks_.back().trainable) feat_ = alloc<T>(NI * g2 * cfeat);
  if (train_backbone_) block_in_ = alloc<T>(NI * hp * hp * 256);

  input_proj_ = lin("input_proj.weight", "input_proj.bias");
  if (desc.loc_in_dim) {
    loc_w_ = P("loc_proj.weight").p; loc_b_ = P("loc_proj.bias").p;
    g_loc_w_ = P("loc_proj.weight").g; g_loc_b_ = P("loc_proj.bias").g;
  }
  if (params_.count("transformer.positional_encoding.pe")) {
    pe_ = P("transformer.positional_encoding.pe").p;
  } else {
    const std::string pk = "transformer.positional_encoding.";
    learned_ = true;
    RETR_CHECK(P(pk + "pos_embed.weight").shape[0] >= S, "sequence longer than the learned position table");
    pos_w_ = P(pk + "pos_embed.weight").p; g_pos_w_ = P(pk + "pos_embed.weight").g;
    pos_g_ = P(pk + "LayerNorm.weight").p; g_pos_g_ = P(pk + "LayerNorm.weight").g;
    pos_b_ = P(pk + "LayerNorm.bias").p; g_pos_b_ = P(pk + "LayerNorm.bias").g;
    pos_ln_ = alloc<float>((int64_t)S * d); dposln_ = alloc<float>((int64_t)S * d);
    posD_ = alloc<float>(B * S * d); dPos_ = alloc<float>(B * S * d);
  }
  if (!learned_) RETR_CHECK(P("transformer.positional_encoding.pe").shape[0] >= S, "sequence longer than the sine table");
  const std::string em = "transformer.embeddings.";
  e_word_ = P(em + "word_embeddings.weight").p; g_e_word_ = P(em + "word_embeddings.weight").g;
  e_pos_ = P(em + "position_embeddings.weight").p; g_e_pos_ = P(em + "position_embeddings.weight").g;
  emb_g_ = P(em + "LayerNorm.weight").p; g_emb_g_ = P(em + "LayerNorm.weight").g;
  emb_b_ = P(em + "LayerNorm.bias").p; g_emb_b_ = P(em + "LayerNorm.bias").g;
  const int64_t Me = B * S, Md = B * Tn;
  for (int l = 0; l < desc.enc_layers; ++l) {
    const std::string p = "transformer.encoder.layers." + std::to_string(l) + ".";
    const std::string sa = p + "self_attn.sublayer.";
    TEnc<T> e;
    e.g1 = P(p + "self_attn.norm.weight").p; e.gg1 = P(p + "self_attn.norm.weight").g;
    e.b1 = P(p + "self_attn.norm.bias").p; e.gb1 = P(p + "self_attn.norm.bias").g;
    e.g2 = P(p + "ff.norm.weight").p; e.gg2 = P(p + "ff.norm.weight").g;
    e.b2 = P(p + "ff.norm.bias").p; e.gb2 = P(p + "ff.norm.bias").g;
    e.qk = lin(sa + "in_proj_weight", sa + "in_proj_bias", 0, 2 * d);
    e.v = lin(sa + "in_proj_weight", sa + "in_proj_bias", 2 * d, d);
    e.gin_b = P(sa + "in_proj_bias").g;
    e.o = lin(sa + "out_proj.weight", sa + "out_proj.bias");
    e.ff1 = lin(p + "ff.sublayer.0.weight", p + "ff.sublayer.0.bias");
    e.ff2 = lin(p + "ff.sublayer.2.weight", p + "ff.sublayer.2.bias");
    e.xa = alloc<float>(Me * d); e.xb = alloc<float>(Me * d);
    e.n1 = alloc<T>(Me * d); e.n1p = alloc<T>(Me * d); e.qkv = alloc<T>(Me * 3 * d);
    e.P = flash_ ? nullptr : alloc<T>(B * H * (int64_t)S * S); e.lse = flash_ ? alloc<float>(B * H * (int64_t)S) : nullptr;
    e.att = alloc<T>(Me * d); e.n2 = alloc<T>(Me * d); e.h = alloc<T>(Me * ff);
    enc_.push_back(e);
  }
  if (desc.pre_norm) {        // the final encoder LayerNorm exists only then (models/ConcatTransformer.py:24)
    enc_ng_ = P("transformer.encoder.norm.weight").p; g_enc_ng_ = P("transformer.encoder.norm.weight").g;
    enc_nb_ = P("transformer.encoder.norm.bias").p; g_enc_nb_ = P("transformer.encoder.norm.bias").g;
  }
  for (int l = 0; l < desc.dec_layers; ++l) {
    const std::string p = "transformer.decoder.layers." + std::to_string(l) + ".";
    const std::string sa = p + "tgt_self_attn.sublayer.", ca = p + "tgt_src_cross_attn.sublayer.";
    TDec<T> e;
    e.g1 = P(p + "tgt_self_attn.norm.weight").p; e.gg1 = P(p + "tgt_self_attn.norm.weight").g;
    e.b1 = P(p + "tgt_self_attn.norm.bias").p; e.gb1 = P(p + "tgt_self_attn.norm.bias").g;
    e.g2 = P(p + "tgt_src_cross_attn.norm.weight").p; e.gg2 = P(p + "tgt_src_cross_attn.norm.weight").g;
    e.b2 = P(p + "tgt_src_cross_attn.norm.bias").p; e.gb2 = P(p + "tgt_src_cross_attn.norm.bias").g;
    e.g3 = P(p + "ff.norm.weight").p; e.gg3 = P(p + "ff.norm.weight").g;
    e.b3 = P(p + "ff.norm.bias").p; e.gb3 = P(p + "ff.norm.bias").g;
    e.s_qk = lin(sa + "in_proj_weight", sa + "in_proj_bias", 0, 2 * d);
    e.s_v = lin(sa + "in_proj_weight", sa + "in_proj_bias", 2 * d, d);
    e.gs_b = P(sa + "in_proj_bias").g;
    e.s_o = lin(sa + "out_proj.weight", sa + "out_proj.bias");
    e.c_q = lin(ca + "in_proj_weight", ca + "in_proj_bias", 0, d);
    e.c_k = lin(ca + "in_proj_weight", ca + "in_proj_bias", d, d);
    e.c_v = lin(ca + "in_proj_weight", ca + "in_proj_bias", 2 * d, d);
    e.gc_b = P(ca + "in_proj_bias").g;
    e.c_o = lin(ca + "out_proj.weight", ca + "out_proj.bias");
    e.ff1 = lin(p + "ff.sublayer.0.weight", p + "ff.sublayer.0.bias");
    e.ff2 = lin(p + "ff.sublayer.2.weight", p + "ff.sublayer.2.bias");
    e.ya = alloc<float>(Md * d); e.yb = alloc<float>(Md * d); e.yc = alloc<float>(Md * d);
    e.n1 = alloc<T>(Md * d); e.n1p = alloc<T>(Md * d); e.qkv = alloc<T>(Md * 3 * d);
    e.Ps = flash_ ? nullptr : alloc<T>(B * H * (int64_t)Tn * Tn); e.att_s = alloc<T>(Md * d);
    e.lse_s = flash_ ? alloc<float>(B * H * (int64_t)Tn) : nullptr; e.lse_c = e.lse_s ? alloc<float>(B * H * (int64_t)Tn) : nullptr;
    e.n2p = alloc<T>(Md * d); e.qc = alloc<T>(Md * d); e.ckv = alloc<T>(Me * 2 * d);
    e.Pc = flash_ ? nullptr : alloc<T>(B * H * (int64_t)Tn * S); e.att_c = alloc<T>(Md * d); e.n3 = alloc<T>(Md * d); e.h = alloc<T>(Md * ff);
    dec_.push_back(e);
  }
  dec_ng_ = P("transformer.decoder.norm.weight").p; g_dec_ng_ = P("transformer.decoder.norm.weight").g;
  dec_nb_ = P("transformer.decoder.norm.bias").p; g_dec_nb_ = P("transformer.decoder.norm.bias").g;
  for (int i = 0; i < 3; ++i)
    mlp_[i] = lin("mlp.layers." + std::to_string(i) + ".weight", "mlp.layers." + std::to_string(i) + ".bias");
  RETR_CHECK(mlp_[2].out == desc.vocab, "vocab size mismatch");

  x_final_ = alloc<float>(Me * d); y_final_ = alloc<float>(Md * d); emb_pre_ = alloc<float>(Md * d);
  memory_ = alloc<T>(Me * d); memp_ = alloc<T>(Me * d);
  hs_ = alloc<T>(Md * d); m1_ = alloc<T>(Md * mlp_[0].out); m2_ = alloc<T>(Md * mlp_[1].out);
  dlg_ = alloc<T>(Md * Vp);
  CUDA_OK(cudaMemset(dlg_, 0, Md * Vp * sizeof(T)));
  tok_mask_ = alloc<uint8_t>(Me); id_mask_ = alloc<uint8_t>(Md); ids_ = alloc<int64_t>(Md);
  loc_ = alloc<float>(B * std::max(1, (int)desc.n_loc));
  // backward scratch
  const int64_t Mx = std::max(Me, Md);
  dX_ = alloc<float>(Me * d); dY_ = alloc<float>(Md * d); dMem_ = alloc<float>(Me * d);
  tmpf_ = alloc<float>(Mx * d); tmpf2_ = alloc<float>(Mx * d); embtmp_ = alloc<float>(2 * Md * d);
  gb_ = alloc<T>(Mx * d); dh_ = alloc<T>(Mx * std::max<int64_t>(ff, 512)); datt_ = alloc<T>(Mx * d);
  dqkv_ = alloc<T>(Mx * 3 * d); dckv_ = alloc<T>(Me * 2 * d);
  dS_ = flash_ ? nullptr : alloc<T>(B * H * (int64_t)std::max(S, Tn) * std::max(S, Tn));
  delta_ = alloc<float>(B * H * (int64_t)std::max(S, Tn));
  dtok_ = alloc<T>(NI * g2 * d);
  const int64_t Mdp = rup(Md, 16), Mep = rup(Me, 16), Mtp = rup(NI * g2, 16);
  tA_cap_ = std::max({max_tA, (int64_t)desc.vocab * Mdp, std::max(ff, (int64_t)3 * d) * std::max(Mdp, Mep), (int64_t)d * Mtp});
  tB_cap_ = std::max({max_tB, std::max(ff, (int64_t)512) * std::max(Mdp, Mep), cfeat * Mtp});
  tA_ = alloc<T>(tA_cap_); tB_ = alloc<T>(tB_cap_);
  if (train_backbone_) {
    dcol_ = alloc<T>(max_col);
    for (auto& b : bb_) b = alloc<T>(maxact);
    dwk_ = alloc<float>(max_dwk);
  }
  ws_floats_ = (int64_t)48 << 20;
  ws_ = alloc<float>(ws_floats_);
  sumsq_ = alloc<float>(4);
  CUDA_OK(cudaDeviceSynchronize());
  build_buckets();
  finalized_ = true;
}

// ---- per-step weight refresh (the optimiser updates the fp32 masters in place) -------------------------------------
template <typename T>
void TrainerT<T>::refresh(const TLin<T>& l, cudaStream_t s) {
  if (!std::is_same<T, float>::value) {
    f32_to_bf16(l.w, reinterpret_cast<bf16*>(l.wT), (int64_t)l.out * l.in, s);
  }
  transpose_pad<float, T>(l.w, l.in, l.wtT, l.out_pad, l.out, l.in, s);
}
template <typename T>
void TrainerT<T>::refresh(const TConv<T>& c, cudaStream_t s) {
  repack_filter<T>(c.w, c.wk, c.cout, c.cin, c.k, c.k, c.kpad, s);
  if (c.wkt) transpose_pad<T, T>(c.wk, c.kpad, c.wkt, c.cout, c.cout, c.kpad, s);
  if (c.wkd) repack_filter_dgrad<T>(c.w, c.wkd, c.cout, c.cin, c.k, c.k, s);
}
template <typename T>
void TrainerT<T>::refresh_all(cudaStream_t s) {
  if (!frozen_ready_) refresh(stem_, s);
  for (auto& b : blocks_) {
    if (!b.trainable && frozen_ready_) continue;       // frozen filters are repacked once
    refresh(b.c1, s); refresh(b.c2, s);
    if (!b.basic) refresh(b.c3, s);
    if (b.has_ds) refresh(b.ds, s);
  }
  frozen_ready_ = true;
  refresh(input_proj_, s);
  for (auto& e : enc_) { refresh(e.qk, s); refresh(e.v, s); refresh(e.o, s); refresh(e.ff1, s); refresh(e.ff2, s); }
  for (auto& e : dec_) {
    refresh(e.s_qk, s); refresh(e.s_v, s); refresh(e.s_o, s); refresh(e.c_q, s); refresh(e.c_k, s); refresh(e.c_v, s);
    refresh(e.c_o, s); refresh(e.ff1, s); refresh(e.ff2, s);
  }
  for (auto& l : mlp_) refresh(l, s);
}

// ---- forward ---------------------------------------------------------------------------------------------------------
template <typename T>
void TrainerT<T>::run_conv(const TConv<T>& c, const T* in, int64_t n, int H, int W, T* out, const T* residual, int relu,
                           int* OH, cudaStream_t s) {
  const int oh = (H + 2 * c.pad - c.dil * (c.k - 1) - 1) / c.stride + 1;
  GemmParams p;
  p.A = in; p.W = c.wk; p.C = out;
  p.M = n * oh * oh; p.N = c.cout; p.K = c.kpad; p.lda = c.cin; p.ldc = c.cout;
  p.scale = c.scale; p.bias = c.shift; p.relu = relu;
  p.residual = residual; p.ldr = c.cout; p.res_fp32 = std::is_same<T, float>::value;
  if (c.k != 1 || c.stride != 1) {
    p.conv.kh = p.conv.kw = c.k; p.conv.stride = c.stride; p.conv.pad = c.pad; p.conv.dil = c.dil;
    p.conv.H = H; p.conv.W = W; p.conv.C = c.cin; p.conv.OH = oh; p.conv.OW = oh;
  }
  gemmT<T>(p, s);
  *OH = oh;
}

template <typename T>
void TrainerT<T>::branch_out(const TLin<T>& l, const T* A, int64_t M, const float* x_in, float* x_out, DropCfg dc,
                             cudaStream_t s) {
  GemmParams p = lp(l, A, l.in, M, x_out, l.out);
  p.out_fp32 = true;
  if (dc.p > 0.f) {
    p.C = tmpf_;
    gemmT<T>(p, s);
    dropout_add(x_in, tmpf_, x_out, M * l.out, dc, s);
  } else {
    p.residual = x_in; p.ldr = l.out; p.res_fp32 = true;
    gemmT<T>(p, s);
  }
}

template <typename T>
void TrainerT<T>::attn_fwd(const T* q, int64_t ldq, const T* k, const T* v, int64_t ldkv, T* out, T* probs, float* lse,
                           int Lq, int Lk, const uint8_t* pad, int causal, unsigned site, cudaStream_t s) {
  AttnTrainParams a{};
  a.q = q; a.k = k; a.v = v; a.ldq = ldq; a.ldk = a.ldv = ldkv; a.out = out; a.ldo = desc.d_model; a.probs = probs; a.lse = lse;
  a.B = cur_B_; a.Lq = Lq; a.Lk = Lk; a.H = desc.n_heads; a.dh = desc.d_model / desc.n_heads;
  a.key_pad = pad; a.causal = causal; a.drop = drop(site);
  attention_train_fwd<T>(a, s);
}
template <typename T>
void TrainerT<T>::attn_bwd(const T* q, int64_t ldq, const T* k, const T* v, int64_t ldkv, const T* probs, const float* lse,
                           const T* out, const T* dout, T* dq, int64_t lddq, T* dk, T* dv, int64_t lddkv, int Lq, int Lk, const uint8_t* pad,
                           int causal, unsigned site, cudaStream_t s) {
  AttnBwdParams b{};
  b.f.q = q; b.f.k = k; b.f.v = v; b.f.ldq = ldq; b.f.ldk = b.f.ldv = ldkv; b.f.probs = const_cast<T*>(probs);
  b.f.lse = const_cast<float*>(lse); b.f.out = const_cast<T*>(out); b.f.ldo = desc.d_model; b.delta = delta_;
  b.f.B = cur_B_; b.f.Lq = Lq; b.f.Lk = Lk; b.f.H = desc.n_heads; b.f.dh = desc.d_model / desc.n_heads;
  b.f.key_pad = pad; b.f.causal = causal; b.f.drop = drop(site);
  b.dout = dout; b.lddo = desc.d_model; b.ds = dS_;
  b.dq = dq; b.dk = dk; b.dv = dv; b.lddq = lddq; b.lddk = b.lddv = lddkv;
  attention_train_bwd<T>(b, s);
}

template <typename T>
void TrainerT<T>::forward_impl(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                               const uint8_t* g_mask, const float* loc, const int64_t* ids, const uint8_t* id_mask,
                               float* logits_out, bool logits_T, float dropout_p, unsigned long long seed,
                               cudaStream_t s) {
  RETR_CHECK(finalized_, "call retr_train_finalize first");
  RETR_CHECK(B > 0 && B <= desc.max_batch, "batch exceeds max_batch");
  RETR_CHECK(t_img && t_mask && ids && id_mask, "null input");
  RETR_CHECK(desc.n_images == 1 || (g_img && g_mask), "context images and masks required");
  RETR_CHECK(desc.loc_in_dim == 0 || loc, "location features required");
  RETR_CHECK(dropout_p >= 0.f && dropout_p < 1.f, "dropout probability");
  const long long l0 = g_launches;
  cur_B_ = B; drop_p_ = dropout_p; seed_ = seed;
  const int d = desc.d_model, Tn = desc.max_pos, Hi = desc.image_size;
  const int64_t NI = (int64_t)B * desc.n_images, g2 = (int64_t)grid_hw * grid_hw, img_elems = 3LL * Hi * Hi;
  const int64_t Me = (int64_t)B * S, Md = (int64_t)B * Tn;
  refresh_all(s);
  CUDA_OK(cudaMemcpyAsync(ids_, ids, Md * 8, cudaMemcpyDeviceToDevice, s));
  CUDA_OK(cudaMemcpyAsync(id_mask_, id_mask, Md, cudaMemcpyDeviceToDevice, s));
  if (loc) CUDA_OK(cudaMemcpyAsync(loc_, loc, (int64_t)B * desc.n_loc * 4, cudaMemcpyDeviceToDevice, s));

  // ---- backbone over all n_images*B crops at once (target crops first) ----
  int Hc = (Hi + 6 - 7) / 2 + 1;
  for (int set = 0; set < desc.n_images; ++set)
    stem_im2col<T>(set == 0 ? t_img : g_img, im2col_ + (int64_t)set * B * Hc * Hc * stem_.kpad, B, Hi, Hi, Hc, Hc,
                   stem_.kpad, s);
  (void)img_elems;
  {
    GemmParams p;
    p.A = im2col_; p.W = stem_.wk; p.C = a0_;
    p.M = NI * Hc * Hc; p.N = stem_.cout; p.K = stem_.kpad; p.lda = stem_.kpad; p.ldc = stem_.cout;
    p.scale = stem_.scale; p.bias = stem_.shift; p.relu = 1;
    gemmT<T>(p, s);
  }
  const int PH = (Hc + 2 - 3) / 2 + 1;
  maxpool_3x3s2<T>(a0_, a1_, NI, Hc, Hc, stem_.cout, PH, PH, s);
  Hc = PH;
  const T* x = a1_;
  int flip = 0;
  for (size_t bi = 0; bi < blocks_.size(); ++bi) {
    TBlock<T>& b = blocks_[bi];
    int h1, h2, h3, hd;
    T* o1 = b.trainable ? b.o1 : fz_[2];
    T* o2 = b.trainable ? b.o2 : fz_[3];
    T* idn = b.trainable ? b.idn : fz_[4];
    T* y = b.trainable ? b.y : ((bi + 1 == blocks_.size()) ? feat_ : fz_[flip]);
    if (!b.trainable && train_backbone_ && (int)bi + 1 == first_trainable_) y = block_in_;
    run_conv(b.c1, x, NI, Hc, Hc, o1, nullptr, 1, &h1, s);
    const T* identity = x;
    if (b.has_ds) {
      run_conv(b.ds, x, NI, Hc, Hc, idn, nullptr, 0, &hd, s);
      identity = idn;
    }
    if (b.basic) {
      run_conv(b.c2, o1, NI, h1, h1, y, identity, 2, &h3, s);
    } else {
      run_conv(b.c2, o1, NI, h1, h1, o2, nullptr, 1, &h2, s);
      run_conv(b.c3, o2, NI, h2, h2, y, identity, 2, &h3, s);
    }
    x = y;
    flip ^= 1;
    Hc = h3;
  }
  const T* feat = x;
  if (blocks_.back().trainable) feat_ = blocks_.back().y;

  // ---- tokens (models/caption.py:119-147) ----
  const int64_t cfeat = blocks_.back().out_channels();
  const int off_loc = (int)g2, off_g = (int)g2 + n_loc_tok;
  float* x0 = enc_[0].xa;
  for (int set = 0; set < desc.n_images; ++set) {
    GemmParams p = lp(input_proj_, feat + (int64_t)set * B * g2 * cfeat, cfeat, (int64_t)B * g2,
                      x0 + (int64_t)(set ? off_g : 0) * d, d);
    p.out_fp32 = true; p.c_group = (int)g2; p.c_group_stride = S;
    gemmT<T>(p, s);
    subsample_mask(set == 0 ? t_mask : g_mask, tok_mask_, B, Hi, Hi, grid_hw, grid_hw, S, set ? off_g : 0, s);
  }
  if (desc.loc_in_dim)
    loc_tokens(loc_, loc_w_, loc_b_, x0, tok_mask_, B, desc.n_loc, desc.loc_in_dim == 1, d, S, off_loc, s);

  if (learned_) {
    layernorm<float>(pos_w_, pos_g_, pos_b_, pos_ln_, S, d, 1e-5f, s);
    pos_expand_kernel<<<ewb(Me * d), 256, 0, s>>>(pos_ln_, posD_, Me * d, (int64_t)S * d, pos_drop());
    LAUNCH_OK();
  }
  // ---- encoder (models/ConcatTransformer.py:171-194, transformer_modules.py:22-46,77-97) ----
  for (int l = 0; l < desc.enc_layers; ++l) {
    TEnc<T>& e = enc_[l];
    float* x_next = l + 1 < desc.enc_layers ? enc_[l + 1].xa : x_final_;
    layernorm_pos<T>(e.xa, e.g1, e.b1, e.n1, e.n1p, pos_ptr(), pos_period(Me), Me, d, 1e-5f, s);
    gemmT<T>(lp(e.qk, e.n1p, d, Me, e.qkv, 3 * d), s);
    gemmT<T>(lp(e.v, e.n1, d, Me, e.qkv + 2 * d, 3 * d), s);
    attn_fwd(e.qkv, 3 * d, e.qkv + d, e.qkv + 2 * d, 3 * d, e.att, e.P, e.lse, S, S, tok_mask_, 0, 10 + 4 * l, s);
    branch_out(e.o, e.att, Me, e.xa, e.xb, drop(11 + 4 * l), s);
    layernorm_pos<T>(e.xb, e.g2, e.b2, e.n2, nullptr, nullptr, 1, Me, d, 1e-5f, s);
    GemmParams p = lp(e.ff1, e.n2, d, Me, e.h, e.ff1.out);
    p.relu = 1;
    gemmT<T>(p, s);
    branch_out(e.ff2, e.h, Me, e.xb, x_next, drop(12 + 4 * l), s);
  }
  if (desc.pre_norm) {
    layernorm_pos<T>(x_final_, enc_ng_, enc_nb_, memory_, memp_, pos_ptr(), pos_period(Me), Me, d, 1e-5f, s);
  } else {                    // pre_norm=False: memory is the last layer's output as it is
    cast_pos_kernel<T><<<ewb(Me * d), 256, 0, s>>>(x_final_, memory_, memp_, pos_ptr(), pos_period(Me), Me * d, d);
    LAUNCH_OK();
  }

  // ---- decoder (models/ConcatTransformer.py:63-70,197-257) ----
  embed_ln_train(ids_, e_word_, e_pos_, emb_g_, emb_b_, emb_pre_, dec_[0].ya, B, Tn, d, desc.embed_ln_eps, drop(1), s);
  for (int l = 0; l < desc.dec_layers; ++l) {
    TDec<T>& e = dec_[l];
    float* y_next = l + 1 < desc.dec_layers ? dec_[l + 1].ya : y_final_;
    layernorm_pos<T>(e.ya, e.g1, e.b1, e.n1, e.n1p, e_pos_, Tn, Md, d, 1e-5f, s);
    gemmT<T>(lp(e.s_qk, e.n1p, d, Md, e.qkv, 3 * d), s);
    gemmT<T>(lp(e.s_v, e.n1, d, Md, e.qkv + 2 * d, 3 * d), s);
    attn_fwd(e.qkv, 3 * d, e.qkv + d, e.qkv + 2 * d, 3 * d, e.att_s, e.Ps, e.lse_s, Tn, Tn, id_mask_, 1, 100 + 8 * l, s);
    branch_out(e.s_o, e.att_s, Md, e.ya, e.yb, drop(101 + 8 * l), s);
    layernorm_pos<T>(e.yb, e.g2, e.b2, nullptr, e.n2p, e_pos_, Tn, Md, d, 1e-5f, s);
    gemmT<T>(lp(e.c_q, e.n2p, d, Md, e.qc, d), s);
    gemmT<T>(lp(e.c_k, memp_, d, Me, e.ckv, 2 * d), s);
    gemmT<T>(lp(e.c_v, memory_, d, Me, e.ckv + d, 2 * d), s);
    attn_fwd(e.qc, d, e.ckv, e.ckv + d, 2 * d, e.att_c, e.Pc, e.lse_c, Tn, S, tok_mask_, 0, 102 + 8 * l, s);
    branch_out(e.c_o, e.att_c, Md, e.yb, e.yc, drop(103 + 8 * l), s);
    layernorm_pos<T>(e.yc, e.g3, e.b3, e.n3, nullptr, nullptr, 1, Md, d, 1e-5f, s);
    GemmParams p = lp(e.ff1, e.n3, d, Md, e.h, e.ff1.out);
    p.relu = 1;
    gemmT<T>(p, s);
    branch_out(e.ff2, e.h, Md, e.yc, y_next, drop(104 + 8 * l), s);
  }
  layernorm_pos<T>(y_final_, dec_ng_, dec_nb_, hs_, nullptr, nullptr, 1, Md, d, 1e-5f, s);
  // ---- MLP head (models/caption.py:161-174) ----
  GemmParams p = lp(mlp_[0], hs_, d, Md, m1_, mlp_[0].out); p.relu = 1; gemmT<T>(p, s);
  p = lp(mlp_[1], m1_, mlp_[0].out, Md, m2_, mlp_[1].out); p.relu = 1; gemmT<T>(p, s);
  if (logits_T) {
    p = lp(mlp_[2], m2_, mlp_[1].out, Md, dlg_, Vp);
    gemmT<T>(p, s);
  }
  if (logits_out) {
    p = lp(mlp_[2], m2_, mlp_[1].out, Md, logits_out, desc.vocab);
    p.out_fp32 = true;
    gemmT<T>(p, s);
  }
  launches += g_launches - l0;
}

template <typename T>
void TrainerT<T>::forward(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                          const float* loc, const int64_t* ids, const uint8_t* id_mask, float* logits_out, float dropout_p,
                          unsigned long long seed, cudaStream_t s) {
  forward_impl(B, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, logits_out, false, dropout_p, seed, s);
}

// RETR_WGRAD_TN=0: weight gradients through transposed copies + the K-major GEMM (the first implementation)
static bool wgrad_in_place() {
  static const bool on = !(getenv("RETR_WGRAD_TN") && atoi(getenv("RETR_WGRAD_TN")) == 0);
  return on;
}

// ---- backward helpers ---------------------------------------------------------------------------------------------
template <typename T>
void TrainerT<T>::wgrad(const T* dy, int64_t ld_dy, int n_out, const T* x, int64_t ld_x, int k_in, int64_t M, float* gw,
                        bool accumulate, cudaStream_t s) {
  if (std::is_same<T, bf16>::value && wgrad_in_place() && ld_dy % 8 == 0 && ld_x % 8 == 0 &&
      reinterpret_cast<uintptr_t>(dy) % 16 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0) {
    // dW = dY^T . X with both operands read in place as MN-major tcgen05 tiles (no transposed copies)
    GemmParams p;
    p.mn_major = true;
    p.A = dy; p.lda = ld_dy; p.W = x; p.ldb = ld_x; p.M = n_out; p.N = k_in; p.K = (int)M; p.C = gw; p.ldc = k_in;
    p.out_fp32 = true; p.split_k = 0; p.split_ws = ws_; p.split_ws_floats = ws_floats_; p.accumulate = accumulate;
    gemmT<T>(p, s);
    return;
  }
  const int64_t Mp = rup(M, 16);
  RETR_CHECK((int64_t)n_out * Mp <= tA_cap_ && (int64_t)k_in * Mp <= tB_cap_, "wgrad scratch too small");
  transpose_pad<T, T>(dy, ld_dy, tA_, Mp, M, n_out, s);
  transpose_pad<T, T>(x, ld_x, tB_, Mp, M, k_in, s);
  GemmParams p;
  p.A = tA_; p.lda = Mp; p.W = tB_; p.M = n_out; p.N = k_in; p.K = (int)Mp; p.C = gw; p.ldc = k_in; p.out_fp32 = true;
  if (std::is_same<T, bf16>::value) {
    p.split_k = 0; p.split_ws = ws_; p.split_ws_floats = ws_floats_; p.accumulate = accumulate;
  } else if (accumulate) {
    p.residual = gw; p.ldr = k_in; p.res_fp32 = true;
  }
  gemmT<T>(p, s);
}

template <typename T>
void TrainerT<T>::lin_bwd(const TLin<T>& l, const T* dy, int64_t ld_dy, const T* x, int64_t ld_x, int64_t M, void* dx,
                          bool dx_fp32, float* bias_grad, cudaStream_t s) {
  if (l.gw) wgrad(dy, ld_dy, l.out, x, ld_x, l.in, M, l.gw, false, s);
  if (bias_grad) colsum<T>(dy, ld_dy, bias_grad, M, l.out, 0, s);
  if (dx) {
    GemmParams p;
    p.A = dy; p.lda = ld_dy; p.W = l.wtT; p.M = M; p.N = l.in; p.K = l.out_pad; p.C = dx; p.ldc = l.in;
    p.out_fp32 = dx_fp32;
    gemmT<T>(p, s);
  }
}

// g = dL/d(conv output) with the FrozenBN scale already applied, [nimg*OH*OH, cout]; x = conv input NHWC
template <typename T>
bool TrainerT<T>::conv_bwd(const TConv<T>& c, const T* g, const T* x, int64_t nimg, int H, int OH, T* dx, const T* dx_add,
                           bool need_dx, cudaStream_t s, const T* relu_out, const float* relu_scale) {
  static const bool fuse_env = getenv("RETR_FUSE_RELU_BWD") && atoi(getenv("RETR_FUSE_RELU_BWD")) != 0;   // off until verified
  const bool can_mask = fuse_env && std::is_same<T, bf16>::value && relu_out && !dx_add && c.cin % 64 == 0;
  bool fused = false;
  const int64_t Mo = nimg * OH * OH, Mp = rup(Mo, 16);
  const int kk = c.k * c.k * c.cin;
  const bool plain = c.k == 1 && c.stride == 1;
  const bool in_place = std::is_same<T, bf16>::value && wgrad_in_place() && c.cout % 8 == 0 && c.cin % 64 == 0 &&
                        c.pad <= 127 && c.stride <= 8;
  if (c.gw && in_place) {
    // dW[C_out, (kh,kw,C_in)] = dY^T . im2col(X): dY and the NHWC input read in place (TMA im2col as the B operand)
    GemmParams p;
    p.mn_major = true;
    p.A = g; p.lda = c.cout; p.W = x; p.ldb = c.cin; p.M = c.cout; p.N = kk; p.K = (int)Mo; p.ldc = kk; p.out_fp32 = true;
    p.C = c.k == 1 ? c.gw : dwk_;
    if (!plain) {
      p.conv.kh = p.conv.kw = c.k; p.conv.stride = c.stride; p.conv.pad = c.pad; p.conv.dil = c.dil;
      p.conv.H = H; p.conv.W = H; p.conv.C = c.cin; p.conv.OH = OH; p.conv.OW = OH;
    }
    p.split_k = 0; p.split_ws = ws_; p.split_ws_floats = ws_floats_;
    gemmT<T>(p, s);
    if (c.k != 1) filter_grad_unpack(dwk_, c.gw, c.cout, c.cin, c.k, c.k, kk, s);
  } else if (c.gw) {
    RETR_CHECK((int64_t)c.cout * Mp <= tA_cap_ && (int64_t)kk * Mp <= tB_cap_, "conv wgrad scratch too small");
    transpose_pad<T, T>(g, c.cout, tA_, Mp, Mo, c.cout, s);
    if (plain) transpose_pad<T, T>(x, c.cin, tB_, Mp, Mo, c.cin, s);
    else im2col_T<T>(x, tB_, Mp, nimg, H, H, c.cin, c.k, c.stride, c.pad, c.dil, OH, OH, s);
    GemmParams p;
    p.A = tA_; p.lda = Mp; p.W = tB_; p.M = c.cout; p.N = kk; p.K = (int)Mp; p.ldc = kk; p.out_fp32 = true;
    p.C = c.k == 1 ? c.gw : dwk_;
    if (std::is_same<T, bf16>::value) { p.split_k = 0; p.split_ws = ws_; p.split_ws_floats = ws_floats_; }
    gemmT<T>(p, s);
    if (c.k != 1) filter_grad_unpack(dwk_, c.gw, c.cout, c.cin, c.k, c.k, kk, s);
  }
  if (need_dx && c.wkd) {
    // input gradient of a stride-1 k x k convolution = convolution of dY with the flipped, transposed filter (TMA
    // im2col over dY, pad' = dil*(k-1) - pad); the other branch's gradient rides in as the epilogue residual.  No
    // [pixels, k*k*C_in] column matrix, no col2im.
    GemmParams p;
    p.A = g; p.lda = c.cout; p.W = c.wkd; p.C = dx; p.M = nimg * H * H; p.N = c.cin; p.K = c.k * c.k * c.cout; p.ldc = c.cin;
    p.conv.kh = p.conv.kw = c.k; p.conv.stride = 1; p.conv.pad = c.dil * (c.k - 1) - c.pad; p.conv.dil = c.dil;
    p.conv.H = OH; p.conv.W = OH; p.conv.C = c.cout; p.conv.OH = H; p.conv.OW = H;
    if (dx_add) { p.residual = dx_add; p.ldr = c.cin; p.res_fp32 = false; }
    if (can_mask) { p.scale = relu_scale; p.residual = relu_out; p.ldr = c.cin; p.res_fp32 = false; p.relu = 3; fused = true; }
    gemmT<T>(p, s);
  } else if (need_dx) {
    GemmParams p;
    p.A = g; p.lda = c.cout; p.W = c.wkt; p.M = Mo; p.N = kk; p.K = c.cout; p.ldc = kk;
    if (plain) {
      p.C = dx;
      if (dx_add) { p.residual = dx_add; p.ldr = kk; p.res_fp32 = std::is_same<T, float>::value; }
      if (can_mask) { p.scale = relu_scale; p.residual = relu_out; p.ldr = kk; p.res_fp32 = false; p.relu = 3; fused = true; }
      gemmT<T>(p, s);
    } else {
      p.C = dcol_;
      gemmT<T>(p, s);
      col2im<T>(dcol_, dx, nimg, H, H, c.cin, c.k, c.stride, c.pad, c.dil, OH, OH, dx_add, s);
    }
  }
  return fused;
}

// ---- backward -------------------------------------------------------------------------------------------------------
template <typename T>
void TrainerT<T>::backward(const float* dlogits, cudaStream_t s) {
  RETR_CHECK(cur_B_ > 0, "forward must run before backward");
  RETR_CHECK(dlogits, "null dlogits");
  const int64_t Md = (int64_t)cur_B_ * desc.max_pos;
  cast_rows_kernel<<<ewb(Md * desc.vocab), 256, 0, s>>>(dlogits, desc.vocab, dlg_, Vp, Md, desc.vocab);
  LAUNCH_OK();
  backward_impl(s);
}

template <typename T>
void TrainerT<T>::loss_and_grads(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                                 const uint8_t* g_mask, const float* loc, const int64_t* ids, const uint8_t* id_mask,
                                 const int64_t* targets, float dropout_p, unsigned long long seed, float* loss_dev,
                                 cudaStream_t s) {
  RETR_CHECK(targets && loss_dev, "null argument");
  forward_impl(B, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, nullptr, true, dropout_p, seed, s);
  const int64_t Md = (int64_t)B * desc.max_pos;
  CUDA_OK(cudaMemsetAsync(loss_dev, 0, 4, s));
  // dL/dlogits of the MEAN over all B*T positions; loss_dev accumulates the SUM (caller divides)
  ce_loss_fwd_bwd<T>(dlg_, Vp, targets, Md, desc.vocab, 1.f / (float)Md, loss_dev, s);
  backward_impl(s);
}

template <typename T>
void TrainerT<T>::backward_impl(cudaStream_t s) {
  const long long l0 = g_launches;
  const int B = cur_B_, d = desc.d_model, Tn = desc.max_pos;
  const int64_t Me = (int64_t)B * S, Md = (int64_t)B * Tn, ff = desc.d_ff;
  const int64_t NI = (int64_t)B * desc.n_images, g2 = (int64_t)grid_hw * grid_hw;
  if (arena_g_) {
    CUDA_OK(cudaMemsetAsync(arena_g_, 0, arena_n_ * 4, s));
  } else {
    for (auto& kv : params_)
      if (kv.second.g) CUDA_OK(cudaMemsetAsync(kv.second.g, 0, kv.second.numel * 4, s));
  }

  if (learned_) CUDA_OK(cudaMemsetAsync(dPos_, 0, Me * d * 4, s));
  // ---- head ----
  lin_bwd(mlp_[2], dlg_, Vp, m2_, mlp_[1].out, Md, dh_, false, mlp_[2].gb, s);          // dh_ = d m2 [Md,512]
  relu_bwd<T>(dh_, m2_, Md * mlp_[1].out, s);
  lin_bwd(mlp_[1], dh_, mlp_[1].out, m1_, mlp_[0].out, Md, dqkv_, false, mlp_[1].gb, s);   // dqkv_ = d m1 [Md,512]
  relu_bwd<T>(dqkv_, m1_, Md * mlp_[0].out, s);
  lin_bwd(mlp_[0], dqkv_, mlp_[0].out, hs_, d, Md, tmpf_, true, mlp_[0].gb, s);           // tmpf_ = d hs (fp32)
  layernorm_bwd(y_final_, dec_ng_, tmpf_, nullptr, dY_, 0, g_dec_ng_, g_dec_nb_, Md, d, 1e-5f, s);
  CUDA_OK(cudaMemsetAsync(dMem_, 0, Me * d * 4, s));
  bucket_ready(0, s);                                      // the 16 M head gradients travel under the whole rest of the pass

  // ---- decoder layers ----
  for (int l = desc.dec_layers - 1; l >= 0; --l) {
    TDec<T>& e = dec_[l];
    // FFN: y_next = yc + drop(W2 relu(W1 LN3(yc)))
    dropout_bwd_cast<T>(dY_, gb_, Md * d, drop(104 + 8 * l), s);
    lin_bwd(e.ff2, gb_, d, e.h, ff, Md, dh_, false, e.ff2.gb, s);
    relu_bwd<T>(dh_, e.h, Md * ff, s);
    lin_bwd(e.ff1, dh_, ff, e.n3, d, Md, tmpf_, true, e.ff1.gb, s);
    layernorm_bwd(e.yc, e.g3, tmpf_, nullptr, dY_, 1, e.gg3, e.gb3, Md, d, 1e-5f, s);
    // cross attention: yc = yb + drop(Wo att(q = Wq(LN2(yb)+E_pos), k = Wk(mem+pos), v = Wv mem))
    dropout_bwd_cast<T>(dY_, gb_, Md * d, drop(103 + 8 * l), s);
    lin_bwd(e.c_o, gb_, d, e.att_c, d, Md, datt_, false, e.c_o.gb, s);
    attn_bwd(e.qc, d, e.ckv, e.ckv + d, 2 * d, e.Pc, e.lse_c, e.att_c, datt_, dqkv_, d, dckv_, dckv_ + d, 2 * d, Tn, S, tok_mask_, 0,
             102 + 8 * l, s);
    lin_bwd(e.c_q, dqkv_, d, e.n2p, d, Md, tmpf_, true, e.gc_b, s);
    if (g_e_pos_) pos_grad_accumulate(tmpf_, g_e_pos_, B, Tn, d, s);
    layernorm_bwd(e.yb, e.g2, tmpf_, nullptr, dY_, 1, e.gg2, e.gb2, Md, d, 1e-5f, s);
    {
      // K / V projections of the memory: weight grads, bias grads, and dMem += dK Wk + dV Wv
      if (e.c_k.gw) wgrad(dckv_, 2 * d, d, memp_, d, d, Me, e.c_k.gw, false, s);
      if (e.c_v.gw) wgrad(dckv_ + d, 2 * d, d, memory_, d, d, Me, e.c_v.gw, false, s);
      if (e.gc_b) colsum<T>(dckv_, 2 * d, e.gc_b + d, Me, 2 * d, 0, s);
      GemmParams p;
      p.A = dckv_; p.lda = 2 * d; p.W = e.c_k.wtT; p.M = Me; p.N = d; p.K = e.c_k.out_pad; p.C = dMem_; p.ldc = d;
      p.out_fp32 = true; p.residual = dMem_; p.ldr = d; p.res_fp32 = true;
      gemmT<T>(p, s);
      if (learned_) {                                       // k = memory + pos: the same term is the gradient of the positions
        GemmParams q = p;
        q.C = dPos_; q.residual = dPos_;
        gemmT<T>(q, s);
      }
      p.A = dckv_ + d; p.W = e.c_v.wtT; p.K = e.c_v.out_pad;
      gemmT<T>(p, s);
    }
    // self attention: yb = ya + drop(Wo att(q = k = LN1(ya)+E_pos, v = LN1(ya)))
    dropout_bwd_cast<T>(dY_, gb_, Md * d, drop(101 + 8 * l), s);
    lin_bwd(e.s_o, gb_, d, e.att_s, d, Md, datt_, false, e.s_o.gb, s);
    attn_bwd(e.qkv, 3 * d, e.qkv + d, e.qkv + 2 * d, 3 * d, e.Ps, e.lse_s, e.att_s, datt_, dqkv_, 3 * d, dqkv_ + d, dqkv_ + 2 * d, 3 * d, Tn,
             Tn, id_mask_, 1, 100 + 8 * l, s);
    if (e.gs_b) colsum<T>(dqkv_, 3 * d, e.gs_b, Md, 3 * d, 0, s);
    lin_bwd(e.s_qk, dqkv_, 3 * d, e.n1p, d, Md, tmpf_, true, nullptr, s);
    lin_bwd(e.s_v, dqkv_ + 2 * d, 3 * d, e.n1, d, Md, tmpf2_, true, nullptr, s);
    if (g_e_pos_) pos_grad_accumulate(tmpf_, g_e_pos_, B, Tn, d, s);
    layernorm_bwd(e.ya, e.g1, tmpf_, tmpf2_, dY_, 1, e.gg1, e.gb1, Md, d, 1e-5f, s);
  }
  if (g_e_word_ && g_e_pos_)
    embed_ln_bwd(ids_, emb_pre_, emb_g_, dY_, g_e_word_, g_e_pos_, g_emb_g_, g_emb_b_, embtmp_, B, Tn, d, desc.embed_ln_eps,
                 drop(1), pad_id_, s);
  // (decoder.norm's gradients were written before bucket 0 was recorded but belong to this bucket's range: fine,
  // every bucket is recorded AFTER its last writer)
  bucket_ready(1, s);

  // ---- encoder ----
  if (desc.pre_norm) layernorm_bwd(x_final_, enc_ng_, dMem_, nullptr, dX_, 0, g_enc_ng_, g_enc_nb_, Me, d, 1e-5f, s);
  else CUDA_OK(cudaMemcpyAsync(dX_, dMem_, Me * d * 4, cudaMemcpyDeviceToDevice, s));
  for (int l = desc.enc_layers - 1; l >= 0; --l) {
    TEnc<T>& e = enc_[l];
    dropout_bwd_cast<T>(dX_, gb_, Me * d, drop(12 + 4 * l), s);
    lin_bwd(e.ff2, gb_, d, e.h, ff, Me, dh_, false, e.ff2.gb, s);
    relu_bwd<T>(dh_, e.h, Me * ff, s);
    lin_bwd(e.ff1, dh_, ff, e.n2, d, Me, tmpf_, true, e.ff1.gb, s);
    layernorm_bwd(e.xb, e.g2, tmpf_, nullptr, dX_, 1, e.gg2, e.gb2, Me, d, 1e-5f, s);
    dropout_bwd_cast<T>(dX_, gb_, Me * d, drop(11 + 4 * l), s);
    lin_bwd(e.o, gb_, d, e.att, d, Me, datt_, false, e.o.gb, s);
    attn_bwd(e.qkv, 3 * d, e.qkv + d, e.qkv + 2 * d, 3 * d, e.P, e.lse, e.att, datt_, dqkv_, 3 * d, dqkv_ + d, dqkv_ + 2 * d, 3 * d, S, S,
             tok_mask_, 0, 10 + 4 * l, s);
    if (e.gin_b) colsum<T>(dqkv_, 3 * d, e.gin_b, Me, 3 * d, 0, s);
    lin_bwd(e.qk, dqkv_, 3 * d, e.n1p, d, Me, tmpf_, true, nullptr, s);
    lin_bwd(e.v, dqkv_ + 2 * d, 3 * d, e.n1, d, Me, tmpf2_, true, nullptr, s);
    if (learned_) { axpy_kernel<<<ewb(Me * d), 256, 0, s>>>(dPos_, tmpf_, Me * d); LAUNCH_OK(); }
    layernorm_bwd(e.xa, e.g1, tmpf_, tmpf2_, dX_, 1, e.gg1, e.gb1, Me, d, 1e-5f, s);
  }
  if (learned_ && g_pos_w_) {
    // sum the masked gradients over the batch, then back through LayerNorm(pos_embed.weight[:S])
    pos_reduce_kernel<<<ewb((int64_t)S * d), 256, 0, s>>>(dPos_, dposln_, B, (int64_t)S * d, pos_drop());
    LAUNCH_OK();
    layernorm_bwd(pos_w_, pos_g_, dposln_, nullptr, g_pos_w_, 0, g_pos_g_, g_pos_b_, S, d, 1e-5f, s);
  }

  // ---- tokens: location projection, input_proj ----
  const int off_loc = (int)g2, off_g = (int)g2 + n_loc_tok;
  if (desc.loc_in_dim && g_loc_w_)
    loc_tokens_bwd(dX_, loc_, g_loc_w_, g_loc_b_, B, desc.n_loc, desc.loc_in_dim == 1, d, S, off_loc, s);
  gather_tok_kernel<T><<<ewb(NI * g2 * d), 256, 0, s>>>(dX_, dtok_, B, (int)g2, S, off_g, desc.n_images, d);
  LAUNCH_OK();
  const int64_t cfeat = blocks_.back().out_channels(), Mt = NI * g2;
  T* dfeat = train_backbone_ ? bb_[0] : nullptr;
  lin_bwd(input_proj_, dtok_, d, feat_, cfeat, Mt, dfeat, false, input_proj_.gb, s);
  bucket_ready(2, s);

  // ---- backbone layers 2-4 (torchvision Bottleneck, resnet.py:143-165) ----
  if (train_backbone_) {
    T* dy = bb_[0];
    T* dxn = bb_[1];
    for (int bi = (int)blocks_.size() - 1; bi >= first_trainable_; --bi) {
      TBlock<T>& b = blocks_[bi];
      const bool need_dx = bi > first_trainable_;
      const T* x_in = bi == first_trainable_ ? block_in_ : blocks_[bi - 1].y;
      const int64_t Mo = NI * b.Hout * b.Hout, Mi = NI * b.Hin * b.Hin;
      T *g3s = bb_[2], *g3 = bb_[3], *t2 = bb_[4], *t1 = bb_[5], *gid = bb_[6];
      const TConv<T>& last = b.basic ? b.c2 : b.c3;
      relu_bn_bwd<T>(dy, b.y, last.scale, g3s, g3, Mo, last.cout, s);
      if (b.basic) {
        if (!conv_bwd(b.c2, g3s, b.o1, NI, b.Hout, b.Hout, t1, nullptr, true, s, b.o1, b.c1.scale))
          relu_bn_bwd<T>(t1, b.o1, b.c1.scale, t1, nullptr, Mo, b.c1.cout, s);
      } else {
        if (!conv_bwd(b.c3, g3s, b.o2, NI, b.Hout, b.Hout, t2, nullptr, true, s, b.o2, b.c2.scale))
          relu_bn_bwd<T>(t2, b.o2, b.c2.scale, t2, nullptr, Mo, b.c2.cout, s);
        if (!conv_bwd(b.c2, t2, b.o1, NI, b.Hin, b.Hout, t1, nullptr, true, s, b.o1, b.c1.scale))
          relu_bn_bwd<T>(t1, b.o1, b.c1.scale, t1, nullptr, Mi, b.c1.cout, s);
      }
      const T* idgrad = g3;
      if (b.has_ds) {
        relu_bn_bwd<T>(g3, nullptr, b.ds.scale, g3s, nullptr, Mo, b.ds.cout, s);
        conv_bwd(b.ds, g3s, x_in, NI, b.Hin, b.Hout, gid, nullptr, need_dx, s);
        idgrad = gid;
      }
      conv_bwd(b.c1, t1, x_in, NI, b.Hin, b.basic ? b.Hout : b.Hin, dxn, idgrad, need_dx, s);
      std::swap(dy, dxn);
      // the first block of a stage is the last one the reverse loop visits: the stage's gradients are final
      int start = 0;
      for (int st = 0; st < 4; ++st) {
        if (bi == start && st >= 1) bucket_ready(6 - st, s);       // layer4 -> bucket 3, layer3 -> 4, layer2 -> 5
        start += desc.blocks[st];
      }
    }
  }
  buckets_join(s);
  launches += g_launches - l0;
}

template <typename T>
void TrainerT<T>::clip_adamw(float max_norm, float lr, float lr_backbone, float beta1, float beta2, float eps,
                             float weight_decay, int step, float* grad_norm_dev, cudaStream_t s) {
  const long long l0 = g_launches;
  CUDA_OK(cudaMemsetAsync(sumsq_, 0, 4, s));
  if (arena_g_) {
    grad_sumsq(arena_g_, arena_n_, sumsq_, s);
    if (grad_norm_dev) CUDA_OK(cudaMemcpyAsync(grad_norm_dev, sumsq_, 4, cudaMemcpyDeviceToDevice, s));
    if (!arena_m_) {
      arena_m_ = alloc<float>(arena_n_); arena_v_ = alloc<float>(arena_n_);
      CUDA_OK(cudaMemsetAsync(arena_m_, 0, arena_n_ * 4, s));
      CUDA_OK(cudaMemsetAsync(arena_v_, 0, arena_n_ * 4, s));
    }
    if (arena_nb_ > 0)
      adamw_step(arena_p_, arena_g_, arena_m_, arena_v_, arena_nb_, sumsq_, max_norm, lr_backbone, beta1, beta2, eps,
                 weight_decay, step, s);
    if (arena_n_ > arena_nb_)
      adamw_step(arena_p_ + arena_nb_, arena_g_ + arena_nb_, arena_m_ + arena_nb_, arena_v_ + arena_nb_,
                 arena_n_ - arena_nb_, sumsq_, max_norm, lr, beta1, beta2, eps, weight_decay, step, s);
    launches += g_launches - l0;
    return;
  }
  for (auto& kv : params_)
    if (kv.second.g) grad_sumsq(kv.second.g, kv.second.numel, sumsq_, s);
  if (grad_norm_dev) CUDA_OK(cudaMemcpyAsync(grad_norm_dev, sumsq_, 4, cudaMemcpyDeviceToDevice, s));
  for (auto& kv : params_) {
    PBind& b = kv.second;
    if (!b.g) continue;
    if (!b.m) {
      b.m = alloc<float>(b.numel); b.v = alloc<float>(b.numel);
      CUDA_OK(cudaMemsetAsync(b.m, 0, b.numel * 4, s));
      CUDA_OK(cudaMemsetAsync(b.v, 0, b.numel * 4, s));
    }
    const bool is_bb = kv.first.find("backbone") != std::string::npos;
    adamw_step(b.p, b.g, b.m, b.v, b.numel, sumsq_, max_norm, is_bb ? lr_backbone : lr, beta1, beta2, eps, weight_decay,
               step, s);
  }
  launches += g_launches - l0;
}

}  // namespace

std::unique_ptr<TrainerBase> make_trainer(const retr_model_desc& d, int pad_token_id) {
  if (d.precision == RETR_FP32) return std::unique_ptr<TrainerBase>(new TrainerT<float>(d, pad_token_id));
  if (d.precision == RETR_BF16) {
    CUDA_OK(cudaSetDevice(d.device));
    gemm_bf16_prepare();
    return std::unique_ptr<TrainerBase>(new TrainerT<bf16>(d, pad_token_id));
  }
  throw Error("unknown precision");
}

}  // namespace retr
