// extern "C" surface declared in include/retr_b200.h.  Every entry point converts C++ exceptions
// into a non-zero status + retr_last_error(); there is no CPU fallback behind any of them.
#include "engine.h"
#include "small_gemm.h"
#include "ffn_fused.h"
#include "train.h"
#include "dp.h"

namespace retr {
thread_local long long g_launches = 0;
static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
bool pdl_enabled() {
  static const bool on = getenv("RETR_NO_PDL") == nullptr;
  return on;
}
}  // namespace retr

using namespace retr;

struct retr_ctx {
  std::unique_ptr<EngineBase> eng;
};
struct retr_trainer {
  std::unique_ptr<TrainerBase> tr;
  int device = 0;
};

#define API_BEGIN try {
#define API_END                                  \
  return 0;                                      \
  }                                              \
  catch (const std::exception& e) {              \
    retr::set_error(e.what());                   \
    return 1;                                    \
  }                                              \
  catch (...) {                                  \
    retr::set_error("unknown native exception"); \
    return 1;                                    \
  }

extern "C" {

const char* retr_last_error(void) { return g_error.c_str(); }
int retr_abi_version(void) { return 1; }

int retr_create(const retr_model_desc* desc, retr_ctx** out) {
  API_BEGIN
  RETR_CHECK(desc && out, "null argument");
  int ndev = 0;
  CUDA_OK(cudaGetDeviceCount(&ndev));
  RETR_CHECK(desc->device >= 0 && desc->device < ndev, "no such CUDA device");
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, desc->device));
  RETR_CHECK(prop.major == 10, std::string("retr_b200 is built for sm_100a only; device is sm_") +
                                   std::to_string(prop.major) + std::to_string(prop.minor));
  retr_ctx* c = new retr_ctx;
  try {
    c->eng = make_engine(*desc);
  } catch (...) {
    delete c;
    throw;
  }
  *out = c;
  API_END
}

void retr_destroy(retr_ctx* ctx) { delete ctx; }

int retr_load_weight(retr_ctx* ctx, const char* name, const float* data, const int64_t* shape, int32_t ndim,
                     int32_t on_device) {
  API_BEGIN
  RETR_CHECK(ctx && name && data && shape, "null argument");
  CUDA_OK(cudaSetDevice(ctx->eng->desc.device));
  ctx->eng->load_weight(name, data, shape, ndim, on_device != 0);
  API_END
}

int retr_finalize_weights(retr_ctx* ctx) {
  API_BEGIN
  RETR_CHECK(ctx, "null context");
  CUDA_OK(cudaSetDevice(ctx->eng->desc.device));
  ctx->eng->finalize();
  API_END
}

int retr_encode(retr_ctx* ctx, int32_t batch, const float* t_img, const uint8_t* t_mask, const float* g_img,
                const uint8_t* g_mask, const float* loc, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx, "null context");
  ctx->eng->encode(batch, t_img, t_mask, g_img, g_mask, loc, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_forward_teacher_forced(retr_ctx* ctx, const int64_t* ids, const uint8_t* id_mask, float* logits,
                                float* att_self, float* att_cross, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && ids && id_mask && logits, "null argument");
  ctx->eng->teacher_forced(ids, id_mask, logits, att_self, att_cross, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_set_encoder_attention_capture(retr_ctx* ctx, float* att_enc) {
  API_BEGIN
  RETR_CHECK(ctx, "null context");
  ctx->eng->att_enc_capture = att_enc;
  API_END
}

int retr_decode_greedy(retr_ctx* ctx, int32_t bos, int32_t eos, int32_t max_len, int64_t* out_ids,
                       int32_t* steps_run_host, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && out_ids, "null argument");
  ctx->eng->decode_greedy(bos, eos, max_len, out_ids, steps_run_host, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_generate_host(retr_ctx* ctx, int32_t batch, const float* t_img_host, const uint8_t* t_mask_host,
                       const float* g_img_host, const uint8_t* g_mask_host, const float* loc_host, int32_t bos,
                       int32_t eos, int32_t max_len, int64_t* out_ids_host, int32_t* steps_run_host, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && t_img_host && t_mask_host && out_ids_host, "null argument");
  ctx->eng->generate_host(batch, t_img_host, t_mask_host, g_img_host, g_mask_host, loc_host, bos, eos, max_len,
                          out_ids_host, steps_run_host, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_prefetch_host(retr_ctx* ctx, int32_t batch, const float* t_img_host, const uint8_t* t_mask_host,
                       const float* g_img_host, const uint8_t* g_mask_host, const float* loc_host) {
  API_BEGIN
  RETR_CHECK(ctx && t_img_host && t_mask_host, "null argument");
  ctx->eng->prefetch_host(batch, t_img_host, t_mask_host, g_img_host, g_mask_host, loc_host);
  API_END
}

int retr_encode_host(retr_ctx* ctx, int32_t batch, const float* t_img_host, const uint8_t* t_mask_host,
                     const float* g_img_host, const uint8_t* g_mask_host, const float* loc_host, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && t_img_host && t_mask_host, "null argument");
  ctx->eng->encode_host(batch, t_img_host, t_mask_host, g_img_host, g_mask_host, loc_host,
                        static_cast<cudaStream_t>(stream), false);
  API_END
}

int retr_decode_host(retr_ctx* ctx, int32_t bos, int32_t eos, int32_t max_len, int64_t* out_ids_host,
                     int32_t* steps_run_host, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && out_ids_host, "null argument");
  ctx->eng->decode_host(bos, eos, max_len, out_ids_host, steps_run_host, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_set_encode_cta_cap(retr_ctx* ctx, int32_t ctas) {
  API_BEGIN
  RETR_CHECK(ctx && ctas >= 0, "bad argument");
  ctx->eng->encode_cta_cap = ctas;
  API_END
}

int retr_encoder_tokens(const retr_ctx* ctx) { return ctx ? ctx->eng->S : -1; }

int retr_read_state(retr_ctx* ctx, const char* what, void* dst, int64_t dst_bytes, void* stream) {
  API_BEGIN
  RETR_CHECK(ctx && what && dst, "null argument");
  ctx->eng->read_state(what, dst, dst_bytes, static_cast<cudaStream_t>(stream));
  API_END
}

int64_t retr_kernel_launches(const retr_ctx* ctx) { return ctx ? ctx->eng->launches : -1; }

int retr_last_phase_ms(const retr_ctx* ctx, float* out3) {
  API_BEGIN
  RETR_CHECK(ctx && out3, "null argument");
  for (int i = 0; i < 3; ++i) out3[i] = ctx->eng->phase_ms[i];
  API_END
}

int retr_op_gemm(int32_t dtype, const retr_gemm_args* a, void* stream) {
  API_BEGIN
  RETR_CHECK(a, "null argument");
  GemmParams p;
  p.A = a->A; p.W = a->W; p.C = a->C; p.M = a->M; p.N = a->N; p.K = a->K; p.lda = a->lda; p.ldc = a->ldc;
  p.scale = a->scale; p.bias = a->bias; p.rowbias = a->rowbias; p.rowbias_period = a->rowbias_period;
  p.residual = a->residual; p.ldr = a->ldr; p.res_fp32 = a->res_fp32 != 0; p.relu = a->relu;
  p.out_fp32 = a->out_fp32 != 0;
  p.conv.kh = a->kh; p.conv.kw = a->kw; p.conv.stride = a->stride; p.conv.pad = a->pad; p.conv.dil = a->dil;
  p.conv.H = a->H; p.conv.W = a->W_in; p.conv.C = a->C_in; p.conv.OH = a->OH; p.conv.OW = a->OW;
  p.argmax_out = a->argmax_out;
  p.mn_major = a->mn_major != 0; p.ldb = a->ldb;
  RETR_CHECK(!p.mn_major || dtype == RETR_BF16, "mn_major GEMM is bf16 only");
  if (dtype == RETR_FP32) gemm_fp32(p, static_cast<cudaStream_t>(stream));
  else if (dtype == RETR_BF16) gemm_bf16(p, static_cast<cudaStream_t>(stream));
  else throw Error("unknown dtype");
  API_END
}

int retr_preprocess(const uint8_t* pixels, const int64_t* offsets, const int32_t* hw, const double* boxes, int32_t batch,
                    int32_t out_size, float* t_img, uint8_t* t_mask, float* g_img, uint8_t* g_mask, float* loc, void* stream) {
  API_BEGIN
  preprocess(pixels, offsets, hw, boxes, batch, out_size, t_img, t_mask, g_img, g_mask, loc, static_cast<cudaStream_t>(stream));
  API_END
}
int retr_preprocess_max_side(int32_t out_size) { return preprocess_max_side(out_size); }

int retr_op_small_gemm_noalloc(const void* A, const void* W, const float* bias, float* C, int32_t M, int32_t N, int32_t K,
                               int32_t split_k, void* stream) {
  API_BEGIN
  small_gemm_splitk_static(static_cast<const bf16*>(A), static_cast<const bf16*>(W), bias, C, M, N, K, split_k,
                           static_cast<cudaStream_t>(stream));
  API_END
}

int retr_op_layernorm(const float* x, const float* gamma, const float* beta, void* y, int64_t rows, int32_t d,
                      float eps, int32_t out_bf16, void* stream) {
  API_BEGIN
  if (out_bf16) layernorm<bf16>(x, gamma, beta, static_cast<bf16*>(y), rows, d, eps, static_cast<cudaStream_t>(stream));
  else layernorm<float>(x, gamma, beta, static_cast<float*>(y), rows, d, eps, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_op_attention(int32_t dtype, const void* q, const void* k, const void* v, void* out, int32_t B, int32_t Lq,
                      int32_t Lk, int32_t n_heads, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo,
                      const uint8_t* key_pad, int32_t causal, float* probs, void* stream) {
  API_BEGIN
  AttnParams p{};
  p.q = q; p.k = k; p.v = v; p.out = out; p.B = B; p.Lq = Lq; p.Lk = Lk; p.H = n_heads;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.ldo = ldo; p.key_pad = key_pad; p.causal = causal; p.probs = probs;
  if (dtype == RETR_FP32) attention<float>(p, static_cast<cudaStream_t>(stream));
  else if (dtype == RETR_BF16) attention<bf16>(p, static_cast<cudaStream_t>(stream));
  else throw Error("unknown dtype");
  API_END
}

int retr_op_decode_ffn(float* y, const float* gamma, const float* beta, float eps, const void* W1, const float* b1,
                       const void* W2, const float* b2, int32_t M, int32_t d_ff, void* stream) {
  API_BEGIN
  RETR_CHECK(ffn_fused_supported(256, d_ff), "retr_op_decode_ffn: d_model 256 and d_ff 2048 only");
  ffn_fused_configure();
  FfnFusedParams p;
  p.Y = y; p.gamma = gamma; p.beta = beta; p.eps = eps; p.W1 = static_cast<const bf16*>(W1); p.b1 = b1;
  p.W2 = static_cast<const bf16*>(W2); p.b2 = b2; p.M = M; p.d_ff = d_ff;
  ffn_fused(p, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_op_decode_out_ln_proj(const void* att, const void* Wo, const float* bo, float* y, const float* gamma,
                               const float* beta, float eps, const void* Wq, const float* bq, const float* rowbias,
                               const int32_t* row, void* q, int32_t M, void* stream) {
  API_BEGIN
  RETR_CHECK((rowbias == nullptr) == (row == nullptr), "rowbias and row go together");
  ffn_fused_configure();
  AttnOutProjParams p;
  p.att = static_cast<const bf16*>(att); p.ld_att = 256; p.Wo = static_cast<const bf16*>(Wo); p.bo = bo; p.Y = y;
  p.gamma = gamma; p.beta = beta; p.eps = eps; p.Wq = static_cast<const bf16*>(Wq); p.bq = bq;
  p.rowbias = rowbias; p.rowbias_dyn = row; p.Q = static_cast<bf16*>(q); p.ld_q = 256; p.M = M;
  attn_out_ln_proj(p, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_op_decode_cross_attention(const void* q, const void* kcache, const void* vcache, int32_t cache_rows, void* out,
                                   int32_t M, int32_t n_heads, int32_t d_model, const int32_t* n_keys, void* stream) {
  API_BEGIN
  decode_attention_configure<bf16>(cache_rows, d_model / n_heads);
  decode_attention<bf16>(static_cast<const bf16*>(q), d_model, nullptr, 0,
                         const_cast<bf16*>(static_cast<const bf16*>(kcache)), const_cast<bf16*>(static_cast<const bf16*>(vcache)),
                         cache_rows, static_cast<bf16*>(out), d_model, M, n_heads, d_model, cache_rows, nullptr, nullptr, 0,
                         static_cast<cudaStream_t>(stream), n_keys);
  API_END
}

// ---- training ------------------------------------------------------------------------------------------
int retr_train_create(const retr_model_desc* desc, int32_t pad_token_id, retr_trainer** out) {
  API_BEGIN
  RETR_CHECK(desc && out, "null argument");
  int ndev = 0;
  CUDA_OK(cudaGetDeviceCount(&ndev));
  RETR_CHECK(desc->device >= 0 && desc->device < ndev, "no such CUDA device");
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, desc->device));
  RETR_CHECK(prop.major == 10, std::string("retr_b200 is built for sm_100a only; device is sm_") +
                                   std::to_string(prop.major) + std::to_string(prop.minor));
  retr_trainer* t = new retr_trainer;
  try {
    t->tr = make_trainer(*desc, pad_token_id);
    t->device = desc->device;
  } catch (...) {
    delete t;
    throw;
  }
  *out = t;
  API_END
}

void retr_train_destroy(retr_trainer* tr) { delete tr; }

int retr_train_bind(retr_trainer* tr, const char* name, float* param, float* grad, const int64_t* shape, int32_t ndim) {
  API_BEGIN
  RETR_CHECK(tr && name && param && shape, "null argument");
  tr->tr->bind(name, param, grad, shape, ndim);
  API_END
}

int retr_train_set_arena(retr_trainer* tr, float* params, float* grads, int64_t n_total, int64_t n_backbone) {
  API_BEGIN
  RETR_CHECK(tr && params && grads && n_total > 0 && n_backbone >= 0 && n_backbone <= n_total, "bad arena");
  tr->tr->set_arena(params, grads, n_total, n_backbone);
  API_END
}

int retr_train_finalize(retr_trainer* tr) {
  API_BEGIN
  RETR_CHECK(tr, "null trainer");
  CUDA_OK(cudaSetDevice(tr->device));
  tr->tr->finalize();
  API_END
}

int retr_train_forward(retr_trainer* tr, int32_t batch, const float* t_img, const uint8_t* t_mask, const float* g_img,
                       const uint8_t* g_mask, const float* loc, const int64_t* ids, const uint8_t* id_mask, float* logits,
                       float dropout_p, uint64_t seed, void* stream) {
  API_BEGIN
  RETR_CHECK(tr && logits, "null argument");
  tr->tr->forward(batch, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, logits, dropout_p, seed,
                  static_cast<cudaStream_t>(stream));
  API_END
}

int retr_train_backward(retr_trainer* tr, const float* dlogits, void* stream) {
  API_BEGIN
  RETR_CHECK(tr, "null trainer");
  tr->tr->backward(dlogits, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_train_loss_and_grads(retr_trainer* tr, int32_t batch, const float* t_img, const uint8_t* t_mask,
                              const float* g_img, const uint8_t* g_mask, const float* loc, const int64_t* ids,
                              const uint8_t* id_mask, const int64_t* targets, float dropout_p, uint64_t seed,
                              float* loss_sum_dev, void* stream) {
  API_BEGIN
  RETR_CHECK(tr, "null trainer");
  tr->tr->loss_and_grads(batch, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, targets, dropout_p, seed, loss_sum_dev,
                         static_cast<cudaStream_t>(stream));
  API_END
}

int retr_train_clip_adamw(retr_trainer* tr, float max_norm, float lr, float lr_backbone, float beta1, float beta2,
                          float eps, float weight_decay, int32_t step, float* grad_sumsq_dev, void* stream) {
  API_BEGIN
  RETR_CHECK(tr && step >= 1, "bad argument");
  tr->tr->clip_adamw(max_norm, lr, lr_backbone, beta1, beta2, eps, weight_decay, step, grad_sumsq_dev,
                     static_cast<cudaStream_t>(stream));
  API_END
}

int retr_nccl_unique_id(void* out128) {
  API_BEGIN
  RETR_CHECK(out128, "null argument");
  nccl_unique_id id;
  nccl_ok(nccl().GetUniqueId(&id), "ncclGetUniqueId");
  memcpy(out128, id.internal, 128);
  API_END
}

int retr_nccl_comm_create(const void* id128, int32_t world, int32_t rank, void** comm_out) {
  API_BEGIN
  RETR_CHECK(id128 && comm_out && world >= 1 && rank >= 0 && rank < world, "bad argument");
  nccl_unique_id id;
  memcpy(id.internal, id128, 128);
  nccl_comm_t c = nullptr;
  nccl_ok(nccl().CommInitRank(&c, world, id, rank), "ncclCommInitRank");
  *comm_out = c;
  API_END
}

void retr_nccl_comm_destroy(void* comm) {
  try {
    if (comm) nccl().CommDestroy(static_cast<nccl_comm_t>(comm));
  } catch (...) {
  }
}

int retr_train_set_comm(retr_trainer* tr, void* nccl_comm, int32_t world) {
  API_BEGIN
  RETR_CHECK(tr, "null trainer");
  tr->tr->set_comm(nccl_comm, world);
  API_END
}

int retr_allreduce_grads(retr_trainer* tr, void* nccl_comm, void* stream) {
  API_BEGIN
  RETR_CHECK(tr, "null trainer");
  tr->tr->allreduce_grads(nccl_comm, static_cast<cudaStream_t>(stream));
  API_END
}

int retr_train_grad_buckets(const retr_trainer* tr, int64_t* lo_hi, int32_t max_buckets) {
  if (!tr) return -1;
  return tr->tr->bucket_info(lo_hi, max_buckets);
}

int64_t retr_train_kernel_launches(const retr_trainer* tr) { return tr ? tr->tr->launches : -1; }

}  // extern "C"
