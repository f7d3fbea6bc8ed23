/* retr_b200 — C ABI of the Blackwell-native (sm_100a) RE⫶TR hot path.
 *
 * The reference (simeonjunker/retr) is pure Python/PyTorch and has no FFI layer; its
 * boundary for this path is the Python API.  This header is the C ABI
 * that sits directly beneath the re-implementation of those Python symbols; every entry
 * point names the reference interface it replaces (paths relative to /root/reference).
 *
 * Conventions: plain pointers and sizes only (no torch types); device pointers unless a
 * name ends in `_host`; `stream` is a cudaStream_t passed as void*; every call returns 0
 * on success, non-zero on error with the message available from retr_last_error().
 * A context is not thread-safe; distinct contexts are independent.
 */
#ifndef RETR_B200_H
#define RETR_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct retr_ctx retr_ctx;

enum { RETR_FP32 = 0, RETR_BF16 = 1 };

/* Hyper-parameters of one model: the attributes build_model reads from Config
 * (configuration_template.py:4-56, models/caption.py:177-212). */
typedef struct retr_model_desc {
  int32_t device;            /* CUDA ordinal */
  int32_t precision;         /* RETR_FP32 (SIMT fp32 parity mode) | RETR_BF16 (tcgen05) */
  int32_t blocks[4];         /* bottlenecks per ResNet stage, e.g. {3,4,23,3} */
  int32_t dilation;          /* layer4 stride replaced by dilation (models/backbone.py:89-91) */
  int32_t n_images;          /* backbone passes per sample: 1 Caption/CaptionLoc, 2 CaptionGlobalLoc */
  int32_t loc_in_dim;        /* 0 none | 7 CaptionLoc (one token) | 1 CaptionGlobalLoc (token per scalar) */
  int32_t n_loc;             /* location scalars per sample (5 for CaptionGlobalLoc, 7 for CaptionLoc) */
  int32_t image_size;        /* H = W of the input crops (224) */
  int32_t d_model, n_heads, d_ff, enc_layers, dec_layers;
  int32_t max_pos;           /* max_position_embeddings == caption length T (128) */
  int32_t vocab;
  int32_t pre_norm;          /* final encoder LayerNorm present (models/ConcatTransformer.py:24) */
  float   embed_ln_eps;      /* DecoderEmbeddings.LayerNorm eps (config.layer_norm_eps) */
  int32_t max_batch;         /* largest number of samples per call; sizes the arena */
} retr_model_desc;

const char* retr_last_error(void);
int         retr_abi_version(void);

/* ---- context ------------------------------------------------------------------------- */
/* replaces: models/caption.py:177-212 build_model (module construction). */
int  retr_create(const retr_model_desc* desc, retr_ctx** out);
void retr_destroy(retr_ctx* ctx);

/* replaces: nn.Module.load_state_dict as used at eval_model.py:27-28.  `name` is the
 * reference state_dict key; `data` is fp32, device or host. */
int retr_load_weight(retr_ctx* ctx, const char* name, const float* data,
                     const int64_t* shape, int32_t ndim, int32_t on_device);
/* Repack everything loaded so far into kernel layouts (KRSC conv filters, bf16 copies,
 * position-bias tables).  Must be called after the last retr_load_weight. */
int retr_finalize_weights(retr_ctx* ctx);

/* ---- hot path ------------------------------------------------------------------------ */
/* replaces: Caption*.forward up to the encoder output (models/caption.py:112-149,
 * models/backbone.py:68-77, models/ConcatTransformer.py:48-67) and the per-layer
 * cross-attention K/V projections.  Images fp32 NCHW [B,3,H,W]; masks uint8 [B,H,W]
 * (non-zero = padding); g_* may be NULL when n_images == 1; loc fp32 [B,n_loc] or NULL. */
int retr_encode(retr_ctx* ctx, int32_t batch, const float* t_img, const uint8_t* t_mask,
                const float* g_img, const uint8_t* g_mask, const float* loc, void* stream);

/* replaces: the decoder + MLP half of Caption*.forward (models/ConcatTransformer.py:63-74,
 * models/caption.py:153) for a full caption buffer.  ids int64 [B,T]; id_mask uint8 [B,T]
 * (non-zero = padding key); logits fp32 [B,T,vocab].  Uses the encoder state of the last
 * retr_encode.  att_* (optional, may be NULL): head-averaged attention stacks
 * self [Ld,B,T,T], cross [Ld,B,T,S] (models/ConcatTransformer.py:72-74). */
int retr_forward_teacher_forced(retr_ctx* ctx, const int64_t* ids, const uint8_t* id_mask,
                                float* logits, float* att_self, float* att_cross, void* stream);
/* return_attention=True support for the encoder half: while att_enc is non-NULL, retr_encode
 * also writes the head-averaged encoder self-attention stack [Le,B,S,S] there. */
int retr_set_encoder_attention_capture(retr_ctx* ctx, float* att_enc);

/* replaces: eval_utils/decode.py:53-81 greedy (the loop after samples are on the device),
 * as a KV-cached device-resident loop captured in a CUDA graph.  out_ids int64
 * [B,max_len] (position 0 = bos, unwritten tail = 0, same early-return semantics);
 * steps_run_host receives the number of decode steps executed. */
int retr_decode_greedy(retr_ctx* ctx, int32_t bos, int32_t eos, int32_t max_len,
                       int64_t* out_ids, int32_t* steps_run_host, void* stream);

/* replaces: eval_utils/decode.py:53-81 greedy end to end with HOST buffers: H2D of the
 * samples (decode.py:62), encode, decode, D2H of the ids. Pinned host memory recommended. */
int retr_generate_host(retr_ctx* ctx, int32_t batch, const float* t_img_host,
                       const uint8_t* t_mask_host, const float* g_img_host,
                       const uint8_t* g_mask_host, const float* loc_host, int32_t bos,
                       int32_t eos, int32_t max_len, int64_t* out_ids_host,
                       int32_t* steps_run_host, void* stream);

/* replaces: the one-batch-at-a-time `samples.to(device)` of the evaluation loop (engine.py:138-146,
 * eval_utils/decode.py:62) with a double-buffered transfer: queues the H2D copies of the NEXT batch on
 * the context's copy stream and returns at once.  The next retr_generate_host call given the same
 * batch and host pointers consumes the staged copy (no second transfer, one backbone pass); any other
 * call simply copies as before.  The host buffers must stay alive and unchanged until that
 * retr_generate_host returns.  Call it between two retr_generate_host calls, never concurrently. */
int retr_prefetch_host(retr_ctx* ctx, int32_t batch, const float* t_img_host,
                       const uint8_t* t_mask_host, const float* g_img_host,
                       const uint8_t* g_mask_host, const float* loc_host);

/* Two contexts in flight (eval_utils.decode.greedy_stream with contexts=2): the reference's eval loop
 * (engine.py:138-146) handles one batch at a time; here the NEXT batch's copies, backbone and encoder
 * (tensor-core bound) run on one context while the CURRENT batch's decode loop (latency / HBM bound) runs
 * on another, each on its own stream.  retr_encode_host is retr_generate_host up to and including the
 * encoder and returns at once; retr_decode_host runs the greedy loop of the batch last encoded on this
 * context (eval_utils/decode.py:64-81) and returns when the ids are in out_ids_host.
 * retr_set_encode_cta_cap bounds the persistent GEMM grids of the encode phase to `ctas` SMs (0 = all)
 * so the decode kernels of the other context always find free SMs. */
int retr_encode_host(retr_ctx* ctx, int32_t batch, const float* t_img_host, const uint8_t* t_mask_host,
                     const float* g_img_host, const uint8_t* g_mask_host, const float* loc_host,
                     void* stream);
int retr_decode_host(retr_ctx* ctx, int32_t bos, int32_t eos, int32_t max_len, int64_t* out_ids_host,
                     int32_t* steps_run_host, void* stream);
int retr_set_encode_cta_cap(retr_ctx* ctx, int32_t ctas);

/* ---- introspection (tests, bench) ------------------------------------------------------ */
int     retr_encoder_tokens(const retr_ctx* ctx);           /* S */
/* Copies of intermediate state as fp32: "memory" [B,S,d], "mask" (uint8 [B,S]),
 * "src" [B,S,d] (encoder input tokens), "feat_t"/"feat_g" [B,14,14,2048] NHWC. */
int     retr_read_state(retr_ctx* ctx, const char* what, void* dst, int64_t dst_bytes, void* stream);
int64_t retr_kernel_launches(const retr_ctx* ctx);          /* kernels launched so far */
/* Per-phase device times (ms) of the last generate/encode/decode measured with CUDA events:
 * out[0]=backbone, out[1]=encoder+crossKV, out[2]=decode loop. */
int     retr_last_phase_ms(const retr_ctx* ctx, float* out3);

/* ---- single-operator entry points (parity tests call kernels one at a time) ------------ */
/* C[M,N] = epilogue(A[M,K] · W[N,K]^T).  dtype: RETR_FP32 -> fp32 SIMT kernel, A/W fp32;
 * RETR_BF16 -> tcgen05 kernel, A/W bf16.  out_fp32 selects the C element type (else the
 * A type).  Optional: scale[N], bias[N], rowbias[period,N] (row m uses m % period),
 * residual[M,N] (res_fp32 selects its type), relu. */
typedef struct retr_gemm_args {
  const void* A; const void* W; void* C;
  int64_t M; int32_t N, K; int64_t lda, ldc;
  const float* scale; const float* bias; const float* rowbias; int32_t rowbias_period;
  const void* residual; int64_t ldr; int32_t res_fp32; int32_t relu; int32_t out_fp32;
  /* implicit-GEMM convolution view of A (NHWC input); kh == 0 -> plain matrix */
  int32_t kh, kw, stride, pad, dil, H, W_in, C_in, OH, OW;
  /* fused argmax over N (vocab head): packed (value,index) keys, uint64 [M]; C may be NULL */
  unsigned long long* argmax_out;
  /* mn_major != 0 (bf16 only): C[M,N] = A^T . W with A given as [K, M] (row pitch lda) and W as [K, N] (row pitch
   * ldb) — the weight-gradient contraction over tokens / pixels, operands read in place.  With a conv view, W is the
   * NHWC layer input, N = kh*kw*C_in and K = images*OH*OW (dW of the convolution as [C_out, (kh,kw,C_in)]). */
  int32_t mn_major; int64_t ldb;
} retr_gemm_args;
int retr_op_gemm(int32_t dtype, const retr_gemm_args* args, void* stream);

/* LayerNorm over the last dim d of fp32 x[rows,d] -> y (fp32 or bf16 by out_bf16). */
int retr_op_layernorm(const float* x, const float* gamma, const float* beta, void* y,
                      int64_t rows, int32_t d, float eps, int32_t out_bf16, void* stream);

/* Multi-head attention core on projected q/k/v (row-major, heads side by side, head dim 32):
 * out[b,i,:] = softmax(q k^T / sqrt(dh) + masks) v.  key_pad uint8 [B,Lk] or NULL.
 * probs (optional fp32 [B,Lq,Lk]) receives head-averaged probabilities. */
int retr_op_attention(int32_t dtype, const void* q, const void* k, const void* v, void* out,
                      int32_t B, int32_t Lq, int32_t Lk, int32_t n_heads, int64_t ldq,
                      int64_t ldk, int64_t ldv, int64_t ldo, const uint8_t* key_pad,
                      int32_t causal, float* probs, void* stream);

/* ---- fused sublayers of one KV-cached decode step (bf16 operands, d_model 256; parity tests) --------
 * retr_op_decode_ffn: y <- y + W2 relu(W1 LN(y) + b1) + b2 in place (the pre-norm FeedForward sublayer,
 *   models/transformer_modules.py:84-97); y fp32 [M,256], W1 bf16 [d_ff,256], W2 bf16 [256,d_ff], d_ff = 2048.
 * retr_op_decode_out_ln_proj: y <- y + Wo att + bo ; q = Wq LN(y) + bq + rowbias[*row]  (attention out-projection +
 *   residual, then the next pre-norm projection, models/transformer_modules.py:58-83); att bf16 [M,256],
 *   Wo/Wq bf16 [256,256], q bf16 [M,256]; bq, rowbias (fp32 [rows,256]) and row (device int) may be NULL together.
 * retr_op_decode_cross_attention: single-query attention of M samples over head-major K/V caches
 *   [M][H][cache_rows][dh] that hold n_keys[m] (device int32 [M], or NULL -> cache_rows) valid rows each. */
int retr_op_decode_ffn(float* y, const float* gamma, const float* beta, float eps, const void* W1, const float* b1,
                       const void* W2, const float* b2, int32_t M, int32_t d_ff, void* stream);
int retr_op_decode_out_ln_proj(const void* att, const void* Wo, const float* bo, float* y, const float* gamma,
                               const float* beta, float eps, const void* Wq, const float* bq, const float* rowbias,
                               const int32_t* row, void* q, int32_t M, void* stream);
int retr_op_decode_cross_attention(const void* q, const void* kcache, const void* vcache, int32_t cache_rows, void* out,
                                   int32_t M, int32_t n_heads, int32_t d_model, const int32_t* n_keys, void* stream);

/* ---- preprocessing: the step immediately before the hot path ------------------------
 * replaces: RefCocoCaption.__getitem__'s image half in validation mode (data_utils/refcoco.py:105-188 with
 * get_transforms('val') :14-46) and its helpers crop_image_to_bb, pad_img_to_max, pad_mask_to_max,
 * compute_position_features (data_utils/utils.py:161-256): crop to the box, context = image with the box zeroed,
 * pad to square (colour 0 / mask True), antialiased bilinear resize to out_size (PIL's fixed-point resampler for the
 * image, torchvision's float resize + .to(bool) for the mask), /255, ImageNet normalisation.  Bit-identical outputs.
 * pixels: packed uint8 RGB images (HWC), image b at byte offsets[b] with hw[2b] rows and hw[2b+1] columns;
 * boxes: double [B,4] = (x, y, w, h) as annotated (unrounded; x, y >= 0, box not empty after clipping to the image).
 * Outputs (device): t_img/g_img fp32 [B,3,out,out], t_mask/g_mask uint8 [B,out,out] (1 = padding / blanked target),
 * loc fp32 [B,5] = [x1, y1, x2, y2, area] relative; g_img+g_mask and loc may be NULL.  All pointers are device
 * pointers.  The padded square side must not exceed retr_preprocess_max_side(out_size) (6944 for 224). */
int retr_preprocess(const uint8_t* pixels, const int64_t* offsets, const int32_t* hw, const double* boxes, int32_t batch,
                    int32_t out_size, float* t_img, uint8_t* t_mask, float* g_img, uint8_t* g_mask, float* loc,
                    void* stream);
int retr_preprocess_max_side(int32_t out_size);

/* ---- allocation-free operator ---------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * C[M,N] (fp32, row pitch N) = A[M,K] . W[N,K]^T + bias[N] (bias may be NULL); A, W bf16 row-major, device memory.
 * The decode step's small-M GEMM (csrc/small_gemm.cu) with a deterministic split-K over `split_k` CTAs per output
 * tile, whose partial tiles and tile tickets live in file-scope static __device__ arrays instead of an arena:
 * the call path performs no cudaMalloc / cudaFree / allocating library call.
 * Limits (a clear error otherwise): split_k * M * N <= 1 048 576 floats (4 MiB of static scratch, held by every
 * GPU that loads the library until exit); ceil(M/64) * ceil(N/32) <= 4096 tiles; N % 8 == 0; K % 256 == 0;
 * split_k divides K/256 (K/512 when K == 512); M <= 2048.
 * All calls on a device SHARE the static arrays, which keep the previous call's contents: tickets are cleared on
 * `stream` before each launch; calls that overlap on the GPU (different streams without a dependency) race. */
int retr_op_small_gemm_noalloc(const void* A, const void* W, const float* bias, float* C, int32_t M, int32_t N,
                               int32_t K, int32_t split_k, void* stream);

/* ---- training step (engine.py:52-87 train_one_epoch; optimiser recipe main.py:30-40) ------------- */
typedef struct retr_trainer retr_trainer;

/* replaces: model.train() + the module tree as seen by loss.backward().  `pad_token_id` is the
 * nn.Embedding padding_idx whose row receives no gradient (models/transformer_modules.py:103-105). */
int  retr_train_create(const retr_model_desc* desc, int32_t pad_token_id, retr_trainer** out);
void retr_train_destroy(retr_trainer* tr);
/* Bind one reference state_dict entry (SURVEY Appendix B key).  `param`: fp32 device memory that the
 * trainer READS at every step (nothing is copied: an optimiser may update it in place).  `grad`: fp32
 * device memory of the same shape that retr_train_backward overwrites, or NULL for a frozen parameter or
 * a buffer (requires_grad False: conv1/layer1, models/backbone.py:58-60; FrozenBatchNorm2d buffers). */
int retr_train_bind(retr_trainer* tr, const char* name, float* param, float* grad, const int64_t* shape,
                    int32_t ndim);
/* Optional (before finalize): declare that every bound trainable parameter / gradient is a slice of two flat
 * fp32 arenas with the same element order, the first n_backbone elements being the "backbone" learning-rate
 * group (main.py:31-36).  The step then zeroes the gradients with one memset, takes the global norm with one
 * kernel and runs AdamW in two launches; without it the same work is done tensor by tensor. */
int retr_train_set_arena(retr_trainer* tr, float* params, float* grads, int64_t n_total, int64_t n_backbone);
int retr_train_finalize(retr_trainer* tr);
/* replaces: outputs = model(*samples, caps[:, :-1], cap_masks[:, :-1]) in train mode (engine.py:70):
 * dropout `dropout_p` at the reference's sites (embeddings, attention probabilities, residual branches)
 * from a counter-based Philox stream keyed by `seed`; activations are kept for the backward pass.
 * logits fp32 [B,T,vocab]. */
int retr_train_forward(retr_trainer* tr, int32_t batch, const float* t_img, const uint8_t* t_mask,
                       const float* g_img, const uint8_t* g_mask, const float* loc, const int64_t* ids,
                       const uint8_t* id_mask, float* logits, float dropout_p, uint64_t seed, void* stream);
/* replaces: loss.backward() (engine.py:80) below the logits: dlogits fp32 [B,T,vocab] -> bound grads. */
int retr_train_backward(retr_trainer* tr, const float* dlogits, void* stream);
/* replaces: engine.py:70-80 in one call — forward, CrossEntropyLoss (mean over all B*T positions, pads
 * included, models/caption.py:210) against targets int64 [B,T], backward.  loss_sum_dev (device fp32)
 * receives the SUM of the per-position losses (divide by B*T); the logits never leave bf16/HBM. */
int retr_train_loss_and_grads(retr_trainer* tr, int32_t batch, const float* t_img, const uint8_t* t_mask,
                              const float* g_img, const uint8_t* g_mask, const float* loc, const int64_t* ids,
                              const uint8_t* id_mask, const int64_t* targets, float dropout_p, uint64_t seed,
                              float* loss_sum_dev, void* stream);
/* replaces: clip_grad_norm_(model.parameters(), max_norm) + AdamW.step() (engine.py:81-83, main.py:30-40):
 * entries whose name contains "backbone" use lr_backbone.  step is the 1-based update count.
 * grad_sumsq_dev (optional, device fp32): squared global gradient norm before clipping. */
int retr_train_clip_adamw(retr_trainer* tr, float max_norm, float lr, float lr_backbone, float beta1,
                          float beta2, float eps, float weight_decay, int32_t step, float* grad_sumsq_dev,
                          void* stream);
int64_t retr_train_kernel_launches(const retr_trainer* tr);

/* ---- data-parallel training (the reference is single-device: main.py:16-24, seed offset main.py:19) ----
 * The gradient arena (retr_train_set_arena) is cut into buckets in REVERSE-EXECUTION order — MLP head, decoder +
 * embeddings, encoder + input/loc projections, backbone layer4, layer3, layer2 — each a contiguous range.
 * retr_train_set_comm: hand the trainer an ncclComm_t (as void*; NULL switches it off).  From then on
 *   retr_train_backward / retr_train_loss_and_grads all-reduce every bucket (ncclAvg = mean over the `world` ranks,
 *   i.e. the gradient of the global batch: every rank averages over the same B*T positions) on a side stream as soon
 *   as the backward pass has produced it, overlapped with the rest of the pass; the caller's stream waits for the last
 *   bucket before the call's work is complete, so retr_train_clip_adamw (clip_grad_norm_ over the REDUCED gradients,
 *   engine.py:81-83) needs nothing else.
 * retr_allreduce_grads: the same buckets as one NCCL group on `stream`, for callers that drive the exchange
 *   themselves (after retr_train_backward without retr_train_set_comm).
 * retr_train_grad_buckets: lo_hi[2i], lo_hi[2i+1] = element range of bucket i in the arena; returns their number.
 * retr_nccl_*: create the communicator without any other NCCL binding (rank 0 makes the 128-byte id, every rank
 *   passes it to retr_nccl_comm_create with its device current).  libnccl.so.2 is loaded at run time. */
int  retr_nccl_unique_id(void* out128);
int  retr_nccl_comm_create(const void* id128, int32_t world, int32_t rank, void** comm_out);
void retr_nccl_comm_destroy(void* comm);
int  retr_train_set_comm(retr_trainer* tr, void* nccl_comm, int32_t world);
int  retr_allreduce_grads(retr_trainer* tr, void* nccl_comm, void* stream);
int  retr_train_grad_buckets(const retr_trainer* tr, int64_t* lo_hi, int32_t max_buckets);

#ifdef __cplusplus
}
#endif
#endif
