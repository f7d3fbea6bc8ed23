// EngineT<T>: the schedule of the hot path.  T = float (fp32 parity mode, SIMT GEMMs) or bf16
// (tcgen05 GEMMs; residual stream, LayerNorm statistics, softmax and accumulators stay fp32).
//
// Reference being replaced (paths relative to /root/reference):
//   backbone     models/backbone.py:54-93 + torchvision ResNet Bottleneck (resnet.py:108-165)
//   tokens       models/caption.py:112-147, models/ConcatTransformer.py:48-61
//   encoder      models/ConcatTransformer.py:77-108,152-194 ; models/transformer_modules.py:22-46,77-97
//   decoder      models/ConcatTransformer.py:111-149,197-257 ; transformer_modules.py:49-74,100-129
//   head         models/caption.py:153,161-174
//   greedy loop  eval_utils/decode.py:53-81
//
// "+pos" trick: every place the reference adds a position table to a projection INPUT
// (q=k=LN(x)+pos, q=LN(y)+query_pos, k=memory+pos) is folded into a per-position bias table,
// W(x+p)+b = Wx + (Wp+b), computed once in fp32 at weight-load time (rowbias epilogue).
#include "engine.h"
#include "decode_stack.h"
#include "ffn_fused.h"
#include "small_gemm.h"

#include <tuple>
#include <type_traits>

namespace retr {
#ifdef RETR_TRACE
void small_gemm_set_trace(unsigned long long* buf);
void decode_ops_set_trace(unsigned long long* buf);
void ffn_fused_set_trace(unsigned long long* buf);
#endif
namespace {

// The decode step is captured on high-priority streams (kernel nodes keep the priority of the stream they were
// captured on): when a second context's encode phase shares the GPU, the short decode kernels are placed ahead of the
// queued CTAs of its convolutions.  RETR_DECODE_PRIORITY=0 turns it off.
// The one-kernel decode step (decode_stack.cu).  RETR_DECODE_STACK=0/1 overrides the default.
static bool stack_decode_wanted() {
  const char* e = getenv("RETR_DECODE_STACK");
  return e ? atoi(e) != 0 : true;    // B = 256: 36.8 ms per 127 steps against 46.8 for the launch chain
}
static int decode_priority() {
  int least = 0, greatest = 0;
  CUDA_OK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
  const char* e = getenv("RETR_DECODE_PRIORITY");
  return (e && atoi(e) == 0) ? least : greatest;
}

template <typename T> void gemm(const GemmParams& p, cudaStream_t s);
template <> void gemm<float>(const GemmParams& p, cudaStream_t s) { gemm_fp32(p, s); }
template <> void gemm<bf16>(const GemmParams& p, cudaStream_t s) { gemm_bf16(p, s); }

__global__ void fill_cols_kernel(float* dst, int rows, int ld, int col0, int ncols, const float* src) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < rows * ncols) dst[(int64_t)(idx / ncols) * ld + col0 + idx % ncols] = src[idx % ncols];
}
template <typename T> __global__ void to_f32_kernel(const T* src, float* dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = to_f(src[i]);
}
template <typename T> __global__ void from_f32_kernel(const float* src, T* dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = from_f<T>(src[i]);
}

template <typename T> struct Conv {
  int cin = 0, cout = 0, k = 1, stride = 1, pad = 0, dil = 1, kpad = 0;
  T* w = nullptr;           // [cout, kpad], k index = (r*kw+s)*cin + c
  float* scale = nullptr;   // folded FrozenBatchNorm2d: y = conv*scale + shift
  float* shift = nullptr;
};
template <typename T> struct Block {
  Conv<T> c1, c2, c3, ds;
  bool has_ds = false;
  bool basic = false;       // torchvision BasicBlock (ResNet-18/34): 3x3 (stride) -> 3x3, no c3
  int out_channels() const { return basic ? c2.cout : c3.cout; }
};
template <typename T> struct Linear {
  T* w = nullptr;           // [out, in]
  float* b = nullptr;
  int out = 0, in = 0;
};
template <typename T> struct EncLayer {
  float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  Linear<T> qkv, out, ff1, ff2;
  float* rb_qkv;            // [S, 3d]: pos.Wqk^T + b_qk | b_v
  float* b_v_tail;          // [3d]: 0 | b_v (with GemmParams::rowbias_cols = 2d the V columns skip the table)
  float* b_in;              // [3d] in_proj_bias (q | k | v): the packed encoder adds the positions before the projection
};
template <typename T> struct DecLayer {
  float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *ln3_g, *ln3_b;
  Linear<T> s_qkv, s_out, c_q, c_kv, c_out, ff1, ff2;
  float* rb_s_qkv;          // [T, 3d]: E_pos.Wqk^T + b_qk | b_v
  float* rb_c_q;            // [T, d] : E_pos.Wq^T + b_q
  float* rb_c_kv;           // [S, 2d]: pos.Wk^T + b_k | b_v
  float* b_cv_tail;         // [2d]: 0 | b_v
  float* b_ca;              // [3d] cross-attention in_proj_bias (q | k | v)
};

template <typename T>
class EngineT final : public EngineBase {
 public:
  explicit EngineT(const retr_model_desc& d) : EngineBase(d) {}
  ~EngineT() override {
    for (auto& kv : graphs_) cudaGraphExecDestroy(kv.second);
    if (done_host_) cudaFreeHost(done_host_);
    if (capture_stream_) cudaStreamDestroy(capture_stream_);
    for (auto st : gstreams_) cudaStreamDestroy(st);
    for (auto e : gjoin_) cudaEventDestroy(e);
    if (gfork_) cudaEventDestroy(gfork_);
  }
  void finalize() override;
  void encode(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
              const float* loc, cudaStream_t s) override;
  void teacher_forced(const int64_t* ids, const uint8_t* id_mask, float* logits, float* att_self, float* att_cross,
                      cudaStream_t s) override;
  void decode_greedy(int bos, int eos, int max_len, int64_t* out_ids, int* steps_host, cudaStream_t s) override;
  void read_state(const std::string& what, void* dst, int64_t bytes, cudaStream_t s) override;

 private:
  T* as_T(const RawWeight& w);                       // device copy of a weight in T
  Linear<T> linear(const std::string& wname, const std::string& bname, int row0 = 0, int rows = -1);
  Conv<T> conv(const std::string& wname, const std::string& bn, int stride, int pad, int dil);
  float* pos_bias_table(const float* pos, int rows, const RawWeight& w, const RawWeight& b, int row0, int n_proj,
                        int n_total);
  float* bias_tail(const RawWeight& b, int row0, int n_proj, int n_total);   // [n_total]: 0 | b[row0+n_proj ...]
  void run_conv(const Conv<T>& c, const T* in, int64_t n, int H, int W, T* out, const T* residual, int relu,
                int* OH, int* OW, cudaStream_t s);
  void backbone(const float* img, int64_t n, T* feat, cudaStream_t s, const float* img2 = nullptr, int64_t n2 = 0);
  void decoder_layers_full(int B, const uint8_t* id_mask, float* att_self, float* att_cross, cudaStream_t s);
  void decode_step(int B, int eos, int max_len, cudaStream_t s);
  void decode_step_fused(int B, int eos, int max_len, cudaStream_t s);
  void decode_rows_fused(int r0, int n, int group, int max_len, cudaStream_t s);
  void dgemm(const Linear<T>& l, const T* A, const float* X, const float* g, const float* b, void* C, bool out_fp32,
             int B, const float* rowbias, int relu, const float* residual, int split_k, cudaStream_t s);
  void alloc_buffers();
  GemmParams lin(const Linear<T>& l, const void* A, int64_t M, void* C) {
    GemmParams p;
    p.A = A; p.W = l.w; p.C = C; p.M = M; p.N = l.out; p.K = l.in; p.lda = l.in; p.ldc = l.out; p.bias = l.b;
    return p;
  }

  // weights
  Conv<T> stem_;
  std::vector<Block<T>> blocks_;
  Linear<T> input_proj_;
  float *loc_w_ = nullptr, *loc_b_ = nullptr;
  std::vector<EncLayer<T>> enc_;
  std::vector<DecLayer<T>> dec_;
  float *enc_norm_g_ = nullptr, *enc_norm_b_ = nullptr, *dec_norm_g_ = nullptr, *dec_norm_b_ = nullptr;
  float *e_word_ = nullptr, *e_pos_ = nullptr, *emb_g_ = nullptr, *emb_b_ = nullptr;
  Linear<T> mlp_[3];
  // activations
  int chunk_ = 512;                                  // images per backbone pass
  T *im2col_ = nullptr, *act_[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}, *feat_ = nullptr;
  float *x_ = nullptr, *y_ = nullptr;                // fp32 residual streams (encoder, decoder)
  T *n_ = nullptr, *qkv_ = nullptr, *att_ = nullptr, *h_ = nullptr, *memory_ = nullptr, *qc_ = nullptr;
  T *hs_ = nullptr, *m1_ = nullptr, *m2_ = nullptr;
  float* src_copy_ = nullptr;                        // encoder input tokens kept for read_state
  std::vector<T*> ckv_;                              // per layer [B*S, 2d] rows of [K|V] (teacher-forced path)
  std::vector<T*> ck_, cv_, sk_, sv_;                // head-major decode caches [B][H][rows][dh]
  int *cross_dst_ = nullptr, *cross_nvalid_ = nullptr;   // cross caches hold only the unpadded keys (decode_ops.cu)
  uint8_t *tok_mask_ = nullptr, *finished_ = nullptr;
  int64_t* caption_ = nullptr;
  int cross_prefetch_ = 0;      // RETR_CROSS_PREFETCH: measured slower (decode 46.7 -> 47.8 ms; 48.5 with the wrap-around), off
  int *step_ = nullptr, *done_ = nullptr;
  int* done_host_ = nullptr;                         // pinned [2]: done, step
  unsigned long long* argmax_keys_ = nullptr;
  unsigned int* vocab_ticket_ = nullptr;             // CTA ticket counter of vocab_argmax_kernel (zero between launches)
  std::map<std::tuple<int, int, int>, cudaGraphExec_t> graphs_;
  cudaStream_t capture_stream_ = nullptr;
  long long step_kernels_ = 0;
  float* splitk_scratch_ = nullptr;                  // [4][B][d] partial sums of the FFN second linear
  unsigned* splitk_tickets_ = nullptr;
  bool fast_decode_ = false;                         // bf16: small_gemm (LN fused) for the decode step
  // Samples are independent through the whole decode step, so the batch is cut into row groups whose
  // kernel chains run as parallel branches (forked streams -> parallel branches of the captured graph):
  // the latency-bound small GEMMs of one group overlap the HBM-bound cross-attention of another.
  int groups_ = 2, group_rows_min_ = 32, cur_group_ = 0;   // measured on B200 at B=256: 1 -> 67.8 ms, 2 -> 63.7, 4 -> 65.0, 8 -> 73.2
  std::vector<cudaStream_t> gstreams_;
  std::vector<cudaEvent_t> gjoin_;
  cudaEvent_t gfork_ = nullptr;
  int cur_G_ = 1;                                    // row groups of the step being captured
  bool ffn_fused_ = false;                           // FFN sublayer of the decode step as one cluster kernel (ffn_fused.cu)
  bf16* stem_wk_ = nullptr;                          // fused stem (stem_fused.cu): weights in its K order; null = three-launch stem
  // the whole decoder stack of a decode step as ONE cluster kernel (decode_stack.cu): its own chunked, swizzled caches
  bool stack_decode_ = false;
  // packed encoder (encode): only the unpadded tokens run through the encoder layers and the cross K/V projections
  bool packed_encoder_ = false, enc_packed_ = false, src_in_x_ = false;    // enabled / memory_ and ckv_ currently hold packed rows
  float* xp_ = nullptr;                                // packed residual stream [<= B*S, d]
  const float* pos_table_ = nullptr;                   // [S, d] fp32 positions of the concatenated sequence
  T *n2_ = nullptr, *mempos_ = nullptr;                // LN(x) + pos, memory + pos (packed encoder)
  T *wk_all_ = nullptr, *wv_all_ = nullptr, *kall_ = nullptr, *vall_ = nullptr;   // cross K / V of ALL decoder layers:
  float *bk_all_ = nullptr, *bv_all_ = nullptr;        // weights [L*d, d], biases [L*d], projections [<= B*S, L*d]
  int *seq_off_ = nullptr, *row_pos_ = nullptr;        // [B + 1] first packed row per sample (+ total); [B*S] position of a row
  void materialize_unpacked(cudaStream_t s);
  int cross_chunks_ = 0, self_chunks_ = 0;
  std::vector<bf16*> ckv2_, skv2_;
  DecStackParams stack_params_;
  void build_stack_params();
  size_t splitk_group_stride_ = 0;                   // floats of split-K scratch per group
};

// ---- weight preparation -------------------------------------------------------------------------
template <> float* EngineT<float>::as_T(const RawWeight& w) { return w.dev; }
template <> bf16* EngineT<bf16>::as_T(const RawWeight& w) {
  bf16* p = static_cast<bf16*>(dalloc(w.numel * 2));
  f32_to_bf16(w.dev, p, w.numel, 0);
  return p;
}

template <typename T>
Linear<T> EngineT<T>::linear(const std::string& wname, const std::string& bname, int row0, int rows) {
  const RawWeight& w = raw(wname);
  RETR_CHECK(w.shape.size() >= 2, "linear weight rank");
  Linear<T> l;
  l.in = (int)(w.numel / w.shape[0]);
  l.out = rows < 0 ? (int)w.shape[0] : rows;
  T* full = as_T(w);
  l.w = full + (int64_t)row0 * l.in;
  l.b = bname.empty() ? nullptr : raw(bname).dev + row0;
  return l;
}

template <typename T>
Conv<T> EngineT<T>::conv(const std::string& wname, const std::string& bn, int stride, int pad, int dil) {
  const RawWeight& w = raw(wname);
  RETR_CHECK(w.shape.size() == 4, "conv weight must be OIHW");
  Conv<T> c;
  c.cout = (int)w.shape[0]; c.cin = (int)w.shape[1]; c.k = (int)w.shape[2];
  c.stride = stride; c.pad = pad; c.dil = dil;
  const int kk = c.k * c.k * c.cin;
  c.kpad = (int)cdiv(kk, 16) * 16;
  c.w = static_cast<T*>(dalloc((size_t)c.cout * c.kpad * sizeof(T)));
  repack_filter<T>(w.dev, c.w, c.cout, c.cin, c.k, c.k, c.kpad, 0);
  // FrozenBatchNorm2d fold (models/backbone.py:41-51), evaluated in fp32 like the reference
  const int n = c.cout;
  std::vector<float> bw(n), bb(n), rm(n), rv(n), sc(n), sh(n);
  CUDA_OK(cudaMemcpy(bw.data(), raw(bn + ".weight").dev, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(bb.data(), raw(bn + ".bias").dev, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(rm.data(), raw(bn + ".running_mean").dev, n * 4, cudaMemcpyDeviceToHost));
  CUDA_OK(cudaMemcpy(rv.data(), raw(bn + ".running_var").dev, n * 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < n; ++i) {
    sc[i] = bw[i] * (1.0f / sqrtf(rv[i] + 1e-5f));
    sh[i] = bb[i] - rm[i] * sc[i];
  }
  c.scale = static_cast<float*>(dalloc(n * 4));
  c.shift = static_cast<float*>(dalloc(n * 4));
  CUDA_OK(cudaMemcpy(c.scale, sc.data(), n * 4, cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy(c.shift, sh.data(), n * 4, cudaMemcpyHostToDevice));
  return c;
}

// table[r, 0:n_proj] = pos[r,:] . W[row0:row0+n_proj,:]^T + b[row0:...]; table[r, n_proj:n_total] = b[row0+n_proj:...]
template <typename T>
float* EngineT<T>::pos_bias_table(const float* pos, int rows, const RawWeight& w, const RawWeight& b, int row0,
                                  int n_proj, int n_total) {
  const int d = desc.d_model;
  float* t = static_cast<float*>(dalloc((size_t)rows * n_total * 4));
  GemmParams p;
  p.A = pos; p.W = w.dev + (int64_t)row0 * d; p.C = t; p.M = rows; p.N = n_proj; p.K = d; p.lda = d; p.ldc = n_total;
  p.bias = b.dev + row0; p.out_fp32 = true;
  gemm_fp32(p, 0);
  if (n_total > n_proj) {
    const int n = rows * (n_total - n_proj);
    fill_cols_kernel<<<(unsigned)cdiv(n, 256), 256>>>(t, rows, n_total, n_proj, n_total - n_proj, b.dev + row0 + n_proj);
    LAUNCH_OK();
  }
  return t;
}

template <typename T>
float* EngineT<T>::bias_tail(const RawWeight& b, int row0, int n_proj, int n_total) {
  float* t = static_cast<float*>(dalloc((size_t)n_total * 4));
  CUDA_OK(cudaMemset(t, 0, (size_t)n_total * 4));
  CUDA_OK(cudaMemcpy(t + n_proj, b.dev + row0 + n_proj, (size_t)(n_total - n_proj) * 4, cudaMemcpyDeviceToDevice));
  return t;
}

template <typename T>
void EngineT<T>::finalize() {
  RETR_CHECK(!finalized_, "finalize called twice");
  const int d = desc.d_model;
  const std::string bb = "backbone.body.";
  stem_ = conv(bb + "conv1.weight", bb + "bn1", 2, 3, 1);
  RETR_CHECK(stem_.cin == 3 && stem_.k == 7, "unexpected stem");
  int dil = 1;
  for (int st = 0; st < 4; ++st) {
    int stride = st == 0 ? 1 : 2;
    const int first_dil = dil;
    if (st == 3 && desc.dilation) { dil *= stride; stride = 1; }
    for (int bi = 0; bi < desc.blocks[st]; ++bi) {
      const std::string p = bb + "layer" + std::to_string(st + 1) + "." + std::to_string(bi) + ".";
      const int s_ = bi == 0 ? stride : 1, dl = bi == 0 ? first_dil : dil;
      Block<T> b;
      b.basic = raw_.count(p + "conv3.weight") == 0;
      if (b.basic) {          // torchvision/models/resnet.py:59-105; dilation > 1 is refused there
        RETR_CHECK(dl == 1, "BasicBlock backbones (ResNet-18/34) need dilation = False");
        b.c1 = conv(p + "conv1.weight", p + "bn1", s_, 1, 1);
        b.c2 = conv(p + "conv2.weight", p + "bn2", 1, 1, 1);
      } else {
        b.c1 = conv(p + "conv1.weight", p + "bn1", 1, 0, 1);
        b.c2 = conv(p + "conv2.weight", p + "bn2", s_, dl, dl);
        b.c3 = conv(p + "conv3.weight", p + "bn3", 1, 0, 1);
      }
      b.has_ds = raw_.count(p + "downsample.0.weight") > 0;
      if (b.has_ds) b.ds = conv(p + "downsample.0.weight", p + "downsample.1", s_, 0, 1);
      blocks_.push_back(b);
    }
  }
  if constexpr (std::is_same<T, bf16>::value) {
    const bool want = !(getenv("RETR_STEM_FUSED") && atoi(getenv("RETR_STEM_FUSED")) == 0);   // RETR_STEM_FUSED=0: im2col + GEMM + max-pool
    if (want && stem_fused_supported(desc.image_size, desc.image_size, stem_.cout, stem_.k, stem_.stride, stem_.pad)) {
      stem_wk_ = static_cast<bf16*>(dalloc((size_t)stem_fused_weight_elems() * 2));
      stem_fused_repack(raw(bb + "conv1.weight").dev, stem_wk_, 0);
    }
  }
  input_proj_ = linear("input_proj.weight", "input_proj.bias");
  RETR_CHECK(input_proj_.out == d, "input_proj width");
  if (desc.loc_in_dim) { loc_w_ = raw("loc_proj.weight").dev; loc_b_ = raw("loc_proj.bias").dev; }

  // position table of the concatenated sequence: the sine buffer ('v2'/'sine', position_encoding.py:7-35) or, for the
  // learned variant ('v3'/'learned', position_encoding.py:38-63), LayerNorm(pos_embed.weight[:S]) — in eval mode a
  // constant of the weights, so it is evaluated once here (its dropout is the identity outside training).
  const float* pe = nullptr;
  const std::string pk = "transformer.positional_encoding.";
  if (raw_.count(pk + "pe")) {
    pe = raw(pk + "pe").dev;                                           // [1024,1,d]
    RETR_CHECK(raw(pk + "pe").shape[0] >= S, "sequence longer than the sine table");
  } else {
    const RawWeight& pw = raw(pk + "pos_embed.weight");                // [1024,d]
    RETR_CHECK(pw.shape[0] >= S && pw.shape[1] == d, "sequence longer than the learned position table");
    float* t = static_cast<float*>(dalloc((size_t)S * d * 4));
    layernorm<float>(pw.dev, raw(pk + "LayerNorm.weight").dev, raw(pk + "LayerNorm.bias").dev, t, S, d, 1e-5f, 0);
    pe = t;
  }
  pos_table_ = pe;
  e_word_ = raw("transformer.embeddings.word_embeddings.weight").dev;
  e_pos_ = raw("transformer.embeddings.position_embeddings.weight").dev;
  emb_g_ = raw("transformer.embeddings.LayerNorm.weight").dev;
  emb_b_ = raw("transformer.embeddings.LayerNorm.bias").dev;
  const int Tn = desc.max_pos;
  for (int l = 0; l < desc.enc_layers; ++l) {
    const std::string p = "transformer.encoder.layers." + std::to_string(l) + ".";
    EncLayer<T> e;
    e.ln1_g = raw(p + "self_attn.norm.weight").dev; e.ln1_b = raw(p + "self_attn.norm.bias").dev;
    e.ln2_g = raw(p + "ff.norm.weight").dev; e.ln2_b = raw(p + "ff.norm.bias").dev;
    e.qkv = linear(p + "self_attn.sublayer.in_proj_weight", "");
    e.out = linear(p + "self_attn.sublayer.out_proj.weight", p + "self_attn.sublayer.out_proj.bias");
    e.ff1 = linear(p + "ff.sublayer.0.weight", p + "ff.sublayer.0.bias");
    e.ff2 = linear(p + "ff.sublayer.2.weight", p + "ff.sublayer.2.bias");
    e.rb_qkv = pos_bias_table(pe, S, raw(p + "self_attn.sublayer.in_proj_weight"),
                              raw(p + "self_attn.sublayer.in_proj_bias"), 0, 2 * d, 3 * d);
    e.b_v_tail = bias_tail(raw(p + "self_attn.sublayer.in_proj_bias"), 0, 2 * d, 3 * d);
    e.b_in = raw(p + "self_attn.sublayer.in_proj_bias").dev;
    enc_.push_back(e);
  }
  if (desc.pre_norm) { enc_norm_g_ = raw("transformer.encoder.norm.weight").dev; enc_norm_b_ = raw("transformer.encoder.norm.bias").dev; }
  for (int l = 0; l < desc.dec_layers; ++l) {
    const std::string p = "transformer.decoder.layers." + std::to_string(l) + ".";
    const std::string sa = p + "tgt_self_attn.sublayer.", ca = p + "tgt_src_cross_attn.sublayer.";
    DecLayer<T> e;
    e.ln1_g = raw(p + "tgt_self_attn.norm.weight").dev; e.ln1_b = raw(p + "tgt_self_attn.norm.bias").dev;
    e.ln2_g = raw(p + "tgt_src_cross_attn.norm.weight").dev; e.ln2_b = raw(p + "tgt_src_cross_attn.norm.bias").dev;
    e.ln3_g = raw(p + "ff.norm.weight").dev; e.ln3_b = raw(p + "ff.norm.bias").dev;
    e.s_qkv = linear(sa + "in_proj_weight", "");
    e.s_out = linear(sa + "out_proj.weight", sa + "out_proj.bias");
    e.c_q = linear(ca + "in_proj_weight", "", 0, d);
    e.c_kv = e.c_q; e.c_kv.w = e.c_q.w + (int64_t)d * d; e.c_kv.out = 2 * d;
    e.c_out = linear(ca + "out_proj.weight", ca + "out_proj.bias");
    e.ff1 = linear(p + "ff.sublayer.0.weight", p + "ff.sublayer.0.bias");
    e.ff2 = linear(p + "ff.sublayer.2.weight", p + "ff.sublayer.2.bias");
    e.rb_s_qkv = pos_bias_table(e_pos_, Tn, raw(sa + "in_proj_weight"), raw(sa + "in_proj_bias"), 0, 2 * d, 3 * d);
    e.rb_c_q = pos_bias_table(e_pos_, Tn, raw(ca + "in_proj_weight"), raw(ca + "in_proj_bias"), 0, d, d);
    e.rb_c_kv = pos_bias_table(pe, S, raw(ca + "in_proj_weight"), raw(ca + "in_proj_bias"), d, d, 2 * d);
    e.b_cv_tail = bias_tail(raw(ca + "in_proj_bias"), d, d, 2 * d);
    e.b_ca = raw(ca + "in_proj_bias").dev;
    dec_.push_back(e);
  }
  dec_norm_g_ = raw("transformer.decoder.norm.weight").dev; dec_norm_b_ = raw("transformer.decoder.norm.bias").dev;
  for (int i = 0; i < 3; ++i)
    mlp_[i] = linear("mlp.layers." + std::to_string(i) + ".weight", "mlp.layers." + std::to_string(i) + ".bias");
  RETR_CHECK(mlp_[2].out == desc.vocab, "vocab size mismatch");
  alloc_buffers();
  CUDA_OK(cudaDeviceSynchronize());
  finalized_ = true;
}


// ---- buffers ---------------------------------------------------------------------------------------
template <typename T>
void EngineT<T>::alloc_buffers() {
  const int64_t B = desc.max_batch, d = desc.d_model, Tn = desc.max_pos, ff = desc.d_ff;
  const int64_t nimg = B * desc.n_images;
  if (const char* e = getenv("RETR_BACKBONE_CHUNK")) chunk_ = std::max(1, atoi(e));
  chunk_ = (int)std::min<int64_t>(chunk_, nimg);
  const int64_t h1 = (desc.image_size + 6 - 7) / 2 + 1;           // stem output side
  const int64_t act = (int64_t)chunk_ * h1 * h1 * 64;              // largest activation of a chunk
  im2col_ = static_cast<T*>(dalloc(chunk_ * h1 * h1 * stem_.kpad * sizeof(T)));
  for (auto& a : act_) a = static_cast<T*>(dalloc(act * sizeof(T)));
  const int64_t g2 = (int64_t)grid_hw * grid_hw, cfeat = blocks_.back().out_channels();
  feat_ = static_cast<T*>(dalloc(nimg * g2 * cfeat * sizeof(T)));
  const int64_t Me = B * S, Md = B * Tn, Mx = std::max(Me, Md);
  x_ = static_cast<float*>(dalloc(Me * d * 4));
  src_copy_ = static_cast<float*>(dalloc(Me * d * 4));
  y_ = static_cast<float*>(dalloc(Md * d * 4));
  n_ = static_cast<T*>(dalloc(Mx * d * sizeof(T)));
  qkv_ = static_cast<T*>(dalloc(Mx * 3 * d * sizeof(T)));
  att_ = static_cast<T*>(dalloc(Mx * d * sizeof(T)));
  h_ = static_cast<T*>(dalloc(Mx * ff * sizeof(T)));
  memory_ = static_cast<T*>(dalloc(Me * d * sizeof(T)));
  qc_ = static_cast<T*>(dalloc(Md * d * sizeof(T)));
  hs_ = static_cast<T*>(dalloc(Md * d * sizeof(T)));
  m1_ = static_cast<T*>(dalloc(Md * mlp_[0].out * sizeof(T)));
  m2_ = static_cast<T*>(dalloc(Md * mlp_[1].out * sizeof(T)));
  for (int l = 0; l < desc.dec_layers; ++l) {
    ckv_.push_back(static_cast<T*>(dalloc(Me * 2 * d * sizeof(T))));
    ck_.push_back(static_cast<T*>(dalloc(Me * d * sizeof(T))));
    cv_.push_back(static_cast<T*>(dalloc(Me * d * sizeof(T))));
    sk_.push_back(static_cast<T*>(dalloc(Md * d * sizeof(T))));
    sv_.push_back(static_cast<T*>(dalloc(Md * d * sizeof(T))));
  }
  tok_mask_ = static_cast<uint8_t*>(dalloc(Me));
  cross_dst_ = static_cast<int*>(dalloc(Me * 4));
  cross_nvalid_ = static_cast<int*>(dalloc((size_t)B * 4));
  seq_off_ = static_cast<int*>(dalloc((size_t)(B + 1) * 4));
  row_pos_ = static_cast<int*>(dalloc(Me * 4));
  finished_ = static_cast<uint8_t*>(dalloc(B));
  caption_ = static_cast<int64_t*>(dalloc(Md * 8));
  step_ = static_cast<int*>(dalloc(16));
  done_ = step_ + 1;
  argmax_keys_ = static_cast<unsigned long long*>(dalloc(Md * 8));
  CUDA_OK(cudaMemset(argmax_keys_, 0, Md * 8));
  vocab_ticket_ = static_cast<unsigned int*>(dalloc(256));
  CUDA_OK(cudaMemset(vocab_ticket_, 0, 256));
  CUDA_OK(cudaMallocHost(&done_host_, 16));
  if (const char* e = getenv("RETR_CROSS_PREFETCH")) cross_prefetch_ = atoi(e);
  if (const char* e = getenv("RETR_DECODE_GROUPS")) groups_ = std::min(16, std::max(1, atoi(e)));
  ffn_fused_ = std::is_same<T, bf16>::value && ffn_fused_supported((int)d, (int)desc.d_ff) &&
               !(getenv("RETR_FFN_FUSED") && atoi(getenv("RETR_FFN_FUSED")) == 0);
  if (ffn_fused_) ffn_fused_configure();
  splitk_group_stride_ = (size_t)8 * B * std::max<int64_t>(d, 512);
  splitk_scratch_ = static_cast<float*>(dalloc((size_t)groups_ * splitk_group_stride_ * 4));
  splitk_tickets_ = static_cast<unsigned*>(dalloc((size_t)groups_ * 1024 * 4));
  CUDA_OK(cudaMemset(splitk_tickets_, 0, (size_t)groups_ * 1024 * 4));
  for (int g = 1; g < groups_; ++g) {
    cudaStream_t st; cudaEvent_t ev;
    CUDA_OK(cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, decode_priority()));
    CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    gstreams_.push_back(st); gjoin_.push_back(ev);
  }
  CUDA_OK(cudaEventCreateWithFlags(&gfork_, cudaEventDisableTiming));
  fast_decode_ = std::is_same<T, bf16>::value && getenv("RETR_NO_FAST_DECODE") == nullptr && (d == 256 || d == 512) &&
                 ff % 256 == 0;
  if (fast_decode_) small_gemm_configure();
  stack_decode_ = fast_decode_ && stack_decode_wanted() &&
                  decode_stack_supported((int)d, desc.n_heads, (int)ff, desc.dec_layers, mlp_[0].out, (int)Tn) &&
                  mlp_[1].out == mlp_[0].out && mlp_[2].in == mlp_[1].out;
  packed_encoder_ = stack_decode_ && !(getenv("RETR_ENCODER_PACKED") && atoi(getenv("RETR_ENCODER_PACKED")) == 0);
  if (packed_encoder_) {
    xp_ = static_cast<float*>(dalloc(Me * d * 4));
    CUDA_OK(cudaMemset(xp_, 0, Me * d * 4));
    // positions are added to the projections' INPUT (a second LayerNorm output) instead of through a per-row bias
    // table in the GEMM epilogue, and the cross K / V projections of all decoder layers become two GEMMs
    const int L = desc.dec_layers;
    n2_ = static_cast<T*>(dalloc(Me * d * sizeof(T)));
    mempos_ = static_cast<T*>(dalloc(Me * d * sizeof(T)));
    kall_ = static_cast<T*>(dalloc(Me * L * d * sizeof(T)));
    vall_ = static_cast<T*>(dalloc(Me * L * d * sizeof(T)));
    wk_all_ = static_cast<T*>(dalloc((size_t)L * d * d * sizeof(T)));
    wv_all_ = static_cast<T*>(dalloc((size_t)L * d * d * sizeof(T)));
    bk_all_ = static_cast<float*>(dalloc((size_t)L * d * 4));
    bv_all_ = static_cast<float*>(dalloc((size_t)L * d * 4));
    for (int l = 0; l < L; ++l) {
      const size_t wb = (size_t)d * d * sizeof(T);
      CUDA_OK(cudaMemcpy(wk_all_ + (size_t)l * d * d, dec_[l].c_kv.w, wb, cudaMemcpyDeviceToDevice));
      CUDA_OK(cudaMemcpy(wv_all_ + (size_t)l * d * d, dec_[l].c_kv.w + (size_t)d * d, wb, cudaMemcpyDeviceToDevice));
      CUDA_OK(cudaMemcpy(bk_all_ + (size_t)l * d, dec_[l].b_ca + d, (size_t)d * 4, cudaMemcpyDeviceToDevice));
      CUDA_OK(cudaMemcpy(bv_all_ + (size_t)l * d, dec_[l].b_ca + 2 * d, (size_t)d * 4, cudaMemcpyDeviceToDevice));
    }
  }
  if (stack_decode_) {
    cross_chunks_ = (int)cdiv(S, DS_CHUNK_KEYS);
    self_chunks_ = (int)Tn / DS_CHUNK_KEYS;
    for (int l = 0; l < desc.dec_layers; ++l) {
      // zero-filled once: rows past a sample's valid keys are multiplied by p = 0 and must stay finite
      const size_t cb = (size_t)B * desc.n_heads * cross_chunks_ * DS_CHUNK_ELEMS * 2;
      const size_t sb = (size_t)B * desc.n_heads * self_chunks_ * DS_CHUNK_ELEMS * 2;
      ckv2_.push_back(static_cast<bf16*>(dalloc(cb)));
      skv2_.push_back(static_cast<bf16*>(dalloc(sb)));
      CUDA_OK(cudaMemset(ckv2_.back(), 0, cb));
      CUDA_OK(cudaMemset(skv2_.back(), 0, sb));
    }
    decode_stack_configure();
    build_stack_params();
  }
  if (getenv("RETR_NO_CARVEOUT_HINT") == nullptr) ops_prefer_max_shared();
  decode_attention_configure<T>(std::max<int>(S, (int)Tn), (int)(d / desc.n_heads));
}

template <> void EngineT<float>::build_stack_params() {}
template <> void EngineT<bf16>::build_stack_params() {
  DecStackParams& p = stack_params_;
  p = DecStackParams{};
  p.n_layers = desc.dec_layers;
  auto frag = [&](const Linear<bf16>& l) {                   // private fragment-ordered copy of a weight matrix
    bf16* out = static_cast<bf16*>(dalloc((size_t)l.out * l.in * 2));
    repack_frag(l.w, out, l.out, l.in, 0);
    return out;
  };
  for (int l = 0; l < desc.dec_layers; ++l) {
    const DecLayer<bf16>& e = dec_[l];
    DecStackLayer& L = p.L[l];
    L.ln1_g = e.ln1_g; L.ln1_b = e.ln1_b; L.ln2_g = e.ln2_g; L.ln2_b = e.ln2_b; L.ln3_g = e.ln3_g; L.ln3_b = e.ln3_b;
    L.w_sqkv = frag(e.s_qkv); L.rb_sqkv = e.rb_s_qkv;
    L.w_so = frag(e.s_out); L.b_so = e.s_out.b;
    L.w_cq = frag(e.c_q); L.rb_cq = e.rb_c_q;
    L.w_co = frag(e.c_out); L.b_co = e.c_out.b;
    L.w_ff1 = frag(e.ff1); L.b_ff1 = e.ff1.b; L.w_ff2 = frag(e.ff2); L.b_ff2 = e.ff2.b;
    L.self_kv = skv2_[l]; L.cross_kv = ckv2_[l];
  }
  p.caption = caption_;
  p.e_word = e_word_; p.e_pos = e_pos_; p.emb_g = emb_g_; p.emb_b = emb_b_; p.emb_eps = desc.embed_ln_eps;
  p.fin_g = dec_norm_g_; p.fin_b = dec_norm_b_;
  p.w_m0 = frag(mlp_[0]); p.b_m0 = mlp_[0].b; p.w_m1 = frag(mlp_[1]); p.b_m1 = mlp_[1].b;
  p.m2_out = m2_;
  p.step = step_; p.n_valid = cross_nvalid_;
  p.self_chunks = self_chunks_; p.cross_chunks = cross_chunks_; p.d_ff = desc.d_ff;
  // measured at B = 256 (decode ms per 127 steps): no stagger 30.0; 2 groups x 8k cycles 29.7, 16k 28.3, 24k 28.8, 32k 30.1;
  // 4 groups x 8k 28.6, 16k 29.6.  L2 prefetch ahead of the ring (RETR_STACK_PREFETCH blocks per warp): 0 -> 30.5, 4 -> 32.6,
  // 8 -> 31.2, 12 -> 31.9, 16 -> 32.6 — not kept.
  p.stagger_cycles = getenv("RETR_STACK_STAGGER") ? atoi(getenv("RETR_STACK_STAGGER")) : 16000;
  p.stagger_groups = getenv("RETR_STACK_STAGGER_GROUPS") ? std::max(1, atoi(getenv("RETR_STACK_STAGGER_GROUPS"))) : 2;
  p.prefetch_blocks = getenv("RETR_STACK_PREFETCH") ? std::min(32, std::max(0, atoi(getenv("RETR_STACK_PREFETCH")))) : 0;
}

// ---- backbone --------------------------------------------------------------------------------------
template <typename T>
void EngineT<T>::run_conv(const Conv<T>& c, const T* in, int64_t n, int H, int W, T* out, const T* residual,
                          int relu, int* OH, int* OW, cudaStream_t s) {
  const int oh = (H + 2 * c.pad - c.dil * (c.k - 1) - 1) / c.stride + 1;
  const int ow = (W + 2 * c.pad - c.dil * (c.k - 1) - 1) / c.stride + 1;
  GemmParams p;
  p.A = in; p.W = c.w; p.C = out;
  p.M = n * oh * ow; p.N = c.cout; p.K = c.kpad; p.lda = c.cin; p.ldc = c.cout;
  p.scale = c.scale; p.bias = c.shift; p.relu = relu;
  p.residual = residual; p.ldr = c.cout; p.res_fp32 = std::is_same<T, float>::value;
  if (c.k != 1 || c.stride != 1) {
    p.conv.kh = p.conv.kw = c.k; p.conv.stride = c.stride; p.conv.pad = c.pad; p.conv.dil = c.dil;
    p.conv.H = H; p.conv.W = W; p.conv.C = c.cin; p.conv.OH = oh; p.conv.OW = ow;
  }
  gemm<T>(p, s);
  *OH = oh; *OW = ow;
}

// img: fp32 NCHW [n,3,Hi,Wi] (n <= chunk_) -> feat NHWC [n, g, g, C4]
template <typename T>
void EngineT<T>::backbone(const float* img, int64_t n, T* feat, cudaStream_t s, const float* img2, int64_t n2) {
  const int Hi = desc.image_size;
  int H = (Hi + 6 - 7) / 2 + 1, W = H;
  const int PH = (H + 2 - 3) / 2 + 1;
  bool fused_stem = false;
  if constexpr (std::is_same<T, bf16>::value) {
    if (stem_wk_) {        // conv + BN + ReLU + max-pool in one kernel: no im2col matrix, no full-resolution conv output
      stem_fused(img, stem_wk_, stem_.scale, stem_.shift, act_[1], n, Hi, Hi, s);
      if (img2) {
        stem_fused(img2, stem_wk_, stem_.scale, stem_.shift, act_[1] + n * PH * PH * stem_.cout, n2, Hi, Hi, s);
        n += n2;
      }
      fused_stem = true;
    }
  }
  if (!fused_stem) {
    stem_im2col<T>(img, im2col_, n, Hi, Hi, H, W, stem_.kpad, s);
    if (img2) {          // a second image set rides in the same pass: every conv launch then covers n + n2 images
      stem_im2col<T>(img2, im2col_ + n * H * W * stem_.kpad, n2, Hi, Hi, H, W, stem_.kpad, s);
      n += n2;
    }
    GemmParams p;
    p.A = im2col_; p.W = stem_.w; p.C = act_[0];
    p.M = n * H * W; p.N = stem_.cout; p.K = stem_.kpad; p.lda = stem_.kpad; p.ldc = stem_.cout;
    p.scale = stem_.scale; p.bias = stem_.shift; p.relu = 1;
    gemm<T>(p, s);
    maxpool_3x3s2<T>(act_[0], act_[1], n, H, W, stem_.cout, PH, PH, s);
  }
  H = W = PH;
  T* x = act_[1];
  T* other = act_[0];
  for (size_t bi = 0; bi < blocks_.size(); ++bi) {
    const Block<T>& b = blocks_[bi];
    int h1, w1, h2, w2, h3, w3;
    T* out = (bi + 1 == blocks_.size()) ? feat : other;
    run_conv(b.c1, x, n, H, W, act_[2], nullptr, 1, &h1, &w1, s);
    const T* identity = x;
    if (b.has_ds) {
      int hd, wd;
      run_conv(b.ds, x, n, H, W, act_[4], nullptr, 0, &hd, &wd, s);
      identity = act_[4];
    }
    if (b.basic) {
      run_conv(b.c2, act_[2], n, h1, w1, out, identity, 2, &h3, &w3, s);      // relu(bn2(conv2) + identity)
    } else {
      run_conv(b.c2, act_[2], n, h1, w1, act_[3], nullptr, 1, &h2, &w2, s);
      run_conv(b.c3, act_[3], n, h2, w2, out, identity, 2, &h3, &w3, s);      // relu(bn3(conv3) + identity)
    }
    other = x; x = out;
    H = h3; W = w3;
  }
  RETR_CHECK(H == grid_hw && W == grid_hw, "backbone output side mismatch");
}

// ---- encode ------------------------------------------------------------------------------------------
template <typename T>
void EngineT<T>::encode(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                        const float* loc, cudaStream_t s) {
  RETR_CHECK(finalized_, "call retr_finalize_weights first");
  RETR_CHECK(B > 0 && B <= desc.max_batch, "batch exceeds max_batch");
  RETR_CHECK(t_img && t_mask, "target images and masks required");
  RETR_CHECK(desc.n_images == 1 || (g_img && g_mask), "context images and masks required");
  RETR_CHECK(desc.loc_in_dim == 0 || loc, "location features required");
  const long long l0 = g_launches;
  const int d = desc.d_model, g2 = grid_hw * grid_hw, Hi = desc.image_size;
  const int64_t cfeat = blocks_.back().out_channels(), img_elems = 3LL * Hi * Hi;
  struct CapGuard { explicit CapGuard(int c) { gemm_set_cta_cap(c); } ~CapGuard() { gemm_set_cta_cap(0); } } cap_guard(encode_cta_cap);
  CUDA_OK(cudaEventRecord(ev_[0], s));
  // Per-launch ramp and tail cost ~18 us x 104 convs per pass, so both image sets go through the backbone in ONE pass
  // when they are already on the device (measured 22.5 -> 20.7 ms at B = 256); generate_host keeps two passes so the
  // context images can still be in flight on the copy engine during the first.
  const bool one_pass = desc.n_images == 2 && !ctx_wait_ && 2 * (int64_t)B <= chunk_;
  if (one_pass) backbone(t_img, B, feat_, s, g_img, B);
  for (int set = 0; set < desc.n_images && !one_pass; ++set) {
    const float* img = set == 0 ? t_img : g_img;
    if (set == 1 && ctx_wait_) CUDA_OK(cudaStreamWaitEvent(s, ctx_ready_, 0));   // generate_host: context copy in flight
    for (int64_t i0 = 0; i0 < B; i0 += chunk_) {
      const int64_t n = std::min<int64_t>(chunk_, B - i0);
      backbone(img + i0 * img_elems, n, feat_ + ((int64_t)set * B + i0) * g2 * cfeat, s);
    }
  }
  CUDA_OK(cudaEventRecord(ev_[1], s));
  // input_proj straight into the concatenated token buffer (models/caption.py:119-123, 139-143)
  const int off_loc = g2, off_g = g2 + n_loc_tok;
  for (int set = 0; set < desc.n_images; ++set) {
    GemmParams p = lin(input_proj_, feat_ + (int64_t)set * B * g2 * cfeat, (int64_t)B * g2, x_ + (int64_t)(set ? off_g : 0) * d);
    p.out_fp32 = true; p.ldc = d; p.c_group = g2; p.c_group_stride = S;
    gemm<T>(p, s);
    subsample_mask(set == 0 ? t_mask : g_mask, tok_mask_, B, Hi, Hi, grid_hw, grid_hw, S, set ? off_g : 0, s);
  }
  if (desc.loc_in_dim)
    loc_tokens(loc, loc_w_, loc_b_, x_, tok_mask_, B, desc.n_loc, desc.loc_in_dim == 1, d, S, off_loc, s);
  const int64_t Me = (int64_t)B * S;
  key_compaction(tok_mask_, B, S, cross_dst_, cross_nvalid_, s);
  // Packed encoder: a padded token is never a key (src_key_padding_mask in every encoder layer, memory_key_padding_mask
  // in every decoder layer: models/ConcatTransformer.py:171-194, 220-257) and every other operation is row-wise, so
  // the rows of padded tokens influence nothing that is returned.  The unpadded rows of all samples are gathered into
  // one dense [sum n_valid, d] matrix (~72 % of B*S on the benchmark's masks); its height exists on the device only, so
  // the GEMMs are sized for B*S rows and stop after ceil(*m_dyn / 128) row blocks, the positional row bias is looked
  // up through row_pos_, and the self-attention runs on variable-length sequences without any key mask.
  const bool packed = packed_encoder_ && !att_enc_capture;
  // the encoder input for read_state("src"): the packed encoder leaves x_ untouched, the padded one works in place
  src_in_x_ = packed;
  if (!packed) CUDA_OK(cudaMemcpyAsync(src_copy_, x_, Me * d * 4, cudaMemcpyDeviceToDevice, s));
  float* xs = x_;
  const int* m_dyn = nullptr;
  const int* ridx = nullptr;
  if (packed) {
    seq_offsets(cross_nvalid_, B, seq_off_, s);
    pack_rows(x_, xp_, cross_dst_, seq_off_, row_pos_, B, S, d, s);
    xs = xp_; m_dyn = seq_off_ + B; ridx = row_pos_;
  }
  // encoder (models/ConcatTransformer.py:171-194)
  for (int l = 0; l < desc.enc_layers; ++l) {
    const EncLayer<T>& e = enc_[l];
    GemmParams p;
    if (packed) {
      // q = k = (LN(x) + pos) W^T, v = LN(x) W^T (transformer_modules.py:33-36): the LayerNorm writes both inputs
      layernorm<T>(xs, e.ln1_g, e.ln1_b, n_, Me, d, 1e-5f, s, m_dyn, n2_, pos_table_, row_pos_);
      p = lin(e.qkv, n2_, Me, qkv_);
      p.N = 2 * d; p.ldc = 3 * d; p.bias = e.b_in; p.m_dyn = m_dyn;
      gemm<T>(p, s);
      p = lin(e.qkv, n_, Me, qkv_ + 2 * d);
      p.W = e.qkv.w + (int64_t)2 * d * d; p.N = d; p.ldc = 3 * d; p.bias = e.b_in + 2 * d; p.m_dyn = m_dyn;
      gemm<T>(p, s);
    } else {
      layernorm<T>(xs, e.ln1_g, e.ln1_b, n_, Me, d, 1e-5f, s, m_dyn);
      p = lin(e.qkv, n_, Me, qkv_);
      p.rowbias = e.rb_qkv; p.rowbias_period = S;
      gemm<T>(p, s);
    }
    AttnParams a{};
    a.q = qkv_; a.k = qkv_ + d; a.v = qkv_ + 2 * d; a.out = att_;
    a.B = B; a.Lq = S; a.Lk = S; a.H = desc.n_heads;
    a.ldq = a.ldk = a.ldv = 3 * d; a.ldo = d; a.key_pad = tok_mask_; a.causal = 0;
    a.probs = att_enc_capture ? att_enc_capture + (int64_t)l * B * S * S : nullptr;
    if (packed) { a.key_pad = nullptr; a.seq_off = seq_off_; a.seq_len = cross_nvalid_; }
    attention<T>(a, s);
    p = lin(e.out, att_, Me, xs);
    p.out_fp32 = true; p.residual = xs; p.ldr = d; p.res_fp32 = true; p.m_dyn = m_dyn;
    gemm<T>(p, s);
    layernorm<T>(xs, e.ln2_g, e.ln2_b, n_, Me, d, 1e-5f, s, m_dyn);
    p = lin(e.ff1, n_, Me, h_);
    p.relu = 1; p.m_dyn = m_dyn;
    gemm<T>(p, s);
    p = lin(e.ff2, h_, Me, xs);
    p.out_fp32 = true; p.residual = xs; p.ldr = d; p.res_fp32 = true; p.m_dyn = m_dyn;
    gemm<T>(p, s);
  }
  const bool kv_batched = packed && desc.pre_norm;
  if (kv_batched) {
    layernorm<T>(xs, enc_norm_g_, enc_norm_b_, memory_, Me, d, 1e-5f, s, m_dyn, mempos_, pos_table_, row_pos_);
  } else if (desc.pre_norm) {
    layernorm<T>(xs, enc_norm_g_, enc_norm_b_, memory_, Me, d, 1e-5f, s, m_dyn);
  } else {
    from_f32_kernel<T><<<(unsigned)cdiv(Me * d, 256), 256, 0, s>>>(xs, memory_, Me * d);
    LAUNCH_OK();
  }
  // cross-attention K/V of every decoder layer, once per sample (k = memory+pos, v = memory)
  if constexpr (std::is_same<T, bf16>::value) {
    if (kv_batched) {
      // all layers at once: K = (memory + pos) . [Wk_0 .. Wk_L]^T, V = memory . [Wv_0 .. Wv_L]^T, then every layer's cache
      const int L = desc.dec_layers;
      GemmParams p;
      p.A = mempos_; p.W = wk_all_; p.C = kall_; p.M = Me; p.N = L * d; p.K = d; p.lda = d; p.ldc = L * d; p.bias = bk_all_;
      p.m_dyn = m_dyn;
      gemm<T>(p, s);
      p.A = memory_; p.W = wv_all_; p.C = vall_; p.bias = bv_all_;
      gemm<T>(p, s);
      kv_all_to_chunked(kall_, vall_, ckv2_.data(), L, B, S, cross_chunks_, cross_dst_, seq_off_, s);
    }
  }
  for (int l = 0; l < desc.dec_layers && !kv_batched; ++l) {
    GemmParams p = lin(dec_[l].c_kv, memory_, Me, ckv_[l]);
    p.rowbias = dec_[l].rb_c_kv; p.rowbias_period = S; p.rowbias_idx = ridx; p.m_dyn = m_dyn;
    if (packed) { p.rowbias_cols = d; p.bias = dec_[l].b_cv_tail; }
    gemm<T>(p, s);
    if constexpr (std::is_same<T, bf16>::value) {
      if (stack_decode_) {
        kv_to_chunked(ckv_[l], ckv2_[l], B, S, cross_chunks_, cross_dst_, s, packed ? seq_off_ : nullptr);
        continue;
      }
    }
    kv_to_head_major<T>(ckv_[l], ck_[l], cv_[l], B, S, desc.n_heads, d / desc.n_heads, s, cross_dst_);
  }
  enc_packed_ = packed;
  CUDA_OK(cudaEventRecord(ev_[2], s));
  cur_B_ = B;
  launches += g_launches - l0;
}


// The full-sequence consumers (teacher-forced decoder, read_state) read memory_ and ckv_ as [B*S] rows: after a packed
// encode the memory is scattered back (padded rows zero: they are masked keys) and the cross K/V projections re-run.
template <typename T>
void EngineT<T>::materialize_unpacked(cudaStream_t s) {
  if (!enc_packed_) return;
  const int B = cur_B_, d = desc.d_model;
  const int64_t Me = (int64_t)B * S;
  unpack_rows<T>(memory_, n_, cross_dst_, seq_off_, B, S, d, s);
  CUDA_OK(cudaMemcpyAsync(memory_, n_, Me * d * sizeof(T), cudaMemcpyDeviceToDevice, s));
  for (int l = 0; l < desc.dec_layers; ++l) {
    GemmParams p = lin(dec_[l].c_kv, memory_, Me, ckv_[l]);
    p.rowbias = dec_[l].rb_c_kv; p.rowbias_period = S;
    gemm<T>(p, s);
  }
  enc_packed_ = false;
}

// ---- teacher-forced decoder + head (models/ConcatTransformer.py:63-74, models/caption.py:153) ------------
template <typename T>
void EngineT<T>::teacher_forced(const int64_t* ids, const uint8_t* id_mask, float* logits, float* att_self,
                                float* att_cross, cudaStream_t s) {
  RETR_CHECK(cur_B_ > 0, "retr_encode must run before the decoder");
  const long long l0 = g_launches;
  materialize_unpacked(s);
  const int B = cur_B_, d = desc.d_model, Tn = desc.max_pos;
  const int64_t Md = (int64_t)B * Tn;
  embed_ln(ids, Tn, e_word_, e_pos_, emb_g_, emb_b_, y_, B, Tn, 0, nullptr, d, desc.embed_ln_eps, s);
  for (int l = 0; l < desc.dec_layers; ++l) {
    const DecLayer<T>& e = dec_[l];
    layernorm<T>(y_, e.ln1_g, e.ln1_b, n_, Md, d, 1e-5f, s);
    GemmParams p = lin(e.s_qkv, n_, Md, qkv_);
    p.rowbias = e.rb_s_qkv; p.rowbias_period = Tn;
    gemm<T>(p, s);
    AttnParams a{};
    a.q = qkv_; a.k = qkv_ + d; a.v = qkv_ + 2 * d; a.out = att_;
    a.B = B; a.Lq = Tn; a.Lk = Tn; a.H = desc.n_heads;
    a.ldq = a.ldk = a.ldv = 3 * d; a.ldo = d; a.key_pad = id_mask; a.causal = 1;
    a.probs = att_self ? att_self + (int64_t)l * B * Tn * Tn : nullptr;
    attention<T>(a, s);
    p = lin(e.s_out, att_, Md, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);

    layernorm<T>(y_, e.ln2_g, e.ln2_b, n_, Md, d, 1e-5f, s);
    p = lin(e.c_q, n_, Md, qc_);
    p.rowbias = e.rb_c_q; p.rowbias_period = Tn;
    gemm<T>(p, s);
    a = AttnParams{};
    a.q = qc_; a.k = ckv_[l]; a.v = ckv_[l] + d; a.out = att_;
    a.B = B; a.Lq = Tn; a.Lk = S; a.H = desc.n_heads;
    a.ldq = d; a.ldk = a.ldv = 2 * d; a.ldo = d; a.key_pad = tok_mask_; a.causal = 0;
    a.probs = att_cross ? att_cross + (int64_t)l * B * Tn * S : nullptr;
    attention<T>(a, s);
    p = lin(e.c_out, att_, Md, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);

    layernorm<T>(y_, e.ln3_g, e.ln3_b, n_, Md, d, 1e-5f, s);
    p = lin(e.ff1, n_, Md, h_);
    p.relu = 1;
    gemm<T>(p, s);
    p = lin(e.ff2, h_, Md, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);
  }
  layernorm<T>(y_, dec_norm_g_, dec_norm_b_, hs_, Md, d, 1e-5f, s);
  GemmParams p = lin(mlp_[0], hs_, Md, m1_); p.relu = 1; gemm<T>(p, s);
  p = lin(mlp_[1], m1_, Md, m2_); p.relu = 1; gemm<T>(p, s);
  p = lin(mlp_[2], m2_, Md, logits); p.out_fp32 = true; gemm<T>(p, s);
  launches += g_launches - l0;
}

// ---- one KV-cached greedy step; position = *step_ (device) --------------------------------------------------
template <typename T>
void EngineT<T>::decode_step(int B, int eos, int max_len, cudaStream_t s) {
  if (fast_decode_) return decode_step_fused(B, eos, max_len, s);
  const int d = desc.d_model, Tn = desc.max_pos, H = desc.n_heads;
  embed_ln(caption_, max_len, e_word_, e_pos_, emb_g_, emb_b_, y_, B, 1, 0, step_, d, desc.embed_ln_eps, s);
  for (int l = 0; l < desc.dec_layers; ++l) {
    const DecLayer<T>& e = dec_[l];
    layernorm<T>(y_, e.ln1_g, e.ln1_b, n_, B, d, 1e-5f, s);
    GemmParams p = lin(e.s_qkv, n_, B, qkv_);
    p.rowbias = e.rb_s_qkv; p.rowbias_dyn = step_;
    gemm<T>(p, s);
    decode_attention<T>(qkv_, 3 * d, qkv_ + d, 3 * d, sk_[l], sv_[l], Tn, att_, d, B, H, d, 0, step_, nullptr, 0, s);
    p = lin(e.s_out, att_, B, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);

    layernorm<T>(y_, e.ln2_g, e.ln2_b, n_, B, d, 1e-5f, s);
    p = lin(e.c_q, n_, B, qc_);
    p.rowbias = e.rb_c_q; p.rowbias_dyn = step_;
    gemm<T>(p, s);
    decode_attention<T>(qc_, d, nullptr, 0, ck_[l], cv_[l], S, att_, d, B, H, d, S, nullptr, nullptr, 0, s, cross_nvalid_);
    p = lin(e.c_out, att_, B, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);

    layernorm<T>(y_, e.ln3_g, e.ln3_b, n_, B, d, 1e-5f, s);
    p = lin(e.ff1, n_, B, h_);
    p.relu = 1;
    gemm<T>(p, s);
    p = lin(e.ff2, h_, B, y_);
    p.out_fp32 = true; p.residual = y_; p.ldr = d; p.res_fp32 = true;
    gemm<T>(p, s);
  }
  layernorm<T>(y_, dec_norm_g_, dec_norm_b_, hs_, B, d, 1e-5f, s);
  GemmParams p = lin(mlp_[0], hs_, B, m1_); p.relu = 1; gemm<T>(p, s);
  p = lin(mlp_[1], m1_, B, m2_); p.relu = 1; gemm<T>(p, s);
  p = lin(mlp_[2], m2_, B, nullptr); p.argmax_out = argmax_keys_; gemm<T>(p, s);     // logits never reach HBM
  greedy_update(argmax_keys_, caption_, max_len, B, eos, step_, done_, finished_, s);
}

// bf16 decode GEMM through small_gemm.cu: A (bf16) or LayerNorm(X) as the left operand.
template <>
void EngineT<float>::dgemm(const Linear<float>&, const float*, const float*, const float*, const float*, void*, bool, int,
                           const float*, int, const float*, int, cudaStream_t) {
  throw Error("fused decode step is a bf16-only path");
}
template <>
void EngineT<bf16>::dgemm(const Linear<bf16>& l, const bf16* A, const float* X, const float* g, const float* b, void* C,
                          bool out_fp32, int B, const float* rowbias, int relu, const float* residual, int split_k,
                          cudaStream_t s) {
  SmallGemmParams p;
  p.A = A; p.lda = l.in; p.X = X; p.ldx = l.in; p.gamma = g; p.beta = b; p.eps = 1e-5f;
  p.W = l.w; p.C = C; p.ldc = l.out; p.out_fp32 = out_fp32;
  p.M = B; p.N = l.out; p.K = l.in;
  p.bias = l.b; p.rowbias = rowbias; p.rowbias_dyn = rowbias ? step_ : nullptr;
  p.residual = residual; p.ldr = l.out; p.relu = relu;
  p.split_k = split_k; p.scratch = splitk_scratch_ + (size_t)cur_group_ * splitk_group_stride_;
  p.tickets = splitk_tickets_ + (size_t)cur_group_ * 1024;
  small_gemm(p, s);
}

// One KV-cached greedy step, bf16, for the samples [r0, r0+n): every pre-norm LayerNorm is the prologue of
// the GEMM it feeds (8 kernels per decoder layer instead of 11), small-M mma.sync GEMMs, deterministic
// split-K for FFN2, vocabulary GEMM with the argmax in its epilogue.
template <typename T>
void EngineT<T>::decode_rows_fused(int r0, int n, int group, int max_len, cudaStream_t s) {
  const int d = desc.d_model, Tn = desc.max_pos, H = desc.n_heads;
  const int64_t ff = desc.d_ff;
  const int kpan = desc.d_ff / 256;
  int split = kpan % 4 == 0 ? 4 : (kpan % 2 == 0 ? 2 : 1);
  if (const char* e = getenv("RETR_FF2_SPLIT")) { const int v = atoi(e); if (v >= 1 && v <= 8 && kpan % v == 0) split = v; }
  cur_group_ = group;
  float* y = y_ + (int64_t)r0 * d;
  T* qkv = qkv_ + (int64_t)r0 * 3 * d;
  T* att = att_ + (int64_t)r0 * d;
  T* qc = qc_ + (int64_t)r0 * d;
  T* h = h_ + (int64_t)r0 * ff;
  T* m1 = m1_ + (int64_t)r0 * mlp_[0].out;
  T* m2 = m2_ + (int64_t)r0 * mlp_[1].out;
  embed_ln(caption_ + (int64_t)r0 * max_len, max_len, e_word_, e_pos_, emb_g_, emb_b_, y, n, 1, 0, step_, d,
           desc.embed_ln_eps, s);
  for (int l = 0; l < desc.dec_layers; ++l) {
    const DecLayer<T>& e = dec_[l];
    T* sk = sk_[l] + (int64_t)r0 * Tn * d;
    T* sv = sv_[l] + (int64_t)r0 * Tn * d;
    T* ck = ck_[l] + (int64_t)r0 * S * d;
    T* cv = cv_[l] + (int64_t)r0 * S * d;
    dgemm(e.s_qkv, nullptr, y, e.ln1_g, e.ln1_b, qkv, false, n, e.rb_s_qkv, 0, nullptr, 1, s);
    decode_attention<T>(qkv, 3 * d, qkv + d, 3 * d, sk, sv, Tn, att, d, n, H, d, 0, step_, nullptr, 0, s);
    bool fused_sq = false;
    if constexpr (std::is_same<T, bf16>::value) {
      if (ffn_fused_) {                              // out-proj + residual + LN + cross-attention query projection, one kernel
        AttnOutProjParams ap;
        ap.att = att; ap.ld_att = d; ap.Wo = e.s_out.w; ap.bo = e.s_out.b; ap.Y = y;
        ap.gamma = e.ln2_g; ap.beta = e.ln2_b; ap.Wq = e.c_q.w; ap.bq = e.c_q.b;
        ap.rowbias = e.rb_c_q; ap.rowbias_dyn = step_; ap.Q = qc; ap.ld_q = d; ap.M = n;
        attn_out_ln_proj(ap, s);
        fused_sq = true;
      }
    }
    if (!fused_sq) {
      dgemm(e.s_out, att, nullptr, nullptr, nullptr, y, true, n, nullptr, 0, y, 1, s);
      dgemm(e.c_q, nullptr, y, e.ln2_g, e.ln2_b, qc, false, n, e.rb_c_q, 0, nullptr, 1, s);
    }
    decode_attention<T>(qc, d, nullptr, 0, ck, cv, S, att, d, n, H, d, S, nullptr, nullptr, 0, s, cross_nvalid_ + r0);
    dgemm(e.c_out, att, nullptr, nullptr, nullptr, y, true, n, nullptr, 0, y, 1, s);
    if (ffn_fused_) {
      if constexpr (std::is_same<T, bf16>::value) {
        FfnFusedParams fp;
        fp.Y = y; fp.gamma = e.ln3_g; fp.beta = e.ln3_b; fp.W1 = e.ff1.w; fp.b1 = e.ff1.b; fp.W2 = e.ff2.w; fp.b2 = e.ff2.b;
        fp.M = n; fp.d_ff = (int)ff;
        // optional: the next layer's cross K/V -> L2 while HBM is idle (=2: also layer 0's for the next step).  Measured
        // SLOWER at B=256 (75 MB per layer next to ~50 MB of weights re-read every step: the prefetch evicts them)
        const int nl = l + 1 < desc.dec_layers ? l + 1 : (cross_prefetch_ >= 2 ? 0 : -1);
        if (cross_prefetch_ && H == 8 && nl >= 0) {
          fp.pf_k = ck_[nl] + (int64_t)r0 * S * d; fp.pf_v = cv_[nl] + (int64_t)r0 * S * d;
          fp.pf_nvalid = cross_nvalid_ + r0; fp.pf_rows = S; fp.pf_dh = d / H;
        }
        ffn_fused(fp, s);
      }
    } else {
      dgemm(e.ff1, nullptr, y, e.ln3_g, e.ln3_b, h, false, n, nullptr, 1, nullptr, 1, s);
      dgemm(e.ff2, h, nullptr, nullptr, nullptr, y, true, n, nullptr, 0, y, split, s);
    }
  }
  dgemm(mlp_[0], nullptr, y, dec_norm_g_, dec_norm_b_, m1, false, n, nullptr, 1, nullptr, 1, s);
  dgemm(mlp_[1], m1, nullptr, nullptr, nullptr, m2, false, n, nullptr, 1, nullptr, 1, s);
  GemmParams p = lin(mlp_[2], m2, n, nullptr); p.argmax_out = argmax_keys_ + r0; p.pdl = false;
  gemm<T>(p, s);                                                   // logits never reach HBM
}

template <typename T>
void EngineT<T>::decode_step_fused(int B, int eos, int max_len, cudaStream_t s) {
  if constexpr (std::is_same<T, bf16>::value) {
    if (stack_decode_) {
      // three launches per step: the decoder stack + first two head layers (one cluster kernel), the vocabulary
      // GEMM with the argmax in its epilogue (tcgen05), the token bookkeeping
      cur_G_ = 1;
      DecStackParams p = stack_params_;
      p.B = B; p.max_len = max_len;
      decode_stack(p, s);
      const bool own_head = !(getenv("RETR_VOCAB_ARGMAX") && atoi(getenv("RETR_VOCAB_ARGMAX")) == 0);
      if (own_head && vocab_argmax_supported(B, mlp_[2].out, mlp_[2].in)) {
        // TWO launches per step: the vocabulary slice kernel also does the token bookkeeping in its last CTA
        VocabArgmaxParams vp;
        vp.A = m2_; vp.lda = mlp_[2].in; vp.W = mlp_[2].w; vp.bias = mlp_[2].b;
        vp.M = B; vp.N = mlp_[2].out; vp.K = mlp_[2].in;
        vp.keys = argmax_keys_; vp.caption = caption_; vp.max_len = max_len; vp.eos = eos;
        vp.step = step_; vp.done = done_; vp.finished = finished_; vp.ticket = vocab_ticket_;
        vocab_argmax(vp, s);
        return;
      }
      GemmParams gp = lin(mlp_[2], m2_, B, nullptr);
      gp.argmax_out = argmax_keys_;
      gemm<T>(gp, s);                                                  // logits never reach HBM
      greedy_update(argmax_keys_, caption_, max_len, B, eos, step_, done_, finished_, s, true);
      return;
    }
  }
  // rows per group: a multiple of the small-GEMM row granularity, at least group_rows_min_
  int rows = (int)cdiv(B, groups_);
  rows = (int)cdiv(std::max(rows, group_rows_min_), 32) * 32;
  const int G = (int)cdiv(B, rows);
  cur_G_ = G;
  if (G > 1) {
    CUDA_OK(cudaEventRecord(gfork_, s));
    for (int g = 1; g < G; ++g) {
      cudaStream_t gs = gstreams_[g - 1];
      CUDA_OK(cudaStreamWaitEvent(gs, gfork_, 0));
      decode_rows_fused(g * rows, std::min(rows, B - g * rows), g, max_len, gs);
      CUDA_OK(cudaEventRecord(gjoin_[g - 1], gs));
    }
  }
  decode_rows_fused(0, std::min(rows, B), 0, max_len, s);
  for (int g = 1; g < G; ++g) CUDA_OK(cudaStreamWaitEvent(s, gjoin_[g - 1], 0));
  greedy_update(argmax_keys_, caption_, max_len, B, eos, step_, done_, finished_, s, G == 1);
}

template <typename T>
void EngineT<T>::decode_greedy(int bos, int eos, int max_len, int64_t* out_ids, int* steps_host, cudaStream_t s) {
  RETR_CHECK(cur_B_ > 0, "retr_encode must run before the decoder");
  RETR_CHECK(max_len >= 2 && max_len <= desc.max_pos, "max_len must be in [2, max_position_embeddings]");
  const int B = cur_B_;
  const long long l0 = g_launches;
#ifdef RETR_TRACE
  static unsigned long long* trace_dev = nullptr;
  const size_t trace_bytes = (size_t)8 * ((1u << 18) + 1) * 8;
  if (!trace_dev) {
    CUDA_OK(cudaMalloc(&trace_dev, trace_bytes));
    small_gemm_set_trace(trace_dev);
    decode_ops_set_trace(trace_dev);
    ffn_fused_set_trace(trace_dev);
  }
  CUDA_OK(cudaMemsetAsync(trace_dev, 0, 64, s));
#endif
  greedy_init(caption_, max_len, B, bos, step_, done_, finished_, argmax_keys_, s);
  CUDA_OK(cudaMemsetAsync(vocab_ticket_, 0, 4, s));     // a decode that was cut short must not leave a partial ticket count
  const bool use_graph = getenv("RETR_NO_GRAPH") == nullptr;
  cudaGraphExec_t exec = nullptr;
  long long per_step = 0;
  if (use_graph) {
    auto key = std::make_tuple(B, eos, max_len);
    auto it = graphs_.find(key);
    if (it == graphs_.end()) {
      cudaGraph_t graph;
      const long long c0 = g_launches;
      // capture on a private stream: the caller's stream may be the legacy default stream,
      // which cannot be captured; nothing executes during capture, so ordering is unaffected
      if (!capture_stream_) CUDA_OK(cudaStreamCreateWithPriority(&capture_stream_, cudaStreamNonBlocking, decode_priority()));
      CUDA_OK(cudaStreamBeginCapture(capture_stream_, cudaStreamCaptureModeThreadLocal));
      try {
        decode_step(B, eos, max_len, capture_stream_);
      } catch (...) {
        cudaGraph_t junk; cudaStreamEndCapture(capture_stream_, &junk);
        throw;
      }
      CUDA_OK(cudaStreamEndCapture(capture_stream_, &graph));
      CUDA_OK(cudaGraphInstantiate(&exec, graph, 0));
      CUDA_OK(cudaGraphDestroy(graph));
      graphs_[key] = exec;
      step_kernels_ = g_launches - c0;
      g_launches = c0;                       // capture is not execution
    } else {
      exec = it->second;
    }
    per_step = step_kernels_;
  }
  CUDA_OK(cudaEventRecord(ev_[2], s));
  // Launch in chunks; poll the device-side done flag one chunk behind so the queue never drains.
  const int total = max_len - 1, chunk = 16;
  int launched = 0;
  bool stop = false;
  cudaEvent_t polled = ev_[3];
  bool poll_pending = false;
  while (launched < total && !stop) {
    const int n = std::min(chunk, total - launched);
    for (int i = 0; i < n; ++i) {
      if (use_graph) { CUDA_OK(cudaGraphLaunch(exec, s)); g_launches += per_step; }
      else decode_step(B, eos, max_len, s);
    }
    launched += n;
    if (poll_pending) {
      CUDA_OK(cudaEventSynchronize(polled));
      if (done_host_[0]) stop = true;
    }
    CUDA_OK(cudaMemcpyAsync(done_host_, done_, 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaEventRecord(polled, s));
    poll_pending = true;
  }
  CUDA_OK(cudaMemcpyAsync(out_ids, caption_, (int64_t)B * max_len * 8, cudaMemcpyDeviceToDevice, s));
  CUDA_OK(cudaMemcpyAsync(done_host_ + 1, step_, 4, cudaMemcpyDeviceToHost, s));
  CUDA_OK(cudaEventRecord(polled, s));
  CUDA_OK(cudaEventSynchronize(polled));
  if (steps_host) *steps_host = done_host_[1];
#ifdef RETR_TRACE
  if (const char* f = getenv("RETR_TRACE_FILE")) {
    std::vector<unsigned long long> host(trace_bytes / 8);
    CUDA_OK(cudaMemcpy(host.data(), trace_dev, trace_bytes, cudaMemcpyDeviceToHost));
    if (FILE* fp = fopen(f, "wb")) { fwrite(host.data(), 8, 8 + 8 * std::min<size_t>(host[0], 1u << 18), fp); fclose(fp); }
  }
#endif
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, ev_[0], ev_[1]) == cudaSuccess) phase_ms[0] = ms;
  if (cudaEventElapsedTime(&ms, ev_[1], ev_[2]) == cudaSuccess) phase_ms[1] = ms;
  if (cudaEventElapsedTime(&ms, ev_[2], polled) == cudaSuccess) phase_ms[2] = ms;
  cudaGetLastError();
  launches += g_launches - l0;
}

template <typename T>
void EngineT<T>::read_state(const std::string& what, void* dst, int64_t bytes, cudaStream_t s) {
  RETR_CHECK(cur_B_ > 0, "nothing encoded yet");
  const int64_t B = cur_B_, d = desc.d_model, g2 = (int64_t)grid_hw * grid_hw, cfeat = blocks_.back().out_channels();
  auto conv_out = [&](const T* src, int64_t n) {
    RETR_CHECK(bytes >= n * 4, "destination too small for '" + what + "'");
    to_f32_kernel<T><<<(unsigned)cdiv(n, 256), 256, 0, s>>>(src, static_cast<float*>(dst), n);
    LAUNCH_OK();
  };
  if (what == "memory") { materialize_unpacked(s); conv_out(memory_, B * S * d); }
  else if (what == "feat_t") conv_out(feat_, B * g2 * cfeat);
  else if (what == "feat_g") { RETR_CHECK(desc.n_images == 2, "no context images"); conv_out(feat_ + B * g2 * cfeat, B * g2 * cfeat); }
  else if (what == "src") {
    RETR_CHECK(bytes >= B * S * d * 4, "destination too small for 'src'");
    CUDA_OK(cudaMemcpyAsync(dst, src_in_x_ ? x_ : src_copy_, B * S * d * 4, cudaMemcpyDeviceToDevice, s));
  } else if (what == "mask") {
    RETR_CHECK(bytes >= B * S, "destination too small for 'mask'");
    CUDA_OK(cudaMemcpyAsync(dst, tok_mask_, B * S, cudaMemcpyDeviceToDevice, s));
  } else {
    throw Error("unknown state '" + what + "'");
  }
}

}  // namespace

std::unique_ptr<EngineBase> make_engine(const retr_model_desc& d) {
  if (d.precision == RETR_FP32) return std::unique_ptr<EngineBase>(new EngineT<float>(d));
  if (d.precision == RETR_BF16) {
    CUDA_OK(cudaSetDevice(d.device));
    gemm_bf16_prepare();
    return std::unique_ptr<EngineBase>(new EngineT<bf16>(d));
  }
  throw Error("unknown precision");
}

}  // namespace retr
