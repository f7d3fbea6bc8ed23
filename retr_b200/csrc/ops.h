// Non-GEMM kernels of the hot path (ops.cu, attention.cu, decode_ops.cu).
#pragma once
#include "common.cuh"

namespace retr {

void ops_prefer_max_shared();   // ops.cu: see the definition

// ---- weight repack ------------------------------------------------------------------------
void f32_to_bf16(const float* src, bf16* dst, int64_t n, cudaStream_t s);
// conv filter OIHW fp32 -> O,(H,W,I) rows of length kpad (zero padded), fp32 or bf16
template <typename T>
void repack_filter(const float* oihw, T* out, int O, int I, int kh, int kw, int kpad, cudaStream_t s);

// ---- backbone -----------------------------------------------------------------------------
// 7x7/2 pad 3 stem patches: NCHW fp32 image -> rows [(n,oh,ow)] x kpad, k = (r*7+s)*3+c
template <typename T>
void stem_im2col(const float* img, T* out, int64_t nimg, int H, int W, int OH, int OW, int kpad, cudaStream_t s);
// the whole stem (7x7/2 conv + FrozenBN + ReLU + 3x3/2 max-pool) as one kernel, bf16 (stem_fused.cu)
bool stem_fused_supported(int H, int W, int cout, int k, int stride, int pad);
int stem_fused_weight_elems();
void stem_fused_repack(const float* w_oihw, bf16* out, cudaStream_t s);
void stem_fused(const float* img, const bf16* wk, const float* scale, const float* shift, bf16* out, int64_t nimg, int H,
                int W, cudaStream_t s);
// 3x3/2 pad 1 max pool, NHWC
template <typename T>
void maxpool_3x3s2(const T* in, T* out, int64_t nimg, int H, int W, int C, int OH, int OW, cudaStream_t s);

// ---- preprocessing (preprocess.cu): raw RGB + box -> normalised 224x224 crops, masks, location features --------
void preprocess(const uint8_t* pixels, const int64_t* offsets, const int32_t* hw, const double* bb, int batch, int out,
                float* t_img, uint8_t* t_mask, float* g_img, uint8_t* g_mask, float* loc, cudaStream_t s);
int preprocess_max_side(int out);          // largest padded square side the resampler's tap arrays cover

// ---- token assembly (models/caption.py:119-147, models/ConcatTransformer.py:48-51) ------------
// mask[b, off + p] = img_mask[b, floor(i*H/gh), floor(j*W/gw)]  (F.interpolate nearest, backbone.py:75)
void subsample_mask(const uint8_t* img_mask, uint8_t* tok_mask, int B, int H, int W, int gh, int gw,
                    int S, int off, cudaStream_t s);
// location tokens: per_scalar ? x[b,off+k,:] = w[:,0]*loc[b,k]+bias  (Linear(1,d), caption.py:126-127)
//                             : x[b,off,:]   = w[:, :n_loc] . loc[b,:] + bias (Linear(7,d), caption.py:78)
void loc_tokens(const float* loc, const float* w, const float* bias, float* x, uint8_t* tok_mask, int B,
                int n_loc, int per_scalar, int d, int S, int off, cudaStream_t s);

// ---- transformer pieces -------------------------------------------------------------------------
// y[r,:] = LN(x[r,:]) ; x fp32, y fp32|bf16; rows_dyn (d = 256 / 512 only): stop at min(rows, *rows_dyn) rows
template <typename T>
void layernorm(const float* x, const float* gamma, const float* beta, T* y, int64_t rows, int d, float eps,
               cudaStream_t s, const int* rows_dyn = nullptr, T* y2 = nullptr, const float* pos = nullptr,
               const int* pos_idx = nullptr);   // y2[r,:] = LN(x[r,:]) + pos[pos_idx[r],:] (d = 256 / 512)
// y[b*rows_per+j, :] = LN_eps(E_word[ids[b*ld_ids + t0 + j]] + E_pos[t0 + j]); t0 = t0_dyn ? *t0_dyn : t0
// (DecoderEmbeddings, models/transformer_modules.py:113-129)
void embed_ln(const int64_t* ids, int64_t ld_ids, const float* e_word, const float* e_pos, const float* gamma,
              const float* beta, float* y, int B, int rows_per, int t0, const int* t0_dyn, int d, float eps,
              cudaStream_t s);

// ---- attention (attention.cu) ---------------------------------------------------------------------
struct AttnParams {
  const void* q; const void* k; const void* v; void* out;
  int B, Lq, Lk, H;
  int64_t ldq, ldk, ldv, ldo;               // row pitches (elements); batch pitch = L * ld
  const uint8_t* key_pad;                   // [B,Lk] non-zero = masked, or null
  int causal;
  float* probs;                             // optional [B,Lq,Lk] head-averaged probabilities
  // packed (variable-length) self-attention, flash_attention_bf16 only: sample b owns rows seq_off[b] .. +seq_len[b] of
  // q/k/v/out (no padded keys inside a sequence); Lq = Lk = the longest possible sequence (sizes the grid)
  const int* seq_off;
  const int* seq_len;
};
template <typename T> void attention(const AttnParams& p, cudaStream_t s);
void flash_attention_bf16(const AttnParams& p, int dh, cudaStream_t s);   // attention_mma.cu
bool attention_tc_supported(const AttnParams& p, int dh);                  // attention_tc.cu: tcgen05 / TMEM, head dim 32
void attention_tc_bf16(const AttnParams& p, cudaStream_t s);

// ---- greedy decode (decode_ops.cu) --------------------------------------------------------------
// Single-query attention over head-major K/V caches ([B][H][cache_rows][dh] each), one CTA per (b, head).
//   self: appends this step's k,v (from qkv rows [q|k|v]) at position *step, attends keys 0..*step
//   cross: attends keys 0..n_keys-1 honouring key_pad
template <typename T>
void decode_attention(const T* q, int64_t ldq, const T* new_kv, int64_t ld_new, T* kcache, T* vcache, int cache_rows,
                      T* out, int64_t ldo, int B, int H, int d, int n_keys, const int* step_dyn,
                      const uint8_t* key_pad, int64_t ld_pad, cudaStream_t s, const int* n_keys_per = nullptr);
// opt in to the dynamic shared memory the kernel needs (call once, outside stream capture)
template <typename T> void decode_attention_configure(int max_rows, int dh);
// rows of [K(d) | V(d)] -> head-major K and V blocks; with dst_row (key_compaction) only the unpadded keys, densely
template <typename T>
void kv_to_head_major(const T* kv, T* kh, T* vh, int B, int rows, int H, int dh, cudaStream_t s,
                      const int* dst_row = nullptr);
// dst_row[b*S + j] = rank of key j among sample b's unpadded keys (-1 if padded); n_valid[b] = their number
// packed encoder: exclusive offsets of the unpadded-token counts (off[B] = total), gather of the unpadded rows (row_pos =
// position of each packed row inside its sequence, 0 past the total), scatter back to [B*S, d] with zero padding
void seq_offsets(const int* n_valid, int B, int* off, cudaStream_t s);
void pack_rows(const float* x, float* xp, const int* dst_row, const int* off, int* row_pos, int B, int S, int d,
               cudaStream_t s);
template <typename T>
void unpack_rows(const T* xp, T* x, const int* dst_row, const int* off, int B, int S, int d, cudaStream_t s);
void key_compaction(const uint8_t* pad, int B, int S, int* dst_row, int* n_valid, cudaStream_t s);
// After the vocab GEMM+argmax of step i=*step: the bookkeeping of eval_utils/decode.py:70-79.
void greedy_update(unsigned long long* argmax_keys, int64_t* caption, int max_len, int B, int eos, int* step,
                   int* done, uint8_t* finished, cudaStream_t s, bool allow_pdl = true);
void greedy_init(int64_t* caption, int max_len, int B, int bos, int* step, int* done, uint8_t* finished,
                 unsigned long long* argmax_keys, cudaStream_t s);

}  // namespace retr
