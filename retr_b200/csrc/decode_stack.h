// One KV-cached greedy step of the WHOLE decoder stack (+ the first two head layers) as a single cluster kernel;
// see decode_stack.cu.  Replaces the ~41-launch chain per step of models/ConcatTransformer.py:220-257 at T = 1.
#pragma once
#include "common.cuh"

namespace retr {

constexpr int DS_MAX_LAYERS = 8;
constexpr int DS_ROWS = 8;           // samples per cluster (one per warp)
constexpr int DS_CHUNK_KEYS = 32;    // keys per K/V cache chunk
constexpr int DS_CHUNK_ELEMS = 2 * DS_CHUNK_KEYS * 32;   // [K 32x32 | V 32x32] bf16 = 4 KB

struct DecStackLayer {
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b, *ln3_g, *ln3_b;
  // every bf16 weight below is in FRAGMENT ORDER (repack_frag), not row-major
  const bf16* w_sqkv; const float* rb_sqkv;      // [3d, d]; [T, 3d] position + bias table (row = step)
  const bf16* w_so; const float* b_so;           // [d, d]
  const bf16* w_cq; const float* rb_cq;          // [d, d]; [T, d]
  const bf16* w_co; const float* b_co;           // [d, d]
  const bf16* w_ff1; const float* b_ff1;         // [d_ff, d]
  const bf16* w_ff2; const float* b_ff2;         // [d, d_ff]
  bf16* self_kv;                                 // [B][4][T/32][2 heads][K 32x32 | V 32x32], rows swizzled
  const bf16* cross_kv;                          // [B][4][cross_chunks][2 heads][K 32x32 | V 32x32]
};

struct DecStackParams {
  DecStackLayer L[DS_MAX_LAYERS];
  int n_layers;
  const int64_t* caption; int max_len;           // ids [B, max_len]; this step reads column *step
  const float *e_word, *e_pos, *emb_g, *emb_b; float emb_eps;
  const float *fin_g, *fin_b;                    // final decoder LayerNorm
  const bf16* w_m0; const float* b_m0;           // [512, d]
  const bf16* w_m1; const float* b_m1;           // [512, 512]
  bf16* m2_out;                                  // [B, 512] input of the vocabulary GEMM
  const int* step;                               // device: position of the token being consumed
  const int* n_valid;                            // [B] unpadded cross keys per sample
  int B, self_chunks, cross_chunks, d_ff;
  int stagger_cycles, stagger_groups;            // start delay of cluster group (blockIdx.y % groups), in SM cycles
  int prefetch_blocks;                           // 8 KB blocks per warp prefetched into L2 ahead of the ring (0 = off)
};

bool decode_stack_supported(int d_model, int n_heads, int d_ff, int n_layers, int mlp_hidden, int max_pos);
void decode_stack_configure();                   // opt-in shared memory; call outside stream capture
void decode_stack(const DecStackParams& p, cudaStream_t s);

// row-major bf16 W[N][K] -> the fragment order the stack kernel streams (decode_stack.cu: load_w)
void repack_frag(const bf16* W, bf16* out, int N, int K, cudaStream_t s);

// encoder-side cache build: rows of [K(d) | V(d)] (d = 256, 8 heads) -> chunked, swizzled, compacted cross cache
// src_off: the source rows are packed as well (row src_off[b] + rank), see engine.cu: encode
void kv_to_chunked(const bf16* kv, bf16* out, int B, int rows, int chunks, const int* dst_row, cudaStream_t s,
                   const int* src_off = nullptr);

// every decoder layer's cache in one launch from the batched K / V projections ([rows, n_layers*256] each, packed rows)
void kv_all_to_chunked(const bf16* k_all, const bf16* v_all, bf16* const* out, int n_layers, int B, int rows, int chunks,
                       const int* dst_row, const int* src_off, cudaStream_t s);

}  // namespace retr
