// tcgen05 / TMA GEMM and implicit-GEMM convolution for sm_100a (bf16 in, fp32 accumulate in TMEM).
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0      TMA producer      W tile (and the A tile for a plain matrix) -> 128B-swizzled smem
//   warp 1      MMA issuer        one thread issues tcgen05.mma (M=128, N=BN, K=16) x4 per k-block
//   warp 2      TMEM allocator    2 accumulator stages x BN columns
//   warp 3      residual producer TMA loads of the residual sub-tiles (4-deep ring) ahead of the epilogue
//   warps 4-11  epilogue          two groups of four warps on alternate 128-byte column sub-tiles: tcgen05.ld (one
//                                 32-column half ahead of the arithmetic) -> scale/bias/rowbias/ReLU/residual
//                                 (specialised forms for the hot-path combinations) -> swizzled smem staging ->
//                                 one 32-row TMA store PER WARP, no CTA-level barrier; a direct register->global
//                                 path remains for argmax-only / remapped / unaligned C
//   warps 8-11  A gather          (AMODE 1 fallback only, instead of the second epilogue group) cp.async im2col
//                                 rows with zero fill, written in the same swizzled layout TMA would produce
// Pipelines: smem full/empty ring (TMA|gather -> MMA), TMEM full/empty pair (MMA -> epilogue) so the
// epilogue of tile i overlaps the main loop of tile i+1.
//
// Replaces, on the reference's path: every nn.Conv2d of the torchvision ResNet body with its
// FrozenBatchNorm2d/ReLU/residual (models/backbone.py:41-51,69; torchvision/models/resnet.py:143-165),
// input_proj (models/caption.py:119), all nn.Linear / MHA in/out projections
// (models/transformer_modules.py:8-10,19; torch/nn/functional.py:5798-5860) and the MLP head
// (models/caption.py:161-174).
#include <cudaTypedefs.h>

#include "engine.h"
#include "gemm.h"
#include "sm100_ptx.cuh"
#include "train_ops.h"

namespace retr {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle row
constexpr int A_TILE_BYTES = BM * BK * 2;
// cp.async groups a gather thread keeps in flight.  The thread signals stage `it - lag` only after it has issued
// group `it`, which needs stage it % S empty, so MMA(j) cannot start before MMA(j + lag - S) has retired plus one
// mbarrier round trip (~1 us): with S = 3, lag = 2 (round-1 first cut) the 3x3 convolutions ran fully serialised at
// 1.48 us per k-block (22 % tensor pipe).  Keep S - lag >= 3: lag 1 up to 5 stages, lag 2 from 6.
constexpr int MAX_GATHER_LAG = 2;

struct KArgs {
  int64_t M;
  int N, K, num_kb, m_tiles, n_tiles;
  void* C;
  int64_t ldc;
  int out_fp32;
  const float* scale;
  const float* bias;
  const float* rowbias;
  int rowbias_period;
  const void* residual;
  int64_t ldr;
  int res_fp32, relu;
  int vecC, vecR, vecRB;
  unsigned long long* argmax_out;
  const bf16* A;
  int kh, kw, stride, pad, dil, H, W, Cin, OH, OW, cblocks;
  int c_group, c_group_stride;
  const int* rowbias_dyn;
  const int* rowbias_idx;
  const int* m_dyn;
  int rowbias_cols;
  int c_dbl;             // warp-store epilogue: two staging buffers per group (short main loops: the epilogue paces the tile)
  int epi_mode;          // bf16 staged epilogue: 0 general, 1-3 the specialised forms (epilogue_chunk_lean)
  int warp_store;        // staged epilogue: every warp stores its own 32 rows (no CTA-level barriers in the epilogue)
  int stages;            // smem ring depth (runtime: depends on the epilogue staging in use)
  int tma_epi;           // epilogue through smem staging + TMA store
  int res_tma;           // residual sub-tiles arrive by TMA (implies tma_epi)
  int r_bufs;            // depth of the residual ring (<= R_BUFS): fewer buffers leave room for one more pipeline stage
  // split-K (weight-gradient GEMMs: tiny M x N, contraction over every token/pixel): tile t covers
  // k-blocks [sp*kb_per_split, ...) of output tile t / splits and writes an fp32 partial tile at row
  // offset sp*M of C (C = workspace [splits*M, N]); splitk_reduce() sums the partials.
  int splits, kb_per_split;
};

constexpr int MAX_STAGES = 8;
constexpr int EPI_TILE_BYTES = BM * 128;     // 128 rows x 128 B (64 bf16 or 32 fp32 columns)
constexpr int C_BUFS = 2, R_BUFS = 6;      // C staging: 2 buffers (1 in the gather variant: one more pipeline stage)
constexpr int SMEM_LIMIT = 232448;           // 227 KB opt-in maximum per CTA
constexpr int BARRIER_BYTES = 512;

template <int BN> struct Cfg {
  static constexpr int B_TILE_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
  static constexpr int SB_BYTES = 2 * BN * 4;
  // no alignment slack: the dynamic window of a kernel without static shared memory starts 1024-byte aligned (the
  // kernel traps if it ever does not), and the slack cost the BN = 256 tiles their fourth pipeline stage
  static int fixed_bytes(bool tma_epi, int r_bufs, bool gather, int c_dbl = 0) {
    return BARRIER_BYTES + SB_BYTES + (tma_epi ? (gather ? 1 : C_BUFS * (1 + c_dbl)) * EPI_TILE_BYTES : 0) +
           r_bufs * EPI_TILE_BYTES;
  }
  static int stages_for(bool tma_epi, int r_bufs, bool gather, int c_dbl = 0) {
    const int n = (SMEM_LIMIT - fixed_bytes(tma_epi, r_bufs, gather, c_dbl)) / STAGE_BYTES;
    return n > MAX_STAGES ? MAX_STAGES : n;
  }
};

struct TileCoord { int m_blk, n_blk, sp, kb0, kb1; };
__device__ __forceinline__ TileCoord tile_coord(const KArgs& a, int tile) {
  TileCoord t;
  t.sp = tile % a.splits;
  const int t2 = tile / a.splits;
  t.m_blk = t2 / a.n_tiles;
  t.n_blk = t2 - t.m_blk * a.n_tiles;
  t.kb0 = t.sp * a.kb_per_split;
  t.kb1 = min(a.num_kb, t.kb0 + a.kb_per_split);
  return t;
}

// One 32-column chunk of one accumulator row.
__device__ __forceinline__ void epilogue_chunk(const KArgs& a, uint32_t (&r)[32], int64_t m, int64_t mc,
                                               int64_t rb_row, int n0, float& best_v, uint32_t& best_i) {
  const int nv = min(32, a.N - n0);
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (a.scale) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (j < nv) v[j] *= __ldg(a.scale + n0 + j);
  }
  if (a.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (j < nv) v[j] += __ldg(a.bias + n0 + j);
  }
  if (a.rowbias && n0 < a.rowbias_cols) {
    const float* rb = a.rowbias + rb_row * a.N + n0;
    if (a.vecRB && nv == 32) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 t = __ldg(reinterpret_cast<const float4*>(rb) + j);
        v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (j < nv) v[j] += __ldg(rb + j);
    }
  }
  if (a.relu == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (a.residual) {
    if (a.res_fp32) {
      const float* rp = static_cast<const float*>(a.residual) + m * a.ldr + n0;
      if (a.vecR && nv == 32) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float4 t = *(reinterpret_cast<const float4*>(rp) + j);
          v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < nv) v[j] += rp[j];
      }
    } else {
      const bf16* rp = static_cast<const bf16*>(a.residual) + m * a.ldr + n0;
      if (a.vecR && nv == 32) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 t = *(reinterpret_cast<const uint4*>(rp) + j);
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float2 f = __bfloat1622float2(h[q]);
            v[8 * j + 2 * q] += f.x; v[8 * j + 2 * q + 1] += f.y;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < nv) v[j] += __bfloat162float(rp[j]);
      }
    }
  }
  if (a.relu == 2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (a.argmax_out) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nv && v[j] > best_v) { best_v = v[j]; best_i = (uint32_t)(n0 + j); }
  }
  if (a.C) {
    if (a.out_fp32) {
      float* cp = static_cast<float*>(a.C) + mc * a.ldc + n0;
      if (a.vecC && nv == 32) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *(reinterpret_cast<float4*>(cp) + j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < nv) cp[j] = v[j];
      }
    } else {
      bf16* cp = static_cast<bf16*>(a.C) + mc * a.ldc + n0;
      if (a.vecC && nv == 32) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 t;
          __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
          for (int q = 0; q < 4; ++q) h[q] = __floats2bfloat162_rn(v[8 * j + 2 * q], v[8 * j + 2 * q + 1]);
          *(reinterpret_cast<uint4*>(cp) + j) = t;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j < nv) cp[j] = __float2bfloat16_rn(v[j]);
      }
    }
  }
}

// Specialised forms of epilogue_chunk_staged<bf16> for the epilogues the hot path is made of; the general function
// carries every feature behind run-time branches (~650 instructions per chunk in the binary), and the epilogue is the
// pacing side of every GEMM with K <= 512: FF1 118 -> 96 us with the lean form.
//   MODE 1: bias, ReLU floor `lo` (0 or -inf)                      linears of the encoder / head
//   MODE 2: scale, bias, ReLU floor                                 conv1 / conv2 / downsample (FrozenBN folded)
//   MODE 3: scale, bias, + bf16 residual sub-tile, ReLU after       conv3 of a bottleneck
//   MODE 4: bias + per-row bias `rowb` (null past rowbias_cols)     q|k|v and k|v projections (positions)
//   MODE 5: scale, kept where the bf16 `residual` tile is > 0       input-gradient GEMMs with the ReLU backward fused
template <int MODE>
__device__ __forceinline__ void epilogue_chunk_lean(const uint32_t (&r)[32], const float* __restrict__ scale,
                                                    const float* __restrict__ bias, float lo, int row, int chunk16,
                                                    const uint8_t* rsm, uint8_t* csm, const float* __restrict__ rowb = nullptr) {
  const uint32_t swz = (uint32_t)(row & 7);
  uint8_t* crow = csm + row * 128;
  const uint8_t* rrow = rsm + row * 128;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[8 * c + j]);
    float4 b0 = make_float4(0.f, 0.f, 0.f, 0.f), b1 = b0;
    if (MODE != 5) {
      b0 = *reinterpret_cast<const float4*>(bias + 8 * c);
      b1 = *reinterpret_cast<const float4*>(bias + 8 * c + 4);
    }
    if (MODE == 5) {
      const float4 s0 = *reinterpret_cast<const float4*>(scale + 8 * c), s1 = *reinterpret_cast<const float4*>(scale + 8 * c + 4);
      v[0] *= s0.x; v[1] *= s0.y; v[2] *= s0.z; v[3] *= s0.w; v[4] *= s1.x; v[5] *= s1.y; v[6] *= s1.z; v[7] *= s1.w;
    } else if (MODE >= 2 && MODE != 4) {
      const float4 s0 = *reinterpret_cast<const float4*>(scale + 8 * c), s1 = *reinterpret_cast<const float4*>(scale + 8 * c + 4);
      // multiply, THEN add, each rounded: what the general epilogues compute (their scale and bias steps sit behind
      // separate branches and are never contracted), so a row's result does not depend on which epilogue its launch
      // shape selects (batch invariance: test_generate_host_matches_device_path)
      const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
      const float bi[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __fadd_rn(__fmul_rn(v[j], sc[j]), bi[j]);
    } else {
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w; v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    if (MODE == 4 && rowb) {
      const float4 r0 = __ldg(reinterpret_cast<const float4*>(rowb + 8 * c)), r1 = __ldg(reinterpret_cast<const float4*>(rowb + 8 * c + 4));
      v[0] += r0.x; v[1] += r0.y; v[2] += r0.z; v[3] += r0.w; v[4] += r1.x; v[5] += r1.y; v[6] += r1.z; v[7] += r1.w;
    }
    const uint32_t off = ((uint32_t)(chunk16 + c) ^ swz) << 4;
    if (MODE == 5) {
      const uint4 t = *reinterpret_cast<const uint4*>(rrow + off);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 f = __bfloat1622float2(h[q]);
        if (!(f.x > 0.f)) v[2 * q] = 0.f;
        if (!(f.y > 0.f)) v[2 * q + 1] = 0.f;
      }
    } else if (MODE == 3) {
      const uint4 t = *reinterpret_cast<const uint4*>(rrow + off);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 f = __bfloat1622float2(h[q]);
        v[2 * q] = fmaxf(v[2 * q] + f.x, 0.f); v[2 * q + 1] = fmaxf(v[2 * q + 1] + f.y, 0.f);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], lo);
    }
    uint4 o;
    __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int q = 0; q < 4; ++q) oh[q] = __floats2bfloat162_rn(v[2 * q], v[2 * q + 1]);
    *reinterpret_cast<uint4*>(crow + off) = o;
  }
}

// Scale/bias (+rowbias) -> ReLU -> residual (from the swizzled smem sub-tile) -> ReLU -> swizzled
// staging for the TMA store.  One 32-column chunk of one accumulator row; `chunk16` is the index of
// the first 16-byte chunk of this row segment inside the 128-byte staging row.
template <typename OutT>
__device__ __forceinline__ void epilogue_chunk_staged(const KArgs& a, uint32_t (&r)[32], const float* sb_scale,
                                                      const float* sb_bias, int64_t rb_row, int n0, int row,
                                                      int chunk16, const uint8_t* rsm, uint8_t* csm) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (a.scale) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = *reinterpret_cast<const float4*>(sb_scale + 4 * j);
      v[4 * j] *= t.x; v[4 * j + 1] *= t.y; v[4 * j + 2] *= t.z; v[4 * j + 3] *= t.w;
    }
  }
  if (a.bias) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 t = *reinterpret_cast<const float4*>(sb_bias + 4 * j);
      v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
    }
  }
  if (a.rowbias && n0 < a.rowbias_cols) {
    const float* rb = a.rowbias + rb_row * a.N + n0;
    if (a.vecRB && n0 + 32 <= a.N) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(rb) + j);
        v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (n0 + j < a.N) v[j] += __ldg(rb + j);
    }
  }
  if (a.relu == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  const uint32_t swz = (uint32_t)(row & 7);
  constexpr int NCH = sizeof(OutT) == 4 ? 8 : 4;          // 16-byte chunks covering 32 columns
  if (rsm) {
    const uint8_t* rrow = rsm + row * 128;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const uint4 t = *reinterpret_cast<const uint4*>(rrow + (((uint32_t)(chunk16 + c) ^ swz) << 4));
      if (a.relu == 3) {
        // ReLU backward: the "residual" tile is the forward activation; keep the gradient where it was positive
        if (sizeof(OutT) == 4) {
          if (!(__uint_as_float(t.x) > 0.f)) v[4 * c] = 0.f;
          if (!(__uint_as_float(t.y) > 0.f)) v[4 * c + 1] = 0.f;
          if (!(__uint_as_float(t.z) > 0.f)) v[4 * c + 2] = 0.f;
          if (!(__uint_as_float(t.w) > 0.f)) v[4 * c + 3] = 0.f;
        } else {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 f = __bfloat1622float2(h[q]);
            if (!(f.x > 0.f)) v[8 * c + 2 * q] = 0.f;
            if (!(f.y > 0.f)) v[8 * c + 2 * q + 1] = 0.f;
          }
        }
      } else if (sizeof(OutT) == 4) {
        v[4 * c] += __uint_as_float(t.x); v[4 * c + 1] += __uint_as_float(t.y);
        v[4 * c + 2] += __uint_as_float(t.z); v[4 * c + 3] += __uint_as_float(t.w);
      } else {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 f = __bfloat1622float2(h[q]);
          v[8 * c + 2 * q] += f.x; v[8 * c + 2 * q + 1] += f.y;
        }
      }
    }
  }
  if (a.relu == 2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  uint8_t* crow = csm + row * 128;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    uint4 t;
    if (sizeof(OutT) == 4) {
      t.x = __float_as_uint(v[4 * c]); t.y = __float_as_uint(v[4 * c + 1]);
      t.z = __float_as_uint(v[4 * c + 2]); t.w = __float_as_uint(v[4 * c + 3]);
    } else {
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
      for (int q = 0; q < 4; ++q) h[q] = __floats2bfloat162_rn(v[8 * c + 2 * q], v[8 * c + 2 * q + 1]);
    }
    *reinterpret_cast<uint4*>(crow + (((uint32_t)(chunk16 + c) ^ swz) << 4)) = t;
  }
}

// AMODE: how the A tile reaches shared memory.  0: plain matrix, tiled TMA.  1: implicit-GEMM convolution, rows gathered
// by cp.async (one thread per tile row; fallback, RETR_CONV_GATHER=1).  2: implicit-GEMM convolution, TMA im2col mode:
// one bulk-tensor instruction per k-block loads 128 consecutive output pixels x 64 channels of one filter tap, zero-
// filling the padding, straight into the 128B-swizzled layout — no gather threads, so the convolution keeps the plain
// variant's two epilogue warp groups.
template <int BN, int AMODE>
__global__ void __launch_bounds__(384, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmR,
                    const KArgs a) {
  using C = Cfg<BN>;
  // epilogue warp groups: a plain GEMM has two (warps 4-7 and 8-11) that take alternate 128-byte column
  // sub-tiles of the accumulator — tiles with K <= 1024 are epilogue-bound with one; the convolution variant
  // keeps one and uses warps 8-11 for the im2col gather (its K >= 576 tiles are main-loop bound).
  constexpr bool GATHER = AMODE == 1;
  constexpr bool IM2COL = AMODE == 2;
  constexpr bool TN = AMODE >= 3;          // both operands MN-major: C[M,N] = A[K,M]^T . B[K,N]   (weight gradients)
  constexpr bool TN_CONV = AMODE == 4;     // ... with B = im2col(NHWC input), loaded by TMA im2col, 64 pixels per k-block
  constexpr int EG = GATHER ? 1 : 2;
  const int STAGES = a.stages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();                  // swizzled tiles need 1024-byte alignment
  const int RB = a.r_bufs;
  uint8_t* cst = smem + STAGES * C::STAGE_BYTES;                         // C staging (TMA store source)
  uint8_t* rst = cst + (a.tma_epi ? (GATHER ? 1 : C_BUFS * (1 + a.c_dbl)) * EPI_TILE_BYTES : 0);   // residual staging (TMA load dest)
  float* sb = reinterpret_cast<float*>(rst + RB * EPI_TILE_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sb) + C::SB_BYTES);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tfull = empty + MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* rfull = tempty + 2;
  uint64_t* rempty = rfull + R_BUFS;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(rempty + R_BUFS);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int EC = a.out_fp32 ? 32 : 64;                 // columns per staged sub-tile (128 B rows)

  if (warp == 0 && lane == 0) {
    if (!GATHER) ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    if (a.tma_epi) ptx::prefetch_tmap(&tmC);
    if (a.res_tma) ptx::prefetch_tmap(&tmR);
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full[s], GATHER ? 129 : 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], 128 * EG);
    }
    for (int s = 0; s < R_BUFS; ++s) {
      ptx::mbar_init(&rfull[s], 1);
      ptx::mbar_init(&rempty[s], 128);
    }
    ptx::fence_barrier_init();
  } else if (warp == 2) {
    ptx::tmem_alloc(tmem_holder, C::TMEM_COLS);
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_holder;
  // Programmatic dependent launch: the next kernel's CTAs may take this SM as soon as this CTA leaves and run their
  // set-up (barriers, TMEM allocation, descriptor prefetch) under the tail of this grid; everything that touches
  // global memory waits for the previous grid here (operands may be its output — activations, refreshed weights).
  pdl_launch_dependents();
  pdl_wait();
  // row blocks to compute: all of them, or (packed encoder) as many as the device-side row count says
  const int m_tiles_run = a.m_dyn ? min(a.m_tiles, (__ldg(a.m_dyn) + BM - 1) / BM) : a.m_tiles;
  const int total_tiles = m_tiles_run * a.n_tiles * a.splits;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer =====
      uint32_t s = 0, ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(a, tile);
        const int m_blk = tc.m_blk, n_blk = tc.n_blk;
        int iw0 = 0, ih0 = 0, img0 = 0;
        if (IM2COL) {
          const int64_t m0 = (int64_t)m_blk * BM;
          const int64_t t = m0 / a.OW;
          iw0 = (int)(m0 - t * a.OW) * a.stride - a.pad;
          ih0 = (int)(t % a.OH) * a.stride - a.pad;
          img0 = (int)(t / a.OH);
        }
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          ptx::mbar_wait(&empty[s], ph ^ 1);
          uint8_t* As = smem + s * C::STAGE_BYTES;
          uint8_t* Bs = As + A_TILE_BYTES;
          ptx::mbar_arrive_expect_tx(&full[s], C::B_TILE_BYTES + (GATHER ? 0 : A_TILE_BYTES));
          if (TN) {
            // MN-major tiles arrive as [64 rows of K] x [64 elements of M or N] boxes of 8 KB
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              ptx::tma_load_2d(As + j * 8192, &tmA, &full[s], m_blk * BM + j * 64, kb * BK);
            int tw = 0, th = 0, tn = 0;
            if (TN_CONV) {                          // first output pixel of this k-block -> base input pixel
              const int64_t m0 = (int64_t)kb * BK;
              const int64_t t = m0 / a.OW;
              tw = (int)(m0 - t * a.OW) * a.stride - a.pad;
              th = (int)(t % a.OH) * a.stride - a.pad;
              tn = (int)(t / a.OH);
            }
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) {
              int n = n_blk * BN + j * 64;
              if (TN_CONV) {
                if (n >= a.N) n = 0;                // columns past N are never stored; keep the coordinates legal
                const int tap = n / a.Cin, c = n - tap * a.Cin;
                const int fr = tap / a.kw, fs = tap - fr * a.kw;
                ptx::tma_load_im2col_4d(Bs + j * 8192, &tmB, &full[s], c, tw, th, tn, (uint16_t)(fs * a.dil),
                                        (uint16_t)(fr * a.dil));
              } else {
                ptx::tma_load_2d(Bs + j * 8192, &tmB, &full[s], n, kb * BK);
              }
            }
          } else if (IM2COL) {
            // k-block kb = (filter tap, 64-channel block); the tile's first output pixel fixes the base input pixel
            const int tap = kb / a.cblocks, cb = kb - tap * a.cblocks;
            const int fr = tap / a.kw, fs = tap - fr * a.kw;
            ptx::tma_load_im2col_4d(As, &tmA, &full[s], cb * BK, iw0, ih0, img0, (uint16_t)(fs * a.dil),
                                    (uint16_t)(fr * a.dil));
          } else if (!GATHER) {
            ptx::tma_load_2d(As, &tmA, &full[s], kb * BK, m_blk * BM);
          }
          if (!TN) ptx::tma_load_2d(Bs, &tmB, &full[s], kb * BK, n_blk * BN);
          if (++s == (uint32_t)STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc = TN ? ptx::idesc_bf16_f32_mn(BM, BN) : ptx::idesc_bf16_f32(BM, BN);
      uint32_t s = 0, ph = 0, tcount = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
        const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
        ptx::mbar_wait(&tempty[as], aph ^ 1);
        ptx::tc_fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        const TileCoord tc = tile_coord(a, tile);
        for (int kb = tc.kb0; kb < tc.kb1; ++kb) {
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after_sync();
          const uint32_t a_addr = ptx::smem_u32(smem + s * C::STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_TILE_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if (TN)       // 16 rows of K = two 8-row groups = 2048 B inside every 64-element box
              ptx::mma_bf16_ss(d_tmem, ptx::smem_desc_mn_sw128(a_addr + k * 2048, 8192),
                               ptx::smem_desc_mn_sw128(b_addr + k * 2048, 8192), idesc, ((kb - tc.kb0) | k) != 0);
            else
              ptx::mma_bf16_ss(d_tmem, ptx::smem_desc_sw128(a_addr + k * 32),
                               ptx::smem_desc_sw128(b_addr + k * 32), idesc, ((kb - tc.kb0) | k) != 0);
          }
          ptx::mma_commit(&empty[s]);      // frees the smem slot when these MMAs retire
          if (++s == (uint32_t)STAGES) { s = 0; ph ^= 1; }
        }
        ptx::mma_commit(&tfull[as]);       // accumulator complete -> epilogue
      }
    }
  } else if (warp == 3) {
    if (lane == 0 && a.res_tma) {
      // ===== residual producer: sub-tiles in the order the epilogue consumes them =====
      uint32_t rb = 0, rph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(a, tile);
        const int m_blk = tc.m_blk, n_blk = tc.n_blk;
        for (int n0 = n_blk * BN; n0 < min(a.N, (n_blk + 1) * BN); n0 += EC) {
          ptx::mbar_wait(&rempty[rb], rph ^ 1);
          ptx::mbar_arrive_expect_tx(&rfull[rb], EPI_TILE_BYTES);
          ptx::tma_load_2d(rst + rb * EPI_TILE_BYTES, &tmR, &rfull[rb], n0, m_blk * BM);
          if (++rb == (uint32_t)RB) { rb = 0; rph ^= 1; }
        }
      }
    }
  } else if (warp >= 4 && warp < 4 + 4 * EG) {
    // ===== epilogue =====
    const int q = warp & 3;
    const int grp = (warp - 4) >> 2;                 // epilogue warp group
    const int row = q * 32 + lane;                   // row inside the tile == TMEM lane
    const int tid_e = (threadIdx.x - 128) & 127;     // thread inside its group
    const int tid_all = threadIdx.x - 128;
    uint32_t tcount = 0, cit = 0, subcnt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++tcount) {
      const TileCoord tc = tile_coord(a, tile);
      const int m_blk = tc.m_blk, n_blk = tc.n_blk;
      const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
      const int64_t m = (int64_t)m_blk * BM + row;
      const int64_t rb_row = a.rowbias_dyn   ? (int64_t)__ldg(a.rowbias_dyn)
                             : a.rowbias_idx ? (m < a.M ? (int64_t)__ldg(a.rowbias_idx + m) : 0)
                                             : (m % a.rowbias_period);
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
      if (a.tma_epi && a.warp_store) {
        // Warp-autonomous staged epilogue.  The accumulator leaves TMEM at 64 B/cycle (a 128 x 256 fp32 tile: 2048
        // cycles, as long as its MMAs with K = 256), so the epilogue is the pacing side of every GEMM with K <= 512 and
        // anything that idles the TMEM port costs directly: here a warp owns rows 32q .. 32q+31 of its group's
        // sub-tiles end to end — tcgen05.ld, scale/bias straight from global (L1 broadcast), swizzled staging slice,
        // its own 32-row TMA store — with no barrier wider than the warp.
        ptx::mbar_wait(&tfull[as], aph);
        ptx::tc_fence_after_sync();
        const int n_end = min(a.N, (n_blk + 1) * BN);
        // staging: one buffer per group, or two used alternately (c_dbl) so that a warp does not wait for the TMA
        // engine to have read its previous sub-tile before it converts the next one
        uint8_t* const cgrp = cst + (EG == 2 ? grp : 0) * (1 + a.c_dbl) * EPI_TILE_BYTES;
        uint8_t* cbuf = cgrp;
        // sub-tile prologue / epilogue shared by both output types
        auto sub_begin = [&](uint32_t rb, uint32_t rph) -> const uint8_t* {
          const uint8_t* rbuf = nullptr;
          if (a.res_tma) {
            ptx::mbar_wait(&rfull[rb], rph);
            rbuf = rst + rb * EPI_TILE_BYTES;
          }
          if (a.c_dbl) {
            cbuf = cgrp + (subcnt & 1u) * EPI_TILE_BYTES;
            ++subcnt;
            if (lane == 0) ptx::bulk_wait_read<1>();            // the store before the previous one has read this buffer
          } else {
            if (lane == 0) ptx::bulk_wait_read<0>();            // this warp's previous store has read its slice
          }
          __syncwarp();
          return rbuf;
        };
        auto sub_end = [&](uint32_t rb, int n0) {
          if (a.res_tma) ptx::mbar_arrive(&rempty[rb]);
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tmC, cbuf + q * 32 * 128, n0, m_blk * BM + q * 32);
            ptx::bulk_commit();
          }
        };
        if (a.out_fp32) {
          for (int n0 = n_blk * BN; n0 < n_end; n0 += EC, ++cit) {
            if (EG == 2 && (int)(cit & 1) != grp) continue;      // the other group's sub-tile
            const uint32_t rb = a.res_tma ? cit % RB : 0, rph = a.res_tma ? (cit / RB) & 1 : 0;
            const uint8_t* rbuf = sub_begin(rb, rph);
            uint32_t r[32];
            ptx::tmem_ld_32x32(t_row + (n0 - n_blk * BN), r);
            ptx::tmem_ld_wait();
            epilogue_chunk_staged<float>(a, r, a.scale ? a.scale + n0 : nullptr, a.bias ? a.bias + n0 : nullptr, rb_row, n0,
                                         row, 0, rbuf, cbuf);
            sub_end(rb, n0);
          }
        } else {
          // bf16: a sub-tile is two 32-column halves.  The TMEM loads run one half AHEAD of the arithmetic (ping-pong
          // register sets, also across the sub-tiles of the tile), so the port has a request from this warp queued
          // while the warp converts, stages and stores — without it all eight warps load, then all eight compute.
          const int first = n_blk * BN + (EG == 2 ? grp : 0) * EC, step = EG * EC;
          const float lo = a.relu == 1 ? 0.f : -INFINITY;
          auto chunk_bf16 = [&](uint32_t (&rr)[32], int nc, int chunk16, const uint8_t* rbuf) {
            switch (a.epi_mode) {                                 // warp-uniform
              case 1: epilogue_chunk_lean<1>(rr, nullptr, a.bias + nc, lo, row, chunk16, rbuf, cbuf); break;
              case 2: epilogue_chunk_lean<2>(rr, a.scale + nc, a.bias + nc, lo, row, chunk16, rbuf, cbuf); break;
              case 3: epilogue_chunk_lean<3>(rr, a.scale + nc, a.bias + nc, lo, row, chunk16, rbuf, cbuf); break;
              case 5: epilogue_chunk_lean<5>(rr, a.scale + nc, nullptr, lo, row, chunk16, rbuf, cbuf); break;
              case 4:
                epilogue_chunk_lean<4>(rr, nullptr, a.bias + nc, lo, row, chunk16, rbuf, cbuf,
                                       nc < a.rowbias_cols ? a.rowbias + rb_row * a.N + nc : nullptr);
                break;
              default:
                epilogue_chunk_staged<bf16>(a, rr, a.scale ? a.scale + nc : nullptr, a.bias ? a.bias + nc : nullptr, rb_row,
                                            nc, row, chunk16, rbuf, cbuf);
            }
          };
          uint32_t ra[32], rc[32];
          if (first < n_end) ptx::tmem_ld_32x32(t_row + (first - n_blk * BN), ra);
          for (int n0 = first; n0 < n_end; n0 += step) {
            const uint32_t sub = cit + (uint32_t)((n0 - n_blk * BN) / EC);
            const uint32_t rb = a.res_tma ? sub % RB : 0, rph = a.res_tma ? (sub / RB) & 1 : 0;
            const int col = n0 - n_blk * BN;
            ptx::tmem_ld_wait();
            ptx::tmem_ld_32x32(t_row + col + 32, rc);
            const uint8_t* rbuf = sub_begin(rb, rph);
            chunk_bf16(ra, n0, 0, rbuf);
            ptx::tmem_ld_wait();
            if (n0 + step < n_end) ptx::tmem_ld_32x32(t_row + col + step, ra);
            chunk_bf16(rc, n0 + 32, 4, rbuf);
            sub_end(rb, n0);
          }
          cit += (uint32_t)((n_end - n_blk * BN + EC - 1) / EC);
          ptx::tmem_ld_wait();
        }
        ptx::tc_fence_before_sync();
        ptx::mbar_arrive(&tempty[as]);
      } else if (a.tma_epi) {
        // per-tile column parameters -> smem (all epilogue threads)
        ptx::named_bar_sync(2, 128 * EG);
        for (int i = tid_all; i < BN; i += 128 * EG) {
          const int n = n_blk * BN + i;
          sb[i] = (a.scale && n < a.N) ? __ldg(a.scale + n) : 1.f;
          sb[BN + i] = (a.bias && n < a.N) ? __ldg(a.bias + n) : 0.f;
        }
        ptx::named_bar_sync(2, 128 * EG);
        ptx::mbar_wait(&tfull[as], aph);
        ptx::tc_fence_after_sync();
        for (int n0 = n_blk * BN; n0 < min(a.N, (n_blk + 1) * BN); n0 += EC, ++cit) {
          if (EG == 2 && (int)(cit & 1) != grp) continue;        // the other group's sub-tile
          const int col = n0 - n_blk * BN;
          uint8_t* cbuf = cst + (EG == 2 ? grp : 0) * EPI_TILE_BYTES;       // one staging buffer per group
          const uint8_t* rbuf = nullptr;
          const uint32_t rb = a.res_tma ? cit % RB : 0, rph = a.res_tma ? (cit / RB) & 1 : 0;
          if (a.res_tma) {
            ptx::mbar_wait(&rfull[rb], rph);
            rbuf = rst + rb * EPI_TILE_BYTES;
          }
          if (tid_e == 0) {                                   // the store that last used cbuf has read it
            ptx::bulk_wait_read<0>();
          }
          ptx::named_bar_sync(1 + 2 * grp, 128);
          for (int c = 0; c < EC / 32; ++c) {
            uint32_t r[32];
            ptx::tmem_ld_32x32(t_row + col + c * 32, r);
            ptx::tmem_ld_wait();
            if (a.out_fp32)
              epilogue_chunk_staged<float>(a, r, sb + col + c * 32, sb + BN + col + c * 32, rb_row, n0 + c * 32, row,
                                           0, rbuf, cbuf);
            else
              epilogue_chunk_staged<bf16>(a, r, sb + col + c * 32, sb + BN + col + c * 32, rb_row, n0 + c * 32, row,
                                          c * 4, rbuf, cbuf);
          }
          if (a.res_tma) ptx::mbar_arrive(&rempty[rb]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(1 + 2 * grp, 128);
          if (tid_e == 0) {
            ptx::tma_store_2d(&tmC, cbuf, n0, m_blk * BM);
            ptx::bulk_commit();
          }
        }
        ptx::tc_fence_before_sync();
        ptx::mbar_arrive(&tempty[as]);
      } else {
        ptx::mbar_wait(&tfull[as], aph);
        ptx::tc_fence_after_sync();
        float best_v = -INFINITY;
        uint32_t best_i = 0;
        const int64_t mc = a.c_group ? (m / a.c_group) * a.c_group_stride + (m % a.c_group) : m + (int64_t)tc.sp * a.M;
#pragma unroll 1
        for (int c = grp; c < BN / 32; c += EG) {     // 32-column chunks alternate between the groups
          const int n0 = n_blk * BN + c * 32;
          if (n0 >= a.N) break;                       // warp-uniform
          uint32_t r[32];
          ptx::tmem_ld_32x32(t_row + c * 32, r);
          ptx::tmem_ld_wait();
          if (m < a.M) epilogue_chunk(a, r, m, mc, rb_row, n0, best_v, best_i);
        }
        ptx::tc_fence_before_sync();
        ptx::mbar_arrive(&tempty[as]);
        if (a.argmax_out && m < a.M && best_v > -INFINITY)
          atomicMax(a.argmax_out + m, pack_argmax(best_v, best_i));
      }
    }
    if (a.tma_epi && (a.warp_store ? lane == 0 : tid_e == 0)) ptx::bulk_wait_read<0>();   // smem must outlive the last store's read
  } else if (GATHER && warp >= 8) {
    // ===== im2col gather producer: one thread per tile row =====
    const int row = threadIdx.x - 256;
    const uint32_t swz = (uint32_t)(row & 7);
    const uint32_t lag = STAGES >= 6 ? 2u : 1u;
    uint32_t it = 0, s = 0, ph = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const TileCoord tc = tile_coord(a, tile);
      const int m_blk = tc.m_blk;
      const int64_t m = (int64_t)m_blk * BM + row;
      const bool valid_m = m < a.M;
      const int ow = (int)(m % a.OW);
      const int64_t t = m / a.OW;
      const int oh = (int)(t % a.OH);
      const int64_t img = t / a.OH;
      const int ih0 = oh * a.stride - a.pad, iw0 = ow * a.stride - a.pad;
      int tap = tc.kb0 / a.cblocks, cb = tc.kb0 - tap * a.cblocks;
      for (int kb = tc.kb0; kb < tc.kb1; ++kb, ++it) {
        ptx::mbar_wait(&empty[s], ph ^ 1);
        const int fr = tap / a.kw, fs = tap - fr * a.kw;
        const int ih = ih0 + fr * a.dil, iw = iw0 + fs * a.dil;
        const bool ok = valid_m && ih >= 0 && ih < a.H && iw >= 0 && iw < a.W;
        const bf16* src = ok ? a.A + ((img * a.H + ih) * a.W + iw) * a.Cin + cb * BK : a.A;
        const uint32_t dst = ptx::smem_u32(smem + s * C::STAGE_BYTES) + row * 128;
        const uint32_t nbytes = ok ? 16u : 0u;
#pragma unroll
        for (uint32_t j = 0; j < 8; ++j) ptx::cp_async_16(dst + ((j ^ swz) << 4), src + j * 8, nbytes);
        ptx::cp_async_commit();
        if (it >= lag) {
          if (lag == 2) ptx::cp_async_wait<2>();
          else ptx::cp_async_wait<1>();
          ptx::fence_proxy_async_smem();
          ptx::mbar_arrive(&full[(it - lag) % STAGES]);
        }
        if (++cb == a.cblocks) { cb = 0; ++tap; }
        if (++s == (uint32_t)STAGES) { s = 0; ph ^= 1; }
      }
    }
    ptx::cp_async_wait<0>();
    ptx::fence_proxy_async_smem();
    for (uint32_t j = (it > lag ? it - lag : 0); j < it; ++j) ptx::mbar_arrive(&full[j % STAGES]);
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ---- host side ------------------------------------------------------------------------------
PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
    RETR_CHECK(p != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled not available");
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

PFN_cuTensorMapEncodeIm2col_v12000 get_encode_im2col_fn() {
  static PFN_cuTensorMapEncodeIm2col_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres));
    RETR_CHECK(p != nullptr && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeIm2col not available");
    fn = reinterpret_cast<PFN_cuTensorMapEncodeIm2col_v12000>(p);
  }
  return fn;
}

// 2-D tensor [rows, cols] with row pitch ld (elements); box {box_cols, box_rows} with 128-byte inner
// extent, 128B swizzle.
CUtensorMap make_tmap_2d(const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows, bool fp32 = false) {
  CUtensorMap m;
  const int esize = fp32 ? 4 : 2;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)(ld * esize)};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esize), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  RETR_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * esize) % 16 == 0,
             "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch");
  CUresult r = get_encode_fn()(&m, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                               const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RETR_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
  return m;
}

// NHWC activation [n, H, W, C] as an im2col tensor map (the A operand of a convolution): the bounding box of base
// pixels starts at -pad and ends (k-1)*dil - pad short of the far edge, traversed with the convolution stride, so
// "128 pixels from base pixel (iw0, ih0, img)" are 128 consecutive output pixels in (img, oh, ow) order; the filter tap
// is the per-instruction offset.  Same corner rule as CUTLASS's fprop (conv/collective/detail.hpp).
CUtensorMap make_tmap_im2col(const void* base, int64_t n, const ConvView& c, int pixels = BM) {
  CUtensorMap m;
  cuuint64_t dims[4] = {(cuuint64_t)c.C, (cuuint64_t)c.W, (cuuint64_t)c.H, (cuuint64_t)n};
  cuuint64_t strides[3] = {(cuuint64_t)c.C * 2, (cuuint64_t)c.W * c.C * 2, (cuuint64_t)c.H * c.W * c.C * 2};
  const int lo = -c.pad, hi_h = c.pad - (c.kh - 1) * c.dil, hi_w = c.pad - (c.kw - 1) * c.dil;
  int lower[2] = {lo, lo}, upper[2] = {hi_w, hi_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)c.stride, (cuuint32_t)c.stride, 1};
  RETR_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0, "im2col operand must be 16-byte aligned");
  CUresult r = get_encode_im2col_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides,
                                      lower, upper, (cuuint32_t)BK, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  RETR_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed (" + std::to_string((int)r) + ")");
  return m;
}

template <int BN, int AMODE>
void launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const CUtensorMap& tmR,
            KArgs& ka, int grid, cudaStream_t stream, bool pdl) {
  constexpr bool GATHER = AMODE == 1;
  static const int want_rb = getenv("RETR_GEMM_RBUFS") ? std::min(R_BUFS, std::max(2, atoi(getenv("RETR_GEMM_RBUFS")) & ~1)) : 4;
  ka.r_bufs = ka.res_tma ? want_rb : 0;
  // short main loops (K <= 256) are paced by the epilogue: a second staging buffer per group is worth a pipeline stage
  static const int cdbl_kb = getenv("RETR_GEMM_CDBL_KB") ? atoi(getenv("RETR_GEMM_CDBL_KB")) : 4;
  ka.c_dbl = ka.warp_store && !GATHER && ka.num_kb <= cdbl_kb && AMODE < 3 &&
             Cfg<BN>::stages_for(ka.tma_epi, ka.r_bufs, GATHER, 1) >= 2;
  ka.stages = Cfg<BN>::stages_for(ka.tma_epi, ka.r_bufs, GATHER, ka.c_dbl);
  RETR_CHECK(ka.stages >= (GATHER ? 3 : 2), "not enough shared memory for the pipeline");
  const int smem = Cfg<BN>::fixed_bytes(ka.tma_epi, ka.r_bufs, GATHER, ka.c_dbl) + ka.stages * Cfg<BN>::STAGE_BYTES;
  if (pdl) {
    launch_pdl(gemm_tcgen05_kernel<BN, AMODE>, dim3((unsigned)grid), dim3(384), (size_t)smem, stream, tmA, tmB, tmC, tmR, ka);
  } else {
    gemm_tcgen05_kernel<BN, AMODE><<<grid, 384, smem, stream>>>(tmA, tmB, tmC, tmR, ka);
    LAUNCH_OK();
  }
}


// ---- vocabulary head of a greedy step: argmax_n( A[M,K] . W[N,K]^T + bias ) with the token bookkeeping fused ----
// The general kernel above runs this shape (M = 256, N = 30522, K = 512) as 240 tiles of 128 x 256 over 148 CTAs: two
// rounds, and every tile re-reads its 128 KB of A next to 256 KB of W, 92 MB through L2 per step (19 us inside the
// decode graph).  Here ONE CTA owns a 208-column slice of the vocabulary (147 CTAs for 30522 words: one round) and
// multiplies it with BOTH 128-row halves of A from the same shared-memory stage (two accumulators = all 512 TMEM
// columns), so W crosses L2 -> SM exactly once per step.  The last CTA to finish (ticket counter) does the work of
// greedy_update_kernel, which removes that launch from the step as well.
constexpr int VA_BN = 208;
constexpr int VA_B_BYTES = VA_BN * BK * 2;                       // 26 KB, a multiple of 1024
constexpr int VA_STAGE_BYTES = 2 * A_TILE_BYTES + VA_B_BYTES;
constexpr int VA_STAGES = 3;
constexpr int VA_THREADS = 320;                                  // producer, MMA, 8 epilogue warps
constexpr int VA_SB = (VA_BN + 31) / 32 * 32;                     // bias slice, padded to whole 32-column chunks
constexpr int VA_SMEM = VA_STAGES * VA_STAGE_BYTES + 128 + VA_SB * 4;

struct VocabArgs {
  int M, N, num_kb;
  const float* bias;
  unsigned long long* keys;
  int64_t* caption;
  int max_len, eos;
  int* step;
  int* done;
  uint8_t* finished;
  unsigned int* ticket;
};

__global__ void __launch_bounds__(VA_THREADS, 1)
vocab_argmax_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const VocabArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + VA_STAGES * VA_STAGE_BYTES);
  uint64_t* empty = full + VA_STAGES;
  uint64_t* tfull = empty + VA_STAGES;
  uint64_t* tempty = tfull + 1;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty + 1);
  int* flags = reinterpret_cast<int*>(tmem_holder + 1);          // [0] this is the last CTA, [1] some row unfinished
  float* sbias = reinterpret_cast<float*>(smem + VA_STAGES * VA_STAGE_BYTES + 128);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * VA_BN;
  const int m_pairs = (a.M + 2 * BM - 1) / (2 * BM);

  if (threadIdx.x == 0) {
    if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < VA_STAGES; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(tfull, 1);
    ptx::mbar_init(tempty, VA_THREADS - 64);
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_holder, 512);
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_holder;
  pdl_launch_dependents();
  pdl_wait();                                   // A is the previous kernel's output; done / step / keys likewise
  const bool skip = a.done != nullptr && *a.done != 0;          // the same for every thread of the grid

  if (!skip) {
    if (warp == 0) {
      if (lane == 0) {
        uint32_t s = 0, ph = 0;
        for (int mp = 0; mp < m_pairs; ++mp) {
          const bool two = mp * 2 * BM + BM < a.M;
          for (int kb = 0; kb < a.num_kb; ++kb) {
            ptx::mbar_wait(&empty[s], ph ^ 1);
            uint8_t* As = smem + s * VA_STAGE_BYTES;
            ptx::mbar_arrive_expect_tx(&full[s], VA_B_BYTES + (two ? 2 : 1) * A_TILE_BYTES);
            ptx::tma_load_2d(As, &tmA, &full[s], kb * BK, mp * 2 * BM);
            if (two) ptx::tma_load_2d(As + A_TILE_BYTES, &tmA, &full[s], kb * BK, mp * 2 * BM + BM);
            ptx::tma_load_2d(As + 2 * A_TILE_BYTES, &tmB, &full[s], kb * BK, n0);
            if (++s == VA_STAGES) { s = 0; ph ^= 1; }
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = ptx::idesc_bf16_f32(BM, VA_BN);
        uint32_t s = 0, ph = 0;
        for (int mp = 0; mp < m_pairs; ++mp) {
          const bool two = mp * 2 * BM + BM < a.M;
          ptx::mbar_wait(tempty, (mp & 1) ^ 1);
          ptx::tc_fence_after_sync();
          for (int kb = 0; kb < a.num_kb; ++kb) {
            ptx::mbar_wait(&full[s], ph);
            ptx::tc_fence_after_sync();
            const uint32_t a_addr = ptx::smem_u32(smem + s * VA_STAGE_BYTES);
            const uint32_t b_addr = a_addr + 2 * A_TILE_BYTES;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t bd = ptx::smem_desc_sw128(b_addr + k * 32);
              ptx::mma_bf16_ss(tmem_base, ptx::smem_desc_sw128(a_addr + k * 32), bd, idesc, (kb | k) != 0);
              if (two)
                ptx::mma_bf16_ss(tmem_base + 256, ptx::smem_desc_sw128(a_addr + A_TILE_BYTES + k * 32), bd, idesc,
                                 (kb | k) != 0);
            }
            ptx::mma_commit(&empty[s]);
            if (++s == VA_STAGES) { s = 0; ph ^= 1; }
          }
          ptx::mma_commit(tfull);
        }
      }
    } else {
      const int e = threadIdx.x - 64;                 // 0..255
      const int h = e >> 7;                           // accumulator = 128-row half of the pair
      const int q = warp & 3;                         // TMEM lane quarter this warp may read
      const int row = q * 32 + lane;
      for (int i = e; i < VA_SB; i += VA_THREADS - 64)
        sbias[i] = (a.bias && i < VA_BN && n0 + i < a.N) ? __ldg(a.bias + n0 + i) : 0.f;
      ptx::named_bar_sync(1, VA_THREADS - 64);
      for (int mp = 0; mp < m_pairs; ++mp) {
        ptx::mbar_wait(tfull, mp & 1);
        ptx::tc_fence_after_sync();
        const int m = mp * 2 * BM + h * BM + row;
        if (mp * 2 * BM + h * BM < a.M) {             // warp-uniform: this accumulator was written
          float best_v = -INFINITY;
          uint32_t best_i = 0;
#pragma unroll 1
          for (int c = 0; c < (VA_BN + 31) / 32; ++c) {
            uint32_t r[32];
            ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + h * 256 + c * 32, r);
            ptx::tmem_ld_wait();
            const int nv = min(32, min(VA_BN - c * 32, a.N - n0 - c * 32));
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float v = __uint_as_float(r[j]) + sbias[c * 32 + j];
              if (j < nv && v > best_v) { best_v = v; best_i = (uint32_t)(n0 + c * 32 + j); }
            }
          }
          if (m < a.M && best_v > -INFINITY) atomicMax(a.keys + m, pack_argmax(best_v, best_i));
        }
        ptx::tc_fence_before_sync();
        ptx::mbar_arrive(tempty);
      }
    }
  }
  __threadfence();
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
  if (skip || a.ticket == nullptr) return;

  // ---- token bookkeeping by the last CTA (eval_utils/decode.py:70-79, as greedy_update_kernel) ----
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(a.ticket, 1u);
    flags[0] = t == gridDim.x - 1;
    flags[1] = 0;
  }
  __syncthreads();
  if (!flags[0]) return;
  __threadfence();
  const int i = *a.step;
  for (int b = threadIdx.x; b < a.M; b += VA_THREADS) {
    const int tok = (int)unpack_argmax_index(__ldcg(a.keys + b));
    const uint8_t fin = a.finished[b] | (uint8_t)(tok == a.eos);
    a.finished[b] = fin;
    if (!fin) flags[1] = 1;
  }
  __syncthreads();
  const bool all_done = flags[1] == 0;
  for (int b = threadIdx.x; b < a.M; b += VA_THREADS) {
    if (!all_done) a.caption[(int64_t)b * a.max_len + i + 1] = (int64_t)unpack_argmax_index(__ldcg(a.keys + b));
    a.keys[b] = 0ull;
  }
  if (threadIdx.x == 0) {
    *a.step = i + 1;
    if (all_done || i + 1 >= a.max_len - 1) *a.done = 1;
    *a.ticket = 0u;
  }
}

}  // namespace

template <int BN> static void prepare_bn() {
  CUDA_OK(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  CUDA_OK(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  CUDA_OK(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  if (BN >= 64) {
    CUDA_OK(cudaFuncSetAttribute(gemm_tcgen05_kernel<(BN < 64 ? 64 : BN), 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    CUDA_OK(cudaFuncSetAttribute(gemm_tcgen05_kernel<(BN < 64 ? 64 : BN), 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  }
}
void gemm_bf16_prepare() {
  prepare_bn<32>(); prepare_bn<64>(); prepare_bn<128>(); prepare_bn<256>();
  get_encode_fn();
  get_encode_im2col_fn();
}

bool vocab_argmax_supported(int64_t M, int N, int K) { return M > 0 && N > 0 && K > 0 && K % BK == 0; }

void vocab_argmax(const VocabArgmaxParams& p, cudaStream_t stream) {
  static bool prepared = false;
  if (!prepared) {
    get_encode_fn();
    CUDA_OK(cudaFuncSetAttribute(vocab_argmax_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, VA_SMEM));
    prepared = true;
  }
  RETR_CHECK(vocab_argmax_supported(p.M, p.N, p.K), "vocab_argmax: K must be a positive multiple of 64");
  RETR_CHECK(!p.ticket || (p.caption && p.step && p.done && p.finished), "vocab_argmax: fused bookkeeping operands");
  VocabArgs a{};
  a.M = (int)p.M; a.N = p.N; a.num_kb = p.K / BK;
  a.bias = p.bias; a.keys = p.keys;
  a.caption = p.caption; a.max_len = p.max_len; a.eos = p.eos;
  a.step = p.step; a.done = p.done; a.finished = p.finished; a.ticket = p.ticket;
  const CUtensorMap tmA = make_tmap_2d(p.A, p.M, p.K, p.lda, BM);
  const CUtensorMap tmB = make_tmap_2d(p.W, p.N, p.K, p.K, VA_BN);
  launch_pdl(vocab_argmax_kernel, dim3((unsigned)cdiv(p.N, VA_BN)), dim3(VA_THREADS), (size_t)VA_SMEM, stream, tmA, tmB, a);
}

int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    CUDA_OK(cudaGetDevice(&dev));
    CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  }
  return n;
}

static thread_local int g_cta_cap = 0;
void gemm_set_cta_cap(int ctas) { g_cta_cap = ctas > 0 ? ctas : 0; }

void gemm_bf16(const GemmParams& p, cudaStream_t stream) {
  static bool prepared = false;
  if (!prepared) { gemm_bf16_prepare(); prepared = true; }
  RETR_CHECK(p.M > 0 && p.N > 0 && p.K > 0, "empty GEMM");
  const bool tn = p.mn_major;
  RETR_CHECK(tn || p.K % 8 == 0, "K must be a multiple of 8 (16-byte rows)");
  const bool gather = p.conv.kh > 0;
  if (tn) {
    RETR_CHECK(p.lda % 8 == 0 && (gather || p.ldb % 8 == 0), "MN-major operands need 16-byte row pitches");
    RETR_CHECK(!gather || p.N == p.conv.kh * p.conv.kw * p.conv.C, "MN-major conv: N must be kh*kw*C_in");
  }
  KArgs ka{};
  ka.M = p.M; ka.N = p.N; ka.K = p.K;
  ka.num_kb = (int)cdiv(p.K, BK);
  ka.C = p.C; ka.ldc = p.ldc; ka.out_fp32 = p.out_fp32;
  ka.scale = p.scale; ka.bias = p.bias; ka.rowbias = p.rowbias;
  ka.rowbias_period = p.rowbias_period > 0 ? p.rowbias_period : 1;
  ka.residual = p.residual; ka.ldr = p.ldr; ka.res_fp32 = p.res_fp32; ka.relu = p.relu;
  ka.argmax_out = p.argmax_out;
  ka.c_group = p.c_group; ka.c_group_stride = p.c_group_stride; ka.rowbias_dyn = p.rowbias_dyn;
  ka.rowbias_idx = p.rowbias_idx; ka.m_dyn = p.m_dyn;
  RETR_CHECK(p.rowbias_cols % 32 == 0 && p.rowbias_cols <= p.N, "rowbias_cols: a multiple of 32, at most N");
  ka.rowbias_cols = p.rowbias_cols > 0 ? p.rowbias_cols : p.N;
  RETR_CHECK(!p.m_dyn || p.split_k == 1, "m_dyn: no split-K");
  const int64_t cbytes = p.out_fp32 ? 4 : 2, rbytes = p.res_fp32 ? 4 : 2;
  ka.vecC = p.C && (reinterpret_cast<uintptr_t>(p.C) % 16 == 0) && (p.ldc * cbytes) % 16 == 0;
  ka.vecR = p.residual && (reinterpret_cast<uintptr_t>(p.residual) % 16 == 0) && (p.ldr * rbytes) % 16 == 0;
  ka.vecRB = p.rowbias && (reinterpret_cast<uintptr_t>(p.rowbias) % 16 == 0) && p.N % 4 == 0;
  ka.A = static_cast<const bf16*>(p.A);
  if (gather) {
    const ConvView& c = p.conv;
    RETR_CHECK(c.C % BK == 0, "implicit-GEMM conv needs C_in % 64 == 0");
    RETR_CHECK(tn || p.K == c.kh * c.kw * c.C, "conv K mismatch");
    RETR_CHECK((tn ? (int64_t)p.K : p.M) % ((int64_t)c.OH * c.OW) == 0, "conv pixel count must be images*OH*OW");
    ka.kh = c.kh; ka.kw = c.kw; ka.stride = c.stride; ka.pad = c.pad; ka.dil = c.dil;
    ka.H = c.H; ka.W = c.W; ka.Cin = c.C; ka.OH = c.OH; ka.OW = c.OW; ka.cblocks = c.C / BK;
  }
  // tile width: widest that still leaves every SM a tile
  const int sms = sm_count();
  int bn = 256;
  int splits = 1;
  const int bn_min = (tn || p.relu == 3) ? 64 : 32;      // MN-major boxes / the staged bf16 epilogue need 64 columns
  if (p.split_k != 1) {
    RETR_CHECK(p.out_fp32 && p.C && p.ldc == p.N && !p.scale && !p.bias && !p.rowbias && !p.residual && !p.relu &&
                   !p.argmax_out && !p.c_group,
               "split-K GEMM takes a plain fp32 C and no epilogue operands");
    while (bn > bn_min && bn / 2 >= p.N) bn /= 2;
    const int64_t tiles = cdiv(p.M, BM) * cdiv(p.N, bn);
    splits = p.split_k > 1 ? p.split_k : (int)std::max<int64_t>(1, (2 * sms) / tiles);
    splits = std::min(splits, std::max(1, ka.num_kb / 4));
    if (p.split_ws) splits = (int)std::min<int64_t>(splits, p.split_ws_floats / std::max<int64_t>(1, p.M * p.N));
    else splits = 1;
    splits = std::max(splits, 1);
  } else {
    while (bn > bn_min && (bn / 2 >= p.N || cdiv(p.M, BM) * cdiv(p.N, bn) < sms)) bn /= 2;
    // a residual GEMM whose whole width is one 256-column tile (the encoder's out-projection / FF2) runs faster as two
    // 128-column tiles: 4 pipeline stages instead of 2 next to the residual ring (encoder 5.90 -> 5.72 ms); the wide
    // conv3 GEMMs of the backbone do not gain (14.84 vs 14.97 ms)
    static const int bn_res = getenv("RETR_GEMM_BN_RES") ? atoi(getenv("RETR_GEMM_BN_RES")) : 0;
    if (p.residual && !gather && !tn && (bn_res > 0 || p.N <= 256)) bn = std::max(bn_min, std::min(bn, bn_res > 0 ? bn_res : 128));
  }
  ka.kb_per_split = (int)cdiv(ka.num_kb, splits);
  splits = (int)cdiv(ka.num_kb, ka.kb_per_split);          // no empty split
  ka.splits = splits;
  const bool acc_in_place = p.split_k != 1 && splits == 1 && p.accumulate;
  if (splits > 1) ka.C = p.split_ws;
  if (acc_in_place) { ka.residual = p.C; ka.ldr = p.ldc; ka.res_fp32 = 1; ka.vecR = ka.vecC; }
  // epilogue route: staged TMA store whenever C is a plain, 16-byte aligned matrix (a staged
  // sub-tile is 128 B wide, so a bf16 C needs BN >= 64)
  const bool force_direct = getenv("RETR_GEMM_DIRECT_EPILOGUE") != nullptr;
  ka.tma_epi = !force_direct && splits == 1 && ka.vecC && !p.argmax_out && !p.c_group && (p.out_fp32 || bn >= 64) &&
               (!ka.residual || (ka.vecR && ka.res_fp32 == p.out_fp32));
  ka.res_tma = ka.tma_epi && ka.residual;
  RETR_CHECK(p.relu != 3 || ka.res_tma, "relu == 3 (mask by the residual operand) needs the staged TMA epilogue");
  ka.m_tiles = (int)cdiv(p.M, BM);
  ka.n_tiles = (int)cdiv(p.N, bn);
  const int grid = (int)std::min<int64_t>((int64_t)ka.m_tiles * ka.n_tiles * splits, g_cta_cap > 0 ? std::min(g_cta_cap, sms) : sms);
  const int64_t n_img = gather ? (tn ? (int64_t)p.K : p.M) / ((int64_t)p.conv.OH * p.conv.OW) : 0;
  CUtensorMap tmB = !tn      ? make_tmap_2d(p.W, p.N, p.K, p.K, bn)
                    : gather ? make_tmap_im2col(p.W, n_img, p.conv, BK)
                             : make_tmap_2d(p.W, p.K, p.N, p.ldb, BK);
  // both spatial corners must fit the descriptor's signed bytes; anything else keeps the cp.async gather
  static const bool force_gather = getenv("RETR_CONV_GATHER") != nullptr && atoi(getenv("RETR_CONV_GATHER")) != 0;
  const bool im2col = !tn && gather && !force_gather && p.conv.pad <= 127 && (p.conv.kh - 1) * p.conv.dil - p.conv.pad <= 128 &&
                      p.conv.stride <= 8 && p.conv.kh == p.conv.kw;
  CUtensorMap tmA = tn       ? make_tmap_2d(p.A, p.K, p.M, p.lda, BK)
                    : im2col ? make_tmap_im2col(p.A, n_img, p.conv)
                    : gather ? tmB
                             : make_tmap_2d(p.A, p.M, p.K, p.lda, BM);
  static const bool warp_store_on = !(getenv("RETR_GEMM_WARP_STORE") && atoi(getenv("RETR_GEMM_WARP_STORE")) == 0);
  ka.warp_store = ka.tma_epi && warp_store_on && p.N % 64 == 0;
  ka.epi_mode = 0;
  static const bool lean_on = !(getenv("RETR_GEMM_LEAN_EPILOGUE") && atoi(getenv("RETR_GEMM_LEAN_EPILOGUE")) == 0);
  if (lean_on && ka.warp_store && !p.out_fp32 && p.bias && p.rowbias && ka.vecRB && !p.scale && !p.residual && p.relu <= 1 &&
      reinterpret_cast<uintptr_t>(p.bias) % 16 == 0)
    ka.epi_mode = 4;
  if (lean_on && ka.warp_store && !p.out_fp32 && !p.bias && !p.rowbias && p.scale && p.relu == 3 && ka.res_tma && !ka.res_fp32 &&
      reinterpret_cast<uintptr_t>(p.scale) % 16 == 0)
    ka.epi_mode = 5;
  if (lean_on && ka.warp_store && !p.out_fp32 && p.bias && !p.rowbias && p.relu <= 2 &&
      reinterpret_cast<uintptr_t>(p.bias) % 16 == 0 && (!p.scale || reinterpret_cast<uintptr_t>(p.scale) % 16 == 0)) {
    if (!p.scale && !p.residual && p.relu <= 1) ka.epi_mode = 1;
    else if (p.scale && !p.residual && p.relu <= 1) ka.epi_mode = 2;
    else if (p.scale && ka.res_tma && !ka.res_fp32 && p.relu == 2) ka.epi_mode = 3;
  }
  CUtensorMap tmC = ka.tma_epi ? make_tmap_2d(p.C, p.M, p.N, p.ldc, ka.warp_store ? 32 : BM, p.out_fp32) : tmB;
  CUtensorMap tmR = ka.res_tma ? make_tmap_2d(ka.residual, p.M, p.N, ka.ldr, BM, ka.res_fp32 != 0) : tmB;
  // a capped grid shares the GPU with another context: an early-launched dependent grid would park its CTAs on
  // exactly the SMs the cap leaves free, so programmatic dependent launch is off then
  const bool pdl = p.pdl && g_cta_cap == 0;
#define RETR_LAUNCH(BN_)                                                   \
  if (tn && gather) launch<(BN_ < 64 ? 64 : BN_), 4>(tmA, tmB, tmC, tmR, ka, grid, stream, pdl);   \
  else if (tn) launch<(BN_ < 64 ? 64 : BN_), 3>(tmA, tmB, tmC, tmR, ka, grid, stream, pdl);        \
  else if (im2col) launch<BN_, 2>(tmA, tmB, tmC, tmR, ka, grid, stream, pdl);          \
  else if (gather) launch<BN_, 1>(tmA, tmB, tmC, tmR, ka, grid, stream, pdl);     \
  else launch<BN_, 0>(tmA, tmB, tmC, tmR, ka, grid, stream, pdl)
  switch (bn) {
    case 256: RETR_LAUNCH(256); break;
    case 128: RETR_LAUNCH(128); break;
    case 64: RETR_LAUNCH(64); break;
    default: RETR_LAUNCH(32); break;
  }
#undef RETR_LAUNCH
  if (splits > 1) splitk_reduce(p.split_ws, static_cast<float*>(p.C), p.M * p.N, splits, p.accumulate ? 1 : 0, stream);
}

}  // namespace retr
