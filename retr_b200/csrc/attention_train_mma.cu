// bf16 flash attention for the TRAINING step on the warp-level tensor path (mma.sync m16n8k16, fp32
// accumulate): forward with dropout on the probabilities and a log-sum-exp per query row, and a
// recompute-style backward (no [Lq,Lk] probability matrix ever reaches HBM):
//   fwd   : O = dropout(softmax(QK^T/sqrt(dh) + masks)) V ;  lse2 = log2(sum exp) per row
//   bwd_dq: per 64-query tile, over all key tiles:  P = exp2(S - lse2), dP = dO V^T (dropout mask),
//           dS = P o (dP - delta), dQ = dS K / sqrt(dh);  delta = rowsum(dO o O) is written out
//   bwd_kv: per 64-key tile, over all query tiles (transposed problem): dV = (P o mask)^T dO, dK = dS^T Q / sqrt(dh)
// Replaces torch's multi_head_attention_forward core + its autograd (torch/nn/functional.py:6630-6664) with the
// reference's masks (key padding, causal) and dropout site (nn.MultiheadAttention(dropout=p),
// models/ConcatTransformer.py:160,204,210).  The fragment layouts are those of attention_mma.cu.
#include "dropout.cuh"
#include "train_ops.h"

namespace retr {
namespace {

constexpr int TQ = 64;     // rows of the tile a CTA owns (4 warps x 16)
constexpr int TK = 64;     // rows of the streamed tile

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp16(uint32_t dst, const void* src, bool ok) {
  const uint32_t n = ok ? 16u : 0u;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t sptr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// one MUFU.EX2 (exp2f adds a denormal-range rescale softmax does not need: tiny probabilities flush to zero)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// rows [r0, r0+64) of a [*, ld] bf16 matrix (head slice of DH columns) -> smem tile [64][DH+8]; rows >= r_end zero
template <int DH>
__device__ __forceinline__ void load_tile(bf16* dst, const bf16* src, int64_t ld, int r0, int r_end) {
  constexpr int LDS = DH + 8, CPR = DH / 8;
  for (int idx = threadIdx.x; idx < 64 * CPR; idx += 128) {
    const int r = idx / CPR, c = idx % CPR;
    const bool ok = r0 + r < r_end;
    cp16(sptr(dst + r * LDS + c * 8), src + (int64_t)(ok ? r0 + r : 0) * ld + c * 8, ok);
  }
}
// A fragments (16 rows of this warp x DH) of a smem tile
template <int DH>
__device__ __forceinline__ void load_a_frags(const bf16* tile, int warp, int lane, uint32_t (&a)[DH / 16][4]) {
  constexpr int LDS = DH + 8;
#pragma unroll
  for (int ks = 0; ks < DH / 16; ++ks)
    ldsm_x4(sptr(tile + (warp * 16 + (lane & 15)) * LDS + ks * 16 + (lane >> 4) * 8), a[ks][0], a[ks][1], a[ks][2], a[ks][3]);
}
// C[16 x 64] = A[16 x DH] . tile[64 x DH]^T      (tile rows are the columns of C)
template <int DH>
__device__ __forceinline__ void mma_a_tileT(float (&c)[8][4], const uint32_t (&a)[DH / 16][4], const bf16* tile, int lane) {
  constexpr int LDS = DH + 8;
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
#pragma unroll
  for (int np = 0; np < 4; ++np) {
#pragma unroll
    for (int ks = 0; ks < DH / 16; ++ks) {
      uint32_t b0, b1, b2, b3;
      const int row = np * 16 + (lane >> 4) * 8 + (lane & 7);
      const int kc = ks * 16 + ((lane >> 3) & 1) * 8;
      ldsm_x4(sptr(tile + row * LDS + kc), b0, b1, b2, b3);
      mma_16816(c[2 * np], a[ks], b0, b1);
      mma_16816(c[2 * np + 1], a[ks], b2, b3);
    }
  }
}
// acc[16 x DH] += P[16 x 64] . tile[64 x DH]     (P given as fp32 C fragments, rounded to bf16)
template <int DH>
__device__ __forceinline__ void mma_p_tile(float (&acc)[DH / 8][4], const float (&pf)[8][4], const bf16* tile, int lane) {
  constexpr int LDS = DH + 8;
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
    uint32_t pa[4];
    pa[0] = pack_bf16(pf[2 * kk][0], pf[2 * kk][1]);
    pa[1] = pack_bf16(pf[2 * kk][2], pf[2 * kk][3]);
    pa[2] = pack_bf16(pf[2 * kk + 1][0], pf[2 * kk + 1][1]);
    pa[3] = pack_bf16(pf[2 * kk + 1][2], pf[2 * kk + 1][3]);
#pragma unroll
    for (int np = 0; np < DH / 16; ++np) {
      uint32_t b0, b1, b2, b3;
      const int row = kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7);
      const int col = np * 16 + (lane >> 4) * 8;
      ldsm_x4_t(sptr(tile + row * LDS + col), b0, b1, b2, b3);
      mma_16816(acc[2 * np], pa, b0, b1);
      mma_16816(acc[2 * np + 1], pa, b2, b3);
    }
  }
}

// ---- forward ---------------------------------------------------------------------------------------------
template <int DH>
__global__ void __launch_bounds__(128) flash_train_fwd_kernel(const AttnTrainParams p) {
  constexpr int LDS = DH + 8, NT = DH / 8;
  __shared__ __align__(16) bf16 Qs[TQ * LDS];
  __shared__ __align__(16) bf16 Ks[2][TK * LDS];
  __shared__ __align__(16) bf16 Vs[2][TK * LDS];
  __shared__ uint8_t Ms[2][TK];
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bf16* qg = static_cast<const bf16*>(p.q) + (int64_t)b * p.Lq * p.ldq + h * DH;
  const bf16* kg = static_cast<const bf16*>(p.k) + (int64_t)b * p.Lk * p.ldk + h * DH;
  const bf16* vg = static_cast<const bf16*>(p.v) + (int64_t)b * p.Lk * p.ldv + h * DH;
  const int k_end = p.causal ? min(p.Lk, q0 + TQ) : p.Lk;
  const int n_tiles = (k_end + TK - 1) / TK;
  auto load_kv = [&](int t, int buf) {
    load_tile<DH>(Ks[buf], kg, p.ldk, t * TK, k_end);
    load_tile<DH>(Vs[buf], vg, p.ldv, t * TK, k_end);
    if (tid < TK) {
      const int j = t * TK + tid;
      Ms[buf][tid] = (j >= k_end) || (p.key_pad && p.key_pad[(int64_t)b * p.Lk + j]);
    }
  };
  load_tile<DH>(Qs, qg, p.ldq, q0, p.Lq);
  load_kv(0, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  const int g = lane >> 2, c4 = lane & 3;
  const int qrow0 = q0 + warp * 16 + g;
  const float sl2 = rsqrtf((float)DH) * 1.4426950408889634f;
  const int64_t prow0 = ((int64_t)b * p.H + h) * p.Lq + qrow0, prow1 = prow0 + 8;     // dropout row ids
  uint32_t qa[DH / 16][4];
  float o[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  for (int t = 0; t < n_tiles; ++t) {
    const int buf = t & 1;
    if (t + 1 < n_tiles) load_kv(t + 1, buf ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    if (t == 0) load_a_frags<DH>(Qs, warp, lane, qa);
    float s[8][4];
    mma_a_tileT<DH>(s, qa, Ks[buf], lane);
    const int kbase = t * TK;
    // padding bytes of the tile as two ballot words (one LDS.U8 per score otherwise; the kernels are issue-bound)
    const unsigned dead_lo = __ballot_sync(0xffffffffu, Ms[buf][lane] != 0) >> (c4 * 2);
    const unsigned dead_hi = __ballot_sync(0xffffffffu, Ms[buf][32 + lane] != 0) >> (c4 * 2);
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int kl = nt * 8 + c4 * 2 + e;
        const bool dead = ((nt < 4 ? dead_lo >> (nt * 8 + e) : dead_hi >> ((nt - 4) * 8 + e)) & 1u) != 0;
        const int j = kbase + kl;
        const bool d0 = dead || (p.causal && j > qrow0), d1 = dead || (p.causal && j > qrow0 + 8);
        s[nt][e] = d0 ? -INFINITY : s[nt][e] * sl2;
        s[nt][2 + e] = d1 ? -INFINITY : s[nt][2 + e] * sl2;
        mx0 = fmaxf(mx0, s[nt][e]);
        mx1 = fmaxf(mx1, s[nt][2 + e]);
      }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;
    const float cr0 = fast_exp2(m0 - mu0), cr1 = fast_exp2(m1 - mu1);
    m0 = mn0; m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = fast_exp2(s[nt][0] - mu0); s[nt][1] = fast_exp2(s[nt][1] - mu0);
      s[nt][2] = fast_exp2(s[nt][2] - mu1); s[nt][3] = fast_exp2(s[nt][3] - mu1);
      rs0 += s[nt][0] + s[nt][1];
      rs1 += s[nt][2] + s[nt][3];
    }
    l0 = l0 * cr0 + rs0; l1 = l1 * cr1 + rs1;          // the normaliser is that of the un-dropped softmax
    if (p.drop.p > 0.f) {
#pragma unroll
      for (int gq = 0; gq < 2; ++gq) {                    // one evaluation per row and 32-key group: four 8-key tiles
        const uint32_t j = (uint32_t)(kbase + gq * 32 + c4 * 2);
        const uint4 r0 = drop_block_row8(p.drop, (uint64_t)prow0, j), r1 = drop_block_row8(p.drop, (uint64_t)prow1, j);
        const uint32_t w0[4] = {r0.x, r0.y, r0.z, r0.w}, w1[4] = {r1.x, r1.y, r1.z, r1.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int nt = gq * 4 + q;
          const float2 k0 = drop_keep_pair(p.drop, w0[q]), k1 = drop_keep_pair(p.drop, w1[q]);
          s[nt][0] *= k0.x; s[nt][1] *= k0.y; s[nt][2] *= k1.x; s[nt][3] *= k1.y;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < NT; ++i) { o[i][0] *= cr0; o[i][1] *= cr0; o[i][2] *= cr1; o[i][3] *= cr1; }
    mma_p_tile<DH>(o, s, Vs[buf], lane);
    __syncthreads();
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  bf16* og = static_cast<bf16*>(p.out) + (int64_t)b * p.Lq * p.ldo + h * DH;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int col = nt * 8 + c4 * 2;
    if (qrow0 < p.Lq)
      *reinterpret_cast<__nv_bfloat162*>(og + (int64_t)qrow0 * p.ldo + col) = __floats2bfloat162_rn(o[nt][0] * i0, o[nt][1] * i0);
    if (qrow0 + 8 < p.Lq)
      *reinterpret_cast<__nv_bfloat162*>(og + (int64_t)(qrow0 + 8) * p.ldo + col) = __floats2bfloat162_rn(o[nt][2] * i1, o[nt][3] * i1);
  }
  if (c4 == 0) {
    float* lse = p.lse + ((int64_t)b * p.H + h) * p.Lq;
    if (qrow0 < p.Lq) lse[qrow0] = m0 + log2f(l0);
    if (qrow0 + 8 < p.Lq) lse[qrow0 + 8] = m1 + log2f(l1);
  }
}

// ---- backward, query-tile view: dQ (and delta) -----------------------------------------------------------------
template <int DH>
__global__ void __launch_bounds__(128) flash_train_bwd_dq_kernel(const AttnBwdParams bp) {
  const AttnTrainParams& p = bp.f;
  constexpr int LDS = DH + 8, NT = DH / 8, TILE = 64 * LDS;
  extern __shared__ __align__(16) uint8_t bwd_smem[];            // 6 tiles + masks (55 KB at DH = 64: dynamic)
  bf16* Qs = reinterpret_cast<bf16*>(bwd_smem);
  bf16* dOs = Qs + TILE;
  bf16* Ks[2] = {dOs + TILE, dOs + 2 * TILE};
  bf16* Vs[2] = {dOs + 3 * TILE, dOs + 4 * TILE};
  uint8_t* Ms[2] = {reinterpret_cast<uint8_t*>(dOs + 5 * TILE), reinterpret_cast<uint8_t*>(dOs + 5 * TILE) + TK};
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * TQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bf16* qg = static_cast<const bf16*>(p.q) + (int64_t)b * p.Lq * p.ldq + h * DH;
  const bf16* kg = static_cast<const bf16*>(p.k) + (int64_t)b * p.Lk * p.ldk + h * DH;
  const bf16* vg = static_cast<const bf16*>(p.v) + (int64_t)b * p.Lk * p.ldv + h * DH;
  const bf16* dog = static_cast<const bf16*>(bp.dout) + (int64_t)b * p.Lq * bp.lddo + h * DH;
  const bf16* og = static_cast<const bf16*>(p.out) + (int64_t)b * p.Lq * p.ldo + h * DH;
  const int k_end = p.causal ? min(p.Lk, q0 + TQ) : p.Lk;
  const int n_tiles = (k_end + TK - 1) / TK;
  auto load_kv = [&](int t, int buf) {
    load_tile<DH>(Ks[buf], kg, p.ldk, t * TK, k_end);
    load_tile<DH>(Vs[buf], vg, p.ldv, t * TK, k_end);
    if (tid < TK) {
      const int j = t * TK + tid;
      Ms[buf][tid] = (j >= k_end) || (p.key_pad && p.key_pad[(int64_t)b * p.Lk + j]);
    }
  };
  load_tile<DH>(Qs, qg, p.ldq, q0, p.Lq);
  load_tile<DH>(dOs, dog, bp.lddo, q0, p.Lq);
  load_kv(0, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  const int g = lane >> 2, c4 = lane & 3;
  const int qrow0 = q0 + warp * 16 + g;
  const float scale = rsqrtf((float)DH), sl2 = scale * 1.4426950408889634f;
  const int64_t bh = (int64_t)b * p.H + h;
  const int64_t prow0 = bh * p.Lq + qrow0, prow1 = prow0 + 8;                          // dropout row ids
  const float lse0 = qrow0 < p.Lq ? p.lse[bh * p.Lq + qrow0] : INFINITY;
  const float lse1 = qrow0 + 8 < p.Lq ? p.lse[bh * p.Lq + qrow0 + 8] : INFINITY;
  uint32_t qa[DH / 16][4], da[DH / 16][4];
  float dq[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
  float delta0 = 0.f, delta1 = 0.f;
  for (int t = 0; t < n_tiles; ++t) {
    const int buf = t & 1;
    if (t + 1 < n_tiles) load_kv(t + 1, buf ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    if (t == 0) {
      load_a_frags<DH>(Qs, warp, lane, qa);
      load_a_frags<DH>(dOs, warp, lane, da);
      // delta = rowsum(dO o O) for this thread's two rows (quad-reduced)
      const int lr = warp * 16 + g;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int col = nt * 8 + c4 * 2;
        const float2 d0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&dOs[lr * LDS + col]));
        const float2 d1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&dOs[(lr + 8) * LDS + col]));
        float2 o0 = make_float2(0.f, 0.f), o1 = o0;
        if (qrow0 < p.Lq) o0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(og + (int64_t)qrow0 * p.ldo + col));
        if (qrow0 + 8 < p.Lq) o1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(og + (int64_t)(qrow0 + 8) * p.ldo + col));
        delta0 += d0.x * o0.x + d0.y * o0.y;
        delta1 += d1.x * o1.x + d1.y * o1.y;
      }
      delta0 += __shfl_xor_sync(0xffffffffu, delta0, 1); delta0 += __shfl_xor_sync(0xffffffffu, delta0, 2);
      delta1 += __shfl_xor_sync(0xffffffffu, delta1, 1); delta1 += __shfl_xor_sync(0xffffffffu, delta1, 2);
      if (c4 == 0) {
        if (qrow0 < p.Lq) bp.delta[bh * p.Lq + qrow0] = delta0;
        if (qrow0 + 8 < p.Lq) bp.delta[bh * p.Lq + qrow0 + 8] = delta1;
      }
    }
    float s[8][4], dp[8][4];
    mma_a_tileT<DH>(s, qa, Ks[buf], lane);
    mma_a_tileT<DH>(dp, da, Vs[buf], lane);
    const int kbase = t * TK;
    uint32_t dw0[8], dw1[8];                              // dropout words of this thread's two rows, one per 8-key tile
    if (p.drop.p > 0.f) {
#pragma unroll
      for (int gq = 0; gq < 2; ++gq) {
        const uint32_t jq = (uint32_t)(kbase + gq * 32 + c4 * 2);
        const uint4 r0 = drop_block_row8(p.drop, (uint64_t)prow0, jq), r1 = drop_block_row8(p.drop, (uint64_t)prow1, jq);
        dw0[gq * 4] = r0.x; dw0[gq * 4 + 1] = r0.y; dw0[gq * 4 + 2] = r0.z; dw0[gq * 4 + 3] = r0.w;
        dw1[gq * 4] = r1.x; dw1[gq * 4 + 1] = r1.y; dw1[gq * 4 + 2] = r1.z; dw1[gq * 4 + 3] = r1.w;
      }
    }
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float2 ka = make_float2(1.f, 1.f), kb = make_float2(1.f, 1.f);
      if (p.drop.p > 0.f) { ka = drop_keep_pair(p.drop, dw0[nt]); kb = drop_keep_pair(p.drop, dw1[nt]); }
      const float ks0[2] = {ka.x, ka.y}, ks1[2] = {kb.x, kb.y};
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int kl = nt * 8 + c4 * 2 + e;
        const bool dead = Ms[buf][kl];
        const int j = kbase + kl;
        const bool d0 = dead || (p.causal && j > qrow0), d1 = dead || (p.causal && j > qrow0 + 8);
        const float p0 = d0 ? 0.f : fast_exp2(fmaf(s[nt][e], sl2, -lse0));
        const float p1 = d1 ? 0.f : fast_exp2(fmaf(s[nt][2 + e], sl2, -lse1));
        s[nt][e] = p0 * (dp[nt][e] * ks0[e] - delta0);
        s[nt][2 + e] = p1 * (dp[nt][2 + e] * ks1[e] - delta1);
      }
    }
    mma_p_tile<DH>(dq, s, Ks[buf], lane);               // dQ += dS K
    __syncthreads();
  }
  bf16* dqg = static_cast<bf16*>(bp.dq) + (int64_t)b * p.Lq * bp.lddq + h * DH;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int col = nt * 8 + c4 * 2;
    if (qrow0 < p.Lq)
      *reinterpret_cast<__nv_bfloat162*>(dqg + (int64_t)qrow0 * bp.lddq + col) = __floats2bfloat162_rn(dq[nt][0] * scale, dq[nt][1] * scale);
    if (qrow0 + 8 < p.Lq)
      *reinterpret_cast<__nv_bfloat162*>(dqg + (int64_t)(qrow0 + 8) * bp.lddq + col) = __floats2bfloat162_rn(dq[nt][2] * scale, dq[nt][3] * scale);
  }
}

// ---- backward, key-tile view: dK, dV (the transposed problem: rows = keys, streamed tiles = queries) -------------
template <int DH>
__global__ void __launch_bounds__(128) flash_train_bwd_kv_kernel(const AttnBwdParams bp) {
  const AttnTrainParams& p = bp.f;
  constexpr int LDS = DH + 8, NT = DH / 8, TILE = 64 * LDS;
  extern __shared__ __align__(16) uint8_t bwd_smem[];
  bf16* Ks = reinterpret_cast<bf16*>(bwd_smem);
  bf16* Vs = Ks + TILE;
  bf16* Qs[2] = {Vs + TILE, Vs + 2 * TILE};
  bf16* dOs[2] = {Vs + 3 * TILE, Vs + 4 * TILE};
  float* Ls[2] = {reinterpret_cast<float*>(Vs + 5 * TILE), reinterpret_cast<float*>(Vs + 5 * TILE) + 64};
  float* Ds[2] = {Ls[0] + 128, Ls[0] + 192};
  const int b = blockIdx.z, h = blockIdx.y, k0 = blockIdx.x * TQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bf16* qg = static_cast<const bf16*>(p.q) + (int64_t)b * p.Lq * p.ldq + h * DH;
  const bf16* kg = static_cast<const bf16*>(p.k) + (int64_t)b * p.Lk * p.ldk + h * DH;
  const bf16* vg = static_cast<const bf16*>(p.v) + (int64_t)b * p.Lk * p.ldv + h * DH;
  const bf16* dog = static_cast<const bf16*>(bp.dout) + (int64_t)b * p.Lq * bp.lddo + h * DH;
  const int64_t bh = (int64_t)b * p.H + h;
  const int t_begin = p.causal ? k0 / TK : 0;            // queries i < j see nothing of key j
  const int n_tiles = (p.Lq + TK - 1) / TK;
  auto load_q = [&](int t, int buf) {
    load_tile<DH>(Qs[buf], qg, p.ldq, t * TK, p.Lq);
    load_tile<DH>(dOs[buf], dog, bp.lddo, t * TK, p.Lq);
    if (tid < TK) {
      const int i = t * TK + tid;
      Ls[buf][tid] = i < p.Lq ? p.lse[bh * p.Lq + i] : INFINITY;     // +inf -> P = 0 for the padding rows
      Ds[buf][tid] = i < p.Lq ? bp.delta[bh * p.Lq + i] : 0.f;
    }
  };
  load_tile<DH>(Ks, kg, p.ldk, k0, p.Lk);
  load_tile<DH>(Vs, vg, p.ldv, k0, p.Lk);
  if (t_begin < n_tiles) load_q(t_begin, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  const int g = lane >> 2, c4 = lane & 3;
  const int krow0 = k0 + warp * 16 + g;                  // this thread's keys: krow0 and krow0 + 8
  const uint8_t* pad = p.key_pad ? p.key_pad + (int64_t)b * p.Lk : nullptr;
  const bool dead0 = krow0 >= p.Lk || (pad && pad[krow0]);
  const bool dead1 = krow0 + 8 >= p.Lk || (pad && pad[min(krow0 + 8, p.Lk - 1)]);
  const float scale = rsqrtf((float)DH), sl2 = scale * 1.4426950408889634f;
  uint32_t ka[DH / 16][4], va[DH / 16][4];
  float dk[NT][4], dv[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) { dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f; dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f; }
  for (int t = t_begin; t < n_tiles; ++t) {
    const int buf = (t - t_begin) & 1;
    if (t + 1 < n_tiles) load_q(t + 1, buf ^ 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    if (t == t_begin) {
      load_a_frags<DH>(Ks, warp, lane, ka);
      load_a_frags<DH>(Vs, warp, lane, va);
    }
    float st[8][4], dpt[8][4];
    mma_a_tileT<DH>(st, ka, Qs[buf], lane);               // S^T  [16 keys x 64 queries]
    mma_a_tileT<DH>(dpt, va, dOs[buf], lane);             // dP^T
    const int qbase = t * TK;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      // queries i and i + 1 of this thread, keys krow0 and krow0 + 8 (same 32-key group, same column pair): lane g and
      // its neighbour g ^ 1 (lane ^ 4) need the same two evaluations, so each computes ONE row and they swap the two
      // words that hold keys krow0.. and krow0 + 8..
      const int64_t prow_e = bh * p.Lq + qbase + nt * 8 + c4 * 2;
      float ms0[2] = {1.f, 1.f}, ms1[2] = {1.f, 1.f};
      if (p.drop.p > 0.f) {
        const uint32_t mine = (uint32_t)g & 1u, t0 = ((uint32_t)krow0 >> 3) & 3u;
        const uint4 blk = drop_block_row8(p.drop, (uint64_t)(prow_e + mine), (uint32_t)krow0);
        const uint32_t wa = drop_word(blk, t0), wb = drop_word(blk, t0 + 1);
        const uint32_t oa = __shfl_xor_sync(0xffffffffu, wa, 4), ob = __shfl_xor_sync(0xffffffffu, wb, 4);
        const uint32_t half = (uint32_t)krow0 & 1u;
        const float ma = drop_keep_half(p.drop, wa, half), mb = drop_keep_half(p.drop, wb, half);
        const float xa = drop_keep_half(p.drop, oa, half), xb = drop_keep_half(p.drop, ob, half);
        ms0[0] = mine ? xa : ma; ms0[1] = mine ? ma : xa;
        ms1[0] = mine ? xb : mb; ms1[1] = mine ? mb : xb;
      }
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int ql = nt * 8 + c4 * 2 + e;
        const int i = qbase + ql;
        const float lse = Ls[buf][ql], dl = Ds[buf][ql];
        const bool d0 = dead0 || (p.causal && krow0 > i), d1 = dead1 || (p.causal && krow0 + 8 > i);
        const float p0 = d0 ? 0.f : fast_exp2(fmaf(st[nt][e], sl2, -lse));
        const float p1 = d1 ? 0.f : fast_exp2(fmaf(st[nt][2 + e], sl2, -lse));
        const float m0 = ms0[e], m1 = ms1[e];
        const float g0 = dpt[nt][e] * m0, g1 = dpt[nt][2 + e] * m1;
        st[nt][e] = p0 * m0;                              // (P o mask)^T  -> dV
        st[nt][2 + e] = p1 * m1;
        dpt[nt][e] = p0 * (g0 - dl);                      // dS^T          -> dK
        dpt[nt][2 + e] = p1 * (g1 - dl);
      }
    }
    mma_p_tile<DH>(dv, st, dOs[buf], lane);
    mma_p_tile<DH>(dk, dpt, Qs[buf], lane);
    __syncthreads();
  }
  bf16* dkg = static_cast<bf16*>(bp.dk) + (int64_t)b * p.Lk * bp.lddk + h * DH;
  bf16* dvg = static_cast<bf16*>(bp.dv) + (int64_t)b * p.Lk * bp.lddv + h * DH;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int col = nt * 8 + c4 * 2;
    if (krow0 < p.Lk) {
      *reinterpret_cast<__nv_bfloat162*>(dkg + (int64_t)krow0 * bp.lddk + col) = __floats2bfloat162_rn(dk[nt][0] * scale, dk[nt][1] * scale);
      *reinterpret_cast<__nv_bfloat162*>(dvg + (int64_t)krow0 * bp.lddv + col) = __floats2bfloat162_rn(dv[nt][0], dv[nt][1]);
    }
    if (krow0 + 8 < p.Lk) {
      *reinterpret_cast<__nv_bfloat162*>(dkg + (int64_t)(krow0 + 8) * bp.lddk + col) = __floats2bfloat162_rn(dk[nt][2] * scale, dk[nt][3] * scale);
      *reinterpret_cast<__nv_bfloat162*>(dvg + (int64_t)(krow0 + 8) * bp.lddv + col) = __floats2bfloat162_rn(dv[nt][2], dv[nt][3]);
    }
  }
}

template <int DH> constexpr int bwd_smem_bytes() { return 6 * 64 * (DH + 8) * 2 + 1024; }

void check_pitches(const AttnTrainParams& p) {
  RETR_CHECK(p.dh == 32 || p.dh == 64, "head dim must be 32 or 64");
  RETR_CHECK(p.ldq % 8 == 0 && p.ldk % 8 == 0 && p.ldv % 8 == 0 && p.ldo % 2 == 0, "attention pitches must be 16-byte multiples");
  RETR_CHECK(p.lse != nullptr, "flash training attention needs the lse buffer");
}

}  // namespace

void flash_train_fwd_bf16(const AttnTrainParams& p, cudaStream_t s) {
  check_pitches(p);
  dim3 grid((unsigned)cdiv(p.Lq, TQ), (unsigned)p.H, (unsigned)p.B);
  if (p.dh == 32) flash_train_fwd_kernel<32><<<grid, 128, 0, s>>>(p);
  else flash_train_fwd_kernel<64><<<grid, 128, 0, s>>>(p);
  LAUNCH_OK();
}

void flash_train_bwd_bf16(const AttnBwdParams& bp, cudaStream_t s) {
  const AttnTrainParams& p = bp.f;
  check_pitches(p);
  RETR_CHECK(bp.delta && p.out && bp.lddo % 8 == 0 && bp.lddq % 2 == 0 && bp.lddk % 2 == 0 && bp.lddv % 2 == 0,
             "flash attention backward operands");
  static bool configured = false;
  if (!configured) {
    CUDA_OK(cudaFuncSetAttribute(flash_train_bwd_dq_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, bwd_smem_bytes<32>()));
    CUDA_OK(cudaFuncSetAttribute(flash_train_bwd_dq_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, bwd_smem_bytes<64>()));
    CUDA_OK(cudaFuncSetAttribute(flash_train_bwd_kv_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, bwd_smem_bytes<32>()));
    CUDA_OK(cudaFuncSetAttribute(flash_train_bwd_kv_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, bwd_smem_bytes<64>()));
    configured = true;
  }
  dim3 gq((unsigned)cdiv(p.Lq, TQ), (unsigned)p.H, (unsigned)p.B);
  dim3 gk((unsigned)cdiv(p.Lk, TQ), (unsigned)p.H, (unsigned)p.B);
  if (p.dh == 32) {
    flash_train_bwd_dq_kernel<32><<<gq, 128, bwd_smem_bytes<32>(), s>>>(bp);
    LAUNCH_OK();
    flash_train_bwd_kv_kernel<32><<<gk, 128, bwd_smem_bytes<32>(), s>>>(bp);
  } else {
    flash_train_bwd_dq_kernel<64><<<gq, 128, bwd_smem_bytes<64>(), s>>>(bp);
    LAUNCH_OK();
    flash_train_bwd_kv_kernel<64><<<gk, 128, bwd_smem_bytes<64>(), s>>>(bp);
  }
  LAUNCH_OK();
}

}  // namespace retr
