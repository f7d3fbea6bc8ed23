// Feed-forward sublayer of one decode step as ONE kernel (pre-norm, models/transformer_modules.py:84-97):
//     y <- y + W2 · relu(W1 · LN(y) + b1) + b2          d_model = 256, d_ff = 256 * CL (CL <= 8)
// In the decode graph this sublayer used to be two dependent launches (LN+FF1, then split-K FF2 with a ticketed
// reduction through global memory): ~22 us of dependent latency per layer for 0.5 GFLOP.  Here a thread-block
// cluster of CL CTAs owns a 32-row block: CTA r takes the r-th 256-wide slice of the hidden dimension, computes
// h_r = relu(LN(y) W1[r]^T + b1[r]) (bf16, stays in shared memory) and its partial y_r = h_r W2[:, r]^T, whose 32-column
// slices it PUSHES into the shared memory of the CTA that owns those columns (st.shared::cluster, fire and forget).
// After one cluster barrier every CTA sums its eight slots in a fixed order (deterministic), adds b2 and the residual
// and writes y in place.  No HBM round trip for h, no atomics, no second launch, no remote-load latency.
// (Also tried: the cross-attention out-projection as a prologue of this kernel — neutral to 0.4 ms slower — and the NEXT
// layer's LN + qkv projection as a tail of this kernel — correct, but 2 ms slower per decode: the
// tail runs on 32 CTAs after two more cluster barriers, while the stand-alone GEMM overlaps its weight fetch with this
// kernel's end through PDL.  Fusing pays when it removes a dependent round trip, not when it serialises one.)
// Both GEMMs are the same shape — A[32 x 256] (LN(y) or h_r) times a 64-row weight panel — so one cp.async ring of
// 64 x 256 panels feeds all eight of them; the first panels are in flight before griddepcontrol.wait (weights do not
// depend on the previous kernel).
#include "ffn_fused.h"

#include <cuda_runtime.h>

namespace retr {
namespace {
RETR_TRACE_DECL

constexpr int FF_BM = 32, FF_D = 256, FF_PANEL = 64, FF_LDS = FF_D + 8, FF_THREADS = 256, FF_NST = 4;
constexpr int FF_PANELS = 2 * FF_D / FF_PANEL;                 // 4 of W1[r], 4 of W2[:, r]

__device__ __forceinline__ void ff_ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ff_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ff_cp16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ uint32_t ff_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void ff_cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
// arrival that orders nothing: "this CTA is running" / "this CTA is done reading" need no fence, and the release form
// costs a membar (~1 us at this kernel's scale)
__device__ __forceinline__ void ff_cluster_arrive_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void ff_cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void ff_st_peer(uint32_t local_addr, uint32_t rank, float2 v) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1,%2};" ::"r"(remote), "f"(v.x), "f"(v.y) : "memory");
}

// one m16n8 output tile per warp (mt = warp & 1, nt = warp >> 1) over K = 256; two accumulator sets over alternate k
__device__ __forceinline__ void ap_tile_mma(const bf16 (*A)[FF_LDS], const bf16 (*W)[FF_LDS], int mt, int nt, int lane,
                                            float (&acc)[4]) {
  float acc2[4] = {0.f, 0.f, 0.f, 0.f};
  acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
  const uint32_t a_base = (uint32_t)__cvta_generic_to_shared(&A[mt * 16 + (lane & 15)][(lane >> 4) * 8]);
  const uint32_t b_base = (uint32_t)__cvta_generic_to_shared(&W[nt * 8 + (lane & 7)][((lane >> 3) & 3) * 8]);
#pragma unroll
  for (int ks = 0; ks < FF_D / 16; ks += 2) {
    uint32_t a[4], a2[4], b0, b1, b2, b3;
    ff_ldsm_x4(a_base + ks * 32, a[0], a[1], a[2], a[3]);
    ff_ldsm_x4(a_base + (ks + 1) * 32, a2[0], a2[1], a2[2], a2[3]);
    ff_ldsm_x4(b_base + ks * 32, b0, b1, b2, b3);               // 8 rows x k [ks*16, +32): fragments of two k-steps
    ff_mma(acc, a, b0, b1);
    ff_mma(acc2, a2, b2, b3);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) acc[k] += acc2[k];
}

struct FfSmem {
  bf16 As[FF_BM][FF_LDS];                    // LN(y) rows
  bf16 Hs[FF_BM][FF_LDS];                    // relu(W1[r] LN(y) + b1[r])
  bf16 Wr[FF_NST][FF_PANEL][FF_LDS];         // weight panel ring
  float Ps[8][FF_BM][FF_D / 8];              // partials of THIS CTA's 32 output columns, one slot per source CTA
};

__global__ void __launch_bounds__(FF_THREADS) ffn_fused_kernel(const FfnFusedParams p) {
  extern __shared__ __align__(128) uint8_t ff_raw[];
  FfSmem& sm = *reinterpret_cast<FfSmem*>(ff_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = (int)ff_cluster_rank(), CL = (int)gridDim.x;
  const int m0 = blockIdx.y * FF_BM;
  const int d_ff = CL * FF_D;

  auto load_panel = [&](int pi, int buf) {
    // pi < 4: rows [rank*256 + pi*64, +64) of W1 [d_ff, 256];  pi >= 4: rows [(pi-4)*64, +64) of W2 [256, d_ff],
    // columns [rank*256, +256)
    const bf16* base = pi < 4 ? p.W1 + ((int64_t)rank * FF_D + pi * FF_PANEL) * FF_D
                              : p.W2 + (int64_t)(pi - 4) * FF_PANEL * d_ff + (int64_t)rank * FF_D;
    const int64_t ld = pi < 4 ? FF_D : d_ff;
#pragma unroll
    for (int k = 0; k < FF_PANEL * (FF_D / 8) / FF_THREADS; ++k) {
      const int idx = tid + k * FF_THREADS, r = idx / (FF_D / 8), ch = idx % (FF_D / 8);
      ff_cp16((uint32_t)__cvta_generic_to_shared(&sm.Wr[buf][r][ch * 8]), base + r * ld + ch * 8);
    }
  };

  TRACE_BEGIN();
  pdl_launch_dependents();
  ff_cluster_arrive_relaxed();                                      // phase 1: "my shared memory exists" (waited on before
                                                                    // the first remote store, five panels from now)
#pragma unroll
  for (int s = 0; s < FF_NST; ++s) {
    load_panel(s, s);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  if (p.pf_k && tid < 2 * FF_BM) {
    // CTA `rank` of the cluster takes head `rank` of its 32 samples: one K and one V block each
    const int b = m0 + (tid >> 1);
    if (b < p.M) {
      const uint32_t bytes = (uint32_t)(p.pf_nvalid ? p.pf_nvalid[b] : p.pf_rows) * p.pf_dh * 2u;
      const bf16* blk = ((tid & 1) ? p.pf_v : p.pf_k) + ((int64_t)b * 8 + rank) * p.pf_rows * p.pf_dh;
      if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(blk), "r"(bytes) : "memory");
    }
  }
  // epilogue operands that do not depend on the previous kernel: this thread's bias values of the four W1 panels
  const int g = lane >> 2, c4 = lane & 3;
  // warp tile: 16 rows (mt) x 16 columns (cg) of the panel output over all of K.  The A fragments (16 k-steps, 64
  // registers) stay in registers for the four panels of a GEMM, so a panel costs one pass over its own 32 KB of shared
  // memory and ONE block barrier — re-reading A per panel was shared-memory-bandwidth bound (1.1 us per panel), and
  // splitting K over two warp groups needed a second barrier and an exchange buffer per panel.
  const int mt = warp & 1, cg = warp >> 1, wc = cg * 16;
  float b1r[4][2][2];
#pragma unroll
  for (int pi = 0; pi < 4; ++pi)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const int col = rank * FF_D + pi * FF_PANEL + wc + nt * 8 + c4 * 2;
      b1r[pi][nt][0] = __ldg(p.b1 + col);
      b1r[pi][nt][1] = __ldg(p.b1 + col + 1);
    }
  float4 gam[2], bet[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    gam[j] = __ldg(reinterpret_cast<const float4*>(p.gamma + j * 128 + lane * 4));
    bet[j] = __ldg(reinterpret_cast<const float4*>(p.beta + j * 128 + lane * 4));
  }
  pdl_wait();                                                       // y is the previous kernel's output
#ifdef RETR_TRACE
  unsigned long long tp[12];
  int tpn = 0;
  if (trace_on_) tp[tpn++] = trace_now();
#endif

  // LayerNorm prologue: warp w normalises rows 4w .. 4w+3 into As (bf16)
  {
    float v[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + warp * 4 + i;
      const float* xr = p.Y + (int64_t)(m < p.M ? m : 0) * FF_D;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(xr + j * 128 + lane * 4);
        v[i][4 * j] = t.x; v[i][4 * j + 1] = t.y; v[i][4 * j + 2] = t.z; v[i][4 * j + 3] = t.w;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[i][j];
      const float mean = warp_sum(s) * (1.f / FF_D);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float t = v[i][j] - mean; q = fmaf(t, t, q); }
      const float rstd = rsqrtf(warp_sum(q) * (1.f / FF_D) + p.eps);
      bf16* dst = &sm.As[warp * 4 + i][0];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 ga = gam[j], be = bet[j];
        __nv_bfloat162 lo = __floats2bfloat162_rn((v[i][4 * j] - mean) * rstd * ga.x + be.x,
                                                  (v[i][4 * j + 1] - mean) * rstd * ga.y + be.y);
        __nv_bfloat162 hi = __floats2bfloat162_rn((v[i][4 * j + 2] - mean) * rstd * ga.z + be.z,
                                                  (v[i][4 * j + 3] - mean) * rstd * ga.w + be.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(dst + j * 128 + lane * 4) = pk;
      }
    }
  }

  uint32_t af[16][4];
#pragma unroll 1
  for (int pi = 0; pi < FF_PANELS; ++pi) {
    const int buf = pi % FF_NST;
    asm volatile("cp.async.wait_group %0;" ::"n"(FF_NST - 2) : "memory");
    __syncthreads();                      // panel pi landed; everyone is done with panel pi-1 (slot free, Hs written)
    if (pi >= 1 && pi - 1 + FF_NST < FF_PANELS) load_panel(pi - 1 + FF_NST, (pi - 1) % FF_NST);
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (pi == 4) ff_cluster_wait();                                 // every CTA of the cluster is running (remote stores ahead)
    if (pi == 0 || pi == 4) {                                       // (re)load this warp's A fragments: 16 rows, all of K
      const bf16(*A)[FF_LDS] = pi < 4 ? sm.As : sm.Hs;
#pragma unroll
      for (int ks = 0; ks < 16; ++ks)
        ff_ldsm_x4((uint32_t)__cvta_generic_to_shared(&A[mt * 16 + (lane & 15)][ks * 16 + (lane >> 4) * 8]),
                   af[ks][0], af[ks][1], af[ks][2], af[ks][3]);
    }
    // two accumulator sets over alternate k-steps: four independent mma chains of length 8
    float acc[2][4], acc2[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
      acc2[nt][0] = acc2[nt][1] = acc2[nt][2] = acc2[nt][3] = 0.f;
    }
    const uint32_t b_base = (uint32_t)__cvta_generic_to_shared(
        &sm.Wr[buf][wc + (lane >> 4) * 8 + (lane & 7)][((lane >> 3) & 1) * 8]);
#pragma unroll
    for (int ks = 0; ks < 16; ks += 2) {
      uint32_t b0, b1, b2, b3, e0, e1, e2, e3;
      ff_ldsm_x4(b_base + ks * 32, b0, b1, b2, b3);
      ff_ldsm_x4(b_base + (ks + 1) * 32, e0, e1, e2, e3);
      ff_mma(acc[0], af[ks], b0, b1);
      ff_mma(acc[1], af[ks], b2, b3);
      ff_mma(acc2[0], af[ks + 1], e0, e1);
      ff_mma(acc2[1], af[ks + 1], e2, e3);
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[nt][k] += acc2[nt][k];
    if (pi < 4) {
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const int col = pi * FF_PANEL + wc + nt * 8 + c4 * 2;
        const float bb0 = pi == 0 ? b1r[0][nt][0] : pi == 1 ? b1r[1][nt][0] : pi == 2 ? b1r[2][nt][0] : b1r[3][nt][0];
        const float bb1 = pi == 0 ? b1r[0][nt][1] : pi == 1 ? b1r[1][nt][1] : pi == 2 ? b1r[2][nt][1] : b1r[3][nt][1];
        *reinterpret_cast<__nv_bfloat162*>(&sm.Hs[mt * 16 + g][col]) =
            __floats2bfloat162_rn(fmaxf(acc[nt][0] + bb0, 0.f), fmaxf(acc[nt][1] + bb1, 0.f));
        *reinterpret_cast<__nv_bfloat162*>(&sm.Hs[mt * 16 + g + 8][col]) =
            __floats2bfloat162_rn(fmaxf(acc[nt][2] + bb0, 0.f), fmaxf(acc[nt][3] + bb1, 0.f));
      }
    } else {
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const int col = (pi - 4) * FF_PANEL + wc + nt * 8 + c4 * 2;  // output column; owner CTA = col / 32
        const uint32_t owner = (uint32_t)(col >> 5);
        const int lc = col & 31;
        ff_st_peer((uint32_t)__cvta_generic_to_shared(&sm.Ps[rank][mt * 16 + g][lc]), owner, make_float2(acc[nt][0], acc[nt][1]));
        ff_st_peer((uint32_t)__cvta_generic_to_shared(&sm.Ps[rank][mt * 16 + g + 8][lc]), owner, make_float2(acc[nt][2], acc[nt][3]));
      }
    }
#ifdef RETR_TRACE
    if (trace_on_) tp[tpn++] = trace_now();
#endif
  }

  // ---- reduction: this CTA owns columns [rank*32, +32) of the 32-row block; its eight partial slots were filled by
  // remote stores, visible after the cluster barrier.  One float4 per thread, slots summed in a fixed order.
  const int row = tid >> 3, lc = (tid & 7) * 4, col = rank * 32 + lc;
  const float4 bias2 = __ldg(reinterpret_cast<const float4*>(p.b2 + col));
  float4 resid = make_float4(0.f, 0.f, 0.f, 0.f);
  if (m0 + row < p.M) resid = *reinterpret_cast<const float4*>(p.Y + (int64_t)(m0 + row) * FF_D + col);
  ff_cluster_arrive();
  ff_cluster_wait();
  {
    float4 s = make_float4(bias2.x + resid.x, bias2.y + resid.y, bias2.z + resid.z, bias2.w + resid.w);
    float4 t[8];
#pragma unroll
    for (int src = 0; src < 8; ++src) t[src] = *reinterpret_cast<const float4*>(&sm.Ps[src][row][lc]);
    float4 part = t[0];
#pragma unroll
    for (int src = 1; src < 8; ++src) { part.x += t[src].x; part.y += t[src].y; part.z += t[src].z; part.w += t[src].w; }
    s.x += part.x; s.y += part.y; s.z += part.z; s.w += part.w;
    if (m0 + row < p.M) *reinterpret_cast<float4*>(p.Y + (int64_t)(m0 + row) * FF_D + col) = s;
  }
#ifdef RETR_TRACE
  if (trace_on_) {
    tp[tpn++] = trace_now();
    const unsigned long long slot_ = atomicAdd(trace_buf_, 2ull);
    if (slot_ + 1 < (1ull << 18)) {
      unsigned long long* r_ = trace_buf_ + 8 + 8 * slot_;
      r_[0] = 9; r_[1] = trace_n_; r_[2] = trace_lin_; r_[3] = trace_t0_;
      r_[4] = tp[0]; r_[5] = tp[1]; r_[6] = tp[2]; r_[7] = tp[3];
      r_[8] = 10; r_[9] = tp[4]; r_[10] = tp[5]; r_[11] = tp[6]; r_[12] = tp[7]; r_[13] = tp[8]; r_[14] = tp[9];
      r_[15] = trace_lin_;
    }
  }
#endif
  TRACE_END(8, p.Y);
}


// ---- attention output projection + residual + LayerNorm + next projection, one cluster kernel ----------------
//     y <- y + Wo · att + bo ;   q = Wq · LN(y) + bq + rowbias[*step]
// (self-attention out-proj followed by the cross-attention query projection of a decoder layer,
// models/transformer_modules.py:58-83 in pre-norm form).  Two dependent 5-8 us launches in the decode graph; here the
// eight CTAs of a cluster each compute a 32-column slice of y for a 32-row block, push it to every peer's shared
// memory (the LayerNorm needs whole rows), and after one cluster barrier each computes a 32-column slice of q.
struct ApSmem {
  bf16 A0[FF_BM][FF_LDS];                    // att rows
  bf16 W0[32][FF_LDS];                       // Wo rows of this CTA's column slice
  bf16 W1[32][FF_LDS];                       // Wq rows of this CTA's column slice
  bf16 As[FF_BM][FF_LDS];                    // LN(y) rows
  float Yfull[FF_BM][FF_D];                  // the block's new y, assembled from the eight slices
};

__global__ void __launch_bounds__(FF_THREADS) attn_out_ln_proj_kernel(const AttnOutProjParams p) {
  extern __shared__ __align__(128) uint8_t ap_raw[];
  ApSmem& sm = *reinterpret_cast<ApSmem*>(ap_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = (int)ff_cluster_rank();
  const int m0 = blockIdx.y * FF_BM;
  const int g = lane >> 2, c4 = lane & 3, mt = warp & 1, nt = warp >> 1;
  TRACE_BEGIN();
  pdl_launch_dependents();
  ff_cluster_arrive_relaxed();                                      // phase 1: this CTA's shared memory exists
  // weights of this CTA's slice: rows [32 rank, +32) of Wo and Wq
#pragma unroll
  for (int k = 0; k < 2 * 32 * (FF_D / 8) / FF_THREADS; ++k) {
    const int idx = tid + k * FF_THREADS, which = idx / (32 * (FF_D / 8)), rem = idx % (32 * (FF_D / 8));
    const int r = rem / (FF_D / 8), ch = rem % (FF_D / 8);
    const bf16* src = (which ? p.Wq : p.Wo) + (int64_t)(rank * 32 + r) * FF_D + ch * 8;
    ff_cp16((uint32_t)__cvta_generic_to_shared(which ? &sm.W1[r][ch * 8] : &sm.W0[r][ch * 8]), src);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  const int col = rank * 32 + nt * 8 + c4 * 2;                      // this thread's output columns col, col + 1
  const float bo0 = __ldg(p.bo + col), bo1 = __ldg(p.bo + col + 1);
  float bq0 = p.bq ? __ldg(p.bq + col) : 0.f, bq1 = p.bq ? __ldg(p.bq + col + 1) : 0.f;
  float4 gam[2], bet[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    gam[j] = __ldg(reinterpret_cast<const float4*>(p.gamma + j * 128 + lane * 4));
    bet[j] = __ldg(reinterpret_cast<const float4*>(p.beta + j * 128 + lane * 4));
  }
  pdl_wait();                                                       // att, y and the step counter come from earlier kernels
#pragma unroll
  for (int k = 0; k < FF_BM * (FF_D / 8) / FF_THREADS; ++k) {
    const int idx = tid + k * FF_THREADS, r = idx / (FF_D / 8), ch = idx % (FF_D / 8);
    const int m = m0 + r < p.M ? m0 + r : 0;
    ff_cp16((uint32_t)__cvta_generic_to_shared(&sm.A0[r][ch * 8]), p.att + (int64_t)m * p.ld_att + ch * 8);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  const int r_lo = m0 + mt * 16 + g, r_hi = r_lo + 8;
  float2 res_lo = make_float2(0.f, 0.f), res_hi = res_lo;
  if (r_lo < p.M) res_lo = *reinterpret_cast<const float2*>(p.Y + (int64_t)r_lo * FF_D + col);
  if (r_hi < p.M) res_hi = *reinterpret_cast<const float2*>(p.Y + (int64_t)r_hi * FF_D + col);
  if (p.rowbias) {
    const int64_t rb = (int64_t)__ldg(p.rowbias_dyn) * FF_D;
    bq0 += __ldg(p.rowbias + rb + col);
    bq1 += __ldg(p.rowbias + rb + col + 1);
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  float acc[4];
  ap_tile_mma(sm.A0, sm.W0, mt, nt, lane, acc);
  const float2 y_lo = make_float2(acc[0] + bo0 + res_lo.x, acc[1] + bo1 + res_lo.y);
  const float2 y_hi = make_float2(acc[2] + bo0 + res_hi.x, acc[3] + bo1 + res_hi.y);
  ff_cluster_wait();                                                // every CTA of the cluster is running
  {
    const uint32_t a_lo = (uint32_t)__cvta_generic_to_shared(&sm.Yfull[mt * 16 + g][col]);
    const uint32_t a_hi = (uint32_t)__cvta_generic_to_shared(&sm.Yfull[mt * 16 + g + 8][col]);
#pragma unroll
    for (int dst = 0; dst < 8; ++dst) {
      ff_st_peer(a_lo, (uint32_t)dst, y_lo);
      ff_st_peer(a_hi, (uint32_t)dst, y_hi);
    }
  }
  ff_cluster_arrive();
  ff_cluster_wait();                                                // whole rows of the new y are here
  // (global stores after the barrier: its release would otherwise have to wait for them)
  if (r_lo < p.M) *reinterpret_cast<float2*>(p.Y + (int64_t)r_lo * FF_D + col) = y_lo;
  if (r_hi < p.M) *reinterpret_cast<float2*>(p.Y + (int64_t)r_hi * FF_D + col) = y_hi;
  {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = warp * 4 + i;
      float v[8];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(&sm.Yfull[r][j * 128 + lane * 4]);
        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
      }
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[j];
      const float mean = warp_sum(s) * (1.f / FF_D);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float t = v[j] - mean; q = fmaf(t, t, q); }
      const float rstd = rsqrtf(warp_sum(q) * (1.f / FF_D) + p.eps);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 ga = gam[j], be = bet[j];
        __nv_bfloat162 lo = __floats2bfloat162_rn((v[4 * j] - mean) * rstd * ga.x + be.x, (v[4 * j + 1] - mean) * rstd * ga.y + be.y);
        __nv_bfloat162 hi = __floats2bfloat162_rn((v[4 * j + 2] - mean) * rstd * ga.z + be.z, (v[4 * j + 3] - mean) * rstd * ga.w + be.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(&sm.As[r][j * 128 + lane * 4]) = pk;
      }
    }
  }
  __syncthreads();
  ap_tile_mma(sm.As, sm.W1, mt, nt, lane, acc);
  if (r_lo < p.M)
    *reinterpret_cast<__nv_bfloat162*>(p.Q + (int64_t)r_lo * p.ld_q + col) = __floats2bfloat162_rn(acc[0] + bq0, acc[1] + bq1);
  if (r_hi < p.M)
    *reinterpret_cast<__nv_bfloat162*>(p.Q + (int64_t)r_hi * p.ld_q + col) = __floats2bfloat162_rn(acc[2] + bq0, acc[3] + bq1);
  TRACE_END(11, p.Y);
}

}  // namespace

#ifdef RETR_TRACE
void ffn_fused_set_trace(unsigned long long* buf) { CUDA_OK(cudaMemcpyToSymbol(trace_buf_, &buf, sizeof(buf))); }
#endif

void ffn_fused_configure() {
  CUDA_OK(cudaFuncSetAttribute(ffn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FfSmem)));
  CUDA_OK(cudaFuncSetAttribute(attn_out_ln_proj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ApSmem)));
}

template <typename P>
static void launch_cluster8(void (*kernel)(P), const P& p, int M, size_t smem, cudaStream_t s) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(8, (unsigned)cdiv(M, FF_BM), 1);
  cfg.blockDim = dim3(FF_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int na = 0;
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = 8; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
  ++na;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, p));
  ++g_launches;
}

void attn_out_ln_proj(const AttnOutProjParams& p, cudaStream_t s) {
  launch_cluster8(attn_out_ln_proj_kernel, p, p.M, sizeof(ApSmem), s);
}

bool ffn_fused_supported(int d_model, int d_ff) {
  if (d_model != FF_D || d_ff % FF_D != 0) return false;
  return d_ff / FF_D == 8;                              // one 32-column slice per CTA and thread-per-float4 reduction
}

void ffn_fused(const FfnFusedParams& p, cudaStream_t s) {
  RETR_CHECK(ffn_fused_supported(FF_D, p.d_ff), "ffn_fused: unsupported shape");
  launch_cluster8(ffn_fused_kernel, p, p.M, sizeof(FfSmem), s);
}

}  // namespace retr
