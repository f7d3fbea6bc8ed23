// bf16 flash attention on the warp-level tensor path (mma.sync m16n8k16, fp32 accumulate) for the
// full-sequence attention of the encoder and of the teacher-forced decoder.  Head dim 32/64 and
// sequence lengths <= 397 make this op softmax/exp-bound, not MMA-bound (1.8% of the forward FLOPs),
// so it keeps scores in registers (FA2 layout reuse: the S accumulator fragment is the A fragment of
// P.V) instead of a TMEM round trip.  Same math as attention.cu: q/sqrt(dh), -inf masks (key padding,
// causal), softmax, P.V (torch/nn/functional.py:6630-6664).
//
// CTA = 4 warps x 16 queries; K/V tiles of 64 keys double-buffered in shared memory via cp.async.
#include "ops.h"

namespace retr {
namespace {

constexpr int FQ = 64;     // queries per CTA
constexpr int FK = 64;     // keys per tile

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

// one MUFU.EX2 (exp2f adds a denormal-range rescale the softmax does not need: tiny probabilities flush to zero)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int DH, bool CAUSAL>
__global__ void __launch_bounds__(128, 6) flash_attention_kernel(const AttnParams p) {
  constexpr int LDS = DH + 8;                       // padded row (elements): conflict-free ldmatrix
  constexpr int KS = DH / 16;                       // k-steps of Q.K^T
  constexpr int NT = DH / 8;                        // n-tiles of the output
  constexpr int CPR = DH / 8;                       // 16-byte chunks per row
  __shared__ __align__(16) bf16 Qs[FQ * LDS];
  __shared__ __align__(16) bf16 Ks[2][FK * LDS];
  __shared__ __align__(16) bf16 Vs[2][FK * LDS];
  __shared__ uint8_t Ms[2][FK];
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * FQ;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // packed sequences: this sample's rows start at seq_off[b] and there are seq_len[b] of them
  const int Lq = p.seq_len ? __ldg(p.seq_len + b) : p.Lq, Lk = p.seq_len ? Lq : p.Lk;
  if (q0 >= Lq) return;
  const int64_t q_row0 = p.seq_off ? (int64_t)__ldg(p.seq_off + b) : (int64_t)b * p.Lq;
  const int64_t k_row0 = p.seq_off ? q_row0 : (int64_t)b * p.Lk;
  const bf16* qg = static_cast<const bf16*>(p.q) + q_row0 * p.ldq + h * DH;
  const bf16* kg = static_cast<const bf16*>(p.k) + k_row0 * p.ldk + h * DH;
  const bf16* vg = static_cast<const bf16*>(p.v) + k_row0 * p.ldv + h * DH;
  const int k_end = CAUSAL ? min(Lk, q0 + FQ) : Lk;
  const int n_tiles = (k_end + FK - 1) / FK;

  auto load_kv = [&](int t, int buf) {
    const int k0 = t * FK;
    for (int idx = tid; idx < FK * CPR; idx += 128) {
      const int r = idx / CPR, c = idx % CPR;
      const bool ok = k0 + r < k_end;
      const int64_t row = ok ? k0 + r : 0;
      cp16((uint32_t)__cvta_generic_to_shared(&Ks[buf][r * LDS + c * 8]), kg + row * p.ldk + c * 8, ok);
      cp16((uint32_t)__cvta_generic_to_shared(&Vs[buf][r * LDS + c * 8]), vg + row * p.ldv + c * 8, ok);
    }
    if (tid < FK) {
      const int j = k0 + tid;
      Ms[buf][tid] = (j >= k_end) || (p.key_pad && p.key_pad[(int64_t)b * Lk + j]);
    }
  };
  // Q tile + first K/V tile
  for (int idx = tid; idx < FQ * CPR; idx += 128) {
    const int r = idx / CPR, c = idx % CPR;
    const bool ok = q0 + r < Lq;
    cp16((uint32_t)__cvta_generic_to_shared(&Qs[r * LDS + c * 8]), qg + (int64_t)(ok ? q0 + r : 0) * p.ldq + c * 8, ok);
  }
  load_kv(0, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");

  const int g = lane >> 2, c4 = lane & 3;
  const int qrow0 = q0 + warp * 16 + g;              // this thread's rows: qrow0 and qrow0 + 8
  const float sl2 = rsqrtf((float)DH) * 1.4426950408889634f;   // scale * log2(e)
  uint32_t qa[KS][4];
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
    if (t == 0) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&Qs[(warp * 16 + (lane & 15)) * LDS + ks * 16 + (lane >> 4) * 8]);
        ldsm_x4(addr, qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
      }
    }
    // S = Q K^T for 64 keys: 8 n-tiles
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int np = 0; np < 4; ++np) {                 // pairs of n-tiles (16 keys)
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        uint32_t b0, b1, b2, b3;
        const int key = np * 16 + (lane >> 4) * 8 + (lane & 7);
        const int kc = ks * 16 + ((lane >> 3) & 1) * 8;
        ldsm_x4((uint32_t)__cvta_generic_to_shared(&Ks[buf][key * LDS + kc]), b0, b1, b2, b3);
        mma_16816(s[2 * np], qa[ks], b0, b1);
        mma_16816(s[2 * np + 1], qa[ks], b2, b3);
      }
    }
    // masks + online softmax (base-2).  The 64 padding bytes of the tile become two ballot words once per warp; the
    // per-score test is then a bit of a register (the kernel is issue-bound: one LDS.U8 + ISETP per score was 15 % of
    // its instructions), and a tile without padding or causal cut skips the masking altogether.
    const int kbase = t * FK;
    const unsigned ball_lo = __ballot_sync(0xffffffffu, Ms[buf][lane] != 0);
    const unsigned ball_hi = __ballot_sync(0xffffffffu, Ms[buf][32 + lane] != 0);
    const unsigned dead_lo = ball_lo >> (c4 * 2), dead_hi = ball_hi >> (c4 * 2);
    float mx0 = -INFINITY, mx1 = -INFINITY;
    if (!CAUSAL && (ball_lo | ball_hi) == 0u) {
      // no masked key in this tile (every full tile of a packed sequence): running maxima only
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
        mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
      }
    } else
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int kl = nt * 8 + c4 * 2 + e;
        const bool dead = ((nt < 4 ? dead_lo >> (nt * 8 + e) : dead_hi >> ((nt - 4) * 8 + e)) & 1u) != 0;
        const int j = kbase + kl;
        const bool d0 = dead || (CAUSAL && j > qrow0), d1 = dead || (CAUSAL && j > qrow0 + 8);
        // scores stay unscaled; scale * log2(e) > 0 is folded into the exponent's FFMA below
        s[nt][e] = d0 ? -INFINITY : s[nt][e];
        s[nt][2 + e] = d1 ? -INFINITY : s[nt][2 + e];
        mx0 = fmaxf(mx0, s[nt][e]);
        mx1 = fmaxf(mx1, s[nt][2 + e]);
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float mu0 = mn0 == -INFINITY ? 0.f : mn0, mu1 = mn1 == -INFINITY ? 0.f : mn1;
    const float cr0 = fast_exp2((m0 - mu0) * sl2), cr1 = fast_exp2((m1 - mu1) * sl2);
    const float ms0 = -mu0 * sl2, ms1 = -mu1 * sl2;
    m0 = mn0; m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = fast_exp2(fmaf(s[nt][0], sl2, ms0)); s[nt][1] = fast_exp2(fmaf(s[nt][1], sl2, ms0));
      s[nt][2] = fast_exp2(fmaf(s[nt][2], sl2, ms1)); s[nt][3] = fast_exp2(fmaf(s[nt][3], sl2, ms1));
      rs0 += s[nt][0] + s[nt][1];
      rs1 += s[nt][2] + s[nt][3];
    }
    l0 = l0 * cr0 + rs0; l1 = l1 * cr1 + rs1;
#pragma unroll
    for (int i = 0; i < NT; ++i) { o[i][0] *= cr0; o[i][1] *= cr0; o[i][2] *= cr1; o[i][3] *= cr1; }
    // O += P V : k-steps of 16 keys, n-tiles of 8 head-dims (two per ldmatrix.x4.trans)
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int np = 0; np < NT / 2; ++np) {
        uint32_t b0, b1, b2, b3;
        const int key = kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7);
        const int col = np * 16 + (lane >> 4) * 8;
        ldsm_x4_t((uint32_t)__cvta_generic_to_shared(&Vs[buf][key * LDS + col]), b0, b1, b2, b3);
        mma_16816(o[2 * np], pa, b0, b1);
        mma_16816(o[2 * np + 1], pa, b2, b3);
      }
    }
    __syncthreads();                                 // everyone done with buf before it is refilled
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  bf16* og = static_cast<bf16*>(p.out) + q_row0 * p.ldo + h * DH;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int col = nt * 8 + c4 * 2;
    if (qrow0 < Lq)
      *reinterpret_cast<__nv_bfloat162*>(og + (int64_t)qrow0 * p.ldo + col) = __floats2bfloat162_rn(o[nt][0] * i0, o[nt][1] * i0);
    if (qrow0 + 8 < Lq)
      *reinterpret_cast<__nv_bfloat162*>(og + (int64_t)(qrow0 + 8) * p.ldo + col) = __floats2bfloat162_rn(o[nt][2] * i1, o[nt][3] * i1);
  }
}

}  // namespace

void flash_attention_bf16(const AttnParams& p, int dh, cudaStream_t s) {
  RETR_CHECK(p.ldq % 8 == 0 && p.ldk % 8 == 0 && p.ldv % 8 == 0 && p.ldo % 2 == 0, "attention pitches must be 16-byte multiples");
  dim3 grid((unsigned)cdiv(p.Lq, FQ), (unsigned)p.H, (unsigned)p.B);
  if (dh == 32) {
    if (p.causal) flash_attention_kernel<32, true><<<grid, 128, 0, s>>>(p);
    else flash_attention_kernel<32, false><<<grid, 128, 0, s>>>(p);
  } else {
    if (p.causal) flash_attention_kernel<64, true><<<grid, 128, 0, s>>>(p);
    else flash_attention_kernel<64, false><<<grid, 128, 0, s>>>(p);
  }
  LAUNCH_OK();
}

}  // namespace retr
