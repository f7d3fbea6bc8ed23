// Full-sequence attention on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM) for head dim 32:
// the encoder's self-attention (S = 397, key-padding mask) and the teacher-forced decoder's (causal).
// Same math as attention.cu / attention_mma.cu: q/sqrt(dh), -inf masks, softmax, P.V
// (torch/nn/functional.py:6630-6664 as called from models/transformer_modules.py:22-74).
//
// A head of 32 dims is half of what one 128-byte swizzled operand row holds, so a CTA takes a PAIR of heads and
// makes both MMAs block-diagonal:
//   S[128 q, 128]  = [Q_a | Q_b] (128 x 64)  x  [K_a 0 ; 0 K_b]^T      columns 0-63: head a's 64 keys, 64-127: head b's
//   O[128 q, 64]   = [P_a | P_b] (128 x 128) x  [V_a 0 ; 0 V_b]        columns 0-31: head a's dims, 32-63: head b's
// Q, K tiles are K-major, 128B-swizzled rows filled by cp.async (the zero blocks by zero-fill copies); the V tile is
// the B operand in MN-major form ([key][dim] rows exactly as V lies in memory: no transpose), P is written by the
// softmax threads as a K-major swizzled A tile.  One thread per query row reads its 128 scores from TMEM
// (tcgen05.ld), keeps one online-softmax state per head, and folds each chunk's partial P.V (read back from TMEM) into
// 64 fp32 registers.  Sequential phases per 64-key chunk (load -> MMA -> softmax -> MMA -> fold); two CTAs per SM
// overlap each other's phases (80 KB shared memory, 256 TMEM columns each).
#include "ops.h"
#include "sm100_ptx.cuh"

namespace retr {
namespace {

constexpr int TQ = 128;                  // queries per CTA = TMEM lanes
constexpr int TK = 64;                   // keys per head per chunk
constexpr int TILE = 128 * 128;          // bytes of one 128-row x 128-byte tile
constexpr int TC_SMEM = 5 * TILE + 1024; // Q, K, V, P (2 k-blocks) + alignment slack

// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, A K-major, B MN-major (bit 16)
__host__ __device__ constexpr uint32_t idesc_bf16_f32_bmn(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void cp16z(uint32_t dst, const void* src, bool ok) {
  const uint32_t n = ok ? 16u : 0u;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ uint32_t swz(uint32_t tile, int row, int chunk) {
  return tile + row * 128 + ((uint32_t)(chunk ^ (row & 7)) << 4);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Block-diagonal fill of a [128 rows x 64 elements] tile from a [L, ld] matrix whose heads lie side by side:
// row r < 64 = key k0 + r of head a in chunks 0-3, zeros in 4-7; row 64 + r = the same key of head b in chunks 4-7.
__device__ __forceinline__ void fill_blockdiag(uint32_t tile, const bf16* base, int64_t ld, int k0, int k_end, int ha,
                                               int tid) {
  for (int idx = tid; idx < 128 * 8; idx += 128) {
    const int row = idx >> 3, c = idx & 7;
    const int hb = row >> 6, key = k0 + (row & 63);
    const bool mine = (c >> 2) == hb && key < k_end;
    const bf16* src = base + (int64_t)(mine ? key : 0) * ld + (ha + hb) * 32 + (c & 3) * 8;
    cp16z(swz(tile, row, c), src, mine);
  }
}

__global__ void __launch_bounds__(128, 2) attention_tc_kernel(const AttnParams p) {
  extern __shared__ uint8_t tc_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t Qs = ptx::smem_u32(smem), Ks = Qs + TILE, Vs = Ks + TILE, Ps = Vs + TILE;     // Ps: two k-blocks
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_holder;
  __shared__ uint32_t dead_w[2];
  const int b = blockIdx.z, hp = blockIdx.y, q0 = blockIdx.x * TQ, ha = 2 * hp;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bf16* qg = static_cast<const bf16*>(p.q) + (int64_t)b * p.Lq * p.ldq;
  const bf16* kg = static_cast<const bf16*>(p.k) + (int64_t)b * p.Lk * p.ldk;
  const bf16* vg = static_cast<const bf16*>(p.v) + (int64_t)b * p.Lk * p.ldv;
  const int k_end = p.causal ? min(p.Lk, q0 + TQ) : p.Lk;
  const int n_chunks = (k_end + TK - 1) / TK;

  if (warp == 0) ptx::tmem_alloc(&tmem_holder, 256);
  if (tid == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  for (int idx = tid; idx < TQ * 8; idx += 128) {            // [Q_a | Q_b]: 64 contiguous elements per query
    const int row = idx >> 3, c = idx & 7;
    const bool ok = q0 + row < p.Lq;
    cp16z(swz(Qs, row, c), qg + (int64_t)(ok ? q0 + row : 0) * p.ldq + ha * 32 + c * 8, ok);
  }
  fill_blockdiag(Ks, kg, p.ldk, 0, k_end, ha, tid);
  fill_blockdiag(Vs, vg, p.ldv, 0, k_end, ha, tid);
  asm volatile("cp.async.commit_group;" ::: "memory");
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = tmem_holder;
  const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16);      // this thread's TMEM lane = its query row
  const uint32_t t_S = t_row, t_O = t_row + 128;

  const float sl2 = rsqrtf(32.f) * 1.4426950408889634f;
  const int qrow = q0 + tid;
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f}, o[2][32];
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int j = 0; j < 32; ++j) o[h][j] = 0.f;
  uint32_t phase = 0;

  for (int c = 0; c < n_chunks; ++c) {
    const int k0 = c * TK;
    if (tid < 64) {                                          // padding / tail bits of this chunk's 64 keys
      const int key = k0 + tid;
      const bool dead = key >= k_end || (p.key_pad && p.key_pad[(int64_t)b * p.Lk + key]);
      const unsigned w = __ballot_sync(0xffffffffu, dead);
      if (lane == 0) dead_w[warp] = w;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      ptx::tc_fence_after_sync();
      constexpr uint32_t idesc = ptx::idesc_bf16_f32(128, 128);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::mma_bf16_ss(tmem_base, ptx::smem_desc_sw128(Qs + k * 32), ptx::smem_desc_sw128(Ks + k * 32), idesc, k != 0);
      ptx::mma_commit(&bar);
    }
    ptx::mbar_wait(&bar, phase);
    phase ^= 1;
    ptx::tc_fence_after_sync();
    if (c + 1 < n_chunks) fill_blockdiag(Ks, kg, p.ldk, k0 + TK, k_end, ha, tid);      // the K tile is free again

    // ---- softmax of this thread's row, one head at a time; P -> shared memory as the A operand of P.V ----
    float corr[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t r0[32], r1[32];
      ptx::tmem_ld_32x32(t_S + h * 64, r0);
      ptx::tmem_ld_32x32(t_S + h * 64 + 32, r1);
      ptx::tmem_ld_wait();
      const uint32_t d0 = dead_w[0], d1 = dead_w[1];
      float s[64], mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const bool x0 = ((d0 >> j) & 1u) || (p.causal && k0 + j > qrow);
        const bool x1 = ((d1 >> j) & 1u) || (p.causal && k0 + 32 + j > qrow);
        s[j] = x0 ? -INFINITY : __uint_as_float(r0[j]);
        s[32 + j] = x1 ? -INFINITY : __uint_as_float(r1[j]);
        mx = fmaxf(mx, fmaxf(s[j], s[32 + j]));
      }
      const float mn = fmaxf(m[h], mx), mu = mn == -INFINITY ? 0.f : mn;
      corr[h] = ex2((m[h] - mu) * sl2);
      const float ms = -mu * sl2;
      m[h] = mn;
      float rs = 0.f;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        float e[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { e[j] = ex2(fmaf(s[ch * 8 + j], sl2, ms)); rs += e[j]; }
        uint4 pk;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int j = 0; j < 4; ++j) hh[j] = __floats2bfloat162_rn(e[2 * j], e[2 * j + 1]);
        const uint32_t dst = swz(Ps + h * TILE, tid, ch);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w) : "memory");
      }
      l[h] = l[h] * corr[h] + rs;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      ptx::tc_fence_after_sync();
      constexpr uint32_t idesc = idesc_bf16_f32_bmn(128, 64);
#pragma unroll
      for (int kb = 0; kb < 2; ++kb)
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::mma_bf16_ss(tmem_base + 128, ptx::smem_desc_sw128(Ps + kb * TILE + k * 32),
                           ptx::smem_desc_mn_sw128(Vs + (kb * 4 + k) * 2048, TILE), idesc, (kb | k) != 0);
      ptx::mma_commit(&bar);
    }
    ptx::mbar_wait(&bar, phase);
    phase ^= 1;
    ptx::tc_fence_after_sync();
    if (c + 1 < n_chunks) fill_blockdiag(Vs, vg, p.ldv, k0 + TK, k_end, ha, tid);      // the V tile is free again
    asm volatile("cp.async.commit_group;" ::: "memory");
    // ---- fold this chunk's P.V into the running output ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t r[32];
      ptx::tmem_ld_32x32(t_O + h * 32, r);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) o[h][j] = fmaf(o[h][j], corr[h], __uint_as_float(r[j]));
    }
  }
  if (qrow < p.Lq) {
    bf16* og = static_cast<bf16*>(p.out) + ((int64_t)b * p.Lq + qrow) * p.ldo + ha * 32;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float inv = 1.f / l[h];
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        uint4 pk;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int j = 0; j < 4; ++j) hh[j] = __floats2bfloat162_rn(o[h][ch * 8 + 2 * j] * inv, o[h][ch * 8 + 2 * j + 1] * inv);
        *reinterpret_cast<uint4*>(og + h * 32 + ch * 8) = pk;
      }
    }
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace

bool attention_tc_supported(const AttnParams& p, int dh) {
  return dh == 32 && p.H % 2 == 0 && p.out && p.ldq % 8 == 0 && p.ldk % 8 == 0 && p.ldv % 8 == 0 && p.ldo % 8 == 0;
}

void attention_tc_bf16(const AttnParams& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CUDA_OK(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
    configured = true;
  }
  dim3 grid((unsigned)cdiv(p.Lq, TQ), (unsigned)(p.H / 2), (unsigned)p.B);
  attention_tc_kernel<<<grid, 128, TC_SMEM, s>>>(p);
  LAUNCH_OK();
}

}  // namespace retr

