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
// softmax threads as a K-major swizzled A tile.  One thread per (query row, head) reads its 64 scores from TMEM
// (tcgen05.ld) and keeps an online-softmax state; the running output accumulates IN TMEM across chunks with a lazy,
// row-private rescale (tcgen05.ld / tcgen05.st only when the row maximum has grown by more than 2^8).  K and V tiles
// are double-buffered; two CTAs per SM (113 KB shared memory, 256 TMEM columns each).
#include "ops.h"
#include "sm100_ptx.cuh"

namespace retr {
namespace {

constexpr int TQ = 128;                  // queries per CTA = TMEM lanes
constexpr int TK = 64;                   // keys per head per chunk
constexpr int TILE = 128 * 128;          // bytes of one 128-row x 128-byte tile
constexpr int TC_SMEM = 7 * TILE + 24 + 128 + 8;   // Q, K x2, V x2, P (2 k-blocks) + barriers: 2 CTAs per SM need <= 113 KB each

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
  for (int idx = tid; idx < 128 * 8; idx += 256) {
    const int row = idx >> 3, c = idx & 7;
    const int hb = row >> 6, key = k0 + (row & 63);
    const bool mine = (c >> 2) == hb && key < k_end;
    const bf16* src = base + (int64_t)(mine ? key : 0) * ld + (ha + hb) * 32 + (c & 3) * 8;
    cp16z(swz(tile, row, c), src, mine);
  }
}

// 32 lanes x 32 consecutive fp32 columns: registers -> TMEM (the row-private rescale of the running output)
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// 256 threads: thread (row, head) = (tid % 128, tid / 128) — warps w and w + 4 share TMEM lane quadrant w.  The running
// output stays in TMEM (P.V accumulates across chunks); its rescale is LAZY: the exponent base of a row only moves
// when the running maximum has grown by more than 8 (log2 units), so P <= 256 and the usual per-chunk read-back of O
// disappears — nothing waits for a P.V MMA except the P buffer's reuse one chunk later.
__global__ void __launch_bounds__(256, 2) attention_tc_kernel(const AttnParams p) {
  // no static shared memory: the dynamic region then starts at the CTA's (1 KB aligned) shared-memory base, which the
  // 128B-swizzled tiles need, without spending a kilobyte of slack that would cost the second CTA per SM
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t Qs = ptx::smem_u32(smem), Ks0 = Qs + TILE, Vs0 = Ks0 + 2 * TILE, Ps = Vs0 + 2 * TILE;   // K, V x2; P: 2 k-blocks
  uint64_t& bar_qk = *reinterpret_cast<uint64_t*>(smem + 7 * TILE);
  uint64_t& bar_pv = *reinterpret_cast<uint64_t*>(smem + 7 * TILE + 8);
  uint32_t& tmem_holder = *reinterpret_cast<uint32_t*>(smem + 7 * TILE + 16);
  uint32_t* dead_w = reinterpret_cast<uint32_t*>(smem + 7 * TILE + 24);      // one bit per key, 32 keys per word (<= 1024 keys)
  if ((Qs & 1023u) != 0) __trap();
  const int b = blockIdx.z, hp = blockIdx.y, q0 = blockIdx.x * TQ, ha = 2 * hp;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = tid & 127, h = tid >> 7;                   // this thread's query row and head (0 = a, 1 = b)
  const bf16* qg = static_cast<const bf16*>(p.q) + (int64_t)b * p.Lq * p.ldq;
  const bf16* kg = static_cast<const bf16*>(p.k) + (int64_t)b * p.Lk * p.ldk;
  const bf16* vg = static_cast<const bf16*>(p.v) + (int64_t)b * p.Lk * p.ldv;
  const int k_end = p.causal ? min(p.Lk, q0 + TQ) : p.Lk;
  const int n_chunks = (k_end + TK - 1) / TK;

  if (warp == 0) ptx::tmem_alloc(&tmem_holder, 256);
  if (tid == 0) { ptx::mbar_init(&bar_qk, 1); ptx::mbar_init(&bar_pv, 1); ptx::fence_barrier_init(); }
  for (int idx = tid; idx < TQ * 8; idx += 256) {            // [Q_a | Q_b]: 64 contiguous elements per query
    const int r = idx >> 3, c = idx & 7;
    const bool ok = q0 + r < p.Lq;
    cp16z(swz(Qs, r, c), qg + (int64_t)(ok ? q0 + r : 0) * p.ldq + ha * 32 + c * 8, ok);
  }
  fill_blockdiag(Ks0, kg, p.ldk, 0, k_end, ha, tid);
  fill_blockdiag(Vs0, vg, p.ldv, 0, k_end, ha, tid);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int w0 = warp; w0 * 32 < n_chunks * TK; w0 += 8) {    // padding / tail bits of every key, once
    const int key = w0 * 32 + lane;
    const bool dead = key >= k_end || (p.key_pad && p.key_pad[(int64_t)b * p.Lk + key]);
    const unsigned w = __ballot_sync(0xffffffffu, dead);
    if (lane == 0) dead_w[w0] = w;
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = tmem_holder;
  const uint32_t t_row = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);    // TMEM lane = query row
  const uint32_t t_S = t_row + h * 64, t_O = t_row + 128 + h * 32;

  const float sl2 = rsqrtf(32.f) * 1.4426950408889634f;
  const int qrow = q0 + row;
  float m_used = -INFINITY, l = 0.f;                         // exponent base in use (lazy), running sum
  uint32_t ph_qk = 0, ph_pv = 0;

  for (int c = 0; c < n_chunks; ++c) {
    const int k0 = c * TK;
    const uint32_t Ks = Ks0 + (c & 1) * TILE, Vs = Vs0 + (c & 1) * TILE;
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
      ptx::mma_commit(&bar_qk);
    }
    ptx::mbar_wait(&bar_qk, ph_qk);
    ph_qk ^= 1;
    ptx::tc_fence_after_sync();

    // ---- this thread's 64 scores of its head ----
    uint32_t r0[32], r1[32];
    ptx::tmem_ld_32x32(t_S, r0);
    ptx::tmem_ld_32x32(t_S + 32, r1);
    ptx::tmem_ld_wait();
    const uint32_t d0 = dead_w[2 * c], d1 = dead_w[2 * c + 1];
    float s[64], mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const bool x0 = ((d0 >> j) & 1u) || (p.causal && k0 + j > qrow);
      const bool x1 = ((d1 >> j) & 1u) || (p.causal && k0 + 32 + j > qrow);
      s[j] = x0 ? -INFINITY : __uint_as_float(r0[j]);
      s[32 + j] = x1 ? -INFINITY : __uint_as_float(r1[j]);
      mx = fmaxf(mx, fmaxf(s[j], s[32 + j]));
    }
    // the previous chunk's P.V has long finished: the P buffer and the other K/V buffers are free, O is quiescent
    if (c > 0) {
      ptx::mbar_wait(&bar_pv, ph_pv);
      ph_pv ^= 1;
      ptx::tc_fence_after_sync();
    }
    if (c + 1 < n_chunks) {
      fill_blockdiag(Ks0 + ((c + 1) & 1) * TILE, kg, p.ldk, k0 + TK, k_end, ha, tid);
      fill_blockdiag(Vs0 + ((c + 1) & 1) * TILE, vg, p.ldv, k0 + TK, k_end, ha, tid);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    // tcgen05.ld / .st are warp-collective (.sync.aligned): the decision to rescale is taken per WARP, and every lane
    // of a rescaling warp then moves its own base up to its own running maximum (factor 1 where nothing changed)
    const bool grow = (mx - m_used) * sl2 > 8.f;             // also the first chunk with a live key (m_used = -inf)
    if (__any_sync(0xffffffffu, grow)) {
      const float m_new = fmaxf(m_used, mx);
      const float f = (m_used == -INFINITY || m_new == m_used) ? 1.f : ex2((m_used - m_new) * sl2);
      if (c > 0) {                                           // uniform: O holds something only after the first P.V
        uint32_t o[32];
        ptx::tmem_ld_32x32(t_O, o);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * f);
        tmem_st_32x32(t_O, o);
      }
      l *= f;
      m_used = m_new;
    }
    const float ms = m_used == -INFINITY ? 0.f : -m_used * sl2;
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
      const uint32_t dst = swz(Ps + h * TILE, row, ch);
      asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(pk.x), "r"(pk.y), "r"(pk.z), "r"(pk.w) : "memory");
    }
    l += rs;
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
                           ptx::smem_desc_mn_sw128(Vs + (kb * 4 + k) * 2048, TILE), idesc, (c | kb | k) != 0);
      ptx::mma_commit(&bar_pv);
    }
  }
  ptx::mbar_wait(&bar_pv, ph_pv);
  ptx::tc_fence_after_sync();
  {
    uint32_t o[32];
    ptx::tmem_ld_32x32(t_O, o);                              // warp-collective: outside the row guard
    ptx::tmem_ld_wait();
    if (qrow < p.Lq) {
      bf16* og = static_cast<bf16*>(p.out) + ((int64_t)b * p.Lq + qrow) * p.ldo + (ha + h) * 32;
      const float inv = 1.f / l;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        uint4 pk;
        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          hh[j] = __floats2bfloat162_rn(__uint_as_float(o[ch * 8 + 2 * j]) * inv, __uint_as_float(o[ch * 8 + 2 * j + 1]) * inv);
        *reinterpret_cast<uint4*>(og + ch * 8) = pk;
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
  return dh == 32 && p.H % 2 == 0 && p.out && p.Lk <= 1024 && p.ldq % 8 == 0 && p.ldk % 8 == 0 && p.ldv % 8 == 0 && p.ldo % 8 == 0;
}

void attention_tc_bf16(const AttnParams& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CUDA_OK(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
    configured = true;
  }
  dim3 grid((unsigned)cdiv(p.Lq, TQ), (unsigned)(p.H / 2), (unsigned)p.B);
  attention_tc_kernel<<<grid, 256, TC_SMEM, s>>>(p);
  LAUNCH_OK();
}

}  // namespace retr

