// Latency-optimised GEMM for the decode step (M = batch <= ~1k rows, K in {256, 512, 2048}):
//   C[M,N] = epi( A · W[N,K]^T ),   A = bf16 matrix  OR  A = LayerNorm(X fp32) computed in the prologue.
// One decode step is a chain of ~40 tiny dependent GEMMs, so what matters is the dependent-latency chain
// of each kernel, not throughput: this kernel has no TMEM allocation, no tensor-map fetch and no
// multi-stage pipeline — one cp.async wave brings the whole K panel of a 64x32 tile into shared memory,
// 4 warps run mma.sync m16n8k16 over it and store.  Fusing the pre-norm LayerNorm
// (models/transformer_modules.py:31,58,89) into the prologue removes one kernel per GEMM.
// K = 2048 (FFN second linear) is split over blockIdx.z; the last CTA of a tile (atomic ticket) adds the
// partial sums in a fixed order, so the result is deterministic.
// The encoder-sized GEMMs use the tcgen05 kernel (gemm_sm100.cu); this one only serves M <= a few hundred.
#include "small_gemm.h"

#include <mutex>
#include <string>

namespace retr {
namespace {
RETR_TRACE_DECL

constexpr int SM_BM = 64, SM_BN = 32, SM_THREADS = 128, SM_PAD = 8;

__device__ __forceinline__ void sg_ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void sg_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void sg_cp16(uint32_t dst, const void* src, bool ok) {
  const uint32_t n = ok ? 16u : 0u;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}

// KC = K panel held in shared memory per iteration (256 or 512); LN requires K == KC.
template <int KC>
__global__ void __launch_bounds__(SM_THREADS) small_gemm_kernel(const SmallGemmParams p) {
  constexpr int LDS = KC + SM_PAD;                  // padded row (elements): conflict-free ldmatrix
  constexpr int CPR = KC / 8;                       // 16-byte chunks per row
  extern __shared__ __align__(16) uint8_t sg_smem[];
  const int chunks = p.K / KC / p.split_k;          // K panels handled by this CTA
  const int nbuf = chunks > 1 ? 2 : 1;              // one panel: no double buffer (half the smem -> co-residency)
  bf16* As = reinterpret_cast<bf16*>(sg_smem);      // [nbuf][64][LDS]
  bf16* Ws = As + nbuf * SM_BM * LDS;               // [nbuf][32][LDS]
  __shared__ int s_ticket;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.y * SM_BM, n0 = blockIdx.x * SM_BN;
  const int kc0 = blockIdx.z * chunks;

  auto load_w = [&](int c, int buf) {
    const bf16* wg = p.W + (int64_t)(kc0 + c) * KC;
    for (int idx = tid; idx < SM_BN * CPR; idx += SM_THREADS) {
      const int r = idx / CPR, ch = idx % CPR;
      const bool ok = n0 + r < p.N;
      sg_cp16((uint32_t)__cvta_generic_to_shared(Ws + ((size_t)buf * SM_BN + r) * LDS + ch * 8),
              wg + (int64_t)(ok ? n0 + r : 0) * p.K + ch * 8, ok);
    }
  };
  auto load_a = [&](int c, int buf) {
    const bf16* ag = p.A + (int64_t)(kc0 + c) * KC;
    for (int idx = tid; idx < SM_BM * CPR; idx += SM_THREADS) {
      const int r = idx / CPR, ch = idx % CPR;
      const bool ok = m0 + r < p.M;
      sg_cp16((uint32_t)__cvta_generic_to_shared(As + ((size_t)buf * SM_BM + r) * LDS + ch * 8),
              ag + (int64_t)(ok ? m0 + r : 0) * p.lda + ch * 8, ok);
    }
  };

  TRACE_BEGIN();
  pdl_launch_dependents();
  load_w(0, 0);                                     // weights do not depend on the previous kernel
  pdl_wait();                                       // ... activations, step counter and residual do
  if (!p.X) load_a(0, 0);
  asm volatile("cp.async.commit_group;" ::: "memory");
  // Epilogue operands are fetched NOW (same memory round trip as the operand panels) so the tail of the
  // kernel has no dependent global loads: bias + rowbias summed per column, residual per element.
  const int g = lane >> 2, c4 = lane & 3;
  const int r0 = m0 + warp * 16 + g;
  float add[4][2], res[4][4];
  {
    const int64_t rb_row = (p.rowbias && p.rowbias_dyn) ? (int64_t)__ldg(p.rowbias_dyn) : 0;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int n = n0 + nt * 8 + c4 * 2;
      add[nt][0] = add[nt][1] = 0.f;
      res[nt][0] = res[nt][1] = res[nt][2] = res[nt][3] = 0.f;
      if (n < p.N) {
        if (p.bias) { add[nt][0] += __ldg(p.bias + n); add[nt][1] += __ldg(p.bias + n + 1); }
        if (p.rowbias) {
          add[nt][0] += __ldg(p.rowbias + rb_row * p.N + n);
          add[nt][1] += __ldg(p.rowbias + rb_row * p.N + n + 1);
        }
        if (p.residual) {
          if (r0 < p.M) {
            const float2 t = *reinterpret_cast<const float2*>(p.residual + (int64_t)r0 * p.ldr + n);
            res[nt][0] = t.x; res[nt][1] = t.y;
          }
          if (r0 + 8 < p.M) {
            const float2 t = *reinterpret_cast<const float2*>(p.residual + (int64_t)(r0 + 8) * p.ldr + n);
            res[nt][2] = t.x; res[nt][3] = t.y;
          }
        }
      }
    }
  }
  if (p.X) {
    // LayerNorm prologue: each warp normalises 16 rows of X (fp32) into the bf16 A panel.
    constexpr int PER = KC / 32;                    // values per lane
    constexpr int RB = KC == 256 ? 16 : 8;          // rows in flight per warp (one memory round trip at d=256)
    float4 gam[PER / 4], bet[PER / 4];              // fetched in the same round trip as X
#pragma unroll
    for (int j = 0; j < PER / 4; ++j) {
      gam[j] = __ldg(reinterpret_cast<const float4*>(p.gamma + j * 128 + lane * 4));
      bet[j] = __ldg(reinterpret_cast<const float4*>(p.beta + j * 128 + lane * 4));
    }
    for (int rb = 0; rb < 16; rb += RB) {
      float v[RB][PER];
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        const int m = m0 + warp * 16 + rb + i;
        const float* xr = p.X + (int64_t)(m < p.M ? m : 0) * p.ldx;
#pragma unroll
        for (int j = 0; j < PER / 4; ++j) {
          const float4 t = *reinterpret_cast<const float4*>(xr + j * 128 + lane * 4);
          v[i][4 * j] = t.x; v[i][4 * j + 1] = t.y; v[i][4 * j + 2] = t.z; v[i][4 * j + 3] = t.w;
        }
      }
#pragma unroll
      for (int i = 0; i < RB; ++i) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < PER; ++j) s += v[i][j];
        const float mean = warp_sum(s) * (1.f / KC);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < PER; ++j) { const float t = v[i][j] - mean; q = fmaf(t, t, q); }
        const float rstd = rsqrtf(warp_sum(q) * (1.f / KC) + p.eps);
        bf16* dst = As + (size_t)(warp * 16 + rb + i) * LDS;
#pragma unroll
        for (int j = 0; j < PER / 4; ++j) {
          const int c = j * 128 + lane * 4;
          const float4 g = gam[j], be = bet[j];
          __nv_bfloat162 lo = __floats2bfloat162_rn((v[i][4 * j] - mean) * rstd * g.x + be.x,
                                                    (v[i][4 * j + 1] - mean) * rstd * g.y + be.y);
          __nv_bfloat162 hi = __floats2bfloat162_rn((v[i][4 * j + 2] - mean) * rstd * g.z + be.z,
                                                    (v[i][4 * j + 3] - mean) * rstd * g.w + be.w);
          uint2 pk;
          pk.x = *reinterpret_cast<uint32_t*>(&lo);
          pk.y = *reinterpret_cast<uint32_t*>(&hi);
          *reinterpret_cast<uint2*>(dst + c) = pk;
        }
      }
    }
  }

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

  for (int c = 0; c < chunks; ++c) {
    const int buf = c & 1;
    if (c + 1 < chunks) { load_w(c + 1, buf ^ 1); load_a(c + 1, buf ^ 1); }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    const bf16* a_s = As + (size_t)buf * SM_BM * LDS;
    const bf16* w_s = Ws + (size_t)buf * SM_BN * LDS;
#pragma unroll 4
    for (int ks = 0; ks < KC / 16; ++ks) {
      uint32_t a[4], b0, b1, b2, b3;
      sg_ldsm_x4((uint32_t)__cvta_generic_to_shared(a_s + (size_t)(warp * 16 + (lane & 15)) * LDS + ks * 16 + (lane >> 4) * 8),
                 a[0], a[1], a[2], a[3]);
#pragma unroll
      for (int np = 0; np < 2; ++np) {
        const int row = np * 16 + (lane >> 4) * 8 + (lane & 7);
        const int kc = ks * 16 + ((lane >> 3) & 1) * 8;
        sg_ldsm_x4((uint32_t)__cvta_generic_to_shared(w_s + (size_t)row * LDS + kc), b0, b1, b2, b3);
        sg_mma(acc[2 * np], a, b0, b1);
        sg_mma(acc[2 * np + 1], a, b2, b3);
      }
    }
    __syncthreads();
  }

  if (p.split_k > 1) {
    // deterministic split-K: partial tiles to scratch, the last CTA of the tile sums them in z order
    float* my = p.scratch + ((size_t)blockIdx.z * p.M) * p.N;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int n = n0 + nt * 8 + c4 * 2;
      if (n < p.N) {
        if (r0 < p.M) *reinterpret_cast<float2*>(my + (size_t)r0 * p.N + n) = make_float2(acc[nt][0], acc[nt][1]);
        if (r0 + 8 < p.M) *reinterpret_cast<float2*>(my + (size_t)(r0 + 8) * p.N + n) = make_float2(acc[nt][2], acc[nt][3]);
      }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = (int)atomicAdd(p.tickets + blockIdx.y * gridDim.x + blockIdx.x, 1u);
    __syncthreads();
    if (s_ticket != p.split_k - 1) { TRACE_END(1, p.C); return; }
    __threadfence();
    if (tid == 0) p.tickets[blockIdx.y * gridDim.x + blockIdx.x] = 0;      // ready for the next launch
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int n = n0 + nt * 8 + c4 * 2;
      acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
      if (n < p.N) {
        for (int z = 0; z < p.split_k; ++z) {
          const float* src = p.scratch + ((size_t)z * p.M) * p.N;
          if (r0 < p.M) {
            const float2 t = __ldcg(reinterpret_cast<const float2*>(src + (size_t)r0 * p.N + n));
            acc[nt][0] += t.x; acc[nt][1] += t.y;
          }
          if (r0 + 8 < p.M) {
            const float2 t = __ldcg(reinterpret_cast<const float2*>(src + (size_t)(r0 + 8) * p.N + n));
            acc[nt][2] += t.x; acc[nt][3] += t.y;
          }
        }
      }
    }
  }
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    const int n = n0 + nt * 8 + c4 * 2;
    if (n >= p.N) continue;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int m = r0 + half * 8;
      if (m >= p.M) continue;
      float v0 = acc[nt][2 * half] + add[nt][0], v1 = acc[nt][2 * half + 1] + add[nt][1];
      if (p.relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
      v0 += res[nt][2 * half]; v1 += res[nt][2 * half + 1];
      if (p.out_fp32) *reinterpret_cast<float2*>(static_cast<float*>(p.C) + (int64_t)m * p.ldc + n) = make_float2(v0, v1);
      else *reinterpret_cast<__nv_bfloat162*>(static_cast<bf16*>(p.C) + (int64_t)m * p.ldc + n) = __floats2bfloat162_rn(v0, v1);
    }
  }
  TRACE_END(1, p.C);
}

template <int KC> size_t sg_smem_bytes(int nbuf = 2) { return (size_t)nbuf * (SM_BM + SM_BN) * (KC + SM_PAD) * sizeof(bf16); }

}  // namespace

#ifdef RETR_TRACE
void small_gemm_set_trace(unsigned long long* buf) { CUDA_OK(cudaMemcpyToSymbol(trace_buf_, &buf, sizeof(buf))); }
#endif

void small_gemm_configure() {
  CUDA_OK(cudaFuncSetAttribute(small_gemm_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sg_smem_bytes<256>()));
  CUDA_OK(cudaFuncSetAttribute(small_gemm_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sg_smem_bytes<512>()));
}

bool small_gemm_supported(int M, int N, int K, bool ln) {
  if (N % 8 != 0 || M > 2048) return false;
  if (ln) return K == 256 || K == 512;
  return K % 256 == 0;
}

// ---- split-K small GEMM on STATIC scratch (no allocation anywhere on the call path) ---------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the
// partial tiles and the tile tickets of the deterministic split-K live in file-scope __device__ arrays that every GPU
// loading the library holds until exit (4 MiB + 16 KiB).  The array NAMES are host-side stand-ins in host code, not
// device addresses: the device address is fetched once per device, with that device current, by
// cudaGetSymbolAddress.  Every call shares the arrays and they keep the previous call's contents: the tickets are
// cleared on the stream before each launch (the partial tiles are fully overwritten before they are read); two calls
// that overlap on the GPU (different streams, no dependency) race on them.
constexpr int64_t NA_MAX_SCRATCH_FLOATS = 1 << 20;          // split_k * M * N <= 1 Mi floats (4 MiB)
constexpr int NA_MAX_TILES = 4096;                           // ceil(M/64) * ceil(N/32) <= 4096
static __device__ __align__(256) float na_scratch[NA_MAX_SCRATCH_FLOATS];
static __device__ __align__(256) unsigned na_tickets[NA_MAX_TILES];

namespace {
struct NaDevice {
  float* scratch = nullptr;
  unsigned* tickets = nullptr;
};
NaDevice na_dev[64];
std::mutex na_mutex;
}  // namespace

void small_gemm_splitk_static(const bf16* A, const bf16* W, const float* bias, float* C, int M, int N, int K,
                              int split_k, cudaStream_t s) {
  RETR_CHECK(A && W && C, "null operand");
  RETR_CHECK(M >= 1 && N >= 1 && K >= 1 && split_k >= 1, "sizes must be positive");
  RETR_CHECK(small_gemm_supported(M, N, K, false), "no-alloc small GEMM: needs N % 8 == 0, K % 256 == 0, M <= 2048");
  const int kc = K == 512 ? 512 : 256;
  RETR_CHECK((K / kc) % split_k == 0, "no-alloc small GEMM: split_k must divide K / 256");
  const int64_t need = (int64_t)split_k * M * N;
  const int64_t tiles = cdiv(M, SM_BM) * cdiv(N, SM_BN);
  RETR_CHECK(need <= NA_MAX_SCRATCH_FLOATS,
             "no-alloc small GEMM: split_k * M * N = " + std::to_string(need) + " floats exceeds the static scratch of " +
                 std::to_string(NA_MAX_SCRATCH_FLOATS) + " floats (4 MiB); use fewer splits or the allocating path");
  RETR_CHECK(tiles <= NA_MAX_TILES, "no-alloc small GEMM: more than " + std::to_string(NA_MAX_TILES) + " output tiles");
  int dev = -1;
  CUDA_OK(cudaGetDevice(&dev));
  RETR_CHECK(dev >= 0 && dev < 64, "device ordinal out of range");
  NaDevice nd;
  {
    std::lock_guard<std::mutex> lock(na_mutex);
    if (!na_dev[dev].scratch) {                              // once per device, with that device current
      float* sp = nullptr;
      unsigned* tp = nullptr;
      CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&sp), na_scratch));
      CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&tp), na_tickets));
      small_gemm_configure();                                // opt-in shared memory: an attribute, not an allocation
      na_dev[dev].tickets = tp;
      na_dev[dev].scratch = sp;
    }
    nd = na_dev[dev];
  }
  // the tickets must start at zero; a previous call that was aborted mid-kernel could have left some behind
  CUDA_OK(cudaMemsetAsync(nd.tickets, 0, (size_t)tiles * sizeof(unsigned), s));
  SmallGemmParams p;
  p.A = A; p.lda = K; p.W = W; p.C = C; p.ldc = N; p.out_fp32 = 1;
  p.M = M; p.N = N; p.K = K; p.bias = bias;
  p.split_k = split_k; p.scratch = nd.scratch; p.tickets = nd.tickets;
  small_gemm(p, s);
}

void small_gemm(const SmallGemmParams& p_in, cudaStream_t s) {
  SmallGemmParams p = p_in;
  RETR_CHECK(small_gemm_supported(p.M, p.N, p.K, p.X != nullptr), "small_gemm: unsupported shape");
  const int kc = (p.K % 512 == 0 && (p.X != nullptr || p.K == 512)) ? 512 : 256;
  if (p.split_k < 1) p.split_k = 1;
  RETR_CHECK((p.K / kc) % p.split_k == 0, "small_gemm: split_k must divide the K panels");
  RETR_CHECK(p.split_k == 1 || (p.scratch && p.tickets), "small_gemm: split-K needs scratch and tickets");
  RETR_CHECK(p.N % 2 == 0 && p.ldc % 2 == 0, "small_gemm: even N / ldc");
  dim3 grid((unsigned)cdiv(p.N, SM_BN), (unsigned)cdiv(p.M, SM_BM), (unsigned)p.split_k);
  const int nbuf = p.K / kc / p.split_k > 1 ? 2 : 1;
  if (kc == 512) launch_pdl(small_gemm_kernel<512>, grid, dim3(SM_THREADS), sg_smem_bytes<512>(nbuf), s, p);
  else launch_pdl(small_gemm_kernel<256>, grid, dim3(SM_THREADS), sg_smem_bytes<256>(nbuf), s, p);
}

}  // namespace retr
