// fp32 SIMT GEMM / implicit-GEMM convolution: the precision="fp32" parity mode
// (BASELINE.json north_star: "with fp32 math, logits within 1e-4 relative and greedy token
// sequences identical").  Same parameter block and epilogue semantics as the tcgen05 kernel
// (gemm_sm100.cu); plain FFMA accumulation in k order, no tensor cores, no split-K.
#include "gemm.h"

namespace retr {
namespace {

constexpr int TM = 64, TN = 64, TK = 16;

struct SArgs {
  const float* A; const float* W; void* C;
  int64_t M; int N, K; int64_t lda, ldc;
  const float* scale; const float* bias; const float* rowbias; int period;
  const void* residual; int64_t ldr; int res_fp32, relu, out_fp32;
  unsigned long long* argmax_out;
  int kh, kw, stride, pad, dil, H, Win, Cin, OH, OW;
  int c_group, c_group_stride;
  const int* rowbias_dyn;
};

template <bool GATHER>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const SArgs a) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  __shared__ int64_t row_img[TM];
  __shared__ int row_ih0[TM], row_iw0[TM];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;
  if (GATHER && tid < TM) {
    const int64_t m = m0 + tid;
    if (m < a.M) {
      const int ow = (int)(m % a.OW);
      const int64_t t = m / a.OW;
      row_img[tid] = t / a.OH;
      row_ih0[tid] = (int)(t % a.OH) * a.stride - a.pad;
      row_iw0[tid] = ow * a.stride - a.pad;
    } else {
      row_img[tid] = -1; row_ih0[tid] = 0; row_iw0[tid] = 0;
    }
  }
  __syncthreads();
  float acc[4][4] = {};
  const int lr = tid >> 2, lk = (tid & 3) * 4;     // loader: row 0..63, k offset 0,4,8,12
  for (int k0 = 0; k0 < a.K; k0 += TK) {
    float4 av = make_float4(0.f, 0.f, 0.f, 0.f), bv = av;
    {
      const int64_t m = m0 + lr;
      const int k = k0 + lk;
      if (GATHER) {
        const int tap = k / a.Cin, c = k - tap * a.Cin;
        const int fr = tap / a.kw, fs = tap - fr * a.kw;
        const int ih = row_ih0[lr] + fr * a.dil, iw = row_iw0[lr] + fs * a.dil;
        if (row_img[lr] >= 0 && ih >= 0 && ih < a.H && iw >= 0 && iw < a.Win)
          av = *reinterpret_cast<const float4*>(a.A + ((row_img[lr] * a.H + ih) * a.Win + iw) * a.Cin + c);
      } else if (m < a.M) {
        av = *reinterpret_cast<const float4*>(a.A + m * a.lda + k);
      }
      const int n = n0 + lr;
      if (n < a.N) bv = *reinterpret_cast<const float4*>(a.W + (int64_t)n * a.K + k);
    }
    As[lk][lr] = av.x; As[lk + 1][lr] = av.y; As[lk + 2][lr] = av.z; As[lk + 3][lr] = av.w;
    Bs[lk][lr] = bv.x; Bs[lk + 1][lr] = bv.y; Bs[lk + 2][lr] = bv.z; Bs[lk + 3][lr] = bv.w;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      const float4 x = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 y = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float xa[4] = {x.x, x.y, x.z, x.w}, yb[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xa[i], yb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= a.M) continue;
    float best_v = -INFINITY;
    uint32_t best_i = 0;
    const int64_t mc = a.c_group ? (m / a.c_group) * a.c_group_stride + (m % a.c_group) : m;
    const int64_t rb_row = a.rowbias_dyn ? (int64_t)*a.rowbias_dyn : (m % a.period);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= a.N) continue;
      float v = acc[i][j];
      if (a.scale) v *= a.scale[n];
      if (a.bias) v += a.bias[n];
      if (a.rowbias) v += a.rowbias[rb_row * a.N + n];
      if (a.relu == 1) v = fmaxf(v, 0.f);
      if (a.residual)
        v += a.res_fp32 ? static_cast<const float*>(a.residual)[m * a.ldr + n]
                        : __bfloat162float(static_cast<const bf16*>(a.residual)[m * a.ldr + n]);
      if (a.relu == 2) v = fmaxf(v, 0.f);
      if (a.argmax_out && v > best_v) { best_v = v; best_i = (uint32_t)n; }
      if (a.C) {
        static_cast<float*>(a.C)[mc * a.ldc + n] = v;          // fp32 mode: every activation is fp32
      }
    }
    if (a.argmax_out && best_v > -INFINITY) atomicMax(a.argmax_out + m, pack_argmax(best_v, best_i));
  }
}

}  // namespace

void gemm_fp32(const GemmParams& p, cudaStream_t stream) {
  RETR_CHECK(p.M > 0 && p.N > 0 && p.K > 0, "empty GEMM");
  RETR_CHECK(p.K % TK == 0, "fp32 GEMM needs K % 16 == 0");
  const bool gather = p.conv.kh > 0;
  SArgs a{};
  a.A = static_cast<const float*>(p.A); a.W = static_cast<const float*>(p.W); a.C = p.C;
  a.M = p.M; a.N = p.N; a.K = p.K; a.lda = p.lda; a.ldc = p.ldc;
  a.scale = p.scale; a.bias = p.bias; a.rowbias = p.rowbias;
  a.period = p.rowbias_period > 0 ? p.rowbias_period : 1;
  a.residual = p.residual; a.ldr = p.ldr; a.res_fp32 = p.res_fp32; a.relu = p.relu; a.out_fp32 = p.out_fp32;
  a.argmax_out = p.argmax_out;
  a.c_group = p.c_group; a.c_group_stride = p.c_group_stride; a.rowbias_dyn = p.rowbias_dyn;
  if (gather) {
    const ConvView& c = p.conv;
    RETR_CHECK(c.C % TK == 0 && p.K == c.kh * c.kw * c.C, "conv view mismatch");
    a.kh = c.kh; a.kw = c.kw; a.stride = c.stride; a.pad = c.pad; a.dil = c.dil;
    a.H = c.H; a.Win = c.W; a.Cin = c.C; a.OH = c.OH; a.OW = c.OW;
  } else {
    RETR_CHECK(p.lda % 4 == 0, "lda must be a multiple of 4");
  }
  dim3 grid((unsigned)cdiv(p.M, TM), (unsigned)cdiv(p.N, TN));
  if (gather) gemm_simt_kernel<true><<<grid, 256, 0, stream>>>(a);
  else gemm_simt_kernel<false><<<grid, 256, 0, stream>>>(a);
  LAUNCH_OK();
}

}  // namespace retr
