// Non-GEMM kernels of the training step (reference engine.py:52-87, models/transformer_modules.py
// dropout sites, torch.nn.functional cross_entropy / layer_norm / multi_head_attention backward).
// Everything here is HBM- or latency-bound streaming work; the contractions of the backward pass go
// through the tcgen05 GEMM (gemm_sm100.cu) after the layout helpers below have made the contraction
// dimension contiguous.
#include "train_ops.h"
#include "dropout.cuh"

#include <type_traits>

namespace retr {
namespace {

constexpr int EW_THREADS = 256;
inline unsigned ew_blocks(int64_t n) { return (unsigned)std::min<int64_t>(cdiv(n, EW_THREADS), 148 * 16); }

template <typename T> struct V16;
template <> struct V16<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float (&o)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
  }
  __device__ static void store(float* p, const float (&o)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
  }
};
template <> struct V16<bf16> {
  static constexpr int N = 8;
  __device__ static void load(const bf16* p, float (&o)[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
  __device__ static void store(bf16* p, const float (&o)[8]) {
    uint4 t;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = t;
  }
};

// ---- layout helpers -----------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void transpose_pad_kernel(const TI* __restrict__ in, int64_t ld_in, TO* __restrict__ out, int64_t ldo,
                                     int64_t M, int N) {
  __shared__ float tile[32][33];
  const int64_t m0 = (int64_t)blockIdx.x * 32;
  const int n0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int64_t m = m0 + i;
    const int n = n0 + threadIdx.x;
    tile[i][threadIdx.x] = (m < M && n < N) ? to_f(in[m * ld_in + n]) : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int n = n0 + i;
    const int64_t m = m0 + threadIdx.x;
    if (n < N && m < ldo) out[(int64_t)n * ldo + m] = from_f<TO>(tile[threadIdx.x][i]);
  }
}

template <typename T>
__global__ void im2col_T_kernel(const T* __restrict__ x, T* __restrict__ colT, int64_t ldm, int64_t M, int H, int W,
                                int C, int k, int stride, int pad, int dil, int OH, int OW) {
  __shared__ float tile[32][33];
  const int tap = blockIdx.z, r = tap / k, sf = tap - r * k;
  const int64_t m0 = (int64_t)blockIdx.x * 32;
  const int c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int64_t m = m0 + i;
    float v = 0.f;
    const int c = c0 + threadIdx.x;
    if (m < M && c < C) {
      const int ow = (int)(m % OW);
      const int64_t t = m / OW;
      const int oh = (int)(t % OH);
      const int64_t img = t / OH;
      const int ih = oh * stride - pad + r * dil, iw = ow * stride - pad + sf * dil;
      if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = to_f(x[((img * H + ih) * W + iw) * C + c]);
    }
    tile[i][threadIdx.x] = v;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i;
    const int64_t m = m0 + threadIdx.x;
    if (c < C && m < ldm) colT[((int64_t)tap * C + c) * ldm + m] = from_f<T>(tile[threadIdx.x][i]);
  }
}

// bf16 fast paths of the two layout helpers: 64x64 tiles, 4-byte (bf16x2) accesses in both directions so every
// warp instruction moves 128 contiguous bytes; smem pitch 66 elements keeps the transposed reads 2-way at worst.
__global__ void __launch_bounds__(256) transpose_bf16x2_kernel(const bf16* __restrict__ in, int64_t ld_in,
                                                               bf16* __restrict__ out, int64_t ldo, int64_t M, int N) {
  __shared__ __align__(4) bf16 t[64][66];
  const int64_t m0 = (int64_t)blockIdx.x * 64;
  const int n0 = blockIdx.y * 64;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int r = w; r < 64; r += 8) {
    const int64_t m = m0 + r;
    const int n = n0 + lane * 2;
    __nv_bfloat162 v = __floats2bfloat162_rn(0.f, 0.f);
    if (m < M && n < N) v = *reinterpret_cast<const __nv_bfloat162*>(in + m * ld_in + n);
    *reinterpret_cast<__nv_bfloat162*>(&t[r][lane * 2]) = v;
  }
  __syncthreads();
  for (int c = w; c < 64; c += 8) {
    const int n = n0 + c;
    const int64_t m = m0 + lane * 2;
    if (n < N && m < ldo) {
      __nv_bfloat162 v;
      v.x = t[lane * 2][c];
      v.y = t[lane * 2 + 1][c];
      *reinterpret_cast<__nv_bfloat162*>(out + (int64_t)n * ldo + m) = v;
    }
  }
}

__global__ void __launch_bounds__(256) im2col_T_bf16x2_kernel(const bf16* __restrict__ x, bf16* __restrict__ colT,
                                                              int64_t ldm, int64_t M, int H, int W, int C, int k,
                                                              int stride, int pad, int dil, int OH, int OW) {
  __shared__ __align__(4) bf16 t[64][66];
  const int tap = blockIdx.z, fr = tap / k, fs = tap - fr * k;
  const int64_t m0 = (int64_t)blockIdx.x * 64;
  const int c0 = blockIdx.y * 64;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int r = w; r < 64; r += 8) {
    const int64_t m = m0 + r;
    const int c = c0 + lane * 2;
    __nv_bfloat162 v = __floats2bfloat162_rn(0.f, 0.f);
    if (m < M && c < C) {
      const int ow = (int)(m % OW);
      const int64_t q = m / OW;
      const int oh = (int)(q % OH);
      const int64_t img = q / OH;
      const int ih = oh * stride - pad + fr * dil, iw = ow * stride - pad + fs * dil;
      if (ih >= 0 && ih < H && iw >= 0 && iw < W)
        v = *reinterpret_cast<const __nv_bfloat162*>(x + ((img * H + ih) * W + iw) * C + c);
    }
    *reinterpret_cast<__nv_bfloat162*>(&t[r][lane * 2]) = v;
  }
  __syncthreads();
  for (int cc = w; cc < 64; cc += 8) {
    const int c = c0 + cc;
    const int64_t m = m0 + lane * 2;
    if (c < C && m < ldm) {
      __nv_bfloat162 v;
      v.x = t[lane * 2][cc];
      v.y = t[lane * 2 + 1][cc];
      *reinterpret_cast<__nv_bfloat162*>(colT + ((int64_t)tap * C + c) * ldm + m) = v;
    }
  }
}

template <typename T>
__global__ void col2im_kernel(const T* __restrict__ dcol, T* __restrict__ dx, const T* __restrict__ add,
                              int64_t total, int H, int W, int C, int k, int stride, int pad, int dil, int OH, int OW) {
  using V = V16<T>;
  constexpr int VN = V::N;
  const int cv = C / VN;
  const int64_t K = (int64_t)k * k * C;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % cv) * VN;
    const int64_t pix = idx / cv;
    const int iw = (int)(pix % W);
    const int64_t t = pix / W;
    const int ih = (int)(t % H);
    const int64_t img = t / H;
    float acc[VN];
    if (add) V::load(add + pix * C + c, acc);
    else {
#pragma unroll
      for (int i = 0; i < VN; ++i) acc[i] = 0.f;
    }
    for (int r = 0; r < k; ++r) {
      const int th = ih + pad - r * dil;
      if (th < 0 || th % stride) continue;
      const int oh = th / stride;
      if (oh >= OH) continue;
      for (int sf = 0; sf < k; ++sf) {
        const int tw = iw + pad - sf * dil;
        if (tw < 0 || tw % stride) continue;
        const int ow = tw / stride;
        if (ow >= OW) continue;
        const int64_t m = (img * OH + oh) * OW + ow;
        float v[VN];
        V::load(dcol + m * K + (int64_t)(r * k + sf) * C + c, v);
#pragma unroll
        for (int i = 0; i < VN; ++i) acc[i] += v[i];
      }
    }
    V::store(dx + pix * C + c, acc);
  }
}

__global__ void filter_grad_unpack_kernel(const float* __restrict__ dwk, float* __restrict__ dw, int O, int C, int kh,
                                          int kw, int kpad) {
  const int64_t n = (int64_t)O * C * kh * kw;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (int64_t)gridDim.x * blockDim.x) {
    const int sf = (int)(idx % kw);
    int64_t t = idx / kw;
    const int r = (int)(t % kh); t /= kh;
    const int c = (int)(t % C);
    const int o = (int)(t / C);
    dw[idx] = dwk[(int64_t)o * kpad + (int64_t)(r * kw + sf) * C + c];
  }
}

__global__ void splitk_reduce_kernel(const float* __restrict__ part, float* __restrict__ out, int64_t n, int splits,
                                     int accumulate) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = accumulate ? out[i] : 0.f;
    for (int k = 0; k < splits; ++k) s += part[(int64_t)k * n + i];
    out[i] = s;
  }
}

// ---- elementwise backward --------------------------------------------------------------------------------
template <typename T>
__global__ void relu_bn_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ out, const float* __restrict__ scale,
                                   T* __restrict__ gs, T* __restrict__ gp, int64_t total, int C) {
  using V = V16<T>;
  constexpr int VN = V::N;
  const int cv = C / VN;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % cv) * VN;
    const int64_t off = idx * VN;
    float g[VN], o[VN];
    V::load(dy + off, g);
    if (out) {
      V::load(out + off, o);
#pragma unroll
      for (int i = 0; i < VN; ++i) g[i] = o[i] > 0.f ? g[i] : 0.f;
    }
    if (gp) V::store(gp + off, g);
    if (gs) {
      if (scale) {
#pragma unroll
        for (int i = 0; i < VN; ++i) g[i] *= __ldg(scale + c + i);
      }
      V::store(gs + off, g);
    }
  }
}

template <typename T>
__global__ void colsum_kernel(const T* __restrict__ dy, int64_t ld, float* __restrict__ out, int64_t M, int N,
                              int64_t rows_per_block) {
  __shared__ float red[8][33];
  const int n = blockIdx.x * 32 + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s = 0.f;
  if (n < N)
    for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) s += to_f(dy[r * ld + n]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(out + n, t);
  }
}

__global__ void dropout_add_kernel(const float* __restrict__ x, const float* __restrict__ br, float* __restrict__ out,
                                   int64_t n, DropCfg dc) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = x[i] + br[i] * drop_scale(dc, (uint64_t)i);
}
template <typename T>
__global__ void dropout_bwd_cast_kernel(const float* __restrict__ dx, T* __restrict__ g, int64_t n, DropCfg dc) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    g[i] = from_f<T>(dx[i] * drop_scale(dc, (uint64_t)i));
}
template <typename T>
__global__ void relu_bwd_kernel(T* __restrict__ g, const T* __restrict__ h, int64_t nvec) {
  using V = V16<T>;
  constexpr int VN = V::N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    float a[VN], b[VN];
    V::load(g + i * VN, a);
    V::load(h + i * VN, b);
#pragma unroll
    for (int j = 0; j < VN; ++j) a[j] = b[j] > 0.f ? a[j] : 0.f;
    V::store(g + i * VN, a);
  }
}

// ---- LayerNorm (warp per row, d a multiple of 32, d <= 1024) --------------------------------------------
constexpr int LN_MAX_PER = 32;       // d <= 1024

template <typename T, int MAXP>
__global__ void layernorm_pos_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                     const float* __restrict__ beta, T* __restrict__ y, T* __restrict__ y2,
                                     const float* __restrict__ pos, int period, int64_t rows, int d, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int per = d >> 5;
  float v[MAXP];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) { v[j] = x[row * d + j * 32 + lane]; s += v[j]; }
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) { const float t = v[j] - mean; q = fmaf(t, t, q); }
  const float rstd = rsqrtf(warp_sum(q) / d + eps);
  const float* pr = pos ? pos + (row % period) * d : nullptr;
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) {
      const int c = j * 32 + lane;
      const float o = (v[j] - mean) * rstd * gamma[c] + beta[c];
      if (y) y[row * d + c] = from_f<T>(o);
      if (y2) y2[row * d + c] = from_f<T>(o + (pr ? pr[c] : 0.f));
    }
}

template <int MAXP>
__global__ void layernorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                     const float* __restrict__ dy1, const float* __restrict__ dy2,
                                     float* __restrict__ dx, int accumulate_dx, float* __restrict__ dgamma,
                                     float* __restrict__ dbeta, int64_t rows, int d, float eps) {
  extern __shared__ float ln_red[];          // [2][d] block partial sums of dgamma / dbeta
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int per = d >> 5;
  for (int i = threadIdx.x; i < 2 * d; i += blockDim.x) ln_red[i] = 0.f;
  __syncthreads();
  float ag[MAXP], ab[MAXP];
#pragma unroll
  for (int j = 0; j < MAXP; ++j) ag[j] = ab[j] = 0.f;
  for (int64_t row = (int64_t)blockIdx.x * nwarp + warp; row < rows; row += (int64_t)gridDim.x * nwarp) {
    float v[MAXP], g[MAXP];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < MAXP; ++j)
      if (j < per) {
        const int64_t o = row * d + j * 32 + lane;
        v[j] = x[o];
        g[j] = dy1[o] + (dy2 ? dy2[o] : 0.f);
        s += v[j];
      }
    const float mean = warp_sum(s) / d;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < MAXP; ++j)
      if (j < per) { const float t = v[j] - mean; q = fmaf(t, t, q); }
    const float rstd = rsqrtf(warp_sum(q) / d + eps);
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int j = 0; j < MAXP; ++j)
      if (j < per) {
        const int c = j * 32 + lane;
        const float xh = (v[j] - mean) * rstd;
        ag[j] += g[j] * xh;
        ab[j] += g[j];
        g[j] *= gamma[c];
        v[j] = xh;
        c1 += g[j];
        c2 += g[j] * xh;
      }
    c1 = warp_sum(c1) / d;
    c2 = warp_sum(c2) / d;
#pragma unroll
    for (int j = 0; j < MAXP; ++j)
      if (j < per) {
        const int64_t o = row * d + j * 32 + lane;
        const float r = rstd * (g[j] - c1 - v[j] * c2);
        dx[o] = accumulate_dx ? dx[o] + r : r;
      }
  }
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) {
      atomicAdd(&ln_red[j * 32 + lane], ag[j]);
      atomicAdd(&ln_red[d + j * 32 + lane], ab[j]);
    }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    if (dgamma) atomicAdd(dgamma + i, ln_red[i]);
    if (dbeta) atomicAdd(dbeta + i, ln_red[d + i]);
  }
}

// d % 128 == 0: one float4 per lane per 128 columns (four times fewer load / store instructions per row, and the
// rows of a warp's loop are independent, so two rows are in flight per iteration)
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_bwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ dy1,
                         const float* __restrict__ dy2, float* __restrict__ dx, int accumulate_dx,
                         float* __restrict__ dgamma, float* __restrict__ dbeta, int64_t rows, int d, float eps) {
  extern __shared__ float ln_red[];          // [2][d] block partial sums of dgamma / dbeta
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int i = threadIdx.x; i < 2 * d; i += blockDim.x) ln_red[i] = 0.f;
  __syncthreads();
  float4 ga[NV], ag[NV], ab[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    ga[j] = __ldg(reinterpret_cast<const float4*>(gamma + j * 128 + lane * 4));
    ag[j] = ab[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float inv_d = 1.f / d;
  for (int64_t row = (int64_t)blockIdx.x * nwarp + warp; row < rows; row += (int64_t)gridDim.x * nwarp) {
    float4 v[NV], g[NV], old[NV];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int64_t o = row * d + j * 128 + lane * 4;
      v[j] = *reinterpret_cast<const float4*>(x + o);
      g[j] = *reinterpret_cast<const float4*>(dy1 + o);
      if (dy2) {
        const float4 t = *reinterpret_cast<const float4*>(dy2 + o);
        g[j].x += t.x; g[j].y += t.y; g[j].z += t.z; g[j].w += t.w;
      }
      if (accumulate_dx) old[j] = *reinterpret_cast<const float4*>(dx + o);
      s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
    }
    const float mean = warp_sum(s) * inv_d;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      v[j].x -= mean; v[j].y -= mean; v[j].z -= mean; v[j].w -= mean;
      q = fmaf(v[j].x, v[j].x, q); q = fmaf(v[j].y, v[j].y, q); q = fmaf(v[j].z, v[j].z, q); q = fmaf(v[j].w, v[j].w, q);
    }
    const float rstd = rsqrtf(warp_sum(q) * inv_d + eps);
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      v[j].x *= rstd; v[j].y *= rstd; v[j].z *= rstd; v[j].w *= rstd;                 // x_hat
      ag[j].x = fmaf(g[j].x, v[j].x, ag[j].x); ag[j].y = fmaf(g[j].y, v[j].y, ag[j].y);
      ag[j].z = fmaf(g[j].z, v[j].z, ag[j].z); ag[j].w = fmaf(g[j].w, v[j].w, ag[j].w);
      ab[j].x += g[j].x; ab[j].y += g[j].y; ab[j].z += g[j].z; ab[j].w += g[j].w;
      g[j].x *= ga[j].x; g[j].y *= ga[j].y; g[j].z *= ga[j].z; g[j].w *= ga[j].w;
      c1 += (g[j].x + g[j].y) + (g[j].z + g[j].w);
      c2 = fmaf(g[j].x, v[j].x, c2); c2 = fmaf(g[j].y, v[j].y, c2); c2 = fmaf(g[j].z, v[j].z, c2); c2 = fmaf(g[j].w, v[j].w, c2);
    }
    c1 = warp_sum(c1) * inv_d;
    c2 = warp_sum(c2) * inv_d;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float4 r;
      r.x = rstd * (g[j].x - c1 - v[j].x * c2); r.y = rstd * (g[j].y - c1 - v[j].y * c2);
      r.z = rstd * (g[j].z - c1 - v[j].z * c2); r.w = rstd * (g[j].w - c1 - v[j].w * c2);
      if (accumulate_dx) { r.x += old[j].x; r.y += old[j].y; r.z += old[j].z; r.w += old[j].w; }
      *reinterpret_cast<float4*>(dx + row * d + j * 128 + lane * 4) = r;
    }
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    float* a = &ln_red[j * 128 + lane * 4];
    atomicAdd(a, ag[j].x); atomicAdd(a + 1, ag[j].y); atomicAdd(a + 2, ag[j].z); atomicAdd(a + 3, ag[j].w);
    float* bb = &ln_red[d + j * 128 + lane * 4];
    atomicAdd(bb, ab[j].x); atomicAdd(bb + 1, ab[j].y); atomicAdd(bb + 2, ab[j].z); atomicAdd(bb + 3, ab[j].w);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    if (dgamma) atomicAdd(dgamma + i, ln_red[i]);
    if (dbeta) atomicAdd(dbeta + i, ln_red[d + i]);
  }
}

// ---- attention with stored probabilities ---------------------------------------------------------------------
constexpr int AT_WARPS = 8, AT_ROWS = 64;     // query rows per CTA

// K and V of one (b, h) as fp32 in shared memory, row pitch DH+1 (conflict-free both by key and by dim)
template <typename T, int DH>
__device__ __forceinline__ void load_kv_tile(const T* k, int64_t ldk, const T* v, int64_t ldv, int b, int h, int Lk,
                                             float* Ks, float* Vs) {
  for (int idx = threadIdx.x; idx < Lk * DH; idx += blockDim.x) {
    const int j = idx / DH, c = idx - j * DH;
    const int64_t row = (int64_t)b * Lk + j;
    Ks[j * (DH + 1) + c] = to_f(k[row * ldk + h * DH + c]);
    Vs[j * (DH + 1) + c] = to_f(v[row * ldv + h * DH + c]);
  }
}

template <typename T, int DH>
__global__ void __launch_bounds__(AT_WARPS * 32) attn_train_fwd_kernel(const AttnTrainParams p) {
  extern __shared__ float at_smem[];
  float* Ks = at_smem;
  float* Vs = Ks + (size_t)p.Lk * (DH + 1);
  float* prow = Vs + (size_t)p.Lk * (DH + 1);            // [AT_WARPS][Lk]
  float* qrow = prow + (size_t)AT_WARPS * p.Lk;            // [AT_WARPS][DH]
  const int b = blockIdx.z, h = blockIdx.y, i0 = blockIdx.x * AT_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* q = static_cast<const T*>(p.q);
  load_kv_tile<T, DH>(static_cast<const T*>(p.k), p.ldk, static_cast<const T*>(p.v), p.ldv, b, h, p.Lk, Ks, Vs);
  __syncthreads();
  const float scale = rsqrtf((float)DH);
  float* pr = prow + (size_t)warp * p.Lk;
  float* qr = qrow + warp * DH;
  T* probs = static_cast<T*>(p.probs);
  T* out = static_cast<T*>(p.out);
  const uint8_t* pad = p.key_pad ? p.key_pad + (int64_t)b * p.Lk : nullptr;
  for (int i = i0 + warp; i < min(p.Lq, i0 + AT_ROWS); i += AT_WARPS) {
    const int64_t qoff = ((int64_t)b * p.Lq + i) * p.ldq + h * DH;
    for (int c = lane; c < DH; c += 32) qr[c] = to_f(q[qoff + c]) * scale;
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j < p.Lk; j += 32) {
      float s = 0.f;
#pragma unroll
      for (int c = 0; c < DH; ++c) s = fmaf(qr[c], Ks[j * (DH + 1) + c], s);
      if ((pad && pad[j]) || (p.causal && j > i)) s = -INFINITY;
      pr[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < p.Lk; j += 32) {
      const float e = __expf(pr[j] - mx);
      pr[j] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    const int64_t prow_off = (((int64_t)b * p.H + h) * p.Lq + i) * p.Lk;
    for (int j = lane; j < p.Lk; j += 32) {
      const float pv = pr[j] * inv;
      if (probs) probs[prow_off + j] = from_f<T>(pv);
      // the stored (rounded) probability is what the backward pass sees: use it here too
      pr[j] = to_f(from_f<T>(pv)) * drop_scale_rc(p.drop, (uint64_t)(prow_off / p.Lk), (uint32_t)j);
    }
    __syncwarp();
    for (int c = lane; c < DH; c += 32) {
      float acc = 0.f;
      for (int j = 0; j < p.Lk; ++j) acc = fmaf(pr[j], Vs[j * (DH + 1) + c], acc);
      out[((int64_t)b * p.Lq + i) * p.ldo + h * DH + c] = from_f<T>(acc);
    }
    __syncwarp();
  }
}

// per query row: dS = P ∘ (dP − Σ dP∘P) (written out), dQ = scale · dS · K
template <typename T, int DH>
__global__ void __launch_bounds__(AT_WARPS * 32) attn_train_bwd_q_kernel(const AttnBwdParams bp) {
  const AttnTrainParams& p = bp.f;
  extern __shared__ float at_smem[];
  float* Ks = at_smem;
  float* Vs = Ks + (size_t)p.Lk * (DH + 1);
  float* prow = Vs + (size_t)p.Lk * (DH + 1);
  float* qrow = prow + (size_t)AT_WARPS * p.Lk;            // holds dO row here
  const int b = blockIdx.z, h = blockIdx.y, i0 = blockIdx.x * AT_ROWS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  load_kv_tile<T, DH>(static_cast<const T*>(p.k), p.ldk, static_cast<const T*>(p.v), p.ldv, b, h, p.Lk, Ks, Vs);
  __syncthreads();
  const float scale = rsqrtf((float)DH);
  float* pr = prow + (size_t)warp * p.Lk;
  float* dor = qrow + warp * DH;
  const T* probs = static_cast<const T*>(p.probs);
  const T* dout = static_cast<const T*>(bp.dout);
  T* ds = static_cast<T*>(bp.ds);
  T* dq = static_cast<T*>(bp.dq);
  for (int i = i0 + warp; i < min(p.Lq, i0 + AT_ROWS); i += AT_WARPS) {
    const int64_t row = (int64_t)b * p.Lq + i;
    for (int c = lane; c < DH; c += 32) dor[c] = to_f(dout[row * bp.lddo + h * DH + c]);
    __syncwarp();
    const int64_t prow_off = (((int64_t)b * p.H + h) * p.Lq + i) * p.Lk;
    float delta = 0.f;
    for (int j = lane; j < p.Lk; j += 32) {
      float dp = 0.f;
#pragma unroll
      for (int c = 0; c < DH; ++c) dp = fmaf(dor[c], Vs[j * (DH + 1) + c], dp);
      dp *= drop_scale_rc(p.drop, (uint64_t)(prow_off / p.Lk), (uint32_t)j);
      const float pv = to_f(probs[prow_off + j]);
      pr[j] = dp;
      delta = fmaf(dp, pv, delta);
    }
    delta = warp_sum(delta);
    for (int j = lane; j < p.Lk; j += 32) {
      const float pv = to_f(probs[prow_off + j]);
      const T dsv = from_f<T>(pv * (pr[j] - delta));
      ds[prow_off + j] = dsv;
      pr[j] = to_f(dsv);
    }
    __syncwarp();
    for (int c = lane; c < DH; c += 32) {
      float acc = 0.f;
      for (int j = 0; j < p.Lk; ++j) acc = fmaf(pr[j], Ks[j * (DH + 1) + c], acc);
      dq[row * bp.lddq + h * DH + c] = from_f<T>(acc * scale);
    }
    __syncwarp();
  }
}

// per tile of 32 keys: dV = (P∘mask)^T · dO, dK = scale · dS^T · Q
template <typename T, int DH>
__global__ void __launch_bounds__(128) attn_train_bwd_kv_kernel(const AttnBwdParams bp) {
  const AttnTrainParams& p = bp.f;
  constexpr int PER = DH / 32;
  __shared__ float Ps[32][33], Ss[32][33], dOs[32][DH + 1], Qs[32][DH + 1];
  const int b = blockIdx.z, h = blockIdx.y, j0 = blockIdx.x * 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* probs = static_cast<const T*>(p.probs);
  const T* ds = static_cast<const T*>(bp.ds);
  const T* dout = static_cast<const T*>(bp.dout);
  const T* q = static_cast<const T*>(p.q);
  float adv[8][PER], adk[8][PER];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int e = 0; e < PER; ++e) adv[a][e] = adk[a][e] = 0.f;
  const int i_begin = p.causal ? (j0 / 32) * 32 : 0;       // rows i < j have P = 0 under the causal mask
  for (int ib = i_begin; ib < p.Lq; ib += 32) {
    __syncthreads();
    for (int idx = threadIdx.x; idx < 32 * 32; idx += 128) {
      const int ii = idx >> 5, jj = idx & 31;
      const int i = ib + ii, j = j0 + jj;
      float pv = 0.f, sv = 0.f;
      if (i < p.Lq && j < p.Lk) {
        const int64_t o = (((int64_t)b * p.H + h) * p.Lq + i) * p.Lk + j;
        pv = to_f(probs[o]) * drop_scale_rc(p.drop, (uint64_t)(o / p.Lk), (uint32_t)j);
        sv = to_f(ds[o]);
      }
      Ps[ii][jj] = pv;
      Ss[ii][jj] = sv;
    }
    for (int idx = threadIdx.x; idx < 32 * DH; idx += 128) {
      const int ii = idx / DH, c = idx - ii * DH;
      const int i = ib + ii;
      float a = 0.f, bq = 0.f;
      if (i < p.Lq) {
        const int64_t row = (int64_t)b * p.Lq + i;
        a = to_f(dout[row * bp.lddo + h * DH + c]);
        bq = to_f(q[row * p.ldq + h * DH + c]);
      }
      dOs[ii][c] = a;
      Qs[ii][c] = bq;
    }
    __syncthreads();
#pragma unroll 4
    for (int ii = 0; ii < 32; ++ii) {
      float dov[PER], qv[PER];
#pragma unroll
      for (int e = 0; e < PER; ++e) { dov[e] = dOs[ii][lane + 32 * e]; qv[e] = Qs[ii][lane + 32 * e]; }
#pragma unroll
      for (int a = 0; a < 8; ++a) {
        const float pv = Ps[ii][warp * 8 + a], sv = Ss[ii][warp * 8 + a];
#pragma unroll
        for (int e = 0; e < PER; ++e) {
          adv[a][e] = fmaf(pv, dov[e], adv[a][e]);
          adk[a][e] = fmaf(sv, qv[e], adk[a][e]);
        }
      }
    }
  }
  const float scale = rsqrtf((float)DH);
  T* dk = static_cast<T*>(bp.dk);
  T* dv = static_cast<T*>(bp.dv);
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int j = j0 + warp * 8 + a;
    if (j >= p.Lk) continue;
    const int64_t row = (int64_t)b * p.Lk + j;
#pragma unroll
    for (int e = 0; e < PER; ++e) {
      dv[row * bp.lddv + h * DH + lane + 32 * e] = from_f<T>(adv[a][e]);
      dk[row * bp.lddk + h * DH + lane + 32 * e] = from_f<T>(adk[a][e] * scale);
    }
  }
}

template <int DH> size_t attn_smem(int Lk) {
  return ((size_t)2 * Lk * (DH + 1) + (size_t)AT_WARPS * Lk + AT_WARPS * DH) * sizeof(float);
}

// ---- embeddings ---------------------------------------------------------------------------------------------
template <int MAXP>
__global__ void embed_ln_train_kernel(const int64_t* __restrict__ ids, const float* __restrict__ e_word,
                                      const float* __restrict__ e_pos, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, float* __restrict__ pre, float* __restrict__ y,
                                      int64_t rows, int Tn, int d, float eps, DropCfg dc) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int per = d >> 5;
  const int t = (int)(row % Tn);
  const int64_t id = ids[row];
  float v[MAXP];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) {
      const int c = j * 32 + lane;
      v[j] = e_word[id * d + c] + e_pos[(int64_t)t * d + c];
      pre[row * d + c] = v[j];
      s += v[j];
    }
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) { const float u = v[j] - mean; q = fmaf(u, u, q); }
  const float rstd = rsqrtf(warp_sum(q) / d + eps);
#pragma unroll
  for (int j = 0; j < MAXP; ++j)
    if (j < per) {
      const int c = j * 32 + lane;
      y[row * d + c] = ((v[j] - mean) * rstd * gamma[c] + beta[c]) * drop_scale(dc, (uint64_t)(row * d + c));
    }
}

__global__ void embed_scatter_kernel(const int64_t* __restrict__ ids, const float* __restrict__ g,
                                     float* __restrict__ d_word, float* __restrict__ d_pos, int64_t rows, int Tn, int d, int pad_id) {
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < rows * d; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = idx / d;
    const int c = (int)(idx - row * d);
    const float v = g[idx];
    if (ids[row] != pad_id) atomicAdd(d_word + ids[row] * d + c, v);       // padding_idx row gets no gradient
    atomicAdd(d_pos + (row % Tn) * d + c, v);
  }
}

__global__ void pos_grad_kernel(const float* __restrict__ g, float* __restrict__ d_pos, int B, int Tn, int d) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Tn * d) return;
  float s = 0.f;
  for (int b = 0; b < B; ++b) s += g[(int64_t)b * Tn * d + idx];
  d_pos[idx] += s;
}

__global__ void loc_tokens_bwd_kernel(const float* __restrict__ dx, const float* __restrict__ loc, float* __restrict__ dw,
                                      float* __restrict__ db, int B, int n_loc, int per_scalar, int d, int S, int off) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d) return;
  if (per_scalar) {
    float sw = 0.f, sb = 0.f;
    for (int b = 0; b < B; ++b)
      for (int k = 0; k < n_loc; ++k) {
        const float g = dx[((int64_t)b * S + off + k) * d + c];
        sw = fmaf(g, loc[b * n_loc + k], sw);
        sb += g;
      }
    dw[c] += sw;
    db[c] += sb;
  } else {
    float sb = 0.f;
    for (int i = 0; i < n_loc; ++i) {
      float sw = 0.f;
      for (int b = 0; b < B; ++b) sw = fmaf(dx[((int64_t)b * S + off) * d + c], loc[b * n_loc + i], sw);
      dw[c * n_loc + i] += sw;
    }
    for (int b = 0; b < B; ++b) sb += dx[((int64_t)b * S + off) * d + c];
    db[c] += sb;
  }
}

// ---- loss -----------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) ce_kernel(T* __restrict__ logits, int64_t ld, const int64_t* __restrict__ targets,
                                                 int V, float grad_scale, float* __restrict__ loss_sum) {
  __shared__ float red[8];
  __shared__ float bc;
  T* row = logits + (int64_t)blockIdx.x * ld;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float mx = -INFINITY;
  for (int j = threadIdx.x; j < V; j += 256) mx = fmaxf(mx, to_f(row[j]));
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  if (threadIdx.x == 0) { float m = red[0]; for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]); bc = m; }
  __syncthreads();
  mx = bc;
  float s = 0.f;
  for (int j = threadIdx.x; j < V; j += 256) s += __expf(to_f(row[j]) - mx);
  s = warp_sum(s);
  __syncthreads();
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) { float t = 0.f; for (int i = 0; i < 8; ++i) t += red[i]; bc = t; }
  __syncthreads();
  const float sum = bc, inv = 1.f / sum;
  const int tgt = (int)targets[blockIdx.x];
  if (threadIdx.x == 0) atomicAdd(loss_sum, logf(sum) + mx - to_f(row[tgt]));
  __syncthreads();
  for (int j = threadIdx.x; j < V; j += 256) {
    const float pj = __expf(to_f(row[j]) - mx) * inv;
    row[j] = from_f<T>((pj - (j == tgt ? 1.f : 0.f)) * grad_scale);
  }
}

// ---- optimiser ------------------------------------------------------------------------------------------------
__global__ void grad_sumsq_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ out) {
  __shared__ float red[8];
  float s = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    s = fmaf(g[i], g[i], s);
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
    atomicAdd(out, t);
  }
}

__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, int64_t n, const float* __restrict__ sumsq, float max_norm, float lr,
                             float beta1, float beta2, float eps, float wd, float bc1, float bc2_sqrt) {
  float coef = 1.f;
  if (max_norm > 0.f) coef = fminf(1.f, max_norm / (sqrtf(*sumsq) + 1e-6f));
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * coef;
    float pi = p[i] * (1.f - lr * wd);
    const float mi = beta1 * m[i] + (1.f - beta1) * gi;
    const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    pi -= (lr / bc1) * mi / (sqrtf(vi) / bc2_sqrt + eps);
    p[i] = pi;
  }
}

}  // namespace

// ================================== host wrappers ===================================================
template <typename TI, typename TO>
void transpose_pad(const TI* in, int64_t ld_in, TO* out, int64_t ldo, int64_t M, int N, cudaStream_t s) {
  if (std::is_same<TI, bf16>::value && std::is_same<TO, bf16>::value && N % 2 == 0 && ld_in % 2 == 0 && ldo % 2 == 0 &&
      reinterpret_cast<uintptr_t>(in) % 4 == 0 && reinterpret_cast<uintptr_t>(out) % 4 == 0) {
    dim3 g2((unsigned)cdiv(ldo, 64), (unsigned)cdiv(N, 64));
    transpose_bf16x2_kernel<<<g2, 256, 0, s>>>(reinterpret_cast<const bf16*>(in), ld_in, reinterpret_cast<bf16*>(out), ldo, M, N);
    LAUNCH_OK();
    return;
  }
  dim3 grid((unsigned)cdiv(ldo, 32), (unsigned)cdiv(N, 32));
  transpose_pad_kernel<TI, TO><<<grid, dim3(32, 8), 0, s>>>(in, ld_in, out, ldo, M, N);
  LAUNCH_OK();
}
template void transpose_pad<float, float>(const float*, int64_t, float*, int64_t, int64_t, int, cudaStream_t);
template void transpose_pad<bf16, bf16>(const bf16*, int64_t, bf16*, int64_t, int64_t, int, cudaStream_t);
template void transpose_pad<float, bf16>(const float*, int64_t, bf16*, int64_t, int64_t, int, cudaStream_t);

template <typename T>
void im2col_T(const T* x, T* colT, int64_t ldm, int64_t nimg, int H, int W, int C, int k, int stride, int pad, int dil,
              int OH, int OW, cudaStream_t s) {
  const int64_t M = nimg * OH * OW;
  if (std::is_same<T, bf16>::value && C % 2 == 0 && ldm % 2 == 0) {
    dim3 g2((unsigned)cdiv(ldm, 64), (unsigned)cdiv(C, 64), (unsigned)(k * k));
    im2col_T_bf16x2_kernel<<<g2, 256, 0, s>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(colT), ldm, M, H, W, C, k,
                                              stride, pad, dil, OH, OW);
    LAUNCH_OK();
    return;
  }
  dim3 grid((unsigned)cdiv(ldm, 32), (unsigned)cdiv(C, 32), (unsigned)(k * k));
  im2col_T_kernel<T><<<grid, dim3(32, 8), 0, s>>>(x, colT, ldm, M, H, W, C, k, stride, pad, dil, OH, OW);
  LAUNCH_OK();
}
template void im2col_T<float>(const float*, float*, int64_t, int64_t, int, int, int, int, int, int, int, int, int, cudaStream_t);
template void im2col_T<bf16>(const bf16*, bf16*, int64_t, int64_t, int, int, int, int, int, int, int, int, int, cudaStream_t);

template <typename T>
void col2im(const T* dcol, T* dx, int64_t nimg, int H, int W, int C, int k, int stride, int pad, int dil, int OH, int OW,
            const T* add, cudaStream_t s) {
  RETR_CHECK(C % V16<T>::N == 0, "col2im: channel count must fill 16-byte vectors");
  const int64_t total = nimg * H * W * (C / V16<T>::N);
  col2im_kernel<T><<<ew_blocks(total), EW_THREADS, 0, s>>>(dcol, dx, add, total, H, W, C, k, stride, pad, dil, OH, OW);
  LAUNCH_OK();
}
template void col2im<float>(const float*, float*, int64_t, int, int, int, int, int, int, int, int, int, const float*, cudaStream_t);
template void col2im<bf16>(const bf16*, bf16*, int64_t, int, int, int, int, int, int, int, int, int, const bf16*, cudaStream_t);

template <typename T>
__global__ void repack_filter_dgrad_kernel(const float* __restrict__ w, T* __restrict__ out, int O, int I, int kh, int kw) {
  const int64_t kk = (int64_t)kh * kw * O, total = (int64_t)I * kk;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx / kk), k = (int)(idx % kk);
    const int tap = k / O, o = k % O, r = kh - 1 - tap / kw, s = kw - 1 - tap % kw;
    out[idx] = from_f<T>(w[(((int64_t)o * I + c) * kh + r) * kw + s]);
  }
}
template <typename T>
void repack_filter_dgrad(const float* oihw, T* out, int O, int I, int kh, int kw, cudaStream_t s) {
  repack_filter_dgrad_kernel<T><<<ew_blocks((int64_t)O * I * kh * kw), EW_THREADS, 0, s>>>(oihw, out, O, I, kh, kw);
  LAUNCH_OK();
}
template void repack_filter_dgrad<float>(const float*, float*, int, int, int, int, cudaStream_t);
template void repack_filter_dgrad<bf16>(const float*, bf16*, int, int, int, int, cudaStream_t);

void filter_grad_unpack(const float* dwk, float* dw, int O, int C, int kh, int kw, int kpad, cudaStream_t s) {
  filter_grad_unpack_kernel<<<ew_blocks((int64_t)O * C * kh * kw), EW_THREADS, 0, s>>>(dwk, dw, O, C, kh, kw, kpad);
  LAUNCH_OK();
}

void splitk_reduce(const float* part, float* out, int64_t n, int splits, int accumulate, cudaStream_t s) {
  splitk_reduce_kernel<<<ew_blocks(n), EW_THREADS, 0, s>>>(part, out, n, splits, accumulate);
  LAUNCH_OK();
}

template <typename T>
void relu_bn_bwd(const T* dy, const T* out, const float* scale, T* g_scaled, T* g_plain, int64_t M, int C,
                 cudaStream_t s) {
  RETR_CHECK(C % V16<T>::N == 0, "relu_bn_bwd: channel count must fill 16-byte vectors");
  const int64_t total = M * (C / V16<T>::N);
  relu_bn_bwd_kernel<T><<<ew_blocks(total), EW_THREADS, 0, s>>>(dy, out, scale, g_scaled, g_plain, total, C);
  LAUNCH_OK();
}
template void relu_bn_bwd<float>(const float*, const float*, const float*, float*, float*, int64_t, int, cudaStream_t);
template void relu_bn_bwd<bf16>(const bf16*, const bf16*, const float*, bf16*, bf16*, int64_t, int, cudaStream_t);

template <typename T>
void colsum(const T* dy, int64_t ld, float* out, int64_t M, int N, int accumulate, cudaStream_t s) {
  if (!accumulate) CUDA_OK(cudaMemsetAsync(out, 0, (size_t)N * 4, s));
  const int64_t chunks = std::min<int64_t>(std::max<int64_t>(cdiv(M, 256), 1), 512);
  const int64_t rpb = cdiv(M, chunks);
  dim3 grid((unsigned)cdiv(N, 32), (unsigned)cdiv(M, rpb));
  colsum_kernel<T><<<grid, dim3(32, 8), 0, s>>>(dy, ld, out, M, N, rpb);
  LAUNCH_OK();
}
template void colsum<float>(const float*, int64_t, float*, int64_t, int, int, cudaStream_t);
template void colsum<bf16>(const bf16*, int64_t, float*, int64_t, int, int, cudaStream_t);

void dropout_add(const float* x, const float* branch, float* out, int64_t n, DropCfg dc, cudaStream_t s) {
  dropout_add_kernel<<<ew_blocks(n), EW_THREADS, 0, s>>>(x, branch, out, n, dc);
  LAUNCH_OK();
}
template <typename T>
void dropout_bwd_cast(const float* dx, T* g, int64_t n, DropCfg dc, cudaStream_t s) {
  dropout_bwd_cast_kernel<T><<<ew_blocks(n), EW_THREADS, 0, s>>>(dx, g, n, dc);
  LAUNCH_OK();
}
template void dropout_bwd_cast<float>(const float*, float*, int64_t, DropCfg, cudaStream_t);
template void dropout_bwd_cast<bf16>(const float*, bf16*, int64_t, DropCfg, cudaStream_t);

template <typename T>
void relu_bwd(T* g, const T* h, int64_t n, cudaStream_t s) {
  RETR_CHECK(n % V16<T>::N == 0, "relu_bwd: size must fill 16-byte vectors");
  relu_bwd_kernel<T><<<ew_blocks(n / V16<T>::N), EW_THREADS, 0, s>>>(g, h, n / V16<T>::N);
  LAUNCH_OK();
}
template void relu_bwd<float>(float*, const float*, int64_t, cudaStream_t);
template void relu_bwd<bf16>(bf16*, const bf16*, int64_t, cudaStream_t);

template <typename T>
void layernorm_pos(const float* x, const float* gamma, const float* beta, T* y, T* y2, const float* pos, int period,
                   int64_t rows, int d, float eps, cudaStream_t s) {
  RETR_CHECK(d % 32 == 0 && d <= 32 * LN_MAX_PER, "layernorm: d must be a multiple of 32, at most 1024");
  const unsigned grid = (unsigned)cdiv(rows, 8);
  const int per = period > 0 ? period : 1;
  if (d <= 256) layernorm_pos_kernel<T, 8><<<grid, 256, 0, s>>>(x, gamma, beta, y, y2, pos, per, rows, d, eps);
  else if (d <= 512) layernorm_pos_kernel<T, 16><<<grid, 256, 0, s>>>(x, gamma, beta, y, y2, pos, per, rows, d, eps);
  else layernorm_pos_kernel<T, 32><<<grid, 256, 0, s>>>(x, gamma, beta, y, y2, pos, per, rows, d, eps);
  LAUNCH_OK();
}
template void layernorm_pos<float>(const float*, const float*, const float*, float*, float*, const float*, int, int64_t,
                                   int, float, cudaStream_t);
template void layernorm_pos<bf16>(const float*, const float*, const float*, bf16*, bf16*, const float*, int, int64_t, int,
                                  float, cudaStream_t);

void layernorm_bwd(const float* x, const float* gamma, const float* dy1, const float* dy2, float* dx, int accumulate_dx,
                   float* dgamma, float* dbeta, int64_t rows, int d, float eps, cudaStream_t s) {
  RETR_CHECK(d % 32 == 0 && d <= 32 * LN_MAX_PER, "layernorm_bwd: d must be a multiple of 32, at most 1024");
  const size_t sm = (size_t)2 * d * 4;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy1) | reinterpret_cast<uintptr_t>(dy2) |
                         reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
  if (aligned && (d == 256 || d == 512) && getenv("RETR_LN_BWD_SCALAR") == nullptr) {
    const unsigned vb = (unsigned)std::min<int64_t>(cdiv(rows, 8), 148 * 8);
    if (d == 256) layernorm_bwd_vec_kernel<2><<<vb, 256, sm, s>>>(x, gamma, dy1, dy2, dx, accumulate_dx, dgamma, dbeta, rows, d, eps);
    else layernorm_bwd_vec_kernel<4><<<vb, 256, sm, s>>>(x, gamma, dy1, dy2, dx, accumulate_dx, dgamma, dbeta, rows, d, eps);
    LAUNCH_OK();
    return;
  }
  const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(rows, 8), 148 * 4);
  if (d <= 256) layernorm_bwd_kernel<8><<<blocks, 256, sm, s>>>(x, gamma, dy1, dy2, dx, accumulate_dx, dgamma, dbeta, rows, d, eps);
  else if (d <= 512) layernorm_bwd_kernel<16><<<blocks, 256, sm, s>>>(x, gamma, dy1, dy2, dx, accumulate_dx, dgamma, dbeta, rows, d, eps);
  else layernorm_bwd_kernel<32><<<blocks, 256, sm, s>>>(x, gamma, dy1, dy2, dx, accumulate_dx, dgamma, dbeta, rows, d, eps);
  LAUNCH_OK();
}

template <typename T>
void attention_train_fwd(const AttnTrainParams& p, cudaStream_t s) {
  if (std::is_same<T, bf16>::value && p.lse) return flash_train_fwd_bf16(p, s);      // tensor-core path
  RETR_CHECK(p.dh == 32 || p.dh == 64, "head dim must be 32 or 64");
  dim3 grid((unsigned)cdiv(p.Lq, AT_ROWS), (unsigned)p.H, (unsigned)p.B);
  if (p.dh == 32) {
    const size_t smem = attn_smem<32>(p.Lk);
    CUDA_OK(cudaFuncSetAttribute(attn_train_fwd_kernel<T, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_train_fwd_kernel<T, 32><<<grid, AT_WARPS * 32, smem, s>>>(p);
  } else {
    const size_t smem = attn_smem<64>(p.Lk);
    RETR_CHECK(smem <= 227 * 1024, "attention: K/V tile does not fit in shared memory");
    CUDA_OK(cudaFuncSetAttribute(attn_train_fwd_kernel<T, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_train_fwd_kernel<T, 64><<<grid, AT_WARPS * 32, smem, s>>>(p);
  }
  LAUNCH_OK();
}
template void attention_train_fwd<float>(const AttnTrainParams&, cudaStream_t);
template void attention_train_fwd<bf16>(const AttnTrainParams&, cudaStream_t);

template <typename T>
void attention_train_bwd(const AttnBwdParams& bp, cudaStream_t s) {
  const AttnTrainParams& p = bp.f;
  if (std::is_same<T, bf16>::value && p.lse) return flash_train_bwd_bf16(bp, s);       // tensor-core path
  RETR_CHECK(p.dh == 32 || p.dh == 64, "head dim must be 32 or 64");
  dim3 gq((unsigned)cdiv(p.Lq, AT_ROWS), (unsigned)p.H, (unsigned)p.B);
  dim3 gk((unsigned)cdiv(p.Lk, 32), (unsigned)p.H, (unsigned)p.B);
  if (p.dh == 32) {
    const size_t smem = attn_smem<32>(p.Lk);
    CUDA_OK(cudaFuncSetAttribute(attn_train_bwd_q_kernel<T, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_train_bwd_q_kernel<T, 32><<<gq, AT_WARPS * 32, smem, s>>>(bp);
    LAUNCH_OK();
    attn_train_bwd_kv_kernel<T, 32><<<gk, 128, 0, s>>>(bp);
  } else {
    const size_t smem = attn_smem<64>(p.Lk);
    RETR_CHECK(smem <= 227 * 1024, "attention: K/V tile does not fit in shared memory");
    CUDA_OK(cudaFuncSetAttribute(attn_train_bwd_q_kernel<T, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attn_train_bwd_q_kernel<T, 64><<<gq, AT_WARPS * 32, smem, s>>>(bp);
    LAUNCH_OK();
    attn_train_bwd_kv_kernel<T, 64><<<gk, 128, 0, s>>>(bp);
  }
  LAUNCH_OK();
}
template void attention_train_bwd<float>(const AttnBwdParams&, cudaStream_t);
template void attention_train_bwd<bf16>(const AttnBwdParams&, cudaStream_t);

void embed_ln_train(const int64_t* ids, const float* e_word, const float* e_pos, const float* gamma, const float* beta,
                    float* pre, float* y, int B, int Tn, int d, float eps, DropCfg dc, cudaStream_t s) {
  RETR_CHECK(d % 32 == 0 && d <= 32 * LN_MAX_PER, "embed_ln: d must be a multiple of 32, at most 1024");
  const int64_t rows = (int64_t)B * Tn;
  const unsigned grid = (unsigned)cdiv(rows, 8);
  if (d <= 256) embed_ln_train_kernel<8><<<grid, 256, 0, s>>>(ids, e_word, e_pos, gamma, beta, pre, y, rows, Tn, d, eps, dc);
  else if (d <= 512) embed_ln_train_kernel<16><<<grid, 256, 0, s>>>(ids, e_word, e_pos, gamma, beta, pre, y, rows, Tn, d, eps, dc);
  else embed_ln_train_kernel<32><<<grid, 256, 0, s>>>(ids, e_word, e_pos, gamma, beta, pre, y, rows, Tn, d, eps, dc);
  LAUNCH_OK();
}

void embed_ln_bwd(const int64_t* ids, const float* pre, const float* gamma, const float* dy, float* d_word, float* d_pos,
                  float* dgamma, float* dbeta, float* tmp, int B, int Tn, int d, float eps, DropCfg dc, int pad_id, cudaStream_t s);

void pos_grad_accumulate(const float* g, float* d_pos, int B, int Tn, int d, cudaStream_t s) {
  pos_grad_kernel<<<(unsigned)cdiv((int64_t)Tn * d, 256), 256, 0, s>>>(g, d_pos, B, Tn, d);
  LAUNCH_OK();
}

void loc_tokens_bwd(const float* dx, const float* loc, float* dw, float* db, int B, int n_loc, int per_scalar, int d,
                    int S, int off, cudaStream_t s) {
  loc_tokens_bwd_kernel<<<(unsigned)cdiv(d, 128), 128, 0, s>>>(dx, loc, dw, db, B, n_loc, per_scalar, d, S, off);
  LAUNCH_OK();
}

template <typename T>
void ce_loss_fwd_bwd(T* logits, int64_t ld, const int64_t* targets, int64_t rows, int V, float grad_scale,
                     float* loss_sum, cudaStream_t s) {
  ce_kernel<T><<<(unsigned)rows, 256, 0, s>>>(logits, ld, targets, V, grad_scale, loss_sum);
  LAUNCH_OK();
}
template void ce_loss_fwd_bwd<float>(float*, int64_t, const int64_t*, int64_t, int, float, float*, cudaStream_t);
template void ce_loss_fwd_bwd<bf16>(bf16*, int64_t, const int64_t*, int64_t, int, float, float*, cudaStream_t);

void grad_sumsq(const float* g, int64_t n, float* sumsq, cudaStream_t s) {
  grad_sumsq_kernel<<<ew_blocks(n), EW_THREADS, 0, s>>>(g, n, sumsq);
  LAUNCH_OK();
}

void adamw_step(float* p, const float* g, float* m, float* v, int64_t n, const float* sumsq, float max_norm, float lr,
                float beta1, float beta2, float eps, float weight_decay, int step, cudaStream_t s) {
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
  adamw_kernel<<<ew_blocks(n), EW_THREADS, 0, s>>>(p, g, m, v, n, sumsq, max_norm, lr, beta1, beta2, eps, weight_decay,
                                                   bc1, sqrtf(bc2));
  LAUNCH_OK();
}

// embedding backward: dropout mask -> LayerNorm backward (statistics from the saved pre-norm sum) -> scatter
void embed_ln_bwd(const int64_t* ids, const float* pre, const float* gamma, const float* dy, float* d_word, float* d_pos,
                  float* dgamma, float* dbeta, float* tmp, int B, int Tn, int d, float eps, DropCfg dc, int pad_id, cudaStream_t s) {
  const int64_t rows = (int64_t)B * Tn;
  float* t1 = tmp;
  float* t2 = tmp + rows * d;
  dropout_bwd_cast<float>(dy, t1, rows * d, dc, s);
  layernorm_bwd(pre, gamma, t1, nullptr, t2, 0, dgamma, dbeta, rows, d, eps, s);
  embed_scatter_kernel<<<ew_blocks(rows * d), EW_THREADS, 0, s>>>(ids, t2, d_word, d_pos, rows, Tn, d, pad_id);
  LAUNCH_OK();
}

}  // namespace retr
