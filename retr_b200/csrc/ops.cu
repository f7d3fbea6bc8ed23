// HBM-bound helper kernels: weight repack, stem im2col, max-pool, masks, location tokens,
// LayerNorm and decoder embeddings.  All are pure streaming kernels (no reuse), vectorised to
// 16-byte accesses where the layout allows; grids are sized from the element count.
#include "ops.h"

#include <algorithm>

namespace retr {
namespace {

inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)cdiv(n, threads); }

__global__ void f32_to_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = __float2bfloat16_rn(src[i]);
}

template <typename T>
__global__ void repack_filter_kernel(const float* __restrict__ w, T* __restrict__ out, int O, int I, int kh,
                                     int kw, int kpad) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)O * kpad) return;
  const int o = (int)(idx / kpad), k = (int)(idx % kpad);
  float v = 0.f;
  if (k < kh * kw * I) {
    const int tap = k / I, c = k % I, r = tap / kw, s = tap % kw;
    v = w[(((int64_t)o * I + c) * kh + r) * kw + s];
  }
  out[idx] = from_f<T>(v);
}

// Each thread produces 8 consecutive k of one output pixel.
template <typename T>
__global__ void stem_im2col_kernel(const float* __restrict__ img, T* __restrict__ out, int64_t rows, int H, int W,
                                   int OH, int OW, int kpad) {
  const int kchunks = kpad / 8;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * kchunks) return;
  const int64_t m = idx / kchunks;
  const int k0 = (int)(idx % kchunks) * 8;
  const int ow = (int)(m % OW);
  const int64_t t = m / OW;
  const int oh = (int)(t % OH);
  const int64_t n = t / OH;
  const float* base = img + n * 3 * H * W;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = k0 + j;
    float x = 0.f;
    if (k < 147) {
      const int tap = k / 3, c = k - tap * 3, r = tap / 7, s = tap - r * 7;
      const int ih = oh * 2 - 3 + r, iw = ow * 2 - 3 + s;
      if (ih >= 0 && ih < H && iw >= 0 && iw < W) x = __ldg(base + ((int64_t)c * H + ih) * W + iw);
    }
    v[j] = x;
  }
  T* o = out + m * kpad + k0;
  if (sizeof(T) == 2) {                              // one 16-byte store per thread
    uint4 t;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    *reinterpret_cast<uint4*>(o) = t;
  } else {
    float* of = reinterpret_cast<float*>(o);
    *reinterpret_cast<float4*>(of) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(of + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
}

// Tiled form: one CTA per (image, output row).  The 7 input rows x 3 channels the row needs are staged in shared
// memory with coalesced float4 loads (zero halo of 4 columns, out-of-image rows zero), then every thread emits 16-byte
// chunks of the im2col matrix from a k -> (channel, row, column) offset table.  The per-thread form above spends its time
// in scattered __ldg + index arithmetic (0.9 TB/s of output); this one is bound by the 1 GB it writes.
template <typename T>
__global__ void __launch_bounds__(256) stem_im2col_tiled_kernel(const float* __restrict__ img, T* __restrict__ out, int H,
                                                                int W, int OH, int OW, int kpad) {
  extern __shared__ __align__(16) float st_smem[];
  const int WP = W + 8;                                     // padded row: 4 zero columns on each side
  float* rows = st_smem;                                    // [3][7][WP]
  int* lut = reinterpret_cast<int*>(st_smem + 21 * WP);     // [kpad] offset of k inside `rows`, -1 = zero padding of K
  const int oh = blockIdx.x, n = blockIdx.y, tid = threadIdx.x;
  const float* base = img + (int64_t)n * 3 * H * W;
  for (int k = tid; k < kpad; k += 256) {
    int off = -1;
    if (k < 147) {
      const int tap = k / 3, c = k - tap * 3, r = tap / 7, sx = tap - r * 7;
      off = (c * 7 + r) * WP + sx + 1;                      // column of iw = ow*2 - 3 + sx is 4 + iw = ow*2 + 1 + sx
    }
    lut[k] = off;
  }
  const int w4 = W / 4;
  for (int idx = tid; idx < 21 * (w4 + 2); idx += 256) {
    const int cr = idx / (w4 + 2), q = idx - cr * (w4 + 2);          // q = 0 and w4+1: the zero halos
    const int c = cr / 7, r = cr - c * 7, ih = oh * 2 - 3 + r;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q >= 1 && q <= w4 && ih >= 0 && ih < H)
      v = __ldg(reinterpret_cast<const float4*>(base + ((int64_t)c * H + ih) * W) + (q - 1));
    *reinterpret_cast<float4*>(rows + cr * WP + q * 4) = v;
  }
  __syncthreads();
  const int kchunks = kpad / 8;
  T* orow = out + ((int64_t)n * OH + oh) * OW * kpad;
  // thread = (pixel lane, k chunk): the eight offsets of its chunk live in registers for the whole row (the first
  // version looked them up in shared memory per element: 16 LDS per 16-byte chunk, shared-memory bound at 1.9 TB/s of
  // output); consecutive threads still write consecutive 16-byte chunks
  const int lanes = 256 / kchunks;                          // pixels in flight per pass (12 for kpad = 160)
  if (tid < lanes * kchunks) {
    const int kc = tid % kchunks, owl = tid / kchunks;
    int off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) off[j] = lut[kc * 8 + j];
    for (int ow = owl; ow < OW; ow += lanes) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = off[j] >= 0 ? rows[off[j] + ow * 2] : 0.f;
      T* o = orow + ((int64_t)ow * kchunks + kc) * 8;
      if (sizeof(T) == 2) {
        uint4 t;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
        for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
        *reinterpret_cast<uint4*>(o) = t;
      } else {
        float* of = reinterpret_cast<float*>(o);
        *reinterpret_cast<float4*>(of) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(of + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
    }
  }
}

// 16 bytes of channels per thread (8 bf16 / 4 fp32): 9 vector loads, one vector store.
template <typename T>
__global__ void maxpool_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t total, int H, int W, int C,
                               int OH, int OW) {
  constexpr int VN = 16 / sizeof(T);
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // over (n, oh, ow, c / VN)
  if (idx >= total) return;
  const int cv = C / VN;
  const int c = (int)(idx % cv) * VN;
  int64_t t = idx / cv;
  const int ow = (int)(t % OW); t /= OW;
  const int oh = (int)(t % OH);
  const int64_t n = t / OH;
  float best[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) best[i] = -INFINITY;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int ih = oh * 2 - 1 + r;
    if (ih < 0 || ih >= H) continue;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int iw = ow * 2 - 1 + s;
      if (iw < 0 || iw >= W) continue;
      const uint4 raw = *reinterpret_cast<const uint4*>(in + ((n * H + ih) * W + iw) * C + c);
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int i = 0; i < VN; ++i) best[i] = fmaxf(best[i], to_f(e[i]));
    }
  }
  uint4 o;
  T* e = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int i = 0; i < VN; ++i) e[i] = from_f<T>(best[i]);
  *reinterpret_cast<uint4*>(out + (((n * OH + oh) * OW + ow) * C + c)) = o;
}

__global__ void subsample_mask_kernel(const uint8_t* __restrict__ m, uint8_t* __restrict__ tok, int B, int H, int W,
                                      int gh, int gw, int S, int off) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * gh * gw) return;
  const int j = idx % gw, i = (idx / gw) % gh, b = idx / (gw * gh);
  // nearest: src = floor(dst * in/out)  (torch upsample_nearest; exact for 224 -> 14)
  const int si = min((int)floorf(i * ((float)H / gh)), H - 1);
  const int sj = min((int)floorf(j * ((float)W / gw)), W - 1);
  tok[(int64_t)b * S + off + i * gw + j] = m[((int64_t)b * H + si) * W + sj] ? 1 : 0;
}

__global__ void loc_tokens_kernel(const float* __restrict__ loc, const float* __restrict__ w,
                                  const float* __restrict__ bias, float* __restrict__ x, uint8_t* __restrict__ tok,
                                  int B, int n_loc, int per_scalar, int d, int S, int off) {
  const int ntok = per_scalar ? n_loc : 1;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * ntok * d) return;
  const int c = idx % d, k = (idx / d) % ntok, b = idx / (d * ntok);
  float v = bias[c];
  if (per_scalar) {
    v = fmaf(loc[b * n_loc + k], w[c], v);               // Linear(1, d): weight [d,1]
  } else {
    for (int i = 0; i < n_loc; ++i) v = fmaf(loc[b * n_loc + i], w[c * n_loc + i], v);
  }
  x[((int64_t)b * S + off + k) * d + c] = v;
  if (c == 0) tok[(int64_t)b * S + off + k] = 0;
}

// One warp per row; d <= 1024 (each lane holds d/32 values, d % 32 == 0).
template <typename T, int MAXPER>
__device__ __forceinline__ void ln_row(float (&v)[MAXPER], int per, int lane, const float* __restrict__ gamma,
                                       const float* __restrict__ beta, T* __restrict__ y, int d, float eps) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXPER; ++i) if (i < per) s += v[i];
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < MAXPER; ++i) if (i < per) { const float t = v[i] - mean; q = fmaf(t, t, q); }
  const float rstd = rsqrtf(warp_sum(q) / d + eps);
#pragma unroll
  for (int i = 0; i < MAXPER; ++i)
    if (i < per) {
      const int c = i * 32 + lane;
      y[c] = from_f<T>((v[i] - mean) * rstd * gamma[c] + beta[c]);
    }
}

// One warp per row, float4 loads (lane holds columns j*128 + lane*4 .. +3), d % 128 == 0.
template <typename T, int D>
__global__ void layernorm_vec_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                     const float* __restrict__ beta, T* __restrict__ y, int64_t rows, float eps,
                                     const int* __restrict__ rows_dyn, T* __restrict__ y2,
                                     const float* __restrict__ pos, const int* __restrict__ pos_idx) {
  constexpr int NV = D / 128;
  const int lane = threadIdx.x & 31;
  if (rows_dyn) rows = min(rows, (int64_t)__ldg(rows_dyn));      // packed encoder: device-side row count
  float4 g[NV], be[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    g[j] = __ldg(reinterpret_cast<const float4*>(gamma + j * 128 + lane * 4));
    be[j] = __ldg(reinterpret_cast<const float4*>(beta + j * 128 + lane * 4));
  }
  const int64_t wstride = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += wstride) {
    float4 v[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] = *reinterpret_cast<const float4*>(x + row * D + j * 128 + lane * 4);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) s += v[j].x + v[j].y + v[j].z + v[j].w;
    const float mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float a = v[j].x - mean, b2 = v[j].y - mean, c = v[j].z - mean, d2 = v[j].w - mean;
      q += a * a + b2 * b2 + c * c + d2 * d2;
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / D) + eps);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float o0 = (v[j].x - mean) * rstd * g[j].x + be[j].x, o1 = (v[j].y - mean) * rstd * g[j].y + be[j].y;
      const float o2 = (v[j].z - mean) * rstd * g[j].z + be[j].z, o3 = (v[j].w - mean) * rstd * g[j].w + be[j].w;
      T* dst = y + row * D + j * 128 + lane * 4;
      if (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(dst) = make_float4(o0, o1, o2, o3);
      } else {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o0, o1), hi = __floats2bfloat162_rn(o2, o3);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(dst) = pk;
      }
      if (y2) {                                  // second output: LN(x) + pos[pos_idx[row]] (with_pos_embed for q / k)
        const float4 pp = __ldg(reinterpret_cast<const float4*>(pos + (int64_t)__ldg(pos_idx + row) * D + j * 128 + lane * 4));
        T* dst2 = y2 + row * D + j * 128 + lane * 4;
        if (sizeof(T) == 4) {
          *reinterpret_cast<float4*>(dst2) = make_float4(o0 + pp.x, o1 + pp.y, o2 + pp.z, o3 + pp.w);
        } else {
          __nv_bfloat162 lo = __floats2bfloat162_rn(o0 + pp.x, o1 + pp.y), hi = __floats2bfloat162_rn(o2 + pp.z, o3 + pp.w);
          uint2 pk;
          pk.x = *reinterpret_cast<uint32_t*>(&lo);
          pk.y = *reinterpret_cast<uint32_t*>(&hi);
          *reinterpret_cast<uint2*>(dst2) = pk;
        }
      }
    }
  }
}

template <typename T>
__global__ void layernorm_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, T* __restrict__ y, int64_t rows, int d, float eps) {
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31, per = d >> 5;
  float v[32];
  const float* xr = x + row * d;
#pragma unroll
  for (int i = 0; i < 32; ++i) if (i < per) v[i] = xr[i * 32 + lane];
  ln_row<T, 32>(v, per, lane, gamma, beta, y + row * d, d, eps);
}

__global__ void embed_ln_kernel(const int64_t* __restrict__ ids, int64_t ld_ids, const float* __restrict__ e_word,
                                const float* __restrict__ e_pos, const float* __restrict__ gamma,
                                const float* __restrict__ beta, float* __restrict__ y, int B, int rows_per, int t0,
                                const int* __restrict__ t0_dyn, int d, float eps) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= (int64_t)B * rows_per) return;
  if (t0_dyn) t0 = *t0_dyn;
  const int b = (int)(row / rows_per), t = t0 + (int)(row % rows_per);
  const int lane = threadIdx.x & 31, per = d >> 5;
  const int64_t id = ids[b * ld_ids + t];
  const float* we = e_word + id * d;
  const float* pe = e_pos + (int64_t)t * d;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) if (i < per) v[i] = we[i * 32 + lane] + pe[i * 32 + lane];
  ln_row<float, 32>(v, per, lane, gamma, beta, y + row * d, d, eps);
}

}  // namespace

// Every kernel of the decode step asks for the maximum shared-memory carveout so that consecutive
// launches (227 KB GEMMs, zero-smem LayerNorms) never force an SM L1/shared reconfiguration.
void ops_prefer_max_shared() {
  const int v = cudaSharedmemCarveoutMaxShared;
  CUDA_OK(cudaFuncSetAttribute(layernorm_kernel<float>, cudaFuncAttributePreferredSharedMemoryCarveout, v));
  CUDA_OK(cudaFuncSetAttribute(layernorm_kernel<bf16>, cudaFuncAttributePreferredSharedMemoryCarveout, v));
  CUDA_OK(cudaFuncSetAttribute(embed_ln_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, v));
}

void f32_to_bf16(const float* src, bf16* dst, int64_t n, cudaStream_t s) {
  f32_to_bf16_kernel<<<blocks_for(n, 256), 256, 0, s>>>(src, dst, n);
  LAUNCH_OK();
}

template <typename T>
void repack_filter(const float* oihw, T* out, int O, int I, int kh, int kw, int kpad, cudaStream_t s) {
  repack_filter_kernel<T><<<blocks_for((int64_t)O * kpad, 256), 256, 0, s>>>(oihw, out, O, I, kh, kw, kpad);
  LAUNCH_OK();
}
template void repack_filter<float>(const float*, float*, int, int, int, int, int, cudaStream_t);
template void repack_filter<bf16>(const float*, bf16*, int, int, int, int, int, cudaStream_t);

template <typename T>
void stem_im2col(const float* img, T* out, int64_t nimg, int H, int W, int OH, int OW, int kpad, cudaStream_t s) {
  RETR_CHECK(kpad % 8 == 0 && kpad >= 147, "stem kpad");
  const int64_t rows = nimg * OH * OW;
  const size_t smem = (size_t)21 * (W + 8) * 4 + (size_t)kpad * 4;
  if (W % 4 == 0 && smem <= 48 * 1024 && nimg <= 65535 && OW == (W + 6 - 7) / 2 + 1 &&
      (reinterpret_cast<uintptr_t>(img) & 15) == 0) {
    stem_im2col_tiled_kernel<T><<<dim3((unsigned)OH, (unsigned)nimg), 256, smem, s>>>(img, out, H, W, OH, OW, kpad);
  } else {
    stem_im2col_kernel<T><<<blocks_for(rows * (kpad / 8), 256), 256, 0, s>>>(img, out, rows, H, W, OH, OW, kpad);
  }
  LAUNCH_OK();
}
template void stem_im2col<float>(const float*, float*, int64_t, int, int, int, int, int, cudaStream_t);
template void stem_im2col<bf16>(const float*, bf16*, int64_t, int, int, int, int, int, cudaStream_t);

template <typename T>
void maxpool_3x3s2(const T* in, T* out, int64_t nimg, int H, int W, int C, int OH, int OW, cudaStream_t s) {
  RETR_CHECK(C % (16 / (int)sizeof(T)) == 0, "maxpool channel count");
  const int64_t total = nimg * OH * OW * (C / (16 / (int)sizeof(T)));
  maxpool_kernel<T><<<blocks_for(total, 256), 256, 0, s>>>(in, out, total, H, W, C, OH, OW);
  LAUNCH_OK();
}
template void maxpool_3x3s2<float>(const float*, float*, int64_t, int, int, int, int, int, cudaStream_t);
template void maxpool_3x3s2<bf16>(const bf16*, bf16*, int64_t, int, int, int, int, int, cudaStream_t);

void subsample_mask(const uint8_t* img_mask, uint8_t* tok_mask, int B, int H, int W, int gh, int gw, int S, int off,
                    cudaStream_t s) {
  subsample_mask_kernel<<<blocks_for((int64_t)B * gh * gw, 256), 256, 0, s>>>(img_mask, tok_mask, B, H, W, gh, gw, S, off);
  LAUNCH_OK();
}

void loc_tokens(const float* loc, const float* w, const float* bias, float* x, uint8_t* tok_mask, int B, int n_loc,
                int per_scalar, int d, int S, int off, cudaStream_t s) {
  const int ntok = per_scalar ? n_loc : 1;
  loc_tokens_kernel<<<blocks_for((int64_t)B * ntok * d, 256), 256, 0, s>>>(loc, w, bias, x, tok_mask, B, n_loc,
                                                                            per_scalar, d, S, off);
  LAUNCH_OK();
}

template <typename T>
void layernorm(const float* x, const float* gamma, const float* beta, T* y, int64_t rows, int d, float eps,
               cudaStream_t s, const int* rows_dyn, T* y2, const float* pos, const int* pos_idx) {
  RETR_CHECK(d % 32 == 0 && d <= 1024, "LayerNorm width must be a multiple of 32, <= 1024");
  RETR_CHECK(!y2 || ((d == 256 || d == 512) && pos && pos_idx), "LayerNorm: the position-added output needs d = 256 / 512");
  if (d == 256 || d == 512) {
    const unsigned grid = (unsigned)std::min<int64_t>(cdiv(rows, 8), 148 * 8);     // grid-stride, 8 warps per CTA
    if (d == 256) layernorm_vec_kernel<T, 256><<<grid, 256, 0, s>>>(x, gamma, beta, y, rows, eps, rows_dyn, y2, pos, pos_idx);
    else layernorm_vec_kernel<T, 512><<<grid, 256, 0, s>>>(x, gamma, beta, y, rows, eps, rows_dyn, y2, pos, pos_idx);
    LAUNCH_OK();
    return;
  }
  layernorm_kernel<T><<<blocks_for(rows, 8), 256, 0, s>>>(x, gamma, beta, y, rows, d, eps);
  LAUNCH_OK();
}
template void layernorm<float>(const float*, const float*, const float*, float*, int64_t, int, float, cudaStream_t,
                               const int*, float*, const float*, const int*);
template void layernorm<bf16>(const float*, const float*, const float*, bf16*, int64_t, int, float, cudaStream_t,
                              const int*, bf16*, const float*, const int*);

void embed_ln(const int64_t* ids, int64_t ld_ids, const float* e_word, const float* e_pos, const float* gamma,
              const float* beta, float* y, int B, int rows_per, int t0, const int* t0_dyn, int d, float eps,
              cudaStream_t s) {
  RETR_CHECK(d % 32 == 0 && d <= 1024, "embedding width");
  launch_pdl(embed_ln_kernel, dim3(blocks_for((int64_t)B * rows_per, 8)), dim3(256), 0, s, ids, ld_ids, e_word, e_pos,
             gamma, beta, y, B, rows_per, t0, t0_dyn, d, eps);
}

}  // namespace retr
