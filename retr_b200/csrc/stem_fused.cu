// ResNet stem in ONE kernel: 7x7/2 convolution (3 -> 64) + FrozenBatchNorm2d + ReLU + 3x3/2 max-pool, from the fp32 NCHW
// image straight to the pooled NHWC bf16 activation.
//
// Reference being replaced: torchvision ResNet.conv1 / bn1 / relu / maxpool (torchvision/models/resnet.py:266-269) as
// kept by IntermediateLayerGetter (models/backbone.py:61-65), bn1 = FrozenBatchNorm2d (models/backbone.py:41-51).
//
// The three-launch form (im2col matrix -> GEMM -> max-pool) writes a 2 GB im2col matrix and a 0.8 GB conv output per
// 512-image pass and reads both back: 1.5 ms of a 15.7 ms backbone pass for 0.5 GB of compulsory traffic.  Here a CTA
// owns an 8x8 tile of POOLED pixels of one image: it stages the 39x39x3 input patch in shared memory as bf16, runs the
// 17x17 convolution outputs it needs as an implicit GEMM on mma.sync (A fragments gathered from the patch with one
// 32-bit shared load per register: K is ordered (channel, row, tap) with the 7 taps padded to 8, so a register pair is
// two adjacent input columns; weights as ldmatrix B fragments), applies scale/shift + ReLU, and max-pools from a
// shared-memory tile.  Nothing but the image is read and nothing but the pooled activation is written.
#include <algorithm>

#include "ops.h"

namespace retr {
namespace {

constexpr int SF_PT = 8;                      // pooled tile side
constexpr int SF_CT = 2 * SF_PT + 1;          // 17 conv rows / cols per tile
constexpr int SF_IN = 2 * SF_CT + 5;          // 39 input rows / cols per tile
constexpr int SF_PW = 48;                     // patch row pitch (elements; cols 0..39 used)
constexpr int SF_KG = 22;                     // (channel, row) groups of 8 taps: 21 real + 1 zero -> K = 176
constexpr int SF_K = SF_KG * 8;
constexpr int SF_WP = SF_K + 8;               // weight row pitch (conflict-free ldmatrix)
constexpr int SF_CP = 72;                     // conv tile pixel pitch (64 channels + 8)
constexpr int SF_NPIX = SF_CT * SF_CT;        // 289
constexpr int SF_MT = (SF_NPIX + 15) / 16;    // 19 m-tiles

struct SfSmem {
  bf16 patch[3][SF_IN][SF_PW];                // 11.2 KB
  bf16 w[64][SF_WP];                          // 23.0 KB
  bf16 conv[SF_NPIX][SF_CP];                  // 40.6 KB
  float scale[64], shift[64];
};

__device__ __forceinline__ void sf_ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void sf_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// wk: [64][SF_K] bf16, k = (c * 7 + r) * 8 + s (s = 7 and group 21: zero)
// Persistent: a CTA keeps the weights in shared memory and walks tiles blockIdx.x, + gridDim.x, ...; the NEXT tile's input
// patch is loaded into registers before the current tile's MMAs start and written to shared memory after them (the
// one-tile-per-CTA version spent 45 % of its time waiting for the patch loads: profiles/r2_ncu_full_stem_fused.txt).
constexpr int SF_FILL = 3 * SF_IN * (SF_PW / 2);            // bf16 pairs of a patch
constexpr int SF_FILL_IT = (SF_FILL + 255) / 256;           // 11 per thread

__global__ void __launch_bounds__(256, 2)
stem_fused_kernel(const float* __restrict__ img, const bf16* __restrict__ wk, const float* __restrict__ scale,
                  const float* __restrict__ shift, bf16* __restrict__ out, int H, int W, int CH, int CW, int PH, int PW,
                  int total_tiles) {
  extern __shared__ __align__(16) uint8_t sf_raw[];
  SfSmem& sm = *reinterpret_cast<SfSmem*>(sf_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, c4 = lane & 3;
  const int tiles_x = PW / SF_PT, tiles_img = tiles_x * (PH / SF_PT);
  // weights + BN: once per CTA
  for (int idx = tid; idx < 64 * (SF_K / 8); idx += 256) {
    const int r = idx / (SF_K / 8), ch = idx - r * (SF_K / 8);
    *reinterpret_cast<uint4*>(&sm.w[r][ch * 8]) = __ldg(reinterpret_cast<const uint4*>(wk + (int64_t)r * SF_K + ch * 8));
  }
  if (tid < 64) { sm.scale[tid] = __ldg(scale + tid); sm.shift[tid] = __ldg(shift + tid); }
  float pv[SF_FILL_IT][2];
  auto load_patch = [&](int tile) {
    const int n = tile / tiles_img, t = tile - n * tiles_img, ty = t / tiles_x, tx = t - ty * tiles_x;
    const int i0y = 2 * (2 * ty * SF_PT - 1) - 3, i0x = 2 * (2 * tx * SF_PT - 1) - 3;
    const float* base = img + (int64_t)n * 3 * H * W;
#pragma unroll
    for (int it = 0; it < SF_FILL_IT; ++it) {
      const int idx = tid + it * 256;
      const int cr = idx / (SF_PW / 2), q = idx - cr * (SF_PW / 2);
      const int c = cr / SF_IN, r = cr - c * SF_IN, iy = i0y + r, ix = i0x + 2 * q;
      float v0 = 0.f, v1 = 0.f;
      if (idx < SF_FILL && iy >= 0 && iy < H) {
        const float* rowp = base + ((int64_t)c * H + iy) * W;
        if (ix >= 0 && ix < W) v0 = __ldg(rowp + ix);
        if (ix + 1 >= 0 && ix + 1 < W) v1 = __ldg(rowp + ix + 1);
      }
      pv[it][0] = v0; pv[it][1] = v1;
    }
  };
  auto store_patch = [&]() {
#pragma unroll
    for (int it = 0; it < SF_FILL_IT; ++it) {
      const int idx = tid + it * 256;
      if (idx < SF_FILL)
        *reinterpret_cast<__nv_bfloat162*>(&sm.patch[0][0][0] + 2 * idx) = __floats2bfloat162_rn(pv[it][0], pv[it][1]);
    }
  };
  if ((int)blockIdx.x < total_tiles) { load_patch(blockIdx.x); store_patch(); }
  __syncthreads();

  const uint32_t w_base = (uint32_t)__cvta_generic_to_shared(&sm.w[(lane >> 4) * 8 + (lane & 7)][((lane >> 3) & 1) * 8]);
  const bf16* patch0 = &sm.patch[0][0][0];
#pragma unroll 1
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
  const int n = tile / tiles_img, t_in = tile - n * tiles_img, ty = t_in / tiles_x, tx = t_in - ty * tiles_x;
  const int p0y = ty * SF_PT, p0x = tx * SF_PT;            // first pooled pixel of the tile
  const int c0y = 2 * p0y - 1, c0x = 2 * p0x - 1;          // first conv pixel (max-pool 3x3/2 pad 1)
  const bool has_next = tile + (int)gridDim.x < total_tiles;
  if (has_next) load_patch(tile + gridDim.x);              // in flight under this tile's MMAs
  // implicit GEMM: M = 289 conv pixels (19 m-tiles over 8 warps), N = 64, K = 176; one m-tile at a time (32 accumulator
  // registers: two CTAs per SM, so that one CTA's fill / epilogue / pooling hides under the other's MMAs)
#pragma unroll 1
  for (int mt = warp; mt < SF_MT; mt += 8) {
    float acc[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    int poff[2];                                            // patch offset of the fragment rows g and g + 8
#pragma unroll
    for (int hrow = 0; hrow < 2; ++hrow) {
      const int p = min(mt * 16 + g + 8 * hrow, SF_NPIX - 1);
      const int py = p / SF_CT, px = p - py * SF_CT;
      poff[hrow] = (2 * py) * SF_PW + 2 * px + 2 * c4;
    }
#pragma unroll
    for (int ks = 0; ks < SF_KG / 2; ++ks) {
      // the two (channel, row) groups of this k-step; the zero group 21 reads group 20's data (weights are zero)
      const int g0 = 2 * ks, g1 = 2 * ks + 1 < 21 ? 2 * ks + 1 : 20;
      const int o0 = ((g0 / 7) * SF_IN + g0 % 7) * SF_PW, o1 = ((g1 / 7) * SF_IN + g1 % 7) * SF_PW;
      uint32_t a[4];
      a[0] = *reinterpret_cast<const uint32_t*>(patch0 + o0 + poff[0]);
      a[1] = *reinterpret_cast<const uint32_t*>(patch0 + o0 + poff[1]);
      a[2] = *reinterpret_cast<const uint32_t*>(patch0 + o1 + poff[0]);
      a[3] = *reinterpret_cast<const uint32_t*>(patch0 + o1 + poff[1]);
#pragma unroll
      for (int np = 0; np < 4; ++np) {                      // n-tile pairs: rows np*16 .. +15 of the weights
        uint32_t b0, b1, b2, b3;
        sf_ldsm_x4(w_base + (uint32_t)(np * 16 * SF_WP + ks * 16) * 2, b0, b1, b2, b3);
        sf_mma(acc[2 * np], a, b0, b1);
        sf_mma(acc[2 * np + 1], a, b2, b3);
      }
    }
    // scale/shift + ReLU -> conv tile (positions outside the convolution's output are max-pool padding: 0 after ReLU)
#pragma unroll
    for (int hrow = 0; hrow < 2; ++hrow) {
      const int p = mt * 16 + g + 8 * hrow;
      if (p < SF_NPIX) {
        const int py = p / SF_CT, px = p - py * SF_CT, cy = c0y + py, cx = c0x + px;
        const bool inside = cy >= 0 && cy < CH && cx >= 0 && cx < CW;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int ch = nt * 8 + 2 * c4;
          const float v0 = inside ? fmaxf(acc[nt][2 * hrow] * sm.scale[ch] + sm.shift[ch], 0.f) : 0.f;
          const float v1 = inside ? fmaxf(acc[nt][2 * hrow + 1] * sm.scale[ch + 1] + sm.shift[ch + 1], 0.f) : 0.f;
          *reinterpret_cast<__nv_bfloat162*>(&sm.conv[p][ch]) = __floats2bfloat162_rn(v0, v1);
        }
      }
    }
  }
  __syncthreads();                                         // conv tile complete, every read of the patch done
  if (has_next) store_patch();
  // 3x3/2 max-pool: pooled (py, px) covers conv tile rows 2py .. 2py+2; one 8-channel chunk per thread item
  for (int idx = tid; idx < SF_PT * SF_PT * 8; idx += 256) {
    const int chunk = idx & 7, pp = idx >> 3, py = pp / SF_PT, px = pp - py * SF_PT;
    __nv_bfloat162 m[4];
    {
      const uint4 t = *reinterpret_cast<const uint4*>(&sm.conv[(2 * py) * SF_CT + 2 * px][chunk * 8]);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
      m[0] = h[0]; m[1] = h[1]; m[2] = h[2]; m[3] = h[3];
    }
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int dx = 0; dx < 3; ++dx) {
        if (dy == 0 && dx == 0) continue;
        const uint4 t = *reinterpret_cast<const uint4*>(&sm.conv[(2 * py + dy) * SF_CT + 2 * px + dx][chunk * 8]);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
        for (int j = 0; j < 4; ++j) m[j] = __hmax2(m[j], h[j]);
      }
    uint4 o;
    __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
    oh[0] = m[0]; oh[1] = m[1]; oh[2] = m[2]; oh[3] = m[3];
    *reinterpret_cast<uint4*>(out + (((int64_t)n * PH + p0y + py) * PW + p0x + px) * 64 + chunk * 8) = o;
  }
  __syncthreads();                                         // conv tile free again, next patch visible
  }
}

// OIHW fp32 [64,3,7,7] -> [64][SF_K] bf16 in the kernel's K order
__global__ void stem_repack_kernel(const float* __restrict__ w, bf16* __restrict__ out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 64 * SF_K) return;
  const int o = idx / SF_K, k = idx - o * SF_K, grp = k >> 3, s = k & 7;
  float v = 0.f;
  if (grp < 21 && s < 7) {
    const int c = grp / 7, r = grp - c * 7;
    v = w[((o * 3 + c) * 7 + r) * 7 + s];
  }
  out[idx] = __float2bfloat16_rn(v);
}

int sm_count_cached() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    CUDA_OK(cudaGetDevice(&dev));
    CUDA_OK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  }
  return n;
}

}  // namespace

bool stem_fused_supported(int H, int W, int cout, int k, int stride, int pad) {
  if (cout != 64 || k != 7 || stride != 2 || pad != 3) return false;
  const int CH = (H + 6 - 7) / 2 + 1, CW = (W + 6 - 7) / 2 + 1;
  const int PH = (CH + 2 - 3) / 2 + 1, PW = (CW + 2 - 3) / 2 + 1;
  return PH % SF_PT == 0 && PW % SF_PT == 0;
}
int stem_fused_weight_elems() { return 64 * SF_K; }
void stem_fused_repack(const float* w_oihw, bf16* out, cudaStream_t s) {
  stem_repack_kernel<<<(64 * SF_K + 255) / 256, 256, 0, s>>>(w_oihw, out);
  LAUNCH_OK();
}
void stem_fused(const float* img, const bf16* wk, const float* scale, const float* shift, bf16* out, int64_t nimg, int H,
                int W, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CUDA_OK(cudaFuncSetAttribute(stem_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SfSmem)));
    configured = true;
  }
  const int CH = (H + 6 - 7) / 2 + 1, CW = (W + 6 - 7) / 2 + 1, PH = (CH + 2 - 3) / 2 + 1, PW = (CW + 2 - 3) / 2 + 1;
  const int64_t total = nimg * (PH / SF_PT) * (PW / SF_PT);
  RETR_CHECK(total < (1ll << 31), "stem_fused: too many tiles per launch");
  const int grid = (int)std::min<int64_t>(total, 2ll * sm_count_cached());
  stem_fused_kernel<<<grid, 256, sizeof(SfSmem), s>>>(img, wk, scale, shift, out, H, W, CH, CW, PH, PW, (int)total);
  LAUNCH_OK();
}

}  // namespace retr
