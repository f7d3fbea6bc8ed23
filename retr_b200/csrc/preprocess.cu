// GPU preprocessing: from raw RGB images + bounding boxes to the encoder inputs of the hot path, bit-identical to the
// reference's PIL / torchvision pipeline in validation mode.
//
// Reference being replaced (per sample, on the host, in DataLoader workers):
//   data_utils/refcoco.py:105-188  RefCocoCaption.__getitem__ (image half) with get_transforms('val') (:14-46)
//   data_utils/utils.py:161-196    crop_image_to_bb: target crop; context = image with the box zeroed, box masked
//   data_utils/utils.py:232-256    pad_img_to_max (ImageOps.pad, colour 0, centred), pad_mask_to_max (True bands)
//   data_utils/utils.py:199-229    compute_position_features
//   torchvision Resize([224], bilinear) on the PIL image = PIL's antialiased two-pass fixed-point resampler
//   (Resample.c: double coefficients -> 22-bit fixed point, horizontal pass to uint8, then vertical pass to uint8),
//   ToTensor (/255) and Normalize (ImageNet mean / std) in float32;
//   Resize on the BOOL mask tensor = aten's antialiased bilinear in float32 followed by .to(bool): an output pixel
//   is masked iff any masked input pixel has a positive weight in its window.
// One thread per output pixel evaluates both passes of the resampler for its 3 channels (the uint8 rounding of the
// horizontal pass is reproduced per tap row) and the mask window; the padded square and the crop are never
// materialised.  Quirk kept: the image is pasted at round-half-even(diff/2), the mask bands start at floor(diff/2).
#include "ops.h"

namespace retr {
namespace {

constexpr int PP_MAX_TAPS = 64;          // 2*ceil(S/out)+1 <= 64  ->  padded side S <= 31 * out
constexpr int PP_PREC = 22;              // PIL: PRECISION_BITS = 32 - 8 - 2

struct Geometry {
  int W, H;                              // source image
  int bx0, by0, bx1, by1;                // bounding box clipped to the image (numpy slicing)
  int rx0, ry0, rw, rh;                  // region that becomes the picture: the crop (target) or the whole image (context)
  int S, ox, oy, mox, moy;               // padded side, paste offset of the picture, start of the unmasked span of the mask
  bool context;
};

__device__ __forceinline__ Geometry geometry(const int32_t* hw, const double* bb, int b, bool context) {
  Geometry g;
  g.H = hw[2 * b]; g.W = hw[2 * b + 1];
  const long long x = (long long)rint(bb[4 * b]), y = (long long)rint(bb[4 * b + 1]);       // Python round(): half to even
  const long long w = (long long)rint(bb[4 * b + 2]), h = (long long)rint(bb[4 * b + 3]);
  auto clampi = [](long long v, int hi) { return (int)(v < 0 ? 0 : (v > hi ? hi : v)); };
  g.bx0 = clampi(x, g.W); g.bx1 = clampi(x + w, g.W); g.by0 = clampi(y, g.H); g.by1 = clampi(y + h, g.H);
  g.context = context;
  if (context) { g.rx0 = 0; g.ry0 = 0; g.rw = g.W; g.rh = g.H; }
  else { g.rx0 = g.bx0; g.ry0 = g.by0; g.rw = g.bx1 - g.bx0; g.rh = g.by1 - g.by0; }
  g.S = max(g.rw, g.rh);
  g.ox = g.oy = g.mox = g.moy = 0;
  if (g.rw != g.S) { g.ox = (int)rint((g.S - g.rw) * 0.5); g.mox = (g.S - g.rw) / 2; }    // ImageOps.pad / pad_mask_to_max
  else if (g.rh != g.S) { g.oy = (int)rint((g.S - g.rh) * 0.5); g.moy = (g.S - g.rh) / 2; }
  return g;
}

// PIL precompute_coeffs for one output index, bilinear filter (support 1), then normalize_coeffs_8bpc
__device__ __forceinline__ void pil_coeffs(int in_size, int out_size, int o, int& lo, int& n, int* k) {
  const double scale = (double)in_size / out_size, fs = scale < 1.0 ? 1.0 : scale, support = fs, ss = 1.0 / fs;
  const double center = (o + 0.5) * scale;
  int xmin = (int)(center - support + 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)(center + support + 0.5);
  if (xmax > in_size) xmax = in_size;
  n = xmax - xmin; lo = xmin;
  double ww = 0.0;
  for (int x = 0; x < n; ++x) {
    double v = (x + xmin - center + 0.5) * ss;
    if (v < 0.0) v = -v;
    ww += v < 1.0 ? 1.0 - v : 0.0;
  }
  for (int x = 0; x < n; ++x) {
    double v = (x + xmin - center + 0.5) * ss;
    if (v < 0.0) v = -v;
    double wgt = v < 1.0 ? 1.0 - v : 0.0;
    if (ww != 0.0) wgt /= ww;
    k[x] = (int)(0.5 + wgt * (double)(1 << PP_PREC));
  }
}

// aten _compute_indices_min_size_weights_aa (float32): the taps of output index o whose bilinear weight is positive
__device__ __forceinline__ void aa_window(int in_size, int out_size, int o, int& lo, int& n, unsigned long long& live) {
  const float scale = (float)in_size / (float)out_size;
  const float support = scale >= 1.f ? scale : 1.f, invscale = scale >= 1.f ? 1.f / scale : 1.f;
  const float center = scale * ((float)o + 0.5f);
  int xmin = (int)(center - support + 0.5f);
  if (xmin < 0) xmin = 0;
  int xmax = (int)(center + support + 0.5f);
  if (xmax > in_size) xmax = in_size;
  lo = xmin; n = xmax - xmin;
  live = 0ull;
  for (int j = 0; j < n && j < 64; ++j) {
    const float v = fabsf(((float)(j + xmin) - center + 0.5f) * invscale);
    if (v < 1.0f) live |= 1ull << j;
  }
}

__device__ __forceinline__ int clip8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ pixels, const int64_t* __restrict__ offsets, const int32_t* __restrict__ hw,
                  const double* __restrict__ bb, int out, float* __restrict__ t_img, uint8_t* __restrict__ t_mask,
                  float* __restrict__ g_img, uint8_t* __restrict__ g_mask, float* __restrict__ loc) {
  const int b = blockIdx.y, which = blockIdx.z, oy_ = blockIdx.x;
  const Geometry g = geometry(hw, bb, b, which == 1);
  float* img_out = which ? g_img : t_img;
  uint8_t* msk_out = which ? g_mask : t_mask;
  if (which == 0 && oy_ == 0 && threadIdx.x == 0 && loc) {
    // compute_position_features: Python float (double) arithmetic on the UNROUNDED box, then float32
    const double x = bb[4 * b], y = bb[4 * b + 1], w = bb[4 * b + 2], h = bb[4 * b + 3], iw = g.W, ih = g.H;
    loc[5 * b] = (float)(x / iw); loc[5 * b + 1] = (float)(y / ih);
    loc[5 * b + 2] = (float)((x + w) / iw); loc[5 * b + 3] = (float)((y + h) / ih);
    loc[5 * b + 4] = (float)((w * h) / (iw * ih));
  }
  const int ox_ = threadIdx.x;
  if (ox_ >= out) return;
  const uint8_t* src = pixels + offsets[b];
  int ky[PP_MAX_TAPS], kx[PP_MAX_TAPS], ylo, yn, xlo, xn;
  pil_coeffs(g.S, out, oy_, ylo, yn, ky);
  pil_coeffs(g.S, out, ox_, xlo, xn, kx);
  int accv[3] = {1 << (PP_PREC - 1), 1 << (PP_PREC - 1), 1 << (PP_PREC - 1)};
  for (int j = 0; j < yn; ++j) {
    const int r = ylo + j - g.oy;                          // row of the picture under padded row ylo + j
    int acch[3] = {1 << (PP_PREC - 1), 1 << (PP_PREC - 1), 1 << (PP_PREC - 1)};
    if (r >= 0 && r < g.rh) {
      const int sy = g.ry0 + r;
      const bool row_in_box = g.context && sy >= g.by0 && sy < g.by1;
      const uint8_t* row = src + ((size_t)sy * g.W) * 3;
      for (int i = 0; i < xn; ++i) {
        const int c = xlo + i - g.ox;
        if (c < 0 || c >= g.rw) continue;                  // padding colour 0
        const int sx = g.rx0 + c;
        if (row_in_box && sx >= g.bx0 && sx < g.bx1) continue;     // the target is blanked out of the context image
        const uint8_t* p = row + (size_t)sx * 3;
        acch[0] += (int)p[0] * kx[i]; acch[1] += (int)p[1] * kx[i]; acch[2] += (int)p[2] * kx[i];
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) accv[c] += clip8(acch[c] >> PP_PREC) * ky[j];      // uint8 result of the horizontal pass
  }
  const float mean[3] = {0.485f, 0.456f, 0.406f}, stdv[3] = {0.229f, 0.224f, 0.225f};
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v = (float)clip8(accv[c] >> PP_PREC) / 255.0f;                      // ToTensor
    img_out[(((size_t)b * 3 + c) * out + oy_) * out + ox_] = (v - mean[c]) / stdv[c];   // Normalize
  }
  // mask: any masked pixel of the padded mask under a positive-weight tap
  int mylo, myn, mxlo, mxn;
  unsigned long long ylive, xlive;
  aa_window(g.S, out, oy_, mylo, myn, ylive);
  aa_window(g.S, out, ox_, mxlo, mxn, xlive);
  bool masked = false;
  for (int j = 0; j < myn && !masked; ++j) {
    if (!((ylive >> j) & 1ull)) continue;
    const int r = mylo + j - g.moy;
    const bool row_pad = r < 0 || r >= g.rh;
    const int sy = g.ry0 + r;
    const bool row_in_box = g.context && !row_pad && sy >= g.by0 && sy < g.by1;
    for (int i = 0; i < mxn; ++i) {
      if (!((xlive >> i) & 1ull)) continue;
      const int c = mxlo + i - g.mox;
      const int sx = g.rx0 + c;
      if (row_pad || c < 0 || c >= g.rw || (row_in_box && sx >= g.bx0 && sx < g.bx1)) { masked = true; break; }
    }
  }
  msk_out[((size_t)b * out + oy_) * out + ox_] = masked ? 1 : 0;
}

}  // namespace

void preprocess(const uint8_t* pixels, const int64_t* offsets, const int32_t* hw, const double* bb, int batch, int out,
                float* t_img, uint8_t* t_mask, float* g_img, uint8_t* g_mask, float* loc, cudaStream_t s) {
  RETR_CHECK(pixels && offsets && hw && bb && t_img && t_mask, "null argument");
  RETR_CHECK((g_img == nullptr) == (g_mask == nullptr), "context image and mask go together");
  RETR_CHECK(batch >= 1 && batch <= 65535 && out >= 1 && out <= 256, "batch in [1, 65535], output side in [1, 256]");
  preprocess_kernel<<<dim3((unsigned)out, (unsigned)batch, g_img ? 2u : 1u), 256, 0, s>>>(pixels, offsets, hw, bb, out, t_img,
                                                                                         t_mask, g_img, g_mask, loc);
  LAUNCH_OK();
}

int preprocess_max_side(int out) { return (PP_MAX_TAPS - 1) / 2 * out; }

}  // namespace retr
