// Counter-based dropout shared by every training kernel: Philox4x32-10 keyed by `seed`, counter =
// (element index / 4, site id); element idx uses word idx % 4.  A pure function of (seed, site, idx), so the
// backward pass regenerates the forward mask instead of storing it.
#pragma once
#include "train_ops.h"

namespace retr {

__device__ __forceinline__ uint4 philox4(uint2 key, uint4 ctr) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0; key.y += W1;
  }
  return ctr;
}
// 0 for a dropped element, 1/(1-p) for a kept one
__device__ __forceinline__ float drop_scale(const DropCfg& dc, uint64_t idx) {
  if (dc.p <= 0.f) return 1.f;
  const uint4 r = philox4(make_uint2((uint32_t)dc.seed, (uint32_t)(dc.seed >> 32)),
                          make_uint4((uint32_t)(idx >> 2), (uint32_t)(idx >> 34), dc.site, 0u));
  const uint32_t sel = (uint32_t)idx & 3u;
  const uint32_t w = sel == 0 ? r.x : (sel == 1 ? r.y : (sel == 2 ? r.z : r.w));
  return ((float)w * 2.3283064365386963e-10f >= dc.p) ? 1.f / (1.f - dc.p) : 0.f;
}

// Attention probabilities are indexed (row = (b*H + h)*Lq + i, key j).  One Philox evaluation covers a 2 x 2 patch
// (rows row&~1..|1, keys j&~1..|1; word = 2*(row&1) + (j&1)), so the two adjacent keys a thread owns in an MMA
// accumulator fragment of the forward / dQ kernels, and the two adjacent queries it owns in the transposed
// dK/dV kernel, each come from ONE evaluation.
__device__ __forceinline__ uint4 drop_block_rc(const DropCfg& dc, uint64_t row, uint32_t col) {
  return philox4(make_uint2((uint32_t)dc.seed, (uint32_t)(dc.seed >> 32)),
                 make_uint4(col >> 1, (uint32_t)(row >> 1), dc.site ^ 0x80000000u, (uint32_t)(row >> 33)));
}
__device__ __forceinline__ float drop_keep(const DropCfg& dc, uint32_t w) {
  return ((float)w * 2.3283064365386963e-10f >= dc.p) ? 1.f / (1.f - dc.p) : 0.f;
}
__device__ __forceinline__ float drop_scale_rc(const DropCfg& dc, uint64_t row, uint32_t col) {
  if (dc.p <= 0.f) return 1.f;
  const uint4 r = drop_block_rc(dc, row, col);
  const uint32_t sel = (((uint32_t)row & 1u) << 1) | (col & 1u);
  return drop_keep(dc, sel == 0 ? r.x : (sel == 1 ? r.y : (sel == 2 ? r.z : r.w)));
}
// (row, col) and (row, col + 1), col even
__device__ __forceinline__ float2 drop_scale_cols2(const DropCfg& dc, uint64_t row, uint32_t col) {
  if (dc.p <= 0.f) return make_float2(1.f, 1.f);
  const uint4 r = drop_block_rc(dc, row, col);
  return (row & 1u) ? make_float2(drop_keep(dc, r.z), drop_keep(dc, r.w)) : make_float2(drop_keep(dc, r.x), drop_keep(dc, r.y));
}
// (row, col) and (row + 1, col), row even
__device__ __forceinline__ float2 drop_scale_rows2(const DropCfg& dc, uint64_t row, uint32_t col) {
  if (dc.p <= 0.f) return make_float2(1.f, 1.f);
  const uint4 r = drop_block_rc(dc, row, col);
  return (col & 1u) ? make_float2(drop_keep(dc, r.y), drop_keep(dc, r.w)) : make_float2(drop_keep(dc, r.x), drop_keep(dc, r.z));
}

}  // namespace retr
