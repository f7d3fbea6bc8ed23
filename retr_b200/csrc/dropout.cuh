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

// Attention probabilities are indexed (row = (b*H + h)*Lq + i, key j).  One Philox evaluation yields EIGHT 16-bit
// decisions: row `row`, the eight keys {32*G + 8*t + 2*c + e : t = 0..3, e = 0..1} — the keys one thread of an MMA
// accumulator fragment owns across four consecutive 8-key tiles (its column pair c).  Field (t, e) = half e of word t;
// an element is kept iff its field >= round(p * 65536), and kept elements are scaled by the exact reciprocal of that
// keep probability.  (The first version spent one evaluation per 2 x 2 patch of which a thread used half: the Philox
// rounds were two thirds of the ALU work of the three training attention kernels — 20 ms of a 77 ms step.)
__device__ __forceinline__ uint4 drop_block_row8(const DropCfg& dc, uint64_t row, uint32_t col) {
  return philox4(make_uint2((uint32_t)dc.seed, (uint32_t)(dc.seed >> 32)),
                 make_uint4(((col >> 5) << 2) | ((col >> 1) & 3u), (uint32_t)row, dc.site ^ 0x80000000u,
                            (uint32_t)(row >> 32)));
}
__device__ __forceinline__ uint32_t drop_thr16(const DropCfg& dc) { return (uint32_t)(dc.p * 65536.f + 0.5f); }
__device__ __forceinline__ uint32_t drop_word(const uint4& r, uint32_t t) {
  return t == 0 ? r.x : (t == 1 ? r.y : (t == 2 ? r.z : r.w));
}
// both halves of one word: (row, col) and (row, col + 1), col even
__device__ __forceinline__ float2 drop_keep_pair(const DropCfg& dc, uint32_t w) {
  const uint32_t thr = drop_thr16(dc);
  const float sc = 65536.f / (65536.f - (float)thr);
  return make_float2((w & 0xFFFFu) >= thr ? sc : 0.f, (w >> 16) >= thr ? sc : 0.f);
}
__device__ __forceinline__ float drop_keep_half(const DropCfg& dc, uint32_t w, uint32_t e) {
  const float2 k = drop_keep_pair(dc, w);
  return e ? k.y : k.x;
}
__device__ __forceinline__ float drop_scale_rc(const DropCfg& dc, uint64_t row, uint32_t col) {
  if (dc.p <= 0.f) return 1.f;
  return drop_keep_half(dc, drop_word(drop_block_row8(dc, row, col), (col >> 3) & 3u), col & 1u);
}

}  // namespace retr
