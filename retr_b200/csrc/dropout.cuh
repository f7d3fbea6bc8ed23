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

}  // namespace retr
