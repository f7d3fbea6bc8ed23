// Multi-head attention core with key-padding and causal masks (flash-style, scores never
// leave the SM).  Restates nn.MultiheadAttention's core as the reference calls it
// (torch/nn/functional.py:6630-6664: q scaled by 1/sqrt(dh), additive -inf masks, softmax,
// P.V) for the encoder self-attention, and the teacher-forced decoder self/cross attention
// (models/transformer_modules.py:22-74).
//
// v1 kernel: fp32 SIMT math for both precisions (inputs/outputs fp32 or bf16).  One thread owns
// one query row (q and the output accumulator live in registers); the CTA stages K/V tiles in
// shared memory as fp32 and every lane reads the same key row (broadcast, conflict-free).
// Online softmax in chunks of 4 keys.
#include "ops.h"

#include <type_traits>

namespace retr {
namespace {

constexpr int QT = 64;    // queries (threads) per CTA
constexpr int KT = 64;    // keys per shared-memory tile

template <typename T, int DH>
__global__ void __launch_bounds__(QT) attention_kernel(const AttnParams p) {
  __shared__ __align__(16) float Ks[KT][DH];
  __shared__ __align__(16) float Vs[KT][DH];
  __shared__ uint8_t Ms[KT];
  const int b = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * QT + threadIdx.x;
  const bool active = qi < p.Lq;
  const T* qp = static_cast<const T*>(p.q) + ((int64_t)b * p.Lq + (active ? qi : 0)) * p.ldq + h * DH;
  const T* kbase = static_cast<const T*>(p.k) + (int64_t)b * p.Lk * p.ldk + h * DH;
  const T* vbase = static_cast<const T*>(p.v) + (int64_t)b * p.Lk * p.ldv + h * DH;
  const float scale = rsqrtf((float)DH);
  float q[DH], acc[DH];
#pragma unroll
  for (int i = 0; i < DH; ++i) { q[i] = to_f(qp[i]) * scale; acc[i] = 0.f; }
  float mrun = -INFINITY, lrun = 0.f;
  // causal: this CTA never needs keys beyond its last query
  const int k_end = p.causal ? min(p.Lk, blockIdx.x * QT + QT) : p.Lk;
  for (int k0 = 0; k0 < k_end; k0 += KT) {
    const int kn = min(KT, k_end - k0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < KT * DH; idx += QT) {
      const int r = idx / DH, c = idx % DH;
      float kv = 0.f, vv = 0.f;
      if (r < kn) {
        kv = to_f(kbase[(int64_t)(k0 + r) * p.ldk + c]);
        vv = to_f(vbase[(int64_t)(k0 + r) * p.ldv + c]);
      }
      Ks[r][c] = kv; Vs[r][c] = vv;
    }
    for (int r = threadIdx.x; r < KT; r += QT)
      Ms[r] = (r >= kn) || (p.key_pad && p.key_pad[(int64_t)b * p.Lk + k0 + r]);
    __syncthreads();
    if (!active) continue;
    for (int j0 = 0; j0 < kn; j0 += 4) {
      float s[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + u;                               // j < KT always (KT % 4 == 0)
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < DH; c += 4) {
          const float4 kk = *reinterpret_cast<const float4*>(&Ks[j][c]);
          d = fmaf(q[c], kk.x, d); d = fmaf(q[c + 1], kk.y, d);
          d = fmaf(q[c + 2], kk.z, d); d = fmaf(q[c + 3], kk.w, d);
        }
        const bool masked = Ms[j] || (p.causal && (k0 + j) > qi);
        s[u] = masked ? -INFINITY : d;
      }
      const float cmax = fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3]));
      if (cmax == -INFINITY) continue;                      // whole chunk masked for this query
      const float mnew = fmaxf(mrun, cmax);
      const float corr = __expf(mrun - mnew);               // exp(-inf) = 0 on the first chunk
      lrun *= corr;
#pragma unroll
      for (int c = 0; c < DH; ++c) acc[c] *= corr;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float pr = __expf(s[u] - mnew);               // masked -> 0
        lrun += pr;
        const int j = j0 + u;
#pragma unroll
        for (int c = 0; c < DH; c += 4) {
          const float4 vv = *reinterpret_cast<const float4*>(&Vs[j][c]);
          acc[c] = fmaf(pr, vv.x, acc[c]); acc[c + 1] = fmaf(pr, vv.y, acc[c + 1]);
          acc[c + 2] = fmaf(pr, vv.z, acc[c + 2]); acc[c + 3] = fmaf(pr, vv.w, acc[c + 3]);
        }
      }
      mrun = mnew;
    }
  }
  if (!active) return;
  const float inv = 1.f / lrun;                             // all keys masked -> NaN, as the reference
  T* op = static_cast<T*>(p.out) + ((int64_t)b * p.Lq + qi) * p.ldo + h * DH;
#pragma unroll
  for (int c = 0; c < DH; ++c) op[c] = from_f<T>(acc[c] * inv);
}

// Head-averaged probabilities [B,Lq,Lk] (need_weights=True path, average_attn_weights=True):
// one warp per (b, query); lanes stride over keys; Lk <= 32*MAXK.
template <typename T, int DH>
__global__ void attention_probs_kernel(const AttnParams p) {
  constexpr int MAXK = 16;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= p.B * p.Lq) return;
  const int b = warp / p.Lq, qi = warp % p.Lq;
  const float scale = rsqrtf((float)DH);
  float avg[MAXK];
#pragma unroll
  for (int i = 0; i < MAXK; ++i) avg[i] = 0.f;
  for (int h = 0; h < p.H; ++h) {
    const T* qp = static_cast<const T*>(p.q) + ((int64_t)b * p.Lq + qi) * p.ldq + h * DH;
    float q[DH];
#pragma unroll
    for (int c = 0; c < DH; ++c) q[c] = to_f(qp[c]) * scale;
    float s[MAXK], mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < MAXK; ++i) {
      const int j = i * 32 + lane;
      s[i] = -INFINITY;
      if (j < p.Lk && !(p.key_pad && p.key_pad[(int64_t)b * p.Lk + j]) && !(p.causal && j > qi)) {
        const T* kp = static_cast<const T*>(p.k) + ((int64_t)b * p.Lk + j) * p.ldk + h * DH;
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < DH; ++c) d = fmaf(q[c], to_f(kp[c]), d);
        s[i] = d;
      }
      mx = fmaxf(mx, s[i]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXK; ++i) { s[i] = __expf(s[i] - mx); sum += s[i]; }
    sum = warp_sum(sum);
    const float inv = 1.f / (sum * p.H);
#pragma unroll
    for (int i = 0; i < MAXK; ++i) avg[i] = fmaf(s[i], inv, avg[i]);
  }
  float* out = p.probs + ((int64_t)b * p.Lq + qi) * p.Lk;
#pragma unroll
  for (int i = 0; i < MAXK; ++i) {
    const int j = i * 32 + lane;
    if (j < p.Lk) out[j] = avg[i];
  }
}

}  // namespace

template <typename T>
void attention(const AttnParams& p, cudaStream_t s) {
  RETR_CHECK(p.B > 0 && p.Lq > 0 && p.Lk > 0 && p.H > 0, "empty attention");
  const int64_t d_model = p.ldo;  // only used for the head-dim check below
  (void)d_model;
  dim3 grid((unsigned)cdiv(p.Lq, QT), (unsigned)p.H, (unsigned)p.B);
  const int dh = (int)(p.ldo / p.H);
  RETR_CHECK(dh == 32 || dh == 64, "head dim must be 32 or 64 (ldo = n_heads * head_dim)");
  if (p.out && std::is_same<T, bf16>::value && getenv("RETR_ATTN_SIMT") == nullptr) {
    // RETR_ATTENTION_TC=1: the tcgen05 / TMEM kernel (attention_tc.cu); default: the warp-level mma.sync flash kernel,
    // which is faster at head dim 32 (measured)
    const char* tce = getenv("RETR_ATTENTION_TC");
    const bool tc = tce && atoi(tce) != 0;
    if (tc && attention_tc_supported(p, dh)) attention_tc_bf16(p, s);
    else flash_attention_bf16(p, dh, s);               // tensor-core path (attention_mma.cu)
  } else if (p.out) {
    if (dh == 32) attention_kernel<T, 32><<<grid, QT, 0, s>>>(p);
    else attention_kernel<T, 64><<<grid, QT, 0, s>>>(p);
    LAUNCH_OK();
  }
  if (p.probs) {
    RETR_CHECK(p.Lk <= 512, "probabilities output supports Lk <= 512");
    const int64_t warps = (int64_t)p.B * p.Lq;
    if (dh == 32) attention_probs_kernel<T, 32><<<(unsigned)cdiv(warps * 32, 128), 128, 0, s>>>(p);
    else attention_probs_kernel<T, 64><<<(unsigned)cdiv(warps * 32, 128), 128, 0, s>>>(p);
    LAUNCH_OK();
  }
}
template void attention<float>(const AttnParams&, cudaStream_t);
template void attention<bf16>(const AttnParams&, cudaStream_t);

}  // namespace retr
