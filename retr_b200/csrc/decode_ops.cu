// Kernels of the KV-cached greedy loop that replaces eval_utils/decode.py:68-79's full-prefix
// recompute: single-query attention over the self / cross K-V caches and the per-step token
// bookkeeping.  The step index lives in device memory so one captured CUDA graph serves every step.
#include "ops.h"

namespace retr {
namespace {

constexpr int DA_WARPS = 4;

template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  typedef float4 Raw;
  __device__ static Raw load_raw(const float* p) { return *reinterpret_cast<const float4*>(p); }
  __device__ static void unpack(const Raw& t, float (&o)[4]) { o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w; }
  __device__ static void load(const float* p, float (&o)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
  }
  __device__ static void store(float* p, const float (&o)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
  }
};
template <> struct Vec16<bf16> {
  static constexpr int N = 8;
  typedef uint4 Raw;
  __device__ static Raw load_raw(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
  __device__ static void unpack(const Raw& t, float (&o)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 f = __bfloat1622float2(h[i]); o[2 * i] = f.x; o[2 * i + 1] = f.y; }
  }
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

// merge two partial softmax states (m, l, acc) -> into the first
template <int N>
__device__ __forceinline__ void merge_state(float& m, float& l, float (&acc)[N], float m2, float l2,
                                            const float (&acc2)[N]) {
  const float mn = fmaxf(m, m2);
  if (mn == -INFINITY) return;
  const float a = __expf(m - mn), b = __expf(m2 - mn);
  l = l * a + l2 * b;
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] = acc[i] * a + acc2[i] * b;
  m = mn;
}

// One warp per (sample, head): no shared memory, no block barrier.  LPK lanes cooperate on one key
// (each holds a 16-byte slice of the head dim); a warp scores 32/LPK keys per load instruction and keeps
// UNROLL instructions (K and V) in flight.  The K/V rows of a (b, h) pair are 2*DH-byte pieces at pitch
// ld_cache; the 8 warps of neighbouring heads read the neighbouring pieces of the same rows.
template <typename T, int DH>
__global__ void __launch_bounds__(DA_WARPS * 32, 4)
decode_attention_kernel(const T* __restrict__ q, int64_t ldq, const T* __restrict__ new_kv, int64_t ld_new,
                        T* __restrict__ cache, int64_t ld_cache, int64_t cache_bp, T* __restrict__ out, int64_t ldo,
                        int d, int H, int n_pairs, int n_keys, const int* __restrict__ step_dyn,
                        const uint8_t* __restrict__ key_pad, int64_t ld_pad) {
  using V = Vec16<T>;
  constexpr int VN = V::N, LPK = DH / VN, KPI = 32 / LPK, UNROLL = sizeof(T) == 2 ? 8 : 4;
  const int gw = blockIdx.x * DA_WARPS + (threadIdx.x >> 5);
  if (gw >= n_pairs) return;                                // warp-uniform
  const int b = gw / H, h = gw - b * H;
  const int lane = threadIdx.x & 31;
  const int sub = lane % LPK, grp = lane / LPK;
  const int e0 = h * DH + sub * VN;                         // this lane's slice of the model dim
  T* kc = cache + (int64_t)b * cache_bp + e0;               // K part; V part at + d
  float qv[VN];
  V::load(q + (int64_t)b * ldq + e0, qv);
  const float scale = rsqrtf((float)DH);
#pragma unroll
  for (int i = 0; i < VN; ++i) qv[i] *= scale;

  int n_cached = n_keys;
  float m = -INFINITY, l = 0.f, acc[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) acc[i] = 0.f;

  if (new_kv) {
    // self-attention: this step's key/value come from the projection output and are appended
    const int step = *step_dyn;
    n_cached = step;
    float kn[VN], vn[VN];
    V::load(new_kv + (int64_t)b * ld_new + e0, kn);
    V::load(new_kv + (int64_t)b * ld_new + d + e0, vn);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VN; ++i) s = fmaf(qv[i], kn[i], s);
#pragma unroll
    for (int o = LPK / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);   // all lanes converged
    if (grp == 0) {
      V::store(kc + (int64_t)step * ld_cache, kn);
      V::store(kc + (int64_t)step * ld_cache + d, vn);
      m = s; l = 1.f;
#pragma unroll
      for (int i = 0; i < VN; ++i) acc[i] = vn[i];
    }
  }

  const uint8_t* pad = key_pad ? key_pad + (int64_t)b * ld_pad : nullptr;
  for (int base = 0; base < n_cached; base += KPI * UNROLL) {      // warp-uniform trip count
    typename V::Raw kraw[UNROLL], vraw[UNROLL];
    bool ok[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int j = base + u * KPI + grp;
      ok[u] = j < n_cached && !(pad && pad[j]);
      if (ok[u]) {
        kraw[u] = V::load_raw(kc + (int64_t)j * ld_cache);
        vraw[u] = V::load_raw(kc + (int64_t)j * ld_cache + d);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      float s = 0.f;
      float kf[VN], vf[VN];
      if (ok[u]) {
        V::unpack(kraw[u], kf);
        V::unpack(vraw[u], vf);
#pragma unroll
        for (int i = 0; i < VN; ++i) s = fmaf(qv[i], kf[i], s);
      }
      // the LPK lanes of a group share j, hence ok[u]: the segment shuffle runs converged
#pragma unroll
      for (int o = LPK / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (ok[u]) {
        const float mn = fmaxf(m, s);
        const float corr = __expf(m - mn), pr = __expf(s - mn);
        l = l * corr + pr;
#pragma unroll
        for (int i = 0; i < VN; ++i) acc[i] = fmaf(pr, vf[i], acc[i] * corr);
        m = mn;
      }
    }
  }
  // merge the KPI key groups of the warp
#pragma unroll
  for (int o = LPK; o < 32; o <<= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
    float a2[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) a2[i] = __shfl_xor_sync(0xffffffffu, acc[i], o);
    merge_state<VN>(m, l, acc, m2, l2, a2);
  }
  if (grp == 0) {
    const float inv = 1.f / l;
#pragma unroll
    for (int i = 0; i < VN; ++i) acc[i] *= inv;
    V::store(out + (int64_t)b * ldo + e0, acc);
  }
}

__global__ void greedy_init_kernel(int64_t* caption, int max_len, int B, int bos, int* step, int* done,
                                   uint8_t* finished, unsigned long long* keys) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < B * max_len) caption[idx] = (idx % max_len == 0) ? bos : 0;
  if (idx < B) { finished[idx] = 0; keys[idx] = 0ull; }
  if (idx == 0) { *step = 0; *done = 0; }
}

// eval_utils/decode.py:70-79 for step i = *step: ids = argmax; finished |= ids == eos; if every
// row is finished return WITHOUT writing this step's tokens; else caption[:, i+1] = ids.
__global__ void greedy_update_kernel(unsigned long long* keys, int64_t* caption, int max_len, int B, int eos,
                                     int* step, int* done, uint8_t* finished) {
  __shared__ int unfinished;
  if (*done) return;
  const int i = *step;
  if (threadIdx.x == 0) unfinished = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const int tok = (int)unpack_argmax_index(keys[b]);
    const uint8_t fin = finished[b] | (uint8_t)(tok == eos);
    finished[b] = fin;
    if (!fin) unfinished = 1;
  }
  __syncthreads();
  const bool all_done = unfinished == 0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    if (!all_done) caption[(int64_t)b * max_len + i + 1] = (int64_t)unpack_argmax_index(keys[b]);
    keys[b] = 0ull;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *step = i + 1;                      // == number of model evaluations so far
    if (all_done || i + 1 >= max_len - 1) *done = 1;
  }
}

}  // namespace

template <typename T>
void decode_attention(const T* q, int64_t ldq, const T* new_kv, int64_t ld_new, T* cache, int64_t ld_cache,
                      int64_t cache_batch_pitch, T* out, int64_t ldo, int B, int H, int d, int n_keys,
                      const int* step_dyn, const uint8_t* key_pad, int64_t ld_pad, cudaStream_t s) {
  const int dh = d / H;
  RETR_CHECK(dh == 32 || dh == 64, "head dim must be 32 or 64");
  RETR_CHECK(!new_kv || step_dyn, "self-attention needs the step counter");
  const int pairs = B * H;
  const unsigned grid = (unsigned)cdiv(pairs, DA_WARPS);
  if (dh == 32)
    decode_attention_kernel<T, 32><<<grid, DA_WARPS * 32, 0, s>>>(q, ldq, new_kv, ld_new, cache, ld_cache,
                                                                   cache_batch_pitch, out, ldo, d, H, pairs, n_keys,
                                                                   step_dyn, key_pad, ld_pad);
  else
    decode_attention_kernel<T, 64><<<grid, DA_WARPS * 32, 0, s>>>(q, ldq, new_kv, ld_new, cache, ld_cache,
                                                                   cache_batch_pitch, out, ldo, d, H, pairs, n_keys,
                                                                   step_dyn, key_pad, ld_pad);
  LAUNCH_OK();
}
template void decode_attention<float>(const float*, int64_t, const float*, int64_t, float*, int64_t, int64_t, float*,
                                      int64_t, int, int, int, int, const int*, const uint8_t*, int64_t, cudaStream_t);
template void decode_attention<bf16>(const bf16*, int64_t, const bf16*, int64_t, bf16*, int64_t, int64_t, bf16*,
                                     int64_t, int, int, int, int, const int*, const uint8_t*, int64_t, cudaStream_t);

void greedy_update(unsigned long long* argmax_keys, int64_t* caption, int max_len, int B, int eos, int* step,
                   int* done, uint8_t* finished, cudaStream_t s) {
  greedy_update_kernel<<<1, 1024, 0, s>>>(argmax_keys, caption, max_len, B, eos, step, done, finished);
  LAUNCH_OK();
}

void greedy_init(int64_t* caption, int max_len, int B, int bos, int* step, int* done, uint8_t* finished,
                 unsigned long long* argmax_keys, cudaStream_t s) {
  greedy_init_kernel<<<(unsigned)cdiv((int64_t)B * max_len, 256), 256, 0, s>>>(caption, max_len, B, bos, step, done,
                                                                              finished, argmax_keys);
  LAUNCH_OK();
}

}  // namespace retr
