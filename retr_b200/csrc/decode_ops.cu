// Kernels of the KV-cached greedy loop that replaces eval_utils/decode.py:68-79's full-prefix
// recompute: single-query attention over the self / cross K-V caches and the per-step token
// bookkeeping.  The step index lives in device memory so one captured CUDA graph serves every step.
#include "gemm.h"
#include "ops.h"
#include "sm100_ptx.cuh"

namespace retr {
namespace {
RETR_TRACE_DECL

constexpr int DA_WARPS = 4;
constexpr int DA_MAX_KEYS = 1024;       // the sine table bounds the encoder length at 1024

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

// One CTA per (sample, head).  The K and V blocks of a (b, h) pair are CONTIGUOUS in the head-major
// caches ([B][H][rows][DH]), so one thread streams them into shared memory with two bulk-async copies
// (cp.async.bulk, completion on an mbarrier) — the whole block is in flight at once, 4 CTAs per SM —
// and the 4 warps then score keys from shared memory: LPK lanes cooperate on one key (16-byte slice
// each, conflict-free), online softmax per lane group, merged by shuffles and one smem pass.
// (Persistent variants measured slower inside the decode graph on B200 at B=256 — a CTA-wide ring with one producer
// warp: 68-74 ms decode vs 62.5; one warp per unit with private rings: 64.5-83 — each needs ~100 KB per SM in flight
// AND several units' dependent load->compute chains overlapped, which is exactly what 4 resident one-unit CTAs give.
// Also measured and dropped: the eight head-CTAs of a sample as a cluster that finishes the out-projection + residual
// itself (DSMEM exchange of the 256 partial sums) — removes a launch but costs +6.7 ms per decode: gang scheduling and
// every CTA held until the slowest head of its sample arrives.)
template <typename T, int DH>
__global__ void __launch_bounds__(DA_WARPS * 32)
decode_attention_kernel(const T* __restrict__ q, int64_t ldq, const T* __restrict__ new_kv, int64_t ld_new,
                        T* __restrict__ kcache, T* __restrict__ vcache, int cache_rows, T* __restrict__ out,
                        int64_t ldo, int d, int H, int n_keys, const int* __restrict__ step_dyn,
                        const uint8_t* __restrict__ key_pad, int64_t ld_pad, const int* __restrict__ n_keys_per) {
  using V = Vec16<T>;
  constexpr int VN = V::N, LPK = DH / VN, KPI = 32 / LPK;
  extern __shared__ __align__(128) uint8_t da_smem[];
  T* Ks = reinterpret_cast<T*>(da_smem);
  T* Vs = Ks + (size_t)cache_rows * DH;
  __shared__ __align__(8) uint64_t bar[2];                  // K landed, V landed
  __shared__ float sm_m[DA_WARPS], sm_l[DA_WARPS], sm_acc[DA_WARPS][DH];
  __shared__ float sc[DA_MAX_KEYS];                         // scores (masked keys: -inf)
  const int bh = blockIdx.x, b = bh / H, h = bh - b * H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = lane % LPK, grp = lane / LPK;
  const int e0 = h * DH + sub * VN;                         // this lane's slice of the model dim
  T* kg = kcache + (size_t)bh * cache_rows * DH;
  T* vg = vcache + (size_t)bh * cache_rows * DH;
  TRACE_BEGIN();
  pdl_launch_dependents();
  if (new_kv) pdl_wait();                                   // self-attention: the new K/V row is the previous kernel's output
  // cross-attention: the caches (and the per-sample key counts of a compacted cache) are older than the previous
  // kernel, so the copies are issued before griddepcontrol.wait; only q depends on it
  const int n_cached = new_kv ? *step_dyn : (n_keys_per ? n_keys_per[b] : n_keys);
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar[0], 1);
    ptx::mbar_init(&bar[1], 1);
    ptx::fence_barrier_init();
    if (n_cached > 0) {
      const uint32_t bytes = (uint32_t)n_cached * DH * sizeof(T);
      ptx::mbar_arrive_expect_tx(&bar[0], bytes);
      ptx::mbar_arrive_expect_tx(&bar[1], bytes);
      if (new_kv) {
        ptx::bulk_load(Ks, kg, bytes, &bar[0]);
        ptx::bulk_load(Vs, vg, bytes, &bar[1]);
      } else {
        // cross-attention K/V: 0.6 GB per step over all layers, each byte read once per step — stream it past the
        // L2 so the ~35 MB of decoder weights and activations every step re-reads stay resident
        const uint64_t pol = ptx::policy_evict_first();
        ptx::bulk_load_hint(Ks, kg, bytes, &bar[0], pol);
        ptx::bulk_load_hint(Vs, vg, bytes, &bar[1], pol);
      }
    }
  }
  // small operands: one global round trip, overlapped with the bulk copies.  The padding bytes go
  // straight into the score array as 0 / -inf.
  if (key_pad) {
    const uint8_t* pad = key_pad + (int64_t)b * ld_pad;
    for (int j = threadIdx.x; j < n_cached; j += DA_WARPS * 32) sc[j] = pad[j] ? -INFINITY : 0.f;
  } else {
    for (int j = threadIdx.x; j < n_cached; j += DA_WARPS * 32) sc[j] = 0.f;
  }
  if (!new_kv) pdl_wait();
  float qv[VN];
  V::load(q + (int64_t)b * ldq + e0, qv);
  const float scale = rsqrtf((float)DH);
#pragma unroll
  for (int i = 0; i < VN; ++i) qv[i] *= scale;
  float s_new = -INFINITY, vn[VN];
  if (new_kv) {
    // self-attention: this step's key/value come from the projection output and are appended
    float kn[VN];
    V::load(new_kv + (int64_t)b * ld_new + e0, kn);
    V::load(new_kv + (int64_t)b * ld_new + d + e0, vn);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VN; ++i) s = fmaf(qv[i], kn[i], s);
#pragma unroll
    for (int o = LPK / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);   // all lanes converged
    s_new = s;
    if (warp == 0 && grp == 0) {
      V::store(kg + (size_t)n_cached * DH + sub * VN, kn);
      V::store(vg + (size_t)n_cached * DH + sub * VN, vn);
    }
  }
  __syncthreads();                                          // barrier init + mask values visible
  // pass 1: scores (independent per key: no serial softmax chain), running max per lane
  float mx = s_new;
  if (n_cached > 0) ptx::mbar_wait(&bar[0], 0);
  for (int base = warp * KPI; base < n_cached; base += DA_WARPS * KPI) {     // warp-uniform trip count
    const int j = base + grp;
    float s = 0.f;
    if (j < n_cached) {
      float kf[VN];
      V::load(Ks + (size_t)j * DH + sub * VN, kf);
#pragma unroll
      for (int i = 0; i < VN; ++i) s = fmaf(qv[i], kf[i], s);
    }
#pragma unroll
    for (int o = LPK / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (j < n_cached) {
      s += sc[j];                                           // -inf for padded keys
      if (sub == 0) sc[j] = s;
      mx = fmaxf(mx, s);
    }
  }
  mx = warp_max(mx);
  if (lane == 0) sm_m[warp] = mx;
  __syncthreads();                                          // scores + per-warp maxima visible
  float M = sm_m[0];
#pragma unroll
  for (int w = 1; w < DA_WARPS; ++w) M = fmaxf(M, sm_m[w]);
  // pass 2: p = exp(s - M), l += p, acc += p v   (no rescaling needed)
  float l = 0.f, acc[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) acc[i] = 0.f;
  if (new_kv && warp == 0 && grp == 0) {
    const float pr = __expf(s_new - M);
    l = pr;
#pragma unroll
    for (int i = 0; i < VN; ++i) acc[i] = pr * vn[i];
  }
  if (n_cached > 0) ptx::mbar_wait(&bar[1], 0);
  for (int base = warp * KPI; base < n_cached; base += DA_WARPS * KPI) {
    const int j = base + grp;
    if (j < n_cached) {
      const float pr = __expf(sc[j] - M);
      float vf[VN];
      V::load(Vs + (size_t)j * DH + sub * VN, vf);
      l += pr;
#pragma unroll
      for (int i = 0; i < VN; ++i) acc[i] = fmaf(pr, vf[i], acc[i]);
    }
  }
  // sum the KPI key groups of the warp, then the warps
#pragma unroll
  for (int o = LPK; o < 32; o <<= 1) {
    l += __shfl_xor_sync(0xffffffffu, l, o);
#pragma unroll
    for (int i = 0; i < VN; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
  }
  if (grp == 0) {
    if (sub == 0) sm_l[warp] = l;
#pragma unroll
    for (int i = 0; i < VN; ++i) sm_acc[warp][sub * VN + i] = acc[i];
  }
  __syncthreads();
  if (threadIdx.x < DH) {
    const int c = threadIdx.x;
    float ll = 0.f, aa = 0.f;
#pragma unroll
    for (int w = 0; w < DA_WARPS; ++w) { ll += sm_l[w]; aa += sm_acc[w][c]; }
    out[(int64_t)b * ldo + h * DH + c] = from_f<T>(aa / ll);
  }
  TRACE_END(new_kv ? 2 : 3, out);
}

// [B*rows, 2d] rows of [K(d) | V(d)]  ->  head-major K[B][H][rows][DH], V[B][H][rows][DH]; 16-byte chunks
// Padded keys contribute exp(-inf) = 0 to every cross-attention the decoder ever computes, so they are dropped ONCE,
// when the caches are built: dst_row[b*rows + r] is the position of key r among sample b's unpadded keys (-1 = padded),
// n_valid[b] their number.  The decode kernels then read n_valid[b] dense rows and need no mask — about a quarter less
// HBM traffic per step on pad-to-square inputs, same result.
__global__ void key_compaction_kernel(const uint8_t* __restrict__ pad, int B, int S, int* __restrict__ dst_row,
                                      int* __restrict__ n_valid) {
  const int b = blockIdx.x, lane = threadIdx.x;             // one warp per sample
  int base = 0;
  for (int j0 = 0; j0 < S; j0 += 32) {
    const int j = j0 + lane;
    const bool valid = j < S && pad[(int64_t)b * S + j] == 0;
    const unsigned m = __ballot_sync(0xffffffffu, valid);
    if (j < S) dst_row[(int64_t)b * S + j] = valid ? base + __popc(m & ((1u << lane) - 1u)) : -1;
    base += __popc(m);
  }
  if (lane == 0) n_valid[b] = base;
}

// ---- packed encoder (engine.cu: encode): only the unpadded tokens of every sample go through the encoder layers ----
// exclusive prefix sum of n_valid over the batch: off[b] = first packed row of sample b, off[B] = total rows
__global__ void seq_offsets_kernel(const int* __restrict__ n_valid, int B, int* __restrict__ off) {
  __shared__ int part[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int carry = 0;
  for (int b0 = 0; b0 < B; b0 += (int)blockDim.x) {
    const int b = b0 + tid;
    const int v = b < B ? n_valid[b] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) part[warp] = x;
    __syncthreads();
    if (warp == 0) {
      int w = lane < (int)(blockDim.x >> 5) ? part[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
      part[lane] = w;
    }
    __syncthreads();
    const int before = carry + (warp ? part[warp - 1] : 0) + x - v;
    if (b < B) off[b] = before;
    carry += part[(blockDim.x >> 5) - 1];
    __syncthreads();
  }
  if (tid == 0) off[B] = carry;
}

// x [B*S, d] fp32 -> xp [off[B], d]: row off[b] + dst_row[b,s] = row (b, s) for every unpadded token; row_pos = s
__global__ void pack_rows_kernel(const float* __restrict__ x, float* __restrict__ xp, const int* __restrict__ dst_row,
                                 const int* __restrict__ off, int* __restrict__ row_pos, int64_t rows, int S, int d) {
  const int64_t br = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (br >= rows) return;
  const int r = dst_row[br];
  if (r < 0) return;
  const int b = (int)(br / S);
  const int64_t to = (int64_t)off[b] + r;
  const float4* src = reinterpret_cast<const float4*>(x + br * d);
  float4* dst = reinterpret_cast<float4*>(xp + to * d);
  for (int i = lane; i < d / 4; i += 32) dst[i] = src[i];
  if (lane == 0) row_pos[to] = (int)(br - (int64_t)b * S);
}

// packed rows -> [B*S, d] (padded rows zero): what the full-sequence consumers of the encoder memory read
template <typename T>
__global__ void unpack_rows_kernel(const T* __restrict__ xp, T* __restrict__ x, const int* __restrict__ dst_row,
                                   const int* __restrict__ off, int64_t rows, int S, int d) {
  constexpr int VN = 16 / sizeof(T);
  const int64_t br = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (br >= rows) return;
  const int r = dst_row[br];
  uint4* dst = reinterpret_cast<uint4*>(x + br * d);
  if (r < 0) {
    for (int i = lane; i < d / VN; i += 32) dst[i] = make_uint4(0u, 0u, 0u, 0u);
    return;
  }
  const uint4* src = reinterpret_cast<const uint4*>(xp + ((int64_t)off[br / S] + r) * d);
  for (int i = lane; i < d / VN; i += 32) dst[i] = src[i];
}

template <typename T>
__global__ void kv_to_head_major_kernel(const T* __restrict__ kv, T* __restrict__ kh, T* __restrict__ vh, int B,
                                        int rows, int H, int dh, const int* __restrict__ dst_row) {
  constexpr int VN = 16 / sizeof(T);
  const int cpr = 2 * H * dh / VN;                          // chunks per source row
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)B * rows * cpr) return;
  const int c = (int)(idx % cpr);
  const int64_t br = idx / cpr;
  int r = (int)(br % rows);
  const int b = (int)(br / rows);
  if (dst_row) {
    r = dst_row[br];
    if (r < 0) return;
  }
  const int e = c * VN, isv = e >= H * dh, eh = isv ? e - H * dh : e;
  const int h = eh / dh, w = eh % dh;
  const uint4 val = *reinterpret_cast<const uint4*>(kv + br * (2 * H * dh) + e);
  T* dst = (isv ? vh : kh) + (((size_t)b * H + h) * rows + r) * dh + w;
  *reinterpret_cast<uint4*>(dst) = val;
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
  TRACE_BEGIN();
  pdl_wait();
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
  TRACE_END(6, 0);
}

}  // namespace

#ifdef RETR_TRACE
void decode_ops_set_trace(unsigned long long* buf) { CUDA_OK(cudaMemcpyToSymbol(trace_buf_, &buf, sizeof(buf))); }
#endif

template <typename T>
void decode_attention(const T* q, int64_t ldq, const T* new_kv, int64_t ld_new, T* kcache, T* vcache, int cache_rows,
                      T* out, int64_t ldo, int B, int H, int d, int n_keys, const int* step_dyn,
                      const uint8_t* key_pad, int64_t ld_pad, cudaStream_t s, const int* n_keys_per) {
  const int dh = d / H;
  RETR_CHECK(dh == 32 || dh == 64, "head dim must be 32 or 64");
  RETR_CHECK(!new_kv || step_dyn, "self-attention needs the step counter");
  const size_t smem = (size_t)2 * cache_rows * dh * sizeof(T);        // K block + V block
  RETR_CHECK(smem <= 220 * 1024 && cache_rows <= DA_MAX_KEYS, "K/V block does not fit in shared memory");
  auto kern = dh == 32 ? decode_attention_kernel<T, 32> : decode_attention_kernel<T, 64>;
  launch_pdl(kern, dim3((unsigned)(B * H)), dim3(DA_WARPS * 32), smem, s, q, ldq, new_kv, ld_new, kcache, vcache,
             cache_rows, out, ldo, d, H, n_keys, step_dyn, key_pad, ld_pad, n_keys_per);
}
template <typename T>
void decode_attention_configure(int max_rows, int dh) {
  auto kern = dh == 32 ? decode_attention_kernel<T, 32> : decode_attention_kernel<T, 64>;
  CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * max_rows * dh * (int)sizeof(T)));
  CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CUDA_OK(cudaFuncSetAttribute(greedy_update_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
}
template void decode_attention_configure<float>(int, int);
template void decode_attention_configure<bf16>(int, int);
template void decode_attention<float>(const float*, int64_t, const float*, int64_t, float*, float*, int, float*, int64_t,
                                      int, int, int, int, const int*, const uint8_t*, int64_t, cudaStream_t, const int*);
template void decode_attention<bf16>(const bf16*, int64_t, const bf16*, int64_t, bf16*, bf16*, int, bf16*, int64_t, int,
                                     int, int, int, const int*, const uint8_t*, int64_t, cudaStream_t, const int*);

void key_compaction(const uint8_t* pad, int B, int S, int* dst_row, int* n_valid, cudaStream_t s) {
  key_compaction_kernel<<<(unsigned)B, 32, 0, s>>>(pad, B, S, dst_row, n_valid);
  LAUNCH_OK();
}

template <typename T>
void kv_to_head_major(const T* kv, T* kh, T* vh, int B, int rows, int H, int dh, cudaStream_t s, const int* dst_row) {
  const int64_t n = (int64_t)B * rows * (2 * H * dh / (16 / (int)sizeof(T)));
  kv_to_head_major_kernel<T><<<(unsigned)cdiv(n, 256), 256, 0, s>>>(kv, kh, vh, B, rows, H, dh, dst_row);
  LAUNCH_OK();
}
template void kv_to_head_major<float>(const float*, float*, float*, int, int, int, int, cudaStream_t, const int*);
template void kv_to_head_major<bf16>(const bf16*, bf16*, bf16*, int, int, int, int, cudaStream_t, const int*);

void seq_offsets(const int* n_valid, int B, int* off, cudaStream_t s) {
  seq_offsets_kernel<<<1, 256, 0, s>>>(n_valid, B, off);
  LAUNCH_OK();
}
void pack_rows(const float* x, float* xp, const int* dst_row, const int* off, int* row_pos, int B, int S, int d,
               cudaStream_t s) {
  RETR_CHECK(d % 4 == 0, "pack_rows: d % 4");
  const int64_t rows = (int64_t)B * S;
  CUDA_OK(cudaMemsetAsync(row_pos, 0, rows * 4, s));             // rows past the packed total index position 0
  pack_rows_kernel<<<(unsigned)cdiv(rows, 8), 256, 0, s>>>(x, xp, dst_row, off, row_pos, rows, S, d);
  LAUNCH_OK();
}
template <typename T>
void unpack_rows(const T* xp, T* x, const int* dst_row, const int* off, int B, int S, int d, cudaStream_t s) {
  RETR_CHECK(d % (16 / (int)sizeof(T)) == 0, "unpack_rows: 16-byte rows");
  const int64_t rows = (int64_t)B * S;
  unpack_rows_kernel<T><<<(unsigned)cdiv(rows, 8), 256, 0, s>>>(xp, x, dst_row, off, rows, S, d);
  LAUNCH_OK();
}
template void unpack_rows<float>(const float*, float*, const int*, const int*, int, int, int, cudaStream_t);
template void unpack_rows<bf16>(const bf16*, bf16*, const int*, const int*, int, int, int, cudaStream_t);

void greedy_update(unsigned long long* argmax_keys, int64_t* caption, int max_len, int B, int eos, int* step,
                   int* done, uint8_t* finished, cudaStream_t s, bool allow_pdl) {
  if (allow_pdl) {
    launch_pdl(greedy_update_kernel, dim3(1), dim3(1024), 0, s, argmax_keys, caption, max_len, B, eos, step, done, finished);
  } else {      // joins several parallel branches: a plain (full) dependency on every one of them
    greedy_update_kernel<<<1, 1024, 0, s>>>(argmax_keys, caption, max_len, B, eos, step, done, finished);
    LAUNCH_OK();
  }
}

void greedy_init(int64_t* caption, int max_len, int B, int bos, int* step, int* done, uint8_t* finished,
                 unsigned long long* argmax_keys, cudaStream_t s) {
  greedy_init_kernel<<<(unsigned)cdiv((int64_t)B * max_len, 256), 256, 0, s>>>(caption, max_len, B, bos, step, done,
                                                                              finished, argmax_keys);
  LAUNCH_OK();
}

}  // namespace retr
