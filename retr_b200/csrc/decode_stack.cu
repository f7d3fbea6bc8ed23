// One KV-cached greedy step through the WHOLE decoder stack in ONE kernel (d_model 256, 8 heads of 32, d_ff 2048).
//
// Reference being replaced: models/ConcatTransformer.py:220-257 (decoder layer: pre-norm self-attention,
// cross-attention, feed-forward, each with a residual), :111-149 (layer loop + final LayerNorm),
// models/transformer_modules.py:113-129 (DecoderEmbeddings) and the first two layers of the MLP head
// (models/caption.py:161-174) evaluated for ONE new position per sample against K/V caches instead of the whole
// 128-long prefix (eval_utils/decode.py:68-79).
//
// Why one kernel: as a chain of launches a decode step was ~41 dependent kernels of 5-25 us for ~70 us of HBM time
// (profiles/r1_decode_timeline.txt).  Samples are independent through the whole stack, so a thread-block CLUSTER of 4
// CTAs takes 8 samples through all six layers with no grid-wide synchronisation at all — only cluster barriers:
//   * CTA r of the cluster owns heads 2r, 2r+1 of every attention and column slice [64r, 64r+64) of every projection:
//     the q/k/v projection gives CTA r exactly the q, k, v of its heads, so attention needs no exchange before it;
//   * every CTA keeps a full fp32 replica of the 8 residual rows in shared memory (LayerNorm needs whole rows); after
//     each sublayer the four 64-column slices are pushed to every peer through distributed shared memory;
//   * the FFN splits the hidden dimension: CTA r computes relu(W1[512r:512r+512] LN(y)) and its partial product with
//     W2[:, 512r:512r+512], pushes 64-column slices of the partial to the owning CTA, which sums them in a fixed order;
//   * GEMMs are one m16n8k16 row block (rows 8-15 are zero registers): weights stream global(L2) -> registers as
//     B fragments with 16-byte loads, double-buffered two n-tiles at a time (each weight byte is used exactly once
//     per CTA, so staging it in shared memory buys nothing); the k index is permuted identically for A and B so that
//     a thread's 16 bytes are 8 consecutive k; the loads of a GEMM are issued BEFORE the cluster barrier / LayerNorm
//     that precedes it;
//   * attention: K/V caches are stored in 4 KB chunks [K 32 keys x 32 | V 32 keys x 32] (bf16, 16-byte units XOR-
//     swizzled by the key index, written that way by the cache builders).  Warp w owns sample w and walks its two
//     heads in lock-step (two independent dependency chains), fed by a private 3-slot ring of 8 KB two-head blocks filled by
//     cp.async.bulk (one elected lane, completion on an mbarrier); the ring runs AHEAD across sublayers and layers
//     (the next cross-attention's chunks arrive while the projections run).  Scores and P.V run on the tensor cores
//     (q as row 0 of the A operand; ldmatrix / ldmatrix.trans on the swizzled chunk), online softmax per chunk:
//     math starts on the first chunk instead of waiting for a whole K block.
// Sizing: 8 samples x 4 CTAs because thread-block clusters of 8 with this much shared memory reach only 15
// co-resident clusters on a B200 (cudaOccupancyMaxActiveClusters), one short of the 16 that 256 samples need — the
// first version (16 samples per 8-CTA cluster) ran the benchmark batch in TWO waves, 500 us per step.  Clusters of 4
// pack the GPCs better: the batch runs as ONE wave of 32 clusters on 128 SMs.
#include "decode_stack.h"

#include "sm100_ptx.cuh"

namespace retr {
namespace {

constexpr int D = 256, DH = 32, NH = 8, CL = 4, HPC = NH / CL, R = DS_ROWS, THREADS = 256, NWARP = 8;
constexpr int SL = D / CL;                   // 64: columns of a projection owned by one CTA
constexpr int LDA = D + 32;                  // bf16 per A row in shared memory: 576 B, rows 16 banks apart
constexpr int FS = 512, LDH = FS + 32;       // FFN hidden slice per CTA
constexpr int MH = 512, LDM1 = MH + 32;      // MLP head hidden width
constexpr int SLOTS = 3, SLOT_BYTES = HPC * DS_CHUNK_ELEMS * 2;   // a slot = one 32-key chunk of BOTH heads (8 KB)
constexpr float LOG2E = 1.4426950408889634f;
static_assert(R == NWARP, "one sample per warp");

// -DDS_PROFILE: thread 0 of cluster 0 / CTA 0 accumulates SM cycles per phase and prints them at step 60
#ifdef DS_PROFILE
#define DS_T(i) do { if (prof_on) { const long long now_ = clock64(); prof[i] += now_ - prof_last; prof_last = now_; } } while (0)
#else
#define DS_T(i) do { } while (0)
#endif

struct Smem {
  uint8_t ring[NWARP][SLOTS][SLOT_BYTES];    // per-warp K/V chunk rings (192 KB)
  float Y[R][D];                             // residual stream rows (full replica in every CTA of the cluster)
  union {                                    // live in disjoint phases, separated by cluster barriers #4 / #6:
    float Ps[CL][R][SL];                     //   FFN partials of this CTA's 64 output columns, one slot per source CTA
    struct {                                 //   this CTA's heads: (scaled) query and the step's new key / value rows
      bf16 Qs[R][SL], Kn[R][SL], Vn[R][SL];
    };
  };
  bf16 Abuf[R][LDA];                         // LayerNorm output (A operand)
  bf16 Att[R][LDA];                          // attention output, heads gathered from the peers
  bf16 Hs[R][LDH];                           // FFN hidden slice; reused as M1[R][LDM1] by the head
  uint64_t bars[NWARP][SLOTS];
};
static_assert(LDM1 <= LDH, "M1 must fit in Hs");
static_assert(sizeof(Smem) <= 227 * 1024, "shared memory budget");

// ---- small PTX wrappers ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t map_peer(uint32_t local_addr, uint32_t rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  return remote;
}
__device__ __forceinline__ void st_peer_f2(uint32_t remote, float2 v) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1,%2};" ::"r"(remote), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void st_peer_b32(uint32_t remote, uint32_t v) {
  asm volatile("st.shared::cluster.b32 [%0], %1;" ::"r"(remote), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

// ---- GEMM pieces: 8 real rows (rows 8-15 of the m16 block are zero registers) -------------------------------------
// A fragments of row g over 256 consecutive k (8 blocks of 32).  Within a block thread (g, c) owns the 8 consecutive
// k [8c, 8c+8): k-step 0 uses elements {0,1 | 2,3}, k-step 1 {4,5 | 6,7} — the same assignment as the 16-byte weight
// loads, so the products pair up correctly (a dot product does not care in which order k is visited).
struct AFrag {
  uint4 x[8];
};
__device__ __forceinline__ void load_a(AFrag& a, const bf16* A, int lda, int lane) {
  const int g = lane >> 2, c = lane & 3;
#pragma unroll
  for (int kb = 0; kb < 8; ++kb) a.x[kb] = *reinterpret_cast<const uint4*>(A + (size_t)g * lda + kb * 32 + 8 * c);
}
// B fragments of NT n-tiles over 256 k.  The weights are stored in FRAGMENT ORDER (repack_frag_kernel, at weight
// load): for every 8-row tile and every 32-wide k block the 32 lanes' 16-byte pieces are contiguous, so one warp-wide
// load reads four whole 128-byte lines (the row-major layout touched eight lines per load and used half of each):
//   Wf[((n / 8) * (K / 32) + k / 32) * 256 + ((n % 8) * 4 + (k % 32) / 8) * 8 + k % 8] = W[n][k]
// wrow[j] = frag_ptr(Wf, tile_j, K / 32, first k block, lane)
template <int NT> struct WFrag {
  uint4 w[NT][8];
};
__device__ __forceinline__ const bf16* frag_ptr(const bf16* Wf, int tile, int kblocks, int kb0, int lane) {
  return Wf + ((size_t)tile * kblocks + kb0) * 256 + lane * 8;
}
template <int NT> __device__ __forceinline__ void load_w(WFrag<NT>& f, const bf16* const (&wrow)[NT]) {
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) f.w[j][kb] = ldg_stream(wrow[j] + kb * 256);
}
// acc[j] += A . W_j^T; k outer / tiles inner so the NT accumulator chains interleave; with one tile the even and odd
// k blocks go to two accumulators
template <int NT> __device__ __forceinline__ void mma_w(const AFrag& a, const WFrag<NT>& f, float (&acc)[NT][4]) {
  if constexpr (NT == 1) {
    float alt[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int kb = 0; kb < 8; kb += 2) {
      mma16816(acc[0], a.x[kb].x, 0u, a.x[kb].y, 0u, f.w[0][kb].x, f.w[0][kb].y);
      mma16816(alt, a.x[kb + 1].x, 0u, a.x[kb + 1].y, 0u, f.w[0][kb + 1].x, f.w[0][kb + 1].y);
      mma16816(acc[0], a.x[kb].z, 0u, a.x[kb].w, 0u, f.w[0][kb].z, f.w[0][kb].w);
      mma16816(alt, a.x[kb + 1].z, 0u, a.x[kb + 1].w, 0u, f.w[0][kb + 1].z, f.w[0][kb + 1].w);
    }
    acc[0][0] += alt[0]; acc[0][1] += alt[1];
  } else {
    // two accumulators per tile (even / odd k blocks): 2 NT independent chains of 8 dependent HMMAs instead of NT of 16 —
    // at M = 16 the GEMM phases are bound by the HMMA issue-to-result latency of those chains, not by throughput
    float alt[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) alt[j][0] = alt[j][1] = alt[j][2] = alt[j][3] = 0.f;
#pragma unroll
    for (int kb = 0; kb < 8; kb += 2) {
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        mma16816(acc[j], a.x[kb].x, 0u, a.x[kb].y, 0u, f.w[j][kb].x, f.w[j][kb].y);
        mma16816(alt[j], a.x[kb + 1].x, 0u, a.x[kb + 1].y, 0u, f.w[j][kb + 1].x, f.w[j][kb + 1].y);
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        mma16816(acc[j], a.x[kb].z, 0u, a.x[kb].w, 0u, f.w[j][kb].z, f.w[j][kb].w);
        mma16816(alt[j], a.x[kb + 1].z, 0u, a.x[kb + 1].w, 0u, f.w[j][kb + 1].z, f.w[j][kb + 1].w);
      }
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) { acc[j][0] += alt[j][0]; acc[j][1] += alt[j][1]; }
  }
}
template <int NT> __device__ __forceinline__ void zero_acc(float (&acc)[NT][4]) {
#pragma unroll
  for (int j = 0; j < NT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
}

// LayerNorm of replica row `warp` -> bf16 A operand
__device__ __forceinline__ void layernorm_row(const float (*Y)[D], bf16 (*A)[LDA], const float* __restrict__ gamma,
                                              const float* __restrict__ beta, float eps, int warp, int lane) {
  float v[8];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float4 t = *reinterpret_cast<const float4*>(&Y[warp][j * 128 + lane * 4]);
    v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += v[j];
  const float mean = warp_sum(s) * (1.f / D);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) { const float t = v[j] - mean; q = fmaf(t, t, q); }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / D) + eps);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const float4 ga = __ldg(reinterpret_cast<const float4*>(gamma + j * 128 + lane * 4));
    const float4 be = __ldg(reinterpret_cast<const float4*>(beta + j * 128 + lane * 4));
    uint2 pk;
    pk.x = pack_bf16((v[4 * j] - mean) * rstd * ga.x + be.x, (v[4 * j + 1] - mean) * rstd * ga.y + be.y);
    pk.y = pack_bf16((v[4 * j + 2] - mean) * rstd * ga.z + be.z, (v[4 * j + 3] - mean) * rstd * ga.w + be.w);
    *reinterpret_cast<uint2*>(&A[warp][j * 128 + lane * 4]) = pk;
  }
}

// ---- per-warp K/V chunk stream ---------------------------------------------------------------------------------------
// The caches hold, per (sample, CTA rank), 8 KB blocks [head A: K 32x32 | V 32x32][head B: K | V] per 32 keys, so one
// bulk copy brings a chunk of BOTH heads of the CTA.  The blocks a warp will consume during the whole kernel form one
// fixed sequence: for layer: [self block 0, 1, ..., cross block 0, 1, ...].  The issue cursor runs SLOTS blocks ahead
// of the consume cursor, across sublayer and layer boundaries.  Issuing costs the consumer ~250 cycles per copy (it
// was 460 with a proxy fence per copy and two copies per iteration), so there is one copy per iteration and the
// fence is kept only where the slot was WRITTEN through the generic proxy (the self-attention patch).
struct Stream {
  int issued, consumed;
  int il, iph, ic;                           // issue cursor: layer, phase (0 self / 1 cross), block
  int n_self, n_cross;                       // blocks in each phase
  int b;                                     // the warp's sample (clamped to B-1)
  uint32_t ring, bars;                       // shared-memory addresses of this warp's ring and barriers
};
__device__ __forceinline__ void stream_issue(const DecStackParams& p, Stream& st, int rank, int lane, uint64_t policy,
                                             bool fence) {
  if (st.il >= p.n_layers) return;
  const int slot = st.issued % SLOTS;
  if (lane == 0) {
    const DecStackLayer& L = p.L[st.il];
    const bf16* src = st.iph == 0
        ? L.self_kv + ((size_t)(st.b * CL + rank) * p.self_chunks + st.ic) * (HPC * DS_CHUNK_ELEMS)
        : L.cross_kv + ((size_t)(st.b * CL + rank) * p.cross_chunks + st.ic) * (HPC * DS_CHUNK_ELEMS);
    const uint32_t bar = st.bars + slot * 8, dst = st.ring + slot * SLOT_BYTES;
    if (fence) ptx::fence_proxy_async_smem();
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(SLOT_BYTES) : "memory");
    if (st.iph == 0) {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(src), "r"(SLOT_BYTES), "r"(bar) : "memory");
    } else {                                 // cross K/V: read once per step, must not evict the weights from L2
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                   ::"r"(dst), "l"(src), "r"(SLOT_BYTES), "r"(bar), "l"(policy) : "memory");
    }
  }
  ++st.issued;
  if (++st.ic == (st.iph == 0 ? st.n_self : st.n_cross)) {
    st.ic = 0;
    if (++st.iph == 2) { st.iph = 0; ++st.il; }
  }
}
// L2 prefetch of the `count` blocks that FOLLOW the issue cursor (lane i takes the i-th): called where the kernel is
// about to spend a long stretch in projections / FFN with HBM idle, so that the ring refills of the next attention hit
// L2 instead of waiting on DRAM.
__device__ __forceinline__ void stream_prefetch(const DecStackParams& p, const Stream& st, int rank, int lane, int count) {
  if (lane >= count) return;
  int il = st.il, iph = st.iph, ic = st.ic;
  for (int i = 0; i < lane && il < p.n_layers; ++i) {
    if (++ic == (iph == 0 ? st.n_self : st.n_cross)) {
      ic = 0;
      if (++iph == 2) { iph = 0; ++il; }
    }
  }
  if (il >= p.n_layers) return;
  const DecStackLayer& L = p.L[il];
  const bf16* src = iph == 0
      ? L.self_kv + ((size_t)(st.b * CL + rank) * p.self_chunks + ic) * (HPC * DS_CHUNK_ELEMS)
      : L.cross_kv + ((size_t)(st.b * CL + rank) * p.cross_chunks + ic) * (HPC * DS_CHUNK_ELEMS);
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(SLOT_BYTES) : "memory");
}
__device__ __forceinline__ void stream_wait(const Stream& st, int idx) {
  const uint32_t bar = st.bars + (idx % SLOTS) * 8, parity = (idx / SLOTS) & 1;
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  }
}

// Per-head running state of the single-query attention (row 0 of the m16 block; only quad g == 0 carries real data)
struct HeadState {
  uint32_t qa[2][2];                         // q fragments: k-step x {dims 2c.., dims 2c+8..}
  float m, l, o[4][4];
};
// One 32-key chunk of one head: scores for all 32 keys (4 n-tiles), one online-softmax update, P.V over both 16-key
// groups.  Keys >= kmax are masked; their V rows are finite (caches are zero-initialised), so p = 0 removes them.
__device__ __forceinline__ void attend_chunk(HeadState& h, uint32_t base, int kmax, int lane) {
  const int c = lane & 3;
  float s[4][4];
  zero_acc<4>(s);
#pragma unroll
  for (int j = 0; j < 4; ++j) {                              // 8 keys each
    const int key = j * 8 + (lane & 7), unit = lane >> 3;
    uint32_t k0, k1, k2, k3;
    ldsm_x4(base + key * 64 + ((unit ^ ((key >> 1) & 3)) << 4), k0, k1, k2, k3);
    mma16816(s[j], h.qa[0][0], 0u, h.qa[0][1], 0u, k0, k1);
    mma16816(s[j], h.qa[1][0], 0u, h.qa[1][1], 0u, k2, k3);
  }
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (j * 8 + 2 * c >= kmax) s[j][0] = -INFINITY;
    if (j * 8 + 2 * c + 1 >= kmax) s[j][1] = -INFINITY;
    mx = fmaxf(mx, fmaxf(s[j][0], s[j][1]));
  }
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
  mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
  const float m_new = fmaxf(h.m, mx);                        // finite: key 0 of the chunk is valid
  const float corr = exp2f(h.m - m_new);
  float ps = 0.f;
  uint32_t pa[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float p0 = exp2f(s[j][0] - m_new), p1 = exp2f(s[j][1] - m_new);
    ps += p0 + p1;
    pa[j] = pack_bf16(p0, p1);
  }
  h.l = fmaf(h.l, corr, ps);
  h.m = m_new;
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) { h.o[nt][0] *= corr; h.o[nt][1] *= corr; }
#pragma unroll
  for (int kg = 0; kg < 2; ++kg)
#pragma unroll
    for (int dp = 0; dp < 2; ++dp) {                         // 16 output dims each
      const int key = kg * 16 + ((lane >> 3) & 1) * 8 + (lane & 7), unit = 2 * dp + (lane >> 4);
      uint32_t v0, v1, v2, v3;
      ldsm_x4_trans(base + 2048 + key * 64 + ((unit ^ ((key >> 1) & 3)) << 4), v0, v1, v2, v3);
      mma16816(h.o[2 * dp], pa[2 * kg], 0u, pa[2 * kg + 1], 0u, v0, v1);
      mma16816(h.o[2 * dp + 1], pa[2 * kg], 0u, pa[2 * kg + 1], 0u, v2, v3);
    }
}

// Single-query attention of sample `warp` for the CTA's two heads over `n_keys` cached keys arriving through the warp's
// ring (chunk c of head A in one slot, of head B in the next).  q (pre-scaled by log2(e)/sqrt(dh)) is row 0 of the
// m16 A operand.  patch_key >= 0: that key's K/V rows (this step's, not yet in the cache copy) are written into the
// chunks first.  The normalised outputs (2 x 32 values) are pushed to columns [64 rank, +64) of row `warp` of every
// peer's Att buffer.
#ifdef DS_PROFILE
#define DS_A(i) do { if (apr) { const long long now_ = clock64(); apr[i] += now_ - alast; alast = now_; } } while (0)
#else
#define DS_A(i) do { } while (0)
#endif
__device__ __forceinline__ void attend2(const DecStackParams& p, Smem& sm, Stream& st, int rank, int warp, int lane,
                                        int n_keys, int patch_key, uint64_t policy, long long* apr = nullptr) {
  const int g = lane >> 2, c = lane & 3;
#ifdef DS_PROFILE
  long long alast = clock64();
#endif
  HeadState h[HPC];
#pragma unroll
  for (int hd = 0; hd < HPC; ++hd) {
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      h[hd].qa[ks][0] = g == 0 ? *reinterpret_cast<const uint32_t*>(&sm.Qs[warp][hd * DH + ks * 16 + 2 * c]) : 0u;
      h[hd].qa[ks][1] = g == 0 ? *reinterpret_cast<const uint32_t*>(&sm.Qs[warp][hd * DH + ks * 16 + 8 + 2 * c]) : 0u;
    }
    h[hd].m = -INFINITY; h[hd].l = 0.f;
    zero_acc<4>(h[hd].o);
  }
  const int n_chunks = (n_keys + DS_CHUNK_KEYS - 1) / DS_CHUNK_KEYS;
  for (int ch = 0; ch < n_chunks; ++ch) {
    const int kmax = min(DS_CHUNK_KEYS, n_keys - ch * DS_CHUNK_KEYS);
    const uint32_t base = st.ring + (st.consumed % SLOTS) * SLOT_BYTES;
    stream_wait(st, st.consumed);
    DS_A(0);
    const bool patched = patch_key >= 0 && (patch_key >> 5) == ch;
    if (patched) {
      if (lane < 16) {
        const int u = lane & 3, isv = (lane >> 2) & 1, hd = lane >> 3, kr = patch_key & 31;
        const uint4 v = *reinterpret_cast<const uint4*>(isv ? &sm.Vn[warp][hd * DH + u * 8] : &sm.Kn[warp][hd * DH + u * 8]);
        const uint32_t dst = base + hd * (DS_CHUNK_ELEMS * 2) + isv * 2048 + kr * 64 + ((u ^ ((kr >> 1) & 3)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
      }
      __syncwarp();
    }
#pragma unroll
    for (int hd = 0; hd < HPC; ++hd) attend_chunk(h[hd], base + hd * (DS_CHUNK_ELEMS * 2), kmax, lane);
    __syncwarp();                                            // every lane is done reading the slot
    DS_A(1);
    ++st.consumed;
    stream_issue(p, st, rank, lane, policy, patched);
    DS_A(2);
  }
  // lanes 0-3 hold the outputs: word (hd, nt) of quad lane c = dims hd*32 + 8 nt + 2c, +1, i.e. word index
  // 16 hd + 4 nt + c of the 32-word row slice.  Lane i sends the four words {(i & 7) + 8k} to peer i >> 3.
  uint32_t w[HPC][4];
#pragma unroll
  for (int hd = 0; hd < HPC; ++hd) {
    float l = h[hd].l;
    l += __shfl_xor_sync(0xffffffffu, l, 1);
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    const float inv = 1.f / l;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) w[hd][nt] = pack_bf16(h[hd].o[nt][0] * inv, h[hd].o[nt][1] * inv);
  }
  const int src = lane & 3, peer = lane >> 3, odd = (lane >> 2) & 1;
  const uint32_t dst = map_peer(ptx::smem_u32(&sm.Att[warp][rank * SL]), (uint32_t)peer) + (lane & 7) * 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) {          // word (lane & 7) + 8k = 16 (k >> 1) + 4 (2 (k & 1) + odd) + src
    const uint32_t va = __shfl_sync(0xffffffffu, w[k >> 1][2 * (k & 1)], src);
    const uint32_t vb = __shfl_sync(0xffffffffu, w[k >> 1][2 * (k & 1) + 1], src);
    st_peer_b32(dst + k * 32, odd ? vb : va);
  }
  DS_A(3);
}

// y slice [8 x 64] of this CTA = A . W[64 rank .. +64]^T + bias + residual (one n-tile per warp, weights already in
// registers), pushed to every peer's Y
__device__ __forceinline__ void proj_residual_push(Smem& sm, const bf16 (*A)[LDA], const WFrag<1>& wf,
                                                   const float* __restrict__ bias, int rank, int warp, int lane) {
  const int g = lane >> 2, c = lane & 3, col = rank * SL + warp * 8 + 2 * c;
  const float b0 = __ldg(bias + col), b1 = __ldg(bias + col + 1);
  AFrag a;
  load_a(a, &A[0][0], LDA, lane);
  float acc[1][4];
  zero_acc<1>(acc);
  mma_w<1>(a, wf, acc);
  const float2 v = make_float2(acc[0][0] + b0 + sm.Y[g][col], acc[0][1] + b1 + sm.Y[g][col + 1]);
  const uint32_t addr = ptx::smem_u32(&sm.Y[g][col]);
#pragma unroll
  for (int dst = 0; dst < CL; ++dst) st_peer_f2(map_peer(addr, (uint32_t)dst), v);
}

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
decode_stack_kernel(const __grid_constant__ DecStackParams p) {
  extern __shared__ __align__(128) uint8_t ds_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(ds_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int rank = (int)cluster_rank();
  const int m0 = blockIdx.y * R;
#ifdef DS_PROFILE
  const bool prof_on = tid == 0 && blockIdx.y == 0 && rank == 0;
  long long prof[20] = {0}, prof_last = clock64(), aprof[8] = {0};
#define DS_APR(i) , (prof_on ? aprof + i : nullptr)
#else
#define DS_APR(i)
#endif
  pdl_launch_dependents();
  cluster_arrive_relaxed();                                  // "my shared memory exists" (waited on before the first push)
  if (tid < NWARP * SLOTS) ptx::mbar_init(&sm.bars[0][0] + tid, 1);
  // (no clearing of the ring: a slot is only ever read after a FULL 8 KB copy from the caches, whose rows beyond a
  // sample's valid keys are finite — zero-initialised at allocation or a previous batch's values — and get p = 0)
  ptx::fence_barrier_init();
  __syncthreads();
  pdl_wait();                                                // ids, step counter and the self caches come from earlier kernels

  if (p.stagger_cycles > 0) {
    // all clusters run the same work and would move through HBM-bound (attention) and L2-bound (projection / FFN)
    // phases in lock-step; a one-off start delay per cluster group spreads the phases over time
    const long long until = clock64() + (long long)(blockIdx.y % p.stagger_groups) * p.stagger_cycles;
    while (clock64() < until) { }
  }
  const int t = *p.step;
  const bool row_ok = m0 + g < p.B;                          // this thread's GEMM output row is a real sample
  Stream st;
  st.issued = st.consumed = 0;
  st.il = st.iph = st.ic = 0;
  st.b = min(m0 + warp, p.B - 1);
  const int nv = max(1, __ldg(p.n_valid + st.b));
  st.n_self = t / DS_CHUNK_KEYS + 1;
  st.n_cross = (nv + DS_CHUNK_KEYS - 1) / DS_CHUNK_KEYS;
  st.ring = ptx::smem_u32(&sm.ring[warp][0][0]);
  st.bars = ptx::smem_u32(&sm.bars[warp][0]);
  const uint64_t policy = ptx::policy_evict_first();
#pragma unroll 1
  for (int i = 0; i < SLOTS; ++i) stream_issue(p, st, rank, lane, policy, false);
  if (p.prefetch_blocks > 0) stream_prefetch(p, st, rank, lane, p.prefetch_blocks);

  // ---- DecoderEmbeddings: y = LN_eps(E_word[id] + E_pos[t]) (models/transformer_modules.py:113-129); warp w = row w ----
  {
    const int64_t id = p.caption[(int64_t)st.b * p.max_len + t];
    float v[8];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float4 we = __ldg(reinterpret_cast<const float4*>(p.e_word + id * D + j * 128 + lane * 4));
      const float4 pe = __ldg(reinterpret_cast<const float4*>(p.e_pos + (int64_t)t * D + j * 128 + lane * 4));
      v[4 * j] = we.x + pe.x; v[4 * j + 1] = we.y + pe.y; v[4 * j + 2] = we.z + pe.z; v[4 * j + 3] = we.w + pe.w;
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[j];
    const float mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float u = v[j] - mean; q = fmaf(u, u, q); }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / D) + p.emb_eps);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float4 ga = __ldg(reinterpret_cast<const float4*>(p.emb_g + j * 128 + lane * 4));
      const float4 be = __ldg(reinterpret_cast<const float4*>(p.emb_b + j * 128 + lane * 4));
      *reinterpret_cast<float4*>(&sm.Y[warp][j * 128 + lane * 4]) =
          make_float4((v[4 * j] - mean) * rstd * ga.x + be.x, (v[4 * j + 1] - mean) * rstd * ga.y + be.y,
                      (v[4 * j + 2] - mean) * rstd * ga.z + be.z, (v[4 * j + 3] - mean) * rstd * ga.w + be.w);
    }
  }
  __syncthreads();
  cluster_wait();                                            // every CTA of the cluster is running
  DS_T(0);

  const float qscale = rsqrtf((float)DH) * LOG2E;
  const int scol = warp * 8 + 2 * c;                         // this thread's column pair within the CTA's 64-column slice
#pragma unroll 1
  for (int l = 0; l < p.n_layers; ++l) {
    const DecStackLayer& L = p.L[l];
    // ---- self-attention sublayer (models/transformer_modules.py:22-46) ----
    {
      // q, k, v columns [64 rank + 8 warp, +8) of each: three n-tiles per warp
      WFrag<3> wf;
      const bf16* wrow[3];
      float rb[3][2];
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        wrow[j] = frag_ptr(L.w_sqkv, j * (D / 8) + rank * (SL / 8) + warp, D / 32, 0, lane);
        const float* r = L.rb_sqkv + (size_t)t * 3 * D + j * D + rank * SL + scol;
        rb[j][0] = __ldg(r); rb[j][1] = __ldg(r + 1);
      }
      load_w<3>(wf, wrow);
      layernorm_row(sm.Y, sm.Abuf, L.ln1_g, L.ln1_b, 1e-5f, warp, lane);
      __syncthreads();
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      float acc[3][4];
      zero_acc<3>(acc);
      mma_w<3>(a, wf, acc);
      *reinterpret_cast<uint32_t*>(&sm.Qs[g][scol]) = pack_bf16((acc[0][0] + rb[0][0]) * qscale, (acc[0][1] + rb[0][1]) * qscale);
      const uint32_t kk = pack_bf16(acc[1][0] + rb[1][0], acc[1][1] + rb[1][1]);
      const uint32_t vv = pack_bf16(acc[2][0] + rb[2][0], acc[2][1] + rb[2][1]);
      *reinterpret_cast<uint32_t*>(&sm.Kn[g][scol]) = kk;
      *reinterpret_cast<uint32_t*>(&sm.Vn[g][scol]) = vv;
      if (row_ok) {
        // append to the cache for the steps to come (chunk t/32, row t%32, swizzled 16-byte units)
        const int kr = t & 31, unit = (warp & 3) ^ ((kr >> 1) & 3);
        bf16* dst = L.self_kv + (((size_t)((m0 + g) * CL + rank) * p.self_chunks + (t >> 5)) * HPC + (warp >> 2)) * DS_CHUNK_ELEMS +
                    kr * 32 + unit * 8 + 2 * c;
        *reinterpret_cast<uint32_t*>(dst) = kk;
        *reinterpret_cast<uint32_t*>(dst + 1024) = vv;
      }
    }
    __syncthreads();
    DS_T(1);
    WFrag<1> wo;
    {
      const bf16* wrow[1] = {frag_ptr(L.w_so, rank * (SL / 8) + warp, D / 32, 0, lane)};
      load_w<1>(wo, wrow);                                   // lands while the attention runs
    }
    attend2(p, sm, st, rank, warp, lane, t + 1, t, policy DS_APR(0));
    DS_T(2);
    cluster_arrive();
    cluster_wait();                                          // #1: Att rows complete in every CTA
    DS_T(3);
    proj_residual_push(sm, sm.Att, wo, L.b_so, rank, warp, lane);
    DS_T(4);
    cluster_arrive();
    float rq[2];
    {
      const bf16* wrow[1] = {frag_ptr(L.w_cq, rank * (SL / 8) + warp, D / 32, 0, lane)};
      load_w<1>(wo, wrow);
      const float* r = L.rb_cq + (size_t)t * D + rank * SL + scol;
      rq[0] = __ldg(r); rq[1] = __ldg(r + 1);
    }
    cluster_wait();                                          // #2: new residual rows complete
    DS_T(5);

    // ---- cross-attention sublayer (models/transformer_modules.py:49-74) ----
    layernorm_row(sm.Y, sm.Abuf, L.ln2_g, L.ln2_b, 1e-5f, warp, lane);
    __syncthreads();
    {
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      float acc[1][4];
      zero_acc<1>(acc);
      mma_w<1>(a, wo, acc);
      *reinterpret_cast<uint32_t*>(&sm.Qs[g][scol]) = pack_bf16((acc[0][0] + rq[0]) * qscale, (acc[0][1] + rq[1]) * qscale);
      const bf16* wrow[1] = {frag_ptr(L.w_co, rank * (SL / 8) + warp, D / 32, 0, lane)};
      load_w<1>(wo, wrow);
    }
    __syncthreads();
    DS_T(6);
    attend2(p, sm, st, rank, warp, lane, nv, -1, policy DS_APR(4));
    DS_T(7);
    cluster_arrive();
    cluster_wait();                                          // #3
    DS_T(8);
    proj_residual_push(sm, sm.Att, wo, L.b_co, rank, warp, lane);
    DS_T(9);
    cluster_arrive();
    // ---- feed-forward sublayer (models/transformer_modules.py:77-97), hidden slice [512 rank, +512) ----
    // FF1: 64 n-tiles, 8 per warp, two at a time through two register buffers (the next pair loads while this one
    // multiplies); FF2: K = 512 in two halves, 4 n-tiles per warp, same ping-pong.
    WFrag<2> wa, wb;
    auto ff1_rows = [&](int bt, const bf16* (&wrow)[2]) {
#pragma unroll
      for (int j = 0; j < 2; ++j) wrow[j] = frag_ptr(L.w_ff1, rank * (FS / 8) + warp * 8 + bt * 2 + j, D / 32, 0, lane);
    };
    auto ff2_rows = [&](int kh, int bt, const bf16* (&wrow)[2]) {
#pragma unroll
      for (int j = 0; j < 2; ++j)
        wrow[j] = frag_ptr(L.w_ff2, warp * 4 + bt * 2 + j, p.d_ff / 32, (rank * FS + kh * D) / 32, lane);
    };
    {
      const bf16* wrow[2];
      ff1_rows(0, wrow);
      load_w<2>(wa, wrow);
    }
    cluster_wait();                                          // #4
    if (p.prefetch_blocks > 0) stream_prefetch(p, st, rank, lane, p.prefetch_blocks);
    DS_T(10);
    layernorm_row(sm.Y, sm.Abuf, L.ln3_g, L.ln3_b, 1e-5f, warp, lane);
    __syncthreads();
    {
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      auto ff1_epilogue = [&](int bt, const float (&acc)[2][4]) {
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int col = (warp * 8 + bt * 2 + j) * 8 + 2 * c;
          const float b0 = __ldg(L.b_ff1 + rank * FS + col), b1 = __ldg(L.b_ff1 + rank * FS + col + 1);
          *reinterpret_cast<uint32_t*>(&sm.Hs[g][col]) = pack_bf16(fmaxf(acc[j][0] + b0, 0.f), fmaxf(acc[j][1] + b1, 0.f));
        }
      };
      const bf16* wrow[2];
      float acc[2][4];
      ff1_rows(1, wrow); load_w<2>(wb, wrow);
      zero_acc<2>(acc); mma_w<2>(a, wa, acc); ff1_epilogue(0, acc);
      ff1_rows(2, wrow); load_w<2>(wa, wrow);
      zero_acc<2>(acc); mma_w<2>(a, wb, acc); ff1_epilogue(1, acc);
      ff1_rows(3, wrow); load_w<2>(wb, wrow);
      zero_acc<2>(acc); mma_w<2>(a, wa, acc); ff1_epilogue(2, acc);
      ff2_rows(0, 0, wrow); load_w<2>(wa, wrow);
      zero_acc<2>(acc); mma_w<2>(a, wb, acc); ff1_epilogue(3, acc);
    }
    __syncthreads();
    DS_T(11);
    {
      AFrag a;
      const bf16* wrow[2];
      float acc0[2][4], acc1[2][4];
      zero_acc<2>(acc0); zero_acc<2>(acc1);
      load_a(a, &sm.Hs[0][0], LDH, lane);
      ff2_rows(0, 1, wrow); load_w<2>(wb, wrow);
      mma_w<2>(a, wa, acc0);
      ff2_rows(1, 0, wrow); load_w<2>(wa, wrow);
      mma_w<2>(a, wb, acc1);
      load_a(a, &sm.Hs[0][D], LDH, lane);
      ff2_rows(1, 1, wrow); load_w<2>(wb, wrow);
      mma_w<2>(a, wa, acc0);
      mma_w<2>(a, wb, acc1);
      // output columns [32 warp, +32) belong to CTA warp >> 1: its slot `rank`
      const uint32_t base = map_peer(ptx::smem_u32(&sm.Ps[rank][g][(warp & 1) * 32 + 2 * c]), (uint32_t)(warp >> 1));
      st_peer_f2(base, make_float2(acc0[0][0], acc0[0][1]));
      st_peer_f2(base + 32, make_float2(acc0[1][0], acc0[1][1]));
      st_peer_f2(base + 64, make_float2(acc1[0][0], acc1[0][1]));
      st_peer_f2(base + 96, make_float2(acc1[1][0], acc1[1][1]));
    }
    DS_T(12);
    cluster_arrive();
    const int rc2 = lane * 2, rcol = rank * SL + rc2;        // reduction: warp = row, lane = column pair of my slice
    float2 s = make_float2(__ldg(L.b_ff2 + rcol) + sm.Y[warp][rcol], __ldg(L.b_ff2 + rcol + 1) + sm.Y[warp][rcol + 1]);
    cluster_wait();                                          // #5: the four partial slots of my columns are here
    DS_T(13);
    {
      float2 part = *reinterpret_cast<const float2*>(&sm.Ps[0][warp][rc2]);
#pragma unroll
      for (int src = 1; src < CL; ++src) {
        const float2 v = *reinterpret_cast<const float2*>(&sm.Ps[src][warp][rc2]);
        part.x += v.x; part.y += v.y;
      }
      s.x += part.x; s.y += part.y;
      const uint32_t a = ptx::smem_u32(&sm.Y[warp][rcol]);
#pragma unroll
      for (int dst = 0; dst < CL; ++dst) st_peer_f2(map_peer(a, (uint32_t)dst), s);
    }
    DS_T(14);
    cluster_arrive();
    cluster_wait();                                          // #6
    DS_T(15);
  }

  // ---- final LayerNorm + MLP head layers 0 and 1 (models/ConcatTransformer.py:146-147, models/caption.py:161-174) ----
  bf16(*M1)[LDH] = sm.Hs;
  {
    // hidden columns [128 rank, +128): two n-tiles per warp, pushed to every peer (layer 1 needs whole rows)
    WFrag<2> wf;
    const int n0 = rank * (MH / CL) + warp * 16;
    const bf16* wrow[2] = {frag_ptr(p.w_m0, n0 / 8, D / 32, 0, lane), frag_ptr(p.w_m0, n0 / 8 + 1, D / 32, 0, lane)};
    load_w<2>(wf, wrow);
    layernorm_row(sm.Y, sm.Abuf, p.fin_g, p.fin_b, 1e-5f, warp, lane);
    __syncthreads();
    AFrag a;
    load_a(a, &sm.Abuf[0][0], LDA, lane);
    float acc[2][4];
    zero_acc<2>(acc);
    mma_w<2>(a, wf, acc);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int col = n0 + j * 8 + 2 * c;
      const float b0 = __ldg(p.b_m0 + col), b1 = __ldg(p.b_m0 + col + 1);
      const uint32_t v = pack_bf16(fmaxf(acc[j][0] + b0, 0.f), fmaxf(acc[j][1] + b1, 0.f));
      const uint32_t addr = ptx::smem_u32(&M1[g][col]);
#pragma unroll
      for (int dst = 0; dst < CL; ++dst) st_peer_b32(map_peer(addr, (uint32_t)dst), v);
    }
  }
  {
    WFrag<2> wf0, wf1;
    const int n0 = rank * (MH / CL) + warp * 16;
    const bf16* wrow0[2] = {frag_ptr(p.w_m1, n0 / 8, MH / 32, 0, lane), frag_ptr(p.w_m1, n0 / 8 + 1, MH / 32, 0, lane)};
    const bf16* wrow1[2] = {frag_ptr(p.w_m1, n0 / 8, MH / 32, D / 32, lane), frag_ptr(p.w_m1, n0 / 8 + 1, MH / 32, D / 32, lane)};
    cluster_arrive();
    load_w<2>(wf0, wrow0);
    load_w<2>(wf1, wrow1);
    cluster_wait();                                          // last remote stores of the kernel are behind this barrier
    float acc[2][4];
    zero_acc<2>(acc);
    AFrag a;
    load_a(a, &M1[0][0], LDH, lane);
    mma_w<2>(a, wf0, acc);
    load_a(a, &M1[0][D], LDH, lane);
    mma_w<2>(a, wf1, acc);
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = n0 + j * 8 + 2 * c;
        const float b0 = __ldg(p.b_m1 + col), b1 = __ldg(p.b_m1 + col + 1);
        *reinterpret_cast<uint32_t*>(p.m2_out + (size_t)(m0 + g) * MH + col) =
            pack_bf16(fmaxf(acc[j][0] + b0, 0.f), fmaxf(acc[j][1] + b1, 0.f));
      }
    }
  }
#ifdef DS_PROFILE
  DS_T(16);
  if (prof_on && t == 60) {
    printf("DS_PROFILE t=60 cycles: prologue %lld | ln1+qkv %lld self %lld bar1 %lld out %lld bar2 %lld | ln2+q %lld cross %lld bar3 "
           "%lld cout %lld bar4 %lld | ln3+ff1 %lld ff2 %lld bar5 %lld reduce %lld bar6 %lld | head %lld\n",
           prof[0], prof[1], prof[2], prof[3], prof[4], prof[5], prof[6], prof[7], prof[8], prof[9], prof[10], prof[11],
           prof[12], prof[13], prof[14], prof[15], prof[16]);
    printf("   attention detail: self wait %lld math %lld issue %lld tail %lld | cross wait %lld math %lld issue %lld tail %lld"
           " (self chunks/head %d, cross chunks/head %d)\n", aprof[0], aprof[1], aprof[2], aprof[3], aprof[4], aprof[5],
           aprof[6], aprof[7], st.n_self, st.n_cross);
  }
#endif
}

// rows of [K(256) | V(256)] -> [B][4 CTA ranks][chunks][2 heads][K 32x32 | V 32x32] with the 16-byte units of a key row XOR-swizzled by
// ((key % 32) >> 1) & 3; dst_row (key compaction) maps a key to its rank among the sample's unpadded keys or -1.
// With src_off (packed encoder) the source rows are packed too: key r of sample b is row src_off[b] + r of kv.
__global__ void kv_to_chunked_kernel(const bf16* __restrict__ kv, bf16* __restrict__ out, int B, int rows, int chunks,
                                     const int* __restrict__ dst_row, const int* __restrict__ src_off) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // one 16-byte unit each
  const int upr = 2 * D / 8;                                                // units per source row
  if (idx >= (int64_t)B * rows * upr) return;
  const int u = (int)(idx % upr);
  const int64_t br = idx / upr;
  const int b = (int)(br / rows);
  int r = (int)(br % rows);
  if (dst_row) {
    r = dst_row[br];
    if (r < 0) return;
  }
  const int isv = u >= D / 8, e = (isv ? u - D / 8 : u) * 8, h = e / DH, unit = (e % DH) / 8, kr = r & 31;
  const int64_t sr = src_off ? (int64_t)src_off[b] + r : br;
  const uint4 val = *reinterpret_cast<const uint4*>(kv + sr * (2 * D) + (size_t)u * 8);
  bf16* dst = out + (((size_t)(b * CL + h / HPC) * chunks + (r >> 5)) * HPC + h % HPC) * DS_CHUNK_ELEMS + isv * 1024 +
              kr * 32 + ((unit ^ ((kr >> 1) & 3)) << 3);
  *reinterpret_cast<uint4*>(dst) = val;
}

// All decoder layers at once from the batched projections: k_all / v_all are [rows, L*256] (layer l in columns
// l*256 ..), rows packed per sample (src_off); out[l] as kv_to_chunked writes it.
struct ChunkedOut { bf16* p[DS_MAX_LAYERS]; };
__global__ void kv_all_to_chunked_kernel(const bf16* __restrict__ k_all, const bf16* __restrict__ v_all, ChunkedOut out,
                                         int B, int rows, int chunks, int n_layers, const int* __restrict__ dst_row,
                                         const int* __restrict__ src_off) {
  // one 16-byte unit per thread; the layer is blockIdx.y, so the blocks of one layer run together (one layer's cache is
  // written as a stream, as six separate launches would) — with the layer as the fast index the launch took 380 us
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int upl = 2 * D / 8;                                                // units per (row, layer): K then V
  if (idx >= (int64_t)B * rows * upl) return;
  const int u = (int)(idx % upl);
  const int64_t br = idx / upl;
  const int l = (int)blockIdx.y;
  const int b = (int)(br / rows);
  const int r = dst_row[br];
  if (r < 0) return;
  const int isv = u >= D / 8, e = (isv ? u - D / 8 : u) * 8, h = e / DH, unit = (e % DH) / 8, kr = r & 31;
  const int64_t sr = (int64_t)src_off[b] + r;
  const uint4 val = *reinterpret_cast<const uint4*>((isv ? v_all : k_all) + sr * ((int64_t)n_layers * D) + l * D + e);
  bf16* dst = out.p[l] + (((size_t)(b * CL + h / HPC) * chunks + (r >> 5)) * HPC + h % HPC) * DS_CHUNK_ELEMS + isv * 1024 +
              kr * 32 + ((unit ^ ((kr >> 1) & 3)) << 3);
  *reinterpret_cast<uint4*>(dst) = val;
}

// row-major W[N][K] -> fragment order (see load_w); one thread per 16-byte piece
__global__ void repack_frag_kernel(const bf16* __restrict__ W, bf16* __restrict__ out, int N, int K) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)N * K / 8) return;
  const int lane = (int)(idx & 31);
  const int64_t blk = idx >> 5;                              // (tile, k block)
  const int kblocks = K / 32, kb = (int)(blk % kblocks), tile = (int)(blk / kblocks);
  const int n = tile * 8 + (lane >> 2), k = kb * 32 + (lane & 3) * 8;
  *reinterpret_cast<uint4*>(out + idx * 8) = *reinterpret_cast<const uint4*>(W + (size_t)n * K + k);
}

}  // namespace

void repack_frag(const bf16* W, bf16* out, int N, int K, cudaStream_t s) {
  RETR_CHECK(N % 8 == 0 && K % 32 == 0, "repack_frag: N % 8, K % 32");
  const int64_t n = (int64_t)N * K / 8;
  repack_frag_kernel<<<(unsigned)cdiv(n, 256), 256, 0, s>>>(W, out, N, K);
  LAUNCH_OK();
}

bool decode_stack_supported(int d_model, int n_heads, int d_ff, int n_layers, int mlp_hidden, int max_pos) {
  return d_model == D && n_heads == NH && d_ff == CL * FS && n_layers >= 1 && n_layers <= DS_MAX_LAYERS &&
         mlp_hidden == MH && max_pos % DS_CHUNK_KEYS == 0;
}

void decode_stack_configure() {
  CUDA_OK(cudaFuncSetAttribute(decode_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  if (getenv("RETR_VERBOSE")) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(CL, 64, 1);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = sizeof(Smem);
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, decode_stack_kernel, &cfg);
    fprintf(stderr, "[retr] decode_stack_kernel: %zu B shared memory per CTA, max co-resident clusters of %d: %d (%s)\n",
            sizeof(Smem), CL, n, cudaGetErrorString(e));
  }
}

void decode_stack(const DecStackParams& p, cudaStream_t s) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(CL, (unsigned)cdiv(p.B, R), 1);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = sizeof(Smem);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  CUDA_OK(cudaLaunchKernelEx(&cfg, decode_stack_kernel, p));
  ++g_launches;
}

void kv_to_chunked(const bf16* kv, bf16* out, int B, int rows, int chunks, const int* dst_row, cudaStream_t s,
                   const int* src_off) {
  RETR_CHECK(!src_off || dst_row, "kv_to_chunked: packed source rows need the compaction map");
  const int64_t n = (int64_t)B * rows * (2 * D / 8);
  kv_to_chunked_kernel<<<(unsigned)cdiv(n, 256), 256, 0, s>>>(kv, out, B, rows, chunks, dst_row, src_off);
  LAUNCH_OK();
}

void kv_all_to_chunked(const bf16* k_all, const bf16* v_all, bf16* const* out, int n_layers, int B, int rows, int chunks,
                       const int* dst_row, const int* src_off, cudaStream_t s) {
  RETR_CHECK(n_layers <= DS_MAX_LAYERS && dst_row && src_off, "kv_all_to_chunked: layers / maps");
  ChunkedOut o{};
  for (int l = 0; l < n_layers; ++l) o.p[l] = out[l];
  const int64_t n = (int64_t)B * rows * (2 * D / 8);
  kv_all_to_chunked_kernel<<<dim3((unsigned)cdiv(n, 256), (unsigned)n_layers), 256, 0, s>>>(k_all, v_all, o, B, rows, chunks,
                                                                                         n_layers, dst_row, src_off);
  LAUNCH_OK();
}

}  // namespace retr
