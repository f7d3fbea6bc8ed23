// One KV-cached greedy step through the WHOLE decoder stack in ONE kernel (d_model 256, 8 heads of 32, d_ff 2048).
//
// Reference being replaced: models/ConcatTransformer.py:220-257 (decoder layer: pre-norm self-attention,
// cross-attention, feed-forward, each with a residual), :111-149 (layer loop + final LayerNorm),
// models/transformer_modules.py:113-129 (DecoderEmbeddings) and the first two layers of the MLP head
// (models/caption.py:161-174) evaluated for ONE new position per sample against K/V caches instead of the whole
// 128-long prefix (eval_utils/decode.py:68-79).
//
// Why one kernel: as a chain of launches a decode step was ~41 dependent kernels of 5-25 us for ~70 us of HBM time
// (profiles/r1_decode_timeline.txt).  Samples are independent through the whole stack, so a thread-block CLUSTER of 8
// CTAs takes 16 samples through all six layers with no grid-wide synchronisation at all — only cluster barriers:
//   * CTA r of the cluster is head r of every attention and column slice r of every projection: the q/k/v projection
//     gives CTA r exactly the q, k, v of head r, so attention needs no exchange before it;
//   * every CTA keeps a full fp32 replica of the 16 residual rows in shared memory (LayerNorm needs whole rows); after
//     each sublayer the eight 32-column slices are pushed to every peer through distributed shared memory;
//   * the FFN splits the hidden dimension: CTA r computes relu(W1[256r:256r+256] LN(y)) and its partial product with
//     W2[:, 256r:256r+256], pushes 32-column slices of the partial to the owning CTA, which sums them in a fixed order;
//   * GEMMs are M = 16 (one m16n8k16 row block): weights stream global(L2) -> registers as B fragments with 16-byte
//     loads (each weight byte is used exactly once per CTA, so staging it in shared memory buys nothing); the k index
//     is permuted identically for A and B so that a thread's 16 bytes are 8 consecutive k;
//   * attention: K/V caches are stored in 4 KB chunks [K 32 keys x 32 | V 32 keys x 32] (bf16, 16-byte units XOR-
//     swizzled by the key index, written that way by the cache builders), each warp owns two samples and a private
//     5-slot ring of chunks filled by cp.async.bulk (one elected lane, completion on an mbarrier); the ring runs
//     AHEAD across sublayers and layers (the next cross-attention's chunks arrive while the projections run).
//     Scores and P.V run on the tensor cores (q as row 0 of the A operand; ldmatrix / ldmatrix.trans on the swizzled
//     chunk), online softmax per 16 keys: math starts on the first chunk instead of waiting for a whole K block.
#include "decode_stack.h"

#include "sm100_ptx.cuh"

namespace retr {
namespace {

constexpr int D = 256, DH = 32, NH = 8, R = DS_ROWS, THREADS = 256, NWARP = 8;
constexpr int LDA = D + 32;                  // bf16 per A row in shared memory: 576 B, rows 16 banks apart
constexpr int MH = 512, LDM1 = MH + 32;      // MLP head hidden width
constexpr int SLOTS = 5, SLOT_BYTES = DS_CHUNK_ELEMS * 2;
constexpr float LOG2E = 1.4426950408889634f;

struct Smem {
  uint8_t ring[NWARP][SLOTS][SLOT_BYTES];    // per-warp K/V chunk rings
  float Y[R][D];                             // residual stream rows (full replica in every CTA of the cluster)
  float Ps[NH][R][32];                       // FFN partials of this CTA's 32 output columns, one slot per source CTA
  bf16 Abuf[R][LDA];                         // LayerNorm output (A operand)
  bf16 Att[R][LDA];                          // attention output, heads gathered from the peers
  bf16 Hs[R][LDA];                           // FFN hidden slice; Att+Hs are reused as M1[R][LDM1] by the head
  bf16 Qs[R][DH], Kn[R][DH], Vn[R][DH];      // this head's (scaled) query and the step's new key / value rows
  uint64_t bars[NWARP][SLOTS];
};
static_assert(sizeof(bf16) * R * LDM1 <= 2 * sizeof(bf16) * R * LDA, "M1 must fit in Att+Hs");

// ---- small PTX wrappers ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync() { cluster_arrive(); cluster_wait(); }
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

// ---- M = 16 GEMM pieces --------------------------------------------------------------------------------------------
// A fragments of the 16 rows over 256 consecutive k (8 blocks of 32).  Within a block thread (g, c) owns the 8
// consecutive k [8c, 8c+8) of rows g and g+8: k-step 0 uses elements {0,1 | 2,3}, k-step 1 {4,5 | 6,7} — the same
// assignment as the 16-byte weight loads below, so the products pair up correctly (a dot product does not care in
// which order k is visited).
struct AFrag {
  uint4 x[8], y[8];
};
__device__ __forceinline__ void load_a(AFrag& a, const bf16* A, int lda, int lane) {
  const int g = lane >> 2, c = lane & 3;
#pragma unroll
  for (int kb = 0; kb < 8; ++kb) {
    a.x[kb] = *reinterpret_cast<const uint4*>(A + (size_t)g * lda + kb * 32 + 8 * c);
    a.y[kb] = *reinterpret_cast<const uint4*>(A + (size_t)(g + 8) * lda + kb * 32 + 8 * c);
  }
}
// acc[j] (+)= A[16 x 256] . W_j[8 x 256]^T for NT n-tiles; wrow[j] = &W[(n0_j + g) * ldw + k0 + 8c]
template <int NT>
__device__ __forceinline__ void mma_tiles(const AFrag& a, const bf16* const (&wrow)[NT], float (&acc)[NT][4]) {
  uint4 w[NT][8];
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) w[j][kb] = ldg_stream(wrow[j] + kb * 32);
#pragma unroll
  for (int j = 0; j < NT; ++j)
#pragma unroll
    for (int kb = 0; kb < 8; ++kb) {
      mma16816(acc[j], a.x[kb].x, a.y[kb].x, a.x[kb].y, a.y[kb].y, w[j][kb].x, w[j][kb].y);
      mma16816(acc[j], a.x[kb].z, a.y[kb].z, a.x[kb].w, a.y[kb].w, w[j][kb].z, w[j][kb].w);
    }
}
template <int NT> __device__ __forceinline__ void zero_acc(float (&acc)[NT][4]) {
#pragma unroll
  for (int j = 0; j < NT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
}

// LayerNorm of the 16 replica rows -> bf16 A operand; warp w takes rows 2w, 2w+1
__device__ __forceinline__ void layernorm_rows(const float (*Y)[D], bf16 (*A)[LDA], const float* __restrict__ gamma,
                                               const float* __restrict__ beta, float eps, int warp, int lane) {
  float4 ga[2], be[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    ga[j] = __ldg(reinterpret_cast<const float4*>(gamma + j * 128 + lane * 4));
    be[j] = __ldg(reinterpret_cast<const float4*>(beta + j * 128 + lane * 4));
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = warp * 2 + i;
    float v[8];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float4 t = *reinterpret_cast<const float4*>(&Y[r][j * 128 + lane * 4]);
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
      uint2 pk;
      pk.x = pack_bf16((v[4 * j] - mean) * rstd * ga[j].x + be[j].x, (v[4 * j + 1] - mean) * rstd * ga[j].y + be[j].y);
      pk.y = pack_bf16((v[4 * j + 2] - mean) * rstd * ga[j].z + be[j].z, (v[4 * j + 3] - mean) * rstd * ga[j].w + be[j].w);
      *reinterpret_cast<uint2*>(&A[r][j * 128 + lane * 4]) = pk;
    }
  }
}

// ---- per-warp K/V chunk stream ---------------------------------------------------------------------------------------
// The chunks a warp will consume during the whole kernel form one fixed sequence:
//   for layer: [self chunks of its sample 0, of its sample 1, cross chunks of sample 0, of sample 1]
// The issue cursor runs SLOTS chunks ahead of the consume cursor, across sublayer and layer boundaries.
struct Stream {
  int issued, consumed;
  int il, iph, isi, ic;                      // issue cursor: layer, phase (0 self / 1 cross), sample 0/1, chunk
  int n_self, n_cross0, n_cross1;            // chunks per (phase, sample)
  int b0, b1;                                // the warp's two samples (clamped to B-1)
  uint32_t ring, bars;                       // shared-memory addresses of this warp's ring and barriers
};
__device__ __forceinline__ void stream_issue(const DecStackParams& p, Stream& st, int rank, int lane, uint64_t policy) {
  if (st.il >= p.n_layers) return;
  const int slot = st.issued % SLOTS;
  if (lane == 0) {
    const DecStackLayer& L = p.L[st.il];
    const int b = st.isi ? st.b1 : st.b0;
    const bf16* src = st.iph == 0
        ? L.self_kv + ((size_t)(b * NH + rank) * p.self_chunks + st.ic) * DS_CHUNK_ELEMS
        : L.cross_kv + ((size_t)(b * NH + rank) * p.cross_chunks + st.ic) * DS_CHUNK_ELEMS;
    const uint32_t bar = st.bars + slot * 8, dst = st.ring + slot * SLOT_BYTES;
    ptx::fence_proxy_async_smem();           // the slot was last touched through the generic proxy
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
  const int n = st.iph == 0 ? st.n_self : (st.isi ? st.n_cross1 : st.n_cross0);
  if (++st.ic == n) {
    st.ic = 0;
    if (++st.isi == 2) {
      st.isi = 0;
      if (++st.iph == 2) { st.iph = 0; ++st.il; }
    }
  }
}
__device__ __forceinline__ void stream_wait(const Stream& st) {
  const uint32_t bar = st.bars + (st.consumed % SLOTS) * 8, parity = (st.consumed / SLOTS) & 1;
  uint32_t ok = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  }
}

// Single-query attention of one (sample, head) over `n_keys` cached keys arriving through the warp's ring.
// q (pre-scaled by log2(e)/sqrt(dh)) is row 0 of the m16 A operand; K chunks give the B fragments of q.K^T through
// ldmatrix, V chunks those of P.V through ldmatrix.trans; only the quad g == 0 (lanes 0-3) carries real rows.
// patch_key >= 0: that key's K/V rows (this step's, not yet in the cache copy) are written into the chunk first.
// The normalised output (32 values) is pushed to column slice `rank` of row `row` of every peer's Att buffer.
__device__ __forceinline__ void attend(const DecStackParams& p, Smem& sm, Stream& st, int rank, int warp, int lane,
                                       int row, int n_keys, int patch_key, uint64_t policy) {
  const int g = lane >> 2, c = lane & 3;
  uint32_t qa[2][2];                                         // k-step x {a0 (dims 2c..), a2 (dims 2c+8..)}; rows 8-15: 0
#pragma unroll
  for (int ks = 0; ks < 2; ++ks) {
    qa[ks][0] = g == 0 ? *reinterpret_cast<const uint32_t*>(&sm.Qs[row][ks * 16 + 2 * c]) : 0u;
    qa[ks][1] = g == 0 ? *reinterpret_cast<const uint32_t*>(&sm.Qs[row][ks * 16 + 8 + 2 * c]) : 0u;
  }
  float m = -INFINITY, l = 0.f, o[4][4];
  zero_acc<4>(o);
  const int n_chunks = (n_keys + DS_CHUNK_KEYS - 1) / DS_CHUNK_KEYS;
  for (int ch = 0; ch < n_chunks; ++ch) {
    const uint32_t base = st.ring + (st.consumed % SLOTS) * SLOT_BYTES;
    stream_wait(st);
    if (patch_key >= 0 && (patch_key >> 5) == ch) {
      if (lane < 8) {
        const int u = lane & 3, isv = lane >> 2, kr = patch_key & 31;
        const uint4 v = *reinterpret_cast<const uint4*>(isv ? &sm.Vn[row][u * 8] : &sm.Kn[row][u * 8]);
        const uint32_t dst = base + isv * 2048 + kr * 64 + ((u ^ ((kr >> 1) & 3)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
      }
      __syncwarp();
    }
    const int kmax = min(DS_CHUNK_KEYS, n_keys - ch * DS_CHUNK_KEYS);
#pragma unroll
    for (int kg = 0; kg < 2; ++kg) {
      if (kg * 16 < kmax) {
        float s[2][4];
        zero_acc<2>(s);
#pragma unroll
        for (int j = 0; j < 2; ++j) {                        // 8 keys each
          const int key = kg * 16 + j * 8 + (lane & 7), unit = lane >> 3;
          uint32_t k0, k1, k2, k3;
          ldsm_x4(base + key * 64 + ((unit ^ ((key >> 1) & 3)) << 4), k0, k1, k2, k3);
          mma16816(s[j], qa[0][0], 0u, qa[0][1], 0u, k0, k1);
          mma16816(s[j], qa[1][0], 0u, qa[1][1], 0u, k2, k3);
        }
        const int key0 = kg * 16 + 2 * c;                    // chunk-local key of s[0][0]; s[1][*] are 8 further
        if (key0 >= kmax) s[0][0] = -INFINITY;
        if (key0 + 1 >= kmax) s[0][1] = -INFINITY;
        if (key0 + 8 >= kmax) s[1][0] = -INFINITY;
        if (key0 + 9 >= kmax) s[1][1] = -INFINITY;
        float mx = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m, mx);                    // finite: the group holds at least one valid key
        const float corr = exp2f(m - m_new);
        const float p00 = exp2f(s[0][0] - m_new), p01 = exp2f(s[0][1] - m_new);
        const float p10 = exp2f(s[1][0] - m_new), p11 = exp2f(s[1][1] - m_new);
        l = fmaf(l, corr, (p00 + p01) + (p10 + p11));
        m = m_new;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) { o[nt][0] *= corr; o[nt][1] *= corr; }
        const uint32_t pa0 = pack_bf16(p00, p01), pa2 = pack_bf16(p10, p11);
#pragma unroll
        for (int dp = 0; dp < 2; ++dp) {                     // 16 output dims each
          const int key = kg * 16 + ((lane >> 3) & 1) * 8 + (lane & 7), unit = 2 * dp + (lane >> 4);
          uint32_t v0, v1, v2, v3;
          ldsm_x4_trans(base + 2048 + key * 64 + ((unit ^ ((key >> 1) & 3)) << 4), v0, v1, v2, v3);
          mma16816(o[2 * dp], pa0, 0u, pa2, 0u, v0, v1);
          mma16816(o[2 * dp + 1], pa0, 0u, pa2, 0u, v2, v3);
        }
      }
    }
    __syncwarp();                                            // every lane is done reading the slot
    ++st.consumed;
    stream_issue(p, st, rank, lane, policy);
  }
  l += __shfl_xor_sync(0xffffffffu, l, 1);
  l += __shfl_xor_sync(0xffffffffu, l, 2);
  const float inv = 1.f / l;
  uint32_t w[4];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) w[nt] = pack_bf16(o[nt][0] * inv, o[nt][1] * inv);   // dims 8nt + 2c, +1 (lanes 0-3)
  // lane i sends quad-lane (i & 3)'s four words to peer i >> 2
  const int src = lane & 3, peer = lane >> 2;
  const uint32_t dst = map_peer(ptx::smem_u32(&sm.Att[row][rank * DH + 2 * src]), (uint32_t)peer);
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) st_peer_b32(dst + nt * 16, __shfl_sync(0xffffffffu, w[nt], src));
}

// y slice [16 x 32] of this CTA = A . W[32 rank .. +32]^T + bias + residual, pushed to every peer's Y (warps 0-3)
__device__ __forceinline__ void proj_residual_push(Smem& sm, const bf16 (*A)[LDA], const bf16* __restrict__ W,
                                                   const float* __restrict__ bias, int rank, int warp, int lane) {
  if (warp < 4) {
    const int g = lane >> 2, c = lane & 3;
    AFrag a;
    load_a(a, &A[0][0], LDA, lane);
    const bf16* wrow[1] = {W + (size_t)(rank * 32 + warp * 8 + g) * D + 8 * c};
    float acc[1][4];
    zero_acc<1>(acc);
    mma_tiles<1>(a, wrow, acc);
    const int col = rank * 32 + warp * 8 + 2 * c;
    const float b0 = __ldg(bias + col), b1 = __ldg(bias + col + 1);
    const float2 lo = make_float2(acc[0][0] + b0 + sm.Y[g][col], acc[0][1] + b1 + sm.Y[g][col + 1]);
    const float2 hi = make_float2(acc[0][2] + b0 + sm.Y[g + 8][col], acc[0][3] + b1 + sm.Y[g + 8][col + 1]);
    const uint32_t a_lo = ptx::smem_u32(&sm.Y[g][col]), a_hi = ptx::smem_u32(&sm.Y[g + 8][col]);
#pragma unroll
    for (int dst = 0; dst < NH; ++dst) {
      st_peer_f2(map_peer(a_lo, (uint32_t)dst), lo);
      st_peer_f2(map_peer(a_hi, (uint32_t)dst), hi);
    }
  }
}

__global__ void __cluster_dims__(NH, 1, 1) __launch_bounds__(THREADS, 1)
decode_stack_kernel(const __grid_constant__ DecStackParams p) {
  extern __shared__ __align__(128) uint8_t ds_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(ds_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int rank = (int)cluster_rank();
  const int m0 = blockIdx.y * R;

  pdl_launch_dependents();
  cluster_arrive_relaxed();                                  // "my shared memory exists" (waited on before the first push)
  if (tid < NWARP * SLOTS) ptx::mbar_init(&sm.bars[0][0] + tid, 1);
  // stale shared memory may hold NaN bit patterns: rows of a chunk beyond the valid keys are multiplied by p = 0
  for (int i = tid; i < (int)(sizeof(sm.ring) / 16); i += THREADS)
    reinterpret_cast<uint4*>(&sm.ring[0][0][0])[i] = make_uint4(0u, 0u, 0u, 0u);
  ptx::fence_barrier_init();
  __syncthreads();
  pdl_wait();                                                // ids, step counter and the self caches come from earlier kernels

  const int t = *p.step;
  Stream st;
  st.issued = st.consumed = 0;
  st.il = st.iph = st.isi = st.ic = 0;
  st.n_self = t / DS_CHUNK_KEYS + 1;
  st.b0 = min(m0 + 2 * warp, p.B - 1);
  st.b1 = min(m0 + 2 * warp + 1, p.B - 1);
  const int nv0 = max(1, __ldg(p.n_valid + st.b0)), nv1 = max(1, __ldg(p.n_valid + st.b1));
  st.n_cross0 = (nv0 + DS_CHUNK_KEYS - 1) / DS_CHUNK_KEYS;
  st.n_cross1 = (nv1 + DS_CHUNK_KEYS - 1) / DS_CHUNK_KEYS;
  st.ring = ptx::smem_u32(&sm.ring[warp][0][0]);
  st.bars = ptx::smem_u32(&sm.bars[warp][0]);
  const uint64_t policy = ptx::policy_evict_first();
#pragma unroll 1
  for (int i = 0; i < SLOTS; ++i) stream_issue(p, st, rank, lane, policy);

  // ---- DecoderEmbeddings: y = LN_eps(E_word[id] + E_pos[t]) (models/transformer_modules.py:113-129) ----
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = warp * 2 + i, b = min(m0 + r, p.B - 1);
    const int64_t id = p.caption[(int64_t)b * p.max_len + t];
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
      *reinterpret_cast<float4*>(&sm.Y[r][j * 128 + lane * 4]) =
          make_float4((v[4 * j] - mean) * rstd * ga.x + be.x, (v[4 * j + 1] - mean) * rstd * ga.y + be.y,
                      (v[4 * j + 2] - mean) * rstd * ga.z + be.z, (v[4 * j + 3] - mean) * rstd * ga.w + be.w);
    }
  }
  __syncthreads();
  cluster_wait();                                            // every CTA of the cluster is running

  const float qscale = rsqrtf((float)DH) * LOG2E;
#pragma unroll 1
  for (int l = 0; l < p.n_layers; ++l) {
    const DecStackLayer& L = p.L[l];
    // ---- self-attention sublayer (models/transformer_modules.py:22-46) ----
    layernorm_rows(sm.Y, sm.Abuf, L.ln1_g, L.ln1_b, 1e-5f, warp, lane);
    __syncthreads();
    {
      // q, k, v of head `rank`: 12 n-tiles (4 each), warp w takes tile w and, for w < 4, tile w + 8
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      const int ntiles = warp < 4 ? 2 : 1;
      float acc[2][4];
      zero_acc<2>(acc);
      auto wrow_of = [&](int nt) {
        return L.w_sqkv + (size_t)((nt >> 2) * D + rank * DH + (nt & 3) * 8 + g) * D + 8 * c;
      };
      if (ntiles == 2) {
        const bf16* wrow[2] = {wrow_of(warp), wrow_of(warp + 8)};
        mma_tiles<2>(a, wrow, acc);
      } else {
        const bf16* wrow[1] = {wrow_of(warp)};
        float acc1[1][4];
        zero_acc<1>(acc1);
        mma_tiles<1>(a, wrow, acc1);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[0][k] = acc1[0][k];
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (j >= ntiles) break;
        const int nt = warp + 8 * j, which = nt >> 2, col = (nt & 3) * 8 + 2 * c;       // col within the head
        const float* rb = L.rb_sqkv + (size_t)t * 3 * D + which * D + rank * DH + col;
        const float r0 = __ldg(rb), r1 = __ldg(rb + 1);
        float v00 = acc[j][0] + r0, v01 = acc[j][1] + r1, v10 = acc[j][2] + r0, v11 = acc[j][3] + r1;
        if (which == 0) {
          *reinterpret_cast<uint32_t*>(&sm.Qs[g][col]) = pack_bf16(v00 * qscale, v01 * qscale);
          *reinterpret_cast<uint32_t*>(&sm.Qs[g + 8][col]) = pack_bf16(v10 * qscale, v11 * qscale);
        } else {
          bf16(*dst)[DH] = which == 1 ? sm.Kn : sm.Vn;
          const uint32_t lo = pack_bf16(v00, v01), hi = pack_bf16(v10, v11);
          *reinterpret_cast<uint32_t*>(&dst[g][col]) = lo;
          *reinterpret_cast<uint32_t*>(&dst[g + 8][col]) = hi;
          // append to the cache for the steps to come (chunk t/32, row t%32, swizzled 16-byte units)
          const int kr = t & 31, unit = (nt & 3) ^ ((kr >> 1) & 3);
          const size_t off = (size_t)(t >> 5) * DS_CHUNK_ELEMS + (which == 2 ? 1024 : 0) + kr * 32 + unit * 8 + 2 * c;
          if (m0 + g < p.B)
            *reinterpret_cast<uint32_t*>(L.self_kv + (size_t)((m0 + g) * NH + rank) * p.self_chunks * DS_CHUNK_ELEMS + off) = lo;
          if (m0 + g + 8 < p.B)
            *reinterpret_cast<uint32_t*>(L.self_kv + (size_t)((m0 + g + 8) * NH + rank) * p.self_chunks * DS_CHUNK_ELEMS + off) = hi;
        }
      }
    }
    __syncthreads();
    attend(p, sm, st, rank, warp, lane, 2 * warp, t + 1, t, policy);
    attend(p, sm, st, rank, warp, lane, 2 * warp + 1, t + 1, t, policy);
    cluster_sync();                                          // #1: Att rows complete in every CTA
    proj_residual_push(sm, sm.Att, L.w_so, L.b_so, rank, warp, lane);
    cluster_sync();                                          // #2: new residual rows complete

    // ---- cross-attention sublayer (models/transformer_modules.py:49-74) ----
    layernorm_rows(sm.Y, sm.Abuf, L.ln2_g, L.ln2_b, 1e-5f, warp, lane);
    __syncthreads();
    if (warp < 4) {
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      const bf16* wrow[1] = {L.w_cq + (size_t)(rank * DH + warp * 8 + g) * D + 8 * c};
      float acc[1][4];
      zero_acc<1>(acc);
      mma_tiles<1>(a, wrow, acc);
      const int col = warp * 8 + 2 * c;
      const float* rb = L.rb_cq + (size_t)t * D + rank * DH + col;
      const float r0 = __ldg(rb), r1 = __ldg(rb + 1);
      *reinterpret_cast<uint32_t*>(&sm.Qs[g][col]) = pack_bf16((acc[0][0] + r0) * qscale, (acc[0][1] + r1) * qscale);
      *reinterpret_cast<uint32_t*>(&sm.Qs[g + 8][col]) = pack_bf16((acc[0][2] + r0) * qscale, (acc[0][3] + r1) * qscale);
    }
    __syncthreads();
    attend(p, sm, st, rank, warp, lane, 2 * warp, nv0, -1, policy);
    attend(p, sm, st, rank, warp, lane, 2 * warp + 1, nv1, -1, policy);
    cluster_sync();                                          // #3
    proj_residual_push(sm, sm.Att, L.w_co, L.b_co, rank, warp, lane);
    cluster_sync();                                          // #4

    // ---- feed-forward sublayer (models/transformer_modules.py:77-97), hidden slice [256 rank, +256) ----
    layernorm_rows(sm.Y, sm.Abuf, L.ln3_g, L.ln3_b, 1e-5f, warp, lane);
    __syncthreads();
    {
      AFrag a;
      load_a(a, &sm.Abuf[0][0], LDA, lane);
      const bf16* wrow[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wrow[j] = L.w_ff1 + (size_t)(rank * D + (warp * 4 + j) * 8 + g) * D + 8 * c;
      float acc[4][4];
      zero_acc<4>(acc);
      mma_tiles<4>(a, wrow, acc);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int col = (warp * 4 + j) * 8 + 2 * c;
        const float b0 = __ldg(L.b_ff1 + rank * D + col), b1 = __ldg(L.b_ff1 + rank * D + col + 1);
        *reinterpret_cast<uint32_t*>(&sm.Hs[g][col]) = pack_bf16(fmaxf(acc[j][0] + b0, 0.f), fmaxf(acc[j][1] + b1, 0.f));
        *reinterpret_cast<uint32_t*>(&sm.Hs[g + 8][col]) = pack_bf16(fmaxf(acc[j][2] + b0, 0.f), fmaxf(acc[j][3] + b1, 0.f));
      }
    }
    __syncthreads();
    {
      AFrag a;
      load_a(a, &sm.Hs[0][0], LDA, lane);
      const bf16* wrow[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wrow[j] = L.w_ff2 + (size_t)((warp * 4 + j) * 8 + g) * p.d_ff + rank * D + 8 * c;
      float acc[4][4];
      zero_acc<4>(acc);
      mma_tiles<4>(a, wrow, acc);
      // output columns [32 warp, +32) belong to CTA `warp`: its slot `rank`
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t a_lo = map_peer(ptx::smem_u32(&sm.Ps[rank][g][j * 8 + 2 * c]), (uint32_t)warp);
        const uint32_t a_hi = map_peer(ptx::smem_u32(&sm.Ps[rank][g + 8][j * 8 + 2 * c]), (uint32_t)warp);
        st_peer_f2(a_lo, make_float2(acc[j][0], acc[j][1]));
        st_peer_f2(a_hi, make_float2(acc[j][2], acc[j][3]));
      }
    }
    cluster_sync();                                          // #5: the eight partial slots of my columns are here
    {
      const int row = tid >> 4, c2 = (tid & 15) * 2, col = rank * 32 + c2;
      float2 s = make_float2(__ldg(L.b_ff2 + col) + sm.Y[row][col], __ldg(L.b_ff2 + col + 1) + sm.Y[row][col + 1]);
      float2 part = *reinterpret_cast<const float2*>(&sm.Ps[0][row][c2]);
#pragma unroll
      for (int src = 1; src < NH; ++src) {
        const float2 v = *reinterpret_cast<const float2*>(&sm.Ps[src][row][c2]);
        part.x += v.x; part.y += v.y;
      }
      s.x += part.x; s.y += part.y;
      const uint32_t a = ptx::smem_u32(&sm.Y[row][col]);
#pragma unroll
      for (int dst = 0; dst < NH; ++dst) st_peer_f2(map_peer(a, (uint32_t)dst), s);
    }
    cluster_sync();                                          // #6
  }

  // ---- final LayerNorm + MLP head layers 0 and 1 (models/ConcatTransformer.py:146-147, models/caption.py:161-174) ----
  layernorm_rows(sm.Y, sm.Abuf, p.fin_g, p.fin_b, 1e-5f, warp, lane);
  __syncthreads();
  bf16(*M1)[LDM1] = reinterpret_cast<bf16(*)[LDM1]>(&sm.Att[0][0]);
  {
    // hidden columns [64 rank, +64): one n-tile per warp, pushed to every peer (layer 1 needs whole rows)
    AFrag a;
    load_a(a, &sm.Abuf[0][0], LDA, lane);
    const int n0 = rank * 64 + warp * 8;
    const bf16* wrow[1] = {p.w_m0 + (size_t)(n0 + g) * D + 8 * c};
    float acc[1][4];
    zero_acc<1>(acc);
    mma_tiles<1>(a, wrow, acc);
    const float b0 = __ldg(p.b_m0 + n0 + 2 * c), b1 = __ldg(p.b_m0 + n0 + 2 * c + 1);
    const uint32_t lo = pack_bf16(fmaxf(acc[0][0] + b0, 0.f), fmaxf(acc[0][1] + b1, 0.f));
    const uint32_t hi = pack_bf16(fmaxf(acc[0][2] + b0, 0.f), fmaxf(acc[0][3] + b1, 0.f));
    const uint32_t a_lo = ptx::smem_u32(&M1[g][n0 + 2 * c]), a_hi = ptx::smem_u32(&M1[g + 8][n0 + 2 * c]);
#pragma unroll
    for (int dst = 0; dst < NH; ++dst) {
      st_peer_b32(map_peer(a_lo, (uint32_t)dst), lo);
      st_peer_b32(map_peer(a_hi, (uint32_t)dst), hi);
    }
  }
  cluster_sync();                                            // last remote stores of the kernel are behind this barrier
  {
    const int n0 = rank * 64 + warp * 8;
    float acc[1][4];
    zero_acc<1>(acc);
#pragma unroll 1
    for (int kh = 0; kh < MH / D; ++kh) {
      AFrag a;
      load_a(a, &M1[0][kh * D], LDM1, lane);
      const bf16* wrow[1] = {p.w_m1 + (size_t)(n0 + g) * MH + kh * D + 8 * c};
      mma_tiles<1>(a, wrow, acc);
    }
    const float b0 = __ldg(p.b_m1 + n0 + 2 * c), b1 = __ldg(p.b_m1 + n0 + 2 * c + 1);
    if (m0 + g < p.B)
      *reinterpret_cast<uint32_t*>(p.m2_out + (size_t)(m0 + g) * MH + n0 + 2 * c) =
          pack_bf16(fmaxf(acc[0][0] + b0, 0.f), fmaxf(acc[0][1] + b1, 0.f));
    if (m0 + g + 8 < p.B)
      *reinterpret_cast<uint32_t*>(p.m2_out + (size_t)(m0 + g + 8) * MH + n0 + 2 * c) =
          pack_bf16(fmaxf(acc[0][2] + b0, 0.f), fmaxf(acc[0][3] + b1, 0.f));
  }
}

// rows of [K(256) | V(256)] -> [B][8][chunks][K 32x32 | V 32x32] with the 16-byte units of a key row XOR-swizzled by
// ((key % 32) >> 1) & 3; dst_row (key compaction) maps a key to its rank among the sample's unpadded keys or -1.
__global__ void kv_to_chunked_kernel(const bf16* __restrict__ kv, bf16* __restrict__ out, int B, int rows, int chunks,
                                     const int* __restrict__ dst_row) {
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
  const uint4 val = *reinterpret_cast<const uint4*>(kv + br * (2 * D) + (size_t)u * 8);
  bf16* dst = out + ((size_t)(b * NH + h) * chunks + (r >> 5)) * DS_CHUNK_ELEMS + isv * 1024 + kr * 32 +
              ((unit ^ ((kr >> 1) & 3)) << 3);
  *reinterpret_cast<uint4*>(dst) = val;
}

}  // namespace

bool decode_stack_supported(int d_model, int n_heads, int d_ff, int n_layers, int mlp_hidden, int max_pos) {
  return d_model == D && n_heads == NH && d_ff == NH * D && n_layers >= 1 && n_layers <= DS_MAX_LAYERS &&
         mlp_hidden == MH && max_pos % DS_CHUNK_KEYS == 0;
}

void decode_stack_configure() {
  CUDA_OK(cudaFuncSetAttribute(decode_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
}

void decode_stack(const DecStackParams& p, cudaStream_t s) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(NH, (unsigned)cdiv(p.B, R), 1);
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

void kv_to_chunked(const bf16* kv, bf16* out, int B, int rows, int chunks, const int* dst_row, cudaStream_t s) {
  const int64_t n = (int64_t)B * rows * (2 * D / 8);
  kv_to_chunked_kernel<<<(unsigned)cdiv(n, 256), 256, 0, s>>>(kv, out, B, rows, chunks, dst_row);
  LAUNCH_OK();
}

}  // namespace retr
