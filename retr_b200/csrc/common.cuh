// Shared host/device helpers for the retr_b200 native library (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

typedef __nv_bfloat16 bf16;

namespace retr {

// ---- error plumbing -------------------------------------------------------------------
void set_error(const std::string& msg);           // capi.cu
extern thread_local long long g_launches;         // kernels launched by this thread's calls

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

#define RETR_CHECK(cond, msg)                                                              \
  do {                                                                                     \
    if (!(cond)) throw ::retr::Error(std::string(msg) + " [" #cond "] at " __FILE__ ":" +  \
                                     std::to_string(__LINE__));                            \
  } while (0)

#define CUDA_OK(expr)                                                                      \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess)                                                                 \
      throw ::retr::Error(std::string("CUDA error ") + cudaGetErrorString(e_) +            \
                          " in " #expr " at " __FILE__ ":" + std::to_string(__LINE__));    \
  } while (0)

#define LAUNCH_OK()                                                                        \
  do {                                                                                     \
    ++::retr::g_launches;                                                                  \
    CUDA_OK(cudaGetLastError());                                                           \
  } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Programmatic dependent launch (PDL): the kernel may start while its stream predecessor is still
// running; it must execute pdl_wait() before touching anything the predecessor produces (or reads).
// Captured into a CUDA graph this becomes a programmatic dependency edge.
bool pdl_enabled();                               // capi.cu (RETR_NO_PDL=1 disables)
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
  ++g_launches;
}

// ---- in-graph timeline (diagnostic builds only: -DRETR_TRACE) --------------------------------------
// The first and the last CTA of a traced kernel record {id, grid, globaltimer at entry/exit, SM, tag}.  ncu serialises
// launches, so this is the only view of how the branches of the decode graph overlap.  Compiled out of the product.
#ifdef RETR_TRACE
#define RETR_TRACE_DECL static __device__ unsigned long long* trace_buf_ = nullptr;
__device__ __forceinline__ unsigned long long trace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TRACE_BEGIN()                                                                                                 \
  const unsigned trace_lin_ = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);                        \
  const unsigned trace_n_ = gridDim.x * gridDim.y * gridDim.z;                                                       \
  const bool trace_on_ = trace_buf_ && threadIdx.x == 0 && (trace_lin_ == 0 || trace_lin_ == trace_n_ - 1);          \
  const unsigned long long trace_t0_ = trace_on_ ? trace_now() : 0ull;
#define TRACE_END(id, tag)                                                                                            \
  if (trace_on_) {                                                                                                    \
    const unsigned long long slot_ = atomicAdd(trace_buf_, 1ull);                                                     \
    if (slot_ < (1ull << 18)) {                                                                                       \
      unsigned long long* r_ = trace_buf_ + 8 + 8 * slot_;                                                            \
      unsigned smid_;                                                                                                 \
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));                                                              \
      r_[0] = (unsigned long long)(id); r_[1] = trace_n_; r_[2] = trace_lin_; r_[3] = trace_t0_;                      \
      r_[4] = trace_now(); r_[5] = smid_; r_[6] = (unsigned long long)(tag);                                          \
    }                                                                                                                 \
  }
#else
#define RETR_TRACE_DECL
#define TRACE_BEGIN()
#define TRACE_END(id, tag)
#endif

// ---- device helpers ---------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Monotone map float -> uint32 (larger float => larger key); NaN-free inputs assumed.
__device__ __forceinline__ uint32_t float_order_key(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
// (value, index) packed so that atomicMax picks the largest value and, among equal values,
// the smallest index — torch.argmax's tie rule.
__device__ __forceinline__ unsigned long long pack_argmax(float v, uint32_t idx) {
  return ((unsigned long long)float_order_key(v) << 32) | (unsigned long long)(0xffffffffu - idx);
}
__host__ __device__ __forceinline__ uint32_t unpack_argmax_index(unsigned long long key) {
  return 0xffffffffu - (uint32_t)(key & 0xffffffffull);
}

}  // namespace retr
