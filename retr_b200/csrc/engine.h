// Native runtime of the RE⫶TR hot path: owns repacked weights, activation buffers, the K/V
// caches and the captured decode-step CUDA graph; schedules every kernel of encode / teacher-
// forced forward / greedy decode.  One Engine per (device, model); see include/retr_b200.h.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/retr_b200.h"
#include "gemm.h"
#include "ops.h"

namespace retr {

struct RawWeight {
  float* dev = nullptr;
  std::vector<int64_t> shape;
  int64_t numel = 0;
};

class EngineBase {
 public:
  explicit EngineBase(const retr_model_desc& d);
  virtual ~EngineBase();
  void load_weight(const std::string& name, const float* data, const int64_t* shape, int ndim, bool on_device);
  virtual void finalize() = 0;
  virtual void encode(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                      const float* loc, cudaStream_t s) = 0;
  virtual void teacher_forced(const int64_t* ids, const uint8_t* id_mask, float* logits, float* att_self,
                              float* att_cross, cudaStream_t s) = 0;
  virtual void decode_greedy(int bos, int eos, int max_len, int64_t* out_ids, int* steps_host, cudaStream_t s) = 0;
  virtual void read_state(const std::string& what, void* dst, int64_t bytes, cudaStream_t s) = 0;
  void generate_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                     const float* loc, int bos, int eos, int max_len, int64_t* out_ids, int* steps_host,
                     cudaStream_t s);
  // The two halves of generate_host, for callers that keep TWO contexts in flight: encode_host is asynchronous
  // (copies + backbone + encoder queued on `s`), decode_host blocks until the ids are in `out_ids`.
  void encode_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                   const float* loc, cudaStream_t s, bool split_context_copy);
  void decode_host(int bos, int eos, int max_len, int64_t* out_ids, int* steps_host, cudaStream_t s);
  // Queue the host->device copies of the NEXT batch on the copy stream (second staging set); the next generate_host
  // call with the same host pointers and batch consumes them instead of copying.
  void prefetch_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img, const uint8_t* g_mask,
                     const float* loc);

  retr_model_desc desc;
  int S = 0;              // encoder tokens per sample
  int grid_hw = 0;        // backbone output side (14 with dilation at 224)
  int n_loc_tok = 0;
  long long launches = 0;
  float phase_ms[3] = {0, 0, 0};
  float* att_enc_capture = nullptr;
  // > 0: persistent GEMM grids of encode() use at most this many CTAs (SMs), leaving the rest to the decode loop
  // of another context running at the same time (retr_set_encode_cta_cap)
  int encode_cta_cap = 0;

 protected:
  void* dalloc(size_t bytes);
  const RawWeight& raw(const std::string& name) const;
  std::map<std::string, RawWeight> raw_;
  std::vector<void*> owned_;
  // host->device staging for generate_host / prefetch_host: two sets, so a batch queued by prefetch_host can wait in
  // one while the batch before it is decoded out of the other
  struct Staging {
    float *t_img = nullptr, *g_img = nullptr, *loc = nullptr;
    uint8_t *t_mask = nullptr, *g_mask = nullptr;
    // what prefetch_host staged here and nobody consumed yet
    bool pending = false;
    int B = 0;
    const void *h_t_img = nullptr, *h_g_img = nullptr, *h_loc = nullptr;
    cudaEvent_t ready = nullptr;
  } st_[2];
  int st_next_ = 0;        // set the next prefetch goes to
  int64_t* st_ids = nullptr;
  void ensure_staging(int set);
  void ensure_copy_stream();
  // generate_host copies the context images on a second stream while the backbone works on the target images;
  // encode() waits for `ctx_ready_` before the second image set when `ctx_wait_` is set
  cudaStream_t copy_stream_ = nullptr;
  cudaEvent_t ctx_ready_ = nullptr, copy_fork_ = nullptr;
  bool ctx_wait_ = false;
  cudaEvent_t ev_[4] = {nullptr, nullptr, nullptr, nullptr};
  bool finalized_ = false;
  int cur_B_ = 0;         // batch of the last encode
};

std::unique_ptr<EngineBase> make_engine(const retr_model_desc& d);
void gemm_bf16_prepare();   // gemm_sm100.cu: set kernel attributes outside any stream capture

}  // namespace retr
