// EngineBase: weight intake, device allocations, host-buffer entry point.
#include "engine.h"

namespace retr {

EngineBase::EngineBase(const retr_model_desc& d) : desc(d) {
  RETR_CHECK(d.n_images == 1 || d.n_images == 2, "n_images must be 1 or 2");
  RETR_CHECK(d.d_model % d.n_heads == 0, "d_model must divide by n_heads");
  RETR_CHECK(d.max_batch > 0, "max_batch must be positive");
  CUDA_OK(cudaSetDevice(d.device));
  // backbone output side: /4 by stem+pool, /2 per later stage except a dilated layer4
  int side = d.image_size;
  side = (side + 2 * 3 - 7) / 2 + 1;     // conv1 7x7/2 pad 3
  side = (side + 2 * 1 - 3) / 2 + 1;     // maxpool 3x3/2 pad 1
  for (int st = 1; st < 4; ++st)
    if (!(st == 3 && d.dilation)) side = (side + 2 * 1 - 3) / 2 + 1;
  grid_hw = side;
  n_loc_tok = d.loc_in_dim == 0 ? 0 : (d.loc_in_dim == 1 ? d.n_loc : 1);
  S = grid_hw * grid_hw * d.n_images + n_loc_tok;
  for (auto& e : ev_) CUDA_OK(cudaEventCreate(&e));
}

EngineBase::~EngineBase() {
  for (void* p : owned_) cudaFree(p);
  for (auto& kv : raw_) cudaFree(kv.second.dev);
  for (auto& e : ev_) if (e) cudaEventDestroy(e);
  if (ctx_ready_) cudaEventDestroy(ctx_ready_);
  if (copy_fork_) cudaEventDestroy(copy_fork_);
  if (copy_stream_) cudaStreamDestroy(copy_stream_);
}

void* EngineBase::dalloc(size_t bytes) {
  void* p = nullptr;
  if (bytes == 0) bytes = 16;
  CUDA_OK(cudaMalloc(&p, bytes));
  owned_.push_back(p);
  return p;
}

const RawWeight& EngineBase::raw(const std::string& name) const {
  auto it = raw_.find(name);
  RETR_CHECK(it != raw_.end(), "missing weight '" + name + "' (call retr_load_weight for every state_dict entry)");
  return it->second;
}

void EngineBase::load_weight(const std::string& name, const float* data, const int64_t* shape, int ndim,
                             bool on_device) {
  RETR_CHECK(!finalized_, "weights already finalized");
  RawWeight w;
  w.numel = 1;
  for (int i = 0; i < ndim; ++i) { w.shape.push_back(shape[i]); w.numel *= shape[i]; }
  CUDA_OK(cudaMalloc(&w.dev, std::max<int64_t>(w.numel, 4) * sizeof(float)));
  CUDA_OK(cudaMemcpy(w.dev, data, w.numel * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice));
  auto it = raw_.find(name);
  if (it != raw_.end()) { cudaFree(it->second.dev); raw_.erase(it); }
  raw_[name] = w;
}

void EngineBase::generate_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                               const uint8_t* g_mask, const float* loc, int bos, int eos, int max_len,
                               int64_t* out_ids, int* steps_host, cudaStream_t s) {
  RETR_CHECK(B > 0 && B <= desc.max_batch, "batch exceeds max_batch");
  const int64_t hw = (int64_t)desc.image_size * desc.image_size;
  if (!st_t_img) {
    const int64_t mb = desc.max_batch;
    st_t_img = static_cast<float*>(dalloc(mb * 3 * hw * 4));
    st_t_mask = static_cast<uint8_t*>(dalloc(mb * hw));
    if (desc.n_images == 2) {
      st_g_img = static_cast<float*>(dalloc(mb * 3 * hw * 4));
      st_g_mask = static_cast<uint8_t*>(dalloc(mb * hw));
    }
    if (desc.n_loc > 0) st_loc = static_cast<float*>(dalloc(mb * desc.n_loc * 4));
    st_ids = static_cast<int64_t*>(dalloc(mb * desc.max_pos * 8));
  }
  // eval_utils/decode.py:62 `samples.to(device)`
  CUDA_OK(cudaMemcpyAsync(st_t_img, t_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, s));
  CUDA_OK(cudaMemcpyAsync(st_t_mask, t_mask, B * hw, cudaMemcpyHostToDevice, s));
  if (desc.n_loc > 0) {
    RETR_CHECK(loc, "location features required");
    CUDA_OK(cudaMemcpyAsync(st_loc, loc, (int64_t)B * desc.n_loc * 4, cudaMemcpyHostToDevice, s));
  }
  if (desc.n_images == 2) {
    // the context images travel on a second stream, behind the target images on the copy engine, while the backbone
    // already runs on the target images; encode() waits for them before the second image set
    RETR_CHECK(g_img && g_mask, "context images required");
    if (!copy_stream_) {
      CUDA_OK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
      CUDA_OK(cudaEventCreateWithFlags(&ctx_ready_, cudaEventDisableTiming));
      CUDA_OK(cudaEventCreateWithFlags(&copy_fork_, cudaEventDisableTiming));
    }
    CUDA_OK(cudaEventRecord(copy_fork_, s));                 // after the target copies and whatever preceded this call
    CUDA_OK(cudaStreamWaitEvent(copy_stream_, copy_fork_, 0));
    CUDA_OK(cudaMemcpyAsync(st_g_img, g_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, copy_stream_));
    CUDA_OK(cudaMemcpyAsync(st_g_mask, g_mask, B * hw, cudaMemcpyHostToDevice, copy_stream_));
    CUDA_OK(cudaEventRecord(ctx_ready_, copy_stream_));
    ctx_wait_ = true;
  }
  try {
    encode(B, st_t_img, st_t_mask, st_g_img, st_g_mask, st_loc, s);
  } catch (...) {
    ctx_wait_ = false;
    throw;
  }
  ctx_wait_ = false;
  RETR_CHECK(max_len <= desc.max_pos, "max_len exceeds max_position_embeddings");
  decode_greedy(bos, eos, max_len, st_ids, steps_host, s);
  CUDA_OK(cudaMemcpyAsync(out_ids, st_ids, (int64_t)B * max_len * 8, cudaMemcpyDeviceToHost, s));
  CUDA_OK(cudaStreamSynchronize(s));
}

}  // namespace retr
