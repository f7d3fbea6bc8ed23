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
  for (auto& st : st_) if (st.ready) cudaEventDestroy(st.ready);
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

void EngineBase::ensure_staging(int set) {
  Staging& st = st_[set];
  if (st.t_img) return;
  const int64_t hw = (int64_t)desc.image_size * desc.image_size;
  const int64_t mb = desc.max_batch;
  st.t_img = static_cast<float*>(dalloc(mb * 3 * hw * 4));
  st.t_mask = static_cast<uint8_t*>(dalloc(mb * hw));
  if (desc.n_images == 2) {
    st.g_img = static_cast<float*>(dalloc(mb * 3 * hw * 4));
    st.g_mask = static_cast<uint8_t*>(dalloc(mb * hw));
  }
  if (desc.n_loc > 0) st.loc = static_cast<float*>(dalloc(mb * desc.n_loc * 4));
  CUDA_OK(cudaEventCreateWithFlags(&st.ready, cudaEventDisableTiming));
  if (!st_ids) st_ids = static_cast<int64_t*>(dalloc(mb * desc.max_pos * 8));
}

void EngineBase::ensure_copy_stream() {
  if (copy_stream_) return;
  CUDA_OK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
  CUDA_OK(cudaEventCreateWithFlags(&ctx_ready_, cudaEventDisableTiming));
  CUDA_OK(cudaEventCreateWithFlags(&copy_fork_, cudaEventDisableTiming));
}

// The DataLoader-side overlap an eval loop gets from pinned memory + non_blocking copies (engine.py:138-146 moves one
// batch at a time): the next batch crosses PCIe on the copy stream while the current one is being decoded.  The caller
// keeps the host buffers alive and unchanged until the generate_host call that consumes them returns.
void EngineBase::prefetch_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                               const uint8_t* g_mask, const float* loc) {
  RETR_CHECK(B > 0 && B <= desc.max_batch, "batch exceeds max_batch");
  RETR_CHECK(t_img && t_mask, "target images required");
  const int64_t hw = (int64_t)desc.image_size * desc.image_size;
  // generate_host synchronises its stream before returning, so no kernel still reads either set here; a set whose
  // prefetch was never consumed is simply overwritten (copy-stream order keeps that safe)
  const int set = st_next_;
  st_next_ ^= 1;
  ensure_staging(set);
  ensure_copy_stream();
  Staging& st = st_[set];
  st.pending = false;
  CUDA_OK(cudaMemcpyAsync(st.t_img, t_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, copy_stream_));
  CUDA_OK(cudaMemcpyAsync(st.t_mask, t_mask, B * hw, cudaMemcpyHostToDevice, copy_stream_));
  if (desc.n_loc > 0) {
    RETR_CHECK(loc, "location features required");
    CUDA_OK(cudaMemcpyAsync(st.loc, loc, (int64_t)B * desc.n_loc * 4, cudaMemcpyHostToDevice, copy_stream_));
  }
  if (desc.n_images == 2) {
    RETR_CHECK(g_img && g_mask, "context images required");
    CUDA_OK(cudaMemcpyAsync(st.g_img, g_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, copy_stream_));
    CUDA_OK(cudaMemcpyAsync(st.g_mask, g_mask, B * hw, cudaMemcpyHostToDevice, copy_stream_));
  }
  CUDA_OK(cudaEventRecord(st.ready, copy_stream_));
  st.pending = true;
  st.B = B;
  st.h_t_img = t_img;
  st.h_g_img = g_img;
  st.h_loc = loc;
}

// Host buffers -> device staging -> encode(), all asynchronous on `s`.  `split_context_copy`: the context images travel
// on the copy stream behind the target images while the backbone already runs on the target images (two backbone
// passes; the serial generate_host path).  Without it everything is copied on `s` and both image sets go through the
// backbone in one pass (the pipelined path: the copies overlap ANOTHER context's decode loop instead).
void EngineBase::encode_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                             const uint8_t* g_mask, const float* loc, cudaStream_t s, bool split_context_copy) {
  RETR_CHECK(B > 0 && B <= desc.max_batch, "batch exceeds max_batch");
  const int64_t hw = (int64_t)desc.image_size * desc.image_size;
  int hit = -1;
  for (int i = 0; i < 2; ++i)
    if (st_[i].pending && st_[i].B == B && st_[i].h_t_img == t_img && st_[i].h_g_img == g_img && st_[i].h_loc == loc)
      hit = i;
  if (hit >= 0) {
    // this batch was staged by prefetch_host: wait for its copy, then one backbone pass over both image sets
    Staging& st = st_[hit];
    st.pending = false;
    CUDA_OK(cudaStreamWaitEvent(s, st.ready, 0));
    encode(B, st.t_img, st.t_mask, st.g_img, st.g_mask, st.loc, s);
    return;
  }
  // a set no prefetched batch is waiting in (the older prefetch loses its slot if both are taken)
  const int set = !st_[0].pending ? 0 : (!st_[1].pending ? 1 : st_next_);
  ensure_staging(set);
  Staging& st = st_[set];
  if (st.pending) {                                        // its copy may still be in flight on the copy stream
    st.pending = false;
    CUDA_OK(cudaStreamWaitEvent(s, st.ready, 0));
  }
  // eval_utils/decode.py:62 `samples.to(device)`
  CUDA_OK(cudaMemcpyAsync(st.t_img, t_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, s));
  CUDA_OK(cudaMemcpyAsync(st.t_mask, t_mask, B * hw, cudaMemcpyHostToDevice, s));
  if (desc.n_loc > 0) {
    RETR_CHECK(loc, "location features required");
    CUDA_OK(cudaMemcpyAsync(st.loc, loc, (int64_t)B * desc.n_loc * 4, cudaMemcpyHostToDevice, s));
  }
  if (desc.n_images == 2) {
    RETR_CHECK(g_img && g_mask, "context images required");
    if (split_context_copy) {
      ensure_copy_stream();
      CUDA_OK(cudaEventRecord(copy_fork_, s));               // after the target copies and whatever preceded this call
      CUDA_OK(cudaStreamWaitEvent(copy_stream_, copy_fork_, 0));
      CUDA_OK(cudaMemcpyAsync(st.g_img, g_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, copy_stream_));
      CUDA_OK(cudaMemcpyAsync(st.g_mask, g_mask, B * hw, cudaMemcpyHostToDevice, copy_stream_));
      CUDA_OK(cudaEventRecord(ctx_ready_, copy_stream_));
      ctx_wait_ = true;
    } else {
      CUDA_OK(cudaMemcpyAsync(st.g_img, g_img, B * 3 * hw * 4, cudaMemcpyHostToDevice, s));
      CUDA_OK(cudaMemcpyAsync(st.g_mask, g_mask, B * hw, cudaMemcpyHostToDevice, s));
    }
  }
  try {
    encode(B, st.t_img, st.t_mask, st.g_img, st.g_mask, st.loc, s);
  } catch (...) {
    ctx_wait_ = false;
    throw;
  }
  ctx_wait_ = false;
}

// Greedy loop over the batch of the last encode / encode_host, ids back to the host buffer; returns when they are there.
void EngineBase::decode_host(int bos, int eos, int max_len, int64_t* out_ids, int* steps_host, cudaStream_t s) {
  RETR_CHECK(max_len <= desc.max_pos, "max_len exceeds max_position_embeddings");
  RETR_CHECK(cur_B_ > 0, "decode_host: nothing encoded");
  ensure_staging(0);
  decode_greedy(bos, eos, max_len, st_ids, steps_host, s);
  CUDA_OK(cudaMemcpyAsync(out_ids, st_ids, (int64_t)cur_B_ * max_len * 8, cudaMemcpyDeviceToHost, s));
  CUDA_OK(cudaStreamSynchronize(s));
}

void EngineBase::generate_host(int B, const float* t_img, const uint8_t* t_mask, const float* g_img,
                               const uint8_t* g_mask, const float* loc, int bos, int eos, int max_len,
                               int64_t* out_ids, int* steps_host, cudaStream_t s) {
  RETR_CHECK(max_len <= desc.max_pos, "max_len exceeds max_position_embeddings");
  encode_host(B, t_img, t_mask, g_img, g_mask, loc, s, true);
  decode_host(bos, eos, max_len, out_ids, steps_host, s);
}

}  // namespace retr
