// NCCL plumbing of the data-parallel training step.  The reference trains on ONE device (main.py:16-24; its only
// multi-process trace is the seed offset at main.py:19), so there is no reference collective to mirror: this is the
// exchange step SURVEY §8(e) scopes — mean of the per-rank gradients of the same B*T-position loss — bucketed in
// reverse-execution order and overlapped with the rest of the backward pass (train.cu: bucket_ready).
#include "dp.h"

#include <dlfcn.h>

#include <mutex>

namespace retr {

const NcclApi& nccl() {
  static NcclApi api{};
  static std::once_flag once;
  static std::string err;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { err = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
    auto sym = [&](const char* n) {
      void* p = dlsym(h, n);
      if (!p && err.empty()) err = std::string("libnccl lacks ") + n;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  });
  if (!err.empty()) throw Error(err);
  return api;
}

void nccl_ok(int rc, const char* what) {
  if (rc != 0) throw Error(std::string("NCCL error in ") + what + ": " + nccl().GetErrorString(rc));
}

}  // namespace retr
