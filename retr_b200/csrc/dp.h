// Data-parallel gradient exchange of the training step: NCCL over NVLink / NVSwitch, loaded at run time
// (the process that hosts the library normally has libnccl.so.2 mapped already — torch's bundled copy).  See dp.cu.
#pragma once
#include "common.cuh"

namespace retr {

// the few stable pieces of nccl.h this library needs (ABI-stable since NCCL 2.10)
typedef struct ncclComm* nccl_comm_t;
struct nccl_unique_id { char internal[128]; };
enum { NCCL_FLOAT32 = 7, NCCL_SUM = 0, NCCL_AVG = 4 };

struct NcclApi {
  int (*GetUniqueId)(nccl_unique_id*);
  int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int);
  int (*CommDestroy)(nccl_comm_t);
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t);
  int (*GroupStart)();
  int (*GroupEnd)();
  const char* (*GetErrorString)(int);
};
const NcclApi& nccl();                      // throws when libnccl.so.2 cannot be loaded
void nccl_ok(int rc, const char* what);

// One contiguous range of the flat gradient arena whose gradients are final at the same point of the backward pass
struct GradBucket {
  const char* name;
  int64_t lo = 0, hi = 0;                   // element range [lo, hi) of the arena
};

}  // namespace retr
