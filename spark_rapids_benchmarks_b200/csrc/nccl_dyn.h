// NCCL, bound at run time.  The engine's exchange operator (exec.cu) all-gathers - or hash-partitions, all-to-all - partial aggregate states between
// the per-GPU processes of one node over NVLink / NVSwitch.  libnccl.so.2 is opened with dlopen when a communicator
// is first requested - inside a torchrun process that is the copy torch already loaded - so the library has no
// link-time NCCL dependency and single-GPU use never touches it.  Only the handful of entry points below are used;
// their signatures have been stable across NCCL 2.x (ncclUniqueId is 128 opaque bytes, passed by value).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace nds {

struct NcclUniqueId { char internal[128]; };

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(void** comm, int nranks, NcclUniqueId id, int rank) = nullptr;
  int (*CommDestroy)(void* comm) = nullptr;
  int (*AllGather)(const void* send, void* recv, size_t count, int datatype, void* comm, cudaStream_t stream) = nullptr;
  // the hash-partition exchange: one grouped send / receive per peer (all-to-all with per-peer sizes)
  int (*Send)(const void* send, size_t count, int datatype, int peer, void* comm, cudaStream_t stream) = nullptr;
  int (*Recv)(void* recv, size_t count, int datatype, int peer, void* comm, cudaStream_t stream) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
};

// Loads the library on first use; throws (common.h fail) when it cannot be found or lacks a symbol.
const NcclApi& nccl_api();
void nccl_check(int rc, const char* what);

constexpr int kNcclChar = 0;      // ncclInt8 / ncclChar: every transfer is counted in bytes

}  // namespace nds
