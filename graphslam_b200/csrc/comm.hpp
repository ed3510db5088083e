// NCCL binding for the multi-GPU path (one process per GPU).  The library is NOT linked against
// NCCL: libnccl.so.2 is resolved at run time with dlopen, so a single-GPU user needs no NCCL at
// all, and inside a PyTorch process the loader hands back the copy torch already mapped (same
// soname) instead of a second one.  Only the handful of entry points the path uses are bound:
// halo exchange = grouped ncclSend/ncclRecv, scalar reductions = ncclAllGather of segment sums
// (summed in fixed order by every rank), pose read-back = grouped ncclBroadcast; ncclCommGetAsyncError is polled
// once per PCG batch so that a failed peer surfaces as GS_ENCCL instead of a hang.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <string>

namespace gs {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t*) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;

  // returns "" on success, else what is missing
  std::string load() {
    if (handle) return "";
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (handle) break;
    }
    if (!handle) return std::string("dlopen(libnccl.so.2) failed: ") + dlerror();
    struct {
      const char* name;
      void** slot;
    } syms[] = {{"ncclGetVersion", (void**)&GetVersion},     {"ncclGetUniqueId", (void**)&GetUniqueId},
                {"ncclCommInitRank", (void**)&CommInitRank}, {"ncclCommDestroy", (void**)&CommDestroy},
                {"ncclCommGetAsyncError", (void**)&CommGetAsyncError},
                {"ncclSend", (void**)&Send},                 {"ncclRecv", (void**)&Recv},
                {"ncclAllGather", (void**)&AllGather},       {"ncclBroadcast", (void**)&Broadcast},
                {"ncclGroupStart", (void**)&GroupStart},     {"ncclGroupEnd", (void**)&GroupEnd},
                {"ncclGetErrorString", (void**)&GetErrorString}};
    for (auto& s : syms) {
      *s.slot = dlsym(handle, s.name);
      if (!*s.slot) {
        handle = nullptr;
        return std::string("libnccl has no symbol ") + s.name;
      }
    }
    return "";
  }
};

inline NcclApi& nccl() {
  static NcclApi api;
  return api;
}

}  // namespace gs
