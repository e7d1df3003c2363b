// sba_comm.h — NCCL all-reduce plumbing for point-sharded global BA.
// The reference is single-process (no counterpart).  NCCL is bound at run time (dlopen of
// libnccl.so.2 — the copy torch already mapped when the host side is Python), so libsba.so has no
// link-time NCCL dependency and single-GPU use needs no NCCL at all.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cstdint>
#include <string>

namespace sba {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool load(std::string* err) {
    if (lib) return true;
    lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { *err = std::string("dlopen libnccl.so.2: ") + dlerror(); return false; }
    GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
    CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
    AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
    GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce) { *err = "libnccl: missing symbols"; return false; }
    return true;
  }
};

inline NcclApi& nccl_api() {
  static NcclApi api;
  return api;
}

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  bool active() const { return comm != nullptr && world > 1; }
  void destroy() {
    if (comm) { nccl_api().CommDestroy(comm); comm = nullptr; }
    world = 1; rank = 0;
  }
  // in-place fp64 all-reduce on `stream`; returns false + err on failure
  bool allreduce(double* buf, size_t n, bool is_max, cudaStream_t stream, std::string* err) {
    if (!active() || n == 0) return true;
    ncclResult_t r = nccl_api().AllReduce(buf, buf, n, ncclFloat64, is_max ? ncclMax : ncclSum, comm, stream);
    if (r != ncclSuccess) { *err = std::string("ncclAllReduce: ") + (nccl_api().GetErrorString ? nccl_api().GetErrorString(r) : "error"); return false; }
    return true;
  }
};

}  // namespace sba
