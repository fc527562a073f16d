// wl_nccl.cuh -- NCCL bound at run time (dlopen), only when a slab-decomposed sim is created.
// Linking libnccl would pin whichever copy the loader finds first; a host process that also uses
// another NCCL client (e.g. PyTorch, which bundles its own libnccl.so.2) must share ONE copy, so
// the library adopts the copy already mapped into the process and loads one only if there is none.
// Override with WL_NCCL_LIB=/path/to/libnccl.so.2.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <string>

namespace wl {

struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  int (*GetVersion_)(int*) = nullptr;
};

inline NcclApi& nccl_api() { static NcclApi a; return a; }

// returns "" on success, else what went wrong
inline std::string nccl_load() {
  NcclApi& a = nccl_api();
  if (a.h) return "";
  void* h = nullptr;
  if (const char* p = getenv("WL_NCCL_LIB")) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // the copy the process already uses
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return std::string("cannot load libnccl.so.2: ") + dlerror();
  bool ok = true;
  auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p) ok = false; return p; };
  a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
  a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
  a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
  a.Send = (decltype(a.Send))sym("ncclSend");
  a.Recv = (decltype(a.Recv))sym("ncclRecv");
  a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
  a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
  a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
  a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
  if (!ok) return "libnccl.so.2 lacks a required symbol";
  a.h = h;
  return "";
}

}  // namespace wl

// the rest of the library spells the calls as usual
#define ncclGetUniqueId wl::nccl_api().GetUniqueId
#define ncclCommInitRank wl::nccl_api().CommInitRank
#define ncclCommDestroy wl::nccl_api().CommDestroy
#define ncclSend wl::nccl_api().Send
#define ncclRecv wl::nccl_api().Recv
#define ncclAllGather wl::nccl_api().AllGather
#define ncclGroupStart wl::nccl_api().GroupStart
#define ncclGroupEnd wl::nccl_api().GroupEnd
#define ncclGetErrorString wl::nccl_api().GetErrorString
