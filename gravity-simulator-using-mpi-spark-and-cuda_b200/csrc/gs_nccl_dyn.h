// gs_nccl_dyn.h -- NCCL is loaded lazily (dlopen) and only when world_size > 1, so the single-GPU
// path and the CPU-only symbol checks never depend on it.  If the process already holds a
// libnccl.so.2 (e.g. torch's bundled copy) the dynamic loader hands back that same library.
#pragma once
#include <dlfcn.h>
#include <mutex>
#include <nccl.h>

namespace gs {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

inline void nccl_api_load(NcclApi& api) {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return;
#define GS_SYM(field, name) *(void**)(&api.field) = dlsym(api.handle, name)
    GS_SYM(GetUniqueId, "ncclGetUniqueId");
    GS_SYM(CommInitRank, "ncclCommInitRank");
    GS_SYM(CommDestroy, "ncclCommDestroy");
    GS_SYM(AllGather, "ncclAllGather");
    GS_SYM(AllReduce, "ncclAllReduce");
    GS_SYM(GetErrorString, "ncclGetErrorString");
#undef GS_SYM
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce;
}

// Loaded once, whichever rank thread gets here first (one thread per GPU in the CLI and the tests).
inline NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] { nccl_api_load(api); });
    return api;
}

}  // namespace gs
