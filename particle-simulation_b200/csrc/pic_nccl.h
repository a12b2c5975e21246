// pic_nccl.h -- NCCL bound at run time (dlopen), so that single-GPU runs carry no NCCL
// dependency and a host process that already loaded an NCCL (e.g. the one bundled with
// PyTorch) shares that copy instead of pulling in a second one.
//
// The reference has no multi-GPU path at all (device 0 is hard-coded, src/pic.cu:394); this
// is the one collective the new engine adds: an int32 sum of the charge box per Poisson step.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>
#include <string>

namespace picb200 {

struct NcclUniqueId { char internal[128]; };       // layout of ncclUniqueId (nccl.h)
constexpr int kNcclInt32 = 2;                      // ncclInt32
constexpr int kNcclSum = 0, kNcclMax = 2;          // ncclSum, ncclMax

struct NcclApi {
    int (*getUniqueId)(NcclUniqueId*);
    int (*commInitRank)(void** comm, int nranks, NcclUniqueId id, int rank);
    int (*commDestroy)(void* comm);
    int (*allReduce)(const void* send, void* recv, size_t count, int dtype, int op, void* comm, cudaStream_t stream);
    const char* (*getErrorString)(int);
};

inline NcclApi* load_nccl(std::string& err) {
    static NcclApi api;
    static bool loaded = false;
    if (loaded) return &api;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* nm : names) {
        h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        err = std::string("CUDA error: cannot load libnccl.so.2 (") + dlerror() + ")\nAt: NCCL init\n";
        return nullptr;
    }
    api.getUniqueId = reinterpret_cast<int (*)(NcclUniqueId*)>(dlsym(h, "ncclGetUniqueId"));
    api.commInitRank = reinterpret_cast<int (*)(void**, int, NcclUniqueId, int)>(dlsym(h, "ncclCommInitRank"));
    api.commDestroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
    api.allReduce = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(h, "ncclAllReduce"));
    api.getErrorString = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
    if (!api.getUniqueId || !api.commInitRank || !api.commDestroy || !api.allReduce || !api.getErrorString) {
        err = "CUDA error: libnccl.so.2 lacks a required symbol\nAt: NCCL init\n";
        return nullptr;
    }
    loaded = true;
    return &api;
}

}  // namespace picb200
