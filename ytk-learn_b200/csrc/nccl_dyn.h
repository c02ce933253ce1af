// NCCL is bound at run time (dlopen) so that libytkgbdt.so loads on hosts without it and a
// single-GPU tree maker never touches it.  Inside a PyTorch process this resolves to the
// libnccl.so.2 torch already loaded; under a JVM to the system one.
#pragma once

#include <dlfcn.h>
#include <nccl.h>

namespace ytk {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                                  cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;

    bool load(const char** why) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) { *why = "libnccl.so.2 not found"; return false; }
#define YTK_NCCL_SYM(field, sym)                                              \
    field = reinterpret_cast<decltype(field)>(dlsym(lib, sym));               \
    if (!field) { *why = "missing NCCL symbol " sym; lib = nullptr; return false; }
        YTK_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        YTK_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        YTK_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        YTK_NCCL_SYM(ReduceScatter, "ncclReduceScatter")
        YTK_NCCL_SYM(AllGather, "ncclAllGather")
        YTK_NCCL_SYM(AllReduce, "ncclAllReduce")
        YTK_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef YTK_NCCL_SYM
        return true;
    }
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    return api;
}

}  // namespace ytk
