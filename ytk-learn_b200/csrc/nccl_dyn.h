// NCCL is bound at run time (dlopen) so that libytkgbdt.so loads on hosts without it and a
// single-GPU tree maker never touches it.  Inside a PyTorch process this resolves to the
// libnccl.so.2 torch already loaded; under a JVM to the system one.
#pragma once

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

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

    // Several threads create their handles at the same time (ncclCommInitRank blocks until every rank
    // has joined), so loading is serialised and `lib` is published only once every symbol is bound.
    std::mutex mu;
    bool load(const char** why) {
        std::lock_guard<std::mutex> lk(mu);
        if (lib) return true;
        void* l = nullptr;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            l = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (l) break;
        }
        if (!l) { *why = "libnccl.so.2 not found"; return false; }
#define YTK_NCCL_SYM(field, sym)                                              \
    field = reinterpret_cast<decltype(field)>(dlsym(l, sym));                 \
    if (!field) { *why = "missing NCCL symbol " sym; return false; }
        YTK_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        YTK_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        YTK_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        YTK_NCCL_SYM(ReduceScatter, "ncclReduceScatter")
        YTK_NCCL_SYM(AllGather, "ncclAllGather")
        YTK_NCCL_SYM(AllReduce, "ncclAllReduce")
        YTK_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef YTK_NCCL_SYM
        lib = l;
        return true;
    }
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    return api;
}

}  // namespace ytk
