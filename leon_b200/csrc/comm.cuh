// NCCL plumbing of the multi-GPU path (one process per GPU, NVLink 5 / NVSwitch).  The library does not link NCCL:
// it binds the copy that is already in the process (the host framework's, e.g. torch's bundled libnccl.so.2) or,
// failing that, the system one, so that a process never ends up with two NCCL runtimes.
#pragma once
#include "common.cuh"
#include <dlfcn.h>
#include <nccl.h>

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;

    bool load(std::string &err)
    {
        if (handle) return true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD);      // the copy already mapped into the process, if any
            if (handle) break;
        }
        for (int i = 0; !handle && i < 2; i++) handle = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (!handle) { err = std::string("NCCL not found: ") + dlerror(); return false; }
#define LB_SYM(field, sym)                                                                  \
    *(void **)(&field) = dlsym(handle, sym);                                                \
    if (!field) { err = std::string("NCCL symbol missing: ") + sym; handle = nullptr; return false; }
        LB_SYM(GetUniqueId, "ncclGetUniqueId")
        LB_SYM(CommInitRank, "ncclCommInitRank")
        LB_SYM(CommDestroy, "ncclCommDestroy")
        LB_SYM(AllGather, "ncclAllGather")
        LB_SYM(AllReduce, "ncclAllReduce")
        LB_SYM(Broadcast, "ncclBroadcast")
        LB_SYM(Send, "ncclSend")
        LB_SYM(Recv, "ncclRecv")
        LB_SYM(GroupStart, "ncclGroupStart")
        LB_SYM(GroupEnd, "ncclGroupEnd")
        LB_SYM(GetErrorString, "ncclGetErrorString")
#undef LB_SYM
        return true;
    }
};

inline NcclApi &nccl_api()
{
    static NcclApi api;
    return api;
}

#define NCCL_TRY(expr)                                                                                   \
    do {                                                                                                 \
        ncclResult_t _r = (expr);                                                                        \
        if (_r != ncclSuccess) {                                                                         \
            char _b[512];                                                                                \
            snprintf(_b, sizeof _b, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, nccl_api().GetErrorString(_r)); \
            throw LeonError(LEON_E_CUDA, _b);                                                            \
        }                                                                                                \
    } while (0)

// communicator of a context: rank r of n, all on one node
struct Comm {
    ncclComm_t comm = nullptr;
    int rank = 0, n = 1;
    bool active() const { return n > 1; }
};
