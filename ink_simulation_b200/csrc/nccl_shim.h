// nccl_shim.h — NCCL entry points resolved at run time (dlopen), so libinkb200.so has no link-time NCCL dependency:
// a single-GPU process never touches NCCL, and a multi-GPU process (one rank per GPU, launched by torchrun) shares the
// libnccl.so.2 that torch already loaded (RTLD_NOLOAD first) instead of pulling in a second copy.
// Types come from <nccl.h> at compile time only.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <string>

namespace ink {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;

    // returns "" on success, else what failed
    std::string load() {
        if (handle) return "";
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD);   // the copy torch (or the host program) already loaded
            if (handle) break;
        }
        if (!handle)
            for (const char* n : names) {
                handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
                if (handle) break;
            }
        if (!handle) return std::string("cannot dlopen libnccl.so.2: ") + dlerror();
#define INK_SYM(field, sym)                                              \
        *(void**)(&field) = dlsym(handle, sym);                          \
        if (!field) return std::string("libnccl is missing symbol ") + sym;
        INK_SYM(GetUniqueId, "ncclGetUniqueId")
        INK_SYM(CommInitRank, "ncclCommInitRank")
        INK_SYM(CommDestroy, "ncclCommDestroy")
        INK_SYM(AllReduce, "ncclAllReduce")
        INK_SYM(AllGather, "ncclAllGather")
        INK_SYM(Broadcast, "ncclBroadcast")
        INK_SYM(Send, "ncclSend")
        INK_SYM(Recv, "ncclRecv")
        INK_SYM(GroupStart, "ncclGroupStart")
        INK_SYM(GroupEnd, "ncclGroupEnd")
        INK_SYM(GetErrorString, "ncclGetErrorString")
#undef INK_SYM
        return "";
    }
};

inline NcclApi& nccl() {
    static NcclApi api;
    return api;
}

}  // namespace ink
