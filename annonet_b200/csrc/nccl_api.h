// libnccl.so.2 loaded at run time (the library ships inside the PyTorch wheel; no link-time dependency): the entry points
// the engine uses for the gradient all-reduce, the start-up broadcast and the seam / label exchange of sharded inference.
#pragma once
#include "engine.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <string>

namespace ab {

struct NcclApi {
    struct UniqueId { char internal[128]; };
    void* lib = nullptr;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(void**, int, UniqueId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    NcclApi()
    {
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) throw state_error(std::string("cannot load libnccl.so.2: ") + dlerror());
        GetUniqueId = (decltype(GetUniqueId))dlsym(lib, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(lib, "ncclCommInitRank");
        AllReduce = (decltype(AllReduce))dlsym(lib, "ncclAllReduce");
        Broadcast = (decltype(Broadcast))dlsym(lib, "ncclBroadcast");
        Send = (decltype(Send))dlsym(lib, "ncclSend");
        Recv = (decltype(Recv))dlsym(lib, "ncclRecv");
        GroupStart = (decltype(GroupStart))dlsym(lib, "ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))dlsym(lib, "ncclGroupEnd");
        CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
        GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !AllReduce || !CommDestroy || !Broadcast || !Send || !Recv || !GroupStart || !GroupEnd)
            throw state_error("libnccl lacks the expected symbols");
    }
    void check(int r, const char* what) const
    {
        if (r != 0) throw cuda_error(std::string("NCCL error in ") + what + ": " + (GetErrorString ? GetErrorString(r) : "?"));
    }
};

}  // namespace ab
