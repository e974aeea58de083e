// parsim_nccl.hpp -- NCCL bound at run time (dlopen), so the single-GPU path has no NCCL
// dependency and a multi-process job shares whatever libnccl.so.2 the process already has
// (e.g. the one torch.distributed loaded).  Only the stable NCCL 2.x C ABI is used.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdlib.h>
#include <strings.h>

#include <string>

namespace parsim {

struct NcclUniqueId { char internal[128]; };
typedef struct ncclComm *ncclComm_t;
enum { kNcclInt32 = 2, kNcclInt64 = 4, kNcclFloat64 = 8, kNcclChar = 0 };
enum { kNcclSum = 0 };

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string error;

    bool load() {
        if (lib) return true;
        // NCCL writes its version banner / warnings to STDOUT unless told otherwise; stdout of the
        // drop-in CLI is exactly two lines (serial/parsim.cpp:450-453), so send them to stderr
        // (read by NCCL at its first call; an explicit NCCL_DEBUG_FILE of the user wins)
        setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);
        // NCCL honours NCCL_DEBUG_FILE only above the VERSION level: WARN prints the same banner
        // plus warnings, into the file
        if (const char *lvl = getenv("NCCL_DEBUG"))
            if (strcasecmp(lvl, "VERSION") == 0) setenv("NCCL_DEBUG", "WARN", 1);
        const char *names[] = {"libnccl.so.2", "libnccl.so", nullptr};
        for (int i = 0; names[i] && !lib; i++) lib = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (!lib) {
            error = std::string("cannot load libnccl.so.2: ") + dlerror();
            return false;
        }
#define PS_SYM(field, name)                                              \
    *(void **)(&field) = dlsym(lib, name);                               \
    if (!field) { error = std::string("libnccl lacks ") + name; return false; }
        PS_SYM(GetUniqueId, "ncclGetUniqueId");
        PS_SYM(CommInitRank, "ncclCommInitRank");
        PS_SYM(CommDestroy, "ncclCommDestroy");
        PS_SYM(Send, "ncclSend");
        PS_SYM(Recv, "ncclRecv");
        PS_SYM(AllReduce, "ncclAllReduce");
        PS_SYM(AllGather, "ncclAllGather");
        PS_SYM(GroupStart, "ncclGroupStart");
        PS_SYM(GroupEnd, "ncclGroupEnd");
        PS_SYM(GetErrorString, "ncclGetErrorString");
#undef PS_SYM
        return true;
    }
};

inline NcclApi &nccl_api() {
    static NcclApi api;
    return api;
}

struct NcclComm {
    ncclComm_t comm = nullptr;
};

}  // namespace parsim
