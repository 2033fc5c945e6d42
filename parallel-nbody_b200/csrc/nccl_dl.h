// nccl_dl.h -- NCCL bound at run time with dlopen, so libnbody_gpu.so has no link-time
// NCCL dependency (single-GPU users never touch it) and, inside a Python process that
// already imported torch, resolves to the very libnccl.so.2 torch loaded.
// Only the handful of entry points the position exchange needs; their C signatures are
// stable across NCCL 2.x (this image: system 2.27.3, torch-bundled 2.28.9).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <cstddef>
#include <cstdlib>

namespace nbody {

struct NcclUniqueId { char internal[128]; };
typedef struct ncclComm *NcclComm;
enum { NCCL_SUCCESS = 0 };
enum { NCCL_DOUBLE = 8 };   // ncclDataType_t: ncclFloat64 / ncclDouble

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int) = nullptr;
    int (*CommInitAll)(NcclComm *, int, const int *) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
};

// Returns nullptr (and fills *why) if NCCL cannot be loaded.
inline const NcclApi *nccl_api(const char **why)
{
    static NcclApi api;
    static bool tried = false;
    static const char *err = "";
    if (!tried) {
        tried = true;
        // stdout belongs to the caller (the host driver prints the final state there): NCCL's own
        // banner/debug lines (NCCL_DEBUG=VERSION|INFO) go to stderr unless the user chose a file
        // (NCCL honours NCCL_DEBUG_FILE only above the VERSION level, so VERSION is raised to WARN,
        // which prints the same banner)
        const char *dbg = getenv("NCCL_DEBUG");
        if (dbg && (dbg[0] == 'V' || dbg[0] == 'v')) setenv("NCCL_DEBUG", "WARN", 1);
        setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            err = "cannot dlopen libnccl.so.2";
        } else {
#define NB_SYM(field, name)                                                      \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name));     \
    if (!api.field) err = "libnccl is missing " name;
            NB_SYM(GetUniqueId, "ncclGetUniqueId")
            NB_SYM(CommInitRank, "ncclCommInitRank")
            NB_SYM(CommInitAll, "ncclCommInitAll")
            NB_SYM(CommDestroy, "ncclCommDestroy")
            NB_SYM(AllGather, "ncclAllGather")
            NB_SYM(GroupStart, "ncclGroupStart")
            NB_SYM(GroupEnd, "ncclGroupEnd")
            NB_SYM(GetErrorString, "ncclGetErrorString")
            NB_SYM(GetVersion, "ncclGetVersion")
#undef NB_SYM
        }
    }
    if (why) *why = err;
    return (api.lib && err[0] == 0) ? &api : nullptr;
}

}  // namespace nbody
