/*
 * nccl_dyn.h -- NCCL bound at run time with dlopen, only when more than one process shares a
 * particle set (or NB_SIM_FLAG_NO_PEER_STORE asks for it).  A single-GPU World never touches
 * NCCL, and libnbody.so has no link-time dependency on it.  When the host program already
 * loaded a libnccl.so.2 (PyTorch bundles one), dlopen returns that same copy.
 *
 * Only the handful of entry points the position exchange needs are declared; types are the
 * NCCL ABI (ncclComm_t is an opaque pointer, ncclUniqueId is 128 bytes, ncclFloat == 7).
 */
#pragma once

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>

#include "util.cuh"

namespace nb {

typedef struct ncclComm *NcclComm;
struct NcclUniqueId {
    char internal[128];
};
enum { kNcclSuccess = 0, kNcclFloat = 7 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId *) = nullptr;
    int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int) = nullptr;
    int (*CommInitAll)(NcclComm *, int, const int *) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};

inline NcclApi &nccl() {
    static NcclApi api;
    if (api.handle) return api;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    NB_ASSERT(api.handle != nullptr, "multi-GPU exchange needs NCCL but libnccl.so.2 cannot be loaded: %s", dlerror());
#define NB_NCCL_SYM(field, sym)                                                   \
    do {                                                                          \
        *(void **)(&api.field) = dlsym(api.handle, sym);                          \
        NB_ASSERT(api.field != nullptr, "libnccl lacks %s", sym);                 \
    } while (0)
    NB_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
    NB_NCCL_SYM(CommInitRank, "ncclCommInitRank");
    NB_NCCL_SYM(CommInitAll, "ncclCommInitAll");
    NB_NCCL_SYM(CommDestroy, "ncclCommDestroy");
    NB_NCCL_SYM(Broadcast, "ncclBroadcast");
    NB_NCCL_SYM(GroupStart, "ncclGroupStart");
    NB_NCCL_SYM(GroupEnd, "ncclGroupEnd");
    NB_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef NB_NCCL_SYM
    return api;
}

#define NB_NCCL(X)                                                                              \
    do {                                                                                        \
        int nb_nccl_err_ = (X);                                                                 \
        if (nb_nccl_err_ != nb::kNcclSuccess)                                                   \
            NB_FAIL("ncclResult = %d, str = %s: %s", nb_nccl_err_,                               \
                    nb::nccl().GetErrorString(nb_nccl_err_), #X);                               \
    } while (0)

}  // namespace nb
