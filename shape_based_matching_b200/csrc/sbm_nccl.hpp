// sbm_nccl.hpp -- NCCL, bound at run time.
//
// The multi-GPU entry points of the C ABI (sbm_group_*) exchange the compact candidate lists with ONE ncclAllGather
// over NVLink / NVSwitch.  libnccl is NOT a link-time dependency of libsbm_b200.so: a single-GPU caller never touches
// it, so the library loads on boxes without NCCL.  The first sbm_group_* call resolves the six functions it needs from
// the libnccl.so.2 already present in the process (a Python caller has torch's) or, failing that, from the system's.
// Only the stable part of the NCCL 2.x ABI is declared here (same declarations as /usr/include/nccl.h).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdlib>
#include <mutex>
#include <string>

namespace sbm
{

struct NcclApi
{
    typedef struct ncclComm *comm_t;
    struct unique_id
    {
        char internal[128]; // NCCL_UNIQUE_ID_BYTES
    };
    enum
    {
        kUint8 = 1 // ncclUint8
    };
    int (*GetVersion)(int *) = nullptr;
    int (*GetUniqueId)(unique_id *) = nullptr;
    int (*CommInitRank)(comm_t *, int, unique_id, int) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, comm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    void *handle = nullptr;
    std::string error;

    // nullptr when libnccl cannot be found; loadError() then says why
    static NcclApi *get()
    {
        NcclApi &api = self();
        static std::once_flag once;
        std::call_once(once, [&api] { api.load(); });
        return api.handle ? &api : nullptr;
    }
    static const std::string &loadError() { return self().error; }

private:
    static NcclApi &self()
    {
        static NcclApi api;
        return api;
    }
    void load()
    {
        const char *names[] = {getenv("SBM_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        void *h = nullptr;
        // the copy already mapped into the process first (RTLD_NOLOAD): two NCCL builds in one process is asking for trouble
        for (const char *n : names)
            if (n && !h)
                h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        for (const char *n : names)
            if (n && !h)
                h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (!h)
        {
            const char *e = dlerror();
            error = std::string("libnccl.so.2 not found (") + (e ? e : "?") + "); set SBM_NCCL_LIB";
            return;
        }
#define SBM_NCCL_SYM(field, name)                                   \
    *(void **)(&field) = dlsym(h, name);                            \
    if (!field)                                                     \
    {                                                               \
        error = std::string("libnccl lacks ") + name;               \
        return;                                                     \
    }
        SBM_NCCL_SYM(GetVersion, "ncclGetVersion")
        SBM_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        SBM_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        SBM_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        SBM_NCCL_SYM(AllGather, "ncclAllGather")
        SBM_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef SBM_NCCL_SYM
        handle = h;
    }
};

} // namespace sbm
