// srt_comm.h -- NCCL entry points for the multi-device calls of the C ABI (srt_comm_*, srt_reduce_accum,
// srt_gather_*), resolved at run time.
//
// The reference spreads a frame over worker threads with WorkQueueCreate / WorkQueueStart(..., ThreadCount)
// (base.h:172-178, main.cpp:660, 851-856); here the workers are GPUs, one handle each, and the only exchange
// step of the path -- combining the per-device images -- is an NCCL reduce issued on the handle's own stream.
//
// libnccl is NOT a link-time dependency of libsrt_cuda.so: a single-GPU user never needs it, and a process that
// already carries an NCCL (PyTorch bundles its own libnccl.so.2) must keep using that one.  The library already
// mapped into the process is preferred (RTLD_NOLOAD), then the loader's search path; SRT_NCCL_LIB=<path> names one
// explicitly.  Without NCCL the srt_comm_* calls fail with SRT_ERR_COMM and a message; nothing falls back to a host copy.
#pragma once
#include <dlfcn.h>
#include <cstdlib>
#include <nccl.h>

#include <mutex>
#include <string>

namespace srt {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    void *handle = nullptr;
    std::string error;
};

inline NcclApi &nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        if (const char *forced = getenv("SRT_NCCL_LIB")) {            // an explicit library path: nothing else is tried
            api.handle = dlopen(forced, RTLD_NOW | RTLD_GLOBAL);
            if (!api.handle) { const char *e = dlerror(); api.error = std::string("SRT_NCCL_LIB=") + forced + " cannot be loaded: " + (e ? e : "?"); return; }
        }
        if (!api.handle)
            for (const char *n : names) { api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (api.handle) break; }
        if (!api.handle)
            for (const char *n : names) { api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (api.handle) break; }
        if (!api.handle) { const char *e = dlerror(); api.error = std::string("libnccl.so.2 not found: ") + (e ? e : "?"); return; }
        auto sym = [&](const char *name) -> void * {
            void *p = dlsym(api.handle, name);
            if (!p && api.error.empty()) api.error = std::string("libnccl: missing symbol ") + name;
            return p;
        };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    });
    return api;
}

} // namespace srt
