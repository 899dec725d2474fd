// NCCL entry points resolved with dlopen at first use, so that libsart_b200.so has no link-time NCCL dependency:
// a single-GPU caller never needs the library, and a process that already carries an NCCL (e.g. one that imported
// torch) shares that copy. Types come from <nccl.h>; only the calls below are used.
#pragma once

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <string>

namespace sart {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;  // why loading failed, empty when usable
};

inline NcclApi& nccl_storage() {
    static NcclApi api;
    return api;
}

// Returns the loaded API, or nullptr with the reason in nccl_storage().error.
inline NcclApi* nccl_api() {
    static std::once_flag once;
    NcclApi& api = nccl_storage();
    std::call_once(once, [&api] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) {
            const char* why = dlerror();
            api.error = std::string("cannot load NCCL (libnccl.so.2): ") + (why ? why : "not found");
            return;
        }
        auto sym = [&api](const char* name) -> void* {
            void* p = dlsym(api.handle, name);
            if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + name;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    });
    return api.error.empty() ? &api : nullptr;
}

}  // namespace sart
