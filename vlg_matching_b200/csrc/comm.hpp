// comm.hpp — the one collective of the sharded search: an NCCL all-gather of 24 bytes per query
// and rank, plus the host logic that turns the gathered rows into every shard's true entry cursor.
//
// NCCL is bound at run time (dlopen "libnccl.so.2": the copy torch has already mapped when the caller is a Python
// process, the system one for the C++ drivers), so libvlgb200.so loads on boxes without it and single-GPU
// callers never touch it.  <nccl.h> is used for its types only.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include "common.cuh"

namespace vlgb {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

inline NcclApi& nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return api;
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(api.handle, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(api.handle, "ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(api.handle, "ncclCommDestroy"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(dlsym(api.handle, "ncclAllGather"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(api.handle, "ncclGetErrorString"));
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString;
    return api;
}

// per-thread communicator (one host thread drives one GPU = one rank)
struct Comm {
    ncclComm_t comm = nullptr;
    int n_ranks = 1, rank = 0;
    DeviceBuffer dev;      // send row | gathered table
    PinnedBuffer host;     // same, pinned
};
inline Comm& comm()
{
    static thread_local Comm c;
    return c;
}

// ---- host logic of the exchange (exported as vlgb_chain_plan so that it is testable without a GPU) --------------
// rows: [n_shards][nq][3] = (match count, exit cursor, first start) as every shard reported them;
// used: [n_shards][nq] = the entry cursor each row was computed with.
// entries[g][q] = cursor the nearest shard to the left of g that reported a match left behind (0 for shard 0) =
// pull_forward state handed over the boundary (vlg_index.hpp:253-264).  A shard must search query q again unless
// its cursor is unchanged, or moved right without passing its first reported start (the orbit of the monotone
// successor depends on the entry only through its first element).
inline uint32_t chain_plan(const uint64_t* rows, const uint64_t* used, uint32_t n_shards, uint32_t nq, uint64_t* entries,
                           uint8_t* redo)
{
    uint32_t n_redo = 0;
    for (uint32_t g = 0; g < n_shards; g++) redo[g] = 0;
    for (uint32_t q = 0; q < nq; q++) {
        uint64_t cursor = 0;
        for (uint32_t g = 0; g < n_shards; g++) {
            const uint64_t* row = rows + ((size_t)g * nq + q) * 3;
            entries[(size_t)g * nq + q] = cursor;
            const uint64_t e_old = used[(size_t)g * nq + q];
            if (cursor != e_old && !(cursor > e_old && (row[0] == 0 || cursor <= row[2]))) redo[g] = 1;
            if (row[0] > 0) cursor = row[1];
        }
    }
    for (uint32_t g = 0; g < n_shards; g++) n_redo += redo[g];
    return n_redo;
}

}  // namespace vlgb
