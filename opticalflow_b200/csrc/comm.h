// ofb200 - NCCL binding for the slab-decomposed solve (host side only).
//
// The library does not link against NCCL: pair-sharded runs never touch it.  For the slab configuration the
// symbols are resolved at run time from the libnccl.so.2 that the host process has already loaded (PyTorch's
// copy), falling back to the system library.  One communicator per process; rank r talks to r-1 and r+1
// (halo rows) and to everyone (scalar all-reduce, coarse-level all-gather) over NVLink / NVSwitch.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <string.h>

namespace ofb {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
    ncclResult_t (*GetVersion)(int*);
    void* handle;
    bool ok;
};

inline NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        memset(&a, 0, sizeof(a));
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);     // the copy the process already uses
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        a.handle = h;
        if (!h) return a;
        bool ok = true;
        auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) ok = false; return p; };
        a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.Send = (decltype(a.Send))sym("ncclSend");
        a.Recv = (decltype(a.Recv))sym("ncclRecv");
        a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
        a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
        a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.GetVersion = (decltype(a.GetVersion))sym("ncclGetVersion");
        a.ok = ok;
        return a;
    }();
    return api;
}

// Mailbox for halo rows written directly by the neighbouring GPU (peer stores over NVLink, CUDA IPC mapping).
// Layout of a rank's mailbox:  4 x u64 arrival flags [dir][parity] | int error | pad to 256 | 4 data slots [dir][parity].
// dir 0 = rows arriving from rank-1, dir 1 = rows arriving from rank+1.  Exchange number `seq` (identical on all
// ranks) uses the slots of parity seq & 1: a rank can be at most one exchange ahead of its neighbour, so two slots
// per direction are enough.
struct PeerBox {
    unsigned char* local;           // this rank's mailbox
    unsigned char* peer[2];         // rank-1's / rank+1's mailbox mapped into this process (nullptr: no such neighbour)
    size_t slot_bytes;
    unsigned long long seq;         // exchanges issued so far
    bool ok;
};
constexpr size_t PEERBOX_HEADER = 256;
constexpr size_t PEERBOX_SLOT_BYTES = 4u << 20;      // per (direction, parity): 7 planes x 32768 columns x fp32 and more

struct Comm {
    ncclComm_t comm;
    int rank, world;
    PeerBox box;
};

template <typename T> struct NcclType;
template <> struct NcclType<float>  { static constexpr ncclDataType_t value = ncclFloat32; };
template <> struct NcclType<double> { static constexpr ncclDataType_t value = ncclFloat64; };

}  // namespace ofb
