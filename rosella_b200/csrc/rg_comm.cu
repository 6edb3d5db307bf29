// rg_comm.cu — NCCL behind the C ABI.  The library is resolved with dlopen on first use (a process that already holds
// libnccl.so.2, e.g. through PyTorch, gets that copy back), so single-GPU callers need no NCCL at all and the types
// come from <nccl.h> at compile time only.
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "rg_comm.h"
#include "rg_internal.cuh"

namespace {

NcclApi g_api;
std::once_flag g_once;
std::string g_load_error;

void load_nccl() {
    // a copy the process already holds wins (PyTorch ships its own libnccl.so.2 and needs exactly that version; loading
    // the system copy first would shadow it for the rest of the process), then the loader's search path
    void *lib = nullptr;
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
        lib = dlopen(name, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (lib) break;
    }
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
        if (lib) break;
        lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!lib) {
        g_load_error = std::string("dlopen(libnccl.so.2) failed: ") + dlerror();
        return;
    }
#define RG_SYM(field, sym)                                             \
    g_api.field = reinterpret_cast<decltype(g_api.field)>(dlsym(lib, #sym)); \
    if (!g_api.field) { g_load_error = "libnccl has no symbol " #sym; return; }
    RG_SYM(GetUniqueId, ncclGetUniqueId)
    RG_SYM(CommInitRank, ncclCommInitRank)
    RG_SYM(CommInitAll, ncclCommInitAll)
    RG_SYM(CommDestroy, ncclCommDestroy)
    RG_SYM(GetErrorString, ncclGetErrorString)
    RG_SYM(GroupStart, ncclGroupStart)
    RG_SYM(GroupEnd, ncclGroupEnd)
    RG_SYM(Send, ncclSend)
    RG_SYM(Recv, ncclRecv)
    RG_SYM(Broadcast, ncclBroadcast)
    RG_SYM(AllReduce, ncclAllReduce)
    RG_SYM(GetVersion, ncclGetVersion)
#undef RG_SYM
    g_api.ok = true;
}

}  // namespace

const NcclApi *rg_nccl(std::string *err) {
    std::call_once(g_once, load_nccl);
    if (!g_api.ok) {
        if (err) *err = g_load_error;
        return nullptr;
    }
    return &g_api;
}

int32_t rg_nccl_fail(rg_handle *h, int r, const char *what) {
    const NcclApi *api = rg_nccl(nullptr);
    return rg_fail(h, RG_ERR_CUDA, "%s failed: %s", what, api ? api->GetErrorString((ncclResult_t)r) : "NCCL not loaded");
}

// rows of an n-row graph owned by `rank`: whole 32-row tiles (the CTAs of the select kernels), sizes differ by at
// most one tile
void rg_shard(uint64_t n, int rank, int world, uint64_t *begin, uint64_t *end) {
    const uint64_t tiles = (n + RG_TILE - 1) / RG_TILE;
    const uint64_t base = tiles / world, extra = tiles % world;
    const uint64_t t0 = (uint64_t)rank * base + std::min<uint64_t>(rank, extra);
    const uint64_t t1 = t0 + base + ((uint64_t)rank < extra ? 1 : 0);
    *begin = std::min(t0 * RG_TILE, n);
    *end = std::min(t1 * RG_TILE, n);
}

// Every rank contributes its local status; all ranks return the minimum (0 = go on).  Entered before collectives so
// that a rank which failed (allocation, bad argument) cannot leave the others waiting inside NCCL.
int32_t rg_comm_agree(rg_handle *h, int32_t local_rc) {
    if (h->world <= 1 || !h->comm) return local_rc;
    const NcclApi *api = rg_nccl(nullptr);
    if (rg_reserve(h, h->dist_status, 2 * sizeof(int32_t)) != RG_OK) return RG_ERR_OOM;   // cannot even vote; NCCL will time out
    int32_t *dev = (int32_t *)h->dist_status.p;
    cudaMemcpyAsync(dev, &local_rc, sizeof(int32_t), cudaMemcpyHostToDevice, h->stream);
    ncclResult_t r = api->AllReduce(dev, dev + 1, 1, ncclInt32, ncclMin, (ncclComm_t)h->comm, h->stream);
    if (r != ncclSuccess) return rg_nccl_fail(h, r, "ncclAllReduce (status)");
    int32_t agreed = 0;
    cudaMemcpyAsync(&agreed, dev + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream);
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) return rg_fail(h, RG_ERR_CUDA, "status agreement: stream error");
    if (agreed != RG_OK && local_rc == RG_OK)
        return rg_fail(h, agreed, "another rank of the communicator failed before the collective (status %d)", agreed);
    return local_rc;
}

extern "C" {

int32_t rg_comm_unique_id(uint8_t id[RG_COMM_ID_BYTES]) {
    if (!id) return rg_fail(nullptr, RG_ERR_INVALID, "rg_comm_unique_id: id is NULL");
    std::string err;
    const NcclApi *api = rg_nccl(&err);
    if (!api) return rg_fail(nullptr, RG_ERR_STATE, "rg_comm_unique_id: %s", err.c_str());
    static_assert(sizeof(ncclUniqueId) == RG_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    ncclResult_t r = api->GetUniqueId(&u);
    if (r != ncclSuccess) return rg_fail(nullptr, RG_ERR_CUDA, "ncclGetUniqueId failed: %s", api->GetErrorString(r));
    memcpy(id, &u, RG_COMM_ID_BYTES);
    return RG_OK;
}

int32_t rg_comm_init(rg_handle *h, int32_t rank, int32_t world, const uint8_t id[RG_COMM_ID_BYTES]) {
    if (!h) return RG_ERR_INVALID;
    if (!id || world < 1 || rank < 0 || rank >= world) return rg_fail(h, RG_ERR_INVALID, "rg_comm_init: rank %d of %d", rank, world);
    if (h->comm) return rg_fail(h, RG_ERR_STATE, "rg_comm_init: the handle already has a communicator");
    std::string err;
    const NcclApi *api = rg_nccl(&err);
    if (!api) return rg_fail(h, RG_ERR_STATE, "rg_comm_init: %s", err.c_str());
    RG_CUDA(h, cudaSetDevice(h->device));
    ncclUniqueId u;
    memcpy(&u, id, RG_COMM_ID_BYTES);
    ncclComm_t comm = nullptr;
    ncclResult_t r = api->CommInitRank(&comm, world, u, rank);
    if (r != ncclSuccess) return rg_nccl_fail(h, r, "ncclCommInitRank");
    h->comm = comm;
    h->rank = rank;
    h->world = world;
    return RG_OK;
}

int32_t rg_comm_destroy(rg_handle *h) {
    if (!h) return RG_ERR_INVALID;
    if (h->comm) {
        const NcclApi *api = rg_nccl(nullptr);
        cudaSetDevice(h->device);
        if (api) api->CommDestroy((ncclComm_t)h->comm);
        h->comm = nullptr;
    }
    h->rank = 0;
    h->world = 1;
    return RG_OK;
}

int32_t rg_comm_shard(uint64_t n, int32_t rank, int32_t world, uint64_t *begin, uint64_t *end) {
    if (!begin || !end || world < 1 || rank < 0 || rank >= world) return RG_ERR_INVALID;
    rg_shard(n, rank, world, begin, end);
    return RG_OK;
}

}  // extern "C"
