// rg_comm.h — the NCCL entry points the library uses, resolved at run time (rg_comm.cu).
#pragma once
#include <nccl.h>

#include <cstdint>
#include <string>

struct rg_handle;

struct NcclApi {
    bool ok = false;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

const NcclApi *rg_nccl(std::string *err);
int32_t rg_nccl_fail(rg_handle *h, int r, const char *what);
void rg_shard(uint64_t n, int rank, int world, uint64_t *begin, uint64_t *end);
int32_t rg_comm_agree(rg_handle *h, int32_t local_rc);

#define RG_NCCL(h, api, call)                                        \
    do {                                                             \
        ncclResult_t r__ = (api)->call;                              \
        if (r__ != ncclSuccess) return rg_nccl_fail((h), r__, #call); \
    } while (0)
