// graph_static.cu — rg_graph_mutual_static: rg_graph_mutual (graph_ops.cu; mutual_knn,
// src/graphs/nearest_neighbour_graph.rs:12-35) for callers that must not allocate device memory (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The device staging that rg_graph_mutual takes from the handle's allocator (the uploaded lists and the keep mask)
// lives here in ONE file-scope `static __device__` array of fixed size: RG_STATIC_MAX_EDGES edges, 5 bytes each
// (5 MiB per GPU that loads the library, held until the process exits).  The entry point calls no cudaMalloc,
// cudaFree or allocating library routine; larger inputs are RG_ERR_INVALID.
//
// In host code `g_scratch` is only a host-side stand-in for the symbol, not a device address: the address is taken
// with cudaGetSymbolAddress once per device (with that device current) and that pointer is what the kernel gets.
// Every call on a device shares the array and finds the previous call's contents in it; the mask region is cleared
// before the kernel.  Two calls that overlap on the same GPU (different handles / streams / threads) race on it.
#include <mutex>

#include "rg_internal.cuh"

namespace {

constexpr uint64_t SCRATCH_IDX_BYTES = (uint64_t)RG_STATIC_MAX_EDGES * sizeof(uint32_t);
constexpr uint64_t SCRATCH_BYTES = SCRATCH_IDX_BYTES + RG_STATIC_MAX_EDGES;          // lists, then the keep mask

static __device__ __align__(256) unsigned char g_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char *g_scratch_addr[MAX_DEVICES] = {};          // device address of g_scratch on device i, once resolved

// set-up, once per device: the device address of the array on h->device (made current by the caller)
int32_t scratch_of(rg_handle *h, unsigned char **out) {
    if (h->device < 0 || h->device >= MAX_DEVICES) return rg_fail(h, RG_ERR_INVALID, "rg_graph_mutual_static: device %d out of range", h->device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[h->device]) {
        unsigned char *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_scratch);
        if (e != cudaSuccess || !p)
            return rg_fail(h, RG_ERR_CUDA, "rg_graph_mutual_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_scratch_addr[h->device] = p;
    }
    *out = g_scratch_addr[h->device];
    return RG_OK;
}

// one thread per edge; the reverse edge is looked up in the same graph (mutual_knn :24)
__global__ void mutual_keep_static_kernel(const uint32_t *__restrict__ idx, uint64_t n, uint32_t k, uint32_t keep_n_edges,
                                          uint8_t *__restrict__ keep) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * k) return;
    const uint32_t node = (uint32_t)(e / k), j = (uint32_t)(e % k);
    const uint32_t nb = idx[e];
    if (nb == 0xffffffffu) return;                             // padding: the cleared mask already says 0
    bool kp = j < keep_n_edges;                                // :19-22
    if (!kp && nb < n) {
        const uint32_t *row = idx + (uint64_t)nb * k;
        for (uint32_t q = 0; q < k && !kp; ++q) {
            const uint32_t v = row[q];
            if (v == 0xffffffffu) break;
            kp = v == node;
        }
    }
    if (kp) keep[e] = 1;
}

}  // namespace

extern "C" int32_t rg_graph_mutual_static(rg_handle *h, const uint32_t *idx, uint64_t n, uint32_t k, uint32_t keep_n_edges,
                                          uint8_t *keep) {
    if (!h) return RG_ERR_INVALID;
    if (!idx || !keep) return rg_fail(h, RG_ERR_INVALID, "rg_graph_mutual_static: NULL pointer");
    if (n == 0 || k == 0) return RG_OK;
    if (n > RG_STATIC_MAX_EDGES || k > RG_STATIC_MAX_EDGES || n * k > RG_STATIC_MAX_EDGES)
        return rg_fail(h, RG_ERR_INVALID, "rg_graph_mutual_static: %llu x %u edges exceed the static scratch (RG_STATIC_MAX_EDGES = %llu); use rg_graph_mutual",
                       (unsigned long long)n, k, (unsigned long long)RG_STATIC_MAX_EDGES);
    RG_CUDA(h, cudaSetDevice(h->device));
    unsigned char *scratch = nullptr;
    RG_TRY(scratch_of(h, &scratch));
    const uint64_t edges = n * k;
    uint32_t *d_idx = (uint32_t *)scratch;
    uint8_t *d_keep = scratch + SCRATCH_IDX_BYTES;
    RG_CUDA(h, cudaMemcpyAsync(d_idx, idx, edges * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemsetAsync(d_keep, 0, edges, h->stream));  // the last call's mask is still there
    mutual_keep_static_kernel<<<(unsigned)((edges + 255) / 256), 256, 0, h->stream>>>(d_idx, n, k, keep_n_edges, d_keep);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    RG_CUDA(h, cudaMemcpyAsync(keep, d_keep, edges, cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    return RG_OK;
}
