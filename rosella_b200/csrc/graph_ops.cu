// graph_ops.cu — what mutual_knn / intersect COMPUTE (src/graphs/nearest_neighbour_graph.rs:12-35,
// :38-72): per edge, keep it if it is among the first keep_n_edges of its node, else only if the
// reverse edge exists in the same / the other graph.  The reference discards these lists
// (SURVEY F5); the caller decides whether to apply the mask.
#include "rg_internal.cuh"

namespace {

// one thread per edge of `base`; `other` is the graph searched for the reverse edge
__global__ void edge_keep_kernel(const uint32_t *__restrict__ base, const uint32_t *__restrict__ other, uint64_t n,
                                 uint32_t k, uint32_t keep_n_edges, uint8_t *__restrict__ keep) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * k) return;
    const uint64_t node = e / k;
    const uint32_t j = (uint32_t)(e % k);
    const uint32_t nb = base[e];
    uint8_t kp = 0;
    if (nb != 0xffffffffu) {
        if (j < keep_n_edges) {
            kp = 1;                                            // :19-22 / :50-53
        } else if (nb < n) {
            const uint32_t *row = other + (uint64_t)nb * k;    // :24 / :57
            for (uint32_t q = 0; q < k; ++q) {
                const uint32_t v = row[q];
                if (v == 0xffffffffu) break;
                if (v == (uint32_t)node) { kp = 1; break; }
            }
        }
    }
    keep[e] = kp;
}

int32_t run(rg_handle *h, const uint32_t *base, const uint32_t *other, uint64_t n, uint32_t k, uint32_t keep_n_edges,
            uint8_t *keep) {
    if (!h) return RG_ERR_INVALID;
    if (!base || !other || !keep) return rg_fail(h, RG_ERR_INVALID, "rg_graph_*: NULL pointer");
    if (n == 0 || k == 0) return RG_OK;
    RG_CUDA(h, cudaSetDevice(h->device));
    const size_t bytes = n * k * sizeof(uint32_t);
    DevBuf db, dother, dk;
    int32_t rc = rg_reserve(h, db, bytes);
    if (rc == RG_OK && other != base) rc = rg_reserve(h, dother, bytes);
    if (rc == RG_OK) rc = rg_reserve(h, dk, n * k);
    cudaError_t e = cudaSuccess;
    if (rc == RG_OK) {
        e = cudaMemcpyAsync(db.p, base, bytes, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess && other != base) e = cudaMemcpyAsync(dother.p, other, bytes, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) {
            const uint64_t edges = n * k;
            edge_keep_kernel<<<(unsigned)((edges + 255) / 256), 256, 0, h->stream>>>(
                (const uint32_t *)db.p, (const uint32_t *)(other != base ? dother.p : db.p), n, k, keep_n_edges,
                (uint8_t *)dk.p);
            e = cudaGetLastError();
            h->tm.kernel_launches += 1;
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(keep, dk.p, n * k, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = rg_fail(h, RG_ERR_CUDA, "rg_graph_*: %s", cudaGetErrorString(e));
    }
    rg_release(db); rg_release(dother); rg_release(dk);
    return rc;
}

}  // namespace

extern "C" int32_t rg_graph_mutual(rg_handle *h, const uint32_t *idx, uint64_t n, uint32_t k, uint32_t keep_n_edges,
                                   uint8_t *keep) {
    return run(h, idx, idx, n, k, keep_n_edges, keep);
}

extern "C" int32_t rg_graph_intersect(rg_handle *h, const uint32_t *base_idx, const uint32_t *other_idx, uint64_t n,
                                      uint32_t k, uint32_t keep_n_edges, uint8_t *keep) {
    return run(h, base_idx, other_idx, n, k, keep_n_edges, keep);
}
