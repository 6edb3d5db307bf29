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

// ---------------------------------------------------------------------------------------------------------------
// rg_knn_build_mutual — EmbedderEngine::build_mutual_kgraph (src/embedding/embedder.rs:31-86) in one call, on the
// device, with the two filters APPLIED (the reference computes them and drops the result,
// nearest_neighbour_graph.rs:33,:70 are commented out):
//   kgraph = intersect(depth_kgraph, tnf_kgraph, max(keep_n_edges, 5))     embedder.rs:68
//   kgraph = mutual_knn(kgraph, keep_n_edges)                              embedder.rs:69
//   disconnected_nodes = nodes left without neighbours                     embedder.rs:72-84
// Result as CSR in host memory: offsets u64[n+1], neighbours / weights in list order.
#include <cub/device/device_scan.cuh>

namespace {

// thread per node: mutual_knn on the lists that survived intersect.  A surviving edge's position in its list is the
// number of surviving edges before it (:18 enumerates the filtered list); the reverse edge must have survived too
// (:24 searches the same filtered graph).
__global__ void mutual_after_intersect_kernel(const uint32_t *__restrict__ base, const uint8_t *__restrict__ keep1,
                                              uint64_t n, uint32_t k, uint32_t keep_n_edges,
                                              uint8_t *__restrict__ keep2, uint64_t *__restrict__ count) {
    const uint64_t node = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n) return;
    uint32_t pos = 0, kept = 0;
    for (uint32_t j = 0; j < k; ++j) {
        const uint64_t e = node * k + j;
        uint8_t kp = 0;
        if (keep1[e]) {
            if (pos < keep_n_edges) {
                kp = 1;
            } else {
                const uint32_t nb = base[e];
                const uint32_t *row = base + (uint64_t)nb * k;
                const uint8_t *rk = keep1 + (uint64_t)nb * k;
                for (uint32_t q = 0; q < k; ++q) {
                    if (row[q] == 0xffffffffu) break;
                    if (row[q] == (uint32_t)node && rk[q]) { kp = 1; break; }
                }
            }
            ++pos;
        }
        keep2[e] = kp;
        kept += kp;
    }
    count[node] = kept;
}

__global__ void compact_kernel(const uint32_t *__restrict__ idx, const float *__restrict__ dist,
                               const uint8_t *__restrict__ keep, const uint64_t *__restrict__ offsets, uint64_t n,
                               uint32_t k, uint32_t *__restrict__ nbr, float *__restrict__ weight) {
    const uint64_t node = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n) return;
    uint64_t o = offsets[node];
    for (uint32_t j = 0; j < k; ++j) {
        const uint64_t e = node * k + j;
        if (keep[e]) {
            nbr[o] = idx[e];
            weight[o] = dist[e];
            ++o;
        }
    }
}

}  // namespace

extern "C" int32_t rg_knn_build_mutual(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                       uint32_t s, uint32_t k, uint32_t variant, uint32_t keep_n_edges,
                                       uint64_t *offsets, uint32_t *neighbours, float *weights,
                                       uint64_t *n_disconnected, uint32_t *disconnected) {
    if (!h) return RG_ERR_INVALID;
    if (!tnf || !cov || !offsets || !neighbours || !weights)
        return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_mutual: NULL pointer");
    RG_TRY(rg_knn_load(h, tnf, cov, n, d, s, variant));
    const uint64_t edges = n * k;
    DevBuf di, dd, ti, td, k1, k2, cnt, off, tmp, on, ow;
    auto done = [&](int32_t rc) {
        for (DevBuf *b : {&di, &dd, &ti, &td, &k1, &k2, &cnt, &off, &tmp, &on, &ow}) rg_release(*b);
        return rc;
    };
#define RG_TRY_D(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define RG_CUDA_D(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, RG_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__))); } while (0)
    RG_TRY_D(rg_reserve(h, di, edges * 4)); RG_TRY_D(rg_reserve(h, dd, edges * 4));
    RG_TRY_D(rg_reserve(h, ti, edges * 4)); RG_TRY_D(rg_reserve(h, td, edges * 4));
    RG_TRY_D(rg_reserve(h, k1, edges)); RG_TRY_D(rg_reserve(h, k2, edges));
    RG_TRY_D(rg_reserve(h, cnt, (n + 1) * 8)); RG_TRY_D(rg_reserve(h, off, (n + 1) * 8));
    RG_TRY_D(rg_knn_search_dev(h, RG_GRAPH_DEPTH, k, 0, n, (uint32_t *)di.p, (float *)dd.p, nullptr));   // embedder.rs:32-65
    RG_TRY_D(rg_knn_search_dev(h, RG_GRAPH_TNF, k, 0, n, (uint32_t *)ti.p, (float *)td.p, nullptr));
    const uint32_t keep_first = keep_n_edges > 5u ? keep_n_edges : 5u;                                    // DEFAULT_KEEP_N_EDGES
    edge_keep_kernel<<<(unsigned)((edges + 255) / 256), 256, 0, h->stream>>>((const uint32_t *)di.p, (const uint32_t *)ti.p,
                                                                            n, k, keep_first, (uint8_t *)k1.p);
    RG_CUDA_D(cudaMemsetAsync(cnt.p, 0, (n + 1) * 8, h->stream));
    mutual_after_intersect_kernel<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(
        (const uint32_t *)di.p, (const uint8_t *)k1.p, n, k, keep_n_edges, (uint8_t *)k2.p, (uint64_t *)cnt.p);
    size_t tmp_bytes = 0;
    RG_CUDA_D(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (const uint64_t *)cnt.p, (uint64_t *)off.p, (int64_t)(n + 1), h->stream));
    RG_TRY_D(rg_reserve(h, tmp, tmp_bytes ? tmp_bytes : 8));
    RG_CUDA_D(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, (const uint64_t *)cnt.p, (uint64_t *)off.p, (int64_t)(n + 1), h->stream));
    RG_CUDA_D(cudaMemcpyAsync(offsets, off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA_D(cudaStreamSynchronize(h->stream));
    const uint64_t total = offsets[n];
    RG_TRY_D(rg_reserve(h, on, (total ? total : 1) * 4)); RG_TRY_D(rg_reserve(h, ow, (total ? total : 1) * 4));
    compact_kernel<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>((const uint32_t *)di.p, (const float *)dd.p,
                                                                      (const uint8_t *)k2.p, (const uint64_t *)off.p, n, k,
                                                                      (uint32_t *)on.p, (float *)ow.p);
    RG_CUDA_D(cudaGetLastError());
    h->tm.kernel_launches += 4;
    if (total) {
        RG_CUDA_D(cudaMemcpyAsync(neighbours, on.p, total * 4, cudaMemcpyDeviceToHost, h->stream));
        RG_CUDA_D(cudaMemcpyAsync(weights, ow.p, total * 4, cudaMemcpyDeviceToHost, h->stream));
    }
    RG_CUDA_D(cudaStreamSynchronize(h->stream));
    h->tm.d2h_bytes += (n + 1) * 8 + total * 8;
    uint64_t nd = 0;
    for (uint64_t i = 0; i < n; ++i)
        if (offsets[i + 1] == offsets[i]) {
            if (disconnected) disconnected[nd] = (uint32_t)i;
            ++nd;
        }
    if (n_disconnected) *n_disconnected = nd;
#undef RG_TRY_D
#undef RG_CUDA_D
    return done(RG_OK);
}
