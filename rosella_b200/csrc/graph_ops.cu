// graph_ops.cu — what mutual_knn / intersect COMPUTE (src/graphs/nearest_neighbour_graph.rs:12-35,
// :38-72): per edge, keep it if it is among the first keep_n_edges of its node, else only if the
// reverse edge exists in the same / the other graph.  The reference discards these lists
// (SURVEY F5); the caller decides whether to apply the mask.
#include <cstdlib>

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

// intersect + mutual_knn + CSR of the lists kept on the handle (cache_di / cache_dd: coverage graph, cache_ti: TNF
// graph, [n][cache_k]); host outputs.  Shared by rg_knn_build_mutual, rg_knn_refilter and rg_multi_knn_build_mutual.
int32_t rg_filter_cached_lists(rg_handle *h, uint32_t keep_n_edges, uint64_t *offsets, uint32_t *neighbours,
                               float *weights, uint64_t *n_disconnected, uint32_t *disconnected) {
    if (h->lists_cached != 3) return rg_fail(h, RG_ERR_STATE, "no kNN lists are kept on the handle (call rg_knn_build_mutual first)");
    if (!offsets || !neighbours || !weights) return rg_fail(h, RG_ERR_INVALID, "graph filter: NULL output pointer");
    RG_CUDA(h, cudaSetDevice(h->device));
    const uint64_t n = h->n;
    const uint32_t k = h->cache_k;
    const uint64_t edges = n * k;
    const uint32_t *di = (const uint32_t *)h->cache_di.p, *ti = (const uint32_t *)h->cache_ti.p;
    const float *dd = (const float *)h->cache_dd.p;
    DevBuf k1, k2, cnt, off, tmp, on, ow;
    auto done = [&](int32_t rc) {
        for (DevBuf *b : {&k1, &k2, &cnt, &off, &tmp, &on, &ow}) rg_release(*b);
        return rc;
    };
#define RG_TRY_D(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define RG_CUDA_D(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, RG_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__))); } while (0)
    RG_TRY_D(rg_reserve(h, k1, edges)); RG_TRY_D(rg_reserve(h, k2, edges));
    RG_TRY_D(rg_reserve(h, cnt, (n + 1) * 8)); RG_TRY_D(rg_reserve(h, off, (n + 1) * 8));
    const uint32_t keep_first = keep_n_edges > 5u ? keep_n_edges : 5u;                                    // DEFAULT_KEEP_N_EDGES
    edge_keep_kernel<<<(unsigned)((edges + 255) / 256), 256, 0, h->stream>>>(di, ti, n, k, keep_first, (uint8_t *)k1.p);
    RG_CUDA_D(cudaGetLastError());
    RG_CUDA_D(cudaMemsetAsync(cnt.p, 0, (n + 1) * 8, h->stream));
    mutual_after_intersect_kernel<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(
        di, (const uint8_t *)k1.p, n, k, keep_n_edges, (uint8_t *)k2.p, (uint64_t *)cnt.p);
    RG_CUDA_D(cudaGetLastError());
    size_t tmp_bytes = 0;
    RG_CUDA_D(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (const uint64_t *)cnt.p, (uint64_t *)off.p, (int64_t)(n + 1), h->stream));
    RG_TRY_D(rg_reserve(h, tmp, tmp_bytes ? tmp_bytes : 8));
    RG_CUDA_D(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, (const uint64_t *)cnt.p, (uint64_t *)off.p, (int64_t)(n + 1), h->stream));
    RG_CUDA_D(cudaMemcpyAsync(offsets, off.p, (n + 1) * 8, cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA_D(cudaStreamSynchronize(h->stream));
    const uint64_t total = offsets[n];
    RG_TRY_D(rg_reserve(h, on, (total ? total : 1) * 4)); RG_TRY_D(rg_reserve(h, ow, (total ? total : 1) * 4));
    compact_kernel<<<(unsigned)((n + 127) / 128), 128, 0, h->stream>>>(di, dd, (const uint8_t *)k2.p, (const uint64_t *)off.p, n, k,
                                                                      (uint32_t *)on.p, (float *)ow.p);
    RG_CUDA_D(cudaGetLastError());
    h->tm.kernel_launches += 3;
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

// reserve the kept lists for n rows x k (both graphs)
int32_t rg_reserve_list_cache(rg_handle *h, uint64_t n, uint32_t k) {
    h->lists_cached = 0;
    RG_TRY(rg_reserve(h, h->cache_di, n * k * 4)); RG_TRY(rg_reserve(h, h->cache_dd, n * k * 4));
    RG_TRY(rg_reserve(h, h->cache_ti, n * k * 4)); RG_TRY(rg_reserve(h, h->cache_td, n * k * 4));
    h->cache_k = k;
    return RG_OK;
}

extern "C" int32_t rg_knn_build_mutual(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                       uint32_t s, uint32_t k, uint32_t variant, uint32_t keep_n_edges,
                                       uint64_t *offsets, uint32_t *neighbours, float *weights,
                                       uint64_t *n_disconnected, uint32_t *disconnected) {
    if (!h) return RG_ERR_INVALID;
    if (!tnf || !cov || !offsets || !neighbours || !weights)
        return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_mutual: NULL pointer");
    RG_TRY(rg_knn_load(h, tnf, cov, n, d, s, variant));
    RG_TRY(rg_reserve_list_cache(h, n, k));
    RG_TRY(rg_knn_search_dev(h, RG_GRAPH_DEPTH, k, 0, n, (uint32_t *)h->cache_di.p, (float *)h->cache_dd.p, nullptr));   // embedder.rs:32-65
    RG_TRY(rg_knn_search_dev(h, RG_GRAPH_TNF, k, 0, n, (uint32_t *)h->cache_ti.p, (float *)h->cache_td.p, nullptr));
    h->lists_cached = 3;
    return rg_filter_cached_lists(h, keep_n_edges, offsets, neighbours, weights, n_disconnected, disconnected);
}

extern "C" int32_t rg_knn_refilter(rg_handle *h, uint32_t keep_n_edges, uint64_t *offsets, uint32_t *neighbours,
                                   float *weights, uint64_t *n_disconnected, uint32_t *disconnected) {
    if (!h) return RG_ERR_INVALID;
    return rg_filter_cached_lists(h, keep_n_edges, offsets, neighbours, weights, n_disconnected, disconnected);
}

// evaluate_subset (recover_engine.rs:410-435): the rows of row_ids are gathered (retrieve_contig_data, :522-543) and
// the graph of those rows alone is built; node numbers are positions in row_ids
extern "C" int32_t rg_knn_build_subset(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                       uint32_t s, const uint64_t *row_ids, uint64_t m, uint32_t k, uint32_t variant,
                                       uint32_t which_mask, uint32_t *tnf_idx, float *tnf_dist, uint32_t *depth_idx,
                                       float *depth_dist) {
    if (!h) return RG_ERR_INVALID;
    if (!row_ids) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_subset: row_ids is NULL");
    const bool want_tnf = which_mask & (1u << RG_GRAPH_TNF), want_depth = which_mask & (1u << RG_GRAPH_DEPTH);
    if ((want_tnf && !tnf) || (want_depth && !cov)) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_subset: a requested graph has a NULL table");
    for (uint64_t i = 0; i < m; ++i)
        if (row_ids[i] >= n) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_subset: row_ids[%llu] = %llu is outside [0,%llu)",
                                            (unsigned long long)i, (unsigned long long)row_ids[i], (unsigned long long)n);
    double *sub_tnf = nullptr, *sub_cov = nullptr;
    const size_t wt = d, wc = 2 * (size_t)s;
    if (want_tnf) sub_tnf = (double *)malloc((m ? m : 1) * wt * sizeof(double));
    if (want_depth) sub_cov = (double *)malloc((m ? m : 1) * wc * sizeof(double));
    if ((want_tnf && !sub_tnf) || (want_depth && !sub_cov)) {
        free(sub_tnf); free(sub_cov);
        return rg_fail(h, RG_ERR_OOM, "rg_knn_build_subset: host allocation for %llu gathered rows failed", (unsigned long long)m);
    }
    for (uint64_t i = 0; i < m; ++i) {
        if (want_tnf) memcpy(sub_tnf + i * wt, tnf + row_ids[i] * wt, wt * sizeof(double));
        if (want_depth) memcpy(sub_cov + i * wc, cov + row_ids[i] * wc, wc * sizeof(double));
    }
    const int32_t rc = rg_knn_build(h, sub_tnf, sub_cov, m, d, s, k, variant, which_mask, tnf_idx, tnf_dist, depth_idx, depth_dist);
    free(sub_tnf); free(sub_cov);
    return rc;
}
