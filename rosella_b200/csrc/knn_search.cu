// knn_search.cu — host-buffer flavours of the kNN entry points (the calls the Rust shim binds):
// rg_knn_search and the one-shot rg_knn_build that stands in for
// EmbedderEngine::build_mutual_kgraph (src/embedding/embedder.rs:31-86).
#include "rg_internal.cuh"

extern "C" int32_t rg_knn_search(rg_handle *h, uint32_t which, uint32_t k, uint64_t q_begin, uint64_t q_end,
                                 uint32_t *idx, float *dist, uint32_t *n_valid) {
    if (!h) return RG_ERR_INVALID;
    if (!idx || !dist) return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: output pointer is NULL");
    if (q_begin > q_end) return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: q_begin > q_end");
    const uint64_t q_count = q_end - q_begin;
    if (q_count == 0) return RG_OK;
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_TRY(rg_reserve(h, h->out_i, q_count * k * sizeof(uint32_t)));
    RG_TRY(rg_reserve(h, h->out_d, q_count * k * sizeof(float)));
    RG_TRY(rg_reserve(h, h->out_nv, q_count * sizeof(uint32_t)));
    RG_TRY(rg_knn_search_dev(h, which, k, q_begin, q_end, (uint32_t *)h->out_i.p, (float *)h->out_d.p,
                             (uint32_t *)h->out_nv.p));
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    RG_CUDA(h, cudaMemcpyAsync(idx, h->out_i.p, q_count * k * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(dist, h->out_d.p, q_count * k * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (n_valid)
        RG_CUDA(h, cudaMemcpyAsync(n_valid, h->out_nv.p, q_count * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    h->tm.d2h_ms = rg_elapsed(h->ev[0], h->ev[1]);
    h->tm.d2h_bytes += q_count * k * 8 + (n_valid ? q_count * 4 : 0);
    return RG_OK;
}

extern "C" int32_t rg_knn_build(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                uint32_t s, uint32_t k, uint32_t variant, uint32_t which_mask, uint32_t *tnf_idx,
                                float *tnf_dist, uint32_t *depth_idx, float *depth_dist) {
    if (!h) return RG_ERR_INVALID;
    const bool want_tnf = which_mask & (1u << RG_GRAPH_TNF), want_depth = which_mask & (1u << RG_GRAPH_DEPTH);
    if (!want_tnf && !want_depth) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build: which_mask selects no graph");
    if (want_tnf && (!tnf || !tnf_idx || !tnf_dist)) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build: TNF graph requested with a NULL pointer");
    if (want_depth && (!cov || !depth_idx || !depth_dist)) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build: depth graph requested with a NULL pointer");
    RG_TRY(rg_knn_load(h, want_tnf ? tnf : nullptr, want_depth ? cov : nullptr, n, d, s, variant));
    if (want_depth) RG_TRY(rg_knn_search(h, RG_GRAPH_DEPTH, k, 0, n, depth_idx, depth_dist, nullptr));
    if (want_tnf) RG_TRY(rg_knn_search(h, RG_GRAPH_TNF, k, 0, n, tnf_idx, tnf_dist, nullptr));
    return RG_OK;
}
