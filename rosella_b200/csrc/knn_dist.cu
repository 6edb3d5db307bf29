// knn_dist.cu — EmbedderEngine::build_mutual_kgraph (src/embedding/embedder.rs:31-86) on the ranks of an NCCL
// communicator, entirely inside the library (the round-1 build needed Python + torch.distributed for this).
//
//   load     only the root rank reads the caller's host tables; the f64 tables reach the other GPUs by ncclBroadcast
//            over NVLink (1.9 GB at 1M x (136 + 50): milliseconds instead of one PCIe upload per rank); every rank
//            then runs the pre-transform and the collision scan on its own copy.
//   coverage graph   the symmetric block build of knn_depth_sym.cu: rank p evaluates its share of the row blocks
//            (every unordered pair on exactly one rank) and holds partial K' lists for ALL rows; one all-to-all
//            (grouped ncclSend / ncclRecv, uneven row shards) hands every row's `world` partial lists to the rank
//            that owns the row; merge by rank counting, f64 re-evaluation, certificate as on one GPU.
//   TNF graphs       query-row shards (the tensor-core select is 2 % of the build; symmetry would not pay).
//   result   written in place into the rank's rows of the [n][k] arrays; gather = 1 adds an in-place all-gather
//            (one grouped ncclBroadcast per owner, slabs are uneven).
// Everything is issued on rg_stream(h); phase times come from CUDA events on that stream (rg_get_dist_timings).
#include "rg_comm.h"
#include "rg_internal.cuh"

int32_t rg_knn_load_begin(rg_handle *h, bool has_tnf, bool has_cov, uint64_t n, uint32_t d, uint32_t s, uint32_t variant);
int32_t rg_knn_load_finish(rg_handle *h, bool has_tnf, bool has_cov, uint64_t n, uint32_t d, uint32_t s, uint32_t variant);
int32_t rg_launch_depth_build_sym(rg_handle *h, uint32_t kprime, float *cd, uint32_t *ci, uint32_t part, uint32_t n_parts);

namespace {

int32_t ensure_events(rg_handle *h) {
    for (auto &ev : h->dist_ev)
        if (!ev) RG_CUDA(h, cudaEventCreate(&ev));
    return RG_OK;
}

// in-place all-gather of uneven row slabs: rows [begin_r, end_r) of `buf` (elem_per_row 4-byte elements per row) are
// valid on rank r; afterwards all rows are valid everywhere
int32_t all_gather_rows(rg_handle *h, const NcclApi *api, void *buf, uint64_t n, uint32_t elem_per_row) {
    RG_NCCL(h, api, GroupStart());
    for (int r = 0; r < h->world; ++r) {
        uint64_t b, e;
        rg_shard(n, r, h->world, &b, &e);
        if (e == b) continue;
        uint32_t *p = (uint32_t *)buf + b * elem_per_row;
        RG_NCCL(h, api, Broadcast(p, p, (e - b) * elem_per_row, ncclUint32, r, (ncclComm_t)h->comm, h->stream));
    }
    RG_NCCL(h, api, GroupEnd());
    return RG_OK;
}

}  // namespace

static int32_t resolve_dist_timings(rg_handle *h) {
    if (h->dtm.partial_ms < 0.0) {
        const bool with_d2h = h->dtm.partial_ms < -1.5;
        RG_CUDA(h, cudaSetDevice(h->device));
        RG_CUDA(h, cudaEventSynchronize(h->dist_ev[with_d2h ? 5 : 4]));
        h->dtm.partial_ms = rg_elapsed(h->dist_ev[0], h->dist_ev[1]);
        h->dtm.exchange_ms = rg_elapsed(h->dist_ev[1], h->dist_ev[2]);
        h->dtm.finish_ms = rg_elapsed(h->dist_ev[2], h->dist_ev[3]);
        h->dtm.gather_ms = rg_elapsed(h->dist_ev[3], h->dist_ev[with_d2h ? 5 : 4]);
        if (h->dtm_which <= RG_GRAPH_TNF_EUCLID) h->dtm_graph[h->dtm_which] = h->dtm;
    }
    return RG_OK;
}

extern "C" {

int32_t rg_knn_load_dist(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                         uint32_t variant, int32_t root) {
    if (!h) return RG_ERR_INVALID;
    if (h->world <= 1 || !h->comm) return rg_knn_load(h, tnf, cov, n, d, s, variant);
    if (root < -1 || root >= h->world) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_dist: root %d of %d", root, h->world);
    const NcclApi *api = rg_nccl(nullptr);
    RG_CUDA(h, cudaSetDevice(h->device));
    int32_t rc = ensure_events(h);
    // which tables exist is the root's call (root == -1: every rank holds the same host tables, rank 0 speaks)
    const int speaker = root < 0 ? 0 : root;
    int32_t have[2] = {tnf != nullptr, cov != nullptr};
    if (rc == RG_OK) rc = rg_reserve(h, h->dist_status, 4 * sizeof(int32_t));
    RG_TRY(rg_comm_agree(h, rc));
    int32_t *dev = (int32_t *)h->dist_status.p + 2;
    RG_CUDA(h, cudaMemcpyAsync(dev, have, sizeof(have), cudaMemcpyHostToDevice, h->stream));
    RG_NCCL(h, api, Broadcast(dev, dev, 2, ncclInt32, speaker, (ncclComm_t)h->comm, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(have, dev, sizeof(have), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    const bool has_tnf = have[0] != 0, has_cov = have[1] != 0;
    rc = rg_knn_load_begin(h, has_tnf, has_cov, n, d, s, variant);
    if (rc == RG_OK && root < 0 && ((has_tnf && !tnf) || (has_cov && !cov)))
        rc = rg_fail(h, RG_ERR_INVALID, "rg_knn_load_dist: root = -1 needs the host tables on every rank");
    RG_TRY(rg_comm_agree(h, rc));
    RG_CUDA(h, cudaEventRecord(h->dist_ev[0], h->stream));
    if (root >= 0) {
        // one PCIe upload by the root, then NVLink
        if (h->rank == root) {
            if (has_tnf) RG_CUDA(h, cudaMemcpyAsync(h->tnf64.p, tnf, n * d * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            if (has_cov) RG_CUDA(h, cudaMemcpyAsync(h->cov64.p, cov, n * 2 * s * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            h->tm.h2d_bytes += (has_tnf ? n * d * 8 : 0) + (has_cov ? n * 2 * s * 8 : 0);
        }
        RG_CUDA(h, cudaEventRecord(h->dist_ev[1], h->stream));
        if (has_tnf) RG_NCCL(h, api, Broadcast(h->tnf64.p, h->tnf64.p, n * d, ncclDouble, root, (ncclComm_t)h->comm, h->stream));
        if (has_cov) RG_NCCL(h, api, Broadcast(h->cov64.p, h->cov64.p, n * 2 * s, ncclDouble, root, (ncclComm_t)h->comm, h->stream));
    } else {
        // every rank uploads ITS row shard over its own PCIe link (1 / world of the bytes each, all links at once), then
        // an in-place all-gather over NVLink completes the tables everywhere
        uint64_t qb, qe;
        rg_shard(n, h->rank, h->world, &qb, &qe);
        if (qe > qb) {
            if (has_tnf) RG_CUDA(h, cudaMemcpyAsync((double *)h->tnf64.p + qb * d, tnf + qb * d, (qe - qb) * d * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            if (has_cov) RG_CUDA(h, cudaMemcpyAsync((double *)h->cov64.p + qb * 2 * s, cov + qb * 2 * s, (qe - qb) * 2 * s * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            h->tm.h2d_bytes += (qe - qb) * ((has_tnf ? d : 0) + (has_cov ? 2 * s : 0)) * 8;
        }
        RG_CUDA(h, cudaEventRecord(h->dist_ev[1], h->stream));
        if (has_tnf) RG_TRY(all_gather_rows(h, api, h->tnf64.p, n, 2 * d));          // rows of d doubles = 2 d 4-byte elements
        if (has_cov) RG_TRY(all_gather_rows(h, api, h->cov64.p, n, 4 * s));
    }
    RG_CUDA(h, cudaEventRecord(h->dist_ev[2], h->stream));
    RG_CUDA(h, cudaEventSynchronize(h->dist_ev[2]));
    h->dtm.load_h2d_ms = rg_elapsed(h->dist_ev[0], h->dist_ev[1]);
    h->dtm.load_bcast_ms = rg_elapsed(h->dist_ev[1], h->dist_ev[2]);
    h->tm.h2d_ms = h->dtm.load_h2d_ms;
    RG_TRY(rg_comm_agree(h, rg_knn_load_finish(h, has_tnf, has_cov, n, d, s, variant)));
    h->dtm.load_transform_ms = h->tm.transform_ms;
    return RG_OK;
}

int32_t rg_knn_build_dist_dev(rg_handle *h, uint32_t which, uint32_t k, uint32_t gather, uint32_t *idx_dev, float *dist_dev) {
    if (!h) return RG_ERR_INVALID;
    if (!idx_dev || !dist_dev) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_dist: output pointer is NULL");
    if (h->dist_ev[0]) resolve_dist_timings(h);      // keep the phases of the previous graph (rg_get_dist_timings_of)
    h->dtm_which = which;
    if (h->world <= 1 || !h->comm) {
        RG_TRY(ensure_events(h));
        RG_CUDA(h, cudaEventRecord(h->dist_ev[0], h->stream));
        RG_TRY(rg_knn_search_dev(h, which, k, 0, h->n, idx_dev, dist_dev, nullptr));
        for (int e = 1; e <= 4; ++e) RG_CUDA(h, cudaEventRecord(h->dist_ev[e], h->stream));
        h->dtm.partial_ms = -1.0;
        return RG_OK;
    }
    const NcclApi *api = rg_nccl(nullptr);
    RG_CUDA(h, cudaSetDevice(h->device));
    // everything that can fail locally happens before the first collective, then the ranks agree
    int32_t rc = ensure_events(h);
    if (rc == RG_OK && h->n == 0) rc = rg_fail(h, RG_ERR_STATE, "rg_knn_build_dist: no tables loaded");
    if (rc == RG_OK && which > RG_GRAPH_TNF_EUCLID) rc = rg_fail(h, RG_ERR_INVALID, "rg_knn_build_dist: unknown graph %u", which);
    if (rc == RG_OK && which != RG_GRAPH_DEPTH && !h->has_tnf) rc = rg_fail(h, RG_ERR_STATE, "rg_knn_build_dist: no TNF table loaded");
    if (rc == RG_OK && which == RG_GRAPH_DEPTH && !h->has_cov) rc = rg_fail(h, RG_ERR_STATE, "rg_knn_build_dist: no coverage table loaded");
    const uint32_t kp = rc == RG_OK ? rg_knn_kprime(h, which, k) : 0;
    if (rc == RG_OK && (k == 0 || kp == 0)) rc = rg_fail(h, RG_ERR_INVALID, "rg_knn_build_dist: k=%u must be in [1,224]", k);
    const uint64_t n = h->n;
    uint64_t qb = 0, qe = 0;
    rg_shard(n, h->rank, h->world, &qb, &qe);
    const uint64_t mine = qe - qb;
    const bool sym = which == RG_GRAPH_DEPTH && h->opt_depth_build != 1;
    if (rc == RG_OK && sym) {
        rc = rg_reserve(h, h->dist_pd, n * kp * sizeof(float));
        if (rc == RG_OK) rc = rg_reserve(h, h->dist_pi, n * kp * sizeof(uint32_t));
        if (rc == RG_OK) rc = rg_reserve(h, h->dist_rd, (uint64_t)h->world * mine * kp * sizeof(float));
        if (rc == RG_OK) rc = rg_reserve(h, h->dist_ri, (uint64_t)h->world * mine * kp * sizeof(uint32_t));
        if (rc == RG_OK) rc = rg_reserve(h, h->cand_d, mine * kp * sizeof(float));
        if (rc == RG_OK) rc = rg_reserve(h, h->cand_i, mine * kp * sizeof(uint32_t));
    }
    RG_TRY(rg_comm_agree(h, rc));
    RG_CUDA(h, cudaEventRecord(h->dist_ev[0], h->stream));
    if (sym) {
        float *pd = (float *)h->dist_pd.p, *rd = (float *)h->dist_rd.p;
        uint32_t *pi = (uint32_t *)h->dist_pi.p, *ri = (uint32_t *)h->dist_ri.p;
        RG_TRY(rg_launch_depth_build_sym(h, kp, pd, pi, (uint32_t)h->rank, (uint32_t)h->world));
        RG_CUDA(h, cudaEventRecord(h->dist_ev[1], h->stream));
        RG_NCCL(h, api, GroupStart());
        for (int r = 0; r < h->world; ++r) {
            uint64_t b, e;
            rg_shard(n, r, h->world, &b, &e);
            if (e > b) {
                RG_NCCL(h, api, Send(pd + b * kp, (e - b) * kp, ncclFloat, r, (ncclComm_t)h->comm, h->stream));
                RG_NCCL(h, api, Send(pi + b * kp, (e - b) * kp, ncclUint32, r, (ncclComm_t)h->comm, h->stream));
            }
            if (mine) {
                RG_NCCL(h, api, Recv(rd + (uint64_t)r * mine * kp, mine * kp, ncclFloat, r, (ncclComm_t)h->comm, h->stream));
                RG_NCCL(h, api, Recv(ri + (uint64_t)r * mine * kp, mine * kp, ncclUint32, r, (ncclComm_t)h->comm, h->stream));
            }
        }
        RG_NCCL(h, api, GroupEnd());
        RG_CUDA(h, cudaEventRecord(h->dist_ev[2], h->stream));
        if (mine)
            RG_TRY(rg_knn_depth_merge_dev(h, k, (uint32_t)h->world, rd, ri, qb, qe, idx_dev + qb * k, dist_dev + qb * k, nullptr));
    } else {
        RG_CUDA(h, cudaEventRecord(h->dist_ev[1], h->stream));
        RG_CUDA(h, cudaEventRecord(h->dist_ev[2], h->stream));
        if (mine) RG_TRY(rg_knn_search_dev(h, which, k, qb, qe, idx_dev + qb * k, dist_dev + qb * k, nullptr));
    }
    RG_CUDA(h, cudaEventRecord(h->dist_ev[3], h->stream));
    if (gather) {
        RG_TRY(all_gather_rows(h, api, idx_dev, n, k));
        RG_TRY(all_gather_rows(h, api, dist_dev, n, k));
    }
    RG_CUDA(h, cudaEventRecord(h->dist_ev[4], h->stream));
    h->dtm.partial_ms = -1.0;     // resolved by rg_get_dist_timings
    return RG_OK;
}

int32_t rg_knn_build_dist(rg_handle *h, uint32_t which, uint32_t k, uint32_t gather, uint32_t *idx, float *dist) {
    if (!h) return RG_ERR_INVALID;
    if (!idx || !dist) return rg_fail(h, RG_ERR_INVALID, "rg_knn_build_dist: output pointer is NULL");
    RG_CUDA(h, cudaSetDevice(h->device));
    const uint64_t n = h->n;
    int32_t rc = rg_reserve(h, h->out_i, (n ? n : 1) * k * sizeof(uint32_t));
    if (rc == RG_OK) rc = rg_reserve(h, h->out_d, (n ? n : 1) * k * sizeof(float));
    RG_TRY(rg_comm_agree(h, rc));
    RG_TRY(rg_knn_build_dist_dev(h, which, k, gather, (uint32_t *)h->out_i.p, (float *)h->out_d.p));
    uint64_t qb = 0, qe = n;
    if (!gather) rg_shard(n, h->rank, h->world, &qb, &qe);
    if (qe > qb) {
        RG_CUDA(h, cudaMemcpyAsync(idx + qb * k, (uint32_t *)h->out_i.p + qb * k, (qe - qb) * k * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
        RG_CUDA(h, cudaMemcpyAsync(dist + qb * k, (float *)h->out_d.p + qb * k, (qe - qb) * k * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    }
    RG_CUDA(h, cudaEventRecord(h->dist_ev[5], h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    h->tm.d2h_bytes += (qe - qb) * k * 8;
    h->dtm.partial_ms = -2.0;     // as -1, and the D2H copy belongs to gather_ms
    return RG_OK;
}

int32_t rg_get_dist_timings(rg_handle *h, rg_dist_timings *out) {
    if (!h || !out) return RG_ERR_INVALID;
    RG_TRY(resolve_dist_timings(h));
    *out = h->dtm;
    return RG_OK;
}

int32_t rg_get_dist_timings_of(rg_handle *h, uint32_t which, rg_dist_timings *out) {
    if (!h || !out || which > RG_GRAPH_TNF_EUCLID) return RG_ERR_INVALID;
    RG_TRY(resolve_dist_timings(h));
    *out = h->dtm_graph[which];
    return RG_OK;
}

}  // extern "C"
