// rg_multi.cu — n GPUs behind ONE caller thread (include/rosella_gpu.h, rg_multi_*).
//
// EmbedderEngine::build_mutual_kgraph is called from Rosella's single pipeline thread
// (src/recover/recover_engine.rs:510-519); a drop-in that wants all the GPUs of the node therefore has to fan out inside
// the library.  rg_multi owns one rg_handle per device and an NCCL communicator over them (ncclCommInitAll); every
// call starts one worker thread per device that runs the collective entry points of knn_dist.cu with its handle, and
// the workers write their row slabs straight into the caller's host arrays (disjoint rows, no gather).
#include <algorithm>
#include <thread>
#include <vector>

#include "rg_comm.h"
#include "rg_internal.cuh"

int32_t rg_filter_cached_lists(rg_handle *h, uint32_t keep_n_edges, uint64_t *offsets, uint32_t *neighbours,
                               float *weights, uint64_t *n_disconnected, uint32_t *disconnected);
int32_t rg_reserve_list_cache(rg_handle *h, uint64_t n, uint32_t k);

struct rg_multi {
    std::vector<rg_handle *> hs;
    std::string err;
    std::vector<uint64_t> skt_base;      // rg_multi_sketch_contigs: first hash of device i in the concatenated sketches
};

namespace {

thread_local std::string g_multi_create_error;

template <class F>
int32_t run_all(rg_multi *m, F fn) {
    const int n = (int)m->hs.size();
    std::vector<int32_t> rc(n, RG_OK);
    if (n == 1) {
        try {
            rc[0] = fn(0, m->hs[0]);
        } catch (...) {
            rc[0] = rg_fail(m->hs[0], RG_ERR_OOM, "host allocation failed");
        }
    } else {
        std::vector<std::thread> th;
        th.reserve(n);
        for (int i = 0; i < n; ++i)
            th.emplace_back([&, i] {
                try {
                    rc[i] = fn(i, m->hs[i]);
                } catch (...) {            // no exception leaves the library (host allocation in a worker)
                    rc[i] = rg_fail(m->hs[i], RG_ERR_OOM, "worker thread: host allocation failed");
                }
            });
        for (auto &t : th) t.join();
    }
    for (int i = 0; i < n; ++i)
        if (rc[i] != RG_OK) {
            // prefer the rank that failed by itself over the ranks that only report "another rank failed"
            int pick = i;
            for (int j = 0; j < n; ++j)
                if (rc[j] != RG_OK && m->hs[j]->err.find("another rank") == std::string::npos) { pick = j; break; }
            m->err = "device " + std::to_string(m->hs[pick]->device) + ": " + m->hs[pick]->err;
            return rc[pick];
        }
    return RG_OK;
}

}  // namespace

extern "C" {

int32_t rg_multi_create(const int32_t *device_ids, int32_t n_dev, rg_multi **out) {
    if (!out) return RG_ERR_INVALID;
    *out = nullptr;
    if (!device_ids || n_dev < 1) {
        g_multi_create_error = "rg_multi_create: need at least one device id";
        return RG_ERR_INVALID;
    }
    rg_multi *m = new (std::nothrow) rg_multi();
    if (!m) return RG_ERR_OOM;
    auto fail = [&](int32_t rc, const std::string &msg) {
        g_multi_create_error = msg;
        rg_multi_destroy(m);
        return rc;
    };
    for (int i = 0; i < n_dev; ++i) {
        for (int j = 0; j < i; ++j)
            if (device_ids[j] == device_ids[i]) return fail(RG_ERR_INVALID, "rg_multi_create: device " + std::to_string(device_ids[i]) + " is listed twice");
        rg_handle *h = nullptr;
        const int32_t rc = rg_create(device_ids[i], &h);
        if (rc != RG_OK) return fail(rc, rg_last_error(nullptr));
        m->hs.push_back(h);
    }
    if (n_dev > 1) {
        std::string err;
        const NcclApi *api = rg_nccl(&err);
        if (!api) return fail(RG_ERR_STATE, "rg_multi_create: " + err);
        std::vector<ncclComm_t> comms(n_dev);
        std::vector<int> devs(device_ids, device_ids + n_dev);
        const ncclResult_t r = api->CommInitAll(comms.data(), n_dev, devs.data());
        if (r != ncclSuccess) return fail(RG_ERR_CUDA, std::string("ncclCommInitAll failed: ") + api->GetErrorString(r));
        for (int i = 0; i < n_dev; ++i) {
            m->hs[i]->comm = comms[i];
            m->hs[i]->rank = i;
            m->hs[i]->world = n_dev;
        }
    }
    *out = m;
    return RG_OK;
}

void rg_multi_destroy(rg_multi *m) {
    if (!m) return;
    for (rg_handle *h : m->hs) rg_destroy(h);
    delete m;
}

const char *rg_multi_last_error(const rg_multi *m) { return m ? m->err.c_str() : g_multi_create_error.c_str(); }

int32_t rg_multi_size(const rg_multi *m) { return m ? (int32_t)m->hs.size() : 0; }

rg_handle *rg_multi_handle(rg_multi *m, int32_t i) {
    return (m && i >= 0 && i < (int32_t)m->hs.size()) ? m->hs[i] : nullptr;
}

int32_t rg_multi_knn_build(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                           uint32_t k, uint32_t variant, uint32_t which_mask, uint32_t *tnf_idx, float *tnf_dist,
                           uint32_t *depth_idx, float *depth_dist) {
    if (!m) return RG_ERR_INVALID;
    const bool want_tnf = which_mask & (1u << RG_GRAPH_TNF), want_depth = which_mask & (1u << RG_GRAPH_DEPTH);
    auto bad = [&](const char *msg) { m->err = msg; return (int32_t)RG_ERR_INVALID; };
    if (!want_tnf && !want_depth) return bad("rg_multi_knn_build: which_mask selects no graph");
    if (want_tnf && (!tnf || !tnf_idx || !tnf_dist)) return bad("rg_multi_knn_build: TNF graph requested with a NULL pointer");
    if (want_depth && (!cov || !depth_idx || !depth_dist)) return bad("rg_multi_knn_build: depth graph requested with a NULL pointer");
    return run_all(m, [&](int, rg_handle *h) -> int32_t {
        RG_TRY(rg_knn_load_dist(h, want_tnf ? tnf : nullptr, want_depth ? cov : nullptr, n, d, s, variant, -1));
        if (want_depth) RG_TRY(rg_knn_build_dist(h, RG_GRAPH_DEPTH, k, 0, depth_idx, depth_dist));
        if (want_tnf) RG_TRY(rg_knn_build_dist(h, RG_GRAPH_TNF, k, 0, tnf_idx, tnf_dist));
        return RG_OK;
    });
}

// rg_multi_knn_build with the MinHash sketches of the rows: the depth graph is DepthDistance in full (embedder.rs:133-171).
// Every device joins the sketches on its own copy (rg_knn_load_sketches is local work).
int32_t rg_multi_knn_build_sketched(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                                    uint32_t k, uint32_t variant, uint32_t which_mask, const uint64_t *hashes,
                                    const uint64_t *offsets, double scale, uint32_t *tnf_idx, float *tnf_dist,
                                    uint32_t *depth_idx, float *depth_dist) {
    if (!m) return RG_ERR_INVALID;
    const bool want_tnf = which_mask & (1u << RG_GRAPH_TNF), want_depth = which_mask & (1u << RG_GRAPH_DEPTH);
    auto bad = [&](const char *msg) { m->err = msg; return (int32_t)RG_ERR_INVALID; };
    if (!want_depth || !cov || !depth_idx || !depth_dist || !offsets) return bad("rg_multi_knn_build_sketched: the depth graph, its table and the sketch offsets are required");
    if (want_tnf && (!tnf || !tnf_idx || !tnf_dist)) return bad("rg_multi_knn_build_sketched: TNF graph requested with a NULL pointer");
    return run_all(m, [&](int, rg_handle *h) -> int32_t {
        RG_TRY(rg_knn_load_dist(h, want_tnf ? tnf : nullptr, cov, n, d, s, variant, -1));
        RG_TRY(rg_comm_agree(h, rg_knn_load_sketches(h, hashes, offsets, n, scale)));
        RG_TRY(rg_knn_build_dist(h, RG_GRAPH_DEPTH, k, 0, depth_idx, depth_dist));
        if (want_tnf) RG_TRY(rg_knn_build_dist(h, RG_GRAPH_TNF, k, 0, tnf_idx, tnf_dist));
        return RG_OK;
    });
}

int32_t rg_multi_knn_build_mutual(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                  uint32_t s, uint32_t k, uint32_t variant, uint32_t keep_n_edges, uint64_t *offsets,
                                  uint32_t *neighbours, float *weights, uint64_t *n_disconnected, uint32_t *disconnected) {
    if (!m) return RG_ERR_INVALID;
    if (!tnf || !cov || !offsets || !neighbours || !weights) {
        m->err = "rg_multi_knn_build_mutual: NULL pointer";
        return RG_ERR_INVALID;
    }
    return run_all(m, [&](int i, rg_handle *h) -> int32_t {
        RG_TRY(rg_knn_load_dist(h, tnf, cov, n, d, s, variant, -1));
        RG_TRY(rg_comm_agree(h, rg_reserve_list_cache(h, n, k)));
        // both graphs complete on every GPU (in-place all-gather); the filters then run on device 0
        RG_TRY(rg_knn_build_dist_dev(h, RG_GRAPH_DEPTH, k, 1, (uint32_t *)h->cache_di.p, (float *)h->cache_dd.p));
        RG_TRY(rg_knn_build_dist_dev(h, RG_GRAPH_TNF, k, 1, (uint32_t *)h->cache_ti.p, (float *)h->cache_td.p));
        RG_TRY(rg_sync(h));
        h->lists_cached = 3;
        if (i != 0) return RG_OK;
        return rg_filter_cached_lists(h, keep_n_edges, offsets, neighbours, weights, n_disconnected, disconnected);
    });
}

// Contiguous ranges of records with equal bases per device: no reordering, no exchange, every device uploads its own
// range over its own PCIe link and writes its rows of the outputs.  (A range boundary is off by at most one record
// from the ideal split; the longest record of a metagenome assembly is far below total / n_dev.)
int32_t rg_multi_tnf_count_ascii(rg_multi *m, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t *counts,
                                 uint64_t *n_kmers, double *freqs) {
    if (!m) return RG_ERR_INVALID;
    if (!offsets || !counts || !n_kmers) {
        m->err = "rg_multi_tnf_count_ascii: NULL pointer";
        return RG_ERR_INVALID;
    }
    const int nd = (int)m->hs.size();
    std::vector<uint64_t> cut(nd + 1, 0);
    const uint64_t base0 = n ? offsets[0] : 0, total = n ? offsets[n] - base0 : 0;
    for (int i = 1; i < nd; ++i) {
        const uint64_t target = base0 + total / nd * i;
        cut[i] = std::max<uint64_t>(cut[i - 1], (uint64_t)(std::lower_bound(offsets, offsets + n + 1, target) - offsets));
        if (cut[i] > n) cut[i] = n;
    }
    cut[nd] = n;
    return run_all(m, [&](int i, rg_handle *h) -> int32_t {
        const uint64_t a = cut[i], b = cut[i + 1];
        if (b == a) return RG_OK;
        const uint64_t base = offsets[a] & ~(uint64_t)31;          // keep the 32-byte alignment of the segments
        std::vector<uint64_t> local(b - a + 1);
        for (uint64_t r = a; r <= b; ++r) local[r - a] = offsets[r] - base;
        return rg_tnf_count_ascii(h, seq + base, local.data(), b - a, counts + a * RG_NBINS, n_kmers + a,
                                  freqs ? freqs + a * RG_NBINS : nullptr);
    });
}

// sketch_contigs on all devices: contiguous ranges of records with equal bytes per device (the split of
// rg_multi_tnf_count_ascii), every device sketches its range from its own upload; no exchange.  The hashes stay on the
// devices until rg_multi_sketch_fetch copies every device's slab to its place in the caller's array.
int32_t rg_multi_sketch_contigs(rg_multi *m, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t kmer_length,
                                double scale, uint64_t seed, uint64_t *sketch_offsets, uint64_t *n_valid_kmers, uint64_t *n_hashes) {
    if (!m) return RG_ERR_INVALID;
    if (!offsets || !sketch_offsets) {
        m->err = "rg_multi_sketch_contigs: NULL pointer";
        return RG_ERR_INVALID;
    }
    const int nd = (int)m->hs.size();
    std::vector<uint64_t> cut(nd + 1, 0), totals(nd, 0);
    const uint64_t base0 = n ? offsets[0] : 0, total = n ? offsets[n] - base0 : 0;
    for (int i = 1; i < nd; ++i) {
        const uint64_t target = base0 + total / nd * i;
        cut[i] = std::max<uint64_t>(cut[i - 1], (uint64_t)(std::lower_bound(offsets, offsets + n + 1, target) - offsets));
        if (cut[i] > n) cut[i] = n;
    }
    cut[nd] = n;
    m->skt_base.assign(nd + 1, 0);
    sketch_offsets[0] = 0;
    const int32_t rc = run_all(m, [&](int i, rg_handle *h) -> int32_t {
        const uint64_t a = cut[i], b = cut[i + 1];
        std::vector<uint64_t> local(b - a + 1);
        for (uint64_t r = a; r <= b; ++r) local[r - a] = offsets[r] - offsets[a];
        std::vector<uint64_t> sk(b - a + 1, 0);
        RG_TRY(rg_sketch_contigs(h, seq ? seq + offsets[a] : nullptr, local.data(), b - a, kmer_length, scale, seed, sk.data(),
                                 n_valid_kmers ? n_valid_kmers + a : nullptr, &totals[i]));
        for (uint64_t r = a + 1; r <= b; ++r) sketch_offsets[r] = sk[r - a];      // relative to the device's slab for now
        return RG_OK;
    });
    if (rc != RG_OK) return rc;
    for (int i = 0; i < nd; ++i) {
        m->skt_base[i + 1] = m->skt_base[i] + totals[i];
        for (uint64_t r = cut[i] + 1; r <= cut[i + 1]; ++r) sketch_offsets[r] += m->skt_base[i];
    }
    if (n_hashes) *n_hashes = m->skt_base[nd];
    return RG_OK;
}

int32_t rg_multi_sketch_fetch(rg_multi *m, uint64_t *hashes, uint64_t capacity) {
    if (!m) return RG_ERR_INVALID;
    const int nd = (int)m->hs.size();
    if ((int)m->skt_base.size() != nd + 1) {
        m->err = "rg_multi_sketch_fetch: call rg_multi_sketch_contigs first";
        return RG_ERR_STATE;
    }
    if (m->skt_base[nd] && (!hashes || capacity < m->skt_base[nd])) {
        m->err = "rg_multi_sketch_fetch: room for " + std::to_string(capacity) + " hashes, the sketches hold " + std::to_string(m->skt_base[nd]);
        return RG_ERR_INVALID;
    }
    return run_all(m, [&](int i, rg_handle *h) -> int32_t {
        const uint64_t cnt = m->skt_base[i + 1] - m->skt_base[i];
        return cnt ? rg_sketch_fetch(h, hashes + m->skt_base[i], cnt) : (int32_t)RG_OK;
    });
}

}  // extern "C"
