// embedder.cpp — EmbedderEngine::build_mutual_kgraph, KGraph, mutual_knn, intersect on top of the C ABI.
//   src/embedding/embedder.rs:15-87, src/graphs/nearest_neighbour_graph.rs:12-72,
//   annembed KGraph (third party; only what the path touches: neighbours, node <-> data-id map, max_nbng).
#include <algorithm>

#include "rosella_host.hpp"

namespace rosella {

static const size_t DEFAULT_KEEP_N_EDGES = 5;        // embedder.rs:13

KGraph::KGraph(std::vector<std::vector<OutEdge>> neighbours, std::vector<size_t> data_ids, size_t max_nbng)
    : neighbours_(std::move(neighbours)), data_ids_(std::move(data_ids)), max_nbng_(max_nbng) {
    if (neighbours_.size() != data_ids_.size()) throw Error("KGraph: one data id per node is required");
}

std::optional<size_t> KGraph::get_data_id_from_idx(size_t idx) const {
    if (idx >= data_ids_.size()) return std::nullopt;
    return data_ids_[idx];
}

std::optional<size_t> KGraph::get_idx_from_dataid(size_t data_id) const {
    // the graphs built here keep rows in insertion order (data id == node index); fall back to a scan otherwise
    if (data_id < data_ids_.size() && data_ids_[data_id] == data_id) return data_id;
    auto it = std::find(data_ids_.begin(), data_ids_.end(), data_id);
    if (it == data_ids_.end()) return std::nullopt;
    return (size_t)(it - data_ids_.begin());
}

// rows of k (idx, f32 distance) entries padded with UINT32_MAX -> KGraph with f64 weights (KGraph<f64>, embedder.rs:64)
static KGraph kgraph_from_lists(const std::vector<uint32_t> &idx, const std::vector<float> &dist, size_t n, size_t k) {
    std::vector<std::vector<OutEdge>> nb(n);
    std::vector<size_t> ids(n);
    for (size_t i = 0; i < n; ++i) {
        ids[i] = i;
        nb[i].reserve(k);
        for (size_t j = 0; j < k; ++j) {
            const uint32_t node = idx[i * k + j];
            if (node == UINT32_MAX) break;
            nb[i].push_back(OutEdge{(size_t)node, (double)dist[i * k + j]});
        }
    }
    return KGraph(std::move(nb), std::move(ids), k);
}

// neighbour lists -> n x k index matrix padded with UINT32_MAX (the layout rg_graph_* take)
static std::vector<uint32_t> lists_to_matrix(const KGraph &g, size_t &k) {
    k = 1;
    for (const auto &l : g.get_neighbours()) k = std::max(k, l.size());
    std::vector<uint32_t> m(g.get_nb_nodes() * k, UINT32_MAX);
    for (size_t i = 0; i < g.get_nb_nodes(); ++i)
        for (size_t j = 0; j < g.get_neighbours()[i].size(); ++j) m[i * k + j] = (uint32_t)g.get_neighbours()[i][j].node;
    return m;
}

static void apply_keep(KGraph &g, const std::vector<uint8_t> &keep, size_t k) {
    auto &nb = g.neighbours_mut();
    for (size_t i = 0; i < nb.size(); ++i) {
        std::vector<OutEdge> kept;
        for (size_t j = 0; j < nb[i].size(); ++j)
            if (keep[i * k + j]) kept.push_back(nb[i][j]);
        nb[i].swap(kept);
    }
}

KGraph mutual_knn(Gpu &gpu, KGraph knn_graph, size_t keep_n_edges, bool apply) {
    size_t k = 0;
    const std::vector<uint32_t> m = lists_to_matrix(knn_graph, k);
    std::vector<uint8_t> keep(m.size());
    gpu.check(rg_graph_mutual(gpu.raw(), m.data(), knn_graph.get_nb_nodes(), (uint32_t)k,
                              (uint32_t)std::min<size_t>(keep_n_edges, UINT32_MAX), keep.data()),
              "rg_graph_mutual");                                        // nearest_neighbour_graph.rs:13-30
    if (apply) apply_keep(knn_graph, keep, k);                           // :33 is commented out in the reference
    return knn_graph;
}

KGraph intersect(Gpu &gpu, KGraph base_knn_graph, const KGraph &intersecting_graph, size_t keep_n_edges, bool apply) {
    if (base_knn_graph.get_nb_nodes() != intersecting_graph.get_nb_nodes())
        throw Error("intersect: the two graphs must be built from the same contig information");
    size_t kb = 0, ko = 0;
    const std::vector<uint32_t> mb = lists_to_matrix(base_knn_graph, kb);
    std::vector<uint32_t> mo = lists_to_matrix(intersecting_graph, ko);
    const size_t n = base_knn_graph.get_nb_nodes(), k = std::max(kb, ko);
    auto widen = [&](const std::vector<uint32_t> &m, size_t from) {
        if (from == k) return m;
        std::vector<uint32_t> w(n * k, UINT32_MAX);
        for (size_t i = 0; i < n; ++i) std::copy(m.begin() + i * from, m.begin() + (i + 1) * from, w.begin() + i * k);
        return w;
    };
    const std::vector<uint32_t> b = widen(mb, kb), o = widen(mo, ko);
    std::vector<uint8_t> keep(n * k);
    gpu.check(rg_graph_intersect(gpu.raw(), b.data(), o.data(), n, (uint32_t)k,
                                 (uint32_t)std::min<size_t>(keep_n_edges, UINT32_MAX), keep.data()),
              "rg_graph_intersect");                                     // nearest_neighbour_graph.rs:42-67
    if (apply) apply_keep(base_knn_graph, keep, k);                      // :70 is commented out in the reference
    return base_knn_graph;
}

std::pair<KGraph, std::vector<size_t>> EmbedderEngine::build_mutual_kgraph(Gpu &gpu, size_t keep_n_edges,
                                                                           size_t n_neighbours, size_t n_contigs,
                                                                           size_t max_layers, size_t ef_construction,
                                                                           bool apply_filters, uint32_t variant) const {
    (void)n_contigs; (void)max_layers; (void)ef_construction;            // HNSW sizing: the lists here are exact
    const size_t n = contig_information.size();
    if (n < 2) throw Error("build_mutual_kgraph: at least 2 contigs are required");
    if (n_neighbours == 0) throw Error("build_mutual_kgraph: n_neighbours must be positive");
    // flatten the borrowed row views (recover_engine.rs:522-543 hands one ContigInformation per contig)
    const ContigInformation &first = contig_information[0].at(0);
    const size_t d = first.tnf_len, c = first.coverage_len;
    if (c % 2) throw Error("build_mutual_kgraph: coverage rows hold (mean, variance) pairs");
    std::vector<double> tnf(n * d), cov(n * c);
    for (size_t i = 0; i < n; ++i) {
        const ContigInformation &ci = contig_information[i].at(0);
        if (ci.tnf_len != d || ci.coverage_len != c) throw Error("build_mutual_kgraph: rows of unequal length");
        std::copy(ci.tnf, ci.tnf + d, tnf.begin() + i * d);
        std::copy(ci.coverage, ci.coverage + c, cov.begin() + i * c);
    }
    const size_t k = n_neighbours;
    std::vector<uint32_t> ti(n * k), di(n * k);
    std::vector<float> td(n * k), dd(n * k);
    bool sketched = false;
    for (size_t i = 0; i < n && !sketched; ++i) sketched = contig_information[i].at(0).sketch != nullptr;
    if (!sketched) {
        gpu.check(rg_knn_build(gpu.raw(), tnf.data(), cov.data(), n, (uint32_t)d, (uint32_t)(c / 2), (uint32_t)k, variant,
                               (1u << RG_GRAPH_TNF) | (1u << RG_GRAPH_DEPTH), ti.data(), td.data(), di.data(), dd.data()),
                  "rg_knn_build");                                       // embedder.rs:32-65
    } else {
        // DepthDistance with its MinHash factor (embedder.rs:147-164): the sketches go in between load and search
        std::vector<uint64_t> hashes, offs(n + 1, 0);
        for (size_t i = 0; i < n; ++i) {
            const ContigInformation &ci = contig_information[i].at(0);
            if (ci.sketch) hashes.insert(hashes.end(), ci.sketch, ci.sketch + ci.sketch_len);
            offs[i + 1] = hashes.size();
        }
        gpu.check(rg_knn_load(gpu.raw(), tnf.data(), cov.data(), n, (uint32_t)d, (uint32_t)(c / 2), variant), "rg_knn_load");
        gpu.check(rg_knn_load_sketches(gpu.raw(), hashes.data(), offs.data(), n, sketch_scale), "rg_knn_load_sketches");
        gpu.check(rg_knn_search(gpu.raw(), RG_GRAPH_DEPTH, (uint32_t)k, 0, n, di.data(), dd.data(), nullptr), "rg_knn_search");
        gpu.check(rg_knn_search(gpu.raw(), RG_GRAPH_TNF, (uint32_t)k, 0, n, ti.data(), td.data(), nullptr), "rg_knn_search");
    }
    KGraph depth_kgraph = kgraph_from_lists(di, dd, n, k);
    last_tnf_kgraph = kgraph_from_lists(ti, td, n, k);
    KGraph kgraph = intersect(gpu, std::move(depth_kgraph), last_tnf_kgraph,
                              std::max(keep_n_edges, DEFAULT_KEEP_N_EDGES), apply_filters);   // :68
    kgraph = mutual_knn(gpu, std::move(kgraph), keep_n_edges, apply_filters);                 // :69
    std::vector<size_t> disconnected_nodes;                              // :72-84
    for (size_t i = 0; i < n; ++i)
        if (kgraph.get_neighbours()[i].empty()) disconnected_nodes.push_back(i);
    return {std::move(kgraph), std::move(disconnected_nodes)};
}

}  // namespace rosella
