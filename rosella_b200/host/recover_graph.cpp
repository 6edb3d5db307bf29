// recover_graph.cpp — the graph driver of RecoverEngine under `no_flight` (src/recover/recover_engine.rs:407-543):
// retrieve_contig_data (:522-543), build_mutual_kgraph with the k clamp (:510-519), prefilter_nodes (:464-482),
// filter_contigs (:485-505), and CoverageTable::filter_by_index (src/coverage/coverage_table.rs:69-131).
// The graphs come from EmbedderEngine (embedder.cpp), i.e. from the GPU; everything else here is host bookkeeping.
#include <algorithm>
#include <filesystem>

#include "rosella_host.hpp"

namespace rosella {

// coverage_table.rs:69-131: rows, average depths, names and lengths lose the same positions; returns the removed names.
// As in the reference, an index past the end makes the reshape fail (:81-82).
std::set<std::string> CoverageTable::filter_by_index(const std::set<size_t> &indices_to_remove) {
    for (size_t i : indices_to_remove)
        if (i >= nrows) throw Error("CoverageTable::filter_by_index: index " + std::to_string(i) + " out of " + std::to_string(nrows) + " rows (ShapeError)");
    std::set<std::string> removed;
    std::vector<double> t, depths;
    std::vector<std::string> names;
    std::vector<size_t> lengths;
    t.reserve((nrows - indices_to_remove.size()) * ncols);
    for (size_t i = 0; i < nrows; ++i) {
        if (indices_to_remove.count(i)) {
            if (i < contig_names.size()) removed.insert(contig_names[i]);
            continue;
        }
        t.insert(t.end(), table.begin() + (std::ptrdiff_t)(i * ncols), table.begin() + (std::ptrdiff_t)((i + 1) * ncols));
        if (i < average_depths.size()) depths.push_back(average_depths[i]);
        if (i < contig_names.size()) names.push_back(contig_names[i]);
        if (i < contig_lengths.size()) lengths.push_back(contig_lengths[i]);
    }
    nrows -= indices_to_remove.size();
    table.swap(t); average_depths.swap(depths); contig_names.swap(names); contig_lengths.swap(lengths);
    return removed;
}

RecoverEngine::RecoverEngine(CoverageTable coverage, KmerFrequencyTable tnf, ContigSketchResult sketches)
    : coverage_table(std::move(coverage)), tnf_table(std::move(tnf)), contig_sketches(std::move(sketches)) {
    n_contigs = coverage_table.nrows;
    if (tnf_table.nrows != n_contigs) throw Error("RecoverEngine: the TNF table and the coverage table differ in rows");
    if (!contig_sketches.contig_sketches.empty() && contig_sketches.contig_sketches.size() != n_contigs)
        throw Error("RecoverEngine: one sketch per contig is required (or none at all)");
}

// recover_engine.rs:93-131
RecoverEngine RecoverEngine::from_files(Gpu &gpu, const RecoverArgs &m) {
    std::error_code ec;
    std::filesystem::create_directories(m.output_directory, ec);                         // :92 create_dir_all
    if (ec) throw Error("RecoverEngine: cannot create " + m.output_directory + ": " + ec.message());
    CoverageTable coverage_table = CoverageTable::from_file(m.coverage_file, m.coverage_mode);   // :95 (coverage_calculator.rs:90)
    const size_t n_contigs = coverage_table.nrows;                                       // :96
    const std::set<std::string> filtered_contigs = coverage_table.filter_by_length(m.min_contig_size);   // :99
    if (coverage_table.nrows != n_contigs - filtered_contigs.size())                     // :104
        throw Error("Coverage table row count and total contigs minus filtered contigs do not match.");
    KmerFrequencyTable tnf_table = m.kmer_frequency_file ? KmerFrequencyTable::read(*m.kmer_frequency_file, true)     // :105-107
                                                         : count_kmers(gpu, CountKmersArgs{m.assembly, m.output_directory}, n_contigs, true);   // :110
    if (n_contigs != tnf_table.nrows) throw Error("Coverage table row count and TNF table row count do not match.");   // :112
    tnf_table.filter_by_name(filtered_contigs);                                          // :114
    if (coverage_table.nrows != tnf_table.nrows) throw Error("Coverage table and TNF table have different number of contigs.");   // :115
    ContigSketchResult contig_sketches = sketch_contigs(gpu, SketchContigsArgs{m.assembly, m.min_contig_size, m.sketch_settings});  // :120
    if (coverage_table.nrows != contig_sketches.contig_sketches.size())                 // :122
        throw Error("Coverage table and contig sketches have different number of contigs.");
    RecoverEngine e(std::move(coverage_table), std::move(tnf_table), std::move(contig_sketches));
    e.n_neighbours = m.n_neighbours; e.max_layers = m.max_layers; e.ef_construction = m.ef_construction;
    e.sketch_scale = m.sketch_settings.sketch_scale;
    return e;                                                                            // :140 filtered_contigs starts empty
}

// :522-543 one ContigInformation per included row, in row order (borrowed views into the three tables)
std::vector<std::vector<ContigInformation>> RecoverEngine::retrieve_contig_data(const std::set<size_t> &indices_to_include) const {
    std::vector<std::vector<ContigInformation>> out;
    out.reserve(indices_to_include.size());
    const bool sketched = !contig_sketches.contig_sketches.empty();
    static const uint64_t none = 0;                                       // an empty sketch still is a sketch
    for (size_t row = 0; row < n_contigs; ++row) {
        if (!indices_to_include.count(row)) continue;                     // :531-533
        ContigInformation ci{coverage_table.row(row), coverage_table.ncols, tnf_table.row(row), tnf_table.ncols};
        if (sketched) {
            const std::vector<uint64_t> &hs = contig_sketches.contig_sketches[row].hashes;
            ci.sketch = hs.empty() ? &none : hs.data();
            ci.sketch_len = hs.size();
        }
        out.push_back({ci});
    }
    return out;
}

// :510-519
std::pair<KGraph, std::vector<size_t>> RecoverEngine::build_mutual_kgraph(Gpu &gpu, size_t keep_n_edges,
                                                                          const std::set<size_t> &indices_to_include) const {
    EmbedderEngine embedder_engine(retrieve_contig_data(indices_to_include));
    embedder_engine.sketch_scale = sketch_scale;
    const size_t k = clamp_n_neighbours(indices_to_include.size(), n_neighbours);       // :513-517
    return embedder_engine.build_mutual_kgraph(gpu, keep_n_edges, k, n_contigs, max_layers, ef_construction, apply_filters, variant);
}

// :410-435, the part before the embedding: subset graph + its disconnected nodes mapped back to original rows (:424-426)
std::pair<KGraph, std::vector<size_t>> RecoverEngine::evaluate_subset_graph(Gpu &gpu, const std::set<size_t> &original_contig_indices,
                                                                            size_t keep_n_edges) const {
    const std::vector<size_t> contig_id_map(original_contig_indices.begin(), original_contig_indices.end());   // position -> original
    auto [subset_kgraph, disconnected_nodes] = build_mutual_kgraph(gpu, keep_n_edges, original_contig_indices);
    std::vector<size_t> disconnected_indices;
    for (size_t i : disconnected_nodes)
        if (auto id = subset_kgraph.get_data_id_from_idx(i)) disconnected_indices.push_back(contig_id_map.at(*id));
    return {std::move(subset_kgraph), std::move(disconnected_indices)};
}

// :485-505 node index -> data id -> the same row of the three tables
void RecoverEngine::filter_contigs(const KGraph &kgraph, const std::vector<size_t> &disconnected_nodes) {
    std::set<size_t> original_indices;
    for (size_t idx : disconnected_nodes)
        if (auto id = kgraph.get_data_id_from_idx(idx)) original_indices.insert(*id);
    for (const std::string &name : coverage_table.filter_by_index(original_indices)) filtered_contigs.insert(name);
    if (!contig_sketches.contig_sketches.empty()) contig_sketches.filter_by_index(original_indices);
    tnf_table.filter_by_index(original_indices);
    n_contigs = coverage_table.nrows;      // the reference reads contig_sketches.contig_names.len() (:502): the same number
}

// :464-482 rebuild until no node is left without neighbours; after `rounds` rounds keep_n_edges grows by n_neighbours,
// which leaves no node disconnected
KGraph RecoverEngine::prefilter_nodes(Gpu &gpu, size_t rounds) {
    auto all = [&] { std::set<size_t> s; for (size_t i = 0; i < n_contigs; ++i) s.insert(s.end(), i); return s; };
    auto [kgraph, disconnected_nodes] = build_mutual_kgraph(gpu, 0, all());
    size_t current_round = 1, keep_n_edges = 0;
    prefilter_rounds = 1;
    while (!disconnected_nodes.empty()) {
        filter_contigs(kgraph, disconnected_nodes);
        current_round += 1;
        if (current_round > rounds) keep_n_edges += n_neighbours;        // :474-478
        std::tie(kgraph, disconnected_nodes) = build_mutual_kgraph(gpu, keep_n_edges, all());
        prefilter_rounds += 1;
    }
    return std::move(kgraph);
}

}  // namespace rosella
