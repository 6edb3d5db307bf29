// rosella_host.hpp — C++17 host-side mirror of the reference's Rust interface for the hot path, on top
// of the C ABI (include/rosella_gpu.h).  Same names, argument meaning and error behaviour as the
// reference so that a maintainer (and the parity tests) can read one against the other:
//
//   count_kmers / KmerCounter::run      src/kmers/kmer_counting.rs:13-16, :41-123
//   KmerFrequencyTable                  src/kmers/kmer_counting.rs:211-360
//   ContigInformation, EmbedderEngine   src/embedding/embedder.rs:15-104
//   sketch_contigs, ContigSketchResult  src/sketch/contig_sketcher.rs:11-92, :152-176
//   KGraph / OutEdge                    annembed::fromhnsw::kgraph (third party; the subset the path uses)
//   mutual_knn / intersect              src/graphs/nearest_neighbour_graph.rs:12-72
//   RecoverEngine graph driver          src/recover/recover_engine.rs:464-543 (prefilter_nodes, filter_contigs,
//                                       build_mutual_kgraph with the k clamp, retrieve_contig_data); CoverageTable subset
//
// Errors: the reference returns anyhow::Result; here every failure throws rosella::Error (the C wrapper
// in capi.cpp turns it into a status + message).  There is no CPU fallback: a Gpu is required.
#pragma once
#include <cstddef>
#include <cstdint>
#include <optional>
#include <set>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/rosella_gpu.h"

namespace rosella {

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// one rg_handle = one GPU; not copyable, one caller thread at a time
class Gpu {
   public:
    explicit Gpu(int device = 0);
    ~Gpu();
    Gpu(const Gpu &) = delete;
    Gpu &operator=(const Gpu &) = delete;
    rg_handle *raw() const { return h_; }
    void check(int32_t rc, const char *what) const;

   private:
    rg_handle *h_ = nullptr;
};

// ------------------------------------------------------------------ FASTA/FASTQ (needletail::parse_fastx_file)
struct FastxRecords {
    std::vector<std::string> ids;      // header line without '>' / '@' (needletail SequenceRecord::id)
    std::vector<uint8_t> seq;          // concatenated raw sequence bytes
    std::vector<uint64_t> offsets;     // ids.size() + 1
};
FastxRecords parse_fastx_file(const std::string &path);   // plain or gzip; throws on malformed input

// ------------------------------------------------------------------ KmerFrequencyTable (kmer_counting.rs:211-360)
struct KmerFrequencyTable {
    size_t kmer_size = 4;
    size_t nrows = 0, ncols = 0;
    std::vector<double> kmer_table;            // row-major nrows x ncols (ndarray standard layout)
    std::vector<std::string> contig_names;
    std::string table_path;

    const double *row(size_t i) const { return kmer_table.data() + i * ncols; }
    // :285-295 header-less CSV "name,v0,v1,..."; floats in shortest round-trip form (what csv + ryu write)
    void write(const std::string &output_file);
    // :298-333; apply_clr = the `no_flight` feature (:322-323)
    static KmerFrequencyTable read(const std::string &input_file, bool apply_clr);
    // :228-282; both return the names that were removed
    std::set<std::string> filter_by_name(const std::set<std::string> &to_filter);
    std::set<std::string> filter_by_index(const std::set<size_t> &indices_to_remove);
    // :340-359  y = ln(x / mean(x)) per row
    static void clr(std::vector<double> &table, size_t nrows, size_t ncols);
};

struct CountKmersArgs {          // the two clap arguments KmerCounter::new reads (:27-28)
    std::string assembly;
    std::string output_directory;
};
// count_kmers(m, n_contigs) (:13-16): reuses <out>/kmer_frequencies.tsv when present (:46-50), otherwise
// reads the assembly, counts on the GPU, writes the table and reads it back (:117-120)
KmerFrequencyTable count_kmers(Gpu &gpu, const CountKmersArgs &m, std::optional<size_t> n_contigs, bool no_flight);

// ------------------------------------------------------------------ sketching (src/sketch/contig_sketcher.rs)
struct SketchSettings {           // :133-150; clap defaults src/cli.rs:798-811
    size_t kmer_size = 31;
    double sketch_scale = 0.1;
    uint64_t seed = 42;
};
// finch::serialization::Sketch, the fields contig_sketcher.rs:80-88 fills and the path reads.  hashes = KmerCount.hash
// of to_vec(): ascending, distinct (the k-mer bytes and counts of KmerCount are not produced; SketchDistance reads
// the hash only, sketch_distances.rs:61-126).
struct Sketch {
    std::string name;
    uint64_t seq_length = 0;        // total_bases
    uint64_t num_valid_kmers = 0;   // total_kmers
    std::vector<uint64_t> hashes;
    SketchSettings sketch_params;   // SketchParams::Scaled { kmer_length, scale, hash_seed }
    size_t kmers_to_sketch = 0;     //                      { kmers_to_sketch } (:65)
};
struct ContigSketchResult {       // :152-176
    std::vector<std::string> contig_names;
    std::vector<Sketch> contig_sketches;
    void filter_by_index(const std::set<size_t> &indices_to_remove);
};
struct SketchContigsArgs {        // the clap arguments ContigSketcher::new reads (:23-26)
    std::string assembly;
    size_t min_contig_size = 1500;
    SketchSettings sketch_settings;
};
// sketch_contigs(m) (:11-14): ContigSketcher::run -> obtain_sketches (:47-92), all records in one device call
ContigSketchResult sketch_contigs(Gpu &gpu, const SketchContigsArgs &m);

// ------------------------------------------------------------------ graph types (annembed KGraph subset)
struct OutEdge {
    size_t node;
    double weight;
};
class KGraph {
   public:
    KGraph() = default;
    KGraph(std::vector<std::vector<OutEdge>> neighbours, std::vector<size_t> data_ids, size_t max_nbng);
    const std::vector<std::vector<OutEdge>> &get_neighbours() const { return neighbours_; }
    size_t get_nb_nodes() const { return neighbours_.size(); }
    size_t get_max_nbng() const { return max_nbng_; }
    std::optional<size_t> get_data_id_from_idx(size_t idx) const;
    std::optional<size_t> get_idx_from_dataid(size_t data_id) const;
    std::vector<std::vector<OutEdge>> &neighbours_mut() { return neighbours_; }

   private:
    std::vector<std::vector<OutEdge>> neighbours_;
    std::vector<size_t> data_ids_;      // IndexSet<DataId>: position = node index
    size_t max_nbng_ = 0;
};

// embedder.rs:89-104.  sketch: finch's Sketch.hashes of the contig (KmerCount.hash, sorted and unique as to_vec()
// returns them), nullptr = no sketch: DepthDistance then has the factor 1 for every pair.
struct ContigInformation {
    const double *coverage;
    size_t coverage_len;      // 2 x samples, interleaved mean,var
    const double *tnf;
    size_t tnf_len;
    const uint64_t *sketch = nullptr;
    size_t sketch_len = 0;
};

class EmbedderEngine {        // embedder.rs:15-87
   public:
    explicit EmbedderEngine(std::vector<std::vector<ContigInformation>> contig_information)
        : contig_information(std::move(contig_information)) {}
    // (kgraph, disconnected_nodes); max_layers / ef_construction / n_contigs are accepted and ignored (exact lists).
    // As in the reference, the graph returned is the DEPTH kNN graph: intersect() and mutual_knn() are
    // evaluated and their result dropped (nearest_neighbour_graph.rs:33,:70) unless apply_filters is set.
    std::pair<KGraph, std::vector<size_t>> build_mutual_kgraph(Gpu &gpu, size_t keep_n_edges, size_t n_neighbours,
                                                               size_t n_contigs, size_t max_layers,
                                                               size_t ef_construction, bool apply_filters = false,
                                                               uint32_t variant = RG_VARIANT_RUST) const;
    std::vector<std::vector<ContigInformation>> contig_information;
    double sketch_scale = 0.1;          // SketchParams::Scaled.scale of the sketches (contig_sketcher.rs:66-71; cli default 0.1)
    // the TNF graph of the last build (the reference builds it as tnf_kgraph, :65)
    mutable KGraph last_tnf_kgraph;
};

// ------------------------------------------------------------------ RecoverEngine's graph driver (recover_engine.rs:407-543)
enum class MappingMode { ShortRead, LongRead, ShortBam, LongBam };   // src/external/coverm_engine.rs:217-222
struct CoverageTable {            // src/coverage/coverage_table.rs:9-17
    size_t nrows = 0, ncols = 0;                 // ncols = 2 x samples; MetabatDistance reads mean, variance interleaved
    std::vector<double> table;                   // row-major
    std::vector<double> average_depths;
    std::vector<std::string> contig_names;
    std::vector<size_t> contig_lengths;
    std::vector<std::string> sample_names;
    std::string output_path;
    const double *row(size_t i) const { return table.data() + i * ncols; }
    std::set<std::string> filter_by_length(size_t min_contig_size);                      // :38-52
    std::set<std::string> get_contig_names(const std::set<size_t> &indices) const;       // :54-67
    std::set<std::string> filter_by_index(const std::set<size_t> &indices_to_remove);   // :69-131, returns the removed names
    // :135-284 coverm's TSV; a row comes out as [cov_1..cov_S, var_1..var_S], long-read tables of S > 1 fail
    static CoverageTable from_file(const std::string &file_path, MappingMode mode);
    void merge(const CoverageTable &other);                                              // :286-308
    static CoverageTable merge_many(std::vector<CoverageTable> coverage_tables);         // :315-329
    void write(const std::string &output_path);                                          // :335-377, values as {:.3}
};

struct RecoverArgs {              // the clap arguments RecoverEngine::new reads for the three tables (recover_engine.rs:76-131)
    std::string assembly, output_directory;
    std::string coverage_file;                   // --coverage-file: calculate_coverage reads it as MappingMode::ShortBam
    MappingMode coverage_mode = MappingMode::ShortBam;   // (coverage_calculator.rs:90); running coverm itself is out of scope
    std::optional<std::string> kmer_frequency_file;
    size_t min_contig_size = 1500;
    SketchSettings sketch_settings;
    size_t n_neighbours = 100, max_layers = 16, ef_construction = 100;
};

class RecoverEngine {             // `no_flight` members of recover_engine.rs:48-75 that the graph driver uses
   public:
    RecoverEngine(CoverageTable coverage, KmerFrequencyTable tnf, ContigSketchResult sketches);
    // RecoverEngine::new (:93-131) without the coverm subprocess: coverage table from its file, filter_by_length, TNF table
    // read or counted on the GPU (CLR applied, `no_flight`), filtered by the removed names, sketches on the GPU; the
    // reference's assert_eq! checks become Errors with the same messages
    static RecoverEngine from_files(Gpu &gpu, const RecoverArgs &m);
    // :522-543
    std::vector<std::vector<ContigInformation>> retrieve_contig_data(const std::set<size_t> &indices_to_include) const;
    // :510-519 (k clamp :513-517); node i of the graph = the i-th included row
    std::pair<KGraph, std::vector<size_t>> build_mutual_kgraph(Gpu &gpu, size_t keep_n_edges,
                                                               const std::set<size_t> &indices_to_include) const;
    // the graph half of evaluate_subset (:410-435; embedding and HDBSCAN are downstream): the graph of the included rows
    // and its disconnected nodes as ORIGINAL row indices.  Node i is the i-th included row in ascending order — that is
    // what retrieve_contig_data produces; the reference builds its position map from the HashSet's iteration order
    // (:413), which only agrees with that by accident.
    std::pair<KGraph, std::vector<size_t>> evaluate_subset_graph(Gpu &gpu, const std::set<size_t> &original_contig_indices,
                                                                 size_t keep_n_edges) const;
    KGraph prefilter_nodes(Gpu &gpu, size_t rounds);                                      // :464-482
    void filter_contigs(const KGraph &kgraph, const std::vector<size_t> &disconnected_nodes);   // :485-505

    CoverageTable coverage_table;
    KmerFrequencyTable tnf_table;
    ContigSketchResult contig_sketches;          // empty = no sketches (DepthDistance factor 1)
    size_t n_contigs = 0;
    size_t n_neighbours = 100, max_layers = 16, ef_construction = 100;   // cli.rs:841-866
    std::set<std::string> filtered_contigs;
    double sketch_scale = 0.1;
    uint32_t variant = RG_VARIANT_RUST;
    // false = the reference as it stands (intersect / mutual_knn results dropped, so an exact graph with n > k never has a
    // disconnected node and prefilter_nodes ends after one build); true = the filters its doc comments describe
    bool apply_filters = false;
    size_t prefilter_rounds = 0;                 // builds the last prefilter_nodes needed
};

// nearest_neighbour_graph.rs:12-35 / :38-72.  apply = false reproduces the reference (result dropped).
KGraph mutual_knn(Gpu &gpu, KGraph knn_graph, size_t keep_n_edges, bool apply = false);
KGraph intersect(Gpu &gpu, KGraph base_knn_graph, const KGraph &intersecting_graph, size_t keep_n_edges, bool apply = false);

// recover_engine.rs:513-517
inline size_t clamp_n_neighbours(size_t n_included, size_t n_neighbours) {
    return n_included < n_neighbours * 10 ? std::min(n_included / 2, n_neighbours) : n_neighbours;
}

std::string format_f64(double v);      // shortest round-trip text as Rust's csv/ryu writes it

}  // namespace rosella
