// capi.cpp — flat C entry points over the C++ host mirror (rosella_host.hpp) so that the Python test-suite can
// drive it through ctypes.  Test plumbing: the product boundary is include/rosella_gpu.h, and a C++ caller uses
// rosella_host.hpp directly.  Every function returns 0 / a handle on success and sets rh_last_error() otherwise.
#include <cstring>
#include <memory>

#include "rosella_host.hpp"

using namespace rosella;

static thread_local std::string g_err;

template <class F>
static int guard(F f) {
    try { f(); return 0; }
    catch (const std::exception &e) { g_err = e.what(); return -1; }
    catch (...) { g_err = "unknown C++ exception"; return -1; }
}

struct rh_graph {
    KGraph g, tnf;
    std::vector<size_t> disconnected;
};

extern "C" {

const char *rh_last_error(void) { return g_err.c_str(); }

int rh_format_f64(double v, char *buf, size_t cap) {
    return guard([&] {
        const std::string s = format_f64(v);
        if (s.size() + 1 > cap) throw Error("rh_format_f64: buffer too small");
        std::memcpy(buf, s.c_str(), s.size() + 1);
    });
}

size_t rh_clamp_n_neighbours(size_t n_included, size_t n_neighbours) { return clamp_n_neighbours(n_included, n_neighbours); }

// ---------------------------------------------------------------- KmerFrequencyTable
KmerFrequencyTable *rh_table_new(const double *data, size_t nrows, size_t ncols, const char *const *names) {
    KmerFrequencyTable *t = nullptr;
    if (guard([&] {
            auto p = std::make_unique<KmerFrequencyTable>();
            p->nrows = nrows; p->ncols = ncols;
            p->kmer_table.assign(data, data + nrows * ncols);
            for (size_t i = 0; i < nrows; ++i) p->contig_names.emplace_back(names[i]);
            t = p.release();
        })) return nullptr;
    return t;
}
void rh_table_free(KmerFrequencyTable *t) { delete t; }
int rh_table_write(KmerFrequencyTable *t, const char *path) { return guard([&] { t->write(path); }); }
KmerFrequencyTable *rh_table_read(const char *path, int apply_clr) {
    KmerFrequencyTable *t = nullptr;
    if (guard([&] { t = new KmerFrequencyTable(KmerFrequencyTable::read(path, apply_clr != 0)); })) return nullptr;
    return t;
}
size_t rh_table_nrows(const KmerFrequencyTable *t) { return t->nrows; }
size_t rh_table_ncols(const KmerFrequencyTable *t) { return t->ncols; }
size_t rh_table_kmer_size(const KmerFrequencyTable *t) { return t->kmer_size; }
const double *rh_table_data(const KmerFrequencyTable *t) { return t->kmer_table.data(); }
const char *rh_table_name(const KmerFrequencyTable *t, size_t i) { return t->contig_names.at(i).c_str(); }
const char *rh_table_path(const KmerFrequencyTable *t) { return t->table_path.c_str(); }
// returns the number of removed rows, or -1
long rh_table_filter_by_index(KmerFrequencyTable *t, const size_t *idx, size_t n) {
    long removed = -1;
    if (guard([&] { removed = (long)t->filter_by_index(std::set<size_t>(idx, idx + n)).size(); })) return -1;
    return removed;
}
long rh_table_filter_by_name(KmerFrequencyTable *t, const char *const *names, size_t n) {
    long removed = -1;
    if (guard([&] {
            std::set<std::string> s;
            for (size_t i = 0; i < n; ++i) s.insert(names[i]);
            removed = (long)t->filter_by_name(s).size();
        })) return -1;
    return removed;
}
int rh_clr(double *table, size_t nrows, size_t ncols) {
    return guard([&] {
        std::vector<double> v(table, table + nrows * ncols);
        KmerFrequencyTable::clr(v, nrows, ncols);
        std::copy(v.begin(), v.end(), table);
    });
}

// ---------------------------------------------------------------- FASTA / FASTQ
FastxRecords *rh_fastx_parse(const char *path) {
    FastxRecords *r = nullptr;
    if (guard([&] { r = new FastxRecords(parse_fastx_file(path)); })) return nullptr;
    return r;
}
void rh_fastx_free(FastxRecords *r) { delete r; }
size_t rh_fastx_n(const FastxRecords *r) { return r->ids.size(); }
const char *rh_fastx_id(const FastxRecords *r, size_t i) { return r->ids.at(i).c_str(); }
const uint8_t *rh_fastx_seq(const FastxRecords *r) { return r->seq.data(); }
const uint64_t *rh_fastx_offsets(const FastxRecords *r) { return r->offsets.data(); }

// ---------------------------------------------------------------- GPU entry points of the mirror
KmerFrequencyTable *rh_count_kmers(int device, const char *assembly, const char *output_directory, int no_flight) {
    KmerFrequencyTable *t = nullptr;
    if (guard([&] {
            Gpu gpu(device);
            t = new KmerFrequencyTable(count_kmers(gpu, CountKmersArgs{assembly, output_directory}, std::nullopt, no_flight != 0));
        })) return nullptr;
    return t;
}

struct rh_sketches {
    ContigSketchResult r;
    std::vector<uint64_t> offsets, hashes, num_valid_kmers, seq_length;
};
static void rh_sketches_flatten(rh_sketches *p);
rh_sketches *rh_sketch_contigs(int device, const char *assembly, size_t min_contig_size, size_t kmer_size, double scale, uint64_t seed) {
    rh_sketches *out = nullptr;
    if (guard([&] {
            Gpu gpu(device);
            SketchContigsArgs m{assembly, min_contig_size, SketchSettings{kmer_size, scale, seed}};
            auto p = std::make_unique<rh_sketches>();
            p->r = sketch_contigs(gpu, m);
            rh_sketches_flatten(p.get());
            out = p.release();
        })) return nullptr;
    return out;
}
// a ContigSketchResult from caller data (tests of filter_by_index need no GPU)
rh_sketches *rh_sketches_new(const char *const *names, const uint64_t *offsets, const uint64_t *hashes, size_t n) {
    rh_sketches *out = nullptr;
    if (guard([&] {
            auto p = std::make_unique<rh_sketches>();
            for (size_t i = 0; i < n; ++i) {
                Sketch sk;
                sk.name = names[i];
                sk.hashes.assign(hashes + offsets[i], hashes + offsets[i + 1]);
                p->r.contig_names.emplace_back(names[i]);
                p->r.contig_sketches.push_back(std::move(sk));
            }
            rh_sketches_flatten(p.get());
            out = p.release();
        })) return nullptr;
    return out;
}
static void rh_sketches_flatten(rh_sketches *p) {
    p->offsets.assign(1, 0); p->hashes.clear(); p->num_valid_kmers.clear(); p->seq_length.clear();
    for (const Sketch &s : p->r.contig_sketches) {
        p->hashes.insert(p->hashes.end(), s.hashes.begin(), s.hashes.end());
        p->offsets.push_back(p->hashes.size());
        p->num_valid_kmers.push_back(s.num_valid_kmers);
        p->seq_length.push_back(s.seq_length);
    }
}
int rh_sketches_filter_by_index(rh_sketches *s, const size_t *idx, size_t n) {
    return guard([&] {
        s->r.filter_by_index(std::set<size_t>(idx, idx + n));
        if (s->r.contig_names.size() != s->r.contig_sketches.size()) throw Error("ContigSketchResult: names and sketches differ in length");
        rh_sketches_flatten(s);
    });
}
void rh_sketches_free(rh_sketches *s) { delete s; }
size_t rh_sketches_n(const rh_sketches *s) { return s->r.contig_names.size(); }
const char *rh_sketches_name(const rh_sketches *s, size_t i) { return s->r.contig_names.at(i).c_str(); }
const uint64_t *rh_sketches_offsets(const rh_sketches *s) { return s->offsets.data(); }
const uint64_t *rh_sketches_hashes(const rh_sketches *s) { return s->hashes.data(); }
const uint64_t *rh_sketches_num_valid_kmers(const rh_sketches *s) { return s->num_valid_kmers.data(); }
const uint64_t *rh_sketches_seq_length(const rh_sketches *s) { return s->seq_length.data(); }

// ---------------------------------------------------------------- CoverageTable
CoverageTable *rh_cov_from_file(const char *path, int mode) {
    CoverageTable *t = nullptr;
    if (guard([&] {
            if (mode < 0 || mode > 3) throw Error("rh_cov_from_file: mode must be 0 ShortRead, 1 LongRead, 2 ShortBam, 3 LongBam");
            t = new CoverageTable(CoverageTable::from_file(path, (MappingMode)mode));
        })) return nullptr;
    return t;
}
CoverageTable *rh_cov_new(const double *table, size_t nrows, size_t ncols, const char *const *names, const size_t *lengths,
                          const char *const *samples) {
    CoverageTable *t = nullptr;
    if (guard([&] {
            auto p = std::make_unique<CoverageTable>();
            p->nrows = nrows; p->ncols = ncols;
            p->table.assign(table, table + nrows * ncols);
            for (size_t i = 0; i < nrows; ++i) {
                p->contig_names.emplace_back(names[i]);
                p->contig_lengths.push_back(lengths[i]);
                double sum = 0.0;
                for (size_t j = 0; j < ncols; j += 2) sum += table[i * ncols + j];
                p->average_depths.push_back(ncols ? sum / (double)(ncols / 2) : 0.0);
            }
            for (size_t j = 0; j < ncols / 2; ++j) p->sample_names.emplace_back(samples[j]);
            t = p.release();
        })) return nullptr;
    return t;
}
void rh_cov_free(CoverageTable *t) { delete t; }
int rh_cov_write(CoverageTable *t, const char *path) { return guard([&] { t->write(path); }); }
int rh_cov_merge(CoverageTable *t, const CoverageTable *other) { return guard([&] { t->merge(*other); }); }
long rh_cov_filter_by_length(CoverageTable *t, size_t min_contig_size) {
    long removed = -1;
    if (guard([&] { removed = (long)t->filter_by_length(min_contig_size).size(); })) return -1;
    return removed;
}
size_t rh_cov_nrows(const CoverageTable *t) { return t->nrows; }
size_t rh_cov_ncols(const CoverageTable *t) { return t->ncols; }
size_t rh_cov_n_samples(const CoverageTable *t) { return t->sample_names.size(); }
const double *rh_cov_data(const CoverageTable *t) { return t->table.data(); }
const double *rh_cov_average_depths(const CoverageTable *t) { return t->average_depths.data(); }
const size_t *rh_cov_lengths(const CoverageTable *t) { return t->contig_lengths.data(); }
const char *rh_cov_name(const CoverageTable *t, size_t i) { return t->contig_names.at(i).c_str(); }
const char *rh_cov_sample(const CoverageTable *t, size_t i) { return t->sample_names.at(i).c_str(); }
const char *rh_cov_path(const CoverageTable *t) { return t->output_path.c_str(); }

// ---------------------------------------------------------------- RecoverEngine's graph driver
RecoverEngine *rh_recover_new(const double *cov, size_t n, size_t c, const double *tnf, size_t d, const char *const *names,
                              const uint64_t *sketch_offsets, const uint64_t *sketch_hashes, size_t n_neighbours, int apply_filters,
                              uint32_t variant, double sketch_scale) {
    RecoverEngine *out = nullptr;
    if (guard([&] {
            CoverageTable ct;
            ct.nrows = n; ct.ncols = c;
            ct.table.assign(cov, cov + n * c);
            KmerFrequencyTable kt;
            kt.nrows = n; kt.ncols = d;
            kt.kmer_table.assign(tnf, tnf + n * d);
            ContigSketchResult sk;
            for (size_t i = 0; i < n; ++i) {
                ct.contig_names.emplace_back(names[i]);
                kt.contig_names.emplace_back(names[i]);
                ct.contig_lengths.push_back(0);
                double m = 0.0;
                for (size_t j = 0; j < c; j += 2) m += cov[i * c + j];
                ct.average_depths.push_back(c ? m / (double)(c / 2) : 0.0);
                if (sketch_offsets) {
                    Sketch one;
                    one.name = names[i];
                    if (sketch_hashes) one.hashes.assign(sketch_hashes + sketch_offsets[i], sketch_hashes + sketch_offsets[i + 1]);
                    sk.contig_names.emplace_back(names[i]);
                    sk.contig_sketches.push_back(std::move(one));
                }
            }
            auto p = std::make_unique<RecoverEngine>(std::move(ct), std::move(kt), std::move(sk));
            p->n_neighbours = n_neighbours;
            p->apply_filters = apply_filters != 0;
            p->variant = variant;
            p->sketch_scale = sketch_scale;
            out = p.release();
        })) return nullptr;
    return out;
}
// kmer_frequency_file may be NULL (count on the GPU)
RecoverEngine *rh_recover_from_files(int device, const char *assembly, const char *output_directory, const char *coverage_file, int mode,
                                     const char *kmer_frequency_file, size_t min_contig_size, size_t n_neighbours, int apply_filters) {
    RecoverEngine *out = nullptr;
    if (guard([&] {
            if (mode < 0 || mode > 3) throw Error("rh_recover_from_files: mode must be 0..3");
            Gpu gpu(device);
            RecoverArgs m;
            m.assembly = assembly; m.output_directory = output_directory; m.coverage_file = coverage_file;
            m.coverage_mode = (MappingMode)mode;
            if (kmer_frequency_file) m.kmer_frequency_file = std::string(kmer_frequency_file);
            m.min_contig_size = min_contig_size;
            m.n_neighbours = n_neighbours;
            auto p = std::make_unique<RecoverEngine>(RecoverEngine::from_files(gpu, m));
            p->apply_filters = apply_filters != 0;
            out = p.release();
        })) return nullptr;
    return out;
}
size_t rh_recover_cov_cols(const RecoverEngine *r) { return r->coverage_table.ncols; }
size_t rh_recover_tnf_cols(const RecoverEngine *r) { return r->tnf_table.ncols; }
void rh_recover_free(RecoverEngine *r) { delete r; }
size_t rh_recover_n_contigs(const RecoverEngine *r) { return r->n_contigs; }
size_t rh_recover_prefilter_rounds(const RecoverEngine *r) { return r->prefilter_rounds; }
size_t rh_recover_n_filtered(const RecoverEngine *r) { return r->filtered_contigs.size(); }
int rh_recover_is_filtered(const RecoverEngine *r, const char *name) { return r->filtered_contigs.count(name) ? 1 : 0; }
const char *rh_recover_name(const RecoverEngine *r, size_t i) { return r->coverage_table.contig_names.at(i).c_str(); }
const char *rh_recover_tnf_name(const RecoverEngine *r, size_t i) { return r->tnf_table.contig_names.at(i).c_str(); }
const double *rh_recover_cov_row(const RecoverEngine *r, size_t i) { return r->coverage_table.row(i); }
const double *rh_recover_tnf_row(const RecoverEngine *r, size_t i) { return r->tnf_table.row(i); }
long rh_recover_sketch_len(const RecoverEngine *r, size_t i) {
    return r->contig_sketches.contig_sketches.empty() ? -1 : (long)r->contig_sketches.contig_sketches.at(i).hashes.size();
}
// retrieve_contig_data(indices): per included row its first coverage / TNF value and sketch length, in order; returns the count
long rh_recover_retrieve(const RecoverEngine *r, const size_t *idx, size_t n, double *cov0, double *tnf0, long *sketch_len) {
    long cnt = -1;
    if (guard([&] {
            const auto data = r->retrieve_contig_data(std::set<size_t>(idx, idx + n));
            for (size_t i = 0; i < data.size(); ++i) {
                const ContigInformation &ci = data[i].at(0);
                cov0[i] = ci.coverage[0]; tnf0[i] = ci.tnf[0];
                sketch_len[i] = ci.sketch ? (long)ci.sketch_len : -1;
            }
            cnt = (long)data.size();
        })) return -1;
    return cnt;
}
// filter_contigs with a graph whose node i has data id i (what build_mutual_kgraph over all rows returns)
int rh_recover_filter_contigs(RecoverEngine *r, const size_t *nodes, size_t n) {
    return guard([&] {
        std::vector<size_t> ids(r->n_contigs);
        for (size_t i = 0; i < ids.size(); ++i) ids[i] = i;
        KGraph g(std::vector<std::vector<OutEdge>>(r->n_contigs), std::move(ids), 0);
        r->filter_contigs(g, std::vector<size_t>(nodes, nodes + n));
    });
}
// the disconnected list of the returned graph holds ORIGINAL row indices
rh_graph *rh_recover_evaluate_subset_graph(int device, const RecoverEngine *r, const size_t *idx, size_t n, size_t keep_n_edges) {
    rh_graph *out = nullptr;
    if (guard([&] {
            Gpu gpu(device);
            auto p = std::make_unique<rh_graph>();
            auto res = r->evaluate_subset_graph(gpu, std::set<size_t>(idx, idx + n), keep_n_edges);
            p->g = std::move(res.first);
            p->disconnected = std::move(res.second);
            out = p.release();
        })) return nullptr;
    return out;
}
rh_graph *rh_recover_prefilter_nodes(int device, RecoverEngine *r, size_t rounds) {
    rh_graph *out = nullptr;
    if (guard([&] {
            Gpu gpu(device);
            auto p = std::make_unique<rh_graph>();
            p->g = r->prefilter_nodes(gpu, rounds);
            out = p.release();
        })) return nullptr;
    return out;
}

// sketch_hashes / sketch_offsets (n + 1 entries) may be NULL: no sketches
rh_graph *rh_build_mutual_kgraph(int device, const double *tnf, size_t d, const double *cov, size_t c, size_t n,
                                 size_t keep_n_edges, size_t n_neighbours, size_t max_layers, size_t ef_construction,
                                 int apply_filters, uint32_t variant, const uint64_t *sketch_hashes,
                                 const uint64_t *sketch_offsets, double sketch_scale) {
    rh_graph *out = nullptr;
    if (guard([&] {
            Gpu gpu(device);
            std::vector<std::vector<ContigInformation>> info(n);
            for (size_t i = 0; i < n; ++i) {
                ContigInformation ci{cov + i * c, c, tnf + i * d, d};
                if (sketch_offsets) {
                    static const uint64_t none = 0;                      // an empty sketch still is a sketch
                    ci.sketch = sketch_hashes ? sketch_hashes + sketch_offsets[i] : &none;
                    ci.sketch_len = (size_t)(sketch_offsets[i + 1] - sketch_offsets[i]);
                }
                info[i].push_back(ci);
            }
            EmbedderEngine engine(std::move(info));
            engine.sketch_scale = sketch_scale;
            auto res = engine.build_mutual_kgraph(gpu, keep_n_edges, n_neighbours, n, max_layers, ef_construction,
                                                  apply_filters != 0, variant);
            out = new rh_graph{std::move(res.first), engine.last_tnf_kgraph, std::move(res.second)};
        })) return nullptr;
    return out;
}
void rh_graph_free(rh_graph *g) { delete g; }
size_t rh_graph_nb_nodes(const rh_graph *g) { return g->g.get_nb_nodes(); }
size_t rh_graph_max_nbng(const rh_graph *g) { return g->g.get_max_nbng(); }
size_t rh_graph_n_disconnected(const rh_graph *g) { return g->disconnected.size(); }
const size_t *rh_graph_disconnected(const rh_graph *g) { return g->disconnected.data(); }
// which: 0 = the returned kgraph, 1 = the TNF kgraph of the same build
size_t rh_graph_degree(const rh_graph *g, int which, size_t node) {
    return (which ? g->tnf : g->g).get_neighbours().at(node).size();
}
int rh_graph_neighbours(const rh_graph *g, int which, size_t node, size_t *nodes, double *weights) {
    return guard([&] {
        const auto &l = (which ? g->tnf : g->g).get_neighbours().at(node);
        for (size_t j = 0; j < l.size(); ++j) { nodes[j] = l[j].node; weights[j] = l[j].weight; }
    });
}
long rh_graph_data_id_from_idx(const rh_graph *g, size_t idx) {
    auto v = g->g.get_data_id_from_idx(idx);
    return v ? (long)*v : -1;
}

}  // extern "C"
