// kmer_table.cpp — KmerFrequencyTable and count_kmers (src/kmers/kmer_counting.rs:13-16, :41-123, :211-360)
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

#include "rosella_host.hpp"

namespace rosella {

Gpu::Gpu(int device) {
    const int32_t rc = rg_create(device, &h_);
    if (rc != RG_OK) throw Error(std::string("rosella_gpu: ") + rg_last_error(nullptr));
}
Gpu::~Gpu() { rg_destroy(h_); }
void Gpu::check(int32_t rc, const char *what) const {
    if (rc != RG_OK) throw Error(std::string(what) + ": " + rg_last_error(h_));
}

// Shortest round-trip decimal, laid out as ryu's pretty printer (what Rust's csv crate emits for f64):
// integers get ".0", decimal notation while -5 < decimal-point position <= 16, else d.ddde[-]x.
std::string format_f64(double v) {
    if (std::isnan(v)) return "NaN";
    if (std::isinf(v)) return v < 0 ? "-inf" : "inf";
    if (v == 0.0) return std::signbit(v) ? "-0.0" : "0.0";
    char buf[64];
    auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    std::string s(buf, res.ptr);                       // [-]d[.ddd]e[+-]XX
    std::string out;
    size_t p = 0;
    if (s[0] == '-') { out = "-"; p = 1; }
    const size_t e = s.find('e');
    std::string digits;
    for (size_t i = p; i < e; ++i)
        if (s[i] != '.') digits.push_back(s[i]);
    const int exp10 = std::atoi(s.c_str() + e + 1);
    const int len = (int)digits.size();
    const int kk = exp10 + 1;                          // position of the decimal point
    if (len <= kk && kk <= 16) {                       // 1234e7 -> 12340000000.0
        out += digits + std::string(kk - len, '0') + ".0";
    } else if (0 < kk && kk <= 16) {                   // 1234e-2 -> 12.34
        out += digits.substr(0, kk) + "." + digits.substr(kk);
    } else if (-5 < kk && kk <= 0) {                   // 1234e-6 -> 0.001234
        out += "0." + std::string(-kk, '0') + digits;
    } else {                                           // 1.234e-7 / 1e30
        out += digits.substr(0, 1);
        if (len > 1) out += "." + digits.substr(1);
        out += "e" + std::to_string(kk - 1);
    }
    return out;
}

static std::string csv_field(const std::string &f) {   // csv crate: quote when the field needs it
    if (f.find_first_of(",\"\r\n") == std::string::npos && !f.empty()) return f;
    std::string q = "\"";
    for (char c : f) { if (c == '"') q.push_back('"'); q.push_back(c); }
    return q + "\"";
}

void KmerFrequencyTable::write(const std::string &output_file) {
    table_path = output_file;                                            // :286
    std::ofstream out(output_file, std::ios::binary);
    if (!out) throw Error("cannot create " + output_file);
    std::string line;
    for (size_t r = 0; r < nrows; ++r) {                                 // :289-291
        line = csv_field(contig_names[r]);
        for (size_t j = 0; j < ncols; ++j) { line.push_back(','); line += format_f64(row(r)[j]); }
        line.push_back('\n');
        out.write(line.data(), (std::streamsize)line.size());
    }
    out.flush();
    if (!out) throw Error("write failed: " + output_file);
}

static std::vector<std::string> split_csv(const std::string &line) {
    std::vector<std::string> f;
    std::string cur;
    bool quoted = false;
    for (size_t i = 0; i < line.size(); ++i) {
        const char c = line[i];
        if (quoted) {
            if (c == '"' && i + 1 < line.size() && line[i + 1] == '"') { cur.push_back('"'); ++i; }
            else if (c == '"') quoted = false;
            else cur.push_back(c);
        } else if (c == '"' && cur.empty()) quoted = true;
        else if (c == ',') { f.push_back(cur); cur.clear(); }
        else cur.push_back(c);
    }
    f.push_back(cur);
    return f;
}

KmerFrequencyTable KmerFrequencyTable::read(const std::string &input_file, bool apply_clr) {
    std::ifstream in(input_file, std::ios::binary);
    if (!in) throw Error("cannot open " + input_file);
    KmerFrequencyTable t;
    std::string line;
    while (std::getline(in, line)) {                                     // :304-308, no header (:300)
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        auto f = split_csv(line);
        if (f.size() < 2) throw Error("CSV deserialize error in " + input_file + ": expected name + values");
        if (t.nrows == 0) t.ncols = f.size() - 1;
        else if (f.size() - 1 != t.ncols) throw Error("CSV error: rows of unequal length in " + input_file);
        t.contig_names.push_back(f[0]);
        for (size_t j = 1; j < f.size(); ++j) {
            char *end = nullptr;
            const double v = std::strtod(f[j].c_str(), &end);            // accepts NaN / inf like f64::from_str
            if (end == f[j].c_str() || *end != '\0') throw Error("CSV deserialize error: invalid float '" + f[j] + "'");
            t.kmer_table.push_back(v);
        }
        t.nrows += 1;
    }
    if (t.nrows == 0) throw Error("empty k-mer table " + input_file);    // reference: index out of bounds panic (:310)
    t.kmer_size = (size_t)std::llround(std::log((double)t.ncols) / std::log(4.0));   // :312
    if (apply_clr) clr(t.kmer_table, t.nrows, t.ncols);                  // :322-323
    t.table_path = input_file;
    return t;
}

// ndarray ^0.15 ArrayBase::sum on a contiguous row (numeric_util::unrolled_fold): eight partial sums over strides
// of 8, folded as acc += (p0+p4), (p1+p5), (p2+p6), (p3+p7), then the < 8 remaining elements in order
static double ndarray_sum(const double *xs, size_t len) {
    double p[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc = 0.0;
    size_t i = 0;
    for (; i + 8 <= len; i += 8)
        for (int l = 0; l < 8; ++l) p[l] = p[l] + xs[i + l];
    acc = acc + (p[0] + p[4]);
    acc = acc + (p[1] + p[5]);
    acc = acc + (p[2] + p[6]);
    acc = acc + (p[3] + p[7]);
    for (; i < len; ++i) acc = acc + xs[i];
    return acc;
}

void KmerFrequencyTable::clr(std::vector<double> &table, size_t nrows, size_t ncols) {
    for (size_t r = 0; r < nrows; ++r) {                                 // :345-355
        double *row = table.data() + r * ncols;
        const double sum = ndarray_sum(row, ncols);                      // row.sum(), :345
        const double mean = sum / (double)ncols;
        for (size_t j = 0; j < ncols; ++j) row[j] = std::log(row[j] / mean);
    }
}

std::set<std::string> KmerFrequencyTable::filter_by_name(const std::set<std::string> &to_filter) {
    std::set<size_t> idx;                                                // :230-241
    for (size_t i = 0; i < contig_names.size(); ++i)
        if (to_filter.count(contig_names[i])) idx.insert(i);
    return filter_by_index(idx);
}

std::set<std::string> KmerFrequencyTable::filter_by_index(const std::set<size_t> &indices_to_remove) {
    std::vector<double> nt;                                              // :246-257
    std::vector<std::string> names;
    std::set<std::string> removed;
    size_t hit = 0;
    for (size_t i = 0; i < nrows; ++i) {
        if (indices_to_remove.count(i)) { removed.insert(contig_names[i]); hit += 1; continue; }
        nt.insert(nt.end(), row(i), row(i) + ncols);
        names.push_back(contig_names[i]);
    }
    if (hit != indices_to_remove.size())                                 // into_shape fails in the reference (:257)
        throw Error("filter_by_index: an index is out of range");
    kmer_table.swap(nt);
    contig_names.swap(names);
    nrows = contig_names.size();
    return removed;
}

static bool file_exists(const std::string &p) {
    struct stat st;
    return ::stat(p.c_str(), &st) == 0;
}

KmerFrequencyTable count_kmers(Gpu &gpu, const CountKmersArgs &m, std::optional<size_t> n_contigs, bool no_flight) {
    (void)n_contigs;   // capacity hint only (:56-62)
    const std::string output_file = m.output_directory + "/kmer_frequencies.tsv";   // :42
    if (file_exists(output_file)) return KmerFrequencyTable::read(output_file, no_flight);   // :46-50
    FastxRecords rec = parse_fastx_file(m.assembly);                     // :54
    const size_t n = rec.ids.size();
    std::vector<uint32_t> counts(n * RG_NBINS);
    std::vector<uint64_t> n_kmers(n);
    KmerFrequencyTable t;
    t.kmer_size = 4; t.nrows = n; t.ncols = RG_NBINS;
    t.kmer_table.assign(n * RG_NBINS, 0.0);
    gpu.check(rg_tnf_count_ascii(gpu.raw(), rec.seq.data(), rec.offsets.data(), n, counts.data(), n_kmers.data(),
                                 t.kmer_table.data()),
              "rg_tnf_count_ascii");                                     // the loop :64-109
    t.contig_names = std::move(rec.ids);
    t.write(output_file);                                                // :117
    return KmerFrequencyTable::read(output_file, no_flight);             // :120
}

}  // namespace rosella
