// coverage_table.cpp — CoverageTable I/O (src/coverage/coverage_table.rs:38-67, :135-387): the table the coverage
// half of the kNN path reads.  Restated with the reference's behaviour as it is, including two things its comments do
// not say: from_file stores a row as [cov_1..cov_S, var_1..var_S] (it pushes the coverage vector, then the
// variance vector, :175-187), not interleaved; and in the long-read layout the repeated length columns are not skipped
// (:245-251), so a long-read table of more than one sample does not reshape and from_file fails.
// TSV through Rust's csv crate: tab-delimited, first row is a header, fields quoted only when they hold a tab, a quote
// or a line break; every record must have as many fields as the header.
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include "rosella_host.hpp"

namespace rosella {

namespace {

// one csv record (tab-delimited, RFC 4180 quoting); returns false at end of input
bool read_record(std::istream &in, std::vector<std::string> &fields) {
    fields.clear();
    std::string cur;
    bool any = false, quoted = false;
    int ch;
    while ((ch = in.get()) != EOF) {
        any = true;
        const char c = (char)ch;
        if (quoted) {
            if (c == '"') {
                if (in.peek() == '"') { cur.push_back('"'); in.get(); }
                else quoted = false;
            } else cur.push_back(c);
            continue;
        }
        if (c == '"' && cur.empty()) { quoted = true; continue; }
        if (c == '\t') { fields.push_back(cur); cur.clear(); continue; }
        if (c == '\r') { if (in.peek() == '\n') in.get(); fields.push_back(cur); return true; }
        if (c == '\n') { fields.push_back(cur); return true; }
        cur.push_back(c);
    }
    if (!any) return false;
    fields.push_back(cur);
    return true;
}

bool blank(const std::vector<std::string> &f) { return f.size() == 1 && f[0].empty(); }   // csv skips empty lines

size_t parse_usize(const std::string &s) {             // str::parse::<usize>: digits with an optional '+', nothing else
    size_t i = (!s.empty() && s[0] == '+') ? 1 : 0;
    if (i == s.size()) throw Error("CoverageTable: cannot parse integer from \"" + s + "\"");
    size_t v = 0;
    for (; i < s.size(); ++i) {
        if (s[i] < '0' || s[i] > '9') throw Error("CoverageTable: invalid digit found in \"" + s + "\"");
        const size_t nv = v * 10 + (size_t)(s[i] - '0');
        if (nv < v) throw Error("CoverageTable: number too large: \"" + s + "\"");
        v = nv;
    }
    return v;
}

double parse_f64(const std::string &s) {               // str::parse::<f64>: decimal / exponent, inf, infinity, nan; no blanks, no hex
    if (s.empty() || s.find_first_of(" \t\r\nxXpP") != std::string::npos) throw Error("CoverageTable: invalid float literal \"" + s + "\"");
    char *end = nullptr;
    errno = 0;
    const double v = std::strtod(s.c_str(), &end);
    if (end != s.c_str() + s.size()) throw Error("CoverageTable: invalid float literal \"" + s + "\"");
    return v;
}

std::string sample_of(std::string header, bool first_word_only) {
    if (first_word_only) {                             // :230 split_whitespace().next(): drops " Mean" / " Variance"
        const size_t b = header.find_first_not_of(" \t\n\r");
        if (b == std::string::npos) throw Error("CoverageTable: empty sample column header");   // unwrap() on None
        const size_t e = header.find_first_of(" \t\n\r", b);
        const std::string word = header.substr(b, e == std::string::npos ? std::string::npos : e - b);
        if (header.find('/') == std::string::npos) return word;
        header = word;
    } else if (header.find('/') == std::string::npos) return header;
    // {reference}/{sample}.bam -> {sample} (:153-158, :231-236)
    std::string name = header.substr(header.rfind('/') + 1);
    for (size_t p; (p = name.find(".bam")) != std::string::npos;) name.erase(p, 4);
    return name;
}

std::string fixed3(double v) {                         // format!("{:.3}")
    if (std::isnan(v)) return "NaN";
    if (std::isinf(v)) return v < 0 ? "-inf" : "inf";
    char buf[400];
    std::snprintf(buf, sizeof(buf), "%.3f", v);
    return buf;
}

std::string quoted_if_needed(const std::string &f) {
    if (f.find_first_of("\t\"\n\r") == std::string::npos) return f;
    std::string out = "\"";
    for (char c : f) { if (c == '"') out.push_back('"'); out.push_back(c); }
    return out + "\"";
}

}  // namespace

CoverageTable CoverageTable::from_file(const std::string &file_path, MappingMode mode) {
    std::ifstream in(file_path, std::ios::binary);
    if (!in) throw Error("CoverageTable::from_file: cannot open " + file_path);
    const bool is_long = mode == MappingMode::LongBam || mode == MappingMode::LongRead;
    std::vector<std::string> f;
    CoverageTable t;
    if (!read_record(in, f)) f.clear();                // an empty file has an empty header
    const size_t n_fields = f.size();
    // :149 skip(3).step_by(2)  /  :228 skip(2).step_by(3)
    for (size_t i = is_long ? 2 : 3; i < n_fields; i += is_long ? 3 : 2) t.sample_names.push_back(sample_of(f[i], is_long));
    std::vector<double> flat;
    while (read_record(in, f)) {
        if (blank(f)) continue;
        if (f.size() != n_fields)
            throw Error("CoverageTable::from_file: found record with " + std::to_string(f.size()) + " fields, but the previous record has " +
                        std::to_string(n_fields) + " fields");
        size_t p = 0;
        if (p >= f.size()) throw Error("CoverageTable::from_file: record without a contig name");
        t.contig_names.push_back(f[p++]);
        if (p >= f.size()) throw Error("CoverageTable::from_file: record without a contig length");
        t.contig_lengths.push_back(parse_usize(f[p++]));
        double average_depth = 0.0;
        if (!is_long) {
            if (p >= f.size()) throw Error("CoverageTable::from_file: record without an average depth");
            average_depth = parse_f64(f[p++]);
        }
        std::vector<double> coverage, variance;
        for (size_t i = 0; p < f.size(); ++p, ++i) (i % 2 == 0 ? coverage : variance).push_back(parse_f64(f[p]));
        if (is_long) {                                  // :257 the mean of whatever landed in `coverage`
            double sum = 0.0;
            for (double v : coverage) sum += v;
            average_depth = sum / (double)coverage.size();
        }
        flat.insert(flat.end(), coverage.begin(), coverage.end());   // :185-186 / :259-260: coverages, then variances
        flat.insert(flat.end(), variance.begin(), variance.end());
        t.average_depths.push_back(average_depth);
    }
    t.nrows = t.contig_names.size();
    t.ncols = t.sample_names.size() * 2;
    if (flat.size() != t.nrows * t.ncols)              // Array2::from_shape_vec (:192-195 / :267-270)
        throw Error("CoverageTable::from_file: ShapeError: " + std::to_string(flat.size()) + " values do not fill " + std::to_string(t.nrows) +
                    " x " + std::to_string(t.ncols));
    t.table.swap(flat);
    t.output_path = file_path;
    return t;
}

// :38-52
std::set<std::string> CoverageTable::filter_by_length(size_t min_contig_size) {
    std::set<size_t> idx;
    for (size_t i = 0; i < contig_lengths.size(); ++i)
        if (contig_lengths[i] < min_contig_size) idx.insert(i);
    return filter_by_index(idx);
}

// :54-67
std::set<std::string> CoverageTable::get_contig_names(const std::set<size_t> &indices) const {
    std::set<std::string> out;
    for (size_t i : indices)
        if (i < contig_names.size()) out.insert(contig_names[i]);
    return out;
}

// :286-308
void CoverageTable::merge(const CoverageTable &other) {
    if (contig_names != other.contig_names) throw Error("Cannot merge coverage tables with different contig names");
    if (other.nrows != nrows) throw Error("CoverageTable::merge: ShapeError: tables differ in rows");
    sample_names.insert(sample_names.end(), other.sample_names.begin(), other.sample_names.end());
    std::vector<double> t(nrows * (ncols + other.ncols));
    for (size_t i = 0; i < nrows; ++i) {                // append along the columns (:296)
        std::copy(row(i), row(i) + ncols, t.begin() + (std::ptrdiff_t)(i * (ncols + other.ncols)));
        std::copy(other.row(i), other.row(i) + other.ncols, t.begin() + (std::ptrdiff_t)(i * (ncols + other.ncols) + ncols));
    }
    table.swap(t);
    ncols += other.ncols;
    for (size_t i = 0; i < nrows; ++i) {                // :300-305 every second value / number of samples
        double sum = 0.0;
        for (size_t j = 0; j < ncols; j += 2) sum += row(i)[j];
        if (i < average_depths.size()) average_depths[i] = sum / (double)sample_names.size();
        else average_depths.push_back(sum / (double)sample_names.size());
    }
}

// :315-329
CoverageTable CoverageTable::merge_many(std::vector<CoverageTable> coverage_tables) {
    if (coverage_tables.empty()) throw Error("CoverageTable::merge_many: no tables (unwrap on None)");
    CoverageTable merged = std::move(coverage_tables[0]);
    for (size_t i = 1; i < coverage_tables.size(); ++i) merged.merge(coverage_tables[i]);
    return merged;
}

// :335-377
void CoverageTable::write(const std::string &path) {
    output_path = path;
    std::ofstream out(path, std::ios::binary);
    if (!out) throw Error("CoverageTable::write: cannot create " + path);
    out << "contigName\tcontigLen\ttotalAvgDepth";
    for (const std::string &s : sample_names) out << '\t' << quoted_if_needed(s) << '\t' << quoted_if_needed(s + "-var");
    out << '\n';
    const size_t n = std::min(std::min(contig_names.size(), contig_lengths.size()), std::min(average_depths.size(), nrows));   // zip
    for (size_t i = 0; i < n; ++i) {
        out << quoted_if_needed(contig_names[i]) << '\t' << contig_lengths[i] << '\t' << fixed3(average_depths[i]);
        for (size_t j = 0; j < ncols; ++j) out << '\t' << fixed3(row(i)[j]);
        out << '\n';
    }
    out.flush();
    if (!out) throw Error("CoverageTable::write: write to " + path + " failed");
}

}  // namespace rosella
