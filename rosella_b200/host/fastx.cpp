// fastx.cpp — the subset of needletail::parse_fastx_file the path uses (kmer_counting.rs:54,:64-68):
// FASTA ('>') and FASTQ ('@') records, plain or gzip, id = header line without its first byte,
// sequence bytes handed on raw (normalize(false) happens on the device).
#include <zlib.h>

#include <cstring>

#include "rosella_host.hpp"

namespace rosella {

namespace {
struct GzReader {
    gzFile f;
    std::vector<char> buf;
    explicit GzReader(const std::string &path) : f(gzopen(path.c_str(), "rb")), buf(1 << 20) {
        if (!f) throw Error("cannot open " + path);
        gzbuffer(f, 1 << 20);
    }
    ~GzReader() { gzclose(f); }
    // reads one line without its terminator; false at EOF
    bool line(std::string &out) {
        out.clear();
        for (;;) {
            if (!gzgets(f, buf.data(), (int)buf.size())) return !out.empty();
            const size_t n = std::strlen(buf.data());
            out.append(buf.data(), n);
            if (n && buf[n - 1] == '\n') break;
            if (gzeof(f)) break;
        }
        while (!out.empty() && (out.back() == '\n' || out.back() == '\r')) out.pop_back();
        return true;
    }
};
}  // namespace

FastxRecords parse_fastx_file(const std::string &path) {
    GzReader in(path);      // zlib reads plain files transparently
    FastxRecords r;
    r.offsets.push_back(0);
    std::string line;
    bool have = in.line(line);
    while (have && line.empty()) have = in.line(line);
    if (!have) throw Error("empty file " + path);                         // needletail: EmptyFile
    if (line[0] != '>' && line[0] != '@')
        throw Error("invalid start byte in " + path + ": expected '>' or '@'");   // needletail: InvalidStart
    const bool fastq = line[0] == '@';
    while (have) {
        if (line.empty()) { have = in.line(line); continue; }
        if (line[0] != (fastq ? '@' : '>')) throw Error("malformed record header in " + path);
        r.ids.emplace_back(line.substr(1));
        if (fastq) {
            std::string seq, plus, qual;
            if (!in.line(seq) || !in.line(plus) || !in.line(qual) || plus.empty() || plus[0] != '+')
                throw Error("truncated FASTQ record in " + path);
            if (qual.size() != seq.size()) throw Error("FASTQ sequence / quality length mismatch in " + path);
            r.seq.insert(r.seq.end(), seq.begin(), seq.end());
            have = in.line(line);
        } else {
            while ((have = in.line(line)) && (line.empty() || line[0] != '>')) {
                r.seq.insert(r.seq.end(), line.begin(), line.end());
                r.seq.push_back('\n');     // multi-line FASTA: the line break is part of the raw record
            }
        }
        r.offsets.push_back(r.seq.size());
    }
    return r;
}

}  // namespace rosella
