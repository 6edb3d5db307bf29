// sketcher.cpp — sketch_contigs / ContigSketcher (src/sketch/contig_sketcher.rs:11-92) and ContigSketchResult (:152-176)
// on top of rg_sketch_contigs: one device call for all records instead of one finch sketcher per record.
#include <algorithm>

#include "rosella_host.hpp"

namespace rosella {

ContigSketchResult sketch_contigs(Gpu &gpu, const SketchContigsArgs &m) {
    FastxRecords rec = parse_fastx_file(m.assembly);                         // :48
    // :53-55 records shorter than min_contig_size are skipped.  seqrec.seq().len(): needletail's seq() is the record's
    // sequence without its line terminators (raw_seq() would keep them), so a 60-column FASTA record counts its bases,
    // not its bytes
    FastxRecords kept;
    kept.offsets.push_back(0);
    for (size_t i = 0; i < rec.ids.size(); ++i) {
        const uint64_t b = rec.offsets[i], e = rec.offsets[i + 1];
        uint64_t seq_len = 0;
        for (uint64_t p = b; p < e; ++p) seq_len += !(rec.seq[p] == '\n' || rec.seq[p] == '\r');
        if (seq_len < m.min_contig_size) continue;
        kept.ids.push_back(rec.ids[i]);
        kept.seq.insert(kept.seq.end(), rec.seq.begin() + (std::ptrdiff_t)b, rec.seq.begin() + (std::ptrdiff_t)e);
        kept.offsets.push_back(kept.seq.size());
    }
    const size_t n = kept.ids.size();
    std::vector<uint64_t> sk_off(n + 1, 0), n_kmers(n ? n : 1, 0);
    uint64_t total = 0;
    gpu.check(rg_sketch_contigs(gpu.raw(), kept.seq.data(), kept.offsets.data(), n, (uint32_t)m.sketch_settings.kmer_size,
                                m.sketch_settings.sketch_scale, m.sketch_settings.seed, sk_off.data(), n_kmers.data(), &total),
              "rg_sketch_contigs");                                          // the loop :51-90
    std::vector<uint64_t> hashes(total ? total : 1);
    gpu.check(rg_sketch_fetch(gpu.raw(), hashes.data(), hashes.size()), "rg_sketch_fetch");
    ContigSketchResult out;
    out.contig_sketches.reserve(n);
    for (size_t i = 0; i < n; ++i) {
        Sketch s;
        s.name = kept.ids[i];                                                // :57,:81
        // :62,:76 total_bases of the normalised record: the bytes normalize(false) keeps
        uint64_t norm = 0;
        for (uint64_t p = kept.offsets[i]; p < kept.offsets[i + 1]; ++p) {
            const uint8_t c = kept.seq[p];
            norm += !(c == ' ' || c == '\t' || c == '\r' || c == '\n');
        }
        s.seq_length = norm;
        s.num_valid_kmers = n_kmers[i];
        s.hashes.assign(hashes.begin() + (std::ptrdiff_t)sk_off[i], hashes.begin() + (std::ptrdiff_t)sk_off[i + 1]);   // :77 to_vec()
        s.sketch_params = m.sketch_settings;
        s.kmers_to_sketch = (size_t)((double)norm * m.sketch_settings.sketch_scale);   // :65
        out.contig_sketches.push_back(std::move(s));
    }
    out.contig_names = std::move(kept.ids);                                  // :58
    return out;
}

// :158-176
void ContigSketchResult::filter_by_index(const std::set<size_t> &indices_to_remove) {
    std::vector<std::string> names;
    std::vector<Sketch> sketches;
    for (size_t i = 0; i < contig_names.size(); ++i) {
        if (indices_to_remove.count(i)) continue;
        names.push_back(std::move(contig_names[i]));
        sketches.push_back(std::move(contig_sketches[i]));
    }
    contig_names = std::move(names);
    contig_sketches = std::move(sketches);
}

}  // namespace rosella
