// tnf_host.cpp — host-side pieces of the TNF path: the canonical 4-mer index
// (src/kmers/kmer_counting.rs:132-169), needletail's normalize(false) byte classes
// (call site :71), and the 2-bit + validity-bit packer that produces the layout the
// counting kernel streams.  No device code here.
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/rosella_gpu.h"
#include "tnf_tables.h"

namespace rg {

// 2-bit code, first base in the HIGH bits (so numeric order == lexicographic order of the string)
static inline uint32_t revcomp4(uint32_t code) {
    uint32_t rc = 0;
    for (int i = 0; i < 4; ++i) {
        const uint32_t b = (code >> (2 * i)) & 3u;   // base at position 3-i
        rc |= (3u - b) << (2 * (3 - i));             // complement, placed at position i
    }
    return rc;
}

void build_bin_tables(uint8_t *lut_be, uint8_t *lut_le, uint8_t *names) {
    // canonical form = lexicographic min of (kmer, reverse complement), :146-150; bins = sorted order, :158-166
    std::vector<uint32_t> canon;
    for (uint32_t c = 0; c < 256; ++c) {
        const uint32_t m = std::min(c, revcomp4(c));
        if (std::find(canon.begin(), canon.end(), m) == canon.end()) canon.push_back(m);
    }
    std::sort(canon.begin(), canon.end());
    static const char B[4] = {'A', 'C', 'G', 'T'};
    for (uint32_t c = 0; c < 256; ++c) {
        const uint32_t m = std::min(c, revcomp4(c));
        const uint8_t bin = (uint8_t)(std::lower_bound(canon.begin(), canon.end(), m) - canon.begin());
        if (lut_be) lut_be[c] = bin;
        if (lut_le) {
            // little-endian window: first base in the LOW bits (how the packed words are laid out)
            const uint32_t le = ((c >> 6) & 3u) | (((c >> 4) & 3u) << 2) | (((c >> 2) & 3u) << 4) | ((c & 3u) << 6);
            lut_le[le] = bin;
        }
    }
    if (names)
        for (size_t i = 0; i < canon.size(); ++i)
            for (int j = 0; j < 4; ++j) names[4 * i + j] = (uint8_t)B[(canon[i] >> (2 * (3 - j))) & 3u];
}

// needletail normalize(false): 0..3 = A C G T (also a c g t, u/U -> T); CLS_INVALID = any other byte that
// is kept (N, '-', '.', '~', IUPAC, ...: all break a k-mer window); CLS_SKIP = bytes that are removed.
void build_class_table(uint8_t *cls) {
    for (int c = 0; c < 256; ++c) cls[c] = CLS_INVALID;
    cls['A'] = cls['a'] = 0;
    cls['C'] = cls['c'] = 1;
    cls['G'] = cls['g'] = 2;
    cls['T'] = cls['t'] = cls['U'] = cls['u'] = 3;
    cls[' '] = cls['\t'] = cls['\r'] = cls['\n'] = CLS_SKIP;
}

}  // namespace rg

extern "C" {

int32_t rg_tnf_canonical_kmers(uint8_t *names) {
    if (!names) return RG_ERR_INVALID;
    rg::build_bin_tables(nullptr, nullptr, names);
    return RG_OK;
}

static uint64_t record_words(const uint8_t *cls, const uint8_t *seq, uint64_t b, uint64_t e) {
    uint64_t kept = 0;
    for (uint64_t i = b; i < e; ++i) kept += cls[seq[i]] != rg::CLS_SKIP;
    return (kept + 31) / 32;
}

uint64_t rg_tnf_packed_words(const uint8_t *seq, const uint64_t *offsets, uint64_t n) {
    if (!seq || !offsets) return 0;
    uint8_t cls[256];
    rg::build_class_table(cls);
    uint64_t words = 0;
    for (uint64_t r = 0; r < n; ++r) words += record_words(cls, seq, offsets[r], offsets[r + 1]);
    return words;
}

int32_t rg_tnf_pack_host(const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint64_t *codes,
                         uint32_t *valid, uint64_t *word_offsets, int32_t threads) {
    if (!seq || !offsets || !codes || !valid || !word_offsets) return RG_ERR_INVALID;
    uint8_t cls[256];
    rg::build_class_table(cls);
    if (threads <= 0) threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
    threads = (int32_t)std::min<uint64_t>((uint64_t)threads, std::max<uint64_t>(n, 1));
    // pass 1: words per record (parallel), then an exclusive scan
    std::vector<uint64_t> w(n);
    auto par = [&](auto fn) {
        std::atomic<uint64_t> next{0};
        std::vector<std::thread> pool;
        auto body = [&]() {
            for (;;) {
                const uint64_t b = next.fetch_add(64);
                if (b >= n) break;
                for (uint64_t r = b; r < std::min(n, b + 64); ++r) fn(r);
            }
        };
        for (int t = 1; t < threads; ++t) pool.emplace_back(body);
        body();
        for (auto &t : pool) t.join();
    };
    par([&](uint64_t r) { w[r] = record_words(cls, seq, offsets[r], offsets[r + 1]); });
    word_offsets[0] = 0;
    for (uint64_t r = 0; r < n; ++r) word_offsets[r + 1] = word_offsets[r] + w[r];
    // pass 2: pack
    par([&](uint64_t r) {
        uint64_t wi = word_offsets[r];
        uint64_t cw = 0;
        uint32_t vw = 0;
        int fill = 0;
        for (uint64_t i = offsets[r]; i < offsets[r + 1]; ++i) {
            const uint8_t c = cls[seq[i]];
            if (c == rg::CLS_SKIP) continue;
            if (c < 4) {
                cw |= (uint64_t)c << (2 * fill);
                vw |= 1u << fill;
            }
            if (++fill == 32) {
                codes[wi] = cw;
                valid[wi] = vw;
                ++wi;
                cw = 0; vw = 0; fill = 0;
            }
        }
        if (fill) { codes[wi] = cw; valid[wi] = vw; }
    });
    return RG_OK;
}

}  // extern "C"
