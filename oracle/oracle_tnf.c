/*
 * oracle_tnf.c — CPU restatement of TNF counting.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 * Follows /root/reference/src/kmers/kmer_counting.rs:41-169 and the needletail ^0.5
 * semantics the call sites rely on (:71 normalize(false), :74 reverse_complement,
 * :82 canonical_kmers(4, &rc)); needletail's source is not under /root/reference, its
 * published behaviour is restated here.
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const char BASES[4] = {'A', 'C', 'G', 'T'};

static uint8_t complement(uint8_t c) {
    /* needletail complement(): ACGT swap (+ IUPAC, which cannot survive normalize(false)). */
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
        default: return c; /* 'N' and '-' map to themselves */
    }
}

/* kmer_counting.rs:176-208 */
static void increment_kmer(uint8_t *kmer, int k) {
    int i = k - 1;
    for (;;) {
        if (kmer[i] == 'A') { kmer[i] = 'C'; break; }
        if (kmer[i] == 'C') { kmer[i] = 'G'; break; }
        if (kmer[i] == 'G') { kmer[i] = 'T'; break; }
        kmer[i] = 'A';
        if (i == 0) break;
        i -= 1;
    }
}

static int cmp4(const void *a, const void *b) { return memcmp(a, b, 4); }

/* kmer_counting.rs:132-169: enumerate 256 4-mers, keep lexicographic min(kmer, rc), sort. */
void orc_canonical_kmers(uint8_t *names) {
    uint8_t set[256 * 4];
    int n = 0;
    uint8_t kmer[4] = {'A', 'A', 'A', 'A'};
    for (int it = 0; it < 256; ++it) {
        uint8_t rc[4];
        for (int j = 0; j < 4; ++j) rc[j] = complement(kmer[3 - j]);
        const uint8_t *m = memcmp(kmer, rc, 4) < 0 ? kmer : rc;
        int found = 0;
        for (int q = 0; q < n && !found; ++q) found = memcmp(set + 4 * q, m, 4) == 0;
        if (!found) { memcpy(set + 4 * n, m, 4); n += 1; }
        increment_kmer(kmer, 4);
    }
    qsort(set, n, 4, cmp4);
    /* n is 136 by construction */
    memcpy(names, set, (size_t)ORC_NBINS * 4);
}

static int lookup_bin(const uint8_t *names, const uint8_t *kmer) {
    int lo = 0, hi = ORC_NBINS - 1;
    while (lo <= hi) {
        int mid = (lo + hi) / 2;
        int c = memcmp(names + 4 * mid, kmer, 4);
        if (c == 0) return mid;
        if (c < 0) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

void orc_code_to_bin(uint8_t *lut) {
    uint8_t names[ORC_NBINS * 4];
    orc_canonical_kmers(names);
    for (int code = 0; code < 256; ++code) {
        uint8_t kmer[4], rc[4];
        for (int j = 0; j < 4; ++j) kmer[j] = BASES[(code >> (2 * (3 - j))) & 3];
        for (int j = 0; j < 4; ++j) rc[j] = complement(kmer[3 - j]);
        const uint8_t *m = memcmp(kmer, rc, 4) < 0 ? kmer : rc;
        lut[code] = (uint8_t)lookup_bin(names, m);
    }
}

/* needletail normalize(seq, iupac=false) */
size_t orc_normalize(const uint8_t *seq, size_t len, uint8_t *out) {
    size_t o = 0;
    for (size_t i = 0; i < len; ++i) {
        uint8_t c = seq[i], r;
        switch (c) {
            case 'A': case 'C': case 'G': case 'T': case 'N': case '-': r = c; break;
            case 'a': r = 'A'; break;
            case 'c': r = 'C'; break;
            case 'g': r = 'G'; break;
            case 't': case 'u': case 'U': r = 'T'; break;
            case '.': case '~': r = '-'; break;
            case ' ': case '\t': case '\r': case '\n': r = ' '; break;
            default: r = 'N'; break;
        }
        if (r != ' ') out[o++] = r;
    }
    return o;
}

static int good_base(uint8_t c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }

void orc_tnf_count(const uint8_t *seq, const uint64_t *offsets, uint64_t n,
                   uint32_t *counts, uint64_t *n_kmers, double *freqs) {
    uint8_t names[ORC_NBINS * 4];
    orc_canonical_kmers(names);
    for (uint64_t r = 0; r < n; ++r) {
        size_t len = (size_t)(offsets[r + 1] - offsets[r]);
        uint8_t *norm = (uint8_t *)malloc(len + 1);
        uint8_t *rc = (uint8_t *)malloc(len + 1);
        size_t L = orc_normalize(seq + offsets[r], len, norm);          /* :71 */
        for (size_t i = 0; i < L; ++i) rc[i] = complement(norm[L - 1 - i]); /* :74 */
        uint32_t *row = counts + r * ORC_NBINS;
        memset(row, 0, sizeof(uint32_t) * ORC_NBINS);
        uint64_t nk = 0;
        /* needletail CanonicalKmers: every start position whose k bases are all ACGT */
        for (size_t pos = 0; pos + 4 <= L; ++pos) {
            if (!(good_base(norm[pos]) && good_base(norm[pos + 1]) && good_base(norm[pos + 2]) &&
                  good_base(norm[pos + 3])))
                continue;
            const uint8_t *fw = norm + pos;
            const uint8_t *rv = rc + (L - pos - 4);
            const uint8_t *km = memcmp(fw, rv, 4) < 0 ? fw : rv;
            int bin = lookup_bin(names, km);                            /* :88 */
            if (bin < 0) continue;                                      /* :96-97 (unreachable) */
            row[bin] += 1;                                              /* :100 */
            nk += 1;                                                    /* :101 */
        }
        n_kmers[r] = nk;
        if (freqs)
            for (int j = 0; j < ORC_NBINS; ++j)
                freqs[r * ORC_NBINS + j] = (double)row[j] / (double)nk; /* :104-107, 0/0 = NaN */
        free(norm);
        free(rc);
    }
}

void orc_tnf_count_fast(const uint8_t *seq, const uint64_t *offsets, uint64_t n,
                        uint32_t *counts, uint64_t *n_kmers, double *freqs, int threads) {
    uint8_t lut[256];
    orc_code_to_bin(lut);
    int8_t code_of[256];
    uint8_t skip[256];
    for (int c = 0; c < 256; ++c) {
        uint8_t in = (uint8_t)c, out;
        size_t o = orc_normalize(&in, 1, &out);
        skip[c] = (o == 0);
        code_of[c] = -1;
        if (o == 1) {
            if (out == 'A') code_of[c] = 0;
            else if (out == 'C') code_of[c] = 1;
            else if (out == 'G') code_of[c] = 2;
            else if (out == 'T') code_of[c] = 3;
        }
    }
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
#endif
    for (int64_t r = 0; r < (int64_t)n; ++r) {
        uint32_t *row = counts + (uint64_t)r * ORC_NBINS;
        memset(row, 0, sizeof(uint32_t) * ORC_NBINS);
        uint32_t code = 0, run = 0;
        uint64_t nk = 0;
        for (uint64_t i = offsets[r]; i < offsets[r + 1]; ++i) {
            uint8_t c = seq[i];
            if (skip[c]) continue;
            int b = code_of[c];
            if (b < 0) { run = 0; continue; }
            code = ((code << 2) | (uint32_t)b) & 0xFF;
            if (++run >= 4) { row[lut[code]] += 1; nk += 1; }
        }
        n_kmers[r] = nk;
        if (freqs)
            for (int j = 0; j < ORC_NBINS; ++j)
                freqs[(uint64_t)r * ORC_NBINS + j] = (double)row[j] / (double)nk;
    }
    (void)threads;
}
