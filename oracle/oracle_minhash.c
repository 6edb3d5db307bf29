/*
 * oracle_minhash.c — CPU restatement of sketch_contigs.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Follows /root/reference/src/sketch/contig_sketcher.rs:47-92: per record, normalize(false) (:59), contig_size = the
 * normalised length (:62), kmers_to_sketch = (contig_size as f64 * sketch_scale) as usize (:65),
 * SketchParams::Scaled { kmers_to_sketch, kmer_length, scale, hash_seed } (:66-71), sketcher.process(&seqrec) (:74),
 * sketcher.to_vec() (:77).  Defaults: kmer_length 31, scale 0.1, seed 42 (src/cli.rs:798-811).
 *
 * finch ^0.6 is NOT under /root/reference (Cargo dependency, no lock file): PARITY UNPINNED for the scheme itself.  Its
 * published behaviour, restated here:
 *   - ScaledSketcher::process: for every canonical k-mer of the normalised record (needletail canonical_kmers(k, &rc):
 *     every window of k A/C/G/T bases, the lexicographically smaller of the window and its reverse complement),
 *     hash = MurmurHash3_x64_128(k-mer bytes, seed), first 64-bit half;
 *   - the sketch keeps every distinct hash <= max_hash = u64::MAX / (1/scale as u64), and while it holds no more than
 *     kmers_to_sketch hashes it also keeps larger ones (the heap is trimmed from the top while it is longer than
 *     kmers_to_sketch and its largest hash exceeds max_hash) — i.e. the result is the distinct hashes <= max_hash,
 *     padded with the next smallest distinct hashes up to kmers_to_sketch entries;
 *   - to_vec(): ascending by hash.
 * MurmurHash3_x64_128 is Austin Appleby's public-domain function, restated below and PINNED in tests/test_oracle.py
 * against the canonical MurmurHash3.cpp that scikit-learn vendors on this image (compiled by the test).
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

static inline uint64_t fmix64(uint64_t k) {
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdULL;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ULL;
    k ^= k >> 33;
    return k;
}

static inline uint64_t load_le64(const uint8_t *p, int nbytes) {
    uint64_t v = 0;
    for (int i = 0; i < nbytes; ++i) v |= (uint64_t)p[i] << (8 * i);
    return v;
}

/* both halves of MurmurHash3_x64_128 (seed: both state words start from it) */
void orc_murmur3_x64_128(const uint8_t *key, uint64_t len, uint64_t seed, uint64_t out[2]) {
    const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
    uint64_t h1 = seed, h2 = seed;
    const uint64_t nblocks = len / 16;
    for (uint64_t b = 0; b < nblocks; ++b) {
        uint64_t k1 = load_le64(key + 16 * b, 8), k2 = load_le64(key + 16 * b + 8, 8);
        k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
        h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729ULL;
        k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
        h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5ULL;
    }
    const uint8_t *tail = key + 16 * nblocks;
    const int t = (int)(len & 15);
    if (t > 8) {
        uint64_t k2 = load_le64(tail + 8, t - 8);
        k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    }
    if (t > 0) {
        uint64_t k1 = load_le64(tail, t > 8 ? 8 : t);
        k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    }
    h1 ^= len; h2 ^= len;
    h1 += h2; h2 += h1;
    h1 = fmix64(h1); h2 = fmix64(h2);
    h1 += h2; h2 += h1;
    out[0] = h1; out[1] = h2;
}

static int cmp_u64(const void *a, const void *b) {
    const uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : x > y;
}

static uint8_t comp(uint8_t c) {
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
        default: return c;
    }
}

uint64_t orc_sketch_max_hash(double scale) { return UINT64_MAX / (uint64_t)(1.0 / scale); }

/* One record -> its sketch.  out may be NULL (count only); at most cap hashes are written.  Returns the sketch length.
 * n_valid_kmers = k-mers hashed (finch total_kmers), norm_len = the normalised length. */
uint64_t orc_sketch_record(const uint8_t *seq, uint64_t len, uint32_t k, double scale, uint64_t seed, uint64_t *out,
                           uint64_t cap, uint64_t *n_valid_kmers, uint64_t *norm_len) {
    uint8_t *norm = (uint8_t *)malloc(len + 1), *rc = (uint8_t *)malloc(len + 1);
    const uint64_t L = orc_normalize(seq, len, norm);                        /* :59 */
    for (uint64_t i = 0; i < L; ++i) rc[i] = comp(norm[L - 1 - i]);
    const uint64_t size = (uint64_t)((double)L * scale);                    /* :65 */
    const uint64_t max_hash = orc_sketch_max_hash(scale);
    uint64_t *hs = (uint64_t *)malloc(sizeof(uint64_t) * (L + 1));
    uint64_t nh = 0, run = 0;
    for (uint64_t i = 0; i < L; ++i) {
        const uint8_t c = norm[i];
        run = (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? run + 1 : 0;
        if (run < k) continue;
        const uint64_t pos = i + 1 - k;
        const uint8_t *fw = norm + pos, *rv = rc + (L - pos - k);
        const uint8_t *km = memcmp(fw, rv, k) < 0 ? fw : rv;
        uint64_t h[2];
        orc_murmur3_x64_128(km, k, seed, h);
        hs[nh++] = h[0];
    }
    if (n_valid_kmers) *n_valid_kmers = nh;
    if (norm_len) *norm_len = L;
    qsort(hs, nh, sizeof(uint64_t), cmp_u64);
    uint64_t u = 0;
    for (uint64_t i = 0; i < nh; ++i)
        if (u == 0 || hs[u - 1] != hs[i]) hs[u++] = hs[i];
    uint64_t below = 0;
    while (below < u && hs[below] <= max_hash) below += 1;
    uint64_t keep = below >= size ? below : (size < u ? size : u);
    if (out)
        for (uint64_t i = 0; i < keep && i < cap; ++i) out[i] = hs[i];
    free(hs); free(rc); free(norm);
    return keep;
}

/* All records: sk_offsets [n + 1]; hashes may be NULL (first call: sizes only) */
void orc_sketch_contigs(const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t k, double scale, uint64_t seed,
                        uint64_t *sk_offsets, uint64_t *hashes, uint64_t *n_valid_kmers, uint64_t *norm_len) {
    if (!hashes) {
        sk_offsets[0] = 0;
#pragma omp parallel for schedule(dynamic, 8)
        for (int64_t r = 0; r < (int64_t)n; ++r)
            sk_offsets[r + 1] = orc_sketch_record(seq + offsets[r], offsets[r + 1] - offsets[r], k, scale, seed, NULL, 0,
                                                  n_valid_kmers ? n_valid_kmers + r : NULL, norm_len ? norm_len + r : NULL);
        for (uint64_t r = 0; r < n; ++r) sk_offsets[r + 1] += sk_offsets[r];
        return;
    }
#pragma omp parallel for schedule(dynamic, 8)
    for (int64_t r = 0; r < (int64_t)n; ++r)
        orc_sketch_record(seq + offsets[r], offsets[r + 1] - offsets[r], k, scale, seed, hashes + sk_offsets[r],
                          sk_offsets[r + 1] - sk_offsets[r], NULL, NULL);
}
