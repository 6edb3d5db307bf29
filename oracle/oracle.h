/*
 * oracle.h — CPU restatement of Rosella's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this.  The product (rosella_b200/) never links or calls it.
 *
 * Parity status:
 *   - flight variant of the coverage metric, rho and tnf_euclidean: PINNED against the
 *     reference's own numba functions (flight/flight/metrics.py) run in the build
 *     container; vectors committed under tests/golden/ (oracle/make_golden.py).
 *   - Rust variant (src/kmers/kmer_counting.rs, src/coverage/coverage_calculator.rs):
 *     "parity unpinned" by the reference itself — it ships no tests and no golden
 *     outputs, and no Rust toolchain exists here.  Anchored on SURVEY.md §4 known-answer
 *     values for the bundled FASTA and on the flight twin where the two agree.
 */
#ifndef ROSELLA_ORACLE_H
#define ROSELLA_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NBINS 136
#define ORC_VARIANT_RUST 0
#define ORC_VARIANT_FLIGHT 1

/* kmer_counting.rs:132-169 — the 136 sorted canonical 4-mers.  names: 136*4 bytes (no NUL). */
void orc_canonical_kmers(uint8_t *names /*136*4*/);
/* 256-entry table: 2-bit code (A=0,C=1,G=2,T=3, first base in the high bits) -> bin 0..135 */
void orc_code_to_bin(uint8_t *lut /*256*/);

/* needletail Sequence::normalize(false) semantics; returns normalised length. */
size_t orc_normalize(const uint8_t *seq, size_t len, uint8_t *out);

/* kmer_counting.rs:64-109, faithful: normalise, reverse-complement copy, per-window
 * lexicographic min(kmer, rc), table lookup by string.  seq holds raw record bytes
 * (may include newlines / lower case / IUPAC); record i = [offsets[i], offsets[i+1]).
 * freqs may be NULL.  n_kmers==0 -> freqs row is NaN (0/0) as in the reference. */
void orc_tnf_count(const uint8_t *seq, const uint64_t *offsets, uint64_t n,
                   uint32_t *counts /*n*136*/, uint64_t *n_kmers /*n*/, double *freqs /*n*136|NULL*/);
/* Same result through a 256->136 LUT + rolling code, OpenMP over records ("best reasonable CPU"). */
void orc_tnf_count_fast(const uint8_t *seq, const uint64_t *offsets, uint64_t n,
                        uint32_t *counts, uint64_t *n_kmers, double *freqs, int threads);

/* kmer_counting.rs:340-359 — CLR in place semantics: y_j = ln(x_j / mean(x)). */
double orc_ndarray_sum(const double *xs, uint32_t len);
void orc_clr(const double *in, double *out, uint64_t n, uint32_t d);
/* flight binning.py:192 — clr(x + 1) (skbio clr: ln x - mean(ln x)). */
void orc_clr_flight(const double *in, double *out, uint64_t n, uint32_t d);

/* kmer_counting.rs:366-414 */
double orc_kmer_correlation(const double *x, const double *y, uint32_t d);
/* flight metrics.py:601-645 (rho) and :64-77 (tnf_euclidean), without fastmath */
double orc_rho_flight(const double *x, const double *y, uint32_t d);
double orc_tnf_euclidean(const double *x, const double *y, uint32_t d);

/* coverage_calculator.rs:588-666 (variant RUST) / metrics.py:445-538 (variant FLIGHT).
 * a, b: 2*S doubles, interleaved mean,var. */
double orc_metabat(const double *a, const double *b, uint32_t s, int variant);

/* embedder.rs:108-127 / :133-171: f64 metric truncated to f32, NaN guards; sketch factor = 1. */
float orc_tnf_distance_f32(const double *x, const double *y, uint32_t d, int variant);
float orc_depth_distance_f32(const double *a, const double *b, uint32_t s, int variant);

/* Exhaustive exact top-k for query rows qrows[0..nq) (NULL = rows 0..nq) against all n rows.
 * which: 0 = TNF metric over tnf[n*d], 1 = depth metric over cov[n*2s], 2 = Euclidean distance over
 * tnf[n*d] (flight/flight/metrics.py:64-77, the index built at flight/flight/rosella/embedding.py:163).
 * Self excluded; ascending by (f32 distance, index); padding idx=UINT32_MAX, dist=+inf.
 * dist64 (optional) receives the untruncated f64 distances. */
void orc_knn(const double *data, uint64_t n, uint32_t width, int which, int variant,
             const uint64_t *qrows, uint64_t nq, uint32_t k,
             uint32_t *idx_out, float *dist_out, double *dist64, int threads);

/* nearest_neighbour_graph.rs:12-35 / :38-72 — what the functions COMPUTE (the reference then
 * drops the result).  Lists are n rows of k entries padded with UINT32_MAX.
 * keep[i*k+j] = 1 if edge j of node i survives. */
void orc_mutual_knn(const uint32_t *idx, uint64_t n, uint32_t k, uint32_t keep_n_edges, uint8_t *keep);
void orc_intersect(const uint32_t *base, const uint32_t *other, uint64_t n, uint32_t k,
                   uint32_t keep_n_edges, uint8_t *keep);

/* MinHash factor of DepthDistance (oracle_sketch.c): raw_distance / SketchDistance::eval
 * (src/sketch/sketch_distances.rs:61-126,:136-154) and DepthDistance::eval (src/embedding/embedder.rs:133-171).
 * Sketches are sorted u64 hash lists given by the caller (finch computes them on the reference side). */
double orc_min_jaccard(const uint64_t *q, uint64_t nq, const uint64_t *r, uint64_t nr, double scale);
float orc_sketch_distance_f32(const uint64_t *q, uint64_t nq, const uint64_t *r, uint64_t nr, double scale);
float orc_depth_distance_sketch_f32(const double *a, const double *b, uint32_t s, int variant,
                                    const uint64_t *qa, uint64_t na, const uint64_t *qb, uint64_t nb, double scale);
void orc_knn_depth_sketch(const double *cov, uint64_t n, uint32_t width, int variant, const uint64_t *hashes,
                          const uint64_t *offsets, double scale, const uint64_t *qrows, uint64_t nq, uint32_t k,
                          uint32_t *idx_out, float *dist_out, int threads);

/* sketch_contigs (src/sketch/contig_sketcher.rs:47-92) with finch's scaled MinHash restated (oracle_minhash.c; the scheme
 * is PARITY UNPINNED — finch is not under /root/reference; MurmurHash3_x64_128 is pinned against Appleby's source). */
void orc_murmur3_x64_128(const uint8_t *key, uint64_t len, uint64_t seed, uint64_t out[2]);
uint64_t orc_sketch_max_hash(double scale);
uint64_t orc_sketch_record(const uint8_t *seq, uint64_t len, uint32_t k, double scale, uint64_t seed, uint64_t *out,
                           uint64_t cap, uint64_t *n_valid_kmers, uint64_t *norm_len);
void orc_sketch_contigs(const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t k, double scale, uint64_t seed,
                        uint64_t *sk_offsets, uint64_t *hashes, uint64_t *n_valid_kmers, uint64_t *norm_len);

void orc_coverage_distance_condensed(const double *cov, uint64_t n, uint32_t s, double *out, int threads);

/* cluster_utils.rs:7-34 and :38-120 (silhouette with `a` over all other members, see oracle_sketch.c) */
void orc_condensed_euclidean(const double *data, uint64_t n, uint32_t d, double *out);
double orc_silhouette(const double *dist, uint64_t n, const int32_t *labels, uint32_t n_clusters, double *per_point);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
