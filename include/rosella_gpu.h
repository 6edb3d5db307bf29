/*
 * rosella_gpu.h — C ABI of librosella_gpu.so: Rosella's data-parallel hot path on one B200.
 *
 * Drop-in boundary.  Each entry point names the reference interface it
 * replaces (paths relative to the rosella repository):
 *
 *   rg_tnf_*   replace the body of KmerCounter::run            src/kmers/kmer_counting.rs:41-123
 *              (per-record loop :64-109; canonical index :132-169)
 *   rg_knn_*   replace EmbedderEngine::build_mutual_kgraph     src/embedding/embedder.rs:31-86
 *              with the metrics TNFDistance::eval :108-127  (KmerCorrelation::distance,
 *              src/kmers/kmer_counting.rs:366-414) and DepthDistance::eval :133-171
 *              (MetabatDistance::distance, src/coverage/coverage_calculator.rs:588-666)
 *   rg_graph_* compute what mutual_knn / intersect compute    src/graphs/nearest_neighbour_graph.rs:12-72
 *   rg_sketch_* replace ContigSketcher::obtain_sketches        src/sketch/contig_sketcher.rs:47-92
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every host buffer; the library never keeps
 *     a caller pointer after a call returns; device memory belongs to the handle.
 *   - every function returns 0 on success or a negative rg_status; rg_last_error() has the text.
 *     No exception, abort() or exit() crosses the boundary.
 *   - NO CPU FALLBACK: without an sm_100 device rg_create fails with RG_ERR_NO_DEVICE.
 *   - a handle serves one caller thread at a time (the reference calls both entry points from
 *     its single pipeline thread, recover_engine.rs:111,:510).
 *   - one handle = one GPU.  More GPUs, all inside the library (NCCL over NVLink, dlopen'ed on first use):
 *       rg_multi_*        one process, one caller thread, n GPUs — what EmbedderEngine::build_mutual_kgraph's single
 *                         pipeline thread (recover_engine.rs:510-519) binds;
 *       rg_comm_* + rg_knn_load_dist / rg_knn_build_dist   one process per GPU (MPI / torchrun style launchers).
 */
#ifndef ROSELLA_GPU_H
#define ROSELLA_GPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RG_NBINS 136u          /* canonical 4-mers, kmer_counting.rs:11,:132-169 */
#define RG_ABI_VERSION 2

typedef struct rg_handle rg_handle;

typedef enum {
    RG_OK = 0,
    RG_ERR_NO_DEVICE = -1,     /* no CUDA device, or not compute capability 10.x */
    RG_ERR_INVALID = -2,       /* bad argument (null pointer, k == 0, shapes, non-finite coverage) */
    RG_ERR_CUDA = -3,          /* CUDA runtime error; text in rg_last_error */
    RG_ERR_OOM = -4,           /* device or host allocation failed */
    RG_ERR_STATE = -5          /* call order (e.g. search before load) */
} rg_status;

/* metric variant: which reference code the arithmetic follows */
#define RG_VARIANT_RUST 0u     /* coverage_calculator.rs:582-666 + kmer_counting.rs:366-414 */
#define RG_VARIANT_FLIGHT 1u   /* flight/flight/metrics.py:445-558 (metabat_distance_nn), :601-645 (rho) */

/* which kNN list */
#define RG_GRAPH_TNF 0u        /* TNFDistance,   embedder.rs:41-48 (tnf_nn) */
#define RG_GRAPH_DEPTH 1u      /* DepthDistance, embedder.rs:32-39 (depth_nn); MinHash factor: rg_knn_load_sketches */
#define RG_GRAPH_TNF_EUCLID 2u /* Euclidean distance of the TNF rows as passed: flight's second TNF index,
                                  flight/flight/rosella/embedding.py:163 (metrics.py:64-77); any variant;
                                  a NaN distance (non-finite row) is reported as FLT_MAX */

/* ------------------------------------------------------------------ lifecycle */
int32_t rg_abi_version(void);
int32_t rg_create(int32_t device_id, rg_handle **out);
void rg_destroy(rg_handle *h);
/* message of the last failing call on h (h == NULL: last failing rg_create on this thread) */
const char *rg_last_error(const rg_handle *h);
int32_t rg_sync(rg_handle *h);
/* the CUDA stream (cudaStream_t) all of the handle's work is issued on */
void *rg_stream(rg_handle *h);
/* Behaviour switches live on the handle; the library never reads the process environment.
 *   "depth_build"      0 automatic (default), 1 batch kernel, 2 symmetric block build for whole-graph searches
 *   "sym_block"        rows per block of the symmetric build (0 = 9472 = one wave of CTAs)
 *   "sym_chunk"        columns per evaluation launch of the symmetric build (0 = 4 blocks)
 *   "tnf_select_impl"  0 automatic (tcgen05 where its shape limits allow), 1 FP32 CUDA-core kernel
 *   "exact_compact"    1 (default) the exhaustive pass collects the rows that pass its FP32 prefilter and evaluates them
 *                      side by side; 0 = in place (A/B only)
 *   "certify"          1 (default) margin certificate + exhaustive f64 pass for uncertified rows; 0 = off (A/B only)
 *   "kprime_depth", "kprime_tnf"   candidate-list length override (0 = automatic; 48 (TNF only), 64, 128 or 256)
 *   "sketch_slice_bytes"  bytes of records rg_sketch_contigs processes per slice (0 = 256 MiB; the upload of the next
 *                      slice overlaps the hashing of the current one) */
int32_t rg_set_option(rg_handle *h, const char *key, int64_t value);

/* ------------------------------------------------------------------ TNF counting */
/* the 136 canonical 4-mers in bin order, 4 bytes each, no terminator (host only) */
int32_t rg_tnf_canonical_kmers(uint8_t *names /* 136*4 */);

/* Raw record bytes exactly as needletail yields them (may hold newlines, lower case, U, IUPAC,
 * gaps); record i = seq[offsets[i] .. offsets[i+1]).  Semantics of kmer_counting.rs:64-109:
 * counts[i*136+b] = occurrences of canonical 4-mer b, n_kmers[i] = counted windows,
 * freqs = count / n_kmers as f64 (NaN row when n_kmers == 0).  freqs may be NULL. */
int32_t rg_tnf_count_ascii(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n,
                           uint32_t *counts, uint64_t *n_kmers, double *freqs);

/* Packed layout streamed by the counting kernel: per record, normalised bases (whitespace
 * removed) 2 bits each, A=0 C=1 G=2 T=3, base j of a 32-base group in bits [2j, 2j+2) of a
 * uint64 word; validity 1 bit per base (1 = A/C/G/T) in a uint32 word; every record starts on a
 * word boundary, padding bases are invalid.  word_offsets has n+1 entries. */
uint64_t rg_tnf_packed_words(const uint8_t *seq, const uint64_t *offsets, uint64_t n);
int32_t rg_tnf_pack_host(const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint64_t *codes,
                         uint32_t *valid, uint64_t *word_offsets, int32_t threads);
int32_t rg_tnf_count_packed(rg_handle *h, const uint64_t *codes, const uint32_t *valid,
                            const uint64_t *word_offsets, uint64_t n, uint32_t *counts,
                            uint64_t *n_kmers, double *freqs);

/* staged form (inputs resident in HBM between calls) */
int32_t rg_tnf_load_packed(rg_handle *h, const uint64_t *codes, const uint32_t *valid,
                           const uint64_t *word_offsets, uint64_t n);
int32_t rg_tnf_load_ascii(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n);
int32_t rg_tnf_count_resident(rg_handle *h);   /* async on rg_stream; result stays on device */
int32_t rg_tnf_fetch(rg_handle *h, uint32_t *counts, uint64_t *n_kmers, double *freqs);

/* ------------------------------------------------------------------ kNN graph */
/* Copy the two tables to the device and pre-transform them.
 *   tnf  n x d f64 row-major, exactly the rows Rust hands to TNFDistance (already CLR'd under
 *        no_flight, kmer_counting.rs:322-323); d must be 136.
 *   cov  n x 2s f64 row-major, interleaved mean,var per sample as MetabatDistance reads them
 *        (coverage_calculator.rs:593-596).
 * Either may be NULL if that graph is never searched.  Non-finite coverage -> RG_ERR_INVALID
 * (the reference panics in Normal::new, coverage_calculator.rs:639); so is a negative mean (no read depth is negative,
 * and the FP32 select phase relies on mean + eps >= eps for :619).  The load also scans every coverage column for
 * f64-distinct values that collide in FP32 and keeps those row pairs as extra candidates of the f64 pass (knn_exact.cu). */
int32_t rg_knn_load(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                    uint32_t s, uint32_t variant);

/* The MinHash factor of DepthDistance (src/embedding/embedder.rs:133-171): with sketches loaded, RG_GRAPH_DEPTH is
 *     metabat x (1 - min_jaccard) as f32,   min_jaccard of raw_distance (src/sketch/sketch_distances.rs:61-126),
 * exactly what DepthDistance::eval returns; without them the factor is 1.  The sketches are the caller's: finch
 * computes them on the reference side (src/sketch/contig_sketcher.rs:47-92; Sketch.hashes, sorted and unique), row i =
 * hashes[offsets[i] .. offsets[i+1]), offsets[0] = 0, n = the loaded rows; scale = SketchParams::Scaled.scale (0.1 by
 * default, 0 = unscaled).  Call after rg_knn_load (a new rg_knn_load drops the sketches).  RG_ERR_INVALID when a list
 * is not strictly increasing or one hash occurs in more than 4096 sketches. 
 * hashes == NULL (with offsets from rg_sketch_contigs): use the sketches that call left on this handle's device —
 * no round trip through the host. */
int32_t rg_knn_load_sketches(rg_handle *h, const uint64_t *hashes, const uint64_t *offsets, uint64_t n, double scale);

/* sketch_contigs (src/sketch/contig_sketcher.rs:47-92): finch's scaled MinHash of every record, the sketches
 * rg_knn_load_sketches takes.  seq / offsets: the raw records exactly as for rg_tnf_count_ascii (normalize(false) of
 * :59 happens on the device); kmer_length <= 32 (reference default 31), scale in (0, 1] (0.1), seed (42)
 * (src/cli.rs:798-811).  Per record: every window of kmer_length A/C/G/T bases, the lexicographically smaller of the
 * window and its reverse complement, MurmurHash3_x64_128(bytes, seed) first half; the sketch = the distinct hashes
 * <= u64::MAX / (1/scale), padded with the next smallest distinct hashes up to kmers_to_sketch = (normalised length *
 * scale) as usize entries (:65-71), ascending (finch ^0.6 ScaledSketcher restated: the crate is not part of the
 * reference tree, see oracle/oracle_minhash.c).  sketch_offsets: n + 1 (host); n_valid_kmers (n, may be NULL): k-mers
 * hashed per record (finch total_kmers, :76); *n_hashes (may be NULL) = sketch_offsets[n].  The hashes stay on the
 * device until rg_sketch_fetch copies them (capacity in hashes, >= sketch_offsets[n]).  Records <= 2 GiB each.
 * rg_timings after the call: select_ms = device time of the whole call, of which h2d_ms = upload of the records,
 * transform_ms = the two hash passes + scan, rescore_ms = the segmented sort (first pass of every slice). */
int32_t rg_sketch_contigs(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t kmer_length,
                          double scale, uint64_t seed, uint64_t *sketch_offsets, uint64_t *n_valid_kmers, uint64_t *n_hashes);
int32_t rg_sketch_fetch(rg_handle *h, uint64_t *hashes, uint64_t capacity);

/* Exact top-k of query rows [q_begin, q_end) against all n loaded rows, self excluded, ascending
 * by (f32 distance, index).  idx/dist: (q_end-q_begin) x k; rows with fewer than k candidates are
 * padded with UINT32_MAX / +inf; n_valid (may be NULL) gets the count per row.
 * Distances are the reference metric evaluated in f64 and truncated to f32 (embedder.rs:125,:170).
 * The FP32 select phase is certified per row against its own f64 values; rows that fail are re-evaluated against all
 * n rows inside the same call (rg_timings.rows_flagged / rows_exact, rg_knn_flagged_rows). */
int32_t rg_knn_search(rg_handle *h, uint32_t which, uint32_t k, uint64_t q_begin, uint64_t q_end,
                      uint32_t *idx, float *dist, uint32_t *n_valid);
/* same, outputs are DEVICE pointers; asynchronous on rg_stream(h) */
int32_t rg_knn_search_dev(rg_handle *h, uint32_t which, uint32_t k, uint64_t q_begin, uint64_t q_end,
                          uint32_t *idx_dev, float *dist_dev, uint32_t *n_valid_dev);

/* A search of ALL rows of the coverage graph (q_begin = 0, q_end = n) evaluates each unordered pair once
 * (the metric is symmetric up to a sign rule that is tracked, knn_depth_sym.cu); the lists are the same.
 *
 * The same whole-graph coverage build split over n_parts processes (one GPU each, tables loaded in all):
 *   1. rg_knn_depth_partial_dev: part p evaluates its share of the row blocks against the rows to their right
 *      and writes, for EVERY row, a K'-entry candidate list sorted by (distance, index) and padded with
 *      (+inf, UINT32_MAX): cand_dist_dev / cand_idx_dev are device arrays [n][K'], K' = rg_knn_kprime(h, RG_GRAPH_DEPTH, k);
 *   2. the host exchanges the lists so that the owner of rows [q_begin, q_end) holds all parts' lists of those
 *      rows, [n_parts][q_end - q_begin][K'] (an all-to-all; rg_knn_build_dist below does all three steps inside the
 *      library with NCCL — these two calls remain for hosts that bring their own transport);
 *   3. rg_knn_depth_merge_dev merges them and finishes as rg_knn_search_dev does (f64 distances, same outputs).
 * Both calls are asynchronous on rg_stream(h).  Replaces nothing in the reference by itself: it is how
 * EmbedderEngine::build_mutual_kgraph (embedder.rs:31-86) scales past one GPU. */
uint32_t rg_knn_kprime(rg_handle *h, uint32_t which, uint32_t k);
int32_t rg_knn_depth_partial_dev(rg_handle *h, uint32_t k, uint32_t part, uint32_t n_parts,
                                 float *cand_dist_dev, uint32_t *cand_idx_dev);
int32_t rg_knn_depth_merge_dev(rg_handle *h, uint32_t k, uint32_t n_parts, const float *parts_dist_dev,
                               const uint32_t *parts_idx_dev, uint64_t q_begin, uint64_t q_end,
                               uint32_t *idx_dev, float *dist_dev, uint32_t *n_valid_dev);

/* one call: load + search of all rows for the graphs in which_mask (bit RG_GRAPH_*);
 * outputs per graph may be NULL when its bit is clear. */
int32_t rg_knn_build(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                     uint32_t s, uint32_t k, uint32_t variant, uint32_t which_mask,
                     uint32_t *tnf_idx, float *tnf_dist, uint32_t *depth_idx, float *depth_dist);

/* EmbedderEngine::build_mutual_kgraph (src/embedding/embedder.rs:31-86) in one call with its two filters APPLIED on
 * the device (the reference computes them and drops the result, nearest_neighbour_graph.rs:33,:70):
 *   kgraph = intersect(depth_kgraph, tnf_kgraph, max(keep_n_edges, 5))   (embedder.rs:68)
 *   kgraph = mutual_knn(kgraph, keep_n_edges)                            (embedder.rs:69)
 * Result as CSR in HOST memory: offsets[n+1]; neighbours / weights (capacity n*k each) hold node i's surviving edges
 * at [offsets[i], offsets[i+1]) in list order (ascending coverage distance, f32 of the f64 metric);
 * disconnected (capacity n, may be NULL) lists the nodes left without edges in ascending order
 * (embedder.rs:72-84), *n_disconnected their number. */
int32_t rg_knn_build_mutual(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                            uint32_t s, uint32_t k, uint32_t variant, uint32_t keep_n_edges,
                            uint64_t *offsets, uint32_t *neighbours, float *weights,
                            uint64_t *n_disconnected, uint32_t *disconnected);

/* The lists of the last rg_knn_build_mutual stay on the handle: the reference calls build_mutual_kgraph again with
 * only keep_n_edges changed (prefilter_nodes, recover_engine.rs:464-482; the embed retry loop, :438-459).  This
 * re-applies intersect + mutual_knn to the kept lists — no load, no search.  RG_ERR_STATE without kept lists. */
int32_t rg_knn_refilter(rg_handle *h, uint32_t keep_n_edges, uint64_t *offsets, uint32_t *neighbours, float *weights,
                        uint64_t *n_disconnected, uint32_t *disconnected);

/* build_mutual_kgraph of a SUBSET of the rows (evaluate_subset, recover_engine.rs:410-435: retrieve_contig_data
 * gathers the rows of the given indices, :522-543).  row_ids: m distinct rows of the n-row tables; the graph is
 * built over those m rows only and its node numbers are POSITIONS in row_ids (the caller maps them back exactly as
 * recover_engine.rs:422 does with get_data_id_from_idx).  Outputs as rg_knn_build (m x k per graph). */
int32_t rg_knn_build_subset(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                            const uint64_t *row_ids, uint64_t m, uint32_t k, uint32_t variant, uint32_t which_mask,
                            uint32_t *tnf_idx, float *tnf_dist, uint32_t *depth_idx, float *depth_dist);

/* ------------------------------------------------------------------ more than one GPU */
/* (a) One process per GPU.  Rank 0 makes an id and hands it to the others by any means (file, MPI, torch.distributed);
 * every rank then joins with its own handle.  All calls below are COLLECTIVE: every rank makes them in the same order.
 * A rank that fails before a collective makes every rank return an error (a status word is agreed on first). */
#define RG_COMM_ID_BYTES 128
int32_t rg_comm_unique_id(uint8_t id[RG_COMM_ID_BYTES]);
int32_t rg_comm_init(rg_handle *h, int32_t rank, int32_t world, const uint8_t id[RG_COMM_ID_BYTES]);
int32_t rg_comm_destroy(rg_handle *h);
/* rows [*begin, *end) of an n-row graph that rank owns (whole 32-row tiles, sizes differ by at most one tile) */
int32_t rg_comm_shard(uint64_t n, int32_t rank, int32_t world, uint64_t *begin, uint64_t *end);
/* rg_knn_load across the communicator: only `root` reads its host tables (the others may pass NULL for a table the
 * root passes); the f64 tables travel by ncclBroadcast over NVLink instead of one PCIe upload per rank.
 * root = -1: every rank holds the same host tables (ranks in one process do): each rank uploads only ITS row shard over
 * its own PCIe link and an in-place all-gather over NVLink completes the tables (rg_multi_* load this way). */
int32_t rg_knn_load_dist(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                         uint32_t variant, int32_t root);
/* Whole graph `which` of the loaded tables on all ranks.  Coverage graph: symmetric block build, every unordered
 * pair evaluated once on one rank, partial K' lists exchanged all-to-all (ncclSend/Recv), merged, re-evaluated in f64
 * on the owner of the row; TNF graphs: query-row shards.  gather = 1: idx / dist (n x k) are complete on every rank
 * (in-place all-gather of the slabs); gather = 0: only this rank's rows [begin, end) are written.
 * *_dev: device pointers, asynchronous on rg_stream(h).  Host flavour: pageable or pinned host arrays, synchronous. */
int32_t rg_knn_build_dist_dev(rg_handle *h, uint32_t which, uint32_t k, uint32_t gather, uint32_t *idx_dev, float *dist_dev);
int32_t rg_knn_build_dist(rg_handle *h, uint32_t which, uint32_t k, uint32_t gather, uint32_t *idx, float *dist);
typedef struct {
    double load_h2d_ms, load_bcast_ms, load_transform_ms;   /* last rg_knn_load_dist */
    double partial_ms;      /* last coverage build: this rank's share of the row blocks (evaluation + list reducers) */
    double exchange_ms;     /* all-to-all of the partial lists */
    double finish_ms;       /* merge + f64 re-evaluation + certificate (TNF graphs: select + re-evaluation) */
    double gather_ms;       /* all-gather of the finished slabs / device->host copy of this rank's slab */
} rg_dist_timings;
int32_t rg_get_dist_timings(rg_handle *h, rg_dist_timings *out);                      /* the last build */
int32_t rg_get_dist_timings_of(rg_handle *h, uint32_t which, rg_dist_timings *out);    /* the last build of graph `which` */

/* (b) One process, one caller thread, n_dev GPUs: one worker thread and one handle per device inside the library,
 * communicator made with ncclCommInitAll.  The calls mirror the single-GPU ones and return the same results;
 * rg_multi_tnf_count_ascii splits the records into n_dev contiguous ranges of equal bases (no exchange, every GPU
 * uploads its own range over its own PCIe link). */
typedef struct rg_multi rg_multi;
int32_t rg_multi_create(const int32_t *device_ids, int32_t n_dev, rg_multi **out);
void rg_multi_destroy(rg_multi *m);
const char *rg_multi_last_error(const rg_multi *m);
int32_t rg_multi_size(const rg_multi *m);
rg_handle *rg_multi_handle(rg_multi *m, int32_t i);     /* options, timings; do not issue work on it while m runs */
int32_t rg_multi_knn_build(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                           uint32_t k, uint32_t variant, uint32_t which_mask, uint32_t *tnf_idx, float *tnf_dist,
                           uint32_t *depth_idx, float *depth_dist);
/* the same with the rows' MinHash sketches (rg_knn_load_sketches): the depth graph is DepthDistance in full */
int32_t rg_multi_knn_build_sketched(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d, uint32_t s,
                                    uint32_t k, uint32_t variant, uint32_t which_mask, const uint64_t *hashes,
                                    const uint64_t *offsets, double scale, uint32_t *tnf_idx, float *tnf_dist,
                                    uint32_t *depth_idx, float *depth_dist);
int32_t rg_multi_knn_build_mutual(rg_multi *m, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                                  uint32_t s, uint32_t k, uint32_t variant, uint32_t keep_n_edges, uint64_t *offsets,
                                  uint32_t *neighbours, float *weights, uint64_t *n_disconnected, uint32_t *disconnected);
int32_t rg_multi_tnf_count_ascii(rg_multi *m, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t *counts,
                                 uint64_t *n_kmers, double *freqs);
/* rg_sketch_contigs / rg_sketch_fetch on all devices: contiguous ranges of records with equal bytes per device, no
 * exchange; the arrays are the caller's, global (contig_sketcher.rs:47-92). */
int32_t rg_multi_sketch_contigs(rg_multi *m, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t kmer_length,
                                double scale, uint64_t seed, uint64_t *sketch_offsets, uint64_t *n_valid_kmers, uint64_t *n_hashes);
int32_t rg_multi_sketch_fetch(rg_multi *m, uint64_t *hashes, uint64_t capacity);

/* ------------------------------------------------------------------ flight's all-pairs rank statistic */
/* ProfileDistanceEngine.makeRankStat (flight/flight/distance.py:92-108, as flight/flight/rosella/rosella.py:337-350
 * calls it: fun = a / max(a)): for every pair i < j
 *     stat = (cov_rank / max cov_rank)^w * (kmer_rank / max kmer_rank),   w = max(1/3, s / (s + 4)), 1/2 when s == 1,
 * cov_rank = fractional rank (ties averaged, distance.py:168-201) of metrics.coverage_distance among all pairs,
 * kmer_rank = the same for the Euclidean TNF distance with pair weights lengths[i] * lengths[j] (distance.py:53-58).
 * tnf: n x d rows as flight holds them (clr(x + 1)); cov: n x 2s interleaved mean, var; lengths: contigLen as f64.
 * Outputs (either may be NULL): stat_condensed, n (n - 1) / 2 values in scipy's pdist order — what the reference
 * returns; idx / stat_knn, n x k: per row the k pairs with the smallest statistic, ascending by (stat, index) — the
 * sparse form a kNN consumer needs instead of squareform(stat).  All f64 on the device; n <= 200000 and
 * 9 x 8 x n (n - 1) / 2 bytes of HBM (RG_ERR_OOM beyond).  Timings: select_ms = distances, rescore_ms = ranks. */
int32_t rg_rank_stat(rg_handle *h, const double *tnf, const double *cov, const double *lengths, uint64_t n, uint32_t d,
                     uint32_t s, uint32_t k, double *stat_condensed, uint32_t *idx, double *stat_knn);

/* ------------------------------------------------------------------ cluster search helpers */
/* condensed_pairwise_distance (src/clustering/cluster_utils.rs:7-34): Euclidean distance of all row pairs i < j of the
 * n x d f64 matrix, condensed order, NaN -> 0. */
int32_t rg_condensed_euclidean(rg_handle *h, const double *data, uint64_t n, uint32_t d, double *out /* n(n-1)/2 */);
/* silhouette_score (cluster_utils.rs:38-120) computed from the points, not from the O(n^2) distance vector:
 * labels[i] in [0, n_clusters) or -1 for a point in no cluster (an outlier: it gets no score and belongs to no cluster,
 * as in the reference's cluster map).  *score = sum of the per-point values / n (:119); per_point (may be NULL) gets s of
 * every clustered point, 0 elsewhere.  One stated deviation: `a` is the mean distance to ALL other members of the point's
 * cluster — the reference averages over `indices.iter().skip(*i)` of a HashSet (:57), which depends on the hash seed of
 * the process and is not reproducible.  select_ms of rg_timings = the kernel. */
int32_t rg_silhouette(rg_handle *h, const double *data, uint64_t n, uint32_t d, const int32_t *labels, uint32_t n_clusters,
                      double *score, double *per_point);

/* ------------------------------------------------------------------ graph post-ops */
/* keep[i*k+j] = 1 when edge j of node i survives mutual_knn / intersect
 * (nearest_neighbour_graph.rs:12-35 / :38-72; the reference computes these lists and then drops
 * them, SURVEY F5 — callers choose whether to apply them).  Host buffers. */
int32_t rg_graph_mutual(rg_handle *h, const uint32_t *idx, uint64_t n, uint32_t k,
                        uint32_t keep_n_edges, uint8_t *keep);
int32_t rg_graph_intersect(rg_handle *h, const uint32_t *base_idx, const uint32_t *other_idx,
                           uint64_t n, uint32_t k, uint32_t keep_n_edges, uint8_t *keep);
/* rg_graph_mutual without any device allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments and the same mask as rg_graph_mutual (host buffers), but the
 * device staging (lists + mask) is a fixed `static __device__` array inside the library: at most
 * RG_STATIC_MAX_EDGES = 2^20 edges (n * k), 5 MiB per GPU held until the process exits; larger inputs return
 * RG_ERR_INVALID ("use rg_graph_mutual").  The call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine.  All calls on one GPU share the array: calls that overlap on the same GPU (other handles, threads or
 * streams) race and must be serialised by the caller. */
#define RG_STATIC_MAX_EDGES ((uint64_t)1 << 20)
int32_t rg_graph_mutual_static(rg_handle *h, const uint32_t *idx, uint64_t n, uint32_t k,
                               uint32_t keep_n_edges, uint8_t *keep);

/* ------------------------------------------------------------------ instrumentation */
typedef struct {
    double h2d_ms;           /* last load: host->device copies */
    double transform_ms;     /* last load: pre-transform kernels */
    double select_ms;        /* last search: fused distance + top-k' kernel (device time) */
    double rescore_ms;       /* last search: f64 re-evaluation + final sort kernel */
    double d2h_ms;           /* last search / fetch: device->host copies */
    double tnf_count_ms;     /* last TNF counting kernel (device time) */
    uint64_t kernel_launches; /* kernels launched by this handle since creation */
    uint64_t h2d_bytes, d2h_bytes; /* since creation */
    /* exactness bookkeeping since the last rg_knn_load (ABI 2) */
    uint64_t rows_flagged;   /* query rows whose FP32 select margin could not be certified (or sit in a dense collision group) */
    uint64_t rows_exact;     /* query rows re-evaluated exhaustively in f64 against all n rows */
    uint64_t extra_pairs;    /* directed (row, candidate) pairs added because an FP32-rounded mean / variance collides */
} rg_timings;
int32_t rg_get_timings(rg_handle *h, rg_timings *out);
/* The rows of the LAST search whose select margin was not certified (they were re-evaluated exhaustively): local row
 * (bit 31: the row's FP32 operands collide wholesale), the k-th f64 distance among its candidates, its measured select
 * error.  Any output may be NULL; at most cap entries are written, *count gets the number found.  Synchronises. */
int32_t rg_knn_flagged_rows(rg_handle *h, uint32_t *rows, float *kth_dist, float *select_eps, uint64_t cap, uint64_t *count);

#ifdef __cplusplus
}
#endif
#endif /* ROSELLA_GPU_H */
