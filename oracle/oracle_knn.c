/*
 * oracle_knn.c — exhaustive exact top-k and the graph post-ops.  TEST INFRASTRUCTURE ONLY.
 *
 * The reference builds its kNN lists with an approximate index (hnsw_rs, embedder.rs:32-65);
 * BASELINE.json's north star makes the metric evaluated exhaustively the ground truth, so the
 * oracle evaluates every (query, reference) pair with the restated f64 metric, truncates to
 * f32 exactly as embedder.rs:125,:170 do, and keeps the k smallest by (distance, index).
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float d; uint32_t i; double d64; } cand_t;

static int cand_less(float da, uint32_t ia, float db, uint32_t ib) {
    return da < db || (da == db && ia < ib);
}

void orc_knn(const double *data, uint64_t n, uint32_t width, int which, int variant,
             const uint64_t *qrows, uint64_t nq, uint32_t k,
             uint32_t *idx_out, float *dist_out, double *dist64, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        cand_t *best = (cand_t *)malloc(sizeof(cand_t) * (k + 1));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t qi = 0; qi < (int64_t)nq; ++qi) {
            uint64_t q = qrows ? qrows[qi] : (uint64_t)qi;
            const double *qrow = data + q * width;
            uint32_t cnt = 0;
            for (uint64_t j = 0; j < n; ++j) {
                if (j == q) continue;
                const double *rrow = data + j * width;
                double d64;
                float d;
                if (which == 0) {
                    d64 = variant == ORC_VARIANT_FLIGHT ? orc_rho_flight(qrow, rrow, width)
                                                        : orc_kmer_correlation(qrow, rrow, width);
                    /* embedder.rs:122-125 (same guards as orc_tnf_distance_f32) */
                    d = isnan(d64) ? (variant == ORC_VARIANT_RUST ? 0.0f : 2.0f) : (float)d64;
                } else if (which == 1) {
                    d64 = orc_metabat(qrow, rrow, width / 2, variant);
                    /* embedder.rs:165-170, MinHash factor 1.0 */
                    d = isnan(d64) ? 1.0f : (float)d64;
                } else {
                    /* flight's second TNF index, rosella/embedding.py:163 (metrics.tnf_euclidean);
                     * a NaN distance (non-finite row) is defined as the largest finite f32 */
                    d64 = orc_tnf_euclidean(qrow, rrow, width);
                    d = isnan(d64) ? 3.402823466e+38f : (float)d64;
                }
                if (cnt == k && !cand_less(d, (uint32_t)j, best[k - 1].d, best[k - 1].i)) continue;
                uint32_t p = cnt < k ? cnt : k - 1;
                while (p > 0 && cand_less(d, (uint32_t)j, best[p - 1].d, best[p - 1].i)) {
                    best[p] = best[p - 1];
                    p -= 1;
                }
                best[p].d = d; best[p].i = (uint32_t)j; best[p].d64 = d64;
                if (cnt < k) cnt += 1;
            }
            for (uint32_t p = 0; p < k; ++p) {
                idx_out[(uint64_t)qi * k + p] = p < cnt ? best[p].i : UINT32_MAX;
                dist_out[(uint64_t)qi * k + p] = p < cnt ? best[p].d : INFINITY;
                if (dist64) dist64[(uint64_t)qi * k + p] = p < cnt ? best[p].d64 : INFINITY;
            }
        }
        free(best);
    }
    (void)threads;
}

static int list_has(const uint32_t *row, uint32_t k, uint32_t node) {
    for (uint32_t j = 0; j < k; ++j) {
        if (row[j] == UINT32_MAX) break;
        if (row[j] == node) return 1;
    }
    return 0;
}

/* nearest_neighbour_graph.rs:12-35: keep the first keep_n_edges, then only mutual edges */
void orc_mutual_knn(const uint32_t *idx, uint64_t n, uint32_t k, uint32_t keep_n_edges, uint8_t *keep) {
    for (uint64_t node = 0; node < n; ++node)
        for (uint32_t j = 0; j < k; ++j) {
            uint32_t nb = idx[node * k + j];
            uint8_t kp = 0;
            if (nb != UINT32_MAX) {
                if (j < keep_n_edges) kp = 1;                                    /* :19-22 */
                else kp = (uint8_t)list_has(idx + (uint64_t)nb * k, k, (uint32_t)node); /* :24 */
            }
            keep[node * k + j] = kp;
        }
}

/* nearest_neighbour_graph.rs:38-72: keep the first keep_n_edges of the base list, then only
 * edges whose reverse direction exists in the intersecting graph (data ids are the row ids
 * in both graphs, so get_idx_from_dataid is the identity) */
void orc_intersect(const uint32_t *base, const uint32_t *other, uint64_t n, uint32_t k,
                   uint32_t keep_n_edges, uint8_t *keep) {
    for (uint64_t node = 0; node < n; ++node)
        for (uint32_t j = 0; j < k; ++j) {
            uint32_t nb = base[node * k + j];
            uint8_t kp = 0;
            if (nb != UINT32_MAX) {
                if (j < keep_n_edges) kp = 1;                                      /* :50-53 */
                else kp = (uint8_t)list_has(other + (uint64_t)nb * k, k, (uint32_t)node); /* :57 */
            }
            keep[node * k + j] = kp;
        }
}
