/*
 * oracle_sketch.c — the MinHash factor of DepthDistance.  TEST INFRASTRUCTURE ONLY.
 *
 * Restates, from the reference's own source (present under /root/reference, no third-party arithmetic):
 *   raw_distance            src/sketch/sketch_distances.rs:61-126
 *   distance (min_scale)    src/sketch/sketch_distances.rs:6-46
 *   SketchDistance::eval    src/sketch/sketch_distances.rs:136-154
 *   DepthDistance::eval     src/embedding/embedder.rs:133-171
 * The sketches themselves (sorted u64 hash lists) are INPUTS here, as they are for the GPU path: finch ^0.6 computes
 * them on the reference side (src/sketch/contig_sketcher.rs:47-92) and its hash stream cannot be pinned in this
 * container (crate source absent) — "parity unpinned" for the sketching, exact restatement for their use.
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* raw_distance (:61-126): (i, j, common) after the merge walk and the scaled tail walk */
static void raw_walk(const uint64_t *q, uint64_t nq, const uint64_t *r, uint64_t nr, double scale,
                     uint64_t *pi, uint64_t *pj, uint64_t *pcommon) {
    uint64_t i = 0, j = 0, common = 0;
    while (i < nq && j < nr) {                                  /* :69-79 */
        if (q[i] < r[j]) i += 1;
        else if (q[i] > r[j]) j += 1;
        else { common += 1; i += 1; j += 1; }
    }
    if (scale > 0.) {                                           /* :83-98 */
        const uint64_t max_hash = UINT64_MAX / (uint64_t)(1.0 / scale);
        while (i < nq && q[i] < max_hash) i += 1;
        while (j < nr && r[j] < max_hash) j += 1;
    }
    *pi = i; *pj = j; *pcommon = common;
}

double orc_min_jaccard(const uint64_t *q, uint64_t nq, const uint64_t *r, uint64_t nr, double scale) {
    uint64_t i, j, common;
    raw_walk(q, nq, r, nr, scale, &i, &j, &common);
    if (i == 0 && j == 0) return 1.;                            /* :109-111 */
    return (double)common / (double)(i < j ? i : j);            /* :112 (0/0 = NaN when one side is 0) */
}

/* SketchDistance::eval on one pair (:136-154): (1 - min_jaccard) as f32, NaN -> 1 */
float orc_sketch_distance_f32(const uint64_t *q, uint64_t nq, const uint64_t *r, uint64_t nr, double scale) {
    float res = (float)(1.0 - orc_min_jaccard(q, nq, r, nr, scale));
    if (isnan(res)) res = 1.0f;
    return res;
}

/* DepthDistance::eval on one pair (embedder.rs:133-171) */
float orc_depth_distance_sketch_f32(const double *a, const double *b, uint32_t s, int variant,
                                    const uint64_t *qa, uint64_t na, const uint64_t *qb, uint64_t nb, double scale) {
    const double metabat = orc_metabat(a, b, s, variant);       /* :134-144 */
    float mj = orc_sketch_distance_f32(qa, na, qb, nb, scale);  /* :147-158 */
    if (isnan(mj)) mj = 1.0f;                                   /* :160-162 */
    const double result = metabat * (double)mj;                 /* :164 */
    if (isnan(result)) return 1.0f;                             /* :165-167 */
    return (float)result;                                       /* :169 */
}

typedef struct { float d; uint32_t i; } scand_t;
static int sless(float da, uint32_t ia, float db, uint32_t ib) { return da < db || (da == db && ia < ib); }

/* exhaustive exact top-k of the depth graph with sketches (hashes: concatenated sorted lists, offsets n+1) */
void orc_knn_depth_sketch(const double *cov, uint64_t n, uint32_t width, int variant, const uint64_t *hashes,
                          const uint64_t *offsets, double scale, const uint64_t *qrows, uint64_t nq, uint32_t k,
                          uint32_t *idx_out, float *dist_out, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel num_threads(threads)
#endif
    {
        scand_t *best = (scand_t *)malloc(sizeof(scand_t) * (k + 1));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
        for (int64_t qi = 0; qi < (int64_t)nq; ++qi) {
            const uint64_t q = qrows ? qrows[qi] : (uint64_t)qi;
            uint32_t cnt = 0;
            for (uint64_t j = 0; j < n; ++j) {
                if (j == q) continue;
                const float d = orc_depth_distance_sketch_f32(cov + q * width, cov + j * width, width / 2, variant,
                                                              hashes + offsets[q], offsets[q + 1] - offsets[q],
                                                              hashes + offsets[j], offsets[j + 1] - offsets[j], scale);
                if (cnt == k && !sless(d, (uint32_t)j, best[k - 1].d, best[k - 1].i)) continue;
                uint32_t p = cnt < k ? cnt : k - 1;
                while (p > 0 && sless(d, (uint32_t)j, best[p - 1].d, best[p - 1].i)) { best[p] = best[p - 1]; p -= 1; }
                best[p].d = d; best[p].i = (uint32_t)j;
                if (cnt < k) cnt += 1;
            }
            for (uint32_t p = 0; p < k; ++p) {
                idx_out[(uint64_t)qi * k + p] = p < cnt ? best[p].i : UINT32_MAX;
                dist_out[(uint64_t)qi * k + p] = p < cnt ? best[p].d : INFINITY;
            }
        }
        free(best);
    }
    (void)threads;
}

/* flight metrics.py:429-443 coverage_distance for all pairs i < j in scipy pdist order (oracle/rankstat.py) */
void orc_coverage_distance_condensed(const double *cov, uint64_t n, uint32_t s, double *out, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 8) num_threads(threads)
#endif
    for (int64_t i = 0; i < (int64_t)n - 1; ++i) {
        const double *a = cov + (uint64_t)i * 2 * s;
        uint64_t p = (uint64_t)i * n - (uint64_t)i * ((uint64_t)i + 1) / 2;
        for (uint64_t j = (uint64_t)i + 1; j < n; ++j, ++p) {
            const double *b = cov + j * 2 * s;
            int same = 1;
            for (uint32_t c = 0; c < 2 * s; ++c) same &= a[c] == b[c];
            out[p] = same ? 0.0 : orc_metabat(a, b, s, ORC_VARIANT_FLIGHT);
        }
    }
    (void)threads;
}

/* src/clustering/cluster_utils.rs:7-34 condensed_pairwise_distance: Euclidean, NaN -> 0 */
void orc_condensed_euclidean(const double *data, uint64_t n, uint32_t d, double *out) {
    uint64_t p = 0;
    for (uint64_t i = 0; i + 1 < n; ++i)
        for (uint64_t j = i + 1; j < n; ++j, ++p) {
            double r = 0.0;
            for (uint32_t c = 0; c < d; ++c) { const double t = data[i * d + c] - data[j * d + c]; r += t * t; }
            r = sqrt(r);
            out[p] = isnan(r) ? 0.0 : r;
        }
}

/* cluster_utils.rs:38-120 silhouette_score from a condensed distance vector, with `a` = mean over ALL other members of
 * the point's cluster (the reference's `indices.iter().skip(*i)` over a HashSet, :57, depends on the hash seed).
 * labels[i] in [0, n_clusters) or -1.  Sums run over members in increasing point id.  per_point may be NULL. */
double orc_silhouette(const double *dist, uint64_t n, const int32_t *labels, uint32_t n_clusters, double *per_point) {
    if (n_clusters <= 1) return 0.0;                                    /* :39-41 */
    double total = 0.0;
    double *sums = (double *)malloc(sizeof(double) * n_clusters);
    uint64_t *cnts = (uint64_t *)malloc(sizeof(uint64_t) * n_clusters);
    for (uint32_t c = 0; c < n_clusters; ++c)
        for (uint64_t i = 0; i < n; ++i) {                              /* clusters, then members: the order of :106-111 */
            if (labels[i] != (int32_t)c) continue;
            for (uint32_t k = 0; k < n_clusters; ++k) { sums[k] = 0.0; cnts[k] = 0; }
            for (uint64_t j = 0; j < n; ++j) {
                if (j == i || labels[j] < 0) continue;
                const uint64_t lo = i < j ? i : j, hi = i < j ? j : i;
                const uint64_t index = n * (n - 1) / 2 - (n - lo) * (n - lo - 1) / 2 + hi - lo - 1;   /* :126-132 */
                sums[labels[j]] += dist[index];
                cnts[labels[j]] += 1;
            }
            double a = sums[c] / (double)(cnts[c] ? cnts[c] : 1);       /* :66-69 */
            double b = -1.0;                                            /* :53 */
            for (uint32_t k = 0; k < n_clusters; ++k) {
                if (k == c) continue;
                const double m = sums[k] / (double)(cnts[k] ? cnts[k] : 1);   /* :84-88 */
                if (m < b || b < 0.0) b = m;                            /* :90-92 */
            }
            const double s = (b - a) / (a > b ? a : b);                 /* :94 */
            if (per_point) per_point[i] = s;
            total += s;
        }
    free(sums); free(cnts);
    return total / (double)n;                                           /* :119 */
}
