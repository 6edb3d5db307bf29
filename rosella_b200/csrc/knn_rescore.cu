// knn_rescore.cu — K2c: re-evaluate the K' selected candidates of every query row with the
// reference's own f64 arithmetic, truncate to f32 as embedder.rs:125,:170 do, order by
// (f32 distance, index) and emit the first k.  Compiled with -fmad=false so that no
// multiply-add is contracted: the operation order below is the reference's, statement by
// statement (src/kmers/kmer_counting.rs:366-414, src/coverage/coverage_calculator.rs:588-666,
// flight/flight/metrics.py:445-558,:601-645).
#include <algorithm>

#include "knn_depth_common.cuh"
#include "metric_f64.cuh"
#include "rg_internal.cuh"

namespace {

using namespace metric64;

// the f32 distance the boundary reports for the pair (qrow, rrow): the reference's f64 metric truncated to f32 with
// the NaN guards of embedder.rs:122-125 / :165-170
template <int VARIANT, int WHICH>
__device__ __forceinline__ float pair_distance_f32(const double *__restrict__ qrow, const double *__restrict__ rrow,
                                                   uint32_t width) {
    if (WHICH == RG_GRAPH_TNF_EUCLID) {
        const double v = tnf_euclidean_f64(qrow, rrow, width);
        return isnan(v) ? 3.402823466e+38f : (float)v;
    }
    if (WHICH == RG_GRAPH_TNF) {
        const double v = VARIANT == RG_VARIANT_FLIGHT ? rho_flight_f64(qrow, rrow, width) : kmer_correlation_f64(qrow, rrow, width);
        // embedder.rs:122-125; FLIGHT: NaN is defined as the metric's maximum
        return isnan(v) ? (VARIANT == RG_VARIANT_RUST ? 0.0f : 2.0f) : (float)v;
    }
    const double v = metabat_f64<VARIANT>(qrow, rrow, width / 2);
    return isnan(v) ? 1.0f : (float)v;                                     // embedder.rs:165-170
}

// MinHash sketches of the rows (rg_knn_load_sketches); hash == nullptr: no sketches, factor 1
struct SketchArgs {
    const uint64_t *hash, *off;
    uint64_t max_hash;
};

__device__ __forceinline__ uint64_t sk_lower_bound(const uint64_t *p, uint64_t n, uint64_t key, bool upper) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if (p[mid] < key || (upper && p[mid] == key)) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// SketchDistance::eval for one pair (sketch_distances.rs:136-154): (1 - min_jaccard) as f32, NaN -> 1, with
// min_jaccard = common / min(i, j) of raw_distance (:61-126).  `common` comes from the hash join (knn_exact.cu); i and
// j are where the reference's merge walk and its scaled tail walk stop: the list that ends first is consumed whole,
// the other one up to that list's last hash, then both advance over their hashes below max_hash.
__device__ float sketch_factor(const SketchArgs &sk, uint32_t q, uint32_t j, uint32_t common) {
    const uint64_t qb = sk.off[q], nq = sk.off[q + 1] - qb, rb = sk.off[j], nr = sk.off[j + 1] - rb;
    if (common == 0 && nq && nr) return 1.0f;               // 0 / min(i, j), or 0 / 0 = NaN -> 1: the factor is exactly 1
    const uint64_t *qh = sk.hash + qb, *rh = sk.hash + rb;
    uint64_t i = 0, jj = 0;
    if (nq && nr) {
        const uint64_t ql = qh[nq - 1], rl = rh[nr - 1];
        if (ql < rl) { i = nq; jj = sk_lower_bound(rh, nr, ql, true); }
        else if (ql > rl) { jj = nr; i = sk_lower_bound(qh, nq, rl, true); }
        else { i = nq; jj = nr; }
    } else {
        common = 0;
    }
    if (sk.max_hash) {                                       // :83-98
        const uint64_t iq = sk_lower_bound(qh, nq, sk.max_hash, false), ir = sk_lower_bound(rh, nr, sk.max_hash, false);
        if (iq > i) i = iq;
        if (ir > jj) jj = ir;
    }
    double mj;
    if (i == 0 && jj == 0) mj = 1.0;                         // :109-111
    else mj = (double)common / (double)(i < jj ? i : jj);    // :112
    float res = (float)(1.0 - mj);                           // :143
    if (isnan(res)) res = 1.0f;                              // :150-152
    return res;
}

// DepthDistance::eval for one pair (embedder.rs:133-171) -> the f32 the boundary reports; *metabat32 gets the f32 of
// the coverage term alone (what the select phase estimated)
template <int VARIANT>
__device__ float depth_pair(const double *__restrict__ qrow, const double *__restrict__ rrow, uint32_t width,
                            const SketchArgs &sk, uint32_t q, uint32_t j, uint32_t common, float *metabat32) {
    const double m = metabat_f64<VARIANT>(qrow, rrow, width / 2);
    *metabat32 = isnan(m) ? 1.0f : (float)m;
    if (!sk.hash) return *metabat32;                         // no sketches: factor 1 (embedder.rs:165-170)
    const double res = m * (double)sketch_factor(sk, q, j, common);   // :164
    return isnan(res) ? 1.0f : (float)res;                   // :165-169
}

// number of sketch hashes rows (grow, j) share: j's entry among the sorted extra candidates of grow
__device__ __forceinline__ uint32_t common_of(const uint32_t *extra_idx, const uint32_t *extra_common, uint64_t eo, uint32_t cnt,
                                              uint32_t j, bool *joined) {
    uint32_t lo = 0, hi = cnt;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (extra_idx[eo + mid] < j) lo = mid + 1; else hi = mid;
    }
    const bool hit = lo < cnt && extra_idx[eo + lo] == j;
    if (joined) *joined = hit;
    return hit && extra_common ? extra_common[eo + lo] : 0u;
}

struct RescoreArgs {
    const double *data;            // f64 table of the graph, row-major [n][width]
    uint32_t n, width;
    const float *cand_d;           // select-phase distances [q_count][kprime] (FP32; squared for the Euclidean graph)
    const uint32_t *cand_i;        // select-phase candidates
    uint32_t kprime, k, q_begin, q_count;
    const uint64_t *extra_off;     // CSR of extra candidates by GLOBAL row (null: none)
    const uint32_t *extra_idx;
    const uint32_t *extra_common;  // shared sketch hashes per extra candidate
    SketchArgs sk;
    const uint8_t *row_flag;       // rows that always take the exhaustive pass (null: none)
    uint32_t ecap;                 // extra candidates evaluated per row at most
    uint32_t *idx_out;
    float *dist_out;
    uint32_t *n_valid;
    uint32_t *flag_list;           // out: local rows that were not certified (null: certificate off); bit 31 = the
                                   // row's FP32 values cannot be trusted at all (dense collision group, too many extras)
    float *flag_dk, *flag_eps;     // out, per flag_list entry: the k-th f64 distance among the candidates and eps
    uint32_t *counters;            // [0] length of flag_list, [1] cumulative flagged
};

// One warp per query row: f64 distances of the K' select candidates (+ the row's extra candidates), final order by
// (f32 distance, index), first k out — and the CERTIFICATE of the select phase:
//   every row that is not a candidate has a select distance >= thr = the K'-th select distance; the select error is
//   measured on the row's own candidates, eps = max |d_select - d_f64|.  If thr - d_k(f64) > max(4 eps, floor) no row
//   outside the list can precede the k-th neighbour unless its error is more than 4x the worst one seen among the K'
//   nearest (and above the floor); otherwise the row is appended to flag_list and re-evaluated by exact_rows_kernel.
//   This is an empirical certificate, not a proof: DESIGN.md 4.4 states what it assumes.
template <int VARIANT, int WHICH>
__global__ void __launch_bounds__(128)
rescore_kernel(const RescoreArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const uint32_t cap = a.kprime + a.ecap;
    float *sd = reinterpret_cast<float *>(smem_raw) + (size_t)warp * cap;
    uint32_t *si = reinterpret_cast<uint32_t *>(smem_raw + (size_t)wpb * cap * 4) + (size_t)warp * cap;
    const uint32_t q = blockIdx.x * wpb + warp;
    if (q >= a.q_count) return;
    const uint32_t grow = a.q_begin + q, kprime = a.kprime, k = a.k, n = a.n;
    const double *qrow = a.data + (uint64_t)grow * a.width;
    bool force = a.row_flag && a.row_flag[grow];
    uint32_t ne = 0, ne_all = 0;
    uint64_t eo = 0;
    if (a.extra_off) {
        eo = a.extra_off[grow];
        const uint64_t cnt = a.extra_off[grow + 1] - eo;
        if (cnt > a.ecap) force = true;
        ne_all = (uint32_t)cnt;
        ne = (uint32_t)(cnt > a.ecap ? a.ecap : cnt);
    }
    const uint32_t total = kprime + ne;
    for (uint32_t e = lane; e < kprime; e += 32) si[e] = a.cand_i[(uint64_t)q * kprime + e];
    __syncwarp();
    for (uint32_t e = lane; e < ne; e += 32) {            // extras are sorted and unique; drop those already selected
        uint32_t j = a.extra_idx[eo + e];
        for (uint32_t o = 0; o < kprime; ++o)
            if (si[o] == j) { j = 0xffffffffu; break; }
        si[kprime + e] = j;
    }
    __syncwarp();
    float eps = 0.f;
    for (uint32_t e = lane; e < total; e += 32) {
        const uint32_t j = si[e];
        float df = INFINITY;
        if (j < n) {
            float dm;                                            // what the select phase estimated: the metric without
            if (WHICH == RG_GRAPH_DEPTH) {                       // the MinHash factor
                const uint32_t common = !a.sk.hash ? 0u
                                        : e >= kprime ? a.extra_common[eo + (e - kprime)]
                                                      : common_of(a.extra_idx, a.extra_common, eo, ne_all, j, nullptr);
                df = depth_pair<VARIANT>(qrow, a.data + (uint64_t)j * a.width, a.width, a.sk, grow, j, common, &dm);
            } else {
                df = dm = pair_distance_f32<VARIANT, WHICH>(qrow, a.data + (uint64_t)j * a.width, a.width);
            }
            if (e < kprime) {
                float ds = a.cand_d[(uint64_t)q * kprime + e];
                if (WHICH == RG_GRAPH_TNF_EUCLID) ds = sqrtf(ds);
                // Structural under-estimates are left out: 1 - 2^-y underflows to exactly 0 for near-equal means
                // (d_select = 0), and a tiny negative equal-variance term is NaN -> 1.0 in f64 (:647,:659-661) where FP32
                // sees -0 or a positive rounding.  Both only cost a candidate slot; counting them would flag ~1.5 % of
                // all rows for nothing (profiles/r2b_flags_1m.json).
                const bool structural = ds == 0.f || (WHICH == RG_GRAPH_DEPTH && dm == 1.0f && ds < 1.0f);
                if (ds < 1e37f && dm < 1e37f && !structural) eps = fmaxf(eps, fabsf(ds - dm));
            }
        }
        sd[e] = df;
    }
    __syncwarp();
    uint32_t valid = 0;
    float dk = -INFINITY;
    for (uint32_t e = lane; e < total; e += 32) {
        const float d = sd[e];
        const uint32_t j = si[e];
        uint32_t rank = 0;
        for (uint32_t o = 0; o < total; ++o) {
            const float od = sd[o];
            const uint32_t oj = si[o];
            rank += (od < d) || (od == d && (oj < j || (oj == j && o < e)));
        }
        if (rank < k) {
            a.idx_out[(uint64_t)q * k + rank] = j;
            a.dist_out[(uint64_t)q * k + rank] = d;
            valid += j < n;
            if (rank == k - 1) dk = d;
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        valid += __shfl_xor_sync(0xffffffffu, valid, o);
        eps = fmaxf(eps, __shfl_xor_sync(0xffffffffu, eps, o));
        dk = fmaxf(dk, __shfl_xor_sync(0xffffffffu, dk, o));
    }
    if (lane == 0) {
        if (a.n_valid) a.n_valid[q] = valid;
        if (a.flag_list) {
            const bool full = a.cand_i[(uint64_t)q * kprime + kprime - 1] < n;   // else every row already is a candidate
            float thr = a.cand_d[(uint64_t)q * kprime + kprime - 1];
            if (WHICH == RG_GRAPH_TNF_EUCLID) thr = sqrtf(thr);
            // floor: the candidates are the rows with the SMALLEST select distances, so their errors are biased low
            // (measured: cluster rows lost a neighbour over-estimated by 2.5e-6 while no candidate showed more than
            // 1e-7); 2e-5 absolute + 1e-3 relative is ~8x the largest systematic FP32 error seen at d ~ 1e-6
            const float floor_ = WHICH == RG_GRAPH_DEPTH ? 2e-5f + 1e-3f * fabsf(thr) : 1e-6f + 1e-5f * fabsf(thr);
            const bool certified = !full || (thr - dk > fmaxf(4.f * eps, floor_));
            if (force || !certified) {
                const uint32_t slot = atomicAdd(&a.counters[0], 1u);
                a.flag_list[slot] = q | (force ? 0x80000000u : 0u);
                a.flag_dk[slot] = valid >= k ? dk : INFINITY;
                a.flag_eps[slot] = eps;
                atomicAdd(&a.counters[1], 1u);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Exhaustive pass for the rows the certificate rejected: one CTA per flagged row walks ALL n rows and keeps the k best
// by (f32 of the f64 distance, index) — the definition of the result.  8 warps take interleaved groups of 32 rows in
// increasing index (the arrival order topk_insert needs for ties) with one list each; the 8 disjoint lists are merged
// by rank counting.
// Coverage graph: the k-th f64 distance among the row's candidates (dk) bounds the true k-th distance from above, so
// a row whose FP32 distance — computed by the very code of the select kernels, depth_sample2 — exceeds
// dk + 8 eps cannot be a neighbour under the same error model the certificate uses, and its f64 evaluation (~20x the
// FP32 cost) is skipped; the row's extra candidates (FP32 collisions, knn_exact.cu) are always evaluated.  Rows
// flagged with bit 31 (their FP32 operands collide wholesale) and the TNF graphs evaluate every pair in f64.
constexpr int XW = 8;
constexpr int XPARTS = 16;     // a flagged row is split over up to 16 CTAs when few rows are flagged
constexpr uint32_t XCAP = 4096; // rows per (flagged row, part) that may pass the FP32 prefilter and still be evaluated side by side

struct ExactArgs {
    const double *data;
    uint32_t n, width, k, q_begin;
    const uint32_t *flag_list;
    const float *flag_dk, *flag_eps;
    uint32_t *counters;
    uint32_t *idx_out;
    float *dist_out;
    uint32_t *n_valid;
    const float4 *cov_t;           // FP32 operand tiles of the coverage graph (null: TNF graphs)
    const uint64_t *extra_off;
    const uint32_t *extra_idx;
    const uint32_t *extra_common;
    SketchArgs sk;
    int compact;                   // 1: collect the rows that pass the prefilter and evaluate them side by side
    float *part_d;                 // scratch [gridDim.x][K]: sorted partial lists of split rows
    uint32_t *part_i;
    uint32_t *part_done;           // [gridDim.x] arrival counters of split rows (zero at launch)
};

// rank of (d, i) among `lists` sorted K-entry lists (disjoint sets): binary search per list
template <int K>
__device__ __forceinline__ uint32_t rank_in_lists(const float *ld, const uint32_t *li, int lists, float d, uint32_t i,
                                                  uint32_t stop) {
    uint32_t rank = 0;
    for (int p = 0; p < lists && rank < stop; ++p) {
        const float *pd = ld + p * K;
        const uint32_t *pi = li + p * K;
        uint32_t lo = 0, hi = K;                  // first entry of list p not before (d, i)
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (pd[mid] < d || (pd[mid] == d && pi[mid] < i)) lo = mid + 1; else hi = mid;
        }
        rank += lo;
    }
    return rank;
}

template <int VARIANT, int WHICH, int K>
__global__ void __launch_bounds__(XW * 32)
exact_rows_kernel(const ExactArgs a) {
    extern __shared__ __align__(16) uint8_t xsm[];
    const uint32_t width = a.width, n = a.n, k = a.k, S = width / 2;
    double *qs = reinterpret_cast<double *>(xsm);                                       // [width]
    float *ld = reinterpret_cast<float *>(qs + width);                                  // [XPARTS][K] (XW used per row part)
    uint32_t *li = reinterpret_cast<uint32_t *>(ld + XPARTS * K);
    ulonglong2 *q4 = reinterpret_cast<ulonglong2 *>(li + XPARTS * K);                   // [S][2] (DEPTH): {x,x,y,y},{z,z,w,w}
    uint32_t *cl_j = reinterpret_cast<uint32_t *>(q4 + 2 * S);                          // [XCAP] (DEPTH): rows that passed the prefilter
    float *cl_d = reinterpret_cast<float *>(cl_j + XCAP);                               // [XCAP] their f64 distances (as f32)
    __shared__ uint32_t s_found, s_last, s_cl;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t count = a.counters[0];
    if (count == 0) return;
    const uint32_t parts = min((uint32_t)XPARTS, max(1u, gridDim.x / count));
    const uint32_t n_groups = (n + 31) / 32;
    for (uint32_t w = blockIdx.x; w < count * parts; w += gridDim.x) {
        const uint32_t f = w / parts, part = w % parts;
        const uint32_t entry = a.flag_list[f], q = entry & 0x7fffffffu, grow = a.q_begin + q;
        bool prefilter = WHICH == RG_GRAPH_DEPTH && !(entry & 0x80000000u);
        float bound = INFINITY;
        uint64_t eo = 0;
        uint32_t ne = 0;
        if (WHICH == RG_GRAPH_DEPTH && prefilter) {
            const float dk = a.flag_dk[f], eps = a.flag_eps[f];
            bound = dk + fmaxf(16.f * eps, 0.25f * fabsf(dk) + 1e-4f);
            if (!(bound < INFINITY)) prefilter = false;
        }
        if (WHICH == RG_GRAPH_DEPTH && a.extra_off) { eo = a.extra_off[grow]; ne = (uint32_t)(a.extra_off[grow + 1] - eo); }
        __syncthreads();
        for (uint32_t c = threadIdx.x; c < width; c += XW * 32) qs[c] = a.data[(uint64_t)grow * width + c];
        if (WHICH == RG_GRAPH_DEPTH && prefilter)
            for (uint32_t c = threadIdx.x; c < S; c += XW * 32) {
                const float4 v = a.cov_t[((uint64_t)(grow / RG_TILE) * S + c) * RG_TILE + grow % RG_TILE];
                q4[2 * c] = make_ulonglong2(mk(v.x, v.x).v, mk(v.y, v.y).v);
                q4[2 * c + 1] = make_ulonglong2(mk(v.z, v.z).v, mk(v.w, v.w).v);
            }
        __syncthreads();
        const uint32_t g_begin = (uint32_t)((uint64_t)n_groups * part / parts), g_end = (uint32_t)((uint64_t)n_groups * (part + 1) / parts);
        // Coverage graph with the FP32 prefilter: the rows that pass are rare and scattered, so evaluating them in place
        // would run the ~60k-instruction f64 metric on one lane of a warp at a time.  Pass 0 therefore only COLLECTS the
        // passing rows in shared memory; all 256 threads then evaluate them side by side and rank them by (distance,
        // index).  If more than XCAP rows pass in this part (a cluster that size), pass 1 falls back to evaluating in
        // place with the per-warp lists.
        bool compact = WHICH == RG_GRAPH_DEPTH && prefilter && a.compact;
        for (int pass = 0; pass < 2; ++pass) {
            if (threadIdx.x == 0) s_cl = 0;
            for (int e = lane; e < K; e += 32) { ld[warp * K + e] = INFINITY; li[warp * K + e] = 0xffffffffu; }
            __syncthreads();
            float thr = INFINITY;
            for (uint32_t g = g_begin + warp; g < g_end; g += XW) {
                const uint32_t j = g * 32 + lane;
                bool want = j < n && j != grow;
                if (WHICH == RG_GRAPH_DEPTH && prefilter) {
                    float acc0 = 0.f, cnt0 = 0.f, acc1 = 0.f, cnt1 = 0.f;
                    depthk::SignTrack unused;
                    const float4 *ref = a.cov_t + (uint64_t)g * S * RG_TILE + lane;
                    // two samples ahead in registers: the loop is a chain of ~150-cycle evaluations fed by global loads,
                    // and a CTA has only 8 warps to hide them behind
                    float4 b0 = ref[0], b1 = S > 1 ? ref[RG_TILE] : b0;
                    for (uint32_t smp = 0; smp < S; ++smp) {
                        const float4 b = b0;
                        b0 = b1;
                        if (smp + 2 < S) b1 = ref[(size_t)(smp + 2) * RG_TILE];
                        const ulonglong2 xy = q4[2 * smp], zw = q4[2 * smp + 1];
                        depthk::depth_sample2<VARIANT, false>(pk{xy.x}, pk{xy.y}, pk{zw.x}, pk{zw.y}, b, acc0, cnt0, acc1, cnt1, unused);
                    }
                    const float d32 = depthk::depth_finish<VARIANT>(acc0, cnt0);
                    if (want && d32 > bound) {                 // not a neighbour by the FP32 value: keep only extra candidates
                        bool joined;
                        common_of(a.extra_idx, nullptr, eo, ne, j, &joined);
                        want = joined;
                    }
                }
                if (compact) {
                    const unsigned m = __ballot_sync(0xffffffffu, want);
                    if (m) {
                        uint32_t base = 0;
                        if (lane == 0) base = atomicAdd(&s_cl, (uint32_t)__popc(m));
                        base = __shfl_sync(0xffffffffu, base, 0);
                        const uint32_t pos = base + __popc(m & ((1u << lane) - 1u));
                        if (want && pos < XCAP) cl_j[pos] = j;
                    }
                    continue;
                }
                float d = INFINITY;
                if (want) {
                    if (WHICH == RG_GRAPH_DEPTH) {
                        float dm;
                        const uint32_t common = a.sk.hash ? common_of(a.extra_idx, a.extra_common, eo, ne, j, nullptr) : 0u;
                        d = depth_pair<VARIANT>(qs, a.data + (uint64_t)j * width, width, a.sk, grow, j, common, &dm);
                    } else {
                        d = pair_distance_f32<VARIANT, WHICH>(qs, a.data + (uint64_t)j * width, width);
                    }
                }
                topk_offer<K>(ld + warp * K, li + warp * K, d, j, thr, lane);
            }
            __syncthreads();
            if (!compact) break;
            const uint32_t L = s_cl;
            if (L > XCAP) { compact = false; __syncthreads(); continue; }     // too many for the buffer: evaluate in place
            if (WHICH == RG_GRAPH_DEPTH)
                for (uint32_t e = threadIdx.x; e < L; e += XW * 32) {
                    const uint32_t j = cl_j[e];
                    float dm;
                    const uint32_t common = a.sk.hash ? common_of(a.extra_idx, a.extra_common, eo, ne, j, nullptr) : 0u;
                    cl_d[e] = depth_pair<VARIANT>(qs, a.data + (uint64_t)j * width, width, a.sk, grow, j, common, &dm);
                }
            __syncthreads();
            for (uint32_t e = threadIdx.x; e < L; e += XW * 32) {             // the K best by (distance, index) -> list 0, sorted
                const float d = cl_d[e];
                const uint32_t j = cl_j[e];
                uint32_t rank = 0;
                for (uint32_t o = 0; o < L && rank < (uint32_t)K; ++o) rank += cl_d[o] < d || (cl_d[o] == d && cl_j[o] < j);
                if (rank < (uint32_t)K) { ld[rank] = d; li[rank] = j; }
            }
            __syncthreads();
            break;
        }
        if (threadIdx.x == 0) { s_found = 0; s_last = parts == 1; }
        if (parts > 1)                                      // this part's sorted list goes to scratch slot w
            for (uint32_t c = threadIdx.x; c < K; c += XW * 32) { a.part_d[(size_t)w * K + c] = INFINITY; a.part_i[(size_t)w * K + c] = 0xffffffffu; }
        __syncthreads();
        uint32_t mine = 0;
        for (uint32_t c = threadIdx.x; c < XW * K; c += XW * 32) {
            const float d = ld[c];
            const uint32_t i = li[c];
            if (i == 0xffffffffu) continue;
            ++mine;
            const uint32_t rank = rank_in_lists<K>(ld, li, XW, d, i, parts == 1 ? k : K);
            if (parts == 1) {
                if (rank < k) { a.idx_out[(uint64_t)q * k + rank] = i; a.dist_out[(uint64_t)q * k + rank] = d; }
            } else if (rank < K) {
                a.part_d[(size_t)w * K + rank] = d;
                a.part_i[(size_t)w * K + rank] = i;
            }
        }
        if (parts > 1) {                                    // the last part of the row to arrive merges all of them
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) s_last = atomicAdd(&a.part_done[f], 1u) == parts - 1;
            __syncthreads();
            if (!s_last) continue;
            __threadfence();
            mine = 0;
            for (uint32_t c = threadIdx.x; c < parts * K; c += XW * 32) {
                ld[c] = __ldcg(a.part_d + (size_t)f * parts * K + c);
                li[c] = __ldcg(a.part_i + (size_t)f * parts * K + c);
            }
            __syncthreads();
            for (uint32_t c = threadIdx.x; c < parts * K; c += XW * 32) {
                const float d = ld[c];
                const uint32_t i = li[c];
                if (i == 0xffffffffu) continue;
                ++mine;
                const uint32_t rank = rank_in_lists<K>(ld, li, (int)parts, d, i, k);
                if (rank < k) { a.idx_out[(uint64_t)q * k + rank] = i; a.dist_out[(uint64_t)q * k + rank] = d; }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if (lane == 0 && mine) atomicAdd(&s_found, mine);
        __syncthreads();
        const uint32_t nv = min(s_found, k);
        for (uint32_t r = nv + threadIdx.x; r < k; r += XW * 32) {      // fewer than k rows exist: pad
            a.idx_out[(uint64_t)q * k + r] = 0xffffffffu;
            a.dist_out[(uint64_t)q * k + r] = INFINITY;
        }
        if (threadIdx.x == 0) {
            if (a.n_valid) a.n_valid[q] = nv;
            atomicAdd(&a.counters[2], 1u);
        }
    }
}

__global__ void reset_flag_count_kernel(uint32_t *counters) { counters[0] = 0; }

template <int VARIANT, int WHICH>
int32_t launch_exact(rg_handle *h, const double *data, uint32_t width, uint32_t k, uint64_t q_begin, uint64_t q_count,
                     uint32_t *idx, float *dist, uint32_t *nv) {
    const int K = k <= 32 ? 32 : k <= 64 ? 64 : k <= 128 ? 128 : 256;
    const bool depth = WHICH == RG_GRAPH_DEPTH;
    const size_t smem = (size_t)width * 8 + (size_t)XPARTS * K * 8 + (depth ? (size_t)(width / 2) * 32 + (size_t)XCAP * 8 : 0);
    const unsigned grid = (unsigned)h->sm_count * 2;
    // scratch of split rows: [grid][K] distances | [grid][K] indices | [grid] arrival counters
    RG_TRY(rg_reserve(h, h->exact_scratch, (size_t)grid * K * 8 + (size_t)grid * 4));
    RG_CUDA(h, cudaMemsetAsync((uint8_t *)h->exact_scratch.p + (size_t)grid * K * 8, 0, (size_t)grid * 4, h->stream));
    ExactArgs a{};
    a.compact = h->opt_exact_compact;
    a.part_d = (float *)h->exact_scratch.p;
    a.part_i = (uint32_t *)h->exact_scratch.p + (size_t)grid * K;
    a.part_done = (uint32_t *)h->exact_scratch.p + (size_t)grid * K * 2;
    a.data = data; a.n = (uint32_t)h->n; a.width = width; a.k = k; a.q_begin = (uint32_t)q_begin;
    a.flag_list = (const uint32_t *)h->flag_list.p;
    a.flag_dk = (const float *)h->flag_list.p + q_count;
    a.flag_eps = (const float *)h->flag_list.p + 2 * q_count;
    a.counters = (uint32_t *)h->cert_ctr.p;
    a.idx_out = idx; a.dist_out = dist; a.n_valid = nv;
    if (depth) {
        a.cov_t = (const float4 *)h->cov_t.p;
        if (h->has_extra) {
            a.extra_off = (const uint64_t *)h->extra_off.p;
            a.extra_idx = (const uint32_t *)h->extra_idx.p;
            a.extra_common = (const uint32_t *)h->extra_common.p;
        }
        if (h->has_sketch) a.sk = SketchArgs{(const uint64_t *)h->sk_hash.p, (const uint64_t *)h->sk_off.p, h->sk_max_hash};
    }
    auto go = [&](auto kern) -> int32_t {
        RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, XW * 32, smem, h->stream>>>(a);
        RG_CUDA(h, cudaGetLastError());
        h->tm.kernel_launches += 1;
        return RG_OK;
    };
    switch (K) {
        case 32: return go(exact_rows_kernel<VARIANT, WHICH, 32>);
        case 64: return go(exact_rows_kernel<VARIANT, WHICH, 64>);
        case 128: return go(exact_rows_kernel<VARIANT, WHICH, 128>);
        default: return go(exact_rows_kernel<VARIANT, WHICH, 256>);
    }
}

// f64 re-evaluation + certificate + exhaustive pass of the uncertified rows; everything asynchronous on h->stream
template <int VARIANT, int WHICH>
int32_t launch_rescore(rg_handle *h, const double *data, uint32_t width, uint32_t kprime, uint32_t k,
                       uint64_t q_begin, uint64_t q_count, uint32_t *idx, float *dist, uint32_t *nv) {
    const int wpb = 4;
    const bool certify = h->opt_certify != 0;
    const bool extras = WHICH == RG_GRAPH_DEPTH && h->has_extra;
    RescoreArgs a{};
    a.data = data; a.n = (uint32_t)h->n; a.width = width;
    a.cand_d = (const float *)h->cand_d.p; a.cand_i = (const uint32_t *)h->cand_i.p;
    a.kprime = kprime; a.k = k; a.q_begin = (uint32_t)q_begin; a.q_count = (uint32_t)q_count;
    a.ecap = extras ? 64u : 0u;
    if (extras) {
        a.extra_off = (const uint64_t *)h->extra_off.p;
        a.extra_idx = (const uint32_t *)h->extra_idx.p;
        a.extra_common = (const uint32_t *)h->extra_common.p;
        a.row_flag = (const uint8_t *)h->row_flag.p;
        if (h->has_sketch) a.sk = SketchArgs{(const uint64_t *)h->sk_hash.p, (const uint64_t *)h->sk_off.p, h->sk_max_hash};
    }
    a.idx_out = idx; a.dist_out = dist; a.n_valid = nv;
    if (certify) {
        RG_TRY(rg_reserve(h, h->flag_list, 3 * q_count * sizeof(uint32_t)));   // rows | k-th f64 distance | eps
        h->flag_rows = q_count;
        if (!h->cert_ctr.p) {
            RG_TRY(rg_reserve(h, h->cert_ctr, 8 * sizeof(uint32_t)));
            RG_CUDA(h, cudaMemsetAsync(h->cert_ctr.p, 0, 8 * sizeof(uint32_t), h->stream));
        }
        reset_flag_count_kernel<<<1, 1, 0, h->stream>>>((uint32_t *)h->cert_ctr.p);
        a.flag_list = (uint32_t *)h->flag_list.p;
        a.flag_dk = (float *)h->flag_list.p + q_count;
        a.flag_eps = (float *)h->flag_list.p + 2 * q_count;
        a.counters = (uint32_t *)h->cert_ctr.p;
    }
    const size_t smem = (size_t)wpb * (kprime + a.ecap) * 8;
    rescore_kernel<VARIANT, WHICH><<<(unsigned)((q_count + wpb - 1) / wpb), wpb * 32, smem, h->stream>>>(a);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1 + (certify ? 1 : 0);
    if (certify) RG_TRY((launch_exact<VARIANT, WHICH>(h, data, width, k, q_begin, q_count, idx, dist, nv)));
    return RG_OK;
}

}  // namespace

int32_t rg_launch_depth_select(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count);
int32_t rg_launch_depth_build_sym(rg_handle *h, uint32_t kprime, float *cd, uint32_t *ci, uint32_t part, uint32_t n_parts);
int32_t rg_launch_depth_merge(rg_handle *h, uint32_t kprime, uint32_t n_parts, const float *pd, const uint32_t *pi, uint64_t q_count);
bool rg_depth_build_sym_wanted(const rg_handle *h);
int32_t rg_launch_tnf_select(rg_handle *h, uint32_t which, uint32_t kprime, uint64_t q_begin, uint64_t q_count);
uint32_t rg_tnf_kprime(const rg_handle *h, uint32_t k, uint32_t general);

// K' = candidates kept by the FP32 select phase: k plus a margin that absorbs FP32 reordering
static uint32_t pick_kprime(uint32_t k) {
    if (k <= 48) return 64;
    if (k <= 104) return 128;
    if (k <= 224) return 256;
    return 0;
}

extern "C" int32_t rg_knn_search_dev(rg_handle *h, uint32_t which, uint32_t k, uint64_t q_begin, uint64_t q_end,
                                     uint32_t *idx_dev, float *dist_dev, uint32_t *n_valid_dev) {
    if (!h) return RG_ERR_INVALID;
    if (h->n == 0) return rg_fail(h, RG_ERR_STATE, "rg_knn_search: no tables loaded (call rg_knn_load first)");
    if (which > RG_GRAPH_TNF_EUCLID) return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: unknown graph %u", which);
    if (which != RG_GRAPH_DEPTH && !h->has_tnf) return rg_fail(h, RG_ERR_STATE, "rg_knn_search: no TNF table loaded");
    if (which == RG_GRAPH_DEPTH && !h->has_cov) return rg_fail(h, RG_ERR_STATE, "rg_knn_search: no coverage table loaded");
    if (!idx_dev || !dist_dev) return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: output pointer is NULL");
    if (q_begin > q_end || q_end > h->n)
        return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: query range [%llu,%llu) outside [0,%llu)",
                       (unsigned long long)q_begin, (unsigned long long)q_end, (unsigned long long)h->n);
    const uint32_t kprime = rg_knn_kprime(h, which, k);
    if (k == 0 || kprime == 0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_search: k=%u must be in [1,224]", k);
    const uint64_t q_count = q_end - q_begin;
    if (q_count == 0) return RG_OK;
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_TRY(rg_reserve(h, h->cand_d, q_count * kprime * sizeof(float)));
    RG_TRY(rg_reserve(h, h->cand_i, q_count * kprime * sizeof(uint32_t)));
    RG_CUDA(h, cudaEventRecord(h->ev[3], h->stream));
    if (which == RG_GRAPH_DEPTH && q_count == h->n && rg_depth_build_sym_wanted(h)) RG_TRY(rg_launch_depth_build_sym(h, kprime, (float *)h->cand_d.p, (uint32_t *)h->cand_i.p, 0, 1));
    else if (which == RG_GRAPH_DEPTH) RG_TRY(rg_launch_depth_select(h, kprime, q_begin, q_count));
    else RG_TRY(rg_launch_tnf_select(h, which, kprime, q_begin, q_count));
    RG_CUDA(h, cudaEventRecord(h->ev[4], h->stream));
    const bool fl = h->variant == RG_VARIANT_FLIGHT;
    if (which == RG_GRAPH_DEPTH) {
        const double *data = (const double *)h->cov64.p;
        RG_TRY(fl ? (launch_rescore<RG_VARIANT_FLIGHT, RG_GRAPH_DEPTH>(h, data, 2 * h->s, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev))
                  : (launch_rescore<RG_VARIANT_RUST, RG_GRAPH_DEPTH>(h, data, 2 * h->s, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev)));
    } else if (which == RG_GRAPH_TNF_EUCLID) {
        RG_TRY((launch_rescore<RG_VARIANT_FLIGHT, RG_GRAPH_TNF_EUCLID>(h, (const double *)h->tnf64.p, h->d, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev)));
    } else {
        const double *data = (const double *)h->tnf64.p;
        RG_TRY(fl ? (launch_rescore<RG_VARIANT_FLIGHT, RG_GRAPH_TNF>(h, data, h->d, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev))
                  : (launch_rescore<RG_VARIANT_RUST, RG_GRAPH_TNF>(h, data, h->d, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev)));
    }
    RG_CUDA(h, cudaEventRecord(h->ev[5], h->stream));
    h->tm.select_ms = -1.0;   // resolved lazily by rg_get_timings (the launches above are asynchronous)
    return RG_OK;
}

extern "C" uint32_t rg_knn_kprime(rg_handle *h, uint32_t which, uint32_t k) {
    if (!h || k == 0) return 0;
    uint32_t kprime = pick_kprime(k);
    if (kprime && which != RG_GRAPH_DEPTH && h->n) kprime = rg_tnf_kprime(h, k, kprime);
    // test hook (rg_set_option "kprime_depth" / "kprime_tnf"): a shorter or longer list, never shorter than k
    const uint32_t ov = which == RG_GRAPH_DEPTH ? h->opt_kprime_depth : h->opt_kprime_tnf;
    if (kprime && ov >= k) kprime = ov;
    return kprime;
}

extern "C" int32_t rg_knn_depth_partial_dev(rg_handle *h, uint32_t k, uint32_t part, uint32_t n_parts,
                                            float *cand_dist_dev, uint32_t *cand_idx_dev) {
    if (!h) return RG_ERR_INVALID;
    if (h->n == 0 || !h->has_cov) return rg_fail(h, RG_ERR_STATE, "rg_knn_depth_partial: no coverage table loaded");
    if (!cand_dist_dev || !cand_idx_dev) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_partial: output pointer is NULL");
    if (n_parts == 0 || part >= n_parts) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_partial: part %u of %u", part, n_parts);
    const uint32_t kprime = rg_knn_kprime(h, RG_GRAPH_DEPTH, k);
    if (k == 0 || kprime == 0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_partial: k=%u must be in [1,224]", k);
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_CUDA(h, cudaEventRecord(h->ev[3], h->stream));
    RG_TRY(rg_launch_depth_build_sym(h, kprime, cand_dist_dev, cand_idx_dev, part, n_parts));
    RG_CUDA(h, cudaEventRecord(h->ev[4], h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[5], h->stream));
    h->tm.select_ms = -1.0;
    return RG_OK;
}

extern "C" int32_t rg_knn_depth_merge_dev(rg_handle *h, uint32_t k, uint32_t n_parts, const float *parts_dist_dev,
                                          const uint32_t *parts_idx_dev, uint64_t q_begin, uint64_t q_end,
                                          uint32_t *idx_dev, float *dist_dev, uint32_t *n_valid_dev) {
    if (!h) return RG_ERR_INVALID;
    if (h->n == 0 || !h->has_cov) return rg_fail(h, RG_ERR_STATE, "rg_knn_depth_merge: no coverage table loaded");
    if (!parts_dist_dev || !parts_idx_dev || !idx_dev || !dist_dev) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_merge: NULL pointer");
    if (q_begin > q_end || q_end > h->n) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_merge: query range outside [0,n)");
    if (n_parts == 0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_merge: n_parts = 0");
    const uint32_t kprime = rg_knn_kprime(h, RG_GRAPH_DEPTH, k);
    if (k == 0 || kprime == 0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_merge: k=%u must be in [1,224]", k);
    const uint64_t q_count = q_end - q_begin;
    if (q_count == 0) return RG_OK;
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_TRY(rg_reserve(h, h->cand_d, q_count * kprime * sizeof(float)));
    RG_TRY(rg_reserve(h, h->cand_i, q_count * kprime * sizeof(uint32_t)));
    RG_CUDA(h, cudaEventRecord(h->ev[3], h->stream));
    RG_TRY(rg_launch_depth_merge(h, kprime, n_parts, parts_dist_dev, parts_idx_dev, q_count));
    RG_CUDA(h, cudaEventRecord(h->ev[4], h->stream));
    const double *data = (const double *)h->cov64.p;
    RG_TRY(h->variant == RG_VARIANT_FLIGHT
               ? (launch_rescore<RG_VARIANT_FLIGHT, RG_GRAPH_DEPTH>(h, data, 2 * h->s, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev))
               : (launch_rescore<RG_VARIANT_RUST, RG_GRAPH_DEPTH>(h, data, 2 * h->s, kprime, k, q_begin, q_count, idx_dev, dist_dev, n_valid_dev)));
    RG_CUDA(h, cudaEventRecord(h->ev[5], h->stream));
    h->tm.select_ms = -1.0;
    return RG_OK;
}

extern "C" int32_t rg_knn_flagged_rows(rg_handle *h, uint32_t *rows, float *kth_dist, float *select_eps, uint64_t cap,
                                       uint64_t *count) {
    if (!h) return RG_ERR_INVALID;
    if (count) *count = 0;
    if (!h->cert_ctr.p || !h->flag_list.p || h->flag_rows == 0) return RG_OK;
    RG_CUDA(h, cudaSetDevice(h->device));
    uint32_t c = 0;
    RG_CUDA(h, cudaMemcpyAsync(&c, h->cert_ctr.p, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    if (count) *count = c;
    const uint64_t m = c < cap ? c : cap;
    const uint32_t *base = (const uint32_t *)h->flag_list.p;
    if (m && rows) RG_CUDA(h, cudaMemcpy(rows, base, m * 4, cudaMemcpyDeviceToHost));
    if (m && kth_dist) RG_CUDA(h, cudaMemcpy(kth_dist, base + h->flag_rows, m * 4, cudaMemcpyDeviceToHost));
    if (m && select_eps) RG_CUDA(h, cudaMemcpy(select_eps, base + 2 * h->flag_rows, m * 4, cudaMemcpyDeviceToHost));
    return RG_OK;
}
