// knn_rescore.cu — K2c: re-evaluate the K' selected candidates of every query row with the
// reference's own f64 arithmetic, truncate to f32 as embedder.rs:125,:170 do, order by
// (f32 distance, index) and emit the first k.  Compiled with -fmad=false so that no
// multiply-add is contracted: the operation order below is the reference's, statement by
// statement (src/kmers/kmer_counting.rs:366-414, src/coverage/coverage_calculator.rs:588-666,
// flight/flight/metrics.py:445-558,:601-645).
#include "rg_internal.cuh"

namespace {

__device__ double kmer_correlation_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
    double sx = 0.0, sy = 0.0;
    for (uint32_t j = 0; j < d; ++j) sx += x[j];
    for (uint32_t j = 0; j < d; ++j) sy += y[j];
    const double mu_x = sx / (double)d, mu_y = sy / (double)d;          // :367-368
    double norm_x = 0.0, norm_y = 0.0, dot = 0.0;
    for (uint32_t j = 0; j < d; ++j) {                                    // :375-381
        const double a = x[j] - mu_x, b = y[j] - mu_y;
        dot += a * b;
        norm_x += a * a;
        norm_y += b * b;
    }
    if (norm_x == 0.0 && norm_y == 0.0) return 0.0;                       // :383
    if (dot == 0.0) return 1.0;                                           // :385
    norm_x = norm_x / ((double)d - 1.0);                                  // :401-403
    norm_y = norm_y / ((double)d - 1.0);
    dot = dot / ((double)d - 1.0);
    const double vlr = -2.0 * dot + norm_x + norm_y;                      // :404
    double rho = 1.0 - vlr / (norm_x + norm_y);                           // :405-408
    rho += 1.0;
    rho = 2.0 - rho;
    rho /= 2.0;
    if (isnan(rho)) return 1.0;                                           // :410
    return rho;
}

__device__ double rho_flight_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
    double mu_x = 0.0, mu_y = 0.0;
    for (uint32_t j = 0; j < d; ++j) { mu_x += x[j]; mu_y += y[j]; }
    mu_x /= (double)d;
    mu_y /= (double)d;
    double norm_x = 0.0, norm_y = 0.0, dot = 0.0;
    for (uint32_t j = 0; j < d; ++j) {
        const double a = x[j] - mu_x, b = y[j] - mu_y;
        norm_x += a * a;
        norm_y += b * b;
        dot += a * b;
    }
    norm_x /= (double)(d - 1);
    norm_y /= (double)(d - 1);
    dot /= (double)(d - 1);
    const double vlr = -2.0 * dot + norm_x + norm_y;
    double rho = 1.0 - vlr / (norm_x + norm_y);
    rho += 1.0;
    return 2.0 - rho;
}

// flight/flight/metrics.py:64-77 (tnf_euclidean)
__device__ double tnf_euclidean_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
    double r = 0.0;
    for (uint32_t j = 0; j < d; ++j) r += (x[j] - y[j]) * (x[j] - y[j]);
    return sqrt(r);
}

// statrs Normal::cdf (call sites coverage_calculator.rs:647,:650)
__device__ __forceinline__ double cdf_statrs(double mean, double sd, double x) {
    return 0.5 * erfc((mean - x) / (sd * 1.4142135623730951));
}
// flight/metrics.py:543-558
__device__ __forceinline__ double cdf_flight(double mean, double sd, double x) {
    const double y = 0.5 * (1.0 + erf((x - mean) / (sd * sqrt(2.0))));
    return y > 1.0 ? 1.0 : y;
}

template <int VARIANT>
__device__ double metabat_f64(const double *__restrict__ a, const double *__restrict__ b, uint32_t s) {
    const double EPS = VARIANT == RG_VARIANT_FLIGHT ? 1e-6 : 0.0000001;
    double lnsum = 0.0;
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < s; ++i) {
        const double a_mean = a[2 * i] + EPS, b_mean = b[2 * i] + EPS;     // :600-603
        double a_var = a[2 * i + 1] + EPS, b_var = b[2 * i + 1] + EPS;
        if (a_var < 1.0) a_var = 1.0;                                      // :604-609
        if (b_var < 1.0) b_var = 1.0;
        if ((a_mean > EPS || b_mean > EPS) && a_mean != b_mean) {          // :619
            double k1, k2;
            if (fabs(a_var - b_var) < 1e-4) {                              // :620
                k1 = k2 = (a_mean + b_mean) / 2.0;
            } else {                                                       // :625-627
                const double tmp = sqrt(a_var * b_var * ((a_mean - b_mean) * (a_mean - b_mean) -
                                        2.0 * (a_var - b_var) * log(sqrt(b_var / a_var))));
                k1 = (tmp - a_mean * b_var + b_mean * a_var) / (a_var - b_var);
                k2 = (tmp + a_mean * b_var - b_mean * a_var) / (b_var - a_var);
            }
            if (k1 > k2) { const double t = k1; k1 = k2; k2 = t; }         // :630
            double m1, s1, m2, s2;
            if (a_var > b_var) { m1 = b_mean; s1 = sqrt(b_var); m2 = a_mean; s2 = sqrt(a_var); }  // :638-644
            else               { m1 = a_mean; s1 = sqrt(a_var); m2 = b_mean; s2 = sqrt(b_var); }
            double dd;
            if (VARIANT == RG_VARIANT_RUST) {
                if (fabs(k1 - k2) < EPS) dd = cdf_statrs(m1, s1, k1) - cdf_statrs(m2, s2, k1);     // :646-648
                else dd = fabs(cdf_statrs(m1, s1, k2) - cdf_statrs(m1, s1, k1) +
                               cdf_statrs(m2, s2, k1) - cdf_statrs(m2, s2, k2));                   // :650
            } else {
                if (k1 == k2) dd = fabs(cdf_flight(m1, s1, k1) - cdf_flight(m2, s2, k1));          // metrics.py:516-519
                else dd = fabs(cdf_flight(m1, s1, k2) - cdf_flight(m1, s1, k1) +
                               cdf_flight(m2, s2, k1) - cdf_flight(m2, s2, k2));
                dd = fmin(fmax(dd, 1e-6), 1.0 - 1e-6);
            }
            lnsum += log(dd);
            cnt += 1;
        }
    }
    if (cnt > 0) {
        double d = exp(lnsum / (double)cnt);                               // :659
        if (VARIANT == RG_VARIANT_RUST && isnan(d)) d = 1.0;               // :660
        return d;
    }
    return 1.0;                                                            // :665
}

// one warp per query row
template <int VARIANT, int WHICH>
__global__ void __launch_bounds__(128)
rescore_kernel(const double *__restrict__ data, uint32_t n, uint32_t width, const uint32_t *__restrict__ cand_i,
               uint32_t kprime, uint32_t k, uint32_t q_begin, uint32_t q_count, uint32_t *__restrict__ idx_out,
               float *__restrict__ dist_out, uint32_t *__restrict__ n_valid) {
    extern __shared__ uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float *sd = reinterpret_cast<float *>(smem_raw) + (size_t)warp * kprime;
    uint32_t *si = reinterpret_cast<uint32_t *>(smem_raw + (size_t)wpb * kprime * 4) + (size_t)warp * kprime;
    const uint32_t q = blockIdx.x * wpb + warp;
    if (q >= q_count) return;
    const double *qrow = data + (uint64_t)(q_begin + q) * width;
    for (uint32_t e = lane; e < kprime; e += 32) {
        const uint32_t j = cand_i[(uint64_t)q * kprime + e];
        float df = INFINITY;
        if (j < n) {
            const double *rrow = data + (uint64_t)j * width;
            double v;
            if (WHICH == RG_GRAPH_TNF_EUCLID) {
                v = tnf_euclidean_f64(qrow, rrow, width);
                df = isnan(v) ? 3.402823466e+38f : (float)v;
            } else if (WHICH == RG_GRAPH_TNF) {
                v = VARIANT == RG_VARIANT_FLIGHT ? rho_flight_f64(qrow, rrow, width) : kmer_correlation_f64(qrow, rrow, width);
                // embedder.rs:122-125; FLIGHT: NaN is defined as the metric's maximum
                df = isnan(v) ? (VARIANT == RG_VARIANT_RUST ? 0.0f : 2.0f) : (float)v;
            } else {
                v = metabat_f64<VARIANT>(qrow, rrow, width / 2);
                df = isnan(v) ? 1.0f : (float)v;                           // embedder.rs:165-170, MinHash factor 1
            }
        }
        sd[e] = df;
        si[e] = j;
    }
    __syncwarp();
    uint32_t valid = 0;
    for (uint32_t e = lane; e < kprime; e += 32) {
        const float d = sd[e];
        const uint32_t j = si[e];
        uint32_t rank = 0;
        for (uint32_t o = 0; o < kprime; ++o) {
            const float od = sd[o];
            const uint32_t oj = si[o];
            rank += (od < d) || (od == d && (oj < j || (oj == j && o < e)));
        }
        if (rank < k) {
            idx_out[(uint64_t)q * k + rank] = j;
            dist_out[(uint64_t)q * k + rank] = d;
            valid += j < n;
        }
    }
    if (n_valid) {
#pragma unroll
        for (int o = 16; o; o >>= 1) valid += __shfl_xor_sync(0xffffffffu, valid, o);
        if (lane == 0) n_valid[q] = valid;
    }
}

template <int VARIANT, int WHICH>
int32_t launch_rescore(rg_handle *h, const double *data, uint32_t width, uint32_t kprime, uint32_t k,
                       uint64_t q_begin, uint64_t q_count, uint32_t *idx, float *dist, uint32_t *nv) {
    const int wpb = 4;
    const size_t smem = (size_t)wpb * kprime * 8;
    rescore_kernel<VARIANT, WHICH><<<(unsigned)((q_count + wpb - 1) / wpb), wpb * 32, smem, h->stream>>>(
        data, (uint32_t)h->n, width, (const uint32_t *)h->cand_i.p, kprime, k, (uint32_t)q_begin, (uint32_t)q_count,
        idx, dist, nv);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
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
    uint32_t kprime = pick_kprime(k);
    if (kprime && which != RG_GRAPH_DEPTH) kprime = rg_tnf_kprime(h, k, kprime);
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
    return kprime;
}

extern "C" int32_t rg_knn_depth_partial_dev(rg_handle *h, uint32_t k, uint32_t part, uint32_t n_parts,
                                            float *cand_dist_dev, uint32_t *cand_idx_dev) {
    if (!h) return RG_ERR_INVALID;
    if (h->n == 0 || !h->has_cov) return rg_fail(h, RG_ERR_STATE, "rg_knn_depth_partial: no coverage table loaded");
    if (!cand_dist_dev || !cand_idx_dev) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_partial: output pointer is NULL");
    if (n_parts == 0 || part >= n_parts) return rg_fail(h, RG_ERR_INVALID, "rg_knn_depth_partial: part %u of %u", part, n_parts);
    const uint32_t kprime = pick_kprime(k);
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
    const uint32_t kprime = pick_kprime(k);
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
