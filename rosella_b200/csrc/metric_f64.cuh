// metric_f64.cuh — the reference's metrics in f64, statement by statement (src/kmers/kmer_counting.rs:366-414,
// src/coverage/coverage_calculator.rs:588-666, flight/flight/metrics.py:64-77,:445-558,:601-645).  Every translation unit
// that includes this file is compiled with -fmad=false (rosella_b200/build.py) so that no multiply-add is contracted:
// the operation order below is the reference's.
#pragma once
#include "rg_internal.cuh"

namespace metric64 {

static __device__ double kmer_correlation_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
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

static __device__ double rho_flight_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
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
static __device__ double tnf_euclidean_f64(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
    double r = 0.0;
    for (uint32_t j = 0; j < d; ++j) r += (x[j] - y[j]) * (x[j] - y[j]);
    return sqrt(r);
}

// statrs Normal::cdf (call sites coverage_calculator.rs:647,:650)
static __device__ __forceinline__ double cdf_statrs(double mean, double sd, double x) {
    return 0.5 * erfc((mean - x) / (sd * 1.4142135623730951));
}
// flight/metrics.py:543-558
static __device__ __forceinline__ double cdf_flight(double mean, double sd, double x) {
    const double y = 0.5 * (1.0 + erf((x - mean) / (sd * sqrt(2.0))));
    return y > 1.0 ? 1.0 : y;
}

template <int VARIANT>
static __device__ double metabat_f64(const double *__restrict__ a, const double *__restrict__ b, uint32_t s) {
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

}  // namespace metric64
