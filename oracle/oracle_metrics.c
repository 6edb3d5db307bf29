/*
 * oracle_metrics.c — CPU f64 restatement of Rosella's distance metrics and an exhaustive
 * exact top-k.  TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Third-party arithmetic restated (crate sources are not under /root/reference, no Cargo.lock):
 *   statrs ^0.16  Normal::cdf(x) = 0.5 * erfc((mean - x) / (std_dev * sqrt(2)))
 *                 (call sites coverage_calculator.rs:639-650)
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ndarray ^0.15 ArrayBase::sum on a contiguous row = numeric_util::unrolled_fold: eight partial sums over
 * strides of 8, combined as acc += (p0+p4); acc += (p1+p5); acc += (p2+p6); acc += (p3+p7); then the (< 8)
 * remaining elements in order.  The crate is not under /root/reference: restated from the
 * published source; the call site is kmer_counting.rs:345 (row.sum()). */
double orc_ndarray_sum(const double *xs, uint32_t len) {
    double p[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc = 0.0;
    uint32_t i = 0;
    for (; i + 8 <= len; i += 8)
        for (int l = 0; l < 8; ++l) p[l] = p[l] + xs[i + l];
    acc = acc + (p[0] + p[4]);
    acc = acc + (p[1] + p[5]);
    acc = acc + (p[2] + p[6]);
    acc = acc + (p[3] + p[7]);
    for (; i < len; ++i) acc = acc + xs[i];
    return acc;
}

/* kmer_counting.rs:340-359 */
void orc_clr(const double *in, double *out, uint64_t n, uint32_t d) {
    for (uint64_t r = 0; r < n; ++r) {
        const double *row = in + r * d;
        double sum = orc_ndarray_sum(row, d);
        double mean = sum / (double)d;
        for (uint32_t j = 0; j < d; ++j) out[r * d + j] = log(row[j] / mean);
    }
}

/* flight/rosella/binning.py:192: skbio.stats.composition.clr(x + 1) = ln(x+1) - mean(ln(x+1)) */
void orc_clr_flight(const double *in, double *out, uint64_t n, uint32_t d) {
    for (uint64_t r = 0; r < n; ++r) {
        const double *row = in + r * d;
        double s = 0.0;
        for (uint32_t j = 0; j < d; ++j) { out[r * d + j] = log(row[j] + 1.0); s += out[r * d + j]; }
        double m = s / (double)d;
        for (uint32_t j = 0; j < d; ++j) out[r * d + j] -= m;
    }
}

/* kmer_counting.rs:366-414 */
double orc_kmer_correlation(const double *x, const double *y, uint32_t d) {
    double sx = 0.0, sy = 0.0;
    for (uint32_t j = 0; j < d; ++j) sx += x[j];
    for (uint32_t j = 0; j < d; ++j) sy += y[j];
    double mu_x = sx / (double)d, mu_y = sy / (double)d;
    double norm_x = 0.0, norm_y = 0.0, dot = 0.0;
    for (uint32_t j = 0; j < d; ++j) {
        double a = x[j] - mu_x, b = y[j] - mu_y;
        dot += a * b;
        norm_x += a * a;
        norm_y += b * b;
    }
    if (norm_x == 0.0 && norm_y == 0.0) return 0.0;   /* :383 */
    if (dot == 0.0) return 1.0;                        /* :385 */
    norm_x = norm_x / ((double)d - 1.0);
    norm_y = norm_y / ((double)d - 1.0);
    dot = dot / ((double)d - 1.0);
    double vlr = -2.0 * dot + norm_x + norm_y;
    double rho = 1.0 - vlr / (norm_x + norm_y);
    rho += 1.0;
    rho = 2.0 - rho;
    rho /= 2.0;
    if (isnan(rho)) return 1.0;                        /* :410 */
    return rho;
}

/* flight/metrics.py:601-645 */
double orc_rho_flight(const double *x, const double *y, uint32_t d) {
    double mu_x = 0.0, mu_y = 0.0;
    for (uint32_t j = 0; j < d; ++j) { mu_x += x[j]; mu_y += y[j]; }
    mu_x /= (double)d; mu_y /= (double)d;
    double norm_x = 0.0, norm_y = 0.0, dot = 0.0;
    for (uint32_t j = 0; j < d; ++j) {
        double a = x[j] - mu_x, b = y[j] - mu_y;
        norm_x += a * a; norm_y += b * b; dot += a * b;
    }
    norm_x /= (double)(d - 1); norm_y /= (double)(d - 1); dot /= (double)(d - 1);
    double vlr = -2.0 * dot + norm_x + norm_y;
    double rho = 1.0 - vlr / (norm_x + norm_y);
    rho += 1.0;
    return 2.0 - rho;
}

/* flight/metrics.py:64-77 */
double orc_tnf_euclidean(const double *x, const double *y, uint32_t d) {
    double r = 0.0;
    for (uint32_t j = 0; j < d; ++j) r += (x[j] - y[j]) * (x[j] - y[j]);
    return sqrt(r);
}

static double normal_cdf_statrs(double mean, double sd, double x) {
    return 0.5 * erfc((mean - x) / (sd * M_SQRT2));
}

/* flight/metrics.py:543-558 */
static double normal_cdf_flight(double mean, double sd, double x) {
    double y = 0.5 * (1.0 + erf((x - mean) / (sd * sqrt(2.0))));
    return y > 1.0 ? 1.0 : y;
}

/* coverage_calculator.rs:588-666 */
static double metabat_rust(const double *a, const double *b, uint32_t s) {
    const double EPSILON = 0.0000001, MIN_VAR_EPSILON = 1e-4, MIN_VAR = 1.0;
    double lnsum = 0.0;
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < s; ++i) {
        double a_mean = a[2 * i] + EPSILON, b_mean = b[2 * i] + EPSILON;
        double a_var = a[2 * i + 1] + EPSILON, b_var = b[2 * i + 1] + EPSILON;
        if (a_var < MIN_VAR) a_var = MIN_VAR;
        if (b_var < MIN_VAR) b_var = MIN_VAR;
        if ((a_mean > EPSILON || b_mean > EPSILON) && a_mean != b_mean) {
            double k1, k2;
            if (fabs(a_var - b_var) < MIN_VAR_EPSILON) {
                k1 = k2 = (a_mean + b_mean) / 2.0;
            } else {
                double tmp = sqrt(a_var * b_var * ((a_mean - b_mean) * (a_mean - b_mean) -
                                  2.0 * (a_var - b_var) * log(sqrt(b_var / a_var))));
                k1 = (tmp - a_mean * b_var + b_mean * a_var) / (a_var - b_var);
                k2 = (tmp + a_mean * b_var - b_mean * a_var) / (b_var - a_var);
            }
            if (k1 > k2) { double t = k1; k1 = k2; k2 = t; }
            double m1, s1, m2, s2;
            if (a_var > b_var) { m1 = b_mean; s1 = sqrt(b_var); m2 = a_mean; s2 = sqrt(a_var); }
            else               { m1 = a_mean; s1 = sqrt(a_var); m2 = b_mean; s2 = sqrt(b_var); }
            double dd;
            if (fabs(k1 - k2) < EPSILON)
                dd = normal_cdf_statrs(m1, s1, k1) - normal_cdf_statrs(m2, s2, k1);
            else
                dd = fabs(normal_cdf_statrs(m1, s1, k2) - normal_cdf_statrs(m1, s1, k1) +
                          normal_cdf_statrs(m2, s2, k1) - normal_cdf_statrs(m2, s2, k2));
            lnsum += log(dd);   /* ln of a negative d is NaN, as f64::ln in Rust */
            cnt += 1;
        }
    }
    if (cnt > 0) {
        double d = exp(lnsum / (double)cnt);
        if (isnan(d)) d = 1.0;
        return d;
    }
    return 1.0;
}

/* flight/metrics.py:445-538 (metabat_distance_nn) */
static double metabat_flight(const double *a, const double *b, uint32_t s) {
    double lnsum = 0.0;
    uint32_t cnt = 0;
    for (uint32_t i = 0; i < s; ++i) {
        double a_mean = a[2 * i] + 1e-6, a_var = a[2 * i + 1] + 1e-6;
        double b_mean = b[2 * i] + 1e-6, b_var = b[2 * i + 1] + 1e-6;
        if ((a_mean > 1e-6 || b_mean > 1e-6) && a_mean != b_mean) {
            if (a_var < 1.0) a_var = 1.0;
            if (b_var < 1.0) b_var = 1.0;
            double k1, k2;
            if (fabs(b_var - a_var) < 1e-4) {
                k1 = k2 = (a_mean + b_mean) / 2.0;
            } else {
                double tmp = sqrt(a_var * b_var * ((a_mean - b_mean) * (a_mean - b_mean) -
                                  2.0 * (a_var - b_var) * log(sqrt(b_var / a_var))));
                k1 = (tmp - a_mean * b_var + b_mean * a_var) / (a_var - b_var);
                k2 = (tmp + a_mean * b_var - b_mean * a_var) / (b_var - a_var);
            }
            if (k1 > k2) { double t = k1; k1 = k2; k2 = t; }
            double m1, s1, m2, s2;
            if (a_var > b_var) { m1 = b_mean; s1 = sqrt(b_var); m2 = a_mean; s2 = sqrt(a_var); }
            else               { m1 = a_mean; s1 = sqrt(a_var); m2 = b_mean; s2 = sqrt(b_var); }
            double dd;
            if (k1 == k2)
                dd = fabs(normal_cdf_flight(m1, s1, k1) - normal_cdf_flight(m2, s2, k1));
            else
                dd = fabs(normal_cdf_flight(m1, s1, k2) - normal_cdf_flight(m1, s1, k1) +
                          normal_cdf_flight(m2, s2, k1) - normal_cdf_flight(m2, s2, k2));
            dd = fmin(fmax(dd, 1e-6), 1.0 - 1e-6);
            lnsum += log(dd);
            cnt += 1;
        }
    }
    if (cnt >= 1) return exp(lnsum / (double)cnt);
    return 1.0;
}

double orc_metabat(const double *a, const double *b, uint32_t s, int variant) {
    return variant == ORC_VARIANT_FLIGHT ? metabat_flight(a, b, s) : metabat_rust(a, b, s);
}

/* embedder.rs:108-127.  FLIGHT variant: rho (= 2x the Rust value), no guards. */
float orc_tnf_distance_f32(const double *x, const double *y, uint32_t d, int variant) {
    double v = variant == ORC_VARIANT_FLIGHT ? orc_rho_flight(x, y, d) : orc_kmer_correlation(x, y, d);
    if (variant == ORC_VARIANT_RUST && isnan(v)) return 0.0f;    /* :122-124 */
    /* flight leaves NaN (0/0 for two constant rows) undefined for NNDescent; this framework
     * defines it as the metric's maximum so that ordering is total. */
    if (isnan(v)) return 2.0f;
    return (float)v;
}

/* embedder.rs:133-171 with the MinHash factor fixed at 1.0 (out of v1 scope) */
float orc_depth_distance_f32(const double *a, const double *b, uint32_t s, int variant) {
    double v = orc_metabat(a, b, s, variant) * 1.0;
    if (isnan(v)) return 1.0f;    /* :166-168 (flight cannot produce NaN here: d is clamped) */
    return (float)v;
}
