// knn_exact.cu — the part of rg_knn_load that makes the FP32 select phase safe for an EXACT coverage graph.
//
// The select kernels (knn_depth2.cu, knn_depth_sym.cu) evaluate MetabatDistance (coverage_calculator.rs:588-666) in
// FP32 and only choose K' candidates; knn_rescore.cu then evaluates the reference's f64 statements on those.  A
// continuous FP32 error is covered by the margin certificate of knn_rescore.cu.  What a margin cannot cover are the
// metric's BRANCHES, where an FP32-rounded operand can flip a predicate that the f64 operands decide the other way and
// move the distance by a finite amount:
//   (1) `a_mean != b_mean` (:619): two means that differ in f64 and collide in FP32 are skipped by the select kernel
//       although the reference counts the sample;
//   (2) `a_var > b_var` (:638) picks p1 / p2 and with it the SIGN the Rust equal-variance branch keeps (:647): two
//       variances that differ in f64 and collide in FP32 may get the other sign (negative -> NaN -> 1.0, :659-661),
//       and above 2^10 one FP32 ulp exceeds MIN_VAR_EPSILON, so they may also land on the other side of (:620);
//   (3) `|a_var - b_var| < 1e-4` (:620) for FP32-distinct variances: handled inside the kernel by a threshold lowered
//       by the rounding error of the difference (knn_depth_common.cuh), so that "FP32 says equal" implies "f64 says
//       equal"; the other direction only moves the value continuously or lowers the select distance.
// (1) and (2) always involve two rows whose values share one FP32 rounding in one column, so they can be enumerated:
// per (sample, field) column the f64 values are sorted, every run of equal FP32 roundings that holds more than one f64
// value yields its cross pairs, and each pair becomes an EXTRA CANDIDATE of both rows — the f64 re-evaluation then
// sees the pair whatever the select kernel made of it.  Runs longer than RUNCAP (near-duplicate clusters) mark their
// rows for the exhaustive f64 pass instead.  Cost: 2 S radix sorts of n (f64, u32) pairs per load.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>

#include "rg_internal.cuh"

namespace {

constexpr uint32_t RUNCAP = 128;

// field 0: mean + eps (:600-601); field 1: max(var + eps, 1) (:602-609) — the values knn_transform.cu rounds to FP32
__global__ void column_keys_kernel(const double *__restrict__ cov, uint32_t n, uint32_t s, uint32_t smp, int field,
                                   double eps, double *__restrict__ keys, uint32_t *__restrict__ rows) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = cov[(uint64_t)i * 2 * s + 2 * smp + field] + eps;
    if (field == 1 && v < 1.0) v = 1.0;
    keys[i] = v;
    rows[i] = i;
}

// keys sorted ascending.  The first element of every run of equal FP32 roundings checks whether the run holds more
// than one f64 value (sorted: first != last) and, if so, emits the cross pairs in both directions.
__global__ void collide_kernel(const double *__restrict__ keys, const uint32_t *__restrict__ rows, uint32_t n,
                               uint64_t *__restrict__ pairs, unsigned long long *__restrict__ n_pairs, uint64_t cap,
                               uint8_t *__restrict__ row_flag) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float k32 = (float)keys[p];
    if (p > 0 && (float)keys[p - 1] == k32) return;          // not the start of a run
    if (p + 1 >= n || (float)keys[p + 1] != k32) return;     // run of one
    uint32_t lo = p + 1, hi = n;                              // end of the run: the FP32 rounding is monotonic
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if ((float)keys[mid] == k32) lo = mid + 1; else hi = mid;
    }
    const uint32_t end = lo;
    if (keys[end - 1] == keys[p]) return;                     // one f64 value: FP32 equality is f64 equality
    if (end - p > RUNCAP) {
        for (uint32_t i = p; i < end; ++i) row_flag[rows[i]] = 1;
        return;
    }
    for (uint32_t i = p; i < end; ++i)
        for (uint32_t j = i + 1; j < end; ++j) {
            if (keys[i] == keys[j]) continue;
            const uint64_t ri = rows[i], rj = rows[j];
            const unsigned long long slot = atomicAdd(n_pairs, 2ull);
            if (slot + 2 <= cap) {
                pairs[slot] = (ri << 32) | rj;
                pairs[slot + 1] = (rj << 32) | ri;
            } else {
                row_flag[ri] = 1;
                row_flag[rj] = 1;
            }
        }
}

// uniq: sorted unique (row << 32 | candidate); off[r] = number of entries of rows < r
__global__ void extra_csr_kernel(const uint64_t *__restrict__ uniq, uint64_t count, uint32_t n,
                                 uint64_t *__restrict__ off, uint32_t *__restrict__ idx) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < count) idx[t] = (uint32_t)uniq[t];
    if (t <= n) {
        const uint64_t key = t << 32;
        uint64_t lo = 0, hi = count;
        while (lo < hi) {
            const uint64_t mid = lo + (hi - lo) / 2;
            if (uniq[mid] < key) lo = mid + 1; else hi = mid;
        }
        off[t] = lo;
    }
}

}  // namespace

// Called by rg_knn_load after the coverage transform (tables resident in h->cov64).  Leaves h->extra_off / extra_idx /
// row_flag and h->n_extra; synchronises the stream.
int32_t rg_build_collision_patches(rg_handle *h, uint64_t n64, uint32_t s, uint32_t variant) {
    const uint32_t n = (uint32_t)n64;
    const double eps = variant == RG_VARIANT_FLIGHT ? 1e-6 : 1e-7;
    const int fields = variant == RG_VARIANT_FLIGHT ? 1 : 2;      // flight takes abs() in both branches: no sign rule
    const uint64_t cap = 64ull * n + (1ull << 20);
    DevBuf keys[2], rows[2], tmp, pairs, pairs2, ctr, uniq, nsel;
    auto done = [&](int32_t rc) {
        for (DevBuf *b : {&keys[0], &keys[1], &rows[0], &rows[1], &tmp, &pairs, &pairs2, &ctr, &uniq, &nsel}) rg_release(*b);
        return rc;
    };
#define XT(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define XC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__)); } while (0)
    for (int b = 0; b < 2; ++b) {
        XT(rg_reserve(h, keys[b], (size_t)n * sizeof(double)));
        XT(rg_reserve(h, rows[b], (size_t)n * sizeof(uint32_t)));
    }
    XT(rg_reserve(h, pairs, cap * sizeof(uint64_t)));
    XT(rg_reserve(h, ctr, sizeof(unsigned long long)));
    XT(rg_reserve(h, h->row_flag, n));
    XT(rg_reserve(h, h->extra_off, ((size_t)n + 1) * sizeof(uint64_t)));
    XC(cudaMemsetAsync(ctr.p, 0, sizeof(unsigned long long), h->stream));
    XC(cudaMemsetAsync(h->row_flag.p, 0, n, h->stream));
    size_t tmp_bytes = 0;
    XC(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const double *)keys[0].p, (double *)keys[1].p,
                                       (const uint32_t *)rows[0].p, (uint32_t *)rows[1].p, (int)n, 0, 64, h->stream));
    XT(rg_reserve(h, tmp, tmp_bytes));
    const unsigned blocks = (n + 255) / 256;
    for (uint32_t smp = 0; smp < s; ++smp)
        for (int field = 0; field < fields; ++field) {
            column_keys_kernel<<<blocks, 256, 0, h->stream>>>((const double *)h->cov64.p, n, s, smp, field, eps,
                                                              (double *)keys[0].p, (uint32_t *)rows[0].p);
            XC(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, (const double *)keys[0].p, (double *)keys[1].p,
                                               (const uint32_t *)rows[0].p, (uint32_t *)rows[1].p, (int)n, 0, 64, h->stream));
            collide_kernel<<<blocks, 256, 0, h->stream>>>((const double *)keys[1].p, (const uint32_t *)rows[1].p, n,
                                                          (uint64_t *)pairs.p, (unsigned long long *)ctr.p, cap,
                                                          (uint8_t *)h->row_flag.p);
            h->tm.kernel_launches += 2;
        }
    XC(cudaGetLastError());
    unsigned long long found = 0;
    XC(cudaMemcpyAsync(&found, ctr.p, sizeof(found), cudaMemcpyDeviceToHost, h->stream));
    XC(cudaStreamSynchronize(h->stream));
    const uint64_t count = found < cap ? found : cap / 2 * 2;
    uint64_t n_uniq = 0;
    if (count) {
        XT(rg_reserve(h, pairs2, count * sizeof(uint64_t)));
        XT(rg_reserve(h, uniq, count * sizeof(uint64_t)));
        XT(rg_reserve(h, nsel, sizeof(uint64_t)));
        size_t b1 = 0, b2 = 0;
        XC(cub::DeviceRadixSort::SortKeys(nullptr, b1, (const uint64_t *)pairs.p, (uint64_t *)pairs2.p, (int64_t)count, 0, 64, h->stream));
        XC(cub::DeviceSelect::Unique(nullptr, b2, (const uint64_t *)pairs2.p, (uint64_t *)uniq.p, (uint64_t *)nsel.p, (int64_t)count, h->stream));
        XT(rg_reserve(h, tmp, b1 > b2 ? b1 : b2));
        size_t tb = tmp.cap;
        XC(cub::DeviceRadixSort::SortKeys(tmp.p, tb, (const uint64_t *)pairs.p, (uint64_t *)pairs2.p, (int64_t)count, 0, 64, h->stream));
        tb = tmp.cap;
        XC(cub::DeviceSelect::Unique(tmp.p, tb, (const uint64_t *)pairs2.p, (uint64_t *)uniq.p, (uint64_t *)nsel.p, (int64_t)count, h->stream));
        XC(cudaMemcpyAsync(&n_uniq, nsel.p, sizeof(n_uniq), cudaMemcpyDeviceToHost, h->stream));
        XC(cudaStreamSynchronize(h->stream));
    }
    XT(rg_reserve(h, h->extra_idx, (n_uniq ? n_uniq : 1) * sizeof(uint32_t)));
    const uint64_t threads = n_uniq > (uint64_t)n + 1 ? n_uniq : (uint64_t)n + 1;
    extra_csr_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, h->stream>>>((const uint64_t *)uniq.p, n_uniq, n,
                                                                             (uint64_t *)h->extra_off.p, (uint32_t *)h->extra_idx.p);
    XC(cudaGetLastError());
    XC(cudaStreamSynchronize(h->stream));
    h->tm.kernel_launches += 1;
    h->n_extra = n_uniq;
    h->has_extra = true;
#undef XT
#undef XC
    return done(RG_OK);
}
