// rank_stat.cu — flight's all-pairs rank statistic on the device.
//
// Reference: ProfileDistanceEngine.makeRanks / makeRankStat, flight/flight/distance.py:47-108, driven by
// flight/flight/rosella/rosella.py:337-350 with fun = a / max(a) and one process:
//     cov_d  = pdist(cov, metrics.coverage_distance)          metrics.py:429-443 (identical rows -> 0, else the geometric
//                                                             mean of metabat_distance, :343-427 == metabat_distance_nn)
//     kmer_d = pdist(tnf, 'euclidean')
//     cov_ranks  = fractional ranks of cov_d   (ties averaged, distance.py:168-201)
//     kmer_ranks = fractional ranks of kmer_d, positions replaced by cumulative pair weights len_i * len_j (:53-58)
//     stat = (cov_ranks / max)^w * (kmer_ranks / max),  w = max(1/3, S / (S + 4)), 1/2 for one sample
// This is the default product's O(N^2) wall: ~11-20 us per pair through scipy + numba on one core
// (profiles/r2_c1_flight_cpu.json), and an O(N^2) squareform handed to UMAP.
//
// Device plan, everything in f64 (ranks are decided by exact order and exact ties, so no FP32 stage):
//   1. pair_distance_kernel: one thread per pair i < j, both distances with the reference's statements (metric_f64.cuh),
//      stored condensed (the order scipy's pdist uses);
//   2. ranks = positions in the sorted distance vector: one radix sort per metric (keys only for the coverage ranks; with
//      the pair weights as payload and an inclusive scan for the weighted ranks), then two binary searches per pair for
//      the ends of its streak of equal values — rank = (end of the previous streak + end of this streak + 1) / 2 exactly
//      as :189-193 computes it;
//   3. the statistic per pair, condensed; optionally the k smallest per row (sparse output instead of the reference's
//      dense squareform).
// Memory: 5 f64 arrays of n (n - 1) / 2 entries (~50k contigs in 50 GB); larger n needs a distributed sort (not built).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/device/device_scan.cuh>

#include "metric_f64.cuh"
#include "rg_internal.cuh"

namespace {

using namespace metric64;

// condensed index of the pair (i, j), i < j, among n rows: the order of scipy.spatial.distance.pdist
__device__ __forceinline__ uint64_t condensed(uint64_t i, uint64_t j, uint64_t n) { return i * n - i * (i + 1) / 2 + (j - i - 1); }

__global__ void pair_distance_kernel(const double *__restrict__ cov, const double *__restrict__ tnf, const double *__restrict__ len,
                                     uint32_t n, uint32_t s, uint32_t d, double *__restrict__ dc, double *__restrict__ dk,
                                     double *__restrict__ w) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= n || j >= n || i >= j) return;
    const double *a = cov + (uint64_t)i * 2 * s, *b = cov + (uint64_t)j * 2 * s;
    bool same = true;                                          // np.array_equal(a, b), metrics.py:430
    for (uint32_t c = 0; c < 2 * s; ++c) same &= a[c] == b[c];
    const uint64_t p = condensed(i, j, n);
    dc[p] = same ? 0.0 : metabat_f64<RG_VARIANT_FLIGHT>(a, b, s);
    dk[p] = tnf_euclidean_f64(tnf + (uint64_t)i * d, tnf + (uint64_t)j * d, d);
    w[p] = len[i] * len[j];                                    // distance.py:57 (integers below 2^53: exact in f64)
}

// fractional rank of v among the sorted values (distance.py:168-201).  pos: unweighted (cum == nullptr) the rank at the end
// of a streak is its last 1-based position; weighted it is the cumulative weight there.
__device__ __forceinline__ double frac_rank(const double *__restrict__ sorted, const double *__restrict__ cum, uint64_t count, double v) {
    uint64_t lo = 0, hi = count;                               // first position with sorted >= v
    while (lo < hi) { const uint64_t mid = lo + (hi - lo) / 2; if (sorted[mid] < v) lo = mid + 1; else hi = mid; }
    const uint64_t first = lo;
    hi = count;                                                // first position with sorted > v
    while (lo < hi) { const uint64_t mid = lo + (hi - lo) / 2; if (sorted[mid] <= v) lo = mid + 1; else hi = mid; }
    const uint64_t last = lo;                                  // the streak is [first, last)
    const double end_prev = first == 0 ? 0.0 : (cum ? cum[first - 1] : (double)first);
    const double end_this = cum ? cum[last - 1] : (double)last;
    // :193 (rflag[:1] + 1, rflag[1:] + rflag[:-1] + 1) * 0.5
    return (end_this + end_prev + 1.0) * 0.5;
}

__global__ void rank_kernel(const double *__restrict__ val, const double *__restrict__ sorted, const double *__restrict__ cum,
                            uint64_t count, double *__restrict__ rank) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < count) rank[p] = frac_rank(sorted, cum, count, val[p]);
}

__global__ void stat_kernel(double *__restrict__ cov_rank, const double *__restrict__ kmer_rank, uint64_t count,
                            const double *__restrict__ maxes, double w) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < count) cov_rank[p] = pow(cov_rank[p] / maxes[0], w) * (kmer_rank[p] / maxes[1]);   // distance.py:106
}

// the k smallest statistics of every row, ascending by (stat, index); one warp per row
__global__ void __launch_bounds__(128)
row_topk_kernel(const double *__restrict__ stat, uint32_t n, uint32_t k, uint32_t *__restrict__ idx, double *__restrict__ out) {
    extern __shared__ __align__(16) uint8_t sm_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    double *ld = reinterpret_cast<double *>(sm_raw) + (size_t)warp * k;
    uint32_t *li = reinterpret_cast<uint32_t *>(sm_raw + (size_t)wpb * k * 8) + (size_t)warp * k;
    const uint32_t r = blockIdx.x * wpb + warp;
    if (r >= n) return;
    for (uint32_t e = lane; e < k; e += 32) { ld[e] = INFINITY; li[e] = 0xffffffffu; }
    __syncwarp();
    double thr = INFINITY;
    for (uint32_t c0 = 0; c0 < n; c0 += 32) {
        const uint32_t c = c0 + lane;
        double v = INFINITY;
        if (c < n && c != r) v = stat[c > r ? condensed(r, c, n) : condensed(c, r, n)];
        unsigned m = __ballot_sync(0xffffffffu, v < thr);
        while (m) {                                            // candidates in increasing index: ties keep index order
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const double cv = __shfl_sync(0xffffffffu, v, src);
            const uint32_t ci = c0 + src;
            if (!(cv < thr)) continue;
            uint32_t pos = 0;                                  // entries <= cv stay in front
            for (uint32_t e = lane; e < k; e += 32) pos += ld[e] <= cv;
#pragma unroll
            for (int o = 16; o; o >>= 1) pos += __shfl_xor_sync(0xffffffffu, pos, o);
            for (uint32_t e0 = (k - 1) / 32 * 32 + 32; e0 > 0; e0 -= 32) {     // shift the tail right by one, back to front
                const uint32_t e = e0 - 32 + lane;
                double mv = 0.0; uint32_t mi = 0;
                const bool move = e < k && e > pos;
                if (move) { mv = ld[e - 1]; mi = li[e - 1]; }
                __syncwarp();
                if (move) { ld[e] = mv; li[e] = mi; }
                __syncwarp();
            }
            if (lane == 0) { ld[pos] = cv; li[pos] = ci; }
            __syncwarp();
            thr = ld[k - 1];
        }
    }
    for (uint32_t e = lane; e < k; e += 32) { idx[(uint64_t)r * k + e] = li[e]; out[(uint64_t)r * k + e] = ld[e]; }
}

}  // namespace

extern "C" int32_t rg_rank_stat(rg_handle *h, const double *tnf, const double *cov, const double *lengths, uint64_t n,
                                uint32_t d, uint32_t s, uint32_t k, double *stat_condensed, uint32_t *idx, double *stat_knn) {
    if (!h) return RG_ERR_INVALID;
    if (!tnf || !cov || !lengths) return rg_fail(h, RG_ERR_INVALID, "rg_rank_stat: NULL input");
    if (n < 2 || n > 200000) return rg_fail(h, RG_ERR_INVALID, "rg_rank_stat: n=%llu outside [2, 200000] (n (n - 1) / 2 pairs are held in HBM)", (unsigned long long)n);
    if (d == 0 || s == 0) return rg_fail(h, RG_ERR_INVALID, "rg_rank_stat: d and s must be positive");
    if ((idx || stat_knn) && (!idx || !stat_knn || k == 0 || k > 1024 || k > n - 1))
        return rg_fail(h, RG_ERR_INVALID, "rg_rank_stat: sparse output needs idx, stat_knn and 1 <= k <= min(1024, n - 1)");
    if (!stat_condensed && !idx) return rg_fail(h, RG_ERR_INVALID, "rg_rank_stat: no output requested");
    RG_CUDA(h, cudaSetDevice(h->device));
    const uint64_t P = n * (n - 1) / 2;
    DevBuf dcov, dtnf, dlen, dc, dk, w, sc, sk, sw, tmp, mx, oi, os;
    auto done = [&](int32_t rc) {
        for (DevBuf *b : {&dcov, &dtnf, &dlen, &dc, &dk, &w, &sc, &sk, &sw, &tmp, &mx, &oi, &os}) rg_release(*b);
        return rc;
    };
#define XT(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define XC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__)); } while (0)
    XT(rg_reserve(h, dcov, n * 2 * s * 8)); XT(rg_reserve(h, dtnf, n * d * 8)); XT(rg_reserve(h, dlen, n * 8));
    XT(rg_reserve(h, dc, P * 8)); XT(rg_reserve(h, dk, P * 8)); XT(rg_reserve(h, w, P * 8));
    XT(rg_reserve(h, sc, P * 8)); XT(rg_reserve(h, sk, P * 8)); XT(rg_reserve(h, sw, P * 8)); XT(rg_reserve(h, mx, 16));
    XC(cudaMemcpyAsync(dcov.p, cov, n * 2 * s * 8, cudaMemcpyHostToDevice, h->stream));
    XC(cudaMemcpyAsync(dtnf.p, tnf, n * d * 8, cudaMemcpyHostToDevice, h->stream));
    XC(cudaMemcpyAsync(dlen.p, lengths, n * 8, cudaMemcpyHostToDevice, h->stream));
    h->tm.h2d_bytes += n * (2 * s + d + 1) * 8;
    XC(cudaEventRecord(h->ev[3], h->stream));
    const dim3 blk(32, 8), grd((unsigned)((n + 31) / 32), (unsigned)((n + 7) / 8));
    pair_distance_kernel<<<grd, blk, 0, h->stream>>>((const double *)dcov.p, (const double *)dtnf.p, (const double *)dlen.p, (uint32_t)n, s, d,
                                                     (double *)dc.p, (double *)dk.p, (double *)w.p);
    XC(cudaGetLastError());
    XC(cudaEventRecord(h->ev[4], h->stream));
    size_t b1 = 0, b2 = 0, b3 = 0, b4 = 0;
    XC(cub::DeviceRadixSort::SortKeys(nullptr, b1, (const double *)dc.p, (double *)sc.p, (int64_t)P, 0, 64, h->stream));
    XC(cub::DeviceRadixSort::SortPairs(nullptr, b2, (const double *)dk.p, (double *)sk.p, (const double *)w.p, (double *)sw.p, (int64_t)P, 0, 64, h->stream));
    XC(cub::DeviceScan::InclusiveSum(nullptr, b3, (const double *)sw.p, (double *)sw.p, (int64_t)P, h->stream));
    XC(cub::DeviceReduce::Max(nullptr, b4, (const double *)dc.p, (double *)mx.p, (int64_t)P, h->stream));
    XT(rg_reserve(h, tmp, std::max(std::max(b1, b2), std::max(b3, b4))));
    size_t tb = tmp.cap;
    XC(cub::DeviceRadixSort::SortKeys(tmp.p, tb, (const double *)dc.p, (double *)sc.p, (int64_t)P, 0, 64, h->stream));
    tb = tmp.cap;
    XC(cub::DeviceRadixSort::SortPairs(tmp.p, tb, (const double *)dk.p, (double *)sk.p, (const double *)w.p, (double *)sw.p, (int64_t)P, 0, 64, h->stream));
    tb = tmp.cap;
    XC(cub::DeviceScan::InclusiveSum(tmp.p, tb, (const double *)sw.p, (double *)sw.p, (int64_t)P, h->stream));
    const unsigned pb = (unsigned)((P + 255) / 256);
    // ranks overwrite the weight / distance arrays they no longer need: w <- cov ranks, dk <- kmer ranks (dk is read per
    // thread before its own slot is written)
    rank_kernel<<<pb, 256, 0, h->stream>>>((const double *)dc.p, (const double *)sc.p, nullptr, P, (double *)w.p);
    rank_kernel<<<pb, 256, 0, h->stream>>>((const double *)dk.p, (const double *)sk.p, (const double *)sw.p, P, (double *)dk.p);
    tb = tmp.cap;
    XC(cub::DeviceReduce::Max(tmp.p, tb, (const double *)w.p, (double *)mx.p, (int64_t)P, h->stream));
    tb = tmp.cap;
    XC(cub::DeviceReduce::Max(tmp.p, tb, (const double *)dk.p, (double *)mx.p + 1, (int64_t)P, h->stream));
    double wexp = (double)s / ((double)s + 4.0);                // distance.py:96-98
    if (wexp < 1.0 / 3.0) wexp = 1.0 / 3.0;
    if (s == 1) wexp = 0.5;
    stat_kernel<<<pb, 256, 0, h->stream>>>((double *)w.p, (const double *)dk.p, P, (const double *)mx.p, wexp);
    XC(cudaGetLastError());
    XC(cudaEventRecord(h->ev[5], h->stream));
    h->tm.kernel_launches += 4;
    if (stat_condensed) {
        XC(cudaMemcpyAsync(stat_condensed, w.p, P * 8, cudaMemcpyDeviceToHost, h->stream));
        h->tm.d2h_bytes += P * 8;
    }
    if (idx) {
        XT(rg_reserve(h, oi, n * k * 4)); XT(rg_reserve(h, os, n * k * 8));
        const int wpb = 4;
        row_topk_kernel<<<(unsigned)((n + wpb - 1) / wpb), wpb * 32, (size_t)wpb * k * 12, h->stream>>>((const double *)w.p, (uint32_t)n, k,
                                                                                                   (uint32_t *)oi.p, (double *)os.p);
        XC(cudaGetLastError());
        XC(cudaMemcpyAsync(idx, oi.p, n * k * 4, cudaMemcpyDeviceToHost, h->stream));
        XC(cudaMemcpyAsync(stat_knn, os.p, n * k * 8, cudaMemcpyDeviceToHost, h->stream));
        h->tm.kernel_launches += 1;
        h->tm.d2h_bytes += n * k * 12;
    }
    XC(cudaStreamSynchronize(h->stream));
    h->tm.select_ms = rg_elapsed(h->ev[3], h->ev[4]);           // all-pairs f64 distances
    h->tm.rescore_ms = rg_elapsed(h->ev[4], h->ev[5]);          // sorts, scans, ranks, statistic
#undef XT
#undef XC
    return done(RG_OK);
}
