// cluster_ops.cu — SURVEY.md 8(f) rank 4: the two O(n^2) helpers the Rust path's cluster search leans on
// (src/clustering/clusterer.rs:76-127 calls them 100 times per run on the embedding).
//
//   rg_condensed_euclidean   condensed_pairwise_distance, src/clustering/cluster_utils.rs:7-34: Euclidean distance of
//                            all row pairs i < j in condensed order, NaN -> 0; f64, summed in the reference's order.
//   rg_silhouette            silhouette_score, cluster_utils.rs:38-120, WITHOUT materialising the n (n - 1) / 2 distances
//                            (4 TB at 1M points): per point i of a cluster, a = mean distance to the other members,
//                            b = the smallest mean distance to another cluster's members, s = (b - a) / max(a, b); the score
//                            is sum(s) / n (:119 divides by the number of rows, outliers included).
//                            Deviation, stated: the reference computes `a` over `indices.iter().skip(*i)` of a HashSet
//                            (:57) — a slice of the set that depends on the process's hash seed, so its value is not
//                            reproducible between two runs of the reference itself.  Here `a` is the mean over ALL other
//                            members (what the comment at :50 says it is).  Everything else follows the reference: a
//                            single cluster scores 0 (:39-41), a cluster of one has a = 0 (:66-69), b starts at -1 (:53,
//                            :90-92), NaN s values are summed as they are.
// Points are grouped by cluster on the host (counting sort), so every per-cluster sum runs over a contiguous range in
// a fixed order: one CTA per point, one warp per cluster at a time, lanes stride the members, tree reduction — no
// floating-point atomics, the result is deterministic.
#include <vector>

#include "rg_internal.cuh"

namespace {

__device__ __forceinline__ double euclid(const double *__restrict__ x, const double *__restrict__ y, uint32_t d) {
    double r = 0.0;
    for (uint32_t c = 0; c < d; ++c) { const double t = x[c] - y[c]; r += t * t; }   // :19-23 (x - y).powi(2), sum, sqrt
    r = sqrt(r);
    return isnan(r) ? 0.0 : r;                                                        // :24-26
}

__global__ void condensed_kernel(const double *__restrict__ data, uint32_t n, uint32_t d, double *__restrict__ out) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= n || j >= n || i >= j) return;
    out[(uint64_t)i * n - (uint64_t)i * (i + 1) / 2 + (j - i - 1)] = euclid(data + (uint64_t)i * d, data + (uint64_t)j * d, d);
}

// order: point ids grouped by cluster; start[c] .. start[c + 1]: the members of cluster c.  One CTA per clustered point.
__global__ void __launch_bounds__(256)
silhouette_kernel(const double *__restrict__ data, uint32_t d, const uint32_t *__restrict__ order, const uint32_t *__restrict__ start,
                  uint32_t n_clusters, uint32_t n_clustered, double *__restrict__ s_out) {
    extern __shared__ double xi[];                       // the point's coordinates
    __shared__ double w_a[8], w_b[8];
    const uint32_t t = blockIdx.x;
    if (t >= n_clustered) return;
    const uint32_t i = order[t];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    for (uint32_t c = threadIdx.x; c < d; c += blockDim.x) xi[c] = data[(uint64_t)i * d + c];
    __syncthreads();
    uint32_t lo = 0, hi = n_clusters;                    // the cluster of position t
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (start[mid] <= t) lo = mid; else hi = mid; }
    const uint32_t own = lo;
    double a = 0.0, b = -1.0;                            // :50-53
    for (uint32_t c = warp; c < n_clusters; c += warps) {
        const uint32_t m0 = start[c], m1 = start[c + 1];
        double sum = 0.0;
        for (uint32_t p = m0 + lane; p < m1; p += 32) {
            const uint32_t j = order[p];
            if (j != i) sum += euclid(xi, data + (uint64_t)j * d, d);
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (c == own) {
            uint32_t cnt = m1 - m0 - 1;                  // the other members
            if (cnt == 0) cnt = 1;                       // :66-68
            a = sum / (double)cnt;
        } else {
            uint32_t cnt = m1 - m0;
            if (cnt == 0) cnt = 1;                       // :84-86
            const double mean = sum / (double)cnt;
            if (mean < b || b < 0.0) b = mean;           // :90-92
        }
    }
    if (lane == 0) { w_a[warp] = a; w_b[warp] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double aa = 0.0, bb = -1.0;
        for (int w = 0; w < warps; ++w) {
            aa += w_a[w];                                // exactly one warp holds the own-cluster mean, the others 0
            if (w_b[w] >= 0.0 && (w_b[w] < bb || bb < 0.0)) bb = w_b[w];
        }
        s_out[i] = (bb - aa) / fmax(aa, bb);             // :94
    }
}

}  // namespace

extern "C" int32_t rg_condensed_euclidean(rg_handle *h, const double *data, uint64_t n, uint32_t d, double *out) {
    if (!h) return RG_ERR_INVALID;
    if (!data || !out || n < 2 || d == 0) return rg_fail(h, RG_ERR_INVALID, "rg_condensed_euclidean: need data, out, n >= 2, d >= 1");
    if (n > 1000000) return rg_fail(h, RG_ERR_INVALID, "rg_condensed_euclidean: n too large for a condensed matrix");
    RG_CUDA(h, cudaSetDevice(h->device));
    const uint64_t P = n * (n - 1) / 2;
    DevBuf dd, dout;
    int32_t rc = rg_reserve(h, dd, n * d * 8);
    if (rc == RG_OK) rc = rg_reserve(h, dout, P * 8);
    cudaError_t e = cudaSuccess;
    if (rc == RG_OK) {
        e = cudaMemcpyAsync(dd.p, data, n * d * 8, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) {
            const dim3 blk(32, 8), grd((unsigned)((n + 31) / 32), (unsigned)((n + 7) / 8));
            condensed_kernel<<<grd, blk, 0, h->stream>>>((const double *)dd.p, (uint32_t)n, d, (double *)dout.p);
            e = cudaGetLastError();
            h->tm.kernel_launches += 1;
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout.p, P * 8, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = rg_fail(h, RG_ERR_CUDA, "rg_condensed_euclidean: %s", cudaGetErrorString(e));
        else { h->tm.h2d_bytes += n * d * 8; h->tm.d2h_bytes += P * 8; }
    }
    rg_release(dd); rg_release(dout);
    return rc;
}

extern "C" int32_t rg_silhouette(rg_handle *h, const double *data, uint64_t n, uint32_t d, const int32_t *labels,
                                 uint32_t n_clusters, double *score, double *per_point) {
    if (!h) return RG_ERR_INVALID;
    if (!data || !labels || !score || n == 0 || d == 0 || d > 4096) return rg_fail(h, RG_ERR_INVALID, "rg_silhouette: bad argument");
    if (n >= 0xffffffffull) return rg_fail(h, RG_ERR_INVALID, "rg_silhouette: n must be < 2^32-1");
    for (uint64_t i = 0; i < n; ++i)
        if (labels[i] < -1 || labels[i] >= (int64_t)n_clusters)
            return rg_fail(h, RG_ERR_INVALID, "rg_silhouette: labels[%llu] = %d outside [-1, %u)", (unsigned long long)i, labels[i], n_clusters);
    if (per_point) for (uint64_t i = 0; i < n; ++i) per_point[i] = 0.0;
    if (n_clusters <= 1) { *score = 0.0; return RG_OK; }                  // :39-41
    std::vector<uint32_t> start, order;
    try { start.assign(n_clusters + 1, 0); order.resize(n); } catch (...) { return rg_fail(h, RG_ERR_OOM, "rg_silhouette: host allocation failed"); }
    for (uint64_t i = 0; i < n; ++i) if (labels[i] >= 0) start[labels[i] + 1] += 1;
    for (uint32_t c = 0; c < n_clusters; ++c) start[c + 1] += start[c];
    const uint32_t n_clustered = start[n_clusters];
    { std::vector<uint32_t> fill(start.begin(), start.end() - 1);
      for (uint64_t i = 0; i < n; ++i) if (labels[i] >= 0) order[fill[labels[i]]++] = (uint32_t)i; }
    RG_CUDA(h, cudaSetDevice(h->device));
    DevBuf dd, dord, dst, ds;
    int32_t rc = rg_reserve(h, dd, n * d * 8);
    if (rc == RG_OK) rc = rg_reserve(h, dord, (n ? n : 1) * 4);
    if (rc == RG_OK) rc = rg_reserve(h, dst, (n_clusters + 1) * 4);
    if (rc == RG_OK) rc = rg_reserve(h, ds, n * 8);
    std::vector<double> s_host;
    if (rc == RG_OK) { try { s_host.resize(n); } catch (...) { rc = rg_fail(h, RG_ERR_OOM, "rg_silhouette: host allocation failed"); } }
    cudaError_t e = cudaSuccess;
    if (rc == RG_OK) {
        e = cudaMemcpyAsync(dd.p, data, n * d * 8, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dord.p, order.data(), n * 4, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dst.p, start.data(), (n_clusters + 1) * 4, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(ds.p, 0, n * 8, h->stream);
        if (e == cudaSuccess && n_clustered) {
            cudaEventRecord(h->ev[3], h->stream);
            silhouette_kernel<<<n_clustered, 256, d * 8, h->stream>>>((const double *)dd.p, d, (const uint32_t *)dord.p, (const uint32_t *)dst.p,
                                                                     n_clusters, n_clustered, (double *)ds.p);
            e = cudaGetLastError();
            cudaEventRecord(h->ev[4], h->stream);
            cudaEventRecord(h->ev[5], h->stream);
            h->tm.kernel_launches += 1;
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(s_host.data(), ds.p, n * 8, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = rg_fail(h, RG_ERR_CUDA, "rg_silhouette: %s", cudaGetErrorString(e));
    }
    if (rc == RG_OK) {
        if (n_clustered) { h->tm.select_ms = rg_elapsed(h->ev[3], h->ev[4]); h->tm.rescore_ms = 0.0; }
        // :106-119 the sum over clusters and points, in cluster / member order, divided by the number of rows
        double sum = 0.0;
        for (uint32_t t = 0; t < n_clustered; ++t) sum += s_host[order[t]];
        *score = sum / (double)n;
        if (per_point) for (uint32_t t = 0; t < n_clustered; ++t) per_point[order[t]] = s_host[order[t]];
    }
    rg_release(dd); rg_release(dord); rg_release(dst); rg_release(ds);
    return rc;
}
