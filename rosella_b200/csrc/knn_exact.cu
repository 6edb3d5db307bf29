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
// sees the pair whatever the select kernel made of it.  A row with more than 64 such partners in one column (a cluster of
// near-duplicates) takes the exhaustive f64 pass instead.  Cost: 2 S radix sorts of n (f64, u32) pairs per load.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/device/device_select.cuh>

#include "rg_internal.cuh"

namespace {

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

// keys sorted ascending.  Every element looks at the run of elements that share its FP32 rounding: the members whose
// f64 value differs from its own are the rows the select kernel may misjudge against this row.  Up to EMIT_CAP of them
// become extra candidates of this row (one directed pair each; the other row emits its own direction); a row with more
// partners takes the exhaustive pass instead.  The decision is per row, so one stray value inside a run of 300 000
// identical ones (variances clamped to exactly 1.0 plus one variance of 1 + 3e-8) costs one extra candidate per member
// and one exhaustive row — not 300 000 exhaustive rows.
constexpr uint32_t EMIT_CAP = 64;

__global__ void collide_kernel(const double *__restrict__ keys, const uint32_t *__restrict__ rows, uint32_t n,
                               uint64_t *__restrict__ pairs, unsigned long long *__restrict__ n_pairs, uint64_t cap,
                               uint8_t *__restrict__ row_flag) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const double v = keys[p];
    const float k32 = (float)v;
    const bool left = p > 0 && (float)keys[p - 1] == k32, right = p + 1 < n && (float)keys[p + 1] == k32;
    if (!left && !right) return;                              // alone in its FP32 rounding
    uint32_t lo = 0, hi = p;                                  // run of the FP32 rounding: [begin, end) (the rounding is monotonic)
    while (lo < hi) { const uint32_t mid = lo + (hi - lo) / 2; if ((float)keys[mid] < k32) lo = mid + 1; else hi = mid; }
    const uint32_t begin = lo;
    lo = p + 1; hi = n;
    while (lo < hi) { const uint32_t mid = lo + (hi - lo) / 2; if ((float)keys[mid] <= k32) lo = mid + 1; else hi = mid; }
    const uint32_t end = lo;
    if (keys[begin] == keys[end - 1]) return;                 // one f64 value: FP32 equality is f64 equality
    lo = begin; hi = p;                                       // this row's f64 value inside the run: [gb, ge)
    while (lo < hi) { const uint32_t mid = lo + (hi - lo) / 2; if (keys[mid] < v) lo = mid + 1; else hi = mid; }
    const uint32_t gb = lo;
    lo = p + 1; hi = end;
    while (lo < hi) { const uint32_t mid = lo + (hi - lo) / 2; if (keys[mid] <= v) lo = mid + 1; else hi = mid; }
    const uint32_t ge = lo;
    const uint32_t others = (gb - begin) + (end - ge);
    const uint64_t me = rows[p];
    if (others > EMIT_CAP) { row_flag[me] = 1; return; }
    const unsigned long long slot = atomicAdd(n_pairs, (unsigned long long)others);
    if (slot + others > cap) { row_flag[me] = 1; return; }    // the host grows the buffer and scans again
    uint64_t w = slot;
    for (uint32_t o = begin; o < gb; ++o) pairs[w++] = (me << 32) | rows[o];
    for (uint32_t o = ge; o < end; ++o) pairs[w++] = (me << 32) | rows[o];
}

}  // namespace

int32_t rg_rebuild_extras(rg_handle *h, const uint64_t *sk_pairs, uint64_t n_sk);

// Called by rg_knn_load after the coverage transform (tables resident in h->cov64).  Leaves h->extra_off / extra_idx /
// row_flag and h->n_extra; synchronises the stream.
int32_t rg_build_collision_patches(rg_handle *h, uint64_t n64, uint32_t s, uint32_t variant) {
    const uint32_t n = (uint32_t)n64;
    const double eps = variant == RG_VARIANT_FLIGHT ? 1e-6 : 1e-7;
    const int fields = variant == RG_VARIANT_FLIGHT ? 1 : 2;      // flight takes abs() in both branches: no sign rule
    // the scratch of the scan lives on the handle (h->ws): a load per call must not pay cudaMalloc / cudaFree of
    // several hundred MB (measured: 0.2 s per rg_knn_load at 1M rows)
    DevBuf &keys0 = h->ws[0], &keys1 = h->ws[1], &rows0 = h->ws[2], &rows1 = h->ws[3], &tmp = h->ws[4], &pairs = h->ws[5], &ctr = h->ws[6];
    RG_TRY(rg_reserve(h, keys0, (size_t)n * sizeof(double)));
    RG_TRY(rg_reserve(h, keys1, (size_t)n * sizeof(double)));
    RG_TRY(rg_reserve(h, rows0, (size_t)n * sizeof(uint32_t)));
    RG_TRY(rg_reserve(h, rows1, (size_t)n * sizeof(uint32_t)));
    RG_TRY(rg_reserve(h, pairs, (4ull * n + (1ull << 20)) * sizeof(uint64_t)));
    RG_TRY(rg_reserve(h, ctr, sizeof(unsigned long long)));
    RG_TRY(rg_reserve(h, h->row_flag, n));
    size_t tmp_bytes = 0;
    RG_CUDA(h, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const double *)keys0.p, (double *)keys1.p,
                                               (const uint32_t *)rows0.p, (uint32_t *)rows1.p, (int)n, 0, 64, h->stream));
    RG_TRY(rg_reserve(h, tmp, tmp_bytes));
    const unsigned blocks = (n + 255) / 256;
    unsigned long long found = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        const uint64_t cap = pairs.cap / sizeof(uint64_t);
        RG_CUDA(h, cudaMemsetAsync(ctr.p, 0, sizeof(unsigned long long), h->stream));
        RG_CUDA(h, cudaMemsetAsync(h->row_flag.p, 0, n, h->stream));
        for (uint32_t smp = 0; smp < s; ++smp)
            for (int field = 0; field < fields; ++field) {
                column_keys_kernel<<<blocks, 256, 0, h->stream>>>((const double *)h->cov64.p, n, s, smp, field, eps,
                                                                  (double *)keys0.p, (uint32_t *)rows0.p);
                size_t tb = tmp.cap;
                RG_CUDA(h, cub::DeviceRadixSort::SortPairs(tmp.p, tb, (const double *)keys0.p, (double *)keys1.p,
                                                           (const uint32_t *)rows0.p, (uint32_t *)rows1.p, (int)n, 0, 64, h->stream));
                collide_kernel<<<blocks, 256, 0, h->stream>>>((const double *)keys1.p, (const uint32_t *)rows1.p, n,
                                                              (uint64_t *)pairs.p, (unsigned long long *)ctr.p, cap,
                                                              (uint8_t *)h->row_flag.p);
                h->tm.kernel_launches += 2;
            }
        RG_CUDA(h, cudaGetLastError());
        RG_CUDA(h, cudaMemcpyAsync(&found, ctr.p, sizeof(found), cudaMemcpyDeviceToHost, h->stream));
        RG_CUDA(h, cudaStreamSynchronize(h->stream));
        if (found <= cap) break;
        // more pairs than the buffer holds (dense data): grow to what was counted and scan again (same count, so it fits)
        if (attempt == 1) return rg_fail(h, RG_ERR_STATE, "collision scan: pair count changed between two scans");
        RG_TRY(rg_reserve(h, pairs, (found + found / 4) * sizeof(uint64_t)));
    }
    const uint64_t count = found;
    // the collision pairs stay on the handle (unsorted, duplicates allowed): rg_rebuild_extras sorts and merges them
    // with the pairs of rows that share sketch hashes (rg_knn_load_sketches)
    RG_TRY(rg_reserve(h, h->coll_pairs, (count ? count : 1) * sizeof(uint64_t)));
    if (count) RG_CUDA(h, cudaMemcpyAsync(h->coll_pairs.p, pairs.p, count * sizeof(uint64_t), cudaMemcpyDeviceToDevice, h->stream));
    h->n_coll_pairs = count;
    return rg_rebuild_extras(h, nullptr, 0);
}

namespace {

__global__ void fill_u32_kernel(uint32_t *p, uint64_t count, uint32_t v) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < count) p[t] = v;
}

// uniq: sorted unique (row << 32 | candidate) with the number of shared sketch hashes per pair
__global__ void extra_csr2_kernel(const uint64_t *__restrict__ uniq, const uint32_t *__restrict__ common, uint64_t count,
                                  uint32_t n, uint64_t *__restrict__ off, uint32_t *__restrict__ idx,
                                  uint32_t *__restrict__ idx_common) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < count) { idx[t] = (uint32_t)uniq[t]; idx_common[t] = common[t]; }
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

// h->extra_off / extra_idx / extra_common <- collision pairs (shared-hash count 0) merged with `sk_pairs` (one entry
// per shared hash of a pair: the count becomes the pair's number of common hashes).  Synchronises the stream.
int32_t rg_rebuild_extras(rg_handle *h, const uint64_t *sk_pairs, uint64_t n_sk) {
    const uint32_t n = (uint32_t)(h->n ? h->n : h->pending_n);
    const uint64_t total = h->n_coll_pairs + n_sk;
    DevBuf &keys = h->ws[7], &vals = h->ws[8], &keys2 = h->ws[9], &vals2 = h->ws[10], &tmp = h->ws[11], &uniq = h->ws[12],
           &sums = h->ws[13], &nsel = h->ws[14];
    auto done = [&](int32_t rc) { return rc; };      // the scratch stays on the handle
#define XT(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define XC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__)); } while (0)
    XT(rg_reserve(h, h->extra_off, ((size_t)n + 1) * sizeof(uint64_t)));
    uint64_t n_uniq = 0;
    if (total) {
        XT(rg_reserve(h, keys, total * 8)); XT(rg_reserve(h, vals, total * 4));
        XT(rg_reserve(h, keys2, total * 8)); XT(rg_reserve(h, vals2, total * 4));
        XT(rg_reserve(h, uniq, total * 8)); XT(rg_reserve(h, sums, total * 4)); XT(rg_reserve(h, nsel, 8));
        if (h->n_coll_pairs) {
            XC(cudaMemcpyAsync(keys.p, h->coll_pairs.p, h->n_coll_pairs * 8, cudaMemcpyDeviceToDevice, h->stream));
            XC(cudaMemsetAsync(vals.p, 0, h->n_coll_pairs * 4, h->stream));
        }
        if (n_sk) {
            XC(cudaMemcpyAsync((uint64_t *)keys.p + h->n_coll_pairs, sk_pairs, n_sk * 8, cudaMemcpyDeviceToDevice, h->stream));
            fill_u32_kernel<<<(unsigned)((n_sk + 255) / 256), 256, 0, h->stream>>>((uint32_t *)vals.p + h->n_coll_pairs, n_sk, 1u);
        }
        size_t b1 = 0, b2 = 0;
        XC(cub::DeviceRadixSort::SortPairs(nullptr, b1, (const uint64_t *)keys.p, (uint64_t *)keys2.p, (const uint32_t *)vals.p,
                                           (uint32_t *)vals2.p, (int64_t)total, 0, 64, h->stream));
        XC(cub::DeviceReduce::ReduceByKey(nullptr, b2, (const uint64_t *)keys2.p, (uint64_t *)uniq.p, (const uint32_t *)vals2.p,
                                          (uint32_t *)sums.p, (uint64_t *)nsel.p, cub::Sum(), (int64_t)total, h->stream));
        XT(rg_reserve(h, tmp, b1 > b2 ? b1 : b2));
        size_t tb = tmp.cap;
        XC(cub::DeviceRadixSort::SortPairs(tmp.p, tb, (const uint64_t *)keys.p, (uint64_t *)keys2.p, (const uint32_t *)vals.p,
                                           (uint32_t *)vals2.p, (int64_t)total, 0, 64, h->stream));
        tb = tmp.cap;
        XC(cub::DeviceReduce::ReduceByKey(tmp.p, tb, (const uint64_t *)keys2.p, (uint64_t *)uniq.p, (const uint32_t *)vals2.p,
                                          (uint32_t *)sums.p, (uint64_t *)nsel.p, cub::Sum(), (int64_t)total, h->stream));
        XC(cudaMemcpyAsync(&n_uniq, nsel.p, sizeof(n_uniq), cudaMemcpyDeviceToHost, h->stream));
        XC(cudaStreamSynchronize(h->stream));
    }
    XT(rg_reserve(h, h->extra_idx, (n_uniq ? n_uniq : 1) * sizeof(uint32_t)));
    XT(rg_reserve(h, h->extra_common, (n_uniq ? n_uniq : 1) * sizeof(uint32_t)));
    const uint64_t threads = n_uniq > (uint64_t)n + 1 ? n_uniq : (uint64_t)n + 1;
    extra_csr2_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, h->stream>>>(
        (const uint64_t *)uniq.p, (const uint32_t *)sums.p, n_uniq, n, (uint64_t *)h->extra_off.p, (uint32_t *)h->extra_idx.p,
        (uint32_t *)h->extra_common.p);
    XC(cudaGetLastError());
    XC(cudaStreamSynchronize(h->stream));
    h->tm.kernel_launches += 2;
    h->n_extra = n_uniq;
    h->has_extra = true;
#undef XT
#undef XC
    return done(RG_OK);
}

// ---------------------------------------------------------------------------------------------------------------
// rg_knn_load_sketches — the MinHash factor of DepthDistance (src/embedding/embedder.rs:133-171):
//     d = metabat x (1 - min_jaccard) as f32,  min_jaccard = common / min(i, j)  (sketch_distances.rs:61-126)
// The sketches are the caller's (finch computes them on the reference side, contig_sketcher.rs:47-92): per row a
// strictly increasing list of u64 hashes.  1 - min_jaccard is exactly 1 unless the two sketches share a hash, and it
// only LOWERS a distance, so the exact top-k of the product metric is the top-k over
//     (top-K' by metabat)  U  (rows that share at least one hash with the query),
// re-evaluated with the product metric.  The second set comes from one sort of all (hash, row) pairs: every run of
// equal hashes yields its row pairs, once per shared hash, so the multiplicity of a pair IS `common`; i and j follow
// from binary searches on the two lists (knn_rescore.cu), no merge walk is needed on the device.
namespace {

__global__ void sketch_rows_kernel(const uint64_t *__restrict__ off, uint32_t n, uint64_t total, uint32_t *__restrict__ rows) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    uint32_t lo = 0, hi = n;                      // last row whose offset is <= t
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo + 1) / 2;
        if (off[mid] <= t) lo = mid; else hi = mid - 1;
    }
    rows[t] = lo;
}

// flags: bit 0 = a list is not strictly increasing
__global__ void sketch_check_kernel(const uint64_t *__restrict__ hashes, const uint32_t *__restrict__ rows, uint64_t total,
                                    uint32_t *__restrict__ flags) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t + 1 >= total) return;
    if (rows[t] == rows[t + 1] && hashes[t] >= hashes[t + 1]) atomicOr(flags, 1u);
}

// sorted by hash: element p pairs its row with every other row of its run (one entry per shared hash and direction)
__global__ void sketch_join_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rows, uint64_t total,
                                   uint64_t *__restrict__ pairs, unsigned long long *__restrict__ n_pairs, uint64_t cap,
                                   uint32_t run_cap, uint32_t *__restrict__ flags) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= total) return;
    const uint64_t key = keys[p];
    const bool left = p > 0 && keys[p - 1] == key, right = p + 1 < total && keys[p + 1] == key;
    if (!left && !right) return;
    uint64_t lo = 0, hi = p;                      // first element of the run
    while (lo < hi) { const uint64_t mid = lo + (hi - lo) / 2; if (keys[mid] < key) lo = mid + 1; else hi = mid; }
    const uint64_t begin = lo;
    lo = p + 1; hi = total;                       // one past the run
    while (lo < hi) { const uint64_t mid = lo + (hi - lo) / 2; if (keys[mid] <= key) lo = mid + 1; else hi = mid; }
    const uint64_t end = lo;
    if (end - begin > run_cap) { atomicOr(flags, 2u); return; }
    const uint64_t me = rows[p];
    for (uint64_t o = begin; o < end; ++o) {
        if (o == p) continue;
        const unsigned long long slot = atomicAdd(n_pairs, 1ull);
        if (slot < cap) pairs[slot] = (me << 32) | rows[o];
        else atomicOr(flags, 4u);
    }
}

// rows with an empty sketch / without a hash below max_hash: raw_distance ends with i == 0 && j == 0 for such a pair
// and returns min_jaccard = 1 (:109-111), i.e. distance 0
__global__ void sketch_empty_kernel(const uint64_t *__restrict__ off, const uint64_t *__restrict__ hashes, uint32_t n,
                                    uint64_t max_hash, uint32_t *__restrict__ empty_rows, uint32_t *__restrict__ nolow_rows,
                                    uint32_t *__restrict__ counts) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint64_t b = off[r], e = off[r + 1];
    if (e == b) empty_rows[atomicAdd(&counts[0], 1u)] = r;
    if (e == b || (max_hash && hashes[b] >= max_hash)) nolow_rows[atomicAdd(&counts[1], 1u)] = r;
}

__global__ void sketch_empty_pairs_kernel(const uint32_t *__restrict__ empty_rows, uint32_t n_empty,
                                          const uint32_t *__restrict__ nolow_rows, uint32_t n_nolow,
                                          uint64_t *__restrict__ pairs, unsigned long long *__restrict__ n_pairs, uint64_t cap,
                                          uint32_t *__restrict__ flags) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (uint64_t)n_empty * n_nolow) return;
    const uint64_t a = empty_rows[t / n_nolow], b = nolow_rows[t % n_nolow];
    if (a == b) return;
    const unsigned long long slot = atomicAdd(n_pairs, 2ull);
    if (slot + 2 <= cap) { pairs[slot] = (a << 32) | b; pairs[slot + 1] = (b << 32) | a; }
    else atomicOr(flags, 4u);
}

}  // namespace

extern "C" int32_t rg_knn_load_sketches(rg_handle *h, const uint64_t *hashes, const uint64_t *offsets, uint64_t n, double scale) {
    if (!h) return RG_ERR_INVALID;
    if (h->n == 0 || !h->has_cov) return rg_fail(h, RG_ERR_STATE, "rg_knn_load_sketches: load the coverage table first (rg_knn_load)");
    if (n != h->n) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: %llu sketches for %llu loaded rows", (unsigned long long)n, (unsigned long long)h->n);
    if (!offsets) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: offsets is NULL");
    if (!(scale >= 0.0) || scale > 1.0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: scale %g outside [0,1]", scale);
    for (uint64_t r = 0; r < n; ++r)
        if (offsets[r + 1] < offsets[r]) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: offsets must be non-decreasing");
    const uint64_t total = offsets[n] - offsets[0];
    // hashes == NULL: the sketches rg_sketch_contigs left on this handle (same total), copied device to device
    const bool resident = total && !hashes && h->skt_hash.p && h->skt_total == total;
    if (total && !hashes && !resident)
        return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: hashes is NULL and the handle holds no sketches of %llu hashes from rg_sketch_contigs",
                       (unsigned long long)total);
    if (offsets[0] != 0) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: offsets[0] must be 0");
    RG_CUDA(h, cudaSetDevice(h->device));
    h->has_sketch = false;
    h->lists_cached = 0;
    // sketch_distances.rs:84: u64::max_value() / scale.recip() as u64
    const uint64_t max_hash = scale > 0.0 ? UINT64_MAX / (uint64_t)(1.0 / scale) : 0;
    const uint64_t cap = 64ull * n + (1ull << 22);
    DevBuf rows, keys2, rows2, tmp, pairs, ctr, er, nr;
    auto done = [&](int32_t rc) {
        for (DevBuf *b : {&rows, &keys2, &rows2, &tmp, &pairs, &ctr, &er, &nr}) rg_release(*b);
        return rc;
    };
#define XT(expr) do { int32_t rc__ = (expr); if (rc__ != RG_OK) return done(rc__); } while (0)
#define XC(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return done(rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__)); } while (0)
    XT(rg_reserve(h, h->sk_hash, (total ? total : 1) * 8));
    XT(rg_reserve(h, h->sk_off, (n + 1) * 8));
    XT(rg_reserve(h, rows, (total ? total : 1) * 4)); XT(rg_reserve(h, keys2, (total ? total : 1) * 8)); XT(rg_reserve(h, rows2, (total ? total : 1) * 4));
    XT(rg_reserve(h, pairs, cap * 8)); XT(rg_reserve(h, ctr, 32)); XT(rg_reserve(h, er, n * 4)); XT(rg_reserve(h, nr, n * 4));
    if (total) XC(cudaMemcpyAsync(h->sk_hash.p, resident ? h->skt_hash.p : (const void *)hashes, total * 8,
                                  resident ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, h->stream));
    XC(cudaMemcpyAsync(h->sk_off.p, offsets, (n + 1) * 8, cudaMemcpyHostToDevice, h->stream));
    XC(cudaMemsetAsync(ctr.p, 0, 32, h->stream));
    h->tm.h2d_bytes += (resident ? 0 : total * 8) + (n + 1) * 8;
    unsigned long long *d_np = (unsigned long long *)ctr.p;
    uint32_t *d_flags = (uint32_t *)ctr.p + 2, *d_cnt = (uint32_t *)ctr.p + 4;
    if (total) {
        const unsigned blocks = (unsigned)((total + 255) / 256);
        sketch_rows_kernel<<<blocks, 256, 0, h->stream>>>((const uint64_t *)h->sk_off.p, (uint32_t)n, total, (uint32_t *)rows.p);
        sketch_check_kernel<<<blocks, 256, 0, h->stream>>>((const uint64_t *)h->sk_hash.p, (const uint32_t *)rows.p, total, d_flags);
        size_t tb = 0;
        XC(cub::DeviceRadixSort::SortPairs(nullptr, tb, (const uint64_t *)h->sk_hash.p, (uint64_t *)keys2.p, (const uint32_t *)rows.p,
                                           (uint32_t *)rows2.p, (int64_t)total, 0, 64, h->stream));
        XT(rg_reserve(h, tmp, tb));
        XC(cub::DeviceRadixSort::SortPairs(tmp.p, tb, (const uint64_t *)h->sk_hash.p, (uint64_t *)keys2.p, (const uint32_t *)rows.p,
                                           (uint32_t *)rows2.p, (int64_t)total, 0, 64, h->stream));
        sketch_join_kernel<<<blocks, 256, 0, h->stream>>>((const uint64_t *)keys2.p, (const uint32_t *)rows2.p, total,
                                                          (uint64_t *)pairs.p, d_np, cap, 4096u, d_flags);
        h->tm.kernel_launches += 3;
    }
    sketch_empty_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>((const uint64_t *)h->sk_off.p, (const uint64_t *)h->sk_hash.p,
                                                                          (uint32_t)n, max_hash, (uint32_t *)er.p, (uint32_t *)nr.p, d_cnt);
    uint32_t cnts[2] = {0, 0};
    XC(cudaMemcpyAsync(cnts, d_cnt, sizeof(cnts), cudaMemcpyDeviceToHost, h->stream));
    XC(cudaStreamSynchronize(h->stream));
    if ((uint64_t)cnts[0] * cnts[1] > (1ull << 24))
        return done(rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: %u empty sketches x %u sketches without a hash below max_hash: "
                                                "every such pair is at distance 0 (sketch_distances.rs:109-111); too many to enumerate", cnts[0], cnts[1]));
    if (cnts[0] && cnts[1]) {
        const uint64_t t = (uint64_t)cnts[0] * cnts[1];
        sketch_empty_pairs_kernel<<<(unsigned)((t + 255) / 256), 256, 0, h->stream>>>((const uint32_t *)er.p, cnts[0], (const uint32_t *)nr.p,
                                                                                    cnts[1], (uint64_t *)pairs.p, d_np, cap, d_flags);
    }
    h->tm.kernel_launches += 2;
    unsigned long long found = 0;
    uint32_t flags = 0;
    XC(cudaMemcpyAsync(&found, d_np, sizeof(found), cudaMemcpyDeviceToHost, h->stream));
    XC(cudaMemcpyAsync(&flags, d_flags, sizeof(flags), cudaMemcpyDeviceToHost, h->stream));
    XC(cudaStreamSynchronize(h->stream));
    if (flags & 1u) return done(rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: a sketch is not strictly increasing (finch's to_vec is sorted and unique)"));
    if (flags & 2u) return done(rg_fail(h, RG_ERR_INVALID, "rg_knn_load_sketches: one hash occurs in more than 4096 sketches"));
    if (flags & 4u) return done(rg_fail(h, RG_ERR_OOM, "rg_knn_load_sketches: more than %llu (row, row, shared hash) triples", (unsigned long long)cap));
    h->sk_scale = scale;
    h->sk_max_hash = max_hash;
    const int32_t rc = rg_rebuild_extras(h, (const uint64_t *)pairs.p, found);
    if (rc == RG_OK) h->has_sketch = true;
#undef XT
#undef XC
    return done(rc);
}
