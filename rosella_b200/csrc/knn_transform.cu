// knn_transform.cu — rg_knn_load: copy the caller's f64 tables to HBM and derive the f32
// tile-major operands streamed by the select kernels (knn_depth.cu, knn_tnf.cu).
//
// Device layouts (RG_TILE = 32 rows per tile, one row per lane):
//   tnf_t  float4 [n_tiles][d/4][32]   centred rows x - mean(x)  (kmer_counting.rs:367-377)
//   tnf_hi / tnf_lo  float4 [n_panels][d/4][128]: the same centred rows split as x = hi + lo with hi = tf32(x),
//                                      lo = tf32(x - hi), in 128-row panels.  Inside a panel every 16-byte K slice
//                                      holds its 128 rows contiguously, i.e. 8-row x 16-byte core matrices 128 B
//                                      apart along M/N and 2048 B apart along K: the K-major no-swizzle canonical
//                                      layout tcgen05.mma reads through a shared-memory descriptor (knn_tnf_tc.cu)
//   cov_t  float4 [n_tiles][s][32]     (mean+eps, max(var+eps,1), ln var', +-1/(sqrt(var')*sqrt2))
//                                      sign of .w = "mean+eps > eps" (coverage_calculator.rs:600-619)
// Each tile is one contiguous block so a single TMA bulk copy stages it into shared memory, and
// for a fixed (j/4 or sample) the 32 lanes read 32 consecutive float4 (conflict-free LDS.128).
#include "rg_internal.cuh"

__device__ __forceinline__ float to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

__global__ void tnf_transform_kernel(const double *__restrict__ tnf, uint64_t n, uint32_t d,
                                     float *__restrict__ out, float *__restrict__ norm, float *__restrict__ mean,
                                     float *__restrict__ hi, float *__restrict__ lo) {
    const int lane = threadIdx.x & 31;
    const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const uint64_t rows_padded = (n + RG_PANEL - 1) / RG_PANEL * RG_PANEL;
    if (row >= rows_padded) return;
    const uint64_t panel = row / RG_PANEL;
    const uint32_t pr = (uint32_t)(row % RG_PANEL);
    const uint64_t tile = row / RG_TILE;
    const uint32_t r = (uint32_t)(row % RG_TILE);
    double x[8];
    double sum = 0.0;
    const int iters = (d + 31) / 32;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        uint32_t j = it * 32 + lane;
        x[it] = (it < iters && j < d && row < n) ? tnf[row * d + j] : 0.0;
        sum += x[it];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const double mu = sum / (double)d;
    double nn = 0.0;
    int bad = 0;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        uint32_t j = it * 32 + lane;
        if (it < iters && j < d) {
            double c = row < n ? x[it] - mu : 0.0;
            if (!isfinite(c)) { bad = 1; c = 0.0; }
            nn += c * c;
            const float cf = (float)c;
            if (out) out[((tile * (d / 4) + j / 4) * RG_TILE + r) * 4 + (j & 3)] = cf;
            if (hi) {
                const float h = to_tf32(cf);
                const uint64_t o = ((panel * (d / 4) + j / 4) * RG_PANEL + pr) * 4 + (j & 3);
                hi[o] = h;
                lo[o] = to_tf32(cf - h);
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) nn += __shfl_xor_sync(0xffffffffu, nn, o);
    bad = __any_sync(0xffffffffu, bad);
    // a row with a non-finite centred value (ln 0 in the CLR, kmer_counting.rs:351) is stored as zeros
    // with a NaN norm: every distance to it becomes NaN -> 1.0 exactly as :410 does in f64
    if (lane == 0 && norm) {
        norm[row] = bad ? __int_as_float(0x7fc00000) : (float)nn;
        mean[row] = row < n && isfinite(mu) ? (float)mu : 0.f;
    }
}

__global__ void cov_transform_kernel(const double *__restrict__ cov, uint64_t n, uint32_t s, uint32_t variant,
                                     float4 *__restrict__ out, uint32_t *__restrict__ err_flag) {
    const uint64_t rows_padded = (n + RG_TILE - 1) / RG_TILE * RG_TILE;
    const uint64_t total = rows_padded * s;
    const double eps = variant == RG_VARIANT_FLIGHT ? 1e-6 : 1e-7;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (uint64_t)gridDim.x * blockDim.x) {
        // t enumerates the OUTPUT layout so that stores are coalesced
        const uint32_t r = (uint32_t)(t % RG_TILE);
        const uint64_t ts = t / RG_TILE;
        const uint32_t smp = (uint32_t)(ts % s);
        const uint64_t row = (ts / s) * RG_TILE + r;
        double mean = 0.0, var = 0.0;
        if (row < n) {
            mean = cov[row * 2 * s + 2 * smp];
            var = cov[row * 2 * s + 2 * smp + 1];
            if (!isfinite(mean) || !isfinite(var)) { atomicOr(err_flag, 1u); mean = 0.0; var = 0.0; }
            if (mean < 0.0) { atomicOr(err_flag, 2u); mean = 0.0; }
        }
        const double m = mean + eps;
        double v = var + eps;
        if (v < 1.0) v = 1.0;
        const double inv = 1.0 / (sqrt(v) * 1.4142135623730951);
        const bool present = m > eps;
        out[t] = make_float4((float)m, (float)v, (float)log(v), (float)(present ? inv : -inv));
    }
}

int32_t rg_build_collision_patches(rg_handle *h, uint64_t n, uint32_t s, uint32_t variant);

// tnf_t (FP32 centred rows, tile-major) is read only by the FP32 CUDA-core TNF select (K' > 64, d % 8 != 0 or forced):
// built on first use so that the tensor-core path does not keep 0.5 GB per million rows for nothing
int32_t rg_ensure_tnf_tiles(rg_handle *h) {
    if (h->tnf_t_valid) return RG_OK;
    const uint64_t rows_padded = h->n_panels * RG_PANEL;
    RG_TRY(rg_reserve(h, h->tnf_t, rows_padded * h->d * sizeof(float)));
    const int wpb = 8;
    tnf_transform_kernel<<<(unsigned)((rows_padded + wpb - 1) / wpb), wpb * 32, 0, h->stream>>>(
        (const double *)h->tnf64.p, h->n, h->d, (float *)h->tnf_t.p, nullptr, nullptr, nullptr, nullptr);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    h->tnf_t_valid = true;
    return RG_OK;
}

// rg_knn_load in three phases, shared with the broadcast-fed load of the multi-GPU build (knn_dist.cu):
//   rg_knn_load_begin   validate, reserve every buffer                        (no table data needed)
//   (fill h->tnf64 / h->cov64: H2D copy here, ncclBroadcast from the root rank there)
//   rg_knn_load_finish  pre-transform kernels, input checks, collision patches; publishes the shape on the handle
int32_t rg_knn_load_begin(rg_handle *h, bool has_tnf, bool has_cov, uint64_t n, uint32_t d, uint32_t s, uint32_t variant) {
    if (!h) return RG_ERR_INVALID;
    if (!has_tnf && !has_cov) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: both tables are NULL");
    if (n < 2) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: need at least 2 rows, got %llu", (unsigned long long)n);
    if (n >= 0xffffffffull) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: n must be < 2^32-1");
    if (has_tnf && (d == 0 || d % 4 != 0 || d > 256))
        return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: d=%u must be a multiple of 4 in (0,256] (TNF has 136 bins)", d);
    if (has_cov && (s == 0 || s > 4096)) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: s=%u out of range [1,4096]", s);
    if (variant > RG_VARIANT_FLIGHT) return rg_fail(h, RG_ERR_INVALID, "rg_knn_load: unknown variant %u", variant);
    RG_CUDA(h, cudaSetDevice(h->device));
    h->n = 0;
    h->has_tnf = h->has_cov = false;
    h->lists_cached = 0;
    const uint64_t rows_padded = (n + RG_PANEL - 1) / RG_PANEL * RG_PANEL;   // whole 128-row panels (a multiple of the 32-row tiles)
    RG_TRY(rg_reserve(h, h->err_flag, sizeof(uint32_t)));
    if (has_tnf) {
        RG_TRY(rg_reserve(h, h->tnf64, n * d * sizeof(double)));
        RG_TRY(rg_reserve(h, h->tnf_norm, rows_padded * sizeof(float)));
        RG_TRY(rg_reserve(h, h->tnf_mean, rows_padded * sizeof(float)));
        RG_TRY(rg_reserve(h, h->tnf_hi, rows_padded * d * sizeof(float)));
        RG_TRY(rg_reserve(h, h->tnf_lo, rows_padded * d * sizeof(float)));
    }
    if (has_cov) {
        RG_TRY(rg_reserve(h, h->cov64, n * 2 * s * sizeof(double)));
        RG_TRY(rg_reserve(h, h->cov_t, rows_padded * s * sizeof(float4)));
    }
    return RG_OK;
}

int32_t rg_knn_load_finish(rg_handle *h, bool has_tnf, bool has_cov, uint64_t n, uint32_t d, uint32_t s, uint32_t variant) {
    const uint64_t n_tiles = (n + RG_TILE - 1) / RG_TILE;
    const uint64_t n_panels = (n + RG_PANEL - 1) / RG_PANEL;
    const uint64_t rows_padded = n_panels * RG_PANEL;
    RG_CUDA(h, cudaMemsetAsync(h->err_flag.p, 0, sizeof(uint32_t), h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    if (has_tnf) {
        const int wpb = 8;
        tnf_transform_kernel<<<(unsigned)((rows_padded + wpb - 1) / wpb), wpb * 32, 0, h->stream>>>(
            (const double *)h->tnf64.p, n, d, nullptr, (float *)h->tnf_norm.p, (float *)h->tnf_mean.p,
            (float *)h->tnf_hi.p, (float *)h->tnf_lo.p);
        RG_CUDA(h, cudaGetLastError());
        h->tm.kernel_launches += 1;
    }
    if (has_cov) {
        const uint64_t total = rows_padded * s;
        unsigned blocks = (unsigned)((total + 255) / 256);
        const unsigned cap = (unsigned)h->sm_count * 32u;
        if (blocks > cap) blocks = cap;
        cov_transform_kernel<<<blocks, 256, 0, h->stream>>>((const double *)h->cov64.p, n, s, variant,
                                                            (float4 *)h->cov_t.p, (uint32_t *)h->err_flag.p);
        RG_CUDA(h, cudaGetLastError());
        h->tm.kernel_launches += 1;
    }
    RG_CUDA(h, cudaEventRecord(h->ev[2], h->stream));
    uint32_t flag = 0;
    RG_CUDA(h, cudaMemcpyAsync(&flag, h->err_flag.p, sizeof(flag), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    h->tm.transform_ms = rg_elapsed(h->ev[1], h->ev[2]);
    if (flag & 1u)
        return rg_fail(h, RG_ERR_INVALID,
                       "rg_knn_load: coverage table holds a non-finite value (the reference panics in "
                       "Normal::new, coverage_calculator.rs:639)");
    if (flag & 2u)
        return rg_fail(h, RG_ERR_INVALID,
                       "rg_knn_load: coverage table holds a negative mean (coverage is a read depth; the select phase "
                       "relies on mean + eps >= eps for coverage_calculator.rs:619)");
    h->has_extra = false;
    h->has_sketch = false;
    h->n_extra = 0;
    h->n_coll_pairs = 0;
    h->pending_n = n;
    h->tnf_t_valid = false;
    if (h->cert_ctr.p) RG_CUDA(h, cudaMemsetAsync(h->cert_ctr.p, 0, 8 * sizeof(uint32_t), h->stream));
    if (has_cov && h->opt_certify) {
        RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
        RG_TRY(rg_build_collision_patches(h, n, s, variant));
        RG_CUDA(h, cudaEventRecord(h->ev[2], h->stream));
        RG_CUDA(h, cudaEventSynchronize(h->ev[2]));
        h->tm.transform_ms += rg_elapsed(h->ev[1], h->ev[2]);
    }
    h->n = n; h->d = d; h->s = s; h->variant = variant; h->n_tiles = n_tiles; h->n_panels = n_panels;
    h->has_tnf = has_tnf;
    h->has_cov = has_cov;
    return RG_OK;
}

extern "C" int32_t rg_knn_load(rg_handle *h, const double *tnf, const double *cov, uint64_t n, uint32_t d,
                               uint32_t s, uint32_t variant) {
    RG_TRY(rg_knn_load_begin(h, tnf != nullptr, cov != nullptr, n, d, s, variant));
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    if (tnf) {
        RG_CUDA(h, cudaMemcpyAsync(h->tnf64.p, tnf, n * d * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        h->tm.h2d_bytes += n * d * sizeof(double);
    }
    if (cov) {
        RG_CUDA(h, cudaMemcpyAsync(h->cov64.p, cov, n * 2 * s * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        h->tm.h2d_bytes += n * 2 * s * sizeof(double);
    }
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_CUDA(h, cudaEventSynchronize(h->ev[1]));
    h->tm.h2d_ms = rg_elapsed(h->ev[0], h->ev[1]);
    return rg_knn_load_finish(h, tnf != nullptr, cov != nullptr, n, d, s, variant);
}
