// knn_depth.cu — K2a: fused coverage ("depth") distance + per-row top-K' selection.
//
// Metric: MetabatDistance::distance, src/coverage/coverage_calculator.rs:588-666 (variant RUST) or
// metabat_distance_nn, flight/flight/metrics.py:445-538 (variant FLIGHT), evaluated in FP32 on the
// CUDA cores with the per-contig terms (ln var', 1/(sd*sqrt2)) precomputed by knn_transform.cu.
// This kernel only SELECTS K' >= k candidates per query row; knn_rescore.cu re-evaluates them with
// the reference's f64 arithmetic, so the FP32 error here can only perturb membership at the K'
// boundary, never a reported distance.
//
// Mapping: CTA = 8 warps = 32 query rows (4 per warp, all four in flight for ILP); the 32 lanes of
// a warp hold 32 consecutive reference rows.  Reference tiles (32 rows x S samples x float4) are
// streamed through a 2-stage shared-memory ring by 1-D TMA bulk copies (cp.async.bulk + mbarrier).
// Each query's sorted K' list lives in shared memory and is owned by its warp (rg_internal.cuh).
//
// Algebra used (derivation in DESIGN.md §4.2).  With D = ma - mb, dv = va - vb,
//   tmp = sqrt(va*vb*(D^2 + dv*(ln va - ln vb))), r = 1/dv, the two intersection points k+- of the
//   normal pdfs give  (k+- - ma)/(sa*sqrt2) = (-va*D +- tmp)*r*ia,  (k+- - mb)/(sb*sqrt2) = (-vb*D +- tmp)*r*ib
//   and d = 1/2 |erf(xa+) - erf(xa-) - erf(xb+) + erf(xb-)|: no sorting of k1,k2 and no choice of the
//   narrower distribution is needed because |.| absorbs both sign flips (coverage_calculator.rs:630-651).
//   Equal-variance branch (:620-623): d = 1/2 (+-)(erf(D*ia/2) + erf(D*ib/2)), sign from :638 (RUST keeps the
//   sign, so ln of a negative d turns the whole pair into NaN -> 1.0, :659-661).
#include <cstdlib>

#include "rg_internal.cuh"

namespace {

__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// erf(x) = sign(x) * (1 - 2^(-t*g(t))), t = min(|x|, 4); g = degree-6 fit of -log2(erfc(t))/t on [0,4],
// max abs error 1.6e-7 in FP32 (rosella_b200/tools/fit_erf.py).  6 FFMA + 1 FMUL + 1 MUFU.EX2 + 1 FADD.
__device__ __forceinline__ float erf_fast(float x) {
    const float t = fminf(fabsf(x), 4.0f);
    float p = -1.0021986768753148e-04f;
    p = fmaf(p, t, 4.6156521905287484e-04f);
    p = fmaf(p, t, 2.302256555731618e-03f);
    p = fmaf(p, t, -2.945253550246694e-02f);
    p = fmaf(p, t, 1.489636801568387e-01f);
    p = fmaf(p, t, 9.183286415185535e-01f);
    p = fmaf(p, t, 1.627913732074033f);
    const float e = mufu_ex2(-(p * t));
    return copysignf(1.0f - e, x);
}

// one (query a, reference b) sample: accumulates log2 of 2d and the contributing-sample count
template <int VARIANT>
__device__ __forceinline__ void depth_sample(const float4 a, const float4 b, float &acc, float &cnt) {
    const float ia = fabsf(a.w), ib = fabsf(b.w);
    const bool contrib = (a.w > 0.f || b.w > 0.f) && (a.x != b.x);   // :619
    const float D = a.x - b.x;
    const float dv = a.y - b.y;
    const bool eq = fabsf(dv) < 1e-4f;                                // :620
    const float t = fmaf(dv, a.z - b.z, D * D);
    const float tmp = mufu_sqrt(fmaxf((a.y * b.y) * t, 0.f));         // :625
    const float r = mufu_rcp(dv);
    const float ra = r * ia, rb = r * ib;
    const float A = a.y * D, B = b.y * D;
    const float h = 0.5f * D;
    const float x1 = eq ? h * ia : (tmp - A) * ra;
    const float x2 = eq ? 0.f : (-tmp - A) * ra;
    const float x3 = eq ? -h * ib : (tmp - B) * rb;
    const float x4 = eq ? 0.f : (-tmp - B) * rb;
    float val = (erf_fast(x1) - erf_fast(x2)) - (erf_fast(x3) - erf_fast(x4));
    if (VARIANT == RG_VARIANT_RUST) {
        // general branch: |.| (:650); equal-variance branch keeps the sign of p1.cdf - p2.cdf (:647)
        const float sgn = dv > 0.f ? val : -val;
        val = eq ? sgn : fabsf(val);
    } else {
        val = fminf(fmaxf(fabsf(val), 2e-6f), 2.0f - 2e-6f);          // metrics.py:519,:523 clamp on d = val/2
    }
    const float l = mufu_lg2(val);                                    // NaN for val < 0, -inf for 0
    acc += contrib ? l : 0.f;
    cnt += contrib ? 1.f : 0.f;
}

template <int VARIANT>
__device__ __forceinline__ float depth_finish(float acc, float cnt) {
    // d = exp(mean ln d_s) = 1/2 * 2^(mean log2(2 d_s))   (:656-663)
    float d = 0.5f * mufu_ex2(__fdividef(acc, cnt));
    if (!(cnt > 0.f)) d = 1.0f;                                       // :665
    if (d != d) d = 1.0f;                                             // :660, embedder.rs:166
    return d;
}

constexpr int QPW = 4;          // query rows per warp
constexpr int WARPS = 8;
constexpr int QB = QPW * WARPS; // query rows per CTA (= RG_TILE)

template <int VARIANT, int K>
__global__ void __launch_bounds__(WARPS * 32, 2)
depth_select_kernel(const float4 *__restrict__ cov_t, uint32_t n, uint32_t S, uint32_t n_tiles,
                    uint32_t q_begin, uint32_t q_count, float *__restrict__ cand_d,
                    uint32_t *__restrict__ cand_i) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tile_f4 = S * RG_TILE;
    float4 *stage[2] = {reinterpret_cast<float4 *>(smem), reinterpret_cast<float4 *>(smem) + tile_f4};
    float4 *qd = stage[1] + tile_f4;
    float *list_d = reinterpret_cast<float *>(qd + tile_f4);
    uint32_t *list_i = reinterpret_cast<uint32_t *>(list_d + QB * K);
    uint64_t *bars = reinterpret_cast<uint64_t *>(list_i + QB * K);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q0 = q_begin + blockIdx.x * QB;
    const uint32_t q_end = q_begin + q_count;

    for (uint32_t i = tid; i < tile_f4; i += WARPS * 32) {
        const uint32_t s = i / RG_TILE, r = i % RG_TILE;
        const uint32_t row = min(q0 + r, n - 1);
        qd[i] = cov_t[((uint64_t)(row / RG_TILE) * S + s) * RG_TILE + row % RG_TILE];
    }
    for (uint32_t i = tid; i < QB * K; i += WARPS * 32) {
        list_d[i] = INFINITY;
        list_i[i] = 0xffffffffu;
    }
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const uint32_t tile_bytes = tile_f4 * sizeof(float4);
    if (tid == 0) {
        mbar_expect_tx(&bars[0], tile_bytes);
        tma_bulk_g2s(stage[0], cov_t, tile_bytes, &bars[0]);
    }

    float thr[QPW];
    uint32_t qrow[QPW];
#pragma unroll
    for (int u = 0; u < QPW; ++u) {
        thr[u] = INFINITY;
        qrow[u] = q0 + warp * QPW + u;
    }

    for (uint32_t t = 0; t < n_tiles; ++t) {
        __syncthreads();   // every warp is done with tile t-1, so its stage may be overwritten
        if (tid == 0 && t + 1 < n_tiles) {
            uint64_t *bar = &bars[(t + 1) & 1];
            mbar_expect_tx(bar, tile_bytes);
            tma_bulk_g2s(stage[(t + 1) & 1], cov_t + (uint64_t)(t + 1) * tile_f4, tile_bytes, bar);
        }
        mbar_wait(&bars[t & 1], (t >> 1) & 1);
        const float4 *ref = stage[t & 1];
        float acc[QPW], cnt[QPW];
#pragma unroll
        for (int u = 0; u < QPW; ++u) acc[u] = cnt[u] = 0.f;
        const float4 *qp = qd + warp * QPW;
#pragma unroll 2
        for (uint32_t s = 0; s < S; ++s) {
            const float4 b = ref[s * RG_TILE + lane];
#pragma unroll
            for (int u = 0; u < QPW; ++u) depth_sample<VARIANT>(qp[s * RG_TILE + u], b, acc[u], cnt[u]);
        }
        const uint32_t ridx = t * RG_TILE + lane;
#pragma unroll
        for (int u = 0; u < QPW; ++u) {
            float d = depth_finish<VARIANT>(acc[u], cnt[u]);
            if (ridx >= n || ridx == qrow[u] || qrow[u] >= q_end) d = INFINITY;
            topk_offer<K>(list_d + (warp * QPW + u) * K, list_i + (warp * QPW + u) * K, d, ridx, thr[u], lane);
        }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < QPW; ++u) {
        if (qrow[u] < q_end) {
            const uint64_t o = (uint64_t)(qrow[u] - q_begin) * K;
            for (int e = lane; e < K; e += 32) {
                cand_d[o + e] = list_d[(warp * QPW + u) * K + e];
                cand_i[o + e] = list_i[(warp * QPW + u) * K + e];
            }
        }
    }
}

template <int VARIANT, int K>
int32_t launch(rg_handle *h, uint64_t q_begin, uint64_t q_count) {
    const size_t tile_bytes = (size_t)h->s * RG_TILE * sizeof(float4);
    const size_t smem = 3 * tile_bytes + (size_t)QB * K * 8 + 2 * sizeof(uint64_t);
    if (smem > 227 * 1024)
        return rg_fail(h, RG_ERR_INVALID,
                       "depth select: %u samples with K'=%d need %zu bytes of shared memory (> 227 KB); "
                       "at most ~128 samples are supported in this version", h->s, K, smem);
    auto kern = depth_select_kernel<VARIANT, K>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)((q_count + QB - 1) / QB);
    kern<<<grid, WARPS * 32, smem, h->stream>>>((const float4 *)h->cov_t.p, (uint32_t)h->n, h->s,
                                                (uint32_t)h->n_tiles, (uint32_t)q_begin, (uint32_t)q_count,
                                                (float *)h->cand_d.p, (uint32_t *)h->cand_i.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

template <int VARIANT>
int32_t launch_k(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    switch (kprime) {
        case 64: return launch<VARIANT, 64>(h, q_begin, q_count);
        case 128: return launch<VARIANT, 128>(h, q_begin, q_count);
        case 256: return launch<VARIANT, 256>(h, q_begin, q_count);
    }
    return rg_fail(h, RG_ERR_INVALID, "depth select: unsupported K'=%u", kprime);
}

}  // namespace

int32_t rg_launch_depth_select_v2(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count);

// cand_d / cand_i must hold q_count * kprime entries.  The packed f32x2 kernel (knn_depth2.cu) is the
// product path; RG_DEPTH_IMPL=1 selects the scalar first-generation kernel for A/B measurements.
int32_t rg_launch_depth_select(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    static const bool use_v1 = [] { const char *e = getenv("RG_DEPTH_IMPL"); return e && e[0] == '1'; }();
    if (!use_v1) return rg_launch_depth_select_v2(h, kprime, q_begin, q_count);
    return h->variant == RG_VARIANT_FLIGHT ? launch_k<RG_VARIANT_FLIGHT>(h, kprime, q_begin, q_count)
                                           : launch_k<RG_VARIANT_RUST>(h, kprime, q_begin, q_count);
}
