// knn_tnf.cu — K2b: fused TNF distance + per-row top-K' selection.
//
// Metric: KmerCorrelation::distance, src/kmers/kmer_counting.rs:366-414 —
//   d = 0.5 - dot(x~, y~) / (|x~|^2 + |y~|^2) on centred rows (guards :383,:385,:410);
// FLIGHT variant: rho = 2d without guards (flight/flight/metrics.py:601-645).
// The centred rows and their squared norms come from knn_transform.cu; here the 136-term dot product
// is an FP32 FFMA register tile (8 query rows x RR reference groups per lane).  BASELINE.json allows
// tensor cores for this term only if shown within tolerance; FP32 CUDA cores are used and the term
// is <= 4% of the path's flops at the headline shape, see DESIGN.md §4.3.
// Selection only — knn_rescore.cu recomputes the kept candidates in f64.
#include "knn_tnf_common.cuh"
#include "rg_internal.cuh"

namespace {

constexpr int QPW = 8;
constexpr int WARPS = 8;
constexpr int QB = QPW * WARPS;   // 64 query rows per CTA

template <int VARIANT, int K, int RR>
__global__ void __launch_bounds__(WARPS * 32, 1)
tnf_select_kernel(const float4 *__restrict__ tnf_t, const float *__restrict__ norm, const float *__restrict__ mean,
                  uint32_t n, uint32_t D4,
                  uint32_t n_tiles, uint32_t q_begin, uint32_t q_count, float *__restrict__ cand_d,
                  uint32_t *__restrict__ cand_i) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tile_f4 = D4 * RG_TILE;            // float4 per 32-row tile
    const uint32_t stage_f4 = tile_f4 * RR;
    float4 *stage[2] = {reinterpret_cast<float4 *>(smem), reinterpret_cast<float4 *>(smem) + stage_f4};
    float4 *qd = stage[1] + stage_f4;                 // [D4][QB]
    float *qn = reinterpret_cast<float *>(qd + D4 * QB);
    float *list_d = qn + QB;
    uint32_t *list_i = reinterpret_cast<uint32_t *>(list_d + QB * K);
    uint64_t *bars = reinterpret_cast<uint64_t *>(list_i + QB * K);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q0 = q_begin + blockIdx.x * QB;
    const uint32_t q_end = q_begin + q_count;

    for (uint32_t i = tid; i < D4 * QB; i += WARPS * 32) {
        const uint32_t j4 = i / QB, r = i % QB;
        const uint32_t row = min(q0 + r, n - 1);
        qd[i] = tnf_t[((uint64_t)(row / RG_TILE) * D4 + j4) * RG_TILE + row % RG_TILE];
    }
    for (uint32_t i = tid; i < QB; i += WARPS * 32) qn[i] = norm[min(q0 + i, n - 1)];
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

    const uint32_t n_stages = (n_tiles + RR - 1) / RR;
    auto issue = [&](uint32_t st) {
        const uint32_t first = st * RR;
        const uint32_t tiles = min((uint32_t)RR, n_tiles - first);
        const uint32_t bytes = tiles * tile_f4 * (uint32_t)sizeof(float4);
        uint64_t *bar = &bars[st & 1];
        mbar_expect_tx(bar, bytes);
        tma_bulk_g2s(stage[st & 1], tnf_t + (uint64_t)first * tile_f4, bytes, bar);
    };
    if (tid == 0) issue(0);

    float thr[QPW], nq[QPW], mq[QPW];
#pragma unroll
    for (int u = 0; u < QPW; ++u) {
        thr[u] = INFINITY;
        nq[u] = qn[warp * QPW + u];
        mq[u] = VARIANT == METRIC_EUCLID ? mean[min(q0 + warp * QPW + u, n - 1)] : 0.f;
    }
    const float dim = (float)(D4 * 4);

    for (uint32_t st = 0; st < n_stages; ++st) {
        __syncthreads();
        if (tid == 0 && st + 1 < n_stages) issue(st + 1);
        mbar_wait(&bars[st & 1], (st >> 1) & 1);
        const float4 *ref = stage[st & 1];
        float acc[QPW][RR];
#pragma unroll
        for (int u = 0; u < QPW; ++u)
#pragma unroll
            for (int g = 0; g < RR; ++g) acc[u][g] = 0.f;
        const float4 *qp = qd + warp * QPW;
#pragma unroll 2
        for (uint32_t j4 = 0; j4 < D4; ++j4) {
            float4 b[RR];
#pragma unroll
            for (int g = 0; g < RR; ++g) b[g] = ref[(g * D4 + j4) * RG_TILE + lane];
#pragma unroll
            for (int u = 0; u < QPW; ++u) {
                const float4 a = qp[j4 * QB + u];
#pragma unroll
                for (int g = 0; g < RR; ++g) {
                    acc[u][g] = fmaf(a.x, b[g].x, acc[u][g]);
                    acc[u][g] = fmaf(a.y, b[g].y, acc[u][g]);
                    acc[u][g] = fmaf(a.z, b[g].z, acc[u][g]);
                    acc[u][g] = fmaf(a.w, b[g].w, acc[u][g]);
                }
            }
        }
#pragma unroll
        for (int g = 0; g < RR; ++g) {
            const uint32_t ridx = (st * RR + g) * RG_TILE + lane;
            const float nr = ridx < n ? norm[ridx] : 0.f;
            const float mr = (VARIANT == METRIC_EUCLID && ridx < n) ? mean[ridx] : 0.f;
#pragma unroll
            for (int u = 0; u < QPW; ++u) {
                const uint32_t qrow = q0 + warp * QPW + u;
                const float dm = mq[u] - mr;
                float d = tnf_finish<VARIANT>(acc[u][g], nq[u], nr, dim * dm * dm);
                if (ridx >= n || ridx == qrow || qrow >= q_end) d = INFINITY;
                topk_offer<K>(list_d + (warp * QPW + u) * K, list_i + (warp * QPW + u) * K, d, ridx, thr[u], lane);
            }
        }
    }
    __syncwarp();
#pragma unroll
    for (int u = 0; u < QPW; ++u) {
        const uint32_t qrow = q0 + warp * QPW + u;
        if (qrow < q_end) {
            const uint64_t o = (uint64_t)(qrow - q_begin) * K;
            for (int e = lane; e < K; e += 32) {
                cand_d[o + e] = list_d[(warp * QPW + u) * K + e];
                cand_i[o + e] = list_i[(warp * QPW + u) * K + e];
            }
        }
    }
}

template <int VARIANT, int K, int RR>
int32_t launch(rg_handle *h, uint64_t q_begin, uint64_t q_count) {
    const uint32_t D4 = h->d / 4;
    const size_t smem = (size_t)2 * RR * D4 * RG_TILE * 16 + (size_t)D4 * QB * 16 + QB * 4 + (size_t)QB * K * 8 + 16;
    if (smem > 227 * 1024)
        return rg_fail(h, RG_ERR_INVALID, "tnf select: d=%u K'=%d needs %zu bytes of shared memory", h->d, K, smem);
    auto kern = tnf_select_kernel<VARIANT, K, RR>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)((q_count + QB - 1) / QB);
    kern<<<grid, WARPS * 32, smem, h->stream>>>((const float4 *)h->tnf_t.p, (const float *)h->tnf_norm.p,
                                                (const float *)h->tnf_mean.p, (uint32_t)h->n, D4, (uint32_t)h->n_tiles, (uint32_t)q_begin,
                                                (uint32_t)q_count, (float *)h->cand_d.p, (uint32_t *)h->cand_i.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

template <int VARIANT>
int32_t launch_k(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    switch (kprime) {
        case 64: return launch<VARIANT, 64, 4>(h, q_begin, q_count);
        case 128: return launch<VARIANT, 128, 2>(h, q_begin, q_count);
        case 256: return launch<VARIANT, 256, 1>(h, q_begin, q_count);
    }
    return rg_fail(h, RG_ERR_INVALID, "tnf select: unsupported K'=%u", kprime);
}

}  // namespace

bool rg_tnf_select_tc_supported(const rg_handle *h, uint32_t kprime);
int32_t rg_ensure_tnf_tiles(rg_handle *h);
int32_t rg_launch_tnf_select_tc(rg_handle *h, uint32_t which, uint32_t kprime, uint64_t q_begin, uint64_t q_count);

// The tcgen05 kernel (knn_tnf_tc.cu) is the product path wherever its shape limits allow (K' = 64, d % 8 == 0);
// larger K' lists do not fit in registers and take the FP32 CUDA-core kernel below.  rg_set_option(h,
// "tnf_select_impl", 1) forces the CUDA-core kernel for A/B measurements.

// K' of the TNF graphs: the tensor-core kernel evaluates the dot product to FP32 accuracy (3xTF32), so a margin of
// 12-16 candidates over k covers the pairs whose FP32 order can differ from the f64 order: k <= 32 takes 48
// candidates and k <= 52 (BASELINE.json configs[4] has k = 50) takes 64; everything else keeps the general K'
// (knn_rescore.cu: pick_kprime) and the FP32 CUDA-core kernel
uint32_t rg_tnf_kprime(const rg_handle *h, uint32_t k, uint32_t general) {
    if (h->opt_tnf_fp32) return general;
    if (k <= 32 && rg_tnf_select_tc_supported(h, 48)) return 48;
    if (k <= 52 && rg_tnf_select_tc_supported(h, 64)) return 64;
    return general;
}

int32_t rg_launch_tnf_select(rg_handle *h, uint32_t which, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    if (!h->opt_tnf_fp32 && rg_tnf_select_tc_supported(h, kprime)) return rg_launch_tnf_select_tc(h, which, kprime, q_begin, q_count);
    RG_TRY(rg_ensure_tnf_tiles(h));
    if (which == RG_GRAPH_TNF_EUCLID) return launch_k<METRIC_EUCLID>(h, kprime, q_begin, q_count);
    return h->variant == RG_VARIANT_FLIGHT ? launch_k<RG_VARIANT_FLIGHT>(h, kprime, q_begin, q_count)
                                           : launch_k<RG_VARIANT_RUST>(h, kprime, q_begin, q_count);
}
