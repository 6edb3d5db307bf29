// knn_depth2.cu — K2a: fused coverage ("depth") distance + per-row top-K' selection.
//
// Metric: MetabatDistance::distance, src/coverage/coverage_calculator.rs:588-666 (variant RUST) or
// metabat_distance_nn, flight/flight/metrics.py:445-538 (variant FLIGHT), evaluated in FP32 on the CUDA cores with
// the per-contig terms (ln var', 1/(sd*sqrt2)) precomputed by knn_transform.cu.  This kernel only SELECTS K' >= k
// candidates per query row; knn_rescore.cu re-evaluates them with the reference's f64 arithmetic, so the FP32 error
// here can only perturb membership at the K' boundary, never a reported distance.
//
// Mapping: CTA = 8 warps = 32 query rows; a warp owns 4 query rows as two f32x2-packed pairs (pk.cuh) and its 32
// lanes hold 32 consecutive reference rows.  Reference tiles (32 rows x SC samples x float4) stream through a
// 2-stage shared-memory ring of 1-D TMA bulk copies (cp.async.bulk + mbarrier).  Each query's sorted K' list lives
// in shared memory and is owned by its warp (rg_internal.cuh: topk_offer / topk_insert).
//
// Algebra (derivation in DESIGN.md §4.2).  With D = ma - mb, n = vb - va, r = 1/n,
//   tmp = sqrt(va vb (D^2 + n (ln vb - ln va))), the two intersection points k+- of the normal pdfs give
//   (k+- - ma)/(sa sqrt2) = (+-tmp - va D) r ia,   (k+- - mb)/(sb sqrt2) = (+-tmp - vb D) r ib
//   and d = 1/2 |erf(xa+) - erf(xa-) - erf(xb+) + erf(xb-)|: no sorting of k1, k2 and no choice of the narrower
//   distribution is needed because |.| absorbs both sign flips (coverage_calculator.rs:630-651).
//   Equal-variance branch (:620-623): d = 1/2 (+-)(erf(D ia / 2) + erf(D ib / 2)), sign from :638 (RUST keeps the
//   sign, so ln of a negative d turns the whole pair into NaN -> 1.0, :659-661).
//
// Issue cost, measured on B200 (profiles/r1_microbench_issue_model.json): FFMA 1 cycle/warp, FFMA2 2 pipe cycles
// for twice the work but one dispatch slot, every ALU-pipe op (FSEL, FSETP, FMNMX, LOP3) 2 cycles, MUFU 8 cycles of
// the XU pipe.  Hence
//     - the FMA-pipe work of two query rows is issued as f32x2,
//     - the erf polynomial needs no clamp: a positive leading coefficient makes t*g(t) grow monotonically past the
//       fitted range, so 2^-(t g) underflows to 0,
//     - the equal-variance branch is blended arithmetically (r = 0 and a signed half-difference he select it)
//       instead of four FSELs, sqrt takes |q| as a source modifier, and the accumulators are updated by predicated
//       FADDs instead of FSEL + FADD.
// Per (pair, sample): 7 MUFU (sqrt, rcp, 4 x ex2, lg2) = 56 XU cycles, ~49 FMA-pipe cycles, ~53 dispatch cycles; ncu
// (profiles/r1c_ncu_full_summary.json) shows the XU pipe at 74 % and the FMA pipe at 66 % of peak.
#include <algorithm>

#include "knn_depth_common.cuh"

namespace {

using namespace depthk;

constexpr int QPW = 4;          // query rows per warp = 2 packed pairs
constexpr int WARPS = 8;
constexpr int QB = QPW * WARPS; // 32 query rows per CTA
constexpr int PAIRS = QB / 2;

// Shared memory: 2 reference stages [SC][32] float4 | query pairs [S][PAIRS] 2 x float4
// ({x0,x1,y0,y1},{z0,z1,w0,w1}) | K' lists | 2 mbarriers.  A stage holds SC <= S consecutive samples of one 32-row
// reference tile (contiguous in cov_t), so the ring stays small when S is large and two CTAs still fit per SM;
// the per-row accumulators live in registers across the chunks of a tile.
template <int VARIANT, int K>
__global__ void __launch_bounds__(WARPS * 32, 2)
depth_select2_kernel(const float4 *__restrict__ cov_t, uint32_t n, uint32_t S, uint32_t SC, uint32_t n_tiles,
                     uint32_t q_begin, uint32_t q_count, float *__restrict__ cand_d, uint32_t *__restrict__ cand_i) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tile_f4 = S * RG_TILE, stage_f4 = SC * RG_TILE;
    float4 *stage[2] = {reinterpret_cast<float4 *>(smem), reinterpret_cast<float4 *>(smem) + stage_f4};
    float *qd = reinterpret_cast<float *>(stage[1] + stage_f4);
    float *list_d = qd + tile_f4 * 4;
    uint32_t *list_i = reinterpret_cast<uint32_t *>(list_d + QB * K);
    uint64_t *bars = reinterpret_cast<uint64_t *>(list_i + QB * K);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q0 = q_begin + blockIdx.x * QB;
    const uint32_t q_end = q_begin + q_count;

    for (uint32_t i = tid; i < tile_f4; i += WARPS * 32) {
        const uint32_t s = i / RG_TILE, r = i % RG_TILE;
        const uint32_t row = min(q0 + r, n - 1);
        const float4 v = cov_t[((uint64_t)(row / RG_TILE) * S + s) * RG_TILE + row % RG_TILE];
        float *dst = qd + ((s * PAIRS + (r >> 1)) * 8) + (r & 1);
        dst[0] = v.x; dst[2] = v.y; dst[4] = v.z; dst[6] = v.w;
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
    const uint32_t n_chunks = (S + SC - 1) / SC;
    const uint32_t n_stages = n_tiles * n_chunks;
    // stage g = (tile g / n_chunks, samples [c * SC, min(S, (c + 1) * SC)) with c = g % n_chunks)
    auto issue = [&](uint32_t g) {
        const uint32_t t = g / n_chunks, c = g - t * n_chunks;
        const uint32_t len = min(SC, S - c * SC);
        const uint32_t bytes = len * RG_TILE * (uint32_t)sizeof(float4);
        uint64_t *bar = &bars[g & 1];
        mbar_expect_tx(bar, bytes);
        tma_bulk_g2s(stage[g & 1], cov_t + ((uint64_t)t * S + (uint64_t)c * SC) * RG_TILE, bytes, bar);
    };
    if (tid == 0) issue(0);
    float thr[QPW];
    uint32_t qrow[QPW];
#pragma unroll
    for (int u = 0; u < QPW; ++u) {
        thr[u] = INFINITY;
        qrow[u] = q0 + warp * QPW + u;
    }
    const ulonglong2 *qp = reinterpret_cast<const ulonglong2 *>(qd) + warp * 4;   // 2 pairs x 2 ulonglong2

    float acc[QPW], cnt[QPW];
#pragma unroll
    for (int u = 0; u < QPW; ++u) acc[u] = cnt[u] = 0.f;
    uint32_t t = 0, c = 0;
    for (uint32_t g = 0; g < n_stages; ++g) {
        __syncthreads();   // every warp is done with stage g-1, so its buffer may be overwritten
        if (tid == 0 && g + 1 < n_stages) issue(g + 1);
        mbar_wait(&bars[g & 1], (g >> 1) & 1);
        const float4 *ref = stage[g & 1];
        const uint32_t s0 = c * SC, len = min(SC, S - s0);
        const ulonglong2 *qs = qp + (size_t)s0 * (PAIRS * 2);
#pragma unroll 2
        for (uint32_t s = 0; s < len; ++s) {
            const float4 b = ref[s * RG_TILE + lane];
#pragma unroll
            for (int pr = 0; pr < 2; ++pr) {
                const ulonglong2 xy = qs[s * (PAIRS * 2) + pr * 2], zw = qs[s * (PAIRS * 2) + pr * 2 + 1];
                SignTrack unused;
                depth_sample2<VARIANT, false>(pk{xy.x}, pk{xy.y}, pk{zw.x}, pk{zw.y}, b, acc[2 * pr], cnt[2 * pr],
                                              acc[2 * pr + 1], cnt[2 * pr + 1], unused);
            }
        }
        if (++c == n_chunks) {          // last chunk of tile t: finish the 4 x 32 distances and offer them
            const uint32_t ridx = t * RG_TILE + lane;
#pragma unroll
            for (int u = 0; u < QPW; ++u) {
                float d = depth_finish<VARIANT>(acc[u], cnt[u]);
                if (ridx >= n || ridx == qrow[u] || qrow[u] >= q_end) d = INFINITY;
                topk_offer<K>(list_d + (warp * QPW + u) * K, list_i + (warp * QPW + u) * K, d, ridx, thr[u], lane);
                acc[u] = cnt[u] = 0.f;
            }
            c = 0;
            ++t;
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
    const size_t fixed = tile_bytes + (size_t)QB * K * 8 + 2 * sizeof(uint64_t);   // query pairs + lists + barriers
    size_t smem = 0;
    const uint32_t sc = pick_chunk(h->s, fixed, smem);
    if (sc == 0)
        return rg_fail(h, RG_ERR_INVALID,
                       "depth select: %u samples with K'=%d need %zu bytes of shared memory for the query rows and "
                       "lists alone (> 227 KB)", h->s, K, fixed);
    auto kern = depth_select2_kernel<VARIANT, K>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)((q_count + QB - 1) / QB);
    kern<<<grid, WARPS * 32, smem, h->stream>>>((const float4 *)h->cov_t.p, (uint32_t)h->n, h->s, sc,
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

// cand_d / cand_i must hold q_count * kprime entries
int32_t rg_launch_depth_select(rg_handle *h, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    return h->variant == RG_VARIANT_FLIGHT ? launch_k<RG_VARIANT_FLIGHT>(h, kprime, q_begin, q_count)
                                           : launch_k<RG_VARIANT_RUST>(h, kprime, q_begin, q_count);
}
