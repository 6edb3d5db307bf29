// knn_depth_sym.cu — K2a': the coverage graph of ALL rows (rg_knn_build, rg_knn_search over [0, n)) with every
// unordered pair evaluated once.
//
// The coverage metric (coverage_calculator.rs:588-666, metrics.py:445-538) is symmetric up to the sign rule of the
// equal-variance branch (:638/:647, tracked apart — knn_depth_common.cuh SignTrack), and its evaluation is what the
// build costs (knn_depth2.cu is bound by the XU / FMA pipes, not by memory).  The rows are cut into blocks of B; for
// each block pair I <= J
//   1. depth_tile_kernel evaluates the distances of rows I against rows J (same CTA mapping, TMA ring and FP32
//      arithmetic as knn_depth2.cu) and writes them twice: Dm[i][j] as row i sees the pair, Dt[j][i] as row j does;
//   2. topk_reduce_kernel streams Dm into the K' lists of the rows of I, and (J > I) Dt into those of the rows of J.
// (Several column blocks J are evaluated per launch; see build().)
// The lists live in the candidate arrays [n][K'] that knn_rescore.cu reads, so the f64 re-evaluation is unchanged.
// The reduction is bandwidth work (one coalesced pass over 2 B^2 floats per pair, ~4 % of the evaluation time) and
// keeps the evaluation kernel free of cross-row traffic: an earlier attempt that pushed each distance into both rows'
// lists from inside the evaluation kernel was slower than evaluating every pair twice.
//
// Order: candidates reach the list of a row in increasing reference index (blocks I' < I deliver their columns while
// the outer loop is at I', then the row's own pass runs over J = I, I+1, ...), which is the arrival order topk_insert
// needs for the (distance, index) tie rule.
#include "knn_depth_common.cuh"

namespace {

using namespace depthk;

constexpr int QPW = 4;
constexpr int WARPS = 8;
constexpr int QB = QPW * WARPS;
constexpr int PAIRS = QB / 2;
constexpr int RWARPS = 4;      // reducer CTA: 128 threads x 32 registers fit beside two evaluation CTAs on an SM

// rows [q0base + 32 b, +32) against reference tiles [t_begin, t_begin + n_tiles_blk); Dm pitch_m floats per row
// (>= 32 n_tiles_blk), Dt (may be null) pitch_t floats per row (>= 32 gridDim.x)
template <int VARIANT>
__global__ void __launch_bounds__(WARPS * 32, 2)
depth_tile_kernel(const float4 *__restrict__ cov_t, uint32_t n, uint32_t S, uint32_t SC, uint32_t q0base,
                  uint32_t t_begin, uint32_t n_tiles_blk, float *__restrict__ Dm, uint32_t pitch_m,
                  float *__restrict__ Dt, uint32_t pitch_t) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t tile_f4 = S * RG_TILE, stage_f4 = SC * RG_TILE;
    float4 *stage[2] = {reinterpret_cast<float4 *>(smem), reinterpret_cast<float4 *>(smem) + stage_f4};
    float *qd = reinterpret_cast<float *>(stage[1] + stage_f4);
    uint64_t *bars = reinterpret_cast<uint64_t *>(qd + tile_f4 * 4);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t q0 = q0base + blockIdx.x * QB;
    for (uint32_t i = tid; i < tile_f4; i += WARPS * 32) {
        const uint32_t s = i / RG_TILE, r = i % RG_TILE;
        const uint32_t row = min(q0 + r, n - 1);
        const float4 v = cov_t[((uint64_t)(row / RG_TILE) * S + s) * RG_TILE + row % RG_TILE];
        float *dst = qd + ((s * PAIRS + (r >> 1)) * 8) + (r & 1);
        dst[0] = v.x; dst[2] = v.y; dst[4] = v.z; dst[6] = v.w;
    }
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const uint32_t n_chunks = (S + SC - 1) / SC;
    const uint32_t n_stages = n_tiles_blk * n_chunks;
    auto issue = [&](uint32_t g) {
        const uint32_t t = t_begin + g / n_chunks, c = g % n_chunks;
        const uint32_t len = min(SC, S - c * SC);
        const uint32_t bytes = len * RG_TILE * (uint32_t)sizeof(float4);
        uint64_t *bar = &bars[g & 1];
        mbar_expect_tx(bar, bytes);
        tma_bulk_g2s(stage[g & 1], cov_t + ((uint64_t)t * S + (uint64_t)c * SC) * RG_TILE, bytes, bar);
    };
    if (tid == 0) issue(0);
    const ulonglong2 *qp = reinterpret_cast<const ulonglong2 *>(qd) + warp * 4;
    const uint32_t ql = blockIdx.x * QB + warp * QPW;      // first of this warp's 4 rows, local to the block

    float acc[QPW], cnt[QPW];
    SignTrack st[2];
#pragma unroll
    for (int u = 0; u < QPW; ++u) acc[u] = cnt[u] = 0.f;
    st[0] = st[1] = SignTrack{0.f, 0.f, 0.f, 0.f};
    uint32_t t = 0, c = 0;
    for (uint32_t g = 0; g < n_stages; ++g) {
        __syncthreads();
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
                depth_sample2<VARIANT, true>(pk{xy.x}, pk{xy.y}, pk{zw.x}, pk{zw.y}, b, acc[2 * pr], cnt[2 * pr],
                                             acc[2 * pr + 1], cnt[2 * pr + 1], st[pr]);
            }
        }
        if (++c == n_chunks) {
            const uint32_t col = t * RG_TILE + lane;
            const float mnf[QPW] = {st[0].mnf0, st[0].mnf1, st[1].mnf0, st[1].mnf1};
            const float mnb[QPW] = {st[0].mnb0, st[0].mnb1, st[1].mnb0, st[1].mnb1};
            float dk[QPW];
#pragma unroll
            for (int u = 0; u < QPW; ++u) {
                const float dc = depth_finish<VARIANT>(acc[u], cnt[u]);
                float df = dc;
                dk[u] = dc;
                if (VARIANT == RG_VARIANT_RUST) {
                    if (mnf[u] < 0.f) df = 1.0f;          // ln of a negative value -> NaN -> 1.0 (:659-661)
                    if (mnb[u] < 0.f) dk[u] = 1.0f;       // the same pair seen from the reference row
                }
                Dm[(size_t)(ql + u) * pitch_m + col] = df;
                acc[u] = cnt[u] = 0.f;
            }
            if (Dt) *reinterpret_cast<float4 *>(Dt + (size_t)col * pitch_t + ql) = make_float4(dk[0], dk[1], dk[2], dk[3]);
            st[0] = st[1] = SignTrack{0.f, 0.f, 0.f, 0.f};
            c = 0;
            ++t;
        }
    }
}

// One warp per list row: merge the row's pitch-padded line of M (columns = global rows col0, col0+1, ...) into its
// sorted K' list in cand_d / cand_i [n][K].
template <int K>
__global__ void __launch_bounds__(RWARPS * 32)
topk_reduce_kernel(const float *__restrict__ M, uint32_t pitch, uint32_t n_cols, uint32_t row0, uint32_t n_rows,
                   uint32_t col0, uint32_t n, float *__restrict__ cand_d, uint32_t *__restrict__ cand_i) {
    __shared__ float s_d[RWARPS][K];
    __shared__ uint32_t s_i[RWARPS][K];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t r = blockIdx.x * RWARPS + warp;
    if (r >= n_rows || row0 + r >= n) return;
    const uint32_t grow = row0 + r;
    float *ld = s_d[warp];
    uint32_t *li = s_i[warp];
    for (int e = lane; e < K; e += 32) {
        ld[e] = cand_d[(uint64_t)grow * K + e];
        li[e] = cand_i[(uint64_t)grow * K + e];
    }
    __syncwarp();
    float thr = ld[K - 1];
    const float *line = M + (size_t)r * pitch;
    for (uint32_t c = 0; c < n_cols; c += 128) {
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t cc = c + j * 32 + lane;
            v[j] = cc < n_cols ? __ldcs(line + cc) : INFINITY;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t gi = col0 + c + j * 32 + lane;
            const float d = (gi >= n || gi == grow) ? INFINITY : v[j];
            topk_offer<K>(ld, li, d, gi, thr, lane);
        }
    }
    __syncwarp();
    for (int e = lane; e < K; e += 32) {
        cand_d[(uint64_t)grow * K + e] = ld[e];
        cand_i[(uint64_t)grow * K + e] = li[e];
    }
}

__global__ void list_init_kernel(float *cand_d, uint32_t *cand_i, uint64_t count) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
        cand_d[i] = INFINITY;
        cand_i[i] = 0xffffffffu;
    }
}

static uint64_t opt_rows(uint64_t opt, uint64_t dflt) {
    if (opt > 0) dflt = opt;
    return (dflt + QB - 1) / QB * QB;
}

// rows per block: 9472 = 296 x 32, one full wave of 148 SMs x 2 CTAs per evaluation launch
uint32_t block_rows(const rg_handle *h) {
    return (uint32_t)std::min<uint64_t>(opt_rows(h->opt_sym_block, 296ull * QB), h->n_tiles * RG_TILE);
}

// columns per evaluation launch (the width of the distance buffers): 4 blocks
uint32_t chunk_cols(const rg_handle *h, uint32_t B) {
    return (uint32_t)std::min<uint64_t>(opt_rows(h->opt_sym_chunk, 4ull * B), h->n_tiles * RG_TILE);
}

// Block I = rows [i0, i0 + ni) is evaluated against the columns [i0, n) in chunks: its own square in full (both
// orders of a pair inside the block are computed, B / 2N of the work) and everything to the right once.  Dm feeds the
// lists of the block's rows, Dt those of the columns right of the block.
//
// part / n_parts: a process that holds one of n_parts GPUs takes every n_parts-th block in boustrophedon order (blocks
// get cheaper to the right, so consecutive rounds run in opposite directions) and leaves PARTIAL lists for all rows.
template <int VARIANT, int K>
int32_t build(rg_handle *h, float *cd, uint32_t *ci, uint32_t part, uint32_t n_parts) {
    const uint32_t n = (uint32_t)h->n, S = h->s, B = block_rows(h), C = chunk_cols(h, B);
    const size_t tile_bytes = (size_t)S * RG_TILE * sizeof(float4);
    size_t smem = 0;
    const uint32_t sc = pick_chunk(S, tile_bytes + 2 * sizeof(uint64_t), smem);
    if (sc == 0)
        return rg_fail(h, RG_ERR_INVALID, "depth build: %u samples need %zu bytes of shared memory for the query rows alone (> 227 KB)",
                       S, tile_bytes);
    const size_t buf_floats = (size_t)B * C;
    RG_TRY(rg_reserve(h, h->sym_d, 2 * buf_floats * sizeof(float)));
    RG_TRY(rg_reserve(h, h->sym_t, 2 * buf_floats * sizeof(float)));
    for (auto &st : h->sym_stream)
        if (!st) RG_CUDA(h, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto &ev : h->sym_ev)
        if (!ev) RG_CUDA(h, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    auto kern = depth_tile_kernel<VARIANT>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaStream_t red = h->sym_stream[2];
    cudaEvent_t *evaluated = h->sym_ev, *reduced = h->sym_ev + 2, fork = h->sym_ev[4];
    list_init_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(cd, ci, (uint64_t)n * K);
    h->tm.kernel_launches += 1;
    RG_CUDA(h, cudaEventRecord(fork, h->stream));
    for (auto &st : h->sym_stream) RG_CUDA(h, cudaStreamWaitEvent(st, fork, 0));
    // Launch m evaluates into buffer m & 1 on evaluation stream m & 1, so the CTAs of launch m + 1 fill the SMs as
    // those of launch m retire; the reducers of launch m follow on their own stream (in launch order: the lists need
    // their candidates by increasing index) and release the buffer for launch m + 2.
    // an error inside the loop must not return with the three side streams still forked from h->stream (they may be
    // writing the lists and the distance buffers a later call frees): record it, leave the loop, join, then return it
    int32_t status = RG_OK;
#define SYM_CUDA(call)                                                                                         \
    do {                                                                                                       \
        cudaError_t e__ = (call);                                                                              \
        if (e__ != cudaSuccess && status == RG_OK)                                                             \
            status = rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, \
                             cudaGetErrorString(e__), __FILE__, __LINE__);                                     \
    } while (0)
    uint32_t m = 0;
    for (uint32_t i0 = 0, blk = 0; i0 < n; i0 += B, ++blk) {
        const uint32_t round = blk / n_parts, pos = blk % n_parts;
        if (((round & 1) ? n_parts - 1 - pos : pos) != part) continue;
        const uint32_t ni = std::min(B, n - i0), ni_pad = (ni + QB - 1) / QB * QB, i1 = i0 + ni;
        for (uint32_t c0 = i0; c0 < n; c0 += C, ++m) {
            const uint32_t nc = std::min(C, n - c0), nc_pad = (nc + RG_TILE - 1) / RG_TILE * RG_TILE, c1 = c0 + nc;
            const uint32_t jb = std::max(c0, i1);          // first column right of the block
            float *Dm = (float *)h->sym_d.p + (m & 1) * buf_floats, *Dt = (float *)h->sym_t.p + (m & 1) * buf_floats;
            cudaStream_t ev_st = h->sym_stream[m & 1];
            if (m >= 2) SYM_CUDA(cudaStreamWaitEvent(ev_st, reduced[m & 1], 0));
            kern<<<ni_pad / QB, WARPS * 32, smem, ev_st>>>((const float4 *)h->cov_t.p, n, S, sc, i0, c0 / RG_TILE,
                                                           nc_pad / RG_TILE, Dm, nc_pad, jb < c1 ? Dt : nullptr, ni_pad);
            SYM_CUDA(cudaEventRecord(evaluated[m & 1], ev_st));
            SYM_CUDA(cudaStreamWaitEvent(red, evaluated[m & 1], 0));
            topk_reduce_kernel<K><<<(ni + RWARPS - 1) / RWARPS, RWARPS * 32, 0, red>>>(Dm, nc_pad, nc_pad, i0, ni, c0, n, cd, ci);
            h->tm.kernel_launches += 2;
            if (jb < c1) {
                topk_reduce_kernel<K><<<(c1 - jb + RWARPS - 1) / RWARPS, RWARPS * 32, 0, red>>>(
                    Dt + (size_t)(jb - c0) * ni_pad, ni_pad, ni_pad, jb, c1 - jb, i0, n, cd, ci);
                h->tm.kernel_launches += 1;
            }
            SYM_CUDA(cudaEventRecord(reduced[m & 1], red));
        }
        SYM_CUDA(cudaGetLastError());
        if (status != RG_OK) break;
    }
    // join: everything the lists depend on ends on the reducer stream
    if (status != RG_OK) {
        for (auto &st : h->sym_stream) cudaStreamSynchronize(st);
        return status;
    }
#undef SYM_CUDA
    RG_CUDA(h, cudaEventRecord(fork, red));
    RG_CUDA(h, cudaStreamWaitEvent(h->stream, fork, 0));
    return RG_OK;
}

template <int VARIANT>
int32_t build_k(rg_handle *h, uint32_t kprime, float *cd, uint32_t *ci, uint32_t part, uint32_t n_parts) {
    switch (kprime) {
        case 64: return build<VARIANT, 64>(h, cd, ci, part, n_parts);
        case 128: return build<VARIANT, 128>(h, cd, ci, part, n_parts);
        case 256: return build<VARIANT, 256>(h, cd, ci, part, n_parts);
    }
    return rg_fail(h, RG_ERR_INVALID, "depth build: unsupported K'=%u", kprime);
}

// Merge the n_parts partial lists of each row (pd / pi: [n_parts][q_count][K], each sorted by (distance, index), padded
// with (inf, 0xffffffff)) into the row's K best.  No pair is held by two parts, so the final position of a candidate
// is the number of candidates that precede it in (distance, index) order: one binary search per part.
template <int K>
__global__ void __launch_bounds__(RWARPS * 32)
merge_lists_kernel(const float *__restrict__ pd, const uint32_t *__restrict__ pi, uint32_t n_parts, uint32_t q_count,
                   float *__restrict__ cd, uint32_t *__restrict__ ci) {
    extern __shared__ __align__(16) uint8_t msm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t r = blockIdx.x * RWARPS + warp, total = n_parts * K;
    if (r >= q_count) return;
    float *sd = reinterpret_cast<float *>(msm) + (size_t)warp * total;
    uint32_t *si = reinterpret_cast<uint32_t *>(msm) + (size_t)RWARPS * total + (size_t)warp * total;
    for (uint32_t c = lane; c < total; c += 32) {
        const uint32_t p = c / K, e = c % K;
        sd[c] = pd[((size_t)p * q_count + r) * K + e];
        si[c] = pi[((size_t)p * q_count + r) * K + e];
    }
    for (int e = lane; e < K; e += 32) {
        cd[(size_t)r * K + e] = INFINITY;
        ci[(size_t)r * K + e] = 0xffffffffu;
    }
    __syncwarp();
    for (uint32_t c = lane; c < total; c += 32) {
        const float d = sd[c];
        const uint32_t i = si[c];
        if (i == 0xffffffffu) continue;
        uint32_t rank = 0;
        for (uint32_t p = 0; p < n_parts && rank < K; ++p) {
            const float *ld = sd + p * K;
            const uint32_t *li = si + p * K;
            uint32_t lo = 0, hi = K;                  // first entry of part p not before (d, i)
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                const bool before = ld[mid] < d || (ld[mid] == d && li[mid] < i);
                if (before) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < K) {
            cd[(size_t)r * K + rank] = d;
            ci[(size_t)r * K + rank] = i;
        }
    }
}

template <int K>
int32_t merge(rg_handle *h, uint32_t n_parts, const float *pd, const uint32_t *pi, uint64_t q_count) {
    const size_t smem = (size_t)RWARPS * n_parts * K * 8;
    if (smem > 227 * 1024)
        return rg_fail(h, RG_ERR_INVALID, "depth merge: %u parts x K'=%d do not fit in shared memory", n_parts, K);
    auto kern = merge_lists_kernel<K>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)((q_count + RWARPS - 1) / RWARPS), RWARPS * 32, smem, h->stream>>>(
        pd, pi, n_parts, (uint32_t)q_count, (float *)h->cand_d.p, (uint32_t *)h->cand_i.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

}  // namespace

// Whole-graph builds take the symmetric path once there is more than one block of rows (below that it evaluates the
// same square as the batch kernel); rg_set_option "depth_build" = 1 keeps the batch kernel, 2 forces the block build,
// "sym_block" / "sym_chunk" set the block rows and the columns per launch.
bool rg_depth_build_sym_wanted(const rg_handle *h) {
    if (h->opt_depth_build == 1) return false;
    if (h->opt_depth_build == 2) return true;
    return h->n > block_rows(h);
}

// cd / ci: n * kprime entries on the device.  part == 0, n_parts == 1: the whole graph.
int32_t rg_launch_depth_build_sym(rg_handle *h, uint32_t kprime, float *cd, uint32_t *ci, uint32_t part, uint32_t n_parts) {
    return h->variant == RG_VARIANT_FLIGHT ? build_k<RG_VARIANT_FLIGHT>(h, kprime, cd, ci, part, n_parts)
                                           : build_k<RG_VARIANT_RUST>(h, kprime, cd, ci, part, n_parts);
}

// h->cand_d / cand_i (q_count * kprime entries) <- merged partial lists
int32_t rg_launch_depth_merge(rg_handle *h, uint32_t kprime, uint32_t n_parts, const float *pd, const uint32_t *pi,
                              uint64_t q_count) {
    switch (kprime) {
        case 64: return merge<64>(h, n_parts, pd, pi, q_count);
        case 128: return merge<128>(h, n_parts, pd, pi, q_count);
        case 256: return merge<256>(h, n_parts, pd, pi, q_count);
    }
    return rg_fail(h, RG_ERR_INVALID, "depth merge: unsupported K'=%u", kprime);
}
