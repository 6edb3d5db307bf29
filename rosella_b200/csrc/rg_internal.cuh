// rg_internal.cuh — handle layout, error plumbing and small device helpers shared by the
// translation units of librosella_gpu.so (sm_100a only; no CPU fallback anywhere).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/rosella_gpu.h"

#define RG_TILE 32u   // rows per tile of the tile-major device layouts (one row per lane)
#define RG_PANEL 128u // rows per panel of the tensor-core operand layout (UMMA M = N = 128)

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct rg_handle {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {};
    std::string err;
    rg_timings tm = {};

    // ---- kNN state (rg_knn_load)
    uint64_t n = 0;
    uint32_t d = 0, s = 0, variant = 0;
    uint64_t n_tiles = 0;
    bool has_tnf = false, has_cov = false;
    DevBuf tnf64, cov64;      // the caller's f64 tables, kept for the f64 re-evaluation
    DevBuf tnf_t;             // f32 centred rows, tile-major [tile][d/4][32] float4 (built on demand: rg_ensure_tnf_tiles)
    bool tnf_t_valid = false;
    DevBuf tnf_norm;          // f32 [n_tiles*32] sum of squares of the centred row (NaN = non-finite row)
    DevBuf tnf_mean;          // f32 [n_tiles*32] mean of the row as passed (Euclidean graph)
    DevBuf tnf_hi, tnf_lo;    // tf32 split of the centred rows, panel-major [panels][d/4][128] float4 (tensor-core select)
    uint64_t n_panels = 0;    // 128-row panels
    DevBuf cov_t;             // float4 (mean, var, ln var, +-1/(sd*sqrt2)) tile-major [tile][s][32]
    DevBuf err_flag;          // u32 device error word (non-finite coverage, ...)
    DevBuf cand_d, cand_i;    // select-phase candidates [rows][kprime]
    DevBuf sym_d, sym_t;      // symmetric depth build: two launches' worth of distances and their transposes (knn_depth_sym.cu)
    cudaStream_t sym_stream[3] = {};   // two alternating evaluation streams + the reducer stream
    cudaEvent_t sym_ev[5] = {};        // evaluated[2], reduced[2], fork
    DevBuf out_i, out_d, out_nv; // device staging of final lists for host-output calls

    // ---- exactness (knn_exact.cu, knn_rescore.cu)
    uint64_t flag_rows = 0;   // rows of the last search (flag_list holds 3 arrays of that length)
    DevBuf flag_list;         // u32 [rows of the search]: local rows whose select margin was not certified
    DevBuf exact_scratch;     // partial lists + arrival counters of flagged rows that are split over several CTAs
    DevBuf cert_ctr;          // u32 [8]: [0] rows in flag_list (reset per search), [1] rows flagged since the load,
                              //          [2] rows re-evaluated exhaustively in f64 since the load
    DevBuf row_flag;          // u8 [n]: 1 = this row always takes the exhaustive f64 pass (dense FP32 collision group)
    DevBuf extra_off;         // u64 [n+1] CSR of the extra candidates of each row ...
    DevBuf extra_idx;         // u32 [...]: rows whose FP32 mean / variance collides with the row's while the f64 values differ
    DevBuf extra_common;      // u32 [...]: number of sketch hashes the pair shares (0: FP32-collision pair only)
    uint64_t n_extra = 0;     // directed extra candidate pairs found by the last load
    bool has_extra = false;
    DevBuf ws[16];            // load-time scratch of the collision scan / extras rebuild, kept between loads
    DevBuf coll_pairs;        // u64 [...]: FP32-collision pairs of the last load (row << 32 | row), unsorted
    uint64_t n_coll_pairs = 0;
    uint64_t pending_n = 0;   // rows of the load in progress (h->n is published at its end)
    // ---- MinHash sketches (rg_knn_load_sketches): the (1 - min_jaccard) factor of DepthDistance
    DevBuf sk_hash, sk_off;   // u64 concatenated sorted hash lists, u64 [n+1] offsets
    double sk_scale = 0.0;
    uint64_t sk_max_hash = 0;
    bool has_sketch = false;

    // ---- MinHash sketching (sketch.cu): the sketches of the last rg_sketch_contigs, concatenated
    DevBuf skt_hash;
    uint64_t skt_total = 0, skt_redone = 0;   // hashes held; records redone with all their hashes (candidate threshold too low)

    // ---- multi-GPU (rg_comm.cu, knn_dist.cu): one handle = one rank of an NCCL communicator
    void *comm = nullptr;             // ncclComm_t (NCCL is dlopen'ed on first use; single-GPU callers never need it)
    int rank = 0, world = 1;
    DevBuf dist_pd, dist_pi;          // this rank's partial lists of ALL rows [n][K'] (symmetric coverage build)
    DevBuf dist_rd, dist_ri;          // all ranks' partial lists of this rank's rows [world][rows][K']
    DevBuf dist_status;               // i32 status word agreed on before a collective is entered
    cudaEvent_t dist_ev[8] = {};
    rg_dist_timings dtm = {};         // last build
    rg_dist_timings dtm_graph[3] = {};   // last build of each graph
    uint32_t dtm_which = 0;

    // ---- kNN lists kept for rg_knn_refilter (graph_ops.cu)
    DevBuf cache_di, cache_dd, cache_ti, cache_td;
    uint32_t cache_k = 0;
    int lists_cached = 0;

    // ---- options (rg_set_option); the library never reads the environment
    int opt_depth_build = 0;          // 0 = automatic, 1 = batch kernel, 2 = symmetric block build
    uint64_t opt_sym_block = 0, opt_sym_chunk = 0;   // rows per block / columns per launch of the symmetric build (0 = default)
    int opt_tnf_fp32 = 0;             // 1 = force the FP32 CUDA-core TNF select
    int opt_exact_compact = 1;        // 0 = the exhaustive pass evaluates passing rows in place (A/B only)
    int opt_certify = 1;              // 0 = skip the margin certificate and the exhaustive pass (A/B measurements only)
    uint64_t opt_sketch_slice = 0;    // bytes of records per slice of rg_sketch_contigs (0 = 256 MiB; test hook)
    uint32_t opt_kprime_depth = 0, opt_kprime_tnf = 0;   // candidate-list override (test hook; 0 = automatic)

    // ---- TNF counting state
    uint64_t tnf_n = 0, tnf_words = 0, tnf_bytes = 0;
    int tnf_mode = 0;         // 0 none, 1 packed resident, 2 ascii resident
    DevBuf codes, valid, woff, ascii, aoff, counts, nkmers, freqs;
    DevBuf tnf_tables;        // 256 B little-endian window -> bin LUT, then 256 B byte-class table
    DevBuf chunk_start;       // u64 [n+1] exclusive scan of chunks per record
    uint64_t tnf_chunks = 0;
    cudaStream_t tnf_stream[2] = {};   // the two lanes of the pipelined rg_tnf_count_ascii
    cudaEvent_t tnf_ev = nullptr;
};

int32_t rg_fail(rg_handle *h, int32_t code, const char *fmt, ...);
int32_t rg_reserve(rg_handle *h, DevBuf &b, size_t bytes);
void rg_release(DevBuf &b);
void rg_set_create_error(const char *msg);

#define RG_CUDA(h, call)                                                                      \
    do {                                                                                      \
        cudaError_t e__ = (call);                                                             \
        if (e__ != cudaSuccess)                                                               \
            return rg_fail((h), e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA,  \
                           "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                           __LINE__);                                                         \
    } while (0)

#define RG_TRY(expr)                 \
    do {                             \
        int32_t rc__ = (expr);       \
        if (rc__ != RG_OK) return rc__; \
    } while (0)

static inline float rg_elapsed(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- mbarrier + TMA bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                             uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------- warp-cooperative top-k list
// A list of K entries (K a multiple of 32) sorted ascending by (dist, arrival order) lives in
// shared memory and is owned by one warp.  Candidates arrive in increasing index order, so
// inserting a tie AFTER the equal entries keeps ties ordered by index.
template <int K>
__device__ __forceinline__ float topk_insert(float *ld, uint32_t *li, float d, uint32_t idx, int lane) {
    constexpr int C = K / 32;
    float pv[C];
    uint32_t pi[C];
    int pos = 0;
#pragma unroll
    for (int c = 0; c < C; ++c) {
        int e = c * 32 + lane;
        float v = ld[e];
        pos += __popc(__ballot_sync(0xffffffffu, v <= d));
        pv[c] = e > 0 ? ld[e - 1] : 0.f;
        pi[c] = e > 0 ? li[e - 1] : 0u;
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < C; ++c) {
        int e = c * 32 + lane;
        if (e == pos) { ld[e] = d; li[e] = idx; }
        else if (e > pos) { ld[e] = pv[c]; li[e] = pi[c]; }
    }
    __syncwarp();
    return ld[K - 1];
}

// Offer the per-lane candidates (d, idx) of one warp step to the list; thr is the current K-th
// distance (warp-uniform).  Lanes are visited in increasing lane (= index) order.
template <int K>
__device__ __forceinline__ void topk_offer(float *ld, uint32_t *li, float d, uint32_t idx, float &thr,
                                           int lane) {
    unsigned m = __ballot_sync(0xffffffffu, d < thr);
    while (m) {
        int src = __ffs(m) - 1;
        m &= m - 1;
        float cd = __shfl_sync(0xffffffffu, d, src);
        uint32_t ci = __shfl_sync(0xffffffffu, idx, src);
        if (cd < thr) thr = topk_insert<K>(ld, li, cd, ci, lane);
    }
}
#endif  // __CUDACC__
