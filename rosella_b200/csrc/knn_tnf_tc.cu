// knn_tnf_tc.cu — K2b on the 5th-generation tensor cores: fused TNF distance + per-row top-K' selection with the
// 136-term dot product issued as tcgen05.mma (kind::tf32) and accumulated in TMEM.
//
// Metric: KmerCorrelation::distance (src/kmers/kmer_counting.rs:366-414), flight's rho (metrics.py:601-645) and
// flight's tnf_euclidean (metrics.py:64-77); all three need only x~.y~, |x~|^2, |y~|^2 (and the row means), so the
// all-pairs work is one GEMM Q (128 x d) x R^T (d x n) per CTA.  BASELINE.json allows tensor cores for this term
// "only if that term is shown to stay within tolerance": a single TF32 pass (10-bit mantissa) is not, so every
// operand is split as x = hi + lo (hi = tf32(x), lo = tf32(x - hi), knn_transform.cu) and the product is evaluated
// as hi.hi + hi.lo + lo.hi into the same FP32 accumulator ("3xTF32": the dropped lo.lo term is 2^-22 relative,
// i.e. FP32 accuracy; tests/test_gpu_knn.py holds it to the same recall and distance bars as the FP32 kernel).
// Selection only — knn_rescore.cu recomputes the kept candidates in f64.
//
// One CTA = 128 query rows (UMMA M) against all reference panels of 128 rows (UMMA N), K = d in steps of 8:
//   warp 0      TMA producer: query hi/lo planes once, then reference K-chunks through a 4-stage mbarrier ring
//               (1-D cp.async.bulk: a K-chunk of a panel is contiguous in the panel-major layout)
//   warp 1      allocates TMEM (2 x 128 FP32 columns), then one lane issues the MMAs: shared-memory descriptors,
//               K-major, no swizzle (8-row x 16-byte core matrices, SBO = 128 B, LBO = 2048 B); tcgen05.commit
//               releases ring stages and publishes a finished accumulator
//   warps 2..5  epilogue: tcgen05.ld 32 columns at a time (thread = query row, registers = reference rows), a
//               division-free test against the row's current K'-th distance, and a sorted K' list per thread
//               held in REGISTERS (compare-exchange chain), so nothing but the operands lives in shared memory.
// The accumulator is double-buffered: the MMAs of panel p+1 run while the epilogue drains panel p.
#include "knn_tnf_common.cuh"
#include "rg_internal.cuh"

namespace {

constexpr int TQ = 128;              // query rows per CTA (UMMA M)
constexpr int TN = 128;              // reference rows per panel (UMMA N)
constexpr int KC = 4;                // 16-byte K slices per ring stage = 16 elements = two K=8 MMA steps
constexpr int STAGES = 4;
constexpr int THREADS = 192;         // warp 0 producer, warp 1 MMA, warps 2..5 epilogue
constexpr uint32_t SLICE_BYTES = TN * 16;                 // one 16-byte K slice of a 128-row panel: 2048 B
constexpr uint32_t STAGE_BYTES = 2 * KC * SLICE_BYTES;    // hi + lo chunk: 16 KB
constexpr uint32_t TMEM_COLS = 256;  // two FP32 accumulators of 128 columns

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem) {   // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {     // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(TMEM_COLS) : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, 128 x 128 x 8, TF32 inputs, FP32 accumulate
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout): start address >> 4 in [0,14),
// leading byte offset >> 4 in [16,30) (distance between the two 16-byte K slices of one K = 8 step), stride byte
// offset >> 4 in [32,46) (distance between 8-row core matrices), version 1 in [46,48), layout type 0 = no swizzle.
__device__ __forceinline__ uint64_t smem_desc(const void *p) {
    return (uint64_t)((smem_u32(p) >> 4) & 0x3fffu) | ((uint64_t)(SLICE_BYTES >> 4) << 16) |
           ((uint64_t)(128u >> 4) << 32) | (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both
// K-major, N >> 3 in [17,23), M >> 4 in [24,29)
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TQ >> 4) << 24);

// 32 consecutive FP32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// one FP32 column of this thread's TMEM lane
__device__ __forceinline__ float tmem_ld1(uint32_t taddr) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n" : "=r"(r) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return __uint_as_float(r);
}

// ---------------------------------------------------------------- per-thread sorted list in registers
// ld ascending; a new entry goes AFTER equal ones (candidates arrive in increasing index order, so ties stay
// ordered by index).  Fully unrolled compare-exchange chain: no dynamic register indexing.
// A newcomer that beats the K'-th entry lands at a uniformly distributed rank, so the 16-entry blocks ahead of its
// position are skipped (2 of 3 blocks run on average for K' = 48); one copy of each block keeps the code small.
template <int K>
__device__ __forceinline__ void list_insert(float (&ld)[K], uint32_t (&li)[K], float d, uint32_t idx) {
    static_assert(K % 16 == 0, "K' list is walked in blocks of 16");
    bool p = false;
#pragma unroll
    for (int blk = 0; blk < K / 16; ++blk) {
        if (blk == K / 16 - 1 || p || d < ld[blk * 16 + 15]) {
#pragma unroll
            for (int e = blk * 16; e < blk * 16 + 16; ++e) {
                p = p || d < ld[e];          // once displaced, everything behind moves down by one
                const float td = ld[e];
                const uint32_t ti = li[e];
                ld[e] = p ? d : td;
                li[e] = p ? idx : ti;
                d = p ? td : d;
                idx = p ? ti : idx;
            }
        }
    }
}

struct TcSmem {
    uint64_t full[STAGES], empty[STAGES];   // ring: TMA -> MMA, MMA -> TMA
    uint64_t tfull[2], tempty[2];           // accumulators: MMA -> epilogue, epilogue -> MMA
    uint64_t qfull;                         // query planes landed
    uint32_t tmem_base;
    float nbuf[2][TN], mbuf[2][TN];         // norms / means of the reference rows of the panel in each accumulator
};

template <int VARIANT, int K>
__global__ void __launch_bounds__(THREADS, 1)
tnf_select_tc_kernel(const float *__restrict__ hi, const float *__restrict__ lo, const float *__restrict__ norm,
                     const float *__restrict__ mean, uint32_t n, uint32_t D4, uint32_t n_panels, uint32_t first_qpanel,
                     uint32_t q_begin, uint32_t q_count, float *__restrict__ cand_d, uint32_t *__restrict__ cand_i) {
    extern __shared__ __align__(128) uint8_t smem[];
    // [query hi plane | query lo plane | STAGES x (hi chunk | lo chunk) | barriers]
    const uint32_t plane_bytes = D4 * SLICE_BYTES;
    uint8_t *q_hi = smem, *q_lo = smem + plane_bytes;
    uint8_t *ring = smem + 2 * plane_bytes;
    TcSmem &sm = *reinterpret_cast<TcSmem *>(ring + STAGES * STAGE_BYTES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t qpanel = first_qpanel + blockIdx.x;
    const uint32_t n_chunks = (D4 + KC - 1) / KC;            // K chunks per panel (the last may be short)
    const uint64_t panel_floats = (uint64_t)D4 * TN * 4;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&sm.tfull[b], 1); mbar_init(&sm.tempty[b], 4); }
        mbar_init(&sm.qfull, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&sm.tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;

    if (warp == 0) {
        // ===================================================== TMA producer
        if (lane == 0) {
            mbar_expect_tx(&sm.qfull, 2 * plane_bytes);
            tma_bulk_g2s(q_hi, hi + (uint64_t)qpanel * panel_floats, plane_bytes, &sm.qfull);
            tma_bulk_g2s(q_lo, lo + (uint64_t)qpanel * panel_floats, plane_bytes, &sm.qfull);
            uint32_t it = 0;
            for (uint32_t p = 0; p < n_panels; ++p) {
                for (uint32_t c = 0; c < n_chunks; ++c, ++it) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    mbar_wait(&sm.empty[s], ph ^ 1);         // passes at once on the first round
                    const uint32_t slices = min((uint32_t)KC, D4 - c * KC);
                    const uint32_t bytes = slices * SLICE_BYTES;
                    uint8_t *dst = ring + s * STAGE_BYTES;
                    const uint64_t off = (uint64_t)p * panel_floats + (uint64_t)c * KC * TN * 4;
                    mbar_expect_tx(&sm.full[s], 2 * bytes);
                    tma_bulk_g2s(dst, hi + off, bytes, &sm.full[s]);
                    tma_bulk_g2s(dst + KC * SLICE_BYTES, lo + off, bytes, &sm.full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one lane)
        if (lane == 0) {
            mbar_wait(&sm.qfull, 0);
            tc_fence_after();
            uint32_t it = 0;
            for (uint32_t p = 0; p < n_panels; ++p) {
                const uint32_t b = p & 1;
                mbar_wait(&sm.tempty[b], ((p >> 1) & 1) ^ 1);   // the epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem + b * TN;
                for (uint32_t c = 0; c < n_chunks; ++c, ++it) {
                    const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
                    mbar_wait(&sm.full[s], ph);
                    tc_fence_after();
                    const uint32_t slices = min((uint32_t)KC, D4 - c * KC);
                    const uint8_t *r_hi = ring + s * STAGE_BYTES, *r_lo = r_hi + KC * SLICE_BYTES;
                    for (uint32_t ks = 0; ks < slices / 2; ++ks) {          // one K = 8 step = two 16-byte slices
                        const uint32_t qoff = (c * KC + ks * 2) * SLICE_BYTES, roff = ks * 2 * SLICE_BYTES;
                        const uint64_t a_hi = smem_desc(q_hi + qoff), a_lo = smem_desc(q_lo + qoff);
                        const uint64_t b_hi = smem_desc(r_hi + roff), b_lo = smem_desc(r_lo + roff);
                        tc_mma_tf32(d_tmem, a_hi, b_hi, IDESC, (c | ks) != 0);
                        tc_mma_tf32(d_tmem, a_hi, b_lo, IDESC, 1);
                        tc_mma_tf32(d_tmem, a_lo, b_hi, IDESC, 1);
                    }
                    tc_commit(&sm.empty[s]);                 // stage free once these MMAs have read it
                }
                tc_commit(&sm.tfull[b]);                     // accumulator of panel p complete
            }
        }
    } else {
        // ===================================================== epilogue: thread = query row
        const uint32_t quad = warp & 3;                      // the TMEM lane quadrant this warp may read
        const uint32_t row = quad * 32 + lane;
        const uint32_t et = tid - 64;                        // 0..127 among the epilogue threads
        const uint32_t qrow = qpanel * TQ + row;
        const bool q_ok = qrow >= q_begin && qrow < q_begin + q_count && qrow < n;
        const float nq = norm[qrow], mq = VARIANT == METRIC_EUCLID ? mean[qrow] : 0.f;
        const float dim = (float)(D4 * 4);
        float ld[K];
        uint32_t li[K];
#pragma unroll
        for (int e = 0; e < K; ++e) { ld[e] = INFINITY; li[e] = 0xffffffffu; }
        // RUST / FLIGHT: d < thr  <=>  dot + coef (nq + nr) > 0 with coef = thr - 1/2 or (thr - 1)/2, tested as
        // !(fma(coef, nr, dot) < c0), c0 = -coef nq; NaN (flagged rows, inf * 0) and equality (two constant rows)
        // pass too and are settled by the exact form.  EUCLID: d^2 = nq + nr - 2 dot + dim (mu_q - mu_r)^2 < thr.
        float thr = INFINITY, coef = INFINITY, c0 = -coef * nq;
        float n_next = norm[et], m_next = VARIANT == METRIC_EUCLID ? mean[et] : 0.f;   // panel 0, this thread's column
        for (uint32_t p = 0; p < n_panels; ++p) {
            const uint32_t b = p & 1;
            // norms (and means) of the panel's 128 reference rows go through shared memory (broadcast reads);
            // buffer b was last read two panels ago and every epilogue thread has passed a barrier since
            sm.nbuf[b][et] = n_next;
            if (VARIANT == METRIC_EUCLID) sm.mbuf[b][et] = m_next;
            if (p + 1 < n_panels) {
                n_next = norm[(p + 1) * TN + et];
                if (VARIANT == METRIC_EUCLID) m_next = mean[(p + 1) * TN + et];
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            mbar_wait(&sm.tfull[b], (p >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                const uint32_t taddr = tmem + ((quad * 32u) << 16) + b * TN + cc * 32;
                const float *nb = &sm.nbuf[b][cc * 32], *mb = &sm.mbuf[b][cc * 32];
                float acc[32];
                tmem_ld32(taddr, acc);
                // pass 1, branch-free: one bit per column that may beat this row's K'-th distance
                uint32_t mask = 0;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    bool pass;
                    if (VARIANT == METRIC_EUCLID) {
                        const float dm = mq - mb[j];
                        pass = !(thr - (fmaf(-2.0f, acc[j], nq + nb[j]) + dim * dm * dm) <= 0.f);
                    } else {
                        pass = !(fmaf(coef, nb[j], acc[j]) < c0);
                    }
                    if (pass) mask |= 1u << j;
                }
                // pass 2, rare: the warp walks the union of its lanes' columns; the column's accumulator is read
                // back from TMEM (no dynamic register indexing) and the lanes that flagged it insert
                uint32_t wm = __reduce_or_sync(0xffffffffu, mask);
                while (wm) {
                    const int j = __ffs(wm) - 1;
                    wm &= wm - 1;
                    const float dot = tmem_ld1(taddr + j);
                    if ((mask >> j) & 1u) {
                        const uint32_t ridx = p * TN + cc * 32 + j;
                        const float dm = VARIANT == METRIC_EUCLID ? mq - mb[j] : 0.f;
                        float d = tnf_finish<VARIANT>(dot, nq, nb[j], dim * dm * dm);
                        if (ridx >= n || ridx == qrow || !q_ok) d = INFINITY;
                        if (d < thr) {
                            list_insert<K>(ld, li, d, ridx);
                            thr = ld[K - 1];
                            coef = VARIANT == RG_VARIANT_RUST ? thr - 0.5f : 0.5f * (thr - 1.0f);
                            c0 = -coef * nq;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.tempty[b]);
        }
        if (q_ok) {
            const uint64_t o = (uint64_t)(qrow - q_begin) * K;
#pragma unroll
            for (int e = 0; e < K; ++e) { cand_d[o + e] = ld[e]; cand_i[o + e] = li[e]; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem);
}

template <int VARIANT, int K>
int32_t launch_tc(rg_handle *h, uint64_t q_begin, uint64_t q_count) {
    const uint32_t D4 = h->d / 4;
    const size_t smem = (size_t)2 * D4 * SLICE_BYTES + (size_t)STAGES * STAGE_BYTES + sizeof(TcSmem);
    auto kern = tnf_select_tc_kernel<VARIANT, K>;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint32_t first = (uint32_t)(q_begin / TQ), last = (uint32_t)((q_begin + q_count - 1) / TQ);
    kern<<<last - first + 1, THREADS, smem, h->stream>>>(
        (const float *)h->tnf_hi.p, (const float *)h->tnf_lo.p, (const float *)h->tnf_norm.p,
        (const float *)h->tnf_mean.p, (uint32_t)h->n, D4, (uint32_t)h->n_panels, first, (uint32_t)q_begin,
        (uint32_t)q_count, (float *)h->cand_d.p, (uint32_t *)h->cand_i.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

}  // namespace

// The tensor-core kernel keeps its K' list in registers, so it serves K' = 48 / 64 (k <= 48, which covers the
// reference's k = 30); K steps of 8 need d % 8 == 0 and the two query planes must fit beside the ring.
bool rg_tnf_select_tc_supported(const rg_handle *h, uint32_t kprime) {
    const size_t smem = (size_t)2 * (h->d / 4) * SLICE_BYTES + (size_t)STAGES * STAGE_BYTES + sizeof(TcSmem);
    return (kprime == 48 || kprime == 64) && h->d % 8 == 0 && smem <= 227 * 1024;
}

template <int K>
static int32_t launch_tc_k(rg_handle *h, uint32_t which, uint64_t q_begin, uint64_t q_count) {
    if (which == RG_GRAPH_TNF_EUCLID) return launch_tc<METRIC_EUCLID, K>(h, q_begin, q_count);
    return h->variant == RG_VARIANT_FLIGHT ? launch_tc<RG_VARIANT_FLIGHT, K>(h, q_begin, q_count)
                                           : launch_tc<RG_VARIANT_RUST, K>(h, q_begin, q_count);
}

int32_t rg_launch_tnf_select_tc(rg_handle *h, uint32_t which, uint32_t kprime, uint64_t q_begin, uint64_t q_count) {
    return kprime == 48 ? launch_tc_k<48>(h, which, q_begin, q_count) : launch_tc_k<64>(h, which, q_begin, q_count);
}
