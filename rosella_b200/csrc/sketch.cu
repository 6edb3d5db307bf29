// sketch.cu — sketch_contigs (src/sketch/contig_sketcher.rs:47-92) on the GPU: finch's scaled MinHash of every record
// (k = 31, scale 0.1, seed 42 by default, src/cli.rs:798-811), the input of SketchDistance / DepthDistance
// (src/sketch/sketch_distances.rs:61-154, src/embedding/embedder.rs:133-171; here rg_knn_load_sketches).
//
// finch ^0.6 is a Cargo dependency that is not under the reference tree; its scheme is restated (oracle/oracle_minhash.c
// states it and says "parity unpinned"): for every window of k A/C/G/T bases of the normalised record, the
// lexicographically smaller of the window and its reverse complement is hashed with MurmurHash3_x64_128(bytes, seed)
// (first half); the sketch is the distinct hashes <= max_hash = u64::MAX / (1/scale), padded with the next smallest
// distinct hashes while it has fewer than kmers_to_sketch = (normalised length x scale) entries; ascending.
//
// Device pipeline per slice of records (256 MiB of bytes by default, so every count fits an int for cub):
//   1. hash kernel, COUNT pass: a thread owns 64 raw bytes of one record, rebuilds the rolling k-mer state from the
//      bytes before them, hashes every k-mer that ENDS in its bytes and counts those with hash <= T (T = 1.5 max_hash);
//   2. exclusive scan of the per-thread counts = where each thread writes: candidates of a record are contiguous, no
//      atomics, no record ids to carry;
//   3. hash kernel, EMIT pass (hashing twice costs less than carrying (hash, record) pairs through a sort);
//   4. cub::DeviceSegmentedSort of the candidates, one segment per record;
//   5. select kernel, one warp per record: distinct hashes, how many lie <= max_hash, the kept length by finch's rule.
//      A record whose candidates do not reach kmers_to_sketch (repeats, long N runs: rare) is flagged and steps 1-5 are
//      repeated for the flagged records alone with T = u64::MAX, so the result never depends on T;
//   6. scan of the kept lengths = the sketch offsets; compact kernel writes the kept distinct hashes.
// The upload of slice i + 1 runs on a copy stream beside steps 1-6 of slice i (two alternating buffers).
// The hash kernel is integer-ALU bound (309 warp instructions per 32 k-mers and the ALU pipe at 82-84 % under ncu, profiles/r2y_*: 12 64-bit multiplies of MurmurHash3 and the
// 2-bit -> ASCII expansion of the canonical k-mer through PRMT), one byte of HBM traffic per k-mer.
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>

#include <algorithm>
#include <vector>

#include "rg_internal.cuh"
#include "tnf_tables.h"

namespace {

constexpr uint32_t SEG = 64;                 // raw bytes per thread
constexpr int HT = 128;                      // threads per CTA of the hash kernel
constexpr uint64_t SLICE_BYTES = 256ull << 20;   // default slice; rg_set_option("sketch_slice_bytes") overrides (test hook)

__device__ __forceinline__ uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
__device__ __forceinline__ uint64_t fmix64(uint64_t k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return k;
}

// 8 bases (2 bits each, first base in the two highest bits of x16) -> their 8 ASCII bytes, first base in the lowest byte
__device__ __forceinline__ uint64_t ascii8(uint32_t x16) {
    auto four = [](uint32_t y) -> uint32_t {                 // y: 8 bits, 4 bases, first base highest
        uint32_t t = (y | (y << 4)) & 0x0F0Fu;
        t = (t | (t << 2)) & 0x3333u;                        // nibble i = base 3 - i
        return __byte_perm(__byte_perm(0x54474341u, 0u, t), 0u, 0x0123u);   // "ACGT" look-up, then first base first
    };
    return (uint64_t)four((x16 >> 8) & 0xffu) | ((uint64_t)four(x16 & 0xffu) << 32);
}

// MurmurHash3_x64_128 (Austin Appleby, public domain) of the k ASCII bytes of a k-mer, first half.
// canon: the k-mer as 2-bit codes, first base most significant, in the low 2k bits.
__device__ __forceinline__ uint64_t murmur_kmer(uint64_t canon, uint32_t k, uint64_t seed) {
    const uint64_t c1 = 0x87c37b91114253d5ull, c2 = 0x4cf5ad432745937full;
    const uint64_t x = canon << (64 - 2 * k);                // first base in bits 62..63
    uint64_t w[4];
    w[0] = ascii8((uint32_t)(x >> 48));
    w[1] = ascii8((uint32_t)(x >> 32) & 0xffffu);
    w[2] = ascii8((uint32_t)(x >> 16) & 0xffffu);
    w[3] = ascii8((uint32_t)x & 0xffffu);
    // bytes past the k-th are zero (the expansion wrote 'A' there)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int have = (int)k - 8 * i;                     // bytes of word i that belong to the k-mer
        w[i] = have >= 8 ? w[i] : (have <= 0 ? 0ull : (w[i] & ((1ull << (8 * have)) - 1ull)));
    }
    uint64_t h1 = seed, h2 = seed;
    auto block = [&](uint64_t k1, uint64_t k2) {
        k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
        h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729ull;
        k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
        h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5ull;
    };
    uint64_t t1, t2;                                         // the tail words (zero-padded: a zero word mixes to zero)
    if (k >= 32) { block(w[0], w[1]); block(w[2], w[3]); t1 = 0; t2 = 0; }
    else if (k >= 16) { block(w[0], w[1]); t1 = w[2]; t2 = w[3]; }
    else { t1 = w[0]; t2 = w[1]; }
    t2 *= c2; t2 = rotl64(t2, 33); t2 *= c1; h2 ^= t2;
    t1 *= c1; t1 = rotl64(t1, 31); t1 *= c2; h1 ^= t1;
    h1 ^= (uint64_t)k; h2 ^= (uint64_t)k;
    h1 += h2; h2 += h1;
    h1 = fmix64(h1); h2 = fmix64(h2);
    return h1 + h2;
}

// record of work item j: rec_ids[j] (a subset pass) or j itself
__device__ __forceinline__ uint64_t rec_of(const uint32_t *__restrict__ rec_ids, uint32_t j) { return rec_ids ? rec_ids[j] : j; }

// seg_start[j] = first thread (segment) of work item j.  COUNT pass: seg_cnt[seg] = candidates of the segment,
// n_kmers / norm_len accumulate per record (may be NULL).  EMIT pass: cand[seg_off[seg] + i] = hash.
template <bool EMIT>
__global__ void __launch_bounds__(HT)
sketch_hash_kernel(const uint8_t *__restrict__ seq, const uint64_t *__restrict__ offsets, const uint32_t *__restrict__ rec_ids,
                   const uint32_t *__restrict__ seg_start, uint32_t m, uint32_t n_seg, uint32_t k, uint64_t seed, uint64_t thr,
                   const uint8_t *__restrict__ cls_table, uint32_t *__restrict__ seg_cnt, const uint32_t *__restrict__ seg_off,
                   uint64_t *__restrict__ cand, unsigned long long *__restrict__ n_kmers, unsigned long long *__restrict__ norm_len) {
    __shared__ uint8_t cls[256];
    for (int i = threadIdx.x; i < 256; i += HT) cls[i] = cls_table[i];
    __syncthreads();
    const uint32_t seg = blockIdx.x * HT + threadIdx.x;
    if (seg >= n_seg) return;
    uint32_t lo = 0, hi = m;                                 // work item: seg_start[lo] <= seg < seg_start[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (seg_start[mid] <= seg) lo = mid; else hi = mid; }
    const uint64_t rec = rec_of(rec_ids, lo);
    const uint64_t rb = offsets[rec], re = offsets[rec + 1];
    const uint64_t p0 = rb + (uint64_t)(seg - seg_start[lo]) * SEG, p1 = min(p0 + SEG, re);
    // the bytes before p0 that can still be inside a window ending at or after p0: walk back over k - 1 bases
    uint64_t q = p0;
    for (uint32_t need = k - 1; need > 0 && q > rb;) {
        const uint8_t c = cls[seq[q - 1]];
        if (c == rg::CLS_INVALID) break;                     // a kept non-base: no window crosses it
        --q;
        if (c < 4) --need;
    }
    const uint64_t kmask = k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1ull);
    const int top = 2 * ((int)k - 1);
    uint64_t fw = 0, rc = 0;
    uint32_t run = 0, emitted = 0, hashed = 0, kept = 0;
    uint64_t *dst = EMIT ? cand + seg_off[seg] : nullptr;
    for (uint64_t p = q; p < p1; ++p) {
        const uint8_t c = cls[seq[p]];
        if (c == rg::CLS_SKIP) continue;                     // removed by normalize(false)
        if (p >= p0) kept += 1;
        if (c >= 4) { run = 0; continue; }
        fw = ((fw << 2) | c) & kmask;
        rc = (rc >> 2) | ((uint64_t)(3u - c) << top);
        if (++run < k || p < p0) continue;
        hashed += 1;
        const uint64_t hsh = murmur_kmer(fw < rc ? fw : rc, k, seed);
        if (hsh <= thr) {
            if (EMIT) dst[emitted] = hsh;
            emitted += 1;
        }
    }
    if (!EMIT) {
        seg_cnt[seg] = emitted;
        if (n_kmers && hashed) atomicAdd(n_kmers + rec, (unsigned long long)hashed);
        if (norm_len && kept) atomicAdd(norm_len + rec, (unsigned long long)kept);
    }
}

// begin[j] = seg_off[seg_start[j]] for j < m, begin[m] = total
__global__ void sketch_begin_kernel(const uint32_t *__restrict__ seg_start, const uint32_t *__restrict__ seg_off, uint32_t m,
                                    uint32_t n_seg, uint32_t total, int *__restrict__ begin) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j > m) return;
    const uint32_t s = j < m ? seg_start[j] : n_seg;
    begin[j] = (int)(s < n_seg ? seg_off[s] : total);
}

// One warp per work item: its sorted candidates -> distinct count, distinct <= max_hash, kept length (finch's rule).
// flag[rec] = 1 when the candidates (all <= thr < u64::MAX) cannot fill kmers_to_sketch although the record has k-mers
// that were not candidates: the record is redone with all its hashes.
__global__ void sketch_select_kernel(const uint64_t *__restrict__ sorted, const int *__restrict__ begin, const uint32_t *__restrict__ rec_ids,
                                     uint32_t m, uint64_t max_hash, uint64_t thr, double scale,
                                     const unsigned long long *__restrict__ norm_len, const unsigned long long *__restrict__ n_kmers,
                                     uint64_t *__restrict__ keep, uint64_t *__restrict__ src_begin, uint32_t *__restrict__ src_cnt,
                                     uint8_t *__restrict__ src_sel, uint8_t sel, uint8_t *__restrict__ flag, uint32_t *__restrict__ n_flagged) {
    const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (j >= m) return;
    const uint64_t rec = rec_of(rec_ids, j);
    const int b = begin[j], e = begin[j + 1];
    uint32_t distinct = 0, below = 0;
    for (int i0 = b; i0 < e; i0 += 32) {
        const int i = i0 + (int)lane;
        const bool in = i < e;
        const uint64_t v = in ? sorted[i] : 0;
        const bool first = in && (i == b || sorted[i - 1] != v);
        distinct += __popc(__ballot_sync(0xffffffffu, first));
        below += __popc(__ballot_sync(0xffffffffu, first && v <= max_hash));
    }
    if (lane != 0) return;
    const uint64_t size = (uint64_t)((double)norm_len[rec] * scale);          // contig_sketcher.rs:65
    const uint64_t kp = below >= size ? below : min(size, (uint64_t)distinct);
    if (thr != ~0ull && kp < size && n_kmers[rec] > (unsigned long long)(e - b)) {
        flag[rec] = 1;
        atomicAdd(n_flagged, 1u);
        return;
    }
    keep[rec] = kp;
    src_begin[rec] = (uint64_t)b;
    src_cnt[rec] = (uint32_t)(e - b);
    src_sel[rec] = sel;
}

// One warp per record: the first keep[rec] distinct hashes of its sorted candidates -> out[off[rec] ..]
__global__ void sketch_compact_kernel(const uint64_t *__restrict__ sorted0, const uint64_t *__restrict__ sorted1,
                                      const uint64_t *__restrict__ src_begin, const uint32_t *__restrict__ src_cnt,
                                      const uint8_t *__restrict__ src_sel, const uint64_t *__restrict__ keep,
                                      const uint64_t *__restrict__ off, uint32_t n, uint64_t *__restrict__ out) {
    const uint32_t rec = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (rec >= n) return;
    const uint64_t kp = keep[rec];
    if (kp == 0) return;
    const uint64_t *src = (src_sel[rec] ? sorted1 : sorted0) + src_begin[rec];
    const uint32_t cnt = src_cnt[rec];
    uint64_t *dst = out + off[rec];
    uint64_t written = 0;
    for (uint32_t i0 = 0; i0 < cnt && written < kp; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool in = i < cnt;
        const uint64_t v = in ? src[i] : 0;
        const bool first = in && (i == 0 || src[i - 1] != v);
        const unsigned bal = __ballot_sync(0xffffffffu, first);
        const uint64_t rank = written + __popc(bal & ((1u << lane) - 1u));
        if (first && rank < kp) dst[rank] = v;
        written += __popc(bal);
    }
}

struct Pass {                                // device buffers of one run of steps 1-4
    DevBuf seg_start, seg_cnt, seg_off, cand, sorted, begin, tmp, rec_ids;
    void release() { for (DevBuf *b : {&seg_start, &seg_cnt, &seg_off, &cand, &sorted, &begin, &tmp, &rec_ids}) rg_release(*b); }
};

}  // namespace

#define SK_CUDA(call)                                                                                              \
    do {                                                                                                           \
        cudaError_t e__ = (call);                                                                                  \
        if (e__ != cudaSuccess) {                                                                                  \
            rc = rg_fail(h, e__ == cudaErrorMemoryAllocation ? RG_ERR_OOM : RG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, \
                         cudaGetErrorString(e__), __FILE__, __LINE__);                                             \
            return rc;                                                                                             \
        }                                                                                                          \
    } while (0)
#define SK_TRY(expr) do { rc = (expr); if (rc != RG_OK) return rc; } while (0)

namespace {

struct SliceCtx {
    rg_handle *h;
    const uint8_t *d_seq;            // the slice's bytes
    const uint64_t *d_off;           // m + 1 offsets relative to the slice
    const uint64_t *off;             // the same on the host
    uint32_t m;
    uint32_t k;
    uint64_t seed;
    const uint8_t *d_cls;
    unsigned long long *d_nk, *d_nl; // per record of the slice
};

// steps 1-4 for the work items `ids` (NULL: all m records of the slice) with candidate threshold thr
// after_small_copies (may be empty) runs once, right after the pass has queued its small host->device copies: the
// host->device copy engine is a FIFO across streams, so the large upload of the NEXT slice must be queued behind them —
// queued before, every slice would wait for the next slice's upload and nothing would overlap (measured).
template <class F>
int32_t run_pass(const SliceCtx &c, Pass &P, const std::vector<uint32_t> *ids, uint64_t thr, bool with_counts, uint32_t *total_out, bool *timed,
                 F after_small_copies) {
    *timed = false;
    rg_handle *h = c.h;
    int32_t rc = RG_OK;
    const uint32_t m = ids ? (uint32_t)ids->size() : c.m;
    std::vector<uint32_t> seg_start((size_t)m + 1);
    uint64_t acc = 0;
    for (uint32_t j = 0; j < m; ++j) {
        const uint64_t r = ids ? (*ids)[j] : j;
        seg_start[j] = (uint32_t)acc;
        acc += (c.off[r + 1] - c.off[r] + SEG - 1) / SEG;
    }
    seg_start[m] = (uint32_t)acc;
    const uint32_t n_seg = (uint32_t)acc;
    *total_out = 0;
    SK_TRY(rg_reserve(h, P.begin, ((size_t)m + 1) * sizeof(int)));
    if (n_seg == 0) {
        SK_CUDA(cudaMemsetAsync(P.begin.p, 0, ((size_t)m + 1) * sizeof(int), h->stream));
        return after_small_copies();
    }
    SK_TRY(rg_reserve(h, P.seg_start, ((size_t)m + 1) * 4));
    SK_TRY(rg_reserve(h, P.seg_cnt, ((size_t)n_seg + 1) * 4));
    SK_TRY(rg_reserve(h, P.seg_off, ((size_t)n_seg + 1) * 4));
    SK_CUDA(cudaMemcpyAsync(P.seg_start.p, seg_start.data(), ((size_t)m + 1) * 4, cudaMemcpyHostToDevice, h->stream));
    const uint32_t *d_ids = nullptr;
    if (ids) {
        SK_TRY(rg_reserve(h, P.rec_ids, (size_t)m * 4));
        SK_CUDA(cudaMemcpyAsync(P.rec_ids.p, ids->data(), (size_t)m * 4, cudaMemcpyHostToDevice, h->stream));
        d_ids = (const uint32_t *)P.rec_ids.p;
    }
    SK_TRY(after_small_copies());
    SK_CUDA(cudaMemsetAsync((uint32_t *)P.seg_cnt.p + n_seg, 0, 4, h->stream));
    const unsigned grid = (n_seg + HT - 1) / HT;
    cudaEventRecord(h->ev[5], h->stream);
    sketch_hash_kernel<false><<<grid, HT, 0, h->stream>>>(c.d_seq, c.d_off, d_ids, (const uint32_t *)P.seg_start.p, m, n_seg, c.k, c.seed,
                                                          thr, c.d_cls, (uint32_t *)P.seg_cnt.p, nullptr, nullptr,
                                                          with_counts ? c.d_nk : nullptr, with_counts ? c.d_nl : nullptr);
    SK_CUDA(cudaGetLastError());
    size_t tb = 0;
    SK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, (const uint32_t *)P.seg_cnt.p, (uint32_t *)P.seg_off.p, (int)(n_seg + 1), h->stream));
    SK_TRY(rg_reserve(h, P.tmp, tb));
    SK_CUDA(cub::DeviceScan::ExclusiveSum(P.tmp.p, tb, (const uint32_t *)P.seg_cnt.p, (uint32_t *)P.seg_off.p, (int)(n_seg + 1), h->stream));
    uint32_t total = 0;
    SK_CUDA(cudaMemcpyAsync(&total, (uint32_t *)P.seg_off.p + n_seg, 4, cudaMemcpyDeviceToHost, h->stream));
    SK_CUDA(cudaStreamSynchronize(h->stream));
    h->tm.kernel_launches += 2;
    if (total > 0x7fffffffu) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: %u candidates in one slice", total);
    sketch_begin_kernel<<<(m + 1 + 255) / 256, 256, 0, h->stream>>>((const uint32_t *)P.seg_start.p, (const uint32_t *)P.seg_off.p, m, n_seg,
                                                                  total, (int *)P.begin.p);
    SK_CUDA(cudaGetLastError());
    h->tm.kernel_launches += 1;
    *total_out = total;
    if (total == 0) return RG_OK;
    SK_TRY(rg_reserve(h, P.cand, (size_t)total * 8));
    SK_TRY(rg_reserve(h, P.sorted, (size_t)total * 8));
    sketch_hash_kernel<true><<<grid, HT, 0, h->stream>>>(c.d_seq, c.d_off, d_ids, (const uint32_t *)P.seg_start.p, m, n_seg, c.k, c.seed, thr,
                                                         c.d_cls, nullptr, (const uint32_t *)P.seg_off.p, (uint64_t *)P.cand.p, nullptr, nullptr);
    SK_CUDA(cudaGetLastError());
    cudaEventRecord(h->ev[6], h->stream);
    size_t sb = 0;
    SK_CUDA(cub::DeviceSegmentedSort::SortKeys(nullptr, sb, (const uint64_t *)P.cand.p, (uint64_t *)P.sorted.p, (int)total, (int)m,
                                               (const int *)P.begin.p, (const int *)P.begin.p + 1, h->stream));
    SK_TRY(rg_reserve(h, P.tmp, sb));
    SK_CUDA(cub::DeviceSegmentedSort::SortKeys(P.tmp.p, sb, (const uint64_t *)P.cand.p, (uint64_t *)P.sorted.p, (int)total, (int)m,
                                               (const int *)P.begin.p, (const int *)P.begin.p + 1, h->stream));
    cudaEventRecord(h->ev[7], h->stream);
    h->tm.kernel_launches += 2;
    *timed = true;
    return RG_OK;
}

// grow b to `bytes`, keeping its first `used` bytes
int32_t grow_preserve(rg_handle *h, DevBuf &b, size_t used, size_t bytes) {
    if (b.cap >= bytes) return RG_OK;
    DevBuf nb;
    int32_t rc = rg_reserve(h, nb, std::max(bytes, b.cap + b.cap / 2));
    if (rc != RG_OK) return rc;
    if (used) {
        const cudaError_t e = cudaMemcpyAsync(nb.p, b.p, used, cudaMemcpyDeviceToDevice, h->stream);
        if (e == cudaSuccess) cudaStreamSynchronize(h->stream);
        else { rg_release(nb); return rg_fail(h, RG_ERR_CUDA, "rg_sketch_contigs: %s", cudaGetErrorString(e)); }
    }
    rg_release(b);
    b = nb;
    return RG_OK;
}

struct Uploader {                 // the copy stream that uploads slice i + 1 while slice i is hashed and sorted
    cudaStream_t cs = nullptr;
    cudaEvent_t t0[2] = {}, done[2] = {};
    int32_t create(rg_handle *h) {
        cudaError_t e = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
        for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
            e = cudaEventCreate(&t0[i]);
            if (e == cudaSuccess) e = cudaEventCreate(&done[i]);
        }
        return e == cudaSuccess ? RG_OK : rg_fail(h, RG_ERR_CUDA, "rg_sketch_contigs: copy stream: %s", cudaGetErrorString(e));
    }
    void destroy() {
        if (cs) cudaStreamSynchronize(cs);
        for (int i = 0; i < 2; ++i) { if (t0[i]) cudaEventDestroy(t0[i]); if (done[i]) cudaEventDestroy(done[i]); }
        if (cs) cudaStreamDestroy(cs);
    }
};

int32_t sketch_impl(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t k, double scale, uint64_t seed,
                    uint64_t *sketch_offsets, uint64_t *n_valid_kmers, Pass &P0, Pass &P1, DevBuf (&B)[13], Uploader &up) {
    int32_t rc = RG_OK;
    DevBuf &d_seq0 = B[0], &d_seq1 = B[12], &d_off = B[1], &d_cls = B[2], &d_nk = B[3], &d_nl = B[4], &d_keep = B[5], &d_sb = B[6], &d_sc = B[7],
           &d_sel = B[8], &d_flag = B[9], &d_soff = B[10], &d_tmp = B[11];
    uint8_t cls[256];
    rg::build_class_table(cls);
    SK_TRY(rg_reserve(h, d_cls, 256 + 16));
    SK_CUDA(cudaMemcpyAsync(d_cls.p, cls, 256, cudaMemcpyHostToDevice, h->stream));
    uint32_t *d_nflag = (uint32_t *)((uint8_t *)d_cls.p + 256);
    const uint64_t max_hash = UINT64_MAX / (uint64_t)(1.0 / scale);
    const uint64_t thr = max_hash > (UINT64_MAX / 3) * 2 ? ~0ull : max_hash + max_hash / 2;
    const uint64_t total_bytes = offsets[n] - offsets[0];
    SK_TRY(rg_reserve(h, h->skt_hash, (size_t)((double)total_bytes * scale * 1.15) * 8 + (1 << 20)));
    h->skt_total = 0;
    sketch_offsets[0] = 0;
    const uint64_t slice_bytes = h->opt_sketch_slice ? h->opt_sketch_slice : SLICE_BYTES;
    std::vector<std::pair<uint64_t, uint64_t>> slices;
    for (uint64_t r0 = 0; r0 < n;) {
        uint64_t r1 = r0 + 1;
        while (r1 < n && offsets[r1 + 1] - offsets[r0] <= slice_bytes && r1 - r0 < 0x7fffff00ull) ++r1;
        if (offsets[r1] - offsets[r0] > 0x7fffffffull)
            return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: record %llu is longer than 2 GiB", (unsigned long long)r0);
        slices.emplace_back(r0, r1);
        r0 = r1;
    }
    // slice si goes to buffer si & 1 on the copy stream; the buffer's previous user (slice si - 2) is complete
    auto upload = [&](size_t si) -> int32_t {
        int32_t rc = RG_OK;
        const uint64_t sb = offsets[slices[si].first], bytes = offsets[slices[si].second] - sb;
        DevBuf &dst = (si & 1) ? d_seq1 : d_seq0;
        SK_TRY(rg_reserve(h, dst, bytes + 16));
        cudaEventRecord(up.t0[si & 1], up.cs);
        if (bytes) SK_CUDA(cudaMemcpyAsync(dst.p, seq + sb, bytes, cudaMemcpyHostToDevice, up.cs));
        cudaEventRecord(up.done[si & 1], up.cs);
        return RG_OK;
    };
    SK_TRY(upload(0));
    for (size_t si = 0; si < slices.size(); ++si) {
        const uint64_t r0 = slices[si].first, r1 = slices[si].second;
        const uint32_t m = (uint32_t)(r1 - r0);
        const uint64_t sb = offsets[r0], bytes = offsets[r1] - sb;
        std::vector<uint64_t> off((size_t)m + 1);
        for (uint32_t j = 0; j <= m; ++j) off[j] = offsets[r0 + j] - sb;
        DevBuf &d_seq = (si & 1) ? d_seq1 : d_seq0;
        SK_TRY(rg_reserve(h, d_off, ((size_t)m + 1) * 8));
        for (DevBuf *b : {&d_nk, &d_nl, &d_keep, &d_sb}) SK_TRY(rg_reserve(h, *b, ((size_t)m + 1) * 8));
        SK_TRY(rg_reserve(h, d_soff, ((size_t)m + 1) * 8));
        SK_TRY(rg_reserve(h, d_sc, (size_t)m * 4));
        SK_TRY(rg_reserve(h, d_sel, m));
        SK_TRY(rg_reserve(h, d_flag, m));
        SK_CUDA(cudaStreamWaitEvent(h->stream, up.done[si & 1], 0));
        SK_CUDA(cudaMemcpyAsync(d_off.p, off.data(), ((size_t)m + 1) * 8, cudaMemcpyHostToDevice, h->stream));
        SK_CUDA(cudaMemsetAsync(d_nk.p, 0, ((size_t)m + 1) * 8, h->stream));
        SK_CUDA(cudaMemsetAsync(d_nl.p, 0, ((size_t)m + 1) * 8, h->stream));
        SK_CUDA(cudaMemsetAsync(d_keep.p, 0, ((size_t)m + 1) * 8, h->stream));
        SK_CUDA(cudaMemsetAsync(d_flag.p, 0, m, h->stream));
        SK_CUDA(cudaMemsetAsync(d_nflag, 0, 4, h->stream));
        h->tm.h2d_bytes += bytes + ((size_t)m + 1) * 8;
        SliceCtx c{h, (const uint8_t *)d_seq.p, (const uint64_t *)d_off.p, off.data(), m, k, seed, (const uint8_t *)d_cls.p,
                   (unsigned long long *)d_nk.p, (unsigned long long *)d_nl.p};
        uint32_t total0 = 0, total1 = 0;
        bool timed0 = false, timed1 = false;
        // the next slice's upload overlaps everything below; it is queued from inside the pass (see run_pass)
        SK_TRY(run_pass(c, P0, nullptr, thr, true, &total0, &timed0,
                        [&]() -> int32_t { return si + 1 < slices.size() ? upload(si + 1) : (int32_t)RG_OK; }));
        const unsigned wgrid = (unsigned)(((uint64_t)m * 32 + 255) / 256);
        sketch_select_kernel<<<wgrid, 256, 0, h->stream>>>((const uint64_t *)P0.sorted.p, (const int *)P0.begin.p, nullptr, m, max_hash, thr, scale,
                                                           (const unsigned long long *)d_nl.p, (const unsigned long long *)d_nk.p,
                                                           (uint64_t *)d_keep.p, (uint64_t *)d_sb.p, (uint32_t *)d_sc.p, (uint8_t *)d_sel.p, 0,
                                                           (uint8_t *)d_flag.p, d_nflag);
        SK_CUDA(cudaGetLastError());
        uint32_t n_flag = 0;
        SK_CUDA(cudaMemcpyAsync(&n_flag, d_nflag, 4, cudaMemcpyDeviceToHost, h->stream));
        SK_CUDA(cudaStreamSynchronize(h->stream));
        h->tm.kernel_launches += 1;
        if (timed0) { h->tm.transform_ms += rg_elapsed(h->ev[5], h->ev[6]); h->tm.rescore_ms += rg_elapsed(h->ev[6], h->ev[7]); }
        if (n_flag) {
            std::vector<uint8_t> flag(m);
            SK_CUDA(cudaMemcpy(flag.data(), d_flag.p, m, cudaMemcpyDeviceToHost));
            std::vector<uint32_t> ids;
            ids.reserve(n_flag);
            for (uint32_t j = 0; j < m; ++j) if (flag[j]) ids.push_back(j);
            SK_TRY(run_pass(c, P1, &ids, ~0ull, false, &total1, &timed1, []() -> int32_t { return RG_OK; }));
            (void)timed1;
            const unsigned fgrid = (unsigned)(((uint64_t)ids.size() * 32 + 255) / 256);
            sketch_select_kernel<<<fgrid, 256, 0, h->stream>>>((const uint64_t *)P1.sorted.p, (const int *)P1.begin.p, (const uint32_t *)P1.rec_ids.p,
                                                               (uint32_t)ids.size(), max_hash, ~0ull, scale, (const unsigned long long *)d_nl.p,
                                                               (const unsigned long long *)d_nk.p, (uint64_t *)d_keep.p, (uint64_t *)d_sb.p,
                                                               (uint32_t *)d_sc.p, (uint8_t *)d_sel.p, 1, (uint8_t *)d_flag.p, d_nflag);
            SK_CUDA(cudaGetLastError());
            h->tm.kernel_launches += 1;
            h->skt_redone += ids.size();
        }
        size_t tb = 0;
        SK_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, (const uint64_t *)d_keep.p, (uint64_t *)d_soff.p, (int)(m + 1), h->stream));
        SK_TRY(rg_reserve(h, d_tmp, tb));
        SK_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tb, (const uint64_t *)d_keep.p, (uint64_t *)d_soff.p, (int)(m + 1), h->stream));
        SK_CUDA(cudaMemcpyAsync(sketch_offsets + r0 + 1, (uint64_t *)d_soff.p + 1, (size_t)m * 8, cudaMemcpyDeviceToHost, h->stream));
        if (n_valid_kmers) SK_CUDA(cudaMemcpyAsync(n_valid_kmers + r0, d_nk.p, (size_t)m * 8, cudaMemcpyDeviceToHost, h->stream));
        SK_CUDA(cudaStreamSynchronize(h->stream));
        const uint64_t base = h->skt_total, slice_total = sketch_offsets[r1];      // still relative to the slice
        for (uint64_t r = r0 + 1; r <= r1; ++r) sketch_offsets[r] += base;
        SK_TRY(grow_preserve(h, h->skt_hash, base * 8, (base + slice_total) * 8 + 16));
        if (slice_total) {
            sketch_compact_kernel<<<wgrid, 256, 0, h->stream>>>((const uint64_t *)P0.sorted.p, (const uint64_t *)P1.sorted.p, (const uint64_t *)d_sb.p,
                                                                (const uint32_t *)d_sc.p, (const uint8_t *)d_sel.p, (const uint64_t *)d_keep.p,
                                                                (const uint64_t *)d_soff.p, m, (uint64_t *)h->skt_hash.p + base);
            SK_CUDA(cudaGetLastError());
            SK_CUDA(cudaStreamSynchronize(h->stream));       // the pass buffers are reused by the next slice
        }
        h->tm.kernel_launches += 2;
        h->tm.d2h_bytes += (size_t)m * 8 * (n_valid_kmers ? 2 : 1);
        h->skt_total = base + slice_total;
        h->tm.h2d_ms += rg_elapsed(up.t0[si & 1], up.done[si & 1]);     // complete: h->stream waited for it and was synchronised
    }
    return RG_OK;
}

}  // namespace

extern "C" int32_t rg_sketch_contigs(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t kmer_length,
                                     double scale, uint64_t seed, uint64_t *sketch_offsets, uint64_t *n_valid_kmers, uint64_t *n_hashes) {
    if (!h) return RG_ERR_INVALID;
    if (!offsets || !sketch_offsets) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: offsets / sketch_offsets is NULL");
    if (kmer_length < 1 || kmer_length > 32) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: kmer_length %u outside [1, 32]", kmer_length);
    if (!(scale > 0.0) || scale > 1.0) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: scale %g outside (0, 1]", scale);
    if (n >= 0xffffffffull) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: n must be < 2^32 - 1");
    for (uint64_t r = 0; r < n; ++r)
        if (offsets[r + 1] < offsets[r]) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: offsets decrease at record %llu", (unsigned long long)r);
    if (n && offsets[n] > offsets[0] && !seq) return rg_fail(h, RG_ERR_INVALID, "rg_sketch_contigs: seq is NULL");
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return rg_fail(h, RG_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    h->skt_total = 0;
    h->skt_redone = 0;
    sketch_offsets[0] = 0;
    if (n_hashes) *n_hashes = 0;
    if (n == 0) return RG_OK;
    Pass P0, P1;
    DevBuf B[13];
    Uploader up;
    int32_t rc = up.create(h);
    h->tm.h2d_ms = h->tm.transform_ms = h->tm.rescore_ms = 0.0;
    cudaEventRecord(h->ev[3], h->stream);
    if (rc == RG_OK) rc = sketch_impl(h, seq, offsets, n, kmer_length, scale, seed, sketch_offsets, n_valid_kmers, P0, P1, B, up);
    cudaEventRecord(h->ev[4], h->stream);
    cudaStreamSynchronize(h->stream);
    up.destroy();
    if (rc == RG_OK) h->tm.select_ms = rg_elapsed(h->ev[3], h->ev[4]);
    P0.release(); P1.release();
    for (DevBuf &b : B) rg_release(b);
    if (rc != RG_OK) { h->skt_total = 0; return rc; }
    if (n_hashes) *n_hashes = h->skt_total;
    return RG_OK;
}

extern "C" int32_t rg_sketch_fetch(rg_handle *h, uint64_t *hashes, uint64_t capacity) {
    if (!h) return RG_ERR_INVALID;
    if (h->skt_total == 0) return RG_OK;
    if (!hashes || capacity < h->skt_total)
        return rg_fail(h, RG_ERR_INVALID, "rg_sketch_fetch: room for %llu hashes, the sketches hold %llu", (unsigned long long)capacity,
                       (unsigned long long)h->skt_total);
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_CUDA(h, cudaMemcpyAsync(hashes, h->skt_hash.p, h->skt_total * 8, cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    h->tm.d2h_bytes += h->skt_total * 8;
    return RG_OK;
}
