// tnf_count.cu — K1: tetranucleotide counting kernels.
//
// Replaces the per-record loop of KmerCounter::run (src/kmers/kmer_counting.rs:64-109): for every
// window of 4 consecutive A/C/G/T bases of the normalised record, increment the bin of the
// lexicographically smaller of (4-mer, reverse complement) (:82-101); n_kmers = counted windows;
// freq = count / n_kmers in f64 (:104-107).  Integer work, bit-exact by construction.
//
//   tnf_count_packed_kernel  streams the 2-bit + validity-bit layout (0.375 B/base): one 32-base
//                            word per thread, coalesced 8 B + 4 B loads, 256->136 LUT and
//                            warp-private 136-bin histograms in shared memory, one flush of
//                            <= 136 global atomics per 8192-base chunk.
//   tnf_count_ascii_kernel   same result straight from the raw record bytes (1 B/base) with
//                            needletail's normalize(false) folded in as a byte-class table.
#include "rg_internal.cuh"
#include "tnf_tables.h"

namespace {

constexpr int THREADS = 256;
constexpr int WARPS = THREADS / 32;
constexpr uint32_t CHUNK_WORDS = THREADS;          // one packed word (32 bases) per thread
constexpr uint32_t BYTES_PER_THREAD = 32;
constexpr uint32_t CHUNK_BYTES = THREADS * BYTES_PER_THREAD;

// largest r in [0, n) with start[r] <= chunk  (start is the exclusive scan of chunks per record)
__device__ __forceinline__ uint64_t find_record(const uint64_t *__restrict__ start, uint64_t n, uint64_t chunk) {
    uint64_t lo = 0, hi = n;   // invariant: start[lo] <= chunk < start[hi]
    while (hi - lo > 1) {
        const uint64_t mid = (lo + hi) >> 1;
        if (start[mid] <= chunk) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ void flush_hist(uint32_t (*hist)[RG_NBINS], uint32_t *__restrict__ counts, uint64_t rec) {
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < RG_NBINS; b += THREADS) {
        uint32_t s = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) { s += hist[w][b]; hist[w][b] = 0; }
        if (s) atomicAdd(&counts[rec * RG_NBINS + b], s);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(THREADS)
tnf_count_packed_kernel(const uint64_t *__restrict__ codes, const uint32_t *__restrict__ valid,
                        const uint64_t *__restrict__ woff, const uint64_t *__restrict__ chunk_start, uint64_t n,
                        uint64_t n_chunks, const uint8_t *__restrict__ tables, uint32_t *__restrict__ counts) {
    __shared__ uint32_t hist[WARPS][RG_NBINS];
    __shared__ uint8_t lut[256];
    const int tid = threadIdx.x, warp = tid >> 5;
    lut[tid] = tables[tid];
    for (uint32_t i = tid; i < WARPS * RG_NBINS; i += THREADS) (&hist[0][0])[i] = 0;
    __syncthreads();
    for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const uint64_t rec = find_record(chunk_start, n, chunk);
        const uint64_t w_in_rec = (chunk - chunk_start[rec]) * CHUNK_WORDS + tid;
        const uint64_t w = woff[rec] + w_in_rec;
        if (w < woff[rec + 1]) {
            const uint64_t c = codes[w];
            const uint32_t v = valid[w];
            uint32_t pc = 0, pv = 0;   // last 3 bases of the previous word of the same record
            if (w_in_rec > 0) { pc = (uint32_t)(codes[w - 1] >> 58); pv = valid[w - 1] >> 29; }
            // 35-base extended word: bits [0,6) = previous 3 bases, then the 32 bases of this word
            const uint32_t e0 = ((uint32_t)c << 6) | pc;
            const uint32_t e1 = (uint32_t)(c >> 26);
            const uint32_t e2 = (uint32_t)(c >> 58);
            const uint64_t vext = ((uint64_t)v << 3) | pv;
#pragma unroll
            for (int j = 0; j < 32; ++j) {          // window ending at base j of this word
                const int o = 2 * j;
                uint32_t win;
                if (o < 32) win = __funnelshift_r(e0, e1, o);
                else win = __funnelshift_r(e1, e2, o - 32);
                win &= 0xffu;
                if (((uint32_t)(vext >> j) & 0xfu) == 0xfu) atomicAdd(&hist[warp][lut[win]], 1u);
            }
        }
        flush_hist(hist, counts, rec);
    }
}

__global__ void __launch_bounds__(THREADS)
tnf_count_ascii_kernel(const uint8_t *__restrict__ seq, const uint64_t *__restrict__ offsets,
                       const uint64_t *__restrict__ chunk_start, uint64_t n, uint64_t n_chunks,
                       const uint8_t *__restrict__ tables, uint32_t *__restrict__ counts) {
    __shared__ uint32_t hist[WARPS][RG_NBINS];
    __shared__ uint8_t lut[256];
    __shared__ uint8_t cls[256];
    const int tid = threadIdx.x, warp = tid >> 5;
    lut[tid] = tables[tid];
    cls[tid] = tables[256 + tid];
    for (uint32_t i = tid; i < WARPS * RG_NBINS; i += THREADS) (&hist[0][0])[i] = 0;
    __syncthreads();
    for (uint64_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const uint64_t rec = find_record(chunk_start, n, chunk);
        const uint64_t rb = offsets[rec], re = offsets[rec + 1];
        const uint64_t b0 = rb + (chunk - chunk_start[rec]) * CHUNK_BYTES + (uint64_t)tid * BYTES_PER_THREAD;
        if (b0 < re) {
            // rebuild the rolling state from the (up to) 3 kept bytes before this segment
            uint8_t back[3];
            int nb = 0;
            for (uint64_t p = b0; p > rb && nb < 3;) {
                --p;
                const uint8_t c = cls[seq[p]];
                if (c != rg::CLS_SKIP) back[nb++] = c;
            }
            uint32_t code = 0, run = 0;
            for (int i = nb - 1; i >= 0; --i) {
                if (back[i] < 4) { code = (code >> 2) | ((uint32_t)back[i] << 6); run += 1; }
                else run = 0;
            }
            const uint64_t b1 = min(b0 + BYTES_PER_THREAD, re);
            for (uint64_t p = b0; p < b1; ++p) {
                const uint8_t c = cls[seq[p]];
                if (c == rg::CLS_SKIP) continue;
                if (c < 4) {
                    code = (code >> 2) | ((uint32_t)c << 6);
                    if (++run >= 4) atomicAdd(&hist[warp][lut[code & 0xffu]], 1u);
                } else {
                    run = 0;
                }
            }
        }
        flush_hist(hist, counts, rec);
    }
}

// one warp per record: n_kmers = sum of the 136 bins (every counted window hits exactly one bin,
// kmer_counting.rs:100-101); freq = count / n_kmers, NaN for 0/0 as in :104-107
__global__ void tnf_freq_kernel(const uint32_t *__restrict__ counts, uint64_t n, uint64_t *__restrict__ n_kmers,
                                double *__restrict__ freqs) {
    const int lane = threadIdx.x & 31;
    const uint64_t rec = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (rec >= n) return;
    uint32_t c[5];
    uint64_t s = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const uint32_t b = i * 32 + lane;
        c[i] = b < RG_NBINS ? counts[rec * RG_NBINS + b] : 0u;
        s += c[i];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) n_kmers[rec] = s;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const uint32_t b = i * 32 + lane;
        if (b < RG_NBINS) freqs[rec * RG_NBINS + b] = (double)c[i] / (double)s;
    }
}

}  // namespace

int32_t rg_tnf_launch_packed(rg_handle *h) {
    unsigned grid = (unsigned)std::min<uint64_t>(h->tnf_chunks, (uint64_t)h->sm_count * 8);
    if (grid == 0) return RG_OK;
    tnf_count_packed_kernel<<<grid, THREADS, 0, h->stream>>>(
        (const uint64_t *)h->codes.p, (const uint32_t *)h->valid.p, (const uint64_t *)h->woff.p,
        (const uint64_t *)h->chunk_start.p, h->tnf_n, h->tnf_chunks, (const uint8_t *)h->tnf_tables.p,
        (uint32_t *)h->counts.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

int32_t rg_tnf_launch_ascii(rg_handle *h) {
    unsigned grid = (unsigned)std::min<uint64_t>(h->tnf_chunks, (uint64_t)h->sm_count * 8);
    if (grid == 0) return RG_OK;
    tnf_count_ascii_kernel<<<grid, THREADS, 0, h->stream>>>(
        (const uint8_t *)h->ascii.p, (const uint64_t *)h->aoff.p, (const uint64_t *)h->chunk_start.p, h->tnf_n,
        h->tnf_chunks, (const uint8_t *)h->tnf_tables.p, (uint32_t *)h->counts.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

int32_t rg_tnf_launch_freq(rg_handle *h) {
    if (h->tnf_n == 0) return RG_OK;
    const int wpb = 8;
    tnf_freq_kernel<<<(unsigned)((h->tnf_n + wpb - 1) / wpb), wpb * 32, 0, h->stream>>>(
        (const uint32_t *)h->counts.p, h->tnf_n, (uint64_t *)h->nkmers.p, (double *)h->freqs.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

uint32_t rg_tnf_chunk_words() { return CHUNK_WORDS; }
uint32_t rg_tnf_chunk_bytes() { return CHUNK_BYTES; }
