// tnf_count.cu — K1 epilogue: n_kmers and frequencies from the 136-bin counts (the counting kernels are in
// tnf_count2.cu).  kmer_counting.rs:100-107: every counted window hits exactly one bin, freq = count / n_kmers.
#include "rg_internal.cuh"

namespace {

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

// records [0, n) of a counts slab, on `stream` (the pipelined ASCII path)
int32_t rg_tnf_launch_freq_range(rg_handle *h, cudaStream_t stream, const uint32_t *counts, uint64_t n, uint64_t *nkmers, double *freqs) {
    if (n == 0) return RG_OK;
    const int wpb = 8;
    tnf_freq_kernel<<<(unsigned)((n + wpb - 1) / wpb), wpb * 32, 0, stream>>>(counts, n, nkmers, freqs);
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

