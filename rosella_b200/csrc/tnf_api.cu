// tnf_api.cu — C ABI of the TNF path (include/rosella_gpu.h, "TNF counting").
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "rg_internal.cuh"
#include "tnf_tables.h"

int32_t rg_tnf_launch_freq(rg_handle *h);
// counting kernels (tnf_count2.cu): thread-private byte counters
int32_t rg_tnf_launch_packed2(rg_handle *h);
int32_t rg_tnf_launch_ascii2(rg_handle *h);
int32_t rg_tnf_launch_ascii2_range(rg_handle *h, cudaStream_t stream, const uint64_t *offsets, const uint64_t *chunk_start,
                                   uint64_t n, uint64_t n_chunks, uint32_t *counts);
int32_t rg_tnf_launch_freq_range(rg_handle *h, cudaStream_t stream, const uint32_t *counts, uint64_t n, uint64_t *nkmers, double *freqs);
uint32_t rg_tnf2_chunk_units();
uint32_t rg_tnf2_seg_bytes();

static int32_t ensure_tables(rg_handle *h) {
    if (h->tnf_tables.p) return RG_OK;
    uint8_t t[512];
    rg::build_bin_tables(nullptr, t, nullptr);
    rg::build_class_table(t + 256);
    RG_TRY(rg_reserve(h, h->tnf_tables, sizeof(t)));
    RG_CUDA(h, cudaMemcpy(h->tnf_tables.p, t, sizeof(t), cudaMemcpyHostToDevice));
    return RG_OK;
}

// exclusive scan of chunks per record -> device; lens given as an offsets array.  align = 0: a record of
// len elements has ceil(len/unit) chunks; align > 0 (ASCII v2): the record owns the align-byte ALIGNED
// segments it touches and a chunk holds `unit` of them.
static int32_t upload_chunks(rg_handle *h, const uint64_t *offsets, uint64_t n, uint64_t unit, uint64_t align = 0) {
    std::vector<uint64_t> start;
    try { start.resize(n + 1); } catch (...) { return rg_fail(h, RG_ERR_OOM, "host allocation of %llu chunk offsets failed", (unsigned long long)(n + 1)); }
    start[0] = 0;
    for (uint64_t r = 0; r < n; ++r) {
        uint64_t len = offsets[r + 1] - offsets[r];
        if (align && len) len = (offsets[r + 1] + align - 1) / align - offsets[r] / align;
        start[r + 1] = start[r] + (len + unit - 1) / unit;
    }
    h->tnf_chunks = start[n];
    RG_TRY(rg_reserve(h, h->chunk_start, (n + 1) * sizeof(uint64_t)));
    RG_CUDA(h, cudaMemcpyAsync(h->chunk_start.p, start.data(), (n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));   // start[] goes out of scope
    return RG_OK;
}

static int32_t reserve_outputs(rg_handle *h, uint64_t n) {
    RG_TRY(rg_reserve(h, h->counts, n * RG_NBINS * sizeof(uint32_t)));
    RG_TRY(rg_reserve(h, h->nkmers, n * sizeof(uint64_t)));
    RG_TRY(rg_reserve(h, h->freqs, n * RG_NBINS * sizeof(double)));
    return RG_OK;
}

extern "C" {

int32_t rg_tnf_load_packed(rg_handle *h, const uint64_t *codes, const uint32_t *valid, const uint64_t *word_offsets,
                           uint64_t n) {
    if (!h) return RG_ERR_INVALID;
    if (!word_offsets || (n && (!codes || !valid))) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_load_packed: NULL pointer");
    RG_CUDA(h, cudaSetDevice(h->device));
    h->tnf_mode = 0;
    for (uint64_t r = 0; r < n; ++r)
        if (word_offsets[r + 1] < word_offsets[r]) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_load_packed: word_offsets must be non-decreasing");
    const uint64_t words = n ? word_offsets[n] : 0;
    RG_TRY(ensure_tables(h));
    RG_TRY(rg_reserve(h, h->codes, words * sizeof(uint64_t)));
    RG_TRY(rg_reserve(h, h->valid, words * sizeof(uint32_t)));
    RG_TRY(rg_reserve(h, h->woff, (n + 1) * sizeof(uint64_t)));
    RG_TRY(reserve_outputs(h, n));
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->codes.p, codes, words * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->valid.p, valid, words * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->woff.p, word_offsets, (n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_TRY(upload_chunks(h, word_offsets, n, rg_tnf2_chunk_units()));
    h->tm.h2d_ms = rg_elapsed(h->ev[0], h->ev[1]);
    h->tm.h2d_bytes += words * 12 + (n + 1) * 16;
    h->tnf_n = n; h->tnf_words = words; h->tnf_bytes = 0; h->tnf_mode = 1;
    return RG_OK;
}

int32_t rg_tnf_load_ascii(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n) {
    if (!h) return RG_ERR_INVALID;
    if (!offsets || (n && !seq && offsets[n] > 0)) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_load_ascii: NULL pointer");
    RG_CUDA(h, cudaSetDevice(h->device));
    h->tnf_mode = 0;
    for (uint64_t r = 0; r < n; ++r)
        if (offsets[r + 1] < offsets[r]) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_load_ascii: offsets must be non-decreasing");
    const uint64_t lo = n ? offsets[0] : 0, hi = n ? offsets[n] : 0;
    RG_TRY(ensure_tables(h));
    RG_TRY(rg_reserve(h, h->ascii, (hi + 63) / 32 * 32));   // the v2 kernel reads whole aligned 32-byte segments
    RG_TRY(rg_reserve(h, h->aoff, (n + 1) * sizeof(uint64_t)));
    RG_TRY(reserve_outputs(h, n));
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    if (hi > lo)
        RG_CUDA(h, cudaMemcpyAsync((uint8_t *)h->ascii.p + lo, seq + lo, hi - lo, cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->aoff.p, offsets, (n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_TRY(upload_chunks(h, offsets, n, rg_tnf2_chunk_units(), rg_tnf2_seg_bytes()));
    h->tm.h2d_ms = rg_elapsed(h->ev[0], h->ev[1]);
    h->tm.h2d_bytes += (hi - lo) + (n + 1) * 16;
    h->tnf_n = n; h->tnf_words = 0; h->tnf_bytes = hi - lo; h->tnf_mode = 2;
    return RG_OK;
}

int32_t rg_tnf_count_resident(rg_handle *h) {
    if (!h) return RG_ERR_INVALID;
    if (h->tnf_mode == 0) return rg_fail(h, RG_ERR_STATE, "rg_tnf_count_resident: nothing loaded (rg_tnf_load_* first)");
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_CUDA(h, cudaMemsetAsync(h->counts.p, 0, h->tnf_n * RG_NBINS * sizeof(uint32_t), h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[6], h->stream));
    RG_TRY(h->tnf_mode == 1 ? rg_tnf_launch_packed2(h) : rg_tnf_launch_ascii2(h));
    RG_CUDA(h, cudaEventRecord(h->ev[7], h->stream));
    RG_TRY(rg_tnf_launch_freq(h));
    h->tm.tnf_count_ms = -1.0;   // resolved lazily by rg_get_timings
    return RG_OK;
}

int32_t rg_tnf_fetch(rg_handle *h, uint32_t *counts, uint64_t *n_kmers, double *freqs) {
    if (!h) return RG_ERR_INVALID;
    if (h->tnf_mode == 0) return rg_fail(h, RG_ERR_STATE, "rg_tnf_fetch: nothing counted");
    RG_CUDA(h, cudaSetDevice(h->device));
    const uint64_t n = h->tnf_n;
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    if (counts) RG_CUDA(h, cudaMemcpyAsync(counts, h->counts.p, n * RG_NBINS * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    if (n_kmers) RG_CUDA(h, cudaMemcpyAsync(n_kmers, h->nkmers.p, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (freqs) RG_CUDA(h, cudaMemcpyAsync(freqs, h->freqs.p, n * RG_NBINS * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    h->tm.d2h_ms = rg_elapsed(h->ev[0], h->ev[1]);
    h->tm.d2h_bytes += (counts ? n * RG_NBINS * 4 : 0) + (n_kmers ? n * 8 : 0) + (freqs ? n * RG_NBINS * 8 : 0);
    return RG_OK;
}

// The one-call form pipelines: the records are cut into slices of equal bytes, slice c is uploaded, counted and fetched
// on stream c & 1, so the kernels and the device->host copies of one slice run under the host->device copy of the next
// (the call is bound by the PCIe upload: 1 byte per base in, 1.6 KB per record out).
int32_t rg_tnf_count_ascii(rg_handle *h, const uint8_t *seq, const uint64_t *offsets, uint64_t n, uint32_t *counts,
                           uint64_t *n_kmers, double *freqs) {
    if (!h) return RG_ERR_INVALID;
    if (!counts || !n_kmers) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_count_ascii: counts / n_kmers is NULL");
    if (!offsets || (n && !seq && offsets[n] > 0)) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_count_ascii: NULL pointer");
    RG_CUDA(h, cudaSetDevice(h->device));
    h->tnf_mode = 0;
    for (uint64_t r = 0; r < n; ++r)
        if (offsets[r + 1] < offsets[r]) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_count_ascii: offsets must be non-decreasing");
    if (n == 0) return RG_OK;
    const uint64_t lo = offsets[0], hi = offsets[n], bytes = hi - lo;
    // slices of >= 64 MB, at most 16: enough to hide the kernels, few enough that launch / copy set-up stays invisible
    uint64_t n_slices = bytes / (64ull << 20);
    if (n_slices < 1) n_slices = 1;
    if (n_slices > 16) n_slices = 16;
    std::vector<uint64_t> cut(n_slices + 1, 0), start;
    try { start.resize(n + n_slices + 1); } catch (...) { return rg_fail(h, RG_ERR_OOM, "rg_tnf_count_ascii: host allocation failed"); }
    cut[n_slices] = n;
    for (uint64_t c = 1; c < n_slices; ++c) {
        const uint64_t target = lo + bytes / n_slices * c;
        uint64_t r = (uint64_t)(std::lower_bound(offsets, offsets + n + 1, target) - offsets);
        cut[c] = std::min(std::max(r, cut[c - 1]), n);
    }
    RG_TRY(ensure_tables(h));
    RG_TRY(rg_reserve(h, h->ascii, (hi + 63) / 32 * 32));
    RG_TRY(rg_reserve(h, h->aoff, (n + 1) * sizeof(uint64_t)));
    RG_TRY(rg_reserve(h, h->chunk_start, (n + n_slices + 1) * sizeof(uint64_t)));
    RG_TRY(reserve_outputs(h, n));
    for (auto &st : h->tnf_stream)
        if (!st) RG_CUDA(h, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    if (!h->tnf_ev) RG_CUDA(h, cudaEventCreateWithFlags(&h->tnf_ev, cudaEventDisableTiming));
    // chunk bookkeeping of every slice (local exclusive scans, one after the other in one array)
    const uint64_t unit = rg_tnf2_chunk_units(), align = rg_tnf2_seg_bytes();
    std::vector<uint64_t> slice_chunks(n_slices, 0);
    uint64_t w = 0;
    for (uint64_t c = 0; c < n_slices; ++c) {
        start[w] = 0;
        for (uint64_t r = cut[c]; r < cut[c + 1]; ++r) {
            uint64_t len = offsets[r + 1] - offsets[r];
            if (len) len = (offsets[r + 1] + align - 1) / align - offsets[r] / align;
            start[w + 1] = start[w] + (len + unit - 1) / unit;
            ++w;
        }
        slice_chunks[c] = start[w];
        ++w;
    }
    RG_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->aoff.p, offsets, (n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemcpyAsync(h->chunk_start.p, start.data(), (n + n_slices + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    RG_CUDA(h, cudaMemsetAsync(h->counts.p, 0, n * RG_NBINS * sizeof(uint32_t), h->stream));
    RG_CUDA(h, cudaEventRecord(h->tnf_ev, h->stream));
    for (auto &st : h->tnf_stream) RG_CUDA(h, cudaStreamWaitEvent(st, h->tnf_ev, 0));
    int32_t rc = RG_OK;
    for (uint64_t c = 0; c < n_slices && rc == RG_OK; ++c) {
        cudaStream_t st = h->tnf_stream[c & 1];
        const uint64_t a = cut[c], b = cut[c + 1], nrec = b - a;
        if (nrec == 0) continue;
        const uint64_t b0 = offsets[a], b1 = offsets[b];
        cudaError_t e = cudaSuccess;
        if (b1 > b0) e = cudaMemcpyAsync((uint8_t *)h->ascii.p + b0, seq + b0, b1 - b0, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) {
            uint32_t *dcounts = (uint32_t *)h->counts.p + a * RG_NBINS;
            uint64_t *dnk = (uint64_t *)h->nkmers.p + a;
            double *dfr = (double *)h->freqs.p + a * RG_NBINS;
            rc = rg_tnf_launch_ascii2_range(h, st, (const uint64_t *)h->aoff.p + a, (const uint64_t *)h->chunk_start.p + a + c, nrec,
                                            slice_chunks[c], dcounts);
            if (rc == RG_OK) rc = rg_tnf_launch_freq_range(h, st, dcounts, nrec, dnk, dfr);
            if (rc == RG_OK) {
                e = cudaMemcpyAsync(counts + a * RG_NBINS, dcounts, nrec * RG_NBINS * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaMemcpyAsync(n_kmers + a, dnk, nrec * sizeof(uint64_t), cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess && freqs) e = cudaMemcpyAsync(freqs + a * RG_NBINS, dfr, nrec * RG_NBINS * sizeof(double), cudaMemcpyDeviceToHost, st);
            }
        }
        if (e != cudaSuccess) rc = rg_fail(h, RG_ERR_CUDA, "rg_tnf_count_ascii: %s", cudaGetErrorString(e));
    }
    // join (also on failure: nothing may still be writing into the caller's buffers when the call returns)
    for (auto &st : h->tnf_stream) cudaStreamSynchronize(st);
    RG_CUDA(h, cudaEventRecord(h->ev[1], h->stream));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    RG_TRY(rc);
    h->tm.h2d_ms = rg_elapsed(h->ev[0], h->ev[1]);      // the whole pipelined call (copies and kernels overlap)
    h->tm.h2d_bytes += bytes + (2 * n + n_slices + 2) * 8;
    h->tm.d2h_bytes += n * RG_NBINS * 4 + n * 8 + (freqs ? n * RG_NBINS * 8 : 0);
    h->tnf_n = n; h->tnf_words = 0; h->tnf_bytes = bytes; h->tnf_mode = 2;   // the inputs stay resident (rg_tnf_count_resident works)
    // rg_tnf_count_resident walks ONE scan over all records: rebuild it for the resident form
    return upload_chunks(h, offsets, n, unit, align);
}

int32_t rg_tnf_count_packed(rg_handle *h, const uint64_t *codes, const uint32_t *valid, const uint64_t *word_offsets,
                            uint64_t n, uint32_t *counts, uint64_t *n_kmers, double *freqs) {
    if (!h) return RG_ERR_INVALID;
    if (!counts || !n_kmers) return rg_fail(h, RG_ERR_INVALID, "rg_tnf_count_packed: counts / n_kmers is NULL");
    RG_TRY(rg_tnf_load_packed(h, codes, valid, word_offsets, n));
    RG_TRY(rg_tnf_count_resident(h));
    return rg_tnf_fetch(h, counts, n_kmers, freqs);
}

}  // extern "C"
