// rg_api.cu — handle lifecycle, error text, timings.  See include/rosella_gpu.h.
#include "rg_internal.cuh"

static thread_local std::string g_create_error;

void rg_set_create_error(const char *msg) { g_create_error = msg; }

int32_t rg_fail(rg_handle *h, int32_t code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_create_error = buf;
    return code;
}

int32_t rg_reserve(rg_handle *h, DevBuf &b, size_t bytes) {
    if (bytes == 0) bytes = 16;
    if (b.cap >= bytes) return RG_OK;
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        b.p = nullptr;
        return rg_fail(h, RG_ERR_OOM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
    }
    b.cap = bytes;
    return RG_OK;
}

void rg_release(DevBuf &b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
}

extern "C" {

int32_t rg_abi_version(void) { return RG_ABI_VERSION; }

int32_t rg_create(int32_t device_id, rg_handle **out) {
    if (!out) return rg_fail(nullptr, RG_ERR_INVALID, "rg_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return rg_fail(nullptr, RG_ERR_NO_DEVICE,
                       "rg_create: no CUDA device visible (%s); this library has no CPU fallback",
                       e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (device_id < 0 || device_id >= count)
        return rg_fail(nullptr, RG_ERR_INVALID, "rg_create: device_id %d out of range [0,%d)", device_id, count);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device_id)) != cudaSuccess)
        return rg_fail(nullptr, RG_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return rg_fail(nullptr, RG_ERR_NO_DEVICE,
                       "rg_create: device %d (%s) is compute capability %d.%d; the kernels are built for "
                       "sm_100a only and there is no fallback",
                       device_id, prop.name, prop.major, prop.minor);
    rg_handle *h = new (std::nothrow) rg_handle();
    if (!h) return rg_fail(nullptr, RG_ERR_OOM, "rg_create: host allocation failed");
    h->device = device_id;
    h->sm_count = prop.multiProcessorCount;
    if ((e = cudaSetDevice(device_id)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete h;
        return rg_fail(nullptr, RG_ERR_CUDA, "rg_create: %s", cudaGetErrorString(e));
    }
    for (auto &ev : h->ev)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) {
            delete h;
            return rg_fail(nullptr, RG_ERR_CUDA, "rg_create: cudaEventCreate: %s", cudaGetErrorString(e));
        }
    *out = h;
    return RG_OK;
}

void rg_destroy(rg_handle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    rg_comm_destroy(h);
    for (auto &ev : h->dist_ev)
        if (ev) cudaEventDestroy(ev);
    DevBuf *bufs[] = {&h->tnf64, &h->cov64, &h->tnf_t, &h->tnf_norm, &h->tnf_mean, &h->tnf_hi, &h->tnf_lo, &h->cov_t, &h->err_flag,
                      &h->cand_d, &h->cand_i, &h->sym_d, &h->sym_t, &h->out_i, &h->out_d, &h->out_nv, &h->flag_list, &h->cert_ctr, &h->exact_scratch, &h->dist_pd, &h->dist_pi, &h->dist_rd, &h->dist_ri,
                      &h->dist_status, &h->cache_di, &h->cache_dd, &h->cache_ti, &h->cache_td,
                      &h->row_flag, &h->extra_off, &h->extra_idx, &h->extra_common, &h->coll_pairs, &h->sk_hash, &h->sk_off, &h->skt_hash, &h->codes, &h->valid,
                      &h->woff, &h->ascii, &h->aoff, &h->counts, &h->nkmers, &h->freqs, &h->tnf_tables,
                      &h->chunk_start};
    for (DevBuf *b : bufs) rg_release(*b);
    for (DevBuf &b : h->ws) rg_release(b);
    for (auto &ev : h->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto &ev : h->sym_ev)
        if (ev) cudaEventDestroy(ev);
    for (auto &st : h->sym_stream)
        if (st) cudaStreamDestroy(st);
    for (auto &st : h->tnf_stream)
        if (st) cudaStreamDestroy(st);
    if (h->tnf_ev) cudaEventDestroy(h->tnf_ev);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

const char *rg_last_error(const rg_handle *h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int32_t rg_sync(rg_handle *h) {
    if (!h) return RG_ERR_INVALID;
    RG_CUDA(h, cudaSetDevice(h->device));
    RG_CUDA(h, cudaStreamSynchronize(h->stream));
    return RG_OK;
}

void *rg_stream(rg_handle *h) { return h ? (void *)h->stream : nullptr; }

int32_t rg_set_option(rg_handle *h, const char *key, int64_t value) {
    if (!h) return RG_ERR_INVALID;
    if (!key) return rg_fail(h, RG_ERR_INVALID, "rg_set_option: key is NULL");
    const std::string k(key);
    auto bad = [&]() { return rg_fail(h, RG_ERR_INVALID, "rg_set_option: value %lld is not valid for \"%s\"", (long long)value, key); };
    if (k == "depth_build") { if (value < 0 || value > 2) return bad(); h->opt_depth_build = (int)value; }
    else if (k == "sym_block") { if (value < 0) return bad(); h->opt_sym_block = (uint64_t)value; }
    else if (k == "sym_chunk") { if (value < 0) return bad(); h->opt_sym_chunk = (uint64_t)value; }
    else if (k == "tnf_select_impl") { if (value < 0 || value > 1) return bad(); h->opt_tnf_fp32 = (int)value; }
    else if (k == "exact_compact") { if (value < 0 || value > 1) return bad(); h->opt_exact_compact = (int)value; }
    else if (k == "certify") { if (value < 0 || value > 1) return bad(); h->opt_certify = (int)value; }
    else if (k == "kprime_depth") { if (value != 0 && value != 64 && value != 128 && value != 256) return bad(); h->opt_kprime_depth = (uint32_t)value; }
    else if (k == "kprime_tnf") { if (value != 0 && value != 48 && value != 64 && value != 128 && value != 256) return bad(); h->opt_kprime_tnf = (uint32_t)value; }
    else if (k == "sketch_slice_bytes") { if (value < 0 || value > (1ll << 31) - 1) return bad(); h->opt_sketch_slice = (uint64_t)value; }
    else return rg_fail(h, RG_ERR_INVALID, "rg_set_option: unknown key \"%s\"", key);
    return RG_OK;
}

int32_t rg_get_timings(rg_handle *h, rg_timings *out) {
    if (!h || !out) return RG_ERR_INVALID;
    if (h->tm.select_ms < 0.0) {   // asynchronous search pending: events 3..5 bracket select / rescore
        RG_CUDA(h, cudaSetDevice(h->device));
        RG_CUDA(h, cudaEventSynchronize(h->ev[5]));
        h->tm.select_ms = rg_elapsed(h->ev[3], h->ev[4]);
        h->tm.rescore_ms = rg_elapsed(h->ev[4], h->ev[5]);
    }
    if (h->tm.tnf_count_ms < 0.0) {   // asynchronous count pending: events 6..7
        RG_CUDA(h, cudaSetDevice(h->device));
        RG_CUDA(h, cudaEventSynchronize(h->ev[7]));
        h->tm.tnf_count_ms = rg_elapsed(h->ev[6], h->ev[7]);
    }
    if (h->cert_ctr.p) {
        uint32_t c[4] = {};
        RG_CUDA(h, cudaSetDevice(h->device));
        RG_CUDA(h, cudaMemcpyAsync(c, h->cert_ctr.p, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
        RG_CUDA(h, cudaStreamSynchronize(h->stream));
        h->tm.rows_flagged = c[1];
        h->tm.rows_exact = c[2];
    }
    h->tm.extra_pairs = h->n_extra;
    *out = h->tm;
    return RG_OK;
}

}  // extern "C"
