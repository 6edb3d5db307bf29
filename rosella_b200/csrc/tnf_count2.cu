// tnf_count2.cu — K1: tetranucleotide counting with thread-private byte counters.
//
// Replaces the per-record loop of KmerCounter::run (src/kmers/kmer_counting.rs:64-109): for every window of 4
// consecutive A/C/G/T bases of the normalised record, increment the bin of the lexicographically smaller of
// (4-mer, reverse complement) (:82-101).  The first generation of this kernel paid a shared-memory atomic plus a LUT
// load per base and was bound by the ALU/LSU pipes at 5 % of the HBM roofline (profiles/r1c_ncu_full_summary.json,
// profiles/r1d_tnf_4gbp_v1.json).  Here
//   * every thread owns one BYTE counter per raw 8-bit window code: counter (code, warp, lane) is byte
//     (warp & 3) of the 32-bit word [code][lane] of a 32 KB plane shared by the 4 warps of a CTA, so a
//     warp's 32 lanes always touch 32 different banks whatever their codes are (no conflicts) and no two
//     threads ever share a byte.  An increment is ONE fire-and-forget ATOMS.ADD of 1 << 8 * warp on the word
//     (atomic only because the four warps share the word; a byte is folded before it can carry into its
//     neighbour): one LSU instruction per base and no load -> add -> store chain (LDS.U8 + IADD + STS.U8
//     measured 2.44 Tbp/s, the atomic 3.23 Tbp/s, profiles/r1r_tnf_4gbp_atoms.json);
//   * the address of the counter is formed straight from the packed word: (ext >> (2j - 7)) & 0x7f80
//     is code * 128, one funnel shift and one LOP3 per base;
//   * a thread counts at most 7 words x 32 windows = 224 <= 255 windows between two flushes, so a byte
//     cannot overflow; the flush sums the 128 bytes of a code with IDP.4A, folds the 256 raw codes to
//     the 136 canonical bins (256 -> 136 LUT, tnf_host.cpp) into a CTA histogram, and that histogram
//     goes to global memory only when the CTA moves on to another record.
//   * the ASCII kernel classifies a unit of 32 bytes in registers (classify4: PRMT table look-ups keyed by byte & 7,
//     one exact all-bases test per unit) and counts it like a packed word: 0.86 -> 2.26 Tbp/s (profiles/r2q_*); units
//     with any other byte take the byte-by-byte path.
// Integer work, bit-exact by construction; tests/test_gpu_tnf.py compares with the oracle.
#include "rg_internal.cuh"
#include "tnf_tables.h"

namespace {

constexpr int T2 = 128;                       // threads per CTA: the 4 warps sharing one counter plane
constexpr int WPT = 7;                        // 32-base units per thread between flushes (7 * 32 <= 255)
constexpr uint32_t UNIT_PER_CHUNK = T2 * WPT; // 896 words (28 672 bases) or 896 x 32 raw bytes per chunk
constexpr uint32_t SEG_BYTES = 32;            // ASCII kernel: raw bytes per unit

struct Smem2 {
    uint32_t cnt[256 * 32];                   // [code][lane] words, byte (warp & 3) = that thread's counter
    uint32_t hist[RG_NBINS];                  // CTA histogram of the record being counted
    uint8_t lut[256];                         // little-endian window code -> canonical bin
    uint8_t cls[256];                         // byte class table (ASCII kernel)
};

// sum the byte counters of the chunk into the CTA histogram and clear them
__device__ __forceinline__ void fold_counters(Smem2 &sm, int tid) {
    __syncthreads();
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int code = tid + half * T2;
        uint4 *row = reinterpret_cast<uint4 *>(sm.cnt + code * 32);
        uint32_t s = 0;
#pragma unroll
        for (int l = 0; l < 8; ++l) {
            const int w = (l + tid) & 7;       // rotate: the 8 lanes of a quarter warp read 8 different 16-byte bank groups
            const uint4 v = row[w];
            row[w] = make_uint4(0u, 0u, 0u, 0u);
            s = __dp4a(v.x, 0x01010101u, s);   // unsigned bytes: a counter may hold up to 224
            s = __dp4a(v.y, 0x01010101u, s);
            s = __dp4a(v.z, 0x01010101u, s);
            s = __dp4a(v.w, 0x01010101u, s);
        }
        if (s) atomicAdd(&sm.hist[sm.lut[code]], s);
    }
    __syncthreads();
}

// add the CTA histogram to the record's global counts and clear it
__device__ __forceinline__ void flush_record(Smem2 &sm, uint32_t *__restrict__ counts, uint64_t rec, int tid) {
    for (int b = tid; b < (int)RG_NBINS; b += T2) {
        const uint32_t s = sm.hist[b];
        if (s) { atomicAdd(&counts[rec * RG_NBINS + b], s); sm.hist[b] = 0; }
    }
    __syncthreads();
}

// counter increment: a = byte address of the word that holds this thread's counter, inc = 1 << (8 * byte).  One fire-and-forget shared-memory atomic on the
// containing word (the four warps of a CTA own the four bytes of a word, and a byte is folded before it can carry)
// instead of a load / add / store of the byte: one LSU instruction per base and no dependent chain.
// tb packs both: bits [2,7) = lane (the word of the row), bits [0,2) = warp (the byte of the word).
__device__ __forceinline__ void bump(uint8_t *__restrict__ cb, uint32_t a, uint32_t inc) {
    atomicAdd(reinterpret_cast<unsigned int *>(cb + a), inc);
}

// ---------------------------------------------------------------- packed input (0.375 B/base)
// One 32-base word: the windows ending at its 32 bases (the first three reach into the previous word).
// cb = the (CTA-uniform) byte base of the counter plane, tb = lane * 4 = this lane's word in a row, inc = 1 << 8 warp
__device__ __forceinline__ void count_word(uint8_t *__restrict__ cb, uint32_t tb, uint32_t inc, uint64_t c, uint32_t v,
                                           uint32_t pc, uint32_t pv) {
    // 70-bit extended word: bits [0,6) = the previous 3 bases, then the 32 bases of this word
    const uint32_t e0 = ((uint32_t)c << 6) | pc;
    const uint32_t e1 = (uint32_t)(c >> 26);
    const uint32_t e2 = (uint32_t)(c >> 58);
    const uint64_t vext = ((uint64_t)v << 3) | pv;
    // window ending at base j = bits [2j, 2j+8) of the extended word; code * 128 = (ext >> (2j-7)) & 0x7f80
    auto addr = [&](int j) -> uint32_t {
        const int sh = 2 * j - 7;
        uint32_t x;
        if (sh < 0) x = e0 << (-sh);
        else if (sh < 32) x = __funnelshift_r(e0, e1, sh);
        else x = __funnelshift_r(e1, e2, sh - 32);
        return (x & 0x7f80u) | tb;             // one LOP3
    };
    if (vext == 0x7ffffffffull) {               // common case: all 35 bases are A/C/G/T
#pragma unroll
        for (int j = 0; j < 32; ++j) bump(cb, addr(j), inc);
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (((uint32_t)(vext >> j) & 0xfu) == 0xfu) bump(cb, addr(j), inc);
    }
}

// chunk -> record bookkeeping shared by both kernels: the CTA walks a contiguous range of chunks, so the
// record index only ever moves forward and the CTA histogram is flushed when it does
struct RecordCursor {
    uint64_t rec;
    __device__ __forceinline__ void seek(const uint64_t *__restrict__ start, uint64_t n, uint64_t chunk) {
        uint64_t lo = 0, hi = n;                // invariant: start[lo] <= chunk < start[hi]
        while (hi - lo > 1) {
            const uint64_t mid = (lo + hi) >> 1;
            if (start[mid] <= chunk) lo = mid; else hi = mid;
        }
        rec = lo;
    }
};

__global__ void __launch_bounds__(T2, 6)
tnf_count_packed2_kernel(const uint64_t *__restrict__ codes, const uint32_t *__restrict__ valid,
                         const uint64_t *__restrict__ woff, const uint64_t *__restrict__ chunk_start, uint64_t n,
                         uint64_t n_chunks, uint64_t chunks_per_cta, const uint8_t *__restrict__ tables,
                         uint32_t *__restrict__ counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    Smem2 &sm = *reinterpret_cast<Smem2 *>(smem_raw);
    const int tid = threadIdx.x;
    for (int i = tid; i < 256 * 32; i += T2) sm.cnt[i] = 0;
    for (int i = tid; i < (int)RG_NBINS; i += T2) sm.hist[i] = 0;
    for (int i = tid; i < 256; i += T2) sm.lut[i] = tables[i];
    __syncthreads();
    uint8_t *cb = reinterpret_cast<uint8_t *>(sm.cnt);
    const uint32_t tb = (tid & 31) * 4, inc = 1u << (8 * (tid >> 5));   // word of the row, byte of the word

    const uint64_t c_begin = (uint64_t)blockIdx.x * chunks_per_cta;
    const uint64_t c_end = min(c_begin + chunks_per_cta, n_chunks);
    if (c_begin >= c_end) return;
    RecordCursor cur;
    cur.seek(chunk_start, n, c_begin);
    for (uint64_t chunk = c_begin; chunk < c_end; ++chunk) {
        if (chunk >= chunk_start[cur.rec + 1]) {                    // next record (records without chunks are skipped)
            flush_record(sm, counts, cur.rec, tid);
            do { ++cur.rec; } while (chunk >= chunk_start[cur.rec + 1]);
        }
        const uint64_t rec_w0 = woff[cur.rec], rec_w1 = woff[cur.rec + 1];
        const uint64_t w0 = rec_w0 + (chunk - chunk_start[cur.rec]) * UNIT_PER_CHUNK + tid;
        // software pipeline of depth 1: the loads of word i+1 are in flight while word i is counted (a word is
        // ~32 dependent shared-memory updates, far longer than a DRAM round trip)
        uint64_t c = 0, cp = 0;
        uint32_t v = 0, vp = 0;
        auto fetch = [&](uint64_t w) {
            if (w < rec_w1) {
                c = codes[w];
                v = valid[w];
                if (w > rec_w0) { cp = codes[w - 1]; vp = valid[w - 1]; }   // last 3 bases of the previous word of this record
                else { cp = 0; vp = 0; }
            }
        };
        fetch(w0);
#pragma unroll 1
        for (int i = 0; i < WPT; ++i) {
            const uint64_t w = w0 + (uint64_t)i * T2;
            if (w >= rec_w1) break;
            const uint64_t c_now = c;
            const uint32_t v_now = v, pc = (uint32_t)(cp >> 58), pv = vp >> 29;
            if (i + 1 < WPT) fetch(w + T2);
            count_word(cb, tb, inc, c_now, v_now, pc, pv);
        }
        fold_counters(sm, tid);
    }
    flush_record(sm, counts, cur.rec, tid);
}

// ---------------------------------------------------------------- raw ASCII input (1 B/base)
// Units are the 32-byte ALIGNED segments of the device buffer (two 16-byte loads per thread); a record
// [rb, re) owns the units floor(rb/32) .. ceil(re/32) and a thread counts the bytes of its unit that lie
// inside the record.  needletail's normalize(false) is the byte-class table (tnf_host.cpp).
struct Roll {
    uint32_t code7;   // rolling window code * 128 (first base in the low bits), garbage below bit 7
    uint32_t run;     // consecutive A/C/G/T bases seen
};

__device__ __forceinline__ void count_byte(uint8_t *__restrict__ cb, uint32_t tb, uint32_t inc,
                                           const uint8_t *__restrict__ cls, uint32_t byte, Roll &st) {
    const uint32_t c = cls[byte];
    if (c == rg::CLS_SKIP) return;              // removed by normalize: the window carries on
    if (c < 4) {
        st.code7 = (st.code7 >> 2) | (c << 13);
        if (++st.run >= 4) bump(cb, (st.code7 & 0x7f80u) | tb, inc);
    } else {
        st.run = 0;                             // kept but not a base: breaks the window
    }
}

// Classification of 4 raw bytes in registers (no class-table LDS): the low three bits of A/C/G/T/U (either case) are
// 1 / 3 / 7 / 4 / 5 — distinct — so `byte & 7` selects, through PRMT, both the byte an A/C/G/T/U with those bits would
// be (table EXP, 0xFF where none exists) and its 2-bit code (table COD).  byte ^ EXP[byte & 7] is 0x00 (upper case) or
// 0x20 (lower case) exactly for the ten bytes needletail's normalize(false) turns into a base; OR-ing it over the unit
// gives one exact all-bases test.  The selector nibbles come out pairwise swapped (b1, b0, b3, b2); the raw word is
// permuted the same way for the compare and the packing multiply puts the codes back in order.
// Returns the four codes in the TOP byte (first base lowest), junk below.
__device__ __forceinline__ uint32_t classify4(uint32_t w, uint32_t &bad) {
    const uint32_t v = w & 0x07070707u;
    const uint32_t t = v * 0x1001u;                              // v | v << 12: byte 1 = (b1, b0), byte 3 = (b3, b2)
    const uint32_t sel = __byte_perm(t, 0u, 0x0031u);
    const uint32_t exp = __byte_perm(0x43FF41FFu, 0x47FF5554u, sel);
    const uint32_t cod = __byte_perm(0x01000000u, 0x02000303u, sel);
    bad |= __byte_perm(w, 0u, 0x2301u) ^ exp;
    return cod * 0x04014010u;                                    // c0 << 24 | c1 << 26 | c2 << 28 | c3 << 30, disjoint partial products
}

__global__ void __launch_bounds__(T2, 6)
tnf_count_ascii2_kernel(const uint8_t *__restrict__ seq, const uint64_t *__restrict__ offsets,
                        const uint64_t *__restrict__ chunk_start, uint64_t n, uint64_t n_chunks,
                        uint64_t chunks_per_cta, const uint8_t *__restrict__ tables, uint32_t *__restrict__ counts) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    Smem2 &sm = *reinterpret_cast<Smem2 *>(smem_raw);
    const int tid = threadIdx.x;
    for (int i = tid; i < 256 * 32; i += T2) sm.cnt[i] = 0;
    for (int i = tid; i < (int)RG_NBINS; i += T2) sm.hist[i] = 0;
    for (int i = tid; i < 256; i += T2) { sm.lut[i] = tables[i]; sm.cls[i] = tables[256 + i]; }
    __syncthreads();
    uint8_t *cb = reinterpret_cast<uint8_t *>(sm.cnt);
    const uint32_t tb = (tid & 31) * 4, inc = 1u << (8 * (tid >> 5));   // word of the row, byte of the word
    const uint8_t *cls = sm.cls;

    const uint64_t c_begin = (uint64_t)blockIdx.x * chunks_per_cta;
    const uint64_t c_end = min(c_begin + chunks_per_cta, n_chunks);
    if (c_begin >= c_end) return;
    RecordCursor cur;
    cur.seek(chunk_start, n, c_begin);
    for (uint64_t chunk = c_begin; chunk < c_end; ++chunk) {
        if (chunk >= chunk_start[cur.rec + 1]) {
            flush_record(sm, counts, cur.rec, tid);
            do { ++cur.rec; } while (chunk >= chunk_start[cur.rec + 1]);
        }
        const uint64_t rb = offsets[cur.rec], re = offsets[cur.rec + 1];
        const uint64_t u_end = (re + SEG_BYTES - 1) / SEG_BYTES;
        const uint64_t u0 = rb / SEG_BYTES + (chunk - chunk_start[cur.rec]) * UNIT_PER_CHUNK + tid;
        // depth-1 software pipeline of the two 16-byte loads of a unit
        uint4 n0 = make_uint4(0, 0, 0, 0), n1 = n0;
        uint32_t npw = 0;                                        // the 4 bytes before the unit
        if (u0 < u_end) {
            n0 = *reinterpret_cast<const uint4 *>(seq + u0 * SEG_BYTES);
            n1 = *reinterpret_cast<const uint4 *>(seq + u0 * SEG_BYTES + 16);
            if (u0 > 0) npw = *reinterpret_cast<const uint32_t *>(seq + u0 * SEG_BYTES - 4);
        }
#pragma unroll 1
        for (int i = 0; i < WPT; ++i) {
            const uint64_t u = u0 + (uint64_t)i * T2;
            if (u >= u_end) break;
            const uint4 q0 = n0, q1 = n1;
            const uint32_t pw = npw;
            if (i + 1 < WPT && u + T2 < u_end) {
                n0 = *reinterpret_cast<const uint4 *>(seq + (u + T2) * SEG_BYTES);
                n1 = *reinterpret_cast<const uint4 *>(seq + (u + T2) * SEG_BYTES + 16);
                npw = *reinterpret_cast<const uint32_t *>(seq + (u + T2) * SEG_BYTES - 4);
            }
            const uint64_t base = u * SEG_BYTES;
            const uint64_t p0 = max(base, rb), p1 = min(base + SEG_BYTES, re);
            // Fast path: the unit and the 4 bytes before it lie inside the record and all 36 bytes are bases.  The 32
            // codes are packed in registers and counted exactly like a word of the packed kernel: no class-table
            // load, no rolling state, two ALU instructions and one ATOMS per base.
            if (base >= rb + 4 && p1 == base + SEG_BYTES) {
                uint32_t bad = 0, lo = 0, hi = 0;
                hi = __funnelshift_l(classify4(q1.w, bad), hi, 8);
                hi = __funnelshift_l(classify4(q1.z, bad), hi, 8);
                hi = __funnelshift_l(classify4(q1.y, bad), hi, 8);
                hi = __funnelshift_l(classify4(q1.x, bad), hi, 8);
                lo = __funnelshift_l(classify4(q0.w, bad), lo, 8);
                lo = __funnelshift_l(classify4(q0.z, bad), lo, 8);
                lo = __funnelshift_l(classify4(q0.y, bad), lo, 8);
                lo = __funnelshift_l(classify4(q0.x, bad), lo, 8);
                const uint32_t pc = classify4(pw, bad) >> 26;    // the last 3 bases before the unit
                if ((bad & 0xDFDFDFDFu) == 0u) {
                    count_word(cb, tb, inc, ((uint64_t)hi << 32) | lo, 0xffffffffu, pc, 7u);
                    continue;
                }
            }
            // rebuild the rolling state from the (up to) 3 kept bytes before p0
            uint8_t back[3];
            int nb = 0;
            for (uint64_t p = p0; p > rb && nb < 3;) {
                --p;
                const uint8_t c = cls[seq[p]];
                if (c != rg::CLS_SKIP) back[nb++] = c;
            }
            Roll st{0u, 0u};
            for (int k = nb - 1; k >= 0; --k) {
                if (back[k] < 4) { st.code7 = (st.code7 >> 2) | ((uint32_t)back[k] << 13); st.run += 1; }
                else st.run = 0;
            }
            const uint32_t wd[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
            if (p0 == base && p1 == base + SEG_BYTES) {
#pragma unroll
                for (int b = 0; b < 32; ++b) count_byte(cb, tb, inc, cls, (wd[b >> 2] >> (8 * (b & 3))) & 0xffu, st);
            } else {
                const int b0 = (int)(p0 - base), b1 = (int)(p1 - base);
#pragma unroll
                for (int b = 0; b < 32; ++b)
                    if (b >= b0 && b < b1) count_byte(cb, tb, inc, cls, (wd[b >> 2] >> (8 * (b & 3))) & 0xffu, st);
            }
        }
        fold_counters(sm, tid);
    }
    flush_record(sm, counts, cur.rec, tid);
}

}  // namespace

uint32_t rg_tnf2_chunk_units() { return UNIT_PER_CHUNK; }
uint32_t rg_tnf2_seg_bytes() { return SEG_BYTES; }

static unsigned plan_grid(rg_handle *h, uint64_t n_chunks, uint64_t &per) {
    // contiguous chunk ranges per CTA; ~4 CTAs' worth of ranges per resident slot evens out uneven fill
    const uint64_t slots = (uint64_t)h->sm_count * 6 * 4;
    per = (n_chunks + slots - 1) / slots;
    if (per == 0) per = 1;
    return (unsigned)((n_chunks + per - 1) / per);
}

int32_t rg_tnf_launch_packed2(rg_handle *h) {
    if (h->tnf_chunks == 0) return RG_OK;
    uint64_t per;
    const unsigned grid = plan_grid(h, h->tnf_chunks, per);
    auto kern = tnf_count_packed2_kernel;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem2)));
    kern<<<grid, T2, sizeof(Smem2), h->stream>>>((const uint64_t *)h->codes.p, (const uint32_t *)h->valid.p,
                                                 (const uint64_t *)h->woff.p, (const uint64_t *)h->chunk_start.p,
                                                 h->tnf_n, h->tnf_chunks, per, (const uint8_t *)h->tnf_tables.p,
                                                 (uint32_t *)h->counts.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

// records [0, n) described by offsets / chunk_start (device), counted into `counts`, on `stream`
int32_t rg_tnf_launch_ascii2_range(rg_handle *h, cudaStream_t stream, const uint64_t *offsets, const uint64_t *chunk_start,
                                   uint64_t n, uint64_t n_chunks, uint32_t *counts) {
    if (n_chunks == 0) return RG_OK;
    uint64_t per;
    const unsigned grid = plan_grid(h, n_chunks, per);
    auto kern = tnf_count_ascii2_kernel;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem2)));
    kern<<<grid, T2, sizeof(Smem2), stream>>>((const uint8_t *)h->ascii.p, offsets, chunk_start, n, n_chunks, per,
                                              (const uint8_t *)h->tnf_tables.p, counts);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}

int32_t rg_tnf_launch_ascii2(rg_handle *h) {
    if (h->tnf_chunks == 0) return RG_OK;
    uint64_t per;
    const unsigned grid = plan_grid(h, h->tnf_chunks, per);
    auto kern = tnf_count_ascii2_kernel;
    RG_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem2)));
    kern<<<grid, T2, sizeof(Smem2), h->stream>>>((const uint8_t *)h->ascii.p, (const uint64_t *)h->aoff.p,
                                                 (const uint64_t *)h->chunk_start.p, h->tnf_n, h->tnf_chunks, per,
                                                 (const uint8_t *)h->tnf_tables.p, (uint32_t *)h->counts.p);
    RG_CUDA(h, cudaGetLastError());
    h->tm.kernel_launches += 1;
    return RG_OK;
}
