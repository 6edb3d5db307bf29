// knn_depth_common.cuh — the FP32 form of the coverage metric shared by knn_depth2.cu (query batch against all rows,
// fused top-K') and knn_depth_sym.cu (symmetric whole-graph build).  See knn_depth2.cu for the algebra.
#pragma once
#include <algorithm>

#include "pk.cuh"
#include "rg_internal.cuh"

namespace depthk {

__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// erf for both halves: sign(x) * (1 - 2^(-t g(t))), t = |x| (no clamp), g = degree-4 fit of
// -log2(erfc(t))/t on [0,4] with a positive leading coefficient: max abs error 6.7e-7 in FP32, t g(t)
// increases monotonically beyond 4 so the result saturates to exactly +-1 (tools/fit_erf.py).
// 4 FFMA2 + FMUL2 + FADD2 per two values, 2 MUFU.EX2, 2 LOP3.  The error budget is set by the K' - k
// candidate margin, not by the reported distances (those are re-evaluated in f64).
__device__ __forceinline__ pk erf2(pk x) {
    const pk t = abs2(x);
    pk p = bc(2.944113406715287e-03f);
    p = fma2(p, t, bc(-2.9589898317605317e-02f));
    p = fma2(p, t, bc(1.4866547752377543e-01f));
    p = fma2(p, t, bc(9.185094237957379e-01f));
    p = fma2(p, t, bc(1.6278889993946153f));
    const pk y = p * t;
    const pk r = bc(1.0f) - mk(mufu_ex2(-lo(y)), mufu_ex2(-hi(y)));
    return mk(copysignf(lo(r), lo(x)), copysignf(hi(r), hi(x)));
}

// two query rows (halves of ax..aw) against one reference row b, one sample
// SYM (symmetric whole-graph build, knn_depth_sym.cu; RUST only): the value is accumulated as |val| and the sign of
// the equal-variance branch is tracked apart, once as the pair (query, reference) sees it (mnf) and once as
// (reference, query) would (mnb): the roles of p1 / p2 (:638) swap only when the two variances are EXACTLY equal, so
// the sign flips only then.  A negative minimum means ln of a negative value -> NaN -> distance 1.0 (:659-661).
struct SignTrack {
    float mnf0, mnb0, mnf1, mnb1;
};

template <int VARIANT, bool SYM>
__device__ __forceinline__ void depth_sample2(const pk ax, const pk ay, const pk az, const pk aw, const float4 b,
                                              float &acc0, float &cnt0, float &acc1, float &cnt1, SignTrack &st) {
    const float ib = fabsf(b.w);
    const pk ia = abs2(aw);
    const pk D = ax - bc(b.x);
    // everything is written in ndv = var_b - var_a = -dv: r = 1/ndv flips the sign of all four erf
    // arguments, hence of val, which |.| absorbs; ndv is +0 for equal variances, so its sign bit is
    // exactly "var_a > var_b" (:638) and no separate difference is needed for the RUST sign below
    const pk ndv = bc(b.y) - ay;
    const pk t = fma2(ndv, bc(b.z) - az, D * D);
    const pk q = (ay * bc(b.y)) * t;                                    // :625
    const pk tmp = mk(mufu_sqrt(fabsf(lo(q))), mufu_sqrt(fabsf(hi(q))));
    // :620 |a_var - b_var| < MIN_VAR_EPSILON, with the threshold lowered by the rounding error of the FP32 difference
    // (2^-24 (va + vb) <= 2^-23 vb here; 2^-21 vb is subtracted) so that "equal" in FP32 implies "equal" in f64 — the
    // branch keeps a sign in the Rust variant (:647), and a sign applied to a pair the reference sends through the
    // general branch would turn a finite distance into 1.0.  Above vb = 209 only exactly equal FP32 variances pass;
    // those are f64-equal or patched as extra candidates (knn_exact.cu).
    const float eq_thr = fmaxf(fmaf(-4.76837158e-7f, b.y, 1e-4f), 1e-30f);
    const bool eq0 = fabsf(lo(ndv)) < eq_thr, eq1 = fabsf(hi(ndv)) < eq_thr;
    const pk r = mk(eq0 ? 0.f : mufu_rcp(lo(ndv)), eq1 ? 0.f : mufu_rcp(hi(ndv)));
    // equal-variance branch: he = +-D/2 (0 otherwise); RUST folds the sign of p1.cdf - p2.cdf (:638,:647)
    // into it: + when var_a > var_b, - otherwise
    const pk hD = D * bc(0.5f);
    float h0 = lo(hD), h1 = hi(hD);
    if (VARIANT == RG_VARIANT_RUST) {
        h0 = __uint_as_float(__float_as_uint(h0) ^ (~__float_as_uint(lo(ndv)) & 0x80000000u));
        h1 = __uint_as_float(__float_as_uint(h1) ^ (~__float_as_uint(hi(ndv)) & 0x80000000u));
    }
    const pk he = mk(eq0 ? h0 : 0.f, eq1 ? h1 : 0.f);
    const pk ra = r * ia, rb = r * bc(ib);
    const pk A = ay * D, B = bc(b.y) * D;
    const pk x1 = fma2(tmp - A, ra, he * ia);
    const pk x2 = neg(tmp + A) * ra;
    const pk x3 = fma2(tmp - B, rb, neg(he * bc(ib)));
    const pk x4 = neg(tmp + B) * rb;
    const pk val = (erf2(x1) - erf2(x2)) - (erf2(x3) - erf2(x4));
    float v0 = lo(val), v1 = hi(val);
    if (VARIANT == RG_VARIANT_RUST && SYM) {
        const float s0 = eq0 ? v0 : 0.f, s1 = eq1 ? v1 : 0.f;           // :647 keeps the sign
        st.mnf0 = fminf(st.mnf0, s0);
        st.mnf1 = fminf(st.mnf1, s1);
        st.mnb0 = fminf(st.mnb0, lo(ndv) == 0.f ? -s0 : s0);
        st.mnb1 = fminf(st.mnb1, hi(ndv) == 0.f ? -s1 : s1);
        v0 = fabsf(v0);
        v1 = fabsf(v1);
    } else if (VARIANT == RG_VARIANT_RUST) {
        v0 = eq0 ? v0 : fabsf(v0);                                       // :650 |.| ; :647 keeps the sign
        v1 = eq1 ? v1 : fabsf(v1);
    } else {
        v0 = fminf(fmaxf(fabsf(v0), 2e-6f), 2.0f - 2e-6f);              // metrics.py:519,:523 on d = val/2
        v1 = fminf(fmaxf(fabsf(v1), 2e-6f), 2.0f - 2e-6f);
    }
    // :619  (a_mean > EPS || b_mean > EPS) && a_mean != b_mean.  Two absent samples both hold exactly
    // mean + EPS = EPS, so they are already excluded by the inequality; a present sample has mean + EPS > EPS.
    // The presence flag carried in the sign of .w is therefore not needed here.
    const bool c0 = lo(ax) != b.x, c1 = hi(ax) != b.x;
    const float l0 = mufu_lg2(v0), l1 = mufu_lg2(v1);                    // NaN for v < 0, -inf for 0
    if (c0) { acc0 += l0; cnt0 += 1.0f; }                                // compiled to predicated FADDs
    if (c1) { acc1 += l1; cnt1 += 1.0f; }
}

template <int VARIANT>
__device__ __forceinline__ float depth_finish(float acc, float cnt) {
    float d = 0.5f * mufu_ex2(__fdividef(acc, cnt));                     // :656-663
    if (!(cnt > 0.f)) d = 1.0f;                                          // :665
    if (d != d) d = 1.0f;                                                // :660, embedder.rs:166
    return d;
}

// samples per stage: the largest chunk that still lets two CTAs share an SM, else the largest that fits at all
static inline uint32_t pick_chunk(uint32_t S, size_t fixed_bytes, size_t &smem) {
    const size_t per_sample = (size_t)RG_TILE * sizeof(float4);             // 512 B per sample per stage
    const size_t budgets[2] = {(size_t)113 * 1024, (size_t)227 * 1024};     // 2 CTAs / SM, 1 CTA / SM
    for (size_t budget : budgets) {
        if (fixed_bytes + 2 * per_sample > budget) continue;
        uint32_t sc = (uint32_t)std::min<size_t>(S, (budget - fixed_bytes) / (2 * per_sample));
        if (sc < S && sc < 8 && budget == budgets[0]) continue;             // tiny stages: one CTA per SM is better
        // even out the chunks (e.g. S = 100, max 30 -> 4 chunks of 25)
        const uint32_t chunks = (S + sc - 1) / sc;
        sc = (S + chunks - 1) / chunks;
        smem = fixed_bytes + 2 * per_sample * sc;
        return sc;
    }
    return 0;
}

}  // namespace depthk
