// knn_tnf_common.cuh — the FP32 form of the TNF metrics shared by the two select kernels
// (knn_tnf.cu: FP32 CUDA cores; knn_tnf_tc.cu: tcgen05 tensor cores for the dot product).
#pragma once
#include "rg_internal.cuh"

// VARIANT 2 = squared Euclidean distance of the rows as passed (RG_GRAPH_TNF_EUCLID; flight/flight/metrics.py:64-77):
// |x - y|^2 = |x~|^2 + |y~|^2 - 2 x~.y~ + d (mean_x - mean_y)^2 on the centred rows; selection only needs the order,
// so the square root is left to the f64 re-evaluation.
constexpr int METRIC_EUCLID = 2;

template <int VARIANT>
__device__ __forceinline__ float tnf_finish(float dot, float nq, float nr, float dmu2) {
    const float den = nq + nr;                      // NaN when either row is flagged non-finite
    float d;
    if (VARIANT == METRIC_EUCLID) {
        d = fmaxf(fmaf(-2.0f, dot, den) + dmu2, 0.f);
        if (den != den) d = 3.402823466e+38f;       // non-finite row: the defined maximum (include/rosella_gpu.h)
    } else if (VARIANT == RG_VARIANT_RUST) {
        d = 0.5f - __fdividef(dot, den);
        if (dot == 0.f) d = 1.0f;                   // :385
        if (nq == 0.f && nr == 0.f) d = 0.f;        // :383
        if (d != d) d = 1.0f;                       // :410
    } else {
        d = 1.0f - 2.0f * __fdividef(dot, den);
        if (d != d) d = 2.0f;                       // defined here as the metric's maximum (include/rosella_gpu.h)
    }
    return d;
}

