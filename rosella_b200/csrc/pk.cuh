// pk.cuh — packed pair of FP32 values driven by the sm_100 f32x2 instructions (SASS FFMA2 / FMUL2 /
// FADD2).  One warp instruction performs the operation for two values per lane, so the FMA-pipe
// work of two (query, reference) evaluations costs one issue slot instead of two.  ptxas folds
// bc(x) (same scalar in both halves) into the instruction's broadcast operand form (R.F32), and
// fabs / negation into source modifiers, so no register moves are generated.
#pragma once
#include <cstdint>

struct pk {
    unsigned long long v;
};

__device__ __forceinline__ pk mk(float lo, float hi) {
    pk r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ pk bc(float x) { return mk(x, x); }
__device__ __forceinline__ void un(pk a, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
}
__device__ __forceinline__ float lo(pk a) { float l, h; un(a, l, h); return l; }
__device__ __forceinline__ float hi(pk a) { float l, h; un(a, l, h); return h; }
__device__ __forceinline__ pk operator+(pk a, pk b) {
    pk r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ pk operator-(pk a, pk b) {
    pk r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ pk operator*(pk a, pk b) {
    pk r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ pk fma2(pk a, pk b, pk c) {
    pk r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
    return r;
}
__device__ __forceinline__ pk neg(pk a) { return mk(-lo(a), -hi(a)); }
__device__ __forceinline__ pk abs2(pk a) { return mk(fabsf(lo(a)), fabsf(hi(a))); }
