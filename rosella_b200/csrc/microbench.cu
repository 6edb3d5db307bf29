// microbench.cu — measures the non-tensor FP32 pipe peaks that bound the kNN select kernels
// (MEASURED_PEAKS.json has no FP32 CUDA-core entry; SURVEY §8d asks the builder to measure one).
// Standalone: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
// Prints one JSON object: warp-instruction issue rates and the derived TFLOP/s.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int ITERS = 4096;
constexpr int CHAINS = 8;

__global__ void k_ffma(float *out, float a, float b) {
    float x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) x[c] = threadIdx.x * 1e-6f + c;
    for (int i = 0; i < ITERS; ++i)
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) x[c] = fmaf(x[c], a, b);
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

__global__ void k_ffma2(float *out, float a, float b) {
    unsigned long long x[CHAINS], aa, bb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        float v = threadIdx.x * 1e-6f + c;
        asm("mov.b64 %0, {%1, %1};" : "=l"(x[c]) : "f"(v));
    }
    for (int i = 0; i < ITERS; ++i)
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[c]) : "l"(aa), "l"(bb));
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[c]));
        s += lo + hi;
    }
    if (s == 12345.678f) out[0] = s;
}

template <int OP>   // 0 ex2, 1 rcp, 2 lg2, 3 sqrt
__global__ void k_mufu(float *out) {
    float x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) x[c] = 0.5f + threadIdx.x * 1e-6f + c * 0.01f;
    for (int i = 0; i < ITERS; ++i)
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[c]));
            if (OP == 1) {   // rcp(rcp(x)) = x lets ptxas fold the chain away: break the identity with one FFMA (FMA pipe, not XU)
                asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[c]));
                x[c] = fmaf(x[c], 1.0000001f, 1e-9f);
            }
            if (OP == 2) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[c]));
            if (OP == 3) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(x[c]));
        }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

// the depth kernel's mix: R FFMA per MUFU (and optionally per ALU op)
template <int R, int ALU>
__global__ void k_mix(float *out, float a, float b) {
    float x[CHAINS], y[CHAINS];
    int z[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { x[c] = threadIdx.x * 1e-6f + c; y[c] = 0.5f + c * 0.01f; z[c] = threadIdx.x + c; }
    for (int i = 0; i < ITERS / 4; ++i)
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
#pragma unroll
            for (int r = 0; r < R; ++r) x[c] = fmaf(x[c], a, b);
#pragma unroll
            for (int r = 0; r < ALU; ++r) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[c]) : "r"(i), "r"(r));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(y[c]));
        }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += x[c] + y[c] + z[c];
    if (s == 12345.678f) out[0] = s;
}

template <typename F>
static double time_ms(F launch) {
    cudaEvent_t a, b;
    CHECK(cudaEventCreate(&a)); CHECK(cudaEventCreate(&b));
    launch(); launch();
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CHECK(cudaEventRecord(a));
        launch();
        CHECK(cudaEventRecord(b));
        CHECK(cudaEventSynchronize(b));
        float ms; CHECK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CHECK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0));
    float *out; CHECK(cudaMalloc(&out, 4));
    const int blocks = p.multiProcessorCount * 8, threads = 256;
    const double warps = (double)blocks * threads / 32.0;
    auto rate = [&](double ms, double instr_per_thread) { return warps * instr_per_thread / (ms * 1e-3); };   // warp-instr/s
    const double n = (double)ITERS * CHAINS;
    double t_ffma = time_ms([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    double t_ffma2 = time_ms([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    double t_ex2 = time_ms([&] { k_mufu<0><<<blocks, threads>>>(out); });
    double t_rcp = time_ms([&] { k_mufu<1><<<blocks, threads>>>(out); });
    double t_lg2 = time_ms([&] { k_mufu<2><<<blocks, threads>>>(out); });
    double t_sqrt = time_ms([&] { k_mufu<3><<<blocks, threads>>>(out); });
    double t_mix8 = time_ms([&] { k_mix<8, 0><<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    double t_mix8a = time_ms([&] { k_mix<8, 3><<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    double t_mix4a = time_ms([&] { k_mix<4, 2><<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    const double nm = (double)(ITERS / 4) * CHAINS;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz_nominal\": %d,\n", p.name, p.multiProcessorCount, p.clockRate / 1000);
    printf(" \"ffma_tflops\": %.2f, \"ffma_winst_per_s\": %.4g,\n", rate(t_ffma, n) * 64 / 1e12, rate(t_ffma, n));
    printf(" \"ffma2_tflops\": %.2f, \"ffma2_winst_per_s\": %.4g,\n", rate(t_ffma2, n) * 128 / 1e12, rate(t_ffma2, n));
    printf(" \"mufu_ex2_winst_per_s\": %.4g, \"mufu_rcp_winst_per_s\": %.4g, \"mufu_lg2_winst_per_s\": %.4g, \"mufu_sqrt_winst_per_s\": %.4g,\n",
           rate(t_ex2, n), rate(t_rcp, n), rate(t_lg2, n), rate(t_sqrt, n));
    printf(" \"mix_8ffma_1mufu_winst_per_s\": %.4g, \"mix_8ffma_3alu_1mufu_winst_per_s\": %.4g, \"mix_4ffma_2alu_1mufu_winst_per_s\": %.4g}\n",
           rate(t_mix8, nm * 9), rate(t_mix8a, nm * 12), rate(t_mix4a, nm * 7));
    return 0;
}
