// microbench2.cu — issue-cost model of the SM sub-partition for the instruction mixes of the kNN
// select kernels: how many cycles does an iteration of {NF FFMA, NF2 FFMA2, NA LOP3, NM MUFU.EX2}
// take per warp when 16 warps share one scheduler?  Prints cycles per iteration per SMSP.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int ITERS = 2048;

template <int NF, int NF2, int NA, int NM>
__global__ void k_mix(float *out, float a, float b) {
    float x[8], y[4];
    unsigned long long p[8], aa, bb;
    int z[8];
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        x[c] = threadIdx.x * 1e-6f + c;
        z[c] = threadIdx.x + c;
        asm("mov.b64 %0, {%1, %1};" : "=l"(p[c]) : "f"(x[c]));
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) y[c] = 0.5f + c * 0.01f;
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int r = 0; r < NF; ++r) x[r & 7] = fmaf(x[r & 7], a, b);
#pragma unroll
        for (int r = 0; r < NF2; ++r) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[r & 7]) : "l"(aa), "l"(bb));
#pragma unroll
        for (int r = 0; r < NA; ++r) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[r & 7]) : "r"(i), "r"(r));
#pragma unroll
        for (int r = 0; r < NM; ++r) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(y[r & 3]));
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[c]));
        s += x[c] + z[c] + lo + hi;
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) s += y[c];
    if (s == 12345.678f) out[0] = s;
}

static float *g_out;
static int g_blocks;
static double g_clock_hz;

template <int NF, int NF2, int NA, int NM>
static void run(const char *name, bool last = false) {
    cudaEvent_t a, b;
    CHECK(cudaEventCreate(&a)); CHECK(cudaEventCreate(&b));
    for (int i = 0; i < 2; ++i) k_mix<NF, NF2, NA, NM><<<g_blocks, 256>>>(g_out, 1.0001f, 0.5f);
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CHECK(cudaEventRecord(a));
        k_mix<NF, NF2, NA, NM><<<g_blocks, 256>>>(g_out, 1.0001f, 0.5f);
        CHECK(cudaEventRecord(b));
        CHECK(cudaEventSynchronize(b));
        float ms; CHECK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    // warps per SMSP = blocks*8 / (SMs*4); cycles per iteration per warp-slot
    const double warps_per_smsp = (double)g_blocks * 8.0 / (148.0 * 4.0);
    const double cyc = best * 1e-3 * g_clock_hz / (ITERS * warps_per_smsp);
    printf(" \"%s\": {\"ffma\": %d, \"ffma2\": %d, \"alu\": %d, \"mufu\": %d, \"cycles_per_iter\": %.2f}%s\n", name, NF, NF2, NA, NM,
           cyc, last ? "" : ",");
}

int main() {
    cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0));
    CHECK(cudaMalloc(&g_out, 4));
    g_blocks = p.multiProcessorCount * 8;
    g_clock_hz = p.clockRate * 1e3;
    printf("{\"note\": \"cycles per iteration per warp at nominal %d MHz, 16 warps per scheduler\",\n", p.clockRate / 1000);
    run<16, 0, 0, 0>("f16");
    run<0, 16, 0, 0>("p16");
    run<0, 0, 16, 0>("a16");
    run<0, 0, 0, 4>("m4");
    run<16, 0, 16, 0>("f16_a16");
    run<0, 16, 16, 0>("p16_a16");
    run<0, 16, 32, 0>("p16_a32");
    run<16, 0, 0, 2>("f16_m2");
    run<0, 16, 0, 2>("p16_m2");
    run<0, 16, 0, 4>("p16_m4");
    run<0, 0, 16, 2>("a16_m2");
    run<0, 0, 16, 4>("a16_m4");
    run<0, 16, 8, 4>("p16_a8_m4");
    run<0, 28, 16, 7>("p28_a16_m7");
    run<57, 0, 19, 7>("f57_a19_m7");
    run<8, 24, 16, 7>("f8_p24_a16_m7", true);
    printf("}\n");
    return 0;
}
