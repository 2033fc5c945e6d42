// fp64_probe.cu -- development probe: what FP64-pipe rate is reachable with register operands?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_probe tools/fp64_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int INNER = 128;

// (a) one register + two immediates/constants per DFMA (what nbody_gpu_fp64_peak measures)
__global__ void __launch_bounds__(256) k_imm(double *sink, int outer)
{
    double a[8];
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    for (int o = 0; o < outer; ++o)
#pragma unroll
        for (int it = 0; it < INNER; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fma(a[k], 1.0000000001, 1e-12);
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;
}
// (b) three distinct register operands per DFMA
__global__ void __launch_bounds__(256) k_rrr(double *sink, int outer, double m0, double c0)
{
    double a[8], m[8], c[8];
    for (int k = 0; k < 8; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k); m[k] = m0 + k * 1e-12; c[k] = c0 * (k + 1); }
    for (int o = 0; o < outer; ++o)
#pragma unroll
        for (int it = 0; it < INNER; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = fma(a[k], m[k], c[(k + it) & 7]);
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;
}
// (c) a mix like the force loop: DADD, DMUL, DFMA with register operands, results feeding each other
__global__ void __launch_bounds__(256) k_mix(double *sink, int outer, double m0, double c0)
{
    double a[8], b[8];
    for (int k = 0; k < 8; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k); b[k] = m0 + k * 1e-3; }
    for (int o = 0; o < outer; ++o)
#pragma unroll
        for (int it = 0; it < INNER / 4; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double d = b[k] - a[k];          // DADD
                double e = d * d;                // DMUL
                double f = fma(d, b[(k + 1) & 7], e);   // DFMA
                a[k] = fma(f, c0, a[k]);         // DFMA
            }
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;
}

// (d) the mix plus integer work: per 4 FP64 ops, ALU integer ops (min on the high word + selects)
template <int NALU>
__global__ void __launch_bounds__(256) k_mix_alu(double *sink, int outer, double m0, double c0)
{
    double a[8], b[8];
    int mn[4] = {0x7fffffff, 0x7ffffff0, 0x7fffff00, 0x7ffff000};
    for (int k = 0; k < 8; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k); b[k] = m0 + k * 1e-3; }
    for (int o = 0; o < outer; ++o)
#pragma unroll
        for (int it = 0; it < INNER / 4; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double d = b[k] - a[k];
                double e = d * d;
                double f = fma(d, b[(k + 1) & 7], e);
                a[k] = fma(f, c0, a[k]);
#pragma unroll
                for (int q = 0; q < NALU; ++q) mn[q] = min(mn[q], __double2hiint(q & 1 ? e : f) + q);
            }
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456 || mn[0] + mn[1] + mn[2] + mn[3] == 77) sink[0] = s;
}
// (e) the mix plus one MUFU.RSQ64H per 4 FP64 ops
__global__ void __launch_bounds__(256) k_mix_mufu(double *sink, int outer, double m0, double c0)
{
    double a[8], b[8];
    for (int k = 0; k < 8; ++k) { a[k] = 1.0 + 1e-9 * (threadIdx.x + k); b[k] = m0 + k * 1e-3; }
    for (int o = 0; o < outer; ++o)
#pragma unroll
        for (int it = 0; it < INNER / 4; ++it)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double d = b[k] - a[k];
                double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
                double e = d * y;
                double f = fma(d, b[(k + 1) & 7], e);
                a[k] = fma(f, c0, a[k]);
            }
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;
}
// (f) the mix plus LDS.128 broadcast loads feeding it (one per 8 FP64 ops)
__global__ void __launch_bounds__(256) k_mix_lds(double *sink, int outer, double m0, double c0)
{
    __shared__ double2 sm[256];
    sm[threadIdx.x] = make_double2(m0 + threadIdx.x * 1e-6, c0);
    __syncthreads();
    double a[8];
    for (int k = 0; k < 8; ++k) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    for (int o = 0; o < outer; ++o)
#pragma unroll 4
        for (int it = 0; it < INNER / 4; ++it) {
            const double2 v0 = sm[(it * 4) & 255], v1 = sm[(it * 4 + 1) & 255], v2 = sm[(it * 4 + 2) & 255], v3 = sm[(it * 4 + 3) & 255];
            const double bb[8] = {v0.x, v0.y, v1.x, v1.y, v2.x, v2.y, v3.x, v3.y};
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                double d = bb[k] - a[k];
                double e = d * d;
                double f = fma(d, bb[(k + 1) & 7], e);
                a[k] = fma(f, c0, a[k]);
            }
        }
    double s = 0; for (int k = 0; k < 8; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;
}

template <typename F> void run(const char *name, F launch, double ops_per_thread_per_outer)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = 148 * 4, outer = 2000;
    launch(blocks, 100);
    cudaEventRecord(e0); launch(blocks, outer); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double inst = (double)blocks * 256 * ops_per_thread_per_outer * outer;
    printf("%-6s %8.3f ms  %.3e FP64 lane-instr/s  (= %.2f TFLOP/s if all were FMA)\n", name, ms, inst / (ms * 1e-3),
           2 * inst / (ms * 1e-3) / 1e12);
}

int main()
{
    double *sink; cudaMalloc(&sink, 8);
    run("imm", [&](int b, int o) { k_imm<<<b, 256>>>(sink, o); }, 8.0 * INNER);
    run("rrr", [&](int b, int o) { k_rrr<<<b, 256>>>(sink, o, 1.0000000001, 1e-12); }, 8.0 * INNER);
    run("mix", [&](int b, int o) { k_mix<<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    run("alu1", [&](int b, int o) { k_mix_alu<1><<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    run("alu2", [&](int b, int o) { k_mix_alu<2><<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    run("alu4", [&](int b, int o) { k_mix_alu<4><<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    run("mufu", [&](int b, int o) { k_mix_mufu<<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    run("lds", [&](int b, int o) { k_mix_lds<<<b, 256>>>(sink, o, 1.5, 1e-12); }, 8.0 * INNER);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
