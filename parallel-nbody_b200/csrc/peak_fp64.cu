// peak_fp64.cu -- measured FP64 (non-tensor) peak: the roofline denominator for the force
// kernel.  MEASURED_PEAKS.json carries HBM and bf16 numbers only, so the library measures
// the DFMA rate itself: every thread runs 8 independent fused multiply-add chains from
// registers, no memory traffic, enough warps per SM to cover the pipe latency.
#include <cuda_runtime.h>

#include <cstdio>

#include "../../include/nbody_gpu.h"

namespace {

constexpr int CHAINS = 8;
constexpr int INNER = 256;

__global__ void __launch_bounds__(256) dfma_kernel(double *sink, int outer, long long *cycles)
{
    double a[CHAINS];
    const double m = 1.0000000001, c = 1e-12;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) a[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    const long long t0 = clock64();
    for (int o = 0; o < outer; ++o) {
#pragma unroll
        for (int it = 0; it < INNER; ++it) {
#pragma unroll
            for (int k = 0; k < CHAINS; ++k) a[k] = fma(a[k], m, c);
        }
    }
    const long long t1 = clock64();
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += a[k];
    if (s == 123.456) sink[0] = s;   // keep the chains alive
    if (blockIdx.x == 0 && threadIdx.x == 0) cycles[0] = t1 - t0;
}

}  // namespace

extern "C" int nbody_gpu_fp64_peak(int device, int ms_budget, double *tflops, double *sm_mhz_est)
{
    if (!tflops) return NBODY_GPU_EINVAL;
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return NBODY_GPU_ECUDA;
    cudaDeviceProp prop;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return NBODY_GPU_ECUDA;
    const int blocks = prop.multiProcessorCount * 4;   // 4 x 256 threads = 32 warps per SM
    double *sink = nullptr;
    long long *cyc = nullptr;
    if (cudaMalloc(&sink, sizeof(double)) != cudaSuccess || cudaMalloc(&cyc, sizeof(long long)) != cudaSuccess)
        return NBODY_GPU_ECUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    if (ms_budget < 5) ms_budget = 5;
    int outer = 64;
    double best = 0.0, mhz = 0.0;
    // warm up, then grow the run until it is about ms_budget long; keep the best rate seen
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        dfma_kernel<<<blocks, 256>>>(sink, outer, cyc);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) return NBODY_GPU_ECUDA;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        long long cycles = 0;
        cudaMemcpy(&cycles, cyc, sizeof cycles, cudaMemcpyDeviceToHost);
        const double fmas = static_cast<double>(blocks) * 256.0 * CHAINS * INNER * outer;
        const double tf = 2.0 * fmas / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) { best = tf; mhz = cycles / (ms * 1e-3) / 1e6; }
        if (ms < ms_budget * 0.5f && outer < (1 << 20)) outer *= 2;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    cudaFree(cyc);
    if (cudaGetLastError() != cudaSuccess) return NBODY_GPU_ECUDA;
    *tflops = best;
    if (sm_mhz_est) *sm_mhz_est = mhz;
    return NBODY_GPU_OK;
}
