// Development aid: error statistics of the FP64 reciprocal-square-root seed (MUFU.RSQ64H, PTX rsqrt.approx.ftz.f64)
// over the mantissa range: e = 1 - q*y0^2.  The far-field pair of forces_sym.cuh corrects the seed to first order,
// y0*(1 + e/2); the dropped term 3e^2/8 is one-sided, and its MEAN over the seed's error distribution is the
// constant SYM_RSQRT_BIAS that the kernel folds into its final multiply by G (so that the pair terms are unbiased).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/rsqrt_seed_probe tools/rsqrt_seed_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void probe(unsigned long long n, double *out, int log_uniform)   // out: sum e, sum e^2, max |e|, sum of true relative error of y0(1+e/2)
{
    double se = 0, se2 = 0, me = 0, sr = 0;
    for (unsigned long long k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (unsigned long long)gridDim.x * blockDim.x) {
        // q over [1, 4): every mantissa/exponent-parity combination the seed table can see, quasi-randomly
        unsigned long long h = k * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29;
        const double u = (h >> 11) * (1.0 / 9007199254740992.0);
        const double q = log_uniform ? exp2(2.0 * u) : 1.0 + 3.0 * u;   // 4^u: log-uniform (products of many factors), or uniform
        double y0;
        asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(q));
        const double t = q * y0;
        const double e = fma(-t, y0, 1.0);
        const double y1 = fma(y0 * e, 0.5, y0);
        const double exact = 1.0 / sqrt(q);
        se += e;
        se2 += e * e;
        me = fmax(me, fabs(e));
        sr += (y1 - exact) / exact;
    }
    atomicAdd(&out[0], se);
    atomicAdd(&out[1], se2);
    atomicAdd(&out[3], sr);
    unsigned long long *m = reinterpret_cast<unsigned long long *>(&out[2]);
    atomicMax(m, (unsigned long long)__double_as_longlong(me));
}

int main()
{
    double *d, h[4] = {0, 0, 0, 0};
    const unsigned long long n = 1ull << 30;
    cudaMalloc(&d, sizeof h);
    cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
    for (int lu = 0; lu < 2; ++lu) {
        double z[4] = {0, 0, 0, 0};
        cudaMemcpy(d, z, sizeof z, cudaMemcpyHostToDevice);
        probe<<<148 * 8, 256>>>(n, d, lu);
        if (cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("CUDA error\n"); return 1; }
        printf("q %s over [1,4): samples %llu  mean e %.4e  rms e %.4e  max |e| %.4e (2^%.2f)\n", lu ? "log-uniform" : "uniform", n,
               h[0] / n, sqrt(h[1] / n), h[2], log2(h[2]));
        printf("  mean of the dropped term 3e^2/8 = %.6e ; measured mean relative error of y0*(1+e/2) = %.6e\n", 0.375 * h[1] / n, h[3] / n);
    }
    return 0;
}
