// kernels_fast.cu -- FAST-mode instantiations (FMA contraction on, algebraic force direction).
#include "step_kernels.cuh"
#include "forces_fast.cuh"

namespace nbody {

void launch_forces_fast(const ForceArgs &a, int nsplit, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    dim3 grid((rows + FAST_ROWS_PER_CTA - 1) / FAST_ROWS_PER_CTA, nsplit);
    forces_fast_kernel<<<grid, FORCE_THREADS, 0, st>>>(a);
}

void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    integrate_kernel<MODE_FAST><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    reduce_partials_kernel<MODE_FAST><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

}  // namespace nbody
