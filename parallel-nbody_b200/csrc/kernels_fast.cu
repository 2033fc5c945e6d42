// kernels_fast.cu -- FAST-mode instantiations (FMA contraction on, algebraic force direction).
#include <cstdlib>

#include "step_kernels.cuh"
#include "forces_fast.cuh"
#include "forces_sym.cuh"
#include "p2p.cuh"

namespace nbody {

void launch_forces_fast(const ForceArgs &a, int nsplit, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    // 2 bodies per thread, 4-way unrolled j loop, integer max for the clamp (the round-1 sweep of ten shapes:
    // profiles/r1_fast_variant_sweep.log)
    dim3 grid((rows + FORCE_THREADS * FAST_IB - 1) / (FORCE_THREADS * FAST_IB), nsplit);
    forces_fast_kernel<FAST_IB, 4, 1, true><<<grid, FORCE_THREADS, 0, st>>>(a);
}

void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    const int threads = a.boxes ? TILE_J : 128;   // fused tile boxes: one CTA per tile of the cell order
    integrate_kernel<MODE_FAST><<<(rows + threads - 1) / threads, threads, 0, st>>>(a);
}

// Two shapes of the symmetric kernel, chosen by n alone (like every other choice that fixes the summation
// order): 64 threads x 8 bodies (2 warps per CTA; one LDS pair, one accumulator rotation and one address update
// serve 8 pairs) from SYM_WIDE_N bodies on, 128 x 4 below (more, shorter CTAs: better tail when the launch has few).
// NBODY_DEV builds can pin either shape with NBODY_SYM_VARIANT=1|2 for A/B timing.
int launch_forces_sym(const SymArgs &a, int nitems, cudaStream_t st)
{
    if (nitems <= 0) return 0;
    bool wide = a.n >= SYM_WIDE_N;
#ifdef NBODY_DEV
    static const int env_variant = getenv("NBODY_SYM_VARIANT") ? atoi(getenv("NBODY_SYM_VARIANT")) : 0;
    if (env_variant == 1) wide = false;
    if (env_variant == 2) wide = true;
#endif
    const dim3 grid(nitems);
    static bool attr_set_dev[64] = {};   // the attribute is per device
    int dev = 0;
    cudaGetDevice(&dev);
    bool &attr_set = attr_set_dev[dev & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(forces_sym_kernel<64, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, sym_smem_bytes(64));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(forces_sym_kernel<128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, sym_smem_bytes(128));
        if (e != cudaSuccess) return static_cast<int>(e);
        attr_set = true;
    }
    if (wide) forces_sym_kernel<64, 8><<<grid, 64, sym_smem_bytes(64), st>>>(a);
    else forces_sym_kernel<128, 4><<<grid, 128, sym_smem_bytes(128), st>>>(a);
    return 0;
}

void launch_sym_reduce(const SymReduceArgs &a, cudaStream_t st)
{
    if (a.n <= 0 || a.nv <= 0) return;
    dim3 grid((a.n + 127) / 128, a.nv);
    sym_reduce_kernel<<<grid, 128, 0, st>>>(a);
}

void launch_sym_reduce_p2p(const SymReduceArgs &a, const PeerTable &t, int rows_per_shard, cudaStream_t st)
{
    if (a.n <= 0 || a.nv <= 0) return;
    dim3 grid((a.n + 127) / 128, a.nv);
    sym_reduce_p2p_kernel<<<grid, 128, 0, st>>>(a, t, rows_per_shard);
}

void launch_p2p_barrier(const PeerTable &t, int phase, int *error_word, cudaStream_t st)
{
    p2p_barrier_kernel<<<1, 32, 0, st>>>(t, phase, error_word);
}

void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    reduce_partials_kernel<MODE_FAST><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

}  // namespace nbody
