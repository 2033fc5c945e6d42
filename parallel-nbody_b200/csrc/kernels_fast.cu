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
    // development knob: NBODY_FAST_VARIANT selects a tuning variant (default 0)
    static const int variant = getenv("NBODY_FAST_VARIANT") ? atoi(getenv("NBODY_FAST_VARIANT")) : 0;
#define NB_LAUNCH(IB, UN, MINB, IMAX)                                                            \
    {                                                                                             \
        dim3 grid((rows + FORCE_THREADS * IB - 1) / (FORCE_THREADS * IB), nsplit);                \
        forces_fast_kernel<IB, UN, MINB, IMAX><<<grid, FORCE_THREADS, 0, st>>>(a);                \
    }
    switch (variant) {
    default: NB_LAUNCH(2, 4, 1, true) break;
    case 10: NB_LAUNCH(2, 4, 1, false) break;
    case 1: NB_LAUNCH(2, 4, 8, false) break;
    case 2: NB_LAUNCH(2, 2, 1, false) break;
    case 3: NB_LAUNCH(4, 2, 1, false) break;
    case 4: NB_LAUNCH(1, 8, 1, false) break;
    case 6: NB_LAUNCH(2, 4, 8, true) break;
    case 7: NB_LAUNCH(4, 2, 1, true) break;
    case 8: NB_LAUNCH(2, 8, 1, false) break;
    case 9: NB_LAUNCH(1, 4, 1, true) break;
    }
#undef NB_LAUNCH
}

void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    integrate_kernel<MODE_FAST><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

int launch_forces_sym(SymArgs a, int nblocks_here, int nsplit, cudaStream_t st)
{
    if (nblocks_here <= 0) return 0;
    a.nblocks_here = nblocks_here;
    a.nsplit = nsplit;
    a.nitems = nblocks_here * nsplit;
    // development knob: NBODY_SYM_VARIANT pins a tuning variant; default: 8 bodies per thread (2 warps per CTA) for
    // large worlds, where halving the per-step LDS/SHFL/address work wins 2.7 %, 4 bodies per thread below
    // (more, shorter CTAs: better tail at N = 65536).  A function of n only, like every other choice that
    // affects the summation order.
    static const int env_variant = getenv("NBODY_SYM_VARIANT") ? atoi(getenv("NBODY_SYM_VARIANT")) : 0;
    const int variant = env_variant ? env_variant : (a.n >= SYM_WIDE_N ? 22 : 0);
    dim3 grid(nblocks_here, nsplit);
    cudaError_t e;
#define NB_SYM_LAUNCH(...)                                                                                          \
    e = cudaFuncSetAttribute(forces_sym_kernel<__VA_ARGS__>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);    \
    if (e != cudaSuccess) return static_cast<int>(e);                                                              \
    forces_sym_kernel<__VA_ARGS__><<<grid, threads, smem, st>>>(a);
    int threads = 128;
    if (variant == 1) threads = 256;
    if (variant == 20 || variant == 21 || variant == 22 || variant == 30) threads = 64;
    const int smem = sym_smem_bytes(threads);
    // persistent variants: one wave of CTAs (as many as are resident at once) pulling items from a.counter
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
#define NB_SYM_LAUNCH_P(CTAS_PER_SM, ...)                                                                           \
    if (!a.counter) return static_cast<int>(cudaErrorInvalidValue);                                                 \
    e = cudaFuncSetAttribute(forces_sym_kernel<__VA_ARGS__>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);    \
    if (e != cudaSuccess) return static_cast<int>(e);                                                              \
    e = cudaMemsetAsync(a.counter, 0, sizeof(int), st);                                                             \
    if (e != cudaSuccess) return static_cast<int>(e);                                                              \
    forces_sym_kernel<__VA_ARGS__><<<min(a.nitems, sms * CTAS_PER_SM), threads, smem, st>>>(a);
    switch (variant) {
    default: NB_SYM_LAUNCH(128, 4) break;             // near flavour 3, far flavour 5; default for n < SYM_WIDE_N
    case 1: NB_SYM_LAUNCH(256, 2) break;
    case 4: NB_SYM_LAUNCH(128, 4, 1, 2) break;
    case 5: NB_SYM_LAUNCH(128, 4, 1, 8) break;
    case 10: NB_SYM_LAUNCH(128, 4, 1, 4, 0, 4) break;  // round-1 pair arithmetic
    case 20: NB_SYM_LAUNCH(64, 8, 1, 1) break;
    case 21: NB_SYM_LAUNCH(64, 8, 1, 2) break;
    case 22: NB_SYM_LAUNCH(64, 8, 1, 4) break;         // default for n >= SYM_WIDE_N
    case 23: NB_SYM_LAUNCH(128, 4) break;              // the 128 x 4 shape at any n
    case 30: NB_SYM_LAUNCH_P(4, 64, 8, 1, 4, 3, 5, true) break;
    case 31: NB_SYM_LAUNCH_P(2, 128, 4, 1, 4, 3, 5, true) break;
    }
#undef NB_SYM_LAUNCH
#undef NB_SYM_LAUNCH_P
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

void launch_p2p_signal(const PeerTable &t, int phase, unsigned long long epoch, cudaStream_t st)
{
    p2p_signal_kernel<<<1, 32, 0, st>>>(t, phase, epoch);
}

void launch_p2p_wait(const PeerTable &t, int phase, unsigned long long epoch, int *error_word, cudaStream_t st)
{
    p2p_wait_kernel<<<1, 32, 0, st>>>(t, phase, epoch, error_word);
}

void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    reduce_partials_kernel<MODE_FAST><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

}  // namespace nbody
