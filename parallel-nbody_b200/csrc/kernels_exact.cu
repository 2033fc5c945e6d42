// kernels_exact.cu -- EXACT-mode instantiations.  Build with -fmad=false (see Makefile):
// bit-identical to reference nbody-seq.c compiled without FMA contraction and linked
// against include/nbody_detmath.h (oracle/_ref/nbody-seq-det-hex).
#include <cstdlib>

#include "step_kernels.cuh"
#include "forces_exact_wide.cuh"
#include "render.cuh"
#include "smalln.cuh"

namespace nbody {

int launch_forces_exact(const ForceArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return 0;
#ifdef NBODY_DEV
    // development knob: NBODY_EXACT_VARIANT=1 selects the one-thread-per-body kernel (same bits)
    static const int variant = getenv("NBODY_EXACT_VARIANT") ? atoi(getenv("NBODY_EXACT_VARIANT")) : 0;
    if (variant == 1) {
        dim3 grid((rows + FORCE_THREADS - 1) / FORCE_THREADS, 1);
        forces_kernel<MODE_EXACT><<<grid, FORCE_THREADS, 0, st>>>(a);
        return 0;
    }
#endif
    static bool attr_set_dev[64] = {};   // the attribute is per device
    int dev = 0;
    cudaGetDevice(&dev);
    bool &attr_set = attr_set_dev[dev & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(forces_exact_wide_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, exw_smem_bytes(32));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(forces_exact_wide_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, exw_smem_bytes(16));
        if (e != cudaSuccess) return static_cast<int>(e);
        attr_set = true;
    }
    // 16 bodies per CTA while 32 would not even give every SM (148 on a B200) two CTAs: N = 4096 launches 256 CTAs
    // instead of 128.  The choice cannot change a bit (see the kernel), so it may depend on the shard size.
    if ((rows + EXW_ROWS - 1) / EXW_ROWS < 2 * 148)
        forces_exact_wide_kernel<16><<<(rows + 15) / 16, EXW_THREADS, exw_smem_bytes(16), st>>>(a);
    else
        forces_exact_wide_kernel<32><<<(rows + 31) / 32, EXW_THREADS, exw_smem_bytes(32), st>>>(a);
    return 0;
}

// Cluster size of the one-launch step loop: 1 CTA up to SMALL_N_ONE_CTA bodies (no cluster barrier at all), else the
// largest of 16 / 8 CTAs the device can co-schedule with this much shared memory (16 needs the non-portable
// opt-in and a GPC with 16 free SMs).  The result does not depend on it: every body's fold order is fixed.
static int small_cluster_for(int n, int device)
{
    if (n <= SMALL_N_ONE_CTA) return 1;
    static int best[64] = {};
    int &b = best[device & 63];
    if (b) return b;
    b = SMALL_CLUSTER;
    if (cudaFuncSetAttribute(small_world_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess &&
        cudaFuncSetAttribute(small_world_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             static_cast<int>(small_smem_bytes(SMALL_N_MAX, 16))) == cudaSuccess) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(16);
        cfg.blockDim = dim3(SMALL_THREADS);
        cfg.dynamicSmemBytes = small_smem_bytes(SMALL_N_MAX, 16);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 16;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int nclusters = 0;
        if (cudaOccupancyMaxActiveClusters(&nclusters, small_world_kernel, &cfg) == cudaSuccess && nclusters >= 1) b = 16;
    }
    cudaGetLastError();
    return b;
}

int launch_small_world(const SmallArgs &a, cudaStream_t st)
{
    if (a.n < 1 || a.n > SMALL_N_MAX) return static_cast<int>(cudaErrorInvalidValue);
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned c = static_cast<unsigned>(small_cluster_for(a.n, dev));
    const size_t smem = small_smem_bytes(a.n, static_cast<int>(c));
    cudaError_t e = cudaFuncSetAttribute(small_world_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return static_cast<int>(e);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(c);
    cfg.blockDim = dim3(SMALL_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = c;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, small_world_kernel, a);
    return static_cast<int>(e);
}

void launch_integrate_exact(const IntegrateArgs &a, cudaStream_t st)
{
    const int rows = a.hi - a.lo;
    if (rows <= 0) return;
    integrate_kernel<MODE_EXACT><<<(rows + 127) / 128, 128, 0, st>>>(a);
}

void launch_render(const double2 *pos, const double2 *mr, int n, int xdim, int ydim, int *owner, unsigned char *rgb,
                   cudaStream_t st)
{
    const int npix = xdim * ydim;
    render_clear_kernel<<<(npix + 255) / 256, 256, 0, st>>>(owner, npix);
    render_cover_kernel<<<n, 128, 0, st>>>(pos, mr, n, xdim, ydim, owner);
    render_paint_kernel<<<(npix + 255) / 256, 256, 0, st>>>(owner, npix, n, rgb);
}

}  // namespace nbody
