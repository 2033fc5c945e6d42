// step_args.h -- kernel argument blocks and host-side launchers of the N-body step.
// No device code: included by nbody_gpu.cu (the C ABI) and by step_kernels.cuh.
#pragma once
#include <cuda_runtime.h>

namespace nbody {

constexpr int MODE_EXACT = 0;
constexpr int MODE_FAST = 1;

constexpr int FORCE_THREADS = 128;  // bodies i per CTA (one per thread)
constexpr int TILE_J = 256;         // bodies j per shared-memory stage
constexpr int STAGES = 2;

struct ForceArgs {
    const double2 *pos;   // current positions, all n bodies
    const double2 *mr;    // mass, radius
    double2 *out;         // EXACT: frc (indexed by i); FAST: part[s * npad + i]
    int n;                // bodies (loop bound for j)
    int npad;             // leading dimension of part
    int lo, hi;           // rows owned by this device: i in [lo, hi)
    int chunk;            // FAST: j-chunk length; chunk s covers [s*chunk, min(n,(s+1)*chunk))
};

constexpr int SYM_RB = 512;         // rows per block of the symmetric (evaluate-each-pair-once) kernel

struct SymArgs {
    const double2 *pos;
    const double2 *mr;
    double2 *rowpart;     // [nsplit][npad]  row-side partial sums
    double2 *colpart;     // [nblocks][npad] reaction (column-side) partial sums, valid for j >= (B+1)*SYM_RB
    int n;
    int npad;
    int chunk;            // j-chunk length, a multiple of SYM_RB
    int block_first;      // this launch handles row blocks block_first + k*block_stride, k = 0..grid.x-1
    int block_stride;
};

struct IntegrateArgs {
    const double2 *pos_old;
    double2 *pos_new;
    const double2 *mr;
    double2 *vel;
    double2 *frc;          // FAST: written (sum of partials); EXACT: read
    const double2 *part;   // FAST: partial sums (row side)
    const double2 *colpart;  // FAST symmetric: reaction partial sums, else nullptr
    int chunk;             // FAST symmetric: j-chunk length
    int nsplit;            // FAST: number of j-chunks
    int npad;
    int lo, hi;
    int xdim, ydim;
};

// launchers, defined in kernels_exact.cu / kernels_fast.cu
void launch_forces_exact(const ForceArgs &a, cudaStream_t st);
void launch_integrate_exact(const IntegrateArgs &a, cudaStream_t st);
void launch_forces_fast(const ForceArgs &a, int nsplit, cudaStream_t st);
void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st);
void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st);
int launch_forces_sym(const SymArgs &a, int nblocks_here, int nsplit, cudaStream_t st);   // returns cudaError_t

}  // namespace nbody
