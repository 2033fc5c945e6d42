// p2p.cuh -- the per-step exchange fused into the kernels over NVLink peer memory (no NCCL in the step).
//
// With the symmetric kernel sharded over P GPUs (P | 8) the step needs two exchanges: every GPU's share of
// the 8 virtual-rank force vectors must reach the GPU that integrates body j, and the new positions must
// reach everybody (the reference: one MPI_Allgatherv of whole structs, nbody-par.c:913-917).  Here both
// are plain stores into peer memory issued by the producing kernels themselves:
//   sym_reduce_p2p_kernel  writes vbuf[v][j] straight into the vbuf of the GPU that owns row j,
//   integrate_kernel       writes pos_new[i] into every GPU's position buffer,
// each followed by a one-warp signal kernel (release store of the step's epoch into every peer's flag
// word) and, on the consumer side, a one-warp wait kernel (acquire loads until all P flags carry the
// epoch).  Positions are double-buffered and the two waits order the reuse of vbuf, so no buffer is
// overwritten while a peer still reads it.  A wait gives up after ~10 s and
// raises an error word instead of hanging the GPU.
#pragma once
#include <cuda_runtime.h>

#include "step_args.h"

namespace nbody {

__global__ void p2p_signal_kernel(const PeerTable t, int phase, unsigned long long epoch)
{
    const int r = threadIdx.x;
    if (r >= t.npeers) return;
    __threadfence_system();
    unsigned long long *dst = t.flags[r] + phase * MAX_PEERS + t.me;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(epoch) : "memory");
}

__global__ void p2p_wait_kernel(const PeerTable t, int phase, unsigned long long epoch, int *error_word)
{
    const int r = threadIdx.x;
    if (r >= t.npeers) return;
    const unsigned long long *src = t.flags[t.me] + phase * MAX_PEERS + r;
    const long long t0 = clock64();
    unsigned long long v = 0;
    if (*reinterpret_cast<volatile int *>(error_word) != 0) return;   // an earlier wait already gave up
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
        if (v >= epoch) break;
        if (clock64() - t0 > 20000000000LL) {   // ~10 s at 1.9 GHz: a peer never arrived
            atomicExch(error_word, 1 + phase);
            break;
        }
        __nanosleep(200);
    }
}

// sym_reduce_kernel with the all-to-all fused in: vbuf[v][j] goes to the GPU that integrates row j.
__global__ void __launch_bounds__(128) sym_reduce_p2p_kernel(const SymReduceArgs a, const PeerTable t, int rows_per_shard)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const int v = a.v0 + blockIdx.y;
    const int blk = j / SYM_RB;
    double xf = 0.0, yf = 0.0;
    for (int g0 = 0; g0 < blk; g0 += 2 * SYM_V) {
        const int Bs[2] = {g0 + v, g0 + 2 * SYM_V - 1 - v};
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            if (Bs[q] < blk) {
                const double2 c = a.colpart[static_cast<size_t>(sym_local_block(Bs[q], a.v0, a.nv)) * a.npad + j];
                xf += c.x;
                yf += c.y;
            }
        }
    }
    if (sym_vrank(blk) == v) {
        for (int s = (blk * SYM_RB) / a.chunk; s < a.nsplit; ++s) {
            const double2 c = a.rowpart[static_cast<size_t>(s) * a.npad + j];
            xf += c.x;
            yf += c.y;
        }
    }
    double2 *dst = t.vbuf[j / rows_per_shard];
    dst[static_cast<size_t>(v) * a.npad + j] = make_double2(xf, yf);
}

}  // namespace nbody
