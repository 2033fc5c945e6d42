// p2p.cuh -- the per-step exchange fused into the kernels over NVLink peer memory (no NCCL in the step).
//
// With the symmetric kernel sharded over P GPUs (P | 8) the step needs two exchanges: every GPU's share of
// the 8 virtual-rank force vectors must reach the GPU that integrates body j, and the new positions must
// reach everybody (the reference: one MPI_Allgatherv of whole structs, nbody-par.c:913-917).  Here both
// are plain stores into peer memory issued by the producing kernels themselves:
//   sym_reduce_p2p_kernel  writes vbuf[v][j] straight into the vbuf of the GPU that owns row j,
//   integrate_kernel       writes pos_new[i] (and the tile boxes) into every GPU's buffers,
// each followed by a one-warp barrier kernel: release store of this device's barrier count into every peer's
// flag word, then acquire loads until all P flags carry that count.  The count lives in device memory
// (PeerTable::epoch) and is advanced by the kernel itself, so the launch arguments never change and the whole
// sharded step loop can sit in a CUDA graph.  Positions are double-buffered and the two barriers order the
// reuse of vbuf, so no buffer is overwritten while a peer still reads it.
// A wait gives up after P2P_WAIT_NS (30 s of wall time), raises an error word and poisons every peer's flags so
// that the other GPUs stop at once instead of waiting out their own limit.
#pragma once
#include <cuda_runtime.h>

#include "step_args.h"

namespace nbody {

constexpr unsigned long long P2P_POISON = ~0ULL;
// How long a GPU waits for a peer at a barrier before it declares the job dead.  The barriers sit in the step, so
// this also bounds the host-side skew between ranks of a multi-process job (one rank instantiating a graph, doing
// I/O or a check between two collective calls): 30 s of wall time (%globaltimer).
constexpr unsigned long long P2P_WAIT_NS = 30000000000ULL;

__global__ void __launch_bounds__(32) p2p_barrier_kernel(const PeerTable t, int phase, int *error_word)
{
    const int r = threadIdx.x;
    unsigned long long e = 0;
    if (r == 0) {
        e = t.epoch[phase] + 1;   // this device's own count: same sequence of barriers on every device
        t.epoch[phase] = e;
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if (r >= t.npeers) return;
    if (*reinterpret_cast<volatile int *>(error_word) != 0) return;   // an earlier wait already gave up
    // signal: everything this device stored before this kernel (stream order) is visible to a peer that sees e
    __threadfence_system();
    unsigned long long *dst = t.flags[r] + phase * MAX_PEERS + t.me;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(e) : "memory");
    // wait for peer r's signal
    const unsigned long long *src = t.flags[t.me] + phase * MAX_PEERS + r;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned long long v = 0;
    for (unsigned spins = 0;; ++spins) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
        if (v >= e) break;
        if ((spins & 1023u) == 1023u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (now - t0 > P2P_WAIT_NS) {   // wall time, whatever the SM clock: a peer never arrived
                v = P2P_POISON;
                break;
            }
        }
        __nanosleep(200);
    }
    if (v == P2P_POISON) {
        atomicExch(error_word, 1 + phase);
        for (int q = 0; q < t.npeers; ++q)       // release everybody who waits for this device, in every phase
            for (int ph = 0; ph < P2P_PHASES; ++ph) {
                unsigned long long *d = t.flags[q] + ph * MAX_PEERS + t.me;
                asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(d), "l"(P2P_POISON) : "memory");
            }
    }
}

// sym_reduce_kernel with the all-to-all fused in: vbuf[v][j] goes to the GPU that integrates row j.
__global__ void __launch_bounds__(128) sym_reduce_p2p_kernel(const SymReduceArgs a, const PeerTable t, int rows_per_shard)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const int v = a.v0 + blockIdx.y;
    const int blk = j / SYM_RB;
    const double2 r = sym_take_reactions(a, j, blockIdx.y);
    double xf = r.x, yf = r.y;
    if (sym_vrank(blk) == v) {
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const double2 c = a.rowsum[static_cast<size_t>(p) * a.npad + j];
            xf += c.x;
            yf += c.y;
        }
    }
    double2 *dst = t.vbuf[j / rows_per_shard];
    dst[static_cast<size_t>(v) * a.npad + j] = make_double2(xf, yf);
}

}  // namespace nbody
