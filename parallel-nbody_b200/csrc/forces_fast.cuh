// forces_fast.cuh -- FAST-mode force kernel (FMA on, algebraic direction, FP64-pipe bound).
//
// Per directed pair (i <- j), reference nbody-seq.c:89-98 rewritten without trig:
//     cos(atan2(dy,dx)) = dx/r, sin = dy/r,  force = m_i m_j G / max(dsqr, (r_i+r_j)^2)
//     => contribution = m_i G * [ m_j * rsqrt(forced^2 * dsqr) ] * (dx, dy)
// The factor m_i G is applied once per body after the loop.  17 FP64-pipe instructions
// per pair (see DESIGN.md for the count and the ncu confirmation):
//     dx dy | dsqr (2) | mindist mindsqr | max | forced^2 | q | t e p | z0 ze s | ax ay
// Everything else in the loop is kept off the FP64 pipe and under one issue slot per
// FP64 instruction: one LDS.128 pair per j serves two bodies i per thread, the
// coincidence test is an integer min on dsqr's high word.
//
// Coincident bodies (dsqr == 0, common on the integer start grid): the hot loop adds
// exactly 0 for them (q is kept positive by a 1e-290 offset that is below half an ulp of
// any real q) and records that one happened; a rare re-scan of the tile then adds the
// reference's convention: atan2(0,0) = 0 gives the lower index (+force, 0) and the
// higher (-force, 0) (nbody-seq.c:91-106).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "ptx_tma.cuh"
#include "step_args.h"

namespace nbody {

constexpr int FAST_IB = 2;                                  // bodies i per thread (default variant)
constexpr int FAST_ROWS_PER_CTA = FORCE_THREADS * FAST_IB;  // 256

__device__ __forceinline__ double rsqrt_seed(double q)
{
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(q));   // MUFU.RSQ64H, rel. error <= 2^-22.9
    return y0;
}

struct FastBody {
    double x, y, r;   // position, radius
    double ax, ay;    // accumulators (without the m_i*G factor)
    int minhi;        // min over j of the high word of dsqr: 0 <=> some dsqr == 0
};

// `self` = 1 for the pair (i, i): it adds exactly 0 like any coincident pair but must not raise the coincidence
// word (every body would otherwise take the re-scan once per step); 0 (a literal) everywhere else.
template <bool IMAX>
__device__ __forceinline__ void pair_fast2(FastBody &b, double xj, double yj, double mj, double rj, int self = 0)
{
    const double dx = xj - b.x;
    const double dy = yj - b.y;
    const double dsqr = fma(dx, dx, dy * dy);
    const double mindist = b.r + rj;
    const double mindsqr = mindist * mindist;
    double forced;
    if (IMAX) {
        // both operands are >= 0: IEEE order == integer order of the bit patterns (keeps the compare off the FP64 pipe)
        forced = __longlong_as_double(max(__double_as_longlong(dsqr), __double_as_longlong(mindsqr)));
    } else {
        forced = fmax(dsqr, mindsqr);
    }
    b.minhi = min(b.minhi, __double2hiint(dsqr) | self);
    const double f2 = forced * forced;
    const double q = fma(f2, dsqr, 1e-290);
    const double y0 = rsqrt_seed(q);
    const double t = q * y0;
    const double e = fma(-t, y0, 1.0);       // 1 - q*y0^2
    const double p = fma(e, 0.375, 0.5);     // 1/2 + 3e/8
    const double z0 = mj * y0;
    const double ze = z0 * e;
    const double s = fma(ze, p, z0);         // m_j * y0 * (1 + e/2 + 3e^2/8) = m_j / sqrt(q)
    b.ax = fma(s, dx, b.ax);
    b.ay = fma(s, dy, b.ay);
}

// Rare path: add the reference's (+-force, 0) for bodies j coincident with body i.
// `orig` (slot -> original body index, or nullptr when the arrays are in creation order): the convention is about
// ORIGINAL indices, also for bodies that became coincident after the last re-sort of a cell-ordered handle.
__device__ __noinline__ double coincident_rescan(const double2 *s_pos, const double2 *s_mr, int cnt, int jb, int i,
                                                 double xi, double yi, double ri, const int *orig = nullptr)
{
    double ax = 0.0;
    for (int jj = 0; jj < cnt; ++jj) {
        const double2 pj = s_pos[jj];
        const int j = jb + jj;
        if (pj.x == xi && pj.y == yi && j != i) {
            const double2 qj = s_mr[jj];
            const double md = ri + qj.y;
            const double f = qj.x / (md * md);   // m_j / mindsqr ; m_i*G applied by the caller
            const bool j_is_higher = orig ? orig[j] > orig[i] : j > i;
            ax += j_is_higher ? f : -f;
        }
    }
    return ax;
}

template <int IB, int UNROLL, int MINB, bool IMAX>
__global__ void __launch_bounds__(FORCE_THREADS, MINB) forces_fast_kernel(const ForceArgs a)
{
    __shared__ alignas(128) double2 s_pos[STAGES][TILE_J];
    __shared__ alignas(128) double2 s_mr[STAGES][TILE_J];
    __shared__ alignas(8) uint64_t s_full[STAGES];

    const int tid = threadIdx.x;
    const int ibase = a.lo + blockIdx.x * (FORCE_THREADS * IB) + tid;
    const int j0 = blockIdx.y * a.chunk;
    const int j1 = min(a.n, j0 + a.chunk);
    const int ntiles = (j1 - j0 + TILE_J - 1) / TILE_J;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&s_full[s], 1);
        mbar_init_fence();
    }
    __syncthreads();

    auto issue = [&](int tile) {
        const int st = tile % STAGES;
        const int jb = j0 + tile * TILE_J;
        const uint32_t bytes = static_cast<uint32_t>(min(TILE_J, j1 - jb)) * sizeof(double2);
        mbar_arrive_expect_tx(&s_full[st], 2 * bytes);
        tma_load_1d(&s_pos[st][0], a.pos + jb, bytes, &s_full[st]);
        tma_load_1d(&s_mr[st][0], a.mr + jb, bytes, &s_full[st]);
    };
    if (tid == 0) {
        for (int t = 0; t < STAGES && t < ntiles; ++t) issue(t);
    }

    FastBody b[IB];
    int idx[IB];
    double gm[IB];
#pragma unroll
    for (int k = 0; k < IB; ++k) {
        idx[k] = ibase + k * FORCE_THREADS;
        const bool live = idx[k] < a.hi;
        const double2 p = live ? a.pos[idx[k]] : make_double2(0.0, 0.0);
        const double2 q = live ? a.mr[idx[k]] : make_double2(1.0, 1.0);
        b[k].x = p.x; b[k].y = p.y; b[k].r = q.y;
        b[k].ax = 0.0; b[k].ay = 0.0;
        b[k].minhi = 0x7fffffff;
        gm[k] = q.x * NB_GRAVITY;
    }

    for (int tile = 0; tile < ntiles; ++tile) {
        const int st = tile % STAGES;
        mbar_wait(&s_full[st], (tile / STAGES) & 1);
        const int jb = j0 + tile * TILE_J;
        const int cnt = min(TILE_J, j1 - jb);
        bool self_tile = false;   // does this tile hold one of the thread's own bodies?
#pragma unroll
        for (int k = 0; k < IB; ++k) self_tile |= (idx[k] >= jb && idx[k] < jb + cnt);
        if (!self_tile) {
#pragma unroll UNROLL
            for (int jj = 0; jj < cnt; ++jj) {
                const double2 pj = s_pos[st][jj];
                const double2 qj = s_mr[st][jj];
#pragma unroll
                for (int k = 0; k < IB; ++k) pair_fast2<IMAX>(b[k], pj.x, pj.y, qj.x, qj.y);
            }
        } else {
#pragma unroll 2
            for (int jj = 0; jj < cnt; ++jj) {
                const double2 pj = s_pos[st][jj];
                const double2 qj = s_mr[st][jj];
#pragma unroll
                for (int k = 0; k < IB; ++k) pair_fast2<IMAX>(b[k], pj.x, pj.y, qj.x, qj.y, (jb + jj == idx[k]) ? 1 : 0);
            }
        }
#pragma unroll
        for (int k = 0; k < IB; ++k) {
            if (b[k].minhi == 0) {   // a body coincident with body i (not i itself) was in this tile
                b[k].ax += coincident_rescan(&s_pos[st][0], &s_mr[st][0], cnt, jb, idx[k], b[k].x, b[k].y, b[k].r);
                b[k].minhi = 0x7fffffff;
            }
        }
        __syncthreads();   // every thread is done with stage st
        if (tid == 0 && tile + STAGES < ntiles) issue(tile + STAGES);
    }

#pragma unroll
    for (int k = 0; k < IB; ++k) {
        if (idx[k] < a.hi)
            a.out[static_cast<size_t>(blockIdx.y) * a.npad + idx[k]] = make_double2(gm[k] * b[k].ax, gm[k] * b[k].ay);
    }
}

}  // namespace nbody
