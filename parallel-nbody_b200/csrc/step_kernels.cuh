// step_kernels.cuh -- the two kernels of the N-body step, as templates on the mode.
//
//   forces_kernel<MODE>     clear_forces + compute_forces (reference nbody-seq.c:67-109)
//   integrate_kernel<MODE>  compute_velocities + compute_positions (nbody-seq.c:114-165), fused
//
// Instantiated for MODE = EXACT in kernels_exact.cu (built with -fmad=false) and for
// MODE = FAST in kernels_fast.cu (FMA contraction on).  Device code is whole-program
// per translation unit, so the two instantiations never share floating-point code.
//
// Data layout in HBM (structure of arrays, 16 B per body per array):
//   pos[2][npad] double2 (x, y)    ping-pong: the reference's x[2]/y[2] slots (nbody-seq.c:31-32)
//   mr[npad]     double2 (mass, radius)  constant after init (nbody-seq.c:37-38)
//   vel[npad]    double2 (xv, yv)
//   frc[npad]    double2 (xf, yf)  last step's pair sums -- what print() shows
//   part[S][npad] double2          fast mode only: per-j-chunk partial sums
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>

#include "../../include/nbody_detmath.h"
#include "ptx_tma.cuh"
#include "step_args.h"

namespace nbody {

// reference nbody-seq.c:22-25 -- the same literal expressions, folded by the compiler
#define NB_GRAVITY 1.1
#define NB_FRICTION 0.01
#define NB_DELTA_T (0.025 / 5000)

// atan2 combination table (include/nbody_detmath_tables.h); copied to shared memory by EXACT kernels
static __constant__ double c_atan_tab[NBODY_DM_ATAN_TAB_LEN] = NBODY_DM_ATAN_TAB_INIT;

// ---------------------------------------------------------------------------------
// Per-pair arithmetic
// ---------------------------------------------------------------------------------

// EXACT: one oriented evaluation, b < c, exactly nbody-seq.c:89-98 with detmath trig.
// Every operation is an explicit round-to-nearest intrinsic: nothing can be contracted.
__device__ __forceinline__ void pair_exact(double xb, double yb, double mb, double rb, double xc, double yc,
                                           double mc, double rc, const double *tab, double &fx, double &fy)
{
    double dx = __dsub_rn(xc, xb);
    double dy = __dsub_rn(yc, yb);
    double angle = nbody_dm_atan2(dy, dx, tab);
    double dsqr = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    double mindist = __dadd_rn(rb, rc);
    double mindsqr = __dmul_rn(mindist, mindist);
    double forced = (dsqr < mindsqr) ? mindsqr : dsqr;
    double force = __ddiv_rn(__dmul_rn(__dmul_rn(mb, mc), NB_GRAVITY), forced);
    double s, c;
    nbody_dm_sincos(angle, &s, &c);
    fx = __dmul_rn(force, c);
    fy = __dmul_rn(force, s);
}

// FAST: 1/sqrt(q) to ~1e-15 relative from the hardware seed (MUFU.RSQ64H) and one
// third-order correction; q > 0 finite.
__device__ __forceinline__ double rsqrt_fast(double q)
{
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(q));
    double t = q * y0;
    double e = fma(-t, y0, 1.0);          // 1 - q*y0^2
    double p = fma(e, 0.375, 0.5);        // 1/2 + 3e/8
    double ye = y0 * e;
    return fma(ye, p, y0);                // y0*(1 + e/2 + 3e^2/8)
}

// FAST: force of body j on body i, accumulated into (ax, ay).
//   direction  cos(atan2(dy,dx)) = dx/r, sin = dy/r   (the reference's trig, algebraically)
//   magnitude  m_i*m_j*G / max(dsqr, (r_i+r_j)^2)
//   => s = m_i*m_j*G / (forced * sqrt(dsqr)) = m_i*m_j*G * rsqrt(forced^2 * dsqr)
// Coincident bodies (dsqr == 0): the reference's atan2(0,0) = 0 gives the lower index
// (+force, 0) and the higher (-force, 0) (nbody-seq.c:91-106); j == i contributes 0.
__device__ __forceinline__ void pair_fast(double xi, double yi, double gmi, double ri, int i, double xj, double yj,
                                          double mj, double rj, int j, double &ax, double &ay)
{
    double dx = xj - xi;
    double dy = yj - yi;
    double dsqr = fma(dx, dx, dy * dy);
    double mindist = ri + rj;
    double mindsqr = mindist * mindist;
    double forced = fmax(dsqr, mindsqr);
    if (dsqr == 0.0) {
        dx = (j > i) ? 1.0 : ((j < i) ? -1.0 : 0.0);
        dsqr = 1.0;
    }
    double q = (forced * forced) * dsqr;
    double s = (gmi * mj) * rsqrt_fast(q);
    ax = fma(s, dx, ax);
    ay = fma(s, dy, ay);
}

// ---------------------------------------------------------------------------------
// forces_kernel
//   grid  = (ceil((hi-lo)/FORCE_THREADS), S)   S = 1 in EXACT mode
//   block = FORCE_THREADS; thread t owns body i = lo + blockIdx.x*FORCE_THREADS + t
// Body j's (x,y) and (mass,radius) are staged through shared memory in tiles of
// TILE_J by two 1-D TMA bulk copies per tile, double-buffered behind mbarriers; every
// lane of a warp reads the same j (LDS.128 broadcast).
// ---------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(FORCE_THREADS) forces_kernel(const ForceArgs a)
{
    __shared__ alignas(128) double2 s_pos[STAGES][TILE_J];
    __shared__ alignas(128) double2 s_mr[STAGES][TILE_J];
    __shared__ alignas(8) uint64_t s_full[STAGES];
    __shared__ double s_tab[(MODE == MODE_EXACT) ? NBODY_DM_ATAN_TAB_LEN : 1];

    const int tid = threadIdx.x;
    const int i = a.lo + blockIdx.x * FORCE_THREADS + tid;
    const bool live = i < a.hi;

    int j0 = 0, j1 = a.n;
    if (MODE == MODE_FAST) {
        j0 = blockIdx.y * a.chunk;
        j1 = min(a.n, j0 + a.chunk);
    }
    const int ntiles = (j1 - j0 + TILE_J - 1) / TILE_J;

    if (MODE == MODE_EXACT) {
        for (int k = tid; k < NBODY_DM_ATAN_TAB_LEN; k += FORCE_THREADS) s_tab[k] = c_atan_tab[k];
    }
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

    double xi = 0.0, yi = 0.0, mi = 1.0, ri = 1.0;
    if (live) {
        const double2 p = a.pos[i];
        const double2 q = a.mr[i];
        xi = p.x; yi = p.y; mi = q.x; ri = q.y;
    }
    const double gmi = mi * NB_GRAVITY;   // FAST only
    double ax = 0.0, ay = 0.0;

    for (int tile = 0; tile < ntiles; ++tile) {
        const int st = tile % STAGES;
        mbar_wait(&s_full[st], (tile / STAGES) & 1);
        const int jb = j0 + tile * TILE_J;
        const int cnt = min(TILE_J, j1 - jb);
        if (MODE == MODE_EXACT) {
#pragma unroll 2
            for (int jj = 0; jj < cnt; ++jj) {
                const int j = jb + jj;
                const double2 pj = s_pos[st][jj];
                const double2 qj = s_mr[st][jj];
                const bool lower = j < i;   // pair (j, i): body i is the reference's "c"
                double fx, fy;
                pair_exact(lower ? pj.x : xi, lower ? pj.y : yi, lower ? qj.x : mi, lower ? qj.y : ri,
                           lower ? xi : pj.x, lower ? yi : pj.y, lower ? mi : qj.x, lower ? ri : qj.y, s_tab, fx,
                           fy);
                // XF(c) -= xf for j < i, XF(b) += xf for j > i; negation is exact
                const double nx = __dadd_rn(ax, lower ? -fx : fx);
                const double ny = __dadd_rn(ay, lower ? -fy : fy);
                if (j != i) { ax = nx; ay = ny; }
            }
        } else {
#pragma unroll 4
            for (int jj = 0; jj < cnt; ++jj) {
                const double2 pj = s_pos[st][jj];
                const double2 qj = s_mr[st][jj];
                pair_fast(xi, yi, gmi, ri, i, pj.x, pj.y, qj.x, qj.y, jb + jj, ax, ay);
            }
        }
        __syncthreads();   // every thread is done with stage st
        if (tid == 0 && tile + STAGES < ntiles) issue(tile + STAGES);
    }

    if (live) {
        if (MODE == MODE_FAST) a.out[static_cast<size_t>(blockIdx.y) * a.npad + i] = make_double2(ax, ay);
        else a.out[i] = make_double2(ax, ay);
    }
}

// ---------------------------------------------------------------------------------
// integrate_kernel: one thread per owned body.
// ---------------------------------------------------------------------------------
// The friction term of compute_velocities (nbody-seq.c:120-125): (force*cos(angle), force*sin(angle)) with
// force = sqrt(xv^2+yv^2)*FRICTION, angle = atan2(yv, xv).  It depends on the velocity alone, not on the
// pair forces, so the one-launch step loop (smalln.cuh) evaluates it beside the pair terms.
__device__ __forceinline__ double2 friction_exact(double2 v, const double *tab)
{
    double force = __dmul_rn(__dsqrt_rn(__dadd_rn(__dmul_rn(v.x, v.x), __dmul_rn(v.y, v.y))), NB_FRICTION);
    double angle = nbody_dm_atan2(v.y, v.x, tab);
    double s, c;
    nbody_dm_sincos(angle, &s, &c);
    return make_double2(__dmul_rn(force, c), __dmul_rn(force, s));
}

// The rest of compute_velocities and compute_positions (nbody-seq.c:126-128,141-163): fric = the friction term
// (EXACT) or unused (FAST: FRICTION*v, formed here); v is updated in place; returns the new position.
template <int MODE>
__device__ __forceinline__ double2 advance_body(double2 p, double2 q, double2 &v, double xf, double yf, double2 fric,
                                                int xdim, int ydim)
{
    double xn, yn;
    if (MODE == MODE_EXACT) {
        double fx = __dsub_rn(xf, fric.x);
        double fy = __dsub_rn(yf, fric.y);
        v.x = __dadd_rn(v.x, __dmul_rn(__ddiv_rn(fx, q.x), NB_DELTA_T));
        v.y = __dadd_rn(v.y, __dmul_rn(__ddiv_rn(fy, q.x), NB_DELTA_T));
        // compute_positions uses the updated velocity
        xn = __dadd_rn(p.x, __dmul_rn(v.x, NB_DELTA_T));
        yn = __dadd_rn(p.y, __dmul_rn(v.y, NB_DELTA_T));
    } else {
        // friction acts along -v: sqrt(|v|^2)*FRICTION*(cos,sin)(atan2(yv,xv)) = FRICTION*(xv,yv)
        double fx = xf - NB_FRICTION * v.x;
        double fy = yf - NB_FRICTION * v.y;
        double im = NB_DELTA_T / q.x;
        v.x = fma(fx, im, v.x);
        v.y = fma(fy, im, v.y);
        xn = fma(v.x, NB_DELTA_T, p.x);
        yn = fma(v.y, NB_DELTA_T, p.y);
    }
    if (xn < 0) {
        xn = 0;
        v.x = -v.x;
    } else if (xn >= xdim) {
        xn = xdim - 1;
        v.x = -v.x;
    }
    if (yn < 0) {
        yn = 0;
        v.y = -v.y;
    } else if (yn >= ydim) {
        yn = ydim - 1;
        v.y = -v.y;
    }
    return make_double2(xn, yn);
}

// compute_velocities + compute_positions for one body (nbody-seq.c:114-165): p = position, q = (mass, radius),
// v = velocity (updated in place), (xf, yf) = the body's pair-force sum; returns the new position.
template <int MODE>
__device__ __forceinline__ double2 integrate_math(double2 p, double2 q, double2 &v, double xf, double yf, int xdim,
                                                  int ydim, const double *tab)
{
    double2 fric = make_double2(0.0, 0.0);
    if (MODE == MODE_EXACT) fric = friction_exact(v, tab);
    return advance_body<MODE>(p, q, v, xf, yf, fric, xdim, ydim);
}

// FAST: total force on body i from the partial sums, in a fixed order.  Directed kernel: the j-chunk
// partials, chunk ascending.  Symmetric kernel: the per-virtual-rank vectors of sym_reduce_kernel.
__device__ __forceinline__ void sum_partials(const IntegrateArgs &a, int i, double &xf, double &yf)
{
    xf = 0.0;
    yf = 0.0;
    if (a.vbuf) {   // symmetric kernel: SYM_V per-virtual-rank sums, v ascending
#pragma unroll
        for (int v = 0; v < SYM_V; ++v) {
            const double2 t = a.vbuf[static_cast<size_t>(v) * a.npad + i];
            xf += t.x;
            yf += t.y;
        }
        return;
    }
    const int s0 = 0;
    for (int s = s0; s < a.nsplit; ++s) {
        const double2 t = a.part[static_cast<size_t>(s) * a.npad + i];
        xf += t.x;
        yf += t.y;
    }
}

// Block size: 128 threads, or TILE_J when the tile boxes of the new positions are fused in (a.boxes != nullptr,
// a.lo % TILE_J == 0: CTA = one tile of the cell order; the force kernel's far-field test reads the boxes next
// step, so the separate tile_boxes_kernel launch is only needed after a re-sort).
template <int MODE>
__global__ void __launch_bounds__(TILE_J) integrate_kernel(const IntegrateArgs a)
{
    __shared__ double s_tab[(MODE == MODE_EXACT) ? NBODY_DM_ATAN_TAB_LEN : 1];
    __shared__ double s_red[(MODE == MODE_FAST) ? 5 : 1][TILE_J / 32];
    // a force-kernel hand-over or a peer barrier gave up: do not integrate garbage (uniform over the CTA)
    if (a.error_word && *reinterpret_cast<volatile const int *>(a.error_word) != 0) return;
    if (MODE == MODE_EXACT) {
        for (int k = threadIdx.x; k < NBODY_DM_ATAN_TAB_LEN; k += blockDim.x) s_tab[k] = c_atan_tab[k];
        __syncthreads();
    }
    const int i = a.lo + blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < a.hi;
    double2 pn = make_double2(0.0, 0.0);
    double radius = 0.0;
    if (live) {
        double xf, yf;
        if (MODE == MODE_FAST) {
            // fixed-order sum of the j-chunk partials: deterministic, independent of the GPU count
            sum_partials(a, i, xf, yf);
            a.frc[i] = make_double2(xf, yf);
        } else {
            const double2 t = a.frc[i];
            xf = t.x;
            yf = t.y;
        }
        const double2 q = a.mr[i];
        double2 v = a.vel[i];
        pn = integrate_math<MODE>(a.pos_old[i], q, v, xf, yf, a.xdim, a.ydim, s_tab);
        radius = q.y;
        a.vel[i] = v;
        if (a.npeers > 0) {
            for (int r = 0; r < a.npeers; ++r) a.peer_pos_new[r][i] = pn;   // every GPU's copy, over NVLink (incl. our own)
        } else {
            a.pos_new[i] = pn;
        }
    }
    if (MODE == MODE_FAST && a.boxes) {   // block-uniform
        // boxes[tile] = {xmin, ymin, xmax, ymax, rmax, 0, 0, 0} of the tile's NEW positions (as tile_boxes_kernel)
        double v[5] = {DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, 0.0};
        if (live) {
            v[0] = pn.x; v[1] = pn.y; v[2] = pn.x; v[3] = pn.y;
            v[4] = radius;
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            v[0] = fmin(v[0], __shfl_xor_sync(0xffffffffu, v[0], off));
            v[1] = fmin(v[1], __shfl_xor_sync(0xffffffffu, v[1], off));
            v[2] = fmax(v[2], __shfl_xor_sync(0xffffffffu, v[2], off));
            v[3] = fmax(v[3], __shfl_xor_sync(0xffffffffu, v[3], off));
            v[4] = fmax(v[4], __shfl_xor_sync(0xffffffffu, v[4], off));
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) {
#pragma unroll
            for (int q = 0; q < 5; ++q) s_red[q][warp] = v[q];
        }
        __syncthreads();
        if (threadIdx.x < SYM_BOX_DOUBLES) {
            const int q = threadIdx.x;
            double r = 0.0;
            if (q < 5) {
                r = s_red[q][0];
                for (int w = 1; w < TILE_J / 32; ++w) r = (q < 2) ? fmin(r, s_red[q][w]) : fmax(r, s_red[q][w]);
            }
            const size_t at = static_cast<size_t>((a.lo + blockIdx.x * TILE_J) / TILE_J) * SYM_BOX_DOUBLES + q;
            if (a.npeers > 0) {
                for (int p = 0; p < a.npeers; ++p) a.peer_boxes[p][at] = r;
            } else {
                a.boxes[at] = r;
            }
        }
    }
}

// FAST: frc[i] = sum over j-chunks of part[s][i], s ascending (the same order integrate_kernel uses);
// used by the forces-only entry point.
template <int MODE>
__global__ void __launch_bounds__(128) reduce_partials_kernel(const IntegrateArgs a)
{
    const int i = a.lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.hi) return;
    double xf, yf;
    sum_partials(a, i, xf, yf);
    a.frc[i] = make_double2(xf, yf);
}

}  // namespace nbody
