// smalln.cuh -- the whole step loop of a SMALL world in one launch (EXACT mode).
//
// The reference's loop body is four calls per step (nbody-seq.c:457-464) and its acceptance runs are tiny worlds
// stepped many times (32 bodies x 100000 steps, 128 x 1000: bin/nbody-sanity-check:16, bin/tests.sh:34).  Two
// kernel launches per step cost more than the arithmetic there, so for n <= SMALL_N_MAX one CTA (n <= 32) or one
// thread-block cluster of SMALL_CLUSTER CTAs keeps the world in shared memory and runs `steps` iterations of
//   terms:  every thread evaluates signed pair terms  -f(j,i) (j<i) / +f(i,j) (j>i)  into shared memory,
//   fold:   one thread per body adds its n terms strictly in j order (the reference's order for XF(i),
//           nbody-seq.c:87-107), then integrates the body (nbody-seq.c:114-165),
//   share:  the new position goes into every CTA's copy of the world through distributed shared memory,
// with one cluster barrier per step.  Pair arithmetic, fold order and integration are the very functions the
// tiled kernels use (pair_exact, integrate_math<EXACT>): the bits are the same whichever path runs.
#pragma once
#include <cooperative_groups.h>

#include "step_kernels.cuh"

namespace nbody {

constexpr int SMALL_THREADS = 1024;

__host__ __device__ inline int small_cluster_size(int n) { return n <= SMALL_N_ONE_CTA ? 1 : SMALL_CLUSTER; }
__host__ __device__ inline int small_rows_per_cta(int n)
{
    const int c = small_cluster_size(n);
    return (n + c - 1) / c;
}
// shared memory: pos[2][n] + mr[n] (whole world) | vel[nloc] + frc[nloc] (own bodies) | term[n][nloc] | atan table
__host__ __device__ inline size_t small_smem_bytes(int n)
{
    const size_t nloc = static_cast<size_t>(small_rows_per_cta(n));
    return 16 * (3 * static_cast<size_t>(n) + 2 * nloc + static_cast<size_t>(n) * nloc) + NBODY_DM_ATAN_TAB_LEN * 8 + 16;
}

__global__ void __launch_bounds__(SMALL_THREADS, 1) small_world_kernel(const SmallArgs a)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const int ncta = static_cast<int>(cluster.num_blocks());
    const int rank = static_cast<int>(cluster.block_rank());
    const int n = a.n;
    const int nloc = (n + ncta - 1) / ncta;          // bodies per CTA
    const int i0 = rank * nloc;                      // this CTA folds and integrates bodies [i0, i0 + cnt)
    const int cnt = max(0, min(nloc, n - i0));

    extern __shared__ __align__(16) unsigned char small_smem[];
    double2 *s_pos = reinterpret_cast<double2 *>(small_smem);   // [2][n]
    double2 *s_mr = s_pos + 2 * n;                              // [n]
    double2 *s_vel = s_mr + n;                                  // [nloc]
    double2 *s_frc = s_vel + nloc;                              // [nloc]
    double2 *s_term = s_frc + nloc;                             // [n][nloc]: term of body j on own body il
    double *s_tab = reinterpret_cast<double *>(s_term + static_cast<size_t>(n) * nloc);

    const int tid = threadIdx.x;
    for (int k = tid; k < NBODY_DM_ATAN_TAB_LEN; k += SMALL_THREADS) s_tab[k] = c_atan_tab[k];
    int cur = a.cur;
    for (int k = tid; k < n; k += SMALL_THREADS) {
        s_pos[cur * n + k] = a.pos[cur][k];
        s_mr[k] = a.mr[k];
    }
    for (int k = tid; k < cnt; k += SMALL_THREADS) {
        s_vel[k] = a.vel[i0 + k];
        s_frc[k] = a.frc[i0 + k];   // steps == 0 leaves the forces as they were
    }
    cluster.sync();   // every CTA's shared memory exists and is initialised before anyone stores into it

    const int nterms = n * cnt;
    for (int step = 0; step < a.steps; ++step) {
        const double2 *pc = s_pos + cur * n;
        // terms: il fastest, so a warp reads one j (broadcast) and writes consecutive slots
        for (int t = tid; t < nterms; t += SMALL_THREADS) {
            const int j = t / cnt;
            const int il = t - j * cnt;
            const int i = i0 + il;
            const double2 pi = pc[i], qi = s_mr[i];
            const double2 pj = pc[j], qj = s_mr[j];
            const bool lower = j < i;   // pair (j, i): body i is the reference's "c"
            double fx, fy;
            pair_exact(lower ? pj.x : pi.x, lower ? pj.y : pi.y, lower ? qj.x : qi.x, lower ? qj.y : qi.y,
                       lower ? pi.x : pj.x, lower ? pi.y : pj.y, lower ? qi.x : qj.x, lower ? qi.y : qj.y, s_tab, fx, fy);
            // XF(c) -= xf for j < i, XF(b) += xf for j > i (negation is exact); the self term is +0
            double2 v = make_double2(lower ? -fx : fx, lower ? -fy : fy);
            if (j == i) v = make_double2(0.0, 0.0);
            s_term[j * nloc + il] = v;
        }
        __syncthreads();
        if (tid < cnt) {
            double ax = 0.0, ay = 0.0;
#pragma unroll 4
            for (int j = 0; j < n; ++j) {
                const double2 v = s_term[j * nloc + tid];
                ax = __dadd_rn(ax, v.x);
                ay = __dadd_rn(ay, v.y);
            }
            const int i = i0 + tid;
            s_frc[tid] = make_double2(ax, ay);
            double2 v = s_vel[tid];
            const double2 pn = integrate_math<MODE_EXACT>(pc[i], s_mr[i], v, ax, ay, a.xdim, a.ydim, s_tab);
            s_vel[tid] = v;
            const int at = (cur ^ 1) * n + i;
            if (ncta == 1) {
                s_pos[at] = pn;
            } else {
                for (int r = 0; r < ncta; ++r) cluster.map_shared_rank(s_pos, r)[at] = pn;
            }
        }
        // the other slot is complete everywhere, and nobody reads slot `cur` any more
        if (ncta == 1) __syncthreads();
        else cluster.sync();
        cur ^= 1;   // the reference's world->old ^= 1 (nbody-seq.c:464)
    }

    for (int k = tid; k < cnt; k += SMALL_THREADS) {
        a.pos[cur][i0 + k] = s_pos[cur * n + i0 + k];
        a.vel[i0 + k] = s_vel[k];
        a.frc[i0 + k] = s_frc[k];
    }
    cluster.sync();   // no CTA exits while a peer could still address its shared memory
}

}  // namespace nbody
