// smalln.cuh -- the whole step loop of a SMALL world in one launch (EXACT mode).
//
// The reference's loop body is four calls per step (nbody-seq.c:457-464) and its acceptance runs are tiny worlds
// stepped many times (32 bodies x 100000 steps, 128 x 1000: bin/nbody-sanity-check:16, bin/tests.sh:34).  Two
// kernel launches per step cost more than the arithmetic there, so for n <= SMALL_N_MAX one CTA (n <= 32) or one
// thread-block cluster (16 CTAs where the device can co-schedule them, else 8) keeps the world in shared memory
// and runs `steps` iterations of
//   terms:  one work item per pair term a CTA needs: for an own body i and a body j of ANOTHER CTA the signed term
//           -f(j,i) (j<i) / +f(i,j) (j>i); for two own bodies a<b one evaluation f(a,b) stored as +f for a and
//           -f for b (negation is exact: XF(b) += xf, XF(c) -= xf, nbody-seq.c:103-106).  The last warp evaluates
//           the friction terms of the own bodies meanwhile (they need the velocity only), so the trig of
//           compute_velocities runs beside the pair trig, not after it, and no warp mixes the two code paths,
//   fold:   one thread per body adds its n terms strictly in j order (the reference's order for XF(i),
//           nbody-seq.c:87-107; the self slot holds +0), then finishes the body (nbody-seq.c:126-128,141-163),
//   share:  the new position goes into every CTA's copy of the world through distributed shared memory,
// with one cluster barrier per step.  Pair arithmetic, fold order and integration are the very functions the
// tiled kernels use (pair_exact, friction_exact + advance_body = integrate_math<EXACT>): the bits are the same
// whichever path runs and whatever the cluster size.
#pragma once
#include <cooperative_groups.h>

#include "step_kernels.cuh"

namespace nbody {

constexpr int SMALL_THREADS = 1024;
constexpr int SMALL_PAIR_THREADS = SMALL_THREADS - 32;   // the last warp evaluates friction terms

// shared memory: pos[2][n] + mr[n] (whole world) | vel[nloc] + frc[nloc] + fric[nloc] (own bodies) | term[n][nloc] | atan table
__host__ __device__ inline size_t small_smem_bytes(int n, int cluster)
{
    const size_t nloc = static_cast<size_t>((n + cluster - 1) / cluster);
    return 16 * (3 * static_cast<size_t>(n) + 3 * nloc + static_cast<size_t>(n) * nloc) + NBODY_DM_ATAN_TAB_LEN * 8 + 16;
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
    double2 *s_fric = s_frc + nloc;                             // [nloc]: friction term of the step
    double2 *s_term = s_fric + nloc;                            // [n][nloc]: term of body j on own body il
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
        s_term[(i0 + k) * nloc + k] = make_double2(0.0, 0.0);   // the self slot: +0 in every fold
    }
    cluster.sync();   // every CTA's shared memory exists and is initialised before anyone stores into it

    // work items of this CTA: nA cross items (j outside the own range, il fastest: a warp reads one j, broadcast,
    // and writes consecutive slots), then nB own-own unordered pairs
    const int nA = (n - cnt) * cnt;
    const int nB = cnt * (cnt - 1) / 2;
    auto decode = [&](int t, int &j, int &il) {   // item t -> (j, il) for a cross item; (b, a - i0) with a < b for an own-own item
        if (t < nA) {
            const int jo = t / cnt;
            il = t - jo * cnt;
            j = jo < i0 ? jo : jo + cnt;
        } else {
            int u = t - nA, a2 = 0;
            while (u >= cnt - 1 - a2) {
                u -= cnt - 1 - a2;
                ++a2;
            }
            il = a2;
            j = i0 + a2 + 1 + u;
        }
    };
    int j_first = 0, il_first = 0;   // the thread's first item is the same every step
    if (tid < SMALL_PAIR_THREADS && tid < nA + nB) decode(tid, j_first, il_first);

    for (int step = 0; step < a.steps; ++step) {
        const double2 *pc = s_pos + cur * n;
        if (tid < SMALL_PAIR_THREADS) {
            for (int t = tid; t < nA + nB; t += SMALL_PAIR_THREADS) {
                int j = j_first, il = il_first;
                if (t != tid) decode(t, j, il);
                const int i = i0 + il;
                const double2 pi = pc[i], qi = s_mr[i];
                const double2 pj = pc[j], qj = s_mr[j];
                const bool lower = j < i;   // pair (j, i): body i is the reference's "c"
                double fx, fy;
                pair_exact(lower ? pj.x : pi.x, lower ? pj.y : pi.y, lower ? qj.x : qi.x, lower ? qj.y : qi.y,
                           lower ? pi.x : pj.x, lower ? pi.y : pj.y, lower ? qi.x : qj.x, lower ? qi.y : qj.y, s_tab, fx, fy);
                // XF(c) -= xf for j < i, XF(b) += xf for j > i (negation is exact)
                s_term[j * nloc + il] = make_double2(lower ? -fx : fx, lower ? -fy : fy);
                if (t >= nA) s_term[i * nloc + (j - i0)] = make_double2(-fx, -fy);   // own-own (i < j): body j's term from i
            }
        } else if (tid - SMALL_PAIR_THREADS < cnt) {
            const int il = tid - SMALL_PAIR_THREADS;
            s_fric[il] = friction_exact(s_vel[il], s_tab);
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
            const double2 pn = advance_body<MODE_EXACT>(pc[i], s_mr[i], v, ax, ay, s_fric[tid], a.xdim, a.ydim);
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
