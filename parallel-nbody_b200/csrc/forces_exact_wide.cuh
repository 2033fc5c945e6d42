// forces_exact_wide.cuh -- EXACT-mode force kernel with term-parallel evaluation and ordered fold.
//
// Exact mode must add body i's contributions in the reference's order (j ascending,
// nbody-seq.c:87-107), which makes the SUM a serial chain -- but the ~130-instruction pair terms
// themselves are independent.  A CTA owns 32 bodies i (lane = body).  16 compute warps evaluate
// the signed terms  -f(j,i) (j<i) / +f(i,j) (j>i)  for a tile of 64 bodies j (4 j's per warp,
// interleaved for ILP) into shared memory; one fold warp then adds the tile's 64 terms to each body's
// accumulator strictly in j order while the compute warps are already on the next tile.  Same bits
// as the one-thread-per-body kernel (and as the reference source + detmath), ~16x the parallelism:
// N = 4096 keeps 128 SMs x 17 warps busy instead of 128 warps in total.
#pragma once
#include "step_kernels.cuh"

namespace nbody {

constexpr int EXW_ROWS = 32;                       // bodies i per CTA (the wide shape); 16 when the launch is small
constexpr int EXW_CWARPS = 16;                     // compute warps
constexpr int EXW_THREADS = (EXW_CWARPS + 1) * 32; // + 1 fold warp
constexpr int EXW_TJ = 64;                         // bodies j per tile
constexpr int EXW_JPW = EXW_TJ / EXW_CWARPS;       // j's per compute warp per tile
constexpr int exw_smem_bytes(int rows)
{
    return 2 * EXW_TJ * 16 * 2 /* pos, mr */ + 2 * EXW_TJ * rows * 16 /* terms */ + NBODY_DM_ATAN_TAB_LEN * 8 + 64;
}

// ROWS = 32: lane = body, a warp instruction evaluates one j for 32 bodies.  ROWS = 16: half as many bodies per
// CTA and twice as many CTAs -- for launches that would otherwise leave SMs empty (N = 4096: 128 CTAs on 148 SMs) --
// with lanes 0..15 and 16..31 working on two different j's for the same 16 bodies.  Which j a lane evaluates does
// not matter: every term lands in its (j, body) slot and the fold warp adds them in j order either way: same bits.
template <int ROWS>
__global__ void __launch_bounds__(EXW_THREADS) forces_exact_wide_kernel(const ForceArgs a)
{
    static_assert(ROWS == 32 || ROWS == 16, "lane = body, or two j's per warp instruction");
    constexpr int JPI = 32 / ROWS;                 // j's per warp instruction
    extern __shared__ __align__(128) unsigned char exw_smem[];
    double2 *s_pos = reinterpret_cast<double2 *>(exw_smem);      // [2][EXW_TJ]
    double2 *s_mr = s_pos + 2 * EXW_TJ;                          // [2][EXW_TJ]
    double2 *s_term = s_mr + 2 * EXW_TJ;                         // [2][EXW_TJ][ROWS]
    double *s_tab = reinterpret_cast<double *>(s_term + 2 * EXW_TJ * ROWS);
    uint64_t *s_full = reinterpret_cast<uint64_t *>(s_tab + NBODY_DM_ATAN_TAB_LEN);

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int row = lane % ROWS;                  // body of this lane
    const int sub = lane / ROWS;                  // which of the JPI j's of a warp instruction
    const int i = a.lo + blockIdx.x * ROWS + row;
    const bool live = i < a.hi;
    const int ntiles = (a.n + EXW_TJ - 1) / EXW_TJ;

    for (int k = tid; k < NBODY_DM_ATAN_TAB_LEN; k += EXW_THREADS) s_tab[k] = c_atan_tab[k];
    if (tid == 0) {
        mbar_init(&s_full[0], 1);
        mbar_init(&s_full[1], 1);
        mbar_init_fence();
    }
    __syncthreads();

    auto issue = [&](int tile) {
        const int st = tile & 1;
        const int jb = tile * EXW_TJ;
        const uint32_t bytes = static_cast<uint32_t>(min(EXW_TJ, a.n - jb)) * sizeof(double2);
        mbar_arrive_expect_tx(&s_full[st], 2 * bytes);
        tma_load_1d(s_pos + st * EXW_TJ, a.pos + jb, bytes, &s_full[st]);
        tma_load_1d(s_mr + st * EXW_TJ, a.mr + jb, bytes, &s_full[st]);
    };
    if (tid == 0) {
        issue(0);
        if (ntiles > 1) issue(1);
    }

    double xi = 0.0, yi = 0.0, mi = 1.0, ri = 1.0;
    if (live) {
        const double2 p = a.pos[i];
        const double2 q = a.mr[i];
        xi = p.x; yi = p.y; mi = q.x; ri = q.y;
    }
    double ax = 0.0, ay = 0.0;   // fold warp only

    // iteration t: compute warps evaluate tile t, the fold warp folds tile t-1
    for (int t = 0; t <= ntiles; ++t) {
        if (warp < EXW_CWARPS) {
            if (t < ntiles) {
                const int st = t & 1;
                mbar_wait(&s_full[st], (t >> 1) & 1);
                const int jb = t * EXW_TJ;
                const int cnt = min(EXW_TJ, a.n - jb);
                double2 *term = s_term + st * EXW_TJ * ROWS;
#pragma unroll
                for (int u = 0; u < EXW_JPW / JPI; ++u) {
                    const int jj = warp + (u * JPI + sub) * EXW_CWARPS;
                    if (jj < cnt) {
                        const int j = jb + jj;
                        const double2 pj = s_pos[st * EXW_TJ + jj];
                        const double2 qj = s_mr[st * EXW_TJ + jj];
                        const bool lower = j < i;   // pair (j, i): body i is the reference's "c"
                        double fx, fy;
                        pair_exact(lower ? pj.x : xi, lower ? pj.y : yi, lower ? qj.x : mi, lower ? qj.y : ri,
                                   lower ? xi : pj.x, lower ? yi : pj.y, lower ? mi : qj.x, lower ? ri : qj.y, s_tab,
                                   fx, fy);
                        // XF(c) -= xf for j < i, XF(b) += xf for j > i (negation is exact); the self term is +0
                        double2 v = make_double2(lower ? -fx : fx, lower ? -fy : fy);
                        if (j == i) v = make_double2(0.0, 0.0);
                        term[jj * ROWS + row] = v;
                    }
                }
            }
        } else if (t > 0) {
            const int tp = t - 1;
            const int cnt = min(EXW_TJ, a.n - tp * EXW_TJ);
            const double2 *term = s_term + (tp & 1) * EXW_TJ * ROWS;
#pragma unroll 8
            for (int jj = 0; jj < cnt; ++jj) {
                const double2 v = term[jj * ROWS + row];   // ROWS = 16: lanes 16..31 fold the same bodies again, unused
                ax = __dadd_rn(ax, v.x);
                ay = __dadd_rn(ay, v.y);
            }
        }
        __syncthreads();
        // stage (t & 1) of pos/mr is free now (tile t computed); tile t+2 goes there
        if (tid == 0 && t + 2 < ntiles) issue(t + 2);
    }

    if (warp == EXW_CWARPS && sub == 0 && live) a.out[i] = make_double2(ax, ay);
}

}  // namespace nbody
