// forces_sym.cuh -- FAST-mode force kernel that evaluates every unordered pair ONCE.
//
// The reference's compute_forces walks the upper triangle and gives +f to body b and -f
// to body c (nbody-seq.c:87-107).  This kernel does the same on the GPU: bodies are cut
// into row blocks of SYM_RB; CTA (block B, j-chunk s) keeps SYM_IB bodies i of block B
// per thread in registers and walks the j tiles of chunk s
//   * below the block: skipped (those pairs belong to the block that owns j),
//   * inside the block (diagonal tiles): directed evaluation, both directions, no reaction,
//   * above the block: one evaluation per pair; +f accumulates in registers for body i
//     (row sum); the reaction on body j travels through the warp in a rotating accumulator,
//     the warps' sums are added in warp order through shared memory and kept there for a
//     segment of SYM_SEG bodies j.  The segment is then ADDED to vacc[v][j] (v = the block's
//     virtual rank) -- not with atomics: the CTAs of v's row blocks take turns in ascending
//     block order (turn[v][segment] counts them, acquire/release at GPU scope, the way a
//     serial split-K GEMM hands its tile from CTA to CTA), so the order of the additions is
//     fixed, the result is deterministic, and the reaction sums take nv * N * 16 B instead
//     of one N-vector per row block.  CTAs are dispatched row block by row block, so the
//     CTA a turn waits for started a row of CTAs earlier and has normally long finished.
// Far field: with the bodies in cell order (kernels_sort.cu) most (row block, j tile) pairs are further apart
// than any r_i + r_j of their bodies.  Every tile carries a bounding box (x/y range, largest radius; one more
// TMA copy per tile); if the gap between the block's box and the tile's box is at least rmax_B + rmax_T the
// clamp max(dsqr, mindsqr) is the identity for every pair of the tile and no two bodies can coincide, so the
// tile runs the far-field form of the pair: 15 FP64-pipe instructions and no integer work.  The test uses the
// same rounded operations as the pair itself (monotone), so it is exact and both forms give the same bits.
// Row sums: the CTAs of one row block leave their sums in a ring of slices and count themselves in; the block's last
// CTA in dispatch order waits for the count and adds the slices in chunk order (rowsum[phase][i]; a small ring that
// stays in the L2 instead of one N-vector per j-chunk in HBM).
// sym_reduce_kernel then adds each body's row sums to its block's virtual rank: SYM_V per-virtual-rank vectors
// (step_args.h), and integrate_kernel adds those in a fixed order: the same bits on 1, 2, 4 or 8 GPUs.
// Per unordered pair: 17 FP64-pipe instructions (15 in far-field tiles) + 1 MUFU + 1.5 .. 6 ALU/LSU/SHFL, against
// 2 x (16 + ...) for the directed kernel.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "forces_fast.cuh"

#ifndef NBODY_SYM_RSQRT_ORDER
#define NBODY_SYM_RSQRT_ORDER 2   // order of the rsqrt correction in this kernel's pairs (2 or 3)
#endif
// Second order: y1 = y0*(1 + e/2), e = 1 - q*y0^2.  The dropped term 3e^2/8 is one-sided (every pair term comes out too
// small); its mean over the seed's error distribution is a property of the seed unit (MUFU.RSQ64H with the low word
// taken from a dead register, rsqrt_seed_donor), measured on both BASELINE workloads as 1.94e-13 (N = 65536:
// 1.92e-13, N = 262144: 1.97e-13; profiles/r2_error_table.log).  The kernel multiplies its sums by G anyway: folding
// (1 + that mean) into the constant makes the pair terms unbiased at no cost, and the error of a body's force
// drops from the bias (2e-13) to the zero-mean scatter.
// The pair computes TWICE that value, y0*(3 - q*y0^2) = 2*y0*(1 + e/2), in three instructions (t = q*y0,
// w = fma(-t, y0, 3), y0*w) instead of four; the sums are multiplied by G/2 at the end (exact halving), and the
// re-scans of coincident bodies add twice their term.
#if NBODY_SYM_RSQRT_ORDER >= 3
#define NBODY_SYM_G NB_GRAVITY
#else
#define NBODY_SYM_G (NB_GRAVITY * (1.0 + 1.94e-13))
#endif
#define NBODY_SYM_GH (0.5 * NBODY_SYM_G)

namespace nbody {

struct SymBody {
    double x, y, r, m;   // position, radius, mass
    double ax, ay;       // row accumulators (without the m_i*G factor)
    int minhi;
};

// MUFU.RSQ64H reads and writes high words only; PTX's rsqrt.approx.f64 also zeroes the low word, which costs a
// move per pair.  The seed is only good to 2^-20 (tools/rsqrt_seed_probe.cu), so any low word will do: take it from a value that dies
// here (the register pair is reused in place, no move).
__device__ __forceinline__ double rsqrt_seed_donor(double q, double donor)
{
    int hi;
    asm("{\n.reg .f64 t;\n.reg .b32 lo;\nrsqrt.approx.ftz.f64 t, %1;\nmov.b64 {lo, %0}, t;\n}" : "=r"(hi) : "d"(q));
    return __hiloint2double(hi, __double2loint(donor));
}

// One unordered pair (i < j): row side into b.ax/ay, reaction side into (rx, ry).
// FLAVOR bit 0: seed low word from a dead value; bit 1: one coincidence word per thread (min3); bit 2: far field
// (no clamp, no coincidence test: the caller guarantees dsqr >= mindsqr > 0).
template <int FLAVOR>
__device__ __forceinline__ void pair_sym(SymBody &b, double xj, double yj, double mj, double rj, double &rx,
                                         double &ry, int &minhi)
{
    const double dx = xj - b.x;
    const double dy = yj - b.y;
    const double dsqr = fma(dx, dx, dy * dy);
    double q, f2;
    if (FLAVOR & 4) {
        f2 = dsqr * dsqr;
        q = f2 * dsqr;
    } else {
        const double mindist = b.r + rj;
        const double mindsqr = mindist * mindist;
        const double forced = __longlong_as_double(max(__double_as_longlong(dsqr), __double_as_longlong(mindsqr)));
        if (FLAVOR & 2) minhi = __double2hiint(dsqr);   // combined by the caller
        else b.minhi = min(b.minhi, __double2hiint(dsqr));
        f2 = forced * forced;
        q = fma(f2, dsqr, 1e-290);
    }
    const double y0 = (FLAVOR & 1) ? rsqrt_seed_donor(q, f2) : rsqrt_seed(q);
    const double t = q * y0;
#if NBODY_SYM_RSQRT_ORDER >= 3
    const double e = fma(-t, y0, 1.0);       // 1 - q*y0^2; |e| < 4e-6 (seed: 2^-19, measured by tools/rsqrt_seed_probe.cu, + the donated low word: 2^-20)
    const double ye = y0 * e;
    const double p = fma(e, 0.75, 1.0);
    const double y1 = fma(ye, p, y0 + y0);   // 2*y0*(1 + e/2 + 3e^2/8) = 2 / (forced * sqrt(dsqr)) to ~1e-16
#else
    const double w = fma(-t, y0, 3.0);       // 2 + e, e = 1 - q*y0^2, |e| < 4e-6
    const double y1 = y0 * w;                // 2*y0*(1 + e/2); the mean of the dropped term is folded into NBODY_SYM_G
#endif
    const double zi = mj * y1;
    b.ax = fma(zi, dx, b.ax);
    b.ay = fma(zi, dy, b.ay);
    const double zj = b.m * y1;
    rx = fma(zj, dx, rx);
    ry = fma(zj, dy, ry);
}

// Directed evaluation (diagonal tiles), same arithmetic as forces_fast.cuh.  `self` = 1 for the pair (i, i): its
// dsqr = 0 adds exactly 0 like any coincident pair, but it must not raise the coincidence word (every body would
// otherwise take the re-scan once per step).
__device__ __forceinline__ void pair_dir(SymBody &b, double xj, double yj, double mj, double rj, int self)
{
    const double dx = xj - b.x;
    const double dy = yj - b.y;
    const double dsqr = fma(dx, dx, dy * dy);
    const double mindist = b.r + rj;
    const double mindsqr = mindist * mindist;
    const double forced = __longlong_as_double(max(__double_as_longlong(dsqr), __double_as_longlong(mindsqr)));
    b.minhi = min(b.minhi, __double2hiint(dsqr) | self);
    const double f2 = forced * forced;
    const double q = fma(f2, dsqr, 1e-290);
    const double y0 = rsqrt_seed_donor(q, f2);   // the same seed as pair_sym: the same error distribution
    const double t = q * y0;
    const double z0 = mj * y0;
#if NBODY_SYM_RSQRT_ORDER >= 3
    const double e = fma(-t, y0, 1.0);
    const double ze = z0 * e;
    const double p = fma(e, 0.75, 1.0);
    const double s = fma(ze, p, z0 + z0);
#else
    const double w = fma(-t, y0, 3.0);       // as in pair_sym: twice the term, G/2 at the end
    const double s = z0 * w;
#endif
    b.ax = fma(s, dx, b.ax);
    b.ay = fma(s, dy, b.ay);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

template <int IB, bool GUARD, int UNROLL, int FLAVOR>
__device__ __forceinline__ void sym_group(SymBody (&b)[IB], const double2 *tp, const double2 *tm, int g, int cnt,
                                          int lane, double2 *col, int &minhi)
{
    double cx = 0.0, cy = 0.0;
    const int src = (lane + 1) & 31;
#pragma unroll UNROLL
    for (int r = 0; r < 32; ++r) {
        const int jj = g + ((lane + r) & 31);
        double2 pj, qj;
        int hi[IB];
        if (GUARD && jj >= cnt) {
            pj = make_double2(-1.0e6, -1.0e6);   // beyond the last body: massless, far away, never stored
            qj = make_double2(0.0, 1.0);
        } else {
            pj = tp[jj];
            qj = tm[jj];
        }
#pragma unroll
        for (int k = 0; k < IB; ++k) {
            if ((FLAVOR & 6) == 2) {
                pair_sym<FLAVOR>(b[k], pj.x, pj.y, qj.x, qj.y, cx, cy, hi[k]);
            } else {
                pair_sym<FLAVOR>(b[k], pj.x, pj.y, qj.x, qj.y, cx, cy, minhi);
            }
        }
        if ((FLAVOR & 6) == 2) {
#pragma unroll
            for (int k = 0; k + 1 < IB; k += 2) minhi = __vimin3_s32(minhi, hi[k], hi[k + 1]);
            if (IB & 1) minhi = min(minhi, hi[IB - 1]);
        }
        cx = __shfl_sync(0xffffffffu, cx, src);
        cy = __shfl_sync(0xffffffffu, cy, src);
    }
    col[g + lane] = make_double2(cx, cy);   // entries >= cnt are never read
}

constexpr int sym_smem_bytes(int threads) { return STAGES * TILE_J * 16 * 2 + (threads / 32) * TILE_J * 16 + SYM_SEG * 16 + (STAGES + 1) * SYM_BOX_DOUBLES * 8 + 64; }
static_assert(SYM_RB % TILE_J == 0, "row blocks must be tile aligned");
static_assert(SYM_SEG % TILE_J == 0 && SYM_SEG % SYM_RB == 0, "segments are whole tiles and whole row blocks");

// A CTA's turn in its (virtual rank, segment) chain: wait until the `seq` blocks before it have added their sums.
// ASSUMPTION (not a CUDA guarantee): CTAs start in ascending blockIdx order, so the CTA waited for -- always one
// with a lower index (tests/test_sym_items.py checks the table) -- is running or done.  Every GPU so far dispatches
// that way; should one not, the wait gives up after SYM_WAIT_NS of wall time (%globaltimer: a debugger, a
// time-sliced GPU or a profiler stretch cycles, not this limit) or as soon as another CTA has given up, and raises
// the error word.  From then on the step is lost: later force CTAs, the fold and integrate_kernel return at once
// (so the state stays that of the last good step), and the host reports the failure at its next synchronisation.
constexpr unsigned long long SYM_WAIT_NS = 4000000000ULL;
__device__ __forceinline__ unsigned long long wall_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void turn_wait(const int *turn, int seq, int *error_word)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(turn) : "memory");
    if (v == seq) return;
    const unsigned long long t0 = wall_ns();
    for (unsigned spins = 0;; ++spins) {
        __nanosleep(spins < 64 ? 32 : 256);
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(turn) : "memory");
        if (v == seq) return;
        if ((spins & 255u) == 255u) {
            if (*reinterpret_cast<volatile int *>(error_word) != 0) return;   // somebody else gave up already
            if (wall_ns() - t0 > SYM_WAIT_NS) break;
        }
    }
    atomicExch(error_word, SYM_ERR_TURN);
}
__device__ __forceinline__ void turn_pass(int *turn, int next)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(turn), "r"(next) : "memory");
}
// count this CTA into its group (release: its slice is visible first): a reduction, nothing comes back
__device__ __forceinline__ void group_arrive(int *cnt)
{
    asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(cnt) : "memory");
}
__device__ __forceinline__ int peek_acquire(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Rare path, row side of a tile above the block (all slots j > slot i): +- m_j / mindsqr for coincident j.  The
// reference gives the body with the lower ORIGINAL index (+force, 0) (atan2(0,0) = 0, nbody-seq.c:91-106): in cell
// order (orig != nullptr) the slot order says nothing about that for bodies that met after the last re-sort.
__device__ __noinline__ double sym_row_rescan(const double2 *s_pos, const double2 *s_mr, int cnt, double xi, double yi,
                                              double ri, const int *orig, int jb, int i)
{
    double ax = 0.0;
    const int oi = orig ? orig[i] : 0;
    for (int jj = 0; jj < cnt; ++jj) {
        const double2 pj = s_pos[jj];
        if (pj.x == xi && pj.y == yi) {
            const double2 qj = s_mr[jj];
            const double md = ri + qj.y;
            const double f = qj.x / (md * md);
            ax += (!orig || orig[jb + jj] > oi) ? f : -f;
        }
    }
    return ax;
}

// Rare path, column side: sum over the block's bodies i (slots below j) coincident with body j of +- m_i / mindsqr
// (+: i has the lower original index, so j takes the reaction -force; the caller multiplies by -G m_j / 2).
__device__ __noinline__ double sym_col_rescan(const double2 *pos, const double2 *mr, int i0, int i1, double xj,
                                              double yj, double rj, const int *orig, int j)
{
    double c = 0.0;
    const int oj = orig ? orig[j] : 0;
    for (int i = i0; i < i1; ++i) {
        const double2 pi = pos[i];
        if (pi.x == xj && pi.y == yj) {
            const double2 qi = mr[i];
            const double md = qi.y + rj;
            const double f = qi.x / (md * md);
            c += (!orig || orig[i] < oj) ? f : -f;
        }
    }
    return c;
}

template <int SYM_THREADS, int SYM_IB, int MINB = 1, int UNROLL = 4, int FLAVOR = 3, int FAR_FLAVOR = 5>
__global__ void __launch_bounds__(SYM_THREADS, MINB) forces_sym_kernel(const SymArgs a)
{
    static_assert(SYM_RB == SYM_THREADS * SYM_IB, "row block = threads x bodies per thread");
    constexpr int SYM_WARPS = SYM_THREADS / 32;
    extern __shared__ __align__(128) unsigned char sym_smem[];
    double2 *s_pos = reinterpret_cast<double2 *>(sym_smem);                 // [STAGES][TILE_J]
    double2 *s_mr = s_pos + STAGES * TILE_J;                                // [STAGES][TILE_J]
    double2 *s_col = s_mr + STAGES * TILE_J;                                // [SYM_WARPS][TILE_J]
    double2 *s_seg = s_col + SYM_WARPS * TILE_J;                            // [SYM_SEG] reaction sums of the current segment
    double *s_box = reinterpret_cast<double *>(s_seg + SYM_SEG);              // [STAGES][SYM_BOX_DOUBLES] tile boxes
    double *s_bbox = s_box + STAGES * SYM_BOX_DOUBLES;                        // [SYM_BOX_DOUBLES] this row block's box
    uint64_t *s_full = reinterpret_cast<uint64_t *>(s_bbox + SYM_BOX_DOUBLES);
    int *s_flag = reinterpret_cast<int *>(s_full + STAGES);                 // [2] coincidence flags, [1] ring slot free

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    // grid = the live (row block, chunk) items in dispatch order (SymArgs::items): no CTA is launched for a chunk that
    // lies entirely below its row block
    if (*reinterpret_cast<volatile const int *>(a.error_word) != 0) return;   // a hand-over failed earlier: the step is lost
    const int4 item = a.items[blockIdx.x];
    const int blk = item.x;   // row block index
    const int by = item.y;    // chunk index
    const int i0 = blk * SYM_RB;
    const int i1 = min(a.n, i0 + SYM_RB);
    const int jc0 = chunk_start(a.cm, by);
    const int j1 = min(a.n, jc0 + chunk_len(a.cm, by));
    const int j0 = max(jc0, i0);            // tiles below the block are not ours (tile aligned: chunks and blocks are)
    if (i0 >= a.n || j0 >= j1) return;      // block beyond the last body, or whole chunk below the block
    const int ntiles = (j1 - j0 + TILE_J - 1) / TILE_J;
    // this CTA's place in the reaction-sum chains of its virtual rank
    const int vloc = sym_vrank(blk) - a.v0;
    const int seq = sym_vrank_seq(blk);
    const int seg0 = by < a.cm.nbig ? by * segs_per_chunk(a.cm) : a.cm.nbig * segs_per_chunk(a.cm) + (by - a.cm.nbig);

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&s_full[s], 1);
        mbar_init_fence();
        s_flag[0] = 0;
        s_flag[1] = 0;
        // the ring slot this item will leave its row sums in: the fold that read its previous contents is a ring of
        // CTAs back -- look now, behind the prologue's latencies, and wait at the end only if it is not over yet
        s_flag[2] = item.w < 0 || peek_acquire(a.gdone + item.w) == 1;
    }
    if (tid == 32 && a.boxes) {
        // bounding box of this row block = union of its tiles' boxes
        const double *bb = a.boxes + static_cast<size_t>(i0 / TILE_J) * SYM_BOX_DOUBLES;
        double bx0 = bb[0], by0 = bb[1], bx1 = bb[2], by1 = bb[3], brmax = bb[4];
#pragma unroll
        for (int t = 1; t < SYM_RB / TILE_J; ++t) {
            if (i0 + t * TILE_J < a.n) {
                bb += SYM_BOX_DOUBLES;
                bx0 = fmin(bx0, bb[0]); by0 = fmin(by0, bb[1]);
                bx1 = fmax(bx1, bb[2]); by1 = fmax(by1, bb[3]);
                brmax = fmax(brmax, bb[4]);
            }
        }
        s_bbox[0] = bx0; s_bbox[1] = by0; s_bbox[2] = bx1; s_bbox[3] = by1; s_bbox[4] = brmax;
    }
    __syncthreads();

    auto issue = [&](int tile) {
        const int st = tile % STAGES;
        const int jb = j0 + tile * TILE_J;
        const uint32_t bytes = static_cast<uint32_t>(min(TILE_J, j1 - jb)) * sizeof(double2);
        const uint32_t box_bytes = a.boxes ? SYM_BOX_DOUBLES * 8 : 0;
        mbar_arrive_expect_tx(&s_full[st], 2 * bytes + box_bytes);
        tma_load_1d(s_pos + st * TILE_J, a.pos + jb, bytes, &s_full[st]);
        tma_load_1d(s_mr + st * TILE_J, a.mr + jb, bytes, &s_full[st]);
        if (a.boxes) tma_load_1d(s_box + st * SYM_BOX_DOUBLES, a.boxes + static_cast<size_t>(jb / TILE_J) * SYM_BOX_DOUBLES, box_bytes, &s_full[st]);
    };
    if (tid == 0) {
        for (int t = 0; t < STAGES && t < ntiles; ++t) issue(t);
    }

    SymBody b[SYM_IB];
    int idx[SYM_IB];
    int minhi = 0x7fffffff;
#pragma unroll
    for (int k = 0; k < SYM_IB; ++k) {
        idx[k] = i0 + tid + k * SYM_THREADS;
        const bool live = idx[k] < a.n;
        // padding rows: massless (no reaction) and away from every body and from the padding columns
        const double2 p = live ? a.pos[idx[k]] : make_double2(-2.0e6, -2.0e6);
        const double2 q = live ? a.mr[idx[k]] : make_double2(0.0, 1.0);
        b[k].x = p.x; b[k].y = p.y; b[k].m = q.x; b[k].r = q.y;
        b[k].ax = 0.0; b[k].ay = 0.0;
        b[k].minhi = 0x7fffffff;
    }

    for (int tile = 0; tile < ntiles; ++tile) {
        const int st = tile % STAGES;
        mbar_wait(&s_full[st], (tile / STAGES) & 1);
        const int jb = j0 + tile * TILE_J;
        const int cnt = min(TILE_J, j1 - jb);
        const double2 *tp = s_pos + st * TILE_J;
        const double2 *tm = s_mr + st * TILE_J;
        const bool diag = jb < i0 + SYM_RB;   // block-uniform

        if (diag) {
#pragma unroll 4
            for (int jj = 0; jj < cnt; ++jj) {
                const double2 pj = tp[jj];
                const double2 qj = tm[jj];
#pragma unroll
                for (int k = 0; k < SYM_IB; ++k) pair_dir(b[k], pj.x, pj.y, qj.x, qj.y, (jb + jj == idx[k]) ? 1 : 0);
            }
#pragma unroll
            for (int k = 0; k < SYM_IB; ++k) {
                if (b[k].minhi == 0) {
                    if (idx[k] < a.n)
                        b[k].ax += 2.0 * coincident_rescan(tp, tm, cnt, jb, idx[k], b[k].x, b[k].y, b[k].r, a.orig);
                    b[k].minhi = 0x7fffffff;
                }
            }
            __syncthreads();
        } else {
            // 32 j's at a time.  Lane l starts on j = g + l and moves to the next j every step; the reaction
            // accumulator of a j travels the other way (one lane down per step), so it meets every lane
            // exactly once and after 32 steps lane l holds the warp's complete reaction sum for j = g + l:
            // no reduction tree, no atomics, fixed order.
            bool far = false;
            if (a.boxes) {
                // gap between the boxes against the largest r_i + r_j, in the pair's own operations:
                // dx >= gx and dy >= gy for every pair, rounding is monotone, so dsqr >= gap2 >= md2 >= mindsqr
                const double *tb = s_box + st * SYM_BOX_DOUBLES;
                const double *rb = s_bbox;
                const double gx = fmax(0.0, fmax(tb[0] - rb[2], rb[0] - tb[2]));
                const double gy = fmax(0.0, fmax(tb[1] - rb[3], rb[1] - tb[3]));
                const double gap2 = fma(gx, gx, gy * gy);
                const double md = rb[4] + tb[4];
                far = gap2 >= md * md;
            }
            if (far) {
                if (cnt == TILE_J) {
                    for (int g = 0; g < TILE_J; g += 32) sym_group<SYM_IB, false, UNROLL, FAR_FLAVOR>(b, tp, tm, g, cnt, lane, s_col + warp * TILE_J, minhi);
                } else {
                    for (int g = 0; g < cnt; g += 32) sym_group<SYM_IB, true, UNROLL, FAR_FLAVOR>(b, tp, tm, g, cnt, lane, s_col + warp * TILE_J, minhi);
                }
            } else if (cnt == TILE_J) {
                for (int g = 0; g < TILE_J; g += 32) sym_group<SYM_IB, false, UNROLL, FLAVOR>(b, tp, tm, g, cnt, lane, s_col + warp * TILE_J, minhi);
            } else {
                for (int g = 0; g < cnt; g += 32) sym_group<SYM_IB, true, UNROLL, FLAVOR>(b, tp, tm, g, cnt, lane, s_col + warp * TILE_J, minhi);
            }
            bool hit = false;
            if ((FLAVOR & 2) && minhi == 0) {
#pragma unroll
                for (int k = 0; k < SYM_IB; ++k) b[k].minhi = 0;
                minhi = 0x7fffffff;
            }
#pragma unroll
            for (int k = 0; k < SYM_IB; ++k) {
                if (b[k].minhi == 0) {
                    if (idx[k] < a.n) {
                        b[k].ax += 2.0 * sym_row_rescan(tp, tm, cnt, b[k].x, b[k].y, b[k].r, a.orig, jb, idx[k]);
                        hit = true;
                    }
                    b[k].minhi = 0x7fffffff;
                }
            }
            if (hit) s_flag[tile & 1] = 1;
            __syncthreads();
            // column side: one thread per j, warps summed in fixed order
            for (int jt = tid; jt < cnt; jt += SYM_THREADS) {
                double cx = 0.0, cy = 0.0;
#pragma unroll
                for (int w = 0; w < SYM_WARPS; ++w) {
                    const double2 c = s_col[w * TILE_J + jt];
                    cx += c.x;
                    cy += c.y;
                }
                const double2 pj = tp[jt];
                const double2 qj = tm[jt];
                if (s_flag[tile & 1]) cx += 2.0 * sym_col_rescan(a.pos, a.mr, i0, i1, pj.x, pj.y, qj.y, a.orig, jb + jt);
                const double gmj = -NBODY_SYM_GH * qj.x;
                s_seg[(jb - jc0) % SYM_SEG + jt] = make_double2(gmj * cx, gmj * cy);
            }
            if (tid == 0) s_flag[(tile + 1) & 1] = 0;
            __syncthreads();
        }
        if (tid == 0 && tile + STAGES < ntiles) issue(tile + STAGES);
        // last tile of a segment: add the segment's reaction sums to vacc when it is this block's turn.  Only the
        // part above the block carries sums (a segment that ends with the block itself has none: nothing to add,
        // and no later block of the chain either).
        {
            const int segi = (jb - jc0) / SYM_SEG;
            const int sstart = jc0 + segi * SYM_SEG;
            const int send = min(j1, sstart + SYM_SEG);
            const int lo = max(sstart, i0 + SYM_RB);
            if (jb + TILE_J >= send && lo < send) {   // block-uniform
                int *turn = a.turn + static_cast<size_t>(vloc) * a.nseg + seg0 + segi;
                if (tid == 0) turn_wait(turn, seq, a.error_word);
                __syncthreads();
                // the additions are done by the L2 (RED.ADD.F64, nothing comes back to the SM): one IEEE add per
                // element like a load/add/store would do, and -- it being this block's turn -- nobody else touches
                // these elements until the fence below has seen them complete
                double *acc = reinterpret_cast<double *>(a.vacc + static_cast<size_t>(vloc) * a.npad);
                for (int j = lo + tid; j < send; j += SYM_THREADS) {
                    const double2 c = s_seg[j - sstart];
                    atomicAdd(acc + 2 * j, c.x);
                    atomicAdd(acc + 2 * j + 1, c.y);
                }
                __syncthreads();
                if (tid == 0) turn_pass(turn, seq + 1);   // release at GPU scope, cumulative over the barrier above
            }
        }
    }

    // Row sums.  Every item of the group but the last leaves its SYM_RB sums in its ring slice and counts itself
    // in; the last item (in dispatch order: the others have lower CTA indices, they are running or done) waits for
    // the count, adds the slices in chunk order, its own sums last, and writes the block's rows of rowsum[phase].
    {
        const int2 grp = a.groups[item.z];   // (first item, items)
        const bool folds = static_cast<int>(blockIdx.x) == grp.x + grp.y - 1;
        if (!folds) {
            if (!s_flag[2]) {   // block-uniform; normally the slot was free when the CTA started
                if (tid == 0) turn_wait(a.gdone + item.w, 1, a.error_word);
                __syncthreads();
            }
            double2 *slice = a.slices + static_cast<size_t>(blockIdx.x % SYM_RING) * SYM_RB;
#pragma unroll
            for (int k = 0; k < SYM_IB; ++k) {
                const double gm = NBODY_SYM_GH * b[k].m;
                __stcg(slice + (idx[k] - i0), make_double2(gm * b[k].ax, gm * b[k].ay));
            }
            __syncthreads();
            if (tid == 0) group_arrive(a.gcnt + item.z);
        } else {
            if (tid == 0) turn_wait(a.gcnt + item.z, grp.y - 1, a.error_word);
            __syncthreads();
            double sx[SYM_IB], sy[SYM_IB];
#pragma unroll
            for (int k = 0; k < SYM_IB; ++k) sx[k] = sy[k] = 0.0;
#pragma unroll 2
            for (int t = 0; t < grp.y - 1; ++t) {
                const double2 *slice = a.slices + static_cast<size_t>((grp.x + t) % SYM_RING) * SYM_RB + tid;
#pragma unroll
                for (int k = 0; k < SYM_IB; ++k) {
                    const double2 c = __ldcg(slice + k * SYM_THREADS);
                    sx[k] += c.x;
                    sy[k] += c.y;
                }
            }
            double2 *out = a.rowsum + (by < a.cm.nbig ? 0 : static_cast<size_t>(a.npad));
#pragma unroll
            for (int k = 0; k < SYM_IB; ++k) {
                if (idx[k] < a.n) {
                    const double gm = NBODY_SYM_GH * b[k].m;
                    out[idx[k]] = make_double2(sx[k] + gm * b[k].ax, sy[k] + gm * b[k].ay);
                }
            }
            __syncthreads();
            if (tid == 0) turn_pass(a.gdone + item.z, 1);
        }
    }
}

// The force kernel's reaction sum for (local virtual rank vl, body j); leaves vacc and the turn counts zero for
// the next launch.
__device__ __forceinline__ double2 sym_take_reactions(const SymReduceArgs &a, int j, int vl)
{
    double2 *p = a.vacc + static_cast<size_t>(vl) * a.npad + j;
    const double2 r = *p;
    *p = make_double2(0.0, 0.0);
    if (j < a.nseg) a.turn[static_cast<size_t>(vl) * a.nseg + j] = 0;
    if (vl == 0)
        for (int q = j; q < a.ngroups; q += a.n) {
            a.gcnt[q] = 0;
            a.gdone[q] = 0;
        }
    return r;
}

// vbuf[v][j] for the virtual ranks this device owns, one thread per (j, v).
__global__ void __launch_bounds__(128) sym_reduce_kernel(const SymReduceArgs a)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const int v = a.v0 + blockIdx.y;
    const int blk = j / SYM_RB;
    const double2 r = sym_take_reactions(a, j, blockIdx.y);   // v's blocks below block(j), B ascending
    double xf = r.x, yf = r.y;
    if (sym_vrank(blk) == v) {                // j's own row sums: over the long chunks, then over the short ones
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const double2 t = a.rowsum[static_cast<size_t>(p) * a.npad + j];
            xf += t.x;
            yf += t.y;
        }
    }
    a.vbuf[static_cast<size_t>(v) * a.npad + j] = make_double2(xf, yf);
}

}  // namespace nbody
