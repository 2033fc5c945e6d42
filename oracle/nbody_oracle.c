/* nbody_oracle.c -- CPU restatement of the reference's all-pairs N-body step.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (parallel-nbody_b200/,
 * nbody/, include/) may call, link or load this file; it is the checker used by
 * tests/, by __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl
 * reference legs.  It restates, in plain C on structure-of-arrays buffers, what
 * /root/reference/nbody/nbody-seq.c computes; each function cites the lines it
 * follows.  Parity status: pinned at text level against the reference's two
 * golden outputs (bin/REF_OUTPUT, tests/128_MY_REF_OUTPUT) and at bit level
 * against the reference source itself compiled here (oracle/_ref, see Makefile),
 * both checked by tests/test_oracle.py.
 *
 * trig selects the atan2/sincos used: 0 = the C library's (what the reference
 * links against), 1 = include/nbody_detmath.h (the fixed evaluation scheme the
 * exact-mode GPU kernels share).
 *
 * Build: gcc -O3 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/Makefile).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "../include/nbody_detmath.h"

/* reference nbody-seq.c:22-27 -- same literal expressions so the compiler folds them identically */
#define GRAVITY  1.1
#define FRICTION 0.01
#define DELTA_T  (0.025/5000)
#define SEED     27102015

static const double dm_tab[NBODY_DM_ATAN_TAB_LEN] = NBODY_DM_ATAN_TAB_INIT;

/* entry points for tests/test_detmath.py */
double oracle_det_atan2(double y, double x) { return nbody_dm_atan2(y, x, dm_tab); }
void oracle_det_sincos(double a, double *s, double *c) { nbody_dm_sincos(a, s, c); }
void oracle_det_atan2_v(int n, const double *y, const double *x, double *out)
{
    for (int i = 0; i < n; ++i) out[i] = nbody_dm_atan2(y[i], x[i], dm_tab);
}
void oracle_det_sincos_v(int n, const double *a, double *s, double *c)
{
    for (int i = 0; i < n; ++i) nbody_dm_sincos(a[i], &s[i], &c[i]);
}
void oracle_libm_atan2_v(int n, const double *y, const double *x, double *out)
{
    for (int i = 0; i < n; ++i) out[i] = atan2(y[i], x[i]);
}
void oracle_libm_sincos_v(int n, const double *a, double *s, double *c)
{
    for (int i = 0; i < n; ++i) { s[i] = sin(a[i]); c[i] = cos(a[i]); }
}

static inline double o_atan2(int trig, double y, double x)
{
    return trig ? nbody_dm_atan2(y, x, dm_tab) : atan2(y, x);
}
static inline void o_sincos(int trig, double a, double *s, double *c)
{
    if (trig) nbody_dm_sincos(a, s, c);
    else { *c = cos(a); *s = sin(a); }
}

/* Body initialisation: nbody-seq.c:440-449 / nbody-par.c:208-222.
 * glibc rand() sequence, 4 draws per body in the order X, Y, XV, YV.  b*b and
 * n*n are formed in double (the reference forms them in int, identical for
 * n <= 46340 and undefined behaviour above). */
void oracle_init_bodies(int n, int xdim, int ydim,
                        double *x, double *y, double *xv, double *yv,
                        double *mass, double *radius)
{
    srand(SEED);
    double diag = sqrt(1.0 * ((xdim * xdim) + (ydim * ydim)));
    for (int b = 0; b < n; ++b) {
        x[b] = (rand() % xdim);
        y[b] = (rand() % ydim);
        radius[b] = 1 + (((double)b * b + 1.0) * diag) / (25.0 * ((double)n * n + 1.0));
        mass[b] = radius[b] * radius[b] * radius[b];
        xv[b] = ((rand() % 20000) - 10000) / 2000.0;
        yv[b] = ((rand() % 20000) - 10000) / 2000.0;
    }
}

/* One oriented pair evaluation, b < c: nbody-seq.c:89-98. */
static inline void pair_force(int trig, double xb, double yb, double mb, double rb,
                              double xc, double yc, double mc, double rc,
                              double *fx, double *fy)
{
    double dx = xc - xb;
    double dy = yc - yb;
    double angle = o_atan2(trig, dy, dx);
    double dsqr = dx*dx + dy*dy;
    double mindist = rb + rc;
    double mindsqr = mindist*mindist;
    double forced = ((dsqr < mindsqr) ? mindsqr : dsqr);
    double force = mb * mc * GRAVITY / forced;
    double s, c;
    o_sincos(trig, angle, &s, &c);
    *fx = force * c;
    *fy = force * s;
}

/* clear_forces + compute_forces exactly as the sequential reference walks them
 * (nbody-seq.c:67-109): triangular loop, +f to b, -f to c. */
void oracle_forces_symmetric(int n, int trig, const double *x, const double *y,
                             const double *mass, const double *radius,
                             double *xf, double *yf)
{
    for (int b = 0; b < n; ++b) { yf[b] = xf[b] = 0; }
    for (int b = 0; b < n; ++b) {
        for (int c = b + 1; c < n; ++c) {
            double fx, fy;
            pair_force(trig, x[b], y[b], mass[b], radius[b], x[c], y[c], mass[c], radius[c], &fx, &fy);
            xf[b] += fx;
            yf[b] += fy;
            xf[c] -= fx;
            yf[c] -= fy;
        }
    }
}

/* A bounded SAMPLE of the triangular walk for CPU-baseline timing: rows b in [lo,hi) only,
 * each against every c > b (nbody-seq.c:87-107).  xf/yf receive partial sums (timing only).
 * Returns the number of pair evaluations done; each is two directed pair-interactions. */
long long oracle_forces_symmetric_sample(int n, int trig, int lo, int hi,
                                         const double *x, const double *y,
                                         const double *mass, const double *radius,
                                         double *xf, double *yf)
{
    long long evals = 0;
    for (int b = lo; b < hi; ++b) {
        for (int c = b + 1; c < n; ++c) {
            double fx, fy;
            pair_force(trig, x[b], y[b], mass[b], radius[b], x[c], y[c], mass[c], radius[c], &fx, &fy);
            xf[b] += fx;
            yf[b] += fy;
            xf[c] -= fx;
            yf[c] -= fy;
        }
        evals += n - 1 - b;
    }
    return evals;
}

/* The same sums, one body at a time: body i receives -f(j,i) for j<i then +f(i,j)
 * for j>i, j ascending -- the order in which the triangular loop above touches
 * XF(i)/YF(i).  Bit-identical to oracle_forces_symmetric (tests check it), twice
 * the evaluations, rows independent: this is the form the GPU kernel has, and the
 * form that can be sampled (rows [lo,hi) only) or threaded for a CPU baseline. */
void oracle_forces_rows(int n, int trig, int lo, int hi, int threads,
                        const double *x, const double *y,
                        const double *mass, const double *radius,
                        double *xf, double *yf)
{
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int i = lo; i < hi; ++i) {
        double ax = 0, ay = 0;
        for (int j = 0; j < i; ++j) {
            double fx, fy;
            pair_force(trig, x[j], y[j], mass[j], radius[j], x[i], y[i], mass[i], radius[i], &fx, &fy);
            ax -= fx;
            ay -= fy;
        }
        for (int j = i + 1; j < n; ++j) {
            double fx, fy;
            pair_force(trig, x[i], y[i], mass[i], radius[i], x[j], y[j], mass[j], radius[j], &fx, &fy);
            ax += fx;
            ay += fy;
        }
        xf[i] = ax;
        yf[i] = ay;
    }
}

/* oracle_forces_rows for a LIST of bodies (any order): the sums of rows[0..nrows) only, each exactly as above.
 * For sampled checks at sizes where all rows would take hours on the CPU. */
void oracle_forces_row_list(int n, int trig, int nrows, const int *rows, int threads,
                            const double *x, const double *y,
                            const double *mass, const double *radius,
                            double *xf, double *yf)
{
    if (threads < 1) threads = 1;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int k = 0; k < nrows; ++k) {
        const int i = rows[k];
        double ax = 0, ay = 0;
        for (int j = 0; j < i; ++j) {
            double fx, fy;
            pair_force(trig, x[j], y[j], mass[j], radius[j], x[i], y[i], mass[i], radius[i], &fx, &fy);
            ax -= fx;
            ay -= fy;
        }
        for (int j = i + 1; j < n; ++j) {
            double fx, fy;
            pair_force(trig, x[i], y[i], mass[i], radius[i], x[j], y[j], mass[j], radius[j], &fx, &fy);
            ax += fx;
            ay += fy;
        }
        xf[i] = ax;
        yf[i] = ay;
    }
}

/* compute_velocities: nbody-seq.c:114-130.  XF/YF are read, not modified. */
void oracle_velocities(int n, int trig, const double *xf, const double *yf,
                       const double *mass, double *xv, double *yv)
{
    for (int b = 0; b < n; ++b) {
        double vx = xv[b];
        double vy = yv[b];
        double force = sqrt(vx*vx + vy*vy) * FRICTION;
        double angle = o_atan2(trig, vy, vx);
        double s, c;
        o_sincos(trig, angle, &s, &c);
        double fx = xf[b] - (force * c);
        double fy = yf[b] - (force * s);
        xv[b] += (fx / mass[b]) * DELTA_T;
        yv[b] += (fy / mass[b]) * DELTA_T;
    }
}

/* compute_positions: nbody-seq.c:135-165.  Uses the already-updated velocity;
 * the bounce reflects the velocity and clamps to 0 or dim-1. */
void oracle_positions(int n, int xdim, int ydim, const double *x, const double *y,
                      double *xv, double *yv, double *xn_out, double *yn_out)
{
    for (int b = 0; b < n; ++b) {
        double xn = x[b] + xv[b] * DELTA_T;
        double yn = y[b] + yv[b] * DELTA_T;
        if (xn < 0) {
            xn = 0;
            xv[b] = -xv[b];
        } else if (xn >= xdim) {
            xn = xdim - 1;
            xv[b] = -xv[b];
        }
        if (yn < 0) {
            yn = 0;
            yv[b] = -yv[b];
        } else if (yn >= ydim) {
            yn = ydim - 1;
            yv[b] = -yv[b];
        }
        xn_out[b] = xn;
        yn_out[b] = yn;
    }
}

/* The step loop, nbody-seq.c:457-464.  x,y hold the current positions on entry and
 * on return (the reference's slot flip is a pointer swap here); xf,yf return the
 * last step's forces (what print(), nbody-seq.c:356-365, shows).
 * threads <= 1: triangular reference walk; threads > 1: row form, same bits. */
int oracle_run(int n, int xdim, int ydim, int steps, int trig, int threads,
               double *x, double *y, double *xf, double *yf,
               double *xv, double *yv, const double *mass, const double *radius)
{
    double *xn = (double *)malloc(sizeof(double) * (size_t)n);
    double *yn = (double *)malloc(sizeof(double) * (size_t)n);
    if (!xn || !yn) { free(xn); free(yn); return 1; }
    for (int step = 0; step < steps; ++step) {
        if (threads > 1) oracle_forces_rows(n, trig, 0, n, threads, x, y, mass, radius, xf, yf);
        else oracle_forces_symmetric(n, trig, x, y, mass, radius, xf, yf);
        oracle_velocities(n, trig, xf, yf, mass, xv, yv);
        oracle_positions(n, xdim, ydim, x, y, xv, yv, xn, yn);
        memcpy(x, xn, sizeof(double) * (size_t)n);
        memcpy(y, yn, sizeof(double) * (size_t)n);
    }
    free(xn);
    free(yn);
    return 0;
}

/* display(): nbody-seq.c:307-354.  For every pixel the first body whose disc (radius R+0.5) covers it
 * gives the tint, else black.  image = xdim*ydim*3 bytes.  (The tint product is formed in 64-bit; the
 * reference's int product is identical for bodyCt <= 524288.) */
void oracle_display(int n, int xdim, int ydim, const double *x, const double *y, const double *radius,
                    unsigned char *image)
{
    double i, j;
    for (j = 0; j < ydim; ++j) {
        for (i = 0; i < xdim; ++i) {
            unsigned char *p = image + (3 * ((long)i + ((long)j * xdim)));
            int b;
            for (b = 0; b < n; ++b) {
                double dy = y[b] - j;
                double dx = x[b] - i;
                double d = sqrt(dx*dx + dy*dy);
                if (d <= radius[b] + 0.5) break;
            }
            if (b < n) {
                int tint = (int)((0xfffLL * (b + 1)) / (n + 2));
                p[0] = (tint & 0xf) << 4;
                p[1] = (tint & 0xf0);
                p[2] = (tint & 0xf00) >> 4;
            } else {
                p[2] = (p[1] = (p[0] = 0));
            }
        }
    }
}

/* The reference's flop model, nbody-seq.c:367-380, in 64-bit. */
long long oracle_nr_flops(long long n, long long steps)
{
    return steps * (20 * (n * (n - 1) / 2) + 18 * n + 4 * n);
}
