/* nbody_gpu.h -- C ABI of the B200 all-pairs N-body step (libnbody_gpu.so).
 *
 * The reference (jfdev001/parallel-nbody) has no plugin or FFI interface: its
 * boundary is the four static calls in the step loop,
 *     clear_forces / compute_forces / compute_velocities / compute_positions
 *     + MPI_Allgatherv + slot flip
 * (nbody/nbody-seq.c:457-464, nbody/nbody-par.c:901-921), all mutating one
 * `struct world *` in place.  This header is the boundary a maintainer would bind
 * instead: plain pointers and sizes, no CUDA or torch types.  Every pointer is
 * host memory owned by the caller unless stated otherwise; the library copies.
 *
 * Return value: 0 = ok, nonzero = failure (NBODY_GPU_E*); the message is in
 * nbody_gpu_last_error().  The reference's error behaviour is fprintf(stderr)+
 * exit(1) (nbody-seq.c:401-404,431-434); the host driver does exactly that with
 * the message.  There is no CPU fallback anywhere behind this interface: without
 * a CUDA device nbody_gpu_init fails.
 *
 * After a CUDA or NCCL failure inside a step, an upload or a fetch the device state is undefined: the
 * handle remembers the first such failure and every later call on it returns that error.
 *
 * Threading: one host thread drives a handle (the reference's step loop has one
 * caller thread per rank).  nbody_gpu_step only enqueues; fetch/sync wait.
 *
 * COLLECTIVE CALLS.  In a multi-process job (cfg.nranks > 1) the ranks are the reference's MPI ranks and, like
 * MPI collectives, these calls must be made by EVERY rank, in the same order, with the same arguments NULL or
 * non-NULL: nbody_gpu_init, nbody_gpu_set_state, nbody_gpu_step (and _timed/_bench/_profiled) with the same step
 * count, nbody_gpu_forces, nbody_gpu_fetch, nbody_gpu_fetch_range, nbody_gpu_destroy.
 * The reference's "rank 0 prints" pattern therefore becomes "every rank fetches, rank 0 prints".  The GPUs meet
 * at barriers inside these calls; a rank that does not turn up within 30 s (wall time) is taken for dead: the
 * waiting GPUs raise an error, every rank's handle is poisoned and the next synchronising call reports
 * "peer exchange timed out".  Host-side work between two collective calls (I/O, a check) may differ between ranks
 * by less than that.  Not collective: nbody_gpu_sync, nbody_gpu_describe, nbody_gpu_launch_count, the host helpers.
 */
#ifndef NBODY_GPU_H
#define NBODY_GPU_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NBODY_GPU_ABI_VERSION 5

enum {
    NBODY_GPU_OK = 0,
    NBODY_GPU_EINVAL = 1,   /* bad argument */
    NBODY_GPU_ECUDA = 2,    /* CUDA runtime error (no device, launch failure, ...) */
    NBODY_GPU_ENCCL = 3,    /* NCCL missing or failed */
    NBODY_GPU_ENOMEM = 4
};

enum {
    NBODY_GPU_MODE_EXACT = 0,  /* no FMA contraction, detmath trig, reference summation order:
                                  bit-identical to nbody-seq.c linked against include/nbody_detmath.h */
    NBODY_GPU_MODE_FAST = 1    /* FMA, algebraic force direction, j-range split into fixed chunks */
};

#define NBODY_GPU_NCCL_ID_BYTES 128

typedef struct nbody_gpu nbody_gpu;   /* opaque, owned by the library */

/* Replaces: struct world's scalar members (nbody-seq.c:42-50) + the MPI rank/size
 * pair (nbody-par.c:775-780).  Zero-initialise, then set what you need. */
typedef struct nbody_gpu_cfg {
    int n;          /* bodies (world->bodyCt); >= 1, no MAXBODIES cap */
    int xdim;       /* world->xdim (from the P6 header, nbody-seq.c:276) */
    int ydim;       /* world->ydim */
    int mode;       /* NBODY_GPU_MODE_* */
    int ngpus;      /* single-process sharding over devices 0..ngpus-1; <= 1 means one device */
    int device;     /* device ordinal when ngpus <= 1; -1 = current device */
    /* One process per GPU (the nbody-par MPI shape): rank/nranks as MPI_Comm_rank/size,
     * nccl_id = NBODY_GPU_NCCL_ID_BYTES bytes obtained from nbody_gpu_nccl_unique_id()
     * on rank 0 and distributed by the caller (MPI_Bcast, torch.distributed, ...).
     * nranks <= 1 disables it.  Mutually exclusive with ngpus > 1. */
    int rank;
    int nranks;
    const void *nccl_id;
    int jsplit;     /* fast mode: number of fixed j-chunks (0 = library default for n) */
    int use_graph;  /* 1 = capture the step loop in a CUDA graph (default when 0: yes), -1 = plain launches */
    int symmetric;  /* fast mode: 1 = evaluate each unordered pair once (+f to i, -f to j, like nbody-seq.c:103-106),
                       -1 = directed evaluation only, 0 = library default (symmetric on one GPU for n >= 8192) */
    int p2p;        /* sharded symmetric fast mode: 0 = fuse the per-step exchange into the kernels over NVLink peer
                       memory when the GPUs can map each other (else ncclAllGather), 1 = require it, -1 = NCCL only */
    int sort;       /* symmetric fast mode: 0 = library default (device arrays kept in cell order for n >= 16384 so that
                       most tiles take the far-field form of the pair; invisible to the caller: every entry point
                       speaks the caller's body order), 1 = always, -1 = keep creation order */
    int sort_interval; /* cell order: steps between re-sorts (0 = library default, 256).  Part of what fixes the summation
                       order: same value, same bits -- however nbody_gpu_step calls are split and on any GPU count */
} nbody_gpu_cfg;

/* Replaces calloc(world) + initialize_simulation_data + MPI_Bcast
 * (nbody-par.c:810,870-874): creates device state from the six initial arrays
 * (length n each).  Forces start at zero like the reference's calloc. */
int nbody_gpu_init(nbody_gpu **out, const nbody_gpu_cfg *cfg,
                   const double *x, const double *y,
                   const double *xv, const double *yv,
                   const double *mass, const double *radius);

/* Re-upload positions and velocities into an existing handle (same n); forces are
 * zeroed.  For harnesses that time many runs without re-creating graphs.  One host-to-device copy per
 * array straight from the caller's memory; returns when the copies have left the host buffers.  In
 * multi-rank mode every rank passes the full arrays (as after the reference's MPI_Bcast, nbody-par.c:874);
 * with the peer-memory exchange each rank uploads only its own rows and stores them into every GPU. */
int nbody_gpu_set_state(nbody_gpu *h, const double *x, const double *y,
                        const double *xv, const double *yv);

/* Replaces the body of the step loop, `steps` times (nbody-seq.c:457-464;
 * nbody-par.c:901-921 including MPI_Allgatherv and old ^= 1).  Asynchronous. */
int nbody_gpu_step(nbody_gpu *h, int steps);

/* As nbody_gpu_step, but bracketed by CUDA events on the library's own stream(s) and
 * synchronised; *ms receives the device time (max over this handle's devices). */
int nbody_gpu_step_timed(nbody_gpu *h, int steps, float *ms);

/* As nbody_gpu_step_timed -- the production path: CUDA-graph launches, or the one-launch loop of a small world --
 * with, if flush_l2 != 0, a 192 MiB scratch buffer (larger than the L2) overwritten before every launch unit,
 * inside the timed span.  What bench.py reports as `value`. */
int nbody_gpu_step_bench(nbody_gpu *h, int steps, int flush_l2, float *ms);

/* As nbody_gpu_step_timed with plain launches (no CUDA graph, no single-launch loop) and a CUDA event
 * between the phases of every step: out_ms[0] = whole span, [1] = sum of force-kernel spans, [2] = integrate kernels,
 * [3] = position exchange (0 when not sharded); each the max over this handle's devices.
 * flush_l2 != 0 overwrites a 256 MiB scratch buffer between steps (inside the whole span,
 * outside the per-phase spans) so that no step starts with the previous step's L2 contents.
 * This is the reference's --measure-comm (nbody-par.c:912,918) done with device timers. */
int nbody_gpu_step_profiled(nbody_gpu *h, int steps, int flush_l2, float out_ms[4]);

/* compute_forces only (clear_forces + compute_forces, nbody-seq.c:458-459) on the
 * current positions, into xf/yf; the handle's state -- including the last step's forces that
 * nbody_gpu_fetch returns -- is untouched.  Synchronous. */
int nbody_gpu_forces(nbody_gpu *h, double *xf, double *yf);

/* No counterpart in the reference (which owns all of its memory and has no accelerator): a force pass for callers
 * that are not allowed to allocate device memory.  Typical ones: a plug-in inside an application whose own pool
 * took the GPU's memory at start-up, so cudaMalloc has nothing to give or is reserved for the host application;
 * and a latency-sensitive loop that cannot afford cudaMalloc/cudaFree, either of which may wait for every stream
 * on the device.  The call itself performs no cudaMalloc, cudaFree or other allocating call.
 *
 * It is compute_forces in fast mode (clear_forces + compute_forces, nbody-seq.c:458-459, evaluated pair by pair in
 * both directions like nbody_gpu_forces on a handle with cfg.symmetric = -1, cfg.sort = -1 and the default jsplit:
 * same kernels, same chunks, same bits), but on the CALLER'S DEVICE arrays, in the library's device layout:
 *     pos_xy      [2n] x0 y0 x1 y1 ...            (read)
 *     mass_radius [2n] m0 R0 m1 R1 ...            (read)
 *     force_xy    [2n] xf0 yf0 xf1 yf1 ...        (written)
 * all three 16-byte aligned device memory of the CURRENT device; `stream` is a cudaStream_t passed as a pointer
 * (NULL = the default stream).  The kernels are only enqueued: force_xy is valid once the stream has caught up.
 *
 * The per-chunk partial sums, which a handle keeps in memory it allocates, live in one fixed array built into the
 * library: NBODY_GPU_STATIC_SCRATCH_BYTES on every GPU that loads the library, for the life of the process.  That
 * is why n is limited to NBODY_GPU_STATIC_MAX_BODIES; a larger n is refused with NBODY_GPU_EINVAL and a message.
 *  - The first call on each device looks up the array's device address (a one-off set-up step): make that first
 *    call during start-up, before any stream capture begins.
 *  - All calls on a device share the one array.  A call reads only what it has itself written, so nothing needs
 *    clearing between calls, but two calls whose kernels overlap on the GPU (different streams without an
 *    ordering between them) race on it and both results are garbage: keep calls on one stream or order them. */
#define NBODY_GPU_STATIC_MAX_BODIES 32768
#define NBODY_GPU_STATIC_SCRATCH_BYTES (16u << 20)
int nbody_gpu_forces_static(int n, const double *pos_xy, const double *mass_radius, double *force_xy, void *stream);

/* Replaces reading world->bodies after the loop (print(), nbody-seq.c:356-365):
 * waits for the queue, then copies the CURRENT positions, the last step's forces and
 * the velocities to the host (length n each; any pointer may be NULL).  In
 * multi-rank mode every rank receives the full state, like the reference after its
 * last MPI_Allgatherv (nbody-par.c:933-936). */
int nbody_gpu_fetch(nbody_gpu *h, double *x, double *y, double *xf, double *yf,
                    double *xv, double *yv);

/* nbody_gpu_fetch for the bodies [first, first + count) only: the arrays have `count` elements.  A rank of a
 * multi-process job that keeps its own rows on the host (nbody_gpu_partition), like the reference's ranks between
 * two MPI_Allgatherv calls, fetches just those -- the same collective rules as nbody_gpu_fetch apply: every rank
 * calls it, with the same arrays NULL or non-NULL. */
int nbody_gpu_fetch_range(nbody_gpu *h, int first, int count, double *x, double *y, double *xf, double *yf,
                          double *xv, double *yv);

/* Replaces display() (nbody-seq.c:326-354; dead code in nbody-par.c:674-702): paints the CURRENT positions
 * into rgb (xdim*ydim*3 bytes, host, row-major, 3 bytes per pixel): the first (lowest-index) body whose
 * disc of radius R+0.5 covers the pixel gives the tint, empty space is black.  Headless: a device buffer,
 * copied to the host; the driver writes it behind the P6 header when secs_per_update > 0. */
int nbody_gpu_render(nbody_gpu *h, unsigned char *rgb);

/* Fraction of the symmetric kernel's pair evaluations (tiles above the diagonal) that currently take the
 * far-field form (no clamp, 15 instead of 17 FP64-pipe instructions); 0 when the handle is not in cell order.
 * For bench.py's instruction accounting.  Synchronous. */
int nbody_gpu_far_fraction(nbody_gpu *h, double *fraction);

int nbody_gpu_sync(nbody_gpu *h);
void nbody_gpu_destroy(nbody_gpu *h);

/* Page-lock a host buffer the caller keeps for the whole run (the reference's `struct world`, allocated once:
 * nbody-par.c:810) so that nbody_gpu_set_state / nbody_gpu_fetch / nbody_gpu_init copy to and from it at full
 * PCIe speed; without it the copies still work (the driver stages pageable memory).  Unpin before freeing. */
int nbody_gpu_pin_host(void *ptr, size_t bytes);
int nbody_gpu_unpin_host(void *ptr);

/* Thread-local message of the last failure on this thread ("" if none). */
const char *nbody_gpu_last_error(void);

/* Kernels launched through this handle so far (force, integrate, reduce and NCCL calls counted
 * separately in out[0..3] if out != NULL); returns the total. */
long long nbody_gpu_launch_count(const nbody_gpu *h, long long out[4]);

/* One-line JSON description of how this handle runs the step (kernel, symmetric or directed
 * evaluation, j-chunks, FP64-pipe instructions issued per directed pair-interaction): written to
 * buf (at most len bytes, NUL-terminated); returns the length needed. For logs and bench.py. */
int nbody_gpu_describe(const nbody_gpu *h, char *buf, int len);

/* Row partition used for sharding: rank r owns rows [lo, hi).  Equal ceil(n/P) blocks
 * (the last ones may be short or empty) so that one ncclAllGather replaces
 * MPI_Allgatherv (nbody-par.c:913-917). */
void nbody_gpu_partition(int n, int nranks, int rank, int *lo, int *hi);

/* How a job of `nranks` GPUs would share n bodies (host arithmetic only, no device needed; the library's own
 * sharding code, for tests and capacity planning): out[0..1] = rows [lo, hi) this rank integrates;
 * out[7] = 1 if the symmetric kernel is used (fast mode, n >= 8192, nranks divides 8) and then out[2..3] = first
 * and number of the 8 virtual ranks owned, out[4] = row blocks (512 rows) whose pairs this rank evaluates,
 * out[5] = j-chunks, out[6] = long chunks among them; *pair_evaluations = pair formulas this rank evaluates per
 * step (symmetric: each unordered pair once, directed inside a row block; else rows x n). */
int nbody_gpu_shard_plan(int n, int mode, int nranks, int rank, int out[8], long long *pair_evaluations);

/* The symmetric kernel's work items for rank `rank` of `nranks` (host arithmetic only; the library's own table):
 * item k is CTA k of the force kernel = (row block blocks[k], j-chunk chunks[k]), in dispatch order.  Writes at most
 * `cap` items, *count = the number there are.  Chunk s covers bodies [chunk_start[s], chunk_start[s+1]) --
 * chunk_start (may be NULL) receives at most chunk_cap entries, *nchunks = the number of chunks.
 * Two hand-overs between CTAs fix the order of the sums, and tests check from this table that every CTA waited for
 * has a lower index (it is dispatched earlier, so no wait can deadlock):
 *  - the reaction sums of one (virtual rank, 1024-body segment) are added by that virtual rank's row blocks in
 *    ascending order, each CTA waiting for the previous one;
 *  - the row sums of a group groups[k] (a row block's long-chunk items, or its short-chunk items: consecutive items)
 *    are left in a ring of *ring slices (slot = k % ring) by all items but the group's last, which adds them in
 *    chunk order; guards[k] = the group whose fold must be over before item k may overwrite its slot (-1: none).
 * Any of blocks/chunks/groups/guards/ring may be NULL.  Returns NBODY_GPU_EINVAL when the symmetric kernel would
 * not be used. */
int nbody_gpu_sym_work_items(int n, int nranks, int rank, int *blocks, int *chunks, int *groups, int *guards, int cap,
                             int *count, int *chunk_start, int chunk_cap, int *nchunks, int *ring);

/* The reference's own partition rule, for comparison (nbody-par.c:225-263):
 * counts n/P, ranks 1..n%P get one more; displacements are the prefix sum. */
void nbody_gpu_partition_reference(int n, int nranks, int *counts, int *displs);

/* The reference's synthetic initial state (initialize_simulation_data, nbody-par.c:208-222; nbody-seq.c:440-449):
 * srand(27102015), four rand() draws per body (x, y, xv, yv), radius and mass from the body index.  Host code
 * only (the C library's rand() sequence is part of parity); b*b and n*n are formed in double so that n > 46340
 * is defined.  For harnesses (bench.py) that need the reference's workload without the reference. */
void nbody_gpu_reference_init(int n, int xdim, int ydim, double *x, double *y, double *xv, double *yv,
                              double *mass, double *radius);

/* The cell-order sort key of a position (host mirror of the device code): Hilbert index, 16 bits per axis, of
 * the cell (x, y) * 65535 / max(xdim, ydim).  The device sorts by (this key, original index).  Host only. */
unsigned nbody_gpu_cell_key(double x, double y, int xdim, int ydim);

/* ncclGetUniqueId through the library's NCCL binding; out = NBODY_GPU_NCCL_ID_BYTES bytes. */
int nbody_gpu_nccl_unique_id(void *out);

/* Number of visible CUDA devices (0 if none / no driver). */
int nbody_gpu_device_count(void);

/* Measured FP64 (non-tensor) peak of `device`: a dependent-chain-free DFMA kernel,
 * `ms_budget` milliseconds long; *tflops = 2 * DFMA/s / 1e12, *sm_mhz_est = the SM clock
 * implied by a clock64() bracket.  The roofline denominator of bench.py. */
int nbody_gpu_fp64_peak(int device, int ms_budget, double *tflops, double *sm_mhz_est);

int nbody_gpu_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_GPU_H */
