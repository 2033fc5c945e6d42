// nbody_gpu.cu -- implementation of the C ABI in include/nbody_gpu.h.
//
// One handle drives either one device, several devices of one process (row shards,
// ncclCommInitAll), or one rank of a multi-process job (row shard, ncclCommInitRank).
// Per step and per device:  forces_kernel -> integrate_kernel -> all-gather of the new
// positions (only when sharded).  Single-device step loops are captured in a CUDA graph.
// There is no CPU fallback: every entry point that computes needs a CUDA device.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/nbody_gpu.h"
#include "nccl_dl.h"
#include "step_args.h"

namespace nbody {

static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define NB_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(NBODY_GPU_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                        __FILE__, __LINE__);                                                       \
    } while (0)

#define NB_NCCL(api, call)                                                                         \
    do {                                                                                           \
        int r_ = (call);                                                                           \
        if (r_ != NCCL_SUCCESS)                                                                    \
            return fail(NBODY_GPU_ENCCL, "%s failed: %s (%s:%d)", #call, (api)->GetErrorString(r_), \
                        __FILE__, __LINE__);                                                       \
    } while (0)

struct Shard {
    int device = 0;
    int lo = 0, hi = 0;          // rows owned
    int part_index = 0;          // position of this shard in the all-gather (rank)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double2 *pos[2] = {nullptr, nullptr};
    double2 *mr = nullptr;
    double2 *vel = nullptr;
    double2 *frc = nullptr;
    double2 *part = nullptr;
    double2 *colpart = nullptr;  // symmetric fast kernel: reaction partial sums [owned blocks][npad]
    double2 *vbuf = nullptr;     // symmetric fast kernel: per-virtual-rank force vectors [SYM_V][npad]
    // cell order + far-field test of the symmetric kernel (kernels_sort.cu)
    int *orig = nullptr;         // original index of slot s (identity until the first sort)
    double *boxes = nullptr;     // [ceil(n/TILE_J)][SYM_BOX_DOUBLES] per-tile bounding boxes, rebuilt every step
    double2 *mr0 = nullptr;      // (mass, radius) in original order, for set_state
    unsigned long long *keys[2] = {nullptr, nullptr};
    int *vals[2] = {nullptr, nullptr};
    int *orig_tmp = nullptr;
    void *sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    double2 *gat[4] = {nullptr, nullptr, nullptr, nullptr};   // gather targets of a re-sort; scratch for unsorting
    int *sym_counter = nullptr;  // work-item counter of the persistent symmetric kernel
    void *flush = nullptr;       // 256 MiB L2-flush scratch of nbody_gpu_step_profiled, kept between calls
    int v0 = 0, nv = SYM_V;      // virtual ranks owned by this shard
    PeerTable peers = {};        // fused peer-memory exchange (npeers = 0: not in use)
    unsigned long long *flags = nullptr;   // [2][MAX_PEERS] epoch words written by the peers
    int *p2p_err = nullptr;      // set by a wait kernel that timed out
    std::vector<void *> ipc_opened;
    int nblocks_owned = 0;       // colpart slots (row blocks of the owned virtual ranks, whole groups)
    NcclComm comm = nullptr;
    cudaGraphExec_t graph[2] = {nullptr, nullptr};   // indexed by the starting slot
};

}  // namespace nbody

using namespace nbody;

constexpr int NSTAGE_BUF = 3;   // pinned staging buffers: the three arrays of a fetch / an upload are in flight together

struct nbody_gpu {
    nbody_gpu_cfg cfg;
    int n = 0;
    int nshards_total = 1;       // P: shards across the whole job
    int rows_per_shard = 0;      // ceil(n / P)
    int npad = 0;                // P * rows_per_shard (leading dimension of every array)
    int nsplit = 1;              // fast mode: j-chunks
    int chunk = 0;               // fast mode: j-chunk length
    bool symmetric = false;      // fast mode: evaluate each unordered pair once (forces_sym_kernel)
    bool sorted = false;         // symmetric: device arrays are kept in cell order (Shard::orig maps slots to bodies)
    int sort_interval = 0;       // steps between re-sorts
    long long since_sort = 0;    // steps since the last sort
    long long sorts = 0;
    bool p2p = false;            // sharded symmetric: exchange fused into the kernels over peer memory (no NCCL per step)
    unsigned long long epoch = 0;
    int nblocks = 0;             // symmetric: number of SYM_RB row blocks
    int cur = 0;                 // slot holding the current positions (the reference's world->old)
    int graph_steps = 0;         // steps captured per graph launch (even), 0 = graphs off
    bool sharded = false;
    std::vector<Shard> shards;   // local shards (1, or ngpus in single-process mode)
    double2 *stage = nullptr;    // pinned host staging, NSTAGE_BUF x npad elements
    long long launches[4] = {0, 0, 0, 0};   // force, integrate, reduce, nccl
};

namespace nbody {

static void partition(int n, int nranks, int rank, int *lo, int *hi)
{
    if (nranks < 1) nranks = 1;
    const int rows = (n + nranks - 1) / nranks;
    long long l = static_cast<long long>(rank) * rows;
    long long h = l + rows;
    if (l > n) l = n;
    if (h > n) h = n;
    *lo = static_cast<int>(l);
    *hi = static_cast<int>(h);
}

// Fast mode: number of fixed j-chunks, a function of n only so that results do not
// depend on how many GPUs share the rows.
static void choose_split(int n, int requested, int *nsplit, int *chunk)
{
    int s = requested;
    if (s <= 0) {
        const long long want = (2664LL * 1024 + n - 1) / n;   // >= ~18 CTAs per SM on an 8-way shard
        s = 8;
        while (s < want && s < 64) s *= 2;
        while (s > 1 && n / s < 1024) s /= 2;
    }
    if (s > n) s = n;
    if (s < 1) s = 1;
    int c = (n + s - 1) / s;
    c = (c + TILE_J - 1) / TILE_J * TILE_J;
    *chunk = c;
    *nsplit = (n + c - 1) / c;
}

// Symmetric kernel: chunks are multiples of SYM_RB; many small chunks keep the triangular work balanced.
static void choose_split_sym(int n, int requested, int *nsplit, int *chunk)
{
    int c;
    if (requested > 0) {
        c = (n + requested - 1) / requested;
    } else {
        // short CTAs (4 tiles up to N = 262144): more waves per launch, less tail -- +0.3 % (N = 262144) to +2.8 %
        // (N = 65536) on one GPU against 8-tile chunks, and an 8-way shard still sees ~14 waves
        c = n / 256;
        if (c < 1024) c = 1024;
    }
    if (c < SYM_RB) c = SYM_RB;
    c = (c + SYM_RB - 1) / SYM_RB * SYM_RB;
    *chunk = c;
    *nsplit = (n + c - 1) / c;
}

static void free_shard(Shard &s)
{
    cudaSetDevice(s.device);
    for (int k = 0; k < 2; ++k) {
        if (s.graph[k]) cudaGraphExecDestroy(s.graph[k]);
        if (s.pos[k]) cudaFree(s.pos[k]);
    }
    if (s.mr) cudaFree(s.mr);
    if (s.vel) cudaFree(s.vel);
    if (s.frc) cudaFree(s.frc);
    if (s.part) cudaFree(s.part);
    if (s.colpart) cudaFree(s.colpart);
    if (s.sym_counter) cudaFree(s.sym_counter);
    if (s.flush) cudaFree(s.flush);
    if (s.orig) cudaFree(s.orig);
    if (s.boxes) cudaFree(s.boxes);
    if (s.mr0) cudaFree(s.mr0);
    if (s.orig_tmp) cudaFree(s.orig_tmp);
    if (s.sort_tmp) cudaFree(s.sort_tmp);
    for (int k = 0; k < 2; ++k) {
        if (s.keys[k]) cudaFree(s.keys[k]);
        if (s.vals[k]) cudaFree(s.vals[k]);
    }
    for (int k = 0; k < 4; ++k)
        if (s.gat[k]) cudaFree(s.gat[k]);
    for (void *m : s.ipc_opened) cudaIpcCloseMemHandle(m);
    s.ipc_opened.clear();
    if (s.vbuf) cudaFree(s.vbuf);
    if (s.flags) cudaFree(s.flags);
    if (s.p2p_err) cudaFree(s.p2p_err);
    if (s.ev0) cudaEventDestroy(s.ev0);
    if (s.ev1) cudaEventDestroy(s.ev1);
    if (s.comm) {
        const NcclApi *api = nccl_api(nullptr);
        if (api) api->CommDestroy(s.comm);
    }
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Shard();
}

static ForceArgs force_args(const nbody_gpu *h, const Shard &s, int slot)
{
    ForceArgs a;
    a.pos = s.pos[slot];
    a.mr = s.mr;
    a.out = (h->cfg.mode == NBODY_GPU_MODE_FAST) ? s.part : s.frc;
    a.n = h->n;
    a.npad = h->npad;
    // symmetric kernel: every device holds the complete force vectors after the exchange and integrates
    // ALL bodies (identical bits everywhere), which saves the position all-gather
    a.lo = h->symmetric ? 0 : s.lo;
    a.hi = h->symmetric ? h->n : s.hi;
    a.chunk = h->chunk;
    return a;
}

static IntegrateArgs integrate_args(const nbody_gpu *h, const Shard &s, int slot)
{
    IntegrateArgs a;
    a.pos_old = s.pos[slot];
    a.pos_new = s.pos[slot ^ 1];
    a.mr = s.mr;
    a.vel = s.vel;
    a.frc = s.frc;
    a.part = s.part;
    a.vbuf = h->symmetric ? s.vbuf : nullptr;
    a.nsplit = h->nsplit;
    a.npad = h->npad;
    // symmetric kernel: every device holds the complete force vectors after the exchange and integrates
    // ALL bodies (identical bits everywhere), which saves the position all-gather
    // (with the fused peer exchange each device integrates its own rows and stores the new positions to all)
    const bool all_rows = h->symmetric && !h->p2p;
    a.lo = all_rows ? 0 : s.lo;
    a.hi = all_rows ? h->n : s.hi;
    a.xdim = h->cfg.xdim;
    a.ydim = h->cfg.ydim;
    a.npeers = 0;
    for (int r = 0; r < MAX_PEERS; ++r) a.peer_pos_new[r] = nullptr;
    if (h->p2p) {
        a.npeers = s.peers.npeers;
        for (int r = 0; r < s.peers.npeers; ++r) a.peer_pos_new[r] = s.peers.pos[slot ^ 1][r];
    }
    return a;
}

// Enqueue the force kernel of shard s on the positions in slot `slot` (for the symmetric kernel also
// the fold of its partial sums into this device's virtual-rank vectors).
static void enqueue_forces(nbody_gpu *h, Shard &s, int slot)
{
    if (h->cfg.mode != NBODY_GPU_MODE_FAST) {
        launch_forces_exact(force_args(h, s, slot), s.stream);
    } else if (h->symmetric) {
        SymArgs a;
        a.pos = s.pos[slot];
        a.mr = s.mr;
        a.rowpart = s.part;
        a.colpart = s.colpart;
        a.boxes = s.boxes;
        a.counter = s.sym_counter;
        a.n = h->n;
        a.npad = h->npad;
        a.chunk = h->chunk;
        a.v0 = s.v0;
        a.nv = s.nv;
        if (s.boxes) launch_tile_boxes(s.pos[slot], s.mr, h->n, s.boxes, s.stream);
        launch_forces_sym(a, s.nblocks_owned, h->nsplit, s.stream);
    } else {
        launch_forces_fast(force_args(h, s, slot), h->nsplit, s.stream);
    }
}

static void enqueue_sym_reduce(nbody_gpu *h, Shard &s, bool to_peers = false)
{
    SymReduceArgs a;
    a.rowpart = s.part;
    a.colpart = s.colpart;
    a.vbuf = s.vbuf;
    a.n = h->n;
    a.npad = h->npad;
    a.chunk = h->chunk;
    a.nsplit = h->nsplit;
    a.v0 = s.v0;
    a.nv = s.nv;
    if (to_peers) launch_sym_reduce_p2p(a, s.peers, h->rows_per_shard, s.stream);
    else launch_sym_reduce(a, s.stream);
}

static void enqueue_integrate(nbody_gpu *h, Shard &s, int slot)
{
    if (h->cfg.mode == NBODY_GPU_MODE_FAST) launch_integrate_fast(integrate_args(h, s, slot), s.stream);
    else launch_integrate_exact(integrate_args(h, s, slot), s.stream);
}

}  // namespace nbody

namespace nbody {

enum { AG_POS0 = 0, AG_POS1 = 1, AG_VEL = 2, AG_FRC = 3, AG_VBUF = 4 };

// In-place all-gather on every local shard.  Row arrays (npad double2, shard r's rows at
// [r*rows_per_shard, ...)) replace MPI_Allgatherv (reference nbody-par.c:913-917); AG_VBUF gathers the
// symmetric kernel's per-virtual-rank force vectors (shard r owns SYM_V/P consecutive ones).
static int allgather(nbody_gpu *h, int which)
{
    if (!h->sharded) return NBODY_GPU_OK;
    const char *why = "";
    const NcclApi *api = nccl_api(&why);
    if (!api) return fail(NBODY_GPU_ENCCL, "NCCL unavailable: %s", why);
    const bool group = h->shards.size() > 1;
    if (group) NB_NCCL(api, api->GroupStart());
    for (Shard &s : h->shards) {
        double2 *buf;
        size_t per_shard;   // double2 elements contributed by each shard
        if (which == AG_VBUF) {
            buf = s.vbuf;
            per_shard = static_cast<size_t>(s.nv) * h->npad;
        } else {
            buf = which == AG_POS0 ? s.pos[0] : which == AG_POS1 ? s.pos[1] : which == AG_VEL ? s.vel : s.frc;
            per_shard = static_cast<size_t>(h->rows_per_shard);
        }
        const double2 *send = buf + static_cast<size_t>(s.part_index) * per_shard;
        NB_NCCL(api, api->AllGather(send, buf, per_shard * 2, NCCL_DOUBLE, s.comm, s.stream));
    }
    if (group) NB_NCCL(api, api->GroupEnd());
    h->launches[3] += static_cast<long long>(h->shards.size());
    return NBODY_GPU_OK;
}

// Put every local shard's arrays into cell order (kernels_sort.cu).  All devices hold the same positions, so
// they all arrive at the same permutation without talking to each other; with the fused peer exchange the
// velocities and last forces live sharded and are gathered first (`gather`).
static int resort(nbody_gpu *h, bool gather)
{
    if (!h->sorted) return NBODY_GPU_OK;
    if (gather && h->sharded && h->p2p) {
        int rc = allgather(h, AG_VEL);
        if (!rc) rc = allgather(h, AG_FRC);
        if (rc) return rc;
    }
    const double scale = 65535.0 / static_cast<double>(h->cfg.xdim > h->cfg.ydim ? h->cfg.xdim : h->cfg.ydim);
    const size_t bytes = sizeof(double2) * static_cast<size_t>(h->n);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        SortArgs a;
        a.pos = s.pos[h->cur];
        a.orig = s.orig;
        a.n = h->n;
        a.scale = scale;
        a.keys_in = s.keys[0];
        a.keys_out = s.keys[1];
        a.vals_in = s.vals[0];
        a.vals_out = s.vals[1];
        a.temp = s.sort_tmp;
        a.temp_bytes = s.sort_tmp_bytes;
        const int e = launch_sort(a, s.stream);
        if (e) return fail(NBODY_GPU_ECUDA, "radix sort failed: %s", cudaGetErrorString(static_cast<cudaError_t>(e)));
        PermuteArgs p;
        p.src_slot = s.vals[1];
        p.pos_in = s.pos[h->cur]; p.mr_in = s.mr; p.vel_in = s.vel; p.frc_in = s.frc; p.orig_in = s.orig;
        p.pos_out = s.gat[0]; p.mr_out = s.gat[1]; p.vel_out = s.gat[2]; p.frc_out = s.gat[3]; p.orig_out = s.orig_tmp;
        p.n = h->n;
        launch_permute(p, s.stream);
        NB_CUDA(cudaMemcpyAsync(s.pos[h->cur], s.gat[0], bytes, cudaMemcpyDeviceToDevice, s.stream));
        NB_CUDA(cudaMemcpyAsync(s.mr, s.gat[1], bytes, cudaMemcpyDeviceToDevice, s.stream));
        NB_CUDA(cudaMemcpyAsync(s.vel, s.gat[2], bytes, cudaMemcpyDeviceToDevice, s.stream));
        NB_CUDA(cudaMemcpyAsync(s.frc, s.gat[3], bytes, cudaMemcpyDeviceToDevice, s.stream));
        NB_CUDA(cudaMemcpyAsync(s.orig, s.orig_tmp, sizeof(int) * static_cast<size_t>(h->n), cudaMemcpyDeviceToDevice, s.stream));
        NB_CUDA(cudaGetLastError());
    }
    h->launches[2] += static_cast<long long>(h->shards.size()) * 3;
    h->since_sort = 0;
    h->sorts += 1;
    return NBODY_GPU_OK;
}

// The re-sort schedule is a function of the step count alone, so results do not depend on how calls to
// nbody_gpu_step are split or on the number of GPUs.
static int maybe_resort(nbody_gpu *h)
{
    if (h->sorted && h->since_sort >= h->sort_interval) return resort(h, true);
    return NBODY_GPU_OK;
}

// Optional per-phase event hooks for nbody_gpu_step_profiled: mark(shard index, phase 0..4).
struct StepMarks {
    virtual int mark(size_t shard, int phase) = 0;
    virtual ~StepMarks() {}
};

// One step on every local shard, reading slot h->cur:
//   forces [+ symmetric fold] | exchange A (symmetric & sharded: the virtual-rank force vectors)
//   | integrate | exchange B (the new positions) | flip.
// Exchanges: fused peer-memory stores + signal/wait kernels (h->p2p), else ncclAllGather.
// Phases for marks: 0 start, 1 after forces, 2 after fold, 3 after exchange A, 4 after integrate, 5 after exchange B.
static int run_one_step(nbody_gpu *h, StepMarks *marks)
{
    const size_t nsh = h->shards.size();
#define NB_MARK(phase)                                          \
    if (marks) {                                                \
        for (size_t k_ = 0; k_ < nsh; ++k_) {                   \
            int rc_ = marks->mark(k_, phase);                   \
            if (rc_) return rc_;                                \
        }                                                       \
    }
    {
        int rc = maybe_resort(h);
        if (rc) return rc;
    }
    NB_MARK(0)
    h->epoch += 1;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        enqueue_forces(h, s, h->cur);
    }
    h->launches[0] += static_cast<long long>(nsh);
    if (h->sorted) h->launches[2] += static_cast<long long>(nsh);   // tile_boxes_kernel
    NB_MARK(1)
    if (h->symmetric) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            enqueue_sym_reduce(h, s, h->p2p);
            if (h->p2p) launch_p2p_signal(s.peers, 0, h->epoch, s.stream);
        }
        h->launches[2] += static_cast<long long>(nsh) * (h->p2p ? 2 : 1);
    }
    NB_MARK(2)
    if (h->p2p) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            launch_p2p_wait(s.peers, 0, h->epoch, s.p2p_err, s.stream);
        }
        h->launches[2] += static_cast<long long>(nsh);
    } else if (h->symmetric) {
        int rc = allgather(h, AG_VBUF);
        if (rc) return rc;
    }
    NB_MARK(3)
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        enqueue_integrate(h, s, h->cur);
        if (h->p2p) launch_p2p_signal(s.peers, 1, h->epoch, s.stream);
    }
    h->launches[1] += static_cast<long long>(nsh);
    NB_MARK(4)
    if (h->p2p) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            launch_p2p_wait(s.peers, 1, h->epoch, s.p2p_err, s.stream);
        }
        h->launches[2] += static_cast<long long>(nsh) * 2;
    } else if (!h->symmetric) {
        int rc = allgather(h, h->cur ^ 1);
        if (rc) return rc;
    }
    NB_MARK(5)
#undef NB_MARK
    h->cur ^= 1;   // the reference's world->old ^= 1 (nbody-seq.c:464)
    h->since_sort += 1;
    return NBODY_GPU_OK;
}

// Interleave two host arrays into pinned staging buffer `buf` and enqueue one H2D copy per local device.
// The caller synchronises (sync_all) before the staging buffer is used again.
static int upload(nbody_gpu *h, double2 *Shard::*field, int slot, const double *a, const double *b, int buf)
{
    double2 *stage = h->stage + static_cast<size_t>(buf) * h->npad;
    for (int i = 0; i < h->n; ++i) stage[i] = make_double2(a ? a[i] : 0.0, b ? b[i] : 0.0);
    for (int i = h->n; i < h->npad; ++i) stage[i] = make_double2(0.0, 0.0);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        double2 *dst = field ? s.*field : s.pos[slot];
        NB_CUDA(cudaMemcpyAsync(dst, stage, sizeof(double2) * h->npad, cudaMemcpyHostToDevice, s.stream));
    }
    return NBODY_GPU_OK;
}

static int build_graph(nbody_gpu *h, Shard &s, int start_slot)
{
    cudaGraph_t g = nullptr;
    NB_CUDA(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
    int slot = start_slot;
    for (int k = 0; k < h->graph_steps; ++k) {
        enqueue_forces(h, s, slot);
        if (h->symmetric) enqueue_sym_reduce(h, s);
        enqueue_integrate(h, s, slot);
        slot ^= 1;
    }
    cudaError_t e = cudaStreamEndCapture(s.stream, &g);
    if (e != cudaSuccess) return fail(NBODY_GPU_ECUDA, "graph capture failed: %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&s.graph[start_slot], g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return fail(NBODY_GPU_ECUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
    return NBODY_GPU_OK;
}

static int sync_all(nbody_gpu *h)
{
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaStreamSynchronize(s.stream));
        if (h->p2p && s.p2p_err) {
            int e = 0;
            NB_CUDA(cudaMemcpy(&e, s.p2p_err, sizeof e, cudaMemcpyDeviceToHost));
            if (e) return fail(NBODY_GPU_ECUDA, "peer exchange %c timed out waiting for another GPU", e == 1 ? 'A' : 'B');
        }
    }
    return NBODY_GPU_OK;
}

static int do_steps(nbody_gpu *h, int steps)
{
    if (steps < 0) return fail(NBODY_GPU_EINVAL, "steps must be >= 0");
    int left = steps;
    if (!h->sharded && h->graph_steps > 0) {
        Shard &s = h->shards[0];
        NB_CUDA(cudaSetDevice(s.device));
        while (left >= h->graph_steps) {
            if (h->sorted && h->since_sort + h->graph_steps > h->sort_interval) {
                // a re-sort is due before or inside the next graph launch: single steps until it has happened
                int rc = run_one_step(h, nullptr);
                if (rc) return rc;
                left -= 1;
                NB_CUDA(cudaSetDevice(s.device));
                continue;
            }
            if (!s.graph[h->cur]) {
                int rc = build_graph(h, s, h->cur);
                if (rc) return rc;
            }
            NB_CUDA(cudaGraphLaunch(s.graph[h->cur], s.stream));
            h->launches[0] += h->graph_steps;
            h->launches[1] += h->graph_steps;
            if (h->symmetric) h->launches[2] += h->graph_steps * (h->sorted ? 2 : 1);
            h->since_sort += h->graph_steps;
            left -= h->graph_steps;   // graph_steps is even: h->cur is unchanged
        }
    }
    for (; left > 0; --left) {
        int rc = run_one_step(h, nullptr);
        if (rc) return rc;
    }
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaGetLastError());
    }
    return NBODY_GPU_OK;
}


// Build the peer-memory tables for the fused exchange.  Single process: cudaDeviceEnablePeerAccess between
// all local devices.  One process per GPU: CUDA IPC handles of vbuf / pos[0] / pos[1] / flags, all-gathered
// through the NCCL communicator that already exists for the bootstrap.  Returns NBODY_GPU_OK with
// h->p2p = false when peer access is not possible (the NCCL path is then used).
static int setup_p2p(nbody_gpu *h)
{
    const int P = h->nshards_total;
    if (P > MAX_PEERS) return NBODY_GPU_OK;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaMalloc(&s.flags, sizeof(unsigned long long) * 2 * MAX_PEERS));
        NB_CUDA(cudaMemset(s.flags, 0, sizeof(unsigned long long) * 2 * MAX_PEERS));
        NB_CUDA(cudaMalloc(&s.p2p_err, sizeof(int)));
        NB_CUDA(cudaMemset(s.p2p_err, 0, sizeof(int)));
    }
    if (h->shards.size() > 1) {   // one process, several devices
        for (Shard &a : h->shards) {
            for (Shard &b : h->shards) {
                if (a.device == b.device) continue;
                int ok = 0;
                NB_CUDA(cudaDeviceCanAccessPeer(&ok, a.device, b.device));
                if (!ok) return NBODY_GPU_OK;
                NB_CUDA(cudaSetDevice(a.device));
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return NBODY_GPU_OK;
                cudaGetLastError();
            }
        }
        for (Shard &a : h->shards) {
            a.peers.npeers = P;
            a.peers.me = a.part_index;
            for (Shard &b : h->shards) {
                a.peers.vbuf[b.part_index] = b.vbuf;
                a.peers.pos[0][b.part_index] = b.pos[0];
                a.peers.pos[1][b.part_index] = b.pos[1];
                a.peers.flags[b.part_index] = b.flags;
            }
        }
        h->p2p = true;
        return NBODY_GPU_OK;
    }
    // one process per GPU
    Shard &s = h->shards[0];
    const NcclApi *api = nccl_api(nullptr);
    if (!api) return NBODY_GPU_OK;
    struct Handles { cudaIpcMemHandle_t m[4]; };
    static_assert(sizeof(Handles) % sizeof(double) == 0, "handles travel as doubles");
    Handles mine;
    void *ptrs[4] = {s.vbuf, s.pos[0], s.pos[1], s.flags};
    NB_CUDA(cudaSetDevice(s.device));
    int ipc_ok = 1;
    for (int k = 0; k < 4; ++k) {
        if (cudaIpcGetMemHandle(&mine.m[k], ptrs[k]) != cudaSuccess) {
            cudaGetLastError();
            ipc_ok = 0;
            memset(&mine.m[k], 0, sizeof mine.m[k]);
        }
    }
    // every rank must take the same decision: gather the handles plus an "ok" marker
    struct Packet { Handles h; double ok; };
    Packet *dev = nullptr;
    std::vector<Packet> all(P);
    NB_CUDA(cudaMalloc(&dev, sizeof(Packet) * P));
    Packet pk;
    pk.h = mine;
    pk.ok = ipc_ok;
    NB_CUDA(cudaMemcpyAsync(dev + s.part_index, &pk, sizeof pk, cudaMemcpyHostToDevice, s.stream));
    NB_NCCL(api, api->AllGather(dev + s.part_index, dev, sizeof(Packet) / sizeof(double), NCCL_DOUBLE, s.comm, s.stream));
    NB_CUDA(cudaMemcpyAsync(all.data(), dev, sizeof(Packet) * P, cudaMemcpyDeviceToHost, s.stream));
    NB_CUDA(cudaStreamSynchronize(s.stream));
    cudaFree(dev);
    for (int r = 0; r < P; ++r)
        if (all[r].ok != 1.0) return NBODY_GPU_OK;
    s.peers.npeers = P;
    s.peers.me = s.part_index;
    int opened_all = 1;
    for (int r = 0; r < P; ++r) {
        void *q[4];
        if (r == s.part_index) {
            for (int k = 0; k < 4; ++k) q[k] = ptrs[k];
        } else {
            for (int k = 0; k < 4; ++k) {
                q[k] = nullptr;
                if (cudaIpcOpenMemHandle(&q[k], all[r].h.m[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    opened_all = 0;
                } else {
                    s.ipc_opened.push_back(q[k]);
                }
            }
        }
        s.peers.vbuf[r] = static_cast<double2 *>(q[0]);
        s.peers.pos[0][r] = static_cast<double2 *>(q[1]);
        s.peers.pos[1][r] = static_cast<double2 *>(q[2]);
        s.peers.flags[r] = static_cast<unsigned long long *>(q[3]);
    }
    // agree on the outcome (a rank that could not map a peer must not leave the others waiting for it)
    double *flag = nullptr;
    std::vector<double> oks(P);
    NB_CUDA(cudaMalloc(&flag, sizeof(double) * P));
    const double my = opened_all;
    NB_CUDA(cudaMemcpyAsync(flag + s.part_index, &my, sizeof my, cudaMemcpyHostToDevice, s.stream));
    NB_NCCL(api, api->AllGather(flag + s.part_index, flag, 1, NCCL_DOUBLE, s.comm, s.stream));
    NB_CUDA(cudaMemcpyAsync(oks.data(), flag, sizeof(double) * P, cudaMemcpyDeviceToHost, s.stream));
    NB_CUDA(cudaStreamSynchronize(s.stream));
    cudaFree(flag);
    for (int r = 0; r < P; ++r)
        if (oks[r] != 1.0) return NBODY_GPU_OK;
    h->p2p = true;
    return NBODY_GPU_OK;
}

}  // namespace nbody

// ---------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------
extern "C" {

int nbody_gpu_abi_version(void) { return NBODY_GPU_ABI_VERSION; }

const char *nbody_gpu_last_error(void) { return g_err; }

int nbody_gpu_device_count(void)
{
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return c;
}

void nbody_gpu_partition(int n, int nranks, int rank, int *lo, int *hi) { partition(n, nranks, rank, lo, hi); }

void nbody_gpu_partition_reference(int n, int nranks, int *counts, int *displs)
{
    // nbody-par.c:225-250: n/P each, then ranks n%P down to 1 get one more (rank 0 never does)
    if (nranks < 1) return;
    for (int r = 0; r < nranks; ++r) counts[r] = n / nranks;
    for (int r = n % nranks; r > 0; --r) counts[r] += 1;
    displs[0] = 0;
    for (int r = 1; r < nranks; ++r) displs[r] = displs[r - 1] + counts[r - 1];
}

void nbody_gpu_reference_init(int n, int xdim, int ydim, double *x, double *y, double *xv, double *yv, double *mass,
                              double *radius)
{
    // nbody-par.c:208-222 / nbody-seq.c:440-449 with the reference's seed (nbody-seq.c:26): four rand() draws per
    // body in the order X, Y, XV, YV; b*b and n*n formed in double (int in the reference: the same for n <= 46340,
    // undefined behaviour above)
    if (n < 1 || xdim < 1 || ydim < 1 || !x || !y || !xv || !yv || !mass || !radius) return;
    srand(27102015);
    const double diag = sqrt(1.0 * ((xdim * xdim) + (ydim * ydim)));
    for (int b = 0; b < n; ++b) {
        x[b] = (rand() % xdim);
        y[b] = (rand() % ydim);
        radius[b] = 1 + ((static_cast<double>(b) * b + 1.0) * diag) / (25.0 * (static_cast<double>(n) * n + 1.0));
        mass[b] = radius[b] * radius[b] * radius[b];
        xv[b] = ((rand() % 20000) - 10000) / 2000.0;
        yv[b] = ((rand() % 20000) - 10000) / 2000.0;
    }
}

unsigned nbody_gpu_cell_key(double x, double y, int xdim, int ydim)
{
    if (xdim < 1 || ydim < 1) return 0;
    return cell_key_host(x, y, 65535.0 / static_cast<double>(xdim > ydim ? xdim : ydim));
}

int nbody_gpu_nccl_unique_id(void *out)
{
    if (!out) return fail(NBODY_GPU_EINVAL, "out is NULL");
    const char *why = "";
    const NcclApi *api = nccl_api(&why);
    if (!api) return fail(NBODY_GPU_ENCCL, "NCCL unavailable: %s", why);
    NcclUniqueId id;
    NB_NCCL(api, api->GetUniqueId(&id));
    memcpy(out, &id, NBODY_GPU_NCCL_ID_BYTES);
    return NBODY_GPU_OK;
}

void nbody_gpu_destroy(nbody_gpu *h)
{
    if (!h) return;
    for (Shard &s : h->shards) free_shard(s);
    if (h->stage) cudaFreeHost(h->stage);
    delete h;
}

static int init_impl(nbody_gpu *h, const nbody_gpu_cfg *cfg, const double *x, const double *y, const double *xv,
                     const double *yv, const double *mass, const double *radius)
{
    h->cfg = *cfg;
    h->n = cfg->n;
    const bool multi_rank = cfg->nranks > 1;
    const bool multi_dev = cfg->ngpus > 1;
    if (multi_rank && multi_dev) return fail(NBODY_GPU_EINVAL, "ngpus > 1 and nranks > 1 are mutually exclusive");
    int ndev = 0;
    NB_CUDA(cudaGetDeviceCount(&ndev));
    if (ndev < 1) return fail(NBODY_GPU_ECUDA, "no CUDA device");
    const int P = multi_rank ? cfg->nranks : (multi_dev ? cfg->ngpus : 1);
    if (multi_dev && cfg->ngpus > ndev)
        return fail(NBODY_GPU_EINVAL, "ngpus = %d but only %d CUDA devices are visible", cfg->ngpus, ndev);
    if (multi_rank && (cfg->rank < 0 || cfg->rank >= cfg->nranks || !cfg->nccl_id))
        return fail(NBODY_GPU_EINVAL, "multi-rank init needs 0 <= rank < nranks and nccl_id");
    h->nshards_total = P;
    h->sharded = P > 1;
    h->rows_per_shard = (h->n + P - 1) / P;
    h->npad = h->rows_per_shard * P;
    h->symmetric = cfg->mode == NBODY_GPU_MODE_FAST && (SYM_V % P == 0) &&
                   (cfg->symmetric > 0 || (cfg->symmetric == 0 && h->n >= 8192));
    h->nblocks = (h->n + SYM_RB - 1) / SYM_RB;
    h->sorted = h->symmetric && (cfg->sort > 0 || (cfg->sort == 0 && h->n >= 16384));
    h->sort_interval = 256;
    if (const char *e = getenv("NBODY_SORT_INTERVAL")) {   // development knob
        if (atoi(e) > 0) h->sort_interval = atoi(e);
    }
    if (h->symmetric) choose_split_sym(h->n, cfg->jsplit, &h->nsplit, &h->chunk);
    else if (cfg->mode == NBODY_GPU_MODE_FAST) choose_split(h->n, cfg->jsplit, &h->nsplit, &h->chunk);
    else { h->nsplit = 1; h->chunk = h->n; }
    if (cfg->use_graph >= 0 && !h->sharded) {
        // small worlds are launch-bound: put more steps behind one graph launch
        h->graph_steps = (h->n <= 2048) ? 32 : (h->n <= 16384 ? 8 : 2);
    }

    const int nlocal = multi_dev ? cfg->ngpus : 1;
    h->shards.resize(nlocal);
    for (int k = 0; k < nlocal; ++k) {
        Shard &s = h->shards[k];
        if (multi_dev) s.device = k;
        else if (cfg->device >= 0) s.device = cfg->device;
        else NB_CUDA(cudaGetDevice(&s.device));
        if (s.device >= ndev) return fail(NBODY_GPU_EINVAL, "device %d not visible (%d devices)", s.device, ndev);
        s.part_index = multi_rank ? cfg->rank : k;
        partition(h->n, P, s.part_index, &s.lo, &s.hi);
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        NB_CUDA(cudaEventCreate(&s.ev0));
        NB_CUDA(cudaEventCreate(&s.ev1));
        const size_t bytes = sizeof(double2) * static_cast<size_t>(h->npad);
        NB_CUDA(cudaMalloc(&s.pos[0], bytes));
        NB_CUDA(cudaMalloc(&s.pos[1], bytes));
        NB_CUDA(cudaMalloc(&s.mr, bytes));
        NB_CUDA(cudaMalloc(&s.vel, bytes));
        NB_CUDA(cudaMalloc(&s.frc, bytes));
        NB_CUDA(cudaMemsetAsync(s.pos[1], 0, bytes, s.stream));
        NB_CUDA(cudaMemsetAsync(s.frc, 0, bytes, s.stream));
        if (cfg->mode == NBODY_GPU_MODE_FAST) {
            NB_CUDA(cudaMalloc(&s.part, bytes * h->nsplit));
            NB_CUDA(cudaMemsetAsync(s.part, 0, bytes * h->nsplit, s.stream));
            if (h->symmetric) {
                s.nv = SYM_V / P;
                s.v0 = s.part_index * s.nv;
                // colpart slots = grid.x of the force kernel: whole groups of 2*SYM_V row blocks; slots whose
                // block lies beyond the last body exit at once
                const size_t slots = static_cast<size_t>((h->nblocks + 2 * SYM_V - 1) / (2 * SYM_V)) * 2 * s.nv;
                s.nblocks_owned = static_cast<int>(slots);
                NB_CUDA(cudaMalloc(&s.colpart, bytes * slots));
                NB_CUDA(cudaMemsetAsync(s.colpart, 0, bytes * slots, s.stream));
                NB_CUDA(cudaMalloc(&s.sym_counter, sizeof(int)));
                NB_CUDA(cudaMalloc(&s.vbuf, bytes * SYM_V));
                NB_CUDA(cudaMemsetAsync(s.vbuf, 0, bytes * SYM_V, s.stream));
            }
            if (h->sorted) {
                const size_t ntiles = static_cast<size_t>((h->n + TILE_J - 1) / TILE_J);
                NB_CUDA(cudaMalloc(&s.boxes, ntiles * SYM_BOX_DOUBLES * sizeof(double)));
                NB_CUDA(cudaMalloc(&s.orig, sizeof(int) * static_cast<size_t>(h->npad)));
                NB_CUDA(cudaMalloc(&s.orig_tmp, sizeof(int) * static_cast<size_t>(h->npad)));
                NB_CUDA(cudaMalloc(&s.mr0, bytes));
                for (int q = 0; q < 2; ++q) {
                    NB_CUDA(cudaMalloc(&s.keys[q], sizeof(unsigned long long) * static_cast<size_t>(h->npad)));
                    NB_CUDA(cudaMalloc(&s.vals[q], sizeof(int) * static_cast<size_t>(h->npad)));
                }
                for (int q = 0; q < 4; ++q) NB_CUDA(cudaMalloc(&s.gat[q], bytes));
                s.sort_tmp_bytes = sort_temp_bytes(h->n);
                NB_CUDA(cudaMalloc(&s.sort_tmp, s.sort_tmp_bytes ? s.sort_tmp_bytes : 16));
            }
        }
    }
    NB_CUDA(cudaMallocHost(&h->stage, sizeof(double2) * static_cast<size_t>(h->npad) * NSTAGE_BUF));

    if (h->sharded) {
        const char *why = "";
        const NcclApi *api = nccl_api(&why);
        if (!api) return fail(NBODY_GPU_ENCCL, "NCCL unavailable: %s", why);
        if (multi_rank) {
            NcclUniqueId id;
            memcpy(&id, cfg->nccl_id, NBODY_GPU_NCCL_ID_BYTES);
            NB_CUDA(cudaSetDevice(h->shards[0].device));
            NB_NCCL(api, api->CommInitRank(&h->shards[0].comm, cfg->nranks, id, cfg->rank));
        } else {
            std::vector<NcclComm> comms(nlocal);
            std::vector<int> devs(nlocal);
            for (int k = 0; k < nlocal; ++k) devs[k] = h->shards[k].device;
            NB_NCCL(api, api->CommInitAll(comms.data(), nlocal, devs.data()));
            for (int k = 0; k < nlocal; ++k) h->shards[k].comm = comms[k];
        }
    }

    int rc = upload(h, nullptr, 0, x, y, 0);
    if (!rc) rc = upload(h, &Shard::mr, 0, mass, radius, 1);
    if (!rc) rc = upload(h, &Shard::vel, 0, xv, yv, 2);
    h->cur = 0;
    if (!rc && h->sorted) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            NB_CUDA(cudaMemcpyAsync(s.mr0, s.mr, sizeof(double2) * static_cast<size_t>(h->npad), cudaMemcpyDeviceToDevice, s.stream));
            launch_iota(s.orig, h->npad, s.stream);
        }
        rc = resort(h, false);
    }
    if (!rc && h->sharded && h->symmetric && cfg->p2p >= 0) {
        rc = setup_p2p(h);
        if (!rc && !h->p2p && cfg->p2p > 0) rc = fail(NBODY_GPU_ECUDA, "peer-memory exchange requested but peer access is unavailable");
    }
    if (!rc) rc = sync_all(h);   // the staging buffers are free again
    return rc;
}

int nbody_gpu_init(nbody_gpu **out, const nbody_gpu_cfg *cfg, const double *x, const double *y, const double *xv,
                   const double *yv, const double *mass, const double *radius)
{
    g_err[0] = 0;
    if (!out) return fail(NBODY_GPU_EINVAL, "out is NULL");
    *out = nullptr;
    if (!cfg || !x || !y || !xv || !yv || !mass || !radius) return fail(NBODY_GPU_EINVAL, "NULL argument");
    if (cfg->n < 1) return fail(NBODY_GPU_EINVAL, "n must be >= 1 (got %d)", cfg->n);
    if (cfg->xdim < 1 || cfg->ydim < 1) return fail(NBODY_GPU_EINVAL, "xdim/ydim must be >= 1");
    if (cfg->mode != NBODY_GPU_MODE_EXACT && cfg->mode != NBODY_GPU_MODE_FAST)
        return fail(NBODY_GPU_EINVAL, "unknown mode %d", cfg->mode);
    nbody_gpu *h = new (std::nothrow) nbody_gpu();
    if (!h) return fail(NBODY_GPU_ENOMEM, "out of host memory");
    int rc = init_impl(h, cfg, x, y, xv, yv, mass, radius);
    if (rc) {
        char keep[sizeof g_err];
        memcpy(keep, g_err, sizeof keep);
        nbody_gpu_destroy(h);
        memcpy(g_err, keep, sizeof keep);
        return rc;
    }
    *out = h;
    return NBODY_GPU_OK;
}

int nbody_gpu_set_state(nbody_gpu *h, const double *x, const double *y, const double *xv, const double *yv)
{
    if (!h || !x || !y || !xv || !yv) return fail(NBODY_GPU_EINVAL, "NULL argument");
    int rc = sync_all(h);
    if (!rc) rc = upload(h, nullptr, h->cur, x, y, 0);
    if (!rc) rc = upload(h, &Shard::vel, 0, xv, yv, 1);
    if (rc) return rc;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaMemsetAsync(s.frc, 0, sizeof(double2) * static_cast<size_t>(h->npad), s.stream));
        if (h->sorted) {   // the upload is in original order: back to the identity permutation, then sort again
            NB_CUDA(cudaMemcpyAsync(s.mr, s.mr0, sizeof(double2) * static_cast<size_t>(h->npad), cudaMemcpyDeviceToDevice, s.stream));
            launch_iota(s.orig, h->npad, s.stream);
        }
    }
    if (h->sorted) {
        rc = resort(h, false);
        if (rc) return rc;
    }
    return sync_all(h);   // the staging buffers are free again; no device runs ahead into a step
}

int nbody_gpu_step(nbody_gpu *h, int steps)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    return do_steps(h, steps);
}

int nbody_gpu_step_timed(nbody_gpu *h, int steps, float *ms)
{
    if (!h || !ms) return fail(NBODY_GPU_EINVAL, "NULL argument");
    int rc = sync_all(h);
    if (rc) return rc;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventRecord(s.ev0, s.stream));
    }
    rc = do_steps(h, steps);
    if (rc) return rc;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventRecord(s.ev1, s.stream));
    }
    float worst = 0.f;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventSynchronize(s.ev1));
        float t = 0.f;
        NB_CUDA(cudaEventElapsedTime(&t, s.ev0, s.ev1));
        if (t > worst) worst = t;
    }
    *ms = worst;
    return NBODY_GPU_OK;
}

namespace {
struct EventMarks : StepMarks {
    nbody_gpu *h;
    std::vector<cudaEvent_t> ev;   // [shard][step][phase 0..5]
    int steps, step = 0;
    EventMarks(nbody_gpu *h_, int steps_) : h(h_), ev(h_->shards.size() * static_cast<size_t>(steps_) * 6, nullptr), steps(steps_) {}
    cudaEvent_t &at(size_t k, int st, int phase) { return ev[(k * steps + st) * 6 + phase]; }
    int mark(size_t k, int phase) override
    {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
        NB_CUDA(cudaEventRecord(at(k, step, phase), h->shards[k].stream));
        return NBODY_GPU_OK;
    }
};
}  // namespace

int nbody_gpu_step_profiled(nbody_gpu *h, int steps, int flush_l2, float out_ms[4])
{
    if (!h || !out_ms || steps < 0) return fail(NBODY_GPU_EINVAL, "bad argument");
    int rc = sync_all(h);
    if (rc) return rc;
    for (int q = 0; q < 4; ++q) out_ms[q] = 0.f;
    if (steps == 0) return NBODY_GPU_OK;
    const size_t flush_bytes = 256u << 20;
    const size_t nsh = h->shards.size();
    EventMarks marks(h, steps);
    for (size_t k = 0; k < nsh; ++k) {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
        for (int st = 0; st < steps; ++st)
            for (int ph = 0; ph < 6; ++ph) NB_CUDA(cudaEventCreate(&marks.at(k, st, ph)));
        if (flush_l2 && !h->shards[k].flush) NB_CUDA(cudaMalloc(&h->shards[k].flush, flush_bytes));   // first call only
        NB_CUDA(cudaEventRecord(h->shards[k].ev0, h->shards[k].stream));
    }
    for (int st = 0; st < steps; ++st) {
        if (flush_l2) {
            for (size_t k = 0; k < nsh; ++k) {
                NB_CUDA(cudaSetDevice(h->shards[k].device));
                NB_CUDA(cudaMemsetAsync(h->shards[k].flush, st & 0xff, flush_bytes, h->shards[k].stream));
            }
        }
        marks.step = st;
        rc = run_one_step(h, &marks);
        if (rc) return rc;
    }
    for (size_t k = 0; k < nsh; ++k) {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
        NB_CUDA(cudaEventRecord(h->shards[k].ev1, h->shards[k].stream));
    }
    for (size_t k = 0; k < nsh; ++k) {
        Shard &s = h->shards[k];
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventSynchronize(s.ev1));
        float acc[4] = {0.f, 0.f, 0.f, 0.f}, t = 0.f;
        NB_CUDA(cudaEventElapsedTime(&acc[0], s.ev0, s.ev1));
        if (getenv("NBODY_PROFILE_SPLIT")) {   // development knob: per-step phase times of this shard
            for (int st = 0; st < steps; ++st) {
                float ph_ms[5];
                for (int ph = 0; ph < 5; ++ph) NB_CUDA(cudaEventElapsedTime(&ph_ms[ph], marks.at(k, st, ph), marks.at(k, st, ph + 1)));
                fprintf(stderr, "[nbody_gpu] shard %d step %d: force %.3f fold %.3f exchA %.3f integrate %.3f exchB %.3f ms\n",
                        s.part_index, st, ph_ms[0], ph_ms[1], ph_ms[2], ph_ms[3], ph_ms[4]);
            }
        }
        for (int st = 0; st < steps; ++st) {
            // phases: 0-1 force kernel | 1-2 symmetric fold | 2-3 exchange A | 3-4 integrate | 4-5 exchange B
            static const int bucket[5] = {1, 2, 3, 2, 3};
            for (int ph = 0; ph < 5; ++ph) {
                NB_CUDA(cudaEventElapsedTime(&t, marks.at(k, st, ph), marks.at(k, st, ph + 1)));
                acc[bucket[ph]] += t;
            }
        }
        if (!h->sharded) acc[3] = 0.f;
        for (int q = 0; q < 4; ++q)
            if (acc[q] > out_ms[q]) out_ms[q] = acc[q];
        for (int st = 0; st < steps; ++st)
            for (int ph = 0; ph < 6; ++ph) cudaEventDestroy(marks.at(k, st, ph));
        NB_CUDA(cudaGetLastError());
    }
    return NBODY_GPU_OK;
}

int nbody_gpu_sync(nbody_gpu *h)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    return sync_all(h);
}

// Device array of shard 0 -> pinned staging buffer `buf` (asynchronous), in the caller's body order: cell order
// is undone on the device first.  download_finish de-interleaves after the stream has been synchronised.
static int download_begin(nbody_gpu *h, const double2 *src, int buf, bool wanted)
{
    if (!wanted) return NBODY_GPU_OK;
    Shard &s = h->shards[0];
    if (h->sorted) {
        launch_unsort(src, s.orig, s.gat[buf], h->n, s.stream);
        src = s.gat[buf];
    }
    NB_CUDA(cudaMemcpyAsync(h->stage + static_cast<size_t>(buf) * h->npad, src, sizeof(double2) * static_cast<size_t>(h->n),
                            cudaMemcpyDeviceToHost, s.stream));
    return NBODY_GPU_OK;
}

static void download_finish(nbody_gpu *h, int buf, double *a, double *b)
{
    const double2 *stage = h->stage + static_cast<size_t>(buf) * h->npad;
    if (a && b) {
        for (int i = 0; i < h->n; ++i) {
            a[i] = stage[i].x;
            b[i] = stage[i].y;
        }
    } else if (a) {
        for (int i = 0; i < h->n; ++i) a[i] = stage[i].x;
    } else if (b) {
        for (int i = 0; i < h->n; ++i) b[i] = stage[i].y;
    }
}

int nbody_gpu_forces(nbody_gpu *h, double *xf, double *yf)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        enqueue_forces(h, s, h->cur);
        if (h->symmetric) enqueue_sym_reduce(h, s);
        h->launches[0] += 1;
        NB_CUDA(cudaGetLastError());
    }
    int rc = NBODY_GPU_OK;
    if (h->symmetric) rc = allgather(h, AG_VBUF);
    if (rc) return rc;
    if (h->cfg.mode == NBODY_GPU_MODE_FAST) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            launch_reduce_fast(integrate_args(h, s, h->cur), s.stream);   // partial sums -> frc
            h->launches[2] += 1;
        }
    }
    if (!h->symmetric || h->p2p) rc = allgather(h, AG_FRC);
    if (rc) return rc;
    Shard &s0 = h->shards[0];
    NB_CUDA(cudaSetDevice(s0.device));
    rc = download_begin(h, s0.frc, 0, xf || yf);
    if (!rc) rc = sync_all(h);
    if (rc) return rc;
    download_finish(h, 0, xf, yf);
    return NBODY_GPU_OK;
}

int nbody_gpu_fetch(nbody_gpu *h, double *x, double *y, double *xf, double *yf, double *xv, double *yv)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    int rc = NBODY_GPU_OK;
    if (h->sharded && (!h->symmetric || h->p2p)) {
        // forces and velocities live sharded during the run; gather them like the reference's
        // whole-struct MPI_Allgatherv leaves them (nbody-par.c:933-936)
        if (xv || yv) rc = allgather(h, AG_VEL);
        if (!rc && (xf || yf)) rc = allgather(h, AG_FRC);
        if (rc) return rc;
    }
    Shard &s0 = h->shards[0];
    NB_CUDA(cudaSetDevice(s0.device));
    rc = download_begin(h, s0.pos[h->cur], 0, x || y);
    if (!rc) rc = download_begin(h, s0.frc, 1, xf || yf);
    if (!rc) rc = download_begin(h, s0.vel, 2, xv || yv);
    if (!rc) rc = sync_all(h);
    if (rc) return rc;
    download_finish(h, 0, x, y);
    download_finish(h, 1, xf, yf);
    download_finish(h, 2, xv, yv);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaGetLastError());
    }
    return NBODY_GPU_OK;
}

int nbody_gpu_describe(const nbody_gpu *h, char *buf, int len)
{
    if (!h) return 0;
    const bool fast = h->cfg.mode == NBODY_GPU_MODE_FAST;
    const char *kernel = !fast ? "forces_kernel<EXACT>" : h->symmetric ? (h->n >= SYM_WIDE_N ? "forces_sym_kernel<64,8>" : "forces_sym_kernel<128,4>") : "forces_fast_kernel<2,4,1,true>";
    // FP64-pipe instructions per directed pair-interaction (SASS counts): the symmetric
    // kernel issues 19 per unordered pair = 2 interactions; exact mode is not a fixed count (divisions, branches)
    // (19 near, 17 in far-field tiles: see nbody_gpu_far_fraction)
    const double ipp = !fast ? 0.0 : h->symmetric ? 9.5 : 16.0;
    char tmp[640];
    const int need = snprintf(tmp, sizeof tmp,
                              "{\"mode\": \"%s\", \"kernel\": \"%s\", \"symmetric\": %s, \"nsplit\": %d, \"chunk\": %d, "
                              "\"row_blocks\": %d, \"shards\": %d, \"graph_steps\": %d, \"exchange\": \"%s\", \"fp64_instr_per_pair\": %.1f, "
                              "\"cell_order\": %s, \"sort_interval\": %d, \"sorts\": %lld}",
                              fast ? "fast" : "exact", kernel, h->symmetric ? "true" : "false", h->nsplit, h->chunk,
                              h->symmetric ? h->nblocks : 0, h->nshards_total, h->sharded ? 0 : h->graph_steps,
                              !h->sharded ? "none" : h->p2p ? "peer-memory stores + signal/wait kernels" : "ncclAllGather", ipp,
                              h->sorted ? "true" : "false", h->sorted ? h->sort_interval : 0, h->sorts);
    if (buf && len > 0) {
        strncpy(buf, tmp, static_cast<size_t>(len) - 1);
        buf[len - 1] = 0;
    }
    return need + 1;
}

int nbody_gpu_far_fraction(nbody_gpu *h, double *fraction)
{
    if (!h || !fraction) return fail(NBODY_GPU_EINVAL, "NULL argument");
    *fraction = 0.0;
    if (!h->sorted) return NBODY_GPU_OK;
    int rc = sync_all(h);
    if (rc) return rc;
    Shard &s = h->shards[0];
    NB_CUDA(cudaSetDevice(s.device));
    const int ntiles = (h->n + TILE_J - 1) / TILE_J;
    std::vector<double> bx(static_cast<size_t>(ntiles) * SYM_BOX_DOUBLES);
    launch_tile_boxes(s.pos[h->cur], s.mr, h->n, s.boxes, s.stream);
    NB_CUDA(cudaMemcpyAsync(bx.data(), s.boxes, bx.size() * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    NB_CUDA(cudaStreamSynchronize(s.stream));
    // the same walk and the same test as forces_sym_kernel: tiles above each row block
    const int tiles_per_block = SYM_RB / TILE_J;
    double far = 0.0, all = 0.0;
    for (int B = 0; B < h->nblocks; ++B) {
        double b[5];
        for (int t = 0; t < tiles_per_block && B * tiles_per_block + t < ntiles; ++t) {
            const double *q = &bx[static_cast<size_t>(B * tiles_per_block + t) * SYM_BOX_DOUBLES];
            if (t == 0) { for (int k = 0; k < 5; ++k) b[k] = q[k]; continue; }
            b[0] = q[0] < b[0] ? q[0] : b[0]; b[1] = q[1] < b[1] ? q[1] : b[1];
            b[2] = q[2] > b[2] ? q[2] : b[2]; b[3] = q[3] > b[3] ? q[3] : b[3];
            b[4] = q[4] > b[4] ? q[4] : b[4];
        }
        for (int T = (B + 1) * tiles_per_block; T < ntiles; ++T) {
            const double *q = &bx[static_cast<size_t>(T) * SYM_BOX_DOUBLES];
            double gx = q[0] - b[2] > b[0] - q[2] ? q[0] - b[2] : b[0] - q[2];
            double gy = q[1] - b[3] > b[1] - q[3] ? q[1] - b[3] : b[1] - q[3];
            if (gx < 0.0) gx = 0.0;
            if (gy < 0.0) gy = 0.0;
            const double md = b[4] + q[4];
            const int cnt = (T + 1) * TILE_J <= h->n ? TILE_J : h->n - T * TILE_J;
            all += cnt;
            if (gx * gx + gy * gy >= md * md) far += cnt;
        }
    }
    if (all > 0.0) *fraction = far / all;
    return NBODY_GPU_OK;
}

int nbody_gpu_render(nbody_gpu *h, unsigned char *rgb)
{
    if (!h || !rgb) return fail(NBODY_GPU_EINVAL, "NULL argument");
    Shard &s = h->shards[0];   // every device holds all current positions
    NB_CUDA(cudaSetDevice(s.device));
    const size_t npix = static_cast<size_t>(h->cfg.xdim) * h->cfg.ydim;
    int *owner = nullptr;
    unsigned char *img = nullptr;
    NB_CUDA(cudaMalloc(&owner, npix * sizeof(int)));
    NB_CUDA(cudaMalloc(&img, npix * 3));
    const double2 *rpos = s.pos[h->cur], *rmr = s.mr;
    if (h->sorted) {   // the lowest ORIGINAL index wins a pixel (nbody-seq.c:336-347): paint from body order
        launch_unsort(s.pos[h->cur], s.orig, s.gat[0], h->n, s.stream);
        rpos = s.gat[0];
        rmr = s.mr0;
    }
    launch_render(rpos, rmr, h->n, h->cfg.xdim, h->cfg.ydim, owner, img, s.stream);
    h->launches[2] += 3;
    cudaError_t e = cudaMemcpyAsync(rgb, img, npix * 3, cudaMemcpyDeviceToHost, s.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
    cudaFree(owner);
    cudaFree(img);
    if (e != cudaSuccess) return fail(NBODY_GPU_ECUDA, "render failed: %s", cudaGetErrorString(e));
    return NBODY_GPU_OK;
}

long long nbody_gpu_launch_count(const nbody_gpu *h, long long out[4])
{
    if (!h) return 0;
    long long t = 0;
    for (int k = 0; k < 4; ++k) {
        if (out) out[k] = h->launches[k];
        t += h->launches[k];
    }
    return t;
}

}  // extern "C"
