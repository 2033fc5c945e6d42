// nbody_gpu.cu -- implementation of the C ABI in include/nbody_gpu.h.
//
// One handle drives either one device, several devices of one process (row shards,
// ncclCommInitAll), or one rank of a multi-process job (row shard, ncclCommInitRank).
// Per step and per device:  force kernel [-> fold] -> exchange A -> integrate_kernel -> exchange B
// (exchanges only when sharded: peer-memory stores + barrier kernels, or ncclAllGather).
// Step loops are captured in CUDA graphs (one device, and sharded with the peer-memory exchange);
// a small exact-mode world runs its whole step loop inside one kernel launch (smalln.cuh).
// There is no CPU fallback: every entry point that computes needs a CUDA device.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <mutex>
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

#define NB_LAUNCH_RC(expr)                                                                         \
    do {                                                                                           \
        const int e_ = (expr);                                                                     \
        if (e_ != 0)                                                                               \
            return fail(NBODY_GPU_ECUDA, "%s failed: %s (%s:%d)", #expr,                           \
                        cudaGetErrorString(static_cast<cudaError_t>(e_)), __FILE__, __LINE__);     \
    } while (0)

constexpr int PROF_CHUNK = 64;   // steps per pass of nbody_gpu_step_profiled (fixed event pool)

// Byte offsets of the arrays other devices may address, inside one allocation per device ("arena"): one CUDA IPC
// handle per rank maps them all, and the layout is the same on every device.
struct ArenaLayout {
    size_t pos[2], vel, frc, vbuf, boxes, flags, epoch, bytes;
};

struct Shard {
    int device = 0;
    int lo = 0, hi = 0;          // rows owned
    int part_index = 0;          // position of this shard in the all-gather (rank)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_copy = nullptr;
    unsigned char *arena = nullptr;
    double2 *pos[2] = {nullptr, nullptr};
    double2 *vel = nullptr;
    double2 *frc = nullptr;
    double2 *vbuf = nullptr;     // symmetric fast kernel: per-virtual-rank force vectors [SYM_V][npad]
    double *boxes = nullptr;     // [ceil(n/TILE_J)][SYM_BOX_DOUBLES] per-tile bounding boxes of pos[cur]
    unsigned long long *flags = nullptr;   // [P2P_PHASES][MAX_PEERS] barrier counts written by the peers
    unsigned long long *epoch = nullptr;   // [P2P_PHASES] this device's own barrier counts
    double2 *mr = nullptr;
    double2 *part = nullptr;
    double2 *vacc = nullptr;     // symmetric fast kernel: reaction sums per owned virtual rank [nv][npad] (SymArgs)
    int *turn = nullptr;         // symmetric fast kernel: [nv][nseg] blocks that have added their reaction sums
    int nseg = 0;
    int *sym_err = nullptr;      // device error word: a force-kernel CTA gave up waiting for its turn (SYM_ERR_TURN)
    // cell order + far-field test of the symmetric kernel (kernels_sort.cu)
    int *orig = nullptr;         // original index of slot s (identity until the first sort)
    double2 *mr0 = nullptr;      // (mass, radius) in original order, for set_state
    unsigned long long *keys[2] = {nullptr, nullptr};
    int *vals[2] = {nullptr, nullptr};
    int *orig_tmp = nullptr;
    void *sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;
    double2 *gat[4] = {nullptr, nullptr, nullptr, nullptr};   // gather targets of a re-sort; scratch for unsorting
    double *plan = nullptr;      // [6][npad] planar staging: what the host sends (x y xv yv m r) / receives (x y xf yf xv yv)
    double2 *fscratch = nullptr; // forces-only entry point: result rows (frc itself stays the last step's forces)
    int *render_owner = nullptr; // nbody_gpu_render: per-pixel owner and the image, kept between calls
    unsigned char *render_img = nullptr;
    void *flush = nullptr;       // 256 MiB L2-flush scratch of nbody_gpu_step_profiled, kept between calls
    std::vector<cudaEvent_t> prof_ev;   // [PROF_CHUNK][6] phase events of nbody_gpu_step_profiled
    int v0 = 0, nv = SYM_V;      // virtual ranks owned by this shard
    PeerTable peers = {};        // fused peer-memory exchange (npeers = 0: not in use)
    int *p2p_err = nullptr;      // = sym_err: ... or a barrier kernel timed out (1 + phase)
    std::vector<void *> ipc_opened;
    int4 *items = nullptr;       // symmetric kernel: (row block, chunk, group, guard) work items in dispatch order (SymArgs)
    int sym_items = 0;           // symmetric kernel: grid size
    int2 *groups = nullptr;      // symmetric kernel: (first item, items) per (row block, long | short chunks) group
    int ngroups = 0;
    int *gcnt = nullptr;         // symmetric kernel: [2 * nblocks] arrivals per group
    int *gdone = nullptr;        // symmetric kernel: [2 * nblocks] group folded
    double2 *slices = nullptr;   // symmetric kernel: [SYM_RING][SYM_RB] row sums per work item
    double2 *rowsum = nullptr;   // symmetric kernel: [2][npad] row sums over the long / the short chunks
    NcclComm comm = nullptr;
    cudaGraphExec_t graph[2] = {nullptr, nullptr};   // indexed by the starting slot
};

}  // namespace nbody

using namespace nbody;

struct nbody_gpu {
    nbody_gpu_cfg cfg;
    int n = 0;
    int nshards_total = 1;       // P: shards across the whole job
    int rows_per_shard = 0;      // ceil(n / P)
    int npad = 0;                // P * rows_per_shard (leading dimension of every array)
    int nsplit = 1;              // fast mode: j-chunks
    int chunk = 0;               // fast mode, directed kernel: j-chunk length
    ChunkMap cm = {};            // fast mode, symmetric kernel: j-chunks (long ones, then one-tile ones)
    bool symmetric = false;      // fast mode: evaluate each unordered pair once (forces_sym_kernel)
    bool sorted = false;         // symmetric: device arrays are kept in cell order (Shard::orig maps slots to bodies)
    bool fused_boxes = false;    // sorted: integrate_kernel also writes the tile boxes of the new positions
    bool small = false;          // exact mode, one device, n <= SMALL_N_MAX: the step loop is ONE launch (smalln.cuh)
    int sort_interval = 0;       // steps between re-sorts
    long long since_sort = 0;    // steps since the last sort
    long long sorts = 0;
    bool p2p = false;            // sharded symmetric: exchange fused into the kernels over peer memory (no NCCL per step)
    int nblocks = 0;             // symmetric: number of SYM_RB row blocks
    int cur = 0;                 // slot holding the current positions (the reference's world->old)
    int graph_steps = 0;         // steps captured per graph launch (even), 0 = graphs off
    bool sharded = false;
    ArenaLayout lay = {};
    std::vector<Shard> shards;   // local shards (1, or ngpus in single-process mode)
    long long launches[4] = {0, 0, 0, 0};   // force, integrate, other kernels, nccl
    int failed_rc = 0;           // a step / upload / fetch failed: the device state is undefined from then on
    char failed_msg[512] = "";
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

// Directed fast kernel: number of fixed j-chunks, a function of n only so that results do not
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

// Symmetric kernel: long chunks of max(1024, n/256) bodies (4-tile CTAs up to N = 262144: many waves per launch,
// also on an 8-way shard) and, over the last ~n/32 bodies, one-tile chunks that are dispatched last and fill the
// tail of the launch (ChunkMap, step_args.h).  A requested split (cfg.jsplit) gives uniform chunks.
static ChunkMap choose_split_sym(int n, int requested)
{
    ChunkMap m;
    int c;
    if (requested > 0) {
        c = (n + requested - 1) / requested;
    } else {
        c = n / 256;
        if (c < 1024) c = 1024;
    }
    if (c < SYM_RB) c = SYM_RB;
    c = (c + SYM_RB - 1) / SYM_RB * SYM_RB;
    m.chunk = c;
    m.csmall = TILE_J;
    if (requested > 0) {
        m.nbig = (n + c - 1) / c;
        m.jsmall0 = m.nbig * c;
        m.nsplit = m.nbig;
    } else {
        m.nbig = (n - n / 32) / c;
        m.jsmall0 = m.nbig * c;
        m.nsplit = m.nbig + (n - m.jsmall0 + m.csmall - 1) / m.csmall;
    }
    return m;
}

// The ring of row-sum slices (SYM_RING slots, slot = item index % SYM_RING) is only safe while no group -- at most
// all long chunks, or all short chunks, of one row block -- has more than SYM_RING items: with a longer one, item
// i + SYM_RING of the group would have to wait for the fold of its OWN group before overwriting item i's slot, and
// that fold waits for it.  With the default split this is n > ~16.7 million bodies, or a requested jsplit > SYM_RING.
static int sym_max_group(const ChunkMap &cm) { return cm.nbig > cm.nsplit - cm.nbig ? cm.nbig : cm.nsplit - cm.nbig; }
static bool sym_ring_fits(const ChunkMap &cm) { return sym_max_group(cm) <= SYM_RING; }

// Work items of the symmetric kernel for a device that owns virtual ranks [v0, v0 + nv): every (owned row block,
// chunk) pair whose chunk reaches above the block's first row, in dispatch order -- the long chunks row block by row
// block, then the short chunks row block by row block.  The row blocks of a group of 2*SYM_V go in the order
// 16g+v (v = v0, v0+1, ...) then 16g+15-v (v = v0, v0+1, ...): every virtual rank's blocks still ascend, which
// puts every item behind its predecessor in the reaction-sum chain of its (virtual rank, segment) -- forces_sym.cuh
// waits on lower CTA indices only -- and the two are nv rows of items apart (8 rows on one GPU: the predecessor
// finished long ago; one row on an 8-way shard: it started about half a CTA lifetime earlier).  The short items
// still come last and fill the tail of the launch.
static void sym_work_items(int n, const ChunkMap &cm, int v0, int nv, std::vector<int4> *items, std::vector<int2> *groups)
{
    const int nblocks = (n + SYM_RB - 1) / SYM_RB;
    std::vector<int> rows;
    for (int g0 = 0; g0 < nblocks; g0 += 2 * SYM_V)
        for (int half = 0; half < 2; ++half)
            for (int v = v0; v < v0 + nv; ++v) {
                const int B = half == 0 ? g0 + v : g0 + 2 * SYM_V - 1 - v;
                if (B < nblocks) rows.push_back(B);
            }
    items->clear();
    groups->clear();
    for (int phase = 0; phase < 2; ++phase) {
        const int s_lo = phase == 0 ? 0 : cm.nbig;
        const int s_hi = phase == 0 ? cm.nbig : cm.nsplit;
        for (int B : rows) {
            int s = chunk_of(cm, B * SYM_RB);   // the chunk that holds the block's first row
            if (s < s_lo) s = s_lo;
            if (s >= s_hi) continue;
            const int g = static_cast<int>(groups->size());
            groups->push_back(make_int2(static_cast<int>(items->size()), s_hi - s));
            for (; s < s_hi; ++s) items->push_back(make_int4(B, s, g, -1));
        }
    }
    // Ring of row-sum slices: item i stores into slot i % SYM_RING unless it is its group's last item (that one
    // folds).  Before it does, the fold of the group that read the slot's previous contents must be over.
    std::vector<int> tenant(SYM_RING, -1);   // group of the last item that stored into the slot
    for (size_t i = 0; i < items->size(); ++i) {
        int4 &it = (*items)[i];
        const int2 grp = (*groups)[it.z];
        if (static_cast<int>(i) == grp.x + grp.y - 1) continue;   // folds, stores nothing
        it.w = tenant[i % SYM_RING];
        tenant[i % SYM_RING] = it.z;
    }
}

static int upload_sym_items(Shard &s, int n, const ChunkMap &cm, int v0, int nv)
{
    std::vector<int4> items;
    std::vector<int2> groups;
    sym_work_items(n, cm, v0, nv, &items, &groups);
    if (s.items) cudaFree(s.items);
    if (s.groups) cudaFree(s.groups);
    s.items = nullptr;
    s.groups = nullptr;
    NB_CUDA(cudaMalloc(&s.items, sizeof(int4) * (items.empty() ? 1 : items.size())));
    NB_CUDA(cudaMalloc(&s.groups, sizeof(int2) * (groups.empty() ? 1 : groups.size())));
    if (!items.empty()) NB_CUDA(cudaMemcpy(s.items, items.data(), sizeof(int4) * items.size(), cudaMemcpyHostToDevice));
    if (!groups.empty()) NB_CUDA(cudaMemcpy(s.groups, groups.data(), sizeof(int2) * groups.size(), cudaMemcpyHostToDevice));
    s.sym_items = static_cast<int>(items.size());
    s.ngroups = static_cast<int>(groups.size());
    return NBODY_GPU_OK;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static ArenaLayout arena_layout(int npad, bool symmetric, bool sorted, int n)
{
    ArenaLayout L;
    const size_t row = sizeof(double2) * static_cast<size_t>(npad);
    size_t at = 0;
    auto take = [&](size_t bytes) {
        const size_t here = at;
        at = align_up(at + bytes, 256);
        return here;
    };
    L.pos[0] = take(row);
    L.pos[1] = take(row);
    L.vel = take(row);
    L.frc = take(row);
    L.vbuf = take(symmetric ? row * SYM_V : 0);
    L.boxes = take(sorted ? static_cast<size_t>((n + TILE_J - 1) / TILE_J) * SYM_BOX_DOUBLES * sizeof(double) : 0);
    L.flags = take(sizeof(unsigned long long) * P2P_PHASES * MAX_PEERS);
    L.epoch = take(sizeof(unsigned long long) * P2P_PHASES);
    L.bytes = at;
    return L;
}

static void free_shard(Shard &s)
{
    cudaSetDevice(s.device);
    for (int k = 0; k < 2; ++k)
        if (s.graph[k]) cudaGraphExecDestroy(s.graph[k]);
    for (void *m : s.ipc_opened) cudaIpcCloseMemHandle(m);
    s.ipc_opened.clear();
    if (s.arena) cudaFree(s.arena);
    if (s.mr) cudaFree(s.mr);
    if (s.part) cudaFree(s.part);
    if (s.vacc) cudaFree(s.vacc);
    if (s.turn) cudaFree(s.turn);
    if (s.sym_err) cudaFree(s.sym_err);
    if (s.items) cudaFree(s.items);
    if (s.groups) cudaFree(s.groups);
    if (s.gcnt) cudaFree(s.gcnt);
    if (s.gdone) cudaFree(s.gdone);
    if (s.slices) cudaFree(s.slices);
    if (s.rowsum) cudaFree(s.rowsum);
    if (s.flush) cudaFree(s.flush);
    if (s.orig) cudaFree(s.orig);
    if (s.mr0) cudaFree(s.mr0);
    if (s.orig_tmp) cudaFree(s.orig_tmp);
    if (s.sort_tmp) cudaFree(s.sort_tmp);
    for (int k = 0; k < 2; ++k) {
        if (s.keys[k]) cudaFree(s.keys[k]);
        if (s.vals[k]) cudaFree(s.vals[k]);
    }
    for (int k = 0; k < 4; ++k)
        if (s.gat[k]) cudaFree(s.gat[k]);
    if (s.plan) cudaFree(s.plan);
    if (s.fscratch) cudaFree(s.fscratch);
    if (s.render_owner) cudaFree(s.render_owner);
    if (s.render_img) cudaFree(s.render_img);
    for (cudaEvent_t e : s.prof_ev)
        if (e) cudaEventDestroy(e);
    if (s.ev0) cudaEventDestroy(s.ev0);
    if (s.ev1) cudaEventDestroy(s.ev1);
    if (s.ev_copy) cudaEventDestroy(s.ev_copy);
    if (s.comm) {
        const NcclApi *api = nccl_api(nullptr);
        if (api) api->CommDestroy(s.comm);
    }
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Shard();
}

// A failed CUDA / NCCL call inside a step, an upload or a fetch leaves the device state undefined: the handle
// remembers the first such failure and every later call returns it instead of computing on garbage.
static int poison(nbody_gpu *h, int rc)
{
    if (rc != NBODY_GPU_OK && rc != NBODY_GPU_EINVAL && !h->failed_rc) {
        h->failed_rc = rc;
        snprintf(h->failed_msg, sizeof h->failed_msg, "%s", g_err);
    }
    return rc;
}

static int check_usable(const nbody_gpu *h)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    if (h->failed_rc) return fail(h->failed_rc, "handle is unusable after an earlier failure: %.400s", h->failed_msg);
    return NBODY_GPU_OK;
}

// Does the velocity / force state live sharded between steps (own rows only on each device)?
static bool state_sharded(const nbody_gpu *h) { return h->sharded && (!h->symmetric || h->p2p); }

static ForceArgs force_args(const nbody_gpu *h, const Shard &s, int slot, double2 *exact_out)
{
    ForceArgs a;
    a.pos = s.pos[slot];
    a.mr = s.mr;
    a.out = (h->cfg.mode == NBODY_GPU_MODE_FAST) ? s.part : exact_out;
    a.n = h->n;
    a.npad = h->npad;
    a.lo = s.lo;
    a.hi = s.hi;
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
    a.n = h->n;
    // symmetric kernel with the NCCL exchange: every device holds the complete force vectors after the
    // all-gather and integrates ALL bodies (identical bits everywhere), which saves the position all-gather;
    // with the fused peer exchange each device integrates its own rows and stores the new positions to all
    const bool all_rows = h->symmetric && !h->p2p;
    a.lo = all_rows ? 0 : s.lo;
    a.hi = all_rows ? h->n : s.hi;
    a.xdim = h->cfg.xdim;
    a.ydim = h->cfg.ydim;
    a.npeers = 0;
    a.error_word = s.sym_err;
    a.boxes = h->fused_boxes ? s.boxes : nullptr;
    for (int r = 0; r < MAX_PEERS; ++r) {
        a.peer_pos_new[r] = nullptr;
        a.peer_boxes[r] = nullptr;
    }
    if (h->p2p) {
        a.npeers = s.peers.npeers;
        for (int r = 0; r < s.peers.npeers; ++r) {
            a.peer_pos_new[r] = s.peers.pos[slot ^ 1][r];
            a.peer_boxes[r] = s.peers.boxes[r];
        }
    }
    return a;
}

// Enqueue the force kernel of shard s on the positions in slot `slot`.
static int enqueue_forces(nbody_gpu *h, Shard &s, int slot, double2 *exact_out)
{
    if (h->cfg.mode != NBODY_GPU_MODE_FAST) {
        NB_LAUNCH_RC(launch_forces_exact(force_args(h, s, slot, exact_out), s.stream));
    } else if (h->symmetric) {
        SymArgs a;
        a.pos = s.pos[slot];
        a.mr = s.mr;
        a.slices = s.slices;
        a.rowsum = s.rowsum;
        a.groups = s.groups;
        a.gcnt = s.gcnt;
        a.gdone = s.gdone;
        a.vacc = s.vacc;
        a.turn = s.turn;
        a.nseg = s.nseg;
        a.error_word = s.sym_err;
        a.boxes = s.boxes;
        a.orig = h->sorted ? s.orig : nullptr;
        a.n = h->n;
        a.npad = h->npad;
        a.cm = h->cm;
        a.v0 = s.v0;
        a.nv = s.nv;
        a.items = s.items;
#ifdef NBODY_DEV
        // development knob (read at init, see init_impl): the work items are those of ONE virtual rank only -- what
        // one GPU of an 8-way shard runs -- for timing and ncu on a single device; the step's results are garbage
        if (const char *e = getenv("NBODY_DEV_VRANK")) {
            a.v0 = atoi(e) % SYM_V;
            a.nv = 1;
        }
#endif
        if (h->sorted && !h->fused_boxes) {   // shard boundaries cut through tiles: the boxes get their own launch
            launch_tile_boxes(s.pos[slot], s.mr, h->n, s.boxes, s.stream);
            h->launches[2] += 1;
        }
        NB_LAUNCH_RC(launch_forces_sym(a, s.sym_items, s.stream));
    } else {
        launch_forces_fast(force_args(h, s, slot, nullptr), h->nsplit, s.stream);
    }
    h->launches[0] += 1;
    return NBODY_GPU_OK;
}

static void enqueue_sym_reduce(nbody_gpu *h, Shard &s, bool to_peers)
{
    SymReduceArgs a;
    a.rowsum = s.rowsum;
    a.gcnt = s.gcnt;
    a.gdone = s.gdone;
    a.ngroups = 2 * h->nblocks;
    a.vacc = s.vacc;
    a.turn = s.turn;
    a.nseg = s.nseg;
    a.vbuf = s.vbuf;
    a.n = h->n;
    a.npad = h->npad;
    a.cm = h->cm;
    a.v0 = s.v0;
    a.nv = s.nv;
    if (to_peers) launch_sym_reduce_p2p(a, s.peers, h->rows_per_shard, s.stream);
    else launch_sym_reduce(a, s.stream);
    h->launches[2] += 1;
}

static void enqueue_integrate(nbody_gpu *h, Shard &s, int slot)
{
    if (h->cfg.mode == NBODY_GPU_MODE_FAST) launch_integrate_fast(integrate_args(h, s, slot), s.stream);
    else launch_integrate_exact(integrate_args(h, s, slot), s.stream);
    h->launches[1] += 1;
}

static void enqueue_barrier(nbody_gpu *h, Shard &s, int phase)
{
    launch_p2p_barrier(s.peers, phase, s.p2p_err, s.stream);
    h->launches[2] += 1;
}

enum { AG_POS0 = 0, AG_POS1 = 1, AG_VEL = 2, AG_FRC = 3, AG_VBUF = 4, AG_SCRATCH = 5 };

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
            buf = which == AG_POS0 ? s.pos[0] : which == AG_POS1 ? s.pos[1] : which == AG_VEL ? s.vel
                : which == AG_FRC ? s.frc : s.fscratch;
            per_shard = static_cast<size_t>(h->rows_per_shard);
        }
        const double2 *send = buf + static_cast<size_t>(s.part_index) * per_shard;
        NB_NCCL(api, api->AllGather(send, buf, per_shard * 2, NCCL_DOUBLE, s.comm, s.stream));
    }
    if (group) NB_NCCL(api, api->GroupEnd());
    h->launches[3] += static_cast<long long>(h->shards.size());
    return NBODY_GPU_OK;
}

// Make the sharded velocity / force arrays complete on every device (before a fetch or a re-sort; what the
// reference's whole-struct MPI_Allgatherv does every step, nbody-par.c:933-936).  Peer-mapped jobs: every device
// stores its rows into all peers between two barriers; otherwise two ncclAllGather calls.
static int complete_state(nbody_gpu *h, bool want_vel, bool want_frc)
{
    if (!state_sharded(h) || (!want_vel && !want_frc)) return NBODY_GPU_OK;
    if (!h->p2p) {
        int rc = want_vel ? allgather(h, AG_VEL) : NBODY_GPU_OK;
        if (!rc && want_frc) rc = allgather(h, AG_FRC);
        return rc;
    }
    for (Shard &s : h->shards) {   // nobody is still working on the arrays we are about to store into
        NB_CUDA(cudaSetDevice(s.device));
        enqueue_barrier(h, s, 2);
    }
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        PublishArgs p;
        p.src[0] = want_vel ? s.vel : nullptr;
        p.src[1] = want_frc ? s.frc : nullptr;
        for (int r = 0; r < MAX_PEERS; ++r) {
            p.dst[0][r] = r < s.peers.npeers ? s.peers.vel[r] : nullptr;
            p.dst[1][r] = r < s.peers.npeers ? s.peers.frc[r] : nullptr;
        }
        p.npeers = s.peers.npeers;
        p.me = s.peers.me;
        p.lo = s.lo;
        p.hi = s.hi;
        launch_publish(p, s.stream);
        h->launches[2] += 1;
        enqueue_barrier(h, s, 2);
        NB_CUDA(cudaGetLastError());
    }
    return NBODY_GPU_OK;
}

// Put every local shard's arrays into cell order (kernels_sort.cu).  All devices hold the same positions, so
// they all arrive at the same permutation without talking to each other; sharded velocities and last forces
// are completed first (`gather`).  Ends with the tile boxes of the new order.
static int resort(nbody_gpu *h, bool gather)
{
    if (!h->sorted) return NBODY_GPU_OK;
    if (gather) {
        int rc = complete_state(h, true, true);
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
        launch_tile_boxes(s.pos[h->cur], s.mr, h->n, s.boxes, s.stream);
        NB_CUDA(cudaGetLastError());
    }
    h->launches[2] += static_cast<long long>(h->shards.size()) * 4;
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

// Optional per-phase event hooks for nbody_gpu_step_profiled: mark(shard index, phase 0..5).
struct StepMarks {
    virtual int mark(size_t shard, int phase) = 0;
    virtual ~StepMarks() {}
};

// One step on every local shard, reading slot h->cur:
//   forces [+ symmetric fold] | exchange A (symmetric & sharded: the virtual-rank force vectors)
//   | integrate | exchange B (the new positions) | flip.
// Exchanges: fused peer-memory stores + barrier kernels (h->p2p), else ncclAllGather.
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
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        int rc = enqueue_forces(h, s, h->cur, s.frc);
        if (rc) return rc;
    }
    NB_MARK(1)
    if (h->symmetric) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            enqueue_sym_reduce(h, s, h->p2p);
        }
    }
    NB_MARK(2)
    if (h->p2p) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            enqueue_barrier(h, s, 0);
        }
    } else if (h->symmetric) {
        int rc = allgather(h, AG_VBUF);
        if (rc) return rc;
    }
    NB_MARK(3)
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        enqueue_integrate(h, s, h->cur);
    }
    NB_MARK(4)
    if (h->p2p) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            enqueue_barrier(h, s, 1);
        }
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

// The step loop as a CUDA graph: h->graph_steps steps (even: the slot parity is preserved) of shard s, starting
// from slot `start_slot`.  One device, or a shard of a peer-mapped job (the barrier kernels keep their own
// counts in device memory, so nothing in the graph changes from launch to launch).
static int build_graph(nbody_gpu *h, Shard &s, int start_slot)
{
    cudaGraph_t g = nullptr;
    long long keep[4];
    memcpy(keep, h->launches, sizeof keep);   // counted at launch time, not at capture time
    NB_CUDA(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
    int slot = start_slot;
    int rc = NBODY_GPU_OK;
    for (int k = 0; k < h->graph_steps && !rc; ++k) {
        rc = enqueue_forces(h, s, slot, s.frc);
        if (rc) break;
        if (h->symmetric) enqueue_sym_reduce(h, s, h->p2p);
        if (h->p2p) enqueue_barrier(h, s, 0);
        enqueue_integrate(h, s, slot);
        if (h->p2p) enqueue_barrier(h, s, 1);
        slot ^= 1;
    }
    cudaError_t e = cudaStreamEndCapture(s.stream, &g);
    memcpy(h->launches, keep, sizeof keep);
    if (rc) {
        if (g) cudaGraphDestroy(g);
        return rc;
    }
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
        if (s.sym_err) {
            int e = 0;
            NB_CUDA(cudaMemcpy(&e, s.sym_err, sizeof e, cudaMemcpyDeviceToHost));
            if (e == SYM_ERR_TURN) return fail(NBODY_GPU_ECUDA, "force kernel: a CTA gave up waiting for its turn to add reaction sums");
            if (e) return fail(NBODY_GPU_ECUDA, "peer exchange %c timed out waiting for another GPU", e == 1 ? 'A' : e == 2 ? 'B' : 'S');
        }
    }
    return NBODY_GPU_OK;
}

// Kernel launches of one step of one shard (for the launch counters of a graph launch).
static void count_graph_launch(nbody_gpu *h)
{
    const long long k = h->graph_steps;
    h->launches[0] += k;
    h->launches[1] += k;
    long long other = 0;
    if (h->symmetric) other += 1;                          // fold
    if (h->sorted && !h->fused_boxes) other += 1;          // tile boxes
    if (h->p2p) other += 2;                                // barriers
    h->launches[2] += k * other;
}

static int do_steps(nbody_gpu *h, int steps)
{
    if (steps < 0) return fail(NBODY_GPU_EINVAL, "steps must be >= 0");
    int left = steps;
    if (h->small && left > 0) {
        // the whole loop in one launch: the world lives in shared memory (smalln.cuh)
        Shard &s = h->shards[0];
        NB_CUDA(cudaSetDevice(s.device));
        SmallArgs a;
        a.pos[0] = s.pos[0];
        a.pos[1] = s.pos[1];
        a.mr = s.mr;
        a.vel = s.vel;
        a.frc = s.frc;
        a.n = h->n;
        a.steps = left;
        a.cur = h->cur;
        a.xdim = h->cfg.xdim;
        a.ydim = h->cfg.ydim;
        NB_LAUNCH_RC(launch_small_world(a, s.stream));
        h->launches[0] += 1;
        h->cur ^= (left & 1);
        left = 0;
    }
    if (h->graph_steps > 0 && (!h->sharded || h->p2p)) {
        while (left >= h->graph_steps) {
            if (h->sorted && h->since_sort + h->graph_steps > h->sort_interval) {
                // a re-sort is due before or inside the next graph launch: single steps until it has happened
                int rc = run_one_step(h, nullptr);
                if (rc) return rc;
                left -= 1;
                continue;
            }
            for (Shard &s : h->shards) {
                NB_CUDA(cudaSetDevice(s.device));
                if (!s.graph[h->cur]) {
                    int rc = build_graph(h, s, h->cur);
                    if (rc) return rc;
                }
            }
            for (Shard &s : h->shards) {
                NB_CUDA(cudaSetDevice(s.device));
                NB_CUDA(cudaGraphLaunch(s.graph[h->cur], s.stream));
                count_graph_launch(h);
            }
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

static void point_into_arena(const ArenaLayout &L, unsigned char *base, double2 **pos0, double2 **pos1, double2 **vel,
                             double2 **frc, double2 **vbuf, double **boxes, unsigned long long **flags)
{
    *pos0 = reinterpret_cast<double2 *>(base + L.pos[0]);
    *pos1 = reinterpret_cast<double2 *>(base + L.pos[1]);
    *vel = reinterpret_cast<double2 *>(base + L.vel);
    *frc = reinterpret_cast<double2 *>(base + L.frc);
    *vbuf = reinterpret_cast<double2 *>(base + L.vbuf);
    *boxes = reinterpret_cast<double *>(base + L.boxes);
    *flags = reinterpret_cast<unsigned long long *>(base + L.flags);
}

static void fill_peer(Shard &s, const ArenaLayout &L, int r, unsigned char *base)
{
    point_into_arena(L, base, &s.peers.pos[0][r], &s.peers.pos[1][r], &s.peers.vel[r], &s.peers.frc[r], &s.peers.vbuf[r],
                     &s.peers.boxes[r], &s.peers.flags[r]);
}

// Build the peer-memory tables for the fused exchange.  Single process: cudaDeviceEnablePeerAccess between
// all local devices.  One process per GPU: one CUDA IPC handle per rank (the arena), all-gathered through the
// NCCL communicator that already exists for the bootstrap.  Returns NBODY_GPU_OK with h->p2p = false when peer
// access is not possible (the NCCL path is then used).
static int setup_p2p(nbody_gpu *h)
{
    const int P = h->nshards_total;
    if (P > MAX_PEERS) return NBODY_GPU_OK;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        s.p2p_err = s.sym_err;   // one error word per device (peer exchange implies the symmetric kernel)
        s.peers.epoch = s.epoch;
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
            for (Shard &b : h->shards) fill_peer(a, h->lay, b.part_index, b.arena);
        }
        // the arenas are zeroed (flags, counts) before any device signals into them
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            NB_CUDA(cudaStreamSynchronize(s.stream));
        }
        h->p2p = true;
        return NBODY_GPU_OK;
    }
    // one process per GPU
    Shard &s = h->shards[0];
    const NcclApi *api = nccl_api(nullptr);
    if (!api) return NBODY_GPU_OK;
    struct Packet { cudaIpcMemHandle_t m; double ok; };
    static_assert(sizeof(Packet) % sizeof(double) == 0, "packets travel as doubles");
    Packet pk;
    NB_CUDA(cudaSetDevice(s.device));
    pk.ok = 1.0;
    if (cudaIpcGetMemHandle(&pk.m, s.arena) != cudaSuccess) {
        cudaGetLastError();
        pk.ok = 0.0;
        memset(&pk.m, 0, sizeof pk.m);
    }
    // every rank must take the same decision: gather the handles plus an "ok" marker
    Packet *dev = nullptr;
    std::vector<Packet> all(P);
    NB_CUDA(cudaMalloc(&dev, sizeof(Packet) * P));
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
        void *q = s.arena;
        if (r != s.part_index) {
            q = nullptr;
            if (cudaIpcOpenMemHandle(&q, all[r].m, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                opened_all = 0;
                q = s.arena;   // never dereferenced: the ranks agree below not to use the table
            } else {
                s.ipc_opened.push_back(q);
            }
        }
        fill_peer(s, h->lay, r, static_cast<unsigned char *>(q));
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

// Upload the caller's planar arrays (body order) into slot `slot` and the velocity array; mass/radius too when
// given (init).  One H2D copy per array straight from the caller's memory -- at full PCIe speed when the caller
// registered it (nbody_gpu_pin_host), staged by the driver when it is pageable -- then a kernel interleaves into
// the double2 arrays.  Peer-mapped jobs: every rank uploads only ITS rows and the kernel stores them into every
// device (between two barriers), like the root's MPI_Bcast but with each rank sending an eighth
// (nbody-par.c:874).  Returns after the copies have left the host buffers; the kernels may still be queued.
static int upload_state(nbody_gpu *h, int slot, const double *x, const double *y, const double *xv, const double *yv,
                        const double *mass, const double *radius)
{
    const size_t np = static_cast<size_t>(h->npad);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        const int ulo = h->p2p ? s.lo : 0, uhi = h->p2p ? s.hi : h->n;
        const size_t cnt = static_cast<size_t>(uhi > ulo ? uhi - ulo : 0) * sizeof(double);
        const double *src[4] = {x, y, xv, yv};
        for (int k = 0; k < 4; ++k)
            if (cnt) NB_CUDA(cudaMemcpyAsync(s.plan + k * np + ulo, src[k] + ulo, cnt, cudaMemcpyHostToDevice, s.stream));
        if (mass && radius) {
            NB_CUDA(cudaMemcpyAsync(s.plan + 4 * np, mass, sizeof(double) * h->n, cudaMemcpyHostToDevice, s.stream));
            NB_CUDA(cudaMemcpyAsync(s.plan + 5 * np, radius, sizeof(double) * h->n, cudaMemcpyHostToDevice, s.stream));
        }
        NB_CUDA(cudaEventRecord(s.ev_copy, s.stream));
    }
    if (h->p2p) {
        for (Shard &s : h->shards) {   // no peer is still reading the arrays we are about to store into
            NB_CUDA(cudaSetDevice(s.device));
            enqueue_barrier(h, s, 2);
        }
    }
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        ScatterArgs a;
        a.lo = h->p2p ? s.lo : 0;
        a.hi = h->p2p ? s.hi : h->n;
        for (int r = 0; r < MAX_PEERS; ++r) a.dst[r] = nullptr;
        a.a = s.plan;
        a.b = s.plan + np;
        a.ndst = h->p2p ? s.peers.npeers : 1;
        for (int r = 0; r < a.ndst; ++r) a.dst[r] = h->p2p ? s.peers.pos[slot][r] : s.pos[slot];
        launch_scatter(a, s.stream);
        a.a = s.plan + 2 * np;
        a.b = s.plan + 3 * np;
        for (int r = 0; r < a.ndst; ++r) a.dst[r] = h->p2p ? s.peers.vel[r] : s.vel;
        launch_scatter(a, s.stream);
        h->launches[2] += 2;
        if (mass && radius) {
            a.lo = 0;
            a.hi = h->n;
            a.a = s.plan + 4 * np;
            a.b = s.plan + 5 * np;
            a.ndst = 1;
            a.dst[0] = s.mr;
            launch_scatter(a, s.stream);
            h->launches[2] += 1;
        }
        NB_CUDA(cudaMemsetAsync(s.frc, 0, sizeof(double2) * np, s.stream));
        if (h->p2p) enqueue_barrier(h, s, 2);
        NB_CUDA(cudaGetLastError());
    }
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventSynchronize(s.ev_copy));
    }
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

int nbody_gpu_shard_plan(int n, int mode, int nranks, int rank, int out[8], long long *pair_evaluations)
{
    if (n < 1 || nranks < 1 || rank < 0 || rank >= nranks || !out) return fail(NBODY_GPU_EINVAL, "bad argument");
    int lo = 0, hi = 0;
    partition(n, nranks, rank, &lo, &hi);
    const bool fast = mode == NBODY_GPU_MODE_FAST;
    const bool symmetric = fast && (SYM_V % nranks == 0) && n >= 8192 && sym_ring_fits(choose_split_sym(n, 0));
    for (int k = 0; k < 8; ++k) out[k] = 0;
    out[0] = lo;
    out[1] = hi;
    long long evals = 0;
    if (symmetric) {
        const ChunkMap cm = choose_split_sym(n, 0);
        const int nv = SYM_V / nranks, v0 = rank * nv;
        const int nblocks = (n + SYM_RB - 1) / SYM_RB;
        int owned = 0;
        for (int B = 0; B < nblocks; ++B) {
            if (sym_vrank(B) < v0 || sym_vrank(B) >= v0 + nv) continue;
            owned += 1;
            // the kernel's walk: directed pairs inside the block's own tiles, one evaluation per pair above it
            const long long rows = (B + 1) * SYM_RB <= n ? SYM_RB : n - B * SYM_RB;
            evals += rows * rows + rows * static_cast<long long>(n - B * SYM_RB - rows);
        }
        out[2] = v0;
        out[3] = nv;
        out[4] = owned;
        out[5] = cm.nsplit;
        out[6] = cm.nbig;
        out[7] = 1;
    } else {
        evals = static_cast<long long>(hi - lo) * n;   // directed: every owned row against all n columns
        int nsplit = 1, chunk = n;
        if (fast) choose_split(n, 0, &nsplit, &chunk);
        out[5] = nsplit;
        out[6] = nsplit;
    }
    if (pair_evaluations) *pair_evaluations = evals;
    return NBODY_GPU_OK;
}

int nbody_gpu_sym_work_items(int n, int nranks, int rank, int *blocks, int *chunks, int *groups_out, int *guards, int cap,
                             int *count, int *chunk_start_out, int chunk_cap, int *nchunks, int *ring)
{
    if (n < 1 || nranks < 1 || rank < 0 || rank >= nranks || cap < 0 || !count) return fail(NBODY_GPU_EINVAL, "bad argument");
    if (SYM_V % nranks != 0) return fail(NBODY_GPU_EINVAL, "the symmetric kernel shards over 1, 2, 4 or 8 GPUs");
    const ChunkMap cm = choose_split_sym(n, 0);
    if (!sym_ring_fits(cm))
        return fail(NBODY_GPU_EINVAL, "n = %d: a row block would have %d items, the slice ring holds %d (the library uses the "
                    "directed kernel for such n)", n, sym_max_group(cm), SYM_RING);
    const int nv = SYM_V / nranks;
    std::vector<int4> items;
    std::vector<int2> groups;
    sym_work_items(n, cm, rank * nv, nv, &items, &groups);
    *count = static_cast<int>(items.size());
    for (int k = 0; k < *count && k < cap; ++k) {
        if (blocks) blocks[k] = items[k].x;
        if (chunks) chunks[k] = items[k].y;
        if (groups_out) groups_out[k] = items[k].z;
        if (guards) guards[k] = items[k].w;
    }
    if (ring) *ring = SYM_RING;
    if (nchunks) *nchunks = cm.nsplit;
    if (chunk_start_out)
        for (int s = 0; s <= cm.nsplit && s < chunk_cap; ++s) {
            long long at = s < cm.nsplit ? chunk_start(cm, s) : n;
            chunk_start_out[s] = static_cast<int>(at < n ? at : n);
        }
    return NBODY_GPU_OK;
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

int nbody_gpu_pin_host(void *ptr, size_t bytes)
{
    if (!ptr || !bytes) return fail(NBODY_GPU_EINVAL, "NULL or empty range");
    cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        return NBODY_GPU_OK;
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(NBODY_GPU_ECUDA, "cudaHostRegister(%p, %zu) failed: %s", ptr, bytes, cudaGetErrorString(e));
    }
    return NBODY_GPU_OK;
}

int nbody_gpu_unpin_host(void *ptr)
{
    if (!ptr) return fail(NBODY_GPU_EINVAL, "NULL pointer");
    cudaError_t e = cudaHostUnregister(ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(NBODY_GPU_ECUDA, "cudaHostUnregister(%p) failed: %s", ptr, cudaGetErrorString(e));
    }
    return NBODY_GPU_OK;
}

void nbody_gpu_destroy(nbody_gpu *h)
{
    if (!h) return;
    for (Shard &s : h->shards) {   // nothing of ours is still running when the memory goes away
        if (s.stream) {
            cudaSetDevice(s.device);
            cudaStreamSynchronize(s.stream);
        }
    }
    for (Shard &s : h->shards) free_shard(s);
    cudaGetLastError();
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
    if (cfg->sort_interval < 0) return fail(NBODY_GPU_EINVAL, "sort_interval must be >= 0");
    h->nshards_total = P;
    h->sharded = P > 1;
    h->rows_per_shard = (h->n + P - 1) / P;
    h->npad = h->rows_per_shard * P;
    const bool fast = cfg->mode == NBODY_GPU_MODE_FAST;
    h->symmetric = fast && (SYM_V % P == 0) && (cfg->symmetric > 0 || (cfg->symmetric == 0 && h->n >= 8192));
    h->nblocks = (h->n + SYM_RB - 1) / SYM_RB;
    const size_t bytes = sizeof(double2) * static_cast<size_t>(h->npad);
    if (h->symmetric) {
        h->cm = choose_split_sym(h->n, cfg->jsplit);
        if (!sym_ring_fits(h->cm)) {   // a row block's items would wrap round the slice ring onto themselves
            if (cfg->symmetric > 0)
                return fail(NBODY_GPU_EINVAL, "the symmetric kernel cannot split n = %d into %d chunks per row block: its ring "
                            "holds %d row-sum slices (use a smaller jsplit, or symmetric = 0 for the directed kernel)",
                            h->n, sym_max_group(h->cm), SYM_RING);
            h->symmetric = false;
        }
    }
    if (h->symmetric) {
        // the symmetric kernel's scratch: reaction sums per owned virtual rank, two row-sum vectors, a fixed ring of
        // slices -- O(n); the check stays for very large n
        const size_t slots = static_cast<size_t>(SYM_V / P);
        size_t free_b = 0, total_b = 0;
        if (multi_dev) NB_CUDA(cudaSetDevice(0));
        else if (cfg->device >= 0) NB_CUDA(cudaSetDevice(cfg->device));
        NB_CUDA(cudaMemGetInfo(&free_b, &total_b));
        const double need = static_cast<double>(bytes) * (static_cast<double>(slots) + 2.0 + 24.0) +
                            static_cast<double>(sizeof(double2)) * SYM_RING * SYM_RB;
        if (need > 0.92 * static_cast<double>(free_b)) {
            if (cfg->symmetric > 0)
                return fail(NBODY_GPU_ENOMEM, "the symmetric kernel needs %.1f GB of partial sums for n = %d on %d GPU(s), %.1f GB are free",
                            need / 1e9, h->n, P, free_b / 1e9);
            h->symmetric = false;
        }
    }
    h->sorted = h->symmetric && (cfg->sort > 0 || (cfg->sort == 0 && h->n >= 16384));
    h->sort_interval = cfg->sort_interval > 0 ? cfg->sort_interval : 256;
    h->fused_boxes = h->sorted && (P == 1 || h->rows_per_shard % TILE_J == 0);
    if (h->symmetric) h->nsplit = h->cm.nsplit;
    else if (fast) choose_split(h->n, cfg->jsplit, &h->nsplit, &h->chunk);
    else { h->nsplit = 1; h->chunk = h->n; }
    h->small = !fast && !h->sharded && h->n <= SMALL_N_MAX && cfg->use_graph >= 0;
    if (cfg->use_graph >= 0) {
        // small worlds are launch-bound: put more steps behind one graph launch
        h->graph_steps = (h->n <= 2048) ? 32 : (h->n <= 16384 ? 8 : 2);
    }
    h->lay = arena_layout(h->npad, h->symmetric, h->sorted, h->n);

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
        NB_CUDA(cudaEventCreateWithFlags(&s.ev_copy, cudaEventDisableTiming));
        NB_CUDA(cudaMalloc(&s.arena, h->lay.bytes));
        NB_CUDA(cudaMemsetAsync(s.arena, 0, h->lay.bytes, s.stream));
        point_into_arena(h->lay, s.arena, &s.pos[0], &s.pos[1], &s.vel, &s.frc, &s.vbuf, &s.boxes, &s.flags);
        s.epoch = reinterpret_cast<unsigned long long *>(s.arena + h->lay.epoch);
        if (!h->symmetric) s.vbuf = nullptr;
        if (!h->sorted) s.boxes = nullptr;
        NB_CUDA(cudaMalloc(&s.mr, bytes));
        NB_CUDA(cudaMemsetAsync(s.mr, 0, bytes, s.stream));
        NB_CUDA(cudaMalloc(&s.plan, sizeof(double) * 6 * static_cast<size_t>(h->npad)));
        if (fast) {
            if (!h->symmetric) {   // directed kernel: one partial-sum vector per j-chunk
                NB_CUDA(cudaMalloc(&s.part, bytes * h->nsplit));
                NB_CUDA(cudaMemsetAsync(s.part, 0, bytes * h->nsplit, s.stream));
            }
            if (h->symmetric) {
                const size_t ring = sizeof(double2) * static_cast<size_t>(SYM_RING) * SYM_RB;
                const size_t gwords = sizeof(int) * 2 * static_cast<size_t>(h->nblocks);
                NB_CUDA(cudaMalloc(&s.slices, ring));
                NB_CUDA(cudaMemsetAsync(s.slices, 0, ring, s.stream));
                NB_CUDA(cudaMalloc(&s.rowsum, bytes * 2));
                NB_CUDA(cudaMemsetAsync(s.rowsum, 0, bytes * 2, s.stream));   // rows without long (short) chunks stay 0
                NB_CUDA(cudaMalloc(&s.gcnt, gwords));
                NB_CUDA(cudaMemsetAsync(s.gcnt, 0, gwords, s.stream));
                NB_CUDA(cudaMalloc(&s.gdone, gwords));
                NB_CUDA(cudaMemsetAsync(s.gdone, 0, gwords, s.stream));
                s.nv = SYM_V / P;
                s.v0 = s.part_index * s.nv;
                s.nseg = seg_count(h->cm);
                NB_CUDA(cudaMalloc(&s.vacc, bytes * s.nv));
                NB_CUDA(cudaMemsetAsync(s.vacc, 0, bytes * s.nv, s.stream));
                NB_CUDA(cudaMalloc(&s.turn, sizeof(int) * static_cast<size_t>(s.nv) * s.nseg));
                NB_CUDA(cudaMemsetAsync(s.turn, 0, sizeof(int) * static_cast<size_t>(s.nv) * s.nseg, s.stream));
                NB_CUDA(cudaMalloc(&s.sym_err, sizeof(int)));
                NB_CUDA(cudaMemsetAsync(s.sym_err, 0, sizeof(int), s.stream));
                int items_v0 = s.v0, items_nv = s.nv;
#ifdef NBODY_DEV
                if (const char *e = getenv("NBODY_DEV_VRANK")) {
                    items_v0 = atoi(e) % SYM_V;
                    items_nv = 1;
                }
#endif
                const int rc_items = upload_sym_items(s, h->n, h->cm, items_v0, items_nv);
                if (rc_items) return rc_items;
            }
            if (h->sorted) {
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
    int rc = NBODY_GPU_OK;
    if (h->sharded && h->symmetric && cfg->p2p >= 0) {
        rc = setup_p2p(h);
        if (!rc && !h->p2p && cfg->p2p > 0) rc = fail(NBODY_GPU_ECUDA, "peer-memory exchange requested but peer access is unavailable");
        if (rc) return rc;
    }

    h->cur = 0;
    rc = upload_state(h, 0, x, y, xv, yv, mass, radius);
    if (!rc && h->sorted) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            NB_CUDA(cudaMemcpyAsync(s.mr0, s.mr, bytes, cudaMemcpyDeviceToDevice, s.stream));
            launch_iota(s.orig, h->npad, s.stream);
        }
        rc = resort(h, false);
    }
    if (!rc) rc = sync_all(h);
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
    int rc = check_usable(h);
    if (rc) return rc;
    if (!x || !y || !xv || !yv) return fail(NBODY_GPU_EINVAL, "NULL argument");
    rc = upload_state(h, h->cur, x, y, xv, yv, nullptr, nullptr);
    if (rc) return poison(h, rc);
    if (h->sorted) {   // the upload is in original order: back to the identity permutation, then sort again
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            NB_CUDA(cudaMemcpyAsync(s.mr, s.mr0, sizeof(double2) * static_cast<size_t>(h->npad), cudaMemcpyDeviceToDevice, s.stream));
            launch_iota(s.orig, h->npad, s.stream);
            h->launches[2] += 1;
        }
        rc = resort(h, false);
        if (rc) return poison(h, rc);
    }
    return NBODY_GPU_OK;   // the host buffers are free again; the sort may still be queued
}

int nbody_gpu_step(nbody_gpu *h, int steps)
{
    int rc = check_usable(h);
    if (rc) return rc;
    return poison(h, do_steps(h, steps));
}

int nbody_gpu_step_timed(nbody_gpu *h, int steps, float *ms)
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (!ms) return fail(NBODY_GPU_EINVAL, "NULL argument");
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventRecord(s.ev0, s.stream));
    }
    rc = do_steps(h, steps);
    if (rc) return poison(h, rc);
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
    rc = sync_all(h);   // also reads the peer-exchange error word
    if (rc) return poison(h, rc);
    *ms = worst;
    return NBODY_GPU_OK;
}

int nbody_gpu_step_bench(nbody_gpu *h, int steps, int flush_l2, float *ms)
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (!ms || steps < 0) return fail(NBODY_GPU_EINVAL, "bad argument");
    const size_t flush_bytes = 192u << 20;   // larger than the 126 MB L2
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        if (flush_l2 && !s.flush) NB_CUDA(cudaMalloc(&s.flush, 256u << 20));   // first call only
    }
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventRecord(s.ev0, s.stream));
    }
    // the production path (graph launches / the one-launch loop), one launch unit at a time with the flush between
    const int unit = h->small ? (steps > 0 ? steps : 1) : ((h->graph_steps > 0 && (!h->sharded || h->p2p)) ? h->graph_steps : 1);
    int tick = 0;
    for (int left = steps; left > 0;) {
        const int k = left < unit ? left : unit;
        if (flush_l2) {
            for (Shard &s : h->shards) {
                NB_CUDA(cudaSetDevice(s.device));
                NB_CUDA(cudaMemsetAsync(s.flush, tick & 0xff, flush_bytes, s.stream));
            }
            ++tick;
        }
        rc = do_steps(h, k);
        if (rc) return poison(h, rc);
        left -= k;
    }
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
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    *ms = worst;
    return NBODY_GPU_OK;
}

namespace {
struct EventMarks : StepMarks {
    nbody_gpu *h;
    int step = 0;   // index inside the current pass, < PROF_CHUNK
    explicit EventMarks(nbody_gpu *h_) : h(h_) {}
    cudaEvent_t &at(size_t k, int st, int phase) { return h->shards[k].prof_ev[static_cast<size_t>(st) * 6 + phase]; }
    int mark(size_t k, int phase) override
    {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
        NB_CUDA(cudaEventRecord(at(k, step, phase), h->shards[k].stream));
        return NBODY_GPU_OK;
    }
};

int profiled_pass(nbody_gpu *h, int steps, int flush_l2, float acc_out[4])
{
    const size_t flush_bytes = 256u << 20;
    const size_t nsh = h->shards.size();
    EventMarks marks(h);
    for (size_t k = 0; k < nsh; ++k) {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
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
        int rc = run_one_step(h, &marks);
        if (rc) return rc;
    }
    for (size_t k = 0; k < nsh; ++k) {
        NB_CUDA(cudaSetDevice(h->shards[k].device));
        NB_CUDA(cudaEventRecord(h->shards[k].ev1, h->shards[k].stream));
    }
    float worst[4] = {0.f, 0.f, 0.f, 0.f};
    for (size_t k = 0; k < nsh; ++k) {
        Shard &s = h->shards[k];
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaEventSynchronize(s.ev1));
        float acc[4] = {0.f, 0.f, 0.f, 0.f}, t = 0.f;
        NB_CUDA(cudaEventElapsedTime(&acc[0], s.ev0, s.ev1));
        for (int st = 0; st < steps; ++st) {
            // phases: 0-1 force kernel | 1-2 symmetric fold | 2-3 exchange A | 3-4 integrate | 4-5 exchange B
            static const int bucket[5] = {1, 2, 3, 2, 3};
            float ph_ms[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            for (int ph = 0; ph < 5; ++ph) {
                NB_CUDA(cudaEventElapsedTime(&t, marks.at(k, st, ph), marks.at(k, st, ph + 1)));
                ph_ms[ph] = t;
            }
            for (int ph = 0; ph < 5; ++ph) acc[bucket[ph]] += ph_ms[ph];
#ifdef NBODY_DEV
            if (getenv("NBODY_PROFILE_SPLIT"))   // development knob: per-step phase times of this shard
                fprintf(stderr, "[nbody_gpu] shard %d step %d: force %.3f fold %.3f exchA %.3f integrate %.3f exchB %.3f ms\n",
                        s.part_index, st, ph_ms[0], ph_ms[1], ph_ms[2], ph_ms[3], ph_ms[4]);
#endif
        }
        if (!h->sharded) acc[3] = 0.f;
        for (int q = 0; q < 4; ++q)
            if (acc[q] > worst[q]) worst[q] = acc[q];
    }
    for (int q = 0; q < 4; ++q) acc_out[q] += worst[q];
    return NBODY_GPU_OK;
}
}  // namespace

int nbody_gpu_step_profiled(nbody_gpu *h, int steps, int flush_l2, float out_ms[4])
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (!out_ms || steps < 0) return fail(NBODY_GPU_EINVAL, "bad argument");
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    for (int q = 0; q < 4; ++q) out_ms[q] = 0.f;
    for (Shard &s : h->shards) {   // first call only: the event pool and the flush scratch stay with the handle
        NB_CUDA(cudaSetDevice(s.device));
        if (s.prof_ev.empty()) {
            s.prof_ev.assign(static_cast<size_t>(PROF_CHUNK) * 6, nullptr);
            for (cudaEvent_t &e : s.prof_ev) NB_CUDA(cudaEventCreate(&e));
        }
        if (flush_l2 && !s.flush) NB_CUDA(cudaMalloc(&s.flush, 256u << 20));
    }
    // bounded passes: a fixed pool of events however long the run (nbody-par --measure-comm passes the whole run)
    for (int done = 0; done < steps;) {
        const int pass = steps - done < PROF_CHUNK ? steps - done : PROF_CHUNK;
        rc = profiled_pass(h, pass, flush_l2, out_ms);
        if (rc) return poison(h, rc);
        done += pass;
    }
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    return NBODY_GPU_OK;
}

int nbody_gpu_sync(nbody_gpu *h)
{
    int rc = check_usable(h);
    if (rc) return rc;
    return poison(h, sync_all(h));
}

// Up to three double2 device arrays of shard 0 (slot order) -> the caller's planar arrays (body order): one
// gather kernel writes planar staging on the device, then one D2H copy per requested array, straight into the
// caller's memory (full PCIe speed if it is registered, driver-staged if pageable).
static int download(nbody_gpu *h, const double2 *const src[3], double *const out[6], int first, int count)
{
    Shard &s = h->shards[0];
    NB_CUDA(cudaSetDevice(s.device));
    const size_t np = static_cast<size_t>(h->npad);
    GatherArgs g;
    bool any = false;
    for (int k = 0; k < 3; ++k) {
        const bool want = out[2 * k] || out[2 * k + 1];
        g.src[k] = want ? src[k] : nullptr;
        g.a[k] = s.plan + (2 * k) * np;
        g.b[k] = s.plan + (2 * k + 1) * np;
        any |= want;
    }
    if (!any) return NBODY_GPU_OK;
    g.orig = h->sorted ? s.orig : nullptr;
    g.n = h->n;
    g.first = first;
    g.count = count;
    launch_gather_planar(g, s.stream);
    h->launches[2] += 1;
    for (int k = 0; k < 6; ++k)
        if (out[k] && count > 0)
            NB_CUDA(cudaMemcpyAsync(out[k], s.plan + k * np, sizeof(double) * count, cudaMemcpyDeviceToHost, s.stream));
    return NBODY_GPU_OK;
}

int nbody_gpu_forces(nbody_gpu *h, double *xf, double *yf)
{
    int rc = check_usable(h);
    if (rc) return rc;
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        if (!s.fscratch) {
            NB_CUDA(cudaMalloc(&s.fscratch, sizeof(double2) * static_cast<size_t>(h->npad)));
            NB_CUDA(cudaMemsetAsync(s.fscratch, 0, sizeof(double2) * static_cast<size_t>(h->npad), s.stream));
        }
        rc = enqueue_forces(h, s, h->cur, s.fscratch);
        if (rc) return poison(h, rc);
        if (h->symmetric) enqueue_sym_reduce(h, s, false);
        NB_CUDA(cudaGetLastError());
    }
    if (h->symmetric && h->sharded) rc = allgather(h, AG_VBUF);
    if (rc) return poison(h, rc);
    if (h->cfg.mode == NBODY_GPU_MODE_FAST) {
        for (Shard &s : h->shards) {
            NB_CUDA(cudaSetDevice(s.device));
            IntegrateArgs a = integrate_args(h, s, h->cur);
            a.frc = s.fscratch;   // the last step's forces in frc stay what print() shows
            launch_reduce_fast(a, s.stream);   // partial sums -> fscratch
            h->launches[2] += 1;
        }
    }
    if (state_sharded(h)) rc = allgather(h, AG_SCRATCH);
    if (rc) return poison(h, rc);
    const double2 *src[3] = {nullptr, h->shards[0].fscratch, nullptr};
    double *out[6] = {nullptr, nullptr, xf, yf, nullptr, nullptr};
    rc = download(h, src, out, 0, h->n);
    if (!rc) rc = sync_all(h);
    return poison(h, rc);
}

int nbody_gpu_fetch_range(nbody_gpu *h, int first, int count, double *x, double *y, double *xf, double *yf, double *xv,
                          double *yv)
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (first < 0 || count < 0 || first > h->n || count > h->n - first)
        return fail(NBODY_GPU_EINVAL, "bodies [%d, %d) are not inside [0, %d)", first, first + count, h->n);
    // forces and velocities live sharded during the run; complete them like the reference's
    // whole-struct MPI_Allgatherv leaves them (nbody-par.c:933-936)
    rc = complete_state(h, xv || yv, xf || yf);
    if (rc) return poison(h, rc);
    Shard &s0 = h->shards[0];
    const double2 *src[3] = {s0.pos[h->cur], s0.frc, s0.vel};
    double *out[6] = {x, y, xf, yf, xv, yv};
    rc = download(h, src, out, first, count);
    if (!rc) rc = sync_all(h);
    if (rc) return poison(h, rc);
    for (Shard &s : h->shards) {
        NB_CUDA(cudaSetDevice(s.device));
        NB_CUDA(cudaGetLastError());
    }
    return NBODY_GPU_OK;
}

int nbody_gpu_fetch(nbody_gpu *h, double *x, double *y, double *xf, double *yf, double *xv, double *yv)
{
    if (!h) return fail(NBODY_GPU_EINVAL, "handle is NULL");
    return nbody_gpu_fetch_range(h, 0, h->n, x, y, xf, yf, xv, yv);
}

int nbody_gpu_describe(const nbody_gpu *h, char *buf, int len)
{
    if (!h) return 0;
    const bool fast = h->cfg.mode == NBODY_GPU_MODE_FAST;
    const char *kernel = !fast ? (h->small ? "small_world_kernel" : "forces_exact_wide_kernel")
                         : h->symmetric ? (h->n >= SYM_WIDE_N ? "forces_sym_kernel<64,8>" : "forces_sym_kernel<128,4>")
                                        : "forces_fast_kernel<2,4,1,true>";
    // FP64-pipe instructions per directed pair-interaction (SASS counts): the symmetric
    // kernel issues 17 per unordered pair = 2 interactions (15 in far-field tiles: see nbody_gpu_far_fraction);
    // exact mode is not a fixed count (divisions, branches)
    const double ipp = !fast ? 0.0 : h->symmetric ? 8.5 : 16.0;   // per directed interaction: 17 per unordered near-field pair (15 far field)
    const int graph = h->small ? 0 : ((!h->sharded || h->p2p) ? h->graph_steps : 0);
    char tmp[768];
    const int need = snprintf(tmp, sizeof tmp,
                              "{\"mode\": \"%s\", \"kernel\": \"%s\", \"symmetric\": %s, \"nsplit\": %d, \"chunk\": %d, "
                              "\"short_chunks\": %d, \"row_blocks\": %d, \"shards\": %d, \"graph_steps\": %d, \"single_launch_loop\": %s, "
                              "\"exchange\": \"%s\", \"fp64_instr_per_pair\": %.1f, "
                              "\"cell_order\": %s, \"fused_tile_boxes\": %s, \"sort_interval\": %d, \"sorts\": %lld, "
                              "\"partial_sum_bytes\": %lld}",
                              fast ? "fast" : "exact", kernel, h->symmetric ? "true" : "false", h->nsplit,
                              h->symmetric ? h->cm.chunk : h->chunk, h->symmetric ? h->cm.nsplit - h->cm.nbig : 0,
                              h->symmetric ? h->nblocks : 0, h->nshards_total, graph, h->small ? "true" : "false",
                              !h->sharded ? "none" : h->p2p ? "peer-memory stores + barrier kernels" : "ncclAllGather", ipp,
                              h->sorted ? "true" : "false", h->fused_boxes ? "true" : "false",
                              h->sorted ? h->sort_interval : 0, h->sorts,
                              // device memory of the fast kernels' partial sums, per GPU: row sums per j-chunk, and the
                              // symmetric kernel's reaction sums per owned virtual rank
                              !fast ? 0LL
                              : h->symmetric ? static_cast<long long>(sizeof(double2)) *
                                                   (static_cast<long long>(h->npad) * (h->shards[0].nv + 2) + static_cast<long long>(SYM_RING) * SYM_RB)
                                             : static_cast<long long>(sizeof(double2)) * h->npad * h->nsplit);
    if (buf && len > 0) {
        strncpy(buf, tmp, static_cast<size_t>(len) - 1);
        buf[len - 1] = 0;
    }
    return need + 1;
}

int nbody_gpu_far_fraction(nbody_gpu *h, double *fraction)
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (!fraction) return fail(NBODY_GPU_EINVAL, "NULL argument");
    *fraction = 0.0;
    if (!h->sorted) return NBODY_GPU_OK;
    rc = sync_all(h);
    if (rc) return poison(h, rc);
    Shard &s = h->shards[0];
    NB_CUDA(cudaSetDevice(s.device));
    const int ntiles = (h->n + TILE_J - 1) / TILE_J;
    std::vector<double> bx(static_cast<size_t>(ntiles) * SYM_BOX_DOUBLES);
    // the boxes always describe pos[cur] (re-sort, integrate_kernel): read them as the force kernel would
    if (!h->fused_boxes) launch_tile_boxes(s.pos[h->cur], s.mr, h->n, s.boxes, s.stream);
    NB_CUDA(cudaMemcpyAsync(bx.data(), s.boxes, bx.size() * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    NB_CUDA(cudaStreamSynchronize(s.stream));
    // the same walk and the same test as forces_sym_kernel: tiles above each row block
    const int tiles_per_block = SYM_RB / TILE_J;
    double far = 0.0, all = 0.0;
    for (int B = 0; B < h->nblocks; ++B) {
        double b[5] = {0, 0, 0, 0, 0};
        for (int t = 0; t < tiles_per_block && B * tiles_per_block + t < ntiles; ++t) {
            const double *q = &bx[static_cast<size_t>(B * tiles_per_block + t) * SYM_BOX_DOUBLES];
            if (t == 0) { for (int k = 0; k < 5; ++k) b[k] = q[k]; continue; }
            b[0] = q[0] < b[0] ? q[0] : b[0]; b[1] = q[1] < b[1] ? q[1] : b[1];
            b[2] = q[2] > b[2] ? q[2] : b[2]; b[3] = q[3] > b[3] ? q[3] : b[3];
            b[4] = q[4] > b[4] ? q[4] : b[4];
        }
        const double rows = (B + 1) * SYM_RB <= h->n ? SYM_RB : h->n - B * SYM_RB;
        for (int T = (B + 1) * tiles_per_block; T < ntiles; ++T) {
            const double *q = &bx[static_cast<size_t>(T) * SYM_BOX_DOUBLES];
            double gx = q[0] - b[2] > b[0] - q[2] ? q[0] - b[2] : b[0] - q[2];
            double gy = q[1] - b[3] > b[1] - q[3] ? q[1] - b[3] : b[1] - q[3];
            if (gx < 0.0) gx = 0.0;
            if (gy < 0.0) gy = 0.0;
            const double md = b[4] + q[4];
            const double cnt = rows * ((T + 1) * TILE_J <= h->n ? TILE_J : h->n - T * TILE_J);
            all += cnt;
            if (gx * gx + gy * gy >= md * md) far += cnt;
        }
    }
    if (all > 0.0) *fraction = far / all;
    return NBODY_GPU_OK;
}

int nbody_gpu_render(nbody_gpu *h, unsigned char *rgb)
{
    int rc = check_usable(h);
    if (rc) return rc;
    if (!rgb) return fail(NBODY_GPU_EINVAL, "NULL argument");
    Shard &s = h->shards[0];   // every device holds all current positions
    NB_CUDA(cudaSetDevice(s.device));
    const size_t npix = static_cast<size_t>(h->cfg.xdim) * h->cfg.ydim;
    if (!s.render_owner) NB_CUDA(cudaMalloc(&s.render_owner, npix * sizeof(int)));   // first call only
    if (!s.render_img) NB_CUDA(cudaMalloc(&s.render_img, npix * 3));
    const double2 *rpos = s.pos[h->cur], *rmr = s.mr;
    if (h->sorted) {   // the lowest ORIGINAL index wins a pixel (nbody-seq.c:336-347): paint from body order
        launch_unsort(s.pos[h->cur], s.orig, s.gat[0], h->n, s.stream);
        rpos = s.gat[0];
        rmr = s.mr0;
    }
    launch_render(rpos, rmr, h->n, h->cfg.xdim, h->cfg.ydim, s.render_owner, s.render_img, s.stream);
    h->launches[2] += 3;
    cudaError_t e = cudaMemcpyAsync(rgb, s.render_img, npix * 3, cudaMemcpyDeviceToHost, s.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
    if (e != cudaSuccess) return poison(h, fail(NBODY_GPU_ECUDA, "render failed: %s", cudaGetErrorString(e)));
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

// ---------------------------------------------------------------------------------------------------------
// nbody_gpu_forces_static: the directed fast-mode force pass for callers that must not allocate.
//
// Who that is: a host program whose device memory was carved up once, at start-up, by somebody else's pool (a
// simulation framework, an inference server sharing the GPU) so that the driver's allocator has nothing left or
// is off limits to plug-ins; and a per-frame or per-request loop in which a cudaMalloc or cudaFree -- both may wait
// for the whole device -- is a stall other streams would feel.  Such a caller already owns its position and force
// buffers; what it cannot supply is the kernel's scratch: the directed kernel leaves one partial force vector per
// j-chunk (forces_fast.cuh) that a second kernel adds up in chunk order.  Here that scratch is g_static_part,
// a fixed array inside the library image.  Every GPU whose context loads the library carries its
// NBODY_GPU_STATIC_SCRATCH_BYTES (16 MiB) until the process ends, used or not -- hence the modest limit of
// NBODY_GPU_STATIC_MAX_BODIES bodies.
//
// g_static_part, named in HOST code, is only the host-side shadow of the symbol, never a device address; the
// kernels are handed what cudaGetSymbolAddress returns, asked once per device with that device current.
// ---------------------------------------------------------------------------------------------------------
constexpr size_t STATIC_PART_ELEMS = NBODY_GPU_STATIC_SCRATCH_BYTES / sizeof(double2);
static __device__ __align__(256) double2 g_static_part[STATIC_PART_ELEMS];

constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mutex;
static double2 *g_static_part_addr[STATIC_MAX_DEVICES] = {};   // device address of g_static_part on each device

// Device address of the scratch array on the CURRENT device `dev` (looked up on the first call per device).
static int static_part_address(int dev, double2 **out)
{
    if (dev < 0 || dev >= STATIC_MAX_DEVICES)
        return fail(NBODY_GPU_EINVAL, "device ordinal %d outside 0..%d", dev, STATIC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_part_addr[dev]) {
        void *p = nullptr;
        NB_CUDA(cudaGetSymbolAddress(&p, g_static_part));
        if (!p) return fail(NBODY_GPU_ECUDA, "cudaGetSymbolAddress returned no address for the static scratch");
        g_static_part_addr[dev] = static_cast<double2 *>(p);
    }
    *out = g_static_part_addr[dev];
    return NBODY_GPU_OK;
}

static int check_device_array(const void *p, int dev, const char *what)
{
    if (!p) return fail(NBODY_GPU_EINVAL, "%s is NULL", what);
    if (reinterpret_cast<uintptr_t>(p) % sizeof(double2))
        return fail(NBODY_GPU_EINVAL, "%s is not 16-byte aligned", what);
    cudaPointerAttributes at;
    NB_CUDA(cudaPointerGetAttributes(&at, p));
    if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
        return fail(NBODY_GPU_EINVAL, "%s is not device memory (this entry point takes DEVICE pointers)", what);
    if (at.type == cudaMemoryTypeDevice && at.device != dev)
        return fail(NBODY_GPU_EINVAL, "%s lives on device %d, the current device is %d", what, at.device, dev);
    return NBODY_GPU_OK;
}

extern "C" int nbody_gpu_forces_static(int n, const double *pos_xy, const double *mass_radius, double *force_xy,
                                       void *stream)
{
    if (n < 1) return fail(NBODY_GPU_EINVAL, "n = %d: need at least one body", n);
    if (n > NBODY_GPU_STATIC_MAX_BODIES)
        return fail(NBODY_GPU_EINVAL,
                    "n = %d exceeds NBODY_GPU_STATIC_MAX_BODIES = %d: the built-in scratch array of %zu bytes cannot "
                    "hold the partial sums; use a handle (nbody_gpu_init + nbody_gpu_forces), which allocates",
                    n, NBODY_GPU_STATIC_MAX_BODIES, static_cast<size_t>(NBODY_GPU_STATIC_SCRATCH_BYTES));
    int nsplit = 1, chunk = n;
    choose_split(n, 0, &nsplit, &chunk);   // the library default of a handle: same chunks, same summation order
    const int npad = (n + 15) / 16 * 16;
    if (static_cast<size_t>(nsplit) * npad > STATIC_PART_ELEMS)
        return fail(NBODY_GPU_EINVAL, "n = %d needs %d partial vectors of %d: more than the built-in scratch holds",
                    n, nsplit, npad);
    int dev = -1;
    NB_CUDA(cudaGetDevice(&dev));
    int rc = check_device_array(pos_xy, dev, "pos_xy");
    if (!rc) rc = check_device_array(mass_radius, dev, "mass_radius");
    if (!rc) rc = check_device_array(force_xy, dev, "force_xy");
    if (rc) return rc;
    double2 *part = nullptr;
    rc = static_part_address(dev, &part);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    // Nothing in the array has to start at zero: every (chunk, row < n) element the fold reads is stored by the
    // force kernel of this same call first (a CTA of an empty chunk stores zeros), so the leftovers of earlier
    // calls are never read.
    ForceArgs fa;
    fa.pos = reinterpret_cast<const double2 *>(pos_xy);
    fa.mr = reinterpret_cast<const double2 *>(mass_radius);
    fa.out = part;
    fa.n = n;
    fa.npad = npad;
    fa.lo = 0;
    fa.hi = n;
    fa.chunk = chunk;
    launch_forces_fast(fa, nsplit, st);
    NB_CUDA(cudaGetLastError());

    IntegrateArgs ia;
    memset(&ia, 0, sizeof ia);
    ia.part = part;
    ia.vbuf = nullptr;
    ia.frc = reinterpret_cast<double2 *>(force_xy);
    ia.nsplit = nsplit;
    ia.npad = npad;
    ia.n = n;
    ia.lo = 0;
    ia.hi = n;
    launch_reduce_fast(ia, st);   // partial vectors, chunk ascending -> force_xy
    NB_CUDA(cudaGetLastError());
    return NBODY_GPU_OK;
}
