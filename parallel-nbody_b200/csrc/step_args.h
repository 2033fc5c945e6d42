// step_args.h -- kernel argument blocks and host-side launchers of the N-body step.
// No device code: included by nbody_gpu.cu (the C ABI) and by step_kernels.cuh.
#pragma once
#include <cuda_runtime.h>

namespace nbody {

constexpr int MODE_EXACT = 0;
constexpr int MODE_FAST = 1;

constexpr int FORCE_THREADS = 128;  // bodies i per CTA (one per thread)
constexpr int TILE_J = 256;         // bodies j per shared-memory stage
constexpr int STAGES = 2;

struct ForceArgs {
    const double2 *pos;   // current positions, all n bodies
    const double2 *mr;    // mass, radius
    double2 *out;         // EXACT: frc (indexed by i); FAST: part[s * npad + i]
    int n;                // bodies (loop bound for j)
    int npad;             // leading dimension of part
    int lo, hi;           // rows owned by this device: i in [lo, hi)
    int chunk;            // FAST: j-chunk length; chunk s covers [s*chunk, min(n,(s+1)*chunk))
};

constexpr int SYM_RB = 512;         // rows per block of the symmetric (evaluate-each-pair-once) kernel
constexpr int SYM_WIDE_N = 49152;   // from this many bodies on: 64 threads x 8 bodies per CTA instead of 128 x 4 (profiles/r2_shape_sweep.log)
constexpr int SYM_BOX_DOUBLES = 8;  // per-tile bounding box: xmin ymin xmax ymax rmax + padding (one 64-byte bulk copy)

// j-chunks of the symmetric kernel: `nbig` chunks of `chunk` bodies cover [0, jsmall0), then chunks of `csmall`
// bodies (one tile) cover [jsmall0, n).  A CTA is one (row block, chunk) item and the hardware dispatches CTAs in
// chunk order, so the short items come last: when the long items run out, every SM slot that falls idle picks
// up a quarter-size item instead of waiting for the slowest long one (the tail of a launch costs about half an
// item per slot: 3.4 % of the step on an 8-way shard of N = 262144 with uniform chunks).  The map is a function
// of n only, like everything else that fixes the summation order.
struct ChunkMap {
    int chunk;     // long chunk length, a multiple of SYM_RB
    int nbig;      // number of long chunks
    int jsmall0;   // nbig * chunk: where the short chunks start (>= n if there are none)
    int csmall;    // short chunk length (TILE_J)
    int nsplit;    // chunks in total
};
__host__ __device__ inline int chunk_start(const ChunkMap &m, int s)
{
    return s < m.nbig ? s * m.chunk : m.jsmall0 + (s - m.nbig) * m.csmall;
}
__host__ __device__ inline int chunk_len(const ChunkMap &m, int s) { return s < m.nbig ? m.chunk : m.csmall; }
__host__ __device__ inline int chunk_of(const ChunkMap &m, int j)
{
    return j < m.jsmall0 ? j / m.chunk : m.nbig + (j - m.jsmall0) / m.csmall;
}

// Reaction sums are handed from CTA to CTA in segments of SYM_SEG bodies j: a long chunk holds segs_per_chunk()
// of them, a short chunk is one.
constexpr int SYM_SEG = 1024;
__host__ __device__ inline int segs_per_chunk(const ChunkMap &m) { return (m.chunk + SYM_SEG - 1) / SYM_SEG; }
__host__ __device__ inline int seg_count(const ChunkMap &m) { return m.nbig * segs_per_chunk(m) + (m.nsplit - m.nbig); }

// Row sums: the CTAs of one row block (one "group": its long-chunk items, or its short-chunk items) leave their
// SYM_RB row sums in a ring of SYM_RING slices (slot = CTA index % SYM_RING) and count themselves in; the group's
// last CTA in dispatch order adds the slices in chunk order and writes the block's rows of rowsum[phase].
#ifndef NBODY_SYM_RING
#define NBODY_SYM_RING 2048   // 16 MB of slices: a ring is 3.5 waves of CTAs long and short enough that a slice is overwritten before the L2 writes it back (profiles/r2_ring_sweep.log)
#endif
constexpr int SYM_RING = NBODY_SYM_RING;
struct SymArgs {
    const double2 *pos;
    const double2 *mr;
    double2 *slices;      // [SYM_RING][SYM_RB] row sums per work item, reused round the ring
    double2 *rowsum;      // [2][npad] row sums over the long chunks / over the short chunks, chunk ascending
    const int2 *groups;   // [ngroups] (first item, items) of every group
    int *gcnt;            // [ngroups] items that have stored their slice       } zero at launch
    int *gdone;           // [ngroups] 1 once the group's slices have been added } (sym_reduce_kernel leaves them so)
    // Reaction (column-side) sums: vacc[v - v0][j] collects, for virtual rank v, the reactions on body j from v's
    // row blocks below block(j), B ascending.  The CTAs of one (v, segment) chain add their share one after the
    // other: turn[(v - v0) * nseg + seg] counts the blocks that have added theirs (forces_sym.cuh).  Both are zero
    // at launch (sym_reduce_kernel leaves them so).
    double2 *vacc;        // [nv][npad]
    int *turn;            // [nv][nseg]
    int nseg;
    int *error_word;      // set to SYM_ERR_TURN when a CTA gave up waiting for its turn
    const double *boxes;  // [ceil(n/TILE_J)][8] per-tile bounding boxes (xmin ymin xmax ymax rmax . . .), or nullptr
    const int *orig;      // cell order: slot -> original body index (decides the sign for coincident bodies), or nullptr
    int n;
    int npad;
    ChunkMap cm;
    int v0, nv;           // this device owns the row blocks B with sym_vrank(B) in [v0, v0+nv): "virtual ranks"
    // The grid holds live work items only, one CTA per item (x = row block, y = chunk), in dispatch order: row block
    // by row block (ascending) over the long chunks, then row block by row block over the short chunks -- the
    // predecessor of an item in its (v, segment) chain is at least a row of items ahead of it, and the short items
    // still fill the tail of the launch.  item = (row block, chunk, group, group whose fold must be over before this
    // item may overwrite its ring slot, or -1).
    const int4 *items;
};
constexpr int SYM_ERR_TURN = 16;

// The symmetric kernel's partial sums are combined through SYM_V fixed "virtual ranks": row block B
// belongs to virtual rank sym_vrank(B); vbuf[v][j] = reactions on j from v's blocks below block(j)
// (B ascending) + (if block(j) belongs to v) j's own row sums (chunk ascending); F(j) = sum over
// v = 0..SYM_V-1 of vbuf[v][j].  A device owns SYM_V/P consecutive virtual ranks, so the sum --
// and therefore every bit of the result -- is the same for P = 1, 2, 4, 8 GPUs.
constexpr int SYM_V = 8;

// Row block -> virtual rank, boustrophedon over groups of 2*SYM_V blocks (0..7 then 7..0): block B's work
// falls linearly with B, so every virtual rank gets the same share and any split of the virtual ranks over
// GPUs is balanced.
__host__ __device__ inline int sym_vrank(int B)
{
    const int r = B % (2 * SYM_V);
    return r < SYM_V ? r : 2 * SYM_V - 1 - r;
}
// position of row block B among the blocks of its virtual rank, ascending (16g+v, 16g+15-v, 16(g+1)+v, ...)
__host__ __device__ inline int sym_vrank_seq(int B)
{
    return 2 * (B / (2 * SYM_V)) + ((B % (2 * SYM_V)) < SYM_V ? 0 : 1);
}
struct SymReduceArgs {
    const double2 *rowsum; // [2][npad] the force kernel's row sums (long chunks, short chunks)
    int *gcnt, *gdone;     // [ngroups] zeroed for the next launch
    int ngroups;
    double2 *vacc;        // [nv][npad] reaction sums of the force kernel; read, then zeroed for the next launch
    int *turn;            // [nv][nseg] zeroed for the next launch
    int nseg;
    double2 *vbuf;        // [SYM_V][npad]
    int n, npad;
    ChunkMap cm;
    int v0, nv;
};

// Peer-memory view of a sharded job (all pointers are device addresses valid on THIS device: own memory,
// cudaDeviceEnablePeerAccess mappings, or cudaIpcOpenMemHandle mappings).
constexpr int MAX_PEERS = 8;
constexpr int P2P_PHASES = 3;   // 0: force vectors (exchange A), 1: integrated positions (B), 2: uploaded / published state
struct PeerTable {
    double2 *vbuf[MAX_PEERS];
    double2 *pos[2][MAX_PEERS];
    double2 *vel[MAX_PEERS];
    double2 *frc[MAX_PEERS];
    double *boxes[MAX_PEERS];
    unsigned long long *flags[MAX_PEERS];   // [P2P_PHASES][MAX_PEERS sources] epoch words on each device
    unsigned long long *epoch;              // [P2P_PHASES] this device's own count of barriers passed per phase
    int npeers;                             // 0 = no peer exchange
    int me;
};

struct IntegrateArgs {
    const double2 *pos_old;
    double2 *pos_new;
    const double2 *mr;
    double2 *vel;
    double2 *frc;          // FAST: written (sum of partials); EXACT: read
    const double2 *part;   // FAST: partial sums (row side)
    const double2 *vbuf;   // FAST symmetric: [SYM_V][npad] per-virtual-rank sums (see SymReduceArgs), else nullptr
    int nsplit;            // FAST: number of j-chunks
    int npad;
    int n;                 // bodies
    int lo, hi;
    int xdim, ydim;
    double2 *peer_pos_new[MAX_PEERS];   // fused position exchange: store pos_new[i] here instead (npeers entries)
    double *boxes;                      // fused tile boxes of the NEW positions (cell order; lo % TILE_J == 0), or nullptr
    double *peer_boxes[MAX_PEERS];      // with the fused exchange: every device's box array
    int npeers;                         // 0 = store to pos_new / boxes only
    const int *error_word;              // device error word of the symmetric kernel / peer barriers (nullptr: none): once it is
                                        // nonzero the step is lost and integrate_kernel leaves the state as it is
};

// Cell order (kernels_sort.cu): key = Hilbert index of the position << 32 | original index, value = slot.
struct SortArgs {
    const double2 *pos;
    const int *orig;
    int n;
    double scale;                   // world coordinate -> 16-bit cell coordinate
    unsigned long long *keys_in, *keys_out;
    int *vals_in, *vals_out;
    void *temp;
    size_t temp_bytes;
};
struct PermuteArgs {
    const int *src_slot;            // sorted values: new slot t takes old slot src_slot[t]
    const double2 *pos_in, *mr_in, *vel_in, *frc_in;
    const int *orig_in;
    double2 *pos_out, *mr_out, *vel_out, *frc_out;
    int *orig_out;
    int n;
};
// Planar host layout <-> the device's double2 arrays (kernels_sort.cu).
struct ScatterArgs {                // uploaded rows [lo, hi) of planar (a, b) -> double2 array(s)
    const double *a, *b;            // planar device staging, indexed by body
    double2 *dst[MAX_PEERS];        // ndst destinations (this device, or every device of a peer-mapped job)
    int ndst;
    int lo, hi;
};
struct GatherArgs {                 // double2 arrays in slot order -> planar arrays in the caller's body order
    const double2 *src[3];          // up to three arrays (nullptr = skip)
    double *a[3], *b[3];
    const int *orig;                // slot -> body (nullptr = identity)
    int n;
    int first, count;               // only bodies [first, first + count) are written, at index body - first
};
struct PublishArgs {                // store this device's rows of up to two sharded arrays into every peer
    const double2 *src[2];
    double2 *dst[2][MAX_PEERS];
    int npeers, me;
    int lo, hi;
};
size_t sort_temp_bytes(int n);
unsigned cell_key_host(double x, double y, double scale);
int launch_sort(const SortArgs &a, cudaStream_t st);   // returns cudaError_t
void launch_permute(const PermuteArgs &a, cudaStream_t st);
void launch_unsort(const double2 *in, const int *orig, double2 *out, int n, cudaStream_t st);
void launch_iota(int *orig, int n, cudaStream_t st);
void launch_tile_boxes(const double2 *pos, const double2 *mr, int n, double *boxes, cudaStream_t st);
void launch_scatter(const ScatterArgs &a, cudaStream_t st);
void launch_gather_planar(const GatherArgs &a, cudaStream_t st);
void launch_publish(const PublishArgs &a, cudaStream_t st);

// The whole step loop of a small world in ONE launch (smalln.cuh, EXACT mode).
constexpr int SMALL_N_MAX = 256;    // bodies
constexpr int SMALL_N_ONE_CTA = 32; // up to here one CTA; above, a cluster of 16 CTAs (SMALL_CLUSTER where 16 cannot be co-scheduled)
constexpr int SMALL_CLUSTER = 8;
struct SmallArgs {
    double2 *pos[2];      // ping-pong positions in HBM; `cur` holds the state at entry
    const double2 *mr;
    double2 *vel;
    double2 *frc;
    int n, steps, cur;
    int xdim, ydim;
};
int launch_small_world(const SmallArgs &a, cudaStream_t st);   // returns cudaError_t

// launchers, defined in kernels_exact.cu / kernels_fast.cu
int launch_forces_exact(const ForceArgs &a, cudaStream_t st);   // returns cudaError_t
void launch_integrate_exact(const IntegrateArgs &a, cudaStream_t st);
void launch_forces_fast(const ForceArgs &a, int nsplit, cudaStream_t st);
void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st);
void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st);
int launch_forces_sym(const SymArgs &a, int nitems, cudaStream_t st);   // returns cudaError_t
void launch_render(const double2 *pos, const double2 *mr, int n, int xdim, int ydim, int *owner, unsigned char *rgb,
                   cudaStream_t st);
void launch_sym_reduce(const SymReduceArgs &a, cudaStream_t st);
void launch_sym_reduce_p2p(const SymReduceArgs &a, const PeerTable &t, int rows_per_shard, cudaStream_t st);
void launch_p2p_barrier(const PeerTable &t, int phase, int *error_word, cudaStream_t st);

}  // namespace nbody
