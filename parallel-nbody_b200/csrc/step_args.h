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
constexpr int SYM_WIDE_N = 131072;  // from this many bodies on: 64 threads x 8 bodies per CTA instead of 128 x 4
constexpr int SYM_BOX_DOUBLES = 8;  // per-tile bounding box: xmin ymin xmax ymax rmax + padding (one 64-byte bulk copy)

struct SymArgs {
    const double2 *pos;
    const double2 *mr;
    double2 *rowpart;     // [nsplit][npad]  row-side partial sums
    double2 *colpart;     // [nblocks][npad] reaction (column-side) partial sums, valid for j >= (B+1)*SYM_RB
    const double *boxes;  // [ceil(n/TILE_J)][8] per-tile bounding boxes (xmin ymin xmax ymax rmax . . .), or nullptr
    int n;
    int npad;
    int chunk;            // j-chunk length, a multiple of SYM_RB
    int v0, nv;           // this device owns the row blocks B with (B % SYM_V) in [v0, v0+nv): "virtual ranks"
    int *counter;         // persistent launch: work-item counter, zeroed before every launch (else nullptr)
    int nitems;           // persistent launch: nblocks_here * nsplit
    int nblocks_here, nsplit;
};

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
// local slot of row block B in a device's colpart array (device owns virtual ranks [v0, v0+nv))
__host__ __device__ inline int sym_local_block(int B, int v0, int nv)
{
    const int r = B % (2 * SYM_V);
    const int half = r < SYM_V ? 0 : 1;
    return (B / (2 * SYM_V)) * 2 * nv + half * nv + (sym_vrank(B) - v0);
}
// inverse: k-th owned block (ascending B within each virtual rank)
__host__ __device__ inline int sym_owned_block(int k, int v0, int nv)
{
    const int g = k / (2 * nv), r = k % (2 * nv);
    return r < nv ? g * 2 * SYM_V + v0 + r : g * 2 * SYM_V + 2 * SYM_V - 1 - (v0 + r - nv);
}
struct SymReduceArgs {
    const double2 *rowpart;
    const double2 *colpart;
    double2 *vbuf;        // [SYM_V][npad]
    int n, npad, chunk, nsplit;
    int v0, nv;
};

// Peer-memory view of a sharded job (all pointers are device addresses valid on THIS device: own memory,
// cudaDeviceEnablePeerAccess mappings, or cudaIpcOpenMemHandle mappings).
constexpr int MAX_PEERS = 8;
struct PeerTable {
    double2 *vbuf[MAX_PEERS];
    double2 *pos[2][MAX_PEERS];
    unsigned long long *flags[MAX_PEERS];   // [2 phases][MAX_PEERS sources] epoch words on each device
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
    int lo, hi;
    int xdim, ydim;
    double2 *peer_pos_new[MAX_PEERS];   // fused position exchange: also store pos_new[i] here (npeers entries)
    int npeers;                         // 0 = store to pos_new only
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
size_t sort_temp_bytes(int n);
unsigned cell_key_host(double x, double y, double scale);
int launch_sort(const SortArgs &a, cudaStream_t st);   // returns cudaError_t
void launch_permute(const PermuteArgs &a, cudaStream_t st);
void launch_unsort(const double2 *in, const int *orig, double2 *out, int n, cudaStream_t st);
void launch_iota(int *orig, int n, cudaStream_t st);
void launch_tile_boxes(const double2 *pos, const double2 *mr, int n, double *boxes, cudaStream_t st);

// launchers, defined in kernels_exact.cu / kernels_fast.cu
void launch_forces_exact(const ForceArgs &a, cudaStream_t st);
void launch_integrate_exact(const IntegrateArgs &a, cudaStream_t st);
void launch_forces_fast(const ForceArgs &a, int nsplit, cudaStream_t st);
void launch_integrate_fast(const IntegrateArgs &a, cudaStream_t st);
void launch_reduce_fast(const IntegrateArgs &a, cudaStream_t st);
int launch_forces_sym(SymArgs a, int nblocks_here, int nsplit, cudaStream_t st);   // returns cudaError_t
void launch_render(const double2 *pos, const double2 *mr, int n, int xdim, int ydim, int *owner, unsigned char *rgb,
                   cudaStream_t st);
void launch_sym_reduce(const SymReduceArgs &a, cudaStream_t st);
void launch_sym_reduce_p2p(const SymReduceArgs &a, const PeerTable &t, int rows_per_shard, cudaStream_t st);
void launch_p2p_signal(const PeerTable &t, int phase, unsigned long long epoch, cudaStream_t st);
void launch_p2p_wait(const PeerTable &t, int phase, unsigned long long epoch, int *error_word, cudaStream_t st);

}  // namespace nbody
