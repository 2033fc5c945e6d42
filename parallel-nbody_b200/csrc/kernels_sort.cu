// kernels_sort.cu -- cell order for the symmetric fast kernel's far-field test (forces_sym.cuh).
//
// The reference keeps bodies in creation order (nbody-seq.c:440-449), which is random in space.  The fast
// symmetric kernel only needs SOME fixed order, so the library keeps its device arrays in "cell order":
// bodies sorted by the Hilbert index of their position (16 bits per axis) with the original index as the
// tie-break, which makes every run of 256 consecutive bodies a compact patch of the world.  orig[s] is the
// original index of slot s; fetch/forces/render undo the permutation, so callers never see it.
//
//   sort_keys_kernel   key = hilbert(cell(x, y)) << 32 | orig, value = slot
//   cub radix sort     (keys, values)                      -- library call, not on the step's hot path
//   permute_kernel     gathers pos / mr / vel / frc / orig through the sorted values
//   tile_boxes_kernel  per TILE_J bodies: x/y range and largest radius (every step: positions move)
//   unsort_kernel      out[orig[s]] = in[s]
//   gather_planar_kernel   up to three double2 arrays in slot order -> planar x[] / y[] arrays in the caller's
//                          body order (what nbody_gpu_fetch copies to the host, one copy per array)
//   scatter_kernel     planar (a[], b[]) uploaded by nbody_gpu_set_state -> double2 rows, on this device or,
//                      with the peer-mapped exchange, on every device (each rank uploads only its own rows)
//   publish_kernel     a device's rows of the sharded velocity / force arrays -> every peer (before a fetch or
//                      a re-sort of a peer-mapped job; replaces two ncclAllGather calls)
//
// Coincident bodies share a cell, so they stay in original-index order and the reference's convention for
// them (lower index gets +force, nbody-seq.c:91-106) reads the same in slot order.
#include <cub/device/device_radix_sort.cuh>

#include <cfloat>

#include "step_args.h"

namespace nbody {

__host__ __device__ __forceinline__ unsigned hilbert_index16(unsigned x, unsigned y)
{
    constexpr unsigned N = 1u << 16;
    unsigned d = 0;
#pragma unroll
    for (unsigned s = N / 2; s > 0; s >>= 1) {
        const unsigned rx = (x & s) ? 1u : 0u;
        const unsigned ry = (y & s) ? 1u : 0u;
        d += s * s * ((3u * rx) ^ ry);
        if (ry == 0) {
            if (rx == 1) {
                x = N - 1 - x;
                y = N - 1 - y;
            }
            const unsigned t = x;
            x = y;
            y = t;
        }
    }
    return d;
}

__global__ void __launch_bounds__(256) sort_keys_kernel(const double2 *pos, const int *orig, int n, double scale,
                                                        unsigned long long *keys, int *vals)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const double2 p = pos[s];
    const double cx = fmin(fmax(p.x * scale, 0.0), 65535.0);   // NaN -> 0: a broken world still sorts
    const double cy = fmin(fmax(p.y * scale, 0.0), 65535.0);
    const unsigned d = hilbert_index16(static_cast<unsigned>(cx), static_cast<unsigned>(cy));
    keys[s] = (static_cast<unsigned long long>(d) << 32) | static_cast<unsigned>(orig[s]);
    vals[s] = s;
}

__global__ void __launch_bounds__(256) permute_kernel(const PermuteArgs a)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n) return;
    const int s = a.src_slot[t];
    a.pos_out[t] = a.pos_in[s];
    a.mr_out[t] = a.mr_in[s];
    a.vel_out[t] = a.vel_in[s];
    a.frc_out[t] = a.frc_in[s];
    a.orig_out[t] = a.orig_in[s];
}

__global__ void __launch_bounds__(256) unsort_kernel(const double2 *in, const int *orig, double2 *out, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) out[orig[s]] = in[s];
}

__global__ void __launch_bounds__(256) gather_planar_kernel(const GatherArgs g)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= g.n) return;
    const int o = (g.orig ? g.orig[s] : s) - g.first;
    if (o < 0 || o >= g.count) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (g.src[k]) {
            const double2 v = g.src[k][s];
            g.a[k][o] = v.x;
            g.b[k][o] = v.y;
        }
    }
}

__global__ void __launch_bounds__(256) scatter_kernel(const ScatterArgs g)
{
    const int i = g.lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.hi) return;
    const double2 v = make_double2(g.a[i], g.b[i]);
    for (int d = 0; d < g.ndst; ++d) g.dst[d][i] = v;
}

__global__ void __launch_bounds__(256) publish_kernel(const PublishArgs g)
{
    const int i = g.lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.hi) return;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        if (g.src[k]) {
            const double2 v = g.src[k][i];
            for (int r = 0; r < g.npeers; ++r)
                if (r != g.me) g.dst[k][r][i] = v;
        }
    }
}

__global__ void __launch_bounds__(256) iota_kernel(int *orig, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) orig[s] = s;
}

// One CTA of TILE_J threads per tile: boxes[tile] = {xmin, ymin, xmax, ymax, rmax, 0, 0, 0} over the live bodies.
__global__ void __launch_bounds__(TILE_J) tile_boxes_kernel(const double2 *pos, const double2 *mr, int n, double *boxes)
{
    __shared__ double red[5][TILE_J / 32];
    const int s = blockIdx.x * TILE_J + threadIdx.x;
    double v[5] = {DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, 0.0};
    if (s < n) {
        const double2 p = pos[s];
        v[0] = p.x; v[1] = p.y; v[2] = p.x; v[3] = p.y;
        v[4] = mr[s].y;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        v[0] = fmin(v[0], __shfl_xor_sync(0xffffffffu, v[0], off));
        v[1] = fmin(v[1], __shfl_xor_sync(0xffffffffu, v[1], off));
        v[2] = fmax(v[2], __shfl_xor_sync(0xffffffffu, v[2], off));
        v[3] = fmax(v[3], __shfl_xor_sync(0xffffffffu, v[3], off));
        v[4] = fmax(v[4], __shfl_xor_sync(0xffffffffu, v[4], off));
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < 5; ++q) red[q][warp] = v[q];
    }
    __syncthreads();
    if (threadIdx.x < SYM_BOX_DOUBLES) {
        const int q = threadIdx.x;
        double r = 0.0;
        if (q < 5) {
            r = red[q][0];
            for (int w = 1; w < TILE_J / 32; ++w) r = (q < 2) ? fmin(r, red[q][w]) : fmax(r, red[q][w]);
        }
        boxes[static_cast<size_t>(blockIdx.x) * SYM_BOX_DOUBLES + q] = r;
    }
}

// Host mirror of sort_keys_kernel's key (without the original-index tie-break), for tests of the curve itself.
unsigned cell_key_host(double x, double y, double scale)
{
    const double cx = fmin(fmax(x * scale, 0.0), 65535.0);
    const double cy = fmin(fmax(y * scale, 0.0), 65535.0);
    return hilbert_index16(static_cast<unsigned>(cx), static_cast<unsigned>(cy));
}

size_t sort_temp_bytes(int n)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, static_cast<const unsigned long long *>(nullptr),
                                    static_cast<unsigned long long *>(nullptr), static_cast<const int *>(nullptr),
                                    static_cast<int *>(nullptr), n);
    return bytes;
}

int launch_sort(const SortArgs &a, cudaStream_t st)
{
    if (a.n <= 0) return 0;
    const int grid = (a.n + 255) / 256;
    sort_keys_kernel<<<grid, 256, 0, st>>>(a.pos, a.orig, a.n, a.scale, a.keys_in, a.vals_in);
    size_t bytes = a.temp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(a.temp, bytes, a.keys_in, a.keys_out, a.vals_in, a.vals_out, a.n, 0, 64, st);
    return static_cast<int>(e);
}

void launch_permute(const PermuteArgs &a, cudaStream_t st)
{
    if (a.n <= 0) return;
    permute_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(a);
}

void launch_unsort(const double2 *in, const int *orig, double2 *out, int n, cudaStream_t st)
{
    if (n <= 0) return;
    unsort_kernel<<<(n + 255) / 256, 256, 0, st>>>(in, orig, out, n);
}

void launch_gather_planar(const GatherArgs &a, cudaStream_t st)
{
    if (a.n <= 0) return;
    gather_planar_kernel<<<(a.n + 255) / 256, 256, 0, st>>>(a);
}

void launch_scatter(const ScatterArgs &a, cudaStream_t st)
{
    if (a.hi <= a.lo) return;
    scatter_kernel<<<(a.hi - a.lo + 255) / 256, 256, 0, st>>>(a);
}

void launch_publish(const PublishArgs &a, cudaStream_t st)
{
    if (a.hi <= a.lo) return;
    publish_kernel<<<(a.hi - a.lo + 255) / 256, 256, 0, st>>>(a);
}

void launch_iota(int *orig, int n, cudaStream_t st)
{
    if (n <= 0) return;
    iota_kernel<<<(n + 255) / 256, 256, 0, st>>>(orig, n);
}

void launch_tile_boxes(const double2 *pos, const double2 *mr, int n, double *boxes, cudaStream_t st)
{
    if (n <= 0) return;
    tile_boxes_kernel<<<(n + TILE_J - 1) / TILE_J, TILE_J, 0, st>>>(pos, mr, n, boxes);
}

}  // namespace nbody
