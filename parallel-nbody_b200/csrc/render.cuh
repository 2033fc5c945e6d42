// render.cuh -- the reference's display() (nbody-seq.c:326-354) as CUDA kernels on a headless device buffer.
//
// Reference: for every pixel (i, j) find the FIRST body b (lowest index) with
//     sqrt((X(b)-i)^2 + (Y(b)-j)^2) <= R(b) + 0.5
// and paint it with tint = 0xfff*(b+1)/(bodyCt+2) (nbody-seq.c:307-316), else black: O(W*H*N) on the CPU.
// Here: one CTA per body walks only the body's bounding box and does atomicMin(owner[pixel], b) with the
// very same predicate (IEEE sqrt, no FMA: this header is compiled in the -fmad=false translation unit), so
// "first covering body" = minimum index, identical bytes; then one thread per pixel paints.
#pragma once
#include <cuda_runtime.h>

#include <climits>

namespace nbody {

__global__ void render_clear_kernel(int *owner, int npix)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < npix) owner[p] = INT_MAX;
}

__global__ void __launch_bounds__(128) render_cover_kernel(const double2 *pos, const double2 *mr, int n, int xdim,
                                                            int ydim, int *owner)
{
    const int b = blockIdx.x;
    if (b >= n) return;
    const double2 p = pos[b];
    const double reach = __dadd_rn(mr[b].y, 0.5);   // R(b) + 0.5
    int x0 = static_cast<int>(floor(p.x - reach)) - 1, x1 = static_cast<int>(ceil(p.x + reach)) + 1;
    int y0 = static_cast<int>(floor(p.y - reach)) - 1, y1 = static_cast<int>(ceil(p.y + reach)) + 1;
    x0 = max(x0, 0); y0 = max(y0, 0);
    x1 = min(x1, xdim - 1); y1 = min(y1, ydim - 1);
    const int w = x1 - x0 + 1, hgt = y1 - y0 + 1;
    if (w <= 0 || hgt <= 0) return;
    for (int t = threadIdx.x; t < w * hgt; t += blockDim.x) {
        const int i = x0 + t % w, j = y0 + t / w;
        const double dy = __dsub_rn(p.y, static_cast<double>(j));
        const double dx = __dsub_rn(p.x, static_cast<double>(i));
        const double d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
        if (d <= reach) atomicMin(&owner[i + j * xdim], b);
    }
}

__global__ void render_paint_kernel(const int *owner, int npix, int n, unsigned char *rgb)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const int b = owner[p];
    unsigned char c0 = 0, c1 = 0, c2 = 0;
    if (b != INT_MAX) {
        // nbody-seq.c:311-315 (the product is formed in 64-bit: the reference's int overflows for b > 524287)
        const int tint = static_cast<int>((0xfffLL * (b + 1)) / (n + 2));
        c0 = static_cast<unsigned char>((tint & 0xf) << 4);
        c1 = static_cast<unsigned char>(tint & 0xf0);
        c2 = static_cast<unsigned char>((tint & 0xf00) >> 4);
    }
    rgb[3 * p + 0] = c0;
    rgb[3 * p + 1] = c1;
    rgb[3 * p + 2] = c2;
}

}  // namespace nbody
