// Conservative footprint culling shared by the forward and backward blend kernels.
//
// The reference evaluates every (pixel, instance) pair of a tile and discards the pair when
// alpha = min(0.99, o * exp(power)) < 1/255 (forward.cu:339-349).  alpha >= 1/255 needs
// q(d) = -power = 0.5 (A dx^2 + C dy^2) + B dx dy <= log(255 o), i.e. the pixel lies inside an ellipse
// around the splat centre.  The blend kernels therefore split each 16x16 tile into eight 8x4 pixel blocks
// (one warp each) and give a warp only the instances whose ellipse can reach its block:
//   stage 1  axis-aligned bounding box of the ellipse against the block rectangle (4 compares);
//   stage 2  exact minimum of q over the block rectangle (q is convex: 0 if the centre is inside, else
//            the smallest of the four edge minima).
// Both stages test against log(255 o) + slack, where slack over-covers every rounding error that the
// reference's float evaluation of `power`, expf and this file's own arithmetic can have on the tile
// (4e-7 * sum of the magnitudes of the three products, bounded over the whole tile, + 1e-4).  A pair that
// survives is then evaluated with the reference's exact expression, so the blended result, n_contrib and
// every gradient are those of the reference; a pair that is dropped would have been discarded there too.
#pragma once
#include <cuda_runtime.h>

namespace rg {

constexpr int kBlkW = 8, kBlkH = 4;  // pixel block of one warp

struct CullRec {     // per staged instance
    float4 bb;       // xmin, xmax, ymin, ymax of the keep region (empty when xmin > xmax)
    float4 q;        // Lq (level incl. slack), -B/C, -B/A, flag (1 = exact stage usable)
};

// Built once per (tile, instance) by the staging thread.  tile_x0/y0: first pixel of the tile.
__device__ __forceinline__ CullRec make_cull(float2 xy, float4 co, float tile_x0, float tile_y0) {
    CullRec r;
    const float A = co.x, B = co.y, C = co.z, o = co.w;
    const float det = A * C - B * B;
    const float big = 3.0e38f;
    if (!(o > 0.f)) {  // alpha <= 0 < 1/255 for every pixel (NaN opacity falls through to "keep")
        if (o <= 0.f) { r.bb = {1.f, 0.f, 1.f, 0.f}; r.q = {0.f, 0.f, 0.f, 0.f}; return r; }
    }
    const float L = logf(255.0f * o);  // alpha >= 1/255  <=>  q <= L
    // bound of |0.5 A dx^2| + |0.5 C dy^2| + |B dx dy| over the tile
    const float dxm = fmaxf(fabsf(xy.x - tile_x0), fabsf(xy.x - (tile_x0 + 15.f)));
    const float dym = fmaxf(fabsf(xy.y - tile_y0), fabsf(xy.y - (tile_y0 + 15.f)));
    const float S = 0.5f * (fabsf(A) * dxm * dxm + fabsf(C) * dym * dym) + fabsf(B) * dxm * dym;
    const float Lq = L + 1e-4f + 1e-6f * S + 1e-6f * fabsf(L);
    if (!(A > 0.f && C > 0.f && det > 0.f) || !(Lq == Lq) || Lq > 1e30f) {  // not a proper ellipse: never cull
        r.bb = {-big, big, -big, big};
        r.q = {big, 0.f, 0.f, 0.f};
        return r;
    }
    if (Lq < 0.f) {  // o < 1/255: no pixel can reach alpha >= 1/255
        r.bb = {1.f, 0.f, 1.f, 0.f}; r.q = {0.f, 0.f, 0.f, 0.f};
        return r;
    }
    const float inv = 1.0f / det;
    const float hx = sqrtf(2.f * Lq * C * inv) * 1.0001f + 1e-3f;
    const float hy = sqrtf(2.f * Lq * A * inv) * 1.0001f + 1e-3f;
    r.bb = {xy.x - hx, xy.x + hx, xy.y - hy, xy.y + hy};
    r.q = {Lq, -B / C, -B / A, 1.f};
    return r;
}

// True when the instance may contribute to some pixel of the block [x0, x0+w-1] x [y0, y0+h-1].
__device__ __forceinline__ bool cull_keep(const CullRec& r, float2 xy, float4 co, float x0, float y0, float x1, float y1) {
    if (r.bb.y < x0 || r.bb.x > x1 || r.bb.w < y0 || r.bb.z > y1) return false;
    if (r.q.w == 0.f) return true;
    const float dxlo = xy.x - x1, dxhi = xy.x - x0, dylo = xy.y - y1, dyhi = xy.y - y0;
    if (dxlo <= 0.f && dxhi >= 0.f && dylo <= 0.f && dyhi >= 0.f) return true;  // centre inside the block
    const float A = co.x, B = co.y, C = co.z;
    float qmin = 3.0e38f;
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const float dx = e ? dxhi : dxlo;
        const float dy = fminf(fmaxf(r.q.y * dx, dylo), dyhi);
        qmin = fminf(qmin, 0.5f * (A * dx * dx + C * dy * dy) + B * dx * dy);
        const float ey = e ? dyhi : dylo;
        const float ex = fminf(fmaxf(r.q.z * ey, dxlo), dxhi);
        qmin = fminf(qmin, 0.5f * (A * ex * ex + C * ey * ey) + B * ex * ey);
    }
    return qmin <= r.q.x;
}

}  // namespace rg
