// Conservative footprint culling shared by the forward and backward blend kernels.
//
// The reference evaluates every (pixel, instance) pair of a tile and discards the pair when
// alpha = min(0.99, o * exp(power)) < 1/255 (forward.cu:339-349).  alpha >= 1/255 needs
// q(d) = -power = 0.5 (A dx^2 + C dy^2) + B dx dy <= L = log(255 o), i.e. the pixel lies inside an ellipse around
// the splat centre.  The blend kernels give each warp an 8x4 pixel block and let it skip every instance whose
// ellipse cannot reach the block:
//   stage 1  axis-aligned bounding box of the ellipse against the block rectangle (4 compares);
//   stage 2  exact minimum of q over the block rectangle (q is convex: 0 if the centre is inside, else the
//            smallest of the four edge minima).
// Both stages test against an inflated level Lq >= L that over-covers every rounding error on the way:
//   * the reference's float evaluation of `power` is off by at most ~4e-7 * S, S = sum of the magnitudes of its three
//     products; with u = sqrt(A)|dx|, v = sqrt(C)|dy|, rho = B / sqrt(AC):
//     q >= (1 - |rho|) (u^2 + v^2)/2 and S <= (1 + |rho|) (u^2 + v^2)/2, hence S <= kappa * q, kappa = (1+|rho|)/(1-|rho|);
//   * a pixel is dropped only if q > Lq = (L + 1e-4) / (1 - 1e-6 kappa), which implies q - 4e-7 S > L + 1e-4: the
//     reference's own float `power` is then below -L by a margin (1e-4) far above the error of expf / logf, so the
//     reference discards that pair too;
//   * kappa so large that 1e-6 kappa >= 0.5 (degenerate conics) switches the culling off for that Gaussian.
// A pair that survives is evaluated with the reference's exact expression, so blended values, n_contrib and all
// gradients are those of the reference; a dropped pair would have been discarded there as well.
//
// The per-Gaussian part is computed once in the preprocess kernel and stored with everything else the blend kernels
// need about a splat in ONE 64-byte record (two 32-byte sectors, four 16-byte loads):
//   rec[0] = {x, y, hx, hy}     pixel centre, half extents of the keep region    -> stage 1
//   rec[1] = {A, B, C, o}       conic, opacity                                   -> stage 2 + blend
//   rec[2] = {r, g, b, depth}                                                    -> blend
//   rec[3] = {Lq, thr, -B/C, -B/A}  inflated level; thr = -L - 1e-4 is the per-pixel pre-test that avoids expf; the two
//                                   slopes locate the minimum of q along a vertical / horizontal edge (stage 2)
// hx = hy = -3e38 marks "cannot be seen at all" (o <= 1/255), hx = Lq = +3e38 "never cull".
#pragma once
#include <cuda_runtime.h>

namespace rg {

constexpr int kBlkW = 8, kBlkH = 4;  // pixel block of one warp

__device__ __forceinline__ void write_splat_rec(float4* rec, float2 xy, float3 conic, float o, float3 color, float depth) {
    const float A = conic.x, B = conic.y, C = conic.z;
    const float big = 3.0e38f;
    float hx, hy, Lq, thr;
    const float L = logf(255.0f * o);
    const float det = A * C - B * B;
    thr = -L - 1e-4f;
    if (o <= 0.f) {  // alpha <= 0 everywhere (a NaN opacity falls through to "never cull")
        hx = hy = -big; Lq = 0.f; thr = big;
    } else if (!(A > 0.f && C > 0.f && det > 0.f) || !(L == L)) {
        hx = hy = Lq = big;
    } else if (L < -1e-3f) {  // o < 1/255: no pixel reaches alpha >= 1/255
        hx = hy = -big; Lq = 0.f;
    } else {
        const float rho = fminf(fabsf(B) * rsqrtf(A * C), 1.0f);
        const float kappa = (1.f + rho) / fmaxf(1.f - rho, 1e-12f);
        if (!(1e-6f * kappa < 0.5f)) {
            hx = hy = Lq = big;
        } else {
            Lq = (fmaxf(L, 0.f) + 1e-4f) / (1.f - 1e-6f * kappa) * 1.00001f;
            const float inv = 1.0f / det;
            hx = sqrtf(2.f * Lq * C * inv) * 1.0001f + 1e-3f;
            hy = sqrtf(2.f * Lq * A * inv) * 1.0001f + 1e-3f;
        }
    }
    rec[0] = {xy.x, xy.y, hx, hy};
    rec[1] = {A, B, C, o};
    rec[2] = {color.x, color.y, color.z, depth};
    rec[3] = {Lq, thr, -B / C, -B / A};
}

// Stage 1: can the keep region's bounding box touch the block [x0, x1] x [y0, y1] (pixel-centre coordinates)?
__device__ __forceinline__ bool cull_aabb(float4 r0, float x0, float y0, float x1, float y1) {
    return !(r0.x + r0.z < x0 || r0.x - r0.z > x1 || r0.y + r0.w < y0 || r0.y - r0.w > y1);
}

// Stage 2: true when the instance may contribute to some pixel of the block.
__device__ __forceinline__ bool cull_exact(float4 r0, float4 co, float4 r3, float x0, float y0, float x1, float y1) {
    const float Lq = r3.x;
    if (Lq > 1e37f) return true;
    const float dxlo = r0.x - x1, dxhi = r0.x - x0, dylo = r0.y - y1, dyhi = r0.y - y0;
    if (dxlo <= 0.f && dxhi >= 0.f && dylo <= 0.f && dyhi >= 0.f) return true;  // centre inside the block
    const float A = co.x, B = co.y, C = co.z;
    const float rBC = r3.z, rBA = r3.w;
    float qmin = 3.0e38f;
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const float dx = e ? dxhi : dxlo;
        const float dy = fminf(fmaxf(rBC * dx, dylo), dyhi);
        qmin = fminf(qmin, 0.5f * (A * dx * dx + C * dy * dy) + B * dx * dy);
        const float ey = e ? dyhi : dylo;
        const float ex = fminf(fmaxf(rBA * ey, dxlo), dxhi);
        qmin = fminf(qmin, 0.5f * (A * ex * ex + C * ey * ey) + B * ex * ey);
    }
    // this file's own rounding in qmin is covered by the same kappa argument (relative 1e-6 kappa) plus the factor
    // 1.00001 folded into Lq
    return qmin <= Lq;
}

}  // namespace rg
