// Dual-quaternion blending of the bone transforms (utils/dual_quaternion.py:135-180: QT2DQ, the weighted sum, DQ2QT),
// forward and backward, shared by the standalone skinning kernels (skin.cu) and the fused preprocess / per-Gaussian
// backward kernels.  A bone contributes  real = q / |q|  (its node_rot, which the reference detaches) and
// dual = 1/2 standardise((0, t) * real)  with t the translation of its global transform; the blend of a point is
//   B_r = sum_k w_k real_k,  B_d = sum_k w_k dual_k,  r0 = B_r / |B_r|,  d0 = B_d / |B_r|  (|B_r| clamped at 1e-8),
//   y = R(r0) x + 2 (d0 * conj(r0)).xyz.
// Gradients flow to the weights and to the bone translations (the rotation enters through a detached quaternion).
#pragma once
#include "bones.cuh"
#include "common.cuh"

namespace rg {

struct DqBone {
    float r[4];  // normalised rotation quaternion (w, x, y, z)
    float d[4];  // dual part, sign-standardised, already halved
    float sgn;   // +1 / -1: the sign the standardisation applied
};

__device__ __forceinline__ DqBone dq_of_bone(const BoneSm& b) {
    DqBone o;
    const float qn = rsqrtf(fmaxf(b.q[0] * b.q[0] + b.q[1] * b.q[1] + b.q[2] * b.q[2] + b.q[3] * b.q[3], 1e-24f));
    const float rw = b.q[0] * qn, rx = b.q[1] * qn, ry = b.q[2] * qn, rz = b.q[3] * qn;
    const float tx = b.G[9], ty = b.G[10], tz = b.G[11];
    float dw = -tx * rx - ty * ry - tz * rz;
    float dx = tx * rw + ty * rz - tz * ry;
    float dy = -tx * rz + ty * rw + tz * rx;
    float dz = tx * ry - ty * rx + tz * rw;
    o.sgn = dw < 0 ? -1.f : 1.f;  // quaternion_multiply standardises the product to a non-negative real part
    o.r[0] = rw; o.r[1] = rx; o.r[2] = ry; o.r[3] = rz;
    o.d[0] = 0.5f * o.sgn * dw; o.d[1] = 0.5f * o.sgn * dx; o.d[2] = 0.5f * o.sgn * dy; o.d[3] = 0.5f * o.sgn * dz;
    return o;
}

struct DqBlend {
    float r0[4], d0[4];  // normalised blend
    float rn;            // |B_r| (clamped)
    float y[3];          // R(r0) x + t
};

// Br / Bd: the weight-normalised sums
__device__ __forceinline__ DqBlend dq_finish(const float (&Br)[4], const float (&Bd)[4], float x, float y, float z) {
    DqBlend o;
    o.rn = fmaxf(sqrtf(Br[0] * Br[0] + Br[1] * Br[1] + Br[2] * Br[2] + Br[3] * Br[3]), 1e-8f);
    const float s = 1.0f / o.rn;
#pragma unroll
    for (int i = 0; i < 4; i++) { o.r0[i] = Br[i] * s; o.d0[i] = Bd[i] * s; }
    const float w0 = o.r0[0], x0 = o.r0[1], y0 = o.r0[2], z0 = o.r0[3];
    const float w1 = o.d0[0], x1 = o.d0[1], y1 = o.d0[2], z1 = o.d0[3];
    const float tx = 2 * (-w1 * x0 + x1 * w0 - y1 * z0 + z1 * y0);
    const float ty = 2 * (-w1 * y0 + x1 * z0 + y1 * w0 - z1 * x0);
    const float tz = 2 * (-w1 * z0 - x1 * y0 + y1 * x0 + z1 * w0);
    o.y[0] = (1 - 2 * y0 * y0 - 2 * z0 * z0) * x + (2 * x0 * y0 - 2 * w0 * z0) * y + (2 * x0 * z0 + 2 * w0 * y0) * z + tx;
    o.y[1] = (2 * x0 * y0 + 2 * w0 * z0) * x + (1 - 2 * x0 * x0 - 2 * z0 * z0) * y + (2 * y0 * z0 - 2 * w0 * x0) * z + ty;
    o.y[2] = (2 * x0 * z0 - 2 * w0 * y0) * x + (2 * y0 * z0 + 2 * w0 * x0) * y + (1 - 2 * x0 * x0 - 2 * y0 * y0) * z + tz;
    return o;
}

// Backward through dq_finish: gy = dL/dy (3), gh = dL/d r0 from the rotation output (4)  ->  gBr, gBd = dL/dB_r, dL/dB_d.
__device__ __forceinline__ void dq_finish_backward(const DqBlend& f, float px, float py, float pz, const float (&gy)[3],
                                                   const float (&gh)[4], float (&gBr)[4], float (&gBd)[4]) {
    const float qw = f.r0[0], qx = f.r0[1], qy = f.r0[2], qz = f.r0[3];
    const float w1 = f.d0[0], x1 = f.d0[1], y1 = f.d0[2], z1 = f.d0[3];
    float gr[4], gd[4];
    // d(R(r0) p)/d r0
    gr[0] = gy[0] * (-2 * qz * py + 2 * qy * pz) + gy[1] * (2 * qz * px - 2 * qx * pz) + gy[2] * (-2 * qy * px + 2 * qx * py);
    gr[1] = gy[0] * (2 * qy * py + 2 * qz * pz) + gy[1] * (2 * qy * px - 4 * qx * py - 2 * qw * pz) +
            gy[2] * (2 * qz * px + 2 * qw * py - 4 * qx * pz);
    gr[2] = gy[0] * (-4 * qy * px + 2 * qx * py + 2 * qw * pz) + gy[1] * (2 * qx * px + 2 * qz * pz) +
            gy[2] * (-2 * qw * px + 2 * qz * py - 4 * qy * pz);
    gr[3] = gy[0] * (-4 * qz * px - 2 * qw * py + 2 * qx * pz) + gy[1] * (2 * qw * px - 4 * qz * py + 2 * qy * pz) +
            gy[2] * (2 * qx * px + 2 * qy * py);
    // d t / d r0 and d t / d d0
    gr[0] += 2 * (gy[0] * x1 + gy[1] * y1 + gy[2] * z1);
    gr[1] += 2 * (-gy[0] * w1 - gy[1] * z1 + gy[2] * y1);
    gr[2] += 2 * (gy[0] * z1 - gy[1] * w1 - gy[2] * x1);
    gr[3] += 2 * (-gy[0] * y1 + gy[1] * x1 - gy[2] * w1);
    gd[0] = 2 * (-gy[0] * qx - gy[1] * qy - gy[2] * qz);
    gd[1] = 2 * (gy[0] * qw + gy[1] * qz - gy[2] * qy);
    gd[2] = 2 * (-gy[0] * qz + gy[1] * qw + gy[2] * qx);
    gd[3] = 2 * (gy[0] * qy - gy[1] * qx + gy[2] * qw);
#pragma unroll
    for (int i = 0; i < 4; i++) gr[i] += gh[i];
    // r0 = B_r / |B_r|, d0 = B_d / |B_r|
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) dot += f.r0[i] * gr[i] + f.d0[i] * gd[i];
    const float s = 1.0f / f.rn;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        gBr[i] = (gr[i] - f.r0[i] * dot) * s;
        gBd[i] = gd[i] * s;
    }
}

// dL/dt of one bone from dL/d(dual_k) = w_k gBd (dual_k = 1/2 sgn (0, t) * real_k is linear in t); returns it scaled by `scale`
__device__ __forceinline__ void dq_bone_dt(const DqBone& b, const float (&gBd)[4], float scale, float (&dt)[3]) {
    const float k = 0.5f * b.sgn * scale;
    const float rw = b.r[0], rx = b.r[1], ry = b.r[2], rz = b.r[3];
    dt[0] = k * (-rx * gBd[0] + rw * gBd[1] - rz * gBd[2] + ry * gBd[3]);
    dt[1] = k * (-ry * gBd[0] + rz * gBd[1] + rw * gBd[2] - rx * gBd[3]);
    dt[2] = k * (-rz * gBd[0] - ry * gBd[1] + rx * gBd[2] + rw * gBd[3]);
}

// Skinning backward of ONE point in dual-quaternion mode (the counterpart of the two linear-blend sweeps of
// skin_backward_kernel / gauss_backward_kernel<true>): a = dL/d(skinned position), h = dL/d(d_rotation), both WITHOUT the
// motion mask m.  Per-bone sums go to the warp's accumulator row s_acc (13 floats per bone: 12 of dG -- only the
// translation entries 9..11 are non-zero here -- and d radius), dL/dglobal_trans to s_acc[13 K ..]; every lane of the warp
// must call it (dead lanes with live = false).  mlp_row: this point's (J-1) MLP weights or null; dmlp_row / dmask: outputs
// or null.  Returns nothing; offsets' gradient (m a) is the caller's.
__device__ __forceinline__ void dq_skin_backward_point(const BoneSm* s_b, int K, int lane, bool live, float x, float y, float z,
                                                       float m, float ax, float ay, float az, float h0, float h1, float h2, float h3,
                                                       const float* mlp_row, float gt0, float gt1, float gt2, float ox, float oy,
                                                       float oz, float* s_acc, float* dmask, float* dmlp_row) {
    // sweep 1: the blend
    float wsum = 0.f, Br[4] = {0, 0, 0, 0}, Bd[4] = {0, 0, 0, 0};
    for (int k = 0; k < K; k++) {
        const BoneSm& b = s_b[k];
        float w = expf(-seg_dist2(b, x, y, z) * b.inv_2r2);
        if (mlp_row && live) w *= mlp_row[k];
        w += 1e-7f;
        const DqBone q = dq_of_bone(b);
        wsum += w;
#pragma unroll
        for (int i = 0; i < 4; i++) { Br[i] += w * q.r[i]; Bd[i] += w * q.d[i]; }
    }
    const float inv = 1.0f / wsum;
#pragma unroll
    for (int i = 0; i < 4; i++) { Br[i] *= inv; Bd[i] *= inv; }
    const DqBlend f = dq_finish(Br, Bd, x, y, z);
    if (dmask && live)
        *dmask = ax * (f.y[0] + gt0 + ox - x) + ay * (f.y[1] + gt1 + oy - y) + az * (f.y[2] + gt2 + oz - z) +
                 h0 * f.r0[0] + h1 * f.r0[1] + h2 * f.r0[2] + h3 * f.r0[3];
    const float mm = live ? m : 0.f;
    const float gy[3] = {mm * ax, mm * ay, mm * az};
    const float gh[4] = {mm * h0, mm * h1, mm * h2, mm * h3};
    float gBr[4], gBd[4];
    dq_finish_backward(f, x, y, z, gy, gh, gBr, gBd);
    float cbar = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) cbar += gBr[i] * Br[i] + gBd[i] * Bd[i];
    {
        const float s0 = warp_sum(gy[0]), s1 = warp_sum(gy[1]), s2 = warp_sum(gy[2]);
        if (lane == 0) { s_acc[K * 13] += s0; s_acc[K * 13 + 1] += s1; s_acc[K * 13 + 2] += s2; }
    }
    // sweep 2: per-bone terms
    const int my_id = (lane >> 1) & 15;
    for (int k = 0; k < K; k++) {
        const BoneSm& b = s_b[k];
        const float d2 = seg_dist2(b, x, y, z);
        const float e = expf(-d2 * b.inv_2r2);
        const float sw = (mlp_row && live) ? mlp_row[k] : 1.0f;
        const float w = (e * sw + 1e-7f) * inv;
        const DqBone q = dq_of_bone(b);
        float c = 0.f;
#pragma unroll
        for (int i = 0; i < 4; i++) c += gBr[i] * q.r[i] + gBd[i] * q.d[i];
        const float dwhat = (c - cbar) * inv;  // dL/d(un-normalised weight); the mask is already inside gBr / gBd
        if (dmlp_row && live) dmlp_row[k] = dwhat * e;
        float dt[3];
        dq_bone_dt(q, gBd, w, dt);
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = 0.f;
        v[9] = dt[0]; v[10] = dt[1]; v[11] = dt[2];
        v[12] = dwhat * sw * e * d2 * (2.0f * b.inv_2r2);
        const float r = transpose_reduce16(v, lane);
        if (!(lane & 1) && my_id < 13 && my_id >= 9) s_acc[k * 13 + my_id] += r;
    }
}

}  // namespace rg
