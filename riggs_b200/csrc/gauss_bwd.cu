// Per-Gaussian backward: 2-D gradients (mean2D, conic, opacity, colour, depth) -> gradients of means3D,
// cov3D, SH, scale, rotation.  One thread per Gaussian, ONE kernel: it replaces both computeCov2DCUDA
// (dgr/cuda_rasterizer/backward.cu:144-274) and the backward preprocessCUDA (:346-412) with their helpers
// (SH backward :20-139, cov3D backward :278-341), which the reference runs as two passes over all P.
//
// Every output row is written (zeros for culled Gaussians), so the caller hands in uninitialised memory;
// the reference's binding zero-fills ~300 B per Gaussian per step instead (dgr/rasterize_points.cu:158-167).
// HBM-bound: reads 44 + 12*M + 40 (+24 with a kept covariance) bytes, writes 88 + 12*M bytes per Gaussian.
#include "common.cuh"
#include "kernels.cuh"
#include "math.cuh"

namespace rg {

__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {  // auxiliary.h:107-117
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward (backward.cu:20-139): writes dL_dsh rows 0..(deg+1)^2-1 and returns the mean gradient that
// flows through the view direction.
template <int MAXC>
__device__ __forceinline__ float3 sh_backward(int deg, int M, const float3 pos, const float3 campos, const float* sh,
                                              uint8_t clamp_bits, float3 dL_dcolor, float* __restrict__ dL_dsh_row,
                                              bool vec_ok) {
    const float3 dir_orig = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    const float g[3] = {dL_dcolor.x * ((clamp_bits & 1) ? 0.f : 1.f), dL_dcolor.y * ((clamp_bits & 2) ? 0.f : 1.f),
                        dL_dcolor.z * ((clamp_bits & 4) ? 0.f : 1.f)};
    float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
    float basis[MAXC];
#pragma unroll
    for (int k = 0; k < MAXC; k++) basis[k] = 0.f;
    basis[0] = kSH_C0;
#define SH(k, c) sh[(k) * 3 + (c)]
    if (deg > 0) {
        basis[1] = -kSH_C1 * y;
        basis[2] = kSH_C1 * z;
        basis[3] = -kSH_C1 * x;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            dx[c] = -kSH_C1 * SH(3, c);
            dy[c] = -kSH_C1 * SH(1, c);
            dz[c] = kSH_C1 * SH(2, c);
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            basis[4] = kSH_C2[0] * xy;
            basis[5] = kSH_C2[1] * yz;
            basis[6] = kSH_C2[2] * (2.f * zz - xx - yy);
            basis[7] = kSH_C2[3] * xz;
            basis[8] = kSH_C2[4] * (xx - yy);
#pragma unroll
            for (int c = 0; c < 3; c++) {
                dx[c] += kSH_C2[0] * y * SH(4, c) + kSH_C2[2] * 2.f * -x * SH(6, c) + kSH_C2[3] * z * SH(7, c) +
                         kSH_C2[4] * 2.f * x * SH(8, c);
                dy[c] += kSH_C2[0] * x * SH(4, c) + kSH_C2[1] * z * SH(5, c) + kSH_C2[2] * 2.f * -y * SH(6, c) +
                         kSH_C2[4] * 2.f * -y * SH(8, c);
                dz[c] += kSH_C2[1] * y * SH(5, c) + kSH_C2[2] * 2.f * 2.f * z * SH(6, c) + kSH_C2[3] * x * SH(7, c);
            }
            if (deg > 2) {
                basis[9] = kSH_C3[0] * y * (3.f * xx - yy);
                basis[10] = kSH_C3[1] * xy * z;
                basis[11] = kSH_C3[2] * y * (4.f * zz - xx - yy);
                basis[12] = kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                basis[13] = kSH_C3[4] * x * (4.f * zz - xx - yy);
                basis[14] = kSH_C3[5] * z * (xx - yy);
                basis[15] = kSH_C3[6] * x * (xx - 3.f * yy);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    dx[c] += (kSH_C3[0] * SH(9, c) * 3.f * 2.f * xy + kSH_C3[1] * SH(10, c) * yz +
                              kSH_C3[2] * SH(11, c) * -2.f * xy + kSH_C3[3] * SH(12, c) * -3.f * 2.f * xz +
                              kSH_C3[4] * SH(13, c) * (-3.f * xx + 4.f * zz - yy) + kSH_C3[5] * SH(14, c) * 2.f * xz +
                              kSH_C3[6] * SH(15, c) * 3.f * (xx - yy));
                    dy[c] += (kSH_C3[0] * SH(9, c) * 3.f * (xx - yy) + kSH_C3[1] * SH(10, c) * xz +
                              kSH_C3[2] * SH(11, c) * (-3.f * yy + 4.f * zz - xx) + kSH_C3[3] * SH(12, c) * -3.f * 2.f * yz +
                              kSH_C3[4] * SH(13, c) * -2.f * xy + kSH_C3[5] * SH(14, c) * -2.f * yz +
                              kSH_C3[6] * SH(15, c) * -3.f * 2.f * xy);
                    dz[c] += (kSH_C3[1] * SH(10, c) * xy + kSH_C3[2] * SH(11, c) * 4.f * 2.f * yz +
                              kSH_C3[3] * SH(12, c) * 3.f * (2.f * zz - xx - yy) + kSH_C3[4] * SH(13, c) * 4.f * 2.f * xz +
                              kSH_C3[5] * SH(14, c) * (xx - yy));
                }
            }
        }
    }
#undef SH
    // dL_dsh[k] = basis_k * dL_dRGB; rows above the active degree stay zero.
    if (vec_ok && M == MAXC) {
        float out[MAXC * 3];
#pragma unroll
        for (int k = 0; k < MAXC; k++) {
            out[3 * k] = basis[k] * g[0];
            out[3 * k + 1] = basis[k] * g[1];
            out[3 * k + 2] = basis[k] * g[2];
        }
        float4* o4 = reinterpret_cast<float4*>(dL_dsh_row);
#pragma unroll
        for (int i = 0; i < MAXC * 3 / 4; i++) o4[i] = {out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]};
    } else {
#pragma unroll
        for (int k = 0; k < MAXC; k++) {
            if (k < M) {
                dL_dsh_row[3 * k] = basis[k] * g[0];
                dL_dsh_row[3 * k + 1] = basis[k] * g[1];
                dL_dsh_row[3 * k + 2] = basis[k] * g[2];
            }
        }
    }
    const float3 dL_ddir = {dx[0] * g[0] + dx[1] * g[1] + dx[2] * g[2], dy[0] * g[0] + dy[1] * g[1] + dy[2] * g[2],
                            dz[0] * g[0] + dz[1] * g[1] + dz[2] * g[2]};
    return dnormvdv(dir_orig, dL_ddir);
}

// cov3D backward (backward.cu:278-341): gradient w.r.t. scale and the UN-normalised quaternion.
__device__ __forceinline__ void cov3d_backward(const float3 scale, float mod, const float4 rot, const float* dc,
                                               float3& dL_dscale, float4& dL_drot) {
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    Mat3 R = quat_rot(r, x, y, z);
    const float3 s = {mod * scale.x, mod * scale.y, mod * scale.z};
    Mat3 S = mat3_cols(s.x, 0.f, 0.f, 0.f, s.y, 0.f, 0.f, 0.f, s.z);
    Mat3 M = mul(S, R);
    Mat3 dL_dSigma = mat3_cols(dc[0], 0.5f * dc[1], 0.5f * dc[2], 0.5f * dc[1], dc[3], 0.5f * dc[4], 0.5f * dc[2],
                               0.5f * dc[4], dc[5]);
    Mat3 M2;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) M2.m[c][k] = 2.0f * M.m[c][k];
    Mat3 dL_dM = mul(M2, dL_dSigma);
    Mat3 Rt = transpose(R);
    Mat3 dL_dMt = transpose(dL_dM);
    dL_dscale.x = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
    dL_dscale.y = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
    dL_dscale.z = Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] + Rt.m[2][2] * dL_dMt.m[2][2];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        dL_dMt.m[0][k] *= s.x;
        dL_dMt.m[1][k] *= s.y;
        dL_dMt.m[2][k] *= s.z;
    }
#define DM(a_, b_) dL_dMt.m[a_][b_]
    dL_drot.x = 2 * z * (DM(0, 1) - DM(1, 0)) + 2 * y * (DM(2, 0) - DM(0, 2)) + 2 * x * (DM(1, 2) - DM(2, 1));
    dL_drot.y = 2 * y * (DM(1, 0) + DM(0, 1)) + 2 * z * (DM(2, 0) + DM(0, 2)) + 2 * r * (DM(1, 2) - DM(2, 1)) -
                4 * x * (DM(2, 2) + DM(1, 1));
    dL_drot.z = 2 * x * (DM(1, 0) + DM(0, 1)) + 2 * r * (DM(2, 0) - DM(0, 2)) + 2 * z * (DM(1, 2) + DM(2, 1)) -
                4 * y * (DM(2, 2) + DM(0, 0));
    dL_drot.w = 2 * r * (DM(0, 1) - DM(1, 0)) + 2 * x * (DM(2, 0) + DM(0, 2)) + 2 * y * (DM(1, 2) + DM(2, 1)) -
                4 * z * (DM(1, 1) + DM(0, 0));
#undef DM
}

__global__ void __launch_bounds__(256) gauss_backward_kernel(GaussBwdArgs a) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ float s_mat[35];
    if (threadIdx.x < 16) s_mat[threadIdx.x] = a.viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_mat[threadIdx.x] = a.projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_mat[threadIdx.x] = a.campos[threadIdx.x - 32];
    __syncthreads();
    if (idx >= a.P) return;
    const float* view = s_mat;
    const float* proj = s_mat + 16;
    const size_t i = (size_t)idx;

    if (!(a.radii[idx] > 0)) {  // culled: all gradients are zero
        a.dL_dmean2D[3 * i] = a.dL_dmean2D[3 * i + 1] = a.dL_dmean2D[3 * i + 2] = 0.f;
        reinterpret_cast<float4*>(a.dL_dconic)[i] = {0.f, 0.f, 0.f, 0.f};
        a.dL_dopacity[i] = 0.f;
        a.dL_dcolor[3 * i] = a.dL_dcolor[3 * i + 1] = a.dL_dcolor[3 * i + 2] = 0.f;
        a.dL_ddepth[i] = 0.f;
        a.dL_dmean3D[3 * i] = a.dL_dmean3D[3 * i + 1] = a.dL_dmean3D[3 * i + 2] = 0.f;
#pragma unroll
        for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * i + k] = 0.f;
        if (a.dL_dsh) {
            float* row = a.dL_dsh + i * a.M * 3;
            if (a.sh_vec_ok && a.M == 16) {
#pragma unroll
                for (int k = 0; k < 12; k++) reinterpret_cast<float4*>(row)[k] = {0.f, 0.f, 0.f, 0.f};
            } else {
                for (int k = 0; k < a.M * 3; k++) row[k] = 0.f;
            }
        }
        if (a.dL_dscale) {
            a.dL_dscale[3 * i] = a.dL_dscale[3 * i + 1] = a.dL_dscale[3 * i + 2] = 0.f;
            reinterpret_cast<float4*>(a.dL_drot)[i] = {0.f, 0.f, 0.f, 0.f};
        }
        return;
    }

    // 2-D gradients accumulated by the backward blend
    const float* acc = a.grad_acc + i * 10;
    const float2 g_mean2D = {acc[0], acc[1]};
    const float3 dL_dconic = {acc[2], acc[3], acc[4]};
    const float g_opacity = acc[5];
    const float3 g_color = {acc[6], acc[7], acc[8]};
    const float g_depth = acc[9];
    a.dL_dmean2D[3 * i] = g_mean2D.x;
    a.dL_dmean2D[3 * i + 1] = g_mean2D.y;
    a.dL_dmean2D[3 * i + 2] = 0.f;
    reinterpret_cast<float4*>(a.dL_dconic)[i] = {dL_dconic.x, dL_dconic.y, 0.f, dL_dconic.z};
    a.dL_dopacity[i] = g_opacity;
    a.dL_dcolor[3 * i] = g_color.x;
    a.dL_dcolor[3 * i + 1] = g_color.y;
    a.dL_dcolor[3 * i + 2] = g_color.z;
    a.dL_ddepth[i] = g_depth;

    const float3 mean = {a.means3D[3 * i], a.means3D[3 * i + 1], a.means3D[3 * i + 2]};
    float3 sc = {0, 0, 0};
    float4 q = {0, 0, 0, 0};
    float cov3D[6];
    if (a.scales) {
        sc = {a.scales[3 * i], a.scales[3 * i + 1], a.scales[3 * i + 2]};
        q = reinterpret_cast<const float4*>(a.rotations)[i];
    }
    if (a.cov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) cov3D[k] = a.cov3D[6 * i + k];
    } else {
        cov3d_from_scale_rot(sc, a.scale_modifier, q, cov3D);
    }

    // ---- conic -> cov2D -> cov3D and the covariance path of the mean gradient (backward.cu:162-273)
    Cov2D c2 = cov2d_project(mean, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, cov3D, view);
    const Mat3& T = c2.T;
    const float3 t = c2.t;
    const float ca = c2.cov.x, cb = c2.cov.y, cc = c2.cov.z;
    const float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6] = {0, 0, 0, 0, 0, 0};
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dL_dconic.x + 2 * cb * cc * dL_dconic.y + (denom - ca * cc) * dL_dconic.z);
        dL_dc = denom2inv * (-ca * ca * dL_dconic.z + 2 * ca * cb * dL_dconic.y + (denom - ca * cc) * dL_dconic.x);
        dL_db = denom2inv * 2 * (cb * cc * dL_dconic.x - (denom + 2 * cb * cb) * dL_dconic.y + ca * cb * dL_dconic.z);
        dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
        dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
        dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
        dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                  2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                  2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                  2 * T.m[1][1] * T.m[1][2] * dL_dc;
    }
#pragma unroll
    for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * i + k] = dcov[k];

    const float V[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
    float dT0[3], dT1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float r0 = T.m[0][0] * V[k][0] + T.m[0][1] * V[k][1] + T.m[0][2] * V[k][2];
        const float r1 = T.m[1][0] * V[k][0] + T.m[1][1] * V[k][1] + T.m[1][2] * V[k][2];
        dT0[k] = 2 * r0 * dL_da + r1 * dL_db;
        dT1[k] = 2 * r1 * dL_dc + r0 * dL_db;
    }
    // W[c][r] = view[4*r + c]
    const float dL_dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
    const float dL_dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
    const float dL_dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
    const float dL_dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;
    const float h_x = a.focal_x, h_y = a.focal_y;
    const float dL_dtx = c2.x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = c2.y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                         (2 * h_y * t.y) * tz3 * dL_dJ12;
    float3 dL_dmean = {view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz,
                       view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz,
                       view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz};

    // ---- projection and depth paths (backward.cu:372-403)
    const float4 m_hom = xform_point_4x4(mean, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    dL_dmean.x += (proj[0] * m_w - proj[3] * mul1) * g_mean2D.x + (proj[1] * m_w - proj[3] * mul2) * g_mean2D.y;
    dL_dmean.y += (proj[4] * m_w - proj[7] * mul1) * g_mean2D.x + (proj[5] * m_w - proj[7] * mul2) * g_mean2D.y;
    dL_dmean.z += (proj[8] * m_w - proj[11] * mul1) * g_mean2D.x + (proj[9] * m_w - proj[11] * mul2) * g_mean2D.y;
    const float mul3 = view[2] * mean.x + view[6] * mean.y + view[10] * mean.z + view[14];
    dL_dmean.x += (view[2] - view[3] * mul3) * g_depth;
    dL_dmean.y += (view[6] - view[7] * mul3) * g_depth;
    dL_dmean.z += (view[10] - view[11] * mul3) * g_depth;

    // ---- SH path
    if (a.shs) {
        float sh[48];
        const float* src = a.shs + i * a.M * 3;
        if (a.sh_vec_ok && a.M == 16) {
#pragma unroll
            for (int k = 0; k < 12; k++) {
                float4 v = __ldg(reinterpret_cast<const float4*>(src) + k);
                sh[4 * k] = v.x; sh[4 * k + 1] = v.y; sh[4 * k + 2] = v.z; sh[4 * k + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 48; k++) sh[k] = k < a.M * 3 ? __ldg(src + k) : 0.f;
        }
        const float3 campos = {s_mat[32], s_mat[33], s_mat[34]};
        float3 dm = sh_backward<16>(a.D, a.M, mean, campos, sh, a.clamped[idx], g_color, a.dL_dsh + i * a.M * 3,
                                    a.sh_vec_ok);
        dL_dmean.x += dm.x;
        dL_dmean.y += dm.y;
        dL_dmean.z += dm.z;
    }
    a.dL_dmean3D[3 * i] = dL_dmean.x;
    a.dL_dmean3D[3 * i + 1] = dL_dmean.y;
    a.dL_dmean3D[3 * i + 2] = dL_dmean.z;

    // ---- scale / rotation path
    if (a.dL_dscale) {
        float3 ds;
        float4 dq;
        cov3d_backward(sc, a.scale_modifier, q, dcov, ds, dq);
        a.dL_dscale[3 * i] = ds.x;
        a.dL_dscale[3 * i + 1] = ds.y;
        a.dL_dscale[3 * i + 2] = ds.z;
        reinterpret_cast<float4*>(a.dL_drot)[i] = dq;
    }
}

int launch_gauss_backward(const GaussBwdArgs& a, cudaStream_t stream) {
    if (a.P == 0) return 0;
    LaunchScope ls("gauss_backward", stream);
    gauss_backward_kernel<<<(a.P + 255) / 256, 256, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
