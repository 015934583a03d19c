// Per-Gaussian backward: 2-D gradients (mean2D, conic, opacity, colour, depth) -> gradients of the 3-D
// parameters.  One thread per Gaussian, ONE kernel.
//
// gauss_backward_kernel<false> (reference-shaped entry) replaces both computeCov2DCUDA
// (dgr/cuda_rasterizer/backward.cu:144-274) and the backward preprocessCUDA (:346-412) with their helpers
// (SH backward :20-139, cov3D backward :278-341), which the reference runs as two passes over all P.
// Every output row is written (zeros for culled Gaussians), so the caller hands in uninitialised memory; the
// reference's binding zero-fills ~300 B per Gaussian per step instead (dgr/rasterize_points.cu:158-167).
//
// gauss_backward_kernel<true> (fused entry) continues in registers through the activations
// (exp / normalize / sigmoid, gaussian_renderer/__init__.py:74-92) and through the skinning
// (SURVEY.md appendix A.2): it writes gradients of the RAW model tensors (_xyz, _scaling, _rotation, _opacity,
// _features_dc, _features_rest) and reduces the per-bone gradients dL/dG (3x4), dL/d_node_radius and
// dL/dglobal_trans -- warp transpose-reduce, shared-memory accumulators per CTA, one atomicAdd per value per CTA.
//
// HBM-bound: reads 44 + 12*M + 40 (+12 kept means3D) bytes, writes 88 + 12*M bytes per Gaussian.
#include "bones.cuh"
#include "bulk_copy.cuh"
#include "common.cuh"
#include "dq.cuh"
#include "kernels.cuh"
#include "math.cuh"
#include "sh.cuh"

namespace rg {

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

// DQ (fused entry only): the skinning backward of the dual-quaternion blend (dq.cuh) instead of the linear one.
template <bool FUSED, bool DQ = false>
__global__ void __launch_bounds__(256, 3) gauss_backward_kernel(GaussBwdArgs a, SkinFuseBwdArgs s) {
    const int idx = a.row_begin + blockIdx.x * blockDim.x + threadIdx.x;
    const int row_end = a.row_end > 0 ? a.row_end : a.P;
    const int lane = threadIdx.x & 31;
    __shared__ float s_mat[35];
    __shared__ __align__(8) uint64_t s_bar;
    extern __shared__ __align__(16) unsigned char s_raw[];
    if (threadIdx.x < 16) s_mat[threadIdx.x] = a.viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_mat[threadIdx.x] = a.projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_mat[threadIdx.x] = a.campos[threadIdx.x - 32];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    const int K = FUSED ? s.f.J - 1 : 0;
    // per-warp accumulator rows (K*13 + 3 floats each): a warp owns its row, so no shared-memory atomics are needed
    const int acc_stride = K * 13 + 3;
    float* s_acc_all = reinterpret_cast<float*>(s_raw + sizeof(BoneSm) * (size_t)(K > 1 ? K : 1));
    float* s_acc = s_acc_all + (threadIdx.x >> 5) * acc_stride;
    if (FUSED) {
        stage_bones(s.f.J, s.f.joints, s.f.parents, s.f.node_radius_raw, s.f.G, s.f.node_rot, s_b);
        for (int k = threadIdx.x; k < 8 * acc_stride; k += blockDim.x) s_acc_all[k] = 0.f;
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    }
    __syncthreads();
    // Fused variant: the CTA's block of features_rest (256 rows, contiguous) comes in through one bulk (TMA) copy and
    // the block of d_features_rest leaves the same way; per-thread accesses would have a 180-byte stride.
    float* s_sh = nullptr;
    size_t sh_row = 0, sh_first = 0;
    uint32_t sh_bytes = 0;
    bool sh_bulk = false;
    if (FUSED) {
        s_sh = reinterpret_cast<float*>(s_raw + align_up_dev(sizeof(BoneSm) * (size_t)(K > 1 ? K : 1) +
                                                             sizeof(float) * 8 * (size_t)acc_stride, 128));
        sh_row = (size_t)s.f.M_rest * 3;
        sh_first = (size_t)a.row_begin + (size_t)blockIdx.x * blockDim.x;
        const int n_valid = max(0, min((int)blockDim.x, row_end - (int)sh_first));
        const float* src = s.f.features_rest + sh_first * sh_row;
        sh_bytes = (uint32_t)(n_valid * sh_row * sizeof(float));
        sh_bulk = (sh_bytes % 16 == 0) && ((reinterpret_cast<size_t>(src) & 15) == 0) && sh_bytes > 0;
        if (sh_bulk) {
            if (threadIdx.x == 0) bulk_load(s_sh, src, sh_bytes, &s_bar);
        } else {
            for (int k = threadIdx.x; k < (int)(sh_bytes / 4); k += blockDim.x) s_sh[k] = __ldg(src + k);
            __syncthreads();
        }
    }
    const float* view = s_mat;
    const float* proj = s_mat + 16;
    const size_t i = (size_t)idx;
    const bool in_range = idx < row_end;
    const bool visible = in_range && a.radii[idx] > 0;
    if (!FUSED && !in_range) return;

    float3 dL_dmean = {0, 0, 0};     // dL/dmeans3D
    float3 ds = {0, 0, 0};           // dL/dscale (activated)
    float4 dq = {0, 0, 0, 0};        // dL/drotation (as given to the rasterizer, i.e. normalised in the fused path)
    float g_opacity = 0.f;
    float2 g_mean2D = {0, 0};
    float dcov[6] = {0, 0, 0, 0, 0, 0};
    float3 g_color = {0, 0, 0};
    float g_depth = 0.f;
    float3 dL_dconic = {0, 0, 0};
    float3 sc = {0, 0, 0};
    float4 q = {0, 0, 0, 0};
    float4 qraw = {0, 0, 0, 0};
    float qnorm = 1.f, opacity_act = 0.f;
    float basis[16], gsh[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 16; k++) basis[k] = 0.f;

    if (visible) {
        const float* acc = a.grad_acc + i * 10;
        g_mean2D = {acc[0], acc[1]};
        dL_dconic = {acc[2], acc[3], acc[4]};
        g_opacity = acc[5];
        g_color = {acc[6], acc[7], acc[8]};
        g_depth = acc[9];
        const float3 mean = {a.means3D[3 * i], a.means3D[3 * i + 1], a.means3D[3 * i + 2]};
        float cov3D[6];
        if (FUSED) {
            // re-derive the activated scale / rotation exactly as the forward did
            qraw = reinterpret_cast<const float4*>(s.f.rotation_raw)[i];
            if (!s.f.zero_d_rotation) {
                const float4 dr = reinterpret_cast<const float4*>(s.d_rot_saved)[i];
                qraw.x += dr.x; qraw.y += dr.y; qraw.z += dr.z; qraw.w += dr.w;
            }
            qnorm = fmaxf(sqrtf(qraw.x * qraw.x + qraw.y * qraw.y + qraw.z * qraw.z + qraw.w * qraw.w), 1e-12f);
            q = {qraw.x / qnorm, qraw.y / qnorm, qraw.z / qnorm, qraw.w / qnorm};
            if (s.f.isotropic) { const float e = expf(s.f.scaling_raw[i]); sc = {e, e, e}; }
            else sc = {expf(s.f.scaling_raw[3 * i]), expf(s.f.scaling_raw[3 * i + 1]), expf(s.f.scaling_raw[3 * i + 2])};
            opacity_act = 1.0f / (1.0f + expf(-s.f.opacity_raw[i]));
            cov3d_from_scale_rot(sc, a.scale_modifier, q, cov3D);
        } else {
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
        }

        // ---- conic -> cov2D -> cov3D and the covariance path of the mean gradient (backward.cu:162-273)
        Cov2D c2 = cov2d_project(mean, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, cov3D, view);
        const Mat3& T = c2.T;
        const float3 t = c2.t;
        const float ca = c2.cov.x, cb = c2.cov.y, cc = c2.cov.z;
        const float denom = ca * cc - cb * cb;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
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
        dL_dmean = {view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz,
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
        if (FUSED || a.shs) {
            float sh[48];
            if (FUSED) {
                if (sh_bulk) mbar_wait(&s_bar, 0);
                load_sh_staged(s.f.features_dc, s_sh + (size_t)threadIdx.x * sh_row, i, s.f.M_rest, sh);
            }
            else load_sh<16>(a.shs, i, a.M, a.sh_vec_ok, sh);
            const float3 campos = {s_mat[32], s_mat[33], s_mat[34]};
            const float3 dm = sh_backward_core(a.D, mean, campos, sh, a.clamped[idx], g_color, basis, gsh);
            dL_dmean.x += dm.x;
            dL_dmean.y += dm.y;
            dL_dmean.z += dm.z;
        }
        // ---- scale / rotation path
        if (FUSED || a.dL_dscale) cov3d_backward(sc, a.scale_modifier, q, dcov, ds, dq);
    }

    if (!FUSED) {
        // ---- outputs of the reference-shaped entry: every row is written
        a.dL_dmean2D[3 * i] = g_mean2D.x;
        a.dL_dmean2D[3 * i + 1] = g_mean2D.y;
        a.dL_dmean2D[3 * i + 2] = 0.f;
        reinterpret_cast<float4*>(a.dL_dconic)[i] = {dL_dconic.x, dL_dconic.y, 0.f, dL_dconic.z};
        a.dL_dopacity[i] = g_opacity;
        a.dL_dcolor[3 * i] = g_color.x;
        a.dL_dcolor[3 * i + 1] = g_color.y;
        a.dL_dcolor[3 * i + 2] = g_color.z;
        a.dL_ddepth[i] = g_depth;
        a.dL_dmean3D[3 * i] = dL_dmean.x;
        a.dL_dmean3D[3 * i + 1] = dL_dmean.y;
        a.dL_dmean3D[3 * i + 2] = dL_dmean.z;
#pragma unroll
        for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * i + k] = dcov[k];
        if (a.dL_dsh) {
            float* row = a.dL_dsh + i * a.M * 3;
            if (a.sh_vec_ok && a.M == 16) {
                float out[48];
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    out[3 * k] = basis[k] * gsh[0]; out[3 * k + 1] = basis[k] * gsh[1]; out[3 * k + 2] = basis[k] * gsh[2];
                }
#pragma unroll
                for (int k = 0; k < 12; k++)
                    reinterpret_cast<float4*>(row)[k] = {out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]};
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    if (k < a.M) {
                        row[3 * k] = basis[k] * gsh[0]; row[3 * k + 1] = basis[k] * gsh[1]; row[3 * k + 2] = basis[k] * gsh[2];
                    }
                }
            }
        }
        if (a.dL_dscale) {
            a.dL_dscale[3 * i] = ds.x;
            a.dL_dscale[3 * i + 1] = ds.y;
            a.dL_dscale[3 * i + 2] = ds.z;
            reinterpret_cast<float4*>(a.dL_drot)[i] = dq;
        }
        return;
    }

    // ================= fused entry: activations + skinning backward =================
    // d(normalize): n = q / |q|  ->  dq_raw = (g - n (n.g)) / |q|
    const float ndg = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;
    const float4 drot_raw = {(dq.x - q.x * ndg) / qnorm, (dq.y - q.y * ndg) / qnorm, (dq.z - q.z * ndg) / qnorm,
                             (dq.w - q.w * ndg) / qnorm};
    if (in_range) {
        a.dL_dmean2D[3 * i] = g_mean2D.x; a.dL_dmean2D[3 * i + 1] = g_mean2D.y; a.dL_dmean2D[3 * i + 2] = 0.f;  // per view
        const float dsc0 = ds.x * sc.x, dsc1 = ds.y * sc.y, dsc2 = ds.z * sc.z;
        const float dop = g_opacity * opacity_act * (1.f - opacity_act);
        if (!s.accumulate) {
            s.d_xyz[3 * i] = dL_dmean.x; s.d_xyz[3 * i + 1] = dL_dmean.y; s.d_xyz[3 * i + 2] = dL_dmean.z;
            if (s.f.isotropic) s.d_scaling_raw[i] = dsc0 + dsc1 + dsc2;
            else { s.d_scaling_raw[3 * i] = dsc0; s.d_scaling_raw[3 * i + 1] = dsc1; s.d_scaling_raw[3 * i + 2] = dsc2; }
            reinterpret_cast<float4*>(s.d_rotation_raw)[i] = drot_raw;
            s.d_opacity_raw[i] = dop;
            s.d_features_dc[3 * i] = basis[0] * gsh[0]; s.d_features_dc[3 * i + 1] = basis[0] * gsh[1]; s.d_features_dc[3 * i + 2] = basis[0] * gsh[2];
        } else if (visible) {  // a row belongs to one thread: plain read-modify-write
            s.d_xyz[3 * i] += dL_dmean.x; s.d_xyz[3 * i + 1] += dL_dmean.y; s.d_xyz[3 * i + 2] += dL_dmean.z;
            if (s.f.isotropic) s.d_scaling_raw[i] += dsc0 + dsc1 + dsc2;
            else { s.d_scaling_raw[3 * i] += dsc0; s.d_scaling_raw[3 * i + 1] += dsc1; s.d_scaling_raw[3 * i + 2] += dsc2; }
            float4 r = reinterpret_cast<float4*>(s.d_rotation_raw)[i];
            r.x += drot_raw.x; r.y += drot_raw.y; r.z += drot_raw.z; r.w += drot_raw.w;
            reinterpret_cast<float4*>(s.d_rotation_raw)[i] = r;
            s.d_opacity_raw[i] += dop;
            s.d_features_dc[3 * i] += basis[0] * gsh[0]; s.d_features_dc[3 * i + 1] += basis[0] * gsh[1]; s.d_features_dc[3 * i + 2] += basis[0] * gsh[2];
        }
    }
    // d_features_rest: every thread drops its row into the staging block (all SH rows have been read into registers
    // by now), then the block leaves the SM as one bulk store (or a cooperative coalesced copy for a ragged tail).
    if (sh_bulk) mbar_wait(&s_bar, 0);  // threads of culled Gaussians never waited: the load must have landed first
    __syncthreads();
    if (in_range) {
        float* rest = s_sh + (size_t)threadIdx.x * sh_row;
#pragma unroll
        for (int k = 1; k < 16; k++) {
            if (k <= s.f.M_rest) {
                rest[3 * (k - 1)] = basis[k] * gsh[0]; rest[3 * (k - 1) + 1] = basis[k] * gsh[1]; rest[3 * (k - 1) + 2] = basis[k] * gsh[2];
            }
        }
    }
    __syncthreads();
    {
        float* dst = s.d_features_rest + sh_first * sh_row;
        const bool st_bulk = (sh_bytes % 16 == 0) && ((reinterpret_cast<size_t>(dst) & 15) == 0) && sh_bytes > 0;
        if (st_bulk) {
            if (threadIdx.x == 0) {
                if (s.accumulate) bulk_reduce_add_f32(dst, s_sh, sh_bytes);
                else bulk_store(dst, s_sh, sh_bytes);
            }
        } else if (s.accumulate) {
            for (int k = threadIdx.x; k < (int)(sh_bytes / 4); k += blockDim.x) dst[k] += s_sh[k];
        } else {
            for (int k = threadIdx.x; k < (int)(sh_bytes / 4); k += blockDim.x) dst[k] = s_sh[k];
        }
    }

    // ---- skinning backward (appendix A.2): a = dL/dmeans3D, h = dL/d(d_rotation)
    float x = 0, y = 0, z = 0, m = 0;
    if (in_range) {
        x = s.f.xyz[3 * i]; y = s.f.xyz[3 * i + 1]; z = s.f.xyz[3 * i + 2];
        m = s.f.motion_mask ? s.f.motion_mask[i] : 1.0f;
    }
    const float gx = dL_dmean.x, gy = dL_dmean.y, gz = dL_dmean.z;
    const float h0 = s.f.zero_d_rotation ? 0.f : drot_raw.x, h1 = s.f.zero_d_rotation ? 0.f : drot_raw.y,
                h2 = s.f.zero_d_rotation ? 0.f : drot_raw.z, h3 = s.f.zero_d_rotation ? 0.f : drot_raw.w;
    const bool contributes = visible && (gx != 0.f || gy != 0.f || gz != 0.f || h0 != 0.f || h1 != 0.f || h2 != 0.f || h3 != 0.f);
    const bool want_elem = s.dmask || s.doffsets || s.dmlp_weight;
    if (!__any_sync(0xffffffffu, contributes)) {
        if (in_range && want_elem) {
            if (s.dmask) s.dmask[i] = 0.f;
            if (s.doffsets) { s.doffsets[3 * i] = 0.f; s.doffsets[3 * i + 1] = 0.f; s.doffsets[3 * i + 2] = 0.f; }
            if (s.dmlp_weight) for (int k = 0; k < K; k++) s.dmlp_weight[i * K + k] = 0.f;
        }
    } else if (DQ) {
        float ox = 0.f, oy = 0.f, oz = 0.f;
        if (s.f.offsets && in_range) { ox = s.f.offsets[3 * i]; oy = s.f.offsets[3 * i + 1]; oz = s.f.offsets[3 * i + 2]; }
        dq_skin_backward_point(s_b, K, lane, in_range, x, y, z, m, gx, gy, gz, h0, h1, h2, h3,
                               s.f.mlp_weight ? s.f.mlp_weight + (in_range ? i : 0) * K : nullptr, s.f.global_trans[0],
                               s.f.global_trans[1], s.f.global_trans[2], ox, oy, oz, s_acc,
                               (s.dmask && in_range) ? s.dmask + i : nullptr,
                               (s.dmlp_weight && in_range) ? s.dmlp_weight + i * K : nullptr);
        if (in_range && s.doffsets) { s.doffsets[3 * i] = m * gx; s.doffsets[3 * i + 1] = m * gy; s.doffsets[3 * i + 2] = m * gz; }
    } else {
        // No first sweep over the bones: the forward kept the normaliser, and m * sum_k w_k c_k follows from what it kept too --
        //   m * sum_k w_k y_k = (p - x) - m (g + o - x)   (p = the skinned position),   m * sum_k w_k q_k = d_rot_saved.
        float inv = 1.f, mcbar = 0.f;
        if (in_range) {
            inv = 1.0f / s.wsum_saved[i];
            float ox = 0.f, oy = 0.f, oz = 0.f;
            if (s.f.offsets) { ox = s.f.offsets[3 * i]; oy = s.f.offsets[3 * i + 1]; oz = s.f.offsets[3 * i + 2]; }
            const float sx = (a.means3D[3 * i] - x) - m * (s.f.global_trans[0] + ox - x);
            const float sy = (a.means3D[3 * i + 1] - y) - m * (s.f.global_trans[1] + oy - y);
            const float sz = (a.means3D[3 * i + 2] - z) - m * (s.f.global_trans[2] + oz - z);
            mcbar = gx * sx + gy * sy + gz * sz;
            if (!s.f.zero_d_rotation) {
                const float4 dr = reinterpret_cast<const float4*>(s.d_rot_saved)[i];
                mcbar += h0 * dr.x + h1 * dr.y + h2 * dr.z + h3 * dr.w;
            }
            if (s.doffsets) { s.doffsets[3 * i] = m * gx; s.doffsets[3 * i + 1] = m * gy; s.doffsets[3 * i + 2] = m * gz; }
        }
        const bool want_mask = s.dmask != nullptr;  // dL/dmask needs the un-masked blend (the kept one is lost where m = 0)
        float ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
        const float mgx = m * gx, mgy = m * gy, mgz = m * gz;
        {
            const float s0 = warp_sum(mgx), s1 = warp_sum(mgy), s2 = warp_sum(mgz);
            if (lane == 0) { s_acc[K * 13] += s0; s_acc[K * 13 + 1] += s1; s_acc[K * 13 + 2] += s2; }
        }
        const int my_id = (lane >> 1) & 15;
        const bool writer = !(lane & 1) && my_id < 13;
        // sweep 2: per-bone terms, transpose-reduced over the warp
        for (int k = 0; k < K; k++) {
            const BoneSm& b = s_b[k];
            const float d2 = seg_dist2(b, x, y, z);
            const float e = expf(-d2 * b.inv_2r2);
            const float sw = (s.f.mlp_weight && in_range) ? s.f.mlp_weight[i * K + k] : 1.0f;
            const float w = (e * sw + 1e-7f) * inv;
            const float yx = b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9];
            const float yy = b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10];
            const float yz = b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11];
            const float c = gx * yx + gy * yy + gz * yz + h0 * b.q[0] + h1 * b.q[1] + h2 * b.q[2] + h3 * b.q[3];
            const float dwhat = (m * c - mcbar) * inv;
            if (want_mask) {
                ax += w * yx; ay += w * yy; az += w * yz;
                q0 += w * b.q[0]; q1 += w * b.q[1]; q2 += w * b.q[2]; q3 += w * b.q[3];
            }
            if (s.dmlp_weight && in_range) s.dmlp_weight[i * K + k] = dwhat * e;
            float v[16];
            v[0] = mgx * w * x; v[1] = mgx * w * y; v[2] = mgx * w * z;
            v[3] = mgy * w * x; v[4] = mgy * w * y; v[5] = mgy * w * z;
            v[6] = mgz * w * x; v[7] = mgz * w * y; v[8] = mgz * w * z;
            v[9] = mgx * w; v[10] = mgy * w; v[11] = mgz * w;
            v[12] = dwhat * sw * e * d2 * (2.0f * b.inv_2r2);
            v[13] = v[14] = v[15] = 0.f;
            const float r = transpose_reduce16(v, lane);
            if (writer) s_acc[k * 13 + my_id] += r;
        }
        if (want_mask && in_range) {
            float ox = 0.f, oy = 0.f, oz = 0.f;
            if (s.f.offsets) { ox = s.f.offsets[3 * i]; oy = s.f.offsets[3 * i + 1]; oz = s.f.offsets[3 * i + 2]; }
            s.dmask[i] = gx * (ax + s.f.global_trans[0] + ox - x) + gy * (ay + s.f.global_trans[1] + oy - y) +
                         gz * (az + s.f.global_trans[2] + oz - z) + h0 * q0 + h1 * q1 + h2 * q2 + h3 * q3;
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K * 13; k += blockDim.x) {
        float val = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) val += s_acc_all[w * acc_stride + k];
        if (val != 0.f) {
            const int bone = k / 13, c = k % 13;
            if (c < 12) atomicAdd(&s.dG[12 * (bone + 1) + c], val);
            else atomicAdd(&s.dnode_radius_raw[bone + 1], val);
        }
    }
    if (threadIdx.x == 0) bulk_store_wait_read();  // shared memory must outlive the bulk store's reads
    if (threadIdx.x < 3) {
        float val = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) val += s_acc_all[w * acc_stride + K * 13 + threadIdx.x];
        if (val != 0.f) atomicAdd(&s.dglobal_trans[threadIdx.x], val);
    }
}

int launch_gauss_backward(const GaussBwdArgs& a, cudaStream_t stream) {
    if (a.P == 0) return 0;
    LaunchScope ls("gauss_backward", stream);
    SkinFuseBwdArgs none = {};
    gauss_backward_kernel<false><<<(a.P + 255) / 256, 256, 0, stream>>>(a, none);
    return 0;
}

int launch_gauss_skin_backward(const GaussBwdArgs& a, const SkinFuseBwdArgs& s, cudaStream_t stream) {
    if (a.P == 0) return 0;
    if (s.f.J > kMaxJoints || s.f.J < 2) { set_error("skinned backward: J must be in [2, 128]"); return -1; }
    LaunchScope ls("gauss_skin_backward", stream);
    const size_t smem = align_up(bone_smem_bytes(s.f.J, false) + 8 * sizeof(float) * ((size_t)(s.f.J - 1) * 13 + 3), 128) +
                        (size_t)256 * s.f.M_rest * 3 * sizeof(float) + 16;
    if (ensure_dyn_smem(reinterpret_cast<const void*>(gauss_backward_kernel<true>), 160 * 1024)) return -1;
    const int rows = (a.row_end > 0 ? a.row_end : a.P) - a.row_begin;
    if (rows <= 0) return 0;
    if (s.f.blend_mode == 1) {
        if (ensure_dyn_smem(reinterpret_cast<const void*>(gauss_backward_kernel<true, true>), 160 * 1024)) return -1;
        gauss_backward_kernel<true, true><<<(rows + 255) / 256, 256, smem, stream>>>(a, s);
    } else {
        gauss_backward_kernel<true><<<(rows + 255) / 256, 256, smem, stream>>>(a, s);
    }
    return 0;
}

}  // namespace rg
