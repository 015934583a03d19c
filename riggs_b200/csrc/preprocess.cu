// Per-Gaussian forward stage: [skeleton skinning + parameter activations ->] frustum cull, projection, 3D->2D
// covariance, conic, radius, tile rectangle, SH->RGB.  One thread per Gaussian.
//
// preprocess_kernel<false> replaces preprocessCUDA (dgr/cuda_rasterizer/forward.cu:155-256) behind the
// reference-shaped entry point (activated means3D / scales / rotations / opacities / shs come from the caller).
//
// preprocess_kernel<true> is the fused entry: it starts from the RAW model parameters and folds in
//   * linear-blend skinning over the bone table held in shared memory
//     (SkeletonWarp.deform_by_pose, skeleton_utils/skeleton_warp.py:130-172),
//   * means3D = xyz + d_xyz, scales = exp(_scaling), rotations = normalize(_rotation + d_rotation),
//     opacity = sigmoid(_opacity), shs = cat(features_dc, features_rest)
//     (render(), gaussian_renderer/__init__.py:74-114; scene/gaussian_model.py:105-132),
// so the ~40 eager torch kernels of the skinning chain, the 6-8 activation kernels and the SH cat (192 B per
// Gaussian read + written every frame) disappear; nothing of size N is materialised except means3D (kept for
// the backward).
//
// HBM-bound streaming kernel: algorithmic bytes per Gaussian = 44 (xyz, rot, scale, opacity) + 12*M (SH)
// read, 48 written (+1 clamp byte, +12 means3D in the fused variant).
#include "bones.cuh"
#include "bulk_copy.cuh"
#include "common.cuh"
#include "cull.cuh"
#include "dq.cuh"
#include "kernels.cuh"
#include "math.cuh"
#include "sh.cuh"

namespace rg {

// DQ (fused entry only): the bone transforms are blended as dual quaternions (dq.cuh) instead of linearly.
template <bool FUSED, bool DQ = false>
__global__ void __launch_bounds__(256, 4) preprocess_kernel(PreprocessArgs a, SkinFuseArgs s) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ float s_mat[35];  // view[16], proj[16], campos[3]
    extern __shared__ __align__(16) unsigned char s_raw[];
    if (threadIdx.x < 16) s_mat[threadIdx.x] = a.viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_mat[threadIdx.x] = a.projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_mat[threadIdx.x] = a.cam_pos[threadIdx.x - 32];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    // Fused variant: the depth sort's four 8-bit digit histograms are counted here, while the key is in a register
    // (saves the sort its histogram launch and a read of the keys)
    __shared__ uint32_t s_hist[FUSED ? 4 * 256 : 1];
    const bool count_digits = FUSED && a.sort_hist != nullptr;
    if (count_digits)
        for (int k = threadIdx.x; k < 4 * 256; k += blockDim.x) s_hist[k] = 0;
    // Fused variant: this CTA's block of features_rest (256 rows of M_rest*12 B, contiguous in memory) is pulled into
    // shared memory by one bulk (TMA) copy that runs while the threads do the skinning loop below.
    __shared__ __align__(8) uint64_t s_bar;
    float* s_sh = nullptr;
    bool sh_bulk = false;
    if (FUSED) {
        stage_bones(s.J, s.joints, s.parents, s.node_radius_raw, s.G, s.node_rot, s_b);
        s_sh = reinterpret_cast<float*>(s_raw + align_up_dev(sizeof(BoneSm) * (size_t)(s.J > 1 ? s.J - 1 : 1), 128));
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    }
    __syncthreads();
    if (FUSED) {
        const size_t row = (size_t)s.M_rest * 3;
        const size_t first = (size_t)blockIdx.x * blockDim.x;
        const int n_valid = min((int)blockDim.x, a.P - (int)first);
        const float* src = s.features_rest + first * row;
        const uint32_t bytes = (uint32_t)(n_valid * row * sizeof(float));
        sh_bulk = (bytes % 16 == 0) && ((reinterpret_cast<size_t>(src) & 15) == 0) && bytes > 0;
        if (sh_bulk) {
            if (threadIdx.x == 0) bulk_load(s_sh, src, bytes, &s_bar);
        } else {  // ragged tail block: plain cooperative copy
            for (int k = threadIdx.x; k < n_valid * (int)row; k += blockDim.x) s_sh[k] = __ldg(src + k);
            __syncthreads();
        }
    }
    // The thread that started the bulk copy is thread 0 and it is always a valid row (first < P).  Threads past P own
    // no row and never touch s_sh; the CTA stays alive until thread 0 has seen the copy land (end of the kernel).
    const bool has_row = idx < a.P;
    if (!has_row && !count_digits) return;
    if (has_row) {
    const float* view = s_mat;
    const float* proj = s_mat + 16;
    const size_t i = (size_t)idx;

    float3 p_orig, sc = {0, 0, 0};
    float4 q = {0, 0, 0, 0};
    float opacity;
    if (FUSED) {
        // ---- skinning (A.1 of SURVEY.md): weights recomputed in registers, bones from shared memory
        const float x = s.xyz[3 * i], y = s.xyz[3 * i + 1], z = s.xyz[3 * i + 2];
        const int K = s.J - 1;
        float wsum = 0.f, ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
        float Br[4] = {0, 0, 0, 0}, Bd[4] = {0, 0, 0, 0};
        for (int k = 0; k < K; k++) {
            const BoneSm& b = s_b[k];
            const float d2 = seg_dist2(b, x, y, z);
            float w = expf(-d2 * b.inv_2r2);
            if (s.mlp_weight) w *= s.mlp_weight[i * K + k];
            w += 1e-7f;
            wsum += w;
            if (!DQ) {
                ax += w * (b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9]);
                ay += w * (b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10]);
                az += w * (b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11]);
                q0 += w * b.q[0]; q1 += w * b.q[1]; q2 += w * b.q[2]; q3 += w * b.q[3];
            } else {
                const DqBone qb = dq_of_bone(b);
#pragma unroll
                for (int c = 0; c < 4; c++) { Br[c] += w * qb.r[c]; Bd[c] += w * qb.d[c]; }
            }
        }
        float inv = 1.0f / wsum;
        s.wsum_out[i] = wsum;
        if (DQ) {  // the blended rigid transform applied to x; its unit quaternion is the rotation output
#pragma unroll
            for (int c = 0; c < 4; c++) { Br[c] *= inv; Bd[c] *= inv; }
            const DqBlend f = dq_finish(Br, Bd, x, y, z);
            ax = f.y[0]; ay = f.y[1]; az = f.y[2];
            q0 = f.r0[0]; q1 = f.r0[1]; q2 = f.r0[2]; q3 = f.r0[3];
            inv = 1.0f;
        }
        const float m = s.motion_mask ? s.motion_mask[i] : 1.0f;
        float ox = 0.f, oy = 0.f, oz = 0.f;
        if (s.offsets) { ox = s.offsets[3 * i]; oy = s.offsets[3 * i + 1]; oz = s.offsets[3 * i + 2]; }
        // d_xyz is rounded before it is added to x (no FMA across the two): the unfused path stores d_xyz as a
        // tensor and adds it in a separate op (gaussian_renderer/__init__.py:74), and the positions must be the same
        const float dx = __fmul_rn(ax * inv + s.global_trans[0] + ox - x, m);
        const float dy = __fmul_rn(ay * inv + s.global_trans[1] + oy - y, m);
        const float dz = __fmul_rn(az * inv + s.global_trans[2] + oz - z, m);
        p_orig = {__fadd_rn(x, dx), __fadd_rn(y, dy), __fadd_rn(z, dz)};
        s.means3D_out[3 * i] = p_orig.x; s.means3D_out[3 * i + 1] = p_orig.y; s.means3D_out[3 * i + 2] = p_orig.z;
        // ---- activations
        float4 r = reinterpret_cast<const float4*>(s.rotation_raw)[i];
        if (!s.zero_d_rotation) {
            const float4 dr = {q0 * inv * m, q1 * inv * m, q2 * inv * m, q3 * inv * m};
            reinterpret_cast<float4*>(s.d_rot_out)[i] = dr;
            r.x += dr.x; r.y += dr.y; r.z += dr.z; r.w += dr.w;
        }
        const float rn = fmaxf(sqrtf(r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w), 1e-12f);  // F.normalize eps
        q = {r.x / rn, r.y / rn, r.z / rn, r.w / rn};
        if (s.isotropic) { const float e = expf(s.scaling_raw[i]); sc = {e, e, e}; }
        else sc = {expf(s.scaling_raw[3 * i]), expf(s.scaling_raw[3 * i + 1]), expf(s.scaling_raw[3 * i + 2])};
        opacity = 1.0f / (1.0f + expf(-s.opacity_raw[i]));
        if (s.dbg_scales) {  // test aid: the activated inputs the reference rasterizer would have been given
            s.dbg_scales[3 * i] = sc.x; s.dbg_scales[3 * i + 1] = sc.y; s.dbg_scales[3 * i + 2] = sc.z;
            reinterpret_cast<float4*>(s.dbg_rotations)[i] = q;
            s.dbg_opacities[i] = opacity;
        }
    } else {
        p_orig = {a.means3D[3 * i], a.means3D[3 * i + 1], a.means3D[3 * i + 2]};
        opacity = a.opacities[i];
    }

    uint32_t my_tiles = 0;
    int radius_out = 0;
    float3 p_view = xform_point_4x3(p_orig, view);
    do {
        if (p_view.z <= 0.2f) {  // near cull (auxiliary.h:154)
            if (a.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
            break;
        }
        float4 p_hom = xform_point_4x4(p_orig, proj);
        float p_w = 1.0f / (p_hom.w + 0.0000001f);
        float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

        float cov3D[6];
        if (!FUSED && a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov3D[k] = a.cov3D_precomp[6 * i + k];
        } else {
            if (!FUSED) {
                sc = {a.scales[3 * i], a.scales[3 * i + 1], a.scales[3 * i + 2]};
                q = reinterpret_cast<const float4*>(a.rotations)[i];
            }
            cov3d_from_scale_rot(sc, a.scale_modifier, q, cov3D);
        }
        Cov2D c2 = cov2d_project(p_orig, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, cov3D, view);
        float3 cov = c2.cov;
        float det = (cov.x * cov.z - cov.y * cov.y);
        if (det == 0.0f) break;
        float det_inv = 1.f / det;
        float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
        float mid = 0.5f * (cov.x + cov.z);
        float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        float2 point_image = {ndc_to_pix(p_proj.x, a.W), ndc_to_pix(p_proj.y, a.H)};
        uint2 rect_min, rect_max;
        tile_rect(point_image, (int)my_radius, rect_min, rect_max, a.grid_x, a.grid_y);
        if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) break;

        float3 color;
        uint8_t clamp_bits = 0;
        if (FUSED || a.colors_precomp == nullptr) {
            float sh[48];
            if (FUSED) {
                if (sh_bulk) mbar_wait(&s_bar, 0);
                load_sh_staged(s.features_dc, s_sh + (size_t)threadIdx.x * s.M_rest * 3, i, s.M_rest, sh);
            }
            else load_sh<16>(a.shs, i, a.M, a.sh_vec_ok, sh);
            const float3 campos = {s_mat[32], s_mat[33], s_mat[34]};
            color = sh_to_rgb(a.D, p_orig, campos, sh, clamp_bits);
        } else {
            color = {a.colors_precomp[3 * i], a.colors_precomp[3 * i + 1], a.colors_precomp[3 * i + 2]};
        }
        a.clamped[idx] = clamp_bits;
        write_splat_rec(a.rec + 4 * (size_t)idx, point_image, conic, opacity, color, p_view.z);
        a.depths[idx] = p_view.z;
        radius_out = (int)my_radius;
        a.means2D[idx] = point_image;
        a.conic_opacity[idx] = {conic.x, conic.y, conic.z, opacity};
        my_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
    } while (false);
    a.radii[idx] = radius_out;
    a.tiles_touched[idx] = my_tiles;
    if (my_tiles == 0) a.depths[idx] = __int_as_float(0x7f800000);  // culled: +inf sorts behind everything
    a.ids[idx] = idx;
    if (count_digits) {
        const uint32_t key = my_tiles == 0 ? 0x7f800000u : __float_as_uint(p_view.z);
#pragma unroll
        for (int d = 0; d < 4; d++) atomicAdd(&s_hist[d * 256 + ((key >> (8 * d)) & 255u)], 1u);
    }
    }  // has_row
    if (count_digits) {
        __syncthreads();
        for (int k = threadIdx.x; k < 4 * 256; k += blockDim.x)
            if (s_hist[k]) atomicAdd(&a.sort_hist[k], s_hist[k]);
    }
    // A culled Gaussian leaves the loop above without reading the staged SH block, so it never waited for the bulk
    // copy.  The CTA must not retire while the copy engine may still write into its shared memory: the issuing thread
    // waits for the barrier's phase here (immediate when it already waited above).
    if (FUSED && sh_bulk && threadIdx.x == 0) mbar_wait(&s_bar, 0);
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    float3 pv = xform_point_4x3(p, view);
    present[idx] = pv.z > 0.2f;
}

int launch_preprocess(const PreprocessArgs& a, cudaStream_t stream) {
    if (a.P == 0) return 0;
    LaunchScope ls("preprocess", stream);
    SkinFuseArgs none = {};
    preprocess_kernel<false><<<(a.P + 255) / 256, 256, 0, stream>>>(a, none);
    return 0;
}

int launch_skin_preprocess(const PreprocessArgs& a, const SkinFuseArgs& s, cudaStream_t stream) {
    if (a.P == 0) return 0;
    if (s.J > kMaxJoints || s.J < 2) { set_error("skinned preprocess: J must be in [2, 128]"); return -1; }
    LaunchScope ls("skin_preprocess", stream);
    const size_t smem = align_up(bone_smem_bytes(s.J, false), 128) + (size_t)256 * s.M_rest * 3 * sizeof(float) + 16;
    if (ensure_dyn_smem(reinterpret_cast<const void*>(preprocess_kernel<true>), 96 * 1024)) return -1;
    if (s.blend_mode == 1) {
        if (ensure_dyn_smem(reinterpret_cast<const void*>(preprocess_kernel<true, true>), 96 * 1024)) return -1;
        preprocess_kernel<true, true><<<(a.P + 255) / 256, 256, smem, stream>>>(a, s);
    } else {
        preprocess_kernel<true><<<(a.P + 255) / 256, 256, smem, stream>>>(a, s);
    }
    return 0;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream) {
    if (P == 0) return 0;
    LaunchScope ls("mark_visible", stream);
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
    return 0;
}

}  // namespace rg
