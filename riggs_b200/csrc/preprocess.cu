// Per-Gaussian forward stage: frustum cull, projection, 3D->2D covariance, conic, radius, tile rectangle,
// SH->RGB.  One thread per Gaussian; replaces preprocessCUDA (dgr/cuda_rasterizer/forward.cu:155-256) and
// checkFrustum (dgr/cuda_rasterizer/rasterizer_impl.cu:54-66).
//
// HBM-bound streaming kernel: algorithmic bytes per Gaussian = 44 (xyz, rot, scale, opacity) + 12*M (SH)
// read, 48 written (+1 clamp byte, +24 when cov3D is kept for a precomputed-covariance backward).
#include "common.cuh"
#include "math.cuh"
#include "kernels.cuh"

namespace rg {

// SH -> RGB for one Gaussian (forward.cu:20-71).  sh points at M*3 floats (coefficient-major, rgb inner).
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float3 pos, const float3 campos, const float* sh,
                                            uint8_t& clamp_bits) {
    float3 dir = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    const float x = dir.x / len, y = dir.y / len, z = dir.z / len;
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
#define SH(k) sh[(k) * 3 + c]
        float r = kSH_C0 * SH(0);
        if (deg > 0) {
            r = r - kSH_C1 * y * SH(1) + kSH_C1 * z * SH(2) - kSH_C1 * x * SH(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z;
                float xy = x * y, yz = y * z, xz = x * z;
                r = r + kSH_C2[0] * xy * SH(4) + kSH_C2[1] * yz * SH(5) + kSH_C2[2] * (2.0f * zz - xx - yy) * SH(6) +
                    kSH_C2[3] * xz * SH(7) + kSH_C2[4] * (xx - yy) * SH(8);
                if (deg > 2) {
                    r = r + kSH_C3[0] * y * (3.0f * xx - yy) * SH(9) + kSH_C3[1] * xy * z * SH(10) +
                        kSH_C3[2] * y * (4.0f * zz - xx - yy) * SH(11) +
                        kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12) +
                        kSH_C3[4] * x * (4.0f * zz - xx - yy) * SH(13) + kSH_C3[5] * z * (xx - yy) * SH(14) +
                        kSH_C3[6] * x * (xx - 3.0f * yy) * SH(15);
                }
            }
        }
#undef SH
        r += 0.5f;
        res[c] = r;
    }
    clamp_bits = (res[0] < 0 ? 1 : 0) | (res[1] < 0 ? 2 : 0) | (res[2] < 0 ? 4 : 0);
    return {fmaxf(res[0], 0.0f), fmaxf(res[1], 0.0f), fmaxf(res[2], 0.0f)};
}

// Loads the M*3 SH floats of Gaussian idx into registers; 128-bit loads when the row is 16-byte aligned.
template <int MAXC>
__device__ __forceinline__ void load_sh(const float* __restrict__ shs, size_t idx, int M, bool vec_ok, float* sh) {
    const float* src = shs + idx * (size_t)M * 3;
    if (vec_ok && M == MAXC) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
        for (int i = 0; i < MAXC * 3 / 4; i++) {
            float4 v = __ldg(s4 + i);
            sh[4 * i] = v.x; sh[4 * i + 1] = v.y; sh[4 * i + 2] = v.z; sh[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < MAXC * 3; i++) sh[i] = i < M * 3 ? __ldg(src + i) : 0.f;
    }
}

__global__ void __launch_bounds__(256) preprocess_kernel(PreprocessArgs a) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ float s_mat[35];  // view[16], proj[16], campos[3]
    if (threadIdx.x < 16) s_mat[threadIdx.x] = a.viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_mat[threadIdx.x] = a.projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_mat[threadIdx.x] = a.cam_pos[threadIdx.x - 32];
    __syncthreads();
    const float* view = s_mat;
    const float* proj = s_mat + 16;
    uint32_t my_tiles = 0;
    if (idx < a.P) {
        int radius_out = 0;
        float3 p_orig = {a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]};
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
            if (a.cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; k++) cov3D[k] = a.cov3D_precomp[6 * (size_t)idx + k];
            } else {
                float3 sc = {a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]};
                float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
                cov3d_from_scale_rot(sc, a.scale_modifier, q, cov3D);
                if (a.cov3D_out) {
#pragma unroll
                    for (int k = 0; k < 6; k++) a.cov3D_out[6 * (size_t)idx + k] = cov3D[k];
                }
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
            if (a.colors_precomp == nullptr) {
                float sh[48];
                load_sh<16>(a.shs, idx, a.M, a.sh_vec_ok, sh);
                const float3 campos = {s_mat[32], s_mat[33], s_mat[34]};
                color = sh_to_rgb(a.D, p_orig, campos, sh, clamp_bits);
            } else {
                color = {a.colors_precomp[3 * idx], a.colors_precomp[3 * idx + 1], a.colors_precomp[3 * idx + 2]};
            }
            a.clamped[idx] = clamp_bits;
            a.rgb[3 * idx] = color.x;
            a.rgb[3 * idx + 1] = color.y;
            a.rgb[3 * idx + 2] = color.z;
            a.depths[idx] = p_view.z;
            radius_out = (int)my_radius;
            a.means2D[idx] = point_image;
            a.conic_opacity[idx] = {conic.x, conic.y, conic.z, a.opacities[idx]};
            my_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
        } while (false);
        a.radii[idx] = radius_out;
        a.tiles_touched[idx] = my_tiles;
        if (my_tiles == 0) a.depths[idx] = __int_as_float(0x7f800000);  // culled: +inf sorts behind everything
        a.ids[idx] = idx;
    }
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
    preprocess_kernel<<<(a.P + 255) / 256, 256, 0, stream>>>(a);
    return 0;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream) {
    if (P == 0) return 0;
    LaunchScope ls("mark_visible", stream);
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
    return 0;
}

}  // namespace rg
