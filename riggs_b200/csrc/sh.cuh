// Spherical-harmonics colour: forward (dgr/cuda_rasterizer/forward.cu:20-71) and backward
// (dgr/cuda_rasterizer/backward.cu:20-139), written per channel in the reference's expression order.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "math.cuh"

namespace rg {

// SH -> RGB for one Gaussian.  sh points at 16*3 floats in registers (coefficient-major, rgb inner).
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

// Loads the M*3 SH floats of Gaussian idx into 48 registers (zero padded); 128-bit loads when rows are aligned.
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

// Same, from the model's two raw tensors features_dc (N,1,3) and features_rest (N,M_rest,3): the fused entry
// never materialises their concatenation.
__device__ __forceinline__ void load_sh_split(const float* __restrict__ dc, const float* __restrict__ rest, size_t idx,
                                              int M_rest, float* sh) {
    sh[0] = __ldg(dc + 3 * idx); sh[1] = __ldg(dc + 3 * idx + 1); sh[2] = __ldg(dc + 3 * idx + 2);
    const float* src = rest + idx * (size_t)M_rest * 3;
#pragma unroll
    for (int i = 0; i < 45; i++) sh[3 + i] = i < M_rest * 3 ? __ldg(src + i) : 0.f;
}

// Same, with this Gaussian's features_rest row already staged in shared memory (row stride M_rest*3 floats; an
// odd stride of 45 words makes the per-lane reads bank-conflict free).
__device__ __forceinline__ void load_sh_staged(const float* __restrict__ dc, const float* s_row, size_t idx, int M_rest,
                                               float* sh) {
    sh[0] = __ldg(dc + 3 * idx); sh[1] = __ldg(dc + 3 * idx + 1); sh[2] = __ldg(dc + 3 * idx + 2);
#pragma unroll
    for (int i = 0; i < 45; i++) sh[3 + i] = i < M_rest * 3 ? s_row[i] : 0.f;
}

__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {  // auxiliary.h:107-117
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward.  Outputs the 16 basis values (zero above the active degree) and the clamp-masked colour
// gradient g, so that dL_dsh[k] = basis[k] * g; returns the part of dL/dmean that flows through the view
// direction.
__device__ __forceinline__ float3 sh_backward_core(int deg, const float3 pos, const float3 campos, const float* sh,
                                                   uint8_t clamp_bits, float3 dL_dcolor, float* basis, float* g) {
    const float3 dir_orig = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    g[0] = dL_dcolor.x * ((clamp_bits & 1) ? 0.f : 1.f);
    g[1] = dL_dcolor.y * ((clamp_bits & 2) ? 0.f : 1.f);
    g[2] = dL_dcolor.z * ((clamp_bits & 4) ? 0.f : 1.f);
    float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 16; k++) basis[k] = 0.f;
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
    const float3 dL_ddir = {dx[0] * g[0] + dx[1] * g[1] + dx[2] * g[2], dy[0] * g[0] + dy[1] * g[1] + dy[2] * g[2],
                            dz[0] * g[0] + dz[1] * g[1] + dz[2] * g[2]};
    return dnormvdv(dir_orig, dL_ddir);
}

}  // namespace rg
