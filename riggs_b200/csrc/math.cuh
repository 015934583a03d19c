// Small fp32 math shared by the per-Gaussian kernels.
//
// The bit-exact gates (tile keys, radii, tile ranges) require the same float expression trees -- and
// therefore the same nvcc FMA contraction -- as the reference's device code.  The reference writes its
// 3x3 algebra with GLM (column-major, element [c][r]; product element = a[0][r]*b[c][0] + a[1][r]*b[c][1]
// + a[2][r]*b[c][2], dgr/third_party/glm/glm/detail/type_mat3x3.inl:486-519).  Mat3 below keeps that
// element order so the generated arithmetic is the same; it shares no code with GLM.
#pragma once
#include <cuda_runtime.h>

namespace rg {

struct Mat3 {
    float m[3][3];  // m[column][row]
};

__device__ __forceinline__ Mat3 mat3_cols(float a0, float a1, float a2, float b0, float b1, float b2, float c0,
                                          float c1, float c2) {
    Mat3 r;
    r.m[0][0] = a0; r.m[0][1] = a1; r.m[0][2] = a2;
    r.m[1][0] = b0; r.m[1][1] = b1; r.m[1][2] = b2;
    r.m[2][0] = c0; r.m[2][1] = c1; r.m[2][2] = c2;
    return r;
}

__device__ __forceinline__ Mat3 mul(const Mat3& a, const Mat3& b) {
    Mat3 r;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++)
            r.m[c][k] = a.m[0][k] * b.m[c][0] + a.m[1][k] * b.m[c][1] + a.m[2][k] * b.m[c][2];
    return r;
}

__device__ __forceinline__ Mat3 transpose(const Mat3& a) {
    Mat3 r;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int k = 0; k < 3; k++) r.m[c][k] = a.m[k][c];
    return r;
}

// Row-vector convention of the reference's matrices: element [4*c + r] (dgr/cuda_rasterizer/auxiliary.h:58-77).
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}

// Pixel coordinate from NDC; evaluated in double like the reference (auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle of a splat (auxiliary.h:46-56): C truncation of float quotients, clamped to the grid.
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rect_min, uint2& rect_max,
                                          int grid_x, int grid_y) {
    rect_min = {(unsigned)min(grid_x, max((int)0, (int)((p.x - max_radius) / RG_TILE))),
                (unsigned)min(grid_y, max((int)0, (int)((p.y - max_radius) / RG_TILE)))};
    rect_max = {(unsigned)min(grid_x, max((int)0, (int)((p.x + max_radius + RG_TILE - 1) / RG_TILE))),
                (unsigned)min(grid_y, max((int)0, (int)((p.y + max_radius + RG_TILE - 1) / RG_TILE)))};
}

// Rotation matrix of an un-normalised quaternion (w,x,y,z) in GLM element order (forward.cu:134-138).
__device__ __forceinline__ Mat3 quat_rot(float r, float x, float y, float z) {
    return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                     2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                     2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// World-space covariance (upper triangle) from scale and quaternion (forward.cu:118-152).
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 q, float* cov) {
    Mat3 S = mat3_cols(mod * scale.x, 0.f, 0.f, 0.f, mod * scale.y, 0.f, 0.f, 0.f, mod * scale.z);
    Mat3 R = quat_rot(q.x, q.y, q.z, q.w);
    Mat3 M = mul(S, R);
    Mat3 Sigma = mul(transpose(M), M);
    cov[0] = Sigma.m[0][0]; cov[1] = Sigma.m[0][1]; cov[2] = Sigma.m[0][2];
    cov[3] = Sigma.m[1][1]; cov[4] = Sigma.m[1][2]; cov[5] = Sigma.m[2][2];
}

// EWA projection of the 3-D covariance (forward.cu:74-113).  Returns (a, b, c) of the 2x2 covariance after
// the +0.3 low-pass; T, the clamped view-space point and the clamp masks are returned for the backward.
struct Cov2D {
    float3 cov;
    Mat3 T;
    float3 t;
    float x_grad_mul, y_grad_mul;
};
__device__ __forceinline__ Cov2D cov2d_project(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                               float tan_fovy, const float* cov3D, const float* viewmatrix) {
    Cov2D o;
    float3 t = xform_point_4x3(mean, viewmatrix);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    o.x_grad_mul = txtz < -limx || txtz > limx ? 0.f : 1.f;
    o.y_grad_mul = tytz < -limy || tytz > limy ? 0.f : 1.f;
    Mat3 J = mat3_cols(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
                       -(focal_y * t.y) / (t.z * t.z), 0.f, 0.f, 0.f);
    Mat3 W = mat3_cols(viewmatrix[0], viewmatrix[4], viewmatrix[8], viewmatrix[1], viewmatrix[5], viewmatrix[9],
                       viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    Mat3 T = mul(W, J);
    Mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    Mat3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
    o.cov = {cov.m[0][0] + 0.3f, cov.m[0][1], cov.m[1][1] + 0.3f};
    o.T = T;
    o.t = t;
    return o;
}

// SH basis constants (auxiliary.h:22-39).
__device__ constexpr float kSH_C0 = 0.28209479177387814f;
__device__ constexpr float kSH_C1 = 0.4886025119029199f;
__device__ constexpr float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                        -1.0925484305920792f, 0.5462742152960396f};
__device__ constexpr float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                        0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                        -0.5900435899266435f};

}  // namespace rg
