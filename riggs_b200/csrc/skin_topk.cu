// Linear-blend skinning by the K NEAREST bones only: SkeletonWarp.cal_nn_weight_skeleton with K > 0
// (skeleton_utils/skeleton_warp.py:46-49: torch.topk of the point-to-segment distances, smallest first, idx + 1)
// followed by the LBS of deform_by_pose (:130-172), forward and backward.  The all-bones case (K = -1, what the
// reference's scripts run) is skin.cu / the fused kernels; this is the same arithmetic restricted to the selected
// bones, so the per-bone gradient sums are scattered (shared-memory atomics per CTA, one global atomic per value per
// CTA) instead of transpose-reduced in lock step.
#include "bones.cuh"
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kTopKMax = 8;

struct SkinTopKArgs {
    int N, J, K;
    const float* xyz;
    const float* joints;
    const int* parents;
    const float* node_radius_raw;
    const float* G;
    const float* node_rot;
    const float* global_trans;
    const float* motion_mask;  // optional
    const float* offsets;      // optional
};

__global__ void __launch_bounds__(256) skin_topk_forward_kernel(SkinTopKArgs a, float* __restrict__ d_xyz,
                                                                float* __restrict__ d_rot, float* __restrict__ nn_weight,
                                                                long long* __restrict__ nn_idx, float* __restrict__ nn_dist2) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    stage_bones(a.J, a.joints, a.parents, a.node_radius_raw, a.G, a.node_rot, s_b);
    __syncthreads();
    const int B = a.J - 1, K = a.K;
    const float g0 = a.global_trans[0], g1 = a.global_trans[1], g2 = a.global_trans[2];
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += gridDim.x * blockDim.x) {
        const float x = a.xyz[3 * (size_t)n], y = a.xyz[3 * (size_t)n + 1], z = a.xyz[3 * (size_t)n + 2];
        float bd[kTopKMax];
        int bi[kTopKMax];
#pragma unroll
        for (int s = 0; s < kTopKMax; s++) { bd[s] = 3.0e38f; bi[s] = 0; }
        for (int k = 0; k < B; k++) {
            const float d2 = seg_dist2(s_b[k], x, y, z);
            if (d2 < bd[kTopKMax - 1]) {  // strict: of two equal distances the lower bone stays in front
                bd[kTopKMax - 1] = d2;
                bi[kTopKMax - 1] = k;
#pragma unroll
                for (int s = kTopKMax - 1; s > 0; s--) {
                    if (bd[s] < bd[s - 1]) {
                        const float td = bd[s]; bd[s] = bd[s - 1]; bd[s - 1] = td;
                        const int ti = bi[s]; bi[s] = bi[s - 1]; bi[s - 1] = ti;
                    }
                }
            }
        }
        float w[kTopKMax], wsum = 0.f;
#pragma unroll
        for (int s = 0; s < kTopKMax; s++) {
            w[s] = 0.f;
            if (s < K) { w[s] = expf(-bd[s] * s_b[bi[s]].inv_2r2) + 1e-7f; wsum += w[s]; }
        }
        const float inv = 1.0f / wsum;
        float ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
        for (int s = 0; s < kTopKMax; s++) {
            if (s < K) {
                const BoneSm& b = s_b[bi[s]];
                const float ws = w[s] * inv;
                ax += ws * (b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9]);
                ay += ws * (b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10]);
                az += ws * (b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11]);
                q0 += ws * b.q[0]; q1 += ws * b.q[1]; q2 += ws * b.q[2]; q3 += ws * b.q[3];
                nn_weight[(size_t)n * K + s] = ws;
                nn_idx[(size_t)n * K + s] = bi[s] + 1;
                nn_dist2[(size_t)n * K + s] = bd[s];
            }
        }
        const float m = a.motion_mask ? a.motion_mask[n] : 1.0f;
        float ox = 0.f, oy = 0.f, oz = 0.f;
        if (a.offsets) { ox = a.offsets[3 * (size_t)n]; oy = a.offsets[3 * (size_t)n + 1]; oz = a.offsets[3 * (size_t)n + 2]; }
        d_xyz[3 * (size_t)n] = (ax + g0 + ox - x) * m;
        d_xyz[3 * (size_t)n + 1] = (ay + g1 + oy - y) * m;
        d_xyz[3 * (size_t)n + 2] = (az + g2 + oz - z) * m;
        d_rot[4 * (size_t)n] = q0 * m; d_rot[4 * (size_t)n + 1] = q1 * m;
        d_rot[4 * (size_t)n + 2] = q2 * m; d_rot[4 * (size_t)n + 3] = q3 * m;
    }
}

// Appendix A.2 of SURVEY.md restricted to the bones nn_idx selected in the forward pass.
__global__ void __launch_bounds__(256) skin_topk_backward_kernel(SkinTopKArgs a, const long long* __restrict__ nn_idx,
                                                                 const float* __restrict__ g_dxyz, const float* __restrict__ g_drot,
                                                                 float* __restrict__ dG, float* __restrict__ dradius,
                                                                 float* __restrict__ dgtrans, float* __restrict__ dmask,
                                                                 float* __restrict__ doffsets) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    const int B = a.J - 1, K = a.K;
    float* s_acc = reinterpret_cast<float*>(s_raw + sizeof(BoneSm) * (size_t)max(B, 1));  // [B][13] + 3
    stage_bones(a.J, a.joints, a.parents, a.node_radius_raw, a.G, a.node_rot, s_b);
    for (int i = threadIdx.x; i < B * 13 + 3; i += blockDim.x) s_acc[i] = 0.f;
    __syncthreads();
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += gridDim.x * blockDim.x) {
        const float x = a.xyz[3 * (size_t)n], y = a.xyz[3 * (size_t)n + 1], z = a.xyz[3 * (size_t)n + 2];
        const float m = a.motion_mask ? a.motion_mask[n] : 1.0f;
        float gx = 0.f, gy = 0.f, gz = 0.f, h0 = 0.f, h1 = 0.f, h2 = 0.f, h3 = 0.f;
        if (g_dxyz) { gx = g_dxyz[3 * (size_t)n]; gy = g_dxyz[3 * (size_t)n + 1]; gz = g_dxyz[3 * (size_t)n + 2]; }
        if (g_drot) { h0 = g_drot[4 * (size_t)n]; h1 = g_drot[4 * (size_t)n + 1]; h2 = g_drot[4 * (size_t)n + 2]; h3 = g_drot[4 * (size_t)n + 3]; }
        int bi[kTopKMax];
        float e[kTopKMax], d2[kTopKMax], c[kTopKMax];
        float wsum = 0.f, scw = 0.f, ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
        for (int s = 0; s < kTopKMax; s++) {
            bi[s] = 0; e[s] = 0.f; d2[s] = 0.f; c[s] = 0.f;
            if (s < K) {
                bi[s] = (int)nn_idx[(size_t)n * K + s] - 1;
                const BoneSm& b = s_b[bi[s]];
                d2[s] = seg_dist2(b, x, y, z);
                e[s] = expf(-d2[s] * b.inv_2r2);
                const float w = e[s] + 1e-7f;
                const float yx = b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9];
                const float yy = b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10];
                const float yz = b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11];
                c[s] = gx * yx + gy * yy + gz * yz + h0 * b.q[0] + h1 * b.q[1] + h2 * b.q[2] + h3 * b.q[3];
                wsum += w;
                scw += w * c[s];
                ax += w * yx; ay += w * yy; az += w * yz;
                q0 += w * b.q[0]; q1 += w * b.q[1]; q2 += w * b.q[2]; q3 += w * b.q[3];
            }
        }
        const float inv = 1.0f / wsum;
        const float cbar = scw * inv;
        float ox = 0.f, oy = 0.f, oz = 0.f;
        if (a.offsets) { ox = a.offsets[3 * (size_t)n]; oy = a.offsets[3 * (size_t)n + 1]; oz = a.offsets[3 * (size_t)n + 2]; }
        if (dmask)
            dmask[n] = gx * (ax * inv + a.global_trans[0] + ox - x) + gy * (ay * inv + a.global_trans[1] + oy - y) +
                       gz * (az * inv + a.global_trans[2] + oz - z) + (h0 * q0 + h1 * q1 + h2 * q2 + h3 * q3) * inv;
        const float mgx = m * gx, mgy = m * gy, mgz = m * gz;
        if (doffsets) { doffsets[3 * (size_t)n] = mgx; doffsets[3 * (size_t)n + 1] = mgy; doffsets[3 * (size_t)n + 2] = mgz; }
        atomicAdd(&s_acc[B * 13], mgx); atomicAdd(&s_acc[B * 13 + 1], mgy); atomicAdd(&s_acc[B * 13 + 2], mgz);
#pragma unroll
        for (int s = 0; s < kTopKMax; s++) {
            if (s < K) {
                const BoneSm& b = s_b[bi[s]];
                const float w = (e[s] + 1e-7f) * inv;
                const float dwhat = m * (c[s] - cbar) * inv;
                float* row = s_acc + bi[s] * 13;
                atomicAdd(row + 0, mgx * w * x); atomicAdd(row + 1, mgx * w * y); atomicAdd(row + 2, mgx * w * z);
                atomicAdd(row + 3, mgy * w * x); atomicAdd(row + 4, mgy * w * y); atomicAdd(row + 5, mgy * w * z);
                atomicAdd(row + 6, mgz * w * x); atomicAdd(row + 7, mgz * w * y); atomicAdd(row + 8, mgz * w * z);
                atomicAdd(row + 9, mgx * w); atomicAdd(row + 10, mgy * w); atomicAdd(row + 11, mgz * w);
                atomicAdd(row + 12, dwhat * e[s] * d2[s] * (2.0f * b.inv_2r2));
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < B * 13; i += blockDim.x) {
        const float v = s_acc[i];
        if (v != 0.f) {
            const int bone = i / 13, cc = i % 13;
            if (cc < 12) atomicAdd(&dG[12 * (bone + 1) + cc], v);
            else atomicAdd(&dradius[bone + 1], v);
        }
    }
    if (threadIdx.x < 3 && dgtrans) {
        const float v = s_acc[B * 13 + threadIdx.x];
        if (v != 0.f) atomicAdd(&dgtrans[threadIdx.x], v);
    }
}

}  // namespace rg

using namespace rg;

static int topk_check(const char* who, int N, int J, int K) {
    if (N < 0 || J < 2 || J > kMaxJoints || K < 1 || K > kTopKMax || K > J - 1) {
        set_error(std::string(who) + ": need 2 <= J <= 128 and 1 <= K <= min(8, J - 1)");
        return -1;
    }
    return 0;
}

extern "C" int rg_skin_topk_forward(int N, int J, int K, const float* xyz, const float* joints, const int* parents,
                                    const float* node_radius_raw, const float* G, const float* node_rot,
                                    const float* global_trans, const float* motion_mask, const float* offsets, float* d_xyz,
                                    float* d_rot, float* nn_weight, long long* nn_idx, float* nn_dist2, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (topk_check("rg_skin_topk_forward", N, J, K)) return -1;
    if (N == 0) return 0;
    if (!xyz || !joints || !parents || !node_radius_raw || !G || !node_rot || !global_trans || !d_xyz || !d_rot || !nn_weight ||
        !nn_idx || !nn_dist2) {
        set_error("rg_skin_topk_forward: null argument");
        return -1;
    }
    SkinTopKArgs a{N, J, K, xyz, joints, parents, node_radius_raw, G, node_rot, global_trans, motion_mask, offsets};
    LaunchScope ls("skin_topk_forward", stream);
    const int blocks = std::min((N + 255) / 256, RG_NUM_SMS * 8);
    skin_topk_forward_kernel<<<blocks, 256, bone_smem_bytes(J, false), stream>>>(a, d_xyz, d_rot, nn_weight, nn_idx, nn_dist2);
    if (cudaGetLastError() != cudaSuccess) { set_error("rg_skin_topk_forward: launch failed"); return -1; }
    return 0;
}

extern "C" int rg_skin_topk_backward(int N, int J, int K, const float* xyz, const float* joints, const int* parents,
                                     const float* node_radius_raw, const float* G, const float* node_rot,
                                     const float* global_trans, const float* motion_mask, const float* offsets,
                                     const long long* nn_idx, const float* g_dxyz, const float* g_drot, float* dG,
                                     float* dnode_radius_raw, float* dglobal_trans, float* dmask, float* doffsets,
                                     rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (topk_check("rg_skin_topk_backward", N, J, K)) return -1;
    if (N == 0) return 0;
    if (!xyz || !joints || !parents || !node_radius_raw || !G || !node_rot || !global_trans || !nn_idx || !dG || !dnode_radius_raw) {
        set_error("rg_skin_topk_backward: null argument");
        return -1;
    }
    SkinTopKArgs a{N, J, K, xyz, joints, parents, node_radius_raw, G, node_rot, global_trans, motion_mask, offsets};
    LaunchScope ls("skin_topk_backward", stream);
    const int blocks = std::min((N + 255) / 256, RG_NUM_SMS * 4);
    const size_t smem = bone_smem_bytes(J, false) + sizeof(float) * ((size_t)(J - 1) * 13 + 3);
    skin_topk_backward_kernel<<<blocks, 256, smem, stream>>>(a, nn_idx, g_dxyz, g_drot, dG, dnode_radius_raw, dglobal_trans, dmask,
                                                            doffsets);
    if (cudaGetLastError() != cudaSuccess) { set_error("rg_skin_topk_backward: launch failed"); return -1; }
    return 0;
}
