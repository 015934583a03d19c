// Skeleton skinning: forward kinematics over the bone tree and linear-blend (or dual-quaternion) skinning
// of the Gaussians.  Replaces the ~40 + 2J eager torch kernels of SkeletonWarp.deform_by_pose
// (skeleton_utils/skeleton_warp.py:130-172) -- bone-segment distances (:207-238), gaussian-kernel weights
// (:41-76), the Python loop of 4x4 chain products (:242-273), the (N,K,3,3) gather + einsum (:149-150) --
// by one tiny FK kernel (one thread per pose) and one streaming kernel over the Gaussians that keeps the
// whole bone table in shared memory and never materialises an (N,K,...) temporary.
//
// Skin kernels are HBM-streaming: 12 B read (+4 mask, +12 offsets, +4(J-1) MLP weights when present) and
// 28 B written per Gaussian in the forward; the ~60 flops per (Gaussian, bone) pair stay in registers.
#include "common.cuh"
#include "kernels.cuh"
#include "dq.cuh"
#include "skin.cuh"
#include "bones.cuh"

namespace rg {

// ---- forward kinematics -------------------------------------------------------------------------------

// Rotation matrix (row-major r[3*a+b]) of an un-normalised quaternion, scaled by 2/|q|^2
// (utils/time_utils.py:115-132).
__device__ __forceinline__ void quat_to_mat_rm(const float* q, float* R) {
    const float r = q[0], i = q[1], j = q[2], k = q[3];
    const float two_s = 2.0f / (r * r + i * i + j * j + k * k);
    R[0] = 1 - two_s * (j * j + k * k);
    R[1] = two_s * (i * j - k * r);
    R[2] = two_s * (i * k + j * r);
    R[3] = two_s * (i * j + k * r);
    R[4] = 1 - two_s * (i * i + k * k);
    R[5] = two_s * (j * k - i * r);
    R[6] = two_s * (i * k - j * r);
    R[7] = two_s * (j * k + i * r);
    R[8] = 1 - two_s * (i * i + j * j);
}

// matrix_to_quaternion (utils/time_utils.py:146-205): four candidates, the one with the largest |component|
// wins (first index on ties, like argmax); denominators floored at 0.1; no sign standardisation.
__device__ __forceinline__ void mat_to_quat_rm(const float* m, float* q) {
    const float m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5], m20 = m[6], m21 = m[7],
                m22 = m[8];
    float t[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22, 1.0f - m00 - m11 + m22};
    float qa[4];
#pragma unroll
    for (int c = 0; c < 4; c++) qa[c] = t[c] > 0 ? sqrtf(t[c]) : 0.f;
    int best = 0;
#pragma unroll
    for (int c = 1; c < 4; c++)
        if (qa[c] > qa[best]) best = c;
    float cand[4];
    if (best == 0) { cand[0] = qa[0] * qa[0]; cand[1] = m21 - m12; cand[2] = m02 - m20; cand[3] = m10 - m01; }
    else if (best == 1) { cand[0] = m21 - m12; cand[1] = qa[1] * qa[1]; cand[2] = m10 + m01; cand[3] = m02 + m20; }
    else if (best == 2) { cand[0] = m02 - m20; cand[1] = m10 + m01; cand[2] = qa[2] * qa[2]; cand[3] = m12 + m21; }
    else { cand[0] = m10 - m01; cand[1] = m20 + m02; cand[2] = m21 + m12; cand[3] = qa[3] * qa[3]; }
    const float den = 2.0f * fmaxf(qa[best], 0.1f);
#pragma unroll
    for (int c = 0; c < 4; c++) q[c] = cand[c] / den;
}

// One warp per pose.  Lanes build the local transforms of all joints in shared memory in parallel, lane 0 walks
// the chain G_i = G_parent(i) * T_i (parents[i] < i) on shared memory, then the lanes derive node_rot / posed
// joints and write everything out coalesced.  A pose costs a few microseconds instead of ~2J eager launches;
// the poses of a sequence (novel-pose rendering) run side by side, one CTA each.
__global__ void __launch_bounds__(32) fk_forward_kernel(int F, int J, const float* __restrict__ local_rot,
                                                        const float* __restrict__ gtrans, const float* __restrict__ joints,
                                                        const int* __restrict__ parents, float* __restrict__ G,
                                                        float* __restrict__ node_rot, float* __restrict__ posed) {
    __shared__ float s_T[kMaxJoints * 12];   // local rotation (row-major) + local translation
    __shared__ float s_G[kMaxJoints * 12];
    __shared__ int s_par[kMaxJoints];
    const int f = blockIdx.x, lane = threadIdx.x;
    const float* q = local_rot + (size_t)f * J * 4;
    for (int i = lane; i < J; i += 32) {
        float R[9];
        quat_to_mat_rm(q + 4 * i, R);
        const int par = parents[i];
        s_par[i] = par;
        const int vp = i == 0 ? 0 : par;  // the root pivots on itself (skeleton_warp.py:246-247)
        const float p[3] = {joints[3 * vp], joints[3 * vp + 1], joints[3 * vp + 2]};
#pragma unroll
        for (int a = 0; a < 9; a++) s_T[12 * i + a] = R[a];
#pragma unroll
        for (int a = 0; a < 3; a++) s_T[12 * i + 9 + a] = p[a] - (R[3 * a] * p[0] + R[3 * a + 1] * p[1] + R[3 * a + 2] * p[2]);
    }
    __syncwarp();
    // The chain itself is serial in i, but the twelve elements of one G_i are independent: lanes 0..11 take one each
    // (the single-lane walk was ~400 cycles per joint, all of it shared-memory round trips).
    if (lane < 12) s_G[lane] = s_T[lane];
    __syncwarp();
    for (int i = 1; i < J; i++) {
        if (lane < 12) {
            const float* Gp = s_G + 12 * s_par[i];
            const float* T = s_T + 12 * i;
            const int a = lane < 9 ? lane / 3 : lane - 9, b = lane < 9 ? lane % 3 : 0;
            float v;
            if (lane < 9) v = Gp[3 * a] * T[b] + Gp[3 * a + 1] * T[3 + b] + Gp[3 * a + 2] * T[6 + b];
            else v = Gp[3 * a] * T[9] + Gp[3 * a + 1] * T[10] + Gp[3 * a + 2] * T[11] + Gp[9 + a];
            s_G[12 * i + lane] = v;
        }
        __syncwarp();
    }
    float* Gf = G + (size_t)f * J * 12;
    for (int k = lane; k < J * 12; k += 32) Gf[k] = s_G[k];
    for (int i = lane; i < J; i += 32) {
        const float* Gi = s_G + 12 * i;
        if (node_rot) mat_to_quat_rm(Gi, node_rot + ((size_t)f * J + i) * 4);
        if (posed) {
            const float jx = joints[3 * i], jy = joints[3 * i + 1], jz = joints[3 * i + 2];
#pragma unroll
            for (int a = 0; a < 3; a++)
                posed[((size_t)f * J + i) * 3 + a] = Gi[3 * a] * jx + Gi[3 * a + 1] * jy + Gi[3 * a + 2] * jz + Gi[9 + a] + gtrans[3 * f + a];
        }
    }
}

// Reverse sweep, same layout: lanes seed the per-joint gradients of the global transforms in shared memory,
// lane 0 walks the chain back from the leaves, lanes finish with the quaternion gradients.
__global__ void __launch_bounds__(32) fk_backward_kernel(int F, int J, const float* __restrict__ local_rot,
                                                         const float* __restrict__ joints, const int* __restrict__ parents,
                                                         const float* __restrict__ G, const float* __restrict__ dG_in,
                                                         const float* __restrict__ dposed, float* __restrict__ dlocal_rot,
                                                         float* __restrict__ dgtrans, const float* __restrict__ dgtrans_add,
                                                         float* __restrict__ radius_dst, const float* __restrict__ radius_src,
                                                         int radius_add) {
    __shared__ float s_W[kMaxJoints * 12];   // running dL/dG_i; afterwards dL/dT_i (local rotation + translation)
    __shared__ float s_G[kMaxJoints * 12];
    __shared__ float s_R[kMaxJoints * 9];
    __shared__ float s_p[kMaxJoints * 3];    // pivot of joint i
    __shared__ int s_par[kMaxJoints];
    const int f = blockIdx.x, lane = threadIdx.x;
    const float* q = local_rot + (size_t)f * J * 4;
    float gt[3] = {0, 0, 0};
    // rides along (fused entry): dL/d node_radius_raw leaves the workspace's accumulator as a copy or, for the later views of
    // a batch, as a sum -- J floats that had a launch of their own
    if (radius_dst && f == 0)
        for (int k = lane; k < J; k += 32) radius_dst[k] = radius_add ? radius_dst[k] + radius_src[k] : radius_src[k];
    for (int k = lane; k < J * 12; k += 32) s_G[k] = G[(size_t)f * J * 12 + k];
    for (int i = lane; i < J; i += 32) {
        float* Wi = s_W + 12 * i;
#pragma unroll
        for (int a = 0; a < 12; a++) Wi[a] = dG_in ? dG_in[((size_t)f * J + i) * 12 + a] : 0.f;
        if (dposed) {  // posed_i = G_R p_i + G_t + g
            const float* dp = dposed + ((size_t)f * J + i) * 3;
            const float jx = joints[3 * i], jy = joints[3 * i + 1], jz = joints[3 * i + 2];
#pragma unroll
            for (int a = 0; a < 3; a++) {
                Wi[3 * a] += dp[a] * jx; Wi[3 * a + 1] += dp[a] * jy; Wi[3 * a + 2] += dp[a] * jz;
                Wi[9 + a] += dp[a];
                gt[a] += dp[a];
            }
        }
        quat_to_mat_rm(q + 4 * i, s_R + 9 * i);
        const int par = parents[i];
        s_par[i] = par;
        const int vp = i == 0 ? 0 : par;
#pragma unroll
        for (int a = 0; a < 3; a++) s_p[3 * i + a] = joints[3 * vp + a];
    }
    if (dgtrans) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float t = warp_sum(gt[a]);
            if (lane == 0) dgtrans[3 * f + a] = t + (dgtrans_add ? dgtrans_add[a] : 0.f);
        }
    }
    __syncwarp();
    // Serial part: children push their gradient into the parent, leaves first; the twelve elements of the parent's
    // accumulator are independent (lanes 0..11).  G_R_i = Gp_R * R ; G_t_i = Gp_R * tl + Gp_t, tl = p - R p.
    for (int i = J - 1; i >= 1; i--) {
        if (lane < 12) {
            const float* Wi = s_W + 12 * i;
            const float* R = s_R + 9 * i;
            const float* p = s_p + 3 * i;
            float* Wp = s_W + 12 * s_par[i];
            const int a = lane < 9 ? lane / 3 : lane - 9, b = lane < 9 ? lane % 3 : 0;
            float add;
            if (lane < 9) {
                const float tlb = p[b] - (R[3 * b] * p[0] + R[3 * b + 1] * p[1] + R[3 * b + 2] * p[2]);
                add = Wi[3 * a] * R[3 * b] + Wi[3 * a + 1] * R[3 * b + 1] + Wi[3 * a + 2] * R[3 * b + 2] + Wi[9 + a] * tlb;
            } else {
                add = Wi[9 + a];
            }
            Wp[lane] += add;
        }
        __syncwarp();
    }
    // Every accumulator is final now: turn dL/dG_i into the gradient of the local transform, one joint per lane
    // (in place: a joint reads only its own row and its parent's G).
    for (int i = 1 + lane; i < J; i += 32) {
        float* Wi = s_W + 12 * i;
        const float* Gp = s_G + 12 * s_par[i];
        float gRl[9], gtl[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
#pragma unroll
            for (int b = 0; b < 3; b++) gRl[3 * a + b] = Gp[a] * Wi[b] + Gp[3 + a] * Wi[3 + b] + Gp[6 + a] * Wi[6 + b];
            gtl[a] = Gp[a] * Wi[9] + Gp[3 + a] * Wi[10] + Gp[6 + a] * Wi[11];
        }
#pragma unroll
        for (int a = 0; a < 9; a++) Wi[a] = gRl[a];
#pragma unroll
        for (int a = 0; a < 3; a++) Wi[9 + a] = gtl[a];
    }
    __syncwarp();
    for (int i = lane; i < J; i += 32) {
        const float* Wi = s_W + 12 * i;   // dL/dR_i (local), dL/dtl_i
        const float* p = s_p + 3 * i;
        float gRl[9];
        // tl = p - R p  ->  dR[a][b] -= gtl[a] p[b]
#pragma unroll
        for (int a = 0; a < 3; a++)
#pragma unroll
            for (int b = 0; b < 3; b++) gRl[3 * a + b] = Wi[3 * a + b] - Wi[9 + a] * p[b];
        // R = I + s*B(q), s = 2/|q|^2, ds/dq_m = -s^2 q_m
        const float r = q[4 * i], x = q[4 * i + 1], y = q[4 * i + 2], z = q[4 * i + 3];
        const float s = 2.0f / (r * r + x * x + y * y + z * z);
        const float B[9] = {-(y * y + z * z), x * y - z * r, x * z + y * r, x * y + z * r, -(x * x + z * z),
                            y * z - x * r, x * z - y * r, y * z + x * r, -(x * x + y * y)};
        float gB = 0.f;
#pragma unroll
        for (int a = 0; a < 9; a++) gB += gRl[a] * B[a];
        const float dBr = -z * gRl[1] + y * gRl[2] + z * gRl[3] - x * gRl[5] - y * gRl[6] + x * gRl[7];
        const float dBx = y * gRl[1] + z * gRl[2] + y * gRl[3] - 2 * x * gRl[4] - r * gRl[5] + z * gRl[6] + r * gRl[7] - 2 * x * gRl[8];
        const float dBy = -2 * y * gRl[0] + x * gRl[1] + r * gRl[2] + x * gRl[3] + z * gRl[5] - r * gRl[6] + z * gRl[7] - 2 * y * gRl[8];
        const float dBz = -2 * z * gRl[0] - r * gRl[1] + x * gRl[2] + r * gRl[3] - 2 * z * gRl[4] + y * gRl[5] + x * gRl[6] + y * gRl[7];
        float* dq = dlocal_rot + ((size_t)f * J + i) * 4;
        const float s2 = s * s;
        dq[0] = s * dBr - s2 * r * gB;
        dq[1] = s * dBx - s2 * x * gB;
        dq[2] = s * dBy - s2 * y * gB;
        dq[3] = s * dBz - s2 * z * gB;
    }
}

// ---- skinning -------------------------------------------------------------------------------------------


__global__ void __launch_bounds__(256) skin_forward_kernel(SkinArgs a) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    stage_bones(a.J, a.joints, a.parents, a.node_radius_raw, a.G, a.node_rot, s_b);
    __syncthreads();
    const int K = a.J - 1;
    const float g0 = a.global_trans[0], g1 = a.global_trans[1], g2 = a.global_trans[2];
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += gridDim.x * blockDim.x) {
        const float x = a.xyz[3 * (size_t)n], y = a.xyz[3 * (size_t)n + 1], z = a.xyz[3 * (size_t)n + 2];
        float wsum = 0.f, ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
        // dual-quaternion accumulators (blend_mode 1)
        float dr[4] = {0, 0, 0, 0}, dd[4] = {0, 0, 0, 0};
        for (int k = 0; k < K; k++) {
            const BoneSm& b = s_b[k];
            const float d2 = seg_dist2(b, x, y, z);
            float w = expf(-d2 * b.inv_2r2);
            if (a.mlp_weight) w *= a.mlp_weight[(size_t)n * K + k];
            w += 1e-7f;
            if (a.nn_weight) a.nn_weight[(size_t)n * K + k] = w;
            wsum += w;
            if (a.blend_mode == 0) {
                ax += w * (b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9]);
                ay += w * (b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10]);
                az += w * (b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11]);
                q0 += w * b.q[0]; q1 += w * b.q[1]; q2 += w * b.q[2]; q3 += w * b.q[3];
            } else {
                // QT2DQ (utils/dual_quaternion.py:135-144): real = normalised q, dual = 0.5 * standardise((0,t) * q)
                float qn = rsqrtf(fmaxf(b.q[0] * b.q[0] + b.q[1] * b.q[1] + b.q[2] * b.q[2] + b.q[3] * b.q[3], 1e-24f));
                const float rw = b.q[0] * qn, rx = b.q[1] * qn, ry = b.q[2] * qn, rz = b.q[3] * qn;
                const float tx = b.G[9], ty = b.G[10], tz = b.G[11];
                float dw = -tx * rx - ty * ry - tz * rz;
                float dx = tx * rw + ty * rz - tz * ry;
                float dy = -tx * rz + ty * rw + tz * rx;
                float dz = tx * ry - ty * rx + tz * rw;
                if (dw < 0) { dw = -dw; dx = -dx; dy = -dy; dz = -dz; }
                dr[0] += w * rw; dr[1] += w * rx; dr[2] += w * ry; dr[3] += w * rz;
                dd[0] += w * 0.5f * dw; dd[1] += w * 0.5f * dx; dd[2] += w * 0.5f * dy; dd[3] += w * 0.5f * dz;
            }
        }
        const float inv = 1.0f / wsum;
        if (a.nn_weight)
            for (int k = 0; k < K; k++) a.nn_weight[(size_t)n * K + k] *= inv;
        const float m = a.motion_mask ? a.motion_mask[n] : 1.0f;
        float ox = 0.f, oy = 0.f, oz = 0.f;
        if (a.offsets) { ox = a.offsets[3 * (size_t)n]; oy = a.offsets[3 * (size_t)n + 1]; oz = a.offsets[3 * (size_t)n + 2]; }
        if (a.blend_mode == 0) {
            a.d_xyz[3 * (size_t)n] = (ax * inv + g0 + ox - x) * m;
            a.d_xyz[3 * (size_t)n + 1] = (ay * inv + g1 + oy - y) * m;
            a.d_xyz[3 * (size_t)n + 2] = (az * inv + g2 + oz - z) * m;
            reinterpret_cast<float4*>(a.d_rot)[n] = {q0 * inv * m, q1 * inv * m, q2 * inv * m, q3 * inv * m};
        } else {
            // DQ2QT (utils/dual_quaternion.py:147-166) of the weight-normalised blend
            float rn = sqrtf(dr[0] * dr[0] + dr[1] * dr[1] + dr[2] * dr[2] + dr[3] * dr[3]) * inv;
            rn = fmaxf(rn, 1e-8f);
            const float sc = inv / rn;
            const float w0 = dr[0] * sc, x0 = dr[1] * sc, y0 = dr[2] * sc, z0 = dr[3] * sc;
            const float w1 = dd[0] * sc, x1 = dd[1] * sc, y1 = dd[2] * sc, z1 = dd[3] * sc;
            const float tx = 2 * (-w1 * x0 + x1 * w0 - y1 * z0 + z1 * y0);
            const float ty = 2 * (-w1 * y0 + x1 * z0 + y1 * w0 - z1 * x0);
            const float tz = 2 * (-w1 * z0 - x1 * y0 + y1 * x0 + z1 * w0);
            const float rx = (1 - 2 * y0 * y0 - 2 * z0 * z0) * x + (2 * x0 * y0 - 2 * w0 * z0) * y + (2 * x0 * z0 + 2 * w0 * y0) * z;
            const float ry = (2 * x0 * y0 + 2 * w0 * z0) * x + (1 - 2 * x0 * x0 - 2 * z0 * z0) * y + (2 * y0 * z0 - 2 * w0 * x0) * z;
            const float rz = (2 * x0 * z0 - 2 * w0 * y0) * x + (2 * y0 * z0 + 2 * w0 * x0) * y + (1 - 2 * x0 * x0 - 2 * y0 * y0) * z;
            a.d_xyz[3 * (size_t)n] = (rx + tx + g0 + ox - x) * m;
            a.d_xyz[3 * (size_t)n + 1] = (ry + ty + g1 + oy - y) * m;
            a.d_xyz[3 * (size_t)n + 2] = (rz + tz + g2 + oz - z) * m;
            reinterpret_cast<float4*>(a.d_rot)[n] = {w0 * m, x0 * m, y0 * m, z0 * m};
        }
    }
}

// Backward of the linear-blend forward.  Per Gaussian two sweeps over the bone table (weights are recomputed,
// nothing was stored): sweep 1 gets the normaliser and sum_k w_k dL/dw_k, sweep 2 emits the per-bone terms.
// Per-bone sums (dG 12 floats, d radius 1 float) are reduced warp-wide with a 16-shuffle transpose-reduce into the
// warp's own accumulator row in shared memory (no shared atomics) and flushed with one global atomicAdd per value per CTA; the grid is persistent (a few CTAs per SM), so
// a frame costs ~600 x 13(J-1) atomics instead of N x 13(J-1).
__global__ void __launch_bounds__(256) skin_backward_kernel(SkinBwdArgs a) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    BoneSm* s_b = reinterpret_cast<BoneSm*>(s_raw);
    const int K = a.J - 1;
    const int acc_stride = K * 13 + 3;  // one row per warp
    float* s_acc_all = reinterpret_cast<float*>(s_raw + sizeof(BoneSm) * (size_t)max(K, 1));
    float* s_acc = s_acc_all + (threadIdx.x >> 5) * acc_stride;
    stage_bones(a.J, a.joints, a.parents, a.node_radius_raw, a.G, a.node_rot, s_b);
    for (int i = threadIdx.x; i < 8 * acc_stride; i += blockDim.x) s_acc_all[i] = 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const float g0 = a.global_trans[0], g1 = a.global_trans[1], g2 = a.global_trans[2];
    const int n_iter = (a.N + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (int it = 0; it < n_iter; it++) {
        const int n = (it * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const bool live = n < a.N;
        float x = 0, y = 0, z = 0, m = 0, gx = 0, gy = 0, gz = 0, h0 = 0, h1 = 0, h2 = 0, h3 = 0;
        if (live) {
            x = a.xyz[3 * (size_t)n]; y = a.xyz[3 * (size_t)n + 1]; z = a.xyz[3 * (size_t)n + 2];
            m = a.motion_mask ? a.motion_mask[n] : 1.0f;
            gx = a.g_dxyz[3 * (size_t)n]; gy = a.g_dxyz[3 * (size_t)n + 1]; gz = a.g_dxyz[3 * (size_t)n + 2];
            if (a.g_drot) {
                const float4 h = reinterpret_cast<const float4*>(a.g_drot)[n];
                h0 = h.x; h1 = h.y; h2 = h.z; h3 = h.w;
            }
        }
        if (a.blend_mode == 1) {  // dual-quaternion blend: its own pair of sweeps (dq.cuh)
            float ox = 0.f, oy = 0.f, oz = 0.f;
            if (a.offsets && live) { ox = a.offsets[3 * (size_t)n]; oy = a.offsets[3 * (size_t)n + 1]; oz = a.offsets[3 * (size_t)n + 2]; }
            dq_skin_backward_point(s_b, K, lane, live, x, y, z, m, gx, gy, gz, h0, h1, h2, h3,
                                   a.mlp_weight ? a.mlp_weight + (size_t)(live ? n : 0) * K : nullptr, g0, g1, g2, ox, oy, oz, s_acc,
                                   (a.dmask && live) ? a.dmask + n : nullptr,
                                   (a.dmlp_weight && live) ? a.dmlp_weight + (size_t)n * K : nullptr);
            if (live && a.doffsets) {
                a.doffsets[3 * (size_t)n] = m * gx; a.doffsets[3 * (size_t)n + 1] = m * gy; a.doffsets[3 * (size_t)n + 2] = m * gz;
            }
            continue;
        }
        // sweep 1
        float wsum = 0.f, sc = 0.f, ax = 0.f, ay = 0.f, az = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
        for (int k = 0; k < K; k++) {
            const BoneSm& b = s_b[k];
            const float d2 = seg_dist2(b, x, y, z);
            float w = expf(-d2 * b.inv_2r2);
            if (a.mlp_weight && live) w *= a.mlp_weight[(size_t)n * K + k];
            w += 1e-7f;
            const float yx = b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9];
            const float yy = b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10];
            const float yz = b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11];
            const float c = gx * yx + gy * yy + gz * yz + h0 * b.q[0] + h1 * b.q[1] + h2 * b.q[2] + h3 * b.q[3];
            wsum += w;
            sc += w * c;
            ax += w * yx; ay += w * yy; az += w * yz;
            q0 += w * b.q[0]; q1 += w * b.q[1]; q2 += w * b.q[2]; q3 += w * b.q[3];
        }
        const float inv = 1.0f / wsum;
        const float cbar = sc * inv;  // sum_k w_k c_k with normalised weights
        if (live) {
            if (a.dmask)
                a.dmask[n] = gx * (ax * inv + g0 + (a.offsets ? a.offsets[3 * (size_t)n] : 0.f) - x) +
                             gy * (ay * inv + g1 + (a.offsets ? a.offsets[3 * (size_t)n + 1] : 0.f) - y) +
                             gz * (az * inv + g2 + (a.offsets ? a.offsets[3 * (size_t)n + 2] : 0.f) - z) +
                             (h0 * q0 + h1 * q1 + h2 * q2 + h3 * q3) * inv;
            if (a.doffsets) {
                a.doffsets[3 * (size_t)n] = m * gx; a.doffsets[3 * (size_t)n + 1] = m * gy; a.doffsets[3 * (size_t)n + 2] = m * gz;
            }
        }
        const float mgx = live ? m * gx : 0.f, mgy = live ? m * gy : 0.f, mgz = live ? m * gz : 0.f;
        // global_trans gradient: sum_n m a
        {
            const float s0 = warp_sum(mgx), s1 = warp_sum(mgy), s2 = warp_sum(mgz);
            if (lane == 0) { s_acc[K * 13] += s0; s_acc[K * 13 + 1] += s1; s_acc[K * 13 + 2] += s2; }
        }
        // sweep 2
        for (int k = 0; k < K; k++) {
            const BoneSm& b = s_b[k];
            const float d2 = seg_dist2(b, x, y, z);
            const float e = expf(-d2 * b.inv_2r2);
            const float s = (a.mlp_weight && live) ? a.mlp_weight[(size_t)n * K + k] : 1.0f;
            const float what = e * s + 1e-7f;
            const float w = what * inv;
            const float yx = b.G[0] * x + b.G[1] * y + b.G[2] * z + b.G[9];
            const float yy = b.G[3] * x + b.G[4] * y + b.G[5] * z + b.G[10];
            const float yz = b.G[6] * x + b.G[7] * y + b.G[8] * z + b.G[11];
            const float c = gx * yx + gy * yy + gz * yz + h0 * b.q[0] + h1 * b.q[1] + h2 * b.q[2] + h3 * b.q[3];
            const float dwhat = live ? m * (c - cbar) * inv : 0.f;  // dL/d what_k
            if (a.dmlp_weight && live) a.dmlp_weight[(size_t)n * K + k] = dwhat * e;
            float v[16];
            v[0] = mgx * w * x; v[1] = mgx * w * y; v[2] = mgx * w * z;
            v[3] = mgy * w * x; v[4] = mgy * w * y; v[5] = mgy * w * z;
            v[6] = mgz * w * x; v[7] = mgz * w * y; v[8] = mgz * w * z;
            v[9] = mgx * w; v[10] = mgy * w; v[11] = mgz * w;
            v[12] = dwhat * s * e * d2 * (2.0f * b.inv_2r2);  // d/d(log r): e * d2 / r^2
            v[13] = v[14] = v[15] = 0.f;
            const float r = transpose_reduce16(v, lane);  // the lane holds the total of v[(lane >> 1) & 15]
            const int my_id = (lane >> 1) & 15;
            if (!(lane & 1) && my_id < 13) s_acc[k * 13 + my_id] += r;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K * 13; i += blockDim.x) {
        const int k = i / 13, c = i % 13;
        float val = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) val += s_acc_all[w * acc_stride + i];
        if (val != 0.f) {
            if (c < 12) atomicAdd(&a.dG[12 * (k + 1) + c], val);
            else atomicAdd(&a.dnode_radius_raw[k + 1], val);
        }
    }
    if (threadIdx.x < 3 && a.dglobal_trans) {
        float val = 0.f;
#pragma unroll
        for (int w = 0; w < 8; w++) val += s_acc_all[w * acc_stride + K * 13 + threadIdx.x];
        atomicAdd(&a.dglobal_trans[threadIdx.x], val);
    }
}

int launch_fk_forward(int F, int J, const float* local_rot, const float* gtrans, const float* joints,
                      const int* parents, float* G, float* node_rot, float* posed, cudaStream_t stream) {
    if (F == 0 || J == 0) return 0;
    if (J > kMaxJoints) { set_error("forward kinematics: at most 128 joints"); return -1; }
    LaunchScope ls("fk_forward", stream);
    fk_forward_kernel<<<F, 32, 0, stream>>>(F, J, local_rot, gtrans, joints, parents, G, node_rot, posed);
    return 0;
}

int launch_fk_backward(int F, int J, const float* local_rot, const float* joints, const int* parents, const float* G,
                       const float* dG, const float* dposed, float* dlocal_rot, float* dgtrans, const float* dgtrans_add,
                       cudaStream_t stream, float* radius_dst, const float* radius_src, int radius_add) {
    if (F == 0 || J == 0) return 0;
    if (J > kMaxJoints) { set_error("forward kinematics: at most 128 joints"); return -1; }
    LaunchScope ls("fk_backward", stream);
    fk_backward_kernel<<<F, 32, 0, stream>>>(F, J, local_rot, joints, parents, G, dG, dposed, dlocal_rot, dgtrans,
                                            dgtrans_add, radius_dst, radius_src, radius_add);
    return 0;
}

int launch_skin_forward(const SkinArgs& a, cudaStream_t stream) {
    if (a.N == 0) return 0;
    if (a.J > kMaxJoints || a.J < 2) { set_error("skin: J must be in [2, 128]"); return -1; }
    int blocks = (a.N + 255) / 256;
    if (blocks > RG_NUM_SMS * 8) blocks = RG_NUM_SMS * 8;
    LaunchScope ls("skin_forward", stream);
    skin_forward_kernel<<<blocks, 256, bone_smem_bytes(a.J, false), stream>>>(a);
    return 0;
}

int launch_skin_backward(const SkinBwdArgs& a, cudaStream_t stream) {
    if (a.N == 0) return 0;
    if (a.J > kMaxJoints || a.J < 2) { set_error("skin: J must be in [2, 128]"); return -1; }
    int blocks = (a.N + 255) / 256;
    if (blocks > RG_NUM_SMS * 4) blocks = RG_NUM_SMS * 4;
    LaunchScope ls("skin_backward", stream);
    // J = 128: 13 KB of bones + 8 accumulator rows of 6.6 KB
    if (ensure_dyn_smem(reinterpret_cast<const void*>(skin_backward_kernel), 96 * 1024)) return -1;
    skin_backward_kernel<<<blocks, 256, bone_smem_bytes(a.J, true), stream>>>(a);
    return 0;
}

}  // namespace rg
