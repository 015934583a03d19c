// Stage-1 control-node warp: the K nearest control nodes of every Gaussian, their
// radial-basis weights and the blend of the nodes' motion onto the Gaussian, forward and backward.
//
// Replaces ControlNodeWarp.cal_nn_weight (utils/time_utils.py:934-967: pytorch3d knn_points over the (N, M)
// distance matrix, exp(-d / 2r^2) * sigmoid(node weight) + 1e-7, normalised) and the blending part of
// ControlNodeWarp.forward (:1133-1198, 1213-1224: local-frame LBS about each node, weighted sums of the nodes'
// rotation / scaling / opacity / colour residuals, motion mask) -- about 25 eager kernels with (N, K, 3, 3) gathers --
// by one streaming kernel per direction.  The node table (M <= a few thousand rows of 3 + hyper floats) is searched
// from shared memory; the K best stay in registers; nothing of size (N, M) or (N, K, 3, 3) is materialised.
//
// Backward: one thread per Gaussian recomputes the un-normalised weights from the saved distances; the per-node
// sums (translation 3, rotation 4, scaling 3, local quaternion 4, radius 1, node weight 1, opacity 1, colour 3,
// hyper coordinates <= 5) are accumulated in a shared-memory table by a persistent grid and flushed with one
// global atomic per value per CTA (global atomics directly when the table does not fit).
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kNwMaxD = 8;       // 3 + hyper dimensions of the search space
constexpr int kNwChunk = 1024;   // nodes staged per pass of the search
constexpr int kNwChannels = 25;  // row of the per-node gradient table (layout below)
// gradient table row: [0..2] d_xyz | [3..6] d_rotation | [7..9] d_scaling | [10..13] local quaternion |
//                     [14] radius_raw | [15] node_weight_raw | [16] d_opacity | [17..19] d_color | [20..24] hyper coords
constexpr int kChT = 0, kChR = 3, kChS = 7, kChQ = 10, kChRad = 14, kChW = 15, kChO = 16, kChC = 17, kChH = 20;

struct NodeWarpArgs {
    int N, M, K, D;  // D = 3 + hyper
    const float* x;            // (N,3)
    const float* feature;      // (N,D-3) or null when D == 3
    const float* nodes;        // (M,D)
    const float* radius_raw;   // (M)
    const float* weight_raw;   // (M) or null
    const float* local_quat;   // (M,4) un-normalised (local_rotation + [1,0,0,0]) or null: no local frame
    const float* node_trans;   // (M,3)
    const float* node_rot;     // (M,4)
    const float* node_scale;   // (M,3)
    const float* node_opacity; // (M) or null
    const float* node_color;   // (M,3) or null
    const float* mask;         // (N) or null
    int add_rot_bias;          // d_rot_as_res == False: rotation = blend * mask + [1,0,0,0]
};

__device__ __forceinline__ void nw_quat_to_mat(const float* q, float* R) {  // utils/time_utils.py:115-132
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

// dL/dq from dL/dR (row-major D[3a+b]) for R = I + two_s * A(q), two_s = 2 / |q|^2.
__device__ __forceinline__ void nw_quat_to_mat_bwd(const float* q, const float* D, float* dq) {
    const float r = q[0], i = q[1], j = q[2], k = q[3];
    const float n = r * r + i * i + j * j + k * k;
    const float two_s = 2.0f / n;
    const float AD = -(j * j + k * k) * D[0] + (i * j - k * r) * D[1] + (i * k + j * r) * D[2] + (i * j + k * r) * D[3] -
                     (i * i + k * k) * D[4] + (j * k - i * r) * D[5] + (i * k - j * r) * D[6] + (j * k + i * r) * D[7] -
                     (i * i + j * j) * D[8];
    const float c = -two_s * two_s * AD;  // d two_s / dq = -two_s^2 q
    dq[0] = two_s * (-k * D[1] + j * D[2] + k * D[3] - i * D[5] - j * D[6] + i * D[7]) + c * r;
    dq[1] = two_s * (j * D[1] + k * D[2] + j * D[3] - 2 * i * D[4] - r * D[5] + k * D[6] + r * D[7] - 2 * i * D[8]) + c * i;
    dq[2] = two_s * (-2 * j * D[0] + i * D[1] + r * D[2] + i * D[3] + k * D[5] - r * D[6] + k * D[7] - 2 * j * D[8]) + c * j;
    dq[3] = two_s * (-2 * k * D[0] - r * D[1] + i * D[2] + r * D[3] - 2 * k * D[4] + j * D[5] + i * D[6] + j * D[7]) + c * k;
}

template <int KMAX>
__global__ void __launch_bounds__(256) node_warp_forward_kernel(NodeWarpArgs a, float* __restrict__ d_xyz,
                                                                float* __restrict__ d_rot, float* __restrict__ d_scale,
                                                                float* __restrict__ d_opacity, float* __restrict__ d_color,
                                                                long long* __restrict__ nn_idx, float* __restrict__ nn_weight,
                                                                float* __restrict__ nn_dist) {
    extern __shared__ float s_nodes[];  // [chunk][D]
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = n < a.N;
    const int D = a.D, K = a.K;
    float p[kNwMaxD];
#pragma unroll
    for (int c = 0; c < kNwMaxD; c++) {
        p[c] = 0.f;
        if (live && c < D) p[c] = c < 3 ? a.x[(size_t)n * 3 + c] : a.feature[(size_t)n * (D - 3) + (c - 3)];
    }
    float bd[KMAX];
    int bi[KMAX];
#pragma unroll
    for (int s = 0; s < KMAX; s++) { bd[s] = 3.0e38f; bi[s] = -1; }
    for (int m0 = 0; m0 < a.M; m0 += kNwChunk) {
        const int mc = min(kNwChunk, a.M - m0);
        __syncthreads();
        for (int k = threadIdx.x; k < mc * D; k += blockDim.x) s_nodes[k] = a.nodes[(size_t)m0 * D + k];
        __syncthreads();
        for (int j = 0; j < mc; j++) {
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < kNwMaxD; c++) {
                if (c < D) {
                    const float t = p[c] - s_nodes[j * D + c];
                    d = fmaf(t, t, d);
                }
            }
            if (d < bd[KMAX - 1]) {  // strict: among equal distances the lower index stays in front
                bd[KMAX - 1] = d;
                bi[KMAX - 1] = m0 + j;
#pragma unroll
                for (int s = KMAX - 1; s > 0; s--) {
                    if (bd[s] < bd[s - 1]) {
                        const float td = bd[s]; bd[s] = bd[s - 1]; bd[s - 1] = td;
                        const int ti = bi[s]; bi[s] = bi[s - 1]; bi[s - 1] = ti;
                    }
                }
            }
        }
    }
    if (!live) return;
    float w[KMAX];
    float sumw = 0.f;
#pragma unroll
    for (int s = 0; s < KMAX; s++) {
        w[s] = 0.f;
        if (s < K) {
            const int m = bi[s];
            const float r = expf(a.radius_raw[m]);
            float e = expf(-bd[s] / (2.0f * r * r));
            if (a.weight_raw) e *= 1.0f / (1.0f + expf(-a.weight_raw[m]));
            w[s] = e + 1e-7f;
            sumw += w[s];
        }
    }
    const float inv = 1.0f / sumw;
    float tx = 0.f, ty = 0.f, tz = 0.f, q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f, op = 0.f,
          c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int s = 0; s < KMAX; s++) {
        if (s < K) {
            const int m = bi[s];
            const float ws = w[s] * inv;
            nn_idx[(size_t)n * K + s] = m;
            nn_weight[(size_t)n * K + s] = ws;
            nn_dist[(size_t)n * K + s] = bd[s];
            float y0 = a.node_trans[3 * m], y1 = a.node_trans[3 * m + 1], y2 = a.node_trans[3 * m + 2];
            if (a.local_quat) {
                float R[9];
                nw_quat_to_mat(a.local_quat + 4 * m, R);
                const float cx = a.nodes[(size_t)m * D], cy = a.nodes[(size_t)m * D + 1], cz = a.nodes[(size_t)m * D + 2];
                const float ux = p[0] - cx, uy = p[1] - cy, uz = p[2] - cz;
                y0 += R[0] * ux + R[1] * uy + R[2] * uz + cx;
                y1 += R[3] * ux + R[4] * uy + R[5] * uz + cy;
                y2 += R[6] * ux + R[7] * uy + R[8] * uz + cz;
            }
            tx = fmaf(ws, y0, tx); ty = fmaf(ws, y1, ty); tz = fmaf(ws, y2, tz);
            q0 = fmaf(ws, a.node_rot[4 * m], q0); q1 = fmaf(ws, a.node_rot[4 * m + 1], q1);
            q2 = fmaf(ws, a.node_rot[4 * m + 2], q2); q3 = fmaf(ws, a.node_rot[4 * m + 3], q3);
            s0 = fmaf(ws, a.node_scale[3 * m], s0); s1 = fmaf(ws, a.node_scale[3 * m + 1], s1);
            s2 = fmaf(ws, a.node_scale[3 * m + 2], s2);
            if (a.node_opacity) op = fmaf(ws, a.node_opacity[m], op);
            if (a.node_color) {
                c0 = fmaf(ws, a.node_color[3 * m], c0); c1 = fmaf(ws, a.node_color[3 * m + 1], c1);
                c2 = fmaf(ws, a.node_color[3 * m + 2], c2);
            }
        }
    }
    if (a.local_quat) { tx -= p[0]; ty -= p[1]; tz -= p[2]; }
    const float mk = a.mask ? a.mask[n] : 1.0f;
    d_xyz[(size_t)n * 3] = tx * mk; d_xyz[(size_t)n * 3 + 1] = ty * mk; d_xyz[(size_t)n * 3 + 2] = tz * mk;
    d_rot[(size_t)n * 4] = q0 * mk + (a.add_rot_bias ? 1.0f : 0.0f);
    d_rot[(size_t)n * 4 + 1] = q1 * mk; d_rot[(size_t)n * 4 + 2] = q2 * mk; d_rot[(size_t)n * 4 + 3] = q3 * mk;
    d_scale[(size_t)n * 3] = s0 * mk; d_scale[(size_t)n * 3 + 1] = s1 * mk; d_scale[(size_t)n * 3 + 2] = s2 * mk;
    if (d_opacity) d_opacity[n] = op * mk;
    if (d_color) { d_color[(size_t)n * 3] = c0 * mk; d_color[(size_t)n * 3 + 1] = c1 * mk; d_color[(size_t)n * 3 + 2] = c2 * mk; }
}

template <int KMAX>
__global__ void __launch_bounds__(256) node_warp_backward_kernel(NodeWarpArgs a, const long long* __restrict__ nn_idx,
                                                                 const float* __restrict__ nn_dist,
                                                                 const float* __restrict__ g_xyz, const float* __restrict__ g_rot,
                                                                 const float* __restrict__ g_scale,
                                                                 const float* __restrict__ g_opacity,
                                                                 const float* __restrict__ g_color, int use_smem,
                                                                 float* __restrict__ dnode,  // (M, kNwChannels), accumulated
                                                                 float* __restrict__ dmask, float* __restrict__ dfeature) {
    extern __shared__ float s_acc[];  // [M][kNwChannels] when use_smem
    const int D = a.D, K = a.K, H = a.D - 3;
    float* acc = use_smem ? s_acc : dnode;
    if (use_smem) {
        for (int i = threadIdx.x; i < a.M * kNwChannels; i += blockDim.x) s_acc[i] = 0.f;
        __syncthreads();
    }
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += gridDim.x * blockDim.x) {
        const float px = a.x[(size_t)n * 3], py = a.x[(size_t)n * 3 + 1], pz = a.x[(size_t)n * 3 + 2];
        const float mk = a.mask ? a.mask[n] : 1.0f;
        float ga[3] = {0.f, 0.f, 0.f}, gh[4] = {0.f, 0.f, 0.f, 0.f}, gs[3] = {0.f, 0.f, 0.f}, go = 0.f, gc[3] = {0.f, 0.f, 0.f};
        if (g_xyz) { ga[0] = g_xyz[(size_t)n * 3]; ga[1] = g_xyz[(size_t)n * 3 + 1]; ga[2] = g_xyz[(size_t)n * 3 + 2]; }
        if (g_rot) { gh[0] = g_rot[(size_t)n * 4]; gh[1] = g_rot[(size_t)n * 4 + 1]; gh[2] = g_rot[(size_t)n * 4 + 2]; gh[3] = g_rot[(size_t)n * 4 + 3]; }
        if (g_scale) { gs[0] = g_scale[(size_t)n * 3]; gs[1] = g_scale[(size_t)n * 3 + 1]; gs[2] = g_scale[(size_t)n * 3 + 2]; }
        if (g_opacity && a.node_opacity) go = g_opacity[n];
        if (g_color && a.node_color) { gc[0] = g_color[(size_t)n * 3]; gc[1] = g_color[(size_t)n * 3 + 1]; gc[2] = g_color[(size_t)n * 3 + 2]; }
        int idx[KMAX];
        float what[KMAX], es[KMAX], sg[KMAX], r2[KMAX], dist[KMAX], dw[KMAX];
        float sumw = 0.f;
#pragma unroll
        for (int s = 0; s < KMAX; s++) {
            idx[s] = 0; what[s] = 0.f; es[s] = 0.f; sg[s] = 1.f; r2[s] = 1.f; dist[s] = 0.f; dw[s] = 0.f;
            if (s < K) {
                const int m = (int)nn_idx[(size_t)n * K + s];
                idx[s] = m;
                dist[s] = nn_dist[(size_t)n * K + s];
                const float r = expf(a.radius_raw[m]);
                r2[s] = r * r;
                float e = expf(-dist[s] / (2.0f * r2[s]));
                if (a.weight_raw) { sg[s] = 1.0f / (1.0f + expf(-a.weight_raw[m])); e *= sg[s]; }
                es[s] = e;
                what[s] = e + 1e-7f;
                sumw += what[s];
            }
        }
        const float inv = 1.0f / sumw;
        float bt[3] = {0.f, 0.f, 0.f};  // blended, un-masked outputs (for d mask)
        float bq = 0.f;                 // sum over the other outputs of cotangent . blended value
        float dot = 0.f;
#pragma unroll
        for (int s = 0; s < KMAX; s++) {
            if (s < K) {
                const int m = idx[s];
                const float ws = what[s] * inv;
                float* row = acc + (size_t)m * kNwChannels;
                float y[3] = {a.node_trans[3 * m], a.node_trans[3 * m + 1], a.node_trans[3 * m + 2]};
                const float wm = ws * mk;
                if (a.local_quat) {
                    float R[9], Dm[9], dq[4];
                    nw_quat_to_mat(a.local_quat + 4 * m, R);
                    const float cx = a.nodes[(size_t)m * D], cy = a.nodes[(size_t)m * D + 1], cz = a.nodes[(size_t)m * D + 2];
                    const float u[3] = {px - cx, py - cy, pz - cz};
                    y[0] += R[0] * u[0] + R[1] * u[1] + R[2] * u[2] + cx;
                    y[1] += R[3] * u[0] + R[4] * u[1] + R[5] * u[2] + cy;
                    y[2] += R[6] * u[0] + R[7] * u[1] + R[8] * u[2] + cz;
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 3; c++) Dm[3 * r + c] = wm * ga[r] * u[c];
                    nw_quat_to_mat_bwd(a.local_quat + 4 * m, Dm, dq);
#pragma unroll
                    for (int c = 0; c < 4; c++) atomicAdd(row + kChQ + c, dq[c]);
                }
                float v = ga[0] * y[0] + ga[1] * y[1] + ga[2] * y[2];
#pragma unroll
                for (int c = 0; c < 3; c++) { bt[c] = fmaf(ws, y[c], bt[c]); atomicAdd(row + kChT + c, wm * ga[c]); }
                float o = 0.f;
#pragma unroll
                for (int c = 0; c < 4; c++) { o = fmaf(gh[c], a.node_rot[4 * m + c], o); atomicAdd(row + kChR + c, wm * gh[c]); }
#pragma unroll
                for (int c = 0; c < 3; c++) { o = fmaf(gs[c], a.node_scale[3 * m + c], o); atomicAdd(row + kChS + c, wm * gs[c]); }
                if (a.node_opacity) { o = fmaf(go, a.node_opacity[m], o); atomicAdd(row + kChO, wm * go); }
                if (a.node_color) {
#pragma unroll
                    for (int c = 0; c < 3; c++) { o = fmaf(gc[c], a.node_color[3 * m + c], o); atomicAdd(row + kChC + c, wm * gc[c]); }
                }
                bq = fmaf(ws, o, bq);
                dw[s] = mk * (v + o);
                dot = fmaf(ws, dw[s], dot);
            }
        }
        float df[kNwMaxD - 3];
#pragma unroll
        for (int h = 0; h < kNwMaxD - 3; h++) df[h] = 0.f;
#pragma unroll
        for (int s = 0; s < KMAX; s++) {
            if (s < K) {
                const int m = idx[s];
                float* row = acc + (size_t)m * kNwChannels;
                const float dwh = (dw[s] - dot) * inv;  // d loss / d (un-normalised weight)
                const float t = dwh * es[s];
                atomicAdd(row + kChRad, t * dist[s] / r2[s]);
                if (a.weight_raw) atomicAdd(row + kChW, t * (1.0f - sg[s]));
                const float dd = -t / (2.0f * r2[s]);  // d loss / d (squared distance)
#pragma unroll
                for (int h = 0; h < kNwMaxD - 3; h++) {
                    if (h < H) {
                        const float diff = a.feature[(size_t)n * H + h] - a.nodes[(size_t)m * D + 3 + h];
                        const float g = 2.0f * dd * diff;
                        df[h] += g;
                        atomicAdd(row + kChH + h, -g);
                    }
                }
            }
        }
        if (dfeature) {
#pragma unroll
            for (int h = 0; h < kNwMaxD - 3; h++)
                if (h < H) dfeature[(size_t)n * H + h] = df[h];
        }
        if (dmask) {
            float t0 = bt[0], t1 = bt[1], t2 = bt[2];
            if (a.local_quat) { t0 -= px; t1 -= py; t2 -= pz; }
            dmask[n] = ga[0] * t0 + ga[1] * t1 + ga[2] * t2 + bq;
        }
    }
    if (use_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < a.M * kNwChannels; i += blockDim.x) {
            const float v = s_acc[i];
            if (v != 0.f) atomicAdd(dnode + i, v);
        }
    }
}

}  // namespace rg

using namespace rg;

static int nw_check(const char* who, int N, int M, int K, int hyper, const void* x, const void* feature, const void* nodes,
                    const void* radius, const void* nt, const void* nr, const void* ns) {
    if (N < 0 || M < 1 || K < 1 || K > 8 || hyper < 0 || hyper > kNwMaxD - 3) {
        set_error(std::string(who) + ": need M >= 1, 1 <= K <= 8, 0 <= hyper <= 5");
        return -1;
    }
    if (K > M) { set_error(std::string(who) + ": K exceeds the number of nodes"); return -1; }
    if (N > 0 && (!x || !nodes || !radius || !nt || !nr || !ns || (hyper > 0 && !feature))) {
        set_error(std::string(who) + ": null argument");
        return -1;
    }
    return 0;
}

extern "C" int rg_node_warp_forward(int N, int M, int K, int hyper, const float* x, const float* feature, const float* nodes,
                                    const float* node_radius_raw, const float* node_weight_raw, const float* local_quat,
                                    const float* node_trans, const float* node_rot, const float* node_scale,
                                    const float* node_opacity, const float* node_color, const float* motion_mask,
                                    int add_rot_bias, float* d_xyz, float* d_rot, float* d_scale, float* d_opacity,
                                    float* d_color, long long* nn_idx, float* nn_weight, float* nn_dist, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (nw_check("rg_node_warp_forward", N, M, K, hyper, x, feature, nodes, node_radius_raw, node_trans, node_rot, node_scale))
        return -1;
    if (N == 0) return 0;
    if (!d_xyz || !d_rot || !d_scale || !nn_idx || !nn_weight || !nn_dist) { set_error("rg_node_warp_forward: null output"); return -1; }
    NodeWarpArgs a{N, M, K, 3 + hyper, x, feature, nodes, node_radius_raw, node_weight_raw, local_quat, node_trans,
                   node_rot, node_scale, node_opacity, node_color, motion_mask, add_rot_bias};
    LaunchScope ls("node_warp_forward", stream);
    const int blocks = (N + 255) / 256;
    const size_t smem = sizeof(float) * (size_t)kNwChunk * a.D;
    if (K <= 4)
        node_warp_forward_kernel<4><<<blocks, 256, smem, stream>>>(a, d_xyz, d_rot, d_scale, node_opacity ? d_opacity : nullptr,
                                                                   node_color ? d_color : nullptr, nn_idx, nn_weight, nn_dist);
    else
        node_warp_forward_kernel<8><<<blocks, 256, smem, stream>>>(a, d_xyz, d_rot, d_scale, node_opacity ? d_opacity : nullptr,
                                                                   node_color ? d_color : nullptr, nn_idx, nn_weight, nn_dist);
    if (cudaGetLastError() != cudaSuccess) { set_error("rg_node_warp_forward: launch failed"); return -1; }
    return 0;
}

extern "C" int rg_node_warp_grad_channels(void) { return kNwChannels; }

extern "C" int rg_node_warp_backward(int N, int M, int K, int hyper, const float* x, const float* feature, const float* nodes,
                                     const float* node_radius_raw, const float* node_weight_raw, const float* local_quat,
                                     const float* node_trans, const float* node_rot, const float* node_scale,
                                     const float* node_opacity, const float* node_color, const float* motion_mask,
                                     const long long* nn_idx, const float* nn_dist, const float* g_xyz, const float* g_rot,
                                     const float* g_scale, const float* g_opacity, const float* g_color, float* dnode,
                                     float* dmask, float* dfeature, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (nw_check("rg_node_warp_backward", N, M, K, hyper, x, feature, nodes, node_radius_raw, node_trans, node_rot, node_scale))
        return -1;
    if (N == 0) return 0;
    if (!nn_idx || !nn_dist || !dnode) { set_error("rg_node_warp_backward: null argument"); return -1; }
    NodeWarpArgs a{N, M, K, 3 + hyper, x, feature, nodes, node_radius_raw, node_weight_raw, local_quat, node_trans,
                   node_rot, node_scale, node_opacity, node_color, motion_mask, 0};
    LaunchScope ls("node_warp_backward", stream);
    const size_t table = sizeof(float) * (size_t)M * kNwChannels;
    const int use_smem = table <= 100 * 1024;
    const size_t smem = use_smem ? table : 0;
    const int blocks = std::min((N + 255) / 256, RG_NUM_SMS * 2);
    auto k4 = node_warp_backward_kernel<4>;
    auto k8 = node_warp_backward_kernel<8>;
    if (use_smem && smem > 48 * 1024)  // per device and function; cheap enough to repeat
        cudaFuncSetAttribute(K <= 4 ? k4 : k8, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (K <= 4)
        k4<<<blocks, 256, smem, stream>>>(a, nn_idx, nn_dist, g_xyz, g_rot, g_scale, g_opacity, g_color, use_smem, dnode, dmask, dfeature);
    else
        k8<<<blocks, 256, smem, stream>>>(a, nn_idx, nn_dist, g_xyz, g_rot, g_scale, g_opacity, g_color, use_smem, dnode, dmask, dfeature);
    if (cudaGetLastError() != cudaSuccess) { set_error("rg_node_warp_backward: launch failed"); return -1; }
    return 0;
}
