// K nearest control nodes of every point: the brute-force search the reference gets from pytorch3d.ops.knn_points
// (utils/time_utils.py:950, :980, :1051, :1061, :1124 -- K = 1..9 of 512-1024 nodes, 3 + hyper dimensions; pytorch3d
// is a git-main dependency of the reference and absent here, so this restates its documented contract: squared
// Euclidean distances, the K smallest in ascending order, ties to the lower index).  SURVEY section 8f, N4.
//
// One thread per query point; the nodes are staged in shared memory in chunks and every thread keeps its K best in
// registers (insertion into a sorted array, K a template parameter).  N x M x D multiply-adds: 300k points x 1024 nodes
// x 3 = 0.9 GFLOP, a few microseconds; the reference path allocates an (N, M) distance matrix per call.
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kKnnChunk = 1024;  // nodes staged per pass
constexpr int kKnnMaxD = 8;

template <int K, int D>
__global__ void __launch_bounds__(256) knn_points_kernel(int N, int M, const float* __restrict__ pts,
                                                         const float* __restrict__ nodes, int k_out,
                                                         float* __restrict__ dists, long long* __restrict__ idx) {
    extern __shared__ float s_nodes[];  // [chunk][D]
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    float p[D];
#pragma unroll
    for (int c = 0; c < D; c++) p[c] = n < N ? pts[(size_t)n * D + c] : 0.f;
    float bd[K];
    int bi[K];
#pragma unroll
    for (int s = 0; s < K; s++) { bd[s] = 3.0e38f; bi[s] = -1; }
    for (int m0 = 0; m0 < M; m0 += kKnnChunk) {
        const int mc = min(kKnnChunk, M - m0);
        __syncthreads();
        for (int k = threadIdx.x; k < mc * D; k += blockDim.x) s_nodes[k] = nodes[(size_t)m0 * D + k];
        __syncthreads();
        for (int j = 0; j < mc; j++) {
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < D; c++) {
                const float t = p[c] - s_nodes[j * D + c];
                d = fmaf(t, t, d);
            }
            if (d < bd[K - 1]) {  // strict: among equal distances the lower index stays in front
                bd[K - 1] = d;
                bi[K - 1] = m0 + j;
#pragma unroll
                for (int s = K - 1; s > 0; s--) {
                    if (bd[s] < bd[s - 1]) {
                        const float td = bd[s]; bd[s] = bd[s - 1]; bd[s - 1] = td;
                        const int ti = bi[s]; bi[s] = bi[s - 1]; bi[s - 1] = ti;
                    }
                }
            }
        }
    }
    if (n < N) {
#pragma unroll
        for (int s = 0; s < K; s++) {
            if (s < k_out) {
                // fewer nodes than K: pytorch3d pads distances with 0 and indices with 0 beyond the node count
                dists[(size_t)n * k_out + s] = bi[s] >= 0 ? bd[s] : 0.f;
                idx[(size_t)n * k_out + s] = bi[s] >= 0 ? bi[s] : 0;
            }
        }
    }
}

template <int K>
static int launch_knn_d(int N, int M, int D, const float* pts, const float* nodes, int k_out, float* dists, long long* idx,
                        cudaStream_t stream) {
    const int blocks = (N + 255) / 256;
    const size_t smem = sizeof(float) * (size_t)kKnnChunk * D;
    switch (D) {
        case 1: knn_points_kernel<K, 1><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 2: knn_points_kernel<K, 2><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 3: knn_points_kernel<K, 3><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 4: knn_points_kernel<K, 4><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 5: knn_points_kernel<K, 5><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 6: knn_points_kernel<K, 6><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 7: knn_points_kernel<K, 7><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        case 8: knn_points_kernel<K, 8><<<blocks, 256, smem, stream>>>(N, M, pts, nodes, k_out, dists, idx); break;
        default: return -1;
    }
    return 0;
}

}  // namespace rg

using namespace rg;

extern "C" int rg_knn_points(int N, int M, int D, int K, const float* points, const float* nodes, float* dists, long long* idx,
                             rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (N < 0 || M < 1 || D < 1 || D > kKnnMaxD || K < 1 || K > 16) {
        set_error("rg_knn_points: need M >= 1, 1 <= D <= 8, 1 <= K <= 16");
        return -1;
    }
    if (N == 0) return 0;
    if (!points || !nodes || !dists || !idx) { set_error("rg_knn_points: null argument"); return -1; }
    LaunchScope ls("knn_points", stream);
    int rc;
    if (K <= 1) rc = launch_knn_d<1>(N, M, D, points, nodes, K, dists, idx, stream);
    else if (K <= 4) rc = launch_knn_d<4>(N, M, D, points, nodes, K, dists, idx, stream);
    else if (K <= 9) rc = launch_knn_d<9>(N, M, D, points, nodes, K, dists, idx, stream);
    else rc = launch_knn_d<16>(N, M, D, points, nodes, K, dists, idx, stream);
    if (rc) { set_error("rg_knn_points: unsupported dimension"); return -1; }
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
