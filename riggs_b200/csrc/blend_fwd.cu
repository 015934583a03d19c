// Forward alpha blend: one CTA per 16x16 tile, one thread per pixel, front-to-back over the tile's depth-sorted
// instance range.  Replaces renderCUDA (dgr/cuda_rasterizer/forward.cu:261-381).
//
// Differences in HOW (results are the same):
//  * colour and depth of each staged instance are gathered once per tile into shared memory together with
//    xy / conic / opacity, instead of being fetched from global memory per contributing (pixel, instance) pair;
//  * a per-instance log-threshold rejects pairs whose alpha would be < 1/255 before evaluating expf.  The test
//    is conservative (margin far above the rounding error of expf/logf); every pair it lets through takes the
//    reference's exact expression, so blended values and n_contrib are unchanged.
#include "common.cuh"
#include "kernels.cuh"

namespace rg {

constexpr float kSkipMargin = 1e-4f;

__global__ void __launch_bounds__(RG_TILE_PIX) blend_forward_kernel(BlendFwdArgs a) {
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const uint2 pix = {blockIdx.x * RG_TILE + tx, blockIdx.y * RG_TILE + ty};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    bool done = !inside;

    const uint2 range = a.ranges[blockIdx.y * a.grid_x + blockIdx.x];
    const int rounds = ((range.y - range.x + RG_TILE_PIX - 1) / RG_TILE_PIX);
    int toDo = range.y - range.x;

    __shared__ float2 s_xy[RG_TILE_PIX];
    __shared__ float4 s_co[RG_TILE_PIX];
    __shared__ float4 s_rgbd[RG_TILE_PIX];
    __shared__ float s_thr[RG_TILE_PIX];

    float T = 1.0f;
    uint32_t contributor = 0, last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;

    for (int i = 0; i < rounds; i++, toDo -= RG_TILE_PIX) {
        int num_done = __syncthreads_count(done);
        if (num_done == RG_TILE_PIX) break;
        int progress = i * RG_TILE_PIX + threadIdx.x;
        if (range.x + progress < range.y) {
            const uint32_t id = a.point_list[range.x + progress];
            s_xy[threadIdx.x] = a.means2D[id];
            const float4 co = a.conic_opacity[id];
            s_co[threadIdx.x] = co;
            s_rgbd[threadIdx.x] = {a.colors[3 * id], a.colors[3 * id + 1], a.colors[3 * id + 2], a.depths[id]};
            // alpha = o*exp(power) < 1/255  <=>  power < -log(255*o)
            s_thr[threadIdx.x] = -logf(255.0f * co.w) - kSkipMargin;
        }
        __syncthreads();
        const int n = min(RG_TILE_PIX, toDo);
        for (int j = 0; !done && j < n; j++) {
            contributor++;
            const float2 xy = s_xy[j];
            const float2 d = {xy.x - pixf.x, xy.y - pixf.y};
            const float4 con_o = s_co[j];
            const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
            if (power > 0.0f) continue;
            if (power < s_thr[j]) continue;  // certainly alpha < 1/255
            const float alpha = fminf(0.99f, con_o.w * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) {
                done = true;
                continue;
            }
            const float4 c = s_rgbd[j];
            C0 += c.x * alpha * T;
            C1 += c.y * alpha * T;
            C2 += c.z * alpha * T;
            weight += alpha * T;
            D += c.w * alpha * T;
            T = test_T;
            last_contributor = contributor;
        }
    }
    if (inside) {
        const size_t HW = (size_t)a.H * a.W;
        a.n_contrib[pix_id] = last_contributor;
        a.out_color[pix_id] = C0 + T * a.bg[0];
        a.out_color[HW + pix_id] = C1 + T * a.bg[1];
        a.out_color[2 * HW + pix_id] = C2 + T * a.bg[2];
        a.out_alpha[pix_id] = weight;  // sum of alpha*T, as the reference (forward.cu:378)
        a.out_depth[pix_id] = D;
    }
}

int launch_blend_forward(const BlendFwdArgs& a, cudaStream_t stream) {
    dim3 grid(a.grid_x, a.grid_y, 1);
    LaunchScope ls("blend_forward", stream);
    blend_forward_kernel<<<grid, RG_TILE_PIX, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
