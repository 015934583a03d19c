// Backward alpha blend: one CTA per tile, one thread per pixel, back-to-front replay.
// Replaces renderCUDA of dgr/cuda_rasterizer/backward.cu:415-601.
//
// The reference issues 10 global float atomicAdds per contributing (pixel, instance) pair.  Here the ten
// per-pair terms are summed across the warp with shuffles (only when some lane of the warp contributes),
// the eight warps of the CTA meet in a shared-memory accumulator, and each staged instance is flushed to
// global memory ONCE per tile: 10 atomics per (tile, instance) instead of up to 2560.
// The replay also starts at the tile's largest n_contrib instead of the end of the range: entries behind
// every pixel's last contributor are skipped by the reference one by one (backward.cu:517-519) and never
// contribute.
#include "common.cuh"
#include "kernels.cuh"

namespace rg {

constexpr float kSkipMarginB = 1e-4f;
constexpr int kAcc = 10;  // mean2D.x, mean2D.y, conic.x, conic.y, conic.w, opacity, r, g, b, depth

__global__ void __launch_bounds__(RG_TILE_PIX) blend_backward_kernel(BlendBwdArgs a) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int tx = tid & 15, ty = tid >> 4;
    const uint2 pix = {blockIdx.x * RG_TILE + tx, blockIdx.y * RG_TILE + ty};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    const uint2 range = a.ranges[blockIdx.y * a.grid_x + blockIdx.x];

    __shared__ uint32_t s_id[RG_TILE_PIX];
    __shared__ float2 s_xy[RG_TILE_PIX];
    __shared__ float4 s_co[RG_TILE_PIX];
    __shared__ float4 s_rgbd[RG_TILE_PIX];
    __shared__ float s_thr[RG_TILE_PIX];
    __shared__ float s_acc[RG_TILE_PIX * kAcc];
    __shared__ int s_max;

    const float T_final = inside ? (1 - a.alphas[pix_id]) : 0;
    float T = T_final;
    const int last_contributor = inside ? a.n_contrib[pix_id] : 0;

    // Tile-wide replay start: the largest last_contributor of the tile.
    if (tid == 0) s_max = 0;
    __syncthreads();
    int wmax = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) atomicMax(&s_max, wmax);
    __syncthreads();
    const int total = s_max;  // instances [range.x, range.x + total) can contribute
    if (total == 0) return;

    float accum_rec[3] = {0, 0, 0}, dL_dpixel[3] = {0, 0, 0};
    float accum_depth_rec = 0, dL_dpixel_depth = 0, accum_alpha_rec = 0, dL_dalpha = 0;
    if (inside) {
        const size_t HW = (size_t)a.H * a.W;
#pragma unroll
        for (int c = 0; c < 3; c++) dL_dpixel[c] = a.dL_dpix[c * HW + pix_id];
        dL_dpixel_depth = a.dL_dpix_depth[pix_id];
        dL_dalpha = a.dL_dalphas[pix_id];
    }
    float last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0;
    const float ddelx_dx = 0.5 * a.W;
    const float ddely_dy = 0.5 * a.H;
    const float bg_dot_dpixel = a.bg[0] * dL_dpixel[0] + a.bg[1] * dL_dpixel[1] + a.bg[2] * dL_dpixel[2];

    const int rounds = (total + RG_TILE_PIX - 1) / RG_TILE_PIX;
    int toDo = total;
    uint32_t contributor = total;
    for (int i = 0; i < rounds; i++, toDo -= RG_TILE_PIX) {
        __syncthreads();
        const int progress = i * RG_TILE_PIX + tid;
        if (progress < total) {
            const uint32_t id = a.point_list[range.x + total - progress - 1];
            s_id[tid] = id;
            s_xy[tid] = a.means2D[id];
            const float4 co = a.conic_opacity[id];
            s_co[tid] = co;
            s_rgbd[tid] = {a.colors[3 * id], a.colors[3 * id + 1], a.colors[3 * id + 2], a.depths[id]};
            s_thr[tid] = -logf(255.0f * co.w) - kSkipMarginB;
        }
#pragma unroll
        for (int k = 0; k < kAcc; k++) s_acc[k * RG_TILE_PIX + tid] = 0.f;
        __syncthreads();
        const int n = min(RG_TILE_PIX, toDo);
        for (int j = 0; j < n; j++) {
            contributor--;
            float v[kAcc];
            bool active = false;
            if (inside && (int)contributor < last_contributor) {
                const float2 xy = s_xy[j];
                const float2 d = {xy.x - pixf.x, xy.y - pixf.y};
                const float4 con_o = s_co[j];
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                if (power <= 0.0f && !(power < s_thr[j])) {
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, con_o.w * G);
                    if (alpha >= 1.0f / 255.0f) {
                        active = true;
                        T = T / (1.f - alpha);
                        const float dchannel_dcolor = alpha * T;
                        const float4 c = s_rgbd[j];
                        const float col[3] = {c.x, c.y, c.z};
                        float dL_dopa = 0.0f;
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = col[ch];
                            dL_dopa += (col[ch] - accum_rec[ch]) * dL_dpixel[ch];
                            v[6 + ch] = dchannel_dcolor * dL_dpixel[ch];
                        }
                        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
                        last_depth = c.w;
                        dL_dopa += (c.w - accum_depth_rec) * dL_dpixel_depth;
                        v[9] = dchannel_dcolor * dL_dpixel_depth;
                        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
                        dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
                        dL_dopa *= T;
                        last_alpha = alpha;
                        dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                        const float dL_dG = con_o.w * dL_dopa;
                        const float gdx = G * d.x;
                        const float gdy = G * d.y;
                        const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
                        const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
                        v[0] = dL_dG * dG_ddelx * ddelx_dx;
                        v[1] = dL_dG * dG_ddely * ddely_dy;
                        v[2] = -0.5f * gdx * d.x * dL_dG;
                        v[3] = -0.5f * gdx * d.y * dL_dG;
                        v[4] = -0.5f * gdy * d.y * dL_dG;
                        v[5] = G * dL_dopa;
                    }
                }
            }
            if (__any_sync(0xffffffffu, active)) {
#pragma unroll
                for (int k = 0; k < kAcc; k++) {
                    float x = active ? v[k] : 0.f;
                    x = warp_sum(x);
                    if (lane == k) atomicAdd(&s_acc[k * RG_TILE_PIX + j], x);
                }
            }
        }
        __syncthreads();
        // one flush per staged instance
        if (progress < total) {
            const uint32_t id = s_id[tid];
            float* dst = a.grad_acc + (size_t)id * kAcc;
#pragma unroll
            for (int k = 0; k < kAcc; k++) {
                const float x = s_acc[k * RG_TILE_PIX + tid];
                if (x != 0.f) atomicAdd(dst + k, x);
            }
        }
    }
}

int launch_blend_backward(const BlendBwdArgs& a, cudaStream_t stream) {
    dim3 grid(a.grid_x, a.grid_y, 1);
    LaunchScope ls("blend_backward", stream);
    blend_backward_kernel<<<grid, RG_TILE_PIX, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
