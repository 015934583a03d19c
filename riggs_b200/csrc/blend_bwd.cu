// Backward alpha blend: one CTA per tile, back-to-front replay.  Replaces renderCUDA of
// dgr/cuda_rasterizer/backward.cu:415-601.
//
// The reference issues 10 global float atomicAdds per contributing (pixel, instance) pair.  Here
//  * a warp owns an 8x4 pixel block and, per batch of 256 staged instances, first builds the list of
//    instances whose alpha >= 1/255 ellipse can reach its block (cull.cuh) and that lie in front of the
//    block's deepest last contributor; only those are replayed;
//  * the ten per-pair terms of an instance are summed across the warp with a 12-shuffle transpose-reduce
//    (10 values x 32 lanes -> one value per lane) and leave the SM as ONE 10-lane reduction per
//    (warp block, instance): ~2 x 10 atomics per (tile, instance) instead of up to 2560;
//  * the replay starts at the tile's largest n_contrib instead of the end of the range: entries behind
//    every pixel's last contributor are skipped by the reference one by one (backward.cu:517-519).
// Per-pair arithmetic is the reference's expression order; sums differ only in association (tolerance 1e-4).
#include "common.cuh"
#include "cull.cuh"
#include "kernels.cuh"

namespace rg {

constexpr float kSkipMarginB = 1e-4f;
constexpr int kAcc = 10;  // mean2D.x, mean2D.y, conic.x, conic.y, conic.w, opacity, r, g, b, depth

// Sum each of v[0..9] over the 32 lanes.  Afterwards the lane with id_of_lane(lane) = k >= 0 holds the
// total of v[k] in the return value.  12 shuffles (5+3+2+1+1) instead of 50.
__device__ __forceinline__ float transpose_reduce10(const float* v, int lane) {
    const unsigned full = 0xffffffffu;
    const bool u4 = lane & 16, u3 = lane & 8, u2 = lane & 4, u1 = lane & 2;
    float w[5];
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const float keep = u4 ? v[5 + i] : v[i];
        const float send = u4 ? v[i] : v[5 + i];
        w[i] = keep + __shfl_xor_sync(full, send, 16);
    }
    float x[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const float hi = i < 2 ? w[3 + i] : 0.f;
        const float keep = u3 ? hi : w[i];
        const float send = u3 ? w[i] : hi;
        x[i] = keep + __shfl_xor_sync(full, send, 8);
    }
    float y[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float hi = i < 1 ? x[2] : 0.f;
        const float keep = u2 ? hi : x[i];
        const float send = u2 ? x[i] : hi;
        y[i] = keep + __shfl_xor_sync(full, send, 4);
    }
    const float keep = u1 ? y[1] : y[0];
    const float send = u1 ? y[0] : y[1];
    float s = keep + __shfl_xor_sync(full, send, 2);
    s += __shfl_xor_sync(full, s, 1);
    return s;
}

__device__ __forceinline__ int id_of_lane(int lane) {
    // bits 3..1 select within a group of five: 000->0 001->1 010->2 100->3 101->4, others hold zeros
    const int t = (lane >> 1) & 7;
    const int sub = t == 0 ? 0 : t == 1 ? 1 : t == 2 ? 2 : t == 4 ? 3 : t == 5 ? 4 : -1;
    if (sub < 0 || (lane & 1)) return -1;
    return ((lane & 16) ? 5 : 0) + sub;
}

__global__ void __launch_bounds__(RG_TILE_PIX) blend_backward_kernel(BlendBwdArgs a) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bx0 = (warp & 1) * kBlkW, by0 = (warp >> 1) * kBlkH;
    const uint2 pix = {blockIdx.x * RG_TILE + bx0 + (lane & 7), blockIdx.y * RG_TILE + by0 + (lane >> 3)};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    const uint2 range = a.ranges[blockIdx.y * a.grid_x + blockIdx.x];
    const float tile_x0 = (float)(blockIdx.x * RG_TILE), tile_y0 = (float)(blockIdx.y * RG_TILE);
    const float blk_x0 = tile_x0 + bx0, blk_y0 = tile_y0 + by0;
    const float blk_x1 = blk_x0 + (kBlkW - 1), blk_y1 = blk_y0 + (kBlkH - 1);

    __shared__ uint32_t s_id[RG_TILE_PIX];
    __shared__ float2 s_xy[RG_TILE_PIX];
    __shared__ float4 s_co[RG_TILE_PIX];
    __shared__ float4 s_rgbd[RG_TILE_PIX];
    __shared__ float s_thr[RG_TILE_PIX];
    __shared__ CullRec s_cull[RG_TILE_PIX];
    __shared__ uint8_t s_list[8][RG_TILE_PIX];
    __shared__ int s_max;

    const float T_final = inside ? (1 - a.alphas[pix_id]) : 0;
    float T = T_final;
    const int last_contributor = inside ? a.n_contrib[pix_id] : 0;

    if (tid == 0) s_max = 0;
    __syncthreads();
    int wmax = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) atomicMax(&s_max, wmax);
    __syncthreads();
    const int total = s_max;  // instances [range.x, range.x + total) can contribute to this tile
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
    const int my_id = id_of_lane(lane);

    const int rounds = (total + RG_TILE_PIX - 1) / RG_TILE_PIX;
    for (int i = 0; i < rounds; i++) {
        __syncthreads();
        const int progress = i * RG_TILE_PIX + tid;  // 0 = deepest candidate
        if (progress < total) {
            const uint32_t id = a.point_list[range.x + total - progress - 1];
            const float2 xy = a.means2D[id];
            const float4 co = a.conic_opacity[id];
            s_id[tid] = id;
            s_xy[tid] = xy;
            s_co[tid] = co;
            s_rgbd[tid] = {a.colors[3 * id], a.colors[3 * id + 1], a.colors[3 * id + 2], a.depths[id]};
            s_thr[tid] = -logf(255.0f * co.w) - kSkipMarginB;
            s_cull[tid] = make_cull(xy, co, tile_x0, tile_y0);
        }
        __syncthreads();
        const int n = min(RG_TILE_PIX, total - i * RG_TILE_PIX);
        // position (0-based, in the tile's range) of staged entry e is  pos(e) = total - 1 - (i*256 + e)
        const int pos0 = total - 1 - i * RG_TILE_PIX;
        if (pos0 - (n - 1) >= wmax) continue;  // the whole batch lies behind this block's last contributors
        int cnt = 0;
#pragma unroll
        for (int k = 0; k < RG_TILE_PIX / 32; k++) {
            const int e = k * 32 + lane;
            bool keep = false;
            if (e < n && pos0 - e < wmax) keep = cull_keep(s_cull[e], s_xy[e], s_co[e], blk_x0, blk_y0, blk_x1, blk_y1);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) s_list[warp][cnt + __popc(m & ((1u << lane) - 1))] = (uint8_t)e;
            cnt += __popc(m);
        }
        __syncwarp();
        for (int l = 0; l < cnt; l++) {
            const int j = s_list[warp][l];
            const int contributor = pos0 - j;
            float v[kAcc];
#pragma unroll
            for (int k = 0; k < kAcc; k++) v[k] = 0.f;
            bool active = false;
            if (contributor < last_contributor) {
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
                const float s = transpose_reduce10(v, lane);
                if (my_id >= 0 && s != 0.f) atomicAdd(a.grad_acc + (size_t)s_id[j] * kAcc + my_id, s);
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
