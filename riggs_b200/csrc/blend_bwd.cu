// Backward alpha blend, back-to-front replay.  Replaces renderCUDA of dgr/cuda_rasterizer/backward.cu:415-601.
//
// The reference issues 10 global float atomicAdds per contributing (pixel, instance) pair from a 256-thread CTA per
// tile that walks the whole list in lock step.  Here (same decomposition as blend_fwd.cu)
//  * one warp = one 8x4 pixel block, independent of the rest of its tile: it replays only the entries in front of
//    ITS deepest last contributor (the reference skips the others one by one, backward.cu:517-519), and of those only
//    the instances whose alpha >= 1/255 footprint can reach the block (ListWalker of blend_common.cuh);
//  * two survivors are replayed per iteration: the Gaussian evaluations are independent, the T / accumulator
//    recurrence between them is a handful of FMAs, and the two warp reductions interleave;
//  * the ten per-pair terms of an instance are summed across the warp with a 12-shuffle transpose-reduce
//    (10 values x 32 lanes -> one value per lane) and leave the SM as ONE 10-lane reduction per
//    (warp block, instance) instead of up to 320 atomics.
// Per-pair arithmetic is the reference's expression order; sums differ only in association (tolerance 1e-4).
#include "blend_common.cuh"
#include "kernels.cuh"

namespace rg {

constexpr int kAcc = 10;  // mean2D.x, mean2D.y, conic.x, conic.y, conic.w, opacity, r, g, b, depth

// Optional per-(tile, warp) cycle counters for performance analysis (rg_debug_tile_profile); null in production.
__device__ uint32_t* g_bwd_tile_prof = nullptr;
void set_bwd_tile_profile(uint32_t* p) { cudaMemcpyToSymbol(g_bwd_tile_prof, &p, sizeof(p)); }

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

__global__ void __launch_bounds__(kBlendWarps * 32) blend_backward_kernel(BlendBwdArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = (int)a.tile_order[blockIdx.x >> 2], strip = blockIdx.x & 3;
    const int tile_x = tile % a.grid_x, tile_y = tile / a.grid_x;
    const int bx0 = warp * kBlkW, by0 = strip * kBlkH;
    const uint2 pix = {(unsigned)(tile_x * RG_TILE + bx0 + (lane & 7)), (unsigned)(tile_y * RG_TILE + by0 + (lane >> 3))};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;

    const int last_contributor = inside ? a.n_contrib[pix_id] : 0;
    int total = last_contributor;  // entries [range.x, range.x + total) can contribute to this block
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total = max(total, __shfl_xor_sync(0xffffffffu, total, o));
    if (total == 0) return;

    constexpr int kPerLane = 2, kGroup = 32 * kPerLane;
    __shared__ WarpStage<kPerLane> s_stage[kBlendWarps];
    WarpStage<kPerLane>& st = s_stage[warp];
    ListWalker<true, kPerLane> walk;
    walk.pl = a.point_list; walk.rec = a.rec; walk.first = a.ranges[tile_y * a.grid_x + tile_x].x; walk.n = total;
    walk.lane = lane;
    walk.x0 = (float)(tile_x * RG_TILE + bx0); walk.y0 = (float)(tile_y * RG_TILE + by0);
    walk.x1 = walk.x0 + (kBlkW - 1); walk.y1 = walk.y0 + (kBlkH - 1);
    const int n_groups = (total + kGroup - 1) / kGroup;
    walk.begin(st);

    const float T_final = inside ? (1 - a.alphas[pix_id]) : 0;
    float T = T_final;
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
    uint32_t* const prof = g_bwd_tile_prof;
    long long t_start = 0;
    uint32_t n_list = 0, n_red = 0;
    if (prof) { t_start = clock64(); walk.timing = true; }

    constexpr int U = 2;
    for (int g = 0; g < n_groups; g++) {
        const int cnt = walk.next(g, st);
        if (prof) n_list += cnt;
        for (int l = 0; l < cnt; l += U) {
            float v[U][kAcc];
            bool active[U];
            float G[U], alpha[U];
            float2 d[U];
            float4 con_o[U];
            int e[U];
            // independent part: Gaussian falloff and alpha of each of the U survivors
#pragma unroll
            for (int u = 0; u < U; u++) {
                e[u] = min(l + u, kGroup - 1);
                const float4 x = st.xyts[e[u]];
                // walk position k is range entry total-1-k; the pixel saw entries [0, last_contributor)
                const int contributor = total - 1 - __float_as_int(x.w);
                con_o[u] = st.co[e[u]];
                d[u] = {x.x - pixf.x, x.y - pixf.y};
                const float power =
                    -0.5f * (con_o[u].x * d[u].x * d[u].x + con_o[u].z * d[u].y * d[u].y) - con_o[u].y * d[u].x * d[u].y;
                active[u] = l + u < cnt && contributor < last_contributor && !(power > 0.0f) && !(power < x.z);
                G[u] = 0.f;
                alpha[u] = 0.f;
                if (active[u]) {
                    G[u] = expf(power);
                    alpha[u] = fminf(0.99f, con_o[u].w * G[u]);
                    active[u] = !(alpha[u] < 1.0f / 255.0f);
                }
            }
            // serial part: the back-to-front recurrence
#pragma unroll
            for (int u = 0; u < U; u++) {
#pragma unroll
                for (int k = 0; k < kAcc; k++) v[u][k] = 0.f;
                if (active[u]) {
                    // one reciprocal (accurate to the last bit or so) serves both quotients of the reference (T / (1 - alpha) and
                    // -T_final / (1 - alpha)): one rounding more than a division, far inside the 1e-4 gate
                    const float one_ma = 1.f - alpha[u];  // in [0.01, 1]
                    float inv_1ma = __fdividef(1.f, one_ma);        // MUFU.RCP, then one Newton step: < 1 ulp
                    inv_1ma = fmaf(inv_1ma, fmaf(-one_ma, inv_1ma, 1.f), inv_1ma);
                    T = T * inv_1ma;
                    const float dchannel_dcolor = alpha[u] * T;
                    const float4 c = st.rgbd[e[u]];
                    const float col[3] = {c.x, c.y, c.z};
                    float dL_dopa = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = col[ch];
                        dL_dopa += (col[ch] - accum_rec[ch]) * dL_dpixel[ch];
                        v[u][6 + ch] = dchannel_dcolor * dL_dpixel[ch];
                    }
                    accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
                    last_depth = c.w;
                    dL_dopa += (c.w - accum_depth_rec) * dL_dpixel_depth;
                    v[u][9] = dchannel_dcolor * dL_dpixel_depth;
                    accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
                    dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
                    dL_dopa *= T;
                    last_alpha = alpha[u];
                    dL_dopa += (-T_final * inv_1ma) * bg_dot_dpixel;
                    const float dL_dG = con_o[u].w * dL_dopa;
                    const float gdx = G[u] * d[u].x;
                    const float gdy = G[u] * d[u].y;
                    const float dG_ddelx = -gdx * con_o[u].x - gdy * con_o[u].y;
                    const float dG_ddely = -gdy * con_o[u].z - gdx * con_o[u].y;
                    v[u][0] = dL_dG * dG_ddelx * ddelx_dx;
                    v[u][1] = dL_dG * dG_ddely * ddely_dy;
                    v[u][2] = -0.5f * gdx * d[u].x * dL_dG;
                    v[u][3] = -0.5f * gdx * d[u].y * dL_dG;
                    v[u][4] = -0.5f * gdy * d[u].y * dL_dG;
                    v[u][5] = G[u] * dL_dopa;
                }
            }
            // one 10-lane reduction per (block, instance)
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (__any_sync(0xffffffffu, active[u])) {
                    const float s = transpose_reduce10(v[u], lane);
                    if (my_id >= 0 && s != 0.f) atomicAdd(a.grad_acc + (size_t)st.id[e[u]] * kAcc + my_id, s);
                }
            }
        }
    }
    if (prof && lane == 0) {
        uint32_t* o = prof + ((size_t)(tile_y * a.grid_x + tile_x) * 8 + strip * 2 + warp) * 8;
        o[0] = (uint32_t)(clock64() - t_start); o[1] = (uint32_t)walk.t_wait; o[2] = (uint32_t)walk.t_stage1; o[3] = (uint32_t)walk.t_stage2; o[4] = n_list; o[5] = n_groups; o[6] = n_red; o[7] = 0;
    }
}

int launch_blend_backward(const BlendBwdArgs& a, cudaStream_t stream) {
    const dim3 grid((unsigned)(a.grid_x * a.grid_y * 4), 1, 1);
    LaunchScope ls("blend_backward", stream);
    blend_backward_kernel<<<grid, kBlendWarps * 32, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
