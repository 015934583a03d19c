// Forward alpha blend.  Replaces renderCUDA (dgr/cuda_rasterizer/forward.cu:261-381).
//
// The reference runs one 256-thread CTA per 16x16 tile: all threads walk the tile's whole depth-sorted list in
// lock step (two __syncthreads per 256 instances) and every pixel evaluates every instance.  On B200 that shape is
// latency-bound: the few heavy tiles of a frame decide the kernel's duration while most SMs idle.  Here the unit of
// work is ONE WARP = one 8x4 pixel block, independent of the other seven blocks of its tile -- no block barrier, a
// finished warp retires and frees its slot:
//   * blend_common.cuh's ListWalker hands the warp, 128 list entries at a time, the instances whose alpha >= 1/255
//     footprint can reach the block (exact-conservative test of cull.cuh), gathered with a two-deep prefetch;
//   * the 32 pixels blend only those, in list order, with the reference's exact expressions (same power / alpha / T
//     tests; a log-threshold pre-test avoids expf for pairs that are certainly below 1/255).  Four survivors are
//     evaluated per iteration: their alphas are independent, only the short T / colour update is serial, so the
//     expf latencies overlap instead of adding up;
//   * the warp stops as soon as its 32 pixels are saturated (T < 1e-4) or the list ends.
// Every warp re-reads the tile's list (8x the L2 reads of a cooperative staging; one frame's records fit in L2 several
// times over) -- L2 bandwidth is plentiful here, latency and balance are not.
#include "blend_common.cuh"
#include "kernels.cuh"

namespace rg {

// Optional per-(tile, warp) cycle counters for performance analysis (rg_debug_tile_profile); null in production.
__device__ uint32_t* g_fwd_tile_prof = nullptr;
void set_fwd_tile_profile(uint32_t* p) { cudaMemcpyToSymbol(g_fwd_tile_prof, &p, sizeof(p)); }

#ifndef RG_FWD_MINBLK
#define RG_FWD_MINBLK 1
#endif
__global__ void __launch_bounds__(kBlendWarps * 32, RG_FWD_MINBLK) blend_forward_kernel(BlendFwdArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kCtasPerTile = 8 / kBlendWarps;
    const int tile = (int)a.tile_order[blockIdx.x / kCtasPerTile];
    const int blk = (blockIdx.x % kCtasPerTile) * kBlendWarps + warp;  // 8x4 block of the tile: 2 across, 4 down
    const int tile_x = tile % a.grid_x, tile_y = tile / a.grid_x;
    if (a.split_n > 1 && tile_y % a.split_n != a.split_i) return;  // another GPU's tile row: nothing is written
    const int bx0 = (blk & 1) * kBlkW, by0 = (blk >> 1) * kBlkH;
    const uint2 pix = {(unsigned)(tile_x * RG_TILE + bx0 + (lane & 7)), (unsigned)(tile_y * RG_TILE + by0 + (lane >> 3))};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    bool done = !inside;
    const uint2 range = a.ranges[tile_y * a.grid_x + tile_x];

    constexpr int kPerLane = 4, kGroup = 32 * kPerLane;
    __shared__ WarpStage<kPerLane> s_stage[kBlendWarps];
    WarpStage<kPerLane>& st = s_stage[warp];
    ListWalker<false, kPerLane> walk;
    walk.pl = a.point_list; walk.rec = a.rec; walk.first = range.x; walk.n = (int)(range.y - range.x); walk.lane = lane;
    walk.x0 = (float)(tile_x * RG_TILE + bx0); walk.y0 = (float)(tile_y * RG_TILE + by0);
    walk.x1 = walk.x0 + (kBlkW - 1); walk.y1 = walk.y0 + (kBlkH - 1);
    walk.surv_bits = a.surv_mask + (size_t)blk * a.surv_stride;
    const int n_groups = (walk.n + kGroup - 1) / kGroup;

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;
    uint32_t* const prof = g_fwd_tile_prof;
    long long t_start = 0;
    uint32_t n_list = 0, n_seen = 0;
    if (prof) { t_start = clock64(); walk.timing = true; }

    if (n_groups > 0) walk.begin(st);
    for (int g = 0; g < n_groups; g++) {
        if (__all_sync(0xffffffffu, done)) break;
        const int cnt = walk.next(g, st);
        if (prof) { n_seen++; n_list += cnt; }
        // Straight-line code, no per-lane branches: the warp that owns a heavy block is alone on the kernel's critical
        // path, and what bounds it is the latency of its own dependent chain, not issue slots.  Entries that the
        // reference would skip take alpha = 0 (T * (1 - 0) and C + 0 are exact), so only T -> test -> select is serial.
        // four slots are evaluated per iteration; make the ones behind the last survivor finite (zero weight, no select)
        if (cnt & 3) {
            if (lane < 4 - (cnt & 3)) {
                const float4 z = {0.f, 0.f, 0.f, 0.f};
                st.xyts[cnt + lane] = z; st.co[cnt + lane] = z; st.rgbd[cnt + lane] = z;
            }
            __syncwarp();
        }
        const int n_it = (cnt + 3) >> 2;
        for (int it = 0; it < n_it; it++) {
            const int l = it * 4;
            float alpha[4];
            float4 c[4];
            int pos[4];
            bool act[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int e = min(l + u, kGroup - 1);
                const float4 x = st.xyts[e];
                const float4 con_o = st.co[e];
                c[u] = st.rgbd[e];
                const float2 d = {x.x - pixf.x, x.y - pixf.y};
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                alpha[u] = fminf(0.99f, con_o.w * expf(power));
                // the reference skips power > 0 and alpha < 1/255; power < thr is certainly alpha < 1/255
                act[u] = l + u < cnt && !(power > 0.0f) && !(power < x.z) && !(alpha[u] < 1.0f / 255.0f);
                pos[u] = __float_as_int(x.w);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float test_T = T * (1 - alpha[u]);
                const bool live = act[u] && !done;
                const bool stop = live && test_T < 0.0001f;
                const bool add = live && !stop;
                done = done || stop;
                // a skipped entry has weight 0: its (finite) colour needs no select of its own
                const float wa = add ? alpha[u] : 0.f;
                C0 = C0 + c[u].x * wa * T;
                C1 = C1 + c[u].y * wa * T;
                C2 = C2 + c[u].z * wa * T;
                weight = weight + wa * T;
                D = D + c[u].w * wa * T;
                // 1-based position in the tile's range, as the reference counts
                last_contributor = add ? (uint32_t)pos[u] + 1 : last_contributor;
                T = add ? test_T : T;
            }
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
    if (prof && lane == 0) {
        uint32_t* o = prof + ((size_t)(tile_y * a.grid_x + tile_x) * 8 + blk) * 8;
        o[0] = (uint32_t)(clock64() - t_start); o[1] = (uint32_t)walk.t_wait; o[2] = (uint32_t)walk.t_stage1; o[3] = (uint32_t)walk.t_stage2; o[4] = n_list; o[5] = n_seen; o[6] = walk.n; o[7] = 0;
    }
}

int launch_blend_forward(const BlendFwdArgs& a, cudaStream_t stream) {
    const dim3 grid((unsigned)(a.grid_x * a.grid_y * (8 / kBlendWarps)), 1, 1);
    LaunchScope ls("blend_forward", stream);
    blend_forward_kernel<<<grid, kBlendWarps * 32, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
