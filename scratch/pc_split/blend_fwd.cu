// Forward alpha blend.  Replaces renderCUDA (dgr/cuda_rasterizer/forward.cu:261-381).
//
// The reference runs one 256-thread CTA per 16x16 tile: all threads walk the tile's whole depth-sorted list in
// lock step (two __syncthreads per 256 instances) and every pixel evaluates every instance.  On B200 that shape is
// latency-bound: the few heavy tiles of a frame decide the kernel's duration while most SMs idle.  Here the unit of
// work is one 8x4 pixel block, independent of the other seven blocks of its tile, served by a CTA of two warps
// (blend_common.cuh):
//   * the filter warp hands over, 64 list entries at a time, the instances whose alpha >= 1/255 footprint can reach
//     the block (exact-conservative test of cull.cuh);
//   * the blend warp's 32 pixels blend only those, in list order, with the reference's exact expressions (same power /
//     alpha / T tests; a log-threshold pre-test avoids expf for pairs that are certainly below 1/255).  Four
//     survivors are evaluated per iteration: their alphas are independent, only the short T / colour update is
//     serial, so the expf latencies overlap instead of adding up;
//   * both stop as soon as the 32 pixels are saturated (T < 1e-4) or the list ends.
// Every block re-reads its tile's list (8x the L2 reads of a cooperative staging; one frame's records fit in L2 several
// times over) -- L2 bandwidth is plentiful here, latency and balance are not.  Tiles are launched longest list first
// (tile_order from the binning pass).
#include "blend_common.cuh"
#include "kernels.cuh"

namespace rg {

// Optional per-(tile, block) cycle counters for performance analysis (rg_debug_tile_profile); null in production.
__device__ uint32_t* g_fwd_tile_prof = nullptr;
void set_fwd_tile_profile(uint32_t* p) { cudaMemcpyToSymbol(g_fwd_tile_prof, &p, sizeof(p)); }

__global__ void __launch_bounds__(kBlendThreads) blend_forward_kernel(BlendFwdArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = (int)a.tile_order[blockIdx.x >> 3], blk = blockIdx.x & 7;
    const int tile_x = tile % a.grid_x, tile_y = tile / a.grid_x;
    const int bx0 = (blk & 1) * kBlkW, by0 = (blk >> 1) * kBlkH;  // 2 blocks across, 4 down
    const uint2 range = a.ranges[tile];
    const int n_total = (int)(range.y - range.x);

    __shared__ BlockPipe bp;
    pipe_init(bp);
    __syncthreads();

    if (warp == 1) {  // ---- filter warp
        if (n_total == 0) return;
        ListWalker<false> walk;
        walk.pl = a.point_list; walk.rec = a.rec; walk.first = range.x; walk.n = n_total; walk.lane = lane;
        walk.x0 = (float)(tile_x * RG_TILE + bx0); walk.y0 = (float)(tile_y * RG_TILE + by0);
        walk.x1 = walk.x0 + (kBlkW - 1); walk.y1 = walk.y0 + (kBlkH - 1);
        run_filter_warp(walk, bp);
        return;
    }

    // ---- blend warp: one pixel per lane
    const uint2 pix = {(unsigned)(tile_x * RG_TILE + bx0 + (lane & 7)), (unsigned)(tile_y * RG_TILE + by0 + (lane >> 3))};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    bool done = !inside;
    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;
    uint32_t* const prof = g_fwd_tile_prof;
    long long t_start = 0, t_wait = 0;
    uint32_t n_list = 0, n_slots = 0;
    if (prof) t_start = clock64();

    RingReader ring;
    while (n_total > 0) {
        long long tw = 0;
        if (prof) tw = clock64();
        const int cnt = ring.acquire(bp);
        if (prof) { t_wait += clock64() - tw; n_slots++; n_list += max(cnt, 0); }
        if (cnt < 0) break;
        const SurvStage& st = bp.stage[ring.slot];
        for (int l = 0; l < cnt; l += 4) {
            if (done) break;
            float alpha[4], power[4], opac[4];
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
                power[u] = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                // the reference skips power > 0 and alpha < 1/255; power < thr is certainly alpha < 1/255
                act[u] = l + u < cnt && !(power[u] > 0.0f) && !(power[u] < x.z);
                opac[u] = con_o.w;
                pos[u] = __float_as_int(x.w);
            }
            if (act[0] | act[1] | act[2] | act[3]) {
                // the four falloffs are independent: evaluated together so that their latencies overlap
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    alpha[u] = fminf(0.99f, opac[u] * expf(power[u]));
                    act[u] = act[u] && !(alpha[u] < 1.0f / 255.0f);
                }
                // the only serial part: transmittance and the accumulators
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const float test_T = T * (1 - alpha[u]);
                    const bool live = act[u] && !done;
                    const bool stop = live && test_T < 0.0001f;
                    const bool add = live && !stop;
                    done = done || stop;
                    C0 = add ? C0 + c[u].x * alpha[u] * T : C0;
                    C1 = add ? C1 + c[u].y * alpha[u] * T : C1;
                    C2 = add ? C2 + c[u].z * alpha[u] * T : C2;
                    weight = add ? weight + alpha[u] * T : weight;
                    D = add ? D + c[u].w * alpha[u] * T : D;
                    // 1-based position in the tile's range, as the reference counts
                    last_contributor = add ? (uint32_t)pos[u] + 1 : last_contributor;
                    T = add ? test_T : T;
                }
            }
        }
        ring.release(bp, lane);
        if (__all_sync(0xffffffffu, done)) {
            if (lane == 0) *reinterpret_cast<volatile int*>(&bp.stop) = 1;
            break;
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
        uint32_t* o = prof + ((size_t)tile * 8 + blk) * 8;
        o[0] = (uint32_t)(clock64() - t_start); o[1] = (uint32_t)t_wait; o[2] = 0; o[3] = 0;
        o[4] = n_list; o[5] = n_slots; o[6] = n_total; o[7] = 0;
    }
}

int launch_blend_forward(const BlendFwdArgs& a, cudaStream_t stream) {
    const dim3 grid((unsigned)(a.grid_x * a.grid_y * 8), 1, 1);
    LaunchScope ls("blend_forward", stream);
    blend_forward_kernel<<<grid, kBlendThreads, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
