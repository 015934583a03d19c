// Forward alpha blend: one CTA per 16x16 tile, front-to-back over the tile's depth-sorted instance range.
// Replaces renderCUDA (dgr/cuda_rasterizer/forward.cu:261-381).
//
// Differences in HOW (results are the same, see cull.cuh for the argument):
//  * a warp owns an 8x4 pixel block of the tile.  Every batch of 256 staged instances is first filtered per
//    warp (ellipse-vs-block test, 8 instances per lane, ballot-compacted into a per-warp list in shared
//    memory); the per-pixel loop then runs over that list only.  On the C0 workload this removes ~3/4 of
//    the (pixel, instance) evaluations and lets warps whose pixels are saturated or outside the silhouette
//    skim through long lists at the cost of the filter alone;
//  * colour and depth of each staged instance are gathered once per tile into shared memory instead of
//    being fetched from global memory per contributing pair;
//  * survivors take the reference's exact expression (same power / alpha / T tests), after a conservative
//    log-threshold pre-test that avoids expf for pairs that are certainly below 1/255.
#include "common.cuh"
#include "cull.cuh"
#include "kernels.cuh"

namespace rg {

constexpr float kSkipMargin = 1e-4f;

__global__ void __launch_bounds__(RG_TILE_PIX) blend_forward_kernel(BlendFwdArgs a) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // warp -> 8x4 block: 2 blocks across, 4 down
    const int bx0 = (warp & 1) * kBlkW, by0 = (warp >> 1) * kBlkH;
    const uint2 pix = {blockIdx.x * RG_TILE + bx0 + (lane & 7), blockIdx.y * RG_TILE + by0 + (lane >> 3)};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;
    bool done = !inside;
    const float tile_x0 = (float)(blockIdx.x * RG_TILE), tile_y0 = (float)(blockIdx.y * RG_TILE);
    const float blk_x0 = tile_x0 + bx0, blk_y0 = tile_y0 + by0;
    const float blk_x1 = blk_x0 + (kBlkW - 1), blk_y1 = blk_y0 + (kBlkH - 1);

    const uint2 range = a.ranges[blockIdx.y * a.grid_x + blockIdx.x];
    const int rounds = ((range.y - range.x + RG_TILE_PIX - 1) / RG_TILE_PIX);
    int toDo = range.y - range.x;

    __shared__ float2 s_xy[RG_TILE_PIX];
    __shared__ float4 s_co[RG_TILE_PIX];
    __shared__ float4 s_rgbd[RG_TILE_PIX];
    __shared__ float s_thr[RG_TILE_PIX];
    __shared__ CullRec s_cull[RG_TILE_PIX];
    __shared__ uint8_t s_list[8][RG_TILE_PIX];

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, weight = 0.f, D = 0.f;

    // Software pipeline: the gathers of batch i+1 (id -> xy, conic, colour, depth: two dependent trips to L2/HBM) are
    // issued right after batch i has been published to shared memory and complete while the warps filter and blend
    // batch i.  On tiles whose pixels never saturate the per-batch latency is the kernel's critical path.
    float2 n_xy = {0.f, 0.f};
    float4 n_co = {0.f, 0.f, 0.f, 0.f}, n_rgbd = {0.f, 0.f, 0.f, 0.f};
    auto fetch = [&](int batch) {
        const uint32_t pos = range.x + (uint32_t)batch * RG_TILE_PIX + tid;
        if (pos < range.y) {
            const uint32_t id = a.point_list[pos];
            n_xy = a.means2D[id];
            n_co = a.conic_opacity[id];
            n_rgbd = {a.colors[3 * id], a.colors[3 * id + 1], a.colors[3 * id + 2], a.depths[id]};
        }
    };
    if (rounds > 0) fetch(0);
    for (int i = 0; i < rounds; i++, toDo -= RG_TILE_PIX) {
        int num_done = __syncthreads_count(done);
        if (num_done == RG_TILE_PIX) break;
        const int progress = i * RG_TILE_PIX + tid;
        if (range.x + progress < range.y) {
            s_xy[tid] = n_xy;
            s_co[tid] = n_co;
            s_rgbd[tid] = n_rgbd;
            s_thr[tid] = -logf(255.0f * n_co.w) - kSkipMargin;  // alpha < 1/255  <=>  power < -log(255 o)
            s_cull[tid] = make_cull(n_xy, n_co, tile_x0, tile_y0);
        }
        __syncthreads();
        if (i + 1 < rounds) fetch(i + 1);
        const int n = min(RG_TILE_PIX, toDo);
        if (__all_sync(0xffffffffu, done)) continue;  // this warp's block is finished; keep helping with loads
        // ---- per-warp filter: which of the n staged instances can reach this 8x4 block?
        int cnt = 0;
#pragma unroll
        for (int k = 0; k < RG_TILE_PIX / 32; k++) {
            const int e = k * 32 + lane;
            bool keep = false;
            if (e < n) keep = cull_keep(s_cull[e], s_xy[e], s_co[e], blk_x0, blk_y0, blk_x1, blk_y1);
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) s_list[warp][cnt + __popc(m & ((1u << lane) - 1))] = (uint8_t)e;
            cnt += __popc(m);
        }
        __syncwarp();
        // ---- per-pixel blend over the surviving instances, in list (= depth) order
        const uint32_t base = (uint32_t)i * RG_TILE_PIX;
        for (int l = 0; !done && l < cnt; l++) {
            const int j = s_list[warp][l];
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
            last_contributor = base + j + 1;  // 1-based position in the tile's range, as the reference counts
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
