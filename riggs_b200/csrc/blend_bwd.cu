// Backward alpha blend, back-to-front replay.  Replaces renderCUDA of dgr/cuda_rasterizer/backward.cu:415-601.
//
// The reference issues 10 global float atomicAdds per contributing (pixel, instance) pair from a 256-thread CTA per
// tile that walks the whole list in lock step.  Here (same decomposition as blend_fwd.cu)
//  * one warp = one 8x4 pixel block, independent of the rest of its tile: it replays only the entries in front of
//    ITS deepest last contributor (the reference skips the others one by one, backward.cu:517-519), and of those only
//    the instances the forward blend marked as able to reach the block (MaskWalker of blend_common.cuh reads the
//    block's survivor bitmap: no culling test is repeated, records are fetched for survivors only);
//  * phase A, one pixel per lane: the reference's recurrence for T; what lies behind the instance (accumulated
//    colour / depth / alpha) is carried as un-normalised suffix sums (see the loop).  Of the ten per-pair gradient terms only TWO scalars depend on
//    that recurrence:  Q = G dL/d(alpha)  and  W = alpha T ; the rest is geometry of the pair and constants of the
//    pixel.  Q and W go to a 32 x 32 (pixel x instance) tile in shared memory;
//  * phase B, one INSTANCE per lane, every 32 instances: the ten sums over the block's pixels are linear in Q and W,
//        sum Q {1, lx, ly, lx^2, lx ly, ly^2}   and   sum W {dL/dr, dL/dg, dL/db, dL/ddepth}
//    (lx, ly = pixel position inside the block), so each lane runs 32 x 10 FMAs down its column, rebuilds the moments
//    about its instance's centre and issues its ten reductions to global memory itself.
// Against a cross-lane reduction per instance (12 shuffles + ~40 selects/adds + one RED instruction, per instance and
// per block) the sums cost 13 instructions per pixel per 32 instances and nothing sits in the per-instance dependent
// chain any more (measured on the C0 frame: 114 M -> 80 M warp instructions, 205 -> 147 us with the bitmap walk).
// Sums differ from the reference's in association only (tolerance 1e-4).
#include "blend_common.cuh"
#include "kernels.cuh"

namespace rg {

constexpr int kAcc = 10;  // mean2D.x, mean2D.y, conic.x, conic.y, conic.w, opacity, r, g, b, depth
constexpr int kChunk = 32;          // instances per phase-B flush: one per lane
constexpr int kPitch = kChunk + 1;  // row pitch of the Q / W tiles: conflict-free for both phases

// Optional per-(tile, warp) cycle counters for performance analysis (rg_debug_tile_profile); null in production.
__device__ uint32_t* g_bwd_tile_prof = nullptr;
void set_bwd_tile_profile(uint32_t* p) { cudaMemcpyToSymbol(g_bwd_tile_prof, &p, sizeof(p)); }

// Shared memory per warp decides how many warps an SM holds here (the kernel is issue-bound and wants all it can get):
// the phase-B results reuse the W tile once every lane has read its column.
struct __align__(16) BwdChunk {
    float Q[32 * kPitch];   // [pixel][instance]
    float W[32 * kPitch];   // ... and, from the middle of a flush on, out[kChunk * kAcc]: phase-B results, instance-major
    float2 xy[kChunk];      // instance centre
    float4 co[kChunk];      // conic + opacity
    float4 cpix[32];        // per pixel: dL/d(r, g, b, depth)
    uint32_t id[kChunk];
};
static_assert(kChunk * kAcc <= 32 * kPitch, "phase-B results must fit into the W tile");

// Phase B: lane e owns instance e of the chunk.
__device__ __noinline__ void flush_chunk(BwdChunk& ch, int fill, int lane, float ox, float oy, float ddelx_dx,
                                            float ddely_dy, float* __restrict__ grad_acc) {
    __syncwarp();
    float r[kAcc];
#pragma unroll
    for (int k = 0; k < kAcc; k++) r[k] = 0.f;
    if (lane < fill) {
        float s0 = 0.f, sx = 0.f, sy = 0.f, sxx = 0.f, sxy = 0.f, syy = 0.f, wr = 0.f, wg = 0.f, wb = 0.f, wd = 0.f;
#pragma unroll
        for (int p = 0; p < 32; p++) {
            const float lx = (float)(p & 7), ly = (float)(p >> 3);
            const float q = ch.Q[p * kPitch + lane], w = ch.W[p * kPitch + lane];
            const float4 c = ch.cpix[p];
            s0 += q;
            sx = fmaf(q, lx, sx);
            sy = fmaf(q, ly, sy);
            sxx = fmaf(q, lx * lx, sxx);
            sxy = fmaf(q, lx * ly, sxy);
            syy = fmaf(q, ly * ly, syy);
            wr = fmaf(w, c.x, wr);
            wg = fmaf(w, c.y, wg);
            wb = fmaf(w, c.z, wb);
            wd = fmaf(w, c.w, wd);
        }
        // moments about the instance centre: d = (X - lx, Y - ly), X = x - block origin
        const float2 e = ch.xy[lane];
        const float4 co = ch.co[lane];
        const float X = e.x - ox, Y = e.y - oy;
        const float qdx = X * s0 - sx, qdy = Y * s0 - sy;
        const float qdxx = X * (X * s0 - 2.f * sx) + sxx;
        const float qdxy = X * (Y * s0 - sy) - Y * sx + sxy;
        const float qdyy = Y * (Y * s0 - 2.f * sy) + syy;
        // dL/dG = o dL/dalpha;  dG/ddelx = -G (A dx + B dy), dG/ddely = -G (C dy + B dx)
        const float o = co.w;
        r[0] = -o * ddelx_dx * (co.x * qdx + co.y * qdy);
        r[1] = -o * ddely_dy * (co.z * qdy + co.y * qdx);
        r[2] = -0.5f * o * qdxx;
        r[3] = -0.5f * o * qdxy;
        r[4] = -0.5f * o * qdyy;
        r[5] = s0;
        r[6] = wr;
        r[7] = wg;
        r[8] = wb;
        r[9] = wd;
    }
    __syncwarp();  // every lane has read its W column: the tile's memory now carries the results
    float* const out_all = ch.W;
    if (lane < fill) {
        float* out = out_all + lane * kAcc;
#pragma unroll
        for (int k = 0; k < kAcc; k++) out[k] = r[k];
    }
    __syncwarp();
    // The ten values of an instance are contiguous in grad_acc: lanes take consecutive (instance, value) pairs so that
    // one instruction's reductions fall into a few 32-byte sectors instead of 32.
#pragma unroll
    for (int j = 0; j < kAcc; j++) {
        const int idx = j * 32 + lane;
        const int ent = idx / kAcc, k = idx - ent * kAcc;
        if (ent < fill) {
            const float v = out_all[idx];
            if (v != 0.f) atomicAdd(grad_acc + (size_t)ch.id[ent] * kAcc + k, v);
        }
    }
    __syncwarp();
}

#ifndef RG_BWD_MINBLK
#define RG_BWD_MINBLK 8  // 8 CTAs (16 warps) per SM: at most 128 registers; measured 0.1351 -> 0.1316 ms at C0 against no bound (133 registers, 6 CTAs)
#endif
#ifndef RG_BWD_PERLANE
#define RG_BWD_PERLANE 2
#endif
__global__ void __launch_bounds__(kBlendWarps * 32, RG_BWD_MINBLK) blend_backward_kernel(BlendBwdArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kCtasPerTile = 8 / kBlendWarps;
    const int tile = (int)a.tile_order[blockIdx.x / kCtasPerTile];
    const int blk = (blockIdx.x % kCtasPerTile) * kBlendWarps + warp;  // 8x4 block of the tile: 2 across, 4 down
    const int tile_x = tile % a.grid_x, tile_y = tile / a.grid_x;
    if (a.split_n > 1 && tile_y % a.split_n != a.split_i) return;  // another GPU's tile row: its n_contrib was never written here
    const int bx0 = (blk & 1) * kBlkW, by0 = (blk >> 1) * kBlkH;
    const uint2 pix = {(unsigned)(tile_x * RG_TILE + bx0 + (lane & 7)), (unsigned)(tile_y * RG_TILE + by0 + (lane >> 3))};
    const uint32_t pix_id = a.W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (unsigned)a.W && pix.y < (unsigned)a.H;

    const int last_contributor = inside ? a.n_contrib[pix_id] : 0;
    int total = last_contributor;  // entries [range.x, range.x + total) can contribute to this block
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total = max(total, __shfl_xor_sync(0xffffffffu, total, o));
    if (total == 0) return;

    constexpr int kPerLane = RG_BWD_PERLANE, kGroup = 32 * kPerLane;
    __shared__ WarpStage<kPerLane, false> s_stage[kBlendWarps];
    __shared__ BwdChunk s_chunk[kBlendWarps];
    WarpStage<kPerLane, false>& st = s_stage[warp];
    BwdChunk& ch = s_chunk[warp];
    MaskWalker<kPerLane> walk;
    walk.pl = a.point_list; walk.rec = a.rec; walk.surv_bits = a.surv_mask + (size_t)blk * a.surv_stride; walk.first = a.ranges[tile].x; walk.n = total;
    walk.lane = lane;
    const float ox = (float)(tile_x * RG_TILE + bx0), oy = (float)(tile_y * RG_TILE + by0);  // block origin (pixel centres)
    const int n_groups = (total + kGroup - 1) / kGroup;
    walk.begin(st);

    const float T_final = inside ? (1 - a.alphas[pix_id]) : 0;
    float T = T_final;
    float S_rec[3] = {0, 0, 0}, dL_dpixel[3] = {0, 0, 0};  // S_*: un-normalised sums over the instances behind
    float S_depth = 0, dL_dpixel_depth = 0, S_alpha = 0, dL_dalpha = 0;
    if (inside) {
        const size_t HW = (size_t)a.H * a.W;
#pragma unroll
        for (int c = 0; c < 3; c++) dL_dpixel[c] = a.dL_dpix[c * HW + pix_id];
        if (a.dL_dpix_depth) dL_dpixel_depth = a.dL_dpix_depth[pix_id];  // null: that output has no cotangent
        if (a.dL_dalphas) dL_dalpha = a.dL_dalphas[pix_id];
    }
    ch.cpix[lane] = {dL_dpixel[0], dL_dpixel[1], dL_dpixel[2], dL_dpixel_depth};
    const float ddelx_dx = 0.5 * a.W;
    const float ddely_dy = 0.5 * a.H;
    const float bg_dot_dpixel = a.bg[0] * dL_dpixel[0] + a.bg[1] * dL_dpixel[1] + a.bg[2] * dL_dpixel[2];
    uint32_t* const prof = g_bwd_tile_prof;
    long long t_start = 0;
    uint32_t n_list = 0, n_flush = 0;
    if (prof) { t_start = clock64(); walk.timing = true; }

    // lanes 0-1 copy the centre, 2-5 conic+opacity, 6 the id (all 32-bit words; ids travel as bit patterns)
    const uint32_t* meta_src = lane < 2 ? reinterpret_cast<const uint32_t*>(st.xyts) + lane
                               : lane < 6 ? reinterpret_cast<const uint32_t*>(st.co) + (lane - 2) : st.id;
    uint32_t* meta_dst = lane < 2 ? reinterpret_cast<uint32_t*>(ch.xy) + lane
                         : lane < 6 ? reinterpret_cast<uint32_t*>(ch.co) + (lane - 2) : ch.id;
    const int meta_sstride = lane < 6 ? 4 : 1, meta_dstride = lane < 2 ? 2 : (lane < 6 ? 4 : 1);
    constexpr int U = 2;
    int fill = 0;  // instances waiting in the chunk
    for (int g = 0; g < n_groups; g++) {
        const int cnt = walk.next(g, st);
        if (prof) n_list += cnt;
        // The loop below evaluates U slots per iteration; the slots behind the last survivor hold whatever an earlier
        // group left there (or nothing yet), and the suffix sums multiply colours by a zero weight instead of selecting:
        // make those slots finite.
        if (cnt % U) {
            if (lane < U - cnt % U) {
                const float4 z = {0.f, 0.f, 0.f, 0.f};
                st.xyts[cnt + lane] = z; st.co[cnt + lane] = z; st.rgbd[cnt + lane] = z;
            }
            __syncwarp();
        }
        // Straight-line code without per-lane branches, U instances per iteration: the warps of the heavy blocks are
        // the kernel's critical path and what bounds them is the latency of their own dependent chain.  A pair the
        // reference skips takes alpha = 0 -- T / (1 - 0) is exact -- and leaves the recurrence state untouched.
        const int n_it = (cnt + U - 1) / U;
        for (int it = 0; it < n_it; it++) {
            const int l = it * U;
            bool active[U];
            float G[U], alpha[U], q[U], w[U];
            float4 xy[U], con_o[U], col[U];
            int e[U];
            // independent part: Gaussian falloff and alpha of each of the U survivors
#pragma unroll
            for (int u = 0; u < U; u++) {
                e[u] = min(l + u, kGroup - 1);
                xy[u] = st.xyts[e[u]];
                con_o[u] = st.co[e[u]];
                col[u] = st.rgbd[e[u]];
                // walk position k is range entry total-1-k; the pixel saw entries [0, last_contributor)
                const int contributor = total - 1 - __float_as_int(xy[u].w);
                const float2 d = {xy[u].x - pixf.x, xy[u].y - pixf.y};
                const float power = -0.5f * (con_o[u].x * d.x * d.x + con_o[u].z * d.y * d.y) - con_o[u].y * d.x * d.y;
                G[u] = expf(power);
                alpha[u] = fminf(0.99f, con_o[u].w * G[u]);
                active[u] = l + u < cnt && contributor < last_contributor && !(power > 0.0f) && !(power < xy[u].z) &&
                            !(alpha[u] < 1.0f / 255.0f);
            }
            // serial part: the back-to-front recurrence of the reference, reduced to the two scalars phase B needs
#pragma unroll
            for (int u = 0; u < U; u++) {
                const bool on = active[u];
                const float a_eff = on ? alpha[u] : 0.f;
                // one reciprocal (accurate to the last bit or so) serves both quotients of the reference (T / (1 - alpha)
                // and -T_final / (1 - alpha)): one rounding more than a division, far inside the 1e-4 gate
                const float one_ma = 1.f - a_eff;  // in [0.01, 1]
                float inv_1ma = __fdividef(1.f, one_ma);        // MUFU.RCP, then one Newton step: < 1 ulp
                inv_1ma = fmaf(inv_1ma, fmaf(-one_ma, inv_1ma, 1.f), inv_1ma);
                T = T * inv_1ma;
                w[u] = a_eff * T;  // dchannel_dcolor
                // The reference carries "what lies behind the instance", normalised by the transmittance behind it, through
                // acc = last_alpha last_c + (1 - last_alpha) acc (backward.cu:543-567).  Here the UN-normalised suffix sums
                // S = sum over the instances behind of value * alpha * T are carried instead: T (c - acc) = c T - S / (1 - alpha),
                // and S += c alpha T is a no-op for a skipped pair (alpha = 0) without any select.  Same value, one rounding
                // order apart (tolerance 1e-4).
                float dL_dopa = 0.0f;
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const float ck = k == 0 ? col[u].x : (k == 1 ? col[u].y : col[u].z);
                    dL_dopa = fmaf(fmaf(-S_rec[k], inv_1ma, ck * T), dL_dpixel[k], dL_dopa);
                    S_rec[k] = fmaf(ck, w[u], S_rec[k]);
                }
                dL_dopa = fmaf(fmaf(-S_depth, inv_1ma, col[u].w * T), dL_dpixel_depth, dL_dopa);
                S_depth = fmaf(col[u].w, w[u], S_depth);
                dL_dopa = fmaf(fmaf(-S_alpha, inv_1ma, T), dL_dalpha, dL_dopa);
                S_alpha += w[u];
                dL_dopa += (-T_final * inv_1ma) * bg_dot_dpixel;
                q[u] = on ? G[u] * dL_dopa : 0.f;
            }
            // park the pair scalars in the chunk; instances no pixel of the block sees are dropped
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (__any_sync(0xffffffffu, active[u])) {
                    ch.Q[lane * kPitch + fill] = q[u];
                    ch.W[lane * kPitch + fill] = w[u];
                    // instance record -> chunk: nine words copied by nine lanes (no divergent single-lane stores)
                    if (lane < 7) meta_dst[fill * meta_dstride] = meta_src[e[u] * meta_sstride];
                    if (++fill == kChunk) {
                        flush_chunk(ch, fill, lane, ox, oy, ddelx_dx, ddely_dy, a.grad_acc);
                        fill = 0;
                        if (prof) n_flush++;
                    }
                }
            }
        }
    }
    if (fill > 0) {
        flush_chunk(ch, fill, lane, ox, oy, ddelx_dx, ddely_dy, a.grad_acc);
        if (prof) n_flush++;
    }
    if (prof && lane == 0) {
        uint32_t* o = prof + ((size_t)tile * 8 + blk) * 8;
        o[0] = (uint32_t)(clock64() - t_start); o[1] = (uint32_t)walk.t_wait; o[2] = (uint32_t)walk.t_stage1;
        o[3] = (uint32_t)walk.t_stage2; o[4] = n_list; o[5] = n_groups; o[6] = n_flush; o[7] = 0;
    }
}

int launch_blend_backward(const BlendBwdArgs& a, cudaStream_t stream) {
    const dim3 grid((unsigned)(a.grid_x * a.grid_y * (8 / kBlendWarps)), 1, 1);
    LaunchScope ls("blend_backward", stream);
    blend_backward_kernel<<<grid, kBlendWarps * 32, 0, stream>>>(a);
    return 0;
}

}  // namespace rg
