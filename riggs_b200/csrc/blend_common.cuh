// Front end shared by the forward and backward blend kernels: one warp walks one tile's instance list on behalf of
// its 8x4 pixel block and reduces it to the few instances whose footprint can reach the block.
//
// The walk is software-pipelined so that the dependent chain  list entry -> record  (two trips to L2) never sits on
// the warp's critical path -- the heavy tiles of a frame are walked by a handful of warps whose latency IS the
// kernel's duration:
//   iteration g   ids of group g+2 requested (coalesced)            ) in flight while group g is filtered and the
//                 rec[0] of group g+1 requested (ids came in at g-1) ) caller blends group g's survivors
//   stage 1       bounding-box test on rec[0] of group g (registers), ballot-compacted into a candidate buffer
//   stage 2       candidates are spread over the lanes; each lane fetches the rest of its candidate's record, runs
//                 the exact ellipse-vs-rectangle test (cull.cuh) and the survivors are ballot-compacted, in list
//                 order, into the buffer the caller's per-pixel loop reads.
#pragma once
#include "common.cuh"
#include "cull.cuh"

namespace rg {

constexpr int kGroup = 64;      // list entries per iteration: two per lane
constexpr int kBlendWarps = 2;  // warps per CTA: a 16x4 pixel strip of one tile

struct __align__(16) WarpStage {
    float4 cand_r0[kGroup];  // stage-1 survivors
    float4 xyts[kGroup];     // stage-2 survivors: {x, y, thr, bits(position in the walk)}
    float4 co[kGroup];       //                    {A, B, C, opacity}
    float4 rgbd[kGroup];     //                    {r, g, b, depth}
    uint32_t cand_id[kGroup];
    uint32_t id[kGroup];  // Gaussian id of each stage-2 survivor
    uint8_t cand_k[kGroup];  // position in the group of each candidate
};

// REVERSE = false: entries first, first+1, ...   (forward blend: front to back)
// REVERSE = true:  entries first+n-1, first+n-2, ...   (backward blend: back to front)
template <bool REVERSE>
struct ListWalker {
    const uint32_t* pl;
    const float4* rec;
    uint32_t first;
    int n, lane;
    float x0, y0, x1, y1;
    uint32_t id_next[2], id_after[2];
    float4 r0_next[2];
    static constexpr uint32_t kNone = 0xffffffffu;
    bool timing = false;  // performance analysis only (rg_debug_tile_profile)
    long long t_wait = 0, t_stage1 = 0, t_stage2 = 0;

    __device__ __forceinline__ uint32_t load_id(int k) const {
        return k < n ? pl[REVERSE ? first + (uint32_t)(n - 1 - k) : first + (uint32_t)k] : kNone;
    }
    __device__ __forceinline__ float4 load_r0(uint32_t id) const {
        return id != kNone ? rec[4 * (size_t)id] : float4{0.f, 0.f, -3.0e38f, -3.0e38f};
    }
    __device__ __forceinline__ void begin() {
#pragma unroll
        for (int j = 0; j < 2; j++) id_next[j] = load_id(j * 32 + lane);
#pragma unroll
        for (int j = 0; j < 2; j++) id_after[j] = load_id(kGroup + j * 32 + lane);
#pragma unroll
        for (int j = 0; j < 2; j++) r0_next[j] = load_r0(id_next[j]);
    }
    // Filters group g (walk positions [64 g, 64 g + 64)) into st; returns the number of survivors.  Position k of
    // the walk is list entry first + k (or first + n - 1 - k when REVERSE).
    __device__ __forceinline__ int next(int g, WarpStage& st) {
        uint32_t id_cur[2];
        float4 r0[2];
        long long tt = 0;
        if (timing) tt = clock64();
#pragma unroll
        for (int j = 0; j < 2; j++) {
            id_cur[j] = id_next[j];
            r0[j] = r0_next[j];
            id_next[j] = id_after[j];
        }
#pragma unroll
        for (int j = 0; j < 2; j++) id_after[j] = load_id((g + 2) * kGroup + j * 32 + lane);
#pragma unroll
        for (int j = 0; j < 2; j++) r0_next[j] = load_r0(id_next[j]);
        const unsigned lt = (1u << lane) - 1;
        if (timing) {
            // force the prefetched values to land so that the wait is attributed here
            asm volatile("" ::"r"(id_cur[0]), "r"(id_cur[1]), "f"(r0[0].x), "f"(r0[1].x), "r"(id_next[0]), "r"(id_next[1]));
            const long long t = clock64();
            t_wait += t - tt;
            tt = t;
        }
        // ---- stage 1
        int ncand = 0;
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const bool pass = id_cur[j] != kNone && cull_aabb(r0[j], x0, y0, x1, y1);
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (pass) {
                const int slot = ncand + __popc(m & lt);
                st.cand_r0[slot] = r0[j];
                st.cand_id[slot] = id_cur[j];
                st.cand_k[slot] = (uint8_t)(j * 32 + lane);
            }
            ncand += __popc(m);
        }
        if (timing) {
            const long long t = clock64();
            t_stage1 += t - tt;
            tt = t;
        }
        if (ncand == 0) return 0;
        __syncwarp();
        // ---- stage 2
        int nsurv = 0;
        for (int c0 = 0; c0 < ncand; c0 += 32) {
            const int c = c0 + lane;
            bool keep = false;
            uint32_t id = 0;
            float4 a = {0.f, 0.f, 0.f, 0.f}, co = a, rgbd = a, tail = a;
            int k = 0;
            if (c < ncand) {
                id = st.cand_id[c];
                a = st.cand_r0[c];
                k = st.cand_k[c];
                const float4* r = rec + 4 * (size_t)id;
                co = r[1];
                rgbd = r[2];
                tail = r[3];
                keep = cull_exact(a, co, tail.x, x0, y0, x1, y1);
            }
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int slot = nsurv + __popc(m & lt);
                st.xyts[slot] = {a.x, a.y, tail.y, __int_as_float(g * kGroup + k)};
                st.co[slot] = co;
                st.rgbd[slot] = rgbd;
                st.id[slot] = id;
            }
            nsurv += __popc(m);
        }
        __syncwarp();
        if (timing) t_stage2 += clock64() - tt;
        return nsurv;
    }
};

}  // namespace rg
