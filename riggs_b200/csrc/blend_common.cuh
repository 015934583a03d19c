// Front end shared by the forward and backward blend kernels: one warp walks one tile's instance list on behalf of
// its 8x4 pixel block and reduces it to the few instances whose footprint can reach the block.
//
// The walk is software-pipelined so that the dependent chain  list entry -> record  (two trips to L2) never sits on
// the warp's critical path -- the heavy tiles of a frame are walked by a handful of warps whose latency IS the
// kernel's duration:
//   stage 1   bounding-box test on rec[0] of a group of 128 entries (four per lane, registers), ballot-compacted into
//             a candidate buffer;
//   stage 2   candidates are spread over the lanes; each lane fetches the rest of its candidate's record, runs the
//             exact ellipse-vs-rectangle test (cull.cuh) and the survivors are ballot-compacted, in list order,
//             into the buffer the caller's per-pixel loop reads.
// While the caller blends the survivors of group g, these requests are in flight: stage-2 records of group g+1
// (its stage 1 already ran), rec[0] of group g+2, list entries of group g+3.
// The forward walker also raises one bit per survivor in its block's bitmap over the instance list; the backward blend
// walks by those bits (MaskWalker) instead of testing again.
#pragma once
#include "common.cuh"
#include "cull.cuh"

namespace rg {

constexpr int kBlendWarps = 2;  // warps per CTA: a 16x4 pixel strip of one tile

// kPerLane list entries per lane and walk step (kGroup = 32 kPerLane per step): the forward kernel, whose walk is a
// third of its instructions, takes 4; the backward kernel, register-bound in its replay loop, takes 2.
// kCull = false (the backward's bitmap walk): no candidate records and no survivor bits are staged.
template <int kPerLane, bool kCull = true>
struct __align__(16) WarpStage {
    static constexpr int kGroup = 32 * kPerLane;
    float4 cand_r0[kCull ? kGroup : 1];  // stage-1 survivors
    float4 xyts[kGroup];     // stage-2 survivors: {x, y, thr, bits(position in the walk)}
    float4 co[kGroup];       //                    {A, B, C, opacity}
    float4 rgbd[kGroup];     //                    {r, g, b, depth}
    uint32_t cand_id[kGroup];
    uint32_t id[kGroup];  // Gaussian id of each stage-2 survivor
    uint8_t cand_k[kGroup];  // position in the group of each candidate
    uint32_t bits[kCull ? kPerLane + 2 : 1];  // survivor bits of the group being filtered (forward)
};

// REVERSE = false: entries first, first+1, ...   (forward blend: front to back)
// REVERSE = true:  entries first+n-1, first+n-2, ...   (backward blend: back to front)
template <bool REVERSE, int kPerLane>
struct ListWalker {
    static constexpr int kGroup = 32 * kPerLane;
    using Stage = WarpStage<kPerLane>;
    const uint32_t* pl;
    const float4* rec;
    uint32_t first;
    int n, lane;
    float x0, y0, x1, y1;
    uint32_t* surv_bits = nullptr;  // if set (forward): this block's bitmap over the instance list, one bit per entry,
                                    // raised for every survivor; the backward blend walks by these bits
    static constexpr uint32_t kNone = 0xffffffffu;
    bool timing = false;  // performance analysis only (rg_debug_tile_profile)
    long long t_wait = 0, t_stage1 = 0, t_stage2 = 0;

    // pipeline registers
    uint32_t id_1[kPerLane], id_2[kPerLane];  // list entries of the next two groups (the second possibly in flight)
    float4 r0_1[kPerLane];             // rec[0] of the next group (possibly in flight)
    int ncand;                           // candidates of the group whose stage 2 is pending
    float4 p_co, p_rgbd, p_tail;         // this lane's pending stage-2 record (in flight)

    __device__ __forceinline__ uint32_t load_id(int k) const {
        return k < n ? pl[REVERSE ? first + (uint32_t)(n - 1 - k) : first + (uint32_t)k] : kNone;
    }
    __device__ __forceinline__ float4 load_r0(uint32_t id) const {
        return id != kNone ? rec[4 * (size_t)id] : float4{0.f, 0.f, -3.0e38f, -3.0e38f};
    }
    // stage 1 of one group: registers -> candidate buffer; then request the first 32 candidates' records
    __device__ __forceinline__ void stage1(const uint32_t (&id)[kPerLane], const float4 (&r0)[kPerLane], Stage& st) {
        const unsigned lt = (1u << lane) - 1;
        ncand = 0;
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            const bool pass = id[j] != kNone && cull_aabb(r0[j], x0, y0, x1, y1);
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (pass) {
                const int slot = ncand + __popc(m & lt);
                st.cand_r0[slot] = r0[j];
                st.cand_id[slot] = id[j];
                st.cand_k[slot] = (uint8_t)(j * 32 + lane);
            }
            ncand += __popc(m);
        }
        __syncwarp();
        if (lane < ncand) {
            const float4* r = rec + 4 * (size_t)st.cand_id[lane];
            p_co = r[1];
            p_rgbd = r[2];
            p_tail = r[3];
        }
    }
    __device__ __forceinline__ void begin(Stage& st) {
        uint32_t id_0[kPerLane];
        float4 r0_0[kPerLane];
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            id_0[j] = load_id(j * 32 + lane);
            id_1[j] = load_id(kGroup + j * 32 + lane);
            id_2[j] = load_id(2 * kGroup + j * 32 + lane);
        }
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            r0_0[j] = load_r0(id_0[j]);
            r0_1[j] = load_r0(id_1[j]);
        }
        stage1(id_0, r0_0, st);
    }
    // Hands out group g (walk positions [128 g, 128 g + 128)): finishes its stage 2 into st and returns the number of
    // survivors; runs stage 1 of group g+1 and issues the loads further down the list before returning.  Position k
    // of the walk is list entry first + k (or first + n - 1 - k when REVERSE).
    __device__ __forceinline__ int next(int g, Stage& st) {
        const unsigned lt = (1u << lane) - 1;
        long long tt = 0;
        if (timing) tt = clock64();
        if (surv_bits && lane < kPerLane + 2) st.bits[lane] = 0;
        __syncwarp();  // every lane is done reading the previous group's survivors
        // ---- finish stage 2 of group g
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
                if (c0 == 0) {
                    co = p_co; rgbd = p_rgbd; tail = p_tail;
                } else {  // more than 32 candidates in one group: not pipelined
                    const float4* r = rec + 4 * (size_t)id;
                    co = r[1]; rgbd = r[2]; tail = r[3];
                }
                keep = cull_exact(a, co, tail, x0, y0, x1, y1);
            }
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int slot = nsurv + __popc(m & lt);
                st.xyts[slot] = {a.x, a.y, tail.y, __int_as_float(g * kGroup + k)};
                st.co[slot] = co;
                st.rgbd[slot] = rgbd;
                st.id[slot] = id;
                if (surv_bits) atomicOr(&st.bits[((first & 31u) + (uint32_t)k) >> 5], 1u << ((first + (uint32_t)k) & 31u));
            }
            nsurv += __popc(m);
        }
        __syncwarp();
        if (surv_bits && nsurv > 0 && lane < kPerLane + 2) {
            // list index of walk position p is first + p (forward walks only); kGroup is a multiple of 32, so the group's
            // bits start at bit (first & 31) of word (first >> 5) + g * kPerLane.  Neighbouring tiles share boundary words.
            const uint32_t wbits = st.bits[lane];
            if (wbits) atomicOr(surv_bits + (first >> 5) + (uint32_t)(g * kPerLane + lane), wbits);
        }
        if (timing) {
            const long long t = clock64();
            t_stage2 += t - tt;
            tt = t;
        }
        // ---- stage 1 of group g+1, then the requests for g+1 (stage 2), g+2 (rec[0]) and g+3 (list entries)
        if ((g + 1) * kGroup < n) {
            if (timing) {
                asm volatile("" ::"r"(id_1[0]), "r"(id_1[1]), "f"(r0_1[0].x), "f"(r0_1[1].x));
                const long long t = clock64();
                t_wait += t - tt;
                tt = t;
            }
            stage1(id_1, r0_1, st);
#pragma unroll
            for (int j = 0; j < kPerLane; j++) {
                id_1[j] = id_2[j];
                r0_1[j] = load_r0(id_2[j]);
                id_2[j] = load_id((g + 3) * kGroup + j * 32 + lane);
            }
            if (timing) t_stage1 += clock64() - tt;
        } else {
            ncand = 0;
        }
        return nsurv;
    }
};

// Backward-direction walker over the survivor bits the forward blend left behind: the entries of the tile's list
// whose byte has bit `blk` set are exactly those the forward walker kept for this block, so neither the bounding-box
// nor the ellipse test is repeated, and records are fetched for survivors only.  Walk position p is list entry
// first + n - 1 - p.  Same interface and the same stage buffer as ListWalker.
template <int kPerLane>
struct MaskWalker {
    static constexpr int kGroup = 32 * kPerLane;
    using Stage = WarpStage<kPerLane, false>;
    const uint32_t* pl;
    const float4* rec;
    const uint32_t* surv_bits;  // this block's bitmap (see ListWalker::surv_bits)
    uint32_t first;
    int n, lane;
    static constexpr uint32_t kNone = 0xffffffffu;
    bool timing = false;
    long long t_wait = 0, t_stage1 = 0, t_stage2 = 0;

    uint32_t id_1[kPerLane];   // list entries of the next group (in flight)
    uint32_t bit_1[kPerLane];  // their survivor bits
    int ncand;
    float4 p_r0, p_co, p_rgbd, p_tail;  // this lane's pending record (in flight)

    __device__ __forceinline__ void load_group(int g) {
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            const int p = g * kGroup + j * 32 + lane;
            id_1[j] = kNone;
            bit_1[j] = 0;
            if (p < n) {
                const uint32_t at = first + (uint32_t)(n - 1 - p);
                id_1[j] = pl[at];
                bit_1[j] = (surv_bits[at >> 5] >> (at & 31u)) & 1u;
            }
        }
    }
    __device__ __forceinline__ void stage1(Stage& st) {
        const unsigned lt = (1u << lane) - 1;
        ncand = 0;
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            const bool pass = bit_1[j] != 0;
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (pass) {
                const int slot = ncand + __popc(m & lt);
                st.cand_id[slot] = id_1[j];
                st.cand_k[slot] = (uint8_t)(j * 32 + lane);
            }
            ncand += __popc(m);
        }
        __syncwarp();
        if (lane < ncand) {
            const float4* r = rec + 4 * (size_t)st.cand_id[lane];
            p_r0 = r[0]; p_co = r[1]; p_rgbd = r[2]; p_tail = r[3];
        }
    }
    __device__ __forceinline__ void begin(Stage& st) {
        load_group(0);
        stage1(st);
        load_group(1);
    }
    __device__ __forceinline__ int next(int g, Stage& st) {
        long long tt = 0;
        if (timing) tt = clock64();
        __syncwarp();  // every lane is done reading the previous group's survivors
        for (int c0 = 0; c0 < ncand; c0 += 32) {
            const int c = c0 + lane;
            if (c < ncand) {
                const uint32_t id = st.cand_id[c];
                const int k = st.cand_k[c];
                float4 a, co, rgbd, tail;
                if (c0 == 0) {
                    a = p_r0; co = p_co; rgbd = p_rgbd; tail = p_tail;
                } else {  // more than 32 survivors in one group: not pipelined
                    const float4* r = rec + 4 * (size_t)id;
                    a = r[0]; co = r[1]; rgbd = r[2]; tail = r[3];
                }
                st.xyts[c] = {a.x, a.y, tail.y, __int_as_float(g * kGroup + k)};
                st.co[c] = co;
                st.rgbd[c] = rgbd;
                st.id[c] = id;
            }
        }
        const int nsurv = ncand;
        __syncwarp();
        if (timing) {
            const long long t = clock64();
            t_stage2 += t - tt;
            tt = t;
        }
        if ((g + 1) * kGroup < n) {
            stage1(st);
            load_group(g + 2);
            if (timing) t_stage1 += clock64() - tt;
        } else {
            ncand = 0;
        }
        return nsurv;
    }
};

}  // namespace rg
