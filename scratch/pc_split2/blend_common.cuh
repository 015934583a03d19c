// Front end shared by the forward and backward blend kernels.
//
// One CTA = one 8x4 pixel block of a tile = two warps:
//   * the FILTER warp walks the tile's instance list and reduces it to the few instances whose footprint can reach
//     the block (ListWalker below), kGroup list entries at a time, into a small ring of survivor buffers in shared
//     memory;
//   * the BLEND warp (one pixel per lane) consumes the ring and runs the per-pixel recurrence.
// The two hand buffers over with a pair of mbarriers per ring slot (full / empty), so walking and blending overlap
// and no CTA-wide barrier is needed after start-up.  The heavy tiles of a frame are served by a handful of warps whose
// latency IS the kernel's duration (measured with rg_debug_tile_profile); walking is 30-45 % of that latency when one
// warp does both jobs.
//
// The walk is software-pipelined so that the dependent chain  list entry -> record  (two trips to L2) is hidden:
//   stage 1   bounding-box test on rec[0] of a group of entries (kPerLane per lane, registers), ballot-compacted
//             into a candidate buffer;
//   stage 2   candidates are spread over the lanes; each lane fetches the rest of its candidate's record, runs the
//             exact ellipse-vs-rectangle test (cull.cuh) and the survivors are ballot-compacted, in list order,
//             into the ring slot.
// A walker keeps in flight: the stage-2 records of its next group (stage 1 already ran), rec[0] of the one after,
// list entries of the one after that.  The filter warp runs TWO walkers on alternating groups, so that while one
// waits for its loads the other computes.
#pragma once
#include "bulk_copy.cuh"
#include "common.cuh"
#include "cull.cuh"

namespace rg {

constexpr int kStages = 2;         // ring slots between the filter warp and the blend warp
constexpr int kBlendThreads = 64;  // blend warp + filter warp

template <int kPerLane>
struct __align__(16) SurvStage {  // what the blend warp needs of each surviving instance
    static constexpr int kGroup = 32 * kPerLane;
    float4 xyts[kGroup];  // {x, y, thr, bits(position in the walk)}
    float4 co[kGroup];    // {A, B, C, opacity}
    float4 rgbd[kGroup];  // {r, g, b, depth}
    uint32_t id[kGroup];  // Gaussian id
};

template <int kPerLane>
struct __align__(16) CandBuf {  // stage-1 survivors, private to one walker
    static constexpr int kGroup = 32 * kPerLane;
    float4 r0[kGroup];
    uint32_t id[kGroup];
    uint8_t k[kGroup];  // position in the group
};

template <int kPerLane>
struct __align__(16) BlockPipe {
    SurvStage<kPerLane> stage[kStages];
    CandBuf<kPerLane> cand[2];
    uint64_t full[kStages], empty[kStages];
    int cnt[kStages];  // survivors in the slot; -1 = end of the walk
    int stop;          // set by the blend warp when its pixels need nothing more
    int total;         // backward: entries to replay (written by the blend warp before the start-up barrier)
};

template <int kPerLane>
__device__ __forceinline__ void pipe_init(BlockPipe<kPerLane>& bp) {
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kStages; s++) {
            mbar_init(&bp.full[s], 1);
            mbar_init(&bp.empty[s], 1);
        }
        bp.stop = 0;
    }
}

// REVERSE = false: entries first, first+1, ...   (forward blend: front to back)
// REVERSE = true:  entries first+n-1, first+n-2, ...   (backward blend: back to front)
// The walker visits groups g0, g0 + kStride, g0 + 2 kStride, ... of kGroup walk positions each.
template <bool REVERSE, int kPerLane, int kStride>
struct ListWalker {
    static constexpr int kGroup = 32 * kPerLane;
    using Cand = CandBuf<kPerLane>;
    using Surv = SurvStage<kPerLane>;
    const uint32_t* pl;
    const float4* rec;
    uint32_t first;
    int n, lane;
    float x0, y0, x1, y1;
    static constexpr uint32_t kNone = 0xffffffffu;

    // pipeline registers
    uint32_t id_1[kPerLane], id_2[kPerLane];  // list entries of this walker's next two groups (the second in flight)
    float4 r0_1[kPerLane];                    // rec[0] of the next group (in flight)
    int ncand;                                // candidates of the group whose stage 2 is pending
    float4 p_co, p_rgbd, p_tail;              // this lane's pending stage-2 record (in flight)

    __device__ __forceinline__ uint32_t load_id(int k) const {
        return k < n ? pl[REVERSE ? first + (uint32_t)(n - 1 - k) : first + (uint32_t)k] : kNone;
    }
    __device__ __forceinline__ float4 load_r0(uint32_t id) const {
        return id != kNone ? rec[4 * (size_t)id] : float4{0.f, 0.f, -3.0e38f, -3.0e38f};
    }
    // stage 1 of one group: registers -> candidate buffer; then request the first 32 candidates' records
    __device__ __forceinline__ void stage1(const uint32_t (&id)[kPerLane], const float4 (&r0)[kPerLane], Cand& cb) {
        const unsigned lt = (1u << lane) - 1;
        ncand = 0;
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            const bool pass = id[j] != kNone && cull_aabb(r0[j], x0, y0, x1, y1);
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (pass) {
                const int slot = ncand + __popc(m & lt);
                cb.r0[slot] = r0[j];
                cb.id[slot] = id[j];
                cb.k[slot] = (uint8_t)(j * 32 + lane);
            }
            ncand += __popc(m);
        }
        __syncwarp();
        if (lane < ncand) {
            const float4* r = rec + 4 * (size_t)cb.id[lane];
            p_co = r[1];
            p_rgbd = r[2];
            p_tail = r[3];
        }
    }
    // Prepares group g0 (stage 1 done, stage-2 records requested) and the loads of the two groups after it.
    __device__ __forceinline__ void begin(int g0, Cand& cb) {
        uint32_t id_0[kPerLane];
        float4 r0_0[kPerLane];
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            id_0[j] = load_id(g0 * kGroup + j * 32 + lane);
            id_1[j] = load_id((g0 + kStride) * kGroup + j * 32 + lane);
            id_2[j] = load_id((g0 + 2 * kStride) * kGroup + j * 32 + lane);
        }
#pragma unroll
        for (int j = 0; j < kPerLane; j++) {
            r0_0[j] = load_r0(id_0[j]);
            r0_1[j] = load_r0(id_1[j]);
        }
        stage1(id_0, r0_0, cb);
    }
    // Group g's candidates are in cb with their records in flight: true when at least one exists, i.e. when next()
    // will need an empty ring slot to write to.
    __device__ __forceinline__ bool pending() const { return ncand > 0; }
    // Finishes stage 2 of group g (walk positions [kGroup g, kGroup g + kGroup)) into `out` and returns the number of
    // survivors; then runs stage 1 of this walker's next group and issues the loads further down the list.  Position k
    // of the walk is list entry first + k (or first + n - 1 - k when REVERSE).
    __device__ __forceinline__ int next(int g, Cand& cb, Surv& out) {
        const unsigned lt = (1u << lane) - 1;
        int nsurv = 0;
        for (int c0 = 0; c0 < ncand; c0 += 32) {
            const int c = c0 + lane;
            bool keep = false;
            uint32_t id = 0;
            float4 a = {0.f, 0.f, 0.f, 0.f}, co = a, rgbd = a, tail = a;
            int k = 0;
            if (c < ncand) {
                id = cb.id[c];
                a = cb.r0[c];
                k = cb.k[c];
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
                out.xyts[slot] = {a.x, a.y, tail.y, __int_as_float(g * kGroup + k)};
                out.co[slot] = co;
                out.rgbd[slot] = rgbd;
                out.id[slot] = id;
            }
            nsurv += __popc(m);
        }
        __syncwarp();
        if ((g + kStride) * kGroup < n) {
            stage1(id_1, r0_1, cb);
#pragma unroll
            for (int j = 0; j < kPerLane; j++) {
                id_1[j] = id_2[j];
                r0_1[j] = load_r0(id_2[j]);
                id_2[j] = load_id((g + 3 * kStride) * kGroup + j * 32 + lane);
            }
        } else {
            ncand = 0;
        }
        return nsurv;
    }
};

// The filter warp's whole job: walk n entries of the list with two interleaved walkers, publish every non-empty group of
// survivors in walk order, then the end marker.  Returns early when the blend warp raises bp.stop.
template <bool REVERSE, int kPerLane>
__device__ __forceinline__ void run_filter_warp(const uint32_t* pl, const float4* rec, uint32_t first, int n, int lane,
                                                float x0, float y0, float x1, float y1, BlockPipe<kPerLane>& bp,
                                                uint32_t* prof_out = nullptr) {
    long long t_we = 0, t_nx = 0, t_all = prof_out ? clock64() : 0;
    using Walker = ListWalker<REVERSE, kPerLane, 2>;
    constexpr int kGroup = Walker::kGroup;
    Walker wa, wb;
    wa.pl = pl; wa.rec = rec; wa.first = first; wa.n = n; wa.lane = lane;
    wa.x0 = x0; wa.y0 = y0; wa.x1 = x1; wa.y1 = y1; wa.ncand = 0;
    wb = wa;
    const int n_groups = (n + kGroup - 1) / kGroup;
    volatile int* stop = &bp.stop;
    int slot = 0;
    uint32_t lap = 0;
    bool have_slot = false;
    auto wait_empty = [&]() -> bool {  // false: the blend warp has stopped
        while (!mbar_try_wait(&bp.empty[slot], (lap & 1) ^ 1))
            if (*stop) return false;
        return !*stop;
    };
    auto step = [&](Walker& w, int g, CandBuf<kPerLane>& cb) -> bool {
        long long t0 = prof_out ? clock64() : 0;
        if (w.pending() && !have_slot) {
            if (!wait_empty()) return false;
            have_slot = true;
        }
        long long t1 = prof_out ? clock64() : 0;
        const int nsurv = w.next(g, cb, bp.stage[slot]);
        if (prof_out) { const long long t2 = clock64(); t_we += t1 - t0; t_nx += t2 - t1; }
        if (nsurv > 0) {
            if (lane == 0) bp.cnt[slot] = nsurv;
            __syncwarp();
            if (lane == 0) mbar_arrive(&bp.full[slot]);
            have_slot = false;
            if (++slot == kStages) { slot = 0; lap++; }
        }
        return true;
    };
    if (n_groups > 0) wa.begin(0, bp.cand[0]);
    if (n_groups > 1) wb.begin(1, bp.cand[1]);
    for (int g = 0; g < n_groups; g += 2) {
        if (*stop) return;
        if (!step(wa, g, bp.cand[0])) return;
        if (g + 1 < n_groups && !step(wb, g + 1, bp.cand[1])) return;
    }
    if (prof_out && lane == 0) { prof_out[2] = (uint32_t)t_we; prof_out[3] = (uint32_t)t_nx; prof_out[7] = (uint32_t)(clock64() - t_all); }
    if (!have_slot && !wait_empty()) return;
    if (lane == 0) {
        bp.cnt[slot] = -1;
        mbar_arrive(&bp.full[slot]);
    }
}

// Blend-warp side of the ring.
struct RingReader {
    int slot = 0;
    uint32_t lap = 0;
    // Blocks until the next slot is published; returns its survivor count (-1: end of the walk).
    template <int kPerLane>
    __device__ __forceinline__ int acquire(BlockPipe<kPerLane>& bp) {
        while (!mbar_try_wait(&bp.full[slot], lap & 1)) {}
        return *reinterpret_cast<volatile int*>(&bp.cnt[slot]);
    }
    template <int kPerLane>
    __device__ __forceinline__ void release(BlockPipe<kPerLane>& bp, int lane) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bp.empty[slot]);
        if (++slot == kStages) { slot = 0; lap++; }
    }
};

}  // namespace rg
