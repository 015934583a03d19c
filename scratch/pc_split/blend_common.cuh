// Front end shared by the forward and backward blend kernels.
//
// One CTA = one 8x4 pixel block of a tile = two warps:
//   * the FILTER warp walks the tile's instance list and reduces it to the few instances whose footprint can reach
//     the block (ListWalker below), 64 list entries at a time, into a small ring of survivor buffers in shared memory;
//   * the BLEND warp (one pixel per lane) consumes the ring and runs the per-pixel recurrence.
// The two hand buffers over with a pair of mbarriers per ring slot (full / empty), so walking and blending overlap
// and no CTA-wide barrier is needed after start-up.  The heavy tiles of a frame are walked by a handful of warps
// whose latency IS the kernel's duration; splitting the walk from the recurrence halves that critical path.
//
// The walk itself is software-pipelined so that the dependent chain  list entry -> record  never stalls it:
//   stage 1   bounding-box test on rec[0] of a group of 64 entries (two per lane, registers), ballot-compacted into
//             a candidate buffer;
//   stage 2   candidates are spread over the lanes; each lane fetches the rest of its candidate's record, runs the
//             exact ellipse-vs-rectangle test (cull.cuh) and the survivors are ballot-compacted, in list order,
//             into the ring slot.
// While stage 2 of group g runs, these requests are in flight: stage-2 records of group g+1 (its stage 1 already
// ran), rec[0] of group g+2, list entries of group g+3.
#pragma once
#include "bulk_copy.cuh"
#include "common.cuh"
#include "cull.cuh"

namespace rg {

constexpr int kGroup = 64;   // list entries per walk step: two per lane
constexpr int kStages = 3;   // ring slots between the filter warp and the blend warp
constexpr int kBlendThreads = 64;

struct __align__(16) SurvStage {  // what the blend warp needs of each surviving instance
    float4 xyts[kGroup];          // {x, y, thr, bits(position in the walk)}
    float4 co[kGroup];            // {A, B, C, opacity}
    float4 rgbd[kGroup];          // {r, g, b, depth}
    uint32_t id[kGroup];          // Gaussian id
};

struct __align__(16) CandBuf {  // stage-1 survivors, private to the filter warp
    float4 r0[kGroup];
    uint32_t id[kGroup];
    uint8_t k[kGroup];  // position in the group
};

struct __align__(16) BlockPipe {
    SurvStage stage[kStages];
    CandBuf cand;
    uint64_t full[kStages], empty[kStages];
    int cnt[kStages];  // survivors in the slot; -1 = end of the walk
    int stop;          // set by the blend warp when its pixels need nothing more
    int total;         // backward: entries to replay (written by the blend warp before the start-up barrier)
};

__device__ __forceinline__ void pipe_init(BlockPipe& bp) {
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
template <bool REVERSE>
struct ListWalker {
    const uint32_t* pl;
    const float4* rec;
    uint32_t first;
    int n, lane;
    float x0, y0, x1, y1;
    static constexpr uint32_t kNone = 0xffffffffu;

    // pipeline registers
    uint32_t id_1[2], id_2[2];    // list entries of the next two groups (the second possibly in flight)
    float4 r0_1[2];               // rec[0] of the next group (possibly in flight)
    int ncand;                    // candidates of the group whose stage 2 is pending
    float4 p_co, p_rgbd, p_tail;  // this lane's pending stage-2 record (in flight)

    __device__ __forceinline__ uint32_t load_id(int k) const {
        return k < n ? pl[REVERSE ? first + (uint32_t)(n - 1 - k) : first + (uint32_t)k] : kNone;
    }
    __device__ __forceinline__ float4 load_r0(uint32_t id) const {
        return id != kNone ? rec[4 * (size_t)id] : float4{0.f, 0.f, -3.0e38f, -3.0e38f};
    }
    // stage 1 of one group: registers -> candidate buffer; then request the first 32 candidates' records
    __device__ __forceinline__ void stage1(const uint32_t (&id)[2], const float4 (&r0)[2], CandBuf& cb) {
        const unsigned lt = (1u << lane) - 1;
        ncand = 0;
#pragma unroll
        for (int j = 0; j < 2; j++) {
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
    __device__ __forceinline__ void begin(CandBuf& cb) {
        uint32_t id_0[2];
        float4 r0_0[2];
#pragma unroll
        for (int j = 0; j < 2; j++) {
            id_0[j] = load_id(j * 32 + lane);
            id_1[j] = load_id(kGroup + j * 32 + lane);
            id_2[j] = load_id(2 * kGroup + j * 32 + lane);
        }
#pragma unroll
        for (int j = 0; j < 2; j++) {
            r0_0[j] = load_r0(id_0[j]);
            r0_1[j] = load_r0(id_1[j]);
        }
        stage1(id_0, r0_0, cb);
    }
    // Group g's candidates (walk positions [64 g, 64 g + 64)) are in cb with their records in flight: true when at
    // least one exists, i.e. when next() will need an empty ring slot to write to.
    __device__ __forceinline__ bool pending() const { return ncand > 0; }
    // Finishes stage 2 of group g into `out` and returns the number of survivors; then runs stage 1 of group g+1 and
    // issues the loads further down the list.  Position k of the walk is list entry first + k (or first + n - 1 - k
    // when REVERSE).
    __device__ __forceinline__ int next(int g, CandBuf& cb, SurvStage& out) {
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
        if ((g + 1) * kGroup < n) {
            stage1(id_1, r0_1, cb);
#pragma unroll
            for (int j = 0; j < 2; j++) {
                id_1[j] = id_2[j];
                r0_1[j] = load_r0(id_2[j]);
                id_2[j] = load_id((g + 3) * kGroup + j * 32 + lane);
            }
        } else {
            ncand = 0;
        }
        return nsurv;
    }
};

// The filter warp's whole job: walk n entries of the list, publish every non-empty group of survivors, then the end
// marker.  Returns early when the blend warp raises bp.stop.
template <bool REVERSE>
__device__ __forceinline__ void run_filter_warp(ListWalker<REVERSE>& walk, BlockPipe& bp) {
    const int lane = walk.lane;
    const int n_groups = (walk.n + kGroup - 1) / kGroup;
    volatile int* stop = &bp.stop;
    int slot = 0;
    uint32_t lap = 0;
    auto wait_empty = [&]() -> bool {  // false: the blend warp has stopped
        while (!mbar_try_wait(&bp.empty[slot], (lap & 1) ^ 1))
            if (*stop) return false;
        return !*stop;
    };
    if (n_groups > 0) walk.begin(bp.cand);
    bool have_slot = false;
    for (int g = 0; g < n_groups; g++) {
        if (*stop) return;
        if (walk.pending() && !have_slot) {
            if (!wait_empty()) return;
            have_slot = true;
        }
        const int nsurv = walk.next(g, bp.cand, bp.stage[slot]);
        if (nsurv > 0) {
            if (lane == 0) bp.cnt[slot] = nsurv;
            __syncwarp();
            if (lane == 0) mbar_arrive(&bp.full[slot]);
            have_slot = false;
            if (++slot == kStages) { slot = 0; lap++; }
        }
    }
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
    __device__ __forceinline__ int acquire(BlockPipe& bp) {
        while (!mbar_try_wait(&bp.full[slot], lap & 1)) {}
        return *reinterpret_cast<volatile int*>(&bp.cnt[slot]);
    }
    __device__ __forceinline__ void release(BlockPipe& bp, int lane) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bp.empty[slot]);
        if (++slot == kStages) { slot = 0; lap++; }
    }
};

}  // namespace rg
