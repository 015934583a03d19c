// The per-Gaussian MLP front end of the skinning path as tensor-core kernels for sm_100a (tcgen05 + TMEM):
//   WeightMLP  (skeleton_utils/network_utils.py:73-107):  x -> PE(10) 63 -> 8 x 256 ReLU (skip after layer 4) -> J-1, sigmoid
//   DeformMLP  (network_utils.py:6-65, "detail_net"):     (x PE(4) 27, pose 4J) -> 8 x 256 ReLU (skip) -> 3
// The reference runs them as ~20 eager torch kernels each way with fp32 cuBLAS GEMMs over all N points and keeps every
// (N, 256) fp32 activation in HBM.  Here ONE kernel carries a tile of 128 points through all nine layers:
//   * activations never leave the SM: the bf16 activation matrix of a tile lives in TENSOR MEMORY and is the A operand
//     of the next layer's tcgen05.mma (A from TMEM, B = weights from shared memory, fp32 accumulators in TMEM);
//   * a layer's 256 outputs are produced in four quarters of 64 columns with two accumulator buffers, so the four
//     epilogue warps (TMEM -> registers, bias, ReLU, bf16, -> TMEM) work on quarter q while the tensor core computes
//     quarter q+1, and layer l+1 starts on the k-slabs whose quarters are already written;
//   * weights stream through a ring of 8 KB chunks (64 outputs x 64 inputs, bf16, pre-swizzled by mlp_pack_kernel) filled
//     by 1-D bulk (TMA) copies from a producer warp, freed by tcgen05.commit;
//   * the positional encoding is computed in registers, and a pose that is the same for every point (detail_net's second
//     input) is folded into the biases of the two layers that see it by the caller.
// Arithmetic: bf16 operands, fp32 accumulation (the reference: fp32 cuBLAS).  Tolerances are stated in DESIGN.md and
// tests/test_gpu_mlp.py.
//
// Training: the forward also stores every layer's post-ReLU activations, bf16, FEATURE-major ([layer][feature][point]),
// which is the K-major operand layout the weight-gradient GEMMs want; mlp_chain_kernel<true> runs the same pipeline
// backwards (dH = dZ W, masked by the stored activations) and stores dZ the same way; mlp_wgrad_kernel contracts the two
// over the points.
#include "../../include/riggs_b200.h"

#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <string>

#include "bulk_copy.cuh"
#include "common.cuh"

namespace rg {

#ifndef RG_MLP_PROF
#define RG_MLP_PROF 0  // 1: CTA 0 records cycle counts of its second tile (tuning builds only)
#endif
#if RG_MLP_PROF
__device__ unsigned long long g_mlp_prof[4096];
#define PROF_ON (blockIdx.x == 0 && t == 1)
#endif
constexpr int kHid = 256;        // hidden width of the reference's MLPs (W = 256)
constexpr int kLayers = 8;       // D = 8 hidden layers, then the head
constexpr int kSkipInto = 5;     // the embedded input is concatenated in front of layer 5's input (skips = [D // 2])
constexpr int kTileM = 128;      // points per tile = TMEM lanes
constexpr int kQuarter = 64;     // output columns per accumulator buffer = input columns per k-slab
constexpr int kChunkBytes = kQuarter * kQuarter * 2;  // one weight chunk: 64 outputs x 64 inputs, bf16 = 8 KB
constexpr int kStageBytes = 2 * kChunkBytes;  // a ring stage holds TWO chunks (consecutive k-slabs of one quarter): one copy, one
                                              // wait, one descriptor and one commit per eight tcgen05.mma instead of per four
constexpr int kStages = 9;       // ring depth: 144 KB of weights in flight / resident
constexpr int kEpiWarps = 8;   // one epilogue group: two warps per TMEM lane quarter, each takes half of a quarter's 64 columns
constexpr int kEpiGroups = 2;  // group 0 takes the quarters that land in accumulator buffer 0, group 1 those in buffer 1:
                               // a quarter's epilogue (TMEM load, bias / ReLU / pack, TMEM store, the training-state
                               // stores) is a ~2000-cycle latency chain per warp, and with ONE group the four quarters of
                               // a layer queued behind each other while the tensor core waited for the last one
constexpr int kMmaWarps = 2;   // issuing warp i feeds accumulator buffer i (the quarters with (g & 1) == i): issuing a weight
                               // chunk's four tcgen05.mma (waits, descriptors, commit) is ~360 cycles of one warp's serial
                               // instruction stream against 128 cycles of tensor-core work -- the kernel's bound with one issuer
constexpr int kStageWarps = 4;  // warps of epilogue group 0 that also stage a tile's input region, one per TMEM lane quarter
constexpr int kMlpThreads = (kEpiGroups * kEpiWarps + kMmaWarps + 1) * 32;  // + the weight producer warp

// TMEM columns (32-bit each, 128 lanes): two fp32 accumulator quarters, two bf16 activation buffers (256 values = 128
// columns each), the embedded input (64 values = 32 columns)
constexpr uint32_t kColAcc = 0, kColAct = 128, kColEmb = 384, kEmbCols = 64, kTmemCols = 512;  // two input regions of 64 columns

struct MlpLayerPlan {
    int n_quarters;  // output quarters (4 for hidden layers; 1 or 2 for the head)
    int n_slabs;     // input k-slabs of 64: [input-region slabs] + [4 activation slabs?]
    int n_emb;       // leading slabs that come from the tile's input region (forward: the encoding; backward: dL/d head)
    int has_act;     // the other slabs are the previous step's activations
    int n_last;      // columns of the last quarter (multiple of 16)
};

struct MlpArgs {
    int N;                 // points
    int n_tiles;
    int n_steps;           // forward: 9 (eight hidden layers + head); backward: 8
    int multires;          // positional-encoding octaves (10: WeightMLP, 4: DeformMLP); embedding = 3 + 6 * multires <= 63
    int out_ch;            // head outputs
    int out_act;           // 0: none, 1: sigmoid
    int total_chunks;      // weight chunks per tile
    const float* x;        // (N, 3)
    const unsigned char* w_packed;  // chunk images in consumption order
    const float* bias;     // (9, 256): hidden layers' biases, then the head's (zero padded)
    float* out;            // (N, out_ch)
    __nv_bfloat16* acts;   // training: (8, 256, n_pad) post-ReLU activations, feature-major; null: not stored
    __nv_bfloat16* emb;    // training: (64, n_pad) embedded input, feature-major; null: not stored
    uint32_t* relu_mask;   // training: (8, 8, n_pad) words: bit i of word [layer][2 q + half][point] = unit 64 q + 32 half + i is on
    long long n_pad;       // row pitch of acts / emb / dz (points, multiple of 128)
    // backward chain (mlp_chain_kernel<true>): dL/d out -> dL/d(pre-activation) of every layer
    const float* g_out;    // (N, out_ch) upstream gradient
    const float* out_fwd;  // (N, out_ch) forward output (for the sigmoid's derivative), or null
    __nv_bfloat16* dz;     // (9, 256, n_pad): slot l = dL/dZ_l of hidden layer l, slot 8 = dL/dZ of the head; feature-major
    MlpLayerPlan plan[kLayers + 1];
};

// ---------------------------------------------------------------------------------------------------------------------
// PTX wrappers (tcgen05 / TMEM)
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// arrives on the mbarrier once every tcgen05 operation issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T, bf16 x bf16 -> fp32, M = 128
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// one lane of a converged warp (the warp stays converged around it: the issuing loop runs on all 32 lanes with
// warp-uniform values, which the compiler keeps in uniform registers -- under `if (lane == 0)` it wrapped every
// tcgen05 instruction in an election loop, ~90 instructions per weight chunk on the kernel's critical path)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
        "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, "
        "%18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
        "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
        "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);  // .x (low half) = lo
    return *reinterpret_cast<uint32_t*>(&v);
}

// Shared-memory matrix descriptor of one weight chunk's k-step: K-major rows of 128 bytes (64 bf16), 128-byte swizzle,
// eight-row groups 1024 bytes apart; the k-step of 16 columns starts 32 bytes further into the (un-swizzled) row.
__device__ __forceinline__ uint64_t b_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3ffff) >> 4);  // start address, bits [0,14)
    d |= (uint64_t)1 << 16;                       // leading byte offset (unused for swizzled K-major), bits [16,30)
    d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset: 8 rows x 128 B, bits [32,46)
    d |= (uint64_t)1 << 46;                       // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
    return d;
}
// instruction descriptor: D fp32, A/B bf16, both K-major, M = 128, N = n
__device__ __forceinline__ uint32_t instr_desc(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

struct MlpSmem {
    uint64_t full[kStages], empty[kStages];
    uint64_t acc_full[2], acc_free[2];
    uint64_t act_ready[2][4];
    uint64_t emb_ready[2];
    uint32_t tmem_base;
    alignas(16) float bias[(kLayers + 1) * kHid];  // read as float4
    // head outputs of one warp's 32 points x 32 columns on their way to global memory (pitch 33: conflict-free both ways)
    float out_stage[kEpiGroups * kEpiWarps][32 * 33];
};

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(
            taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}

// 16 packed bf16 pairs = 32 consecutive features of one point -> feature-major global memory (coalesced over the warp's points)
__device__ __forceinline__ void store_features(__nv_bfloat16* dst, long long pitch, const uint32_t (&pk)[16]) {
    unsigned short* p = reinterpret_cast<unsigned short*>(dst);
#pragma unroll
    for (int i = 0; i < 16; i++) {
        p[0] = (unsigned short)(pk[i] & 0xffffu);
        p[pitch] = (unsigned short)(pk[i] >> 16);
        p += 2 * pitch;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// The chain kernel.  Persistent: CTA b handles tiles b, b + gridDim.x, ...
// BWD = false: forward (encoding -> 8 hidden layers -> head).
// BWD = true:  the same pipeline run backwards: step 0 multiplies dL/d(head pre-activation) by the head's weights, step j by
//              layer 8-j's; the epilogue applies ReLU' (one bit per unit, left behind by the forward) instead of adding a
//              bias, and stores dL/dZ of every layer for the weight-gradient kernel.  No gradient is wanted for x (it is
//              detached, skeleton_warp.py:131), so the chain stops at dZ_0.
// Warps 0..7: epilogue -- warp e owns TMEM lanes 32 (e & 3) .. +31 (a warp reaches only the lane quarter warp % 4) and
// the column half e >> 2 of every 64-column quarter; warp 8: MMA issuer; warp 9: weight producer.
template <bool BWD>
__global__ void __launch_bounds__(kMlpThreads, 1) mlp_chain_kernel(MlpArgs a) {
    extern __shared__ __align__(1024) unsigned char s_dyn[];
    unsigned char* ring = reinterpret_cast<unsigned char*>(align_up_dev(reinterpret_cast<size_t>(s_dyn), 1024));
    MlpSmem& sm = *reinterpret_cast<MlpSmem*>(ring + (size_t)kStages * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
        // epilogue -> tensor-core barriers count WARPS: every thread fences its TMEM accesses, the warp converges and one
        // lane arrives (256 single-thread arrivals on one shared-memory word serialise: three such barriers per quarter
        // were ~1500 cycles of the ~2300 a quarter took)
        for (int i = 0; i < 2; i++) { mbar_init(&sm.acc_full[i], 1); mbar_init(&sm.acc_free[i], kEpiWarps); }
        for (int i = 0; i < 8; i++) mbar_init(&sm.act_ready[i >> 2][i & 3], kEpiWarps);
        for (int i = 0; i < 2; i++) mbar_init(&sm.emb_ready[i], kStageWarps);
    }
    if (!BWD)
        for (int i = threadIdx.x; i < (kLayers + 1) * kHid; i += blockDim.x) sm.bias[i] = a.bias[i];
    constexpr int kMmaWarp = kEpiGroups * kEpiWarps, kProducerWarp = kMmaWarp + kMmaWarps;
    if (warp == kMmaWarp) tmem_alloc(&sm.tmem_base, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;
    const int my_tiles = a.n_tiles > (int)blockIdx.x ? (a.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

    if (warp == kProducerWarp) {
        // ===================== producer: weight chunks, the same sequence for every tile =====================
        if (lane == 0) {
            uint32_t c = 0;  // ring stages handed out so far
            for (int t = 0; t < my_tiles; t++) {
                int j = 0;   // chunk index inside the packed image (consumption order)
                for (int l = 0; l < a.n_steps; l++) {
                    const int n_slabs = a.plan[l].n_slabs, n_q = a.plan[l].n_quarters;
                    for (int q = 0; q < n_q; q++) {
                        for (int s = 0; s < n_slabs; s += 2, c++) {
                            const int n = min(2, n_slabs - s);  // the quarter's slabs s, s + 1 are adjacent chunks of the image
                            const uint32_t st = c % kStages;
                            mbar_wait(&sm.empty[st], ((c / kStages) & 1) ^ 1);  // passes at once the first time round
                            bulk_load(ring + (size_t)st * kStageBytes, a.w_packed + (size_t)j * kChunkBytes, (uint32_t)(n * kChunkBytes),
                                      &sm.full[st]);
                            j += n;
                        }
                    }
                }
            }
        }
    } else if (warp >= kMmaWarp && warp < kProducerWarp) {
        // ===================== MMA issuers: the whole warp runs the loop, one elected lane issues =====================
        {
            const uint32_t iw = (uint32_t)(warp - kMmaWarp);
            uint32_t c = 0, g = 0;
            for (int t = 0; t < my_tiles; t++) {
                mbar_wait(&sm.emb_ready[t & 1], (t >> 1) & 1);  // the tile's input region: written while the tile before it ran
                tc_fence_after();
#pragma unroll 1
                for (int l = 0; l < a.n_steps; l++) {
                    const MlpLayerPlan p = a.plan[l];
                    const uint32_t in_col = kColAct + (uint32_t)(l & 1) * 128;  // step l reads buffer l & 1 (l >= 1)
                    const int n_use = (l - 1) >> 1;  // how often this tile has written that buffer before (parity of the wait)
                    bool slabs_seen = false;  // this warp has waited for the step's input slabs
#pragma unroll 1
                    for (int q = 0; q < p.n_quarters; q++, g++) {
                        const uint32_t ab = g & 1;
                        if (ab != iw) { c += (uint32_t)((p.n_slabs + 1) >> 1); continue; }  // the other issuing warp's quarter
#if RG_MLP_PROF
                        long long pt0 = clock64(), w_act = 0, w_full = 0;
#endif
                        mbar_wait(&sm.acc_free[ab], ((g >> 1) & 1) ^ 1);
                        tc_fence_after();
#if RG_MLP_PROF
                        long long pt1 = clock64();
#endif
                        const int n = (q == p.n_quarters - 1) ? p.n_last : kQuarter;
                        const uint32_t idesc = instr_desc(n);
                        const uint32_t d_col = tmem + kColAcc + ab * kQuarter;
                        for (int s0 = 0; s0 < p.n_slabs; s0 += 2, c++) {
                            const int n_here = min(2, p.n_slabs - s0);  // slabs in this ring stage
                            uint32_t a_cols[2] = {0u, 0u};
#pragma unroll
                            for (int h = 0; h < 2; h++) {
                                const int s = s0 + h;
                                if (h >= n_here) break;
                                if (s < p.n_emb) {
                                    a_cols[h] = tmem + kColEmb + (uint32_t)(t & 1) * kEmbCols + (uint32_t)s * 32;
                                } else {
                                    const int sa = s - p.n_emb;  // activation slab = output quarter of the previous step
                                    if (!slabs_seen) {  // this warp's later quarters of the step read slabs it has already waited for
#if RG_MLP_PROF
                                        long long x0 = clock64();
#endif
                                        mbar_wait(&sm.act_ready[l & 1][sa], n_use & 1);
                                        tc_fence_after();
#if RG_MLP_PROF
                                        w_act += clock64() - x0;
#endif
                                    }
                                    a_cols[h] = tmem + in_col + (uint32_t)sa * 32;
                                }
                            }
                            const uint32_t st = c % kStages;
#if RG_MLP_PROF
                            long long y0 = clock64();
#endif
                            mbar_wait(&sm.full[st], (c / kStages) & 1);
                            tc_fence_after();
#if RG_MLP_PROF
                            w_full += clock64() - y0;
#endif
                            const uint64_t bd = b_desc(smem_u32(ring + (size_t)st * kStageBytes));
                            if (elect_one()) {
#pragma unroll
                                for (int k = 0; k < 4; k++)  // the k-step of 16 columns starts 32 bytes (2 address units) further on
                                    mma_ts(d_col, a_cols[0] + (uint32_t)k * 8, bd + (uint64_t)(k * 2), idesc, (s0 | k) != 0);
                                if (n_here == 2) {
#pragma unroll
                                    for (int k = 0; k < 4; k++)  // the second chunk of the stage: 8 KB = 512 address units further on
                                        mma_ts(d_col, a_cols[1] + (uint32_t)k * 8, bd + (uint64_t)(kChunkBytes / 16 + k * 2), idesc, 1u);
                                }
                                tc_commit(&sm.empty[st]);  // the stage is free once these MMAs have read it
                            }
                            __syncwarp();
                        }
                        if (elect_one()) tc_commit(&sm.acc_full[ab]);
                        __syncwarp();
                        slabs_seen = true;
#if RG_MLP_PROF
                        if (PROF_ON && lane == 0) {
                            unsigned long long* o = g_mlp_prof + (size_t)(l * 4 + q) * 8;
                            o[0] = (unsigned long long)pt0; o[1] = (unsigned long long)(pt1 - pt0); o[2] = (unsigned long long)w_act;
                            o[3] = (unsigned long long)w_full; o[4] = (unsigned long long)(clock64() - pt0);
                        }
#endif
                    }
                }
            }
        }
    } else {
        // ===================== epilogue warps: thread = point (TMEM lane) x half of a quarter's columns =====================
        const int lq = warp & 3, ch = (warp >> 2) & 1;
        const uint32_t group = (uint32_t)(warp >> 3);  // its quarters are those with (g & 1) == group
        const uint32_t lane_base = (uint32_t)(lq * 32) << 16;
        const int row = lq * 32 + lane;
        uint32_t g = 0;
        const int n_emb = 3 + 6 * a.multires;
        // The input region of a tile (forward: the positional encoding; backward: dL/d head) is written one tile AHEAD, into
        // the other of two TMEM regions, by four warps of group 0 (one per TMEM lane quarter) before they turn to the
        // current tile's quarters: the tensor core never waits for it.
        auto stage_input = [&](int t) {
            const long long pt = ((long long)blockIdx.x + (long long)t * gridDim.x) * kTileM + row;
            const bool valid = pt < a.N;
            const bool in_buf = pt < a.n_pad;
            const uint32_t emb_col = kColEmb + (uint32_t)(t & 1) * kEmbCols;
            if (!BWD) {
                // ---- positional encoding (utils/time_utils.py:208-256: x, then sin / cos per octave), bf16, into TMEM:
                //      this thread packs its point's 64 features.  One sincosf per coordinate, the higher octaves by
                //      angle doubling (sin 2t = 2 sin t cos t, cos 2t = 1 - 2 sin^2 t): the error doubles per octave, ~1e-4
                //      at the tenth against the 4e-3 of the bf16 operand it becomes; thirty accurate sincosf per thread were
                //      a ~10 000-cycle block in front of every tile
                float e[64];
                float xv[3] = {0.f, 0.f, 0.f};
                if (valid) { xv[0] = a.x[3 * pt]; xv[1] = a.x[3 * pt + 1]; xv[2] = a.x[3 * pt + 2]; }
#pragma unroll
                for (int i = 0; i < 64; i++) e[i] = 0.f;
                e[0] = xv[0]; e[1] = xv[1]; e[2] = xv[2];
                float sv[3], cv[3];
#pragma unroll
                for (int d = 0; d < 3; d++) sincosf(xv[d], &sv[d], &cv[d]);
#pragma unroll
                for (int o = 0; o < 10; o++) {
                    if (o < a.multires) {
#pragma unroll
                        for (int d = 0; d < 3; d++) {
                            e[3 + 6 * o + d] = sv[d];
                            e[6 + 6 * o + d] = cv[d];
                            const float s2 = 2.f * sv[d] * cv[d], c2 = fmaf(-2.f * sv[d], sv[d], 1.f);
                            sv[d] = s2; cv[d] = c2;
                        }
                    }
                }
                uint32_t pk[32];
#pragma unroll
                for (int i = 0; i < 32; i++) pk[i] = pack_bf16(e[2 * i], e[2 * i + 1]);
                tmem_st32(tmem + lane_base + emb_col, pk);
                if (a.emb && in_buf) {  // rows of points past N are written as zeros: the weight-gradient GEMM sums over them
#pragma unroll
                    for (int i = 0; i < 64; i++)
                        if (i < n_emb) a.emb[(size_t)i * a.n_pad + pt] = __float2bfloat16_rn(valid ? e[i] : 0.f);
                }
            } else {
                // ---- dL/d(head pre-activation) of this point: the upstream gradient (times the sigmoid's derivative)
                const int n_in = a.plan[0].n_emb;  // slabs of 64 head outputs
#pragma unroll 1
                for (int sc = 0; sc < 2 * n_in; sc++) {
                    const int sl = sc >> 1, ch = sc & 1;
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        float g2[2];
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int col = sl * 64 + ch * 32 + 2 * i + h;
                            float gv = 0.f;
                            if (valid && col < a.out_ch) {
                                gv = a.g_out[(size_t)pt * a.out_ch + col];
                                if (a.out_act == 1) {
                                    const float sg = a.out_fwd[(size_t)pt * a.out_ch + col];
                                    gv *= sg * (1.f - sg);
                                }
                            }
                            g2[h] = gv;
                        }
                        pk[i] = pack_bf16(g2[0], g2[1]);
                    }
                    tmem_st16(tmem + lane_base + emb_col + (uint32_t)sl * 32 + (uint32_t)ch * 16, pk);
                    if (in_buf) store_features(a.dz + ((size_t)kLayers * kHid + (size_t)sl * 64 + (size_t)ch * 32) * a.n_pad + pt, a.n_pad, pk);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.emb_ready[t & 1]);
        };
        const bool stages = group == 0 && ch == 0;
        if (stages && my_tiles > 0) stage_input(0);
        // deferred part of the head's epilogue (see there)
        float* const stg = sm.out_stage[warp];
        long long pend_pt0 = 0;
        int pend_r = 0, pend_rows = 0, pend_cols = 0, pend_col0 = 0;
        float pend_bias = 0.f;
        auto head_piece = [&]() {
            float z[8];
#pragma unroll
            for (int u = 0; u < 8; u++) z[u] = stg[min(pend_r + u, 31) * 33 + lane] + pend_bias;
            if (a.out_act == 1) {
#pragma unroll
                for (int u = 0; u < 8; u++) z[u] = __fdividef(1.0f, 1.0f + __expf(-z[u]));
            }
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (pend_r + u < pend_rows && lane < pend_cols)
                    a.out[(size_t)(pend_pt0 + pend_r + u) * a.out_ch + pend_col0 + lane] = z[u];
            pend_r += 8;
        };
        for (int t = 0; t < my_tiles; t++) {
            const long long pt = ((long long)blockIdx.x + (long long)t * gridDim.x) * kTileM + row;
            const bool valid = pt < a.N;
            const bool in_buf = pt < a.n_pad;
            if (stages && t + 1 < my_tiles) stage_input(t + 1);
#pragma unroll 1
            for (int l = 0; l < a.n_steps; l++) {
                const MlpLayerPlan p = a.plan[l];
                const uint32_t out_col = kColAct + (uint32_t)((l + 1) & 1) * 128;  // step l writes buffer (l + 1) & 1
                const float* bias = sm.bias + l * kHid;
#pragma unroll 1
                for (int q = 0; q < p.n_quarters; q++, g++) {
                    const uint32_t ab = g & 1;
                    if (ab != group) continue;
#if RG_MLP_PROF
                    long long e0 = clock64();
#endif
                    mbar_wait(&sm.acc_full[ab], (g >> 1) & 1);
                    tc_fence_after();
#if RG_MLP_PROF
                    long long e1 = clock64();
#endif
                    uint32_t v[32];
                    tmem_ld32(tmem + lane_base + kColAcc + ab * kQuarter + (uint32_t)ch * 32, v);
                    tmem_ld_wait();
                    // the accumulator is in registers: hand the buffer back before doing anything else (the tensor core is
                    // waiting for it two quarters ahead)
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.acc_free[ab]);
                    const int col0 = q * kQuarter + ch * 32;  // first of this thread's 32 columns
                    if (BWD || l < kLayers) {
                        uint32_t pk[16];
                        // forward: this layer's slot; backward step l produces dZ of hidden layer 7 - l
                        const int slot = BWD ? (kLayers - 1 - l) : l;
                        uint32_t* mword = a.relu_mask + (((size_t)slot * 8 + (size_t)(q * 2 + ch)) * a.n_pad + (size_t)pt);
                        uint32_t bits = 0;
                        if (BWD) bits = valid ? *mword : 0u;  // one bit per unit: was its activation positive?
                        if (!BWD) {
                            const float4* b4 = reinterpret_cast<const float4*>(bias + col0);
#pragma unroll
                            for (int i = 0; i < 8; i++) {
                                const float4 bb = b4[i];
                                const float z0 = __uint_as_float(v[4 * i]) + bb.x, z1 = __uint_as_float(v[4 * i + 1]) + bb.y;
                                const float z2 = __uint_as_float(v[4 * i + 2]) + bb.z, z3 = __uint_as_float(v[4 * i + 3]) + bb.w;
                                // one bit per unit (a positive fp32 value never rounds to a bf16 zero: same exponent range)
                                bits |= (z0 > 0.f ? 1u : 0u) << (4 * i) | (z1 > 0.f ? 1u : 0u) << (4 * i + 1) |
                                        (z2 > 0.f ? 1u : 0u) << (4 * i + 2) | (z3 > 0.f ? 1u : 0u) << (4 * i + 3);
                                pk[2 * i] = pack_bf16(fmaxf(z0, 0.f), fmaxf(z1, 0.f));
                                pk[2 * i + 1] = pack_bf16(fmaxf(z2, 0.f), fmaxf(z3, 0.f));
                            }
                            if (!valid) {  // rows past N (last tile only): zeros, so that the weight gradients sum nothing
                                bits = 0;
#pragma unroll
                                for (int i = 0; i < 16; i++) pk[i] = 0u;
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; i++) {
                                const float h0 = (bits >> (2 * i)) & 1u ? __uint_as_float(v[2 * i]) : 0.f;
                                const float h1 = (bits >> (2 * i + 1)) & 1u ? __uint_as_float(v[2 * i + 1]) : 0.f;
                                pk[i] = pack_bf16(h0, h1);
                            }
                        }
                        tmem_st16(tmem + lane_base + out_col + (uint32_t)q * 32 + (uint32_t)ch * 16, pk);
                        tmem_st_wait();
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&sm.act_ready[(l + 1) & 1][q]);  // the next step may start on this k-slab
#if RG_MLP_PROF
                        if (PROF_ON && (warp & 7) == 0 && lane == 0) {
                            unsigned long long* o = g_mlp_prof + 1024 + (size_t)(l * 4 + q) * 8;
                            o[0] = (unsigned long long)e0; o[1] = (unsigned long long)(e1 - e0); o[2] = (unsigned long long)(clock64() - e1);
                        }
#endif
                        __nv_bfloat16* store = BWD ? a.dz : a.acts;  // training state, feature-major
                        if (store && in_buf) {
                            store_features(store + ((size_t)slot * kHid + (size_t)col0) * a.n_pad + (size_t)pt, a.n_pad, pk);
                            if (!BWD) *mword = bits;
                        }
                        if (!BWD && pend_r < pend_rows) head_piece();
                    } else {
                        // The head's outputs leave through shared memory, transposed: a thread holds one point's columns; a
                        // lane of head_piece() takes one COLUMN, so that a store instruction writes one point's columns side
                        // by side.  Only the staging happens here: the loop over the points (exp, divide, store: ~6000 cycles
                        // for one warp) is cut into pieces of eight points that run behind the warp's NEXT quarters, where it
                        // has slack -- in one block it held up the first quarters of every following tile.
                        while (pend_r < pend_rows) head_piece();  // (the previous tile's, if any is left)
                        const int n_cols = min(32, a.out_ch - col0);  // this warp's columns of the head (<= 0: none)
                        if (n_cols > 0) {
#pragma unroll
                            for (int i = 0; i < 32; i++) stg[lane * 33 + i] = __uint_as_float(v[i]);
                            __syncwarp();
                            pend_pt0 = pt - lane;  // the warp's first point
                            pend_rows = (int)min((long long)32, a.N - pend_pt0);
                            pend_r = 0;
                            pend_cols = n_cols;
                            pend_col0 = col0;
                            pend_bias = lane < n_cols ? bias[col0 + lane] : 0.f;
                        }
#if RG_MLP_PROF
                        if (PROF_ON && (warp & 7) == 0 && lane == 0) {
                            unsigned long long* o = g_mlp_prof + 1024 + (size_t)(l * 4 + q) * 8;
                            o[0] = (unsigned long long)e0; o[1] = (unsigned long long)(e1 - e0); o[2] = (unsigned long long)(clock64() - e1);
                        }
#endif
                    }
                }
            }
        }
        while (pend_r < pend_rows) head_piece();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem, kTmemCols);
}

// ---------------------------------------------------------------------------------------------------------------------
// Weight packing: fp32 (out, in) matrices of torch.nn.Linear -> bf16 chunk images in the order the chain kernel consumes
// them.  Forward image: chunk (layer, quarter q, slab s) holds rows n = 64 q + r of W_l and 64 of its input columns.
// Backward image (transposed): step j multiplies by W^T -- chunk (step, q, s) holds rows k = 64 q + r (an INPUT feature of
// the forward layer) and columns n = 64 s + kk (its outputs): B[k][n] = W[n][act_col0 + k].
struct PackLayer {
    const float* w;   // (n_out, ld) row-major
    int n_out, ld;
    int emb_col0, emb_cols;  // forward: columns of the embedded-input part (-1: none)
    int act_col0;            // first column of the 256 activation inputs (-1: none)
    int chunk0;              // index of the step's first chunk
    int n_quarters, n_slabs;
};
struct PackArgs {
    PackLayer layer[kLayers + 1];
    unsigned char* out;
    int total_chunks, n_steps, transposed;
};

__global__ void mlp_pack_kernel(PackArgs a) {
    // one thread per 16-byte piece: 8 columns of one row of one chunk
    const long long piece = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_pieces = (long long)a.total_chunks * (kChunkBytes / 16);
    if (piece >= n_pieces) return;
    const int chunk = (int)(piece / (kChunkBytes / 16));
    const int in_chunk = (int)(piece % (kChunkBytes / 16));
    const int r = in_chunk >> 3, cchunk = in_chunk & 7;  // row of the chunk, logical 16-byte column group
    int l = 0;
    while (l + 1 < a.n_steps && chunk >= a.layer[l + 1].chunk0) l++;
    const PackLayer L = a.layer[l];
    const int rel = chunk - L.chunk0;
    const int q = rel / L.n_slabs, s = rel % L.n_slabs;
    const int row = q * kQuarter + r;
    __nv_bfloat16 v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int kk = cchunk * 8 + j;  // column inside the slab
        float w = 0.f;
        if (!a.transposed) {
            if (row < L.n_out) {
                if (L.emb_col0 >= 0 && s == 0) {
                    if (kk < L.emb_cols) w = L.w[(size_t)row * L.ld + L.emb_col0 + kk];
                } else {
                    const int sa = s - (L.emb_col0 >= 0 ? 1 : 0);
                    w = L.w[(size_t)row * L.ld + L.act_col0 + sa * kQuarter + kk];
                }
            }
        } else {
            const int n = s * kQuarter + kk;  // forward output = contraction index of the backward step
            if (n < L.n_out) w = L.w[(size_t)n * L.ld + L.act_col0 + row];
        }
        v[j] = __float2bfloat16_rn(w);
    }
    // 128-byte swizzle: the 16-byte piece c of row r sits at piece position c ^ (r & 7)
    unsigned char* dst = a.out + (size_t)chunk * kChunkBytes + (size_t)r * 128 + (size_t)((cchunk ^ (r & 7)) * 16);
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
}

// forward plan: 8 hidden layers + head
static void plan_forward(int out_ch, MlpLayerPlan* plan, int* chunk0, int* total) {
    int c = 0;
    for (int l = 0; l <= kLayers; l++) {
        MlpLayerPlan p;
        p.n_emb = (l == 0 || l == kSkipInto) ? 1 : 0;
        p.has_act = l > 0 ? 1 : 0;
        p.n_slabs = p.n_emb + 4 * p.has_act;
        if (l < kLayers) { p.n_quarters = 4; p.n_last = kQuarter; }
        else {
            const int out_pad = (out_ch + 15) / 16 * 16;
            p.n_quarters = (out_pad + kQuarter - 1) / kQuarter;
            p.n_last = out_pad - (p.n_quarters - 1) * kQuarter;
        }
        plan[l] = p;
        chunk0[l] = c;
        c += p.n_quarters * p.n_slabs;
    }
    *total = c;
}
// backward plan: step 0 = through the head (its input region holds dL/d head, 1 or 2 slabs), steps 1..7 = through layers 7..1
static void plan_backward(int out_ch, MlpLayerPlan* plan, int* chunk0, int* total) {
    int c = 0;
    for (int j = 0; j < kLayers; j++) {
        MlpLayerPlan p;
        p.n_emb = j == 0 ? (out_ch + kQuarter - 1) / kQuarter : 0;
        p.has_act = j > 0 ? 1 : 0;
        p.n_slabs = p.n_emb + 4 * p.has_act;
        p.n_quarters = 4;
        p.n_last = kQuarter;
        plan[j] = p;
        chunk0[j] = c;
        c += p.n_quarters * p.n_slabs;
    }
    *total = c;
}

// ---------------------------------------------------------------------------------------------------------------------
// Weight gradients: dW_l[n][k] = sum over the points of dZ_l[pt][n] * H_{l-1}[pt][k]  (H_{-1} = the encoding; layer 5 sees
// both), db_l[n] = sum dZ_l[pt][n].  Both operands were stored feature-major, i.e. K-major for this contraction, so they come
// in as plain 2-D TMA tiles (128-byte swizzle) and go straight into tcgen05.mma from shared memory.  One CTA owns one
// (layer, half of the outputs) pair and one slice of the points; its 128 x (256 + 64 + 16) fp32 result lives in TMEM and is
// added to the fp32 gradient tensors (torch layout) with atomics at the end.  The bias gradient is the product with a
// tile of ones.
constexpr int kWgStages = 3;
constexpr int kWgATile = 128 * 128;   // dZ^T tile: 128 outputs x 64 points, bf16
constexpr int kWgBTile = 256 * 128;   // H^T tile: 256 inputs x 64 points
constexpr int kWgETile = 64 * 128;    // encoding^T tile: 64 x 64 points
constexpr int kWgStageBytes = kWgATile + kWgBTile + kWgETile;
constexpr int kWgThreads = 6 * 32;    // producer, MMA issuer, four epilogue warps
constexpr uint32_t kWgColAct = 0, kWgColEmb = 256, kWgColBias = 320;

struct WgItem {
    int layer;      // 0..8 (8 = head)
    int m0;         // first output row of this CTA's half
    int rows;       // valid output rows from m0 on (<= 128)
    int has_act, has_emb;
    float* dw;      // (n_out, ld) fp32
    int ld, emb_col0, emb_cols, act_col0;
    float* db;      // (n_out)
};
struct WgArgs {
    WgItem item[2 * (kLayers + 1)];
    int n_items, n_splits;
    long long n_slabs;  // n_pad / 64
};

struct WgSmem {
    uint64_t full[kWgStages], empty[kWgStages], done;
    uint32_t tmem_base;
};

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(tmap), "r"(x), "r"(y), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc_, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc_), "r"(idesc), "r"(accumulate)
        : "memory");
}

__global__ void __launch_bounds__(kWgThreads, 1)
mlp_wgrad_kernel(WgArgs a, const __grid_constant__ CUtensorMap tm_dz, const __grid_constant__ CUtensorMap tm_act,
                 const __grid_constant__ CUtensorMap tm_emb) {
    extern __shared__ __align__(1024) unsigned char s_dyn[];
    unsigned char* ring = reinterpret_cast<unsigned char*>(align_up_dev(reinterpret_cast<size_t>(s_dyn), 1024));
    unsigned char* ones = ring + (size_t)kWgStages * kWgStageBytes;  // 16 rows x 128 bytes of bf16 1.0
    WgSmem& sm = *reinterpret_cast<WgSmem*>(ones + 16 * 128);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const WgItem it = a.item[blockIdx.x / a.n_splits];
    const int split = blockIdx.x % a.n_splits;
    const long long per = (a.n_slabs + a.n_splits - 1) / a.n_splits;
    const long long s0 = per * split, s1 = s0 + per < a.n_slabs ? s0 + per : a.n_slabs;
    const int n_my = s1 > s0 ? (int)(s1 - s0) : 0;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kWgStages; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], 1); }
        mbar_init(&sm.done, 1);
    }
    for (int i = threadIdx.x; i < 16 * 64; i += blockDim.x) reinterpret_cast<__nv_bfloat16*>(ones)[i] = __float2bfloat16_rn(1.0f);
    if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the ones tile is read by the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;
    const uint32_t stage_bytes = kWgATile + (it.has_act ? kWgBTile : 0) + (it.has_emb ? kWgETile : 0);

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < n_my; i++) {
                const int st = i % kWgStages;
                mbar_wait(&sm.empty[st], ((i / kWgStages) & 1) ^ 1);
                unsigned char* base = ring + (size_t)st * kWgStageBytes;
                const int x = (int)((s0 + i) * 64);
                mbar_expect_tx(&sm.full[st], stage_bytes);
                tma_load_2d(base, &tm_dz, x, it.layer * kHid + it.m0, &sm.full[st]);
                if (it.has_act) tma_load_2d(base + kWgATile, &tm_act, x, (it.layer - 1) * kHid, &sm.full[st]);
                if (it.has_emb) tma_load_2d(base + kWgATile + kWgBTile, &tm_emb, x, 0, &sm.full[st]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t id_act = instr_desc(256), id_emb = instr_desc(64), id_one = instr_desc(16);
            const uint32_t ones_addr = smem_u32(ones);
            for (int i = 0; i < n_my; i++) {
                const int st = i % kWgStages;
                mbar_wait(&sm.full[st], (i / kWgStages) & 1);
                tc_fence_after();
                const uint32_t base = smem_u32(ring + (size_t)st * kWgStageBytes);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint64_t ad = b_desc(base + (uint32_t)k * 32);
                    const uint32_t acc = (i | k) != 0;
                    if (it.has_act) mma_ss(tmem + kWgColAct, ad, b_desc(base + kWgATile + (uint32_t)k * 32), id_act, acc);
                    if (it.has_emb) mma_ss(tmem + kWgColEmb, ad, b_desc(base + kWgATile + kWgBTile + (uint32_t)k * 32), id_emb, acc);
                    mma_ss(tmem + kWgColBias, ad, b_desc(ones_addr), id_one, acc);  // every point weighs 1: the k-step does not matter
                }
                tc_commit(&sm.empty[st]);
            }
            tc_commit(&sm.done);
        }
    } else if (n_my > 0) {
        // ---- epilogue: thread = output row; add the tile to the fp32 gradient tensors
        const int ew = warp & 3;  // a warp reaches the TMEM lanes 32 (warp % 4) .. +31 only: warps 2, 3, 4, 5 -> quarters 2, 3, 0, 1
        const uint32_t lane_base = (uint32_t)(ew * 32) << 16;
        const int r = ew * 32 + lane;
        mbar_wait(&sm.done, 0);
        tc_fence_after();
        const bool row_ok = r < it.rows;
        float* dwrow = it.dw + (size_t)(it.m0 + r) * it.ld;
        if (it.has_act) {
#pragma unroll 1
            for (int c0 = 0; c0 < 256; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem + lane_base + kWgColAct + (uint32_t)c0, v);
                tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < 32; i++) atomicAdd(dwrow + it.act_col0 + c0 + i, __uint_as_float(v[i]));
                }
            }
        }
        if (it.has_emb) {
#pragma unroll 1
            for (int c0 = 0; c0 < 64; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem + lane_base + kWgColEmb + (uint32_t)c0, v);
                tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < 32; i++)
                        if (c0 + i < it.emb_cols) atomicAdd(dwrow + it.emb_col0 + c0 + i, __uint_as_float(v[i]));
                }
            }
        }
        {
            uint32_t v[32];
            tmem_ld32(tmem + lane_base + kWgColBias, v);  // 16 columns are meaningful, all equal
            tmem_ld_wait();
            if (row_ok) atomicAdd(it.db + it.m0 + r, __uint_as_float(v[0]));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// 2-D bf16 tensor map over a feature-major (rows, n_pad) array: box = 64 points x box_rows features, 128-byte swizzle
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}
static int make_tmap(CUtensorMap* tm, void* base, long long rows, long long n_pad, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) { set_error("rg_mlp_backward: cuTensorMapEncodeTiled is not available from this driver"); return -1; }
    cuuint64_t dims[2] = {(cuuint64_t)n_pad, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)n_pad * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("rg_mlp_backward: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")"); return -1; }
    return 0;
}

}  // namespace rg

using namespace rg;

extern "C" {

size_t rg_mlp_packed_bytes(int out_ch) {
    MlpLayerPlan plan[kLayers + 1];
    int chunk0[kLayers + 1], tf = 0, tb = 0;
    if (out_ch < 1 || out_ch > 128) return 0;
    plan_forward(out_ch, plan, chunk0, &tf);
    plan_backward(out_ch, plan, chunk0, &tb);
    return (size_t)(tf + tb) * kChunkBytes;  // forward image, then the transposed image of the backward chain
}

int rg_mlp_pack(int emb_cols, int out_ch, const float* const* weights, const int* lds, const int* emb_col0, const int* act_col0,
                void* packed, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (emb_cols < 1 || emb_cols > 63 || out_ch < 1 || out_ch > 128 || !weights || !lds || !emb_col0 || !act_col0 || !packed) {
        set_error("rg_mlp_pack: bad argument (embedding of at most 63 columns, at most 128 outputs)");
        return -1;
    }
    for (int l = 0; l <= kLayers; l++)
        if (!weights[l]) { set_error("rg_mlp_pack: null weight matrix"); return -1; }
    MlpLayerPlan plan[kLayers + 1];
    int chunk0[kLayers + 1], total = 0;
    LaunchScope ls("mlp_pack", stream, 2);
    // ---- forward image
    plan_forward(out_ch, plan, chunk0, &total);
    PackArgs a;
    for (int l = 0; l <= kLayers; l++) {
        PackLayer& L = a.layer[l];
        L.w = weights[l]; L.ld = lds[l]; L.n_out = l < kLayers ? kHid : out_ch;
        L.emb_col0 = plan[l].n_emb ? emb_col0[l] : -1; L.emb_cols = emb_cols;
        L.act_col0 = plan[l].has_act ? act_col0[l] : -1;
        L.chunk0 = chunk0[l]; L.n_quarters = plan[l].n_quarters; L.n_slabs = plan[l].n_slabs;
    }
    a.out = static_cast<unsigned char*>(packed);
    a.total_chunks = total; a.n_steps = kLayers + 1; a.transposed = 0;
    long long pieces = (long long)total * (kChunkBytes / 16);
    mlp_pack_kernel<<<(unsigned)((pieces + 255) / 256), 256, 0, stream>>>(a);
    // ---- backward image: step 0 uses the head's weights, step j layer 8 - j's (activation columns only)
    const size_t fwd_bytes = (size_t)total * kChunkBytes;
    plan_backward(out_ch, plan, chunk0, &total);
    for (int j = 0; j < kLayers; j++) {
        const int l = kLayers - j;  // 8 = head, 7, ..., 1
        PackLayer& L = a.layer[j];
        L.w = weights[l]; L.ld = lds[l]; L.n_out = l < kLayers ? kHid : out_ch;
        L.emb_col0 = -1; L.emb_cols = 0; L.act_col0 = act_col0[l];
        L.chunk0 = chunk0[j]; L.n_quarters = plan[j].n_quarters; L.n_slabs = plan[j].n_slabs;
    }
    a.out = static_cast<unsigned char*>(packed) + fwd_bytes;
    a.total_chunks = total; a.n_steps = kLayers; a.transposed = 1;
    pieces = (long long)total * (kChunkBytes / 16);
    mlp_pack_kernel<<<(unsigned)((pieces + 255) / 256), 256, 0, stream>>>(a);
    return check_launch("rg_mlp_pack", stream, false);
}

static int launch_chain(bool bwd, MlpArgs& a, cudaStream_t stream) {
    const size_t smem = (size_t)kStages * kStageBytes + sizeof(MlpSmem) + 1024;
    const void* fn = bwd ? reinterpret_cast<const void*>(mlp_chain_kernel<true>) : reinterpret_cast<const void*>(mlp_chain_kernel<false>);
    if (ensure_dyn_smem(fn, (int)smem)) return -1;
    const int grid = std::min(a.n_tiles, RG_NUM_SMS);
    if (bwd) mlp_chain_kernel<true><<<grid, kMlpThreads, smem, stream>>>(a);
    else mlp_chain_kernel<false><<<grid, kMlpThreads, smem, stream>>>(a);
    return 0;
}

#if RG_MLP_PROF
extern "C" int rg_debug_mlp_prof_read(unsigned long long* host, int n) {
    return cudaMemcpyFromSymbol(host, g_mlp_prof, sizeof(unsigned long long) * (size_t)n) == cudaSuccess ? 0 : -1;
}
#endif

int rg_mlp_forward(int N, int multires, int out_ch, int out_act, const float* x, const void* packed, const float* bias,
                   float* out, void* acts, void* emb, void* relu_mask, long long n_pad, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (N < 0 || multires < 0 || multires > 10 || out_ch < 1 || out_ch > 128 || !x || !packed || !bias || !out) {
        set_error("rg_mlp_forward: bad argument");
        return -1;
    }
    if ((acts || emb || relu_mask) && (n_pad < N || (n_pad & 127) || !acts || !emb || !relu_mask)) {
        set_error("rg_mlp_forward: training state needs acts, emb and relu_mask, n_pad a multiple of 128 and >= N");
        return -1;
    }
    if (N == 0) return 0;
    MlpArgs a = {};
    int chunk0[kLayers + 1];
    plan_forward(out_ch, a.plan, chunk0, &a.total_chunks);
    a.n_steps = kLayers + 1;
    a.N = N; a.n_tiles = (N + kTileM - 1) / kTileM; a.multires = multires; a.out_ch = out_ch; a.out_act = out_act;
    a.x = x; a.w_packed = static_cast<const unsigned char*>(packed); a.bias = bias; a.out = out;
    a.acts = static_cast<__nv_bfloat16*>(acts); a.emb = static_cast<__nv_bfloat16*>(emb); a.n_pad = n_pad;
    a.relu_mask = static_cast<uint32_t*>(relu_mask);
    LaunchScope ls("mlp_forward", stream);
    if (launch_chain(false, a, stream)) return -1;
    return check_launch("rg_mlp_forward", stream, false);
}

int rg_mlp_backward(int N, int out_ch, int out_act, int emb_cols, const float* g_out, const float* out_fwd, const void* packed,
                    const void* acts, const void* emb, const void* relu_mask, void* dz, long long n_pad, float* const* d_weights, const int* lds,
                    const int* emb_col0, const int* act_col0, float* d_bias, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (N < 0 || out_ch < 1 || out_ch > 128 || !g_out || !packed || !acts || !emb || !relu_mask || !dz || !d_weights || !lds || !emb_col0 ||
        !act_col0 || !d_bias || (out_act == 1 && !out_fwd) || n_pad < N || (n_pad & 127)) {
        set_error("rg_mlp_backward: bad argument");
        return -1;
    }
    if (N == 0) return 0;
    // ---- 1. the chain, backwards: dL/dZ of every layer
    MlpArgs a = {};
    MlpLayerPlan fplan[kLayers + 1];
    int chunk0[kLayers + 1], fwd_chunks = 0;
    plan_forward(out_ch, fplan, chunk0, &fwd_chunks);
    plan_backward(out_ch, a.plan, chunk0, &a.total_chunks);
    a.n_steps = kLayers;
    a.N = N; a.n_tiles = (N + kTileM - 1) / kTileM; a.out_ch = out_ch; a.out_act = out_act;
    a.w_packed = static_cast<const unsigned char*>(packed) + (size_t)fwd_chunks * kChunkBytes;
    a.acts = static_cast<__nv_bfloat16*>(const_cast<void*>(acts));
    a.n_pad = n_pad; a.g_out = g_out; a.out_fwd = out_fwd; a.dz = static_cast<__nv_bfloat16*>(dz);
    a.relu_mask = static_cast<uint32_t*>(const_cast<void*>(relu_mask));
    {
        LaunchScope ls("mlp_backward_chain", stream);
        if (launch_chain(true, a, stream)) return -1;
    }
    // ---- 2. weight and bias gradients
    CUtensorMap tm_dz, tm_act, tm_emb;
    if (make_tmap(&tm_dz, dz, (long long)(kLayers + 1) * kHid, n_pad, 128)) return -1;
    if (make_tmap(&tm_act, const_cast<void*>(acts), (long long)kLayers * kHid, n_pad, 256)) return -1;
    if (make_tmap(&tm_emb, const_cast<void*>(emb), 64, n_pad, 64)) return -1;
    WgArgs w = {};
    int n = 0;
    for (int l = 0; l <= kLayers; l++) {
        const int n_out = l < kLayers ? kHid : out_ch;
        for (int m0 = 0; m0 < n_out; m0 += 128) {
            WgItem& it = w.item[n++];
            it.layer = l; it.m0 = m0; it.rows = std::min(128, n_out - m0);
            it.has_act = l > 0; it.has_emb = (l == 0 || l == kSkipInto);
            it.dw = d_weights[l]; it.ld = lds[l]; it.emb_col0 = emb_col0[l]; it.emb_cols = emb_cols; it.act_col0 = act_col0[l];
            it.db = d_bias + (size_t)l * kHid;
            if (!it.dw) { set_error("rg_mlp_backward: null gradient tensor"); return -1; }
        }
    }
    w.n_items = n;
    w.n_slabs = n_pad / 64;
    w.n_splits = (int)std::max<long long>(1, std::min<long long>(w.n_slabs, (2 * RG_NUM_SMS) / n));
    const size_t smem = (size_t)kWgStages * kWgStageBytes + 16 * 128 + sizeof(WgSmem) + 1024;
    if (ensure_dyn_smem(reinterpret_cast<const void*>(mlp_wgrad_kernel), (int)smem)) return -1;
    {
        LaunchScope ls("mlp_wgrad", stream);
        mlp_wgrad_kernel<<<n * w.n_splits, kWgThreads, smem, stream>>>(w, tm_dz, tm_act, tm_emb);
    }
    return check_launch("rg_mlp_backward", stream, false);
}

}  // extern "C"
