// Gradient all-reduce over NVLink / NVSwitch without NCCL: one kernel per rank on a buffer that the ranks have mapped
// into one multicast address range (NVLS "multimem" addressing; the mapping itself is made by the host, e.g. with
// torch.distributed._symmetric_memory).  Rank r owns the r-th slice of the buffer:
//   multimem.ld_reduce  one load returns the SUM over all ranks' copies of a 16-byte word (the switch reduces),
//   multimem.st         one store writes the result into every rank's copy.
// So every byte crosses each GPU's link once in and once out, and there is no intermediate buffer or second pass.
// The caller brackets the kernel with two cross-rank barriers on the stream (all ranks have finished writing their
// gradients / all slices have been written back).  The reference has no distributed code; this replaces the
// ncclAllReduce a data-parallel trainer would issue after the backward.
#include <cstdlib>

#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

__device__ __forceinline__ float4 mm_ld_reduce_add(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void mm_st(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

// words [w0, w1) of 16 bytes belong to this rank
__global__ void __launch_bounds__(512) allreduce_multimem_kernel(float* __restrict__ mc, size_t w0, size_t w1) {
    constexpr int kUnroll = 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = w0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (kUnroll - 1) * stride < w1; i += kUnroll * stride) {
        float4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; u++) v[u] = mm_ld_reduce_add(mc + 4 * (i + u * stride));
#pragma unroll
        for (int u = 0; u < kUnroll; u++) mm_st(mc + 4 * (i + u * stride), v[u]);
    }
    for (; i < w1; i += stride) mm_st(mc + 4 * i, mm_ld_reduce_add(mc + 4 * i));
}

// Same slice ownership over plain peer-to-peer mappings (every rank's buffer mapped into this process): rank r loads
// word i of its slice from all ranks, adds, and stores the sum into all ranks' copies.
struct PeerPtrs {
    float* p[8];
};
template <int WORLD, int kUnroll = (WORLD <= 2 ? 4 : 2)>
__global__ void __launch_bounds__(512) allreduce_p2p_kernel(PeerPtrs bufs, size_t w0, size_t w1) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = w0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (kUnroll - 1) * stride < w1; i += kUnroll * stride) {
        float4 v[kUnroll][WORLD];
#pragma unroll
        for (int u = 0; u < kUnroll; u++)
#pragma unroll
            for (int r = 0; r < WORLD; r++) v[u][r] = __ldcv(reinterpret_cast<const float4*>(bufs.p[r]) + i + u * stride);
#pragma unroll
        for (int u = 0; u < kUnroll; u++) {
            float4 s = v[u][0];
#pragma unroll
            for (int r = 1; r < WORLD; r++) { s.x += v[u][r].x; s.y += v[u][r].y; s.z += v[u][r].z; s.w += v[u][r].w; }
#pragma unroll
            for (int r = 0; r < WORLD; r++) reinterpret_cast<float4*>(bufs.p[r])[i + u * stride] = s;
        }
    }
    for (; i < w1; i += stride) {
        float4 s = __ldcv(reinterpret_cast<const float4*>(bufs.p[0]) + i);
#pragma unroll
        for (int r = 1; r < WORLD; r++) {
            const float4 v = __ldcv(reinterpret_cast<const float4*>(bufs.p[r]) + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
#pragma unroll
        for (int r = 0; r < WORLD; r++) reinterpret_cast<float4*>(bufs.p[r])[i] = s;
    }
}

// ---- the same two kernels over several word ranges of the buffer (one launch): used when the all-reduce of a chunk of
// gradient rows (one range per parameter tensor) runs beside the backward of the next chunk
constexpr int kMaxRanges = 8;
struct WordRanges {
    int n;
    size_t w0[kMaxRanges], w1[kMaxRanges];  // this rank's slice of each range, in 16-byte words
};

template <int WORLD>
__global__ void __launch_bounds__(512) allreduce_p2p_ranges_kernel(PeerPtrs bufs, WordRanges rr) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int q = 0; q < rr.n; q++) {
        for (size_t i = rr.w0[q] + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rr.w1[q]; i += stride) {
            float4 v[WORLD];
#pragma unroll
            for (int r = 0; r < WORLD; r++) v[r] = __ldcv(reinterpret_cast<const float4*>(bufs.p[r]) + i);
            float4 s = v[0];
#pragma unroll
            for (int r = 1; r < WORLD; r++) { s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w; }
#pragma unroll
            for (int r = 0; r < WORLD; r++) reinterpret_cast<float4*>(bufs.p[r])[i] = s;
        }
    }
}

__global__ void __launch_bounds__(512) allreduce_multimem_ranges_kernel(float* __restrict__ mc, WordRanges rr) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int q = 0; q < rr.n; q++)
        for (size_t i = rr.w0[q] + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rr.w1[q]; i += stride)
            mm_st(mc + 4 * i, mm_ld_reduce_add(mc + 4 * i));
}

// false: bad ranges
static bool slice_ranges(int n_ranges, const size_t* offsets, const size_t* lengths, int rank, int world, WordRanges& rr) {
    if (n_ranges < 1 || n_ranges > kMaxRanges) return false;
    rr.n = n_ranges;
    for (int q = 0; q < n_ranges; q++) {
        if ((offsets[q] & 3) || (lengths[q] & 3)) return false;
        const size_t words = lengths[q] / 4, base = offsets[q] / 4;
        const size_t per = (words + world - 1) / world;
        const size_t a = per * (size_t)rank < words ? per * (size_t)rank : words;
        const size_t b = a + per < words ? a + per : words;
        rr.w0[q] = base + a;
        rr.w1[q] = base + b;
    }
    return true;
}

}  // namespace rg

using namespace rg;

// CTAs of the whole-buffer kernels: two per SM unless RG_AR_BLOCKS says otherwise (tuning aid)
static int allreduce_blocks() {
    static const int v = [] {
        const char* e = std::getenv("RG_AR_BLOCKS");
        const int n = e ? std::atoi(e) : 0;
        return n > 0 ? n : RG_NUM_SMS * 2;
    }();
    return v;
}

extern "C" int rg_allreduce_ranges(float* const* peer_bufs, float* multicast_base, int n_ranges, const size_t* offsets,
                                   const size_t* lengths, int rank, int world, int blocks, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    WordRanges rr;
    if (rank < 0 || rank >= world || !slice_ranges(n_ranges, offsets, lengths, rank, world, rr)) {
        set_error("rg_allreduce_ranges: 1..8 ranges, offsets and lengths multiples of 4 floats, rank inside world");
        return -1;
    }
    if (blocks <= 0) blocks = RG_NUM_SMS;
    if (multicast_base) {
        LaunchScope ls("allreduce_multimem", stream);
        allreduce_multimem_ranges_kernel<<<blocks, 512, 0, stream>>>(multicast_base, rr);
        return cudaGetLastError() == cudaSuccess ? 0 : -1;
    }
    if (!peer_bufs || !(world == 2 || world == 4 || world == 8)) {
        set_error("rg_allreduce_ranges: peer buffers need a world of 2, 4 or 8");
        return -1;
    }
    PeerPtrs pp;
    for (int r = 0; r < world; r++) {
        if (!peer_bufs[r] || (reinterpret_cast<uintptr_t>(peer_bufs[r]) & 15)) { set_error("rg_allreduce_ranges: null or misaligned peer buffer"); return -1; }
        pp.p[r] = peer_bufs[r];
    }
    LaunchScope ls("allreduce_p2p", stream);
    if (world == 2) allreduce_p2p_ranges_kernel<2><<<blocks, 512, 0, stream>>>(pp, rr);
    else if (world == 4) allreduce_p2p_ranges_kernel<4><<<blocks, 512, 0, stream>>>(pp, rr);
    else allreduce_p2p_ranges_kernel<8><<<blocks, 512, 0, stream>>>(pp, rr);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}


extern "C" int rg_allreduce_p2p(float* const* peer_bufs, size_t n_floats, int rank, int world, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!peer_bufs || rank < 0 || rank >= world || !(world == 2 || world == 4 || world == 8)) {
        set_error("rg_allreduce_p2p: world must be 2, 4 or 8 and rank inside it");
        return -1;
    }
    if (n_floats & 3) { set_error("rg_allreduce_p2p: the buffer must be a multiple of 4 floats long"); return -1; }
    PeerPtrs pp;
    // the sum starts with rank 0's copy on every rank: same association everywhere, bit-identical results
    for (int r = 0; r < world; r++) {
        if (!peer_bufs[r] || (reinterpret_cast<uintptr_t>(peer_bufs[r]) & 15)) { set_error("rg_allreduce_p2p: null or misaligned peer buffer"); return -1; }
        pp.p[r] = peer_bufs[r];
    }
    const size_t words = n_floats / 4;
    const size_t per = (words + world - 1) / world;
    const size_t w0 = per * (size_t)rank < words ? per * (size_t)rank : words;
    const size_t w1 = w0 + per < words ? w0 + per : words;
    if (w1 == w0) return 0;
    LaunchScope ls("allreduce_p2p", stream);
    const int nb = allreduce_blocks();
    static const int unroll = [] { const char* e = std::getenv("RG_AR_UNROLL"); return e ? std::atoi(e) : 0; }();  // tuning aid
    if (world == 2 && unroll == 8) allreduce_p2p_kernel<2, 8><<<nb, 512, 0, stream>>>(pp, w0, w1);
    else if (world == 2 && unroll == 2) allreduce_p2p_kernel<2, 2><<<nb, 512, 0, stream>>>(pp, w0, w1);
    else if (world == 2) allreduce_p2p_kernel<2><<<nb, 512, 0, stream>>>(pp, w0, w1);
    else if (world == 4 && unroll == 4) allreduce_p2p_kernel<4, 4><<<nb, 512, 0, stream>>>(pp, w0, w1);
    else if (world == 4) allreduce_p2p_kernel<4><<<nb, 512, 0, stream>>>(pp, w0, w1);
    else allreduce_p2p_kernel<8><<<nb, 512, 0, stream>>>(pp, w0, w1);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}


extern "C" int rg_allreduce_multimem(float* multicast_base, size_t n_floats, int rank, int world, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!multicast_base || world < 1 || rank < 0 || rank >= world) { set_error("rg_allreduce_multimem: bad argument"); return -1; }
    if ((reinterpret_cast<uintptr_t>(multicast_base) & 15) || (n_floats & 3)) {
        set_error("rg_allreduce_multimem: the buffer must be 16-byte aligned and a multiple of 4 floats long");
        return -1;
    }
    const size_t words = n_floats / 4;
    const size_t per = (words + world - 1) / world;
    const size_t w0 = per * (size_t)rank < words ? per * (size_t)rank : words;
    const size_t w1 = w0 + per < words ? w0 + per : words;
    if (w1 == w0) return 0;
    LaunchScope ls("allreduce_multimem", stream);
    allreduce_multimem_kernel<<<allreduce_blocks(), 512, 0, stream>>>(multicast_base, w0, w1);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
