// Shared helpers for the sm_100a kernels of libriggs_b200.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

#define RG_TILE 16                 // tile edge in pixels (reference: BLOCK_X = BLOCK_Y = 16, dgr/cuda_rasterizer/config.h)
#define RG_TILE_PIX (RG_TILE * RG_TILE)
#define RG_NUM_SMS 148

namespace rg {

void set_error(const std::string& msg);
// Returns 0 when the last launch (and, in debug mode, its execution) succeeded.
int check_launch(const char* what, cudaStream_t stream, bool debug);

// Counts kernel launches and, when profiling is enabled (rg_profile_enable), brackets them with CUDA events on
// the launching stream.
class LaunchScope {
public:
    LaunchScope(const char* name, cudaStream_t stream, int n_kernels = 1);
    ~LaunchScope();
private:
    const char* name_;
    cudaStream_t stream_;
    cudaEvent_t start_;
    bool on_;
};

// Raises a kernel's dynamic shared-memory limit.  The attribute is per (function, device): remembered per pair, so a
// process that drives several GPUs gets it on each of them.  Returns 0, or -1 with the error text set.
int ensure_dyn_smem(const void* func, int bytes);

// A second stream (and three events) per (host thread, device) for work that need not sit in the launching stream's
// chain: fork with cudaEventRecord(fork, main) + cudaStreamWaitEvent(side, fork), join with an event recorded on `side`.
// Under stream capture the fork / join become parallel branches of the graph.  Null when the stream cannot be made.
struct SideStream {
    cudaStream_t stream;
    cudaEvent_t fork, join_a, join_b;
};
SideStream* side_stream();

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
__device__ __forceinline__ size_t align_up_dev(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Carves typed arrays out of one opaque byte buffer, 256-byte aligned.
struct Carver {
    char* base;
    size_t off = 0;
    explicit Carver(char* b) : base(b) {}
    template <typename T>
    T* take(size_t count) {
        off = align_up(off, 256);
        T* p = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        return p;
    }
};

// Per-Gaussian state kept between forward and backward ("geometry buffer"), plus the binning scratch.
struct GeomState {
    float* depths;            // P (+inf for culled Gaussians: they sort last)
    uint8_t* clamped;         // P, bit c set when colour channel c was clamped at 0
    float2* means2D;          // P
    float4* conic_opacity;    // P
    float4* rec;              // 4P: 64-byte splat record read by the blend kernels (cull.cuh: SplatRec layout)
    uint32_t* tiles_touched;  // P
    float* means3D;           // 3P, skinned positions kept by the fused entry point
    uint32_t* ids;            // P iota
    uint32_t* sorted_keys;    // P depth bits, ascending
    uint32_t* sorted_ids;     // P Gaussian ids in depth order (stable)
    uint2* rects;             // P tile rectangles in depth order: {x0 | y0 << 16, width | rows << 16}, tile_count -> tile_scatter
    char* sort_temp;
    size_t sort_temp_bytes;
    uint16_t* M;              // NW x Tpad per (chunk, tile): instances of the earlier chunks of the same group
    uint16_t* Mg;             // NG x Tpad instance counts per (group, tile)
    uint32_t* E;              // NG x T first slot of (group, tile) relative to the tile's first slot
    uint32_t* S;              // nseg x T
    uint32_t* tile_base;      // T
    uint32_t* tile_order;     // T tile indices, longest instance lists first: launch order of the blend kernels
    uint32_t* status;         // [0] num_rendered, [1] overflow flag
    float* grad_acc;          // 10P floats: per-Gaussian 2-D gradient accumulators of the backward blend
    float* d_rot;             // 4P blended bone quaternion * mask (fused entry, kept for the backward)
    float* skin_wsum;         // P: sum of the un-normalised skinning weights (fused entry): spares the backward a sweep over the bones
    float* fk;                // bone tables of the fused entry: G 128x12 | node_rot 128x4 | posed 128x3 | dG 128x12 | dradius 128 | dtrans 4
    static GeomState carve(char* base, size_t P, int W, int H, size_t* bytes);
};

struct ImageState {
    uint32_t* n_contrib;  // H*W
    uint2* ranges;        // tiles
    static ImageState carve(char* base, int W, int H, size_t* bytes);
};

struct BinningState {
    uint32_t* point_list;  // R instance ids: grouped by tile, depth-ordered inside a tile
    uint32_t* surv_mask;   // eight bitmaps over the instance list, one per 8x4 block of a tile (surv_stride words each):
                           // bit set when the forward blend found that the instance can reach that block of its
                           // tile; the backward blend walks by these bits
    size_t surv_mask_bytes, surv_stride;
    static BinningState carve(char* base, size_t R, size_t* bytes);
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum each of v[0..15] over the 32 lanes with 16 shuffles; the lane holds the total of v[(lane >> 1) & 15].
__device__ __forceinline__ float transpose_reduce16(const float* v, int lane) {
    const unsigned full = 0xffffffffu;
    const bool u4 = lane & 16, u3 = lane & 8, u2 = lane & 4, u1 = lane & 2;
    float w[8], x[4], y[2];
#pragma unroll
    for (int i = 0; i < 8; i++) w[i] = (u4 ? v[8 + i] : v[i]) + __shfl_xor_sync(full, u4 ? v[i] : v[8 + i], 16);
#pragma unroll
    for (int i = 0; i < 4; i++) x[i] = (u3 ? w[4 + i] : w[i]) + __shfl_xor_sync(full, u3 ? w[i] : w[4 + i], 8);
#pragma unroll
    for (int i = 0; i < 2; i++) y[i] = (u2 ? x[2 + i] : x[i]) + __shfl_xor_sync(full, u2 ? x[i] : x[2 + i], 4);
    float s = (u1 ? y[1] : y[0]) + __shfl_xor_sync(full, u1 ? y[0] : y[1], 2);
    s += __shfl_xor_sync(full, s, 1);
    return s;  // value id = 8*u4 + 4*u3 + 2*u2 + u1
}

}  // namespace rg
