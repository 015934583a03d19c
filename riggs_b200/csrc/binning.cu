// Binning: offsets scan, (tile | depth) key emission, stable radix sort, per-tile ranges.
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + identifyTileRanges and the call to
// cub::DeviceRadixSort::SortPairs of dgr/cuda_rasterizer/rasterizer_impl.cu:70-138, 278-319.
// Pure integer work: results must be bit-identical to the reference's.
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"
#include "kernels.cuh"
#include "math.cuh"

namespace rg {

// ---- level 2 of the offsets scan: exclusive scan of <= a few thousand block sums, one CTA ----------
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(uint32_t* sums, int n, uint32_t* total_out) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < n; base += 1024) {
        int i = base + threadIdx.x;
        uint32_t v = i < n ? sums[i] : 0;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        uint32_t warp_excl = warp == 0 ? 0 : s_warp[warp - 1];
        uint32_t carry = s_carry;
        if (i < n) sums[i] = carry + warp_excl + x - v;  // exclusive
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        sums[n] = s_carry;
        if (total_out) *total_out = s_carry;
    }
}

int launch_scan_block_sums(uint32_t* block_sums, int nblocks, uint32_t* total_out, cudaStream_t stream) {
    LaunchScope ls("scan_block_sums", stream);
    scan_block_sums_kernel<<<1, 1024, 0, stream>>>(block_sums, nblocks, total_out);
    return 0;
}

// ---- key emission ---------------------------------------------------------------------------------
// One CTA per 256 Gaussians (the blocking of preprocess, whose per-block sums were scanned above).  The CTA
// re-scans its 256 tile counts in shared memory, then its instances are emitted by all 256 threads in
// flat order (instance s of the CTA -> owning Gaussian by binary search), so every store is coalesced and
// a Gaussian covering 400 tiles costs its neighbours nothing.  Emission order (Gaussian id ascending,
// tiles row-major inside the rectangle) equals the reference's, which fixes the order of equal keys.
__global__ void __launch_bounds__(256) emit_keys_kernel(int P, int grid_x, int grid_y, const float2* __restrict__ means2D,
                                                        const float* __restrict__ depths, const int* __restrict__ radii,
                                                        const uint32_t* __restrict__ tiles_touched,
                                                        const uint32_t* __restrict__ block_excl,
                                                        uint32_t* __restrict__ point_offsets, uint64_t* __restrict__ keys,
                                                        uint32_t* __restrict__ vals) {
    __shared__ uint32_t s_off[257];   // exclusive offsets inside the CTA, s_off[256] = CTA total
    __shared__ uint32_t s_warp[8];
    __shared__ uint32_t s_depth[256];
    __shared__ ushort4 s_rect[256];  // min.x, min.y, width, unused
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int idx = blockIdx.x * 256 + tid;
    uint32_t cnt = 0;
    if (idx < P) cnt = tiles_touched[idx];
    uint32_t x = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) wbase += (w < warp) ? s_warp[w] : 0;
    const uint32_t incl = wbase + x;
    const uint32_t base = block_excl[blockIdx.x];
    s_off[tid] = incl - cnt;
    if (tid == 255) s_off[256] = incl;
    if (idx < P) point_offsets[idx] = base + incl;
    if (cnt > 0) {
        uint2 rmin, rmax;
        tile_rect(means2D[idx], radii[idx], rmin, rmax, grid_x, grid_y);
        s_rect[tid] = make_ushort4((unsigned short)rmin.x, (unsigned short)rmin.y, (unsigned short)(rmax.x - rmin.x), 0);
        s_depth[tid] = __float_as_uint(depths[idx]);
    }
    __syncthreads();
    const uint32_t total = s_off[256];
    for (uint32_t s = tid; s < total; s += 256) {
        // largest g with s_off[g] <= s  (counts may be zero, so take the last such g)
        int lo = 0, hi = 255;
#pragma unroll
        for (int it = 0; it < 8; it++) {
            int mid = (lo + hi + 1) >> 1;
            if (s_off[mid] <= s) lo = mid; else hi = mid - 1;
        }
        const uint32_t local = s - s_off[lo];
        const ushort4 r = s_rect[lo];
        const uint32_t ty = r.y + local / r.z;
        const uint32_t tx = r.x + local % r.z;
        uint64_t key = (uint64_t)(ty * (uint32_t)grid_x + tx);
        key <<= 32;
        key |= s_depth[lo];
        keys[base + s] = key;
        vals[base + s] = (uint32_t)(blockIdx.x * 256 + lo);
    }
}

int launch_emit_keys(int P, int grid_x, int grid_y, const float2* means2D, const float* depths, const int* radii,
                     const uint32_t* tiles_touched, const uint32_t* block_sums_excl, uint32_t* point_offsets,
                     uint64_t* keys, uint32_t* vals, cudaStream_t stream) {
    if (P == 0) return 0;
    LaunchScope ls("emit_keys", stream);
    emit_keys_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, grid_x, grid_y, means2D, depths, radii, tiles_touched,
                                                          block_sums_excl, point_offsets, keys, vals);
    return 0;
}

// ---- sort -----------------------------------------------------------------------------------------
size_t radix_sort_temp_bytes(size_t R) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                    (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)R, 0, 64);
    return bytes;
}

int launch_sort_pairs(char* temp, size_t temp_bytes, const uint64_t* keys_in, uint64_t* keys_out,
                      const uint32_t* vals_in, uint32_t* vals_out, int64_t R, int end_bit, cudaStream_t stream) {
    if (R <= 0) return 0;
    LaunchScope ls("radix_sort", stream, 0);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, R, 0,
                                                    end_bit, stream);
    if (e != cudaSuccess) {
        set_error(std::string("radix sort: ") + cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

// ---- per-tile ranges ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tile_ranges_kernel(int64_t L, const uint64_t* __restrict__ keys,
                                                          uint2* __restrict__ ranges) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    uint32_t cur = (uint32_t)(keys[idx] >> 32);
    if (idx == 0) ranges[cur].x = 0;
    else {
        uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
        if (cur != prev) {
            ranges[prev].y = (uint32_t)idx;
            ranges[cur].x = (uint32_t)idx;
        }
    }
    if (idx == L - 1) ranges[cur].y = (uint32_t)L;
}

int launch_tile_ranges(int64_t R, const uint64_t* keys, uint2* ranges, int n_tiles, cudaStream_t stream) {
    LaunchScope ls("tile_ranges", stream, R > 0 ? 1 : 0);
    cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)n_tiles, stream);
    if (R > 0) tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, stream>>>(R, keys, ranges);
    return 0;
}

}  // namespace rg
