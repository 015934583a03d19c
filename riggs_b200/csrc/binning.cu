// Binning: from per-Gaussian (depth, tile rectangle) to the per-tile, depth-ordered instance lists.
//
// The reference materialises one 64-bit key (tile << 32 | depth bits) + one 32-bit id per (Gaussian, tile)
// instance and runs a 44..45-bit radix sort over all R instances (duplicateWithKeys + cub SortPairs +
// identifyTileRanges, dgr/cuda_rasterizer/rasterizer_impl.cu:70-138, 278-319): ~12 B written and
// 6 x 24 B moved per instance.  The result it defines -- instances grouped by tile, ordered by depth bits
// inside a tile, ties in Gaussian-id order -- is produced here without ever building the keys:
//
//   1. stable radix sort of the P Gaussians by depth bits (32-bit keys, P items: small, L2 resident);
//   2. tile_count:   the depth-ordered Gaussians are cut into NW contiguous chunks, one warp each; a warp
//                    counts the instances of its chunk per tile in shared memory (one lane per Gaussian, shared
//                    atomics) -> row of M[NW][T], and keeps the tile rectangles in depth order;
//   3. column scans: per tile, prefix over the chunks (two levels) and prefix over the tiles ->
//                    E[w][t] = first output slot of chunk w in tile t, ranges[t], num_rendered;
//   4. tile_scatter: every warp walks its chunk again, one Gaussian after the other, and drops each instance id
//                    straight into its final slot (cursor per tile in shared memory; the tiles of one Gaussian are
//                    distinct, so its lanes need no ranking among themselves).
//
// The instance list is written exactly once (4 B per instance) and nothing instance-sized is ever read
// back; num_rendered stays on the device (the sync-free entry points never look at it on the host).
// Order proof: slots of tile t are handed out by (chunk index, position inside the chunk's walk), chunks
// are contiguous ranges of the depth-sorted order, and the depth sort is stable over ascending ids -- the
// same total order as the reference's stable sort on (tile, depth) over id-ordered emission.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "kernels.cuh"
#include "math.cuh"

namespace rg {

// ---- 1. depth sort: csrc/radix_sort.cu; the library sort below is kept as the A/B partner (RG_DEPTH_SORT=cub) ----
size_t depth_sort_temp_bytes_cub(size_t P) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (int)P, 0, 32);
    return bytes;
}

int launch_depth_sort_cub(char* temp, size_t temp_bytes, const float* depths, uint32_t* keys_out, const uint32_t* ids_in,
                          uint32_t* ids_out, int P, cudaStream_t stream) {
    if (P <= 0) return 0;
    LaunchScope ls("depth_sort", stream, 0);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, reinterpret_cast<const uint32_t*>(depths), keys_out,
                                                    ids_in, ids_out, P, 0, 32, stream);
    if (e != cudaSuccess) {
        set_error(std::string("depth sort: ") + cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

// ---- 2. per-chunk, per-tile counts ------------------------------------------------------------------
// Counting needs no order inside a chunk: every lane takes one Gaussian of the chunk and adds one to the counter of
// each tile of its rectangle (two 16-bit counters per shared-memory word, a chunk holds at most 65535 Gaussians so a
// half never carries into its neighbour).  Rectangles of more than kWideRect tiles are left to the whole warp.
// blockIdx.y = band: the CTA counts only the tile rows [band * band_rows, (band + 1) * band_rows) and its counters
// span those rows only (BinPlan).
// split_n > 1: only the tile rows y with y % split_n == split_i are binned (one view rendered by several GPUs, each
// owning every split_n-th tile row); a rectangle then keeps its owned rows, split_n rows apart.
constexpr uint32_t kWideRect = 48;

// Rows of [y0, y1) that belong to this part of a split view: first owned row and the number of owned rows.
__device__ __forceinline__ void owned_rows(uint32_t y0, uint32_t y1, uint32_t split_i, uint32_t split_n, uint32_t& first,
                                           uint32_t& rows) {
    if (split_n <= 1) { first = y0; rows = y1 > y0 ? y1 - y0 : 0u; return; }
    first = y0 + (split_i + split_n - y0 % split_n) % split_n;
    rows = first < y1 ? (y1 - first + split_n - 1) / split_n : 0u;
}

__global__ void __launch_bounds__(256) tile_count_kernel(int P, int T, int NW, int L, int grid_x, int grid_y, int band_rows,
                                                         int split_i, int split_n,
                                                         const uint32_t* __restrict__ sorted_ids,
                                                         const uint32_t* __restrict__ tiles_touched,
                                                         const float2* __restrict__ means2D, const int* __restrict__ radii,
                                                         uint2* __restrict__ rects, uint16_t* __restrict__ M,
                                                         uint16_t* __restrict__ Mg) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
    const int gw = blockIdx.x * wpc + warp;
    const int Tpad = (T + 1) & ~1;
    const uint32_t by0 = blockIdx.y * (uint32_t)band_rows, by1 = min((uint32_t)grid_y, by0 + (uint32_t)band_rows);
    const uint32_t t_lo = by0 * (uint32_t)grid_x;              // even: band_rows is even
    const int Tb = (int)((by1 - by0) * (uint32_t)grid_x), Tbpad = (Tb + 1) & ~1;
    uint16_t* hist = reinterpret_cast<uint16_t*>(s_raw) + (size_t)warp * Tbpad;
    uint32_t* h32 = reinterpret_cast<uint32_t*>(hist);
    for (int k = lane; k < Tbpad / 2; k += 32) h32[k] = 0;
    __syncwarp();
    const int lo = min(P, gw * L), hi = gw < NW ? min(P, lo + L) : lo;  // warps past the last chunk count nothing
    for (int base = lo; base < hi; base += 32) {
        const int i = base + lane;
        uint32_t cnt = 0, x0 = 0, w = 0, y0 = 0, rows = 0;
        if (i < hi) {
            const uint32_t id = sorted_ids[i];
            uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
            if (tiles_touched[id] > 0) tile_rect(means2D[id], radii[id], rmin, rmax, grid_x, grid_y);
            x0 = rmin.x; w = rmax.x - rmin.x;
            if (blockIdx.y == 0) {
                // kept for the scatter pass, which then reads one packed word pair per Gaussian in depth order
                uint32_t fy, fr;
                owned_rows(rmin.y, rmax.y, (uint32_t)split_i, (uint32_t)split_n, fy, fr);
                rects[i] = fr && w ? make_uint2(x0 | (fy << 16), w | (fr << 16)) : make_uint2(0, 0);
            }
            owned_rows(max(rmin.y, by0), min(rmax.y, by1), (uint32_t)split_i, (uint32_t)split_n, y0, rows);
            cnt = rows * w;
        }
        const uint32_t step = (uint32_t)split_n * (uint32_t)grid_x;
        if (cnt > 0 && cnt <= kWideRect) {
            uint32_t row = y0 * (uint32_t)grid_x - t_lo + x0;
            for (uint32_t r = 0; r < rows; r++, row += step) {
                for (uint32_t t = row; t < row + w; t++) atomicAdd(&h32[t >> 1], 1u << ((t & 1u) * 16u));
            }
        }
        unsigned wide = __ballot_sync(0xffffffffu, cnt > kWideRect);
        while (wide) {
            const int o = __ffs(wide) - 1;
            wide &= wide - 1;
            const uint32_t bx0 = __shfl_sync(0xffffffffu, x0, o), yy0 = __shfl_sync(0xffffffffu, y0, o);
            const uint32_t bw = __shfl_sync(0xffffffffu, w, o), n = __shfl_sync(0xffffffffu, cnt, o);
            const float rw = __frcp_rn((float)bw);  // s / w through a reciprocal, exact here (see tile_scatter_kernel)
            const uint32_t t0 = yy0 * (uint32_t)grid_x - t_lo + bx0;
            for (uint32_t s = lane; s < n; s += 32) {
                const uint32_t q = __float2uint_rz(((float)s + 0.5f) * rw);
                const uint32_t t = t0 + q * step + (s - q * bw);
                atomicAdd(&h32[t >> 1], 1u << ((t & 1u) * 16u));
            }
        }
    }
    // The warps of a CTA hold consecutive chunks (a group): per tile, turn their counts into offsets inside the
    // group (what the scatter pass starts its cursors from) and keep one total per (group, tile) -- the column
    // scans then run over the groups, eight times fewer rows than chunks.
    __syncthreads();
    const uint16_t* all = reinterpret_cast<const uint16_t*>(s_raw);
    uint16_t* allw = reinterpret_cast<uint16_t*>(s_raw);
    for (int t = threadIdx.x; t < Tbpad; t += blockDim.x) {
        uint32_t run = 0;
        for (int w = 0; w < wpc; w++) {
            const uint32_t c = all[(size_t)w * Tbpad + t];
            allw[(size_t)w * Tbpad + t] = (uint16_t)run;
            run += c;
        }
        Mg[(size_t)blockIdx.x * Tpad + t_lo + t] = (uint16_t)run;
    }
    __syncthreads();
    if (gw >= NW) return;
    uint32_t* row = reinterpret_cast<uint32_t*>(M + (size_t)gw * Tpad + t_lo);
    for (int k = lane; k < Tbpad / 2; k += 32) row[k] = h32[k];
}

// ---- 3a. sums over segments of 32 group rows -----------------------------------------------------------
constexpr int kSegRows = 32;
__global__ void __launch_bounds__(256) seg_sum_kernel(int T, int Tpad, int NW, const uint16_t* __restrict__ M,  // rows = groups
                                                      uint32_t* __restrict__ S, uint32_t* __restrict__ tile_total) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int seg = blockIdx.y;
    if (t >= T) return;
    uint32_t s = 0;
    const int w0 = seg * kSegRows, w1 = min(NW, w0 + kSegRows);
#pragma unroll 8
    for (int w = w0; w < w1; w++) s += M[(size_t)w * Tpad + t];
    S[(size_t)seg * T + t] = s;
    if (s) atomicAdd(tile_total + t, s);  // zeroed by the launcher; becomes tile_base in the next kernel
}

constexpr int kOrderBuckets = 128;
__device__ __forceinline__ int order_bucket(uint32_t n) {
    if (n < 4) return (int)n;
    const int e = 31 - __clz(n);
    return e * 4 + (int)((n >> (e - 2)) & 3u);
}

// ---- 3b. per-tile totals, prefix over tiles, ranges, launch order, num_rendered -----
template <int NT>
__device__ __forceinline__ void tile_base_body(int T, uint32_t* __restrict__ tile_base, uint2* __restrict__ ranges,
                                               uint32_t* __restrict__ tile_order, uint32_t* __restrict__ status,
                                               uint32_t capacity) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_bucket[kOrderBuckets];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < T; base += NT) {
        const int t = base + threadIdx.x;
        const uint32_t tot = t < T ? tile_base[t] : 0u;  // per-tile totals accumulated by seg_sum_kernel
        uint32_t x = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = lane < NT / 32 ? s_warp[lane] : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = s_carry;
        const uint32_t excl = carry + (warp ? s_warp[warp - 1] : 0) + x - tot;
        if (t < T) {
            tile_base[t] = excl;
            // clamped to the caller's capacity so an overflowing frame (flagged in status[1]) stays memory-safe
            ranges[t] = tot ? make_uint2(min(excl, capacity), min(excl + tot, capacity)) : make_uint2(0, 0);
        }
        __syncthreads();
        if (threadIdx.x == NT - 1) s_carry = carry + s_warp[NT / 32 - 1];
        __syncthreads();
    }
    // Launch order of the blend kernels: tiles with the longest lists first (counting sort on a 2-bit-mantissa log
    // scale, descending), so that the frame's few heavy tiles start at once instead of wherever the raster order puts
    // them, and the short ones fill in behind.
    for (int b = threadIdx.x; b < kOrderBuckets; b += NT) s_bucket[b] = 0;
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += NT) {
        const uint2 r = ranges[t];
        atomicAdd(&s_bucket[order_bucket(r.y - r.x)], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 32) {  // exclusive prefix over the buckets in DESCENDING order, four buckets per lane
        constexpr int kPer = kOrderBuckets / 32;
        uint32_t c[kPer], tot = 0;
#pragma unroll
        for (int i = 0; i < kPer; i++) { c[i] = s_bucket[kOrderBuckets - 1 - (lane * kPer + i)]; tot += c[i]; }
        uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        uint32_t run = incl - tot;
#pragma unroll
        for (int i = 0; i < kPer; i++) { s_bucket[kOrderBuckets - 1 - (lane * kPer + i)] = run; run += c[i]; }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += NT) {
        const uint2 r = ranges[t];
        tile_order[atomicAdd(&s_bucket[order_bucket(r.y - r.x)], 1u)] = (uint32_t)t;
    }
    if (threadIdx.x == 0) {
        status[0] = s_carry;                       // num_rendered
        status[1] = s_carry > capacity ? 1u : 0u;  // instance list would overflow the caller's buffer
    }
}

__global__ void __launch_bounds__(1024) tile_base_kernel(int T, uint32_t* __restrict__ tile_base,
                                                         uint2* __restrict__ ranges, uint32_t* __restrict__ tile_order,
                                                         uint32_t* __restrict__ status, uint32_t capacity) {
    tile_base_body<1024>(T, tile_base, ranges, tile_order, status, capacity);
}

// ---- 3c. first slot of every (group, tile) -----------------------------------------------------------
// E holds offsets RELATIVE to the tile's first slot, so this kernel does not wait for the prefix over the tiles:
// with base_in_grid the block row y == nseg is one extra CTA that computes that prefix (tile_base, ranges, launch
// order, num_rendered) next to the column scans instead of in a launch of its own before them.
__global__ void __launch_bounds__(1024) col_scan_kernel(int T, int Tpad, int NW, int nseg, const uint16_t* __restrict__ M,
                                                       const uint32_t* __restrict__ S, uint32_t* __restrict__ E,
                                                       uint32_t* __restrict__ tile_base, uint2* __restrict__ ranges,
                                                       uint32_t* __restrict__ tile_order, uint32_t* __restrict__ status,
                                                       uint32_t capacity) {
    const int seg = blockIdx.y;
    if (seg == nseg) {  // only launched when the prefix over the tiles is still pending
        if (blockIdx.x == 0) tile_base_body<1024>(T, tile_base, ranges, tile_order, status, capacity);
        return;
    }
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    uint32_t run = 0;
#pragma unroll 8
    for (int s = 0; s < seg; s++) run += S[(size_t)s * T + t];  // prefix over the earlier segments
    const int w0 = seg * kSegRows, w1 = min(NW, w0 + kSegRows);
    for (int w = w0; w < w1; w++) {
        E[(size_t)w * T + t] = run;
        run += M[(size_t)w * Tpad + t];
    }
}

// ---- 4. scatter ---------------------------------------------------------------------------------------
// A warp hands out the slots of its chunk one Gaussian at a time, in depth order: the tiles of ONE Gaussian are
// distinct, so its lanes read, use and bump the per-tile cursors without any ranking among themselves, and the
// order between Gaussians is the order of the loop.  (The earlier walk of 32 instances of mixed Gaussians per step
// needed a MATCH.ANY per step to rank equal tiles: 36 % of that kernel's stall samples.)
// blockIdx.y = band, as in tile_count_kernel: the CTA hands out only the slots of its band's tiles.
__global__ void __launch_bounds__(256) tile_scatter_kernel(int P, int T, int NW, int L, int grid_x, int grid_y, int band_rows,
                                                           int split_n, const uint32_t* __restrict__ sorted_ids,
                                                           const uint2* __restrict__ rects, const uint16_t* __restrict__ M,
                                                           const uint32_t* __restrict__ E,
                                                           const uint32_t* __restrict__ tile_base,
                                                           uint32_t* __restrict__ point_list, uint32_t capacity) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
    const int gw = blockIdx.x * wpc + warp;
    const int Tpad = (T + 1) & ~1;
    const uint32_t by0 = blockIdx.y * (uint32_t)band_rows, by1 = min((uint32_t)grid_y, by0 + (uint32_t)band_rows);
    const uint32_t t_lo = by0 * (uint32_t)grid_x;
    const int Tb = (int)((by1 - by0) * (uint32_t)grid_x), Tbpad = (Tb + 1) & ~1;
    // slot of an instance = first slot of (group, tile) [32-bit, shared by the CTA's warps] + the warp's running
    // 16-bit cursor, which starts at the chunk's offset inside the group (tile_count's M rows)
    uint32_t* base = reinterpret_cast<uint32_t*>(s_raw);
    uint16_t* cur = reinterpret_cast<uint16_t*>(base + Tbpad) + (size_t)warp * Tbpad;
    const uint32_t* grow = E + (size_t)blockIdx.x * T + t_lo;
    for (int k = threadIdx.x; k < Tb; k += blockDim.x) base[k] = grow[k] + tile_base[t_lo + k];
    if (gw < NW) {
        const uint32_t* row = reinterpret_cast<const uint32_t*>(M + (size_t)gw * Tpad + t_lo);
        uint32_t* c32 = reinterpret_cast<uint32_t*>(cur);
        for (int k = lane; k < Tbpad / 2; k += 32) c32[k] = row[k];
    }
    __syncthreads();
    if (gw >= NW) return;
    const int lo = min(P, gw * L), hi = min(P, lo + L);
    if (lo >= hi) return;
    const unsigned full = 0xffffffffu;
    const uint32_t pitch = (uint32_t)grid_x * (uint32_t)(split_n > 1 ? split_n : 1);  // distance between two rows of a rectangle
    const uint32_t off_a = (uint32_t)(lane >> 3) * pitch + (uint32_t)(lane & 7);  // lane's tile in an 8 x 4 layout
    const uint32_t off_b = (uint32_t)(lane >> 2) * pitch + (uint32_t)(lane & 3);  // ... in a 4 x 8 layout
    const bool banded = gridDim.y > 1;
    // 32 Gaussians per block: one coalesced load of their rectangles and ids (the next block's is in flight while
    // this one is handed out), then lane j's values are broadcast in turn
    uint2 rc_next = lo + lane < hi ? rects[lo + lane] : make_uint2(0, 0);
    uint32_t id_next = lo + lane < hi ? sorted_ids[lo + lane] : 0u;
    for (int b0 = lo; b0 < hi; b0 += 32) {
        uint2 rc_mine = rc_next;
        const uint32_t id_mine = id_next;
        const int nx = b0 + 32 + lane;
        rc_next = nx < hi ? rects[nx] : make_uint2(0, 0);
        id_next = nx < hi ? sorted_ids[nx] : 0u;
        if (banded && rc_mine.y != 0) {  // keep the rectangle's rows that lie in this band (rows are split_n apart)
            const uint32_t sn = (uint32_t)(split_n > 1 ? split_n : 1);
            uint32_t y0 = rc_mine.x >> 16, h = rc_mine.y >> 16;
            const uint32_t skip = y0 < by0 ? (by0 - y0 + sn - 1) / sn : 0u;
            y0 += skip * sn;
            h = h > skip ? h - skip : 0u;
            if (h) h = y0 < by1 ? min(h, (by1 - y0 + sn - 1) / sn) : 0u;
            rc_mine = h ? make_uint2((rc_mine.x & 0xffffu) | (y0 << 16), (rc_mine.y & 0xffffu) | (h << 16)) : make_uint2(0, 0);
        }
        unsigned live = __ballot_sync(full, rc_mine.y != 0);  // culled Gaussians (and those outside the band) have no tiles
        while (live) {
            const int j = __ffs(live) - 1;
            live &= live - 1;
            const uint32_t xy = __shfl_sync(full, rc_mine.x, j), wh = __shfl_sync(full, rc_mine.y, j);
            const uint32_t id = __shfl_sync(full, id_mine, j);
            const uint32_t w = wh & 0xffffu, h = wh >> 16;
            const uint32_t t0 = (xy >> 16) * (uint32_t)grid_x + (xy & 0xffffu) - t_lo;  // band-local tile index
            if (w <= 8 && h <= 4) {  // the usual case: the rectangle fits the warp laid out 8 x 4 ...
                if ((uint32_t)(lane & 7) < w && (uint32_t)(lane >> 3) < h) {
                    const uint32_t t = t0 + off_a;
                    const uint32_t k = cur[t];
                    const uint32_t c = base[t] + k;
                    if (c < capacity) point_list[c] = id;
                    cur[t] = (uint16_t)(k + 1);
                }
            } else if (w <= 4 && h <= 8) {  // ... or 4 x 8
                if ((uint32_t)(lane & 3) < w && (uint32_t)(lane >> 2) < h) {
                    const uint32_t t = t0 + off_b;
                    const uint32_t k = cur[t];
                    const uint32_t c = base[t] + k;
                    if (c < capacity) point_list[c] = id;
                    cur[t] = (uint16_t)(k + 1);
                }
            } else {
                // s / w through a reciprocal: (s + 0.5) / w stays 0.5 / w away from every integer, and three roundings of
                // fp32 move it by less than that for s < 2^21 (a rectangle has at most 65535 tiles)
                const uint32_t n = w * h;
                const float rw = __frcp_rn((float)w);
                for (uint32_t s = lane; s < n; s += 32) {
                    const uint32_t q = __float2uint_rz(((float)s + 0.5f) * rw);
                    const uint32_t t = t0 + q * pitch + (s - q * w);
                    const uint32_t k = cur[t];
                    const uint32_t c = base[t] + k;
                    if (c < capacity) point_list[c] = id;
                    cur[t] = (uint16_t)(k + 1);
                }
            }
            __syncwarp();  // the next Gaussian may touch the same tiles from other lanes
        }
    }
}

// ---- launch plumbing ----------------------------------------------------------------------------------
BinPlan make_bin_plan(int P, int grid_x, int grid_y) {
    BinPlan b;
    const int T = grid_x * grid_y;
    b.T = T;
    b.Tpad = (T + 1) & ~1;
    static const int envL = [] { const char* e = std::getenv("RG_BIN_L"); return e ? atoi(e) : 0; }();
    static const int envRows = [] { const char* e = std::getenv("RG_BIN_MAXROWS"); return e ? atoi(e) : 0; }();
    static const int envBand = [] { const char* e = std::getenv("RG_BAND_TILES"); return e ? atoi(e) : 0; }();  // tuning aid
    int L = envL > 0 ? envL : 128;
    int NW = (P + L - 1) / L;
    const int kMaxRows = envRows > 0 ? envRows : 8192;
    if (NW > kMaxRows) {
        NW = kMaxRows;
        L = (P + NW - 1) / NW;
        NW = (P + L - 1) / L;
    }
    if (NW < 1) NW = 1;
    b.L = L;
    b.NW = NW;
    // Bands: the per-warp 16-bit counters / cursors of the two walking kernels cover one band of tile rows.  Up to
    // 4400 tiles (800 x 800: 2500, 1024 x 1024: 4096) one band is the whole image; a 1920 x 1080 grid (8160 tiles) is
    // counted in two bands and scattered in four, which brings the scatter CTA from 163 KB of shared memory (one CTA =
    // 8 warps per SM) down to 43 KB.  Every band walks all the chunks and skips the Gaussians that do not reach its rows
    // (measured at C4, 1 M Gaussians: count 0.182 -> 0.111 ms with two bands, 0.142 with four; scatter 0.377 -> 0.296 ms
    // with four; at 4096 tiles bands cost more than they give).  band_rows is even so that a band's first tile index is
    // (the 16-bit rows are moved as 32-bit words).
    auto bands = [&](int max_tiles, int& nb_out, int& rows_out) {
        int nb = T <= 4400 ? 1 : (T + max_tiles - 1) / max_tiles;
        if (envBand > 0) nb = (T + envBand - 1) / envBand;
        if (nb < 1) nb = 1;
        int rows = (grid_y + nb - 1) / nb;
        rows = (rows + 1) & ~1;
        if (rows < 2) rows = 2;
        rows_out = rows;
        nb_out = (grid_y + rows - 1) / rows;
        if (nb_out < 1) nb_out = 1;
    };
    bands(4400, b.nb_count, b.band_rows_count);
    bands(2560, b.nb, b.band_rows);
    const int Tb = ((b.band_rows < grid_y ? b.band_rows : grid_y) * grid_x + 1) & ~1;
    const int Tbc = ((b.band_rows_count < grid_y ? b.band_rows_count : grid_y) * grid_x + 1) & ~1;
    // a group = the chunks of one CTA of the two walking kernels; its size is bounded by their shared memory
    // (scatter: 4 bytes per tile for the group + 2 per tile and warp; count: 2 per tile and warp)
    const size_t smem_budget = 200 * 1024;
    int wpc = std::min(((int)smem_budget - 4 * Tb) / (2 * Tb), (int)smem_budget / (2 * Tbc));
    b.wpc = wpc > 8 ? 8 : wpc;
    if (b.wpc < 1) b.wpc = 0;
    b.NG = b.wpc ? (NW + b.wpc - 1) / b.wpc : 0;
    b.nseg = (b.NG + kSegRows - 1) / kSegRows;
    return b;
}

int launch_tile_count(const BinPlan& b, int P, int grid_x, int grid_y, const uint32_t* sorted_ids,
                      const uint32_t* tiles_touched, const float2* means2D, const int* radii, uint2* rects, uint16_t* M,
                      uint16_t* Mg, cudaStream_t stream, int split_i, int split_n) {
    if (split_n < 1) split_n = 1;
    if (b.wpc < 1 || (size_t)b.L * b.wpc > 65535 || b.T > 65535 || grid_x > 65535 || grid_y > 65535) {
        set_error("binning: image has too many tiles (or chunk too long) for the shared-memory tile counters");
        return -1;
    }
    const size_t Tb = (size_t)((std::min(b.band_rows_count, grid_y) * grid_x + 1) & ~1);
    const size_t smem = (size_t)b.wpc * Tb * 2;
    if (ensure_dyn_smem(reinterpret_cast<const void*>(tile_count_kernel), 220 * 1024) ||
        ensure_dyn_smem(reinterpret_cast<const void*>(tile_scatter_kernel), 220 * 1024))
        return -1;
    LaunchScope ls("tile_count", stream);
    tile_count_kernel<<<dim3(b.NG, b.nb_count), b.wpc * 32, smem, stream>>>(P, b.T, b.NW, b.L, grid_x, grid_y, b.band_rows_count, split_i, split_n, sorted_ids,
                                                                    tiles_touched, means2D, radii, rects, M, Mg);
    return 0;
}

int launch_tile_offsets_a(const BinPlan& b, const uint16_t* Mg, uint32_t* S, uint32_t* tile_base, uint2* ranges,
                          uint32_t* tile_order, uint32_t* status, uint32_t capacity, cudaStream_t stream, bool with_base,
                          bool base_zeroed) {
    LaunchScope ls("tile_offsets", stream, with_base ? 2 : 1);
    dim3 g((b.T + 255) / 256, b.nseg);
    if (!base_zeroed) cudaMemsetAsync(tile_base, 0, sizeof(uint32_t) * b.T, stream);  // (the fused entry zeroes it on its side stream)
    seg_sum_kernel<<<g, 256, 0, stream>>>(b.T, b.Tpad, b.NG, Mg, S, tile_base);
    if (with_base) tile_base_kernel<<<1, 1024, 0, stream>>>(b.T, tile_base, ranges, tile_order, status, capacity);
    return 0;
}

int launch_tile_scatter(const BinPlan& b, int P, int grid_x, int grid_y, const uint32_t* sorted_ids,
                        const uint2* rects, const uint16_t* M, const uint16_t* Mg,
                        const uint32_t* S, uint32_t* tile_base, uint32_t* E, uint32_t* point_list,
                        uint32_t capacity, cudaStream_t stream, bool base_pending, uint2* ranges, uint32_t* tile_order,
                        uint32_t* status, int split_n) {
    LaunchScope ls("tile_scatter", stream, 2);
    dim3 g((b.T + 1023) / 1024, b.nseg + (base_pending ? 1 : 0));
    col_scan_kernel<<<g, 1024, 0, stream>>>(b.T, b.Tpad, b.NG, b.nseg, Mg, S, E, tile_base, ranges, tile_order, status, capacity);
    const size_t Tb = (size_t)((std::min(b.band_rows, grid_y) * grid_x + 1) & ~1);
    const size_t smem = Tb * 4 + (size_t)b.wpc * Tb * 2;
    tile_scatter_kernel<<<dim3(b.NG, b.nb), b.wpc * 32, smem, stream>>>(P, b.T, b.NW, b.L, grid_x, grid_y, b.band_rows, split_n, sorted_ids, rects,
                                                                      M, E, tile_base, point_list, capacity);
    return 0;
}

// ---- debug / test aid: rebuild the reference's sorted 64-bit keys from (ranges, point_list, depths) --------
__global__ void rebuild_keys_kernel(int T, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                    const float* __restrict__ depths, uint64_t* __restrict__ keys) {
    const int t = blockIdx.x;
    const uint2 r = ranges[t];
    for (uint32_t p = r.x + threadIdx.x; p < r.y; p += blockDim.x)
        keys[p] = ((uint64_t)t << 32) | __float_as_uint(depths[point_list[p]]);
}

int launch_rebuild_keys(int T, const uint2* ranges, const uint32_t* point_list, const float* depths, uint64_t* keys,
                        cudaStream_t stream) {
    rebuild_keys_kernel<<<T, 256, 0, stream>>>(T, ranges, point_list, depths, keys);
    return 0;
}

}  // namespace rg
