// Stable LSD radix sort of the P Gaussians by depth bits (32-bit keys, value = Gaussian id), sized for the few
// hundred thousand keys of one frame.  Same result as the stable cub::DeviceRadixSort::SortPairs it replaces
// (which itself stands in for the depth half of the reference's 64-bit key sort, rasterizer_impl.cu:304-309):
// ascending key bits, equal keys in ascending id order.
//
// Why not the library sort: its one-sweep kernel is tuned for huge arrays (384 threads x 23 keys per CTA), so 300 k
// keys run on 34 of the 148 SMs and every one of the four digit passes is a 12-15 us latency chain (ncu,
// profiles/r1_z_summary.md).  Here a CTA takes 2048 keys (147 CTAs for 300 k: one per SM):
//   sort_hist_kernel   all four 8-bit digit histograms in one read of the keys;
//   sort_pass_kernel   one sweep per digit: keys of a tile are ranked per warp with __match_any_sync (lane order =
//                      key order, so the pass is stable), the digit counts of the tile are chained to the tiles
//                      before it by decoupled look-back (one thread per digit, 32 predecessors fetched per
//                      round), the tile is laid out sorted in shared memory and written out in runs of consecutive
//                      addresses.  Tile numbers are handed out by an atomic counter, so a tile never waits for one
//                      that has not started.  A pass whose digit is the same for every key copies its tile.
// The first pass takes the ids from the key positions (no iota array is read).
#include <cstdlib>

#include "common.cuh"
#include "kernels.cuh"

namespace rg {

constexpr int kSortBits = 8;
constexpr int kSortBins = 1 << kSortBits;
constexpr int kSortPasses = 32 / kSortBits;
constexpr int kSortThreads = 512;
constexpr int kSortItems = 4;
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048 keys per CTA
constexpr int kSortWarps = kSortThreads / 32;
constexpr uint32_t kFlagPartial = 1u << 30, kFlagInclusive = 2u << 30, kValueMask = (1u << 30) - 1;

struct SortCtl {        // all zero before a sort starts (one memset)
    uint32_t* hist;     // [passes][bins] digit totals
    uint32_t* counter;  // [passes] next tile number of each pass
    uint32_t* status;   // [passes][tiles][bins] look-back words: flag | count
    size_t words;
};

static SortCtl sort_ctl(char* base, size_t tiles) {
    SortCtl c;
    c.hist = reinterpret_cast<uint32_t*>(base);
    c.counter = c.hist + kSortPasses * kSortBins;
    c.status = c.counter + 64;
    c.words = (size_t)kSortPasses * kSortBins + 64 + (size_t)kSortPasses * tiles * kSortBins;
    return c;
}

__global__ void __launch_bounds__(256) sort_hist_kernel(const uint32_t* __restrict__ keys, int P,
                                                                 uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_h[kSortPasses * kSortBins];
    for (int i = threadIdx.x; i < kSortPasses * kSortBins; i += blockDim.x) s_h[i] = 0;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const uint32_t k = keys[i];
#pragma unroll
        for (int p = 0; p < kSortPasses; p++) atomicAdd(&s_h[p * kSortBins + ((k >> (p * kSortBits)) & (kSortBins - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kSortPasses * kSortBins; i += blockDim.x)
        if (s_h[i]) atomicAdd(&hist[i], s_h[i]);
}

__device__ __forceinline__ uint32_t ld_status(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_status(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// vals_in == nullptr: the value of a key is its position (first pass).
__global__ void __launch_bounds__(kSortThreads) sort_pass_kernel(const uint32_t* __restrict__ keys_in,
                                                                 const uint32_t* __restrict__ vals_in,
                                                                 uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                 int P, int shift, const uint32_t* __restrict__ hist,
                                                                 uint32_t* __restrict__ counter, uint32_t* __restrict__ status) {
    __shared__ uint32_t s_cnt[kSortWarps][kSortBins];  // per warp: running digit counts, then exclusive offsets
    __shared__ uint32_t s_local[kSortBins];            // first slot of the digit inside the sorted tile
    __shared__ uint32_t s_delta[kSortBins];            // (first output slot of (this tile, digit)) - s_local
    __shared__ uint32_t s_scan[2][kSortBins / 32];
    __shared__ uint32_t s_keys[kSortTile], s_vals[kSortTile];  // the tile in sorted order (coalesced write-out)
    __shared__ uint32_t s_tile;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned full = 0xffffffffu;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    for (int i = tid; i < kSortWarps * kSortBins; i += kSortThreads) (&s_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const int tile_base = (int)tile * kSortTile;
    const int base = tile_base + warp * (32 * kSortItems);

    // All keys share this digit (typically the top byte: every depth of the frame inside one factor-of-four range and
    // nothing culled): the pass is the identity permutation -- copy the tile and leave.
    if (__syncthreads_or(tid < kSortBins && hist[tid] == (uint32_t)P)) {
#pragma unroll
        for (int j = 0; j < kSortItems; j++) {
            const int i = tile_base + j * kSortThreads + tid;
            if (i < P) {
                keys_out[i] = keys_in[i];
                vals_out[i] = vals_in ? vals_in[i] : (uint32_t)i;
            }
        }
        return;
    }

    uint32_t key[kSortItems], val[kSortItems], rank[kSortItems];
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        const int i = base + j * 32 + lane;
        key[j] = i < P ? keys_in[i] : 0xffffffffu;
        val[j] = (vals_in && i < P) ? vals_in[i] : (uint32_t)i;
    }
    // ---- rank inside the warp, key order = (item, lane) -------------------------------------------------
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        const bool valid = base + j * 32 + lane < P;
        const uint32_t d = (key[j] >> shift) & (kSortBins - 1);
        const unsigned peers = __match_any_sync(full, valid ? d : kSortBins + lane);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = s_cnt[warp][d];
            s_cnt[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(full, old, leader);
        rank[j] = old + __popc(peers & ((1u << lane) - 1));
        __syncwarp();
    }
    __syncthreads();
    // ---- one thread per digit (warps 0..7): offsets of the warps, look-back over the earlier tiles -----------
    uint32_t prefix = 0, total = 0, h = 0, incl_h = 0, incl_t = 0;
    if (tid < kSortBins) {
        const int d = tid;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t c = s_cnt[w][d];
            s_cnt[w][d] = total;
            total += c;
        }
        uint32_t* mine = status + (size_t)tile * kSortBins + d;
        if (tile == 0) {
            st_status(mine, kFlagInclusive | total);
        } else {
            st_status(mine, kFlagPartial | total);
            int t = (int)tile - 1;
            while (t >= 0) {
                constexpr int W = 32;
                uint32_t v[W];
#pragma unroll
                for (int i = 0; i < W; i++) v[i] = t - i >= 0 ? ld_status(status + (size_t)(t - i) * kSortBins + d) : kFlagInclusive;
                bool done = false;
#pragma unroll
                for (int i = 0; i < W; i++) {
                    if (done) continue;
                    if ((v[i] >> 30) == 0) { done = true; continue; }  // not published yet: fetch again from here
                    prefix += v[i] & kValueMask;
                    t--;
                    if (v[i] & kFlagInclusive) { done = true; t = -1; }
                }
            }
            st_status(mine, kFlagInclusive | (prefix + total));
        }
        // exclusive scans over the 256 digits: of the array-wide digit totals (first slot of the digit in the
        // output) and of this tile's totals (first slot of the digit in the sorted tile)
        h = hist[d];
        incl_h = h;
        incl_t = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(full, incl_h, o), z = __shfl_up_sync(full, incl_t, o);
            if (lane >= o) { incl_h += y; incl_t += z; }
        }
        if (lane == 31) { s_scan[0][warp] = incl_h; s_scan[1][warp] = incl_t; }
    }
    __syncthreads();
    if (tid < kSortBins) {
        uint32_t before_h = 0, before_t = 0;
#pragma unroll
        for (int w = 0; w < kSortBins / 32; w++)
            if (w < warp) { before_h += s_scan[0][w]; before_t += s_scan[1][w]; }
        const uint32_t local = before_t + incl_t - total;
        s_local[tid] = local;
        s_delta[tid] = before_h + incl_h - h + prefix - local;
    }
    __syncthreads();
    // ---- the tile in sorted order in shared memory, then out in runs of consecutive addresses -----------------
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        if (base + j * 32 + lane < P) {
            const uint32_t d = (key[j] >> shift) & (kSortBins - 1);
            const uint32_t p = s_local[d] + s_cnt[warp][d] + rank[j];
            s_keys[p] = key[j];
            s_vals[p] = val[j];
        }
    }
    __syncthreads();
    const int n_valid = min(kSortTile, P - tile_base);
#pragma unroll
    for (int j = 0; j < kSortItems; j++) {
        const int p = tid + j * kSortThreads;
        if (p < n_valid) {
            const uint32_t k = s_keys[p];
            const uint32_t pos = s_delta[(k >> shift) & (kSortBins - 1)] + (uint32_t)p;
            keys_out[pos] = k;
            vals_out[pos] = s_vals[p];
        }
    }
}

size_t depth_sort_temp_bytes(size_t P) {
    const size_t tiles = (P + kSortTile - 1) / kSortTile;
    SortCtl c = sort_ctl(nullptr, tiles);
    const size_t own = align_up(c.words * 4, 256) + 2 * align_up(P * 4, 256);
    return std::max(own, depth_sort_temp_bytes_cub(P));
}

static bool use_cub_sort() {
    static const bool v = [] {
        const char* e = std::getenv("RG_DEPTH_SORT");
        return e && std::string(e) == "cub";
    }();
    return v;
}

// Zeroes the sort's control words and returns the address of its four digit histograms, for a producer of the keys that
// counts the digits itself while it has them in registers (the fused preprocess kernel): the sort then starts with its
// first pass instead of a histogram launch.  Null when the library sort is selected (it counts for itself).
uint32_t* depth_sort_prepare(char* temp, int P, cudaStream_t stream) {
    if (P <= 0 || use_cub_sort() || P >= (1 << 30)) return nullptr;
    const size_t tiles = ((size_t)P + kSortTile - 1) / kSortTile;
    SortCtl c = sort_ctl(temp, tiles);
    if (cudaMemsetAsync(temp, 0, c.words * 4, stream) != cudaSuccess) return nullptr;
    return c.hist;
}

int launch_depth_sort(char* temp, size_t temp_bytes, const float* depths, uint32_t* keys_out, const uint32_t* ids_in,
                      uint32_t* ids_out, int P, cudaStream_t stream, bool hist_ready) {
    if (P <= 0) return 0;
    if (use_cub_sort() || P >= (1 << 30)) return launch_depth_sort_cub(temp, temp_bytes, depths, keys_out, ids_in, ids_out, P, stream);
    LaunchScope ls("depth_sort", stream, (hist_ready ? 0 : 1) + kSortPasses);
    const size_t tiles = ((size_t)P + kSortTile - 1) / kSortTile;
    SortCtl c = sort_ctl(temp, tiles);
    uint32_t* tmp_keys = reinterpret_cast<uint32_t*>(temp + align_up(c.words * 4, 256));
    uint32_t* tmp_ids = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(tmp_keys) + align_up((size_t)P * 4, 256));
    const uint32_t* keys = reinterpret_cast<const uint32_t*>(depths);
    if (!hist_ready) {
        if (cudaMemsetAsync(temp, 0, c.words * 4, stream) != cudaSuccess) { set_error("depth sort: memset failed"); return -1; }
        sort_hist_kernel<<<std::min<int>((int)((P + 2047) / 2048), RG_NUM_SMS * 2), 256, 0, stream>>>(keys, P, c.hist);
    }
    // depths -> tmp -> out -> tmp -> out
    const uint32_t* kin = keys;
    const uint32_t* vin = nullptr;
    for (int p = 0; p < kSortPasses; p++) {
        uint32_t* kout = (p & 1) ? keys_out : tmp_keys;
        uint32_t* vout = (p & 1) ? ids_out : tmp_ids;
        sort_pass_kernel<<<(unsigned)tiles, kSortThreads, 0, stream>>>(kin, vin, kout, vout, P, p * kSortBits, c.hist + p * kSortBins,
                                                                       c.counter + p, c.status + (size_t)p * tiles * kSortBins);
        kin = kout;
        vin = vout;
    }
    if (cudaGetLastError() != cudaSuccess) { set_error("depth sort: launch failed"); return -1; }
    return 0;
}

}  // namespace rg

// Test aid: the depth sort on caller-provided keys (P x u32 -> sorted keys + original positions); temp must hold
// rg_debug_depth_sort_bytes(P) bytes.
extern "C" size_t rg_debug_depth_sort_bytes(int P) { return rg::depth_sort_temp_bytes((size_t)std::max(P, 1)); }
extern "C" int rg_debug_depth_sort(int P, const uint32_t* keys, uint32_t* keys_out, uint32_t* ids_out, char* temp,
                                   void* stream) {
    return rg::launch_depth_sort(temp, rg::depth_sort_temp_bytes((size_t)std::max(P, 1)), reinterpret_cast<const float*>(keys),
                                 keys_out, nullptr, ids_out, P, static_cast<cudaStream_t>(stream));
}
