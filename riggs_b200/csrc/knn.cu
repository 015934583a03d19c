// distCUDA2: mean squared distance of every point to its three nearest other points.
// Replaces SimpleKNN::knn (knn/simple_knn.cu:186-221): bounding box -> 30-bit Morton codes -> sort ->
// per-box AABBs -> exact pruned search.  The answer is the exact 3-NN mean (the boxes only prune), so it
// equals the reference's up to floating-point ties.  Cold path (called once when a model is created,
// scene/gaussian_model.py:170); all scratch comes from the caller, nothing is allocated per call.
#include <cfloat>

#include "common.cuh"
#include "kernels.cuh"
#include "knn.cuh"

namespace rg {

constexpr int kBox = 256;

struct KnnScratch {
    float* minmax;      // 6 * nblk partials, then final 6 at [0..5] of `bbox`
    float* bbox;        // 6
    uint32_t* codes;    // P
    uint32_t* codes_s;  // P
    uint32_t* idx;      // P
    uint32_t* idx_s;    // P
    float* boxes;       // 6 * nbox
    char* sort_temp;
    size_t sort_temp_bytes;
    size_t total;
};

// The Morton codes are ordered by the library's own stable radix sort (radix_sort.cu, the depth sort of the forward
// pass: 32-bit keys, ids from the key positions); the 30-bit codes simply have an all-zero top digit.
static size_t knn_sort_temp(int P) { return depth_sort_temp_bytes((size_t)P); }

static KnnScratch carve_knn(char* base, int P) {
    KnnScratch s;
    Carver c(base);
    const int nblk = (P + 1023) / 1024;
    const int nbox = (P + kBox - 1) / kBox;
    s.minmax = c.take<float>(6 * (size_t)nblk);
    s.bbox = c.take<float>(8);
    s.codes = c.take<uint32_t>(P);
    s.codes_s = c.take<uint32_t>(P);
    s.idx = c.take<uint32_t>(P);
    s.idx_s = c.take<uint32_t>(P);
    s.boxes = c.take<float>(6 * (size_t)nbox);
    s.sort_temp_bytes = knn_sort_temp(P);
    s.sort_temp = c.take<char>(s.sort_temp_bytes);
    s.total = align_up(c.off, 256);
    return s;
}

size_t knn_scratch_bytes(int P) {
    if (P <= 0) return 256;
    return carve_knn(nullptr, P).total;
}

__device__ __forceinline__ float blk_reduce(float v, bool is_min, float* s_w) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float u = __shfl_xor_sync(0xffffffffu, v, o);
        v = is_min ? fminf(v, u) : fmaxf(v, u);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s_w[warp] = v;
    __syncthreads();
    const int nw = blockDim.x >> 5;
    float r = s_w[0];
    for (int w = 1; w < nw; w++) r = is_min ? fminf(r, s_w[w]) : fmaxf(r, s_w[w]);
    return r;
}

// stage 1: per-1024-point partial bounding boxes; stage 2 (one CTA): final box
__global__ void __launch_bounds__(1024) bbox_partial_kernel(int P, const float* __restrict__ pts, float* __restrict__ part) {
    __shared__ float s_w[32];
    const int i = blockIdx.x * 1024 + threadIdx.x;
    float v[3];
    for (int c = 0; c < 3; c++) v[c] = i < P ? pts[3 * (size_t)i + c] : 0.f;
    for (int c = 0; c < 3; c++) {
        float mn = blk_reduce(i < P ? v[c] : FLT_MAX, true, s_w);
        float mx = blk_reduce(i < P ? v[c] : -FLT_MAX, false, s_w);
        if (threadIdx.x == 0) { part[6 * blockIdx.x + c] = mn; part[6 * blockIdx.x + 3 + c] = mx; }
    }
}
__global__ void __launch_bounds__(1024) bbox_final_kernel(int nblk, const float* __restrict__ part, float* __restrict__ bbox) {
    __shared__ float s_w[32];
    for (int c = 0; c < 3; c++) {
        float mn = FLT_MAX, mx = -FLT_MAX;
        for (int b = threadIdx.x; b < nblk; b += 1024) { mn = fminf(mn, part[6 * b + c]); mx = fmaxf(mx, part[6 * b + 3 + c]); }
        mn = blk_reduce(mn, true, s_w);
        mx = blk_reduce(mx, false, s_w);
        if (threadIdx.x == 0) { bbox[c] = mn; bbox[3 + c] = mx; }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {  // knn/simple_knn.cu:46-53
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void morton_kernel(int P, const float* __restrict__ pts, const float* __restrict__ bbox,
                              uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float3 mn = {bbox[0], bbox[1], bbox[2]}, mx = {bbox[3], bbox[4], bbox[5]};
    const float3 p = {pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]};
    const uint32_t x = spread10(((p.x - mn.x) / (mx.x - mn.x)) * ((1 << 10) - 1));
    const uint32_t y = spread10(((p.y - mn.y) / (mx.y - mn.y)) * ((1 << 10) - 1));
    const uint32_t z = spread10(((p.z - mn.z) / (mx.z - mn.z)) * ((1 << 10) - 1));
    codes[i] = x | (y << 1) | (z << 2);
    idx[i] = i;
}

__global__ void __launch_bounds__(kBox) box_aabb_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ idx_s,
                                                        float* __restrict__ boxes) {
    __shared__ float s_w[32];
    const int i = blockIdx.x * kBox + threadIdx.x;
    float v[3] = {0, 0, 0};
    if (i < P) { const uint32_t id = idx_s[i]; for (int c = 0; c < 3; c++) v[c] = pts[3 * (size_t)id + c]; }
    for (int c = 0; c < 3; c++) {
        float mn = blk_reduce(i < P ? v[c] : FLT_MAX, true, s_w);
        float mx = blk_reduce(i < P ? v[c] : -FLT_MAX, false, s_w);
        if (threadIdx.x == 0) { boxes[6 * blockIdx.x + c] = mn; boxes[6 * blockIdx.x + 3 + c] = mx; }
    }
}

__device__ __forceinline__ void keep3(float dist, float* best) {  // knn/simple_knn.cu:120-134
#pragma unroll
    for (int j = 0; j < 3; j++)
        if (best[j] > dist) { float t = best[j]; best[j] = dist; dist = t; }
}

__device__ __forceinline__ float box_dist2(const float* bx, const float3& p) {  // knn/simple_knn.cu:106-117
    float dx = 0, dy = 0, dz = 0;
    if (p.x < bx[0] || p.x > bx[3]) dx = fminf(fabsf(p.x - bx[0]), fabsf(p.x - bx[3]));
    if (p.y < bx[1] || p.y > bx[4]) dy = fminf(fabsf(p.y - bx[1]), fabsf(p.y - bx[4]));
    if (p.z < bx[2] || p.z > bx[5]) dz = fminf(fabsf(p.z - bx[2]), fabsf(p.z - bx[5]));
    return dx * dx + dy * dy + dz * dz;
}

__global__ void __launch_bounds__(kBox) knn3_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ idx_s,
                                                    const float* __restrict__ boxes, float* __restrict__ out) {
    const int i = blockIdx.x * kBox + threadIdx.x;
    if (i >= P) return;
    const uint32_t me = idx_s[i];
    const float3 p = {pts[3 * (size_t)me], pts[3 * (size_t)me + 1], pts[3 * (size_t)me + 2]};
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    for (int j = max(0, i - 3); j <= min(P - 1, i + 3); j++) {
        if (j == i) continue;
        const uint32_t o = idx_s[j];
        const float dx = pts[3 * (size_t)o] - p.x, dy = pts[3 * (size_t)o + 1] - p.y, dz = pts[3 * (size_t)o + 2] - p.z;
        keep3(dx * dx + dy * dy + dz * dz, best);
    }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    const int nbox = (P + kBox - 1) / kBox;
    for (int b = 0; b < nbox; b++) {
        const float d = box_dist2(boxes + 6 * b, p);
        if (d > reject || d > best[2]) continue;
        const int end = min(P, (b + 1) * kBox);
        for (int j = b * kBox; j < end; j++) {
            if (j == i) continue;
            const uint32_t o = idx_s[j];
            const float dx = pts[3 * (size_t)o] - p.x, dy = pts[3 * (size_t)o + 1] - p.y, dz = pts[3 * (size_t)o + 2] - p.z;
            keep3(dx * dx + dy * dy + dz * dz, best);
        }
    }
    out[me] = (best[0] + best[1] + best[2]) / 3.0f;
}

int launch_dist2_knn3(int P, const float* points, float* out, char* scratch, size_t scratch_bytes, cudaStream_t stream) {
    if (P <= 0) return 0;
    KnnScratch s = carve_knn(scratch, P);
    if (scratch_bytes < s.total) { set_error("rg_dist_cuda2: scratch buffer too small"); return -1; }
    const int nblk = (P + 1023) / 1024, nbox = (P + kBox - 1) / kBox;
    LaunchScope ls("dist_cuda2", stream, 5);  // the sort counts its own launches
    bbox_partial_kernel<<<nblk, 1024, 0, stream>>>(P, points, s.minmax);
    bbox_final_kernel<<<1, 1024, 0, stream>>>(nblk, s.minmax, s.bbox);
    morton_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, points, s.bbox, s.codes, s.idx);
    if (launch_depth_sort(s.sort_temp, s.sort_temp_bytes, reinterpret_cast<const float*>(s.codes), s.codes_s, nullptr, s.idx_s, P, stream))
        return -1;
    box_aabb_kernel<<<nbox, kBox, 0, stream>>>(P, points, s.idx_s, s.boxes);
    knn3_kernel<<<nbox, kBox, 0, stream>>>(P, points, s.idx_s, s.boxes, out);
    return 0;
}

}  // namespace rg

// ---- rg_dist_cuda2_static: distCUDA2 for callers that cannot allocate device memory --------------------------
//
// Some hosts of this library have nowhere to get scratch memory from at the moment they need the 3-NN distances:
// a viewer or inference plug-in that is handed a fixed device-memory budget at start-up and must not call the CUDA
// allocator afterwards (cudaMalloc synchronises the whole device and can fail late, in the middle of a session), or a
// plain C / C++ host that links neither torch nor any other caching allocator and therefore has no cheap way to
// produce the rg_knn_scratch_bytes(P) bytes that rg_dist_cuda2 asks for.  For them the scratch lives in the array
// below, inside the library image.  The price is stated up front: every GPU that loads the library holds
// kStaticKnnBytes (2 MiB) until the process exits, and one call handles at most kStaticKnnMaxPoints points.
#include <mutex>

#include "../../include/riggs_b200.h"

namespace {

constexpr int kStaticKnnMaxPoints = 65536;
constexpr size_t kStaticKnnBytes = 2097152;  // 2 MiB >= knn_scratch_bytes(65536) = 1,593,856 B on a B200 (the sort's share depends on the device)
static __device__ __align__(256) unsigned char g_knn_static_scratch[kStaticKnnBytes];

// In host code `g_knn_static_scratch` is only a handle for the symbol, not an address a kernel can use; the device
// address has to be asked for, per device, with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
char* g_static_ptr[kMaxDevices] = {nullptr};

int static_scratch_for_current_device(char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        rg::set_error(std::string("rg_dist_cuda2_static: cudaGetDevice failed: ") + cudaGetErrorString(e));
        return -1;
    }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_knn_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            rg::set_error(std::string("rg_dist_cuda2_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return -1;
        }
        g_static_ptr[dev] = static_cast<char*>(p);
    }
    *out = g_static_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int rg_dist_cuda2_static(int P, const float* points, float* mean_dists, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P <= 0) return 0;
    if (P > kStaticKnnMaxPoints) {
        rg::set_error("rg_dist_cuda2_static: at most 65536 points per call (built-in scratch is 2 MiB); "
                      "use rg_dist_cuda2 with caller-provided scratch for larger clouds");
        return -1;
    }
    if (rg::knn_scratch_bytes(P) > kStaticKnnBytes) {
        rg::set_error("rg_dist_cuda2_static: built-in scratch too small for this point count");
        return -1;
    }
    char* scratch = nullptr;
    if (static_scratch_for_current_device(&scratch)) return -1;  // first call on a device looks the address up
    // Nothing in the scratch has to start at zero: the bounding-box partials, Morton codes, index arrays and box AABBs
    // are each fully written by one kernel before the next one reads them, and the radix sort clears its own counters.
    if (rg::launch_dist2_knn3(P, points, mean_dists, scratch, kStaticKnnBytes, stream)) return -1;
    return rg::check_launch("dist_cuda2_static", stream, false);
}
