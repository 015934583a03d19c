// Row compaction of every per-Gaussian tensor at once: GaussianModel.prune_points / _prune_optimizer
// (scene/gaussian_model.py:356-391).  The reference applies a boolean mask to 6 parameters, their 12 Adam moments and 3
// statistics one tensor at a time (each: nonzero + gather, plus re-wrapping in nn.Parameter).  Here: one exclusive scan of
// the mask (cub), one kernel that inverts it into the list of surviving rows, one kernel that copies the surviving rows
// of ALL tensors (4-byte words, coalesced writes).  Order of the survivors is preserved, as with boolean indexing.
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kMaxRowTensors = 32;
struct RowTable {
    int n;
    const uint32_t* src[kMaxRowTensors];
    uint32_t* dst[kMaxRowTensors];
    uint32_t words[kMaxRowTensors];  // 4-byte words per row
    size_t end[kMaxRowTensors];      // exclusive end in the concatenated (words per row) space
};

struct KeepToInt {
    __host__ __device__ int operator()(uint8_t k) const { return k ? 1 : 0; }
};

__global__ void __launch_bounds__(256) survivors_kernel(int P, const uint8_t* __restrict__ keep, const int* __restrict__ excl,
                                                        int* __restrict__ src_row, int* __restrict__ count_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (keep[i]) src_row[excl[i]] = i;
    if (i == P - 1) *count_out = excl[i] + (keep[i] ? 1 : 0);
}

// blockIdx.y = tensor; threads run over (surviving row, word) of that tensor: coalesced writes, row-gathered reads
__global__ void __launch_bounds__(256) gather_rows_kernel(RowTable t, const int* __restrict__ src_row,
                                                          const int* __restrict__ count) {
    const int q = blockIdx.y;
    const uint32_t words = t.words[q];
    const uint32_t* __restrict__ src = t.src[q];
    uint32_t* __restrict__ dst = t.dst[q];
    const size_t total = (size_t)(*count) * words;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t r = i / words, w = i - r * words;
        dst[i] = src[(size_t)src_row[r] * words + w];
    }
}

// dst row r = src row src_row[r], or zeros where src_row[r] < 0 (rows that start from nothing: the Adam moments of a
// cloned or split Gaussian).  n_out is known on the host here.
__global__ void __launch_bounds__(256) gather_rows_indexed_kernel(RowTable t, const int* __restrict__ src_row, size_t n_out) {
    const int q = blockIdx.y;
    const uint32_t words = t.words[q];
    const uint32_t* __restrict__ src = t.src[q];
    uint32_t* __restrict__ dst = t.dst[q];
    const size_t total = n_out * words;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const size_t r = i / words, w = i - r * words;
        const int s = src_row[r];
        dst[i] = s >= 0 ? src[(size_t)s * words + w] : 0u;
    }
}

}  // namespace rg

using namespace rg;

extern "C" {

size_t rg_compact_scratch_bytes(int P) {
    if (P <= 0) return 256;
    size_t temp = 0;
    cub::TransformInputIterator<int, KeepToInt, const uint8_t*> it(nullptr, KeepToInt());
    cub::DeviceScan::ExclusiveSum(nullptr, temp, it, (int*)nullptr, P);
    return align_up(temp, 256) + 2 * align_up(sizeof(int) * (size_t)P, 256) + 256;
}

int rg_compact_rows(int P, const uint8_t* keep, int n_tensors, const RgRowTensor* tensors, void* scratch, size_t scratch_bytes,
                    int* count_out, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P <= 0 || !keep || !tensors || n_tensors < 1 || n_tensors > kMaxRowTensors || !scratch || !count_out ||
        scratch_bytes < rg_compact_scratch_bytes(P)) {
        set_error("rg_compact_rows: bad argument (1..32 tensors, scratch of rg_compact_scratch_bytes(P))");
        return -1;
    }
    RowTable t;
    t.n = n_tensors;
    size_t acc = 0;
    for (int q = 0; q < n_tensors; q++) {
        if (!tensors[q].src || !tensors[q].dst || tensors[q].row_bytes == 0 || (tensors[q].row_bytes & 3) ||
            (reinterpret_cast<uintptr_t>(tensors[q].src) & 3) || (reinterpret_cast<uintptr_t>(tensors[q].dst) & 3)) {
            set_error("rg_compact_rows: rows must be non-empty multiples of 4 bytes, 4-byte aligned");
            return -1;
        }
        t.src[q] = static_cast<const uint32_t*>(tensors[q].src);
        t.dst[q] = static_cast<uint32_t*>(tensors[q].dst);
        t.words[q] = (uint32_t)(tensors[q].row_bytes / 4);
        acc += t.words[q];
        t.end[q] = acc;
    }
    size_t temp = 0;
    cub::TransformInputIterator<int, KeepToInt, const uint8_t*> it(keep, KeepToInt());
    cub::DeviceScan::ExclusiveSum(nullptr, temp, it, (int*)nullptr, P);
    char* base = static_cast<char*>(scratch);
    int* excl = reinterpret_cast<int*>(base + align_up(temp, 256));
    int* src_row = reinterpret_cast<int*>(base + align_up(temp, 256) + align_up(sizeof(int) * (size_t)P, 256));
    LaunchScope ls("compact_rows", stream, 0);
    if (cub::DeviceScan::ExclusiveSum(base, temp, it, excl, P, stream) != cudaSuccess) { set_error("rg_compact_rows: scan failed"); return -1; }
    survivors_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, keep, excl, src_row, count_out);
    size_t blocks = ((size_t)P * 4 + 255) / 256;  // grid-stride: about four words per thread for a 1-word row
    if (blocks > (size_t)RG_NUM_SMS * 8) blocks = (size_t)RG_NUM_SMS * 8;
    gather_rows_kernel<<<dim3((unsigned)blocks, (unsigned)n_tensors), 256, 0, stream>>>(t, src_row, count_out);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // extern "C"

// Rows of several tensors regrouped by ONE explicit index list: what densify_and_clone / densify_and_split /
// cat_tensors_to_optimizer + prune_points do with a torch.cat and a boolean-mask index per tensor and per Adam moment
// (scene/gaussian_model.py:356-518).  dst tensor q gets n_out rows; row r = src row src_row[r] (any order, repeats
// allowed), or zeros where src_row[r] < 0.
extern "C" int rg_gather_rows(long long n_out, const int* src_row, int n_tensors, const RgRowTensor* tensors, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n_out < 0 || n_tensors < 1 || n_tensors > kMaxRowTensors || !tensors || (n_out > 0 && !src_row)) {
        set_error("rg_gather_rows: bad argument (1..32 tensors)");
        return -1;
    }
    if (n_out == 0) return 0;
    RowTable t;
    t.n = n_tensors;
    size_t acc = 0, max_words = 1;
    for (int q = 0; q < n_tensors; q++) {
        if (!tensors[q].src || !tensors[q].dst || tensors[q].row_bytes == 0 || (tensors[q].row_bytes & 3) ||
            (reinterpret_cast<uintptr_t>(tensors[q].src) & 3) || (reinterpret_cast<uintptr_t>(tensors[q].dst) & 3)) {
            set_error("rg_gather_rows: rows must be non-empty multiples of 4 bytes, 4-byte aligned");
            return -1;
        }
        t.src[q] = static_cast<const uint32_t*>(tensors[q].src);
        t.dst[q] = static_cast<uint32_t*>(tensors[q].dst);
        t.words[q] = (uint32_t)(tensors[q].row_bytes / 4);
        acc += t.words[q];
        t.end[q] = acc;
        max_words = std::max<size_t>(max_words, t.words[q]);
    }
    LaunchScope ls("gather_rows", stream);
    size_t blocks = ((size_t)n_out * std::min<size_t>(max_words, 8) + 255) / 256;
    blocks = std::min<size_t>(std::max<size_t>(blocks, 1), (size_t)RG_NUM_SMS * 16);
    gather_rows_indexed_kernel<<<dim3((unsigned)blocks, (unsigned)n_tensors), 256, 0, stream>>>(t, src_row, (size_t)n_out);
    if (cudaGetLastError() != cudaSuccess) { set_error("rg_gather_rows: launch failed"); return -1; }
    return 0;
}
