#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace rg {
size_t knn_scratch_bytes(int P);
int launch_dist2_knn3(int P, const float* points, float* out, char* scratch, size_t scratch_bytes, cudaStream_t stream);
}  // namespace rg
