// L1 image loss with its gradient in one pass.  Replaces l1_loss of utils/loss_utils.py:
//   torch.abs(network_output - gt).mean()
// which, with autograd, costs the trainer seven element-wise / reduction launches over the image per iteration
// (sub, abs, mean | div, sign, mul, neg/add) -- at 800x800 they add up to the cost of the whole preprocess stage.
// Here one kernel reads both images once, writes d loss / d output = sign(output - gt) / n and reduces the loss:
// per-CTA partial sums go to a scratch array and the last CTA to finish adds them up in index order, so the result
// is deterministic for a given n (it differs from torch's reduction tree by rounding only).
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kLossThreads = 256;

__global__ void __launch_bounds__(kLossThreads) l1_loss_kernel(size_t n, const float* __restrict__ a, const float* __restrict__ b,
                                                              float* __restrict__ grad, float* __restrict__ loss,
                                                              float* __restrict__ partial, unsigned int* __restrict__ ticket) {
    const float inv_n = 1.0f / (float)n;
    float acc = 0.f;
    const size_t n4 = n >> 2;
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* b4 = reinterpret_cast<const float4*>(b);
    float4* g4 = reinterpret_cast<float4*>(grad);
    auto sgn = [&](float d) { return d > 0.f ? inv_n : (d < 0.f ? -inv_n : 0.f); };
    // four independent 16-byte pairs in flight per thread: the grid is small on purpose (every CTA ends with an atomic on
    // the same ticket word, and those serialise in L2), so a thread covers several words and must not wait for each in turn
    constexpr int kU = 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (kU - 1) * stride < n4; i += kU * stride) {
        float4 x[kU], y[kU];
#pragma unroll
        for (int u = 0; u < kU; u++) { x[u] = a4[i + u * stride]; y[u] = b4[i + u * stride]; }
#pragma unroll
        for (int u = 0; u < kU; u++) {
            const float4 d = {x[u].x - y[u].x, x[u].y - y[u].y, x[u].z - y[u].z, x[u].w - y[u].w};
            acc += (fabsf(d.x) + fabsf(d.y)) + (fabsf(d.z) + fabsf(d.w));
            g4[i + u * stride] = {sgn(d.x), sgn(d.y), sgn(d.z), sgn(d.w)};
        }
    }
    for (; i < n4; i += stride) {
        const float4 x = a4[i], y = b4[i];
        const float4 d = {x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w};
        acc += (fabsf(d.x) + fabsf(d.y)) + (fabsf(d.z) + fabsf(d.w));
        g4[i] = {sgn(d.x), sgn(d.y), sgn(d.z), sgn(d.w)};
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const size_t i = (n4 << 2) + threadIdx.x;
        const float d = a[i] - b[i];
        acc += fabsf(d);
        grad[i] = sgn(d);
    }
    __shared__ float s_part[kLossThreads / 32];
    __shared__ bool s_last;
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kLossThreads / 32; w++) s += s_part[w];
        partial[blockIdx.x] = s;
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        float s = 0.f;
        for (unsigned i = threadIdx.x; i < gridDim.x; i += 32) s += __ldcg(partial + i);
        s = warp_sum(s);
        if (threadIdx.x == 0) {
            *loss = s * inv_n;
            *ticket = 0;  // ready for the next launch
        }
    }
}

}  // namespace rg

using namespace rg;

extern "C" {

size_t rg_l1_loss_scratch_bytes(void) { return sizeof(float) * 1024 + 256; }

int rg_l1_loss(size_t n, const float* output, const float* gt, float* loss, float* grad, void* scratch, size_t scratch_bytes,
               rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n == 0) { set_error("rg_l1_loss: empty input"); return -1; }
    if (!output || !gt || !loss || !grad || !scratch || scratch_bytes < rg_l1_loss_scratch_bytes()) {
        set_error("rg_l1_loss: null argument or scratch too small");
        return -1;
    }
    if ((reinterpret_cast<uintptr_t>(output) | reinterpret_cast<uintptr_t>(gt) | reinterpret_cast<uintptr_t>(grad)) & 15) {
        set_error("rg_l1_loss: output, gt and grad must be 16-byte aligned");
        return -1;
    }
    // scratch: [0, 4096) per-CTA partial sums, then the completion ticket (must be zero before the first launch;
    // the kernel resets it)
    float* partial = static_cast<float*>(scratch);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(static_cast<char*>(scratch) + sizeof(float) * 1024);
    size_t want = ((n >> 2) + kLossThreads - 1) / kLossThreads;
    const int blocks = (int)(want < 1 ? 1 : (want > 148 ? 148 : want));  // one CTA per SM: 148 ticket atomics, not 592
    LaunchScope ls("l1_loss", stream);
    l1_loss_kernel<<<blocks, kLossThreads, 0, stream>>>(n, output, gt, grad, loss, partial, ticket);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // extern "C"
