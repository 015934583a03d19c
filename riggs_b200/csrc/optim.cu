// Optimiser step and densification bookkeeping that follow the backward in every training iteration
// (SURVEY section 8f, N2).
//
// adam_kernel: torch.optim.Adam(l, lr=0.0, eps=1e-15) of scene/gaussian_model.py:203-217 -- six (seven) parameter
// groups with their own learning rates, betas (0.9, 0.999), no weight decay, no amsgrad -- as ONE launch over all groups
// instead of the per-group foreach launches: each element reads p, g, m, v and writes p, m, v once (28 B per element,
// HBM-bound).  Arithmetic follows torch's single-tensor Adam:
//     m <- m + (1 - b1)(g - m);   v <- b2 v + (1 - b2) g g;   p <- p - (lr / (1 - b1^t)) m / (sqrt(v) / sqrt(1 - b2^t) + eps).
//
// densify_stats_kernel: add_densification_stats (scene/gaussian_model.py:516-518) and the max_radii2D update
// (train_rig.py:335) for the Gaussians that were visible in the frame, one pass over P instead of ~10 masked launches.
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kMaxAdamGroups = 16;
struct AdamTable {
    int n;
    float* p[kMaxAdamGroups];
    const float* g[kMaxAdamGroups];
    float* m[kMaxAdamGroups];
    float* v[kMaxAdamGroups];
    size_t end[kMaxAdamGroups];  // exclusive end of the group in the concatenated index space
    float step_size[kMaxAdamGroups];  // lr / (1 - b1^t)
};

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, float one_minus_b1, float b2,
                                          float one_minus_b2, float inv_sqrt_bc2, float eps, float step_size) {
    m = m + one_minus_b1 * (g - m);
    v = v * b2 + one_minus_b2 * g * g;
    const float denom = sqrtf(v) * inv_sqrt_bc2 + eps;
    p = p - step_size * (m / denom);
}

// VEC: every group is a multiple of four elements long and 16-byte aligned; t.end then counts 16-byte words
template <bool VEC>
__global__ void __launch_bounds__(256) adam_kernel(AdamTable t, float one_minus_b1, float b2, float one_minus_b2,
                                                   float inv_sqrt_bc2, float eps) {
    const size_t total = t.end[t.n - 1];
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        int q = 0;
        while (i >= t.end[q]) q++;
        const size_t k = i - (q ? t.end[q - 1] : 0);
        const float ss = t.step_size[q];
        if (VEC) {
            const float4 g = reinterpret_cast<const float4*>(t.g[q])[k];
            float4 m = reinterpret_cast<float4*>(t.m[q])[k], v = reinterpret_cast<float4*>(t.v[q])[k];
            float4 p = reinterpret_cast<float4*>(t.p[q])[k];
            adam_elem(p.x, g.x, m.x, v.x, one_minus_b1, b2, one_minus_b2, inv_sqrt_bc2, eps, ss);
            adam_elem(p.y, g.y, m.y, v.y, one_minus_b1, b2, one_minus_b2, inv_sqrt_bc2, eps, ss);
            adam_elem(p.z, g.z, m.z, v.z, one_minus_b1, b2, one_minus_b2, inv_sqrt_bc2, eps, ss);
            adam_elem(p.w, g.w, m.w, v.w, one_minus_b1, b2, one_minus_b2, inv_sqrt_bc2, eps, ss);
            reinterpret_cast<float4*>(t.m[q])[k] = m;
            reinterpret_cast<float4*>(t.v[q])[k] = v;
            reinterpret_cast<float4*>(t.p[q])[k] = p;
        } else {
            float m = t.m[q][k], v = t.v[q][k], p = t.p[q][k];
            adam_elem(p, t.g[q][k], m, v, one_minus_b1, b2, one_minus_b2, inv_sqrt_bc2, eps, ss);
            t.m[q][k] = m;
            t.v[q][k] = v;
            t.p[q][k] = p;
        }
    }
}

__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const float* __restrict__ viewspace_grad,
                                                            const int* __restrict__ radii, float* __restrict__ accum,
                                                            float* __restrict__ denom, float* __restrict__ max_radii) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;  // update_filter = visibility_filter = radii > 0
    const float gx = viewspace_grad[3 * (size_t)i], gy = viewspace_grad[3 * (size_t)i + 1];
    accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.0f;
    if (max_radii) max_radii[i] = fmaxf(max_radii[i], (float)r);
}

}  // namespace rg

using namespace rg;

extern "C" {

int rg_adam_step(int n_groups, const RgAdamGroup* groups, int step, double beta1, double beta2, double eps,
                 rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n_groups < 1 || n_groups > kMaxAdamGroups || !groups || step < 1) {
        set_error("rg_adam_step: 1..16 groups and step >= 1");
        return -1;
    }
    AdamTable t;
    t.n = 0;
    size_t total = 0;
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    bool vec = true;
    for (int q = 0; q < n_groups; q++) {
        const RgAdamGroup& gq = groups[q];
        if (gq.n == 0) continue;
        if (!gq.param || !gq.grad || !gq.exp_avg || !gq.exp_avg_sq) { set_error("rg_adam_step: null tensor in a group"); return -1; }
        const uintptr_t bits = reinterpret_cast<uintptr_t>(gq.param) | reinterpret_cast<uintptr_t>(gq.grad) |
                               reinterpret_cast<uintptr_t>(gq.exp_avg) | reinterpret_cast<uintptr_t>(gq.exp_avg_sq);
        if ((gq.n & 3) || (bits & 15)) vec = false;
    }
    for (int q = 0; q < n_groups; q++) {
        const RgAdamGroup& gq = groups[q];
        if (gq.n == 0) continue;
        total += vec ? gq.n / 4 : gq.n;
        t.p[t.n] = gq.param; t.g[t.n] = gq.grad; t.m[t.n] = gq.exp_avg; t.v[t.n] = gq.exp_avg_sq;
        t.end[t.n] = total;
        t.step_size[t.n] = (float)((double)gq.lr / bc1);
        t.n++;
    }
    if (total == 0) return 0;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)RG_NUM_SMS * 16) blocks = (size_t)RG_NUM_SMS * 16;
    LaunchScope ls("adam", stream);
    // 1 - beta in double, rounded once, as torch does (1.0f - 0.999f is off by 1.3e-5 relative)
    const float omb1 = (float)(1.0 - beta1), omb2 = (float)(1.0 - beta2), isb = (float)(1.0 / sqrt(bc2));
    if (vec) adam_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(t, omb1, (float)beta2, omb2, isb, (float)eps);
    else adam_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(t, omb1, (float)beta2, omb2, isb, (float)eps);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int rg_densification_stats(int P, const float* viewspace_grad, const int* radii, float* xyz_gradient_accum, float* denom,
                           float* max_radii2D, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P <= 0) return 0;
    if (!viewspace_grad || !radii || !xyz_gradient_accum || !denom) { set_error("rg_densification_stats: null argument"); return -1; }
    LaunchScope ls("densify_stats", stream);
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, viewspace_grad, radii, xyz_gradient_accum, denom, max_radii2D);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // extern "C"
