// Bone table shared by the skinning kernels and by the fused skin+preprocess / per-Gaussian-backward kernels.
#pragma once
#include <cuda_runtime.h>

namespace rg {

constexpr int kMaxJoints = 128;

struct BoneSm {            // per bone k (child joint k = 1..J-1), staged once per CTA
    float a[3];            // segment start = rest joint of the parent
    float ab[3];           // segment vector
    float inv_len2;        // 1 / max(|ab|^2, 1e-6)
    float inv_2r2;         // 1 / (2 r_k^2), r_k = exp(node_radius_raw[k])
    float G[12];           // global transform of joint k
    float q[4];            // node_rot of joint k
};

__device__ __forceinline__ void stage_bones(int J, const float* joints, const int* parents, const float* radius_raw,
                                            const float* G, const float* node_rot, BoneSm* s_b) {
    for (int k = threadIdx.x + 1; k < J; k += blockDim.x) {
        BoneSm b;
        const int p = parents[k];
        float l2 = 0.f;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            b.a[c] = joints[3 * p + c];
            b.ab[c] = joints[3 * k + c] - b.a[c];
            l2 += b.ab[c] * b.ab[c];
        }
        b.inv_len2 = 1.0f / fmaxf(l2, 1e-6f);
        const float r = expf(radius_raw[k]);
        b.inv_2r2 = 1.0f / (2 * r * r);
#pragma unroll
        for (int c = 0; c < 12; c++) b.G[c] = G[12 * k + c];
#pragma unroll
        for (int c = 0; c < 4; c++) b.q[c] = node_rot ? node_rot[4 * k + c] : 0.f;
        s_b[k - 1] = b;
    }
}

// squared distance of x to the bone segment (skeleton_warp.py:215-238)
__device__ __forceinline__ float seg_dist2(const BoneSm& b, float x, float y, float z) {
    const float px = x - b.a[0], py = y - b.a[1], pz = z - b.a[2];
    float t = (px * b.ab[0] + py * b.ab[1] + pz * b.ab[2]) * b.inv_len2;
    t = fminf(fmaxf(t, 0.0f), 1.0f);
    const float sx = t * b.ab[0] - px, sy = t * b.ab[1] - py, sz = t * b.ab[2] - pz;
    return sx * sx + sy * sy + sz * sz;
}

// bytes of dynamic shared memory for the bone table (+ the per-bone gradient accumulators of a backward kernel)
inline size_t bone_smem_bytes(int J, bool with_grad_acc) {
    size_t b = sizeof(BoneSm) * (size_t)(J > 1 ? J - 1 : 1);
    if (with_grad_acc) b += 8 * sizeof(float) * ((size_t)(J - 1) * 13 + 3);  // one accumulator row per warp of a 256-thread CTA
    return b;
}

}  // namespace rg
