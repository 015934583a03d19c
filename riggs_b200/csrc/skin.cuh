// Launch interfaces of the skinning kernels.
#pragma once
#include <cuda_runtime.h>

namespace rg {

struct SkinArgs {
    int N, J;
    const float* xyz;
    const float* joints;
    const int* parents;
    const float* node_radius_raw;
    const float* G;         // (J,12) of one pose
    const float* node_rot;  // (J,4)
    const float* global_trans;
    const float* motion_mask;  // optional
    const float* mlp_weight;   // optional (N,J-1)
    const float* offsets;      // optional (N,3)
    int blend_mode;
    float* d_xyz;
    float* d_rot;
    float* nn_weight;  // optional (N,J-1)
};

struct SkinBwdArgs {
    int N, J;
    const float* xyz;
    const float* joints;
    const int* parents;
    const float* node_radius_raw;
    const float* G;
    const float* node_rot;
    const float* global_trans;
    const float* motion_mask;
    const float* mlp_weight;
    const float* offsets;
    const float* g_dxyz;
    const float* g_drot;  // optional
    int blend_mode;       // 0: linear blend, 1: dual-quaternion blend (dq.cuh)
    float* dG;                // (J,12) accumulated
    float* dnode_radius_raw;  // (J) accumulated
    float* dglobal_trans;     // (3) accumulated, optional
    float* dmask;             // optional
    float* dmlp_weight;       // optional
    float* doffsets;          // optional
};

int launch_fk_forward(int F, int J, const float* local_rot, const float* gtrans, const float* joints,
                      const int* parents, float* G, float* node_rot, float* posed, cudaStream_t stream);
int launch_fk_backward(int F, int J, const float* local_rot, const float* joints, const int* parents, const float* G,
                       const float* dG, const float* dposed, float* dlocal_rot, float* dgtrans, const float* dgtrans_add,
                       cudaStream_t stream, float* radius_dst = nullptr, const float* radius_src = nullptr, int radius_add = 0);
int launch_skin_forward(const SkinArgs& a, cudaStream_t stream);
int launch_skin_backward(const SkinBwdArgs& a, cudaStream_t stream);

}  // namespace rg
