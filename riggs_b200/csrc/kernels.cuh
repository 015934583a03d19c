// Internal launch interfaces between the translation units of libriggs_b200.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace rg {

struct PreprocessArgs {
    int P, D, M;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* opacities;
    const float* shs;
    const float* cov3D_precomp;
    const float* colors_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int grid_x, grid_y;
    bool prefiltered;
    bool sh_vec_ok;  // SH rows are 16-byte aligned -> 128-bit loads
    // outputs
    int* radii;
    float2* means2D;
    float* depths;
    float* cov3D_out;  // may be null
    float* rgb;
    float4* conic_opacity;
    uint32_t* tiles_touched;
    uint8_t* clamped;
    uint32_t* block_sums;  // one partial sum per 256-Gaussian block
};

int launch_preprocess(const PreprocessArgs& a, cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream);

// Exclusive scan of the per-block sums in place; total written to block_sums[nblocks] and to *total_out.
int launch_scan_block_sums(uint32_t* block_sums, int nblocks, uint32_t* total_out, cudaStream_t stream);

// Balanced key emission + inclusive offsets (one block per 256 Gaussians, same blocking as preprocess).
int launch_emit_keys(int P, int grid_x, int grid_y, const float2* means2D, const float* depths, const int* radii,
                     const uint32_t* tiles_touched, const uint32_t* block_sums_excl, uint32_t* point_offsets,
                     uint64_t* keys, uint32_t* vals, cudaStream_t stream);

size_t radix_sort_temp_bytes(size_t R);
int launch_sort_pairs(char* temp, size_t temp_bytes, const uint64_t* keys_in, uint64_t* keys_out,
                      const uint32_t* vals_in, uint32_t* vals_out, int64_t R, int end_bit, cudaStream_t stream);

int launch_tile_ranges(int64_t R, const uint64_t* keys, uint2* ranges, int n_tiles, cudaStream_t stream);

struct BlendFwdArgs {
    int W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const float2* means2D;
    const float* colors;  // 3 per Gaussian
    const float* depths;
    const float4* conic_opacity;
    const float* bg;
    float* out_color;
    float* out_depth;
    float* out_alpha;
    uint32_t* n_contrib;
};
int launch_blend_forward(const BlendFwdArgs& a, cudaStream_t stream);

struct BlendBwdArgs {
    int P, W, H, grid_x, grid_y;
    const uint2* ranges;
    const uint32_t* point_list;
    const float* bg;
    const float2* means2D;
    const float4* conic_opacity;
    const float* colors;
    const float* depths;
    const float* alphas;
    const uint32_t* n_contrib;
    const float* dL_dpix;
    const float* dL_dpix_depth;
    const float* dL_dalphas;
    float* grad_acc;  // 10 floats per Gaussian, zeroed: mean2D.xy, conic.xyw, opacity, color.rgb, depth
};
int launch_blend_backward(const BlendBwdArgs& a, cudaStream_t stream);

struct GaussBwdArgs {
    int P, D, M;
    const float* means3D;
    const int* radii;
    const float* shs;
    const uint8_t* clamped;
    const float* scales;
    const float* rotations;
    float scale_modifier;
    const float* cov3D;  // precomputed covariance (or the one kept by the forward); null -> recompute from scale/rot
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    int W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    bool sh_vec_ok;
    const float* grad_acc;  // 10P
    // outputs, all fully written
    float* dL_dmean2D;   // P*3
    float* dL_dconic;    // P*4
    float* dL_dopacity;  // P
    float* dL_dcolor;    // P*3
    float* dL_ddepth;    // P
    float* dL_dmean3D;   // P*3
    float* dL_dcov3D;    // P*6
    float* dL_dsh;       // P*M*3 or null
    float* dL_dscale;    // P*3 or null
    float* dL_drot;      // P*4 or null
};
int launch_gauss_backward(const GaussBwdArgs& a, cudaStream_t stream);

}  // namespace rg
