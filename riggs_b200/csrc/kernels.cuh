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
    float4* rec;  // 4 x float4 per Gaussian: the packed splat record of cull.cuh, all the blend kernels read
    float4* conic_opacity;
    uint32_t* tiles_touched;
    uint8_t* clamped;
    uint32_t* ids;  // iota, value array of the depth sort
    uint32_t* sort_hist;  // optional: the depth sort's four 8-bit digit histograms (4 x 256), counted here (fused entry)
};

// Extra inputs of the fused skin + activation + preprocess kernel: raw model tensors and one pose's bone table.
struct SkinFuseArgs {
    int J, M_rest;
    bool isotropic;        // _scaling has one column
    bool zero_d_rotation;  // the trainer zeroes d_rotation for isotropic Gaussians (train_rig.py:422-423)
    int blend_mode;        // 0: linear blend skinning (what the reference runs), 1: dual-quaternion blending (dq.cuh)
    const float* xyz;
    const float* scaling_raw;
    const float* rotation_raw;
    const float* opacity_raw;
    const float* features_dc;
    const float* features_rest;
    const float* motion_mask;  // optional (N)
    const float* mlp_weight;   // optional (N,J-1)
    const float* offsets;      // optional (N,3)
    const float* joints;
    const int* parents;
    const float* node_radius_raw;
    const float* G;         // (J,12)
    const float* node_rot;  // (J,4)
    const float* global_trans;
    float* means3D_out;  // (N,3) kept for the backward
    float* d_rot_out;    // (N,4) blended bone quaternion * mask, kept for the backward
    float* wsum_out;     // (N) sum of the un-normalised skinning weights, kept for the backward
    float* dbg_scales;   // optional test outputs: the activated tensors the reference rasterizer would be given
    float* dbg_rotations;
    float* dbg_opacities;
};

struct SkinFuseBwdArgs {
    SkinFuseArgs f;
    const float* d_rot_saved;  // (N,4) from the forward
    const float* wsum_saved;   // (N) from the forward
    // gradients of the raw model tensors (fully written)
    float* d_xyz;
    float* d_scaling_raw;
    float* d_rotation_raw;
    float* d_opacity_raw;
    float* d_features_dc;
    float* d_features_rest;
    // skeleton gradients (accumulated with atomics; zeroed by the launcher's caller)
    float* dG;                // (J,12)
    float* dnode_radius_raw;  // (J)
    float* dglobal_trans;     // (3)
    // optional element-wise outputs
    float* dmask;
    float* dmlp_weight;
    float* doffsets;
    // multi-view batches: ADD into the six parameter-gradient tensors instead of overwriting them (a culled Gaussian
    // then adds nothing); d_features_rest goes through the copy engine's bulk reduction
    bool accumulate;
};

int launch_preprocess(const PreprocessArgs& a, cudaStream_t stream);
int launch_skin_preprocess(const PreprocessArgs& a, const SkinFuseArgs& s, cudaStream_t stream);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream);

// ---- binning (binning.cu) ----
struct BinPlan {
    int T, Tpad;      // tiles, tiles padded to even (u16 rows are copied as u32 words)
    int NW, L;        // chunks (one warp each) of L depth-ordered Gaussians
    int wpc, NG;      // chunks per group (= warps per CTA of the two walking kernels, shared-memory bound), groups
    int nseg;         // segments of 32 group rows
    int nb, band_rows;  // the walking kernels cover the tile grid in nb bands of band_rows tile rows (grid.y = band): their
                        // per-warp counters span one band, not the image, so a 1080p grid keeps 4-5 CTAs per SM instead of 1
    int nb_count, band_rows_count;  // ... tile_count's own band split (nb / band_rows are tile_scatter's)
};
BinPlan make_bin_plan(int P, int grid_x, int grid_y);
size_t depth_sort_temp_bytes(size_t P);      // radix_sort.cu (covers the library sort's need as well)
size_t depth_sort_temp_bytes_cub(size_t P);  // binning.cu
int launch_depth_sort_cub(char* temp, size_t temp_bytes, const float* depths, uint32_t* keys_out, const uint32_t* ids_in,
                          uint32_t* ids_out, int P, cudaStream_t stream);
int launch_depth_sort(char* temp, size_t temp_bytes, const float* depths, uint32_t* keys_out, const uint32_t* ids_in,
                      uint32_t* ids_out, int P, cudaStream_t stream, bool hist_ready = false);
uint32_t* depth_sort_prepare(char* temp, int P, cudaStream_t stream);  // see radix_sort.cu
int launch_tile_count(const BinPlan& b, int P, int grid_x, int grid_y, const uint32_t* sorted_ids,
                      const uint32_t* tiles_touched, const float2* means2D, const int* radii, uint2* rects, uint16_t* M,
                      uint16_t* Mg, cudaStream_t stream, int split_i = 0, int split_n = 1);  // split: see tile_count_kernel
// segment sums + tile bases + ranges + tile launch order (longest lists first) + status[0] = num_rendered, status[1] = overflow flag
int launch_tile_offsets_a(const BinPlan& b, const uint16_t* Mg, uint32_t* S, uint32_t* tile_base, uint2* ranges,
                          uint32_t* tile_order, uint32_t* status, uint32_t capacity, cudaStream_t stream, bool with_base = true,
                          bool base_zeroed = false);
int launch_tile_scatter(const BinPlan& b, int P, int grid_x, int grid_y, const uint32_t* sorted_ids,
                        const uint2* rects, const uint16_t* M, const uint16_t* Mg,
                        const uint32_t* S, uint32_t* tile_base, uint32_t* E, uint32_t* point_list,
                        uint32_t capacity, cudaStream_t stream, bool base_pending = false, uint2* ranges = nullptr,
                        uint32_t* tile_order = nullptr, uint32_t* status = nullptr, int split_n = 1);
int launch_rebuild_keys(int T, const uint2* ranges, const uint32_t* point_list, const float* depths, uint64_t* keys,
                        cudaStream_t stream);

struct BlendFwdArgs {
    int W, H, grid_x, grid_y;
    int split_i = 0, split_n = 1;  // split_n > 1: tile rows with y % split_n != split_i belong to another GPU and are left untouched
    const uint2* ranges;
    const uint32_t* point_list;
    const float4* rec;  // packed splat records (cull.cuh)
    const uint32_t* tile_order;
    uint32_t* surv_mask;  // written: eight bitmaps over the instance list (one per 8x4 block of a tile, surv_stride words
                          // apart): bit set = the instance can reach that block (zeroed by the launcher's caller)
    size_t surv_stride;
    const float* bg;
    float* out_color;
    float* out_depth;
    float* out_alpha;
    uint32_t* n_contrib;
};
int launch_blend_forward(const BlendFwdArgs& a, cudaStream_t stream);

struct BlendBwdArgs {
    int P, W, H, grid_x, grid_y;
    int split_i = 0, split_n = 1;
    const uint2* ranges;
    const uint32_t* point_list;
    const float* bg;
    const float4* rec;
    const uint32_t* tile_order;
    const uint32_t* surv_mask;  // the forward blend's survivor bitmaps
    size_t surv_stride;
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
    int row_begin, row_end;  // rows handled by this launch (row_end = 0: all P rows); row_begin a multiple of 256
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
int launch_gauss_skin_backward(const GaussBwdArgs& a, const SkinFuseBwdArgs& s, cudaStream_t stream);

}  // namespace rg
