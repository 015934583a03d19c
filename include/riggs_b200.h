/*
 * riggs_b200.h -- C ABI of libriggs_b200.so (sm_100a): the skin + splat-rasterize hot path of RigGS.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream; no torch types cross this
 * boundary.  All float data is fp32, all pointers are DEVICE pointers unless a comment says "host".
 * Matrices `viewmatrix` / `projmatrix` are 16 floats in the reference's transposed storage
 * (element [4*c + r] = M[r][c], scene/cameras.py:61-71).  A NULL optional pointer selects the other
 * code path exactly like the reference's null tensors do (dgr/rasterize_points.cu:96-115).
 *
 * Return value: >= 0 on success (rg_rasterize_forward returns num_rendered), < 0 on error with the
 * message available from rg_last_error().  With debug != 0 every launch is followed by a stream
 * synchronise + error check, like the reference's CHECK_CUDA (dgr/cuda_rasterizer/auxiliary.h:166-173).
 *
 * Paths below are relative to the reference tree; dgr/ = submodules/diff-gaussian-rasterization/,
 * knn/ = submodules/simple-knn/.
 */
#ifndef RIGGS_B200_H_INCLUDED
#define RIGGS_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* rg_stream_t; /* a cudaStream_t */

/* Growable state buffers: the callee asks the caller for `bytes` of device memory and gets the base
 * address back.  Replaces the three std::function<char*(size_t)> of
 * dgr/cuda_rasterizer/rasterizer.h:33-35 (bound to torch resize_ lambdas at dgr/rasterize_points.cu:27-33). */
typedef char* (*rg_alloc_fn)(void* ctx, size_t bytes);

int rg_version(void);
const char* rg_last_error(void);

/* ---- rasterizer ------------------------------------------------------------------------------ */

/* Sizes of the three opaque state buffers (replaces CudaRasterizer::required<T>,
 * dgr/cuda_rasterizer/rasterizer_impl.h:64-71). */
size_t rg_geom_bytes(int P, int width, int height);
size_t rg_image_bytes(int width, int height);
size_t rg_binning_bytes(int64_t num_rendered);

/* Replaces CudaRasterizer::Rasterizer::forward (dgr/cuda_rasterizer/rasterizer.h:32-58,
 * rasterizer_impl.cu:197-339; called from RasterizeGaussiansCUDA, dgr/rasterize_points.cu:91-116).
 * Stages: preprocess -> depth sort of the Gaussians -> per-tile instance counts and ranges ->
 * [one 4-byte D2H read of num_rendered, as the reference does, to size the binning buffer] ->
 * in-order scatter of the instance ids into their tiles -> alpha blend.  The instance list equals the
 * reference's sorted point_list (same tiles, same depth order, same tie order); no 64-bit keys are built.
 * out_color (3,H,W), out_depth (1,H,W), out_alpha (1,H,W), radii (P) int32 are fully written. */
int rg_rasterize_forward(rg_alloc_fn geom_alloc, void* geom_ctx, rg_alloc_fn binning_alloc, void* binning_ctx,
                         rg_alloc_fn image_alloc, void* image_ctx, int P, int D, int M,
                         const float* background, int width, int height, const float* means3D,
                         const float* shs, const float* colors_precomp, const float* opacities,
                         const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                         const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered,
                         float* out_color, float* out_depth, float* out_alpha, int* radii, int debug,
                         rg_stream_t stream);

/* Replaces CudaRasterizer::Rasterizer::backward (dgr/cuda_rasterizer/rasterizer.h:60-89,
 * rasterizer_impl.cu:343-447; called from RasterizeGaussiansBackwardCUDA, dgr/rasterize_points.cu:171-204).
 * Every output is fully written by the kernels (no zero-initialisation needed by the caller):
 * dL_dmean2D (P,3) [z = 0], dL_dconic (P,4) [.z = 0], dL_dopacity (P), dL_dcolor (P,3), dL_ddepth (P),
 * dL_dmean3D (P,3), dL_dcov3D (P,6), dL_dsh (P,M,3) or NULL, dL_dscale (P,3) / dL_drot (P,4) or NULL.
 * dL_dpix_depth / dL_dalphas may be NULL (no cotangent for that output). */
int rg_rasterize_backward(int P, int D, int M, int R, const float* background, int width, int height,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* alphas, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                          const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                          const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
                          const float* dL_dpix, const float* dL_dpix_depth, const float* dL_dalphas,
                          float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                          float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                          float* dL_dscale, float* dL_drot, int debug, rg_stream_t stream);

/* Replaces CudaRasterizer::Rasterizer::markVisible (dgr/cuda_rasterizer/rasterizer.h:25-30,
 * rasterizer_impl.cu:141-153).  present: P bytes (bool). */
int rg_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    uint8_t* present, rg_stream_t stream);

/* Addresses of the arrays inside the opaque buffers, for tests and debuggers (the reference keeps the
 * same information in GeometryState / BinningState / ImageState, dgr/cuda_rasterizer/rasterizer_impl.h:30-61).
 * out[0]=depths f32[P] (+inf where culled)  out[1]=clamped u8[P] (bit c = channel c clamped)
 * out[2]=means2D f32x2[P]         out[3]=conic_opacity f32x4[P]      out[4]=splat records f32x16[P] (rgb at 8..10)
 * out[5]=tiles_touched u32[P]     out[6]=ids in depth order u32[P]   out[7]=means3D f32[3P] (fused entry only)
 * out[8]=point_list u32[R]        out[9]=n_contrib u32[H*W]          out[10]=ranges u32x2[tiles]
 * out[11]=status u32[2]: num_rendered, overflow flag
 * out[12]=the backward's per-Gaussian screen-space accumulators f32[10P] (mean2D.xy, conic.xyw, opacity, colour.rgb,
 *         depth), array-major: zeroed and filled by phase 1 of rg_skinned_backward_part, read by phase 2 */
int rg_state_ptrs(int P, int64_t R, int width, int height, char* geom_buffer, char* binning_buffer,
                  char* image_buffer, void** out /* host, 12 entries */);

/* Test aid: rebuilds the reference's SORTED 64-bit keys ((tile << 32) | depth bits,
 * dgr/cuda_rasterizer/rasterizer_impl.cu:102-105) for the R instances of the last forward, from ranges,
 * point_list and depths, so the bit-exact gate on keys can be applied although the kernels never build them. */
int rg_debug_sorted_keys(int P, int64_t R, int width, int height, char* geom_buffer, char* binning_buffer,
                         char* image_buffer, uint64_t* keys_out, rg_stream_t stream);

/* Test aid: the stable depth sort of the forward pass (csrc/radix_sort.cu; replaces the depth half of the reference's
 * cub::DeviceRadixSort::SortPairs, dgr/cuda_rasterizer/rasterizer_impl.cu:304-309) on caller-provided 32-bit keys:
 * keys_out = keys ascending, ids_out = their original positions, equal keys in ascending position order.
 * temp: rg_debug_depth_sort_bytes(P) bytes of device memory. */
size_t rg_debug_depth_sort_bytes(int P);
int rg_debug_depth_sort(int P, const uint32_t* keys, uint32_t* keys_out, uint32_t* ids_out, char* temp, rg_stream_t stream);

/* ---- simple-knn -------------------------------------------------------------------------------- */

/* Replaces SimpleKNN::knn (knn/simple_knn.h:18, knn/simple_knn.cu:186-221; bound by distCUDA2,
 * knn/spatial.cu:15-26): mean squared distance of every point to its 3 nearest other points.
 * scratch: device memory of rg_knn_scratch_bytes(P) bytes, caller-owned. */
size_t rg_knn_scratch_bytes(int P);
int rg_dist_cuda2(int P, const float* points, float* mean_dists, char* scratch, size_t scratch_bytes,
                  rg_stream_t stream);

/* No counterpart in the reference.  Same result as rg_dist_cuda2, for callers that have no way to provide the
 * scratch buffer: a viewer / inference plug-in living inside a fixed device-memory budget that must not touch the
 * CUDA allocator once it is running (cudaMalloc stalls the whole device and may fail late), or a plain C host that
 * carries no allocator of its own.  The scratch is a 2 MiB array built into the library, so
 *   - every GPU that loads the library holds those 2 MiB until the process ends;
 *   - P is limited to 65536 points per call; a larger P is refused with an error (rg_last_error);
 *   - the function itself calls no cudaMalloc / cudaFree and no allocating library routine; points and mean_dists
 *     are still the caller's buffers;
 *   - all calls on one GPU share the array: calls whose kernels can overlap on the GPU (different streams, or
 *     different host threads) race on it and must be serialised by the caller.  Nothing in it needs clearing
 *     between calls;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress) before launching;
 *     make that first call outside any stream capture. */
int rg_dist_cuda2_static(int P, const float* points, float* mean_dists, rg_stream_t stream);

/* ---- skeleton skinning -------------------------------------------------------------------------- */

/* Per-frame bone table produced by forward kinematics and consumed by the skinning kernels:
 *   G        (F,J,12)  global transform of every joint: row-major 3x3 rotation then translation
 *   node_rot (F,J,4)   matrix_to_quaternion(G_R) (utils/time_utils.py:146-205), no sign standardisation
 *   posed    (F,J,3)   posed joints + global_trans  (= d_nodes, skeleton_utils/skeleton_warp.py:146)
 * Replaces quaternion_to_matrix (utils/time_utils.py:115-132), SkeletonWarp.chain_product_transform
 * (skeleton_utils/skeleton_warp.py:242-273) and matrix_to_quaternion for F poses at once.
 * local_rot (F,J,4) un-normalised quaternions, global_trans (F,3), joints (J,3) rest pose,
 * parents (J) int32 with parents[0] = -1 and parents[i] < i. */
int rg_fk_forward(int F, int J, const float* local_rot, const float* global_trans, const float* joints,
                  const int* parents, float* G, float* node_rot, float* posed, rg_stream_t stream);

/* Reverse of rg_fk_forward (the reference relies on torch autograd through the same ops):
 * dG (F,J,12), dposed (F,J,3) or NULL  ->  dlocal_rot (F,J,4), dglobal_trans_from_posed (F,3) or NULL. */
int rg_fk_backward(int F, int J, const float* local_rot, const float* joints, const int* parents,
                   const float* G, const float* dG, const float* dposed, float* dlocal_rot,
                   float* dglobal_trans, rg_stream_t stream);

/* Linear-blend skinning of N points by the J-1 bones of one pose.  Replaces
 * SkeletonWarp.cal_nn_weight_skeleton with K = -1 and the LBS part of SkeletonWarp.deform_by_pose
 * (skeleton_utils/skeleton_warp.py:41-76, 130-172, 207-238):
 *   d_xyz = mask * (sum_k w_k (R_k x + t_k) + global_trans + offsets - x),  d_rot = mask * sum_k w_k q_k.
 * blend_mode 0 = linear blend (what the reference runs); 1 = dual-quaternion blend of the bone
 * transforms (utils/dual_quaternion.py:135-180) for d_xyz, d_rot then holds the blended unit quaternion.
 * Optional (NULL = absent): motion_mask (N), mlp_weight (N,J-1) = sigmoid(WeightMLP(x)),
 * offsets (N,3) = detail_net output, nn_weight (N,J-1) output. */
int rg_skin_forward(int N, int J, const float* xyz, const float* joints, const int* parents,
                    const float* node_radius_raw, const float* G, const float* node_rot,
                    const float* global_trans, const float* motion_mask, const float* mlp_weight,
                    const float* offsets, int blend_mode, float* d_xyz, float* d_rot, float* nn_weight,
                    rg_stream_t stream);

/* Backward of rg_skin_forward (blend_mode 0).  Inputs: g_dxyz (N,3), g_drot (N,4).
 * Outputs: dG (J,12) and dnode_radius_raw (J) and dglobal_trans (3) are ACCUMULATED with atomics and
 * must be zeroed by the caller; dmask (N), dmlp_weight (N,J-1), doffsets (N,3) are written (NULL = skip).
 * No gradient flows to xyz (x.detach(), skeleton_warp.py:131) nor through node_rot (:144). */
int rg_skin_backward(int N, int J, const float* xyz, const float* joints, const int* parents,
                     const float* node_radius_raw, const float* G, const float* node_rot,
                     const float* global_trans, const float* motion_mask, const float* mlp_weight,
                     const float* offsets, const float* g_dxyz, const float* g_drot, float* dG,
                     float* dnode_radius_raw, float* dglobal_trans, float* dmask, float* dmlp_weight,
                     float* doffsets, rg_stream_t stream);

/* rg_skin_backward for either blend mode of rg_skin_forward.  blend_mode 1 (dual-quaternion blending,
 * utils/dual_quaternion.py:135-180 differentiated: the reference would rely on torch autograd through QT2DQ / DQ2QT):
 * the bone rotations enter through node_rot, which carries no gradient (skeleton_warp.py:144), so of dG only the
 * translation entries 9..11 of every bone receive one; everything else as rg_skin_backward. */
int rg_skin_backward_blend(int N, int J, const float* xyz, const float* joints, const int* parents,
                           const float* node_radius_raw, const float* G, const float* node_rot,
                           const float* global_trans, const float* motion_mask, const float* mlp_weight,
                           const float* offsets, int blend_mode, const float* g_dxyz, const float* g_drot, float* dG,
                           float* dnode_radius_raw, float* dglobal_trans, float* dmask, float* dmlp_weight,
                           float* doffsets, rg_stream_t stream);

/* Linear-blend skinning by the K nearest bones of every point: SkeletonWarp.cal_nn_weight_skeleton with K > 0
 * (skeleton_utils/skeleton_warp.py:46-49: torch.topk of the squared point-to-segment distances, smallest first,
 * nn_idx = bone + 1) followed by the LBS of deform_by_pose (:130-172).  Same arguments as rg_skin_forward without
 * mlp_weight (the reference's K > 0 branch indexes the MLP output out of range, :59) and with linear blending only.
 * Outputs: d_xyz (N,3), d_rot (N,4), nn_weight (N,K), nn_idx (N,K) int64 joint numbers 1..J-1, nn_dist2 (N,K).
 * 1 <= K <= min(8, J - 1). */
int rg_skin_topk_forward(int N, int J, int K, const float* xyz, const float* joints, const int* parents,
                         const float* node_radius_raw, const float* G, const float* node_rot, const float* global_trans,
                         const float* motion_mask, const float* offsets, float* d_xyz, float* d_rot, float* nn_weight,
                         long long* nn_idx, float* nn_dist2, rg_stream_t stream);

/* Backward of rg_skin_topk_forward for the bones nn_idx selected (the selection itself has no gradient).  g_dxyz (N,3),
 * g_drot (N,4) (NULL = none); dG (J,12), dnode_radius_raw (J), dglobal_trans (3) are ACCUMULATED (zero them first);
 * dmask (N), doffsets (N,3) are written (NULL = skip). */
int rg_skin_topk_backward(int N, int J, int K, const float* xyz, const float* joints, const int* parents,
                          const float* node_radius_raw, const float* G, const float* node_rot, const float* global_trans,
                          const float* motion_mask, const float* offsets, const long long* nn_idx, const float* g_dxyz,
                          const float* g_drot, float* dG, float* dnode_radius_raw, float* dglobal_trans, float* dmask,
                          float* doffsets, rg_stream_t stream);

/* ---- stage-1 control-node warp ------------------------------------------------------------------ */

/* K nearest control nodes, their weights and the blend of the per-node motion onto N points.  Replaces
 * ControlNodeWarp.cal_nn_weight (utils/time_utils.py:934-967, non-skinning branch with gs_kernel=True; the search
 * itself is pytorch3d.ops.knn_points there) and the blending of ControlNodeWarp.forward (:1133-1198, 1213-1224;
 * the animation branches with node_trans_bias are not covered):
 *   search space: [x | feature[:, :hyper]] against nodes (M, 3+hyper), squared distances, K smallest ascending;
 *   w_k = exp(-dist_k / (2 exp(radius_raw)^2)) [* sigmoid(weight_raw)] + 1e-7, normalised over the K neighbours;
 *   d_xyz = mask * (sum_k w_k (R_k (x - c_k) + c_k + t_k) - x)   with R_k = quaternion_to_matrix(local_quat_k),
 *           c_k = nodes[k, :3]  (local_quat != NULL: "local_frame"),   else   mask * sum_k w_k t_k;
 *   d_rot / d_scale / d_opacity / d_color = mask * sum_k w_k (node value); add_rot_bias != 0 adds [1,0,0,0] to
 *   d_rot after masking (d_rot_as_res = False).
 * local_quat is the un-normalised quaternion local_rotation + [1,0,0,0] (M,4).  Optional (NULL = absent):
 * feature (needed when hyper > 0), node_weight_raw, local_quat, node_opacity (M) with d_opacity (N),
 * node_color (M,3) with d_color (N,3), motion_mask (N).  nn_idx (N,K) int64, nn_weight (N,K), nn_dist (N,K) are
 * always written (they are what cal_nn_weight returns and what the backward reads).  1 <= K <= 8, hyper <= 5. */
int rg_node_warp_forward(int N, int M, int K, int hyper, const float* x, const float* feature, const float* nodes,
                         const float* node_radius_raw, const float* node_weight_raw, const float* local_quat,
                         const float* node_trans, const float* node_rot, const float* node_scale,
                         const float* node_opacity, const float* node_color, const float* motion_mask, int add_rot_bias,
                         float* d_xyz, float* d_rot, float* d_scale, float* d_opacity, float* d_color,
                         long long* nn_idx, float* nn_weight, float* nn_dist, rg_stream_t stream);

/* Backward of rg_node_warp_forward (the reference relies on torch autograd through the same ops).  Cotangents
 * g_* (NULL = none).  dnode (M, rg_node_warp_grad_channels()) is ACCUMULATED and must be zeroed by the caller; a
 * row holds the gradients of [node_trans 0..2 | node_rot 3..6 | node_scale 7..9 | local_quat 10..13 |
 * node_radius_raw 14 | node_weight_raw 15 | node_opacity 16 | node_color 17..19 | nodes[:, 3:3+hyper] 20..24].
 * dmask (N) and dfeature (N,hyper) are written (NULL = skip).  No gradient flows to x or to nodes[:, :3]
 * (both detached, utils/time_utils.py:946-949, 1136, 1153). */
int rg_node_warp_grad_channels(void);
int rg_node_warp_backward(int N, int M, int K, int hyper, const float* x, const float* feature, const float* nodes,
                          const float* node_radius_raw, const float* node_weight_raw, const float* local_quat,
                          const float* node_trans, const float* node_rot, const float* node_scale,
                          const float* node_opacity, const float* node_color, const float* motion_mask,
                          const long long* nn_idx, const float* nn_dist, const float* g_xyz, const float* g_rot,
                          const float* g_scale, const float* g_opacity, const float* g_color, float* dnode,
                          float* dmask, float* dfeature, rg_stream_t stream);

/* ---- fused skin + rasterize (the hot path as ONE call per direction) ------------------------------ */

/* What gaussian_renderer.render() + SkeletonWarp.deform_by_pose() are given, as raw device tensors: the model's
 * un-activated parameters (scene/gaussian_model.py:105-132), the skeleton, one pose and one camera.  The
 * fused entry points fold the whole chain  pose -> FK -> LBS -> means3D / exp / normalize / sigmoid / SH cat ->
 * rasterize  (skeleton_utils/skeleton_warp.py:130-172, gaussian_renderer/__init__.py:37-151) into the kernels;
 * they have no counterpart function in the reference (it runs ~70 eager torch kernels + the rasterizer). */
typedef struct RgSkinnedScene {
    int P, J;               /* Gaussians, joints */
    int sh_degree, M_rest;  /* active SH degree; rows of features_rest (15 for degree 3) */
    int isotropic;          /* scaling_raw is (P,1) and is repeated over the three axes */
    int zero_d_rotation;    /* drop the blended bone rotation (trainer convention for isotropic Gaussians) */
    int width, height;
    float tan_fovx, tan_fovy, scale_modifier;
    const float *xyz, *scaling_raw, *rotation_raw, *opacity_raw, *features_dc, *features_rest;
    const float *motion_mask;   /* (P) or NULL = 1 */
    const float *mlp_weight;    /* (P,J-1) sigmoid(WeightMLP(x)) or NULL */
    const float *offsets;       /* (P,3) detail-net output or NULL */
    const float *joints;        /* (J,3) rest pose */
    const int *parents;         /* (J) */
    const float *node_radius_raw; /* (J) */
    const float *local_rot;     /* (J,4) un-normalised quaternions of the pose */
    const float *global_trans;  /* (3) */
    const float *viewmatrix, *projmatrix, *campos, *background;
    int blend_mode;             /* 0: linear blend skinning (skeleton_utils/skeleton_warp.py:149-150, what the reference runs);
                                   1: the bone transforms blended as dual quaternions (utils/dual_quaternion.py:135-180) */
    int split_index, split_count; /* one view rendered by split_count GPUs (last bullet; no counterpart
                                   in the reference): with split_count > 1 this call bins and blends only the 16-pixel tile
                                   rows y with y % split_count == split_index -- pixels of other rows are left untouched in
                                   out_color / out_depth / out_alpha -- and phase 1 of the backward adds only those rows'
                                   share to the per-Gaussian screen-space accumulators (out[12] of rg_state_ptrs), which the
                                   caller sums over the GPUs before phases 2 and 4.  0 or 1: the whole image. */
} RgSkinnedScene;

/* Caller-owned device scratch, reused across frames; nothing is allocated and the host never waits.
 * geom: rg_geom_bytes(P, width, height); image: rg_image_bytes(width, height); binning: any size -- it bounds
 * the number of (Gaussian, tile) instances a frame may produce (rg_binning_capacity; 5 bytes per instance: the list
 * and the forward blend's survivor bitmaps, laid out for the CAPACITY -- pass rg_binning_capacity(bytes) as num_rendered
 * to the reference-shaped calls that are to read a workspace's binning buffer). */
typedef struct RgWorkspace {
    char* geom;    size_t geom_bytes;
    char* image;   size_t image_bytes;
    char* binning; size_t binning_bytes;
} RgWorkspace;

uint32_t rg_binning_capacity(size_t binning_bytes);

/* Forward: out_color (3,H,W), out_depth (1,H,W), out_alpha (1,H,W), radii (P); optional d_nodes (J,3) = posed
 * joints + global_trans; optional dbg_* (P,3)/(P,4)/(P) = the activated scales / rotations / opacities the
 * kernel derived (test aid: these plus out[7] of rg_state_ptrs are what the reference rasterizer would be fed). */
int rg_skinned_forward(const RgSkinnedScene* scene, const RgWorkspace* ws, float* out_color, float* out_depth,
                       float* out_alpha, int* radii, float* d_nodes, float* dbg_scales, float* dbg_rotations,
                       float* dbg_opacities, rg_stream_t stream);

/* Queues an asynchronous copy of {num_rendered, overflow flag} of the last forward into host_out2 (pinned host
 * memory, 2 x uint32).  overflow != 0 means the binning buffer was too small: the frame is invalid (but memory
 * safe); grow the buffer and render again. */
int rg_skinned_status(const RgSkinnedScene* scene, const RgWorkspace* ws, uint32_t* host_out2, rg_stream_t stream);

/* Backward of rg_skinned_forward on the same scene/workspace.  Inputs: radii and alphas (= out_alpha) of the
 * forward, dL/dcolor (3,H,W), dL/ddepth (1,H,W) and dL/dalpha (1,H,W) (either may be NULL: that output has no
 * cotangent, which saves the caller a zero-filled image), optional dL/d(d_nodes) (J,3).
 * Outputs (fully written): gradients of the raw tensors d_xyz (P,3), d_scaling_raw (P,3|1), d_rotation_raw (P,4),
 * d_opacity_raw (P), d_features_dc (P,3), d_features_rest (P,M_rest,3), d_mean2D (P,3) [screen-space, z = 0],
 * d_local_rot (J,4), d_global_trans (3), d_node_radius_raw (J); optional d_mask (P), d_mlp_weight (P,J-1),
 * d_offsets (P,3). */
int rg_skinned_backward(const RgSkinnedScene* scene, const RgWorkspace* ws, const int* radii, const float* alphas,
                        const float* dL_dcolor, const float* dL_ddepth, const float* dL_dalpha,
                        const float* dL_dnodes, float* d_xyz, float* d_scaling_raw, float* d_rotation_raw,
                        float* d_opacity_raw, float* d_features_dc, float* d_features_rest, float* d_mean2D,
                        float* d_local_rot, float* d_global_trans, float* d_node_radius_raw, float* d_mask,
                        float* d_mlp_weight, float* d_offsets, rg_stream_t stream);

/* The trainer's image loss and its gradient in two launches (train_rig.py:508-509):
 *   loss = (1 - lambda_dssim) * mean|output - gt| + lambda_dssim * (1 - ssim(output, gt)),
 * ssim = utils/loss_utils.py:36-77 (11x11 Gaussian window, sigma 1.5, zero padding, per channel, mean over all).
 * output, gt, grad: (channels, height, width) fp32; loss3 receives {loss, l1, ssim}; scratch:
 * rg_ssim_loss_scratch_bytes(channels, height, width) bytes, no initialisation needed.  lambda_dssim = 1 gives
 * grad = -d ssim / d output.  Replaces l1_loss + ssim + their autograd (five grouped conv2d and ~20 element-wise
 * launches forward, as many backward). */
size_t rg_ssim_loss_scratch_bytes(int channels, int height, int width);
int rg_ssim_l1_loss(int channels, int height, int width, const float* output, const float* gt, float lambda_dssim,
                    float* loss3, float* grad, void* scratch, size_t scratch_bytes, rg_stream_t stream);

/* ---- what follows the backward in a training iteration --------------------------- */

/* One parameter group of torch.optim.Adam as the reference sets it up (scene/gaussian_model.py:203-217). */
typedef struct RgAdamGroup {
    float* param;        /* updated in place */
    const float* grad;
    float* exp_avg;      /* Adam state, zero before the first step */
    float* exp_avg_sq;
    size_t n;            /* elements */
    float lr;
} RgAdamGroup;

/* torch.optim.Adam.step() for up to 16 groups in one launch (no weight decay, no amsgrad); `step` = 1 for the first
 * call.  The reference uses betas (0.9, 0.999) and eps 1e-15; they are doubles because torch forms 1 - beta in double
 * before rounding to fp32. */
int rg_adam_step(int n_groups, const RgAdamGroup* groups, int step, double beta1, double beta2, double eps,
                 rg_stream_t stream);

/* add_densification_stats (scene/gaussian_model.py:516-518) with update_filter = radii > 0, plus the max_radii2D update of
 * train_rig.py:335 (max_radii2D may be NULL): for every visible Gaussian
 * xyz_gradient_accum += |viewspace_grad.xy|, denom += 1, max_radii2D = max(max_radii2D, radii).
 * viewspace_grad (P,3), radii (P) int32, the three statistics (P) fp32. */
int rg_densification_stats(int P, const float* viewspace_grad, const int* radii, float* xyz_gradient_accum, float* denom,
                           float* max_radii2D, rg_stream_t stream);

/* Boolean-mask compaction of every per-Gaussian tensor in one go: GaussianModel.prune_points / _prune_optimizer
 * (scene/gaussian_model.py:356-391) apply `tensor[valid_points_mask]` to the parameters, their Adam moments and the
 * densification statistics one by one.  keep: P bytes (non-zero = the row survives); every tensor has P rows of
 * row_bytes (a multiple of 4) and a destination with room for the surviving rows (P rows is always enough); survivors
 * keep their order; *count_out (device int) receives their number.  scratch: rg_compact_scratch_bytes(P). */
typedef struct RgRowTensor {
    const void* src;
    void* dst;
    size_t row_bytes;
} RgRowTensor;
size_t rg_compact_scratch_bytes(int P);
int rg_compact_rows(int P, const uint8_t* keep, int n_tensors, const RgRowTensor* tensors, void* scratch, size_t scratch_bytes,
                    int* count_out, rg_stream_t stream);

/* Rows of up to 32 tensors regrouped by one explicit index list: dst tensor q gets n_out rows, row r = src row
 * src_row[r] (any order, repeats allowed) or zeros where src_row[r] < 0.  Replaces the torch.cat + boolean-index pairs
 * of densification_postfix / cat_tensors_to_optimizer / prune_points (scene/gaussian_model.py:356-440) when points are
 * cloned or split: parameters are gathered from their parents, the Adam moments of new points start at zero. */
int rg_gather_rows(long long n_out, const int* src_row, int n_tensors, const RgRowTensor* tensors, rg_stream_t stream);

/* K nearest of M nodes for each of N points, brute force: the search the reference's stage-1 control-node code gets from
 * pytorch3d.ops.knn_points (utils/time_utils.py:950,980,1051,1061,1124).  points (N,D), nodes (M,D), 1 <= D <= 8,
 * 1 <= K <= 16; dists (N,K) squared Euclidean distances in ascending order, idx (N,K) int64; ties go to the lower index;
 * with M < K the tail is padded with distance 0 / index 0 as pytorch3d does. */
int rg_knn_points(int N, int M, int D, int K, const float* points, const float* nodes, float* dists, long long* idx,
                  rg_stream_t stream);

/* rg_skinned_backward in parts, for callers that overlap the gradient all-reduce with the backward: `phases` is a bit
 * set -- 1: zero the accumulators and run the blend backward, 2: per-Gaussian backward (parameter gradients) of rows
 * [row_begin, row_end) only (row_begin a multiple of 256), 4: skeleton backward (d_local_rot, d_global_trans,
 * d_node_radius_raw).  Running 1, then 2 over consecutive row chunks, then 4 equals rg_skinned_backward.
 * Bit 8 (with 2 and/or 4): the gradients of the PARAMETERS -- d_xyz, d_scaling_raw, d_rotation_raw, d_opacity_raw,
 * d_features_dc, d_features_rest, d_node_radius_raw -- are ADDED to what the buffers hold instead of overwriting it
 * (the second and later views of a multi-view batch: configs of several views per optimiser step; the reference
 * gets the same sum from autograd accumulating into .grad).  Per-view outputs (d_mean2D, d_local_rot, d_global_trans,
 * d_mask, d_mlp_weight, d_offsets) are always overwritten. */
int rg_skinned_backward_part(const RgSkinnedScene* scene, const RgWorkspace* ws, int phases, int row_begin, int row_end,
                             const int* radii, const float* alphas, const float* dL_dcolor, const float* dL_ddepth,
                             const float* dL_dalpha, const float* dL_dnodes, float* d_xyz, float* d_scaling_raw,
                             float* d_rotation_raw, float* d_opacity_raw, float* d_features_dc, float* d_features_rest,
                             float* d_mean2D, float* d_local_rot, float* d_global_trans, float* d_node_radius_raw,
                             float* d_mask, float* d_mlp_weight, float* d_offsets, rg_stream_t stream);

/* ---- the per-Gaussian MLP front end of the skinning path, on the tensor cores ---------- */

/* The reference's two per-Gaussian MLPs -- WeightMLP (skeleton_utils/network_utils.py:73-107; called at
 * skeleton_utils/skeleton_warp.py:56-61) and DeformMLP = detail_net (network_utils.py:6-65; skeleton_warp.py:98-104,
 * 152-153) -- share one shape: positional encoding of x (3 + 6 * multires <= 63 values), eight 256-wide ReLU layers, the
 * encoding concatenated in FRONT of the input of layer 5, a linear head (sigmoid for WeightMLP).  The reference runs each
 * as ~20 eager torch kernels per direction with fp32 cuBLAS GEMMs over all N points; here a tile of 128 points goes
 * through all nine layers inside one kernel (tcgen05.mma with the activations in tensor memory, csrc/mlp.cu).  bf16
 * operands, fp32 accumulation.
 *
 * rg_mlp_pack: torch.nn.Linear weights (fp32, (n_out, ld) row-major; weights[0..7] the hidden layers, weights[8] the
 * head) -> the bf16 chunk images the kernels stream.  For layer l, emb_col0[l] is the first column that multiplies the
 * encoding (layers 0 and 5; ignored elsewhere) and act_col0[l] the first of the 256 columns that multiply the previous
 * layer's activations (layers 1..8).  Columns in between (detail_net's pose input, the same for every point) are NOT
 * packed: the caller folds W[:, pose columns] @ pose into that layer's bias.  packed: rg_mlp_packed_bytes(out_ch) bytes.
 * Call it again whenever the weights change. */
size_t rg_mlp_packed_bytes(int out_ch);
int rg_mlp_pack(int emb_cols, int out_ch, const float* const* weights /* host array of 9 device pointers */,
                const int* lds /* host, 9 */, const int* emb_col0 /* host, 9 */, const int* act_col0 /* host, 9 */,
                void* packed, rg_stream_t stream);

/* Forward: x (N,3) -> out (N,out_ch); out_act 0 = linear head (detail_net), 1 = sigmoid (WeightMLP).  bias: (9,256)
 * fp32, the hidden layers' biases then the head's, zero padded.  Training (acts != NULL): also stores the post-ReLU
 * activations of the eight hidden layers as bf16, feature-major (8, 256, n_pad), the encoding (64, n_pad) into emb and one
 * bit per hidden unit ("was it positive") into relu_mask (8 * 8 * n_pad 32-bit words);
 * n_pad >= N, a multiple of 128 (columns of points past N are written as zeros).  These are what rg_mlp_backward reads. */
int rg_mlp_forward(int N, int multires, int out_ch, int out_act, const float* x, const void* packed, const float* bias,
                   float* out, void* acts, void* emb, void* relu_mask, long long n_pad, rg_stream_t stream);

/* Backward of rg_mlp_forward (the reference: torch autograd through the same ~20 ops).  g_out (N,out_ch) is dL/d out;
 * out_fwd the forward's output (needed for the sigmoid head, else NULL); acts / emb the state the forward stored; dz
 * scratch of 9 * 256 * n_pad bf16 (dL/d pre-activation of every layer, written here).  No gradient flows to x (it is
 * detached, skeleton_utils/skeleton_warp.py:131).  d_weights[l] (host array of 9 device pointers, fp32, the shapes and
 * column conventions of rg_mlp_pack) and d_bias (9,256) are ACCUMULATED into with atomics: zero them first.  Columns of
 * a weight that rg_mlp_pack skips (the pose columns of detail_net's layers 0 and 5) get no gradient here: theirs is the
 * outer product d_bias[l] x pose, which the caller forms.
 * Two kernels: the chain run backwards on the tensor cores (dH = dZ W with the transposed weight image, ReLU' from the
 * stored activations), then one tcgen05 GEMM per (layer, half of its outputs, slice of the points) that contracts dZ
 * with the layer's inputs over the points, operands arriving as 2-D TMA tiles. */
int rg_mlp_backward(int N, int out_ch, int out_act, int emb_cols, const float* g_out, const float* out_fwd, const void* packed,
                    const void* acts, const void* emb, const void* relu_mask, void* dz, long long n_pad,
                    float* const* d_weights /* host, 9 */,
                    const int* lds /* host, 9 */, const int* emb_col0 /* host, 9 */, const int* act_col0 /* host, 9 */,
                    float* d_bias, rg_stream_t stream);

/* ---- multi-GPU (no counterpart in the reference, which is single-GPU) -------------------------------- */

/* In-place sum over `world` ranks of a gradient buffer that every rank has mapped at the multicast (NVLS multimem)
 * address `multicast_base` (the host maps it, e.g. torch.distributed._symmetric_memory).  Each rank calls this with
 * its own rank on its own stream, between two cross-rank barriers: all ranks have finished writing the buffer /
 * all ranks' kernels have finished.  n_floats must be a multiple of 4 and the base 16-byte aligned.  Rank r reduces
 * the r-th slice through the switch (multimem.ld_reduce) and broadcasts it to every copy (multimem.st). */
int rg_allreduce_multimem(float* multicast_base, size_t n_floats, int rank, int world, rg_stream_t stream);

/* The same in-place sum over plain peer-to-peer mappings: peer_bufs[r] (host array of `world` device pointers, world
 * in {2, 4, 8}) is rank r's copy of the buffer as mapped into THIS process.  Rank r loads its slice from every copy,
 * adds in rank order (bit-identical results on all ranks) and stores the sum into every copy.  Same barrier contract. */
int rg_allreduce_p2p(float* const* peer_bufs, size_t n_floats, int rank, int world, rg_stream_t stream);

/* Both kernels over several ranges of the buffer in one launch (offsets / lengths in floats, multiples of 4, at most 8
 * ranges): multicast_base != NULL selects the NVLS kernel, else peer_bufs.  `blocks` CTAs of 512 threads (0: one per
 * SM) -- a small grid leaves room for a kernel running beside it.  Same barrier contract. */
int rg_allreduce_ranges(float* const* peer_bufs, float* multicast_base, int n_ranges, const size_t* offsets,
                        const size_t* lengths, int rank, int world, int blocks, rg_stream_t stream);

/* ---- measurement aids (no counterpart in the reference; used by bench.py) ------------------------ */

/* Kernels of this library launched since it was loaded (library sort passes are not counted). */
long long rg_launch_count(void);
/* L1 image loss and its gradient in one pass: *loss = mean |output - gt|, grad = sign(output - gt) / n.
 * Replaces l1_loss of utils/loss_utils.py (torch.abs(network_output - gt).mean()) together with the element-wise
 * kernels autograd runs for its backward.  output, gt, grad: n floats, 16-byte aligned.  scratch:
 * rg_l1_loss_scratch_bytes() bytes of device memory, zero-filled before the first call and then reused (the kernel
 * leaves it ready for the next call); one scratch per stream.  Deterministic for a given n. */
size_t rg_l1_loss_scratch_bytes(void);
int rg_l1_loss(size_t n, const float* output, const float* gt, float* loss, float* grad, void* scratch,
               size_t scratch_bytes, rg_stream_t stream);

/* Per-kernel device timing: when enabled every launch is bracketed by CUDA events on its stream.
 * rg_profile_get synchronises on the recorded events and returns the summed milliseconds per kernel name. */
void rg_profile_enable(int on);
void rg_profile_reset(void);
int rg_profile_count(void);
int rg_profile_get(int index, char* name_out, int name_cap, float* total_ms, int* launches);
/* Per-(tile, warp) cycle counters of the two blend kernels: fwd / bwd point to tiles*8*8 uint32 of device memory
 * (or NULL to switch the counters off): total cycles, prefetch-wait cycles, stage-1 cycles, stage-2 cycles, surviving
 * list entries, groups walked, aux (fwd: list length, bwd: reductions issued), 0. */
int rg_debug_tile_profile(uint32_t* fwd, uint32_t* bwd);

#ifdef __cplusplus
}
#endif
#endif /* RIGGS_B200_H_INCLUDED */
