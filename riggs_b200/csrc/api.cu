// extern "C" surface of libriggs_b200.so (declared in include/riggs_b200.h).
#include "../../include/riggs_b200.h"

#include <mutex>
#include <string>

#include "common.cuh"
#include "kernels.cuh"
#include "knn.cuh"
#include "skin.cuh"

namespace rg {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

int check_launch(const char* what, cudaStream_t stream, bool debug) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && debug) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        set_error(std::string(what) + ": " + cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

GeomState GeomState::carve(char* base, size_t P, int W, int H, size_t* bytes) {
    Carver c(base);
    GeomState g;
    const BinPlan bp = make_bin_plan((int)P, (W + RG_TILE - 1) / RG_TILE, (H + RG_TILE - 1) / RG_TILE);
    g.depths = c.take<float>(P);
    g.clamped = c.take<uint8_t>(P);
    g.means2D = c.take<float2>(P);
    g.conic_opacity = c.take<float4>(P);
    g.rec = c.take<float4>(4 * P);
    g.tiles_touched = c.take<uint32_t>(P);
    g.means3D = c.take<float>(3 * P);
    g.ids = c.take<uint32_t>(P);
    g.sorted_keys = c.take<uint32_t>(P);
    g.sorted_ids = c.take<uint32_t>(P);
    g.rects = c.take<uint2>(P);
    g.sort_temp_bytes = depth_sort_temp_bytes(P);
    g.sort_temp = c.take<char>(g.sort_temp_bytes);
    g.M = c.take<uint16_t>((size_t)bp.NW * bp.Tpad);
    g.Mg = c.take<uint16_t>((size_t)bp.NG * bp.Tpad);
    g.E = c.take<uint32_t>((size_t)bp.NG * bp.T);
    g.S = c.take<uint32_t>((size_t)bp.nseg * bp.T);
    g.tile_base = c.take<uint32_t>(bp.T + 1);
    g.tile_order = c.take<uint32_t>(bp.T);
    g.status = c.take<uint32_t>(8);
    g.grad_acc = c.take<float>(10 * P);
    g.d_rot = c.take<float>(4 * P);
    g.skin_wsum = c.take<float>(P);
    g.fk = c.take<float>(128 * (12 + 4 + 3 + 12 + 1) + 4);
    if (bytes) *bytes = align_up(c.off, 256) + 256;
    return g;
}

ImageState ImageState::carve(char* base, int W, int H, size_t* bytes) {
    Carver c(base);
    ImageState s;
    const size_t tiles = (size_t)((W + RG_TILE - 1) / RG_TILE) * ((H + RG_TILE - 1) / RG_TILE);
    s.n_contrib = c.take<uint32_t>((size_t)W * H);
    s.ranges = c.take<uint2>(tiles);
    if (bytes) *bytes = align_up(c.off, 256) + 256;
    return s;
}

BinningState BinningState::carve(char* base, size_t R, size_t* bytes) {
    Carver c(base);
    BinningState b;
    b.point_list = c.take<uint32_t>(R);
    b.surv_stride = (R + 31) / 32 + 2;
    b.surv_mask_bytes = 8 * b.surv_stride * sizeof(uint32_t);
    b.surv_mask = c.take<uint32_t>(b.surv_mask_bytes / 4);
    if (bytes) *bytes = align_up(c.off, 256) + 256;
    return b;
}

// The buffers handed over by torch are 512-byte aligned; be strict so carve() offsets stay valid.
static char* aligned_base(char* p) { return reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256)); }

// Pinned landing slot for the 4-byte num_rendered read-back, one per host thread.
static uint32_t* pinned_slot() {
    static thread_local uint32_t* slot = nullptr;
    if (!slot) {
        if (cudaHostAlloc(&slot, 64, cudaHostAllocDefault) != cudaSuccess) slot = nullptr;
    }
    return slot;
}

static bool is_aligned16(const void* p) { return (reinterpret_cast<size_t>(p) & 15) == 0; }

}  // namespace rg

using namespace rg;

extern "C" {

int rg_version(void) { return 1; }
const char* rg_last_error(void) { return g_error.c_str(); }

size_t rg_geom_bytes(int P, int width, int height) { size_t b; GeomState::carve(nullptr, (size_t)P, width, height, &b); return b; }
size_t rg_image_bytes(int width, int height) { size_t b; ImageState::carve(nullptr, width, height, &b); return b; }
size_t rg_binning_bytes(int64_t R) { size_t b; BinningState::carve(nullptr, (size_t)(R > 0 ? R : 0), &b); return b; }

int rg_rasterize_forward(rg_alloc_fn geom_alloc, void* geom_ctx, rg_alloc_fn binning_alloc, void* binning_ctx,
                         rg_alloc_fn image_alloc, void* image_ctx, int P, int D, int M, const float* background,
                         int width, int height, const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                         const float* cam_pos, float tan_fovx, float tan_fovy, int prefiltered, float* out_color,
                         float* out_depth, float* out_alpha, int* radii, int debug, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool dbg = debug != 0;
    if (P <= 0 || width <= 0 || height <= 0) { set_error("rg_rasterize_forward: empty problem"); return -1; }
    if (!shs && !colors_precomp) { set_error("rg_rasterize_forward: need shs or colors_precomp"); return -1; }
    if (!cov3D_precomp && (!scales || !rotations)) { set_error("rg_rasterize_forward: need scales+rotations or cov3D_precomp"); return -1; }
    if (shs && M > 16) { set_error("rg_rasterize_forward: at most 16 SH coefficients per channel"); return -1; }
    if (rotations && !is_aligned16(rotations)) { set_error("rg_rasterize_forward: rotations must be 16-byte aligned"); return -1; }
    const float focal_y = height / (2.0f * tan_fovy);
    const float focal_x = width / (2.0f * tan_fovx);
    const int grid_x = (width + RG_TILE - 1) / RG_TILE, grid_y = (height + RG_TILE - 1) / RG_TILE;

    char* gptr = geom_alloc(geom_ctx, rg_geom_bytes(P, width, height));
    char* iptr = image_alloc(image_ctx, rg_image_bytes(width, height));
    if (!gptr || !iptr) { set_error("rg_rasterize_forward: state buffer allocation failed"); return -1; }
    GeomState geom = GeomState::carve(aligned_base(gptr), P, width, height, nullptr);
    ImageState img = ImageState::carve(aligned_base(iptr), width, height, nullptr);
    const BinPlan bp = make_bin_plan(P, grid_x, grid_y);

    PreprocessArgs pa;
    pa.P = P; pa.D = D; pa.M = M;
    pa.means3D = means3D; pa.scales = scales; pa.scale_modifier = scale_modifier; pa.rotations = rotations;
    pa.opacities = opacities; pa.shs = shs; pa.cov3D_precomp = cov3D_precomp; pa.colors_precomp = colors_precomp;
    pa.viewmatrix = viewmatrix; pa.projmatrix = projmatrix; pa.cam_pos = cam_pos;
    pa.W = width; pa.H = height; pa.tan_fovx = tan_fovx; pa.tan_fovy = tan_fovy; pa.focal_x = focal_x; pa.focal_y = focal_y;
    pa.grid_x = grid_x; pa.grid_y = grid_y; pa.prefiltered = prefiltered != 0;
    pa.sh_vec_ok = shs && M == 16 && is_aligned16(shs);
    pa.radii = radii; pa.means2D = geom.means2D; pa.depths = geom.depths; pa.cov3D_out = nullptr;
    pa.rec = geom.rec; pa.conic_opacity = geom.conic_opacity; pa.tiles_touched = geom.tiles_touched;
    pa.clamped = geom.clamped; pa.ids = geom.ids;
    launch_preprocess(pa, stream);
    if (check_launch("preprocess", stream, dbg)) return -1;

    if (launch_depth_sort(geom.sort_temp, geom.sort_temp_bytes, geom.depths, geom.sorted_keys, geom.ids, geom.sorted_ids,
                          P, stream))
        return -1;
    if (check_launch("depth_sort", stream, dbg)) return -1;
    if (launch_tile_count(bp, P, grid_x, grid_y, geom.sorted_ids, geom.tiles_touched, geom.means2D, radii, geom.rects, geom.M, geom.Mg, stream))
        return -1;
    launch_tile_offsets_a(bp, geom.Mg, geom.S, geom.tile_base, img.ranges, geom.tile_order, geom.status, 0xffffffffu, stream);
    if (check_launch("tile_count/offsets", stream, dbg)) return -1;

    // The one host round trip of the forward (the reference does the same, rasterizer_impl.cu:282): the
    // binning buffer is sized by the caller's allocator from num_rendered.  rg_skinned_forward has no such trip.
    uint32_t* slot = pinned_slot();
    if (!slot) { set_error("rg_rasterize_forward: pinned host allocation failed"); return -1; }
    cudaMemcpyAsync(slot, geom.status, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) { set_error(std::string("num_rendered read-back: ") + cudaGetErrorString(e)); return -1; }
    const int64_t R = *slot;
    if (R > 0x7fffffff) { set_error("rg_rasterize_forward: more than 2^31 tile instances"); return -1; }

    char* bptr = binning_alloc(binning_ctx, rg_binning_bytes(R));
    if (!bptr) { set_error("rg_rasterize_forward: binning buffer allocation failed"); return -1; }
    BinningState bin = BinningState::carve(aligned_base(bptr), (size_t)R, nullptr);
    if (R > 0) {
        launch_tile_scatter(bp, P, grid_x, grid_y, geom.sorted_ids, geom.rects, geom.M, geom.Mg,
                            geom.S, geom.tile_base, geom.E, bin.point_list, (uint32_t)R, stream);
        if (check_launch("tile_scatter", stream, dbg)) return -1;
    }

    BlendFwdArgs fa;
    fa.W = width; fa.H = height; fa.grid_x = grid_x; fa.grid_y = grid_y;
    cudaMemsetAsync(bin.surv_mask, 0, bin.surv_mask_bytes, stream);
    fa.ranges = img.ranges; fa.point_list = bin.point_list; fa.surv_mask = bin.surv_mask; fa.surv_stride = bin.surv_stride;
    fa.rec = geom.rec; fa.bg = background; fa.tile_order = geom.tile_order;
    fa.out_color = out_color; fa.out_depth = out_depth; fa.out_alpha = out_alpha; fa.n_contrib = img.n_contrib;
    launch_blend_forward(fa, stream);
    if (check_launch("blend_forward", stream, dbg)) return -1;
    return (int)R;
}

int rg_rasterize_backward(int P, int D, int M, int R, const float* background, int width, int height,
                          const float* means3D, const float* shs, const float* colors_precomp, const float* alphas,
                          const float* scales, float scale_modifier, const float* rotations,
                          const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tan_fovx, float tan_fovy, const int* radii, char* geom_buffer,
                          char* binning_buffer, char* image_buffer, const float* dL_dpix,
                          const float* dL_dpix_depth, const float* dL_dalphas, float* dL_dmean2D, float* dL_dconic,
                          float* dL_dopacity, float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                          float* dL_dsh, float* dL_dscale, float* dL_drot, int debug, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const bool dbg = debug != 0;
    if (P <= 0) { set_error("rg_rasterize_backward: empty problem"); return -1; }
    if (shs && !dL_dsh) { set_error("rg_rasterize_backward: dL_dsh required when shs is given"); return -1; }
    if (scales && (!dL_dscale || !dL_drot || !rotations)) { set_error("rg_rasterize_backward: dL_dscale/dL_drot required when scales are given"); return -1; }
    if (!scales && !cov3D_precomp) { set_error("rg_rasterize_backward: need scales+rotations or cov3D_precomp"); return -1; }
    if (!is_aligned16(dL_dconic) || (dL_drot && !is_aligned16(dL_drot)) || (rotations && !is_aligned16(rotations))) {
        set_error("rg_rasterize_backward: dL_dconic, dL_drot and rotations must be 16-byte aligned");
        return -1;
    }
    const float focal_y = height / (2.0f * tan_fovy);
    const float focal_x = width / (2.0f * tan_fovx);
    const int grid_x = (width + RG_TILE - 1) / RG_TILE, grid_y = (height + RG_TILE - 1) / RG_TILE;
    GeomState geom = GeomState::carve(aligned_base(geom_buffer), P, width, height, nullptr);
    ImageState img = ImageState::carve(aligned_base(image_buffer), width, height, nullptr);
    BinningState bin = BinningState::carve(aligned_base(binning_buffer), (size_t)(R > 0 ? R : 0), nullptr);

    float* acc = geom.grad_acc;
    cudaMemsetAsync(acc, 0, sizeof(float) * 10 * (size_t)P, stream);
    BlendBwdArgs ba;
    ba.P = P; ba.W = width; ba.H = height; ba.grid_x = grid_x; ba.grid_y = grid_y;
    ba.ranges = img.ranges; ba.point_list = bin.point_list; ba.surv_mask = bin.surv_mask; ba.surv_stride = bin.surv_stride; ba.bg = background;
    ba.rec = geom.rec; ba.tile_order = geom.tile_order;
    ba.alphas = alphas; ba.n_contrib = img.n_contrib; ba.dL_dpix = dL_dpix;
    ba.dL_dpix_depth = dL_dpix_depth; ba.dL_dalphas = dL_dalphas; ba.grad_acc = acc;
    if (R > 0) launch_blend_backward(ba, stream);
    if (check_launch("blend_backward", stream, dbg)) return -1;

    GaussBwdArgs ga;
    ga.row_begin = 0; ga.row_end = 0;
    ga.P = P; ga.D = D; ga.M = M; ga.means3D = means3D; ga.radii = radii; ga.shs = shs; ga.clamped = geom.clamped;
    ga.scales = scales; ga.rotations = rotations; ga.scale_modifier = scale_modifier;
    ga.cov3D = cov3D_precomp;  // with scales/rotations the covariance is recomputed in registers
    ga.viewmatrix = viewmatrix; ga.projmatrix = projmatrix; ga.campos = campos; ga.W = width; ga.H = height;
    ga.tan_fovx = tan_fovx; ga.tan_fovy = tan_fovy; ga.focal_x = focal_x; ga.focal_y = focal_y;
    ga.sh_vec_ok = shs && M == 16 && is_aligned16(shs) && is_aligned16(dL_dsh);
    ga.grad_acc = acc;
    ga.dL_dmean2D = dL_dmean2D; ga.dL_dconic = dL_dconic; ga.dL_dopacity = dL_dopacity; ga.dL_dcolor = dL_dcolor;
    ga.dL_ddepth = dL_ddepth; ga.dL_dmean3D = dL_dmean3D; ga.dL_dcov3D = dL_dcov3D; ga.dL_dsh = dL_dsh;
    ga.dL_dscale = scales ? dL_dscale : nullptr; ga.dL_drot = scales ? dL_drot : nullptr;
    launch_gauss_backward(ga, stream);
    if (check_launch("gauss_backward", stream, dbg)) return -1;
    return 0;
}

int rg_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                    rg_stream_t stream_) {
    (void)projmatrix;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_mark_visible(P, means3D, viewmatrix, present, stream);
    return check_launch("mark_visible", stream, false);
}

int rg_state_ptrs(int P, int64_t R, int width, int height, char* geom_buffer, char* binning_buffer, char* image_buffer,
                  void** out) {
    GeomState g = GeomState::carve(aligned_base(geom_buffer), P, width, height, nullptr);
    out[0] = g.depths; out[1] = g.clamped; out[2] = g.means2D; out[3] = g.conic_opacity; out[4] = g.rec;
    out[5] = g.tiles_touched; out[6] = g.sorted_ids; out[7] = g.means3D;
    if (binning_buffer) {
        BinningState b = BinningState::carve(aligned_base(binning_buffer), (size_t)(R > 0 ? R : 0), nullptr);
        out[8] = b.point_list;
    } else {
        out[8] = nullptr;
    }
    ImageState im = ImageState::carve(aligned_base(image_buffer), width, height, nullptr);
    out[9] = im.n_contrib; out[10] = im.ranges; out[11] = g.status; out[12] = g.grad_acc;
    return 0;
}

int rg_debug_sorted_keys(int P, int64_t R, int width, int height, char* geom_buffer, char* binning_buffer,
                         char* image_buffer, uint64_t* keys_out, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (R <= 0) return 0;
    GeomState g = GeomState::carve(aligned_base(geom_buffer), P, width, height, nullptr);
    BinningState b = BinningState::carve(aligned_base(binning_buffer), (size_t)R, nullptr);
    ImageState im = ImageState::carve(aligned_base(image_buffer), width, height, nullptr);
    const int T = ((width + RG_TILE - 1) / RG_TILE) * ((height + RG_TILE - 1) / RG_TILE);
    launch_rebuild_keys(T, im.ranges, b.point_list, g.depths, keys_out, stream);
    return check_launch("debug_sorted_keys", stream, false);
}

size_t rg_knn_scratch_bytes(int P) { return knn_scratch_bytes(P) + 256; }

int rg_dist_cuda2(int P, const float* points, float* mean_dists, char* scratch, size_t scratch_bytes,
                  rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (P <= 0) return 0;
    char* base = aligned_base(scratch);
    size_t lost = (size_t)(base - scratch);
    if (!scratch || scratch_bytes < lost) { set_error("rg_dist_cuda2: scratch buffer missing"); return -1; }
    if (launch_dist2_knn3(P, points, mean_dists, base, scratch_bytes - lost, stream)) return -1;
    return check_launch("dist_cuda2", stream, false);
}

int rg_fk_forward(int F, int J, const float* local_rot, const float* global_trans, const float* joints,
                  const int* parents, float* G, float* node_rot, float* posed, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (launch_fk_forward(F, J, local_rot, global_trans, joints, parents, G, node_rot, posed, stream)) return -1;
    return check_launch("fk_forward", stream, false);
}

int rg_fk_backward(int F, int J, const float* local_rot, const float* joints, const int* parents, const float* G,
                   const float* dG, const float* dposed, float* dlocal_rot, float* dglobal_trans, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (launch_fk_backward(F, J, local_rot, joints, parents, G, dG, dposed, dlocal_rot, dglobal_trans, nullptr, stream))
        return -1;
    return check_launch("fk_backward", stream, false);
}

int rg_skin_forward(int N, int J, const float* xyz, const float* joints, const int* parents,
                    const float* node_radius_raw, const float* G, const float* node_rot, const float* global_trans,
                    const float* motion_mask, const float* mlp_weight, const float* offsets, int blend_mode,
                    float* d_xyz, float* d_rot, float* nn_weight, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!is_aligned16(d_rot)) { set_error("rg_skin_forward: d_rot must be 16-byte aligned"); return -1; }
    SkinArgs a;
    a.N = N; a.J = J; a.xyz = xyz; a.joints = joints; a.parents = parents; a.node_radius_raw = node_radius_raw;
    a.G = G; a.node_rot = node_rot; a.global_trans = global_trans; a.motion_mask = motion_mask;
    a.mlp_weight = mlp_weight; a.offsets = offsets; a.blend_mode = blend_mode; a.d_xyz = d_xyz; a.d_rot = d_rot;
    a.nn_weight = nn_weight;
    if (launch_skin_forward(a, stream)) return -1;
    return check_launch("skin_forward", stream, false);
}

int rg_skin_backward_blend(int N, int J, const float* xyz, const float* joints, const int* parents,
                           const float* node_radius_raw, const float* G, const float* node_rot, const float* global_trans,
                           const float* motion_mask, const float* mlp_weight, const float* offsets, int blend_mode,
                           const float* g_dxyz, const float* g_drot, float* dG, float* dnode_radius_raw, float* dglobal_trans,
                           float* dmask, float* dmlp_weight, float* doffsets, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (g_drot && !is_aligned16(g_drot)) { set_error("rg_skin_backward: g_drot must be 16-byte aligned"); return -1; }
    if (blend_mode != 0 && blend_mode != 1) { set_error("rg_skin_backward: blend_mode must be 0 (linear) or 1 (dual quaternion)"); return -1; }
    SkinBwdArgs a;
    a.N = N; a.J = J; a.xyz = xyz; a.joints = joints; a.parents = parents; a.node_radius_raw = node_radius_raw;
    a.G = G; a.node_rot = node_rot; a.global_trans = global_trans; a.motion_mask = motion_mask;
    a.mlp_weight = mlp_weight; a.offsets = offsets; a.g_dxyz = g_dxyz; a.g_drot = g_drot; a.dG = dG;
    a.blend_mode = blend_mode;
    a.dnode_radius_raw = dnode_radius_raw; a.dglobal_trans = dglobal_trans; a.dmask = dmask;
    a.dmlp_weight = dmlp_weight; a.doffsets = doffsets;
    if (launch_skin_backward(a, stream)) return -1;
    return check_launch("skin_backward", stream, false);
}

int rg_skin_backward(int N, int J, const float* xyz, const float* joints, const int* parents,
                     const float* node_radius_raw, const float* G, const float* node_rot, const float* global_trans,
                     const float* motion_mask, const float* mlp_weight, const float* offsets, const float* g_dxyz,
                     const float* g_drot, float* dG, float* dnode_radius_raw, float* dglobal_trans, float* dmask,
                     float* dmlp_weight, float* doffsets, rg_stream_t stream) {
    return rg_skin_backward_blend(N, J, xyz, joints, parents, node_radius_raw, G, node_rot, global_trans, motion_mask,
                                  mlp_weight, offsets, 0, g_dxyz, g_drot, dG, dnode_radius_raw, dglobal_trans, dmask,
                                  dmlp_weight, doffsets, stream);
}

}  // extern "C"
