// extern "C" surface, part 2: the fused skin + rasterize entry points (no host round trip).
#include "../../include/riggs_b200.h"

#include <string>

#include "common.cuh"
#include "kernels.cuh"
#include "skin.cuh"

using namespace rg;

namespace {
char* abase(char* p) { return reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256)); }
bool al16(const void* p) { return (reinterpret_cast<size_t>(p) & 15) == 0; }

struct FkSlots {
    float *G, *node_rot, *posed, *dG, *dradius, *dtrans;
    explicit FkSlots(float* b) {
        G = b; node_rot = G + 128 * 12; posed = node_rot + 128 * 4; dG = posed + 128 * 3; dradius = dG + 128 * 12;
        dtrans = dradius + 128;
    }
};

int check_scene(const RgSkinnedScene* s, const RgWorkspace* w, const char* who) {
    if (!s || !w) { set_error(std::string(who) + ": null scene/workspace"); return -1; }
    if (s->P <= 0 || s->width <= 0 || s->height <= 0) { set_error(std::string(who) + ": empty problem"); return -1; }
    if (s->J < 2 || s->J > 128) { set_error(std::string(who) + ": J must be in [2, 128]"); return -1; }
    if (s->M_rest < 0 || s->M_rest > 15) { set_error(std::string(who) + ": at most 15 rest SH coefficients"); return -1; }
    if (!al16(s->rotation_raw)) { set_error(std::string(who) + ": rotation_raw must be 16-byte aligned"); return -1; }
    if (!w->geom || !w->image || !w->binning) { set_error(std::string(who) + ": workspace buffers missing"); return -1; }
    if (w->geom_bytes < rg_geom_bytes(s->P, s->width, s->height) || w->image_bytes < rg_image_bytes(s->width, s->height)) {
        set_error(std::string(who) + ": workspace geom/image buffer too small");
        return -1;
    }
    if (w->binning_bytes < 1024) { set_error(std::string(who) + ": workspace binning buffer too small"); return -1; }
    if (s->blend_mode != 0 && s->blend_mode != 1) { set_error(std::string(who) + ": blend_mode must be 0 (linear) or 1 (dual quaternion)"); return -1; }
    if (s->split_count < 0 || (s->split_count > 1 && (s->split_index < 0 || s->split_index >= s->split_count))) {
        set_error(std::string(who) + ": split_index must lie in [0, split_count)");
        return -1;
    }
    return 0;
}

SkinFuseArgs fuse_args(const RgSkinnedScene* s, const GeomState& geom, const FkSlots& fk) {
    SkinFuseArgs f = {};
    f.J = s->J; f.M_rest = s->M_rest; f.isotropic = s->isotropic != 0; f.zero_d_rotation = s->zero_d_rotation != 0;
    f.blend_mode = s->blend_mode;
    f.xyz = s->xyz; f.scaling_raw = s->scaling_raw; f.rotation_raw = s->rotation_raw; f.opacity_raw = s->opacity_raw;
    f.features_dc = s->features_dc; f.features_rest = s->features_rest; f.motion_mask = s->motion_mask;
    f.mlp_weight = s->mlp_weight; f.offsets = s->offsets; f.joints = s->joints; f.parents = s->parents;
    f.node_radius_raw = s->node_radius_raw; f.G = fk.G; f.node_rot = fk.node_rot; f.global_trans = s->global_trans;
    f.means3D_out = geom.means3D; f.d_rot_out = geom.d_rot; f.wsum_out = geom.skin_wsum;
    return f;
}
}  // namespace

extern "C" {

uint32_t rg_binning_capacity(size_t binning_bytes) {
    if (binning_bytes < 1024) return 0;
    size_t n = (binning_bytes - 1024) / 5;  // 4 bytes of list + 1 byte of survivor mask per instance
    return n > 0xfffffff0u ? 0xfffffff0u : (uint32_t)n;
}

int rg_skinned_forward(const RgSkinnedScene* s, const RgWorkspace* w, float* out_color, float* out_depth,
                       float* out_alpha, int* radii, float* d_nodes, float* dbg_scales, float* dbg_rotations,
                       float* dbg_opacities, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (check_scene(s, w, "rg_skinned_forward")) return -1;
    const int P = s->P, W = s->width, H = s->height;
    const int grid_x = (W + RG_TILE - 1) / RG_TILE, grid_y = (H + RG_TILE - 1) / RG_TILE;
    GeomState geom = GeomState::carve(abase(w->geom), P, W, H, nullptr);
    ImageState img = ImageState::carve(abase(w->image), W, H, nullptr);
    const uint32_t cap = rg_binning_capacity(w->binning_bytes);
    BinningState bin = BinningState::carve(abase(w->binning), cap, nullptr);
    FkSlots fk(geom.fk);
    const BinPlan bp = make_bin_plan(P, grid_x, grid_y);

    // The forward's three zero-fills run on a side stream (a parallel branch of a captured graph) instead of between
    // the kernels of the chain: the sort's control words must be clean before the preprocess kernel counts digits into
    // them, the tile totals before the column sums, the survivor bitmaps before the blend.
    SideStream* side = side_stream();
    PreprocessArgs pa = {};
    if (side) {
        cudaEventRecord(side->fork, stream);
        cudaStreamWaitEvent(side->stream, side->fork, 0);
        pa.sort_hist = depth_sort_prepare(geom.sort_temp, P, side->stream);
        cudaEventRecord(side->join_a, side->stream);
        cudaMemsetAsync(geom.tile_base, 0, sizeof(uint32_t) * bp.T, side->stream);
        cudaMemsetAsync(bin.surv_mask, 0, bin.surv_mask_bytes, side->stream);
        cudaEventRecord(side->join_b, side->stream);
    }

    if (launch_fk_forward(1, s->J, s->local_rot, s->global_trans, s->joints, s->parents, fk.G, fk.node_rot,
                          d_nodes ? d_nodes : fk.posed, stream))
        return -1;

    pa.P = P; pa.D = s->sh_degree; pa.M = s->M_rest + 1;
    pa.scale_modifier = s->scale_modifier;
    pa.viewmatrix = s->viewmatrix; pa.projmatrix = s->projmatrix; pa.cam_pos = s->campos;
    pa.W = W; pa.H = H; pa.tan_fovx = s->tan_fovx; pa.tan_fovy = s->tan_fovy;
    pa.focal_y = H / (2.0f * s->tan_fovy); pa.focal_x = W / (2.0f * s->tan_fovx);
    pa.grid_x = grid_x; pa.grid_y = grid_y; pa.prefiltered = false; pa.sh_vec_ok = false;
    pa.radii = radii; pa.means2D = geom.means2D; pa.depths = geom.depths; pa.rec = geom.rec;
    pa.conic_opacity = geom.conic_opacity; pa.tiles_touched = geom.tiles_touched; pa.clamped = geom.clamped;
    pa.ids = geom.ids;
    SkinFuseArgs f = fuse_args(s, geom, fk);
    f.dbg_scales = dbg_scales; f.dbg_rotations = dbg_rotations; f.dbg_opacities = dbg_opacities;
    if (dbg_rotations && !al16(dbg_rotations)) { set_error("rg_skinned_forward: dbg_rotations must be 16-byte aligned"); return -1; }
    if (side) cudaStreamWaitEvent(stream, side->join_a, 0);
    else pa.sort_hist = depth_sort_prepare(geom.sort_temp, P, stream);  // the preprocess kernel counts the sort's digits
    if (launch_skin_preprocess(pa, f, stream)) return -1;

    if (launch_depth_sort(geom.sort_temp, geom.sort_temp_bytes, geom.depths, geom.sorted_keys, geom.ids, geom.sorted_ids, P, stream,
                          pa.sort_hist != nullptr))
        return -1;
    const int split_n = s->split_count > 1 ? s->split_count : 1, split_i = split_n > 1 ? s->split_index : 0;
    if (launch_tile_count(bp, P, grid_x, grid_y, geom.sorted_ids, geom.tiles_touched, geom.means2D, radii, geom.rects, geom.M, geom.Mg, stream,
                          split_i, split_n))
        return -1;
    // nobody reads num_rendered on the host here, so the prefix over the tiles rides along with the column scans
    if (side) cudaStreamWaitEvent(stream, side->join_b, 0);
    launch_tile_offsets_a(bp, geom.Mg, geom.S, geom.tile_base, img.ranges, geom.tile_order, geom.status, cap, stream, false, side != nullptr);
    launch_tile_scatter(bp, P, grid_x, grid_y, geom.sorted_ids, geom.rects, geom.M, geom.Mg, geom.S,
                        geom.tile_base, geom.E, bin.point_list, cap, stream, true, img.ranges, geom.tile_order, geom.status, split_n);

    BlendFwdArgs fa;
    fa.W = W; fa.H = H; fa.grid_x = grid_x; fa.grid_y = grid_y; fa.split_i = split_i; fa.split_n = split_n;
    if (!side) cudaMemsetAsync(bin.surv_mask, 0, bin.surv_mask_bytes, stream);
    fa.ranges = img.ranges; fa.point_list = bin.point_list; fa.surv_mask = bin.surv_mask; fa.surv_stride = bin.surv_stride; fa.rec = geom.rec; fa.tile_order = geom.tile_order;
    fa.bg = s->background;
    fa.out_color = out_color; fa.out_depth = out_depth; fa.out_alpha = out_alpha; fa.n_contrib = img.n_contrib;
    launch_blend_forward(fa, stream);
    return check_launch("rg_skinned_forward", stream, false);
}

int rg_skinned_status(const RgSkinnedScene* s, const RgWorkspace* w, uint32_t* host_out2, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!s || !w || !host_out2) { set_error("rg_skinned_status: null argument"); return -1; }
    GeomState geom = GeomState::carve(abase(w->geom), s->P, s->width, s->height, nullptr);
    cudaError_t e = cudaMemcpyAsync(host_out2, geom.status, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) { set_error(std::string("rg_skinned_status: ") + cudaGetErrorString(e)); return -1; }
    return 0;
}

// phases: 1 = zero the accumulators + blend backward, 2 = per-Gaussian backward of rows [row_begin, row_end),
// 4 = skeleton backward, 8 = ADD into the parameter gradients (d_xyz .. d_features_rest, d_node_radius_raw) instead of
// overwriting them: the second and later views of a multi-view batch.  rg_skinned_backward runs 1|2|4 over all rows; a data-parallel caller runs 1, then 2 chunk by
// chunk (all-reducing each chunk's finished gradient rows on another stream meanwhile), then 4.
int rg_skinned_backward_part(const RgSkinnedScene* s, const RgWorkspace* w, int phases, int row_begin, int row_end,
                             const int* radii, const float* alphas, const float* dL_dcolor, const float* dL_ddepth,
                             const float* dL_dalpha, const float* dL_dnodes, float* d_xyz, float* d_scaling_raw,
                             float* d_rotation_raw, float* d_opacity_raw, float* d_features_dc, float* d_features_rest,
                             float* d_mean2D, float* d_local_rot, float* d_global_trans, float* d_node_radius_raw,
                             float* d_mask, float* d_mlp_weight, float* d_offsets, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (check_scene(s, w, "rg_skinned_backward")) return -1;
    if (!al16(d_rotation_raw)) { set_error("rg_skinned_backward: d_rotation_raw must be 16-byte aligned"); return -1; }
    const int P = s->P, W = s->width, H = s->height;
    if ((phases & 2) && (row_begin < 0 || row_end > P || row_begin > row_end || (row_begin & 255))) {
        set_error("rg_skinned_backward_part: rows must lie in [0, P] and row_begin must be a multiple of 256");
        return -1;
    }
    const int grid_x = (W + RG_TILE - 1) / RG_TILE, grid_y = (H + RG_TILE - 1) / RG_TILE;
    GeomState geom = GeomState::carve(abase(w->geom), P, W, H, nullptr);
    ImageState img = ImageState::carve(abase(w->image), W, H, nullptr);
    const uint32_t cap = rg_binning_capacity(w->binning_bytes);
    BinningState bin = BinningState::carve(abase(w->binning), cap, nullptr);
    FkSlots fk(geom.fk);

    if (phases & 1) {
        cudaMemsetAsync(geom.grad_acc, 0, sizeof(float) * 10 * (size_t)P, stream);
        cudaMemsetAsync(fk.dG, 0, sizeof(float) * (128 * 12 + 128 + 4), stream);

        BlendBwdArgs ba;
        ba.P = P; ba.W = W; ba.H = H; ba.grid_x = grid_x; ba.grid_y = grid_y;
        ba.split_n = s->split_count > 1 ? s->split_count : 1; ba.split_i = ba.split_n > 1 ? s->split_index : 0;
        ba.ranges = img.ranges; ba.point_list = bin.point_list; ba.surv_mask = bin.surv_mask; ba.surv_stride = bin.surv_stride;
        ba.bg = s->background;
        ba.rec = geom.rec; ba.tile_order = geom.tile_order; ba.alphas = alphas;
        ba.n_contrib = img.n_contrib; ba.dL_dpix = dL_dcolor; ba.dL_dpix_depth = dL_ddepth; ba.dL_dalphas = dL_dalpha;
        ba.grad_acc = geom.grad_acc;
        launch_blend_backward(ba, stream);
    }
    if ((phases & 2) && row_end > row_begin) {
        GaussBwdArgs ga = {};
        ga.P = P; ga.row_begin = row_begin; ga.row_end = row_end;
        ga.D = s->sh_degree; ga.M = s->M_rest + 1; ga.means3D = geom.means3D; ga.radii = radii;
        ga.clamped = geom.clamped; ga.scale_modifier = s->scale_modifier;
        ga.viewmatrix = s->viewmatrix; ga.projmatrix = s->projmatrix; ga.campos = s->campos; ga.W = W; ga.H = H;
        ga.tan_fovx = s->tan_fovx; ga.tan_fovy = s->tan_fovy;
        ga.focal_y = H / (2.0f * s->tan_fovy); ga.focal_x = W / (2.0f * s->tan_fovx);
        ga.grad_acc = geom.grad_acc; ga.dL_dmean2D = d_mean2D;
        SkinFuseBwdArgs sb = {};
        sb.f = fuse_args(s, geom, fk);
        sb.d_rot_saved = geom.d_rot;
        sb.wsum_saved = geom.skin_wsum;
        sb.d_xyz = d_xyz; sb.d_scaling_raw = d_scaling_raw; sb.d_rotation_raw = d_rotation_raw; sb.d_opacity_raw = d_opacity_raw;
        sb.d_features_dc = d_features_dc; sb.d_features_rest = d_features_rest;
        sb.dG = fk.dG; sb.dnode_radius_raw = fk.dradius; sb.dglobal_trans = fk.dtrans;
        sb.dmask = d_mask; sb.dmlp_weight = d_mlp_weight; sb.doffsets = d_offsets;
        sb.accumulate = (phases & 8) != 0;
        if (launch_gauss_skin_backward(ga, sb, stream)) return -1;
    }
    if (phases & 4) {
        // d_node_radius_raw leaves the workspace's accumulator inside the same launch: a copy or, for the later views of a batch, a sum
        if (launch_fk_backward(1, s->J, s->local_rot, s->joints, s->parents, fk.G, fk.dG, dL_dnodes, d_local_rot,
                               d_global_trans, fk.dtrans, stream, d_node_radius_raw, fk.dradius, (phases & 8) != 0))
            return -1;
    }
    return check_launch("rg_skinned_backward", stream, false);
}

int rg_skinned_backward(const RgSkinnedScene* s, const RgWorkspace* w, const int* radii, const float* alphas,
                        const float* dL_dcolor, const float* dL_ddepth, const float* dL_dalpha, const float* dL_dnodes,
                        float* d_xyz, float* d_scaling_raw, float* d_rotation_raw, float* d_opacity_raw,
                        float* d_features_dc, float* d_features_rest, float* d_mean2D, float* d_local_rot,
                        float* d_global_trans, float* d_node_radius_raw, float* d_mask, float* d_mlp_weight,
                        float* d_offsets, rg_stream_t stream) {
    return rg_skinned_backward_part(s, w, 7, 0, s ? s->P : 0, radii, alphas, dL_dcolor, dL_ddepth, dL_dalpha, dL_dnodes, d_xyz,
                                    d_scaling_raw, d_rotation_raw, d_opacity_raw, d_features_dc, d_features_rest, d_mean2D,
                                    d_local_rot, d_global_trans, d_node_radius_raw, d_mask, d_mlp_weight, d_offsets, stream);
}

}  // extern "C"
