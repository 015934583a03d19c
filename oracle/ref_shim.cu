/*
 * ref_shim.cu -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A thin extern "C" face over the UNMODIFIED reference CUDA sources, so the tests and the
 * "reference CUDA on one B200" baseline can drive the reference's own kernels without its torch
 * binding.  oracle/build_ref.py compiles the reference translation units straight from
 * /root/reference/submodules/{diff-gaussian-rasterization,simple-knn} (nothing is copied into this
 * repository) and links them with this file into oracle/_ref/libref_riggs.so.
 *
 * It plays the role of dgr/rasterize_points.cu:35-229 (allocate the three growable state buffers,
 * call CudaRasterizer::Rasterizer::{forward,backward,markVisible}) and knn/spatial.cu:15-26, with
 * cudaMalloc'd buffers instead of torch tensors, and additionally exposes the addresses of the
 * state arrays inside those opaque buffers (layouts of dgr/cuda_rasterizer/rasterizer_impl.cu:155-193)
 * so keys / sorted ids / ranges / n_contrib can be compared bit for bit.
 *
 * The reference launches on the legacy default stream; callers must be on it too (torch's default
 * stream is).
 */
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <stdexcept>

#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"

namespace {
struct Buf {
    char* p = nullptr;
    size_t cap = 0;
    size_t used = 0;
    char* get(size_t n) {
        used = n;
        if (n > cap) {
            if (p) cudaFree(p);
            size_t want = n + n / 4 + 256;
            if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; cap = 0; throw std::runtime_error("ref_shim: cudaMalloc failed"); }
            cap = want;
        }
        return p;
    }
    ~Buf() { if (p) cudaFree(p); }
};
}  // namespace

struct RefState {
    Buf geom, binning, img;
    int P = 0, R = 0, W = 0, H = 0;
};

extern "C" {

RefState* ref_state_new() { return new RefState(); }
void ref_state_free(RefState* s) { delete s; }

int ref_forward(RefState* s, int P, int D, int M, const float* bg, int W, int H, const float* means3D,
                const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                float tan_fovy, int prefiltered, float* out_color, float* out_depth, float* out_alpha,
                int* radii, int debug) {
    try {
        s->P = P; s->W = W; s->H = H;
        std::function<char*(size_t)> g = [s](size_t n) { return s->geom.get(n); };
        std::function<char*(size_t)> b = [s](size_t n) { return s->binning.get(n); };
        std::function<char*(size_t)> i = [s](size_t n) { return s->img.get(n); };
        s->R = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, bg, W, H, means3D, shs, colors_precomp,
                                                   opacities, scales, scale_modifier, rotations, cov3D_precomp,
                                                   viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                                                   prefiltered != 0, out_color, out_depth, out_alpha, radii,
                                                   debug != 0);
        return s->R;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_forward: %s\n", e.what());
        return -1;
    }
}

int ref_backward(RefState* s, int P, int D, int M, const float* bg, int W, int H, const float* means3D,
                 const float* shs, const float* colors_precomp, const float* alphas, const float* scales,
                 float scale_modifier, const float* rotations, const float* cov3D_precomp,
                 const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                 float tan_fovy, const int* radii, const float* dL_dpix, const float* dL_dpix_depth,
                 const float* dL_dalphas, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscale, float* dL_drot, int debug) {
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, s->R, bg, W, H, means3D, shs, colors_precomp, alphas, scales,
                                             scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
                                             campos, tan_fovx, tan_fovy, radii, s->geom.p, s->binning.p, s->img.p,
                                             dL_dpix, dL_dpix_depth, dL_dalphas, dL_dmean2D, dL_dconic, dL_dopacity,
                                             dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
                                             dL_drot, debug != 0);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_backward: %s\n", e.what());
        return -1;
    }
}

/* Addresses of the reference's internal arrays for the last forward call.
 * out[0..9]  geom: depths, clamped, internal_radii, means2D, cov3D, conic_opacity, rgb, tiles_touched,
 *                  point_offsets, (unused)
 * out[10..13] binning: point_list, point_list_unsorted, point_list_keys, point_list_keys_unsorted
 * out[14..15] image: n_contrib, ranges */
void ref_state_ptrs(RefState* s, void** out) {
    char* c = s->geom.p;
    CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c, s->P);
    out[0] = g.depths; out[1] = g.clamped; out[2] = g.internal_radii; out[3] = g.means2D; out[4] = g.cov3D;
    out[5] = g.conic_opacity; out[6] = g.rgb; out[7] = g.tiles_touched; out[8] = g.point_offsets; out[9] = nullptr;
    c = s->binning.p;
    if (c && s->R > 0) {
        CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c, s->R);
        out[10] = b.point_list; out[11] = b.point_list_unsorted; out[12] = b.point_list_keys;
        out[13] = b.point_list_keys_unsorted;
    } else {
        out[10] = out[11] = out[12] = out[13] = nullptr;
    }
    c = s->img.p;
    CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(c, (size_t)s->W * s->H);
    out[14] = im.n_contrib; out[15] = im.ranges;
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

void ref_dist_cuda2(int P, float* points, float* mean_dists) { SimpleKNN::knn(P, (float3*)points, mean_dists); }

}  // extern "C"
