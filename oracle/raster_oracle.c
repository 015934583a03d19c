/*
 * raster_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, CPU restatement of the arithmetic of the reference's splat rasterizer
 * (RigGS, submodules/diff-gaussian-rasterization, abbreviated dgr/ below) and of
 * simple-knn's distCUDA2.  Only tests/, __graft_entry__.smoke() and the CPU-baseline /
 * reference legs of bench.py may load this library; the product path never does.
 *
 * Parity pinning: the reference ships no golden vectors for this path (SURVEY.md
 * section 4).  This restatement is pinned (a) on the GPU box against the reference's
 * own CUDA sources compiled unmodified into oracle/_ref/ (tests/test_gpu_vs_reference.py)
 * and (b) in the CPU-only suite against tests/golden/ref_cuda_*.npz, which are outputs
 * of that same compiled reference captured on a B200 by tests/golden/make_ref_cuda_golden.py.
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (see oracle/build_oracle.py).
 * A second build with -DORC_DOUBLE (liboracle_f64.so) evaluates the same expressions in double on
 * double arrays: the arbiter when two fp32 implementations disagree about an ill-conditioned
 * gradient.
 * -ffp-contract=off keeps every float expression exactly as written; the reference is
 * compiled by nvcc with FMA contraction, so float results agree to rounding, not bitwise.
 * Discrete outputs derived from floats (radius, tile rectangle) can therefore differ for
 * a handful of Gaussians per million; all pure-integer stages (keys, sort, ranges) are
 * bit-exact given the same float inputs.
 *
 * Matrix convention (dgr/cuda_rasterizer/auxiliary.h:58-77): viewmatrix / projmatrix are
 * 16 floats with element [4*c + r] = M[r][c] (i.e. the transposed, row-vector storage of
 * scene/cameras.py:61-71).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef ORC_DOUBLE
typedef double real;
#define R_SQRT sqrt
#define R_FMAX fmax
#define R_FMIN fmin
#define R_CEIL ceil
#define R_EXP exp
#else
typedef float real;
#define R_SQRT sqrtf
#define R_FMAX fmaxf
#define R_FMIN fminf
#define R_CEIL ceilf
#define R_EXP expf
#endif
int orc_real_bytes(void) { return (int)sizeof(real); }

#define TILE 16 /* dgr/cuda_rasterizer/config.h:15-16 */

/* SH basis constants, dgr/cuda_rasterizer/auxiliary.h:22-39 */
static const real C0 = 0.28209479177387814f;
static const real C1 = 0.4886025119029199f;
static const real C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                            -1.0925484305920792f, 0.5462742152960396f};
static const real C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                            0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                            -0.5900435899266435f};

/* column-major 3x3, m[c][r], the layout GLM uses (type_mat3x3.inl:486-519) */
typedef struct { real m[3][3]; } M3;

static M3 m3_mul(const M3* a, const M3* b) {
    M3 r;
    for (int c = 0; c < 3; c++)
        for (int k = 0; k < 3; k++)
            r.m[c][k] = a->m[0][k] * b->m[c][0] + a->m[1][k] * b->m[c][1] + a->m[2][k] * b->m[c][2];
    return r;
}
static M3 m3_t(const M3* a) {
    M3 r;
    for (int c = 0; c < 3; c++)
        for (int k = 0; k < 3; k++) r.m[c][k] = a->m[k][c];
    return r;
}

/* auxiliary.h:58-77 */
static void xform4x3(const real* p, const real* m, real* o) {
    o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
    o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
    o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static void xform4x4(const real* p, const real* m, real* o) {
    o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
    o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
    o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
    o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}

/* auxiliary.h:41-44 -- evaluated in double, then rounded to real */
static real ndc2pix(real v, int S) { return (real)(((v + 1.0) * S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:46-56 */
static void get_rect(real px, real py, int max_radius, int gx, int gy, uint32_t* mn, uint32_t* mx) {
    mn[0] = (uint32_t)imin(gx, imax(0, (int)((px - max_radius) / TILE)));
    mn[1] = (uint32_t)imin(gy, imax(0, (int)((py - max_radius) / TILE)));
    mx[0] = (uint32_t)imin(gx, imax(0, (int)((px + max_radius + TILE - 1) / TILE)));
    mx[1] = (uint32_t)imin(gy, imax(0, (int)((py + max_radius + TILE - 1) / TILE)));
}

/* dgr/cuda_rasterizer/forward.cu:118-152 (quaternion used as given, not normalised) */
static void cov3d_from_scale_rot(const real* s, real mod, const real* q, real* cov) {
    M3 S;
    memset(&S, 0, sizeof S);
    S.m[0][0] = mod * s[0];
    S.m[1][1] = mod * s[1];
    S.m[2][2] = mod * s[2];
    real r = q[0], x = q[1], y = q[2], z = q[3];
    M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    M3 Mm = m3_mul(&S, &R);
    M3 Mt = m3_t(&Mm);
    M3 Sig = m3_mul(&Mt, &Mm);
    cov[0] = Sig.m[0][0]; cov[1] = Sig.m[0][1]; cov[2] = Sig.m[0][2];
    cov[3] = Sig.m[1][1]; cov[4] = Sig.m[1][2]; cov[5] = Sig.m[2][2];
}

/* forward.cu:74-113; also returns T (for the backward) when Tout != NULL */
static void cov2d(const real* mean, real fx, real fy, real tfx, real tfy, const real* c3,
                  const real* view, real* out3, M3* Tout, real* tout, real* gmul) {
    real t[3];
    xform4x3(mean, view, t);
    const real limx = 1.3f * tfx, limy = 1.3f * tfy;
    const real txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = R_FMIN(limx, R_FMAX(-limx, txtz)) * t[2];
    t[1] = R_FMIN(limy, R_FMAX(-limy, tytz)) * t[2];
    if (gmul) {
        gmul[0] = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        gmul[1] = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    }
    M3 J = {{{fx / t[2], 0.f, -(fx * t[0]) / (t[2] * t[2])},
             {0.f, fy / t[2], -(fy * t[1]) / (t[2] * t[2])},
             {0.f, 0.f, 0.f}}};
    M3 Wm = {{{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}}};
    M3 T = m3_mul(&Wm, &J);
    M3 V = {{{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}}};
    M3 Tt = m3_t(&T), Vt = m3_t(&V);
    M3 a = m3_mul(&Tt, &Vt);
    M3 cov = m3_mul(&a, &T);
    out3[0] = cov.m[0][0] + 0.3f;
    out3[1] = cov.m[0][1];
    out3[2] = cov.m[1][1] + 0.3f;
    if (Tout) *Tout = T;
    if (tout) { tout[0] = t[0]; tout[1] = t[1]; tout[2] = t[2]; }
}

/* forward.cu:20-71 */
static void sh_to_rgb(int deg, int M, const real* pos, const real* campos, const real* sh,
                      real* rgb, uint8_t* clamped) {
    real d[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    real len = R_SQRT(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    real x = d[0] / len, y = d[1] / len, z = d[2] / len;
    (void)M;
    for (int c = 0; c < 3; c++) {
#define SH(k) sh[(k) * 3 + c]
        real r = C0 * SH(0);
        if (deg > 0) {
            r = r - C1 * y * SH(1) + C1 * z * SH(2) - C1 * x * SH(3);
            if (deg > 1) {
                real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                r = r + C2[0] * xy * SH(4) + C2[1] * yz * SH(5) + C2[2] * (2.0f * zz - xx - yy) * SH(6) +
                    C2[3] * xz * SH(7) + C2[4] * (xx - yy) * SH(8);
                if (deg > 2) {
                    r = r + C3[0] * y * (3.0f * xx - yy) * SH(9) + C3[1] * xy * z * SH(10) +
                        C3[2] * y * (4.0f * zz - xx - yy) * SH(11) +
                        C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12) +
                        C3[4] * x * (4.0f * zz - xx - yy) * SH(13) + C3[5] * z * (xx - yy) * SH(14) +
                        C3[6] * x * (xx - 3.0f * yy) * SH(15);
                }
            }
        }
#undef SH
        r += 0.5f;
        clamped[c] = (r < 0);
        rgb[c] = r > 0.f ? r : 0.f;
    }
}

/* forward.cu:155-256.  Optional pointers (NULL selects the other path): shs | colors_precomp,
 * (scales, rots) | cov3D_precomp.  Outputs are fully written (zeros for culled Gaussians). */
void orc_preprocess(int P, int D, int M, const real* means3D, const real* scales, real scale_mod,
                    const real* rots, const real* opac, const real* shs, const real* cov3D_precomp,
                    const real* colors_precomp, const real* view, const real* proj, const real* campos,
                    int W, int H, real tfx, real tfy, int* radii, real* means2D, real* depths,
                    real* cov3D, real* rgb, real* conic_opacity, uint32_t* tiles_touched,
                    uint8_t* clamped) {
    const real fy = H / (2.0f * tfy), fx = W / (2.0f * tfx); /* rasterizer_impl.cu:223-224 */
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        means2D[2 * i] = means2D[2 * i + 1] = 0.f;
        depths[i] = 0.f;
        for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = 0.f;
        for (int k = 0; k < 3; k++) { rgb[3 * i + k] = 0.f; clamped[3 * i + k] = 0; }
        for (int k = 0; k < 6; k++) cov3D[6 * i + k] = 0.f;
        const real* p = means3D + 3 * i;
        real pv[3], ph[4];
        xform4x3(p, view, pv);
        if (pv[2] <= 0.2f) continue; /* auxiliary.h:154 */
        xform4x4(p, proj, ph);
        real pw = 1.0f / (ph[3] + 0.0000001f);
        real pp[3] = {ph[0] * pw, ph[1] * pw, ph[2] * pw};
        const real* c3;
        if (cov3D_precomp) c3 = cov3D_precomp + 6 * i;
        else { cov3d_from_scale_rot(scales + 3 * i, scale_mod, rots + 4 * i, cov3D + 6 * i); c3 = cov3D + 6 * i; }
        real cv[3];
        cov2d(p, fx, fy, tfx, tfy, c3, view, cv, NULL, NULL, NULL);
        real det = cv[0] * cv[2] - cv[1] * cv[1];
        if (det == 0.0f) continue;
        real det_inv = 1.f / det;
        real conic[3] = {cv[2] * det_inv, -cv[1] * det_inv, cv[0] * det_inv};
        real mid = 0.5f * (cv[0] + cv[2]);
        real l1 = mid + R_SQRT(R_FMAX(0.1f, mid * mid - det));
        real l2 = mid - R_SQRT(R_FMAX(0.1f, mid * mid - det));
        real my_radius = R_CEIL(3.f * R_SQRT(R_FMAX(l1, l2)));
        real px = ndc2pix(pp[0], W), py = ndc2pix(pp[1], H);
        uint32_t mn[2], mx[2];
        get_rect(px, py, (int)my_radius, gx, gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        if (!colors_precomp) sh_to_rgb(D, M, p, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
        depths[i] = pv[2];
        radii[i] = (int)my_radius;
        means2D[2 * i] = px;
        means2D[2 * i + 1] = py;
        conic_opacity[4 * i + 0] = conic[0];
        conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2];
        conic_opacity[4 * i + 3] = opac[i];
        tiles_touched[i] = (mx[1] - mn[1]) * (mx[0] - mn[0]);
    }
}

/* rasterizer_impl.cu:54-66 */
void orc_mark_visible(int P, const real* means3D, const real* view, const real* proj, uint8_t* present) {
    (void)proj;
    for (int i = 0; i < P; i++) {
        real pv[3];
        xform4x3(means3D + 3 * i, view, pv);
        present[i] = pv[2] > 0.2f;
    }
}

/* rasterizer_impl.cu:35-50 */
uint32_t orc_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Inclusive scan of tiles_touched (rasterizer_impl.cu:278); returns the total. */
int64_t orc_scan(int P, const uint32_t* tiles_touched, uint32_t* offsets) {
    uint32_t acc = 0;
    for (int i = 0; i < P; i++) { acc += tiles_touched[i]; offsets[i] = acc; }
    return P ? (int64_t)acc : 0;
}

/* rasterizer_impl.cu:70-111 */
void orc_duplicate_with_keys(int P, int W, int H, const real* means2D, const real* depths,
                             const uint32_t* offsets, const int* radii, uint64_t* keys, uint32_t* vals) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0 : offsets[i - 1];
        uint32_t mn[2], mx[2];
        get_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, mn, mx);
        uint32_t dbits;
        const float d32 = (float)depths[i]; /* the key holds the fp32 bits in both builds */
        memcpy(&dbits, &d32, 4);
        for (uint32_t y = mn[1]; y < mx[1]; y++)
            for (uint32_t x = mn[0]; x < mx[0]; x++) {
                uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
                key <<= 32;
                key |= dbits;
                keys[off] = key;
                vals[off] = (uint32_t)i;
                off++;
            }
    }
}

/* Stable LSD radix sort on bits [0, end_bit) -- the contract of cub::DeviceRadixSort::SortPairs
 * as called at rasterizer_impl.cu:304-309. */
void orc_sort_pairs(int64_t n, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout,
                    int end_bit) {
    if (n <= 0) return;
    uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * n);
    uint64_t* kb = (uint64_t*)malloc(sizeof(uint64_t) * n);
    uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * n);
    uint32_t* vb = (uint32_t*)malloc(sizeof(uint32_t) * n);
    memcpy(ka, kin, sizeof(uint64_t) * n);
    memcpy(va, vin, sizeof(uint32_t) * n);
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << bits) - 1;
        int64_t cnt[257];
        memset(cnt, 0, sizeof cnt);
        for (int64_t i = 0; i < n; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int b = 0; b < 256; b++) cnt[b + 1] += cnt[b];
        for (int64_t i = 0; i < n; i++) {
            int64_t d = cnt[(ka[i] >> shift) & mask]++;
            kb[d] = ka[i];
            vb[d] = va[i];
        }
        uint64_t* tk = ka; ka = kb; kb = tk;
        uint32_t* tv = va; va = vb; vb = tv;
    }
    memcpy(kout, ka, sizeof(uint64_t) * n);
    memcpy(vout, va, sizeof(uint32_t) * n);
    free(ka); free(kb); free(va); free(vb);
}

/* rasterizer_impl.cu:116-138 with the memset of :311 (ranges has n_tiles uint2 entries) */
void orc_tile_ranges(int64_t L, const uint64_t* keys, int n_tiles, uint32_t* ranges) {
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)n_tiles);
    for (int64_t i = 0; i < L; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == L - 1) ranges[2 * cur + 1] = (uint32_t)L;
    }
}

/* forward.cu:261-381, one pixel at a time over the tile's sorted range. */
void orc_blend_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                       const real* means2D, const real* colors, const real* depths,
                       const real* conic_opacity, const real* bg, real* out_color, real* out_depth,
                       real* out_alpha, uint32_t* n_contrib) {
    const int gx = (W + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 8)
    for (int py = 0; py < H; py++)
        for (int px = 0; px < W; px++) {
            int tile = (py / TILE) * gx + px / TILE;
            uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
            real T = 1.0f, C[3] = {0, 0, 0}, wsum = 0, Dp = 0;
            uint32_t contributor = 0, last = 0;
            const real pxf = (real)px, pyf = (real)py;
            for (uint32_t k = r0; k < r1; k++) {
                contributor++;
                uint32_t id = point_list[k];
                real dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                const real* co = conic_opacity + 4 * id;
                real power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                if (power > 0.0f) continue;
                real alpha = R_FMIN(0.99f, co[3] * R_EXP(power));
                if (alpha < 1.0f / 255.0f) continue;
                real test_T = T * (1 - alpha);
                if (test_T < 0.0001f) break; /* done: entry not blended (forward.cu:351-355) */
                for (int c = 0; c < 3; c++) C[c] += colors[3 * id + c] * alpha * T;
                wsum += alpha * T;
                Dp += depths[id] * alpha * T;
                T = test_T;
                last = contributor;
            }
            size_t pid = (size_t)py * W + px;
            n_contrib[pid] = last;
            for (int c = 0; c < 3; c++) out_color[(size_t)c * H * W + pid] = C[c] + T * bg[c];
            out_alpha[pid] = wsum; /* sum of alpha*T, not 1-T (forward.cu:378) */
            out_depth[pid] = Dp;
        }
}

/* backward.cu:415-601.  Per-pair terms are computed in real exactly as written there; the
 * per-Gaussian sums (global atomicAdd in the reference, order-nondeterministic) are kept in
 * double here so the oracle is the better-rounded side of any disagreement.
 * Outputs (all double, zero-initialised by the caller): dmean2D[P*2], dconic[P*3] (x,y,w),
 * dopacity[P], dcolor[P*3], ddepth[P]. */
void orc_blend_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                        const real* bg, const real* means2D, const real* conic_opacity,
                        const real* colors, const real* depths, const real* alphas,
                        const uint32_t* n_contrib, const real* dL_dpix, const real* dL_dpix_depth,
                        const real* dL_dpix_alpha, double* dmean2D, double* dconic, double* dopacity,
                        double* dcolor, double* ddepth) {
    (void)P;
    const int gx = (W + TILE - 1) / TILE;
    const real ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 4)
    for (int py = 0; py < H; py++)
        for (int px = 0; px < W; px++) {
            int tile = (py / TILE) * gx + px / TILE;
            uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
            size_t pid = (size_t)py * W + px;
            const real T_final = 1 - alphas[pid];
            real T = T_final;
            const uint32_t last_contributor = n_contrib[pid];
            real accum_rec[3] = {0, 0, 0}, dpix[3], accum_depth = 0, accum_alpha = 0;
            for (int c = 0; c < 3; c++) dpix[c] = dL_dpix[(size_t)c * H * W + pid];
            const real dpd = dL_dpix_depth[pid], dpa = dL_dpix_alpha[pid];
            real last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0;
            const real pxf = (real)px, pyf = (real)py;
            uint32_t contributor = r1 - r0;
            for (uint32_t k = r1; k-- > r0;) {
                contributor--;
                if (contributor >= last_contributor) continue;
                uint32_t id = point_list[k];
                real dx = means2D[2 * id] - pxf, dy = means2D[2 * id + 1] - pyf;
                const real* co = conic_opacity + 4 * id;
                real power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                if (power > 0.0f) continue;
                const real G = R_EXP(power);
                const real alpha = R_FMIN(0.99f, co[3] * G);
                if (alpha < 1.0f / 255.0f) continue;
                T = T / (1.f - alpha);
                const real w = alpha * T;
                real dL_dopa = 0.0f;
                for (int c = 0; c < 3; c++) {
                    const real col = colors[3 * id + c];
                    accum_rec[c] = last_alpha * last_color[c] + (1.f - last_alpha) * accum_rec[c];
                    last_color[c] = col;
                    dL_dopa += (col - accum_rec[c]) * dpix[c];
                    _Pragma("omp atomic")
                    dcolor[3 * (size_t)id + c] += (double)(w * dpix[c]);
                }
                const real cd = depths[id];
                accum_depth = last_alpha * last_depth + (1.f - last_alpha) * accum_depth;
                last_depth = cd;
                dL_dopa += (cd - accum_depth) * dpd;
                _Pragma("omp atomic")
                ddepth[id] += (double)(w * dpd);
                accum_alpha = last_alpha + (1.f - last_alpha) * accum_alpha;
                dL_dopa += (1 - accum_alpha) * dpa;
                dL_dopa *= T;
                last_alpha = alpha;
                real bgdot = 0;
                for (int c = 0; c < 3; c++) bgdot += bg[c] * dpix[c];
                dL_dopa += (-T_final / (1.f - alpha)) * bgdot;
                const real dL_dG = co[3] * dL_dopa;
                const real gdx = G * dx, gdy = G * dy;
                const real dG_ddelx = -gdx * co[0] - gdy * co[1];
                const real dG_ddely = -gdy * co[2] - gdx * co[1];
                _Pragma("omp atomic")
                dmean2D[2 * (size_t)id] += (double)(dL_dG * dG_ddelx * ddelx_dx);
                _Pragma("omp atomic")
                dmean2D[2 * (size_t)id + 1] += (double)(dL_dG * dG_ddely * ddely_dy);
                _Pragma("omp atomic")
                dconic[3 * (size_t)id] += (double)(-0.5f * gdx * dx * dL_dG);
                _Pragma("omp atomic")
                dconic[3 * (size_t)id + 1] += (double)(-0.5f * gdx * dy * dL_dG);
                _Pragma("omp atomic")
                dconic[3 * (size_t)id + 2] += (double)(-0.5f * gdy * dy * dL_dG);
                _Pragma("omp atomic")
                dopacity[id] += (double)(G * dL_dopa);
            }
        }
}

/* auxiliary.h:107-117 */
static void dnormvdv3(const real* v, const real* dv, real* o) {
    real sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    real inv = 1.0f / R_SQRT(sum2 * sum2 * sum2);
    o[0] = ((+sum2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * inv;
    o[1] = (-v[0] * v[1] * dv[0] + (sum2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * inv;
    o[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (sum2 - v[2] * v[2]) * dv[2]) * inv;
}

/* backward.cu:20-139 */
static void sh_backward(int deg, int M, const real* pos, const real* campos, const real* sh,
                        const uint8_t* clamped, const real* dL_dcolor, real* dL_dmean_acc, real* dL_dsh) {
    real dorig[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    real len = R_SQRT(dorig[0] * dorig[0] + dorig[1] * dorig[1] + dorig[2] * dorig[2]);
    real x = dorig[0] / len, y = dorig[1] / len, z = dorig[2] / len;
    real g[3] = {dL_dcolor[0] * (clamped[0] ? 0.f : 1.f), dL_dcolor[1] * (clamped[1] ? 0.f : 1.f),
                  dL_dcolor[2] * (clamped[2] ? 0.f : 1.f)};
    real dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
    (void)M;
#define SH(k, c) sh[(k) * 3 + (c)]
#define DSH(k, v) for (int c = 0; c < 3; c++) dL_dsh[(k) * 3 + c] = (v) * g[c]
    DSH(0, C0);
    if (deg > 0) {
        DSH(1, -C1 * y); DSH(2, C1 * z); DSH(3, -C1 * x);
        for (int c = 0; c < 3; c++) { dx[c] = -C1 * SH(3, c); dy[c] = -C1 * SH(1, c); dz[c] = C1 * SH(2, c); }
        if (deg > 1) {
            real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            DSH(4, C2[0] * xy); DSH(5, C2[1] * yz); DSH(6, C2[2] * (2.f * zz - xx - yy));
            DSH(7, C2[3] * xz); DSH(8, C2[4] * (xx - yy));
            for (int c = 0; c < 3; c++) {
                dx[c] += C2[0] * y * SH(4, c) + C2[2] * 2.f * -x * SH(6, c) + C2[3] * z * SH(7, c) + C2[4] * 2.f * x * SH(8, c);
                dy[c] += C2[0] * x * SH(4, c) + C2[1] * z * SH(5, c) + C2[2] * 2.f * -y * SH(6, c) + C2[4] * 2.f * -y * SH(8, c);
                dz[c] += C2[1] * y * SH(5, c) + C2[2] * 2.f * 2.f * z * SH(6, c) + C2[3] * x * SH(7, c);
            }
            if (deg > 2) {
                DSH(9, C3[0] * y * (3.f * xx - yy)); DSH(10, C3[1] * xy * z);
                DSH(11, C3[2] * y * (4.f * zz - xx - yy));
                DSH(12, C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                DSH(13, C3[4] * x * (4.f * zz - xx - yy)); DSH(14, C3[5] * z * (xx - yy));
                DSH(15, C3[6] * x * (xx - 3.f * yy));
                for (int c = 0; c < 3; c++) {
                    dx[c] += (C3[0] * SH(9, c) * 3.f * 2.f * xy + C3[1] * SH(10, c) * yz + C3[2] * SH(11, c) * -2.f * xy +
                              C3[3] * SH(12, c) * -3.f * 2.f * xz + C3[4] * SH(13, c) * (-3.f * xx + 4.f * zz - yy) +
                              C3[5] * SH(14, c) * 2.f * xz + C3[6] * SH(15, c) * 3.f * (xx - yy));
                    dy[c] += (C3[0] * SH(9, c) * 3.f * (xx - yy) + C3[1] * SH(10, c) * xz +
                              C3[2] * SH(11, c) * (-3.f * yy + 4.f * zz - xx) + C3[3] * SH(12, c) * -3.f * 2.f * yz +
                              C3[4] * SH(13, c) * -2.f * xy + C3[5] * SH(14, c) * -2.f * yz +
                              C3[6] * SH(15, c) * -3.f * 2.f * xy);
                    dz[c] += (C3[1] * SH(10, c) * xy + C3[2] * SH(11, c) * 4.f * 2.f * yz +
                              C3[3] * SH(12, c) * 3.f * (2.f * zz - xx - yy) + C3[4] * SH(13, c) * 4.f * 2.f * xz +
                              C3[5] * SH(14, c) * (xx - yy));
                }
            }
        }
    }
#undef SH
#undef DSH
    real ddir[3] = {dx[0] * g[0] + dx[1] * g[1] + dx[2] * g[2], dy[0] * g[0] + dy[1] * g[1] + dy[2] * g[2],
                     dz[0] * g[0] + dz[1] * g[1] + dz[2] * g[2]};
    real dm[3];
    dnormvdv3(dorig, ddir, dm);
    dL_dmean_acc[0] += dm[0]; dL_dmean_acc[1] += dm[1]; dL_dmean_acc[2] += dm[2];
}

/* backward.cu:278-341 */
static void cov3d_backward(const real* scale, real mod, const real* q, const real* dc, real* dscale, real* drot) {
    real r = q[0], x = q[1], y = q[2], z = q[3];
    M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    real s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    M3 S;
    memset(&S, 0, sizeof S);
    S.m[0][0] = s[0]; S.m[1][1] = s[1]; S.m[2][2] = s[2];
    M3 Mm = m3_mul(&S, &R);
    M3 dSig = {{{dc[0], 0.5f * dc[1], 0.5f * dc[2]}, {0.5f * dc[1], dc[3], 0.5f * dc[4]}, {0.5f * dc[2], 0.5f * dc[4], dc[5]}}};
    M3 twoM;
    for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) twoM.m[c][k] = 2.0f * Mm.m[c][k];
    M3 dM = m3_mul(&twoM, &dSig);
    M3 Rt = m3_t(&R), dMt = m3_t(&dM);
    for (int k = 0; k < 3; k++)
        dscale[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
    for (int k = 0; k < 3; k++) for (int e = 0; e < 3; e++) dMt.m[k][e] *= s[k];
#define D(a, b) dMt.m[a][b]
    drot[0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
    drot[1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
    drot[2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
    drot[3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
}

/* backward.cu:144-274 (computeCov2DCUDA) followed by :346-412 (preprocessCUDA backward).
 * Inputs dmean2D[P*2], dconic[P*3] (x,y,w of the reference's float4), dcolor[P*3], ddepth[P]
 * are the (real) outputs of the blend backward.  Outputs are fully written; Gaussians with
 * radii <= 0 get zeros (the reference leaves its zero-initialised tensors untouched there). */
void orc_preprocess_backward(int P, int D, int M, const real* means3D, const int* radii, const real* shs,
                             const uint8_t* clamped, const real* scales, const real* rots, real scale_mod,
                             const real* cov3Ds, const real* view, const real* proj, const real* campos,
                             int W, int H, real tfx, real tfy, const real* dmean2D, const real* dconic,
                             const real* dcolor, const real* ddepth, real* dmean3D, real* dcov3D,
                             real* dsh, real* dscale, real* drot) {
    const real fy = H / (2.0f * tfy), fx = W / (2.0f * tfx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        for (int k = 0; k < 3; k++) dmean3D[3 * i + k] = 0.f;
        for (int k = 0; k < 6; k++) dcov3D[6 * i + k] = 0.f;
        if (dsh) for (int k = 0; k < 3 * M; k++) dsh[(size_t)i * 3 * M + k] = 0.f;
        if (dscale) { for (int k = 0; k < 3; k++) dscale[3 * i + k] = 0.f; for (int k = 0; k < 4; k++) drot[4 * i + k] = 0.f; }
        if (!(radii[i] > 0)) continue;
        const real* mean = means3D + 3 * i;
        const real* c3 = cov3Ds + 6 * i;
        real cv[3], t[3], gm[2];
        M3 T;
        cov2d(mean, fx, fy, tfx, tfy, c3, view, cv, &T, t, gm);
        const real a = cv[0], b = cv[1], c = cv[2];
        const real dcx = dconic[3 * i], dcy = dconic[3 * i + 1], dcz = dconic[3 * i + 2];
        real denom = a * c - b * b;
        real dL_da = 0, dL_db = 0, dL_dc = 0;
        real denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        real* dc = dcov3D + 6 * i;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
            dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
            dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
#define TT(c_, r_) T.m[c_][r_]
            dc[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
            dc[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
            dc[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
            dc[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
            dc[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
            dc[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
        }
        real V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        real dT[2][3];
        for (int k = 0; k < 3; k++) {
            dT[0][k] = 2 * (TT(0, 0) * V[k][0] + TT(0, 1) * V[k][1] + TT(0, 2) * V[k][2]) * dL_da +
                       (TT(1, 0) * V[k][0] + TT(1, 1) * V[k][1] + TT(1, 2) * V[k][2]) * dL_db;
            dT[1][k] = 2 * (TT(1, 0) * V[k][0] + TT(1, 1) * V[k][1] + TT(1, 2) * V[k][2]) * dL_dc +
                       (TT(0, 0) * V[k][0] + TT(0, 1) * V[k][1] + TT(0, 2) * V[k][2]) * dL_db;
        }
#undef TT
        /* W (GLM, column-major): W[c][r] = view[4*r + c]  (backward.cu:182-185) */
#define WW(c_, r_) view[4 * (r_) + (c_)]
        real dJ00 = WW(0, 0) * dT[0][0] + WW(0, 1) * dT[0][1] + WW(0, 2) * dT[0][2];
        real dJ02 = WW(2, 0) * dT[0][0] + WW(2, 1) * dT[0][1] + WW(2, 2) * dT[0][2];
        real dJ11 = WW(1, 0) * dT[1][0] + WW(1, 1) * dT[1][1] + WW(1, 2) * dT[1][2];
        real dJ12 = WW(2, 0) * dT[1][0] + WW(2, 1) * dT[1][1] + WW(2, 2) * dT[1][2];
#undef WW
        real tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        real dtx = gm[0] * -fx * tz2 * dJ02;
        real dty = gm[1] * -fy * tz2 * dJ12;
        real dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t[0]) * tz3 * dJ02 + (2 * fy * t[1]) * tz3 * dJ12;
        real dm[3] = {view[0] * dtx + view[1] * dty + view[2] * dtz, view[4] * dtx + view[5] * dty + view[6] * dtz,
                       view[8] * dtx + view[9] * dty + view[10] * dtz}; /* auxiliary.h:89-97 */
        /* preprocess backward, backward.cu:372-403 */
        real mh[4];
        xform4x4(mean, proj, mh);
        real mw = 1.0f / (mh[3] + 0.0000001f);
        real mul1 = (proj[0] * mean[0] + proj[4] * mean[1] + proj[8] * mean[2] + proj[12]) * mw * mw;
        real mul2 = (proj[1] * mean[0] + proj[5] * mean[1] + proj[9] * mean[2] + proj[13]) * mw * mw;
        const real g2x = dmean2D[2 * i], g2y = dmean2D[2 * i + 1];
        real d1[3];
        d1[0] = (proj[0] * mw - proj[3] * mul1) * g2x + (proj[1] * mw - proj[3] * mul2) * g2y;
        d1[1] = (proj[4] * mw - proj[7] * mul1) * g2x + (proj[5] * mw - proj[7] * mul2) * g2y;
        d1[2] = (proj[8] * mw - proj[11] * mul1) * g2x + (proj[9] * mw - proj[11] * mul2) * g2y;
        for (int k = 0; k < 3; k++) dm[k] += d1[k];
        real mul3 = view[2] * mean[0] + view[6] * mean[1] + view[10] * mean[2] + view[14];
        real d2[3] = {(view[2] - view[3] * mul3) * ddepth[i], (view[6] - view[7] * mul3) * ddepth[i],
                       (view[10] - view[11] * mul3) * ddepth[i]};
        for (int k = 0; k < 3; k++) dm[k] += d2[k];
        if (shs) sh_backward(D, M, mean, campos, shs + (size_t)i * 3 * M, clamped + 3 * i, dcolor + 3 * i, dm, dsh + (size_t)i * 3 * M);
        for (int k = 0; k < 3; k++) dmean3D[3 * i + k] = dm[k];
        if (scales) cov3d_backward(scales + 3 * i, scale_mod, rots + 4 * i, dc, dscale + 3 * i, drot + 4 * i);
    }
}

/* simple-knn: the result of knn/simple_knn.cu:186-221 is the exact mean of the three smallest squared
 * distances to other points (the Morton boxes are only an acceleration structure), so a brute-force scan
 * is a valid restatement.  updateKBest (:120-134) keeps the three smallest, duplicates included. */
void orc_dist2_knn3(int P, const real* pts, real* out) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        real best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        const real* r = pts + 3 * i;
        for (int j = 0; j < P; j++) {
            if (j == i) continue;
            real dx = pts[3 * j] - r[0], dy = pts[3 * j + 1] - r[1], dz = pts[3 * j + 2] - r[2];
            real dist = dx * dx + dy * dy + dz * dz;
            for (int k = 0; k < 3; k++)
                if (best[k] > dist) { real t = best[k]; best[k] = dist; dist = t; }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
