// Image loss of the trainer with its gradient, two launches:
//     loss = (1 - lambda) * mean|x - y| + lambda * (1 - mean SSIM(x, y))          (train_rig.py:508-509)
// SSIM as utils/loss_utils.py:36-77: 11x11 Gaussian window (sigma 1.5, separable, zero padding 5), per channel
//     mu1 = w*x, mu2 = w*y, s11 = w*x^2 - mu1^2, s22 = w*y^2 - mu2^2, s12 = w*xy - mu1 mu2,
//     S = (2 mu1 mu2 + C1)(2 s12 + C2) / ((mu1^2 + mu2^2 + C1)(s11 + s22 + C2)),  C1 = 0.01^2, C2 = 0.03^2.
// The reference runs five grouped conv2d plus ~20 element-wise launches and autograd runs them backwards again.
// Here
//   pass 1 (ssim_map_kernel): a CTA stages a 16x16 tile of x and y with its 5-pixel halo in shared memory, runs the
//     five separable convolutions, evaluates S and the three partial derivatives the gradient needs,
//         Dm  = dS/d(w*x)   (through mu1 in all four factors),   Dxx = dS/d(w*x^2) = -S / B2,   Dxy = dS/d(w*xy) = 2 A1 / (B1 B2)
//     writes them to scratch, and reduces sum S and sum |x - y| (per-CTA partials, last CTA adds them up in index order);
//   pass 2 (ssim_grad_kernel): the window is symmetric, so  dL/dx(p) = w*Dm + 2 x(p) (w*Dxx) + y(p) (w*Dxy)  -- three
//     more separable convolutions on the scratch maps -- scaled by -lambda / n, plus the L1 term (1 - lambda) sign(x - y) / n.
// Tolerance against the torch restatement of the reference: 1e-5 on the loss, 1e-4 on the gradient (tests).
#include "common.cuh"
#include "../../include/riggs_b200.h"

namespace rg {

constexpr int kTile = 16, kHalo = 5, kWin = 11, kIn = kTile + 2 * kHalo;  // 26

struct SsimWindow {
    float w[kWin];
};

// gaussian(11, 1.5) of utils/loss_utils.py:31-33, evaluated in double and rounded once
static SsimWindow make_window() {
    SsimWindow g;
    double v[kWin], s = 0.0;
    for (int i = 0; i < kWin; i++) { v[i] = exp(-(double)((i - kWin / 2) * (i - kWin / 2)) / (2.0 * 1.5 * 1.5)); s += v[i]; }
    for (int i = 0; i < kWin; i++) g.w[i] = (float)(v[i] / s);
    return g;
}

constexpr int kPitch = 28;  // row pitch of the staged tiles: 26 used, rows start on 16-byte boundaries
constexpr int kPitchH = 20;  // pitch of the filtered tiles (16 used): rows four apart fall into different banks

// Separable 11-tap convolution of N_MAPS maps held as s_in[map][26][28] -> s_v[map][16][16].  Both passes are register
// blocked: a thread produces four neighbouring outputs from fourteen inputs, so the shared-memory loads per FMA drop
// from 1 to 1/11 (horizontal pass, 128-bit loads) and 1/3 (vertical pass); a one-load-per-FMA version of this function was
// bound by the shared-memory pipe (72 us for the five maps of an 800x800x3 image; now ~45 us).
template <int N_MAPS>
__device__ __forceinline__ void separable(const float (*s_in)[kIn][kPitch], float (*s_h)[kIn][kPitchH],
                                          float (*s_v)[kTile][kPitchH], const SsimWindow& g) {
    const int tid = threadIdx.x;
    constexpr int kSeg = kTile / 4;  // four outputs per work item
    for (int it = tid; it < N_MAPS * kIn * kSeg; it += kTile * kTile) {
        // consecutive threads take consecutive ROWS of one column segment: with a pitch of 28 words the eight 128-bit
        // loads of a quarter warp then fall into eight different groups of four banks
        const int m = it / (kIn * kSeg), rem = it - m * (kIn * kSeg), r = rem % kIn, c0 = (rem / kIn) * 4;
        float in[16];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float4 v = *reinterpret_cast<const float4*>(&s_in[m][r][c0 + 4 * q]);
            in[4 * q] = v.x; in[4 * q + 1] = v.y; in[4 * q + 2] = v.z; in[4 * q + 3] = v.w;
        }
        float4 o = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int t = 0; t < kWin; t++) {
            o.x = fmaf(g.w[t], in[t], o.x);
            o.y = fmaf(g.w[t], in[t + 1], o.y);
            o.z = fmaf(g.w[t], in[t + 2], o.z);
            o.w = fmaf(g.w[t], in[t + 3], o.w);
        }
        *reinterpret_cast<float4*>(&s_h[m][r][c0]) = o;
    }
    __syncthreads();
    for (int it = tid; it < N_MAPS * kTile * kSeg; it += kTile * kTile) {
        const int m = it / (kTile * kSeg), rem = it - m * (kTile * kSeg), r0 = (rem / kTile) * 4, c = rem % kTile;
        float in[14];
#pragma unroll
        for (int q = 0; q < 14; q++) in[q] = s_h[m][r0 + q][c];
        float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int t = 0; t < kWin; t++) {
#pragma unroll
            for (int q = 0; q < 4; q++) o[q] = fmaf(g.w[t], in[t + q], o[q]);
        }
#pragma unroll
        for (int q = 0; q < 4; q++) s_v[m][r0 + q][c] = o[q];
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kTile* kTile) ssim_map_kernel(int C, int H, int W, const float* __restrict__ x,
                                                                const float* __restrict__ y, SsimWindow g,
                                                                float* __restrict__ dmap,  // [3][C][H][W]: Dm, Dxx, Dxy
                                                                float* __restrict__ partial) {  // [blocks][2]: sum S, sum |x-y|
    __shared__ __align__(16) float s_in[5][kIn][kPitch];
    __shared__ __align__(16) float s_h[5][kIn][kPitchH];
    __shared__ float s_v[5][kTile][kPitchH];
    __shared__ float s_red[2][kTile * kTile / 32];
    const int ch = blockIdx.z, tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile;
    const size_t plane = (size_t)H * W;
    const float* xc = x + ch * plane;
    const float* yc = y + ch * plane;
    // one warp per staged row (26 of its 32 lanes), rows r, r + 8, ...
    for (int r = threadIdx.x >> 5, c = threadIdx.x & 31; r < kIn; r += kTile * kTile / 32) {
        if (c >= kIn) continue;
        const int gy = ty0 + r - kHalo, gx = tx0 + c - kHalo;
        float a = 0.f, b = 0.f;  // zero padding, as conv2d(padding=5)
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) { a = xc[(size_t)gy * W + gx]; b = yc[(size_t)gy * W + gx]; }
        s_in[0][r][c] = a; s_in[1][r][c] = b; s_in[2][r][c] = a * a; s_in[3][r][c] = b * b; s_in[4][r][c] = a * b;
    }
    __syncthreads();
    separable<5>(s_in, s_h, s_v, g);
    const int py = threadIdx.x / kTile, px = threadIdx.x % kTile;
    const float v[5] = {s_v[0][py][px], s_v[1][py][px], s_v[2][py][px], s_v[3][py][px], s_v[4][py][px]};
    const int gy = ty0 + py, gx = tx0 + px;
    float S = 0.f, l1 = 0.f;
    if (gy < H && gx < W) {
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float mu1 = v[0], mu2 = v[1];
        const float s11 = v[2] - mu1 * mu1, s22 = v[3] - mu2 * mu2, s12 = v[4] - mu1 * mu2;
        const float A1 = 2.f * mu1 * mu2 + C1, A2 = 2.f * s12 + C2, B1 = mu1 * mu1 + mu2 * mu2 + C1, B2 = s11 + s22 + C2;
        const float inv = 1.0f / (B1 * B2);
        S = A1 * A2 * inv;
        // dS/dmu1 with s11 = cxx - mu1^2 and s12 = cxy - mu1 mu2 held at fixed cxx, cxy
        const float Dm = 2.f * mu2 * (A2 - A1) * inv - S * 2.f * mu1 * (1.0f / B1 - 1.0f / B2);
        const size_t o = ch * plane + (size_t)gy * W + gx, maps = (size_t)C * plane;
        dmap[o] = Dm;
        dmap[maps + o] = -S / B2;
        dmap[2 * maps + o] = 2.f * A1 * inv;
        l1 = fabsf(s_in[0][py + kHalo][px + kHalo] - s_in[1][py + kHalo][px + kHalo]);
    }
    S = warp_sum(S);
    l1 = warp_sum(l1);
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = S; s_red[1][threadIdx.x >> 5] = l1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < kTile * kTile / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
        const size_t blk = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[2 * blk] = a;
        partial[2 * blk + 1] = b;
    }
}

__global__ void __launch_bounds__(kTile* kTile) ssim_grad_kernel(int C, int H, int W, const float* __restrict__ x,
                                                                 const float* __restrict__ y, SsimWindow g,
                                                                 const float* __restrict__ dmap, float lambda,
                                                                 float* __restrict__ grad, const float* __restrict__ partial,
                                                                 int n_partial, float* __restrict__ loss_out) {  // [3]: loss, l1, ssim
    __shared__ __align__(16) float s_in[3][kIn][kPitch];
    __shared__ __align__(16) float s_h[3][kIn][kPitchH];
    __shared__ float s_v[3][kTile][kPitchH];
    const int ch = blockIdx.z, tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile;
    const size_t plane = (size_t)H * W, maps = (size_t)C * plane;
    for (int r = threadIdx.x >> 5, c = threadIdx.x & 31; r < kIn; r += kTile * kTile / 32) {
        if (c >= kIn) continue;
        const int gy = ty0 + r - kHalo, gx = tx0 + c - kHalo;
        float a = 0.f, b = 0.f, d = 0.f;
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
            const size_t o = ch * plane + (size_t)gy * W + gx;
            a = dmap[o]; b = dmap[maps + o]; d = dmap[2 * maps + o];
        }
        s_in[0][r][c] = a; s_in[1][r][c] = b; s_in[2][r][c] = d;
    }
    __syncthreads();
    separable<3>(s_in, s_h, s_v, g);
    const int py = threadIdx.x / kTile, px = threadIdx.x % kTile;
    const float v[3] = {s_v[0][py][px], s_v[1][py][px], s_v[2][py][px]};
    const int gy = ty0 + py, gx = tx0 + px;
    const float inv_n = 1.0f / (float)maps;
    if (gy < H && gx < W) {
        const size_t o = ch * plane + (size_t)gy * W + gx;
        const float a = x[o], b = y[o];
        const float dssim = v[0] + 2.f * a * v[1] + b * v[2];  // d(sum S)/dx(p)
        const float d = a - b;
        const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        grad[o] = ((1.f - lambda) * sgn - lambda * dssim) * inv_n;
    }
    // the first CTA also folds the pass-1 partial sums, in index order (deterministic)
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x < 32) {
        float s = 0.f, l = 0.f;
        for (int i = threadIdx.x; i < n_partial; i += 32) { s += partial[2 * i]; l += partial[2 * i + 1]; }
        s = warp_sum(s);
        l = warp_sum(l);
        if (threadIdx.x == 0) {
            const float ssim = s * inv_n, l1 = l * inv_n;
            loss_out[0] = (1.f - lambda) * l1 + lambda * (1.f - ssim);
            loss_out[1] = l1;
            loss_out[2] = ssim;
        }
    }
}

}  // namespace rg

using namespace rg;

extern "C" {

size_t rg_ssim_loss_scratch_bytes(int channels, int height, int width) {
    if (channels <= 0 || height <= 0 || width <= 0) return 0;
    const size_t blocks = (size_t)channels * ((height + kTile - 1) / kTile) * ((width + kTile - 1) / kTile);
    return sizeof(float) * (3 * (size_t)channels * height * width + 2 * blocks) + 256;
}

int rg_ssim_l1_loss(int channels, int height, int width, const float* output, const float* gt, float lambda_dssim,
                    float* loss3, float* grad, void* scratch, size_t scratch_bytes, rg_stream_t stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (channels <= 0 || height <= 0 || width <= 0 || channels > 65535) { set_error("rg_ssim_l1_loss: empty image"); return -1; }
    if (!output || !gt || !loss3 || !grad || !scratch || scratch_bytes < rg_ssim_loss_scratch_bytes(channels, height, width)) {
        set_error("rg_ssim_l1_loss: null argument or scratch too small");
        return -1;
    }
    static const SsimWindow g = make_window();
    const dim3 grid((width + kTile - 1) / kTile, (height + kTile - 1) / kTile, channels);
    const int blocks = (int)(grid.x * grid.y * grid.z);
    float* dmap = static_cast<float*>(scratch);
    float* partial = dmap + 3 * (size_t)channels * height * width;
    {
        LaunchScope ls("ssim_map", stream);
        ssim_map_kernel<<<grid, kTile * kTile, 0, stream>>>(channels, height, width, output, gt, g, dmap, partial);
    }
    {
        LaunchScope ls("ssim_grad", stream);
        ssim_grad_kernel<<<grid, kTile * kTile, 0, stream>>>(channels, height, width, output, gt, g, dmap, lambda_dssim, grad,
                                                            partial, blocks, loss3);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // extern "C"
