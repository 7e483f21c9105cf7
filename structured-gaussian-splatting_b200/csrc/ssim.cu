// ssim.cu - fused L1 + D-SSIM image loss of the training step.
//
// Replaces, around the rasterizer (reference train.py:104-105, utils/loss_utils.py:17-63):
//   Ll1  = l1_loss(image, gt)
//   loss = (1 - lambda_dssim) * Ll1 + lambda_dssim * (1 - ssim(image, gt))
// where ssim() is five depthwise 11x11 Gaussian convolutions (sigma 1.5, zero padding 5) of x, y, x*x, y*y, x*y,
// ~15 elementwise launches on 3xHxW temporaries, and the autograd backward of all of it.  Here: two kernels.
//   ssim_fwd_kernel  one CTA per 16x16 pixel tile and channel: stages the 26x26 halo tile of both images in shared memory,
//                    runs the separable window (11 horizontal taps, then 11 vertical), evaluates the SSIM map, adds the
//                    tile's sum(map) and sum|x-y| to two device scalars and writes three partial-derivative maps
//                    dm/dmu1, dm/dE[x^2], dm/dE[xy] (the chain rule through sigma1_sq = E[x^2] - mu1^2 etc. applied).
//   ssim_bwd_kernel  the adjoint of a zero-padded symmetric convolution is the same convolution: the three maps are
//                    windowed again and combined,  dL/dx = k_ssim (W*A + 2 x W*B + y W*C) + k_l1 sign(x - y).
// The loss is a mean, so the upstream factor is a constant known at launch: value and gradient come from one call.
// HBM: forward reads 2 and writes 3 floats per element, backward reads 5 and writes 1 (44 B per element in all).
#include "gsr_common.cuh"
#include <cmath>

namespace gsr {

constexpr int SS_T  = 16;              // output tile edge
constexpr int SS_R  = 5;               // window radius (window_size 11, loss_utils.py:39)
constexpr int SS_IN = SS_T + 2 * SS_R; // 26: staged edge

struct SsimWindow { float w[2 * SS_R + 1]; };

// loss_utils.py:23-25: gauss[i] = exp(-(i - 5)^2 / (2 sigma^2)) as float32, divided by its float32 sum.
static SsimWindow make_window()
{
    SsimWindow win;
    float sum = 0.f;
    for (int i = 0; i <= 2 * SS_R; i++) {
        win.w[i] = (float)std::exp(-(double)((i - SS_R) * (i - SS_R)) / (2.0 * 1.5 * 1.5));
        sum += win.w[i];
    }
    for (int i = 0; i <= 2 * SS_R; i++) win.w[i] /= sum;
    return win;
}

__device__ __forceinline__ float block_sum_256(float v, float* wsum)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x < 8) {
        r = wsum[threadIdx.x];
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) r += __shfl_xor_sync(0xffu, r, o);
    }
    __syncthreads();
    return r;                          // valid in thread 0
}

__global__ void __launch_bounds__(256)
ssim_fwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W, SsimWindow win,
                float* __restrict__ maps, size_t plane, float* __restrict__ sums)
{
    __shared__ float sx[SS_IN][SS_IN + 1];          // odd row stride: a column of 26 rows hits 26 banks
    __shared__ float sy[SS_IN][SS_IN + 1];
    __shared__ float hz[5][SS_IN][SS_T + 1];
    __shared__ float wsum[8];

    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * SS_T - SS_R, y0 = blockIdx.y * SS_T - SS_R;

    for (int i = tid; i < SS_IN * SS_IN; i += 256) {
        const int r = i / SS_IN, c = i - r * SS_IN;
        const int gx = x0 + c, gy = y0 + r;
        float a = 0.f, b = 0.f;                    // zero padding (F.conv2d padding=5)
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t o = base + (size_t)gy * W + gx;
            a = img1[o]; b = img2[o];
        }
        sx[r][c] = a; sy[r][c] = b;
    }
    __syncthreads();

    // horizontal taps: 26 rows x 16 columns; lanes run down a column (conflict-free with the odd stride)
    for (int i = tid; i < SS_IN * SS_T; i += 256) {
        const int c = i / SS_IN, r = i - c * SS_IN;
        float m1 = 0.f, m2 = 0.f, xx = 0.f, yy = 0.f, xy = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * SS_R; k++) {
            const float a = sx[r][c + k], b = sy[r][c + k], w = win.w[k];
            const float wa = w * a, wb = w * b;
            m1 += wa; m2 += wb;
            xx = fmaf(wa, a, xx); yy = fmaf(wb, b, yy); xy = fmaf(wa, b, xy);
        }
        hz[0][r][c] = m1; hz[1][r][c] = m2; hz[2][r][c] = xx; hz[3][r][c] = yy; hz[4][r][c] = xy;
    }
    __syncthreads();

    const int tx = tid & 15, ty = tid >> 4;
    const int gx = blockIdx.x * SS_T + tx, gy = blockIdx.y * SS_T + ty;
    float mu1 = 0.f, mu2 = 0.f, xx = 0.f, yy = 0.f, xy = 0.f;
#pragma unroll
    for (int k = 0; k <= 2 * SS_R; k++) {
        const float w = win.w[k];
        mu1 = fmaf(w, hz[0][ty + k][tx], mu1);
        mu2 = fmaf(w, hz[1][ty + k][tx], mu2);
        xx  = fmaf(w, hz[2][ty + k][tx], xx);
        yy  = fmaf(w, hz[3][ty + k][tx], yy);
        xy  = fmaf(w, hz[4][ty + k][tx], xy);
    }
    float m = 0.f, l1 = 0.f;
    if (gx < W && gy < H) {
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;      // loss_utils.py:56-57
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
        const float s1 = xx - mu1_sq, s2 = yy - mu2_sq, s12 = xy - mu12;
        const float a = 2.f * mu12 + C1, b = 2.f * s12 + C2;
        const float c = mu1_sq + mu2_sq + C1, d = s1 + s2 + C2;
        const float rc = 1.f / c, rd = 1.f / d;
        m = (a * b) * (rc * rd);                                 // loss_utils.py:59
        const float px = sx[ty + SS_R][tx + SS_R], py = sy[ty + SS_R][tx + SS_R];
        l1 = fabsf(px - py);
        if (maps) {
            const size_t o = base + (size_t)gy * W + gx;
            // total derivative wrt mu1 (through a, b = 2(E[xy] - mu1 mu2) + C2, c, d = E[x^2] - mu1^2 + ...)
            maps[o]             = 2.f * mu2 * (b - a) * (rc * rd) + 2.f * mu1 * m * (rd - rc);
            maps[plane + o]     = -m * rd;                       // dm/dE[x^2]
            maps[2 * plane + o] = 2.f * a * (rc * rd);           // dm/dE[xy]
        }
    }
    const float bl1 = block_sum_256(l1, wsum);
    const float bm  = block_sum_256(m, wsum);
    if (tid == 0) { atomicAdd(&sums[0], bl1); atomicAdd(&sums[1], bm); }
}

__global__ void __launch_bounds__(256)
ssim_bwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, const float* __restrict__ maps, size_t plane,
                int H, int W, SsimWindow win, float k_l1, float k_ssim, float* __restrict__ grad)
{
    __shared__ float sm[3][SS_IN][SS_IN + 1];
    __shared__ float hz[3][SS_IN][SS_T + 1];

    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * SS_T - SS_R, y0 = blockIdx.y * SS_T - SS_R;

    for (int i = tid; i < SS_IN * SS_IN; i += 256) {
        const int r = i / SS_IN, c = i - r * SS_IN;
        const int gx = x0 + c, gy = y0 + r;
        float a = 0.f, b = 0.f, cc = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t o = base + (size_t)gy * W + gx;
            a = maps[o]; b = maps[plane + o]; cc = maps[2 * plane + o];
        }
        sm[0][r][c] = a; sm[1][r][c] = b; sm[2][r][c] = cc;
    }
    __syncthreads();
    for (int i = tid; i < SS_IN * SS_T; i += 256) {
        const int c = i / SS_IN, r = i - c * SS_IN;
        float a = 0.f, b = 0.f, cc = 0.f;
#pragma unroll
        for (int k = 0; k <= 2 * SS_R; k++) {
            const float w = win.w[k];
            a = fmaf(w, sm[0][r][c + k], a); b = fmaf(w, sm[1][r][c + k], b); cc = fmaf(w, sm[2][r][c + k], cc);
        }
        hz[0][r][c] = a; hz[1][r][c] = b; hz[2][r][c] = cc;
    }
    __syncthreads();
    const int tx = tid & 15, ty = tid >> 4;
    const int gx = blockIdx.x * SS_T + tx, gy = blockIdx.y * SS_T + ty;
    if (gx >= W || gy >= H) return;
    float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
    for (int k = 0; k <= 2 * SS_R; k++) {
        const float w = win.w[k];
        a = fmaf(w, hz[0][ty + k][tx], a); b = fmaf(w, hz[1][ty + k][tx], b); c = fmaf(w, hz[2][ty + k][tx], c);
    }
    const size_t o = base + (size_t)gy * W + gx;
    const float x = img1[o], y = img2[o];
    const float d = x - y;
    const float s = d > 0.f ? k_l1 : (d < 0.f ? -k_l1 : 0.f);
    grad[o] = fmaf(k_ssim, a + 2.f * x * b + y * c, s);
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_l1_ssim_loss(const float* image, const float* target, int C, int H, int W, float k_l1, float k_ssim,
                                float* sums, float* maps, float* grad, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!image || !target || !sums || C < 0 || H < 0 || W < 0) return GSR_EINVAL;
    if (grad && !maps) return GSR_EINVAL;
    GSR_CUDA(cudaMemsetAsync(sums, 0, 2 * sizeof(float), s));
    if (C == 0 || H == 0 || W == 0) return GSR_OK;
    if (C > 65535) return GSR_EINVAL;
    static const SsimWindow win = make_window();
    const size_t plane = (size_t)C * H * W;
    const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C);
    GSR_LAUNCH(ssim_fwd_kernel, grid, 256, 0, s, image, target, H, W, win, grad ? maps : nullptr, plane, sums);
    if (grad)
        GSR_LAUNCH(ssim_bwd_kernel, grid, 256, 0, s, image, target, (const float*)maps, plane, H, W, win, k_l1, k_ssim, grad);
    return GSR_OK;
}
