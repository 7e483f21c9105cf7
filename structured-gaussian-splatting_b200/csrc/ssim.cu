// ssim.cu - fused L1 + D-SSIM image loss of the training step.
//
// Replaces, around the rasterizer (reference train.py:104-105, utils/loss_utils.py:17-63):
//   Ll1  = l1_loss(image, gt)
//   loss = (1 - lambda_dssim) * Ll1 + lambda_dssim * (1 - ssim(image, gt))
// where ssim() is five depthwise 11x11 Gaussian convolutions (sigma 1.5, zero padding 5) of x, y, x*x, y*y, x*y,
// ~15 elementwise launches on 3xHxW temporaries, and the autograd backward of all of it.  Here: two kernels.
//   ssim_fwd_kernel  one CTA per 32x16 pixel tile and channel: stages the 42x26 halo tile of both images in shared memory,
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

constexpr int SS_TX = 32, SS_TY = 16;     // output tile
constexpr int SS_R  = 5;                  // window radius (window_size 11, loss_utils.py:39)
constexpr int SS_IW = SS_TX + 2 * SS_R;   // 42: staged width
constexpr int SS_IH = SS_TY + 2 * SS_R;   // 26: staged height
constexpr int SS_S  = 43;                 // staged row stride: odd and 3 mod 4, see the horizontal pass
constexpr int SS_HS = 33;                 // row stride of the horizontally filtered planes
constexpr int SS_HW = SS_IH * (SS_TX / 4);   // 208 horizontal work items: (row, group of 4 columns)

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

// Both kernels share one structure.  (1) The 42x26 halo tile goes to shared memory (zero outside the image = the
// convolution's zero padding).  (2) Horizontal taps with a register sliding window: a thread owns one staged row and 4
// adjacent output columns, loads the 14 values under them once and forms the 4 x 11-tap sums from registers (28 shared
// loads per 4 outputs instead of 88).  Lanes run over the 8 column groups of a row (bank 4 g + j) and 4 rows per warp whose
// offsets 43 r are 0, 3, 2, 1 mod 4: all 32 lanes hit different banks.  (3) Vertical taps: a thread owns 2 vertically
// adjacent pixels and slides over the 12 rows under them (lanes = 32 consecutive columns: conflict-free).

__global__ void __launch_bounds__(256)
ssim_fwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W, SsimWindow win,
                float* __restrict__ maps, size_t plane, float* __restrict__ sums)
{
    __shared__ float sx[SS_IH][SS_S];
    __shared__ float sy[SS_IH][SS_S];
    __shared__ float hz[5][SS_IH][SS_HS];
    __shared__ float wsum[8];

    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * SS_TX - SS_R, y0 = blockIdx.y * SS_TY - SS_R;

    for (int i = tid; i < SS_IH * SS_IW; i += 256) {
        const int r = i / SS_IW, c = i - r * SS_IW;
        const int gx = x0 + c, gy = y0 + r;
        float a = 0.f, b = 0.f;                    // zero padding (F.conv2d padding=5)
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t o = base + (size_t)gy * W + gx;
            a = img1[o]; b = img2[o];
        }
        sx[r][c] = a; sy[r][c] = b;
    }
    __syncthreads();

    if (tid < SS_HW) {
        const int r = tid >> 3, c0 = (tid & 7) * 4;
        float a[14], b[14];
#pragma unroll
        for (int j = 0; j < 14; j++) { a[j] = sx[r][c0 + j]; b[j] = sy[r][c0 + j]; }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            float m1 = 0.f, m2 = 0.f, xx = 0.f, yy = 0.f, xy = 0.f;
#pragma unroll
            for (int k = 0; k <= 2 * SS_R; k++) {
                const float w = win.w[k], wa = w * a[u + k], wb = w * b[u + k];
                m1 += wa; m2 += wb;
                xx = fmaf(wa, a[u + k], xx); yy = fmaf(wb, b[u + k], yy); xy = fmaf(wa, b[u + k], xy);
            }
            hz[0][r][c0 + u] = m1; hz[1][r][c0 + u] = m2; hz[2][r][c0 + u] = xx; hz[3][r][c0 + u] = yy; hz[4][r][c0 + u] = xy;
        }
    }
    __syncthreads();

    const int tx = tid & 31, ty = (tid >> 5) * 2;
    float acc[2][5];
#pragma unroll
    for (int u = 0; u < 2; u++)
#pragma unroll
        for (int q = 0; q < 5; q++) acc[u][q] = 0.f;
#pragma unroll
    for (int k = 0; k < 12; k++) {
#pragma unroll
        for (int q = 0; q < 5; q++) {
            const float v = hz[q][ty + k][tx];
            if (k < 11) acc[0][q] = fmaf(win.w[k < 11 ? k : 0], v, acc[0][q]);
            if (k > 0) acc[1][q] = fmaf(win.w[k > 0 ? k - 1 : 0], v, acc[1][q]);
        }
    }
    const int gx = blockIdx.x * SS_TX + tx;
    float m_sum = 0.f, l1 = 0.f;
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int gy = blockIdx.y * SS_TY + ty + u;
        if (gx < W && gy < H) {
            const float mu1 = acc[u][0], mu2 = acc[u][1], xx = acc[u][2], yy = acc[u][3], xy = acc[u][4];
            const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;      // loss_utils.py:56-57
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = xx - mu1_sq, s2 = yy - mu2_sq, s12 = xy - mu12;
            const float a = 2.f * mu12 + C1, b = 2.f * s12 + C2;
            const float c = mu1_sq + mu2_sq + C1, d = s1 + s2 + C2;
            const float rc = 1.f / c, rd = 1.f / d;
            const float m = (a * b) * (rc * rd);                     // loss_utils.py:59
            m_sum += m;
            l1 += fabsf(sx[ty + u + SS_R][tx + SS_R] - sy[ty + u + SS_R][tx + SS_R]);
            if (maps) {
                const size_t o = base + (size_t)gy * W + gx;
                // total derivative wrt mu1 (through a, b = 2(E[xy] - mu1 mu2) + C2, c, d = E[x^2] - mu1^2 + ...)
                maps[o]             = 2.f * mu2 * (b - a) * (rc * rd) + 2.f * mu1 * m * (rd - rc);
                maps[plane + o]     = -m * rd;                       // dm/dE[x^2]
                maps[2 * plane + o] = 2.f * a * (rc * rd);           // dm/dE[xy]
            }
        }
    }
    const float bl1 = block_sum_256(l1, wsum);
    const float bm  = block_sum_256(m_sum, wsum);
    if (tid == 0) { atomicAdd(&sums[0], bl1); atomicAdd(&sums[1], bm); }
}

__global__ void __launch_bounds__(256)
ssim_bwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, const float* __restrict__ maps, size_t plane,
                int H, int W, SsimWindow win, float k_l1, float k_ssim, float* __restrict__ grad)
{
    __shared__ float sm[3][SS_IH][SS_S];
    __shared__ float hz[3][SS_IH][SS_HS];

    const int tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * SS_TX - SS_R, y0 = blockIdx.y * SS_TY - SS_R;

    for (int i = tid; i < SS_IH * SS_IW; i += 256) {
        const int r = i / SS_IW, c = i - r * SS_IW;
        const int gx = x0 + c, gy = y0 + r;
        float a = 0.f, b = 0.f, cc = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t o = base + (size_t)gy * W + gx;
            a = maps[o]; b = maps[plane + o]; cc = maps[2 * plane + o];
        }
        sm[0][r][c] = a; sm[1][r][c] = b; sm[2][r][c] = cc;
    }
    __syncthreads();
    if (tid < SS_HW) {
        const int r = tid >> 3, c0 = (tid & 7) * 4;
#pragma unroll
        for (int q = 0; q < 3; q++) {
            float v[14];
#pragma unroll
            for (int j = 0; j < 14; j++) v[j] = sm[q][r][c0 + j];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                float a = 0.f;
#pragma unroll
                for (int k = 0; k <= 2 * SS_R; k++) a = fmaf(win.w[k], v[u + k], a);
                hz[q][r][c0 + u] = a;
            }
        }
    }
    __syncthreads();
    const int tx = tid & 31, ty = (tid >> 5) * 2;
    float acc[2][3];
#pragma unroll
    for (int u = 0; u < 2; u++)
#pragma unroll
        for (int q = 0; q < 3; q++) acc[u][q] = 0.f;
#pragma unroll
    for (int k = 0; k < 12; k++) {
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const float v = hz[q][ty + k][tx];
            if (k < 11) acc[0][q] = fmaf(win.w[k < 11 ? k : 0], v, acc[0][q]);
            if (k > 0) acc[1][q] = fmaf(win.w[k > 0 ? k - 1 : 0], v, acc[1][q]);
        }
    }
    const int gx = blockIdx.x * SS_TX + tx;
#pragma unroll
    for (int u = 0; u < 2; u++) {
        const int gy = blockIdx.y * SS_TY + ty + u;
        if (gx < W && gy < H) {
            const size_t o = base + (size_t)gy * W + gx;
            const float x = img1[o], y = img2[o];
            const float d = x - y;
            const float s = d > 0.f ? k_l1 : (d < 0.f ? -k_l1 : 0.f);
            grad[o] = fmaf(k_ssim, acc[u][0] + 2.f * x * acc[u][1] + y * acc[u][2], s);
        }
    }
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
    const dim3 grid((W + SS_TX - 1) / SS_TX, (H + SS_TY - 1) / SS_TY, C);
    GSR_LAUNCH(ssim_fwd_kernel, grid, 256, 0, s, image, target, H, W, win, grad ? maps : nullptr, plane, sums);
    if (grad)
        GSR_LAUNCH(ssim_bwd_kernel, grid, 256, 0, s, image, target, (const float*)maps, plane, H, W, win, k_l1, k_ssim, grad);
    return GSR_OK;
}
