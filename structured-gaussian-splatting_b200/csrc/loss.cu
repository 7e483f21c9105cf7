// loss.cu - fused L1 image loss, forward and gradient in one pass (first piece).
//
// Replaces, for the training step around the rasterizer (reference utils/loss_utils.py:17-18, train.py:104):
//   l1_loss(network_output, gt) = torch.abs(network_output - gt).mean()
// and its autograd backward (sub, abs, mean, then sign * grad / n: six elementwise launches and three 25 MB
// temporaries at 1080p) by one kernel that reads both images once, writes dL/dimage = sign(image - gt) / n and
// adds the block sums of |image - gt| to one device scalar.  HBM-bound: 12 B per element.
#include "gsr_common.cuh"

namespace gsr {

__global__ void __launch_bounds__(256)
l1_loss_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ grad, size_t n, float scale,
               float* __restrict__ sum)
{
    __shared__ float wsum[8];
    float acc = 0.f;
    const size_t n4 = n >> 2;
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* b4 = reinterpret_cast<const float4*>(b);
    float4* g4 = reinterpret_cast<float4*>(grad);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 x = a4[i], y = b4[i];
        const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
        acc += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3);
        if (grad) {
            float4 g;                         // torch: d|x|/dx = sign(x), 0 at 0
            g.x = d0 > 0.f ? scale : (d0 < 0.f ? -scale : 0.f);
            g.y = d1 > 0.f ? scale : (d1 < 0.f ? -scale : 0.f);
            g.z = d2 > 0.f ? scale : (d2 < 0.f ? -scale : 0.f);
            g.w = d3 > 0.f ? scale : (d3 < 0.f ? -scale : 0.f);
            g4[i] = g;
        }
    }
    if (blockIdx.x == 0) {                    // tail (n not a multiple of 4)
        for (size_t i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) {
            const float d = a[i] - b[i];
            acc += fabsf(d);
            if (grad) grad[i] = d > 0.f ? scale : (d < 0.f ? -scale : 0.f);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 8) {
        float v = wsum[threadIdx.x];
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
        if (threadIdx.x == 0) atomicAdd(sum, v);
    }
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_l1_loss(const float* image, const float* target, size_t n, float grad_scale, float* loss_sum, float* grad,
                           gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!image || !target || !loss_sum) return GSR_EINVAL;
    if ((((uintptr_t)image | (uintptr_t)target | (uintptr_t)grad) & 15) != 0) return GSR_EINVAL;
    GSR_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), s));
    if (n == 0) return GSR_OK;
    const size_t n4 = n >> 2;
    int blocks = (int)((n4 + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    GSR_LAUNCH(l1_loss_kernel, blocks, 256, 0, s, image, target, grad, n, grad_scale, loss_sum);
    return GSR_OK;
}
