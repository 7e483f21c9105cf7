// activate.cu - the model's parameter activations, forward and backward, one launch each.
//
// Replaces GaussianModel.get_scaling / get_rotation / get_opacity (reference scene/gaussian_model.py:33-40,100-125):
//   scaling  = torch.exp(_scaling)                         [P,3]
//   rotation = torch.nn.functional.normalize(_rotation)    [P,4]   q / max(|q|_2, 1e-12)
//   opacity  = torch.sigmoid(_opacity)                     [P,1]   1 / (1 + exp(-x))
// (exp, sigmoid, norm, clamp_min, expand, div: six launches forward and about ten backward on P-sized arrays) by one
// kernel per direction: 32 B read + 32 B written per Gaussian forward; the backward re-derives the activations from the raw
// parameters (no saved outputs) and writes - or adds to - the raw parameters' gradients:
//   dL/d_scaling = g_s * exp(s);   dL/d_opacity = g_o * y (1 - y);   dL/d_rotation = (g_q - q^ (q^ . g_q)) / max(|q|, eps)
#include "gsr_common.cuh"

namespace gsr {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(256)
activate_fwd_kernel(int P, const float* __restrict__ s_raw, const float* __restrict__ q_raw, const float* __restrict__ o_raw,
                    float* __restrict__ scales, float* __restrict__ rots, float* __restrict__ opac)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    if (scales) {
        scales[3 * (size_t)i] = expf(s_raw[3 * (size_t)i]);
        scales[3 * (size_t)i + 1] = expf(s_raw[3 * (size_t)i + 1]);
        scales[3 * (size_t)i + 2] = expf(s_raw[3 * (size_t)i + 2]);
    }
    if (rots) {
        const float4 q = reinterpret_cast<const float4*>(q_raw)[i];
        const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
        reinterpret_cast<float4*>(rots)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    }
    if (opac) opac[i] = sigmoidf_(o_raw[i]);
}

__global__ void __launch_bounds__(256)
activate_bwd_kernel(int P, const float* __restrict__ s_raw, const float* __restrict__ q_raw, const float* __restrict__ o_raw,
                    const float* __restrict__ g_scales, const float* __restrict__ g_rots, const float* __restrict__ g_opac,
                    float* __restrict__ d_s, float* __restrict__ d_q, float* __restrict__ d_o, int accumulate)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    if (d_s) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const size_t k = 3 * (size_t)i + c;
            const float v = g_scales[k] * expf(s_raw[k]);
            d_s[k] = accumulate ? d_s[k] + v : v;
        }
    }
    if (d_q) {
        const float4 q = reinterpret_cast<const float4*>(q_raw)[i];
        const float4 g = reinterpret_cast<const float4*>(g_rots)[i];
        const float nr = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        const float n = fmaxf(nr, 1e-12f);
        const float inv = 1.0f / n;
        // y = q / n.  n = |q| when |q| >= eps: dy/dq = (I - y y^T) / n;  clamped (|q| < eps): n is a constant, dy/dq = I / n
        const float dot = (nr >= 1e-12f) ? (q.x * g.x + q.y * g.y + q.z * g.z + q.w * g.w) * inv * inv : 0.f;
        float4 v = make_float4((g.x - q.x * dot) * inv, (g.y - q.y * dot) * inv, (g.z - q.z * dot) * inv, (g.w - q.w * dot) * inv);
        float4* out = reinterpret_cast<float4*>(d_q) + i;
        if (accumulate) { const float4 o = *out; v = make_float4(o.x + v.x, o.y + v.y, o.z + v.z, o.w + v.w); }
        *out = v;
    }
    if (d_o) {
        const float y = sigmoidf_(o_raw[i]);
        const float v = g_opac[i] * ((1.0f - y) * y);
        d_o[i] = accumulate ? d_o[i] + v : v;
    }
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_activate_forward(int P, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                                    float* scales, float* rotations, float* opacities, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (P < 0) return GSR_EINVAL;
    if (P == 0) return GSR_OK;
    if ((scales && !scaling_raw) || (rotations && !rotation_raw) || (opacities && !opacity_raw)) return GSR_EINVAL;
    if ((((uintptr_t)rotation_raw | (uintptr_t)rotations) & 15) != 0) return GSR_EINVAL;
    GSR_LAUNCH(activate_fwd_kernel, (P + 255) / 256, 256, 0, s, P, scaling_raw, rotation_raw, opacity_raw, scales, rotations, opacities);
    return GSR_OK;
}

extern "C" int gsr_activate_backward(int P, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                                     const float* dL_dscales, const float* dL_drotations, const float* dL_dopacities,
                                     float* dL_dscaling_raw, float* dL_drotation_raw, float* dL_dopacity_raw, int accumulate,
                                     gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (P < 0) return GSR_EINVAL;
    if (P == 0) return GSR_OK;
    if ((dL_dscaling_raw && (!scaling_raw || !dL_dscales)) || (dL_drotation_raw && (!rotation_raw || !dL_drotations)) ||
        (dL_dopacity_raw && (!opacity_raw || !dL_dopacities)))
        return GSR_EINVAL;
    if ((((uintptr_t)rotation_raw | (uintptr_t)dL_drotations | (uintptr_t)dL_drotation_raw) & 15) != 0) return GSR_EINVAL;
    GSR_LAUNCH(activate_bwd_kernel, (P + 255) / 256, 256, 0, s, P, scaling_raw, rotation_raw, opacity_raw, dL_dscales, dL_drotations,
               dL_dopacities, dL_dscaling_raw, dL_drotation_raw, dL_dopacity_raw, accumulate);
    return GSR_OK;
}
