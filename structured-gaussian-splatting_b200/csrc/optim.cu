// optim.cu - fused Adam step over all parameter groups + densification statistics.
//
// Replaces, after the rasterizer's backward (reference train.py:129-142, scene/gaussian_model.py:159-177,415-417):
//   gaussians.optimizer.step()            torch.optim.Adam(l, lr=0.0, eps=1e-15) over six groups (xyz, f_dc, f_rest,
//                                         opacity, scaling, rotation), each with its own lr: seven foreach passes
//                                         (lerp, mul, addcmul, sqrt, div, add, addcdiv) over every group
//   max_radii2D[vis] = max(max_radii2D[vis], radii[vis]); add_densification_stats(viewspace_points, vis)
// by ONE launch for the Adam step of all groups (28 B per element: read p, g, m, v, write p, m, v) and one for the
// statistics.  The arithmetic is torch's, operation for operation, in fp32:
//   m = m + (1 - b1) (g - m)                       _foreach_lerp_
//   v = v b2 + (1 - b2) g g                        _foreach_mul_, _foreach_addcmul_
//   p = p - step_size * m / (sqrt(v) / sqrt(1 - b2^t) + eps),   step_size = lr / (1 - b1^t)
// with the bias corrections evaluated in double on the host like torch does.  `visible` (optional) restricts the update
// to rows whose flag is set - the "sparse Adam" of later 3DGS code bases; NULL is the reference's dense step.
#include "gsr_common.cuh"
#include <cmath>

namespace gsr {

constexpr int ADAM_MAX_GROUPS = 8;
constexpr int ADAM_ELEMS_PER_BLOCK = 256 * 4 * 4;       // 256 threads x 4 float4

struct AdamGroupDev {
    float* p; const float* g; float* m; float* v;
    unsigned long long n;
    unsigned first_block;
    int row;
    float lr_neg_step, bc2_sqrt;
};
struct AdamArgs {
    AdamGroupDev grp[ADAM_MAX_GROUPS];
    int n_groups;
    float b2, omb1, omb2, eps;
    const unsigned char* visible;
};

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const AdamArgs& a, float neg_step, float bc2s)
{
    m = fmaf(a.omb1, g - m, m);
    v = fmaf(a.omb2 * g, g, v * a.b2);
    const float denom = __fdiv_rn(sqrtf(v), bc2s) + a.eps;
    p = fmaf(neg_step, __fdiv_rn(m, denom), p);
}

__global__ void __launch_bounds__(256)
adam_step_kernel(const __grid_constant__ AdamArgs a)
{
    int gi = 0;
#pragma unroll
    for (int i = 1; i < ADAM_MAX_GROUPS; i++)
        if (i < a.n_groups && blockIdx.x >= a.grp[i].first_block) gi = i;
    const AdamGroupDev& G = a.grp[gi];
    const unsigned long long base = (unsigned long long)(blockIdx.x - G.first_block) * ADAM_ELEMS_PER_BLOCK;
    const float ns = G.lr_neg_step, bc = G.bc2_sqrt;
    const bool vec = ((((uintptr_t)G.p | (uintptr_t)G.g | (uintptr_t)G.m | (uintptr_t)G.v) & 15) == 0);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const unsigned long long e = base + ((unsigned long long)k * 256 + threadIdx.x) * 4;
        if (e >= G.n) break;
        if (vec && e + 4 <= G.n) {
            bool live[4] = {true, true, true, true};
            if (a.visible) {
#pragma unroll
                for (int j = 0; j < 4; j++) live[j] = a.visible[(e + j) / (unsigned)G.row] != 0;
                if (!(live[0] | live[1] | live[2] | live[3])) continue;
            }
            float4 p = *reinterpret_cast<const float4*>(G.p + e);
            const float4 g = *reinterpret_cast<const float4*>(G.g + e);
            float4 m = *reinterpret_cast<const float4*>(G.m + e);
            float4 v = *reinterpret_cast<const float4*>(G.v + e);
            if (live[0]) adam_elem(p.x, g.x, m.x, v.x, a, ns, bc);
            if (live[1]) adam_elem(p.y, g.y, m.y, v.y, a, ns, bc);
            if (live[2]) adam_elem(p.z, g.z, m.z, v.z, a, ns, bc);
            if (live[3]) adam_elem(p.w, g.w, m.w, v.w, a, ns, bc);
            *reinterpret_cast<float4*>(G.p + e) = p;
            *reinterpret_cast<float4*>(G.m + e) = m;
            *reinterpret_cast<float4*>(G.v + e) = v;
        } else {
            for (unsigned long long i = e; i < e + 4 && i < G.n; i++) {
                if (a.visible && !a.visible[i / (unsigned)G.row]) continue;
                float p = G.p[i], m = G.m[i], v = G.v[i];
                adam_elem(p, G.g[i], m, v, a, ns, bc);
                G.p[i] = p; G.m[i] = m; G.v[i] = v;
            }
        }
    }
}

// train.py:131-132 + gaussian_model.py:415-417, for update_filter = radii > 0
__global__ void __launch_bounds__(256)
densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dmean2D, float* __restrict__ max_radii2D,
                     float* __restrict__ grad_accum, float* __restrict__ denom)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    const float gx = dmean2D[3 * (size_t)i], gy = dmean2D[3 * (size_t)i + 1];
    grad_accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.f;
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_adam_step(const GsrAdamGroup* groups, int n_groups, double beta1, double beta2, double eps,
                             const unsigned char* visible, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (n_groups < 0 || n_groups > ADAM_MAX_GROUPS || (n_groups > 0 && !groups)) return GSR_EINVAL;
    AdamArgs a{};
    a.n_groups = 0;
    a.b2 = (float)beta2; a.omb1 = (float)(1.0 - beta1); a.omb2 = (float)(1.0 - beta2); a.eps = (float)eps;
    a.visible = visible;
    unsigned long long blocks = 0;
    for (int i = 0; i < n_groups; i++) {
        const GsrAdamGroup& g = groups[i];
        if (g.n == 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq || g.step < 1 || g.row < 1) return GSR_EINVAL;
        const double bc1 = 1.0 - std::pow(beta1, (double)g.step);
        const double bc2 = 1.0 - std::pow(beta2, (double)g.step);
        AdamGroupDev& d = a.grp[a.n_groups++];
        d.p = g.param; d.g = g.grad; d.m = g.exp_avg; d.v = g.exp_avg_sq; d.n = g.n; d.row = g.row;
        d.first_block = (unsigned)blocks;
        d.lr_neg_step = (float)(-((double)g.lr / bc1));
        d.bc2_sqrt = (float)std::sqrt(bc2);
        blocks += (g.n + ADAM_ELEMS_PER_BLOCK - 1) / ADAM_ELEMS_PER_BLOCK;
    }
    if (blocks == 0) return GSR_OK;
    if (blocks > 0x7fffffffull) return GSR_EINVAL;
    GSR_LAUNCH(adam_step_kernel, (unsigned)blocks, 256, 0, s, a);
    return GSR_OK;
}

extern "C" int gsr_densify_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D, float* xyz_gradient_accum,
                                 float* denom, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (P < 0) return GSR_EINVAL;
    if (P == 0) return GSR_OK;
    if (!radii || !dL_dmean2D || !max_radii2D || !xyz_gradient_accum || !denom) return GSR_EINVAL;
    GSR_LAUNCH(densify_stats_kernel, (P + 255) / 256, 256, 0, s, P, radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom);
    return GSR_OK;
}
