// math3.cuh - tiny 3x3 helpers with a FIXED scalar evaluation order.
//
// The bit-exact integer contract (radii, tile rects, depth bits) depends on fp32 rounding of
// the projection / covariance arithmetic.  The reference evaluates it through GLM column-major
// matrices (third_party/glm/glm/detail/type_mat3x3.inl:486-516): every product entry is the
// left-to-right sum  a[0][j]*b[i][0] + a[1][j]*b[i][1] + a[2][j]*b[i][2]  with literal-zero
// terms KEPT, and nvcc's default FMA contraction then fixes the rounding.  These helpers keep
// that expression order (own code, same arithmetic); they are compiled with the same nvcc
// defaults (-fmad=true, IEEE div/sqrt, no fast-math).
#pragma once
#include <cuda_runtime.h>

namespace gsr {

// column-major: c[i][j] = column i, row j
struct M3 {
    float c[3][3];
};

__device__ __forceinline__ M3 m3_cols(float a0, float a1, float a2, float b0, float b1, float b2,
                                      float c0, float c1, float c2)
{
    M3 r;
    r.c[0][0] = a0; r.c[0][1] = a1; r.c[0][2] = a2;
    r.c[1][0] = b0; r.c[1][1] = b1; r.c[1][2] = b2;
    r.c[2][0] = c0; r.c[2][1] = c1; r.c[2][2] = c2;
    return r;
}

__device__ __forceinline__ M3 m3_mul(const M3& a, const M3& b)
{
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            r.c[i][j] = a.c[0][j] * b.c[i][0] + a.c[1][j] * b.c[i][1] + a.c[2][j] * b.c[i][2];
    return r;
}

__device__ __forceinline__ M3 m3_t(const M3& a)
{
    M3 r;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) r.c[i][j] = a.c[j][i];
    return r;
}

// world->view / full projection, row-vector storage m[4*c + r]  (auxiliary.h:58-77)
__device__ __forceinline__ float3 xform4x3(const float3 p, const float* m)
{
    float3 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
    };
    return t;
}
__device__ __forceinline__ float4 xform4x4(const float3 p, const float* m)
{
    float4 t = {
        m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
        m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
        m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
        m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
    return t;
}

// SH basis constants (auxiliary.h:22-39)
#define GSR_SH_C0 0.28209479177387814f
#define GSR_SH_C1 0.4886025119029199f
#define GSR_SH_C2_0 1.0925484305920792f
#define GSR_SH_C2_1 -1.0925484305920792f
#define GSR_SH_C2_2 0.31539156525252005f
#define GSR_SH_C2_3 -1.0925484305920792f
#define GSR_SH_C2_4 0.5462742152960396f
#define GSR_SH_C3_0 -0.5900435899266435f
#define GSR_SH_C3_1 2.890611442640554f
#define GSR_SH_C3_2 -0.4570457994644658f
#define GSR_SH_C3_3 0.3731763325901154f
#define GSR_SH_C3_4 -0.4570457994644658f
#define GSR_SH_C3_5 1.445305721320277f
#define GSR_SH_C3_6 -0.5900435899266435f

}  // namespace gsr
