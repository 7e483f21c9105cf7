// preprocess_bwd.cu - per-Gaussian backward of the preprocess stage, ONE fused kernel.
//
// Replaces (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer/backward.cu):
//   computeCov2DCUDA            :144-274   dL/dconic -> dL/dcov3D, dL/dmean3D (through J, clamp masks)
//   preprocessCUDA<3> (bwd)     :346-396   dL/dmean2D -> dL/dmean3D, SH backward (:20-139), cov3D -> scale/rot (:278-341)
// plus the nine torch::zeros of the reference binding (rasterize_points.cu:151-159): this kernel WRITES
// every element of every gradient it owns (zeros for culled Gaussians), so the binding allocates with
// empty() and no 300 B/Gaussian memset is needed.
//
// Input is the packed accumulator row written by the blending backward (12 floats per Gaussian:
// 0-2 dL/dcolour, 3-4 sum w*(dx,dy), 5-7 sum w*(dx^2, dx dy, dy^2), 8 sum w, with w = dL/dG * G), turned
// into dL/dmean2D, dL/dconic, dL/dopacity here with the Gaussian's conic and opacity.  cov3D is recomputed from
// scale/rotation instead of being saved by the forward.  Gradients are fp32 with a 1e-3 relative
// tolerance, so the arithmetic here is free to use a cheaper operation order than the reference.
#include "gsr_common.cuh"
#include "math3.cuh"

namespace gsr {

struct PreBwdArgs {
    int P, D, M;
    const float* means;
    const int* radii;
    const float* shs;
    const uint8_t* clamped;
    const float* scales;
    const float* rots;
    const float* cov_pre;
    float mod;
    const float* accum;      // [P,12] moment rows written by the blending backward
    const SplatRec* rec;     // forward records: conic and opacity turn the moments into gradients
    float* dL_dmean2D;       // [P,3]
    float* dL_dopacity;      // [P]
    float* dL_dcolor;        // [P,3] or null
    float* dL_dmean3D;       // [P,3]
    float* dL_dcov3D;        // [P,6] or null
    float* dL_dsh;           // [P,M,3] or null
    float* dL_dscale;        // [P,3] or null
    float* dL_drot;          // [P,4] or null
};

__device__ __forceinline__ void store_sh_grads(float* __restrict__ dst, const float* g, int nfloats, int total_floats)
{
    // g holds nfloats live values, the rest of the row is zero
    const bool vec = ((((size_t)dst) & 15) == 0) && (total_floats % 4 == 0);
    if (vec) {
        float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
        for (int i = 0; i < 12; i++) {
            if (4 * i < total_floats) {
                float4 v;
                v.x = 4 * i + 0 < nfloats ? g[4 * i + 0] : 0.f;
                v.y = 4 * i + 1 < nfloats ? g[4 * i + 1] : 0.f;
                v.z = 4 * i + 2 < nfloats ? g[4 * i + 2] : 0.f;
                v.w = 4 * i + 3 < nfloats ? g[4 * i + 3] : 0.f;
                d4[i] = v;
            }
        }
        for (int i = 48; i < total_floats; i++) dst[i] = 0.f;
    } else {
#pragma unroll
        for (int i = 0; i < 48; i++)
            if (i < total_floats) dst[i] = i < nfloats ? g[i] : 0.f;
        for (int i = 48; i < total_floats; i++) dst[i] = 0.f;
    }
}

__global__ void __launch_bounds__(256) preprocess_bwd_kernel(PreBwdArgs a, CameraArgs ca)
{
    __shared__ Camera cam;
    load_camera(cam, ca);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    const int M3f = 3 * a.M;

    if (!(a.radii[idx] > 0)) {
        // culled: every owned gradient row is zero
        a.dL_dmean2D[3 * idx] = 0.f; a.dL_dmean2D[3 * idx + 1] = 0.f; a.dL_dmean2D[3 * idx + 2] = 0.f;
        a.dL_dopacity[idx] = 0.f;
        a.dL_dmean3D[3 * idx] = 0.f; a.dL_dmean3D[3 * idx + 1] = 0.f; a.dL_dmean3D[3 * idx + 2] = 0.f;
        if (a.dL_dcolor) { a.dL_dcolor[3 * idx] = 0.f; a.dL_dcolor[3 * idx + 1] = 0.f; a.dL_dcolor[3 * idx + 2] = 0.f; }
        if (a.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = 0.f;
        }
        if (a.dL_dsh) {
            float z[1] = {0.f};
            store_sh_grads(a.dL_dsh + (size_t)M3f * idx, z, 0, M3f);
        }
        if (a.dL_dscale) { a.dL_dscale[3 * idx] = 0.f; a.dL_dscale[3 * idx + 1] = 0.f; a.dL_dscale[3 * idx + 2] = 0.f; }
        if (a.dL_drot) {
            a.dL_drot[4 * idx] = 0.f; a.dL_drot[4 * idx + 1] = 0.f; a.dL_drot[4 * idx + 2] = 0.f; a.dL_drot[4 * idx + 3] = 0.f;
        }
        return;
    }

    // accumulator row: three 16-byte loads
    const float4* ar = reinterpret_cast<const float4*>(a.accum + (size_t)12 * idx);
    const float4 A0 = ar[0], A1 = ar[1], A2 = ar[2];
    const float dcol[3] = {A0.x, A0.y, A0.z};
    // moments -> gradients (backward.cu:536-551 with the per-Gaussian constants factored out of the pixel sums)
    const float4 r0 = a.rec[idx].q0, r1 = a.rec[idx].q1;
    const float cA = r0.z, cB = r0.w, cC = r1.x, opac = r1.y;
    const float g2x = -0.5f * cam.W * (cA * A0.w + cB * A1.x);
    const float g2y = -0.5f * cam.H * (cC * A1.x + cB * A0.w);
    const float dcon_x = -0.5f * A1.y, dcon_y = -0.5f * A1.z, dcon_w = -0.5f * A1.w;
    a.dL_dmean2D[3 * idx] = g2x; a.dL_dmean2D[3 * idx + 1] = g2y; a.dL_dmean2D[3 * idx + 2] = 0.f;
    a.dL_dopacity[idx] = opac > 0.f ? A2.x / opac : 0.f;
    if (a.dL_dcolor) { a.dL_dcolor[3 * idx] = dcol[0]; a.dL_dcolor[3 * idx + 1] = dcol[1]; a.dL_dcolor[3 * idx + 2] = dcol[2]; }

    const float3 p = {a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]};

    // ---- 3D covariance (recomputed) and the rotation / scale factors it came from
    float c6[6];
    float3 sc = {0.f, 0.f, 0.f};
    float4 q = {1.f, 0.f, 0.f, 0.f};
    M3 R;        // R.c[i][j] as in the forward (GLM column-major ctor order)
    if (a.cov_pre) {
#pragma unroll
        for (int k = 0; k < 6; k++) c6[k] = a.cov_pre[6 * (size_t)idx + k];
    } else {
        sc = make_float3(a.mod * a.scales[3 * idx], a.mod * a.scales[3 * idx + 1], a.mod * a.scales[3 * idx + 2]);
        q = make_float4(a.rots[4 * idx], a.rots[4 * idx + 1], a.rots[4 * idx + 2], a.rots[4 * idx + 3]);
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                    2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                    2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
        // M = S * R  (M.c[i][j] = s_j * R.c[i][j]);  Sigma = M^T M
        M3 Mm;
#pragma unroll
        for (int i = 0; i < 3; i++) { Mm.c[i][0] = sc.x * R.c[i][0]; Mm.c[i][1] = sc.y * R.c[i][1]; Mm.c[i][2] = sc.z * R.c[i][2]; }
        M3 Sig = m3_mul(m3_t(Mm), Mm);
        c6[0] = Sig.c[0][0]; c6[1] = Sig.c[0][1]; c6[2] = Sig.c[0][2];
        c6[3] = Sig.c[1][1]; c6[4] = Sig.c[1][2]; c6[5] = Sig.c[2][2];
    }

    // ---- cov2D backward (backward.cu:144-274)
    float3 t = xform4x3(p, cam.view);
    const float limx = 1.3f * cam.tanx, limy = 1.3f * cam.tany;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;

    M3 J = m3_cols(cam.fx / t.z, 0.0f, -(cam.fx * t.x) / (t.z * t.z),
                   0.0f, cam.fy / t.z, -(cam.fy * t.y) / (t.z * t.z),
                   0.0f, 0.0f, 0.0f);
    M3 Wm = m3_cols(cam.view[0], cam.view[4], cam.view[8], cam.view[1], cam.view[5], cam.view[9],
                    cam.view[2], cam.view[6], cam.view[10]);
    M3 Vrk = m3_cols(c6[0], c6[1], c6[2], c6[1], c6[3], c6[4], c6[2], c6[4], c6[5]);
    M3 T = m3_mul(Wm, J);
    M3 cov2D = m3_mul(m3_mul(m3_t(T), m3_t(Vrk)), T);

    const float ca_ = cov2D.c[0][0] + 0.3f, cb = cov2D.c[0][1], cc = cov2D.c[1][1] + 0.3f;
    const float denom = ca_ * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dcon_x + 2 * cb * cc * dcon_y + (denom - ca_ * cc) * dcon_w);
        dL_dc = denom2inv * (-ca_ * ca_ * dcon_w + 2 * ca_ * cb * dcon_y + (denom - ca_ * cc) * dcon_x);
        dL_db = denom2inv * 2 * (cb * cc * dcon_x - (denom + 2 * cb * cb) * dcon_y + ca_ * cb * dcon_w);
        dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
    }
    if (a.dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    }

    // dL/dT (upper 2x3), then dL/dJ, then dL/dt
    float tv0[3], tv1[3];   // T row products with Vrk: tvk[j] = sum_m T[k][m] * Vrk[j][m]
#pragma unroll
    for (int j = 0; j < 3; j++) {
        tv0[j] = T.c[0][0] * Vrk.c[j][0] + T.c[0][1] * Vrk.c[j][1] + T.c[0][2] * Vrk.c[j][2];
        tv1[j] = T.c[1][0] * Vrk.c[j][0] + T.c[1][1] * Vrk.c[j][1] + T.c[1][2] * Vrk.c[j][2];
    }
    const float dT00 = 2 * tv0[0] * dL_da + tv1[0] * dL_db, dT01 = 2 * tv0[1] * dL_da + tv1[1] * dL_db,
                dT02 = 2 * tv0[2] * dL_da + tv1[2] * dL_db;
    const float dT10 = 2 * tv1[0] * dL_dc + tv0[0] * dL_db, dT11 = 2 * tv1[1] * dL_dc + tv0[1] * dL_db,
                dT12 = 2 * tv1[2] * dL_dc + tv0[2] * dL_db;
    const float dJ00 = Wm.c[0][0] * dT00 + Wm.c[0][1] * dT01 + Wm.c[0][2] * dT02;
    const float dJ02 = Wm.c[2][0] * dT00 + Wm.c[2][1] * dT01 + Wm.c[2][2] * dT02;
    const float dJ11 = Wm.c[1][0] * dT10 + Wm.c[1][1] * dT11 + Wm.c[1][2] * dT12;
    const float dJ12 = Wm.c[2][0] * dT10 + Wm.c[2][1] * dT11 + Wm.c[2][2] * dT12;
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dtx = x_grad_mul * -cam.fx * tz2 * dJ02;
    const float dty = y_grad_mul * -cam.fy * tz2 * dJ12;
    const float dtz = -cam.fx * tz2 * dJ00 - cam.fy * tz2 * dJ11 + (2 * cam.fx * t.x) * tz3 * dJ02 + (2 * cam.fy * t.y) * tz3 * dJ12;
    // transformVec4x3Transpose (auxiliary.h:89-97); this is the ASSIGNMENT of backward.cu:273
    float dmx = cam.view[0] * dtx + cam.view[1] * dty + cam.view[2] * dtz;
    float dmy = cam.view[4] * dtx + cam.view[5] * dty + cam.view[6] * dtz;
    float dmz = cam.view[8] * dtx + cam.view[9] * dty + cam.view[10] * dtz;

    // ---- projection backward (backward.cu:369-387)
    {
        const float* PR = cam.proj;
        const float4 mh = xform4x4(p, PR);
        const float mw = 1.0f / (mh.w + 0.0000001f);
        const float mul1 = (PR[0] * p.x + PR[4] * p.y + PR[8] * p.z + PR[12]) * mw * mw;
        const float mul2 = (PR[1] * p.x + PR[5] * p.y + PR[9] * p.z + PR[13]) * mw * mw;
        dmx += (PR[0] * mw - PR[3] * mul1) * g2x + (PR[1] * mw - PR[3] * mul2) * g2y;
        dmy += (PR[4] * mw - PR[7] * mul1) * g2x + (PR[5] * mw - PR[7] * mul2) * g2y;
        dmz += (PR[8] * mw - PR[11] * mul1) * g2x + (PR[9] * mw - PR[11] * mul2) * g2y;
    }

    // ---- SH backward (backward.cu:20-139)
    if (a.shs) {
        const float ox = p.x - cam.cam[0], oy = p.y - cam.cam[1], oz = p.z - cam.cam[2];
        const float s2 = ox * ox + oy * oy + oz * oz;
        const float len = sqrtf(s2);
        const float x = ox / len, y = oy / len, z = oz / len;
        const uint32_t cl = a.clamped[idx];
        const float dRGB[3] = {(cl & 1) ? 0.f : dcol[0], (cl & 2) ? 0.f : dcol[1], (cl & 4) ? 0.f : dcol[2]};
        const float* shp = a.shs + (size_t)M3f * idx;
        const bool vec = (((size_t)shp) & 15) == 0 && (M3f % 4 == 0);
        float f[48];
        const int ncoef = (a.D + 1) * (a.D + 1);
        // load only the live coefficients
        if (a.D > 0) {
            if (vec && a.D == 3) {
                const float4* p4 = reinterpret_cast<const float4*>(shp);
#pragma unroll
                for (int i = 0; i < 12; i++) { float4 v = __ldg(p4 + i); f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w; }
            } else {
#pragma unroll
                for (int i = 0; i < 48; i++) f[i] = i < 3 * ncoef ? __ldg(shp + i) : 0.f;
            }
        }
        float w[16];         // basis value per coefficient
        float gx[3] = {0, 0, 0}, gy[3] = {0, 0, 0}, gz[3] = {0, 0, 0};
#pragma unroll
        for (int k = 0; k < 16; k++) w[k] = 0.f;
        w[0] = GSR_SH_C0;
#define S_(k, ch) f[3 * (k) + (ch)]
        if (a.D > 0) {
            w[1] = -GSR_SH_C1 * y; w[2] = GSR_SH_C1 * z; w[3] = -GSR_SH_C1 * x;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) { gx[ch] = -GSR_SH_C1 * S_(3, ch); gy[ch] = -GSR_SH_C1 * S_(1, ch); gz[ch] = GSR_SH_C1 * S_(2, ch); }
            if (a.D > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                w[4] = GSR_SH_C2_0 * xy; w[5] = GSR_SH_C2_1 * yz; w[6] = GSR_SH_C2_2 * (2.f * zz - xx - yy);
                w[7] = GSR_SH_C2_3 * xz; w[8] = GSR_SH_C2_4 * (xx - yy);
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    gx[ch] += GSR_SH_C2_0 * y * S_(4, ch) + GSR_SH_C2_2 * 2.f * -x * S_(6, ch) + GSR_SH_C2_3 * z * S_(7, ch) + GSR_SH_C2_4 * 2.f * x * S_(8, ch);
                    gy[ch] += GSR_SH_C2_0 * x * S_(4, ch) + GSR_SH_C2_1 * z * S_(5, ch) + GSR_SH_C2_2 * 2.f * -y * S_(6, ch) + GSR_SH_C2_4 * 2.f * -y * S_(8, ch);
                    gz[ch] += GSR_SH_C2_1 * y * S_(5, ch) + GSR_SH_C2_2 * 2.f * 2.f * z * S_(6, ch) + GSR_SH_C2_3 * x * S_(7, ch);
                }
                if (a.D > 2) {
                    w[9] = GSR_SH_C3_0 * y * (3.f * xx - yy); w[10] = GSR_SH_C3_1 * xy * z;
                    w[11] = GSR_SH_C3_2 * y * (4.f * zz - xx - yy); w[12] = GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                    w[13] = GSR_SH_C3_4 * x * (4.f * zz - xx - yy); w[14] = GSR_SH_C3_5 * z * (xx - yy);
                    w[15] = GSR_SH_C3_6 * x * (xx - 3.f * yy);
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        gx[ch] += (GSR_SH_C3_0 * S_(9, ch) * 3.f * 2.f * xy + GSR_SH_C3_1 * S_(10, ch) * yz + GSR_SH_C3_2 * S_(11, ch) * -2.f * xy +
                                   GSR_SH_C3_3 * S_(12, ch) * -3.f * 2.f * xz + GSR_SH_C3_4 * S_(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                                   GSR_SH_C3_5 * S_(14, ch) * 2.f * xz + GSR_SH_C3_6 * S_(15, ch) * 3.f * (xx - yy));
                        gy[ch] += (GSR_SH_C3_0 * S_(9, ch) * 3.f * (xx - yy) + GSR_SH_C3_1 * S_(10, ch) * xz +
                                   GSR_SH_C3_2 * S_(11, ch) * (-3.f * yy + 4.f * zz - xx) + GSR_SH_C3_3 * S_(12, ch) * -3.f * 2.f * yz +
                                   GSR_SH_C3_4 * S_(13, ch) * -2.f * xy + GSR_SH_C3_5 * S_(14, ch) * -2.f * yz +
                                   GSR_SH_C3_6 * S_(15, ch) * -3.f * 2.f * xy);
                        gz[ch] += (GSR_SH_C3_1 * S_(10, ch) * xy + GSR_SH_C3_2 * S_(11, ch) * 4.f * 2.f * yz +
                                   GSR_SH_C3_3 * S_(12, ch) * 3.f * (2.f * zz - xx - yy) + GSR_SH_C3_4 * S_(13, ch) * 4.f * 2.f * xz +
                                   GSR_SH_C3_5 * S_(14, ch) * (xx - yy));
                    }
                }
            }
        }
#undef S_
        // dL/dsh[k][ch] = basis_k * dL/dRGB[ch]; reuse f as the output row
#pragma unroll
        for (int k = 0; k < 16; k++) { f[3 * k] = w[k] * dRGB[0]; f[3 * k + 1] = w[k] * dRGB[1]; f[3 * k + 2] = w[k] * dRGB[2]; }
        store_sh_grads(a.dL_dsh + (size_t)M3f * idx, f, 3 * ncoef, M3f);

        const float ddx = gx[0] * dRGB[0] + gx[1] * dRGB[1] + gx[2] * dRGB[2];
        const float ddy = gy[0] * dRGB[0] + gy[1] * dRGB[1] + gy[2] * dRGB[2];
        const float ddz = gz[0] * dRGB[0] + gz[1] * dRGB[1] + gz[2] * dRGB[2];
        // dnormvdv (auxiliary.h:107-117)
        const float inv = 1.0f / sqrtf(s2 * s2 * s2);
        dmx += ((+s2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv;
        dmy += (-ox * oy * ddx + (s2 - oy * oy) * ddy - oz * oy * ddz) * inv;
        dmz += (-ox * oz * ddx - oy * oz * ddy + (s2 - oz * oz) * ddz) * inv;
    }
    a.dL_dmean3D[3 * idx] = dmx; a.dL_dmean3D[3 * idx + 1] = dmy; a.dL_dmean3D[3 * idx + 2] = dmz;

    // ---- cov3D -> scale / rotation (backward.cu:278-341; no quaternion-normalisation Jacobian)
    if (a.dL_dscale) {
        // dL/dSigma symmetric with halved off-diagonals; dL/dM = 2 M dL/dSigma
        M3 dSig = m3_cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                          0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
        M3 M2;
#pragma unroll
        for (int i = 0; i < 3; i++) { M2.c[i][0] = 2.0f * sc.x * R.c[i][0]; M2.c[i][1] = 2.0f * sc.y * R.c[i][1]; M2.c[i][2] = 2.0f * sc.z * R.c[i][2]; }
        M3 dM = m3_mul(M2, dSig);
        M3 Rt = m3_t(R), dMt = m3_t(dM);
        float dsv[3];
#pragma unroll
        for (int k = 0; k < 3; k++) dsv[k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
        // parity quirk: like the reference (backward.cu:322-325) this is the gradient w.r.t. mod*scale, NOT multiplied by mod
        a.dL_dscale[3 * idx] = dsv[0]; a.dL_dscale[3 * idx + 1] = dsv[1]; a.dL_dscale[3 * idx + 2] = dsv[2];
        const float sv[3] = {sc.x, sc.y, sc.z};
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int j = 0; j < 3; j++) dMt.c[k][j] *= sv[k];
        const float r = q.x, x = q.y, y = q.z, z = q.w;
#define D_(i, j) dMt.c[i][j]
        a.dL_drot[4 * idx + 0] = 2 * z * (D_(0, 1) - D_(1, 0)) + 2 * y * (D_(2, 0) - D_(0, 2)) + 2 * x * (D_(1, 2) - D_(2, 1));
        a.dL_drot[4 * idx + 1] = 2 * y * (D_(1, 0) + D_(0, 1)) + 2 * z * (D_(2, 0) + D_(0, 2)) + 2 * r * (D_(1, 2) - D_(2, 1)) - 4 * x * (D_(2, 2) + D_(1, 1));
        a.dL_drot[4 * idx + 2] = 2 * x * (D_(1, 0) + D_(0, 1)) + 2 * r * (D_(2, 0) - D_(0, 2)) + 2 * z * (D_(1, 2) + D_(2, 1)) - 4 * y * (D_(2, 2) + D_(0, 0));
        a.dL_drot[4 * idx + 3] = 2 * r * (D_(0, 1) - D_(1, 0)) + 2 * x * (D_(2, 0) + D_(0, 2)) + 2 * y * (D_(1, 2) + D_(2, 1)) - 4 * z * (D_(1, 1) + D_(0, 0));
#undef D_
    }
}

int launch_preprocess_backward(const GsrBwdArgs& a, const GeomView& g, const float* accum, cudaStream_t s)
{
    const CameraArgs cam = make_camera_args(a.viewmatrix, a.projmatrix, a.cam_pos, a.tan_fovx, a.tan_fovy, a.W, a.H);
    PreBwdArgs k;
    k.P = a.P; k.D = a.D; k.M = a.M;
    k.means = a.means3D; k.radii = a.radii; k.shs = a.shs; k.clamped = g.clamped;
    k.scales = a.scales; k.rots = a.rotations; k.cov_pre = a.cov3D_precomp; k.mod = a.scale_modifier;
    k.accum = accum; k.rec = g.rec;
    k.dL_dmean2D = a.dL_dmean2D; k.dL_dopacity = a.dL_dopacity; k.dL_dcolor = a.dL_dcolor;
    k.dL_dmean3D = a.dL_dmean3D; k.dL_dcov3D = a.dL_dcov3D; k.dL_dsh = a.shs ? a.dL_dsh : nullptr;
    k.dL_dscale = a.cov3D_precomp ? nullptr : a.dL_dscale;
    k.dL_drot = a.cov3D_precomp ? nullptr : a.dL_drot;
    GSR_LAUNCH(preprocess_bwd_kernel, cdiv(a.P, 256), 256, 0, s, k, cam);
    return GSR_OK;
}

}  // namespace gsr
