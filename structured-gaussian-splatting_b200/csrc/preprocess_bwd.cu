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
    unsigned acc;            // GSR_ACC_* bits: add to the existing contents (visible rows only) instead of overwriting
    int V;                   // views of a batched call (<= GSR_MAX_VIEWS); per-view arrays are [V][P]
    int sh_done;             // the SH part (dL_dsh, its share of dL_dmean3D) was done by sh_bwd_m16_kernel
};

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// out = v, or out += v when accumulating
__device__ __forceinline__ void put(float* p, float v, bool acc) { *p = acc ? *p + v : v; }
// dL/dmean3D gets a share from the SH kernel and one from preprocess_bwd_kernel; when the two run concurrently
// (launch_preprocess_backward) both add atomically onto a buffer that was zeroed (or already holds earlier views' sums)
constexpr unsigned ACC_ATOMIC_M3 = 0x80000000u;
__device__ __forceinline__ void put_m3(float* p, float v, bool acc, bool atomic)
{
    if (atomic) atomicAdd(p, v);
    else put(p, v, acc);
}

__device__ __forceinline__ void store_sh_grads(float* __restrict__ dst, const float* g, int nfloats, int total_floats, bool acc = false)
{
    if (acc) {      // read-modify-write of the live coefficients only
        for (int i = 0; i < nfloats; i++) dst[i] += g[i];
        return;
    }
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

template <bool WITH_SH, bool MULTI>
__global__ void __launch_bounds__(256, WITH_SH ? 1 : 4) preprocess_bwd_kernel(PreBwdArgs a, CameraArgs ca,
                                                                                            const __grid_constant__ ViewTab vt)
{
    const int NV = MULTI ? a.V : 1;            // compile-time 1 for the single-view call: the view loops fold away
    // a.V views (1 for the reference's single-view call): view v's per-view arrays (radii, records, clamp flags, accumulator
    // rows, dL/dmean2D, camera) are reached through the pointer table vt - slices v * P of one allocation for a batched call,
    // separate allocations for the deferred backward of single-view calls; the parameter gradients are summed over the
    // views in registers and written once.
    __shared__ Camera cams[MULTI ? GSR_MAX_VIEWS : 1];
    for (int v = 0; v < NV; v++) {
        CameraArgs cv = ca;
        if (MULTI) { cv.view = vt.view[v]; cv.proj = vt.proj[v]; cv.cam = vt.cam[v]; }
        load_camera(cams[v], cv);
    }
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    const int M3f = 3 * a.M;

    const bool acc_m = a.acc & GSR_ACC_MEAN3D, acc_o = a.acc & GSR_ACC_OPACITY, acc_sh = a.acc & GSR_ACC_SH,
               acc_s = a.acc & GSR_ACC_SCALE, acc_r = a.acc & GSR_ACC_ROT;
    bool any_visible = false;
    for (int v = 0; v < NV; v++) any_visible = any_visible || (MULTI ? vt.radii[v][idx] : a.radii[idx]) > 0;
    if (!any_visible) {
        // culled in every view: every owned gradient row is zero (left untouched when accumulating)
        for (int v = 0; v < NV; v++) {
            float* m2 = (MULTI ? vt.m2d[v] : a.dL_dmean2D) + 3 * (size_t)idx;
            m2[0] = 0.f; m2[1] = 0.f; m2[2] = 0.f;
        }
        if (!acc_o) a.dL_dopacity[idx] = 0.f;
        if (!acc_m && !a.sh_done) { a.dL_dmean3D[3 * idx] = 0.f; a.dL_dmean3D[3 * idx + 1] = 0.f; a.dL_dmean3D[3 * idx + 2] = 0.f; }
        if (a.dL_dcolor) { a.dL_dcolor[3 * idx] = 0.f; a.dL_dcolor[3 * idx + 1] = 0.f; a.dL_dcolor[3 * idx + 2] = 0.f; }
        if (a.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * (size_t)idx + k] = 0.f;
        }
        if (WITH_SH && a.dL_dsh && !acc_sh && !a.sh_done) {
            float z[1] = {0.f};
            store_sh_grads(a.dL_dsh + (size_t)M3f * idx, z, 0, M3f);
        }
        if (a.dL_dscale && !acc_s) { a.dL_dscale[3 * idx] = 0.f; a.dL_dscale[3 * idx + 1] = 0.f; a.dL_dscale[3 * idx + 2] = 0.f; }
        if (a.dL_drot && !acc_r) {
            a.dL_drot[4 * idx] = 0.f; a.dL_drot[4 * idx + 1] = 0.f; a.dL_drot[4 * idx + 2] = 0.f; a.dL_drot[4 * idx + 3] = 0.f;
        }
        return;
    }

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

    // ---- per view
    float s_dop = 0.f, s_dm[3] = {0.f, 0.f, 0.f}, s_dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool sh_written = false;
    (void)sh_written;
    for (int v = 0; v < NV; v++) {
    if (MULTI && v + 1 < NV) {
        // the next view's rows on their way into L2 while this view is computed (the loop is a chain of dependent loads)
        prefetch_l2(vt.radii[v + 1] + idx);
        prefetch_l2(vt.accum[v + 1] + (size_t)12 * idx);
        prefetch_l2(&vt.rec[v + 1][idx]);
    }
    const Camera& cam = cams[v];
    const int* v_radii = MULTI ? vt.radii[v] : a.radii;
    float* v_m2d = MULTI ? vt.m2d[v] : a.dL_dmean2D;
    const float* v_accum = MULTI ? vt.accum[v] : a.accum;
    const SplatRec* v_rec = MULTI ? vt.rec[v] : a.rec;
    const uint8_t* v_clamped = MULTI ? vt.clamped[v] : a.clamped;
    (void)v_clamped;
    const size_t vi = (size_t)idx;            // index inside view v's arrays
    if (!(v_radii[vi] > 0)) {
        v_m2d[3 * vi] = 0.f; v_m2d[3 * vi + 1] = 0.f; v_m2d[3 * vi + 2] = 0.f;
        continue;
    }
    // accumulator row: three 16-byte loads
    const float4* ar = reinterpret_cast<const float4*>(v_accum + (size_t)12 * vi);
    const float4 A0 = ar[0], A1 = ar[1], A2 = ar[2];
    const float dcol[3] = {A0.x, A0.y, A0.z};
    // moments -> gradients (backward.cu:536-551 with the per-Gaussian constants factored out of the pixel sums)
    const float4 r0 = v_rec[vi].q0, r1 = v_rec[vi].q1;
    const float cA = r0.z, cB = r0.w, cC = r1.x, opac = r1.y;
    const float g2x = -0.5f * cam.W * (cA * A0.w + cB * A1.x);
    const float g2y = -0.5f * cam.H * (cC * A1.x + cB * A0.w);
    const float dcon_x = -0.5f * A1.y, dcon_y = -0.5f * A1.z, dcon_w = -0.5f * A1.w;
    v_m2d[3 * vi] = g2x; v_m2d[3 * vi + 1] = g2y; v_m2d[3 * vi + 2] = 0.f;
    s_dop += opac > 0.f ? A2.x / opac : 0.f;
    if (a.dL_dcolor) { a.dL_dcolor[3 * idx] = dcol[0]; a.dL_dcolor[3 * idx + 1] = dcol[1]; a.dL_dcolor[3 * idx + 2] = dcol[2]; }

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
    if (WITH_SH && a.shs && !a.sh_done) {
        const float ox = p.x - cam.cam[0], oy = p.y - cam.cam[1], oz = p.z - cam.cam[2];
        const float s2 = ox * ox + oy * oy + oz * oz;
        const float len = sqrtf(s2);
        const float x = ox / len, y = oy / len, z = oz / len;
        const uint32_t cl = v_clamped[vi];
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
        store_sh_grads(a.dL_dsh + (size_t)M3f * idx, f, 3 * ncoef, M3f, acc_sh || sh_written);
        sh_written = true;

        const float ddx = gx[0] * dRGB[0] + gx[1] * dRGB[1] + gx[2] * dRGB[2];
        const float ddy = gy[0] * dRGB[0] + gy[1] * dRGB[1] + gy[2] * dRGB[2];
        const float ddz = gz[0] * dRGB[0] + gz[1] * dRGB[1] + gz[2] * dRGB[2];
        // dnormvdv (auxiliary.h:107-117)
        const float inv = 1.0f / sqrtf(s2 * s2 * s2);
        dmx += ((+s2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv;
        dmy += (-ox * oy * ddx + (s2 - oy * oy) * ddy - oz * oy * ddz) * inv;
        dmz += (-ox * oz * ddx - oy * oz * ddy + (s2 - oz * oz) * ddz) * inv;
    }
    s_dm[0] += dmx; s_dm[1] += dmy; s_dm[2] += dmz;
#pragma unroll
    for (int k = 0; k < 6; k++) s_dcov[k] += dcov[k];
    }   // views

    put(a.dL_dopacity + idx, s_dop, acc_o);
    {
        const bool add = acc_m || a.sh_done;       // sh_bwd_m16_kernel has already written its share
        const bool atom = a.acc & ACC_ATOMIC_M3;   // ... or is writing it right now
        put_m3(a.dL_dmean3D + 3 * idx, s_dm[0], add, atom); put_m3(a.dL_dmean3D + 3 * idx + 1, s_dm[1], add, atom);
        put_m3(a.dL_dmean3D + 3 * idx + 2, s_dm[2], add, atom);
    }

    // ---- cov3D -> scale / rotation (backward.cu:278-341; no quaternion-normalisation Jacobian)
    if (a.dL_dscale) {
        // dL/dSigma symmetric with halved off-diagonals; dL/dM = 2 M dL/dSigma
        const float* dcov = s_dcov;
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
        put(a.dL_dscale + 3 * idx, dsv[0], acc_s); put(a.dL_dscale + 3 * idx + 1, dsv[1], acc_s); put(a.dL_dscale + 3 * idx + 2, dsv[2], acc_s);
        const float sv[3] = {sc.x, sc.y, sc.z};
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int j = 0; j < 3; j++) dMt.c[k][j] *= sv[k];
        const float r = q.x, x = q.y, y = q.z, z = q.w;
#define D_(i, j) dMt.c[i][j]
        put(a.dL_drot + 4 * idx + 0, 2 * z * (D_(0, 1) - D_(1, 0)) + 2 * y * (D_(2, 0) - D_(0, 2)) + 2 * x * (D_(1, 2) - D_(2, 1)), acc_r);
        put(a.dL_drot + 4 * idx + 1, 2 * y * (D_(1, 0) + D_(0, 1)) + 2 * z * (D_(2, 0) + D_(0, 2)) + 2 * r * (D_(1, 2) - D_(2, 1)) - 4 * x * (D_(2, 2) + D_(1, 1)), acc_r);
        put(a.dL_drot + 4 * idx + 2, 2 * x * (D_(1, 0) + D_(0, 1)) + 2 * r * (D_(2, 0) - D_(0, 2)) + 2 * z * (D_(1, 2) + D_(2, 1)) - 4 * y * (D_(2, 2) + D_(0, 0)), acc_r);
        put(a.dL_drot + 4 * idx + 3, 2 * r * (D_(0, 1) - D_(1, 0)) + 2 * x * (D_(2, 0) + D_(0, 2)) + 2 * y * (D_(1, 2) + D_(2, 1)) - 4 * z * (D_(1, 1) + D_(0, 0)), acc_r);
#undef D_
    }
}

// ------------------------------------------------------------------ SH backward, 16 coefficients per Gaussian
// One lane per Gaussian for the arithmetic, but the 192-byte SH rows travel through shared memory: a warp loads the
// 32 rows it owns as 12 fully coalesced float4 loads per lane (6 KB contiguous), each lane then walks ITS row in
// four chunks of 4 coefficients (3 x LDS.128, rows padded to 13 float4 so the 8 lanes of a phase hit distinct
// banks), overwrites the chunk with dL/dsh = basis_k * dL/dRGB, and the rows go back out coalesced (added to
// the destination when accumulating).  The direction gradient sum_k dbasis_k/d(x,y,z) * (sh_k . dL/dRGB) stays in
// registers; dnormvdv (auxiliary.h:107-117) turns it into this part of dL/dmean3D.  Restates backward.cu:20-139.
constexpr int SHB_WARPS = 4;
constexpr int SHB_ROW4 = 13;            // padded row length in float4

// basis and derivative of coefficients 4*J .. 4*J+3 (constants auxiliary.h:22-39, signs forward.cu:30-59)
template <int J>
__device__ __forceinline__ void sh_chunk(int D, float x, float y, float z, float* w, float* bx, float* by, float* bz)
{
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
    for (int k = 0; k < 4; k++) { w[k] = 0.f; bx[k] = 0.f; by[k] = 0.f; bz[k] = 0.f; }
    if (J == 0) {
        w[0] = GSR_SH_C0;
        if (D > 0) {
            w[1] = -GSR_SH_C1 * y; by[1] = -GSR_SH_C1;
            w[2] = GSR_SH_C1 * z; bz[2] = GSR_SH_C1;
            w[3] = -GSR_SH_C1 * x; bx[3] = -GSR_SH_C1;
        }
    } else if (J == 1) {
        if (D > 1) {
            w[0] = GSR_SH_C2_0 * xy; bx[0] = GSR_SH_C2_0 * y; by[0] = GSR_SH_C2_0 * x;
            w[1] = GSR_SH_C2_1 * yz; by[1] = GSR_SH_C2_1 * z; bz[1] = GSR_SH_C2_1 * y;
            w[2] = GSR_SH_C2_2 * (2.f * zz - xx - yy); bx[2] = GSR_SH_C2_2 * -2.f * x; by[2] = GSR_SH_C2_2 * -2.f * y; bz[2] = GSR_SH_C2_2 * 4.f * z;
            w[3] = GSR_SH_C2_3 * xz; bx[3] = GSR_SH_C2_3 * z; bz[3] = GSR_SH_C2_3 * x;
        }
    } else if (J == 2) {
        if (D > 1) { w[0] = GSR_SH_C2_4 * (xx - yy); bx[0] = GSR_SH_C2_4 * 2.f * x; by[0] = GSR_SH_C2_4 * -2.f * y; }
        if (D > 2) {
            w[1] = GSR_SH_C3_0 * y * (3.f * xx - yy); bx[1] = GSR_SH_C3_0 * 6.f * xy; by[1] = GSR_SH_C3_0 * 3.f * (xx - yy);
            w[2] = GSR_SH_C3_1 * xy * z; bx[2] = GSR_SH_C3_1 * yz; by[2] = GSR_SH_C3_1 * xz; bz[2] = GSR_SH_C3_1 * xy;
            w[3] = GSR_SH_C3_2 * y * (4.f * zz - xx - yy); bx[3] = GSR_SH_C3_2 * -2.f * xy;
            by[3] = GSR_SH_C3_2 * (-3.f * yy + 4.f * zz - xx); bz[3] = GSR_SH_C3_2 * 8.f * yz;
        }
    } else {
        if (D > 2) {
            w[0] = GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy); bx[0] = GSR_SH_C3_3 * -6.f * xz; by[0] = GSR_SH_C3_3 * -6.f * yz;
            bz[0] = GSR_SH_C3_3 * 3.f * (2.f * zz - xx - yy);
            w[1] = GSR_SH_C3_4 * x * (4.f * zz - xx - yy); bx[1] = GSR_SH_C3_4 * (-3.f * xx + 4.f * zz - yy); by[1] = GSR_SH_C3_4 * -2.f * xy;
            bz[1] = GSR_SH_C3_4 * 8.f * xz;
            w[2] = GSR_SH_C3_5 * z * (xx - yy); bx[2] = GSR_SH_C3_5 * 2.f * xz; by[2] = GSR_SH_C3_5 * -2.f * yz; bz[2] = GSR_SH_C3_5 * (xx - yy);
            w[3] = GSR_SH_C3_6 * x * (xx - 3.f * yy); bx[3] = GSR_SH_C3_6 * 3.f * (xx - yy); by[3] = GSR_SH_C3_6 * -6.f * xy;
        }
    }
}

// a lane's SH row in shared memory: padded float4 rows (one [P,16,3] array) or the flat copies of the two split arrays
struct RowPadded {
    float4* row;
    template <int J> __device__ __forceinline__ void load(float* f) const
    {
        const float4 s0 = row[3 * J], s1 = row[3 * J + 1], s2 = row[3 * J + 2];
        f[0] = s0.x; f[1] = s0.y; f[2] = s0.z; f[3] = s0.w; f[4] = s1.x; f[5] = s1.y; f[6] = s1.z; f[7] = s1.w;
        f[8] = s2.x; f[9] = s2.y; f[10] = s2.z; f[11] = s2.w;
    }
    template <int J> __device__ __forceinline__ void store(const float* g) const
    {
        row[3 * J] = make_float4(g[0], g[1], g[2], g[3]);
        row[3 * J + 1] = make_float4(g[4], g[5], g[6], g[7]);
        row[3 * J + 2] = make_float4(g[8], g[9], g[10], g[11]);
    }
};
struct RowFlat {
    float* dc;      // 3 floats of this lane (stride 3 between lanes)
    float* rest;    // 45 floats of this lane (stride 45: odd, conflict-free scalar accesses)
    template <int J> __device__ __forceinline__ void load(float* f) const
    {
#pragma unroll
        for (int i = 0; i < 12; i++) f[i] = (12 * J + i < 3) ? dc[12 * J + i] : rest[12 * J + i - 3];
    }
    template <int J> __device__ __forceinline__ void store(const float* g) const
    {
#pragma unroll
        for (int i = 0; i < 12; i++) {
            if (12 * J + i < 3) dc[12 * J + i] = g[i];
            else rest[12 * J + i - 3] = g[i];
        }
    }
};

template <int J, class Row>
__device__ __forceinline__ void sh_chunk_apply(const Row& row, int D, float x, float y, float z, float dR, float dG, float dB, float& ddx,
                                               float& ddy, float& ddz)
{
    float w[4], bx[4], by[4], bz[4];
    sh_chunk<J>(D, x, y, z, w, bx, by, bz);
    float f[12];
    row.template load<J>(f);
    float g[12];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float sd = f[3 * k] * dR + f[3 * k + 1] * dG + f[3 * k + 2] * dB;
        ddx += bx[k] * sd; ddy += by[k] * sd; ddz += bz[k] * sd;
        g[3 * k] = w[k] * dR; g[3 * k + 1] = w[k] * dG; g[3 * k + 2] = w[k] * dB;
    }
    row.template store<J>(g);
}

// ---- staging for SPLIT SH storage (GsrBwdArgs.shs_rest): the warp's 32 rows are two contiguous runs, 32 x 3 floats of DC
// coefficients and 32 x 45 floats of the others (both 16-byte aligned at a multiple of 32 rows).  They are copied to shared
// memory AS THEY LIE (128-bit accesses, no index arithmetic); a lane then walks its own row with scalar accesses - the row
// strides 3 and 45 are odd, so those are bank-conflict free (RowFlat).  Lanes of culled Gaussians zero their row, so the way
// out is a plain copy (or a plain add in accumulating mode).
constexpr int SHB_FLAT_DC = 32 * 3;          // float offset of the f_rest copy inside the warp's buffer (96 + 1440 <= 32 * 52)

__device__ __forceinline__ void flat_in(float* sm, const float* __restrict__ src, int n, int lane)
{
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(sm);
    const int n4 = n >> 2;
    for (int e = lane; e < n4; e += 32) d4[e] = __ldg(s4 + e);
    for (int f = (n4 << 2) + lane; f < n; f += 32) sm[f] = __ldg(src + f);
}

__device__ __forceinline__ void flat_out(const float* sm, float* __restrict__ dst, int n, bool acc, int lane)
{
    const float4* s4 = reinterpret_cast<const float4*>(sm);
    float4* d4 = reinterpret_cast<float4*>(dst);
    const int n4 = n >> 2;
    if (acc) {
        for (int e0 = 0; e0 < n4; e0 += 128) {          // four loads in flight before the dependent stores
            float4 o[4];
#pragma unroll
            for (int u = 0; u < 4; u++) { const int e = e0 + u * 32 + lane; if (e < n4) o[u] = d4[e]; }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int e = e0 + u * 32 + lane;
                if (e < n4) { const float4 v = s4[e]; d4[e] = make_float4(o[u].x + v.x, o[u].y + v.y, o[u].z + v.z, o[u].w + v.w); }
            }
        }
        for (int f = (n4 << 2) + lane; f < n; f += 32) dst[f] += sm[f];
    } else {
        for (int e = lane; e < n4; e += 32) d4[e] = s4[e];
        for (int f = (n4 << 2) + lane; f < n; f += 32) dst[f] = sm[f];
    }
}

__device__ __forceinline__ void flat_zero_row(const RowFlat& r)
{
    r.dc[0] = 0.f; r.dc[1] = 0.f; r.dc[2] = 0.f;
#pragma unroll
    for (int i = 0; i < 45; i++) r.rest[i] = 0.f;
}

template <bool SPLIT>
__global__ void __launch_bounds__(32 * SHB_WARPS, 7) sh_bwd_m16_kernel(int P, int D, const float* __restrict__ means,
                                                                    const int* __restrict__ radii, const float* __restrict__ shs,
                                                                    const uint8_t* __restrict__ clamped, const float* __restrict__ accum,
                                                                    const float* __restrict__ campos, float* __restrict__ dL_dsh,
                                                                    float* __restrict__ dL_dmean3D, unsigned acc,
                                                                    const float* __restrict__ shs_rest, float* __restrict__ dL_dsh_rest)
{
    __shared__ float4 rows[SHB_WARPS][32 * SHB_ROW4];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int g0 = (blockIdx.x * SHB_WARPS + w) * 32;
    if (g0 >= P) return;
    const int nrow = min(32, P - g0);
    const bool acc_sh = acc & GSR_ACC_SH, acc_m = acc & GSR_ACC_MEAN3D;
    float4* my = rows[w];
    const int gid = g0 + lane;
    const bool visible = lane < nrow && radii[gid] > 0;
    const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
    // stage in (rows of culled Gaussians are not needed)
    float* const flat = reinterpret_cast<float*>(my);
    const RowFlat frow{flat + lane * 3, flat + SHB_FLAT_DC + lane * 45};
    if (SPLIT) {
        flat_in(flat, shs + (size_t)g0 * 3, nrow * 3, lane);
        flat_in(flat + SHB_FLAT_DC, shs_rest + (size_t)g0 * 45, nrow * 45, lane);
    } else {
        const float4* src = reinterpret_cast<const float4*>(shs) + (size_t)g0 * 12;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if ((vis_mask >> r) & 1u) my[r * SHB_ROW4 + c] = __ldg(src + e);
        }
    }
    __syncwarp();
    if (visible) {
        const float4 A0 = *reinterpret_cast<const float4*>(accum + (size_t)12 * gid);
        const uint32_t cl = clamped[gid];
        const float dR = (cl & 1) ? 0.f : A0.x, dG = (cl & 2) ? 0.f : A0.y, dB = (cl & 4) ? 0.f : A0.z;
        const float ox = means[3 * (size_t)gid] - campos[0], oy = means[3 * (size_t)gid + 1] - campos[1],
                    oz = means[3 * (size_t)gid + 2] - campos[2];
        const float sq = ox * ox + oy * oy + oz * oz;
        const float rl = rsqrtf(sq);
        const float x = ox * rl, y = oy * rl, z = oz * rl;
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;
        if (SPLIT) {
            sh_chunk_apply<0>(frow, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<1>(frow, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<2>(frow, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<3>(frow, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
        } else {
            const RowPadded row{my + lane * SHB_ROW4};
            sh_chunk_apply<0>(row, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<1>(row, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<2>(row, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
            sh_chunk_apply<3>(row, D, x, y, z, dR, dG, dB, ddx, ddy, ddz);
        }
        const float inv = rl * rl * rl;          // 1 / |o|^3
        const bool atom = acc & ACC_ATOMIC_M3;
        put_m3(dL_dmean3D + 3 * (size_t)gid, ((sq - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv, acc_m, atom);
        put_m3(dL_dmean3D + 3 * (size_t)gid + 1, (-ox * oy * ddx + (sq - oy * oy) * ddy - oz * oy * ddz) * inv, acc_m, atom);
        put_m3(dL_dmean3D + 3 * (size_t)gid + 2, (-ox * oz * ddx - oy * oz * ddy + (sq - oz * oz) * ddz) * inv, acc_m, atom);
    } else if (lane < nrow) {
        if (!acc_m && !(acc & ACC_ATOMIC_M3)) { dL_dmean3D[3 * (size_t)gid] = 0.f; dL_dmean3D[3 * (size_t)gid + 1] = 0.f; dL_dmean3D[3 * (size_t)gid + 2] = 0.f; }
        if (SPLIT) flat_zero_row(frow);
    }
    __syncwarp();
    // stage out
    if (SPLIT) {
        flat_out(flat, dL_dsh + (size_t)g0 * 3, nrow * 3, acc_sh, lane);
        flat_out(flat + SHB_FLAT_DC, dL_dsh_rest + (size_t)g0 * 45, nrow * 45, acc_sh, lane);
        return;
    }
    float4* dst = reinterpret_cast<float4*>(dL_dsh) + (size_t)g0 * 12;
    if (acc_sh) {
        // all twelve reads of the destination are issued before the first dependent store
        float4 o[12];
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12;
            o[i] = (r < nrow && ((vis_mask >> r) & 1u)) ? dst[e] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if (r < nrow && ((vis_mask >> r) & 1u)) {
                const float4 v = my[r * SHB_ROW4 + c];
                dst[e] = make_float4(o[i].x + v.x, o[i].y + v.y, o[i].z + v.z, o[i].w + v.w);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if (r < nrow) dst[e] = ((vis_mask >> r) & 1u) ? my[r * SHB_ROW4 + c] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}

// ------------------------------------------------------------------ SH backward over the V views of a batched call
// Same staging as sh_bwd_m16_kernel, but a row is read once and written once for ALL views: per chunk of 4 coefficients
// the lane loops over the views it is visible in, adds basis_k(dir_v) * dL/dRGB_v into the chunk's output and applies
// the direction gradient of that chunk to dL/dmean3D straight away (dnormvdv is linear in it), so no per-view state
// survives the inner loop.  Replaces V read-modify-write passes over the 192-byte gradient rows.
template <int J, class Row>
__device__ __forceinline__ void sh_chunk_views(const Row& row, int V, int D, int gid, float px, float py, float pz, const ViewTab& vt,
                                               float* dm)
{
    float f[12];
    row.template load<J>(f);
    float g[12];
#pragma unroll
    for (int k = 0; k < 12; k++) g[k] = 0.f;
    for (int v = 0; v < V; v++) {
        if (!(vt.radii[v][gid] > 0)) continue;
        const float4 A0 = *reinterpret_cast<const float4*>(vt.accum[v] + (size_t)12 * gid);
        const uint32_t cl = vt.clamped[v][gid];
        const float dR = (cl & 1) ? 0.f : A0.x, dG = (cl & 2) ? 0.f : A0.y, dB = (cl & 4) ? 0.f : A0.z;
        const float* campos = vt.cam[v];
        const float ox = px - campos[0], oy = py - campos[1], oz = pz - campos[2];
        const float sq = ox * ox + oy * oy + oz * oz;
        const float rl = rsqrtf(sq);
        float w[4], bx[4], by[4], bz[4];
        sh_chunk<J>(D, ox * rl, oy * rl, oz * rl, w, bx, by, bz);
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float sd = f[3 * k] * dR + f[3 * k + 1] * dG + f[3 * k + 2] * dB;
            ddx += bx[k] * sd; ddy += by[k] * sd; ddz += bz[k] * sd;
            g[3 * k] += w[k] * dR; g[3 * k + 1] += w[k] * dG; g[3 * k + 2] += w[k] * dB;
        }
        const float inv = rl * rl * rl;
        dm[0] += ((sq - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv;
        dm[1] += (-ox * oy * ddx + (sq - oy * oy) * ddy - oz * oy * ddz) * inv;
        dm[2] += (-ox * oz * ddx - oy * oz * ddy + (sq - oz * oz) * ddz) * inv;
    }
    row.template store<J>(g);
}

template <bool SPLIT>
__global__ void __launch_bounds__(32 * SHB_WARPS, 7) sh_bwd_m16_views_kernel(int P, int V, int D, const float* __restrict__ means,
                                                                          const float* __restrict__ shs, float* __restrict__ dL_dsh,
                                                                          float* __restrict__ dL_dmean3D, unsigned acc,
                                                                          const float* __restrict__ shs_rest,
                                                                          float* __restrict__ dL_dsh_rest,
                                                                          const __grid_constant__ ViewTab vt)
{
    __shared__ float4 rows[SHB_WARPS][32 * SHB_ROW4];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int g0 = (blockIdx.x * SHB_WARPS + w) * 32;
    if (g0 >= P) return;
    const int nrow = min(32, P - g0);
    const bool acc_sh = acc & GSR_ACC_SH, acc_m = acc & GSR_ACC_MEAN3D;
    float4* my = rows[w];
    const int gid = g0 + lane;
    bool visible = false;
    if (lane < nrow)
        for (int v = 0; v < V; v++) visible = visible || vt.radii[v][gid] > 0;
    const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
    if (visible) {
        // every view's accumulator row and clamp byte on their way into L2 while the SH rows are staged: the chunk loops
        // below read them view after view, each read a dependent miss otherwise
        for (int v = 0; v < V; v++) {
            prefetch_l2(vt.accum[v] + (size_t)12 * gid);
            if (lane == __ffs(vis_mask) - 1) prefetch_l2(vt.clamped[v] + gid);
        }
    }
    float* const flat = reinterpret_cast<float*>(my);
    const RowFlat frow{flat + lane * 3, flat + SHB_FLAT_DC + lane * 45};
    if (SPLIT) {
        flat_in(flat, shs + (size_t)g0 * 3, nrow * 3, lane);
        flat_in(flat + SHB_FLAT_DC, shs_rest + (size_t)g0 * 45, nrow * 45, lane);
    } else {
        const float4* src = reinterpret_cast<const float4*>(shs) + (size_t)g0 * 12;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if ((vis_mask >> r) & 1u) my[r * SHB_ROW4 + c] = __ldg(src + e);
        }
    }
    __syncwarp();
    if (visible) {
        const float px = means[3 * (size_t)gid], py = means[3 * (size_t)gid + 1], pz = means[3 * (size_t)gid + 2];
        float dm[3] = {0.f, 0.f, 0.f};
        if (SPLIT) {
            sh_chunk_views<0>(frow, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<1>(frow, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<2>(frow, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<3>(frow, V, D, gid, px, py, pz, vt, dm);
        } else {
            const RowPadded row{my + lane * SHB_ROW4};
            sh_chunk_views<0>(row, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<1>(row, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<2>(row, V, D, gid, px, py, pz, vt, dm);
            sh_chunk_views<3>(row, V, D, gid, px, py, pz, vt, dm);
        }
        const bool atom = acc & ACC_ATOMIC_M3;
        put_m3(dL_dmean3D + 3 * (size_t)gid, dm[0], acc_m, atom);
        put_m3(dL_dmean3D + 3 * (size_t)gid + 1, dm[1], acc_m, atom);
        put_m3(dL_dmean3D + 3 * (size_t)gid + 2, dm[2], acc_m, atom);
    } else if (lane < nrow) {
        if (!acc_m && !(acc & ACC_ATOMIC_M3)) { dL_dmean3D[3 * (size_t)gid] = 0.f; dL_dmean3D[3 * (size_t)gid + 1] = 0.f; dL_dmean3D[3 * (size_t)gid + 2] = 0.f; }
        if (SPLIT) flat_zero_row(frow);
    }
    __syncwarp();
    if (SPLIT) {
        flat_out(flat, dL_dsh + (size_t)g0 * 3, nrow * 3, acc_sh, lane);
        flat_out(flat + SHB_FLAT_DC, dL_dsh_rest + (size_t)g0 * 45, nrow * 45, acc_sh, lane);
        return;
    }
    float4* dst = reinterpret_cast<float4*>(dL_dsh) + (size_t)g0 * 12;
    if (acc_sh) {
        float4 o[12];
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12;
            o[i] = (r < nrow && ((vis_mask >> r) & 1u)) ? dst[e] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if (r < nrow && ((vis_mask >> r) & 1u)) {
                const float4 v = my[r * SHB_ROW4 + c];
                dst[e] = make_float4(o[i].x + v.x, o[i].y + v.y, o[i].z + v.z, o[i].w + v.w);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane, r = e / 12, c = e - r * 12;
            if (r < nrow) dst[e] = ((vis_mask >> r) & 1u) ? my[r * SHB_ROW4 + c] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}

int launch_preprocess_backward(const GsrBwdArgs& a, int V, const GeomView& g, const float* accum, cudaStream_t s, const ViewTab* tab)
{
    // a: the batched call's arguments (per-view arrays [V][...], cameras [V][16] / [V][3]); V <= GSR_MAX_VIEWS.
    // tab != nullptr: the views' arrays are reached through this pointer table instead (deferred backward of single-view
    // calls: every view has blobs of its own); a.radii / a.dL_dmean2D / g / accum are then ignored for V > 1.
    ViewTab vt = {};
    if (tab) vt = *tab;
    else if (V > 1) {
        for (int v = 0; v < V; v++) {
            const size_t o = (size_t)v * a.P;
            vt.radii[v] = a.radii + o; vt.rec[v] = g.rec + o; vt.clamped[v] = g.clamped + o; vt.accum[v] = accum + 12 * o;
            vt.m2d[v] = a.dL_dmean2D + 3 * o;
            vt.view[v] = a.viewmatrix + 16 * v; vt.proj[v] = a.projmatrix ? a.projmatrix + 16 * v : nullptr;
            vt.cam[v] = a.cam_pos ? a.cam_pos + 3 * v : nullptr;
        }
    }
    const CameraArgs cam = make_camera_args(a.viewmatrix, a.projmatrix, a.cam_pos, a.tan_fovx, a.tan_fovy, a.W, a.H);
    PreBwdArgs k;
    k.P = a.P; k.D = a.D; k.M = a.M; k.V = V;
    k.means = a.means3D; k.radii = a.radii; k.shs = a.shs; k.clamped = g.clamped;
    k.scales = a.scales; k.rots = a.rotations; k.cov_pre = a.cov3D_precomp; k.mod = a.scale_modifier;
    k.accum = accum; k.rec = g.rec;
    k.dL_dmean2D = a.dL_dmean2D; k.dL_dopacity = a.dL_dopacity; k.dL_dcolor = a.dL_dcolor;
    k.dL_dmean3D = a.dL_dmean3D; k.dL_dcov3D = a.dL_dcov3D; k.dL_dsh = a.shs ? a.dL_dsh : nullptr;
    k.dL_dscale = a.cov3D_precomp ? nullptr : a.dL_dscale;
    k.dL_drot = a.cov3D_precomp ? nullptr : a.dL_drot;
    k.acc = a.accumulate;
    k.sh_done = 0;
    const bool split = a.shs && a.shs_rest && a.M > 1;      // DC and the other coefficients in two arrays (M == 16, checked by the API)
    const bool sh_fast = split || (a.shs && a.M == 16 && ((((size_t)a.shs) | ((size_t)a.dL_dsh)) & 15) == 0 && options().sh_split);
    // The SH kernel and preprocess_bwd_kernel are both HBM-bound and share one output only (dL/dmean3D): with that output
    // turned into atomic adds they run side by side on two streams (the pair reaches a higher fraction of the HBM peak than
    // either alone) and are joined before this function returns.
    // (the side stream and its events belong to the device context, one set per caller stream)
    const bool forked = sh_fast && options().bwd_side != 0;
    cudaStream_t main_s = s;
    unsigned acc_flags = a.accumulate;
    SideStream ss;
    if (forked) {
        DeviceCtx* ctx = nullptr;
        int rc0;
        if ((rc0 = device_ctx(&ctx)) || (rc0 = ctx_side_stream(ctx, 1, main_s, &ss))) return rc0;
    }
    if (forked) {
        if (!(a.accumulate & GSR_ACC_MEAN3D)) GSR_CUDA(cudaMemsetAsync(a.dL_dmean3D, 0, sizeof(float) * 3 * (size_t)a.P, main_s));
        acc_flags |= ACC_ATOMIC_M3;
        k.acc = acc_flags;
        GSR_CUDA(cudaEventRecord(ss.fork, main_s));
        GSR_CUDA(cudaStreamWaitEvent(ss.side, ss.fork, 0));
        s = ss.side;
    }
    if (split) {
        if (((((size_t)a.shs) | ((size_t)a.dL_dsh) | ((size_t)a.shs_rest) | ((size_t)a.dL_dsh_rest)) & 15) != 0) return GSR_EINVAL;
        if (V == 1)
            GSR_LAUNCH(sh_bwd_m16_kernel<true>, cdiv(a.P, 32 * SHB_WARPS), 32 * SHB_WARPS, 0, s, a.P, a.D, a.means3D, a.radii, a.shs,
                       g.clamped, accum, a.cam_pos, a.dL_dsh, a.dL_dmean3D, acc_flags, a.shs_rest, a.dL_dsh_rest);
        else
            GSR_LAUNCH(sh_bwd_m16_views_kernel<true>, cdiv(a.P, 32 * SHB_WARPS), 32 * SHB_WARPS, 0, s, a.P, V, a.D, a.means3D,
                       a.shs, a.dL_dsh, a.dL_dmean3D, acc_flags, a.shs_rest, a.dL_dsh_rest, vt);
        k.sh_done = 1;
    } else if (a.shs && a.M == 16 && ((((size_t)a.shs) | ((size_t)a.dL_dsh)) & 15) == 0 && options().sh_split) {
        if (V == 1)
            GSR_LAUNCH(sh_bwd_m16_kernel<false>, cdiv(a.P, 32 * SHB_WARPS), 32 * SHB_WARPS, 0, s, a.P, a.D, a.means3D, a.radii, a.shs,
                       g.clamped, accum, a.cam_pos, a.dL_dsh, a.dL_dmean3D, acc_flags, (const float*)nullptr, (float*)nullptr);
        else
            GSR_LAUNCH(sh_bwd_m16_views_kernel<false>, cdiv(a.P, 32 * SHB_WARPS), 32 * SHB_WARPS, 0, s, a.P, V, a.D, a.means3D,
                       a.shs, a.dL_dsh, a.dL_dmean3D, acc_flags, (const float*)nullptr, (float*)nullptr, vt);
        k.sh_done = 1;
    }
    if (forked) {
        GSR_CUDA(cudaEventRecord(ss.join, ss.side));
        s = main_s;
    }
    const bool with_sh = !(k.sh_done || !a.shs);
    if (V == 1) {
        if (!with_sh) GSR_LAUNCH((preprocess_bwd_kernel<false, false>), cdiv(a.P, 256), 256, 0, s, k, cam, vt);
        else GSR_LAUNCH((preprocess_bwd_kernel<true, false>), cdiv(a.P, 256), 256, 0, s, k, cam, vt);
    } else {
        if (!with_sh) GSR_LAUNCH((preprocess_bwd_kernel<false, true>), cdiv(a.P, 256), 256, 0, s, k, cam, vt);
        else GSR_LAUNCH((preprocess_bwd_kernel<true, true>), cdiv(a.P, 256), 256, 0, s, k, cam, vt);
    }
    if (forked) GSR_CUDA(cudaStreamWaitEvent(main_s, ss.join, 0));
    return GSR_OK;
}

}  // namespace gsr
