// preprocess.cu - per-Gaussian forward preprocess, frustum marking.
//
// Replaces (reference paths relative to submodules/diff-gaussian-rasterization/cuda_rasterizer):
//   preprocess_fwd_kernel  <- preprocessCUDA<3>            forward.cu:155-256 (+ :20-71, :74-113, :118-152)
//   mark_visible_kernel    <- checkFrustum                 rasterizer_impl.cu:54-66
//
// One thread per Gaussian.  Output is ONE packed 48-byte record per Gaussian (SplatRec) instead of
// the reference's seven separate arrays (79 B/Gaussian): the blending kernels gather exactly this
// record per tile-list entry, and the binning stages read radius / tile count / depth out of it.
// cov3D is not stored: the backward recomputes it from scale/rotation (24 B/Gaussian saved twice).
#include "gsr_common.cuh"
#include "math3.cuh"

namespace gsr {

// double arithmetic on purpose: the reference's literals are doubles (auxiliary.h:41-44)
__device__ __forceinline__ float ndc_to_pix_d(float v, int S)
{
    return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5);
}

__device__ __forceinline__ void tile_rect(float px, float py, int rad, int gx, int gy, int& x0, int& y0, int& x1, int& y1)
{
    // auxiliary.h:46-56: float divide by the tile size, C truncation, clamp to [0, grid]
    x0 = min(gx, max(0, (int)((px - rad) / GSR_TILE)));
    y0 = min(gy, max(0, (int)((py - rad) / GSR_TILE)));
    x1 = min(gx, max(0, (int)((px + rad + GSR_TILE - 1) / GSR_TILE)));
    y1 = min(gy, max(0, (int)((py + rad + GSR_TILE - 1) / GSR_TILE)));
}

// scale / rotation -> world covariance, upper triangle (forward.cu:118-152; quaternion used un-normalised)
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 s, float mod, const float4 q, float* cov6)
{
    M3 S = m3_cols(mod * s.x, 0.0f, 0.0f, 0.0f, mod * s.y, 0.0f, 0.0f, 0.0f, mod * s.z);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    M3 M = m3_mul(S, R);
    M3 Sig = m3_mul(m3_t(M), M);
    cov6[0] = Sig.c[0][0]; cov6[1] = Sig.c[0][1]; cov6[2] = Sig.c[0][2];
    cov6[3] = Sig.c[1][1]; cov6[4] = Sig.c[1][2]; cov6[5] = Sig.c[2][2];
}

// EWA projection of the 3D covariance (forward.cu:74-113); returns (a, b, c) with the 0.3 low-pass added
__device__ __forceinline__ float3 cov2d_ewa(const float3 mean, const Camera& cam, const float* cov6)
{
    float3 t = xform4x3(mean, cam.view);
    const float limx = 1.3f * cam.tanx;
    const float limy = 1.3f * cam.tany;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;

    M3 J = m3_cols(cam.fx / t.z, 0.0f, -(cam.fx * t.x) / (t.z * t.z),
                   0.0f, cam.fy / t.z, -(cam.fy * t.y) / (t.z * t.z),
                   0.0f, 0.0f, 0.0f);
    M3 Wm = m3_cols(cam.view[0], cam.view[4], cam.view[8],
                    cam.view[1], cam.view[5], cam.view[9],
                    cam.view[2], cam.view[6], cam.view[10]);
    M3 T = m3_mul(Wm, J);
    M3 Vrk = m3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
    M3 cov = m3_mul(m3_mul(m3_t(T), m3_t(Vrk)), T);
    float3 r = {cov.c[0][0] + 0.3f, cov.c[0][1], cov.c[1][1] + 0.3f};
    return r;
}

// SH -> RGB for one Gaussian (forward.cu:20-71).  f = the Gaussian's coefficients, [k][channel] flattened.
template <int D>
__device__ __forceinline__ void sh_to_rgb(const float* f, float x, float y, float z, float* rgb, uint32_t& clamp_bits)
{
    clamp_bits = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ch++) {
#define S_(k) f[3 * (k) + ch]
        float res = GSR_SH_C0 * S_(0);
        if (D > 0) {
            res = res - GSR_SH_C1 * y * S_(1) + GSR_SH_C1 * z * S_(2) - GSR_SH_C1 * x * S_(3);
            if (D > 1) {
                float xx = x * x, yy = y * y, zz = z * z;
                float xy = x * y, yz = y * z, xz = x * z;
                res = res + GSR_SH_C2_0 * xy * S_(4) + GSR_SH_C2_1 * yz * S_(5) +
                      GSR_SH_C2_2 * (2.0f * zz - xx - yy) * S_(6) + GSR_SH_C2_3 * xz * S_(7) +
                      GSR_SH_C2_4 * (xx - yy) * S_(8);
                if (D > 2) {
                    res = res + GSR_SH_C3_0 * y * (3.0f * xx - yy) * S_(9) + GSR_SH_C3_1 * xy * z * S_(10) +
                          GSR_SH_C3_2 * y * (4.0f * zz - xx - yy) * S_(11) +
                          GSR_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S_(12) +
                          GSR_SH_C3_4 * x * (4.0f * zz - xx - yy) * S_(13) + GSR_SH_C3_5 * z * (xx - yy) * S_(14) +
                          GSR_SH_C3_6 * x * (xx - 3.0f * yy) * S_(15);
                }
            }
        }
#undef S_
        res += 0.5f;
        if (res < 0) clamp_bits |= 1u << ch;
        rgb[ch] = fmaxf(res, 0.0f);
    }
}

// loads the first NF floats of a Gaussian's SH block; 128-bit loads when the block is 16-byte aligned
template <int NF>
__device__ __forceinline__ void load_sh(const float* __restrict__ p, bool vec, float* f)
{
    if (vec) {
        const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
        for (int i = 0; i < NF / 4; i++) {
            float4 v = __ldg(p4 + i);
            f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w;
        }
#pragma unroll
        for (int i = NF / 4 * 4; i < NF; i++) f[i] = __ldg(p + i);
    } else {
#pragma unroll
        for (int i = 0; i < NF; i++) f[i] = __ldg(p + i);
    }
}

// split storage: DC coefficient at dc, coefficients 1.. in rows of 45 floats at rest_base (row idx starts 4-byte aligned, at
// float offset 45 idx = idx mod 4 past a 16-byte boundary).  The row is fetched with the 128-bit loads that cover it
// (12 of them for 45 floats, as many as the unsplit 48-float row needs) and shifted into place by a two-stage select
// network on idx mod 4 - 45 scalar loads would ask the L1 for four times as many sectors.
template <int NREST>
__device__ __forceinline__ void load_sh_split(const float* __restrict__ dc, const float* __restrict__ rest_base, size_t idx,
                                              size_t P, bool vec, float* f)
{
    f[0] = __ldg(dc); f[1] = __ldg(dc + 1); f[2] = __ldg(dc + 2);
    if (NREST == 0) return;
    constexpr int NV = (NREST + 3 + 3) / 4;            // float4s covering [row - 3, row + NREST)
    const float* row = rest_base + 45 * idx;
    const unsigned o = (unsigned)(idx & 3);            // floats between the 16-byte boundary below the row and the row
    if (vec && 45 * idx - o + 4 * NV <= 45 * P) {
        const float4* p4 = reinterpret_cast<const float4*>(row - o);
        float t[4 * NV + 4];
#pragma unroll
        for (int i = 0; i < NV; i++) {
            const float4 v = __ldg(p4 + i);
            t[4 * i] = v.x; t[4 * i + 1] = v.y; t[4 * i + 2] = v.z; t[4 * i + 3] = v.w;
        }
#pragma unroll
        for (int i = 4 * NV; i < 4 * NV + 4; i++) t[i] = 0.f;
        const bool s1 = o & 1u, s2 = o & 2u;
        float u[NREST + 2];
#pragma unroll
        for (int i = 0; i < NREST + 2; i++) u[i] = s1 ? t[i + 1] : t[i];
#pragma unroll
        for (int i = 0; i < NREST; i++) f[3 + i] = s2 ? u[i + 2] : u[i];
    } else {
#pragma unroll
        for (int i = 0; i < NREST; i++) f[3 + i] = __ldg(row + i);
    }
}

struct PreArgs {
    int P, D, M;
    const float* means;
    const float* shs;
    const float* shs_rest;
    const float* colors_pre;
    const float* opac;
    const float* scales;
    const float* rots;
    const float* cov_pre;
    float mod;
    int prefiltered;
    int* radii;
    SplatRec* rec;
    uint8_t* clamped;
    uint32_t* depthkey; // [P] sort key of the depth pass
    ushort4* rect;      // [P] tile rect
    uint32_t* total;    // sum of tiles touched (R)
    int* err;          // device flag: prefiltered violated
    int view_base;     // first view of this launch inside the batched call (scene s of a multi-scene call: s * views_per_scene)
    uint32_t key_bias; // depth key of a Gaussian that emits instances = bits(depth) - key_bias (gsr_common.cuh: GSR_DEPTH_KEY_BIAS or 0)
    uint32_t key_cull; // depth key of a culled Gaussian
};

constexpr int PRE_THREADS = 128;
constexpr int PRE_ROW4 = 13;        // float4 per staged SH row: 12 + 1 padding (odd: conflict-free LDS.128 by row)

// where one view's results go (a slice of a batched call's arrays, or a view's own blob)
struct PreOut {
    int* radii;
    SplatRec* rec;
    uint8_t* clamped;
    uint32_t* depthkey;
    ushort4* rect;
    uint32_t* total;    // sum of tiles touched (R)
    int* err;           // device flag: prefiltered violated
};

// One Gaussian under one camera (forward.cu:155-259): cull, covariance, conic, tile rect, colour; writes row `oidx` of the
// view's outputs.  `band`: the view's band of tile rows in a batched call's stacked image.  The whole warp calls (the
// staged SH rows were copied by all of its lanes).
// what does not depend on the camera, loaded / computed once when a thread walks several views (PRE)
struct GaussIn {
    float3 p;
    float c6[6];
    float opac;
};

template <bool STAGE, bool PRE>
__device__ __forceinline__ void preprocess_gaussian(const PreArgs& a, const Camera& cam, const PreOut& o, size_t oidx, int band, int idx,
                                                    int lane, const float4* st, bool inb, const GaussIn& gi)
{
    SplatRec out;
    out.q0 = make_float4(0.f, 0.f, 0.f, 0.f);
    out.q1 = make_float4(0.f, 0.f, 0.f, 0.f);
    out.q2 = make_float4(0.f, __uint_as_float(0xFFFFFFFFu), __int_as_float(0), __uint_as_float(0u));
    int radius_out = 0;
    uint32_t clamp_bits = 0;
    ushort4 rect_out = make_ushort4(0, 0, 0, 0);
    bool vis = false;                           // survives every cull: gets a colour
    float3 p = {0.f, 0.f, 0.f};

    if (inb) {
    p = PRE ? gi.p : make_float3(a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]);
    const float3 p_view = xform4x3(p, cam.view);
    if (!(p_view.z <= 0.2f)) {   // near cull only (auxiliary.h:154); NaN passes like the reference
        const float4 p_hom = xform4x4(p, cam.proj);
        const float p_w = 1.0f / (p_hom.w + 0.0000001f);
        const float projx = p_hom.x * p_w, projy = p_hom.y * p_w;

        float c6[6];
        if (PRE) {
#pragma unroll
            for (int k = 0; k < 6; k++) c6[k] = gi.c6[k];
        } else if (a.cov_pre != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c6[k] = a.cov_pre[6 * (size_t)idx + k];
        } else {
            const float3 s = {a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]};
            const float4 q = {a.rots[4 * idx], a.rots[4 * idx + 1], a.rots[4 * idx + 2], a.rots[4 * idx + 3]};
            cov3d_from_scale_rot(s, a.mod, q, c6);
        }
        const float3 cov = cov2d_ewa(p, cam, c6);
        const float det = (cov.x * cov.z - cov.y * cov.y);
        if (det != 0.0f) {
            const float det_inv = 1.f / det;
            const float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
            const float mid = 0.5f * (cov.x + cov.z);
            const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
            const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
            const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
            const float px = ndc_to_pix_d(projx, cam.W), py = ndc_to_pix_d(projy, cam.H);
            int x0, y0, x1, y1;
            tile_rect(px, py, (int)my_radius, cam.gx, cam.gy, x0, y0, x1, y1);
            const uint32_t area = (uint32_t)(x1 - x0) * (uint32_t)(y1 - y0);
            if (area != 0) {
                vis = true;
                // (the colour is filled in below)
                out.q0 = make_float4(px, py, conic.x, conic.y);
                out.q1 = make_float4(conic.z, PRE ? gi.opac : a.opac[idx], 0.f, 0.f);
                out.q2 = make_float4(0.f, p_view.z, __int_as_float((int)my_radius), __uint_as_float(area));
                radius_out = (int)my_radius;
                const int yoff = band * cam.gy;
                rect_out = make_ushort4((unsigned short)x0, (unsigned short)(y0 + yoff), (unsigned short)x1, (unsigned short)(y1 + yoff));
            }
        }
    } else if (a.prefiltered) {
        *o.err = 1;   // reference: printf + __trap (auxiliary.h:156-160); we flag it and report GSR_EINVAL in debug mode
    }
    }

    if (STAGE) {
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncwarp();                           // the rows were copied by all lanes of the warp
    }
    if (vis) {
        float rgb[3];
        if (a.colors_pre != nullptr) {
            rgb[0] = a.colors_pre[3 * (size_t)idx];
            rgb[1] = a.colors_pre[3 * (size_t)idx + 1];
            rgb[2] = a.colors_pre[3 * (size_t)idx + 2];
        } else {
            // view direction (forward.cu:26-28): dir / length(dir), IEEE divide per component
            float dx = p.x - cam.cam[0], dy = p.y - cam.cam[1], dz = p.z - cam.cam[2];
            const float len = sqrtf(dx * dx + dy * dy + dz * dz);
            dx = dx / len; dy = dy / len; dz = dz / len;
            float f[48];
            if (STAGE) {
                const float4* row = st + lane * PRE_ROW4;
#pragma unroll
                for (int i = 0; i < 12; i++) {
                    const float4 v = row[i];
                    f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w;
                }
                sh_to_rgb<3>(f, dx, dy, dz, rgb, clamp_bits);
            } else if (a.shs_rest != nullptr) {
                // M == 16 here (the API rejects other split layouts): rows of 45 floats
                const float* dc = a.shs + (size_t)3 * idx;
                const bool vec = (((size_t)a.shs_rest) & 15) == 0;
                switch (a.D) {
                case 0: load_sh_split<0>(dc, a.shs_rest, idx, a.P, vec, f); sh_to_rgb<0>(f, dx, dy, dz, rgb, clamp_bits); break;
                case 1: load_sh_split<9>(dc, a.shs_rest, idx, a.P, vec, f); sh_to_rgb<1>(f, dx, dy, dz, rgb, clamp_bits); break;
                case 2: load_sh_split<24>(dc, a.shs_rest, idx, a.P, vec, f); sh_to_rgb<2>(f, dx, dy, dz, rgb, clamp_bits); break;
                default: load_sh_split<45>(dc, a.shs_rest, idx, a.P, vec, f); sh_to_rgb<3>(f, dx, dy, dz, rgb, clamp_bits); break;
                }
            } else {
                const float* shp = a.shs + (size_t)3 * a.M * idx;
                const bool vec = (((size_t)shp) & 15) == 0;
                switch (a.D) {
                case 0: load_sh<3>(shp, false, f); sh_to_rgb<0>(f, dx, dy, dz, rgb, clamp_bits); break;
                case 1: load_sh<12>(shp, vec, f); sh_to_rgb<1>(f, dx, dy, dz, rgb, clamp_bits); break;
                case 2: load_sh<27>(shp, vec, f); sh_to_rgb<2>(f, dx, dy, dz, rgb, clamp_bits); break;
                default: load_sh<48>(shp, vec, f); sh_to_rgb<3>(f, dx, dy, dz, rgb, clamp_bits); break;
                }
            }
        }
        out.q1.z = rgb[0]; out.q1.w = rgb[1]; out.q2.x = rgb[2];
    }
    if (!inb) return;
    o.radii[oidx] = radius_out;
    o.rec[oidx] = out;
    o.clamped[oidx] = (uint8_t)clamp_bits;
    o.depthkey[oidx] = vis ? __float_as_uint(out.q2.y) - a.key_bias : a.key_cull;
    o.rect[oidx] = rect_out;
    // R = sum of tiles touched: one atomic per warp that has any
    uint32_t tiles = __float_as_uint(out.q2.w);
    const unsigned am = __activemask();
    tiles = __reduce_add_sync(am, tiles);
    if (tiles && (int)(threadIdx.x & 31) == __ffs(am) - 1) atomicAdd(o.total, tiles);
}

// STAGE (SH degree 3 on [P,16,3] storage, 16-byte aligned): the 192-byte SH rows are 60 % of this kernel's traffic.  Read
// per thread (12 x LDG.128 at a 192-byte lane stride) every load instruction touches 32 different lines and the kernel is
// bound by the L1's line rate, not by HBM (measured 4.3 TB/s).  Instead the warp copies the 6 KB its 32 Gaussians occupy
// with fully coalesced 16-byte cp.async into shared memory at the top of the kernel - the copy overlaps the cull /
// covariance / rect arithmetic and costs no registers - and every visible lane then reads its own row from there.
// Rows of Gaussians that turn out culled are fetched for nothing (15 % at C3).
template <bool STAGE>
__global__ void __launch_bounds__(PRE_THREADS) preprocess_fwd_kernel(PreArgs a, CameraArgs ca)
{
    __shared__ Camera cam;
    __shared__ float4 stage[STAGE ? (PRE_THREADS / 32) * 32 * PRE_ROW4 : 1];
    // blockIdx.y = view of a batched call (gsr_forward_batched): its camera, its slice v * P + idx of the per-view
    // outputs, its band of tile rows in the stacked image
    // (the launch's camera arrays and Gaussian parameters start at its first view / its scene; outputs are indexed by the
    // global view a.view_base + blockIdx.y)
    const int lview = blockIdx.y;
    const int view = a.view_base + lview;
    const int lane = threadIdx.x & 31;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    float4* const st = stage + (STAGE ? (threadIdx.x >> 5) * 32 * PRE_ROW4 : 0);
    if (STAGE) {
        const int wbase = idx - lane;                                   // first Gaussian of this warp
        const int n4 = max(0, min(32, a.P - wbase)) * 12;               // float4s of the warp's rows
        const float4* src = reinterpret_cast<const float4*>(a.shs) + (size_t)wbase * 12;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane;
            if (e < n4) {
                const int r = e / 12, c = e - r * 12;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(st + r * PRE_ROW4 + c);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src + e) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    ca.view += 16 * lview;
    if (ca.proj) ca.proj += 16 * lview;
    if (ca.cam) ca.cam += 3 * lview;
    load_camera(cam, ca);
    const bool inb = idx < a.P;
    if (!STAGE && !inb) return;
    const size_t oidx = (size_t)view * a.P + idx;

    PreOut o;
    o.radii = a.radii; o.rec = a.rec; o.clamped = a.clamped; o.depthkey = a.depthkey; o.rect = a.rect; o.total = a.total; o.err = a.err;
    preprocess_gaussian<STAGE, false>(a, cam, o, oidx, view, idx, lane, st, inb, GaussIn{});
}

// ------------------------------------------------------------------ several views of ONE Gaussian set, parameters read once
// gsr_preprocess_views: a step that renders V views through single-view calls reads the Gaussian parameters V times - at SH
// degree 3 that is 236 of the ~300 bytes per Gaussian a view's preprocess moves.  Here a thread keeps its Gaussian (the SH
// row stays staged in shared memory) and walks the views' cameras; every view's results go to that view's own blob, exactly
// as preprocess_fwd_kernel writes them (same device function).
struct PreViews {
    PreOut out[GSR_MAX_VIEWS];
    const float* view[GSR_MAX_VIEWS];
    const float* proj[GSR_MAX_VIEWS];
    const float* cam[GSR_MAX_VIEWS];
};

template <bool STAGE>
__global__ void __launch_bounds__(PRE_THREADS) preprocess_fwd_views_kernel(PreArgs a, CameraArgs ca, const __grid_constant__ PreViews pv, int V)
{
    __shared__ Camera cams[GSR_MAX_VIEWS];
    __shared__ float4 stage[STAGE ? (PRE_THREADS / 32) * 32 * PRE_ROW4 : 1];
    const int lane = threadIdx.x & 31;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    float4* const st = stage + (STAGE ? (threadIdx.x >> 5) * 32 * PRE_ROW4 : 0);
    if (STAGE) {
        const int wbase = idx - lane;                                   // first Gaussian of this warp
        const int n4 = max(0, min(32, a.P - wbase)) * 12;               // float4s of the warp's rows
        const float4* src = reinterpret_cast<const float4*>(a.shs) + (size_t)wbase * 12;
#pragma unroll
        for (int i = 0; i < 12; i++) {
            const int e = i * 32 + lane;
            if (e < n4) {
                const int r = e / 12, c = e - r * 12;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(st + r * PRE_ROW4 + c);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src + e) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    }
    for (int v = 0; v < V; v++) {
        CameraArgs cv = ca;
        cv.view = pv.view[v]; cv.proj = pv.proj[v]; cv.cam = pv.cam[v];
        load_camera(cams[v], cv);
    }
    const bool inb = idx < a.P;
    if (!STAGE && !inb) return;
    GaussIn gi = {};
    if (inb) {
        gi.p = make_float3(a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]);
        const float3 s = {a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]};
        const float4 q = {a.rots[4 * idx], a.rots[4 * idx + 1], a.rots[4 * idx + 2], a.rots[4 * idx + 3]};
        cov3d_from_scale_rot(s, a.mod, q, gi.c6);
        gi.opac = a.opac[idx];
    }
    for (int v = 0; v < V; v++) preprocess_gaussian<STAGE, true>(a, cams[v], pv.out[v], (size_t)idx, 0, idx, lane, st, inb, gi);
}

int launch_preprocess(const GsrFwdArgs& a, int V, int view_base, int V_total, const GeomView& g, cudaStream_t s)
{
    const CameraArgs cam = make_camera_args(a.viewmatrix, a.projmatrix, a.cam_pos, a.tan_fovx, a.tan_fovy, a.W, a.H);
    PreArgs k;
    k.P = a.P; k.D = a.D; k.M = a.M;
    k.means = a.means3D; k.shs = a.shs; k.shs_rest = (a.shs && a.M > 1) ? a.shs_rest : nullptr; k.colors_pre = a.colors_precomp; k.opac = a.opacities;
    k.scales = a.scales; k.rots = a.rotations; k.cov_pre = a.cov3D_precomp; k.mod = a.scale_modifier;
    k.prefiltered = a.prefiltered; k.radii = a.radii; k.rec = g.rec; k.clamped = g.clamped;
    const GeomScratch sc = carve_geom_scratch(g.scratch, a.P * V_total);
    k.view_base = view_base;
    k.depthkey = sc.depthkey; k.rect = sc.rect; k.total = sc.counters;
    k.err = reinterpret_cast<int*>(sc.counters + 8);
    const bool biased = options().radix_depth == 3;
    k.key_bias = biased ? GSR_DEPTH_KEY_BIAS : 0u; k.key_cull = biased ? 0u : 0xFFFFFFFFu;
    const bool stage = options().pre_stage && a.shs && !a.colors_precomp && !k.shs_rest && a.M == 16 && a.D == 3 && (((size_t)a.shs) & 15) == 0;
    if (stage) GSR_LAUNCH(preprocess_fwd_kernel<true>, dim3(cdiv(a.P, PRE_THREADS), V), PRE_THREADS, 0, s, k, cam);
    else GSR_LAUNCH(preprocess_fwd_kernel<false>, dim3(cdiv(a.P, PRE_THREADS), V), PRE_THREADS, 0, s, k, cam);
    return GSR_OK;
}

int launch_preprocess_views(const GsrFwdArgs& a, int n_views, const GsrViewFront* views, const GeomView* g, cudaStream_t s)
{
    const CameraArgs cam = make_camera_args(views[0].viewmatrix, views[0].projmatrix, views[0].cam_pos, a.tan_fovx, a.tan_fovy, a.W, a.H);
    PreArgs k = {};
    k.P = a.P; k.D = a.D; k.M = a.M;
    k.means = a.means3D; k.shs = a.shs; k.shs_rest = (a.shs && a.M > 1) ? a.shs_rest : nullptr; k.colors_pre = nullptr; k.opac = a.opacities;
    k.scales = a.scales; k.rots = a.rotations; k.cov_pre = nullptr; k.mod = a.scale_modifier;
    k.prefiltered = a.prefiltered;
    const bool biased = options().radix_depth == 3;
    k.key_bias = biased ? GSR_DEPTH_KEY_BIAS : 0u; k.key_cull = biased ? 0u : 0xFFFFFFFFu;
    PreViews pv = {};
    for (int v = 0; v < n_views; v++) {
        const GeomScratch sc = carve_geom_scratch(g[v].scratch, a.P);
        PreOut& o = pv.out[v];
        o.radii = views[v].radii; o.rec = g[v].rec; o.clamped = g[v].clamped; o.depthkey = sc.depthkey; o.rect = sc.rect;
        o.total = sc.counters; o.err = reinterpret_cast<int*>(sc.counters + 8);
        pv.view[v] = views[v].viewmatrix; pv.proj[v] = views[v].projmatrix; pv.cam[v] = views[v].cam_pos;
    }
    const bool stage = options().pre_stage && a.shs && !k.shs_rest && a.M == 16 && a.D == 3 && (((size_t)a.shs) & 15) == 0;
    if (stage) GSR_LAUNCH(preprocess_fwd_views_kernel<true>, cdiv(a.P, PRE_THREADS), PRE_THREADS, 0, s, k, cam, pv, n_views);
    else GSR_LAUNCH(preprocess_fwd_views_kernel<false>, cdiv(a.P, PRE_THREADS), PRE_THREADS, 0, s, k, cam, pv, n_views);
    return GSR_OK;
}

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means,
                                                           CameraArgs ca, uint8_t* present)
{
    __shared__ Camera cam;
    load_camera(cam, ca);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
    present[idx] = !(xform4x3(p, cam.view).z <= 0.2f);
}

int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t s)
{
    const CameraArgs cam = make_camera_args(view, nullptr, nullptr, 1.f, 1.f, GSR_TILE, GSR_TILE);
    GSR_LAUNCH(mark_visible_kernel, cdiv(P, 256), 256, 0, s, P, means, cam, present);
    return GSR_OK;
}

}  // namespace gsr
