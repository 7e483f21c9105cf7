/*
 * gsr_oracle.c - CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C, fp32 restatement of the reference's Gaussian rasterizer hot path and
 * of simple-knn's distCUDA2, written from the normative description in SURVEY.md
 * Appendix A.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product path never does.
 *
 * Parity status: the reference ships NO tests, golden vectors or CPU path for this
 * code, so this oracle is pinned by (a) outputs of the
 * unmodified reference CUDA extension (oracle/_ref, built by oracle/build_ref.sh)
 * run on a B200 and committed under tests/golden/ref_*.npz, and (b) the
 * reference's own torch restatements of two sub-steps (utils/sh_utils.py eval_sh,
 * scene/gaussian_model.py build_covariance_from_scaling_rotation), committed as
 * tests/golden/sub_*.npz.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/submodules/).  DGR = diff-gaussian-rasterization/cuda_rasterizer,
 * SK = simple-knn.
 *
 * Arithmetic: IEEE fp32, compiled with -ffp-contract=off, i.e. NO fused
 * multiply-adds; the GPU code contracts, so floats agree to a few ulp, integers
 * agree except where an ulp flips a ceil()/truncation (counted in the tests).
 * Matrix convention follows GLM: m.c[i][j] is column i, row j.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16               /* DGR/config.h:16-17 */

typedef struct { float c[3][3]; } m3;

/* GLM mat3*mat3 (third_party/glm/glm/detail/type_mat3x3.inl:486-516):
 * out[i][j] = a[0][j]*b[i][0] + a[1][j]*b[i][1] + a[2][j]*b[i][2], summed left to right */
static m3 m3_mul(const m3 a, const m3 b)
{
    m3 r;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            r.c[i][j] = a.c[0][j] * b.c[i][0] + a.c[1][j] * b.c[i][1] + a.c[2][j] * b.c[i][2];
    return r;
}
static m3 m3_t(const m3 a)
{
    m3 r;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            r.c[i][j] = a.c[j][i];
    return r;
}
/* GLM constructor order: nine scalars fill column 0, then column 1, then column 2 */
static m3 m3_make(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1, float c2)
{
    m3 r = {{{a0, a1, a2}, {b0, b1, b2}, {c0, c1, c2}}};
    return r;
}

/* DGR/auxiliary.h:58-77 */
static void xform4x3(const float* p, const float* m, float* o)
{
    o[0] = m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12];
    o[1] = m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13];
    o[2] = m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14];
}
static void xform4x4(const float* p, const float* m, float* o)
{
    xform4x3(p, m, o);
    o[3] = m[3] * p[0] + m[7] * p[1] + m[11] * p[2] + m[15];
}

/* DGR/auxiliary.h:41-44: double arithmetic because the literals are doubles */
static float ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* DGR/auxiliary.h:46-56 */
static void tile_rect(float px, float py, int rad, int gx, int gy, int* x0, int* y0, int* x1, int* y1)
{
    *x0 = imin(gx, imax(0, (int)((px - rad) / TILE)));
    *y0 = imin(gy, imax(0, (int)((py - rad) / TILE)));
    *x1 = imin(gx, imax(0, (int)((px + rad + TILE - 1) / TILE)));
    *y1 = imin(gy, imax(0, (int)((py + rad + TILE - 1) / TILE)));
}

/* SH constants, DGR/auxiliary.h:22-39 */
static const float SHC0 = 0.28209479177387814f;
static const float SHC1 = 0.4886025119029199f;
static const float SHC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                              -1.0925484305920792f, 0.5462742152960396f};
static const float SHC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                              -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

/* scale/rotation -> 3D covariance, DGR/forward.cu:118-152 (quaternion NOT normalised, :127) */
static void cov3d_from_scale_rot(const float* s, float mod, const float* q, float* cov6)
{
    m3 S = m3_make(mod * s[0], 0, 0, 0, mod * s[1], 0, 0, 0, mod * s[2]);
    float r = q[0], x = q[1], y = q[2], z = q[3];
    m3 R = m3_make(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    m3 M = m3_mul(S, R);
    m3 Sig = m3_mul(m3_t(M), M);
    cov6[0] = Sig.c[0][0]; cov6[1] = Sig.c[0][1]; cov6[2] = Sig.c[0][2];
    cov6[3] = Sig.c[1][1]; cov6[4] = Sig.c[1][2]; cov6[5] = Sig.c[2][2];
}

/* shared by forward (DGR/forward.cu:74-113) and backward (DGR/backward.cu:160-192) */
static void ewa_setup(const float* mean, float fx, float fy, float tanx, float tany, const float* cov6,
                      const float* V, float* t, float* xmul, float* ymul, m3* W, m3* T, m3* Vrk, m3* cov)
{
    xform4x3(mean, V, t);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    *xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    *ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    m3 J = m3_make(fx / t[2], 0.f, -(fx * t[0]) / (t[2] * t[2]),
                   0.f, fy / t[2], -(fy * t[1]) / (t[2] * t[2]),
                   0.f, 0.f, 0.f);
    *W = m3_make(V[0], V[4], V[8], V[1], V[5], V[9], V[2], V[6], V[10]);
    *T = m3_mul(*W, J);
    *Vrk = m3_make(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
    *cov = m3_mul(m3_mul(m3_t(*T), m3_t(*Vrk)), *T);
}

/* SH -> RGB, DGR/forward.cu:20-71.  sh: [M][3] of this Gaussian */
static void sh_to_rgb(int D, const float* sh, const float* pos, const float* cam, float* rgb, uint8_t* clamped)
{
    float dx = pos[0] - cam[0], dy = pos[1] - cam[1], dz = pos[2] - cam[2];
    float len = sqrtf(dx * dx + dy * dy + dz * dz);
    float x = dx / len, y = dy / len, z = dz / len;
    for (int ch = 0; ch < 3; ch++) {
#define S(k) sh[3 * (k) + ch]
        float res = SHC0 * S(0);
        if (D > 0) {
            res = res - SHC1 * y * S(1) + SHC1 * z * S(2) - SHC1 * x * S(3);
            if (D > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SHC2[0] * xy * S(4) + SHC2[1] * yz * S(5) + SHC2[2] * (2.0f * zz - xx - yy) * S(6) +
                      SHC2[3] * xz * S(7) + SHC2[4] * (xx - yy) * S(8);
                if (D > 2) {
                    res = res + SHC3[0] * y * (3.0f * xx - yy) * S(9) + SHC3[1] * xy * z * S(10) +
                          SHC3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                          SHC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                          SHC3[4] * x * (4.0f * zz - xx - yy) * S(13) + SHC3[5] * z * (xx - yy) * S(14) +
                          SHC3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        res += 0.5f;
        clamped[ch] = res < 0;
        rgb[ch] = fmaxf(res, 0.0f);
    }
}

/*
 * Per-Gaussian preprocess, DGR/forward.cu:155-256 (+ auxiliary.h:139-164 near cull).
 * Outputs are sized [P] (xy: [P,2], cov3d: [P,6], conic_o: [P,4], rgb: [P,3], clamped: [P,3]).
 * For culled Gaussians radii = tiles = 0 and the other outputs are left untouched
 * (the reference leaves them uninitialised).
 */
void orc_preprocess(int P, int D, int M, const float* means, const float* scales, float mod, const float* rots,
                    const float* opac, const float* shs, const float* cov3d_pre, const float* colors_pre,
                    const float* V, const float* PR, const float* cam, int W, int H, float tanx, float tany,
                    int* radii, float* xy, float* depth, float* cov3d, float* conic_o, float* rgb,
                    uint8_t* clamped, uint32_t* tiles)
{
    const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);       /* rasterizer_impl.cu:222-223 */
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0; tiles[i] = 0;
        const float* p = means + 3 * i;
        float pv[3], ph[4];
        xform4x4(p, PR, ph);
        xform4x3(p, V, pv);
        if (pv[2] <= 0.2f) continue;                                   /* auxiliary.h:154 */
        float pw = 1.0f / (ph[3] + 0.0000001f);
        float projx = ph[0] * pw, projy = ph[1] * pw;
        float c6[6];
        const float* cv;
        if (cov3d_pre) cv = cov3d_pre + 6 * i;
        else { cov3d_from_scale_rot(scales + 3 * i, mod, rots + 4 * i, c6); memcpy(cov3d + 6 * i, c6, sizeof c6); cv = c6; }
        float t[3], xm, ym; m3 Wm, T, Vrk, cov;
        ewa_setup(p, fx, fy, tanx, tany, cv, V, t, &xm, &ym, &Wm, &T, &Vrk, &cov);
        float a = cov.c[0][0] + 0.3f, b = cov.c[0][1], c = cov.c[1][1] + 0.3f;
        float det = a * c - b * b;
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float con[3] = {c * det_inv, -b * det_inv, a * det_inv};
        float mid = 0.5f * (a + c);
        float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float rad = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        float px = ndc2pix(projx, W), py = ndc2pix(projy, H);
        int x0, y0, x1, y1;
        tile_rect(px, py, (int)rad, gx, gy, &x0, &y0, &x1, &y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!colors_pre) sh_to_rgb(D, shs + (size_t)3 * M * i, p, cam, rgb + 3 * i, clamped + 3 * i);
        depth[i] = pv[2];
        radii[i] = (int)rad;
        xy[2 * i] = px; xy[2 * i + 1] = py;
        conic_o[4 * i] = con[0]; conic_o[4 * i + 1] = con[1]; conic_o[4 * i + 2] = con[2]; conic_o[4 * i + 3] = opac[i];
        tiles[i] = (uint32_t)((y1 - y0) * (x1 - x0));
    }
}

/* stable LSD radix sort of (key,value) on the low `bits` bits - the semantics of
 * cub::DeviceRadixSort::SortPairs(begin_bit=0,end_bit=bits), rasterizer_impl.cu:303-308 */
static void sort_pairs(uint64_t* k, uint32_t* v, size_t n, int bits)
{
    uint64_t* k2 = (uint64_t*)malloc(n * sizeof *k2);
    uint32_t* v2 = (uint32_t*)malloc(n * sizeof *v2);
    for (int sh = 0; sh < bits; sh += 8) {
        int w = bits - sh < 8 ? bits - sh : 8;
        size_t cnt[257] = {0};
        uint64_t mask = ((uint64_t)1 << w) - 1;
        for (size_t i = 0; i < n; i++) cnt[((k[i] >> sh) & mask) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (size_t i = 0; i < n; i++) { size_t o = cnt[(k[i] >> sh) & mask]++; k2[o] = k[i]; v2[o] = v[i]; }
        memcpy(k, k2, n * sizeof *k); memcpy(v, v2, n * sizeof *v);
    }
    free(k2); free(v2);
}

/* DGR/rasterizer_impl.cu:35-50 */
static int higher_msb(uint32_t n)
{
    int msb = 16, step = 16;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

/* total number of (Gaussian, tile) instances = last element of the inclusive scan, rasterizer_impl.cu:277-281 */
long long orc_count_instances(int P, const uint32_t* tiles)
{
    long long r = 0;
    for (int i = 0; i < P; i++) r += tiles[i];
    return r;
}

/*
 * duplicateWithKeys + SortPairs + identifyTileRanges, DGR/rasterizer_impl.cu:70-138,300-318.
 * keys/vals: [R]; ranges: [2*tiles] zero-initialised here.
 */
void orc_bin(int P, int W, int H, const float* xy, const float* depth, const int* radii,
             long long R, uint64_t* keys, uint32_t* vals, uint32_t* ranges)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    size_t off = 0;
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        int x0, y0, x1, y1;
        tile_rect(xy[2 * i], xy[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
        uint32_t dbits; memcpy(&dbits, depth + i, 4);
        for (int y = y0; y < y1; y++)
            for (int x = x0; x < x1; x++) {
                keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
                vals[off] = (uint32_t)i;
                off++;
            }
    }
    sort_pairs(keys, vals, (size_t)R, 32 + higher_msb((uint32_t)(gx * gy)));
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (long long i = 0; i < R; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
}

/*
 * Forward blending, DGR/forward.cu:261-374, restated per pixel (the tile-lockstep of
 * the CUDA block does not change any per-pixel result).
 * tile_stride/tile_first select a subset of tiles (bench sampling): tiles
 * tile_first, tile_first+tile_stride, ... are processed; others are untouched.
 * pair counters (may be NULL): visited (entries walked before the pixel stopped),
 * blended (entries that contributed).
 */
void orc_blend_forward(int W, int H, const uint32_t* ranges, const uint32_t* list, const float* xy,
                       const float* feat, const float* conic_o, const float* bg,
                       float* final_T, uint32_t* n_contrib, float* out,
                       int tile_first, int tile_stride, long long* visited, long long* blended)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    long long nv = 0, nb = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : nv, nb)
    for (int t = tile_first; t < gx * gy; t += tile_stride) {
        int tx = t % gx, ty = t / gx;
        uint32_t beg = ranges[2 * t], end = ranges[2 * t + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                if (pxi >= W || pyi >= H) continue;
                float pxf = (float)pxi, pyf = (float)pyi;
                float T = 1.0f, C[3] = {0, 0, 0};
                uint32_t contributor = 0, last = 0;
                for (uint32_t j = beg; j < end; j++) {
                    contributor++; nv++;
                    uint32_t g = list[j];
                    float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
                    const float* co = conic_o + 4 * g;
                    float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > 0.0f) continue;
                    float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break;                /* done: this entry is not blended */
                    for (int ch = 0; ch < 3; ch++) C[ch] += feat[3 * g + ch] * alpha * T;
                    T = test_T;
                    last = contributor; nb++;
                }
                size_t pid = (size_t)W * pyi + pxi;
                final_T[pid] = T;
                n_contrib[pid] = last;
                for (int ch = 0; ch < 3; ch++) out[(size_t)ch * H * W + pid] = C[ch] + T * bg[ch];
            }
    }
    if (visited) *visited = nv;
    if (blended) *blended = nb;
}

/*
 * Backward blending, DGR/backward.cu:399-557.  Per-Gaussian sums are accumulated
 * in DOUBLE (acc arrays) so the oracle value is the order-independent "true" sum;
 * GPU float atomics are compared against it with the 1e-3 relative tolerance.
 * d_mean2d: [P,3] (z unused), d_conic: [P,4] (.z unused), d_opac: [P], d_color: [P,3]; all double, zeroed by caller.
 */
void orc_blend_backward(int W, int H, const uint32_t* ranges, const uint32_t* list, const float* bg,
                        const float* xy, const float* conic_o, const float* feat, const float* final_T,
                        const uint32_t* n_contrib, const float* dL_dpix,
                        double* d_mean2d, double* d_conic, double* d_opac, double* d_color,
                        int tile_first, int tile_stride)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);     /* backward.cu:460-461 */
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = tile_first; t < gx * gy; t += tile_stride) {
        int tx = t % gx, ty = t / gx;
        uint32_t beg = ranges[2 * t], end = ranges[2 * t + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                if (pxi >= W || pyi >= H) continue;
                size_t pid = (size_t)W * pyi + pxi;
                float pxf = (float)pxi, pyf = (float)pyi;
                const float T_final = final_T[pid];
                float T = T_final;
                uint32_t contributor = end - beg;
                const uint32_t last_contributor = n_contrib[pid];
                float accum_rec[3] = {0, 0, 0}, dpix[3], last_alpha = 0, last_color[3] = {0, 0, 0};
                for (int ch = 0; ch < 3; ch++) dpix[ch] = dL_dpix[(size_t)ch * H * W + pid];
                for (uint32_t jj = end; jj-- > beg;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    uint32_t g = list[jj];
                    float dx = xy[2 * g] - pxf, dy = xy[2 * g + 1] - pyf;
                    const float* co = conic_o + 4 * g;
                    float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                    if (power > 0.0f) continue;
                    float G = expf(power);
                    float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    for (int ch = 0; ch < 3; ch++) {
                        float c = feat[3 * g + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dalpha += (c - accum_rec[ch]) * dpix[ch];
                        double add = dchannel_dcolor * dpix[ch];
#pragma omp atomic
                        d_color[3 * (size_t)g + ch] += add;
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    float bg_dot = 0;
                    for (int ch = 0; ch < 3; ch++) bg_dot += bg[ch] * dpix[ch];
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                    float dL_dG = co[3] * dL_dalpha;
                    float gdx = G * dx, gdy = G * dy;
                    float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    float dG_ddely = -gdy * co[2] - gdx * co[1];
                    double v0 = dL_dG * dG_ddelx * ddelx_dx, v1 = dL_dG * dG_ddely * ddely_dy;
                    double v2 = -0.5f * gdx * dx * dL_dG, v3 = -0.5f * gdx * dy * dL_dG, v4 = -0.5f * gdy * dy * dL_dG;
                    double v5 = G * dL_dalpha;
#pragma omp atomic
                    d_mean2d[3 * (size_t)g] += v0;
#pragma omp atomic
                    d_mean2d[3 * (size_t)g + 1] += v1;
#pragma omp atomic
                    d_conic[4 * (size_t)g] += v2;
#pragma omp atomic
                    d_conic[4 * (size_t)g + 1] += v3;
#pragma omp atomic
                    d_conic[4 * (size_t)g + 3] += v4;
#pragma omp atomic
                    d_opac[g] += v5;
                }
            }
    }
}

/* SH backward, DGR/backward.cu:20-139.  Adds the direction term into d_mean[3]. */
static void sh_backward(int D, int M, const float* sh, const float* pos, const float* cam, const uint8_t* clamped,
                        const float* dL_dcolor, float* d_mean, float* d_sh)
{
    float ox = pos[0] - cam[0], oy = pos[1] - cam[1], oz = pos[2] - cam[2];
    float len = sqrtf(ox * ox + oy * oy + oz * oz);
    float x = ox / len, y = oy / len, z = oz / len;
    float dRGB[3];
    for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[ch] * (clamped[ch] ? 0.f : 1.f);
    float gx[3] = {0, 0, 0}, gy[3] = {0, 0, 0}, gz[3] = {0, 0, 0};
    (void)M;
#define S(k, ch) sh[3 * (k) + (ch)]
#define OUT(k, w) do { for (int ch = 0; ch < 3; ch++) d_sh[3 * (k) + ch] = (w) * dRGB[ch]; } while (0)
    OUT(0, SHC0);
    if (D > 0) {
        OUT(1, -SHC1 * y); OUT(2, SHC1 * z); OUT(3, -SHC1 * x);
        for (int ch = 0; ch < 3; ch++) { gx[ch] = -SHC1 * S(3, ch); gy[ch] = -SHC1 * S(1, ch); gz[ch] = SHC1 * S(2, ch); }
        if (D > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            OUT(4, SHC2[0] * xy); OUT(5, SHC2[1] * yz); OUT(6, SHC2[2] * (2.f * zz - xx - yy));
            OUT(7, SHC2[3] * xz); OUT(8, SHC2[4] * (xx - yy));
            for (int ch = 0; ch < 3; ch++) {
                gx[ch] += SHC2[0] * y * S(4, ch) + SHC2[2] * 2.f * -x * S(6, ch) + SHC2[3] * z * S(7, ch) + SHC2[4] * 2.f * x * S(8, ch);
                gy[ch] += SHC2[0] * x * S(4, ch) + SHC2[1] * z * S(5, ch) + SHC2[2] * 2.f * -y * S(6, ch) + SHC2[4] * 2.f * -y * S(8, ch);
                gz[ch] += SHC2[1] * y * S(5, ch) + SHC2[2] * 2.f * 2.f * z * S(6, ch) + SHC2[3] * x * S(7, ch);
            }
            if (D > 2) {
                OUT(9, SHC3[0] * y * (3.f * xx - yy)); OUT(10, SHC3[1] * xy * z);
                OUT(11, SHC3[2] * y * (4.f * zz - xx - yy)); OUT(12, SHC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                OUT(13, SHC3[4] * x * (4.f * zz - xx - yy)); OUT(14, SHC3[5] * z * (xx - yy));
                OUT(15, SHC3[6] * x * (xx - 3.f * yy));
                for (int ch = 0; ch < 3; ch++) {
                    gx[ch] += (SHC3[0] * S(9, ch) * 3.f * 2.f * xy + SHC3[1] * S(10, ch) * yz + SHC3[2] * S(11, ch) * -2.f * xy +
                               SHC3[3] * S(12, ch) * -3.f * 2.f * xz + SHC3[4] * S(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                               SHC3[5] * S(14, ch) * 2.f * xz + SHC3[6] * S(15, ch) * 3.f * (xx - yy));
                    gy[ch] += (SHC3[0] * S(9, ch) * 3.f * (xx - yy) + SHC3[1] * S(10, ch) * xz +
                               SHC3[2] * S(11, ch) * (-3.f * yy + 4.f * zz - xx) + SHC3[3] * S(12, ch) * -3.f * 2.f * yz +
                               SHC3[4] * S(13, ch) * -2.f * xy + SHC3[5] * S(14, ch) * -2.f * yz +
                               SHC3[6] * S(15, ch) * -3.f * 2.f * xy);
                    gz[ch] += (SHC3[1] * S(10, ch) * xy + SHC3[2] * S(11, ch) * 4.f * 2.f * yz +
                               SHC3[3] * S(12, ch) * 3.f * (2.f * zz - xx - yy) + SHC3[4] * S(13, ch) * 4.f * 2.f * xz +
                               SHC3[5] * S(14, ch) * (xx - yy));
                }
            }
        }
    }
#undef S
#undef OUT
    float ddir[3] = {gx[0] * dRGB[0] + gx[1] * dRGB[1] + gx[2] * dRGB[2],
                     gy[0] * dRGB[0] + gy[1] * dRGB[1] + gy[2] * dRGB[2],
                     gz[0] * dRGB[0] + gz[1] * dRGB[1] + gz[2] * dRGB[2]};
    /* dnormvdv, DGR/auxiliary.h:107-117 */
    float s2 = ox * ox + oy * oy + oz * oz;
    float inv = 1.0f / sqrtf(s2 * s2 * s2);
    d_mean[0] += ((+s2 - ox * ox) * ddir[0] - oy * ox * ddir[1] - oz * ox * ddir[2]) * inv;
    d_mean[1] += (-ox * oy * ddir[0] + (s2 - oy * oy) * ddir[1] - oz * oy * ddir[2]) * inv;
    d_mean[2] += (-ox * oz * ddir[0] - oy * oz * ddir[1] + (s2 - oz * oz) * ddir[2]) * inv;
}

/* cov3D -> scale / rotation gradients, DGR/backward.cu:278-341 (no quaternion-normalisation Jacobian, :340) */
static void cov3d_backward(const float* sc, float mod, const float* q, const float* dcov, float* d_scale, float* d_rot)
{
    float r = q[0], x = q[1], y = q[2], z = q[3];
    m3 R = m3_make(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    float s[3] = {mod * sc[0], mod * sc[1], mod * sc[2]};
    m3 S = m3_make(s[0], 0, 0, 0, s[1], 0, 0, 0, s[2]);
    m3 M = m3_mul(S, R);
    m3 dSig = m3_make(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                      0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
    m3 M2 = M;                                   /* 2.0f * M (GLM scalar*mat: every element times 2) */
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) M2.c[i][j] = 2.0f * M.c[i][j];
    m3 dM = m3_mul(M2, dSig);
    m3 Rt = m3_t(R), dMt = m3_t(dM);
    for (int k = 0; k < 3; k++)
        d_scale[k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
    for (int k = 0; k < 3; k++) for (int j = 0; j < 3; j++) dMt.c[k][j] *= s[k];
#define D_(i, j) dMt.c[i][j]
    d_rot[0] = 2 * z * (D_(0, 1) - D_(1, 0)) + 2 * y * (D_(2, 0) - D_(0, 2)) + 2 * x * (D_(1, 2) - D_(2, 1));
    d_rot[1] = 2 * y * (D_(1, 0) + D_(0, 1)) + 2 * z * (D_(2, 0) + D_(0, 2)) + 2 * r * (D_(1, 2) - D_(2, 1)) - 4 * x * (D_(2, 2) + D_(1, 1));
    d_rot[2] = 2 * x * (D_(1, 0) + D_(0, 1)) + 2 * r * (D_(2, 0) - D_(0, 2)) + 2 * z * (D_(1, 2) + D_(2, 1)) - 4 * y * (D_(2, 2) + D_(0, 0));
    d_rot[3] = 2 * r * (D_(0, 1) - D_(1, 0)) + 2 * x * (D_(2, 0) + D_(0, 2)) + 2 * y * (D_(1, 2) + D_(2, 1)) - 4 * z * (D_(1, 1) + D_(0, 0));
#undef D_
}

/*
 * Backward of preprocess = computeCov2DCUDA (DGR/backward.cu:144-274) followed by
 * preprocessCUDA (DGR/backward.cu:346-396), per visible Gaussian (radii > 0).
 * Inputs d_mean2d [P,3], d_conic [P,4], d_color [P,3] in fp32.
 * Outputs (all zero for culled Gaussians): d_mean3d [P,3], d_cov3d [P,6], d_sh [P,M,3], d_scale [P,3], d_rot [P,4].
 * d_color is only read (it is an output of the blending backward).
 */
void orc_preprocess_backward(int P, int D, int M, const float* means, const int* radii, const float* shs,
                             const uint8_t* clamped, const float* scales, const float* rots, float mod,
                             const float* cov3d, const float* V, const float* PR, const float* cam,
                             int W, int H, float tanx, float tany,
                             const float* d_mean2d, const float* d_conic, const float* d_color,
                             float* d_mean3d, float* d_cov3d, float* d_sh, float* d_scale, float* d_rot)
{
    const float fy = H / (2.0f * tany), fx = W / (2.0f * tanx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float* dm = d_mean3d + 3 * (size_t)i;
        float* dc = d_cov3d + 6 * (size_t)i;
        dm[0] = dm[1] = dm[2] = 0; for (int k = 0; k < 6; k++) dc[k] = 0;
        if (d_sh) memset(d_sh + (size_t)3 * M * i, 0, sizeof(float) * 3 * M);
        if (d_scale) { d_scale[3 * i] = d_scale[3 * i + 1] = d_scale[3 * i + 2] = 0; }
        if (d_rot) { d_rot[4 * i] = d_rot[4 * i + 1] = d_rot[4 * i + 2] = d_rot[4 * i + 3] = 0; }
        if (!(radii[i] > 0)) continue;
        const float* p = means + 3 * i;
        const float* cv = cov3d + 6 * (size_t)i;
        /* --- computeCov2DCUDA --- */
        float dcon[3] = {d_conic[4 * i], d_conic[4 * i + 1], d_conic[4 * i + 3]};
        float t[3], xm, ym; m3 Wm, T, Vrk, cov;
        ewa_setup(p, fx, fy, tanx, tany, cv, V, t, &xm, &ym, &Wm, &T, &Vrk, &cov);
        float a = cov.c[0][0] + 0.3f, b = cov.c[0][1], c = cov.c[1][1] + 0.3f;
        float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcon[0] + 2 * b * c * dcon[1] + (denom - a * c) * dcon[2]);
            dL_dc = denom2inv * (-a * a * dcon[2] + 2 * a * b * dcon[1] + (denom - a * c) * dcon[0]);
            dL_db = denom2inv * 2 * (b * c * dcon[0] - (denom + 2 * b * b) * dcon[1] + a * b * dcon[2]);
#define T_(i, j) T.c[i][j]
            dc[0] = (T_(0, 0) * T_(0, 0) * dL_da + T_(0, 0) * T_(1, 0) * dL_db + T_(1, 0) * T_(1, 0) * dL_dc);
            dc[3] = (T_(0, 1) * T_(0, 1) * dL_da + T_(0, 1) * T_(1, 1) * dL_db + T_(1, 1) * T_(1, 1) * dL_dc);
            dc[5] = (T_(0, 2) * T_(0, 2) * dL_da + T_(0, 2) * T_(1, 2) * dL_db + T_(1, 2) * T_(1, 2) * dL_dc);
            dc[1] = 2 * T_(0, 0) * T_(0, 1) * dL_da + (T_(0, 0) * T_(1, 1) + T_(0, 1) * T_(1, 0)) * dL_db + 2 * T_(1, 0) * T_(1, 1) * dL_dc;
            dc[2] = 2 * T_(0, 0) * T_(0, 2) * dL_da + (T_(0, 0) * T_(1, 2) + T_(0, 2) * T_(1, 0)) * dL_db + 2 * T_(1, 0) * T_(1, 2) * dL_dc;
            dc[4] = 2 * T_(0, 2) * T_(0, 1) * dL_da + (T_(0, 1) * T_(1, 2) + T_(0, 2) * T_(1, 1)) * dL_db + 2 * T_(1, 1) * T_(1, 2) * dL_dc;
        }
#define V_(i, j) Vrk.c[i][j]
        float dT00 = 2 * (T_(0, 0) * V_(0, 0) + T_(0, 1) * V_(0, 1) + T_(0, 2) * V_(0, 2)) * dL_da + (T_(1, 0) * V_(0, 0) + T_(1, 1) * V_(0, 1) + T_(1, 2) * V_(0, 2)) * dL_db;
        float dT01 = 2 * (T_(0, 0) * V_(1, 0) + T_(0, 1) * V_(1, 1) + T_(0, 2) * V_(1, 2)) * dL_da + (T_(1, 0) * V_(1, 0) + T_(1, 1) * V_(1, 1) + T_(1, 2) * V_(1, 2)) * dL_db;
        float dT02 = 2 * (T_(0, 0) * V_(2, 0) + T_(0, 1) * V_(2, 1) + T_(0, 2) * V_(2, 2)) * dL_da + (T_(1, 0) * V_(2, 0) + T_(1, 1) * V_(2, 1) + T_(1, 2) * V_(2, 2)) * dL_db;
        float dT10 = 2 * (T_(1, 0) * V_(0, 0) + T_(1, 1) * V_(0, 1) + T_(1, 2) * V_(0, 2)) * dL_dc + (T_(0, 0) * V_(0, 0) + T_(0, 1) * V_(0, 1) + T_(0, 2) * V_(0, 2)) * dL_db;
        float dT11 = 2 * (T_(1, 0) * V_(1, 0) + T_(1, 1) * V_(1, 1) + T_(1, 2) * V_(1, 2)) * dL_dc + (T_(0, 0) * V_(1, 0) + T_(0, 1) * V_(1, 1) + T_(0, 2) * V_(1, 2)) * dL_db;
        float dT12 = 2 * (T_(1, 0) * V_(2, 0) + T_(1, 1) * V_(2, 1) + T_(1, 2) * V_(2, 2)) * dL_dc + (T_(0, 0) * V_(2, 0) + T_(0, 1) * V_(2, 1) + T_(0, 2) * V_(2, 2)) * dL_db;
#undef V_
#undef T_
#define W_(i, j) Wm.c[i][j]
        float dJ00 = W_(0, 0) * dT00 + W_(0, 1) * dT01 + W_(0, 2) * dT02;
        float dJ02 = W_(2, 0) * dT00 + W_(2, 1) * dT01 + W_(2, 2) * dT02;
        float dJ11 = W_(1, 0) * dT10 + W_(1, 1) * dT11 + W_(1, 2) * dT12;
        float dJ12 = W_(2, 0) * dT10 + W_(2, 1) * dT11 + W_(2, 2) * dT12;
#undef W_
        float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        float dtx = xm * -fx * tz2 * dJ02;
        float dty = ym * -fy * tz2 * dJ12;
        float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t[0]) * tz3 * dJ02 + (2 * fy * t[1]) * tz3 * dJ12;
        /* transformVec4x3Transpose, auxiliary.h:89-97; ASSIGNMENT (backward.cu:273) */
        dm[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
        dm[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
        dm[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
        /* --- preprocessCUDA (backward) --- */
        float mh[4]; xform4x4(p, PR, mh);
        float mw = 1.0f / (mh[3] + 0.0000001f);
        float mul1 = (PR[0] * p[0] + PR[4] * p[1] + PR[8] * p[2] + PR[12]) * mw * mw;
        float mul2 = (PR[1] * p[0] + PR[5] * p[1] + PR[9] * p[2] + PR[13]) * mw * mw;
        float g2x = d_mean2d[3 * i], g2y = d_mean2d[3 * i + 1];
        dm[0] += (PR[0] * mw - PR[3] * mul1) * g2x + (PR[1] * mw - PR[3] * mul2) * g2y;
        dm[1] += (PR[4] * mw - PR[7] * mul1) * g2x + (PR[5] * mw - PR[7] * mul2) * g2y;
        dm[2] += (PR[8] * mw - PR[11] * mul1) * g2x + (PR[9] * mw - PR[11] * mul2) * g2y;
        if (shs) sh_backward(D, M, shs + (size_t)3 * M * i, p, cam, clamped + 3 * i, d_color + 3 * i, dm, d_sh + (size_t)3 * M * i);
        if (scales) cov3d_backward(scales + 3 * i, mod, rots + 4 * i, dc, d_scale + 3 * i, d_rot + 4 * i);
    }
}

/* checkFrustum, DGR/rasterizer_impl.cu:54-66 */
void orc_mark_visible(int P, const float* means, const float* V, uint8_t* present)
{
    for (int i = 0; i < P; i++) { float pv[3]; xform4x3(means + 3 * i, V, pv); present[i] = pv[2] > 0.2f; }
}

/* ------------------------------------------------------------------ distCUDA2 */

static void kbest3(const float* a, const float* b, float* best)    /* SK/simple_knn.cu:131-145 */
{
    float dx = b[0] - a[0], dy = b[1] - a[1], dz = b[2] - a[2];
    float d = dx * dx + dy * dy + dz * dz;
    for (int j = 0; j < 3; j++) if (best[j] > d) { float t = best[j]; best[j] = d; d = t; }
}

/* definition-level oracle: exact mean of the three smallest squared distances to other indices */
void orc_knn_brute(int P, const float* pts, float* out)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        for (int j = 0; j < P; j++) if (j != i) kbest3(pts + 3 * i, pts + 3 * j, best);
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}

static uint32_t spread10(uint32_t x)                                /* SK/simple_knn.cu:45-52 */
{
    x = (x | (x << 16)) & 0x030000FF; x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3; x = (x | (x << 2)) & 0x09249249;
    return x;
}

/* the reference algorithm itself, SK/simple_knn.cu:185-221: bbox with init 0, Morton order,
 * boxes of 1024, reject radius from the +-3 Morton neighbours, then scan every box closer than that */
void orc_knn_morton(int P, const float* pts, float* out, uint32_t* morton_out /* may be NULL, [P] */)
{
    float mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
    for (int i = 0; i < P; i++) for (int k = 0; k < 3; k++) { mn[k] = fminf(mn[k], pts[3 * i + k]); mx[k] = fmaxf(mx[k], pts[3 * i + k]); }
    uint64_t* kk = (uint64_t*)malloc(sizeof *kk * (size_t)P);
    uint32_t* idx = (uint32_t*)malloc(sizeof *idx * (size_t)P);
    for (int i = 0; i < P; i++) {
        uint32_t c[3];
        for (int k = 0; k < 3; k++) c[k] = spread10((uint32_t)(((pts[3 * i + k] - mn[k]) / (mx[k] - mn[k])) * 1023));
        uint32_t code = c[0] | (c[1] << 1) | (c[2] << 2);
        if (morton_out) morton_out[i] = code;
        kk[i] = code; idx[i] = (uint32_t)i;
    }
    sort_pairs(kk, idx, (size_t)P, 32);
    const int BOX = 1024, nb = (P + BOX - 1) / BOX;
    float* bmin = (float*)malloc(sizeof(float) * 3 * nb); float* bmax = (float*)malloc(sizeof(float) * 3 * nb);
    for (int b = 0; b < nb; b++) {
        for (int k = 0; k < 3; k++) { bmin[3 * b + k] = FLT_MAX; bmax[3 * b + k] = -FLT_MAX; }
        for (int i = b * BOX; i < imin(P, (b + 1) * BOX); i++)
            for (int k = 0; k < 3; k++) { float v = pts[3 * idx[i] + k]; bmin[3 * b + k] = fminf(bmin[3 * b + k], v); bmax[3 * b + k] = fmaxf(bmax[3 * b + k], v); }
    }
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < P; i++) {
        const float* p = pts + 3 * idx[i];
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        for (int j = imax(0, i - 3); j <= imin(P - 1, i + 3); j++) if (j != i) kbest3(p, pts + 3 * idx[j], best);
        float reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
        for (int b = 0; b < nb; b++) {
            float d = 0;                                            /* distBoxPoint, SK/simple_knn.cu:119-129 */
            for (int k = 0; k < 3; k++)
                if (p[k] < bmin[3 * b + k] || p[k] > bmax[3 * b + k]) {
                    float e = fminf(fabsf(p[k] - bmin[3 * b + k]), fabsf(p[k] - bmax[3 * b + k]));
                    d += e * e;
                }
            if (d > reject || d > best[2]) continue;
            for (int j = b * BOX; j < imin(P, (b + 1) * BOX); j++) if (j != i) kbest3(p, pts + 3 * idx[j], best);
        }
        out[idx[i]] = (best[0] + best[1] + best[2]) / 3.0f;
    }
    free(kk); free(idx); free(bmin); free(bmax);
}

void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
