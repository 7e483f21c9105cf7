// blend.cu - per-tile alpha blending, forward and backward.
//
// Replaces (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer):
//   blend_fwd_kernel  <- renderCUDA<3> (forward)   forward.cu:261-374
//   blend_bwd_kernel  <- renderCUDA<3> (backward)  backward.cu:399-557
//
// One CTA per 16x16 tile, one thread per pixel.  Tile-list entries are staged in shared memory as
// packed records gathered with 16-byte cp.async copies (position+conic+opacity+colour: everything
// the inner loop touches, so the loop never goes to global memory), double buffered so the gather
// of batch b+1 overlaps the blending of batch b with a single barrier per batch.
// The per-pixel arithmetic keeps the reference's expression order: skip decisions (power > 0,
// alpha < 1/255, T < 1e-4) must be identical or a pixel can differ by up to 1/255.
//
// Backward: every lane computes its pixel's contribution to the 9 per-Gaussian gradient terms, a
// transposing warp butterfly (9 + 5 shuffles instead of 45) reduces them, and 9 lanes issue one
// red.global.add each into a packed 48-byte accumulator row (one or two L2 sectors per warp-entry
// instead of the reference's 9 atomics per pixel-entry into four arrays).
#include "gsr_common.cuh"

namespace gsr {

constexpr int BL_THREADS = GSR_TILE * GSR_TILE;   // 256

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem)
{
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

struct BlendStage {
    float4 q0[BL_THREADS];   // pix.x pix.y conic.x conic.y
    float4 q1[BL_THREADS];   // conic.z opacity r g
    float  b[BL_THREADS];    // b
};

// issue the gather of one batch: entry `pos` of the sorted list (if in range) into slot threadIdx
__device__ __forceinline__ void stage_gather(BlendStage& st, const SplatRec* __restrict__ rec, uint32_t id, bool valid)
{
    if (valid) {
        const SplatRec* r = rec + id;
        cp_async16(&st.q0[threadIdx.x], &r->q0);
        cp_async16(&st.q1[threadIdx.x], &r->q1);
        cp_async4(&st.b[threadIdx.x], &r->q2);
    }
    cp_async_commit();
}

__global__ void __launch_bounds__(BL_THREADS)
blend_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                 int W, int H, const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                 float* __restrict__ out)
{
    __shared__ BlendStage stage[2];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const uint32_t px = blockIdx.x * GSR_TILE + tx, py = blockIdx.y * GSR_TILE + ty;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    bool done = !inside;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;

    float T = 1.0f;
    uint32_t contributor = 0, last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f;

    // prologue: ids of batches 0 and 1, gather of batch 0
    uint32_t id_next = 0;
    {
        const bool v0 = (int)threadIdx.x < total;
        const uint32_t id0 = v0 ? vals[range.x + threadIdx.x] : 0;
        stage_gather(stage[0], rec, id0, v0);
        const int p1 = BL_THREADS + threadIdx.x;
        if (p1 < total) id_next = vals[range.x + p1];
    }
    int todo = total;
    for (int i = 0; i < rounds; i++, todo -= BL_THREADS) {
        cp_async_wait<0>();
        // barrier: batch i is visible to everyone, everyone has left batch i-1 (its buffer is reused below)
        const int num_done = __syncthreads_count(done);
        if (num_done == BL_THREADS) break;
        // gather batch i+1 into the other buffer, fetch the ids of batch i+2
        {
            const int p1 = (i + 1) * BL_THREADS + threadIdx.x;
            stage_gather(stage[(i + 1) & 1], rec, id_next, p1 < total);
            const int p2 = p1 + BL_THREADS;
            if (p2 < total) id_next = vals[range.x + p2];
        }
        const BlendStage& st = stage[i & 1];
        const int n = min(BL_THREADS, todo);
        for (int j = 0; !done && j < n; j++) {
            contributor++;
            const float4 a = st.q0[j];
            const float4 c = st.q1[j];
            const float dx = a.x - pxf, dy = a.y - pyf;
            const float power = -0.5f * (a.z * dx * dx + c.x * dy * dy) - a.w * dx * dy;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, c.y * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) { done = true; continue; }
            const float at = alpha * T;     // colour weight; (c * alpha) * T in the reference, same value to 1 ulp
            C0 = fmaf(c.z, at, C0);
            C1 = fmaf(c.w, at, C1);
            C2 = fmaf(st.b[j], at, C2);
            T = test_T;
            last_contributor = contributor;
        }
    }
    cp_async_wait<0>();
    if (inside) {
        const size_t pix = (size_t)W * py + px;
        final_T[pix] = T;
        n_contrib[pix] = last_contributor;
        const size_t HW = (size_t)H * W;
        out[pix] = C0 + T * bg[0];
        out[HW + pix] = C1 + T * bg[1];
        out[2 * HW + pix] = C2 + T * bg[2];
    }
}

// ------------------------------------------------------------------------------------------ sub-tile culling
// Each warp owns an 8x4 pixel sub-tile (2 x 4 sub-tiles per 16x16 tile).  Before a batch is blended, the
// thread that gathered entry t decides for which of the 8 sub-tiles the entry can possibly be blended:
// alpha >= 1/255 needs  q(d) = 0.5*(A dx^2 + C dy^2) + B dx dy <= ln(255*opacity) =: tau.  The test is exact up
// to rounding: q is a convex quadratic, so its minimum over the sub-tile's pixel-centre rectangle is attained
// either on the vertical edge nearest the centre (u = cu) at the clamped 1-D minimiser v* = -B cu / C, or on the
// nearest horizontal edge (v = cv) at u* = -B cv / A (both candidates are points of the rectangle, so taking
// the smaller never under-estimates; when the centre is inside both are 0).  A sub-tile whose minimum exceeds
// tau cannot contain a pixel the reference would blend, so dropping the entry for that warp leaves every
// result bit-identical.  tau is inflated for the rounding of the per-pixel fp32 power (error <= ~1e-6 * q *
// AC/det), of the approximate log / reciprocals and of this evaluation; the rectangle grows by 1e-3 px;
// ill-conditioned conics (AC/det > 1e4) are never culled.
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ uint32_t subtile_mask(const float4 a, const float4 c, float X0, float Y0)
{
    const float o = c.y;
    if (o < 1.0f / 255.0f) return 0u;                 // alpha = min(0.99, o*exp(p<=0)) <= o < 1/255 everywhere
    const float A = a.z, B = a.w, C = c.x;
    const float det = A * C - B * B;
    const float kappa = A * C * rcp_fast(det);
    if (!(kappa <= 1.0e4f) || !(det > 0.f) || !(A > 0.f) || !(C > 0.f)) return 0xFFu;
    const float tau = (__logf(255.0f * o) + 1.0e-3f) * (1.0f + 6.0e-6f * kappa) * 1.0004f;
    const float nBC = -B * rcp_fast(C), nBA = -B * rcp_fast(A);
    const float hA = 0.5f * A, hC = 0.5f * C;
    float u0[2], u1[2], vst[2], qa[2], bu[2];
#pragma unroll
    for (int sx = 0; sx < 2; sx++) {
        u0[sx] = (X0 + 8.0f * sx - 1.0e-3f) - a.x; u1[sx] = u0[sx] + 7.002f;
        const float cu = fminf(fmaxf(0.0f, u0[sx]), u1[sx]);
        vst[sx] = nBC * cu; qa[sx] = hA * cu * cu; bu[sx] = B * cu;
    }
    uint32_t m = 0;
#pragma unroll
    for (int sy = 0; sy < 4; sy++) {
        const float v0 = (Y0 + 4.0f * sy - 1.0e-3f) - a.y, v1 = v0 + 3.002f;
        const float cv = fminf(fmaxf(0.0f, v0), v1);
        const float ust = nBA * cv, qc = hC * cv * cv, bv = B * cv;
#pragma unroll
        for (int sx = 0; sx < 2; sx++) {
            const float vs = fminf(fmaxf(vst[sx], v0), v1);
            const float q1 = fmaf(vs, fmaf(hC, vs, bu[sx]), qa[sx]);
            const float us = fminf(fmaxf(ust, u0[sx]), u1[sx]);
            const float q2 = fmaf(us, fmaf(hA, us, bv), qc);
            if (fminf(q1, q2) <= tau) m |= 1u << (sy * 2 + sx);       // warp w: sx = w & 1, sy = w >> 1
        }
    }
    return m;
}

// ------------------------------------------------------------------------------------------ forward with sub-tile culling
// Staging as in blend_fwd_kernel plus the sub-tile masks; the per-warp entry loop takes up to UMAX surviving entries per trip:
// their alphas (independent of the pixel's running state) are evaluated side by side, which doubles the
// instruction-level parallelism of the latency chain LDS -> power -> expf, halves the loop overhead, and the
// sequential part (T test, colour accumulation) is predicated instead of branched.  Arithmetic per pixel is
// unchanged (same expressions, same order), so every output is bit-identical to variants 1 and 2.
__device__ __forceinline__ float entry_alpha(const float4 a, const float4 c, float pxf, float pyf, bool& ok)
{
    const float dx = a.x - pxf, dy = a.y - pyf;
    const float power = -0.5f * (a.z * dx * dx + c.x * dy * dy) - a.w * dx * dy;
    const float alpha = fminf(0.99f, c.y * expf(power));
    ok = !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
    return alpha;
}

__device__ __forceinline__ void blend_step(float alpha, bool ok, float r, float g, float b, uint32_t pos, bool& done, float& T,
                                           float& C0, float& C1, float& C2, uint32_t& last)
{
    const float test_T = T * (1 - alpha);
    const bool v = ok && !done;
    const bool stop = v && (test_T < 0.0001f);
    done = done || stop;
    if (v && !stop) {
        const float at = alpha * T;         // colour weight; (c * alpha) * T in the reference, same value to 1 ulp
        C0 = fmaf(r, at, C0);
        C1 = fmaf(g, at, C1);
        C2 = fmaf(b, at, C2);
        T = test_T;
        last = pos;
    }
}

// U surviving entries of chunk `jb` (taken from the low set bits of m) in one trip
template <int U>
__device__ __forceinline__ void blend_trip(const BlendStage& st, uint32_t& m, int jb, uint32_t pos_base, float pxf, float pyf,
                                           bool& done, float& T, float& C0, float& C1, float& C2, uint32_t& last)
{
    int j[U];
    float4 a[U], c[U];
    float b[U], al[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; u++) { j[u] = jb + __ffs(m) - 1; m &= m - 1; }
#pragma unroll
    for (int u = 0; u < U; u++) { a[u] = st.q0[j[u]]; c[u] = st.q1[j[u]]; b[u] = st.b[j[u]]; }
#pragma unroll
    for (int u = 0; u < U; u++) al[u] = entry_alpha(a[u], c[u], pxf, pyf, ok[u]);
#pragma unroll
    for (int u = 0; u < U; u++) blend_step(al[u], ok[u], c[u].z, c[u].w, b[u], pos_base + j[u], done, T, C0, C1, C2, last);
}

template <int UMAX, int MINB>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_fwd_v3_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    int W, int H, const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                    float* __restrict__ out, uint32_t* __restrict__ planes, uint32_t plane_words, int gyv,
                    const uint32_t* __restrict__ tile_order)
{
    __shared__ BlendStage stage[2];
    __shared__ uint32_t bits[2][8][8];     // [buffer][consumer warp][chunk of 32 entries]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // the CTA's tile: the launch order (longest lists first) when given, else the grid position
    const uint32_t lin = blockIdx.y * gridDim.x + blockIdx.x;
    const uint32_t tile = tile_order ? tile_order[lin] : lin;
    const uint32_t trow = tile / gridDim.x, bx = tile - trow * gridDim.x;
    // batched calls stack the views' tile rows: tile row trow = view * gyv + by; pixel coordinates stay per view
    const uint32_t view = trow / (uint32_t)gyv, by = trow - view * (uint32_t)gyv;
    const float X0 = (float)(bx * GSR_TILE), Y0 = (float)(by * GSR_TILE);
    const uint32_t px = bx * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = by * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    bool done = !inside;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    // classification words are kept for the backward: plane k (sub-tile k), word (range.x >> 5) + 16 tile + chunk
    uint32_t* const plane = planes + (size_t)lane * plane_words + (range.x >> 5) + 16u * tile + wid;
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C0 = 0.f, C1 = 0.f, C2 = 0.f;

    uint32_t id_next = 0;
    {
        const bool v0 = (int)threadIdx.x < total;
        const uint32_t id0 = v0 ? vals[range.x + threadIdx.x] : 0;
        stage_gather(stage[0], rec, id0, v0);
        const int p1 = BL_THREADS + threadIdx.x;
        if (p1 < total) id_next = vals[range.x + p1];
    }
    for (int i = 0; i < rounds; i++) {
        const int buf = i & 1;
        cp_async_wait<0>();       // my own entry of batch i has landed: classify it
        {
            uint32_t m = 0;
            if (i * BL_THREADS + (int)threadIdx.x < total) m = subtile_mask(stage[buf].q0[threadIdx.x], stage[buf].q1[threadIdx.x], X0, Y0);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t bk = __ballot_sync(0xffffffffu, (m >> k) & 1u);
                if (lane == k) { bits[buf][k][wid] = bk; plane[i * 8] = bk; }
            }
        }
        const int num_done = __syncthreads_count(done);
        if (num_done == BL_THREADS) break;
        {
            const int p1 = (i + 1) * BL_THREADS + threadIdx.x;
            stage_gather(stage[buf ^ 1], rec, id_next, p1 < total);
            const int p2 = p1 + BL_THREADS;
            if (p2 < total) id_next = vals[range.x + p2];
        }
        if (__all_sync(0xffffffffu, done)) continue;
        const BlendStage& st = stage[buf];
        const uint32_t pos_base = (uint32_t)(i * BL_THREADS) + 1u;
#pragma unroll 1
        for (int cch = 0; cch < 8; cch++) {
            uint32_t m = bits[buf][wid][cch];
            if (m == 0) continue;
            const int jb = cch * 32;
            do {
                const int n = __popc(m);
                if (UMAX >= 4 && n >= 4) blend_trip<4>(st, m, jb, pos_base, pxf, pyf, done, T, C0, C1, C2, last_contributor);
                else if (UMAX >= 3 && n == 3) blend_trip<3>(st, m, jb, pos_base, pxf, pyf, done, T, C0, C1, C2, last_contributor);
                else if (UMAX >= 2 && n >= 2) blend_trip<2>(st, m, jb, pos_base, pxf, pyf, done, T, C0, C1, C2, last_contributor);
                else blend_trip<1>(st, m, jb, pos_base, pxf, pyf, done, T, C0, C1, C2, last_contributor);
            } while (m);
            if (__all_sync(0xffffffffu, done)) break;
        }
    }
    cp_async_wait<0>();
    if (inside) {
        const size_t HW = (size_t)H * W;
        const size_t pix = (size_t)W * py + px;
        final_T[view * HW + pix] = T;
        n_contrib[view * HW + pix] = last_contributor;
        float* o = out + 3 * view * HW + pix;
        o[0] = C0 + T * bg[0];
        o[HW] = C1 + T * bg[1];
        o[2 * HW] = C2 + T * bg[2];
    }
}

// ------------------------------------------------------------------------------------------ statistics (debug hook)
// Counts what the culled forward does, to size the kernels' work (scripts/exp_blend.py): not on the product path.
//  0 entries staged (until the tile is done)   1 (warp, entry) pairs evaluated   2 lanes not done at evaluation
//  3 lanes with power <= 0   4 lanes with alpha >= 1/255   5 lanes blended   6 (warp, entry) pairs with >= 1 lane passing alpha
//  7 (half-warp 8x2, entry) pairs with >= 1 lane passing alpha   8 (quarter 4x2.. lanes 8-groups) likewise   9 batches
__device__ unsigned long long g_blend_stats[16];

__global__ void __launch_bounds__(BL_THREADS)
blend_stats_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec, int W, int H)
{
    __shared__ BlendStage stage;
    __shared__ uint32_t bits[8][8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const float X0 = (float)(blockIdx.x * GSR_TILE), Y0 = (float)(blockIdx.y * GSR_TILE);
    const uint32_t px = blockIdx.x * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = blockIdx.y * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    bool done = !inside;
    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;
    float T = 1.0f;
    unsigned long long c[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    __shared__ uint32_t wcnt[8];
    uint32_t tile_cnt = 0;
    for (int i = 0; i < rounds; i++) {
        uint32_t n44[2] = {0, 0}, n82[2] = {0, 0}, n42[4] = {0, 0, 0, 0}, nw = 0;
        const int num_done = __syncthreads_count(done);
        if (num_done == BL_THREADS) break;
        const int p = i * BL_THREADS + (int)threadIdx.x;
        uint32_t m = 0;
        if (p < total) {
            const SplatRec r = rec[vals[range.x + p]];
            stage.q0[threadIdx.x] = r.q0; stage.q1[threadIdx.x] = r.q1; stage.b[threadIdx.x] = r.q2.x;
            m = subtile_mask(r.q0, r.q1, X0, Y0);
            c[0]++;
        }
        if (threadIdx.x == 0) c[9]++;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t bk = __ballot_sync(0xffffffffu, (m >> k) & 1u);
            if (lane == k) bits[k][wid] = bk;
        }
        __syncthreads();
        for (int cch = 0; cch < 8; cch++) {
            uint32_t mm = bits[wid][cch];
            if (__all_sync(0xffffffffu, done)) mm = 0;
            while (mm) {
                const int j = cch * 32 + __ffs(mm) - 1;
                mm &= mm - 1;
                if (__all_sync(0xffffffffu, done)) break;
                if (lane == 0) c[1]++;
                nw++;
                bool pass = false;
                if (!done) {
                    c[2]++;
                    const float4 a = stage.q0[j];
                    const float4 q = stage.q1[j];
                    const float dx = a.x - pxf, dy = a.y - pyf;
                    const float power = -0.5f * (a.z * dx * dx + q.x * dy * dy) - a.w * dx * dy;
                    if (!(power > 0.0f)) {
                        c[3]++;
                        const float alpha = fminf(0.99f, q.y * expf(power));
                        if (!(alpha < 1.0f / 255.0f)) {
                            c[4]++;
                            pass = true;
                            const float test_T = T * (1 - alpha);
                            if (test_T < 0.0001f) done = true;
                            else { c[5]++; T = test_T; }
                        }
                    }
                }
                const uint32_t pb = __ballot_sync(0xffffffffu, pass);
                if (lane == 0) {
                    c[6] += pb != 0;
                    c[7] += ((pb & 0xFFFFu) != 0) + ((pb >> 16) != 0);
                    c[8] += ((pb & 0xFFu) != 0) + (((pb >> 8) & 0xFFu) != 0) + (((pb >> 16) & 0xFFu) != 0) + ((pb >> 24) != 0);
                }
                n44[0] += (pb & 0x0F0F0F0Fu) != 0; n44[1] += (pb & 0xF0F0F0F0u) != 0;
                n82[0] += (pb & 0x0000FFFFu) != 0; n82[1] += (pb & 0xFFFF0000u) != 0;
                n42[0] += (pb & 0x00000F0Fu) != 0; n42[1] += (pb & 0x0000F0F0u) != 0;
                n42[2] += (pb & 0x0F0F0000u) != 0; n42[3] += (pb & 0xF0F00000u) != 0;
            }
        }
        if (lane == 0) {
            c[10] += max(n44[0], n44[1]); c[11] += max(n82[0], n82[1]);
            c[12] += max(max(n42[0], n42[1]), max(n42[2], n42[3]));
            c[13] += n44[0] + n44[1];
            wcnt[wid] = nw;
        }
        tile_cnt += nw;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t mx = 0;
            for (int k = 0; k < 8; k++) mx = max(mx, wcnt[k]);
            c[14] += mx;                      // per-batch critical path (slowest warp)
        }
    }
    if (lane == 0) wcnt[wid] = tile_cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t mx = 0;
        for (int k = 0; k < 8; k++) mx = max(mx, wcnt[k]);
        c[15] += mx;                          // per-tile critical path if warps never waited for each other
    }
#pragma unroll
    for (int k = 0; k < 16; k++) {
        unsigned long long v = c[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&g_blend_stats[k], v);
    }
}

// classification words kept for the backward (in the binning blob's sort scratch, dead once the tile split is done):
// 8 planes (one per sub-tile of a tile) of (R >> 5) + 16 tiles + 16 words
size_t blend_plane_words(long long R, int W, int H)
{
    const size_t tiles = (size_t)((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE);
    return (size_t)((R > 0 ? R : 0) >> 5) + 16 * tiles + 16;
}

int launch_blend_forward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                         float* out_color, cudaStream_t s)
{
    const int gyv = (H + GSR_TILE - 1) / GSR_TILE;
    const dim3 grid((W + GSR_TILE - 1) / GSR_TILE, gyv * V);
    const uint32_t pw = (uint32_t)blend_plane_words(R, W, gyv * V * GSR_TILE);
    if (V > 1 && (env_int("GSR_BLEND", 0) != 0 || env_int("GSR_BLEND_STATS", 0) != 0)) return GSR_EINVAL;   // comparison arms are single-view
    const uint32_t* lpt = env_int("GSR_LPT", 1) ? im.tile_order : nullptr;
    if (env_int("GSR_BLEND_STATS", 0) == 1)
        GSR_LAUNCH(blend_stats_kernel, grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H);
    if (env_int("GSR_BLEND", 0) == 1)
        GSR_LAUNCH(blend_fwd_kernel, grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib, out_color);
    else if (env_int("GSR_FWD_MINB", 4) == 3)
        GSR_LAUNCH((blend_fwd_v3_kernel<4, 3>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    else
        GSR_LAUNCH((blend_fwd_v3_kernel<4, 4>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    return GSR_OK;
}

// ------------------------------------------------------------------------------------------ backward

// Per-pixel contribution of one blended entry to the per-Gaussian sums (restates backward.cu:494-556).
// Instead of the reference's nine final gradient terms the lanes produce MOMENTS of w = dL/dG * G:
//   g0..g2 = alpha*T*dL/dpix (colour), gmx = w dx, gmy = w dy, gcx = w dx^2, gcy = w dx dy, gcw = w dy^2, gop = w
// The conic / opacity factors are constant per Gaussian, so they are applied once per Gaussian in
// preprocess_bwd (dL/dmean2D.x = -W/2 (A gmx + B gmy), dL/dconic = -1/2 (gcx, gcy, gcw), dL/dopacity = gop / o).
// 1/(1-alpha) is one MUFU.RCP (alpha <= 0.99) instead of two IEEE divisions: gradients carry a 1e-3 tolerance.
__device__ __forceinline__ void bwd_live_math(float alpha, float araw, float dx, float dy, float c0, float c1, float c2,
                                              float dp0, float dp1, float dp2, float tfb, float& T, float& last_alpha,
                                              float& acc0, float& acc1, float& acc2, float& lc0, float& lc1, float& lc2,
                                              float& g0, float& g1, float& g2, float& gmx, float& gmy, float& gcx, float& gcy,
                                              float& gcw, float& gop)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f - alpha));
    T *= r;
    const float at = alpha * T;
    acc0 = fmaf(last_alpha, lc0 - acc0, acc0); lc0 = c0;
    acc1 = fmaf(last_alpha, lc1 - acc1, acc1); lc1 = c1;
    acc2 = fmaf(last_alpha, lc2 - acc2, acc2); lc2 = c2;
    float dL_dalpha = (c0 - acc0) * dp0 + (c1 - acc1) * dp1 + (c2 - acc2) * dp2;
    g0 = at * dp0; g1 = at * dp1; g2 = at * dp2;
    last_alpha = alpha;
    dL_dalpha = fmaf(dL_dalpha, T, -tfb * r);
    const float w = araw * dL_dalpha;          // dL/dG * G = opacity * dL/dalpha * G  (no zeroing when alpha was clamped)
    gmx = w * dx; gmy = w * dy;
    gcx = gmx * dx; gcy = gmx * dy; gcw = gmy * dy;
    gop = w;
}

// reduces 8 per-lane values over the warp with a transposing butterfly: afterwards lane L holds the
// warp total of value ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1)  (replicated over the low two lane bits)
__device__ __forceinline__ float butterfly8(float v0, float v1, float v2, float v3, float v4, float v5, float v6, float v7,
                                            int lane)
{
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
    // offset 16: 8 -> 4
    float a0 = (b4 ? v4 : v0) + __shfl_xor_sync(0xffffffffu, b4 ? v0 : v4, 16);
    float a1 = (b4 ? v5 : v1) + __shfl_xor_sync(0xffffffffu, b4 ? v1 : v5, 16);
    float a2 = (b4 ? v6 : v2) + __shfl_xor_sync(0xffffffffu, b4 ? v2 : v6, 16);
    float a3 = (b4 ? v7 : v3) + __shfl_xor_sync(0xffffffffu, b4 ? v3 : v7, 16);
    // offset 8: 4 -> 2
    float c0 = (b3 ? a2 : a0) + __shfl_xor_sync(0xffffffffu, b3 ? a0 : a2, 8);
    float c1 = (b3 ? a3 : a1) + __shfl_xor_sync(0xffffffffu, b3 ? a1 : a3, 8);
    // offset 4: 2 -> 1
    float e = (b2 ? c1 : c0) + __shfl_xor_sync(0xffffffffu, b2 ? c0 : c1, 4);
    e += __shfl_xor_sync(0xffffffffu, e, 2);
    e += __shfl_xor_sync(0xffffffffu, e, 1);
    return e;
}

struct BlendStageBwd {
    float4 q0[BL_THREADS];
    float4 q1[BL_THREADS];
    float  b[BL_THREADS];
    uint32_t id[BL_THREADS];
};

// accumulator row layout (12 floats per Gaussian): 0-2 dL/dcolour, 3-4 sum w*(dx,dy), 5-7 sum w*(dx^2, dx dy, dy^2), 8 sum w
__global__ void __launch_bounds__(BL_THREADS)
blend_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                 int W, int H, const float* __restrict__ bg, const float* __restrict__ final_T,
                 const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix, float* __restrict__ accum)
{
    __shared__ BlendStageBwd stage[2];
    const int lane = threadIdx.x & 31;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const uint32_t px = blockIdx.x * GSR_TILE + tx, py = blockIdx.y * GSR_TILE + ty;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)W * py + px;
    const size_t HW = (size_t)H * W;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int total = (int)(range.y - range.x);
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;

    const float T_final = inside ? final_T[pix] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix] : 0;
    float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
    if (inside) { dp0 = dL_dpix[pix]; dp1 = dL_dpix[HW + pix]; dp2 = dL_dpix[2 * HW + pix]; }
    const float tfb = T_final * (bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2);
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f;       // accum_rec
    float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;

    // the warp only has work in list positions below the largest n_contrib of its lanes
    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));

    // which accumulator column this lane commits after the butterfly
    const int col8 = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);

    // entries are walked back to front: batch i covers list positions total-1-i*256 .. (descending)
    auto gather = [&](BlendStageBwd& st, int batch) {
        const int p = batch * BL_THREADS + threadIdx.x;
        if (p < total) {
            const uint32_t id = vals[range.y - 1 - p];
            st.id[threadIdx.x] = id;
            const SplatRec* r = rec + id;
            cp_async16(&st.q0[threadIdx.x], &r->q0);
            cp_async16(&st.q1[threadIdx.x], &r->q1);
            cp_async4(&st.b[threadIdx.x], &r->q2);
        }
        cp_async_commit();
    };
    gather(stage[0], 0);
    int contributor = total;
    for (int i = 0; i < rounds; i++) {
        cp_async_wait<0>();
        __syncthreads();
        if (i + 1 < rounds) gather(stage[(i + 1) & 1], i + 1);
        const BlendStageBwd& st = stage[i & 1];
        const int n = min(BL_THREADS, total - i * BL_THREADS);
        for (int j = 0; j < n; j++) {
            contributor--;
            if (contributor >= warp_last) continue;      // warp-uniform
            const float4 a = st.q0[j];
            const float4 c = st.q1[j];
            const float dx = a.x - pxf, dy = a.y - pyf;
            const float power = -0.5f * (a.z * dx * dx + c.x * dy * dy) - a.w * dx * dy;
            const float araw = c.y * expf(power);
            const float alpha = fminf(0.99f, araw);
            const bool live = contributor < last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(0xffffffffu, live)) continue;

            float g0 = 0.f, g1 = 0.f, g2 = 0.f, gmx = 0.f, gmy = 0.f, gcx = 0.f, gcy = 0.f, gcw = 0.f, gop = 0.f;
            if (live) {
                bwd_live_math(alpha, araw, dx, dy, c.z, c.w, st.b[j], dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1, acc2,
                              lc0, lc1, lc2, g0, g1, g2, gmx, gmy, gcx, gcy, gcw, gop);
            }
            const float r8 = butterfly8(g0, g1, g2, gmx, gmy, gcx, gcy, gcw, lane);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) gop += __shfl_xor_sync(0xffffffffu, gop, o);
            float* row = accum + (size_t)st.id[j] * 12;
            if ((lane & 3) == 0) atomicAdd(row + col8, r8);
            else if (lane == 1) atomicAdd(row + 8, gop);
        }
    }
}

// ------------------------------------------------------------------------------------------ backward with sub-tile culling
// Same tiles / sub-tiles / culling as the forward; the list is truncated to the largest n_contrib of the tile, a warp
// additionally ignores entries at or beyond its own largest n_contrib.
// Two surviving entries per trip: both alphas are evaluated side by side (instruction-level parallelism on the
// LDS -> power -> expf chain), the per-pixel state update stays sequential, and the two entries' 8 + 8 values
// go through ONE 16-value transposing butterfly (16 + 5 shuffles for two entries instead of 2 x 14).
struct PixState {
    float T, last_alpha, acc0, acc1, acc2, lc0, lc1, lc2;
};
struct EntryHead {
    float dx, dy, araw, alpha;
    bool live;
};
__device__ __forceinline__ EntryHead bwd_head(const float4 a, const float4 c, float pxf, float pyf, int pos, int last_contributor)
{
    EntryHead h;
    h.dx = a.x - pxf; h.dy = a.y - pyf;
    const float power = -0.5f * (a.z * h.dx * h.dx + c.x * h.dy * h.dy) - a.w * h.dx * h.dy;
    h.araw = c.y * expf(power);
    h.alpha = fminf(0.99f, h.araw);
    h.live = pos < last_contributor && !(power > 0.0f) && !(h.alpha < 1.0f / 255.0f);
    return h;
}

// lane L ends up with the warp total of v[((L>>4)&1)*8 + ((L>>3)&1)*4 + ((L>>2)&1)*2 + ((L>>1)&1)]
__device__ __forceinline__ float butterfly16(const float* v, int lane)
{
    const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
    float a[8], c[4], d[2];
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = (b4 ? v[k + 8] : v[k]) + __shfl_xor_sync(0xffffffffu, b4 ? v[k] : v[k + 8], 16);
#pragma unroll
    for (int k = 0; k < 4; k++) c[k] = (b3 ? a[k + 4] : a[k]) + __shfl_xor_sync(0xffffffffu, b3 ? a[k] : a[k + 4], 8);
#pragma unroll
    for (int k = 0; k < 2; k++) d[k] = (b2 ? c[k + 2] : c[k]) + __shfl_xor_sync(0xffffffffu, b2 ? c[k] : c[k + 2], 4);
    float e = (b1 ? d[1] : d[0]) + __shfl_xor_sync(0xffffffffu, b1 ? d[0] : d[1], 2);
    e += __shfl_xor_sync(0xffffffffu, e, 1);
    return e;
}

template <int MINB>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_bwd_v3_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    int W, int H, const float* __restrict__ bg, const float* __restrict__ final_T,
                    const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix, float* __restrict__ accum)
{
    __shared__ BlendStageBwd stage[2];
    __shared__ uint32_t bits[2][8][8];
    __shared__ int s_warp_last[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const float X0 = (float)(blockIdx.x * GSR_TILE), Y0 = (float)(blockIdx.y * GSR_TILE);
    const uint32_t px = blockIdx.x * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = blockIdx.y * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)W * py + px;
    const size_t HW = (size_t)H * W;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

    const float T_final = inside ? final_T[pix] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix] : 0;
    float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
    if (inside) { dp0 = dL_dpix[pix]; dp1 = dL_dpix[HW + pix]; dp2 = dL_dpix[2 * HW + pix]; }
    const float tfb = T_final * (bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2);
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f;
    float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;

    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    if (lane == 0) s_warp_last[wid] = warp_last;
    __syncthreads();
    int total = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) total = max(total, s_warp_last[k]);
    total = min(total, (int)(range.y - range.x));
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;
    const uint32_t list_end = range.x + total;

    const int col8 = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
    const int col16 = (lane >> 1) & 7, ent16 = lane >> 4;

    auto gather = [&](BlendStageBwd& st, int batch) {
        const int p = batch * BL_THREADS + threadIdx.x;
        if (p < total) {
            const uint32_t id = vals[list_end - 1 - p];
            st.id[threadIdx.x] = id;
            const SplatRec* r = rec + id;
            cp_async16(&st.q0[threadIdx.x], &r->q0);
            cp_async16(&st.q1[threadIdx.x], &r->q1);
            cp_async4(&st.b[threadIdx.x], &r->q2);
        }
        cp_async_commit();
    };
    gather(stage[0], 0);
    for (int i = 0; i < rounds; i++) {
        const int buf = i & 1;
        cp_async_wait<0>();
        {
            uint32_t m = 0;
            const int p = i * BL_THREADS + (int)threadIdx.x;
            if (p < total) {
                m = subtile_mask(stage[buf].q0[threadIdx.x], stage[buf].q1[threadIdx.x], X0, Y0);
                const int pos = total - 1 - p;
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (pos >= s_warp_last[k]) m &= ~(1u << k);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t bk = __ballot_sync(0xffffffffu, (m >> k) & 1u);
                if (lane == k) bits[buf][k][wid] = bk;
            }
        }
        __syncthreads();
        if (i + 1 < rounds) gather(stage[buf ^ 1], i + 1);
        const BlendStageBwd& st = stage[buf];
        const int pos_top = total - 1 - i * BL_THREADS;
#pragma unroll 1
        for (int cch = 0; cch < 8; cch++) {
            uint32_t m = bits[buf][wid][cch];
            const int jb = cch * 32;
            while (m) {
                const int j0 = jb + __ffs(m) - 1;
                m &= m - 1;
                if (m) {
                    const int j1 = jb + __ffs(m) - 1;
                    m &= m - 1;
                    const float4 a0 = st.q0[j0], c0 = st.q1[j0], a1 = st.q0[j1], c1 = st.q1[j1];
                    const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, pos_top - j0, last_contributor);
                    const EntryHead h1 = bwd_head(a1, c1, pxf, pyf, pos_top - j1, last_contributor);
                    const bool any0 = __any_sync(0xffffffffu, h0.live), any1 = __any_sync(0xffffffffu, h1.live);
                    if (!any0 && !any1) continue;
                    float v[16], gop0 = 0.f, gop1 = 0.f;
#pragma unroll
                    for (int k = 0; k < 16; k++) v[k] = 0.f;
                    if (h0.live)
                        bwd_live_math(h0.alpha, h0.araw, h0.dx, h0.dy, c0.z, c0.w, st.b[j0], dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1,
                                      acc2, lc0, lc1, lc2, v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], gop0);
                    if (h1.live)
                        bwd_live_math(h1.alpha, h1.araw, h1.dx, h1.dy, c1.z, c1.w, st.b[j1], dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1,
                                      acc2, lc0, lc1, lc2, v[8], v[9], v[10], v[11], v[12], v[13], v[14], v[15], gop1);
                    const float e = butterfly16(v, lane);
                    float g = (ent16 ? gop1 : gop0) + __shfl_xor_sync(0xffffffffu, ent16 ? gop0 : gop1, 16);
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) g += __shfl_xor_sync(0xffffffffu, g, o);
                    const bool mine = ent16 ? any1 : any0;
                    float* row = accum + (size_t)st.id[ent16 ? j1 : j0] * 12;
                    if (mine) {
                        if ((lane & 1) == 0) atomicAdd(row + col16, e);
                        else if ((lane & 15) == 1) atomicAdd(row + 8, g);
                    }
                } else {
                    const float4 a0 = st.q0[j0], c0 = st.q1[j0];
                    const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, pos_top - j0, last_contributor);
                    if (!__any_sync(0xffffffffu, h0.live)) continue;
                    float g0 = 0.f, g1 = 0.f, g2 = 0.f, gmx = 0.f, gmy = 0.f, gcx = 0.f, gcy = 0.f, gcw = 0.f, gop = 0.f;
                    if (h0.live)
                        bwd_live_math(h0.alpha, h0.araw, h0.dx, h0.dy, c0.z, c0.w, st.b[j0], dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1,
                                      acc2, lc0, lc1, lc2, g0, g1, g2, gmx, gmy, gcx, gcy, gcw, gop);
                    const float r8 = butterfly8(g0, g1, g2, gmx, gmy, gcx, gcy, gcw, lane);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) gop += __shfl_xor_sync(0xffffffffu, gop, o);
                    float* row = accum + (size_t)st.id[j0] * 12;
                    if ((lane & 3) == 0) atomicAdd(row + col8, r8);
                    else if (lane == 1) atomicAdd(row + 8, gop);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------ backward, tensor-core reduction
// The per-Gaussian sums of blend_bwd_v3 cost a 16-value butterfly (21 shuffles, 30 selects, 21 adds) per pair of
// entries: 40 % of the kernel's instructions.  They are linear in two per-lane scalars of the entry,
//   w  = dL/dG * G   and   at = alpha * T,
// with coefficients that do not depend on the entry: the pixel's offset inside the warp's 8x4 sub-tile
// (lx, ly in 0..7 x 0..3) for the six moments of w, and the pixel's dL/dpixel for the three colour sums.  So a warp
// parks (w, at) of each surviving entry in a shared-memory slot (two STS) and, every 8 entries, reduces all of them
// over the 32 lanes with the tensor cores (mma.sync m16n8k8 tf32, fp32 accumulate, K = the 32 lanes in 4 steps):
//   A[16 x 32]  rows 0-7 = w of slots 0-7, rows 8-15 = at of slots 0-7          (split hi + lo: two MMAs)
//   B1[32 x 8]  columns (1, lx, lx', ...) = the moment matrix (small integers: exact in tf32)   -> D1 rows 0-7
//   B2[32 x 8]  columns (hi, lo) of dL/dpixel r, g, b                                           -> D2 rows 8-15
// The hi/lo splits keep ~21 mantissa bits (error < 1e-6; the gradients carry a 1e-3 tolerance).  The slot array is
// stored in A-fragment order, so one LDS.128 per k-step is the whole A operand.  The moments about the sub-tile origin
// are shifted to the Gaussian's centre (dx = u - lx, u = mean.x - origin.x) and committed with the same RED.ADDs
// into the same accumulator rows as before (two of them as 8-byte vector REDs).
constexpr int RED_SLOTS = 8;
constexpr int RED_KSTRIDE = 144;      // floats per k-step of the slot array: 8 slots x 16 + 16 (STS of one slot: 2-way conflicts only)
struct __align__(16) WarpRed {
    float a[4 * RED_KSTRIDE];         // [k-step s][slot][t][b][w | at]   (lane = 8t + 2s + b)
    float2 b2[4][32];                 // B2 fragments: [k-step][lane (g, t)] = column g of dL/dpixel at lanes 8t+2s, 8t+2s+1
    float4 meta[RED_SLOTS];           // mean.x, mean.y, Gaussian id (bits)
};
struct BlendStageBwd4 {
    float4 q0[BL_THREADS];
    float4 q1[BL_THREADS];
    float2 bi[BL_THREADS];            // colour b, Gaussian id (bits): one LDS.64 at the top of a trip
};
struct __align__(16) BwdSmemV4 {
    BlendStageBwd4 stage[2];
    WarpRed red[8];
    float2 b1[4][32];                 // B1 fragments (moment matrix), same for all warps
    uint32_t bits[2][8][8];
    int warp_last[8];
};

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float fsel(bool c, float a, float b)       // branch-free select (the compiler may not turn it into a jump)
{
    float r;
    asm("{ .reg .pred p; setp.ne.s32 p, %3, 0; selp.f32 %0, %1, %2, p; }" : "=f"(r) : "f"(a), "f"(b), "r"((int)c));
    return r;
}
__device__ __forceinline__ void red_add_v2(float* addr, float x, float y)
{
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(x), "f"(y) : "memory");
}

// Reduces slots [0, cnt) of the warp over its 32 lanes and commits them.  Fragment layout of m16n8k8 (g = lane >> 2,
// t = lane & 3): A rows g / g+8 at k = t / t+4, B column g at k = t / t+4, D rows g / g+8, columns 2t, 2t+1.
// k-step s maps k = t to lane 8t + 2s and k = t+4 to lane 8t + 2s + 1.
__device__ __forceinline__ void red_flush(WarpRed& R, const float2 (*b1)[32], int cnt, int lane, float X0s, float Y0s,
                                          float* __restrict__ accum)
{
    __syncwarp();
    const int g = lane >> 2, t = lane & 3;
    float d1[4] = {0.f, 0.f, 0.f, 0.f}, d2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int s = 0; s < 4; s++) {
        const float4 av = *reinterpret_cast<const float4*>(&R.a[s * RED_KSTRIDE + 4 * lane]);
        const float2 m = b1[s][lane];
        const float2 q = R.b2[s][lane];
        const float x[4] = {av.x, av.y, av.z, av.w};
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {      // x = hi + lo, hi exact in tf32 (the tensor core ignores the 13 low mantissa bits)
            hi[k] = __float_as_uint(x[k]) & 0xffffe000u;
            lo[k] = __float_as_uint(x[k] - __uint_as_float(hi[k]));
        }
        mma_tf32(d1, hi, __float_as_uint(m.x), __float_as_uint(m.y));
        mma_tf32(d1, lo, __float_as_uint(m.x), __float_as_uint(m.y));
        mma_tf32(d2, hi, __float_as_uint(q.x), __float_as_uint(q.y));
        mma_tf32(d2, lo, __float_as_uint(q.x), __float_as_uint(q.y));
    }
    // lane (g, t): d1[0], d1[1] = moments 2t, 2t+1 of slot g about the sub-tile origin (S0 Sx | Sy Sxx | Sxy Syy | - -),
    //              d2[2] + d2[3] = colour sum t of slot g
    const float4 me = R.meta[g];
    const int q0 = lane & ~3;
    const float S0 = __shfl_sync(0xffffffffu, d1[0], q0);
    const float Sx = __shfl_sync(0xffffffffu, d1[1], q0);
    const float Sy = __shfl_sync(0xffffffffu, d1[0], q0 + 1);
    const float u = me.x - X0s, v = me.y - Y0s;          // dx = u - lx, dy = v - ly
    const float a1 = fmaf(u, S0, -Sx);                   // sum w dx
    const float a2 = fmaf(v, S0, -Sy);                   // sum w dy
    const float o3 = fmaf(u, a1 - Sx, d1[1]);            // sum w dx^2  = Sxx - 2u Sx + u^2 S0
    const float o4 = fmaf(-v, Sx, fmaf(u, a2, d1[0]));   // sum w dx dy = Sxy - u Sy - v Sx + uv S0
    const float o5 = fmaf(v, a2 - Sy, d1[1]);            // sum w dy^2  = Syy - 2v Sy + v^2 S0
    const float out0 = fsel(t == 0, d1[0], fsel(t == 1, a2, o4));
    const float out1 = fsel(t == 0, a1, fsel(t == 1, o3, o5));
    const float col = d2[2] + d2[3];
    if (g < cnt && t < 3) {
        float* row = accum + (size_t)__float_as_uint(me.z) * 12;
        if (t == 0) { atomicAdd(row + 8, out0); atomicAdd(row + 3, out1); }
        else red_add_v2(row + 2 + 2 * t, out0, out1);    // t = 1: columns 4, 5   t = 2: columns 6, 7
        atomicAdd(row + t, col);
    }
    __syncwarp();
}

// per-pixel state update of one blended entry (restates backward.cu:494-556); returns w = dL/dG * G and at = alpha * T
__device__ __forceinline__ void bwd_live_wa(float alpha, float araw, float c0, float c1, float c2, float dp0, float dp1, float dp2,
                                            float tfb, float& T, float& last_alpha, float& acc0, float& acc1, float& acc2,
                                            float& lc0, float& lc1, float& lc2, float& w, float& at)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f - alpha));
    T *= r;
    at = alpha * T;
    acc0 = fmaf(last_alpha, lc0 - acc0, acc0); lc0 = c0;
    acc1 = fmaf(last_alpha, lc1 - acc1, acc1); lc1 = c1;
    acc2 = fmaf(last_alpha, lc2 - acc2, acc2); lc2 = c2;
    float dL_dalpha = (c0 - acc0) * dp0 + (c1 - acc1) * dp1 + (c2 - acc2) * dp2;
    last_alpha = alpha;
    dL_dalpha = fmaf(dL_dalpha, T, -tfb * r);
    w = araw * dL_dalpha;
}

template <int MINB>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_bwd_v4_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    int W, int H, const float* __restrict__ bg, const float* __restrict__ final_T,
                    const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix, float* __restrict__ accum)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmemV4& S = *reinterpret_cast<BwdSmemV4*>(smem_raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpRed& R = S.red[wid];
    const float X0 = (float)(blockIdx.x * GSR_TILE), Y0 = (float)(blockIdx.y * GSR_TILE);
    const float X0s = X0 + (float)((wid & 1) * 8), Y0s = Y0 + (float)((wid >> 1) * 4);
    const uint32_t px = blockIdx.x * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = blockIdx.y * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)W * py + px;
    const size_t HW = (size_t)H * W;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

    const float T_final = inside ? final_T[pix] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix] : 0;
    float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
    if (inside) { dp0 = dL_dpix[pix]; dp1 = dL_dpix[HW + pix]; dp2 = dL_dpix[2 * HW + pix]; }
    const float tfb = T_final * (bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2);
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f;
    float last_alpha = 0.f, lc0 = 0.f, lc1 = 0.f, lc2 = 0.f;

    // constant operands of the reduction in B-fragment order: the moment matrix (whole block) and this warp's
    // dL/dpixel columns (r_hi, r_lo, g_hi, g_lo, b_hi, b_lo, 0, 0)
    if (threadIdx.x < 128) {
        const int s = threadIdx.x >> 5, l = threadIdx.x & 31, n = l >> 2;
        const float ly = (float)(l & 3), lx0 = (float)(2 * s), lx1 = lx0 + 1.f;
        float2 m;
        m.x = n == 0 ? 1.f : n == 1 ? lx0 : n == 2 ? ly : n == 3 ? lx0 * lx0 : n == 4 ? lx0 * ly : n == 5 ? ly * ly : 0.f;
        m.y = n == 0 ? 1.f : n == 1 ? lx1 : n == 2 ? ly : n == 3 ? lx1 * lx1 : n == 4 ? lx1 * ly : n == 5 ? ly * ly : 0.f;
        S.b1[s][l] = m;
    }
    const int tL = lane >> 3, sL = (lane >> 1) & 3, bL = lane & 1;      // this pixel is k index tL + 4 bL of k-step sL
    {
        const float d[3] = {dp0, dp1, dp2};
        float* col = reinterpret_cast<float*>(&R.b2[sL][tL]) + bL;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float hi = __uint_as_float(__float_as_uint(d[c]) & 0xffffe000u);
            col[(2 * c) * 8] = hi;                 // column n lives at lane index 4n + tL: 8 floats apart
            col[(2 * c + 1) * 8] = d[c] - hi;
        }
        col[6 * 8] = 0.f;
        col[7 * 8] = 0.f;
    }
    float2* const slot0 = reinterpret_cast<float2*>(&R.a[sL * RED_KSTRIDE + tL * 4 + bL * 2]);   // slot c: slot0 + 8 c

    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    if (lane == 0) S.warp_last[wid] = warp_last;
    __syncthreads();
    int total = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) total = max(total, S.warp_last[k]);
    total = min(total, (int)(range.y - range.x));
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;
    const uint32_t list_end = range.x + total;

    auto gather = [&](BlendStageBwd4& st, int batch) {
        const int p = batch * BL_THREADS + threadIdx.x;
        if (p < total) {
            const uint32_t id = vals[list_end - 1 - p];
            st.bi[threadIdx.x].y = __uint_as_float(id);
            const SplatRec* r = rec + id;
            cp_async16(&st.q0[threadIdx.x], &r->q0);
            cp_async16(&st.q1[threadIdx.x], &r->q1);
            cp_async4(&st.bi[threadIdx.x].x, &r->q2);
        }
        cp_async_commit();
    };
    gather(S.stage[0], 0);
    int cnt = 0;                  // filled reduction slots of this warp (warp-uniform)
    for (int i = 0; i < rounds; i++) {
        const int buf = i & 1;
        cp_async_wait<0>();
        {
            uint32_t m = 0;
            const int p = i * BL_THREADS + (int)threadIdx.x;
            if (p < total) {
                m = subtile_mask(S.stage[buf].q0[threadIdx.x], S.stage[buf].q1[threadIdx.x], X0, Y0);
                const int pos = total - 1 - p;
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (pos >= S.warp_last[k]) m &= ~(1u << k);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t bk = __ballot_sync(0xffffffffu, (m >> k) & 1u);
                if (lane == k) S.bits[buf][k][wid] = bk;
            }
        }
        __syncthreads();
        if (i + 1 < rounds) gather(S.stage[buf ^ 1], i + 1);
        const BlendStageBwd4& st = S.stage[buf];
        const int pos_top = total - 1 - i * BL_THREADS;
#pragma unroll 1
        for (int cch = 0; cch < 8; cch++) {
            uint32_t m = S.bits[buf][wid][cch];
            const int jb = cch * 32;
            while (m) {
                if (cnt >= RED_SLOTS - 1) { red_flush(R, S.b1, cnt, lane, X0s, Y0s, accum); cnt = 0; }
                const int j0 = jb + __ffs(m) - 1;
                m &= m - 1;
                if (m) {
                    const int j1 = jb + __ffs(m) - 1;
                    m &= m - 1;
                    const float4 a0 = st.q0[j0], c0 = st.q1[j0], a1 = st.q0[j1], c1 = st.q1[j1];
                    const float2 e0 = st.bi[j0], e1 = st.bi[j1];
                    const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, pos_top - j0, last_contributor);
                    const EntryHead h1 = bwd_head(a1, c1, pxf, pyf, pos_top - j1, last_contributor);
                    const bool any0 = __any_sync(0xffffffffu, h0.live), any1 = __any_sync(0xffffffffu, h1.live);
                    if (any0) {
                        float wv = 0.f, atv = 0.f;
                        if (h0.live)
                            bwd_live_wa(h0.alpha, h0.araw, c0.z, c0.w, e0.x, dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1, acc2,
                                        lc0, lc1, lc2, wv, atv);
                        slot0[8 * cnt] = make_float2(wv, atv);
                        if (lane == 0) R.meta[cnt] = make_float4(a0.x, a0.y, e0.y, 0.f);
                        cnt++;
                    }
                    if (any1) {
                        float wv = 0.f, atv = 0.f;
                        if (h1.live)
                            bwd_live_wa(h1.alpha, h1.araw, c1.z, c1.w, e1.x, dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1, acc2,
                                        lc0, lc1, lc2, wv, atv);
                        slot0[8 * cnt] = make_float2(wv, atv);
                        if (lane == 0) R.meta[cnt] = make_float4(a1.x, a1.y, e1.y, 0.f);
                        cnt++;
                    }
                } else {
                    const float4 a0 = st.q0[j0], c0 = st.q1[j0];
                    const float2 e0 = st.bi[j0];
                    const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, pos_top - j0, last_contributor);
                    if (!__any_sync(0xffffffffu, h0.live)) continue;
                    float wv = 0.f, atv = 0.f;
                    if (h0.live)
                        bwd_live_wa(h0.alpha, h0.araw, c0.z, c0.w, e0.x, dp0, dp1, dp2, tfb, T, last_alpha, acc0, acc1, acc2,
                                    lc0, lc1, lc2, wv, atv);
                    slot0[8 * cnt] = make_float2(wv, atv);
                    if (lane == 0) R.meta[cnt] = make_float4(a0.x, a0.y, e0.y, 0.f);
                    cnt++;
                }
            }
        }
    }
    if (cnt) red_flush(R, S.b1, cnt, lane, X0s, Y0s, accum);
}

// ------------------------------------------------------------------------------------------ backward, warp-autonomous
// blend_bwd_v4 spends 19 % of its warp-time at the per-batch block barrier (the eight sub-tiles of a tile do not carry
// the same number of surviving entries) and re-runs the sub-tile classification the forward already did.  Here the
// forward exports its classification - one word per (sub-tile, 32 list entries), `planes` - and every warp walks the
// tile list on its own, back to front, 32 entries per step: lanes whose bit is set gather their record with cp.async
// into a warp-private ring (so a record is fetched once per sub-tile that needs it, as before), NS - 1 steps ahead of the
// blending; there is no block barrier after the prologue and no classification.  Reduction (red_flush) as in v4.
// exp stays the full-precision expf of the forward: with ex2.approx (2 instructions instead of 8) a pixel's alpha can
// land on the other side of the 1/255 skip threshold than it did in the forward (measured: about one pixel per 3 M pairs),
// which puts a 0.4 % error on that pixel's transmittance chain - gradient errors of 3e-4 instead of 1e-6.
// Per-pixel state update of one blended entry in suffix-sum form.  With e = dL/dpixel . colour and
// Sd = sum over the entries behind this one of e_k alpha_k T_k (+ T_final dL/dpixel . background),
//   dL/dalpha = T e - Sd / (1 - alpha)
// which is the reference's (colour - accum_rec) . dL/dpixel * T - T_final/(1 - alpha) bg . dL/dpixel (backward.cu:494-533)
// with accum_rec written out as Sd / T_next: 11 operations and two state registers (T, Sd) instead of 23 and eight.
__device__ __forceinline__ void bwd_live_sd(float alpha, float araw, float c0, float c1, float c2, float dp0, float dp1, float dp2,
                                            float& T, float& Sd, float& w, float& at)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f - alpha));
    T *= r;
    const float e = fmaf(c2, dp2, fmaf(c1, dp1, c0 * dp0));
    at = alpha * T;
    const float dL_dalpha = fmaf(T, e, -(Sd * r));
    Sd = fmaf(e, at, Sd);
    w = araw * dL_dalpha;
}

template <int NS>
struct __align__(16) WarpRing {
    float4 q0[NS][32];
    float4 q1[NS][32];
    float2 bi[NS][32];                // colour b, Gaussian id (bits)
};
template <int NS>
struct __align__(16) BwdSmemV5 {
    WarpRing<NS> ring[8];
    WarpRed red[8];
    float2 b1[4][32];
};

template <int NS, int MINB>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_bwd_v5_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    const uint32_t* __restrict__ planes, uint32_t plane_words, int W, int H, const float* __restrict__ bg,
                    const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                    float* __restrict__ accum, int gyv, const uint32_t* __restrict__ tile_order)
{
    constexpr int D = NS - 1;          // gather distance in steps
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmemV5<NS>& S = *reinterpret_cast<BwdSmemV5<NS>*>(smem_raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpRed& R = S.red[wid];
    WarpRing<NS>& Q = S.ring[wid];
    const uint32_t lin = blockIdx.y * gridDim.x + blockIdx.x;                                  // see blend_fwd_v3_kernel
    const uint32_t tile = tile_order ? tile_order[lin] : lin;
    const uint32_t trow = tile / gridDim.x, bx = tile - trow * gridDim.x;
    const uint32_t view = trow / (uint32_t)gyv, by = trow - view * (uint32_t)gyv;
    const float X0s = (float)(bx * GSR_TILE + (wid & 1) * 8), Y0s = (float)(by * GSR_TILE + (wid >> 1) * 4);
    const uint32_t px = bx * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = by * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t HW = (size_t)H * W;
    const size_t pix = view * HW + (size_t)W * py + px;
    dL_dpix += 2 * view * HW;                   // [V][3][H][W]: channel c of this pixel is dL_dpix[c * HW + pix]
    const uint2 range = ranges[tile];

    const float T_final = inside ? final_T[pix] : 0.f;
    float T = T_final;
    const int last_contributor = inside ? (int)n_contrib[pix] : 0;
    float dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
    if (inside) { dp0 = dL_dpix[pix]; dp1 = dL_dpix[HW + pix]; dp2 = dL_dpix[2 * HW + pix]; }
    float Sd = T_final * (bg[0] * dp0 + bg[1] * dp1 + bg[2] * dp2);     // see bwd_live_sd

    // constant operands of the reduction (see blend_bwd_v4_kernel)
    if (threadIdx.x < 128) {
        const int s = threadIdx.x >> 5, l = threadIdx.x & 31, n = l >> 2;
        const float ly = (float)(l & 3), lx0 = (float)(2 * s), lx1 = lx0 + 1.f;
        float2 m;
        m.x = n == 0 ? 1.f : n == 1 ? lx0 : n == 2 ? ly : n == 3 ? lx0 * lx0 : n == 4 ? lx0 * ly : n == 5 ? ly * ly : 0.f;
        m.y = n == 0 ? 1.f : n == 1 ? lx1 : n == 2 ? ly : n == 3 ? lx1 * lx1 : n == 4 ? lx1 * ly : n == 5 ? ly * ly : 0.f;
        S.b1[s][l] = m;
    }
    const int tL = lane >> 3, sL = (lane >> 1) & 3, bL = lane & 1;
    {
        const float d[3] = {dp0, dp1, dp2};
        float* col = reinterpret_cast<float*>(&R.b2[sL][tL]) + bL;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float hi = __uint_as_float(__float_as_uint(d[c]) & 0xffffe000u);
            col[(2 * c) * 8] = hi;
            col[(2 * c + 1) * 8] = d[c] - hi;
        }
        col[6 * 8] = 0.f;
        col[7 * 8] = 0.f;
    }
    float2* const slot0 = reinterpret_cast<float2*>(&R.a[sL * RED_KSTRIDE + tL * 4 + bL * 2]);

    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    warp_last = min(warp_last, (int)(range.y - range.x));
    __syncthreads();                       // b1 is visible; the only block barrier
    const int nch = (warp_last + 31) >> 5; // chunks of 32 list entries this warp has work in
    if (nch == 0) return;
    const uint32_t* plane = planes + (size_t)wid * plane_words + (range.x >> 5) + 16u * tile;
    const uint32_t* tvals = vals + range.x;
    // step s = 0 .. nch-1 handles chunk nch-1-s (back to front).  The classification words of 32 consecutive steps sit
    // in one register per lane: wA = block of the step whose mask is extracted next, wB = the block after it.
    auto load_words = [&](int blk) -> uint32_t {
        const int s = blk * 32 + lane;
        return s < nch ? plane[nch - 1 - s] : 0u;
    };
    uint32_t wA = load_words(0), wB = load_words(1);
    int blkA = 0;
    auto load_id = [&](int s, uint32_t m) -> uint32_t { return ((m >> lane) & 1u) ? tvals[(nch - 1 - s) * 32 + lane] : 0u; };
    auto issue = [&](int s, uint32_t m, uint32_t id) {
        if ((m >> lane) & 1u) {
            const int k = s % NS;
            const SplatRec* r = rec + id;
            cp_async16(&Q.q0[k][lane], &r->q0);
            cp_async16(&Q.q1[k][lane], &r->q1);
            cp_async4(&Q.bi[k][lane].x, &r->q2);
            Q.bi[k][lane].y = __uint_as_float(id);
        }
        cp_async_commit();
    };
    uint32_t mq[D + 2];                // masks of steps s .. s+D+1
#pragma unroll
    for (int k = 0; k < D + 2; k++) {
        const uint32_t w = __shfl_sync(0xffffffffu, wA, k);
        mq[k] = k < nch ? w : 0u;
    }
    uint32_t idc = load_id(0, mq[0]);
#pragma unroll
    for (int k = 0; k < D; k++) {
        const uint32_t idn = load_id(k + 1, mq[k + 1]);
        issue(k, mq[k], idc);
        idc = idn;
    }
    int cnt = 0;                       // filled reduction slots (warp-uniform)
    for (int s = 0; s < nch; s++) {
        const uint32_t idn = load_id(s + D + 1, mq[D + 1]);     // ids of step s+D+1 (mask is 0 beyond the end)
        issue(s + D, mq[D], idc);                               // records of step s+D
        idc = idn;
        cp_async_wait<D>();                                     // my records of step s have landed
        __syncwarp();
        const int kq = s % NS;
        const int cbase = (nch - 1 - s) * 32;
        uint32_t m = mq[0];
        while (m) {
            if (cnt >= RED_SLOTS - 1) { red_flush(R, S.b1, cnt, lane, X0s, Y0s, accum); cnt = 0; }
            const int j0 = 31 - __clz(m);
            m ^= 1u << j0;
            if (m) {
                const int j1 = 31 - __clz(m);
                m ^= 1u << j1;
                const float4 a0 = Q.q0[kq][j0], c0 = Q.q1[kq][j0], a1 = Q.q0[kq][j1], c1 = Q.q1[kq][j1];
                const float2 e0 = Q.bi[kq][j0], e1 = Q.bi[kq][j1];
                const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, cbase + j0, last_contributor);
                const EntryHead h1 = bwd_head(a1, c1, pxf, pyf, cbase + j1, last_contributor);
                const bool any0 = __any_sync(0xffffffffu, h0.live), any1 = __any_sync(0xffffffffu, h1.live);
                if (any0) {
                    float wv = 0.f, atv = 0.f;
                    if (h0.live)
                        bwd_live_sd(h0.alpha, h0.araw, c0.z, c0.w, e0.x, dp0, dp1, dp2, T, Sd, wv, atv);
                    slot0[8 * cnt] = make_float2(wv, atv);
                    if (lane == 0) R.meta[cnt] = make_float4(a0.x, a0.y, e0.y, 0.f);
                    cnt++;
                }
                if (any1) {
                    float wv = 0.f, atv = 0.f;
                    if (h1.live)
                        bwd_live_sd(h1.alpha, h1.araw, c1.z, c1.w, e1.x, dp0, dp1, dp2, T, Sd, wv, atv);
                    slot0[8 * cnt] = make_float2(wv, atv);
                    if (lane == 0) R.meta[cnt] = make_float4(a1.x, a1.y, e1.y, 0.f);
                    cnt++;
                }
            } else {
                const float4 a0 = Q.q0[kq][j0], c0 = Q.q1[kq][j0];
                const float2 e0 = Q.bi[kq][j0];
                const EntryHead h0 = bwd_head(a0, c0, pxf, pyf, cbase + j0, last_contributor);
                if (!__any_sync(0xffffffffu, h0.live)) continue;
                float wv = 0.f, atv = 0.f;
                if (h0.live)
                    bwd_live_sd(h0.alpha, h0.araw, c0.z, c0.w, e0.x, dp0, dp1, dp2, T, Sd, wv, atv);
                slot0[8 * cnt] = make_float2(wv, atv);
                if (lane == 0) R.meta[cnt] = make_float4(a0.x, a0.y, e0.y, 0.f);
                cnt++;
            }
        }
        // masks move up one step; extract the mask of step s+D+2 (its block becomes wA when it starts a new one)
#pragma unroll
        for (int k = 0; k < D + 1; k++) mq[k] = mq[k + 1];
        const int sn = s + D + 2;
        if ((sn >> 5) != blkA) { wA = wB; blkA++; wB = load_words(blkA + 1); }
        const uint32_t wn = __shfl_sync(0xffffffffu, wA, sn & 31);
        mq[D + 1] = sn < nch ? wn : 0u;
        __syncwarp();                  // everyone has left ring slot kq before the next step's gather overwrites it
    }
    if (cnt) red_flush(R, S.b1, cnt, lane, X0s, Y0s, accum);
}

int launch_blend_backward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                          const float* dL_dpix, float* accum, cudaStream_t s)
{
    const int gyv = (H + GSR_TILE - 1) / GSR_TILE;
    const dim3 grid((W + GSR_TILE - 1) / GSR_TILE, gyv * V);
    const uint32_t pw = (uint32_t)blend_plane_words(R, W, gyv * V * GSR_TILE);
    if (V > 1 && env_int("GSR_BLEND", 0) != 0) return GSR_EINVAL;        // comparison arms are single-view
    const int variant = env_int("GSR_BLEND", 0);       // 0: v5 (needs the planes of blend_fwd_v3), 1 / 3 / 4: comparison arms
    const uint32_t* lpt = env_int("GSR_LPT", 1) ? im.tile_order : nullptr;
    if (env_int("GSR_BLEND", 0) == 1)
        GSR_LAUNCH(blend_bwd_kernel, grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib, dL_dpix,
                   accum);
    else if (env_int("GSR_BLEND", 0) == 3)
        GSR_LAUNCH((blend_bwd_v3_kernel<4>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   dL_dpix, accum);
    else if (variant == 0) {
        static bool attr5_set = false;
        if (!attr5_set) {
            GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v5_kernel<2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV5<2>)));
            GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v5_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV5<2>)));
            GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v5_kernel<4, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV5<4>)));
            attr5_set = true;
        }
        const int ns = env_int("GSR_BWD_NS", 2), minb = env_int("GSR_BWD_MINB", 4);
        if (ns == 4)
            GSR_LAUNCH((blend_bwd_v5_kernel<4, 3>), grid, BL_THREADS, sizeof(BwdSmemV5<4>), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H,
                       bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        else if (minb == 3)
            GSR_LAUNCH((blend_bwd_v5_kernel<2, 3>), grid, BL_THREADS, sizeof(BwdSmemV5<2>), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H,
                       bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        else
            GSR_LAUNCH((blend_bwd_v5_kernel<2, 4>), grid, BL_THREADS, sizeof(BwdSmemV5<2>), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H,
                       bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
    } else {
        static bool attr_set = false;
        if (!attr_set) {
            GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v4_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV4)));
            GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v4_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV4)));
            attr_set = true;
        }
        if (env_int("GSR_BWD_MINB", 4) == 3)
            GSR_LAUNCH((blend_bwd_v4_kernel<3>), grid, BL_THREADS, sizeof(BwdSmemV4), s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T,
                       im.n_contrib, dL_dpix, accum);
        else
            GSR_LAUNCH((blend_bwd_v4_kernel<4>), grid, BL_THREADS, sizeof(BwdSmemV4), s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T,
                       im.n_contrib, dL_dpix, accum);
    }
    return GSR_OK;
}

}  // namespace gsr

// debug hook (not declared in include/gsr.h): reads and optionally clears the counters of blend_stats_kernel
extern "C" int gsr_debug_blend_stats(unsigned long long* out16, int reset)
{
    if (out16 && cudaMemcpyFromSymbol(out16, gsr::g_blend_stats, sizeof(unsigned long long) * 16) != cudaSuccess) return GSR_ECUDA;
    if (reset) {
        unsigned long long z[16] = {0};
        if (cudaMemcpyToSymbol(gsr::g_blend_stats, z, sizeof z) != cudaSuccess) return GSR_ECUDA;
    }
    return GSR_OK;
}
