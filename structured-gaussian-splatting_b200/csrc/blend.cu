// blend.cu - per-tile alpha blending, forward and backward (product kernels).
//
// Replaces (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer):
//   blend_fwd_v3_kernel          <- renderCUDA<3> (forward)   forward.cu:261-374
//   blend_bwd_v6_kernel (/ v5)   <- renderCUDA<3> (backward)  backward.cu:399-557
//
// One CTA per 16x16 tile, eight warps, each warp owns an 8x4 pixel sub-tile.  Tile-list entries are staged in shared
// memory as packed records gathered with 16-byte cp.async copies (position + conic + opacity + colour: everything the
// inner loop touches), double buffered.  The per-pixel arithmetic keeps the reference's expression order: the skip
// decisions (power > 0, alpha < 1/255, T < 1e-4) must be identical or a pixel can differ by up to 1/255.
// The plain kernels without sub-tile culling, the statistics kernel and the cp.async.bulk staging arm live in
// arms/blend_arms.cu (test-only library).
#include "blend_common.cuh"

namespace gsr {

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

// ------------------------------------------------------------------------------------------ forward, packed pairs
// Same staging, classification and exported planes as blend_fwd_v3_kernel.  A warp no longer walks the set bits of its
// classification words entry by entry (seven index instructions per entry): it queues the batch's surviving entries (one
// byte each), then 32 at a time one lane per entry copies them into a warp-private array of pair records (PairRec; about
// two instructions per entry all told), and the blending loop is a counted loop over pairs whose pixel-independent
// arithmetic runs as FADD2 / FMUL2 / FFMA2 (blend_common.cuh).  The sequential part (blend_step) is the scalar code of v3.
struct __align__(16) FwdPairs { PairRec p[16]; };

template <int MINB, int UNR = 1>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_fwd_v4_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    int W, int H, const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                    float* __restrict__ out, uint32_t* __restrict__ planes, uint32_t plane_words, int gyv,
                    const uint32_t* __restrict__ tile_order)
{
    __shared__ BlendStage stage[2];
    __shared__ uint32_t bits[2][8][8];     // [buffer][consumer warp][chunk of 32 entries]
    __shared__ FwdPairs pairs[8];
    __shared__ uint8_t queue[8][BL_THREADS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t lin = blockIdx.y * gridDim.x + blockIdx.x;                  // see blend_fwd_v3_kernel
    const uint32_t tile = tile_order ? tile_order[lin] : lin;
    const uint32_t trow = tile / gridDim.x, bx = tile - trow * gridDim.x;
    const uint32_t view = trow / (uint32_t)gyv, by = trow - view * (uint32_t)gyv;
    const float X0 = (float)(bx * GSR_TILE), Y0 = (float)(by * GSR_TILE);
    const uint32_t px = bx * GSR_TILE + (wid & 1) * 8 + (lane & 7);
    const uint32_t py = by * GSR_TILE + (wid >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const float pxf = (float)px, pyf = (float)py;
    bool done = !inside;

    const uint2 range = ranges[tile];
    const int total = (int)(range.y - range.x);
    uint32_t* const plane = planes + (size_t)lane * plane_words + (range.x >> 5) + 16u * tile + wid;
    const int rounds = (total + BL_THREADS - 1) / BL_THREADS;
    const uint32_t below = (1u << lane) - 1u;
    float* const pw = reinterpret_cast<float*>(pairs[wid].p);                 // this warp's pair records, as words
    const PairRec* const pr = pairs[wid].p;
    uint8_t* const qu = queue[wid];

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float2 C01 = make_float2(0.f, 0.f);         // red and green accumulate as one register pair (FFMA2)
    float C2 = 0.f;

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
        // the warp's surviving entries of this batch, in list order: one byte per entry
        int cnt = 0;
#pragma unroll
        for (int cch = 0; cch < 8; cch++) {
            const uint32_t m = bits[buf][wid][cch];
            if ((m >> lane) & 1u) qu[cnt + __popc(m & below)] = (uint8_t)(cch * 32 + lane);
            cnt += __popc(m);
        }
        __syncwarp();
#pragma unroll 1
        for (int g = 0; g < cnt; g += 32) {
            const int n = min(32, cnt - g);
            if (lane < n) {                             // entry g + lane becomes slot lane & 1 of pair lane >> 1
                const int j = qu[g + lane];
                const float4 a = st.q0[j], c = st.q1[j];
                const float b = st.b[j];
                float* w = pw + (lane >> 1) * (int)(sizeof(PairRec) / 4) + (lane & 1);
                w[0] = a.x; w[2] = a.y; w[4] = a.z; w[6] = a.w; w[8] = c.x; w[10] = c.y;
                w[12 + (lane & 1)] = c.z; w[13 + (lane & 1)] = c.w;          // colours: (r0 g0 r1 g1), see blend_step_rg
               
                w[16] = b; w[18] = __uint_as_float(pos_base + (uint32_t)j);
            } else if (lane == n) {                     // an odd group's last pair: its second slot blends nothing (opacity 0)
                float* w = pw + (lane >> 1) * (int)(sizeof(PairRec) / 4) + (lane & 1);
                w[0] = 0.f; w[2] = 0.f; w[4] = 0.f; w[6] = 0.f; w[8] = 0.f; w[10] = 0.f;
            }
            __syncwarp();
#pragma unroll UNR
            for (int q = 0; 2 * q < n; q++) {
                const PairRec r = pr[q];
                float2 dx, dy, power, araw;
                pair_head(r, pxf, pyf, dx, dy, power, araw);
                const float al0 = fminf(0.99f, araw.x), al1 = fminf(0.99f, araw.y);
                const bool ok0 = !(power.x > 0.0f) && !(al0 < 1.0f / 255.0f);
                const bool ok1 = !(power.y > 0.0f) && !(al1 < 1.0f / 255.0f);
                blend_step_rg(al0, ok0, make_float2(r.rg.x, r.rg.y), r.bp.x, __float_as_uint(r.bp.z), done, T, C01, C2, last_contributor);
                blend_step_rg(al1, ok1, make_float2(r.rg.z, r.rg.w), r.bp.y, __float_as_uint(r.bp.w), done, T, C01, C2, last_contributor);
            }
            if (__all_sync(0xffffffffu, done)) break;    // (also orders this group's reads before the next group's writes)
        }
    }
    cp_async_wait<0>();
    if (inside) {
        const size_t HW = (size_t)H * W;
        const size_t pix = (size_t)W * py + px;
        final_T[view * HW + pix] = T;
        n_contrib[view * HW + pix] = last_contributor;
        float* o = out + 3 * view * HW + pix;
        o[0] = C01.x + T * bg[0];
        o[HW] = C01.y + T * bg[1];
        o[2 * HW] = C2 + T * bg[2];
    }
}

// classification words kept for the backward (in the binning blob's sort scratch, dead once the tile split is done):
// 8 planes (one per sub-tile of a tile) of (R >> 5) + 16 tiles + 16 words
size_t blend_plane_words(long long R, int W, int H)
{
    const size_t tiles = (size_t)((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE);
    return (size_t)((R > 0 ? R : 0) >> 5) + 16 * tiles + 16;
}

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

constexpr int RED_SLOTS = 8;
constexpr int RED_KSTRIDE = 144;      // floats per k-step of the slot array: 8 slots x 16 + 16 (STS of one slot: 2-way conflicts only)
struct __align__(16) WarpRed {
    float a[4 * RED_KSTRIDE];         // [k-step s][slot][t][b][w | at]   (lane = 8t + 2s + b)
    float2 b2[4][32];                 // B2 fragments: [k-step][lane (g, t)] = column g of dL/dpixel at lanes 8t+2s, 8t+2s+1
    float4 meta[RED_SLOTS];           // mean.x, mean.y, Gaussian id (bits)
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

// ------------------------------------------------------------------------------------------ backward, compacted work queue (v6)
// blend_bwd_v5 walks the tile list 32 entries per step and finds ~7 of them set in its classification word: the step
// overhead (mask bookkeeping, id loads, gather issue: ~70 instructions) is paid per ~7 entries, every entry costs a
// find-highest-bit sequence and a warp vote, and the tensor-core reduction runs with half of its rows idle (8 entries x
// (w | at) share one 16-row A operand).  v6 keeps the arithmetic and changes the bookkeeping:
//  * work queue: the warp compacts the set bits of its classification words into a queue of list positions in shared
//    memory (one STS per set bit, ~14 instructions per 32-entry word) and then works on DENSE groups of 32 queued entries:
//    every lane gathers one record per group (ids one group ahead, records one group ahead through cp.async into a
//    two-buffer ring), and the blending loop is a plain counted loop over the ring - no bit scans, no votes;
//  * 16-row reduction: w and at of an entry go to two separate A operands (16 entries x 32 lanes each), so each of the
//    16 HMMAs of a flush does useful work in all rows: one flush per 16 entries instead of one per 8, the entry's
//    centre and id are read from the ring at flush time (no per-entry meta store);
//  * entries none of whose lanes is live simply contribute zero sums, which the epilogue does not commit (a RED of 0 is
//    a no-op, so skipping it is exact): no warp vote per entry.
// Shared-memory layout of the A operands (floats): [k-step s][row g = entry & 7][t][pair slot p][w | at], pair slot
// p = 2 b + (hi ^ (s & 1)), hi = entry >> 3, rows padded to 36 floats.  A pixel lane L is K index (t, b) of k-step s with
// s = L >> 3, t = (L >> 1) & 3, b = L & 1: its STS.64 of (w, at) is bank-conflict free (the 16 lanes of a store phase
// have s in {0,1} or {2,3}; the parity swap spreads them over all 32 banks) and so is the reader's LDS.128 (row stride
// 36 = 4 mod 32).  The k-step loop is unrolled, so the parity swap costs no instruction on the reading side.
constexpr int V6_QCAP = 128;            // queue capacity (power of two; <= 95 live entries, see the fill loop)
// SLOTS = 16: w and at in two A operands of 16 entries each (rows padded to 36 floats, 288 per k-step);
// SLOTS = 8:  one A operand, rows 0-7 = w, rows 8-15 = at of 8 entries ([k-step][entry][t][b][w | at], 16 floats per
//             entry, 144 per k-step): half the shared memory (4 CTAs per SM instead of 3) for twice the HMMAs per entry
template <int SLOTS> struct V6Layout;
template <> struct V6Layout<16> { static constexpr int GS = 36, KS = 8 * 36; };
template <> struct V6Layout<8>  { static constexpr int GS = 16, KS = 144; };
template <int SLOTS>
struct __align__(16) WarpV6 {
    float4 q0[2][32];                   // ring: pix.x pix.y conic.x conic.y
    float4 q1[2][32];                   //       conic.z opacity r g
    float4 e[2][32];                    //       Gaussian id (bits), list position (bits), colour b, unused
    float  a[4 * V6Layout<SLOTS>::KS];
    float2 b2[4][32];                   // B2 fragments of this warp's dL/dpixel (hi, lo per channel)
    uint32_t queue[V6_QCAP];
};
template <int SLOTS>
struct __align__(16) BwdSmemV6 {
    WarpV6<SLOTS> w[8];
    float2 b1[4][32];                   // B1 fragments (moment matrix), same for all warps
};

// x = hi + lo with hi exact in tf32 (the tensor core ignores the 13 low mantissa bits)
__device__ __forceinline__ void split_tf32(const float (&x)[4], uint32_t (&hi)[4], uint32_t (&lo)[4])
{
#pragma unroll
    for (int k = 0; k < 4; k++) {
        hi[k] = __float_as_uint(x[k]) & 0xffffe000u;
        lo[k] = __float_as_uint(x[k] - __uint_as_float(hi[k]));
    }
}

// one reduced row (entry j of ring buffer kq): moments m0, m1 = columns 2t, 2t+1 of D1 (S0 Sx | Sy Sxx | Sxy Syy | - -,
// about the sub-tile origin), col = colour sum t; shifted to the Gaussian's centre and committed
template <class Fetch>
__device__ __forceinline__ void v6_commit(const Fetch& fetch, int j, bool valid, int lane, float m0, float m1, float col,
                                          float X0s, float Y0s, float* __restrict__ accum)
{
    const int t = lane & 3, q0l = lane & ~3;
    const float S0 = __shfl_sync(0xffffffffu, m0, q0l);
    const float Sx = __shfl_sync(0xffffffffu, m1, q0l);
    const float Sy = __shfl_sync(0xffffffffu, m0, q0l + 1);
    if (valid && t < 3) {
        float2 xy;
        uint32_t id;
        fetch(j, xy, id);                                    // the entry's centre and Gaussian id
        const float u = xy.x - X0s, v = xy.y - Y0s;         // dx = u - lx, dy = v - ly
        const float a1 = fmaf(u, S0, -Sx);                   // sum w dx
        const float a2 = fmaf(v, S0, -Sy);                   // sum w dy
        const float o3 = fmaf(u, a1 - Sx, m1);               // sum w dx^2  = Sxx - 2u Sx + u^2 S0      (lane t = 1)
        const float o4 = fmaf(-v, Sx, fmaf(u, a2, m0));      // sum w dx dy = Sxy - u Sy - v Sx + uv S0 (lane t = 2)
        const float o5 = fmaf(v, a2 - Sy, m1);               // sum w dy^2  = Syy - 2v Sy + v^2 S0      (lane t = 2)
        const float out0 = fsel(t == 0, m0, fsel(t == 1, a2, o4));
        const float out1 = fsel(t == 0, a1, fsel(t == 1, o3, o5));
        float* rowp = accum + (size_t)id * 12;
        // an entry without a live lane has exactly zero sums: adding them is a no-op, so they are not sent
        if (out0 != 0.f || out1 != 0.f) {
            if (t == 0) { atomicAdd(rowp + 8, out0); atomicAdd(rowp + 3, out1); }
            else red_add_v2(rowp + 2 + 2 * t, out0, out1);   // t = 1: columns 4, 5   t = 2: columns 6, 7
        }
        if (col != 0.f) atomicAdd(rowp + t, col);
    }
}

// v6_commit with every lane computing and sending: the moments of an entry go out as ONE two-float reduction per lane of
// its quad - lane 0: columns 8, 9 = (S0, 0), lane 1: 4, 5, lane 2: 6, 7, lane 3: columns 2, 3 = (0, sum w dx); column 9 is
// padding and a zero added to the blue column changes nothing - and the colour sums as one scalar reduction from lanes 0-2:
// two guarded instructions, the arithmetic outside any divergent region (the branchy form spends 15 % of
// blend_bwd_v7_kernel's time, a third of it resolving its four nested branches).  A lane whose row is not valid reads a
// stale ring entry; nothing of it is sent.
template <class Fetch>
__device__ __forceinline__ void v6_commit_quad(const Fetch& fetch, int j, bool valid, int lane, float m0, float m1, float col,
                                               float X0s, float Y0s, float* __restrict__ accum)
{
    const int t = lane & 3, q0l = lane & ~3;
    const float S0 = __shfl_sync(0xffffffffu, m0, q0l);
    const float Sx = __shfl_sync(0xffffffffu, m1, q0l);
    const float Sy = __shfl_sync(0xffffffffu, m0, q0l + 1);
    float2 xy;
    uint32_t id;
    fetch(j, xy, id);
    const float u = xy.x - X0s, v = xy.y - Y0s;
    const float a1 = fmaf(u, S0, -Sx);                   // sum w dx
    const float a2 = fmaf(v, S0, -Sy);                   // sum w dy
    const float o3 = fmaf(u, a1 - Sx, m1);               // sum w dx^2      (lane 1)
    const float o4 = fmaf(-v, Sx, fmaf(u, a2, m0));      // sum w dx dy     (lane 2)
    const float o5 = fmaf(v, a2 - Sy, m1);               // sum w dy^2      (lane 2)
    const float out0 = fsel(t == 0, m0, fsel(t == 1, a2, fsel(t == 2, o4, 0.f)));
    const float out1 = fsel(t == 0, 0.f, fsel(t == 1, o3, fsel(t == 2, o5, a1)));
    const int off = (0x2648 >> (4 * t)) & 15;            // t = 0, 1, 2, 3 -> columns 8, 4, 6, 2
    float* rowp = accum + (size_t)id * 12;
    // an entry without a live lane has exactly zero sums: adding them is a no-op, so they are not sent
    if (valid && (out0 != 0.f || out1 != 0.f)) red_add_v2(rowp + off, out0, out1);
    if (valid && t < 3 && col != 0.f) atomicAdd(rowp + t, col);
}

// Reduces rows [0, cnt) of the warp's A operand(s) (entries jbase .. jbase + cnt - 1 of ring buffer kq) over the 32 lanes
// and commits them.  Fragments of m16n8k8 (g = lane >> 2, t = lane & 3): A rows g / g+8 at k = t / t+4, B column g at
// k = t / t+4, D rows g / g+8, columns 2t, 2t+1.  Columns of D2: r_hi r_lo | g_hi g_lo | b_hi b_lo | - -.
template <int SLOTS, bool PRED = false, class Fetch>
__device__ __forceinline__ void v6_flush(const float* __restrict__ Sa, const float2 (*Sb2)[32], const float2 (*b1)[32], const Fetch& fetch,
                                         int jbase, int cnt, int lane, float X0s, float Y0s, float* __restrict__ accum)
{
    constexpr int GS = V6Layout<SLOTS>::GS, KS = V6Layout<SLOTS>::KS;
    __syncwarp();
    const int g = lane >> 2, t = lane & 3;
    float d1[4] = {0.f, 0.f, 0.f, 0.f}, d2[4] = {0.f, 0.f, 0.f, 0.f};
    if (SLOTS == 16) {
        const float* ab = Sa + g * GS + t * 8;
#pragma unroll
        for (int s = 0; s < 4; s++) {
            const float4 Q1 = *reinterpret_cast<const float4*>(ab + s * KS);          // k = t:   (w, at) of the two rows
            const float4 Q2 = *reinterpret_cast<const float4*>(ab + s * KS + 4);      // k = t+4
            const float2 m = b1[s][lane];
            const float2 q = Sb2[s][lane];
            float xw[4], xa[4];
            if ((s & 1) == 0) {
                xw[0] = Q1.x; xw[1] = Q1.z; xw[2] = Q2.x; xw[3] = Q2.z;
                xa[0] = Q1.y; xa[1] = Q1.w; xa[2] = Q2.y; xa[3] = Q2.w;
            } else {
                xw[0] = Q1.z; xw[1] = Q1.x; xw[2] = Q2.z; xw[3] = Q2.x;
                xa[0] = Q1.w; xa[1] = Q1.y; xa[2] = Q2.w; xa[3] = Q2.y;
            }
            uint32_t whi[4], wlo[4], ahi[4], alo[4];
            split_tf32(xw, whi, wlo);
            split_tf32(xa, ahi, alo);
            mma_tf32(d1, whi, __float_as_uint(m.x), __float_as_uint(m.y));
            mma_tf32(d1, wlo, __float_as_uint(m.x), __float_as_uint(m.y));
            mma_tf32(d2, ahi, __float_as_uint(q.x), __float_as_uint(q.y));
            mma_tf32(d2, alo, __float_as_uint(q.x), __float_as_uint(q.y));
        }
#pragma unroll
        for (int hi = 0; hi < 2; hi++) {
            if (PRED)
                v6_commit_quad(fetch, jbase + g + 8 * hi, g + 8 * hi < cnt, lane, d1[2 * hi], d1[2 * hi + 1], d2[2 * hi] + d2[2 * hi + 1],
                               X0s, Y0s, accum);
            else
                v6_commit(fetch, jbase + g + 8 * hi, g + 8 * hi < cnt, lane, d1[2 * hi], d1[2 * hi + 1], d2[2 * hi] + d2[2 * hi + 1],
                          X0s, Y0s, accum);
        }
    } else {
        const float* ab = Sa + g * GS + t * 4;
#pragma unroll
        for (int s = 0; s < 4; s++) {
            const float4 av = *reinterpret_cast<const float4*>(ab + s * KS);           // w, at at k = t | w, at at k = t+4
            const float2 m = b1[s][lane];
            const float2 q = Sb2[s][lane];
            const float x[4] = {av.x, av.y, av.z, av.w};      // rows g (w) / g+8 (at) at k = t, then at k = t+4
            uint32_t hi[4], lo[4];
            split_tf32(x, hi, lo);
            mma_tf32(d1, hi, __float_as_uint(m.x), __float_as_uint(m.y));
            mma_tf32(d1, lo, __float_as_uint(m.x), __float_as_uint(m.y));
            mma_tf32(d2, hi, __float_as_uint(q.x), __float_as_uint(q.y));
            mma_tf32(d2, lo, __float_as_uint(q.x), __float_as_uint(q.y));
        }
        v6_commit(fetch, jbase + g, g < cnt, lane, d1[0], d1[1], d2[2] + d2[3], X0s, Y0s, accum);
    }
    __syncwarp();
}

// head of one entry for this lane's pixel, and the branch-free state update: a lane that does not blend the entry (dead)
// runs the same instructions with alpha = 0, which leaves T and Sd unchanged and yields w = at = 0
__device__ __forceinline__ void v6_entry(const float4 a, const float4 c, const float4 e, float pxf, float pyf, int last_contributor,
                                         bool slot_valid, float dp0, float dp1, float dp2, float& T, float& Sd, float& w, float& at)
{
    const float dx = a.x - pxf, dy = a.y - pyf;
    const float power = -0.5f * (a.z * dx * dx + c.x * dy * dy) - a.w * dx * dy;
    const float araw = c.y * expf(power);
    const float alpha = fminf(0.99f, araw);
    const bool live = slot_valid && __float_as_int(e.y) < last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
    const float araw_s = live ? araw : 0.f, alpha_s = live ? alpha : 0.f;
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f - alpha_s));        // rcp(1) == 1 exactly
    T *= r;
    const float ec = fmaf(e.z, dp2, fmaf(c.w, dp1, c.z * dp0));
    at = alpha_s * T;
    const float dL_dalpha = fmaf(T, ec, -(Sd * r));
    Sd = fmaf(ec, at, Sd);
    w = araw_s * dL_dalpha;
}

template <int SLOTS, int TRIP, int MINB>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_bwd_v6_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    const uint32_t* __restrict__ planes, uint32_t plane_words, int W, int H, const float* __restrict__ bg,
                    const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                    float* __restrict__ accum, int gyv, const uint32_t* __restrict__ tile_order)
{
    constexpr int GS = V6Layout<SLOTS>::GS, KS = V6Layout<SLOTS>::KS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmemV6<SLOTS>& SM = *reinterpret_cast<BwdSmemV6<SLOTS>*>(smem_raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpV6<SLOTS>& S = SM.w[wid];
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

    // this pixel lane as a K index of the reduction: k-step sK, index tK + 4 bK
    const int sK = lane >> 3, tK = (lane >> 1) & 3, bK = lane & 1;
    // B1: the moment matrix (1, lx, ly, lx^2, lx ly, ly^2, 0, 0) in B-fragment order; reading lane l = (n = l >> 2, t = l & 3)
    // of k-step s takes column n at the pixels (lx, ly) = (2t, s) and (2t + 1, s)
    if (threadIdx.x < 128) {
        const int s = threadIdx.x >> 5, l = threadIdx.x & 31, n = l >> 2;
        const float ly = (float)s, lx0 = (float)(2 * (l & 3)), lx1 = lx0 + 1.f;
        float2 m;
        m.x = n == 0 ? 1.f : n == 1 ? lx0 : n == 2 ? ly : n == 3 ? lx0 * lx0 : n == 4 ? lx0 * ly : n == 5 ? ly * ly : 0.f;
        m.y = n == 0 ? 1.f : n == 1 ? lx1 : n == 2 ? ly : n == 3 ? lx1 * lx1 : n == 4 ? lx1 * ly : n == 5 ? ly * ly : 0.f;
        SM.b1[s][l] = m;
    }
    {
        // B2: columns (r_hi, r_lo, g_hi, g_lo, b_hi, b_lo, 0, 0) of this warp's dL/dpixel; column n of this pixel lives at
        // reading lane 4 n + tK, component bK
        const float d[3] = {dp0, dp1, dp2};
        float* col = reinterpret_cast<float*>(&S.b2[sK][tK]) + bK;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float hi = __uint_as_float(__float_as_uint(d[c]) & 0xffffe000u);
            col[(2 * c) * 8] = hi;
            col[(2 * c + 1) * 8] = d[c] - hi;
        }
        col[6 * 8] = 0.f;
        col[7 * 8] = 0.f;
    }
    // this lane's (w, at) slot of entry c of a flush group (see the layouts above):
    //   SLOTS 16: (c < 8 ? a_lo : a_hi) + (c & 7) * GS       SLOTS 8: a_lo + c * GS
    float* const a_lo = SLOTS == 16 ? S.a + sK * KS + tK * 8 + 4 * bK + 2 * (sK & 1) : S.a + sK * KS + tK * 4 + 2 * bK;
    float* const a_hi = S.a + sK * KS + tK * 8 + 4 * bK + 2 - 2 * (sK & 1);

    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    warp_last = min(warp_last, (int)(range.y - range.x));
    __syncthreads();                       // b1 is visible; the only block barrier
    const int nch = (warp_last + 31) >> 5; // chunks of 32 list entries this warp has work in
    if (nch == 0) return;
    const uint32_t* plane = planes + (size_t)wid * plane_words + (range.x >> 5) + 16u * tile;
    const uint32_t* tvals = vals + range.x;

    // ---- work queue: step st = 0 .. nch-1 handles chunk nch-1-st (back to front); the words of 32 consecutive steps sit in
    // one register per lane (wA: block blkA, wB: the block after it)
    auto load_words = [&](int blk) -> uint32_t {
        const int s = blk * 32 + lane;
        return s < nch ? plane[nch - 1 - s] : 0u;
    };
    uint32_t wA = load_words(0), wB = load_words(1);
    int blkA = 0, st_next = 0;
    uint32_t qtail = 0;                                    // entries queued so far (warp-uniform)
    const uint32_t above = ~(0xffffffffu >> lane);         // the bits lane `lane` ranks behind: it owns bit 31 - lane
    {
        // the top chunk may reach beyond the warp's last contributor
        const int rem = warp_last - (nch - 1) * 32;        // 1 .. 32
        if (lane == 0 && rem < 32) wA &= (1u << rem) - 1u;
    }
    auto fill = [&](uint32_t want) {                       // queue entries until `want` are there or the list is exhausted
        while (qtail < want && st_next < nch) {
            const uint32_t m = __shfl_sync(0xffffffffu, wA, st_next & 31);
            const uint32_t base = (uint32_t)(nch - 1 - st_next) * 32u;
            if ((m >> (31 - lane)) & 1u) S.queue[(qtail + __popc(m & above)) & (V6_QCAP - 1)] = base + 31u - (uint32_t)lane;
            qtail += __popc(m);
            st_next++;
            if ((st_next & 31) == 0) { wA = wB; blkA++; wB = load_words(blkA + 1); }
        }
        __syncwarp();
    };
    // group k = queue entries [32k, 32k + 32); its size is final once `fill(32 (k + 1))` has returned
    auto group_size = [&](uint32_t k) -> int { return (int)min(32u, qtail - min(qtail, 32u * k)); };
    auto load_pos = [&](uint32_t k, uint32_t& pos, uint32_t& id) {      // lane's entry of group k: list position and Gaussian id
        const bool v = lane < group_size(k);
        pos = v ? S.queue[(32u * k + (uint32_t)lane) & (V6_QCAP - 1)] : 0u;
        id = v ? tvals[pos] : 0u;
    };
    auto issue = [&](uint32_t k, uint32_t pos, uint32_t id) {           // gather group k into ring buffer k & 1
        const int kb = k & 1, gs = group_size(k);
        if (lane < gs) {
            const SplatRec* r = rec + id;
            cp_async16(&S.q0[kb][lane], &r->q0);
            cp_async16(&S.q1[kb][lane], &r->q1);
            cp_async4(&S.e[kb][lane].z, &r->q2);
            *reinterpret_cast<uint2*>(&S.e[kb][lane]) = make_uint2(id, pos);
        } else if (lane < gs + TRIP - 1) {
            // the slots a trip may read behind the last entry get a record that blends nothing and holds no NaN
            S.q0[kb][lane] = make_float4(0.f, 0.f, 0.f, 0.f);
            S.q1[kb][lane] = make_float4(0.f, 0.f, 0.f, 0.f);
            S.e[kb][lane] = make_float4(0.f, __int_as_float(0x7fffffff), 0.f, 0.f);
        }
        cp_async_commit();
    };

    uint32_t posc, idc;
    fill(64);
    load_pos(0, posc, idc);
    issue(0, posc, idc);
    load_pos(1, posc, idc);
#pragma unroll 1
    for (uint32_t k = 0;; k++) {
        const int n = group_size(k);
        if (n == 0) break;                                  // (one empty cp.async group may still be pending: harmless)
        fill(32u * (k + 3));
        uint32_t posn, idn;
        load_pos(k + 2, posn, idn);                         // ids two groups ahead
        issue(k + 1, posc, idc);                            // records one group ahead
        posc = posn; idc = idn;
        cp_async_wait<1>();                                 // my records of group k have landed
        __syncwarp();
        const int kq = k & 1;
        const float4* rq0 = S.q0[kq];
        const float4* rq1 = S.q1[kq];
        const float4* re = S.e[kq];
#pragma unroll 1
        for (int jb = 0; jb < n; jb += SLOTS) {
            const int cnt = min(SLOTS, n - jb);             // entries of this flush group (warp-uniform)
#pragma unroll
            for (int c = 0; c < SLOTS; c += TRIP) {
                if (c < cnt) {
                    float4 a[TRIP], q[TRIP], e[TRIP];
                    float w[TRIP], at[TRIP];
#pragma unroll
                    for (int u = 0; u < TRIP; u++) { a[u] = rq0[jb + c + u]; q[u] = rq1[jb + c + u]; e[u] = re[jb + c + u]; }
#pragma unroll
                    for (int u = 0; u < TRIP; u++)
                        v6_entry(a[u], q[u], e[u], pxf, pyf, last_contributor, u == 0 || c + u < cnt, dp0, dp1, dp2, T, Sd, w[u], at[u]);
#pragma unroll
                    for (int u = 0; u < TRIP; u++) {
                        const int cc = c + u;
                        float* const sp = SLOTS == 16 ? ((cc >> 3) ? a_hi : a_lo) + (cc & 7) * GS : a_lo + cc * GS;
                        *reinterpret_cast<float2*>(sp) = make_float2(w[u], at[u]);
                    }
                }
            }
            v6_flush<SLOTS>(S.a, S.b2, SM.b1, [&](int j, float2& xy, uint32_t& id) {
                xy = *reinterpret_cast<const float2*>(&rq0[j]);
                id = __float_as_uint(re[j].x);
            }, jb, cnt, lane, X0s, Y0s, accum);
        }
        __syncwarp();                  // everyone has left ring buffer kq before the gather of group k + 2 overwrites it
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------ backward, packed pairs
// blend_bwd_v6_kernel<16, 2> with the pixel-independent arithmetic of the two entries of a trip on register pairs
// (FADD2 / FMUL2 / FFMA2, blend_common.cuh): the gathered group is rewritten in place as pair records, the per-entry
// state chain (T, Sd) stays scalar.  Same operations in the same order as v6_entry: results are bit-identical to v6.
struct __align__(16) WarpV7 {
    float4 ring[2][96];                 // as gathered: [0, 32) q0 of entry = lane, [32, 64) q1, [64, 96) (id, pos, b, -)
                                        // as blended:  PairRec[16] in [0, 80), the 32 Gaussian ids in [80, 88)
    float  a[4 * V6Layout<16>::KS];
    float2 b2[4][32];
    uint32_t queue[V6_QCAP];
};
struct __align__(16) BwdSmemV7 {
    WarpV7 w[8];
    float2 b1[4][32];
};

__device__ __forceinline__ void v7_pair(const PairRec r, float pxf, float pyf, int last_contributor, float dp0, float dp1, float dp2,
                                        float& T, float& Sd, float& w0, float& at0, float& w1, float& at1)
{
    float2 dx, dy, power, araw;
    pair_head(r, pxf, pyf, dx, dy, power, araw);
    const float al0 = fminf(0.99f, araw.x), al1 = fminf(0.99f, araw.y);
    const bool live0 = __float_as_int(r.bp.z) < last_contributor && !(power.x > 0.0f) && !(al0 < 1.0f / 255.0f);
    const bool live1 = __float_as_int(r.bp.w) < last_contributor && !(power.y > 0.0f) && !(al1 < 1.0f / 255.0f);
    const float2 araw_s = make_float2(live0 ? araw.x : 0.f, live1 ? araw.y : 0.f);
    const float2 alpha_s = make_float2(live0 ? al0 : 0.f, live1 ? al1 : 0.f);
    const float2 om = __fadd2_rn(bc2(1.f), neg2(alpha_s));
    float r0, r1;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(om.x));              // rcp(1) == 1 exactly
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(om.y));
    const float2 ec = __ffma2_rn(make_float2(r.bp.x, r.bp.y), bc2(dp2),
                                 __ffma2_rn(make_float2(r.rg.z, r.rg.w), bc2(dp1), __fmul2_rn(make_float2(r.rg.x, r.rg.y), bc2(dp0))));
    T *= r0;
    at0 = alpha_s.x * T;
    const float dL0 = fmaf(T, ec.x, -(Sd * r0));
    Sd = fmaf(ec.x, at0, Sd);
    w0 = araw_s.x * dL0;
    T *= r1;
    at1 = alpha_s.y * T;
    const float dL1 = fmaf(T, ec.y, -(Sd * r1));
    Sd = fmaf(ec.y, at1, Sd);
    w1 = araw_s.y * dL1;
}

// IDQ: the Gaussian ids of a group travel list -> shared memory by cp.async, in the same commit group as the records of the
// group before it, and are read from shared memory one iteration later.  (Loaded into a register two groups ahead, the
// compiler rotates the register with a MOV right behind the load and the warp waits there for the load: 3 % of the
// kernel's samples.)  The ids land in the eight float4 the pair-record form of the ring buffer leaves unused.
template <int MINB, bool FULL, bool PRED = false, bool IDQ = false>
__global__ void __launch_bounds__(BL_THREADS, MINB)
blend_bwd_v7_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals, const SplatRec* __restrict__ rec,
                    const uint32_t* __restrict__ planes, uint32_t plane_words, int W, int H, const float* __restrict__ bg,
                    const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
                    float* __restrict__ accum, int gyv, const uint32_t* __restrict__ tile_order)
{
    constexpr int SLOTS = 16, GS = V6Layout<16>::GS, KS = V6Layout<16>::KS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BwdSmemV7& SM = *reinterpret_cast<BwdSmemV7*>(smem_raw);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpV7& S = SM.w[wid];
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

    // this pixel lane as a K index of the reduction: k-step sK, index tK + 4 bK
    const int sK = lane >> 3, tK = (lane >> 1) & 3, bK = lane & 1;
    // B1: the moment matrix (1, lx, ly, lx^2, lx ly, ly^2, 0, 0) in B-fragment order; reading lane l = (n = l >> 2, t = l & 3)
    // of k-step s takes column n at the pixels (lx, ly) = (2t, s) and (2t + 1, s)
    if (threadIdx.x < 128) {
        const int s = threadIdx.x >> 5, l = threadIdx.x & 31, n = l >> 2;
        const float ly = (float)s, lx0 = (float)(2 * (l & 3)), lx1 = lx0 + 1.f;
        float2 m;
        m.x = n == 0 ? 1.f : n == 1 ? lx0 : n == 2 ? ly : n == 3 ? lx0 * lx0 : n == 4 ? lx0 * ly : n == 5 ? ly * ly : 0.f;
        m.y = n == 0 ? 1.f : n == 1 ? lx1 : n == 2 ? ly : n == 3 ? lx1 * lx1 : n == 4 ? lx1 * ly : n == 5 ? ly * ly : 0.f;
        SM.b1[s][l] = m;
    }
    {
        // B2: columns (r_hi, r_lo, g_hi, g_lo, b_hi, b_lo, 0, 0) of this warp's dL/dpixel; column n of this pixel lives at
        // reading lane 4 n + tK, component bK
        const float d[3] = {dp0, dp1, dp2};
        float* col = reinterpret_cast<float*>(&S.b2[sK][tK]) + bK;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float hi = __uint_as_float(__float_as_uint(d[c]) & 0xffffe000u);
            col[(2 * c) * 8] = hi;
            col[(2 * c + 1) * 8] = d[c] - hi;
        }
        col[6 * 8] = 0.f;
        col[7 * 8] = 0.f;
    }
    // this lane's (w, at) slot of entry c of a flush group (see the layouts above):
    //   (c < 8 ? a_lo : a_hi) + (c & 7) * GS
    float* const a_lo = S.a + sK * KS + tK * 8 + 4 * bK + 2 * (sK & 1);
    float* const a_hi = S.a + sK * KS + tK * 8 + 4 * bK + 2 - 2 * (sK & 1);

    int warp_last = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) warp_last = max(warp_last, __shfl_xor_sync(0xffffffffu, warp_last, o));
    warp_last = min(warp_last, (int)(range.y - range.x));
    __syncthreads();                       // b1 is visible; the only block barrier
    const int nch = (warp_last + 31) >> 5; // chunks of 32 list entries this warp has work in
    if (nch == 0) return;
    const uint32_t* plane = planes + (size_t)wid * plane_words + (range.x >> 5) + 16u * tile;
    const uint32_t* tvals = vals + range.x;

    // ---- work queue: step st = 0 .. nch-1 handles chunk nch-1-st (back to front); the words of 32 consecutive steps sit in
    // one register per lane (wA: block blkA, wB: the block after it)
    auto load_words = [&](int blk) -> uint32_t {
        const int s = blk * 32 + lane;
        return s < nch ? plane[nch - 1 - s] : 0u;
    };
    uint32_t wA = load_words(0), wB = load_words(1);
    int blkA = 0, st_next = 0;
    uint32_t qtail = 0;                                    // entries queued so far (warp-uniform)
    const uint32_t above = ~(0xffffffffu >> lane);         // the bits lane `lane` ranks behind: it owns bit 31 - lane
    {
        // the top chunk may reach beyond the warp's last contributor
        const int rem = warp_last - (nch - 1) * 32;        // 1 .. 32
        if (lane == 0 && rem < 32) wA &= (1u << rem) - 1u;
    }
    auto fill = [&](uint32_t want) {                       // queue entries until `want` are there or the list is exhausted
        while (qtail < want && st_next < nch) {
            const uint32_t m = __shfl_sync(0xffffffffu, wA, st_next & 31);
            const uint32_t base = (uint32_t)(nch - 1 - st_next) * 32u;
            if ((m >> (31 - lane)) & 1u) S.queue[(qtail + __popc(m & above)) & (V6_QCAP - 1)] = base + 31u - (uint32_t)lane;
            qtail += __popc(m);
            st_next++;
            if ((st_next & 31) == 0) { wA = wB; blkA++; wB = load_words(blkA + 1); }
        }
        __syncwarp();
    };
    // group k = queue entries [32k, 32k + 32); its size is final once `fill(32 (k + 1))` has returned
    auto group_size = [&](uint32_t k) -> int { return (int)min(32u, qtail - min(qtail, 32u * k)); };
    auto load_pos = [&](uint32_t k, uint32_t& pos, uint32_t& id) {      // lane's entry of group k: list position and Gaussian id
        const bool v = lane < group_size(k);
        pos = v ? S.queue[(32u * k + (uint32_t)lane) & (V6_QCAP - 1)] : 0u;
        id = v ? tvals[pos] : 0u;
    };
    auto issue = [&](uint32_t k, uint32_t pos, uint32_t id) {           // gather group k into ring buffer k & 1 (lane-indexed form)
        float4* R = S.ring[k & 1];
        if (lane < group_size(k)) {
            const SplatRec* r = rec + id;
            cp_async16(&R[lane], &r->q0);
            cp_async16(&R[32 + lane], &r->q1);
            cp_async4(&R[64 + lane].z, &r->q2);
            *reinterpret_cast<uint2*>(&R[64 + lane]) = make_uint2(id, pos);
        }
        cp_async_commit();
    };

    uint32_t posc, idc;
    fill(64);
    load_pos(0, posc, idc);
    issue(0, posc, idc);
    load_pos(1, posc, idc);
    if (IDQ) reinterpret_cast<uint32_t*>(S.ring[1] + 88)[lane] = idc;        // where iteration 0 looks for the ids of group 1
#pragma unroll 1
    for (uint32_t k = 0;; k++) {
        const int n = group_size(k);
        if (n == 0) break;                                  // (one empty cp.async group may still be pending: harmless)
        fill(32u * (k + 3));
        if (IDQ) {
            cp_async_wait<0>();                             // records of group k and ids of group k + 1 have landed
            __syncwarp();
            const bool v1 = lane < group_size(k + 1);
            posc = v1 ? S.queue[(32u * (k + 1) + (uint32_t)lane) & (V6_QCAP - 1)] : 0u;
            idc = v1 ? reinterpret_cast<const uint32_t*>(S.ring[(k + 1) & 1] + 88)[lane] : 0u;
            __syncwarp();                                   // (lanes 24-31 write their record slot over the id words)
            // records one group ahead, into the buffer the ids were read from (not committed yet: with the ids below)
            float4* Rn = S.ring[(k + 1) & 1];
            if (v1) {
                const SplatRec* r = rec + idc;
                cp_async16(&Rn[lane], &r->q0);
                cp_async16(&Rn[32 + lane], &r->q1);
                cp_async4(&Rn[64 + lane].z, &r->q2);
                *reinterpret_cast<uint2*>(&Rn[64 + lane]) = make_uint2(idc, posc);
            }
        } else {
            uint32_t posn, idn;
            load_pos(k + 2, posn, idn);                     // ids two groups ahead
            issue(k + 1, posc, idc);                        // records one group ahead
            posc = posn; idc = idn;
            cp_async_wait<1>();                             // my records of group k have landed
            __syncwarp();
        }
        float4* const R = S.ring[k & 1];
        {
            // lane-indexed -> pair records, in place: every lane takes its entry into registers, then writes it back
            // interleaved with its neighbour's.  Slots behind the group's last entry get a record that blends nothing
            // and holds no NaN (an odd group reads one of them).
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a, e = make_float4(0.f, __int_as_float(0x7fffffff), 0.f, 0.f);
            if (lane < n) { a = R[lane]; c = R[32 + lane]; e = R[64 + lane]; }
            __syncwarp();
            float* w = reinterpret_cast<float*>(R) + (lane >> 1) * (int)(sizeof(PairRec) / 4) + (lane & 1);
            w[0] = a.x; w[2] = a.y; w[4] = a.z; w[6] = a.w; w[8] = c.x; w[10] = c.y; w[12] = c.z; w[14] = c.w;
            w[16] = e.z; w[18] = e.y;
            reinterpret_cast<float*>(R + 80)[lane] = e.x;
            __syncwarp();
        }
        if (IDQ) {
            // ids two groups ahead: list -> the free tail of this buffer, committed together with the records issued above
            if (lane < group_size(k + 2)) {
                const uint32_t p2 = S.queue[(32u * (k + 2) + (uint32_t)lane) & (V6_QCAP - 1)];
                cp_async4(reinterpret_cast<uint32_t*>(R + 88) + lane, tvals + p2);
            }
            cp_async_commit();
        }
        const PairRec* const P = reinterpret_cast<const PairRec*>(R);
        const float* const Pw = reinterpret_cast<const float*>(R);
#pragma unroll 1
        for (int jb = 0; jb < n; jb += SLOTS) {
            const int cnt = min(SLOTS, n - jb);             // entries of this flush group (warp-uniform)
            if (FULL && cnt == SLOTS) {
                // a full flush group (all but a warp's last one): straight-line code, the eight pairs' loads and heads can be
                // scheduled across each other
#pragma unroll
                for (int c = 0; c < SLOTS; c += 2) {
                    float w0, at0, w1, at1;
                    v7_pair(P[(jb + c) >> 1], pxf, pyf, last_contributor, dp0, dp1, dp2, T, Sd, w0, at0, w1, at1);
                    float* const sp = ((c >> 3) ? a_hi : a_lo) + (c & 7) * GS;
                    *reinterpret_cast<float2*>(sp) = make_float2(w0, at0);
                    *reinterpret_cast<float2*>(sp + GS) = make_float2(w1, at1);
                }
            } else {
#pragma unroll
                for (int c = 0; c < SLOTS; c += 2) {
                    if (c < cnt) {
                        float w0, at0, w1, at1;
                        v7_pair(P[(jb + c) >> 1], pxf, pyf, last_contributor, dp0, dp1, dp2, T, Sd, w0, at0, w1, at1);
                        float* const sp = ((c >> 3) ? a_hi : a_lo) + (c & 7) * GS;
                        *reinterpret_cast<float2*>(sp) = make_float2(w0, at0);
                        *reinterpret_cast<float2*>(sp + GS) = make_float2(w1, at1);
                    }
                }
            }
            v6_flush<SLOTS, PRED>(S.a, S.b2, SM.b1, [&](int j, float2& xy, uint32_t& id) {
                const float* q = Pw + (j >> 1) * (int)(sizeof(PairRec) / 4) + (j & 1);
                xy = make_float2(q[0], q[2]);
                id = __float_as_uint(Pw[320 + j]);
            }, jb, cnt, lane, X0s, Y0s, accum);
        }
        __syncwarp();                  // everyone has left ring buffer k & 1 before the gather of group k + 2 overwrites it
    }
    cp_async_wait<0>();
}

// ------------------------------------------------------------------------------------------ launchers
int launch_blend_forward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                         float* out_color, cudaStream_t s)
{
    const Options& o = options();
    const int gyv = (H + GSR_TILE - 1) / GSR_TILE;
    const dim3 grid((W + GSR_TILE - 1) / GSR_TILE, gyv * V);
    const uint32_t pw = (uint32_t)blend_plane_words(R, W, gyv * V * GSR_TILE);
    const uint32_t* lpt = o.lpt ? im.tile_order : nullptr;
    // comparison arms (test-only library): never taken unless a test or benchmark asked for them
    if (o.blend_stats == 1) {
        if (V > 1 || !g_arms.blend_stats) return GSR_EINVAL;
        int rc = g_arms.blend_stats(R, W, H, g, b, im, s);
        if (rc) return rc;
    }
    if (o.blend_arm == 1) {
        if (V > 1 || !g_arms.blend_forward_plain) return GSR_EINVAL;
        return g_arms.blend_forward_plain(R, W, H, bg, g, b, im, out_color, s);
    }
    if (o.blend_arm == 2) {
        if (!g_arms.blend_forward_bulk) return GSR_EINVAL;
        return g_arms.blend_forward_bulk(R, W, H, V, bg, g, b, im, out_color, s);
    }
    if (o.fwd_variant == 4) {
#define COMMA ,
#define GSR_F4(MINB)                                                                                                        \
    GSR_LAUNCH((blend_fwd_v4_kernel<MINB>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib, \
               out_color, b.scratch, pw, gyv, lpt)
        if (o.fwd_minb == 6) GSR_F4(6);
        else if (o.fwd_minb == 4) GSR_F4(4);
        else if (o.fwd_minb == 5) GSR_F4(5);
        else GSR_F4(4 COMMA 2);                            // 42, the default: four CTAs per SM, pair loop unrolled twice
#undef GSR_F4
        return GSR_OK;
    }
    if (o.fwd_minb == 5)
        GSR_LAUNCH((blend_fwd_v3_kernel<4, 5>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    else if (o.fwd_minb == 52)
        GSR_LAUNCH((blend_fwd_v3_kernel<2, 5>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    else if (o.fwd_minb == 3)
        GSR_LAUNCH((blend_fwd_v3_kernel<4, 3>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    else
        GSR_LAUNCH((blend_fwd_v3_kernel<4, 4>), grid, BL_THREADS, 0, s, im.ranges, b.vals, g.rec, W, H, bg, im.final_T, im.n_contrib,
                   out_color, b.scratch, pw, gyv, lpt);
    return GSR_OK;
}

int launch_blend_backward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                          const float* dL_dpix, float* accum, cudaStream_t s)
{
    const Options& o = options();
    const int gyv = (H + GSR_TILE - 1) / GSR_TILE;
    const dim3 grid((W + GSR_TILE - 1) / GSR_TILE, gyv * V);
    const uint32_t pw = (uint32_t)blend_plane_words(R, W, gyv * V * GSR_TILE);
    const uint32_t* lpt = o.lpt ? im.tile_order : nullptr;
    if (o.blend_arm == 1) {        // plain kernel of the test-only library (does not need the forward's classification words)
        if (V > 1 || !g_arms.blend_backward_plain) return GSR_EINVAL;
        return g_arms.blend_backward_plain(R, W, H, bg, g, b, im, dL_dpix, accum, s);
    }
    DeviceCtx* ctx = nullptr;
    int rc;
    if ((rc = device_ctx(&ctx))) return rc;
    rc = ctx_once(ctx, 2u, [] {    // per device: opt in to the kernels' dynamic shared memory
        GSR_CUDA(cudaFuncSetAttribute((blend_bwd_v7_kernel<3, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV7)));
        GSR_CUDA(cudaFuncSetAttribute((blend_bwd_v7_kernel<3, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV7)));
        GSR_CUDA(cudaFuncSetAttribute((blend_bwd_v7_kernel<3, true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV7)));
        GSR_CUDA(cudaFuncSetAttribute((blend_bwd_v7_kernel<3, true, true, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV7)));
        GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v6_kernel<16, 2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV6<16>)));
        GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v6_kernel<16, 4, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV6<16>)));
        GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v6_kernel<8, 2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV6<8>)));
        GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v5_kernel<2, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV5<2>)));
        GSR_CUDA(cudaFuncSetAttribute(blend_bwd_v5_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmemV5<2>)));
        return (int)GSR_OK;
    });
    if (rc) return rc;
    if (o.bwd_variant == 5) {
        if (o.bwd_minb == 3)
            GSR_LAUNCH((blend_bwd_v5_kernel<2, 3>), grid, BL_THREADS, sizeof(BwdSmemV5<2>), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H,
                       bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        else
            GSR_LAUNCH((blend_bwd_v5_kernel<2, 4>), grid, BL_THREADS, sizeof(BwdSmemV5<2>), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H,
                       bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
    } else {
#define GSR_V6(SLOTS, TRIP, MINB)                                                                                                  \
    GSR_LAUNCH((blend_bwd_v6_kernel<SLOTS, TRIP, MINB>), grid, BL_THREADS, sizeof(BwdSmemV6<SLOTS>), s, im.ranges, b.vals, g.rec,   \
               b.scratch, pw, W, H, bg, im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt)
        // bwd_variant 6: 16 reduction rows, 3 CTAs per SM;  68: 8 rows, 4 CTAs per SM, trips of 2 entries;
        // 64: 16 rows, trips of 4
        if (o.bwd_variant == 7) {                 // v7 with the quad commit and the ids through cp.async: the product kernel
            GSR_LAUNCH((blend_bwd_v7_kernel<3, true, true, true>), grid, BL_THREADS, sizeof(BwdSmemV7), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H, bg,
                       im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        } else if (o.bwd_variant == 73) {         // ids loaded into a register two groups ahead (2.0235 vs 2.011 ms per view in the step)
            GSR_LAUNCH((blend_bwd_v7_kernel<3, true, true>), grid, BL_THREADS, sizeof(BwdSmemV7), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H, bg,
                       im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        } else if (o.bwd_variant == 72) {         // v7 with the branchy commit of v6 (2.039 vs 2.009 ms per view in the step)
            GSR_LAUNCH((blend_bwd_v7_kernel<3, true>), grid, BL_THREADS, sizeof(BwdSmemV7), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H, bg,
                       im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        } else if (o.bwd_variant == 70) {
            GSR_LAUNCH((blend_bwd_v7_kernel<3, false>), grid, BL_THREADS, sizeof(BwdSmemV7), s, im.ranges, b.vals, g.rec, b.scratch, pw, W, H, bg,
                       im.final_T, im.n_contrib, dL_dpix, accum, gyv, lpt);
        } else if (o.bwd_variant == 68) GSR_V6(8, 2, 4);
        else if (o.bwd_variant == 64) GSR_V6(16, 4, 3);
        else GSR_V6(16, 2, 3);
#undef GSR_V6
    }
    return GSR_OK;
}

}  // namespace gsr
