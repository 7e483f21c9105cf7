// sort.cu - binning: depth ordering of Gaussians, instance emission, stable split by tile, tile ranges.
//
// Replaces (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu):
//   cub::DeviceScan::InclusiveSum            :277      -> scan_* kernels (over the depth-ordered tile counts)
//   duplicateWithKeys                        :70-111   -> emit_instances_kernel
//   cub::DeviceRadixSort::SortPairs (u64)    :303-308  -> radix_* kernels, see below
//   identifyTileRanges (+ cudaMemset)        :116-138,310-318 -> tile_ranges_kernel
//
// The reference sorts R (tile<<32 | depth) 64-bit keys with 6 radix passes over 12 B pairs (~152 B per
// instance of HBM traffic).  We produce the IDENTICAL sequence with far less traffic by trimming the
// key to the bits that are live at each stage:
//   (1) stable LSD radix sort of the P Gaussians by their 32 depth bits (4 passes over P items),
//   (2) instances are emitted in that order (still row-major over each Gaussian's tile rect),
//   (3) a stable LSD radix split by tile id only (ceil(tile_bits/8) passes over R items, 2 at 1080p).
// A stable sort by (tile) of a sequence already ordered by (depth, id) is ordered by (tile, depth, id),
// which is exactly what a stable sort on tile<<32|depth of the id-major emission gives.
//
// Each radix pass is three launches: per-block digit histogram, per-digit scan over blocks, ranked
// scatter (warp match-any ranking keeps the pass stable).
#include "gsr_common.cuh"

namespace gsr {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_IPT = 16;                         // items per thread
constexpr int RS_TILE = RS_THREADS * RS_IPT;       // items per block (4096)
constexpr int RS_NB = 256;                         // max bins (8-bit digits)

// ------------------------------------------------------------------ radix pass: histogram
__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, int shift,
                                                                int bits, uint32_t* __restrict__ hist, int nblocks)
{
    __shared__ uint32_t sh[RS_NB];
    const uint32_t nb = 1u << bits, mask = nb - 1;
    for (uint32_t i = threadIdx.x; i < nb; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * RS_TILE;
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = base + i * RS_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&sh[(keys[idx] >> shift) & mask], 1u);
    }
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < nb; d += RS_THREADS) hist[(size_t)d * nblocks + blockIdx.x] = sh[d];
}

// ------------------------------------------------------------------ radix pass: scan each digit row over blocks
// one CTA per digit; hist[d][*] becomes the exclusive prefix over blocks, totals[d] the row sum
__global__ void __launch_bounds__(RS_THREADS) radix_rowscan_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ totals,
                                                                   int nblocks)
{
    __shared__ uint32_t warp_sum[RS_WARPS];
    uint32_t* row = hist + (size_t)blockIdx.x * nblocks;
    const int per = (nblocks + RS_THREADS - 1) / RS_THREADS;
    const int j0 = threadIdx.x * per, j1 = min(nblocks, j0 + per);
    uint32_t s = 0;
    for (int j = j0; j < j1; j++) s += row[j];
    // block exclusive scan of s
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sum[w] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (int i = 0; i < w; i++) wbase += warp_sum[i];
    uint32_t run = wbase + inc - s;
    for (int j = j0; j < j1; j++) { uint32_t v = row[j]; row[j] = run; run += v; }
    if (threadIdx.x == RS_THREADS - 1) totals[blockIdx.x] = wbase + inc;
}

// ------------------------------------------------------------------ radix pass: ranked scatter
// vals_in == nullptr means "identity" (value = item index).  keys_out may be nullptr.
// One warp match-any per item gives its stable rank inside the warp's run; per-warp digit counts are
// scanned across the CTA; the 4096 items are first placed in digit order in shared memory and then
// written out so that consecutive threads write consecutive addresses inside each digit run
// (full 32-byte sectors instead of 4-byte scattered stores).
__global__ void __launch_bounds__(RS_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                   const uint32_t* __restrict__ vals_in,
                                                                   uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                   uint32_t n, int shift, int bits,
                                                                   const uint32_t* __restrict__ hist, const uint32_t* __restrict__ totals,
                                                                   int nblocks)
{
    __shared__ uint32_t wcount[RS_WARPS][RS_NB];
    __shared__ uint32_t lbase[RS_NB];       // CTA-local exclusive offset of each digit
    __shared__ uint32_t gdelta[RS_NB];      // global position of local slot lp of digit d = gdelta[d] + lp
    __shared__ uint32_t warp_sum[RS_WARPS];
    __shared__ uint32_t skey[RS_TILE];
    __shared__ uint32_t sval[RS_TILE];
    const uint32_t nb = 1u << bits, mask = nb - 1;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t t = threadIdx.x;

    for (uint32_t i = t; i < RS_WARPS * RS_NB; i += RS_THREADS) (&wcount[0][0])[i] = 0;
    // global digit bases for this CTA: exclusive scan of totals + this CTA's row prefix
    uint32_t gscan = 0;
    {
        const uint32_t v = t < nb ? totals[t] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) warp_sum[w] = inc;
        __syncthreads();
        uint32_t wb = 0;
        for (int i = 0; i < w; i++) wb += warp_sum[i];
        if (t < nb) gscan = wb + inc - v + hist[(size_t)t * nblocks + blockIdx.x];
    }
    __syncthreads();

    // phase 1: each warp walks its contiguous run of 32*IPT items in order (i, lane)
    const uint32_t wbase_idx = blockIdx.x * RS_TILE + w * (32 * RS_IPT);
    uint32_t key[RS_IPT];
    uint32_t loff[RS_IPT];      // rank of the item among the warp's items with the same digit
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        key[i] = idx < n ? keys_in[idx] : 0xFFFFFFFFu;
    }
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        const bool valid = idx < n;
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
        loff[i] = 0;
        if (valid) {
            const uint32_t d = (key[i] >> shift) & mask;
            // lanes holding the same digit, from 8 ballots (match.any is microcoded and far slower)
            uint32_t peers = vmask;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                const bool bit = (d >> b) & 1u;
                const uint32_t bal = __ballot_sync(vmask, bit);
                peers &= bit ? bal : ~bal;
            }
            const uint32_t rank = __popc(peers & ((1u << lane) - 1));
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if (lane == leader) { old = wcount[w][d]; wcount[w][d] = old + __popc(peers); }
            old = __shfl_sync(peers, old, leader);
            loff[i] = old + rank;
        }
        __syncwarp();
    }
    __syncthreads();
    // phase 2: per digit, exclusive prefix over warps; then exclusive scan over digits
    {
        uint32_t tot = 0;
        if (t < nb) {
#pragma unroll
            for (int i = 0; i < RS_WARPS; i++) { uint32_t c = wcount[i][t]; wcount[i][t] = tot; tot += c; }
        }
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) warp_sum[w] = inc;
        __syncthreads();
        uint32_t wb = 0;
        for (int i = 0; i < w; i++) wb += warp_sum[i];
        if (t < nb) { const uint32_t lb = wb + inc - tot; lbase[t] = lb; gdelta[t] = gscan - lb; }
    }
    __syncthreads();
    // phase 3: place items in digit order in shared memory
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        if (idx < n) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t lp = lbase[d] + wcount[w][d] + loff[i];
            skey[lp] = key[i];
            sval[lp] = vals_in ? vals_in[idx] : idx;
        }
    }
    __syncthreads();
    // phase 4: coalesced write-out
    const uint32_t n_blk = min((uint32_t)RS_TILE, n - blockIdx.x * RS_TILE);
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t lp = i * RS_THREADS + t;
        if (lp < n_blk) {
            const uint32_t k = skey[lp];
            const uint32_t gpos = gdelta[(k >> shift) & mask] + lp;
            if (keys_out) keys_out[gpos] = k;
            vals_out[gpos] = sval[lp];
        }
    }
}

// one LSD pass over (keys, vals); hist needs 256 * nblocks u32, totals 256 u32
static int radix_pass(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t n,
                      int shift, int bits, uint32_t* hist, uint32_t* totals, cudaStream_t s)
{
    const int nblocks = cdiv(n, RS_TILE);
    GSR_LAUNCH(radix_hist_kernel, nblocks, RS_THREADS, 0, s, keys_in, n, shift, bits, hist, nblocks);
    GSR_LAUNCH(radix_rowscan_kernel, 1 << bits, RS_THREADS, 0, s, hist, totals, nblocks);
    GSR_LAUNCH(radix_scatter_kernel, nblocks, RS_THREADS, 0, s, keys_in, vals_in, keys_out, vals_out, n, shift, bits, hist,
               totals, nblocks);
    return GSR_OK;
}

// stable LSD sort of n (key, value) pairs on key bits [0, 8*npass); values start as identity.
// K0, K1, V0: n-word ping-pong buffers; result values land in vals_out (keys are not kept).
size_t radix_hist_words(uint32_t n) { return align_up((size_t)RS_NB * cdiv(n > 0 ? n : 1, RS_TILE), 64) + RS_NB; }
int radix_sort_by_key_u32(const uint32_t* keys, uint32_t* K0, uint32_t* K1, uint32_t* V0, uint32_t* vals_out, uint32_t n, int npass,
                          uint32_t* hist_and_totals, cudaStream_t s)
{
    uint32_t* hist = hist_and_totals;
    uint32_t* totals = hist + align_up((size_t)RS_NB * cdiv(n > 0 ? n : 1, RS_TILE), 64);
    // choose the value ping-pong so that the LAST pass writes vals_out
    uint32_t* Vb[2] = {(npass & 1) ? vals_out : V0, (npass & 1) ? V0 : vals_out};
    uint32_t* Kb[2] = {K0, K1};
    const uint32_t* kin = keys;
    const uint32_t* vin = nullptr;
    for (int p = 0; p < npass; p++) {
        const bool last = p == npass - 1;
        int rc = radix_pass(kin, vin, last ? nullptr : Kb[p & 1], Vb[p & 1], n, 8 * p, 8, hist, totals, s);
        if (rc) return rc;
        kin = Kb[p & 1]; vin = Vb[p & 1];
    }
    return GSR_OK;
}

// ------------------------------------------------------------------ exclusive scan of tile counts in depth order
constexpr int SC_TILE = 4096;   // items per block, 256 threads x 16

__global__ void __launch_bounds__(256) scan_reduce_kernel(const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles,
                                                          uint32_t n, uint32_t* __restrict__ block_sums)
{
    __shared__ uint32_t ws[8];
    uint32_t s = 0;
    const uint32_t base = blockIdx.x * SC_TILE;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t idx = base + i * 256 + threadIdx.x;
        if (idx < n) s += tiles[order[idx]];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < 8; i++) t += ws[i];
        block_sums[blockIdx.x] = t;
    }
}

// single CTA: exclusive scan of block_sums in place; writes the grand total to *total and *offs_end
__global__ void __launch_bounds__(1024) scan_spine_kernel(uint32_t* __restrict__ block_sums, int nb, uint32_t* __restrict__ total,
                                                          uint32_t* __restrict__ offs_end)
{
    __shared__ uint32_t ws[32];
    __shared__ uint32_t carry_s, chunk_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int base = 0; base < nb; base += 1024) {
        const int j = base + threadIdx.x;
        const uint32_t v = j < nb ? block_sums[j] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) ws[w] = inc;
        __syncthreads();
        if (w == 0) {
            const uint32_t x = ws[lane];
            uint32_t xi = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, xi, o);
                if (lane >= o) xi += t;
            }
            ws[lane] = xi - x;               // exclusive warp bases
            if (lane == 31) chunk_s = xi;    // chunk total
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        if (j < nb) block_sums[j] = carry + ws[w] + inc - v;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + chunk_s;
        __syncthreads();
    }
    if (threadIdx.x == 0) { *total = carry_s; *offs_end = carry_s; }
}

__global__ void __launch_bounds__(256) scan_apply_kernel(const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles,
                                                         uint32_t n, const uint32_t* __restrict__ block_sums,
                                                         uint32_t* __restrict__ offs)
{
    __shared__ uint32_t ws[8];
    // thread t owns 16 consecutive items so the scan order is the item order
    const uint32_t base = blockIdx.x * SC_TILE + threadIdx.x * 16;
    uint32_t v[16];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t idx = base + i;
        v[i] = idx < n ? tiles[order[idx]] : 0;
        s += v[i];
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) ws[w] = inc;
    __syncthreads();
    uint32_t run = block_sums[blockIdx.x] + inc - s;
    for (int i = 0; i < w; i++) run += ws[i];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint32_t idx = base + i;
        if (idx < n) offs[idx] = run;
        run += v[i];
    }
}

// ------------------------------------------------------------------ depth ordering
// scratch layout (u32 words): depthkey[P] tiles[P] K0[P] K1[P] V0[P] hist[256*nblk] totals[256] bsum[nscan] total[1]
struct GeomScratch {
    uint32_t *depthkey, *tiles, *K0, *K1, *V0, *hist, *totals, *bsum, *total;
};
static GeomScratch carve_geom_scratch(uint32_t* p, int P)
{
    GeomScratch g;
    const size_t Pa = align_up((size_t)P, 64);
    g.depthkey = p; p += Pa;
    g.tiles = p; p += Pa;
    g.K0 = p; p += Pa;
    g.K1 = p; p += Pa;
    g.V0 = p; p += Pa;
    g.hist = p; p += align_up((size_t)RS_NB * cdiv(P, RS_TILE), 64);
    g.totals = p; p += RS_NB;
    g.bsum = p; p += align_up((size_t)cdiv(P, SC_TILE) + 1, 64);
    g.total = p; p += 64;
    return g;
}
size_t geom_scratch_bytes(int P)
{
    const size_t Pa = align_up((size_t)(P > 0 ? P : 1), 64);
    return 4 * (5 * Pa + align_up((size_t)RS_NB * cdiv(P > 0 ? P : 1, RS_TILE), 64) + RS_NB +
                align_up((size_t)cdiv(P > 0 ? P : 1, SC_TILE) + 1, 64) + 64);
}

// pulls the depth bits and tile count out of the packed records (coalesced 16 B loads of q2)
__global__ void __launch_bounds__(256) extract_keys_kernel(int P, const SplatRec* __restrict__ rec, uint32_t* __restrict__ depthkey,
                                                           uint32_t* __restrict__ tiles)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float4 q2 = rec[idx].q2;
    depthkey[idx] = __float_as_uint(q2.y);
    tiles[idx] = __float_as_uint(q2.w);
}

int launch_depth_order(int P, const GeomView& g, cudaStream_t s)
{
    GeomScratch sc = carve_geom_scratch(g.scratch, P);
    GSR_LAUNCH(extract_keys_kernel, cdiv(P, 256), 256, 0, s, P, g.rec, sc.depthkey, sc.tiles);
    // 4 x 8-bit stable passes over the 32 depth bits; values start as identity
    int rc;
    if ((rc = radix_pass(sc.depthkey, nullptr, sc.K0, sc.V0, P, 0, 8, sc.hist, sc.totals, s))) return rc;
    if ((rc = radix_pass(sc.K0, sc.V0, sc.K1, g.order, P, 8, 8, sc.hist, sc.totals, s))) return rc;
    if ((rc = radix_pass(sc.K1, g.order, sc.K0, sc.V0, P, 16, 8, sc.hist, sc.totals, s))) return rc;
    if ((rc = radix_pass(sc.K0, sc.V0, nullptr, g.order, P, 24, 8, sc.hist, sc.totals, s))) return rc;
    // exclusive scan of tile counts in that order -> offs[0..P], total
    const int nb = cdiv(P, SC_TILE);
    GSR_LAUNCH(scan_reduce_kernel, nb, 256, 0, s, g.order, sc.tiles, (uint32_t)P, sc.bsum);
    GSR_LAUNCH(scan_spine_kernel, 1, 1024, 0, s, sc.bsum, nb, sc.total, g.offs + P);
    GSR_LAUNCH(scan_apply_kernel, nb, 256, 0, s, g.order, sc.tiles, (uint32_t)P, sc.bsum, g.offs);
    return GSR_OK;
}

// ------------------------------------------------------------------ instance emission (in depth order)
// Warp-cooperative: a warp takes 32 consecutive Gaussians of the depth order, and its lanes write the
// warp's instances with consecutive lanes on consecutive output slots (coalesced), finding the owner of
// slot t with a 5-step binary search over the lanes' offsets (shuffles).  Row-major over each
// Gaussian's tile rect, exactly the order of duplicateWithKeys (rasterizer_impl.cu:98-109).
__global__ void __launch_bounds__(256) emit_instances_kernel(int P, int gx, int gy, const uint32_t* __restrict__ order,
                                                             const uint32_t* __restrict__ offs, const SplatRec* __restrict__ rec,
                                                             uint32_t* __restrict__ tile_ids, uint32_t* __restrict__ gauss_ids)
{
    const int lane = threadIdx.x & 31;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t g = 0, off = 0xFFFFFFFFu;
    int x0 = 0, y0 = 0, wdt = 1, cnt = 0;
    if (k < P) {
        g = order[k];
        off = offs[k];
        const float4 q2 = rec[g].q2;
        const int rad = __float_as_int(q2.z);
        if (rad > 0) {
            const float4 q0 = rec[g].q0;
            // same rect as preprocess (auxiliary.h:46-56)
            x0 = min(gx, max(0, (int)((q0.x - rad) / GSR_TILE)));
            y0 = min(gy, max(0, (int)((q0.y - rad) / GSR_TILE)));
            const int x1 = min(gx, max(0, (int)((q0.x + rad + GSR_TILE - 1) / GSR_TILE)));
            const int y1 = min(gy, max(0, (int)((q0.y + rad + GSR_TILE - 1) / GSR_TILE)));
            wdt = max(1, x1 - x0);
            cnt = (x1 - x0) * (y1 - y0);
        }
    }
    const uint32_t off0 = __shfl_sync(0xffffffffu, off, 0);
    if (off0 == 0xFFFFFFFFu) return;                       // whole warp beyond P
    const uint32_t rel = off - off0;                       // lanes beyond P carry a huge value: never owners
    // total instances of this warp = end offset of the last lane
    const uint32_t my_end = (k < P) ? rel + (uint32_t)cnt : 0u;
    uint32_t total = my_end;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total = max(total, __shfl_xor_sync(0xffffffffu, total, o));
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        int lo = 0;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            const uint32_t v = __shfl_sync(0xffffffffu, rel, lo + s);     // lo + s <= 31 always
            if (v <= t) lo += s;
        }
        const uint32_t orel = __shfl_sync(0xffffffffu, rel, lo);
        const int ox0 = __shfl_sync(0xffffffffu, x0, lo), oy0 = __shfl_sync(0xffffffffu, y0, lo);
        const int ow = __shfl_sync(0xffffffffu, wdt, lo);
        const uint32_t og = __shfl_sync(0xffffffffu, g, lo);
        if (t < total) {
            const uint32_t j = t - orel;
            const uint32_t row = j / (uint32_t)ow, col = j - row * (uint32_t)ow;
            tile_ids[off0 + t] = (uint32_t)((oy0 + (int)row) * gx + ox0 + (int)col);
            gauss_ids[off0 + t] = og;
        }
    }
}

// ------------------------------------------------------------------ tile ranges from the sorted tile ids
__global__ void __launch_bounds__(256) tile_ranges_kernel(uint32_t R, const uint32_t* __restrict__ tile_sorted, uint2* __restrict__ ranges)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t cur = tile_sorted[i];
    if (i == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = tile_sorted[i - 1];
        if (cur != prev) { ranges[prev].y = i; ranges[cur].x = i; }
    }
    if (i == R - 1) ranges[cur].y = R;
}

// host: position of the highest set bit + 1 (what the reference's getHigherMsb computes, rasterizer_impl.cu:35-50)
static int tile_bits(uint32_t tiles)
{
    int b = 0;
    while (b < 32 && (tiles >> b)) b++;
    return b;
}

// scratch layout (u32 words): T0[R] G0[R] T1[R] G1[R] hist[256*nblk] totals[256]
size_t bin_scratch_bytes(long long R)
{
    const size_t Ra = align_up((size_t)(R > 0 ? R : 1), 64);
    return 4 * (4 * Ra + align_up((size_t)RS_NB * cdiv(R > 0 ? R : 1, RS_TILE), 64) + RS_NB);
}

int launch_tile_split(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s)
{
    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
    GSR_CUDA(cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * (size_t)gx * gy, s));
    if (R <= 0) return GSR_OK;
    const size_t Ra = align_up((size_t)R, 64);
    uint32_t* T[2] = {b.scratch, b.scratch + 2 * Ra};
    uint32_t* G[2] = {b.scratch + Ra, b.scratch + 3 * Ra};
    uint32_t* hist = b.scratch + 4 * Ra;
    uint32_t* totals = hist + align_up((size_t)RS_NB * cdiv(R, RS_TILE), 64);

    GSR_LAUNCH(emit_instances_kernel, cdiv(P, 256), 256, 0, s, P, gx, gy, g.order, g.offs, g.rec, T[0], G[0]);

    const int tb = tile_bits((uint32_t)(gx * gy));
    const int npass = tb <= 8 ? 1 : (tb + 7) / 8;
    int shift = 0, cur = 0;
    for (int p = 0; p < npass; p++) {
        const int bits = (tb - shift + (npass - p) - 1) / (npass - p);    // balanced digits, e.g. 13 -> 7 + 6
        const bool last = p == npass - 1;
        int rc = radix_pass(T[cur], G[cur], T[cur ^ 1], last ? b.vals : G[cur ^ 1], (uint32_t)R, shift, bits, hist, totals, s);
        if (rc) return rc;
        shift += bits; cur ^= 1;
    }
    GSR_LAUNCH(tile_ranges_kernel, cdiv(R, 256), 256, 0, s, (uint32_t)R, T[cur], im.ranges);
    return GSR_OK;
}

}  // namespace gsr
