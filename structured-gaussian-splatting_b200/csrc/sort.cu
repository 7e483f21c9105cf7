// sort.cu - binning: depth ordering of Gaussians, instance emission, stable split by tile, tile ranges.
//
// Replaces (reference: submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu):
//   cub::DeviceScan::InclusiveSum            :277      -> fused into emit_instances_kernel (decoupled look-back)
//   duplicateWithKeys                        :70-111   -> emit_instances_kernel
//   cub::DeviceRadixSort::SortPairs (u64)    :303-308  -> radix_onesweep_kernel, see below
//   identifyTileRanges (+ cudaMemset)        :116-138,310-318 -> tile_count_kernel + tile_prefix_kernel
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
// Every radix pass is ONE launch ("onesweep"): the digit histograms of all passes are known up front
// (radix_hist4_kernel for the depth bits, tile_prefix_kernel for the tile digits), and the per-block digit offsets come from a decoupled look-back over per-block status words
// instead of a separate histogram + scan launch per pass.  Blocks take their tile by atomic ticket, so
// a block only ever waits for blocks that are already running.
#include "gsr_common.cuh"

namespace gsr {

constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_IPT = 8;                          // items per thread
constexpr int RS_TILE = RS_THREADS * RS_IPT;       // items per block (4096)
constexpr int RS_NB = 256;                         // max bins (8-bit digits)

constexpr uint32_t ST_AGG = 1u << 30, ST_INC = 2u << 30, ST_VAL = (1u << 30) - 1;   // look-back status word

__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ void st_volatile(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }
__device__ __forceinline__ unsigned long long ld_volatile64(const unsigned long long* p)
{
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ void st_volatile64(unsigned long long* p, unsigned long long v)
{
    *reinterpret_cast<volatile unsigned long long*>(p) = v;
}

// exclusive scan over the block's NW warps (one value per thread); warp_sum is NW words of shared memory
template <int NW>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* warp_sum, uint32_t* total = nullptr)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    __syncthreads();                       // warp_sum may still be read from a previous use
    if (lane == 31) warp_sum[w] = inc;
    __syncthreads();
    uint32_t wb = 0, all = 0;
#pragma unroll
    for (int i = 0; i < NW; i++) { const uint32_t s = warp_sum[i]; if (i < w) wb += s; all += s; }
    if (total) *total = all;
    return wb + inc - v;
}

// ------------------------------------------------------------------ digit histograms of all four depth passes
// one read of the keys; shared-memory atomics; a few hundred blocks flush 4 x 256 counters each
__global__ void __launch_bounds__(RS_THREADS) radix_hist4_kernel(const uint32_t* __restrict__ keys, uint32_t n,
                                                                 uint32_t* __restrict__ ghist)
{
    __shared__ uint32_t sh[4 * RS_NB];
    for (int i = threadIdx.x; i < 4 * RS_NB; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t n4 = n >> 2;
    const uint4* k4 = reinterpret_cast<const uint4*>(keys);
    for (uint32_t i = blockIdx.x * RS_THREADS + threadIdx.x; i < n4; i += gridDim.x * RS_THREADS) {
        const uint4 k = k4[i];
        const uint32_t kk[4] = {k.x, k.y, k.z, k.w};
#pragma unroll
        for (int c = 0; c < 4; c++) {
            atomicAdd(&sh[kk[c] & 255u], 1u);
            atomicAdd(&sh[256 + ((kk[c] >> 8) & 255u)], 1u);
            atomicAdd(&sh[512 + ((kk[c] >> 16) & 255u)], 1u);
            atomicAdd(&sh[768 + (kk[c] >> 24)], 1u);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u)) {
        const uint32_t k = keys[(n4 << 2) + threadIdx.x];
        atomicAdd(&sh[k & 255u], 1u);
        atomicAdd(&sh[256 + ((k >> 8) & 255u)], 1u);
        atomicAdd(&sh[512 + ((k >> 16) & 255u)], 1u);
        atomicAdd(&sh[768 + (k >> 24)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 4 * RS_NB; i += RS_THREADS)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// ------------------------------------------------------------------ one LSD pass in one launch
// keys_in[n] (+ vals_in[n], nullptr = identity) -> keys_out (may be nullptr) / vals_out, stable on digit
// (key >> shift) & (2^BITS - 1).  ghist[2^BITS] = global digit counts of this pass, status[nblocks][2^BITS] and
// *ticket zeroed by the caller.
//  phase 1  each warp walks its contiguous run of 32*IPT items; 8 ballots give the lanes holding the same digit,
//           hence the item's stable rank inside the warp's run (match.any is microcoded and far slower)
//  phase 2  thread d owns digit d: prefix over warps, publish the block's count, look back over the preceding
//           blocks' status words for the exclusive prefix, publish the inclusive one
//  phase 3  items are placed in digit order in shared memory, phase 4 writes them out so that consecutive
//           threads write consecutive addresses inside each digit run (full sectors, not 4-byte scatters)
template <int BITS, int THREADS, int IPT>
__global__ void __launch_bounds__(THREADS) radix_onesweep_kernel(const uint32_t* __restrict__ keys_in,
                                                                    const uint32_t* __restrict__ vals_in,
                                                                    uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                    uint32_t n, int shift, const uint32_t* __restrict__ ghist,
                                                                    uint32_t* __restrict__ status, uint32_t* __restrict__ ticket,
                                                                    uint32_t val_base, int skip_if_single)
{
    constexpr uint32_t NB = 1u << BITS, MASK = NB - 1;
    constexpr int RS_THREADS = THREADS, RS_IPT = IPT, RS_WARPS = THREADS / 32, RS_TILE = THREADS * IPT;     // (shadow the defaults)
    static_assert(NB <= (uint32_t)THREADS, "one thread per digit in phase 2");
    // skip_if_single: every item carries the same digit (one bin of ghist holds all n) - the pass is the identity and the
    // pass before it has already written its values to vals_out (radix_onesweep_wide_kernel, vals_out_alt): nothing to do
    if (skip_if_single && __syncthreads_or(threadIdx.x < NB && ghist[threadIdx.x] == n)) return;
    __shared__ uint32_t skey[RS_TILE];
    __shared__ uint32_t sval[RS_TILE];
    uint32_t (*wcount)[NB] = reinterpret_cast<uint32_t (*)[NB]>(sval);     // [RS_WARPS][NB], dead before sval is written
    static_assert(RS_WARPS * NB <= RS_TILE, "wcount must fit in sval");
    __shared__ uint32_t lbase[NB];          // CTA-local exclusive offset of each digit
    __shared__ uint32_t gdelta[NB];         // global position of local slot lp of digit d = gdelta[d] + lp
    __shared__ uint32_t warp_sum[RS_WARPS];
    __shared__ uint32_t s_blk;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t t = threadIdx.x;

    if (t == 0) s_blk = atomicAdd(ticket, 1u);
    for (uint32_t i = t; i < RS_WARPS * NB; i += RS_THREADS) sval[i] = 0;
    __syncthreads();
    const uint32_t blk = s_blk;

    // phase 1
    const uint32_t wbase_idx = blk * RS_TILE + w * (32 * RS_IPT);
    const uint32_t lt_mask = (1u << lane) - 1;
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
        const uint32_t d = (key[i] >> shift) & MASK;
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < BITS; b++) {
            const uint32_t bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            peers &= ~(bal ^ (0u - bit));          // bit set: keep the lanes that voted 1, else those that voted 0 (one LOP3)
        }
        const uint32_t old = wcount[w][d];
        __syncwarp();
        const uint32_t cnt = __popc(peers), rank = __popc(peers & lt_mask);
        if (valid && rank == cnt - 1) wcount[w][d] = old + cnt;
        __syncwarp();
        loff[i] = old + rank;
    }
    __syncthreads();

    // phase 2
    uint32_t tot = 0, excl = 0;
    if (t < NB) {
#pragma unroll
        for (int i = 0; i < RS_WARPS; i++) { const uint32_t c = wcount[i][t]; wcount[i][t] = tot; tot += c; }
        uint32_t* mine = status + (size_t)blk * NB + t;
        if (blk == 0) st_volatile(mine, ST_INC | tot);
        else {
            st_volatile(mine, ST_AGG | tot);
            // four predecessors per round trip to L2; consumed nearest first
            bool found = false;
            for (int k = (int)blk - 1; !found; k -= 4) {
                uint32_t sv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) sv[j] = k - j >= 0 ? ld_volatile(status + (size_t)(k - j) * NB + t) : ST_INC;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (found) break;
                    uint32_t v = sv[j];
                    while ((v >> 30) == 0) v = ld_volatile(status + (size_t)(k - j) * NB + t);
                    excl += v & ST_VAL;
                    found = (v & ST_INC) != 0;
                }
            }
            st_volatile(mine, ST_INC | (excl + tot));
        }
    }
    const uint32_t gbase = block_excl_scan<RS_WARPS>(t < NB ? ghist[t] : 0u, warp_sum);
    const uint32_t lb = block_excl_scan<RS_WARPS>(tot, warp_sum);
    if (t < NB) { lbase[t] = lb; gdelta[t] = gbase + excl - lb; }
    __syncthreads();

    // phase 3: place items in digit order in shared memory (all value loads are issued before the first store;
    // the slots are computed before wcount's storage is reused for the values)
    uint32_t val[RS_IPT];
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        val[i] = (vals_in && idx < n) ? vals_in[idx] : idx + val_base;     // identity values of a segment start at its offset
        const uint32_t d = (key[i] >> shift) & MASK;
        loff[i] += lbase[d] + wcount[w][d];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        if (idx < n) { skey[loff[i]] = key[i]; sval[loff[i]] = val[i]; }
    }
    __syncthreads();
    // phase 4: coalesced write-out
    const uint32_t n_blk = min((uint32_t)RS_TILE, n - blk * RS_TILE);
#pragma unroll
    for (int i = 0; i < RS_IPT; i++) {
        const uint32_t lp = i * RS_THREADS + t;
        if (lp < n_blk) {
            const uint32_t k = skey[lp];
            const uint32_t gpos = gdelta[(k >> shift) & MASK] + lp;
            if (keys_out) keys_out[gpos] = k;
            vals_out[gpos] = sval[lp];
        }
    }
}

// one pass; bits in 1..8
static int radix_pass(const uint32_t* keys_in, const uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t n,
                      int shift, int bits, const uint32_t* ghist, uint32_t* status, uint32_t* ticket, cudaStream_t s,
                      uint32_t val_base = 0, int shape = 0, int skip = 0)
{
    // Block shape (both 4096 items per block): 0 = 512 threads x 8 items, 1 = 256 threads x 16 items.  Measured at C3: the
    // R-sized tile-split passes gain from four independent blocks per SM instead of two (0.299 -> 0.260 ms for the stage:
    // a block stalled in its look-back or at a barrier leaves three others issuing), the P-sized depth passes do not
    // (0.148 vs 0.151 ms: they are bound by the look-back chain itself).
    if (shape == 2) {
        // 256 threads x 8 items: 2048 items per block, twice as many (shorter-lived) blocks, eight resident per SM
#define GSR_PASS_C(B)                                                                                                    \
    case B:                                                                                                              \
        GSR_LAUNCH((radix_onesweep_kernel<B, 256, 8>), cdiv(n, 2048), 256, 0, s, keys_in, vals_in, keys_out, vals_out, n, shift, ghist, \
                   status, ticket, val_base, skip);                                                                      \
        return GSR_OK;
        switch (bits) {
            GSR_PASS_C(1) GSR_PASS_C(2) GSR_PASS_C(3) GSR_PASS_C(4) GSR_PASS_C(5) GSR_PASS_C(6) GSR_PASS_C(7) GSR_PASS_C(8)
        default: return GSR_EINVAL;
        }
#undef GSR_PASS_C
    }
    const int nblocks = cdiv(n, RS_TILE);
    if (shape == 1) {
#define GSR_PASS_B(B)                                                                                                    \
    case B:                                                                                                              \
        GSR_LAUNCH((radix_onesweep_kernel<B, 256, 16>), nblocks, 256, 0, s, keys_in, vals_in, keys_out, vals_out, n, shift, ghist, \
                   status, ticket, val_base, skip);                                                                      \
        return GSR_OK;
        switch (bits) {
            GSR_PASS_B(1) GSR_PASS_B(2) GSR_PASS_B(3) GSR_PASS_B(4) GSR_PASS_B(5) GSR_PASS_B(6) GSR_PASS_B(7) GSR_PASS_B(8)
        default: return GSR_EINVAL;
        }
#undef GSR_PASS_B
    }
#define GSR_PASS(B)                                                                                                      \
    case B:                                                                                                              \
        GSR_LAUNCH((radix_onesweep_kernel<B, RS_THREADS, RS_IPT>), nblocks, RS_THREADS, 0, s, keys_in, vals_in, keys_out, vals_out, n, shift, ghist, \
                   status, ticket, val_base, skip);                                                                      \
        break;
    switch (bits) {
        GSR_PASS(1) GSR_PASS(2) GSR_PASS(3) GSR_PASS(4) GSR_PASS(5) GSR_PASS(6) GSR_PASS(7) GSR_PASS(8)
    default: return GSR_EINVAL;
    }
#undef GSR_PASS
    return GSR_OK;
}

// ------------------------------------------------------------------ depth sort in three 9-bit passes (+ a conditional fourth)
// The depth keys preprocess writes under radix_depth = 3 are float bits minus the bits of the near plane (0.2f), culled
// Gaussians 0: monotone in the depth, so the order of the visible Gaussians is the one the raw float bits give, and every
// depth below 0.2 * 2^16 = 13107 leaves the five top bits zero.  The sort then runs digits of 9 + 9 + 9 + 5 bits: the
// histogram kernel sees (on the device) whether the fourth digit is the same for all n items; if so the third pass
// writes its values straight to the sort's output and the fourth pass returns at once - three P-sized passes instead of
// four.  Deeper scenes take all four passes; the result is the same stable order either way.
constexpr int RW_BITS = 9, RW_NB = 1 << RW_BITS, RW_TOP_BITS = 32 - 3 * RW_BITS, RW_TOP_NB = 1 << RW_TOP_BITS;
constexpr int RW_HIST = 3 * RW_NB + RW_TOP_NB;        // 1568 counters
constexpr int RW_CTL_HEAD = RW_HIST + 64;             // + tickets; even, so the 64-bit status words behind it stay aligned

__global__ void __launch_bounds__(RS_THREADS) radix_hist_wide_kernel(const uint32_t* __restrict__ keys, uint32_t n,
                                                                     uint32_t* __restrict__ ghist)
{
    __shared__ uint32_t sh[RW_HIST];
    for (int i = threadIdx.x; i < RW_HIST; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t n4 = n >> 2;
    const int lane = threadIdx.x & 31;
    const uint4* k4 = reinterpret_cast<const uint4*>(keys);
    // (warp-uniform trip count: the two high digits are nearly constant, so a warp whose lanes agree adds once)
    for (uint32_t base = blockIdx.x * RS_THREADS; base < n4; base += gridDim.x * RS_THREADS) {
        const uint32_t i = base + threadIdx.x;
        const bool valid = i < n4;
        const uint4 k = valid ? k4[i] : make_uint4(0u, 0u, 0u, 0u);
        const uint32_t kk[4] = {k.x, k.y, k.z, k.w};
#pragma unroll
        for (int c = 0; c < 4; c++) {
            if (valid) {
                atomicAdd(&sh[kk[c] & (RW_NB - 1)], 1u);
                atomicAdd(&sh[RW_NB + ((kk[c] >> RW_BITS) & (RW_NB - 1))], 1u);
            }
            const uint32_t hi = kk[c] >> (2 * RW_BITS);                    // third and top digit together (14 bits)
            const uint32_t hi0 = __shfl_sync(0xffffffffu, hi, 0);
            if (__all_sync(0xffffffffu, valid && hi == hi0)) {
                if (lane == 0) {
                    atomicAdd(&sh[2 * RW_NB + (hi0 & (RW_NB - 1))], 32u);
                    atomicAdd(&sh[3 * RW_NB + (hi0 >> RW_BITS)], 32u);
                }
            } else if (valid) {
                atomicAdd(&sh[2 * RW_NB + (hi & (RW_NB - 1))], 1u);
                atomicAdd(&sh[3 * RW_NB + (hi >> RW_BITS)], 1u);
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u)) {
        const uint32_t k = keys[(n4 << 2) + threadIdx.x];
        atomicAdd(&sh[k & (RW_NB - 1)], 1u);
        atomicAdd(&sh[RW_NB + ((k >> RW_BITS) & (RW_NB - 1))], 1u);
        atomicAdd(&sh[2 * RW_NB + ((k >> (2 * RW_BITS)) & (RW_NB - 1))], 1u);
        atomicAdd(&sh[3 * RW_NB + (k >> (3 * RW_BITS))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RW_HIST; i += RS_THREADS)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// radix_onesweep_kernel for digits of 2 x THREADS bins: thread t owns the digits 2t and 2t + 1, whose look-back status
// words travel as ONE 64-bit word (low half: digit 2t), published and read with single 8-byte accesses, so the look-back
// makes as many round trips as the narrow kernel's.  vals_out_alt / next_hist: when one bin of next_hist[0 .. next_nb)
// holds all n items the NEXT pass is the identity; this pass then writes its values to vals_out_alt (the sort's output)
// and no keys.
template <int BITS, int THREADS, int IPT>
__global__ void __launch_bounds__(THREADS) radix_onesweep_wide_kernel(const uint32_t* __restrict__ keys_in,
                                                                         const uint32_t* __restrict__ vals_in,
                                                                         uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                                         uint32_t* __restrict__ vals_out_alt,
                                                                         uint32_t n, int shift, const uint32_t* __restrict__ ghist,
                                                                         const uint32_t* __restrict__ next_hist, int next_nb,
                                                                         unsigned long long* __restrict__ status,
                                                                         uint32_t* __restrict__ ticket, uint32_t val_base)
{
    constexpr uint32_t NB = 1u << BITS, MASK = NB - 1;
    constexpr int WARPS = THREADS / 32, TILE = THREADS * IPT;
    static_assert(NB == 2u * THREADS, "two digits per thread in phase 2");
    __shared__ uint32_t skey[TILE];
    __shared__ __align__(8) uint32_t sval[TILE];
    uint32_t (*wcount)[NB] = reinterpret_cast<uint32_t (*)[NB]>(sval);     // [WARPS][NB], dead before sval is written
    static_assert(WARPS * NB <= TILE, "wcount must fit in sval");
    __shared__ uint32_t lbase[NB];          // CTA-local exclusive offset of each digit
    __shared__ uint32_t gdelta[NB];         // global position of local slot lp of digit d = gdelta[d] + lp
    __shared__ uint32_t warp_sum[WARPS];
    __shared__ uint32_t s_blk;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t t = threadIdx.x;

    if (t == 0) s_blk = atomicAdd(ticket, 1u);
    for (uint32_t i = t; i < WARPS * NB; i += THREADS) sval[i] = 0;
    const bool next_single = next_hist != nullptr && (int)t < next_nb && next_hist[t] == n;
    const bool alt = __syncthreads_or(next_single) != 0;
    const uint32_t blk = s_blk;

    // phase 1: stable rank of every item among the warp's items with the same digit (BITS ballots)
    const uint32_t wbase_idx = blk * TILE + w * (32 * IPT);
    const uint32_t lt_mask = (1u << lane) - 1;
    uint32_t key[IPT];
    uint32_t loff[IPT];
#pragma unroll
    for (int i = 0; i < IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        key[i] = idx < n ? keys_in[idx] : 0xFFFFFFFFu;
    }
#pragma unroll
    for (int i = 0; i < IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        const bool valid = idx < n;
        const uint32_t d = (key[i] >> shift) & MASK;
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < BITS; b++) {
            const uint32_t bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            peers &= ~(bal ^ (0u - bit));
        }
        const uint32_t old = wcount[w][d];
        __syncwarp();
        const uint32_t cnt = __popc(peers), rank = __popc(peers & lt_mask);
        if (valid && rank == cnt - 1) wcount[w][d] = old + cnt;
        __syncwarp();
        loff[i] = old + rank;
    }
    __syncthreads();

    // phase 2: digits 2t, 2t + 1
    uint32_t tot0 = 0, tot1 = 0, excl0 = 0, excl1 = 0;
    {
#pragma unroll
        for (int i = 0; i < WARPS; i++) {
            uint2* wc = reinterpret_cast<uint2*>(&wcount[i][2 * t]);
            const uint2 c = *wc;
            *wc = make_uint2(tot0, tot1);
            tot0 += c.x; tot1 += c.y;
        }
        auto pack = [](uint32_t flag, uint32_t a, uint32_t b) -> unsigned long long {
            return ((unsigned long long)(flag | b) << 32) | (unsigned long long)(flag | a);
        };
        unsigned long long* mine = status + (size_t)blk * THREADS + t;
        if (blk == 0) st_volatile64(mine, pack(ST_INC, tot0, tot1));
        else {
            st_volatile64(mine, pack(ST_AGG, tot0, tot1));
            bool found = false;
            for (int k = (int)blk - 1; !found; k -= 4) {
                unsigned long long sv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) sv[j] = k - j >= 0 ? ld_volatile64(status + (size_t)(k - j) * THREADS + t) : pack(ST_INC, 0u, 0u);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (found) break;
                    unsigned long long v = sv[j];
                    while (((uint32_t)v >> 30) == 0) v = ld_volatile64(status + (size_t)(k - j) * THREADS + t);
                    excl0 += (uint32_t)v & ST_VAL;
                    excl1 += (uint32_t)(v >> 32) & ST_VAL;
                    found = ((uint32_t)v & ST_INC) != 0;
                }
            }
            st_volatile64(mine, pack(ST_INC, excl0 + tot0, excl1 + tot1));
        }
    }
    const uint2 gh = *reinterpret_cast<const uint2*>(ghist + 2 * t);
    const uint32_t gbase = block_excl_scan<WARPS>(gh.x + gh.y, warp_sum);
    const uint32_t lb = block_excl_scan<WARPS>(tot0 + tot1, warp_sum);
    lbase[2 * t] = lb; lbase[2 * t + 1] = lb + tot0;
    gdelta[2 * t] = gbase + excl0 - lb; gdelta[2 * t + 1] = gbase + gh.x + excl1 - (lb + tot0);
    __syncthreads();

    // phase 3: place items in digit order in shared memory
    uint32_t val[IPT];
#pragma unroll
    for (int i = 0; i < IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        val[i] = (vals_in && idx < n) ? vals_in[idx] : idx + val_base;
        const uint32_t d = (key[i] >> shift) & MASK;
        loff[i] += lbase[d] + wcount[w][d];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < IPT; i++) {
        const uint32_t idx = wbase_idx + i * 32 + lane;
        if (idx < n) { skey[loff[i]] = key[i]; sval[loff[i]] = val[i]; }
    }
    __syncthreads();
    // phase 4: coalesced write-out
    const uint32_t n_blk = min((uint32_t)TILE, n - blk * TILE);
    uint32_t* const vo = alt ? vals_out_alt : vals_out;
    uint32_t* const ko = alt ? nullptr : keys_out;
#pragma unroll
    for (int i = 0; i < IPT; i++) {
        const uint32_t lp = i * THREADS + t;
        if (lp < n_blk) {
            const uint32_t k = skey[lp];
            const uint32_t gpos = gdelta[(k >> shift) & MASK] + lp;
            if (ko) ko[gpos] = k;
            vo[gpos] = sval[lp];
        }
    }
}

size_t radix_ctl_words_wide(uint32_t n) { return RW_CTL_HEAD + (size_t)RW_HIST * cdiv(n > 0 ? n : 1, RS_TILE); }

// ctl: radix_ctl_words_wide(n) zeroed words.  The key array is used as a value buffer once the first pass has read it
// (keys are DESTROYED): the third pass must be able to write the sort's output, so neither of the buffers it reads may be
// vals_out, and K0 / K1 / V0 are taken.
static int radix_sort_by_key_u32_wide(uint32_t* keys, uint32_t* K0, uint32_t* K1, uint32_t* V0, uint32_t* vals_out, uint32_t n,
                                      uint32_t* ctl, cudaStream_t s, uint32_t val_base)
{
    const int nblocks = cdiv(n, RS_TILE);
    uint32_t* ghist = ctl;
    uint32_t* tickets = ctl + RW_HIST;
    uint32_t* status = ctl + RW_CTL_HEAD;
    GSR_LAUNCH(radix_hist_wide_kernel, min(cdiv(n, 4 * RS_THREADS * 4), 148 * 4), RS_THREADS, 0, s, keys, n, ghist);
    uint32_t* Vb[3] = {V0, keys, V0};       // values: pass 0 -> V0, 1 -> keys, 2 -> V0 (vals_out when pass 3 is the identity), 3 -> vals_out
    uint32_t* Kb[3] = {K0, K1, K0};
    const uint32_t* kin = keys;
    const uint32_t* vin = nullptr;
    for (int p = 0; p < 3; p++) {
        const bool third = p == 2;
        GSR_LAUNCH((radix_onesweep_wide_kernel<RW_BITS, 256, 16>), nblocks, 256, 0, s, kin, vin, Kb[p], Vb[p],
                   third ? vals_out : nullptr, n, RW_BITS * p, ghist + p * RW_NB, third ? ghist + 3 * RW_NB : nullptr, RW_TOP_NB,
                   reinterpret_cast<unsigned long long*>(status + (size_t)p * RW_NB * nblocks), tickets + p, val_base);
        kin = Kb[p]; vin = Vb[p];
    }
    return radix_pass(kin, vin, nullptr, vals_out, n, 3 * RW_BITS, RW_TOP_BITS, ghist + 3 * RW_NB, status + (size_t)3 * RW_NB * nblocks,
                      tickets + 3, s, val_base, 1, 1);
}

// stable LSD sort of n (key, value) pairs on all 32 key bits; values start as identity.
// K0, K1, V0: n-word ping-pong buffers; result values land in vals_out (keys are not kept).
// ctl: radix_ctl_words(n) words (digit histograms, tickets, look-back status), zeroed here.
size_t radix_ctl_words(uint32_t n) { return 4 * RS_NB + 64 + (size_t)4 * RS_NB * cdiv(n > 0 ? n : 1, RS_TILE); }
// ... for any block shape (the smallest block takes 2048 items)
static size_t radix_ctl_words_any(uint32_t n) { return 4 * RS_NB + 64 + (size_t)4 * RS_NB * cdiv(n > 0 ? n : 1, 2048); }
int radix_sort_by_key_u32(const uint32_t* keys, uint32_t* K0, uint32_t* K1, uint32_t* V0, uint32_t* vals_out, uint32_t n,
                          uint32_t* ctl, bool ctl_zeroed, cudaStream_t s, uint32_t val_base, int shape)
{
    // shape 3: three 9-bit passes + a conditional fourth (needs radix_ctl_words_wide(n) zeroed words; the depth sort of the
    // geometry scratch has them)
    if (shape == 3) {
        if (!ctl_zeroed) GSR_CUDA(cudaMemsetAsync(ctl, 0, radix_ctl_words_wide(n) * 4, s));
        return radix_sort_by_key_u32_wide(const_cast<uint32_t*>(keys), K0, K1, V0, vals_out, n, ctl, s, val_base);
    }
    // shape 2 needs radix_ctl_words_any(n) words of control block (the geometry scratch has them; radix_ctl_words(n) is what
    // shapes 0 and 1 need)
    const int nblocks = cdiv(n, shape == 2 ? 2048 : RS_TILE);
    uint32_t* ghist = ctl;
    uint32_t* tickets = ctl + 4 * RS_NB;
    uint32_t* status = tickets + 64;
    if (!ctl_zeroed) GSR_CUDA(cudaMemsetAsync(ctl, 0, (shape == 2 ? radix_ctl_words_any(n) : radix_ctl_words(n)) * 4, s));
    GSR_LAUNCH(radix_hist4_kernel, min(cdiv(n, 4 * RS_THREADS * 4), 148 * 4), RS_THREADS, 0, s, keys, n, ghist);
    uint32_t* Vb[2] = {V0, vals_out};       // 4 passes: the last one writes vals_out
    uint32_t* Kb[2] = {K0, K1};
    const uint32_t* kin = keys;
    const uint32_t* vin = nullptr;
    for (int p = 0; p < 4; p++) {
        const bool last = p == 3;
        int rc = radix_pass(kin, vin, last ? nullptr : Kb[p & 1], Vb[p & 1], n, 8 * p, 8, ghist + p * RS_NB,
                            status + (size_t)p * RS_NB * nblocks, tickets + p, s, val_base, shape);
        if (rc) return rc;
        kin = Kb[p & 1]; vin = Vb[p & 1];
    }
    return GSR_OK;
}

// A batched call's depth sort may run as several independent segment sorts (launch_depth_order): every segment needs
// its own histograms / tickets / look-back words inside the (already zeroed) control block.
constexpr int DEPTH_MAX_SEGMENTS = 64;
constexpr int DEPTH_MAX_VIEWS = 4095;          // status words of tile_prefix_kernel (one block per view) in the control block

// ------------------------------------------------------------------ geometry scratch layout
GeomScratch carve_geom_scratch(uint32_t* p, int P)
{
    const size_t Pp = (size_t)(P > 0 ? P : 1);
    const size_t Pa = align_up(Pp, 64);
    GeomScratch g;
    g.ctl = p;
    g.counters = p;
    g.tile_hist = p + 64;
    g.emit_status = reinterpret_cast<unsigned long long*>(p + 64 + 4 * RS_NB);
    g.view_status = p + 64 + 4 * RS_NB + 2 * align_up((size_t)cdiv(Pp, 256), 32);          // ticket + one word per view
    g.radix_ctl = g.view_status + (DEPTH_MAX_VIEWS + 1);
    g.ctl_words = align_up(64 + 4 * RS_NB + 2 * align_up((size_t)cdiv(Pp, 256), 32) + (DEPTH_MAX_VIEWS + 1) +
                               max(radix_ctl_words_any((uint32_t)Pp), radix_ctl_words_wide((uint32_t)Pp)) +
                               (size_t)DEPTH_MAX_SEGMENTS * (8 * RS_NB + 64),
                           64);                                                           // + slack for the segmented depth sort
    p += g.ctl_words;
    g.depthkey = p; p += Pa;
    g.K0 = p; p += Pa;
    g.K1 = p; p += Pa;
    g.V0 = p; p += Pa;
    g.rect = reinterpret_cast<ushort4*>(p); p += 2 * Pa;
    g.total_words = g.ctl_words + 6 * Pa;
    return g;
}
size_t geom_scratch_bytes(int P) { return carve_geom_scratch(nullptr, P).total_words * 4; }

// ------------------------------------------------------------------ tile-id digits of the split passes
struct TileDigits {
    int npass;
    int shift[4], bits[4];
};
static TileDigits tile_digits(uint32_t tiles)
{
    // bits = position of the highest set bit + 1 (what the reference's getHigherMsb computes, rasterizer_impl.cu:35-50)
    int tb = 0;
    while (tb < 32 && (tiles >> tb)) tb++;
    if (tb == 0) tb = 1;
    TileDigits d;
    d.npass = (tb + 7) / 8;
    int shift = 0;
    for (int p = 0; p < 4; p++) { d.shift[p] = 0; d.bits[p] = 0; }
    for (int p = 0; p < d.npass; p++) {
        d.bits[p] = (tb - shift + (d.npass - p) - 1) / (d.npass - p);      // balanced digits, e.g. 13 -> 7 + 6
        d.shift[p] = shift;
        shift += d.bits[p];
    }
    return d;
}

// ------------------------------------------------------------------ instances per tile -> tile ranges, split histograms
// Every visible Gaussian adds +1 over its tile rect: four corner updates of a 2-D difference array (shared-memory
// atomics, one flush per block); the 2-D prefix sum of that array is the number of instances of every tile.  Its
// exclusive scan in tile-id order IS the reference's tile ranges (identifyTileRanges, rasterizer_impl.cu:116-138:
// empty tiles stay (0,0)), and the digit histograms the split passes need are marginals of it - no pass over
// the R instances is needed for either.
template <bool SMEM>
__global__ void __launch_bounds__(1024) tile_count_kernel(int Pper, int gyv, const ushort4* __restrict__ rect, int gx1, int band,
                                                          int* __restrict__ diff)
{
    // blockIdx.y = view: its Pper virtual Gaussians, its band of (gyv + 1) x gx1 cells of the difference array (every view
    // has a band of its own, so the bands are independent and one band fits in shared memory whatever the number of views)
    extern __shared__ int sd[];
    const int view = blockIdx.y;
    rect += (size_t)view * Pper;
    diff += (size_t)view * band;
    const int row0 = view * gyv;
    if (SMEM) {
        for (int i = threadIdx.x; i < band; i += 1024) sd[i] = 0;
        __syncthreads();
    }
    int* d = SMEM ? sd : diff;
    for (int k = blockIdx.x * 1024 + threadIdx.x; k < Pper; k += gridDim.x * 1024) {
        const ushort4 r = rect[k];
        if (r.x == r.z || r.y == r.w) continue;
        const int y0 = r.y - row0, y1 = r.w - row0;
        atomicAdd(&d[y0 * gx1 + r.x], 1); atomicAdd(&d[y0 * gx1 + r.z], -1);
        atomicAdd(&d[y1 * gx1 + r.x], -1); atomicAdd(&d[y1 * gx1 + r.z], 1);
    }
    if (SMEM) {
        __syncthreads();
        for (int i = threadIdx.x; i < band; i += 1024)
            if (sd[i]) atomicAdd(&diff[i], sd[i]);
    }
}

// one block per view (taken by ticket, so a block only waits for views that are already running): 2-D prefix of the view's
// band of the difference array = instances per tile, exclusive scan over the tile ids (the views' totals chained through
// status words), ranges, and the view's share of the digit histograms of the tile split
template <bool SMEM>
__global__ void __launch_bounds__(1024) tile_prefix_kernel(int gx, int gyv, int* __restrict__ diff, uint2* __restrict__ ranges,
                                                           uint32_t* __restrict__ tile_hist, TileDigits td,
                                                           uint32_t* __restrict__ vstatus, uint32_t* __restrict__ tile_order)
{
    extern __shared__ int sd[];
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t sh[4 * RS_NB];
    __shared__ uint32_t s_view, s_base, s_max;
    __shared__ uint32_t bucket[64];
    const int gx1 = gx + 1, band = gx1 * (gyv + 1), t = threadIdx.x;
    if (t == 0) s_view = atomicAdd(vstatus, 1u);
    for (int i = t; i < 4 * RS_NB; i += 1024) sh[i] = 0;
    __syncthreads();
    const int view = (int)s_view;
    diff += (size_t)view * band;
    int* d = SMEM ? sd : diff;
    if (SMEM)
        for (int i = t; i < band; i += 1024) sd[i] = diff[i];
    __syncthreads();
    for (int x = t; x <= gx; x += 1024) {            // prefix down the columns
        int acc = 0;
        for (int y = 0; y <= gyv; y++) { acc += d[y * gx1 + x]; d[y * gx1 + x] = acc; }
    }
    __syncthreads();
    for (int y = t; y <= gyv; y += 1024) {           // prefix along the rows: d[y][x] = instances of tile (x, y)
        int acc = 0;
        for (int x = 0; x <= gx; x++) { acc += d[y * gx1 + x]; d[y * gx1 + x] = acc; }
    }
    __syncthreads();
    const int ntiles = gx * gyv, per = (ntiles + 1023) / 1024;
    const int i0 = min(ntiles, t * per), i1 = min(ntiles, i0 + per);
    uint32_t mine = 0;
    for (int i = i0; i < i1; i++) mine += (uint32_t)d[(i / gx) * gx1 + (i % gx)];
    uint32_t total;
    uint32_t start = block_excl_scan<32>(mine, warp_sum, &total);
    if (t == 0) {
        // instances of all earlier views: look back over their status words (aggregate / inclusive, like the radix passes)
        uint32_t base = 0;
        uint32_t* st = vstatus + 1;
        if (view == 0) st_volatile(st, ST_INC | total);
        else {
            st_volatile(st + view, ST_AGG | total);
            for (int j = view - 1;; j--) {
                uint32_t v;
                do { v = ld_volatile(st + j); } while ((v >> 30) == 0);
                base += v & ST_VAL;
                if (v & ST_INC) break;
            }
            st_volatile(st + view, ST_INC | (base + total));
        }
        s_base = base;
    }
    __syncthreads();
    start += s_base;
    const uint32_t tile0 = (uint32_t)view * (uint32_t)ntiles;
    for (int i = i0; i < i1; i++) {
        const uint32_t c = (uint32_t)d[(i / gx) * gx1 + (i % gx)];
        const uint32_t tile = tile0 + (uint32_t)i;
        ranges[tile] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        start += c;
        if (c)
            for (int p = 0; p < td.npass; p++) atomicAdd(&sh[p * RS_NB + ((tile >> td.shift[p]) & ((1u << td.bits[p]) - 1))], c);
    }
    __syncthreads();
    for (int i = t; i < 4 * RS_NB; i += 1024)
        if (sh[i]) atomicAdd(&tile_hist[i], sh[i]);          // tile_hist lies in the zeroed control block

    // launch order of the view's tiles for the blending kernels: longest lists first (64 buckets of list length, counting
    // sort), so the last CTAs of a view are short ones and the tail of the grid stays busy
    if (t == 0) s_max = 0;
    if (t < 64) bucket[t] = 0;
    __syncthreads();
    uint32_t mx = 0;
    for (int i = i0; i < i1; i++) mx = max(mx, (uint32_t)d[(i / gx) * gx1 + (i % gx)]);
    if (mx) atomicMax(&s_max, mx);
    __syncthreads();
    const uint32_t scale = s_max + 1;
    for (int i = i0; i < i1; i++) {
        const uint32_t c = (uint32_t)d[(i / gx) * gx1 + (i % gx)];
        atomicAdd(&bucket[63u - min(63u, (c * 64u) / scale)], 1u);
    }
    __syncthreads();
    if (t < 32) {                                            // exclusive scan of the 64 bucket sizes
        const uint32_t a = bucket[2 * t], b = bucket[2 * t + 1];
        uint32_t inc = a + b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
            if (t >= o) inc += u;
        }
        bucket[2 * t] = inc - a - b;
        bucket[2 * t + 1] = inc - b;
    }
    __syncthreads();
    for (int i = i0; i < i1; i++) {
        const uint32_t c = (uint32_t)d[(i / gx) * gx1 + (i % gx)];
        const uint32_t pos = atomicAdd(&bucket[63u - min(63u, (c * 64u) / scale)], 1u);
        tile_order[tile0 + pos] = tile0 + (uint32_t)i;
    }
}

// seg_items > 0: the P items are seg_items-sized runs (whole views of a batched call) that never share a tile, so each
// run may be sorted on its own (comparison arm, GSR_DEPTH_SEG_ITEMS; one sort over everything measured as fast or faster).
int launch_depth_order(int P, int W, int H, const GeomView& g, const ImgView& im, cudaStream_t s, int seg_items, int V)
{
    // P virtual Gaussians = V views x P / V; H = the stacked image height (V x whole tiles)
    const GeomScratch sc = carve_geom_scratch(g.scratch, P);
    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
    if (V < 1 || gy % V != 0 || P % V != 0 || V > DEPTH_MAX_VIEWS) return GSR_EINVAL;
    const int gyv = gy / V, Pper = P / V;
    const int band = (gx + 1) * (gyv + 1);
    const size_t smem = (size_t)band * sizeof(int);
    const TileDigits td = tile_digits((uint32_t)(gx * gy));
    // instances per tile (depends on preprocess only): ranges + histograms of the tile split
    const int cblocks = max(1, min(cdiv(Pper, 4096), max(1, 296 / V)));
    // The two tile kernels depend on preprocess only and the radix passes do not need them: they run on a side stream next to
    // the depth sort (both are small latency-bound launches) and are joined before this function returns.
    // (the side stream and its fork / join events belong to the device context, one set per caller stream)
    const bool forked = options().side_stream != 0;
    cudaStream_t main_s = s;
    DeviceCtx* ctx = nullptr;
    SideStream ss;
    {
        int rc0;
        if ((rc0 = device_ctx(&ctx))) return rc0;
        if (forked && (rc0 = ctx_side_stream(ctx, 0, main_s, &ss))) return rc0;
    }
    if (forked) {
        GSR_CUDA(cudaEventRecord(ss.fork, main_s));
        GSR_CUDA(cudaStreamWaitEvent(ss.side, ss.fork, 0));
        s = ss.side;
    }
    if (smem <= 160 * 1024) {
        if (smem > 40 * 1024) {
            // once per device: the largest band these launches accept
            int rc0 = ctx_once(ctx, 1u, [] {
                GSR_CUDA(cudaFuncSetAttribute(tile_count_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
                GSR_CUDA(cudaFuncSetAttribute(tile_prefix_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
                return (int)GSR_OK;
            });
            if (rc0) return rc0;
        }
        GSR_LAUNCH(tile_count_kernel<true>, dim3(cblocks, V), 1024, smem, s, Pper, gyv, sc.rect, gx + 1, band, im.tile_diff);
        GSR_LAUNCH(tile_prefix_kernel<true>, V, 1024, smem, s, gx, gyv, im.tile_diff, im.ranges, sc.tile_hist, td, sc.view_status, im.tile_order);
    } else {
        GSR_LAUNCH(tile_count_kernel<false>, dim3(max(cblocks, min(cdiv(Pper, 4096), 592)), V), 1024, 0, s, Pper, gyv, sc.rect, gx + 1, band,
                   im.tile_diff);
        GSR_LAUNCH(tile_prefix_kernel<false>, V, 1024, 0, s, gx, gyv, im.tile_diff, im.ranges, sc.tile_hist, td, sc.view_status, im.tile_order);
    }
    if (forked) {
        GSR_CUDA(cudaEventRecord(ss.join, ss.side));
        s = main_s;
    }
    // depth keys were written by preprocess; the control block was zeroed before it
    int rc = GSR_OK;
    if (seg_items <= 0 || seg_items >= P || (seg_items & 3) != 0 || cdiv(P, seg_items) > DEPTH_MAX_SEGMENTS) {
        rc = radix_sort_by_key_u32(sc.depthkey, sc.K0, sc.K1, sc.V0, g.order, (uint32_t)P, sc.radix_ctl, true, s, 0, options().radix_depth);
    } else {
        uint32_t* ctl = sc.radix_ctl;
        for (int off = 0; off < P && rc == GSR_OK; off += seg_items) {
            const int n = min(seg_items, P - off);
            rc = radix_sort_by_key_u32(sc.depthkey + off, sc.K0 + off, sc.K1 + off, sc.V0 + off, g.order + off, (uint32_t)n, ctl, true, s,
                                       (uint32_t)off, options().radix_depth == 3 ? 1 : options().radix_depth);
            ctl += radix_ctl_words_any((uint32_t)n);
        }
    }
    if (forked) GSR_CUDA(cudaStreamWaitEvent(main_s, ss.join, 0));      // join even when a launch failed
    return rc;
}

// ------------------------------------------------------------------ instance emission (in depth order)
// Chunks of 256 depth-ordered Gaussians, taken by ticket.  The exclusive prefix of the tile counts (the reference's
// InclusiveSum) is fused in: block scan + decoupled look-back over 64-bit per-chunk status words.  Emission is
// warp-cooperative: a warp takes 32 consecutive Gaussians and its lanes write the warp's instances with consecutive
// lanes on consecutive output slots (coalesced), finding the owner of slot t with a 5-step binary search over
// the lanes' offsets (shuffles).  Row-major over each Gaussian's tile rect, exactly the order of duplicateWithKeys
// (rasterizer_impl.cu:98-109).
constexpr unsigned long long ES_AGG = 1ull << 62, ES_INC = 2ull << 62, ES_VAL = (1ull << 62) - 1;

constexpr int EM_SUB = 4;                       // sub-rounds of 256 Gaussians per chunk (one look-back per chunk)
constexpr int EM_CHUNK = 256 * EM_SUB;
constexpr int EM_WIN = 256;                     // instance slots of a warp's window in shared memory (2 KB per warp)

__global__ void __launch_bounds__(256) emit_instances_kernel(int P, int gx, const uint32_t* __restrict__ order,
                                                             const ushort4* __restrict__ rect, uint32_t* __restrict__ tile_ids,
                                                             uint32_t* __restrict__ gauss_ids,
                                                             unsigned long long* __restrict__ estatus, uint32_t* __restrict__ ticket)
{
    __shared__ uint32_t s_chunk, s_base;
    __shared__ uint2 em_win[8][EM_WIN];
    __shared__ uint32_t em_ws[EM_SUB][8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nchunks = (P + EM_CHUNK - 1) / EM_CHUNK;
    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) s_chunk = atomicAdd(ticket, 1u);
        __syncthreads();
        const int c = (int)s_chunk;
        if (c >= nchunks) break;
        uint32_t g[EM_SUB];
        int x0[EM_SUB], y0[EM_SUB], wdt[EM_SUB], cnt[EM_SUB];
#pragma unroll
        for (int r = 0; r < EM_SUB; r++) {
            const int k = c * EM_CHUNK + r * 256 + threadIdx.x;
            g[r] = k < P ? order[k] : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int r = 0; r < EM_SUB; r++) {
            const ushort4 q = g[r] != 0xFFFFFFFFu ? rect[g[r]] : make_ushort4(0, 0, 0, 0);
            x0[r] = q.x; y0[r] = q.y;
            wdt[r] = max(1, (int)q.z - (int)q.x);
            cnt[r] = ((int)q.z - (int)q.x) * ((int)q.w - (int)q.y);
        }
        // exclusive offsets inside the chunk (sub-round major = depth order), chunk total: warp scans of all sub-rounds, ONE
        // barrier, then every thread adds the warp sums in front of it (the kernel was bound by its block barriers: 11 per
        // chunk with one block scan per sub-round and a single-thread look-back)
        uint32_t excl[EM_SUB], run = 0;
        {
            uint32_t inc[EM_SUB];
#pragma unroll
            for (int r = 0; r < EM_SUB; r++) {
                inc[r] = (uint32_t)cnt[r];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t u = __shfl_up_sync(0xffffffffu, inc[r], o);
                    if (lane >= o) inc[r] += u;
                }
                if (lane == 31) em_ws[r][w] = inc[r];
            }
            __syncthreads();
#pragma unroll
            for (int r = 0; r < EM_SUB; r++) {
                uint32_t wb = 0, all = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) { const uint32_t v = em_ws[r][i]; if (i < w) wb += v; all += v; }
                excl[r] = run + wb + inc[r] - (uint32_t)cnt[r];
                run += all;
            }
        }
        // look-back over the preceding chunks' status words, 32 at a time, by the first warp
        if (w == 0) {
            unsigned long long base = 0;
            if (c == 0) {
                if (lane == 0) st_volatile64(estatus, ES_INC | run);
            } else {
                if (lane == 0) st_volatile64(estatus + c, ES_AGG | run);
                for (int j = c - 1;; j -= 32) {
                    const int idx = j - lane;
                    unsigned long long v = ES_INC;                    // before chunk 0: an inclusive prefix of 0
                    if (idx >= 0) {
                        do { v = ld_volatile64(estatus + idx); } while ((v >> 62) == 0);
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, (v & ES_INC) != 0);
                    const int first = m ? __ffs(m) - 1 : 32;          // nearest chunk with an inclusive prefix
                    unsigned long long val = lane <= first ? (v & ES_VAL) : 0ull;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                    base += val;
                    if (m) break;
                }
                if (lane == 0) st_volatile64(estatus + c, ES_INC | (base + run));
            }
            if (lane == 0) s_base = (uint32_t)base;
        }
        __syncthreads();
        const uint32_t chunk_base = s_base;
#pragma unroll
        for (int r = 0; r < EM_SUB; r++) {
            const uint32_t off = chunk_base + excl[r];
            const uint32_t off0 = __shfl_sync(0xffffffffu, off, 0);
            const uint32_t rel = off - off0;
            // total instances of this warp = end offset of the last lane
            const uint32_t total = __shfl_sync(0xffffffffu, rel + (uint32_t)cnt[r], 31);
            // Push form: every lane writes the instances of ITS Gaussian (row-major over the rect, one running (row, col)
            // pair: no search, no division) into a window of EM_WIN slots of the warp's run in shared memory; the warp then
            // copies the window out with consecutive lanes on consecutive slots.  The push loop runs as long as the largest
            // rect of the 32 Gaussians (mean 5 tiles at C3), against 5 shuffle-search steps + 5 fetch shuffles + a division
            // per 32 instances of the pull form it replaces.
            uint2* const win = em_win[threadIdx.x >> 5];
            for (uint32_t base = 0; base < total; base += EM_WIN) {
                // this lane's instances inside [base, base + EM_WIN): k in [k0, k1)
                const uint32_t k0 = rel < base ? min(base - rel, (uint32_t)cnt[r]) : 0u;
                const uint32_t k1 = rel + (uint32_t)cnt[r] > base + EM_WIN ? (base + EM_WIN > rel ? base + EM_WIN - rel : 0u) : (uint32_t)cnt[r];
                if (k0 < k1) {
                    uint32_t row = 0, col = k0;
                    if (k0 >= (uint32_t)wdt[r]) { row = k0 / (uint32_t)wdt[r]; col = k0 - row * (uint32_t)wdt[r]; }    // only when a rect straddles windows
                    uint32_t tile = (uint32_t)((y0[r] + (int)row) * gx + x0[r] + (int)col);
                    uint2* dst = win + (rel + k0 - base);
                    for (uint32_t k = k0; k < k1; k++) {
                        *dst++ = make_uint2(tile, g[r]);
                        tile++; col++;
                        if (col == (uint32_t)wdt[r]) { col = 0; tile += (uint32_t)(gx - wdt[r]); }
                    }
                }
                __syncwarp();
                const uint32_t nwin = min((uint32_t)EM_WIN, total - base);
                for (uint32_t t = lane; t < nwin; t += 32) {
                    const uint2 v = win[t];
                    tile_ids[off0 + base + t] = v.x;
                    gauss_ids[off0 + base + t] = v.y;
                }
                __syncwarp();
            }
        }
    }
}

// scratch layout (u32 words): T0[R] G0[R] T1[R] G1[R] status[sum over passes of nblk * 2^bits] tickets[64]
static size_t bin_status_words(long long R)
{
    return (size_t)4 * RS_NB * cdiv(R > 0 ? R : 1, 2048) + 64;          // worst case: 4 passes of 8 bits in 2048-item blocks
}
size_t bin_scratch_bytes(long long R)
{
    const size_t Ra = align_up((size_t)(R > 0 ? R : 1), 64);
    return 4 * (4 * Ra + bin_status_words(R));
}

int launch_tile_split(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s)
{
    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
    (void)im;                            // ranges were written by tile_prefix_kernel
    if (R <= 0) return GSR_OK;
    if (R >= (1ll << 30)) return GSR_ETOOBIG;
    const GeomScratch sc = carve_geom_scratch(g.scratch, P);
    const size_t Ra = align_up((size_t)R, 64);
    uint32_t* T[2] = {b.scratch, b.scratch + 2 * Ra};
    uint32_t* G[2] = {b.scratch + Ra, b.scratch + 3 * Ra};
    uint32_t* status = b.scratch + 4 * Ra;
    const TileDigits td = tile_digits((uint32_t)(gx * gy));
    const int nblk = cdiv(R, options().radix_split == 2 ? 2048 : RS_TILE);
    size_t status_words = 64;
    for (int p = 0; p < td.npass; p++) status_words += (size_t)nblk << td.bits[p];
    GSR_CUDA(cudaMemsetAsync(status, 0, status_words * 4, s));
    uint32_t* tickets = status;          // first 64 words
    uint32_t* st = status + 64;

    GSR_LAUNCH(emit_instances_kernel, min(cdiv(P, EM_CHUNK), 148 * 8), 256, 0, s, P, gx, g.order, sc.rect, T[0], G[0], sc.emit_status,
               sc.counters + 1);
    int cur = 0;
    for (int p = 0; p < td.npass; p++) {
        const bool last = p == td.npass - 1;
        int rc = radix_pass(T[cur], G[cur], last ? nullptr : T[cur ^ 1], last ? b.vals : G[cur ^ 1], (uint32_t)R, td.shift[p], td.bits[p],
                            sc.tile_hist + p * RS_NB, st, tickets + p, s, 0, options().radix_split);
        if (rc) return rc;
        st += (size_t)nblk << td.bits[p];
        cur ^= 1;
    }
    return GSR_OK;
}

}  // namespace gsr
