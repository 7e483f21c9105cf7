// arms/sort_cub.cu - TEST-ONLY COMPARISON ARM (option sort_cub = 1, libgsr_b200_arms.so): the reference's binning formulation on the same
// inputs - id-major emission of 64-bit tile|depth keys, cub::DeviceRadixSort::SortPairs over
// 32 + tile_bits bits (CUB 2.8 picks its sm_100 tuning), identifyTileRanges
// (rasterizer_impl.cu:70-138,277,300-318).  It exists so tests can assert that our trimmed-key sort
// yields the identical sequence and so bench.py can time "the bar to beat" on the same GPU.
// It is never the default path.
#include "../gsr_common.cuh"
#include <cub/cub.cuh>

namespace gsr {

__global__ void __launch_bounds__(256) cub_tiles_kernel(int P, const SplatRec* __restrict__ rec, uint32_t* __restrict__ tiles)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) tiles[i] = __float_as_uint(rec[i].q2.w);
}

__global__ void __launch_bounds__(256) cub_emit_kernel(int P, int gx, int gy, const SplatRec* __restrict__ rec,
                                                       const uint32_t* __restrict__ offs_incl, unsigned long long* __restrict__ keys,
                                                       uint32_t* __restrict__ vals)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float4 q2 = rec[i].q2;
    const int rad = __float_as_int(q2.z);
    if (rad <= 0) return;
    const float4 q0 = rec[i].q0;
    const int x0 = min(gx, max(0, (int)((q0.x - rad) / GSR_TILE)));
    const int y0 = min(gy, max(0, (int)((q0.y - rad) / GSR_TILE)));
    const int x1 = min(gx, max(0, (int)((q0.x + rad + GSR_TILE - 1) / GSR_TILE)));
    const int y1 = min(gy, max(0, (int)((q0.y + rad + GSR_TILE - 1) / GSR_TILE)));
    uint32_t off = i == 0 ? 0 : offs_incl[i - 1];
    const unsigned long long dbits = __float_as_uint(q2.y);
    for (int y = y0; y < y1; y++)
        for (int x = x0; x < x1; x++) {
            keys[off] = ((unsigned long long)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off] = (uint32_t)i;
            off++;
        }
}

__global__ void __launch_bounds__(256) cub_ranges_kernel(uint32_t R, const unsigned long long* __restrict__ keys, uint2* __restrict__ ranges)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    const uint32_t cur = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
        if (cur != prev) { ranges[prev].y = i; ranges[cur].x = i; }
    }
    if (i == R - 1) ranges[cur].y = R;
}

static size_t cub_temp_bytes(long long R, int P)
{
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (unsigned long long*)nullptr, (unsigned long long*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (int)R, 0, 64);
    cub::DeviceScan::InclusiveSum(nullptr, b, (uint32_t*)nullptr, (uint32_t*)nullptr, P);
    return align_up(a > b ? a : b);
}

size_t arm_cub_sort_scratch_bytes(long long R)
{
    const size_t Rp = (size_t)(R > 0 ? R : 1);
    // keys_unsorted, keys_sorted (u64), vals_unsorted (u32), temp (sized for the worst case of P <= 2^27 scan)
    return 2 * align_up(Rp * 8) + align_up(Rp * 4) + cub_temp_bytes(R > 0 ? R : 1, 1 << 27);
}

int arm_tile_split_cub(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s)
{
    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
    GSR_CUDA(cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * (size_t)gx * gy, s));
    if (R <= 0) return GSR_OK;
    const size_t Rp = (size_t)R;
    char* p = (char*)b.scratch;
    unsigned long long* keys_u = (unsigned long long*)p; p += align_up(Rp * 8);
    unsigned long long* keys_s = (unsigned long long*)p; p += align_up(Rp * 8);
    uint32_t* vals_u = (uint32_t*)p; p += align_up(Rp * 4);
    void* temp = p;
    size_t temp_bytes = cub_temp_bytes(R, 1 << 27);
    // per-id tile counts and their inclusive scan live in the geometry scratch (free at this point)
    uint32_t* tiles = g.scratch;
    uint32_t* offs_incl = g.scratch + align_up((size_t)P, 64);
    GSR_LAUNCH(cub_tiles_kernel, cdiv(P, 256), 256, 0, s, P, g.rec, tiles);
    GSR_CUDA(cub::DeviceScan::InclusiveSum(temp, temp_bytes, tiles, offs_incl, P, s));
    ++g_launches;
    GSR_LAUNCH(cub_emit_kernel, cdiv(P, 256), 256, 0, s, P, gx, gy, g.rec, offs_incl, keys_u, vals_u);
    int bit = 0;
    while (bit < 32 && ((uint32_t)(gx * gy) >> bit)) bit++;
    GSR_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_u, keys_s, vals_u, b.vals, (int)R, 0, 32 + bit, s));
    g_launches += 8;
    GSR_LAUNCH(cub_ranges_kernel, cdiv(R, 256), 256, 0, s, (uint32_t)R, keys_s, im.ranges);
    return GSR_OK;
}

}  // namespace gsr
