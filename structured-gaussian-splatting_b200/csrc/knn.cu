// knn.cu - distCUDA2: mean squared distance of every point to its 3 nearest neighbours.
//
// Replaces SimpleKNN::knn and its kernels (reference: submodules/simple-knn/simple_knn.cu:185-221;
// coord2Morton :54-70, boxMinMax :78-117, boxMeanDist :147-183, updateKBest :131-145).
//
// The result is the exact 3-NN (by index: coincident points count as distance 0), so it does not
// depend on the search order - only on the fp32 distance expression, which is kept as
// d.x*d.x + d.y*d.y + d.z*d.z.  That leaves the search free to be organised for the GPU:
//   * bounding box with the reference's init (0,0,0) by float atomics, no host read-back (the
//     reference copies min and max to the host, two blocking syncs);
//   * Morton codes, our stable 8-bit radix sort (4 passes), points gathered into Morton order as float4;
//   * boxes of 256 Morton-consecutive points; ONE CTA per query box.  Candidate boxes are pruned at
//     CTA level (box-to-box distance against the CTA's worst current 3rd-best), their points are
//     staged in shared memory and scanned by all 256 queries with broadcast reads - instead of the
//     reference's per-thread walk over all boxes with divergent global gathers.
// Pruning is conservative (a box is skipped only if its distance bound exceeds the current 3rd best,
// computed with the same monotone fp32 expression), so the output equals brute force.
#include "gsr_common.cuh"
#include <cfloat>
#include <mutex>

namespace gsr {

constexpr int KB = 256;   // points per box = threads per CTA

struct KnnBox { float mn[3], mx[3], pad[2]; };   // 32 B

__global__ void __launch_bounds__(256) knn_bounds_kernel(int P, const float* __restrict__ pts, unsigned int* __restrict__ mm)
{
    // mm[0..2]: min as "most negative" via unsigned max on negative floats; mm[3..5]: max via signed max. init 0 = reference init (0,0,0)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float v[3] = {0.f, 0.f, 0.f};
    if (i < P) { v[0] = pts[3 * i]; v[1] = pts[3 * i + 1]; v[2] = pts[3 * i + 2]; }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float lo = fminf(v[k], 0.f), hi = fmaxf(v[k], 0.f);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if ((threadIdx.x & 31) == 0) {
            if (lo < 0.f) atomicMax(&mm[k], __float_as_uint(lo));
            if (hi > 0.f) atomicMax(reinterpret_cast<int*>(&mm[3 + k]), __float_as_int(hi));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x)
{
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ pts, const unsigned int* __restrict__ mm,
                                                         uint32_t* __restrict__ codes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t c[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float mn = __uint_as_float(mm[k]), mx = __uint_as_float(mm[3 + k]);
        c[k] = spread10((uint32_t)(((pts[3 * i + k] - mn) / (mx - mn)) * ((1 << 10) - 1)));
    }
    codes[i] = c[0] | (c[1] << 1) | (c[2] << 2);
}

__global__ void __launch_bounds__(KB) knn_gather_boxes_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                                                              float4* __restrict__ sorted, KnnBox* __restrict__ boxes)
{
    __shared__ float red[6][KB / 32];
    const int i = blockIdx.x * KB + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t id = order[i];
        const float x = pts[3 * id], y = pts[3 * id + 1], z = pts[3 * id + 2];
        sorted[i] = make_float4(x, y, z, __uint_as_float(id));
        lo[0] = hi[0] = x; lo[1] = hi[1] = y; lo[2] = hi[2] = z;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31) == 0) { red[k][threadIdx.x >> 5] = lo[k]; red[3 + k][threadIdx.x >> 5] = hi[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        float a = FLT_MAX, b = -FLT_MAX;
        for (int w = 0; w < KB / 32; w++) { a = fminf(a, red[threadIdx.x][w]); b = fmaxf(b, red[3 + threadIdx.x][w]); }
        boxes[blockIdx.x].mn[threadIdx.x] = a;
        boxes[blockIdx.x].mx[threadIdx.x] = b;
    }
}

__device__ __forceinline__ void kbest3(float dist, float& b0, float& b1, float& b2)
{
    // insertion keeping the three smallest (updateKBest<3>, simple_knn.cu:131-145)
    if (b0 > dist) { float t = b0; b0 = dist; dist = t; }
    if (b1 > dist) { float t = b1; b1 = dist; dist = t; }
    if (b2 > dist) { b2 = dist; }
}

__device__ __forceinline__ float block_max(float v, float* sh)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = sh[0];
#pragma unroll
    for (int w = 1; w < KB / 32; w++) r = fmaxf(r, sh[w]);
    return r;
}

__global__ void __launch_bounds__(KB) knn_search_kernel(int P, int nbox, const float4* __restrict__ sorted,
                                                        const KnnBox* __restrict__ boxes, float* __restrict__ out)
{
    __shared__ float4 cand[KB];
    __shared__ int list[KB];
    __shared__ int list_n;
    __shared__ float red[KB / 32];
    const int own = blockIdx.x;
    const int i = own * KB + threadIdx.x;
    const bool valid = i < P;
    const int lane = threadIdx.x & 31;

    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) q = sorted[i];
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;

    // own box first: gives every query a finite bound quickly
    cand[threadIdx.x] = q;
    __syncthreads();
    {
        const int n = min(KB, P - own * KB);
        if (valid)
            for (int j = 0; j < n; j++) {
                if (j == (int)threadIdx.x) continue;
                const float4 c = cand[j];
                const float dx = c.x - q.x, dy = c.y - q.y, dz = c.z - q.z;
                kbest3(dx * dx + dy * dy + dz * dz, b0, b1, b2);
            }
    }
    const KnnBox me = boxes[own];
    float bound = block_max(valid ? b2 : 0.f, red);

    for (int base = 0; base < nbox; base += KB) {
        // CTA-level candidate selection among boxes [base, base+KB)
        if (threadIdx.x == 0) list_n = 0;
        __syncthreads();
        const int b = base + threadIdx.x;
        bool keep = false;
        if (b < nbox && b != own) {
            const KnnBox o = boxes[b];
            float d2;
            {
                const float gx = fmaxf(0.f, fmaxf(me.mn[0] - o.mx[0], o.mn[0] - me.mx[0]));
                const float gy = fmaxf(0.f, fmaxf(me.mn[1] - o.mx[1], o.mn[1] - me.mx[1]));
                const float gz = fmaxf(0.f, fmaxf(me.mn[2] - o.mx[2], o.mn[2] - me.mx[2]));
                d2 = gx * gx + gy * gy + gz * gz;
            }
            keep = !(d2 > bound);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, keep);
        int wbase = 0;
        if (lane == 0 && m) wbase = atomicAdd(&list_n, __popc(m));
        wbase = __shfl_sync(0xffffffffu, wbase, 0);
        if (keep) list[wbase + __popc(m & ((1u << lane) - 1))] = b;
        __syncthreads();
        const int n_list = list_n;
        for (int li = 0; li < n_list; li++) {
            const int cb = list[li];
            const int n = min(KB, P - cb * KB);
            __syncthreads();                       // previous candidate fully consumed
            if ((int)threadIdx.x < n) cand[threadIdx.x] = sorted[cb * KB + threadIdx.x];
            __syncthreads();
            if (valid) {
                // per-query test against the candidate's bounds (distBoxPoint, simple_knn.cu:119-129)
                const KnnBox o = boxes[cb];
                float ex = 0.f, ey = 0.f, ez = 0.f;
                if (q.x < o.mn[0] || q.x > o.mx[0]) ex = fminf(fabsf(q.x - o.mn[0]), fabsf(q.x - o.mx[0]));
                if (q.y < o.mn[1] || q.y > o.mx[1]) ey = fminf(fabsf(q.y - o.mn[1]), fabsf(q.y - o.mx[1]));
                if (q.z < o.mn[2] || q.z > o.mx[2]) ez = fminf(fabsf(q.z - o.mn[2]), fabsf(q.z - o.mx[2]));
                const float dbox = ex * ex + ey * ey + ez * ez;
                if (!(dbox > b2)) {
                    for (int j = 0; j < n; j++) {
                        const float4 c = cand[j];
                        const float dx = c.x - q.x, dy = c.y - q.y, dz = c.z - q.z;
                        kbest3(dx * dx + dy * dy + dz * dz, b0, b1, b2);
                    }
                }
            }
        }
        if (n_list > 0) bound = block_max(valid ? b2 : 0.f, red);
        else __syncthreads();
    }
    if (valid) out[__float_as_uint(q.w)] = (b0 + b1 + b2) / 3.0f;
}

int launch_dist2(int P, const float* points, float* out, GsrAlloc scratch, cudaStream_t s)
{
    const int nbox = cdiv(P, KB);
    const size_t Pa = align_up((size_t)P, 64);
    const size_t words = 64 + 5 * Pa + radix_ctl_words((uint32_t)P) + 4 * Pa /*float4 sorted*/ + 8 * (size_t)nbox + 64;
    uint32_t* base = (uint32_t*)scratch.fn(scratch.ctx, words * 4);
    if (!base) return GSR_ENOMEM;
    uint32_t* mm = base;
    uint32_t* codes = base + 64;
    uint32_t* K0 = codes + Pa;
    uint32_t* K1 = K0 + Pa;
    uint32_t* V0 = K1 + Pa;
    uint32_t* order = V0 + Pa;
    uint32_t* hist = order + Pa;
    float4* sorted = reinterpret_cast<float4*>(hist + radix_ctl_words((uint32_t)P));
    KnnBox* boxes = reinterpret_cast<KnnBox*>(reinterpret_cast<uint32_t*>(sorted) + 4 * Pa);

    GSR_CUDA(cudaMemsetAsync(mm, 0, 64 * 4, s));
    GSR_LAUNCH(knn_bounds_kernel, cdiv(P, 256), 256, 0, s, P, points, mm);
    GSR_LAUNCH(knn_morton_kernel, cdiv(P, 256), 256, 0, s, P, points, mm, codes);
    int rc = radix_sort_by_key_u32(codes, K0, K1, V0, order, (uint32_t)P, hist, false, s);
    if (rc) return rc;
    GSR_LAUNCH(knn_gather_boxes_kernel, nbox, KB, 0, s, P, points, order, sorted, boxes);
    GSR_LAUNCH(knn_search_kernel, nbox, KB, 0, s, P, nbox, sorted, boxes, out);
    return GSR_OK;
}

}  // namespace gsr

// ------------------------------------------------------------------------------------------ distCUDA2 without an allocator
// gsr_dist2_static (include/gsr.h): same kernels as gsr_dist2, scratch in the file-scope device array below instead of
// an allocator callback, for callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in the
// reference.  Every GPU that loads the library holds the array until exit.
// All calls on a device share it: it keeps the last call's contents (launch_dist2 clears what it needs zeroed: the
// bounding-box words and the sort's control block), and calls that overlap on the GPU race.
namespace {

constexpr size_t kDist2StaticWords = 64 + 5 * (size_t)GSR_DIST2_STATIC_MAX_POINTS + 4 * (size_t)GSR_DIST2_STATIC_MAX_POINTS
                                   + 8 * ((size_t)GSR_DIST2_STATIC_MAX_POINTS / gsr::KB) + 64
                                   + 128 * 1024 /* control block of the sort: radix_ctl_words(max points), checked at run time */;
static __device__ __align__(256) uint32_t g_dist2_static_scratch[kDist2StaticWords];

constexpr int kMaxDevices = 64;
std::mutex g_dist2_static_mu;
void* g_dist2_static_ptr[kMaxDevices];          // device address of the array on each device, looked up once per device

struct StaticArena { void* base; size_t bytes; };
void* static_arena_alloc(void* ctx, size_t bytes)      // a GsrAlloc over the array: hands it out if the request fits
{
    const StaticArena* a = static_cast<const StaticArena*>(ctx);
    return bytes <= a->bytes ? a->base : nullptr;
}

}  // namespace

extern "C" int gsr_dist2_static(int P, const float* points, float* out, gsr_stream_t stream)
{
    if (P < 0 || (P > 0 && (!points || !out))) return GSR_EINVAL;
    if (P > GSR_DIST2_STATIC_MAX_POINTS)
        return gsr::record_error_text(GSR_ETOOBIG, "gsr_dist2_static: P exceeds GSR_DIST2_STATIC_MAX_POINTS (262144); use gsr_dist2");
    if (P == 0) return GSR_OK;
    int dev = -1;
    GSR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return gsr::record_error_text(GSR_EINVAL, "gsr_dist2_static: device ordinal out of range");
    void* base = nullptr;
    {
        // the array's name is a host-side stand-in here, not a device address: ask the runtime, with the device current
        std::lock_guard<std::mutex> lk(g_dist2_static_mu);
        if (!g_dist2_static_ptr[dev]) {
            void* p = nullptr;
            GSR_CUDA(cudaGetSymbolAddress(&p, g_dist2_static_scratch));
            g_dist2_static_ptr[dev] = p;
        }
        base = g_dist2_static_ptr[dev];
    }
    StaticArena arena{base, sizeof(uint32_t) * kDist2StaticWords};
    const int rc = gsr::launch_dist2(P, points, out, GsrAlloc{static_arena_alloc, &arena}, (cudaStream_t)stream);
    if (rc == GSR_ENOMEM) return gsr::record_error_text(GSR_ETOOBIG, "gsr_dist2_static: scratch array too small for this P");
    return rc;
}
