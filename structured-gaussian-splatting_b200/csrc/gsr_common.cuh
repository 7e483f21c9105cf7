// gsr_common.cuh - shared definitions of the B200-native rasterizer core (sm_100a only).
//
// Private to csrc/: blob layouts, launch bookkeeping, small device helpers.
// The public boundary is include/gsr.h (plain C).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/gsr.h"

#define GSR_TILE 16                 // reference config.h:16-17 (BLOCK_X = BLOCK_Y = 16), part of the bit-exact contract
#define GSR_MAGIC 0x47535242u       // "GSRB"
#define GSR_MAX_VIEWS 16            // views the fused per-Gaussian backward of a batched call handles in one launch

namespace gsr {

// ---------------------------------------------------------------- error / launch bookkeeping
extern unsigned long long g_launches;          // api.cu
extern char g_last_cuda_error[256];            // api.cu
int record_cuda_error(cudaError_t e, const char* where);

#define GSR_CUDA(call)                                                           \
    do {                                                                         \
        cudaError_t _e = (call);                                                 \
        if (_e != cudaSuccess) return gsr::record_cuda_error(_e, #call);         \
    } while (0)

// every kernel launch goes through this so gsr_launch_count() is an honest count
#define GSR_LAUNCH(kernel, grid, block, smem, stream, ...)                       \
    do {                                                                         \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);              \
        ++gsr::g_launches;                                                       \
        cudaError_t _e = cudaPeekAtLastError();                                  \
        if (_e != cudaSuccess) return gsr::record_cuda_error(_e, #kernel);       \
    } while (0)

// CHECK_CUDA semantics of the reference (auxiliary.h:166-173): in debug mode synchronise after each stage
#define GSR_STAGE(debug, stream, name)                                           \
    do {                                                                         \
        if (debug) {                                                             \
            cudaError_t _e = cudaStreamSynchronize(stream);                      \
            if (_e != cudaSuccess) return gsr::record_cuda_error(_e, name);      \
        }                                                                        \
    } while (0)

// optional per-stage event timing (api.cu)
void stage_begin(int stage, cudaStream_t s);
void stage_end(int stage, cudaStream_t s);

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }
static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------- per-Gaussian record (48 B, three float4)
// Written by preprocess for every Gaussian; gathered by tile-list entry in the blending kernels.
//   q0 = { pix.x, pix.y, conic.x, conic.y }
//   q1 = { conic.z, opacity, r, g }
//   q2 = { b, depth, radius (int bits), tiles_touched (uint bits) }
// Culled Gaussians have q2.z = q2.w = 0 and depth bits 0xFFFFFFFF (sorts last).
struct __align__(16) SplatRec { float4 q0, q1, q2; };

// ---------------------------------------------------------------- blob layouts (private; only fwd<->bwd consistency matters)
struct BlobHeader {            // first 256 bytes of every blob
    uint32_t magic, kind;      // kind: 1 geom, 2 binning, 3 image
    int32_t  P, R, W, H;
    int32_t  n_sorted;         // geom: number of entries in `order`
    uint32_t pad[57];
};
static_assert(sizeof(BlobHeader) == 256, "header is 256 B");

struct GeomView {              // views into the geometry blob
    BlobHeader* hdr;
    SplatRec*   rec;           // [P]
    uint8_t*    clamped;       // [P] bit c set <=> SH colour channel c was clamped to 0
    uint32_t*   order;         // [P] Gaussian ids sorted by (depth bits, id); culled last
    uint32_t*   scratch;       // sort / scan scratch (GeomScratch), only live inside forward
    size_t      scratch_bytes;
};
struct BinView {
    BlobHeader* hdr;
    uint32_t*   vals;          // [R] Gaussian id per instance, sorted by (tile, depth bits, id)
    uint32_t*   scratch;       // only live inside forward
    size_t      scratch_bytes;
};
struct ImgView {
    BlobHeader* hdr;
    uint2*      ranges;        // [tiles]
    float*      final_T;       // [W*H]
    uint32_t*   n_contrib;     // [W*H]
    int*        tile_diff;     // [V][(gy+1)*(gx+1)] 2-D difference array of the tile rects, one band per view (zeroed before preprocess), only live inside forward
    uint32_t*   tile_order;    // [tiles] the tiles of every view by decreasing list length (blending launch order), written by tile_prefix_kernel
};

// layout of GeomView::scratch (sort.cu).  The control block [ctl, ctl + ctl_words) must be zero before preprocess runs.
struct GeomScratch {
    uint32_t* ctl;
    size_t    ctl_words;
    uint32_t* counters;        // [64]: 0 = R (sum of tiles touched, accumulated by preprocess), 1 = emit ticket, 8 = prefiltered-violated flag
    uint32_t* tile_hist;       // [4][256] digit histograms of the tile-split passes (counted by emit_instances_kernel)
    unsigned long long* emit_status;   // [ceil(P/256)] look-back words of the fused tile-count scan
    uint32_t* view_status;     // tile_prefix_kernel: ticket + one look-back word per view of a batched call
    uint32_t* radix_ctl;       // digit histograms, tickets and look-back words of the four depth passes
    uint32_t* depthkey;        // [P] float bits of the view-space depth, 0xFFFFFFFF for culled Gaussians (written by preprocess)
    ushort4*  rect;            // [P] tile rect (x0, y0, x1, y1) of every Gaussian, all zero when culled (written by preprocess)
    uint32_t *K0, *K1, *V0;    // [P] ping-pong buffers of the depth sort
    size_t    total_words;
};
GeomScratch carve_geom_scratch(uint32_t* scratch, int P);
size_t geom_scratch_bytes(int P);
size_t bin_scratch_bytes(long long R);
GeomView geom_view(void* blob, int P);
BinView  bin_view(void* blob, long long R);
ImgView  img_view(void* blob, int W, int H, int V = 1);

// ---------------------------------------------------------------- stage entry points (one per .cu)
struct Camera {                // lives in shared memory of the per-Gaussian kernels
    float view[16], proj[16];
    float cam[3];
    float tanx, tany, fx, fy;
    int W, H, gx, gy;
};
struct CameraArgs {            // kernel argument: device pointers to the matrices (no host read-back)
    const float *view, *proj, *cam;
    float tanx, tany, fx, fy;
    int W, H, gx, gy;
};
static inline CameraArgs make_camera_args(const float* view, const float* proj, const float* cam, float tanx, float tany,
                                          int W, int H)
{
    CameraArgs c;
    c.view = view; c.proj = proj; c.cam = cam;
    c.tanx = tanx; c.tany = tany;
    c.fy = H / (2.0f * tany);            // rasterizer_impl.cu:222-223
    c.fx = W / (2.0f * tanx);
    c.W = W; c.H = H;
    c.gx = (W + GSR_TILE - 1) / GSR_TILE; c.gy = (H + GSR_TILE - 1) / GSR_TILE;
    return c;
}
#ifdef __CUDACC__
// whole block must call; ends with a barrier
__device__ __forceinline__ void load_camera(Camera& cam, const CameraArgs& ca)
{
    const int t = threadIdx.x;
    if (t < 16) cam.view[t] = ca.view[t];
    else if (t < 32) cam.proj[t - 16] = ca.proj ? ca.proj[t - 16] : 0.f;
    else if (t < 35) cam.cam[t - 32] = ca.cam ? ca.cam[t - 32] : 0.f;
    else if (t == 35) {
        cam.tanx = ca.tanx; cam.tany = ca.tany; cam.fx = ca.fx; cam.fy = ca.fy;
        cam.W = ca.W; cam.H = ca.H; cam.gx = ca.gx; cam.gy = ca.gy;
    }
    __syncthreads();
}
#endif

// preprocess.cu
int launch_preprocess(const GsrFwdArgs& a, int V, int view_base, int V_total, const GeomView& g, cudaStream_t s);
int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t s);
int launch_preprocess_backward(const GsrBwdArgs& a, int V, const GeomView& g, const float* accum, cudaStream_t s);

// sort.cu
// sorts Gaussian ids by depth bits (stable) into g.order; counts instances per tile -> tile ranges, split-pass histograms
int launch_depth_order(int P, int W, int H, const GeomView& g, const ImgView& im, cudaStream_t s, int seg_items = 0, int V = 1);
// emits instances in depth order, splits them stably by tile id, writes vals + tile ranges
int launch_tile_split(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s);
// comparison arm only (bench / tests): the reference's formulation, cub::DeviceRadixSort on 64-bit keys
int launch_tile_split_cub(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s);
size_t cub_sort_scratch_bytes(long long R);

// generic stable radix sort of (u32 key, identity value) pairs on all 32 key bits (sort.cu); ctl = radix_ctl_words(n)
// words of control state, zeroed by the call unless ctl_zeroed
size_t radix_ctl_words(uint32_t n);
int radix_sort_by_key_u32(const uint32_t* keys, uint32_t* K0, uint32_t* K1, uint32_t* V0, uint32_t* vals_out, uint32_t n,
                          uint32_t* ctl, bool ctl_zeroed, cudaStream_t s, uint32_t val_base = 0);   // values = val_base + index

// blend.cu
size_t blend_plane_words(long long R, int W, int H);
int launch_blend_forward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                         float* out_color, cudaStream_t s);
int launch_blend_backward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                          const float* dL_dpix, float* accum, cudaStream_t s);

// knn.cu
int launch_dist2(int P, const float* points, float* out, GsrAlloc scratch, cudaStream_t s);

int env_int(const char* name, int dflt);

}  // namespace gsr
