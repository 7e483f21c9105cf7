// gsr_common.cuh - shared definitions of the B200-native rasterizer core (sm_100a only).
//
// Private to csrc/: blob layouts, launch bookkeeping, small device helpers.
// The public boundary is include/gsr.h (plain C).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "../../include/gsr.h"

#define GSR_TILE 16                 // reference config.h:16-17 (BLOCK_X = BLOCK_Y = 16), part of the bit-exact contract
#define GSR_MAGIC 0x47535242u       // "GSRB"
#define GSR_MAX_VIEWS 16            // views the fused per-Gaussian backward of a batched call handles in one launch

namespace gsr {

// ---------------------------------------------------------------- error / launch bookkeeping
extern std::atomic<unsigned long long> g_launches;      // api.cu
int record_cuda_error(cudaError_t e, const char* where);    // the message is kept per host thread
int record_error_text(int code, const char* text);           // same slot (gsr_last_cuda_error()); returns `code`

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
    uint32_t* depthkey;        // [P] sort key of the depth pass (written by preprocess): float bits of the view-space depth, 0xFFFFFFFF for
                               // culled Gaussians; under radix_depth = 3 the bits minus those of the near plane, culled 0 (depth_key_bias)
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
// n_views <= GSR_MAX_VIEWS single-view blobs filled in one pass over the parameters (gsr_preprocess_views)
int launch_preprocess_views(const GsrFwdArgs& a, int n_views, const GsrViewFront* views, const GeomView* g, cudaStream_t s);
int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t s);
// per-view base pointers of a multi-view per-Gaussian backward (kernel parameter; <= GSR_MAX_VIEWS views)
struct ViewTab {
    const int*      radii[GSR_MAX_VIEWS];
    const SplatRec* rec[GSR_MAX_VIEWS];
    const uint8_t*  clamped[GSR_MAX_VIEWS];
    const float*    accum[GSR_MAX_VIEWS];
    float*          m2d[GSR_MAX_VIEWS];
    const float*    view[GSR_MAX_VIEWS];
    const float*    proj[GSR_MAX_VIEWS];
    const float*    cam[GSR_MAX_VIEWS];
};
int launch_preprocess_backward(const GsrBwdArgs& a, int V, const GeomView& g, const float* accum, cudaStream_t s,
                               const ViewTab* tab = nullptr);

// sort.cu
// sorts Gaussian ids by depth bits (stable) into g.order; counts instances per tile -> tile ranges, split-pass histograms
int launch_depth_order(int P, int W, int H, const GeomView& g, const ImgView& im, cudaStream_t s, int seg_items = 0, int V = 1);
// emits instances in depth order, splits them stably by tile id, writes vals + tile ranges
int launch_tile_split(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s);

// generic stable radix sort of (u32 key, identity value) pairs on all 32 key bits (sort.cu); ctl = radix_ctl_words(n)
// words of control state, zeroed by the call unless ctl_zeroed.  shape 3 (three 9-bit passes + a conditional fourth) needs
// radix_ctl_words_wide(n) words and DESTROYS the keys (the array serves as a value buffer).
size_t radix_ctl_words(uint32_t n);
size_t radix_ctl_words_wide(uint32_t n);
// depth keys under radix_depth = 3: bits(depth) - bits(0.2f) for Gaussians that emit instances (their depth is > 0.2 or NaN,
// so the difference never wraps and the unsigned order is unchanged), 0 for culled ones (they emit nothing: their place in
// the order is immaterial)
constexpr uint32_t GSR_DEPTH_KEY_BIAS = 0x3E4CCCCDu;
int radix_sort_by_key_u32(const uint32_t* keys, uint32_t* K0, uint32_t* K1, uint32_t* V0, uint32_t* vals_out, uint32_t n,
                          uint32_t* ctl, bool ctl_zeroed, cudaStream_t s, uint32_t val_base = 0,     // values = val_base + index
                          int shape = 0);                                                           // block shape, see radix_pass

// blend.cu
size_t blend_plane_words(long long R, int W, int H);
int launch_blend_forward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                         float* out_color, cudaStream_t s);
int launch_blend_backward(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                          const float* dL_dpix, float* accum, cudaStream_t s);

// knn.cu
int launch_dist2(int P, const float* points, float* out, GsrAlloc scratch, cudaStream_t s);

// ---------------------------------------------------------------- tuning options (context.cu)
// Read from the environment ONCE, at the first call into the library (GSR_<NAME>); tests and benchmarks change them through
// gsr_debug_set_option().  The product defaults are the values below; nothing calls getenv on the hot path.
struct Options {
    int bwd_variant = 7;        // blending backward kernel: 7 = blend_bwd_v7_kernel (packed pairs, quad commit, ids through cp.async),
                                // 73 = ids through a register, 72 = with v6's branchy commit (and ids through a register),
                                // 70 = without the straight-line path, 6 = blend_bwd_v6_kernel, 5 = blend_bwd_v5_kernel (round 1)
    int fwd_variant = 4;        // blending forward kernel: 4 = blend_fwd_v4_kernel (packed pairs), 3 = blend_fwd_v3_kernel
    int lpt = 1;                // blending kernels take their tiles longest list first
    int fwd_minb = 42;          // forward blending kernel: v4: 42 = four CTAs per SM (64 registers), pair loop unrolled twice; 4, 5, 6 = that many
                                // CTAs per SM, no unrolling; v3: 3, 4 or 5, 52 = trips of 2 at 5.  In the pipelined step, ms per view: 5: 1.995,
                                // 6: 1.992, 42: 1.974-1.979; measured and dropped: 62: 1.984, 52: 1.980, 44: 1.975-2.01, 32: 2.014, 34: 2.003
                                // (profiles/round2_zl_ab_fwd_unroll.json)
    int bwd_minb = 3;           // same for the backward (v5: 3 or 4; v6: 3)
    int side_stream = 1;        // tile-count kernels on a side stream next to the depth sort
    int bwd_side = 1;           // SH backward on a side stream next to the cov/projection backward
    int sh_split = 1;           // dedicated M = 16 SH backward kernel
    int pre_stage = 1;          // preprocess moves the SH rows of a warp's 32 Gaussians through shared memory (coalesced cp.async)
    int radix_split = 1;        // block shape of the tile-split radix passes: 0 = 512 threads x 8 items, 1 = 256 x 16
    int radix_depth = 3;        // depth-sort passes: 0 / 1 = four 8-bit passes in those block shapes (alone 0.150 vs 0.155 ms; next to the
                                // blending kernels of a pipelined step the smaller blocks find room sooner: 2.042 -> 2.021 ms per view);
                                // 3 = biased keys, three 9-bit passes in 256 x 16 blocks + a fourth only for depths >= 13107 (sort.cu)
    int depth_seg_items = -1;   // >= 0: batched calls sort the depth keys in runs of whole views (comparison arm)
    // comparison arms: honoured only when the test-only library libgsr_b200_arms.so is loaded (else GSR_EINVAL)
    int blend_arm = 0;          // 1: plain blending kernels (no culling)   2: forward staging through cp.async.bulk + mbarrier
    int sort_cub = 0;           // 1: cub::DeviceRadixSort on the reference's 64-bit keys
    int blend_stats = 0;        // 1: run the statistics kernel next to the forward
};
Options& options();

// entry points of the test-only comparison arms (arms/*.cu -> libgsr_b200_arms.so); all null in the product library
struct ArmHooks {
    int (*blend_forward_plain)(long long R, int W, int H, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                               float* out_color, cudaStream_t s);
    int (*blend_backward_plain)(long long R, int W, int H, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                                const float* dL_dpix, float* accum, cudaStream_t s);
    int (*blend_forward_bulk)(long long R, int W, int H, int V, const float* bg, const GeomView& g, const BinView& b, const ImgView& im,
                              float* out_color, cudaStream_t s);
    int (*blend_stats)(long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s);
    int (*tile_split_cub)(int P, long long R, int W, int H, const GeomView& g, const BinView& b, const ImgView& im, cudaStream_t s);
    size_t (*cub_sort_scratch_bytes)(long long R);
};
extern ArmHooks g_arms;
void register_arms(const ArmHooks& h);

// ---------------------------------------------------------------- per-device state (context.cu)
// Everything the library keeps between calls lives here, one instance per CUDA device, created on first use under a
// mutex: side streams with their fork / join events (one set per caller stream, so calls issued from different streams
// do not serialise on a shared side stream), pinned landing slots of the R read-back (one per forward call in flight),
// the gradient fence, and the per-device "function attributes set" flags.  Host threads may call into the library
// concurrently, on the same or on different devices.
struct SideStream {
    cudaStream_t user = nullptr, side = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
struct RSlot {
    uint32_t* host = nullptr;       // 64 pinned bytes
    cudaEvent_t ev = nullptr;
    bool busy = false;
};
struct DeviceCtx {
    int device = -1;
    std::mutex mu;                  // guards the vectors and flags below
    std::vector<SideStream> sides[2];   // 0: forward (tile kernels), 1: backward (SH kernel)
    std::vector<RSlot*> rslots;
    unsigned attrs_set = 0;         // bit per kernel family whose cudaFuncSetAttribute calls were made on this device
    std::mutex fence_mu;            // serialises stage 5 (read-modify-write of the caller's gradient buffers) in host call order
    cudaEvent_t grad_fence = nullptr;
    bool fence_recorded = false;
};
int device_ctx(DeviceCtx** out);                                                   // of the current device
int ctx_side_stream(DeviceCtx* c, int which, cudaStream_t user, SideStream* out);  // created on first use
int ctx_acquire_rslot(DeviceCtx* c, RSlot** out);
void ctx_release_rslot(DeviceCtx* c, RSlot* r);
// runs fn() once per device and kernel family (cudaFuncSetAttribute is per device)
template <class F>
static inline int ctx_once(DeviceCtx* c, unsigned bit, F fn)
{
    std::lock_guard<std::mutex> lk(c->mu);
    if (c->attrs_set & bit) return GSR_OK;
    const int rc = fn();
    if (rc == GSR_OK) c->attrs_set |= bit;
    return rc;
}

}  // namespace gsr
