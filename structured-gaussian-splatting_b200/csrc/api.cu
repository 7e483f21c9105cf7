// api.cu - the C-ABI entry points of include/gsr.h: argument validation, blob carving, stage orchestration.
//
// Replaces the host orchestration of the reference core:
//   gsr_forward   <- CudaRasterizer::Rasterizer::forward   rasterizer_impl.cu:198-336
//   gsr_backward  <- CudaRasterizer::Rasterizer::backward  rasterizer_impl.cu:340-434
//   blob carving  <- GeometryState/ImageState/BinningState::fromChunk, rasterizer_impl.cu:155-194
// Every launch is on the caller's stream (the reference uses the legacy default stream).
#include "gsr_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace gsr {

std::atomic<unsigned long long> g_launches{0};
static thread_local char t_last_cuda_error[256] = "";

int record_cuda_error(cudaError_t e, const char* where)
{
    snprintf(t_last_cuda_error, sizeof t_last_cuda_error, "%s: %s", where, cudaGetErrorString(e));
    cudaGetLastError();   // clear the sticky flag of non-fatal errors
    return GSR_ECUDA;
}

int record_error_text(int code, const char* text)
{
    snprintf(t_last_cuda_error, sizeof t_last_cuda_error, "%s", text);
    return code;
}

// ------------------------------------------------------------------ per-stage event timing
// A measurement facility for bench.py (off by default): one global table guarded by a mutex; the events belong to the
// device that was current when the stage was opened, and a stage is open on one stream at a time per host thread.
static std::atomic<bool> g_timing{false};
struct Pending { int stage; cudaEvent_t a, b; };
static std::mutex g_timing_mu;
static std::vector<Pending> g_pending;
static std::vector<cudaEvent_t> g_pool;
static double g_stage_ms[GSR_NUM_STAGES];
static unsigned long long g_stage_calls[GSR_NUM_STAGES];
static thread_local cudaEvent_t t_open[GSR_NUM_STAGES];

static cudaEvent_t get_event()
{
    {
        std::lock_guard<std::mutex> lk(g_timing_mu);
        if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
void stage_begin(int stage, cudaStream_t s)
{
    if (!g_timing.load(std::memory_order_relaxed)) return;
    t_open[stage] = get_event();
    cudaEventRecord(t_open[stage], s);
}
void stage_end(int stage, cudaStream_t s)
{
    if (!g_timing.load(std::memory_order_relaxed) || !t_open[stage]) return;
    cudaEvent_t b = get_event();
    cudaEventRecord(b, s);
    std::lock_guard<std::mutex> lk(g_timing_mu);
    g_pending.push_back(Pending{stage, t_open[stage], b});
    t_open[stage] = nullptr;
}

// ------------------------------------------------------------------ blob layouts
static size_t geom_fixed_bytes(int P)
{
    const size_t Pp = (size_t)(P > 0 ? P : 1);
    return sizeof(BlobHeader) + align_up(Pp * sizeof(SplatRec)) + align_up(Pp) + align_up(Pp * 4);
}
GeomView geom_view(void* blob, int P)
{
    const size_t Pp = (size_t)(P > 0 ? P : 1);
    GeomView g;
    char* p = (char*)blob;
    g.hdr = (BlobHeader*)p; p += sizeof(BlobHeader);
    g.rec = (SplatRec*)p; p += align_up(Pp * sizeof(SplatRec));
    g.clamped = (uint8_t*)p; p += align_up(Pp);
    g.order = (uint32_t*)p; p += align_up(Pp * 4);
    g.scratch = (uint32_t*)p;
    g.scratch_bytes = geom_scratch_bytes(P);
    return g;
}
BinView bin_view(void* blob, long long R)
{
    const size_t Rp = (size_t)(R > 0 ? R : 1);
    BinView b;
    char* p = (char*)blob;
    b.hdr = (BlobHeader*)p; p += sizeof(BlobHeader);
    b.vals = (uint32_t*)p; p += align_up(Rp * 4);
    b.scratch = (uint32_t*)p;
    b.scratch_bytes = bin_scratch_bytes(R);
    return b;
}
ImgView img_view(void* blob, int W, int H, int V)
{
    // V views are stacked vertically: tile row v * gy + ty belongs to view v (every view is padded to whole tiles)
    const size_t gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (size_t)((H + GSR_TILE - 1) / GSR_TILE) * (V > 0 ? V : 1);
    const size_t tiles = gx * gy;
    const size_t N = (size_t)W * H * (V > 0 ? V : 1);
    ImgView v;
    char* p = (char*)blob;
    v.hdr = (BlobHeader*)p; p += sizeof(BlobHeader);
    v.ranges = (uint2*)p; p += align_up((tiles ? tiles : 1) * sizeof(uint2));
    v.final_T = (float*)p; p += align_up((N ? N : 1) * 4);
    v.n_contrib = (uint32_t*)p; p += align_up((N ? N : 1) * 4);
    v.tile_diff = (int*)p; p += align_up((gx + 1) * (gy / (V > 0 ? V : 1) + 1) * (V > 0 ? V : 1) * sizeof(int));
    v.tile_order = (uint32_t*)p;
    return v;
}

__global__ void init_headers_kernel(BlobHeader* g, BlobHeader* b, BlobHeader* im, int P, int R, int W, int H)
{
    BlobHeader* h[3] = {g, b, im};
    const int k = threadIdx.x;
    if (k < 3) {
        h[k]->magic = GSR_MAGIC; h[k]->kind = k + 1;
        h[k]->P = P; h[k]->R = R; h[k]->W = W; h[k]->H = H;
    }
}

// test hook: rebuild the reference's 64-bit keys from (tile of the range, depth bits of the value)
__global__ void export_keys_kernel(int tiles, const uint2* __restrict__ ranges, const uint32_t* __restrict__ vals,
                                   const SplatRec* __restrict__ rec, unsigned long long* __restrict__ keys)
{
    const int t = blockIdx.x;
    if (t >= tiles) return;
    const uint2 r = ranges[t];
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
        keys[i] = ((unsigned long long)(uint32_t)t << 32) | __float_as_uint(rec[vals[i]].q2.y);
}

}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_geom_bytes(int P) { return geom_fixed_bytes(P) + geom_scratch_bytes(P); }
size_t gsr_img_bytes_batched(int W, int H, int n_views)
{
    const size_t V = n_views > 0 ? n_views : 1;
    const size_t gx = (W + GSR_TILE - 1) / GSR_TILE, gy1 = (size_t)((H + GSR_TILE - 1) / GSR_TILE), gy = gy1 * V;
    const size_t tiles = gx * gy;
    const size_t N = (size_t)W * H * V;
    return sizeof(BlobHeader) + align_up((tiles ? tiles : 1) * sizeof(uint2)) + 2 * align_up((N ? N : 1) * 4) +
           align_up((gx + 1) * (gy1 + 1) * V * sizeof(int)) +        // the difference array has one band per view
           align_up((tiles ? tiles : 1) * sizeof(uint32_t));         // tile launch order
}
size_t gsr_img_bytes(int W, int H) { return gsr_img_bytes_batched(W, H, 1); }
size_t gsr_binning_bytes(int P, int R, int W, int H)
{
    (void)P;
    const size_t Rp = (size_t)(R > 0 ? R : 1);
    size_t scratch = bin_scratch_bytes(R);
    if (options().sort_cub == 1 && g_arms.cub_sort_scratch_bytes) { size_t c = g_arms.cub_sort_scratch_bytes(R); if (c > scratch) scratch = c; }
    const size_t planes = 8 * 4 * blend_plane_words(R, W, H);      // written by the forward blending into the dead sort scratch
    if (planes > scratch) scratch = planes;
    return sizeof(BlobHeader) + align_up(Rp * 4) + scratch;
}
size_t gsr_bwd_accum_bytes(int P) { return (size_t)(P > 0 ? P : 1) * 12 * sizeof(float); }

// V views of ONE Gaussian set in one pass: the views are stacked into a tall image (tile row v * gy + ty), every
// (Gaussian, view) pair is a "virtual Gaussian" v * P + g with its own record, and depth sort, tile split and blending
// run once over all of them.  V = 1 is the reference's single-view call.
// S > 1: S scenes of P Gaussians each (parameter arrays [S][P][...]) with V / S cameras per scene: preprocess runs once per scene
// on that scene's parameters and cameras, everything after it once over all S * (V / S) * P virtual Gaussians.
// front_ (gsr_forward_front): the stream of everything before the blending; the blending stays on stream_.
// prepared (gsr_forward_prepared): a geometry blob gsr_preprocess_views has already filled - stage 0 is skipped.
static int forward_impl(const GsrFwdArgs* a, int V, int S, GsrAlloc geom, GsrAlloc binning, GsrAlloc img, gsr_stream_t stream_,
                        int* num_rendered, gsr_stream_t front_ = nullptr, void* prepared = nullptr)
{
    cudaStream_t const blend_s = (cudaStream_t)stream_;
    cudaStream_t s = front_ ? (cudaStream_t)front_ : blend_s;       // the front's stream until the blending starts
    if (!a || !num_rendered || (!geom.fn && !prepared) || !binning.fn || !img.fn) return GSR_EINVAL;
    if (prepared && (V != 1 || S != 1 || a->colors_precomp || a->cov3D_precomp)) return GSR_EINVAL;
    *num_rendered = 0;
    if (a->P < 0 || a->W <= 0 || a->H <= 0 || a->D < 0 || a->D > 3 || V < 1 || S < 1 || V % S != 0) return GSR_EINVAL;
    if (!a->out_color || !a->background || !a->viewmatrix || !a->projmatrix) return GSR_EINVAL;
    const int gx = (a->W + GSR_TILE - 1) / GSR_TILE, gy = (a->H + GSR_TILE - 1) / GSR_TILE;
    // P * V < 2^30: the look-back status words of the depth sort keep 30 value bits (sort.cu ST_VAL)
    if (gx > 65535 || (long long)gy * V > 65535 || (long long)a->P * V >= (1ll << 30) || V > 4095) return GSR_ETOOBIG;
    const int Pv = a->P * V;                    // virtual Gaussians
    const int gyT = gy * V, HT = gyT * GSR_TILE; // the tall image in whole tiles
    if (a->P == 0) {
        // reference binding returns the zero-filled image without calling the core (rasterize_points.cu:81)
        GSR_CUDA(cudaMemsetAsync(a->out_color, 0, sizeof(float) * 3 * (size_t)a->W * a->H * V, blend_s));
        return GSR_OK;
    }
    if (!a->means3D || !a->opacities || (!a->radii && !prepared)) return GSR_EINVAL;
    if (!a->shs && !a->colors_precomp) return GSR_ECHANNELS;
    if (a->shs && !a->colors_precomp && (!a->cam_pos || a->M < (a->D + 1) * (a->D + 1))) return GSR_EINVAL;
    if (a->shs && a->shs_rest && a->M != 16 && a->M != 1) return GSR_EINVAL;       // split SH storage: M = 16 (or DC only)
    if (!a->cov3D_precomp && (!a->scales || !a->rotations)) return GSR_ECOV;

    void* gblob = prepared ? prepared : geom.fn(geom.ctx, gsr_geom_bytes(Pv));
    void* iblob = img.fn(img.ctx, gsr_img_bytes_batched(a->W, a->H, V));
    if (!gblob || !iblob) return GSR_ENOMEM;
    GeomView g = geom_view(gblob, Pv);
    ImgView im = img_view(iblob, a->W, a->H, V);
    const GeomScratch sc = carve_geom_scratch(g.scratch, Pv);
    if (!prepared) {
        GSR_CUDA(cudaMemsetAsync(g.hdr, 0, sizeof(BlobHeader), s));
        GSR_CUDA(cudaMemsetAsync(sc.ctl, 0, sc.ctl_words * 4, s));       // counters, histograms, look-back words
    }
    GSR_CUDA(cudaMemsetAsync(im.tile_diff, 0, sizeof(int) * (size_t)(gx + 1) * (gy + 1) * V, s));

    int rc;
    if (!prepared) {
    stage_begin(0, s);
    {
        const int Vs = V / S;
        const int Mdc = (a->shs_rest && a->M > 1) ? 1 : a->M;
        for (int sc_i = 0; sc_i < S; sc_i++) {
            GsrFwdArgs as = *a;
            const size_t o = (size_t)sc_i * a->P;
            as.means3D = a->means3D + 3 * o; as.opacities = a->opacities + o;
            if (a->shs) as.shs = a->shs + 3 * (size_t)Mdc * o;
            if (a->shs_rest) as.shs_rest = a->shs_rest + 3 * (size_t)(a->M - 1) * o;
            if (a->colors_precomp) as.colors_precomp = a->colors_precomp + 3 * o;
            if (a->scales) as.scales = a->scales + 3 * o;
            if (a->rotations) as.rotations = a->rotations + 4 * o;
            if (a->cov3D_precomp) as.cov3D_precomp = a->cov3D_precomp + 6 * o;
            as.viewmatrix = a->viewmatrix + 16 * (size_t)sc_i * Vs; as.projmatrix = a->projmatrix + 16 * (size_t)sc_i * Vs;
            if (a->cam_pos) as.cam_pos = a->cam_pos + 3 * (size_t)sc_i * Vs;
            if ((rc = launch_preprocess(as, Vs, sc_i * Vs, V, g, s))) return rc;
        }
    }
    stage_end(0, s);
    GSR_STAGE(a->debug, s, "preprocess");
    }

    // R (sum of tiles touched) is accumulated by preprocess: start its read-back now and wait for it only after
    // the depth sort has been queued, so the host's launch work overlaps the device
    // (every call in flight has a pinned slot and an event of its own, taken from the device context)
    DeviceCtx* ctx = nullptr;
    if ((rc = device_ctx(&ctx))) return rc;
    RSlot* slot = nullptr;
    if ((rc = ctx_acquire_rslot(ctx, &slot))) return rc;
    struct SlotGuard { DeviceCtx* c; RSlot* r; ~SlotGuard() { ctx_release_rslot(c, r); } } slot_guard{ctx, slot};
    GSR_CUDA(cudaMemcpyAsync(slot->host, sc.counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    GSR_CUDA(cudaEventRecord(slot->ev, s));

    stage_begin(1, s);
    {
        // views never share a tile, so the depth order is only needed inside a view: GSR_DEPTH_SEG_ITEMS = n sorts runs of
        // whole views of at most n items instead of all V * P items at once.  Off by default: measured at C3 with 8 views
        // (24 M items) one sort takes 0.119 ms per view, two runs of 12 M 0.122, eight runs of 3 M 0.151.
        int seg = 0;
        if (V > 1 && (a->P & 3) == 0 && options().depth_seg_items >= 0) {
            const long long per = (long long)options().depth_seg_items / a->P;     // 0 items = one view per run
            const long long views_per_seg = per < 1 ? 1 : per;
            if (views_per_seg < V) seg = (int)(views_per_seg * a->P);
        }
        if ((rc = launch_depth_order(Pv, a->W, HT, g, im, s, seg, V))) return rc;
    }
    stage_end(1, s);
    GSR_STAGE(a->debug, s, "depth_order");

    // the binning blob is sized from R: the one blocking wait of the forward, like rasterizer_impl.cu:280-281
    GSR_CUDA(cudaEventSynchronize(slot->ev));
    const uint32_t R_host = *slot->host;
    if (R_host >= 0x40000000u) return GSR_ETOOBIG;
    const int R = (int)R_host;
    *num_rendered = R;

    void* bblob = binning.fn(binning.ctx, gsr_binning_bytes(Pv, R, a->W, HT));
    if (!bblob) return GSR_ENOMEM;
    BinView b = bin_view(bblob, R);
    GSR_LAUNCH(init_headers_kernel, 1, 32, 0, s, g.hdr, b.hdr, im.hdr, Pv, R, a->W, a->H);

    stage_begin(2, s);
    if (options().sort_cub == 1) rc = g_arms.tile_split_cub ? g_arms.tile_split_cub(Pv, R, a->W, HT, g, b, im, s) : GSR_EINVAL;
    else rc = launch_tile_split(Pv, R, a->W, HT, g, b, im, s);
    if (rc) return rc;
    stage_end(2, s);
    GSR_STAGE(a->debug, s, "tile_split");

    if (s != blend_s) {
        // two-stream form: the blending stream waits for the front (the call's own event is free again after the wait above)
        GSR_CUDA(cudaEventRecord(slot->ev, s));
        GSR_CUDA(cudaStreamWaitEvent(blend_s, slot->ev, 0));
        s = blend_s;
    }
    stage_begin(3, s);
    if ((rc = launch_blend_forward(R, a->W, a->H, V, a->background, g, b, im, a->out_color, s))) return rc;
    stage_end(3, s);
    GSR_STAGE(a->debug, s, "blend_forward");
    if (a->debug && a->prefiltered) {
        int err = 0;
        GSR_CUDA(cudaMemcpy(&err, sc.counters + 8, sizeof(int), cudaMemcpyDeviceToHost));
        if (err) return GSR_EINVAL;
    }
    return GSR_OK;
}

int gsr_forward(const GsrFwdArgs* a, GsrAlloc geom, GsrAlloc binning, GsrAlloc img, gsr_stream_t stream, int* num_rendered)
{
    return forward_impl(a, 1, 1, geom, binning, img, stream, num_rendered);
}
int gsr_forward_front(const GsrFwdArgs* a, GsrAlloc geom, GsrAlloc binning, GsrAlloc img, gsr_stream_t front_stream, gsr_stream_t stream,
                      int* num_rendered)
{
    return forward_impl(a, 1, 1, geom, binning, img, stream, num_rendered, front_stream == stream ? nullptr : front_stream);
}
int gsr_preprocess_views(const GsrFwdArgs* a, int n_views, const GsrViewFront* views, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!a || !views || n_views < 1 || n_views > GSR_MAX_VIEWS) return GSR_EINVAL;
    if (a->P < 0 || a->W <= 0 || a->H <= 0 || a->D < 0 || a->D > 3) return GSR_EINVAL;
    const int gx = (a->W + GSR_TILE - 1) / GSR_TILE, gy = (a->H + GSR_TILE - 1) / GSR_TILE;
    if (gx > 65535 || gy > 65535 || (long long)a->P >= (1ll << 30)) return GSR_ETOOBIG;
    if (a->P == 0) return GSR_OK;
    if (!a->means3D || !a->opacities || !a->scales || !a->rotations || a->colors_precomp || a->cov3D_precomp) return GSR_EINVAL;
    if (!a->shs) return GSR_ECHANNELS;
    if (a->M < (a->D + 1) * (a->D + 1)) return GSR_EINVAL;
    if (a->shs_rest && a->M != 16 && a->M != 1) return GSR_EINVAL;
    GeomView g[GSR_MAX_VIEWS];
    for (int v = 0; v < n_views; v++) {
        if (!views[v].viewmatrix || !views[v].projmatrix || !views[v].cam_pos || !views[v].radii || !views[v].geom_blob) return GSR_EINVAL;
        g[v] = geom_view(views[v].geom_blob, a->P);
        const GeomScratch sc = carve_geom_scratch(g[v].scratch, a->P);
        GSR_CUDA(cudaMemsetAsync(g[v].hdr, 0, sizeof(BlobHeader), s));
        GSR_CUDA(cudaMemsetAsync(sc.ctl, 0, sc.ctl_words * 4, s));       // counters, histograms, look-back words
    }
    stage_begin(0, s);
    int rc = launch_preprocess_views(*a, n_views, views, g, s);
    if (rc) return rc;
    stage_end(0, s);
    GSR_STAGE(a->debug, s, "preprocess");
    return GSR_OK;
}
int gsr_forward_prepared(const GsrFwdArgs* a, void* prepared_geom_blob, GsrAlloc binning, GsrAlloc img, gsr_stream_t front_stream,
                         gsr_stream_t stream, int* num_rendered)
{
    if (!prepared_geom_blob) return GSR_EBLOB;
    return forward_impl(a, 1, 1, GsrAlloc{nullptr, nullptr}, binning, img, stream, num_rendered,
                        (front_stream == stream) ? nullptr : front_stream, prepared_geom_blob);
}
int gsr_forward_batched(const GsrFwdArgs* a, int n_views, GsrAlloc geom, GsrAlloc binning, GsrAlloc img, gsr_stream_t stream,
                        int* num_rendered)
{
    if (n_views > 1 && a && (a->colors_precomp || a->cov3D_precomp)) return GSR_EINVAL;   // per-view accumulation of their gradients is not wired
    return forward_impl(a, n_views, 1, geom, binning, img, stream, num_rendered);
}
int gsr_forward_scenes(const GsrFwdArgs* a, int n_scenes, int views_per_scene, GsrAlloc geom, GsrAlloc binning, GsrAlloc img,
                       gsr_stream_t stream, int* num_rendered)
{
    if (n_scenes < 1 || views_per_scene < 1) return GSR_EINVAL;
    if (a && (a->colors_precomp || a->cov3D_precomp) && (long long)n_scenes * views_per_scene > 1) return GSR_EINVAL;
    if ((long long)n_scenes * views_per_scene > 65535) return GSR_ETOOBIG;
    return forward_impl(a, n_scenes * views_per_scene, n_scenes, geom, binning, img, stream, num_rendered);
}

// stages: bit 0 = blending backward (fills the accumulator rows of all Gaussians), bit 1 = per-Gaussian backward.
// [g_begin, g_end) restricts the per-Gaussian backward to a range of Gaussians (single-view, single-scene calls only): a
// data-parallel trainer runs the last view's per-Gaussian backward range by range and starts the all-reduce of each range's
// gradients while the next range is still being computed (gsr_backward_staged).
static int backward_impl(const GsrBwdArgs* a, int V, int S, gsr_stream_t stream_, int stages = 3, int g_begin = 0, int g_end = -1)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!a) return GSR_EINVAL;
    if (a->P < 0 || a->W <= 0 || a->H <= 0 || a->R < 0 || a->D < 0 || a->D > 3 || V < 1 || S < 1 || V % S != 0) return GSR_EINVAL;
    if (a->P == 0) return GSR_OK;
    if (!a->means3D || !a->radii || !a->dL_dpix || !a->background || !a->viewmatrix || !a->projmatrix) return GSR_EINVAL;
    if (!a->geom_blob || !a->binning_blob || !a->img_blob || !a->accum) return GSR_EBLOB;
    if (stages & 2) {          // the gradient outputs belong to the per-Gaussian backward
        if (!a->dL_dmean2D || !a->dL_dopacity || !a->dL_dmean3D) return GSR_EINVAL;
        if (a->shs && (!a->dL_dsh || !a->cam_pos)) return GSR_EINVAL;
        if (a->shs && a->shs_rest && a->M > 1 && (a->M != 16 || !a->dL_dsh_rest)) return GSR_EINVAL;
        if (!a->cov3D_precomp && (!a->scales || !a->rotations || !a->dL_dscale || !a->dL_drot)) return GSR_ECOV;
    }

    if (g_end < 0) g_end = a->P;
    if (g_begin < 0 || g_end > a->P || g_begin > g_end || (stages & ~3) != 0) return GSR_EINVAL;
    const bool ranged = g_begin != 0 || g_end != a->P;
    if (ranged && (V != 1 || S != 1 || (g_begin & 127) != 0)) return GSR_EINVAL;   // ranges start on multiples of 128 (alignment of the SH rows)
    const int Pv = a->P * V;
    GeomView g = geom_view(const_cast<void*>(a->geom_blob), Pv);
    BinView b = bin_view(const_cast<void*>(a->binning_blob), a->R);
    ImgView im = img_view(const_cast<void*>(a->img_blob), a->W, a->H, V);
    if (a->debug) {
        BlobHeader h[3];
        GSR_CUDA(cudaMemcpyAsync(&h[0], g.hdr, sizeof(BlobHeader), cudaMemcpyDeviceToHost, s));
        GSR_CUDA(cudaMemcpyAsync(&h[1], b.hdr, sizeof(BlobHeader), cudaMemcpyDeviceToHost, s));
        GSR_CUDA(cudaMemcpyAsync(&h[2], im.hdr, sizeof(BlobHeader), cudaMemcpyDeviceToHost, s));
        GSR_CUDA(cudaStreamSynchronize(s));
        for (int k = 0; k < 3; k++)
            if (h[k].magic != GSR_MAGIC || h[k].kind != (uint32_t)(k + 1) || h[k].P != Pv || h[k].R != a->R || h[k].W != a->W ||
                h[k].H != a->H)
                return GSR_EBLOB;
    }
    float* accum = (float*)a->accum;
    int rc;
    if (stages & 1) {
        stage_begin(4, s);
        if (!a->accum_zeroed) GSR_CUDA(cudaMemsetAsync(accum, 0, gsr_bwd_accum_bytes(Pv), s));
        if ((rc = launch_blend_backward(a->R, a->W, a->H, V, a->background, g, b, im, a->dL_dpix, accum, s))) return rc;
        stage_end(4, s);
        GSR_STAGE(a->debug, s, "blend_backward");
    }
    if (!(stages & 2) || g_begin == g_end) return GSR_OK;
    stage_begin(5, s);
    // Gradient fence: the per-Gaussian kernels read-modify-write the caller's gradient buffers when they accumulate.  Calls
    // issued from different streams (a trainer keeping two views in flight, gsr_b200.dp.backward_views(n_streams=2)) are
    // ordered here, in host call order, through one event: everything before this point (blending) still overlaps.
    // The fence belongs to the device; its mutex keeps wait -> launches -> record of one call together when several host
    // threads drive the same device.
    DeviceCtx* ctx = nullptr;
    if ((rc = device_ctx(&ctx))) return rc;
    std::lock_guard<std::mutex> fence_lock(ctx->fence_mu);
    if (ctx->fence_recorded) GSR_CUDA(cudaStreamWaitEvent(s, ctx->grad_fence, 0));
    // scene by scene (S = 1 for the single-scene calls): the per-Gaussian backward of scene sc_i sees its own parameters,
    // cameras and gradient rows and the slice of the per-view arrays that belongs to its Vs views
    const int Vs = V / S;
    const int Mdc = (a->shs_rest && a->M > 1) ? 1 : a->M;
    for (int sc_i = 0; sc_i < S; sc_i++) {
        GsrBwdArgs as = *a;
        GeomView gs = g;
        as.P = g_end - g_begin;                            // (== a->P unless a range was asked for)
        const size_t o = (size_t)sc_i * a->P + (size_t)g_begin;     // first Gaussian of the scene (of the range)
        const size_t ov = o * Vs;                          // first virtual Gaussian of the scene (= sc_i * Vs views * P)
        as.means3D = a->means3D + 3 * o;
        if (a->shs) { as.shs = a->shs + 3 * (size_t)Mdc * o; as.dL_dsh = a->dL_dsh + 3 * (size_t)Mdc * o; }
        if (a->shs_rest) { as.shs_rest = a->shs_rest + 3 * (size_t)(a->M - 1) * o; if (a->dL_dsh_rest) as.dL_dsh_rest = a->dL_dsh_rest + 3 * (size_t)(a->M - 1) * o; }
        if (a->colors_precomp) as.colors_precomp = a->colors_precomp + 3 * o;
        if (a->scales) as.scales = a->scales + 3 * o;
        if (a->rotations) as.rotations = a->rotations + 4 * o;
        if (a->cov3D_precomp) as.cov3D_precomp = a->cov3D_precomp + 6 * o;
        as.viewmatrix = a->viewmatrix + 16 * (size_t)sc_i * Vs; as.projmatrix = a->projmatrix + 16 * (size_t)sc_i * Vs;
        if (a->cam_pos) as.cam_pos = a->cam_pos + 3 * (size_t)sc_i * Vs;
        as.radii = a->radii + ov;
        as.dL_dmean2D = a->dL_dmean2D + 3 * ov;
        as.dL_dopacity = a->dL_dopacity + o; as.dL_dmean3D = a->dL_dmean3D + 3 * o;
        if (a->dL_dcolor) as.dL_dcolor = a->dL_dcolor + 3 * o;
        if (a->dL_dcov3D) as.dL_dcov3D = a->dL_dcov3D + 6 * o;
        if (a->dL_dscale) as.dL_dscale = a->dL_dscale + 3 * o;
        if (a->dL_drot) as.dL_drot = a->dL_drot + 4 * o;
        gs.rec = g.rec + ov; gs.clamped = g.clamped + ov;
        float* accum_s = accum + 12 * ov;
        if (Vs <= GSR_MAX_VIEWS) {
            // one launch pair over the scene's views: the per-view contributions are summed in registers (preprocess_bwd.cu)
            if ((rc = launch_preprocess_backward(as, Vs, gs, accum_s, s))) return rc;
        } else {
            for (int v = 0; v < Vs; v++) {
                // view by view over its slice of the virtual Gaussians; the views after the first add to what the first wrote
                GsrBwdArgs av = as;
                GeomView gv = gs;
                const size_t ovv = (size_t)v * a->P;
                gv.rec = gs.rec + ovv; gv.clamped = gs.clamped + ovv;
                av.radii = as.radii + ovv;
                av.viewmatrix = as.viewmatrix + 16 * v; av.projmatrix = as.projmatrix + 16 * v;
                if (as.cam_pos) av.cam_pos = as.cam_pos + 3 * v;
                av.dL_dmean2D = as.dL_dmean2D + 3 * ovv;
                if (v > 0) av.accumulate = GSR_ACC_MEAN3D | GSR_ACC_OPACITY | GSR_ACC_SH | GSR_ACC_SCALE | GSR_ACC_ROT;
                if ((rc = launch_preprocess_backward(av, 1, gv, accum_s + 12 * ovv, s))) return rc;
            }
        }
    }
    GSR_CUDA(cudaEventRecord(ctx->grad_fence, s));
    ctx->fence_recorded = true;
    stage_end(5, s);
    GSR_STAGE(a->debug, s, "preprocess_backward");
    return GSR_OK;
}

int gsr_backward(const GsrBwdArgs* a, gsr_stream_t stream) { return backward_impl(a, 1, 1, stream); }
int gsr_backward_staged(const GsrBwdArgs* a, int stages, int g_begin, int g_end, gsr_stream_t stream)
{
    return backward_impl(a, 1, 1, stream, stages, g_begin, g_end);
}
// Per-Gaussian backward of n_views single-view calls at once (the deferred form of a multi-view step): every view brings
// the blobs of its own forward and the accumulator rows its blending backward filled (gsr_backward_staged, stages = 1); the
// parameters are read once and every gradient row is written once for all views, instead of once per view.
int gsr_backward_views(const GsrBwdArgs* a, int n_views, const GsrViewState* views, int g_begin, int g_end, gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!a || !views || n_views < 1 || n_views > GSR_MAX_VIEWS) return GSR_EINVAL;
    if (a->P < 0 || a->W <= 0 || a->H <= 0 || a->D < 0 || a->D > 3) return GSR_EINVAL;
    if (a->P == 0) return GSR_OK;
    if (a->colors_precomp || a->cov3D_precomp) return GSR_EINVAL;           // like the batched entry points
    if (!a->means3D || !a->shs || !a->scales || !a->rotations) return GSR_EINVAL;
    if (!a->dL_dopacity || !a->dL_dmean3D || !a->dL_dsh || !a->dL_dscale || !a->dL_drot) return GSR_EINVAL;
    if (a->shs_rest && a->M > 1 && (a->M != 16 || !a->dL_dsh_rest)) return GSR_EINVAL;
    if (g_end < 0) g_end = a->P;
    if (g_begin < 0 || g_end > a->P || g_begin > g_end || (g_begin & 127) != 0) return GSR_EINVAL;
    if (g_begin == g_end) return GSR_OK;
    const size_t o = (size_t)g_begin;
    const int Mdc = (a->shs_rest && a->M > 1) ? 1 : a->M;
    ViewTab vt = {};
    for (int v = 0; v < n_views; v++) {
        const GsrViewState& w = views[v];
        if (!w.viewmatrix || !w.projmatrix || !w.cam_pos || !w.radii || !w.geom_blob || !w.accum || !w.dL_dmean2D) return GSR_EINVAL;
        const GeomView gv = geom_view(const_cast<void*>(w.geom_blob), a->P);
        vt.radii[v] = w.radii + o; vt.rec[v] = gv.rec + o; vt.clamped[v] = gv.clamped + o;
        vt.accum[v] = (const float*)w.accum + 12 * o; vt.m2d[v] = w.dL_dmean2D + 3 * o;
        vt.view[v] = w.viewmatrix; vt.proj[v] = w.projmatrix; vt.cam[v] = w.cam_pos;
    }
    GsrBwdArgs as = *a;
    as.P = g_end - g_begin;
    as.means3D = a->means3D + 3 * o;
    as.shs = a->shs + 3 * (size_t)Mdc * o; as.dL_dsh = a->dL_dsh + 3 * (size_t)Mdc * o;
    if (a->shs_rest) { as.shs_rest = a->shs_rest + 3 * (size_t)(a->M - 1) * o; if (a->dL_dsh_rest) as.dL_dsh_rest = a->dL_dsh_rest + 3 * (size_t)(a->M - 1) * o; }
    as.scales = a->scales + 3 * o; as.rotations = a->rotations + 4 * o;
    as.dL_dopacity = a->dL_dopacity + o; as.dL_dmean3D = a->dL_dmean3D + 3 * o;
    as.dL_dscale = a->dL_dscale + 3 * o; as.dL_drot = a->dL_drot + 4 * o;
    as.dL_dcolor = nullptr; as.dL_dcov3D = nullptr;
    as.viewmatrix = views[0].viewmatrix; as.projmatrix = views[0].projmatrix; as.cam_pos = views[0].cam_pos;
    as.radii = vt.radii[0]; as.dL_dmean2D = vt.m2d[0];
    GeomView g0v = {};
    g0v.rec = const_cast<SplatRec*>(vt.rec[0]); g0v.clamped = const_cast<uint8_t*>(vt.clamped[0]);
    int rc;
    DeviceCtx* ctx = nullptr;
    if ((rc = device_ctx(&ctx))) return rc;
    stage_begin(5, s);
    std::lock_guard<std::mutex> fence_lock(ctx->fence_mu);
    if (ctx->fence_recorded) GSR_CUDA(cudaStreamWaitEvent(s, ctx->grad_fence, 0));
    if ((rc = launch_preprocess_backward(as, n_views, g0v, vt.accum[0], s, n_views > 1 ? &vt : nullptr))) return rc;
    GSR_CUDA(cudaEventRecord(ctx->grad_fence, s));
    ctx->fence_recorded = true;
    stage_end(5, s);
    GSR_STAGE(a->debug, s, "preprocess_backward (views)");
    return GSR_OK;
}

int gsr_backward_batched(const GsrBwdArgs* a, int n_views, gsr_stream_t stream)
{
    if (n_views > 1 && a && (a->colors_precomp || a->cov3D_precomp)) return GSR_EINVAL;
    return backward_impl(a, n_views, 1, stream);
}
int gsr_backward_scenes(const GsrBwdArgs* a, int n_scenes, int views_per_scene, gsr_stream_t stream)
{
    if (n_scenes < 1 || views_per_scene < 1 || (long long)n_scenes * views_per_scene > 65535) return GSR_EINVAL;
    if (a && (a->colors_precomp || a->cov3D_precomp) && (long long)n_scenes * views_per_scene > 1) return GSR_EINVAL;
    return backward_impl(a, n_scenes * views_per_scene, n_scenes, stream);
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     gsr_stream_t stream)
{
    (void)projmatrix;
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return GSR_EINVAL;
    if (P == 0) return GSR_OK;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

int gsr_dist2(int P, const float* points, float* out, GsrAlloc scratch, gsr_stream_t stream)
{
    if (P < 0 || (P > 0 && (!points || !out || !scratch.fn))) return GSR_EINVAL;
    if (P == 0) return GSR_OK;
    stage_begin(6, (cudaStream_t)stream);
    int rc = launch_dist2(P, points, out, scratch, (cudaStream_t)stream);
    stage_end(6, (cudaStream_t)stream);
    return rc;
}

int gsr_export_binning(const void* geom_blob, const void* binning_blob, const void* img_blob, int P, int R, int W, int H,
                       uint64_t* keys, uint32_t* values, uint32_t* ranges, uint32_t* n_contrib, float* final_T,
                       gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (!geom_blob || !binning_blob || !img_blob) return GSR_EBLOB;
    GeomView g = geom_view(const_cast<void*>(geom_blob), P);
    BinView b = bin_view(const_cast<void*>(binning_blob), R);
    ImgView im = img_view(const_cast<void*>(img_blob), W, H);
    const int tiles = ((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE);
    const size_t N = (size_t)W * H;
    if (keys && R > 0) GSR_LAUNCH(export_keys_kernel, tiles, 128, 0, s, tiles, im.ranges, b.vals, g.rec, (unsigned long long*)keys);
    if (values && R > 0) GSR_CUDA(cudaMemcpyAsync(values, b.vals, 4 * (size_t)R, cudaMemcpyDeviceToDevice, s));
    if (ranges) GSR_CUDA(cudaMemcpyAsync(ranges, im.ranges, 8 * (size_t)tiles, cudaMemcpyDeviceToDevice, s));
    if (n_contrib) GSR_CUDA(cudaMemcpyAsync(n_contrib, im.n_contrib, 4 * N, cudaMemcpyDeviceToDevice, s));
    if (final_T) GSR_CUDA(cudaMemcpyAsync(final_T, im.final_T, 4 * N, cudaMemcpyDeviceToDevice, s));
    return GSR_OK;
}

const char* gsr_strerror(int code)
{
    switch (code) {
    case GSR_OK: return "ok";
    case GSR_EINVAL: return "invalid argument";
    case GSR_ECUDA: return "CUDA error";
    case GSR_ENOMEM: return "scratch allocator returned NULL";
    case GSR_ECHANNELS: return "For non-RGB, provide precomputed Gaussian colors!";
    case GSR_ECOV: return "provide either scales+rotations or a precomputed 3D covariance";
    case GSR_EBLOB: return "state blob does not belong to this forward call";
    case GSR_ETOOBIG: return "image or instance count too large";
    default: return "unknown error";
    }
}
const char* gsr_last_cuda_error(void) { return t_last_cuda_error; }
int gsr_abi_version(void) { return GSR_ABI_VERSION; }

void gsr_timing_enable(int on) { g_timing.store(on != 0); }
void gsr_timing_reset(void)
{
    std::lock_guard<std::mutex> lk(g_timing_mu);
    for (int i = 0; i < GSR_NUM_STAGES; i++) { g_stage_ms[i] = 0; g_stage_calls[i] = 0; }
}
int gsr_timing_collect(void)
{
    std::vector<Pending> todo;
    {
        std::lock_guard<std::mutex> lk(g_timing_mu);
        todo.swap(g_pending);
    }
    for (auto& p : todo) {
        GSR_CUDA(cudaEventSynchronize(p.b));
        float ms = 0.f;
        GSR_CUDA(cudaEventElapsedTime(&ms, p.a, p.b));
        std::lock_guard<std::mutex> lk(g_timing_mu);
        g_stage_ms[p.stage] += ms; g_stage_calls[p.stage]++;
        g_pool.push_back(p.a); g_pool.push_back(p.b);
    }
    return GSR_OK;
}
int gsr_timing_read(int stage, double* ms_total, unsigned long long* calls)
{
    if (stage < 0 || stage >= GSR_NUM_STAGES) return GSR_EINVAL;
    std::lock_guard<std::mutex> lk(g_timing_mu);
    if (ms_total) *ms_total = g_stage_ms[stage];
    if (calls) *calls = g_stage_calls[stage];
    return GSR_OK;
}
const char* gsr_stage_name(int stage)
{
    static const char* n[GSR_NUM_STAGES] = {"preprocess", "depth_order", "tile_split", "blend_forward", "blend_backward",
                                            "preprocess_backward", "dist2"};
    return stage >= 0 && stage < GSR_NUM_STAGES ? n[stage] : "?";
}
void gsr_reset_launch_count(void) { g_launches.store(0); }
unsigned long long gsr_launch_count(void) { return g_launches.load(); }

}  // extern "C"
