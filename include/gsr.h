/*
 * gsr.h - C ABI of the B200-native Gaussian-splat rasterizer hot path.
 *
 * This is the drop-in boundary.  Every entry point replaces one static method /
 * free function of the reference's C++ core, which the reference's torch binding
 * (submodules/diff-gaussian-rasterization/rasterize_points.cu, simple-knn/spatial.cu)
 * calls with raw device pointers.  Signatures are plain C: pointers, sizes, POD
 * structs; no torch / C++ types cross this boundary.  All pointers are DEVICE
 * pointers unless a comment says "host".  Every call is asynchronous on `stream`
 * except where stated; nothing is ever launched on the legacy default stream.
 *
 * Return value: 0 on success, a negative GSR_E* code otherwise (never throws).
 * gsr_strerror() maps a code to text; gsr_last_cuda_error() gives the CUDA
 * error string behind GSR_ECUDA (kept per host thread).
 *
 * Threads and devices: like the reference core (static methods without state,
 * rasterizer.h:24-84) every entry point may be called from several host threads
 * at once, on the same or on different CUDA devices; a call works on the device
 * that is current in the calling thread (the pointers must belong to it).  The
 * library keeps only reusable CUDA resources between calls (side streams, events,
 * pinned read-back slots), one set per device, created on first use.  Gradient
 * accumulation into caller-owned buffers (GsrBwdArgs.accumulate) from concurrent
 * calls on one device is ordered in host call order.  No environment variable is
 * read after the first call into the library.
 *
 * reference interface replaced (file:line, relative to /root/reference/submodules):
 *   gsr_forward        <- CudaRasterizer::Rasterizer::forward      diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:31-57,  rasterizer_impl.cu:198-336
 *   gsr_backward       <- CudaRasterizer::Rasterizer::backward     rasterizer.h:59-84,  rasterizer_impl.cu:340-434
 *   gsr_mark_visible   <- CudaRasterizer::Rasterizer::markVisible  rasterizer.h:24-29,  rasterizer_impl.cu:141-153
 *   gsr_dist2          <- SimpleKNN::knn                           simple-knn/simple_knn.h:17-19, simple_knn.cu:185-221
 *   gsr_export_binning <- (test hook) layout of BinningState/ImageState, rasterizer_impl.cu:172-194
 */
#ifndef GSR_B200_H_INCLUDED
#define GSR_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_ABI_VERSION 5

/* error codes */
#define GSR_OK          0
#define GSR_EINVAL     -1   /* bad argument (null required pointer, negative size, ...) */
#define GSR_ECUDA      -2   /* a CUDA runtime call or kernel failed; see gsr_last_cuda_error() */
#define GSR_ENOMEM     -3   /* an allocator callback returned NULL */
#define GSR_ECHANNELS  -4   /* neither shs nor colors_precomp given (reference: "For non-RGB, provide precomputed Gaussian colors!") */
#define GSR_ECOV       -5   /* neither (scales, rotations) nor cov3D_precomp given */
#define GSR_EBLOB      -6   /* a state blob handed to backward does not carry our magic / sizes */
#define GSR_ETOOBIG    -7   /* image has more than 65535 tiles per axis, P * n_views >= 2^30 or R >= 2^30 */

/* CUDA stream handle (== cudaStream_t) without dragging in cuda headers. */
typedef struct CUstream_st* gsr_stream_t;

/*
 * Scratch allocator callback: must return a device pointer to at least `bytes`
 * bytes, 256-byte aligned, that stays valid until the caller (the binding) frees
 * it - the three state blobs of gsr_forward are saved for gsr_backward.
 * Mirrors std::function<char*(size_t)> of rasterizer.h:32-34 (implemented in the
 * reference binding by torch::Tensor::resize_, rasterize_points.cu:27-33).
 * Each callback is invoked exactly once per gsr_forward call.
 */
typedef void* (*gsr_alloc_fn)(void* ctx, size_t bytes);
typedef struct GsrAlloc { gsr_alloc_fn fn; void* ctx; } GsrAlloc;

typedef struct GsrFwdArgs {
    int P;                       /* number of Gaussians */
    int D;                       /* active SH degree 0..3 */
    int M;                       /* SH coefficients per Gaussian in memory ((Dmax+1)^2), 0 if no shs */
    int W, H;                    /* image size in pixels */
    const float* background;     /* [3] */
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3]   or NULL (exactly one of shs / colors_precomp) */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    float        scale_modifier;
    const float* rotations;      /* [P,4] (r,x,y,z), used un-normalised, or NULL */
    const float* cov3D_precomp;  /* [P,6] (xx,xy,xz,yy,yz,zz) or NULL */
    const float* viewmatrix;     /* [16] world->view, row-vector convention: m[4*c+r] */
    const float* projmatrix;     /* [16] full projection, same convention */
    const float* cam_pos;        /* [3] */
    float        tan_fovx, tan_fovy;
    int          prefiltered;    /* reference traps when a prefiltered point is culled; we report GSR_EINVAL in debug */
    int          debug;          /* 1: synchronise + check after every stage (CHECK_CUDA semantics, auxiliary.h:166-173) */
    float*       out_color;      /* [3,H,W] written for every pixel */
    int*         radii;          /* [P] written for every Gaussian (0 = culled) */
    /* Split SH storage (ABI 2): when non-NULL, `shs` holds only the DC coefficient [P,1,3] and `shs_rest` the others
       [P,M-1,3] - the reference model's two parameters _features_dc / _features_rest (scene/gaussian_model.py:151-152)
       without the torch.cat of get_features (:119-122).  M stays the total count; M must be 16 (or 1: shs_rest ignored). */
    const float* shs_rest;
} GsrFwdArgs;

/*
 * Forward.  Blocks the calling host thread once (like rasterizer_impl.cu:280-281) to learn
 * num_rendered, because the binning blob is sized from it (the wait is placed after the
 * depth sort has been queued; other host threads are not blocked).
 * *num_rendered (host) receives R = sum of tiles touched.
 */
int gsr_forward(const GsrFwdArgs* args, GsrAlloc geom, GsrAlloc binning, GsrAlloc img,
                gsr_stream_t stream, int* num_rendered);

/*
 * gsr_forward on two streams (ABI 5).  Everything up to and including the tile split (preprocess, depth order, instance
 * emission, tile split: rasterizer_impl.cu:216-318) is issued on `front_stream`, the blending (FORWARD::render,
 * rasterizer_impl.cu:320-333) on `stream`, which waits for the front through an event of the library's own.  Results are
 * those of gsr_forward.  What it is for: the front of a view is bound by memory latency and bandwidth, the blending kernels by
 * instruction issue; a trainer that renders several views per step calls the forward of view i+1 BEFORE the backward of
 * view i and gives `front_stream` a higher priority than `stream`, so the front of the next view runs inside the blending
 * kernels of the current one (gsr_b200/dp.py `deferred_views(pipeline=True)`; the reference runs one view at a time on the
 * legacy stream).  The caller's side of the contract: (1) every input (parameters, camera) must be valid ON `front_stream` -
 * the library does not make `front_stream` wait for `stream`; (2) radii and the three blobs are written on `front_stream`
 * and read on `stream`: an allocator that recycles memory by stream must be told about both (the torch binding allocates
 * them under the front stream and calls record_stream for `stream`).  The host is blocked once, for the front's
 * num_rendered only.  front_stream == stream (or NULL) is gsr_forward.
 */
int gsr_forward_front(const GsrFwdArgs* args, GsrAlloc geom, GsrAlloc binning, GsrAlloc img,
                      gsr_stream_t front_stream, gsr_stream_t stream, int* num_rendered);

/*
 * Preprocess of several views at once (ABI 5).  A step that renders n_views <= 16 views of ONE Gaussian set through
 * single-view calls reads the Gaussian parameters once per view - at SH degree 3 that is 236 of the ~300 bytes per Gaussian
 * a view's preprocess (preprocessCUDA, forward.cu:155-259) moves.  gsr_preprocess_views runs that stage for all the views in
 * one pass over the parameters: `common` carries what the views share (P, D, M, W, H, tan_fov*, scale_modifier, prefiltered,
 * means3D, shs / shs_rest, opacities, scales, rotations; colors_precomp / cov3D_precomp must be NULL, out_color / radii /
 * camera fields are ignored), every view brings its camera, its radii output and a geometry blob of gsr_geom_bytes(P) bytes.
 * Each blob is then handed to gsr_forward_prepared, which runs the rest of gsr_forward (depth order, tile split, blending) on
 * it; results are those of gsr_forward with that camera.  A prepared blob is consumed by exactly one gsr_forward_prepared
 * call.  (The reference preprocesses one view per call: rasterizer_impl.cu:232-260.)
 */
typedef struct GsrViewFront {
    const float* viewmatrix;     /* [16] */
    const float* projmatrix;     /* [16] */
    const float* cam_pos;        /* [3]  */
    int*         radii;          /* [P], written */
    void*        geom_blob;      /* gsr_geom_bytes(P) bytes, written */
} GsrViewFront;
int gsr_preprocess_views(const GsrFwdArgs* common, int n_views, const GsrViewFront* views, gsr_stream_t stream);

/*
 * The rest of gsr_forward on a blob prepared by gsr_preprocess_views.  `args` as for gsr_forward with the view's camera
 * (args->radii is not written: gsr_preprocess_views did); front_stream as in gsr_forward_front (NULL or == stream: one stream).
 * The work of gsr_preprocess_views must be complete on (or ordered before) the stream the front runs on.
 */
int gsr_forward_prepared(const GsrFwdArgs* args, void* prepared_geom_blob, GsrAlloc binning, GsrAlloc img,
                         gsr_stream_t front_stream, gsr_stream_t stream, int* num_rendered);

/*
 * Batched forward: n_views cameras over ONE Gaussian set in a single pass (the reference
 * renders one view per call, train_lgm.py:111-120).  Same argument struct with these per-view extensions:
 *   viewmatrix [n_views,16], projmatrix [n_views,16], cam_pos [n_views,3]  (W, H, tan_fov*, background shared)
 *   out_color  [n_views,3,H,W], radii [n_views,P]
 * Every view's result is identical to a gsr_forward call with that camera; depth sort, tile split and blending run once
 * over all (Gaussian, view) pairs, and the host is blocked once.  colors_precomp / cov3D_precomp must be NULL when
 * n_views > 1.  *num_rendered is the total over the views.  The blobs belong to gsr_backward_batched with the same n_views;
 * the image blob is obtained with gsr_img_bytes_batched bytes.
 */
int gsr_forward_batched(const GsrFwdArgs* args, int n_views, GsrAlloc geom, GsrAlloc binning, GsrAlloc img,
                        gsr_stream_t stream, int* num_rendered);

typedef struct GsrBwdArgs {
    int P, D, M, R, W, H;
    const float* background;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* scales;
    float        scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float        tan_fovx, tan_fovy;
    const int*   radii;          /* [P] as returned by forward */
    const void*  geom_blob;      /* the three blobs obtained through the forward allocators */
    const void*  binning_blob;
    const void*  img_blob;
    const float* dL_dpix;        /* [3,H,W] */
    /* scratch: gsr_bwd_accum_bytes(P) bytes, 16-byte aligned.  Zeroed and filled by gsr_backward itself
       (packed per-Gaussian accumulator rows of the blending backward; replaces the reference's separate
       dL_dconic [P,2,2] / dL_dcolors [P,3] atomics targets, rasterize_points.cu:151-159). */
    void*  accum;
    /* outputs.  EVERY element of every non-NULL output is written (zeros for culled Gaussians): the
       caller allocates them uninitialised, no pre-zeroing (the reference needs nine torch::zeros). */
    float* dL_dmean2D;           /* [P,3]  (x,y in NDC units, z == 0); required */
    float* dL_dopacity;          /* [P];   required */
    float* dL_dmean3D;           /* [P,3]; required */
    float* dL_dcolor;            /* [P,3]  or NULL (wanted when colors_precomp is the input) */
    float* dL_dcov3D;            /* [P,6]  or NULL (wanted when cov3D_precomp is the input) */
    float* dL_dsh;               /* [P,M,3]; required iff shs != NULL */
    float* dL_dscale;            /* [P,3];   required iff scales != NULL */
    float* dL_drot;              /* [P,4];   required iff rotations != NULL */
    int    debug;
    /* GSR_ACC_* bits.  A flagged output is ACCUMULATED: rows of visible Gaussians are added to what the buffer
       holds, rows of culled Gaussians are left untouched (multi-view steps sum parameter gradients in place
       instead of through one extra elementwise add per output and view).  0 = overwrite everything (reference). */
    unsigned accumulate;
    /* split SH storage as in GsrFwdArgs (ABI 2): dL_dsh is then [P,1,3] and dL_dsh_rest [P,M-1,3] */
    const float* shs_rest;
    float*       dL_dsh_rest;
    /* ABI 5: non-zero = `accum` already holds zeros (the caller cleared it off the critical path, e.g. on the front stream
       of gsr_forward_front while the forward's blending ran); 0 = the blending backward clears it itself (144 MB at C3). */
    int          accum_zeroed;
} GsrBwdArgs;

#define GSR_ACC_MEAN3D   1u
#define GSR_ACC_OPACITY  2u
#define GSR_ACC_SH       4u
#define GSR_ACC_SCALE    8u
#define GSR_ACC_ROT     16u

int gsr_backward(const GsrBwdArgs* args, gsr_stream_t stream);

/*
 * gsr_backward in separately callable pieces (ABI 3).  `stages` bit 0: the blending backward (fills args->accum for all
 * Gaussians); bit 1: the per-Gaussian backward (cov2D / cov3D / SH / projection, rasterizer_impl.cu:404-433) restricted to
 * the Gaussians [g_begin, g_end) - g_begin a multiple of 128, g_end = -1 for P.  Every output row inside the range is written
 * exactly as gsr_backward writes it, rows outside are not touched; gsr_backward(args) ==
 * gsr_backward_staged(args, 3, 0, -1).  A view-sharded data-parallel trainer calls bit 0 once and bit 1 range by range for
 * the last view of its step, starting the all-reduce of a range's gradients while the next range is computed
 * (gsr_b200/dp.py; the reference has no multi-GPU path).
 */
int gsr_backward_staged(const GsrBwdArgs* args, int stages, int g_begin, int g_end, gsr_stream_t stream);

/*
 * Deferred per-Gaussian backward (ABI 4): a step that renders several views of ONE Gaussian set through single-view calls
 * runs, per view, only the blending backward (gsr_backward_staged(args_v, 1, 0, 0, stream) into that view's accum scratch) and,
 * once per step, the per-Gaussian backward of all n_views <= 16 views together: the parameters are read once and every
 * gradient row is written once (summed over the views) instead of once per view.  `common` carries what the views share
 * (P, D, M, W, H, tan_fov*, scale_modifier, means3D, shs / shs_rest, scales, rotations, the summed outputs dL_dopacity,
 * dL_dmean3D, dL_dsh / dL_dsh_rest, dL_dscale, dL_drot and `accumulate`); colors_precomp / cov3D_precomp must be NULL.
 * [g_begin, g_end) restricts the call to a range of Gaussians as in gsr_backward_staged (g_end = -1: P).
 * The result equals n_views gsr_backward calls of which the first overwrites and the others accumulate.
 */
typedef struct GsrViewState {
    const float* viewmatrix;     /* [16] */
    const float* projmatrix;     /* [16] */
    const float* cam_pos;        /* [3]  */
    const int*   radii;          /* [P] of this view's forward */
    const void*  geom_blob;      /* of this view's forward */
    const void*  accum;          /* gsr_bwd_accum_bytes(P) bytes filled by this view's blending backward */
    float*       dL_dmean2D;     /* [P,3], written for this view */
} GsrViewState;
int gsr_backward_views(const GsrBwdArgs* common, int n_views, const GsrViewState* views, int g_begin, int g_end,
                       gsr_stream_t stream);

/*
 * Backward of gsr_forward_batched: radii [n_views,P], dL_dpix [n_views,3,H,W], camera arrays as in the forward,
 * accum = gsr_bwd_accum_bytes(n_views * P) bytes; dL_dmean2D is per view [n_views,P,3]; every other gradient is the
 * SUM over the views ([P,...]: view 0 overwrites - or accumulates when `accumulate` says so - and the others add).
 */
int gsr_backward_batched(const GsrBwdArgs* args, int n_views, gsr_stream_t stream);

/*
 * Several scenes per call (many small scenes, a few views each; the reference renders
 * one view of one scene per call, train_lgm.py:111-120).  n_scenes scenes of args->P Gaussians EACH, views_per_scene cameras
 * per scene.  Same structs with
 *   means3D [S,P,3], opacities [S,P], scales [S,P,3], rotations [S,P,4], shs [S,P,M,3] (or split: shs [S,P,1,3], shs_rest [S,P,M-1,3])
 *   viewmatrix / projmatrix [S,Vs,16], cam_pos [S,Vs,3];  out_color [S,Vs,3,H,W], radii [S,Vs,P]
 * and for the backward dL_dpix [S,Vs,3,H,W], dL_dmean2D [S,Vs,P,3], parameter gradients [S,P,...] (summed over the scene's views),
 * accum = gsr_bwd_accum_bytes(S*Vs*P).  Preprocess and the per-Gaussian backward run once per scene, depth sort, tile split
 * and both blending kernels ONCE over all S*Vs*P (Gaussian, view) pairs; the host is blocked once.  Every view's result is
 * identical to a single-view call on its scene.  gsr_forward_batched(args, V) == gsr_forward_scenes(args, 1, V).
 * Blob sizes: gsr_geom_bytes(S*Vs*P), gsr_img_bytes_batched(W, H, S*Vs).  Split SH storage needs P % 4 == 0 when S > 1.
 */
int gsr_forward_scenes(const GsrFwdArgs* args, int n_scenes, int views_per_scene, GsrAlloc geom, GsrAlloc binning,
                       GsrAlloc img, gsr_stream_t stream, int* num_rendered);
int gsr_backward_scenes(const GsrBwdArgs* args, int n_scenes, int views_per_scene, gsr_stream_t stream);

/* present[i] = (view-space z of means3D[i] > 0.2).  `present` is uint8 (bool). */
int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, gsr_stream_t stream);

/*
 * distCUDA2: out[i] = mean of the 3 smallest squared distances from points[i] to
 * the other points (by index; coincident points count as distance 0).
 * Scratch comes from `scratch` (one call).  Synchronises the stream once
 * internally (bounding box -> host), like simple_knn.cu:197,200.
 */
int gsr_dist2(int P, const float* points, float* out, GsrAlloc scratch, gsr_stream_t stream);

/*
 * distCUDA2 for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): the same kernels and the same result as gsr_dist2, but the scratch lives in a file-scope __device__ array
 * of the library (about 10 MB, held by every GPU that loads the library until exit) instead of coming from an
 * allocator callback.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
 * This entry point has no counterpart in the reference.
 *   P <= GSR_DIST2_STATIC_MAX_POINTS, else GSR_ETOOBIG (gsr_last_cuda_error() holds the text); use gsr_dist2 beyond it.
 * All calls on one device share the array, which keeps the last call's contents (the call clears what must start at
 * zero): calls that overlap on the GPU - two streams, two host threads on one device - race; order them on one stream.
 * Additive: GSR_ABI_VERSION is unchanged.
 */
#define GSR_DIST2_STATIC_MAX_POINTS 262144
int gsr_dist2_static(int P, const float* points, float* out, gsr_stream_t stream);

/*
 * Fused L1 image loss (forward + gradient in one pass): replaces utils/loss_utils.py:17-18
 * `torch.abs(network_output - gt).mean()` and its autograd backward.
 *   *loss_sum = sum_i |image[i] - target[i]|         (device scalar, zeroed by the call)
 *   grad[i]   = grad_scale * sign(image[i] - target[i])   (0 where equal; grad may be NULL)
 * image, target and grad must be 16-byte aligned device pointers of n floats.
 */
int gsr_l1_loss(const float* image, const float* target, size_t n, float grad_scale, float* loss_sum,
                float* grad, gsr_stream_t stream);

/*
 * Fused L1 + D-SSIM image loss (value and gradient from one call): replaces train.py:104-105
 *   loss = (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))
 * with ssim() of utils/loss_utils.py:27-63 (11x11 Gaussian window, sigma 1.5, zero padding 5, depthwise, mean of the map)
 * and the autograd backward of both.  image, target, grad: [C, H, W] float device arrays.
 *   sums[0] = sum |image - target|,  sums[1] = sum ssim_map        (device, zeroed by the call)
 *   grad    = k_l1 * sign(image - target) + k_ssim * d(sum ssim_map)/d image     (grad may be NULL: value only)
 *   maps    = scratch of 3*C*H*W floats (the three partial-derivative maps), required when grad != NULL
 * For the loss above: n = C*H*W, k_l1 = (1 - lambda)/n, k_ssim = -lambda/n, loss = (1-lambda) sums[0]/n + lambda (1 - sums[1]/n).
 */
int gsr_l1_ssim_loss(const float* image, const float* target, int C, int H, int W, float k_l1, float k_ssim,
                     float* sums, float* maps, float* grad, gsr_stream_t stream);

/*
 * Parameter activations of the model, one launch per direction (scene/gaussian_model.py:33-40,100-125):
 *   scales = exp(scaling_raw) [P,3];  rotations = rotation_raw / max(|rotation_raw|, 1e-12) [P,4];
 *   opacities = 1 / (1 + exp(-opacity_raw)) [P]
 * Any output (with its input) may be NULL.  The backward recomputes the activations from the raw parameters and writes
 * (accumulate = 0) or adds (accumulate = 1) the raw parameters' gradients; a NULL gradient output skips that parameter.
 * The [P,4] arrays must be 16-byte aligned.
 */
int gsr_activate_forward(int P, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                         float* scales, float* rotations, float* opacities, gsr_stream_t stream);
int gsr_activate_backward(int P, const float* scaling_raw, const float* rotation_raw, const float* opacity_raw,
                          const float* dL_dscales, const float* dL_drotations, const float* dL_dopacities,
                          float* dL_dscaling_raw, float* dL_drotation_raw, float* dL_dopacity_raw, int accumulate,
                          gsr_stream_t stream);

/*
 * Fused Adam step over all parameter groups in one launch: replaces gaussians.optimizer.step() of train.py:138-140,
 * i.e. torch.optim.Adam(l, lr=0.0, eps=1e-15) over the groups built at scene/gaussian_model.py:164-175
 * (no weight decay, no amsgrad).  Same fp32 arithmetic as torch's foreach implementation:
 *   m += (1-beta1)(g-m);  v = v*beta2 + (1-beta2) g*g;  p -= lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps)
 * `step` is the 1-based step count of that group AFTER this update (torch increments before use).
 * `row` = elements per Gaussian in this group (3, 3, 45, 1, 3, 4 ...), used only with `visible`:
 * visible == NULL is the reference's dense update; otherwise visible[P] (bytes, non-zero = update) restricts the
 * update to flagged Gaussians (moments and parameters of the others are left untouched).
 */
typedef struct GsrAdamGroup {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    size_t n;                    /* number of floats */
    int row;                     /* floats per Gaussian */
    int step;                    /* >= 1 */
    float lr;
} GsrAdamGroup;
int gsr_adam_step(const GsrAdamGroup* groups, int n_groups, double beta1, double beta2, double eps,
                  const unsigned char* visible, gsr_stream_t stream);

/*
 * Densification statistics of one view (train.py:131-132, scene/gaussian_model.py:415-417) for the filter radii > 0:
 *   max_radii2D[i] = max(max_radii2D[i], radii[i]);  xyz_gradient_accum[i] += |dL_dmean2D[i, :2]|;  denom[i] += 1
 * dL_dmean2D is the [P,3] gradient of the rasterizer's means2D input (viewspace_point_tensor.grad).
 */
int gsr_densify_stats(int P, const int* radii, const float* dL_dmean2D, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, gsr_stream_t stream);

/*
 * Column gather between the PLY record layout and the parameter tensors (scene/gaussian_model.py:201-266 save_ply / load_ply):
 *   out[r * out_cols + j] = col_of_out[j] >= 0 ? in[r * in_cols + col_of_out[j]] : 0      for r < rows, j < out_cols <= 64
 * in / out are device arrays; col_of_out is a HOST array of out_cols entries.  Loading: in = the file's vertex payload
 * (62 floats per Gaussian at SH-3), one call per parameter tensor; saving: in = the parameters concatenated per Gaussian,
 * out = the 62-column records (negative entries produce the zero normals); see gsr_b200/ply.py.
 */
int gsr_gather_columns(const float* in, int in_cols, float* out, int out_cols, const int* col_of_out, long long rows,
                       gsr_stream_t stream);

/*
 * Test hook for the bit-exact integer contract: reconstructs, from our private
 * blobs, the arrays the reference keeps in BinningState / ImageState:
 *   keys[R]   = (tile << 32) | float_bits(depth)   sorted
 *   values[R] = Gaussian index                     sorted
 *   ranges[2*tiles] = (start, end) per tile, (0,0) for empty tiles
 *   n_contrib[W*H], final_T[W*H]
 * Any output pointer may be NULL.
 */
int gsr_export_binning(const void* geom_blob, const void* binning_blob, const void* img_blob,
                       int P, int R, int W, int H,
                       uint64_t* keys, uint32_t* values, uint32_t* ranges,
                       uint32_t* n_contrib, float* final_T, gsr_stream_t stream);

/* bytes the three blobs need (geom, img are known up-front; binning needs R) */
size_t gsr_geom_bytes(int P);
size_t gsr_img_bytes(int W, int H);
size_t gsr_img_bytes_batched(int W, int H, int n_views);
size_t gsr_binning_bytes(int P, int R, int W, int H);
size_t gsr_bwd_accum_bytes(int P);

const char* gsr_strerror(int code);
const char* gsr_last_cuda_error(void);
int gsr_abi_version(void);

/*
 * Per-stage device timing (bench.py's roofline leg).  When enabled, every stage of gsr_forward /
 * gsr_backward is bracketed by cudaEvents recorded on the caller's stream; gsr_timing_collect()
 * synchronises those events and adds the elapsed times to per-stage totals.
 * Stage ids: 0 preprocess, 1 depth_order, 2 tile_split, 3 blend_forward, 4 blend_backward,
 * 5 preprocess_backward, 6 dist2.   ms / calls may be NULL.
 */
#define GSR_NUM_STAGES 7
void gsr_timing_enable(int on);
int  gsr_timing_collect(void);
int  gsr_timing_read(int stage, double* ms_total, unsigned long long* calls);
void gsr_timing_reset(void);
const char* gsr_stage_name(int stage);

/* kernel-launch counter (all launches made by this library since the last reset) */
void gsr_reset_launch_count(void);
unsigned long long gsr_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H_INCLUDED */
