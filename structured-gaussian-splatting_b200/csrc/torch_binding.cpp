// torch_binding.cpp - thin PyTorch extension over the C ABI of include/gsr.h (host-only C++, no device code).
//
// Mirrors the reference bindings' module-level functions (same names, argument order, returns):
//   rasterize_gaussians            <- RasterizeGaussiansCUDA          diff-gaussian-rasterization/rasterize_points.cu:35-115
//   rasterize_gaussians_backward   <- RasterizeGaussiansBackwardCUDA  rasterize_points.cu:117-196
//   mark_visible                   <- markVisible                     rasterize_points.cu:198-217
//   distCUDA2                      <- distCUDA2                       simple-knn/spatial.cu:15-26
// Differences a caller can observe: work is issued on torch's CURRENT stream (the reference uses the
// legacy default stream); gradient tensors nobody asked for (dL_dcolors without colors_precomp,
// dL_dcov3D without cov3D_precomp) come back empty instead of as P-sized temporaries; errors from the
// core are raised as RuntimeError carrying gsr_strerror().
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <memory>
#include <functional>
#include <tuple>
#include <string>
#include <vector>
#include "../../include/gsr.h"

// debug hooks of the core library (csrc/context.cu; not part of the public C ABI): tuning options and the comparison arms
extern "C" int gsr_debug_set_option(const char* name, int value);
extern "C" int gsr_debug_get_option(const char* name, int* value);
extern "C" int gsr_debug_arms_loaded(void);

namespace {

// The binning blob's size follows the view's instance count, which differs from camera to camera: handed to torch's caching
// allocator byte-exact, a step keeps meeting sizes no cached block fits and falls through to cudaMalloc in the middle of the
// pipelined step (3-4 times in 20 steps at C3, the odd run 5 % slower).  Large blobs are rounded up to eight size classes
// per power of two (at most 12.5 % more), so that after the first steps every request finds a cached block.
void* resize_cb(void* ctx, size_t bytes)
{
    auto* t = static_cast<torch::Tensor*>(ctx);
    size_t cap = bytes;
    if (bytes > (size_t(1) << 20)) {
        int top = 0;
        while ((bytes >> (top + 1)) != 0) top++;
        const size_t g = size_t(1) << (top - 3);
        cap = (bytes + g - 1) / g * g;
    }
    t->resize_({(long long)cap});
    return t->data_ptr();
}

void check(int rc, const char* what)
{
    if (rc == GSR_OK) return;
    std::string msg = std::string(what) + ": " + gsr_strerror(rc);
    if (rc == GSR_ECUDA) msg += std::string(" (") + gsr_last_cuda_error() + ")";
    throw std::runtime_error(msg);
}

// contiguous fp32 view of an optional input; empty tensor => nullptr (the reference's "absent" sentinel)
struct F32 {
    torch::Tensor t;
    explicit F32(const torch::Tensor& x) { if (x.numel() > 0) t = x.contiguous(); }
    const float* ptr() const { return t.defined() ? t.data_ptr<float>() : nullptr; }
};

gsr_stream_t current_stream() { return reinterpret_cast<gsr_stream_t>(at::cuda::getCurrentCUDAStream().stream()); }

}  // namespace

// `front_stream` != 0 (a cudaStream_t of means3D's device, as an integer): gsr_forward_front - preprocess and binning are issued
// on that stream, the blending on torch's current stream.  radii and the state blobs, which the front writes and the current
// stream reads, are allocated under the front stream and recorded for the current one (the caching allocator recycles memory
// per stream).  The caller has made the inputs valid on the front stream (diff_gaussian_rasterization.set_front_stream).
static std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
forward_core(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                          const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                          const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                          const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                          const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                          const bool prefiltered, const bool debug, const torch::Tensor& sh_rest, int64_t front_stream,
                          const torch::Tensor* prepared_geom = nullptr, const torch::Tensor* prepared_radii = nullptr)
{
    if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const int H = image_height, W = image_width;
    auto f32 = means3D.options().dtype(torch::kFloat32);
    auto i32 = means3D.options().dtype(torch::kInt32);
    auto u8 = means3D.options().dtype(torch::kByte);

    // inputs the front would read through a .contiguous() copy made on the current stream: plain single-stream call
    for (const torch::Tensor* t : {&background, &means3D, &colors, &opacity, &scales, &rotations, &cov3D_precomp, &viewmatrix, &projmatrix,
                                   &sh, &campos, &sh_rest})
        if (t->numel() > 0 && !t->is_contiguous()) front_stream = 0;
    const c10::cuda::CUDAStream main_stream = at::cuda::getCurrentCUDAStream();
    if (P == 0 || (cudaStream_t)front_stream == main_stream.stream()) front_stream = 0;
    torch::Tensor out_color = P == 0 ? torch::zeros({3, H, W}, f32) : torch::empty({3, H, W}, f32);
    std::unique_ptr<c10::cuda::CUDAStreamGuard> front_guard;
    if (front_stream)
        front_guard.reset(new c10::cuda::CUDAStreamGuard(
            c10::cuda::getStreamFromExternal(reinterpret_cast<cudaStream_t>(front_stream), means3D.device().index())));
    const bool prepared = prepared_geom != nullptr && P != 0;
    torch::Tensor radii = prepared ? *prepared_radii : (P == 0 ? torch::zeros({P}, i32) : torch::empty({P}, i32));
    torch::Tensor geomBuffer = prepared ? *prepared_geom : torch::empty({0}, u8);
    torch::Tensor binningBuffer = torch::empty({0}, u8), imgBuffer = torch::empty({0}, u8);
    int rendered = 0;
    if (P != 0) {
        int M = 0;
        if (sh.numel() != 0) M = (int)sh.size(1);
        if (sh_rest.numel() != 0) {
            TORCH_CHECK(sh.numel() != 0 && sh.size(1) == 1 && sh_rest.dim() == 3 && sh_rest.size(0) == P, "split SH: sh [P,1,3] and sh_rest [P,K,3] expected");
            M = 1 + (int)sh_rest.size(1);
        }
        F32 bg(background), m3(means3D), col(colors), op(opacity), sc(scales), rot(rotations), cov(cov3D_precomp), vm(viewmatrix),
            pm(projmatrix), shs(sh), cam(campos), shr(sh_rest);
        GsrFwdArgs a = {};
        a.P = P; a.D = degree; a.M = M; a.W = W; a.H = H;
        a.shs_rest = shr.ptr();
        a.background = bg.ptr(); a.means3D = m3.ptr(); a.shs = shs.ptr(); a.colors_precomp = col.ptr(); a.opacities = op.ptr();
        a.scales = sc.ptr(); a.scale_modifier = scale_modifier; a.rotations = rot.ptr(); a.cov3D_precomp = cov.ptr();
        a.viewmatrix = vm.ptr(); a.projmatrix = pm.ptr(); a.cam_pos = cam.ptr();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.prefiltered = prefiltered; a.debug = debug;
        a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int>();
        if (prepared) {
            TORCH_CHECK(prepared_geom->is_cuda() && prepared_geom->numel() >= (int64_t)gsr_geom_bytes(P) && prepared_radii->numel() == P,
                        "prepared view: geometry blob / radii of another size");
            check(gsr_forward_prepared(&a, geomBuffer.data_ptr(), GsrAlloc{resize_cb, &binningBuffer}, GsrAlloc{resize_cb, &imgBuffer},
                                       front_stream ? current_stream() : nullptr, reinterpret_cast<gsr_stream_t>(main_stream.stream()),
                                       &rendered),
                  "rasterize_gaussians");
            front_guard.reset();
            // (the blob and radii may have been allocated under another stream than the one that reads them from here on)
            for (torch::Tensor* t : {&radii, &geomBuffer, &binningBuffer, &imgBuffer})
                if (t->numel() > 0) t->record_stream(main_stream);
        } else if (front_stream) {
            // (torch's current stream is the front stream here: the blobs the callbacks resize belong to it)
            check(gsr_forward_front(&a, GsrAlloc{resize_cb, &geomBuffer}, GsrAlloc{resize_cb, &binningBuffer}, GsrAlloc{resize_cb, &imgBuffer},
                                    current_stream(), reinterpret_cast<gsr_stream_t>(main_stream.stream()), &rendered),
                  "rasterize_gaussians");
            front_guard.reset();
            for (torch::Tensor* t : {&radii, &geomBuffer, &binningBuffer, &imgBuffer})
                if (t->numel() > 0) t->record_stream(main_stream);
        } else {
            check(gsr_forward(&a, GsrAlloc{resize_cb, &geomBuffer}, GsrAlloc{resize_cb, &binningBuffer}, GsrAlloc{resize_cb, &imgBuffer},
                              current_stream(), &rendered),
                  "rasterize_gaussians");
        }
    }
    return std::make_tuple(rendered, out_color, radii, geomBuffer, binningBuffer, imgBuffer);
}

// `sh_rest` non-empty: split SH storage - `sh` is the model's _features_dc [P,1,3] and `sh_rest` its _features_rest [P,M-1,3]
// (scene/gaussian_model.py:151-152), read in place instead of through the torch.cat of get_features (:119-122)
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_split(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                          const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                          const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                          const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                          const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                          const bool prefiltered, const bool debug, const torch::Tensor& sh_rest)
{
    return forward_core(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                        tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, sh_rest, 0);
}

// rasterize_gaussians with the front of the view on a stream of its own (gsr_forward_front)
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_front(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                          const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                          const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                          const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                          const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                          const bool prefiltered, const bool debug, int64_t front_stream)
{
    return forward_core(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                        tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, torch::empty({0}),
                        front_stream);
}

// the reference's entry point (rasterize_points.h:18-38)
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                    const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                    const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                    const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                    const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                    const bool prefiltered, const bool debug)
{
    return rasterize_gaussians_split(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                                     projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug,
                                     torch::empty({0}));
}

// gsr_preprocess_views: the preprocess stage of n views (lists of camera tensors) in one pass over the parameters; returns every
// view's geometry blob and radii, to be handed to rasterize_gaussians_prepared.  Issued on torch's current stream.
std::tuple<std::vector<torch::Tensor>, std::vector<torch::Tensor>>
prepare_views(const torch::Tensor& means3D, const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
              const float scale_modifier, const torch::Tensor& sh, const int degree, const float tan_fovx, const float tan_fovy,
              const int image_height, const int image_width, const std::vector<torch::Tensor>& viewmatrices,
              const std::vector<torch::Tensor>& projmatrices, const std::vector<torch::Tensor>& camposs, const bool prefiltered)
{
    TORCH_CHECK(means3D.is_cuda() && means3D.ndimension() == 2 && means3D.size(1) == 3, "means3D must be a CUDA tensor (num_points, 3)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const size_t V = viewmatrices.size();
    TORCH_CHECK(V >= 1 && V <= 16 && projmatrices.size() == V && camposs.size() == V, "prepare_views: 1..16 cameras");
    TORCH_CHECK(sh.dim() == 3 && sh.size(0) == P, "prepare_views: sh [P,M,3] expected");
    for (const torch::Tensor* t : {&means3D, &opacity, &scales, &rotations, &sh})
        TORCH_CHECK(t->is_cuda() && t->is_contiguous() && t->scalar_type() == torch::kFloat32, "prepare_views: contiguous fp32 CUDA parameters");
    auto i32 = means3D.options().dtype(torch::kInt32);
    auto u8 = means3D.options().dtype(torch::kByte);
    std::vector<torch::Tensor> geoms, radii;
    std::vector<GsrViewFront> vf(V);
    std::vector<F32> keep;
    keep.reserve(3 * V);
    for (size_t v = 0; v < V; v++) {
        geoms.push_back(torch::empty({(int64_t)gsr_geom_bytes(P)}, u8));
        radii.push_back(P == 0 ? torch::zeros({P}, i32) : torch::empty({P}, i32));
        for (const torch::Tensor* t : {&viewmatrices[v], &projmatrices[v], &camposs[v]})
            TORCH_CHECK(t->is_cuda() && t->is_contiguous() && t->scalar_type() == torch::kFloat32, "prepare_views: contiguous fp32 CUDA cameras");
        vf[v].viewmatrix = viewmatrices[v].data_ptr<float>(); vf[v].projmatrix = projmatrices[v].data_ptr<float>();
        vf[v].cam_pos = camposs[v].data_ptr<float>();
        vf[v].radii = radii[v].data_ptr<int>(); vf[v].geom_blob = geoms[v].data_ptr();
    }
    if (P != 0) {
        GsrFwdArgs a = {};
        a.P = P; a.D = degree; a.M = (int)sh.size(1); a.W = image_width; a.H = image_height;
        a.means3D = means3D.data_ptr<float>(); a.shs = sh.data_ptr<float>(); a.opacities = opacity.data_ptr<float>();
        a.scales = scales.data_ptr<float>(); a.scale_modifier = scale_modifier; a.rotations = rotations.data_ptr<float>();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.prefiltered = prefiltered;
        check(gsr_preprocess_views(&a, (int)V, vf.data(), current_stream()), "prepare_views");
    }
    return std::make_tuple(geoms, radii);
}

// rasterize_gaussians on a view prepare_views has preprocessed (gsr_forward_prepared); front_stream as in rasterize_gaussians_front
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_prepared(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                             const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                             const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                             const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                             const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                             const bool prefiltered, const bool debug, const torch::Tensor& geom, const torch::Tensor& radii,
                             int64_t front_stream)
{
    TORCH_CHECK(colors.numel() == 0 && cov3D_precomp.numel() == 0, "prepared views take SH colours and scale / rotation covariances");
    return forward_core(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                        tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, torch::empty({0}),
                        front_stream, &geom, &radii);
}

// returns the reference's eight gradients plus dL_dsh_rest (empty unless sh_rest is given)
using BwdResult = std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
                             torch::Tensor, torch::Tensor>;

// n_chunks > 0: the per-Gaussian backward runs in n_chunks ranges of Gaussians (gsr_backward_staged) and `after_chunk(g0, g1)`
// is called on the host after each range has been queued (the data-parallel step starts that range's all-reduce there)
static BwdResult
backward_core(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                  const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                                  const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                                  const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                                  const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                                  const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                                  const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                                  const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                  const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate,
                                  const torch::Tensor& sh_rest, const torch::Tensor& into_sh_rest, int n_chunks,
                                  const std::function<void(int, int)>& after_chunk)
{
    // Extension of the reference entry point: a non-empty `into_*` tensor (contiguous fp32 CUDA, same element count as
    // the gradient) receives that gradient directly - overwritten, or added to when `accumulate` - and the
    // corresponding slot of the returned tuple is an empty tensor.
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const int H = (int)dL_dout_color.size(1), W = (int)dL_dout_color.size(2);
    int M = 0;
    if (sh.numel() != 0) M = (int)sh.size(1);
    const bool split = sh_rest.numel() != 0;
    const int Mdc = M;                                        // coefficients held by `sh` itself
    if (split) {
        TORCH_CHECK(M == 1 && sh_rest.dim() == 3 && sh_rest.size(0) == P, "split SH: sh [P,1,3] and sh_rest [P,K,3] expected");
        M = 1 + (int)sh_rest.size(1);
    }
    auto f32 = means3D.options().dtype(torch::kFloat32);
    const bool want_colors = colors.numel() != 0, want_cov = cov3D_precomp.numel() != 0;
    const bool has_sr = scales.numel() != 0 && rotations.numel() != 0;

    // the core writes every element of every gradient it is given: no zero-fill (reference: nine torch::zeros)
    auto alloc = [&](std::initializer_list<int64_t> shape) { return P == 0 ? torch::zeros(shape, f32) : torch::empty(shape, f32); };
    unsigned acc_bits = 0;
    auto target = [&](const torch::Tensor& into, int64_t numel, unsigned bit) -> float* {
        if (into.numel() == 0) return nullptr;
        TORCH_CHECK(into.is_cuda() && into.scalar_type() == torch::kFloat32 && into.is_contiguous() && into.numel() == numel,
                    "gradient target must be a contiguous fp32 CUDA tensor with the gradient's element count");
        if (accumulate) acc_bits |= bit;
        return into.data_ptr<float>();
    };
    float* t_m3 = target(into_means3D, (int64_t)P * 3, GSR_ACC_MEAN3D);
    float* t_op = target(into_opacity, (int64_t)P, GSR_ACC_OPACITY);
    float* t_sh = target(into_sh, (int64_t)P * Mdc * 3, GSR_ACC_SH);
    float* t_shr = split ? target(into_sh_rest, (int64_t)P * (M - 1) * 3, GSR_ACC_SH) : nullptr;
    TORCH_CHECK(!split || !accumulate || ((t_sh != nullptr) == (t_shr != nullptr)),
                "split SH: accumulate needs targets for both f_dc and f_rest or for neither");
    float* t_sc = has_sr ? target(into_scales, (int64_t)P * 3, GSR_ACC_SCALE) : nullptr;
    float* t_rt = has_sr ? target(into_rotations, (int64_t)P * 4, GSR_ACC_ROT) : nullptr;
    auto none = [&]() { return torch::empty({0}, f32); };
    torch::Tensor dL_dmeans3D = t_m3 ? none() : alloc({P, 3});
    torch::Tensor dL_dmeans2D = alloc({P, 3});
    torch::Tensor dL_dcolors = want_colors ? alloc({P, 3}) : none();
    torch::Tensor dL_dopacity = t_op ? none() : alloc({P, 1});
    torch::Tensor dL_dcov3D = want_cov ? alloc({P, 6}) : none();
    torch::Tensor dL_dsh = t_sh ? none() : alloc({P, Mdc, 3});
    torch::Tensor dL_dsh_rest = (split && !t_shr) ? alloc({P, M - 1, 3}) : none();
    torch::Tensor dL_dscales = (has_sr && !t_sc) ? alloc({P, 3}) : none();
    torch::Tensor dL_drotations = (has_sr && !t_rt) ? alloc({P, 4}) : none();

    if (P != 0) {
        torch::Tensor accum = torch::empty({(long long)(gsr_bwd_accum_bytes(P) / sizeof(float))}, f32);
        F32 bg(background), m3(means3D), col(colors), sc(scales), rot(rotations), cov(cov3D_precomp), vm(viewmatrix), pm(projmatrix),
            shs(sh), cam(campos), dpix(dL_dout_color), shr(sh_rest);
        torch::Tensor radii_c = radii.contiguous();
        GsrBwdArgs a = {};
        a.shs_rest = shr.ptr();
        a.dL_dsh_rest = split ? (t_shr ? t_shr : dL_dsh_rest.data_ptr<float>()) : nullptr;
        a.P = P; a.D = degree; a.M = M; a.R = R; a.W = W; a.H = H;
        a.background = bg.ptr(); a.means3D = m3.ptr(); a.shs = shs.ptr(); a.colors_precomp = col.ptr();
        a.scales = sc.ptr(); a.scale_modifier = scale_modifier; a.rotations = rot.ptr(); a.cov3D_precomp = cov.ptr();
        a.viewmatrix = vm.ptr(); a.projmatrix = pm.ptr(); a.cam_pos = cam.ptr();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.radii = radii_c.data_ptr<int>();
        a.geom_blob = geomBuffer.data_ptr(); a.binning_blob = binningBuffer.data_ptr(); a.img_blob = imageBuffer.data_ptr();
        a.dL_dpix = dpix.ptr();
        a.accum = accum.data_ptr();
        a.dL_dmean2D = dL_dmeans2D.data_ptr<float>();
        a.dL_dopacity = t_op ? t_op : dL_dopacity.data_ptr<float>();
        a.dL_dmean3D = t_m3 ? t_m3 : dL_dmeans3D.data_ptr<float>();
        a.dL_dcolor = want_colors ? dL_dcolors.data_ptr<float>() : nullptr;
        a.dL_dcov3D = want_cov ? dL_dcov3D.data_ptr<float>() : nullptr;
        a.dL_dsh = M > 0 ? (t_sh ? t_sh : dL_dsh.data_ptr<float>()) : nullptr;
        a.dL_dscale = has_sr ? (t_sc ? t_sc : dL_dscales.data_ptr<float>()) : nullptr;
        a.dL_drot = has_sr ? (t_rt ? t_rt : dL_drotations.data_ptr<float>()) : nullptr;
        a.debug = debug;
        a.accumulate = acc_bits;
        if (n_chunks <= 0) {
            check(gsr_backward(&a, current_stream()), "rasterize_gaussians_backward");
        } else {
            check(gsr_backward_staged(&a, 1, 0, 0, current_stream()), "rasterize_gaussians_backward (blending)");
            const int rows = (((P + n_chunks - 1) / n_chunks) + 127) / 128 * 128;      // ranges start on multiples of 128
            for (int g0 = 0; g0 < P; g0 += rows) {
                const int g1 = std::min(P, g0 + rows);
                check(gsr_backward_staged(&a, 2, g0, g1, current_stream()), "rasterize_gaussians_backward (per-Gaussian range)");
                if (after_chunk) after_chunk(g0, g1);
            }
        }
    }
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, dL_dsh_rest);
}

BwdResult
rasterize_gaussians_backward_split(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                  const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                                  const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                                  const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                                  const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                                  const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                                  const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                                  const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                  const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate,
                                  const torch::Tensor& sh_rest, const torch::Tensor& into_sh_rest)
{
    return backward_core(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                         tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug,
                         into_means3D, into_opacity, into_sh, into_scales, into_rotations, accumulate, sh_rest, into_sh_rest, 0, nullptr);
}

// rasterize_gaussians_backward_into with the per-Gaussian backward in `n_chunks` ranges; `after_chunk(g0, g1)` (a Python
// callable) runs after each range has been queued on the current stream.  Holds the GIL (it calls back into Python).
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_backward_chunked(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                     const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                                     const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                                     const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                                     const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                                     const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                                     const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                                     const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                     const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate,
                                     const int n_chunks, py::object after_chunk)
{
    const torch::Tensor e = torch::empty({0});
    std::function<void(int, int)> cb;
    if (!after_chunk.is_none()) cb = [&after_chunk](int g0, int g1) { after_chunk(g0, g1); };
    auto r = backward_core(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                           tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug,
                           into_means3D, into_opacity, into_sh, into_scales, into_rotations, accumulate, e, e, n_chunks < 1 ? 1 : n_chunks, cb);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r), std::get<3>(r), std::get<4>(r), std::get<5>(r),
                           std::get<6>(r), std::get<7>(r));
}

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_backward_into(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                  const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                                  const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                                  const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                                  const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                                  const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                                  const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                                  const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                  const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate)
{
    const torch::Tensor e = torch::empty({0});
    auto r = rasterize_gaussians_backward_split(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                                viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer,
                                                R, binningBuffer, imageBuffer, debug, into_means3D, into_opacity, into_sh, into_scales,
                                                into_rotations, accumulate, e, e);
    return std::make_tuple(std::get<0>(r), std::get<1>(r), std::get<2>(r), std::get<3>(r), std::get<4>(r), std::get<5>(r),
                           std::get<6>(r), std::get<7>(r));
}

// the reference's entry point (rasterize_points.h:40-62): every gradient freshly allocated
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_backward(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                             const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                             const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                             const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                             const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                             const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                             const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug)
{
    const torch::Tensor e = torch::empty({0});
    return rasterize_gaussians_backward_into(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                             viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer,
                                             R, binningBuffer, imageBuffer, debug, e, e, e, e, e, false);
}

// Deferred backward, step 1 (per view): the blending backward alone.  Returns the view's accumulator rows [P,12]; the
// per-Gaussian backward of all pending views follows in rasterize_gaussians_backward_views.
torch::Tensor rasterize_gaussians_backward_blend(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                                 const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                                                 const float tan_fovy, const torch::Tensor& dL_dout_color, const int M, const int degree,
                                                 const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                                                 const torch::Tensor& imageBuffer, const bool debug, const torch::Tensor& zeroed_accum)
{
    // zeroed_accum non-empty: a [P,12] fp32 buffer of zeros the caller cleared off the critical path (GsrBwdArgs.accum_zeroed)
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const int H = (int)dL_dout_color.size(1), W = (int)dL_dout_color.size(2);
    auto f32 = means3D.options().dtype(torch::kFloat32);
    const bool given = zeroed_accum.numel() != 0;
    if (given)
        TORCH_CHECK(zeroed_accum.is_cuda() && zeroed_accum.is_contiguous() && zeroed_accum.scalar_type() == torch::kFloat32 &&
                    zeroed_accum.numel() == (int64_t)P * 12, "zeroed_accum: contiguous fp32 CUDA [P,12] expected");
    torch::Tensor accum = given ? zeroed_accum : torch::empty({(long long)P, 12}, f32);
    if (P != 0) {
        F32 bg(background), m3(means3D), vm(viewmatrix), pm(projmatrix), dpix(dL_dout_color);
        torch::Tensor radii_c = radii.contiguous();
        GsrBwdArgs a = {};
        a.P = P; a.D = degree; a.M = M; a.R = R; a.W = W; a.H = H;
        a.background = bg.ptr(); a.means3D = m3.ptr(); a.viewmatrix = vm.ptr(); a.projmatrix = pm.ptr();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.radii = radii_c.data_ptr<int>();
        a.geom_blob = geomBuffer.data_ptr(); a.binning_blob = binningBuffer.data_ptr(); a.img_blob = imageBuffer.data_ptr();
        a.dL_dpix = dpix.ptr();
        a.accum = accum.data_ptr();
        a.accum_zeroed = given ? 1 : 0;
        a.debug = debug;
        check(gsr_backward_staged(&a, 1, 0, 0, current_stream()), "rasterize_gaussians_backward_blend");
    }
    return accum;
}

// Deferred backward, step 2 (once per step): the per-Gaussian backward of all pending views (<= 16) of ONE parameter set,
// gradients written (or, with `accumulate`, added) into the given buffers; returns dL_dmeans2D [P,3] per view.
// n_chunks > 0: in ranges of Gaussians with `after_chunk(g0, g1)` called after each range has been queued.
std::vector<torch::Tensor>
rasterize_gaussians_backward_views(const torch::Tensor& means3D, const torch::Tensor& scales, const torch::Tensor& rotations,
                                   const float scale_modifier, const torch::Tensor& sh, const int degree, const float tan_fovx,
                                   const float tan_fovy, const int image_height, const int image_width,
                                   std::vector<torch::Tensor> viewmatrices, std::vector<torch::Tensor> projmatrices,
                                   std::vector<torch::Tensor> camposs, std::vector<torch::Tensor> radii,
                                   std::vector<torch::Tensor> geomBuffers, std::vector<torch::Tensor> accums,
                                   const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                   const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate,
                                   const int n_chunks, py::object after_chunk)
{
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    const size_t V = viewmatrices.size();
    TORCH_CHECK(V >= 1 && V <= 16 && projmatrices.size() == V && camposs.size() == V && radii.size() == V && geomBuffers.size() == V &&
                accums.size() == V, "rasterize_gaussians_backward_views: 1..16 views, lists of equal length");
    TORCH_CHECK(sh.dim() == 3 && sh.size(0) == P && scales.numel() == (int64_t)P * 3 && rotations.numel() == (int64_t)P * 4,
                "rasterize_gaussians_backward_views takes shs [P,M,3], scales and rotations");
    const int M = (int)sh.size(1);
    auto ok = [&](const torch::Tensor& t, int64_t n) {
        return t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == n;
    };
    TORCH_CHECK(ok(into_means3D, (int64_t)P * 3) && ok(into_opacity, P) && ok(into_sh, (int64_t)P * M * 3) && ok(into_scales, (int64_t)P * 3) &&
                ok(into_rotations, (int64_t)P * 4), "gradient targets must be contiguous fp32 CUDA tensors with the gradients' element counts");
    auto f32 = means3D.options().dtype(torch::kFloat32);
    std::vector<torch::Tensor> m2d;
    for (size_t v = 0; v < V; v++) m2d.push_back(P == 0 ? torch::zeros({P, 3}, f32) : torch::empty({P, 3}, f32));
    if (P == 0) return m2d;
    F32 m3(means3D), sc(scales), rot(rotations), shs(sh);
    std::vector<torch::Tensor> keep;
    std::vector<GsrViewState> vs(V);
    for (size_t v = 0; v < V; v++) {
        torch::Tensor vm = viewmatrices[v].contiguous(), pm = projmatrices[v].contiguous(), cp = camposs[v].contiguous(),
                      rd = radii[v].contiguous();
        TORCH_CHECK(accums[v].is_cuda() && accums[v].is_contiguous() && accums[v].numel() == (int64_t)P * 12 && rd.numel() == P,
                    "rasterize_gaussians_backward_views: bad per-view state");
        keep.insert(keep.end(), {vm, pm, cp, rd});
        vs[v].viewmatrix = vm.data_ptr<float>(); vs[v].projmatrix = pm.data_ptr<float>(); vs[v].cam_pos = cp.data_ptr<float>();
        vs[v].radii = rd.data_ptr<int>(); vs[v].geom_blob = geomBuffers[v].data_ptr(); vs[v].accum = accums[v].data_ptr();
        vs[v].dL_dmean2D = m2d[v].data_ptr<float>();
    }
    GsrBwdArgs a = {};
    a.P = P; a.D = degree; a.M = M; a.W = image_width; a.H = image_height;
    a.means3D = m3.ptr(); a.shs = shs.ptr(); a.scales = sc.ptr(); a.scale_modifier = scale_modifier; a.rotations = rot.ptr();
    a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
    a.dL_dopacity = into_opacity.data_ptr<float>(); a.dL_dmean3D = into_means3D.data_ptr<float>(); a.dL_dsh = into_sh.data_ptr<float>();
    a.dL_dscale = into_scales.data_ptr<float>(); a.dL_drot = into_rotations.data_ptr<float>();
    if (accumulate) a.accumulate = GSR_ACC_MEAN3D | GSR_ACC_OPACITY | GSR_ACC_SH | GSR_ACC_SCALE | GSR_ACC_ROT;
    if (n_chunks <= 0) {
        check(gsr_backward_views(&a, (int)V, vs.data(), 0, -1, current_stream()), "rasterize_gaussians_backward_views");
    } else {
        const int rows = (((P + n_chunks - 1) / n_chunks) + 127) / 128 * 128;
        for (int g0 = 0; g0 < P; g0 += rows) {
            const int g1 = std::min(P, g0 + rows);
            check(gsr_backward_views(&a, (int)V, vs.data(), g0, g1, current_stream()), "rasterize_gaussians_backward_views (range)");
            if (!after_chunk.is_none()) after_chunk(g0, g1);
        }
    }
    return m2d;
}

torch::Tensor mark_visible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix)
{
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = (int)means3D.size(0);
    torch::Tensor present = torch::full({P}, false, means3D.options().dtype(at::kBool));
    if (P != 0) {
        F32 m3(means3D), vm(viewmatrix), pm(projmatrix);
        check(gsr_mark_visible(P, m3.ptr(), vm.ptr(), pm.ptr(), reinterpret_cast<uint8_t*>(present.data_ptr<bool>()), current_stream()),
              "mark_visible");
    }
    return present;
}

torch::Tensor distCUDA2(const torch::Tensor& points)
{
    TORCH_CHECK(points.is_cuda(), "points must be a CUDA tensor");
    const c10::cuda::CUDAGuard guard(points.device());
    const int P = (int)points.size(0);
    auto f32 = points.options().dtype(torch::kFloat32);
    torch::Tensor means = torch::full({P}, 0.0, f32);
    if (P != 0) {
        F32 pts(points);
        torch::Tensor scratch = torch::empty({0}, points.options().dtype(torch::kByte));
        check(gsr_dist2(P, pts.ptr(), means.data_ptr<float>(), GsrAlloc{resize_cb, &scratch}, current_stream()), "distCUDA2");
    }
    return means;
}

// test / bench hook: the reference's BinningState / ImageState arrays rebuilt from our blobs
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
export_binning(const torch::Tensor& geomBuffer, const torch::Tensor& binningBuffer, const torch::Tensor& imgBuffer, int P, int R,
               int W, int H)
{
    const c10::cuda::CUDAGuard guard(geomBuffer.device());
    auto o = geomBuffer.options();
    const int tiles = ((W + 15) / 16) * ((H + 15) / 16);
    torch::Tensor keys = torch::empty({R}, o.dtype(torch::kInt64));
    torch::Tensor vals = torch::empty({R}, o.dtype(torch::kInt32));
    torch::Tensor ranges = torch::empty({tiles, 2}, o.dtype(torch::kInt32));
    torch::Tensor ncontrib = torch::empty({H, W}, o.dtype(torch::kInt32));
    torch::Tensor finalT = torch::empty({H, W}, o.dtype(torch::kFloat32));
    check(gsr_export_binning(geomBuffer.data_ptr(), binningBuffer.data_ptr(), imgBuffer.data_ptr(), P, R, W, H,
                             reinterpret_cast<uint64_t*>(keys.data_ptr<int64_t>()), reinterpret_cast<uint32_t*>(vals.data_ptr<int>()),
                             reinterpret_cast<uint32_t*>(ranges.data_ptr<int>()), reinterpret_cast<uint32_t*>(ncontrib.data_ptr<int>()),
                             finalT.data_ptr<float>(), current_stream()),
          "export_binning");
    return std::make_tuple(keys, vals, ranges, ncontrib, finalT);
}

// Batched multi-view entry points (extension): V cameras over one Gaussian set per call.
//   viewmatrices [V,4,4], projmatrices [V,4,4], camposs [V,3]  ->  colors [V,3,H,W], radii [V,P]
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_batched(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& opacity,
                            const torch::Tensor& scales, const torch::Tensor& rotations, const float scale_modifier,
                            const torch::Tensor& viewmatrices, const torch::Tensor& projmatrices, const float tan_fovx,
                            const float tan_fovy, const int image_height, const int image_width, const torch::Tensor& sh,
                            const int degree, const torch::Tensor& camposs, const bool debug, const torch::Tensor& sh_rest)
{
    // one scene: means3D [P,3], cameras [V,4,4] / [V,3].  Several scenes: means3D [S,P,3] (and every parameter with the
    // leading S), cameras [S,Vs,4,4] / [S,Vs,3]; outputs get the leading S as well.
    const bool scenes = means3D.ndimension() == 3;
    if (!scenes && (means3D.ndimension() != 2 || means3D.size(1) != 3)) AT_ERROR("means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    if (scenes) {
        TORCH_CHECK(means3D.size(2) == 3 && viewmatrices.ndimension() == 4 && viewmatrices.size(0) == means3D.size(0) &&
                        projmatrices.sizes() == viewmatrices.sizes() && camposs.ndimension() == 3 && camposs.size(0) == means3D.size(0) &&
                        camposs.size(1) == viewmatrices.size(1),
                    "scenes: means3D [S,P,3], viewmatrices / projmatrices [S,Vs,4,4], camposs [S,Vs,3]");
    } else {
        TORCH_CHECK(viewmatrices.ndimension() == 3 && projmatrices.sizes() == viewmatrices.sizes() && camposs.ndimension() == 2 &&
                        camposs.size(0) == viewmatrices.size(0),
                    "viewmatrices / projmatrices must be [V,4,4] and camposs [V,3]");
    }
    TORCH_CHECK(sh.numel() != 0 && scales.numel() != 0 && rotations.numel() != 0, "batched rasterization takes shs, scales and rotations");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int S = scenes ? (int)means3D.size(0) : 1, Vs = (int)viewmatrices.size(scenes ? 1 : 0);
    const int P = (int)means3D.size(scenes ? 1 : 0), V = S * Vs;
    const int shdim = scenes ? 2 : 1;                     // the coefficient dimension of sh / sh_rest
    const int H = image_height, W = image_width;
    auto f32 = means3D.options().dtype(torch::kFloat32);
    auto i32 = means3D.options().dtype(torch::kInt32);
    auto u8 = means3D.options().dtype(torch::kByte);
    torch::Tensor out_color = P == 0 ? torch::zeros({V, 3, H, W}, f32) : torch::empty({V, 3, H, W}, f32);
    torch::Tensor radii = P == 0 ? torch::zeros({V, P}, i32) : torch::empty({V, P}, i32);
    if (scenes) { out_color = out_color.view({S, Vs, 3, H, W}); radii = radii.view({S, Vs, P}); }
    torch::Tensor geomBuffer = torch::empty({0}, u8), binningBuffer = torch::empty({0}, u8), imgBuffer = torch::empty({0}, u8);
    int rendered = 0;
    if (P != 0 && V != 0) {
        F32 bg(background), m3(means3D), op(opacity), sc(scales), rot(rotations), vm(viewmatrices), pm(projmatrices), shs(sh), cam(camposs),
            shr(sh_rest);
        GsrFwdArgs a = {};
        a.P = P; a.D = degree; a.M = (int)sh.size(shdim); a.W = W; a.H = H;
        if (sh_rest.numel() != 0) {                              // split SH storage: sh = features_dc, sh_rest = features_rest
            TORCH_CHECK(sh.size(shdim) == 1 && sh_rest.dim() == shdim + 2 && sh_rest.size(shdim - 1) == P,
                        "split SH: sh [P,1,3] and sh_rest [P,K,3] expected");
            a.M = 1 + (int)sh_rest.size(shdim);
            a.shs_rest = shr.ptr();
        }
        a.background = bg.ptr(); a.means3D = m3.ptr(); a.shs = shs.ptr(); a.opacities = op.ptr();
        a.scales = sc.ptr(); a.scale_modifier = scale_modifier; a.rotations = rot.ptr();
        a.viewmatrix = vm.ptr(); a.projmatrix = pm.ptr(); a.cam_pos = cam.ptr();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.prefiltered = 0; a.debug = debug;
        a.out_color = out_color.data_ptr<float>(); a.radii = radii.data_ptr<int>();
        check(gsr_forward_scenes(&a, S, Vs, GsrAlloc{resize_cb, &geomBuffer}, GsrAlloc{resize_cb, &binningBuffer},
                                 GsrAlloc{resize_cb, &imgBuffer}, current_stream(), &rendered),
              "rasterize_gaussians_batched");
    }
    return std::make_tuple(rendered, out_color, radii, geomBuffer, binningBuffer, imgBuffer);
}

// returns (dL_dmeans2D [V,P,3], dL_dopacity [P,1], dL_dmeans3D [P,3], dL_dsh [P,M,3], dL_dscales [P,3], dL_drotations [P,4],
// dL_dsh_rest), all but the first summed over the views; with split SH storage dL_dsh is [P,1,3] and dL_dsh_rest [P,M-1,3]
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians_batched_backward(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                     const torch::Tensor& scales, const torch::Tensor& rotations, const float scale_modifier,
                                     const torch::Tensor& viewmatrices, const torch::Tensor& projmatrices, const float tan_fovx,
                                     const float tan_fovy, const torch::Tensor& dL_dout_color, const torch::Tensor& sh, const int degree,
                                     const torch::Tensor& camposs, const torch::Tensor& geomBuffer, const int R,
                                     const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug,
                                     const torch::Tensor& into_means3D, const torch::Tensor& into_opacity, const torch::Tensor& into_sh,
                                     const torch::Tensor& into_scales, const torch::Tensor& into_rotations, const bool accumulate,
                                     const torch::Tensor& sh_rest, const torch::Tensor& into_sh_rest)
{
    // non-empty `into_*` (all five or none): the summed gradients are written (or, with `accumulate`, added) straight into
    // these buffers - the parameters' .grad - and the corresponding returned tensors are empty
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor");
    const bool scenes = means3D.ndimension() == 3;
    TORCH_CHECK(dL_dout_color.ndimension() == (scenes ? 5 : 4), "dL_dout_color must be [V,3,H,W] ([S,Vs,3,H,W] with scenes)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int S = scenes ? (int)means3D.size(0) : 1, Vs = (int)dL_dout_color.size(scenes ? 1 : 0);
    const int P = (int)means3D.size(scenes ? 1 : 0), V = S * Vs;
    const int H = (int)dL_dout_color.size(scenes ? 3 : 2), W = (int)dL_dout_color.size(scenes ? 4 : 3);
    const int shdim = scenes ? 2 : 1;
    const bool split = sh_rest.numel() != 0;
    const int Mdc = (int)sh.size(shdim);
    const int M = split ? 1 + (int)sh_rest.size(shdim) : Mdc;
    TORCH_CHECK(!split || (Mdc == 1 && sh_rest.dim() == shdim + 2 && sh_rest.size(shdim - 1) == P),
                "split SH: sh [P,1,3] and sh_rest [P,K,3] expected");
    auto f32 = means3D.options().dtype(torch::kFloat32);
    const bool live = P != 0 && V != 0;
    auto alloc = [&](std::initializer_list<int64_t> shape) { return live ? torch::empty(shape, f32) : torch::zeros(shape, f32); };
    const bool into = into_means3D.numel() != 0;
    if (into) {
        auto ok = [&](const torch::Tensor& t, int64_t n) {
            return t.is_cuda() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() && t.numel() == n;
        };
        const int64_t SP = (int64_t)S * P;
        TORCH_CHECK(ok(into_means3D, SP * 3) && ok(into_opacity, SP) && ok(into_sh, SP * Mdc * 3) &&
                        ok(into_scales, SP * 3) && ok(into_rotations, SP * 4) && (!split || ok(into_sh_rest, SP * (M - 1) * 3)),
                    "gradient targets must be contiguous fp32 CUDA tensors with the gradients' element counts");
    }
    auto none = [&]() { return torch::empty({0}, f32); };
    torch::Tensor dL_dmeans2D = alloc({V, P, 3});
    const int64_t SPn = (int64_t)S * P;                    // gradients are allocated flat over the scenes and reshaped at the end
    torch::Tensor dL_dopacity = into ? none() : alloc({SPn, 1}), dL_dmeans3D = into ? none() : alloc({SPn, 3}),
                  dL_dsh = into ? none() : alloc({SPn, Mdc, 3}), dL_dscales = into ? none() : alloc({SPn, 3}),
                  dL_drotations = into ? none() : alloc({SPn, 4}), dL_dsh_rest = (split && !into) ? alloc({SPn, M - 1, 3}) : none();
    if (live) {
        torch::Tensor accum = torch::empty({(long long)(gsr_bwd_accum_bytes(P * V) / sizeof(float))}, f32);
        F32 bg(background), m3(means3D), sc(scales), rot(rotations), vm(viewmatrices), pm(projmatrices), shs(sh), cam(camposs),
            dpix(dL_dout_color), shr(sh_rest);
        torch::Tensor radii_c = radii.contiguous();
        GsrBwdArgs a = {};
        if (split) { a.shs_rest = shr.ptr(); a.dL_dsh_rest = (into ? into_sh_rest : dL_dsh_rest).data_ptr<float>(); }
        a.P = P; a.D = degree; a.M = M; a.R = R; a.W = W; a.H = H;
        a.background = bg.ptr(); a.means3D = m3.ptr(); a.shs = shs.ptr();
        a.scales = sc.ptr(); a.scale_modifier = scale_modifier; a.rotations = rot.ptr();
        a.viewmatrix = vm.ptr(); a.projmatrix = pm.ptr(); a.cam_pos = cam.ptr();
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.radii = radii_c.data_ptr<int>();
        a.geom_blob = geomBuffer.data_ptr(); a.binning_blob = binningBuffer.data_ptr(); a.img_blob = imageBuffer.data_ptr();
        a.dL_dpix = dpix.ptr();
        a.accum = accum.data_ptr();
        a.dL_dmean2D = dL_dmeans2D.data_ptr<float>();
        a.dL_dopacity = (into ? into_opacity : dL_dopacity).data_ptr<float>();
        a.dL_dmean3D = (into ? into_means3D : dL_dmeans3D).data_ptr<float>();
        a.dL_dsh = (into ? into_sh : dL_dsh).data_ptr<float>();
        a.dL_dscale = (into ? into_scales : dL_dscales).data_ptr<float>();
        a.dL_drot = (into ? into_rotations : dL_drotations).data_ptr<float>();
        a.debug = debug;
        if (into && accumulate) a.accumulate = GSR_ACC_MEAN3D | GSR_ACC_OPACITY | GSR_ACC_SH | GSR_ACC_SCALE | GSR_ACC_ROT;
        check(gsr_backward_scenes(&a, S, Vs, current_stream()), "rasterize_gaussians_batched_backward");
    }
    if (scenes) {
        dL_dmeans2D = dL_dmeans2D.view({S, Vs, P, 3});
        if (!into) {
            dL_dopacity = dL_dopacity.view({S, P, 1}); dL_dmeans3D = dL_dmeans3D.view({S, P, 3}); dL_dsh = dL_dsh.view({S, P, Mdc, 3});
            dL_dscales = dL_dscales.view({S, P, 3}); dL_drotations = dL_drotations.view({S, P, 4});
            if (split) dL_dsh_rest = dL_dsh_rest.view({S, P, M - 1, 3});
        }
    }
    return std::make_tuple(dL_dmeans2D, dL_dopacity, dL_dmeans3D, dL_dsh, dL_dscales, dL_drotations, dL_dsh_rest);
}

// fused L1 loss: returns (sum |image - target| as a 1-element tensor, dL/dimage for loss = mean |image - target|)
std::tuple<torch::Tensor, torch::Tensor> l1_loss(const torch::Tensor& image, const torch::Tensor& target)
{
    TORCH_CHECK(image.is_cuda() && target.is_cuda(), "l1_loss: CUDA tensors expected");
    TORCH_CHECK(image.numel() == target.numel(), "l1_loss: size mismatch");
    const c10::cuda::CUDAGuard guard(image.device());
    F32 a(image), b(target);
    const size_t n = (size_t)image.numel();
    auto f32 = image.options().dtype(torch::kFloat32);
    torch::Tensor sum = torch::empty({1}, f32);
    torch::Tensor grad = torch::empty(image.sizes(), f32);
    check(gsr_l1_loss(a.ptr(), b.ptr(), n, n ? 1.0f / (float)n : 0.f, sum.data_ptr<float>(), grad.data_ptr<float>(), current_stream()),
          "l1_loss");
    return std::make_tuple(sum, grad);
}

// fused L1 + D-SSIM: returns (sums[2] = {sum |x-y|, sum ssim_map}, dL/dimage) for
// loss = (1 - lambda) mean|x-y| + lambda (1 - mean ssim_map); with_grad = false returns an empty gradient
std::tuple<torch::Tensor, torch::Tensor> l1_ssim_loss(const torch::Tensor& image, const torch::Tensor& target, double lambda,
                                                     bool with_grad)
{
    TORCH_CHECK(image.is_cuda() && target.is_cuda(), "l1_ssim_loss: CUDA tensors expected");
    TORCH_CHECK(image.dim() == 3 && image.sizes() == target.sizes(), "l1_ssim_loss: two [C,H,W] images expected");
    const c10::cuda::CUDAGuard guard(image.device());
    F32 a(image), b(target);
    const int C = (int)image.size(0), H = (int)image.size(1), W = (int)image.size(2);
    const double n = (double)image.numel();
    auto f32 = image.options().dtype(torch::kFloat32);
    torch::Tensor sums = torch::empty({2}, f32);
    torch::Tensor grad = with_grad ? torch::empty(image.sizes(), f32) : torch::empty({0}, f32);
    torch::Tensor maps = with_grad ? torch::empty({3, C, H, W}, f32) : torch::empty({0}, f32);
    check(gsr_l1_ssim_loss(a.ptr(), b.ptr(), C, H, W, n > 0 ? (float)((1.0 - lambda) / n) : 0.f, n > 0 ? (float)(-lambda / n) : 0.f,
                           sums.data_ptr<float>(), with_grad ? maps.data_ptr<float>() : nullptr,
                           with_grad ? grad.data_ptr<float>() : nullptr, current_stream()),
          "l1_ssim_loss");
    return std::make_tuple(sums, grad);
}

// fused Adam over several parameter groups (one launch); everything is updated in place
void adam_step(std::vector<torch::Tensor> params, std::vector<torch::Tensor> grads, std::vector<torch::Tensor> exp_avgs,
               std::vector<torch::Tensor> exp_avg_sqs, std::vector<int64_t> steps, std::vector<double> lrs, std::vector<int64_t> rows,
               double beta1, double beta2, double eps, const torch::Tensor& visible)
{
    const size_t n = params.size();
    TORCH_CHECK(grads.size() == n && exp_avgs.size() == n && exp_avg_sqs.size() == n && steps.size() == n && lrs.size() == n &&
                rows.size() == n, "adam_step: list lengths differ");
    if (n == 0) return;
    const c10::cuda::CUDAGuard guard(params[0].device());
    std::vector<torch::Tensor> keep;                       // contiguous copies of gradients, if any were needed
    const unsigned char* vis = nullptr;
    torch::Tensor visc;
    if (visible.numel() > 0) {
        TORCH_CHECK(visible.is_cuda() && (visible.scalar_type() == torch::kBool || visible.scalar_type() == torch::kUInt8),
                    "adam_step: visible must be a CUDA bool / uint8 tensor");
        visc = visible.contiguous();
        vis = reinterpret_cast<const unsigned char*>(visc.data_ptr());
    }
    for (size_t base = 0; base < n; base += 8) {
        GsrAdamGroup g[8];
        const int cnt = (int)std::min<size_t>(8, n - base);
        for (int i = 0; i < cnt; i++) {
            const size_t k = base + i;
            for (const torch::Tensor* t : {&params[k], &exp_avgs[k], &exp_avg_sqs[k]})
                TORCH_CHECK(t->is_cuda() && t->scalar_type() == torch::kFloat32 && t->is_contiguous() && t->numel() == params[k].numel(),
                            "adam_step: parameters and moments must be contiguous CUDA float32 tensors of one size");
            TORCH_CHECK(grads[k].is_cuda() && grads[k].scalar_type() == torch::kFloat32 && grads[k].numel() == params[k].numel(),
                        "adam_step: gradient must be a CUDA float32 tensor of the parameter's size");
            torch::Tensor gc = grads[k].contiguous();
            keep.push_back(gc);
            if (vis) TORCH_CHECK(rows[k] >= 1 && params[k].numel() == visible.numel() * rows[k], "adam_step: visible does not match rows");
            g[i].param = params[k].data_ptr<float>();
            g[i].grad = gc.data_ptr<float>();
            g[i].exp_avg = exp_avgs[k].data_ptr<float>();
            g[i].exp_avg_sq = exp_avg_sqs[k].data_ptr<float>();
            g[i].n = (size_t)params[k].numel();
            g[i].row = (int)std::max<int64_t>(1, rows[k]);
            g[i].step = (int)steps[k];
            g[i].lr = (float)lrs[k];
        }
        check(gsr_adam_step(g, cnt, beta1, beta2, eps, vis, current_stream()), "adam_step");
    }
}

void densify_stats(const torch::Tensor& radii, const torch::Tensor& dmean2D, torch::Tensor max_radii2D, torch::Tensor grad_accum,
                   torch::Tensor denom)
{
    const int64_t P = radii.numel();
    TORCH_CHECK(radii.is_cuda() && radii.scalar_type() == torch::kInt32 && radii.is_contiguous(), "densify_stats: radii must be CUDA int32");
    TORCH_CHECK(dmean2D.is_cuda() && dmean2D.scalar_type() == torch::kFloat32 && dmean2D.is_contiguous() && dmean2D.numel() == 3 * P,
                "densify_stats: dL/dmeans2D must be a contiguous CUDA float32 [P,3] tensor");
    for (const torch::Tensor* t : {&max_radii2D, &grad_accum, &denom})
        TORCH_CHECK(t->is_cuda() && t->scalar_type() == torch::kFloat32 && t->is_contiguous() && t->numel() == P,
                    "densify_stats: statistics must be contiguous CUDA float32 tensors of P elements");
    const c10::cuda::CUDAGuard guard(radii.device());
    check(gsr_densify_stats((int)P, radii.data_ptr<int>(), dmean2D.data_ptr<float>(), max_radii2D.data_ptr<float>(),
                            grad_accum.data_ptr<float>(), denom.data_ptr<float>(), current_stream()), "densify_stats");
}

// out[r, j] = in[r, cols[j]] (cols[j] < 0: 0) for a [rows, in_cols] CUDA float tensor; returns [rows, len(cols)]
torch::Tensor gather_columns(const torch::Tensor& in, std::vector<int> cols)
{
    TORCH_CHECK(in.is_cuda() && in.scalar_type() == torch::kFloat32 && in.dim() == 2 && in.is_contiguous(),
                "gather_columns: contiguous CUDA float32 [rows, cols] tensor expected");
    TORCH_CHECK(cols.size() <= 64, "gather_columns: at most 64 output columns");
    const c10::cuda::CUDAGuard guard(in.device());
    torch::Tensor out = torch::empty({in.size(0), (int64_t)cols.size()}, in.options());
    check(gsr_gather_columns(in.data_ptr<float>(), (int)in.size(1), out.data_ptr<float>(), (int)cols.size(), cols.data(),
                             (long long)in.size(0), current_stream()), "gather_columns");
    return out;
}

// (exp(scaling), normalize(rotation), sigmoid(opacity)) in one launch
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor> activate(const torch::Tensor& scaling, const torch::Tensor& rotation,
                                                                 const torch::Tensor& opacity)
{
    TORCH_CHECK(scaling.is_cuda() && rotation.is_cuda() && opacity.is_cuda(), "activate: CUDA tensors expected");
    const int64_t P = scaling.size(0);
    TORCH_CHECK(scaling.numel() == 3 * P && rotation.numel() == 4 * P && opacity.numel() == P, "activate: [P,3], [P,4], [P,1] expected");
    const c10::cuda::CUDAGuard guard(scaling.device());
    F32 s(scaling), q(rotation), o(opacity);
    auto f32 = scaling.options().dtype(torch::kFloat32);
    torch::Tensor so = torch::empty({P, 3}, f32), qo = torch::empty({P, 4}, f32), oo = torch::empty({P, 1}, f32);
    check(gsr_activate_forward((int)P, s.ptr(), q.ptr(), o.ptr(), so.data_ptr<float>(), qo.data_ptr<float>(), oo.data_ptr<float>(),
                               current_stream()), "activate");
    return std::make_tuple(so, qo, oo);
}

// gradients of the raw parameters; a non-empty `into_*` (contiguous fp32, the leaf's .grad) is written / added to in place
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor>
activate_backward(const torch::Tensor& scaling, const torch::Tensor& rotation, const torch::Tensor& opacity, const torch::Tensor& g_s,
                  const torch::Tensor& g_q, const torch::Tensor& g_o, const torch::Tensor& into_s, const torch::Tensor& into_q,
                  const torch::Tensor& into_o, bool accumulate)
{
    const int64_t P = scaling.size(0);
    const c10::cuda::CUDAGuard guard(scaling.device());
    F32 s(scaling), q(rotation), o(opacity), gs(g_s), gq(g_q), go(g_o);
    auto f32 = scaling.options().dtype(torch::kFloat32);
    const bool all_into = into_s.numel() && into_q.numel() && into_o.numel();
    TORCH_CHECK(all_into || !(into_s.numel() || into_q.numel() || into_o.numel()), "activate_backward: give all three targets or none");
    torch::Tensor ds, dq, dop;
    if (all_into) {
        for (const torch::Tensor* t : {&into_s, &into_q, &into_o})
            TORCH_CHECK(t->is_cuda() && t->scalar_type() == torch::kFloat32 && t->is_contiguous(), "activate_backward: bad gradient target");
        TORCH_CHECK(into_s.numel() == 3 * P && into_q.numel() == 4 * P && into_o.numel() == P, "activate_backward: target sizes");
        ds = into_s; dq = into_q; dop = into_o;
    } else {
        ds = torch::empty({P, 3}, f32); dq = torch::empty({P, 4}, f32); dop = torch::empty({P, 1}, f32);
    }
    check(gsr_activate_backward((int)P, s.ptr(), q.ptr(), o.ptr(), gs.ptr(), gq.ptr(), go.ptr(), ds.data_ptr<float>(),
                                dq.data_ptr<float>(), dop.data_ptr<float>(), (all_into && accumulate) ? 1 : 0, current_stream()),
          "activate_backward");
    if (all_into) { torch::Tensor e = torch::empty({0}, f32); return std::make_tuple(e, e, e); }
    return std::make_tuple(ds, dq, dop);
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
    m.def("activate", &activate);
    m.def("activate_backward", &activate_backward);
    m.def("gather_columns", &gather_columns);
    m.def("adam_step", &adam_step);
    m.def("densify_stats", &densify_stats);
    m.def("l1_ssim_loss", &l1_ssim_loss);
    m.def("l1_loss", &l1_loss);
    // The rasterizer entry points release the GIL (they touch no Python object): the forward blocks the host once for
    // num_rendered, and a trainer driving two CUDA streams from two Python threads must not serialise on that wait.
    // (The reference binding holds the GIL throughout: SURVEY 8(b).)
    using nogil = py::call_guard<py::gil_scoped_release>;
    m.def("rasterize_gaussians_batched", &rasterize_gaussians_batched, nogil());
    m.def("rasterize_gaussians_batched_backward", &rasterize_gaussians_batched_backward, nogil());
    m.def("rasterize_gaussians", &rasterize_gaussians, nogil());
    m.def("rasterize_gaussians_split", &rasterize_gaussians_split, nogil());
    m.def("rasterize_gaussians_front", &rasterize_gaussians_front, nogil());
    m.def("prepare_views", &prepare_views, nogil());
    m.def("rasterize_gaussians_prepared", &rasterize_gaussians_prepared, nogil());
    m.def("rasterize_gaussians_backward_split", &rasterize_gaussians_backward_split, nogil());
    m.def("rasterize_gaussians_backward", &rasterize_gaussians_backward, nogil());
    m.def("rasterize_gaussians_backward_into", &rasterize_gaussians_backward_into, nogil());
    m.def("rasterize_gaussians_backward_chunked", &rasterize_gaussians_backward_chunked);
    m.def("rasterize_gaussians_backward_blend", &rasterize_gaussians_backward_blend, nogil());
    m.def("rasterize_gaussians_backward_views", &rasterize_gaussians_backward_views);
    m.def("set_option", [](const std::string& name, int value) {
        int prev = 0;
        if (gsr_debug_get_option(name.c_str(), &prev) != GSR_OK || gsr_debug_set_option(name.c_str(), value) != GSR_OK)
            throw std::runtime_error("unknown gsr option: " + name);
        return prev;
    });
    m.def("get_option", [](const std::string& name) {
        int v = 0;
        if (gsr_debug_get_option(name.c_str(), &v) != GSR_OK) throw std::runtime_error("unknown gsr option: " + name);
        return v;
    });
    m.def("arms_loaded", []() { return gsr_debug_arms_loaded() != 0; });
    m.def("mark_visible", &mark_visible);
    m.def("distCUDA2", &distCUDA2);
    m.def("export_binning", &export_binning);
    m.def("launch_count", []() { return (unsigned long long)gsr_launch_count(); });
    m.def("reset_launch_count", []() { gsr_reset_launch_count(); });
    m.def("abi_version", []() { return gsr_abi_version(); });
    m.def("timing_enable", [](bool on) { gsr_timing_enable(on ? 1 : 0); });
    m.def("timing_reset", []() { gsr_timing_reset(); });
    m.def("timing_collect", []() {
        check(gsr_timing_collect(), "timing_collect");
        std::vector<std::tuple<std::string, double, unsigned long long>> out;
        for (int i = 0; i < GSR_NUM_STAGES; i++) {
            double ms = 0; unsigned long long n = 0;
            gsr_timing_read(i, &ms, &n);
            out.emplace_back(gsr_stage_name(i), ms, n);
        }
        return out;
    });
}
