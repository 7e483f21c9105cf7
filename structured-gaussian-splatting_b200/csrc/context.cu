// context.cu - the state the library keeps between calls: tuning options (read from the environment once), the hook
// table of the test-only comparison arms, and one context per CUDA device (side streams, pinned read-back slots, the
// gradient fence).  The reference core keeps no state at all (static methods, rasterizer.h:24-84: "re-entrant per
// device"); ours keeps only reusable CUDA resources, per device and guarded for concurrent host threads.
#include "gsr_common.cuh"
#include <cstdlib>
#include <cstring>

namespace gsr {

static int env_int(const char* name, int dflt)
{
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

struct OptionName { const char* name; int Options::*field; };
static const OptionName kOptionNames[] = {
    {"BWD_VARIANT", &Options::bwd_variant}, {"FWD_VARIANT", &Options::fwd_variant}, {"LPT", &Options::lpt}, {"FWD_MINB", &Options::fwd_minb},
    {"BWD_MINB", &Options::bwd_minb}, {"SIDE_STREAM", &Options::side_stream}, {"BWD_SIDE", &Options::bwd_side},
    {"SH_SPLIT", &Options::sh_split}, {"PRE_STAGE", &Options::pre_stage}, {"DEPTH_SEG_ITEMS", &Options::depth_seg_items}, {"RADIX_SPLIT", &Options::radix_split}, {"RADIX_DEPTH", &Options::radix_depth}, {"BLEND_ARM", &Options::blend_arm},
    {"SORT_CUB", &Options::sort_cub}, {"BLEND_STATS", &Options::blend_stats},
};

Options& options()
{
    static Options o;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const OptionName& n : kOptionNames) {
            char env[64] = "GSR_";
            strncat(env, n.name, sizeof env - 5);
            o.*(n.field) = env_int(env, o.*(n.field));
        }
    });
    return o;
}

ArmHooks g_arms = {};
void register_arms(const ArmHooks& h) { g_arms = h; }

// ------------------------------------------------------------------ per-device contexts
constexpr int MAX_DEVICES = 64;
static DeviceCtx* g_ctx[MAX_DEVICES];
static std::mutex g_ctx_mu;

int device_ctx(DeviceCtx** out)
{
    int dev = 0;
    GSR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return GSR_EINVAL;
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    if (!g_ctx[dev]) {
        DeviceCtx* c = new DeviceCtx();
        c->device = dev;
        cudaError_t e = cudaEventCreateWithFlags(&c->grad_fence, cudaEventDisableTiming);
        if (e != cudaSuccess) { delete c; return record_cuda_error(e, "cudaEventCreateWithFlags(grad_fence)"); }
        g_ctx[dev] = c;
    }
    *out = g_ctx[dev];
    return GSR_OK;
}

int ctx_side_stream(DeviceCtx* c, int which, cudaStream_t user, SideStream* out)
{
    std::lock_guard<std::mutex> lk(c->mu);
    for (const SideStream& s : c->sides[which])
        if (s.user == user) { *out = s; return GSR_OK; }
    SideStream s;
    s.user = user;
    if (c->sides[which].size() >= 16) {
        // many short-lived caller streams: share the oldest entry's resources (correct, merely less concurrent)
        s = c->sides[which][0];
        *out = s;
        return GSR_OK;
    }
    // the helper stream runs pieces of the caller's own work: it inherits the caller stream's priority
    int prio = 0;
    if (cudaStreamGetPriority(user, &prio) != cudaSuccess) { cudaGetLastError(); prio = 0; }
    GSR_CUDA(cudaStreamCreateWithPriority(&s.side, cudaStreamNonBlocking, prio));
    GSR_CUDA(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
    GSR_CUDA(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
    c->sides[which].push_back(s);
    *out = s;
    return GSR_OK;
}

int ctx_acquire_rslot(DeviceCtx* c, RSlot** out)
{
    std::lock_guard<std::mutex> lk(c->mu);
    for (RSlot* r : c->rslots)
        if (!r->busy) { r->busy = true; *out = r; return GSR_OK; }
    RSlot* r = new RSlot();
    cudaError_t e = cudaHostAlloc((void**)&r->host, 64, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev, cudaEventDisableTiming);
    if (e != cudaSuccess) { delete r; return record_cuda_error(e, "R read-back slot"); }
    r->busy = true;
    c->rslots.push_back(r);
    *out = r;
    return GSR_OK;
}

void ctx_release_rslot(DeviceCtx* c, RSlot* r)
{
    std::lock_guard<std::mutex> lk(c->mu);
    r->busy = false;
}

}  // namespace gsr

// debug hooks (not part of include/gsr.h): tests and benchmarks switch comparison arms / tuning values at run time
extern "C" int gsr_debug_set_option(const char* name, int value)
{
    if (!name) return GSR_EINVAL;
    gsr::Options& o = gsr::options();
    for (const gsr::OptionName& n : gsr::kOptionNames)
        if (strcasecmp(n.name, name) == 0) { o.*(n.field) = value; return GSR_OK; }
    return GSR_EINVAL;
}
extern "C" int gsr_debug_get_option(const char* name, int* value)
{
    if (!name || !value) return GSR_EINVAL;
    gsr::Options& o = gsr::options();
    for (const gsr::OptionName& n : gsr::kOptionNames)
        if (strcasecmp(n.name, name) == 0) { *value = o.*(n.field); return GSR_OK; }
    return GSR_EINVAL;
}
extern "C" int gsr_debug_arms_loaded(void) { return gsr::g_arms.blend_forward_plain != nullptr; }
