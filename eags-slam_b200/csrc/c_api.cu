// c_api.cu -- the extern "C" boundary declared in include/eags_raster.h.
// Stateless apart from two process-wide tuning knobs and a launch counter; never allocates, never
// synchronises; every launch goes to the stream handed in.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "kernels.h"

namespace {

thread_local char g_err[512] = "";
std::atomic<int> g_opt_tma{1};
std::atomic<int> g_opt_agg{1};
std::atomic<int> g_opt_radix{2};
std::atomic<int> g_opt_stats{0};
std::atomic<unsigned long long> g_launches{0};

// Optional per-kernel timing ("kernel_timing" option): CUDA events recorded on the caller's stream
// around every launcher, read back (and synchronised) by eags_kernel_times().  Off by default: the
// product path never records events.
std::atomic<int> g_opt_timing{0};
struct TimedSpan { const char* name; cudaEvent_t beg, end; };
// process-wide (mutex-guarded), not thread-local: autograd runs the backward launchers on its own device
// thread, and the caller of eags_kernel_times() wants those spans too
std::mutex g_span_mu;
std::vector<TimedSpan> g_spans;
std::vector<cudaEvent_t> g_event_pool;
cudaEvent_t take_event() {
    {
        std::lock_guard<std::mutex> lk(g_span_mu);
        if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
struct ScopedSpan {
    cudaStream_t s; bool on; TimedSpan span;
    ScopedSpan(const char* name, cudaStream_t st) : s(st), on(g_opt_timing.load() != 0) {
        if (on) { span.name = name; span.beg = take_event(); span.end = take_event(); cudaEventRecord(span.beg, s); }
    }
    ~ScopedSpan() {
        if (on) {
            cudaEventRecord(span.end, s);
            std::lock_guard<std::mutex> lk(g_span_mu);
            g_spans.push_back(span);
        }
    }
};

int fail(int code, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s", what);
    return code;
}
int fail_cuda(const char* where) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaErrorUnknown;
    snprintf(g_err, sizeof(g_err), "%s: %s", where, cudaGetErrorString(e));
    return (int)e;
}
#define CHECK_LAUNCH(expr, where)                 \
    do {                                          \
        int n__;                                  \
        { ScopedSpan span__(where, s); n__ = (expr); } \
        if (n__ < 0) return fail_cuda(where);     \
        g_launches.fetch_add((unsigned)n__);      \
    } while (0)
#define CHECK_CUDA(expr, where)                              \
    do {                                                     \
        if ((expr) != cudaSuccess) return fail_cuda(where);  \
    } while (0)

bool view_ok(const EagsView* v) {
    return v && v->image_height > 0 && v->image_width > 0 && v->bg && v->viewmatrix && v->projmatrix &&
           v->campos && v->sh_degree >= 0 && v->sh_degree <= 3 &&
           (v->variant == EAGS_VARIANT_A || v->variant == EAGS_VARIANT_B);
}
bool gauss_ok(const EagsView* v, const EagsGaussians* g) {
    if (!g || g->P < 0) return false;
    if (g->P == 0) return true;
    if (!g->means3D || !g->opacities) return false;
    if ((g->shs != nullptr) == (g->colors_precomp != nullptr)) return false;
    const bool sr = g->scales && g->rotations;
    if (sr == (g->cov3D_precomp != nullptr)) return false;
    if (g->shs && g->M < (v->sh_degree + 1) * (v->sh_degree + 1)) return false;
    return true;
}

}  // namespace

extern "C" {

int eags_abi_version(void) { return EAGS_ABI_VERSION; }
const char* eags_last_error(void) { return g_err; }
uint64_t eags_launch_count(void) { return g_launches.load(); }

size_t eags_geom_state_bytes(int32_t P) { return geom_layout(P < 0 ? 0 : P).total; }
size_t eags_image_state_bytes(int32_t W, int32_t H) { return image_layout(W < 1 ? 1 : W, H < 1 ? 1 : H).total; }
size_t eags_binning_state_bytes(int64_t capacity, int32_t with_keys) {
    return binning_layout(capacity, with_keys).total;
}
size_t eags_backward_scratch_bytes(int32_t P) { return scratch_layout(P < 0 ? 0 : P).total; }
size_t eags_backward_stats_offset(int32_t P) { return scratch_layout(P < 0 ? 0 : P).pose + EAGS_SCRATCH_STATS_OFF; }

int eags_set_option(const char* name, int32_t value) {
    if (name && !strcmp(name, "tma")) { g_opt_tma.store(value ? 1 : 0); return 0; }
    if (name && !strcmp(name, "kernel_timing")) { g_opt_timing.store(value ? 1 : 0); return 0; }
    if (name && !strcmp(name, "preprocess_agg")) { g_opt_agg.store(value ? 1 : 0); return 0; }
    if (name && !strcmp(name, "radix_sort")) { g_opt_radix.store(value < 0 ? 0 : (value > 2 ? 2 : value)); return 0; }
    if (name && !strcmp(name, "kernel_stats")) { g_opt_stats.store(value ? 1 : 0); return 0; }
    return fail(EAGS_ERR_INVALID, "eags_set_option: unknown option");
}
int eags_get_option(const char* name, int32_t* value) {
    if (name && value && !strcmp(name, "tma")) { *value = g_opt_tma.load(); return 0; }
    if (name && value && !strcmp(name, "kernel_timing")) { *value = g_opt_timing.load(); return 0; }
    if (name && value && !strcmp(name, "preprocess_agg")) { *value = g_opt_agg.load(); return 0; }
    if (name && value && !strcmp(name, "radix_sort")) { *value = g_opt_radix.load(); return 0; }
    if (name && value && !strcmp(name, "kernel_stats")) { *value = g_opt_stats.load(); return 0; }
    return fail(EAGS_ERR_INVALID, "eags_get_option: unknown option");
}

int eags_kernel_times(int32_t max_spans, const char** names, float* ms) {
    std::vector<TimedSpan> spans;
    {
        std::lock_guard<std::mutex> lk(g_span_mu);
        spans.swap(g_spans);
    }
    int n = 0;
    for (TimedSpan& sp : spans) {
        float t = 0.f;
        cudaEventSynchronize(sp.end);
        cudaEventElapsedTime(&t, sp.beg, sp.end);
        if (n < max_spans && names && ms) { names[n] = sp.name; ms[n] = t; n++; }
    }
    std::lock_guard<std::mutex> lk(g_span_mu);
    for (TimedSpan& sp : spans) { g_event_pool.push_back(sp.beg); g_event_pool.push_back(sp.end); }
    return n;
}

// Shared body of the two forward flavours.  tile_cap == 0: exact path (count -> scan -> duplicate ->
// sort).  tile_cap > 0: bucketed path -- preprocess_fwd scatters every key into its tile's fixed-stride
// bucket of `tile_cap` slots inside `binning_state` and there is no duplicate kernel; if a tile
// outgrows its bucket a device flag makes every later kernel exit and the caller re-issues stage 2 on
// the exact path (the tile scan has already produced the exact cursors).
static int forward_geom_impl(const EagsView* view, const EagsGaussians* g, void* geom_state,
                             void* image_state, int32_t* radii, void* ctrl_pinned, size_t ctrl_bytes,
                             void* binning_state, int64_t capacity, int32_t with_keys, int32_t tile_cap,
                             void* sticky_dev, void* stream) {
    if (!view_ok(view)) return fail(EAGS_ERR_INVALID, "eags_forward_geom: bad view settings");
    if (!gauss_ok(view, g)) return fail(EAGS_ERR_INVALID, "eags_forward_geom: bad Gaussian inputs");
    if (!image_state || (g->P > 0 && (!geom_state || !radii)))
        return fail(EAGS_ERR_INVALID, "eags_forward_geom: null state buffer");
    if (tile_cap < 0 || (tile_cap > 0 && (!binning_state || capacity < 0)))
        return fail(EAGS_ERR_INVALID, "eags_forward_geom_bucketed: bad bucket arguments");
    cudaStream_t s = (cudaStream_t)stream;
    const int W = view->image_width, H = view->image_height;
    ImageLayout IL = image_layout(W, H);
    ImagePtrs img = image_ptrs(image_state, W, H);
    GeomPtrs geom = geom_ptrs(geom_state, g->P);
    unsigned long long* bucket = nullptr;
    if (tile_cap > 0) bucket = binning_ptrs(binning_state, capacity, with_keys, (int64_t)IL.G * tile_cap).unsorted;
    // ctrl + tile_count are contiguous: one memset node
    CHECK_CUDA(cudaMemsetAsync(image_state, 0, IL.tile_count + (size_t)IL.G * 4 * EAGS_CTR_STRIDE, s), "memset(ctrl)");
    CHECK_LAUNCH(launch_preprocess_fwd(*view, *g, geom, img, radii, bucket, (unsigned)tile_cap, g_opt_agg.load(), s), "preprocess_fwd");
    // The status block goes to the host from inside tile_scan when the caller's buffer is pinned host memory the
    // device can address (every cudaHostAlloc / cudaHostRegister'ed block under unified addressing; torch's
    // pin_memory()); anything else (pageable memory) keeps the copy node.
    void* status_dev = nullptr;
    if (ctrl_pinned && ctrl_bytes % 8 == 0 && ctrl_bytes <= sizeof(EagsCtrl)) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, ctrl_pinned) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
            status_dev = at.devicePointer;
        else
            cudaGetLastError();
    }
    CHECK_LAUNCH(launch_tile_scan(img, (EagsCtrl*)sticky_dev, status_dev, ctrl_bytes, s), "tile_scan");
    if (ctrl_pinned && !status_dev)
        CHECK_CUDA(cudaMemcpyAsync(ctrl_pinned, sticky_dev ? sticky_dev : (void*)img.ctrl, ctrl_bytes,
                                   cudaMemcpyDeviceToHost, s), "memcpy(ctrl)");
    return 0;
}

static int forward_render_impl(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                               void* geom_state, void* image_state, void* binning_state, int64_t capacity,
                               int32_t with_keys, int32_t tile_cap, const EagsImages* out, void* stream) {
    if (!view_ok(view)) return fail(EAGS_ERR_INVALID, "eags_forward_render: bad view settings");
    if (!g || g->P < 0 || !image_state || !binning_state || capacity < 0 || !out || !out->color ||
        !out->depth || !out->third || (g->P > 0 && (!geom_state || !radii)) || tile_cap < 0)
        return fail(EAGS_ERR_INVALID, "eags_forward_render: null buffer");
    if (capacity >= (1ll << 32)) return fail(EAGS_ERR_INVALID, "eags_forward_render: capacity must be < 2^32");
    cudaStream_t s = (cudaStream_t)stream;
    const int W = view->image_width, H = view->image_height;
    ImagePtrs img = image_ptrs(image_state, W, H);
    GeomPtrs geom = geom_ptrs(geom_state, g->P);
    BinningPtrs bin = binning_ptrs(binning_state, capacity, with_keys, (int64_t)img.G * tile_cap);
    if (out->n_touched && g->P > 0)
        CHECK_CUDA(cudaMemsetAsync(out->n_touched, 0, (size_t)g->P * 4, s), "memset(n_touched)");
    if (tile_cap == 0) CHECK_LAUNCH(launch_duplicate(*view, g->P, radii, geom, img, bin, s), "duplicate");
    CHECK_LAUNCH(launch_sort_pack(geom, g->P, img, bin, (unsigned)tile_cap, g_opt_radix.load(), s), "sort_pack");
    CHECK_LAUNCH(launch_blend_fwd(*view, img, bin, *out, g_opt_tma.load(), g_opt_stats.load(), (unsigned)tile_cap, s), "blend_fwd");
    return 0;
}

int eags_forward_geom(const EagsView* view, const EagsGaussians* g, void* geom_state,
                      void* image_state, int32_t* radii, uint64_t* num_rendered_pinned, void* stream) {
    return forward_geom_impl(view, g, geom_state, image_state, radii, num_rendered_pinned, sizeof(uint64_t),
                             nullptr, 0, 0, 0, nullptr, stream);
}

int eags_forward_render(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                        void* geom_state, void* image_state, void* binning_state, int64_t capacity,
                        int32_t with_keys, const EagsImages* out, void* stream) {
    return forward_render_impl(view, g, radii, geom_state, image_state, binning_state, capacity, with_keys, 0,
                               out, stream);
}

size_t eags_binning_state_bytes_bucketed(int64_t capacity, int32_t with_keys, int32_t width, int32_t height,
                                         int32_t tile_cap) {
    ImageLayout IL = image_layout(width, height);
    return binning_layout(capacity, with_keys, (int64_t)IL.G * (tile_cap > 0 ? tile_cap : 0)).total;
}

int eags_forward_geom_bucketed(const EagsView* view, const EagsGaussians* g, void* geom_state,
                               void* image_state, int32_t* radii, EagsForwardStatus* status_pinned,
                               void* binning_state, int64_t capacity, int32_t with_keys, int32_t tile_cap,
                               EagsForwardStatus* sticky_status_dev, void* stream) {
    static_assert(sizeof(EagsForwardStatus) == sizeof(EagsCtrl), "status block mirrors the device control block");
    return forward_geom_impl(view, g, geom_state, image_state, radii, status_pinned, sizeof(EagsCtrl),
                             binning_state, capacity, with_keys, tile_cap, sticky_status_dev, stream);
}

int eags_forward_render_bucketed(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                                 void* geom_state, void* image_state, void* binning_state, int64_t capacity,
                                 int32_t with_keys, int32_t tile_cap, const EagsImages* out, void* stream) {
    return forward_render_impl(view, g, radii, geom_state, image_state, binning_state, capacity, with_keys,
                               tile_cap, out, stream);
}

static bool route_ok(const EagsPeerRoute* r, int64_t P) {
    if (!r || r->world < 1 || r->world > EAGS_MAX_PEERS || r->rank < 0 || r->rank >= r->world || r->P != P)
        return false;
    for (int k = 0; k < r->world; k++)
        if (!r->recv[k] || !r->reduced[k] || !r->flags[k]) return false;
    const int n_acc = (r->accum[0] != nullptr) + (r->accum[1] != nullptr) + (r->accum[2] != nullptr) + (r->accum[3] != nullptr);
    if (n_acc != 0 && n_acc != 4) return false;                         // all four or none
    if (n_acc && (reinterpret_cast<uintptr_t>(r->accum[2]) & 15)) return false;   // rotation rows are read as float4
    return true;
}

int eags_backward(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                  const void* geom_state, const void* image_state, const void* binning_state,
                  int64_t capacity, const EagsImageGrads* gin, const EagsInputGrads* gout,
                  void* scratch, void* stream) {
    return eags_backward_routed(view, g, radii, geom_state, image_state, binning_state, capacity, gin, gout,
                                scratch, nullptr, stream);
}

size_t eags_peer_flags_bytes(void) { return peer_flags_bytes(); }

int eags_peer_reduce(const EagsPeerRoute* route, uint32_t epoch, float scale, int32_t wait, void* stream) {
    if (!route || !route_ok(route, route->P) || route->P < 0)
        return fail(EAGS_ERR_INVALID, "eags_peer_reduce: bad route");
    if (route->P == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_peer_reduce(*route, epoch, scale, wait, s), "peer_reduce");
    return 0;
}

int eags_peer_adam_step(const EagsPeerRoute* route, const EagsPeerAdam* opt, uint32_t epoch, float scale,
                        int32_t wait, void* stream) {
    if (!route || !route_ok(route, route->P) || route->P < 0)
        return fail(EAGS_ERR_INVALID, "eags_peer_adam_step: bad route");
    if (!opt || !opt->m || !opt->v || !opt->step || !(opt->beta1 >= 0.f && opt->beta1 < 1.f) ||
        !(opt->beta2 >= 0.f && opt->beta2 < 1.f) || !(opt->eps >= 0.f))
        return fail(EAGS_ERR_INVALID, "eags_peer_adam_step: bad optimiser state");
    for (int k = 0; k < route->world; k++)
        if (!opt->params[k] || (reinterpret_cast<uintptr_t>(opt->params[k]) & 15))
            return fail(EAGS_ERR_INVALID, "eags_peer_adam_step: parameter replicas must be 16-byte aligned");
    if (route->P == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_peer_adam(*route, *opt, epoch, scale, wait, s), "peer_adam");
    return 0;
}

int eags_backward_routed(const EagsView* view, const EagsGaussians* g, const int32_t* radii,
                         const void* geom_state, const void* image_state, const void* binning_state,
                         int64_t capacity, const EagsImageGrads* gin, const EagsInputGrads* gout,
                         void* scratch, const EagsPeerRoute* route, void* stream) {
    if (route && (!g || !route_ok(route, g->P) || g->cov3D_precomp))
        return fail(EAGS_ERR_INVALID, "eags_backward_routed: bad route (or cov3D_precomp inputs)");
    if (!view_ok(view)) return fail(EAGS_ERR_INVALID, "eags_backward: bad view settings");
    if (!gauss_ok(view, g)) return fail(EAGS_ERR_INVALID, "eags_backward: bad Gaussian inputs");
    if (!gin || !gin->dL_dcolor || !gin->dL_ddepth || !gout)
        return fail(EAGS_ERR_INVALID, "eags_backward: null gradient struct");
    cudaStream_t s = (cudaStream_t)stream;
    if (g->P == 0) {
        if (gout->dL_dtau) CHECK_CUDA(cudaMemsetAsync(gout->dL_dtau, 0, 6 * sizeof(float), s), "memset(dtau)");
        return 0;
    }
    if (!radii || !geom_state || !image_state || !binning_state || !scratch)
        return fail(EAGS_ERR_INVALID, "eags_backward: null buffer");
    const int W = view->image_width, H = view->image_height;
    ImagePtrs img = image_ptrs(const_cast<void*>(image_state), W, H);
    GeomPtrs geom = geom_ptrs(const_cast<void*>(geom_state), g->P);
    BinningPtrs bin = binning_ptrs(const_cast<void*>(binning_state), capacity, 0);
    ScratchLayout SL = scratch_layout(g->P);
    float* grad_acc = (float*)((char*)scratch + SL.grad_acc);
    double* pose_acc = (double*)((char*)scratch + SL.pose);
    CHECK_CUDA(cudaMemsetAsync(scratch, 0, SL.pose + 256, s), "memset(scratch)");
    CHECK_LAUNCH(launch_blend_bwd(*view, img, bin, *gin, grad_acc, (unsigned*)((char*)scratch + SL.pose + EAGS_SCRATCH_WORK_OFF),
                                  g_opt_stats.load() ? (unsigned long long*)((char*)scratch + SL.pose + EAGS_SCRATCH_STATS_OFF) : nullptr,
                                  g_opt_tma.load(), s), "blend_bwd");
    CHECK_LAUNCH(launch_preprocess_bwd(*view, *g, radii, geom, grad_acc, *gout, pose_acc, route, s), "preprocess_bwd");
    return 0;
}

int eags_pose_compose(const float* q, const float* t, const float* viewmatrix, const float* projmatrix,
                      const float* campos, float* viewmatrix_out, float* projmatrix_out,
                      float* campos_out, void* stream) {
    if (!q || !t || !viewmatrix || !projmatrix || !campos || !viewmatrix_out || !projmatrix_out || !campos_out)
        return fail(EAGS_ERR_INVALID, "eags_pose_compose: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_pose_compose(q, t, viewmatrix, projmatrix, campos, viewmatrix_out, projmatrix_out,
                                     campos_out, s), "pose_compose");
    return 0;
}

int eags_pose_chain(const float* q, const float* t, const float* viewmatrix, const float* dL_dtau,
                    float* dL_dq, float* dL_dt, void* stream) {
    if (!q || !t || !viewmatrix || !dL_dtau || !dL_dq || !dL_dt)
        return fail(EAGS_ERR_INVALID, "eags_pose_chain: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_pose_chain(q, t, viewmatrix, dL_dtau, dL_dq, dL_dt, s), "pose_chain");
    return 0;
}

int eags_pose_se3_update(float* theta, float* rho, float* viewmatrix, const float* projmatrix_raw,
                         float* projmatrix_out, float* campos_out, float* tau_norm, void* stream) {
    if (!theta || !rho || !viewmatrix || !projmatrix_raw || !projmatrix_out || !campos_out)
        return fail(EAGS_ERR_INVALID, "eags_pose_se3_update: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_pose_se3_update(theta, rho, viewmatrix, projmatrix_raw, projmatrix_out, campos_out, tau_norm, s),
                 "pose_se3_update");
    return 0;
}

int eags_pose_adam_se3_update(const float* grad_theta, const float* grad_rho, float* state13, float lr_rot,
                              float lr_trans, float beta1, float beta2, float eps, float* viewmatrix,
                              const float* projmatrix_raw, float* projmatrix_out, float* campos_out, float* tau_norm,
                              void* stream) {
    if (!grad_theta || !grad_rho || !state13 || !viewmatrix || !projmatrix_raw || !projmatrix_out || !campos_out)
        return fail(EAGS_ERR_INVALID, "eags_pose_adam_se3_update: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_pose_adam_se3(grad_theta, grad_rho, state13, lr_rot, lr_trans, beta1, beta2, eps, viewmatrix,
                                      projmatrix_raw, projmatrix_out, campos_out, tau_norm, s), "pose_adam_se3");
    return 0;
}

size_t eags_image_l1_scratch_bytes(int32_t W, int32_t H) { return image_l1_scratch_bytes(W, H); }

int eags_image_l1_forward(const EagsImageL1* in, float* g_color, float* g_depth, float* g_alpha,
                          void* scratch, float* sums4, void* stream) {
    if (!in || in->width <= 0 || in->height <= 0 || !in->color || !in->depth || !in->gt_color ||
        !in->gt_depth || !g_color || !g_depth || !scratch || !sums4 || (in->soft_alpha && !in->alpha) ||
        (long long)in->width * in->height >= (1ll << 30))
        return fail(EAGS_ERR_INVALID, "eags_image_l1_forward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_image_l1_fwd(in->width, in->height, in->soft_alpha, in->mask_color, in->norm_mode,
                                     in->eps, in->color, in->depth, in->alpha, in->gt_color, in->gt_depth,
                                     in->mask, in->mask_depth, in->alpha_min, g_color, g_depth, g_alpha, scratch, sums4, s),
                 "image_l1_fwd");
    return 0;
}

int eags_image_l1_forward_static(const EagsImageL1* in, float* g_color, float* g_depth, float* g_alpha,
                                 float* sums4, void* stream) {
    if (!in || in->width <= 0 || in->height <= 0 || !in->color || !in->depth || !in->gt_color ||
        !in->gt_depth || !g_color || !g_depth || !sums4 || (in->soft_alpha && !in->alpha))
        return fail(EAGS_ERR_INVALID, "eags_image_l1_forward_static: bad argument");
    if ((long long)in->width * in->height > (long long)EAGS_IMAGE_L1_STATIC_MAX_PIXELS)
        return fail(EAGS_ERR_INVALID, "eags_image_l1_forward_static: image larger than EAGS_IMAGE_L1_STATIC_MAX_PIXELS "
                                      "(4194304 pixels); use eags_image_l1_forward with caller-owned scratch");
    cudaStream_t s = (cudaStream_t)stream;
    // (no ScopedSpan: the kernel_timing option creates CUDA events, and this entry point creates nothing)
    const int n = launch_image_l1_fwd_static(in->width, in->height, in->soft_alpha, in->mask_color, in->norm_mode,
                                             in->eps, in->color, in->depth, in->alpha, in->gt_color, in->gt_depth,
                                             in->mask, in->mask_depth, in->alpha_min, g_color, g_depth, g_alpha, sums4, s);
    if (n == -2) {
        cudaGetLastError();
        return fail(EAGS_ERR_INVALID, "eags_image_l1_forward_static: cudaGetSymbolAddress of the library's scratch array "
                                      "failed on the current device");
    }
    if (n < 0) return fail_cuda("image_l1_fwd (static scratch)");
    g_launches.fetch_add((unsigned)n);
    return 0;
}

int eags_image_l1_backward(int32_t W, int32_t H, const float* sums4, const float* up_color,
                           const float* up_depth, float* g_color, float* g_depth, float* g_alpha,
                           void* stream) {
    if (W <= 0 || H <= 0 || !sums4 || !up_color || !up_depth || !g_color || !g_depth)
        return fail(EAGS_ERR_INVALID, "eags_image_l1_backward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_image_l1_scale(W, H, sums4, up_color, up_depth, g_color, g_depth, g_alpha, s),
                 "image_l1_scale");
    return 0;
}

size_t eags_neighbors_workspace_bytes(int64_t n_ref) { return neighbors_workspace_bytes(n_ref); }

int eags_knn_mean_dist2(const float* points, int32_t n, float* out, void* workspace,
                        size_t workspace_bytes, void* stream) {
    if (n < 0 || (n > 0 && (!points || !out || !workspace)) ||
        workspace_bytes < neighbors_workspace_bytes(n))
        return fail(EAGS_ERR_INVALID, "eags_knn_mean_dist2: bad argument or workspace too small");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_knn_mean_dist2(points, n, out, workspace, s), "knn_mean_dist2");
    return 0;
}

int eags_radius_has_neighbor(const float* ref_points, int32_t n_ref, const float* query_points,
                             int32_t n_query, float radius, uint8_t* has_neighbor, void* workspace,
                             size_t workspace_bytes, void* stream) {
    if (n_ref < 0 || n_query < 0 || (n_query > 0 && (!query_points || !has_neighbor)) ||
        (n_ref > 0 && n_query > 0 && (!ref_points || !workspace)) ||
        (n_ref > 0 && workspace_bytes < neighbors_workspace_bytes(n_ref)))
        return fail(EAGS_ERR_INVALID, "eags_radius_has_neighbor: bad argument or workspace too small");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_radius_has_neighbor(ref_points, n_ref, query_points, n_query, radius, has_neighbor,
                                            workspace, s),
                 "radius_has_neighbor");
    return 0;
}

size_t eags_ssim_scratch_bytes(int32_t C, int32_t H, int32_t W) {
    return eags_align256((size_t)3 * C * H * W * sizeof(float)) + 256;   // three partial maps + the fp64 sum
}

int eags_ssim_forward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2, void* scratch,
                      float* out_mean, void* stream) {
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !scratch || !out_mean)
        return fail(EAGS_ERR_INVALID, "eags_ssim_forward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    double* sum = (double*)((char*)scratch + eags_align256((size_t)3 * C * H * W * sizeof(float)));
    CHECK_LAUNCH(launch_ssim_fwd(C, H, W, img1, img2, (float*)scratch, sum, out_mean, s), "ssim_fwd");
    return 0;
}

int eags_ssim_backward(int32_t C, int32_t H, int32_t W, const float* img1, const float* img2,
                       const void* scratch, const float* upstream, float* grad_img1, void* stream) {
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !scratch || !upstream || !grad_img1)
        return fail(EAGS_ERR_INVALID, "eags_ssim_backward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_ssim_bwd(C, H, W, img1, img2, (const float*)scratch, upstream, grad_img1, s), "ssim_bwd");
    return 0;
}

int eags_gaussian_activations(int32_t P, const float* raw_scaling, const float* raw_rotation,
                              const float* raw_opacity, float* scales, float* rotations, float* opacities,
                              void* stream) {
    if (P <= 0 || !raw_scaling || !raw_rotation || !raw_opacity || !scales || !rotations || !opacities ||
        ((uintptr_t)raw_rotation & 15) || ((uintptr_t)rotations & 15))
        return fail(EAGS_ERR_INVALID, "eags_gaussian_activations: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_activations(P, raw_scaling, raw_rotation, raw_opacity, scales, rotations, opacities, s),
                 "activations");
    return 0;
}

size_t eags_mapping_loss_scratch_bytes(int32_t W, int32_t H) { return mapping_loss_scratch_bytes(W, H); }

static bool mapping_loss_ok(const EagsMappingLoss* in) {
    return in && in->width > 0 && in->height > 0 && in->image && in->depth && in->gt_image && in->gt_depth &&
           (long long)in->width * in->height < (1ll << 30);
}

int eags_mapping_loss_forward(const EagsMappingLoss* in, void* scratch, float* out8, void* stream) {
    if (!mapping_loss_ok(in) || !scratch || !out8)
        return fail(EAGS_ERR_INVALID, "eags_mapping_loss_forward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_mapping_loss_fwd(in->width, in->height, in->lambda_dssim, in->image, in->depth, in->gt_image,
                                         in->gt_depth, in->mask, in->valid_depth_mask, scratch, out8, s), "mapping_loss_fwd");
    return 0;
}

int eags_mapping_loss_backward(const EagsMappingLoss* in, const void* scratch, const float* out8,
                               const float* up_color, const float* up_depth, float* grad_image,
                               float* grad_depth, void* stream) {
    if (!mapping_loss_ok(in) || !scratch || !out8 || !up_color || !up_depth || !grad_image || !grad_depth)
        return fail(EAGS_ERR_INVALID, "eags_mapping_loss_backward: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_mapping_loss_bwd(in->width, in->height, in->lambda_dssim, in->image, in->depth, in->gt_image,
                                         in->gt_depth, in->mask, in->valid_depth_mask, scratch, out8, up_color, up_depth,
                                         grad_image, grad_depth, s), "mapping_loss_bwd");
    return 0;
}

int eags_isotropic_loss(int32_t P, const float* scaling, float* raw_grad, void* sum_scratch8, float* out_mean,
                        void* stream) {
    if (P <= 0 || !scaling || !raw_grad || !sum_scratch8 || !out_mean)
        return fail(EAGS_ERR_INVALID, "eags_isotropic_loss: bad argument");
    cudaStream_t s = (cudaStream_t)stream;
    CHECK_LAUNCH(launch_isotropic(P, scaling, raw_grad, (double*)sum_scratch8, out_mean, s), "isotropic_loss");
    return 0;
}

int eags_state_view(const char* what, int32_t P, int32_t W, int32_t H, int64_t capacity,
                    int32_t with_keys, const void* geom_state, const void* image_state,
                    const void* binning_state, const void** ptr, int64_t* count, int32_t* elem_bytes) {
    if (!what || !ptr || !count || !elem_bytes) return fail(EAGS_ERR_INVALID, "eags_state_view: null argument");
    GeomLayout GL = geom_layout(P);
    ImageLayout IL = image_layout(W, H);
    BinningLayout BL = binning_layout(capacity, with_keys);
    const char* gb = (const char*)geom_state;
    const char* ib = (const char*)image_state;
    const char* bb = (const char*)binning_state;
    struct Row { const char* name; const char* base; size_t off; int64_t n; int32_t eb; };
    const Row rows[] = {
        {"splat_rec", gb, GL.splat_rec, (int64_t)P * 3, 16},
        {"tiles_touched", gb, GL.tiles_touched, P, 4},
        {"clamped", gb, GL.clamped, P, 1},
        {"ctrl", ib, IL.ctrl, 1, 64},
        {"tile_count", ib, IL.tile_count, (int64_t)IL.G * EAGS_CTR_STRIDE, 4},
        {"ranges", ib, IL.ranges, IL.G, 8},
        {"tile_max_contrib", ib, IL.tile_max, IL.G, 4},
        {"final_T", ib, IL.final_T, IL.N, 4},
        {"n_contrib", ib, IL.n_contrib, IL.N, 4},
        {"point_list", bb, BL.point_list, capacity, 4},
        {"point_list_keys", bb, BL.keys, with_keys ? capacity : 0, 8},
        {"records", bb, BL.records, capacity * 3, 16},
    };
    for (const Row& r : rows) {
        if (!strcmp(what, r.name)) {
            if (!r.base) return fail(EAGS_ERR_INVALID, "eags_state_view: that state buffer was not given");
            *ptr = r.base + r.off;
            *count = r.n;
            *elem_bytes = r.eb;
            return 0;
        }
    }
    return fail(EAGS_ERR_INVALID, "eags_state_view: unknown array name");
}

}  // extern "C"
