// C-ABI of libwgs_rasterizer.so (see include/wgs_rasterizer.h): argument checks, workspace
// carving and stage sequencing.  Replaces RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA
// (rasterize_points.cu:35-211) and CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (rasterizer_impl.cu:141-455).
#include <atomic>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/wgs_rasterizer.h"
#include "common.cuh"
#include "kernels.h"

using namespace wgs;

struct WgsContext {
    int device = 0;
    int num_sms = 0;
    unsigned int* host_R = nullptr;  // pinned + mapped
    unsigned int* dev_R = nullptr;   // device alias of host_R
    cudaEvent_t ev_R = nullptr;
    unsigned int* sticky = nullptr;  // device: {forwards that overflowed their capacity, largest R among them}
    // stage timing (wgs_stage_timing): fwd boundaries 0..5, bwd boundaries 6..8
    bool async = false;              // wgs_set_async: wgs_forward never waits for R
    long long last_capacity = 0;     // capacity of the most recent forward (for wgs_last_num_rendered)
    bool timing = false;
    bool fwd_timed = false, bwd_timed = false;
    cudaEvent_t ev_stage[9] = {};
};

namespace wgs {
// -1: not yet read from the environment (WGS_PDL=0 turns the chained launches off for the process)
static std::atomic<int> g_chained{-1};
bool pdl_enabled() {
    int v = g_chained.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("WGS_PDL");
        v = (e && e[0] == '0') ? 0 : 1;
        int expected = -1;
        g_chained.compare_exchange_strong(expected, v);
        v = g_chained.load(std::memory_order_relaxed);
    }
    return v != 0;
}
}  // namespace wgs

namespace {
thread_local char g_err[512] = "";
void set_err(const char* what, cudaError_t e) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, e == cudaSuccess ? "" : cudaGetErrorString(e));
}
#define CK(call)                                  \
    do {                                          \
        cudaError_t e_ = (call);                  \
        if (e_ != cudaSuccess) {                  \
            set_err(#call, e_);                   \
            return WGS_ERR_CUDA;                  \
        }                                         \
    } while (0)
// CHECK_CUDA(A, debug) of the reference (auxiliary.h:166-173)
#define MARK(ctx, i, stream)                                       \
    do {                                                           \
        if ((ctx)->timing && !capturing) CK(cudaEventRecord((ctx)->ev_stage[i], stream)); \
    } while (0)
#define STAGE(debug, stream, name)                                 \
    do {                                                           \
        cudaError_t e_ = cudaGetLastError();                       \
        if (e_ == cudaSuccess && (debug)) e_ = cudaStreamSynchronize(stream); \
        if (e_ != cudaSuccess) {                                   \
            set_err(name, e_);                                     \
            return WGS_ERR_CUDA;                                   \
        }                                                          \
    } while (0)

bool aligned256(const void* p) { return ((uintptr_t)p & 255u) == 0; }
// tile-row strip of a call: (0,0) means the whole image; returns false if out of range
bool strip_rows(int begin, int end, int gy, int* r0, int* r1) {
    if (begin == 0 && end == 0) { *r0 = 0; *r1 = gy; return true; }
    if (begin < 0 || end <= begin || end > gy) return false;
    *r0 = begin; *r1 = end;
    return true;
}
}  // namespace

extern "C" {

int wgs_abi_version(void) { return WGS_ABI_VERSION; }
const char* wgs_last_error(void) { return g_err; }
// used by the other translation units of the library (loss.cu, ...)
void wgs_set_error_message(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg ? msg : ""); }

WgsContext* wgs_context_create(int device) {
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        set_err("wgs_context_create: no such CUDA device", cudaGetLastError());
        return nullptr;
    }
    if (prop.major != 10) {
        snprintf(g_err, sizeof(g_err), "wgs_context_create: device is sm_%d%d, this library is sm_100a only",
                 prop.major, prop.minor);
        return nullptr;
    }
    WgsContext* c = new WgsContext();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    cudaError_t e = cudaHostAlloc((void**)&c->host_R, 64, cudaHostAllocMapped);
    if (e == cudaSuccess) e = cudaHostGetDevicePointer((void**)&c->dev_R, c->host_R, 0);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_R, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->sticky, 2 * sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMemset(c->sticky, 0, 2 * sizeof(unsigned int));
    if (e == cudaSuccess) e = (cudaError_t)configure_binning();
    if (e == cudaSuccess) e = (cudaError_t)configure_render_bwd();
    if (e == cudaSuccess) e = (cudaError_t)configure_preprocess_bwd();
    if (e == cudaSuccess) e = (cudaError_t)configure_preprocess_fwd();
    if (e != cudaSuccess) {
        set_err("wgs_context_create", e);
        delete c;
        return nullptr;
    }
    return c;
}

void wgs_context_destroy(WgsContext* c) {
    if (!c) return;
    if (c->ev_R) cudaEventDestroy(c->ev_R);
    for (cudaEvent_t e : c->ev_stage)
        if (e) cudaEventDestroy(e);
    if (c->host_R) cudaFreeHost(c->host_R);
    if (c->sticky) cudaFree(c->sticky);
    delete c;
}

int wgs_workspace_sizes(int32_t P, int32_t width, int32_t height, int64_t capacity, WgsWorkspaceSizes* out) {
    if (!out || P < 0 || width <= 0 || height <= 0 || capacity < 0) {
        set_err("wgs_workspace_sizes: invalid argument", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    const size_t tiles = (size_t)((width + 15) / 16) * ((height + 15) / 16);
    out->geom_bytes = geom_bytes((size_t)(P > 0 ? P : 1));
    out->tile_bytes = tile_bytes(tiles);
    out->binning_bytes = align_up((size_t)(capacity > 0 ? capacity : 1) * 8);
    out->image_bytes = image_bytes((size_t)width * height);
    out->grad_bytes = grad_bytes((size_t)(P > 0 ? P : 1));
    return WGS_OK;
}

int wgs_forward(WgsContext* ctx, const WgsForwardArgs* a, void* cuda_stream, int64_t* num_rendered) {
    if (!ctx || !a || !num_rendered) { set_err("wgs_forward: null argument", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    *num_rendered = 0;
    if (a->struct_size != sizeof(WgsForwardArgs)) {
        snprintf(g_err, sizeof(g_err), "wgs_forward: args->struct_size is %u, this library's WgsForwardArgs has %zu bytes "
                 "(ABI %d): rebuild the binding against include/wgs_rasterizer.h", a->struct_size, sizeof(WgsForwardArgs),
                 WGS_ABI_VERSION);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    // Inside a CUDA-graph capture (or in async mode) the host must not wait on the device: R is not read
    // back here, *num_rendered = -1, and the caller checks wgs_last_num_rendered() after a synchronisation.
    cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
    if (ctx && cudaStreamIsCapturing(s, &cap_status) != cudaSuccess) cap_status = cudaStreamCaptureStatusNone;
    const bool capturing = cap_status == cudaStreamCaptureStatusActive;
    if (a->P < 0 || a->width <= 0 || a->height <= 0 || a->width > 65535 * 16 || a->height > 65535 * 16 ||
        a->sh_degree < 0 || a->sh_degree > 3) {
        set_err("wgs_forward: bad P / image size / sh_degree", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (!a->background || !a->viewmatrix || !a->projmatrix || !a->out_color || !a->out_depth || !a->out_opacity ||
        !a->tile || !a->image || !aligned256(a->tile) || !aligned256(a->image)) {
        set_err("wgs_forward: null/misaligned required pointer", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->P > 0) {
        if (!a->means3D || !a->opacities || !a->radii || !a->n_touched || !a->geom || !a->binning || !a->campos ||
            !aligned256(a->geom) || !aligned256(a->binning)) {
            set_err("wgs_forward: null/misaligned required pointer", cudaSuccess);
            return WGS_ERR_INVALID_ARGUMENT;
        }
        if (!a->sh && !a->colors_precomp) {
            set_err("wgs_forward: provide sh or colors_precomp", cudaSuccess);
            return WGS_ERR_NO_COLOR_SOURCE;
        }
        if (a->sh && !a->colors_precomp && a->M < (a->sh_degree + 1) * (a->sh_degree + 1)) {
            set_err("wgs_forward: sh has fewer coefficients than sh_degree needs", cudaSuccess);
            return WGS_ERR_INVALID_ARGUMENT;
        }
        if (!a->cov3D_precomp && (!a->scales || !a->rotations)) {
            set_err("wgs_forward: provide scales+rotations or cov3D_precomp", cudaSuccess);
            return WGS_ERR_INVALID_ARGUMENT;
        }
    }
    const int gx = (a->width + 15) / 16, gy = (a->height + 15) / 16, tiles = gx * gy;
    int row0, row1;
    if (!strip_rows(a->tile_row_begin, a->tile_row_end, gy, &row0, &row1)) {
        set_err("wgs_forward: tile_row_begin/end out of range", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    const size_t Pn = (size_t)(a->P > 0 ? a->P : 1);
    GeomView geom = geom_view(a->geom, Pn);
    TileView tile = tile_view(a->tile, (size_t)tiles);
    ImageView image = image_view(a->image, (size_t)a->width * a->height);

    CK(cudaMemsetAsync(tile.count, 0, tile_clear_bytes((size_t)tiles), s));

    PreFwdParams pp;
    pp.P = a->P; pp.D = a->sh_degree; pp.M = a->M; pp.W = a->width; pp.H = a->height; pp.gx = gx; pp.gy = gy;
    pp.prefiltered = a->prefiltered; pp.copies = hist_copies((size_t)tiles); pp.row0 = row0; pp.row1 = row1;
    pp.tanfovx = a->tanfovx; pp.tanfovy = a->tanfovy;
    pp.focal_y = a->height / (2.0f * a->tanfovy);  // rasterizer_impl.cu:225-226
    pp.focal_x = a->width / (2.0f * a->tanfovx);
    pp.scale_modifier = a->scale_modifier;
    pp.means3D = a->means3D; pp.sh = a->colors_precomp ? nullptr : a->sh; pp.colors_precomp = a->colors_precomp;
    pp.opacities = a->opacities; pp.scales = a->scales; pp.rotations = a->rotations; pp.cov3D_precomp = a->cov3D_precomp;
    pp.viewmatrix = a->viewmatrix; pp.projmatrix = a->projmatrix; pp.campos = a->campos;
    pp.radii = a->radii; pp.n_touched = a->n_touched; pp.geom = geom; pp.tile = tile;
    pp.act = a->activations; pp.sh_rest = a->colors_precomp ? nullptr : a->sh_rest;
    MARK(ctx, 0, s);
    launch_preprocess_fwd(pp, s);
    STAGE(a->debug, s, "preprocess_fwd");
    MARK(ctx, 1, s);

    BinParams bp;
    bp.P = a->P; bp.gx = gx; bp.gy = gy; bp.tiles = tiles; bp.copies = hist_copies((size_t)tiles); bp.capacity = a->capacity;
    bp.geom = geom; bp.tile = tile; bp.bucket = (unsigned long long*)a->binning; bp.host_R = ctx->dev_R;
    const bool no_wait = capturing || ctx->async;
    bp.sticky = no_wait ? ctx->sticky : nullptr;  // a blocking forward reports its overflow itself (and is retried)
    launch_tile_scan(bp, s);
    if (!no_wait) CK(cudaEventRecord(ctx->ev_R, s));
    ctx->last_capacity = a->capacity;
    STAGE(a->debug, s, "tile_scan");
    MARK(ctx, 2, s);
    launch_scatter(bp, s);
    MARK(ctx, 3, s);
    launch_tile_sorts(bp, ctx->num_sms, s);
    STAGE(a->debug, s, "binning");
    MARK(ctx, 4, s);

    RenderFwdParams rp;
    rp.W = a->width; rp.H = a->height; rp.gx = gx; rp.gy = gy; rp.bg = a->background;
    rp.tile0 = row0 * gx; rp.ntiles = (row1 - row0) * gx;
    rp.geom = geom; rp.tile = tile; rp.image = image; rp.bucket = bp.bucket;
    rp.out_color = a->out_color; rp.out_depth = a->out_depth; rp.out_opacity = a->out_opacity;
    rp.n_touched = a->n_touched;
    launch_render_fwd(rp, s);
    STAGE(a->debug, s, "render_fwd");
    MARK(ctx, 5, s);
    ctx->fwd_timed = ctx->timing && !capturing;
    if (no_wait) {
        *num_rendered = -1;
        return WGS_OK;
    }

    // Everything is enqueued; now wait for R only (the GPU keeps running the later stages).
    CK(cudaEventSynchronize(ctx->ev_R));
    const int64_t R = (int64_t)*(volatile unsigned int*)ctx->host_R;
    *num_rendered = R;
    if (R > a->capacity) {
        snprintf(g_err, sizeof(g_err), "wgs_forward: %lld instances exceed capacity %lld", (long long)R,
                 (long long)a->capacity);
        return WGS_ERR_CAPACITY;
    }
    return WGS_OK;
}

int wgs_backward(WgsContext* ctx, const WgsBackwardArgs* a, void* cuda_stream) {
    if (!ctx || !a) { set_err("wgs_backward: null argument", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    if (a->struct_size != sizeof(WgsBackwardArgs)) {
        snprintf(g_err, sizeof(g_err), "wgs_backward: args->struct_size is %u, this library's WgsBackwardArgs has %zu bytes "
                 "(ABI %d): rebuild the binding against include/wgs_rasterizer.h", a->struct_size, sizeof(WgsBackwardArgs),
                 WGS_ABI_VERSION);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap_status) != cudaSuccess) cap_status = cudaStreamCaptureStatusNone;
    const bool capturing = cap_status == cudaStreamCaptureStatusActive;
    if (a->P < 0 || a->width <= 0 || a->height <= 0 || a->accumulate < 0 || a->accumulate > 2) {
        set_err("wgs_backward: bad P / image size / accumulate mode", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->P == 0) {
        if (a->dL_dtau) CK(cudaMemsetAsync(a->dL_dtau, 0, 6 * sizeof(float), s));
        return WGS_OK;
    }
    if (!a->background || !a->means3D || !a->viewmatrix || !a->projmatrix || !a->projmatrix_raw || !a->campos ||
        !a->radii || !a->dL_dout_color || !a->dL_dout_depth || !a->geom || !a->tile || !a->binning || !a->image ||
        !a->grad || !aligned256(a->grad)) {
        set_err("wgs_backward: null/misaligned required pointer", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (!a->cov3D_precomp && (!a->scales || !a->rotations)) {
        set_err("wgs_backward: provide scales+rotations or cov3D_precomp", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    int row0, row1;
    if (!strip_rows(a->tile_row_begin, a->tile_row_end, (a->height + 15) / 16, &row0, &row1)) {
        set_err("wgs_backward: tile_row_begin/end out of range", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    const bool use_sh = a->sh && !a->colors_precomp;
    if (use_sh && !a->dL_dsh) { set_err("wgs_backward: dL_dsh is required on the SH path", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    const int gx = (a->width + 15) / 16, gy = (a->height + 15) / 16, tiles = gx * gy;
    GeomView geom = geom_view(const_cast<void*>(a->geom), (size_t)a->P);
    TileView tile = tile_view(const_cast<void*>(a->tile), (size_t)tiles);
    ImageView image = image_view(const_cast<void*>(a->image), (size_t)a->width * a->height);
    GradView grad = grad_view(a->grad, (size_t)a->P);

    RenderBwdParams rb;
    rb.W = a->width; rb.H = a->height; rb.gx = gx; rb.gy = gy; rb.bg = a->background;
    rb.tile0 = row0 * gx; rb.ntiles = (row1 - row0) * gx;
    rb.geom = geom; rb.tile = tile; rb.image = image; rb.bucket = (const unsigned long long*)a->binning;
    rb.dL_dcolor = a->dL_dout_color; rb.dL_ddepth = a->dL_dout_depth; rb.grad = grad;
    MARK(ctx, 6, s);
    launch_render_bwd(rb, s);
    STAGE(a->debug, s, "render_bwd");
    MARK(ctx, 7, s);

    PreBwdParams pb;
    pb.P = a->P; pb.D = a->sh_degree; pb.M = a->M; pb.W = a->width; pb.H = a->height; pb.accumulate = a->accumulate;
    pb.strip = (row0 != 0 || row1 != gy) ? 1 : 0;
    pb.tanfovx = a->tanfovx; pb.tanfovy = a->tanfovy;
    pb.focal_y = a->height / (2.0f * a->tanfovy);
    pb.focal_x = a->width / (2.0f * a->tanfovx);
    pb.scale_modifier = a->scale_modifier;
    pb.means3D = a->means3D; pb.sh = use_sh ? a->sh : nullptr; pb.colors_precomp = a->colors_precomp;
    pb.scales = a->scales; pb.rotations = a->rotations; pb.cov3D_precomp = a->cov3D_precomp;
    pb.viewmatrix = a->viewmatrix; pb.projmatrix = a->projmatrix; pb.projmatrix_raw = a->projmatrix_raw;
    pb.campos = a->campos; pb.radii = a->radii; pb.geom = geom; pb.grad = grad;
    pb.dL_dmeans3D = a->dL_dmeans3D; pb.dL_dmeans2D = a->dL_dmeans2D; pb.dL_dsh = a->dL_dsh;
    pb.dL_dcolors = a->dL_dcolors; pb.dL_dopacity = a->dL_dopacity; pb.dL_dscales = a->dL_dscales;
    pb.dL_drotations = a->dL_drotations; pb.dL_dcov3D = a->dL_dcov3D; pb.dL_dtau = a->dL_dtau;
    pb.act = a->activations; pb.sh_rest = use_sh ? a->sh_rest : nullptr; pb.dL_dsh_rest = a->dL_dsh_rest;
    pb.stat_grad_accum = a->stat_grad_accum; pb.stat_denom = a->stat_denom; pb.stat_max_radii = a->stat_max_radii;
    pb.misc = tile.misc;
    if ((pb.stat_grad_accum || pb.stat_denom || pb.stat_max_radii) && (row0 != 0 || row1 != gy)) {
        set_err("wgs_backward: densification statistics need a whole-image backward (the gradient norm is per view)", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (pb.sh_rest && pb.dL_dsh && !pb.dL_dsh_rest && a->M > 1) {
        set_err("wgs_backward: sh_rest given without dL_dsh_rest", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    launch_preprocess_bwd(pb, s);
    STAGE(a->debug, s, "preprocess_bwd");
    MARK(ctx, 8, s);
    ctx->bwd_timed = ctx->timing && !capturing;
    return WGS_OK;
}

int wgs_set_chained_launch(int enable) {
    const int before = wgs::pdl_enabled() ? 1 : 0;
    wgs::g_chained.store(enable ? 1 : 0, std::memory_order_relaxed);
    return before;
}

int wgs_set_async(WgsContext* ctx, int enable) {
    if (!ctx) { set_err("wgs_set_async: null context", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    ctx->async = enable != 0;
    return WGS_OK;
}

int wgs_last_num_rendered(WgsContext* ctx, int64_t* num_rendered, int64_t* capacity) {
    if (!ctx || !num_rendered) { set_err("wgs_last_num_rendered: null argument", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    *num_rendered = (int64_t)*(volatile unsigned int*)ctx->host_R;
    if (capacity) *capacity = ctx->last_capacity;
    return *num_rendered > ctx->last_capacity ? WGS_ERR_CAPACITY : WGS_OK;
}

int wgs_overflow_status(WgsContext* ctx, int64_t* overflowed_forwards, int64_t* max_num_rendered, int clear) {
    if (!ctx) { set_err("wgs_overflow_status: null context", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    unsigned int h[2] = {0u, 0u};
    CK(cudaMemcpy(h, ctx->sticky, sizeof(h), cudaMemcpyDeviceToHost));  // the caller has synchronised its streams
    if (clear && h[0]) CK(cudaMemset(ctx->sticky, 0, sizeof(h)));
    if (overflowed_forwards) *overflowed_forwards = (int64_t)h[0];
    if (max_num_rendered) *max_num_rendered = (int64_t)h[1];
    if (h[0]) {
        snprintf(g_err, sizeof(g_err), "%u forward(s) exceeded their instance capacity (largest needed %u)", h[0], h[1]);
        return WGS_ERR_CAPACITY;
    }
    return WGS_OK;
}

int wgs_count_pairs(const WgsForwardArgs* a, uint64_t* counts, void* cuda_stream) {
    if (!a || !counts || a->struct_size != sizeof(WgsForwardArgs) || !a->tile || !a->image || (a->P > 0 && (!a->geom || !a->binning))) {
        set_err("wgs_count_pairs: invalid argument", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    const int gx = (a->width + 15) / 16, gy = (a->height + 15) / 16;
    int row0, row1;
    if (!strip_rows(a->tile_row_begin, a->tile_row_end, gy, &row0, &row1)) {
        set_err("wgs_count_pairs: tile_row_begin/end out of range", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    RenderFwdParams rp = {};
    rp.W = a->width; rp.H = a->height; rp.gx = gx; rp.gy = gy; rp.tile0 = row0 * gx; rp.ntiles = (row1 - row0) * gx;
    rp.geom = geom_view(a->geom, (size_t)(a->P > 0 ? a->P : 1));
    rp.tile = tile_view(a->tile, (size_t)gx * gy);
    rp.image = image_view(a->image, (size_t)a->width * a->height);
    rp.bucket = (const unsigned long long*)a->binning;
    launch_count_pairs(rp, (unsigned long long*)counts, (cudaStream_t)cuda_stream);
    STAGE(0, (cudaStream_t)cuda_stream, "count_pairs");
    return WGS_OK;
}

int wgs_stage_timing(WgsContext* ctx, int enable) {
    if (!ctx) { set_err("wgs_stage_timing: null context", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    if (enable) {
        for (cudaEvent_t& e : ctx->ev_stage)
            if (!e) CK(cudaEventCreate(&e));
    }
    ctx->timing = enable != 0;
    ctx->fwd_timed = ctx->bwd_timed = false;
    return WGS_OK;
}

int wgs_stage_times(WgsContext* ctx, float* ms) {
    if (!ctx || !ms) { set_err("wgs_stage_times: null argument", cudaSuccess); return WGS_ERR_INVALID_ARGUMENT; }
    for (int i = 0; i < WGS_NUM_STAGES; i++) ms[i] = 0.f;
    if (ctx->fwd_timed) {
        CK(cudaEventSynchronize(ctx->ev_stage[5]));
        for (int i = 0; i < 5; i++) CK(cudaEventElapsedTime(ms + i, ctx->ev_stage[i], ctx->ev_stage[i + 1]));
    }
    if (ctx->bwd_timed) {
        CK(cudaEventSynchronize(ctx->ev_stage[8]));
        CK(cudaEventElapsedTime(ms + 5, ctx->ev_stage[6], ctx->ev_stage[7]));
        CK(cudaEventElapsedTime(ms + 6, ctx->ev_stage[7], ctx->ev_stage[8]));
    }
    ctx->fwd_timed = ctx->bwd_timed = false;
    return WGS_OK;
}

int wgs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* cuda_stream) {
    (void)projmatrix;  // the reference's test only uses the view-space depth (auxiliary.h:152)
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) {
        set_err("wgs_mark_visible: invalid argument", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)cuda_stream);
    STAGE(0, (cudaStream_t)cuda_stream, "mark_visible");
    return WGS_OK;
}

int wgs_export_binning(const WgsForwardArgs* a, int64_t num_rendered, uint64_t* point_list_keys,
                       uint32_t* point_list, uint32_t* ranges, uint32_t* n_contrib, float* final_T,
                       void* cuda_stream) {
    if (!a || a->struct_size != sizeof(WgsForwardArgs) || !a->tile || !a->image || num_rendered > a->capacity) {
        set_err("wgs_export_binning: invalid argument", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    const int gx = (a->width + 15) / 16, gy = (a->height + 15) / 16, tiles = gx * gy;
    BinParams bp;
    bp.P = a->P; bp.gx = gx; bp.gy = gy; bp.tiles = tiles; bp.copies = hist_copies((size_t)tiles); bp.capacity = a->capacity;
    bp.geom = geom_view(a->geom, (size_t)(a->P > 0 ? a->P : 1));
    bp.tile = tile_view(a->tile, (size_t)tiles);
    bp.bucket = (unsigned long long*)a->binning; bp.host_R = nullptr; bp.sticky = nullptr;
    launch_export_binning(bp, image_view(a->image, (size_t)a->width * a->height), a->width, a->height,
                          (unsigned long long*)point_list_keys, point_list, ranges, n_contrib, final_T,
                          (cudaStream_t)cuda_stream);
    STAGE(0, (cudaStream_t)cuda_stream, "export_binning");
    return WGS_OK;
}

int wgs_export_geometry(const WgsForwardArgs* a, float* depths, float* means2D, float* conic_opacity, float* rgb,
                        uint8_t* clamped, void* cuda_stream) {
    if (!a || a->struct_size != sizeof(WgsForwardArgs) || (a->P > 0 && (!a->geom || !a->radii))) {
        set_err("wgs_export_geometry: invalid argument", cudaSuccess);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    launch_export_geometry(a->P, geom_view(a->geom, (size_t)(a->P > 0 ? a->P : 1)), a->radii, depths, means2D,
                           conic_opacity, rgb, clamped, (cudaStream_t)cuda_stream);
    STAGE(0, (cudaStream_t)cuda_stream, "export_geometry");
    return WGS_OK;
}

}  // extern "C"
