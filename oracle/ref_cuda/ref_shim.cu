// TEST INFRASTRUCTURE — not part of the product path.
//
// C-ABI shim over the UNMODIFIED reference rasterizer
// (/root/reference/thirdparty/diff-gaussian-rasterization-w-pose/cuda_rasterizer,
//  CudaRasterizer::Rasterizer::{forward,backward,markVisible}, rasterizer.h:20-91).
// The reference sources are compiled where they lie (see build_ref.py); this file only
// adds extern "C" entry points with raw device pointers so the parity tests can run the
// reference's own CUDA kernels side by side with ours and read back its internal
// state (GeometryState / BinningState / ImageState, rasterizer_impl.h:21-73) for
// stage-wise bit-exact comparison.  Nothing here is linked into libwgs_rasterizer.so.
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
// Absolute paths are injected by build_ref.py: adding the reference directory to the
// include path would shadow <math.h> with the reference's own cuda_rasterizer/math.h.
#include REF_RASTERIZER_H
#include REF_RASTERIZER_IMPL_H

namespace {

struct Buf {
    char* ptr = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + (n >> 2) + 256;
            if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
            cap = want;
        }
        return ptr;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = 0; }
};

struct RefCtx {
    Buf geom, binning, img;
    int P = 0, R = 0, W = 0, H = 0;
};

}  // namespace

extern "C" {

void* ref_create() { return new RefCtx(); }

void ref_destroy(void* h) {
    RefCtx* c = (RefCtx*)h;
    c->geom.release(); c->binning.release(); c->img.release();
    delete c;
}

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-122) minus the torch allocations:
// outputs are caller-allocated and must be pre-filled with zeros like torch::full does.
int ref_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                const float* projmatrix, const float* campos, float tanfovx, float tanfovy,
                int prefiltered, float* out_color, float* out_depth, float* out_opacity,
                int* radii, int* n_touched, int debug) {
    RefCtx* c = (RefCtx*)h;
    std::function<char*(size_t)> g = [c](size_t n) { return c->geom.get(n); };
    std::function<char*(size_t)> b = [c](size_t n) { return c->binning.get(n); };
    std::function<char*(size_t)> i = [c](size_t n) { return c->img.get(n); };
    int R = 0;
    try {
        R = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, bg, W, H, means3D, shs,
                                                colors_precomp, opacities, scales, scale_modifier,
                                                rotations, cov3D_precomp, viewmatrix, projmatrix,
                                                campos, tanfovx, tanfovy, prefiltered != 0,
                                                out_color, out_depth, out_opacity, radii, n_touched,
                                                debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_forward: %s\n", e.what());
        return -1;
    }
    c->P = P; c->R = R; c->W = W; c->H = H;
    return R;
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:124-211); the eleven gradient
// buffers are caller-allocated and must be zero-filled (torch::zeros in the reference).
int ref_backward(void* h, int P, int D, int M, int R, const float* bg, int W, int H,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* projmatrix_raw, const float* campos, float tanfovx, float tanfovy,
                 const int* radii, const float* dL_dpix, const float* dL_dpix_depth,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                 float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscale, float* dL_drot, float* dL_dtau, int debug) {
    RefCtx* c = (RefCtx*)h;
    try {
        CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp,
                                             scales, scale_modifier, rotations, cov3D_precomp,
                                             viewmatrix, projmatrix, projmatrix_raw, campos,
                                             tanfovx, tanfovy, radii, c->geom.ptr, c->binning.ptr,
                                             c->img.ptr, dL_dpix, dL_dpix_depth, dL_dmean2D,
                                             dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth,
                                             dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
                                             dL_dtau, debug != 0);
    } catch (const std::exception& e) {
        fprintf(stderr, "ref_backward: %s\n", e.what());
        return -1;
    }
    return 0;
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

// Device-to-device copies of the reference's internal state after ref_forward.
// Any destination may be null.  Layout knowledge comes from the reference's own
// fromChunk() (rasterizer_impl.cu:155-194), not from a re-derivation.
int ref_export_state(void* h,
                     float* depths, float* means2D /*2P*/, float* cov3D /*6P*/,
                     float* conic_opacity /*4P*/, float* rgb /*3P*/, uint32_t* tiles_touched,
                     uint32_t* point_offsets, unsigned char* clamped /*3P*/,
                     uint64_t* point_list_keys /*R*/, uint32_t* point_list /*R*/,
                     uint64_t* point_list_keys_unsorted, uint32_t* point_list_unsorted,
                     uint32_t* ranges /*2*tiles*/, uint32_t* n_contrib /*W*H*/,
                     float* accum_alpha /*W*H*/) {
    RefCtx* c = (RefCtx*)h;
    const size_t P = c->P, R = c->R, N = (size_t)c->W * c->H;
    const size_t tiles = (size_t)((c->W + 15) / 16) * ((c->H + 15) / 16);
    char* gp = c->geom.ptr;
    CudaRasterizer::GeometryState gs = CudaRasterizer::GeometryState::fromChunk(gp, P);
    char* ip = c->img.ptr;
    CudaRasterizer::ImageState is = CudaRasterizer::ImageState::fromChunk(ip, N);
    auto cp = [](void* dst, const void* src, size_t n) {
        if (dst && n) cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice);
    };
    cp(depths, gs.depths, P * 4);
    cp(means2D, gs.means2D, P * 8);
    cp(cov3D, gs.cov3D, P * 24);
    cp(conic_opacity, gs.conic_opacity, P * 16);
    cp(rgb, gs.rgb, P * 12);
    cp(tiles_touched, gs.tiles_touched, P * 4);
    cp(point_offsets, gs.point_offsets, P * 4);
    cp(clamped, gs.clamped, P * 3);
    if (R > 0) {
        char* bp = c->binning.ptr;
        CudaRasterizer::BinningState bs = CudaRasterizer::BinningState::fromChunk(bp, R);
        cp(point_list_keys, bs.point_list_keys, R * 8);
        cp(point_list, bs.point_list, R * 4);
        cp(point_list_keys_unsorted, bs.point_list_keys_unsorted, R * 8);
        cp(point_list_unsorted, bs.point_list_unsorted, R * 4);
    }
    cp(ranges, is.ranges, tiles * 8);
    cp(n_contrib, is.n_contrib, N * 4);
    cp(accum_alpha, is.accum_alpha, N * 4);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : -2;
}

}  // extern "C"
