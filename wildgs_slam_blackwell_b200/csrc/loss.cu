// Fused photometric loss + cotangent producer (include/wgs_loss.h).  Replaces the torch chains of
// get_loss_tracking_rgb (src/utils/slam_utils.py:55-87) and get_loss_mapping_rgbd (:116-143,
// ssim_loss off) together with their autograd backward: one pass over the pixels reads the
// rendered and ground-truth images once, writes dL/dcolor and dL/ddepth in the layout
// render_bwd consumes, and reduces the loss and the exposure gradients deterministically.
// HBM-bound: 28..44 B read + 16 B written per pixel.
#include "../../include/wgs_loss.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

constexpr int LOSS_THREADS = 256;
constexpr int LOSS_PIX_PER_THREAD = 4;  // one float4 per plane and thread

struct LossParams {
    WgsLossArgs a;
    size_t N;
    int blocks;
    double* partials;    // [blocks][5]: sum l_rgb, sum l_depth, dL/da, dL/db (unscaled by 1/(3N), 1/N), sum w
    unsigned int* ticket;
};

__device__ __forceinline__ float sgn(float x) { return (float)(x > 0.f) - (float)(x < 0.f); }

__device__ __forceinline__ float4 ld4(const float* p, size_t i, size_t N, float fill) {
    if (i + 3 < N && ((reinterpret_cast<uintptr_t>(p + i) & 15u) == 0)) return __ldg(reinterpret_cast<const float4*>(p + i));
    float4 v = make_float4(fill, fill, fill, fill);
    if (i < N) v.x = __ldg(p + i);
    if (i + 1 < N) v.y = __ldg(p + i + 1);
    if (i + 2 < N) v.z = __ldg(p + i + 2);
    if (i + 3 < N) v.w = __ldg(p + i + 3);
    return v;
}
__device__ __forceinline__ void st4(float* p, size_t i, size_t N, float4 v) {
    if (i + 3 < N && ((reinterpret_cast<uintptr_t>(p + i) & 15u) == 0)) { *reinterpret_cast<float4*>(p + i) = v; return; }
    if (i < N) p[i] = v.x;
    if (i + 1 < N) p[i + 1] = v.y;
    if (i + 2 < N) p[i + 2] = v.z;
    if (i + 3 < N) p[i + 3] = v.w;
}

__global__ void __launch_bounds__(LOSS_THREADS) photometric_loss_kernel(const __grid_constant__ LossParams p) {
    pdl_wait();     // first statement: see common.cuh (the loss sits between render_fwd and render_bwd)
    pdl_trigger();
    const WgsLossArgs& a = p.a;
    const size_t N = p.N;
    const size_t i0 = ((size_t)blockIdx.x * LOSS_THREADS + threadIdx.x) * LOSS_PIX_PER_THREAD;
    float ea = 1.f, eb = 0.f;
    if (a.exposure_a) { ea = expf(__ldg(a.exposure_a)); eb = __ldg(a.exposure_b); }
    const float kc = a.color_weight / (3.0f * (float)N), kd = a.depth_weight / (float)N;
    float s_rgb = 0.f, s_dep = 0.f, s_da = 0.f, s_db = 0.f, s_w = 0.f;
    if (i0 < N) {
        float c[3][4], g[3][4], op[4], pm[4], un[4];
        {
            float4 v;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                v = ld4(a.color + ch * N, i0, N, 0.f); c[ch][0] = v.x; c[ch][1] = v.y; c[ch][2] = v.z; c[ch][3] = v.w;
                v = ld4(a.gt_color + ch * N, i0, N, 0.f); g[ch][0] = v.x; g[ch][1] = v.y; g[ch][2] = v.z; g[ch][3] = v.w;
            }
            v = a.opacity ? ld4(a.opacity, i0, N, 0.f) : make_float4(1.f, 1.f, 1.f, 1.f);
            op[0] = v.x; op[1] = v.y; op[2] = v.z; op[3] = v.w;
            v = a.pixel_mask ? ld4(a.pixel_mask, i0, N, 0.f) : make_float4(1.f, 1.f, 1.f, 1.f);
            pm[0] = v.x; pm[1] = v.y; pm[2] = v.z; pm[3] = v.w;
            if (a.uncertainty && a.patch_w > 0) {
                // patch-resolution network output -> per-pixel uncertainty (mapping_utils.py:270-277):
                // bilinear(clip(u, 0.1) + 1e-3) (torch's align_corners=False formulas), then (r - 0.1) * rate + 0.1
                const float sy = (float)a.patch_h / (float)a.height, sx = (float)a.patch_w / (float)a.width;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const size_t i = min(i0 + j, N - 1);
                    const int py = (int)(i / a.width), px = (int)(i - (size_t)py * a.width);
                    const float fy = fmaxf(sy * ((float)py + 0.5f) - 0.5f, 0.f), fx = fmaxf(sx * ((float)px + 0.5f) - 0.5f, 0.f);
                    const int y0 = min((int)fy, a.patch_h - 1), y1 = min(y0 + 1, a.patch_h - 1);
                    const int x0 = min((int)fx, a.patch_w - 1), x1 = min(x0 + 1, a.patch_w - 1);
                    const float ly = fminf(fmaxf(fy - (float)y0, 0.f), 1.f), lx = fminf(fmaxf(fx - (float)x0, 0.f), 1.f);
                    auto u = [&](int y, int x) { return fmaxf(__ldg(a.uncertainty + y * a.patch_w + x), 0.1f) + 1e-3f; };
                    const float r = (1.f - ly) * ((1.f - lx) * u(y0, x0) + lx * u(y0, x1)) + ly * ((1.f - lx) * u(y1, x0) + lx * u(y1, x1));
                    un[j] = (r - 0.1f) * a.uncertainty_rate + 0.1f;
                }
            } else {
                v = a.uncertainty ? ld4(a.uncertainty, i0, N, 1.f) : make_float4(1.f, 1.f, 1.f, 1.f);
                un[0] = v.x; un[1] = v.y; un[2] = v.z; un[3] = v.w;
            }
        }
        float dc[3][4], wj[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const bool in = i0 + j < N;
            // rgb_pixel_mask = gt.sum(dim=0) > threshold, times grad_mask (slam_utils.py:72-73,132)
            const float m = ((g[0][j] + g[1][j]) + g[2][j] > a.rgb_boundary_threshold) ? pm[j] : 0.f;
            float w = 1.f;
            if (a.uncertainty) {  // slam_utils.py:81-83
                w = 0.5f / (un[j] * un[j]);
                w = w < 0.1f ? 0.f : w;
            }
            wj[j] = w;
            if (in) s_w += w;
            const float scale = in ? op[j] * w : 0.f;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                const float iab = ea * c[ch][j] + eb;
                const float r = iab * m - g[ch][j] * m;
                s_rgb += scale * fabsf(r);
                const float gab = kc * scale * sgn(r) * m;  // dL/d image_ab
                dc[ch][j] = ea * gab;
                s_da += gab * (ea * c[ch][j]);
                s_db += gab;
            }
        }
#pragma unroll
        for (int ch = 0; ch < 3; ch++)
            st4(a.dL_dcolor + ch * N, i0, N, make_float4(dc[ch][0], dc[ch][1], dc[ch][2], dc[ch][3]));
        if (a.dL_ddepth) {
            float dd[4] = {0.f, 0.f, 0.f, 0.f};
            if (a.depth_weight != 0.f) {
                const float4 dv = ld4(a.depth, i0, N, 0.f), gv = ld4(a.gt_depth, i0, N, 0.f);
                const float d[4] = {dv.x, dv.y, dv.z, dv.w}, gd[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    // slam_utils.py:140-141; mapping_utils.py:262-266 adds the upper bound
                    const float dm = (gd[j] > a.depth_min && !(a.depth_max > 0.f && !(gd[j] < a.depth_max))) ? 1.f : 0.f;
                    const float r = d[j] * dm - gd[j] * dm;
                    // slam_utils.py:238-242: weighted only where the reference depth is in front of the rendered one (+ margin)
                    const float wd = (a.depth_weighted && gd[j] < d[j] + a.depth_margin) ? wj[j] : 1.f;
                    s_dep += wd * fabsf(r);
                    dd[j] = kd * wd * sgn(r) * dm;
                }
            }
            st4(a.dL_ddepth, i0, N, make_float4(dd[0], dd[1], dd[2], dd[3]));
        }
    }
    // block reduction (fixed order) -> partials; the last block sums the partials in block order
    __shared__ double red[5][LOSS_THREADS / 32];
    __shared__ bool last;
    double v[5] = {(double)s_rgb, (double)s_dep, (double)s_da, (double)s_db, (double)s_w};
#pragma unroll
    for (int k = 0; k < 5; k++)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < 5; k++) red[k][warp] = v[k];
    __syncthreads();
    if (threadIdx.x < 5) {
        double s = 0.0;
        for (int wq = 0; wq < LOSS_THREADS / 32; wq++) s += red[threadIdx.x][wq];
        p.partials[(size_t)blockIdx.x * 5 + threadIdx.x] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(p.ticket, 1u) == (unsigned)(p.blocks - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    if (warp < 5) {
        double s = 0.0;
        for (int b = lane; b < p.blocks; b += 32) s += __ldcg(p.partials + (size_t)b * 5 + warp);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) red[warp][0] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double l = (double)a.color_weight * red[0][0] / (3.0 * (double)N) + (double)a.depth_weight * red[1][0] / (double)N;
        *a.loss = (float)l;
        if (a.dL_dexposure) { a.dL_dexposure[0] = (float)red[2][0]; a.dL_dexposure[1] = (float)red[3][0]; }
        if (a.weight_mean) *a.weight_mean = (float)(red[4][0] / (double)N);
        *p.ticket = 0u;  // leave the workspace ready for the next call
    }
}

}  // namespace wgs

using namespace wgs;

extern "C" {

static int loss_blocks(size_t N) {
    return (int)((N + (size_t)LOSS_THREADS * LOSS_PIX_PER_THREAD - 1) / ((size_t)LOSS_THREADS * LOSS_PIX_PER_THREAD));
}

size_t wgs_loss_workspace_bytes(int32_t width, int32_t height) {
    if (width <= 0 || height <= 0) return 0;
    return align_up(256 + (size_t)loss_blocks((size_t)width * height) * 5 * sizeof(double));
}

void wgs_set_error_message(const char* msg);  // api.cu

int wgs_photometric_loss(const WgsLossArgs* a, void* cuda_stream) {
    if (!a || a->width <= 0 || a->height <= 0 || !a->color || !a->gt_color || !a->loss || !a->dL_dcolor ||
        !a->workspace || ((uintptr_t)a->workspace & 255u)) {
        wgs_set_error_message("wgs_photometric_loss: null/misaligned required pointer or bad size");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->patch_w < 0 || a->patch_h < 0 || (a->patch_w > 0) != (a->patch_h > 0) || (a->patch_w > 0 && !a->uncertainty)) {
        wgs_set_error_message("wgs_photometric_loss: patch_w / patch_h describe the uncertainty map: both > 0 or both 0");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->depth_weighted && !a->uncertainty) {
        wgs_set_error_message("wgs_photometric_loss: depth_weighted needs the uncertainty map");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->depth_weight != 0.f && (!a->depth || !a->gt_depth || !a->dL_ddepth)) {
        wgs_set_error_message("wgs_photometric_loss: depth_weight != 0 needs depth, gt_depth and dL_ddepth");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->exposure_a && (!a->exposure_b || !a->dL_dexposure)) {
        wgs_set_error_message("wgs_photometric_loss: exposure_a given without exposure_b / dL_dexposure");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    LossParams p;
    p.a = *a;
    p.N = (size_t)a->width * a->height;
    p.blocks = loss_blocks(p.N);
    p.ticket = (unsigned int*)a->workspace;
    p.partials = (double*)((char*)a->workspace + 256);
    wgs::launch_chained(photometric_loss_kernel, p.blocks, LOSS_THREADS, 0, (cudaStream_t)cuda_stream, p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}

}  // extern "C"
