// Fused uncertainty loss of the mapper (include/wgs_uncertainty.h): the patch-resolution half of
// compute_mapping_loss_components (src/utils/dyn_uncertainty/mapping_utils.py:270-321).  Kernel 1 resamples the
// four full-resolution maps to the patch grid (each patch pixel gathers 4 bilinear / 16 bicubic taps, the
// per-pixel clipped SSIM and depth losses formed on the fly -- no full-resolution intermediate is written);
// kernel 2 (one block) takes the k x k reflect-padded median, evaluates the loss and its gradient w.r.t. the
// uncertainty map, and reduces in a fixed order.  A few thousand elements: latency-, not bandwidth-bound; the point
// is 2 launches instead of ~50.
#include "../../include/wgs_uncertainty.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

struct UncParams {
    WgsUncertaintyLossArgs a;
    int n;
    float* s_ssim; float* s_op; float* s_dl; float* s_dr;   // workspace, n floats each
};

// torch area_pixel_compute_source_index + guard_index_and_lambda (UpSample.cuh), align_corners = false
__device__ __forceinline__ void lin_src(float scale, int dst, int size, int& i0, int& i1, float& l1) {
    const float src = fmaxf(scale * ((float)dst + 0.5f) - 0.5f, 0.f);
    i0 = min((int)src, size - 1);
    i1 = min(i0 + 1, size - 1);
    l1 = fminf(fmaxf(src - (float)i0, 0.f), 1.f);
}
// get_cubic_upsampling_coefficients, A = -0.75
__device__ __forceinline__ void cubic_coeff(float t, float c[4]) {
    const float A = -0.75f;
    auto c1 = [&](float x) { return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f; };
    auto c2 = [&](float x) { return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A; };
    c[0] = c2(t + 1.f); c[1] = c1(t); c[2] = c1(1.f - t); c[3] = c2((1.f - t) + 1.f);
}

__global__ void __launch_bounds__(128) unc_resample_kernel(const __grid_constant__ UncParams p) {
    const WgsUncertaintyLossArgs& a = p.a;
    const int q = blockIdx.x * 128 + threadIdx.x;
    if (q >= p.n) return;
    const int i = q / a.patch_w, j = q - i * a.patch_w, W = a.width, H = a.height;
    const float sy = (float)H / (float)a.patch_h, sx = (float)W / (float)a.patch_w;
    {   // bilinear: clipped SSIM loss and opacity
        int y0, y1, x0, x1;
        float ly, lx;
        lin_src(sy, i, H, y0, y1, ly);
        lin_src(sx, j, W, x0, x1, lx);
        auto tap = [&](int y, int x, float& s, float& o) {
            const size_t k = (size_t)y * W + x;
            o = __ldg(a.opacity + k);
            s = fminf(o * a.ssim_weight * (1.f - __ldg(a.luminance + k)) * (1.f - __ldg(a.structure + k)) *
                          (1.f - __ldg(a.contrast + k)), 5.f);
        };
        float s00, s01, s10, s11, o00, o01, o10, o11;
        tap(y0, x0, s00, o00); tap(y0, x1, s01, o01); tap(y1, x0, s10, o10); tap(y1, x1, s11, o11);
        const float hy = 1.f - ly, hx = 1.f - lx;
        p.s_ssim[q] = hy * (hx * s00 + lx * s01) + ly * (hx * s10 + lx * s11);
        p.s_op[q] = hy * (hx * o00 + lx * o01) + ly * (hx * o10 + lx * o11);
    }
    {   // bicubic: clipped depth loss and reference depth
        const float fy = sy * ((float)i + 0.5f) - 0.5f, fx = sx * ((float)j + 0.5f) - 0.5f;
        const int iy = (int)floorf(fy), ix = (int)floorf(fx);
        float cy[4], cx[4];
        cubic_coeff(fy - (float)iy, cy);
        cubic_coeff(fx - (float)ix, cx);
        float accl = 0.f, accr = 0.f;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int y = min(max(iy - 1 + r, 0), H - 1);
            float rowl = 0.f, rowr = 0.f;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const int x = min(max(ix - 1 + c, 0), W - 1);
                const size_t k = (size_t)y * W + x;
                const float ref = __ldg(a.ref_depth + k), d = __ldg(a.depth + k);
                const float dm = (ref > 0.01f && ref < a.depth_threshold) ? 1.f : 0.f;
                rowl += cx[c] * fminf(fabsf(d * dm - ref * dm), 5.f);
                rowr += cx[c] * ref;
            }
            accl += cy[r] * rowl;
            accr += cy[r] * rowr;
        }
        p.s_dl[q] = accl;
        p.s_dr[q] = accr;
    }
}

__device__ __forceinline__ int reflect(int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

__global__ void __launch_bounds__(1024) unc_loss_kernel(const __grid_constant__ UncParams p) {
    const WgsUncertaintyLossArgs& a = p.a;
    __shared__ double red[32];
    const int k = a.median, r = k / 2, kk = k * k;
    double sum = 0.0;
    for (int q = threadIdx.x; q < p.n; q += 1024) {
        const int i = q / a.patch_w, j = q - i * a.patch_w;
        float v[49];
        int m = 0;
        for (int dy = -r; dy <= r; dy++)
            for (int dx = -r; dx <= r; dx++) {   // insertion sort of the window
                const float x = p.s_ssim[reflect(i + dy, a.patch_h) * a.patch_w + reflect(j + dx, a.patch_w)];
                int t = m++;
                while (t > 0 && v[t - 1] > x) { v[t] = v[t - 1]; t--; }
                v[t] = x;
            }
        const float f = v[(kk - 1) / 2];
        const float u0 = __ldg(a.uncertainty + q);
        const float pr = fmaxf(u0, 0.1f) + 1e-3f;
        const float dl = p.s_dr[q] > a.depth_threshold ? 0.f : p.s_dl[q];
        const bool on = !(p.s_op[q] < a.opacity_th);
        const float ip2 = 1.f / (pr * pr);
        const float u = f * ip2 + 0.5f * logf(pr) + a.depth_mult * dl * ip2;
        if (on) sum += (double)u;
        if (a.dL_duncertainty) {
            const float g = (-2.f * f * ip2 / pr + 0.5f / pr - 2.f * a.depth_mult * dl * ip2 / pr) * (a.weight / (float)p.n);
            a.dL_duncertainty[q] = (on && u0 >= 0.1f) ? g : 0.f;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 32; w++) s += red[w];
        *a.loss = (float)((double)a.weight * s / (double)p.n);
    }
}

}  // namespace wgs

using namespace wgs;
extern "C" {
void wgs_set_error_message(const char* msg);

size_t wgs_uncertainty_workspace_bytes(int32_t patch_w, int32_t patch_h) {
    if (patch_w <= 0 || patch_h <= 0) return 256;
    return align_up((size_t)4 * patch_w * patch_h * sizeof(float));
}

int wgs_uncertainty_loss(const WgsUncertaintyLossArgs* a, void* cuda_stream) {
    if (!a || a->width <= 0 || a->height <= 0 || a->patch_w <= 0 || a->patch_h <= 0 || a->median < 1 || a->median > 7 ||
        !(a->median & 1) || a->median / 2 >= a->patch_w || a->median / 2 >= a->patch_h || !a->uncertainty ||
        !a->luminance || !a->contrast || !a->structure || !a->opacity || !a->depth || !a->ref_depth || !a->loss ||
        !a->workspace || ((uintptr_t)a->workspace & 255u)) {
        wgs_set_error_message("wgs_uncertainty_loss: bad size, median (odd, <= 7, smaller than the patch grid) or null/misaligned pointer");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    UncParams p;
    p.a = *a;
    p.n = a->patch_w * a->patch_h;
    float* ws = (float*)a->workspace;
    p.s_ssim = ws; p.s_op = ws + p.n; p.s_dl = ws + 2 * (size_t)p.n; p.s_dr = ws + 3 * (size_t)p.n;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    unc_resample_kernel<<<(p.n + 127) / 128, 128, 0, s>>>(p);
    unc_loss_kernel<<<1, 1024, 0, s>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}
}
