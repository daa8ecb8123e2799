// Fused SSIM term of the mapping loss (include/wgs_ssim.h).  Replaces loss_utils.ssim / _ssim
// (thirdparty/gaussian_splatting/utils/loss_utils.py:61-101) with its autograd backward, and
// compute_ssim_components (src/utils/dyn_uncertainty/mapping_utils.py:99-204).
//
// One CTA per 16x16 pixel tile of one plane: the (16+2R)^2 input patches of both images go to shared memory
// once (zero outside the image = conv2d's zero padding), the five windowed moments are formed separably
// (horizontal pass into shared memory, vertical pass in registers), and the per-pixel SSIM, its three
// partial derivatives and the block's partial sum leave the CTA.  The backward convolves the three derivative
// maps the same way.  Traffic per pixel and plane: 8 B read + 12 B written forward, 20 B read + 4 B written
// backward (halo re-reads hit L2); the five conv2d + elementwise chain of the reference moves > 150 B.
#include "../../include/wgs_ssim.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

constexpr int ST = 16;                           // tile edge
constexpr int SR = WGS_SSIM_MAX_WINDOW / 2;      // largest radius
constexpr int SN = ST + 2 * SR;                  // 26: patch edge at the largest radius
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;

struct SsimParams {
    int W, H, planes, ppm, R, gx, gy, C;
    float w[WGS_SSIM_MAX_WINDOW];
    const float* x1;
    const float* x2;
    float* maps;       // [3][planes][H][W]
    float* out;        // backward: dL_dimg1
    const float* dmean;
    double* partial;   // [planes][gy*gx]
    float* mean;
    float* lum; float* con; float* str;
    const float* ea; const float* eb;   // exposure (device scalars) or null
    int exp_on;                         // which image is compensated: 1 = x1, 2 = x2, 0 = none
    float* dexp;                        // backward: {dL/da, dL/db} or null
};

// gain/bias: exposure compensation of the pixels inside the image; the zero padding stays zero
__device__ __forceinline__ void load_patch(float (*a)[SN + 1], const float* __restrict__ src, int W, int H, int x0,
                                           int y0, int n, float gain = 1.f, float bias = 0.f) {
    for (int i = threadIdx.x; i < n * n; i += 256) {
        const int r = i / n, c = i - r * n, gx = x0 + c, gy = y0 + r;
        a[r][c] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? fmaf(gain, __ldg(src + (size_t)gy * W + gx), bias) : 0.f;
    }
}

// windowed moments of one plane at this thread's pixel: m = {mu1, mu2, s11, s22, s12}
// RT >= 0: radius known at compile time (tap loops unrolled; 11- and 7-tap windows are the reference's), else p.R
template <int RT>
__device__ __forceinline__ void moments(const SsimParams& p, float (*a)[SN + 1], float (*b)[SN + 1],
                                        float (*h)[SN][ST + 1], const float* __restrict__ x1,
                                        const float* __restrict__ x2, float m[5]) {
    const int R = RT >= 0 ? RT : p.R, n = ST + 2 * R, taps = 2 * R + 1;
    const int x0 = blockIdx.x * ST - R, y0 = blockIdx.y * ST - R;
    __syncthreads();   // previous user of the buffers is done
    const float gain = p.exp_on ? expf(__ldg(p.ea)) : 1.f, bias = p.exp_on ? __ldg(p.eb) : 0.f;
    load_patch(a, x1, p.W, p.H, x0, y0, n, p.exp_on == 1 ? gain : 1.f, p.exp_on == 1 ? bias : 0.f);
    load_patch(b, x2, p.W, p.H, x0, y0, n, p.exp_on == 2 ? gain : 1.f, p.exp_on == 2 ? bias : 0.f);
    __syncthreads();
    for (int i = threadIdx.x; i < n * ST; i += 256) {
        const int r = i / ST, c = i - r * ST;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
        for (int k = 0; k < taps; k++) {
            const float w = p.w[k], u = a[r][c + k], v = b[r][c + k];
            s0 = fmaf(w, u, s0); s1 = fmaf(w, v, s1); s2 = fmaf(w, u * u, s2); s3 = fmaf(w, v * v, s3);
            s4 = fmaf(w, u * v, s4);
        }
        h[0][r][c] = s0; h[1][r][c] = s1; h[2][r][c] = s2; h[3][r][c] = s3; h[4][r][c] = s4;
    }
    __syncthreads();
    const int tx = threadIdx.x & (ST - 1), ty = threadIdx.x / ST;
#pragma unroll
    for (int q = 0; q < 5; q++) m[q] = 0.f;
#pragma unroll
    for (int k = 0; k < taps; k++) {
        const float w = p.w[k];
#pragma unroll
        for (int q = 0; q < 5; q++) m[q] = fmaf(w, h[q][ty + k][tx], m[q]);
    }
}

template <int RT>
__global__ void __launch_bounds__(256) ssim_fwd_kernel(const __grid_constant__ SsimParams p) {
    __shared__ float a[SN][SN + 1], b[SN][SN + 1];
    __shared__ float h[5][SN][ST + 1];
    __shared__ float red[8];
    const int plane = blockIdx.z;
    const size_t N = (size_t)p.W * p.H;
    float m[5];
    moments<RT>(p, a, b, h, p.x1 + plane * N, p.x2 + plane * N, m);
    const int px = blockIdx.x * ST + (threadIdx.x & (ST - 1)), py = blockIdx.y * ST + threadIdx.x / ST;
    float val = 0.f;
    if (px < p.W && py < p.H) {
        const float mu1 = m[0], mu2 = m[1];
        const float sig1 = m[2] - mu1 * mu1, sig2 = m[3] - mu2 * mu2, sig12 = m[4] - mu1 * mu2;
        const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * sig12 + SSIM_C2;
        const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = sig1 + sig2 + SSIM_C2;
        const float iB1 = 1.f / B1, iB2 = 1.f / B2;
        val = A1 * A2 * iB1 * iB2;
        if (p.maps) {
            const size_t o = plane * N + (size_t)py * p.W + px, PN = (size_t)p.planes * N;
            // total derivative w.r.t. mu1 (s11, s12 held), then w.r.t. the raw second moments
            p.maps[o] = 2.f * mu2 * (A2 - A1) * iB1 * iB2 + 2.f * mu1 * val * (iB2 - iB1);
            p.maps[PN + o] = -val * iB2;
            p.maps[2 * PN + o] = 2.f * A1 * iB1 * iB2;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = val;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < 8; i++) s += (double)red[i];
        p.partial[(size_t)plane * (p.gx * p.gy) + blockIdx.y * p.gx + blockIdx.x] = s;
    }
}

// one block per mean: fixed-order sum of the group's partials
__global__ void __launch_bounds__(256) ssim_mean_kernel(const __grid_constant__ SsimParams p) {
    __shared__ double red[256];
    const size_t n = (size_t)p.ppm * p.gx * p.gy;
    const double* __restrict__ src = p.partial + blockIdx.x * n;
    double s = 0.0;
    for (size_t i = threadIdx.x; i < n; i += 256) s += src[i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) p.mean[blockIdx.x] = (float)(red[0] / ((double)p.ppm * p.W * p.H));
}

template <int RT>
__global__ void __launch_bounds__(256) ssim_bwd_kernel(const __grid_constant__ SsimParams p) {
    __shared__ float a[3][SN][SN + 1];
    __shared__ float h[3][SN][ST + 1];
    const int plane = blockIdx.z, R = RT >= 0 ? RT : p.R, n = ST + 2 * R, taps = 2 * R + 1;
    const size_t N = (size_t)p.W * p.H, PN = (size_t)p.planes * N;
    const int x0 = blockIdx.x * ST - R, y0 = blockIdx.y * ST - R;
#pragma unroll
    for (int q = 0; q < 3; q++) load_patch(a[q], p.maps + q * PN + plane * N, p.W, p.H, x0, y0, n);
    __syncthreads();
    for (int i = threadIdx.x; i < n * ST; i += 256) {
        const int r = i / ST, c = i - r * ST;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < taps; k++) {
            const float w = p.w[k];
            s0 = fmaf(w, a[0][r][c + k], s0); s1 = fmaf(w, a[1][r][c + k], s1); s2 = fmaf(w, a[2][r][c + k], s2);
        }
        h[0][r][c] = s0; h[1][r][c] = s1; h[2][r][c] = s2;
    }
    __syncthreads();
    const int tx = threadIdx.x & (ST - 1), ty = threadIdx.x / ST;
    const int px = blockIdx.x * ST + tx, py = blockIdx.y * ST + ty;
    float ga = 0.f, gb = 0.f;   // this pixel's share of dL/da, dL/db
    if (px < p.W && py < p.H) {
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < taps; k++) {
            const float w = p.w[k];
            s0 = fmaf(w, h[0][ty + k][tx], s0); s1 = fmaf(w, h[1][ty + k][tx], s1); s2 = fmaf(w, h[2][ty + k][tx], s2);
        }
        const size_t o = plane * N + (size_t)py * p.W + px;
        const float k = __ldg(p.dmean + plane / p.ppm) / ((float)p.ppm * (float)p.W * (float)p.H);
        const float gain = p.exp_on ? expf(__ldg(p.ea)) : 1.f, bias = p.exp_on ? __ldg(p.eb) : 0.f;
        const float x = __ldg(p.x1 + o), xc = fmaf(gain, x, bias);
        const float g = k * (s0 + 2.f * xc * s1 + __ldg(p.x2 + o) * s2);   // dL / d(compensated img1)
        p.out[o] = gain * g;
        ga = g * gain * x; gb = g;
    }
    if (p.dexp) {   // block partials; ssim_exposure_kernel sums them in a fixed order
        __shared__ float red[2][8];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { ga += __shfl_xor_sync(0xffffffffu, ga, o); gb += __shfl_xor_sync(0xffffffffu, gb, o); }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ga; red[1][threadIdx.x >> 5] = gb; }
        __syncthreads();
        if (threadIdx.x < 2) {
            double s = 0.0;
            for (int i = 0; i < 8; i++) s += (double)red[threadIdx.x][i];
            p.partial[2 * ((size_t)plane * (p.gx * p.gy) + blockIdx.y * p.gx + blockIdx.x) + threadIdx.x] = s;
        }
    }
}

__global__ void __launch_bounds__(256) ssim_exposure_kernel(const __grid_constant__ SsimParams p) {
    __shared__ double red[256];
    const size_t n = (size_t)p.planes * p.gx * p.gy;
    double s = 0.0;
    for (size_t i = threadIdx.x; i < n; i += 256) s += p.partial[2 * i + blockIdx.x];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) p.dexp[blockIdx.x] = (float)red[0];
}

// compute_ssim_components (mapping_utils.py:125-204): grid.z = batch item, channels looped inside
template <int RT>
__global__ void __launch_bounds__(256) ssim_components_kernel(const __grid_constant__ SsimParams p) {
    __shared__ float a[SN][SN + 1], b[SN][SN + 1];
    __shared__ float h[5][SN][ST + 1];
    const size_t N = (size_t)p.W * p.H;
    const float eps = 1.1920928955078125e-07f, C3 = 0.5f * SSIM_C2;   // torch.finfo(float32).eps
    float L = 0.f, Cn = 0.f, S = 0.f;
    for (int ch = 0; ch < p.C; ch++) {
        const size_t plane = (size_t)blockIdx.z * p.C + ch;
        float m[5];
        moments<RT>(p, a, b, h, p.x1 + plane * N, p.x2 + plane * N, m);
        const float mu1 = m[0], mu2 = m[1];
        const float v1 = fmaxf(eps, m[2] - mu1 * mu1), v2 = fmaxf(eps, m[3] - mu2 * mu2);
        float s12 = m[4] - mu1 * mu2;
        const float sd = sqrtf(v1) * sqrtf(v2);
        const float sg = (float)(s12 > 0.f) - (float)(s12 < 0.f);
        s12 = sg * fminf(sqrtf(v1 * v2), fabsf(s12));
        L += (2.f * mu1 * mu2 + SSIM_C1) / (mu1 * mu1 + mu2 * mu2 + SSIM_C1);
        Cn += fminf((2.f * sd + SSIM_C2) / (v1 + v2 + SSIM_C2), 0.98f);
        S += fminf((s12 + C3) / (sd + C3), 0.98f);
    }
    const int px = blockIdx.x * ST + (threadIdx.x & (ST - 1)), py = blockIdx.y * ST + threadIdx.x / ST;
    if (px < p.W && py < p.H) {
        const size_t o = blockIdx.z * N + (size_t)py * p.W + px;
        const float ic = 1.f / (float)p.C;
        p.lum[o] = L * ic; p.con[o] = Cn * ic; p.str[o] = S * ic;
    }
}

// the reference builds the window in float32: torch.Tensor([exp(-(x - ws//2)^2 / (2 sigma^2))]) / sum
static bool fill_window(SsimParams& p, int window) {
    if (window < 1 || window > WGS_SSIM_MAX_WINDOW || !(window & 1)) return false;
    float g[WGS_SSIM_MAX_WINDOW], sum = 0.f;
    for (int x = 0; x < window; x++) {
        const int d = x - window / 2;
        g[x] = (float)exp(-(double)(d * d) / (2.0 * 1.5 * 1.5));
        sum += g[x];
    }
    for (int x = 0; x < WGS_SSIM_MAX_WINDOW; x++) p.w[x] = x < window ? g[x] / sum : 0.f;
    p.R = window / 2;
    return true;
}

}  // namespace wgs

using namespace wgs;
extern "C" {
void wgs_set_error_message(const char* msg);

size_t wgs_ssim_workspace_bytes(int32_t width, int32_t height, int32_t planes) {
    if (width <= 0 || height <= 0 || planes <= 0) return 256;
    const size_t blocks = (size_t)((width + ST - 1) / ST) * ((height + ST - 1) / ST) * planes;
    return align_up(2 * blocks * sizeof(double));
}

static int ssim_setup(const WgsSsimArgs* a, SsimParams& p, const char* who) {
    if (!a || a->width <= 0 || a->height <= 0 || a->planes <= 0 || a->planes_per_mean <= 0 ||
        a->planes % a->planes_per_mean || a->planes > 65535 || !a->img1 || !a->img2 || !fill_window(p, a->window)) {
        wgs_set_error_message(who);
        return WGS_ERR_INVALID_ARGUMENT;
    }
    p.W = a->width; p.H = a->height; p.planes = a->planes; p.ppm = a->planes_per_mean;
    p.gx = (p.W + ST - 1) / ST; p.gy = (p.H + ST - 1) / ST; p.C = 0;
    p.x1 = a->img1; p.x2 = a->img2; p.maps = a->maps; p.out = a->dL_dimg1; p.dmean = a->dL_dmean;
    p.partial = (double*)a->workspace; p.mean = a->mean; p.lum = p.con = p.str = nullptr;
    if ((a->exposure_a == nullptr) != (a->exposure_b == nullptr)) {
        wgs_set_error_message("wgs_ssim: exposure_a and exposure_b go together");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    p.ea = a->exposure_a; p.eb = a->exposure_b; p.exp_on = a->exposure_a ? 1 : 0; p.dexp = nullptr;
    return WGS_OK;
}

int wgs_ssim_forward(const WgsSsimArgs* a, void* cuda_stream) {
    SsimParams p;
    int rc = ssim_setup(a, p, "wgs_ssim_forward: bad size, window (odd, <= 11) or null image pointer");
    if (rc != WGS_OK) return rc;
    if (!a->mean || !a->workspace || ((uintptr_t)a->workspace & 255u)) {
        wgs_set_error_message("wgs_ssim_forward: null mean or null/misaligned workspace");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const dim3 grid(p.gx, p.gy, p.planes);
    if (p.R == 5) ssim_fwd_kernel<5><<<grid, 256, 0, s>>>(p);
    else if (p.R == 3) ssim_fwd_kernel<3><<<grid, 256, 0, s>>>(p);
    else ssim_fwd_kernel<-1><<<grid, 256, 0, s>>>(p);
    ssim_mean_kernel<<<p.planes / p.ppm, 256, 0, s>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}

int wgs_ssim_backward(const WgsSsimArgs* a, void* cuda_stream) {
    SsimParams p;
    int rc = ssim_setup(a, p, "wgs_ssim_backward: bad size, window (odd, <= 11) or null image pointer");
    if (rc != WGS_OK) return rc;
    if (!a->maps || !a->dL_dmean || !a->dL_dimg1) {
        wgs_set_error_message("wgs_ssim_backward: maps, dL_dmean and dL_dimg1 are required");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->dL_dexposure) {
        if (!p.exp_on || !a->workspace || ((uintptr_t)a->workspace & 255u)) {
            wgs_set_error_message("wgs_ssim_backward: dL_dexposure needs exposure_a/b and the workspace");
            return WGS_ERR_INVALID_ARGUMENT;
        }
        p.dexp = a->dL_dexposure;
    }
    const dim3 grid(p.gx, p.gy, p.planes);
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (p.R == 5) ssim_bwd_kernel<5><<<grid, 256, 0, s>>>(p);
    else if (p.R == 3) ssim_bwd_kernel<3><<<grid, 256, 0, s>>>(p);
    else ssim_bwd_kernel<-1><<<grid, 256, 0, s>>>(p);
    if (p.dexp) ssim_exposure_kernel<<<2, 256, 0, s>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}

int wgs_ssim_components(int32_t width, int32_t height, int32_t batch, int32_t channels, int32_t window,
                        const float* img1, const float* img2, const float* exposure_a, const float* exposure_b,
                        float* luminance, float* contrast, float* structure, void* cuda_stream) {
    SsimParams p;
    if (width <= 0 || height <= 0 || batch <= 0 || batch > 65535 || channels <= 0 || !img1 || !img2 || !luminance ||
        !contrast || !structure || (exposure_a == nullptr) != (exposure_b == nullptr) || !fill_window(p, window)) {
        wgs_set_error_message("wgs_ssim_components: bad size, window (odd, <= 11) or null pointer");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    p.W = width; p.H = height; p.planes = batch * channels; p.ppm = channels; p.C = channels;
    p.gx = (p.W + ST - 1) / ST; p.gy = (p.H + ST - 1) / ST;
    p.x1 = img1; p.x2 = img2; p.maps = nullptr; p.out = nullptr; p.dmean = nullptr; p.partial = nullptr; p.mean = nullptr;
    p.lum = luminance; p.con = contrast; p.str = structure;
    p.ea = exposure_a; p.eb = exposure_b; p.exp_on = exposure_a ? 2 : 0; p.dexp = nullptr;
    const dim3 grid(p.gx, p.gy, batch);
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (p.R == 5) ssim_components_kernel<5><<<grid, 256, 0, s>>>(p);
    else if (p.R == 3) ssim_components_kernel<3><<<grid, 256, 0, s>>>(p);
    else ssim_components_kernel<-1><<<grid, 256, 0, s>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wgs_set_error_message(cudaGetErrorString(e)); return WGS_ERR_CUDA; }
    return WGS_OK;
}
}
