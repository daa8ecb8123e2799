// Fused multi-tensor Adam (include/wgs_adam.h): one launch for all parameter tensors of the Gaussian
// model.  HBM-bound: 16 B read + 12 B written per element.
#include "../../include/wgs_adam.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_EPT = 4;

struct AdamParams {
    WgsAdamTensor t[WGS_ADAM_MAX_TENSORS];
    uint32_t first_block[WGS_ADAM_MAX_TENSORS + 1];
    int count;
    float beta2, w1, w2, eps;   // w = 1 - beta, formed in double on the host
    double beta1d, beta2d;
};

__global__ void __launch_bounds__(ADAM_THREADS) adam_kernel(const __grid_constant__ AdamParams p) {
    int e = 0;
    while (e + 1 < p.count && blockIdx.x >= p.first_block[e + 1]) e++;
    const WgsAdamTensor& t = p.t[e];
    const size_t base = ((size_t)(blockIdx.x - p.first_block[e]) * ADAM_THREADS + threadIdx.x) * ADAM_EPT;
    const float w1 = p.w1, w2 = p.w2;
    float step_size = t.step_size, bc2s = t.bias_correction2_sqrt;
    if (t.step_dev) {   // capturable: corrections from the device-side step count (one thread per block, double like torch)
        __shared__ float s_corr[2];
        if (threadIdx.x == 0) {
            const double step = (double)(*t.step_dev) + 1.0;
            const float lr = t.lr_dev ? *t.lr_dev : t.step_size;
            s_corr[0] = (float)((double)lr / (1.0 - pow(p.beta1d, step)));
            s_corr[1] = (float)sqrt(1.0 - pow(p.beta2d, step));
        }
        __syncthreads();
        step_size = s_corr[0]; bc2s = s_corr[1];
    }
    auto update = [&](float g, float& m, float& v, float& w) {
        m = m + w1 * (g - m);                       // exp_avg.lerp_(grad, 1 - beta1)
        v = v * p.beta2 + w2 * g * g;               // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
        const float denom = sqrtf(v) / bc2s + p.eps;
        w = w - step_size * (m / denom);            // param.addcdiv_(exp_avg, denom, value=-step_size)
    };
    const bool vec = base + ADAM_EPT <= t.n &&
                     (((uintptr_t)t.param | (uintptr_t)t.grad | (uintptr_t)t.exp_avg | (uintptr_t)t.exp_avg_sq) & 15u) == 0;
    if (vec) {   // one 16-byte access per array
        const float4 g = __ldg(reinterpret_cast<const float4*>(t.grad + base));
        float4 m = *reinterpret_cast<float4*>(t.exp_avg + base), v = *reinterpret_cast<float4*>(t.exp_avg_sq + base);
        float4 w = *reinterpret_cast<float4*>(t.param + base);
        update(g.x, m.x, v.x, w.x); update(g.y, m.y, v.y, w.y); update(g.z, m.z, v.z, w.z); update(g.w, m.w, v.w, w.w);
        *reinterpret_cast<float4*>(t.exp_avg + base) = m;
        *reinterpret_cast<float4*>(t.exp_avg_sq + base) = v;
        *reinterpret_cast<float4*>(t.param + base) = w;
        return;
    }
#pragma unroll
    for (int k = 0; k < ADAM_EPT; k++) {
        const size_t i = base + k;
        if (i < t.n) {
            float m = t.exp_avg[i], v = t.exp_avg_sq[i], w = t.param[i];
            update(__ldg(t.grad + i), m, v, w);
            t.exp_avg[i] = m;
            t.exp_avg_sq[i] = v;
            t.param[i] = w;
        }
    }
}

// after the update: advance the device-side step counters of the capturable tensors
__global__ void adam_advance_kernel(const __grid_constant__ AdamParams p) {
    const int e = threadIdx.x;
    if (e < p.count && p.t[e].step_dev) {
        bool first = true;   // descriptors may share one counter (segments of one arena): advance it once
        for (int j = 0; j < e; j++) first = first && p.t[j].step_dev != p.t[e].step_dev;
        if (first) *p.t[e].step_dev += 1.f;
    }
}

}  // namespace wgs

extern "C" {
void wgs_set_error_message(const char* msg);

int wgs_adam_step(const WgsAdamTensor* tensors, int32_t count, double beta1, double beta2, double eps,
                  void* cuda_stream) {
    using namespace wgs;
    if (count < 0 || count > WGS_ADAM_MAX_TENSORS || (count > 0 && !tensors)) {
        wgs_set_error_message("wgs_adam_step: bad tensor count or null table");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    AdamParams p;
    p.count = 0; p.beta2 = (float)beta2; p.w1 = (float)(1.0 - beta1); p.w2 = (float)(1.0 - beta2); p.eps = (float)eps;
    p.beta1d = beta1; p.beta2d = beta2;
    bool capturable = false;
    uint32_t blocks = 0;
    for (int i = 0; i < count; i++) {
        const WgsAdamTensor& t = tensors[i];
        if (t.n == 0) continue;
        if (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq) {
            wgs_set_error_message("wgs_adam_step: null tensor pointer");
            return WGS_ERR_INVALID_ARGUMENT;
        }
        capturable |= t.step_dev != nullptr;
        p.t[p.count] = t;
        p.first_block[p.count] = blocks;
        blocks += (uint32_t)((t.n + (size_t)ADAM_THREADS * ADAM_EPT - 1) / ((size_t)ADAM_THREADS * ADAM_EPT));
        p.count++;
    }
    p.first_block[p.count] = blocks;
    if (blocks == 0) return WGS_OK;
    adam_kernel<<<blocks, ADAM_THREADS, 0, (cudaStream_t)cuda_stream>>>(p);
    if (capturable) adam_advance_kernel<<<1, 32, 0, (cudaStream_t)cuda_stream>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}
}
