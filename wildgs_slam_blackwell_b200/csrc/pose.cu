// Fused pose step (include/wgs_pose.h): Adam on the pose/exposure deltas + update_pose (SE3_exp, compose,
// convergence test, zeroing) + camera matrices, one single-warp launch.  Replaces
// pose_optimizer.step() + update_pose() of the pose-refinement loop (src/mapper.py:905-907,
// src/utils/pose_utils.py:17-100).  Pure latency: ~300 flops.
#include "../../include/wgs_pose.h"
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {

struct PoseParams {
    WgsPoseStepArgs a;
    WgsPoseAdam adam;
    int has_adam;
};

__device__ __forceinline__ void mat3_mul(const float* A, const float* B, float* C) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}

__global__ void pose_step_kernel(const __grid_constant__ PoseParams p) {
    if (threadIdx.x != 0) return;
    const WgsPoseStepArgs& a = p.a;
    float theta[3] = {a.cam_rot_delta[0], a.cam_rot_delta[1], a.cam_rot_delta[2]};
    float rho[3] = {a.cam_trans_delta[0], a.cam_trans_delta[1], a.cam_trans_delta[2]};
    if (p.has_adam) {
        const WgsPoseAdam& o = p.adam;
        const int step = *o.step + 1;
        *o.step = step;
        const double bc1 = 1.0 - pow(o.beta1, (double)step), bc2 = 1.0 - pow(o.beta2, (double)step);
        const float bc2s = (float)sqrt(bc2), w1 = (float)(1.0 - o.beta1), w2 = (float)(1.0 - o.beta2);
        const float b2 = (float)o.beta2, eps = (float)o.eps;
        float* prm[8] = {&theta[0], &theta[1], &theta[2], &rho[0], &rho[1], &rho[2], a.exposure_a, a.exposure_b};
        const float* grd[8] = {o.grad_rot, o.grad_rot + 1, o.grad_rot + 2, o.grad_trans, o.grad_trans + 1,
                               o.grad_trans + 2, o.grad_exposure_a, o.grad_exposure_b};
        const double lr[8] = {o.lr_rot, o.lr_rot, o.lr_rot, o.lr_trans, o.lr_trans, o.lr_trans, o.lr_exposure_a,
                             o.lr_exposure_b};
        for (int k = 0; k < 8; k++) {
            if (!prm[k] || !grd[k]) continue;
            const float g = *grd[k];
            float m = o.exp_avg[k], v = o.exp_avg_sq[k];
            m = m + w1 * (g - m);
            v = v * b2 + w2 * g * g;
            o.exp_avg[k] = m;
            o.exp_avg_sq[k] = v;
            const float denom = sqrtf(v) / bc2s + eps;
            *prm[k] = *prm[k] - (float)(lr[k] / bc1) * (m / denom);
        }
    }
    // ---- update_pose (pose_utils.py:81-100) ----
    const float W[9] = {0.f, -theta[2], theta[1], theta[2], 0.f, -theta[0], -theta[1], theta[0], 0.f};   // skew_sym_mat
    float W2[9];
    mat3_mul(W, W, W2);
    const float angle = sqrtf(theta[0] * theta[0] + theta[1] * theta[1] + theta[2] * theta[2]);
    float ca, cb, va, vb;   // R = I + ca W + cb W2 ; V = I + va W + vb W2
    if (angle < 1e-5f) {
        ca = 1.f; cb = 0.5f; va = 0.5f; vb = 1.0f / 6.0f;
    } else {
        const float s = sinf(angle), c = cosf(angle), a2 = angle * angle;
        ca = s / angle; cb = (1.f - c) / a2;
        va = (1.0f - c) / a2; vb = (angle - s) / (a2 * angle);
    }
    float Re[9], Ve[9];
    for (int i = 0; i < 9; i++) {
        const float I = (i % 4 == 0) ? 1.f : 0.f;
        Re[i] = I + ca * W[i] + cb * W2[i];
        Ve[i] = I + va * W[i] + vb * W2[i];
    }
    float te[3];
    for (int i = 0; i < 3; i++) te[i] = Ve[3 * i] * rho[0] + Ve[3 * i + 1] * rho[1] + Ve[3 * i + 2] * rho[2];
    float R0[9], T0[3], Rn[9], Tn[3];
    for (int i = 0; i < 9; i++) R0[i] = a.R[i];
    for (int i = 0; i < 3; i++) T0[i] = a.T[i];
    mat3_mul(Re, R0, Rn);
    for (int i = 0; i < 3; i++) Tn[i] = Re[3 * i] * T0[0] + Re[3 * i + 1] * T0[1] + Re[3 * i + 2] * T0[2] + te[i];
    for (int i = 0; i < 9; i++) a.R[i] = Rn[i];
    for (int i = 0; i < 3; i++) a.T[i] = Tn[i];
    const float tn = sqrtf(rho[0] * rho[0] + rho[1] * rho[1] + rho[2] * rho[2] + theta[0] * theta[0] +
                           theta[1] * theta[1] + theta[2] * theta[2]);
    *a.converged = tn < a.converged_threshold ? 1 : 0;
    for (int i = 0; i < 3; i++) { a.cam_rot_delta[i] = 0.f; a.cam_trans_delta[i] = 0.f; }
    // ---- camera matrices of the new pose (camera.cu) ----
    if (a.viewmatrix) {
        float V[16];
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) V[4 * r + c] = c < 3 ? (r < 3 ? Rn[c * 3 + r] : Tn[c]) : (r < 3 ? 0.f : 1.f);
        for (int i = 0; i < 16; i++) a.viewmatrix[i] = V[i];
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) {
                float s = 0.f;
                for (int k = 0; k < 4; k++) s = fmaf(V[4 * r + k], a.projection_matrix[4 * k + c], s);
                a.projmatrix[4 * r + c] = s;
            }
        for (int i = 0; i < 3; i++) a.campos[i] = -(Rn[i] * Tn[0] + Rn[3 + i] * Tn[1] + Rn[6 + i] * Tn[2]);
    }
}

}  // namespace wgs

extern "C" {
void wgs_set_error_message(const char* msg);

int wgs_pose_step(const WgsPoseStepArgs* a, void* cuda_stream) {
    using namespace wgs;
    if (!a || !a->cam_rot_delta || !a->cam_trans_delta || !a->R || !a->T || !a->converged) {
        wgs_set_error_message("wgs_pose_step: null required pointer");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    if (a->viewmatrix && (!a->projection_matrix || !a->projmatrix || !a->campos)) {
        wgs_set_error_message("wgs_pose_step: viewmatrix given without projection_matrix / projmatrix / campos");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    PoseParams p;
    p.a = *a;
    p.has_adam = a->adam != nullptr;
    if (p.has_adam) {
        p.adam = *a->adam;
        if (!p.adam.exp_avg || !p.adam.exp_avg_sq || !p.adam.step || !p.adam.grad_rot || !p.adam.grad_trans) {
            wgs_set_error_message("wgs_pose_step: incomplete Adam state");
            return WGS_ERR_INVALID_ARGUMENT;
        }
    }
    p.a.adam = nullptr;
    pose_step_kernel<<<1, 32, 0, (cudaStream_t)cuda_stream>>>(p);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}
}
