// Camera matrices for one view in one launch.  Replaces the ~15 tiny torch kernels (two
// torch.linalg.inv, bmm, inverse, transposes) behind the reference Camera's per-access properties
// world_view_transform / full_proj_transform / camera_center (src/utils/camera_utils.py:137-151,
// thirdparty/gaussian_splatting/utils/graphics_utils.py:33-46) that render() reads before
// every rasterizer call (gaussian_renderer/__init__.py:58-72).  SURVEY.md §8(f) row 1.
//   W2C = [R t; 0 1]           (getWorld2View2 with translate = 0, scale = 1 is inv(inv(Rt)) = Rt)
//   viewmatrix = W2C^T,  projmatrix = viewmatrix @ projection_matrix,  campos = -R^T t
#include "../../include/wgs_rasterizer.h"
#include "common.cuh"

namespace wgs {
__global__ void camera_matrices_kernel(const float* __restrict__ R, const float* __restrict__ T,
                                       const float* __restrict__ Pt, float* __restrict__ view,
                                       float* __restrict__ full, float* __restrict__ campos) {
    __shared__ float V[16];
    const int i = threadIdx.x;  // 16 threads
    const int r = i >> 2, c = i & 3;
    // V = W2C^T : V[r][c] = W2C[c][r]
    float v;
    if (c < 3) v = r < 3 ? R[c * 3 + r] : T[c];
    else v = r < 3 ? 0.f : 1.f;
    V[i] = v;
    view[i] = v;
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) s = fmaf(V[r * 4 + k], Pt[k * 4 + c], s);
    full[i] = s;
    if (i < 3) campos[i] = -(R[0 * 3 + i] * T[0] + R[1 * 3 + i] * T[1] + R[2 * 3 + i] * T[2]);
}
}  // namespace wgs

extern "C" {
void wgs_set_error_message(const char* msg);

int wgs_camera_matrices(const float* R, const float* T, const float* projection_matrix, float* viewmatrix,
                        float* projmatrix, float* campos, void* cuda_stream) {
    if (!R || !T || !projection_matrix || !viewmatrix || !projmatrix || !campos) {
        wgs_set_error_message("wgs_camera_matrices: null pointer");
        return WGS_ERR_INVALID_ARGUMENT;
    }
    wgs::camera_matrices_kernel<<<1, 16, 0, (cudaStream_t)cuda_stream>>>(R, T, projection_matrix, viewmatrix,
                                                                         projmatrix, campos);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        wgs_set_error_message(cudaGetErrorString(e));
        return WGS_ERR_CUDA;
    }
    return WGS_OK;
}
}
