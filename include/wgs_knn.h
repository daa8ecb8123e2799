/*
 * wgs_knn.h — C ABI of distCUDA2 (libwgs_rasterizer.so): for every point the mean of the squared
 * distances to its 3 nearest OTHER points.  SURVEY.md §8(f) row 4.
 * Replaces SimpleKNN::knn (thirdparty/simple-knn/simple_knn.cu:182-222) behind
 * simple_knn._C.distCUDA2 (thirdparty/simple-knn/spatial.cu:14-35), which GaussianModel uses to
 * initialise the scales of inserted points (thirdparty/gaussian_splatting/scene/gaussian_model.py, dist2 =
 * clamp_min(distCUDA2(points), 1e-7)).
 * Same result definition: exact 3-NN (the reference's Morton boxes only prune), squared distances
 * evaluated as dx*dx + dy*dy + dz*dz in fp32, mean = (d0 + d1 + d2) / 3.0f, missing neighbours
 * (P < 4) count as FLT_MAX like the reference.  No host synchronisation (the reference blocks twice on
 * D2H copies of the bounding box); all pointers are DEVICE pointers; work goes to the caller's stream.
 */
#ifndef WGS_KNN_H_
#define WGS_KNN_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

size_t wgs_knn_workspace_bytes(int32_t P);
/* points (P,3) float32, mean_dists (P,) float32 out, workspace: wgs_knn_workspace_bytes(P) bytes, 256-B aligned
 * (contents need not be initialised).  Returns WGS_OK or a negative status of wgs_rasterizer.h. */
int wgs_dist_cuda2(int32_t P, const float* points, float* mean_dists, void* workspace, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* WGS_KNN_H_ */
