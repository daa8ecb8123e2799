// TEST INFRASTRUCTURE: raw-pointer C entry into the UNMODIFIED reference simple-knn
// (thirdparty/simple-knn/simple_knn.cu:182-222, SimpleKNN::knn), compiled from the sources where they lie.
#include <cuda_runtime.h>
#include REF_SIMPLE_KNN_H
extern "C" int ref_dist_cuda2(int P, const float* points, float* mean_dists) {
    SimpleKNN::knn(P, (float3*)points, mean_dists);
    return (int)cudaDeviceSynchronize();
}
