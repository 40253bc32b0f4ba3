// oracle/ref_knn_shim.cu -- TEST INFRASTRUCTURE ONLY.  A C entry point over the reference's own
// SimpleKNN::knn (/root/reference/submodules/simple-knn/simple_knn.h:15-19, simple_knn.cu:186-221), which is
// compiled from where it lies into oracle/_ref/libsimple_knn_ref.so (oracle/Makefile, target `ref`).
// Mirrors what the reference's torch binding does (spatial.cu:15-26): device points in, device means out.
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" __attribute__((visibility("default"))) int ref_dist_cuda2(int P, const float* points_dev, float* mean_dists_dev) {
    cudaMemset(mean_dists_dev, 0, sizeof(float) * (size_t)P);
    SimpleKNN::knn(P, (float3*)points_dev, mean_dists_dev);
    return (int)cudaDeviceSynchronize();
}
