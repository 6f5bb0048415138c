// oracle/ref_knn_shim.cu -- extern "C" door to the UNMODIFIED reference simple_knn (test infrastructure only).
// Linked with /root/reference/submodules/simple-knn/simple_knn.cu compiled where it lies (oracle/Makefile `ref`).
#include <cuda_runtime.h>

#include "simple_knn.h"

extern "C" int ref_knn(int P, const float* points_dev, float* mean_dist2_dev) {
  SimpleKNN::knn(P, (float3*)points_dev, mean_dist2_dev);
  return (int)cudaDeviceSynchronize();
}
