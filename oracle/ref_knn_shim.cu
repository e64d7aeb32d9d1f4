// TEST INFRASTRUCTURE ONLY. C-ABI shim around the unmodified reference simple-knn
// (/root/reference/submodules/simple-knn/simple_knn.h:15-19, simple_knn.cu:185-221).
#include <cuda_runtime.h>
#include "simple_knn.h"
extern "C" void ref_knn(int P, float* points, float* mean_dists) {
  SimpleKNN::knn(P, reinterpret_cast<float3*>(points), mean_dists);
}
