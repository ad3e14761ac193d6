// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI wrapper around the UNMODIFIED reference simple-knn (compiled in place
// from /root/reference/submodules/simple-knn/simple_knn.cu by oracle/build_ref.sh
// into oracle/_ref/libref_knn.so). Replaces the torch glue KNN/spatial.cu:15-26.
#include <cuda_runtime.h>
#include "simple_knn.h"

extern "C" int refknn_dist2(int P, const float* points, float* mean_dists) {
  if (P <= 0) return 0;
  SimpleKNN::knn(P, (float3*)points, mean_dists);
  return (int)cudaGetLastError();
}
