// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI wrapper around the UNMODIFIED reference cuda_utils kernels (compiled in place from
// /root/reference/soar/threestudio-soar/utils/cuda_utils.cu by oracle/build_ref.sh into
// oracle/_ref/libref_utils.so). Replaces the torch glue TS/utils/ext.cpp:74-107 (contour_padding)
// and :201-255 (gaussian2occgrid): same zero-initialised outputs, same argument order.
// All pointers are device pointers unless noted.
#include <cuda_runtime.h>
#include "cuda_utils.h"

extern "C" int refutils_gaussian2occgrid(int n_point, float* pos_min, float* pos_max, float* grid_len,
                                         int* grid_dim, float* grid, size_t grid_cells, float* pos,
                                         float* rot, float* scale, float* opac, float* cutoff) {
  cudaMemset(grid, 0, grid_cells * sizeof(float));  // ext.cpp:214 torch::zeros
  gaussians2occgrid_cuda(&n_point, pos_min, pos_max, grid_len, grid_dim, grid, pos, rot, scale, opac,
                         cutoff);
  cudaError_t e = cudaDeviceSynchronize();
  return (int)(e != cudaSuccess ? e : cudaGetLastError());
}

// image [C,H,W] float, mask [H,W] bool (1 byte), res [C,H,W]; `size` is the host int of ext.cpp:77.
extern "C" int refutils_contour_padding(float* image, bool* mask, int C, int H, int W, float* res,
                                        int size) {
  cudaMemset(res, 0, (size_t)C * H * W * sizeof(float));  // ext.cpp:83 zeros_like
  if (size == 0) return 0;                                 // ext.cpp:84
  int reso[3] = {H, W, C};                                 // ext.cpp:82
  contour_padding_cuda(image, mask, reso, res, &size);
  cudaError_t e = cudaDeviceSynchronize();
  return (int)(e != cudaSuccess ? e : cudaGetLastError());
}
