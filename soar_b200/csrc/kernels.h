// Argument blocks and launchers of the rasterizer kernels (one definition, shared
// by the kernels' translation units and the C-ABI orchestration in api.cu).
#pragma once
#include "common.cuh"

namespace soar {

struct PreprocessArgs {
  int P, D, M, W, H;
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  int prefiltered, render_front;
  int grid_x, grid_y;
  const float* means3D;
  const float* scales;
  const float* rotations;
  const float* opacities;
  const float* shs;
  const float* cov3D_precomp;
  const float* colors_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* prcppoint;
  const float* patch_bbox;
  const float* campos;
  const float* config;
  int* radii;
  float* records;
  uint32_t* tiles_touched;
  uint32_t* error_flag;
};
int launch_preprocess_fwd(const PreprocessArgs& a, cudaStream_t stream);

int launch_scan(const uint32_t* in, uint32_t* out, int n, void* status_region,
                size_t status_bytes, cudaStream_t stream);
int launch_duplicate_keys(int P, const float* records, const uint32_t* tiles_touched,
                          const uint32_t* offsets, unsigned long long* keys, uint32_t* values,
                          uint32_t grid_x, cudaStream_t stream);
int launch_radix_sort(unsigned long long* keys[2], uint32_t* values[2], uint32_t n, int end_bit,
                      bool descending, uint32_t* ghist, uint32_t* status, int n_sort_tiles,
                      int* result_buf, cudaStream_t stream);
int launch_tile_ranges(uint32_t L, const unsigned long long* keys, uint2* ranges, int n_tiles,
                       cudaStream_t stream);

struct BlendFwdArgs {
  int W, H, grid_x;
  const uint2* ranges;
  const uint32_t* point_list;
  const float* records;
  const float* bg;
  const float* config;
  float* final_T;
  uint32_t* n_contrib;
  float* final_D;
  float* out_color;
  float* out_normal;
  float* out_depth;
  float* out_opac;
};
int launch_blend_fwd(const BlendFwdArgs& a, int grid_y, cudaStream_t stream);

struct BlendBwdArgs {
  int W, H, grid_x;
  const uint2* ranges;
  const uint32_t* point_list;
  const float* records;
  const float* bg;
  const float* config;
  const float* final_T;
  const uint32_t* n_contrib;
  const float* final_D;
  const float* dL_dpixcolor;
  const float* dL_dpixnormal;
  const float* dL_dpixdepth;
  const float* dL_dpixopac;
  float* gacc;
};
int launch_blend_bwd(const BlendBwdArgs& a, int grid_y, cudaStream_t stream);

struct GeomBwdArgs {
  int P, D, M;
  float tanfovx, tanfovy, focal_x, focal_y, scale_modifier;
  const float* means3D;
  const int* radii;
  const float* shs;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const float* config;
  const float* gacc;
  SoarRastGrads g;
};
int launch_geom_bwd(const GeomBwdArgs& a, cudaStream_t stream);

}  // namespace soar
