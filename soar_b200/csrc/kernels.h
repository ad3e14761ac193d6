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
  uint32_t* depth_keys;  // depth bits of visible Gaussians, 0xffffffff for culled ones
  uint32_t* error_flag;
  int rot_vec;  // rotations is 16-byte aligned (set by the launcher)
};
int launch_preprocess_fwd(const PreprocessArgs& a, cudaStream_t stream);

int launch_scan(const uint32_t* in, const uint32_t* index, uint32_t* out, int n,
                void* status_region, cudaStream_t stream);
int launch_duplicate_tiles(int P, const float* records, const uint32_t* perm,
                           const uint32_t* offsets_incl, uint32_t* keys, uint32_t* values,
                           uint32_t grid_x, bool descending, int tile_bits, uint32_t* ghist,
                           cudaStream_t stream);
int launch_radix_sort(uint32_t* keys[2], uint32_t* values[2], uint32_t n, int end_bit,
                      bool descending, bool iota_values, bool have_hist, uint32_t* ghist,
                      uint32_t* status, int n_sort_tiles, int* result_buf, cudaStream_t stream);
int launch_tile_ranges(uint32_t L, const uint32_t* keys, uint2* ranges, int n_tiles,
                       cudaStream_t stream);
int launch_debug_keys(uint32_t L, const uint32_t* tiles, const uint32_t* ids, const float* records,
                      unsigned long long* out, cudaStream_t stream);

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
  uint8_t* pair_mask;  // [R] block mask of every staged list entry (read back by the backward)
  uint32_t cfg_flags;  // SoarRastView.config_host_flags
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
  const uint8_t* pair_mask;  // [R] written by the forward blend of the same view
  uint32_t cfg_flags;  // SoarRastView.config_host_flags
};
// -1: the packed product kernel; >= 0: the scalar twin's MODE bits (SOAR_BLEND_BWD_SCALAR, blend_bwd.cu)
int blend_bwd_scalar_mode();
int launch_blend_bwd(const BlendBwdArgs& a, int grid_y, cudaStream_t stream);

struct GeomBwdArgs {
  int P, D, M, W, H;
  const float* records;  // the forward's per-Gaussian records (conic, Jinv)
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
  int rot_vec;      // rotations and dL_drotations are 16-byte aligned (set by the launcher)
  int mean_direct;  // GA_MX / GA_MY hold sum(dd (A dx + B dy)), sum(dd (C dy + B dx)) instead of the moments
  SoarRastGrads g;
};
int launch_geom_bwd(const GeomBwdArgs& a, cudaStream_t stream);

}  // namespace soar
