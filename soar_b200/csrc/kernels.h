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
  uint32_t num_rendered;
  uint32_t n_tiles;        // ranges[n_tiles].x = non-empty tiles of the frame (tile_ranges_kernel)
  uint32_t sparse_below;   // set by the launcher: 0 = run; else run only if (non-empty <= this) == one-pixel layout
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
  uint32_t num_rendered;
  uint32_t n_tiles;
  uint32_t sparse_below;
};
// Which blend kernel a launch takes. Two layouts exist of each: two pixels per thread with packed f32x2
// arithmetic (4 warps per tile, fewest instructions: best when the launch fills the GPU) and one pixel
// per thread (8 warps per tile: a long tile list is walked by twice the warps, each with a shorter
// chain; measured 25-35 % faster on launches too small to fill the SMs, where one tile's serial walk is
// the critical path — 32 views of 100k avatar surfels at 512x512: 336 -> 240 us backward, 183 -> 142 us
// forward — and 30 % more instructions when the GPU is full). Without an override the launch picks by
// its number of (tile, Gaussian) pairs. Environment `var`: "0" = packed, n >= 1 = one pixel per thread
// (for the backward n - 1 = its MODE bits). Returns -1 for packed, else the MODE bits.
constexpr uint32_t BLEND_SMALL_LAUNCH_PAIRS = 1u << 20;
int blend_prefers_one_pixel_threads(const char* var, uint32_t num_rendered);
// Larger launches decide ON THE DEVICE, by how many tiles of the frame hold anything: an avatar in front of
// an empty background is a few hundred long tiles however many pairs it has (2M surfels at 1080p: 11.8M
// pairs in 1267 of 8160 tiles), the same latency-bound shape as a small launch. Both layouts are launched;
// every CTA of the one that does not fit the frame returns on its first instruction (a few microseconds
// of empty launch per frame). The threshold: the frame's non-empty tiles do not fill the packed
// kernel's resident CTA slots twice over. Returns 0 when an environment override is in force.
uint32_t blend_sparse_threshold(const char* var);
// (`exit` through inline PTX rather than `return`: an early return edge made the compiler give up hoisting
// the backward's lane-constant reduction bookkeeping out of its inner loop — 343 -> 364 us at c3.)
__device__ __forceinline__ void blend_exit_if_layout_declines(const uint2* ranges, uint32_t n_tiles,
                                                              uint32_t sparse_below, bool one_pixel_layout) {
  if (sparse_below != 0) {
    const bool sparse = ranges[n_tiles].x <= sparse_below;
    if (sparse != one_pixel_layout) asm volatile("exit;");
  }
}
// the backward's choice for a launch (SOAR_BLEND_BWD_SCALAR)
int blend_bwd_scalar_mode(uint32_t num_rendered);
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
