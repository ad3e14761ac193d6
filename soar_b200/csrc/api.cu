// C-ABI entry points of libsoar_b200 (declared in include/soar_b200.h).
// Orchestration only: argument checks, carving the caller-owned scratch
// buffers (layout.h), kernel launches on the caller's stream. Plays the role of
// CudaRasterizer::Rasterizer::forward/backward
// (DGR/cuda_rasterizer/rasterizer_impl.cu:188-379) without any torch type.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"
#include "layout.h"
#include "kernels.h"

namespace soar {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

// ---- helpers --------------------------------------------------------------------------
static inline char* at(void* base, size_t off) { return reinterpret_cast<char*>(base) + off; }
static inline const char* at(const void* base, size_t off) {
  return reinterpret_cast<const char*>(base) + off;
}

struct ViewDims {
  int grid_x, grid_y, n_tiles, tile_bits, passes;  // passes of the tile-id sort
};
static inline ViewDims view_dims(int W, int H) {
  ViewDims d;
  d.grid_x = ceil_div(W, TILE_X);
  d.grid_y = ceil_div(H, TILE_Y);
  d.n_tiles = d.grid_x * d.grid_y;
  d.tile_bits = tile_id_bits((uint32_t)d.n_tiles);
  d.passes = sort_passes(d.tile_bits);
  return d;
}

static int check_view(const SoarRastView* v, const char* who) {
  if (!v) { set_error("%s: view is NULL", who); return SOAR_E_INVALID; }
  if (v->P < 0 || v->width <= 0 || v->height <= 0) {
    set_error("%s: bad dimensions P=%d W=%d H=%d", who, v->P, v->width, v->height);
    return SOAR_E_INVALID;
  }
  if (v->width > 16 * 65535 || v->height > 16 * 65535) {
    set_error("%s: image larger than the 16-bit tile rect encoding", who);
    return SOAR_E_INVALID;
  }
  if (v->P > 0) {
    if (!v->means3D || !v->viewmatrix || !v->projmatrix || !v->bg ||
        !v->prcppoint || !v->patch_bbox || !v->campos || !v->config) {
      set_error("%s: a required pointer is NULL", who);
      return SOAR_E_INVALID;
    }
    if ((v->shs == nullptr) == (v->colors_precomp == nullptr)) {
      set_error("%s: provide exactly one of shs / colors_precomp", who);
      return SOAR_E_INVALID;
    }
    const bool have_sr = v->scales != nullptr && v->rotations != nullptr;
    if (have_sr == (v->cov3D_precomp != nullptr)) {
      set_error("%s: provide exactly one of (scales, rotations) / cov3D_precomp", who);
      return SOAR_E_INVALID;
    }
    if (v->shs && (v->sh_degree < 0 || v->sh_degree > 3 ||
                   v->M < (v->sh_degree + 1) * (v->sh_degree + 1))) {
      set_error("%s: sh_degree %d needs M >= %d (got %d)", who, v->sh_degree,
                (v->sh_degree + 1) * (v->sh_degree + 1), v->M);
      return SOAR_E_INVALID;
    }
  }
  return SOAR_OK;
}

}  // namespace soar

using namespace soar;

extern "C" {

const char* soar_last_error(void) { return g_error; }
int soar_version(void) { return 100; }

size_t soar_rast_geom_bytes(int32_t P) { return geom_layout(P < 0 ? 0 : (size_t)P).total; }
size_t soar_rast_image_bytes(int32_t width, int32_t height) {
  return image_layout(width, height).total;
}
size_t soar_rast_binning_bytes(uint32_t num_rendered, int32_t width, int32_t height) {
  return binning_layout(num_rendered, view_dims(width, height).passes).total;
}
size_t soar_rast_backward_scratch_bytes(int32_t P) {
  return align_up((size_t)(P < 0 ? 0 : P) * GACC_FLOATS * sizeof(float), 256) + 256;
}

int soar_rast_forward_bin_async(const SoarRastView* v, void* geom, size_t geom_bytes, int32_t* radii,
                                uint32_t* status_host, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = check_view(v, "soar_rast_forward_bin");
  if (rc) return rc;
  if (!status_host) { set_error("forward_bin: the host status pointer is NULL"); return SOAR_E_INVALID; }
  status_host[0] = 0; status_host[1] = 0;
  if (v->P == 0) return SOAR_OK;
  const GeomLayout L = geom_layout(v->P);
  if (!geom || geom_bytes < L.total) {
    set_error("forward_bin: geom buffer holds %zu bytes, need %zu", geom_bytes, L.total);
    return SOAR_E_WORKSPACE;
  }
  if (!radii || !v->opacities) {
    set_error("forward_bin: radii / opacities is NULL");
    return SOAR_E_INVALID;
  }
  const ViewDims d = view_dims(v->width, v->height);

  SOAR_CUDA_TRY(cudaMemsetAsync(at(geom, L.zero_begin), 0, L.zero_bytes, stream));
  const int n_scan_tiles = ceil_div(v->P, SCAN_TILE);
  uint32_t* tail = reinterpret_cast<uint32_t*>(at(geom, L.scan_status) +
                                               (size_t)n_scan_tiles * sizeof(uint64_t));
  uint32_t* dkeys[2] = {reinterpret_cast<uint32_t*>(at(geom, L.depth_keys[0])),
                        reinterpret_cast<uint32_t*>(at(geom, L.depth_keys[1]))};
  uint32_t* perm[2] = {reinterpret_cast<uint32_t*>(at(geom, L.perm[0])),
                       reinterpret_cast<uint32_t*>(at(geom, L.perm[1]))};
  PreprocessArgs a;
  a.P = v->P; a.D = v->sh_degree; a.M = v->M; a.W = v->width; a.H = v->height;
  a.tanfovx = v->tanfovx; a.tanfovy = v->tanfovy;
  a.focal_y = v->height / (2.0f * v->tanfovy);
  a.focal_x = v->width / (2.0f * v->tanfovx);
  a.scale_modifier = v->scale_modifier;
  a.prefiltered = v->prefiltered; a.render_front = v->render_front;
  a.grid_x = d.grid_x; a.grid_y = d.grid_y;
  a.means3D = v->means3D; a.scales = v->scales; a.rotations = v->rotations;
  a.opacities = v->opacities; a.shs = v->shs; a.cov3D_precomp = v->cov3D_precomp;
  a.colors_precomp = v->colors_precomp; a.viewmatrix = v->viewmatrix;
  a.projmatrix = v->projmatrix; a.prcppoint = v->prcppoint; a.patch_bbox = v->patch_bbox;
  a.campos = v->campos; a.config = v->config;
  a.radii = radii;
  a.records = reinterpret_cast<float*>(at(geom, L.records));
  a.tiles_touched = reinterpret_cast<uint32_t*>(at(geom, L.tiles_touched));
  a.depth_keys = dkeys[0];
  a.error_flag = tail + 1;
  rc = launch_preprocess_fwd(a, stream);
  if (rc) return rc;
  // depth order of the Gaussians (stable: ties stay in index order); 4 passes -> buffer 0
  int dsorted = 0;
  rc = launch_radix_sort(dkeys, perm, (uint32_t)v->P, 32, v->sort_descending != 0,
                         /*iota_values=*/true, /*have_hist=*/false,
                         reinterpret_cast<uint32_t*>(at(geom, L.ghist)),
                         reinterpret_cast<uint32_t*>(at(geom, L.sort_status)), L.n_sort_tiles,
                         &dsorted, stream);
  if (rc) return rc;
  uint32_t* offsets = reinterpret_cast<uint32_t*>(at(geom, L.point_offsets));
  rc = launch_scan(a.tiles_touched, perm[dsorted], offsets, v->P, at(geom, L.scan_status), stream);
  if (rc) return rc;
  SOAR_CUDA_TRY(cudaMemcpyAsync(status_host, offsets + (v->P - 1), sizeof(uint32_t),
                                cudaMemcpyDeviceToHost, stream));
  if (v->prefiltered)
    SOAR_CUDA_TRY(cudaMemcpyAsync(status_host + 1, tail + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                  stream));
  return SOAR_OK;
}

int soar_rast_forward_bin(const SoarRastView* v, void* geom, size_t geom_bytes, int32_t* radii,
                          uint32_t* num_rendered_host, soar_stream_t stream_) {
  if (!num_rendered_host) { set_error("forward_bin: num_rendered_host is NULL"); return SOAR_E_INVALID; }
  uint32_t status[2] = {0u, 0u};
  const int rc = soar_rast_forward_bin_async(v, geom, geom_bytes, radii, status, stream_);
  if (rc) return rc;
  // the one host sync of the forward (reference: rasterizer_impl.cu:250)
  SOAR_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream_));
  *num_rendered_host = status[0];
  if (status[1]) {
    set_error("Point is filtered although prefiltered is set. This shouldn't happen!");
    return SOAR_E_PREFILTER;
  }
  return SOAR_OK;
}


int soar_rast_forward_render(const SoarRastView* v, void* geom, void* binning,
                             size_t binning_bytes, void* img, size_t img_bytes,
                             uint32_t num_rendered, float* out_color, float* out_normal,
                             float* out_depth, float* out_opac, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = check_view(v, "soar_rast_forward_render");
  if (rc) return rc;
  if (!out_color || !out_normal || !out_depth || !out_opac) {
    set_error("forward_render: an output image pointer is NULL");
    return SOAR_E_INVALID;
  }
  const ViewDims d = view_dims(v->width, v->height);
  const ImageLayout IL = image_layout(v->width, v->height);
  if (!img || img_bytes < IL.total) {
    set_error("forward_render: image buffer holds %zu bytes, need %zu", img_bytes, IL.total);
    return SOAR_E_WORKSPACE;
  }
  const BinningLayout BL = binning_layout(num_rendered, d.passes);
  if (!binning || binning_bytes < BL.total) {
    set_error("forward_render: binning buffer holds %zu bytes, need %zu", binning_bytes, BL.total);
    return SOAR_E_WORKSPACE;
  }
  if (v->P > 0 && !geom) { set_error("forward_render: geom is NULL"); return SOAR_E_INVALID; }
  const GeomLayout GL = geom_layout(v->P);
  const float* records = reinterpret_cast<const float*>(at(geom, GL.records));

  uint32_t* keys[2] = {reinterpret_cast<uint32_t*>(at(binning, BL.keys[0])),
                       reinterpret_cast<uint32_t*>(at(binning, BL.keys[1]))};
  uint32_t* values[2] = {reinterpret_cast<uint32_t*>(at(binning, BL.values[0])),
                         reinterpret_cast<uint32_t*>(at(binning, BL.values[1]))};
  uint2* ranges = reinterpret_cast<uint2*>(at(img, IL.ranges));
  int sorted = 0;
  if (num_rendered > 0) {
    SOAR_CUDA_TRY(cudaMemsetAsync(at(binning, BL.zero_begin), 0, BL.zero_bytes, stream));
    uint32_t* ghist = reinterpret_cast<uint32_t*>(at(binning, BL.ghist));
    rc = launch_duplicate_tiles(v->P, records,
                                reinterpret_cast<const uint32_t*>(at(geom, GL.perm[0])),
                                reinterpret_cast<const uint32_t*>(at(geom, GL.point_offsets)),
                                keys[0], values[0], d.grid_x, v->sort_descending != 0, d.tile_bits,
                                ghist, stream);
    if (rc) return rc;
    rc = launch_radix_sort(keys, values, num_rendered, d.tile_bits, v->sort_descending != 0,
                           /*iota_values=*/false, /*have_hist=*/true, ghist,
                           reinterpret_cast<uint32_t*>(at(binning, BL.status)), BL.n_sort_tiles,
                           &sorted, stream);
    if (rc) return rc;
  }
  rc = launch_tile_ranges(num_rendered, keys[sorted], ranges, d.n_tiles, stream);
  if (rc) return rc;

  BlendFwdArgs b;
  b.W = v->width; b.H = v->height; b.grid_x = d.grid_x;
  b.ranges = ranges; b.point_list = values[sorted]; b.records = records;
  b.bg = v->bg; b.config = v->config;
  b.final_T = reinterpret_cast<float*>(at(img, IL.final_T));
  b.n_contrib = reinterpret_cast<uint32_t*>(at(img, IL.n_contrib));
  b.final_D = reinterpret_cast<float*>(at(img, IL.final_D));
  b.out_color = out_color; b.out_normal = out_normal; b.out_depth = out_depth; b.out_opac = out_opac;
  // the ping-pong half the sort left free holds one block-mask byte per list entry
  b.pair_mask = reinterpret_cast<uint8_t*>(keys[sorted ^ 1]);
  b.cfg_flags = v->config_host_flags;
  b.num_rendered = num_rendered;
  b.n_tiles = (uint32_t)d.n_tiles;
  b.sparse_below = 0;
  return launch_blend_fwd(b, d.grid_y, stream);
}

int soar_rast_backward(const SoarRastView* v, const int32_t* radii, const void* geom,
                       const void* binning, const void* img, uint32_t num_rendered,
                       const float* dL_dout_color, const float* dL_dout_normal,
                       const float* dL_dout_depth, const float* dL_dout_opac, void* scratch,
                       size_t scratch_bytes, const SoarRastGrads* grads, soar_stream_t stream_) {
  return soar_rast_backward_ordered(v, radii, geom, binning, img, num_rendered, dL_dout_color,
                                    dL_dout_normal, dL_dout_depth, dL_dout_opac, scratch,
                                    scratch_bytes, grads, nullptr, nullptr, stream_);
}

int soar_rast_backward_ordered(const SoarRastView* v, const int32_t* radii, const void* geom,
                               const void* binning, const void* img, uint32_t num_rendered,
                               const float* dL_dout_color, const float* dL_dout_normal,
                               const float* dL_dout_depth, const float* dL_dout_opac, void* scratch,
                               size_t scratch_bytes, const SoarRastGrads* grads,
                               soar_event_t accumulate_after, soar_event_t accumulated,
                               soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = check_view(v, "soar_rast_backward");
  if (rc) return rc;
  if (!grads || !grads->dL_dviewmat || !grads->dL_dprojmat || !grads->dL_dcampos) {
    set_error("backward: grads / camera gradient pointers are NULL");
    return SOAR_E_INVALID;
  }
  if (v->P > 0) {
    if (!grads->dL_dmeans2D || !grads->dL_dcolors || !grads->dL_dopacity || !grads->dL_dmeans3D ||
        !grads->dL_dcov3D || !grads->dL_dscales || !grads->dL_drotations ||
        (v->shs && !grads->dL_dsh)) {
      set_error("backward: a gradient output pointer is NULL");
      return SOAR_E_INVALID;
    }
    if (!radii || !geom || !img || !binning || !dL_dout_color || !dL_dout_normal ||
        !dL_dout_depth || !dL_dout_opac) {
      set_error("backward: a required input pointer is NULL");
      return SOAR_E_INVALID;
    }
    if (!scratch || scratch_bytes < soar_rast_backward_scratch_bytes(v->P)) {
      set_error("backward: scratch holds %zu bytes, need %zu", scratch_bytes,
                soar_rast_backward_scratch_bytes(v->P));
      return SOAR_E_WORKSPACE;
    }
  }
  const ViewDims d = view_dims(v->width, v->height);
  const ImageLayout IL = image_layout(v->width, v->height);
  const BinningLayout BL = binning_layout(num_rendered, d.passes);
  const GeomLayout GL = geom_layout(v->P);
  float* gacc = reinterpret_cast<float*>(scratch);

  if (v->P > 0) {
    SOAR_CUDA_TRY(cudaMemsetAsync(gacc, 0, (size_t)v->P * GACC_FLOATS * sizeof(float), stream));
    if (num_rendered > 0) {
      const int sorted = d.passes & 1;
      BlendBwdArgs b;
      b.W = v->width; b.H = v->height; b.grid_x = d.grid_x;
      b.ranges = reinterpret_cast<const uint2*>(at(img, IL.ranges));
      b.point_list = reinterpret_cast<const uint32_t*>(at(binning, BL.values[sorted]));
      b.records = reinterpret_cast<const float*>(at(geom, GL.records));
      b.bg = v->bg; b.config = v->config;
      b.final_T = reinterpret_cast<const float*>(at(img, IL.final_T));
      b.n_contrib = reinterpret_cast<const uint32_t*>(at(img, IL.n_contrib));
      b.final_D = reinterpret_cast<const float*>(at(img, IL.final_D));
      b.dL_dpixcolor = dL_dout_color; b.dL_dpixnormal = dL_dout_normal;
      b.dL_dpixdepth = dL_dout_depth; b.dL_dpixopac = dL_dout_opac;
      b.gacc = gacc;
      b.pair_mask = reinterpret_cast<const uint8_t*>(at(binning, BL.keys[sorted ^ 1]));
      b.cfg_flags = v->config_host_flags;
      b.num_rendered = num_rendered;
      b.n_tiles = (uint32_t)d.n_tiles;
      b.sparse_below = 0;
      rc = launch_blend_bwd(b, d.grid_y, stream);
      if (rc) return rc;
    }
  }
  GeomBwdArgs g;
  g.P = v->P; g.D = v->sh_degree; g.M = v->M; g.W = v->width; g.H = v->height;
  g.records = v->P > 0 ? reinterpret_cast<const float*>(at(geom, GL.records)) : nullptr;
  g.tanfovx = v->tanfovx; g.tanfovy = v->tanfovy;
  g.focal_y = v->height / (2.0f * v->tanfovy);
  g.focal_x = v->width / (2.0f * v->tanfovx);
  g.scale_modifier = v->scale_modifier;
  g.means3D = v->means3D; g.radii = radii; g.shs = v->shs; g.scales = v->scales;
  g.rotations = v->rotations; g.cov3D_precomp = v->cov3D_precomp;
  g.viewmatrix = v->viewmatrix; g.projmatrix = v->projmatrix; g.campos = v->campos;
  g.config = v->config; g.gacc = gacc; g.g = *grads;
  {
    const int mode = blend_bwd_scalar_mode(num_rendered);
    g.mean_direct = mode >= 0 && (mode & 4);
  }
  // the only launch that touches the gradient tensors: ordered against the other views of the step
  if (accumulate_after) SOAR_CUDA_TRY(cudaStreamWaitEvent(stream, (cudaEvent_t)accumulate_after, 0));
  rc = launch_geom_bwd(g, stream);
  if (rc) return rc;
  if (accumulated) SOAR_CUDA_TRY(cudaEventRecord((cudaEvent_t)accumulated, stream));
  return SOAR_OK;
}

int soar_rast_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                           const float* projmatrix, uint8_t* present, soar_stream_t stream) {
  // The reference kernel body is commented out (rasterizer_impl.cu:52-62): nothing is written.
  (void)P; (void)means3D; (void)viewmatrix; (void)projmatrix; (void)present; (void)stream;
  return SOAR_OK;
}

const void* soar_rast_debug_ptr(int what, int32_t P, int32_t width, int32_t height,
                                uint32_t num_rendered, const void* geom, const void* binning,
                                const void* img) {
  const ViewDims d = view_dims(width, height);
  const GeomLayout GL = geom_layout(P);
  const ImageLayout IL = image_layout(width, height);
  const BinningLayout BL = binning_layout(num_rendered, d.passes);
  const int sorted = num_rendered > 0 ? (d.passes & 1) : 0;
  switch (what) {
    case 0: return binning ? at(binning, BL.keys[sorted]) : nullptr;  // u32 tile ids
    case 1: return binning ? at(binning, BL.values[sorted]) : nullptr;
    case 2: return img ? at(img, IL.ranges) : nullptr;
    case 3: return img ? at(img, IL.n_contrib) : nullptr;
    case 4: return img ? at(img, IL.final_T) : nullptr;
    case 5: return geom ? at(geom, GL.tiles_touched) : nullptr;
    case 6: return geom ? at(geom, GL.point_offsets) : nullptr;
    case 7: return geom ? at(geom, GL.records) : nullptr;
    case 8: return img ? at(img, IL.final_D) : nullptr;
    case 9: return geom ? at(geom, GL.perm[0]) : nullptr;
  }
  return nullptr;
}

int soar_rast_debug_sorted_keys(int32_t P, int32_t width, int32_t height, uint32_t num_rendered,
                                const void* geom, const void* binning, uint64_t* keys_out,
                                soar_stream_t stream_) {
  if (num_rendered == 0) return SOAR_OK;
  if (!geom || !binning || !keys_out) { set_error("debug_sorted_keys: NULL pointer"); return SOAR_E_INVALID; }
  const ViewDims d = view_dims(width, height);
  const GeomLayout GL = geom_layout(P);
  const BinningLayout BL = binning_layout(num_rendered, d.passes);
  const int sorted = d.passes & 1;
  return launch_debug_keys(num_rendered, reinterpret_cast<const uint32_t*>(at(binning, BL.keys[sorted])),
                           reinterpret_cast<const uint32_t*>(at(binning, BL.values[sorted])),
                           reinterpret_cast<const float*>(at(geom, GL.records)),
                           reinterpret_cast<unsigned long long*>(keys_out), (cudaStream_t)stream_);
}

}  // extern "C"
