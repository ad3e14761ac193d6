// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// C-ABI wrapper that lets tests/ and `bench.py --impl reference` drive the
// UNMODIFIED reference rasterizer (sources stay under /root/reference; they are
// compiled in place by oracle/build_ref.sh into oracle/_ref/libref_dgr.so).
//
// It replaces the reference's torch glue (DGR/rasterize_points.cu:35-205) with
// plain pointers so the build does not need the torch headers: the three
// resize-callback scratch buffers (rasterize_points.cu:27-33,68-75) become
// grow-only cudaMalloc'd buffers owned by a "view context" handle, which plays
// the role the autograd ctx plays in DGR/diff_gaussian_rasterization/__init__.py:141-154.
//
// The debug accessors decode the reference's private buffer layout with the
// reference's own fromChunk() (DGR/cuda_rasterizer/rasterizer_impl.cu:134-184).
#include <cstdint>
#include <cstdio>
#include <functional>
#include <cuda_runtime.h>
#include "rasterizer.h"
#include "rasterizer_impl.h"

namespace {
struct GrowBuf {
  char* ptr = nullptr;
  size_t cap = 0;
  size_t used = 0;
  char* get(size_t n) {
    if (n > cap) {
      if (ptr) cudaFree(ptr);
      size_t want = n + n / 4 + 256;
      if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
      cap = want;
    }
    used = n;
    return ptr;
  }
  void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = used = 0; }
};
struct ViewCtx {
  GrowBuf geom, binning, img;
  int P = 0, R = 0, W = 0, H = 0;
};
}  // namespace

extern "C" {

void* refdgr_create() { return new ViewCtx(); }

void refdgr_destroy(void* h) {
  ViewCtx* c = (ViewCtx*)h;
  if (!c) return;
  c->geom.release(); c->binning.release(); c->img.release();
  delete c;
}

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-105). Outputs must be
// zero-filled by the caller exactly as torch::full(...,0) does there.
int refdgr_forward(void* h, int P, int D, int M, const float* bg, int W, int H,
                   const float* means3D, const float* shs, const float* colors_precomp,
                   const float* opacities, const float* scales, float scale_modifier,
                   const float* rotations, const float* cov3D_precomp,
                   const float* viewmatrix, const float* projmatrix,
                   const float* prcppoint, const float* patchbbox, const float* campos,
                   float tan_fovx, float tan_fovy, int prefiltered, int render_front,
                   int sort_descending, float* config, float* out_color, float* out_normal,
                   float* out_depth, float* out_opac, int* radii, int debug) {
  ViewCtx* c = (ViewCtx*)h;
  c->P = P; c->W = W; c->H = H; c->R = 0;
  if (P == 0) return 0;
  std::function<char*(size_t)> g = [c](size_t n) { return c->geom.get(n); };
  std::function<char*(size_t)> b = [c](size_t n) { return c->binning.get(n); };
  std::function<char*(size_t)> i = [c](size_t n) { return c->img.get(n); };
  try {
    c->R = CudaRasterizer::Rasterizer::forward(
        g, b, i, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities, scales,
        scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, prcppoint,
        patchbbox, campos, tan_fovx, tan_fovy, prefiltered != 0, render_front != 0,
        sort_descending != 0, config, out_color, out_normal, out_depth, out_opac, radii,
        debug != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "refdgr_forward: %s\n", e.what());
    return -1;
  }
  return c->R;
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:107-187). All
// gradient outputs must be zero-filled by the caller (torch::zeros there).
int refdgr_backward(void* h, int P, int D, int M, int R, const float* bg, int W, int H,
                    const float* means3D, const float* shs, const float* colors_precomp,
                    const float* scales, float scale_modifier, const float* rotations,
                    const float* cov3D_precomp, const float* viewmatrix,
                    const float* projmatrix, const float* campos, const float* prcppoint,
                    const float* patchbbox, float tan_fovx, float tan_fovy, const int* radii,
                    const float* dL_dpixcolor, const float* dL_dpixnormal,
                    const float* dL_dpixdepth, const float* dL_dpixopac, float* dL_dmean2D,
                    float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dnormal,
                    float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                    float* dL_dscale, float* dL_drot, float* dL_dviewmat, float* dL_dprojmat,
                    float* dL_dcampos, int debug, float* config) {
  ViewCtx* c = (ViewCtx*)h;
  if (P == 0) return 0;
  try {
    CudaRasterizer::Rasterizer::backward(
        P, D, M, R, bg, W, H, means3D, shs, colors_precomp, scales, scale_modifier,
        rotations, cov3D_precomp, viewmatrix, projmatrix, campos, prcppoint, patchbbox,
        tan_fovx, tan_fovy, radii, c->geom.ptr, c->binning.ptr, c->img.ptr, dL_dpixcolor,
        dL_dpixnormal, dL_dpixdepth, dL_dpixopac, dL_dmean2D, dL_dconic, dL_dopacity,
        dL_dcolor, dL_dnormal, dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
        dL_dviewmat, dL_dprojmat, dL_dcampos, debug != 0, config);
  } catch (const std::exception& e) {
    fprintf(stderr, "refdgr_backward: %s\n", e.what());
    return -1;
  }
  return 0;
}

// ---- debug accessors (device pointers into the reference's scratch buffers) ----
// what: 0 point_list_keys (u64[R])  1 point_list (u32[R])  2 keys_unsorted (u64[R])
//       3 point_list_unsorted (u32[R])
const void* refdgr_binning_ptr(void* h, int what) {
  ViewCtx* c = (ViewCtx*)h;
  if (!c->binning.ptr) return nullptr;
  char* p = c->binning.ptr;
  CudaRasterizer::BinningState s = CudaRasterizer::BinningState::fromChunk(p, c->R);
  switch (what) {
    case 0: return s.point_list_keys;
    case 1: return s.point_list;
    case 2: return s.point_list_keys_unsorted;
    case 3: return s.point_list_unsorted;
  }
  return nullptr;
}

// what: 0 ranges (uint2[T])  1 n_contrib (u32[W*H])  2 accum_alpha=final_T (f32[W*H])
//       3 accum_depth=final_D (f32[W*H])
const void* refdgr_image_ptr(void* h, int what) {
  ViewCtx* c = (ViewCtx*)h;
  if (!c->img.ptr) return nullptr;
  char* p = c->img.ptr;
  CudaRasterizer::ImageState s = CudaRasterizer::ImageState::fromChunk(p, (size_t)c->W * c->H);
  switch (what) {
    case 0: return s.ranges;
    case 1: return s.n_contrib;
    case 2: return s.accum_alpha;
    case 3: return s.accum_depth;
  }
  return nullptr;
}

// what: 0 depths f32[P]  1 means2D float2[P]  2 cov3D f32[6P]  3 conic_opacity float4[P]
//       4 rgb f32[3P]  5 normal f32[3P]  6 Jinv f32[10P]  7 tiles_touched u32[P]
//       8 point_offsets u32[P]  9 clamped bool[3P]  10 viewCos f32[P]
const void* refdgr_geom_ptr(void* h, int what) {
  ViewCtx* c = (ViewCtx*)h;
  if (!c->geom.ptr) return nullptr;
  char* p = c->geom.ptr;
  CudaRasterizer::GeometryState s = CudaRasterizer::GeometryState::fromChunk(p, c->P);
  switch (what) {
    case 0: return s.depths;
    case 1: return s.means2D;
    case 2: return s.cov3D;
    case 3: return s.conic_opacity;
    case 4: return s.rgb;
    case 5: return s.normal;
    case 6: return s.Jinv;
    case 7: return s.tiles_touched;
    case 8: return s.point_offsets;
    case 9: return s.clamped;
    case 10: return s.viewCos;
  }
  return nullptr;
}

int refdgr_num_rendered(void* h) { return ((ViewCtx*)h)->R; }

}  // extern "C"
