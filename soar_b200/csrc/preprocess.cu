// K1 — per-Gaussian preprocess (projection, surfel normal + local homography,
// 3D->2D covariance, conic, radius, tile rect, SH->RGB) for sm_100a.
//
// Behavioural contract: DGR/cuda_rasterizer/forward.cu:205-385 with its helpers
// (auxiliary.h:42-63 ndc2Pix/getRect, :146-208 in_frustum/front_facing,
// :291-388 local_homo; forward.cu:20-71 SH, :74-139 cov2D, :141-202 quat/cov3D).
// The arithmetic below keeps the reference's expression shapes (including the
// places where it silently computes in double) because radii and tile rects
// are compared bit-exactly. What is ours: one 96-byte record per Gaussian
// instead of eight arrays, no clamped[]/cov3D[] side arrays (the backward
// recomputes them), rect stored so key duplication never re-derives it.
#include "common.cuh"
#include "layout.h"
#include "kernels.h"
#include "gauss_math.cuh"

namespace soar {


// SHV: SH coefficients per Gaussian fetched with 16-byte loads (16 or 4); 0 = scalar path.
// 128 threads x 5 CTAs/SM (no register cap bites: the build takes 89 registers, nothing spills).
// Measured at c3 (tools/gpu_bounds_sweep.sh): 5 CTAs 42.3 us, 6 CTAs under an 85-register cap with 28 B
// of spill 45.4 us, 8 CTAs 51.7 us.
// (Fetching the thread's SH row with cp.async into shared memory at kernel start, awaited only where
// the colour is evaluated, was measured: 51 vs 46 us — not done.)
constexpr int PP_THREADS = 128;
template <int SHV>
__global__ void __launch_bounds__(PP_THREADS, 5) preprocess_fwd_kernel(const PreprocessArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P) return;

  a.radii[idx] = 0;
  a.tiles_touched[idx] = 0;
  a.depth_keys[idx] = 0xffffffffu;

  const float* __restrict__ view = a.viewmatrix;
  const float* __restrict__ proj = a.projmatrix;

  float3 p_orig = {a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]};
  // the other small per-Gaussian inputs are requested together with the mean, ahead of the culling
  // tests that decide whether they are needed: one DRAM round trip instead of three
  float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  if (a.rotations != nullptr) q_in = ld_row4(a.rotations, idx, a.rot_vec);
  float3 sc_in = make_float3(0.f, 0.f, 0.f);
  if (a.cov3D_precomp == nullptr)
    sc_in = make_float3(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]);
  const float opac_in = a.opacities[idx];
  float4 p_hom = xform_point_4x4(p_orig, proj);
  float p_w = 1.0f / (p_hom.w + 0.0000001f);
  float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
  float3 p_view = xform_point_4x3(p_orig, view);

  float2 point_image = {ndc_to_pix(p_proj.x, a.W, a.prcppoint[0]),
                        ndc_to_pix(p_proj.y, a.H, a.prcppoint[1])};

  // frustum test against the patch bbox grown by 20 % (auxiliary.h:146-171)
  {
    float x0 = a.patch_bbox[1], y0 = a.patch_bbox[0], x1 = a.patch_bbox[3], y1 = a.patch_bbox[2];
    float w = x1 - x0, h = y1 - y0;
    float expand = 0.2;
    if (p_view.z < 0 || point_image.x < x0 - w * expand || point_image.x >= x1 + w * expand ||
        point_image.y < y0 - h * expand || point_image.y >= y1 + h * expand) {
      if (a.prefiltered) atomicExch(a.error_flag, 1u);  // reference: printf + __trap()
      return;
    }
  }

  Mat3 R = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  if (a.rotations != nullptr) R = quat_to_rot(q_in.x, q_in.y, q_in.z, q_in.w);

  const bool surface = a.config[0] > 0, pix_depth = a.config[2] > 0;
  float3 n_view = {0.f, 0.f, 0.f};
  float view_cos = 0.f;
  float homo[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) homo[i] = 0.f;

  if (surface) {
    n_view = xform_vec_4x3(make_float3(R.m[0][2], R.m[1][2], R.m[2][2]), view);
    float3 ax0_view = xform_vec_4x3(make_float3(R.m[0][0], R.m[1][0], R.m[2][0]), view);
    float3 ax1_view = xform_vec_4x3(make_float3(R.m[0][1], R.m[1][1], R.m[2][1]), view);
    // facing test (auxiliary.h:173-208); the compare is in double
    float dot = p_view.x * n_view.x + p_view.y * n_view.y + p_view.z * n_view.z;
    bool front = !(dot > -0.01);
    if (front) view_cos = dot;
    else if (a.prefiltered) atomicExch(a.error_flag, 1u);
    if (a.render_front && !front) return;
    if (pix_depth) {
      bool grazing = local_homography(p_view, n_view, a.focal_x, a.focal_y, ax0_view, ax1_view, homo);
      if (grazing) return;
    }
  }

  float cov3D_local[6];
  const float* cov3D;
  if (a.cov3D_precomp != nullptr) {
    cov3D = a.cov3D_precomp + idx * 6;
  } else {
    const float3 sc = sc_in;
    scale_rot_to_cov3d(sc, a.scale_modifier, R, surface, cov3D_local);
    cov3D = cov3D_local;
  }

  float3 cov = project_cov(p_view, a.focal_x, a.focal_y, a.tanfovx, a.tanfovy, cov3D, view);

  float det = (cov.x * cov.z - cov.y * cov.y);
  if (det == 0.0f) return;
  float det_inv = 1.f / det;
  float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};

  float mid = 0.5f * (cov.x + cov.z);
  float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
  float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
  float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));

  uint2 rect_min, rect_max;
  tile_rect(point_image, my_radius, rect_min, rect_max, a.grid_x, a.grid_y);
  if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) return;

  float rgb[3];
  if (a.colors_precomp == nullptr) {
    float3 cam = {a.campos[0], a.campos[1], a.campos[2]};
    if (SHV > 0) {
      float sh_local[SHV > 0 ? SHV * 3 : 4];
      const float4* s4 = reinterpret_cast<const float4*>(a.shs + (size_t)idx * SHV * 3);
#pragma unroll
      for (int i = 0; i < SHV * 3 / 4; ++i) {
        const float4 t = __ldg(s4 + i);
        sh_local[4 * i + 0] = t.x; sh_local[4 * i + 1] = t.y;
        sh_local[4 * i + 2] = t.z; sh_local[4 * i + 3] = t.w;
      }
      sh_to_rgb(SHV == 4 ? min(a.D, 1) : a.D, sh_local, p_orig, cam, rgb);
    } else {
      sh_to_rgb(a.D, a.shs + (size_t)idx * a.M * 3, p_orig, cam, rgb);
    }
  } else {
    rgb[0] = a.colors_precomp[3 * idx];
    rgb[1] = a.colors_precomp[3 * idx + 1];
    rgb[2] = a.colors_precomp[3 * idx + 2];
  }

  const int radius_i = my_radius;
  a.radii[idx] = radius_i;
  a.tiles_touched[idx] = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
  a.depth_keys[idx] = __float_as_uint(p_view.z);

  // Cull threshold for the blend kernels: a pixel contributes only when
  // min(0.99, o*exp(power)) >= 1/255, i.e. power >= log(1/(255 o)). REC_THR sits 0.02 below
  // that (2 % in alpha, far above any rounding of expf or of the power polynomial), so
  // "power < REC_THR" proves alpha < 1/255 without evaluating expf; anything else still takes
  // the reference's exact test. o <= 0 (or NaN) never contributes: +inf rejects every pixel
  // only when the exact test would too, so keep such Gaussians on the exact path with -inf.
  const float opac = opac_in;
  float thr = -INFINITY;
  if (opac > 0.f && opac < INFINITY) thr = -logf(255.0f * opac) - 0.02f;

  float4* rec = reinterpret_cast<float4*>(a.records + (size_t)idx * REC_FLOATS);
  rec[0] = make_float4(point_image.x, point_image.y, conic.x, conic.y);
  rec[1] = make_float4(conic.z, opac, thr, p_view.z);
  rec[2] = make_float4(rgb[0], rgb[1], rgb[2], homo[6]);
  rec[3] = make_float4(n_view.x, n_view.y, n_view.z, homo[9]);
  rec[4] = make_float4(homo[0], homo[1], homo[2], homo[3]);
  rec[5] = make_float4(__uint_as_float(rect_min.x | (rect_max.x << 16)),
                       __uint_as_float(rect_min.y | (rect_max.y << 16)),
                       __int_as_float(radius_i), view_cos);
}

int launch_preprocess_fwd(const PreprocessArgs& a_in, cudaStream_t stream) {
  if (a_in.P <= 0) return SOAR_OK;
  PreprocessArgs a = a_in;
  a.rot_vec = aligned16(a.rotations);
  {
    KernelScope ks(K_PREPROCESS, stream);
    const bool aligned = a.shs != nullptr && (reinterpret_cast<uintptr_t>(a.shs) & 15u) == 0;
    if (aligned && a.M == 16) preprocess_fwd_kernel<16><<<ceil_div(a.P, PP_THREADS), PP_THREADS, 0, stream>>>(a);
    else if (aligned && a.M == 4) preprocess_fwd_kernel<4><<<ceil_div(a.P, PP_THREADS), PP_THREADS, 0, stream>>>(a);
    else preprocess_fwd_kernel<0><<<ceil_div(a.P, PP_THREADS), PP_THREADS, 0, stream>>>(a);
  }
  SOAR_LAUNCH_CHECK("preprocess_fwd_kernel");
  return SOAR_OK;
}

}  // namespace soar
