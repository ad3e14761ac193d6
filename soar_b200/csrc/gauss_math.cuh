// Per-Gaussian math shared by the forward preprocess (K1) and the geometry
// backward (K8+K9): SH basis, quaternion->rotation, cov3D, EWA projection,
// tile rect, local homography. Expression shapes follow the reference
// (DGR/cuda_rasterizer/forward.cu:20-202, auxiliary.h:42-63,291-388) so that
// radii / tile rects round identically.
#pragma once
#include "common.cuh"

namespace soar {

__device__ const float kSH_C0 = 0.28209479177387814f;
__device__ const float kSH_C1 = 0.4886025119029199f;
__device__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                -0.5900435899266435f};

// pixel centre from NDC; the literals are double on purpose (auxiliary.h:42-46)
__device__ __forceinline__ float ndc_to_pix(float v, int S, float prcp) {
  return ((v + 1.0) * S - 1.0) * 0.5 + S * (prcp - 0.5);
}

// Local homography between the screen and the surfel's tangent plane
// (behaviour of auxiliary.h:291-388). Two rays through the pixel centre shifted
// by 1/1000 of a focal length along +x and +y are intersected with the tangent
// plane; the two hit offsets expressed in the surfel's own in-plane axes give
// the 2x2 pixel->tangent Jacobian (rescaled to pixels). Returns true when either
// ray grazes the plane (|cos| < 0.01) and the surfel must be culled.
// out = {J00, J01, J10, J11, axis0.xyz, axis1.xyz}.
struct Ray3 {
  float v[3];
  float len;
};

__device__ __forceinline__ Ray3 unit_ray(float x, float y) {
  Ray3 r;
  r.v[0] = x;
  r.v[1] = y;
  r.v[2] = 1;
  r.len = fmaxf(sqrtf(r.v[0] * r.v[0] + r.v[1] * r.v[1] + r.v[2] * r.v[2]), 0.00000001);
  r.v[0] /= r.len;
  r.v[1] /= r.len;
  r.v[2] /= r.len;
  return r;
}

__device__ __forceinline__ bool local_homography(float3 p_view, float3 n_view, float fx, float fy,
                                                 float3 ax0, float3 ax1, float* out) {
  const float unit = 1000;            // screen step = 1/unit of a focal length
  const float focal = (fx + fy) / 2;  // mean focal length in pixels
  const float px = p_view.x / p_view.z, py = p_view.y / p_view.z;
  const Ray3 rx = unit_ray(px + 1 / unit, py);
  const Ray3 ry = unit_ray(px, py + 1 / unit);

  const float graze = 0.01;
  const float cx = rx.v[0] * n_view.x + rx.v[1] * n_view.y + rx.v[2] * n_view.z;
  const float cy = ry.v[0] * n_view.x + ry.v[1] * n_view.y + ry.v[2] * n_view.z;
  if ((fabsf(cx / rx.len) < graze) || (fabsf(cy / ry.len) < graze)) return true;

  // ray/plane hits, relative to the surfel centre
  const float plane = p_view.x * n_view.x + p_view.y * n_view.y + p_view.z * n_view.z;
  const float tx = plane / cx;
  const float ty = plane / cy;
  const float hx[3] = {rx.v[0] * tx - p_view.x, rx.v[1] * tx - p_view.y, rx.v[2] * tx - p_view.z};
  const float hy[3] = {ry.v[0] * ty - p_view.x, ry.v[1] * ty - p_view.y, ry.v[2] * ty - p_view.z};

  const float px_scale = focal / unit;
  out[0] = (hx[0] * ax0.x + hx[1] * ax0.y + hx[2] * ax0.z) / px_scale;
  out[1] = (hy[0] * ax0.x + hy[1] * ax0.y + hy[2] * ax0.z) / px_scale;
  out[2] = (hx[0] * ax1.x + hx[1] * ax1.y + hx[2] * ax1.z) / px_scale;
  out[3] = (hy[0] * ax1.x + hy[1] * ax1.y + hy[2] * ax1.z) / px_scale;
  out[4] = ax0.x; out[5] = ax0.y; out[6] = ax0.z;
  out[7] = ax1.x; out[8] = ax1.y; out[9] = ax1.z;
  return false;
}

// Rotation from the (un-normalised) quaternion, column-major (forward.cu:141-156).
__device__ __forceinline__ Mat3 quat_to_rot(float r, float x, float y, float z) {
  return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// SH -> RGB for one Gaussian (forward.cu:20-71). `sh` points at this Gaussian's
// [M][3] coefficients. Writes rgb (clamped at 0); returns the clamp mask.
__device__ __forceinline__ unsigned sh_to_rgb(int deg, const float* __restrict__ sh, float3 pos,
                                              float3 campos, float* rgb) {
  float dx = pos.x - campos.x, dy = pos.y - campos.y, dz = pos.z - campos.z;
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float x = dx / len, y = dy / len, z = dz / len;
  unsigned clamp_mask = 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float result = kSH_C0 * sh[0 * 3 + c];
    if (deg > 0) {
      result = result - kSH_C1 * y * sh[1 * 3 + c] + kSH_C1 * z * sh[2 * 3 + c] -
               kSH_C1 * x * sh[3 * 3 + c];
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z;
        float xy = x * y, yz = y * z, xz = x * z;
        result = result + kSH_C2[0] * xy * sh[4 * 3 + c] + kSH_C2[1] * yz * sh[5 * 3 + c] +
                 kSH_C2[2] * (2.0f * zz - xx - yy) * sh[6 * 3 + c] +
                 kSH_C2[3] * xz * sh[7 * 3 + c] + kSH_C2[4] * (xx - yy) * sh[8 * 3 + c];
        if (deg > 2) {
          result = result + kSH_C3[0] * y * (3.0f * xx - yy) * sh[9 * 3 + c] +
                   kSH_C3[1] * xy * z * sh[10 * 3 + c] +
                   kSH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11 * 3 + c] +
                   kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12 * 3 + c] +
                   kSH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13 * 3 + c] +
                   kSH_C3[5] * z * (xx - yy) * sh[14 * 3 + c] +
                   kSH_C3[6] * x * (xx - 3.0f * yy) * sh[15 * 3 + c];
        }
      }
    }
    result += 0.5f;
    if (result < 0) clamp_mask |= 1u << c;
    rgb[c] = fmaxf(result, 0.0f);
  }
  return clamp_mask;
}

// cov3D = (S R)^T (S R), upper triangle (forward.cu:162-202). Note the surfel
// quirk (Q2): the third scale is 0 when mod*surface != 0, else the UNSCALED z.
__device__ __forceinline__ void scale_rot_to_cov3d(float3 scale, float mod, const Mat3& R,
                                                   bool surface, float* cov3D) {
  Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  S.m[0][0] = mod * scale.x;
  S.m[1][1] = mod * scale.y;
  S.m[2][2] = mod * surface ? 0 : scale.z;
  Mat3 Mx = mat3_mul(S, R);
  Mat3 Sigma = mat3_mul(mat3_transpose(Mx), Mx);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
}

// EWA projection of the 3D covariance (forward.cu:74-139): returns (a, b, c) of
// the 2D covariance with the 0.3 px low-pass already added.
__device__ __forceinline__ float3 project_cov(const float3& mean_view, float focal_x, float focal_y,
                                              float tan_fovx, float tan_fovy, const float* cov3D,
                                              const float* viewmatrix) {
  float3 t = mean_view;
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;

  Mat3 J = mat3_cols(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
                     0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
                     0, 0, 0);
  Mat3 Wm = mat3_cols(viewmatrix[0], viewmatrix[4], viewmatrix[8],
                      viewmatrix[1], viewmatrix[5], viewmatrix[9],
                      viewmatrix[2], viewmatrix[6], viewmatrix[10]);
  Mat3 T = mat3_mul(Wm, J);
  Mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2],
                       cov3D[1], cov3D[3], cov3D[4],
                       cov3D[2], cov3D[4], cov3D[5]);
  Mat3 cov = mat3_mul(mat3_mul(mat3_transpose(T), mat3_transpose(Vrk)), T);
  cov.m[0][0] += 0.3f;
  cov.m[1][1] += 0.3f;
  return make_float3(cov.m[0][0], cov.m[0][1], cov.m[1][1]);
}

__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rect_min,
                                          uint2& rect_max, unsigned grid_x, unsigned grid_y) {
  rect_min = {min(grid_x, max((int)0, (int)((p.x - max_radius) / TILE_X))),
              min(grid_y, max((int)0, (int)((p.y - max_radius) / TILE_Y)))};
  rect_max = {min(grid_x, max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X))),
              min(grid_y, max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)))};
}

}  // namespace soar
