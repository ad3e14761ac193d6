// K8 + K9 fused — per-Gaussian geometry backward for sm_100a.
//
// Behavioural contract: DGR/cuda_rasterizer/backward.cu:163-322 (cov2D backward,
// which ASSIGNS dL_dmeans, Q8), :437-526 (projection / depth terms, then SH and
// cov3D backward), :20-158 (SH backward), :326-432 (scale / rotation backward
// with S.z = mod*scale.z, Q7; no quaternion-normalisation Jacobian, Q3).
//
// What is ours: one kernel instead of two; it consumes the 64-byte accumulator
// written by the blend backward and writes EVERY element of every gradient
// tensor (zeros for culled Gaussians), so the caller never zero-fills 15
// tensors (DGR/rasterize_points.cu:133-147). cov3D and the SH clamp mask are
// recomputed from the inputs instead of being carried in side arrays.
#include "common.cuh"
#include "layout.h"
#include "kernels.h"
#include "gauss_math.cuh"

namespace soar {


// 4x4 / 3-vector camera-gradient sums: warp shuffle, then one atomic per warp.
__device__ __forceinline__ void warp_atomic_add(float* dst, float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(dst, v);
}

// Gradient stores. With the tensor's bit set in SoarRastGrads.accumulate_mask the value is added to
// what the caller's buffer already holds (multi-view steps accumulate every view straight into the
// flat all-reduce buffer instead of going through a separate add kernel per tensor per view) and
// culled Gaussians write nothing at all.
#ifdef SOAR_GEOM_RMW
__device__ __forceinline__ void put(float* p, float v, bool acc) { *p = acc ? *p + v : v; }
__device__ __forceinline__ void put4(float4* p, float4 v, bool acc) {
  if (acc) {
    const float4 o = *p;
    v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
  }
  *p = v;
}
#else
// Accumulation is a fire-and-forget reduction at the L2 (red.global.add, 16 bytes at a time): the
// same single rounding as load + add + store, without the DRAM round trip of the load at the tail of
// every thread. One add per element and launch, launches ordered by the stream: deterministic.
__device__ __forceinline__ void put(float* p, float v, bool acc) {
  if (acc) atomicAdd(p, v);
  else *p = v;
}
__device__ __forceinline__ void put4(float4* p, float4 v, bool acc) {
  if (acc) atomicAdd(p, v);
  else *p = v;
}
#endif

// SHV: SH coefficients per Gaussian moved with 16-byte loads/stores (16 or 4; rows are then
// 192 / 48 bytes, 16-byte aligned); 0 = generic scalar path for any M.
// 128 threads x 4 CTAs/SM under a 128-register cap (16 B of spill): twice the loads in flight of
// the uncapped 154-register build, which ran one 256-thread CTA per SM and stalled on the long scoreboard.
constexpr int GB_THREADS = 128;
// (Requesting the CTA's row blocks with prefetch.global.L2 up front was measured: 125 vs 114 us —
// the extra requests compete with the demand loads; not done.)
template <bool LRN_CAM, int SHV>
__device__ __forceinline__ void geom_bwd_body(const GeomBwdArgs& a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = idx < a.P;
  // Everything that depends on idx alone is requested before the visibility test is known, so
  // the radius, the accumulator, the record and the parameters share one DRAM round trip
  // (culled Gaussians, ~6 %, fetch them for nothing).
  float4 A0 = make_float4(0.f, 0.f, 0.f, 0.f), A1 = A0, A2 = A0, A3 = A0, r0 = A0, r1 = A0;
  float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  float3 mean_in = make_float3(0.f, 0.f, 0.f), scale_in = make_float3(0.f, 0.f, 0.f);
  int radius = 0;
  if (in_range) {
    radius = a.radii[idx];
    const float4* acc4 = reinterpret_cast<const float4*>(a.gacc + (size_t)idx * GACC_FLOATS);
    A0 = acc4[0]; A1 = acc4[1]; A2 = acc4[2]; A3 = acc4[3];
    const float4* rec = reinterpret_cast<const float4*>(a.records + (size_t)idx * REC_FLOATS);
    r0 = __ldg(rec + 0); r1 = __ldg(rec + 1);
    mean_in = make_float3(a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]);
    if (a.rotations != nullptr) q_in = ld_row4(a.rotations, idx, a.rot_vec);
    if (a.scales != nullptr)
      scale_in = make_float3(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]);
  }
  const bool visible = in_range && radius > 0;
  const uint32_t accm = a.g.accumulate_mask;
  const bool acc_m2d = accm & SOAR_ACC_MEANS2D, acc_col = accm & SOAR_ACC_COLORS,
             acc_opa = accm & SOAR_ACC_OPACITY, acc_m3d = accm & SOAR_ACC_MEANS3D,
             acc_cov = accm & SOAR_ACC_COV3D, acc_sh = accm & SOAR_ACC_SH,
             acc_scl = accm & SOAR_ACC_SCALES, acc_rot = accm & SOAR_ACC_ROTATIONS;
  const float* __restrict__ view = a.viewmatrix;
  const float* __restrict__ proj = a.projmatrix;

  float cam_view[16], cam_proj[16], cam_pos[3];
  if (LRN_CAM) {
#pragma unroll
    for (int i = 0; i < 16; ++i) cam_view[i] = cam_proj[i] = 0.f;
    cam_pos[0] = cam_pos[1] = cam_pos[2] = 0.f;
  }

  if (in_range && !visible) {
    // culled: every gradient is exactly zero
    float* p;
    if (!acc_m2d) { p = a.g.dL_dmeans2D + 3 * (size_t)idx; p[0] = p[1] = p[2] = 0.f; }
    if (!acc_col) { p = a.g.dL_dcolors + 3 * (size_t)idx;  p[0] = p[1] = p[2] = 0.f; }
    if (!acc_opa) a.g.dL_dopacity[idx] = 0.f;
    if (!acc_m3d) { p = a.g.dL_dmeans3D + 3 * (size_t)idx; p[0] = p[1] = p[2] = 0.f; }
    if (!acc_cov) {
      p = a.g.dL_dcov3D + 6 * (size_t)idx;
#pragma unroll
      for (int i = 0; i < 6; ++i) p[i] = 0.f;
    }
    if (!acc_scl) { p = a.g.dL_dscales + 3 * (size_t)idx;  p[0] = p[1] = p[2] = 0.f; }
    if (!acc_rot) st_row4(a.g.dL_drotations, idx, make_float4(0.f, 0.f, 0.f, 0.f), a.rot_vec);
    if (a.M > 0 && a.g.dL_dsh && !acc_sh) {
      if (SHV > 0) {
        float4* p4 = reinterpret_cast<float4*>(a.g.dL_dsh + (size_t)idx * SHV * 3);
#pragma unroll
        for (int i = 0; i < SHV * 3 / 4; ++i) p4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
        p = a.g.dL_dsh + (size_t)idx * a.M * 3;
        for (int i = 0; i < a.M * 3; ++i) p[i] = 0.f;
      }
    }
  }

  if (visible) {
    const bool surface = a.config[0] > 0;
    // accumulator (layout.h GaccSlot): dcolor[3] ddepth dopacity | six dL/ddist moments |
    // dnormal[3] and the per-pixel-depth sum
    const float dL_dcolor[3] = {A0.x, A0.y, A0.z};
    const float dL_ddepth = A0.w;
    const float dL_dopac = A1.x;
    const float m_x = A1.y, m_y = A1.z, m_xx = A1.w, m_xy = A2.x, m_yy = A2.y;
    const float dL_dnormal[3] = {A2.z, A2.w, A3.x};
    // dL/dmean2D (NDC-scaled, Q10) and dL/dconic from the moments (backward.cu:820-855), with the
    // conic the blend kernels used (the forward's record)
    const float4* rec = reinterpret_cast<const float4*>(a.records + (size_t)idx * REC_FLOATS);
    const float cA = r0.z, cB = r0.w, cC = r1.x;
    const float ddelx_dx = 0.5 * a.W;
    const float ddely_dy = 0.5 * a.H;
    float2 dL_dmean2d = {2 * (cA * m_x + cB * m_y) * ddelx_dx, 2 * (cC * m_y + cB * m_x) * ddely_dy};
    if (a.mean_direct) dL_dmean2d = make_float2(2 * m_x * ddelx_dx, 2 * m_y * ddely_dy);
    if (surface && a.config[2] > 0) {
      const float4 r2 = __ldg(rec + 2), r3 = __ldg(rec + 3), r4 = __ldg(rec + 4);
      const float J6 = r2.w, J9 = r3.w, gd = A3.y;
      dL_dmean2d.x += -gd * (J6 * r4.x + J9 * r4.z);
      dL_dmean2d.y += -gd * (J6 * r4.y + J9 * r4.w);
    }
    const float3 dL_dconic = {m_xx, m_xy, m_yy};

    {
      float* p = a.g.dL_dmeans2D + 3 * (size_t)idx;
      put(p, dL_dmean2d.x, acc_m2d); put(p + 1, dL_dmean2d.y, acc_m2d);
      if (!acc_m2d) p[2] = 0.f;
      p = a.g.dL_dcolors + 3 * (size_t)idx;
      put(p, dL_dcolor[0], acc_col); put(p + 1, dL_dcolor[1], acc_col); put(p + 2, dL_dcolor[2], acc_col);
      put(a.g.dL_dopacity + idx, dL_dopac, acc_opa);
    }

    const float3 mean = mean_in;
    Mat3 R = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    const float4 q = q_in;
    if (a.rotations != nullptr) R = quat_to_rot(q.x, q.y, q.z, q.w);
    const float3 scale = scale_in;

    // ---- cov3D as the forward produced it ----------------------------------------
    float cov3D[6];
    if (a.cov3D_precomp != nullptr) {
#pragma unroll
      for (int i = 0; i < 6; ++i) cov3D[i] = a.cov3D_precomp[6 * (size_t)idx + i];
    } else {
      scale_rot_to_cov3d(scale, a.scale_modifier, R, surface, cov3D);
    }

    // ---- 2D covariance backward (backward.cu:163-322) --------------------------------
    float dL_dcov[6];
    float3 dL_dmean_total;
    {
      float3 t = xform_point_4x3(mean, view);
      const float limx = 1.3f * a.tanfovx;
      const float limy = 1.3f * a.tanfovy;
      const float txtz = t.x / t.z;
      const float tytz = t.y / t.z;
      t.x = min(limx, max(-limx, txtz)) * t.z;
      t.y = min(limy, max(-limy, tytz)) * t.z;
      const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
      const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
      const float h_x = a.focal_x, h_y = a.focal_y;

      const float J0 = h_x / t.z, J1 = -(h_x * t.x) / (t.z * t.z), J2 = h_y / t.z,
                  J3 = -(h_y * t.y) / (t.z * t.z);
      Mat3 J = mat3_cols(J0, 0.0f, J1, 0.0f, J2, J3, 0, 0, 0);
      Mat3 Wm = mat3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6],
                          view[10]);
      Mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2],
                           cov3D[4], cov3D[5]);
      Mat3 T = mat3_mul(Wm, J);
      Mat3 cov2D = mat3_mul(mat3_mul(mat3_transpose(T), mat3_transpose(Vrk)), T);

      const float ca = cov2D.m[0][0] += 0.3f;
      const float cb = cov2D.m[0][1];
      const float cc = cov2D.m[1][1] += 0.3f;

      const float denom = ca * cc - cb * cb;
      float dL_da = 0, dL_db = 0, dL_dc = 0;
      const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);

      if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dL_dconic.x + 2 * cb * cc * dL_dconic.y +
                             (denom - ca * cc) * dL_dconic.z);
        dL_dc = denom2inv * (-ca * ca * dL_dconic.z + 2 * ca * cb * dL_dconic.y +
                             (denom - ca * cc) * dL_dconic.x);
        dL_db = denom2inv * 2 *
                (cb * cc * dL_dconic.x - (denom + 2 * cb * cb) * dL_dconic.y + ca * cb * dL_dconic.z);

        dL_dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db +
                      T.m[1][0] * T.m[1][0] * dL_dc);
        dL_dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db +
                      T.m[1][1] * T.m[1][1] * dL_dc);
        dL_dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db +
                      T.m[1][2] * T.m[1][2] * dL_dc);
        dL_dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da +
                     (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][1] * dL_dc;
        dL_dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da +
                     (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
                     2 * T.m[1][0] * T.m[1][2] * dL_dc;
        dL_dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da +
                     (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
                     2 * T.m[1][1] * T.m[1][2] * dL_dc;
      } else {
#pragma unroll
        for (int i = 0; i < 6; i++) dL_dcov[i] = 0;
      }

      // rows of T dotted with the columns of Vrk
      float TV0[3], TV1[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        TV0[c] = T.m[0][0] * Vrk.m[c][0] + T.m[0][1] * Vrk.m[c][1] + T.m[0][2] * Vrk.m[c][2];
        TV1[c] = T.m[1][0] * Vrk.m[c][0] + T.m[1][1] * Vrk.m[c][1] + T.m[1][2] * Vrk.m[c][2];
      }
      const float dL_dT00 = 2 * TV0[0] * dL_da + TV1[0] * dL_db;
      const float dL_dT01 = 2 * TV0[1] * dL_da + TV1[1] * dL_db;
      const float dL_dT02 = 2 * TV0[2] * dL_da + TV1[2] * dL_db;
      const float dL_dT10 = 2 * TV1[0] * dL_dc + TV0[0] * dL_db;
      const float dL_dT11 = 2 * TV1[1] * dL_dc + TV0[1] * dL_db;
      const float dL_dT12 = 2 * TV1[2] * dL_dc + TV0[2] * dL_db;

      const float dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
      const float dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
      const float dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
      const float dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;

      const float tz = 1.f / t.z;
      const float tz2 = tz * tz;
      const float tz3 = tz2 * tz;

      if (LRN_CAM) {
        cam_view[0] += dL_dT00 * J0; cam_view[1] += dL_dT10 * J2; cam_view[2] += dL_dT00 * J1 + dL_dT10 * J3;
        cam_view[4] += dL_dT01 * J0; cam_view[5] += dL_dT11 * J2; cam_view[6] += dL_dT01 * J1 + dL_dT11 * J3;
        cam_view[8] += dL_dT02 * J0; cam_view[9] += dL_dT12 * J2; cam_view[10] += dL_dT02 * J1 + dL_dT12 * J3;
      }

      const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
      const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
      const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 +
                           (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
      dL_dmean_total = xform_vec_4x3_T(make_float3(dL_dtx, dL_dty, dL_dtz), view);
    }
    {
      float* p = a.g.dL_dcov3D + 6 * (size_t)idx;
#pragma unroll
      for (int i = 0; i < 6; ++i) put(p + i, dL_dcov[i], acc_cov);
    }

    // ---- projection + depth terms (backward.cu:464-496) ------------------------------
    {
      const float4 m_hom = xform_point_4x4(mean, proj);
      const float m_w = 1.0f / (m_hom.w + 0.0000001f);
      const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
      const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
      float3 dm;
      dm.x = (proj[0] * m_w - proj[3] * mul1) * dL_dmean2d.x + (proj[1] * m_w - proj[3] * mul2) * dL_dmean2d.y;
      dm.y = (proj[4] * m_w - proj[7] * mul1) * dL_dmean2d.x + (proj[5] * m_w - proj[7] * mul2) * dL_dmean2d.y;
      dm.z = (proj[8] * m_w - proj[11] * mul1) * dL_dmean2d.x + (proj[9] * m_w - proj[11] * mul2) * dL_dmean2d.y;
      const float3 dd = {dL_ddepth * view[2], dL_ddepth * view[6], dL_ddepth * view[10]};
      dL_dmean_total.x += dm.x + dd.x;
      dL_dmean_total.y += dm.y + dd.y;
      dL_dmean_total.z += dm.z + dd.z;

      if (LRN_CAM) {
        const float mx[4] = {mean.x, mean.y, mean.z, 1.f};
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          cam_proj[4 * r + 0] += dL_dmean2d.x * mx[r] * m_w;
          cam_proj[4 * r + 1] += dL_dmean2d.y * mx[r] * m_w;
          cam_proj[4 * r + 3] += dL_dmean2d.x * -mul1 * mx[r] + dL_dmean2d.y * -mul2 * mx[r];
          cam_view[4 * r + 2] += dL_ddepth * mx[r];
        }
        cam_proj[12] += dL_dmean2d.x * m_w;
        cam_proj[13] += dL_dmean2d.y * m_w;
        cam_proj[15] += dL_dmean2d.x * -mul1 + dL_dmean2d.y * -mul2;
        cam_view[14] += dL_ddepth;
      }
    }

    // ---- SH backward (backward.cu:20-158) ----------------------------------------------
    if (a.shs != nullptr) {
      float sh_local[SHV > 0 ? SHV * 3 : 4];
      if (SHV > 0) {
        const float4* s4 = reinterpret_cast<const float4*>(a.shs + (size_t)idx * SHV * 3);
#pragma unroll
        for (int i = 0; i < SHV * 3 / 4; ++i) {
          const float4 t = __ldg(s4 + i);
          sh_local[4 * i + 0] = t.x; sh_local[4 * i + 1] = t.y;
          sh_local[4 * i + 2] = t.z; sh_local[4 * i + 3] = t.w;
        }
      }
      const float* __restrict__ sh = SHV > 0 ? sh_local : a.shs + (size_t)idx * a.M * 3;
      float* dsh = a.g.dL_dsh + (size_t)idx * a.M * 3;
      const int Deg = SHV == 4 ? min(a.D, 1) : a.D;  // SHV == 4 holds degree <= 1 only
      const float3 cam = {a.campos[0], a.campos[1], a.campos[2]};
      const float3 dir_orig = {mean.x - cam.x, mean.y - cam.y, mean.z - cam.z};
      float rgb[3];
      const unsigned clamp_mask = sh_to_rgb(Deg, sh, mean, cam, rgb);
      const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
      const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
      float dRGB[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) dRGB[c] = dL_dcolor[c] * ((clamp_mask >> c) & 1u ? 0 : 1);

      float basis[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) basis[i] = 0.f;
      float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
      basis[0] = kSH_C0;
      if (Deg > 0) {
        basis[1] = -kSH_C1 * y;
        basis[2] = kSH_C1 * z;
        basis[3] = -kSH_C1 * x;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          dx[c] = -kSH_C1 * sh[3 * 3 + c];
          dy[c] = -kSH_C1 * sh[1 * 3 + c];
          dz[c] = kSH_C1 * sh[2 * 3 + c];
        }
        if (Deg > 1) {
          const float xx = x * x, yy = y * y, zz = z * z;
          const float xy = x * y, yz = y * z, xz = x * z;
          basis[4] = kSH_C2[0] * xy;
          basis[5] = kSH_C2[1] * yz;
          basis[6] = kSH_C2[2] * (2.f * zz - xx - yy);
          basis[7] = kSH_C2[3] * xz;
          basis[8] = kSH_C2[4] * (xx - yy);
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const float s4 = sh[4 * 3 + c], s5 = sh[5 * 3 + c], s6 = sh[6 * 3 + c],
                        s7 = sh[7 * 3 + c], s8 = sh[8 * 3 + c];
            dx[c] += kSH_C2[0] * y * s4 + kSH_C2[2] * 2.f * -x * s6 + kSH_C2[3] * z * s7 +
                     kSH_C2[4] * 2.f * x * s8;
            dy[c] += kSH_C2[0] * x * s4 + kSH_C2[1] * z * s5 + kSH_C2[2] * 2.f * -y * s6 +
                     kSH_C2[4] * 2.f * -y * s8;
            dz[c] += kSH_C2[1] * y * s5 + kSH_C2[2] * 2.f * 2.f * z * s6 + kSH_C2[3] * x * s7;
          }
          if (Deg > 2) {
            basis[9] = kSH_C3[0] * y * (3.f * xx - yy);
            basis[10] = kSH_C3[1] * xy * z;
            basis[11] = kSH_C3[2] * y * (4.f * zz - xx - yy);
            basis[12] = kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
            basis[13] = kSH_C3[4] * x * (4.f * zz - xx - yy);
            basis[14] = kSH_C3[5] * z * (xx - yy);
            basis[15] = kSH_C3[6] * x * (xx - 3.f * yy);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float s9 = sh[9 * 3 + c], s10 = sh[10 * 3 + c], s11 = sh[11 * 3 + c],
                          s12 = sh[12 * 3 + c], s13 = sh[13 * 3 + c], s14 = sh[14 * 3 + c],
                          s15 = sh[15 * 3 + c];
              dx[c] += (kSH_C3[0] * s9 * 3.f * 2.f * xy + kSH_C3[1] * s10 * yz +
                        kSH_C3[2] * s11 * -2.f * xy + kSH_C3[3] * s12 * -3.f * 2.f * xz +
                        kSH_C3[4] * s13 * (-3.f * xx + 4.f * zz - yy) + kSH_C3[5] * s14 * 2.f * xz +
                        kSH_C3[6] * s15 * 3.f * (xx - yy));
              dy[c] += (kSH_C3[0] * s9 * 3.f * (xx - yy) + kSH_C3[1] * s10 * xz +
                        kSH_C3[2] * s11 * (-3.f * yy + 4.f * zz - xx) +
                        kSH_C3[3] * s12 * -3.f * 2.f * yz + kSH_C3[4] * s13 * -2.f * xy +
                        kSH_C3[5] * s14 * -2.f * yz + kSH_C3[6] * s15 * -3.f * 2.f * xy);
              dz[c] += (kSH_C3[1] * s10 * xy + kSH_C3[2] * s11 * 4.f * 2.f * yz +
                        kSH_C3[3] * s12 * 3.f * (2.f * zz - xx - yy) +
                        kSH_C3[4] * s13 * 4.f * 2.f * xz + kSH_C3[5] * s14 * (xx - yy));
            }
          }
        }
      }
      if (SHV > 0) {
        // basis[] is zero beyond the active degree
        float o[SHV > 0 ? SHV * 3 : 4];
#pragma unroll
        for (int i = 0; i < SHV; ++i) {
          o[3 * i + 0] = basis[i] * dRGB[0];
          o[3 * i + 1] = basis[i] * dRGB[1];
          o[3 * i + 2] = basis[i] * dRGB[2];
        }
        float4* d4 = reinterpret_cast<float4*>(dsh);
#pragma unroll
        for (int i = 0; i < SHV * 3 / 4; ++i)
          put4(d4 + i, make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]), acc_sh);
      } else {
        const int active = (Deg + 1) * (Deg + 1);
        for (int i = 0; i < a.M; ++i) {
          const float b = (i < active && i < 16) ? basis[i] : 0.f;
          put(dsh + 3 * i + 0, b * dRGB[0], acc_sh);
          put(dsh + 3 * i + 1, b * dRGB[1], acc_sh);
          put(dsh + 3 * i + 2, b * dRGB[2], acc_sh);
        }
      }
      const float3 dL_ddir = {dx[0] * dRGB[0] + dx[1] * dRGB[1] + dx[2] * dRGB[2],
                              dy[0] * dRGB[0] + dy[1] * dRGB[1] + dy[2] * dRGB[2],
                              dz[0] * dRGB[0] + dz[1] * dRGB[1] + dz[2] * dRGB[2]};
      // gradient through the direction normalisation (auxiliary.h:114-124)
      const float sum2 = dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z;
      const float invsum32 = 1.0f / sqrt(sum2 * sum2 * sum2);
      float3 dmn;
      dmn.x = ((+sum2 - dir_orig.x * dir_orig.x) * dL_ddir.x - dir_orig.y * dir_orig.x * dL_ddir.y -
               dir_orig.z * dir_orig.x * dL_ddir.z) * invsum32;
      dmn.y = (-dir_orig.x * dir_orig.y * dL_ddir.x + (sum2 - dir_orig.y * dir_orig.y) * dL_ddir.y -
               dir_orig.z * dir_orig.y * dL_ddir.z) * invsum32;
      dmn.z = (-dir_orig.x * dir_orig.z * dL_ddir.x - dir_orig.y * dir_orig.z * dL_ddir.y +
               (sum2 - dir_orig.z * dir_orig.z) * dL_ddir.z) * invsum32;
      if (LRN_CAM) {
        cam_pos[0] += -dmn.x;
        cam_pos[1] += -dmn.y;
        cam_pos[2] += -dmn.z;
      }
      dL_dmean_total.x += dmn.x;
      dL_dmean_total.y += dmn.y;
      dL_dmean_total.z += dmn.z;
    }
    {
      float* p = a.g.dL_dmeans3D + 3 * (size_t)idx;
      put(p, dL_dmean_total.x, acc_m3d); put(p + 1, dL_dmean_total.y, acc_m3d);
      put(p + 2, dL_dmean_total.z, acc_m3d);
    }

    // ---- scale / rotation backward (backward.cu:326-432) -------------------------------
    if (a.scales != nullptr) {
      const float r = q.x, x = q.y, y = q.z, z = q.w;
      const float3 s = {a.scale_modifier * scale.x, a.scale_modifier * scale.y, a.scale_modifier * scale.z};
      Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
      S.m[0][0] = s.x;
      S.m[1][1] = s.y;
      S.m[2][2] = s.z;
      Mat3 Mx = mat3_mul(S, R);
      Mat3 dSigma = mat3_cols(dL_dcov[0], 0.5f * dL_dcov[1], 0.5f * dL_dcov[2],
                              0.5f * dL_dcov[1], dL_dcov[3], 0.5f * dL_dcov[4],
                              0.5f * dL_dcov[2], 0.5f * dL_dcov[4], dL_dcov[5]);
      Mat3 M2;
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int w = 0; w < 3; ++w) M2.m[c][w] = Mx.m[c][w] * 2.0f;
      Mat3 dL_dM = mat3_mul(M2, dSigma);
      Mat3 Rt = mat3_transpose(R);
      Mat3 dL_dMt = mat3_transpose(dL_dM);

      float3 ds;
      ds.x = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
      ds.y = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
      ds.z = surface ? 0
                     : Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] +
                           Rt.m[2][2] * dL_dMt.m[2][2];
      {
        float* p = a.g.dL_dscales + 3 * (size_t)idx;
        put(p, ds.x, acc_scl); put(p + 1, ds.y, acc_scl); put(p + 2, ds.z, acc_scl);
      }

      Mat3 dL_dRt = dL_dMt;
#pragma unroll
      for (int w = 0; w < 3; ++w) {
        dL_dRt.m[0][w] *= s.x;
        dL_dRt.m[1][w] *= s.y;
        dL_dRt.m[2][w] *= s.z;
      }
      const float3 dN = {dL_dnormal[0], dL_dnormal[1], dL_dnormal[2]};
      const float3 dL_dwrdN = {dN.x * view[0] + dN.y * view[1] + dN.z * view[2],
                               dN.x * view[4] + dN.y * view[5] + dN.z * view[6],
                               dN.x * view[8] + dN.y * view[9] + dN.z * view[10]};
      dL_dRt.m[2][0] += dL_dwrdN.x;
      dL_dRt.m[2][1] += dL_dwrdN.y;
      dL_dRt.m[2][2] += dL_dwrdN.z;

      if (LRN_CAM) {
        const float wn[3] = {R.m[0][2], R.m[1][2], R.m[2][2]};
#pragma unroll
        for (int r3 = 0; r3 < 3; ++r3) {
          cam_view[4 * r3 + 0] += dN.x * wn[r3];
          cam_view[4 * r3 + 1] += dN.y * wn[r3];
          cam_view[4 * r3 + 2] += dN.z * wn[r3];
        }
      }

      const Mat3& G = dL_dRt;
      float4 dq;
      dq.x = 2 * z * (G.m[0][1] - G.m[1][0]) + 2 * y * (G.m[2][0] - G.m[0][2]) +
             2 * x * (G.m[1][2] - G.m[2][1]);
      dq.y = 2 * y * (G.m[1][0] + G.m[0][1]) + 2 * z * (G.m[2][0] + G.m[0][2]) +
             2 * r * (G.m[1][2] - G.m[2][1]) - 4 * x * (G.m[2][2] + G.m[1][1]);
      dq.z = 2 * x * (G.m[1][0] + G.m[0][1]) + 2 * r * (G.m[2][0] - G.m[0][2]) +
             2 * z * (G.m[1][2] + G.m[2][1]) - 4 * y * (G.m[2][2] + G.m[0][0]);
      dq.w = 2 * r * (G.m[0][1] - G.m[1][0]) + 2 * x * (G.m[2][0] + G.m[0][2]) +
             2 * y * (G.m[1][2] + G.m[2][1]) - 4 * z * (G.m[1][1] + G.m[0][0]);
      if (a.rot_vec) {
        put4(reinterpret_cast<float4*>(a.g.dL_drotations) + idx, dq, acc_rot);
      } else {
        float* p = a.g.dL_drotations + 4 * (size_t)idx;
        put(p, dq.x, acc_rot); put(p + 1, dq.y, acc_rot); put(p + 2, dq.z, acc_rot); put(p + 3, dq.w, acc_rot);
      }
    } else {
      if (!acc_scl) {
        float* p = a.g.dL_dscales + 3 * (size_t)idx;
        p[0] = p[1] = p[2] = 0.f;
      }
      if (!acc_rot) st_row4(a.g.dL_drotations, idx, make_float4(0.f, 0.f, 0.f, 0.f), a.rot_vec);
    }
  }

  if (LRN_CAM) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      warp_atomic_add(a.g.dL_dviewmat + i, cam_view[i]);
      warp_atomic_add(a.g.dL_dprojmat + i, cam_proj[i]);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) warp_atomic_add(a.g.dL_dcampos + i, cam_pos[i]);
  }
}

// config lives on the device (the reference passes a float tensor): one kernel, the lrn_cam body
// selected by a uniform branch (a second launch that retires immediately still costs ~6 us).
template <int SHV>
__global__ void __launch_bounds__(GB_THREADS, 4) geom_bwd_kernel(const GeomBwdArgs a) {
  if (a.config[3] > 0) geom_bwd_body<true, SHV>(a);
  else geom_bwd_body<false, SHV>(a);
}

int launch_geom_bwd(const GeomBwdArgs& a_in, cudaStream_t stream) {
  GeomBwdArgs a = a_in;
  a.rot_vec = aligned16(a.rotations) && aligned16(a.g.dL_drotations);
  // camera gradients are sums: clear them here (the reference gets zeros from torch)
  SOAR_CUDA_TRY(cudaMemsetAsync(a.g.dL_dviewmat, 0, 16 * sizeof(float), stream));
  SOAR_CUDA_TRY(cudaMemsetAsync(a.g.dL_dprojmat, 0, 16 * sizeof(float), stream));
  SOAR_CUDA_TRY(cudaMemsetAsync(a.g.dL_dcampos, 0, 3 * sizeof(float), stream));
  if (a.P <= 0) return SOAR_OK;
  const int grid = ceil_div(a.P, GB_THREADS);
  const bool aligned = a.shs != nullptr && a.g.dL_dsh != nullptr &&
                       ((reinterpret_cast<uintptr_t>(a.shs) | reinterpret_cast<uintptr_t>(a.g.dL_dsh)) & 15u) == 0;
  const int shv = (aligned && a.M == 16) ? 16 : (aligned && a.M == 4 ? 4 : 0);
  {
    KernelScope ks(K_GEOM_BWD, stream);
    if (shv == 16) geom_bwd_kernel<16><<<grid, GB_THREADS, 0, stream>>>(a);
    else if (shv == 4) geom_bwd_kernel<4><<<grid, GB_THREADS, 0, stream>>>(a);
    else geom_bwd_kernel<0><<<grid, GB_THREADS, 0, stream>>>(a);
  }
  SOAR_LAUNCH_CHECK("geom_bwd_kernel");
  return SOAR_OK;
}

}  // namespace soar
