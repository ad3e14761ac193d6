/*
 * rast_oracle.c — TEST INFRASTRUCTURE ONLY (never linked, imported or executed by the product).
 *
 * Plain-C, float32, CPU restatement of the reference's differentiable surfel/Gaussian
 * rasterizer. It exists so that (a) the `-m "not gpu"` tests can check the algorithm against the
 * golden vectors under tests/golden/ (produced by the UNMODIFIED reference CUDA code on a B200,
 * tests/golden/make_golden.py), (b) `smoke()` and the GPU tests have a second, independent
 * checker next to oracle/_ref, and (c) bench.py's `cpu_baseline` leg has a CPU number.
 *
 * Parity status: PINNED against the reference's own outputs (tests/golden/, rast_ files); the
 * reference ships no golden vectors of its own.
 *
 * Every function cites the reference lines it follows. Paths are relative to
 * submodules/diff-gaussian-rasterization/cuda_rasterizer/ of the reference tree.
 * Kernels become loops over idx / (tile, pixel); blocks of 256 threads become a serial walk
 * of each pixel over its tile's list, which is what one CUDA thread does.
 *
 * Floating point: same expression trees in float; the double promotions of the reference
 * (ndc2Pix, front_facing's -0.01, normalize's 1e-8, 0.5*W, 1-1e-6) are kept. nvcc contracts
 * a*b+c into FMAs and gcc (-ffp-contract=fast -mfma) does so too but not always at the same
 * places, so integer outputs (radii, tiles) may differ from the GPU for a ~1e-5 fraction of
 * Gaussians whose value sits on a rounding boundary; tests bound that rate.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define BLOCK_X 16 /* config.h:14-16 */
#define BLOCK_Y 16
#define NUM_CHANNELS 3

/* auxiliary.h:22-40 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                              -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                              0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                              -0.5900435899266435f};

typedef struct {
  int32_t P, D, M, W, H;
  float tanfovx, tanfovy, scale_modifier;
  int32_t prefiltered, render_front, sort_descending, pad_;
  const float *bg, *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations,
      *cov3D_precomp, *viewmatrix, *projmatrix, *prcppoint, *patch_bbox, *campos, *config;
} OView;

/* GeometryState (rasterizer_impl.h:22-45), one array per field, caller-allocated. */
typedef struct {
  float* depths;            /* [P]    */
  uint8_t* clamped;         /* [P,3]  */
  int32_t* radii;           /* [P]    */
  float* means2D;           /* [P,2]  */
  float* cov3D;             /* [P,6]  */
  float* conic_opacity;     /* [P,4]  */
  float* rgb;               /* [P,3]  */
  float* normal;            /* [P,3]  */
  float* Jinv;              /* [P,10] */
  float* viewCos;           /* [P]    */
  uint32_t* tiles_touched;  /* [P]    */
  uint32_t* point_offsets;  /* [P]    */
} OGeom;

typedef struct { float x, y, z; } f3;

/* ---- glm-style column-major 3x3: m[c][r] ---------------------------------------------- */
typedef struct { float m[3][3]; } mat3;
static mat3 mat3_cols(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
  mat3 r = {{{a, b, c}, {d, e, f}, {g, h, i}}};
  return r;
}
/* glm operator*: result[c][r] = sum_k A[k][r] * B[c][k], accumulated k = 0,1,2 */
static mat3 mat3_mul(mat3 A, mat3 B) {
  mat3 r;
  for (int c = 0; c < 3; ++c)
    for (int w = 0; w < 3; ++w)
      r.m[c][w] = A.m[0][w] * B.m[c][0] + A.m[1][w] * B.m[c][1] + A.m[2][w] * B.m[c][2];
  return r;
}
static mat3 mat3_T(mat3 A) {
  mat3 r;
  for (int c = 0; c < 3; ++c)
    for (int w = 0; w < 3; ++w) r.m[c][w] = A.m[w][c];
  return r;
}

/* auxiliary.h:65-103 */
static f3 transformPoint4x3(f3 p, const float* m) {
  f3 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
  return t;
}
static void transformPoint4x4(f3 p, const float* m, float out[4]) {
  out[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  out[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  out[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  out[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}
static f3 transformVec4x3(f3 p, const float* m) {
  f3 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z, m[1] * p.x + m[5] * p.y + m[9] * p.z,
          m[2] * p.x + m[6] * p.y + m[10] * p.z};
  return t;
}
static f3 transformVec4x3Transpose(f3 p, const float* m) {
  f3 t = {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
          m[8] * p.x + m[9] * p.y + m[10] * p.z};
  return t;
}

/* auxiliary.h:42-46 — the literals are doubles, so the whole expression is evaluated in double */
static float ndc2Pix(float v, int S, float prcp) {
  return (float)(((v + 1.0) * S - 1.0) * 0.5 + S * (prcp - 0.5));
}

/* auxiliary.h:53-63 */
static void getRect(float px, float py, int max_radius, int gx, int gy, uint32_t rmin[2], uint32_t rmax[2]) {
  int v;
  v = (int)((px - max_radius) / BLOCK_X); v = v < 0 ? 0 : v; rmin[0] = (uint32_t)(v > gx ? gx : v);
  v = (int)((py - max_radius) / BLOCK_Y); v = v < 0 ? 0 : v; rmin[1] = (uint32_t)(v > gy ? gy : v);
  v = (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X); v = v < 0 ? 0 : v; rmax[0] = (uint32_t)(v > gx ? gx : v);
  v = (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y); v = v < 0 ? 0 : v; rmax[1] = (uint32_t)(v > gy ? gy : v);
}

/* auxiliary.h:114-124 */
static f3 dnormvdv(f3 v, f3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  f3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

/* auxiliary.h:233-239; fmaxf's second argument 0.00000001 is a double literal converted to float */
static float normalize3(float* v) {
  float mod = fmaxf(sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), (float)0.00000001);
  v[0] /= mod; v[1] /= mod; v[2] /= mod;
  return mod;
}

/* auxiliary.h:291-388. Returns 1 when the surfel is seen at a grazing angle (culled). */
static int local_homo(f3 p_view, f3 n_view, float fx, float fy, f3 ax0, f3 ax1, float* res) {
  float ppx = p_view.x / p_view.z, ppy = p_view.y / p_view.z;
  float S_fix = 1000, Svp = (fx + fy) / 2;
  float dir_x0[3], dir_x1[3];
  dir_x0[0] = ppx + 1 / S_fix; dir_x0[1] = ppy; dir_x0[2] = 1;
  float dir_x0_mod = normalize3(dir_x0);
  dir_x1[0] = ppx; dir_x1[1] = ppy + 1 / S_fix; dir_x1[2] = 1;
  float dir_x1_mod = normalize3(dir_x1);
  float thrsh_prj = 0.01;
  float prj_x0 = dir_x0[0] * n_view.x + dir_x0[1] * n_view.y + dir_x0[2] * n_view.z;
  float prj_x1 = dir_x1[0] * n_view.x + dir_x1[1] * n_view.y + dir_x1[2] * n_view.z;
  if ((fabsf(prj_x0 / dir_x0_mod) < thrsh_prj) || (fabsf(prj_x1 / dir_x1_mod) < thrsh_prj)) return 1;
  float t_temp = p_view.x * n_view.x + p_view.y * n_view.y + p_view.z * n_view.z;
  float t_x0 = t_temp / prj_x0, t_x1 = t_temp / prj_x1;
  float xu0[3], xu1[3];
  xu0[0] = dir_x0[0] * t_x0 - p_view.x; xu0[1] = dir_x0[1] * t_x0 - p_view.y; xu0[2] = dir_x0[2] * t_x0 - p_view.z;
  xu1[0] = dir_x1[0] * t_x1 - p_view.x; xu1[1] = dir_x1[1] * t_x1 - p_view.y; xu1[2] = dir_x1[2] * t_x1 - p_view.z;
  /* the reference overwrites the surface-splatting tangent frame by R's axes in view space (:348-354) */
  float u0[3] = {ax0.x, ax0.y, ax0.z}, u1[3] = {ax1.x, ax1.y, ax1.z};
  float J_inv[4];
  J_inv[0] = xu0[0] * u0[0] + xu0[1] * u0[1] + xu0[2] * u0[2];
  J_inv[1] = xu1[0] * u0[0] + xu1[1] * u0[1] + xu1[2] * u0[2];
  J_inv[2] = xu0[0] * u1[0] + xu0[1] * u1[1] + xu0[2] * u1[2];
  J_inv[3] = xu1[0] * u1[0] + xu1[1] * u1[1] + xu1[2] * u1[2];
  for (int i = 0; i < 4; ++i) { J_inv[i] /= (Svp / S_fix); res[i] = J_inv[i]; }
  for (int i = 0; i < 3; ++i) { res[4 + i] = u0[i]; res[7 + i] = u1[i]; }
  return 0;
}

/* forward.cu:141-156 */
static mat3 quaternion2rotmat(const float* q) {
  float r = q[0], x = q[1], y = q[2], z = q[3];
  return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

/* forward.cu:162-202 (incl. the precedence quirk Q2: (mod*surface) ? 0 : scale.z) */
static void computeCov3D(const float* scale, float mod, mat3 R, float* cov3D, int surface) {
  mat3 S = mat3_cols(1, 0, 0, 0, 1, 0, 0, 0, 1);
  S.m[0][0] = mod * scale[0];
  S.m[1][1] = mod * scale[1];
  S.m[2][2] = ((mod * (float)surface) != 0.0f) ? 0 : scale[2];
  mat3 M = mat3_mul(S, R);
  mat3 Sigma = mat3_mul(mat3_T(M), M);
  cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
}

/* forward.cu:74-139 */
static void computeCov2D(f3 t, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                         const float* cov3D, const float* vm, float cov[3]) {
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  mat3 J = mat3_cols(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
                     -(focal_y * t.y) / (t.z * t.z), 0, 0, 0);
  mat3 W = mat3_cols(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
  mat3 T = mat3_mul(W, J);
  mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
  mat3 c = mat3_mul(mat3_mul(mat3_T(T), mat3_T(Vrk)), T);
  cov[0] = c.m[0][0] + 0.3f; cov[1] = c.m[0][1]; cov[2] = c.m[1][1] + 0.3f;
}

/* forward.cu:20-71 */
static void computeColorFromSH(int idx, int deg, int max_coeffs, const float* means, const float* campos,
                               const float* shs, uint8_t* clamped, float out[3]) {
  float dx = means[3 * idx] - campos[0], dy = means[3 * idx + 1] - campos[1], dz = means[3 * idx + 2] - campos[2];
  float len = sqrtf(dx * dx + dy * dy + dz * dz);
  float x = dx / len, y = dy / len, z = dz / len;
  const float* sh = shs + (size_t)idx * max_coeffs * 3;
  for (int c = 0; c < 3; ++c) {
#define SH(k) sh[(k) * 3 + c]
    float result = SH_C0 * SH(0);
    if (deg > 0) {
      result = result - SH_C1 * y * SH(1) + SH_C1 * z * SH(2) - SH_C1 * x * SH(3);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        result = result + SH_C2[0] * xy * SH(4) + SH_C2[1] * yz * SH(5) + SH_C2[2] * (2.0f * zz - xx - yy) * SH(6) +
                 SH_C2[3] * xz * SH(7) + SH_C2[4] * (xx - yy) * SH(8);
        if (deg > 2) {
          result = result + SH_C3[0] * y * (3.0f * xx - yy) * SH(9) + SH_C3[1] * xy * z * SH(10) +
                   SH_C3[2] * y * (4.0f * zz - xx - yy) * SH(11) +
                   SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12) +
                   SH_C3[4] * x * (4.0f * zz - xx - yy) * SH(13) + SH_C3[5] * z * (xx - yy) * SH(14) +
                   SH_C3[6] * x * (xx - 3.0f * yy) * SH(15);
        }
      }
    }
#undef SH
    result += 0.5f;
    clamped[3 * idx + c] = (result < 0);
    out[c] = fmaxf(result, 0.0f);
  }
}

/* rasterizer_impl.cu:35-48 */
static uint32_t getHigherMsb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

int rast_oracle_sort_bits(int W, int H) {
  int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  return 32 + (int)getHigherMsb((uint32_t)(gx * gy));
}

/* ---- K1: preprocessCUDA forward (forward.cu:205-385) + K2 inclusive sum
 *      (rasterizer_impl.cu:242-245). Returns num_rendered, or -1 when prefiltered is set and a
 *      Gaussian was culled (reference: device __trap, auxiliary.h:163-167). */
long rast_oracle_preprocess(const OView* v, OGeom* g) {
  const int P = v->P, W = v->W, H = v->H;
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const float focal_y = H / (2.0f * v->tanfovy), focal_x = W / (2.0f * v->tanfovx); /* rasterizer_impl.cu:201-202 */
  const int surface = v->config[0] > 0, pix_depth = v->config[2] > 0;
  int trap = 0;
#pragma omp parallel for schedule(static) reduction(| : trap)
  for (int idx = 0; idx < P; ++idx) {
    g->radii[idx] = 0;
    g->tiles_touched[idx] = 0;
    f3 p_orig = {v->means3D[3 * idx], v->means3D[3 * idx + 1], v->means3D[3 * idx + 2]};
    float p_hom[4];
    transformPoint4x4(p_orig, v->projmatrix, p_hom);
    float p_w = 1.0f / (p_hom[3] + 0.0000001f);
    f3 p_proj = {p_hom[0] * p_w, p_hom[1] * p_w, p_hom[2] * p_w};
    f3 p_view = transformPoint4x3(p_orig, v->viewmatrix);
    float pix_x = ndc2Pix(p_proj.x, W, v->prcppoint[0]), pix_y = ndc2Pix(p_proj.y, H, v->prcppoint[1]);
    { /* in_frustum, auxiliary.h:146-171; `expand = 0.2` is a double literal stored to float */
      float x0 = v->patch_bbox[1], y0 = v->patch_bbox[0], x1 = v->patch_bbox[3], y1 = v->patch_bbox[2];
      float w = x1 - x0, h = y1 - y0, expand = 0.2;
      if (p_view.z < 0 || pix_x < x0 - w * expand || pix_x >= x1 + w * expand || pix_y < y0 - h * expand ||
          pix_y >= y1 + h * expand) {
        if (v->prefiltered) trap |= 1;
        continue;
      }
    }
    mat3 R = v->rotations ? quaternion2rotmat(v->rotations + 4 * idx) : mat3_cols(1, 0, 0, 0, 1, 0, 0, 0, 1);
    if (surface) {
      f3 a2 = {R.m[0][2], R.m[1][2], R.m[2][2]}, a0 = {R.m[0][0], R.m[1][0], R.m[2][0]}, a1 = {R.m[0][1], R.m[1][1], R.m[2][1]};
      f3 n_view = transformVec4x3(a2, v->viewmatrix);
      f3 ax0_view = transformVec4x3(a0, v->viewmatrix);
      f3 ax1_view = transformVec4x3(a1, v->viewmatrix);
      /* front_facing, auxiliary.h:173-208: compare in double against -0.01 */
      float dot = p_view.x * n_view.x + p_view.y * n_view.y + p_view.z * n_view.z;
      int front;
      if ((double)dot > -0.01) {
        if (v->prefiltered) trap |= 1;
        front = 0;
      } else {
        g->viewCos[idx] = dot;
        front = 1;
      }
      if (v->render_front && !front) continue;
      g->normal[idx * 3 + 0] = n_view.x; g->normal[idx * 3 + 1] = n_view.y; g->normal[idx * 3 + 2] = n_view.z;
      if (pix_depth) {
        float J10[10];
        if (local_homo(p_view, n_view, focal_x, focal_y, ax0_view, ax1_view, J10)) continue;
        for (int i = 0; i < 10; ++i) g->Jinv[idx * 10 + i] = J10[i];
      }
    }
    const float* cov3D;
    if (v->cov3D_precomp) {
      cov3D = v->cov3D_precomp + idx * 6;
    } else {
      computeCov3D(v->scales + 3 * idx, v->scale_modifier, R, g->cov3D + idx * 6, surface);
      cov3D = g->cov3D + idx * 6;
    }
    float cov[3];
    computeCov2D(p_view, focal_x, focal_y, v->tanfovx, v->tanfovy, cov3D, v->viewmatrix, cov);
    float det = (cov[0] * cov[2] - cov[1] * cov[1]);
    if (det == 0.0f) continue;
    float det_inv = 1.f / det;
    float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    float mid = 0.5f * (cov[0] + cov[2]);
    float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    uint32_t rmin[2], rmax[2];
    getRect(pix_x, pix_y, (int)my_radius, gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
    if (!v->colors_precomp) {
      float rgb[3];
      computeColorFromSH(idx, v->D, v->M, v->means3D, v->campos, v->shs, g->clamped, rgb);
      g->rgb[idx * 3 + 0] = rgb[0]; g->rgb[idx * 3 + 1] = rgb[1]; g->rgb[idx * 3 + 2] = rgb[2];
    }
    g->depths[idx] = p_view.z;
    g->radii[idx] = (int32_t)my_radius;
    g->means2D[2 * idx] = pix_x; g->means2D[2 * idx + 1] = pix_y;
    g->conic_opacity[4 * idx + 0] = conic[0]; g->conic_opacity[4 * idx + 1] = conic[1];
    g->conic_opacity[4 * idx + 2] = conic[2]; g->conic_opacity[4 * idx + 3] = v->opacities[idx];
    g->tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
  }
  if (trap) return -1;
  uint32_t run = 0;
  for (int i = 0; i < P; ++i) { run += g->tiles_touched[i]; g->point_offsets[i] = run; }
  return P > 0 ? (long)g->point_offsets[P - 1] : 0;
}

/* Stable LSD radix sort of (u64 key, u32 value) pairs over all 64 bits — equivalent to
 * cub::DeviceRadixSort::SortPairs[Descending] over bits [0, 32+bit) (rasterizer_impl.cu:266-285)
 * because the bits above are zero. */
static void sort_pairs(uint64_t* keys, uint32_t* vals, size_t n, int descending) {
  uint64_t* k2 = (uint64_t*)malloc(n * sizeof(uint64_t));
  uint32_t* v2 = (uint32_t*)malloc(n * sizeof(uint32_t));
  size_t* hist = (size_t*)malloc(65536 * sizeof(size_t));
  const uint64_t flip = descending ? ~(uint64_t)0 : 0;
  uint64_t *ka = keys, *kb = k2;
  uint32_t *va = vals, *vb = v2;
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = pass * 16;
    memset(hist, 0, 65536 * sizeof(size_t));
    for (size_t i = 0; i < n; ++i) hist[((ka[i] ^ flip) >> shift) & 0xffff]++;
    size_t run = 0;
    for (int d = 0; d < 65536; ++d) { size_t c = hist[d]; hist[d] = run; run += c; }
    for (size_t i = 0; i < n; ++i) {
      size_t dst = hist[((ka[i] ^ flip) >> shift) & 0xffff]++;
      kb[dst] = ka[i]; vb[dst] = va[i];
    }
    uint64_t* tk = ka; ka = kb; kb = tk;
    uint32_t* tv = va; va = vb; vb = tv;
  }
  /* 4 passes: result is back in keys/vals */
  free(k2); free(v2); free(hist);
}

/* K3 duplicateWithKeys (rasterizer_impl.cu:66-99), K4 sort, K5 identifyTileRanges (:104-124). */
void rast_oracle_bin(const OView* v, const OGeom* g, long R, uint64_t* keys, uint32_t* point_list, uint32_t* ranges /*[T,2]*/) {
  const int P = v->P, W = v->W, H = v->H;
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
#pragma omp parallel for schedule(dynamic, 1024)
  for (int idx = 0; idx < P; ++idx) {
    if (g->radii[idx] > 0) {
      uint32_t off = (idx == 0) ? 0 : g->point_offsets[idx - 1];
      uint32_t rmin[2], rmax[2];
      getRect(g->means2D[2 * idx], g->means2D[2 * idx + 1], g->radii[idx], gx, gy, rmin, rmax);
      uint32_t dbits;
      memcpy(&dbits, &g->depths[idx], 4);
      for (uint32_t y = rmin[1]; y < rmax[1]; y++)
        for (uint32_t x = rmin[0]; x < rmax[0]; x++) {
          uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
          key <<= 32;
          key |= dbits;
          keys[off] = key;
          point_list[off] = (uint32_t)idx;
          off++;
        }
    }
  }
  sort_pairs(keys, point_list, (size_t)R, v->sort_descending);
  memset(ranges, 0, (size_t)gx * gy * 2 * sizeof(uint32_t));
  for (long idx = 0; idx < R; ++idx) {
    uint32_t cur = (uint32_t)(keys[idx] >> 32);
    if (idx == 0) ranges[2 * cur] = 0;
    else {
      uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
      if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)idx; ranges[2 * cur] = (uint32_t)idx; }
    }
    if (idx == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
  }
}

/* K6 renderCUDA forward (forward.cu:390-692): each pixel walks its tile's list. The block-level
 * early exit only skips work; per pixel the loop ends at `done`. */
void rast_oracle_render(const OView* v, const OGeom* g, const uint32_t* point_list, const uint32_t* ranges,
                        float* final_T, uint32_t* n_contrib, float* final_D, float* out_color,
                        float* out_normal, float* out_depth, float* out_opac) {
  const int W = v->W, H = v->H;
  const int gx = (W + BLOCK_X - 1) / BLOCK_X;
  const int surface = v->config[0] > 0, per_pixel_depth = v->config[2] > 0, normalize_depth = v->config[1] > 0;
  const float* features = v->colors_precomp ? v->colors_precomp : g->rgb; /* rasterizer_impl.cu:298 */
  const size_t HW = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 4) collapse(2)
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const uint32_t tile = (uint32_t)(py / BLOCK_Y) * gx + (px / BLOCK_X);
      const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
      const size_t pix_id = (size_t)W * py + px;
      const float pixfx = (float)px, pixfy = (float)py;
      float T = 1.0f, C[3] = {0, 0, 0}, N[3] = {0, 0, 0}, D = 0;
      uint32_t contributor = 0, last_contributor = 0;
      for (uint32_t k = r0; k < r1; ++k) {
        contributor++;
        const uint32_t id = point_list[k];
        const float dx = g->means2D[2 * id] - pixfx, dy = g->means2D[2 * id + 1] - pixfy;
        const float* co = g->conic_opacity + 4 * id;
        const float dist = (co[0] * dx * dx + co[2] * dy * dy) + 2 * co[1] * dx * dy;
        const float power = -0.5f * dist;
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, co[3] * expf(power));
        if (alpha < 1.0f / 255.0f) continue;
        const float test_T = T * (1 - alpha);
        if (test_T < 0.0001f) break; /* done = true */
        const float w = alpha * T;
        float depth_temp = g->depths[id];
        if (surface && per_pixel_depth) { /* depth_differencing, auxiliary.h:390-397 (.z only is used) */
          const float* Ji = g->Jinv + 10 * id;
          const float u0 = dx * Ji[0] + dy * Ji[1], u1 = dx * Ji[2] + dy * Ji[3];
          depth_temp -= u0 * Ji[6] + u1 * Ji[9];
        }
        D += depth_temp * w;
        for (int ch = 0; ch < 3; ++ch) C[ch] += features[3 * id + ch] * w;
        if (surface) for (int ch = 0; ch < 3; ++ch) N[ch] += g->normal[3 * id + ch] * w;
        T = test_T;
        last_contributor = contributor;
      }
      T = fminf((float)(1 - 0.000001), T); /* forward.cu:620 — double literal */
      final_T[pix_id] = T;
      n_contrib[pix_id] = last_contributor;
      for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix_id] = C[ch] + T * v->bg[ch];
      for (int ch = 0; ch < 3; ++ch) out_normal[ch * HW + pix_id] = surface ? N[ch] : 0;
      out_depth[pix_id] = normalize_depth ? D / (1 - T) : D + T * 10;
      out_opac[pix_id] = 1 - T;
      if (normalize_depth) final_D[pix_id] = D;
    }
}

/* ---- backward ----------------------------------------------------------------------------- */
typedef struct {
  float *dL_dmeans2D /*[P,3]*/, *dL_dconic /*[P,4]*/, *dL_dopacity /*[P]*/, *dL_dcolors /*[P,3]*/,
      *dL_dnormal /*[P,3]*/, *dL_ddepth /*[P]*/, *dL_dmeans3D /*[P,3]*/, *dL_dcov3D /*[P,6]*/,
      *dL_dsh /*[P,M,3]*/, *dL_dscales /*[P,3]*/, *dL_drotations /*[P,4]*/, *dL_dviewmat /*[16]*/,
      *dL_dprojmat /*[16]*/, *dL_dcampos /*[3]*/;
} OGrads; /* all zero-filled by the caller, like rasterize_points.cu:133-147 */

/* K7 renderCUDA backward (backward.cu:529-858). The reference's atomicAdds become plain sums in a
 * fixed order (tile-major, then pixel, then back-to-front), one accumulator set per thread. */
static void render_backward(const OView* v, const OGeom* g, const uint32_t* point_list, const uint32_t* ranges,
                            const float* final_Ts, const uint32_t* n_contrib, const float* final_D,
                            const float* dL_dpixcolor, const float* dL_dpixnormal, const float* dL_dpixdepth,
                            const float* dL_dpixopac, OGrads* o) {
  const int W = v->W, H = v->H, P = v->P;
  const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const int surface = v->config[0] > 0, per_pixel_depth = v->config[2] > 0, normalize_depth = v->config[1] > 0;
  const float* colors = v->colors_precomp ? v->colors_precomp : g->rgb;
  const size_t HW = (size_t)H * W;
  const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  /* per-thread accumulators in double: 13 values per Gaussian */
  double* acc = (double*)calloc((size_t)nthreads * P * 13, sizeof(double));
#pragma omp parallel
  {
    int tid = 0;
#ifdef _OPENMP
    tid = omp_get_thread_num();
#endif
    double* A = acc + (size_t)tid * P * 13;
#pragma omp for schedule(dynamic, 1)
    for (int tile = 0; tile < gx * gy; ++tile) {
      const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
      if (r0 == r1) continue;
      const int tx = tile % gx, ty = tile / gx;
      for (int t = 0; t < BLOCK_X * BLOCK_Y; ++t) {
        const int px = tx * BLOCK_X + (t % BLOCK_X), py = ty * BLOCK_Y + (t / BLOCK_X);
        if (px >= W || py >= H) continue;
        const size_t pix_id = (size_t)W * py + px;
        const float pixfx = (float)px, pixfy = (float)py;
        const float T_final = final_Ts[pix_id];
        const float D_final = normalize_depth ? final_D[pix_id] : 0;
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const uint32_t last_contributor = n_contrib[pix_id];
        float accum_rec[3] = {0, 0, 0}, accum_rec_n[3] = {0, 0, 0}, accum_rec_d = 0;
        float dL_dpixC[3], dL_dpixN[3];
        for (int i = 0; i < 3; ++i) { dL_dpixC[i] = dL_dpixcolor[i * HW + pix_id]; dL_dpixN[i] = dL_dpixnormal[i * HW + pix_id]; }
        const float dL_dpixD = dL_dpixdepth[pix_id] * 1, dL_dpixO = dL_dpixopac[pix_id];
        float last_alpha = 0, last_color[3] = {0, 0, 0}, last_normal[3] = {0, 0, 0}, last_depth = 0;
        for (uint32_t k = r1; k-- > r0;) {
          contributor--;
          if (contributor >= last_contributor) continue;
          const uint32_t id = point_list[k];
          const float dx = g->means2D[2 * id] - pixfx, dy = g->means2D[2 * id + 1] - pixfy;
          const float* co = g->conic_opacity + 4 * id;
          const float dist = (co[0] * dx * dx + co[2] * dy * dy) + 2 * co[1] * dx * dy;
          const float power = -0.5f * dist;
          if (power > 0.0f) continue;
          const float G = expf(power);
          const float alpha = fminf(0.99f, co[3] * G);
          if (alpha < 1.0f / 255.0f) continue;
          T = T / (1.f - alpha);
          const float dchannel_dcolor = alpha * T;
          float Ji[10] = {0};
          if (surface && per_pixel_depth) for (int ch = 0; ch < 10; ++ch) Ji[ch] = g->Jinv[10 * id + ch];
          double* a = A + (size_t)id * 13;
          float dL_dalpha = 0.0f;
          for (int ch = 0; ch < 3; ++ch) {
            const float c_cur = colors[3 * id + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = c_cur;
            const float dL_dchannel = dL_dpixC[ch];
            dL_dalpha += (c_cur - accum_rec[ch]) * dL_dchannel;
            a[0 + ch] += dchannel_dcolor * dL_dchannel;
          }
          if (surface) {
            for (int ch = 0; ch < 3; ++ch) {
              const float n_cur = g->normal[3 * id + ch];
              accum_rec_n[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_rec_n[ch];
              last_normal[ch] = n_cur;
              const float dL_dchannel = dL_dpixN[ch];
              dL_dalpha += (n_cur - accum_rec_n[ch]) * dL_dchannel;
              a[3 + ch] += dchannel_dcolor * dL_dchannel * 10; /* Q4 */
            }
          }
          {
            float d_cur = g->depths[id];
            if (surface && per_pixel_depth) {
              const float u0 = dx * Ji[0] + dy * Ji[1], u1 = dx * Ji[2] + dy * Ji[3];
              d_cur -= u0 * Ji[6] + u1 * Ji[9];
            }
            accum_rec_d = last_alpha * last_depth + (1.f - last_alpha) * accum_rec_d;
            last_depth = d_cur;
            float dL_dchannel = dL_dpixD, dL_dalpha_depth = 0;
            if (normalize_depth) {
              dL_dchannel /= (1.f - T_final);
              dL_dalpha_depth += dL_dpixD * D_final / (1.f - T_final) / (1.f - T_final) * -T_final / (1 - alpha) / T;
            }
            dL_dalpha_depth += (d_cur - accum_rec_d) * dL_dchannel;
            a[6] += dchannel_dcolor * dL_dchannel * 1;
            dL_dalpha += dL_dalpha_depth;
          }
          dL_dalpha *= T;
          dL_dalpha += dL_dpixO * T_final / (1 - alpha);
          last_alpha = alpha;
          float bg_dot_dpixel = 0;
          for (int i = 0; i < 3; ++i) bg_dot_dpixel += v->bg[i] * dL_dpixC[i];
          dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
          if (!normalize_depth) dL_dalpha += (-T_final / (1.f - alpha)) * (10 * dL_dpixD);
          const float dL_ddist = dL_dalpha * co[3] * -0.5f * G;
          float ndc_x = dL_ddist * 2 * (co[0] * dx + co[1] * dy) * ddelx_dx;
          float ndc_y = dL_ddist * 2 * (co[2] * dy + co[1] * dx) * ddely_dy;
          if (surface && per_pixel_depth) {
            ndc_x += 1 * -dL_dpixD * (Ji[6] * Ji[0] + Ji[9] * Ji[2]);
            ndc_y += 1 * -dL_dpixD * (Ji[6] * Ji[1] + Ji[9] * Ji[3]);
          }
          a[7] += ndc_x; a[8] += ndc_y;
          a[9] += dL_ddist * (dx * dx); a[10] += dL_ddist * (1 * dx * dy); a[11] += dL_ddist * (dy * dy);
          a[12] += G * dL_dalpha;
        }
      }
    }
  }
#pragma omp parallel for schedule(static)
  for (int id = 0; id < P; ++id) {
    double s[13] = {0};
    for (int t = 0; t < nthreads; ++t) {
      const double* a = acc + ((size_t)t * P + id) * 13;
      for (int k = 0; k < 13; ++k) s[k] += a[k];
    }
    for (int ch = 0; ch < 3; ++ch) { o->dL_dcolors[3 * id + ch] = (float)s[ch]; o->dL_dnormal[3 * id + ch] = (float)s[3 + ch]; }
    o->dL_ddepth[id] = (float)s[6];
    o->dL_dmeans2D[3 * id] = (float)s[7]; o->dL_dmeans2D[3 * id + 1] = (float)s[8];
    o->dL_dconic[4 * id] = (float)s[9]; o->dL_dconic[4 * id + 1] = (float)s[10]; o->dL_dconic[4 * id + 3] = (float)s[11];
    o->dL_dopacity[id] = (float)s[12];
  }
  free(acc);
}

/* backward.cu:20-158 */
static void sh_backward(int idx, int deg, int max_coeffs, const float* means, const float* campos, const float* shs,
                        const uint8_t* clamped, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs,
                        int lrn_cam, double* dcampos) {
  f3 dir_orig = {means[3 * idx] - campos[0], means[3 * idx + 1] - campos[1], means[3 * idx + 2] - campos[2]};
  float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  const float* sh = shs + (size_t)idx * max_coeffs * 3;
  float* dsh = dL_dshs + (size_t)idx * max_coeffs * 3;
  float dRGB[3], ddir[3] = {0, 0, 0};
  for (int c = 0; c < 3; ++c) dRGB[c] = dL_dcolor[3 * idx + c] * (clamped[3 * idx + c] ? 0 : 1);
  for (int c = 0; c < 3; ++c) {
#define SH(k) sh[(k) * 3 + c]
#define DSH(k) dsh[(k) * 3 + c]
    float dRGBdx = 0, dRGBdy = 0, dRGBdz = 0;
    DSH(0) = SH_C0 * dRGB[c];
    if (deg > 0) {
      DSH(1) = (-SH_C1 * y) * dRGB[c]; DSH(2) = (SH_C1 * z) * dRGB[c]; DSH(3) = (-SH_C1 * x) * dRGB[c];
      dRGBdx = -SH_C1 * SH(3); dRGBdy = -SH_C1 * SH(1); dRGBdz = SH_C1 * SH(2);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        DSH(4) = (SH_C2[0] * xy) * dRGB[c]; DSH(5) = (SH_C2[1] * yz) * dRGB[c];
        DSH(6) = (SH_C2[2] * (2.f * zz - xx - yy)) * dRGB[c]; DSH(7) = (SH_C2[3] * xz) * dRGB[c];
        DSH(8) = (SH_C2[4] * (xx - yy)) * dRGB[c];
        dRGBdx += SH_C2[0] * y * SH(4) + SH_C2[2] * 2.f * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * 2.f * x * SH(8);
        dRGBdy += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * 2.f * -y * SH(6) + SH_C2[4] * 2.f * -y * SH(8);
        dRGBdz += SH_C2[1] * y * SH(5) + SH_C2[2] * 2.f * 2.f * z * SH(6) + SH_C2[3] * x * SH(7);
        if (deg > 2) {
          DSH(9) = (SH_C3[0] * y * (3.f * xx - yy)) * dRGB[c]; DSH(10) = (SH_C3[1] * xy * z) * dRGB[c];
          DSH(11) = (SH_C3[2] * y * (4.f * zz - xx - yy)) * dRGB[c];
          DSH(12) = (SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dRGB[c];
          DSH(13) = (SH_C3[4] * x * (4.f * zz - xx - yy)) * dRGB[c]; DSH(14) = (SH_C3[5] * z * (xx - yy)) * dRGB[c];
          DSH(15) = (SH_C3[6] * x * (xx - 3.f * yy)) * dRGB[c];
          dRGBdx += (SH_C3[0] * SH(9) * 3.f * 2.f * xy + SH_C3[1] * SH(10) * yz + SH_C3[2] * SH(11) * -2.f * xy +
                     SH_C3[3] * SH(12) * -3.f * 2.f * xz + SH_C3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) +
                     SH_C3[5] * SH(14) * 2.f * xz + SH_C3[6] * SH(15) * 3.f * (xx - yy));
          dRGBdy += (SH_C3[0] * SH(9) * 3.f * (xx - yy) + SH_C3[1] * SH(10) * xz +
                     SH_C3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12) * -3.f * 2.f * yz +
                     SH_C3[4] * SH(13) * -2.f * xy + SH_C3[5] * SH(14) * -2.f * yz + SH_C3[6] * SH(15) * -3.f * 2.f * xy);
          dRGBdz += (SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * 4.f * 2.f * yz +
                     SH_C3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13) * 4.f * 2.f * xz +
                     SH_C3[5] * SH(14) * (xx - yy));
        }
      }
    }
#undef SH
#undef DSH
    /* glm::dot(dRGBd?, dL_dRGB) sums the three channels in order 0,1,2 */
    ddir[0] += dRGBdx * dRGB[c]; ddir[1] += dRGBdy * dRGB[c]; ddir[2] += dRGBdz * dRGB[c];
  }
  f3 dd = {ddir[0], ddir[1], ddir[2]};
  f3 dm = dnormvdv(dir_orig, dd);
  if (lrn_cam) { dcampos[0] += -dm.x; dcampos[1] += -dm.y; dcampos[2] += -dm.z; }
  dL_dmeans[3 * idx] += dm.x; dL_dmeans[3 * idx + 1] += dm.y; dL_dmeans[3 * idx + 2] += dm.z;
}

/* K8 computeCov2DCUDA (backward.cu:163-322), K9 preprocessCUDA backward (:437-526) with
 * computeCov3D backward (:326-432). Camera-gradient atomics become serial sums in double. */
static void geom_backward(const OView* v, const OGeom* g, OGrads* o) {
  const int P = v->P;
  const float h_x = v->W / (2.0f * v->tanfovx), h_y = v->H / (2.0f * v->tanfovy);
  const float* vm = v->viewmatrix;
  const float* proj = v->projmatrix;
  const int surface = v->config[0] > 0, lrn_cam = v->config[3] > 0;
  double dview[16] = {0}, dproj[16] = {0}, dcam[3] = {0};
  /* serial over Gaussians when the camera gradients are wanted (deterministic order) */
  for (int idx = 0; idx < P; ++idx) {
    if (!(g->radii[idx] > 0)) continue;
    /* ---- K8 ---- */
    const float* cov3D = (v->cov3D_precomp ? v->cov3D_precomp : g->cov3D) + 6 * idx;
    f3 mean = {v->means3D[3 * idx], v->means3D[3 * idx + 1], v->means3D[3 * idx + 2]};
    float dcx = o->dL_dconic[4 * idx], dcy = o->dL_dconic[4 * idx + 1], dcz = o->dL_dconic[4 * idx + 3];
    f3 t = transformPoint4x3(mean, vm);
    const float limx = 1.3f * v->tanfovx, limy = 1.3f * v->tanfovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
    const float x_grad_mul = txtz < -limx || txtz > limx ? 0 : 1;
    const float y_grad_mul = tytz < -limy || tytz > limy ? 0 : 1;
    float J0 = h_x / t.z, J1 = -(h_x * t.x) / (t.z * t.z), J2 = h_y / t.z, J3 = -(h_y * t.y) / (t.z * t.z);
    mat3 J = mat3_cols(J0, 0.0f, J1, 0.0f, J2, J3, 0, 0, 0);
    mat3 Wm = mat3_cols(vm[0], vm[4], vm[8], vm[1], vm[5], vm[9], vm[2], vm[6], vm[10]);
    mat3 Vrk = mat3_cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    mat3 T = mat3_mul(Wm, J);
    mat3 cov2D = mat3_mul(mat3_mul(mat3_T(T), mat3_T(Vrk)), T);
    float a = cov2D.m[0][0] + 0.3f, b = cov2D.m[0][1], c = cov2D.m[1][1] + 0.3f;
    float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float* dcov = o->dL_dcov3D + 6 * idx;
#define TT(i, j) T.m[i][j]
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
      dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
      dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
      dcov[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
      dcov[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
      dcov[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
      dcov[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
      dcov[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
      dcov[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
    } else {
      for (int i = 0; i < 6; ++i) dcov[i] = 0;
    }
#define VV(i, j) Vrk.m[i][j]
    float dL_dT00 = 2 * (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_da + (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_db;
    float dL_dT01 = 2 * (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_da + (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_db;
    float dL_dT02 = 2 * (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_da + (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_db;
    float dL_dT10 = 2 * (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_dc + (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_db;
    float dL_dT11 = 2 * (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_dc + (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_db;
    float dL_dT12 = 2 * (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_dc + (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_db;
#undef VV
#undef TT
    float dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
    float dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
    float dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
    float dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;
    float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    if (lrn_cam) {
      float dW[16] = {dL_dT00 * J0, dL_dT10 * J2, dL_dT00 * J1 + dL_dT10 * J3, 0,
                      dL_dT01 * J0, dL_dT11 * J2, dL_dT01 * J1 + dL_dT11 * J3, 0,
                      dL_dT02 * J0, dL_dT12 * J2, dL_dT02 * J1 + dL_dT12 * J3, 0, 0, 0, 0, 0};
      for (int i = 0; i < 16; ++i) dview[i] += dW[i];
    }
    float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
    f3 dt = {dL_dtx, dL_dty, dL_dtz};
    f3 dmean_cov = transformVec4x3Transpose(dt, vm);
    float* dmeans = o->dL_dmeans3D + 3 * idx;
    dmeans[0] = dmean_cov.x; dmeans[1] = dmean_cov.y; dmeans[2] = dmean_cov.z; /* assignment, Q8 */

    /* ---- K9 ---- */
    f3 m = mean;
    float m_hom[4];
    transformPoint4x4(m, proj, m_hom);
    float m_w = 1.0f / (m_hom[3] + 0.0000001f);
    float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
    float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
    float d2x = o->dL_dmeans2D[3 * idx], d2y = o->dL_dmeans2D[3 * idx + 1];
    float dmx = (proj[0] * m_w - proj[3] * mul1) * d2x + (proj[1] * m_w - proj[3] * mul2) * d2y;
    float dmy = (proj[4] * m_w - proj[7] * mul1) * d2x + (proj[5] * m_w - proj[7] * mul2) * d2y;
    float dmz = (proj[8] * m_w - proj[11] * mul1) * d2x + (proj[9] * m_w - proj[11] * mul2) * d2y;
    float dL_dd = o->dL_ddepth[idx];
    dmeans[0] += dmx + dL_dd * vm[2]; dmeans[1] += dmy + dL_dd * vm[6]; dmeans[2] += dmz + dL_dd * vm[10];
    if (lrn_cam) {
      float dp[16] = {d2x * m.x * m_w, d2y * m.x * m_w, 0, d2x * -mul1 * m.x + d2y * -mul2 * m.x,
                      d2x * m.y * m_w, d2y * m.y * m_w, 0, d2x * -mul1 * m.y + d2y * -mul2 * m.y,
                      d2x * m.z * m_w, d2y * m.z * m_w, 0, d2x * -mul1 * m.z + d2y * -mul2 * m.z,
                      d2x * m_w,       d2y * m_w,       0, d2x * -mul1 + d2y * -mul2};
      for (int i = 0; i < 16; ++i) dproj[i] += dp[i];
      dview[2] += dL_dd * m.x; dview[6] += dL_dd * m.y; dview[10] += dL_dd * m.z; dview[14] += dL_dd;
    }
    if (v->shs)
      sh_backward(idx, v->D, v->M, v->means3D, v->campos, v->shs, g->clamped, o->dL_dcolors, o->dL_dmeans3D,
                  o->dL_dsh, lrn_cam, dcam);
    if (v->scales) { /* computeCov3D backward, backward.cu:326-432 */
      const float* q = v->rotations + 4 * idx;
      float r = q[0], x = q[1], y = q[2], z = q[3];
      mat3 R = quaternion2rotmat(q);
      float s[3] = {v->scale_modifier * v->scales[3 * idx], v->scale_modifier * v->scales[3 * idx + 1],
                    v->scale_modifier * v->scales[3 * idx + 2]}; /* Q7: s.z not zeroed here */
      mat3 S = mat3_cols(s[0], 0, 0, 0, s[1], 0, 0, 0, s[2]);
      mat3 M = mat3_mul(S, R);
      mat3 dSig = mat3_cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                            0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
      mat3 twoM = M;
      for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) twoM.m[i][j] = 2.0f * M.m[i][j];
      mat3 dL_dM = mat3_mul(twoM, dSig);
      mat3 Rt = mat3_T(R), dL_dMt = mat3_T(dL_dM);
      float* dsc = o->dL_dscales + 3 * idx;
      dsc[0] = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
      dsc[1] = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
      dsc[2] = surface ? 0 : Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] + Rt.m[2][2] * dL_dMt.m[2][2];
      mat3 dRt = dL_dMt;
      for (int j = 0; j < 3; ++j) { dRt.m[0][j] *= s[0]; dRt.m[1][j] *= s[1]; dRt.m[2][j] *= s[2]; }
      const float* dn = o->dL_dnormal + 3 * idx;
      float wn[3] = {dn[0] * vm[0] + dn[1] * vm[1] + dn[2] * vm[2], dn[0] * vm[4] + dn[1] * vm[5] + dn[2] * vm[6],
                     dn[0] * vm[8] + dn[1] * vm[9] + dn[2] * vm[10]};
      dRt.m[2][0] += wn[0]; dRt.m[2][1] += wn[1]; dRt.m[2][2] += wn[2];
      if (lrn_cam) {
        float wrdN[3] = {R.m[0][2], R.m[1][2], R.m[2][2]};
        for (int rr = 0; rr < 3; ++rr) for (int cc = 0; cc < 3; ++cc) dview[4 * rr + cc] += dn[cc] * wrdN[rr];
      }
#define D(i, j) dRt.m[i][j]
      float* dq = o->dL_drotations + 4 * idx;
      dq[0] = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
      dq[1] = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
      dq[2] = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
      dq[3] = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
    }
  }
  for (int i = 0; i < 16; ++i) { o->dL_dviewmat[i] = (float)dview[i]; o->dL_dprojmat[i] = (float)dproj[i]; }
  for (int i = 0; i < 3; ++i) o->dL_dcampos[i] = (float)dcam[i];
}

/* Rasterizer::backward (rasterizer_impl.cu:316-379): K7 -> K8 -> K9. */
void rast_oracle_backward(const OView* v, const OGeom* g, long R, const uint32_t* point_list, const uint32_t* ranges,
                          const float* final_T, const uint32_t* n_contrib, const float* final_D,
                          const float* dL_dpixcolor, const float* dL_dpixnormal, const float* dL_dpixdepth,
                          const float* dL_dpixopac, OGrads* o) {
  (void)R;
  render_backward(v, g, point_list, ranges, final_T, n_contrib, final_D, dL_dpixcolor, dL_dpixnormal,
                  dL_dpixdepth, dL_dpixopac, o);
  geom_backward(v, g, o);
}

int rast_oracle_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
