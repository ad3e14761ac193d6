// SMPL-X body-model forward (`lbs()`) for one pose on sm_100a.
//
// Behavioural contract (all torch ops in the reference, batch size 1):
//   * lbs()                  TS/utils/smplx/lbs.py:147-246
//   * blend_shapes           :269-290   v_shaped = v_template + shapedirs . betas
//   * vertices2joints        :249-266   J = J_regressor . v_shaped
//   * batch_rodrigues        :293-328   angle = |v + 1e-8| (eps per component, Q16)
//   * pose blend shapes      :209-214   (R[1:] - I).flatten() . posedirs
//   * batch_rigid_transform  :343-396   chain over parents; A = G - pad(G . [J;0])
//   * vertex skinning        :225-239   T = W . A;  v = T . [v_posed; 1]
//   * SMPLX.forward tail     TS/utils/smplx/body_models.py:1380-1383: transl is added to the
//     vertices, the joints and A[:, :3, 3] (Q17)
//   * bones                  TS/utils/smpl.py:609  cano2live = A . inv_mats
//
// On the GPU the reference runs this as ~100 tiny torch launches. Here it is five:
//   1 shape blend           one thread per vertex coordinate (80-byte shapedirs rows)
//   2 pose-corrective GEMV  the only real traffic: posedirs is (J-1)*9 x 3V floats (61 MB for
//                           SMPL-X) streamed ONCE, coalesced, split over row chunks so that
//                           ~8k warps are in flight; partial sums are kept per chunk and added in
//                           a fixed order by kernel 5 (deterministic, no float atomics)
//   3 joint regression      one CTA per joint, fixed-order tree reduction over the vertices
//   4 kinematic chain       one CTA; joints of the same tree depth are composed in parallel
//   5 vertex skinning       one thread per vertex, the 55 3x4 transforms in shared memory
// HBM-bound on posedirs; nothing here is a dense contraction worth the tensor pipe.
#include <math.h>
#include "common.cuh"

namespace soar {

constexpr int SMPLX_MAX_J = 64;
constexpr int SMPLX_MAX_NB = 64;
constexpr int SMPLX_PD_SPLIT = 8;   // row chunks of the pose-corrective GEMV
constexpr int SMPLX_THREADS = 256;

struct SmplxLayout {
  size_t v_shaped;  // f32[3V]
  size_t partial;   // f32[SMPLX_PD_SPLIT][3V]
  size_t joints;    // f32[3J] rest-pose joints of the shaped body
  size_t A;         // f32[12J] rows 0..2 of A WITHOUT transl (what the skinning uses)
  size_t total;
};

inline SmplxLayout smplx_layout(size_t V, size_t J) {
  SmplxLayout L;
  size_t off = 0;
  L.v_shaped = off; off = align_up(off + 3 * V * 4, 256);
  L.partial = off;  off = align_up(off + SMPLX_PD_SPLIT * 3 * V * 4, 256);
  L.joints = off;   off = align_up(off + 3 * J * 4, 256);
  L.A = off;        off = align_up(off + 12 * J * 4, 256);
  L.total = off + 256;
  return L;
}

// Rodrigues with the reference's operation order (lbs.py:311-327): R = I + sin K + (1-cos) K K.
__device__ __forceinline__ void rodrigues(const float* __restrict__ rv, float (&R)[9]) {
  const float ex = rv[0] + 1e-8f, ey = rv[1] + 1e-8f, ez = rv[2] + 1e-8f;
  const float angle = sqrtf(ex * ex + ey * ey + ez * ez);
  const float rx = rv[0] / angle, ry = rv[1] / angle, rz = rv[2] / angle;
  const float s = sinf(angle), c1 = 1.f - cosf(angle);
  const float K[9] = {0.f, -rz, ry, rz, 0.f, -rx, -ry, rx, 0.f};
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float kk = K[r * 3 + 0] * K[0 * 3 + c] + K[r * 3 + 1] * K[1 * 3 + c] + K[r * 3 + 2] * K[2 * 3 + c];
      R[r * 3 + c] = (r == c ? 1.f : 0.f) + s * K[r * 3 + c] + c1 * kk;
    }
}

// ---- 1. shape blend -----------------------------------------------------------------------
__global__ void __launch_bounds__(SMPLX_THREADS)
smplx_shape_kernel(const float* __restrict__ v_template, const float* __restrict__ shapedirs,
                   const float* __restrict__ betas, int C, int NB, float* __restrict__ v_shaped) {
  __shared__ float s_b[SMPLX_MAX_NB];
  if (threadIdx.x < NB) s_b[threadIdx.x] = betas[threadIdx.x];
  __syncthreads();
  const int c = blockIdx.x * SMPLX_THREADS + threadIdx.x;
  if (c >= C) return;
  const float* row = shapedirs + (size_t)c * NB;
  float acc = 0.f;
  for (int l = 0; l < NB; ++l) acc += s_b[l] * row[l];
  v_shaped[c] = v_template[c] + acc;
}

// ---- 2. pose-corrective GEMV: partial[s][c] = sum_{p in chunk s} feat[p] * posedirs[p][c] ----
__global__ void __launch_bounds__(SMPLX_THREADS)
smplx_posedirs_kernel(const float* __restrict__ posedirs, const float* __restrict__ full_pose,
                      int C, int n_feat, int rows_per_chunk, float* __restrict__ partial) {
  extern __shared__ float s_feat[];  // rows_per_chunk
  const int p0 = blockIdx.y * rows_per_chunk;
  const int p1 = min(p0 + rows_per_chunk, n_feat);
  // feature p belongs to joint 1 + p / 9, entry p % 9 of (R - I)
  for (int p = p0 + threadIdx.x; p < p1; p += SMPLX_THREADS) {
    float R[9];
    rodrigues(full_pose + 3 * (1 + p / 9), R);
    const int e = p % 9;
    s_feat[p - p0] = R[e] - ((e == 0 || e == 4 || e == 8) ? 1.f : 0.f);
  }
  __syncthreads();
  const int c = blockIdx.x * SMPLX_THREADS + threadIdx.x;
  if (c >= C) return;
  const float* col = posedirs + (size_t)p0 * C + c;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
  int p = p0;
  for (; p + 4 <= p1; p += 4) {  // four independent loads in flight per thread
    const float a = col[0], b = col[(size_t)C], d = col[2 * (size_t)C], e = col[3 * (size_t)C];
    acc0 += s_feat[p - p0] * a;
    acc1 += s_feat[p - p0 + 1] * b;
    acc2 += s_feat[p - p0 + 2] * d;
    acc3 += s_feat[p - p0 + 3] * e;
    col += 4 * (size_t)C;
  }
  for (; p < p1; ++p) {
    acc0 += s_feat[p - p0] * col[0];
    col += C;
  }
  partial[(size_t)blockIdx.y * C + c] = (acc0 + acc1) + (acc2 + acc3);
}

// ---- 3. joint regression: J[j][k] = sum_v Jreg[j][v] * v_shaped[v][k] ------------------------
__global__ void __launch_bounds__(SMPLX_THREADS)
smplx_joints_kernel(const float* __restrict__ J_regressor, const float* __restrict__ v_shaped,
                    int V, float* __restrict__ joints) {
  __shared__ float s_red[3][SMPLX_THREADS];
  const int j = blockIdx.x;
  const float* row = J_regressor + (size_t)j * V;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f;
  for (int v = threadIdx.x; v < V; v += SMPLX_THREADS) {
    const float w = row[v];
    if (w != 0.f) {  // the regressor is sparse
      a0 += w * v_shaped[3 * v];
      a1 += w * v_shaped[3 * v + 1];
      a2 += w * v_shaped[3 * v + 2];
    }
  }
  s_red[0][threadIdx.x] = a0; s_red[1][threadIdx.x] = a1; s_red[2][threadIdx.x] = a2;
  __syncthreads();
  for (int o = SMPLX_THREADS / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      s_red[0][threadIdx.x] += s_red[0][threadIdx.x + o];
      s_red[1][threadIdx.x] += s_red[1][threadIdx.x + o];
      s_red[2][threadIdx.x] += s_red[2][threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x < 3) joints[3 * j + threadIdx.x] = s_red[threadIdx.x][0];
}
