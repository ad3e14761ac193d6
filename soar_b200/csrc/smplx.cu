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

// ---- 4. kinematic chain (one CTA of SMPLX_MAX_J threads, thread j = joint j) ------------------
// G[j] = G[parent] . [R_j | J_j - J_parent]; joints at the same depth are composed together.
__global__ void __launch_bounds__(SMPLX_MAX_J)
smplx_chain_kernel(const float* __restrict__ full_pose, const int32_t* __restrict__ parents,
                   const float* __restrict__ joints, const float* __restrict__ transl,
                   const float* __restrict__ inv_mats, int J, float* __restrict__ A_skin,
                   float* __restrict__ out_joints, float* __restrict__ out_A,
                   float* __restrict__ out_bones) {
  __shared__ float s_G[SMPLX_MAX_J][12];  // rows 0..2 of the 4x4, row-major
  __shared__ int s_depth[SMPLX_MAX_J];
  __shared__ int s_maxdepth;
  const int j = threadIdx.x;
  float L[12];  // local transform
  int parent = -1, depth = 0;
  if (j < J) {
    float R[9];
    rodrigues(full_pose + 3 * j, R);
    parent = j == 0 ? -1 : parents[j];
    float t[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      t[k] = joints[3 * j + k] - (parent >= 0 ? joints[3 * parent + k] : 0.f);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      L[r * 4 + 0] = R[r * 3 + 0]; L[r * 4 + 1] = R[r * 3 + 1]; L[r * 4 + 2] = R[r * 3 + 2];
      L[r * 4 + 3] = t[r];
    }
    for (int p = parent; p >= 0 && depth < SMPLX_MAX_J; p = (p == 0 ? -1 : parents[p])) ++depth;
    s_depth[j] = depth;
    if (depth == 0) {
#pragma unroll
      for (int k = 0; k < 12; ++k) s_G[j][k] = L[k];
    }
  }
  if (j == 0) s_maxdepth = 0;
  __syncthreads();
  if (j < J) atomicMax(&s_maxdepth, depth);
  __syncthreads();
  const int maxdepth = s_maxdepth;
  for (int lvl = 1; lvl <= maxdepth; ++lvl) {
    if (j < J && depth == lvl) {
      const float* P = s_G[parent];
      float G[12];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float acc = P[r * 4 + 0] * L[0 * 4 + c] + P[r * 4 + 1] * L[1 * 4 + c] + P[r * 4 + 2] * L[2 * 4 + c];
          if (c == 3) acc += P[r * 4 + 3];  // times the local's bottom-right 1
          G[r * 4 + c] = acc;
        }
      }
#pragma unroll
      for (int k = 0; k < 12; ++k) s_G[j][k] = G[k];
    }
    __syncthreads();
  }
  if (j >= J) return;
  float G[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) G[k] = s_G[j][k];
  const float tx = transl ? transl[0] : 0.f, ty = transl ? transl[1] : 0.f, tz = transl ? transl[2] : 0.f;
  const float tr[3] = {tx, ty, tz};
  float A[12];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float gj = G[r * 4 + 0] * joints[3 * j] + G[r * 4 + 1] * joints[3 * j + 1] + G[r * 4 + 2] * joints[3 * j + 2];
    A[r * 4 + 0] = G[r * 4 + 0]; A[r * 4 + 1] = G[r * 4 + 1]; A[r * 4 + 2] = G[r * 4 + 2];
    A[r * 4 + 3] = G[r * 4 + 3] - gj;
    if (out_joints) out_joints[3 * j + r] = G[r * 4 + 3] + tr[r];
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) A_skin[12 * j + k] = A[k];
#pragma unroll
  for (int r = 0; r < 3; ++r) A[r * 4 + 3] += tr[r];
  if (out_A) {
#pragma unroll
    for (int k = 0; k < 12; ++k) out_A[16 * j + k] = A[k];
    out_A[16 * j + 12] = 0.f; out_A[16 * j + 13] = 0.f; out_A[16 * j + 14] = 0.f; out_A[16 * j + 15] = 1.f;
  }
  if (out_bones && inv_mats) {
    const float* B = inv_mats + 16 * j;
    const float Afull[16] = {A[0], A[1], A[2], A[3], A[4], A[5], A[6], A[7],
                             A[8], A[9], A[10], A[11], 0.f, 0.f, 0.f, 1.f};
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        out_bones[16 * j + 4 * r + c] = Afull[4 * r + 0] * B[0 * 4 + c] + Afull[4 * r + 1] * B[1 * 4 + c] +
                                        Afull[4 * r + 2] * B[2 * 4 + c] + Afull[4 * r + 3] * B[3 * 4 + c];
  }
}

// ---- 5. vertex skinning --------------------------------------------------------------------
__global__ void __launch_bounds__(SMPLX_THREADS)
smplx_skin_kernel(const float* __restrict__ v_shaped, const float* __restrict__ partial,
                  const float* __restrict__ lbs_weights, const float* __restrict__ A_skin,
                  const float* __restrict__ transl, int V, int J, float* __restrict__ out_verts) {
  __shared__ float s_A[SMPLX_MAX_J * 12];
  for (int i = threadIdx.x; i < J * 12; i += SMPLX_THREADS) s_A[i] = A_skin[i];
  __syncthreads();
  const int v = blockIdx.x * SMPLX_THREADS + threadIdx.x;
  if (v >= V) return;
  const size_t C = (size_t)3 * V;
  float p[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float off = 0.f;
#pragma unroll
    for (int s = 0; s < SMPLX_PD_SPLIT; ++s) off += partial[s * C + 3 * v + k];
    p[k] = off + v_shaped[3 * v + k];
  }
  float T[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) T[k] = 0.f;
  const float* w = lbs_weights + (size_t)v * J;
  for (int j = 0; j < J; ++j) {
    const float wj = w[j];
    if (wj != 0.f) {
#pragma unroll
      for (int k = 0; k < 12; ++k) T[k] += wj * s_A[12 * j + k];
    }
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float o = T[r * 4 + 0] * p[0] + T[r * 4 + 1] * p[1] + T[r * 4 + 2] * p[2] + T[r * 4 + 3];
    out_verts[3 * v + r] = o + (transl ? transl[r] : 0.f);
  }
}

}  // namespace soar

using namespace soar;

extern "C" {

size_t soar_smplx_workspace_bytes(int32_t V, int32_t J) {
  return smplx_layout(V < 0 ? 0 : (size_t)V, J < 0 ? 0 : (size_t)J).total;
}

int soar_smplx_forward(const SoarSmplxModel* m, const float* betas, const float* full_pose,
                       const float* transl, const float* inv_mats, float* out_verts,
                       float* out_joints, float* out_A, float* out_bones, void* workspace,
                       size_t workspace_bytes, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (!m) { set_error("soar_smplx_forward: model is NULL"); return SOAR_E_INVALID; }
  const int V = m->V, J = m->J, NB = m->NB;
  if (V <= 0 || J <= 0 || J > SMPLX_MAX_J || NB < 0 || NB > SMPLX_MAX_NB) {
    set_error("soar_smplx_forward: bad sizes V=%d J=%d (<= %d) NB=%d (<= %d)", V, J, SMPLX_MAX_J, NB,
              SMPLX_MAX_NB);
    return SOAR_E_INVALID;
  }
  if (!m->v_template || !m->posedirs || !m->J_regressor || !m->parents || !m->lbs_weights ||
      (NB > 0 && (!m->shapedirs || !betas)) || !full_pose) {
    set_error("soar_smplx_forward: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  if (out_bones && !inv_mats) {
    set_error("soar_smplx_forward: out_bones needs inv_mats");
    return SOAR_E_INVALID;
  }
  const SmplxLayout L = smplx_layout(V, J);
  if (!workspace || workspace_bytes < L.total) {
    set_error("soar_smplx_forward: workspace holds %zu bytes, need %zu", workspace_bytes, L.total);
    return SOAR_E_WORKSPACE;
  }
  char* ws = reinterpret_cast<char*>(workspace);
  float* v_shaped = reinterpret_cast<float*>(ws + L.v_shaped);
  float* partial = reinterpret_cast<float*>(ws + L.partial);
  float* joints = reinterpret_cast<float*>(ws + L.joints);
  float* A_skin = reinterpret_cast<float*>(ws + L.A);
  const int C = 3 * V;
  const int n_feat = (J - 1) * 9;
  const int rows = ceil_div(n_feat, SMPLX_PD_SPLIT);
  {
    KernelScope ks(K_SMPLX_SHAPE, stream);
    smplx_shape_kernel<<<ceil_div(C, SMPLX_THREADS), SMPLX_THREADS, 0, stream>>>(
        m->v_template, m->shapedirs, betas, C, NB, v_shaped);
  }
  SOAR_LAUNCH_CHECK("smplx_shape_kernel");
  {
    KernelScope ks(K_SMPLX_POSEDIRS, stream);
    dim3 grid(ceil_div(C, SMPLX_THREADS), SMPLX_PD_SPLIT);
    smplx_posedirs_kernel<<<grid, SMPLX_THREADS, (size_t)max(rows, 1) * sizeof(float), stream>>>(
        m->posedirs, full_pose, C, n_feat, rows, partial);
  }
  SOAR_LAUNCH_CHECK("smplx_posedirs_kernel");
  {
    KernelScope ks(K_SMPLX_JOINTS, stream);
    smplx_joints_kernel<<<J, SMPLX_THREADS, 0, stream>>>(m->J_regressor, v_shaped, V, joints);
  }
  SOAR_LAUNCH_CHECK("smplx_joints_kernel");
  {
    KernelScope ks(K_SMPLX_CHAIN, stream);
    smplx_chain_kernel<<<1, SMPLX_MAX_J, 0, stream>>>(full_pose, m->parents, joints, transl, inv_mats,
                                                      J, A_skin, out_joints, out_A, out_bones);
  }
  SOAR_LAUNCH_CHECK("smplx_chain_kernel");
  if (out_verts) {
    KernelScope ks(K_SMPLX_SKIN, stream);
    smplx_skin_kernel<<<ceil_div(V, SMPLX_THREADS), SMPLX_THREADS, 0, stream>>>(
        v_shaped, partial, m->lbs_weights, A_skin, transl, V, J, out_verts);
  }
  SOAR_LAUNCH_CHECK("smplx_skin_kernel");
  return SOAR_OK;
}

}  // extern "C"
