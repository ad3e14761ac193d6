// SMPL-X linear-blend-skinning of canonical Gaussians for sm_100a.
//
// Behavioural contract (all torch ops in the reference):
//   * per-point transform = sum_j weights[n,j] * bones[j]      TS/utils/smpl.py:613
//   * x' = M[:3,:3] x + M[:3,3]                                 TS/renderer/diff_gaussian_rasterizer.py:103-106
//   * SDS views: (x,y,z) -> (z,x,y) and R' = T^T M3 R(q)        :109-112, :321-352
//   * gt views: no swizzle                                       :138-149
//   * q' = normalize(matrix_to_quaternion(R'))                   :113-114 (pytorch3d, unvendored:
//     4 candidates from q_abs = sqrt(max(0, 1 +- m00 +- m11 +- m22)), divided by
//     2*max(q_abs, 0.1), the candidate with the largest q_abs is selected, real
//     part made non-negative; quaternion_to_matrix uses two_s = 2/|q|^2 — a copy of
//     that one is vendored at TS/data/uncond_multiview.py:2422-2451)
//   * gradients flow to points and quaternions only (weights are detached,
//     TS/utils/smpl.py:611; bones are plain buffers).
//
// K = 55 is not a dense contraction worth the tensor pipe: the kernel is bound
// by the 220-byte weight row per point. A CTA stages its 128 rows with
// coalesced 16-byte loads into shared memory (row stride 55 words is odd, so
// the per-thread row walk is bank-conflict free) and the 55 bone matrices are
// read as shared-memory broadcasts.
#include "common.cuh"

namespace soar {

constexpr int LBS_THREADS = 128;
constexpr int LBS_MAX_J = 64;

struct Rot9 {
  float m[9];  // row-major
};

__device__ __forceinline__ Rot9 quat_to_matrix_p3d(const float4 q) {
  const float r = q.x, i = q.y, j = q.z, k = q.w;
  const float two_s = 2.0f / (r * r + i * i + j * j + k * k);
  Rot9 o;
  o.m[0] = 1 - two_s * (j * j + k * k);
  o.m[1] = two_s * (i * j - k * r);
  o.m[2] = two_s * (i * k + j * r);
  o.m[3] = two_s * (i * j + k * r);
  o.m[4] = 1 - two_s * (i * i + k * k);
  o.m[5] = two_s * (j * k - i * r);
  o.m[6] = two_s * (i * k - j * r);
  o.m[7] = two_s * (j * k + i * r);
  o.m[8] = 1 - two_s * (i * i + j * j);
  return o;
}

// candidate selection state shared by forward and backward
struct M2Q {
  int sel;
  float a;        // q_abs[sel]
  float u;        // argument of the sqrt
  float den;      // 2*max(a, 0.1)
  float n[4];     // numerators
  float c[4];     // candidate = n/den
  float norm;     // |c| (clamped at 1e-12)
  float sign;     // standardisation sign
};

__device__ __forceinline__ M2Q matrix_to_quat_p3d(const Rot9& R, float* out) {
  const float m00 = R.m[0], m01 = R.m[1], m02 = R.m[2], m10 = R.m[3], m11 = R.m[4], m12 = R.m[5],
              m20 = R.m[6], m21 = R.m[7], m22 = R.m[8];
  const float u[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22,
                      1.0f - m00 - m11 + m22};
  float qa[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) qa[t] = u[t] > 0.f ? sqrtf(u[t]) : 0.f;
  int sel = 0;  // argmax, first maximum wins
#pragma unroll
  for (int t = 1; t < 4; ++t)
    if (qa[t] > qa[sel]) sel = t;
  M2Q s;
  s.sel = sel;
  s.a = qa[sel];
  s.u = u[sel];
  s.den = 2.0f * fmaxf(s.a, 0.1f);
  const float a2 = s.a * s.a;
  if (sel == 0) { s.n[0] = a2; s.n[1] = m21 - m12; s.n[2] = m02 - m20; s.n[3] = m10 - m01; }
  else if (sel == 1) { s.n[0] = m21 - m12; s.n[1] = a2; s.n[2] = m10 + m01; s.n[3] = m02 + m20; }
  else if (sel == 2) { s.n[0] = m02 - m20; s.n[1] = m10 + m01; s.n[2] = a2; s.n[3] = m12 + m21; }
  else { s.n[0] = m10 - m01; s.n[1] = m20 + m02; s.n[2] = m21 + m12; s.n[3] = a2; }
#pragma unroll
  for (int t = 0; t < 4; ++t) s.c[t] = s.n[t] / s.den;
  s.sign = s.c[0] < 0.f ? -1.f : 1.f;
  s.norm = fmaxf(sqrtf(s.c[0] * s.c[0] + s.c[1] * s.c[1] + s.c[2] * s.c[2] + s.c[3] * s.c[3]), 1e-12f);
#pragma unroll
  for (int t = 0; t < 4; ++t) out[t] = s.sign * s.c[t] / s.norm;
  return s;
}

// rows of (T^T A) for the SDS swizzle "+z,+x,+y": new rows = (row2, row0, row1)
__device__ __forceinline__ int perm_row(int axis_perm, int r) {
  return axis_perm ? (r == 0 ? 2 : r - 1) : r;
}

__global__ void __launch_bounds__(LBS_THREADS)
lbs_fwd_kernel(const float* __restrict__ weights, const float* __restrict__ bones, int J,
               const float* __restrict__ points, const float* __restrict__ quats, int N,
               int axis_perm, float* __restrict__ out_points, float* __restrict__ out_quats,
               float* __restrict__ out_mats) {
  extern __shared__ float smem[];
  float* s_bones = smem;            // [J][12]
  float* s_w = smem + LBS_MAX_J * 12;  // [LBS_THREADS][J]
  const int tid = threadIdx.x;
  const size_t base = (size_t)blockIdx.x * LBS_THREADS;
  const int n_here = (int)min((size_t)LBS_THREADS, (size_t)N - base);

  for (int e = tid; e < J * 12; e += LBS_THREADS) {
    const int j = e / 12, k = e % 12;
    s_bones[e] = bones[j * 16 + k];  // first three rows of the 4x4
  }
  {
    // coalesced copy of this CTA's [n_here x J] block of weights
    const float* src = weights + base * J;
    const int total = n_here * J;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
    const int head = min(total, (int)(((16 - (addr & 15)) & 15) / 4));
    if (tid < head) s_w[tid] = src[tid];
    const int n_vec = (total - head) / 4;
    const float4* src4 = reinterpret_cast<const float4*>(src + head);
    for (int v = tid; v < n_vec; v += LBS_THREADS) {
      const float4 x = __ldg(src4 + v);
      float* d = s_w + head + 4 * v;
      d[0] = x.x; d[1] = x.y; d[2] = x.z; d[3] = x.w;
    }
    for (int e = head + 4 * n_vec + tid; e < total; e += LBS_THREADS) s_w[e] = src[e];
  }
  __syncthreads();
  if (tid >= n_here) return;
  const size_t n = base + tid;

  float M[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) M[k] = 0.f;
  const float* w = s_w + tid * J;
  for (int j = 0; j < J; ++j) {
    const float wj = w[j];
    const float4* b = reinterpret_cast<const float4*>(s_bones + j * 12);
    const float4 b0 = b[0], b1 = b[1], b2 = b[2];
    M[0] += wj * b0.x; M[1] += wj * b0.y; M[2] += wj * b0.z; M[3] += wj * b0.w;
    M[4] += wj * b1.x; M[5] += wj * b1.y; M[6] += wj * b1.z; M[7] += wj * b1.w;
    M[8] += wj * b2.x; M[9] += wj * b2.y; M[10] += wj * b2.z; M[11] += wj * b2.w;
  }
  if (out_mats) {
    float4* o = reinterpret_cast<float4*>(out_mats + n * 12);
    o[0] = make_float4(M[0], M[1], M[2], M[3]);
    o[1] = make_float4(M[4], M[5], M[6], M[7]);
    o[2] = make_float4(M[8], M[9], M[10], M[11]);
  }

  const float px = points[3 * n], py = points[3 * n + 1], pz = points[3 * n + 2];
  float x[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) x[r] = M[4 * r] * px + M[4 * r + 1] * py + M[4 * r + 2] * pz + M[4 * r + 3];
#pragma unroll
  for (int r = 0; r < 3; ++r) out_points[3 * n + r] = x[perm_row(axis_perm, r)];

  const float4 q = reinterpret_cast<const float4*>(quats)[n];
  const Rot9 R = quat_to_matrix_p3d(q);
  Rot9 A;  // M3 * R, rows permuted
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int sr = perm_row(axis_perm, r);
#pragma unroll
    for (int c = 0; c < 3; ++c)
      A.m[3 * r + c] = M[4 * sr] * R.m[c] + M[4 * sr + 1] * R.m[3 + c] + M[4 * sr + 2] * R.m[6 + c];
  }
  float qo[4];
  matrix_to_quat_p3d(A, qo);
  reinterpret_cast<float4*>(out_quats)[n] = make_float4(qo[0], qo[1], qo[2], qo[3]);
}

__global__ void __launch_bounds__(256)
lbs_bwd_kernel(const float* __restrict__ mats, const float* __restrict__ quats, int N,
               int axis_perm, const float* __restrict__ g_points, const float* __restrict__ g_quats,
               float* __restrict__ d_points, float* __restrict__ d_quats) {
  const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (size_t)N) return;
  const float4* mp = reinterpret_cast<const float4*>(mats + n * 12);
  const float4 r0 = mp[0], r1 = mp[1], r2 = mp[2];
  const float M[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};

  // ---- points: x' = P (M3 x + t)  ->  dx = M3^T P^T g ---------------------------------
  float gx[3];  // gradient w.r.t. the un-permuted M3 x + t
#pragma unroll
  for (int r = 0; r < 3; ++r) gx[perm_row(axis_perm, r)] = g_points[3 * n + r];
#pragma unroll
  for (int c = 0; c < 3; ++c) d_points[3 * n + c] = M[c] * gx[0] + M[4 + c] * gx[1] + M[8 + c] * gx[2];

  // ---- quaternion: forward replay --------------------------------------------------------
  const float4 q = reinterpret_cast<const float4*>(quats)[n];
  const Rot9 R = quat_to_matrix_p3d(q);
  Rot9 A;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int sr = perm_row(axis_perm, r);
#pragma unroll
    for (int c = 0; c < 3; ++c)
      A.m[3 * r + c] = M[4 * sr] * R.m[c] + M[4 * sr + 1] * R.m[3 + c] + M[4 * sr + 2] * R.m[6 + c];
  }
  float qo[4];
  const M2Q s = matrix_to_quat_p3d(A, qo);

  const float4 g4 = reinterpret_cast<const float4*>(g_quats)[n];
  const float g[4] = {g4.x, g4.y, g4.z, g4.w};
  // normalize + sign: dL/dc = sign * (g - q'(q'.g)) / |c|
  const float qg = qo[0] * g[0] + qo[1] * g[1] + qo[2] * g[2] + qo[3] * g[3];
  float dc[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) dc[t] = s.sign * (g[t] - qo[t] * qg) / s.norm;
  // c = n / den
  float dn[4], dden = 0.f;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    dn[t] = dc[t] / s.den;
    dden -= dc[t] * s.c[t] / s.den;
  }
  float da = dn[s.sel] * 2.0f * s.a + (s.a > 0.1f ? 2.0f * dden : 0.f);
  const float du = s.u > 0.f ? da / (2.0f * s.a) : 0.f;

  float dA[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) dA[t] = 0.f;
  // off-diagonal numerators; index helpers: m(r,c) = 3r+c
  if (s.sel == 0) {
    dA[7] += dn[1]; dA[5] -= dn[1];   // m21 - m12
    dA[2] += dn[2]; dA[6] -= dn[2];   // m02 - m20
    dA[3] += dn[3]; dA[1] -= dn[3];   // m10 - m01
    dA[0] += du; dA[4] += du; dA[8] += du;
  } else if (s.sel == 1) {
    dA[7] += dn[0]; dA[5] -= dn[0];   // m21 - m12
    dA[3] += dn[2]; dA[1] += dn[2];   // m10 + m01
    dA[2] += dn[3]; dA[6] += dn[3];   // m02 + m20
    dA[0] += du; dA[4] -= du; dA[8] -= du;
  } else if (s.sel == 2) {
    dA[2] += dn[0]; dA[6] -= dn[0];   // m02 - m20
    dA[3] += dn[1]; dA[1] += dn[1];   // m10 + m01
    dA[5] += dn[3]; dA[7] += dn[3];   // m12 + m21
    dA[0] -= du; dA[4] += du; dA[8] -= du;
  } else {
    dA[3] += dn[0]; dA[1] -= dn[0];   // m10 - m01
    dA[6] += dn[1]; dA[2] += dn[1];   // m20 + m02
    dA[7] += dn[2]; dA[5] += dn[2];   // m21 + m12
    dA[0] -= du; dA[4] -= du; dA[8] += du;
  }
  // A = P M3 R  ->  dR = M3^T P^T dA
  float dR[9];
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float acc = 0.f;
#pragma unroll
      for (int r = 0; r < 3; ++r) acc += M[4 * perm_row(axis_perm, r) + k] * dA[3 * r + c];
      dR[3 * k + c] = acc;
    }
  // R = quaternion_to_matrix(q) with two_s = 2/|q|^2
  const float qr = q.x, qi = q.y, qj = q.z, qk = q.w;
  const float nn = qr * qr + qi * qi + qj * qj + qk * qk;
  const float two_s = 2.0f / nn;
  // R = I' + two_s * B(q): entries e_t = const_t + two_s * b_t
  const float b[9] = {-(qj * qj + qk * qk), qi * qj - qk * qr, qi * qk + qj * qr,
                      qi * qj + qk * qr, -(qi * qi + qk * qk), qj * qk - qi * qr,
                      qi * qk - qj * qr, qj * qk + qi * qr, -(qi * qi + qj * qj)};
  float dtwo_s = 0.f;
#pragma unroll
  for (int t = 0; t < 9; ++t) dtwo_s += dR[t] * b[t];
  float db[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) db[t] = dR[t] * two_s;
  const float dnn = -dtwo_s * two_s / nn;  // d(2/nn) = -2/nn^2
  float dqr = 2.0f * qr * dnn, dqi = 2.0f * qi * dnn, dqj = 2.0f * qj * dnn, dqk = 2.0f * qk * dnn;
  // b0 = -(j^2+k^2)
  dqj += -2.0f * qj * db[0]; dqk += -2.0f * qk * db[0];
  // b1 = ij - kr
  dqi += qj * db[1]; dqj += qi * db[1]; dqk += -qr * db[1]; dqr += -qk * db[1];
  // b2 = ik + jr
  dqi += qk * db[2]; dqk += qi * db[2]; dqj += qr * db[2]; dqr += qj * db[2];
  // b3 = ij + kr
  dqi += qj * db[3]; dqj += qi * db[3]; dqk += qr * db[3]; dqr += qk * db[3];
  // b4 = -(i^2+k^2)
  dqi += -2.0f * qi * db[4]; dqk += -2.0f * qk * db[4];
  // b5 = jk - ir
  dqj += qk * db[5]; dqk += qj * db[5]; dqi += -qr * db[5]; dqr += -qi * db[5];
  // b6 = ik - jr
  dqi += qk * db[6]; dqk += qi * db[6]; dqj += -qr * db[6]; dqr += -qj * db[6];
  // b7 = jk + ir
  dqj += qk * db[7]; dqk += qj * db[7]; dqi += qr * db[7]; dqr += qi * db[7];
  // b8 = -(i^2+j^2)
  dqi += -2.0f * qi * db[8]; dqj += -2.0f * qj * db[8];
  reinterpret_cast<float4*>(d_quats)[n] = make_float4(dqr, dqi, dqj, dqk);
}

}  // namespace soar

using namespace soar;

extern "C" {

int soar_lbs_forward(const float* weights, const float* bones, int32_t J, const float* points,
                     const float* quats, int32_t N, int32_t axis_perm, float* out_points,
                     float* out_quats, float* out_mats, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0 || J <= 0 || J > LBS_MAX_J) {
    set_error("soar_lbs_forward: bad sizes N=%d J=%d (J <= %d)", N, J, LBS_MAX_J);
    return SOAR_E_INVALID;
  }
  if (N == 0) return SOAR_OK;
  if (!weights || !bones || !points || !quats || !out_points || !out_quats) {
    set_error("soar_lbs_forward: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  if (!aligned16(quats) || !aligned16(out_quats) || (out_mats && !aligned16(out_mats))) {
    set_error("soar_lbs_forward: quats / out_quats / out_mats must be 16-byte aligned");
    return SOAR_E_INVALID;
  }
  const size_t smem = (size_t)(LBS_MAX_J * 12 + LBS_THREADS * J) * sizeof(float);
  const int grid = ceil_div(N, LBS_THREADS);
  {
    KernelScope ks(K_LBS_FWD, stream);
    lbs_fwd_kernel<<<grid, LBS_THREADS, smem, stream>>>(weights, bones, J, points, quats, N,
                                                        axis_perm, out_points, out_quats, out_mats);
  }
  SOAR_LAUNCH_CHECK("lbs_fwd_kernel");
  return SOAR_OK;
}

int soar_lbs_backward(const float* mats, const float* quats, int32_t N, int32_t axis_perm,
                      const float* dL_dout_points, const float* dL_dout_quats, float* dL_dpoints,
                      float* dL_dquats, soar_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (N < 0) { set_error("soar_lbs_backward: N < 0"); return SOAR_E_INVALID; }
  if (N == 0) return SOAR_OK;
  if (!mats || !quats || !dL_dout_points || !dL_dout_quats || !dL_dpoints || !dL_dquats) {
    set_error("soar_lbs_backward: a required pointer is NULL");
    return SOAR_E_INVALID;
  }
  if (!aligned16(mats) || !aligned16(quats) || !aligned16(dL_dout_quats) || !aligned16(dL_dquats)) {
    set_error("soar_lbs_backward: mats / quats / dL_dout_quats / dL_dquats must be 16-byte aligned");
    return SOAR_E_INVALID;
  }
  {
    KernelScope ks(K_LBS_BWD, stream);
    lbs_bwd_kernel<<<ceil_div(N, 256), 256, 0, stream>>>(mats, quats, N, axis_perm, dL_dout_points,
                                                         dL_dout_quats, dL_dpoints, dL_dquats);
  }
  SOAR_LAUNCH_CHECK("lbs_bwd_kernel");
  return SOAR_OK;
}

}  // extern "C"
