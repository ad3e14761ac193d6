// Shared helpers for the sm_100a kernels: error plumbing, tile constants,
// column-major 3x3 algebra with a fixed summation order, relaxed loads/stores
// for the decoupled look-back protocols.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/soar_b200.h"

namespace soar {

// Tile geometry is part of the bit-exact contract (DGR/cuda_rasterizer/config.h:14-16).
constexpr int TILE_X = 16;
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;

void set_error(const char* fmt, ...);

// Kernel ids for launch accounting / event timing (prof.cu). Every launch site
// opens a KernelScope right before its <<<>>>.
enum KernelId {
  K_PREPROCESS = 0, K_SCAN, K_DUPLICATE, K_RADIX_HIST, K_RADIX_HIST_SCAN, K_RADIX_ONESWEEP,
  K_TILE_RANGES, K_BLEND_FWD, K_BLEND_BWD, K_GEOM_BWD, K_KNN_BBOX, K_KNN_MORTON, K_KNN_GATHER,
  K_KNN_BOXES, K_KNN_SEARCH, K_LBS_FWD, K_LBS_BWD, K_LBS_QUERY, K_SMPLX_SHAPE, K_SMPLX_POSEDIRS,
  K_SMPLX_JOINTS, K_SMPLX_CHAIN, K_SMPLX_SKIN, K_POST_FWD, K_POST_BWD, K_LBS_QUERY_BUILD, K_ADAM_STEP, K_DENSIFY_STATS, K_DENSIFY_SELECT,
  K_MASK_INDEX, K_ROWS_GATHER, K_SPLIT_ROWS, K_OCCGRID, K_CONTOUR_PAD, K_LBS_QUERY_KEYS, K_LBS_QUERY_TAIL,
  K_COUNT
};

class KernelScope {
 public:
  KernelScope(int id, cudaStream_t stream);
  ~KernelScope();
  // a scope that holds more than one <<<>>> (the blend's two layouts) counts the extra launches here
  static void count_extra_launch();

 private:
  int id_;
  cudaStream_t stream_;
  void* e1_;
};

#define SOAR_CUDA_TRY(expr)                                                          \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::soar::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                        __FILE__, __LINE__);                                         \
      return SOAR_E_CUDA;                                                            \
    }                                                                                \
  } while (0)

#define SOAR_LAUNCH_CHECK(name)                                                      \
  do {                                                                               \
    cudaError_t _e = cudaGetLastError();                                             \
    if (_e != cudaSuccess) {                                                         \
      ::soar::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));    \
      return SOAR_E_CUDA;                                                            \
    }                                                                                \
  } while (0)

__host__ __device__ inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Number of tile-id bits the pair sort runs over: the bit length of the LARGEST tile id,
// n_tiles - 1 (at least 1 bit). This deliberately differs from the reference's getHigherMsb(n_tiles)
// (DGR/cuda_rasterizer/rasterizer_impl.cu:35-48), which returns the bit length of n_tiles itself and
// so spends one bit too many whenever n_tiles is a power of two (4096 tiles: 13 there, 12 here).
// Ids stay below n_tiles in both sort directions (the descending flip is applied to the digit, not
// the id), so the sorted order is the same; tests/test_host_logic.py pins 1, 2, 4096 and 4097.
inline int tile_id_bits(uint32_t n_tiles) {
  const uint32_t top = n_tiles > 1 ? n_tiles - 1 : 1;
  int b = 0;
  while (b < 32 && (top >> b) != 0) ++b;
  return b;
}

// ---- column-major 3x3 (m[c][r]) with the summation order k = 0,1,2 -----------
// The per-Gaussian algebra must round like the reference's (which goes through
// a column-major header library); products are therefore always accumulated
// over the inner index in ascending order, left-associated.
struct Mat3 {
  float m[3][3];  // m[col][row]
};

__device__ __forceinline__ Mat3 mat3_cols(float a, float b, float c, float d, float e, float f,
                                          float g, float h, float i) {
  Mat3 r;
  r.m[0][0] = a; r.m[0][1] = b; r.m[0][2] = c;
  r.m[1][0] = d; r.m[1][1] = e; r.m[1][2] = f;
  r.m[2][0] = g; r.m[2][1] = h; r.m[2][2] = i;
  return r;
}

__device__ __forceinline__ Mat3 mat3_mul(const Mat3& A, const Mat3& B) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int w = 0; w < 3; ++w)
      r.m[c][w] = A.m[0][w] * B.m[c][0] + A.m[1][w] * B.m[c][1] + A.m[2][w] * B.m[c][2];
  return r;
}

__device__ __forceinline__ Mat3 mat3_transpose(const Mat3& A) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int w = 0; w < 3; ++w) r.m[c][w] = A.m[w][c];
  return r;
}

// ---- matrices in the row-vector convention, read column-major -----------------
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m) {
  return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                     m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                     m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]);
}
__device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* m) {
  return make_float4(m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12],
                     m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
                     m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14],
                     m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]);
}
__device__ __forceinline__ float3 xform_vec_4x3(const float3& p, const float* m) {
  return make_float3(m[0] * p.x + m[4] * p.y + m[8] * p.z,
                     m[1] * p.x + m[5] * p.y + m[9] * p.z,
                     m[2] * p.x + m[6] * p.y + m[10] * p.z);
}
__device__ __forceinline__ float3 xform_vec_4x3_T(const float3& p, const float* m) {
  return make_float3(m[0] * p.x + m[1] * p.y + m[2] * p.z,
                     m[4] * p.x + m[5] * p.y + m[6] * p.z,
                     m[8] * p.x + m[9] * p.y + m[10] * p.z);
}

// ---- relaxed gpu-scope accesses for single-word status protocols ---------------
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- 16-byte rows behind caller-owned pointers ------------------------------------------------
// [P,4] tensors (rotations and their gradient) are moved as float4 when the base pointer is 16-byte
// aligned — always true for a torch allocation, NOT for a slice of a flat buffer at an odd float
// offset (FlatGrads slices are padded for that reason). The launchers test the pointers and hand the
// kernels a uniform flag; the scalar path is the fallback.
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
__device__ __forceinline__ float4 ld_row4(const float* base, size_t row, bool vec) {
  if (vec) return reinterpret_cast<const float4*>(base)[row];
  const float* p = base + 4 * row;
  return make_float4(p[0], p[1], p[2], p[3]);
}
__device__ __forceinline__ void st_row4(float* base, size_t row, float4 v, bool vec) {
  if (vec) { reinterpret_cast<float4*>(base)[row] = v; return; }
  float* p = base + 4 * row;
  p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

}  // namespace soar
